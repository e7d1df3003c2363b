// sparse_chol.cuh — numeric phase of the exact block-sparse Cholesky of the reduced camera matrix (sm_100a, fp64).
//
// The reference factors the reduced camera system exactly (Ceres SPARSE_SCHUR + CHOLMOD, src/bundle_adjuster.cpp:
// 26-29).  For the global BA calls (src/slam.cpp:156, src/pose_graph_optimizer.cpp:95: thousands of free cameras,
// banded co-visibility + loop-closure blocks) the whole solve S x = rhs is ONE cooperative launch:
//   0  scatter the blocks of S into the (nested-dissection ordered) pattern of L
//   1  elimination-tree levels, leaves first; the unit of work is a CHAIN (a run of columns each the only child of the
//      next: the separators and the top of the tree).  Per level:
//        phase A, whole grid   a warp per block / column of the level's chains gathers what comes from OUTSIDE its
//                              chain — finished columns of lower levels, through L2: the static list of products
//                              L_ik L_jk^T on the FP64 tensor cores (two m8n8k4 MMAs per product), and the
//                              right-hand-side terms  sum_k L_jk y_k  — and subtracts it in place;  grid barrier
//        phase B, CTA / chain  the chain's panel comes to shared memory and is factored column by column: in-chain
//                              products from the panel, 6x6 diagonal factor + inverse, panel  L_ij = A_ij Linv_jj^T,
//                              y_j = Linv_jj t_j;  written back;  grid barrier
//      (chains too large for the panel buffer go column by column through L2)
//   2  the same levels downwards:  x_j = Linv_jj^T (y_j - sum_i L_ij^T x_i); a chain first takes the contribution of
//      the rows outside it for all its columns at once, then one warp walks its columns
// One writer per block, every sum in list order: bit-reproducible, no atomics.  Symbolic phase: sparse_chol.cpp.
#pragma once
#include <cooperative_groups.h>
#include "kernels.cuh"

namespace sba {

struct SpCholDev {
  int32_t n, nlev, nnzL, panel_cap;
  int32_t profile;   // SBA_DEBUG_LS=2: block 0 prints the time of every tree level (tuning aid)
  const int32_t *perm, *col_ptr, *row_idx, *init_src;
  const int64_t *upd_ptr, *upd_ext_end;
  const int32_t *upd_a, *upd_b;
  const int32_t *rowl_ptr, *rowl_ext_end, *rowl_blk, *rowl_col, *lvl_ptr, *lvl_cols, *lvl_len;
  const int32_t *pa_blk_ptr, *pa_blk, *pa_col_ptr, *pa_col;   // per level: blocks with outside products / columns of the shared-memory chains
  double* Lv;     // [nnzL * 36] blocks of L, row-major 6x6 (block (i, j): rows of i, columns of j)
  double* Li;     // [n * 36]    inverse diagonal factors (lower)
  double* y;      // [n * 6]     forward-substituted right-hand side, then the solution (new order)
};

constexpr int kSpThreads = 512;
constexpr int kSpWarps = kSpThreads / 32;
constexpr int kSpStage = 64;       // fallback path: blocks of one column staged in shared memory between gather and panel
constexpr int kSpPanelMax = 400;   // blocks of one chain held in shared memory (115 KB)
constexpr int kSpChainMax = 64;    // columns of such a chain

// dynamic shared memory of k_spchol for chains of up to `cap` blocks: panel, external x, row indices, t / y / x of the
// chain's columns, their inverse factors
__host__ __device__ __forceinline__ size_t spchol_smem_bytes(int cap) {
  return ((size_t)cap * 36 + (size_t)cap * 6 + (size_t)(cap + 1) / 2 + kSpChainMax * 6 + kSpChainMax * 36 + 8) * 8;
}

// sum over the products [u0, u1) of one block's list of L_a L_b^T, on the FP64 tensor cores: two m8n8k4 MMAs per
// product, operands straight from L2 (four products = 16 loads in flight per lane before the first MMA), even and
// odd products on two accumulator chains.  Lane (g, t) = (lane / 4, lane % 4) ends up with elements 6 g + 2 t, + 1.
__device__ __forceinline__ void sp_gather(const SpCholDev& sc, int64_t u0, int64_t u1, int lane, double& r0, double& r1) {
  const int g = lane >> 2, t4 = lane & 3;
  const bool f0 = g < 6, f1 = g < 6 && t4 < 2;
  const int o0 = 6 * (f0 ? g : 0) + t4, o1 = 6 * (f1 ? g : 0) + (f1 ? t4 + 4 : 0);
  double c0 = 0.0, c1 = 0.0, e0 = 0.0, e1 = 0.0;
  for (int64_t ub = u0; ub < u1; ub += 32) {
    const int m = (int)min((int64_t)32, u1 - ub);
    const int ia = (lane < m) ? __ldg(sc.upd_a + ub + lane) : 0;
    const int ib = (lane < m) ? __ldg(sc.upd_b + ub + lane) : 0;
    int u = 0;
    for (; u + 3 < m; u += 4) {
      double a0[4], a1[4], b0[4], b1[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const double* A = sc.Lv + (size_t)__shfl_sync(0xffffffffu, ia, u + q) * 36;
        const double* B = sc.Lv + (size_t)__shfl_sync(0xffffffffu, ib, u + q) * 36;
        a0[q] = f0 ? __ldcg(A + o0) : 0.0; b0[q] = f0 ? __ldcg(B + o0) : 0.0;
        a1[q] = f1 ? __ldcg(A + o1) : 0.0; b1[q] = f1 ? __ldcg(B + o1) : 0.0;
      }
      dmma884(c0, c1, a0[0], b0[0]); dmma884(c0, c1, a1[0], b1[0]);
      dmma884(e0, e1, a0[1], b0[1]); dmma884(e0, e1, a1[1], b1[1]);
      dmma884(c0, c1, a0[2], b0[2]); dmma884(c0, c1, a1[2], b1[2]);
      dmma884(e0, e1, a0[3], b0[3]); dmma884(e0, e1, a1[3], b1[3]);
    }
    for (; u < m; ++u) {
      const double* A = sc.Lv + (size_t)__shfl_sync(0xffffffffu, ia, u) * 36;
      const double* B = sc.Lv + (size_t)__shfl_sync(0xffffffffu, ib, u) * 36;
      const double a00 = f0 ? __ldcg(A + o0) : 0.0, b00 = f0 ? __ldcg(B + o0) : 0.0;
      const double a01 = f1 ? __ldcg(A + o1) : 0.0, b01 = f1 ? __ldcg(B + o1) : 0.0;
      if (u & 1) { dmma884(e0, e1, a00, b00); dmma884(e0, e1, a01, b01); }
      else { dmma884(c0, c1, a00, b00); dmma884(c0, c1, a01, b01); }
    }
  }
  r0 = c0 + e0; r1 = c1 + e1;
}

// the same sum for products whose operands sit in the chain's shared-memory panel P (block ids relative to blk0)
__device__ __forceinline__ void sp_gather_panel(const SpCholDev& sc, const double* P, int blk0, int64_t u0, int64_t u1, int lane,
                                                double& r0, double& r1) {
  const int g = lane >> 2, t4 = lane & 3;
  const bool f0 = g < 6, f1 = g < 6 && t4 < 2;
  const int o0 = 6 * (f0 ? g : 0) + t4, o1 = 6 * (f1 ? g : 0) + (f1 ? t4 + 4 : 0);
  double c0 = 0.0, c1 = 0.0, e0 = 0.0, e1 = 0.0;
  for (int64_t ub = u0; ub < u1; ub += 32) {
    const int m = (int)min((int64_t)32, u1 - ub);
    const int ia = (lane < m) ? __ldg(sc.upd_a + ub + lane) - blk0 : 0;
    const int ib = (lane < m) ? __ldg(sc.upd_b + ub + lane) - blk0 : 0;
    for (int u = 0; u < m; ++u) {
      const double* A = P + (size_t)__shfl_sync(0xffffffffu, ia, u) * 36;
      const double* B = P + (size_t)__shfl_sync(0xffffffffu, ib, u) * 36;
      const double a00 = f0 ? A[o0] : 0.0, b00 = f0 ? B[o0] : 0.0;
      const double a01 = f1 ? A[o1] : 0.0, b01 = f1 ? B[o1] : 0.0;
      if (u & 1) { dmma884(e0, e1, a00, b00); dmma884(e0, e1, a01, b01); }
      else { dmma884(c0, c1, a00, b00); dmma884(c0, c1, a01, b01); }
    }
  }
  r0 = c0 + e0; r1 = c1 + e1;
}

__global__ void __launch_bounds__(kSpThreads, 1) k_spchol(Dev d, SpCholDev sc) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  extern __shared__ double dyn[];
  __shared__ double col_st[kSpStage * 36];
  __shared__ double part[kSpWarps * 36];      // slice partials when a block's product list is split over several warps
  __shared__ double akk[36], li_s[36], tj[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (d.nwin == 1 && (d.w_status[0] != 0 || d.w_fail[0] != 0)) {     // uniform over the grid; leaves x untouched (unused)
    if (blockIdx.x == 0 && tid == 0) d.w_pcg_it[0] = 0;
    return;
  }
  // chain panel in shared memory (chains of <= panel_cap blocks and <= kSpChainMax columns; others go column by column through L2)
  const int cap = sc.panel_cap;
  double* P = dyn;                                  // [cap * 36]
  double* xe = P + (size_t)cap * 36;                // [cap * 6]   backward: x of rows outside the chain
  int* rowi = reinterpret_cast<int*>(xe + (size_t)cap * 6);   // [cap] backward: block rows
  double* ty = xe + (size_t)cap * 6 + (cap + 1) / 2;          // [kSpChainMax * 6] t_j -> y_j -> x_j of the chain's columns
  double* LIc = ty + kSpChainMax * 6;               // [kSpChainMax * 36] their inverse diagonal factors
  // ---- 0: scatter S into the pattern of L
  {
    const int64_t tot = (int64_t)sc.nnzL * 36;
    for (int64_t idx = (int64_t)blockIdx.x * kSpThreads + tid; idx < tot; idx += (int64_t)gridDim.x * kSpThreads) {
      const int b = (int)(idx / 36), e = (int)(idx - 36 * (int64_t)b);
      const int src = __ldg(sc.init_src + b);
      double v = 0.0;
      if (src >= 0) v = d.Su[(size_t)src * 36 + e];
      else if (src != INT32_MIN) { const int r = e / 6, c = e - 6 * r; v = d.Su[(size_t)(~src) * 36 + 6 * c + r]; }
      __stcg(sc.Lv + idx, v);
    }
    for (int w = blockIdx.x * kSpThreads + tid; w < d.nwin; w += gridDim.x * kSpThreads) d.w_pcg_it[w] = 0;
  }
  grid.sync();
  const bool prof = sc.profile && blockIdx.x == 0 && tid == 0;
  unsigned long long tstamp[48];                    // local memory, touched in profile mode only
  int nstamp = 0;
  if (prof) tstamp[nstamp++] = gtimer();
  auto lap = [&]() { if (prof && nstamp < 48) tstamp[nstamp++] = gtimer(); };
  // ---- 1: factorisation + forward substitution: chains of one level in parallel, a chain's columns in sequence
  const int g = lane >> 2, t4 = lane & 3;
  const bool fout = g < 6 && t4 < 3;
  const int ps = lane / 6, ri = lane - 6 * ps;
  for (int lvl = 0; lvl < sc.nlev; ++lvl) {
    const int c_begin = __ldg(sc.lvl_ptr + lvl), c_end = __ldg(sc.lvl_ptr + lvl + 1);
    // ---- phase A, whole grid: everything that comes from OUTSIDE a chain — finished columns of lower levels, read
    //      through L2 — for all blocks and columns of the level at once, a warp per block / column.  In the upper
    //      levels a few chains own thousands of products; spread over the grid they cost one L2 round instead of dozens.
    {
      const int gw = blockIdx.x * kSpWarps + warp, nw = gridDim.x * kSpWarps;
      const int b_begin = __ldg(sc.pa_blk_ptr + lvl), b_end = __ldg(sc.pa_blk_ptr + lvl + 1);
      for (int w = b_begin + gw; w < b_end; w += nw) {
        const int blk = __ldg(sc.pa_blk + w);
        double r0, r1;
        sp_gather(sc, sc.upd_ptr[blk], sc.upd_ext_end[blk], lane, r0, r1);
        if (fout) {
          const int e = 6 * g + 2 * t4;
          double2* dst = reinterpret_cast<double2*>(sc.Lv + (size_t)blk * 36 + e);
          const double2 a = __ldcg(dst);
          __stcg(dst, make_double2(a.x - r0, a.y - r1));
        }
      }
      const int q_begin = __ldg(sc.pa_col_ptr + lvl), q_end = __ldg(sc.pa_col_ptr + lvl + 1);
      for (int w = q_begin + (nw - 1 - gw); w < q_end; w += nw) {       // columns from the other end of the grid
        const int j = __ldg(sc.pa_col + w);
        double acc = 0.0;
        if (lane < 30) {
          const int r1 = __ldg(sc.rowl_ext_end + j);
          for (int e = __ldg(sc.rowl_ptr + j) + ps; e < r1; e += 5) {
            const double* L = sc.Lv + (size_t)__ldg(sc.rowl_blk + e) * 36 + 6 * ri;
            const double* yk = sc.y + (size_t)__ldg(sc.rowl_col + e) * 6;
            acc += (__ldcg(L) * __ldcg(yk) + __ldcg(L + 1) * __ldcg(yk + 1) + __ldcg(L + 2) * __ldcg(yk + 2)) +
                   (__ldcg(L + 3) * __ldcg(yk + 3) + __ldcg(L + 4) * __ldcg(yk + 4) + __ldcg(L + 5) * __ldcg(yk + 5));
          }
        }
        double tot = acc;
#pragma unroll
        for (int qq = 1; qq < 5; ++qq) tot += __shfl_sync(0xffffffffu, acc, (lane + 6 * qq) & 31);
        if (lane < 6) __stcg(sc.y + (size_t)j * 6 + lane, d.rhs[(size_t)__ldg(sc.perm + j) * 6 + lane] - tot);
      }
    }
    grid.sync();
    for (int ci = c_begin + blockIdx.x; ci < c_end; ci += gridDim.x) {
      const int j0 = __ldg(sc.lvl_cols + ci), len = __ldg(sc.lvl_len + ci), jn = j0 + len;
      const int blk0 = __ldg(sc.col_ptr + j0), nblk = __ldg(sc.col_ptr + jn) - blk0;
      if (nblk <= cap && len <= kSpChainMax) {
        // ======== the chain's panel lives in shared memory
        // S1: the panel (outside products already subtracted by phase A) and t of its columns come in coalesced
        for (int i = tid; i < nblk * 36; i += kSpThreads) P[i] = __ldcg(sc.Lv + (size_t)blk0 * 36 + i);
        for (int i = tid; i < len * 6; i += kSpThreads) ty[i] = __ldcg(sc.y + (size_t)j0 * 6 + i);
        __syncthreads();
        // S2: the chain's columns in sequence, all operands in shared memory
        for (int q = 0; q < len; ++q) {
          const int j = j0 + q;
          const int cb0 = __ldg(sc.col_ptr + j) - blk0, nb = __ldg(sc.col_ptr + j + 1) - blk0 - cb0;
          if (q > 0) {
            for (int bi = warp; bi < nb; bi += kSpWarps) {
              const int blk = blk0 + cb0 + bi;
              const int64_t u0 = sc.upd_ext_end[blk], u1 = sc.upd_ptr[blk + 1];
              if (u1 > u0) {
                double r0, r1;
                sp_gather_panel(sc, P, blk0, u0, u1, lane, r0, r1);
                if (fout) { const int e = 6 * g + 2 * t4; P[(cb0 + bi) * 36 + e] -= r0; P[(cb0 + bi) * 36 + e + 1] -= r1; }
              }
            }
            if (warp == kSpWarps - 1) {
              double acc = 0.0;
              if (lane < 30) {
                const int r1 = __ldg(sc.rowl_ptr + j + 1);
                for (int e = __ldg(sc.rowl_ext_end + j) + ps; e < r1; e += 5) {
                  const double* L = P + (size_t)(__ldg(sc.rowl_blk + e) - blk0) * 36 + 6 * ri;
                  const double* yk = ty + (__ldg(sc.rowl_col + e) - j0) * 6;
                  acc += (L[0] * yk[0] + L[1] * yk[1] + L[2] * yk[2]) + (L[3] * yk[3] + L[4] * yk[4] + L[5] * yk[5]);
                }
              }
              double tot = acc;
#pragma unroll
              for (int qq = 1; qq < 5; ++qq) tot += __shfl_sync(0xffffffffu, acc, (lane + 6 * qq) & 31);
              if (lane < 6) ty[q * 6 + lane] -= tot;
            }
            __syncthreads();
          }
          if (tid == 0) {
            double li[36];
            if (!chol6_inverse_factor_fast(P + cb0 * 36, li)) d.w_fail[d.fcam_win[__ldg(sc.perm + j)]] = 1;   // benign race: every writer stores 1
#pragma unroll
            for (int i = 0; i < 36; ++i) { li_s[i] = li[i]; __stcg(sc.Li + (size_t)j * 36 + i, li[i]); }
          }
          __syncthreads();
          for (int idx = tid; idx < (nb - 1) * 6; idx += kSpThreads) {      // panel: a thread per block row, in place
            double* row = P + (size_t)(cb0 + 1 + idx / 6) * 36 + 6 * (idx % 6);
            double av[6], o[6];
#pragma unroll
            for (int t = 0; t < 6; ++t) av[t] = row[t];
#pragma unroll
            for (int c = 0; c < 6; ++c) {
              double acc = 0.0;
#pragma unroll
              for (int t = 0; t < 6; ++t) if (t <= c) acc += av[t] * li_s[6 * c + t];
              o[c] = acc;
            }
#pragma unroll
            for (int t = 0; t < 6; ++t) row[t] = o[t];
          }
          if (warp == kSpWarps - 1) {
            const int c = lane - 26;                         // lanes 26..31: y_j = Linv t_j, in place behind a warp barrier
            double yk = 0.0;
            if (c >= 0) {
#pragma unroll
              for (int t = 0; t < 6; ++t) if (t <= c) yk += li_s[6 * c + t] * ty[q * 6 + t];
            }
            __syncwarp();
            if (c >= 0) ty[q * 6 + c] = yk;
          }
          __syncthreads();
        }
        // S3: the finished panel and y go to L2
        for (int i = tid; i < nblk * 36; i += kSpThreads) __stcg(sc.Lv + (size_t)blk0 * 36 + i, P[i]);
        for (int i = tid; i < len * 6; i += kSpThreads) __stcg(sc.y + (size_t)j0 * 6 + i, ty[i]);
        __syncthreads();
        continue;
      }
      // ======== long chains (dense patterns): column by column through L2
      for (int j = j0; j < jn; ++j) {
        const int b0 = __ldg(sc.col_ptr + j), nb = __ldg(sc.col_ptr + j + 1) - b0;
        const bool staged = nb <= kSpStage;
        // (a) gather: block = A_ij - sum_k L_ik L_jk^T.  Long columns: a warp per block.  Short ones: every block's
        //     list is cut into S contiguous slices, a warp per slice, slices summed in order afterwards.
        const int S = nb >= kSpWarps ? 1 : kSpWarps / nb;
        if (S == 1) {
          for (int bi = warp; bi < nb; bi += kSpWarps) {
            const int blk = b0 + bi;
            double r0, r1;
            sp_gather(sc, sc.upd_ptr[blk], sc.upd_ptr[blk + 1], lane, r0, r1);
            if (fout) {
              const int e = 6 * g + 2 * t4;
              const double2 a = __ldcg(reinterpret_cast<const double2*>(sc.Lv + (size_t)blk * 36 + e));
              const double2 v = make_double2(a.x - r0, a.y - r1);
              if (bi == 0) { akk[e] = v.x; akk[e + 1] = v.y; }
              else if (staged) { col_st[bi * 36 + e] = v.x; col_st[bi * 36 + e + 1] = v.y; }
              else __stcg(reinterpret_cast<double2*>(sc.Lv + (size_t)blk * 36 + e), v);
            }
          }
        } else {
          const int bi = warp / S, sl = warp - bi * S;
          if (bi < nb) {
            const int64_t u0 = sc.upd_ptr[b0 + bi], ulen = sc.upd_ptr[b0 + bi + 1] - u0;
            double r0, r1;
            sp_gather(sc, u0 + ulen * sl / S, u0 + ulen * (sl + 1) / S, lane, r0, r1);
            if (fout) { part[warp * 36 + 6 * g + 2 * t4] = r0; part[warp * 36 + 6 * g + 2 * t4 + 1] = r1; }
          }
        }
        // (b) right-hand side: t_j = rhs_j - sum_k L_jk y_k  (last warp; five entry groups of six rows, summed in lane order)
        if (warp == kSpWarps - 1) {
          double acc = 0.0;
          if (lane < 30) {
            const int r1 = __ldg(sc.rowl_ptr + j + 1);
            for (int e = __ldg(sc.rowl_ptr + j) + ps; e < r1; e += 5) {
              const double* L = sc.Lv + (size_t)__ldg(sc.rowl_blk + e) * 36 + 6 * ri;
              const double* yk = sc.y + (size_t)__ldg(sc.rowl_col + e) * 6;
              acc += (__ldcg(L) * __ldcg(yk) + __ldcg(L + 1) * __ldcg(yk + 1) + __ldcg(L + 2) * __ldcg(yk + 2)) +
                     (__ldcg(L + 3) * __ldcg(yk + 3) + __ldcg(L + 4) * __ldcg(yk + 4) + __ldcg(L + 5) * __ldcg(yk + 5));
            }
          }
          double tot = acc;
#pragma unroll
          for (int q = 1; q < 5; ++q) tot += __shfl_sync(0xffffffffu, acc, (lane + 6 * q) & 31);
          if (lane < 6) tj[lane] = d.rhs[(size_t)__ldg(sc.perm + j) * 6 + lane] - tot;
        }
        __syncthreads();
        if (S > 1) {
          for (int idx = tid; idx < nb * 36; idx += kSpThreads) {
            const int bi = idx / 36, e = idx - 36 * bi;
            double sum = 0.0;
            for (int sl = 0; sl < S; ++sl) sum += part[(bi * S + sl) * 36 + e];
            const double v = __ldcg(sc.Lv + (size_t)(b0 + bi) * 36 + e) - sum;
            if (bi == 0) akk[e] = v; else col_st[bi * 36 + e] = v;
          }
          __syncthreads();
        }
        // (c) diagonal block: 6x6 Cholesky across six lanes of warp 0 + the inverse of the factor
        if (tid == 0) {
          double li[36];
          if (!chol6_inverse_factor_fast(akk, li)) d.w_fail[d.fcam_win[__ldg(sc.perm + j)]] = 1;   // benign race: every writer stores 1
#pragma unroll
          for (int i = 0; i < 36; ++i) { li_s[i] = li[i]; __stcg(sc.Li + (size_t)j * 36 + i, li[i]); }
        }
        __syncthreads();
        // (d) panel L_ij = A_ij Linv^T (i > j) and y_j = Linv t_j
        if (staged) {
          for (int idx = tid; idx < (nb - 1) * 36; idx += kSpThreads) {
            const int bi = 1 + idx / 36, e = idx % 36, r = e / 6, c = e - 6 * r;
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < 6; ++q) if (q <= c) acc += col_st[bi * 36 + 6 * r + q] * li_s[6 * c + q];
            __stcg(sc.Lv + (size_t)(b0 + bi) * 36 + e, acc);
          }
        } else {
          for (int idx = tid; idx < (nb - 1) * 6; idx += kSpThreads) {
            double* row = sc.Lv + (size_t)(b0 + 1 + idx / 6) * 36 + 6 * (idx % 6);
            double a[6], o[6];
#pragma unroll
            for (int q = 0; q < 6; ++q) a[q] = __ldcg(row + q);
#pragma unroll
            for (int c = 0; c < 6; ++c) {
              double acc = 0.0;
#pragma unroll
              for (int q = 0; q < 6; ++q) if (q <= c) acc += a[q] * li_s[6 * c + q];
              o[c] = acc;
            }
#pragma unroll
            for (int q = 0; q < 6; ++q) __stcg(row + q, o[q]);
          }
        }
        if (tid >= kSpThreads - 6) {
          const int c = tid - (kSpThreads - 6);
          double acc = 0.0;
#pragma unroll
          for (int q = 0; q < 6; ++q) if (q <= c) acc += li_s[6 * c + q] * tj[q];
          __stcg(sc.y + (size_t)j * 6 + c, acc);
        }
        __syncthreads();                          // the column is in L2 for the chain's next one; shared buffers are reused
      }
    }
    grid.sync();
    lap();
  }
  // ---- 2: back substitution, levels downwards; a chain's columns in sequence (last first) by its CTA:
  //         x_j = Linv_jj^T (y_j - sum_i L_ij^T x_i), a thread per (block, column), partials summed in block order
  double* xp = col_st;                            // [kSpThreads] partials
  constexpr int G = kSpThreads / 6;               // block groups: thread (gq, c) takes blocks gq, gq + G, ...
  const int gq = tid / 6, cc = tid - 6 * gq;
  for (int lvl = sc.nlev - 1; lvl >= 0; --lvl) {
    const int c_begin = __ldg(sc.lvl_ptr + lvl), c_end = __ldg(sc.lvl_ptr + lvl + 1);
    for (int ci = c_begin + blockIdx.x; ci < c_end; ci += gridDim.x) {
      const int j0 = __ldg(sc.lvl_cols + ci), len = __ldg(sc.lvl_len + ci), jn = j0 + len;
      const int blk0 = __ldg(sc.col_ptr + j0), nblk = __ldg(sc.col_ptr + jn) - blk0;
      if (nblk <= cap && len <= kSpChainMax) {
        // ======== shared-memory chain.  The whole chain comes in by one round of independent loads: panel, block rows,
        // x of the rows outside it, y and the inverse factors of its columns ...
        for (int i = tid; i < nblk * 36; i += kSpThreads) P[i] = __ldcg(sc.Lv + (size_t)blk0 * 36 + i);
        for (int i = tid; i < nblk * 6; i += kSpThreads) {
          const int b = i / 6, row = __ldg(sc.row_idx + blk0 + b);
          if (i == 6 * b) rowi[b] = row;
          xe[i] = (row >= jn) ? __ldcg(sc.y + (size_t)row * 6 + (i - 6 * b)) : 0.0;
        }
        for (int i = tid; i < len * 6; i += kSpThreads) ty[i] = __ldcg(sc.y + (size_t)j0 * 6 + i);
        for (int i = tid; i < len * 36; i += kSpThreads) LIc[i] = __ldcg(sc.Li + (size_t)j0 * 36 + i);
        __syncthreads();
        // ... then what the rows OUTSIDE the chain contribute, for all its columns at once: a thread per (block, component)
        // forms L_b^T x_b in place of x_b (values held in registers across the barrier), a thread per (column, component)
        // adds its column's blocks in block order ...
        {
          constexpr int kMaxIt = (kSpPanelMax * 6 + kSpThreads - 1) / kSpThreads;
          double v[kMaxIt];
#pragma unroll
          for (int k = 0; k < kMaxIt; ++k) {
            const int i = tid + k * kSpThreads;
            v[k] = 0.0;
            if (i < nblk * 6) {
              const int bb = i / 6, c = i - 6 * bb;
              if (rowi[bb] >= jn) {
                const double* L = P + (size_t)bb * 36 + c;
                const double* xi = xe + bb * 6;
                v[k] = (L[0] * xi[0] + L[6] * xi[1] + L[12] * xi[2]) + (L[18] * xi[3] + L[24] * xi[4] + L[30] * xi[5]);
              }
            }
          }
          __syncthreads();
#pragma unroll
          for (int k = 0; k < kMaxIt; ++k) { const int i = tid + k * kSpThreads; if (i < nblk * 6) xe[i] = v[k]; }
          __syncthreads();
          if (tid < len * 6) {
            const int q = tid / 6, c = tid - 6 * q;
            const int cb0 = __ldg(sc.col_ptr + j0 + q) - blk0, cb1 = __ldg(sc.col_ptr + j0 + q + 1) - blk0;
            double e = 0.0;
            for (int bb = cb0 + 1; bb < cb1; ++bb) e += xe[bb * 6 + c];      // in-chain rows hold 0
            ty[tid] -= e;
          }
          __syncthreads();
        }
        // ... and the chain's own columns in sequence (last first) by warp 0 alone: only the in-chain blocks are left
        if (warp == 0) {
          for (int j = jn - 1; j >= j0; --j) {
            const int q = j - j0;
            const int cb0 = __ldg(sc.col_ptr + j) - blk0, cb1 = __ldg(sc.col_ptr + j + 1) - blk0;
            double acc = 0.0;
            if (lane < 30) {
              for (int bb = cb0 + 1 + ps; bb < cb1; bb += 5) {
                const int row = rowi[bb];
                if (row >= jn) break;                       // rows ascend: the rest of the column lies outside the chain
                const double* L = P + (size_t)bb * 36 + ri;
                const double* xi = ty + (row - j0) * 6;
                acc += (L[0] * xi[0] + L[6] * xi[1] + L[12] * xi[2]) + (L[18] * xi[3] + L[24] * xi[4] + L[30] * xi[5]);
              }
            }
            double tot = acc;
#pragma unroll
            for (int qq = 1; qq < 5; ++qq) tot += __shfl_sync(0xffffffffu, acc, (lane + 6 * qq) & 31);
            const double sj = (lane < 6) ? ty[q * 6 + lane] - tot : 0.0;
            double vv = 0.0;
#pragma unroll
            for (int t = 0; t < 6; ++t) {
              const double st = __shfl_sync(0xffffffffu, sj, t);
              if (lane < 6 && t >= lane) vv += LIc[q * 36 + 6 * t + lane] * st;
            }
            __syncwarp();
            if (lane < 6) {
              ty[q * 6 + lane] = vv;
              __stcg(sc.y + (size_t)j * 6 + lane, vv);
              const int f = __ldg(sc.perm + j);
              const int w = d.fcam_win[f];
              const bool ok = isfinite(vv) && __ldcg(d.w_fail + w) == 0;
              if (!isfinite(vv)) d.w_fail[w] = 1;                                   // benign race: every writer stores 1
              d.x[(size_t)f * 6 + lane] = ok ? vv : 0.0;
              d.r[(size_t)f * 6 + lane] = 0.0;                                      // exact solve: no linear residual in the model change
            }
            __syncwarp();
          }
        }
        __syncthreads();
        continue;
      }
      // ======== long chains: column by column through L2, a thread per (block, column), partials summed in block order
      for (int j = jn - 1; j >= j0; --j) {
        const int b0 = __ldg(sc.col_ptr + j), nbl = __ldg(sc.col_ptr + j + 1) - b0 - 1;
        double acc = 0.0;
        double yv = 0.0, lic[6] = {0, 0, 0, 0, 0, 0};
        if (gq < G)
          for (int bb = gq; bb < nbl; bb += G) {
            const double* L = sc.Lv + (size_t)(b0 + 1 + bb) * 36 + cc;               // column cc of L_ij
            const double* xi = sc.y + (size_t)__ldg(sc.row_idx + b0 + 1 + bb) * 6;
            acc += (__ldcg(L) * __ldcg(xi) + __ldcg(L + 6) * __ldcg(xi + 1) + __ldcg(L + 12) * __ldcg(xi + 2)) +
                   (__ldcg(L + 18) * __ldcg(xi + 3) + __ldcg(L + 24) * __ldcg(xi + 4) + __ldcg(L + 30) * __ldcg(xi + 5));
          }
        if (tid < 6) {
          yv = __ldcg(sc.y + (size_t)j * 6 + tid);
#pragma unroll
          for (int q = 0; q < 6; ++q) if (q >= tid) lic[q] = __ldcg(sc.Li + (size_t)j * 36 + 6 * q + tid);     // Linv^T row
        }
        xp[tid] = acc;
        __syncthreads();
        if (warp == 0) {
          double s = 0.0;
          if (lane < 6) {
            const int ng = min(nbl, G);
            for (int q = 0; q < ng; ++q) s += xp[6 * q + lane];
            s = yv - s;
          }
          double v = 0.0;
#pragma unroll
          for (int q = 0; q < 6; ++q) {
            const double sq = __shfl_sync(0xffffffffu, s, q);
            if (q >= lane) v += lic[q] * sq;
          }
          if (lane < 6) {
            __stcg(sc.y + (size_t)j * 6 + lane, v);
            const int f = __ldg(sc.perm + j);
            const int w = d.fcam_win[f];
            const bool ok = isfinite(v) && __ldcg(d.w_fail + w) == 0;
            if (!isfinite(v)) d.w_fail[w] = 1;                                    // benign race: every writer stores 1
            d.x[(size_t)f * 6 + lane] = ok ? v : 0.0;
            d.r[(size_t)f * 6 + lane] = 0.0;                                      // exact solve: no linear residual in the model change
          }
        }
        __syncthreads();
      }
    }
    if (lvl > 0) grid.sync();
    lap();
  }
  if (prof) {
    printf("[spchol] us per level, fwd:");
    for (int i = 1; i <= sc.nlev && i < nstamp; ++i) printf(" %.1f", (double)(tstamp[i] - tstamp[i - 1]) * 1e-3);
    printf(" | bwd (top first):");
    for (int i = sc.nlev + 1; i < nstamp; ++i) printf(" %.1f", (double)(tstamp[i] - tstamp[i - 1]) * 1e-3);
    printf(" | total %.1f\n", (double)(tstamp[nstamp - 1] - tstamp[0]) * 1e-3);
  }
}

}  // namespace sba
