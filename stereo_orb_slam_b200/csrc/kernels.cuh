// kernels.cuh — hand-written sm_100a kernels of the stereo BA hot path (fp64, HBM-bound).
//
// Stage map:  K0 k_campre (camera table) · K1/K2 k_lin_chunk (residual, analytic Jacobian algebra,
// Huber, U/V/W/g accumulation per chunk; Jacobians never reach HBM, W is written once) + k_assemble_cams /
// k_lin_finish · K3 k_schur_chunk + k_assemble_schur · K1-K3 in one kernel: k_fused_chunk · K4 k_chol_band /
// k_chol_window / k_chol_cluster (exact, windows of <= 64 free cameras), k_spchol (sparse_chol.cuh, above that),
// k_pcg_* · K5 + K6 + trust-region control: k_step_loop (single GPU) or k_backsub / k_trial_* / k_ls_* / k_lm_*.
// No global data-path atomics anywhere: every reduction is a fixed-order tree over statically
// assigned chunks, so results are bit-reproducible run to run.
#pragma once
#include <cuda_runtime.h>
#include <cooperative_groups.h>
#include "sba_device.cuh"
#include "sba_linesearch.cuh"
#include "sba_internal.h"

namespace sba {

// Sum N per-thread values over a 128-thread CTA.  Thread t < N returns the total of element t
// (fixed order: xor-butterfly inside each warp, then warps 0..3 in order).
template <int N>
__device__ __forceinline__ double block_sum(double (&v)[N], double* sm /* [4*N] */) {
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double x = v[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    v[i] = x;
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();   // protect sm against a previous use
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < N; ++i) sm[w * N + i] = v[i];
  }
  __syncthreads();
  double out = 0.0;
  if (threadIdx.x < N) out = ((sm[threadIdx.x] + sm[N + threadIdx.x]) + sm[2 * N + threadIdx.x]) + sm[3 * N + threadIdx.x];
  return out;
}

__device__ __forceinline__ double block_max1(double v, double* sm /* [4] */) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) sm[w] = v;
  __syncthreads();
  return fmax(fmax(sm[0], sm[1]), fmax(sm[2], sm[3]));
}

// (.cg loads: the trial camera table is rewritten every line-search round inside one cooperative kernel)
__device__ __forceinline__ void load_campre(const CamPre* base, int c, CamPre& out) {
  const double* s = reinterpret_cast<const double*>(base + c);
#pragma unroll
  for (int i = 0; i < 9; ++i) out.R[i] = __ldcg(s + i);
#pragma unroll
  for (int i = 0; i < 9; ++i) out.Jl[i] = __ldcg(s + 9 + i);
#pragma unroll
  for (int i = 0; i < 3; ++i) out.t[i] = __ldcg(s + 18 + i);
}
// R and t only (cost evaluation)
__device__ __forceinline__ void load_campre_Rt(const CamPre* base, int c, CamPre& out) {
  const double* s = reinterpret_cast<const double*>(base + c);
#pragma unroll
  for (int i = 0; i < 9; ++i) out.R[i] = __ldcg(s + i);
#pragma unroll
  for (int i = 0; i < 3; ++i) out.t[i] = __ldcg(s + 18 + i);
}

// clamp points into the box (Ceres projects x onto the bounds at iteration 0; reference bounds:
// src/bundle_adjuster.cpp:104-108)
__global__ void k_project_points(int64_t n, double* __restrict__ pts, double lo, double hi) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pts[i] = clampd(pts[i], lo, hi);
}

// ------------------------------------------------------------------------------- K1/K2 per chunk
// One CTA per chunk (<= kObsChunk = 256 observations of <= kPointChunk consecutive points), one THREAD PER
// OBSERVATION.  Everything an observation needs is staged in shared memory first with independent, coalesced
// loads — the chunk's points, and R / J_l / t of the chunk's few cameras ("slots") from the per-linearisation camera
// table (K0 = k_campre) — so the only per-thread global reads are idx (1 + 4 B) and uv (16 B).  W (144 B) goes straight
// to HBM; the point-side sums (V, gp: 9 doubles) and the camera-side operands (A, h, G: 18 doubles) are reduced
// inside the CTA — per point by its thread, per camera slot by one warp.  The per-(chunk, slot) camera partials go
// to `cspart`, one 27-double row per slot at its position in the owning camera's list; k_assemble_cams sums each
// camera's (contiguous) rows in list order afterwards: fixed order, no data atomics.
constexpr int kLinThreads = 256;
constexpr int kCamStride = 27;     // doubles per observation in the camera staging area: its U (21) and gc (6) terms

__host__ __device__ __forceinline__ size_t lin_smem_bytes(int n, int np, int ns) {
  return ((size_t)n * (kCamStride + 9) + (size_t)ns * 21 + (size_t)np * 3) * 8;
}

__device__ __forceinline__ void point_contrib(const CamPre& c, const double A[6], const double h[3], const double G[9],
                                              double Vc[9], double W[18]) {
  double Af[9]; sym3_full(A, Af);
  double AR[9]; mul33(Af, c.R, AR);
  Vc[0] = c.R[0] * AR[0] + c.R[3] * AR[3] + c.R[6] * AR[6];
  Vc[1] = c.R[0] * AR[1] + c.R[3] * AR[4] + c.R[6] * AR[7];
  Vc[2] = c.R[0] * AR[2] + c.R[3] * AR[5] + c.R[6] * AR[8];
  Vc[3] = c.R[1] * AR[1] + c.R[4] * AR[4] + c.R[7] * AR[7];
  Vc[4] = c.R[1] * AR[2] + c.R[4] * AR[5] + c.R[7] * AR[8];
  Vc[5] = c.R[2] * AR[2] + c.R[5] * AR[5] + c.R[8] * AR[8];
#pragma unroll
  for (int i = 0; i < 3; ++i) Vc[6 + i] = c.R[i] * h[0] + c.R[3 + i] * h[1] + c.R[6 + i] * h[2];
  mulT33(G, AR, W);
#pragma unroll
  for (int i = 0; i < 9; ++i) W[9 + i] = AR[i];
}

__global__ void __launch_bounds__(kLinThreads, 3) k_lin_chunk(Dev d, DevParams P, int dbg) {
  extern __shared__ double smem[];
  __shared__ double sm1[kLinThreads / 32];
  __shared__ double smx[kLinThreads / 32];
  const int chunk = blockIdx.x;
  const int4 c0 = __ldg(d.cdesc + 4 * (size_t)chunk), c1 = __ldg(d.cdesc + 4 * (size_t)chunk + 1);   // one record, one round trip
  const int p0 = c0.x, np = c0.y, e0 = c0.z, n = c0.w;
  if (n > kMaxSlots) return;                               // a long track: k_lin_long's
  const int cs0 = c1.y, ns = c1.z, cs1 = cs0 + ns;
  double* cam_st = smem;                                   // [n * kCamStride]  A(6) h(3) G(9)
  double* pt_st = cam_st + (size_t)n * kCamStride;         // [n * 9]           V(6) gp(3)
  double* cpre = pt_st + (size_t)n * 9;                    // [ns * 21]         R Jl t of the chunk's cameras
  double* xs = cpre + (size_t)ns * 21;                     // [np * 3]          the chunk's points
  // ---- stage: independent loads only
  const int i = threadIdx.x;
  const bool act = i < n;
  int sl = 0, pl = 0;
  float4 m = make_float4(0.f, 0.f, 0.f, 0.f);
  if (act) {
    sl = d.oslot[e0 + i];
    pl = __ldg(d.ppt + e0 + i) - p0;
    m = __ldg(d.uv_p + e0 + i);
  }
  for (int k = threadIdx.x; k < np * 3; k += kLinThreads) xs[k] = d.pts[(size_t)p0 * 3 + k];
  // the slot cameras' R / J_l / t come from the per-linearisation table (k_campre): one instance of the sin / cos code
  // serves every linearisation kernel, and no CTA waits behind a handful of threads doing trigonometry
  for (int k = threadIdx.x; k < ns * 21; k += kLinThreads) {
    const int s = k / 21;
    cpre[k] = reinterpret_cast<const double*>(d.campre + __ldg(d.cs_cam + cs0 + s))[k - 21 * s];
  }
  __syncthreads();
  double cost = 0.0;
  if (act) {
    const int e = e0 + i;
    const CamPre& cp = *reinterpret_cast<const CamPre*>(cpre + 21 * sl);
    const float uv[4] = {m.x, m.y, m.z, m.w};
    const double X[3] = {xs[3 * pl], xs[3 * pl + 1], xs[3 * pl + 2]};
    double q[3], r[4], D[12];
    const double s = project(P.pl, P.pr, cp, X, uv, P.rdim, q, r, D);
    double rho, w;
    huber(P.huber_delta, s, rho, w);
    cost = 0.5 * rho;
    double A[6], h[3], G[9];
    form_A_h(D, r, w, A, h);
    form_G(cp, q, G);
    {
      double cu[27];
#pragma unroll
      for (int k = 0; k < 27; ++k) cu[k] = 0.0;
      accum_camera(A, h, G, cu, cu + 21);
      double* cs = cam_st + (size_t)i * kCamStride;
#pragma unroll
      for (int k = 0; k < 27; ++k) cs[k] = cu[k];
    }
    double Vc[9], W[18];
    point_contrib(cp, A, h, G, Vc, W);
#pragma unroll
    for (int k = 0; k < 9; ++k) pt_st[(size_t)i * 9 + k] = Vc[k];
    double2* Wo = reinterpret_cast<double2*>(d.W + (size_t)e * 18);
#pragma unroll
    for (int k = 0; k < 9; ++k) Wo[k] = make_double2(W[2 * k], W[2 * k + 1]);
    if (d.r_dbg) {
      const double sq = sqrt(w);
      double* ro = d.r_dbg + (size_t)d.perm_p[e] * 4;
      ro[0] = sq * r[0]; ro[1] = sq * r[1]; ro[2] = sq * r[2]; ro[3] = sq * r[3];
    }
  }
  __syncthreads();
  // ---- per point: V, gp (thread per point, observations of a point are consecutive)
  double gmax = 0.0;
  for (int q = threadIdx.x; q < np && !(dbg & 4); q += kLinThreads) {
    const int p = p0 + q;
    double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    const int i1 = d.pt_off[p + 1] - e0;
    for (int o = d.pt_off[p] - e0; o < i1; ++o) {
#pragma unroll
      for (int k = 0; k < 9; ++k) acc[k] += pt_st[(size_t)o * 9 + k];
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) d.V[(size_t)p * 6 + k] = acc[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      d.gp[(size_t)p * 3 + k] = acc[6 + k];
      const double x = xs[3 * q + k];
      gmax = fmax(gmax, fabs(x - clampd(x - acc[6 + k], P.lo, P.hi)));   // projected-gradient norm
    }
  }
  // ---- per camera slot: U (21) + gc (6), one warp per slot over the slot's observations.  The point sums above
  // keep the first warps busy, so the slots are dealt from the last warp down.
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = kLinThreads / 32;
  for (int cs = cs0 + (NW - 1 - warp); cs < cs1 && !(dbg & 1); cs += NW) {
    const int c = d.cs_cam[cs];
    if (d.fc[c] < 0) continue;        // a fixed camera is not an unknown (src/bundle_adjuster.cpp:113)
    // lane e sums element e of the slot's observations in list order, even and odd list positions apart (two
    // independent add chains, joined at the end: a fixed order); the observation indices come 32 at a time by one
    // coalesced load and are handed round by shuffles, so nothing in the loop waits on global memory
    double s0 = 0.0, s1 = 0.0;
    const int j0 = d.cs_obs_off[cs], j1 = d.cs_obs_off[cs + 1];
    const double* col = cam_st + (lane < 27 ? lane : 0);
    for (int jb = j0; jb < j1; jb += 32) {
      const int mine_o = (jb + lane < j1) ? (int)__ldg(d.sobs + jb + lane) : 0;
      const int m = min(32, j1 - jb);
      int t = 0;
      for (; t + 3 < m; t += 4) {
        const int o0 = __shfl_sync(0xffffffffu, mine_o, t), o1 = __shfl_sync(0xffffffffu, mine_o, t + 1);
        const int o2 = __shfl_sync(0xffffffffu, mine_o, t + 2), o3 = __shfl_sync(0xffffffffu, mine_o, t + 3);
        const double v0 = col[o0 * kCamStride], v1 = col[o1 * kCamStride], v2 = col[o2 * kCamStride], v3 = col[o3 * kCamStride];
        s0 += v0; s1 += v1; s0 += v2; s1 += v3;
      }
      for (; t < m; ++t) {
        const int o = __shfl_sync(0xffffffffu, mine_o, t);
        const double v = col[o * kCamStride];
        if (t & 1) s1 += v; else s0 += v;
      }
    }
    if (lane < 27) d.cspart[(size_t)d.cs_pos[cs] * 27 + lane] = s0 + s1;
  }
  // ---- chunk cost and projected-gradient max
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { cost += __shfl_xor_sync(0xffffffffu, cost, o); gmax = fmax(gmax, __shfl_xor_sync(0xffffffffu, gmax, o)); }
  if (lane == 0) { sm1[warp] = cost; smx[warp] = gmax; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0, mx = 0.0;
#pragma unroll
    for (int k = 0; k < kLinThreads / 32; ++k) { t += sm1[k]; mx = fmax(mx, smx[k]); }
    d.lin_part[2 * chunk] = t; d.lin_part[2 * chunk + 1] = mx;
  }
}

// Jacobi column scaling, fixed at iteration 0 (Ceres jacobi_scaling): s = 1/(1+sqrt(H_jj)), stored squared
__global__ void k_jacobi_scale(Dev d) {
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t nc = (int64_t)d.Nc * 6, np = (int64_t)d.Np * 3;
  if (t < nc) {
    const int c = (int)(t / 6), i = (int)(t - 6 * (int64_t)c);
    const double s = 1.0 / (1.0 + sqrt(d.U[(size_t)c * 21 + up6(i, i)]));
    d.sc2c[t] = s * s;
  } else if (t < nc + np) {
    const int64_t u = t - nc;
    const int64_t p = u / 3; const int i = (int)(u - 3 * p);
    const int di = (i == 0) ? 0 : (i == 1 ? 3 : 5);
    const double s = 1.0 / (1.0 + sqrt(d.V[(size_t)p * 6 + di]));
    d.sc2p[u] = s * s;
  }
}

// per window: cost = sum of chunk costs, gmax = max(chunk gmax, |gc| of free cameras);
// gradient-tolerance test (Ceres: after every accepted step and at iteration 0).
__device__ __forceinline__ void dev_lin_finish(const Dev& d, const DevParams& P, int w, int keep_cost) {
  __shared__ double sm[4];
  __shared__ double smx[4];
  double cost[1] = {0.0};
  double gmax = 0.0;
  for (int ch = d.win_pchunk_off[w] + threadIdx.x; ch < d.win_pchunk_off[w + 1]; ch += kThreads) {
    cost[0] += d.lin_part[2 * ch];
    gmax = fmax(gmax, d.lin_part[2 * ch + 1]);
  }
  for (int f = d.win_fc_off[w] + threadIdx.x; f < d.win_fc_off[w + 1]; f += kThreads) {
    const int c = d.free2cam[f];
#pragma unroll
    for (int i = 0; i < 6; ++i) gmax = fmax(gmax, fabs(__ldcg(d.gc + (size_t)c * 6 + i)));      // (may come from other CTAs of this launch)
  }
  const double tot = block_sum<1>(cost, sm);
  const double mx = block_max1(gmax, smx);
  if (threadIdx.x == 0) {
    // After an accepted step the window's cost stays the accepted candidate's (Ceres: x_cost = candidate_cost, and
    // its re-evaluation returns the same bits).  The line search compares trial costs against it at step sizes
    // where the decrease is a few ulps, so both must come from the same summation (k_trial_points), not this one.
    if (!keep_cost) d.w_cost[w] = tot;
    d.w_gmax[w] = mx;
  }
}

__global__ void __launch_bounds__(kThreads) k_lin_finish(Dev d, DevParams P, int keep_cost) { dev_lin_finish(d, P, blockIdx.x, keep_cost); }

// Point-sharded solve (multi-GPU): the same per-window finish in two halves around ONE packed all-reduce (sum) of
// [U | gc | cost_red | gslots].  A maximum travels through the sum: every rank writes its value into its own slot of
// an otherwise zero array, so the reduced array holds all ranks' values and the maximum is taken locally — on the
// same bits on every rank.
__global__ void __launch_bounds__(kThreads) k_lin_finish_local(Dev d, int rank, int world) {
  __shared__ double sm[4];
  __shared__ double smx[4];
  const int w = blockIdx.x;
  double cost[1] = {0.0};
  double gmax = 0.0;
  for (int ch = d.win_pchunk_off[w] + threadIdx.x; ch < d.win_pchunk_off[w + 1]; ch += kThreads) {
    cost[0] += d.lin_part[2 * ch];
    gmax = fmax(gmax, d.lin_part[2 * ch + 1]);
  }
  const double tot = block_sum<1>(cost, sm);
  const double mx = block_max1(gmax, smx);
  if (threadIdx.x == 0) d.cost_red[w] = tot;
  for (int r = threadIdx.x; r < world; r += kThreads) d.gslots[(size_t)r * d.nwin + w] = (r == rank) ? mx : 0.0;
}

__global__ void __launch_bounds__(kThreads) k_lin_finish_global(Dev d, DevParams P, int keep_cost, int world) {
  __shared__ double smx[4];
  const int w = blockIdx.x;
  double gmax = 0.0;
  for (int r = threadIdx.x; r < world; r += kThreads) gmax = fmax(gmax, d.gslots[(size_t)r * d.nwin + w]);
  for (int f = d.win_fc_off[w] + threadIdx.x; f < d.win_fc_off[w + 1]; f += kThreads) {
    const int c = d.free2cam[f];
#pragma unroll
    for (int i = 0; i < 6; ++i) gmax = fmax(gmax, fabs(d.gc[(size_t)c * 6 + i]));
  }
  const double mx = block_max1(gmax, smx);
  if (threadIdx.x == 0) {
    if (!keep_cost) d.w_cost[w] = d.cost_red[w];
    d.w_gmax[w] = mx;
  }
}

// max-by-sum exchange of a per-window scalar: before the all-reduce every rank fills its own slot, afterwards the
// maximum over the slots is the global one
__global__ void k_slots_put(double* __restrict__ slots, const double* __restrict__ v, int nwin, int rank, int world) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nwin * world) return;
  const int r = i / nwin, w = i - r * nwin;
  slots[i] = (r == rank) ? v[w] : 0.0;
}
__global__ void k_slots_max(double* __restrict__ v, const double* __restrict__ slots, int nwin, int world) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= nwin) return;
  double m = 0.0;
  for (int r = 0; r < world; ++r) m = fmax(m, slots[(size_t)r * nwin + w]);
  v[w] = m;
}

// --------------------------------------------------------------------------- K3: Schur complement
// LM damping in Ceres's Jacobi-scaled space mapped back to the unscaled one:
//   D_j = clamp(s_j^2 H_jj, 1e-6, 1e32) / (s_j^2 radius)        (levenberg_marquardt_strategy)
__device__ __forceinline__ double lm_damp(double sc2, double hjj, double radius) {
  return clampd(sc2 * hjj, 1e-6, 1e32) / (sc2 * radius);
}

__device__ __forceinline__ void load_W(const double* __restrict__ W, size_t pos, double out[18]) {
  const double2* s = reinterpret_cast<const double2*>(W + pos * 18);
#pragma unroll
  for (int i = 0; i < 9; ++i) { const double2 v = __ldg(s + i); out[2 * i] = v.x; out[2 * i + 1] = v.y; }
}

// One CTA per chunk, the whole of K3 for the chunk's points in one pass over W:
//   1. thread per point: Cholesky of V + D (3x3): Linv, Vinv = Linv^T Linv, tp = Vinv gp (Vinv, tp -> HBM for K5),
//      u = Linv gp
//   2. the chunk's W (contiguous in point order) staged ONCE into shared memory with coalesced loads, then each
//      observation's thread turns its row into Z = W Linv^T in place, so that  W_a Vinv W_b^T = Z_a Z_b^T
//   3. one warp per local block (slot pair sa, sb) walks the block's static pair list (a, b): one FP64 tensor-core
//      MMA (m8n8k4) per pair, D += Z_a Z_b^T with rows / columns 6-7 and k = 3 as zero padding; the accumulator
//      fragment IS the 6x6 partial (no cross-lane reduction) and goes straight to `spart`;  one warp per free slot:
//      sum of Z_e u[point(e)] = W_e tp -> rpart
//   4. k_assemble_schur sums every block's / camera's partial rows (contiguous by construction) in list order.
// ---- TMA 1-D bulk copy global -> shared (cp.async.bulk, completion on an mbarrier): one elected thread moves a
// whole tile without spending load/store instructions or registers, and the CTA computes while it is in flight.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src_global, uint32_t bytes, uint64_t* bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
               "l"(src_global), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// D = A B + D on the FP64 tensor cores: A 8x4 (lane holds A[lane/4][lane%4]), B 4x8 (lane holds B[lane%4][lane/4]),
// D 8x8 (lane holds D[lane/4][2 (lane%4)], D[lane/4][2 (lane%4) + 1])
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(done)
                 : "r"(smem_u32(bar)), "r"(parity)
                 : "memory");
  } while (!done);
}

constexpr int kSchurThreads = 256;
constexpr int kZStride = 18;       // doubles per observation in the Z staging area (rows stay 16-byte aligned: LDS.128)
constexpr int kSchurPairCap = 8192;  // block products of one chunk staged in shared memory (16 KB); larger lists are read through L1

__host__ __device__ __forceinline__ size_t schur_smem_bytes(int n, int np, int npairs) {
  return ((size_t)n * kZStride + (size_t)np * 9) * 8 + (size_t)((n + 15) & ~15) + (size_t)(npairs <= kSchurPairCap ? npairs : 0) * 2;
}

__global__ void __launch_bounds__(kSchurThreads, 4) k_schur_chunk(Dev d, int max_pairs, int dbg) {
  extern __shared__ __align__(16) double smem[];
  const int chunk = blockIdx.x;
  const int4 c0 = __ldg(d.cdesc + 4 * (size_t)chunk), c1 = __ldg(d.cdesc + 4 * (size_t)chunk + 1), c2 = __ldg(d.cdesc + 4 * (size_t)chunk + 2);
  const int p0 = c0.x, np = c0.y, e0 = c0.z, n = c0.w;
  if (n > kMaxSlots) return;                               // a long track: k_schur_long's
  const int w = c1.x, cs0 = c1.y, cs1 = c1.y + c1.z, lb0 = c1.w, lb1 = c1.w + c2.x;
  const int pq0 = c2.y, npq = c2.z;
  double* Z = smem;                                     // [n * kZStride]
  double* Li = Z + (size_t)n * kZStride;                // [np * 6]  Linv: 00 10 11 20 21 22
  double* us = Li + (size_t)np * 6;                     // [np * 3]  u = Linv gp
  uint8_t* opl = reinterpret_cast<uint8_t*>(us + (size_t)np * 3);   // [n] chunk-local point of each observation
  uint16_t* prs = reinterpret_cast<uint16_t*>(opl + ((n + 15) & ~15));   // the chunk's pair lists, when they fit
  const bool pairs_staged = npq <= kSchurPairCap && max_pairs <= kSchurPairCap;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = kSchurThreads / 32;
  // one point per thread (np <= kPointChunk <= kSchurThreads): its V, scaling and gradient are requested first ...
  static_assert(kPointChunk <= kSchurThreads, "one point per thread");
  const int q = threadIdx.x;
  const bool has_pt = q < np && !(dbg & 16);
  const int p = p0 + q;
  double V[6] = {0, 0, 0, 0, 0, 0}, s2[3] = {1, 1, 1}, gv[3] = {0, 0, 0};
  if (has_pt) {
#pragma unroll
    for (int i = 0; i < 6; ++i) V[i] = d.V[(size_t)p * 6 + i];
#pragma unroll
    for (int i = 0; i < 3; ++i) { s2[i] = d.sc2p[(size_t)p * 3 + i]; gv[i] = d.gp[(size_t)p * 3 + i]; }
  }
  const double radius = d.w_radius[w];
  // ... then the chunk's W tile (n x 144 B, contiguous in point order, 16-byte aligned) straight into Z by one bulk
  // copy: 34 KB per CTA, queued behind the small latency-critical loads above instead of in front of them.  The
  // other threads meet the barrier only after the __syncthreads below.
  __shared__ __align__(8) uint64_t wbar;
  if (threadIdx.x == 0) {
    mbar_init(&wbar, 1);
    if (n > 0 && !(dbg & 32)) bulk_load(Z, d.W + (size_t)e0 * 18, (uint32_t)n * 144u, &wbar);
  }
  if (has_pt) {
    const double a = V[0] + lm_damp(s2[0], V[0], radius);
    const double dd = V[3] + lm_damp(s2[1], V[3], radius);
    const double f = V[5] + lm_damp(s2[2], V[5], radius);
    const double b = V[1], c = V[2], e = V[4];
    // lower Cholesky factor and its inverse
    double i00 = 0, i10 = 0, i11 = 0, i20 = 0, i21 = 0, i22 = 0;
    bool ok = a > 0.0;
    if (ok) {
      const double l00 = sqrt(a);
      i00 = 1.0 / l00;
      const double l10 = b * i00, l20 = c * i00;
      const double r1 = dd - l10 * l10;
      ok = r1 > 0.0;
      if (ok) {
        const double l11 = sqrt(r1);
        i11 = 1.0 / l11;
        const double l21 = (e - l20 * l10) * i11;
        const double r2 = f - l20 * l20 - l21 * l21;
        ok = r2 > 0.0 && isfinite(r2);
        if (ok) {
          i22 = 1.0 / sqrt(r2);
          i10 = -l10 * i00 * i11;
          i21 = -l21 * i11 * i22;
          i20 = -(l20 * i00 + l21 * i10) * i22;
        }
      }
    }
    if (!ok) { d.w_fail[w] = 1; i00 = i10 = i11 = i20 = i21 = i22 = 0.0; }   // benign race: every writer stores 1
    const double g0 = gv[0], g1 = gv[1], g2 = gv[2];
    const double u0 = i00 * g0, u1 = i10 * g0 + i11 * g1, u2 = i20 * g0 + i21 * g1 + i22 * g2;
    double* vo = d.Vinv + (size_t)p * 6;
    vo[0] = i00 * i00 + i10 * i10 + i20 * i20; vo[1] = i10 * i11 + i20 * i21; vo[2] = i20 * i22;
    vo[3] = i11 * i11 + i21 * i21; vo[4] = i21 * i22; vo[5] = i22 * i22;
    d.tp[(size_t)p * 3] = i00 * u0 + i10 * u1 + i20 * u2;
    d.tp[(size_t)p * 3 + 1] = i11 * u1 + i21 * u2;
    d.tp[(size_t)p * 3 + 2] = i22 * u2;
    Li[q * 6] = i00; Li[q * 6 + 1] = i10; Li[q * 6 + 2] = i11; Li[q * 6 + 3] = i20; Li[q * 6 + 4] = i21; Li[q * 6 + 5] = i22;
    us[q * 3] = u0; us[q * 3 + 1] = u1; us[q * 3 + 2] = u2;
  }
  if (pairs_staged)
    for (int i = threadIdx.x; i < npq; i += kSchurThreads) prs[i] = __ldg(d.pairs + pq0 + i);
  for (int i = threadIdx.x; i < n; i += kSchurThreads) opl[i] = (uint8_t)(__ldg(d.ppt + e0 + i) - p0);
  __syncthreads();
  if (n > 0 && !(dbg & 32)) mbar_wait(&wbar, 0);
  // ---- Z = W Linv^T, each observation's thread on its own row
  for (int i = threadIdx.x; i < n && !(dbg & 64); i += kSchurThreads) {
    const double* L = Li + (int)opl[i] * 6;
    const double i00 = L[0], i10 = L[1], i11 = L[2], i20 = L[3], i21 = L[4], i22 = L[5];
    double* z = Z + (size_t)i * kZStride;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const double w0 = z[3 * r], w1 = z[3 * r + 1], w2 = z[3 * r + 2];
      z[3 * r] = w0 * i00;
      z[3 * r + 1] = w0 * i10 + w1 * i11;
      z[3 * r + 2] = w0 * i20 + w1 * i21 + w2 * i22;
    }
  }
  __syncthreads();
  if (dbg & 8) return;
  // ---- local blocks: one warp per block, one FP64 tensor-core MMA (m8n8k4) per pair: D(8x8) += A(8x4) B(4x8)
  // with A[i][k] = Z_a[i][k], B[k][j] = Z_b[j][k] (rows/cols 6-7 and k = 3 are zero padding), so lane
  // (g, t) = (lane / 4, lane % 4) feeds element 3 g + t of BOTH rows and ends up holding S[g][2t], S[g][2t+1]:
  // two 8-byte shared loads per pair and lane, no cross-lane reduction, partial rows written straight from the
  // accumulator fragment.  Even and odd pairs go to two accumulators (independent MMA chains, fixed order).
  const int g = lane >> 2, t = lane & 3;
  const bool feeds = g < 6 && t < 3;
  const double* Zl = Z + (feeds ? 3 * g + t : 0);
  for (int lb = lb0 + warp; lb < lb1; lb += NW) {
    const int q0 = d.lb_pair_off[lb] - pq0, q1 = d.lb_pair_off[lb + 1] - pq0;
    double c0 = 0.0, c1 = 0.0, e0_ = 0.0, e1_ = 0.0;
    auto mma_pair = [&](unsigned pr, double& x0, double& x1) {
      double a = 0.0, b = 0.0;
      if (feeds) { a = Zl[(pr & 255u) * kZStride]; b = Zl[(pr >> 8) * kZStride]; }
      dmma884(x0, x1, a, b);
    };
    if (dbg & 1) {
    } else if (pairs_staged) {
      // staged list: an odd first entry alone, then two entries per 4-byte shared load (prs is 16-byte aligned)
      int q = q0;
      if ((q & 1) && q < q1) { mma_pair(prs[q], e0_, e1_); ++q; }
      const uint32_t* pp = reinterpret_cast<const uint32_t*>(prs);
      for (; q + 1 < q1; q += 2) {
        const uint32_t two = pp[q >> 1];
        mma_pair(two & 0xffffu, c0, c1);
        mma_pair(two >> 16, e0_, e1_);
      }
      if (q < q1) mma_pair(prs[q], c0, c1);
    } else {
      const uint16_t* gp = d.pairs + pq0;
      int q = q0;
      for (; q + 1 < q1; q += 2) { mma_pair(__ldg(gp + q), c0, c1); mma_pair(__ldg(gp + q + 1), e0_, e1_); }
      if (q < q1) mma_pair(__ldg(gp + q), c0, c1);
    }
    if (feeds) *reinterpret_cast<double2*>(d.spart + (size_t)d.lb_pos[lb] * 36 + 6 * g + 2 * t) = make_double2(c0 + e0_, c1 + e1_);
  }
  // ---- right-hand side partials: sum over the slot's observations of W_e tp[point(e)] = Z_e u[point(e)]
  for (int cs = cs0 + warp; cs < cs1 && !(dbg & 4); cs += NW) {
    const int c = d.cs_cam[cs];
    const int f = d.fc[c];
    if (f < 0) continue;
    double acc[6] = {0, 0, 0, 0, 0, 0};
    const int j1 = d.cs_obs_off[cs + 1];
    for (int j = d.cs_obs_off[cs] + lane; j < j1; j += 32) {
      const int o = d.sobs[j];
      const double* zz = Z + (size_t)o * kZStride;
      const double* u = us + (int)opl[o] * 3;
      const double t0 = u[0], t1 = u[1], t2 = u[2];
#pragma unroll
      for (int r = 0; r < 6; ++r) acc[r] += zz[3 * r] * t0 + zz[3 * r + 1] * t1 + zz[3 * r + 2] * t2;
    }
    double mine = 0.0;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      double x = acc[r];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if (lane == r) mine = x;
    }
    if (lane < 6) d.rpart[(size_t)d.cs_pos[cs] * 6 + lane] = mine;
  }
}

// ------------------------------------------------------------------------ K0: per-linearisation camera table
// R / J_l / t of every camera at the current state, written once per linearisation; the fused chunk kernel stages
// its slot cameras from here instead of recomputing sin / cos behind a barrier in every CTA.
__global__ void k_campre(Dev d) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= d.Nc) return;
  CamPre cp;
  cam_precompute(d.poses + 6 * (size_t)c, cp);
  d.campre[c] = cp;
}

// ------------------------------------------------------------------- K1 + K2 + K3 fused, one CTA per chunk
// A chunk holds ALL observations of its points, so V_p is complete inside the CTA and the Schur step of the chunk
// can follow its linearisation without W ever coming back from HBM (it is still written once: back-substitution
// reads it).  Used from the second linearisation on (the damping needs the Jacobi scaling fixed at iteration 0 and
// the trust-region radius, both known by then); a rejected step re-runs k_schur_chunk alone with the new radius.
//   A  thread per observation: residual, Huber, block algebra; W -> HBM and kept in registers; camera-side terms
//      (27) and point-side terms (9) -> shared memory
//   B  thread per point: V, g_p -> HBM, damped 3x3 Cholesky -> Linv, u (shared), Vinv, t_p (HBM)
//      || warps from the last one down: per-slot sums of the camera-side terms -> cspart
//   C  thread per observation: Z = W Linv^T into the (now dead) camera staging area; the chunk's pair lists are
//      staged into the (now dead) point staging area
//   D  warp per local block: sum over its pairs of Z_a Z_b^T on the FP64 tensor cores -> spart;
//      warp per slot: sum of Z u -> rpart
constexpr int kFusedThreads = 256;
constexpr int kFZStride = 19;      // odd: observation rows of Z fall into different banks
constexpr int kFIdxSlots = 64;     // k_fused_chunk stages the index tables of chunks with at most this many camera slots
constexpr int kFIdxLb = 192;       // ... and at most this many local blocks (others read them from global memory)

__host__ __device__ __forceinline__ size_t fused_smem_bytes(int n, int np, int ns) {
  return ((size_t)n * (kCamStride + 9) + (size_t)ns * 21 + (size_t)np * 12) * 8 + (size_t)((n + 15) & ~15);
}

__device__ __forceinline__ unsigned long long gtimer();
// NT threads per CTA: 256 for chunks of up to 255 observations, 128 for chunks of up to 128 (twice the CTAs per SM)
template <int OCC, bool PROF = false, int NT = kFusedThreads>
__global__ void __launch_bounds__(NT, OCC) k_fused_chunk(Dev d, DevParams P) {
  // PROF (SBA_DEBUG_LS=2, sba_time_stage 11): thread 0 of a few CTAs prints the nanoseconds between the phase boundaries
  unsigned long long tph[8]; int nph = 0;
  auto mark = [&]() { if constexpr (PROF) { if (threadIdx.x == 0 && nph < 8) tph[nph++] = gtimer(); } };
  mark();
  extern __shared__ __align__(16) double smem[];
  __shared__ double sm1[NT / 32];
  __shared__ double smx[NT / 32];
  const int chunk = blockIdx.x;
  const int4 c0 = __ldg(d.cdesc + 4 * (size_t)chunk), c1 = __ldg(d.cdesc + 4 * (size_t)chunk + 1), c2 = __ldg(d.cdesc + 4 * (size_t)chunk + 2);
  const int p0 = c0.x, np = c0.y, e0 = c0.z, n = c0.w;
  if (n > kMaxSlots) return;                               // a long track: k_lin_long + k_schur_long
  const int w = c1.x, cs0 = c1.y, ns = c1.z, lb0 = c1.w, lb1 = c1.w + c2.x;
  const int pq0 = c2.y, npq = c2.z;
  double* cam_st = smem;                                   // [n * 27]  A: camera-side terms;  C/D: Z [n * 19]
  double* pt_st = cam_st + (size_t)n * kCamStride;         // [n * 9]   A: V, gp terms;        C/D: pair lists
  double* cpre = pt_st + (size_t)n * 9;                    // [ns * 21] R Jl t of the chunk's cameras
  double* xs = cpre + (size_t)ns * 21;                     // [np * 3]  the chunk's points
  double* Li = xs + (size_t)np * 3;                        // [np * 6]  Linv: 00 10 11 20 21 22
  double* us = Li + (size_t)np * 6;                        // [np * 3]  u = Linv gp
  uint8_t* opl = reinterpret_cast<uint8_t*>(us + (size_t)np * 3);   // [n] chunk-local point of each observation
  double* Z = cam_st + 1;                                  // one zero element in front of row 0 (see phase D)
  // pair lists staged as BYTE offsets of the two Z rows (a * 152 | b * 152 << 16): the inner loop spends its issue
  // slots on the two loads and the MMA, not on unpacking indices
  uint32_t* prs = reinterpret_cast<uint32_t*>(pt_st);
  const bool pairs_staged = (size_t)npq * 4 <= (size_t)n * 72;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = NT / 32;
  __shared__ int s_off[kFIdxSlots + 1];      // slot -> first entry of s_sobs
  __shared__ int s_pos[kFIdxSlots];          // slot -> row of cspart / rpart, -1 for a fixed camera
  __shared__ int s_lbq[kFIdxLb + 1];         // local block -> first pair (relative to the chunk's list)
  __shared__ int s_lbpos[kFIdxLb];           // local block -> row of spart
  __shared__ uint8_t s_sobs[256];            // chunk-local observation indices grouped by slot
  const bool idx_staged = ns <= kFIdxSlots && lb1 - lb0 <= kFIdxLb;
  // range of slot sl in the grouped observation list (chunk-local), its output row (-1: fixed camera), an entry of the list
  auto slot_of = [&](int sl, int& j0, int& j1) -> int {
    if (idx_staged) { j0 = s_off[sl]; j1 = s_off[sl + 1]; return s_pos[sl]; }
    const int cs = cs0 + sl;
    j0 = d.cs_obs_off[cs] - e0; j1 = d.cs_obs_off[cs + 1] - e0;
    return (d.fc[d.cs_cam[cs]] >= 0) ? d.cs_pos[cs] : -1;
  };
  auto sobs_at = [&](int j) -> int { return idx_staged ? (int)s_sobs[j] : (int)__ldg(d.sobs + e0 + j); };
  // ---- stage.  ALL global loads first, into registers, in two dependency levels; the shared-memory stores after them.
  // (The pointers in Dev are generic: the compiler will not move a global load above a shared store, so a sequence of
  // "load, store" loops costs one round trip to L2 per loop.)
  const int i = threadIdx.x;
  const bool act = i < n;
  const int nlb = lb1 - lb0;
  // level 1: everything addressed by the chunk record alone
  int sl = 0, pl = 0;
  float4 m = make_float4(0.f, 0.f, 0.f, 0.f);
  if (act) {
    sl = d.oslot[e0 + i];
    pl = __ldg(d.ppt + e0 + i) - p0;
    m = __ldg(d.uv_p + e0 + i);
  }
  constexpr int KX = (kPointChunk * 3 + NT - 1) / NT, KL = (kFIdxLb + NT) / NT;    // point coordinates / block-table entries per thread
  double x_r[KX];
#pragma unroll
  for (int u = 0; u < KX; ++u) x_r[u] = (i + u * NT < np * 3) ? d.pts[(size_t)p0 * 3 + i + u * NT] : 0.0;
  const int cp_s = i / 21;                                          // camera-table element i of the staged block
  const int cp_cam = (i < ns * 21) ? __ldg(d.cs_cam + cs0 + cp_s) : -1;
  static_assert(kFIdxSlots < NT, "one slot entry per thread");
  int r_sobs = 0, r_off = 0, r_cam = -1, r_pos = 0, r_lbq[KL], r_lbpos[KL];
#pragma unroll
  for (int u = 0; u < KL; ++u) { r_lbq[u] = 0; r_lbpos[u] = 0; }
  if (idx_staged) {
    if (act) r_sobs = (int)__ldg(d.sobs + e0 + i);
    if (i <= ns) r_off = __ldg(d.cs_obs_off + cs0 + i) - e0;
    if (i < ns) { r_cam = __ldg(d.cs_cam + cs0 + i); r_pos = __ldg(d.cs_pos + cs0 + i); }
#pragma unroll
    for (int u = 0; u < KL; ++u) {
      if (i + u * NT <= nlb) r_lbq[u] = __ldg(d.lb_pair_off + lb0 + i + u * NT) - pq0;
      if (i + u * NT < nlb) r_lbpos[u] = __ldg(d.lb_pos + lb0 + i + u * NT);
    }
  }
  // the point threads request what phase B needs now (it arrives while phase A computes)
  const int q = threadIdx.x;
  const bool has_pt = q < np;
  double s2[3] = {1, 1, 1};
  if (has_pt) {
#pragma unroll
    for (int k = 0; k < 3; ++k) s2[k] = d.sc2p[(size_t)(p0 + q) * 3 + k];
  }
  const double radius = d.w_radius[w];
  const int pe0 = has_pt ? d.pt_off[p0 + q] - e0 : 0, pe1 = has_pt ? d.pt_off[p0 + q + 1] - e0 : 0;
  // level 2: what a level-1 index addresses
  const double cp_v = (cp_cam >= 0) ? reinterpret_cast<const double*>(d.campre + cp_cam)[i - 21 * cp_s] : 0.0;
  const int r_fc = (r_cam >= 0) ? d.fc[r_cam] : -1;
  // stores
  if (act) opl[i] = (uint8_t)pl;
#pragma unroll
  for (int u = 0; u < KX; ++u) if (i + u * NT < np * 3) xs[i + u * NT] = x_r[u];
  if (cp_cam >= 0) cpre[i] = cp_v;
  if (idx_staged) {
    if (act) s_sobs[i] = (uint8_t)r_sobs;
    if (i <= ns) s_off[i] = r_off;
    if (i < ns) s_pos[i] = (r_fc >= 0) ? r_pos : -1;
#pragma unroll
    for (int u = 0; u < KL; ++u) {
      if (i + u * NT <= nlb) s_lbq[i + u * NT] = r_lbq[u];
      if (i + u * NT < nlb) s_lbpos[i + u * NT] = r_lbpos[u];
    }
  }
  // chunks with more than 12 camera slots: the rest of the camera table
  for (int k = i + NT; k < ns * 21; k += NT) {
    const int s = k / 21;
    cpre[k] = reinterpret_cast<const double*>(d.campre + __ldg(d.cs_cam + cs0 + s))[k - 21 * s];
  }
  __syncthreads();
  mark();
  // ---- A
  double cost = 0.0;
  double W[18];
#pragma unroll
  for (int k = 0; k < 18; ++k) W[k] = 0.0;
  if (act) {
    const CamPre& cp = *reinterpret_cast<const CamPre*>(cpre + 21 * sl);
    const float uv[4] = {m.x, m.y, m.z, m.w};
    const double X[3] = {xs[3 * pl], xs[3 * pl + 1], xs[3 * pl + 2]};
    double qv[3], r[4], D[12];
    const double sq = project(P.pl, P.pr, cp, X, uv, P.rdim, qv, r, D);
    double rho, wt;
    huber(P.huber_delta, sq, rho, wt);
    cost = 0.5 * rho;
    double A[6], h[3], G[9];
    form_A_h(D, r, wt, A, h);
    form_G(cp, qv, G);
    {
      double cu[27];
#pragma unroll
      for (int k = 0; k < 27; ++k) cu[k] = 0.0;
      accum_camera(A, h, G, cu, cu + 21);
      double* cs = cam_st + (size_t)i * kCamStride;
#pragma unroll
      for (int k = 0; k < 27; ++k) cs[k] = cu[k];
    }
    double Vc[9];
    point_contrib(cp, A, h, G, Vc, W);
#pragma unroll
    for (int k = 0; k < 9; ++k) pt_st[(size_t)i * 9 + k] = Vc[k];
    double2* Wo = reinterpret_cast<double2*>(d.W + (size_t)(e0 + i) * 18);
#pragma unroll
    for (int k = 0; k < 9; ++k) Wo[k] = make_double2(W[2 * k], W[2 * k + 1]);
  }
  __syncthreads();
  mark();
  // ---- B: per point V, gp, damped Cholesky
  double gmax = 0.0;
  if (has_pt) {
    const int p = p0 + q;
    double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int o = pe0; o < pe1; ++o) {
#pragma unroll
      for (int k = 0; k < 9; ++k) acc[k] += pt_st[(size_t)o * 9 + k];
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) d.V[(size_t)p * 6 + k] = acc[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      d.gp[(size_t)p * 3 + k] = acc[6 + k];
      const double x = xs[3 * q + k];
      gmax = fmax(gmax, fabs(x - clampd(x - acc[6 + k], P.lo, P.hi)));   // projected-gradient norm
    }
    const double a = acc[0] + lm_damp(s2[0], acc[0], radius);
    const double dd = acc[3] + lm_damp(s2[1], acc[3], radius);
    const double f = acc[5] + lm_damp(s2[2], acc[5], radius);
    const double b = acc[1], c = acc[2], e = acc[4];
    double i00 = 0, i10 = 0, i11 = 0, i20 = 0, i21 = 0, i22 = 0;
    bool ok = a > 0.0;
    if (ok) {
      const double l00 = sqrt(a);
      i00 = 1.0 / l00;
      const double l10 = b * i00, l20 = c * i00;
      const double r1 = dd - l10 * l10;
      ok = r1 > 0.0;
      if (ok) {
        const double l11 = sqrt(r1);
        i11 = 1.0 / l11;
        const double l21 = (e - l20 * l10) * i11;
        const double r2 = f - l20 * l20 - l21 * l21;
        ok = r2 > 0.0 && isfinite(r2);
        if (ok) {
          i22 = 1.0 / sqrt(r2);
          i10 = -l10 * i00 * i11;
          i21 = -l21 * i11 * i22;
          i20 = -(l20 * i00 + l21 * i10) * i22;
        }
      }
    }
    if (!ok) { d.w_fail[w] = 1; i00 = i10 = i11 = i20 = i21 = i22 = 0.0; }   // benign race: every writer stores 1
    const double g0 = acc[6], g1 = acc[7], g2 = acc[8];
    const double u0 = i00 * g0, u1 = i10 * g0 + i11 * g1, u2 = i20 * g0 + i21 * g1 + i22 * g2;
    double* vo = d.Vinv + (size_t)p * 6;
    vo[0] = i00 * i00 + i10 * i10 + i20 * i20; vo[1] = i10 * i11 + i20 * i21; vo[2] = i20 * i22;
    vo[3] = i11 * i11 + i21 * i21; vo[4] = i21 * i22; vo[5] = i22 * i22;
    d.tp[(size_t)p * 3] = i00 * u0 + i10 * u1 + i20 * u2;
    d.tp[(size_t)p * 3 + 1] = i11 * u1 + i21 * u2;
    d.tp[(size_t)p * 3 + 2] = i22 * u2;
    Li[q * 6] = i00; Li[q * 6 + 1] = i10; Li[q * 6 + 2] = i11; Li[q * 6 + 3] = i20; Li[q * 6 + 4] = i21; Li[q * 6 + 5] = i22;
    us[q * 3] = u0; us[q * 3 + 1] = u1; us[q * 3 + 2] = u2;
  }
  mark();
  // ---- per camera slot: U (21) + gc (6), one warp per slot, dealt from the last warp down (the point threads sit
  //      in the first warps); same fixed order as k_lin_chunk
  for (int sl = NW - 1 - warp; sl < ns; sl += NW) {
    int j0, j1;
    const int row = slot_of(sl, j0, j1);
    if (row < 0) continue;            // a fixed camera is not an unknown (src/bundle_adjuster.cpp:113)
    double s0 = 0.0, s1 = 0.0;
    const double* col = cam_st + (lane < 27 ? lane : 0);
    for (int jb = j0; jb < j1; jb += 32) {
      const int mine_o = (jb + lane < j1) ? sobs_at(jb + lane) : 0;
      const int mm = min(32, j1 - jb);
      int t = 0;
      for (; t + 3 < mm; t += 4) {
        const int o0 = __shfl_sync(0xffffffffu, mine_o, t), o1 = __shfl_sync(0xffffffffu, mine_o, t + 1);
        const int o2 = __shfl_sync(0xffffffffu, mine_o, t + 2), o3 = __shfl_sync(0xffffffffu, mine_o, t + 3);
        const double v0 = col[o0 * kCamStride], v1 = col[o1 * kCamStride], v2 = col[o2 * kCamStride], v3 = col[o3 * kCamStride];
        s0 += v0; s1 += v1; s0 += v2; s1 += v3;
      }
      for (; t < mm; ++t) {
        const int o = __shfl_sync(0xffffffffu, mine_o, t);
        const double v = col[o * kCamStride];
        if (t & 1) s1 += v; else s0 += v;
      }
    }
    if (lane < 27) d.cspart[(size_t)row * 27 + lane] = s0 + s1;
  }
  // chunk cost and projected-gradient max: warp partials now, summed after the next barrier
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { cost += __shfl_xor_sync(0xffffffffu, cost, o); gmax = fmax(gmax, __shfl_xor_sync(0xffffffffu, gmax, o)); }
  if (lane == 0) { sm1[warp] = cost; smx[warp] = gmax; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0, mx = 0.0;
#pragma unroll
    for (int k = 0; k < NW; ++k) { t += sm1[k]; mx = fmax(mx, smx[k]); }
    d.lin_part[2 * chunk] = t; d.lin_part[2 * chunk + 1] = mx;
  }
  mark();
  // ---- C: Z = W Linv^T from the registers into the dead camera staging area; pair lists into the dead point staging area
  if (act) {
    const double* L = Li + pl * 6;
    const double i00 = L[0], i10 = L[1], i11 = L[2], i20 = L[3], i21 = L[4], i22 = L[5];
    double* z = Z + (size_t)i * kFZStride;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const double w0 = W[3 * r], w1 = W[3 * r + 1], w2 = W[3 * r + 2];
      z[3 * r] = w0 * i00;
      z[3 * r + 1] = w0 * i10 + w1 * i11;
      z[3 * r + 2] = w0 * i20 + w1 * i21 + w2 * i22;
    }
    z[18] = 0.0;                       // the pad element: what the lanes outside the 6x3 operand feed to the MMA
  }
  if (threadIdx.x == 0) cam_st[0] = 0.0;
  constexpr unsigned kRowBytes = kFZStride * 8;
  static_assert(255u * kRowBytes < 65536u, "row offsets fit 16 bits");
  if (pairs_staged)
    for (int k = threadIdx.x; k < npq; k += NT) {
      const unsigned pr = __ldg(d.pairs + pq0 + k);
      prs[k] = (pr & 255u) * kRowBytes | ((pr >> 8) * kRowBytes) << 16;
    }
  __syncthreads();
  mark();
  // ---- D: local blocks, one warp per block, one FP64 tensor-core MMA (m8n8k4) per pair (see k_schur_chunk).  Lane
  //      (g, t) feeds element 3 g + t of both rows; the lanes outside the 6x3 operand read the rows' zero pad.
  const int g = lane >> 2, t = lane & 3;
  const bool feeds = g < 6 && t < 3;
  // The idle lanes read a zero: the pad behind the row (element 18) in the upper half-warp, the pad in front of it
  // (element -1 = the previous row's pad) in the lower one.  A 64-bit load is served half-warp by half-warp; in the lower
  // half elements 0..11 occupy 24 consecutive banks and element 18 falls on the bank of element 2 (a 2-way conflict on
  // every operand load), element -1 does not; in the upper half (elements 12..17) it is the other way round.
  const char* Zl = reinterpret_cast<const char*>(Z + (feeds ? 3 * g + t : (lane < 16 ? -1 : 18)));
  for (int lb = lb0 + warp; lb < lb1; lb += NW) {
    const int q0 = idx_staged ? s_lbq[lb - lb0] : d.lb_pair_off[lb] - pq0, q1 = idx_staged ? s_lbq[lb - lb0 + 1] : d.lb_pair_off[lb + 1] - pq0;
    double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
    auto mma_off = [&](unsigned en, double& x0, double& x1) {
      const double a = *reinterpret_cast<const double*>(Zl + (en & 0xffffu));
      const double b = *reinterpret_cast<const double*>(Zl + (en >> 16));
      dmma884(x0, x1, a, b);
    };
    if (pairs_staged) {
      // even pairs of the list go to one accumulator chain, odd pairs to the other (fixed order)
      int qq = q0;
      if ((qq & 1) && qq < q1) { mma_off(prs[qq], b0, b1); ++qq; }
      for (; qq + 3 < q1; qq += 4) {
        const uint2 e01 = *reinterpret_cast<const uint2*>(prs + qq), e23 = *reinterpret_cast<const uint2*>(prs + qq + 2);
        mma_off(e01.x, a0, a1); mma_off(e01.y, b0, b1);
        mma_off(e23.x, a0, a1); mma_off(e23.y, b0, b1);
      }
      for (; qq + 1 < q1; qq += 2) {
        const uint2 e01 = *reinterpret_cast<const uint2*>(prs + qq);
        mma_off(e01.x, a0, a1); mma_off(e01.y, b0, b1);
      }
      if (qq < q1) mma_off(prs[qq], a0, a1);
    } else {
      const uint16_t* gp = d.pairs + pq0;
      auto mma_idx = [&](unsigned pr, double& x0, double& x1) { mma_off((pr & 255u) * kRowBytes | ((pr >> 8) * kRowBytes) << 16, x0, x1); };
      int qq = q0;
      if ((qq & 1) && qq < q1) { mma_idx(__ldg(gp + qq), b0, b1); ++qq; }
      for (; qq + 1 < q1; qq += 2) { mma_idx(__ldg(gp + qq), a0, a1); mma_idx(__ldg(gp + qq + 1), b0, b1); }
      if (qq < q1) mma_idx(__ldg(gp + qq), a0, a1);
    }
    if (feeds) *reinterpret_cast<double2*>(d.spart + (size_t)(idx_staged ? s_lbpos[lb - lb0] : d.lb_pos[lb]) * 36 + 6 * g + 2 * t) = make_double2(a0 + b0, a1 + b1);
  }
  mark();
  // ---- right-hand side partials: sum over the slot's observations of W_e tp[point(e)] = Z_e u[point(e)]
  for (int sl = NW - 1 - warp; sl < ns; sl += NW) {
    int j0, j1;
    const int row = slot_of(sl, j0, j1);
    if (row < 0) continue;
    double acc[6] = {0, 0, 0, 0, 0, 0};
    for (int j = j0 + lane; j < j1; j += 32) {
      const int o = sobs_at(j);
      const double* zz = Z + (size_t)o * kFZStride;
      const double* u = us + (int)opl[o] * 3;
      const double t0 = u[0], t1 = u[1], t2 = u[2];
#pragma unroll
      for (int r = 0; r < 6; ++r) acc[r] += zz[3 * r] * t0 + zz[3 * r + 1] * t1 + zz[3 * r + 2] * t2;
    }
    double mine = 0.0;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      double x = acc[r];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if (lane == r) mine = x;
    }
    if (lane < 6) d.rpart[(size_t)row * 6 + lane] = mine;
  }
  if constexpr (PROF) {
    mark();
    if (threadIdx.x == 0 && (blockIdx.x % 211) == 0)
      printf("[fused] cta %d n %d np %d ns %d pairs %d | stage %llu A %llu B %llu slots %llu C %llu D %llu rhs %llu | total %llu ns\n", (int)blockIdx.x, n, np, ns, npq,
             tph[1] - tph[0], tph[2] - tph[1], tph[3] - tph[2], tph[4] - tph[3], tph[5] - tph[4], tph[6] - tph[5], tph[7] - tph[6], tph[7] - tph[0]);
  }
}

// ------------------------------------------------------------------ long tracks (points with > 255 observations)
// Loop-closure re-targeting (src/camera_frame.h:92-109) can give one map point hundreds of observations; the
// reference has no limit.  Such a point is a chunk of its own with one slot per observation and one local block
// per ordered pair of free observations; three small kernels do for it what the chunk kernels do for the rest, in
// the same fixed orders (per-thread strided partials, then a fixed tree), feeding the same partial lists.
constexpr int kLongThreads = 256;

// sum of nine per-thread values over the CTA in a fixed order; every thread returns all nine totals
__device__ __forceinline__ void long_sum9(double (&v)[9], double* sm /* [9 * 8] */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    double x = v[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane == 0) sm[i * 8 + warp] = x;
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < kLongThreads / 32; ++k) t += sm[i * 8 + k];
    v[i] = t;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kLongThreads) k_lin_long(Dev d, DevParams P) {
  __shared__ double sm[9 * 8];
  const int chunk = d.long_chunks[blockIdx.x];
  const int4 c0 = __ldg(d.cdesc + 4 * (size_t)chunk), c1 = __ldg(d.cdesc + 4 * (size_t)chunk + 1);
  const int p = c0.x, e0 = c0.z, n = c0.w, cs0 = c1.y;
  const double X[3] = {d.pts[(size_t)p * 3], d.pts[(size_t)p * 3 + 1], d.pts[(size_t)p * 3 + 2]};
  double acc[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};     // V (6), gp (3) over this thread's observations, in order
  double cost = 0.0;
  for (int i = threadIdx.x; i < n; i += kLongThreads) {
    const int e = e0 + i;
    const int cam = __ldg(d.pcam + e);
    CamPre cp;
    load_campre(d.campre, cam, cp);
    const float4 m = __ldg(d.uv_p + e);
    const float uv[4] = {m.x, m.y, m.z, m.w};
    double q[3], r[4], D[12];
    const double s = project(P.pl, P.pr, cp, X, uv, P.rdim, q, r, D);
    double rho, wt;
    huber(P.huber_delta, s, rho, wt);
    cost += 0.5 * rho;
    double A[6], h[3], G[9];
    form_A_h(D, r, wt, A, h);
    form_G(cp, q, G);
    if (d.fc[cam] >= 0) {            // the observation is its own slot: its camera-side terms ARE the slot's partial row
      double cu[27];
#pragma unroll
      for (int k = 0; k < 27; ++k) cu[k] = 0.0;
      accum_camera(A, h, G, cu, cu + 21);
      double* row = d.cspart + (size_t)d.cs_pos[cs0 + i] * 27;
#pragma unroll
      for (int k = 0; k < 27; ++k) row[k] = cu[k];
    }
    double Vc[9], W[18];
    point_contrib(cp, A, h, G, Vc, W);
#pragma unroll
    for (int k = 0; k < 9; ++k) acc[k] += Vc[k];
    double2* Wo = reinterpret_cast<double2*>(d.W + (size_t)e * 18);
#pragma unroll
    for (int k = 0; k < 9; ++k) Wo[k] = make_double2(W[2 * k], W[2 * k + 1]);
    if (d.r_dbg) {
      const double sq = sqrt(wt);
      double* ro = d.r_dbg + (size_t)d.perm_p[e] * 4;
      ro[0] = sq * r[0]; ro[1] = sq * r[1]; ro[2] = sq * r[2]; ro[3] = sq * r[3];
    }
  }
  long_sum9(acc, sm);
  double cv[9] = {cost, 0, 0, 0, 0, 0, 0, 0, 0};
  long_sum9(cv, sm);
  if (threadIdx.x == 0) {
    double gmax = 0.0;
#pragma unroll
    for (int k = 0; k < 6; ++k) d.V[(size_t)p * 6 + k] = acc[k];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      d.gp[(size_t)p * 3 + k] = acc[6 + k];
      gmax = fmax(gmax, fabs(X[k] - clampd(X[k] - acc[6 + k], P.lo, P.hi)));
    }
    d.lin_part[2 * chunk] = cv[0]; d.lin_part[2 * chunk + 1] = gmax;
  }
}

// damped 3x3 Cholesky of the point, Z = W Linv^T of every observation -> zlong, per-slot right-hand-side rows
__global__ void __launch_bounds__(kLongThreads) k_schur_long(Dev d) {
  __shared__ double sL[6], sU3[3];
  const int li = blockIdx.x, chunk = d.long_chunks[li];
  const int4 c0 = __ldg(d.cdesc + 4 * (size_t)chunk), c1 = __ldg(d.cdesc + 4 * (size_t)chunk + 1);
  const int p = c0.x, e0 = c0.z, n = c0.w, w = c1.x, cs0 = c1.y;
  if (threadIdx.x == 0) {
    const double* V = d.V + (size_t)p * 6;
    const double radius = d.w_radius[w];
    const double a = V[0] + lm_damp(d.sc2p[(size_t)p * 3], V[0], radius);
    const double dd = V[3] + lm_damp(d.sc2p[(size_t)p * 3 + 1], V[3], radius);
    const double f = V[5] + lm_damp(d.sc2p[(size_t)p * 3 + 2], V[5], radius);
    const double b = V[1], c = V[2], e = V[4];
    double i00 = 0, i10 = 0, i11 = 0, i20 = 0, i21 = 0, i22 = 0;
    bool ok = a > 0.0;
    if (ok) {
      const double l00 = sqrt(a);
      i00 = 1.0 / l00;
      const double l10 = b * i00, l20 = c * i00;
      const double r1 = dd - l10 * l10;
      ok = r1 > 0.0;
      if (ok) {
        const double l11 = sqrt(r1);
        i11 = 1.0 / l11;
        const double l21 = (e - l20 * l10) * i11;
        const double r2 = f - l20 * l20 - l21 * l21;
        ok = r2 > 0.0 && isfinite(r2);
        if (ok) {
          i22 = 1.0 / sqrt(r2);
          i10 = -l10 * i00 * i11;
          i21 = -l21 * i11 * i22;
          i20 = -(l20 * i00 + l21 * i10) * i22;
        }
      }
    }
    if (!ok) { d.w_fail[w] = 1; i00 = i10 = i11 = i20 = i21 = i22 = 0.0; }
    const double g0 = d.gp[(size_t)p * 3], g1 = d.gp[(size_t)p * 3 + 1], g2 = d.gp[(size_t)p * 3 + 2];
    const double u0 = i00 * g0, u1 = i10 * g0 + i11 * g1, u2 = i20 * g0 + i21 * g1 + i22 * g2;
    double* vo = d.Vinv + (size_t)p * 6;
    vo[0] = i00 * i00 + i10 * i10 + i20 * i20; vo[1] = i10 * i11 + i20 * i21; vo[2] = i20 * i22;
    vo[3] = i11 * i11 + i21 * i21; vo[4] = i21 * i22; vo[5] = i22 * i22;
    d.tp[(size_t)p * 3] = i00 * u0 + i10 * u1 + i20 * u2;
    d.tp[(size_t)p * 3 + 1] = i11 * u1 + i21 * u2;
    d.tp[(size_t)p * 3 + 2] = i22 * u2;
    sL[0] = i00; sL[1] = i10; sL[2] = i11; sL[3] = i20; sL[4] = i21; sL[5] = i22;
    sU3[0] = u0; sU3[1] = u1; sU3[2] = u2;
  }
  __syncthreads();
  const double i00 = sL[0], i10 = sL[1], i11 = sL[2], i20 = sL[3], i21 = sL[4], i22 = sL[5];
  const double t0 = sU3[0], t1 = sU3[1], t2 = sU3[2];
  double* Zb = d.zlong + (size_t)d.long_zoff[li] * 18;
  for (int i = threadIdx.x; i < n; i += kLongThreads) {
    double W[18];
    load_W(d.W, (size_t)(e0 + i), W);
    double* z = Zb + (size_t)i * 18;
    double acc[6];
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      const double w0 = W[3 * r], w1 = W[3 * r + 1], w2 = W[3 * r + 2];
      const double z0 = w0 * i00, z1 = w0 * i10 + w1 * i11, z2 = w0 * i20 + w1 * i21 + w2 * i22;
      z[3 * r] = z0; z[3 * r + 1] = z1; z[3 * r + 2] = z2;
      acc[r] = z0 * t0 + z1 * t1 + z2 * t2;
    }
    if (d.fc[__ldg(d.pcam + e0 + i)] >= 0) {
      double* row = d.rpart + (size_t)d.cs_pos[cs0 + i] * 6;
#pragma unroll
      for (int r = 0; r < 6; ++r) row[r] = acc[r];
    }
  }
}

// every local block of the long chunks is ONE product Z_a Z_b^T: a thread per element, grid-stride over all of them
__global__ void k_schur_long_pairs(Dev d) {
  const int64_t tot = (int64_t)d.long_lb_off[d.nlong] * 36;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < tot; t += (int64_t)gridDim.x * blockDim.x) {
    const int l = (int)(t / 36), e = (int)(t - 36 * (int64_t)l), r = e / 6, c = e - 6 * r;
    int li = 0;                                              // few long chunks: linear search
    while (l >= d.long_lb_off[li + 1]) ++li;
    const int chunk = d.long_chunks[li];
    const int4 c1 = __ldg(d.cdesc + 4 * (size_t)chunk + 1);
    const int lb = c1.w + (l - d.long_lb_off[li]);
    const int q = d.lb_pair_off[lb];
    const int a = d.pairs[q], b = d.pairs[q + 1];
    const double* Za = d.zlong + ((size_t)d.long_zoff[li] + a) * 18 + 3 * r;
    const double* Zb = d.zlong + ((size_t)d.long_zoff[li] + b) * 18 + 3 * c;
    d.spart[(size_t)d.lb_pos[lb] * 36 + e] = Za[0] * Zb[0] + Za[1] * Zb[1] + Za[2] * Zb[2];
  }
}

// Sum of the contiguous partial rows [l0, l1) of one block / camera by a whole CTA of kAsmThreads: warp w takes
// rows w*8 + pg, + 8*kAsmWarps, ..., so that even a camera seen by > 100 chunks
// needs one or two batches of independent loads; the warps' totals are then added in warp order through shared
// memory.  Thread e < E returns element e (all threads pass the barrier).  Fixed order: bit-reproducible.
constexpr int kAsmThreads = 128;
constexpr int kAsmWarps = kAsmThreads / 32;

template <int E, int EG>
__device__ __forceinline__ double cta_gather(const double* part, int l0, int l1, double* sm /* [kAsmWarps * 4 * EG] */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int pg = lane >> 2, e0 = (lane & 3) * EG;
  double acc[EG];
#pragma unroll
  for (int k = 0; k < EG; ++k) acc[k] = 0.0;
  constexpr int RS = 8 * kAsmWarps;                 // row groups over the CTA
  for (int j = l0 + warp * 8 + pg; j < l1; j += 4 * RS) {
    double v[4][EG];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const bool in = j + RS * u < l1;
      const double* row = part + (size_t)(in ? j + RS * u : l0) * E + e0;
#pragma unroll
      for (int k = 0; k < EG; ++k) v[u][k] = (in && e0 + k < E) ? __ldcg(row + k) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u)
#pragma unroll
      for (int k = 0; k < EG; ++k) acc[k] += v[u][k];
  }
#pragma unroll
  for (int k = 0; k < EG; ++k) {
    double x = acc[k];
    x += __shfl_xor_sync(0xffffffffu, x, 4);
    x += __shfl_xor_sync(0xffffffffu, x, 8);
    x += __shfl_xor_sync(0xffffffffu, x, 16);
    if (lane < 4) sm[warp * 4 * EG + e0 + k] = x;
  }
  __syncthreads();
  double tot = 0.0;
  if (threadIdx.x < E) {
#pragma unroll
    for (int w = 0; w < kAsmWarps; ++w) tot += sm[w * 4 * EG + threadIdx.x];
  }
  return tot;
}

// One CTA per free camera: U (21) and gc (6) = sum of the camera's chunk-slot partials in list order.
__global__ void __launch_bounds__(kAsmThreads) k_assemble_cams(Dev d) {
  __shared__ double sm[kAsmWarps * 28];
  const int f = blockIdx.x;
  const int c = d.free2cam[f];
  const double tot = cta_gather<27, 7>(d.cspart, d.cam_part_off[c], d.cam_part_off[c + 1], sm);
  const int e = threadIdx.x;
  if (e < 21) d.U[(size_t)c * 21 + e] = tot;
  else if (e < 27) d.gc[(size_t)c * 6 + (e - 21)] = tot;
}

// One CTA per upper block of S and one per free camera:
//   block CTA    Su[b] = - sum of the block's local partials in list order
//   camera CTA   rsum[f] = sum of the camera's W tp partials; with `prepare` (single GPU: no all-reduce in
//                between) it also assembles its own diagonal block and does the work of k_cam_prepare:
//                S_ff += U + D, Minv = S_ff^-1, rhs = rsum - gc
// mode bits: 1 prepare · 2 the camera CTA also sums its own U / gc partials (the work of k_assemble_cams: after a fused
// linearisation nothing else is between the chunk kernel and the solve) · 4 no block-Jacobi inverse (exact solvers)
__device__ __forceinline__ void dev_lm_begin(const Dev& d, const DevParams& P, int w, double initial_radius, int first, int clear_fail);
__device__ __forceinline__ void dev_lm_begin(const Dev& d, const DevParams& P, int w, double initial_radius, int first, int clear_fail);
// Ticket of k_assemble_schur (mode bit 8): a window's camera CTAs and its chunk-sum CTA each draw one; whoever draws the
// last one sees everybody's results and finishes the window: cost, gradient maximum (= k_lin_finish), gradient test.
__device__ __forceinline__ void assemble_ticket(const Dev& d, const DevParams& P, int w, int keep_cost, int lm_begin) {
  __shared__ int s_last;
  __shared__ double smx[4];
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int ntick = d.win_fc_off[w + 1] - d.win_fc_off[w] + 1;
    const int t = atomicAdd(d.win_done + w, 1);
    s_last = (t == ntick - 1) ? 1 : 0;
    if (s_last) d.win_done[w] = 0;                           // ready for the next launch
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double gmax = 0.0;
  for (int f = d.win_fc_off[w] + threadIdx.x; f < d.win_fc_off[w + 1]; f += kThreads) {
    const int c = d.free2cam[f];
#pragma unroll
    for (int i = 0; i < 6; ++i) gmax = fmax(gmax, fabs(__ldcg(d.gc + (size_t)c * 6 + i)));
  }
  const double mx = block_max1(gmax, smx);
  if (threadIdx.x == 0) {
    if (!keep_cost) d.w_cost[w] = __ldcg(d.win_part + 2 * w);
    d.w_gmax[w] = fmax(mx, __ldcg(d.win_part + 2 * w + 1));
    if (lm_begin) dev_lm_begin(d, P, w, 0.0, 0, 0);
  }
}

// mode bit 8 (single GPU, every window has a free camera): the LAST camera CTA of a window to finish also runs the
// window's finish (cost, gradient maximum: k_lin_finish) and, with lm_begin != 0, the gradient-tolerance test
// (k_lm_begin, first = 0) — a ticket per window instead of two more launches on a side stream.
__global__ void __launch_bounds__(kAsmThreads) k_assemble_schur(Dev d, int mode, DevParams P, int keep_cost, int lm_begin) {
  const int prepare = mode & 1;
  __shared__ double sm[kAsmWarps * 36];
  __shared__ double sM[36];
  __shared__ double sU[28];
  // mode bit 16 (point-sharded ranks): the grid covers only the blocks / cameras this rank has partials for (asm_ids);
  // the rest of U | gc | S | rsum was zeroed by the caller
  const int bid = (mode & 16) ? __ldg(d.asm_ids + blockIdx.x) : (int)blockIdx.x;
  if (bid < d.nbu) {
    const int b = bid;
    if (prepare && d.blk_row[b] == d.blk_col[b]) return;     // diagonal blocks belong to the camera CTAs
    const double tot = cta_gather<36, 9>(d.spart, d.blk_part_off[b], d.blk_part_off[b + 1], sm);
    if (threadIdx.x < 36) d.Su[(size_t)b * 36 + threadIdx.x] = -tot;
    return;
  }
  const int f = bid - d.nbu;
  if (f >= d.Nf) {
    // mode bit 8: one more CTA per window sums the window's chunk costs and gradient maxima while the camera CTAs work
    const int ww = f - d.Nf;
    if (!(mode & 8) || ww >= d.nwin) return;
    __shared__ double smw[4];
    __shared__ double smxw[4];
    double cost[1] = {0.0};
    double gmax = 0.0;
    for (int ch = d.win_pchunk_off[ww] + threadIdx.x; ch < d.win_pchunk_off[ww + 1]; ch += kThreads) {
      cost[0] += d.lin_part[2 * ch];
      gmax = fmax(gmax, d.lin_part[2 * ch + 1]);
    }
    const double tot = block_sum<1>(cost, smw);            // the same sum, in the same order, as k_lin_finish
    const double mx = block_max1(gmax, smxw);
    if (threadIdx.x == 0) { d.win_part[2 * ww] = tot; d.win_part[2 * ww + 1] = mx; }
    assemble_ticket(d, P, ww, keep_cost, lm_begin);
    return;
  }
  // every index and scalar the two gathers and the prepare step need is requested up front, so that their
  // latencies overlap instead of chaining behind the barriers of cta_gather
  const int c = d.free2cam[f];
  const int db = d.diag_blk[f], w = d.fcam_win[f];
  const int r0 = d.cam_part_off[c], r1 = d.cam_part_off[c + 1];
  int b0 = 0, b1 = 0;
  double u = 0.0, damp = 0.0, gcv = 0.0;
  const int ti = threadIdx.x / 6, tj = threadIdx.x - 6 * ti;
  if (prepare) { b0 = d.blk_part_off[db]; b1 = d.blk_part_off[db + 1]; }
  if (mode & 2) {
    double s2 = 1.0, rad = 1.0;
    if (threadIdx.x < 6) { s2 = d.sc2c[(size_t)c * 6 + threadIdx.x]; rad = d.w_radius[w]; }
    const double t27 = cta_gather<27, 7>(d.cspart, r0, r1, sm);
    if (threadIdx.x < 21) d.U[(size_t)c * 21 + threadIdx.x] = t27;
    else if (threadIdx.x < 27) d.gc[(size_t)c * 6 + (threadIdx.x - 21)] = t27;
    if (threadIdx.x < 27) sU[threadIdx.x] = t27;
    __syncthreads();                                         // sU complete; sm is reused by the next gather
    if (threadIdx.x < 36) u = sU[(ti <= tj) ? up6(ti, tj) : up6(tj, ti)];
    if (threadIdx.x < 6) { gcv = sU[21 + threadIdx.x]; sM[threadIdx.x] = lm_damp(s2, sU[up6(threadIdx.x, threadIdx.x)], rad); }
    __syncthreads();
    if (threadIdx.x < 36 && ti == tj) damp = sM[ti];
    __syncthreads();                                         // sM is reused below
  } else if (prepare) {
    if (threadIdx.x < 36) {
      u = d.U[(size_t)c * 21 + ((ti <= tj) ? up6(ti, tj) : up6(tj, ti))];
      if (ti == tj) damp = lm_damp(d.sc2c[(size_t)c * 6 + ti], u, d.w_radius[w]);
    }
    if (threadIdx.x < 6) gcv = d.gc[(size_t)c * 6 + threadIdx.x];
  }
  const double rs = cta_gather<6, 2>(d.rpart, r0, r1, sm);
  if (threadIdx.x < 6) { d.rsum[(size_t)f * 6 + threadIdx.x] = rs; if (prepare) d.rhs[(size_t)f * 6 + threadIdx.x] = rs - gcv; }
  if (prepare) {
    __syncthreads();                                           // sm is reused
    const double tot = cta_gather<36, 9>(d.spart, b0, b1, sm);
    if (threadIdx.x < 36) {
      const double v = -tot + u + damp;
      sM[threadIdx.x] = v;
      d.Su[(size_t)db * 36 + threadIdx.x] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0 && !(mode & 4)) {
      // one thread per camera factorises (all cameras in parallel, registers only)
      double M[36];
#pragma unroll
      for (int i = 0; i < 36; ++i) M[i] = sM[i];
      if (!inv_spd6(M)) {
        d.w_fail[w] = 1;
#pragma unroll
        for (int i = 0; i < 36; ++i) M[i] = (i % 7 == 0) ? 1.0 : 0.0;
      }
#pragma unroll
      for (int i = 0; i < 36; ++i) d.Minv[(size_t)f * 36 + i] = M[i];
    }
  }
  if (!(mode & 8)) return;
  assemble_ticket(d, P, w, keep_cost, lm_begin);
}

// thread per free camera: add U + D to the diagonal block, rhs = -gc + rsum, Minv = (S_ii)^-1
__global__ void k_cam_prepare(Dev d) {
  const double* __restrict__ rsum = d.rsum;
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= d.Nf) return;
  const int c = d.free2cam[f];
  const int w = d.fcam_win[f];
  const double radius = d.w_radius[w];
  const int db = d.diag_blk[f];
  double* S = d.Su + (size_t)db * 36;
  double M[36];
  const double* U = d.U + (size_t)c * 21;
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const double u = (i <= j) ? U[up6(i, j)] : U[up6(j, i)];
      double v = S[6 * i + j] + u;
      if (i == j) v += lm_damp(d.sc2c[(size_t)c * 6 + i], u, radius);
      M[6 * i + j] = v;
    }
#pragma unroll
  for (int i = 0; i < 36; ++i) S[i] = M[i];
  if (!inv_spd6(M)) {
    d.w_fail[w] = 1;
#pragma unroll
    for (int i = 0; i < 36; ++i) M[i] = (i % 7 == 0) ? 1.0 : 0.0;
  }
#pragma unroll
  for (int i = 0; i < 36; ++i) d.Minv[(size_t)f * 36 + i] = M[i];
#pragma unroll
  for (int i = 0; i < 6; ++i) d.rhs[(size_t)f * 6 + i] = rsum[(size_t)f * 6 + i] - d.gc[(size_t)c * 6 + i];
}

// ------------------------------------------------------------------------------------- K4: PCG
// Block-Jacobi preconditioned CG on the reduced camera system, one independent solve per window,
// all windows advanced by the same launches; scalars (alpha, beta, done flags) live on the device.

// thread per free camera: x = 0, r = rhs, z = Minv r, p = z
__global__ void k_pcg_init(Dev d) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= d.Nf) return;
  double r[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) r[i] = d.rhs[(size_t)f * 6 + i];
  const double* M = d.Minv + (size_t)f * 36;
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    double z = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) z += M[6 * i + j] * r[j];
    d.x[(size_t)f * 6 + i] = 0.0;
    d.r[(size_t)f * 6 + i] = r[i];
    d.z[(size_t)f * 6 + i] = z;
    d.p[(size_t)f * 6 + i] = z;
  }
}

// one CTA per window: out = a . b over the window's free-camera rows (fixed order), then the
// scalar recurrences.  mode 0: rz0 (init)   mode 1: pAp -> alpha   mode 2: rz_new -> beta, done test
__global__ void __launch_bounds__(kThreads) k_pcg_dot(Dev d, DevParams P, const double* __restrict__ a,
                                                      const double* __restrict__ b, int mode, int max_it) {
  __shared__ double sm[4];
  const int w = blockIdx.x;
  double s[1] = {0.0};
  const int i1 = d.win_fc_off[w + 1] * 6;
  for (int i = d.win_fc_off[w] * 6 + threadIdx.x; i < i1; i += kThreads) s[0] += a[i] * b[i];
  const double tot = block_sum<1>(s, sm);
  if (threadIdx.x != 0) return;
  if (mode == 0) {
    d.w_rz[w] = tot; d.w_rz0[w] = tot; d.w_pcg_it[w] = 0; d.w_alpha[w] = 0.0; d.w_beta[w] = 0.0;
    int done = 0;
    if (d.w_status[w] != 0 || d.w_fail[w] != 0) done = 1;
    else if (!(tot > 0.0)) { done = 1; if (!(tot == 0.0)) d.w_fail[w] = 1; }   // rhs == 0: x = 0 is the solution; NaN: fail
    d.w_pcg_done[w] = done;
    if (!done) atomicAdd(d.counters + 1, 1);
  } else if (mode == 1) {
    if (d.w_pcg_done[w]) { d.w_alpha[w] = 0.0; return; }
    if (!(tot > 0.0)) { d.w_fail[w] = 1; d.w_pcg_done[w] = 1; d.w_alpha[w] = 0.0; return; }   // S not SPD
    d.w_pAp[w] = tot;
    d.w_alpha[w] = d.w_rz[w] / tot;
  } else {
    if (d.w_pcg_done[w]) { d.w_beta[w] = 0.0; d.w_alpha[w] = 0.0; return; }
    const double rz_old = d.w_rz[w];
    d.w_beta[w] = tot / rz_old;
    d.w_rz[w] = tot;
    const int it = ++d.w_pcg_it[w];
    if (!isfinite(tot)) { d.w_fail[w] = 1; d.w_pcg_done[w] = 1; }
    else if (tot <= P.pcg_tol2 * d.w_rz0[w] || it >= max_it) d.w_pcg_done[w] = 1;
    else atomicAdd(d.counters + 1, 1);     // control-plane counter only (windows still iterating)
  }
}

// thread per scalar row: Ap = S p  (block-CSR over the full pattern; upper blocks stored once)
__global__ void k_pcg_matvec(Dev d) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= d.Nf * 6) return;
  const int f = t / 6, r = t - 6 * f;
  double s = 0.0;
  const int j1 = d.row_ptr[f + 1];
  for (int j = d.row_ptr[f]; j < j1; ++j) {
    const int col = __ldg(d.bsr_col + j);
    const int bi = __ldg(d.bsr_blk + j);
    const double* x = d.p + (size_t)col * 6;
    if (bi >= 0) {
      const double* B = d.Su + (size_t)bi * 36 + 6 * r;
#pragma unroll
      for (int c = 0; c < 6; ++c) s += B[c] * x[c];
    } else {
      const double* B = d.Su + (size_t)(~bi) * 36 + r;   // transposed
#pragma unroll
      for (int c = 0; c < 6; ++c) s += B[6 * c] * x[c];
    }
  }
  d.Ap[t] = s;
}

// thread per free camera: x += alpha p ; r -= alpha Ap ; z = Minv r
__global__ void k_pcg_update_xr(Dev d) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= d.Nf) return;
  const int w = d.fcam_win[f];
  if (d.w_pcg_done[w]) return;
  const double alpha = d.w_alpha[w];
  double r[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const size_t k = (size_t)f * 6 + i;
    d.x[k] += alpha * d.p[k];
    r[i] = d.r[k] - alpha * d.Ap[k];
    d.r[k] = r[i];
  }
  const double* M = d.Minv + (size_t)f * 36;
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    double z = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) z += M[6 * i + j] * r[j];
    d.z[(size_t)f * 6 + i] = z;
  }
}

// thread per scalar row: p = z + beta p
__global__ void k_pcg_update_p(Dev d) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= d.Nf * 6) return;
  const int w = d.fcam_win[t / 6];
  if (d.w_pcg_done[w]) return;
  d.p[t] = d.z[t] + d.w_beta[w] * d.p[t];
}

// Whole PCG solve in ONE launch: one CTA per window, vectors, block pattern and (when it fits) the window's
// blocks of S and the block-Jacobi inverses in shared memory.  Used when every window's reduced system is small
// enough (local / sliding windows, batched windows); the multi-launch kernels above remain for global problems.
// Same recurrences and stopping rule as k_pcg_dot.  Per iteration: three CTA barriers.
//   mat-vec   one warp per block-row; lane (ps, i) = (lane / 6, lane % 6) owns scalar row i of every fifth entry
//             of the block-row (6 FMAs per entry), the five partial rows are summed in lane order
//   dots      fp64, fixed order: warp butterfly -> one partial per warp in shared memory -> every warp re-reduces
//             the 32 partials with the same butterfly (all threads hold the same bits)
//   r         ping-pong buffers, so z = Minv r_new needs no barrier between the r update and its use
constexpr int kPcgThreads = 1024;
constexpr int kPcgWarps = kPcgThreads / 32;

__host__ __device__ __forceinline__ size_t pcg_smem_bytes(int nf, int nb_upper, int nb_full, bool s_in_smem) {
  size_t dbl = (size_t)6 * 6 * nf + 2 * kPcgWarps;                      // x r0 r1 z p Ap + two reduction buffers
  if (s_in_smem) dbl += (size_t)36 * nf + (size_t)36 * nb_upper;
  return dbl * 8 + ((size_t)(nf + 1) + 2 * (size_t)nb_full) * 4;
}

__device__ __forceinline__ double pcg_reduce(double part, double* red, int lane, int warp) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if (lane == 0) red[warp] = part;
  __syncthreads();
  double t = (lane < (int)(blockDim.x >> 5)) ? red[lane] : 0.0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  return t;
}

__global__ void __launch_bounds__(kPcgThreads, 1) k_pcg_window(Dev d, DevParams P, int max_it, int s_in_smem) {
  extern __shared__ double smem[];
  const int w = blockIdx.x;
  const int f0 = d.win_fc_off[w], f1 = d.win_fc_off[w + 1];
  const int nf = f1 - f0, n = 6 * nf;
  if (threadIdx.x == 0) d.w_pcg_it[w] = 0;
  if (nf <= 0) return;
  if (d.w_status[w] != 0 || d.w_fail[w] != 0) return;     // uniform per CTA; leaves x untouched (unused)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nthr = blockDim.x, nwarps = blockDim.x >> 5;     // 64..1024 threads, sized to the window by the host
  double* x = smem; double* ra = x + n; double* rb = ra + n; double* z = rb + n; double* pv = z + n; double* Ap = pv + n;
  double* redA = Ap + n; double* redB = redA + kPcgWarps;
  const int b0 = d.diag_blk[f0];
  const int b1 = (f1 < d.Nf) ? d.diag_blk[f1] : d.nbu;
  const int e0 = d.row_ptr[f0], ne = d.row_ptr[f1] - e0;
  const double* S = d.Su + (size_t)b0 * 36;
  const double* M = d.Minv + (size_t)f0 * 36;
  double* next = redB + kPcgWarps;
  if (s_in_smem) {
    double* Ms = next; double* Ss = Ms + (size_t)nf * 36;
    for (int i = threadIdx.x; i < nf * 36; i += nthr) Ms[i] = M[i];
    for (int i = threadIdx.x; i < (b1 - b0) * 36; i += nthr) Ss[i] = S[i];
    S = Ss; M = Ms;
    next = Ss + (size_t)(b1 - b0) * 36;
  }
  int* rp = reinterpret_cast<int*>(next);                  // [nf + 1] entry offsets, window-local
  int* ecol = rp + nf + 1;                                 // [ne] 6 * local block-column
  int* eblk = ecol + ne;                                   // [ne] element offset of the block in S; < 0: ~offset, transposed
  for (int i = threadIdx.x; i <= nf; i += nthr) rp[i] = d.row_ptr[f0 + i] - e0;
  for (int i = threadIdx.x; i < ne; i += nthr) {
    ecol[i] = 6 * (d.bsr_col[e0 + i] - f0);
    const int bi = d.bsr_blk[e0 + i];
    eblk[i] = (bi >= 0) ? (bi - b0) * 36 : ~((~bi - b0) * 36);
  }
  for (int i = threadIdx.x; i < n; i += nthr) { x[i] = 0.0; ra[i] = d.rhs[(size_t)f0 * 6 + i]; }
  __syncthreads();
  double* r = ra; double* rn = rb;
  // z = Minv r, p = z, rz
  double part = 0.0;
  for (int i = threadIdx.x; i < n; i += nthr) {
    const int f = i / 6, k = i - 6 * f;
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) acc += M[(size_t)f * 36 + 6 * k + j] * r[6 * f + j];
    z[i] = acc; pv[i] = acc;
    part += r[i] * acc;
  }
  double rz = pcg_reduce(part, redB, lane, warp);
  const double rz0 = rz;
  int it = 0;
  bool fail = false;
  const int ps = lane / 6, ri = lane - 6 * ps;
  if (!(rz > 0.0)) { if (!(rz == 0.0)) fail = true; }      // rhs == 0: x = 0 is the solution; NaN: fail
  else {
    for (;;) {
      // ---- Ap = S p  (block-CSR over the full pattern; upper blocks stored once), pAp
      part = 0.0;
      for (int f = warp; f < nf; f += nwarps) {
        double acc = 0.0;
        if (lane < 30) {
          const int j1 = rp[f + 1];
          for (int j = rp[f] + ps; j < j1; j += 5) {
            const double* xv = pv + ecol[j];
            const int off = eblk[j];
            if (off >= 0) {
              const double* B = S + off + 6 * ri;
              acc += (B[0] * xv[0] + B[1] * xv[1] + B[2] * xv[2]) + (B[3] * xv[3] + B[4] * xv[4] + B[5] * xv[5]);
            } else {
              const double* B = S + (~off) + ri;
              acc += (B[0] * xv[0] + B[6] * xv[1] + B[12] * xv[2]) + (B[18] * xv[3] + B[24] * xv[4] + B[30] * xv[5]);
            }
          }
        }
        double tot = acc;                                  // rows of the five entry groups, in lane order
#pragma unroll
        for (int g = 1; g < 5; ++g) tot += __shfl_sync(0xffffffffu, acc, (lane + 6 * g) & 31);
        if (lane < 6) { Ap[6 * f + lane] = tot; part += pv[6 * f + lane] * tot; }
      }
      const double pAp = pcg_reduce(part, redA, lane, warp);
      if (!(pAp > 0.0)) { fail = true; break; }             // S not SPD
      const double alpha = rz / pAp;
      // ---- x += alpha p ; r_new = r - alpha Ap ; z = Minv r_new ; rz_new
      part = 0.0;
      for (int i = threadIdx.x; i < n; i += nthr) {
        const int f = i / 6, k = i - 6 * f;
        x[i] += alpha * pv[i];
        double acc = 0.0, mine = 0.0;
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          const double rj = r[6 * f + j] - alpha * Ap[6 * f + j];
          if (j == k) mine = rj;
          acc += M[(size_t)f * 36 + 6 * k + j] * rj;
        }
        rn[i] = mine; z[i] = acc;
        part += mine * acc;
      }
      const double rz_new = pcg_reduce(part, redB, lane, warp);
      { double* t = r; r = rn; rn = t; }
      ++it;
      if (!isfinite(rz_new)) { fail = true; break; }
      if (rz_new <= P.pcg_tol2 * rz0 || it >= max_it) { rz = rz_new; break; }
      const double beta = rz_new / rz;
      rz = rz_new;
      for (int i = threadIdx.x; i < n; i += nthr) pv[i] = z[i] + beta * pv[i];
      __syncthreads();
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += nthr) { d.x[(size_t)f0 * 6 + i] = x[i]; d.r[(size_t)f0 * 6 + i] = r[i]; }
  if (threadIdx.x == 0) { d.w_pcg_it[w] = it; if (fail) d.w_fail[w] = 1; atomicMax(d.counters + 4, it); }   // control-plane counter
}

// ------------------------------------------------------------------ K4 (small windows): direct solve
// The reference factors the reduced camera matrix exactly (Ceres SPARSE_SCHUR + CHOLMOD, src/bundle_adjuster.cpp:
// 26-29).  For a local-BA window (<= kCholMaxCams free cameras, a dense-ish S of a few hundred unknowns) the same
// is cheaper on the GPU than 100+ CG iterations: one CTA per window runs a right-looking BLOCK Cholesky (6x6
// blocks) on the dense lower block-triangle of S, kept in an L2-resident workspace; the panel column of each step
// and the inverses of the diagonal factors live in shared memory.  The right-hand side rides along as an extra
// block row (forward substitution for free); the back substitution walks the rows of L upwards.  Everything is
// done in a fixed order by one CTA: bit-reproducible.  Three CTA barriers per block column.
constexpr int kCholThreads = 512;
constexpr int kCholMaxCams = 64;

constexpr int kCholSmemCams = 20;    // 210 lower blocks = 60 KB
__host__ __device__ __forceinline__ size_t chol_smem_bytes(int nf) {
  return ((size_t)nf * (36 + 36 + 6) + 36 + 16 + (nf <= kCholSmemCams ? (size_t)nf * (nf + 1) / 2 * 36 : 0)) * 8;
}
__host__ __device__ __forceinline__ int64_t chol_ws_elems(int nf) { return (int64_t)nf * (nf + 1) / 2 * 36; }

// In: a = SPD 6x6 (row-major, lower triangle read).  Out: li = inverse of its lower Cholesky factor (row-major, lower).
// 1/L_jj comes from rsqrt (one dependent step per column instead of sqrt + divide); L_jj = s * rsqrt(s).
__device__ __forceinline__ bool chol6_inverse_factor(const double* a, double* li) {
  double L[6][6], inv[6];
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double s = a[6 * j + j];
#pragma unroll
    for (int t = 0; t < 6; ++t) if (t < j) s -= L[j][t] * L[j][t];
    ok = ok && (s > 0.0) && isfinite(s);
    inv[j] = rsqrt(ok ? s : 1.0);
#pragma unroll
    for (int i = 0; i < 6; ++i) if (i > j) {
      double v = a[6 * i + j];
#pragma unroll
      for (int t = 0; t < 6; ++t) if (t < j) v -= L[i][t] * L[j][t];
      L[i][j] = v * inv[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      double v = 0.0;
      if (i == j) v = inv[j];
      else if (i > j) {
        double acc = 0.0;
#pragma unroll
        for (int t = 0; t < 6; ++t) if (t >= j && t < i) acc += L[i][t] * li[6 * t + j];
        v = -acc * inv[i];
      }
      li[6 * i + j] = v;
    }
  }
  return ok;
}

// 1 / sqrt(s) to double precision without the library routine's latency: single-precision MUFU seed (22 bits), two
// Newton steps in fp64 (44, then 88 bits).  Arguments outside the float range take the library path.
__device__ __forceinline__ double fast_rsqrt(double s) {
  if (!(s > 1e-30 && s < 1e30)) return rsqrt(s);
  double y = (double)rsqrtf((float)s);
  const double h = 0.5 * s;
  y = y * (1.5 - h * y * y);
  y = y * (1.5 - h * y * y);
  return y;
}

// chol6_inverse_factor with the short-latency reciprocal square root: ONE thread, everything in registers — the rows'
// updates are independent, so the thread's own instruction-level parallelism covers what a warp-wide version pays
// in shuffles (the banded and block-sparse factorisations call this once per block column, on their critical path).
__device__ __forceinline__ bool chol6_inverse_factor_fast(const double* a, double* li) {
  double L[6][6], inv[6];
  bool ok = true;
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) if (j <= i) L[i][j] = a[6 * i + j];
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    const double s = L[j][j];
    ok = ok && (s > 0.0) && isfinite(s);
    inv[j] = fast_rsqrt(ok ? s : 1.0);
#pragma unroll
    for (int i = 0; i < 6; ++i) if (i > j) L[i][j] *= inv[j];
#pragma unroll
    for (int i = 0; i < 6; ++i) if (i > j) {
#pragma unroll
      for (int c = 0; c < 6; ++c) if (c > j && c <= i) L[i][c] -= L[i][j] * L[c][j];
    }
  }
#pragma unroll
  for (int j = 0; j < 6; ++j) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      double v = 0.0;
      if (i == j) v = inv[j];
      else if (i > j) {
        double acc = 0.0;
#pragma unroll
        for (int t = 0; t < 6; ++t) if (t >= j && t < i) acc += L[i][t] * li[6 * t + j];
        v = -acc * inv[i];
      }
      li[6 * i + j] = v;
    }
  }
  return ok;
}

// WS_SMEM: the lower block-triangle itself sits in shared memory (windows of <= kCholSmemCams free cameras: no L2
// round trip anywhere in the factorisation); otherwise in the L2-resident workspace.
template <bool WS_SMEM>
__global__ void __launch_bounds__(kCholThreads, 1) k_chol_window(Dev d) {
  const int T = blockDim.x;                // 128 .. kCholThreads, sized to the window by the host
  extern __shared__ double smem[];
  const int w = blockIdx.x;
  const int f0 = d.win_fc_off[w], f1 = d.win_fc_off[w + 1];
  const int nf = f1 - f0, n = 6 * nf;
  if (threadIdx.x == 0) d.w_pcg_it[w] = 0;
  if (nf <= 0) return;
  if (d.w_status[w] != 0 || d.w_fail[w] != 0) return;     // uniform per CTA; leaves x untouched (unused)
  const int tid = threadIdx.x;
  double* P = smem;                         // [nf * 36] panel: L_ik of the current column k (rows i > k)
  double* LI = P + (size_t)nf * 36;         // [nf * 36] inverse diagonal factors
  double* y = LI + (size_t)nf * 36;         // [n]       right-hand side -> forward-substituted -> solution
  double* akk = y + n;                      // [36]
  double* xk = akk + 36;                    // [6]
  __shared__ int bad;
  double* A = WS_SMEM ? (xk + 8) : (d.chol_ws + d.chol_off[w]);    // lower block (i, j), i >= j, at (i (i + 1) / 2 + j) * 36, row-major
  const int nblk = nf * (nf + 1) / 2;
  auto ld = [](const double* p) { return WS_SMEM ? *p : __ldcg(p); };
  auto ld2 = [](const double2* p) { return WS_SMEM ? *p : __ldcg(p); };
  auto st2 = [](double2* p, double2 v) { if (WS_SMEM) *p = v; else __stcg(p, v); };
  const int b0 = d.diag_blk[f0];
  const int b1 = (f1 < d.Nf) ? d.diag_blk[f1] : d.nbu;
  if (tid == 0) bad = 0;
  for (int i = tid; i < nblk * 36; i += T) A[i] = 0.0;
  for (int i = tid; i < n; i += T) y[i] = d.rhs[(size_t)f0 * 6 + i];
  __syncthreads();
  // upper block (r, c), r <= c, of S  ->  lower block (c, r) = its transpose
  for (int i = tid; i < (b1 - b0) * 36; i += T) {
    const int b = b0 + i / 36, e = i % 36, er = e / 6, ec = e - 6 * er;
    const int r = d.blk_row[b] - f0, c = d.blk_col[b] - f0;
    A[(size_t)(c * (c + 1) / 2 + r) * 36 + 6 * ec + er] = d.Su[(size_t)b * 36 + e];
  }
  __syncthreads();
  for (int k = 0; k < nf; ++k) {
    const int m = nf - 1 - k;               // rows below the diagonal
    // (a) diagonal block: factor, keep the inverse of the factor
    if (tid < 36) akk[tid] = ld(A + (size_t)(k * (k + 1) / 2 + k) * 36 + tid);
    __syncthreads();
    if (tid == 0) {
      double li[36];
      if (!chol6_inverse_factor(akk, li)) bad = 1;
#pragma unroll
      for (int i = 0; i < 36; ++i) LI[(size_t)k * 36 + i] = li[i];
    }
    __syncthreads();
    // (b) panel L_ik = A_ik Linv^T for i > k (kept in shared memory and written back for the back substitution);
    //     right-hand side block: y_k <- Linv y_k
    const double* li = LI + (size_t)k * 36;
    for (int idx = tid; idx < m * 36; idx += T) {
      const int i = k + 1 + idx / 36, e = idx % 36, r = e / 6, c = e - 6 * r;
      double* blk = A + (size_t)(i * (i + 1) / 2 + k) * 36;
      double acc = 0.0;
#pragma unroll
      for (int t = 0; t < 6; ++t) if (t <= c) acc += ld(blk + 6 * r + t) * li[6 * c + t];
      P[(size_t)i * 36 + e] = acc;
    }
    if (tid >= T - 6) {
      const int c = tid - (T - 6);
      double acc = 0.0;
#pragma unroll
      for (int t = 0; t < 6; ++t) if (t <= c) acc += li[6 * c + t] * y[6 * k + t];
      xk[c] = acc;
    }
    __syncthreads();
    for (int idx = tid; idx < m * 36; idx += T) {
      const int i = k + 1 + idx / 36, e = idx % 36;
      A[(size_t)(i * (i + 1) / 2 + k) * 36 + e] = P[(size_t)i * 36 + e];
    }
    if (tid < 6) y[6 * k + tid] = xk[tid];
    // (c) trailing update A_ij -= L_ik L_jk^T (i >= j > k) and y_j -= L_jk y_k
    //     one thread per (block, row): its row of L_ik in registers, L_jk streamed with 16-byte shared loads
    const int nupd = m * (m + 1) / 2;
    for (int idx = tid; idx < nupd * 6; idx += T) {
      const int bi = idx / 6, r = idx - 6 * bi;
      int ii = (int)((sqrtf(8.f * (float)bi + 1.f) - 1.f) * 0.5f);
      while ((ii + 1) * (ii + 2) / 2 <= bi) ++ii;
      while (ii * (ii + 1) / 2 > bi) --ii;
      const int jj = bi - ii * (ii + 1) / 2;
      const int i = k + 1 + ii, j = k + 1 + jj;
      const double2* Li = reinterpret_cast<const double2*>(P + (size_t)i * 36 + 6 * r);
      const double2 l0 = Li[0], l1 = Li[1], l2 = Li[2];
      const double2* Lj = reinterpret_cast<const double2*>(P + (size_t)j * 36);
      double acc[6];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        const double2 m0 = Lj[3 * c], m1 = Lj[3 * c + 1], m2 = Lj[3 * c + 2];
        acc[c] = (l0.x * m0.x + l0.y * m0.y + l1.x * m1.x) + (l1.y * m1.y + l2.x * m2.x + l2.y * m2.y);
      }
      double2* dst = reinterpret_cast<double2*>(A + (size_t)(i * (i + 1) / 2 + j) * 36 + 6 * r);
#pragma unroll
      for (int h = 0; h < 3; ++h) {
        double2 v = ld2(dst + h);
        v.x -= acc[2 * h]; v.y -= acc[2 * h + 1];
        st2(dst + h, v);
      }
    }
    for (int idx = tid; idx < m * 6; idx += T) {
      const int j = k + 1 + idx / 6, r = idx % 6;
      const double* Lj = P + (size_t)j * 36 + 6 * r;
      double acc = 0.0;
#pragma unroll
      for (int t = 0; t < 6; ++t) acc += Lj[t] * xk[t];
      y[6 * j + r] -= acc;
    }
    __syncthreads();
  }
  // back substitution L^T x = y, rows of L upwards: x_k = Linv_kk^T y_k, then y_j -= L_kj^T x_k for j < k
  for (int k = nf - 1; k >= 0; --k) {
    const double* li = LI + (size_t)k * 36;
    if (tid < 6) {
      double acc = 0.0;
#pragma unroll
      for (int t = 0; t < 6; ++t) if (t >= tid) acc += li[6 * t + tid] * y[6 * k + t];
      xk[tid] = acc;
    }
    __syncthreads();
    if (tid < 6) y[6 * k + tid] = xk[tid];
    for (int idx = tid; idx < k * 6; idx += T) {
      const int j = idx / 6, c = idx - 6 * j;
      const double* blk = A + (size_t)(k * (k + 1) / 2 + j) * 36;
      double acc = 0.0;
#pragma unroll
      for (int r = 0; r < 6; ++r) acc += ld(blk + 6 * r + c) * xk[r];
      y[6 * j + c] -= acc;
    }
    __syncthreads();
  }
  const bool fail = bad != 0;
  for (int i = tid; i < n; i += T) {
    const double v = y[i];
    d.x[(size_t)f0 * 6 + i] = (fail || !isfinite(v)) ? 0.0 : v;
    d.r[(size_t)f0 * 6 + i] = 0.0;                         // exact solve: no linear residual in the model change
    if (!isfinite(v)) bad = 1;                             // benign race: every writer stores 1
  }
  __syncthreads();
  if (tid == 0 && bad != 0) d.w_fail[w] = 1;
}

// ---- banded windows: the whole factorisation in ONE CTA's shared memory ------------------------------------
// A sliding window's reduced matrix is banded in the reference's frame order (co-visibility reaches a few keyframes
// back: C2 has 322 of 1225 lower blocks), and Cholesky fill stays inside the monotone column envelope: column k holds
// block rows k..last[k], last[] made non-decreasing over each column's reach.  When the envelope of every window fits
// shared memory (C2: 343 blocks = 99 KB; a dense 9-camera window: 45 blocks) one CTA per window runs the right-looking
// block Cholesky with nothing but CTA barriers on the critical path:
//   (a) warp 0: 6x6 Cholesky of the diagonal block across six lanes (rows in registers, columns handed round by
//       shuffles, 1/L_jj by rsqrt) and the inverse of the factor, one column per lane
//   (b) panel L_ik = A_ik Linv^T for the m = last[k] - k rows below, y_k = Linv y_k
//   (c) trailing update A_ij -= L_ik L_jk^T over the m (m + 1) / 2 blocks of the envelope it touches, a thread per
//       element; y_j -= L_jk y_k
// and the back substitution by warp 0 alone.  Fixed order throughout: bit-reproducible.
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

constexpr int kBandThreads = 512;

// envelope blocks, inverse diagonal factors, right-hand side, the (ii, jj) table of the trailing-update pairs, ints
__host__ __device__ __forceinline__ size_t chol_band_smem_bytes(int nf, int env_blocks) {
  return ((size_t)env_blocks * 36 + (size_t)nf * 36 + (size_t)nf * 6 + 48) * 8 + ((size_t)2 * nf + 2) * 4 + (size_t)nf * (nf + 1);
}

// Lanes 0..5 of the calling warp hold row `lane` of a symmetric positive definite 6x6 block in a[0..5] (lower part
// read).  On return every lane j < 6 holds COLUMN j of the inverse of its lower Cholesky factor in x[0..5] (x[i] =
// Linv[i][j], zero above the diagonal).  Returns false (on all lanes) when a pivot is not positive.
__device__ __forceinline__ bool warp_chol6_inverse(double (&a)[6], double (&x)[6], int lane) {
  double inv[6];
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    const double djj = __shfl_sync(0xffffffffu, a[j], j);
    ok = ok && (djj > 0.0) && isfinite(djj);
    inv[j] = rsqrt(ok ? djj : 1.0);
    const double lij = a[j] * inv[j];             // lane i: L[i][j] (meaningful for i >= j; L[j][j] = d * rsqrt(d))
    a[j] = lij;
#pragma unroll
    for (int c = 0; c < 6; ++c) if (c > j) {
      const double lcj = __shfl_sync(0xffffffffu, lij, c);
      if (lane >= c) a[c] -= lij * lcj;
    }
  }
  // every lane gets the strictly lower part of L, then lane j solves L x = e_j
  double L[6][6];
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int t = 0; t < 6; ++t) if (t < i) L[i][t] = __shfl_sync(0xffffffffu, a[t], i);
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    double acc = 0.0;
#pragma unroll
    for (int t = 0; t < 6; ++t) if (t < i) acc += L[i][t] * x[t];
    // x[i] = (e_j[i] - acc) / L[i][i];  rows above the lane's own index stay zero
    x[i] = (i < lane) ? 0.0 : ((i == lane) ? inv[i] : -acc * inv[i]);
  }
  return ok;
}

__global__ void __launch_bounds__(kBandThreads) k_chol_band(Dev d, const int32_t* __restrict__ band_last, int profile) {
  extern __shared__ double smem[];
  constexpr int T = kBandThreads;
  // profile != 0 (SBA_DEBUG_LS=2): thread 0 of window 0 prints the nanoseconds per phase, summed over the columns (tuning aid)
  const bool prof = profile != 0 && blockIdx.x == 0 && threadIdx.x == 0;
  unsigned long long tq0 = prof ? gtimer() : 0, tq[6] = {0, 0, 0, 0, 0, 0};
  auto lap = [&](int i) { if (prof) { const unsigned long long t = gtimer(); tq[i] += t - tq0; tq0 = t; } };
  const int w = blockIdx.x;
  const int f0 = d.win_fc_off[w], f1 = d.win_fc_off[w + 1];
  const int nf = f1 - f0, n = 6 * nf;
  if (threadIdx.x == 0) d.w_pcg_it[w] = 0;
  if (nf <= 0) return;
  if (d.w_status[w] != 0 || d.w_fail[w] != 0) return;     // uniform per CTA; leaves x untouched (unused)
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ int bad;
  __shared__ int env_total, m_max;
  // per-window envelope: last[k] (window-local) and the block offset of every column
  int* lastk = reinterpret_cast<int*>(smem);                 // [nf]
  int* coff = lastk + nf;                                    // [nf + 1]
  double* base = smem + (2 * nf + 2 + 1) / 2;                // 8-byte aligned past the 2 nf + 1 ints
  if (tid == 0) {
    bad = 0;
    int acc = 0, mm = 0;
    for (int k = 0; k < nf; ++k) { const int l = band_last[f0 + k] - f0; lastk[k] = l; coff[k] = acc; acc += l - k + 1; mm = max(mm, l - k); }
    coff[nf] = acc; env_total = acc; m_max = mm;
  }
  __syncthreads();
  const int nenv = env_total, mmax = m_max;
  double* A = base;                                          // [nenv * 36] column envelopes, block (k + ii, k) at coff[k] + ii
  double* LI = A + (size_t)nenv * 36;                        // [nf * 36] inverse diagonal factors
  double* y = LI + (size_t)nf * 36;                          // [n]
  uint16_t* ptab = reinterpret_cast<uint16_t*>(y + n + 8);   // [mmax (mmax + 1) / 2] trailing-update pair p -> ii | jj << 8
  for (int i = tid; i < nenv * 36; i += T) A[i] = 0.0;
  for (int i = tid; i < n; i += T) y[i] = d.rhs[(size_t)f0 * 6 + i];
  for (int ii = tid; ii < mmax; ii += T)
    for (int jj = 0; jj <= ii; ++jj) ptab[ii * (ii + 1) / 2 + jj] = (uint16_t)(ii | (jj << 8));
  __syncthreads();
  {
    // upper block (r, c), r <= c, of S  ->  lower block (c, r) = its transpose, in column r
    const int b0 = d.diag_blk[f0];
    const int b1 = (f1 < d.Nf) ? d.diag_blk[f1] : d.nbu;
    for (int i = tid; i < (b1 - b0) * 36; i += T) {
      const int b = b0 + i / 36, e = i % 36, er = e / 6, ec = e - 6 * er;
      const int r = d.blk_row[b] - f0, c = d.blk_col[b] - f0;
      A[(size_t)(coff[r] + (c - r)) * 36 + 6 * ec + er] = d.Su[(size_t)b * 36 + e];
    }
  }
  __syncthreads();
  // 6x6 Cholesky + inverse of the diagonal block of column k (warp 0)
  auto factor_column = [&](int k) {       // lane 0 of warp 0
    if (lane == 0) {
      double li[36];
      if (!chol6_inverse_factor_fast(A + (size_t)coff[k] * 36, li)) bad = 1;
#pragma unroll
      for (int i = 0; i < 36; ++i) LI[(size_t)k * 36 + i] = li[i];
    }
  };
  lap(0);
  if (warp == 0) factor_column(0);
  __syncthreads();
  lap(2);
  for (int k = 0; k < nf; ++k) {
    const int m = lastk[k] - k;
    double* colk = A + (size_t)coff[k] * 36;
    const double* li = LI + (size_t)k * 36;
    // (b) panel, a thread per block row (it reads its six entries before it writes them: in place), and the
    //     right-hand side block by the last six threads (one warp: the barrier between reads and writes is a warp's)
    for (int idx = tid; idx < m * 6; idx += T) {
      double* row = colk + 36 + (size_t)(idx / 6) * 36 + 6 * (idx % 6);
      double av[6], o[6];
#pragma unroll
      for (int t = 0; t < 6; ++t) av[t] = row[t];
#pragma unroll
      for (int c = 0; c < 6; ++c) {
        double acc = 0.0;
#pragma unroll
        for (int t = 0; t < 6; ++t) if (t <= c) acc += av[t] * li[6 * c + t];
        o[c] = acc;
      }
#pragma unroll
      for (int t = 0; t < 6; ++t) row[t] = o[t];
    }
    if (warp == T / 32 - 1) {
      const int c = lane - 26;                               // lanes 26..31
      double yk = 0.0;
      if (c >= 0) {
#pragma unroll
        for (int t = 0; t < 6; ++t) if (t <= c) yk += li[6 * c + t] * y[6 * k + t];
      }
      __syncwarp();
      if (c >= 0) y[6 * k + c] = yk;
    }
    __syncthreads();
    lap(1);
    // (c) trailing update inside the envelope: block (k + 1 + ii, k + 1 + jj), 0 <= jj <= ii < m, lives in column
    //     k + 1 + jj.  Look-ahead: warp 0 takes pair 0 = the next diagonal block first and factors it straight away,
    //     while the other warps run the rest of the update.
    const int npair = m * (m + 1) / 2;
    if (warp == 0) {
      if (m > 0) {
        const double* L1 = colk + 36;
        double* dst = A + (size_t)coff[k + 1] * 36;
#pragma unroll
        for (int pass = 0; pass < 2; ++pass) {
          const int e = lane + 32 * pass;
          if (e < 36) {
            const int r = e / 6, c = e - 6 * r;
            const double* Li = L1 + 6 * r; const double* Lj = L1 + 6 * c;
            dst[e] -= (Li[0] * Lj[0] + Li[1] * Lj[1] + Li[2] * Lj[2]) + (Li[3] * Lj[3] + Li[4] * Lj[4] + Li[5] * Lj[5]);
          }
        }
        __syncwarp();
      }
      if (k + 1 < nf) factor_column(k + 1);
      lap(2);
    } else {
      for (int idx = tid - 32; idx < (npair - 1) * 36; idx += T - 32) {
        const int p = 1 + idx / 36, e = idx - 36 * (p - 1), r = e / 6, c = e - 6 * r;
        const unsigned pr = ptab[p];
        const int ii = pr & 255u, jj = pr >> 8;
        const double* Li = colk + (size_t)(1 + ii) * 36 + 6 * r;
        const double* Lj = colk + (size_t)(1 + jj) * 36 + 6 * c;
        const double acc = (Li[0] * Lj[0] + Li[1] * Lj[1] + Li[2] * Lj[2]) + (Li[3] * Lj[3] + Li[4] * Lj[4] + Li[5] * Lj[5]);
        A[(size_t)(coff[k + 1 + jj] + (ii - jj)) * 36 + e] -= acc;
      }
      for (int idx = tid - 32; idx < m * 6; idx += T - 32) {
        const int ii = idx / 6, r = idx - 6 * ii;
        const double* Lr = colk + (size_t)(1 + ii) * 36 + 6 * r;
        double acc = 0.0;
#pragma unroll
        for (int t = 0; t < 6; ++t) acc += Lr[t] * y[6 * k + t];
        y[6 * (k + 1 + ii) + r] -= acc;
      }
    }
    __syncthreads();
    lap(3);
  }
  // back substitution by warp 0, right-looking: as soon as x_k is known it is taken out of every y_j that sees it
  // (the columns j < k whose envelope reaches row k: y_j -= L_kj^T x_k), so that a column's critical path is two 6x6
  // matrix-vector products — x_k = Linv_kk^T y_k, then the update of y_{k-1} — instead of a gather over its whole
  // envelope.  Lane group q (six lanes) updates column k - 1 - q, columns further away in a second pass.
  if (warp == 0) {
    const int ps = lane / 6, c = lane - 6 * ps;
    for (int k = nf - 1; k >= 0; --k) {
      // x_k: lane c < 6 holds s_c = y_k[c]; v_c = sum_{q >= c} Linv[q][c] s_q
      const double s = (lane < 6) ? y[6 * k + lane] : 0.0;
      double v = 0.0;
#pragma unroll
      for (int q = 0; q < 6; ++q) {
        const double sq = __shfl_sync(0xffffffffu, s, q);
        if (lane < 6 && q >= lane) v += LI[(size_t)k * 36 + 6 * q + lane] * sq;
      }
      if (lane < 6) y[6 * k + lane] = v;
      double xk[6];
#pragma unroll
      for (int q = 0; q < 6; ++q) xk[q] = __shfl_sync(0xffffffffu, v, q);
      // y_j -= L_kj^T x_k for j = k - 1 - jj, jj = ps, ps + 5, ...; block (k, j) sits in column j at row offset k - j
      if (lane < 30)
        for (int jj = ps; jj < min(k, mmax); jj += 5) {
          const int j = k - 1 - jj;
          if (lastk[j] < k) continue;                                       // column j's envelope ends above row k
          const double* L = A + (size_t)(coff[j] + (k - j)) * 36 + c;       // column c of L_kj = row c of its transpose
          y[6 * j + c] -= (L[0] * xk[0] + L[6] * xk[1] + L[12] * xk[2]) + (L[18] * xk[3] + L[24] * xk[4] + L[30] * xk[5]);
        }
      __syncwarp();
    }
  }
  __syncthreads();
  lap(4);
  const bool fail = bad != 0;
  for (int i = tid; i < n; i += T) {
    const double v = y[i];
    d.x[(size_t)f0 * 6 + i] = (fail || !isfinite(v)) ? 0.0 : v;
    d.r[(size_t)f0 * 6 + i] = 0.0;                         // exact solve: no linear residual in the model change
    if (!isfinite(v)) bad = 1;                             // benign race: every writer stores 1
  }
  __syncthreads();
  if (tid == 0 && bad != 0) d.w_fail[w] = 1;
  lap(5);
  if (prof) printf("[band] nf %d  setup %llu  panel+barrier %llu  diag update+factor (warp 0) %llu  wait for the update %llu  back substitution %llu  write %llu ns\n", nf, tq[0], tq[1], tq[2], tq[3], tq[4], tq[5]);
}

// The same factorisation for a mid-size window (a 50-keyframe sliding window: 294 unknowns, 1225 lower blocks =
// 353 KB) on one thread-block CLUSTER of kCholCluster CTAs: block column j lives in the shared memory of CTA
// j % kCholCluster, so every trailing update is a local shared-memory update.  Per block column k:
//   owner        factors the diagonal block, forms the panel L_ik (i > k) and y_k in its `pan` buffer
//   cluster barrier (ONE per column; `pan` is double-buffered, the owner rotates)
//   every CTA    pulls the panel through distributed shared memory into its local copy and updates its own
//                columns j > k (A_ij -= L_ik L_jk^T, y_j -= L_jk y_k)
// Back substitution: the owner of column k holds all L_ik (i > k): x_k = Linv_kk^T (y_k - sum L_ik^T x_i), pushed
// to every CTA's copy of x, one cluster barrier per column.  Fixed order throughout: bit-reproducible.

constexpr int kCholCluster = 8;
constexpr int kCholClThreads = 256;

__host__ __device__ __forceinline__ int chol_cl_cols(int nf, int rank) { return (nf > rank) ? (nf - rank + kCholCluster - 1) / kCholCluster : 0; }
// blocks of the first lj local columns of CTA `rank` (column j = rank + kCholCluster * l holds rows j..nf-1)
__host__ __device__ __forceinline__ int chol_cl_coloff(int nf, int rank, int lj) {
  return lj * (nf - rank) - kCholCluster * (lj * (lj - 1) / 2);
}
__host__ __device__ __forceinline__ size_t chol_cluster_smem_bytes(int nf) {
  const int maxcols = chol_cl_cols(nf, 0);
  const size_t blocks = (size_t)chol_cl_coloff(nf, 0, maxcols);
  // own columns, pan[2] (panel + y_k), local panel copy (+ y_k), own inverse factors, own y, x, reduction scratch
  return (blocks * 36 + 2 * ((size_t)nf * 36 + 8) + 2 * ((size_t)nf * 36 + 8) + (size_t)maxcols * 36 + (size_t)maxcols * 6 +
          (size_t)nf * 6 + 64 + kCholClThreads) * 8;
}

__global__ void __launch_bounds__(kCholClThreads, 1) k_chol_cluster(Dev d, int profile) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ double smem[];
  constexpr int CL = kCholCluster, T = kCholClThreads;
  const int rank = (int)cluster.block_rank();
  const int w = blockIdx.x / CL;
  const int f0 = d.win_fc_off[w], f1 = d.win_fc_off[w + 1];
  const int nf = f1 - f0, n = 6 * nf;
  if (rank == 0 && threadIdx.x == 0) d.w_pcg_it[w] = 0;
  if (nf <= 0) return;                                      // uniform over the cluster
  if (d.w_status[w] != 0 || d.w_fail[w] != 0) return;       // idem; leaves x untouched (unused)
  const int tid = threadIdx.x;
  const int my_cols = chol_cl_cols(nf, rank), maxcols = chol_cl_cols(nf, 0);
  const int pstride = nf * 36 + 8;                          // panel rows (indexed by i) + y_k at [nf * 36 .. + 6)
  double* Acol = smem;                                                       // own columns
  double* pan = Acol + (size_t)chol_cl_coloff(nf, 0, maxcols) * 36;         // [2][pstride] written by the owner, pulled by all
  double* lpb = pan + 2 * (size_t)pstride;                                   // [2][pstride] local copies of the panels (k & 1)
  double* LIm = lpb + 2 * (size_t)pstride;                                                // [maxcols * 36] inverse diagonal factors (own columns)
  double* ym = LIm + (size_t)maxcols * 36;                                   // [maxcols * 6]  right-hand side of the own columns
  double* xall = ym + (size_t)maxcols * 6;                                   // [n]            solution, every CTA has all of it
  double* red = xall + n;                                                    // [64 + T]
  __shared__ int bad;
  if (tid == 0) bad = 0;
  const int myblocks = chol_cl_coloff(nf, rank, my_cols);
  for (int i = tid; i < myblocks * 36; i += T) Acol[i] = 0.0;
  for (int i = tid; i < my_cols * 6; i += T) { const int lj = i / 6; ym[i] = d.rhs[(size_t)(f0 + rank + CL * lj) * 6 + (i - 6 * lj)]; }
  __syncthreads();
  {
    // upper block (r, c), r <= c, of S -> lower block (c, r) = its transpose, kept by the owner of column r
    const int b0 = d.diag_blk[f0];
    const int b1 = (f1 < d.Nf) ? d.diag_blk[f1] : d.nbu;
    for (int i = tid; i < (b1 - b0) * 36; i += T) {
      const int b = b0 + i / 36, e = i % 36, er = e / 6, ec = e - 6 * er;
      const int r = d.blk_row[b] - f0, c = d.blk_col[b] - f0;
      if (r % CL != rank) continue;
      Acol[(size_t)(chol_cl_coloff(nf, rank, r / CL) + (c - r)) * 36 + 6 * ec + er] = d.Su[(size_t)b * 36 + e];
    }
  }
  cluster.sync();          // every CTA of the cluster is running (and initialised) before anybody reads a peer's shared memory
  // profile != 0 (SBA_DEBUG_LS=2): thread 0 of CTA 1 prints the nanoseconds per phase, summed over the columns (tuning aid)
  const bool prof = profile != 0 && rank == 1 && tid == 0 && w == 0;
  unsigned long long t0 = prof ? gtimer() : 0, tp[6] = {0, 0, 0, 0, 0, 0};
  auto lap = [&](int i) { if (prof) { const unsigned long long t = gtimer(); tp[i] += t - t0; t0 = t; } };
  // factor the diagonal block of the own column k and form its panel + y_k in pan[k & 1] (whole CTA, owner only)
  auto factor_and_panel = [&](int k) {
    const int lk = k / CL, m = nf - 1 - k;
    double* pb = pan + (size_t)(k & 1) * pstride;
    // the owner of column k + 1 is on the critical path: it gets the panel pushed into its local copy, the others pull
    double* nx = (k + 1 < nf) ? cluster.map_shared_rank(lpb + (size_t)(k & 1) * pstride, (k + 1) % CL) : pb;
    double* colk = Acol + (size_t)chol_cl_coloff(nf, rank, lk) * 36;       // rows k..nf-1 of column k
    if (tid == 0) {
      double li[36];
      if (!chol6_inverse_factor(colk, li)) bad = 1;
#pragma unroll
      for (int i = 0; i < 36; ++i) LIm[(size_t)lk * 36 + i] = li[i];
    }
    __syncthreads();
    const double* li = LIm + (size_t)lk * 36;
    for (int idx = tid; idx < m * 36; idx += T) {
      const int ii = 1 + idx / 36, e = idx % 36, r = e / 6, c = e - 6 * r;
      const double* blk = colk + (size_t)ii * 36 + 6 * r;
      double acc = 0.0;
#pragma unroll
      for (int t = 0; t < 6; ++t) if (t <= c) acc += blk[t] * li[6 * c + t];
      pb[(size_t)(k + ii) * 36 + e] = acc;
      nx[(size_t)(k + ii) * 36 + e] = acc;
    }
    if (tid >= T - 6) {
      const int c = tid - (T - 6);
      double acc = 0.0;
#pragma unroll
      for (int t = 0; t < 6; ++t) if (t <= c) acc += li[6 * c + t] * ym[6 * lk + t];
      pb[(size_t)nf * 36 + c] = acc;
      nx[(size_t)nf * 36 + c] = acc;
    }
  };
  // A_ij -= L_ik L_jk^T (i >= j) and y_j -= L_jk y_k for the own local column lj, from the local panel copy:
  // one thread per (block, row)
  auto update_column = [&](int lj, const double* lp) {
    const int j = rank + CL * lj;
    const double* yk = lp + (size_t)nf * 36;
    double* colj = Acol + (size_t)chol_cl_coloff(nf, rank, lj) * 36;
    const double2* Lj = reinterpret_cast<const double2*>(lp + (size_t)j * 36);
    for (int it = tid; it < (nf - j) * 6; it += T) {
      const int ii = it / 6, r = it - 6 * ii;
      const double2* Li = reinterpret_cast<const double2*>(lp + (size_t)(j + ii) * 36 + 6 * r);
      const double2 l0 = Li[0], l1 = Li[1], l2 = Li[2];
      double2* dst = reinterpret_cast<double2*>(colj + (size_t)ii * 36 + 6 * r);
#pragma unroll
      for (int h = 0; h < 3; ++h) {
        const double2 a0 = Lj[6 * h], a1 = Lj[6 * h + 1], a2 = Lj[6 * h + 2];
        const double2 b0 = Lj[6 * h + 3], b1 = Lj[6 * h + 4], b2 = Lj[6 * h + 5];
        double2 v = dst[h];
        v.x -= (l0.x * a0.x + l0.y * a0.y + l1.x * a1.x) + (l1.y * a1.y + l2.x * a2.x + l2.y * a2.y);
        v.y -= (l0.x * b0.x + l0.y * b0.y + l1.x * b1.x) + (l1.y * b1.y + l2.x * b2.x + l2.y * b2.y);
        dst[h] = v;
      }
    }
    if (tid >= T - 6) {
      const int r = tid - (T - 6);
      const double* Lr = lp + (size_t)j * 36 + 6 * r;
      double acc = 0.0;
#pragma unroll
      for (int t = 0; t < 6; ++t) acc += Lr[t] * yk[t];
      ym[6 * lj + r] -= acc;
    }
  };
  if (rank == 0) factor_and_panel(0);
  lap(0);
  cluster.sync();
  lap(1);
  for (int k = 0; k < nf; ++k) {
    const int owner = k % CL, lk = k / CL, m = nf - 1 - k;
    double* pb = pan + (size_t)(k & 1) * pstride;
    double* lp = lpb + (size_t)(k & 1) * pstride;
    const bool next_owner = (k + 1 < nf) && ((k + 1) % CL == rank);
    // pull the panel (rows k+1..nf-1) and y_k from the owner (the next owner already has them)
    if (!next_owner) {
      const double2* src = reinterpret_cast<const double2*>(cluster.map_shared_rank(pb, owner) + (size_t)(k + 1) * 36);
      double2* dst = reinterpret_cast<double2*>(lp + (size_t)(k + 1) * 36);
      // all remote loads first (independent round trips through the cluster network), then the local stores
      constexpr int NP = ((kCholMaxCams - 1) * 18 + T - 1) / T;
      double2 v[NP];
#pragma unroll
      for (int u = 0; u < NP; ++u) { const int i = tid + u * T; if (i < m * 18) v[u] = src[i]; }
#pragma unroll
      for (int u = 0; u < NP; ++u) { const int i = tid + u * T; if (i < m * 18) dst[i] = v[u]; }
      if (tid < 6) lp[(size_t)nf * 36 + tid] = cluster.map_shared_rank(pb, owner)[(size_t)nf * 36 + tid];
    }
    __syncthreads();
    lap(2);
    // look-ahead: the owner of column k + 1 brings that column up to date first, factors it and forms the next
    // panel while the other CTAs (and then itself) do the rest of the trailing update
    const int first = (k + 1 > rank) ? (k + 1 - rank + CL - 1) / CL : 0;    // first own local column with j > k
    if (next_owner) {
      update_column(first, lp);           // rank + CL * first == k + 1
      __syncthreads();
      factor_and_panel(k + 1);
    }
    lap(0);
    if (rank == owner) {
      // keep L_ik (for the back substitution) and y_k in the own column
      double* colk = Acol + (size_t)chol_cl_coloff(nf, rank, lk) * 36;
      for (int i = tid; i < m * 36; i += T) colk[36 + i] = lp[(size_t)(k + 1) * 36 + i];
      if (tid < 6) ym[6 * lk + tid] = lp[(size_t)nf * 36 + tid];
    }
    for (int lj = first + (next_owner ? 1 : 0); lj < my_cols; ++lj) update_column(lj, lp);
    __syncthreads();
    lap(3);
    cluster.sync();                       // panel k + 1 is in its owner's pan; every CTA is done with lp
    lap(1);
  }
  // Back substitution by CTA 0 alone, columns downwards: x_k = Linv_kk^T (y_k - sum over i > k of L_ik^T x_i).  Column
  // k, its inverse factor and y_k are read straight out of their owner's shared memory; none of that depends on x,
  // so the loads of column k - 1 are issued before the reduction of column k finishes (registers `nxt`): the
  // critical path per column is two CTA barriers, no cluster barrier.
  cluster.sync();                                            // all columns final
  if (rank == 0) {
    constexpr int G = 42, PER = (kCholMaxCams * 6 + G - 1) / G;   // 42 groups of 6 threads over the (i, r) pairs of a column
    const int c = tid % 6, g = tid / 6;
    double cur[PER], nxt[PER], li_c[6], li_n[6], y_c = 0.0, y_n = 0.0;
    auto fetch = [&](int k, double* col, double* li, double& yv) {
      const int owner = k % CL, lk = k / CL, m = nf - 1 - k;
      const double* colk = cluster.map_shared_rank(Acol, owner) + (size_t)chol_cl_coloff(nf, owner, lk) * 36;
#pragma unroll
      for (int u = 0; u < PER; ++u) {
        const int q = g + G * u;
        col[u] = (g < G && q < m * 6) ? colk[(size_t)(1 + q / 6) * 36 + 6 * (q % 6) + c] : 0.0;
      }
      if (tid < 6) {
        const double* lik = cluster.map_shared_rank(LIm, owner) + (size_t)lk * 36;
#pragma unroll
        for (int t = 0; t < 6; ++t) li[t] = lik[6 * t + tid];          // column tid of Linv = row tid of Linv^T
        yv = cluster.map_shared_rank(ym, owner)[6 * lk + tid];
      }
    };
    fetch(nf - 1, cur, li_c, y_c);
    for (int k = nf - 1; k >= 0; --k) {
      if (k > 0) fetch(k - 1, nxt, li_n, y_n);
      const int m = nf - 1 - k;
      double acc = 0.0;
#pragma unroll
      for (int u = 0; u < PER; ++u) {
        const int q = g + G * u;
        if (g < G && q < m * 6) acc += cur[u] * xall[6 * (k + 1 + q / 6) + (q % 6)];
      }
      if (g < G) red[64 + tid] = acc;
      __syncthreads();
      if (tid < 6) {
        double s_ = 0.0;
        for (int gg = 0; gg < G; ++gg) s_ += red[64 + 6 * gg + tid];
        red[tid] = y_c - s_;
      }
      __syncthreads();
      if (tid < 6) {
        double v = 0.0;
#pragma unroll
        for (int t = 0; t < 6; ++t) if (t >= tid) v += li_c[t] * red[t];
        xall[6 * k + tid] = v;
      }
      __syncthreads();
#pragma unroll
      for (int u = 0; u < PER; ++u) cur[u] = nxt[u];
#pragma unroll
      for (int t = 0; t < 6; ++t) li_c[t] = li_n[t];
      y_c = y_n;
    }
  }
  cluster.sync();                                            // peers keep their shared memory alive until CTA 0 is done
  lap(4);
  if (prof) printf("[chol] nf %d  own factor+panel %llu  barrier %llu  pull %llu  trailing %llu  back substitution (CTA 0) %llu ns\n", nf, tp[0], tp[1], tp[2], tp[3], tp[4]);
  if (bad != 0) d.w_fail[w] = 1;                             // benign race: every writer stores 1
  if (rank == 0) {
    for (int i = tid; i < n; i += T) {
      const double v = xall[i];
      d.x[(size_t)f0 * 6 + i] = isfinite(v) ? v : 0.0;
      d.r[(size_t)f0 * 6 + i] = 0.0;                         // exact solve: no linear residual in the model change
      if (!isfinite(v)) d.w_fail[w] = 1;
    }
  }
}

// The same solve for windows whose reduced system is too big for one SM's shared-memory bandwidth and issue
// rate (a 50-keyframe sliding window: ~600 blocks of S re-read every iteration): one thread-block CLUSTER of
// kPcgCluster CTAs per window.  Block-rows are dealt round-robin to the CTAs; S, Minv and the pattern are read
// through L1 (each CTA's slice stays resident), the vectors live in shared memory.  Per iteration there are two
// cluster barriers and one CTA barrier:
//   A  Ap = S p on the CTA's rows; every warp leader stores its p.Ap partial into ALL CTAs' reduction buffers
//      through distributed shared memory                                            -> cluster barrier
//   B  alpha; x, r, z = Minv r on the CTA's rows; z and the r.z partials are stored into all CTAs  -> cluster barrier
//   C  beta; every CTA forms the whole p = z + beta p locally                       -> CTA barrier
// Partials are summed in (CTA, warp) order by every thread: all CTAs hold bit-identical scalars.
constexpr int kPcgCluster = 8;
constexpr int kPcgClThreads = 256;
constexpr int kPcgClSlots = kPcgCluster * (kPcgClThreads / 32);   // 64 partials per reduction

// rows_cap / ent_cap: most block-rows / block entries any CTA owns (0, 0: S and Minv stay in global memory)
__host__ __device__ __forceinline__ size_t pcg_cluster_smem_bytes(int nf, int rows_cap, int ent_cap) {
  return ((size_t)6 * 6 * nf + 2 * kPcgClSlots + (size_t)36 * rows_cap + (size_t)36 * ent_cap) * 8 +   // x r0 r1 z p Ap, red, Minv, S
         ((size_t)rows_cap + 1 + (size_t)ent_cap) * 4;
}

__device__ __forceinline__ double pcg_cluster_sum(const double* red, int lane) {
  double t = red[lane] + red[lane + 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  return t;
}

__global__ void __launch_bounds__(kPcgClThreads, 1) k_pcg_cluster(Dev d, DevParams P, int max_it, int rows_cap, int ent_cap) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ double smem[];
  const int rank = (int)cluster.block_rank();
  const int w = blockIdx.x / kPcgCluster;
  const int f0 = d.win_fc_off[w], f1 = d.win_fc_off[w + 1];
  const int nf = f1 - f0, n = 6 * nf;
  if (rank == 0 && threadIdx.x == 0) d.w_pcg_it[w] = 0;
  if (nf <= 0) return;                                      // uniform over the cluster
  if (d.w_status[w] != 0 || d.w_fail[w] != 0) return;       // idem; leaves x untouched (unused)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int NW = kPcgClThreads / 32;
  double* x = smem; double* ra = x + n; double* rb = ra + n; double* z = rb + n; double* pv = z + n; double* Ap = pv + n;
  double* redA = Ap + n; double* redB = redA + kPcgClSlots;
  const double* __restrict__ S = d.Su;
  const double* __restrict__ M = d.Minv + (size_t)f0 * 36;
  // rows of this CTA: local block-rows lf = rank, rank + CL, ...; its scalar rows are dealt to threads
  const int my_rows = (nf > rank) ? (nf - rank + kPcgCluster - 1) / kPcgCluster : 0;
  for (int i = threadIdx.x; i < n; i += kPcgClThreads) { x[i] = 0.0; ra[i] = d.rhs[(size_t)f0 * 6 + i]; }
  // The CTA's slice of S, one 6x6 copy per pattern entry of its rows (transposed blocks already turned), and of
  // Minv: cluster.sync invalidates L1, so what is re-read every iteration has to sit in shared memory.
  const bool staged = ent_cap > 0;
  double* Ms = redB + kPcgClSlots;                           // [rows_cap * 36]
  double* Ss = Ms + (size_t)rows_cap * 36;                   // [ent_cap * 36]
  int* lrp = reinterpret_cast<int*>(Ss + (size_t)ent_cap * 36);   // [rows_cap + 1]
  int* lcol = lrp + rows_cap + 1;                            // [ent_cap] 6 * local block-column
  if (staged) {
    if (threadIdx.x == 0) {
      int acc = 0;
      for (int m = 0; m < my_rows; ++m) {
        const int f = f0 + rank + m * kPcgCluster;
        lrp[m] = acc;
        acc += d.row_ptr[f + 1] - d.row_ptr[f];
      }
      lrp[my_rows] = acc;
    }
    __syncthreads();
    for (int m = 0; m < my_rows; ++m) {
      const int f = f0 + rank + m * kPcgCluster;
      const int g0 = d.row_ptr[f], cnt = d.row_ptr[f + 1] - g0, l0 = lrp[m];
      for (int i = threadIdx.x; i < cnt; i += kPcgClThreads) lcol[l0 + i] = 6 * (d.bsr_col[g0 + i] - f0);
      for (int i = threadIdx.x; i < cnt * 36; i += kPcgClThreads) {
        const int e = i / 36, k = i - 36 * e, rr = k / 6, cc = k - 6 * rr;
        const int bi = d.bsr_blk[g0 + e];
        Ss[(size_t)(l0 + e) * 36 + k] = (bi >= 0) ? S[(size_t)bi * 36 + k] : S[(size_t)(~bi) * 36 + 6 * cc + rr];
      }
      for (int i = threadIdx.x; i < 36; i += kPcgClThreads) Ms[m * 36 + i] = M[(size_t)(rank + m * kPcgCluster) * 36 + i];
    }
  }
  cluster.sync();          // every CTA of the cluster is running before anybody writes into a peer's shared memory
  double* r = ra; double* rn = rb;
  // z = Minv r, p = z over the WHOLE vector in every CTA (cheap, saves a broadcast); rz from the own rows only
  double part = 0.0;
  for (int i = threadIdx.x; i < n; i += kPcgClThreads) {
    const int f = i / 6, k = i - 6 * f;
    double acc = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) acc += __ldg(M + (size_t)f * 36 + 6 * k + j) * r[6 * f + j];
    z[i] = acc; pv[i] = acc;
    if (f % kPcgCluster == rank) part += r[i] * acc;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if (lane < kPcgCluster) cluster.map_shared_rank(redB, lane)[rank * NW + warp] = part;
  cluster.sync();
  double rz = pcg_cluster_sum(redB, lane);
  const double rz0 = rz;
  int it = 0;
  bool fail = false;
  const int ps = lane / 6, ri = lane - 6 * ps;
  if (!(rz > 0.0)) { if (!(rz == 0.0)) fail = true; }      // rhs == 0: x = 0 is the solution; NaN: fail
  else {
    for (;;) {
      // ---- A: Ap = S p on the own block-rows, p.Ap partials to everybody
      part = 0.0;
      for (int m = warp; m < my_rows; m += NW) {
        const int lf = rank + m * kPcgCluster;
        double acc = 0.0;
        if (staged && lane < 30) {
          const int j1 = lrp[m + 1];
          for (int j = lrp[m] + ps; j < j1; j += 5) {
            const double* xv = pv + lcol[j];
            const double* B = Ss + (size_t)j * 36 + 6 * ri;
            acc += (B[0] * xv[0] + B[1] * xv[1] + B[2] * xv[2]) + (B[3] * xv[3] + B[4] * xv[4] + B[5] * xv[5]);
          }
        } else if (lane < 30) {
          const int j1 = __ldg(d.row_ptr + f0 + lf + 1);
          for (int j = __ldg(d.row_ptr + f0 + lf) + ps; j < j1; j += 5) {
            const double* xv = pv + 6 * (__ldg(d.bsr_col + j) - f0);
            const int bi = __ldg(d.bsr_blk + j);
            if (bi >= 0) {
              const double* B = S + (size_t)bi * 36 + 6 * ri;
              acc += (__ldg(B) * xv[0] + __ldg(B + 1) * xv[1] + __ldg(B + 2) * xv[2]) +
                     (__ldg(B + 3) * xv[3] + __ldg(B + 4) * xv[4] + __ldg(B + 5) * xv[5]);
            } else {
              const double* B = S + (size_t)(~bi) * 36 + ri;
              acc += (__ldg(B) * xv[0] + __ldg(B + 6) * xv[1] + __ldg(B + 12) * xv[2]) +
                     (__ldg(B + 18) * xv[3] + __ldg(B + 24) * xv[4] + __ldg(B + 30) * xv[5]);
            }
          }
        }
        double tot = acc;                                  // rows of the five entry groups, in lane order
#pragma unroll
        for (int g = 1; g < 5; ++g) tot += __shfl_sync(0xffffffffu, acc, (lane + 6 * g) & 31);
        if (lane < 6) { Ap[6 * lf + lane] = tot; part += pv[6 * lf + lane] * tot; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      if (lane < kPcgCluster) cluster.map_shared_rank(redA, lane)[rank * NW + warp] = part;
      cluster.sync();
      const double pAp = pcg_cluster_sum(redA, lane);
      if (!(pAp > 0.0)) { fail = true; break; }             // S not SPD
      const double alpha = rz / pAp;
      // ---- B: x, r, z on the own rows (the warp that wrote Ap[row] is not the thread that reads it: CTA barrier
      //         is implied by cluster.sync above); z and the r.z partials to everybody
      part = 0.0;
      for (int t = threadIdx.x; t < my_rows * 6; t += kPcgClThreads) {
        const int m = t / 6, k = t - 6 * m;
        const int f = rank + m * kPcgCluster, i = 6 * f + k;
        x[i] += alpha * pv[i];
        double acc = 0.0, mine = 0.0;
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          const double rj = r[6 * f + j] - alpha * Ap[6 * f + j];
          if (j == k) mine = rj;
          acc += (staged ? Ms[m * 36 + 6 * k + j] : __ldg(M + (size_t)f * 36 + 6 * k + j)) * rj;
        }
        rn[i] = mine;
        part += mine * acc;
#pragma unroll
        for (int c = 0; c < kPcgCluster; ++c) cluster.map_shared_rank(z, c)[i] = acc;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      if (lane < kPcgCluster) cluster.map_shared_rank(redB, lane)[rank * NW + warp] = part;
      cluster.sync();
      const double rz_new = pcg_cluster_sum(redB, lane);
      { double* t = r; r = rn; rn = t; }
      ++it;
      if (!isfinite(rz_new)) { fail = true; break; }
      if (rz_new <= P.pcg_tol2 * rz0 || it >= max_it) { rz = rz_new; break; }
      const double beta = rz_new / rz;
      rz = rz_new;
      // ---- C: the whole p in every CTA
      for (int i = threadIdx.x; i < n; i += kPcgClThreads) pv[i] = z[i] + beta * pv[i];
      __syncthreads();
    }
  }
  cluster.sync();          // nobody leaves while a peer may still write into its shared memory
  for (int t = threadIdx.x; t < my_rows * 6; t += kPcgClThreads) {
    const int m = t / 6, k = t - 6 * m;
    const int i = 6 * (rank + m * kPcgCluster) + k;
    d.x[(size_t)f0 * 6 + i] = x[i]; d.r[(size_t)f0 * 6 + i] = r[i];
  }
  if (rank == 0 && threadIdx.x == 0) { d.w_pcg_it[w] = it; if (fail) d.w_fail[w] = 1; atomicMax(d.counters + 4, it); }   // control-plane counter
}

// Global problems (one window, thousands of free cameras): the whole PCG solve in ONE cooperative launch over
// the block-CSR in global memory, two grid barriers per iteration.
//   A  warp per block-row: Ap = S p with p = z + beta p_old formed on the fly from the previous iteration's
//      vectors (the row's own p goes to the other p buffer: no barrier between "update p" and "use p");
//      per-row p.Ap                                                                  -> grid barrier
//      every CTA sums the per-row dots in the same fixed order -> the same alpha everywhere, no broadcast
//   B  warp per block-row: x += alpha p, r -= alpha Ap, z = Minv r, per-row r.z     -> grid barrier
// counters[4] returns the iteration count.  Vectors: d.x, d.r, d.z, d.Ap, p ping-pong in d.p / pbuf.
__device__ __forceinline__ double coop_sum(const double* v, int n, double* sm /* [4] */) {
  double s[1] = {0.0};
  for (int i = threadIdx.x; i < n; i += kThreads) s[0] += __ldcg(v + i);
  const double t = block_sum<1>(s, sm);
  __shared__ double bc;
  if (threadIdx.x == 0) bc = t;
  __syncthreads();
  return bc;
}

__global__ void __launch_bounds__(kThreads) k_pcg_coop(Dev d, DevParams P, int max_it, double* pbuf, double* rowdot) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  __shared__ double sm[4];
  const int lane = threadIdx.x & 31;
  const int gw = (blockIdx.x * kThreads + threadIdx.x) >> 5, nw = (gridDim.x * kThreads) >> 5;
  const int nf = d.Nf;
  // Two per-row dot buffers: a CTA that is through with a reduction may start writing the NEXT phase's dots while a
  // slower CTA is still summing this phase's, so consecutive phases must not share a buffer.
  double* dotA = rowdot; double* dotB = rowdot + nf;
  if (d.w_status[0] != 0 || d.w_fail[0] != 0) { if (blockIdx.x == 0 && threadIdx.x == 0) d.w_pcg_it[0] = 0; return; }
  const int ps = lane / 6, ri = lane - 6 * ps;
  // init: x = 0, r = rhs, z = Minv r, p_old = 0 (so that p = z + beta * 0 in the first mat-vec), per-row r.z
  for (int f = gw; f < nf; f += nw) {
    double rj[6], z = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) rj[j] = d.rhs[(size_t)f * 6 + j];
    if (lane < 6) {
#pragma unroll
      for (int j = 0; j < 6; ++j) z += d.Minv[(size_t)f * 36 + 6 * lane + j] * rj[j];
      d.x[(size_t)f * 6 + lane] = 0.0; d.r[(size_t)f * 6 + lane] = rj[lane]; d.z[(size_t)f * 6 + lane] = z;
      d.p[(size_t)f * 6 + lane] = 0.0; pbuf[(size_t)f * 6 + lane] = 0.0;
    }
    double dot = (lane < 6) ? rj[lane < 6 ? lane : 0] * z : 0.0;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);   // lanes 0..7 (6, 7 hold zero)
    if (lane == 0) dotB[f] = dot;
  }
  grid.sync();
  double rz = coop_sum(dotB, nf, sm);
  const double rz0 = rz;
  int it = 0;
  bool fail = false;
  double beta = 0.0;
  double* pold = d.p; double* pnew = pbuf;
  if (!(rz > 0.0)) { if (!(rz == 0.0)) fail = true; }
  else {
    for (;;) {
      // ---- A
      for (int f = gw; f < nf; f += nw) {
        double acc = 0.0;
        if (lane < 30) {
          const int j1 = d.row_ptr[f + 1];
          for (int j = d.row_ptr[f] + ps; j < j1; j += 5) {
            const int col = __ldg(d.bsr_col + j), bi = __ldg(d.bsr_blk + j);
            double xv[6];
#pragma unroll
            for (int c = 0; c < 6; ++c) xv[c] = __ldcg(d.z + (size_t)col * 6 + c) + beta * __ldcg(pold + (size_t)col * 6 + c);
            if (bi >= 0) {
              const double* B = d.Su + (size_t)bi * 36 + 6 * ri;
              acc += (__ldg(B) * xv[0] + __ldg(B + 1) * xv[1] + __ldg(B + 2) * xv[2]) + (__ldg(B + 3) * xv[3] + __ldg(B + 4) * xv[4] + __ldg(B + 5) * xv[5]);
            } else {
              const double* B = d.Su + (size_t)(~bi) * 36 + ri;
              acc += (__ldg(B) * xv[0] + __ldg(B + 6) * xv[1] + __ldg(B + 12) * xv[2]) + (__ldg(B + 18) * xv[3] + __ldg(B + 24) * xv[4] + __ldg(B + 30) * xv[5]);
            }
          }
        }
        double tot = acc;
#pragma unroll
        for (int g = 1; g < 5; ++g) tot += __shfl_sync(0xffffffffu, acc, (lane + 6 * g) & 31);
        double dot = 0.0;
        if (lane < 6) {
          const double pv = __ldcg(d.z + (size_t)f * 6 + lane) + beta * __ldcg(pold + (size_t)f * 6 + lane);
          pnew[(size_t)f * 6 + lane] = pv;
          d.Ap[(size_t)f * 6 + lane] = tot;
          dot = pv * tot;
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
        if (lane == 0) dotA[f] = dot;
      }
      grid.sync();
      const double pAp = coop_sum(dotA, nf, sm);
      if (!(pAp > 0.0)) { fail = true; break; }
      const double alpha = rz / pAp;
      // ---- B (the warp that owns a row in A owns it here too: its own pnew / Ap writes are visible to it)
      for (int f = gw; f < nf; f += nw) {
        double rj[6];
#pragma unroll
        for (int j = 0; j < 6; ++j) rj[j] = __ldcg(d.r + (size_t)f * 6 + j) - alpha * __ldcg(d.Ap + (size_t)f * 6 + j);
        double dot = 0.0;
        __syncwarp();                      // every lane has read the old r before lanes 0..5 overwrite it
        if (lane < 6) {
          double z = 0.0;
#pragma unroll
          for (int j = 0; j < 6; ++j) z += d.Minv[(size_t)f * 36 + 6 * lane + j] * rj[j];
          d.x[(size_t)f * 6 + lane] += alpha * __ldcg(pnew + (size_t)f * 6 + lane);
          double mine = 0.0;
#pragma unroll
          for (int j = 0; j < 6; ++j) if (j == lane) mine = rj[j];
          d.r[(size_t)f * 6 + lane] = mine; d.z[(size_t)f * 6 + lane] = z;
          dot = mine * z;
        }
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, o);
        if (lane == 0) dotB[f] = dot;
      }
      grid.sync();
      const double rz_new = coop_sum(dotB, nf, sm);
      { double* t = pold; pold = pnew; pnew = t; }
      ++it;
      if (!isfinite(rz_new)) { fail = true; break; }
      if (rz_new <= P.pcg_tol2 * rz0 || it >= max_it) { rz = rz_new; break; }
      beta = rz_new / rz;
      rz = rz_new;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) { d.w_pcg_it[0] = it; if (fail) d.w_fail[0] = 1; d.counters[4] = it; }
}

// ------------------------------------------------------------------------- K5: back-substitution
// dp = -Vinv (gp + sum_obs W_e^T dc[cam(e)])   (stored; the candidate is formed by the line-search kernels).
// A thread per OBSERVATION forms W_e^T dc (its row of W is one independent load; the camera steps of the chunk's slots
// are staged in shared memory meanwhile), then a thread per point adds its observations in order.  A thread per point
// walking its observations paid three dependent L2 round trips per observation.
// chunk partials: g.dp, dp^T D dp, |x|^2, max|dp|
constexpr int kBackObs = 256;       // observations of a regular chunk (<= 240, structure.cpp)
constexpr int kBackSlots = 64;      // camera slots staged; chunks with more read the camera steps through L2
// TMA = true (k_backsub, one CTA per chunk): the chunk's rows of W (contiguous: n x 144 bytes) come to shared memory by
// ONE cp.async.bulk issued by thread 0 before anything else — no load instructions, no staging registers, the copy
// engine keeps HBM busy while the CTA fetches its indices; Ws / bar are the CTA's dynamic buffer and mbarrier.
template <bool TMA = false>
__device__ __forceinline__ void dev_backsub(const Dev& d, const DevParams& P, int chunk, double* Ws = nullptr, uint64_t* bar = nullptr) {
  __shared__ double sm[4 * 3];
  __shared__ double smx[4];
  __shared__ double sobs[kBackObs * 3];
  __shared__ double dcs[kBackSlots * 6];
  const int4 cd0 = __ldg(d.cdesc + 4 * (size_t)chunk), cd1 = __ldg(d.cdesc + 4 * (size_t)chunk + 1);
  const int p0 = cd0.x, np = cd0.y, e0 = cd0.z, n = cd0.w, w = cd1.x, cs0 = cd1.y, ns = cd1.z;
  const bool regular = n <= kBackObs;
  const bool staged = regular && ns <= kBackSlots;
  if constexpr (TMA) {
    if (staged && n > 0 && threadIdx.x == 0) bulk_load(Ws, d.W + (size_t)e0 * 18, (uint32_t)n * 144u, bar);
  }
  if (staged) {
    for (int i = threadIdx.x; i < ns * 6; i += kThreads) {
      const int sl = i / 6;
      const int f = __ldg(d.fc + __ldg(d.cs_cam + cs0 + sl));
      dcs[i] = (f >= 0) ? d.x[(size_t)f * 6 + (i - 6 * sl)] : 0.0;     // a fixed camera's (absent) columns of W are never read
    }
  }
  // this thread's point: everything that depends only on p is requested now, ahead of both barriers
  const int p = p0 + threadIdx.x;
  const bool has_p = (int)threadIdx.x < np;
  double gq[3] = {0, 0, 0}, Vi[6] = {0, 0, 0, 0, 0, 0}, Vd[3] = {0, 0, 0}, xq[3] = {0, 0, 0}, sq[3] = {0, 0, 0};
  int i0 = 0, i1 = 0;
  if (has_p) {
    i0 = d.pt_off[p] - e0; i1 = d.pt_off[p + 1] - e0;
#pragma unroll
    for (int i = 0; i < 3; ++i) { gq[i] = d.gp[(size_t)p * 3 + i]; xq[i] = d.pts[(size_t)p * 3 + i]; sq[i] = d.sc2p[(size_t)p * 3 + i]; }
#pragma unroll
    for (int i = 0; i < 6; ++i) Vi[i] = d.Vinv[(size_t)p * 6 + i];
    Vd[0] = d.V[(size_t)p * 6]; Vd[1] = d.V[(size_t)p * 6 + 3]; Vd[2] = d.V[(size_t)p * 6 + 5];
  }
  const double radius = d.w_radius[w];
  double tl[3] = {0, 0, 0};          // long chunk (one point, thousands of observations): this thread's share
  if (staged) {
    // the rows of W first: independent loads, issued before the barrier the staged camera steps need
    double Wr[TMA ? 1 : 2][18];
    int sl[2] = {0, 0};
    static_assert(kBackObs <= 2 * kThreads, "two observations per thread");
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int i = threadIdx.x + k * kThreads;
      if (i < n) {
        if constexpr (!TMA) load_W(d.W, (size_t)(e0 + i), Wr[k]);
        sl[k] = d.oslot[e0 + i];
      }
    }
    __syncthreads();
    if constexpr (TMA) { if (n > 0) mbar_wait(bar, 0); }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int i = threadIdx.x + k * kThreads;
      if (i < n) {
        double s0 = 0.0, s1 = 0.0, s2 = 0.0;
        const double* dc = dcs + 6 * sl[k];
        double (&Wk)[18] = Wr[TMA ? 0 : k];
        if constexpr (TMA) {       // LDS.128 at a 144-byte stride: conflict-free per quarter-warp
          const double2* src = reinterpret_cast<const double2*>(Ws + (size_t)i * 18);
#pragma unroll
          for (int q = 0; q < 9; ++q) { const double2 v = src[q]; Wk[2 * q] = v.x; Wk[2 * q + 1] = v.y; }
        }
#pragma unroll
        for (int q = 0; q < 6; ++q) { s0 += Wk[3 * q] * dc[q]; s1 += Wk[3 * q + 1] * dc[q]; s2 += Wk[3 * q + 2] * dc[q]; }
        sobs[3 * i] = s0; sobs[3 * i + 1] = s1; sobs[3 * i + 2] = s2;
      }
    }
  } else {
    for (int i = threadIdx.x; i < n; i += kThreads) {
      const int f = __ldg(d.fc + __ldg(d.pcam + e0 + i));
      double s0 = 0.0, s1 = 0.0, s2 = 0.0;
      if (f >= 0) {
        double W[18];
        load_W(d.W, (size_t)(e0 + i), W);
#pragma unroll
        for (int q = 0; q < 6; ++q) { const double dc = d.x[(size_t)f * 6 + q]; s0 += W[3 * q] * dc; s1 += W[3 * q + 1] * dc; s2 += W[3 * q + 2] * dc; }
      }
      if (regular) { sobs[3 * i] = s0; sobs[3 * i + 1] = s1; sobs[3 * i + 2] = s2; }
      else { tl[0] += s0; tl[1] += s1; tl[2] += s2; }
    }
  }
  __syncthreads();
  if (!regular) {                    // fixed-order block sum of the shares; the chunk's one point belongs to thread 0
    (void)block_sum<3>(tl, sm);      // leaves the four warp sums in sm (the next block_sum starts with a barrier)
    if (threadIdx.x == 0) {
#pragma unroll
      for (int i = 0; i < 3; ++i) tl[i] = ((sm[i] + sm[3 + i]) + sm[6 + i]) + sm[9 + i];
    }
  }
  double part[3] = {0, 0, 0};
  double dmax = 0.0;
  if (has_p) {
    double t0 = gq[0], t1 = gq[1], t2 = gq[2];
    if (regular) {
      for (int i = i0; i < i1; ++i) { t0 += sobs[3 * i]; t1 += sobs[3 * i + 1]; t2 += sobs[3 * i + 2]; }
    } else { t0 += tl[0]; t1 += tl[1]; t2 += tl[2]; }
    const double dp[3] = {-(Vi[0] * t0 + Vi[1] * t1 + Vi[2] * t2), -(Vi[1] * t0 + Vi[3] * t1 + Vi[4] * t2),
                          -(Vi[2] * t0 + Vi[4] * t1 + Vi[5] * t2)};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      d.dp[(size_t)p * 3 + i] = dp[i];
      part[0] += gq[i] * dp[i];
      part[1] += lm_damp(sq[i], Vd[i], radius) * dp[i] * dp[i];
      part[2] += xq[i] * xq[i];
      dmax = fmax(dmax, fabs(dp[i]));
    }
  }
  const double tot = block_sum<3>(part, sm);
  const double mx = block_max1(dmax, smx);
  if (threadIdx.x < 3) d.step_part[(size_t)chunk * 4 + threadIdx.x] = tot;
  if (threadIdx.x == 3) d.step_part[(size_t)chunk * 4 + 3] = mx;
}

#ifndef SBA_BACKSUB_MIN_BLOCKS
#define SBA_BACKSUB_MIN_BLOCKS 5
#endif
__global__ void __launch_bounds__(kThreads, SBA_BACKSUB_MIN_BLOCKS) k_backsub(Dev d, DevParams P) {
  extern __shared__ __align__(16) double Ws[];             // [max regular chunk observations * 18]
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) mbar_init(&bar, 1);
  __syncthreads();
  dev_backsub<true>(d, P, blockIdx.x, Ws, &bar);
}

// one CTA per window: the step scalars (cameras + point chunk partials) and the line-search start state.
// (multi-GPU: camera terms are counted on rank 0 only; the scalars are all-reduced afterwards)
__device__ __forceinline__ void cam_step_finish(const Dev& d, int w, double (&part)[4], double dmax, double* sm, double* smx);

__device__ __forceinline__ void dev_cam_step(const Dev& d, int w, int rank) {
  __shared__ double sm[4 * 4];
  __shared__ double smx[4];
  double part[4] = {0, 0, 0, 0};   // g.d, d D d, |x|^2, dc . r_pcg
  double dmax = 0.0;
  const double radius = d.w_radius[w];
  for (int c = d.win_cam_off[w] + threadIdx.x; c < d.win_cam_off[w + 1]; c += kThreads) {
    const int f = d.fc[c];
    if (f < 0) continue;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const double x = d.poses[(size_t)c * 6 + i];
      const double dc = d.x[(size_t)f * 6 + i];
      dmax = fmax(dmax, fabs(dc));
      if (rank == 0) {
        part[0] += d.gc[(size_t)c * 6 + i] * dc;
        part[1] += lm_damp(d.sc2c[(size_t)c * 6 + i], d.U[(size_t)c * 21 + up6(i, i)], radius) * dc * dc;
        part[2] += x * x;
        part[3] += dc * d.r[(size_t)f * 6 + i];
      }
    }
  }
  // (a global window has tens of thousands of chunk partials for this one CTA: eight rows in flight per thread, added in
  // the same order as one at a time)
  {
    const int c1 = d.win_pchunk_off[w + 1];
    int ch = d.win_pchunk_off[w] + threadIdx.x;
    for (; ch + 7 * kThreads < c1; ch += 8 * kThreads) {
      double2 va[8], vb[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const double2* q = reinterpret_cast<const double2*>(d.step_part + (size_t)(ch + u * kThreads) * 4);
        va[u] = __ldcg(q); vb[u] = __ldcg(q + 1);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) { part[0] += va[u].x; part[1] += va[u].y; part[2] += vb[u].x; dmax = fmax(dmax, vb[u].y); }
    }
    for (; ch < c1; ch += kThreads) {
      const double2* q = reinterpret_cast<const double2*>(d.step_part + (size_t)ch * 4);
      const double2 va = __ldcg(q), vb = __ldcg(q + 1);
      part[0] += va.x; part[1] += va.y; part[2] += vb.x; dmax = fmax(dmax, vb.y);
    }
  }
  cam_step_finish(d, w, part, dmax, sm, smx);
}

// the window's step scalars from per-thread partials (fixed order: block tree), and the line-search start state
__device__ __forceinline__ void cam_step_finish(const Dev& d, int w, double (&part)[4], double dmax, double* sm, double* smx) {
  double tot = block_sum<4>(part, sm);
  const double mx = block_max1(dmax, smx);
  if (threadIdx.x == 0 && d.w_fail[w] != 0) tot = nan("");   // poison the model change: invalid step on every rank
  if (threadIdx.x == 0) {
    d.w_gd[w] = tot;
    d.w_dmax[w] = mx;
    d.w_accept[w] = 0;
    d.w_ls_state[w] = (d.w_status[w] == 0) ? 1 : 0;     // 1: evaluate the full step (cost only)
    d.w_ls_x[w] = 1.0; d.w_ls_it[w] = 0; d.w_ls_pvalid[w] = 0;
  }
  if (threadIdx.x == 1) d.w_dDd[w] = tot;
  if (threadIdx.x == 2) d.w_x2[w] = tot;
  if (threadIdx.x == 3) d.w_dr[w] = tot;
}

__global__ void __launch_bounds__(kThreads) k_cam_step(Dev d, int rank) { dev_cam_step(d, blockIdx.x, rank); }

// ----------------------------------------------- K6: trial point of the projected line search + its cost
// Ceres evaluates x (+) s*delta with the bounds projection inside Plus (parameter_block.h), first at s = 1.
// thread per camera of a searching window: trial pose and its precompute
__device__ __forceinline__ void dev_trial_cam(const Dev& d, int c) {
  if (c >= d.Nc) return;
  const int w = d.cam_win[c];
  const int st = __ldcg(d.w_ls_state + w);
  if (st < 1 || st > 3) return;
  const double step = __ldcg(d.w_ls_x + w);
  const int f = d.fc[c];
  double pose[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    pose[i] = d.poses[(size_t)c * 6 + i];
    if (f >= 0) pose[i] += step * d.x[(size_t)f * 6 + i];
    d.poses_new[(size_t)c * 6 + i] = pose[i];
  }
  CamPre cp;
  cam_precompute(pose, cp);
  d.campre_new[c] = cp;
}

__global__ void k_trial_cams(Dev d) { dev_trial_cam(d, blockIdx.x * blockDim.x + threadIdx.x); }

// thread per point: trial point, cost, |x_trial - x|^2 and (state 2) the directional derivative
//   d/ds cost = sum_obs  rho' (D^T r) . (G dw + dt + R dp)     evaluated at the trial point
constexpr int kTrialSlots = 32;    // chunks with at most this many cameras compute their trial cameras in shared memory

// from_table: the trial cameras of the chunk's slots come from campre_new (written by dev_trial_cam for this trial) instead of
// being recomputed here — the same bits; on a problem with dozens of chunks per CTA the sin / cos of every slot of
// every chunk, by a handful of threads while the rest of the CTA waits, was the critical path of the trial evaluation.
__device__ __forceinline__ void dev_trial_points(const Dev& d, const DevParams& P, int chunk, bool from_table = false) {
  __shared__ double sm[4 * 3];
  __shared__ double cps[kTrialSlots * 21];    // trial R / J_l / t of the chunk's cameras
  __shared__ double Xs[kPointChunk * 3];      // the chunk's trial points
  __shared__ double dps[kPointChunk * 3];     // and their steps
  __shared__ int cfs[kTrialSlots];            // their free-camera index (or -1)
  const int4 cd0 = __ldg(d.cdesc + 4 * (size_t)chunk), cd1 = __ldg(d.cdesc + 4 * (size_t)chunk + 1);
  const int p0 = cd0.x, np = cd0.y, e0 = cd0.z, n = cd0.w, w = cd1.x, cs0 = cd1.y, ns = cd1.z;
  const int st = __ldcg(d.w_ls_state + w);
  if (st < 1 || st > 3) return;
  const bool grad = (st == 2);
  const double step = __ldcg(d.w_ls_x + w);
  // Trial cameras of the chunk, computed here exactly as dev_trial_cam does (same expression, same bits) instead of
  // read back from campre_new: no L2 round trip per observation, and no dependence on the trial-camera phase.
  const bool staged = ns <= kTrialSlots;
  if (staged && from_table) {
    for (int k = threadIdx.x; k < ns * 21; k += kThreads) {
      const int sl = k / 21;
      cps[k] = __ldcg(reinterpret_cast<const double*>(d.campre_new + __ldg(d.cs_cam + cs0 + sl)) + (k - 21 * sl));
    }
    for (int sl = threadIdx.x; sl < ns; sl += kThreads) cfs[sl] = d.fc[__ldg(d.cs_cam + cs0 + sl)];
  } else if (staged) {
    for (int sl = threadIdx.x; sl < ns; sl += kThreads) {
      const int c = d.cs_cam[cs0 + sl];
      const int f = d.fc[c];
      double pose[6];
#pragma unroll
      for (int i = 0; i < 6; ++i) {
        pose[i] = d.poses[(size_t)c * 6 + i];
        if (f >= 0) pose[i] += step * d.x[(size_t)f * 6 + i];
      }
      CamPre cp;
      cam_precompute(pose, cp);
      *reinterpret_cast<CamPre*>(cps + 21 * sl) = cp;
      cfs[sl] = f;
    }
  }
  double part[3] = {0, 0, 0};
  // thread per point: the projected trial point (kept in shared memory for the chunk's observations), |x_trial - x|^2
  static_assert(kPointChunk <= kThreads, "one point per thread");
  if ((int)threadIdx.x < np) {
    const int p = p0 + threadIdx.x;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const double x = d.pts[3 * (size_t)p + i];
      const double dv = d.dp[3 * (size_t)p + i];
      const double X = clampd(x + step * dv, P.lo, P.hi);
      d.pts_new[3 * (size_t)p + i] = X;
      Xs[3 * threadIdx.x + i] = X; dps[3 * threadIdx.x + i] = dv;
      part[2] += (X - x) * (X - x);
    }
  }
  // a regular chunk has at most two observations per thread: their index / pixel loads do not depend on the staged
  // cameras and go out before the barrier (one L2 round trip less on the chunk's critical path)
  const bool pre = staged && n <= 2 * kThreads;
  int pre_pl[2] = {0, 0}, pre_sl[2] = {0, 0};
  float4 pre_m[2] = {make_float4(0.f, 0.f, 0.f, 0.f), make_float4(0.f, 0.f, 0.f, 0.f)};
  if (pre) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int i = threadIdx.x + k * kThreads;
      if (i < n) { pre_pl[k] = __ldg(d.ppt + e0 + i) - p0; pre_m[k] = __ldg(d.uv_p + e0 + i); pre_sl[k] = d.oslot[e0 + i]; }
    }
  }
  __syncthreads();
  // thread per observation: its cost and (state 2) its share of the directional derivative
  auto one = [&](int e, int pl, const float4 m, int sl) {
    const float uv[4] = {m.x, m.y, m.z, m.w};
    const double X[3] = {Xs[3 * pl], Xs[3 * pl + 1], Xs[3 * pl + 2]};
    CamPre cp;
    int f;
    if (staged) {
      const double* src = cps + 21 * sl;
#pragma unroll
      for (int k = 0; k < 9; ++k) cp.R[k] = src[k];
#pragma unroll
      for (int k = 0; k < 3; ++k) cp.t[k] = src[18 + k];
      if (grad) {
#pragma unroll
        for (int k = 0; k < 9; ++k) cp.Jl[k] = src[9 + k];
      }
      f = cfs[sl];
    } else {
      const int c = __ldg(d.pcam + e);
      if (grad) load_campre(d.campre_new, c, cp); else load_campre_Rt(d.campre_new, c, cp);
      f = __ldg(d.fc + c);
    }
    if (!grad) {
      part[0] += obs_cost(P.pl, P.pr, cp, X, uv, P.rdim, P.huber_delta);
    } else {
      const double dpv[3] = {dps[3 * pl], dps[3 * pl + 1], dps[3 * pl + 2]};
      double q[3], r[4], D[12];
      const double s = project(P.pl, P.pr, cp, X, uv, P.rdim, q, r, D);
      double rho, wt;
      huber(P.huber_delta, s, rho, wt);
      part[0] += 0.5 * rho;
      double A[6], h[3], G[9];
      form_A_h(D, r, wt, A, h);
      form_G(cp, q, G);
      double v[3] = {cp.R[0] * dpv[0] + cp.R[1] * dpv[1] + cp.R[2] * dpv[2],
                     cp.R[3] * dpv[0] + cp.R[4] * dpv[1] + cp.R[5] * dpv[2],
                     cp.R[6] * dpv[0] + cp.R[7] * dpv[1] + cp.R[8] * dpv[2]};
      if (f >= 0) {
        const double* dc = d.x + (size_t)f * 6;
#pragma unroll
        for (int k = 0; k < 3; ++k) v[k] += G[3 * k] * dc[0] + G[3 * k + 1] * dc[1] + G[3 * k + 2] * dc[2] + dc[3 + k];
      }
      part[1] += h[0] * v[0] + h[1] * v[1] + h[2] * v[2];
    }
  };
  if (pre) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int i = threadIdx.x + k * kThreads;
      if (i < n) one(e0 + i, pre_pl[k], pre_m[k], pre_sl[k]);
    }
  } else {
    for (int i = threadIdx.x; i < n; i += kThreads) {
      const int e = e0 + i;
      one(e, __ldg(d.ppt + e) - p0, __ldg(d.uv_p + e), staged ? (int)d.oslot[e] : 0);
    }
  }
  const double tot = block_sum<3>(part, sm);
  if (threadIdx.x < 3) d.ls_part[(size_t)chunk * 3 + threadIdx.x] = tot;
}

__global__ void __launch_bounds__(kThreads) k_trial_points(Dev d, DevParams P) { dev_trial_points(d, P, blockIdx.x, true); }   // k_trial_cams ran before

// one CTA per searching window: raw sums (cost, directional derivative, |step|^2 incl. the cameras on rank 0)
__device__ __forceinline__ void dev_ls_reduce(const Dev& d, int w, int rank) {
  __shared__ double sm[4 * 3];
  const int st = __ldcg(d.w_ls_state + w);
  if (st < 1 || st > 3) return;
  double part[3] = {0, 0, 0};
  {
    const int c1 = d.win_pchunk_off[w + 1];
    int ch = d.win_pchunk_off[w] + threadIdx.x;
    for (; ch + 7 * kThreads < c1; ch += 8 * kThreads) {     // eight rows in flight per thread, same order of additions
      double v[8][3];
#pragma unroll
      for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int i = 0; i < 3; ++i) v[u][i] = __ldcg(d.ls_part + (size_t)(ch + u * kThreads) * 3 + i);
#pragma unroll
      for (int u = 0; u < 8; ++u)
#pragma unroll
        for (int i = 0; i < 3; ++i) part[i] += v[u][i];
    }
    for (; ch < c1; ch += kThreads) {
#pragma unroll
      for (int i = 0; i < 3; ++i) part[i] += __ldcg(d.ls_part + (size_t)ch * 3 + i);
    }
  }
  if (rank == 0)
    for (int c = d.win_cam_off[w] + threadIdx.x; c < d.win_cam_off[w + 1]; c += kThreads) {
#pragma unroll
      for (int i = 0; i < 6; ++i) { const double dd = __ldcg(d.poses_new + (size_t)c * 6 + i) - d.poses[(size_t)c * 6 + i]; part[2] += dd * dd; }
    }
  const double tot = block_sum<3>(part, sm);
  if (threadIdx.x < 3) d.w_ls_red[(size_t)threadIdx.x * d.nwin + w] = tot;
}

__global__ void __launch_bounds__(kThreads) k_ls_reduce(Dev d, int rank) { dev_ls_reduce(d, blockIdx.x, rank); }

// -------------------------------------------------------------------------------- LM control
// Ceres trust-region rules (trust_region_minimizer.cc / levenberg_marquardt_strategy.cc restated;
// options from the reference: src/bundle_adjuster.cpp:14-36).  Thread per window.

// after a linearisation: gradient-tolerance test; `first` also initialises the LM state
__device__ __forceinline__ void dev_lm_begin(const Dev& d, const DevParams& P, int w, double initial_radius, int first, int clear_fail) {
  if (first) {
    d.w_radius[w] = initial_radius; d.w_decf[w] = 2.0; d.w_status[w] = 0; d.w_iter[w] = 0; d.w_succ[w] = 0;
    d.w_invalid[w] = 0; d.w_accept[w] = 0;
  }
  if (clear_fail) d.w_fail[w] = 0;
  if (d.w_status[w] == 0) {
    if (!isfinite(d.w_cost[w])) d.w_status[w] = -2;                       // numerical failure
    else if (d.w_gmax[w] <= P.gtol) d.w_status[w] = 3;                    // gradient tolerance
  }
}
__global__ void k_lm_begin(Dev d, DevParams P, double initial_radius, int first, int clear_fail) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w < d.nwin) dev_lm_begin(d, P, w, initial_radius, first, clear_fail);
}

// Projected Armijo line search, one thread per window (ArmijoLineSearch::DoSearch restated; see
// sba_linesearch.cuh).  States: 1 full step evaluated (cost only) · 2 trial evaluated with its directional
// derivative · 3 full step re-evaluated after a failed search · 4 candidate ready for k_lm_control · 0 idle.
__device__ __forceinline__ void dev_ls_control(const Dev& d, const DevParams& P, int w, int32_t* searching) {
  if (w >= d.nwin) return;
  const int st = d.w_ls_state[w];
  if (st < 1 || st > 3) return;
  const double f = d.w_ls_red[w], g = d.w_ls_red[(size_t)d.nwin + w], step2 = d.w_ls_red[2 * (size_t)d.nwin + w];
  const double x = d.w_ls_x[w];
  const double cost0 = d.w_cost[w], gd = d.w_gd[w];
  if (P.debug == 1 && w == 0) printf("[ls] it %d st %d x %.6e f-cost0 %.6e armijo_rhs %.6e g %.6e gd %.6e\n", d.w_iter[w] + 1, st, x, f - cost0, 1e-4 * gd * x, g, gd);
  if (st == 3) { d.w_cand[w] = f; d.w_step2[w] = step2; d.w_ls_state[w] = 4; return; }
  if (st == 1) {
    // an invalid trust-region step never reaches the line search (k_lm_control treats it)
    const double model = -0.5 * gd + 0.5 * d.w_dDd[w] + 0.5 * d.w_dr[w];
    if (d.w_fail[w] != 0 || !(model > 0.0) || !isfinite(model)) { d.w_cand[w] = f; d.w_step2[w] = step2; d.w_ls_state[w] = 4; return; }
  }
  const bool valid = isfinite(f) && (st == 1 || isfinite(g));
  if (valid && f <= cost0 + 1e-4 * gd * x) { d.w_cand[w] = f; d.w_step2[w] = step2; d.w_ls_state[w] = 4; return; }
  if (st == 1 && valid) {          // sufficient decrease failed: the same step again, now with the derivative
    d.w_ls_state[w] = 2;
    atomicAdd(searching, 1);
    return;
  }
  bool failed = false;
  double nx = x;
  const int it = ++d.w_ls_it[w];
  if (it >= 20) failed = true;      // max_num_line_search_step_size_iterations
  else {
    LsSample cur; cur.x = x; cur.f = f; cur.g = g; cur.has_g = 1;
    LsSample prev; prev.x = d.w_ls_px[w]; prev.f = d.w_ls_pf[w]; prev.g = d.w_ls_pg[w]; prev.has_g = 1;
    nx = ls_next_step(cost0, gd, prev, d.w_ls_pvalid[w], cur, valid ? 1 : 0);
    if (nx * d.w_dmax[w] < 1e-9) failed = true;     // min_line_search_step_size
  }
  if (failed) {                     // Ceres keeps the whole step when the search fails
    d.w_ls_x[w] = 1.0; d.w_ls_state[w] = 3;
  } else {
    d.w_ls_px[w] = x; d.w_ls_pf[w] = f; d.w_ls_pg[w] = g; d.w_ls_pvalid[w] = valid ? 1 : 0;
    d.w_ls_x[w] = nx; d.w_ls_state[w] = 2;
  }
  atomicAdd(searching, 1);
}
__global__ void k_ls_control(Dev d, DevParams P) { dev_ls_control(d, P, blockIdx.x * blockDim.x + threadIdx.x, d.counters + 3); }

__device__ __forceinline__ void dev_lm_control(const Dev& d, const DevParams& P, int w) {
  if (w >= d.nwin) return;
  if (d.w_ls_state[w] != 4) return;      // idle, or its line search is still running
  d.w_ls_state[w] = 0;
  if (d.w_status[w] != 0) return;
  const int iter = ++d.w_iter[w];
  const double cost = d.w_cost[w], cand = d.w_cand[w];
  const double model = -0.5 * d.w_gd[w] + 0.5 * d.w_dDd[w] + 0.5 * d.w_dr[w];
  bool stop = false;
  if (d.w_fail[w] != 0 || !(model > 0.0) || !isfinite(model)) {
    // invalid step: Ceres halves the radius; 5 in a row is a failure
    if (++d.w_invalid[w] >= 5) { d.w_status[w] = 7; stop = true; }
    else d.w_radius[w] *= 0.5;
  } else {
    d.w_invalid[w] = 0;
    const double step_norm = sqrt(d.w_step2[w]), x_norm = sqrt(d.w_x2[w]);
    if (step_norm <= P.ptol * (x_norm + P.ptol)) { d.w_status[w] = 1; stop = true; }
    else {
      const double cost_change = cost - cand;
      if (fabs(cost_change) <= P.ftol * cost) { d.w_status[w] = 2; stop = true; }
      else {
        const double rho = cost_change / model;
        if (isfinite(cand) && rho > 1e-3) {
          d.w_accept[w] = 1; d.w_succ[w]++;
          atomicAdd(d.counters + 2, 1);     // control-plane counter: windows that moved
          d.w_cost[w] = cand;
          const double t = 2.0 * rho - 1.0;
          d.w_radius[w] = fmin(1e16, d.w_radius[w] / fmax(1.0 / 3.0, 1.0 - t * t * t));
          d.w_decf[w] = 2.0;
        } else {
          d.w_radius[w] = d.w_radius[w] / d.w_decf[w];
          d.w_decf[w] *= 2.0;
        }
      }
    }
  }
  if (!stop) {
    if (iter >= P.max_iterations) d.w_status[w] = 4;
    else if (d.w_radius[w] < 1e-32) d.w_status[w] = 6;
    else atomicAdd(d.counters, 1);     // control-plane counter: windows still running
  }
}

__global__ void k_lm_control(Dev d, DevParams P) { dev_lm_control(d, P, blockIdx.x * blockDim.x + threadIdx.x); }

// wall-clock limit hit on the host (reference: max_solver_time_in_seconds, src/bundle_adjuster.cpp:18)
__global__ void k_lm_timeout(Dev d) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w < d.nwin && d.w_status[w] == 0) d.w_status[w] = 5;
}

// accepted windows: candidate -> current
__device__ __forceinline__ void dev_apply_points(const Dev& d, int chunk) {
  if (!__ldcg(d.w_accept + d.pchunk_win[chunk])) return;
  const int p = d.pchunk_begin[chunk] + threadIdx.x;
  if (p >= d.pchunk_end[chunk]) return;
#pragma unroll
  for (int i = 0; i < 3; ++i) d.pts[(size_t)p * 3 + i] = d.pts_new[(size_t)p * 3 + i];
}
__global__ void __launch_bounds__(kThreads) k_apply_points(Dev d) { dev_apply_points(d, blockIdx.x); }
__device__ __forceinline__ void dev_apply_cam(const Dev& d, int t) {
  if (t >= d.Nc * 6) return;
  if (__ldcg(d.w_accept + d.cam_win[t / 6])) d.poses[t] = __ldcg(d.poses_new + t);
}
__global__ void k_apply_cams(Dev d) { dev_apply_cam(d, blockIdx.x * blockDim.x + threadIdx.x); }

// K5 + K6 + the trust-region decision of every window in ONE cooperative launch (single GPU): back-substitution,
// step scalars, then the projected line search round by round — trial cameras | grid barrier | trial points and
// their costs | grid barrier | per-window reduction and the Armijo / LM decision | grid barrier — until no window
// is still searching, then accepted candidates become the state.  Replaces 5 launches, a memset and a host round
// trip per line-search round.  The searching counter alternates between counters[3] and [5] so that resetting
// one never races with reading the other; counters[6] returns the number of rounds.
// ---- one global window (thousands of cameras, 10^4..10^5 chunks): the window's reductions in two levels.  Every CTA of
// the step loop first sums the rows it produced itself (chunk order) and its slice of the cameras into one row of
// `cta_part`; the window's CTA then adds a few hundred rows instead of walking 10^5 chunk partials alone.
//   row layout: [0..2] chunk g.dp, dp D dp, |x|^2  [3] chunk max|dp|  [4..7] camera g.d, d D d, |x|^2, dc.r  [8] camera max|dc|
//               [9..11] trial cost, directional derivative, |x_trial - x|^2 (points)  [12] |x_trial - x|^2 (cameras)
__device__ __forceinline__ void dev_step_partials_one(const Dev& d, int nb) {
  __shared__ double sm[4 * 4];
  __shared__ double smx[4];
  double* row = d.cta_part + (size_t)blockIdx.x * 16;
  if (threadIdx.x < 4) {
    double a = 0.0;
    for (int ch = blockIdx.x; ch < d.npchunk; ch += nb) {
      const double v = __ldcg(d.step_part + (size_t)ch * 4 + threadIdx.x);
      a = (threadIdx.x < 3) ? a + v : fmax(a, v);
    }
    row[threadIdx.x] = a;
  }
  double part[4] = {0, 0, 0, 0};
  double dmax = 0.0;
  const double radius = d.w_radius[0];
  for (int c = blockIdx.x * kThreads + threadIdx.x; c < d.Nc; c += nb * kThreads) {
    const int f = d.fc[c];
    if (f < 0) continue;
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const double x = d.poses[(size_t)c * 6 + i];
      const double dc = d.x[(size_t)f * 6 + i];
      dmax = fmax(dmax, fabs(dc));
      part[0] += d.gc[(size_t)c * 6 + i] * dc;
      part[1] += lm_damp(d.sc2c[(size_t)c * 6 + i], d.U[(size_t)c * 21 + up6(i, i)], radius) * dc * dc;
      part[2] += x * x;
      part[3] += dc * d.r[(size_t)f * 6 + i];
    }
  }
  const double tot = block_sum<4>(part, sm);
  const double mx = block_max1(dmax, smx);
  if (threadIdx.x < 4) row[4 + threadIdx.x] = tot;
  if (threadIdx.x == 4) row[8] = mx;
}

__device__ __forceinline__ void dev_cam_step_one(const Dev& d, int nb) {
  __shared__ double sm[4 * 4];
  __shared__ double smx[4];
  double part[4] = {0, 0, 0, 0};
  double dmax = 0.0;
  for (int r = threadIdx.x; r < nb; r += kThreads) {
    const double* row = d.cta_part + (size_t)r * 16;
    part[0] += __ldcg(row + 4) + __ldcg(row); part[1] += __ldcg(row + 5) + __ldcg(row + 1); part[2] += __ldcg(row + 6) + __ldcg(row + 2);
    part[3] += __ldcg(row + 7);
    dmax = fmax(dmax, fmax(__ldcg(row + 3), __ldcg(row + 8)));
  }
  cam_step_finish(d, 0, part, dmax, sm, smx);
}

__device__ __forceinline__ void dev_ls_partials_one(const Dev& d, int nb, double cam2) {
  __shared__ double sm[4];
  double* row = d.cta_part + (size_t)blockIdx.x * 16;
  if (threadIdx.x < 3) {
    double a = 0.0;
    for (int ch = blockIdx.x; ch < d.npchunk; ch += nb) a += __ldcg(d.ls_part + (size_t)ch * 3 + threadIdx.x);
    row[9 + threadIdx.x] = a;
  }
  double c2[1] = {cam2};
  const double tot = block_sum<1>(c2, sm);
  if (threadIdx.x == 0) row[12] = tot;
}

__device__ __forceinline__ void dev_ls_reduce_one(const Dev& d, int nb) {
  __shared__ double sm[4 * 3];
  const int st = __ldcg(d.w_ls_state);
  if (st < 1 || st > 3) return;
  double part[3] = {0, 0, 0};
  for (int r = threadIdx.x; r < nb; r += kThreads) {
    const double* row = d.cta_part + (size_t)r * 16;
    part[0] += __ldcg(row + 9); part[1] += __ldcg(row + 10); part[2] += __ldcg(row + 11) + __ldcg(row + 12);
  }
  const double tot = block_sum<3>(part, sm);
  if (threadIdx.x < 3) d.w_ls_red[(size_t)threadIdx.x * d.nwin] = tot;
}

// (4 CTAs per SM: the step's phases are bound by L2 latency, not by registers; measured 2 -> 4: C4 step 1.08 -> 0.64 ms)
#ifndef SBA_STEP_MIN_BLOCKS
#define SBA_STEP_MIN_BLOCKS 4
#endif
__global__ void __launch_bounds__(kThreads, SBA_STEP_MIN_BLOCKS) k_step_loop(Dev d, DevParams P, int fused_cams, int skip_backsub) {
  cooperative_groups::grid_group grid = cooperative_groups::this_grid();
  const int nb = gridDim.x;
  const bool one = d.nwin == 1 && d.npchunk > 4 * nb;      // one global window: two-level reductions
  // P.debug == 2 (SBA_DEBUG_LS=2): block 0 prints the nanoseconds spent in each phase (tuning aid)
  const bool prof = P.debug == 2 && blockIdx.x == 0 && threadIdx.x == 0;
  unsigned long long t0 = prof ? gtimer() : 0, tp[6] = {0, 0, 0, 0, 0, 0};
  auto lap = [&](int k) { if (prof) { const unsigned long long t = gtimer(); tp[k] += t - t0; t0 = t; } };
  for (int ch = blockIdx.x; ch < d.npchunk && !skip_backsub; ch += nb) { dev_backsub<false>(d, P, ch); __syncthreads(); }
  if (blockIdx.x == 0 && threadIdx.x == 0) { d.counters[0] = 0; d.counters[2] = 0; d.counters[3] = 0; d.counters[5] = 0; }
  if (one) dev_step_partials_one(d, nb);
  grid.sync();
  lap(0);
  if (one) { if (blockIdx.x == 0) dev_cam_step_one(d, nb); }
  else for (int w = blockIdx.x; w < d.nwin; w += nb) { dev_cam_step(d, w, 0); __syncthreads(); }
  grid.sync();
  lap(1);
  int round = 0;
  for (;;) {
    int32_t* cnt = d.counters + ((round & 1) ? 5 : 3);
    int32_t* other = d.counters + ((round & 1) ? 3 : 5);
    double cam2 = 0.0;
    for (int c = blockIdx.x * kThreads + threadIdx.x; c < d.Nc; c += nb * kThreads) {
      dev_trial_cam(d, c);
      if (one) {           // this thread's share of |x_trial - x|^2 over the cameras (it wrote poses_new[c] itself)
#pragma unroll
        for (int i = 0; i < 6; ++i) { const double dd = __ldcg(d.poses_new + (size_t)c * 6 + i) - d.poses[(size_t)c * 6 + i]; cam2 += dd * dd; }
      }
    }
    if (!fused_cams) grid.sync();         // the chunks read campre_new only when one of them has too many cameras to stage
    lap(2);
    for (int ch = blockIdx.x; ch < d.npchunk; ch += nb) { dev_trial_points(d, P, ch, !fused_cams); __syncthreads(); }
    if (one) dev_ls_partials_one(d, nb, cam2);
    grid.sync();
    lap(3);
    if (blockIdx.x == 0 && threadIdx.x == 0) *other = 0;
    if (one) {
      if (blockIdx.x == 0) {
        dev_ls_reduce_one(d, nb);
        __syncthreads();
        if (threadIdx.x == 0) { dev_ls_control(d, P, 0, cnt); dev_lm_control(d, P, 0); }
      }
    } else
    for (int w = blockIdx.x; w < d.nwin; w += nb) {
      dev_ls_reduce(d, w, 0);
      __syncthreads();
      if (threadIdx.x == 0) { dev_ls_control(d, P, w, cnt); dev_lm_control(d, P, w); }
      __syncthreads();
    }
    grid.sync();
    lap(4);
    ++round;
    if (__ldcg(cnt) <= 0 || round > 64) break;
  }
  for (int ch = blockIdx.x; ch < d.npchunk; ch += nb) dev_apply_points(d, ch);
  for (int t = blockIdx.x * kThreads + threadIdx.x; t < d.Nc * 6; t += nb * kThreads) dev_apply_cam(d, t);
  if (blockIdx.x == 0 && threadIdx.x == 0) d.counters[6] = round;
  lap(5);
  if (prof) printf("[step] rounds %d  backsub %llu  cam_step %llu  trial_cam %llu  trial_points %llu  reduce+control %llu  apply %llu ns\n", round, tp[0], tp[1], tp[2], tp[3], tp[4], tp[5]);
}

// FP64 FMA peak of the device (bench.py reports the fp64 fraction beside the HBM fraction): eight
// independent DFMA chains per thread, no memory traffic.  2 * 8 * iters flop per thread.
__global__ void __launch_bounds__(256) k_fp64_peak(double* out, int iters, double a, double b) {
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 4
  for (int i = 0; i < iters; ++i) {
    x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
    x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
  }
  const double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;     // never true: keeps the chains alive
}

// ---- device-resident map (sba_map_*): the window's points are gathered from / scattered to the map's f32 storage by
// their global ids, exactly as the pack / unpack loops do on the host (reference: src/bundle_adjuster.cpp:83-87, :128-132)
__global__ void k_map_gather_points(int np, const uint32_t* __restrict__ ids, const float* __restrict__ map_points, double* __restrict__ pts) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= np * 3) return;
  const int i = t / 3, c = t - 3 * i;
  pts[t] = (double)map_points[(size_t)ids[i] * 3 + c];
}
__global__ void k_map_scatter_points(int np, const uint32_t* __restrict__ ids, const double* __restrict__ pts, float* __restrict__ map_points) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= np * 3) return;
  const int i = t / 3, c = t - 3 * i;
  map_points[(size_t)ids[i] * 3 + c] = (float)pts[t];
}

// permute the caller's observations into the point-sorted view
__global__ void k_permute_uv(Dev d, const float4* __restrict__ uv_in, float4* __restrict__ uv_p) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= d.Nobs) return;
  uv_p[e] = uv_in[d.perm_p[e]];
}

}  // namespace sba
