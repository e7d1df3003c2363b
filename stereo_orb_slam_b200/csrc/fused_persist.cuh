// k_fused_persist — the fused K1-K3 chunk kernel (k_fused_chunk, kernels.cuh) as PERSISTENT CTAs that fetch the next
// chunk's inputs while the current chunk computes.
//
// k_fused_chunk spends 3-4 of its ~15 us per chunk waiting for three dependent levels of global loads (chunk record ->
// indices / pixels / points -> camera rows), and reads its pair list from global memory in the middle of the chunk.  Here
// two CTAs per SM walk the chunks blockIdx.x, blockIdx.x + gridDim.x, ...; every input of chunk k + 1 is copied to shared
// memory by cp.async while chunk k runs, into the place of the same input of chunk k as soon as that is dead:
//   group 1 (after phase A)      pixels, point index and slot of every observation; the index tables (double-buffered:
//                                they are live until the end of the chunk); the record of chunk k + 2
//   group 2 (after phase B)      points, their Jacobi scaling, their observation offsets
//   group 3 (after phase C)      second level: the camera rows and free-camera flags of the slots copied by group 1;
//                                the pair list
// The arithmetic is k_fused_chunk's, phase by phase and operation by operation: results are bit-identical
// (tests/test_gpu_parity.py).  Used when every chunk fits the fixed buffers (kFP*: <= 16 camera slots, <= 160 local
// blocks, <= 2048 pairs that also fit the pair staging area) and the structure has no long-track chunk; otherwise
// k_fused_chunk runs.
#pragma once
#include <cuda_pipeline.h>
#include "kernels.cuh"

namespace sba {

constexpr int kFPSlots = 16;
constexpr int kFPLb = 160;
constexpr int kFPPairs = 2048;

__global__ void __launch_bounds__(kFusedThreads, 2) k_fused_persist(Dev d, DevParams P) {
  constexpr int NT = kFusedThreads, NW = NT / 32;
  extern __shared__ __align__(16) double smem[];
  __shared__ double sm1[NW];
  __shared__ double smx[NW];
  // single-buffered inputs (rewritten for the next chunk once the current one has consumed them)
  __shared__ __align__(16) float4 b_uv[256];
  __shared__ __align__(16) int b_ppt[256];
  __shared__ __align__(16) uint8_t b_oslot[264];             // byte k of the chunk at (e0 & 3) + k
  __shared__ __align__(16) double b_pts[kPointChunk * 3];
  __shared__ __align__(16) double b_sc2[kPointChunk * 3];
  __shared__ __align__(16) int b_ptoff[kPointChunk + 4];
  __shared__ __align__(16) double b_cpre[kFPSlots * 21];
  __shared__ __align__(16) uint16_t b_pairs[kFPPairs + 8];   // pair k of the chunk at (pq0 & 1) + k
  // double-buffered index tables
  __shared__ __align__(16) uint16_t t_sobs[2][264];          // entry j of the chunk at (e0 & 1) + j
  __shared__ int t_off[2][kFPSlots + 1], t_pos[2][kFPSlots], t_cam[2][kFPSlots], t_fc[2][kFPSlots];
  __shared__ int t_lbq[2][kFPLb + 1], t_lbpos[2][kFPLb];
  __shared__ __align__(16) int4 ring[3][3];                  // chunk records of this, the next and the one after
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = threadIdx.x;
  const int stride = gridDim.x;
  if ((int)blockIdx.x >= d.npchunk) return;
  // ---- prologue: the first chunk's inputs the slow way (three dependent levels, once per CTA)
  if (i < 3) {
    ring[0][i] = __ldg(d.cdesc + 4 * (size_t)blockIdx.x + i);
    if ((int)blockIdx.x + stride < d.npchunk) ring[1][i] = __ldg(d.cdesc + 4 * ((size_t)blockIdx.x + stride) + i);
  }
  __syncthreads();
  int it = 0;
  // copies of chunk `slot`'s inputs; cur / nxt select the halves of the double-buffered tables
  auto g1 = [&](int slot, int tb) {
    const int4 c0 = ring[slot][0], c1 = ring[slot][1], c2 = ring[slot][2];
    const int e0 = c0.z, n = c0.w, cs0 = c1.y, ns = c1.z, lb0 = c1.w, nlb = c2.x;
    if (i < n) {
      __pipeline_memcpy_async(b_uv + i, d.uv_p + e0 + i, 16);
      __pipeline_memcpy_async(b_ppt + i, d.ppt + e0 + i, 4);
    }
    if (i < ((e0 & 3) + n + 3) / 4) __pipeline_memcpy_async(reinterpret_cast<uint32_t*>(b_oslot) + i, reinterpret_cast<const uint32_t*>(d.oslot + (e0 & ~3)) + i, 4);
    if (i < ((e0 & 1) + n + 1) / 2) __pipeline_memcpy_async(reinterpret_cast<uint32_t*>(t_sobs[tb]) + i, reinterpret_cast<const uint32_t*>(d.sobs + (e0 & ~1)) + i, 4);
    if (i <= ns) __pipeline_memcpy_async(t_off[tb] + i, d.cs_obs_off + cs0 + i, 4);
    if (i < ns) { __pipeline_memcpy_async(t_pos[tb] + i, d.cs_pos + cs0 + i, 4); __pipeline_memcpy_async(t_cam[tb] + i, d.cs_cam + cs0 + i, 4); }
    if (i <= nlb) __pipeline_memcpy_async(t_lbq[tb] + i, d.lb_pair_off + lb0 + i, 4);
    if (i < nlb) __pipeline_memcpy_async(t_lbpos[tb] + i, d.lb_pos + lb0 + i, 4);
  };
  auto g2 = [&](int slot) {
    const int4 c0 = ring[slot][0];
    const int p0 = c0.x, np = c0.y;
#pragma unroll
    for (int u = 0; u < (kPointChunk * 3 + NT - 1) / NT; ++u) {
      const int k = i + u * NT;
      if (k < np * 3) { __pipeline_memcpy_async(b_pts + k, d.pts + (size_t)p0 * 3 + k, 8); __pipeline_memcpy_async(b_sc2 + k, d.sc2p + (size_t)p0 * 3 + k, 8); }
    }
    if (i <= np) __pipeline_memcpy_async(b_ptoff + i, d.pt_off + p0 + i, 4);
  };
  auto g3 = [&](int slot, int tb) {
    const int4 c1 = ring[slot][1], c2 = ring[slot][2];
    const int ns = c1.z, pq0 = c2.y, npq = c2.z;
    for (int k = i; k < ns * 21; k += NT) {
      const int s = k / 21;
      __pipeline_memcpy_async(b_cpre + k, reinterpret_cast<const double*>(d.campre + t_cam[tb][s]) + (k - 21 * s), 8);
    }
    if (i < ns) __pipeline_memcpy_async(t_fc[tb] + i, d.fc + t_cam[tb][i], 4);
    for (int k = i; k < ((pq0 & 1) + npq + 1) / 2; k += NT)
      __pipeline_memcpy_async(reinterpret_cast<uint32_t*>(b_pairs) + k, reinterpret_cast<const uint32_t*>(d.pairs + (pq0 & ~1)) + k, 4);
  };
  g1(0, 0); g2(0);
  __pipeline_commit();
  __pipeline_wait_prior(0);
  __syncthreads();
  g3(0, 0);
  __pipeline_commit();
  __pipeline_wait_prior(0);
  __syncthreads();
  for (int chunk = blockIdx.x; chunk < d.npchunk; chunk += stride, ++it) {
    const int cur = it & 1, nxt = cur ^ 1;
    const int rs = it % 3, rs1 = (it + 1) % 3, rs2 = (it + 2) % 3;
    const bool has_next = chunk + stride < d.npchunk;
    const int4 c0 = ring[rs][0], c1 = ring[rs][1], c2 = ring[rs][2];
    const int p0 = c0.x, np = c0.y, e0 = c0.z, n = c0.w;
    const int w = c1.x, ns = c1.z, lb0 = c1.w, lb1 = c1.w + c2.x;
    const int pq0 = c2.y, npq = c2.z;
    double* cam_st = smem;                                   // [n * 27]  A: camera-side terms;  C/D: Z [n * 19]
    double* pt_st = cam_st + (size_t)n * kCamStride;         // [n * 9]   A: V, gp terms;        C/D: pair lists
    double* Li = pt_st + (size_t)n * 9;                      // [np * 6]  Linv: 00 10 11 20 21 22
    double* us = Li + (size_t)np * 6;                        // [np * 3]  u = Linv gp
    uint8_t* opl = reinterpret_cast<uint8_t*>(us + (size_t)np * 3);   // [n] chunk-local point of each observation
    double* Z = cam_st + 1;                                  // one zero element in front of row 0 (see phase D)
    uint32_t* prs = reinterpret_cast<uint32_t*>(pt_st);      // pair lists as byte offsets of the two Z rows
    const double radius = d.w_radius[w];
    auto slot_of = [&](int sl, int& j0, int& j1) -> int {
      j0 = t_off[cur][sl] - e0; j1 = t_off[cur][sl + 1] - e0;
      return (t_fc[cur][sl] >= 0) ? t_pos[cur][sl] : -1;
    };
    auto sobs_at = [&](int j) -> int { return (int)t_sobs[cur][(e0 & 1) + j]; };
    auto prefetch_g1 = [&]() {
      g1(rs1, nxt);
      if (i < 3 && chunk + 2 * stride < d.npchunk) __pipeline_memcpy_async(&ring[rs2][i], d.cdesc + 4 * ((size_t)chunk + 2 * stride) + i, 16);
    };
    auto prefetch_g2 = [&]() { g2(rs1); };
    auto prefetch_g3 = [&]() { g3(rs1, nxt); };
    const bool act = i < n;
    int sl = 0, pl = 0;
    float4 m = make_float4(0.f, 0.f, 0.f, 0.f);
    if (act) {
      sl = b_oslot[(e0 & 3) + i];
      pl = b_ppt[i] - p0;
      m = b_uv[i];
      opl[i] = (uint8_t)pl;
    }
    // ---- A
    double cost = 0.0;
    double W[18];
#pragma unroll
    for (int k = 0; k < 18; ++k) W[k] = 0.0;
    if (act) {
      const CamPre& cp = *reinterpret_cast<const CamPre*>(b_cpre + 21 * sl);
      const float uv[4] = {m.x, m.y, m.z, m.w};
      const double X[3] = {b_pts[3 * pl], b_pts[3 * pl + 1], b_pts[3 * pl + 2]};
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
    // uv / ppt / oslot of this chunk are dead: group 1 of the next chunk goes into their place, its tables and the record
    // of the chunk after it into the other halves of theirs
    if (has_next) prefetch_g1();
    __pipeline_commit();
    // ---- B: per point V, gp, damped Cholesky
    double gmax = 0.0;
    const int q = threadIdx.x;
    const bool has_pt = q < np;
    double s2[3] = {1, 1, 1};
    int pe0 = 0, pe1 = 0;
    if (has_pt) {
#pragma unroll
      for (int k = 0; k < 3; ++k) s2[k] = b_sc2[3 * q + k];
      pe0 = b_ptoff[q] - e0; pe1 = b_ptoff[q + 1] - e0;
    }
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
        const double x = b_pts[3 * q + k];
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
    if (has_next) prefetch_g2();              // points, scaling and point offsets of this chunk are dead
    __pipeline_commit();
    if (threadIdx.x == 0) {
      double t = 0.0, mx = 0.0;
#pragma unroll
      for (int k = 0; k < NW; ++k) { t += sm1[k]; mx = fmax(mx, smx[k]); }
      d.lin_part[2 * chunk] = t; d.lin_part[2 * chunk + 1] = mx;
    }
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
    for (int k = threadIdx.x; k < npq; k += NT) {
      const unsigned pr = b_pairs[(pq0 & 1) + k];
      prs[k] = (pr & 255u) * kRowBytes | ((pr >> 8) * kRowBytes) << 16;
    }
    __pipeline_wait_prior(0);                 // this thread's groups 1 and 2 have landed ...
    __syncthreads();                          // ... and everybody's: the next chunk's slot cameras are readable
    if (has_next) prefetch_g3();              // its camera rows and free-camera flags (second level), its pairs
    __pipeline_commit();
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
      const int q0 = t_lbq[cur][lb - lb0] - pq0, q1 = t_lbq[cur][lb - lb0 + 1] - pq0;
      double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
      auto mma_off = [&](unsigned en, double& x0, double& x1) {
        const double a = *reinterpret_cast<const double*>(Zl + (en & 0xffffu));
        const double b = *reinterpret_cast<const double*>(Zl + (en >> 16));
        dmma884(x0, x1, a, b);
      };
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
      if (feeds) *reinterpret_cast<double2*>(d.spart + (size_t)t_lbpos[cur][lb - lb0] * 36 + 6 * g + 2 * t) = make_double2(a0 + b0, a1 + b1);
    }
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
    // the next chunk's phase A overwrites what this chunk's last phases read, and reads what group 3 brings
    __pipeline_wait_prior(0);
    __syncthreads();
  }
}

}  // namespace sba
