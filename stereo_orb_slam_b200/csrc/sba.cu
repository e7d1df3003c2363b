// sba.cu — handle, device memory, launch sequencing and the C-ABI of libsba.so (see include/sba.h).
//
// Replaces the Ceres solve inside BundleAdjuster::Optimize (reference: src/bundle_adjuster.cpp:116-117)
// with an on-device Levenberg-Marquardt loop: linearise (K0-K2) -> damped Schur complement (K3)
// -> block-Jacobi PCG (K4) -> back-substitution (K5) -> candidate cost (K6) -> per-window
// trust-region control.  The host only sequences launches and reads one counter per iteration.
#include <cuda_runtime.h>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/sba.h"
#include "kernels.cuh"
#include "sba_comm.h"
#include "sba_internal.h"

using namespace sba;

namespace {
thread_local std::string g_last_error;
void set_error(const std::string& s) { g_last_error = s; }

#define CUDA_TRY(expr)                                                                              \
  do {                                                                                              \
    cudaError_t _e = (expr);                                                                        \
    if (_e != cudaSuccess) {                                                                        \
      set_error(std::string(#expr) + ": " + cudaGetErrorString(_e));                                \
      return SBA_CUDA_ERROR;                                                                        \
    }                                                                                               \
  } while (0)

inline int cdiv(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
}  // namespace

struct sba_handle_s {
  int device = -1;            // -1: host-only handle (structure build / introspection only)
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[8] = {};
  HostStructure H;
  bool has_structure = false, has_state = false, has_meas = false, jacobi_done = false;
  int32_t shard_begin = -1, shard_end = -1;
  Dev d{};
  DevParams P{};
  std::vector<void*> allocs;          // structure + work buffers (freed on re-structure)
  float4* uv_in = nullptr;            // staging for the caller-ordered observations
  double* rsum = nullptr;
  double* r_dbg = nullptr;
  int32_t* h_counters = nullptr;      // pinned
  int64_t launches = 0;
  int64_t nobs_all = 0;
  sba::Comm comm;                     // NCCL (multi-GPU), inactive on 1 GPU
  double time_structure_s = 0.0;
  double last_radius = 1e4;

  template <typename T>
  int alloc(T** p, size_t n) {
    *p = nullptr;
    if (n == 0) n = 1;
    void* q = nullptr;
    cudaError_t e = cudaMalloc(&q, n * sizeof(T));
    if (e != cudaSuccess) { set_error(std::string("cudaMalloc: ") + cudaGetErrorString(e)); return SBA_CUDA_ERROR; }
    allocs.push_back(q);
    *p = static_cast<T*>(q);
    return SBA_OK;
  }
  template <typename T>
  int upload(const T** dst, const std::vector<T>& v) {
    T* p = nullptr;
    int rc = alloc(&p, v.size());
    if (rc) return rc;
    if (!v.empty()) {
      cudaError_t e = cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, stream);
      if (e != cudaSuccess) { set_error(std::string("H2D: ") + cudaGetErrorString(e)); return SBA_CUDA_ERROR; }
    }
    *dst = p;
    return SBA_OK;
  }
  void free_all() {
    for (void* p : allocs) cudaFree(p);
    allocs.clear();
    uv_in = nullptr; rsum = nullptr; r_dbg = nullptr;
    d = Dev{};
  }
};

#define LAUNCH(h, kernel, grid, block, ...)                                  \
  do {                                                                       \
    if ((grid) > 0) {                                                        \
      kernel<<<(grid), (block), 0, (h)->stream>>>(__VA_ARGS__);              \
      (h)->launches++;                                                       \
    }                                                                        \
  } while (0)

extern "C" {

void sba_default_options(sba_options* o) {
  o->max_iterations = 50; o->max_time_s = 1.0; o->huber_delta = 1.0;
  o->point_lower_bound = -10000.0; o->point_upper_bound = 10000.0;
  o->parameter_tolerance = 1e-8; o->function_tolerance = 1e-16; o->gradient_tolerance = 1e-16;
  o->residual_dim = 4; o->initial_radius = 1e4; o->max_pcg_iterations = 500; o->pcg_tolerance = 1e-10;
  o->verbose = 0;
}

const char* sba_status_string(int s) {
  switch (s) {
    case SBA_OK: return "OK";
    case SBA_NO_CONVERGENCE: return "NO_CONVERGENCE";
    case SBA_NUMERICAL_FAILURE: return "NUMERICAL_FAILURE";
    case SBA_INVALID_ARGUMENT: return "INVALID_ARGUMENT";
    case SBA_CUDA_ERROR: return "CUDA_ERROR";
    case SBA_NCCL_ERROR: return "NCCL_ERROR";
  }
  return "UNKNOWN";
}
const char* sba_last_error(void) { return g_last_error.c_str(); }

int sba_create(int device, sba_handle* out) {
  if (!out) return SBA_INVALID_ARGUMENT;
  *out = nullptr;
  sba_handle h = new sba_handle_s();
  h->device = device;
  if (device >= 0) {
    // No CPU fallback: a handle on a device needs a working CUDA device, loudly.
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    for (int i = 0; i < 8 && e == cudaSuccess; ++i) e = cudaEventCreate(&h->ev[i]);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&h->h_counters, 4 * sizeof(int32_t));
    if (e != cudaSuccess) {
      set_error(std::string("sba_create: no usable CUDA device: ") + cudaGetErrorString(e));
      delete h;
      return SBA_CUDA_ERROR;
    }
  }
  *out = h;
  return SBA_OK;
}

void sba_destroy(sba_handle h) {
  if (!h) return;
  if (h->device >= 0) {
    cudaSetDevice(h->device);
    h->comm.destroy();
    h->free_all();
    for (int i = 0; i < 8; ++i) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    if (h->h_counters) cudaFreeHost(h->h_counters);
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
}

int sba_set_point_shard(sba_handle h, int32_t pt_begin, int32_t pt_end) {
  if (!h || pt_begin < 0 || pt_end < pt_begin) return SBA_INVALID_ARGUMENT;
  h->shard_begin = pt_begin; h->shard_end = pt_end;
  return SBA_OK;
}

int sba_set_structure(sba_handle h, int32_t Nc, int32_t Np, int64_t Nobs, const int32_t* obs_cam,
                      const int32_t* obs_pt, const uint8_t* cam_fixed, int32_t nwin,
                      const int32_t* win_cam_off, const int32_t* win_pt_off) {
  if (!h || (Nobs > 0 && (!obs_cam || !obs_pt))) { set_error("null argument"); return SBA_INVALID_ARGUMENT; }
  auto t0 = std::chrono::steady_clock::now();
  h->has_structure = h->has_state = h->has_meas = h->jacobi_done = false;
  const int32_t pb = h->shard_begin >= 0 ? h->shard_begin : 0;
  const int32_t pe = h->shard_begin >= 0 ? std::min(h->shard_end, Np) : Np;
  std::string err = build_structure(Nc, Np, Nobs, obs_cam, obs_pt, cam_fixed, nwin, win_cam_off, win_pt_off, pb, pe, &h->H);
  if (!err.empty()) { set_error("sba_set_structure: " + err); return SBA_INVALID_ARGUMENT; }
  h->nobs_all = Nobs;
  h->has_structure = true;
  h->time_structure_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (h->device < 0) return SBA_OK;

  CUDA_TRY(cudaSetDevice(h->device));
  h->free_all();
  const HostStructure& H = h->H;
  Dev& d = h->d;
  d.Nc = H.Nc; d.Np = H.Np; d.Nf = H.Nf; d.nwin = H.nwin; d.Nobs = H.Nobs;
  d.nbu = (int32_t)H.blk_row.size(); d.npchunk = (int32_t)H.pchunk_begin.size();
  d.ncchunk = (int32_t)H.cchunk_cam.size(); d.nschunk = (int32_t)H.schunk_blk.size();
  int rc = 0;
#define UP(field) if ((rc = h->upload(&d.field, H.field))) return rc;
  UP(fc) UP(free2cam) UP(cam_win) UP(fcam_win) UP(win_cam_off) UP(win_fc_off) UP(win_pchunk_off)
  UP(pt_off) UP(perm_p) UP(pcam) UP(ppt) UP(cam_off) UP(cpos) UP(cpt)
  UP(pchunk_begin) UP(pchunk_end) UP(pchunk_win)
  UP(cchunk_cam) UP(cchunk_begin) UP(cchunk_end) UP(cam_chunk_off)
  UP(blk_row) UP(blk_col) UP(blk_chunk_off) UP(schunk_blk) UP(diag_blk)
  UP(schunk_begin) UP(schunk_end) UP(ent_a) UP(ent_b) UP(row_ptr) UP(bsr_col) UP(bsr_blk)
#undef UP
  const size_t nobs = (size_t)H.Nobs, nc = (size_t)H.Nc, np = (size_t)H.Np, nf = (size_t)H.Nf, nw = (size_t)H.nwin;
#define AL(ptr, n) if ((rc = h->alloc(&ptr, (n)))) return rc;
  float4 *uvp, *uvc;
  AL(uvp, nobs) AL(uvc, nobs) d.uv_p = uvp; d.uv_c = uvc;
  AL(h->uv_in, (size_t)Nobs)
  AL(d.poses, nc * 6) AL(d.poses_new, nc * 6) AL(d.pts, np * 3) AL(d.pts_new, np * 3)
  AL(d.campre, nc) AL(d.campre_new, nc)
  AL(d.W, nobs * 18) AL(d.V, np * 6) AL(d.gp, np * 3) AL(d.Vinv, np * 6) AL(d.tp, np * 3)
  AL(d.U, nc * 21) AL(d.gc, nc * 6) AL(d.cpart, (size_t)d.ncchunk * 27)
  AL(d.sc2c, nc * 6) AL(d.sc2p, np * 3) AL(d.lin_part, (size_t)d.npchunk * 2)
  AL(d.spart, (size_t)d.nschunk * 36) AL(d.Su, (size_t)d.nbu * 36) AL(d.rpart, (size_t)d.ncchunk * 6)
  AL(d.rhs, nf * 6) AL(h->rsum, nf * 6) AL(d.Minv, nf * 36)
  AL(d.x, nf * 6) AL(d.r, nf * 6) AL(d.z, nf * 6) AL(d.p, nf * 6) AL(d.Ap, nf * 6)
  AL(d.step_part, (size_t)d.npchunk * 4) AL(d.cost_part, (size_t)d.npchunk)
  AL(d.w_radius, nw) AL(d.w_decf, nw) AL(d.w_cost, nw) AL(d.w_cand, nw) AL(d.w_gmax, nw)
  AL(d.w_gd, nw) AL(d.w_dDd, nw) AL(d.w_dr, nw) AL(d.w_step2, nw) AL(d.w_x2, nw)
  AL(d.w_rz, nw) AL(d.w_rz0, nw) AL(d.w_pAp, nw) AL(d.w_alpha, nw) AL(d.w_beta, nw)
  AL(d.w_status, nw) AL(d.w_iter, nw) AL(d.w_succ, nw) AL(d.w_invalid, nw) AL(d.w_accept, nw)
  AL(d.w_fail, nw) AL(d.w_pcg_done, nw) AL(d.w_pcg_it, nw) AL(d.counters, 4)
#undef AL
  // everything the kernels may read before it is first written starts at zero
  CUDA_TRY(cudaMemsetAsync(d.x, 0, std::max<size_t>(1, nf * 6) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.r, 0, std::max<size_t>(1, nf * 6) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.w_status, 0, nw * 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.w_fail, 0, nw * 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.w_accept, 0, nw * 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.counters, 0, 16, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.pts, 0, std::max<size_t>(1, np * 3) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.pts_new, 0, std::max<size_t>(1, np * 3) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.V, 0, std::max<size_t>(1, np * 6) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.gp, 0, std::max<size_t>(1, np * 3) * 8, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  h->time_structure_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  return SBA_OK;
}
