// sba.cu — handle, device memory, launch sequencing and the C-ABI of libsba.so (see include/sba.h).
//
// Replaces the Ceres solve inside BundleAdjuster::Optimize (reference: src/bundle_adjuster.cpp:116-117)
// with an on-device Levenberg-Marquardt loop: linearise (K0-K2) -> damped Schur complement (K3)
// -> block-Jacobi PCG (K4) -> back-substitution (K5) -> candidate cost (K6) -> per-window
// trust-region control.  The host only sequences launches and reads one counter per iteration.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <limits>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/sba.h"
#include "kernels.cuh"
#include "sba_comm.h"
#include "sba_internal.h"
#include "pgo.cuh"
#include "ransac.cuh"
#include "fused_persist.cuh"
#include "sparse_chol.cuh"
#include "sparse_chol.h"

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
// SBA_DEBUG_MASK: phase knock-outs for timing experiments (results are then wrong); 0 in any real run
inline int dbg_mask() { const char* e = std::getenv("SBA_DEBUG_MASK"); return e ? std::atoi(e) : 0; }
}  // namespace

// Device-resident map (sba_map_*): what the reference keeps in m_cam_frames / m_ldm_points
// (src/bundle_adjuster.h:21-22).  Floats that only the solver reads live on the GPU (map points, stereo pixels); the
// integer side (which point every observation sees) and the poses stay mirrored on the host, where the index build
// and the f32 pose (de)parameterisation of the reference run.
struct DeviceMap {
  std::vector<float> T_wc;              // [N*16] world-from-camera, host mirror (Frame::GlobalPose)
  std::vector<int64_t> frame_off{0};    // [N+1]
  std::vector<uint32_t> obs_point_id;   // [K]
  float* d_points = nullptr; size_t n_points = 0, cap_points = 0;   // [M*3] f32 (MapPoint::Position)
  float4* d_uv = nullptr; size_t n_obs = 0, cap_obs = 0;            // [K] (u_l, v_l, u_r, v_r)
  uint32_t* d_ids = nullptr; size_t cap_ids = 0;                    // global ids of the last window's points
  // scratch of the last window (host)
  std::vector<uint32_t> ids; std::vector<int32_t> cam, pt; std::vector<double> poses;
  int32_t last_start = 0, last_end = 0;
  double ms_pack = 0.0, ms_structure = 0.0, ms_upload = 0.0, ms_solve = 0.0, ms_unpack = 0.0;
};

struct sba_handle_s {
  DeviceMap map;
  int device = -1;            // -1: host-only handle (structure build / introspection only)
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[12] = {};     // [8..11]: the second pair of linearisation events of the run-ahead LM loop
  // The camera-side tail of a linearisation (k_assemble_cams, k_lin_finish, k_lm_begin: a few dozen small CTAs)
  // runs on `side`, forked after k_lin_chunk, while the next k_schur_chunk fills the SMs on `stream`;
  // join_tail() makes `stream` wait for it before the first consumer of U / gc / the window scalars.
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool tail_pending = false;
  HostStructure H;
  bool has_structure = false, has_state = false, has_meas = false, jacobi_done = false;
  int32_t shard_begin = -1, shard_end = -1;
  Dev d{};
  DevParams P{};
  cudaMemPool_t pool = nullptr;       // the handle's own stream-ordered pool (no process-wide side effects)
  std::vector<void*> allocs;          // structure + work buffers (freed on re-structure)
  float4* uv_in = nullptr;            // staging for the caller-ordered observations
  double* r_dbg = nullptr;
  unsigned char* flush = nullptr;     // 256 MiB L2-flush scratch (sba_time_stage)
  int step_launches = 0;
  cudaGraphExec_t step_graph = nullptr;   // one K1-K3 step (stage 8) captured once per structure: a single launch per step
  int32_t* h_counters = nullptr;      // pinned
  unsigned long long* stamps = nullptr;   // SBA_DEBUG_LS=2
  uint8_t* stage = nullptr; size_t stage_bytes = 0; cudaEvent_t ev_stage = nullptr; bool stage_busy = false;   // (also the packed upload of sba_set_structure)   // pinned staging of sba_ransac_stereo (grown on demand)
  int64_t launches = 0;
  int64_t nobs_all = 0;
  sba::Comm comm;                     // NCCL (multi-GPU), inactive on 1 GPU
  double time_structure_s = 0.0;
  double last_radius = 1e4;
  // single-launch PCG (k_pcg_window): usable when every window's system fits one CTA
  bool pcg_window_ok = false; int pcg_s_in_smem = 0; size_t pcg_smem = 0; int pcg_threads = 1024;
  int coop_blocks = 0;        // co-resident CTAs of k_step_loop (0: cooperative launch unavailable)
  int coop_pcg_blocks = 0;    // idem for k_pcg_coop
  double* pcg_pbuf = nullptr; double* pcg_rowdot = nullptr;
  bool pcg_cluster_ok = false; size_t pcg_cluster_smem = 0; int pcg_rows_cap = 0, pcg_ent_cap = 0;   // one 8-CTA cluster per window (k_pcg_cluster)
  bool chol_ok = false, chol_in_smem = false; size_t chol_smem = 0; int chol_threads = kCholThreads;   // one CTA per window factors S exactly (k_chol_window)
  bool chol_cluster_ok = false; size_t chol_cluster_smem = 0;   // mid-size windows: one 8-CTA cluster per window (k_chol_cluster)
  bool chol_band_ok = false; size_t chol_band_smem = 0; const int32_t* band_last = nullptr;   // banded windows: envelope in one CTA's shared memory (k_chol_band)
  size_t lin_smem = 0, schur_smem = 0;   // dynamic shared memory of the per-chunk kernels (largest chunk)
  size_t fused_smem = 0; bool fused_ok = false; int fused_occ = 2; int fused_nt = 256;
  size_t backsub_smem = 0;     // k_backsub: W rows of the largest regular chunk
  bool persist_ok = false; int persist_grid = 0;      // k_fused_persist (persistent CTAs, inputs of the next chunk prefetched)   // k_fused_chunk (K1-K3 of a chunk in one CTA)
  bool schur_fresh = false;              // the last linearisation was fused: spart / rpart already hold this radius' Schur partials
  // ... and on a single GPU its camera-side assembly and finish are left to the next do_schur (one assemble kernel for
  // U / gc / S / rhs; cost + gradient test on the side stream while the solve runs)
  std::vector<int32_t> asm_ids_host;
  bool every_window_free = false;     // every window has at least one free camera (k_assemble_schur can finish the windows itself)
  bool cams_pending = false; bool pend_keep_cost = false; int pend_lm_begin = -1; double pend_radius = 0.0;
  // exact block-sparse Cholesky of the reduced system for windows of more than kCholMaxCams free cameras (k_spchol)
  SparseCholSymbolic sym; bool sym_ok = false; SpCholDev sp{}; bool spchol_ok = false; int spchol_blocks = 0; size_t spchol_smem = 0; double time_symbolic_s = 0.0;

  template <typename T>
  int alloc(T** p, size_t n) {
    *p = nullptr;
    if (n == 0) n = 1;
    void* q = nullptr;
    // stream-ordered pool: after the first solve a re-structure (every BundleAdjuster::Optimize call) reuses the
    // pool's cached blocks instead of paying ~100 cudaMalloc/cudaFree round trips
    // (+16 bytes: k_fused_persist copies byte / 16-bit arrays in aligned 32-bit words, the last of which may straddle the end)
    cudaError_t e = pool ? cudaMallocFromPoolAsync(&q, n * sizeof(T) + 16, pool, stream) : cudaMallocAsync(&q, n * sizeof(T) + 16, stream);
    if (e != cudaSuccess) { set_error(std::string("cudaMallocAsync: ") + cudaGetErrorString(e)); return SBA_CUDA_ERROR; }
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
    for (void* p : allocs) cudaFreeAsync(p, stream);
    allocs.clear();
    uv_in = nullptr; r_dbg = nullptr; pcg_pbuf = nullptr; pcg_rowdot = nullptr; flush = nullptr;
    if (step_graph) { cudaGraphExecDestroy(step_graph); step_graph = nullptr; }
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

// SBA_DEBUG_LS=2: device-side time stamps between the kernels of an LM iteration, on both streams (tuning aid)
__global__ void k_stamp(unsigned long long* buf, int idx) { buf[idx] = gtimer(); }
static void stamp(sba_handle h, cudaStream_t st, int idx) {
  if (h->P.debug != 2) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(h->stream, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return;
  if (!h->stamps) { if (cudaMalloc((void**)&h->stamps, 32 * 8) != cudaSuccess) return; cudaMemset(h->stamps, 0, 32 * 8); }
  k_stamp<<<1, 1, 0, st>>>(h->stamps, idx);
}

extern "C" {

void sba_default_options(sba_options* o) {
  o->max_iterations = 50; o->max_time_s = 1.0; o->huber_delta = 1.0;
  o->point_lower_bound = -10000.0; o->point_upper_bound = 10000.0;
  o->parameter_tolerance = 1e-8; o->function_tolerance = 1e-16; o->gradient_tolerance = 1e-16;
  o->residual_dim = 4; o->initial_radius = 1e4; o->max_pcg_iterations = 500; o->pcg_tolerance = 1e-12; o->linear_solver = 0;
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
    for (int i = 0; i < 12 && e == cudaSuccess; ++i) e = cudaEventCreate(&h->ev[i]);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_stage, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
    if (e == cudaSuccess && !std::getenv("SBA_NO_COOP")) {
      int coop = 0, per_sm = 0, sms = 0;
      cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, device);
      cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
      if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_step_loop, kThreads, 0) == cudaSuccess)
        h->coop_blocks = per_sm * sms;
      if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg_coop, kThreads, 0) == cudaSuccess)
        h->coop_pcg_blocks = per_sm * sms;
      if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_spchol, kSpThreads, 0) == cudaSuccess)
        h->spchol_blocks = std::min(per_sm, 1) * sms;
    }
    if (e == cudaSuccess) {      // a private pool that keeps its freed blocks (no trim at synchronisation points): a
                                 // re-structure on the next Optimize call reuses them; the device's default pool is untouched
      cudaMemPoolProps props = {};
      props.allocType = cudaMemAllocationTypePinned;
      props.handleTypes = cudaMemHandleTypeNone;
      props.location.type = cudaMemLocationTypeDevice;
      props.location.id = device;
      if (cudaMemPoolCreate(&h->pool, &props) == cudaSuccess) {
        unsigned long long keep = ~0ull;
        cudaMemPoolSetAttribute(h->pool, cudaMemPoolAttrReleaseThreshold, &keep);
      } else {
        h->pool = nullptr;       // falls back to the default pool with its default (trimming) threshold
        cudaGetLastError();
      }
    }
    if (e == cudaSuccess) e = cudaMallocHost((void**)&h->h_counters, 8 * sizeof(int32_t));
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
    if (h->side) cudaStreamSynchronize(h->side);
    h->comm.destroy();
    h->free_all();
    if (h->map.d_points) cudaFree(h->map.d_points);
    if (h->map.d_uv) cudaFree(h->map.d_uv);
    if (h->map.d_ids) cudaFree(h->map.d_ids);
    if (h->stream) cudaStreamSynchronize(h->stream);     // the frees above are stream-ordered
    if (h->pool) cudaMemPoolDestroy(h->pool);
    for (int i = 0; i < 12; ++i) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
    if (h->h_counters) cudaFreeHost(h->h_counters);
    if (h->stage) cudaFreeHost(h->stage);
    if (h->ev_stage) cudaEventDestroy(h->ev_stage);
    if (h->stamps) cudaFree(h->stamps);
    if (h->side) cudaStreamDestroy(h->side);
    if (h->ev_fork) cudaEventDestroy(h->ev_fork);
    if (h->ev_join) cudaEventDestroy(h->ev_join);
    if (h->stream) cudaStreamDestroy(h->stream);
  }
  delete h;
}

int sba_set_point_shard(sba_handle h, int32_t pt_begin, int32_t pt_end) {
  if (!h || pt_begin < 0 || pt_end < pt_begin) return SBA_INVALID_ARGUMENT;
  h->shard_begin = pt_begin; h->shard_end = pt_end;
  return SBA_OK;
}

// Device side of a symbolic factorisation: index tables, L / Linv / y storage (through the handle's allocation list),
// the chain-panel capacity and the cooperative grid of k_spchol.
static int setup_spchol(sba_handle h, const SparseCholSymbolic& Y, SpCholDev* spo, size_t* smem, int* blocks) {
  SpCholDev& sp = *spo;
  int rc;
  sp.n = Y.n; sp.nlev = Y.nlev; sp.nnzL = (int32_t)Y.nnzL();
  if ((rc = h->upload(&sp.perm, Y.perm))) return rc;
  if ((rc = h->upload(&sp.col_ptr, Y.col_ptr))) return rc;
  if ((rc = h->upload(&sp.row_idx, Y.row_idx))) return rc;
  if ((rc = h->upload(&sp.init_src, Y.init_src))) return rc;
  if ((rc = h->upload(&sp.upd_ptr, Y.upd_ptr))) return rc;
  if ((rc = h->upload(&sp.upd_ext_end, Y.upd_ext_end))) return rc;
  if ((rc = h->upload(&sp.rowl_ext_end, Y.rowl_ext_end))) return rc;
  if ((rc = h->upload(&sp.upd_a, Y.upd_a))) return rc;
  if ((rc = h->upload(&sp.upd_b, Y.upd_b))) return rc;
  if ((rc = h->upload(&sp.rowl_ptr, Y.rowl_ptr))) return rc;
  if ((rc = h->upload(&sp.rowl_blk, Y.rowl_blk))) return rc;
  if ((rc = h->upload(&sp.rowl_col, Y.rowl_col))) return rc;
  if ((rc = h->upload(&sp.lvl_ptr, Y.lvl_ptr))) return rc;
  if ((rc = h->upload(&sp.lvl_cols, Y.lvl_cols))) return rc;
  if ((rc = h->upload(&sp.lvl_len, Y.lvl_len))) return rc;
  if ((rc = h->alloc(&sp.Lv, (size_t)Y.nnzL() * 36))) return rc;
  if ((rc = h->alloc(&sp.Li, (size_t)Y.n * 36))) return rc;
  if ((rc = h->alloc(&sp.y, (size_t)Y.n * 6))) return rc;
  // chains of up to panel_cap blocks are factored out of shared memory; size the launch for the largest that fits
  int cap = 1;
  for (size_t c = 0; c < Y.lvl_cols.size(); ++c) {
    const int nb = Y.col_ptr[Y.lvl_cols[c] + Y.lvl_len[c]] - Y.col_ptr[Y.lvl_cols[c]];
    if (nb <= kSpPanelMax && Y.lvl_len[c] <= kSpChainMax) cap = std::max(cap, nb);
  }
  sp.panel_cap = cap;
  *smem = spchol_smem_bytes(cap);
  // phase A of every level (products and right-hand-side terms that come from OUTSIDE a chain) is spread over the whole
  // grid: flat lists of the blocks / columns of the level's shared-memory chains
  {
    std::vector<int32_t> bptr((size_t)Y.nlev + 1, 0), cptr((size_t)Y.nlev + 1, 0), blks, cols;
    for (int l = 0; l < Y.nlev; ++l) {
      for (int c = Y.lvl_ptr[l]; c < Y.lvl_ptr[l + 1]; ++c) {
        const int j0 = Y.lvl_cols[c], jn = j0 + Y.lvl_len[c];
        const int b0 = Y.col_ptr[j0], b1 = Y.col_ptr[jn];
        if (b1 - b0 > cap || Y.lvl_len[c] > kSpChainMax) continue;       // long chains keep the column-by-column path
        for (int b = b0; b < b1; ++b) if (Y.upd_ext_end[b] > Y.upd_ptr[b]) blks.push_back(b);
        for (int j = j0; j < jn; ++j) cols.push_back(j);
      }
      bptr[l + 1] = (int32_t)blks.size(); cptr[l + 1] = (int32_t)cols.size();
    }
    if ((rc = h->upload(&sp.pa_blk_ptr, bptr))) return rc;
    if ((rc = h->upload(&sp.pa_blk, blks))) return rc;
    if ((rc = h->upload(&sp.pa_col_ptr, cptr))) return rc;
    if ((rc = h->upload(&sp.pa_col, cols))) return rc;
  }
  // (the attribute is per function, not per launch: always the largest panel, so that handles and PGO calls with
  // different capacities never lower it under each other)
  CUDA_TRY(cudaFuncSetAttribute(k_spchol, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)spchol_smem_bytes(kSpPanelMax)));
  int per_sm = 0, sms = 0;
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
  CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_spchol, kSpThreads, *smem));
  *blocks = std::min(per_sm, 1) * sms;
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
  h->sym_ok = false; h->sym = SparseCholSymbolic();
  std::future<void> sym_job;
  {
    // windows of more than kCholMaxCams free cameras (global BA): symbolic phase of the exact block-sparse Cholesky —
    // nested-dissection ordering, elimination-tree levels, static update lists — once per structure
    int nf_max = 0;
    for (int w = 0; w < h->H.nwin; ++w) nf_max = std::max(nf_max, h->H.win_fc_off[w + 1] - h->H.win_fc_off[w]);
    h->every_window_free = h->H.nwin > 0;
    for (int w = 0; w < h->H.nwin; ++w) if (h->H.win_fc_off[w + 1] == h->H.win_fc_off[w]) h->every_window_free = false;
    if (h->H.Nf > 0 && (nf_max > kCholMaxCams || std::getenv("SBA_FORCE_SPARSE_CHOL")) && !std::getenv("SBA_NO_SPARSE_CHOL")) {
      auto symbolic = [h]() {
        const auto ts = std::chrono::steady_clock::now();
        const std::string serr = sparse_chol_symbolic(h->H.Nf, h->H.blk_row, h->H.blk_col, (int64_t)256 << 20, &h->sym);
        h->time_symbolic_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - ts).count();
        h->sym_ok = serr.empty();
        if (!h->sym_ok && std::getenv("SBA_TIMING")) std::fprintf(stderr, "[sba] sparse Cholesky unavailable: %s\n", serr.c_str());
      };
      // one host thread, 7 ms at KITTI-00 scale: on a device handle it runs beside the upload of the structure tables
      // (joined before the factorisation's own tables are set up; an early return joins in the future's destructor)
      if (h->device >= 0) sym_job = std::async(std::launch::async, symbolic); else symbolic();
    }
  }
  h->time_structure_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (h->device < 0) return SBA_OK;
  const double t_host = h->time_structure_s;

  CUDA_TRY(cudaSetDevice(h->device));
  h->free_all();
  const HostStructure& H = h->H;
  Dev& d = h->d;
  d.Nc = H.Nc; d.Np = H.Np; d.Nf = H.Nf; d.nwin = H.nwin; d.Nobs = H.Nobs;
  d.nbu = (int32_t)H.blk_row.size(); d.npchunk = (int32_t)H.pchunk_begin.size();
  d.ncs = (int32_t)H.cs_cam.size(); d.nlb = (int32_t)H.lb_blk.size();
  int rc = 0;
  // The ~45 index tables go up in ONE copy when they are small (windows: a few MB): they are packed into the handle's
  // pinned staging buffer and land in one device allocation — a pageable cudaMemcpyAsync per table costs more than
  // the bytes it moves, and the per-frame calls of the reference's pattern (src/slam.cpp:122-129) pay it every frame.
  // Large structures (global BA: hundreds of MB) keep one copy per table straight from the host vectors.
  struct UpItem { const void** dst; const void* src; size_t bytes, off; };
  std::vector<UpItem> items;
  size_t up_total = 0;
  auto add = [&](auto** dst, const auto& vec) {
    const size_t bytes = vec.size() * sizeof(vec[0]);
    items.push_back({(const void**)dst, vec.data(), bytes, up_total});
    up_total += (bytes + 16 + 255) & ~(size_t)255;        // (+16: see alloc)
  };
  const int32_t* cdesc_i32 = nullptr;
#define UP(field) add(&d.field, H.field);
  UP(fc) UP(free2cam) UP(cam_win) UP(fcam_win) UP(win_cam_off) UP(win_fc_off) UP(win_pchunk_off)
  UP(pt_off) UP(perm_p) UP(pcam) UP(ppt)
  UP(pchunk_begin) UP(pchunk_end) UP(pchunk_win)
  add(&cdesc_i32, H.cdesc);                                                       // slots are 256-byte aligned: int4 loads
  UP(slot_off) UP(cs_cam) UP(cs_obs_off) UP(lb_off) UP(lb_blk) UP(lb_pair_off) UP(pairs) UP(oslot) UP(lb_sa) UP(lb_sb) UP(sobs)
  UP(blk_row) UP(blk_col) UP(diag_blk) UP(blk_part_off) UP(blk_part) UP(cam_part_off) UP(cam_part)
  UP(row_ptr) UP(bsr_col) UP(bsr_blk) UP(lb_pos) UP(cs_pos)
  d.nlong = (int32_t)H.long_chunks.size();
  if (d.nlong > 0) { UP(long_chunks) UP(long_zoff) UP(long_lb_off) }
#undef UP
  if (up_total <= ((size_t)16 << 20) && !std::getenv("SBA_NO_PACKED_UPLOAD")) {
    uint8_t* arena = nullptr;
    if ((rc = h->alloc(&arena, up_total))) return rc;
    if (h->stage_busy) { CUDA_TRY(cudaEventSynchronize(h->ev_stage)); h->stage_busy = false; }
    if (h->stage_bytes < up_total) {
      if (h->stage) cudaFreeHost(h->stage);
      h->stage = nullptr; h->stage_bytes = 0;
      const size_t want = std::max<size_t>(up_total * 2, (size_t)1 << 20);
      CUDA_TRY(cudaMallocHost((void**)&h->stage, want));
      h->stage_bytes = want;
    }
    for (const UpItem& it : items) {
      if (it.bytes) std::memcpy(h->stage + it.off, it.src, it.bytes);
      *it.dst = arena + it.off;
    }
    CUDA_TRY(cudaMemcpyAsync(arena, h->stage, up_total, cudaMemcpyHostToDevice, h->stream));
    CUDA_TRY(cudaEventRecord(h->ev_stage, h->stream));
    h->stage_busy = true;
  } else {
    for (const UpItem& it : items) {
      uint8_t* p = nullptr;
      if ((rc = h->alloc(&p, it.bytes))) return rc;
      if (it.bytes) CUDA_TRY(cudaMemcpyAsync(p, it.src, it.bytes, cudaMemcpyHostToDevice, h->stream));
      *it.dst = p;
    }
  }
  d.cdesc = reinterpret_cast<const int4*>(cdesc_i32);
  {
    // point-sharded ranks assemble only what they have partials for (a rank of 8 touches about an eighth of a
    // trajectory's cameras); built always, used on the sharded path
    std::vector<int32_t>& ids = h->asm_ids_host;          // (a member: the copy may read it after this function returns)
    ids.clear();
    for (int32_t b = 0; b < d.nbu; ++b) if (H.blk_part_off[b + 1] > H.blk_part_off[b]) ids.push_back(b);
    for (int32_t f = 0; f < d.Nf; ++f) { const int32_t c = H.free2cam[f]; if (H.cam_part_off[c + 1] > H.cam_part_off[c]) ids.push_back(d.nbu + f); }
    d.n_asm = (int32_t)ids.size();
    if ((rc = h->upload(&d.asm_ids, ids))) return rc;
  }
  const size_t nobs = (size_t)H.Nobs, nc = (size_t)H.Nc, np = (size_t)H.Np, nf = (size_t)H.Nf, nw = (size_t)H.nwin;
#define AL(ptr, n) if ((rc = h->alloc(&ptr, (n)))) return rc;
  float4* uvp;
  AL(uvp, nobs) d.uv_p = uvp;
  AL(h->uv_in, (size_t)Nobs)
  AL(d.poses, nc * 6) AL(d.poses_new, nc * 6) AL(d.pts, np * 3) AL(d.pts_new, np * 3)
  AL(d.campre_new, nc) AL(d.campre, nc)
  if (d.nlong > 0) AL(d.zlong, (size_t)H.long_zoff.back() * 18)
  AL(d.W, nobs * 18) AL(d.V, np * 6) AL(d.gp, np * 3) AL(d.Vinv, np * 6) AL(d.tp, np * 3)
  // U | gc | cost_red | gslots | (pad) | Su | rsum: ONE block.  Multi-GPU: one all-reduce per linearisation over the
  // first four, one per Schur step over the last two — or, after a fused linearisation, a single one over everything
  // (the never-written gap between them stays zero)
  const size_t lin_len = ((nc * 27 + nw + nw * kMaxRanks) + 31) & ~(size_t)31;
  AL(d.U, lin_len + (size_t)d.nbu * 36 + nf * 6)
  CUDA_TRY(cudaMemsetAsync(d.U, 0, (lin_len + (size_t)d.nbu * 36 + nf * 6) * sizeof(double), h->stream));
  d.gc = d.U + nc * 21; d.cost_red = d.gc + nc * 6; d.gslots = d.cost_red + nw;
  d.Su = d.U + lin_len;
  AL(d.cspart, (size_t)d.ncs * 27)
  AL(d.sc2c, nc * 6) AL(d.sc2p, np * 3) AL(d.lin_part, (size_t)d.npchunk * 2)
  AL(d.spart, (size_t)d.nlb * 36) AL(d.rpart, (size_t)d.ncs * 6)
  d.rsum = d.Su + (size_t)d.nbu * 36;      // Su | rsum: one block, one all-reduce per Schur step (multi-GPU)
  AL(d.rhs, nf * 6) AL(d.Minv, nf * 36)
  AL(d.x, nf * 6) AL(d.r, nf * 6) AL(d.z, nf * 6) AL(d.p, nf * 6) AL(d.Ap, nf * 6) AL(h->pcg_pbuf, nf * 6) AL(h->pcg_rowdot, 2 * nf)
  AL(d.dp, np * 3) AL(d.step_part, (size_t)d.npchunk * 4) AL(d.ls_part, (size_t)d.npchunk * 3)
  AL(d.cta_part, (size_t)std::max(h->coop_blocks, 1) * 16)
  AL(d.win_done, nw) AL(d.win_part, nw * 2)
  CUDA_TRY(cudaMemsetAsync(d.win_done, 0, nw * sizeof(int32_t), h->stream));
  AL(d.w_radius, nw) AL(d.w_decf, nw) AL(d.w_cost, nw) AL(d.w_gmax, nw)
  AL(d.w_gd, nw * 4 + nw * kMaxRanks)   // one block: a multi-GPU solve all-reduces the four step scalars and the max-norm slots in one call
  d.w_dDd = d.w_gd + nw; d.w_x2 = d.w_gd + 2 * nw; d.w_dr = d.w_gd + 3 * nw; d.dslots = d.w_gd + 4 * nw;
  AL(d.w_dmax, nw + 1)   // + one slot: the ranks' agreed time-out flag
  AL(d.w_ls_red, nw * 3) AL(d.w_step2, nw) AL(d.w_cand, nw)
  AL(d.w_ls_x, nw) AL(d.w_ls_px, nw) AL(d.w_ls_pf, nw) AL(d.w_ls_pg, nw)
  AL(d.w_ls_state, nw) AL(d.w_ls_it, nw) AL(d.w_ls_pvalid, nw)
  AL(d.w_rz, nw) AL(d.w_rz0, nw) AL(d.w_pAp, nw) AL(d.w_alpha, nw) AL(d.w_beta, nw)
  AL(d.w_status, nw) AL(d.w_iter, nw) AL(d.w_succ, nw) AL(d.w_invalid, nw) AL(d.w_accept, nw)
  AL(d.w_fail, nw) AL(d.w_pcg_done, nw) AL(d.w_pcg_it, nw) AL(d.counters, 8)
#undef AL
  // everything the kernels may read before it is first written starts at zero
  CUDA_TRY(cudaMemsetAsync(d.x, 0, std::max<size_t>(1, nf * 6) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.r, 0, std::max<size_t>(1, nf * 6) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.w_status, 0, nw * 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.w_fail, 0, nw * 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.w_accept, 0, nw * 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.w_ls_state, 0, nw * 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.counters, 0, 32, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.pts, 0, std::max<size_t>(1, np * 3) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.pts_new, 0, std::max<size_t>(1, np * 3) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.V, 0, std::max<size_t>(1, np * 6) * 8, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.gp, 0, std::max<size_t>(1, np * 3) * 8, h->stream));
  // camera-side sums are only written where a chunk contributes: the rest stays zero for good
  CUDA_TRY(cudaMemsetAsync(d.U, 0, std::max<size_t>(1, nc * 27 + nw + nw * kMaxRanks) * 8, h->stream));   // U | gc | cost_red | gslots
  CUDA_TRY(cudaMemsetAsync(d.Su, 0, std::max<size_t>(1, (size_t)d.nbu * 36 + nf * 6) * 8, h->stream));   // Su | rsum
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  {
    h->lin_smem = lin_smem_bytes(H.max_chunk_obs, H.max_chunk_pts, H.max_slots);
    h->schur_smem = schur_smem_bytes(H.max_chunk_obs, H.max_chunk_pts, H.max_chunk_pairs);
    CUDA_TRY(cudaFuncSetAttribute(k_lin_chunk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->lin_smem, 1024)));
    CUDA_TRY(cudaFuncSetAttribute(k_schur_chunk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->schur_smem, 1024)));
    h->fused_smem = fused_smem_bytes(H.max_chunk_obs, H.max_chunk_pts, H.max_slots);
    h->backsub_smem = (size_t)std::max(H.max_chunk_obs, 1) * 144;
    CUDA_TRY(cudaFuncSetAttribute(k_backsub, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->backsub_smem, 1024)));
    h->fused_ok = h->fused_smem <= (size_t)220 * 1024 && !std::getenv("SBA_NO_FUSED");
    h->schur_fresh = false;
    if (h->fused_ok)
    {
      h->fused_occ = (std::getenv("SBA_FUSED_OCC") && std::atoi(std::getenv("SBA_FUSED_OCC")) == 3) ? 3 : 2;   // tuning aid: 3 CTAs / SM (85 registers, spills)
      CUDA_TRY(cudaFuncSetAttribute(k_fused_chunk<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->fused_smem, 1024)));
      CUDA_TRY(cudaFuncSetAttribute(k_fused_chunk<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->fused_smem, 1024)));
      CUDA_TRY(cudaFuncSetAttribute((k_fused_chunk<4, false, 128>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->fused_smem, 1024)));
      h->fused_nt = (H.max_chunk_obs <= 128 && H.max_chunk_pts <= 128 && !std::getenv("SBA_FUSED_NT256")) ? 128 : 256;
      // persistent variant: every chunk must fit its fixed input buffers
      h->persist_ok = H.long_chunks.empty() && H.max_slots <= kFPSlots && H.max_chunk_obs <= 255 && !std::getenv("SBA_NO_PERSIST") && h->fused_nt == 256;
      for (size_t ch = 0; h->persist_ok && ch < H.pchunk_begin.size(); ++ch) {
        const int32_t* D = H.cdesc.data() + ch * 16;
        if (D[8] > kFPLb || D[10] > kFPPairs || (int64_t)D[10] * 4 > (int64_t)D[3] * 72) h->persist_ok = false;
      }
      if (h->persist_ok) {
        CUDA_TRY(cudaFuncSetAttribute(k_fused_persist, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->fused_smem, 1024)));
        int per_sm = 0, sms = 0;
        CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_fused_persist, kFusedThreads, h->fused_smem));
        h->persist_grid = per_sm * sms;
        if (per_sm < 1) h->persist_ok = false;
      }
    }
  }
  {
    // per-window size of the reduced system -> can one CTA hold it?
    int nf_max = 0; int64_t nb_max = 0;
    for (int w = 0; w < H.nwin; ++w) {
      const int f0 = H.win_fc_off[w], f1 = H.win_fc_off[w + 1];
      if (f1 <= f0) continue;
      nf_max = std::max(nf_max, f1 - f0);
      const int64_t b0 = H.diag_blk[f0], b1 = (f1 < H.Nf) ? H.diag_blk[f1] : (int64_t)H.blk_row.size();
      nb_max = std::max(nb_max, b1 - b0);
    }
    int64_t ne_max = 0;
    for (int w = 0; w < H.nwin; ++w) ne_max = std::max<int64_t>(ne_max, H.row_ptr[H.win_fc_off[w + 1]] - H.row_ptr[H.win_fc_off[w]]);
    const size_t lean = pcg_smem_bytes(nf_max, (int)nb_max, (int)ne_max, false);
    const size_t full = pcg_smem_bytes(nf_max, (int)nb_max, (int)ne_max, true);
    h->pcg_window_ok = (nf_max > 0 && lean <= (size_t)220 * 1024 && !std::getenv("SBA_FORCE_MULTILAUNCH"));   // multi-GPU: S is all-reduced, so every rank runs the same replicated solve
    h->pcg_s_in_smem = (full <= (size_t)220 * 1024) ? 1 : 0;
    h->pcg_smem = h->pcg_s_in_smem ? full : lean;
    h->pcg_threads = 32 * std::min(32, std::max(2, nf_max));
    // mid-size windows: a cluster per window (8 SMs' issue rate and L1 for the mat-vec)
    {
      int rows_cap = 0; int64_t ent_cap = 0;
      for (int w = 0; w < H.nwin; ++w) {
        const int g0 = H.win_fc_off[w], g1 = H.win_fc_off[w + 1];
        for (int r = 0; r < kPcgCluster; ++r) {
          int rows = 0; int64_t ent = 0;
          for (int f = g0 + r; f < g1; f += kPcgCluster) { ++rows; ent += H.row_ptr[f + 1] - H.row_ptr[f]; }
          rows_cap = std::max(rows_cap, rows); ent_cap = std::max(ent_cap, ent);
        }
      }
      h->pcg_rows_cap = rows_cap; h->pcg_ent_cap = (int)std::min<int64_t>(ent_cap, 1 << 30);
      if (pcg_cluster_smem_bytes(nf_max, rows_cap, h->pcg_ent_cap) > (size_t)200 * 1024) { h->pcg_rows_cap = 0; h->pcg_ent_cap = 0; }
    }
    h->pcg_cluster_smem = pcg_cluster_smem_bytes(nf_max, h->pcg_rows_cap, h->pcg_ent_cap);
    h->pcg_cluster_ok = (nf_max >= (std::getenv("SBA_FORCE_CLUSTER") ? 1 : 24) && H.nwin <= 64 && h->pcg_cluster_smem <= (size_t)200 * 1024 && !std::getenv("SBA_NO_CLUSTER") && !std::getenv("SBA_FORCE_MULTILAUNCH"));
    if (h->pcg_cluster_ok)
      CUDA_TRY(cudaFuncSetAttribute(k_pcg_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->pcg_cluster_smem, 1024)));
    if (h->pcg_window_ok)
      CUDA_TRY(cudaFuncSetAttribute(k_pcg_window, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->pcg_smem));
  }
  {
    // direct solve: every window small enough for the dense lower block-triangle workspace
    int nf_max = 0;
    std::vector<int64_t> off((size_t)H.nwin, 0);
    int64_t tot = 0;
    for (int w = 0; w < H.nwin; ++w) {
      const int nfw = H.win_fc_off[w + 1] - H.win_fc_off[w];
      nf_max = std::max(nf_max, nfw);
      off[w] = tot; tot += chol_ws_elems(nfw);
    }
    h->chol_ok = nf_max > 0 && nf_max <= kCholMaxCams && tot <= ((int64_t)1 << 31) && !std::getenv("SBA_NO_DIRECT");
    d.chol_ws = nullptr; d.chol_off = nullptr;
    if (h->chol_ok) {
      if ((rc = h->upload(&d.chol_off, off))) return rc;
      if ((rc = h->alloc(&d.chol_ws, (size_t)std::max<int64_t>(tot, 1)))) return rc;
      h->chol_smem = chol_smem_bytes(nf_max);
      h->chol_threads = nf_max <= 12 ? 128 : (nf_max <= 24 ? 256 : kCholThreads);   // trailing update of the first column: nf (nf - 1) / 2 * 6 rows
      h->chol_in_smem = nf_max <= kCholSmemCams;
      CUDA_TRY(cudaFuncSetAttribute(k_chol_window<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->chol_smem, 1024)));
      CUDA_TRY(cudaFuncSetAttribute(k_chol_window<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->chol_smem, 1024)));
      CUDA_TRY(cudaStreamSynchronize(h->stream));      // `off` is a local
    }
    // banded windows: monotone column envelope of every window (Cholesky fill stays inside it); when the largest fits
    // one CTA's shared memory the whole factorisation runs there (k_chol_band)
    h->chol_band_ok = false; h->band_last = nullptr;
    if (h->chol_ok && !std::getenv("SBA_NO_BAND")) {
      std::vector<int32_t> last((size_t)H.Nf);
      for (int f = 0; f < H.Nf; ++f) last[f] = f;
      for (size_t b = 0; b < H.blk_row.size(); ++b) last[H.blk_row[b]] = std::max(last[H.blk_row[b]], H.blk_col[b]);
      int64_t env_max = 0;
      for (int w = 0; w < H.nwin; ++w) {
        const int g0 = H.win_fc_off[w], g1 = H.win_fc_off[w + 1];
        int64_t env = 0;
        for (int k = g0; k < g1; ++k) {
          if (k > g0) last[k] = std::max(last[k], std::min(last[k - 1], g1 - 1));   // a column reaches at least as far as the previous one (it receives its fill)
          env += last[k] - k + 1;
        }
        env_max = std::max(env_max, env);
      }
      h->chol_band_smem = chol_band_smem_bytes(nf_max, (int)env_max);
      if (h->chol_band_smem <= (size_t)200 * 1024) {
        if ((rc = h->upload(&h->band_last, last))) return rc;
        CUDA_TRY(cudaFuncSetAttribute(k_chol_band, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->chol_band_smem, 1024)));
        CUDA_TRY(cudaStreamSynchronize(h->stream));      // `last` is a local
        h->chol_band_ok = true;
      }
    }
    h->chol_cluster_smem = chol_cluster_smem_bytes(std::max(nf_max, 1));
    h->chol_cluster_ok = h->chol_ok && nf_max >= (std::getenv("SBA_FORCE_CLUSTER") ? 1 : 16) && H.nwin <= 64 &&
                         h->chol_cluster_smem <= (size_t)220 * 1024 && !std::getenv("SBA_NO_CLUSTER");
    if (h->chol_cluster_ok)
      CUDA_TRY(cudaFuncSetAttribute(k_chol_cluster, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)std::max<size_t>(h->chol_cluster_smem, 1024)));
  }
  h->spchol_ok = false; h->sp = SpCholDev{};
  if (sym_job.valid()) sym_job.get();
  if (h->sym_ok && h->spchol_blocks > 0 && (!h->chol_ok || std::getenv("SBA_FORCE_SPARSE_CHOL"))) {
    {
      const SparseCholSymbolic& Y = h->sym;
      if ((rc = setup_spchol(h, Y, &h->sp, &h->spchol_smem, &h->spchol_blocks))) return rc;
      CUDA_TRY(cudaStreamSynchronize(h->stream));
      h->spchol_ok = h->spchol_blocks > 0;
      if (std::getenv("SBA_TIMING"))
        std::fprintf(stderr, "[sba] sparse Cholesky symbolic: %.3f ms, %d block columns, %lld blocks in L (%lld in S), %lld products, %lld chains in %d levels\n",
                     1e3 * h->time_symbolic_s, Y.n, (long long)Y.nnzL(), (long long)H.blk_row.size(), (long long)Y.nupd(), (long long)Y.nchain(), Y.nlev);
    }
  }
  h->time_structure_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (std::getenv("SBA_TIMING"))
    std::fprintf(stderr, "[sba] structure: host build %.3f ms, device alloc + upload %.3f ms (%zu allocations)\n", 1e3 * t_host,
                 1e3 * (h->time_structure_s - t_host), h->allocs.size());
  return SBA_OK;
}

// ---- helpers (still inside extern "C": static, not exported) ------------------------------------
static void fill_params(sba_handle h, const sba_options* o) {
  DevParams& P = h->P;
  P.huber_delta = o->huber_delta; P.lo = o->point_lower_bound; P.hi = o->point_upper_bound;
  P.ptol = o->parameter_tolerance; P.ftol = o->function_tolerance; P.gtol = o->gradient_tolerance;
  P.pcg_tol2 = o->pcg_tolerance * o->pcg_tolerance;
  P.rdim = (o->residual_dim == 3) ? 3 : 4;
  P.max_iterations = o->max_iterations;
  P.debug = std::getenv("SBA_DEBUG_LS") ? std::max(1, std::atoi(std::getenv("SBA_DEBUG_LS"))) : 0;
}

static int check_ready(sba_handle h, bool need_state) {
  if (!h) { set_error("null handle"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) { set_error("host-only handle: no CUDA device bound (there is no CPU fallback)"); return SBA_CUDA_ERROR; }
  if (!h->has_structure) { set_error("sba_set_structure has not been called"); return SBA_INVALID_ARGUMENT; }
  if (need_state && (!h->has_state || !h->has_meas)) { set_error("state / measurements not uploaded"); return SBA_INVALID_ARGUMENT; }
  return SBA_OK;
}

int sba_set_state(sba_handle h, const double* poses, const double* points) {
  int rc = check_ready(h, false);
  if (rc) return rc;
  const Dev& d = h->d;
  if ((d.Nc > 0 && !poses) || (d.Np > 0 && !points)) { set_error("null state"); return SBA_INVALID_ARGUMENT; }
  CUDA_TRY(cudaSetDevice(h->device));
  if (d.Nc > 0) CUDA_TRY(cudaMemcpyAsync(d.poses, poses, (size_t)d.Nc * 48, cudaMemcpyHostToDevice, h->stream));
  if (d.Np > 0) CUDA_TRY(cudaMemcpyAsync(d.pts, points, (size_t)d.Np * 24, cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  h->has_state = true; h->jacobi_done = false;
  return SBA_OK;
}

int sba_set_measurements(sba_handle h, const float* obs_uv, const double* p_l, const double* p_r) {
  int rc = check_ready(h, false);
  if (rc) return rc;
  if ((h->nobs_all > 0 && !obs_uv) || !p_l || !p_r) { set_error("null measurements"); return SBA_INVALID_ARGUMENT; }
  CUDA_TRY(cudaSetDevice(h->device));
  const Dev& d = h->d;
  if (h->nobs_all > 0)
    CUDA_TRY(cudaMemcpyAsync(h->uv_in, obs_uv, (size_t)h->nobs_all * 16, cudaMemcpyHostToDevice, h->stream));
  LAUNCH(h, k_permute_uv, cdiv(d.Nobs, 256), 256, d, h->uv_in, const_cast<float4*>(d.uv_p));
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  std::memcpy(h->P.pl, p_l, 96); std::memcpy(h->P.pr, p_r, 96);
  h->has_meas = true; h->jacobi_done = false;
  return SBA_OK;
}

int sba_get_state(sba_handle h, double* poses, double* points) {
  int rc = check_ready(h, false);
  if (rc) return rc;
  if (!h->has_state) { set_error("no state"); return SBA_INVALID_ARGUMENT; }
  const Dev& d = h->d;
  CUDA_TRY(cudaSetDevice(h->device));
  if (h->comm.active() && d.Np > 0) {
    // every rank holds the updated values of its own points only: zero the rest and sum over ranks
    const size_t b = (size_t)h->H.pt_begin * 3, e = (size_t)h->H.pt_end * 3, n = (size_t)d.Np * 3;
    if (b > 0) CUDA_TRY(cudaMemsetAsync(d.pts, 0, b * 8, h->stream));
    if (e < n) CUDA_TRY(cudaMemsetAsync(d.pts + e, 0, (n - e) * 8, h->stream));
    std::string err;
    if (!h->comm.allreduce(d.pts, n, false, h->stream, &err)) { set_error(err); return SBA_NCCL_ERROR; }
  }
  if (poses && d.Nc > 0) CUDA_TRY(cudaMemcpyAsync(poses, d.poses, (size_t)d.Nc * 48, cudaMemcpyDeviceToHost, h->stream));
  if (points && d.Np > 0) CUDA_TRY(cudaMemcpyAsync(points, d.pts, (size_t)d.Np * 24, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return SBA_OK;
}

#define NCCL_TRY(call)                                           \
  do {                                                           \
    std::string _err;                                            \
    if (!(call)) { set_error(_err); return SBA_NCCL_ERROR; }     \
  } while (0)

// K0 + K1 + K2 at the current state (d.poses / d.pts).
// `stream` waits for the forked tail of the last linearisation (no-op when none is pending)
// the fused K1-K3 chunk kernel: 128-thread CTAs (4 per SM) when every regular chunk has at most 128 observations
static void launch_fused(sba_handle h) {
  const Dev& d = h->d; const DevParams& P = h->P;
  if (h->persist_ok) k_fused_persist<<<std::min(h->persist_grid, (int)d.npchunk), kFusedThreads, h->fused_smem, h->stream>>>(d, P);
  else if (h->fused_nt == 128) k_fused_chunk<4, false, 128><<<d.npchunk, 128, h->fused_smem, h->stream>>>(d, P);
  else if (h->fused_occ == 3) k_fused_chunk<3><<<d.npchunk, kFusedThreads, h->fused_smem, h->stream>>>(d, P);
  else k_fused_chunk<2><<<d.npchunk, kFusedThreads, h->fused_smem, h->stream>>>(d, P);
  h->launches++;
}

// K5: one CTA per chunk, the chunk's rows of W by one bulk copy into its dynamic shared memory
static void launch_backsub(sba_handle h) {
  const Dev& d = h->d;
  if (d.npchunk <= 0) return;
  k_backsub<<<d.npchunk, kThreads, h->backsub_smem, h->stream>>>(d, h->P);
  h->launches++;
}

static int join_tail(sba_handle h) {
  if (!h->tail_pending) return SBA_OK;
  CUDA_TRY(cudaEventRecord(h->ev_join, h->side));
  CUDA_TRY(cudaStreamWaitEvent(h->stream, h->ev_join, 0));
  h->tail_pending = false;
  return SBA_OK;
}

// `fork_tail`: single GPU, Jacobi scaling already fixed: the camera-side tail goes to the side stream and the caller
// (the LM loop, the K1-K3 timing stage) joins after launching k_schur_chunk, which needs none of its results.
// `lm_begin` >= 0 appends k_lm_begin(first = lm_begin) to the tail.
// `fuse`: K3 of every chunk rides along (k_fused_chunk at the radii in d.w_radius; needs the Jacobi scaling): the next
// do_schur only assembles.
static int do_linearize(sba_handle h, bool keep_cost = false, bool fork_tail = false, int lm_begin = -1,
                        double initial_radius = 0.0, bool fuse = false) {
  Dev& d = h->d;
  const DevParams& P = h->P;
  int rc = join_tail(h);
  if (rc) return rc;
  const bool fused = fuse && h->fused_ok && h->jacobi_done && d.npchunk > 0;
  h->schur_fresh = false;
  LAUNCH(h, k_campre, cdiv(d.Nc, 128), 128, d);      // K0: camera table of this linearisation
  if (fused) {
    CUDA_TRY(cudaMemsetAsync(d.w_fail, 0, (size_t)d.nwin * 4, h->stream));      // the chunk kernel reports singular point blocks
    launch_fused(h);
    h->schur_fresh = true;
  } else if (d.npchunk > 0) { k_lin_chunk<<<d.npchunk, kLinThreads, h->lin_smem, h->stream>>>(d, P, dbg_mask() >> 8); h->launches++; }
  if (d.nlong > 0) {        // long tracks (> 255 observations of one point): their own kernels, same partial lists
    LAUNCH(h, k_lin_long, d.nlong, kLongThreads, d, P);
    if (fused) { LAUNCH(h, k_schur_long, d.nlong, kLongThreads, d); LAUNCH(h, k_schur_long_pairs, 296, 256, d); }
  }
  h->cams_pending = false;
  if (fused && fork_tail && !std::getenv("SBA_NO_FORK")) {      // single GPU and point-sharded alike
    h->cams_pending = true; h->pend_keep_cost = keep_cost; h->pend_lm_begin = lm_begin; h->pend_radius = initial_radius;
    CUDA_TRY(cudaGetLastError());
    return SBA_OK;
  }
  if (fork_tail && !h->comm.active() && h->jacobi_done && !std::getenv("SBA_NO_FORK")) {
    CUDA_TRY(cudaEventRecord(h->ev_fork, h->stream));
    CUDA_TRY(cudaStreamWaitEvent(h->side, h->ev_fork, 0));
    if (d.Nf > 0) { k_assemble_cams<<<d.Nf, kAsmThreads, 0, h->side>>>(d); h->launches++; }
    if (d.nwin > 0) { k_lin_finish<<<d.nwin, kThreads, 0, h->side>>>(d, P, keep_cost ? 1 : 0); h->launches++; }
    if (lm_begin >= 0) {
      // w_fail is cleared on `stream` (k_schur_chunk may set it concurrently with the tail)
      if (!fused) CUDA_TRY(cudaMemsetAsync(d.w_fail, 0, (size_t)d.nwin * 4, h->stream));
      k_lm_begin<<<cdiv(d.nwin, 128), 128, 0, h->side>>>(d, P, initial_radius, lm_begin, 0); h->launches++;
    }
    h->tail_pending = true;
    CUDA_TRY(cudaGetLastError());
    return SBA_OK;
  }
  LAUNCH(h, k_assemble_cams, d.Nf, kAsmThreads, d);
  if (h->comm.active()) {
    // ONE packed all-reduce per linearisation: U | gc | window cost | projected-gradient max (a slot per rank)
    const int world = h->comm.world;
    LAUNCH(h, k_lin_finish_local, d.nwin, kThreads, d, h->comm.rank, world);
    std::string _err;
    if (!h->comm.allreduce(d.U, (size_t)d.Nc * 27 + (size_t)d.nwin + (size_t)d.nwin * world, false, h->stream, &_err)) { set_error(_err); return SBA_NCCL_ERROR; }
    LAUNCH(h, k_lin_finish_global, d.nwin, kThreads, d, P, keep_cost ? 1 : 0, world);
  } else {
    LAUNCH(h, k_lin_finish, d.nwin, kThreads, d, P, keep_cost ? 1 : 0);
  }
  if (!h->jacobi_done) {
    LAUNCH(h, k_jacobi_scale, cdiv((int64_t)d.Nc * 6 + (int64_t)d.Np * 3, 256), 256, d);
    h->jacobi_done = true;
  }
  if (lm_begin >= 0) LAUNCH(h, k_lm_begin, cdiv(d.nwin, 128), 128, d, P, initial_radius, lm_begin, fused ? 0 : 1);
  CUDA_TRY(cudaGetLastError());
  return SBA_OK;
}

// K3: damped Schur complement at the per-window radii in d.w_radius.
static int do_schur(sba_handle h, bool exact_solver = false) {
  Dev& d = h->d;
  if (h->schur_fresh) h->schur_fresh = false;      // the fused linearisation already left this radius' partials
  else if (d.npchunk > 0) {
    k_schur_chunk<<<d.npchunk, kSchurThreads, h->schur_smem, h->stream>>>(d, h->H.max_chunk_pairs, dbg_mask()); h->launches++;
    if (d.nlong > 0) { LAUNCH(h, k_schur_long, d.nlong, kLongThreads, d); LAUNCH(h, k_schur_long_pairs, 296, 256, d); }
  }
  { const int jrc = join_tail(h); if (jrc) return jrc; }
  if (h->cams_pending) {
    // after a fused linearisation (single GPU): ONE assemble kernel for U / gc / S / rhs, then the window cost and the
    // gradient test on the side stream while the linear solve runs (do_step joins)
    h->cams_pending = false;
    if (h->comm.active()) {
      // point-sharded: this rank's share of U | gc | S | rsum by ONE assemble kernel, its cost and gradient maximum,
      // then ONE all-reduce over the whole block instead of one per linearisation plus one per Schur step
      const int world = h->comm.world;
      if (!std::getenv("SBA_NO_COMPACT_ASM")) {
        CUDA_TRY(cudaMemsetAsync(d.U, 0, (size_t)((d.rsum + (size_t)d.Nf * 6) - d.U) * sizeof(double), h->stream));
        LAUNCH(h, k_assemble_schur, d.n_asm, kAsmThreads, d, 2 | 16, h->P, 0, 0);
      } else {
        LAUNCH(h, k_assemble_schur, d.nbu + d.Nf, kAsmThreads, d, 2, h->P, 0, 0);
      }
      LAUNCH(h, k_lin_finish_local, d.nwin, kThreads, d, h->comm.rank, world);
      std::string _err;
      if (!h->comm.allreduce(d.U, (size_t)((d.rsum + (size_t)d.Nf * 6) - d.U), false, h->stream, &_err)) { set_error(_err); return SBA_NCCL_ERROR; }
      LAUNCH(h, k_lin_finish_global, d.nwin, kThreads, d, h->P, h->pend_keep_cost ? 1 : 0, world);
      if (h->pend_lm_begin >= 0) LAUNCH(h, k_lm_begin, cdiv(d.nwin, 128), 128, d, h->P, h->pend_radius, h->pend_lm_begin, 0);
      LAUNCH(h, k_cam_prepare, cdiv(d.Nf, 64), 64, d);
      CUDA_TRY(cudaGetLastError());
      return SBA_OK;
    }
    stamp(h, h->stream, 0);
    if (h->every_window_free && h->pend_lm_begin <= 0 && !std::getenv("SBA_NO_TICKET")) {
      // the last camera CTA of every window finishes the window (cost, gradient maximum, gradient-tolerance test): no
      // side stream, two launches fewer
      LAUNCH(h, k_assemble_schur, d.nbu + d.Nf + d.nwin, kAsmThreads, d, 3 | 8 | (exact_solver ? 4 : 0), h->P, h->pend_keep_cost ? 1 : 0, h->pend_lm_begin == 0 ? 1 : 0);
      stamp(h, h->stream, 1);
      CUDA_TRY(cudaGetLastError());
      return SBA_OK;
    }
    LAUNCH(h, k_assemble_schur, d.nbu + d.Nf, kAsmThreads, d, 3 | (exact_solver ? 4 : 0), h->P, 0, 0);
    stamp(h, h->stream, 1);
    CUDA_TRY(cudaEventRecord(h->ev_fork, h->stream));
    CUDA_TRY(cudaStreamWaitEvent(h->side, h->ev_fork, 0));
    stamp(h, h->side, 2);
    if (d.nwin > 0) { k_lin_finish<<<d.nwin, kThreads, 0, h->side>>>(d, h->P, h->pend_keep_cost ? 1 : 0); h->launches++; }
    if (h->pend_lm_begin >= 0) { k_lm_begin<<<cdiv(d.nwin, 128), 128, 0, h->side>>>(d, h->P, h->pend_radius, h->pend_lm_begin, 0); h->launches++; }
    stamp(h, h->side, 3);
    h->tail_pending = true;
    CUDA_TRY(cudaGetLastError());
    return SBA_OK;
  }
  const int fused = h->comm.active() ? 0 : 1;     // multi-GPU: the partial sums are all-reduced before the cameras are prepared
  LAUNCH(h, k_assemble_schur, d.nbu + d.Nf, kAsmThreads, d, fused, h->P, 0, 0);
  if (fused) { CUDA_TRY(cudaGetLastError()); return SBA_OK; }
  if (h->comm.active()) {
    std::string _err;
    if (!h->comm.allreduce(d.Su, (size_t)d.nbu * 36 + (size_t)d.Nf * 6, false, h->stream, &_err)) { set_error(_err); return SBA_NCCL_ERROR; }   // Su | rsum
  }
  LAUNCH(h, k_cam_prepare, cdiv(d.Nf, 64), 64, d);
  CUDA_TRY(cudaGetLastError());
  return SBA_OK;
}

// K4: block-Jacobi PCG on the reduced camera system of every window (replicated on all ranks after
// the all-reduce of S).  The host reads one control counter every `kPcgCheck` iterations.
static constexpr int kPcgCheck = 4;
static int do_pcg(sba_handle h, const sba_options* opt, int32_t* iters_out) {
  Dev& d = h->d;
  const DevParams& P = h->P;
  int it = 0;
  const bool force_coop = std::getenv("SBA_FORCE_COOP_PCG") != nullptr && d.nwin == 1 && h->coop_pcg_blocks > 0;
  const bool force_sparse = std::getenv("SBA_FORCE_SPARSE_CHOL") != nullptr && h->spchol_ok;
  if (opt->linear_solver == 2 && !h->chol_ok && !h->spchol_ok) { set_error("linear_solver = Cholesky: no exact factorisation available for this structure"); return SBA_INVALID_ARGUMENT; }
  if (d.Nf > 0 && opt->linear_solver != 1 && h->spchol_ok && (force_sparse || !h->chol_ok)) {
    // more than 64 free cameras in a window: the exact block-sparse Cholesky, one cooperative launch
    CUDA_TRY(cudaMemsetAsync(d.counters + 4, 0, 4, h->stream));
    int grid = std::max(1, h->spchol_blocks);
    if (const char* e = std::getenv("SBA_SPCHOL_GRID")) grid = std::max(1, std::min(grid, std::atoi(e)));   // tuning aid
    Dev dd = d; SpCholDev ss = h->sp;
    ss.profile = (h->P.debug == 2) ? 1 : 0;
    void* args[] = {&dd, &ss};
    stamp(h, h->stream, 4);
    CUDA_TRY(cudaLaunchCooperativeKernel((const void*)k_spchol, dim3(grid), dim3(kSpThreads), args, h->spchol_smem, h->stream));
    h->launches++;
    stamp(h, h->stream, 5);
    it = -1;
  } else if (d.Nf > 0 && h->chol_ok && opt->linear_solver != 1) {
    CUDA_TRY(cudaMemsetAsync(d.counters + 4, 0, 4, h->stream));
    if (h->chol_band_ok && !std::getenv("SBA_FORCE_CLUSTER")) {
      k_chol_band<<<d.nwin, kBandThreads, h->chol_band_smem, h->stream>>>(d, h->band_last, (int)(h->P.debug == 2));
    } else if (h->chol_cluster_ok) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)d.nwin * kCholCluster); cfg.blockDim = dim3(kCholClThreads);
      cfg.dynamicSmemBytes = h->chol_cluster_smem; cfg.stream = h->stream;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = kCholCluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      CUDA_TRY(cudaLaunchKernelEx(&cfg, k_chol_cluster, d, (int)(h->P.debug == 2)));
    } else {
      if (h->chol_in_smem) k_chol_window<true><<<d.nwin, h->chol_threads, h->chol_smem, h->stream>>>(d);
      else k_chol_window<false><<<d.nwin, h->chol_threads, h->chol_smem, h->stream>>>(d);
    }
    h->launches++;
    it = -1;
  } else if (d.Nf > 0 && h->pcg_cluster_ok && !force_coop) {
    CUDA_TRY(cudaMemsetAsync(d.counters + 4, 0, 4, h->stream));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)d.nwin * kPcgCluster); cfg.blockDim = dim3(kPcgClThreads);
    cfg.dynamicSmemBytes = h->pcg_cluster_smem; cfg.stream = h->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = kPcgCluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, k_pcg_cluster, d, P, (int)opt->max_pcg_iterations, h->pcg_rows_cap, h->pcg_ent_cap));
    h->launches++;
    it = -1;
  } else if (d.Nf > 0 && h->pcg_window_ok && !force_coop) {
    // one launch, no host round trip; the iteration count comes back with the LM counters (counters[4])
    CUDA_TRY(cudaMemsetAsync(d.counters + 4, 0, 4, h->stream));
    k_pcg_window<<<d.nwin, h->pcg_threads, h->pcg_smem, h->stream>>>(d, P, opt->max_pcg_iterations, h->pcg_s_in_smem);
    h->launches++;
    it = -1;
  } else if (d.Nf > 0 && d.nwin == 1 && h->coop_pcg_blocks > 0 && !std::getenv("SBA_FORCE_MULTILAUNCH")) {
    // one big window: cooperative single-launch PCG over the block-CSR in global memory
    const int grid = std::max(1, std::min(h->coop_pcg_blocks, cdiv((int64_t)d.Nf * 32, kThreads)));
    Dev dd = d; DevParams pp = P; int mi = opt->max_pcg_iterations; double* pb = h->pcg_pbuf; double* rd = h->pcg_rowdot;
    void* args[] = {&dd, &pp, &mi, &pb, &rd};
    CUDA_TRY(cudaLaunchCooperativeKernel((const void*)k_pcg_coop, dim3(grid), dim3(kThreads), args, 0, h->stream));
    h->launches++;
    it = -1;
  } else if (d.Nf > 0) {
    const int gf = cdiv(d.Nf, 128), gs = cdiv((int64_t)d.Nf * 6, 128);
    CUDA_TRY(cudaMemsetAsync(d.counters + 1, 0, 4, h->stream));
    LAUNCH(h, k_pcg_init, gf, 128, d);
    LAUNCH(h, k_pcg_dot, d.nwin, kThreads, d, P, d.r, d.z, 0, opt->max_pcg_iterations);
    CUDA_TRY(cudaMemcpyAsync(h->h_counters + 1, d.counters + 1, 4, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    int active = h->h_counters[1];
    while (active > 0 && it < opt->max_pcg_iterations) {
      const int n = std::min(kPcgCheck, opt->max_pcg_iterations - it);
      for (int k = 0; k < n; ++k) {
        LAUNCH(h, k_pcg_matvec, gs, 128, d);
        LAUNCH(h, k_pcg_dot, d.nwin, kThreads, d, P, d.p, d.Ap, 1, opt->max_pcg_iterations);
        LAUNCH(h, k_pcg_update_xr, gf, 128, d);
        if (k == n - 1) CUDA_TRY(cudaMemsetAsync(d.counters + 1, 0, 4, h->stream));
        LAUNCH(h, k_pcg_dot, d.nwin, kThreads, d, P, d.r, d.z, 2, opt->max_pcg_iterations);
        LAUNCH(h, k_pcg_update_p, gs, 128, d);
      }
      it += n;
      CUDA_TRY(cudaMemcpyAsync(h->h_counters + 1, d.counters + 1, 4, cudaMemcpyDeviceToHost, h->stream));
      CUDA_TRY(cudaStreamSynchronize(h->stream));
      active = h->h_counters[1];
    }
  }
  CUDA_TRY(cudaGetLastError());
  if (iters_out) *iters_out = it;
  return SBA_OK;
}

// One round of the projected line search: trial poses / points at every searching window's current step
// size, cost (+ directional derivative), then the per-window Armijo decision.
static int do_ls_round(sba_handle h) {
  Dev& d = h->d;
  const DevParams& P = h->P;
  const int rank = (int)(h->comm.active() ? h->comm.rank : 0);
  CUDA_TRY(cudaMemsetAsync(d.counters + 3, 0, 4, h->stream));
  LAUNCH(h, k_trial_cams, cdiv(d.Nc, 128), 128, d);
  LAUNCH(h, k_trial_points, d.npchunk, kThreads, d, P);
  LAUNCH(h, k_ls_reduce, d.nwin, kThreads, d, rank);
  if (h->comm.active()) {
    std::string _err;
    if (!h->comm.allreduce(d.w_ls_red, (size_t)d.nwin * 3, false, h->stream, &_err)) { set_error(_err); return SBA_NCCL_ERROR; }
  }
  LAUNCH(h, k_ls_control, cdiv(d.nwin, 128), 128, d, P);
  return SBA_OK;
}

// K5 + K6 + trust-region control: step, line search (Ceres: bounds => projected Armijo search after every
// trust-region step), accept / reject per window.  Leaves counters [0] running [2] accepted on the host.
// `defer`: single-GPU cooperative path only — enqueue the step, the read-back of the counters and ev[3], and return
// without waiting; the caller synchronises on ev[3] before it looks at h_counters.
static int do_step(sba_handle h, int* ls_rounds, bool defer = false) {
  Dev& d = h->d;
  const DevParams& P = h->P;
  { const int jrc = join_tail(h); if (jrc) return jrc; }
  const int rank = (int)(h->comm.active() ? h->comm.rank : 0);
  if (!h->comm.active() && h->coop_blocks > 0) {
    // single GPU: the whole step in one cooperative launch (k_step_loop), one host read at the end
    const int want = std::max(std::max(d.npchunk, d.nwin), cdiv(d.Nc, kThreads));
    int grid = std::max(1, std::min(h->coop_blocks, want));
    if (const char* e = std::getenv("SBA_STEP_GRID")) grid = std::max(1, std::min(grid, std::atoi(e)));   // tuning aid
    Dev dd = d; DevParams pp = P;
    // (small problems: every chunk computes its own trial cameras and the grid barrier behind the trial-camera phase is
    // skipped; large ones read the table that phase writes)
    int fused_cams = (h->H.max_slots <= kTrialSlots && h->H.long_chunks.empty() && d.npchunk < 4 * grid) ? 1 : 0;
    // large problems: the back substitution as its own kernel (a CTA per chunk, W by bulk copy, more CTAs per SM than the
    // register-heavy cooperative kernel allows); small ones keep it inside the cooperative launch
    int skip_backsub = (d.npchunk >= 4 * grid && !std::getenv("SBA_BACKSUB_IN_LOOP")) ? 1 : 0;
    if (skip_backsub) launch_backsub(h);
    void* args[] = {&dd, &pp, &fused_cams, &skip_backsub};
    cudaEvent_t dbg[3] = {nullptr, nullptr, nullptr};
    const bool dbg_ev = P.debug == 2 && !defer;           // SBA_DEBUG_LS=2: where the step's wall time goes
    const auto th0 = std::chrono::steady_clock::now();
    if (dbg_ev) { for (auto& e : dbg) cudaEventCreate(&e); cudaEventRecord(dbg[0], h->stream); }
    stamp(h, h->stream, 6);
    CUDA_TRY(cudaLaunchCooperativeKernel((const void*)k_step_loop, dim3(grid), dim3(kThreads), args, 0, h->stream));
    h->launches++;
    stamp(h, h->stream, 7);
    if (dbg_ev) cudaEventRecord(dbg[1], h->stream);
    CUDA_TRY(cudaMemcpyAsync(h->h_counters, d.counters, 32, cudaMemcpyDeviceToHost, h->stream));
    if (dbg_ev) cudaEventRecord(dbg[2], h->stream);
    const auto th1 = std::chrono::steady_clock::now();
    if (defer) { CUDA_TRY(cudaEventRecord(h->ev[3], h->stream)); return SBA_OK; }
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (dbg_ev) {
      const auto th2 = std::chrono::steady_clock::now();
      float k = 0.f, c = 0.f;
      cudaEventElapsedTime(&k, dbg[0], dbg[1]); cudaEventElapsedTime(&c, dbg[1], dbg[2]);
      std::fprintf(stderr, "[step host] enqueue %.1f us, wait %.1f us | device: kernel %.1f us, d2h %.1f us\n",
                   std::chrono::duration<double, std::micro>(th1 - th0).count(), std::chrono::duration<double, std::micro>(th2 - th1).count(), k * 1e3, c * 1e3);
      for (auto& e : dbg) cudaEventDestroy(e);
      if (h->stamps) {
        unsigned long long t[8];
        cudaMemcpy(t, h->stamps, sizeof(t), cudaMemcpyDeviceToHost);
        auto us = [&](int i) { return (double)((long long)t[i] - (long long)t[0]) * 1e-3; };
        std::fprintf(stderr, "[stamps us] assemble 0..%.1f | side: tail %.1f..%.1f | solve %.1f..%.1f | step %.1f..%.1f\n",
                     us(1), us(2), us(3), us(4), us(5), us(6), us(7));
      }
    }
    if (ls_rounds) *ls_rounds += h->h_counters[6];
    return SBA_OK;
  }
  int rc;
  CUDA_TRY(cudaMemsetAsync(d.counters, 0, 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(d.counters + 2, 0, 4, h->stream));
  launch_backsub(h);
  LAUNCH(h, k_cam_step, d.nwin, kThreads, d, rank);
  if (h->comm.active()) {
    std::string _err;   // w_gd, w_dDd, w_x2, w_dr and the max-norm slots are one contiguous block: one all-reduce
    const int world = h->comm.world;
    LAUNCH(h, k_slots_put, cdiv((int64_t)d.nwin * world, 128), 128, d.dslots, d.w_dmax, d.nwin, h->comm.rank, world);
    if (!h->comm.allreduce(d.w_gd, (size_t)d.nwin * 4 + (size_t)d.nwin * world, false, h->stream, &_err)) { set_error(_err); return SBA_NCCL_ERROR; }
    LAUNCH(h, k_slots_max, cdiv(d.nwin, 128), 128, d.w_dmax, d.dslots, d.nwin, world);
  }
  int rounds = 0;
  for (;;) {
    if ((rc = do_ls_round(h))) return rc;
    ++rounds;
    LAUNCH(h, k_lm_control, cdiv(d.nwin, 128), 128, d, P);
    CUDA_TRY(cudaMemcpyAsync(h->h_counters, d.counters, 32, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    if (h->h_counters[3] <= 0 || rounds > 64) break;
  }
  if (ls_rounds) *ls_rounds += rounds;
  LAUNCH(h, k_apply_points, d.npchunk, kThreads, d);
  LAUNCH(h, k_apply_cams, cdiv((int64_t)d.Nc * 6, 256), 256, d);
  CUDA_TRY(cudaGetLastError());
  return SBA_OK;
}

int sba_linearize(sba_handle h, const sba_options* opt) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  sba_options o; if (opt) o = *opt; else sba_default_options(&o);
  CUDA_TRY(cudaSetDevice(h->device));
  fill_params(h, &o);
  if ((rc = do_linearize(h))) return rc;
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return SBA_OK;
}

static int set_radius(sba_handle h, double radius) {
  std::vector<double> r((size_t)h->d.nwin, radius);
  CUDA_TRY(cudaMemcpyAsync(h->d.w_radius, r.data(), r.size() * 8, cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaMemsetAsync(h->d.w_status, 0, (size_t)h->d.nwin * 4, h->stream));
  CUDA_TRY(cudaMemsetAsync(h->d.w_fail, 0, (size_t)h->d.nwin * 4, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));   // r is a host temporary
  h->last_radius = radius;
  return SBA_OK;
}

int sba_schur(sba_handle h, const sba_options* opt, double radius) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  if (!h->jacobi_done) { set_error("sba_schur needs a linearisation (sba_linearize)"); return SBA_INVALID_ARGUMENT; }
  if (!(radius > 0.0)) { set_error("radius must be positive"); return SBA_INVALID_ARGUMENT; }
  sba_options o; if (opt) o = *opt; else sba_default_options(&o);
  CUDA_TRY(cudaSetDevice(h->device));
  fill_params(h, &o);
  if ((rc = set_radius(h, radius))) return rc;
  if ((rc = do_schur(h))) return rc;
  if ((rc = join_tail(h))) return rc;
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return SBA_OK;
}

int sba_pcg(sba_handle h, const sba_options* opt, int32_t* iterations) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  sba_options o; if (opt) o = *opt; else sba_default_options(&o);
  CUDA_TRY(cudaSetDevice(h->device));
  fill_params(h, &o);
  if ((rc = do_pcg(h, &o, iterations))) return rc;
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  if (iterations && *iterations < 0) {     // single-launch path: the count lives on the device
    CUDA_TRY(cudaMemcpy(iterations, h->d.counters + 4, 4, cudaMemcpyDeviceToHost));
  }
  return SBA_OK;
}

int sba_time_linearize_schur(sba_handle h, const sba_options* opt, double radius, int32_t repeats,
                             double* ms_linearize, double* ms_schur) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  if (repeats < 1) repeats = 1;
  sba_options o; if (opt) o = *opt; else sba_default_options(&o);
  CUDA_TRY(cudaSetDevice(h->device));
  fill_params(h, &o);
  if ((rc = do_linearize(h))) return rc;           // makes sure the Jacobi scaling exists
  if ((rc = set_radius(h, radius))) return rc;
  float ml = 0.f, ms = 0.f;
  for (int i = 0; i < repeats; ++i) {
    CUDA_TRY(cudaEventRecord(h->ev[0], h->stream));
    if ((rc = do_linearize(h))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev[1], h->stream));
    if ((rc = do_schur(h))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev[2], h->stream));
    CUDA_TRY(cudaEventSynchronize(h->ev[2]));
    float a = 0.f, b = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&a, h->ev[0], h->ev[1]));
    CUDA_TRY(cudaEventElapsedTime(&b, h->ev[1], h->ev[2]));
    ml += a; ms += b;
  }
  if (ms_linearize) *ms_linearize = ml / repeats;
  if (ms_schur) *ms_schur = ms / repeats;
  return SBA_OK;
}

int sba_time_stage(sba_handle h, const sba_options* opt, double radius, int32_t stage, int32_t repeats,
                   int32_t flush_l2, double* ms_out) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  if (!h->jacobi_done) { set_error("sba_time_stage needs a linearisation (sba_linearize)"); return SBA_INVALID_ARGUMENT; }
  if (stage < 0 || stage > 11) { set_error("unknown stage"); return SBA_INVALID_ARGUMENT; }
  if (repeats < 1) repeats = 1;
  sba_options o; if (opt) o = *opt; else sba_default_options(&o);
  CUDA_TRY(cudaSetDevice(h->device));
  fill_params(h, &o);
  if (radius > 0.0 && (rc = set_radius(h, radius))) return rc;
  Dev& d = h->d;
  const DevParams& P = h->P;
  const size_t flush_bytes = (size_t)256 << 20;
  if (flush_l2 && !h->flush) { if ((rc = h->alloc(&h->flush, flush_bytes))) return rc; }
  double total = 0.0;
  for (int i = 0; i < repeats; ++i) {
    if (flush_l2) CUDA_TRY(cudaMemsetAsync(h->flush, i & 0xff, flush_bytes, h->stream));
    CUDA_TRY(cudaEventRecord(h->ev[6], h->stream));
    switch (stage) {
      case 0: if (i == 0) { LAUNCH(h, k_campre, cdiv(d.Nc, 128), 128, d); CUDA_TRY(cudaEventRecord(h->ev[6], h->stream)); }
              if (d.npchunk > 0) { k_lin_chunk<<<d.npchunk, kLinThreads, h->lin_smem, h->stream>>>(d, P, dbg_mask() >> 8); h->launches++; } break;
      case 1: LAUNCH(h, k_assemble_schur, d.nbu + d.Nf, kAsmThreads, d, 1, h->P, 0, 0); break;
      case 2: if (d.npchunk > 0) { k_schur_chunk<<<d.npchunk, kSchurThreads, h->schur_smem, h->stream>>>(d, h->H.max_chunk_pairs, dbg_mask()); h->launches++; } break;
      case 3: if ((rc = do_linearize(h))) return rc; break;
      case 4: if ((rc = do_schur(h))) return rc; break;
      case 8:
        // one bench step = linearisation + Schur step, launched as ONE CUDA graph on a single GPU (the five launches
        // and the fork / join of the side stream are captured once): the host cannot open gaps inside a timed step
        if (!h->comm.active() && !std::getenv("SBA_NO_GRAPH")) {
          if (!h->step_graph) {
            if ((rc = join_tail(h))) return rc;
            cudaGraph_t g = nullptr;
            CUDA_TRY(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
            const int64_t l0 = h->launches;
            int r1 = do_linearize(h, false, true, -1, 0.0, true);
            if (!r1) r1 = do_schur(h);
            if (!r1) r1 = join_tail(h);
            const cudaError_t ce = cudaStreamEndCapture(h->stream, &g);
            h->step_launches = (int)(h->launches - l0); h->launches = l0;
            if (r1) return r1;
            CUDA_TRY(ce);
            CUDA_TRY(cudaGraphInstantiate(&h->step_graph, g, 0));
            cudaGraphDestroy(g);
            CUDA_TRY(cudaEventRecord(h->ev[6], h->stream));      // the capture swallowed the start event of this repeat
          }
          CUDA_TRY(cudaGraphLaunch(h->step_graph, h->stream));
          h->launches += h->step_launches;
        } else {
          if ((rc = do_linearize(h, false, true, -1, 0.0, true))) return rc;
          if ((rc = do_schur(h))) return rc;
          if ((rc = join_tail(h))) return rc;
        }
        break;
      case 11:     // the fused K1-K3 chunk kernel alone (camera table from the last k_campre)
        if (!h->fused_ok) { set_error("fused chunk kernel unavailable for this structure"); return SBA_INVALID_ARGUMENT; }
        if (i == 0) { LAUNCH(h, k_campre, cdiv(d.Nc, 128), 128, d); CUDA_TRY(cudaEventRecord(h->ev[6], h->stream)); }
        if (d.npchunk > 0) {
          if (P.debug == 2 && i == 0) {
            cudaFuncSetAttribute(k_fused_chunk<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->fused_smem);
            k_fused_chunk<2, true><<<d.npchunk, kFusedThreads, h->fused_smem, h->stream>>>(d, P);
          }
          else { launch_fused(h); h->launches--; }
          h->launches++;
        }
        break;
      case 5: launch_backsub(h); break;
      case 9: LAUNCH(h, k_assemble_cams, d.Nf, kAsmThreads, d); break;
      case 10: LAUNCH(h, k_lin_finish, d.nwin, kThreads, d, P, 0); break;
      case 6:      // one line-search evaluation of the full step (needs sba_schur + sba_pcg for a meaningful step)
        launch_backsub(h);
        LAUNCH(h, k_cam_step, d.nwin, kThreads, d, 0);
        LAUNCH(h, k_trial_cams, cdiv(d.Nc, 128), 128, d);
        CUDA_TRY(cudaEventRecord(h->ev[6], h->stream));   // time the cost kernel alone
        LAUNCH(h, k_trial_points, d.npchunk, kThreads, d, P);
        break;
      case 7: {
        const int gf = cdiv(d.Nf, 128), gs = cdiv((int64_t)d.Nf * 6, 128);
        LAUNCH(h, k_pcg_matvec, gs, 128, d);
        LAUNCH(h, k_pcg_dot, d.nwin, kThreads, d, P, d.p, d.Ap, 1, 1 << 30);
        LAUNCH(h, k_pcg_update_xr, gf, 128, d);
        LAUNCH(h, k_pcg_dot, d.nwin, kThreads, d, P, d.r, d.z, 2, 1 << 30);
        LAUNCH(h, k_pcg_update_p, gs, 128, d);
        break;
      }
    }
    CUDA_TRY(cudaEventRecord(h->ev[7], h->stream));
    CUDA_TRY(cudaEventSynchronize(h->ev[7]));
    float a = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&a, h->ev[6], h->ev[7]));
    total += a;
  }
  CUDA_TRY(cudaGetLastError());
  if (ms_out) *ms_out = total / repeats;
  return SBA_OK;
}

int sba_run(sba_handle h, const sba_options* opt, sba_summary* sum) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  sba_options o; if (opt) o = *opt; else sba_default_options(&o);
  if (o.max_iterations < 0 || !(o.huber_delta > 0.0) || !(o.initial_radius > 0.0) ||
      !(o.point_lower_bound <= o.point_upper_bound)) { set_error("bad options"); return SBA_INVALID_ARGUMENT; }
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  auto secs = [&](clk::time_point a) { return std::chrono::duration<double>(clk::now() - a).count(); };
  CUDA_TRY(cudaSetDevice(h->device));
  fill_params(h, &o);
  Dev& d = h->d;
  const DevParams& P = h->P;
  sba_summary S; std::memset(&S, 0, sizeof(S));
  const int64_t launches0 = h->launches;

  // iteration 0 (Ceres): project onto the bounds, evaluate, fix the Jacobi scaling, gradient test
  LAUNCH(h, k_project_points, cdiv((int64_t)d.Np * 3, 256), 256, (int64_t)d.Np * 3, d.pts, P.lo, P.hi);
  h->jacobi_done = false;
  CUDA_TRY(cudaEventRecord(h->ev[4], h->stream));
  if ((rc = do_linearize(h, false, false, 1, o.initial_radius))) return rc;
  CUDA_TRY(cudaEventRecord(h->ev[5], h->stream));
  S.linearizations = 1;
  std::vector<double> wcost((size_t)d.nwin);
  CUDA_TRY(cudaMemcpyAsync(wcost.data(), d.w_cost, wcost.size() * 8, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  for (double c : wcost) S.initial_cost += c;
  bool lin_pending = true;
  cudaEvent_t* lin_ev = h->ev + 4;
  int lin_slot = 1;
  auto take_lin_time = [&]() {
    if (!lin_pending) return;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, lin_ev[0], lin_ev[1]) == cudaSuccess) S.ms_linearize += ms;
    lin_pending = false;
  };
  take_lin_time();
  // single GPU, cooperative step, fused K1-K3: the host runs one linearisation ahead of the step's outcome
  const bool run_ahead = !h->comm.active() && h->coop_blocks > 0 && h->fused_ok && !o.verbose && !std::getenv("SBA_NO_RUN_AHEAD");

  int running = d.nwin;
  int ls_rounds = 0;
  for (int iter = 0; running > 0 && iter < o.max_iterations; ++iter) {
    bool timed_out = o.max_time_s > 0.0 && secs(t0) >= o.max_time_s;
    if (o.max_time_s > 0.0 && h->comm.active()) {
      // every other LM decision is bit-identical on all ranks; the wall clock is not, and a rank that stops alone
      // would leave its peers inside an all-reduce: the ranks agree on the flag (max) before anybody acts on it
      double flag = timed_out ? 1.0 : 0.0;
      CUDA_TRY(cudaMemcpyAsync(d.w_dmax + d.nwin, &flag, 8, cudaMemcpyHostToDevice, h->stream));
      { std::string _err; if (!h->comm.allreduce(d.w_dmax + d.nwin, 1, true, h->stream, &_err)) { set_error(_err); return SBA_NCCL_ERROR; } }
      CUDA_TRY(cudaMemcpyAsync(&flag, d.w_dmax + d.nwin, 8, cudaMemcpyDeviceToHost, h->stream));
      CUDA_TRY(cudaStreamSynchronize(h->stream));
      timed_out = flag != 0.0;
    }
    if (timed_out) {
      if ((rc = join_tail(h))) return rc;
      LAUNCH(h, k_lm_timeout, cdiv(d.nwin, 128), 128, d);
      break;
    }
    CUDA_TRY(cudaEventRecord(h->ev[0], h->stream));
    if ((rc = do_schur(h, o.linear_solver != 1 && (h->chol_ok || h->spchol_ok)))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev[1], h->stream));
    int32_t pit = 0;
    if ((rc = do_pcg(h, &o, &pit))) return rc;
    CUDA_TRY(cudaEventRecord(h->ev[2], h->stream));
    if (run_ahead) {
      // The next linearisation does not depend on what the host learns from this step: a rejected step re-linearises
      // at the same point (same bits; the fused kernel rebuilds the Schur partials at the reduced radius), a finished
      // window ignores it.  So it goes out BEFORE the host waits for the step's counters, and the ~25 us of
      // wake-up + launch latency per iteration overlap with it instead of idling the GPU.
      if ((rc = do_step(h, &ls_rounds, true))) return rc;
      cudaEvent_t* le = h->ev + (lin_slot ? 8 : 4);      // the previous linearisation's events stay readable
      CUDA_TRY(cudaEventRecord(le[0], h->stream));
      if ((rc = do_linearize(h, true, true, 0, o.initial_radius, true))) return rc;
      CUDA_TRY(cudaEventRecord(le[1], h->stream));
      CUDA_TRY(cudaEventSynchronize(h->ev[3]));
      take_lin_time();                                   // the previous one finished before this step began
      lin_ev = le; lin_slot ^= 1; lin_pending = true;
      ls_rounds += h->h_counters[6];
    } else {
      if ((rc = do_step(h, &ls_rounds))) return rc;      // syncs; h_counters holds [running, -, accepted, -, pcg its]
      CUDA_TRY(cudaEventRecord(h->ev[3], h->stream));
      CUDA_TRY(cudaEventSynchronize(h->ev[3]));
      take_lin_time();
    }
    S.pcg_iterations += (pit >= 0) ? pit : h->h_counters[4];
    float a = 0.f, b = 0.f, c = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&a, h->ev[0], h->ev[1]));
    CUDA_TRY(cudaEventElapsedTime(&b, h->ev[1], h->ev[2]));
    CUDA_TRY(cudaEventElapsedTime(&c, h->ev[2], h->ev[3]));
    S.ms_schur += a; S.ms_pcg += b; S.ms_update += c;
    running = h->h_counters[0];
    const int accepted = h->h_counters[2];
    if (run_ahead) {
      if (accepted > 0) S.linearizations++;
    } else if (accepted > 0) {
      CUDA_TRY(cudaEventRecord(h->ev[4], h->stream));
      if ((rc = do_linearize(h, true, true, 0, o.initial_radius, true))) return rc;
      CUDA_TRY(cudaEventRecord(h->ev[5], h->stream));
      lin_pending = true;
      S.linearizations++;
    } else {
      CUDA_TRY(cudaMemsetAsync(d.w_fail, 0, (size_t)d.nwin * 4, h->stream));
    }
    if (o.verbose) {
      if ((rc = join_tail(h))) return rc;
      CUDA_TRY(cudaMemcpyAsync(wcost.data(), d.w_cost, 8, cudaMemcpyDeviceToHost, h->stream));
      CUDA_TRY(cudaStreamSynchronize(h->stream));
      double dbg[6] = {0, 0, 0, 0, 0, 0};
      cudaMemcpy(&dbg[0], d.w_radius, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&dbg[1], d.w_step2, 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(&dbg[2], d.w_x2, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&dbg[3], d.w_ls_x, 8, cudaMemcpyDeviceToHost);
      cudaMemcpy(&dbg[4], d.w_cand, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&dbg[5], d.w_gd, 8, cudaMemcpyDeviceToHost);
      std::printf("[sba] it %d  cost(w0) %.12e  running %d  accepted %d  pcg %d  ls_rounds %d | radius %.3e |step| %.3e |x| %.3e ls_x %.3e cand %.12e gd %.3e\n",
                  iter + 1, wcost[0], running, accepted, (pit >= 0) ? pit : h->h_counters[4], ls_rounds, dbg[0], std::sqrt(dbg[1]), std::sqrt(dbg[2]), dbg[3], dbg[4], dbg[5]);
    }
  }
  // windows still running here hit the iteration limit
  if ((rc = join_tail(h))) return rc;
  if (h->cams_pending) {      // the last linearisation's camera-side assembly was left to a Schur step that never came
    h->cams_pending = false;
    LAUNCH(h, k_assemble_cams, d.Nf, kAsmThreads, d);
    if (h->comm.active()) {
      const int world = h->comm.world;
      LAUNCH(h, k_lin_finish_local, d.nwin, kThreads, d, h->comm.rank, world);
      std::string _err;
      if (!h->comm.allreduce(d.U, (size_t)d.Nc * 27 + (size_t)d.nwin + (size_t)d.nwin * world, false, h->stream, &_err)) { set_error(_err); return SBA_NCCL_ERROR; }
      LAUNCH(h, k_lin_finish_global, d.nwin, kThreads, d, P, h->pend_keep_cost ? 1 : 0, world);
    } else {
      LAUNCH(h, k_lin_finish, d.nwin, kThreads, d, P, h->pend_keep_cost ? 1 : 0);
    }
    if (h->pend_lm_begin >= 0) LAUNCH(h, k_lm_begin, cdiv(d.nwin, 128), 128, d, P, h->pend_radius, h->pend_lm_begin, 0);
  }
  std::vector<int32_t> st((size_t)d.nwin), its((size_t)d.nwin), succ((size_t)d.nwin);
  CUDA_TRY(cudaMemcpyAsync(st.data(), d.w_status, st.size() * 4, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaMemcpyAsync(its.data(), d.w_iter, its.size() * 4, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaMemcpyAsync(succ.data(), d.w_succ, succ.size() * 4, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaMemcpyAsync(wcost.data(), d.w_cost, wcost.size() * 8, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  take_lin_time();
  bool any_fail = false, any_limit = false;
  for (int w = 0; w < d.nwin; ++w) {
    int s = st[w];
    if (s == 0) s = SBA_TERM_MAX_ITERATIONS;
    if (w == 0) S.termination = s;
    S.final_cost += wcost[w];
    S.iterations = std::max(S.iterations, its[w]);
    S.successful_steps += succ[w];
    if (s == SBA_TERM_INVALID_STEPS || s < 0) any_fail = true;
    else if (s == SBA_TERM_MAX_ITERATIONS || s == SBA_TERM_MAX_TIME) any_limit = true;
    else S.num_windows_converged++;
  }
  S.time_structure_s = h->time_structure_s;
  S.time_solve_s = secs(t0);
  S.time_total_s = S.time_solve_s;
  (void)launches0;
  if (sum) *sum = S;
  if (any_fail) { set_error("numerical failure (invalid steps / non-finite cost) in at least one window"); return SBA_NUMERICAL_FAILURE; }
  return any_limit ? SBA_NO_CONVERGENCE : SBA_OK;
}

int sba_solve(sba_handle h, int32_t Nc, int32_t Np, int64_t Nobs, double* poses, double* points,
              const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv, const double* p_l,
              const double* p_r, const uint8_t* cam_fixed, int32_t nwin, const int32_t* win_cam_off,
              const int32_t* win_pt_off, const sba_options* opt, sba_summary* sum) {
  if (!h) { set_error("null handle"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) { set_error("host-only handle: no CUDA device bound (there is no CPU fallback)"); return SBA_CUDA_ERROR; }
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
  int rc = sba_set_structure(h, Nc, Np, Nobs, obs_cam, obs_pt, cam_fixed, nwin, win_cam_off, win_pt_off);
  if (rc) return rc;
  const auto t1 = clk::now();
  if ((rc = sba_set_state(h, poses, points))) return rc;
  if ((rc = sba_set_measurements(h, obs_uv, p_l, p_r))) return rc;
  const auto t2 = clk::now();
  sba_summary S;
  const int run_rc = sba_run(h, opt, &S);
  // The reference never looks at Ceres' termination type and always writes back the last accepted state, also after
  // FAILURE (src/bundle_adjuster.cpp:116-132): a numerical failure (5 invalid steps in a row, non-finite cost) still
  // returns the last accepted state (candidates are only applied on acceptance) together with its status.
  if (run_rc != SBA_OK && run_rc != SBA_NO_CONVERGENCE && run_rc != SBA_NUMERICAL_FAILURE) return run_rc;   // inputs untouched
  const auto t3 = clk::now();
  if ((rc = sba_get_state(h, poses, points))) return rc;
  const auto t4 = clk::now();
  S.time_structure_s = secs(t0, t1); S.time_h2d_s = secs(t1, t2); S.time_d2h_s = secs(t3, t4);
  S.time_total_s = secs(t0, t4);
  if (sum) *sum = S;
  return run_rc;
}

// ---- introspection --------------------------------------------------------------------------------
extern "C++" {
template <typename T>
static int d2h(sba_handle h, std::vector<T>& dst, const T* src, size_t n) {
  dst.resize(n);
  if (n) CUDA_TRY(cudaMemcpyAsync(dst.data(), src, n * sizeof(T), cudaMemcpyDeviceToHost, h->stream));
  return SBA_OK;
}
}

int sba_get_linearization(sba_handle h, double* r, double* U, double* V, double* W, double* gc,
                          double* gp, double* cost) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  CUDA_TRY(cudaSetDevice(h->device));
  Dev& d = h->d;
  const HostStructure& H = h->H;
  if (r) {   // corrected residuals are not kept by the hot path: re-run the linearisation with the debug sink
    if (!h->r_dbg) { if ((rc = h->alloc(&h->r_dbg, (size_t)h->nobs_all * 4))) return rc; }
    CUDA_TRY(cudaMemsetAsync(h->r_dbg, 0, std::max<size_t>(1, (size_t)h->nobs_all * 4) * 8, h->stream));
    d.r_dbg = h->r_dbg;
    rc = do_linearize(h);
    d.r_dbg = nullptr;
    if (rc) return rc;
    if (h->nobs_all) CUDA_TRY(cudaMemcpyAsync(r, h->r_dbg, (size_t)h->nobs_all * 32, cudaMemcpyDeviceToHost, h->stream));
  }
  std::vector<double> hU, hV, hW;
  if (U) { if ((rc = d2h(h, hU, d.U, (size_t)d.Nc * 21))) return rc; }
  if (V) { if ((rc = d2h(h, hV, d.V, (size_t)d.Np * 6))) return rc; }
  if (W) { if ((rc = d2h(h, hW, d.W, (size_t)d.Nobs * 18))) return rc; }
  if (gc && d.Nc) CUDA_TRY(cudaMemcpyAsync(gc, d.gc, (size_t)d.Nc * 48, cudaMemcpyDeviceToHost, h->stream));
  if (gp && d.Np) CUDA_TRY(cudaMemcpyAsync(gp, d.gp, (size_t)d.Np * 24, cudaMemcpyDeviceToHost, h->stream));
  if (cost) CUDA_TRY(cudaMemcpyAsync(cost, d.w_cost, (size_t)d.nwin * 8, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  if (U)
    for (int c = 0; c < d.Nc; ++c)
      for (int i = 0; i < 6; ++i)
        for (int j = 0; j < 6; ++j) U[(size_t)c * 36 + 6 * i + j] = hU[(size_t)c * 21 + (i <= j ? up6(i, j) : up6(j, i))];
  if (V)
    for (int p = 0; p < d.Np; ++p) {
      const double* s = &hV[(size_t)p * 6];
      double* o = V + (size_t)p * 9;
      o[0] = s[0]; o[1] = s[1]; o[2] = s[2]; o[3] = s[1]; o[4] = s[3]; o[5] = s[4]; o[6] = s[2]; o[7] = s[4]; o[8] = s[5];
    }
  if (W) {
    std::memset(W, 0, (size_t)h->nobs_all * 18 * 8);   // observations of unowned points (multi-GPU) stay zero
    for (int64_t e = 0; e < d.Nobs; ++e) std::memcpy(W + (size_t)H.perm_p[e] * 18, &hW[(size_t)e * 18], 18 * 8);
  }
  return SBA_OK;
}

int64_t sba_get_schur_pattern(sba_handle h, int32_t* rows, int32_t* cols, int64_t capacity) {
  if (!h || !h->has_structure) return -1;
  const HostStructure& H = h->H;
  const int64_t n = H.row_ptr.empty() ? 0 : H.row_ptr[H.Nf];
  if (rows && cols) {
    // row entries are stored lower part first then upper part, each ascending in column: already sorted
    int64_t k = 0;
    for (int f = 0; f < H.Nf; ++f)
      for (int j = H.row_ptr[f]; j < H.row_ptr[f + 1] && k < capacity; ++j, ++k) { rows[k] = f; cols[k] = H.bsr_col[j]; }
  }
  return n;
}

int sba_get_schur_values(sba_handle h, double* blocks, double* rhs, double* solution) {
  int rc = check_ready(h, true);
  if (rc) return rc;
  CUDA_TRY(cudaSetDevice(h->device));
  const Dev& d = h->d;
  const HostStructure& H = h->H;
  std::vector<double> Su;
  if (blocks) { if ((rc = d2h(h, Su, d.Su, (size_t)d.nbu * 36))) return rc; }
  if (rhs && d.Nf) CUDA_TRY(cudaMemcpyAsync(rhs, d.rhs, (size_t)d.Nf * 48, cudaMemcpyDeviceToHost, h->stream));
  if (solution && d.Nf) CUDA_TRY(cudaMemcpyAsync(solution, d.x, (size_t)d.Nf * 48, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  if (blocks) {
    const int64_t n = H.row_ptr[H.Nf];
    for (int64_t k = 0; k < n; ++k) {
      const int32_t bi = H.bsr_blk[k];
      double* o = blocks + (size_t)k * 36;
      if (bi >= 0) std::memcpy(o, &Su[(size_t)bi * 36], 36 * 8);
      else {
        const double* s = &Su[(size_t)(~bi) * 36];
        for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) o[6 * i + j] = s[6 * j + i];
      }
    }
  }
  return SBA_OK;
}

int sba_get_sorted_views(sba_handle h, int32_t* by_point, int32_t* by_camera) {
  if (!h || !h->has_structure) { set_error("no structure"); return SBA_INVALID_ARGUMENT; }
  const HostStructure& H = h->H;
  if (by_point) std::memcpy(by_point, H.perm_p.data(), (size_t)H.Nobs * 4);
  if (by_camera) {   // camera order: grouped by camera, ascending point-order position inside a camera
    std::vector<int64_t> off((size_t)H.Nc + 1, 0);
    for (int64_t e = 0; e < H.Nobs; ++e) off[H.pcam[e] + 1]++;
    for (int c = 0; c < H.Nc; ++c) off[c + 1] += off[c];
    for (int64_t e = 0; e < H.Nobs; ++e) by_camera[off[H.pcam[e]]++] = H.perm_p[e];
  }
  return SBA_OK;
}

int sba_get_factor_symbolic(sba_handle h, int64_t info[5], int32_t* perm, int32_t* col_ptr, int32_t* row_idx,
                            int32_t* init_src, int64_t* upd_ptr, int32_t* upd_a, int32_t* upd_b, int32_t* lvl_ptr,
                            int32_t* lvl_cols, int32_t* lvl_len) {
  if (!h || !h->has_structure) { set_error("no structure"); return SBA_INVALID_ARGUMENT; }
  if (!h->sym_ok) { set_error("this structure has no sparse factorisation (every window has <= 64 free cameras, or the pattern is too dense)"); return SBA_INVALID_ARGUMENT; }
  const SparseCholSymbolic& Y = h->sym;
  if (info) { info[0] = Y.n; info[1] = Y.nnzL(); info[2] = Y.nupd(); info[3] = Y.nlev; info[4] = Y.nchain(); }
  auto cp = [](auto* dst, const auto& v) { if (dst && !v.empty()) std::memcpy(dst, v.data(), v.size() * sizeof(v[0])); };
  cp(perm, Y.perm); cp(col_ptr, Y.col_ptr); cp(row_idx, Y.row_idx); cp(init_src, Y.init_src); cp(upd_ptr, Y.upd_ptr);
  cp(upd_a, Y.upd_a); cp(upd_b, Y.upd_b); cp(lvl_ptr, Y.lvl_ptr); cp(lvl_cols, Y.lvl_cols); cp(lvl_len, Y.lvl_len);
  return SBA_OK;
}

int sba_measure_fp64_peak(sba_handle h, double* tflops) {
  if (!h || h->device < 0 || !tflops) { set_error("sba_measure_fp64_peak needs a device handle"); return SBA_INVALID_ARGUMENT; }
  CUDA_TRY(cudaSetDevice(h->device));
  int sms = 0;
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
  const int grid = sms * 8, iters = 1 << 14;
  double* sink = nullptr;
  CUDA_TRY(cudaMallocAsync((void**)&sink, (size_t)grid * 256 * 8, h->stream));
  double best = 0.0;
  for (int rep = 0; rep < 6; ++rep) {
    CUDA_TRY(cudaEventRecord(h->ev[6], h->stream));
    k_fp64_peak<<<grid, 256, 0, h->stream>>>(sink, iters, 1.0000001, 1e-9);
    CUDA_TRY(cudaEventRecord(h->ev[7], h->stream));
    CUDA_TRY(cudaEventSynchronize(h->ev[7]));
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, h->ev[6], h->ev[7]));
    if (rep > 0 && ms > 0.f) best = std::max(best, 2.0 * 8.0 * iters * (double)grid * 256.0 / (ms * 1e-3) / 1e12);
  }
  CUDA_TRY(cudaFreeAsync(sink, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  *tflops = best;
  return SBA_OK;
}

int32_t sba_num_free_cameras(sba_handle h) { return (h && h->has_structure) ? h->H.Nf : -1; }
int64_t sba_launch_count(sba_handle h) { return h ? h->launches : -1; }
int32_t sba_fused_kernel_variant(sba_handle h) {
  if (!h || !h->has_structure || !h->fused_ok) return 0;
  return h->persist_ok ? 3 : (h->fused_nt == 128 ? 2 : 1);
}

// ---- device-resident map -------------------------------------------------------------
extern "C++" {
template <typename T>
static int grow_device(T** p, size_t* cap, size_t used, size_t need, cudaStream_t stream) {
  if (need <= *cap) return SBA_OK;
  size_t ncap = std::max<size_t>(need, std::max<size_t>(*cap * 2, 1024));
  T* q = nullptr;
  CUDA_TRY(cudaMalloc((void**)&q, ncap * sizeof(T)));
  if (*p && used) CUDA_TRY(cudaMemcpyAsync(q, *p, used * sizeof(T), cudaMemcpyDeviceToDevice, stream));
  if (*p) { CUDA_TRY(cudaStreamSynchronize(stream)); CUDA_TRY(cudaFree(*p)); }
  *p = q; *cap = ncap;
  return SBA_OK;
}
}

int sba_map_reset(sba_handle h) {
  if (!h) { set_error("null handle"); return SBA_INVALID_ARGUMENT; }
  DeviceMap& M = h->map;
  M.T_wc.clear(); M.frame_off.assign(1, 0); M.obs_point_id.clear(); M.n_points = 0; M.n_obs = 0;
  return SBA_OK;
}

int sba_map_append_frame(sba_handle h, const float* T_wc16, int64_t nobs, const uint32_t* obs_point_id, const float* obs_uv,
                         int32_t num_new_points, const float* new_points) {
  if (!h || !T_wc16 || nobs < 0 || num_new_points < 0 || (nobs > 0 && (!obs_point_id || !obs_uv)) || (num_new_points > 0 && !new_points)) {
    set_error("sba_map_append_frame: bad arguments"); return SBA_INVALID_ARGUMENT;
  }
  if (h->device < 0) { set_error("host-only handle: no CUDA device bound (there is no CPU fallback)"); return SBA_CUDA_ERROR; }
  DeviceMap& M = h->map;
  const size_t np_after = M.n_points + (size_t)num_new_points;
  for (int64_t k = 0; k < nobs; ++k)
    if (obs_point_id[k] >= np_after) { set_error("sba_map_append_frame: observation of an unknown map point"); return SBA_INVALID_ARGUMENT; }
  CUDA_TRY(cudaSetDevice(h->device));
  int rc;
  if ((rc = grow_device(&M.d_points, &M.cap_points, M.n_points * 3, np_after * 3, h->stream))) return rc;
  if ((rc = grow_device(&M.d_uv, &M.cap_obs, M.n_obs, M.n_obs + (size_t)nobs, h->stream))) return rc;
  if (num_new_points > 0)
    CUDA_TRY(cudaMemcpyAsync(M.d_points + M.n_points * 3, new_points, (size_t)num_new_points * 12, cudaMemcpyHostToDevice, h->stream));
  if (nobs > 0) CUDA_TRY(cudaMemcpyAsync(M.d_uv + M.n_obs, obs_uv, (size_t)nobs * 16, cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));        // the caller's buffers may be pageable and short-lived
  M.T_wc.insert(M.T_wc.end(), T_wc16, T_wc16 + 16);
  M.obs_point_id.insert(M.obs_point_id.end(), obs_point_id, obs_point_id + nobs);
  M.frame_off.push_back(M.frame_off.back() + nobs);
  M.n_points = np_after; M.n_obs += (size_t)nobs;
  return SBA_OK;
}

int64_t sba_map_num_frames(sba_handle h) { return h ? (int64_t)h->map.frame_off.size() - 1 : -1; }
int64_t sba_map_num_points(sba_handle h) { return h ? (int64_t)h->map.n_points : -1; }

int sba_map_optimize(sba_handle h, int32_t start, int32_t end, const double* p_l, const double* p_r, const sba_options* opt,
                     sba_summary* sum) {
  if (!h || !p_l || !p_r) { set_error("null argument"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) { set_error("host-only handle: no CUDA device bound (there is no CPU fallback)"); return SBA_CUDA_ERROR; }
  if (h->comm.active()) { set_error("sba_map_optimize is a single-GPU path"); return SBA_INVALID_ARGUMENT; }
  DeviceMap& M = h->map;
  const int32_t N = (int32_t)M.frame_off.size() - 1;
  if (start < 0 || end < start || end > N) { set_error("sba_map_optimize: bad window"); return SBA_INVALID_ARGUMENT; }
  if (sum) std::memset(sum, 0, sizeof(*sum));
  if (end == start) return SBA_OK;
  using clk = std::chrono::steady_clock;
  auto ms_since = [](clk::time_point a) { return 1e3 * std::chrono::duration<double>(clk::now() - a).count(); };
  const auto t0 = clk::now();
  // ---- pack, integer side + poses (reference: src/bundle_adjuster.cpp:57-102; same scan order, same numbering)
  const int32_t nc = end - start;
  const int64_t k0 = M.frame_off[start], nobs = M.frame_off[end] - k0;
  M.poses.resize((size_t)nc * 6); M.cam.resize((size_t)nobs); M.pt.resize((size_t)nobs); M.ids.clear();
  {
    std::unordered_map<uint32_t, int32_t> idx;
    idx.reserve((size_t)nobs);
    int64_t ko = 0;
    for (int32_t f = start; f < end; ++f) {
      float T_cw[16];
      if (sba_invert_pose(M.T_wc.data() + (size_t)f * 16, T_cw) != SBA_OK) { set_error("singular frame pose"); return SBA_INVALID_ARGUMENT; }
      sba_matrix_to_pose(T_cw, M.poses.data() + (size_t)(f - start) * 6);
      for (int64_t k = M.frame_off[f]; k < M.frame_off[f + 1]; ++k, ++ko) {
        const uint32_t gid = M.obs_point_id[k];
        auto it = idx.find(gid);
        int32_t pi;
        if (it == idx.end()) { pi = (int32_t)M.ids.size(); M.ids.push_back(gid); idx.emplace(gid, pi); }
        else pi = it->second;
        M.cam[ko] = f - start; M.pt[ko] = pi;
      }
    }
  }
  const int32_t np = (int32_t)M.ids.size();
  M.ms_pack = ms_since(t0);
  const auto t1 = clk::now();
  int rc = sba_set_structure(h, nc, np, nobs, M.cam.data(), M.pt.data(), nullptr, 1, nullptr, nullptr);
  if (rc) return rc;
  M.ms_structure = ms_since(t1);
  const auto t2 = clk::now();
  // ---- state and measurements never leave the device: points gathered by id, pixels are a slice of the map's array
  Dev& d = h->d;
  CUDA_TRY(cudaSetDevice(h->device));
  if ((rc = grow_device(&M.d_ids, &M.cap_ids, 0, (size_t)np, h->stream))) return rc;
  CUDA_TRY(cudaMemcpyAsync(d.poses, M.poses.data(), (size_t)nc * 48, cudaMemcpyHostToDevice, h->stream));
  if (np > 0) {
    CUDA_TRY(cudaMemcpyAsync(M.d_ids, M.ids.data(), (size_t)np * 4, cudaMemcpyHostToDevice, h->stream));
    LAUNCH(h, k_map_gather_points, cdiv((int64_t)np * 3, 256), 256, np, M.d_ids, M.d_points, d.pts);
  }
  LAUNCH(h, k_permute_uv, cdiv(d.Nobs, 256), 256, d, M.d_uv + k0, const_cast<float4*>(d.uv_p));
  CUDA_TRY(cudaGetLastError());
  std::memcpy(h->P.pl, p_l, 96); std::memcpy(h->P.pr, p_r, 96);
  h->has_state = h->has_meas = true; h->jacobi_done = false;
  M.ms_upload = ms_since(t2);
  const auto t3 = clk::now();
  sba_summary S;
  const int run_rc = sba_run(h, opt, &S);
  if (run_rc != SBA_OK && run_rc != SBA_NO_CONVERGENCE && run_rc != SBA_NUMERICAL_FAILURE) return run_rc;
  M.ms_solve = ms_since(t3);
  const auto t4 = clk::now();
  // ---- unpack (reference :121-132): points back into the map on the device, poses through the host's f32 path
  if (np > 0) LAUNCH(h, k_map_scatter_points, cdiv((int64_t)np * 3, 256), 256, np, M.d_ids, d.pts, M.d_points);
  CUDA_TRY(cudaMemcpyAsync(M.poses.data(), d.poses, (size_t)nc * 48, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  if ((rc = sba_unpack_window(M.T_wc.data(), nullptr, start, end, M.poses.data(), nullptr, nullptr, 0))) return rc;
  M.ms_unpack = ms_since(t4);
  M.last_start = start; M.last_end = end;
  S.time_structure_s = 1e-3 * (M.ms_pack + M.ms_structure); S.time_h2d_s = 1e-3 * M.ms_upload; S.time_d2h_s = 1e-3 * M.ms_unpack;
  S.time_total_s = 1e-3 * ms_since(t0);
  if (sum) *sum = S;
  return run_rc;
}

int sba_map_get_poses(sba_handle h, int32_t first, int32_t count, float* T_wc16) {
  if (!h || !T_wc16 || first < 0 || count < 0 || (size_t)(first + count) * 16 > h->map.T_wc.size()) { set_error("bad range"); return SBA_INVALID_ARGUMENT; }
  std::memcpy(T_wc16, h->map.T_wc.data() + (size_t)first * 16, (size_t)count * 64);
  return SBA_OK;
}
int sba_map_set_pose(sba_handle h, int32_t frame, const float* T_wc16) {
  if (!h || !T_wc16 || frame < 0 || (size_t)(frame + 1) * 16 > h->map.T_wc.size()) { set_error("bad frame"); return SBA_INVALID_ARGUMENT; }
  std::memcpy(h->map.T_wc.data() + (size_t)frame * 16, T_wc16, 64);
  return SBA_OK;
}
int sba_map_get_points(sba_handle h, int64_t first, int64_t count, float* xyz) {
  if (!h || !xyz || first < 0 || count < 0 || (size_t)(first + count) > h->map.n_points) { set_error("bad range"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) return SBA_CUDA_ERROR;
  CUDA_TRY(cudaSetDevice(h->device));
  if (count > 0) CUDA_TRY(cudaMemcpyAsync(xyz, h->map.d_points + (size_t)first * 3, (size_t)count * 12, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return SBA_OK;
}
int sba_map_set_points(sba_handle h, int64_t first, int64_t count, const float* xyz) {
  if (!h || !xyz || first < 0 || count < 0 || (size_t)(first + count) > h->map.n_points) { set_error("bad range"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) return SBA_CUDA_ERROR;
  CUDA_TRY(cudaSetDevice(h->device));
  if (count > 0) CUDA_TRY(cudaMemcpyAsync(h->map.d_points + (size_t)first * 3, xyz, (size_t)count * 12, cudaMemcpyHostToDevice, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  return SBA_OK;
}
/* the last window's point ids (what Optimize wrote back) and the wall milliseconds of its phases:
 * ms[5] = {pack (index build + poses), structure, state on device, solve, unpack} */
int64_t sba_map_last_window(sba_handle h, uint32_t* point_ids, int64_t capacity, double* ms) {
  if (!h) return -1;
  const DeviceMap& M = h->map;
  if (point_ids) std::memcpy(point_ids, M.ids.data(), (size_t)std::min<int64_t>(capacity, (int64_t)M.ids.size()) * 4);
  if (ms) { ms[0] = M.ms_pack; ms[1] = M.ms_structure; ms[2] = M.ms_upload; ms[3] = M.ms_solve; ms[4] = M.ms_unpack; }
  return (int64_t)M.ids.size();
}

// ---- pose-graph optimisation ----------------------------------------------------------
static int pgo_run(sba_handle h, int32_t n, double* T_wc, int32_t m, const int32_t* edge_i, const int32_t* edge_j, const double* Z,
                   const uint8_t* fixed, const double* information6, double huber_delta, int32_t iterations, sba_pgo_summary* out) {
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  int rc;
  // ---- structure: free vertices, upper block pattern, partial lists (fixed summation order)
  std::vector<int32_t> fv((size_t)n);
  int32_t nf = 0;
  for (int v = 0; v < n; ++v) fv[v] = ((fixed ? fixed[v] != 0 : v == 0)) ? -1 : nf++;
  if (nf == 0 || m == 0) return SBA_OK;
  std::vector<int64_t> keys;
  keys.reserve((size_t)nf + m);
  for (int f = 0; f < nf; ++f) keys.push_back((int64_t)f * nf + f);
  for (int k = 0; k < m; ++k) {
    const int32_t i = edge_i[k], j = edge_j[k];
    if (i < 0 || j < 0 || i >= n || j >= n || i == j) { set_error("sba_pgo_optimize: bad edge"); return SBA_INVALID_ARGUMENT; }
    const int64_t a = fv[i], b = fv[j];
    if (a >= 0 && b >= 0) keys.push_back(std::min(a, b) * nf + std::max(a, b));
  }
  std::sort(keys.begin(), keys.end());
  keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
  const int32_t nbu = (int32_t)keys.size();
  std::vector<int32_t> blk_row(nbu), blk_col(nbu);
  for (int b = 0; b < nbu; ++b) { blk_row[b] = (int32_t)(keys[b] / nf); blk_col[b] = (int32_t)(keys[b] % nf); }
  auto blk_of = [&](int64_t r, int64_t c) { return (int32_t)(std::lower_bound(keys.begin(), keys.end(), r * nf + c) - keys.begin()); };
  std::vector<std::vector<int32_t>> bl((size_t)nbu), vl((size_t)nf);
  for (int k = 0; k < m; ++k) {
    const int64_t a = fv[edge_i[k]], b = fv[edge_j[k]];
    if (a >= 0) { bl[blk_of(a, a)].push_back(k * 4 + 0); vl[a].push_back(k * 2 + 0); }
    if (b >= 0) { bl[blk_of(b, b)].push_back(k * 4 + 1); vl[b].push_back(k * 2 + 1); }
    if (a >= 0 && b >= 0) bl[blk_of(std::min(a, b), std::max(a, b))].push_back(k * 4 + (a < b ? 2 : 3));
  }
  std::vector<int32_t> bpo((size_t)nbu + 1, 0), bp, vpo((size_t)nf + 1, 0), vp;
  for (int b = 0; b < nbu; ++b) { bpo[b + 1] = bpo[b] + (int32_t)bl[b].size(); bp.insert(bp.end(), bl[b].begin(), bl[b].end()); }
  for (int f = 0; f < nf; ++f) { vpo[f + 1] = vpo[f] + (int32_t)vl[f].size(); vp.insert(vp.end(), vl[f].begin(), vl[f].end()); }
  SparseCholSymbolic Y;
  const auto ts = clk::now();
  const std::string serr = sparse_chol_symbolic(nf, blk_row, blk_col, (int64_t)256 << 20, &Y);
  if (!serr.empty()) { set_error("sba_pgo_optimize: " + serr); return SBA_INVALID_ARGUMENT; }
  out->time_symbolic_s = std::chrono::duration<double>(clk::now() - ts).count();
  // ---- device state
  std::vector<double> X((size_t)n * 12), Zc((size_t)m * 12);
  auto to12 = [](const double* T, double* o) {
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) o[3 * r + c] = T[4 * r + c]; o[9 + r] = T[4 * r + 3]; }
  };
  for (int v = 0; v < n; ++v) to12(T_wc + (size_t)v * 16, &X[(size_t)v * 12]);
  for (int k = 0; k < m; ++k) to12(Z + (size_t)k * 16, &Zc[(size_t)k * 12]);
  std::vector<int32_t> ei(edge_i, edge_i + m), ej(edge_j, edge_j + m);
  PgoDev g{};
  g.n = n; g.nf = nf; g.m = m; g.nbu = nbu; g.huber_delta = huber_delta;
  static const double ref_info[6] = {0.01, 0.01, 0.01, 1.0, 1.0, 1.0};      // src/pose_graph_optimizer.cpp:23-26
  for (int a = 0; a < 6; ++a) g.info[a] = information6 ? information6[a] : ref_info[a];
  if ((rc = h->upload(&g.fv, fv))) return rc;
  if ((rc = h->upload(&g.ei, ei))) return rc;
  if ((rc = h->upload(&g.ej, ej))) return rc;
  if ((rc = h->upload(&g.Z, Zc))) return rc;
  { const double* xp = nullptr; if ((rc = h->upload(&xp, X))) return rc; g.X = const_cast<double*>(xp); }
  if ((rc = h->upload(&g.blk_part_off, bpo))) return rc;
  if ((rc = h->upload(&g.blk_part, bp))) return rc;
  if ((rc = h->upload(&g.vtx_part_off, vpo))) return rc;
  if ((rc = h->upload(&g.vtx_part, vp))) return rc;
  if ((rc = h->upload(&g.blk_row, blk_row))) return rc;
  if ((rc = h->upload(&g.blk_col, blk_col))) return rc;
  if ((rc = h->alloc(&g.Xn, (size_t)n * 12))) return rc;
  if ((rc = h->alloc(&g.epart, (size_t)m * 120))) return rc;
  if ((rc = h->alloc(&g.echi, (size_t)m))) return rc;
  if ((rc = h->alloc(&g.H0, (size_t)nbu * 36))) return rc;
  if ((rc = h->alloc(&g.b0, (size_t)nf * 6))) return rc;
  if ((rc = h->alloc(&g.Su, (size_t)nbu * 36))) return rc;
  if ((rc = h->alloc(&g.rhs, (size_t)nf * 6))) return rc;
  if ((rc = h->alloc(&g.x, (size_t)nf * 6))) return rc;
  if ((rc = h->alloc(&g.red, 4))) return rc;
  // the factorisation kernel sees the system through a Dev: one window, every column in it
  Dev dd{};
  dd.nwin = 1; dd.Nf = nf; dd.nbu = nbu; dd.Su = g.Su; dd.rhs = g.rhs; dd.x = g.x;
  int32_t* flags = nullptr; int32_t* fwin = nullptr;
  if ((rc = h->alloc(&dd.r, (size_t)nf * 6))) return rc;
  if ((rc = h->alloc(&flags, 4))) return rc;
  if ((rc = h->alloc(&fwin, (size_t)nf))) return rc;
  CUDA_TRY(cudaMemsetAsync(flags, 0, 16, h->stream));
  CUDA_TRY(cudaMemsetAsync(fwin, 0, (size_t)nf * 4, h->stream));
  dd.w_status = flags; dd.w_fail = flags + 1; dd.w_pcg_it = flags + 2; dd.fcam_win = fwin;
  SpCholDev sp{}; size_t sp_smem = 0; int sp_blocks = 0;
  if ((rc = setup_spchol(h, Y, &sp, &sp_smem, &sp_blocks))) return rc;
  if (sp_blocks <= 0) { set_error("sba_pgo_optimize: cooperative launch unavailable"); return SBA_CUDA_ERROR; }
  CUDA_TRY(cudaStreamSynchronize(h->stream));          // the host vectors above are locals
  auto read_red = [&](double* r3) -> int {
    CUDA_TRY(cudaMemcpyAsync(r3, g.red, 24, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    return SBA_OK;
  };
  const int ge = cdiv(m, 128), gv = cdiv(n, 128);
  // ---- g2o OptimizationAlgorithmLevenberg restated (oracle/pgo_oracle.py optimize)
  double lambda = -1.0, ni = 2.0, cur = 0.0, red[3];
  for (int it = 0; it < iterations; ++it) {
    LAUNCH(h, k_pgo_linearize, ge, 128, g);
    LAUNCH(h, k_pgo_assemble, nbu + nf, 64, g);
    LAUNCH(h, k_pgo_reduce, 1, 256, g, 0.0);
    if ((rc = read_red(red))) return rc;
    cur = red[0];
    if (it == 0) out->initial_chi2 = cur;
    if (lambda < 0.0) lambda = 1e-5 * std::max(red[2], 0.0);
    double rho = 0.0;
    int qmax = 0;
    do {
      LAUNCH(h, k_pgo_damp, cdiv(std::max((int64_t)nbu * 36, (int64_t)nf * 6), 256), 256, g, lambda);
      CUDA_TRY(cudaMemsetAsync(flags + 1, 0, 4, h->stream));
      { Dev d2 = dd; SpCholDev s2 = sp; void* args[] = {&d2, &s2};
        CUDA_TRY(cudaLaunchCooperativeKernel((const void*)k_spchol, dim3(sp_blocks), dim3(kSpThreads), args, sp_smem, h->stream)); h->launches++; }
      LAUNCH(h, k_pgo_update, gv, 128, g);
      LAUNCH(h, k_pgo_chi2, ge, 128, g);
      LAUNCH(h, k_pgo_reduce, 1, 256, g, lambda);
      int32_t failed = 0;
      CUDA_TRY(cudaMemcpyAsync(&failed, flags + 1, 4, cudaMemcpyDeviceToHost, h->stream));
      if ((rc = read_red(red))) return rc;
      const double tmp = (failed || !std::isfinite(red[0])) ? std::numeric_limits<double>::infinity() : red[0];
      const double scale = red[1] + 1e-3;
      rho = (cur - tmp) / scale;
      out->trials++;
      if (rho > 0.0 && std::isfinite(tmp)) {
        const double t = 2.0 * rho - 1.0;
        const double alpha = std::min(1.0 - t * t * t, 2.0 / 3.0);
        lambda *= std::max(1.0 / 3.0, alpha); ni = 2.0;
        cur = tmp;
        LAUNCH(h, k_pgo_accept, cdiv((int64_t)n * 12, 256), 256, g);
      } else {
        lambda *= ni; ni *= 2.0;
        if (!std::isfinite(lambda)) break;
      }
      ++qmax;
    } while (rho < 0.0 && qmax < 10);
    out->iterations = it + 1;
    if (qmax == 10 || rho == 0.0 || !std::isfinite(lambda)) { out->termination = 1; break; }
  }
  out->final_chi2 = cur; out->final_lambda = lambda;
  CUDA_TRY(cudaMemcpyAsync(X.data(), g.X, X.size() * 8, cudaMemcpyDeviceToHost, h->stream));
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  for (int v = 0; v < n; ++v) {
    double* T = T_wc + (size_t)v * 16; const double* o = &X[(size_t)v * 12];
    for (int r = 0; r < 3; ++r) { for (int c = 0; c < 3; ++c) T[4 * r + c] = o[3 * r + c]; T[4 * r + 3] = o[9 + r]; }
    T[12] = T[13] = T[14] = 0.0; T[15] = 1.0;
  }
  out->time_total_s = std::chrono::duration<double>(clk::now() - t0).count();
  return SBA_OK;
}

int sba_pgo_optimize(sba_handle h, int32_t num_vertices, double* T_wc, int32_t num_edges, const int32_t* edge_i, const int32_t* edge_j,
                     const double* Z, const uint8_t* fixed, const double* information6, double huber_delta, int32_t iterations,
                     sba_pgo_summary* summary) {
  if (!h || num_vertices < 0 || num_edges < 0 || (num_vertices > 0 && !T_wc) || (num_edges > 0 && (!edge_i || !edge_j || !Z)) ||
      !(huber_delta > 0.0) || iterations < 0) { set_error("sba_pgo_optimize: bad arguments"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) { set_error("host-only handle: no CUDA device bound (there is no CPU fallback)"); return SBA_CUDA_ERROR; }
  sba_pgo_summary S; std::memset(&S, 0, sizeof(S));
  CUDA_TRY(cudaSetDevice(h->device));
  // the graph's buffers live in their own allocation list, released when the call returns (the BA structure of the
  // handle, if any, is untouched)
  std::vector<void*> keep;
  keep.swap(h->allocs);
  const int rc = pgo_run(h, num_vertices, T_wc, num_edges, edge_i, edge_j, Z, fixed, information6, huber_delta, iterations, &S);
  for (void* p : h->allocs) cudaFreeAsync(p, h->stream);
  cudaStreamSynchronize(h->stream);
  h->allocs.swap(keep);
  if (summary) *summary = S;
  return rc;
}

// ---- N4: stereo RANSAC alignment (ransac.cuh) --------------------------------------------------------
void sba_ransac_default_options(sba_ransac_options* o) {
  if (!o) return;
  o->max_iterations = 100;          // src/params.h:27
  o->final_model_fitting = 0;       // :18
  o->early_termination = 0;         // :21
  o->concensus_ratio = 0.8f;        // :24
  o->inlier_threshold = 3.0f;       // ransac_solver/include/stereo_reprojection.h:126
}

static int ransac_run(sba_handle h, int32_t nprob, const int64_t* off, const float* a, const float* b, const float* P_l, const float* P_r,
                      const sba_ransac_options& o, const int32_t* samples, uint64_t seed, float* T, uint8_t* labels, float* losses,
                      int32_t* best_it, int32_t* counts) {
  int rc;
  const int64_t n = off[nprob];
  const int iters = o.max_iterations;
  const size_t nh = (size_t)nprob * std::max(iters, 1);
  RansacDev d{};
  d.nprob = nprob; d.iters = iters; d.seed = seed; d.thr_px = o.inlier_threshold; d.ratio = o.concensus_ratio;
  d.final_fit = o.final_model_fitting ? 1 : 0; d.early = o.early_termination ? 1 : 0;
  std::memcpy(d.Pl, P_l, 48); std::memcpy(d.Pr, P_r, 48);
  // one device arena, one packed upload, one packed download: a single frame pair (the visual odometer's call) is
  // latency-bound, every separate allocation / copy costs more than the kernels
  auto up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
  const bool have_s = samples && iters > 0;
  const size_t in_a = 0, in_b = in_a + (size_t)n * 16, in_off = in_b + (size_t)n * 16, in_s = up16(in_off + ((size_t)nprob + 1) * 8);
  const size_t in_bytes = up16(in_s + (have_s ? nh * 12 : 0));
  const size_t o_T = 0, o_loss = o_T + (size_t)nprob * 64, o_best = up16(o_loss + (size_t)n * 4), o_cnt = up16(o_best + (size_t)nprob * 4);
  const size_t o_lab = up16(o_cnt + nh * 4), out_bytes = up16(o_lab + (size_t)n);
  const size_t w_src = in_bytes + out_bytes, w_dst = w_src + (size_t)n * 16, w_T = w_dst + (size_t)n * 16, total = w_T + nh * 48;
  uint8_t* arena = nullptr;
  if ((rc = h->alloc(&arena, total))) return rc;
  if (h->stage_busy) { CUDA_TRY(cudaEventSynchronize(h->ev_stage)); h->stage_busy = false; }      // the buffer is shared with sba_set_structure
  if (h->stage_bytes < std::max(in_bytes, out_bytes)) {
    if (h->stage) cudaFreeHost(h->stage);
    h->stage = nullptr; h->stage_bytes = 0;
    const size_t want = std::max<size_t>(std::max(in_bytes, out_bytes) * 2, 1 << 16);
    CUDA_TRY(cudaMallocHost((void**)&h->stage, want));
    h->stage_bytes = want;
  }
  uint8_t* hs = h->stage;
  if (n > 0) { std::memcpy(hs + in_a, a, (size_t)n * 16); std::memcpy(hs + in_b, b, (size_t)n * 16); }
  std::memcpy(hs + in_off, off, ((size_t)nprob + 1) * 8);
  if (have_s) std::memcpy(hs + in_s, samples, nh * 12);
  CUDA_TRY(cudaMemcpyAsync(arena, hs, in_bytes, cudaMemcpyHostToDevice, h->stream));
  d.first = (const float4*)(arena + in_a); d.second = (const float4*)(arena + in_b); d.off = (const int64_t*)(arena + in_off);
  d.samples = have_s ? (const int32_t*)(arena + in_s) : nullptr;
  uint8_t* ob = arena + in_bytes;
  d.T_out = (float*)(ob + o_T); d.losses = (float*)(ob + o_loss); d.best_it = (int32_t*)(ob + o_best); d.counts = (int32_t*)(ob + o_cnt);
  d.labels = ob + o_lab;
  d.src = (float4*)(arena + w_src); d.dst = (float4*)(arena + w_dst); d.Thyp = (float*)(arena + w_T);
  LAUNCH(h, k_ransac_triangulate, (unsigned)cdiv(2 * n, 128), 128, d, n, 2 * n);
  if (iters > 0) {
    k_ransac_hypotheses<<<dim3((unsigned)iters, (unsigned)nprob), kRansacThreads, 0, h->stream>>>(d);
    h->launches++;
  }
  LAUNCH(h, k_ransac_finish, nprob, kRansacFinishThreads, d);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpyAsync(hs, ob, out_bytes, cudaMemcpyDeviceToHost, h->stream));   // the upload has been consumed: stream order
  CUDA_TRY(cudaStreamSynchronize(h->stream));
  std::memcpy(T, hs + o_T, (size_t)nprob * 64);
  if (n > 0) { std::memcpy(labels, hs + o_lab, (size_t)n); std::memcpy(losses, hs + o_loss, (size_t)n * 4); }
  if (best_it) std::memcpy(best_it, hs + o_best, (size_t)nprob * 4);
  if (counts && iters > 0) std::memcpy(counts, hs + o_cnt, nh * 4);
  return SBA_OK;
}

int sba_ransac_stereo(sba_handle h, int32_t num_problems, const int64_t* offsets, const float* obs_first, const float* obs_second,
                      const float* P_l, const float* P_r, const sba_ransac_options* options, const int32_t* samples, uint64_t seed,
                      float* T, uint8_t* labels, float* losses, int32_t* best_iteration, int32_t* counts) {
  if (!h || num_problems < 0 || !offsets || !P_l || !P_r || (num_problems > 0 && !T)) { set_error("sba_ransac_stereo: bad arguments"); return SBA_INVALID_ARGUMENT; }
  sba_ransac_options o;
  if (options) o = *options; else sba_ransac_default_options(&o);
  if (o.max_iterations < 0 || o.max_iterations > 65535 || !(o.inlier_threshold > 0.f)) { set_error("sba_ransac_stereo: bad options (0 <= max_iterations <= 65535, inlier_threshold > 0)"); return SBA_INVALID_ARGUMENT; }
  if (num_problems > 65535) { set_error("sba_ransac_stereo: at most 65535 problems per call"); return SBA_INVALID_ARGUMENT; }
  if (offsets[0] != 0) { set_error("sba_ransac_stereo: offsets[0] must be 0"); return SBA_INVALID_ARGUMENT; }
  for (int b = 0; b < num_problems; ++b) {
    const int64_t nb = offsets[b + 1] - offsets[b];
    if (nb < 0) { set_error("sba_ransac_stereo: offsets must be non-decreasing"); return SBA_INVALID_ARGUMENT; }
    if (samples)
      for (int64_t k = 0; k < (int64_t)o.max_iterations * 3 && nb > 0; ++k) {
        const int32_t s = samples[((int64_t)b * o.max_iterations) * 3 + k];
        if (s < 0 || s >= nb) { set_error("sba_ransac_stereo: sample index out of range"); return SBA_INVALID_ARGUMENT; }
      }
  }
  const int64_t n = offsets[num_problems];
  if (n > 0 && (!obs_first || !obs_second || !labels || !losses)) { set_error("sba_ransac_stereo: null element / output arrays"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) { set_error("host-only handle: no CUDA device bound (there is no CPU fallback)"); return SBA_CUDA_ERROR; }
  if (num_problems == 0) return SBA_OK;
  CUDA_TRY(cudaSetDevice(h->device));
  std::vector<void*> keep;                 // call-local buffers, released on return (the handle's BA structure is untouched)
  keep.swap(h->allocs);
  const int rc = ransac_run(h, num_problems, offsets, obs_first, obs_second, P_l, P_r, o, samples, seed, T, labels, losses, best_iteration, counts);
  for (void* p : h->allocs) cudaFreeAsync(p, h->stream);
  cudaStreamSynchronize(h->stream);
  h->allocs.swap(keep);
  return rc;
}

int sba_triangulate_stereo(sba_handle h, int64_t n, const float* obs, const float* P_l, const float* P_r, float* xyz) {
  if (!h || n < 0 || !P_l || !P_r || (n > 0 && (!obs || !xyz))) { set_error("sba_triangulate_stereo: bad arguments"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) { set_error("host-only handle: no CUDA device bound (there is no CPU fallback)"); return SBA_CUDA_ERROR; }
  if (n == 0) return SBA_OK;
  CUDA_TRY(cudaSetDevice(h->device));
  std::vector<void*> keep;
  keep.swap(h->allocs);
  auto run = [&]() -> int {
    int rc;
    RansacDev d{};
    std::memcpy(d.Pl, P_l, 48); std::memcpy(d.Pr, P_r, 48);
    float4* d_a = nullptr;
    if ((rc = h->alloc(&d_a, (size_t)n))) return rc;
    if ((rc = h->alloc(&d.src, (size_t)n))) return rc;
    CUDA_TRY(cudaMemcpyAsync(d_a, obs, (size_t)n * 16, cudaMemcpyHostToDevice, h->stream));
    d.first = d_a;
    LAUNCH(h, k_ransac_triangulate, (unsigned)cdiv(n, 128), 128, d, n, n);
    CUDA_TRY(cudaGetLastError());
    std::vector<float> tmp((size_t)n * 4);
    CUDA_TRY(cudaMemcpyAsync(tmp.data(), d.src, (size_t)n * 16, cudaMemcpyDeviceToHost, h->stream));
    CUDA_TRY(cudaStreamSynchronize(h->stream));
    for (int64_t i = 0; i < n; ++i) { xyz[3 * i] = tmp[4 * i]; xyz[3 * i + 1] = tmp[4 * i + 1]; xyz[3 * i + 2] = tmp[4 * i + 2]; }
    return SBA_OK;
  };
  const int rc = run();
  for (void* p : h->allocs) cudaFreeAsync(p, h->stream);
  cudaStreamSynchronize(h->stream);
  h->allocs.swap(keep);
  return rc;
}

// ---- allocation-free RANSAC (no counterpart in the reference) ----------------
// Scratch of sba_ransac_stereo_noalloc: one fixed array per GPU that loads this library.  In HOST code the name below
// is only a stand-in for the symbol — its device address comes from cudaGetSymbolAddress, once per device.
namespace {
constexpr size_t kNaSrc = 0;                                                              // float4 [MAX_ELEMENTS]
constexpr size_t kNaDst = kNaSrc + (size_t)SBA_NOALLOC_MAX_ELEMENTS * 16;                 // float4 [MAX_ELEMENTS]
constexpr size_t kNaThyp = kNaDst + (size_t)SBA_NOALLOC_MAX_ELEMENTS * 16;                // float  [MAX_HYPOTHESES * 12]
constexpr size_t kNaCounts = kNaThyp + (size_t)SBA_NOALLOC_MAX_HYPOTHESES * 48;           // int32  [MAX_HYPOTHESES]
constexpr size_t kNaBest = kNaCounts + (size_t)SBA_NOALLOC_MAX_HYPOTHESES * 4;            // int32  [MAX_HYPOTHESES] (one per problem)
constexpr size_t kNaBytes = kNaBest + (size_t)SBA_NOALLOC_MAX_HYPOTHESES * 4;
constexpr int kNaMaxDevices = 64;
}  // namespace
static __device__ __align__(256) unsigned char g_ransac_noalloc_scratch[kNaBytes];
static unsigned char* g_na_scratch_ptr[kNaMaxDevices] = {};      // device addresses of the array, by device ordinal
static std::mutex g_na_mutex;

// device address of the scratch array on h's device (resolved on first use, with that device current)
static int noalloc_scratch(sba_handle h, unsigned char** out) {
  if (h->device < 0 || h->device >= kNaMaxDevices) { set_error("sba_ransac_stereo_noalloc: device ordinal out of range"); return SBA_INVALID_ARGUMENT; }
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_scratch_ptr[h->device]) {
    CUDA_TRY(cudaSetDevice(h->device));
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_ransac_noalloc_scratch);
    if (e != cudaSuccess || !p) { set_error(std::string("cudaGetSymbolAddress(scratch): ") + cudaGetErrorString(e)); return SBA_CUDA_ERROR; }
    g_na_scratch_ptr[h->device] = static_cast<unsigned char*>(p);
  }
  *out = g_na_scratch_ptr[h->device];
  return SBA_OK;
}

int sba_ransac_stereo_noalloc(sba_handle h, int32_t num_problems, int64_t num_elements, const int64_t* d_offsets,
                              const float* d_obs_first, const float* d_obs_second, const float* P_l, const float* P_r,
                              const sba_ransac_options* options, const int32_t* d_samples, uint64_t seed, float* d_T,
                              uint8_t* d_labels, float* d_losses, int32_t* d_best_iteration, int32_t* d_counts, void* stream) {
  if (!h || num_problems < 0 || num_elements < 0 || !d_offsets || !P_l || !P_r || (num_problems > 0 && !d_T) ||
      (num_elements > 0 && (!d_obs_first || !d_obs_second || !d_labels || !d_losses))) { set_error("sba_ransac_stereo_noalloc: bad arguments"); return SBA_INVALID_ARGUMENT; }
  if (h->device < 0) { set_error("host-only handle: no CUDA device bound (there is no CPU fallback)"); return SBA_CUDA_ERROR; }
  sba_ransac_options o;
  if (options) o = *options; else sba_ransac_default_options(&o);
  if (o.max_iterations < 0 || o.max_iterations > 65535 || !(o.inlier_threshold > 0.f)) { set_error("sba_ransac_stereo_noalloc: bad options (0 <= max_iterations <= 65535, inlier_threshold > 0)"); return SBA_INVALID_ARGUMENT; }
  if (num_problems > 65535) { set_error("sba_ransac_stereo_noalloc: at most 65535 problems per call"); return SBA_INVALID_ARGUMENT; }
  const int64_t nhyp = (int64_t)num_problems * std::max(o.max_iterations, 1);
  if (num_elements > SBA_NOALLOC_MAX_ELEMENTS) {
    set_error("sba_ransac_stereo_noalloc: " + std::to_string(num_elements) + " elements exceed the fixed scratch (SBA_NOALLOC_MAX_ELEMENTS = " +
              std::to_string(SBA_NOALLOC_MAX_ELEMENTS) + "); use sba_ransac_stereo");
    return SBA_INVALID_ARGUMENT;
  }
  if (nhyp > SBA_NOALLOC_MAX_HYPOTHESES) {
    set_error("sba_ransac_stereo_noalloc: " + std::to_string(nhyp) + " hypotheses (problems x iterations) exceed the fixed scratch (SBA_NOALLOC_MAX_HYPOTHESES = " +
              std::to_string(SBA_NOALLOC_MAX_HYPOTHESES) + "); use sba_ransac_stereo");
    return SBA_INVALID_ARGUMENT;
  }
  if ((reinterpret_cast<uintptr_t>(d_obs_first) & 15) || (reinterpret_cast<uintptr_t>(d_obs_second) & 15)) { set_error("sba_ransac_stereo_noalloc: observation arrays must be 16-byte aligned"); return SBA_INVALID_ARGUMENT; }
  if (num_problems == 0) return SBA_OK;
  CUDA_TRY(cudaSetDevice(h->device));
  unsigned char* scratch = nullptr;
  { const int rc = noalloc_scratch(h, &scratch); if (rc) return rc; }
  cudaStream_t st = stream ? static_cast<cudaStream_t>(stream) : h->stream;
  const int iters = o.max_iterations;
  RansacDev d{};
  d.nprob = num_problems; d.iters = iters; d.seed = seed; d.thr_px = o.inlier_threshold; d.ratio = o.concensus_ratio;
  d.final_fit = o.final_model_fitting ? 1 : 0; d.early = o.early_termination ? 1 : 0;
  std::memcpy(d.Pl, P_l, 48); std::memcpy(d.Pr, P_r, 48);
  d.off = d_offsets; d.first = reinterpret_cast<const float4*>(d_obs_first); d.second = reinterpret_cast<const float4*>(d_obs_second);
  d.samples = iters > 0 ? d_samples : nullptr;
  d.src = reinterpret_cast<float4*>(scratch + kNaSrc); d.dst = reinterpret_cast<float4*>(scratch + kNaDst);
  d.Thyp = reinterpret_cast<float*>(scratch + kNaThyp);
  d.counts = d_counts ? d_counts : reinterpret_cast<int32_t*>(scratch + kNaCounts);
  d.best_it = d_best_iteration ? d_best_iteration : reinterpret_cast<int32_t*>(scratch + kNaBest);
  d.T_out = d_T; d.labels = d_labels; d.losses = d_losses;
  // the array keeps the previous call's contents: a problem without elements writes no model, and one without
  // iterations no count — both regions are read by the finish kernel and start from zero
  CUDA_TRY(cudaMemsetAsync(d.Thyp, 0, (size_t)nhyp * 48, st));
  if (iters > 0) CUDA_TRY(cudaMemsetAsync(d.counts, 0, (size_t)num_problems * iters * 4, st));
  const int64_t n = num_elements;
  if (n > 0) { k_ransac_triangulate<<<(unsigned)cdiv(2 * n, 128), 128, 0, st>>>(d, n, 2 * n); h->launches++; }
  if (iters > 0) { k_ransac_hypotheses<<<dim3((unsigned)iters, (unsigned)num_problems), kRansacThreads, 0, st>>>(d); h->launches++; }
  k_ransac_finish<<<num_problems, kRansacFinishThreads, 0, st>>>(d); h->launches++;
  CUDA_TRY(cudaGetLastError());
  if (!stream) CUDA_TRY(cudaStreamSynchronize(st));
  return SBA_OK;
}

// ---- multi-GPU --------------------------------------------------------------------------------------
int sba_comm_unique_id(uint8_t id[128]) {
  std::string err;
  if (!nccl_api().load(&err)) { set_error(err); return SBA_NCCL_ERROR; }
  ncclUniqueId uid;
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  if (nccl_api().GetUniqueId(&uid) != ncclSuccess) { set_error("ncclGetUniqueId failed"); return SBA_NCCL_ERROR; }
  std::memcpy(id, &uid, 128);
  return SBA_OK;
}

int sba_comm_init(sba_handle h, int32_t rank, int32_t world_size, const uint8_t id[128]) {
  if (!h || h->device < 0 || world_size < 1 || world_size > kMaxRanks || rank < 0 || rank >= world_size) { set_error("bad comm arguments (1 <= world_size <= 16)"); return SBA_INVALID_ARGUMENT; }
  std::string err;
  if (!nccl_api().load(&err)) { set_error(err); return SBA_NCCL_ERROR; }
  CUDA_TRY(cudaSetDevice(h->device));
  h->comm.destroy();
  ncclUniqueId uid; std::memcpy(&uid, id, 128);
  ncclComm_t c = nullptr;
  const ncclResult_t r = nccl_api().CommInitRank(&c, world_size, uid, rank);
  if (r != ncclSuccess) { set_error(std::string("ncclCommInitRank: ") + (nccl_api().GetErrorString ? nccl_api().GetErrorString(r) : "error")); return SBA_NCCL_ERROR; }
  h->comm.comm = c; h->comm.rank = rank; h->comm.world = world_size;
  return SBA_OK;
}

}  // extern "C"
