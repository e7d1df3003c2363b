// structure.cpp — one-off host build of the index tables the kernels run on: the point-sorted observation
// view, the chunks with their camera slots and local blocks, and the block pattern of the reduced camera
// matrix with the static partial lists that make every reduction fixed-order and atomic-free.
//
// reference: the observation -> (camera, point) indexing comes from the pack loop
// (src/bundle_adjuster.cpp:57-110); the gauge (first camera constant, :113) decides which cameras
// are free.  The pattern rule is SURVEY §8 R10: block (i,j) of the reduced camera matrix is
// structurally non-zero iff some point is observed by both free cameras i and j.
#include <cstdio>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "sba_internal.h"

namespace sba {

namespace {

struct ChunkRange { int32_t p0, p1, win; bool owned; bool is_long; };
struct LB { uint8_t sa, sb; int64_t key; };
struct ChunkTab {
  std::vector<int32_t> cams, slot_obs_off;
  std::vector<uint8_t> oslot;
  std::vector<uint16_t> sobs;
  std::vector<LB> lbs;
  std::vector<int32_t> lb_pair_cnt;     // [nlb]
  std::vector<uint16_t> pairs;          // grouped by lb
};
struct Work {
  std::vector<ChunkRange> cr;
  std::vector<ChunkTab> tab;
  std::vector<int32_t> gpt_off, gperm, pt_win, tmp;
  std::vector<int64_t> hist;
};

// chunk boundaries over ALL points (unowned ones only feed the block pattern): never across a window, never
// across the owned range, <= kPointChunk points, <= obs_cap observations
bool make_chunks(const HostStructure& H, const std::vector<int32_t>& gpt_off, int obs_cap, std::vector<ChunkRange>* out,
                 std::string* err) {
  out->clear();
  for (int w = 0; w < H.nwin; ++w) {
    int p = H.win_pt_off[w];
    const int pend = H.win_pt_off[w + 1];
    while (p < pend) {
      const bool owned = (p >= H.pt_begin && p < H.pt_end);
      int limit = pend;
      if (owned) limit = std::min(limit, H.pt_end);
      else if (p < H.pt_begin) limit = std::min(limit, H.pt_begin);
      int q = p;
      int nobs = 0;
      bool is_long = false;
      while (q < limit && q - p < kPointChunk) {
        const int k = gpt_off[q + 1] - gpt_off[q];
        if (k > kMaxLongObs) { *err = "a point has more than 4096 observations (unsupported track length)"; return false; }
        if (k > kMaxSlots) {       // long track: a chunk of its own for the k_*_long kernels
          if (q > p) break;
          is_long = true; nobs = k; ++q;
          break;
        }
        // a point never straddles chunks; one whose track alone exceeds the tuning cap (obs_cap < k <= 255) gets a
        // chunk of its own — chunk-local indices are bytes and the per-observation kernels run 256 threads
        if (nobs + k > obs_cap && q > p) break;
        nobs += k; ++q;
      }
      if (q == p) { *err = "internal: empty chunk"; return false; }
      out->push_back(ChunkRange{p, q, w, owned, is_long});
      p = q;
    }
  }
  return true;
}

}  // namespace

std::string build_structure(int32_t Nc, int32_t Np, int64_t NobsAll, const int32_t* obs_cam, const int32_t* obs_pt,
                            const uint8_t* cam_fixed, int32_t nwin, const int32_t* win_cam_off_in,
                            const int32_t* win_pt_off_in, int32_t pt_begin, int32_t pt_end, HostStructure* S) {
  if (Nc < 0 || Np < 0 || NobsAll < 0) return "negative size";
  if (NobsAll > 2000000000LL) return "too many observations for 32-bit positions";
  if (nwin < 1) nwin = 1;
  if (pt_begin < 0 || pt_end > Np || pt_begin > pt_end) return "bad point shard";
  HostStructure& H = *S;
  // vectors keep their capacity from the previous build on this handle (every one is re-sized / re-assigned below)
  H.Nf = 0; H.Nobs = 0; H.obs_cap = kObsChunk;
  H.max_slots = H.max_chunk_obs = H.max_chunk_pts = H.max_chunk_pairs = 0;
  H.long_chunks.clear(); H.long_zoff.clear(); H.long_lb_off.clear();
  if (!H.work) H.work = std::make_shared<Work>();
  Work& Wk = *static_cast<Work*>(H.work.get());
  const bool timing = std::getenv("SBA_TIMING") != nullptr;
  // host threads of the structure build: SBA_HOST_THREADS, else the OpenMP default (launchers such as torchrun set
  // OMP_NUM_THREADS=1 for every rank, which is the wrong default for this short, embarrassingly parallel build)
  int nthreads = omp_get_max_threads();
  if (const char* e = std::getenv("SBA_HOST_THREADS")) nthreads = std::max(1, std::atoi(e));
  auto tprev = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!timing) return;
    const auto t = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[sba]   structure %-28s %.3f ms\n", what, 1e3 * std::chrono::duration<double>(t - tprev).count());
    tprev = t;
  };
  H.Nc = Nc; H.Np = Np; H.nwin = nwin; H.pt_begin = pt_begin; H.pt_end = pt_end;

  // ---- windows
  H.win_cam_off.resize(nwin + 1); H.win_pt_off.resize(nwin + 1);
  if (win_cam_off_in && win_pt_off_in) {
    for (int w = 0; w <= nwin; ++w) { H.win_cam_off[w] = win_cam_off_in[w]; H.win_pt_off[w] = win_pt_off_in[w]; }
    if (H.win_cam_off[0] != 0 || H.win_cam_off[nwin] != Nc || H.win_pt_off[0] != 0 || H.win_pt_off[nwin] != Np)
      return "window offsets must cover [0,Nc) and [0,Np)";
    for (int w = 0; w < nwin; ++w)
      if (H.win_cam_off[w] > H.win_cam_off[w + 1] || H.win_pt_off[w] > H.win_pt_off[w + 1]) return "window offsets not monotone";
  } else {
    if (nwin != 1) return "num_windows > 1 needs window offsets";
    H.win_cam_off[0] = 0; H.win_cam_off[1] = Nc; H.win_pt_off[0] = 0; H.win_pt_off[1] = Np;
  }
  H.cam_win.resize(Nc);
  std::vector<int32_t>& pt_win = Wk.pt_win;
  pt_win.resize(Np);
  for (int w = 0; w < nwin; ++w) {
    for (int c = H.win_cam_off[w]; c < H.win_cam_off[w + 1]; ++c) H.cam_win[c] = w;
    for (int p = H.win_pt_off[w]; p < H.win_pt_off[w + 1]; ++p) pt_win[p] = w;
  }
  // ---- gauge: free cameras
  H.fc.resize(Nc); H.win_fc_off.assign(nwin + 1, 0);
  int Nf = 0;
  for (int w = 0; w < nwin; ++w) {
    H.win_fc_off[w] = Nf;
    for (int c = H.win_cam_off[w]; c < H.win_cam_off[w + 1]; ++c) {
      const bool fixed = cam_fixed ? (cam_fixed[c] != 0) : (c == H.win_cam_off[w]);
      H.fc[c] = fixed ? -1 : Nf++;
    }
  }
  H.win_fc_off[nwin] = Nf; H.Nf = Nf;
  H.free2cam.resize(Nf); H.fcam_win.resize(Nf);
  for (int c = 0; c < Nc; ++c) if (H.fc[c] >= 0) { H.free2cam[H.fc[c]] = c; H.fcam_win[H.fc[c]] = H.cam_win[c]; }

  // ---- validate + global point-order counting sort (stable in input order)
  std::vector<int32_t>& gpt_off = Wk.gpt_off;
  gpt_off.assign((size_t)Np + 1, 0);
  std::vector<int32_t>& gperm = Wk.gperm;
  gperm.resize((size_t)NobsAll);
  if (NobsAll < 65536 || nthreads < 2) {
    for (int64_t k = 0; k < NobsAll; ++k) {
      const int32_t c = obs_cam[k], p = obs_pt[k];
      if (c < 0 || c >= Nc) return "obs_cam out of range";
      if (p < 0 || p >= Np) return "obs_pt out of range";
      if (H.cam_win[c] != pt_win[p]) return "observation links a camera and a point of different windows";
      gpt_off[p + 1]++;
    }
    for (int p = 0; p < Np; ++p) gpt_off[p + 1] += gpt_off[p];
    std::vector<int32_t> cur(gpt_off.begin(), gpt_off.end() - 1);
    for (int64_t k = 0; k < NobsAll; ++k) gperm[cur[obs_pt[k]]++] = (int32_t)k;
  } else {
    // two stable levels, both parallel: (1) every thread scatters its contiguous slice of the input into buckets of
    // 2^shift consecutive points (per-thread histograms, offsets ordered bucket-major then thread-major, so equal keys
    // keep their input order); (2) a counting sort inside every bucket, buckets in parallel
    int shift = 0;
    while ((Np >> shift) > 2048) ++shift;
    const int B = (Np >> shift) + 1, T = nthreads;
    std::vector<int64_t>& hist = Wk.hist; hist.assign((size_t)T * B, 0);
    std::vector<int32_t>& tmp = Wk.tmp; tmp.resize((size_t)NobsAll);
    std::vector<int64_t> bstart((size_t)B + 1, 0);
    int bad = 0; int64_t bad_at = -1;      // the EARLIEST offending observation decides the message, as in the serial scan
#pragma omp parallel num_threads(T)
    {
      const int t = omp_get_thread_num();
      const int64_t k0 = NobsAll * t / T, k1 = NobsAll * (t + 1) / T;
      int64_t* h = hist.data() + (size_t)t * B;
      int mybad = 0; int64_t my_at = -1;
      for (int64_t k = k0; k < k1; ++k) {
        const int32_t c = obs_cam[k], p = obs_pt[k];
        if (c < 0 || c >= Nc) { mybad = 1; my_at = k; break; }
        if (p < 0 || p >= Np) { mybad = 2; my_at = k; break; }
        if (H.cam_win[c] != pt_win[p]) { mybad = 3; my_at = k; break; }
        h[p >> shift]++;
      }
      if (mybad) {
#pragma omp critical
        { if (!bad || my_at < bad_at) { bad = mybad; bad_at = my_at; } }
      }
#pragma omp barrier
#pragma omp single
      {
        if (!bad) {
          int64_t run = 0;
          for (int bb = 0; bb < B; ++bb) {
            bstart[bb] = run;
            for (int tt = 0; tt < T; ++tt) { const int64_t c = hist[(size_t)tt * B + bb]; hist[(size_t)tt * B + bb] = run; run += c; }
          }
          bstart[B] = run;
        }
      }     // implicit barrier
      if (!bad) {
        for (int64_t k = k0; k < k1; ++k) tmp[h[obs_pt[k] >> shift]++] = (int32_t)k;
#pragma omp barrier
#pragma omp for schedule(dynamic, 8)
        for (int bb = 0; bb < B; ++bb) {
          const int32_t q0 = (int32_t)std::min<int64_t>((int64_t)bb << shift, Np), q1 = (int32_t)std::min<int64_t>(((int64_t)bb + 1) << shift, Np);
          const int64_t s0 = bstart[bb], s1 = bstart[bb + 1];
          // counts of the bucket's points into gpt_off[p + 1] (each bucket owns its own slots), then local offsets
          for (int64_t j = s0; j < s1; ++j) gpt_off[obs_pt[tmp[j]] + 1]++;
          int64_t run = s0;
          for (int32_t q = q0; q < q1; ++q) { const int32_t c = gpt_off[q + 1]; gpt_off[q + 1] = (int32_t)run; run += c; }   // start of q, kept in slot q + 1
          for (int64_t j = s0; j < s1; ++j) { const int32_t k = tmp[j]; gperm[gpt_off[obs_pt[k] + 1]++] = k; }               // slot q + 1 ends as end of q
        }
      }
    }
    if (bad == 1) return "obs_cam out of range";
    if (bad == 2) return "obs_pt out of range";
    if (bad == 3) return "observation links a camera and a point of different windows";
    gpt_off[0] = 0;
  }
  lap("validate + counting sort");
  // ---- local (owned) slice of the point order
  const int32_t base = gpt_off[pt_begin];
  const int64_t Nobs = (int64_t)gpt_off[pt_end] - base;
  H.Nobs = Nobs;
  H.pt_off.resize(Np + 1);
  for (int p = 0; p <= Np; ++p) {
    int64_t v = (int64_t)gpt_off[p] - base;
    if (v < 0) v = 0;
    if (v > Nobs) v = Nobs;
    H.pt_off[p] = (int32_t)v;
  }
  H.perm_p.resize(Nobs); H.pcam.resize(Nobs); H.ppt.resize(Nobs);
#pragma omp parallel for schedule(static) num_threads(nthreads)
  for (int64_t e = 0; e < Nobs; ++e) {
    const int32_t k = gperm[base + e];
    H.perm_p[e] = k; H.pcam[e] = obs_cam[k]; H.ppt[e] = obs_pt[k];
  }

  lap("point-order views");
  // ---- chunks, camera slots, local blocks
  std::vector<ChunkRange>& cr = Wk.cr;
  std::vector<ChunkTab>& tab = Wk.tab;
  int obs_cap = kObsChunk;
  if (const char* e = std::getenv("SBA_OBS_CAP")) obs_cap = std::min(kObsChunk, std::max(8, std::atoi(e)));   // tuning aid
  for (int attempt = 0; attempt < 2; ++attempt) {
    std::string err;
    if (!make_chunks(H, gpt_off, obs_cap, &cr, &err)) return err;
    const int nch = (int)cr.size();
    if ((int)tab.size() < nch) tab.resize(nch);        // tables (and their capacity) of earlier builds are reused
    bool overflow = false;
#pragma omp parallel num_threads(nthreads)
    {
      std::vector<int32_t> slot_of_cam(Nc, -1);
      std::vector<uint8_t> bits;
      std::vector<int32_t> cnt, lb_of, cur;
#pragma omp for schedule(dynamic, 16)
      for (int ch = 0; ch < nch; ++ch) {
        const ChunkRange& R = cr[ch];
        ChunkTab& T = tab[ch];
        T.cams.clear(); T.lbs.clear(); T.pairs.clear(); T.lb_pair_cnt.clear(); T.oslot.clear(); T.sobs.clear(); T.slot_obs_off.clear();
        const int32_t g0 = gpt_off[R.p0], g1 = gpt_off[R.p1];
        const int n = g1 - g0;
        if (R.is_long) {
          // one slot per observation, one local block per ordered pair (a, b) of free observations with
          // fc[cam(b)] > fc[cam(a)], or the same camera (then both orders: Z_a Z_b^T and Z_b Z_a^T land on (f, f))
          T.cams.resize(n);
          for (int i = 0; i < n; ++i) T.cams[i] = obs_cam[gperm[g0 + i]];
          struct LL { int64_t key; uint16_t a, b; };
          std::vector<LL> ll;
          for (int a = 0; a < n; ++a) {
            const int64_t fa = H.fc[T.cams[a]];
            if (fa < 0) continue;
            for (int b = 0; b < n; ++b) {
              const int64_t fb = H.fc[T.cams[b]];
              if (fb < fa) continue;
              if (fb == fa && T.cams[a] != T.cams[b]) continue;
              ll.push_back(LL{fa * (int64_t)Nf + fb, (uint16_t)a, (uint16_t)b});
            }
          }
          std::stable_sort(ll.begin(), ll.end(), [](const LL& x, const LL& y) { return x.key < y.key; });
          T.lbs.resize(ll.size()); T.pairs.resize(ll.size() * 2); T.lb_pair_cnt.assign(ll.size(), 2);
          for (size_t l = 0; l < ll.size(); ++l) { T.lbs[l] = LB{0, 0, ll[l].key}; T.pairs[2 * l] = ll[l].a; T.pairs[2 * l + 1] = ll[l].b; }
          if (R.owned) {
            T.oslot.assign(n, 0);
            T.slot_obs_off.resize(n + 1); T.sobs.resize(n);
            for (int i = 0; i <= n; ++i) T.slot_obs_off[i] = i;
            for (int i = 0; i < n; ++i) T.sobs[i] = (uint16_t)i;
          }
          continue;
        }
        T.oslot.resize(n);
        bool bad = false;
        for (int i = 0; i < n; ++i) {
          const int32_t cam = obs_cam[gperm[g0 + i]];
          int32_t s = slot_of_cam[cam];
          if (s < 0) {
            s = (int32_t)T.cams.size();
            if (s >= kMaxSlots) { bad = true; break; }
            slot_of_cam[cam] = s; T.cams.push_back(cam);
          }
          T.oslot[i] = (uint8_t)s;
        }
        for (int32_t cam : T.cams) slot_of_cam[cam] = -1;
        if (bad) {
#pragma omp atomic write
          overflow = true;
          continue;
        }
        const int ns = (int)T.cams.size();
        // co-visibility inside the chunk: pairs of free slots sharing a point
        bits.assign((size_t)ns * ns, 0);
        for (int p = R.p0; p < R.p1; ++p) {
          const int a0 = gpt_off[p] - g0, a1 = gpt_off[p + 1] - g0;
          for (int a = a0; a < a1; ++a) {
            const int sa = T.oslot[a]; const int32_t fa = H.fc[T.cams[sa]];
            if (fa < 0) continue;
            for (int b = a0; b < a1; ++b) {
              const int sb = T.oslot[b]; const int32_t fb = H.fc[T.cams[sb]];
              if (fb < fa) continue;
              bits[(size_t)sa * ns + sb] = 1;
            }
          }
        }
        for (int sa = 0; sa < ns; ++sa)
          for (int sb = 0; sb < ns; ++sb)
            if (bits[(size_t)sa * ns + sb]) {
              const int64_t fa = H.fc[T.cams[sa]], fb = H.fc[T.cams[sb]];
              if (sa != sb && fa == fb) continue;   // cannot happen (slot <-> camera is 1:1); keep the diagonal once
              T.lbs.push_back(LB{(uint8_t)sa, (uint8_t)sb, fa * (int64_t)Nf + fb});
            }
        std::sort(T.lbs.begin(), T.lbs.end(), [](const LB& x, const LB& y) { return x.key < y.key; });
        if (!R.owned) { T.oslot.clear(); continue; }
        // observations grouped by slot (stable)
        cnt.assign(ns + 1, 0);
        for (int i = 0; i < n; ++i) cnt[T.oslot[i] + 1]++;
        for (int s = 0; s < ns; ++s) cnt[s + 1] += cnt[s];
        T.slot_obs_off.assign(cnt.begin(), cnt.end());
        T.sobs.resize(n);
        for (int i = 0; i < n; ++i) T.sobs[cnt[T.oslot[i]]++] = (uint16_t)i;
        // pair lists: for every local block the (a, b) observation pairs of one point, points ascending
        const int nl = (int)T.lbs.size();
        lb_of.assign((size_t)ns * ns, -1);
        for (int l = 0; l < nl; ++l) lb_of[(size_t)T.lbs[l].sa * ns + T.lbs[l].sb] = l;
        T.lb_pair_cnt.assign(nl, 0);
        for (int pass = 0; pass < 2; ++pass) {
          if (pass == 1) {
            cur.assign(nl + 1, 0);
            for (int l = 0; l < nl; ++l) cur[l + 1] = cur[l] + T.lb_pair_cnt[l];
            T.pairs.resize(cur[nl]);
          }
          for (int p = R.p0; p < R.p1; ++p) {
            const int a0 = gpt_off[p] - g0, a1 = gpt_off[p + 1] - g0;
            for (int a = a0; a < a1; ++a) {
              const int sa = T.oslot[a];
              if (H.fc[T.cams[sa]] < 0) continue;
              for (int b = a0; b < a1; ++b) {
                const int l = lb_of[(size_t)sa * ns + T.oslot[b]];
                if (l < 0) continue;
                if (pass == 0) T.lb_pair_cnt[l]++;
                else T.pairs[cur[l]++] = (uint16_t)(a | (b << 8));
              }
            }
          }
        }
      }
    }
    if (!overflow) break;
    if (attempt == 1) return "chunk with more than 255 distinct cameras";
    obs_cap = kMaxSlots;          // <= 255 observations per chunk guarantees <= 255 slots
  }
  H.obs_cap = obs_cap;
  const int nch_all = (int)cr.size();

  lap("chunk tables (parallel)");
  // ---- block pattern: union of all chunks' local blocks + the diagonal of every free camera
  std::vector<int64_t> keys;
  std::vector<int32_t> key_blk;        // small Nf: block id by key (dense Nf x Nf map)
  std::vector<uint64_t> bit_words;     // mid-size Nf: bitmap of the blocks + rank per word
  std::vector<int32_t> bit_rank;
  const bool dense_map = (int64_t)Nf * Nf <= ((int64_t)1 << 22);
  if (dense_map) {
    std::vector<uint8_t> mark((size_t)Nf * Nf, 0);
    for (int f = 0; f < Nf; ++f) mark[(size_t)f * Nf + f] = 1;
    for (int ch = 0; ch < nch_all; ++ch) for (const LB& b : tab[ch].lbs) mark[(size_t)b.key] = 1;
    key_blk.assign((size_t)Nf * Nf, -1);
    for (size_t k = 0; k < mark.size(); ++k) if (mark[k]) { key_blk[k] = (int32_t)keys.size(); keys.push_back((int64_t)k); }
  } else if ((int64_t)Nf * Nf <= ((int64_t)1 << 32)) {
    // up to 65 536 free cameras: one bit per block of the Nf x Nf grid (512 MB at most, 2.6 MB at KITTI-00 scale), marked
    // from the chunks' local blocks and read back in key order with a rank (prefix popcount) per word for the look-ups
    // below.  Sorting the ~650 000 local-block keys of a KITTI-00-sized map took 45 ms, a third of the whole build.
    const size_t nbits = (size_t)Nf * Nf, nwords = (nbits + 63) / 64;
    bit_words.assign(nwords, 0);
    for (int f = 0; f < Nf; ++f) { const size_t k = (size_t)f * Nf + f; bit_words[k >> 6] |= (uint64_t)1 << (k & 63); }
    for (int ch = 0; ch < nch_all; ++ch) for (const LB& b : tab[ch].lbs) bit_words[(size_t)b.key >> 6] |= (uint64_t)1 << ((size_t)b.key & 63);
    bit_rank.resize(nwords + 1);
    size_t tot = 0;
    for (size_t wd = 0; wd < nwords; ++wd) { bit_rank[wd] = (int32_t)tot; tot += (size_t)__builtin_popcountll(bit_words[wd]); }
    bit_rank[nwords] = (int32_t)tot;
    keys.resize(tot);
    size_t o = 0;
    for (size_t wd = 0; wd < nwords; ++wd) {
      uint64_t v = bit_words[wd];
      while (v) { keys[o++] = (int64_t)(wd * 64 + (size_t)__builtin_ctzll(v)); v &= v - 1; }
    }
  } else {
    size_t tot = (size_t)Nf;
    for (int ch = 0; ch < nch_all; ++ch) tot += tab[ch].lbs.size();
    keys.reserve(tot);
    for (int f = 0; f < Nf; ++f) keys.push_back((int64_t)f * Nf + f);
    for (int ch = 0; ch < nch_all; ++ch) for (const LB& b : tab[ch].lbs) keys.push_back(b.key);
    std::sort(keys.begin(), keys.end());
    keys.erase(std::unique(keys.begin(), keys.end()), keys.end());
  }
  const int32_t nbu = (int32_t)keys.size();
  H.blk_row.resize(nbu); H.blk_col.resize(nbu); H.diag_blk.assign(Nf, -1);
  for (int b = 0; b < nbu; ++b) {
    H.blk_row[b] = (int32_t)(keys[b] / std::max(Nf, 1)); H.blk_col[b] = (int32_t)(keys[b] % std::max(Nf, 1));
    if (H.blk_row[b] == H.blk_col[b]) H.diag_blk[H.blk_row[b]] = b;
  }

  lap("block pattern");
  // ---- flatten the owned chunks: offsets by one serial prefix pass, contents filled in parallel
  H.win_pchunk_off.assign(nwin + 1, 0);
  std::vector<int> own;                                   // owned chunks, in order
  for (int ch = 0; ch < nch_all; ++ch) if (cr[ch].owned) own.push_back(ch);
  const int nch = (int)own.size();
  H.pchunk_begin.resize(nch); H.pchunk_end.resize(nch); H.pchunk_win.resize(nch);
  H.slot_off.assign(nch + 1, 0); H.lb_off.assign(nch + 1, 0);
  std::vector<int64_t> pair_off((size_t)nch + 1, 0);
  {
    int w_next = 0;
    for (int i = 0; i < nch; ++i) {
      const ChunkRange& R = cr[own[i]];
      const ChunkTab& T = tab[own[i]];
      while (w_next <= R.win) H.win_pchunk_off[w_next++] = i;
      H.pchunk_begin[i] = R.p0; H.pchunk_end[i] = R.p1; H.pchunk_win[i] = R.win;
      H.slot_off[i + 1] = H.slot_off[i] + (int32_t)T.cams.size();
      H.lb_off[i + 1] = H.lb_off[i] + (int32_t)T.lbs.size();
      pair_off[i + 1] = pair_off[i] + (int64_t)T.pairs.size();
      if (R.is_long) {
        if (H.long_chunks.empty()) { H.long_zoff.assign(1, 0); H.long_lb_off.assign(1, 0); }
        H.long_chunks.push_back(i);
        H.long_zoff.push_back(H.long_zoff.back() + (int32_t)T.cams.size());
        H.long_lb_off.push_back(H.long_lb_off.back() + (int32_t)T.lbs.size());
      } else {
        H.max_slots = std::max(H.max_slots, (int32_t)T.cams.size());
        H.max_chunk_obs = std::max(H.max_chunk_obs, (int32_t)T.oslot.size());
        H.max_chunk_pts = std::max(H.max_chunk_pts, R.p1 - R.p0);
        H.max_chunk_pairs = std::max(H.max_chunk_pairs, (int32_t)std::min<size_t>(T.pairs.size(), 1u << 30));
      }
    }
    while (w_next <= nwin) H.win_pchunk_off[w_next++] = nch;
  }
  if (pair_off[nch] > 2000000000LL) return "too many block products for 32-bit offsets";
  const int32_t ncs_all = H.slot_off[nch], nlb_all = H.lb_off[nch];
  H.oslot.resize(Nobs); H.sobs.resize(Nobs);
  H.cs_cam.resize(ncs_all); H.cs_obs_off.assign((size_t)ncs_all + 1, 0);
  H.lb_sa.resize(nlb_all); H.lb_sb.resize(nlb_all); H.lb_blk.resize(nlb_all); H.lb_pair_off.assign((size_t)nlb_all + 1, 0);
  H.pairs.resize((size_t)pair_off[nch]);
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads)
  for (int i = 0; i < nch; ++i) {
    const ChunkRange& R = cr[own[i]];
    const ChunkTab& T = tab[own[i]];
    const int32_t e0 = H.pt_off[R.p0];
    const int n = (int)T.oslot.size();
    std::memcpy(H.oslot.data() + e0, T.oslot.data(), (size_t)n);
    std::memcpy(H.sobs.data() + e0, T.sobs.data(), (size_t)n * sizeof(uint16_t));
    const int32_t s0 = H.slot_off[i];
    for (size_t sl = 0; sl < T.cams.size(); ++sl) { H.cs_cam[s0 + sl] = T.cams[sl]; H.cs_obs_off[s0 + sl + 1] = e0 + T.slot_obs_off[sl + 1]; }
    const int32_t l0 = H.lb_off[i];
    int64_t q = pair_off[i];
    for (size_t l = 0; l < T.lbs.size(); ++l) {
      const LB& b = T.lbs[l];
      H.lb_sa[l0 + l] = b.sa; H.lb_sb[l0 + l] = b.sb;
      if (dense_map) H.lb_blk[l0 + l] = key_blk[(size_t)b.key];
      else if (!bit_words.empty()) {
        const size_t k = (size_t)b.key;
        H.lb_blk[l0 + l] = bit_rank[k >> 6] + (int32_t)__builtin_popcountll(bit_words[k >> 6] & (((uint64_t)1 << (k & 63)) - 1));
      } else H.lb_blk[l0 + l] = (int32_t)(std::lower_bound(keys.begin(), keys.end(), b.key) - keys.begin());
      q += T.lb_pair_cnt[l];
      H.lb_pair_off[l0 + l + 1] = (int32_t)q;
    }
    if (!T.pairs.empty()) std::memcpy(H.pairs.data() + pair_off[i], T.pairs.data(), T.pairs.size() * sizeof(uint16_t));
  }
  if (NobsAll > 4000000) { std::vector<ChunkTab>().swap(tab); std::vector<int32_t>().swap(gperm); }   // global problems: do not sit on hundreds of MB
  {
    const int nch = (int)H.pchunk_begin.size();
    H.cdesc.assign((size_t)nch * 16, 0);
    for (int ch = 0; ch < nch; ++ch) {
      int32_t* D = H.cdesc.data() + (size_t)ch * 16;
      const int32_t p0 = H.pchunk_begin[ch], p1 = H.pchunk_end[ch];
      D[0] = p0; D[1] = p1 - p0; D[2] = H.pt_off[p0]; D[3] = H.pt_off[p1] - H.pt_off[p0];
      D[4] = H.pchunk_win[ch]; D[5] = H.slot_off[ch]; D[6] = H.slot_off[ch + 1] - H.slot_off[ch]; D[7] = H.lb_off[ch];
      D[8] = H.lb_off[ch + 1] - H.lb_off[ch]; D[9] = H.lb_pair_off[H.lb_off[ch]];
      D[10] = H.lb_pair_off[H.lb_off[ch + 1]] - H.lb_pair_off[H.lb_off[ch]];
    }
    for (size_t li = 0; li < H.long_chunks.size(); ++li) H.cdesc[(size_t)H.long_chunks[li] * 16 + 11] = 1 + (int32_t)li;   // long-track chunk (1-based)
  }

  lap("flatten");
  // ---- static partial lists (ascending => fixed summation order)
  const int32_t nlb = (int32_t)H.lb_blk.size(), ncs = (int32_t)H.cs_cam.size();
  H.blk_part_off.assign(nbu + 1, 0);
  for (int i = 0; i < nlb; ++i) H.blk_part_off[H.lb_blk[i] + 1]++;
  for (int b = 0; b < nbu; ++b) H.blk_part_off[b + 1] += H.blk_part_off[b];
  H.blk_part.resize(nlb);
  { std::vector<int32_t> cur(H.blk_part_off.begin(), H.blk_part_off.end() - 1); for (int i = 0; i < nlb; ++i) H.blk_part[cur[H.lb_blk[i]]++] = i; }
  H.cam_part_off.assign(Nc + 1, 0);
  for (int i = 0; i < ncs; ++i) H.cam_part_off[H.cs_cam[i] + 1]++;
  for (int c = 0; c < Nc; ++c) H.cam_part_off[c + 1] += H.cam_part_off[c];
  H.cam_part.resize(ncs);
  { std::vector<int32_t> cur(H.cam_part_off.begin(), H.cam_part_off.end() - 1); for (int i = 0; i < ncs; ++i) H.cam_part[cur[H.cs_cam[i]]++] = i; }

  H.lb_pos.resize(nlb); H.cs_pos.resize(ncs);
  for (int i = 0; i < nlb; ++i) H.lb_pos[H.blk_part[i]] = i;
  for (int i = 0; i < ncs; ++i) H.cs_pos[H.cam_part[i]] = i;

  // ---- full block-CSR pattern (both triangles) for the mat-vec
  H.row_ptr.assign(Nf + 1, 0);
  for (int b = 0; b < nbu; ++b) {
    H.row_ptr[H.blk_row[b] + 1]++;
    if (H.blk_row[b] != H.blk_col[b]) H.row_ptr[H.blk_col[b] + 1]++;
  }
  for (int f = 0; f < Nf; ++f) H.row_ptr[f + 1] += H.row_ptr[f];
  const int32_t nbf = H.row_ptr[Nf];
  H.bsr_col.resize(nbf); H.bsr_blk.resize(nbf);
  {
    // lower part first (transposed upper blocks, sorted by col because upper blocks are sorted by
    // (row, col) and we visit them in that order), then the upper part of the row.
    std::vector<int32_t> cur(H.row_ptr.begin(), H.row_ptr.end() - 1);
    for (int b = 0; b < nbu; ++b)
      if (H.blk_row[b] != H.blk_col[b]) { const int32_t r = H.blk_col[b]; const int32_t at = cur[r]++; H.bsr_col[at] = H.blk_row[b]; H.bsr_blk[at] = ~b; }
    for (int b = 0; b < nbu; ++b) { const int32_t r = H.blk_row[b]; const int32_t at = cur[r]++; H.bsr_col[at] = H.blk_col[b]; H.bsr_blk[at] = b; }
  }
  lap("partial lists + block-CSR");
  return std::string();
}

}  // namespace sba
