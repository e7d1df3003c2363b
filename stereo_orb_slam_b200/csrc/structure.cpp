// structure.cpp — one-off host build of the index tables the kernels run on:
// point-sorted and camera-sorted observation views, CTA work chunks, the pair list and the block
// pattern of the reduced camera matrix.
//
// reference: the observation -> (camera, point) indexing comes from the pack loop
// (src/bundle_adjuster.cpp:57-110); the gauge (first camera constant, :113) decides which cameras
// are free.  The pattern rule is SURVEY §8 R10: block (i,j) of the reduced camera matrix is
// structurally non-zero iff some point is observed by both free cameras i and j.
#include <algorithm>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "sba_internal.h"

namespace sba {

namespace {
struct Ent {
  int32_t col, a, b;  // a,b: local point-order positions, or -1 when the point is not owned (pattern only)
};
}  // namespace

std::string build_structure(int32_t Nc, int32_t Np, int64_t NobsAll, const int32_t* obs_cam, const int32_t* obs_pt,
                            const uint8_t* cam_fixed, int32_t nwin, const int32_t* win_cam_off_in,
                            const int32_t* win_pt_off_in, int32_t pt_begin, int32_t pt_end, HostStructure* S) {
  if (Nc < 0 || Np < 0 || NobsAll < 0) return "negative size";
  if (NobsAll > 2000000000LL) return "too many observations for 32-bit positions";
  if (nwin < 1) nwin = 1;
  if (pt_begin < 0 || pt_end > Np || pt_begin > pt_end) return "bad point shard";
  HostStructure& H = *S;
  H = HostStructure();
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
  std::vector<int32_t> pt_win(Np);
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
  std::vector<int32_t> gpt_off(Np + 1, 0);
  for (int64_t k = 0; k < NobsAll; ++k) {
    const int32_t c = obs_cam[k], p = obs_pt[k];
    if (c < 0 || c >= Nc) return "obs_cam out of range";
    if (p < 0 || p >= Np) return "obs_pt out of range";
    if (H.cam_win[c] != pt_win[p]) return "observation links a camera and a point of different windows";
    gpt_off[p + 1]++;
  }
  for (int p = 0; p < Np; ++p) gpt_off[p + 1] += gpt_off[p];
  std::vector<int32_t> gperm(NobsAll);
  {
    std::vector<int32_t> cur(gpt_off.begin(), gpt_off.end() - 1);
    for (int64_t k = 0; k < NobsAll; ++k) gperm[cur[obs_pt[k]]++] = (int32_t)k;
  }
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
#pragma omp parallel for schedule(static)
  for (int64_t e = 0; e < Nobs; ++e) {
    const int32_t k = gperm[base + e];
    H.perm_p[e] = k; H.pcam[e] = obs_cam[k]; H.ppt[e] = obs_pt[k];
  }
  // ---- camera order over local positions (inside a camera: ascending point-order position)
  H.cam_off.assign(Nc + 1, 0);
  for (int64_t e = 0; e < Nobs; ++e) H.cam_off[H.pcam[e] + 1]++;
  for (int c = 0; c < Nc; ++c) H.cam_off[c + 1] += H.cam_off[c];
  H.cpos.resize(Nobs); H.cpt.resize(Nobs);
  {
    std::vector<int32_t> cur(H.cam_off.begin(), H.cam_off.end() - 1);
    for (int64_t e = 0; e < Nobs; ++e) { const int32_t i = cur[H.pcam[e]]++; H.cpos[i] = (int32_t)e; H.cpt[i] = H.ppt[e]; }
  }
  // ---- CTA chunks
  H.win_pchunk_off.assign(nwin + 1, 0);
  for (int w = 0; w < nwin; ++w) {
    H.win_pchunk_off[w] = (int32_t)H.pchunk_begin.size();
    const int lo = std::max(H.win_pt_off[w], pt_begin), hi = std::min(H.win_pt_off[w + 1], pt_end);
    for (int b = lo; b < hi; b += kPointChunk) {
      H.pchunk_begin.push_back(b); H.pchunk_end.push_back(std::min(b + kPointChunk, hi)); H.pchunk_win.push_back(w);
    }
  }
  H.win_pchunk_off[nwin] = (int32_t)H.pchunk_begin.size();
  H.cam_chunk_off.assign(Nc + 1, 0);
  for (int c = 0; c < Nc; ++c) {
    H.cam_chunk_off[c] = (int32_t)H.cchunk_cam.size();
    if (H.fc[c] < 0) continue;
    for (int b = H.cam_off[c]; b < H.cam_off[c + 1]; b += kCamChunk) {
      H.cchunk_cam.push_back(c); H.cchunk_begin.push_back(b); H.cchunk_end.push_back(std::min(b + kCamChunk, H.cam_off[c + 1]));
    }
  }
  H.cam_chunk_off[Nc] = (int32_t)H.cchunk_cam.size();

  // ---- pair entries, bucketed by block row (free camera), then sorted by column inside a row
  // ordered pairs (a,b) of one point's observations with fc[a] < fc[b], or fc[a] == fc[b] (the
  // diagonal takes every ordered pair, incl. a == b and both orders of duplicated (cam,pt) pairs).
  std::vector<int64_t> row_cnt(Nf + 1, 0);
  for (int p = 0; p < Np; ++p) {
    const int32_t g0 = gpt_off[p], g1 = gpt_off[p + 1];
    for (int32_t ga = g0; ga < g1; ++ga) {
      const int32_t fa = H.fc[obs_cam[gperm[ga]]];
      if (fa < 0) continue;
      int64_t n = 0;
      for (int32_t gb = g0; gb < g1; ++gb) { const int32_t fb = H.fc[obs_cam[gperm[gb]]]; if (fb >= fa) ++n; }
      row_cnt[fa + 1] += n;
    }
  }
  for (int f = 0; f < Nf; ++f) row_cnt[f + 1] += row_cnt[f];
  const int64_t Etot = row_cnt[Nf];
  std::vector<Ent> ents((size_t)Etot);
  {
    std::vector<int64_t> cur(row_cnt.begin(), row_cnt.end() - 1);
    for (int p = 0; p < Np; ++p) {
      const int32_t g0 = gpt_off[p], g1 = gpt_off[p + 1];
      const bool owned = (p >= pt_begin && p < pt_end);
      for (int32_t ga = g0; ga < g1; ++ga) {
        const int32_t fa = H.fc[obs_cam[gperm[ga]]];
        if (fa < 0) continue;
        for (int32_t gb = g0; gb < g1; ++gb) {
          const int32_t fb = H.fc[obs_cam[gperm[gb]]];
          if (fb < fa) continue;
          Ent& e = ents[(size_t)cur[fa]++];
          e.col = fb; e.a = owned ? ga - base : -1; e.b = owned ? gb - base : -1;
        }
      }
    }
  }
#pragma omp parallel for schedule(dynamic, 16)
  for (int f = 0; f < Nf; ++f)
    std::stable_sort(ents.begin() + row_cnt[f], ents.begin() + row_cnt[f + 1],
                     [](const Ent& x, const Ent& y) { return x.col < y.col; });
  // blocks (the diagonal block of every free camera always exists, first in its row)
  H.diag_blk.assign(Nf, -1);
  H.blk_ent_off.push_back(0);
  H.ent_a.reserve((size_t)Etot); H.ent_b.reserve((size_t)Etot);
  for (int f = 0; f < Nf; ++f) {
    int64_t i = row_cnt[f];
    const int64_t end = row_cnt[f + 1];
    if (i == end || ents[(size_t)i].col != f) {   // camera without (owned or unowned) observations
      H.diag_blk[f] = (int32_t)H.blk_row.size();
      H.blk_row.push_back(f); H.blk_col.push_back(f); H.blk_ent_off.push_back((int64_t)H.ent_a.size());
    }
    while (i < end) {
      const int32_t col = ents[(size_t)i].col;
      if (col == f) H.diag_blk[f] = (int32_t)H.blk_row.size();
      H.blk_row.push_back(f); H.blk_col.push_back(col);
      for (; i < end && ents[(size_t)i].col == col; ++i)
        if (ents[(size_t)i].a >= 0) { H.ent_a.push_back(ents[(size_t)i].a); H.ent_b.push_back(ents[(size_t)i].b); }
      H.blk_ent_off.push_back((int64_t)H.ent_a.size());
    }
  }
  std::vector<Ent>().swap(ents);
  const int32_t nbu = (int32_t)H.blk_row.size();
  H.blk_chunk_off.assign(nbu + 1, 0);
  for (int b = 0; b < nbu; ++b) {
    H.blk_chunk_off[b] = (int32_t)H.schunk_blk.size();
    for (int64_t s = H.blk_ent_off[b]; s < H.blk_ent_off[b + 1]; s += kPairChunk) {
      H.schunk_blk.push_back(b); H.schunk_begin.push_back(s); H.schunk_end.push_back(std::min<int64_t>(s + kPairChunk, H.blk_ent_off[b + 1]));
    }
  }
  H.blk_chunk_off[nbu] = (int32_t)H.schunk_blk.size();

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
  return std::string();
}

}  // namespace sba
