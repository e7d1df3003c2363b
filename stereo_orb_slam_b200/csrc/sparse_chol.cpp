// sparse_chol.cpp — symbolic phase of the exact block-sparse Cholesky (see sparse_chol.h).
//
// reference: Ceres hands the reduced camera matrix to CHOLMOD (src/bundle_adjuster.cpp:26-29), which orders
// (AMD / nested dissection), builds the elimination tree and a supernodal pattern once per structure and then
// refactors numerically every LM iteration.  Same split here, at 6x6-block granularity; the numeric phase runs on
// the GPU.
#include "sparse_chol.h"

#include <algorithm>
#include <climits>
#include <numeric>

namespace sba {

namespace {

constexpr int kLeaf = 12;     // regions of at most this many block columns are not dissected further

struct NdWork {
  const std::vector<int32_t>& ap;
  const std::vector<int32_t>& adj;
  std::vector<int32_t> region;   // region id of every vertex (the id of the innermost nd() call that owns it)
  std::vector<int32_t> dist;     // BFS level inside the current region
  std::vector<int32_t> queue;
  int32_t next_region = 1;
  std::vector<int32_t>* out;
};

// BFS inside region `rid` from `start`; fills W.queue (visit order) and W.dist; returns the number of levels
int bfs(NdWork& W, int32_t rid, int32_t start, size_t* visited) {
  W.queue.clear();
  W.queue.push_back(start);
  W.dist[start] = 0;
  // dist doubles as the visited mark of this BFS: it is reset to -1 for the region's vertices by the caller
  size_t head = 0;
  int last = 0;
  while (head < W.queue.size()) {
    const int32_t v = W.queue[head++];
    last = W.dist[v];
    for (int32_t e = W.ap[v]; e < W.ap[v + 1]; ++e) {
      const int32_t u = W.adj[e];
      if (W.region[u] != rid || W.dist[u] >= 0) continue;
      W.dist[u] = W.dist[v] + 1;
      W.queue.push_back(u);
    }
  }
  *visited = W.queue.size();
  return last + 1;
}

void nd(NdWork& W, std::vector<int32_t>& verts) {
  if (verts.empty()) return;
  if ((int)verts.size() <= kLeaf) {
    std::sort(verts.begin(), verts.end());
    W.out->insert(W.out->end(), verts.begin(), verts.end());
    return;
  }
  const int32_t rid = W.next_region++;
  for (int32_t v : verts) { W.region[v] = rid; W.dist[v] = -1; }
  // connected components of the region: each is ordered on its own (no separator needed between them)
  {
    size_t got = 0;
    bfs(W, rid, verts[0], &got);
    if (got < verts.size()) {
      std::vector<std::vector<int32_t>> comps;
      comps.emplace_back(W.queue.begin(), W.queue.end());
      for (int32_t v : verts) {
        if (W.dist[v] >= 0) continue;
        size_t g2 = 0;
        bfs(W, rid, v, &g2);
        comps.emplace_back(W.queue.begin(), W.queue.end());
      }
      for (auto& c : comps) nd(W, c);
      return;
    }
  }
  // pseudo-peripheral start: repeat BFS from the farthest vertex while the eccentricity grows
  int32_t start = verts[0];
  int nlev = 0;
  for (int it = 0; it < 4; ++it) {
    for (int32_t v : verts) W.dist[v] = -1;
    size_t got = 0;
    const int l = bfs(W, rid, start, &got);
    if (l <= nlev) { nlev = l; break; }
    nlev = l;
    // farthest vertex of smallest degree
    int32_t best = W.queue.back(); int bestdeg = INT_MAX;
    for (size_t i = W.queue.size(); i-- > 0;) {
      const int32_t v = W.queue[i];
      if (W.dist[v] != nlev - 1) break;
      const int deg = W.ap[v + 1] - W.ap[v];
      if (deg < bestdeg) { bestdeg = deg; best = v; }
    }
    if (best == start) break;
    start = best;
  }
  // final level structure from `start`
  for (int32_t v : verts) W.dist[v] = -1;
  { size_t got = 0; nlev = bfs(W, rid, start, &got); }
  if (nlev < 3) {     // (near-)clique: no vertex separator worth taking
    std::sort(verts.begin(), verts.end());
    W.out->insert(W.out->end(), verts.begin(), verts.end());
    return;
  }
  std::vector<int64_t> lsize(nlev, 0);
  for (int32_t v : verts) lsize[W.dist[v]]++;
  std::vector<int64_t> below(nlev + 1, 0);
  for (int l = 0; l < nlev; ++l) below[l + 1] = below[l] + lsize[l];
  const int64_t tot = (int64_t)verts.size();
  int s = 1; int64_t best_cost = INT64_MAX;
  for (int l = 1; l <= nlev - 2; ++l) {
    const int64_t a = below[l], b = tot - below[l + 1];
    const int64_t cost = std::llabs(a - b) + 2 * lsize[l];
    if (cost < best_cost) { best_cost = cost; s = l; }
  }
  std::vector<int32_t> A, B, S;
  for (int32_t v : verts) {
    const int l = W.dist[v];
    if (l < s) A.push_back(v); else if (l > s) B.push_back(v); else S.push_back(v);
  }
  std::vector<int32_t>().swap(verts);
  nd(W, A);
  nd(W, B);
  std::sort(S.begin(), S.end());
  W.out->insert(W.out->end(), S.begin(), S.end());
}

}  // namespace

void nested_dissection_order(int32_t n, const std::vector<int32_t>& adj_ptr, const std::vector<int32_t>& adj,
                             std::vector<int32_t>* order) {
  order->clear();
  order->reserve(n);
  NdWork W{adj_ptr, adj, std::vector<int32_t>(n, 0), std::vector<int32_t>(n, -1), {}, 1, order};
  std::vector<int32_t> all(n);
  std::iota(all.begin(), all.end(), 0);
  nd(W, all);
}

std::string sparse_chol_symbolic(int32_t n, const std::vector<int32_t>& blk_row, const std::vector<int32_t>& blk_col,
                                 int64_t max_updates, SparseCholSymbolic* out) {
  SparseCholSymbolic& C = *out;
  C = SparseCholSymbolic();
  C.n = n;
  if (n <= 0) return std::string();
  const int64_t nbu = (int64_t)blk_row.size();
  // ---- symmetric block graph (no self loops)
  std::vector<int32_t> ap(n + 1, 0);
  for (int64_t b = 0; b < nbu; ++b) {
    const int32_t r = blk_row[b], c = blk_col[b];
    if (r < 0 || c < r || c >= n) return "sparse_chol_symbolic: bad block pattern";
    if (r != c) { ap[r + 1]++; ap[c + 1]++; }
  }
  for (int i = 0; i < n; ++i) ap[i + 1] += ap[i];
  std::vector<int32_t> adj(ap[n]);
  {
    std::vector<int32_t> cur(ap.begin(), ap.end() - 1);
    for (int64_t b = 0; b < nbu; ++b) {
      const int32_t r = blk_row[b], c = blk_col[b];
      if (r != c) { adj[cur[r]++] = c; adj[cur[c]++] = r; }
    }
  }
  nested_dissection_order(n, ap, adj, &C.perm);
  if ((int)C.perm.size() != n) return "sparse_chol_symbolic: ordering lost vertices";
  C.iperm.assign(n, -1);
  for (int i = 0; i < n; ++i) C.iperm[C.perm[i]] = i;

  // ---- elimination tree (Liu) on the permuted pattern: for every column j the neighbours i < j
  std::vector<std::vector<int32_t>> low(n), up(n);     // new indices
  for (int v = 0; v < n; ++v) {
    const int32_t j = C.iperm[v];
    for (int32_t e = ap[v]; e < ap[v + 1]; ++e) {
      const int32_t i = C.iperm[adj[e]];
      if (i < j) low[j].push_back(i); else up[j].push_back(i);
    }
  }
  C.parent.assign(n, -1);
  {
    std::vector<int32_t> anc(n, -1);
    for (int j = 0; j < n; ++j)
      for (int32_t i : low[j]) {
        int32_t r = i;
        while (anc[r] != -1 && anc[r] != j) { const int32_t nx = anc[r]; anc[r] = j; r = nx; }
        if (anc[r] == -1) { anc[r] = j; C.parent[r] = j; }
      }
  }
  // ---- column structures: struct(j) = {i > j : A_ij != 0}  U  union over children c of struct(c) \ {j}
  std::vector<std::vector<int32_t>> st(n);
  {
    std::vector<std::vector<int32_t>> kids(n);
    for (int j = 0; j < n; ++j) if (C.parent[j] >= 0) kids[C.parent[j]].push_back(j);
    std::vector<int32_t> mark(n, -1);
    for (int j = 0; j < n; ++j) {
      std::vector<int32_t>& s = st[j];
      mark[j] = j;
      for (int32_t i : up[j]) if (mark[i] != j) { mark[i] = j; s.push_back(i); }
      for (int32_t c : kids[j])
        for (int32_t i : st[c]) if (mark[i] != j) { mark[i] = j; s.push_back(i); }
      std::sort(s.begin(), s.end());
    }
  }
  C.col_ptr.assign(n + 1, 0);
  for (int j = 0; j < n; ++j) C.col_ptr[j + 1] = C.col_ptr[j] + 1 + (int32_t)st[j].size();
  const int64_t nnzL = C.col_ptr[n];
  if (nnzL > (int64_t)INT32_MAX / 40) return "sparse_chol_symbolic: factor too large";
  C.row_idx.resize(nnzL);
  for (int j = 0; j < n; ++j) {
    int32_t at = C.col_ptr[j];
    C.row_idx[at++] = j;
    for (int32_t i : st[j]) C.row_idx[at++] = i;
  }
  std::vector<std::vector<int32_t>>().swap(st);
  auto find_block = [&](int32_t i, int32_t j) -> int32_t {
    const int32_t* b = C.row_idx.data() + C.col_ptr[j];
    const int32_t* e = C.row_idx.data() + C.col_ptr[j + 1];
    const int32_t* p = std::lower_bound(b, e, i);
    return (p != e && *p == i) ? (int32_t)(p - C.row_idx.data()) : -1;
  };
  // ---- where the blocks of S land
  C.init_src.assign(nnzL, INT32_MIN);
  for (int64_t b = 0; b < nbu; ++b) {
    const int32_t pr = C.iperm[blk_row[b]], pc = C.iperm[blk_col[b]];
    const int32_t i = std::max(pr, pc), j = std::min(pr, pc);
    const int32_t at = find_block(i, j);
    if (at < 0) return "sparse_chol_symbolic: pattern block missing from the factor";
    C.init_src[at] = (pr >= pc) ? (int32_t)b : ~(int32_t)b;
  }
  // ---- update lists (left-looking): block (i, j) receives L_ik L_jk^T for every k with i, j in struct(k)
  C.upd_ptr.assign(nnzL + 1, 0);
  for (int pass = 0; pass < 2; ++pass) {
    std::vector<int64_t> cur;
    if (pass == 1) {
      for (int64_t b = 0; b < nnzL; ++b) C.upd_ptr[b + 1] += C.upd_ptr[b];
      if (C.upd_ptr[nnzL] > max_updates) return "sparse_chol_symbolic: update lists exceed the cap (pattern too dense)";
      C.upd_a.resize(C.upd_ptr[nnzL]); C.upd_b.resize(C.upd_ptr[nnzL]);
      cur.assign(C.upd_ptr.begin(), C.upd_ptr.end() - 1);
    }
    for (int k = 0; k < n; ++k) {
      const int32_t c0 = C.col_ptr[k] + 1, c1 = C.col_ptr[k + 1];
      for (int32_t y = c0; y < c1; ++y) {
        const int32_t j = C.row_idx[y];
        // rows of struct(k) at or below j all appear in column j (fill-path theorem): merge walk
        int32_t t = C.col_ptr[j];
        const int32_t t1 = C.col_ptr[j + 1];
        for (int32_t x = y; x < c1; ++x) {
          const int32_t i = C.row_idx[x];
          while (t < t1 && C.row_idx[t] < i) ++t;
          if (t >= t1 || C.row_idx[t] != i) return "sparse_chol_symbolic: inconsistent fill pattern";
          if (pass == 0) C.upd_ptr[t + 1]++;
          else { const int64_t at = cur[t]++; C.upd_a[at] = x; C.upd_b[at] = y; }
        }
      }
    }
  }
  // ---- row lists for the forward substitution
  C.rowl_ptr.assign(n + 1, 0);
  for (int k = 0; k < n; ++k)
    for (int32_t y = C.col_ptr[k] + 1; y < C.col_ptr[k + 1]; ++y) C.rowl_ptr[C.row_idx[y] + 1]++;
  for (int j = 0; j < n; ++j) C.rowl_ptr[j + 1] += C.rowl_ptr[j];
  C.rowl_blk.resize(C.rowl_ptr[n]); C.rowl_col.resize(C.rowl_ptr[n]);
  {
    std::vector<int32_t> cur(C.rowl_ptr.begin(), C.rowl_ptr.end() - 1);
    for (int k = 0; k < n; ++k)
      for (int32_t y = C.col_ptr[k] + 1; y < C.col_ptr[k + 1]; ++y) {
        const int32_t at = cur[C.row_idx[y]]++;
        C.rowl_blk[at] = y; C.rowl_col[at] = k;
      }
  }
  // ---- chains and their level sets.  A chain is a run of consecutive columns j, j+1, ... in which every column is the
  // ONLY child of the next one (the columns of a separator, and the whole top of the tree): its columns depend on each
  // other in sequence, so one CTA factors the chain without a grid barrier in between.  Only the first column of a
  // chain has children outside it; chains of one level are independent.
  std::vector<int32_t> nkids(n, 0);
  for (int j = 0; j < n; ++j) if (C.parent[j] >= 0) nkids[C.parent[j]]++;
  std::vector<int32_t> chain_of(n, -1), chain_start, chain_len, chain_level;
  for (int j = 0; j < n; ++j) {
    const bool cont = j > 0 && C.parent[j - 1] == j && nkids[j] == 1;
    if (cont) { chain_of[j] = chain_of[j - 1]; chain_len[chain_of[j]]++; }
    else { chain_of[j] = (int32_t)chain_start.size(); chain_start.push_back(j); chain_len.push_back(1); chain_level.push_back(0); }
  }
  // children of a chain's first column belong to earlier chains (smaller indices): one ascending pass
  for (int j = 0; j < n; ++j) {
    const int32_t p = C.parent[j];
    if (p >= 0 && chain_of[p] != chain_of[j])
      chain_level[chain_of[p]] = std::max(chain_level[chain_of[p]], chain_level[chain_of[j]] + 1);
  }
  // product lists and row lists are sorted by source column k: the sources OUTSIDE the block's own chain come first
  // (k < first column of the chain).  The kernel gathers those from L2 for a whole chain at once and runs the
  // in-chain part out of shared memory.
  {
    std::vector<int32_t> col_of_blk(nnzL);
    for (int j = 0; j < n; ++j) for (int32_t b = C.col_ptr[j]; b < C.col_ptr[j + 1]; ++b) col_of_blk[b] = j;
    C.upd_ext_end.resize(nnzL);
    for (int j = 0; j < n; ++j) {
      const int32_t j0 = chain_start[chain_of[j]];
      for (int32_t b = C.col_ptr[j]; b < C.col_ptr[j + 1]; ++b) {
        int64_t u = C.upd_ptr[b];
        while (u < C.upd_ptr[b + 1] && col_of_blk[C.upd_a[u]] < j0) ++u;
        C.upd_ext_end[b] = u;
      }
    }
    C.rowl_ext_end.resize(n);
    for (int j = 0; j < n; ++j) {
      const int32_t j0 = chain_start[chain_of[j]];
      int32_t e = C.rowl_ptr[j];
      while (e < C.rowl_ptr[j + 1] && C.rowl_col[e] < j0) ++e;
      C.rowl_ext_end[j] = e;
    }
  }
  const int nchain = (int)chain_start.size();
  int nlev = 1;
  for (int c = 0; c < nchain; ++c) nlev = std::max(nlev, chain_level[c] + 1);
  C.nlev = nlev;
  C.lvl_ptr.assign(nlev + 1, 0);
  for (int c = 0; c < nchain; ++c) C.lvl_ptr[chain_level[c] + 1]++;
  for (int l = 0; l < nlev; ++l) C.lvl_ptr[l + 1] += C.lvl_ptr[l];
  C.lvl_cols.resize(nchain); C.lvl_len.resize(nchain);
  {
    std::vector<int32_t> cur(C.lvl_ptr.begin(), C.lvl_ptr.end() - 1);
    for (int c = 0; c < nchain; ++c) { const int32_t at = cur[chain_level[c]]++; C.lvl_cols[at] = chain_start[c]; C.lvl_len[at] = chain_len[c]; }
  }
  return std::string();
}

}  // namespace sba
