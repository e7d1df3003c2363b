// sparse_chol.h — symbolic phase of the exact block-sparse Cholesky of the reduced camera matrix (host side).
//
// The reference factors the reduced camera system exactly (Ceres SPARSE_SCHUR + CHOLMOD,
// src/bundle_adjuster.cpp:26-29).  Windows of <= 64 free cameras use the dense block Cholesky kernels; anything
// larger (the global BA calls of src/slam.cpp:156 and src/pose_graph_optimizer.cpp:95: thousands of cameras, a
// banded pattern plus loop-closure blocks) goes through this: a nested-dissection ordering of the 6x6-block
// graph, the elimination tree and the block pattern of L, level sets of the tree (columns of one level are
// independent), and for every block of L the static list of products L_ik L_jk^T it receives (left-looking:
// one writer per block, fixed order, no atomics).  The numeric phase is k_spchol (sparse_chol.cuh).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace sba {

struct SparseCholSymbolic {
  int32_t n = 0;                       // block columns (= free cameras)
  std::vector<int32_t> perm;           // [n] perm[new] = old (free-camera index)
  std::vector<int32_t> iperm;          // [n] iperm[old] = new
  std::vector<int32_t> parent;         // [n] elimination tree (new indices), -1 = root
  std::vector<int32_t> col_ptr;        // [n+1] blocks of column j: [col_ptr[j], col_ptr[j+1]), the diagonal first, rows ascending
  std::vector<int32_t> row_idx;        // [nnzL] block row (new index)
  std::vector<int32_t> init_src;       // [nnzL] upper block of S that initialises the block: b >= 0 as stored, ~b transposed,
                                       //        INT32_MIN: fill-in (starts at zero)
  std::vector<int64_t> upd_ptr;        // [nnzL+1] products received by each block
  std::vector<int32_t> upd_a, upd_b;   // [nupd] block ids of L_ik and L_jk, k ascending
  std::vector<int64_t> upd_ext_end;    // [nnzL] end of the products whose source column lies outside the block's chain (they come first)
  std::vector<int32_t> rowl_ext_end;   // [n] idem for the row lists
  std::vector<int32_t> rowl_ptr;       // [n+1] row lists (forward substitution): blocks (j, k), k < j, k ascending
  std::vector<int32_t> rowl_blk, rowl_col;   // [nnzL - n]
  std::vector<int32_t> lvl_ptr;        // [nlev+1] level sets of the CHAINS of the elimination tree (leaves = level 0)
  std::vector<int32_t> lvl_cols;       // [nchain] first column of each chain, sorted by (level, column)
  std::vector<int32_t> lvl_len;        // [nchain] its length: columns lvl_cols[c] .. + lvl_len[c] - 1, each the only child of the next
  int32_t nlev = 0;
  int64_t nchain() const { return (int64_t)lvl_cols.size(); }
  int64_t nnzL() const { return (int64_t)row_idx.size(); }
  int64_t nupd() const { return (int64_t)upd_a.size(); }
};

// Upper block pattern (blk_row[b] <= blk_col[b], every diagonal block present) of an n x n block matrix.
// max_updates: give up (returns an error string) when the update lists would exceed this many products — the caller
// then falls back to PCG.  Returns an empty string on success.
std::string sparse_chol_symbolic(int32_t n, const std::vector<int32_t>& blk_row, const std::vector<int32_t>& blk_col,
                                 int64_t max_updates, SparseCholSymbolic* out);

// nested-dissection ordering alone (exposed for the tests): order[new] = old
void nested_dissection_order(int32_t n, const std::vector<int32_t>& adj_ptr, const std::vector<int32_t>& adj,
                             std::vector<int32_t>* order);

}  // namespace sba
