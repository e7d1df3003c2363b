// sba_internal.h — host-side structure tables of the stereo bundle adjuster (not part of the ABI).
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace sba {

constexpr int kThreads = 128;        // CTA size of the thread-per-point kernels
constexpr int kPointChunk = 128;     // max points per chunk
constexpr int kObsChunk = 240;       // max observations per chunk (= CTA size of the thread-per-observation kernels;
                                     // chunk-local observation indices fit a byte)
constexpr int kMaxSlots = 255;       // distinct cameras per chunk (slot index is a byte)
constexpr int kMaxLongObs = 4096;    // observations of one point in the long-track path (its cameras form a dense k x k block of S)

// Everything that depends only on (obs_cam, obs_pt, cam_fixed, windows): built once on the host
// (reference: the index build of src/bundle_adjuster.cpp:57-110 fixes obs->cam/pt; the rest is new).
//
// Unit of work = CHUNK: a run of consecutive points (<= kPointChunk points, <= kObsChunk observations, one
// window, one rank).  In the reference's first-appearance point numbering neighbouring points are seen by the
// same few cameras, so a chunk touches a handful of cameras ("slots") and a handful of blocks of the reduced
// camera matrix ("local blocks"); every cross-observation reduction (per point, per camera, per block) is first
// done inside the chunk's CTA and then summed over chunks through static lists, in a fixed order, without atomics.
struct HostStructure {
  int32_t Nc = 0, Np = 0, Nf = 0, nwin = 0;
  int64_t Nobs = 0;          // observations held by this rank (all of them on 1 GPU)
  int32_t pt_begin = 0, pt_end = 0;   // owned point range (multi-GPU shard); [0,Np) on 1 GPU
  int32_t obs_cap = kObsChunk;

  std::vector<int32_t> fc;            // [Nc] free-camera index or -1
  std::vector<int32_t> free2cam;      // [Nf]
  std::vector<int32_t> cam_win;       // [Nc]
  std::vector<int32_t> fcam_win;      // [Nf]
  std::vector<int32_t> win_cam_off, win_pt_off, win_fc_off;  // [nwin+1]

  // point order (positions e = 0..Nobs-1, grouped by point, stable in input order)
  std::vector<int32_t> pt_off;        // [Np+1]   (local obs offsets; empty segments for unowned points)
  std::vector<int32_t> perm_p;        // [Nobs]   input observation index at position e
  std::vector<int32_t> pcam;          // [Nobs]   camera at position e
  std::vector<int32_t> ppt;           // [Nobs]   point at position e
  // chunks (owned points only)
  std::vector<int32_t> pchunk_begin, pchunk_end, pchunk_win;   // point range and window of chunk c
  // one 64-byte record per chunk, everything the per-chunk kernels need to start (one load instead of a chain):
  // [0] p0 [1] np [2] e0 [3] n  [4] win [5] cs0 [6] ns [7] lb0  [8] nlb [9] pq0 [10] npq [11..15] 0
  std::vector<int32_t> cdesc;                                   // [nchunk * 16]
  std::vector<int32_t> win_pchunk_off;                          // [nwin+1]
  // camera slots: chunk c owns slots cs in [slot_off[c], slot_off[c+1])
  std::vector<int32_t> slot_off;                                // [nchunk+1]
  std::vector<int32_t> cs_cam;                                  // [ncs] camera of chunk-slot cs
  std::vector<int32_t> cs_obs_off;                              // [ncs+1] range of sobs (absolute positions)
  std::vector<uint8_t> oslot;                                   // [Nobs] slot (local to its chunk) of position e
  std::vector<uint16_t> sobs;                                   // [Nobs] chunk-local obs index, grouped by slot
  // local blocks: chunk c owns lb in [lb_off[c], lb_off[c+1]); slots sa, sb with fc[sa] <= fc[sb], both free
  std::vector<int32_t> lb_off;                                  // [nchunk+1]
  std::vector<uint8_t> lb_sa, lb_sb;                            // [nlb]
  std::vector<int32_t> lb_blk;                                  // [nlb] upper block id
  // products of local block lb: pairs [lb_pair_off[lb], lb_pair_off[lb+1]) of chunk-local observations (a | b << 8)
  // of one point with slot(a) = sa, slot(b) = sb, points ascending; its partial is sum_pairs W_a Vinv W_b^T
  std::vector<int32_t> lb_pair_off;                             // [nlb+1]
  std::vector<uint16_t> pairs;                                  // [npairs]
  int32_t max_slots = 0, max_chunk_obs = 0, max_chunk_pts = 0, max_chunk_pairs = 0;  // over owned REGULAR chunks (shared-memory sizing)
  // long tracks: a point with more than kMaxSlots observations (loop-closure re-targeting, src/camera_frame.h:92-109)
  // is a chunk of its own, handled by the k_*_long kernels: one slot per observation, one local block per ordered
  // observation pair (its `pairs` entry is the two chunk-local observation indices, 16 bits each)
  std::vector<int32_t> long_chunks;                             // owned chunk ids
  std::vector<int32_t> long_zoff;                               // [nlong+1] offsets (observations) into the Z scratch
  std::vector<int32_t> long_lb_off;                             // [nlong+1] prefix of their local-block counts
  // build workspace kept between calls on one handle (per-chunk tables, sort buffers): BundleAdjuster::Optimize
  // re-structures on every call, and fresh vectors cost more in page faults and allocator calls than the build itself
  std::shared_ptr<void> work;
  // reduced camera matrix: upper blocks (row <= col, free indices), sorted (row, col); partial lists
  std::vector<int32_t> blk_row, blk_col;                        // [nbu]
  std::vector<int32_t> blk_part_off;                            // [nbu+1]
  std::vector<int32_t> blk_part;                                // [nlb] lb indices, ascending
  std::vector<int32_t> cam_part_off;                            // [Nc+1]
  std::vector<int32_t> cam_part;                                // [ncs] chunk-slot indices, ascending
  // where the kernels store their partials: position in the block's / camera's list, so that the partials of one
  // block / camera are CONTIGUOUS rows for the assemble kernels
  std::vector<int32_t> lb_pos;                                  // [nlb] inverse of blk_part
  std::vector<int32_t> cs_pos;                                  // [ncs] inverse of cam_part
  std::vector<int32_t> diag_blk;                                // [Nf] upper-block index of (i,i)
  // full pattern in block-CSR for the PCG mat-vec: entry -> (col, upper block, transposed?)
  std::vector<int32_t> row_ptr;                                 // [Nf+1]
  std::vector<int32_t> bsr_col, bsr_blk;                        // [nbf]; bsr_blk < 0 => ~blk, transposed
};

// returns empty string on success, else an error message
std::string build_structure(int32_t Nc, int32_t Np, int64_t Nobs, const int32_t* obs_cam, const int32_t* obs_pt,
                            const uint8_t* cam_fixed, int32_t nwin, const int32_t* win_cam_off,
                            const int32_t* win_pt_off, int32_t pt_begin, int32_t pt_end, HostStructure* S);

}  // namespace sba
