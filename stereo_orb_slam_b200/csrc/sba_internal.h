// sba_internal.h — host-side structure tables of the stereo bundle adjuster (not part of the ABI).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace sba {

constexpr int kThreads = 128;        // CTA size of every chunked kernel
constexpr int kPointChunk = 128;     // points per CTA (thread per point)
constexpr int kCamChunk = 1024;      // camera-ordered observations per CTA
constexpr int kPairChunk = 1024;     // pair entries per CTA

// Everything that depends only on (obs_cam, obs_pt, cam_fixed, windows): built once on the host
// (reference: the index build of src/bundle_adjuster.cpp:57-110 fixes obs->cam/pt; the rest is new).
struct HostStructure {
  int32_t Nc = 0, Np = 0, Nf = 0, nwin = 0;
  int64_t Nobs = 0;          // observations held by this rank (all of them on 1 GPU)
  int32_t pt_begin = 0, pt_end = 0;   // owned point range (multi-GPU shard); [0,Np) on 1 GPU

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
  // camera order (positions i, grouped by camera, stable)
  std::vector<int32_t> cam_off;       // [Nc+1]
  std::vector<int32_t> cpos;          // [Nobs]   point-order position of camera-order position i
  std::vector<int32_t> cpt;           // [Nobs]   point at camera-order position i
  // CTA work tables
  std::vector<int32_t> pchunk_begin, pchunk_end, pchunk_win;   // point chunks (never span windows)
  std::vector<int32_t> win_pchunk_off;                          // [nwin+1]
  std::vector<int32_t> cchunk_cam, cchunk_begin, cchunk_end;    // camera chunks (free cameras only)
  std::vector<int32_t> cam_chunk_off;                           // [Nc+1]
  // reduced camera matrix: upper blocks (row <= col, free indices), sorted (row, col)
  std::vector<int32_t> blk_row, blk_col;                        // [nbu]
  std::vector<int64_t> blk_ent_off;                             // [nbu+1]
  std::vector<int32_t> ent_a, ent_b;                            // [E] point-order positions
  std::vector<int32_t> schunk_blk;                              // [nsch]
  std::vector<int64_t> schunk_begin, schunk_end;
  std::vector<int32_t> blk_chunk_off;                           // [nbu+1]
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
