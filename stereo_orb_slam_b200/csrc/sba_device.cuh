// sba_device.cuh — device-resident tables and state of one handle; passed by value to kernels.
#pragma once
#include <cstdint>
#include "sba_math.cuh"

namespace sba {

struct DevParams {   // solver constants, by value
  double pl[12], pr[12];
  double huber_delta, lo, hi;
  double ptol, ftol, gtol;
  double pcg_tol2;     // pcg_tolerance^2
  int32_t rdim, max_iterations;
  int32_t debug;       // SBA_DEBUG_LS: trace the line search of window 0 (device printf)
};

constexpr int kMaxRanks = 16;    // ranks of one point-sharded solve (slots of the max-by-sum exchange)

struct Dev {
  int32_t Nc, Np, Nf, nwin, nbu, npchunk, ncs, nlb;
  int64_t Nobs;
  // ---- structure (const): see sba_internal.h
  const int32_t *fc, *free2cam, *cam_win, *fcam_win, *win_cam_off, *win_fc_off, *win_pchunk_off;
  const int32_t *pt_off, *perm_p, *pcam, *ppt;
  const int32_t *pchunk_begin, *pchunk_end, *pchunk_win;
  const int4* cdesc;            // [npchunk * 4]  chunk records (HostStructure::cdesc)
  const int32_t *slot_off, *cs_cam, *cs_obs_off, *lb_off, *lb_blk, *lb_pair_off;
  const uint8_t *oslot, *lb_sa, *lb_sb;
  const uint16_t *sobs, *pairs;
  const int32_t *blk_row, *blk_col, *diag_blk, *blk_part_off, *blk_part, *cam_part_off, *cam_part;
  const int32_t *row_ptr, *bsr_col, *bsr_blk;
  const int32_t *lb_pos, *cs_pos;
  // long tracks (points with > 255 observations: one chunk each, k_*_long kernels)
  const int32_t *long_chunks, *long_zoff, *long_lb_off;   // [nlong], [nlong+1], [nlong+1]
  int32_t nlong;
  double* zlong;                 // [long_zoff[nlong] * 18] Z = W Linv^T of their observations
  // ---- measurements (const during a solve)
  const float4 *uv_p;   // [Nobs] point order
  // ---- state
  double *poses, *poses_new;     // [Nc*6]
  double *pts, *pts_new;         // [Np*3]
  CamPre *campre_new;            // [Nc] trial cameras of the line search
  CamPre *campre;                // [Nc] cameras of the current state, written once per linearisation (k_campre) for k_fused_chunk
  // ---- linearisation
  double *W;                     // [Nobs*18] point order
  double *V, *gp;                // [Np*6], [Np*3]
  double *Vinv, *tp;             // [Np*6], [Np*3]
  double *U, *gc;                // [Nc*21], [Nc*6]   one block with cost_red and gslots behind: ONE all-reduce per linearisation
  double *cost_red, *gslots;     // [nwin] window cost partial of this rank; [kMaxRanks*nwin] projected-gradient max, a slot per rank
  double *cspart;                // [ncs*27] per chunk-slot partial of U (21) and gc (6)
  double *sc2c, *sc2p;           // Jacobi scale^2 [Nc*6], [Np*3]
  double *lin_part;              // [npchunk*2] cost, gmax
  double *r_dbg;                 // [Nobs*4] or null
  // ---- Schur
  double *spart;                 // [nlb*36] per local-block partial of sum W Vinv W^T
  double *Su;                    // [nbu*36]
  double *rpart;                 // [ncs*6] per chunk-slot partial of sum W tp
  double *rsum;                  // [Nf*6] sum over observations of W tp per free camera (directly behind Su: one all-reduce)
  double *rhs;                   // [Nf*6]
  double *Minv;                  // [Nf*36]
  // ---- PCG vectors [Nf*6]
  double *x, *r, *z, *p, *Ap;
  // ---- direct solve of small windows (k_chol_window): dense lower block-triangle of each window's S
  double *chol_ws;               // [sum over windows of nf (nf + 1) / 2 * 36]
  const int64_t *chol_off;       // [nwin] element offset of window w in chol_ws
  // ---- step
  double *dp;                    // [Np*3] point step (back-substitution)
  double *step_part;             // [npchunk*4]  g.dp, dp D dp, |x|^2, max|dp|
  double *ls_part;               // [npchunk*3]  trial cost, directional derivative, |x_trial - x|^2
  double *win_part;              // [nwin*2] cost and gradient maximum over a window's chunks (k_assemble_schur's window CTAs)
  const int32_t *asm_ids; int32_t n_asm;   // blocks (< nbu) and cameras (nbu + f) with a non-empty partial list on this rank
  int32_t *win_done;             // [nwin] tickets of k_assemble_schur's camera CTAs (zero between launches)
  double *cta_part;              // [step-loop CTAs * 16] one global window: per-CTA partials of the rows above and of the
                                 //   camera terms, so that the window's reductions sum a few hundred rows, not 10^5
  // ---- per-window scalars [nwin]
  double *w_radius, *w_decf, *w_cost, *w_cand, *w_gmax;
  double *w_gd, *w_dDd, *w_dr, *w_x2;     // one block [4*nwin] (all-reduced together on multi-GPU)
  double *w_dmax;                          // max-norm of the step
  double *dslots;                          // [kMaxRanks*nwin] behind w_gd..w_dr: every rank's max-norm, summed into its own slot
  double *w_ls_red;                        // [3*nwin] trial cost, directional derivative, |step|^2 (raw sums)
  double *w_step2;
  // projected Armijo line search state (sba_linesearch.cuh)
  double *w_ls_x, *w_ls_px, *w_ls_pf, *w_ls_pg;
  int32_t *w_ls_state, *w_ls_it, *w_ls_pvalid;
  double *w_rz, *w_rz0, *w_pAp, *w_alpha, *w_beta;
  int32_t *w_status, *w_iter, *w_succ, *w_invalid, *w_accept, *w_fail, *w_pcg_done, *w_pcg_it;
  int32_t *counters;             // [0] windows still running, [1] pcg windows not done, [2] accepted, [3] line searches pending
};

}  // namespace sba
