// bundle_adjuster.h — C++ drop-in for the reference's BundleAdjuster, header-only over the C-ABI (sba.h).
//
// Replaces   src/bundle_adjuster.h:13-25  (class)  and  src/bundle_adjuster.cpp:10-133  (ctor + Optimize).
// Same constructor, same `void Optimize(unsigned start_frame_id, unsigned end_frame_id)` with the same
// half-open window [start, end) and the same write-back through Frame::GlobalPose(Matrix4f) and
// MapPoint::Position(std::array<double,3>), so the five call sites (src/slam.cpp:94,123,128,156 and
// src/pose_graph_optimizer.cpp:95) compile unchanged.  What changes is the middle: the ceres::Problem /
// ceres::Solve of src/bundle_adjuster.cpp:59-117 becomes one sba_solve() call on the GPU.
//
// The class is a template over the map types so that it needs neither Eigen nor OpenCV itself:
//   Frame     must offer  GlobalPose() -> M (4x4, element access M(r,c), default-constructible),
//                         GlobalPose(const M&), Observations() -> range of {point_id,u_l,v_l,u_r,v_r}
//   MapPoint  must offer  Position() -> std::array<float,3>,  Position(std::array<double,3>)
// which is exactly the surface of src/camera_frame.h:29-32,75 / src/map_point.h:30-42 / src/observation.h:6-17.
// In the reference tree:   using BundleAdjuster = sba::BundleAdjusterT<Frame, MapPoint>;   
//
// Projections: the reference keeps them in process globals (ReprojectionError::SetLeftProjection /
// SetRightProjection, src/reprojection_error.h:43-51).  sba::ReprojectionError below offers the same two
// static setters; the values travel to the library per call (the library itself has no global state).
#ifndef SBA_BUNDLE_ADJUSTER_H_
#define SBA_BUNDLE_ADJUSTER_H_

#include <array>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

#include "sba.h"

namespace sba {

// src/reprojection_error.h:43-51,65-66 — process-global row-major 3x4 projections
struct ReprojectionError {
  static std::array<double, 12>& p_l() { static std::array<double, 12> v{}; return v; }
  static std::array<double, 12>& p_r() { static std::array<double, 12> v{}; return v; }
  static void SetLeftProjection(const std::array<double, 12>& p) { p_l() = p; }
  static void SetRightProjection(const std::array<double, 12>& p) { p_r() = p; }
};

template <class Frame, class MapPoint>
class BundleAdjusterT {
 public:
  // src/bundle_adjuster.cpp:10-37: holds references to the caller's vectors; options = the reference's
  BundleAdjusterT(std::vector<Frame*>& cam_frames, std::vector<MapPoint*>& ldm_points, int device = 0)
      : m_cam_frames(cam_frames), m_ldm_points(ldm_points), m_device(device) {
    sba_default_options(&m_options);
  }
  ~BundleAdjusterT() { if (m_handle) sba_destroy(m_handle); }
  BundleAdjusterT(const BundleAdjusterT&) = delete;
  BundleAdjusterT& operator=(const BundleAdjusterT&) = delete;

  sba_options& Options() { return m_options; }
  const sba_summary& Summary() const { return m_summary; }
  int Status() const { return m_status; }
  // packed window of the last Pack()/Optimize() (test and debugging aid)
  const std::vector<double>& PackedPoses() const { return m_poses; }
  const std::vector<double>& PackedPoints() const { return m_points; }
  const std::vector<uint32_t>& PackedPointIds() const { return m_ids; }
  const std::vector<int32_t>& PackedObsCam() const { return m_cam; }
  const std::vector<int32_t>& PackedObsPt() const { return m_pt; }
  const std::vector<float>& PackedObsUv() const { return m_uvp; }

  // The pack phase alone (src/bundle_adjuster.cpp:41-110): fills the flat arrays below and returns the number
  // of window points.  Host-only; Optimize() calls it first.
  int32_t Pack(unsigned int start_frame_id, unsigned int end_frame_id) {
    if (end_frame_id < start_frame_id || end_frame_id > m_cam_frames.size())
      throw std::out_of_range("BundleAdjuster: bad window");
    const int32_t nc = (int32_t)(end_frame_id - start_frame_id);
    // ---- gather the window into flat arrays (the layout sba_pack_window reads)
    m_T.resize((size_t)nc * 16);
    m_off.assign((size_t)nc + 1, 0);
    m_pid.clear(); m_uv.clear();
    for (int32_t i = 0; i < nc; ++i) {
      Frame* frame = m_cam_frames[start_frame_id + i];
      const auto pose = frame->GlobalPose();                                    // :65-66
      for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m_T[(size_t)i * 16 + 4 * r + c] = pose(r, c);
      const auto& observations = frame->Observations();                         // :72
      for (const auto& obs : observations) {
        m_pid.push_back(obs.point_id);
        m_uv.push_back(obs.u_l); m_uv.push_back(obs.v_l); m_uv.push_back(obs.u_r); m_uv.push_back(obs.v_r);
      }
      m_off[(size_t)i + 1] = (int64_t)m_pid.size();
    }
    const size_t nobs = m_pid.size();
    // id-indexed staging of the WINDOW's points: only they are read and written here (:81-87); the buffer persists
    // across calls, so growing it costs O(points created since the last call), not O(map)
    if (m_map.size() < m_ldm_points.size() * 3) m_map.resize(m_ldm_points.size() * 3);
    for (uint32_t id : m_pid) {
      if (id >= m_ldm_points.size()) throw std::out_of_range("BundleAdjuster::Optimize: observation of an unknown map point");
      const std::array<float, 3> p = m_ldm_points[id]->Position();
      m_map[(size_t)id * 3] = p[0]; m_map[(size_t)id * 3 + 1] = p[1]; m_map[(size_t)id * 3 + 2] = p[2];
    }
    // ---- pack: camera / point / residual-block numbering bit-exact with :57-102
    m_poses.resize((size_t)nc * 6);
    const size_t cap = nobs ? nobs : 1;
    m_points.resize(cap * 3); m_ids.resize(cap); m_cam.resize(cap); m_pt.resize(cap); m_uvp.resize(cap * 4);
    int32_t np = 0;
    check(sba_pack_window(m_T.data(), m_map.data(), m_off.data(), m_pid.data(), m_uv.data(), 0, nc, m_poses.data(),
                          m_points.data(), m_ids.data(), m_cam.data(), m_pt.data(), m_uvp.data(), &np), "sba_pack_window");
    return np;
  }

  // src/bundle_adjuster.cpp:39-133.  The reference returns void and never looks at the termination type: results
  // are written back whatever the solver says — also after a numerical failure (the last accepted state; see
  // Status() / Summary()).  Only hard errors (no CUDA device, NCCL, invalid indices) throw: there is no CPU fallback.
  void Optimize(unsigned int start_frame_id, unsigned int end_frame_id) {
    if (end_frame_id <= start_frame_id) return;
    const int32_t np = Pack(start_frame_id, end_frame_id);
    const int32_t nc = (int32_t)(end_frame_id - start_frame_id);
    const size_t nobs = m_pid.size();
    // ---- solve (:112-117): first camera of the window constant (cam_fixed = NULL), one window
    if (!m_handle) check(sba_create(m_device, &m_handle), "sba_create");
    m_status = sba_solve(m_handle, nc, np, (int64_t)nobs, m_poses.data(), m_points.data(), m_cam.data(), m_pt.data(),
                         m_uvp.data(), ReprojectionError::p_l().data(), ReprojectionError::p_r().data(), nullptr, 1,
                         nullptr, nullptr, &m_options, &m_summary);
    if (m_status != SBA_OK && m_status != SBA_NO_CONVERGENCE && m_status != SBA_NUMERICAL_FAILURE) check(m_status, "sba_solve");
    if (m_options.verbose)
      std::printf("[sba] cost %.6e -> %.6e, %d iterations, termination %d, %.3f ms\n", m_summary.initial_cost,
                  m_summary.final_cost, m_summary.iterations, m_summary.termination, 1e3 * m_summary.time_total_s);
    // ---- write back (:121-132): every window camera (the fixed one included), then every window point
    for (int32_t i = 0; i < nc; ++i) {
      float T_cw[16], T_wc[16];
      sba_pose_to_matrix(m_poses.data() + (size_t)i * 6, T_cw);                 // :124
      check(sba_invert_pose(T_cw, T_wc), "sba_invert_pose");                    // :125 trans.inverse()
      auto M = m_cam_frames[start_frame_id + i]->GlobalPose();
      for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) M(r, c) = T_wc[4 * r + c];
      m_cam_frames[start_frame_id + i]->GlobalPose(M);                          // setter side effects are the Frame's own
    }
    for (int32_t p = 0; p < np; ++p) {
      const std::array<double, 3> pos = {m_points[(size_t)p * 3], m_points[(size_t)p * 3 + 1], m_points[(size_t)p * 3 + 2]};
      m_ldm_points[m_ids[p]]->Position(pos);                                    // :131
    }
  }

 private:
  static void check(int status, const char* what) {
    if (status != SBA_OK)
      throw std::runtime_error(std::string(what) + ": " + sba_status_string(status) + ": " + sba_last_error());
  }

  std::vector<Frame*>& m_cam_frames;
  std::vector<MapPoint*>& m_ldm_points;
  int m_device;
  sba_options m_options;
  sba_summary m_summary{};
  sba_handle m_handle = nullptr;
  int m_status = SBA_OK;
  // scratch kept across calls (the reference re-allocates per call, :41-57)
  std::vector<float> m_T, m_map, m_uv, m_uvp;
  std::vector<int64_t> m_off;
  std::vector<uint32_t> m_pid, m_ids;
  std::vector<int32_t> m_cam, m_pt;
  std::vector<double> m_poses, m_points;
};

}  // namespace sba

#endif  // SBA_BUNDLE_ADJUSTER_H_
