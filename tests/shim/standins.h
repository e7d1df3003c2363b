// standins.h — TEST-ONLY stand-ins for the reference's map types, with the same public surface the
// BundleAdjuster touches (src/camera_frame.h:29-49,75,111-120; src/map_point.h:14-42; src/observation.h:6-17)
// but without Eigen / OpenCV (absent from this image).  The 4x4 float algebra goes through libsba.so's helpers.
#pragma once
#include <array>
#include <vector>

#include "../../include/sba.h"

struct Mat4 {                       // row-major 4x4 float with Eigen-like element access
  float v[16];
  Mat4() { for (int i = 0; i < 16; ++i) v[i] = (i % 5 == 0) ? 1.0f : 0.0f; }
  float& operator()(int r, int c) { return v[4 * r + c]; }
  float operator()(int r, int c) const { return v[4 * r + c]; }
  Mat4 inverse() const { Mat4 o; sba_invert_pose(v, o.v); return o; }
  Mat4 operator*(const Mat4& b) const { Mat4 o; sba_compose_pose(v, b.v, o.v); return o; }
};

struct Observation {
  Observation(unsigned id, float ul, float vl, float ur, float vr, float s) : point_id(id), u_l(ul), v_l(vl), u_r(ur), v_r(vr), sigma(s) {}
  unsigned point_id;
  float u_l, v_l, u_r, v_r, sigma;
};

class MapPoint {
 public:
  MapPoint(float x, float y, float z) : m_position{x, y, z} {}
  void Transform(const Mat4& T) {   // homogeneous transform in float, like src/map_point.h:21-28
    const float x = m_position[0], y = m_position[1], z = m_position[2];
    for (int i = 0; i < 3; ++i) m_position[i] = T(i, 0) * x + T(i, 1) * y + T(i, 2) * z + T(i, 3) * 1.0f;
  }
  void Position(std::array<double, 3> pos) { for (int i = 0; i < 3; ++i) m_position[i] = (float)pos[i]; }
  std::array<float, 3> Position() const { return m_position; }
 private:
  std::array<float, 3> m_position;
};

class Frame {
 public:
  Frame(Frame* prev, const Mat4& pose_glb) : m_prev(prev), m_glb(pose_glb) {}
  Mat4 GlobalPose() const { return m_glb; }
  Mat4 RelativePose() const { return m_rel; }
  // src/camera_frame.h:32-49: drag first-observed points by the pose change, normalise, refresh the relative pose
  void GlobalPose(const Mat4& pose) {
    const Mat4 diff = pose * m_glb.inverse();
    for (size_t i = 0; i < m_refs.size(); ++i) if (m_first[i]) m_refs[i]->Transform(diff);
    m_glb = pose;
    sba_normalize_pose(m_glb.v);
    Mat4 prev;
    if (m_prev) prev = m_prev->GlobalPose();
    m_rel = prev.inverse() * m_glb;
    sba_normalize_pose(m_rel.v);
  }
  void AddObservation(const Observation& o) { m_obs.push_back(o); }
  std::vector<Observation> Observations() const { return m_obs; }
  void AddMapPoints(MapPoint* p, bool first_observed) { m_refs.push_back(p); m_first.push_back(first_observed); }
 private:
  Frame* m_prev;
  Mat4 m_glb, m_rel;
  std::vector<MapPoint*> m_refs;
  std::vector<bool> m_first;
  std::vector<Observation> m_obs;
};
