// host_pack.cpp — host side of BundleAdjuster::Optimize around the solve: the pack loop, the pose
// (de)parameterisation and the write-back, on flat arrays (see include/sba.h).
//
// reference:
//   pack loop, first-appearance point numbering     src/bundle_adjuster.cpp:41-110
//   MatrixToPose / PoseToMatrix / Normalize         src/math_utils.h:6-41   (all float)
//   write-back                                      src/bundle_adjuster.cpp:121-132
//   Frame::GlobalPose setter (Normalize after set)  src/camera_frame.h:32-49
// The reference leans on Eigen (AngleAxisf, Quaternionf, Matrix4f::inverse); Eigen is absent here,
// so the published algorithms of those classes are restated in float.
#include <cmath>
#include <cstring>
#include <unordered_map>
#include <vector>

#include "../../include/sba.h"

namespace {

// Quaternionf(Matrix3f): coefficient order (x, y, z, w)
void quat_from_rot(const float R[9], float q[4]) {
  float t = R[0] + R[4] + R[8];
  if (t > 0.0f) {
    t = std::sqrt(t + 1.0f);
    q[3] = 0.5f * t;
    t = 0.5f / t;
    q[0] = (R[7] - R[5]) * t;
    q[1] = (R[2] - R[6]) * t;
    q[2] = (R[3] - R[1]) * t;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0f);
    q[i] = 0.5f * t;
    t = 0.5f / t;
    q[3] = (R[3 * k + j] - R[3 * j + k]) * t;
    q[j] = (R[3 * j + i] + R[3 * i + j]) * t;
    q[k] = (R[3 * k + i] + R[3 * i + k]) * t;
  }
}

void quat_to_rot(const float q[4], float R[9]) {
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  const float tx = 2.0f * x, ty = 2.0f * y, tz = 2.0f * z;
  const float twx = tx * w, twy = ty * w, twz = tz * w;
  const float txx = tx * x, txy = ty * x, txz = tz * x;
  const float tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1.0f - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1.0f - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1.0f - (txx + tyy);
}

void rot_of(const float T[16], float R[9]) {
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R[3 * i + j] = T[4 * i + j];
}
void set_rot(float T[16], const float R[9]) {
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) T[4 * i + j] = R[3 * i + j];
}

// Normalize(Matrix4f&), src/math_utils.h:6-10
void normalize_pose(float T[16]) {
  float R[9], q[4];
  rot_of(T, R);
  quat_from_rot(R, q);
  const float n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n > 0.0f) for (int i = 0; i < 4; ++i) q[i] /= n;
  quat_to_rot(q, R);
  set_rot(T, R);
}

// general 4x4 inverse by cofactors (what Matrix4f::inverse() computes), float
bool inverse4(const float m[16], float inv[16]) {
  const float s0 = m[0] * m[5] - m[4] * m[1], s1 = m[0] * m[6] - m[4] * m[2], s2 = m[0] * m[7] - m[4] * m[3];
  const float s3 = m[1] * m[6] - m[5] * m[2], s4 = m[1] * m[7] - m[5] * m[3], s5 = m[2] * m[7] - m[6] * m[3];
  const float c5 = m[10] * m[15] - m[14] * m[11], c4 = m[9] * m[15] - m[13] * m[11], c3 = m[9] * m[14] - m[13] * m[10];
  const float c2 = m[8] * m[15] - m[12] * m[11], c1 = m[8] * m[14] - m[12] * m[10], c0 = m[8] * m[13] - m[12] * m[9];
  const float det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
  if (det == 0.0f || !std::isfinite(det)) return false;
  const float id = 1.0f / det;
  inv[0] = (m[5] * c5 - m[6] * c4 + m[7] * c3) * id;
  inv[1] = (-m[1] * c5 + m[2] * c4 - m[3] * c3) * id;
  inv[2] = (m[13] * s5 - m[14] * s4 + m[15] * s3) * id;
  inv[3] = (-m[9] * s5 + m[10] * s4 - m[11] * s3) * id;
  inv[4] = (-m[4] * c5 + m[6] * c2 - m[7] * c1) * id;
  inv[5] = (m[0] * c5 - m[2] * c2 + m[3] * c1) * id;
  inv[6] = (-m[12] * s5 + m[14] * s2 - m[15] * s1) * id;
  inv[7] = (m[8] * s5 - m[10] * s2 + m[11] * s1) * id;
  inv[8] = (m[4] * c4 - m[5] * c2 + m[7] * c0) * id;
  inv[9] = (-m[0] * c4 + m[1] * c2 - m[3] * c0) * id;
  inv[10] = (m[12] * s4 - m[13] * s2 + m[15] * s0) * id;
  inv[11] = (-m[8] * s4 + m[9] * s2 - m[11] * s0) * id;
  inv[12] = (-m[4] * c3 + m[5] * c1 - m[6] * c0) * id;
  inv[13] = (m[0] * c3 - m[1] * c1 + m[2] * c0) * id;
  inv[14] = (-m[12] * s3 + m[13] * s1 - m[14] * s0) * id;
  inv[15] = (m[8] * s3 - m[9] * s1 + m[10] * s0) * id;
  return true;
}

}  // namespace

extern "C" {

// MatrixToPose, src/math_utils.h:12-25: AngleAxisf(R) -> axis * angle, widened
void sba_matrix_to_pose(const float* T, double* pose) {
  float R[9], q[4];
  rot_of(T, R);
  quat_from_rot(R, q);
  float n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
  float angle, axis[3];
  if (n != 0.0f) {
    angle = 2.0f * std::atan2(n, std::fabs(q[3]));
    if (q[3] < 0.0f) n = -n;
    axis[0] = q[0] / n; axis[1] = q[1] / n; axis[2] = q[2] / n;
  } else {
    angle = 0.0f; axis[0] = 1.0f; axis[1] = 0.0f; axis[2] = 0.0f;
  }
  pose[0] = (double)(axis[0] * angle);
  pose[1] = (double)(axis[1] * angle);
  pose[2] = (double)(axis[2] * angle);
  pose[3] = (double)T[3]; pose[4] = (double)T[7]; pose[5] = (double)T[11];
}

// PoseToMatrix, src/math_utils.h:27-41: pose narrowed to float, AngleAxisf(|r|, r/|r|).toRotationMatrix()
void sba_pose_to_matrix(const double* pose, float* T) {
  const float r[3] = {(float)pose[0], (float)pose[1], (float)pose[2]};
  const float f = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  float a[3] = {r[0], r[1], r[2]};
  if (f > 0.0f) { a[0] = r[0] / f; a[1] = r[1] / f; a[2] = r[2] / f; }   // normalized(): zero stays zero
  const float s = std::sin(f), c = std::cos(f);
  const float sa[3] = {s * a[0], s * a[1], s * a[2]};
  const float ca[3] = {(1.0f - c) * a[0], (1.0f - c) * a[1], (1.0f - c) * a[2]};
  float R[9];
  float tmp;
  tmp = ca[0] * a[1]; R[1] = tmp - sa[2]; R[3] = tmp + sa[2];
  tmp = ca[0] * a[2]; R[2] = tmp + sa[1]; R[6] = tmp - sa[1];
  tmp = ca[1] * a[2]; R[5] = tmp - sa[0]; R[7] = tmp + sa[0];
  R[0] = ca[0] * a[0] + c; R[4] = ca[1] * a[1] + c; R[8] = ca[2] * a[2] + c;
  for (int i = 0; i < 16; ++i) T[i] = (i % 5 == 0) ? 1.0f : 0.0f;
  set_rot(T, R);
  T[3] = (float)pose[3]; T[7] = (float)pose[4]; T[11] = (float)pose[5];
}

int sba_invert_pose(const float* T, float* out) { return inverse4(T, out) ? SBA_OK : SBA_NUMERICAL_FAILURE; }
void sba_normalize_pose(float* T) { normalize_pose(T); }
void sba_compose_pose(const float* A, const float* B, float* out) {
  float tmp[16];
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      float s = 0.0f;
      for (int k = 0; k < 4; ++k) s += A[4 * i + k] * B[4 * k + j];
      tmp[4 * i + j] = s;
    }
  std::memcpy(out, tmp, 64);
}

int sba_pack_window(const float* T_wc, const float* map_points, const int64_t* frame_off,
                    const uint32_t* obs_point_id, const float* obs_uv_in, int32_t start, int32_t end,
                    double* poses, double* points, uint32_t* point_ids, int32_t* obs_cam,
                    int32_t* obs_pt, float* obs_uv, int32_t* num_points) {
  if (!T_wc || !frame_off || start < 0 || end < start || !num_points) return SBA_INVALID_ARGUMENT;
  *num_points = 0;
  if (end == start) return SBA_OK;
  if (!poses || !obs_point_id || !obs_uv_in || !map_points || !points || !point_ids || !obs_cam || !obs_pt || !obs_uv)
    return SBA_INVALID_ARGUMENT;
  std::unordered_map<uint32_t, int32_t> glb_id_to_pt_idx;                      // :57
  glb_id_to_pt_idx.reserve((size_t)(frame_off[end] - frame_off[start]));
  int32_t np = 0;
  int64_t k_out = 0;
  for (int32_t frame_id = start; frame_id < end; ++frame_id) {                 // :62
    float T_cw[16];
    if (!inverse4(T_wc + (size_t)frame_id * 16, T_cw)) return SBA_INVALID_ARGUMENT;   // :66
    sba_matrix_to_pose(T_cw, poses + (size_t)(frame_id - start) * 6);          // :67-69
    for (int64_t k = frame_off[frame_id]; k < frame_off[frame_id + 1]; ++k, ++k_out) {   // :72-74
      const uint32_t gid = obs_point_id[k];
      auto it = glb_id_to_pt_idx.find(gid);                                    // :79
      int32_t pt_idx;
      if (it == glb_id_to_pt_idx.end()) {
        pt_idx = np++;
        points[(size_t)pt_idx * 3 + 0] = (double)map_points[(size_t)gid * 3 + 0];   // :83-87
        points[(size_t)pt_idx * 3 + 1] = (double)map_points[(size_t)gid * 3 + 1];
        points[(size_t)pt_idx * 3 + 2] = (double)map_points[(size_t)gid * 3 + 2];
        point_ids[pt_idx] = gid;                                               // :90
        glb_id_to_pt_idx.emplace(gid, pt_idx);                                 // :93
      } else {
        pt_idx = it->second;
      }
      obs_cam[k_out] = frame_id - start;                                       // :102
      obs_pt[k_out] = pt_idx;
      std::memcpy(obs_uv + (size_t)k_out * 4, obs_uv_in + (size_t)k * 4, 16);  // :99
    }
  }
  *num_points = np;
  return SBA_OK;
}

int sba_unpack_window(float* T_wc, float* map_points, int32_t start, int32_t end, const double* poses,
                      const double* points, const uint32_t* point_ids, int32_t num_points) {
  if (start < 0 || end < start || num_points < 0) return SBA_INVALID_ARGUMENT;
  if (end > start && (!T_wc || !poses)) return SBA_INVALID_ARGUMENT;
  if (num_points > 0 && (!map_points || !points || !point_ids)) return SBA_INVALID_ARGUMENT;
  for (int32_t i = start; i < end; ++i) {                                      // :121-126
    float T_cw[16], T[16];
    sba_pose_to_matrix(poses + (size_t)(i - start) * 6, T_cw);
    if (!inverse4(T_cw, T)) return SBA_NUMERICAL_FAILURE;
    normalize_pose(T);                                                         // camera_frame.h:37-38
    std::memcpy(T_wc + (size_t)i * 16, T, 64);
  }
  // The setter's rigid drag of first-observed points (camera_frame.h:35) is overwritten here for
  // every point it touches inside the window.
  for (int32_t p = 0; p < num_points; ++p) {                                   // :128-132
    const size_t g = (size_t)point_ids[p] * 3;
    map_points[g + 0] = (float)points[(size_t)p * 3 + 0];
    map_points[g + 1] = (float)points[(size_t)p * 3 + 1];
    map_points[g + 2] = (float)points[(size_t)p * 3 + 2];
  }
  return SBA_OK;
}

}  // extern "C"
