"""ORACLE — test infrastructure only.  numpy (float32) restatement of the host side of
BundleAdjuster::Optimize: the pack loop, the pose (de)parameterisation and the write-back.

PARITY UNPINNED for the Eigen float kernels (Eigen is absent here); the *index build* is pure
integer logic restated line by line and is exact.

reference:
  pack loop / first-appearance point numbering   src/bundle_adjuster.cpp:41-110
  MatrixToPose / PoseToMatrix / Normalize        src/math_utils.h:6-41
  unpack + Frame::GlobalPose setter              src/bundle_adjuster.cpp:121-132, src/camera_frame.h:32-49
  Eigen::AngleAxisf(Matrix3f), Quaternionf(Matrix3f), toRotationMatrix  (Eigen 3.3+/3.4 published algorithm)
"""
import numpy as np

F = np.float32


def quat_from_matrix(m):
    """Eigen::Quaternionf(Matrix3f) (Shepperd's method as in Eigen/Quaternion.h), float32. -> (w,x,y,z)"""
    m = np.asarray(m, F)
    t = F(m[0, 0] + m[1, 1] + m[2, 2])
    q = np.zeros(4, F)  # x y z w storage like Eigen coeffs()
    if t > F(0):
        t = np.sqrt(F(t + F(1.0)))
        q[3] = F(0.5) * t
        t = F(0.5) / t
        q[0] = (m[2, 1] - m[1, 2]) * t
        q[1] = (m[0, 2] - m[2, 0]) * t
        q[2] = (m[1, 0] - m[0, 1]) * t
    else:
        i = 0
        if m[1, 1] > m[0, 0]:
            i = 1
        if m[2, 2] > m[i, i]:
            i = 2
        j = (i + 1) % 3
        k = (j + 1) % 3
        t = np.sqrt(F(m[i, i] - m[j, j] - m[k, k] + F(1.0)))
        q[i] = F(0.5) * t
        t = F(0.5) / t
        q[3] = (m[k, j] - m[j, k]) * t
        q[j] = (m[j, i] + m[i, j]) * t
        q[k] = (m[k, i] + m[i, k]) * t
    return q


def quat_to_matrix(q):
    """Eigen QuaternionBase::toRotationMatrix, float32; q = (x,y,z,w)."""
    x, y, z, w = (F(v) for v in q)
    tx, ty, tz = F(2) * x, F(2) * y, F(2) * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    R = np.zeros((3, 3), F)
    R[0, 0] = F(1) - (tyy + tzz); R[0, 1] = txy - twz; R[0, 2] = txz + twy
    R[1, 0] = txy + twz; R[1, 1] = F(1) - (txx + tzz); R[1, 2] = tyz - twx
    R[2, 0] = txz - twy; R[2, 1] = tyz + twx; R[2, 2] = F(1) - (txx + tyy)
    return R


def normalize(T):
    """Normalize(Matrix4f&): re-orthonormalise R through a normalised quaternion (src/math_utils.h:6-10)."""
    T = np.array(T, F)
    q = quat_from_matrix(T[:3, :3])
    n = np.sqrt(F(np.sum(q * q, dtype=F)))
    if n > 0:
        q = q / n
    T[:3, :3] = quat_to_matrix(q)
    return T


def matrix_to_pose(T):
    """MatrixToPose (src/math_utils.h:12-25): AngleAxisf(R) -> [axis*angle, t] widened to f64."""
    T = np.asarray(T, F)
    q = quat_from_matrix(T[:3, :3])
    vec = q[:3]
    n = np.sqrt(F(np.sum(vec * vec, dtype=F)))
    if n != 0:
        angle = F(2) * np.arctan2(n, np.abs(q[3])).astype(F)
        if q[3] < 0:
            n = -n
        axis = vec / n
    else:
        angle = F(0); axis = np.array([1, 0, 0], F)
    pose = np.zeros(6, np.float64)
    pose[:3] = (axis * angle).astype(np.float64)
    pose[3:] = T[:3, 3].astype(np.float64)
    return pose


def pose_to_matrix(pose):
    """PoseToMatrix (src/math_utils.h:27-41): f64 pose narrowed to f32, AngleAxisf(angle, axis)."""
    r = np.asarray(pose[:3], np.float64).astype(F)
    f = np.sqrt(F(np.sum(r * r, dtype=F)))
    rn = r / f if f > 0 else r
    s, c = np.sin(f).astype(F), np.cos(f).astype(F)
    sin_axis = s * rn
    cos1_axis = (F(1) - c) * rn
    R = np.zeros((3, 3), F)
    tmp = cos1_axis[0] * rn[1]; R[0, 1] = tmp - sin_axis[2]; R[1, 0] = tmp + sin_axis[2]
    tmp = cos1_axis[0] * rn[2]; R[0, 2] = tmp + sin_axis[1]; R[2, 0] = tmp - sin_axis[1]
    tmp = cos1_axis[1] * rn[2]; R[1, 2] = tmp - sin_axis[0]; R[2, 1] = tmp + sin_axis[0]
    R[0, 0] = cos1_axis[0] * rn[0] + c; R[1, 1] = cos1_axis[1] * rn[1] + c; R[2, 2] = cos1_axis[2] * rn[2] + c
    T = np.eye(4, dtype=F)
    T[:3, :3] = R
    T[:3, 3] = np.asarray(pose[3:6], np.float64).astype(F)
    return T


def inverse4(T):
    """Matrix4f::inverse(): general 4x4 inverse in float32."""
    return np.linalg.inv(np.asarray(T, F)).astype(F)


def pack(smap, start, end):
    """The pack loop (src/bundle_adjuster.cpp:41-110).  Returns dict with poses [Nc,6] f64,
    points [Np,3] f64, point_ids [Np] u32, obs_cam/obs_pt [Nobs] i32, obs_uv [Nobs,4] f32."""
    poses, points, point_ids, obs_cam, obs_pt, obs_uv = [], [], [], [], [], []
    glb_id_to_pt_idx = {}
    for frame_id in range(start, end):
        poses.append(matrix_to_pose(inverse4(smap.T_wc[frame_id])))
        for k in range(int(smap.frame_off[frame_id]), int(smap.frame_off[frame_id + 1])):
            gid = int(smap.obs_point_id[k])
            if gid not in glb_id_to_pt_idx:
                points.append(smap.points[gid].astype(np.float64))
                point_ids.append(gid)
                glb_id_to_pt_idx[gid] = len(points) - 1
            obs_cam.append(frame_id - start)
            obs_pt.append(glb_id_to_pt_idx[gid])
            obs_uv.append(smap.obs_uv[k])
    return dict(poses=np.array(poses, np.float64).reshape(-1, 6), points=np.array(points, np.float64).reshape(-1, 3),
                point_ids=np.array(point_ids, np.uint32), obs_cam=np.array(obs_cam, np.int32),
                obs_pt=np.array(obs_pt, np.int32), obs_uv=np.array(obs_uv, F).reshape(-1, 4))


def unpack(smap, start, end, poses, points, point_ids):
    """Write-back (src/bundle_adjuster.cpp:121-132).  The Frame::GlobalPose setter's rigid drag of
    first-observed points (src/camera_frame.h:35,111-120) is overwritten by the final point loop for
    every point it touches inside the window, so only Normalize() survives."""
    T = smap.T_wc.copy(); P = smap.points.copy()
    for i in range(start, end):
        T[i] = normalize(inverse4(pose_to_matrix(poses[i - start])))
    for i, gid in enumerate(point_ids):
        P[int(gid)] = np.asarray(points[i], np.float64).astype(F)
    return T, P
