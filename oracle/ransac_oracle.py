"""ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU (numpy) restatement of the stereo RANSAC alignment the reference's visual odometer and loop closer run before
bundle adjustment (reference: ransac_solver/include/ransac_solver.h:51-166 Solver::Solve;
ransac_solver/include/stereo_reprojection.h:28-134 StereoTransModel; callers src/visual_odometer.cpp:93-103, :157-177
and src/pose_graph_optimizer.cpp:189).  Only tests/ and bench legs may import this.

PARITY UNPINNED for the arithmetic that lives outside the reference: cv::triangulatePoints (OpenCV, not vendored,
not version-pinned) and Eigen::umeyama (Eigen, same) are absent from this image.  Pinned by reference source and
restated here line by line:
  * hypotheses: 3 elements drawn with replacement, `rand() % size` each            ransac_solver.h:84-90
    (the draw itself is an input here — `samples[iteration, 3]` — so that any generator, glibc's rand() included,
    reproduces; glibc's generator is restated in stereo_orb_slam_b200/ransac.py and pinned against libc in the tests)
  * best model = FIRST hypothesis with the strictly largest inlier count             :111-121
  * early termination when count > ceil(size * concensus_ratio) (float product)    :62, :123-125
  * optional final fit over the inliers of the best model (all elements if <= 3)   :141-151
  * labels / losses of every element under the final model                         :153-161
  * inlier test: triangulate element.first, transform, project with both 3x4 matrices, 4-term pixel distance
    sqrt(sum of squares) < 3.0, float arithmetic                                  stereo_reprojection.h:96-134
Restated from the libraries' published algorithms:
  * cv::triangulatePoints: per point the 4x4 DLT matrix  [u P_2 - P_0; v P_2 - P_1] for both views in double, its
    right singular vector of the smallest singular value, stored as float (inputs are Point2f and CV_32F matrices);
    the reference then divides by the 4th component in float (stereo_reprojection.h:83-92, :106-110)
  * Eigen::umeyama(src, dst, false): means, covariance sigma = dst_c src_c^T / n, SVD, R = U diag(1,1,s) V^T with
    s = sign(det U det V), t = mean_dst - R mean_src.  The reference runs it in float; here (and on the GPU) the
    3x3 problem is solved in double from the float points and the result is rounded to float — a float SVD is
    accurate to ~1e-6 relative at best, so the two agree to that level and no closer (tests/test_ransac.py checks
    the double fit against a float32 SVD formulation at 1e-4).
Float operations whose order Eigen leaves to the compiler (4x4 * 4-vector, 3x4 * 4-vector, the sum of four squares)
are fixed here as left-to-right sums without fused multiply-add; the CUDA path uses the same order
(__fmul_rn/__fadd_rn), so labels and losses are bit-comparable given the same model.
"""
import numpy as np

F = np.float32
INLIER_THRESHOLD = 3.0          # stereo_reprojection.h:126
MIN_PAIRS = 10                  # src/visual_odometer.cpp:100


class Option:
    """RANSAC::Option (ransac_solver.h:13-19) with the values src/params.h:18-27 sets."""
    def __init__(self, max_iterations=100, final_model_fitting=False, early_termination=False, concensus_ratio=0.8,
                 inlier_threshold=INLIER_THRESHOLD):
        self.max_iterations = int(max_iterations)
        self.final_model_fitting = bool(final_model_fitting)
        self.early_termination = bool(early_termination)
        self.concensus_ratio = float(concensus_ratio)
        self.inlier_threshold = float(inlier_threshold)


def triangulate(P_l, P_r, uv_l, uv_r):
    """cv::triangulatePoints + the reference's division by w.  P_* [3,4] f32, uv_* [n,2] f32 -> [n,3] f32."""
    P_l = np.asarray(P_l, F).reshape(3, 4).astype(np.float64)
    P_r = np.asarray(P_r, F).reshape(3, 4).astype(np.float64)
    uv_l = np.asarray(uv_l, F).reshape(-1, 2).astype(np.float64)
    uv_r = np.asarray(uv_r, F).reshape(-1, 2).astype(np.float64)
    n = uv_l.shape[0]
    A = np.empty((n, 4, 4))
    A[:, 0] = uv_l[:, 0:1] * P_l[2] - P_l[0]
    A[:, 1] = uv_l[:, 1:2] * P_l[2] - P_l[1]
    A[:, 2] = uv_r[:, 0:1] * P_r[2] - P_r[0]
    A[:, 3] = uv_r[:, 1:2] * P_r[2] - P_r[1]
    if n == 0:
        return np.zeros((0, 3), F)
    _, _, Vt = np.linalg.svd(A)
    X = Vt[:, 3, :].astype(F)                     # points4D is CV_32F
    return (X[:, :3] / X[:, 3:4]).astype(F)       # float division, stereo_reprojection.h:85-90


def umeyama(src, dst):
    """Eigen::umeyama(src, dst, with_scaling=false) solved in double from float points; 4x4 f32 out."""
    src = np.asarray(src, F).reshape(-1, 3).astype(np.float64)
    dst = np.asarray(dst, F).reshape(-1, 3).astype(np.float64)
    n = src.shape[0]
    ms = src.sum(0) / n
    md = dst.sum(0) / n
    sigma = (dst - md).T @ (src - ms) / n
    U, s, Vt = np.linalg.svd(sigma)
    if s[0] <= 1e-20 * (ms @ ms + md @ md) or s[0] == 0.0:
        R = np.eye(3)                               # all picks coincide
    elif s[1] <= 1e-12 * s[0]:
        # rank 1 (a repeated or collinear pick: `rand() % size` draws with replacement).  Eigen's answer depends on how
        # its JacobiSVD completes the basis, which the reference's source does not define; stated completion instead
        # (the same on the GPU): u2 / v2 = the coordinate axis of the smallest component, orthogonalised.
        u1, v1 = U[:, 0], Vt[0]
        u2, v2 = _ortho(u1), _ortho(v1)
        R = np.outer(u1, v1) + np.outer(u2, v2) + np.outer(np.cross(u1, u2), np.cross(v1, v2))
    else:
        S = np.ones(3)
        if np.linalg.det(U) * np.linalg.det(Vt) < 0:
            S[2] = -1.0
        R = U @ np.diag(S) @ Vt
    T = np.eye(4)
    T[:3, :3] = R
    T[:3, 3] = md - R @ ms
    return T.astype(F)


def _ortho(a):
    m = 0
    if abs(a[1]) < abs(a[m]):
        m = 1
    if abs(a[2]) < abs(a[m]):
        m = 2
    e = np.zeros(3); e[m] = 1.0
    o = e - (a @ e) * a
    return o / np.linalg.norm(o)


def umeyama_float_svd(src, dst):
    """The same formula carried out in float32 throughout, as the reference's Eigen::MatrixXf call does — used only
    to show the double fit above agrees with a float formulation."""
    src = np.asarray(src, F).reshape(-1, 3)
    dst = np.asarray(dst, F).reshape(-1, 3)
    n = F(src.shape[0])
    ms = (src.sum(0, dtype=F) / n).astype(F)
    md = (dst.sum(0, dtype=F) / n).astype(F)
    sigma = ((dst - md).T @ (src - ms) / n).astype(F)
    U, _, Vt = np.linalg.svd(sigma)
    S = np.ones(3, F)
    if np.linalg.det(U.astype(np.float64)) * np.linalg.det(Vt.astype(np.float64)) < 0:
        S[2] = -1.0
    R = (U * S) @ Vt
    T = np.eye(4, dtype=F)
    T[:3, :3] = R
    T[:3, 3] = md - R @ ms
    return T


def _mat_vec(M, x):
    """rows of M (k x 4, f32) times the columns of x (4 x n, f32): ((m0 x0 + m1 x1) + m2 x2) + m3 x3, no FMA."""
    M = np.asarray(M, F)
    return np.stack([(((M[r, 0] * x[0]).astype(F) + (M[r, 1] * x[1]).astype(F)).astype(F)
                      + (M[r, 2] * x[2]).astype(F)).astype(F) + (M[r, 3] * x[3]).astype(F) for r in range(M.shape[0])]).astype(F)


def reprojection_distance(T, P_l, P_r, src_xyz, obs_dst):
    """StereoTransModel::IsInlier's `dist` for every element (stereo_reprojection.h:112-124).  f32 [n]."""
    T = np.asarray(T, F).reshape(4, 4)
    P_l = np.asarray(P_l, F).reshape(3, 4)
    P_r = np.asarray(P_r, F).reshape(3, 4)
    src_xyz = np.asarray(src_xyz, F).reshape(-1, 3)
    obs_dst = np.asarray(obs_dst, F).reshape(-1, 4)
    h = np.stack([src_xyz[:, 0], src_xyz[:, 1], src_xyz[:, 2], np.ones(src_xyz.shape[0], F)]).astype(F)
    with np.errstate(all="ignore"):
        s = _mat_vec(T, h)
        l = _mat_vec(P_l, s)
        r = _mat_vec(P_r, s)
        ul = (l[0] / l[2]).astype(F); vl = (l[1] / l[2]).astype(F)
        ur = (r[0] / r[2]).astype(F); vr = (r[1] / r[2]).astype(F)
        d0 = (obs_dst[:, 0] - ul).astype(F); d1 = (obs_dst[:, 1] - vl).astype(F)
        d2 = (obs_dst[:, 2] - ur).astype(F); d3 = (obs_dst[:, 3] - vr).astype(F)
        acc = (d0 * d0).astype(F)
        acc = (acc + (d1 * d1).astype(F)).astype(F)
        acc = (acc + (d2 * d2).astype(F)).astype(F)
        acc = (acc + (d3 * d3).astype(F)).astype(F)
        return np.sqrt(acc).astype(F)


def consensus_threshold(size, ratio):
    """unsigned int concensus_count_threshold = ceil(size * m_options.concensus_ratio)  (float product)."""
    return int(np.ceil(F(F(size) * F(ratio))))


def solve(obs_first, obs_second, P_l, P_r, samples, option=None):
    """RANSAC::Solver<ObservationPair, Matrix4f>::Solve with StereoTransModel.
    obs_first / obs_second [n,4] f32 (u_l, v_l, u_r, v_r): the model maps points triangulated from `first` onto
    `second`.  samples [max_iterations, 3] element indices (the reference draws rand() % size).
    Returns (T [4,4] f32, labels [n] bool, losses [n] f32, info)."""
    opt = option or Option()
    a = np.asarray(obs_first, F).reshape(-1, 4)
    b = np.asarray(obs_second, F).reshape(-1, 4)
    n = a.shape[0]
    samples = np.asarray(samples, np.int64).reshape(-1, 3)
    assert samples.shape[0] >= opt.max_iterations and n > 0
    src = triangulate(P_l, P_r, a[:, 0:2], a[:, 2:4])
    dst = triangulate(P_l, P_r, b[:, 0:2], b[:, 2:4])
    thr = consensus_threshold(n, opt.concensus_ratio)
    thr_px = F(opt.inlier_threshold)
    best_count, best_it = -1, -1
    T_final = np.zeros((4, 4), F)
    inl_final = np.zeros(n, bool)
    counts = np.full(opt.max_iterations, -1, np.int64)
    for it in range(opt.max_iterations):
        idx = samples[it]
        T = umeyama(src[idx], dst[idx])
        inl = reprojection_distance(T, P_l, P_r, src, b) < thr_px
        count = int(inl.sum())
        counts[it] = count
        if count > best_count:
            best_count, best_it, T_final, inl_final = count, it, T, inl
        if opt.early_termination and count > thr:
            break
    if opt.final_model_fitting:
        if int(inl_final.sum()) > 3:
            T_final = umeyama(src[inl_final], dst[inl_final])
        else:
            T_final = umeyama(src, dst)
    losses = reprojection_distance(T_final, P_l, P_r, src, b)
    labels = losses < thr_px
    return T_final, labels, losses, dict(best_iteration=best_it, best_count=best_count, counts=counts, src=src, dst=dst)


def calc_transformation(obs_first, obs_second, P_l, P_r, samples, option=None):
    """VisualOdometer::CalcTransformation (src/visual_odometer.cpp:93-103): fewer than 10 pairs -> failure."""
    if np.asarray(obs_first).reshape(-1, 4).shape[0] < MIN_PAIRS:
        return False, None, None, None
    T, labels, losses, _ = solve(obs_first, obs_second, P_l, P_r, samples, option)
    return True, T, labels, losses
