"""ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU (numpy) restatement of the pose-graph optimisation the reference hands to g2o
(reference: src/pose_graph_optimizer.cpp:34-96 Optimize, :98-170 AddOdometryConstraints, :172-217
AddLoopClosureConstraints).  Only tests/ and bench legs may import this.

PARITY UNPINNED: the arithmetic lives in g2o (g2o::VertexSE3, g2o::EdgeSE3, RobustKernelHuber,
OptimizationAlgorithmLevenberg over BlockSolverX + LinearSolverEigen), which is neither vendored nor version-pinned
by the reference (CMakeLists.txt) and is absent from this image.  Pinned by reference source and restated here:
  * vertices = frames' world-from-camera poses (t, q)            src/pose_graph_optimizer.cpp:103-141
  * odometry edges (i-1 -> i), measurement = Frame::RelativePose  :143-168
  * loop edges, vertex 0 = id_2, vertex 1 = id_1                  :192-211
  * information diag(0.01, 0.01, 0.01, 1, 1, 1)  (translation first)   :23-26
  * Huber kernel, default delta 1                                 :162, :207 ; src/pose_graph_optimizer.h:33
  * vertex 0 fixed, 10 Levenberg iterations                       :123-126, :66
Restated from g2o's published algorithm:
  * EdgeSE3 error  e = toVectorMQT(Z^-1 X_i^-1 X_j)  (translation, then the vector part of the unit quaternion with
    w >= 0);  VertexSE3 update  X <- X * fromVectorMQT(delta)
  * robustified quadratic form: H += rho' J^T Omega J, b -= rho' J^T Omega e, chi2 = rho(e^T Omega e)
  * OptimizationAlgorithmLevenberg: lambda_0 = 1e-5 max diag(H); per iteration up to 10 trials of
    (H + lambda I) dx = b; gain rho = (chi2 - chi2_new) / (dx.(lambda dx + b) + 1e-3); accepted when rho > 0 with
    lambda *= max(1/3, min(2/3, 1 - (2 rho - 1)^3)), else lambda *= ni, ni *= 2
The Jacobians are the exact derivatives of that error at delta = 0 (g2o's computeEdgeSE3Gradient computes the same
derivative analytically); a central-difference check is part of tests/test_pgo.py.
"""
import numpy as np

INFORMATION = np.diag([0.01, 0.01, 0.01, 1.0, 1.0, 1.0])
HUBER_DELTA = 1.0


# ---------------------------------------------------------------- quaternions (x, y, z, w), unit
def quat_from_R(R):
    """Eigen::Quaterniond(Matrix3d) (Shepperd), then normalised with w >= 0 (g2o toCompactQuaternion)."""
    t = R[0, 0] + R[1, 1] + R[2, 2]
    q = np.zeros(4)
    if t > 0:
        s = np.sqrt(t + 1.0)
        q[3] = 0.5 * s
        s = 0.5 / s
        q[0] = (R[2, 1] - R[1, 2]) * s; q[1] = (R[0, 2] - R[2, 0]) * s; q[2] = (R[1, 0] - R[0, 1]) * s
    else:
        i = 0
        if R[1, 1] > R[0, 0]:
            i = 1
        if R[2, 2] > R[i, i]:
            i = 2
        j = (i + 1) % 3; k = (j + 1) % 3
        s = np.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0)
        q[i] = 0.5 * s
        s = 0.5 / s
        q[3] = (R[k, j] - R[j, k]) * s; q[j] = (R[j, i] + R[i, j]) * s; q[k] = (R[k, i] + R[i, k]) * s
    q /= np.linalg.norm(q)
    if q[3] < 0:
        q = -q
    return q


def R_from_quat(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def from_vector_mqt(v):
    """g2o internal::fromVectorMQT: translation v[0:3], rotation from the compact quaternion v[3:6]."""
    w2 = 1.0 - v[3:6] @ v[3:6]
    R = np.eye(3) if w2 < 0 else R_from_quat(np.array([v[3], v[4], v[5], np.sqrt(w2)]))
    return R, v[0:3].copy()


def edge_error(Ri, ti, Rj, tj, Rz, tz):
    """e = toVectorMQT(Z^-1 X_i^-1 X_j), plus the pieces the Jacobians need."""
    Rb = Ri.T @ Rj; tb = Ri.T @ (tj - ti)            # B = X_i^-1 X_j
    Re = Rz.T @ Rb; te = Rz.T @ (tb - tz)            # E = Z^-1 B
    q = quat_from_R(Re)
    return np.concatenate([te, q[:3]]), (Re, q, Rb, tb)


def skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]])


def quat_mul(a, b):
    ax, ay, az, aw = a; bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by - ax * bz + ay * bw + az * bx,
                     aw * bz + ax * by - ay * bx + az * bw, aw * bw - ax * bx - ay * by - az * bz])


def edge_jacobians(Ri, ti, Rj, tj, Rz, tz):
    """d e / d delta_i, d e / d delta_j at delta = 0 for X <- X * fromVectorMQT(delta)  (6x6 each)."""
    e, (Re, q, Rb, tb) = edge_error(Ri, ti, Rj, tj, Rz, tz)
    v, w = q[:3], q[3]
    Jj = np.zeros((6, 6)); Ji = np.zeros((6, 6))
    # X_j <- X_j D:  E' = E D  ->  t' = t_e + R_e dt ;  q' = q_e (x) (dq, 1)
    Jj[0:3, 0:3] = Re
    Jj[3:6, 3:6] = w * np.eye(3) + skew(v)
    # X_i <- X_i D:  E' = Z^-1 D^-1 B  ->  t' = R_z^T ((I - 2 [dq]x) t_b - dt) - R_z^T t_z ;  q' = q_z^-1 (x) (-dq, 1) (x) q_b
    Ji[0:3, 0:3] = -Rz.T
    Ji[0:3, 3:6] = Rz.T @ (2.0 * skew(tb))
    qz = quat_from_R(Rz); qzi = np.array([-qz[0], -qz[1], -qz[2], qz[3]])
    qb = quat_from_R(Rb)
    # keep the sign convention of the error's quaternion (w >= 0): q_e = s * (q_z^-1 (x) q_b)
    s = 1.0 if quat_mul(qzi, qb) @ q > 0 else -1.0
    for k in range(3):
        ek = np.zeros(4); ek[k] = -1.0
        Ji[3:6, 3 + k] = s * quat_mul(quat_mul(qzi, ek), qb)[:3]
    return e, Ji, Jj


def huber(e2, delta=HUBER_DELTA):
    """g2o RobustKernelHuber::robustify: (rho, rho')."""
    d2 = delta * delta
    if e2 <= d2:
        return e2, 1.0
    s = np.sqrt(e2)
    return 2.0 * s * delta - d2, delta / s


# ---------------------------------------------------------------- graph, chi2, normal equations
def poses_to_Rt(T_wc):
    T = np.asarray(T_wc, np.float64).reshape(-1, 4, 4)
    return T[:, :3, :3].copy(), T[:, :3, 3].copy()


def chi2(R, t, edges, Zr, Zt, info=INFORMATION, delta=HUBER_DELTA):
    tot = 0.0
    for k, (i, j) in enumerate(edges):
        e, _ = edge_error(R[i], t[i], R[j], t[j], Zr[k], Zt[k])
        tot += huber(e @ info @ e, delta)[0]
    return tot


def build_system(R, t, edges, Zr, Zt, fixed, info=INFORMATION, delta=HUBER_DELTA):
    """Dense H (6n x 6n over ALL vertices; fixed ones get identity rows) and b, robustified as g2o does."""
    n = R.shape[0]
    H = np.zeros((6 * n, 6 * n)); b = np.zeros(6 * n); tot = 0.0
    for k, (i, j) in enumerate(edges):
        e, Ji, Jj = edge_jacobians(R[i], t[i], R[j], t[j], Zr[k], Zt[k])
        rho, rho1 = huber(e @ info @ e, delta)
        tot += rho
        W = rho1 * info
        for (a, Ja) in ((i, Ji), (j, Jj)):
            if fixed[a]:
                continue
            b[6 * a:6 * a + 6] -= Ja.T @ W @ e
            for (c, Jc) in ((i, Ji), (j, Jj)):
                if fixed[c]:
                    continue
                H[6 * a:6 * a + 6, 6 * c:6 * c + 6] += Ja.T @ W @ Jc
    return H, b, tot


def optimize(T_wc, edges, Z, fixed=None, iterations=10, info=INFORMATION, verbose=False, delta=HUBER_DELTA):
    """g2o SparseOptimizer::optimize(iterations) with OptimizationAlgorithmLevenberg restated.
    T_wc [n,4,4]; edges [(i, j)]; Z [m,4,4] measured X_i^-1 X_j.  Returns (T_wc optimised [n,4,4] f64, summary)."""
    R, t = poses_to_Rt(T_wc)
    Zr, Zt = poses_to_Rt(Z)
    n = R.shape[0]
    fixed = np.zeros(n, bool) if fixed is None else np.asarray(fixed, bool)
    free = np.repeat(~fixed, 6)
    lam, ni = None, 2.0
    cur = chi2(R, t, edges, Zr, Zt, info, delta)
    summary = dict(initial_chi2=cur, iterations=0, trials=0)
    for it in range(iterations):
        H, b, cur = build_system(R, t, edges, Zr, Zt, fixed, info, delta)
        if lam is None:
            lam = 1e-5 * max(np.diag(H)[free].max(), 0.0) if free.any() else 0.0
        rho, qmax = 0.0, 0
        while True:
            A = H[np.ix_(free, free)] + lam * np.eye(int(free.sum()))
            try:
                dx_f = np.linalg.solve(A, b[free]); ok = np.all(np.isfinite(dx_f))
            except np.linalg.LinAlgError:
                dx_f = np.zeros(int(free.sum())); ok = False
            dx = np.zeros(6 * n); dx[free] = dx_f
            Rn, tn = R.copy(), t.copy()
            for a in range(n):
                if fixed[a]:
                    continue
                dR, dt = from_vector_mqt(dx[6 * a:6 * a + 6])
                tn[a] = t[a] + R[a] @ dt; Rn[a] = R[a] @ dR
            tmp = chi2(Rn, tn, edges, Zr, Zt, info, delta) if ok else np.inf
            scale = dx @ (lam * dx + b) + 1e-3
            rho = (cur - tmp) / scale
            summary["trials"] += 1
            if rho > 0 and np.isfinite(tmp):
                alpha = min(1.0 - (2.0 * rho - 1.0) ** 3, 2.0 / 3.0)
                lam *= max(1.0 / 3.0, alpha); ni = 2.0
                R, t, cur = Rn, tn, tmp
            else:
                lam *= ni; ni *= 2.0
                if not np.isfinite(lam):
                    break
            qmax += 1
            if not (rho < 0 and qmax < 10):
                break
        summary["iterations"] = it + 1
        if verbose:
            print("[pgo oracle] it %d chi2 %.9e lambda %.3e trials %d" % (it + 1, cur, lam, qmax))
        if qmax == 10 or rho == 0 or not np.isfinite(lam):
            break
    out = np.zeros((n, 4, 4)); out[:, 3, 3] = 1
    out[:, :3, :3] = R; out[:, :3, 3] = t
    summary.update(final_chi2=cur, final_lambda=lam)
    return out, summary
