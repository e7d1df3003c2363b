"""Generate known-answer vectors for the stereo reprojection residual and its Jacobians.

The reference holds no golden vectors for this path, so these are produced by an
independent 50-digit mpmath restatement of ReprojectionError::operator()
(reference: src/reprojection_error.h:12-41) with the exact Rodrigues formula; Jacobians by
mpmath's high-order numerical differentiation (mp.diff) at 50 digits — no code shared with oracle/ or
csrc/.  Run:  python tests/golden/make_kats.py  ->  tests/golden/reprojection_kats.json
"""
import json
import os

import mpmath as mp
import numpy as np

mp.mp.dps = 50


def rot(w):
    th = mp.sqrt(w[0] ** 2 + w[1] ** 2 + w[2] ** 2)
    K = mp.matrix([[0, -w[2], w[1]], [w[2], 0, -w[0]], [-w[1], w[0], 0]])
    if th == 0:
        return mp.eye(3)
    return mp.eye(3) + (mp.sin(th) / th) * K + ((1 - mp.cos(th)) / th ** 2) * (K * K)


def residual(params, uv, pl, pr):
    w, t, X = params[0:3], params[3:6], params[6:9]
    R = rot(w)
    p = [sum(R[i, j] * X[j] for j in range(3)) + t[i] for i in range(3)]
    out = []
    for P, (u, v) in ((pl, uv[0:2]), (pr, uv[2:4])):
        s = 1 / (P[8] * p[0] + P[9] * p[1] + P[10] * p[2] + P[11])
        out.append((P[0] * p[0] + P[1] * p[1] + P[2] * p[2] + P[3]) * s - u)
        out.append((P[4] * p[0] + P[5] * p[1] + P[6] * p[2] + P[7]) * s - v)
    return out


def jacobian(params, uv, pl, pr):
    J = [[None] * 9 for _ in range(4)]
    for k in range(4):
        for i in range(9):
            def f(x, k=k, i=i):
                q = list(params); q[i] = x
                return residual(q, uv, pl, pr)[k]
            J[k][i] = mp.diff(f, params[i])
    return J


def main():
    rng = np.random.default_rng(20241104)
    f32 = np.float32
    fx, cx, cy, b = f32(718.856), f32(607.1928), f32(185.2157), f32(0.53716)
    pl = [float(fx), 0.0, float(cx), 0.0, 0.0, float(fx), float(cy), 0.0, 0.0, 0.0, 1.0, 0.0]
    pr = list(pl); pr[3] = float(f32(-fx * b))
    # a general (non-rectified) pair as well: the reference does not specialise
    pl_g = [700.0, 1.5, 600.0, 3.0, -0.7, 705.0, 190.0, -2.0, 1e-3, -2e-3, 1.0, 0.01]
    pr_g = [698.0, -1.0, 610.0, -380.0, 0.4, 702.0, 188.0, 1.0, -1e-3, 1e-3, 0.999, -0.02]
    cases = []
    thetas = [0.0, 1e-9, 1e-7, 1e-4, 2.4e-3, 1e-2, 0.37, 1.0, float(mp.pi) - 1e-3]
    for ci, th in enumerate(thetas):
        for proj, (Pl, Pr) in (("rectified", (pl, pr)), ("general", (pl_g, pr_g))):
            axis = rng.normal(size=3); axis /= np.linalg.norm(axis)
            w = (axis * th).tolist()
            t = rng.normal(0, 0.5, 3).tolist()
            X = [float(rng.uniform(-10, 10)), float(rng.uniform(-3, 3)), float(rng.uniform(5, 60))]
            # put the measurement near the true projection (plus a few px), stored as f32 like the reference
            params = [mp.mpf(v) for v in (w + t + X)]
            r0 = residual(params, [0, 0, 0, 0], [mp.mpf(v) for v in Pl], [mp.mpf(v) for v in Pr])
            uv = [float(f32(float(r0[k]) + rng.normal(0, 2.0))) for k in range(4)]
            mpl, mpr, muv = [mp.mpf(v) for v in Pl], [mp.mpf(v) for v in Pr], [mp.mpf(v) for v in uv]
            r = residual(params, muv, mpl, mpr)
            J = jacobian(params, muv, mpl, mpr)
            cases.append(dict(name="theta_%d_%s" % (ci, proj), theta=th, cam=w + t, point=X, uv=uv, p_l=Pl, p_r=Pr,
                              r=[float(v) for v in r],
                              Jc=[[float(J[k][i]) for i in range(6)] for k in range(4)],
                              Jp=[[float(J[k][6 + i]) for i in range(3)] for k in range(4)]))
    # Huber rho / sqrt(rho') (ceres::HuberLoss(1.0) published formula), incl. the switch point
    hub = []
    for s in [0.0, 0.25, 1.0, 1.0000001, 2.0, 100.0, 1e6]:
        ms = mp.mpf(s)
        if ms <= 1:
            rho, r1 = ms, mp.mpf(1)
        else:
            rho, r1 = 2 * mp.sqrt(ms) - 1, 1 / mp.sqrt(ms)
        hub.append(dict(s=s, rho=float(rho), sqrt_rho1=float(mp.sqrt(r1))))
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reprojection_kats.json")
    with open(out, "w") as f:
        json.dump(dict(generator="tests/golden/make_kats.py (mpmath, 50 digits)", cases=cases, huber=hub), f, indent=1)
    print("wrote", out, len(cases), "cases")


if __name__ == "__main__":
    main()
