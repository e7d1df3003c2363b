"""Stereo RANSAC rigid alignment on the GPU — host-side mirror of the reference's
RANSAC::Solver<ObservationPair, Matrix4f> over StereoTransModel (reference: ransac_solver/include/ransac_solver.h:51-166,
ransac_solver/include/stereo_reprojection.h:28-134) and of VisualOdometer::CalcTransformation
(src/visual_odometer.cpp:93-103), the call that gives bundle adjustment its initial relative poses and inlier masks.

    solver = StereoRansac(P_l, P_r)                      # StereoTransModel::SetCameraParams + Solver(&model)
    solver.SetOptions(RansacOption())                    # src/params.h:18-27: 100 iterations, no final fit, no early exit
    T, labels, losses = solver.Solve(obs_pairs)          # obs_pairs [n, 2, 4] = vector<ObservationPair>
    ok, T, labels, losses = CalcTransformation(solver, obs_pairs)

The three element draws of every iteration come from `rand() % size` in the reference (ransac_solver.h:84-90).  They
are an input of the C-ABI; this module makes them with a restatement of glibc's rand() (GlibcRand, pinned against
libc in tests/test_ransac.py), so a process that seeds like the reference's (never: glibc starts at srand(1)) draws
the same elements.  The computation runs in libsba.so; there is no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from ._capi import ptr
from .solver import Solver

F32, I32, I64, U8 = C.c_float, C.c_int32, C.c_int64, C.c_uint8
MIN_PAIRS = 10                      # src/visual_odometer.cpp:100


class GlibcRand:
    """glibc's rand() (TYPE_3 additive feedback generator, r[i] = r[i-3] + r[i-31], output >> 1) — the generator behind
    the reference's `rand() % size`.  srand is never called there, so its stream is the one of seed 1."""

    def __init__(self, seed=1):
        self.srand(seed)

    def srand(self, seed):
        seed = int(seed) & 0xFFFFFFFF
        if seed == 0:
            seed = 1
        r = [0] * 34
        r[0] = seed
        word = seed if seed < 0x80000000 else seed - (1 << 32)           # glibc keeps it in an int32_t
        for i in range(1, 31):
            hi, lo = int(word / 127773), 0                               # C division truncates toward zero
            lo = word - hi * 127773
            word = 16807 * lo - 2836 * hi
            if word < 0:
                word += 2147483647
            r[i] = word & 0xFFFFFFFF
        for i in range(31, 34):
            r[i] = r[i - 31]
        self._r = r
        for _ in range(310):
            self._next()

    def _next(self):
        r = self._r
        v = (r[-31] + r[-3]) & 0xFFFFFFFF
        r.append(v)
        if len(r) > 64:
            del r[:len(r) - 34]
        return v

    def rand(self):
        return self._next() >> 1

    def draws(self, iterations, size):
        """The [iterations, 3] element indices Solver::Solve draws for a problem of `size` elements."""
        out = np.empty((iterations, 3), np.int32)
        for i in range(iterations):
            for k in range(3):
                out[i, k] = self.rand() % size
        return out


class RansacOption:
    """RANSAC::Option (ransac_solver.h:13-19); the defaults are what VisualOdometer sets from src/params.h:18-27."""

    def __init__(self, max_iterations=100, final_model_fitting=False, early_termination=False, concensus_ratio=0.8,
                 inlier_threshold=3.0):
        self.max_iterations = int(max_iterations)
        self.final_model_fitting = bool(final_model_fitting)
        self.early_termination = bool(early_termination)
        self.concensus_ratio = float(concensus_ratio)
        self.inlier_threshold = float(inlier_threshold)

    def as_struct(self):
        return capi.RansacOptions(self.max_iterations, int(self.final_model_fitting), int(self.early_termination),
                                  self.concensus_ratio, self.inlier_threshold)


def _pairs(obs_pairs):
    a = np.asarray(obs_pairs, np.float32)
    if a.ndim != 3 or a.shape[1:] != (2, 4):
        raise ValueError("obs_pairs must be [n, 2, 4]: (first, second) x (u_l, v_l, u_r, v_r)")
    return np.ascontiguousarray(a[:, 0, :]), np.ascontiguousarray(a[:, 1, :])


class StereoRansac:
    """Solver<ObservationPair, Matrix4f> bound to a StereoTransModel with the given 3x4 projection matrices."""

    def __init__(self, P_l, P_r, device=0, solver=None, rng=None):
        self.P_l = np.ascontiguousarray(np.asarray(P_l, np.float32).reshape(3, 4))
        self.P_r = np.ascontiguousarray(np.asarray(P_r, np.float32).reshape(3, 4))
        self._sv = solver if solver is not None else Solver(device)
        self.options = RansacOption()
        self.rng = rng if rng is not None else GlibcRand(1)
        self.last = None

    def SetOptions(self, options):
        self.options = options

    def Solve(self, obs_pairs, weights=None, samples=None):
        """-> (T [4,4] f32, labels [n] bool, losses [n] f32).  `weights` is accepted and ignored, as the reference's
        model ignores it.  samples [max_iterations, 3] overrides the generator's draws."""
        a, b = _pairs(obs_pairs)
        T, labels, losses, info = self.solve_batch([a], [b], None if samples is None else [samples])
        self.last = {k: v[0] for k, v in info.items()}
        return T[0], labels[0], losses[0]

    def solve_batch(self, firsts, seconds, samples=None, seed=None):
        """Independent problems in one call.  firsts / seconds: lists of [n_b, 4] arrays.  samples: list of
        [max_iterations, 3] draws (default: this solver's generator, problem after problem, as consecutive Solve calls
        would consume it); seed: use the library's counter-based draw instead."""
        o = self.options
        nb = len(firsts)
        sizes = [int(np.asarray(x).reshape(-1, 4).shape[0]) for x in firsts]
        off = np.zeros(nb + 1, np.int64)
        off[1:] = np.cumsum(sizes)
        n = int(off[-1])
        A = np.ascontiguousarray(np.concatenate([np.asarray(x, np.float32).reshape(-1, 4) for x in firsts]) if nb else np.zeros((0, 4), np.float32))
        B = np.ascontiguousarray(np.concatenate([np.asarray(x, np.float32).reshape(-1, 4) for x in seconds]) if nb else np.zeros((0, 4), np.float32))
        if B.shape != A.shape:
            raise ValueError("first / second sizes differ")
        S = None
        if seed is None:
            if samples is None:
                samples = [self.rng.draws(o.max_iterations, sz) if sz > 0 else np.zeros((o.max_iterations, 3), np.int32) for sz in sizes]
            S = np.ascontiguousarray(np.stack([np.asarray(s, np.int32).reshape(-1, 3)[:o.max_iterations] for s in samples]) if nb else np.zeros((0, o.max_iterations, 3), np.int32))
            if S.shape != (nb, o.max_iterations, 3):
                raise ValueError("samples must be [max_iterations, 3] per problem")
        T = np.zeros((nb, 4, 4), np.float32)
        labels = np.zeros(n, np.uint8)
        losses = np.zeros(n, np.float32)
        best = np.zeros(nb, np.int32)
        counts = np.zeros((nb, max(o.max_iterations, 1)), np.int32)
        st = o.as_struct()
        sv = self._sv
        rc = sv._L.sba_ransac_stereo(sv._h, I32(nb), ptr(off, I64), ptr(A, F32), ptr(B, F32), ptr(self.P_l, F32), ptr(self.P_r, F32),
                                     C.byref(st), ptr(S, I32), C.c_uint64(0 if seed is None else int(seed)), ptr(T, F32), ptr(labels, U8),
                                     ptr(losses, F32), ptr(best, I32), ptr(counts, I32))
        capi.check(rc)
        sl = [slice(int(off[i]), int(off[i + 1])) for i in range(nb)]
        info = dict(best_iteration=list(best), counts=[counts[i, :o.max_iterations] for i in range(nb)], samples=[None if S is None else S[i] for i in range(nb)])
        return T, [labels[s].astype(bool) for s in sl], [losses[s] for s in sl], info


def CalcTransformation(solver, obs_pairs, weights=None, samples=None):
    """VisualOdometer::CalcTransformation (src/visual_odometer.cpp:93-103): (success, T, mask, losses); fewer than 10
    pairs -> (False, None, None, None) without touching the generator, as in the reference."""
    if np.asarray(obs_pairs).shape[0] < MIN_PAIRS:
        return False, None, None, None
    T, labels, losses = solver.Solve(obs_pairs, weights, samples)
    return True, T, labels, losses


def triangulate_stereo(obs, P_l, P_r, device=0, solver=None):
    """cv::triangulatePoints + division by w for [n,4] stereo observations (stereo_reprojection.h:79-92) -> [n,3] f32."""
    o = np.ascontiguousarray(np.asarray(obs, np.float32).reshape(-1, 4))
    Pl = np.ascontiguousarray(np.asarray(P_l, np.float32).reshape(12))
    Pr = np.ascontiguousarray(np.asarray(P_r, np.float32).reshape(12))
    out = np.zeros((o.shape[0], 3), np.float32)
    sv = solver if solver is not None else Solver(device)
    capi.check(sv._L.sba_triangulate_stereo(sv._h, I64(o.shape[0]), ptr(o, F32), ptr(Pl, F32), ptr(Pr, F32), ptr(out, F32)))
    return out
