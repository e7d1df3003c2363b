"""Handle-level wrapper over the C-ABI (include/sba.h): one Solver = one sba_handle = one GPU.

Everything numeric happens inside libsba.so on the device; this file only marshals numpy arrays.
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from ._capi import f32, f64, i32, ptr, u8

D, I32, I64, F, U8 = C.c_double, C.c_int32, C.c_int64, C.c_float, C.c_uint8


class Solver:
    """Device-resident stereo BA problem.  device=-1 builds structure on the host only (no compute)."""

    def __init__(self, device=0):
        self._L = capi.lib()
        self._h = C.c_void_p()
        capi.check(self._L.sba_create(int(device), C.byref(self._h)))
        self.device = device
        self.Nc = self.Np = self.Nobs = self.nwin = 0

    def close(self):
        if self._h:
            self._L.sba_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- multi-GPU plumbing
    @staticmethod
    def comm_unique_id():
        buf = (C.c_uint8 * 128)()
        capi.check(capi.lib().sba_comm_unique_id(buf))
        return bytes(buf)

    def comm_init(self, rank, world_size, uid):
        buf = (C.c_uint8 * 128).from_buffer_copy(uid)
        capi.check(self._L.sba_comm_init(self._h, int(rank), int(world_size), buf))

    def set_point_shard(self, begin, end):
        capi.check(self._L.sba_set_point_shard(self._h, int(begin), int(end)))

    # ---- problem
    def set_structure(self, num_cameras, num_points, obs_cam, obs_pt, cam_fixed=None, win_cam_off=None,
                      win_pt_off=None):
        obs_cam, obs_pt = i32(obs_cam), i32(obs_pt)
        if obs_cam.shape != obs_pt.shape:
            raise ValueError("obs_cam / obs_pt length mismatch")
        cam_fixed = u8(cam_fixed)
        nwin = 1
        if win_cam_off is not None:
            win_cam_off, win_pt_off = i32(win_cam_off), i32(win_pt_off)
            nwin = win_cam_off.size - 1
        capi.check(self._L.sba_set_structure(self._h, I32(num_cameras), I32(num_points), I64(obs_cam.size),
                                             ptr(obs_cam, I32), ptr(obs_pt, I32), ptr(cam_fixed, U8), I32(nwin),
                                             ptr(win_cam_off, I32), ptr(win_pt_off, I32)))
        self.Nc, self.Np, self.Nobs, self.nwin = int(num_cameras), int(num_points), int(obs_cam.size), nwin
        return self

    def set_state(self, poses, points):
        poses, points = f64(poses), f64(points)
        if poses.size != 6 * self.Nc or points.size != 3 * self.Np:
            raise ValueError("state size does not match the structure")
        capi.check(self._L.sba_set_state(self._h, ptr(poses, D), ptr(points, D)))

    def set_measurements(self, obs_uv, p_l, p_r):
        obs_uv, p_l, p_r = f32(obs_uv), f64(p_l), f64(p_r)
        if obs_uv.size != 4 * self.Nobs or p_l.size != 12 or p_r.size != 12:
            raise ValueError("measurement size does not match the structure")
        capi.check(self._L.sba_set_measurements(self._h, ptr(obs_uv, F), ptr(p_l, D), ptr(p_r, D)))

    def get_state(self):
        poses = np.zeros((self.Nc, 6)); points = np.zeros((self.Np, 3))
        capi.check(self._L.sba_get_state(self._h, ptr(poses, D), ptr(points, D)))
        return poses, points

    def run(self, options=None):
        s = capi.Summary()
        o = options if options is not None else capi.default_options()
        rc = self._L.sba_run(self._h, C.byref(o), C.byref(s))
        capi.check(rc, ok=(capi.SBA_OK, capi.SBA_NO_CONVERGENCE, capi.SBA_NUMERICAL_FAILURE))
        out = s.as_dict(); out["status"] = rc
        return out

    def solve(self, poses, points, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed=None, win_cam_off=None,
              win_pt_off=None, options=None):
        """One-shot sba_solve on host buffers.  Returns (poses, points, summary); inputs are not modified."""
        poses = f64(poses).copy(); points = f64(points).copy()
        obs_cam, obs_pt, obs_uv, p_l, p_r = i32(obs_cam), i32(obs_pt), f32(obs_uv), f64(p_l), f64(p_r)
        cam_fixed = u8(cam_fixed)
        nwin = 1
        if win_cam_off is not None:
            win_cam_off, win_pt_off = i32(win_cam_off), i32(win_pt_off)
            nwin = win_cam_off.size - 1
        Nc, Np = poses.size // 6, points.size // 3
        o = options if options is not None else capi.default_options()
        s = capi.Summary()
        rc = self._L.sba_solve(self._h, I32(Nc), I32(Np), I64(obs_cam.size), ptr(poses, D), ptr(points, D),
                               ptr(obs_cam, I32), ptr(obs_pt, I32), ptr(obs_uv, F), ptr(p_l, D), ptr(p_r, D),
                               ptr(cam_fixed, U8), I32(nwin), ptr(win_cam_off, I32), ptr(win_pt_off, I32),
                               C.byref(o), C.byref(s))
        # like the reference, a numerical failure still returns the last accepted state (summary["status"] says so)
        capi.check(rc, ok=(capi.SBA_OK, capi.SBA_NO_CONVERGENCE, capi.SBA_NUMERICAL_FAILURE))
        self.Nc, self.Np, self.Nobs, self.nwin = Nc, Np, int(obs_cam.size), nwin
        out = s.as_dict(); out["status"] = rc
        return poses.reshape(-1, 6), points.reshape(-1, 3), out

    # ---- stages (parity tests + bench)
    def linearize(self, options=None):
        o = options if options is not None else capi.default_options()
        capi.check(self._L.sba_linearize(self._h, C.byref(o)))

    def schur(self, radius, options=None):
        o = options if options is not None else capi.default_options()
        capi.check(self._L.sba_schur(self._h, C.byref(o), D(radius)))

    def pcg(self, options=None):
        o = options if options is not None else capi.default_options()
        it = I32(0)
        capi.check(self._L.sba_pcg(self._h, C.byref(o), C.byref(it)))
        return it.value

    def time_linearize_schur(self, radius=1e4, repeats=10, options=None):
        o = options if options is not None else capi.default_options()
        a, b = D(0), D(0)
        capi.check(self._L.sba_time_linearize_schur(self._h, C.byref(o), D(radius), I32(repeats), C.byref(a), C.byref(b)))
        return a.value, b.value

    def time_stage(self, stage, repeats=10, flush_l2=True, radius=0.0, options=None):
        o = options if options is not None else capi.default_options()
        ms = D(0)
        capi.check(self._L.sba_time_stage(self._h, C.byref(o), D(radius), I32(stage), I32(repeats),
                                          I32(1 if flush_l2 else 0), C.byref(ms)))
        return ms.value

    def measure_fp64_peak(self):
        """FP64 FMA peak of the device in TFLOP/s (own micro-benchmark, register-only DFMA chains)."""
        t = D(0)
        capi.check(self._L.sba_measure_fp64_peak(self._h, C.byref(t)))
        return t.value

    # ---- introspection
    def get_linearization(self, residuals=True):
        r = np.zeros((self.Nobs, 4)) if residuals else None
        out = dict(U=np.zeros((self.Nc, 6, 6)), V=np.zeros((self.Np, 3, 3)), W=np.zeros((self.Nobs, 6, 3)),
                   gc=np.zeros((self.Nc, 6)), gp=np.zeros((self.Np, 3)), cost=np.zeros(self.nwin))
        capi.check(self._L.sba_get_linearization(self._h, ptr(r, D), ptr(out["U"], D), ptr(out["V"], D),
                                                 ptr(out["W"], D), ptr(out["gc"], D), ptr(out["gp"], D),
                                                 ptr(out["cost"], D)))
        if residuals:
            out["r"] = r
        return out

    def schur_pattern(self):
        n = self._L.sba_get_schur_pattern(self._h, None, None, 0)
        if n < 0:
            raise RuntimeError("no structure")
        rows = np.zeros(n, np.int32); cols = np.zeros(n, np.int32)
        self._L.sba_get_schur_pattern(self._h, ptr(rows, I32), ptr(cols, I32), n)
        return rows, cols

    def schur_values(self):
        n = self._L.sba_get_schur_pattern(self._h, None, None, 0)
        nf = self.num_free_cameras
        blocks = np.zeros((n, 6, 6)); rhs = np.zeros(6 * nf); x = np.zeros(6 * nf)
        capi.check(self._L.sba_get_schur_values(self._h, ptr(blocks, D), ptr(rhs, D), ptr(x, D)))
        return blocks, rhs, x

    def schur_dense(self):
        """Reduced camera matrix assembled densely on the host from the block values (tests only)."""
        rows, cols = self.schur_pattern()
        blocks, rhs, x = self.schur_values()
        n = 6 * self.num_free_cameras
        S = np.zeros((n, n))
        for k in range(rows.size):
            S[6 * rows[k]:6 * rows[k] + 6, 6 * cols[k]:6 * cols[k] + 6] = blocks[k]
        return S, rhs, x

    def sorted_views(self):
        bp = np.zeros(self.Nobs, np.int32); bc = np.zeros(self.Nobs, np.int32)
        capi.check(self._L.sba_get_sorted_views(self._h, ptr(bp, I32), ptr(bc, I32)))
        return bp, bc

    def factor_symbolic(self):
        """Symbolic phase of the block-sparse Cholesky (structures with a window of > 64 free cameras)."""
        info = (C.c_int64 * 5)()
        capi.check(self._L.sba_get_factor_symbolic(self._h, info, None, None, None, None, None, None, None, None, None, None))
        n, nnz, nupd, nlev, nchain = (int(v) for v in info)
        out = dict(n=n, nnzL=nnz, nupd=nupd, nlev=nlev, nchain=nchain, perm=np.zeros(n, np.int32), col_ptr=np.zeros(n + 1, np.int32),
                   row_idx=np.zeros(nnz, np.int32), init_src=np.zeros(nnz, np.int32), upd_ptr=np.zeros(nnz + 1, np.int64),
                   upd_a=np.zeros(nupd, np.int32), upd_b=np.zeros(nupd, np.int32), lvl_ptr=np.zeros(nlev + 1, np.int32),
                   lvl_cols=np.zeros(nchain, np.int32), lvl_len=np.zeros(nchain, np.int32))
        capi.check(self._L.sba_get_factor_symbolic(
            self._h, info, ptr(out["perm"], I32), ptr(out["col_ptr"], I32), ptr(out["row_idx"], I32), ptr(out["init_src"], I32),
            ptr(out["upd_ptr"], I64), ptr(out["upd_a"], I32), ptr(out["upd_b"], I32), ptr(out["lvl_ptr"], I32),
            ptr(out["lvl_cols"], I32), ptr(out["lvl_len"], I32)))
        return out

    @property
    def num_free_cameras(self):
        return int(self._L.sba_num_free_cameras(self._h))

    @property
    def fused_kernel(self):
        """Name of the kernel that runs K1-K3 of the chunks for the current structure (None: the two-kernel path)."""
        return {0: None, 1: "k_fused_chunk", 2: "k_fused_chunk", 3: "k_fused_persist"}[int(self._L.sba_fused_kernel_variant(self._h))]

    @property
    def launch_count(self):
        return int(self._L.sba_launch_count(self._h))
