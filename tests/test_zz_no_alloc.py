"""sba_ransac_stereo_noalloc — the allocation-free variant of the stereo RANSAC alignment (no counterpart in the
reference).  Its scratch is a fixed file-scope device array shared by all
calls, inputs and outputs are device pointers of the caller.  Checked against the existing entry point
(sba_ransac_stereo, host buffers) through the C-ABI: same kernels, same arithmetic, so the comparison is bit-exact —
on two calls in a row, the second one smaller and with an empty problem, so stale scratch contents would show.

This file is named to be collected last."""
import ctypes as C

import numpy as np
import pytest

from stereo_orb_slam_b200 import SbaError, Solver, synthetic, _capi
from stereo_orb_slam_b200.ransac import RansacOption, StereoRansac

pytestmark = pytest.mark.gpu
F = np.float32
MAX_ELEMENTS, MAX_HYPOTHESES = 65536, 16384          # SBA_NOALLOC_MAX_* (include/sba.h)


def _problems(sizes, seed0):
    probs = [synthetic.make_stereo_pairs(s, seed0 + i) if s > 0 else None for i, s in enumerate(sizes)]
    ref = next(p for p in probs if p is not None)
    firsts = [p[0][:, 0] if p else np.zeros((0, 4), F) for p in probs]
    seconds = [p[0][:, 1] if p else np.zeros((0, 4), F) for p in probs]
    rng = np.random.default_rng(seed0)
    samples = [rng.integers(0, max(s, 1), (100, 3)).astype(np.int32) for s in sizes]
    return firsts, seconds, samples, ref[3], ref[4]


def _call_noalloc(torch, sv, firsts, seconds, samples, P_l, P_r, opt, want_counts=True):
    dev = torch.device("cuda", 0)
    sizes = [f.shape[0] for f in firsts]
    nb, n = len(sizes), int(sum(sizes))
    off = np.zeros(nb + 1, np.int64); off[1:] = np.cumsum(sizes)
    cat = lambda xs, shape, dt: np.ascontiguousarray(np.concatenate(xs).astype(dt)) if n or shape is None else np.zeros(shape, dt)
    t_off = torch.from_numpy(off).to(dev)
    t_a = torch.from_numpy(cat(firsts, (0, 4), F)).to(dev); t_b = torch.from_numpy(cat(seconds, (0, 4), F)).to(dev)
    t_s = torch.from_numpy(np.ascontiguousarray(np.stack(samples))).to(dev)
    # outputs pre-filled with garbage: everything the entry point promises must be overwritten
    t_T = torch.full((nb, 16), 7.0, dtype=torch.float32, device=dev)
    t_lab = torch.full((max(n, 1),), 9, dtype=torch.uint8, device=dev)
    t_loss = torch.full((max(n, 1),), -1.0, dtype=torch.float32, device=dev)
    t_best = torch.full((nb,), -5, dtype=torch.int32, device=dev)
    t_cnt = torch.full((nb, opt.max_iterations), -7, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()                              # torch's stream is not the handle's
    st = opt.as_struct()
    p = lambda t: C.c_void_p(t.data_ptr())
    Pl = np.ascontiguousarray(P_l, F).reshape(12); Pr = np.ascontiguousarray(P_r, F).reshape(12)
    rc = sv._L.sba_ransac_stereo_noalloc(sv._h, C.c_int32(nb), C.c_int64(n), p(t_off), p(t_a), p(t_b),
                                         Pl.ctypes.data_as(C.c_void_p), Pr.ctypes.data_as(C.c_void_p), C.byref(st), p(t_s), C.c_uint64(0),
                                         p(t_T), p(t_lab), p(t_loss), p(t_best) if want_counts else None,
                                         p(t_cnt) if want_counts else None, None)
    _capi.check(rc)                                       # stream NULL: the call has waited for completion
    sl = [slice(int(off[i]), int(off[i + 1])) for i in range(nb)]
    lab, loss = t_lab.cpu().numpy(), t_loss.cpu().numpy()
    return (t_T.cpu().numpy().reshape(nb, 4, 4), [lab[s].astype(bool) for s in sl], [loss[s] for s in sl],
            t_best.cpu().numpy(), t_cnt.cpu().numpy())


def _same(got, want, info, counts=True):
    T, labels, losses, best, cnt = got
    Tw, lw, sw = want
    assert np.array_equal(T, Tw)
    for a, b in zip(labels, lw):
        assert np.array_equal(a, b)
    for a, b in zip(losses, sw):
        assert np.array_equal(a, b)
    if counts:
        assert np.array_equal(best, np.asarray(info["best_iteration"]))
        assert np.array_equal(cnt, np.stack(info["counts"]))


def test_noalloc_entry_point_equals_the_allocating_one_on_two_calls_in_a_row():
    torch = pytest.importorskip("torch")
    sv = Solver(0)
    for final_fit in (False, True):
        opt = RansacOption(final_model_fitting=final_fit)
        # call 1: a batch that fills a good part of the scratch; call 2: fewer, smaller problems and an empty one
        for k, sizes in enumerate(([700, 64, 1500, 10, 333], [40, 0, 250])):
            firsts, seconds, samples, P_l, P_r = _problems(sizes, 900 + 10 * k)
            ref = StereoRansac(P_l, P_r, solver=sv); ref.SetOptions(opt)
            Tw, lw, sw, info = ref.solve_batch(firsts, seconds, samples)
            got = _call_noalloc(torch, sv, firsts, seconds, samples, P_l, P_r, opt)
            _same(got, (Tw, lw, sw), info)
            if 0 in sizes:
                e = sizes.index(0)
                assert np.all(got[0][e][:3] == 0)          # a problem without elements: T = 0, not the last call's model
            # counts / chosen iteration kept in the library's own scratch when the caller passes NULL
            got2 = _call_noalloc(torch, sv, firsts, seconds, samples, P_l, P_r, opt, want_counts=False)
            _same(got2, (Tw, lw, sw), info, counts=False)
    sv.close()


def test_noalloc_entry_point_rejects_what_does_not_fit_the_fixed_scratch():
    torch = pytest.importorskip("torch")
    sv = Solver(0)
    dev = torch.device("cuda", 0)
    buf = torch.zeros(64, dtype=torch.float32, device=dev)
    P = np.zeros(12, F)
    p = C.c_void_p(buf.data_ptr())
    Pp = P.ctypes.data_as(C.c_void_p)

    def call(nb, n, iters):
        st = RansacOption(max_iterations=iters).as_struct()
        return sv._L.sba_ransac_stereo_noalloc(sv._h, C.c_int32(nb), C.c_int64(n), p, p, p, Pp, Pp, C.byref(st), None, C.c_uint64(1),
                                               p, p, p, None, None, None)
    for nb, n, iters, word in ((1, MAX_ELEMENTS + 1, 100, "elements"), (MAX_HYPOTHESES // 100 + 1, 10, 100, "hypotheses")):
        rc = call(nb, n, iters)
        assert rc == _capi.SBA_INVALID_ARGUMENT
        with pytest.raises(SbaError) as ei:
            _capi.check(rc)
        assert word in str(ei.value) and "exceed the fixed scratch" in str(ei.value)
    # misaligned observation pointer
    st = RansacOption().as_struct()
    rc = sv._L.sba_ransac_stereo_noalloc(sv._h, C.c_int32(1), C.c_int64(4), p, C.c_void_p(buf.data_ptr() + 4), p, Pp, Pp, C.byref(st), None,
                                         C.c_uint64(1), p, p, p, None, None, None)
    assert rc == _capi.SBA_INVALID_ARGUMENT
    sv.close()
