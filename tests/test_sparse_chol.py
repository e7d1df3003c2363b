"""Symbolic phase of the block-sparse Cholesky (csrc/sparse_chol.cpp) checked on the CPU: a numpy emulation of the
numeric phase that follows the SAME static lists the kernel follows (scatter map, per-block product lists, level
sets) must reproduce the reduced matrix, L L^T = P S P^T, on banded, loop-closure and batched patterns.
(reference: CHOLMOD's analyse step behind src/bundle_adjuster.cpp:26-29 has no test of its own.)"""
import numpy as np
import pytest

from stereo_orb_slam_b200 import Solver, pack_window, synthetic

INT32_MIN = -2 ** 31


def host_structure(smap, monkeypatch, force=True):
    if force:
        monkeypatch.setenv("SBA_FORCE_SPARSE_CHOL", "1")
    w = pack_window(smap, 0, smap.num_frames)
    s = Solver(-1).set_structure(w["poses"].shape[0], w["points"].shape[0], w["obs_cam"], w["obs_pt"])
    return s


def random_spd(rows, cols, nf, rng):
    """Symmetric positive definite block matrix on the given full pattern; returns the UPPER blocks in pattern order."""
    up = [(r, c) for r, c in zip(rows, cols) if r <= c]
    B = {}
    deg = np.zeros(nf)
    for r, c in up:
        if r != c:
            B[(r, c)] = rng.normal(0, 1, (6, 6)); deg[r] += 1; deg[c] += 1
    for r, c in up:
        if r == c:
            A = rng.normal(0, 1, (6, 6))
            B[(r, c)] = A @ A.T + (6.0 * (deg[r] + 1)) * 3 * np.eye(6)
    return up, B


def emulate(sym, up, B):
    """Left-looking block factorisation by levels, exactly as k_spchol schedules it."""
    nnz = sym["nnzL"]
    Lv = np.zeros((nnz, 6, 6))
    for b in range(nnz):
        src = int(sym["init_src"][b])
        if src >= 0:
            Lv[b] = B[up[src]]
        elif src != INT32_MIN:
            Lv[b] = B[up[~src]].T
    done = np.zeros(sym["n"], bool)
    col_of = np.repeat(np.arange(sym["n"]), np.diff(sym["col_ptr"]))
    chain_of = np.zeros(sym["n"], np.int64)
    for c in range(sym["nchain"]):
        chain_of[sym["lvl_cols"][c]:sym["lvl_cols"][c] + sym["lvl_len"][c]] = c
    for l in range(sym["nlev"]):
        chains = range(sym["lvl_ptr"][l], sym["lvl_ptr"][l + 1])
        cols = np.concatenate([np.arange(sym["lvl_cols"][c], sym["lvl_cols"][c] + sym["lvl_len"][c]) for c in chains])
        for j in cols:          # a chain's columns in sequence; operands come from finished levels or earlier in the chain
            b0, b1 = sym["col_ptr"][j], sym["col_ptr"][j + 1]
            for b in range(b0, b1):
                for u in range(sym["upd_ptr"][b], sym["upd_ptr"][b + 1]):
                    a, bb = sym["upd_a"][u], sym["upd_b"][u]
                    k = col_of[a]
                    assert col_of[bb] == k and (done[k] or (k < j and k in cols and chain_of[k] == chain_of[j]))
                    Lv[b] -= Lv[a] @ Lv[bb].T
            Ljj = np.linalg.cholesky(Lv[b0])
            Li = np.linalg.inv(Ljj)
            Lv[b0] = Ljj
            for b in range(b0 + 1, b1):
                Lv[b] = Lv[b] @ Li.T
        done[cols] = True
    return Lv


def dense_L(sym, Lv):
    n = sym["n"]
    L = np.zeros((6 * n, 6 * n))
    for j in range(n):
        for b in range(sym["col_ptr"][j], sym["col_ptr"][j + 1]):
            i = sym["row_idx"][b]
            L[6 * i:6 * i + 6, 6 * j:6 * j + 6] = np.tril(Lv[b]) if i == j else Lv[b]
    return L


def check(s, rng):
    sym = s.factor_symbolic()
    nf = s.num_free_cameras
    rows, cols = s.schur_pattern()
    up, B = random_spd(rows, cols, nf, rng)
    assert sym["n"] == nf and sorted(sym["perm"]) == list(range(nf))
    assert int(sym["lvl_len"].sum()) == nf and sym["lvl_ptr"][-1] == sym["nchain"]
    Lv = emulate(sym, up, B)
    L = dense_L(sym, Lv)
    S = np.zeros((6 * nf, 6 * nf))
    for (r, c), v in B.items():
        S[6 * r:6 * r + 6, 6 * c:6 * c + 6] = v
        S[6 * c:6 * c + 6, 6 * r:6 * r + 6] = v.T
    p = np.repeat(6 * sym["perm"], 6) + np.tile(np.arange(6), nf)
    PSP = S[np.ix_(p, p)]
    assert np.abs(L @ L.T - PSP).max() <= 1e-10 * np.abs(PSP).max()
    return sym


def test_banded_chain_orders_into_a_shallow_tree(monkeypatch):
    smap = synthetic.make_config("C5", scale=0.01)          # 200-frame open chain
    s = host_structure(smap, monkeypatch, force=False)      # > 64 free cameras: built without forcing
    sym = check(s, np.random.default_rng(1))
    nf = s.num_free_cameras
    assert sym["nlev"] <= 12                                 # natural order would give one chain of nf columns
    nS = (s.schur_pattern()[0] <= s.schur_pattern()[1]).sum()
    assert sym["nnzL"] <= 3 * nS                             # modest fill


def test_loop_closure_pattern(monkeypatch):
    smap = synthetic.make_config("C4", scale=0.04)          # ~180 frames, the last ones re-observe the first
    s = host_structure(smap, monkeypatch, force=False)
    rows, cols = s.schur_pattern()
    assert (np.abs(rows - cols) > s.num_free_cameras // 2).any()       # off-band (loop closure) blocks exist
    check(s, np.random.default_rng(2))


def test_small_dense_window_and_batched_windows(monkeypatch):
    smap = synthetic.make_map(num_frames=9, num_points=400, track_len=6, seed=5)
    check(host_structure(smap, monkeypatch), np.random.default_rng(3))
    # two independent windows in one structure: the factor of the block-diagonal system has no coupling blocks
    monkeypatch.setenv("SBA_FORCE_SPARSE_CHOL", "1")
    ws = [pack_window(synthetic.make_map(num_frames=7 + i, num_points=200, track_len=4, seed=50 + i), 0, 7 + i) for i in range(2)]
    cam_off = np.cumsum([0] + [w["poses"].shape[0] for w in ws]).astype(np.int32)
    pt_off = np.cumsum([0] + [w["points"].shape[0] for w in ws]).astype(np.int32)
    oc = np.concatenate([w["obs_cam"] + cam_off[i] for i, w in enumerate(ws)])
    op = np.concatenate([w["obs_pt"] + pt_off[i] for i, w in enumerate(ws)])
    s = Solver(-1).set_structure(int(cam_off[-1]), int(pt_off[-1]), oc, op, win_cam_off=cam_off, win_pt_off=pt_off)
    sym = check(s, np.random.default_rng(4))
    nf0 = ws[0]["poses"].shape[0] - 1
    first = sym["perm"] < nf0                                 # new index -> belongs to window 0?
    for j in range(sym["n"]):
        rows_j = sym["row_idx"][sym["col_ptr"][j]:sym["col_ptr"][j + 1]]
        assert (first[rows_j] == first[j]).all()


def test_windows_of_at_most_64_cameras_have_no_sparse_factor(monkeypatch):
    monkeypatch.delenv("SBA_FORCE_SPARSE_CHOL", raising=False)
    smap = synthetic.make_map(num_frames=9, num_points=400, track_len=6, seed=5)
    s = host_structure(smap, monkeypatch, force=False)
    with pytest.raises(Exception):
        s.factor_symbolic()
