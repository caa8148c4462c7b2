"""`evo_bsc_solve_static`: the BSC parameter solve for callers that cannot allocate device memory (no counterpart in
the reference; include/evo_b200.h).  Its scratch is a file-scope device array inside the library, shared by every
call, so the test calls it twice in a row -- a larger system first, then a smaller one that would see the first
call's leftovers -- and requires the outputs of `evo_bsc_solve` (caller workspace) each time: same kernels, same
order of operations, so equal to 1e-12 relative (the f64 tolerance of tests/test_solve_gpu.py is 1e-9).
Collected last on purpose (pytest -x at the end of a round)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _stats(rng, D, H):
    Q, _ = np.linalg.qr(rng.normal(size=(H, H)))
    ev = np.exp(rng.uniform(0, np.log(1e3), size=H))
    Wq = (Q * ev) @ Q.T * 50
    Wp = rng.normal(size=(H, D)) * 30
    pies = rng.uniform(1, 100, size=H)
    return np.concatenate([Wp.ravel(), np.tril(Wq).ravel(), pies, [12345.678]]), Wq, Wp


def _solve(lib, _lib, torch, stats, N, D, H, static, n_obs):
    dev = torch.device("cuda")
    sd = torch.as_tensor(stats, device=dev)
    W = torch.zeros((D, H), dtype=torch.float64, device=dev)
    p = torch.zeros(H, dtype=torch.float64, device=dev)
    scal = torch.zeros(4, dtype=torch.float64, device=dev)
    if static:
        rc = lib.evo_bsc_solve_static(_lib.ptr(sd), N, D, H, 1, 0, n_obs, 0.9, _lib.ptr(W), _lib.ptr(p), _lib.ptr(scal),
                                      _lib.stream())
    else:
        ws = torch.empty(int(lib.evo_bsc_solve_workspace_bytes(D, H)), dtype=torch.uint8, device=dev)
        rc = lib.evo_bsc_solve(_lib.ptr(sd), N, D, H, 1, 0, n_obs, 0.9, _lib.ptr(W), _lib.ptr(p), _lib.ptr(scal),
                               _lib.ptr(ws), _lib.stream())
    _lib.check(rc, "evo_bsc_solve_static" if static else "evo_bsc_solve")
    torch.cuda.synchronize()
    return W.cpu().numpy(), p.cpu().numpy(), scal.cpu().numpy()


def test_static_scratch_solve_equals_the_workspace_solve_on_two_calls_in_a_row():
    import torch

    from evo_b200 import _lib

    lib = _lib.load()
    rng = np.random.RandomState(20)
    N = 5000.0
    # second call smaller than the first (stale contents beyond its own extent), third one larger again
    for D, H, n_obs in [(256, 300, -1.0), (40, 33, 777.0), (144, 512, -1.0)]:
        stats, Wq, Wp = _stats(rng, D, H)
        want = _solve(lib, _lib, torch, stats, N, D, H, False, n_obs)
        got = _solve(lib, _lib, torch, stats, N, D, H, True, n_obs)
        assert got[2][0] == 0 and got[2][3] == 1
        for g, w in zip(got, want):
            assert np.abs(g - w).max() <= 1e-12 * np.abs(w).max()
        ref = np.linalg.lstsq(Wq, Wp, rcond=-1)[0].T
        assert np.abs(got[0] - ref).max() <= 1e-9 * np.abs(ref).max()


def test_static_scratch_solve_rejects_sizes_beyond_its_maximum():
    import torch

    from evo_b200 import _lib

    lib = _lib.load()
    dev = torch.device("cuda")
    D = H = 1024
    assert int(lib.evo_bsc_solve_workspace_bytes(D, H)) > 6 * 1024 * 1024
    sd = torch.zeros(H * D + H * H + H + 1, dtype=torch.float64, device=dev)
    W = torch.zeros((D, H), dtype=torch.float64, device=dev)
    p = torch.zeros(H, dtype=torch.float64, device=dev)
    scal = torch.zeros(4, dtype=torch.float64, device=dev)
    rc = lib.evo_bsc_solve_static(_lib.ptr(sd), 10.0, D, H, 1, 0, -1.0, 0.9, _lib.ptr(W), _lib.ptr(p), _lib.ptr(scal),
                                  _lib.stream())
    assert rc != 0
    with pytest.raises(_lib.EvoError, match="EVO_BSC_SOLVE_STATIC_BYTES"):
        _lib.check(rc, "evo_bsc_solve_static")
