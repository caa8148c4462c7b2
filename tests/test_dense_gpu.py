"""GPU parity of the dense tensor-core log-joint (`evo_bsc_lpj_gemm`, tcgen05 + TMA) against the NumPy
oracle (oracle/bsc.py:lpj_states, the reference's bsc.py:79-97 in float64) and against the sparse
Gram kernel, on seeded inputs.

Stated tolerance: 1e-5 relative on lpj (W is an fp16 hi+lo split with 22 significant bits, the
accumulator and the residual are fp32, the row sum is f64; measured <= 2e-6)."""
import os

import numpy as np
import pytest

import parity as P

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _case(D, H, S, N, act, seed, missing=0.0, wscale=1.0):
    rng = np.random.RandomState(seed)
    W = rng.normal(size=(D, H)) * wscale
    st = rng.random_sample((N, S, H)) < act / H
    st[0, 0] = False  # an all-zero state
    st[N - 1, S - 1] = True  # and an all-one state
    gen = rng.random_sample((N, H)) < act / H
    Y = (gen @ W.T + rng.normal(size=(N, D)) * wscale).astype(np.float32)
    X = rng.random_sample((N, D)) >= missing if missing else np.ones((N, D), dtype=bool)
    return W, Y, X, st


def _oracle(W, Y, X, st, pre1, pil_bar):
    from oracle import bsc as OB

    return np.stack([OB.lpj_states(st[n], W, Y[n].astype(np.float64), X[n], pre1, pil_bar) for n in range(st.shape[0])])


@pytest.mark.parametrize(
    "D,H,S,N,act,missing,wscale",
    [
        (25, 10, 20, 37, 2.0, 0.0, 10.0),     # cfg1 shape: rows not a multiple of 128, S not a multiple of 32
        (256, 300, 64, 300, 5.0, 0.0, 1.0),   # cfg2 shape: H not a multiple of 64
        (144, 70, 48, 33, 4.0, 0.3, 1e-3),    # masked entries, tiny W (power-of-two rescaling)
        (300, 130, 64, 20, 6.0, 0.0, 1.0),    # D > 256 with a ragged second column tile
        (1024, 1024, 128, 12, 8.0, 0.0, 1.0),  # cfg5 shape: 4 column tiles, 16 k-blocks
        (1100, 40, 32, 9, 4.0, 0.0, 1.0),     # D > 1024: y is read from global memory, 5 column tiles
    ],
)
def test_dense_lpj_matches_oracle_and_gram(D, H, S, N, act, missing, wscale):
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import pack_states

    W, Y, X, st = _case(D, H, S, N, act, seed=D + H, missing=missing, wscale=wscale)
    sigma, pi = 1.3 * wscale, act / H
    pre1, pil_bar = -0.5 / sigma ** 2, float(np.log(pi / (1 - pi)))
    ref = _oracle(W, Y, X, st, pre1, pil_bar)
    m = BSC(D, H, S)
    th = {"W": W, "pi": pi, "sigma": sigma, "pre1": pre1, "pil_bar": pil_bar}
    data = {"y": Y.astype(np.float64), "x_infr": X}
    sd = pack_states(st, m.device)
    got = m.lpj_all_dense(th, data, sd).cpu().numpy()
    err = np.max(np.abs(got - ref) / np.maximum(np.abs(ref), 1e-300))
    assert err < TOL, err
    sparse = m.lpj_all(th, data, sd).cpu().numpy()
    assert P.maxrel(sparse, ref) < 1e-9
    # one shared state table for every datapoint (free_energy(full=True) layout)
    got_sh = m.lpj_all_dense(th, data, sd[0], shared=True).cpu().numpy()
    ref_sh = _oracle(W, Y, X, np.broadcast_to(st[0], st.shape), pre1, pil_bar)
    assert np.max(np.abs(got_sh - ref_sh) / np.abs(ref_sh)) < TOL


def test_dense_lpj_linearity_at_full_tile_count():
    """Size-independent property at a size the oracle cannot reach: with y = 0 and W -> c W the
    residual scales by c^2 exactly (power-of-two c), for every row of many persistent tiles."""
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    D, H, S, N = 256, 300, 64, 20000
    rng = np.random.RandomState(7)
    W = rng.normal(size=(D, H))
    m = BSC(D, H, S)
    cont = init_states(N, S, H, "fit", "randflip", 8, 8, 2, device=m.device, seed=3)
    data = {"y": torch.zeros((N, D), dtype=torch.float32, device=m.device), "x_infr": None}
    th = {"W": W, "pre1": -0.5, "pil_bar": 0.0}
    a = m.lpj_all_dense(th, data, cont["ss_dev"]).cpu().numpy()
    th4 = {"W": 4.0 * W, "pre1": -0.5, "pil_bar": 0.0}
    b = m.lpj_all_dense(th4, data, cont["ss_dev"]).cpu().numpy()
    assert np.array_equal(16.0 * a, b)
    s = m.lpj_all(th, data, cont["ss_dev"]).cpu().numpy()
    assert np.max(np.abs(a - s) / np.maximum(np.abs(s), 1e-12)) < TOL


@pytest.mark.parametrize("D,H,N,wscale", [(25, 10, 300, 10.0), (256, 300, 1000, 1.0), (144, 70, 129, 1e-3), (1024, 1024, 500, 1.0), (40, 33, 130, 1.0)])
def test_tc_table_matches_f64_table(D, H, N, wscale):
    """B = Y W from the tcgen05 split-precision GEMM vs the f64 SIMT GEMM (evo_bsc_prepare) and NumPy.
    Stated bound: 1e-5 of max|B| (fp16 hi+lo operands, three passes, fp32 tensor-core accumulation over
    D; measured 1e-7 at D=25 ... 5e-6 at D=1024)."""
    import torch

    from evo_b200 import _lib
    from evo_b200.bsc import BSC

    rng = np.random.RandomState(D + H)
    W = rng.normal(size=(D, H)) * wscale
    Y = (rng.normal(size=(N, D)) * 3.0 * wscale).astype(np.float32)
    m = BSC(D, H, 8)
    Wd = torch.as_tensor(W, device=m.device)
    Yd = torch.as_tensor(Y, device=m.device)
    B64 = torch.empty((N, H), dtype=torch.float64, device=m.device)
    G = torch.empty((H, H), dtype=torch.float64, device=m.device)
    m._call("evo_bsc_prepare", _lib.ptr(Wd), _lib.ptr(Yd), N, D, H, _lib.ptr(G), _lib.ptr(B64), None)
    ws = torch.empty(int(m.lib.evo_bsc_tables_tc_workspace_bytes(N, D, H)), dtype=torch.uint8, device=m.device)
    Btc = torch.full((N, H), float("nan"), dtype=torch.float64, device=m.device)
    m._call("evo_bsc_tables_tc_set_data", _lib.ptr(Yd), N, D, H, _lib.ptr(ws))
    m._call("evo_bsc_tables_tc", _lib.ptr(Wd), N, D, H, _lib.ptr(ws), _lib.ptr(Btc))
    ref = Y.astype(np.float64) @ W
    scale = np.abs(ref).max()
    assert np.abs(B64.cpu().numpy() - ref).max() < 1e-12 * scale
    err = np.abs(Btc.cpu().numpy() - ref).max() / scale
    print('tc table err / max|B|:', D, H, err)
    assert err < 1e-5, err


def test_tc_tables_em_trajectory_matches_f64_tables():
    """Three EM iterations of BSC (RNG mode, same Philox streams) with the tensor-core table vs the f64
    table: free energies within 1e-5 relative, W within 1e-4 of its scale (selection may differ on
    near-ties below the table's 2e-6 resolution -- documented)."""
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    D, H, K, N = 64, 40, 32, 4000
    rng = np.random.RandomState(3)
    Wg = rng.normal(size=(D, H)) * 2.0
    s = rng.random_sample((N, H)) < 3.0 / H
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)
    out = {}
    for mode in ("f64", "tc"):
        m = BSC(D, H, K)
        m.tables = mode
        data = {"y": torch.as_tensor(Y, device=m.device), "x_infr": None}
        np.random.seed(1)
        th = m.check_params(m.standard_init(data))
        cont = init_states(N, K, H, "fit", "randflip", 6, 4, 2, device=m.device, seed=11)
        Fs = []
        for _ in range(3):
            F, _, _, th = m.step(th, cont, data)
            Fs.append(F)
        W = th["W"].cpu().numpy() if isinstance(th["W"], torch.Tensor) else np.asarray(th["W"])
        out[mode] = (np.array(Fs), W)
    # trajectory-level bound (per-call bounds are 1e-5, above): a log-joint difference of 1e-6 can swap two near-tied
    # candidates, after which the two runs hold slightly different K sets
    ferr = np.max(np.abs(out["tc"][0] - out["f64"][0]) / np.abs(out["f64"][0]))
    werr = np.max(np.abs(out["tc"][1] - out["f64"][1])) / np.abs(out["f64"][1]).max()
    print("tc vs f64 trajectory: F", ferr, "W", werr)
    assert ferr < 1e-4, out
    assert werr < 1e-3, werr


@pytest.mark.parametrize("D,H,S,N,act,bound", [(256, 300, 64, 400, 5.0, 1e-5), (1024, 1024, 128, 64, 8.0, 1e-5)])
def test_gram_lpj_from_tc_table(D, H, S, N, act, bound):
    """lpj of the sparse Gram kernel fed by the tensor-core table vs the f64 oracle (stated bound 1e-5 relative; measured 1.3e-6 (D=256) and 1.5e-6 (D=1024):
    the table's fp32 accumulation error grows with D)."""
    from evo_b200.bsc import BSC
    from evo_b200.variational import pack_states

    W, Y, X, st = _case(D, H, S, N, act, seed=5 * D + H)
    sigma, pi = 1.3, act / H
    pre1, pil_bar = -0.5 / sigma ** 2, float(np.log(pi / (1 - pi)))
    ref = _oracle(W, Y, X, st, pre1, pil_bar)
    m = BSC(D, H, S)
    m.tables = "tc"
    th = {"W": W, "pi": pi, "sigma": sigma, "pre1": pre1, "pil_bar": pil_bar}
    data = {"y": Y.astype(np.float64), "x_infr": X}
    got = m.lpj_all(th, data, pack_states(st, m.device)).cpu().numpy()
    err = np.max(np.abs(got - ref) / np.abs(ref))
    print("gram lpj from tc table, max rel err:", D, H, err)
    assert err < bound, err


@pytest.mark.parametrize("D,H,K,N", [(25, 10, 20, 700), (256, 300, 64, 5000), (144, 70, 48, 3001), (1024, 1024, 32, 600), (300, 130, 16, 129)])
def test_tc_wp_statistics_match_f64_statistics(D, H, K, N):
    """M-step statistics with Wp = <S>' Y from the split-K tcgen05 GEMM (`evo_bsc_suffstats_tc`) vs the f64
    path (`evo_bsc_suffstats`) on the same K sets and lpj.  Stated bound: 1e-5 of max|Wp| on Wp (fp16 hi+lo
    operands, three passes, fp32 accumulation over <= 4096 datapoints, f64 across work items); Wq, pies
    and sigma come from the same f64 code (equal to summation order)."""
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    rng = np.random.RandomState(D + H)
    Wg = rng.normal(size=(D, H)) * 2.0
    s = rng.random_sample((N, H)) < 3.0 / H
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)
    m = BSC(D, H, K)
    data = {"y": torch.as_tensor(Y, device=m.device), "x_infr": None}
    np.random.seed(1)
    th = m.check_params(m.standard_init(data))
    cont = init_states(N, K, H, "fit", "randflip", 4, 3, 2, device=m.device, seed=11)
    m.E_step(th, cont, data)
    dd = m.device_data(data)
    out = {}
    for mode in ("f64", "tc"):
        m.stats = mode
        out[mode] = m._stats_enqueue(th, cont, dd).cpu().numpy()
    m.stats = "f64"
    nWp = H * D
    a, b = out["f64"], out["tc"]
    err = np.abs(a[:nWp] - b[:nWp]).max() / np.abs(a[:nWp]).max()
    print("tc Wp err / max|Wp|:", D, H, N, err)
    assert err < 1e-5, err
    rest = np.abs(a[nWp:] - b[nWp:]).max() / np.abs(a[nWp:]).max()
    assert rest < 1e-12, rest


def test_tc_statistics_em_trajectory_matches_f64():
    """Three EM iterations with tables and statistics both on the tensor cores vs both f64: F within 1e-5
    relative, W within 1e-4 of its scale."""
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    D, H, K, N = 64, 40, 32, 4000
    rng = np.random.RandomState(3)
    Wg = rng.normal(size=(D, H)) * 2.0
    s = rng.random_sample((N, H)) < 3.0 / H
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)
    out = {}
    for mode in ("f64", "tc"):
        m = BSC(D, H, K)
        m.tables = m.stats = mode
        data = {"y": torch.as_tensor(Y, device=m.device), "x_infr": None}
        np.random.seed(1)
        th = m.check_params(m.standard_init(data))
        cont = init_states(N, K, H, "fit", "randflip", 6, 4, 2, device=m.device, seed=11)
        Fs = []
        for _ in range(3):
            F, _, _, th = m.step(th, cont, data)
            Fs.append(F)
        W = th["W"].cpu().numpy() if isinstance(th["W"], torch.Tensor) else np.asarray(th["W"])
        out[mode] = (np.array(Fs), W)
    assert np.max(np.abs(out["tc"][0] - out["f64"][0]) / np.abs(out["f64"][0])) < 1e-5, out
    # W: a 1e-6 log-joint difference swaps near-tied candidates on a few of the 4000 datapoints, whose K sets (and hence
    # their share of Wp / Wq) then differ between the two runs: 1e-2 of the scale of W over three iterations
    assert np.max(np.abs(out["tc"][1] - out["f64"][1])) < 1e-2 * np.abs(out["f64"][1]).max()


def test_tc_statistics_with_missing_data_match_f64():
    """Incomplete data: the M-step sums over the reconstructed y, which changes with every iteration, so the
    transposed Y planes of the tensor-core Wp are rebuilt per call (`evo_bsc_wp_tc_set_data`).  Three EM
    iterations with reconstruction, statistics tc vs f64 (tables stay f64: the direct evaluator is used)."""
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    D, H, K, N = 48, 30, 24, 3000
    rng = np.random.RandomState(5)
    Wg = rng.normal(size=(D, H)) * 2.0
    s = rng.random_sample((N, H)) < 3.0 / H
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)
    X = rng.random_sample((N, D)) >= 0.25
    out = {}
    for mode in ("f64", "tc"):
        m = BSC(D, H, K)
        m.stats = mode
        data = {"y": torch.as_tensor(Y, device=m.device), "x_infr": torch.as_tensor(X, device=m.device),
                "x": torch.as_tensor(X, device=m.device)}
        np.random.seed(1)
        th = m.check_params(m.standard_init(data))
        cont = init_states(N, K, H, "fit", "randflip", 6, 4, 2, device=m.device, seed=11)
        Fs = []
        for _ in range(3):
            F, _, _, th = m.step(th, cont, data, do_reconstruction=True)
            Fs.append(F)
        W = th["W"].cpu().numpy() if isinstance(th["W"], torch.Tensor) else np.asarray(th["W"])
        out[mode] = (np.array(Fs), W, float(th["sigma"]))
    assert np.max(np.abs(out["tc"][0] - out["f64"][0]) / np.abs(out["f64"][0])) < 1e-5, out
    assert np.max(np.abs(out["tc"][1] - out["f64"][1])) < 1e-4 * np.abs(out["f64"][1]).max()
    assert abs(out["tc"][2] - out["f64"][2]) < 1e-5 * out["f64"][2]


@pytest.mark.parametrize("D,H,K,N,act,ea,perm", [
    (64, 40, 32, 3000, 3.0, (6, 4, 2), False),    # everything through the fused tail
    (32, 96, 24, 1500, 24.0, (6, 4, 2), False),   # states around the 31-latent list capacity: a mix of tail and rest list
    (48, 300, 64, 2000, 5.0, (8, 8, 2), False),   # cfg2-like shape
    (40, 1030, 32, 8000, 4.0, (8, 4, 2), False),  # 33 words per packed state (N large enough for a regular Wq)
    (64, 40, 32, 1000, 3.0, (6, 4, 2), True),     # permanent all-zero state
])
def test_fused_statistics_equal_the_separate_pass(D, H, K, N, act, ea, perm):
    """BSC.step with the statistics formed in the tail of the E-step kernel (evo_estep_args.fs_*, the bench's path)
    against the same step with the separate statistics pass: same K sets and F, parameters equal up to the order of
    the f64 atomics."""
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    rng = np.random.RandomState(D + H)
    Wg = rng.normal(size=(D, H)) * 2.0
    s = rng.random_sample((N, H)) < act / H
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)
    permanent = {"background": False, "allzero": True, "singletons": False} if perm else None
    out = {}
    for fuse in (True, False):
        m = BSC(D, H, K)
        m.tables = m.stats = "tc"
        m.fuse_stats = fuse
        data = {"y": torch.as_tensor(Y, device=m.device), "x_infr": None}
        np.random.seed(1)
        th = m.check_params(m.standard_init(data))
        cont = init_states(N, K, H, "fit", "randflip", *ea, p_init_Kn=act / H, permanent=permanent, device=m.device, seed=11)
        Fs = []
        for _ in range(2):
            F, _, _, th = m.step(th, cont, data)
            Fs.append(F)
        rest = int(data and m.device_data(data).cache["fs_rest"][0].item()) if fuse else None
        out[fuse] = (np.array(Fs), np.asarray(th["W"]), th["pi"], th["sigma"], cont["ss_dev"].cpu().numpy(), rest)
    a, b = out[True], out[False]
    assert np.array_equal(a[4], b[4])
    # same sums in another order of the f64 atomics (and the order varies from run to run): observed <= 1e-14 on F and
    # <= 1e-11 on W; the bounds leave room for the solve's conditioning
    assert np.max(np.abs(a[0] - b[0]) / np.abs(b[0])) < 1e-11
    assert np.abs(a[1] - b[1]).max() < 1e-8 * np.abs(b[1]).max()
    assert abs(a[2] - b[2]) < 1e-11 * b[2] and abs(a[3] - b[3]) < 1e-11 * b[3]
    if act == 24.0:
        assert 0 < a[5] < N, a[5]  # both routes took part
    elif act == 3.0:
        assert a[5] < N // 10, a[5]  # a few datapoints replace more slots than there are scratch rows
