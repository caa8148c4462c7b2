"""GPU parity tests for the BSC path: CUDA (through the C-ABI) vs the fixtures recorded from the
unmodified reference and vs the NumPy oracle replaying the same injected candidate sets.

Tolerances: K sets and slots bit-exact; F, lpj, theta within 1e-9 relative (all device arithmetic on
this path is float64; Y and W0 of the fixtures are exactly float32-representable)."""
import os
import warnings

import numpy as np
import pytest

import parity as P

pytestmark = pytest.mark.gpu
warnings.simplefilter("ignore")

CASES = ["bsc_bars_replay", "bsc_exact", "bsc_allzero_sparseflip", "bsc_cross", "bsc_missing"]


@pytest.mark.parametrize("name", CASES)
def test_trajectory_matches_reference_and_oracle(name):
    res = P.run_bsc_parity_case(os.path.join(P.GOLDEN, name + ".npz"))
    assert res["ss_equal"], res
    assert res["ss_equal_ref"], res
    assert res["count_err"] == 0.0, res
    # y_reconstructed lives on the device as float32 (like Y), so the missing-data trajectory carries
    # a 6e-8 relative rounding of the estimated entries into the M-step; everything else is float64
    tol = 1e-6 if name == "bsc_missing" else 1e-9
    for k in ("F_err", "F_err_ref", "pi_err", "sigma_err", "lpj_err", "lpj_err_ref"):
        assert res[k] < tol, (k, res)
    assert res["W_err"] < max(tol, 1e-7) and res["W_err_ref"] < max(tol, 1e-7), res
    if "yrec_err_ref" in res:
        assert res["yrec_err_ref"] < 1e-6, res  # y_reconstructed is stored as float32 on the device


@pytest.mark.parametrize("name", ["shape_cfg2", "shape_cfg5"])
def test_trajectory_at_baseline_shape(name):
    """Replayed reference runs at the BASELINE shapes (K = 64 / 128, 10 / 32 words per state, 128 / 256 children
    per datapoint; W fixed, pi and sigma learned -- tests/golden/make_golden_shapes.py)."""
    res = P.run_bsc_parity_case(os.path.join(P.GOLDEN, name + ".npz"))
    assert res["ss_equal"] and res["ss_equal_ref"], res
    assert res["count_err"] == 0.0, res
    for k in ("F_err", "F_err_ref", "pi_err", "sigma_err", "lpj_err", "lpj_err_ref"):
        assert res[k] < 1e-9, (k, res)


@pytest.mark.parametrize("name", ["bsc_bars_replay", "bsc_cross", "shape_cfg2", "shape_cfg5"])
def test_trajectory_with_tensor_core_tables(name):
    """The configuration bench.py times (`tables="tc"`, `stats="tc"`: B = Y W and Wp through fp16 hi+lo split
    tcgen05 GEMMs) on the recorded trajectories.  Stated bound: 1e-5 on lpj / F / parameters for ONE EM iteration
    from identical parameters (per-call bounds: tests/test_dense_gpu.py); over a recorded trajectory of n <= 6
    iterations the deviations compound through the parameters, bound 1e-4.  K sets may differ from the f64 path
    only where two candidates' lpj are closer than that, so lpj is compared when they agree."""
    path = os.path.join(P.GOLDEN, name + ".npz")
    one = P.run_bsc_parity_case(path, iters=1, tables="tc", stats="tc")
    for k in ("F_err", "F_err_ref", "pi_err", "sigma_err", "W_err", "W_err_ref"):
        assert one[k] < 1e-5, (k, one)
    if one["ss_equal"]:
        assert one["lpj_err"] < 1e-5, one
    res = P.run_bsc_parity_case(path, tables="tc", stats="tc")
    for k in ("F_err", "F_err_ref", "pi_err", "sigma_err", "W_err", "W_err_ref"):
        assert res[k] < 1e-4, (k, res)
    if res["ss_equal"]:
        assert res["lpj_err"] < 1e-4, res


def test_lpj_fixture_direct_and_gram():
    """bsc_lpj.npz: masked + complete rows, an all-zero and an all-one state (reference values)."""
    import torch

    from evo_b200 import _lib
    from evo_b200.bsc import BSC
    from evo_b200.variational import pack_states

    z = np.load(os.path.join(P.GOLDEN, "bsc_lpj.npz"))
    W, Y, X, states = z["W"], z["Y"], z["X"], z["states"]
    N, S, H = states.shape
    D = W.shape[0]
    m = BSC(D, H, S)
    th = {"W": W, "pi": float(z["pi"]), "sigma": float(z["sigma"]), "pre1": float(z["pre1"]), "pil_bar": float(z["pil_bar"])}
    # per-datapoint reference signature
    for n in range(N):
        got = m.log_pseudo_joint(th, {"this_states": states[n]}, {"this_y": Y[n], "this_x_infr": X[n]})
        # the fixture's y is float64; the device stores y as float32 (6e-8 relative input rounding)
        assert P.maxrel(got, z["lpj"][n]) < 1e-6
        got0 = m.log_pseudo_joint_permanent_states(th, {}, {"this_y": Y[n], "this_x_infr": X[n]})
        assert abs(got0[0] - z["lpj_allzero"][n]) <= 1e-6 * abs(z["lpj_allzero"][n])
    # batched: Gram form on the complete rows equals the direct form
    rows = [n for n in range(N) if X[n].all()]
    data = {"y": Y[rows].astype(np.float32).astype(np.float64), "x_infr": X[rows]}
    m.E_step_precompute(th, {}, data)
    st = pack_states(states[rows], m.device)
    gram = m.lpj_all(th, data, st).cpu().numpy()
    from oracle import bsc as OB

    for i, n in enumerate(rows):
        want = OB.lpj_states(states[n], W, data["y"][i], X[n], th["pre1"], th["pil_bar"])
        assert P.maxrel(gram[i], want) < 1e-11


def test_lpj_reset_check_semantics():
    """NaN -> f64 lowest, +-inf -> 0.0 (the reference's B_max), counters one per call."""
    from evo_b200.bsc import BSC

    D, H = 4, 6
    m = BSC(D, H, 3)
    W = np.zeros((D, H))
    W[0, 0] = np.inf  # state with latent 0 on -> inf residual -> -inf lpj -> reset to 0.0
    W[1, 1] = np.nan  # latent 1 on -> NaN
    th = {"W": W, "pi": 0.2, "sigma": 1.0, "pre1": -0.5, "pil_bar": float(np.log(0.25))}
    st = np.zeros((3, H), dtype=bool)
    st[0, 0] = True
    st[1, 1] = True
    st[2, 2] = True
    ss = {"this_states": st, "reset_lpj_isnan": 0, "reset_lpj_smaller_eps_lpj": 0, "reset_lpj_isinf": 0}
    out = m.log_pseudo_joint(th, ss, {"this_y": np.ones(D), "this_x_infr": np.ones(D, dtype=bool)})
    assert out[0] == 0.0 and out[1] == np.finfo(np.float64).min and np.isfinite(out[2])
    assert ss["reset_lpj_isnan"] == 1 and ss["reset_lpj_isinf"] == 0  # NaN wins the elif chain


def test_vary_kn_golden_cases():
    """300 random vary_Kn cases recorded from the reference: slots, values and return counts."""
    from evo_b200.variational import vary_Kn

    z = np.load(os.path.join(P.GOLDEN, "vary_kn_cases.npz"))
    ncases = 1 + max(int(k.split("_")[0][1:]) for k in z.files)
    for i in range(ncases):
        g = lambda k: z["c%d_%s" % (i, k)]
        H, S, S_perm, Mp = int(g("H")), int(g("S")), int(g("S_perm")), int(g("Mprime"))
        states, new = g("states").copy(), g("new")
        lpj_old, lpj_new = g("lpj_old").copy(), g("lpj_new")
        lpj = np.zeros(S)
        if new.shape[0] == 0:
            continue
        ret = vary_Kn(lpj_old, lpj_new, lpj, states, new, H, S, S_perm, np.zeros((S_perm, H), bool), Mp)
        assert tuple(ret) == tuple(int(v) for v in g("ret")), i
        assert np.array_equal(states, g("out_states")), i
        assert np.array_equal(lpj, g("out_lpj")), i


def test_free_energy_full():
    """Exact log-likelihood over all 2^H states (Model.free_energy(full=True)) vs the reference value."""
    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    z = np.load(os.path.join(P.GOLDEN, "bsc_full_F.npz"))
    Y, W = z["Y"], z["W"]
    N, D = Y.shape
    H = W.shape[1]
    m = BSC(D, H, 20)
    cont = init_states(N, 20, H, "fit", "randflip", 5, 2, 2)
    th = {"W": W, "pi": float(z["pi"]), "sigma": float(z["sigma"])}
    L = m.free_energy({"y": Y, "x_infr": np.ones_like(Y, dtype=bool)}, th, cont, full=True)
    assert abs(L - float(z["L"])) < 1e-10 * abs(float(z["L"]))


def test_pack_unpack_roundtrip_and_init_states():
    import torch

    from evo_b200.variational import init_states, pack_states, unpack_states
    import oracle.variational as OV

    rng = np.random.RandomState(0)
    for H in (1, 10, 32, 33, 70, 300):
        a = rng.rand(7, 5, H) < 0.3
        p = pack_states(a, torch.device("cuda"))
        assert np.array_equal(p.cpu().numpy().view(np.uint32), OV.pack_states(a))
        assert np.array_equal(unpack_states(p, H).cpu().numpy(), a)
    # initial K sets: S distinct states per datapoint, sparse, never the permanent state
    for (N, S, H, perm) in [(50, 20, 10, None), (40, 64, 300, None),
                            (30, 16, 40, {"background": False, "allzero": True, "singletons": False}),
                            (20, 6, 9, {"background": True, "allzero": False, "singletons": False})]:
        c = init_states(N, S, H, "fit", "randflip", 2, 1, 1, permanent=perm)
        ss = c["ss"]
        assert ss.shape == (N, S, H)
        for n in range(N):
            assert len({r.tobytes() for r in ss[n]}) == S
            if perm and perm["allzero"]:
                assert ss[n].any(axis=1).all()
            if perm and perm["background"]:
                assert ss[n][:, -1].all()
        mean_on = ss[:, :, : (H - 1 if perm and perm["background"] else H)].mean()
        assert mean_on < 6.0 / H


@pytest.mark.parametrize("model", ["bsc", "bsc_missing", "es3c", "es3c_missing"])
def test_free_energy_full_enumerated_equals_table_path(model):
    """free_energy(full=True) enumerates the 2^H states on the device (evo_*_full_logsumexp); for H = 13 -- beyond the
    reference's H < 12 cap -- it must equal the log-likelihood computed from an explicit state table through the
    stand-alone lpj kernels (the route validated against the reference at H = 10 above)."""
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.sssc import SSSC
    from evo_b200.variational import all_states, init_states, pack_states

    D, H, N = 9, 13, 40
    rng = np.random.RandomState(5)
    Y = (rng.normal(size=(N, D)) * 2).astype(np.float32).astype(np.float64)
    X = rng.random_sample((N, D)) > 0.25 if "missing" in model else None
    data = {"y": np.where(X, Y, np.nan) if X is not None else Y, "x_infr": X}
    if model.startswith("bsc"):
        m = BSC(D, H, 8)
        th = m.check_params({"W": rng.normal(size=(D, H)), "pi": 0.15, "sigma": 1.1})
    else:
        m = SSSC(D, H, 8)
        A = rng.normal(size=(H, H)) * 0.2
        th = m.check_params({"W": rng.normal(size=(D, H)), "pies": rng.uniform(0.05, 0.3, size=H), "mus": rng.normal(size=H) + 1,
                             "Psi": np.eye(H) + A @ A.T, "sigma2": np.float64(1.2)})
    cont = init_states(N, 8, H, "fit", "randflip", 2, 1, 1)
    L = m.free_energy(data, th, cont, full=True)
    tab = pack_states(all_states(H), m.device)
    lpj = m.lpj_all(th, data, tab, shared=True)
    want = float(th["ljc"] + m._sum_logsumexp(lpj).item() / N)
    assert abs(L - want) < 1e-11 * abs(want), (L, want)
