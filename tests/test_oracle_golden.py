"""CPU tests: the NumPy oracle against the fixtures recorded from the unmodified reference
(tests/golden/make_golden.py).  No GPU, no /root/reference needed."""
import os
import warnings

import numpy as np
import pytest

import parity as P
from oracle import bsc as OB
from oracle import variational as OV

warnings.simplefilter("ignore")


@pytest.mark.parametrize("name", ["bsc_bars_replay", "bsc_exact", "bsc_allzero_sparseflip", "bsc_cross", "bsc_missing"])
def test_oracle_replay_reproduces_reference_trajectory(name):
    c = P.load_case(os.path.join(P.GOLDEN, name + ".npz"))
    iters = int(c["iters"])
    out, cont, data = P.run_oracle_case(c, iters)
    for it in range(iters):
        F, a, b, th = out[it]
        assert abs(F - c["F"][it]) <= 1e-11 * abs(c["F"][it])
        assert abs(a - c["S_nunique"][it]) < 1e-12 and abs(b - c["S_sub"][it]) < 1e-12
        assert P.maxrel(th["W"], c["W"][it]) < 1e-8
        assert abs(th["pi"] - c["pi"][it]) <= 1e-10 * c["pi"][it]
        assert abs(th["sigma"] - c["sigma"][it]) <= 1e-10 * c["sigma"][it]
    assert np.array_equal(cont["ss"], c["ss_final_bool"])
    assert P.maxrel(cont["lpj"], c["lpj_final"]) < 1e-11
    if float(c["missing"]) > 0:
        assert P.maxrel(data["y_reconstructed"], c["y_reconstructed"]) < 1e-12


def test_oracle_lpj_fixture():
    z = np.load(os.path.join(P.GOLDEN, "bsc_lpj.npz"))
    W, Y, X, states = z["W"], z["Y"], z["X"], z["states"]
    th = OB.precompute({"W": W, "pi": float(z["pi"]), "sigma": float(z["sigma"])}, W.shape[0], W.shape[1], Y.shape[0],
                       int(X.sum()))
    assert abs(th["pre1"] - z["pre1"]) < 1e-15 and abs(th["pil_bar"] - z["pil_bar"]) < 1e-15
    assert abs(th["ljc"] - z["ljc"]) < 1e-12 * abs(z["ljc"])
    for n in range(Y.shape[0]):
        got = OB.lpj_states(states[n], W, Y[n], X[n], th["pre1"], th["pil_bar"])
        assert P.maxrel(got, z["lpj"][n]) < 1e-13
        assert abs(OB.lpj_allzero(Y[n], X[n], th["pre1"])[0] - z["lpj_allzero"][n]) <= 1e-13 * abs(z["lpj_allzero"][n])
    assert np.array_equal(OB.lpj_reset_check(z["reset_in"].copy()), z["reset_out"])


def test_oracle_vary_kn_cases():
    z = np.load(os.path.join(P.GOLDEN, "vary_kn_cases.npz"))
    ncases = 1 + max(int(k.split("_")[0][1:]) for k in z.files)
    assert ncases == 300
    for i in range(ncases):
        g = lambda k: z["c%d_%s" % (i, k)]
        H, S, S_perm, Mp = int(g("H")), int(g("S")), int(g("S_perm")), int(g("Mprime"))
        states, lpj_old, lpj = g("states").copy(), g("lpj_old").copy(), np.zeros(S)
        ret = OV.vary_Kn(lpj_old, g("lpj_new").copy(), lpj, states, g("new").copy(), H, S, S_perm,
                         np.zeros((S_perm, H), bool), Mp)
        if g("new").shape[0] == 0:
            continue
        assert tuple(ret) == tuple(int(v) for v in g("ret"))
        assert np.array_equal(states, g("out_states")) and np.array_equal(lpj, g("out_lpj"))


def test_oracle_full_free_energy():
    z = np.load(os.path.join(P.GOLDEN, "bsc_full_F.npz"))
    Y, W = z["Y"], z["W"]
    L = OB.free_energy_full({"W": W, "pi": float(z["pi"]), "sigma": float(z["sigma"])},
                            {"y": Y, "x_infr": np.ones_like(Y, dtype=bool)}, W.shape[0], W.shape[1],
                            OV.all_states(W.shape[1]))
    assert abs(L - float(z["L"])) <= 1e-12 * abs(float(z["L"]))


def test_pack_unpack_oracle():
    rng = np.random.RandomState(1)
    for H in (1, 31, 32, 33, 64, 65, 300):
        a = rng.rand(3, 4, H) < 0.4
        p = OV.pack_states(a)
        assert p.shape == (3, 4, (H + 31) // 32)
        assert np.array_equal(OV.unpack_states(p, H), a)
        # bit h -> word h // 32, bit h % 32
        h = H - 1
        one = np.zeros((1, H), bool)
        one[0, h] = True
        assert OV.pack_states(one)[0, h // 32] == np.uint32(1 << (h % 32))


# ---- ES3C (SSSC) oracle vs reference fixtures ------------------------------------------------------
@pytest.mark.parametrize("name", ["sssc_exact", "sssc_replay", "sssc_missing", "sssc_allzero_h40"])
def test_sssc_oracle_replay_reproduces_reference(name):
    c = P.load_case(os.path.join(P.GOLDEN, name + ".npz"))
    iters = int(c["iters"])
    out, cont, data = P.run_oracle_sssc_case(c, iters)
    for it in range(iters):
        F, a, b, th = out[it]
        assert abs(F - c["F"][it]) <= 1e-10 * abs(c["F"][it])
        assert abs(a - c["S_nunique"][it]) < 1e-12 and abs(b - c["S_sub"][it]) < 1e-12
        for k in P.PKEYS:
            assert P.maxrel(th[k], c[k][it]) < 1e-7, k
    assert np.array_equal(cont["ss"], c["ss_final_bool"])
    assert P.maxrel(cont["lpj"], c["lpj_final"]) < 1e-9
    assert P.maxrel(data["y_reconstructed"], c["y_reconstructed"]) < 1e-9


def test_sssc_oracle_lpj_fixture():
    from oracle import sssc as OS

    z = np.load(os.path.join(P.GOLDEN, "sssc_lpj.npz"))
    Y, X, states = z["Y"], z["X"], z["states"]
    th = OS.precompute({k: np.array(z[k]) for k in P.PKEYS}, Y.shape[1], Y.shape[0], int(X.sum()))
    assert abs(th["ljc"] - z["ljc"]) <= 1e-12 * abs(z["ljc"])
    for n in range(Y.shape[0]):
        got = OS.lpj_states(states[n], th, Y[n], X[n])[0]
        assert P.maxrel(got, z["lpj"][n]) < 1e-11
        assert abs(OS.lpj_allzero(Y[n], X[n], th["sigma2_inv"])[0] - z["lpj_allzero"][n]) <= 1e-13 * abs(z["lpj_allzero"][n])


# ---- fixtures at the BASELINE shapes (tests/golden/make_golden_shapes.py) ---------------------------------
@pytest.mark.parametrize("name", ["shape_cfg2", "shape_cfg5"])
def test_oracle_replay_at_baseline_shape(name):
    c = P.load_case(os.path.join(P.GOLDEN, name + ".npz"))  # regenerates Y / W0 from the seed and checks their SHA-1
    iters = int(c["iters"])
    out, cont, data = P.run_oracle_case(c, iters)
    for it in range(iters):
        F, a, b, th = out[it]
        assert abs(F - c["F"][it]) <= 1e-11 * abs(c["F"][it])
        assert abs(a - c["S_nunique"][it]) < 1e-12 and abs(b - c["S_sub"][it]) < 1e-12
        assert abs(th["pi"] - c["pi"][it]) <= 1e-10 * c["pi"][it]
        assert abs(th["sigma"] - c["sigma"][it]) <= 1e-10 * c["sigma"][it]
    assert np.array_equal(cont["ss"], c["ss_final_bool"])
    assert P.maxrel(cont["lpj"], c["lpj_final"]) < 1e-11


@pytest.mark.parametrize("name", ["shape_cfg3", "shape_cfg4"])
def test_sssc_oracle_replay_at_baseline_shape(name):
    c = P.load_case(os.path.join(P.GOLDEN, name + ".npz"))
    iters = int(c["iters"])
    out, cont, data = P.run_oracle_sssc_case(c, iters)
    for it in range(iters):
        F, a, b, th = out[it]
        assert abs(F - c["F"][it]) <= 1e-10 * abs(c["F"][it])
        assert abs(a - c["S_nunique"][it]) < 1e-12 and abs(b - c["S_sub"][it]) < 1e-12
        for k in ("pies", "mus", "sigma2"):
            assert P.maxrel(th[k], c[k][it]) < 1e-7, k
    assert np.array_equal(cont["ss"], c["ss_final_bool"])
    assert P.maxrel(cont["lpj"], c["lpj_final"]) < 1e-9
    assert P.maxrel(data["y_reconstructed"], c["y_reconstructed"]) < 1e-6  # stored as float32


def test_sssc_oracle_lpj_larger_systems():
    from oracle import sssc as OS

    z = np.load(os.path.join(P.GOLDEN, "sssc_lpj_dense.npz"))
    H = int(z["H"])
    Y, X = z["Y"], z["X"]
    states = np.unpackbits(z["states"], axis=-1)[..., :H].astype(bool)
    assert states.sum(-1).max() > 32
    th = OS.precompute({k: np.array(z[k]) for k in P.PKEYS}, Y.shape[1], Y.shape[0], int(X.sum()))
    for n in range(Y.shape[0]):
        assert P.maxrel(OS.lpj_states(states[n], th, Y[n], X[n])[0], z["lpj"][n]) < 1e-9
