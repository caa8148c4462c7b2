"""Data generation.  CPU: Philox known-answer vectors and the NumPy dataset definition
(oracle/datagen.py) against the moments of the reference's generate_data (tests/golden/datagen_moments.npz,
recorded from the unmodified reference).  GPU: `evo_generate_data` == the NumPy definition (s bit-exact, y to
float32 rounding), independent of the sharding, and the same moments as the reference."""
import os

import numpy as np
import pytest

import parity as P
from oracle import datagen as DG

Z = np.load(os.path.join(P.GOLDEN, "datagen_moments.npz"))


def test_philox_known_answers():
    """Random123 kat_vectors for philox4x32-10."""
    kat = [((0, 0, 0, 0), 0, (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, 0xffffffffffffffff, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), 0x299f31d0a4093822, (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = DG.philox4x32(key, *ctr)
        assert tuple(int(x) for x in got) == want


def _check_moments(prefix, gen, n, es3c):
    s, y = gen["s"], gen["y"]
    g = lambda k: Z[prefix + k]
    n_ref = float(g("n"))
    se = lambda var: 6.0 * np.sqrt(var * (1.0 / n + 1.0 / n_ref))
    assert (np.abs(s.mean(0) - g("s_mean")) < se(0.25) + 1e-9).all()
    yc = np.cov(y.T)
    assert (np.abs(y.mean(0) - g("y_mean")) < se(np.diag(yc))).all()
    scale = np.sqrt(np.outer(np.diag(yc), np.diag(yc)))
    assert (np.abs(yc - g("y_cov")) < 12.0 * scale * np.sqrt(3.0 * (1.0 / n + 1.0 / n_ref))).all()
    rv = (y - gen["y_mean"]).var(0)
    assert (np.abs(rv - g("resid_var")) < 0.03 * g("resid_var")).all()
    if es3c:
        z = gen["z"]
        H = s.shape[1]
        zm = np.array([z[s[:, h], h].mean() for h in range(H)])
        n_h = np.minimum(s.sum(0), g("zz_cnt").diagonal())
        assert (np.abs(zm - g("z_mean")) < 6.0 * np.sqrt(2.0 * np.diag(g("zz")) / n_h)).all()
        for h in range(H):
            for k in range(H):
                both = s[:, h] & s[:, k]
                m = min(both.sum(), g("zz_cnt")[h, k])
                got = (z[both, h] * z[both, k]).mean()
                tol = 8.0 * np.sqrt(2.0 * 3.0 * g("zz")[h, h] * g("zz")[k, k] / m)
                assert abs(got - g("zz")[h, k]) < tol, (h, k, got, g("zz")[h, k], tol)


def test_numpy_definition_has_the_reference_moments():
    n = 60000
    gen = DG.generate(Z["bsc_W"], Z["bsc_pies"], float(Z["bsc_sigma"]), np.arange(n), 17)
    _check_moments("bsc_", gen, n, False)
    n = 30000
    gen = DG.generate(Z["es3c_W"], Z["es3c_pies"], float(np.sqrt(Z["es3c_sigma2"])), np.arange(n), 18, mus=Z["es3c_mus"],
                      Psi=Z["es3c_Psi"], psi_diag=False)
    _check_moments("es3c_", gen, n, True)


@pytest.mark.gpu
def test_device_generator_equals_the_numpy_definition_and_is_shard_independent():
    from evo_b200.bsc import BSC
    from evo_b200.sssc import SSSC

    rng = np.random.RandomState(2)
    for D, H in ((16, 40), (70, 300)):
        W = rng.normal(size=(D, H))
        pies = rng.uniform(0.01, 0.2, size=H)
        m = BSC(D, H, 4)
        th = {"W": W, "pies": pies, "sigma": 0.8}
        n = 700
        full = m.generate_data(th, n, seed=99)
        ref = DG.generate(W, pies, 0.8, np.arange(n), 99)
        assert np.array_equal(full["s"], ref["s"])
        assert np.abs(full["y_mean"] - ref["y_mean"]).max() <= 1e-6 * np.abs(ref["y_mean"]).max()
        assert np.abs(full["y"] - ref["y"]).max() <= 2e-6 * np.abs(ref["y"]).max()
        # two shards of the same dataset
        a = m.generate_data(th, 300, seed=99, n_offset=0)
        b = m.generate_data(th, 400, seed=99, n_offset=300)
        assert np.array_equal(np.concatenate([a["y"], b["y"]]), full["y"])
        assert np.array_equal(np.concatenate([a["s"], b["s"]]), full["s"])
        th2 = {"W": W, "pi": 0.07, "sigma": 0.8}  # scalar prior (BSC's learned parameter)
        assert np.array_equal(m.generate_data(th2, 100, seed=5)["s"], DG.generate(W, 0.07, 0.8, np.arange(100), 5)["s"])
    D, H = 12, 24
    W = rng.normal(size=(D, H))
    A = rng.normal(size=(H, H)) * 0.3
    Psi = 0.5 * np.eye(H) + A @ A.T
    mus = rng.normal(size=H)
    ms = SSSC(D, H, 4)
    for psi, diag in ((Psi, False), (np.diag(np.diag(Psi)), True)):
        th = {"W": W, "pies": np.full(H, 0.2), "mus": mus, "Psi": psi, "sigma2": np.float64(0.64)}
        got = ms.generate_data(th, 500, seed=7)
        ref = DG.generate(W, np.full(H, 0.2), 0.8, np.arange(500), 7, mus=mus, Psi=psi, psi_diag=diag)
        assert np.array_equal(got["s"], ref["s"])
        assert np.abs(got["z"] - ref["z"]).max() <= 1e-6 * np.abs(ref["z"]).max()
        assert np.abs(got["y"] - ref["y"]).max() <= 2e-6 * np.abs(ref["y"]).max()


@pytest.mark.gpu
def test_device_generator_has_the_reference_moments():
    from evo_b200.bsc import BSC
    from evo_b200.sssc import SSSC

    n = 400000
    D, H = Z["bsc_W"].shape
    gen = BSC(D, H, 4).generate_data({"W": Z["bsc_W"], "pies": Z["bsc_pies"], "sigma": float(Z["bsc_sigma"])}, n, seed=3)
    _check_moments("bsc_", gen, n, False)
    n = 200000
    D, H = Z["es3c_W"].shape
    th = {k: Z["es3c_" + k] for k in ("W", "pies", "mus", "Psi")}
    th["sigma2"] = np.float64(Z["es3c_sigma2"])
    gen = SSSC(D, H, 4).generate_data(th, n, seed=4)
    _check_moments("es3c_", gen, n, True)


@pytest.mark.gpu
def test_standard_init_statistics_and_noise_on_device():
    """standard_init (_models.py:206-283, sssc.py:104-197): data mean / spread from `evo_column_moments`, dictionary
    = data mean + N(0, (sigma/4)^2) from `evo_normal_fill` -- against the reference's formulas in NumPy."""
    from evo_b200.bsc import BSC
    from evo_b200.sssc import SSSC

    rng = np.random.RandomState(4)
    N, D, H = 3000, 20, 50
    Y = (rng.normal(size=(N, D)) * rng.uniform(0.5, 3, size=D) + rng.normal(size=D) * 4).astype(np.float32).astype(np.float64)
    X = rng.random_sample((N, D)) > 0.3
    # BSC, complete data (_models.py:225-240): mean over n, sigma = sqrt(mean_d var_d)
    np.random.seed(3)
    th = BSC(D, H, 8).standard_init({"y": Y, "x_infr": None})
    mean = Y.mean(0)
    sig = np.sqrt(Y.var(0).mean())
    assert abs(th["sigma"] - sig) < 1e-9 * sig and th["pi"] == 1.0 / H
    noise = th["W"] - mean[:, None]
    assert abs(noise.mean()) < 5 * (sig / 4) / np.sqrt(D * H) and abs(noise.std() - sig / 4) < 0.1 * sig / 4
    np.random.seed(3)
    th2 = BSC(D, H, 8).standard_init({"y": Y, "x_infr": None})
    # same Philox seed from np.random -> same noise; the data mean is summed with unordered f64 atomics (last bits)
    assert np.abs(th["W"] - th2["W"]).max() < 1e-12 * np.abs(th["W"]).max(), "np.random.seed makes the initialisation reproducible"
    # BSC, incomplete data (_models.py:241-262): masked mean, sigma over the observed entries
    Yn = np.where(X, Y, np.nan)
    th = BSC(D, H, 8).standard_init({"y": Yn, "x_infr": X})
    mean = np.nanmean(Yn, 0)
    sig = np.sqrt(np.nansum((Yn - mean) ** 2) / X.sum())
    assert abs(th["sigma"] - sig) < 1e-9 * sig
    assert np.abs((th["W"] - mean[:, None]).mean(1)).max() < 6 * (sig / 4) / np.sqrt(H)
    # ES3C (sssc.py:143-181): complete -> mean diagonal of the unbiased covariance + 1e-3
    th = SSSC(D, H, 8).standard_init({"y": Y, "x_infr": None})
    want = np.diag(np.cov(Y.T)).mean() + 0.001
    assert abs(float(th["sigma2"]) - want) < 1e-9 * want
    th = SSSC(D, H, 8).standard_init({"y": Yn, "x_infr": X})
    ym = np.nansum(Yn, 0) / N
    want = np.nansum((Yn - ym) ** 2) / X.sum() + 0.001
    assert abs(float(th["sigma2"]) - want) < 1e-9 * want
