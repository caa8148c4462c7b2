"""Moments of the reference's `generate_data` (evo/models/_models.py:73-99 with bsc.py:27-57 / sssc.py:66-102),
recorded by RUNNING THE UNMODIFIED REFERENCE: the device generator (`evo_generate_data`) and its NumPy definition
(oracle/datagen.py) use a different random stream (Philox instead of the global MT19937), so parity of the
generative model is distributional.

    python tests/golden/make_golden_datagen.py     # needs /root/reference; writes tests/golden/datagen_moments.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

from make_golden import BSC, SSSC  # noqa: E402  (the reference, through the mpi4py stand-in)


def moments(out, es3c):
    y, s = out["y"], out["s"]
    m = {"s_mean": s.mean(0), "y_mean": y.mean(0), "y_cov": np.cov(y.T), "n": np.float64(y.shape[0]),
         "resid_var": (y - out["y_mean"]).var(0)}
    if es3c:
        z = out["z"]
        H = s.shape[1]
        zm = np.array([z[s[:, h], h].mean() for h in range(H)])
        zz = np.zeros((H, H))
        cnt = np.zeros((H, H))
        for h in range(H):
            for g in range(H):
                both = s[:, h] & s[:, g]
                cnt[h, g] = both.sum()
                zz[h, g] = (z[both, h] * z[both, g]).mean() if both.any() else 0.0
        m.update({"z_mean": zm, "zz": zz, "zz_cnt": cnt})
    return m


def main():
    rng = np.random.RandomState(11)
    out = {}
    D, H, N = 8, 12, 200000
    th = {"W": rng.normal(size=(D, H)), "pies": rng.uniform(0.05, 0.4, size=H), "sigma": 0.7}
    np.random.seed(0)
    for k, v in moments(BSC(D, H, 4).generate_data(th, N), False).items():
        out["bsc_" + k] = v
    out.update({"bsc_W": th["W"], "bsc_pies": th["pies"], "bsc_sigma": th["sigma"]})
    D, H, N = 6, 8, 60000
    A = rng.normal(size=(H, H)) * 0.4
    th = {"W": rng.normal(size=(D, H)), "pies": np.full(H, 0.35), "mus": rng.normal(size=H) + 0.5,
          "Psi": 0.5 * np.eye(H) + A @ A.T, "sigma2": np.float64(0.49)}
    np.random.seed(1)
    for k, v in moments(SSSC(D, H, 4).generate_data(th, N), True).items():
        out["es3c_" + k] = v
    out.update({"es3c_" + k: np.asarray(v) for k, v in th.items()})
    np.savez_compressed(os.path.join(HERE, "datagen_moments.npz"), **out)
    print("written", os.path.join(HERE, "datagen_moments.npz"))


if __name__ == "__main__":
    main()
