"""Free-running bars-test trajectories of the UNMODIFIED reference (examples/bars-test/main.py:61-176 with
its default hyper-parameters, params.py: fit / randflip, 10 parents, 1 child, 1 generation, 40 epochs;
H = 10, D = 25, N = 1000, K = 20 = BASELINE.json configs[0]).

The evolutionary operators draw from NumPy's legacy generator, the CUDA path from a Philox counter
stream, so a free run cannot be compared draw for draw.  This script records what CAN be compared:
the same data, the same initial parameters, and the reference's free-energy trajectory under several
operator seeds -- their spread is the tolerance `tests/test_bars_gpu.py` holds the CUDA path to.

Runs only in the build container (needs /root/reference and the mpi4py stand-in).
    python tests/golden/make_golden_bars_free.py          # writes tests/golden/bsc_bars_free.npz
    python tests/golden/make_golden_bars_free.py es3c     # writes tests/golden/sssc_bars_free.npz (--algo es3c
                                                          # of the example: W, pies, sigma2, mus, Psi)
"""
import copy
import os
import sys
import time
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("EVO_REFERENCE", "/root/reference")
sys.path.insert(0, os.path.join(HERE, "_standin"))
sys.path.insert(0, REF)
sys.path.insert(0, ROOT)
warnings.simplefilter("ignore")

from evo.models import BSC, SSSC  # noqa: E402  (the reference)
from evo.variational.utils import init_states  # noqa: E402

H, N, K, EPOCHS = 10, 1000, 20, 40
EA = ("fit", "randflip", 10, 1, 1)
SEEDS = (11, 12, 13, 14)


def bars_dict(H):
    R = H // 2
    W = np.zeros((R, R, H))
    for i in range(R):
        W[i, :, i] = 1.0
        W[:, i, R + i] = 1.0
    return W.reshape(R * R, H)


def main():
    D = (H // 2) ** 2
    model = BSC(D, H, K)
    np.random.seed(7)
    theta_gen = {"W": 10.0 * bars_dict(H), "pi": 2.0 / H, "sigma": 1.0}
    data = model.generate_data(theta_gen, N)
    # float32-representable data and initial W: the device holds Y as float32
    Y = data["y"].astype(np.float32).astype(np.float64)
    my_data = {"y": Y, "x_infr": np.ones_like(Y, dtype=bool)}
    theta0 = model.check_params(model.standard_init(my_data))
    theta0["W"] = theta0["W"].astype(np.float32).astype(np.float64)
    cont = init_states(N, K, H, *EA)
    L_gen = model.free_energy(my_data, model.check_params(copy.deepcopy(theta_gen)), cont, full=True)

    F = np.zeros((len(SEEDS), EPOCHS))
    L_end = np.zeros(len(SEEDS))
    W_end = np.zeros((len(SEEDS), D, H))
    pi_end, sigma_end = np.zeros(len(SEEDS)), np.zeros(len(SEEDS))
    for i, seed in enumerate(SEEDS):
        np.random.seed(seed)
        cont = init_states(N, K, H, *EA)
        theta = copy.deepcopy(theta0)
        t0 = time.time()
        for e in range(EPOCHS):
            F[i, e], _, _, theta = model.step(theta, cont, my_data)
        L_end[i] = model.free_energy(my_data, copy.deepcopy(theta), cont, full=True)
        W_end[i], pi_end[i], sigma_end[i] = theta["W"], theta["pi"], theta["sigma"]
        print("seed %d: F[0]=%.4f F[9]=%.4f F[-1]=%.4f L(theta_end)=%.4f L_gen=%.4f  (%.1f s)"
              % (seed, F[i, 0], F[i, 9], F[i, -1], L_end[i], L_gen, time.time() - t0))
    np.savez_compressed(
        os.path.join(HERE, "bsc_bars_free.npz"),
        Y=Y.astype(np.float32), W_gen=theta_gen["W"], pi_gen=theta_gen["pi"], sigma_gen=theta_gen["sigma"],
        W0=theta0["W"], pi0=float(theta0["pi"]), sigma0=float(theta0["sigma"]),
        F_ref=F, L_end_ref=L_end, L_gen=float(L_gen), W_end_ref=W_end, pi_end_ref=pi_end, sigma_end_ref=sigma_end,
        ea=np.array(EA[2:], dtype=np.int64), K=np.int64(K), seeds=np.array(SEEDS))
    print("spread of F over seeds per epoch (max-min):", np.round(F.max(0) - F.min(0), 4))


def main_es3c():
    D = (H // 2) ** 2
    model = SSSC(D, H, K)
    np.random.seed(7)
    theta_gen = {"W": 10.0 * bars_dict(H), "pies": np.ones(H) * 2.0 / H, "sigma2": np.array(1.0),
                 "mus": np.zeros(H), "Psi": np.eye(H)}
    data = model.generate_data(theta_gen, N)
    Y = data["y"].astype(np.float32).astype(np.float64)
    my_data = {"y": Y, "x_infr": np.ones_like(Y, dtype=bool)}
    theta0 = model.check_params(model.standard_init(my_data))
    theta0["W"] = theta0["W"].astype(np.float32).astype(np.float64)
    cont = init_states(N, K, H, *EA)
    L_gen = model.free_energy(my_data, model.check_params(copy.deepcopy(theta_gen)), cont, full=True)
    seeds = SEEDS[:3]
    F = np.zeros((len(seeds), EPOCHS))
    L_end = np.zeros(len(seeds))
    end = {k: [] for k in ("W", "pies", "sigma2", "mus", "Psi")}
    for i, seed in enumerate(seeds):
        np.random.seed(seed)
        cont = init_states(N, K, H, *EA)
        theta = copy.deepcopy(theta0)
        t0 = time.time()
        for e in range(EPOCHS):
            F[i, e], _, _, theta = model.step(theta, cont, my_data)
        L_end[i] = model.free_energy(my_data, copy.deepcopy(theta), cont, full=True)
        for k in end:
            end[k].append(np.array(theta[k], dtype=np.float64))
        print("seed %d: F[0]=%.4f F[9]=%.4f F[-1]=%.4f L(theta_end)=%.4f L_gen=%.4f  (%.1f s)"
              % (seed, F[i, 0], F[i, 9], F[i, -1], L_end[i], L_gen, time.time() - t0), flush=True)
    out = {"%s_end_ref" % k: np.stack(v) for k, v in end.items()}
    out.update({"%s0" % k: np.asarray(theta0[k], dtype=np.float64) for k in end})
    np.savez_compressed(
        os.path.join(HERE, "sssc_bars_free.npz"), Y=Y.astype(np.float32), W_gen=theta_gen["W"],
        F_ref=F, L_end_ref=L_end, L_gen=float(L_gen), ea=np.array(EA[2:], dtype=np.int64), K=np.int64(K),
        seeds=np.array(seeds), **out)
    print("spread of F over seeds per epoch (max-min):", np.round(F.max(0) - F.min(0), 4))


if __name__ == "__main__":
    main_es3c() if sys.argv[1:] == ["es3c"] else main()
