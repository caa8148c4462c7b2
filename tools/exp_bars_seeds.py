"""How often does a free bars run reach the optimum?  (GPU; compare with the reference's own rate in
tests/golden/bsc_bars_free.npz)   usage: python tools/exp_bars_seeds.py [n_seeds]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from evo_b200 import eem  # noqa: E402
from evo_b200.bsc import BSC  # noqa: E402
from evo_b200.variational import init_states  # noqa: E402

n_seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 64
z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "bsc_bars_free.npz"))
Y = z["Y"].astype(np.float64)
N, D = Y.shape
H, K = z["W0"].shape[1], int(z["K"])
model = BSC(D, H, K)
data = {"y": Y, "x_infr": np.ones_like(Y, dtype=bool)}
finals, traj = [], []
for seed in range(n_seeds):
    th = model.check_params({"W": z["W0"].copy(), "pi": float(z["pi0"]), "sigma": float(z["sigma0"])})
    cont = init_states(N, K, H, "fit", "randflip", 10, 1, 1, device=model.device, seed=5000 + seed)
    th, hist = eem.train(model, th, cont, data, 40)
    finals.append(hist["F"][-1])
    traj.append(hist["F"])
finals = np.array(finals)
traj = np.array(traj)
print("reached -40.2: %d of %d" % ((np.abs(finals + 40.1999) < 2e-3).sum(), n_seeds))
print("finals sorted:", np.round(np.sort(finals), 3))
print("mean trajectory:", np.round(traj.mean(0), 3))
print("reference trajectories:", np.round(z["F_ref"], 3))

# ---- the same for --algo es3c (fewer seeds; prints whole trajectories) ---------------------------
from evo_b200.sssc import SSSC  # noqa: E402

z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", "sssc_bars_free.npz"))
Y = z["Y"].astype(np.float64)
model = SSSC(D, H, K)
data = {"y": Y, "x_infr": np.ones_like(Y, dtype=bool)}
traj = []
for seed in range(min(n_seeds, 16)):
    th = model.check_params({"W": z["W0"].copy(), "pies": z["pies0"].copy(), "sigma2": np.float64(z["sigma20"]),
                             "mus": z["mus0"].copy(), "Psi": z["Psi0"].copy()})
    cont = init_states(N, K, H, "fit", "randflip", 10, 1, 1, device=model.device, seed=7000 + seed)
    th, hist = eem.train(model, th, cont, data, 40)
    traj.append(hist["F"])
traj = np.array(traj)
np.set_printoptions(linewidth=220, precision=2, suppress=True)
print("ES3C trajectories (CUDA):")
print(traj)
print("ES3C reference:")
print(z["F_ref"])
