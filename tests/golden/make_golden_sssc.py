"""Generate the ES3C (SSSC) golden fixtures by RUNNING THE UNMODIFIED REFERENCE and pin
`oracle/sssc.py` against it (same technique as make_golden.py: recorded parent/children tapes are
replayed into the oracle, which must reproduce K sets, F and every parameter).

    python tests/golden/make_golden_sssc.py     # needs /root/reference; writes tests/golden/sssc_*.npz
"""
import copy
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

from make_golden import FAIL, Replay, SSSC, Tape, check, close, ref_init_states  # noqa: E402

from oracle import sssc as OS  # noqa: E402
from oracle import variational as OV  # noqa: E402

PKEYS = ("W", "pies", "mus", "Psi", "sigma2")


def f32(a):
    return np.asarray(a, dtype=np.float64).astype(np.float32).astype(np.float64)


def random_theta(rng, D, H, full_psi=True):
    A = rng.normal(size=(H, H)) * 0.2
    Psi = np.eye(H) + (A @ A.T if full_psi else 0.0)
    return {"W": f32(rng.normal(size=(D, H)) * 2), "pies": rng.uniform(0.1, 0.4, size=H), "mus": rng.normal(size=H) + 1.0,
            "Psi": Psi, "sigma2": np.float64(0.6)}


def golden_lpj():
    rng = np.random.RandomState(20)
    D, H, S = 14, 12, 30
    th = random_theta(rng, D, H)
    Y = f32(rng.normal(size=(5, D)) * 2)
    X = rng.rand(5, D) > 0.3
    X[0] = True
    X[1] = True
    states = rng.rand(5, S, H) < 0.25
    states[2, 0] = False
    states[3, 1] = True
    m = SSSC(D, H, S)
    ss = ref_init_states(5, S, H, "fit", "randflip", 2, 1, 1,
                         permanent={"background": False, "allzero": True, "singletons": False})
    data = {"y": Y, "x_infr": X}
    m.E_step_precompute(th, ss, data)
    out, out0 = np.zeros((5, S)), np.zeros(5)
    for n in range(5):
        data["this_y"], data["this_x_infr"] = Y[n], X[n]
        ss["this_states"] = states[n]
        ss["storage"] = {"storagekeys": (), "counts_norm": 0, "counts": 0}
        out[n] = m.log_pseudo_joint(th, ss, data)
        out0[n] = m.log_pseudo_joint_permanent_states(th, ss, data)[0]
    oth = OS.precompute({k: np.array(th[k]) for k in PKEYS}, D, 5, int(X.sum()))
    mine = np.array([OS.lpj_states(states[n], oth, Y[n], X[n])[0] for n in range(5)])
    mine0 = np.array([OS.lpj_allzero(Y[n], X[n], oth["sigma2_inv"])[0] for n in range(5)])
    check("sssc lpj (masked + complete rows)", close(mine, out, 1e-11) and close(mine0, out0))
    check("sssc precompute", close(oth["ljc"], th["ljc"]) and close(oth["pil_bar"], th["pil_bar"])
          and close(oth["sigma2_inv"], th["sigma2_inv"]))
    np.savez_compressed(os.path.join(HERE, "sssc_lpj.npz"), Y=Y, X=X, states=states, lpj=out, lpj_allzero=out0,
                        ljc=th["ljc"], **{k: np.array(th[k]) for k in PKEYS})


def run_sssc(seed, D, H, N, S, ea, iters, name, to_learn=PKEYS, missing=0.0, permanent=None, exact=False):
    np.random.seed(seed)
    rng = np.random.RandomState(seed + 100)
    model = SSSC(D, H, S, to_learn=list(to_learn))
    gen = random_theta(rng, D, H, full_psi=False)
    gen["pies"] = np.full(H, 2.0 / H)
    Y = f32(model.generate_data(gen, N)["y"])
    if missing > 0:
        Y[np.random.random(Y.shape) < missing] = np.nan
    X = ~np.isnan(Y)
    data = {"y": Y.copy(), "x_infr": X.copy(), "x": X.copy() if missing > 0 else np.zeros_like(X)}
    theta0 = model.check_params(model.standard_init(data))
    theta0["W"] = f32(theta0["W"])
    theta0 = {k: np.array(theta0[k]) for k in PKEYS}
    sel, mut, P, C, G = ea
    bf = 1.0 / H if "sparse" in mut else None
    cont = ref_init_states(N, S, H, sel, mut, P, C, G, bitflip_prob=bf, permanent=copy.deepcopy(permanent))
    ss0 = cont["ss"].copy()
    tape = Tape(cont)
    theta = copy.deepcopy(theta0)
    recon = True
    Fs, nun, nsub, thetas = [], [], [], []
    for it in range(iters):
        F, a, b, theta = model.step(theta, cont, data, do_reconstruction=recon)
        Fs.append(F), nun.append(a), nsub.append(b)
        thetas.append({k: np.array(theta[k]) for k in PKEYS})
        tape.next_iteration()

    ocont = OV.init_states(1, S, H, sel, mut, P, C, G, bitflip_prob=bf, permanent=copy.deepcopy(permanent))
    ocont["ss"] = ss0.copy()
    ocont["lpj"] = np.empty((N, ss0.shape[1] + ocont["S_perm"]))
    rep = Replay(tape, ocont)
    odata = {"y": Y.copy(), "x_infr": X.copy(), "x": data["x"].copy()}
    oth = copy.deepcopy(theta0)
    ok = True
    for it in range(iters):
        F, a, b, oth = OS.em_step(oth, ocont, odata, D, H, ss0.shape[1], to_learn=to_learn, do_reconstruction=recon)
        ok &= close(F, Fs[it], 1e-10) and close(a, nun[it]) and close(b, nsub[it])
        for k in PKEYS:
            ok &= close(oth[k], thetas[it][k], 1e-7)
    ok &= rep.ok and np.array_equal(ocont["ss"], cont["ss"]) and close(ocont["lpj"], cont["lpj"], 1e-9)
    ok &= close(odata["y_reconstructed"], data["y_reconstructed"], 1e-9)
    check("oracle replay == reference: %s" % name, bool(ok))

    ch, cnt = (tape.children_tensor(iters, N, G, H) if tape.mut else
               (np.zeros((iters, N, G, 1, H), bool), np.zeros((iters, N, G), np.int32)))
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"),
        D=D, H=H, N=N, S=ss0.shape[1], P=P, C=C, G=G, iters=iters, selection=sel, mutation=mut,
        allzero=bool(permanent and permanent["allzero"]), missing=missing, to_learn=",".join(to_learn),
        Y=Y, X=X, Xrec=data["x"], ss0=np.packbits(ss0, axis=-1),
        children=np.packbits(ch, axis=-1), child_count=cnt,
        F=np.array(Fs), S_nunique=np.array(nun), S_sub=np.array(nsub),
        ss_final=np.packbits(cont["ss"], axis=-1), lpj_final=cont["lpj"], y_reconstructed=data["y_reconstructed"],
        **{k + "0": theta0[k] for k in PKEYS}, **{k: np.array([t[k] for t in thetas]) for k in PKEYS},
    )


def main():
    golden_lpj()
    run_sssc(0, 10, 6, 60, 2 ** 6, ("fit", "randflip", 3, 1, 1), 4, "sssc_exact", exact=True)
    run_sssc(1, 12, 10, 80, 12, ("fit", "randflip", 4, 2, 2), 5, "sssc_replay")
    run_sssc(2, 12, 10, 80, 12, ("fit", "randflip", 4, 2, 2), 4, "sssc_missing", to_learn=("W", "pies", "sigma2"),
             missing=0.25)
    run_sssc(3, 9, 40, 50, 10, ("rand", "sparseflip", 3, 2, 2), 3, "sssc_allzero_h40",
             permanent={"background": False, "allzero": True, "singletons": False})
    if FAIL:
        print("MISMATCHES:", FAIL)
        sys.exit(1)
    print("all ES3C oracle pins hold; fixtures written to", HERE)


if __name__ == "__main__":
    main()
