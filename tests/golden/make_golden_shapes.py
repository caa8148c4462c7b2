"""Golden fixtures AT THE BASELINE SHAPES (cfg2 / cfg5 BSC, cfg3 / cfg4 ES3C) and for the ES3C evaluator's
larger systems, produced by RUNNING THE UNMODIFIED REFERENCE (same recording technique as make_golden.py).

    python tests/golden/make_golden_shapes.py     # needs /root/reference; writes tests/golden/shape_*.npz,
                                                  # sssc_lpj_dense.npz

What differs from the small fixtures: K = 64 / 128 states of 10 / 16 / 32 words, 64-128 children per
generation, two generations -- the hash-table sizes, thread counts and shared-memory layouts the bench
runs.  N is small (the reference needs 8 ms .. 0.2 s per datapoint at these shapes), so the dictionary W
(and Psi) cannot be re-estimated (H > N: the reference's lstsq / inv would be rank deficient and hit its
random-noise fallbacks); `to_learn` is restricted to the scalar / per-latent parameters and W stays at its
initial value.  Large inputs are regenerated from a legacy `np.random.RandomState` seed by
`shape_inputs()` below (bit-stable across NumPy versions) and pinned by a SHA-1 stored in the fixture,
so the files stay small.
"""
import copy
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

from make_golden import BSC, FAIL, Replay, SSSC, Tape, check, close, ref_init_states  # noqa: E402

from oracle import bsc as OB  # noqa: E402
from oracle import sssc as OS  # noqa: E402
from oracle import variational as OV  # noqa: E402

PKEYS = ("W", "pies", "mus", "Psi", "sigma2")


from shape_inputs import f32, sha, shape_inputs  # noqa: E402,F401


def shape_bsc(name, seed, D, H, N, S, ea, iters, act, p_init):
    inp = shape_inputs("bsc", seed, D, H, N, act)
    Y, X, W0 = inp["Y"], inp["X"], inp["W0"]
    to_learn = ("pi", "sigma")
    np.random.seed(seed)
    model = BSC(D, H, S, to_learn=list(to_learn))
    data = {"y": Y.copy(), "x_infr": X.copy()}
    theta0 = model.check_params({"W": W0.copy(), "pi": act / H, "sigma": 1.2})
    sel, mut, P, C, G = ea
    cont = ref_init_states(N, S, H, sel, mut, P, C, G, p_init_Kn=p_init)
    ss0 = cont["ss"].copy()
    tape = Tape(cont)
    theta = copy.deepcopy(theta0)
    Fs, nun, nsub, pis, sigmas = [], [], [], [], []
    for it in range(iters):
        F, a, b, theta = model.step(theta, cont, data)
        Fs.append(F), nun.append(a), nsub.append(b), pis.append(float(theta["pi"])), sigmas.append(float(theta["sigma"]))
        tape.next_iteration()
    # oracle replay
    ocont = OV.init_states(1, S, H, sel, mut, P, C, G)
    ocont["ss"] = ss0.copy()
    ocont["lpj"] = np.empty((N, S))
    rep = Replay(tape, ocont)
    oth = copy.deepcopy(theta0)
    odata = {"y": Y.copy(), "x_infr": X.copy()}
    ok = True
    for it in range(iters):
        F, a, b, oth = OB.step(oth, ocont, odata, D, H, S, to_learn=to_learn)
        ok &= close(F, Fs[it], 1e-11) and close(a, nun[it]) and close(b, nsub[it])
        ok &= close(oth["pi"], pis[it], 1e-10) and close(oth["sigma"], sigmas[it], 1e-10)
    ok &= rep.ok and np.array_equal(ocont["ss"], cont["ss"]) and close(ocont["lpj"], cont["lpj"], 1e-11)
    check("oracle replay == reference: %s" % name, bool(ok))
    ch, cnt = tape.children_tensor(iters, N, G, H)
    act_mean = float(cont["ss"].sum() / (N * S))
    print("   %s: mean |s| of the final K sets %.2f, max %d, S_nunique %s" % (name, act_mean, cont["ss"].sum(-1).max(), nun))
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"),
        model="bsc", seed=seed, act=act, D=D, H=H, N=N, S=S, P=P, C=C, G=G, iters=iters, selection=sel, mutation=mut,
        allzero=False, missing=0.0, to_learn=",".join(to_learn), p_init=p_init,
        sha_Y=sha(Y), sha_W0=sha(W0), pi0=theta0["pi"], sigma0=theta0["sigma"], ss0=np.packbits(ss0, axis=-1),
        children=np.packbits(ch, axis=-1), child_count=cnt,
        F=np.array(Fs), S_nunique=np.array(nun), S_sub=np.array(nsub), pi=np.array(pis), sigma=np.array(sigmas),
        ss_final=np.packbits(cont["ss"], axis=-1), lpj_final=cont["lpj"],
    )


def shape_sssc(name, seed, D, H, N, S, ea, iters, act, p_init, missing, to_learn):
    inp = shape_inputs("es3c", seed, D, H, N, act, missing)
    Y, X, W0 = inp["Y"].copy(), inp["X"], inp["W0"]
    Yn = Y.copy()
    Yn[~X] = np.nan
    np.random.seed(seed)
    model = SSSC(D, H, S, to_learn=list(to_learn))
    xrec = X.copy() if missing > 0 else np.zeros_like(X)
    data = {"y": Yn.copy(), "x_infr": X.copy(), "x": xrec.copy()}
    theta0 = {"W": W0.copy(), "pies": inp["pies0"].copy(), "mus": inp["mus0"].copy() if "mus" in to_learn else np.ones(H),
              "Psi": inp["Psi0"].copy() if missing == 0 else np.eye(H), "sigma2": np.float64(1.3)}
    theta0 = model.check_params(theta0)
    theta0 = {k: np.array(theta0[k]) for k in PKEYS}
    sel, mut, P, C, G = ea
    cont = ref_init_states(N, S, H, sel, mut, P, C, G, p_init_Kn=p_init)
    ss0 = cont["ss"].copy()
    tape = Tape(cont)
    theta = copy.deepcopy(theta0)
    Fs, nun, nsub, thetas = [], [], [], []
    for it in range(iters):
        F, a, b, theta = model.step(theta, cont, data, do_reconstruction=True)
        Fs.append(F), nun.append(a), nsub.append(b)
        thetas.append({k: np.array(theta[k]) for k in ("pies", "mus", "sigma2")})
        tape.next_iteration()
    ocont = OV.init_states(1, S, H, sel, mut, P, C, G)
    ocont["ss"] = ss0.copy()
    ocont["lpj"] = np.empty((N, S))
    rep = Replay(tape, ocont)
    odata = {"y": Yn.copy(), "x_infr": X.copy(), "x": xrec.copy()}
    oth = copy.deepcopy(theta0)
    ok = True
    for it in range(iters):
        F, a, b, oth = OS.em_step(oth, ocont, odata, D, H, S, to_learn=to_learn, do_reconstruction=True)
        ok &= close(F, Fs[it], 1e-10) and close(a, nun[it]) and close(b, nsub[it])
        for k in ("pies", "mus", "sigma2"):
            ok &= close(oth[k], thetas[it][k], 1e-7)
    ok &= rep.ok and np.array_equal(ocont["ss"], cont["ss"]) and close(ocont["lpj"], cont["lpj"], 1e-9)
    ok &= close(odata["y_reconstructed"], data["y_reconstructed"], 1e-9)
    check("oracle replay == reference: %s" % name, bool(ok))
    ch, cnt = tape.children_tensor(iters, N, G, H)
    print("   %s: mean |s| of the final K sets %.2f, max %d, S_nunique %s" % (
        name, cont["ss"].sum() / (N * S), cont["ss"].sum(-1).max(), nun))
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"),
        model="es3c", seed=seed, act=act, D=D, H=H, N=N, S=S, P=P, C=C, G=G, iters=iters, selection=sel, mutation=mut,
        allzero=False, missing=missing, to_learn=",".join(to_learn), p_init=p_init,
        sha_Y=sha(inp["Y"]), sha_W0=sha(W0), sha_Psi0=sha(theta0["Psi"]), sigma20=theta0["sigma2"],
        pies0=theta0["pies"], mus0=theta0["mus"], ss0=np.packbits(ss0, axis=-1),
        children=np.packbits(ch, axis=-1), child_count=cnt,
        F=np.array(Fs), S_nunique=np.array(nun), S_sub=np.array(nsub),
        ss_final=np.packbits(cont["ss"], axis=-1), lpj_final=cont["lpj"],
        y_reconstructed=np.asarray(data["y_reconstructed"], dtype=np.float32),
        **{k: np.array([t[k] for t in thetas]) for k in ("pies", "mus", "sigma2")},
    )


def golden_lpj_dense():
    """SSSC.log_pseudo_joint (sssc.py:242-326) on states with 1..8, 9..16, 17..32 and 33..40 active latents,
    H = 80 (3 words), complete and incomplete rows, dense non-symmetric Psi."""
    rng = np.random.RandomState(31)
    D, H, S = 48, 80, 40
    A = rng.normal(size=(H, H)) * 0.08
    Psi = np.eye(H) + A @ A.T + 0.02 * rng.normal(size=(H, H))  # not symmetric: the reference's update does not keep symmetry
    th = {"W": f32(rng.normal(size=(D, H))), "pies": rng.uniform(0.05, 0.3, size=H), "mus": rng.normal(size=H) * 0.5 + 1.0,
          "Psi": Psi, "sigma2": np.float64(0.8)}
    N = 4
    Y = f32(rng.normal(size=(N, D)) * 3)
    X = rng.rand(N, D) > 0.3
    X[0] = True
    X[1] = True
    sizes = np.concatenate([rng.randint(1, 9, 10), rng.randint(9, 17, 10), rng.randint(17, 33, 12), rng.randint(33, 41, 8)])
    states = np.zeros((N, S, H), dtype=bool)
    for n in range(N):
        for k in range(S):
            states[n, k, rng.choice(H, size=sizes[k], replace=False)] = True
    m = SSSC(D, H, S)
    ss = ref_init_states(N, S, H, "fit", "randflip", 2, 1, 1)
    data = {"y": Y, "x_infr": X}
    m.E_step_precompute(th, ss, data)
    out = np.zeros((N, S))
    for n in range(N):
        data["this_y"], data["this_x_infr"] = Y[n], X[n]
        ss["this_states"] = states[n]
        ss["storage"] = {"storagekeys": (), "counts_norm": 0, "counts": 0}
        out[n] = m.log_pseudo_joint(th, ss, data)
    oth = OS.precompute({k: np.array(th[k]) for k in PKEYS}, D, N, int(X.sum()))
    mine = np.array([OS.lpj_states(states[n], oth, Y[n], X[n])[0] for n in range(N)])
    check("sssc lpj, 1..40 active latents", close(mine, out, 1e-9))
    np.savez_compressed(os.path.join(HERE, "sssc_lpj_dense.npz"), Y=Y, X=X, states=np.packbits(states, axis=-1), H=H,
                        sizes=sizes, lpj=out, **{k: np.array(th[k]) for k in PKEYS})


def main():
    only = sys.argv[1:]
    if not only or "lpj" in only:
        golden_lpj_dense()
    if not only or "cfg2" in only:
        shape_bsc("shape_cfg2", 42, 256, 300, 32, 64, ("fit", "randflip", 8, 8, 2), 3, act=5.0, p_init=4.0 / 300)
    if not only or "cfg5" in only:
        shape_bsc("shape_cfg5", 45, 1024, 1024, 8, 128, ("fit", "randflip", 16, 8, 2), 2, act=8.0, p_init=8.0 / 1024)
    if not only or "cfg3" in only:
        shape_sssc("shape_cfg3", 43, 144, 512, 16, 64, ("fit", "randflip", 8, 8, 2), 3, act=4.0, p_init=3.0 / 512,
                   missing=0.0, to_learn=("pies", "mus", "sigma2"))
    if not only or "cfg4" in only:
        shape_sssc("shape_cfg4", 44, 144, 512, 16, 64, ("fit", "randflip", 8, 8, 2), 3, act=4.0, p_init=3.0 / 512,
                   missing=0.3, to_learn=("pies", "sigma2"))
    if FAIL:
        print("MISMATCHES:", FAIL)
        sys.exit(1)
    print("all pins hold; fixtures written to", HERE)


if __name__ == "__main__":
    main()
