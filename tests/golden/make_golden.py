"""Generate the golden fixtures in this directory by RUNNING THE UNMODIFIED REFERENCE, and pin the
CPU oracle (`oracle/`) against it while doing so.

Runs only in the build container: needs /root/reference (read-only) and the single-rank mpi4py
stand-in under tests/golden/_standin (real mpi4py is not installed).  The GPU box
never runs this; it only reads the committed *.npz files.

    python tests/golden/make_golden.py            # writes tests/golden/*.npz, exits non-zero on any
                                                  # oracle-vs-reference mismatch

Technique: the reference fetches `parent_selection` and `mutation_algorithm`
as callables from the K-state container, so they are wrapped by recorders; the oracle is then driven
by replaying callables that also assert their *inputs* (the parent pools) equal the recorded ones.
"""
import copy
import os
import sys
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
from evo.variational.utils import init_states as ref_init_states, vary_Kn as ref_vary_Kn  # noqa: E402

from oracle import bsc as OB  # noqa: E402
from oracle import variational as OV  # noqa: E402

FAIL = []


def check(name, ok):
    print(("ok   " if ok else "FAIL ") + name)
    if not ok:
        FAIL.append(name)


def close(a, b, tol=1e-12):
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    return a.shape == b.shape and np.all(np.abs(a - b) <= tol * (1.0 + np.abs(b)))


def bars_dict(H):
    """Horizontal + vertical bars dictionary, (R*R, H) with R = H // 2 -- the generating W of
    examples/bars-test (utils.py:7-36), restated."""
    R = H // 2
    W = np.zeros((R, R, H))
    for i in range(R):
        W[i, :, i] = 1.0
        W[:, i, R + i] = 1.0
    return W.reshape(R * R, H)


# ------------------------------------------------------------------------------------------
class Tape:
    """Records every parent_selection / mutation_algorithm call of one reference run."""

    def __init__(self, container):
        self.sel, self.mut = [], []
        self.it = 0
        self.n = -1
        self.g = 0
        self._sel, self._mut = container["parent_selection"], container["mutation_algorithm"]
        self.container = container
        container["parent_selection"] = self.select
        container["mutation_algorithm"] = self.mutate

    def select(self, cands, n_parents, lpj):
        if np.shares_memory(cands, self.container["ss"]):
            self.n += 1
            self.g = 0
        else:
            self.g += 1
        out = self._sel(cands, n_parents, lpj)
        self.sel.append((self.it, self.n, self.g, np.array(cands), int(n_parents), np.array(lpj), np.array(out)))
        return out

    def mutate(self, parents, n_children, sparseness, p_bf):
        out = self._mut(parents, n_children, sparseness, p_bf)
        self.mut.append((self.it, self.n, self.g, np.array(out)))
        return out

    def next_iteration(self):
        self.it += 1
        self.n = -1

    def children_tensor(self, iters, N, G, H):
        cmax = max(m[3].shape[0] for m in self.mut)
        ch = np.zeros((iters, N, G, cmax, H), dtype=bool)
        cnt = np.zeros((iters, N, G), dtype=np.int32)
        for it, n, g, c in self.mut:
            ch[it, n, g, : c.shape[0], : c.shape[1]] = c
            if c.shape[1] < H:  # background unit is appended after mutation (eas.py:236-239)
                ch[it, n, g, : c.shape[0], -1] = True
            cnt[it, n, g] = c.shape[0]
        return ch, cnt


class Replay:
    """Feeds a recorded tape to the oracle and asserts the oracle asks the same questions."""

    def __init__(self, tape, container, check_inputs=True):
        self.sel, self.mut = list(tape.sel), list(tape.mut)
        self.i = self.j = 0
        self.ok = True
        self.check_inputs = check_inputs
        container["parent_selection"] = self.select
        container["mutation_algorithm"] = self.mutate

    def select(self, cands, n_parents, lpj):
        rec = self.sel[self.i]
        self.i += 1
        if self.check_inputs:
            same = (rec[3].shape == np.asarray(cands).shape and np.array_equal(rec[3], cands)
                    and rec[4] == n_parents and close(rec[5], lpj))
            self.ok &= bool(same)
        return rec[6].copy()

    def mutate(self, parents, n_children, sparseness, p_bf):
        rec = self.mut[self.j]
        self.j += 1
        return rec[3].copy()


def strip(container):
    """Deep copy of a container without the callables."""
    return {k: copy.deepcopy(v) for k, v in container.items()
            if k not in ("parent_selection", "mutation_algorithm")}


# ------------------------------------------------------------------------------------------
def golden_bsc_lpj():
    rng = np.random.RandomState(10)
    D, H, S = 24, 40, 50
    W = rng.normal(size=(D, H))
    Y = rng.normal(size=(6, D)) * 3
    X = rng.rand(6, D) > 0.25
    X[0] = True
    X[1] = True
    states = rng.rand(6, S, H) < 0.12
    states[2, 0] = False  # an all-zero row
    states[3, 1] = True  # an all-one row
    theta = {"W": W, "pi": 0.07, "sigma": 1.3}
    m = BSC(D, H, S)
    data = {"y": Y, "x_infr": X}
    ss = ref_init_states(6, S, H, "fit", "randflip", 2, 1, 1, permanent={"background": False, "allzero": True, "singletons": False})
    m.E_step_precompute(theta, ss, data)
    out = np.zeros((6, S))
    out0 = np.zeros(6)
    for n in range(6):
        data["this_y"], data["this_x_infr"] = Y[n], X[n]
        ss["this_states"] = states[n]
        out[n] = m.log_pseudo_joint(theta, ss, data)
        out0[n] = m.log_pseudo_joint_permanent_states(theta, ss, data)[0]
    th = OB.precompute({"W": W, "pi": 0.07, "sigma": 1.3}, D, H, 6, int(X.sum()))
    mine = np.array([OB.lpj_states(states[n], W, Y[n], X[n], th["pre1"], th["pil_bar"]) for n in range(6)])
    mine0 = np.array([OB.lpj_allzero(Y[n], X[n], th["pre1"])[0] for n in range(6)])
    check("bsc lpj (masked + complete rows)", close(mine, out) and close(mine0, out0))
    check("bsc precompute (incomplete data ljc)", close([th[k] for k in ("pre1", "pil_bar", "piH", "ljc")],
                                                      [theta[k] for k in ("pre1", "pil_bar", "piH", "ljc")]))
    # lpj_reset_check semantics
    bad = np.array([1.0, np.nan, -np.inf, np.inf, -1e308 * 10])
    cnt = {"reset_lpj_isnan": 0, "reset_lpj_smaller_eps_lpj": 0, "reset_lpj_isinf": 0}
    ss.update(cnt)
    r_ref = m.lpj_reset_check(bad.copy(), ss)
    r_me = OB.lpj_reset_check(bad.copy(), cnt)
    check("lpj_reset_check", np.array_equal(r_ref, r_me) and all(ss[k] == cnt[k] for k in cnt))
    np.savez_compressed(os.path.join(HERE, "bsc_lpj.npz"), W=W, Y=Y, X=X, states=states, pi=0.07, sigma=1.3,
                        lpj=out, lpj_allzero=out0, pre1=theta["pre1"], pil_bar=theta["pil_bar"],
                        ljc=theta["ljc"], reset_in=bad, reset_out=r_ref)


def run_bsc(seed, D, H, N, S, ea, iters, W_gen, pi_gen, sigma_gen, permanent=None, missing=0.0,
            exact=False, name=None, to_learn=("W", "pi", "sigma")):
    """Reference run with recording; oracle replay; fixture."""
    np.random.seed(seed)
    model = BSC(D, H, S, to_learn=list(to_learn))
    gen = {"W": W_gen, "pi": pi_gen, "sigma": sigma_gen}
    Y = model.generate_data(gen, N)["y"]
    Y = Y.astype(np.float32).astype(np.float64)  # the CUDA side stores Y as fp32: make that exact
    if missing > 0:
        Y[np.random.random(Y.shape) < missing] = np.nan
    X = ~np.isnan(Y)
    data = {"y": Y.copy(), "x_infr": X.copy()}
    if missing > 0:
        data["x"] = X.copy()
    theta0 = model.check_params(model.standard_init(data))
    theta0["W"] = theta0["W"].astype(np.float32).astype(np.float64)
    sel, mut, P, C, G = ea
    cont = ref_init_states(N, S, H, sel, mut, P, C, G, bitflip_prob=(1.0 / H if "sparse" in mut else None),
                           permanent=copy.deepcopy(permanent))
    ss0 = cont["ss"].copy()
    tape = Tape(cont)
    theta = copy.deepcopy(theta0)
    Fs, nun, nsub, thetas = [], [], [], []
    for it in range(iters):
        F, a, b, theta = model.step(theta, cont, data, do_reconstruction=missing > 0)
        Fs.append(F), nun.append(a), nsub.append(b)
        thetas.append({k: np.array(theta[k]) for k in ("W", "pi", "sigma")})
        tape.next_iteration()

    # ---- oracle replay (quirks on: must reproduce pools, K sets, F, theta)
    ocont = OV.init_states(1, S, H, sel, mut, P, C, G, bitflip_prob=cont["bitflip_prob"],
                           permanent=copy.deepcopy(permanent)) if not exact else OV.init_states(
        1, S, H, sel, mut, P, C, G, permanent=copy.deepcopy(permanent))
    ocont["ss"] = ss0.copy()
    ocont["lpj"] = np.empty((N, S + ocont["S_perm"]))
    rep = Replay(tape, ocont)
    odata = {"y": Y.copy(), "x_infr": X.copy()}
    if missing > 0:
        odata["x"] = X.copy()
    oth = copy.deepcopy(theta0)
    oF = []
    ok = True
    for it in range(iters):
        F, a, b, oth = OB.step(oth, ocont, odata, D, H, S, to_learn=to_learn, do_reconstruction=missing > 0)
        oF.append(F)
        ok &= close(F, Fs[it], 1e-11) and close(a, nun[it]) and close(b, nsub[it])
        ok &= close(oth["W"], thetas[it]["W"], 1e-8) and close(oth["pi"], thetas[it]["pi"], 1e-10)
        ok &= close(oth["sigma"], thetas[it]["sigma"], 1e-10)
    ok &= rep.ok and np.array_equal(ocont["ss"], cont["ss"]) and close(ocont["lpj"], cont["lpj"], 1e-11)
    check("oracle replay == reference: %s" % name, bool(ok))

    # ---- oracle with intended (non-quirk) re-discovery bookkeeping: same children => same K sets
    ocont2 = OV.init_states(1, S, H, sel, mut, P, C, G, bitflip_prob=cont["bitflip_prob"],
                            permanent=copy.deepcopy(permanent))
    ocont2["ss"] = ss0.copy()
    ocont2["lpj"] = np.empty((N, S + ocont2["S_perm"]))
    Replay(tape, ocont2, check_inputs=False)
    oth2 = copy.deepcopy(theta0)
    odata2 = {"y": Y.copy(), "x_infr": X.copy()}
    if missing > 0:
        odata2["x"] = X.copy()
    ok2 = True
    for it in range(iters):
        F, a, b, oth2 = OB.step(oth2, ocont2, odata2, D, H, S, to_learn=to_learn,
                                do_reconstruction=missing > 0, reference_quirks=False)
        ok2 &= close(F, Fs[it], 1e-11)
    ok2 &= np.array_equal(ocont2["ss"], cont["ss"])
    check("injected children make the pool bookkeeping irrelevant: %s" % name, bool(ok2))

    ch, cnt = (tape.children_tensor(iters, N, G, H) if tape.mut else (np.zeros((iters, N, G, 1, H), bool), np.zeros((iters, N, G), np.int32)))
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"),
        D=D, H=H, N=N, S=S, P=P, C=C, G=G, iters=iters, selection=sel, mutation=mut,
        allzero=bool(permanent and permanent["allzero"]), missing=missing,
        Y=Y, X=X, W0=theta0["W"], pi0=theta0["pi"], sigma0=theta0["sigma"], ss0=np.packbits(ss0, axis=-1),
        children=np.packbits(ch, axis=-1), child_count=cnt,
        F=np.array(Fs), S_nunique=np.array(nun), S_sub=np.array(nsub),
        W=np.array([t["W"] for t in thetas]), pi=np.array([t["pi"] for t in thetas]),
        sigma=np.array([t["sigma"] for t in thetas]),
        ss_final=np.packbits(cont["ss"], axis=-1), lpj_final=cont["lpj"],
        y_reconstructed=data.get("y_reconstructed", np.zeros(0)),
    )
    return model, data, cont, theta0


def golden_vary_kn():
    rng = np.random.RandomState(3)
    cases = []
    ok = True
    for t in range(300):
        H = int(rng.randint(6, 40))
        S = int(rng.randint(2, 12))
        S_perm = int(rng.randint(0, 2))
        U = int(rng.randint(0, 15))
        Mprime = int(rng.randint(1, S + 1)) if t % 3 else S
        # unique old states
        pool = {}
        while len(pool) < S:
            r = rng.rand(H) < 0.3
            if S_perm and not r.any():
                continue
            pool[r.tobytes()] = r
        states = np.array(list(pool.values()))
        new = rng.rand(U, H) < 0.3
        if U > 2:
            new[1] = new[0]
            new[2] = states[0]
        if U > 3 and S_perm:
            new[3] = False
        lpj_old = rng.normal(size=S) * 5
        lpj_new = rng.normal(size=U) * 5
        for i in range(U):  # identical states carry identical lpj
            for j in range(i):
                if np.array_equal(new[i], new[j]):
                    lpj_new[i] = lpj_new[j]
        incl = np.zeros((S_perm, H), dtype=bool)
        a_s, a_l, a_o = states.copy(), np.zeros(S), lpj_old.copy()
        ra = ref_vary_Kn(a_o, lpj_new.copy(), a_l, a_s, new.copy(), H, S, S_perm, incl, Mprime) if U else (0, 0)
        if not U:
            a_l[:] = a_o
        b_s, b_l, b_o = states.copy(), np.zeros(S), lpj_old.copy()
        rb = OV.vary_Kn(b_o, lpj_new.copy(), b_l, b_s, new.copy(), H, S, S_perm, incl, Mprime)
        ok &= tuple(ra) == tuple(rb) and np.array_equal(a_s, b_s) and np.array_equal(a_l, b_l)
        cases.append(dict(H=H, S=S, S_perm=S_perm, Mprime=Mprime, states=states, new=new, lpj_old=lpj_old,
                          lpj_new=lpj_new, out_states=a_s, out_lpj=a_l, ret=np.array(ra)))
    check("vary_Kn 300 random cases (slots + values)", bool(ok))
    np.savez_compressed(os.path.join(HERE, "vary_kn_cases.npz"),
                        **{"c%d_%s" % (i, k): v for i, c in enumerate(cases) for k, v in c.items()})


def golden_init_states():
    ok = True
    for seed, (N, S, H, perm) in enumerate([(20, 8, 10, None), (10, 16, 40, {"background": False, "allzero": True, "singletons": False}),
                                             (5, 8, 3, None), (4, 6, 9, {"background": True, "allzero": False, "singletons": False})]):
        np.random.seed(seed)
        a = ref_init_states(N, S, H, "fit", "randflip", 2, 1, 1, permanent=copy.deepcopy(perm))
        np.random.seed(seed)
        b = OV.init_states(N, S, H, "fit", "randflip", 2, 1, 1, permanent=copy.deepcopy(perm))
        ok &= np.array_equal(a["ss"], b["ss"]) and a["S_perm"] == b["S_perm"] and a["incl"].shape == b["incl"].shape
        ok &= (a["sm"] is None) == (b["sm"] is None) and (a["sm"] is None or np.array_equal(a["sm"], b["sm"]))
        ok &= a["lpj"].shape == b["lpj"].shape
    check("init_states draw-for-draw", bool(ok))


def golden_operators():
    ok = True
    rng = np.random.RandomState(0)
    from evo.variational import eas as R
    parents = rng.rand(6, 23) < 0.2
    lpj = rng.normal(size=30) * 4 - 10
    cands = rng.rand(30, 23) < 0.2
    for name, args in [("randflip", (parents, 3, 2.0, None)), ("sparseflip", (parents, 3, 2.0, 0.05)),
                       ("cross", (parents,)), ("cross_randflip", (parents, 5, 2.0, None)),
                       ("cross_sparseflip", (parents, 5, 2.0, 0.05))]:
        np.random.seed(5)
        a = getattr(R, name)(*[x.copy() if isinstance(x, np.ndarray) else x for x in args])
        np.random.seed(5)
        b = OV.MUTATION[name](*[x.copy() if isinstance(x, np.ndarray) else x for x in args])
        ok &= np.array_equal(a, b)
    for name, fn in [("fitparents", R.fitparents), ("randparents", R.randparents)]:
        np.random.seed(6)
        a = fn(cands, 5, lpj)
        np.random.seed(6)
        b = getattr(OV, name)(cands, 5, lpj)
        ok &= np.array_equal(a, b)
    check("operators draw-for-draw under the same seed", bool(ok))


def main():
    golden_bsc_lpj()
    golden_vary_kn()
    golden_init_states()
    golden_operators()

    # (ii) exact E-step trajectory, RNG independent: S = 2^H
    H, D = 8, 16
    rng = np.random.RandomState(1)
    run_bsc(0, D, H, 200, 2 ** H, ("fit", "randflip", 4, 1, 1), 5, rng.normal(size=(D, H)) * 3, 0.2, 1.0,
            exact=True, name="bsc_exact")

    # (iii) bars test, replayed candidates, 2 generations
    model, data, cont, theta0 = run_bsc(0, 25, 10, 300, 20, ("fit", "randflip", 5, 2, 2), 6, 10 * bars_dict(10),
                                        0.2, 1.0, name="bsc_bars_replay")
    # (iv) exact log-likelihood of the generating parameters
    gen = model.check_params({"W": 10 * bars_dict(10), "pi": 0.2, "sigma": 1.0})
    L_ref = model.free_energy(data, copy.deepcopy(gen), cont, full=True)
    L_me = OB.free_energy_full(copy.deepcopy(gen), data, 25, 10, OV.all_states(10))
    check("free_energy(full=True)", close(L_ref, L_me, 1e-12))
    np.savez_compressed(os.path.join(HERE, "bsc_full_F.npz"), Y=data["y"], W=gen["W"], pi=gen["pi"],
                        sigma=gen["sigma"], L=L_ref)

    # all-zero permanent state, sparseflip, H not a multiple of 32
    rng = np.random.RandomState(2)
    run_bsc(1, 30, 40, 120, 16, ("rand", "sparseflip", 4, 2, 2), 4, rng.normal(size=(30, 40)) * 2, 0.05, 0.8,
            permanent={"background": False, "allzero": True, "singletons": False}, name="bsc_allzero_sparseflip")
    # crossover + Mprime default, H > 32
    run_bsc(2, 20, 70, 100, 12, ("fit", "cross_randflip", 4, 3, 2), 4, rng.normal(size=(20, 70)) * 2, 0.04, 0.7,
            name="bsc_cross")
    # missing data with reconstruction
    run_bsc(3, 25, 10, 150, 20, ("fit", "randflip", 5, 2, 2), 4, 10 * bars_dict(10), 0.2, 1.0, missing=0.2,
            name="bsc_missing")

    if FAIL:
        print("MISMATCHES:", FAIL)
        sys.exit(1)
    print("all oracle pins hold; fixtures written to", HERE)


if __name__ == "__main__":
    main()
