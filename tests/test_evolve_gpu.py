"""The device's OWN evolutionary operators (RNG mode), through `evo_evolve` (include/evo_b200.h):

(i)   structural, bit-exact: the selected parents, flip positions / cut points and children the kernel
      emits obey the operator definitions of evo/variational/eas.py:10-150 -- checked against the formulas
      and against `oracle.variational`'s operators fed the recorded positions / cut points;
(ii)  distributional: sparseflip on/off flip frequencies against p1 / p0 of eas.py:74-98, fitness-proportional
      selection against the reference's `np.random.choice(replace=False, p=w)` (eas.py:139-146; successive
      sampling = what the Efraimidis-Spirakis keys of the kernel implement), uniform selection;
(iii) the chain that ties the benchmarked kernels to the oracle: FAST / LIST kernels in RNG mode ==
      generic kernel in RNG mode (trace on) == generic kernel with the emitted children injected ==
      NumPy oracle with the same children injected -- K sets bit-exact, lpj / F to 1e-9.
"""
import copy
import os
import warnings

import numpy as np
import pytest

import parity as P

pytestmark = pytest.mark.gpu
warnings.simplefilter("ignore")

SELS = ["fit", "rand"]
MUTS = ["randflip", "sparseflip", "cross", "cross_randflip", "cross_sparseflip"]


def _bsc_problem(D, H, N, act, seed):
    rng = np.random.RandomState(seed)
    Wg = rng.normal(size=(D, H)) * 2.0
    s = rng.random_sample((N, H)) < act / H
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32).astype(np.float64)
    W0 = (Wg + 0.5 * rng.normal(size=(D, H))).astype(np.float32).astype(np.float64)
    return Y, W0


def _trace_np(cont, H):
    from evo_b200.variational import unpack_states

    tr = cont["trace_out"]
    return {
        "parents": unpack_states(tr["parents"], H).cpu().numpy(), "npar": tr["npar"].cpu().numpy(),
        "children": unpack_states(tr["children"], H).cpu().numpy(), "count": tr["count"].cpu().numpy(),
        "pos": tr["pos"].cpu().numpy(),
    }


@pytest.mark.parametrize("background", [False, True])
@pytest.mark.parametrize("mut", MUTS)
@pytest.mark.parametrize("sel", SELS)
def test_operator_structure(sel, mut, background):
    """(i): every child is what the operator definition makes of its recorded parent(s) and position(s)."""
    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states
    import oracle.variational as OV

    D, H, N, S, Pn, C, G = 12, 70, 96, 12, 4, 3, 2
    Y, W0 = _bsc_problem(D, H, N, 3.0, 7)
    perm = {"background": True, "allzero": False, "singletons": False} if background else None
    m = BSC(D, H, S)
    cont = init_states(N, S, H, sel, mut, Pn, C, G, bitflip_prob=1.5 / H, permanent=perm, seed=123)
    ss0 = cont["ss"].copy()
    cont["trace"] = True
    th = {"W": W0, "pi": 3.0 / H, "sigma": 1.0}
    m.E_step(m.check_params(th), cont, {"y": Y, "x_infr": None})
    t = _trace_np(cont, H)
    Hm = H - 1 if background else H
    assert (t["npar"][:, 0] == Pn).all()
    for n in range(N):
        kn = {r.tobytes() for r in ss0[n]}
        for g in range(G):
            npar, nc = int(t["npar"][n, g]), int(t["count"][n, g])
            if npar == 0:
                assert nc == 0
                continue
            par = t["parents"][n, g, :npar]
            ch = t["children"][n, g, :nc]
            pos = t["pos"][n, g, :nc]
            assert len({r.tobytes() for r in par}) == npar, "parents are drawn without replacement"
            if g == 0:
                assert all(r.tobytes() in kn for r in par), "generation 0 selects from K^n"
            if background:
                assert par[:, -1].all() and ch[:, -1].all(), "the background unit stays on (eas.py:215-218,233-239)"
            if mut in ("randflip", "sparseflip"):
                assert nc == npar * C
                rep = np.repeat(par, C, axis=0)
                diff = ch ^ rep
                assert not diff[:, Hm:].any()
                if mut == "randflip":
                    assert (pos[:, 1] == -1).all() and (pos[:, 0] >= 0).all() and (pos[:, 0] < Hm).all()
                    want = rep.copy()
                    want[np.arange(nc), pos[:, 0]] ^= True
                    assert np.array_equal(ch, want), "child j = parent with position j flipped (eas.py:27-42)"
                    pp = pos[:, 0].reshape(npar, C)
                    assert all(len(set(r)) == C for r in pp), "positions of one parent are distinct (eas.py:27-33)"
                    # the oracle's randflip fed scores that make it choose the recorded positions (child order within
                    # a parent follows argpartition, so rows are compared per parent as sets)
                    sc = np.ones((npar, Hm))
                    for p_ in range(npar):
                        sc[p_, pp[p_]] = np.arange(C) * 1e-3
                    keep = np.random.rand
                    np.random.rand = lambda *a: sc
                    try:
                        och = OV.randflip(par[:, :Hm].copy(), C)
                    finally:
                        np.random.rand = keep
                    for p_ in range(npar):
                        a = {r.tobytes() for r in och[p_ * C:(p_ + 1) * C]}
                        b = {r[:Hm].tobytes() for r in ch[p_ * C:(p_ + 1) * C]}
                        assert a == b
            else:
                assert nc == npar * (npar - 1)
                cps = []
                q = 0
                for i in range(npar):
                    for j in range(i + 1, npar):
                        cp = int(pos[2 * q, 0])
                        assert 1 <= cp <= Hm - 1 and pos[2 * q + 1, 0] == cp, "cut point in {1..H-1} (eas.py:119)"
                        cps.append(cp)
                        c0 = np.concatenate([par[i, :cp], par[j, cp:]])
                        c1 = np.concatenate([par[j, :cp], par[i, cp:]])
                        for k, cw in enumerate((c0, c1)):
                            got = ch[2 * q + k]
                            d = got ^ cw
                            assert not d[Hm:].any()
                            if mut == "cross":
                                assert not d.any(), "one-point crossover (eas.py:120-123)"
                            elif mut == "cross_randflip":
                                fp = int(pos[2 * q + k, 1])
                                assert 0 <= fp < Hm and d.sum() == 1 and d[fp], "cross then one flip (eas.py:128-130)"
                        q += 1
                if mut == "cross":  # the oracle's cross() replaying the recorded cut points
                    it = iter(cps)
                    keep = np.random.randint
                    np.random.randint = lambda low, high=None, **k: next(it)
                    try:
                        och = OV.cross(par[:, :Hm].copy())
                    finally:
                        np.random.randint = keep
                    assert np.array_equal(och, ch[:, :Hm])


def test_sparseflip_flip_rates():
    """(ii): per parent size |s|, on-bits flip with p1 and off-bits with p0 of eas.py:74-98 (z-test, |z| < 5)."""
    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    D, H, N, S, Pn, C = 8, 64, 6000, 16, 8, 8
    p_bf = 1.0 / H
    Y, W0 = _bsc_problem(D, H, N, 5.0, 3)
    m = BSC(D, H, S)
    cont = init_states(N, S, H, "rand", "sparseflip", Pn, C, 1, bitflip_prob=p_bf, p_init_Kn=5.0 / H, seed=5)
    cont["trace"] = True
    th = m.check_params({"W": W0, "pi": 6.0 / H, "sigma": 1.0})
    m.E_step(th, cont, {"y": Y, "x_infr": None})
    piH = th["piH"]
    t = _trace_np(cont, H)
    par = np.repeat(t["parents"][:, 0], C, axis=1).reshape(-1, H)  # (N * Pn * C, H)
    ch = t["children"][:, 0].reshape(-1, H)
    assert (t["count"][:, 0] == Pn * C).all()
    flips = par ^ ch
    a = par.sum(1)
    total = 0
    for size in np.unique(a):
        rows = a == size
        s_ = float(size)
        alpha = (H - s_) * (H * p_bf - (piH - s_)) / ((piH - s_ + H * p_bf) * s_ + 1e-100)
        p0 = H * p_bf / (H + (alpha - 1.0) * s_ + 1e-100)
        p1 = alpha * p0
        p0, p1 = min(max(p0, 0.0), 1.0), min(max(p1, 0.0), 1.0)
        n_on, n_off = int(par[rows].sum()), int((~par[rows]).sum())
        f_on, f_off = int((flips[rows] & par[rows]).sum()), int((flips[rows] & ~par[rows]).sum())
        for nn, ff, pp in ((n_on, f_on, p1), (n_off, f_off, p0)):
            if nn == 0:
                continue
            if pp == 0.0:
                assert ff == 0, (size, nn, ff)
            elif nn * pp < 20:
                continue  # too few expected flips for the normal approximation
            else:
                assert abs(ff - nn * pp) < 5.0 * np.sqrt(nn * pp * (1 - pp)) + 1.0, (size, nn, ff, pp)
            total += nn
    assert total > 1e6, "at least 10^6 bit trials entered the test"


def _selection_counts(sel, N, S, Pn, lpj_spread, seed):
    """Identical K^n and data for every datapoint (different Philox streams): parents of generation 0."""
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    D, H = 6, 20
    rng = np.random.RandomState(seed)
    W0 = rng.normal(size=(D, H)) * lpj_spread
    states = np.zeros((S, H), dtype=bool)
    for k in range(S):
        states[k, rng.choice(H, size=1 + k % 3, replace=False)] = True
    states = np.unique(states, axis=0)
    while states.shape[0] < S:
        r = rng.random_sample(H) < 0.15
        if r.any() and not (states == r).all(1).any():
            states = np.concatenate([states, r[None]])
    states = states[:S]
    y = rng.normal(size=D)
    Y = np.tile(y.astype(np.float32).astype(np.float64)[None], (N, 1))
    m = BSC(D, H, S)
    cont = init_states(N, S, H, sel, "randflip", Pn, 1, 1, seed=77)
    cont["ss"] = np.tile(states[None], (N, 1, 1))
    cont["trace"] = True
    th = m.check_params({"W": W0, "pi": 0.1, "sigma": 1.0})
    ss_before = cont["ss_dev"].clone()
    m.E_step(th, cont, {"y": Y, "x_infr": None})
    t = _trace_np(cont, H)
    par = t["parents"][:, 0]  # (N, Pn, H)
    lpj = m.lpj_all(th, {"y": Y[:1], "x_infr": None}, ss_before[:1]).cpu().numpy()[0]
    # index of each parent in K^n
    key = {states[k].tobytes(): k for k in range(S)}
    idx = np.array([[key[r.tobytes()] for r in par[n]] for n in range(N)])
    return idx, lpj


def test_fit_selection_matches_reference_distribution():
    """(ii): first pick ~ w / sum(w) (chi-square, 15 dof) and inclusion frequencies equal those of the reference's
    fitparents (np.random.choice(replace=False, p=w), eas.py:139-146) within sampling error."""
    import oracle.variational as OV

    N, S, Pn = 200000, 16, 4
    idx, lpj = _selection_counts("fit", N, S, Pn, 0.6, 1)
    w = lpj - 2.0 * min(lpj.min(), 0.0)
    w = w / w.sum()
    assert w.max() / w.min() > 1.5, "the weights are not flat, or the test says nothing"
    first = np.bincount(idx[:, 0], minlength=S)
    chi2 = float(((first - N * w) ** 2 / (N * w)).sum())
    assert chi2 < 60.0, (chi2, first / N, w)  # chi^2_15: P(> 60) ~ 2e-7
    incl = np.bincount(idx.reshape(-1), minlength=S) / N
    np.random.seed(0)
    M = 40000
    ref = np.zeros(S)
    cands = np.arange(S)[:, None]
    for _ in range(M):
        ref[OV.fitparents(cands, Pn, lpj)[:, 0]] += 1
    ref /= M
    sd = np.sqrt(ref * (1 - ref) * (1.0 / M + 1.0 / N))
    assert (np.abs(incl - ref) < 5.0 * sd).all(), (incl, ref)
    assert all(len(set(r)) == Pn for r in idx[:2000])


def test_rand_selection_uniform():
    N, S, Pn = 100000, 16, 4
    idx, _ = _selection_counts("rand", N, S, Pn, 0.6, 2)
    cnt = np.bincount(idx.reshape(-1), minlength=S)
    exp = N * Pn / S
    chi2 = float(((cnt - exp) ** 2 / exp).sum())
    assert chi2 < 60.0 * (1 - Pn / S) + 20.0, (chi2, cnt)
    first = np.bincount(idx[:, 0], minlength=S)
    assert float(((first - N / S) ** 2 / (N / S)).sum()) < 60.0


# ---------------------------------------------------------------------------------------------
# (iii) benchmarked kernels -> generic kernel -> injected children -> oracle
# ---------------------------------------------------------------------------------------------
def _chain_bsc(D, H, N, S, ea, act, p_init, iters, to_learn, missing=0.0, permanent=None, tables="f64"):
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states
    from oracle import bsc as OB
    from oracle import variational as OV

    sel, mut, Pn, C, G = ea
    Y, W0 = _bsc_problem(D, H, N, act, D + H)
    X = np.ones_like(Y, dtype=bool)
    if missing > 0:
        X = np.random.RandomState(5).random_sample(Y.shape) >= missing
    Yn = np.where(X, Y, np.nan)
    th0 = {"W": W0, "pi": act / H, "sigma": 1.1}
    bf = 1.0 / H if "sparse" in mut else None

    def fresh():
        c = init_states(N, S, H, sel, mut, Pn, C, G, bitflip_prob=bf, p_init_Kn=p_init, permanent=copy.deepcopy(permanent),
                        seed=99)
        return c

    def data():
        d = {"y": Yn.copy(), "x_infr": X.copy() if missing > 0 else None}
        if missing > 0:
            d["x"] = X.copy()
        return d

    def run(cont, inject=None, trace=False, generic=False):
        m = BSC(D, H, S, to_learn=to_learn)
        m.tables = tables
        th = copy.deepcopy(th0)
        d = data()
        Fs, kids = [], []
        old = os.environ.pop("EVO_ESTEP_GENERIC", None)
        if generic:
            os.environ["EVO_ESTEP_GENERIC"] = "1"
        try:
            for it in range(iters):
                if inject is not None:
                    cont["inject"] = inject[it]
                cont["trace"] = trace
                F, a, b, th = m.step(th, cont, d, do_reconstruction=missing > 0)
                Fs.append((F, a, b))
                if trace:
                    t = _trace_np(cont, H)
                    kids.append((t["children"], t["count"]))
        finally:
            os.environ.pop("EVO_ESTEP_GENERIC", None)
            if old is not None:
                os.environ["EVO_ESTEP_GENERIC"] = old
        torch.cuda.synchronize()
        return cont["ss"], cont["lpj"], Fs, kids, th

    c0 = fresh()
    ss0 = c0["ss"].copy()
    ssA, lpjA, FA, _, thA = run(c0)                                  # what the bench runs (FAST / LIST when eligible)
    ssG, lpjG, FG, _, _ = run(fresh(), generic=True)                 # generic kernel, RNG mode
    ssX, lpjX, FX, kids, _ = run(fresh(), trace=True)                # generic kernel through evo_evolve, children emitted
    ssB, lpjB, FB, _, _ = run(fresh(), inject=kids)                  # the emitted children injected
    # oracle with the same children
    ocont = OV.init_states(1, S, H, sel, mut, Pn, C, G, bitflip_prob=bf, permanent=copy.deepcopy(permanent))
    ocont["ss"] = ss0.copy()
    ocont["lpj"] = np.empty((N, S + ocont["S_perm"]))
    ch = np.stack([k[0] for k in kids])
    cn = np.stack([k[1] for k in kids])
    rep = P.OracleReplay(ocont, ch, cn)
    od = {"y": Yn.copy(), "x_infr": X.copy()}
    if missing > 0:
        od["x"] = X.copy()
    oth = copy.deepcopy(th0)
    FO = []
    for it in range(iters):
        F, a, b, oth = OB.step(oth, ocont, od, D, H, S, to_learn=to_learn, do_reconstruction=missing > 0,
                               reference_quirks=False)
        FO.append((F, a, b))
        rep.next_iteration()
    assert (cn.sum(axis=(1, 2)) > 0).all(), "the device produced children in every iteration"
    for name, ss_, lpj_, F_ in (("generic", ssG, lpjG, FG), ("trace", ssX, lpjX, FX), ("injected", ssB, lpjB, FB),
                                ("oracle", ocont["ss"], ocont["lpj"], FO)):
        assert np.array_equal(ssA, ss_), "K sets: bench kernels vs %s" % name
        assert P.maxrel(lpjA, lpj_) < 1e-9, name
        for (f1, a1, b1), (f2, a2, b2) in zip(FA, F_):
            assert abs(f1 - f2) < 1e-9 * abs(f2) and abs(a1 - a2) < 1e-12 and abs(b1 - b2) < 1e-12, (name, f1, f2, a1, a2, b1, b2)
    for k in ("pi", "sigma"):
        assert abs(float(thA[k]) - float(oth[k])) < 1e-9 * abs(float(oth[k]))
    if "W" in to_learn:
        assert P.maxrel(np.asarray(thA["W"]), oth["W"]) < 1e-7


@pytest.mark.parametrize("sel,mut", [(s, m) for s in SELS for m in MUTS])
def test_chain_small_every_operator(sel, mut):
    _chain_bsc(20, 40, 200, 16, (sel, mut, 4, 3, 2), 3.0, 2.0 / 40, 3, ("W", "pi", "sigma"))


def test_chain_background_and_allzero():
    _chain_bsc(20, 40, 150, 16, ("fit", "sparseflip", 4, 3, 2), 3.0, 2.0 / 40, 2, ("pi", "sigma"),
               permanent={"background": True, "allzero": False, "singletons": False})
    _chain_bsc(20, 40, 150, 16, ("fit", "randflip", 4, 3, 2), 3.0, 2.0 / 40, 2, ("pi", "sigma"),
               permanent={"background": False, "allzero": True, "singletons": False})


def test_chain_missing_data():
    _chain_bsc(20, 40, 150, 16, ("fit", "randflip", 4, 3, 2), 3.0, 2.0 / 40, 2, ("pi", "sigma"), missing=0.2)


def test_chain_cfg2_shape():
    """D=256, H=300, K=64, fit / randflip 8/8/2 -- the list kernel the headline bench times (T=64, 512-slot table)."""
    _chain_bsc(256, 300, 48, 64, ("fit", "randflip", 8, 8, 2), 5.0, 4.0 / 300, 3, ("pi", "sigma"))


def test_chain_cfg5_shape():
    """D=H=1024, K=128, fit / randflip 16/8/2 (T=128, 1024-slot table, 32-word states; dense initial states so
    that some datapoints leave the 15-entry list form and are redone by the packed-bits kernel)."""
    _chain_bsc(1024, 1024, 12, 128, ("fit", "randflip", 16, 8, 2), 8.0, 9.0 / 1024, 2, ("pi", "sigma"))


def _chain_sssc(D, H, N, S, ea, act, p_init, iters, to_learn, missing=0.0):
    import torch

    from evo_b200.sssc import SSSC
    from evo_b200.variational import init_states
    from oracle import sssc as OS
    from oracle import variational as OV
    import sys
    if P.GOLDEN not in sys.path:
        sys.path.insert(0, P.GOLDEN)
    from shape_inputs import shape_inputs

    sel, mut, Pn, C, G = ea
    inp = shape_inputs("es3c", 1000 + D + H, D, H, N, act, missing)
    Y, X = inp["Y"], inp["X"]
    Yn = np.where(X, Y, np.nan)
    xrec = X.copy() if missing > 0 else np.zeros_like(X)
    th0 = {"W": inp["W0"], "pies": inp["pies0"], "mus": inp["mus0"], "Psi": inp["Psi0"], "sigma2": np.float64(1.3)}

    def run(cont, inject=None, trace=False):
        m = SSSC(D, H, S, to_learn=to_learn)
        th = copy.deepcopy(th0)
        d = {"y": Yn.copy(), "x_infr": X.copy(), "x": xrec.copy()}
        Fs, kids = [], []
        for it in range(iters):
            if inject is not None:
                cont["inject"] = inject[it]
            cont["trace"] = trace
            F, a, b, th = m.step(th, cont, d, do_reconstruction=True)
            Fs.append((F, a, b))
            if trace:
                t = _trace_np(cont, H)
                kids.append((t["children"], t["count"]))
        torch.cuda.synchronize()
        return cont["ss"], cont["lpj"], Fs, kids, th

    fresh = lambda: init_states(N, S, H, sel, mut, Pn, C, G, p_init_Kn=p_init, seed=31)
    c0 = fresh()
    ss0 = c0["ss"].copy()
    ssA, lpjA, FA, _, thA = run(c0)
    ssX, lpjX, FX, kids, _ = run(fresh(), trace=True)
    ssB, lpjB, FB, _, _ = run(fresh(), inject=kids)
    ocont = OV.init_states(1, S, H, sel, mut, Pn, C, G)
    ocont["ss"] = ss0.copy()
    ocont["lpj"] = np.empty((N, S))
    rep = P.OracleReplay(ocont, np.stack([k[0] for k in kids]), np.stack([k[1] for k in kids]))
    od = {"y": Yn.copy(), "x_infr": X.copy(), "x": xrec.copy()}
    oth = copy.deepcopy(th0)
    FO = []
    for it in range(iters):
        F, a, b, oth = OS.em_step(oth, ocont, od, D, H, S, to_learn=to_learn, do_reconstruction=True, reference_quirks=False)
        FO.append((F, a, b))
        rep.next_iteration()
    for name, ss_, lpj_, F_ in (("trace", ssX, lpjX, FX), ("injected", ssB, lpjB, FB), ("oracle", ocont["ss"], ocont["lpj"], FO)):
        assert np.array_equal(ssA, ss_), "K sets: fused RNG-mode kernel vs %s" % name
        assert P.maxrel(lpjA, lpj_) < 1e-7, name
        for (f1, a1, b1), (f2, a2, b2) in zip(FA, F_):
            assert abs(f1 - f2) < 1e-8 * abs(f2) and abs(a1 - a2) < 1e-12 and abs(b1 - b2) < 1e-12, (name, f1, f2)
    for k in to_learn:
        assert P.maxrel(np.asarray(thA[k]), np.asarray(oth[k])) < 1e-6, k


def test_chain_es3c_small_and_cfg3_cfg4_shapes():
    _chain_sssc(12, 10, 120, 12, ("fit", "randflip", 4, 2, 2), 2.0, 0.2, 3, ("W", "pies", "mus", "sigma2", "Psi"))
    _chain_sssc(144, 512, 16, 64, ("fit", "randflip", 8, 8, 2), 4.0, 3.0 / 512, 2, ("pies", "mus", "sigma2"))
    _chain_sssc(144, 512, 12, 64, ("fit", "randflip", 8, 8, 2), 4.0, 3.0 / 512, 2, ("pies", "sigma2"), missing=0.3)


# ---------------------------------------------------------------------------------------------
# reference-signature entry points
# ---------------------------------------------------------------------------------------------
def test_operator_callables_have_the_reference_signatures():
    import evo_b200.variational as V

    rng = np.random.RandomState(0)
    H = 23
    parents = np.unique(rng.random_sample((6, H)) < 0.2, axis=0)
    Pn = parents.shape[0]
    V.seed_operators(4)
    ch = V.randflip(parents, 3, 2.0, None)
    assert ch.shape == (Pn * 3, H) and ((ch ^ np.repeat(parents, 3, axis=0)).sum(1) == 1).all()
    ch = V.sparseflip(parents, 3, 2.0, 0.05)
    assert ch.shape == (Pn * 3, H)
    ch = V.cross(parents)
    assert ch.shape == (Pn * (Pn - 1), H)
    tot = ch.reshape(-1, 2, H).sum(1)  # the two children of a pair carry the bits of their two parents
    q = 0
    for i in range(Pn):
        for j in range(i + 1, Pn):
            assert np.array_equal(tot[q], parents[i].astype(int) + parents[j].astype(int))
            q += 1
    assert V.cross_randflip(parents, 5, 2.0, None).shape == (Pn * (Pn - 1), H)
    assert V.cross_sparseflip(parents, 5, 2.0, 0.05).shape == (Pn * (Pn - 1), H)
    cands = np.unique(rng.random_sample((30, H)) < 0.2, axis=0)
    lpj = rng.normal(size=cands.shape[0]) * 4 - 10
    for fn in (V.fitparents, V.randparents):
        par = fn(cands, 5, lpj)
        assert par.shape == (5, H) and len({r.tobytes() for r in par}) == 5
        assert all((cands == r).all(1).any() for r in par)
    V.seed_operators(4)
    b = V.randflip(parents, 3, 2.0, None)
    c = V.randflip(parents, 3, 2.0, None)
    V.seed_operators(4)
    assert np.array_equal(b, V.randflip(parents, 3, 2.0, None)), "same seed and call index -> same children"
    assert not np.array_equal(b, c), "consecutive calls draw from different streams"


def test_evolve_states_reference_signature():
    """`evolve_states(my_suff_stat, model_params, eval_lpj)` (eas.py:153): new states are unique, not in K^n or
    `incl`, lexicographically ordered within a generation, and carry the callback's values; with one generation
    it returns exactly the states the fused `model.evolve` finds for the same Philox keys."""
    import evo_b200.variational as V
    from evo_b200.bsc import BSC
    from oracle import bsc as OB

    D, H, S = 10, 30, 12
    Y, W0 = _bsc_problem(D, H, 1, 3.0, 9)
    m = BSC(D, H, S)
    th = m.check_params({"W": W0, "pi": 0.1, "sigma": 1.0})
    for G, perm in ((1, None), (3, None), (2, {"background": False, "allzero": True, "singletons": False})):
        cont = V.init_states(1, S, H, "fit", "randflip", 4, 3, G, permanent=copy.deepcopy(perm), seed=8)
        d = {"y": Y, "x_infr": None}
        m.E_step_precompute(th, cont, d)
        ss = cont["ss"][0]
        x = np.ones(D, dtype=bool)
        calls = []

        def eval_lpj(states):
            calls.append(states.copy())
            return OB.lpj_states(states, W0, Y[0], x, th["pre1"], th["pil_bar"])

        lpj0 = eval_lpj(ss)
        calls.clear()
        one = dict(cont)
        one.update({"this_states": ss, "this_lpj": lpj0})
        V.seed_operators(21)
        new, new_lpj = V.evolve_states(one, th, eval_lpj)
        assert new.shape[0] == new_lpj.shape[0] and new.shape[0] > 0 and len(calls) <= G
        keys = [r.tobytes() for r in new]
        assert len(set(keys)) == len(keys) and not (set(keys) & {r.tobytes() for r in ss})
        if perm:
            assert new.any(1).all(), "the permanent all-zero state is never returned"
        for c in calls:
            assert [r.tobytes() for r in c] == sorted(r.tobytes() for r in c), "lexicographic within a generation"
        assert P.maxrel(new_lpj, eval_lpj(new)) < 1e-12
        if G == 1:
            # the fused form with the same keys: seed, iteration = call index, datapoint 0, generation 0
            cont2 = V.init_states(1, S, H, "fit", "randflip", 4, 3, 1, permanent=copy.deepcopy(perm), seed=8)
            cont2["iteration"] = V._CALLS[0] - 1
            st, lp, cnt = m.evolve(th, cont2, d)
            U = int(cnt[0, 0].item())
            got = V.unpack_states(st, H).cpu().numpy()[0, :U]
            assert {r.tobytes() for r in got} == set(keys)
            want = OB.lpj_states(got, W0, Y[0], x, th["pre1"], th["pil_bar"])
            assert P.maxrel(lp.cpu().numpy()[0, :U], want) < 1e-10
            assert np.array_equal(cont2["ss"][0], ss), "evolve leaves K^n untouched"
