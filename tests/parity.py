"""Shared parity harness (test infrastructure): runs the CUDA BSC path on a golden fixture with the
recorded candidate sets injected, and compares with (a) the reference outputs stored in the fixture
and (b) the NumPy oracle replaying the same candidates.  Used by tests/ and __graft_entry__.smoke()."""
import copy
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def relerr(a, b):
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    return float(np.max(np.abs(a - b) / (1e-300 + np.maximum(np.abs(b), 1e-12 * (1 + np.abs(b).max())))) if a.size else 0.0)


def maxrel(a, b):
    """max |a-b| / max|b| -- scale-relative error for parameter arrays."""
    a, b = np.asarray(a, dtype=float), np.asarray(b, dtype=float)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)) if a.size else 0.0


def load_case(path):
    z = np.load(path, allow_pickle=False)
    c = {k: z[k] for k in z.files}
    H, N, S = int(c["H"]), int(c["N"]), int(c["S"])
    if "sha_W0" in c:
        # BASELINE-shape fixture (tests/golden/make_golden_shapes.py): the large inputs are regenerated from the
        # recorded seed and must hash to what the reference run was fed
        if GOLDEN not in sys.path:
            sys.path.insert(0, GOLDEN)
        from shape_inputs import sha, shape_inputs

        model, missing = str(c["model"]), float(c["missing"])
        inp = shape_inputs(model, int(c["seed"]), int(c["D"]), H, N, float(c["act"]), missing)
        assert sha(inp["Y"]) == str(c["sha_Y"]) and sha(inp["W0"]) == str(c["sha_W0"]), \
            "regenerated inputs differ from the ones the fixture was recorded with"
        X = inp["X"]
        c["X"] = X
        c["Y"] = np.where(X, inp["Y"], np.nan)
        c["W0"] = inp["W0"]
        if model == "es3c":
            c["Psi0"] = inp["Psi0"] if missing == 0 else np.eye(H)
            assert sha(c["Psi0"]) == str(c["sha_Psi0"])
            c["Xrec"] = X.copy() if missing > 0 else np.zeros_like(X)
    c["ss0_bool"] = np.unpackbits(c["ss0"], axis=-1)[..., :H].astype(bool)
    c["ss_final_bool"] = np.unpackbits(c["ss_final"], axis=-1)[..., :H].astype(bool)
    c["children_bool"] = np.unpackbits(c["children"], axis=-1)[..., :H].astype(bool)
    return c


def permanent_of(c):
    return {"background": False, "allzero": bool(c["allzero"]), "singletons": False} if bool(c["allzero"]) else None


class OracleReplay:
    """Feeds the fixture's recorded children to the oracle's evolve_states."""

    def __init__(self, cont, children, counts):
        self.cont, self.children, self.counts = cont, children, counts
        self.it, self.n, self.g = 0, -1, 0
        cont["parent_selection"] = self.select
        cont["mutation_algorithm"] = self.mutate

    def select(self, cands, n_parents, lpj):
        if np.shares_memory(cands, self.cont["ss"]):
            self.n += 1
            self.g = 0
        else:
            self.g += 1
        return cands[:n_parents]

    def mutate(self, parents, n_children, sparseness, p_bf):
        k = int(self.counts[self.it, self.n, self.g])
        # with a background unit the operators see the first H-1 latents only (eas.py:215-218,233-239)
        return self.children[self.it, self.n, self.g, :k, :parents.shape[1]].copy()

    def next_iteration(self):
        self.it += 1
        self.n = -1


def learned(c, default):
    return tuple(str(c["to_learn"]).split(",")) if "to_learn" in c else default


def run_oracle_case(c, iters):
    from oracle import bsc as OB
    from oracle import variational as OV

    D, H, N, S = int(c["D"]), int(c["H"]), int(c["N"]), int(c["S"])
    perm = permanent_of(c)
    to_learn = learned(c, ("W", "pi", "sigma"))
    cont = OV.init_states(1, S, H, str(c["selection"]), str(c["mutation"]), int(c["P"]), int(c["C"]), int(c["G"]),
                          bitflip_prob=1.0 / H, permanent=copy.deepcopy(perm))
    cont["ss"] = c["ss0_bool"].copy()
    cont["lpj"] = np.empty((N, cont["ss"].shape[1] + cont["S_perm"]))
    rep = OracleReplay(cont, c["children_bool"], c["child_count"])
    data = {"y": c["Y"].copy(), "x_infr": c["X"].copy()}
    missing = float(c["missing"]) > 0
    if missing:
        data["x"] = c["X"].copy()
    th = {"W": c["W0"].copy(), "pi": float(c["pi0"]), "sigma": float(c["sigma0"])}
    out = []
    for it in range(iters):
        F, a, b, th = OB.step(th, cont, data, D, H, cont["ss"].shape[1], to_learn=to_learn, do_reconstruction=missing,
                              reference_quirks=False)
        out.append((F, a, b, {k: np.array(th[k]) for k in ("W", "pi", "sigma")}))
        rep.next_iteration()
    return out, cont, data


def run_gpu_case(c, iters, inject=True, tables="f64", stats="f64"):
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    D, H, N, S = int(c["D"]), int(c["H"]), int(c["N"]), int(c["S"])
    perm = permanent_of(c)
    model = BSC(D, H, S, to_learn=learned(c, ("W", "pi", "sigma")))
    model.tables, model.stats = tables, stats
    cont = init_states(N, S, H, str(c["selection"]), str(c["mutation"]), int(c["P"]), int(c["C"]), int(c["G"]),
                       bitflip_prob=1.0 / H, permanent=copy.deepcopy(perm))
    cont["ss"] = c["ss0_bool"]
    data = {"y": c["Y"].copy(), "x_infr": c["X"].copy()}
    missing = float(c["missing"]) > 0
    if missing:
        data["x"] = c["X"].copy()
    th = {"W": c["W0"].copy(), "pi": float(c["pi0"]), "sigma": float(c["sigma0"])}
    out = []
    for it in range(iters):
        if inject:
            cont["inject"] = (c["children_bool"][it], c["child_count"][it])
        F, a, b, th = model.step(th, cont, data, do_reconstruction=missing)
        out.append((F, a, b, {k: np.array(th[k]) for k in ("W", "pi", "sigma")}))
    torch.cuda.synchronize()
    return out, cont, data


def run_bsc_parity_case(path, iters=None, tables="f64", stats="f64"):
    """GPU vs fixture (reference outputs) vs oracle; returns error summary."""
    c = load_case(path)
    total = int(c["iters"])
    iters = total if iters is None else min(iters, total)
    gpu, gcont, gdata = run_gpu_case(c, iters, tables=tables, stats=stats)
    ora, ocont, odata = run_oracle_case(c, iters)
    res = {"F_err": 0.0, "W_err": 0.0, "pi_err": 0.0, "sigma_err": 0.0, "count_err": 0.0,
           "F_err_ref": 0.0, "W_err_ref": 0.0}
    for it in range(iters):
        Fg, ag, bg, tg = gpu[it]
        Fo, ao, bo, to = ora[it]
        res["F_err"] = max(res["F_err"], abs(Fg - Fo) / abs(Fo))
        res["W_err"] = max(res["W_err"], maxrel(tg["W"], to["W"]))
        res["pi_err"] = max(res["pi_err"], abs(tg["pi"] - to["pi"]) / abs(to["pi"]))
        res["sigma_err"] = max(res["sigma_err"], abs(tg["sigma"] - to["sigma"]) / abs(to["sigma"]))
        res["count_err"] = max(res["count_err"], abs(ag - ao), abs(bg - bo))
        res["F_err_ref"] = max(res["F_err_ref"], abs(Fg - c["F"][it]) / abs(c["F"][it]))
        if "W" in c:
            res["W_err_ref"] = max(res["W_err_ref"], maxrel(tg["W"], c["W"][it]))
        res["pi_err"] = max(res["pi_err"], abs(tg["pi"] - c["pi"][it]) / abs(c["pi"][it]))
        res["sigma_err"] = max(res["sigma_err"], abs(tg["sigma"] - c["sigma"][it]) / abs(c["sigma"][it]))
    gss = gcont["ss"]
    res["ss_equal"] = bool(np.array_equal(gss, ocont["ss"]))
    res["lpj_err"] = maxrel(gcont["lpj"], ocont["lpj"])
    if iters == total:
        res["ss_equal_ref"] = bool(np.array_equal(gss, c["ss_final_bool"]))
        res["lpj_err_ref"] = maxrel(gcont["lpj"], c["lpj_final"])
        if float(c["missing"]) > 0:
            res["yrec_err_ref"] = maxrel(gdata["y_reconstructed"], c["y_reconstructed"])
    res["iters"] = iters
    return res


# ---------------------------------------------------------------------------------------------
# ES3C (SSSC)
# ---------------------------------------------------------------------------------------------
PKEYS = ("W", "pies", "mus", "Psi", "sigma2")


def _sssc_theta0(c):
    if "sigma20" in c:  # BASELINE-shape fixture
        return {"W": np.array(c["W0"]), "pies": np.array(c["pies0"]), "mus": np.array(c["mus0"]), "Psi": np.array(c["Psi0"]),
                "sigma2": np.float64(c["sigma20"])}
    return {k: np.array(c[k + "0"]) for k in PKEYS}


def run_oracle_sssc_case(c, iters):
    from oracle import sssc as OS
    from oracle import variational as OV

    D, H, N, S = int(c["D"]), int(c["H"]), int(c["N"]), int(c["S"])
    perm = permanent_of(c)
    cont = OV.init_states(1, S, H, str(c["selection"]), str(c["mutation"]), int(c["P"]), int(c["C"]), int(c["G"]),
                          bitflip_prob=1.0 / H, permanent=copy.deepcopy(perm))
    cont["ss"] = c["ss0_bool"].copy()
    cont["lpj"] = np.empty((N, S + cont["S_perm"]))
    rep = OracleReplay(cont, c["children_bool"], c["child_count"])
    data = {"y": c["Y"].copy(), "x_infr": c["X"].copy(), "x": c["Xrec"].copy()}
    th = _sssc_theta0(c)
    to_learn = tuple(str(c["to_learn"]).split(","))
    out = []
    for it in range(iters):
        F, a, b, th = OS.em_step(th, cont, data, D, H, S, to_learn=to_learn, do_reconstruction=True,
                                 reference_quirks=False)
        out.append((F, a, b, {k: np.array(th[k]) for k in PKEYS}))
        rep.next_iteration()
    return out, cont, data


def run_gpu_sssc_case(c, iters):
    import torch

    from evo_b200.sssc import SSSC
    from evo_b200.variational import init_states

    D, H, N, S = int(c["D"]), int(c["H"]), int(c["N"]), int(c["S"])
    perm = permanent_of(c)
    to_learn = tuple(str(c["to_learn"]).split(","))
    model = SSSC(D, H, S, to_learn=to_learn)
    cont = init_states(N, S, H, str(c["selection"]), str(c["mutation"]), int(c["P"]), int(c["C"]), int(c["G"]),
                       bitflip_prob=1.0 / H, permanent=copy.deepcopy(perm))
    cont["ss"] = c["ss0_bool"]
    data = {"y": c["Y"].copy(), "x_infr": c["X"].copy(), "x": c["Xrec"].copy()}
    th = _sssc_theta0(c)
    out = []
    for it in range(iters):
        cont["inject"] = (c["children_bool"][it], c["child_count"][it])
        F, a, b, th = model.step(th, cont, data, do_reconstruction=True)
        out.append((F, a, b, {k: np.array(th[k]) for k in PKEYS}))
    torch.cuda.synchronize()
    return out, cont, data


def run_sssc_parity_case(path, iters=None):
    c = load_case(path)
    total = int(c["iters"])
    iters = total if iters is None else min(iters, total)
    gpu, gcont, gdata = run_gpu_sssc_case(c, iters)
    ora, ocont, odata = run_oracle_sssc_case(c, iters)
    res = {"F_err": 0.0, "F_err_ref": 0.0, "count_err": 0.0}
    for k in PKEYS:
        res[k + "_err"] = 0.0
        res[k + "_err_ref"] = 0.0
    for it in range(iters):
        Fg, ag, bg, tg = gpu[it]
        Fo, ao, bo, to = ora[it]
        res["F_err"] = max(res["F_err"], abs(Fg - Fo) / abs(Fo))
        res["F_err_ref"] = max(res["F_err_ref"], abs(Fg - c["F"][it]) / abs(c["F"][it]))
        res["count_err"] = max(res["count_err"], abs(ag - ao), abs(bg - bo))
        for k in PKEYS:
            res[k + "_err"] = max(res[k + "_err"], maxrel(tg[k], to[k]))
            if k in c:
                res[k + "_err_ref"] = max(res[k + "_err_ref"], maxrel(tg[k], c[k][it]))
    gss = gcont["ss"]
    res["ss_equal"] = bool(np.array_equal(gss, ocont["ss"]))
    res["lpj_err"] = maxrel(gcont["lpj"], ocont["lpj"])
    res["yrec_err"] = maxrel(gdata["y_reconstructed"], odata["y_reconstructed"])
    if iters == total and "y_reconstructed" in c and np.asarray(c["y_reconstructed"]).size:
        res["yrec_err_ref"] = maxrel(gdata["y_reconstructed"], c["y_reconstructed"])
    if iters == total:
        res["ss_equal_ref"] = bool(np.array_equal(gss, c["ss_final_bool"]))
        res["lpj_err_ref"] = maxrel(gcont["lpj"], c["lpj_final"])
    res["iters"] = iters
    return res
