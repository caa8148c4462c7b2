"""Evolutionary K-state container and its operations -- host mirror of `evo/variational`
(utils.py: init_states, vary_Kn; eas.py: evolve_states and the operator names).

The K sets live bit-packed on the GPU (`int32 [N][S][ceil(H/32)]`, latent h -> word h//32 bit h%32)
next to their float64 log-pseudo-joints.  The container is a dict with the reference's keys; `ss`
and `lpj` are materialised from / written to the device copies on access so reference-style code
(`container["ss"]`, `container["lpj"]`) keeps working.
"""
import itertools

import numpy as np
import torch

from . import _lib

_HOST_VIEWS = ("ss", "lpj")


class Operator:
    """The reference's operator callables (eas.py:10-150) as named objects.  Inside `model.step` selection
    and mutation run in the fused CUDA E-step, so the container stores WHICH operator, not a Python function;
    called directly -- `mutation(parents, n_children, sparseness, p_bf) -> children`,
    `selection(candidates, n_parents, lpj) -> parents`, `cross(parents)` -- they run the same device code on
    the arrays given (one `evo_evolve` launch) and return NumPy bool arrays like the reference's functions.
    Random streams: Philox keyed by `OPERATOR_SEED` and a per-process call counter (`seed_operators`)."""

    def __init__(self, kind, name):
        self.kind, self.name = kind, name

    def __call__(self, *args):
        if self.kind == "parent_selection":
            candidates, n_parents, lpj = args
            candidates = np.asarray(candidates, dtype=bool)
            r = _evolve_call(candidates, np.asarray(lpj, dtype=np.float64), None, self.name, "randflip", int(n_parents), 1, 1,
                             False, 0, None, 0.0, 0, OPERATOR_SEED, _next_call(), want=("parents",))
            return r["parents"][0]
        if self.name == "cross":
            (parents,) = args
            n_children, sparseness, p_bf = 1, 0.0, None
        else:
            parents, n_children, sparseness, p_bf = args
        parents = np.asarray(parents, dtype=bool)
        r = _evolve_call(parents, np.zeros(parents.shape[0]), None, "all", self.name, parents.shape[0], int(n_children), 1,
                         False, 0, p_bf, float(sparseness if sparseness is not None else 0.0), 0, OPERATOR_SEED,
                         _next_call(), want=("children",))
        return r["children"][0]

    def __repr__(self):
        return "Operator(%s:%s)" % (self.kind, self.name)


fitparents = Operator("parent_selection", "fit")
randparents = Operator("parent_selection", "rand")
randflip = Operator("mutation_algorithm", "randflip")
sparseflip = Operator("mutation_algorithm", "sparseflip")
cross = Operator("mutation_algorithm", "cross")
cross_randflip = Operator("mutation_algorithm", "cross_randflip")
cross_sparseflip = Operator("mutation_algorithm", "cross_sparseflip")
_SELECTION = {"fit": fitparents, "rand": randparents}
_MUTATION = {"randflip": randflip, "sparseflip": sparseflip, "cross": cross,
             "cross_randflip": cross_randflip, "cross_sparseflip": cross_sparseflip}

OPERATOR_SEED = 0x5EED
_CALLS = [0]


def seed_operators(seed):
    """Seed of the Philox streams behind direct operator / `evolve_states` calls; resets the call counter."""
    global OPERATOR_SEED
    OPERATOR_SEED = int(seed)
    _CALLS[0] = 0


def _next_call():
    _CALLS[0] += 1
    return _CALLS[0]


def words_for(H):
    return (int(H) + 31) // 32


def pack_states(ss, device):
    """(..., H) bool (numpy or torch) -> (..., Hw) int32 CUDA tensor, through the pack kernel."""
    lib = _lib.load()
    t = torch.as_tensor(np.ascontiguousarray(ss) if isinstance(ss, np.ndarray) else ss)
    H = t.shape[-1]
    b = t.to(device=device, dtype=torch.uint8).contiguous()
    rows = b.numel() // H
    out = torch.empty(t.shape[:-1] + (words_for(H),), dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        _lib.check(lib.evo_states_pack(_lib.ptr(b), _lib.ptr(out), rows, H, _lib.stream()), "evo_states_pack")
    return out


def unpack_states(packed, H):
    """(..., Hw) int32 CUDA tensor -> (..., H) bool CUDA tensor, through the unpack kernel."""
    lib = _lib.load()
    packed = packed.contiguous()
    rows = packed.numel() // packed.shape[-1]
    out = torch.empty(packed.shape[:-1] + (H,), dtype=torch.uint8, device=packed.device)
    with torch.cuda.device(packed.device):
        _lib.check(lib.evo_states_unpack(_lib.ptr(packed), _lib.ptr(out), rows, H, _lib.stream()), "evo_states_unpack")
    return out.bool()


def all_states(H):
    """Table of all 2^H states ordered by |s| then combination order (utils.py:58-69)."""
    rows = []
    for g in range(H + 1):
        for comb in itertools.combinations(range(H), g):
            r = np.zeros(H, dtype=bool)
            r[list(comb)] = True
            rows.append(r)
    return np.array(rows, dtype=bool)


class KStates(dict):
    """The `my_suff_stat` dict of the reference (utils.py:143-150,210-226) with device-resident
    `ss_dev` (packed) / `lpj_dev`; `["ss"]` and `["lpj"]` are host views synchronised on access."""

    def __getitem__(self, key):
        if key == "ss":
            return unpack_states(dict.__getitem__(self, "ss_dev"), dict.__getitem__(self, "H")).cpu().numpy()
        if key == "lpj":
            return dict.__getitem__(self, "lpj_dev").cpu().numpy()
        return dict.__getitem__(self, key)

    def __setitem__(self, key, value):
        if key == "ss":
            dev = dict.__getitem__(self, "ss_dev").device
            new = pack_states(np.asarray(value, dtype=bool), dev)
            assert new.shape == dict.__getitem__(self, "ss_dev").shape, "ss shape mismatch"
            dict.__setitem__(self, "ss_dev", new)
            return
        if key == "lpj":
            cur = dict.__getitem__(self, "lpj_dev")
            cur.copy_(torch.as_tensor(np.asarray(value, dtype=np.float64)).to(cur.device).reshape(cur.shape))
            return
        dict.__setitem__(self, key, value)

    def __contains__(self, key):
        return key in _HOST_VIEWS or dict.__contains__(self, key)

    def get(self, key, default=None):
        return self[key] if key in self else default

    def keys(self):
        return list(dict.keys(self)) + list(_HOST_VIEWS)


def init_states(N, S, H, parent_selection, mutation_algorithm, no_parents, no_children, no_generations,
                bitflip_prob=None, Mprime=None, p_init_Kn=None, permanent=None, device=None, seed=0x5EED,
                n_offset=0):
    """K-state container, same arguments and keys as the reference's `init_states`
    (evo/variational/utils.py:155-228).  Extra keyword arguments: `device`, RNG `seed`, and
    `n_offset` = global index of this rank's first datapoint (keeps the Philox streams independent
    of the number of ranks).

    Initial K^n: S distinct Bernoulli(p) states per datapoint drawn on the device (utils.py:100-138);
    S == 2^H enumerates all states (exact E-step, utils.py:71-86)."""
    lib = _lib.load()
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    if permanent is None:
        permanent = {"background": False, "allzero": False, "singletons": False}
    if permanent["background"]:
        S_perm, Hm = 0, H - 1
    else:
        Hm = H
        S_perm = 1 if (permanent["allzero"] == 1 and permanent["singletons"] == 0) else 0
    incl = np.zeros((S_perm, H), dtype=bool)
    sm = all_states(Hm) if Hm < 12 else None

    if S == 2 ** Hm:
        assert Hm < 12, "Exact E-steps too expensive for H={})".format(Hm)
        if permanent["background"]:
            tab = np.concatenate([sm, np.ones((sm.shape[0], 1), dtype=bool)], axis=1)
        else:
            tab = sm[1:] if S_perm == 1 else sm  # the all-zero state is permanent, not a member of K^n
        S_rows = tab.shape[0]
        cols = S_rows + S_perm
        ss_dev = pack_states(tab, device).unsqueeze(0).expand(N, S_rows, words_for(H)).contiguous()
        lpj_dev = torch.zeros((N, cols), dtype=torch.float64, device=device)
    else:
        p0 = 1.0 / H if p_init_Kn is None else float(p_init_Kn)
        ss_dev = torch.empty((N, S, words_for(H)), dtype=torch.int32, device=device)
        lpj_dev = torch.zeros((N, S + S_perm), dtype=torch.float64, device=device)
        with torch.cuda.device(device):
            _lib.check(lib.evo_states_init(_lib.ptr(ss_dev), N, S, H, S_perm, int(bool(permanent["background"])), p0,
                                           int(seed), int(n_offset), _lib.stream()), "evo_states_init")

    if "cross" in mutation_algorithm:
        no_children = no_parents - 1  # utils.py:202-207
    assert no_parents <= S, "no_parents must not exceed S"
    if Mprime is None:
        Mprime = S
    else:
        assert Mprime <= S
    c = KStates()
    c.update({
        "ss_dev": ss_dev, "lpj_dev": lpj_dev, "H": H, "S": ss_dev.shape[1],
        "permanent": permanent, "incl": incl, "S_perm": S_perm, "sm": sm,
        "n_parents": no_parents, "n_children": no_children, "n_generations": no_generations,
        "parent_selection": _SELECTION[parent_selection], "mutation_algorithm": _MUTATION[mutation_algorithm],
        "bitflip_prob": bitflip_prob, "Mprime": Mprime,
        "seed": int(seed), "n_offset": int(n_offset), "iteration": 0, "inject": None,
    })
    return c


def vary_Kn(lpj_old, lpj_new, lpj, states, states_new, H, S, S_perm, incl, Mprime, unification=True,
            reject_worse=True):
    """Top-K replacement for ONE datapoint with the reference's signature (utils.py:231-244); runs
    the `evo_select_topk` kernel on the arrays given and writes `states` / `lpj` back in place."""
    if not unification:
        raise NotImplementedError("unification=False is never used by the reference's callers (SURVEY 8a row 8)")
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    U = int(np.asarray(states_new).shape[0])
    if U == 0:
        lpj[:] = lpj_old
        return 0, 0
    st = pack_states(np.asarray(states, dtype=bool)[None], dev)
    nw = pack_states(np.asarray(states_new, dtype=bool)[None], dev)
    l = torch.zeros((1, S + S_perm), dtype=torch.float64, device=dev)
    l[0, S_perm:] = torch.as_tensor(np.asarray(lpj_old, dtype=np.float64), device=dev)
    ln = torch.as_tensor(np.asarray(lpj_new, dtype=np.float64), device=dev).reshape(1, U).contiguous()
    cnt = torch.zeros((1, 2), dtype=torch.int32, device=dev)
    _lib.check(lib.evo_select_topk(_lib.ptr(st), _lib.ptr(l), 1, S, S_perm, H, int(Mprime), _lib.ptr(nw), _lib.ptr(ln),
                                   None, U, _lib.ptr(cnt), _lib.stream()), "evo_select_topk")
    states[:, :] = unpack_states(st, H)[0].cpu().numpy()
    out = l[0, S_perm:].cpu().numpy()
    lpj[:] = out
    lpj_old[:] = out
    c = cnt.cpu().numpy()[0]
    return int(c[0]), int(c[1])


def pack_injection(children, counts, device):
    """(N,G,C,H) bool + (N,G) int -> device tensors for `evo_estep_args.inj_*`."""
    ch = pack_states(np.asarray(children, dtype=bool), device)
    cnt = torch.as_tensor(np.asarray(counts, dtype=np.int32), device=device).contiguous()
    return ch, cnt


def alloc_trace(lib, ea, N, device, evolve=False):
    """Allocate the trace (and, with `evolve`, the evolve_only output) buffers of an `evo_evolve_args` and point
    the struct at them.  Returns the dict of device tensors (kept alive by the caller)."""
    import ctypes as C

    Cg, Ctot, PK = C.c_int32(), C.c_int32(), C.c_int32()
    _lib.check(lib.evo_evolve_dims(C.byref(ea.e), C.byref(Cg), C.byref(Ctot), C.byref(PK)), "evo_evolve_dims")
    Cg, Ctot, PK = Cg.value, Ctot.value, PK.value
    G, P, Hw = max(int(ea.e.n_generations), 1), max(int(ea.e.n_parents), 1), words_for(ea.e.H)
    i32 = dict(dtype=torch.int32, device=device)
    tr = {
        "parents": torch.zeros((N, G, P, Hw), **i32), "npar": torch.zeros((N, G), **i32),
        "children": torch.zeros((N, G, Cg, Hw), **i32), "count": torch.zeros((N, G), **i32),
        "pos": torch.full((N, G, Cg, 2), -1, **i32), "Cg": Cg, "Ctot": Ctot, "PK": PK,
    }
    ea.tr_parents, ea.tr_npar = tr["parents"].data_ptr(), tr["npar"].data_ptr()
    ea.tr_children, ea.tr_count, ea.tr_pos = tr["children"].data_ptr(), tr["count"].data_ptr(), tr["pos"].data_ptr()
    if evolve:
        tr.update({
            "ev_states": torch.zeros((N, Ctot, Hw), **i32), "ev_lpj": torch.zeros((N, Ctot), dtype=torch.float64, device=device),
            "ev_count": torch.zeros((N, 1 + G), **i32), "ev_pool": torch.zeros((N, PK), **i32),
            "ev_pool_n": torch.zeros((N,), **i32),
        })
        ea.ev_states, ea.ev_lpj, ea.ev_count = tr["ev_states"].data_ptr(), tr["ev_lpj"].data_ptr(), tr["ev_count"].data_ptr()
        ea.ev_pool, ea.ev_pool_n = tr["ev_pool"].data_ptr(), tr["ev_pool_n"].data_ptr()
    return tr


def _evolve_call(states, lpj, pool, selection, mutation, n_parents, n_children, n_generations, background, S_perm, p_bf,
                 sparseness, gen0, seed, iteration, want=("new",)):
    """ONE datapoint through `evo_evolve` (evolve_only, lpj given, new states returned un-evaluated):
    `states` (S,H) bool is the de-duplication set and the index space of `pool`.  Host arrays in / out."""
    import ctypes as C

    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device())
    S, H = states.shape
    ea = _lib.EvolveArgs()
    a = ea.e
    a.N, a.n_offset, a.D, a.H, a.S, a.S_perm, a.Mprime = 1, 0, 1, H, S, int(S_perm), S
    a.n_parents, a.n_children, a.n_generations = int(n_parents), int(n_children), int(n_generations)
    a.selection, a.mutation, a.background = _lib.SEL[selection], _lib.MUT[mutation], int(bool(background))
    a.bitflip_prob = float(p_bf) if p_bf is not None else 0.0
    a.sparseness = float(sparseness)
    a.seed, a.iteration = int(seed), int(iteration)
    st = pack_states(states[None], dev)
    lp = torch.zeros((1, S + int(S_perm)), dtype=torch.float64, device=dev)
    lp[0, int(S_perm):] = torch.as_tensor(np.asarray(lpj, dtype=np.float64), device=dev)
    a.states, a.lpj = st.data_ptr(), lp.data_ptr()
    # the evaluator is never consulted (lpj given, new states un-evaluated): a one-column dummy table
    Wd = torch.zeros((1, H), dtype=torch.float64, device=dev)
    Yd = torch.zeros((1, 1), dtype=torch.float32, device=dev)
    a.eval, a.W, a.Y = _lib.EVAL_BSC_DIRECT, Wd.data_ptr(), Yd.data_ptr()
    a.pre1, a.pil_bar = -0.5, 0.0
    ea.evolve_only, ea.lpj_given, ea.skip_eval, ea.gen0 = 1, 1, 1, int(gen0)
    keep = None
    if pool is not None:
        pl = torch.zeros((1, S), dtype=torch.int32, device=dev)
        pl[0, :len(pool)] = torch.as_tensor(np.asarray(pool, dtype=np.int32), device=dev)
        pn = torch.tensor([len(pool)], dtype=torch.int32, device=dev)
        ea.pool0, ea.pool0_n = pl.data_ptr(), pn.data_ptr()
        keep = (pl, pn)
    with torch.cuda.device(dev):
        tr = alloc_trace(lib, ea, 1, dev, evolve=True)
        _lib.check(lib.evo_evolve(C.byref(ea), _lib.stream()), "evo_evolve")
    out = {}
    if "children" in want:
        cnt = tr["count"].cpu().numpy()[0]
        ch = unpack_states(tr["children"], H).cpu().numpy()[0]
        out["children"] = [ch[g, :cnt[g]] for g in range(ch.shape[0])]
    if "parents" in want:
        npar = tr["npar"].cpu().numpy()[0]
        pa = unpack_states(tr["parents"], H).cpu().numpy()[0]
        out["parents"] = [pa[g, :npar[g]] for g in range(pa.shape[0])]
    if "new" in want:
        U = int(tr["ev_count"][0, 0].item())
        out["new"] = unpack_states(tr["ev_states"], H).cpu().numpy()[0, :U]
        out["pool"] = tr["ev_pool"].cpu().numpy()[0, :int(tr["ev_pool_n"][0].item())]
    del keep, Wd, Yd
    return out


def evolve_states(my_suff_stat, model_params, eval_lpj):
    """The reference's `evolve_states(my_suff_stat, model_params, eval_lpj) -> (new_states, new_lpj)` for ONE
    datapoint (evo/variational/eas.py:153-313): reads `this_states` (K,H), `this_lpj` (K,), `incl`,
    `permanent`, the EA hyper-parameters and the operator names from the container, `piH` from
    `model_params`, and calls `eval_lpj(states) -> lpj` (any Python callable, e.g. the model's
    `log_pseudo_joint` wrapped like `_models.py:515-521`) on the genuinely new states of every generation.

    Selection, mutation, de-duplication and the next parent pool run on the device, one `evo_evolve` launch per
    generation (the callback lives on the host; `model.step` / `model.evolve` use the fused form where the
    evaluator is on the device too).  New states come back generation by generation in lexicographic order
    (latent 0 most significant), like the reference's `np.unique`; re-discovered old states re-enter the pool
    with their own lpj (documented difference (ii))."""
    import warnings

    ss = my_suff_stat
    s = np.asarray(ss["this_states"], dtype=bool)
    lpj = np.asarray(ss["this_lpj"], dtype=np.float64)
    K, H = s.shape
    S_perm = int(np.asarray(ss["incl"]).shape[0])
    background = bool(ss["permanent"]["background"])
    sel, mut = ss["parent_selection"].name, ss["mutation_algorithm"].name
    G = int(ss["n_generations"])
    call = _next_call()
    uniq, uniq_lpj = s.copy(), lpj.copy()
    pool = np.arange(K, dtype=np.int32)
    out_s, out_l = [], []
    for g in range(G):
        r = _evolve_call(uniq, uniq_lpj, pool, sel, mut, int(ss["n_parents"]), int(ss["n_children"]), 1, background,
                         S_perm, ss["bitflip_prob"], float(model_params["piH"]), g, int(ss.get("seed", OPERATOR_SEED)),
                         call)
        new, nxt = r["new"], r["pool"]
        n_old = uniq.shape[0]
        if new.shape[0]:
            order = np.lexsort(new.T[::-1])  # rows ascending with latent 0 most significant (eas.py:252-256)
            inv = np.empty_like(order)
            inv[order] = np.arange(order.size)
            new = new[order]
            nxt = np.where(nxt >= n_old, n_old + inv[np.clip(nxt - n_old, 0, order.size - 1)], nxt).astype(np.int32)
            new_lpj = np.asarray(eval_lpj(new), dtype=np.float64)
            uniq = np.concatenate([uniq, new])
            uniq_lpj = np.concatenate([uniq_lpj, new_lpj])
            out_s.append(new)
            out_l.append(new_lpj)
        if nxt.size == 0:
            warnings.warn("No new and unique states. Skipping evolutionary loop.")  # eas.py:306-308
            break
        pool = nxt
    if not out_s:
        return np.zeros((0, H), dtype=bool), np.zeros(0)
    return np.concatenate(out_s), np.concatenate(out_l)
