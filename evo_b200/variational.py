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
    """Named stand-in for the reference's operator callables (eas.py:10-150).  Selection and
    mutation run inside the fused E-step kernel, so the container stores which one, not a Python
    function.  Calling it explains how to inject candidates instead."""

    def __init__(self, kind, name):
        self.kind, self.name = kind, name

    def __call__(self, *a, **k):
        raise NotImplementedError(
            "%s '%s' runs inside the fused CUDA E-step; to supply your own candidates set "
            "container['inject'] = (children[N,G,C,H] bool, counts[N,G])" % (self.kind, self.name)
        )

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
