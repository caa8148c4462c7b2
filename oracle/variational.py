"""CPU oracle -- evolutionary K-state machinery (TEST INFRASTRUCTURE, not product code).

Restates, in plain NumPy/Python, what the reference's `evo/variational/eas.py` and
`evo/variational/utils.py` compute.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import this package; the product
(`evo_b200/`) never does.

Parity status: PINNED.  `tests/golden/make_golden.py` runs the unmodified reference (imported from
/root/reference with a single-rank mpi4py stand-in) with recording wrappers around its injectable
`parent_selection` / `mutation_algorithm` callables and checks this restatement against it
bit-for-bit (states, slots) and to 1e-12 (floats); the recorded tapes and outputs are committed as
`tests/golden/*.npz` and re-checked by `tests/test_oracle_golden.py`.

The mechanics differ from the reference on purpose (hash-set de-duplication instead of
`np.unique` on void views, explicit rank pairing instead of argpartition/unravel_index), the
results do not.
"""
import warnings

import numpy as np


# --------------------------------------------------------------------------------------------
# packed representation shared with the CUDA side: bit h -> word h // 32, bit h % 32, pad bits 0
# --------------------------------------------------------------------------------------------
def words_for(H):
    return (int(H) + 31) // 32


def pack_states(states):
    """(..., H) bool -> (..., ceil(H/32)) uint32."""
    states = np.asarray(states, dtype=bool)
    H = states.shape[-1]
    Hw = words_for(H)
    padded = np.zeros(states.shape[:-1] + (Hw * 32,), dtype=np.uint8)
    padded[..., :H] = states
    by = np.packbits(padded, axis=-1, bitorder="little")  # byte b holds bits 8b..8b+7
    return np.ascontiguousarray(by).view("<u4").reshape(states.shape[:-1] + (Hw,))


def unpack_states(words, H):
    """(..., Hw) uint32 -> (..., H) bool."""
    words = np.ascontiguousarray(np.asarray(words, dtype="<u4"))
    by = words.view(np.uint8).reshape(words.shape[:-1] + (words.shape[-1] * 4,))
    bits = np.unpackbits(by, axis=-1, bitorder="little")
    return bits[..., :H].astype(bool)


def _key(row):
    # bytes of a bool row compare lexicographically with h=0 most significant, which is the order
    # np.unique gives on the reference's void view (eas.py:252-256) [SURVEY 4.4]
    return np.asarray(row, dtype=bool).tobytes()


# --------------------------------------------------------------------------------------------
# operators (eas.py:10-150) -- RNG draws mirror the reference call-for-call so that seeding the
# legacy global np.random state gives the same children (used for CPU-baseline timing)
# --------------------------------------------------------------------------------------------
def randflip(parents, n_children, sparseness=None, p_bf=None):
    """eas.py:10-43: every parent -> n_children copies, copy j flips the j-th of n_children
    distinct uniformly drawn positions."""
    P, H = parents.shape
    scores = np.random.rand(P, H)
    pos = np.argpartition(scores, n_children - 1, axis=1)[:, :n_children]
    children = np.repeat(parents, n_children, axis=0)
    rows = np.arange(P * n_children)
    children[rows, pos.reshape(-1)] ^= True
    return children


def sparseflip(parents, n_children, sparseness, p_bf):
    """eas.py:46-100: independent per-bit flips, p1 for on-bits, p0 for off-bits."""
    assert p_bf is not None, "Please specify the bitflip probability"
    P, H = parents.shape
    a = parents.sum(axis=1).astype(float)
    tiny = 1e-100
    alpha = (H - a) * (H * p_bf - (sparseness - a)) / ((sparseness - a + H * p_bf) * a + tiny)
    p0 = H * p_bf / (H + (alpha - 1.0) * a + tiny)
    p1 = alpha * p0
    children = np.repeat(parents, n_children, axis=0)
    prob = np.where(children, np.repeat(p1, n_children)[:, None], np.repeat(p0, n_children)[:, None])
    flips = np.random.random((P * n_children, H)) < prob
    return children ^ flips


def cross(parents):
    """eas.py:103-125: one-point crossover of every unordered parent pair, 2 children per pair."""
    P, H = parents.shape
    out = np.empty((P * (P - 1), H), dtype=bool)
    c = 0
    for i in range(P):
        for j in range(i + 1, P):
            cp = np.random.randint(low=1, high=H)
            out[c, :cp], out[c, cp:] = parents[i, :cp], parents[j, cp:]
            out[c + 1, :cp], out[c + 1, cp:] = parents[j, :cp], parents[i, cp:]
            c += 2
    return out


def cross_randflip(parents, n_children, sparseness, p_bf):
    return randflip(cross(parents), 1, sparseness, p_bf)


def cross_sparseflip(parents, n_children, sparseness, p_bf):
    return sparseflip(cross(parents), 1, sparseness, p_bf)


def fitparents(candidates, n_parents, lpj):
    """eas.py:138-146: fitness-proportional sampling without replacement."""
    w = lpj - 2.0 * min(float(np.min(lpj)), 0.0)
    w = w / w.sum()
    return candidates[np.random.choice(candidates.shape[0], size=n_parents, replace=False, p=w)]


def randparents(candidates, n_parents, lpj=None):
    """eas.py:149-150."""
    return candidates[np.random.choice(candidates.shape[0], size=n_parents, replace=False)]


SELECTION = {"fit": fitparents, "rand": randparents}
MUTATION = {
    "randflip": randflip,
    "sparseflip": sparseflip,
    "cross": cross,
    "cross_randflip": cross_randflip,
    "cross_sparseflip": cross_sparseflip,
}


# --------------------------------------------------------------------------------------------
# evolve_states (eas.py:153-313)
# --------------------------------------------------------------------------------------------
def evolve_states(container, theta, eval_lpj, reference_quirks=True, return_pools=False):
    """One datapoint's evolutionary loop.

    `reference_quirks=True` reproduces the reference's bookkeeping of re-discovered old states
    exactly (eas.py:271-295): the index window is built from round 0's child count, the paired lpj
    is shifted by one when S_perm == 0, s_unique[0] can never re-enter the pool, and nothing is
    appended when every candidate was re-discovered.  `False` gives the intended behaviour (each
    re-discovered non-permanent state re-enters the pool with its own lpj) -- that is what the
    CUDA path implements; it only changes later generations' parent pools.
    """
    lpj = container["this_lpj"]
    s = container["this_states"]
    incl = container["incl"]
    background = container["permanent"]["background"]
    n_parents = container["n_parents"]
    n_children = container["n_children"]
    n_generations = container["n_generations"]
    select = container["parent_selection"]
    mutate = container["mutation_algorithm"]
    p_bf = container["bitflip_prob"]
    sparseness = theta["piH"]

    K, H = s.shape
    S_perm = incl.shape[0]
    mutable = H - 1 if background else H

    uniq_states = [np.asarray(r, dtype=bool) for r in incl] + [np.asarray(r, dtype=bool) for r in s]
    uniq_lpj = list(np.asarray(lpj, dtype=float))
    index_of = {}
    for i, r in enumerate(uniq_states):
        index_of.setdefault(_key(r), i)

    out_states, out_lpj = [], []
    pools = []
    pool_states, pool_lpj = s, lpj
    cpg = None
    for g in range(n_generations):
        take = min(pool_states.shape[0], n_parents)
        parents = select(pool_states, take, pool_lpj)
        children = mutate(parents[:, :mutable], n_children, sparseness, p_bf)
        if background:
            children = np.concatenate([children, np.ones((children.shape[0], 1), dtype=bool)], axis=1)
        children = np.asarray(children, dtype=bool)
        c = children.shape[0]
        if cpg is None:
            cpg = c

        child_keys = [_key(r) for r in children]
        fresh = {}
        for i, k in enumerate(child_keys):
            if k not in index_of and k not in fresh:
                fresh[k] = i
        order = sorted(fresh)  # lexicographic, h=0 most significant
        new_g = children[[fresh[k] for k in order]] if order else np.zeros((0, H), dtype=bool)
        lpj_g = np.asarray(eval_lpj(new_g), dtype=float) if order else np.zeros(0)

        n_unique = len(uniq_states)
        produced = set(child_keys)
        if reference_quirks:
            delta = cpg - c
            lo, hi = max(0, 1 - delta), n_unique - 1 - delta
            missed = {i - 1 + delta for i in range(lo, hi + 1) if _key(uniq_states[i]) not in produced}
            if missed:
                again = [t for t in range(n_unique - 1) if t not in missed]
            else:
                again = []
            re_states = [uniq_states[t + 1] for t in again]
            re_lpj = [uniq_lpj[t] for t in again]
        else:
            again = [i for i in range(S_perm, n_unique) if _key(uniq_states[i]) in produced]
            re_states = [uniq_states[i] for i in again]
            re_lpj = [uniq_lpj[i - S_perm] for i in again]

        for r, v in zip(new_g, lpj_g):
            index_of[_key(r)] = len(uniq_states)
            uniq_states.append(r)
            uniq_lpj.append(float(v))
        out_states.extend(new_g)
        out_lpj.extend(lpj_g)

        if len(new_g) + len(re_states) == 0:
            warnings.warn("No new and unique states. Skipping evolutionary loop.")
            break
        pool_states = np.array(list(new_g) + re_states, dtype=bool).reshape(-1, H)
        pool_lpj = np.array(list(lpj_g) + re_lpj, dtype=float)
        pools.append((pool_states, pool_lpj))

    new_states = np.array(out_states, dtype=bool).reshape(-1, H)
    new_lpj = np.array(out_lpj, dtype=float)
    if return_pools:
        return new_states, new_lpj, pools
    return new_states, new_lpj


# --------------------------------------------------------------------------------------------
# vary_Kn (variational/utils.py:231-337)
# --------------------------------------------------------------------------------------------
def vary_Kn(lpj_old, lpj_new, lpj, states, states_new, H, S, S_perm, incl, Mprime,
            unification=True, reject_worse=True):
    """Replace the worst old states of K^n by the best new ones, in place.

    Drops new rows equal to a permanent state, a member of K^n or an earlier new row (survivors in
    lexicographic order); with M' = min(#survivors, Mprime), the j-th best survivor is accepted iff
    it beats the j-th worst old state and then takes that state's slot.  Ties at
    the acceptance boundary keep the old state (NumPy's order there is unspecified; A.6(i)).
    """
    seen = {_key(r) for r in incl} | {_key(r) for r in states}
    fresh = {}
    for i, r in enumerate(states_new):
        k = _key(r)
        if k not in seen and k not in fresh:
            fresh[k] = i
    idx = np.array([fresh[k] for k in sorted(fresh)], dtype=int)

    if not unification:
        if reject_worse and (np.sum(lpj_new) < np.sum(lpj_old)):
            lpj[:] = lpj_old
            return 0, 0
        lpj[:] = lpj_new
        states[:, :] = states_new
        return idx.size, idx.size

    cand_states = states_new[idx] if idx.size else np.zeros((0, H), dtype=bool)
    cand_lpj = np.asarray(lpj_new)[idx] if idx.size else np.zeros(0)
    M = min(idx.size, Mprime)
    n_sub = 0
    if M > 0:
        best_new = np.argsort(-cand_lpj, kind="stable")[:M]
        worst_old = np.argsort(lpj_old, kind="stable")[:M]
        for j in range(M):
            gi, bi = best_new[j], worst_old[j]
            if cand_lpj[gi] > lpj_old[bi]:
                states[bi] = cand_states[gi]
                lpj_old[bi] = cand_lpj[gi]
                n_sub += 1
            else:
                break
    lpj[:] = lpj_old
    return int(idx.size), n_sub


# --------------------------------------------------------------------------------------------
# init_states (variational/utils.py:19-228)
# --------------------------------------------------------------------------------------------
def all_states(H):
    """Table of all 2^H states ordered by |s| then by combination order (utils.py:58-69)."""
    from itertools import combinations

    rows = []
    for g in range(H + 1):
        for comb in combinations(range(H), g):
            r = np.zeros(H, dtype=bool)
            r[list(comb)] = True
            rows.append(r)
    return np.array(rows, dtype=bool)


def init_states(N, S, H, parent_selection, mutation_algorithm, no_parents, no_children,
                no_generations, bitflip_prob=None, Mprime=None, p_init_Kn=None, permanent=None):
    """K-state container (dict) with the reference's keys.  Draw-for-draw identical to the reference
    under the same legacy np.random seed."""
    if permanent is None:
        permanent = {"background": False, "allzero": False, "singletons": False}
    if permanent["background"]:
        S_perm, Hm = 0, H - 1
    else:
        Hm = H
        S_perm = 1 if (permanent["allzero"] == 1 and permanent["singletons"] == 0) else 0
    incl = np.zeros((S_perm, Hm), dtype=bool)
    sm = all_states(Hm) if Hm < 12 else None

    if S == 2 ** Hm:
        assert Hm < 12, "Exact E-steps too expensive for H={})".format(Hm)
        if permanent["background"]:
            ss = np.concatenate([sm, np.ones((sm.shape[0], 1), dtype=bool)], axis=1)
            lpj = np.empty((N, 2 ** Hm))
        else:
            lpj = np.empty((N, S + S_perm))
            ss = sm[1:].copy() if S_perm == 1 else sm.copy()
        ss = np.tile(ss[None], (N, 1, 1))
    else:
        p0 = 1.0 / H if p_init_Kn is None else p_init_Kn
        lpj = np.empty((N, S + S_perm))
        ss = np.empty((N, S, H), dtype=bool)
        if permanent["background"]:
            ss[:, :, -1] = True
        for n in range(N):
            have, seen = [], {_key(r) for r in incl}
            first = True
            while len(have) < S:
                draw = np.random.random(size=(S, Hm)) < p0
                fresh = {}
                for i, r in enumerate(draw):
                    k = _key(r)
                    if k not in seen and k not in fresh:
                        fresh[k] = i
                for k in sorted(fresh):
                    have.append(draw[fresh[k]])
                    seen.add(k)
                first = False
            ss[n, :, :Hm] = np.array(have[:S])
    if permanent["background"]:
        incl = np.zeros((S_perm, H), dtype=bool)

    if "cross" in mutation_algorithm:
        no_children = no_parents - 1
    assert no_parents <= S
    if Mprime is None:
        Mprime = S
    else:
        assert Mprime <= S
    return {
        "ss": ss, "lpj": lpj, "permanent": permanent, "incl": incl, "S_perm": S_perm, "sm": sm,
        "n_parents": no_parents, "n_children": no_children, "n_generations": no_generations,
        "parent_selection": SELECTION[parent_selection],
        "mutation_algorithm": MUTATION[mutation_algorithm],
        "bitflip_prob": bitflip_prob, "Mprime": Mprime,
    }


def split_rows(N, R):
    """Row counts per rank by the np.array_split rule (parallel.py:88-114)."""
    base, extra = divmod(int(N), int(R))
    return [base + (1 if r < extra else 0) for r in range(R)]
