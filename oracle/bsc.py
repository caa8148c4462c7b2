"""CPU oracle -- Binary Sparse Coding E-/M-step (TEST INFRASTRUCTURE, not product code).

NumPy float64 restatement of `evo/models/_models.py` (E_step, free_energy, lpj_reset_check,
reconstruct, check_params) and `evo/models/bsc.py` (precompute, log-pseudo-joint, M-step).
Parity status: PINNED against the unmodified reference by `tests/golden/make_golden.py`
(see oracle/variational.py header).  Never imported by `evo_b200/`.
"""
import numpy as np

from . import variational as V

EPS_LPJ = np.finfo(np.float64).min  # bsc.py:23-24
B_MAX = 0.0  # _models.py:55
TOL = 1e-5  # _models.py:47


def check_params(theta):
    """_models.py:102-159 with BSC's noise policy (:47-52): pi in [1e-5, 1-1e-5] (upper test is >=),
    sigma >= 1e-5."""
    pi, sigma = theta["pi"], theta["sigma"]
    if pi < TOL:
        pi = TOL
    if pi >= 1.0 - TOL:
        pi = 1.0 - TOL
    if sigma < TOL:
        sigma = TOL
    theta["pi"], theta["sigma"] = pi, sigma
    return theta


def precompute(theta, D, H, N_total, n_obs_total=None):
    """bsc.py:100-125.  `n_obs_total` = number of observed entries over all ranks when data is
    incomplete, else None."""
    pi, sigma = theta["pi"], theta["sigma"]
    theta["piH"] = pi * H
    theta["pre1"] = -1.0 / 2.0 / sigma / sigma
    theta["pil_bar"] = np.log(pi / (1.0 - pi))
    if n_obs_total is not None:
        theta["ljc"] = H * np.log(1.0 - pi) - np.log(2 * np.pi * sigma * sigma) * n_obs_total / N_total / 2
    else:
        theta["ljc"] = H * np.log(1.0 - pi) - D / 2 * np.log(2 * np.pi * sigma * sigma)
    return theta


def lpj_reset_check(lpj, counters=None):
    """_models.py:567-596: NaN -> f64 min, < f64 min -> f64 min, +-inf -> B_max = 0.0; the masks are
    taken before any write, so -inf ends at 0.0.  One counter per call."""
    nan = np.isnan(lpj)
    small = lpj < EPS_LPJ
    inf = np.isinf(lpj)
    if counters is not None:
        if nan.any():
            counters["reset_lpj_isnan"] += 1
        elif small.any():
            counters["reset_lpj_smaller_eps_lpj"] += 1
        elif inf.any():
            counters["reset_lpj_isinf"] += 1
    lpj[nan] = EPS_LPJ
    lpj[small] = EPS_LPJ
    lpj[inf] = B_MAX
    return lpj


def lpj_states(states, W, y, x_infr, pre1, pil_bar, counters=None):
    """bsc.py:79-97: pre1 * sum_obs ((s W^T)_d - y_d)^2 + pil_bar * |s| for a batch of states."""
    states = np.asarray(states, dtype=bool)
    Wo = W[x_infr, :]
    mean = states.astype(np.float64) @ Wo.T
    lpj = pre1 * ((mean - y[x_infr]) ** 2).sum(axis=1) + pil_bar * states.sum(axis=1)
    return lpj_reset_check(lpj, counters)


def lpj_allzero(y, x_infr, pre1, counters=None):
    """bsc.py:60-76."""
    return lpj_reset_check(np.array([pre1 * (y[x_infr] ** 2).sum()]), counters)


def free_energy_from_lpj(lpj, ljc, N_total):
    """_models.py:544-547 on one rank: returns (sum_n logsumexp, F)."""
    m = lpj.max(axis=1)
    Fs = (m + np.log(np.exp(lpj - m[:, None]).sum(axis=1))).sum()
    return Fs, ljc + Fs / N_total


def e_step(theta, container, data, D, H, S, N_total=None, reference_quirks=True):
    """_models.py:454-565 on one rank.  Mutates container['ss'], container['lpj'] in place and
    returns (Fs = sum_n logsumexp_n, sum of new-unique counts, sum of substitutions)."""
    y, x_infr = data["y"], data["x_infr"]
    my_N = y.shape[0]
    lpj, ss, incl = container["lpj"], container["ss"], container["incl"]
    S_perm, permanent, Mprime = container["S_perm"], container["permanent"], container["Mprime"]
    counters = container.setdefault("_counters", {})
    for k in ("reset_lpj_isnan", "reset_lpj_smaller_eps_lpj", "reset_lpj_isinf"):
        counters[k] = 0
    W, pre1, pil_bar = theta["W"], theta["pre1"], theta["pil_bar"]
    n_uni = n_sub = 0
    for n in range(my_N):
        yn, xn, states = y[n], x_infr[n], ss[n]
        if not permanent["background"] and S_perm > 0:
            lpj[n, :S_perm] = lpj_allzero(yn, xn, pre1, counters)
        this_lpj = lpj_states(states, W, yn, xn, pre1, pil_bar, counters)
        container["this_states"], container["this_lpj"] = states, this_lpj
        new_s, new_l = V.evolve_states(
            container, theta, lambda st: lpj_states(st, W, yn, xn, pre1, pil_bar, counters),
            reference_quirks=reference_quirks)
        a, b = V.vary_Kn(this_lpj, new_l, lpj[n, S_perm:], states, new_s, H, S, S_perm, incl, Mprime)
        n_uni += a
        n_sub += b
    Fs, _ = free_energy_from_lpj(lpj, 0.0, 1)
    return Fs, n_uni, n_sub


def posterior(lpj):
    """q_nk = exp(lpj - max) / sum (bsc.py:176-177, 219)."""
    p = np.exp(lpj - lpj.max(axis=1, keepdims=True))
    return p / p.sum(axis=1, keepdims=True)


def reconstruct(theta, container, data):
    """_models.py:614-665 + bsc.py:279-287: y_rec[n, ~x] = W[~x] <s>_n; rows with no usable entry
    are skipped."""
    y = data["y"].copy()
    q = posterior(container["lpj"])[:, container["S_perm"]:]
    es = np.einsum("nk,nkh->nh", q, container["ss"].astype(np.float64))
    est = es @ theta["W"].T
    for n in range(y.shape[0]):
        if not data["x_infr"][n].any():
            continue
        miss = ~data["x"][n]
        y[n, miss] = est[n, miss]
    data["y_reconstructed"] = y
    return y


def m_step_stats(theta, container, data):
    """bsc.py:176-223 on one rank: (Wp (H,D), Wq (H,H), pies (H,), sigma_sum)."""
    lpj, ss = container["lpj"], container["ss"]
    S_perm, permanent = container["S_perm"], container["permanent"]
    x_infr = data["x_infr"]
    incomplete = not x_infr.all()
    y = data["y_reconstructed"] if incomplete else data["y"]
    W = theta["W"]
    N, S, H = ss.shape
    qall = posterior(lpj)
    q = qall[:, S_perm:]
    es = np.einsum("nk,nkh->nh", q, ss.astype(np.float64))
    pies = es.sum(axis=0)
    Wp = es.T @ y
    Wq = np.zeros((H, H))
    sig = 0.0
    for n in range(N):
        sn = ss[n].astype(np.float64)
        Wq += (sn * q[n][:, None]).T @ sn
        obs = x_infr[n]
        yo = data["y"][n][obs] if not incomplete else y[n][obs]
        r = ((yo - sn @ W[obs, :].T) ** 2).sum(axis=1)
        sig += (q[n] * r).sum()
        if permanent["allzero"]:
            sig += qall[n, 0] * (yo ** 2).sum()
    return Wp, Wq, pies, sig


def m_step_solve(theta, Wp, Wq, pies, sig, N_total, D, H, to_learn=("W", "pi", "sigma"),
                 n_obs_total=None, background=False):
    """bsc.py:226-275 (replicated solve).  rcond=-1 is what the reference's version test selects
    under NumPy 2.x (bsc.py:232-235)."""
    sigma_old = theta["sigma"]
    if "W" in to_learn:
        theta["W"] = np.linalg.lstsq(Wq, Wp, rcond=-1)[0].T
    if "pi" in to_learn:
        p = pies / N_total
        if background:
            p[-1] = 1.0 - 1.1e-5
        theta["pi"] = p.sum() / H
        theta["pies"] = p
    if "sigma" in to_learn:
        if n_obs_total is not None:
            theta["sigma"] = np.sqrt((sig + n_obs_total * sigma_old ** 2) / N_total / D)
        else:
            theta["sigma"] = np.sqrt(sig / N_total / D)
    return theta


def step(theta, container, data, D, H, S, to_learn=("W", "pi", "sigma"), do_reconstruction=False,
         reference_quirks=True):
    """One single-rank EM iteration, Model.step (_models.py:162-203).  Returns
    (F, S_nunique, S_sub, theta)."""
    N = data["y"].shape[0]
    incomplete = not data["x_infr"].all()
    n_obs = int(data["x_infr"].sum()) if incomplete else None
    theta = check_params(theta)
    precompute(theta, D, H, N, n_obs)
    Fs, n_uni, n_sub = e_step(theta, container, data, D, H, S, reference_quirks=reference_quirks)
    F = theta["ljc"] + Fs / N
    if do_reconstruction:
        reconstruct(theta, container, data)
    if len(to_learn):
        Wp, Wq, pies, sig = m_step_stats(theta, container, data)
        theta = m_step_solve(theta, Wp, Wq, pies, sig, N, D, H, to_learn, n_obs,
                             container["permanent"]["background"])
    return F, n_uni / N, n_sub / N, theta


def free_energy_full(theta, data, D, H, sm):
    """Exact log-likelihood over all 2^H states, Model.free_energy(full=True) (_models.py:334-451):
    the all-zero state is evaluated as a permanent state, the other 2^H - 1 rows as ordinary ones."""
    y, x_infr = data["y"], data["x_infr"]
    N = y.shape[0]
    incomplete = not x_infr.all()
    precompute(theta, D, H, N, int(x_infr.sum()) if incomplete else None)
    rest = sm[1:]
    lpj = np.zeros((N, 1 + rest.shape[0]))
    for n in range(N):
        lpj[n, 0] = lpj_allzero(y[n], x_infr[n], theta["pre1"])[0]
        lpj[n, 1:] = lpj_states(rest, theta["W"], y[n], x_infr[n], theta["pre1"], theta["pil_bar"])
    return free_energy_from_lpj(lpj, theta["ljc"], N)[1]
