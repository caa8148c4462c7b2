"""CPU oracle -- spike-and-slab sparse coding (ES3C, class `SSSC`) E-/M-step.
TEST INFRASTRUCTURE, not product code: only tests/, __graft_entry__.smoke() and bench.py's CPU legs
import it; `evo_b200/` never does.

NumPy float64 restatement of `evo/models/sssc.py` (check_params :199, log_pseudo_joint :242-326,
E_step_precompute :329-366, modelmean :369-405, EM_step :420-813).  The per-state D_obs x D_obs
matrix C^-1 of the reference (:307-309) is never formed: r'C^-1 r = |r|^2/sigma2 - v' Lambda v with
v = W_A' r / sigma2 (Woodbury), which is the same number to 1e-15.

Parity status: PINNED against the unmodified reference by `tests/golden/make_golden_sssc.py`
(recorded candidate tapes replayed; fixtures `tests/golden/sssc_*.npz`).
"""
import numpy as np

from . import variational as V
from .bsc import lpj_reset_check, posterior  # noqa: F401  (same sanitising rule, _models.py:567-596)

TOL = 1e-5          # sssc.py:40 (noise policy bound)
EPS_W = 5e-5        # sssc.py:39,44 (noise added to a non-invertible W update)
EPS_PIES = 5e-5     # sssc.py:39,45
EPS_MUS = np.finfo(np.float64).eps   # sssc.py:46
EPS_PSI = 1e-5      # sssc.py:47
EPS_SIGMA2 = 1e-5   # sssc.py:48
EPS_PJC_SUM = np.finfo(np.float64).tiny  # sssc.py:43


def check_params(theta):
    """_models.py:102-159 with SSSC's noise policy (sssc.py:51-58)."""
    theta["pies"] = np.minimum(1.0 - TOL, np.maximum(TOL, np.asarray(theta["pies"], dtype=float)))
    Psi = np.array(theta["Psi"], dtype=float)
    d = np.diag(Psi).copy()
    d[d < TOL] = TOL
    Psi[np.arange(Psi.shape[0]), np.arange(Psi.shape[0])] = d
    theta["Psi"] = Psi
    theta["sigma2"] = np.float64(max(float(theta["sigma2"]), TOL))
    return theta


def precompute(theta, D, N_total, n_obs_total=None):
    """sssc.py:329-366."""
    pies, sigma2 = theta["pies"], np.float64(theta["sigma2"])
    theta["ljc"] = np.log(1.0 - pies).sum() - D / 2 * np.log(2 * np.pi)
    theta["piH"] = pies.sum()
    theta["pil_bar"] = np.log(pies / (1.0 - pies))
    s128 = sigma2.astype("longdouble")
    theta["sigma2_inv"] = np.float64(1.0 / s128)
    theta["ljc"] -= 0.5 * np.float64(D * np.log(s128))
    if n_obs_total is not None:
        theta["ljc"] = np.log(1.0 - pies).sum() + (-np.log(2 * np.pi) - np.log(sigma2)) * n_obs_total / N_total / 2
    return theta


def state_terms(state, W_obs, Psi, mus, sigma2_inv):
    """Per-state quantities of sssc.py:270-318 (without C_inv): W_A mu_A, C_det, Lambda, Lambda W_A'/sigma2."""
    A = np.flatnonzero(state)
    W_s = W_obs[:, A]
    Psi_s = Psi[np.ix_(A, A)]
    try:
        Psi_s_inv = np.linalg.inv(Psi_s)
    except np.linalg.LinAlgError:
        Psi_s_inv = np.linalg.pinv(Psi_s)
    log_det_Psi = np.linalg.slogdet(Psi_s)[1] if A.size else 0.0
    M = W_s.T @ (sigma2_inv * W_s) + Psi_s_inv
    log_det_M = np.linalg.slogdet(M)[1] if A.size else 0.0
    try:
        Lam = np.linalg.inv(M)
    except np.linalg.LinAlgError:
        Lam = np.linalg.pinv(M)
    return {"A": A, "Wmu": W_s @ mus[A], "C_det": log_det_M + log_det_Psi, "Lam": Lam, "W_s": W_s,
            "LW": (Lam @ W_s.T) * sigma2_inv}


def lpj_states(states, theta, y, x_infr, counters=None, cache=None):
    """sssc.py:242-326 for one datapoint; returns (lpj, list of per-state terms)."""
    W_obs = theta["W"][x_infr, :]
    yo = y[x_infr]
    s2i = theta["sigma2_inv"]
    out = np.zeros(states.shape[0])
    terms = []
    for k, st in enumerate(np.asarray(states, dtype=bool)):
        t = state_terms(st, W_obs, theta["Psi"], theta["mus"], s2i)
        r = yo - t["Wmu"]
        v = (t["W_s"].T @ r) * s2i
        quad = s2i * (r @ r) - v @ t["Lam"] @ v
        out[k] = -0.5 * (t["C_det"] + quad) + theta["pil_bar"][t["A"]].sum()
        terms.append(t)
    return lpj_reset_check(out, counters), terms


def lpj_allzero(y, x_infr, sigma2_inv, counters=None):
    """sssc.py:225-239."""
    return lpj_reset_check(np.array([-0.5 * (y[x_infr] ** 2).sum() * sigma2_inv]), counters)


def em_step(theta, container, data, D, H, S, to_learn=("W", "pies", "mus", "sigma2", "Psi"),
            do_reconstruction=False, reference_quirks=True):
    """One single-rank SSSC.step (sssc.py:408-813).  Mutates container / theta in place; returns
    (F, S_nunique, S_sub, theta).  `stats` of the last call are kept in container['_stats']."""
    y_all, x_infr_all = data["y"], data["x_infr"]
    N = y_all.shape[0]
    incomplete = not x_infr_all.all()
    theta = check_params(theta)
    precompute(theta, D, N, int(x_infr_all.sum()) if incomplete else None)
    W, mus, sigma2 = theta["W"], theta["mus"], np.float64(theta["sigma2"])
    lpj, ss, incl = container["lpj"], container["ss"], container["incl"]
    S_perm, permanent, Mprime = container["S_perm"], container["permanent"], container["Mprime"]
    counters = container.setdefault("_counters", {})
    for k in ("reset_lpj_isnan", "reset_lpj_smaller_eps_lpj", "reset_lpj_isinf"):
        counters[k] = 0
    sum_s, sum_ss = np.zeros(H), np.zeros((H, H))
    sum_sz, sum_szsz = np.zeros(H), np.zeros((H, H))
    Wp = np.zeros((D, H))
    sum_s_sz = np.zeros((H, H))
    sum_szsz_outer = np.zeros((H, H))
    trace_masked = 0.0
    n_uni = n_sub = 0
    if do_reconstruction:
        data["y_reconstructed"] = y_all.copy()
    for n in range(N):
        y, xi, states = y_all[n], x_infr_all[n], ss[n]
        if not permanent["background"] and S_perm > 0:
            lpj[n, :S_perm] = lpj_allzero(y, xi, theta["sigma2_inv"], counters)
        this_lpj, _ = lpj_states(states, theta, y, xi, counters)
        container["this_states"], container["this_lpj"] = states, this_lpj
        new_s, new_l = V.evolve_states(container, theta, lambda st: lpj_states(st, theta, y, xi, counters)[0],
                                       reference_quirks=reference_quirks)
        a, b = V.vary_Kn(this_lpj, new_l, lpj[n, S_perm:], states, new_s, H, S, S_perm, incl, Mprime)
        n_uni += a
        n_sub += b
        # ---- sufficient statistics over the post-selection K^n (sssc.py:553-611)
        pjc = np.exp(lpj[n] - lpj[n].max())
        _, terms = lpj_states(states, theta, y, xi)
        yo = y[xi]
        xs, xss = np.zeros(H), np.zeros((H, H))
        xsz, xszsz = np.zeros(H), np.zeros((H, H))
        for k, t in enumerate(terms):
            A, p = t["A"], pjc[k + S_perm]
            kappa = t["LW"] @ (yo - t["Wmu"]) + mus[A]
            xsz[A] += p * kappa
            xszsz[np.ix_(A, A)] += p * (t["Lam"] + np.outer(kappa, kappa))
        sf = states.astype(np.float64)
        xs += (pjc[S_perm:, None] * sf).sum(axis=0)
        xss += (pjc[S_perm:, None] * sf).T @ sf
        tot = pjc.sum() + EPS_PJC_SUM
        xs, xss, xsz, xszsz = xs / tot, xss / tot, xsz / tot, xszsz / tot
        sum_s += xs
        sum_ss += xss
        sum_sz += xsz
        sum_szsz += xszsz
        # ---- estimator (sssc.py:614-627): y_rec[~x] = W[~x] <sz>  (normalised without the tiny eps)
        y_rec = None
        if do_reconstruction:
            y_rec = data["y_reconstructed"][n]
            miss = ~data["x"][n]
            y_rec[miss] = W[miss, :] @ (xsz * tot / pjc.sum())
        if "W" in to_learn:
            Wp += np.outer(y_rec if incomplete else y, xsz)
        if "Psi" in to_learn:
            sum_s_sz += np.outer(xs, xsz)
        if "sigma2" in to_learn:
            if incomplete:
                wv = W[xi, :] @ xsz
                trace_masked += wv @ wv  # only the trace of the reference's D x D accumulator is used (:752)
            else:
                sum_szsz_outer += np.outer(xsz, xsz)
    container["_stats"] = dict(sum_s=sum_s, sum_ss=sum_ss, sum_sz=sum_sz, sum_szsz=sum_szsz, Wp=Wp,
                               sum_s_sz=sum_s_sz, sum_szsz_outer=sum_szsz_outer, trace_masked=trace_masked)
    # ---- parameter updates (sssc.py:688-770)
    if "W" in to_learn:
        # sssc.py:692-709: singular second moments -> pinv with rank-one noise -> else W + noise
        try:
            theta["W"] = Wp @ np.linalg.inv(sum_szsz)
        except np.linalg.LinAlgError:
            try:
                noise = np.random.normal(0, EPS_W, H)
                theta["W"] = Wp @ np.linalg.pinv(sum_szsz + np.outer(noise, noise))
            except np.linalg.LinAlgError:
                theta["W"] = W + EPS_W * np.random.normal(0, 1, [D, H])
    if "pies" in to_learn:
        p = sum_s / N
        p[p <= EPS_PIES] = EPS_PIES
        p[p >= 1 - EPS_PIES] = 1 - EPS_PIES
        if permanent["background"]:
            p[H - 1] = 1.0 - 1.1e-5
        theta["pies"] = p
    if "mus" in to_learn:
        theta["mus"] = sum_sz * 1.0 / (sum_s + EPS_MUS)
    if "Psi" in to_learn:
        m = theta["mus"]
        Psi = np.outer(m, m) * sum_ss + sum_szsz - 2 * m[:, None] * sum_s_sz
        # element-wise product with the inverse, as the reference does (sssc.py:738)
        theta["Psi"] = Psi * np.linalg.inv(sum_ss + EPS_PSI * np.eye(H))
    if "sigma2" in to_learn:
        if incomplete:
            yo2 = (y_all[x_infr_all] ** 2).sum()
            theta["sigma2"] = np.float64(((yo2 - trace_masked) + x_infr_all.sum() * sigma2) / N / D + EPS_SIGMA2)
        else:
            WtW = theta["W"].T @ theta["W"]
            theta["sigma2"] = np.float64(((y_all ** 2).sum() - np.trace(sum_szsz_outer @ WtW)) / N / D + EPS_SIGMA2)
    m = lpj.max(axis=1)
    Fs = (m + np.log(np.exp(lpj - m[:, None]).sum(axis=1))).sum()
    return theta["ljc"] + Fs / N, n_uni / N, n_sub / N, theta
