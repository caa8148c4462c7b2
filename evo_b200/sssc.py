"""Spike-and-slab sparse coding (ES3C) on the GPU -- drop-in for the reference's `SSSC`
(evo/models/sssc.py:17-813).

One EM iteration = `evo_bsc_prepare` (G = W'W, B = Y W; complete data) -> fused `evo_estep` with
the warp-per-state ES3C evaluator -> `evo_es3c_suffstats` (posterior slab moments, data estimator)
-> one all-reduce of the packed statistics -> replicated parameter updates (sssc.py:688-770).
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from .model import ACC_LEN, Model, _dev

_SCRATCH = {}


def _ensure_scratch(lib, device):
    """Global workspaces for states with 17..32 active latents (3168 doubles per warp)."""
    key = str(device)
    if key not in _SCRATCH:
        warps = lib.evo_device_sm_count() * 64
        buf = torch.empty(warps * 3168, dtype=torch.float64, device=device)
        _SCRATCH[key] = (buf, warps)
    buf, warps = _SCRATCH[key]
    _lib.check(lib.evo_es3c_set_scratch(_lib.ptr(buf), warps), "evo_es3c_set_scratch")


class SSSC(Model):
    def __init__(self, D, H, S, use_storage=True, precision=np.float64,
                 to_learn=("W", "pies", "mus", "sigma2", "Psi"), comm=None, device=None):
        Model.__init__(self, D, H, S, to_learn, comm, device)
        tol = 1e-5
        self.eps_lpj = np.finfo(np.float64).min
        self.eps_pjc_sum = np.finfo(np.float64).tiny
        self.eps_W = self.eps_pies = 5e-5
        self.eps_mus = np.finfo(np.float64).eps
        self.eps_Psi = self.eps_sigma2 = tol
        self.dtype_precision = precision
        self.use_storage = use_storage  # the reference's per-state cache; nothing to cache here
        # statistics pass: states with posterior weight below this are not expanded (include/evo_b200.h, q_skip)
        self.q_skip = 1e-30
        self.noise_policy = {
            "W": (-np.inf, +np.inf, False, None),
            "pies": (tol, 1.0 - tol, False, None),
            "mus": (-np.inf, +np.inf, False, None),
            "Psi": (-np.inf, +np.inf, False, tol),
            "sigma2": (tol, +np.inf, False, None),
        }
        with torch.cuda.device(self.device):
            _ensure_scratch(self.lib, self.device)

    # ---- generative model (sssc.py:66-102), cold path ---------------------------------------------
    def generate_from_hidden(self, model_params, my_hdata):
        th = {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in model_params.items()}
        W, mus, Psi = np.asarray(th["W"]), np.asarray(th["mus"]), np.asarray(th["Psi"])
        s = np.asarray(my_hdata["s"], dtype=bool)
        N, D = s.shape[0], W.shape[0]
        z = np.zeros((N, W.shape[1]))
        y_mean = np.zeros((N, D))
        for n in range(N):
            if s[n].any():
                z[n, s[n]] = np.random.multivariate_normal(mus[s[n]], Psi[np.ix_(s[n], s[n])], 1).flatten()
                y_mean[n] = W[:, s[n]] @ z[n, s[n]]
        y = y_mean + np.sqrt(float(th["sigma2"])) * np.random.randn(N, D)
        return {"y": y, "s": s, "z": z, "y_mean": y_mean}

    # ---- standard_init (sssc.py:104-197) -------------------------------------------------------------
    def standard_init(self, my_data, W_init=None, pi_init=None, sigma_init=None):
        dd = self.device_data(my_data)
        D, H = self.D, self.H
        th = {}
        th["pies"] = self.comm.bcast(np.random.uniform(low=0.1, high=0.5, size=[H]))
        th["mus"] = self.comm.bcast(np.random.normal(0, 1, [H])) if "mus" in self.to_learn else np.ones(H)
        th["Psi"] = np.diag(np.ones(H))
        ssum, ssq, cnt = self._column_moments(dd)
        y_mean = ssum / dd.N_total  # the reference divides the masked sums by N in both cases (sssc.py:150)
        if sigma_init is None:
            r = float((ssq - 2.0 * y_mean * ssum + cnt * y_mean * y_mean).sum().item())  # sum_obs (y - y_mean_d)^2
            if dd.incomplete:
                th["sigma2"] = np.float64(r / dd.n_obs_total + 0.001)
            else:  # mean of the diagonal of the unbiased covariance (np.cov), sssc.py:172
                th["sigma2"] = np.float64(r / D / max(dd.N_total - 1, 1) + 0.001)
        else:
            th["sigma2"] = np.float64(sigma_init)
        if not isinstance(W_init, np.ndarray):
            if W_init == "normal":
                th["W"] = self._noisy_init(None, 5.0).cpu().numpy()
            else:
                th["W"] = self._noisy_init(y_mean, float(np.sqrt(th["sigma2"])) / 4.0).cpu().numpy()
        else:
            th["W"] = W_init
        return th

    def check_params(self, model_params):
        model_params = Model.check_params(self, model_params)
        s2 = model_params["sigma2"]
        s2 = float(s2.item()) if isinstance(s2, (torch.Tensor, np.ndarray)) else float(s2)
        assert np.isfinite(s2) and s2 > 0
        model_params["sigma2"] = np.float64(s2)
        return model_params

    # ---- state-independent terms (sssc.py:329-366) -----------------------------------------------------
    def E_step_precompute(self, model_params, my_suff_stat, my_data):
        dd = self.device_data(my_data)
        D = self.D
        pies = _dev(model_params["pies"], self.device, torch.float64)
        sigma2 = np.float64(model_params["sigma2"])
        log1m = float(torch.log(1.0 - pies).sum().item())
        model_params["piH"] = float(pies.sum().item())
        pil = torch.log(pies / (1.0 - pies))
        model_params["pil_bar"] = pil if isinstance(model_params["pies"], torch.Tensor) else pil.cpu().numpy()
        s128 = sigma2.astype("longdouble")
        model_params["sigma2_inv"] = np.float64(1.0 / s128)
        if dd.incomplete:
            model_params["ljc"] = float(log1m + (-np.log(2 * np.pi) - np.log(sigma2)) * dd.n_obs_total / dd.N_total / 2)
        else:
            model_params["ljc"] = float(log1m - D / 2 * np.log(2 * np.pi) - 0.5 * np.float64(D * np.log(s128)))
        for k in ("reset_lpj_isnan", "reset_lpj_smaller_eps_lpj", "reset_lpj_isinf", "Psi_s_pinv"):
            dict.__setitem__(my_suff_stat, k, 0)
        return pil

    def _tables(self, model_params, dd, pil=None):
        c = dd.cache
        H = self.H
        c["W"] = _dev(model_params["W"], self.device, torch.float64)
        c["Wt"] = c["W"].T.contiguous()
        c["Psi"] = _dev(model_params["Psi"], self.device, torch.float64)
        c["mus"] = _dev(model_params["mus"], self.device, torch.float64)
        c["pil"] = _dev(pil if pil is not None else model_params["pil_bar"], self.device, torch.float64)
        # incomplete data use the same tables (Y is 0 where missing, so B = Y W sums over the observed entries) plus a
        # per-datapoint correction of G over the MISSING observables inside the evaluator (es3c.cuh, masked == 2)
        if "G" not in c:
            c["G"] = torch.empty((H, H), dtype=torch.float64, device=self.device)
            c["B"] = torch.empty((dd.N, H), dtype=torch.float64, device=self.device)
        yn = None
        if "ynorm2" not in c:
            c["ynorm2"] = torch.empty((dd.N,), dtype=torch.float64, device=self.device)
            yn = c["ynorm2"]
        self._call("evo_bsc_prepare", _lib.ptr(c["W"]), _lib.ptr(dd.Y), dd.N, self.D, H, _lib.ptr(c["G"]),
                   _lib.ptr(c["B"]), _lib.ptr(yn))

    def _fill_eval(self, a, model_params, dd):
        c = dd.cache
        a.eval = _lib.EVAL_ES3C
        a.sparseness = float(model_params["piH"])
        a.sigma2 = float(model_params["sigma2"])
        a.pil_bar_h, a.mus, a.Psi = c["pil"].data_ptr(), c["mus"].data_ptr(), c["Psi"].data_ptr()
        a.Y = dd.Y.data_ptr()
        if dd.incomplete:
            a.W, a.mask = c["Wt"].data_ptr(), dd.mask.data_ptr()
        a.G, a.B, a.ynorm2 = c["G"].data_ptr(), c["B"].data_ptr(), c["ynorm2"].data_ptr()

    # ---- E-step: fused kernel over the shard (sssc.py:510-548) -------------------------------------------
    def E_step(self, model_params, my_suff_stat, my_data):
        dd = self.device_data(my_data)
        pil = self.E_step_precompute(model_params, my_suff_stat, my_data)
        self._tables(model_params, dd, pil)
        acc = torch.zeros(ACC_LEN, dtype=torch.float64, device=self.device)
        self._estep_launch(my_suff_stat, dd, acc, lambda a: self._fill_eval(a, model_params, dd))
        dd.cache["acc"] = acc
        return acc

    def _finish_acc(self, acc, model_params, my_suff_stat, dd):
        r = acc.cpu().numpy() if isinstance(acc, torch.Tensor) else acc
        for i, k in enumerate(("reset_lpj_isnan", "reset_lpj_smaller_eps_lpj", "reset_lpj_isinf")):
            dict.__setitem__(my_suff_stat, k, int(r[3 + i]))
        self.last_counts = {"lpj_evaluated": float(r[6])}
        return float(model_params["ljc"] + r[0] / dd.N_total), float(r[1] / dd.N_total), float(r[2] / dd.N_total)

    # ---- one EM iteration (sssc.py:408-813) -----------------------------------------------------------------
    def step(self, model_params, my_suff_stat, my_data, do_reconstruction=False):
        model_params = self.check_params(model_params)
        return self.EM_step(model_params, my_suff_stat, my_data, do_reconstruction)

    def EM_step(self, model_params, my_suff_stat, my_data, do_reconstruction=False):
        dd = self.device_data(my_data)
        D, H = self.D, self.H
        ss = my_suff_stat
        acc = self.E_step(model_params, ss, my_data)
        c = dd.cache
        H2 = H * H
        stats = torch.zeros(2 * H + 4 * H2 + H * D + 1 + ACC_LEN, dtype=torch.float64, device=self.device)
        if do_reconstruction and dd.Yrec is None:
            dd.Yrec = torch.empty_like(dd.Y)
        if dd.incomplete and "W" in self.to_learn:
            assert do_reconstruction, "incomplete data: the W update needs the data estimator (sssc.py:630-632)"
        a = _lib.Es3cStatsArgs()
        a.N, a.D, a.H, a.S, a.S_perm = dd.N, D, H, int(ss["ss_dev"].shape[1]), int(ss["S_perm"])
        a.learn_psi = int("Psi" in self.to_learn)
        a.states, a.lpj, a.Y = ss["ss_dev"].data_ptr(), ss["lpj_dev"].data_ptr(), dd.Y.data_ptr()
        a.mask = dd.mask.data_ptr() if dd.incomplete else None
        a.recon_keep = dd.keep.data_ptr() if dd.keep is not None else None
        a.G, a.B, a.ynorm2 = c["G"].data_ptr(), c["B"].data_ptr(), c["ynorm2"].data_ptr()
        a.Wt, a.Psi, a.mus = c["Wt"].data_ptr(), c["Psi"].data_ptr(), c["mus"].data_ptr()
        a.sigma2 = float(model_params["sigma2"])
        a.stats = stats.data_ptr()
        a.Yrec = dd.Yrec.data_ptr() if do_reconstruction else None
        a.q_skip = float(self.q_skip)
        with torch.cuda.device(self.device):
            ev = self._tick()
            _lib.check(self.lib.evo_es3c_suffstats(C.byref(a), _lib.stream()), "evo_es3c_suffstats")
            self._tock("evo_es3c_suffstats", ev)
        if do_reconstruction:
            my_data["y_reconstructed_dev"] = dd.Yrec
            if not isinstance(my_data["y"], torch.Tensor):
                my_data["y_reconstructed"] = dd.Yrec.double().cpu().numpy()
        # ---- ONE all-reduce: statistics + free-energy / diagnostic scalars in the tail ------------------
        stats[-ACC_LEN:] = acc
        ev = self._tick()
        self.comm.allreduce_(stats)
        self._tock("allreduce", ev)
        # the parameter updates are enqueued before anything is read back: ONE device->host read per EM iteration
        ev = self._tick()
        pend = self._update_enqueue(model_params, ss, dd, stats)
        self._tock("solve", ev)
        host = torch.cat([stats[-ACC_LEN:], pend["scal"]]).cpu().numpy()
        F, S_nunique, S_sub = self._finish_acc(host[:ACC_LEN], model_params, ss, dd)
        self.last_counts["stats_bytes"] = stats.numel() * 8
        self._update_finish(model_params, ss, dd, stats, pend, host[ACC_LEN:])
        return F, S_nunique, S_sub, model_params

    # ---- parameter updates (sssc.py:688-770), replicated on every rank ----------------------------------------
    def _update_enqueue(self, th, ss, dd, stats):
        """`evo_es3c_update`: every learned parameter from the reduced statistics, on the device, no host read."""
        D, H = self.D, self.H
        c = dd.cache
        ws = getattr(self, "_update_ws", None)
        if ws is None:
            nb = int(self.lib.evo_es3c_update_workspace_bytes(D, H))
            ws = self._update_ws = torch.empty(nb, dtype=torch.uint8, device=self.device)
        if "sigma2" in self.to_learn and "ysq" not in c:
            ysq = (dd.Y.double() ** 2).sum().reshape(1)  # once per dataset; Y is 0 where masked
            self.comm.allreduce_(ysq)
            c["ysq"] = float(ysq.item())
        new = {"W": c["W"].clone(), "pies": _dev(th["pies"], self.device, torch.float64).clone(), "mus": c["mus"].clone(),
               "Psi": c["Psi"].clone(), "scal": torch.empty(4, dtype=torch.float64, device=self.device)}
        a = _lib.Es3cUpdateArgs()
        a.D, a.H = D, H
        L = self.to_learn
        a.learn_W, a.learn_pies, a.learn_mus = int("W" in L), int("pies" in L), int("mus" in L)
        a.learn_Psi, a.learn_sigma2 = int("Psi" in L), int("sigma2" in L)
        a.background, a.incomplete = int(bool(ss["permanent"]["background"])), int(dd.incomplete)
        a.N_total, a.n_obs_total = float(dd.N_total), float(dd.n_obs_total)
        a.ysq, a.sigma2_old = float(c.get("ysq", 0.0)), float(th["sigma2"])
        a.eps_pies, a.eps_mus, a.eps_Psi, a.eps_sigma2 = self.eps_pies, self.eps_mus, self.eps_Psi, self.eps_sigma2
        a.stats = stats.data_ptr()
        a.W, a.pies, a.mus, a.Psi = new["W"].data_ptr(), new["pies"].data_ptr(), new["mus"].data_ptr(), new["Psi"].data_ptr()
        a.scal, a.workspace = new["scal"].data_ptr(), ws.data_ptr()
        with torch.cuda.device(self.device):
            _lib.check(self.lib.evo_es3c_update(C.byref(a), _lib.stream()), "evo_es3c_update")
        return new

    def _update_finish(self, th, ss, dd, stats, new, scal):
        D, H = self.D, self.H
        H2 = H * H
        as_numpy = not isinstance(th["W"], torch.Tensor)
        out = (lambda t: t.cpu().numpy()) if as_numpy else (lambda t: t)
        singular_W = "W" in self.to_learn and (int(scal[0]) != 0 or scal[3] < 0.5)
        singular_P = "Psi" in self.to_learn and (int(scal[1]) != 0 or scal[3] < 0.5)
        if singular_W or singular_P:
            # (numerically) singular second moments: the reference's route (sssc.py:695-709) -- pseudo-inverse of the
            # moments plus rank-one noise.  The noise is drawn on rank 0 and broadcast, so that every rank keeps
            # bit-identical parameters (there is no per-iteration Bcast of theta here).  Rare path, device library calls.
            sum_s, sum_sz = stats[:H], stats[H:2 * H]
            sum_ss = stats[2 * H:2 * H + H2].view(H, H)
            sum_szsz = stats[2 * H + H2:2 * H + 2 * H2].view(H, H)
            sum_s_sz = stats[2 * H + 2 * H2:2 * H + 3 * H2].view(H, H)
            sum_outer = stats[2 * H + 3 * H2:2 * H + 4 * H2].view(H, H)
            WpT = stats[2 * H + 4 * H2:2 * H + 4 * H2 + H * D].view(H, D)
            if singular_W:
                noise = torch.as_tensor(self.comm.bcast(np.random.normal(0, self.eps_W, H)), device=self.device)
                new["W"] = (WpT.T @ torch.linalg.pinv(sum_szsz + torch.outer(noise, noise))).contiguous()
                if "sigma2" in self.to_learn and not dd.incomplete:
                    WtW = new["W"].T @ new["W"]
                    scal = scal.copy()
                    scal[2] = (dd.cache["ysq"] - float((sum_outer * WtW.T).sum().item())) / dd.N_total / D + self.eps_sigma2
            if singular_P:
                mu = new["mus"]
                Psi = torch.outer(mu, mu) * sum_ss + sum_szsz - 2 * mu[:, None] * sum_s_sz
                eye = torch.eye(H, dtype=torch.float64, device=self.device)
                new["Psi"] = Psi * torch.linalg.inv(sum_ss + self.eps_Psi * eye)
            self.last_counts["solve_fallback"] = self.last_counts.get("solve_fallback", 0) + 1
        for k in ("W", "pies", "mus", "Psi"):
            if k in self.to_learn:
                th[k] = out(new[k])
        if "sigma2" in self.to_learn:
            th["sigma2"] = np.float64(scal[2])
        return th

    def M_step(self, model_params, my_suff_stat, my_data):
        raise NotImplementedError("SSSC fuses E- and M-step (EM_step), like the reference (sssc.py:408-417)")

    # ---- stand-alone log-pseudo-joints (sssc.py:225-326) ------------------------------------------------------
    def lpj_all(self, model_params, my_data, states_dev, shared=False):
        """lpj of `states_dev` ([N][S'][Hw] or one shared [S'][Hw] table) for every datapoint."""
        dd = self.device_data(my_data)
        if "sigma2_inv" not in model_params or "pil_bar" not in model_params:
            self.E_step_precompute(model_params, {}, my_data)
        self._tables(model_params, dd)
        c = dd.cache
        Sp = states_dev.shape[-2]
        out = torch.empty((dd.N, Sp), dtype=torch.float64, device=self.device)
        self._call("evo_es3c_lpj", _lib.ptr(states_dev), int(shared), dd.N, Sp, self.D, self.H, _lib.ptr(c["Wt"]),
                   _lib.ptr(dd.Y), _lib.ptr(dd.mask), _lib.ptr(c.get("G")), _lib.ptr(c.get("B")),
                   _lib.ptr(c.get("ynorm2")), _lib.ptr(c["Psi"]), _lib.ptr(c["mus"]), _lib.ptr(c["pil"]),
                   float(model_params["sigma2"]), _lib.ptr(out), Sp, None)
        return out

    def log_pseudo_joint(self, model_params, my_suff_stat, my_data):
        """Per-datapoint signature of the reference: states from my_suff_stat["this_states"] (S',H) bool,
        datapoint from my_data["this_y"], my_data["this_x_infr"]."""
        from .variational import pack_states

        states = np.asarray(my_suff_stat["this_states"], dtype=bool)
        y = np.asarray(my_data["this_y"], dtype=np.float64)[None]
        x = np.asarray(my_data["this_x_infr"], dtype=bool)[None].copy()
        one = {"y": np.where(x, y, 0.0), "x_infr": x}
        sub = SSSC(self.D, self.H, states.shape[0], to_learn=self.to_learn, comm=self.comm, device=self.device)
        th = dict(model_params)
        sub.E_step_precompute(th, {}, one)
        out = sub.lpj_all(th, one, pack_states(states[None], self.device))
        return out[0].cpu().numpy()

    def log_pseudo_joint_permanent_states(self, model_params, my_suff_stat, my_data):
        y = np.asarray(my_data["this_y"], dtype=np.float64)
        x = np.asarray(my_data["this_x_infr"], dtype=bool)
        s2i = float(np.float64(1.0) / np.float64(model_params["sigma2"]))
        return np.array([-0.5 * float((y[x].astype(np.float32).astype(np.float64) ** 2).sum()) * s2i])

    def free_energy(self, my_data, model_params, my_suff_stat, full=True, compute_lpj=True):
        """Free energy over the current K sets or exact log-likelihood over all 2^H states (H < 12)
        (_models.py:334-451)."""
        from .variational import pack_states

        dd = self.device_data(my_data)
        self.E_step_precompute(model_params, my_suff_stat, my_data)
        if full:
            # all 2^H states enumerated on the device from their index (H <= 24; the reference stops at H < 12)
            self._tables(model_params, dd)
            c = dd.cache
            Fs = torch.zeros(1, dtype=torch.float64, device=self.device)
            self._call("evo_es3c_full_logsumexp", dd.N, self.D, self.H, int(bool(my_suff_stat["permanent"]["background"])),
                       _lib.ptr(c["Wt"]), _lib.ptr(dd.Y), _lib.ptr(dd.mask), _lib.ptr(c.get("G")), _lib.ptr(c.get("B")),
                       _lib.ptr(c.get("ynorm2")), _lib.ptr(c["Psi"]), _lib.ptr(c["mus"]), _lib.ptr(c["pil"]),
                       float(model_params["sigma2"]), _lib.ptr(Fs))
            self.comm.allreduce_(Fs)
            return float(model_params["ljc"] + Fs.item() / dd.N_total)
        elif compute_lpj:
            lpj = self.lpj_all(model_params, my_data, my_suff_stat["ss_dev"])
            if my_suff_stat["S_perm"] > 0:
                z = torch.zeros((1, my_suff_stat["ss_dev"].shape[-1]), dtype=torch.int32, device=self.device)
                lpj = torch.cat([self.lpj_all(model_params, my_data, z, shared=True), lpj], dim=1).contiguous()
        else:
            lpj = my_suff_stat["lpj_dev"]
        Fs = self._sum_logsumexp(lpj)
        self.comm.allreduce_(Fs)
        return float(model_params["ljc"] + Fs.item() / dd.N_total)

    def reconstruct(self, my_data, my_suff_stat, model_params):
        raise NotImplementedError("SSSC estimates data inside EM_step(do_reconstruction=True), like the reference")
