"""Binary Sparse Coding on the GPU -- drop-in for the reference's `BSC` (evo/models/bsc.py:14-287).

E-step: `evo_bsc_prepare` (G = W'W, B = Y W) then the fused `evo_estep` kernel; complete data uses
the sparse quadratic form, incomplete data the masked residual.  M-step: `evo_bsc_suffstats`, one
all-reduce of the packed statistics, replicated solve.
"""
import numpy as np
import torch

from . import _lib
from .model import ACC_LEN, Model, _dev


class BSC(Model):
    def __init__(self, D, H, S, to_learn=("W", "pi", "sigma"), comm=None, device=None):
        Model.__init__(self, D, H, S, to_learn, comm, device)
        self.eps_lpj = np.finfo(np.float64).min  # bsc.py:23-24
        # table B = Y W of the Gram form: "f64" (SIMT GEMM, bit-reproducible parity path) or "tc"
        # (tcgen05 split-precision GEMM, lpj within 1e-5 of the f64 reference)
        self.tables = "f64"
        self.stats = "f64"  # "tc": Wp = <S>' Y by the split-precision tensor-core GEMM; "auto": tc

    # ---- generative model (bsc.py:27-57), cold path ---------------------------------------------
    def generate_from_hidden(self, model_params, my_hdata):
        W = np.asarray(model_params["W"].cpu() if isinstance(model_params["W"], torch.Tensor) else model_params["W"])
        sigma = float(model_params["sigma"])
        s = np.asarray(my_hdata["s"])
        my_N = s.shape[0]
        D = W.shape[0]
        y_mean = s.astype(np.float64) @ W.T
        y = y_mean + np.random.normal(scale=sigma, size=(my_N, D))
        return {"y": y, "s": s, "y_mean": y_mean}

    # ---- standard_init (_models.py:206-283): data mean + noise, pi = 1/H, sigma = data std -----
    def standard_init(self, my_data, W_init=None, pi_init=None, sigma_init=None):
        dd = self.device_data(my_data)
        D, H = self.D, self.H
        ssum, ssq, cnt = self._column_moments(dd)
        if not dd.incomplete:
            y_mean = ssum / dd.N_total
            var = ssq / dd.N_total - y_mean * y_mean
            sig = float(torch.sqrt(var.clamp_min(0).sum() / D).item())
        else:
            y_mean = ssum / cnt.clamp_min(1.0)
            r = (ssq - 2.0 * y_mean * ssum + cnt * y_mean * y_mean).sum()  # sum over observed entries of (y - mean_d)^2
            sig = float(torch.sqrt(r.clamp_min(0) / dd.n_obs_total).item())
        if sigma_init is None:
            sigma_init = sig
            assert sigma_init > 0.0
        if not isinstance(W_init, np.ndarray):
            if W_init == "random_uniform":
                W_init = self.comm.bcast(np.random.random((D, H)))
            elif W_init == "normal":
                W_init = self._noisy_init(None, 5.0).cpu().numpy()
            elif W_init == "data_mean":
                W_init = np.tile(y_mean.cpu().numpy()[:, None], (1, H))
            else:
                W_init = self._noisy_init(y_mean, sigma_init / 4.0).cpu().numpy()
        if pi_init is None:
            pi_init = 1.0 / H
        return {"W": W_init, "pi": pi_init, "sigma": sigma_init}

    # ---- state-independent terms (bsc.py:100-125) ---------------------------------------------------
    def E_step_precompute(self, model_params, my_suff_stat, my_data):
        dd = self.device_data(my_data)
        D, H = self.D, self.H
        pi, sigma = float(model_params["pi"]), float(model_params["sigma"])
        model_params["piH"] = pi * H
        model_params["pre1"] = -1.0 / 2.0 / sigma / sigma
        model_params["pil_bar"] = float(np.log(pi / (1.0 - pi)))
        if dd.incomplete:
            model_params["ljc"] = float(H * np.log(1.0 - pi)
                                        - np.log(2 * np.pi * sigma * sigma) * dd.n_obs_total / dd.N_total / 2)
        else:
            model_params["ljc"] = float(H * np.log(1.0 - pi) - D / 2 * np.log(2 * np.pi * sigma * sigma))
        for k in ("reset_lpj_isnan", "reset_lpj_smaller_eps_lpj", "reset_lpj_isinf"):
            dict.__setitem__(my_suff_stat, k, 0)

    def _tables(self, model_params, dd):
        """W on the device; for complete data also G = W'W and B = Y W (|y|^2 once per dataset)."""
        Wd = _dev(model_params["W"], self.device, torch.float64)
        c = dd.cache
        c["W"] = Wd
        if dd.incomplete:
            return
        H = self.H
        if "G" not in c:
            c["G"] = torch.empty((H, H), dtype=torch.float64, device=self.device)
            c["B"] = torch.empty((dd.N, H), dtype=torch.float64, device=self.device)
        yn = None
        fresh = "ynorm2" not in c  # first use of this shard (or its values were replaced)
        if fresh:
            c["ynorm2"] = torch.empty((dd.N,), dtype=torch.float64, device=self.device)
            yn = c["ynorm2"]
        if self.tables == "tc":
            # B = Y W on the tensor cores (fp16 hi+lo split, <= 2e-6 of max|B|); G and |y|^2 stay f64
            if "tc_ws" not in c:
                nb = int(self.lib.evo_bsc_tables_tc_workspace_bytes(dd.N, self.D, H))
                c["tc_ws"] = torch.empty(nb, dtype=torch.uint8, device=self.device)
                fresh = True
            self._call("evo_bsc_prepare", _lib.ptr(Wd), _lib.ptr(dd.Y), dd.N, self.D, H, _lib.ptr(c["G"]), None,
                       _lib.ptr(yn))
            if fresh:
                self._call("evo_bsc_tables_tc_set_data", _lib.ptr(dd.Y), dd.N, self.D, H, _lib.ptr(c["tc_ws"]))
            self._call("evo_bsc_tables_tc", _lib.ptr(Wd), dd.N, self.D, H, _lib.ptr(c["tc_ws"]), _lib.ptr(c["B"]))
            return
        self._call("evo_bsc_prepare", _lib.ptr(Wd), _lib.ptr(dd.Y), dd.N, self.D, H, _lib.ptr(c["G"]), _lib.ptr(c["B"]),
                   _lib.ptr(yn))

    def _fill_eval(self, a, model_params, dd):
        c = dd.cache
        a.pre1, a.pil_bar = float(model_params["pre1"]), float(model_params["pil_bar"])
        a.sparseness = float(model_params["piH"])
        a.W, a.Y = c["W"].data_ptr(), dd.Y.data_ptr()
        if dd.incomplete:
            a.eval = _lib.EVAL_BSC_DIRECT
            a.mask = dd.mask.data_ptr()
        else:
            a.eval = _lib.EVAL_BSC_GRAM
            a.G, a.B, a.ynorm2 = c["G"].data_ptr(), c["B"].data_ptr(), c["ynorm2"].data_ptr()
        if self._theta_dev is not None:  # CUDA-graph replay (emgraph.py): the scalars above come from device memory
            a.theta_dev = self._theta_dev.data_ptr()
        return None

    _theta_dev = None

    # ---- E-step (_models.py:454-565) -----------------------------------------------------------------
    def _estep_enqueue(self, model_params, my_suff_stat, my_data, fused=None, precompute=True):
        dd = self.device_data(my_data)
        if precompute:
            self.E_step_precompute(model_params, my_suff_stat, my_data)
        self._tables(model_params, dd)
        acc = torch.zeros(ACC_LEN, dtype=torch.float64, device=self.device)

        def fill(a):
            self._fill_eval(a, model_params, dd)
            if fused is not None:  # statistics in the tail of the E-step kernel (evo_estep_args.fs_*)
                a.fs_stats, a.fs_wp_workspace, a.fs_rest = fused["stats"].data_ptr(), fused["ws"].data_ptr(), fused["rest"].data_ptr()

        self._estep_launch(my_suff_stat, dd, acc, fill)
        dd.cache["acc"] = acc
        return dd, acc

    # fuse the statistics pass into the E-step kernel when Wp goes through the tensor-core GEMM and the data are complete
    fuse_stats = True

    def _fused_begin(self, dd):
        """Statistics buffer, Wp workspace (Y planes valid, <s> planes cleared) and the rest list for a fused E-step."""
        D, H = self.D, self.H
        c = dd.cache
        stats = torch.zeros(H * D + H * H + H + 1 + ACC_LEN, dtype=torch.float64, device=self.device)
        if "wp_tc_ws" not in c:
            nb = int(self.lib.evo_bsc_wp_tc_workspace_bytes(dd.N, D, H))
            c["wp_tc_ws"] = torch.empty(nb, dtype=torch.uint8, device=self.device)
            c.pop("wp_tc_valid", None)
        if not c.get("wp_tc_valid"):
            self._call("evo_bsc_wp_tc_set_data", _lib.ptr(dd.Y), dd.N, D, H, _lib.ptr(c["wp_tc_ws"]))
            c["wp_tc_valid"] = True
        rest = c.get("fs_rest")
        if rest is None or rest.numel() < dd.N + 1:
            rest = c["fs_rest"] = torch.empty(dd.N + 1, dtype=torch.int32, device=self.device)
        self._call("evo_bsc_wp_tc_begin", dd.N, D, H, _lib.ptr(c["wp_tc_ws"]))
        return {"stats": stats, "ws": c["wp_tc_ws"], "rest": rest}

    def _fused_finish(self, model_params, my_suff_stat, dd, fused):
        ss = my_suff_stat
        self._call("evo_bsc_suffstats_tc_finish", _lib.ptr(ss["ss_dev"]), _lib.ptr(ss["lpj_dev"]), _lib.ptr(dd.Y), dd.N,
                   self.D, self.H, int(ss["ss_dev"].shape[1]), int(ss["S_perm"]), float(model_params["pre1"]),
                   float(model_params["pil_bar"]), _lib.ptr(fused["stats"]), _lib.ptr(fused["ws"]), _lib.ptr(fused["rest"]),
                   _lib.ptr(self._theta_dev))
        return fused["stats"]

    def E_step(self, model_params, my_suff_stat, my_data):
        dd, acc = self._estep_enqueue(model_params, my_suff_stat, my_data)
        red = acc.clone()
        self.comm.allreduce_(red)
        return self._estep_finish(red, model_params, my_suff_stat, dd)

    def _estep_finish(self, red, model_params, my_suff_stat, dd):
        r = red.cpu().numpy() if isinstance(red, torch.Tensor) else red
        for i, k in enumerate(("reset_lpj_isnan", "reset_lpj_smaller_eps_lpj", "reset_lpj_isinf")):
            dict.__setitem__(my_suff_stat, k, int(r[3 + i]))
        self.last_counts = {"lpj_evaluated": float(r[6])}
        F = model_params["ljc"] + r[0] / dd.N_total
        return float(F), float(r[1] / dd.N_total), float(r[2] / dd.N_total)

    # ---- stand-alone lpj with the reference's per-datapoint signature (bsc.py:60-97) -----------------
    def log_pseudo_joint(self, model_params, my_suff_stat, my_data):
        states = np.asarray(my_suff_stat["this_states"], dtype=bool)
        y = np.asarray(my_data["this_y"], dtype=np.float64)[None]
        x = np.asarray(my_data["this_x_infr"], dtype=bool)[None]
        from .variational import pack_states

        st = pack_states(states[None], self.device)
        Wd = _dev(model_params["W"], self.device, torch.float64)
        Yd = _dev(np.where(x, y, 0.0), self.device, torch.float32)
        md = _dev(x, self.device, torch.uint8)
        out = torch.empty((1, states.shape[0]), dtype=torch.float64, device=self.device)
        cnt = torch.zeros(3, dtype=torch.int64, device=self.device)
        self._call("evo_bsc_lpj", _lib.EVAL_BSC_DIRECT, _lib.ptr(st), 0, 1, states.shape[0], self.D, self.H, _lib.ptr(Wd),
                   _lib.ptr(Yd), _lib.ptr(md), None, None, None, float(model_params["pre1"]),
                   float(model_params["pil_bar"]), _lib.ptr(out), states.shape[0], _lib.ptr(cnt))
        c = cnt.cpu().numpy()
        for i, k in enumerate(("reset_lpj_isnan", "reset_lpj_smaller_eps_lpj", "reset_lpj_isinf")):
            if k in my_suff_stat:
                dict.__setitem__(my_suff_stat, k, my_suff_stat[k] + int(c[i]))
        return out[0].cpu().numpy()

    def log_pseudo_joint_permanent_states(self, model_params, my_suff_stat, my_data):
        y = np.asarray(my_data["this_y"], dtype=np.float64)
        x = np.asarray(my_data["this_x_infr"], dtype=bool)
        saved = my_suff_stat["this_states"] if "this_states" in my_suff_stat else None
        dict.__setitem__(my_suff_stat, "this_states", np.zeros((1, self.H), dtype=bool))
        out = self.log_pseudo_joint(model_params, my_suff_stat, {"this_y": y, "this_x_infr": x})
        if saved is not None:
            dict.__setitem__(my_suff_stat, "this_states", saved)
        return out

    def lpj_all(self, model_params, my_data, states_dev, shared=False):
        """lpj of `states_dev` ([N][S'][Hw], or one [S'][Hw] table for every datapoint) under
        `model_params`; batched form of log_pseudo_joint used by free_energy."""
        dd = self.device_data(my_data)
        self._tables(model_params, dd)
        c = dd.cache
        Sp = states_dev.shape[-2]
        out = torch.empty((dd.N, Sp), dtype=torch.float64, device=self.device)
        ev = _lib.EVAL_BSC_DIRECT if dd.incomplete else _lib.EVAL_BSC_GRAM
        self._call("evo_bsc_lpj", ev, _lib.ptr(states_dev), int(shared), dd.N, Sp, self.D, self.H, _lib.ptr(c["W"]),
                   _lib.ptr(dd.Y), _lib.ptr(dd.mask), _lib.ptr(c.get("G")), _lib.ptr(c.get("B")),
                   _lib.ptr(c.get("ynorm2")), float(model_params["pre1"]), float(model_params["pil_bar"]),
                   _lib.ptr(out), Sp, None)
        return out

    def dense_workspace(self, model_params):
        """fp16 hi/lo planes of W for the tensor-core log-joint (`evo_bsc_lpj_gemm_prepare`)."""
        Wd = _dev(model_params["W"], self.device, torch.float64)
        nbytes = int(self.lib.evo_bsc_lpj_gemm_workspace_bytes(self.D, self.H))
        ws = getattr(self, "_dense_ws", None)
        if ws is None or ws.numel() != nbytes:
            ws = self._dense_ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        self._call("evo_bsc_lpj_gemm_prepare", _lib.ptr(Wd), self.D, self.H, _lib.ptr(ws))
        return ws

    def lpj_all_dense(self, model_params, my_data, states_dev, shared=False, workspace=None):
        """`lpj_all` through the fused tcgen05 GEMM S W' with the squared-residual epilogue
        (`evo_bsc_lpj_gemm`; fp16 hi+lo split of W, fp32 accumulation: <= 1e-5 relative)."""
        dd = self.device_data(my_data)
        ws = workspace if workspace is not None else self.dense_workspace(model_params)
        Sp = states_dev.shape[-2]
        out = torch.empty((dd.N, Sp), dtype=torch.float64, device=self.device)
        self._call("evo_bsc_lpj_gemm", _lib.ptr(states_dev), int(shared), dd.N, Sp, self.D, self.H, _lib.ptr(ws),
                   _lib.ptr(dd.Y), _lib.ptr(dd.mask), float(model_params["pre1"]), float(model_params["pil_bar"]),
                   _lib.ptr(out), Sp, None)
        return out

    def free_energy(self, my_data, model_params, my_suff_stat, full=True, compute_lpj=True):
        """Free energy over the current K sets, or the exact log-likelihood over all 2^H states
        (`full=True`, H < 12) (_models.py:334-451)."""
        from .variational import pack_states

        dd = self.device_data(my_data)
        self.E_step_precompute(model_params, my_suff_stat, my_data)
        if full:
            # all 2^H states enumerated on the device from their index (the reference tabulates them and stops at
            # H < 12, utils.py:58-69); H <= 24 here
            self._tables(model_params, dd)
            c = dd.cache
            Fs = torch.zeros(1, dtype=torch.float64, device=self.device)
            ev = _lib.EVAL_BSC_DIRECT if dd.incomplete else _lib.EVAL_BSC_GRAM
            self._call("evo_bsc_full_logsumexp", ev, dd.N, self.D, self.H, int(bool(my_suff_stat["permanent"]["background"])),
                       _lib.ptr(c["W"]), _lib.ptr(dd.Y), _lib.ptr(dd.mask), _lib.ptr(c.get("G")), _lib.ptr(c.get("B")),
                       _lib.ptr(c.get("ynorm2")), float(model_params["pre1"]), float(model_params["pil_bar"]), _lib.ptr(Fs))
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

    # ---- estimator (_models.py:614-665, bsc.py:279-287) --------------------------------------------------
    def reconstruct(self, my_data, my_suff_stat, model_params):
        dd = self.device_data(my_data)
        Wd = _dev(model_params["W"], self.device, torch.float64)
        if dd.Yrec is None:
            dd.Yrec = torch.empty_like(dd.Y)
        ss = my_suff_stat
        self._call("evo_bsc_reconstruct", _lib.ptr(ss["ss_dev"]), _lib.ptr(ss["lpj_dev"]), _lib.ptr(Wd), _lib.ptr(dd.Y),
                   _lib.ptr(dd.keep), _lib.ptr(dd.mask), dd.N, self.D, self.H, int(ss["ss_dev"].shape[1]),
                   int(ss["S_perm"]), _lib.ptr(dd.Yrec))
        my_data["y_reconstructed_dev"] = dd.Yrec
        if not isinstance(my_data["y"], torch.Tensor):
            my_data["y_reconstructed"] = dd.Yrec.double().cpu().numpy()

    # ---- M-step (bsc.py:128-277) ------------------------------------------------------------------------
    def step(self, model_params, my_suff_stat, my_data, do_reconstruction=False):
        """Model.step (_models.py:162-203) with the E-step scalars riding in the tail of the statistics
        buffer: ONE all-reduce and one host read per EM iteration."""
        model_params = self.check_params(model_params)
        if len(self.to_learn) == 0:
            return Model.step(self, model_params, my_suff_stat, my_data, do_reconstruction)
        dd0 = self.device_data(my_data)
        fused = None
        if self.fuse_stats and self.stats in ("tc", "auto") and not dd0.incomplete:
            fused = self._fused_begin(dd0)
        dd, acc = self._estep_enqueue(model_params, my_suff_stat, my_data, fused)
        if do_reconstruction:
            self.reconstruct(my_data, my_suff_stat, model_params)
        if fused is not None:
            stats = self._fused_finish(model_params, my_suff_stat, dd, fused)
        else:
            stats = self._stats_enqueue(model_params, my_suff_stat, dd)
        stats[-ACC_LEN:] = acc
        ev = self._tick()
        self.comm.allreduce_(stats)
        self._tock("allreduce", ev)
        # the solve is enqueued before anything is read back: ONE device->host read per EM iteration
        ev = self._tick()
        pend = self._solve_enqueue(model_params, my_suff_stat, dd, stats)
        self._tock("solve", ev)
        host = torch.cat([stats[-ACC_LEN:], pend["scal"]]).cpu().numpy()
        F, S_nunique, S_sub = self._estep_finish(host[:ACC_LEN], model_params, my_suff_stat, dd)
        self.last_counts["stats_bytes"] = stats.numel() * 8
        return F, S_nunique, S_sub, self._solve_finish(model_params, my_suff_stat, dd, pend, host[ACC_LEN:])

    def _stats_enqueue(self, model_params, my_suff_stat, dd):
        D, H = self.D, self.H
        ss = my_suff_stat
        if dd.incomplete:
            assert dd.Yrec is not None, "missing data: call reconstruct() before the M-step (bsc.py:186)"
            Ym = dd.Yrec
        else:
            Ym = dd.Y
        stats = torch.zeros(H * D + H * H + H + 1 + ACC_LEN, dtype=torch.float64, device=self.device)
        S, S_perm = int(ss["ss_dev"].shape[1]), int(ss["S_perm"])
        mode = self.stats
        if mode == "auto":
            # measured on B200: the f64 path costs ~ (active latents per datapoint) x D atomics
            # and swings 20x with the posterior's spread (cfg5, N = 1M: 69 ms); the GEMM is bounded by
            # N H D (22 ms there, 1.0 vs 2.0 ms at cfg2)
            mode = "tc"
        if mode == "tc":
            # Wp = <S>' Y on the tensor cores (fp16 hi+lo planes, <= 1e-5 of max|Wp|); Wq, pies, sigma stay f64
            c = dd.cache
            if "wp_tc_ws" not in c:
                nb = int(self.lib.evo_bsc_wp_tc_workspace_bytes(dd.N, D, H))
                c["wp_tc_ws"] = torch.empty(nb, dtype=torch.uint8, device=self.device)
                c.pop("wp_tc_valid", None)
            if dd.incomplete or not c.get("wp_tc_valid"):  # reconstructed data change with every iteration
                self._call("evo_bsc_wp_tc_set_data", _lib.ptr(Ym), dd.N, D, H, _lib.ptr(c["wp_tc_ws"]))
                c["wp_tc_valid"] = not dd.incomplete
            self._call("evo_bsc_suffstats_tc", _lib.ptr(ss["ss_dev"]), _lib.ptr(ss["lpj_dev"]), _lib.ptr(Ym), dd.N, D, H,
                       S, S_perm, float(model_params["pre1"]), float(model_params["pil_bar"]), _lib.ptr(stats), None,
                       _lib.ptr(c["wp_tc_ws"]))
            return stats
        ws = dd.cache.get("stats_ws")
        if ws is None:
            nb = int(self.lib.evo_bsc_suffstats_workspace_bytes(dd.N, D, H))
            ws = dd.cache["stats_ws"] = torch.empty(nb, dtype=torch.uint8, device=self.device) if nb > 0 else False
        self._call("evo_bsc_suffstats", _lib.ptr(ss["ss_dev"]), _lib.ptr(ss["lpj_dev"]), _lib.ptr(Ym), dd.N, D, H,
                   S, S_perm, int(bool(ss["permanent"]["allzero"])),
                   float(model_params["pre1"]), float(model_params["pil_bar"]), _lib.ptr(stats), None,
                   _lib.ptr(ws) if ws is not False else None, int(ws.numel()) if ws is not False else 0)
        return stats

    def em_graph(self, model_params, my_suff_stat, my_data, max_iterations=4096):
        """Whole EM iterations as one replayed CUDA graph: `g = model.em_graph(theta, ss, data); F, S_nunique, S_sub =
        g.run(n); theta = g.theta_out()` equals n calls of `step` (emgraph.py)."""
        from .emgraph import EMGraph

        return EMGraph(self, model_params, my_suff_stat, my_data, max_iterations)

    def M_step(self, model_params, my_suff_stat, my_data):
        dd = self.device_data(my_data)
        stats = self._stats_enqueue(model_params, my_suff_stat, dd)
        self.comm.allreduce_(stats)
        return self._solve_update(model_params, my_suff_stat, dd, stats)

    def _solve_update(self, model_params, my_suff_stat, dd, stats):
        pend = self._solve_enqueue(model_params, my_suff_stat, dd, stats)
        return self._solve_finish(model_params, my_suff_stat, dd, pend, pend["scal"].cpu().numpy())

    def _solve_enqueue(self, model_params, my_suff_stat, dd, stats, out=None):
        """Parameter update from the reduced statistics (bsc.py:226-275) by `evo_bsc_solve` (block Gaussian
        elimination of [Wq | Wp] on the device), enqueued without a host read; `scal` = (info, pi, sigma, W finite) is
        read once by the caller together with the E-step scalars."""
        D, H = self.D, self.H
        ws = getattr(self, "_solve_ws", None)
        if ws is None:
            nb = int(self.lib.evo_bsc_solve_workspace_bytes(D, H))
            ws = self._solve_ws = torch.empty(nb, dtype=torch.uint8, device=self.device)
        learn_W = "W" in self.to_learn
        if out is not None:  # caller-owned static buffers (emgraph.py); W is not an input of the solve: in place is safe
            W_out, pies, scal = out
            W_out = W_out if learn_W else None
        else:
            W_out = torch.empty((D, H), dtype=torch.float64, device=self.device) if learn_W else None
            pies = torch.empty(H, dtype=torch.float64, device=self.device)
            scal = torch.empty(4, dtype=torch.float64, device=self.device)
        self._call("evo_bsc_solve", _lib.ptr(stats), float(dd.N_total), D, H, int(learn_W),
                   int(bool(my_suff_stat["permanent"]["background"])), float(dd.n_obs_total) if dd.incomplete else -1.0,
                   float(model_params["sigma"]), _lib.ptr(W_out), _lib.ptr(pies), _lib.ptr(scal), _lib.ptr(ws))
        return {"W": W_out, "p": pies, "scal": scal, "stats": stats}

    def _solve_finish(self, model_params, my_suff_stat, dd, pend, scal):
        D, H = self.D, self.H
        as_numpy = not isinstance(model_params["W"], torch.Tensor)
        if "W" in self.to_learn:
            if int(scal[0]) == 0 and scal[3] > 0.5:
                Wn = pend["W"]
            else:
                # Wq is (numerically) singular -- e.g. a latent that carries no posterior mass.  The reference's
                # lstsq(rcond=-1) returns the minimum-norm solution (bsc.py:233-235): eigenvalues below eps x the
                # largest are dropped.  Same thing on the device through the Hermitian pseudo-inverse (rare path).
                stats = pend["stats"]
                Wp = stats[: H * D].view(H, D)
                Wq = stats[H * D: H * D + H * H].view(H, H)
                Wq = torch.tril(Wq) + torch.tril(Wq, -1).T
                Wn = (torch.linalg.pinv(Wq, hermitian=True, rtol=float(np.finfo(np.float64).eps)) @ Wp).T.contiguous()
                self.last_counts["solve_fallback"] = self.last_counts.get("solve_fallback", 0) + 1
            model_params["W"] = Wn.cpu().numpy() if as_numpy else Wn
        if "pi" in self.to_learn:
            model_params["pi"] = float(scal[1])
            model_params["pies"] = pend["p"].cpu().numpy() if as_numpy else pend["p"]
        if "sigma" in self.to_learn:
            sig = float(scal[2])
            model_params["sigma"] = sig if sig == sig else 0.0  # a negative statistic (NaN root) is clamped by check_params
        return model_params
