"""`Model` -- EM skeleton with the reference's interface (evo/models/_models.py:17-665), driving the
CUDA kernels of libevo_b200.so.  Subclasses: `BSC` (bsc.py), `SSSC` (sssc.py).

Per EM iteration and rank: a few kernel launches on the current stream, ONE all-reduce of the packed
sufficient statistics (+ free-energy / diagnostic scalars in its tail), one small replicated solve.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from . import comm as _comm
from . import variational as V

ACC_LEN = 8  # [sum logsumexp, #new unique, #substituted, reset nan, reset small, reset inf, #lpj evaluated, spare]


def _dev(x, device, dtype):
    """numpy / torch (any device) -> contiguous tensor on `device`."""
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=dtype).contiguous()
    return torch.as_tensor(np.ascontiguousarray(x)).to(device=device, dtype=dtype).contiguous()


class DeviceData:
    """Device-resident shard of the dataset (`my_data` of the reference):
    Y float32 [N][D] (0 where missing), inference mask / reconstruction mask as uint8 or None."""

    def __init__(self, my_data, device, comm):
        y = my_data["y"]
        self.N, self.D = int(y.shape[0]), int(y.shape[1])
        x_infr = my_data.get("x_infr") if hasattr(my_data, "get") else my_data["x_infr"]
        if x_infr is None:  # complete data
            xi, self.incomplete = None, False
        else:
            xi = _dev(x_infr, device, torch.uint8)
            self.incomplete = not bool(xi.all().item())
        self.mask = xi if self.incomplete else None
        Y = _dev(y, device, torch.float32)
        if self.incomplete:
            Y = torch.where(xi.bool(), Y, torch.zeros_like(Y))  # NaN placeholders never reach a kernel
        self.Y = Y.contiguous()
        x = my_data["x"] if "x" in my_data else None
        self.keep = _dev(x, device, torch.uint8) if x is not None else None
        self.n_obs = int(xi.sum().item()) if self.incomplete else self.N * self.D
        # constants reduced once (the reference re-reduces them 3-4x per iteration)
        tot = comm.allreduce(np.array([float(self.N), float(self.n_obs)]))
        tot = np.atleast_1d(tot)
        self.N_total, self.n_obs_total = int(round(tot[0])), int(round(tot[1]))
        self.Yrec = None
        self.cache = {}
        self.y_ref = None


class Model:
    """Abstract EM model.  Same constructor and methods as the reference's `Model`
    (_models.py:17-56); `comm` is an `evo_b200.comm.Comm` (or anything with its surface)."""

    def __init__(self, D, H, S, to_learn=("W", "pi", "sigma"), comm=None, device=None):
        self.comm = comm if comm is not None else _comm.world()
        self.to_learn = list(to_learn)
        self.D, self.H, self.S = int(D), int(H), int(S)
        tol = 1e-5
        self.noise_policy = {
            "W": (-np.inf, +np.inf, False, None),
            "pi": (tol, 1.0 - tol, False, None),
            "sigma": (tol, +np.inf, False, None),
        }
        self.B_max = 0.0
        self.B_max_shft = np.inf
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.lib = _lib.load()  # raises if the CUDA library has not been built
        self._data = {}
        self.last_counts = {}

    # ---- plumbing ---------------------------------------------------------------------------
    def device_data(self, my_data):
        """Resident device shard of `my_data` (uploaded once).  Keyed by the identity of `my_data["y"]`; the entry
        holds a reference to that object, so its id cannot be recycled by another array while it is cached."""
        y = my_data["y"]
        dd = self._data.get(id(y))
        if dd is None or dd.y_ref is not y or dd.N != y.shape[0]:
            dd = DeviceData(my_data, self.device, self.comm)
            dd.y_ref = y
            self._data = {id(y): dd}  # one resident dataset per model
        return dd

    def rebind_data(self, my_data, my_data_new):
        """The resident shard of `my_data` now answers to `my_data_new["y"]` (same rows arriving from another buffer,
        e.g. pinned host memory): nothing is copied."""
        dd = self.device_data(my_data)
        dd.y_ref = my_data_new["y"]
        self._data = {id(dd.y_ref): dd}
        return dd

    @staticmethod
    def _values_changed(dd):
        """The shard's VALUES were replaced: drop everything derived from them (|y|^2, the transposed fp16 planes of
        the tensor-core statistics, ES3C's sum of y^2); buffers that only depend on the shape stay."""
        for k in ("ynorm2", "wp_tc_valid", "ysq"):
            dd.cache.pop(k, None)

    def refresh_data(self, my_data):
        """Copy `my_data["y"]` (host buffer, ideally pinned) into the resident device shard again --
        the host->device leg of a step whose batch arrives from the host -- and drop what was
        derived from the old values."""
        dd = self.device_data(my_data)
        y = my_data["y"]
        src = y if isinstance(y, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(y))
        if src.dtype == dd.Y.dtype:
            dd.Y.copy_(src, non_blocking=True)
        else:
            dd.Y.copy_(src.to(dd.Y.dtype))
        if dd.incomplete:
            dd.Y.copy_(torch.where(dd.mask.bool(), dd.Y, torch.zeros_like(dd.Y)))
        self._values_changed(dd)
        return dd

    def stage_data(self, my_data):
        """Start the host->device copy of `my_data["y"]` (pinned host memory) into the SPARE device
        buffer on a side stream and return at once, so the copy of the next batch runs under the
        kernels of the current step.  `adopt_staged` makes it the resident shard."""
        dd = self.device_data(my_data)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.device)
        if getattr(dd, "Y_spare", None) is None:
            dd.Y_spare = torch.empty_like(dd.Y)
            dd.spare_free = None
        y = my_data["y"]
        src = y if isinstance(y, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(y))
        cs = self._copy_stream
        if dd.spare_free is not None:
            cs.wait_event(dd.spare_free)  # kernels of the step that last read the spare buffer are done
        with torch.cuda.stream(cs):
            if src.dtype == dd.Y_spare.dtype:
                dd.Y_spare.copy_(src, non_blocking=True)
            else:
                dd.Y_spare.copy_(src.to(dd.Y_spare.dtype))
            if dd.incomplete:
                dd.Y_spare.copy_(torch.where(dd.mask.bool(), dd.Y_spare, torch.zeros_like(dd.Y_spare)))
            dd.staged = torch.cuda.Event()
            dd.staged.record(cs)
        return dd

    def adopt_staged(self, my_data):
        """The staged buffer becomes the resident shard: the compute stream waits for its copy, the
        previous shard becomes the spare (free once everything enqueued so far has run)."""
        dd = self.device_data(my_data)
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(dd.staged)
        dd.spare_free = torch.cuda.Event()
        dd.spare_free.record(cur)
        dd.Y, dd.Y_spare = dd.Y_spare, dd.Y
        self._values_changed(dd)
        return dd

    def _call(self, name, *args):
        with torch.cuda.device(self.device):
            ev = self._tick()
            _lib.check(getattr(self.lib, name)(*args, _lib.stream()), name)
            self._tock(name, ev)

    # optional per-call CUDA-event timing on the launching stream (bench.py sets `timing = []`)
    timing = None

    def _tick(self):
        if self.timing is None:
            return None
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def _tock(self, name, e0):
        if e0 is None:
            return
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record()
        self.timing.append((name, e0, e1))

    def timing_summary(self):
        """{call: (count, mean ms)} of the recorded calls (synchronises)."""
        torch.cuda.synchronize(self.device)
        out = {}
        for name, e0, e1 in self.timing or []:
            c, t = out.get(name, (0, 0.0))
            out[name] = (c + 1, t + e0.elapsed_time(e1))
        return {k: (c, t / c) for k, (c, t) in out.items()}

    # ---- parameter sanity (_models.py:102-159) -------------------------------------------------
    def check_params(self, model_params):
        """Clamp to the noise policy.  Every rank holds bit-identical parameters (same reduced
        statistics, same deterministic solve), so the reference's per-iteration Bcast is dropped."""
        for param, (lo, hi, absify, lo_diag) in self.noise_policy.items():
            if param not in model_params:
                continue
            v = model_params[param]
            if np.isscalar(v) or (isinstance(v, np.ndarray) and v.ndim == 0):
                v = float(v)
                assert np.isfinite(v), "scalar %s not finite" % param
                if v < lo:
                    v = lo
                if v >= hi:
                    v = hi
                model_params[param] = v
            elif isinstance(v, torch.Tensor):
                if np.isfinite(lo) or np.isfinite(hi):
                    v.clamp_(min=lo if np.isfinite(lo) else None, max=hi if np.isfinite(hi) else None)
                if lo_diag is not None:
                    d = torch.diagonal(v)
                    d.clamp_(min=lo_diag)
            else:
                v = np.asarray(v)
                assert np.isfinite(v).all(), "%s not finite" % param
                if np.isfinite(lo):
                    v[v < lo] = lo
                if np.isfinite(hi):
                    v[v >= hi] = hi
                if lo_diag is not None:
                    idx = np.arange(v.shape[0])
                    dg = v[idx, idx]
                    dg[dg < lo_diag] = lo_diag
                    v[idx, idx] = dg
                model_params[param] = v
        return model_params

    # ---- EM step (_models.py:162-203) -----------------------------------------------------------
    def step(self, model_params, my_suff_stat, my_data, do_reconstruction=False):
        model_params = self.check_params(model_params)
        F, S_nunique, S_sub = self.E_step(model_params, my_suff_stat, my_data)
        if do_reconstruction:
            self.reconstruct(my_data, my_suff_stat, model_params)
        new_params = self.M_step(model_params, my_suff_stat, my_data) if len(self.to_learn) > 0 else model_params
        return F, S_nunique, S_sub, new_params

    # ---- fused E-step launcher --------------------------------------------------------------------
    def _fill_estep_args(self, a, my_suff_stat, dd):
        """Container fields -> `evo_estep_args` (sizes, EA hyper-parameters, RNG keys, K sets); validates the
        device buffers against the shard (the kernels index them unchecked)."""
        ss = my_suff_stat
        perm = ss["permanent"]
        sd, ld = ss["ss_dev"], ss["lpj_dev"]
        S, S_perm = int(sd.shape[1]), int(ss["S_perm"])
        assert sd.dtype == torch.int32 and tuple(sd.shape) == (dd.N, S, V.words_for(self.H)) and sd.is_contiguous(), \
            "ss_dev must be contiguous int32 [%d][S][%d], got %s" % (dd.N, V.words_for(self.H), tuple(sd.shape))
        assert ld.dtype == torch.float64 and tuple(ld.shape) == (dd.N, S + S_perm) and ld.is_contiguous(), \
            "lpj_dev must be contiguous float64 [%d][%d], got %s" % (dd.N, S + S_perm, tuple(ld.shape))
        assert sd.device == self.device and ld.device == self.device, "K sets live on another device than the model"
        a.N, a.n_offset = dd.N, int(ss.get("n_offset", 0) or 0)
        a.D, a.H, a.S, a.S_perm = self.D, self.H, S, S_perm
        a.Mprime = int(ss["Mprime"])
        a.n_parents, a.n_children, a.n_generations = int(ss["n_parents"]), int(ss["n_children"]), int(ss["n_generations"])
        a.selection = _lib.SEL[ss["parent_selection"].name]
        a.mutation = _lib.MUT[ss["mutation_algorithm"].name]
        a.background = int(bool(perm["background"]))
        a.bitflip_prob = float(ss["bitflip_prob"]) if ss["bitflip_prob"] is not None else 0.0
        a.seed, a.iteration = int(ss.get("seed", 0x5EED)), int(ss.get("iteration", 0)) + 1
        a.states, a.lpj = sd.data_ptr(), ld.data_ptr()

    def _estep_launch(self, my_suff_stat, dd, acc, fill):
        """Fill an `evo_estep_args`, let the subclass add its evaluator tables (`fill(args)`), launch.
        With `my_suff_stat["trace"]` set, the same E-step runs through `evo_evolve` and the selected parents,
        the children as mutated and their flip positions / cut points are left in `my_suff_stat["trace_out"]`."""
        ss = my_suff_stat
        trace = bool(ss.get("trace"))
        ea = _lib.EvolveArgs() if trace else None
        a = ea.e if trace else _lib.EstepArgs()
        self._fill_estep_args(a, ss, dd)
        inj = ss.get("inject")
        keep = None
        if inj is not None:
            ch, cnt = inj
            if not isinstance(ch, torch.Tensor) or ch.dtype != torch.int32:
                ch, cnt = V.pack_injection(ch, cnt, self.device)
            keep = (ch, cnt)
            assert ch.shape[0] == dd.N and ch.shape[1] == a.n_generations
            assert ch.shape[3] == V.words_for(self.H) and tuple(cnt.shape) == (dd.N, a.n_generations)
            a.inj_children, a.inj_count, a.inj_cmax = ch.data_ptr(), cnt.data_ptr(), int(ch.shape[2])
        a.acc = acc.data_ptr()
        ovf = getattr(self, "_ovf_list", None)
        if ovf is None or ovf.numel() < dd.N + 1:
            ovf = self._ovf_list = torch.empty(dd.N + 1, dtype=torch.int32, device=self.device)
        a.overflow_list = ovf.data_ptr()
        hold = fill(a)
        with torch.cuda.device(self.device):
            ev = self._tick()
            if trace:
                tr = V.alloc_trace(self.lib, ea, dd.N, self.device)
                _lib.check(self.lib.evo_evolve(C.byref(ea), _lib.stream()), "evo_evolve")
                dict.__setitem__(ss, "trace_out", tr)
            else:
                _lib.check(self.lib.evo_estep(C.byref(a), _lib.stream()), "evo_estep")
            self._tock("evo_estep", ev)
        ss["iteration"] = a.iteration
        return keep, hold

    def evolve(self, model_params, my_suff_stat, my_data):
        """`evolve_states` (evo/variational/eas.py:153-313) for EVERY datapoint of the shard in one launch
        (`evo_evolve`, evolve_only): the K sets are left as they are, the new unique states of all generations
        and their log-pseudo-joints come back.  Returns `(new_states int32 [N][Ctot][Hw] packed, new_lpj
        float64 [N][Ctot], counts int32 [N][1+G])` as device tensors, plus the operator trace in
        `my_suff_stat["trace_out"]`."""
        ss = my_suff_stat
        dd = self.device_data(my_data)
        self._evolve_tables(model_params, ss, my_data, dd)
        ea = _lib.EvolveArgs()
        self._fill_estep_args(ea.e, ss, dd)
        ea.e.acc = None
        hold = self._fill_eval(ea.e, model_params, dd)  # noqa: F841 (keeps evaluator tables alive)
        ea.evolve_only = 1
        with torch.cuda.device(self.device):
            tr = V.alloc_trace(self.lib, ea, dd.N, self.device, evolve=True)
            _lib.check(self.lib.evo_evolve(C.byref(ea), _lib.stream()), "evo_evolve")
        dict.__setitem__(ss, "trace_out", tr)
        ss["iteration"] = ea.e.iteration
        return tr["ev_states"], tr["ev_lpj"], tr["ev_count"]

    def _evolve_tables(self, model_params, my_suff_stat, my_data, dd):
        """State-independent terms + evaluator tables under `model_params` (what E_step does before its launch)."""
        pil = self.E_step_precompute(model_params, my_suff_stat, my_data)
        if pil is None:
            self._tables(model_params, dd)
        else:
            self._tables(model_params, dd, pil)

    # ---- free energy (_models.py:334-451) -------------------------------------------------------
    def _sum_logsumexp(self, lpj):
        out = torch.zeros(1, dtype=torch.float64, device=self.device)
        self._call("evo_free_energy", _lib.ptr(lpj), lpj.shape[0], lpj.shape[1], lpj.stride(0), _lib.ptr(out))
        return out

    # ---- standard_init helpers (_models.py:206-283, sssc.py:104-197): statistics and noise on the device ----------
    def _column_moments(self, dd):
        """(sum_n m y, sum_n m y^2, sum_n m) per observable over ALL ranks: one kernel, one all-reduce."""
        D = self.D
        tot = torch.zeros(3 * D, dtype=torch.float64, device=self.device)
        self._call("evo_column_moments", _lib.ptr(dd.Y), _lib.ptr(dd.mask), dd.N, D, _lib.ptr(tot))
        self.comm.allreduce_(tot)
        return tot[:D], tot[D:2 * D], tot[2 * D:]

    def _noisy_init(self, row_mean, scale):
        """[D][H] = row_mean[:, None] + scale N(0,1), filled on every rank from one Philox seed (taken from NumPy's
        global generator on rank 0, so `np.random.seed` keeps runs reproducible; the reference broadcasts the noise)."""
        seed = int(self.comm.bcast(int(np.random.randint(0, 2 ** 31 - 1))))
        W = torch.empty((self.D, self.H), dtype=torch.float64, device=self.device)
        self._call("evo_normal_fill", _lib.ptr(W), self.D, self.H, _lib.ptr(row_mean.contiguous()) if row_mean is not None else None,
                   float(scale), seed)
        return W

    # ---- data generation (_models.py:73-99) on the device ------------------------------------------------
    def generate_data(self, model_params, my_N, seed=None, n_offset=0, device_only=False):
        """`Model.generate_data(model_params, my_N)` of the reference (_models.py:73-99): draws the hidden states
        s ~ Bernoulli(pies) and the observations of the model (`generate_from_hidden`, bsc.py:27-57 /
        sssc.py:66-102) -- here by ONE kernel (`evo_generate_data`), row n a pure function of
        (seed, n_offset + n), so every rank can generate its own shard of one dataset.  `seed=None` takes a
        seed from NumPy's global generator (the reference draws from it directly).  Returns the reference's dict
        {"y", "s", "y_mean"[, "z"]} as NumPy arrays; with `device_only` {"y_dev" f32, "s_dev" packed} only."""
        assert "pies" in model_params or "pi" in model_params
        H, D = self.H, self.D
        dev = self.device
        W = _dev(model_params["W"], dev, torch.float64)
        assert tuple(W.shape) == (D, H)
        Wt = W.T.contiguous()
        pies = model_params["pies"] if "pies" in model_params else model_params["pi"]
        pies_d, pi = None, 0.0
        if np.ndim(pies) == 0:
            pi = float(pies)
        else:
            pies_d = _dev(pies, dev, torch.float64)
        mus = Psi = None
        diag = 1
        if "Psi" in model_params:  # ES3C
            mus = _dev(model_params["mus"], dev, torch.float64)
            Psi = _dev(model_params["Psi"], dev, torch.float64)
            diag = int(bool((Psi == torch.diag(torch.diagonal(Psi))).all().item()))
            sigma = float(np.sqrt(float(model_params["sigma2"])))
        else:
            sigma = float(model_params["sigma"])
        if seed is None:
            seed = int(np.random.randint(0, 2 ** 31 - 1))
        my_N = int(my_N)
        Y = torch.empty((my_N, D), dtype=torch.float32, device=dev)
        S = torch.empty((my_N, V.words_for(H)), dtype=torch.int32, device=dev)
        Z = torch.empty((my_N, H), dtype=torch.float32, device=dev) if (mus is not None and not device_only) else None
        Ym = torch.empty((my_N, D), dtype=torch.float32, device=dev) if not device_only else None
        ovf = torch.zeros(1, dtype=torch.int32, device=dev)
        self._call("evo_generate_data", _lib.ptr(Wt), _lib.ptr(pies_d), pi, _lib.ptr(mus), _lib.ptr(Psi), diag, sigma,
                   my_N, int(n_offset), D, H, int(seed), _lib.ptr(Y), _lib.ptr(S), _lib.ptr(Z), _lib.ptr(Ym), _lib.ptr(ovf))
        if device_only:
            return {"y_dev": Y, "s_dev": S}
        out = {"y": Y.double().cpu().numpy(), "s": V.unpack_states(S, H).cpu().numpy(), "y_mean": Ym.double().cpu().numpy(),
               "y_dev": Y, "s_dev": S}
        if Z is not None:
            out["z"] = Z.double().cpu().numpy()
        return out
