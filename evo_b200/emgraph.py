"""Whole EM iterations of BSC as ONE CUDA graph.

`BSC.step` (bsc.py) needs the host once per iteration: it clamps the new parameters (`check_params`,
_models.py:102-159), derives the E-step scalars (`E_step_precompute`, bsc.py:100-125) and passes them to the kernels as
arguments.  Here those few scalar operations run on the device (`evo_bsc_em_advance`), the kernels read them from
device memory (`evo_estep_args.theta_dev`), and the launch sequence of one iteration -- tables, fused E-step +
statistics, Wp GEMM, all-reduce, parameter solve, advance -- is captured once with stream capture
(`torch.cuda.CUDAGraph` is only the capture / replay handle) and replayed without a host round trip.  Per-iteration
values (free energy, counters, solve status) are logged on the device and read back after the last replay.

Same kernels, same Philox streams, same order of operations as `step`: K sets bit-identical, F and parameters equal
up to the last-ulp difference of `log` on the device and in NumPy (tests/test_emgraph_gpu.py).  Complete data,
W / pi / sigma learned or not; singular statistics (solve info != 0), which `step` handles by a minimum-norm solve
on the host's decision, end the run with an error that names the iteration.
"""
import numpy as np
import torch

from . import _lib
from .model import ACC_LEN

REC = 16  # doubles per logged iteration: acc[8], scal[4], ljc, pi, sigma, iteration


class EMGraph:
    def __init__(self, model, model_params, my_suff_stat, my_data, max_iterations=4096):
        m = self.model = model
        self.ss, self.data = my_suff_stat, my_data
        dd = self.dd = m.device_data(my_data)
        if dd.incomplete:
            raise ValueError("EMGraph: complete data only (use BSC.step for missing values)")
        if m.timing is not None:
            raise ValueError("EMGraph: per-call event timing (model.timing) cannot be captured")
        if my_suff_stat.get("inject") is not None or my_suff_stat.get("trace"):
            raise ValueError("EMGraph: RNG mode only")
        if not (m.fuse_stats and m.stats in ("tc", "auto")):
            raise ValueError("EMGraph: needs the fused statistics path (model.stats 'tc' / 'auto', fuse_stats)")
        dev = m.device
        self.as_numpy = not isinstance(model_params["W"], torch.Tensor)
        th = m.check_params(dict(model_params))
        m.E_step_precompute(th, my_suff_stat, my_data)
        # static parameter buffers: the graph reads and writes these addresses
        self.W = torch.as_tensor(np.asarray(th["W"].cpu() if isinstance(th["W"], torch.Tensor) else th["W"]),
                                 dtype=torch.float64).to(dev).contiguous().clone()
        self.pies = torch.zeros(m.H, dtype=torch.float64, device=dev)
        self.scal = torch.zeros(4, dtype=torch.float64, device=dev)
        it0 = int(my_suff_stat.get("iteration", 0)) + 1
        self.theta = torch.tensor([th["pre1"], th["pil_bar"], th["piH"], float(it0), float(th["pi"]), float(th["sigma"]),
                                   th["ljc"], 0.0], dtype=torch.float64, device=dev)
        self.max_rec = int(max_iterations)
        self.rec = torch.zeros((self.max_rec, REC), dtype=torch.float64, device=dev)
        self.params = th
        self.graph = None
        self.done = 0  # iterations run through this object

    # ---- one iteration, every scalar from device memory -------------------------------------------------------
    def _body(self):
        m, ss, dd = self.model, self.ss, self.dd
        lo_pi, hi_pi = m.noise_policy["pi"][:2]
        lo_sg, hi_sg = m.noise_policy["sigma"][:2]
        m._theta_dev = self.theta
        try:
            th = self.params
            th["W"] = self.W
            fused = m._fused_begin(dd)
            dd, acc = m._estep_enqueue(th, ss, self.data, fused, precompute=False)
            stats = m._fused_finish(th, ss, dd, fused)
            stats[-ACC_LEN:] = acc
            m.comm.allreduce_(stats)
            m._solve_enqueue(th, ss, dd, stats, out=(self.W, self.pies, self.scal))
            m._call("evo_bsc_em_advance", _lib.ptr(stats[-ACC_LEN:]), _lib.ptr(self.scal), _lib.ptr(self.theta),
                    _lib.ptr(self.rec), self.max_rec, m.D, m.H, int("pi" in m.to_learn), int("sigma" in m.to_learn),
                    float(lo_pi), float(min(hi_pi, 1e300)), float(lo_sg), float(min(hi_sg, 1e300)))
        finally:
            m._theta_dev = None

    def _capture(self):
        dev = self.model.device
        g = torch.cuda.CUDAGraph()
        with torch.cuda.device(dev):
            torch.cuda.synchronize(dev)
            it = self.ss.get("iteration", 0)
            with torch.cuda.graph(g):
                self._body()
            self.ss["iteration"] = it  # capturing ran the host bookkeeping of one iteration, not the iteration
        self.graph = g

    def run(self, iterations):
        """Run `iterations` EM iterations; returns `(F, S_nunique, S_sub)` lists (one entry per iteration).  The first
        iteration ever run through this object is launched eagerly (it allocates and fills the per-dataset caches);
        every later one is a graph replay."""
        n = int(iterations)
        if n <= 0:
            return [], [], []
        if self.done + n > self.max_rec:
            raise ValueError("EMGraph: more than max_iterations = %d iterations" % self.max_rec)
        first = self.done
        todo = n
        with torch.cuda.device(self.model.device):
            if self.done == 0:
                self._body()
                todo -= 1
            if todo > 0 and self.graph is None:
                self._capture()
            for _ in range(todo):
                self.graph.replay()
            self.ss["iteration"] = int(self.ss.get("iteration", 0)) + todo  # the eager iteration counted itself
        self.done += n
        rec = self.rec[first:first + n].cpu().numpy()  # the ONE host read of the run
        N_total = self.dd.N_total
        bad = np.nonzero((rec[:, 8] != 0) | (rec[:, 11] < 0.5))[0]
        if "W" in self.model.to_learn and bad.size:
            raise RuntimeError("EMGraph: singular or non-finite statistics in iteration %d of this run (solve info %d); "
                               "use BSC.step, which takes the reference's minimum-norm route" % (bad[0] + 1, int(rec[bad[0], 8])))
        F = (rec[:, 12] + rec[:, 0] / N_total).tolist()
        for i, k in enumerate(("reset_lpj_isnan", "reset_lpj_smaller_eps_lpj", "reset_lpj_isinf")):
            dict.__setitem__(self.ss, k, int(rec[-1, 3 + i]))
        self.model.last_counts = {"lpj_evaluated": float(rec[-1, 6])}
        self.last, self.records = rec[-1], rec
        return F, (rec[:, 1] / N_total).tolist(), (rec[:, 2] / N_total).tolist()

    def theta_out(self):
        """The parameters after the last iteration, as `BSC.step` would return them (NumPy or device tensors,
        whichever the caller passed in)."""
        th = dict(self.params)
        W, pies = self.W.clone(), self.pies.clone()
        th["W"] = W.cpu().numpy() if self.as_numpy else W
        if self.done > 0:
            if "pi" in self.model.to_learn:
                th["pi"] = float(self.last[9])
                th["pies"] = pies.cpu().numpy() if self.as_numpy else pies
            if "sigma" in self.model.to_learn:
                sig = float(self.last[10])
                th["sigma"] = sig if sig == sig else 0.0
        return th
