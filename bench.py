#!/usr/bin/env python
"""Benchmark of the evo hot path (truncated-variational E-step + M-step) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg5] [--impl ours|reference]

A "step" is one EM iteration (`model.step`: E-step over every datapoint of the shard, sufficient
statistics, all-reduce, parameter update) on synthetic data of the shape BASELINE.json names.
Metric: log-joints evaluated per second (whole job, all ranks); EM iterations/s rides along.
Default workload: cfg5 (the large BSC config BASELINE.json quotes the scaling metric on), N = 1 000 000
datapoints IN TOTAL, sharded over the ranks by the reference's array_split rule -- STRONG scaling.  The
dataset is one dataset whatever the rank count: row n is a pure function of (seed, n) (Philox streams,
`evo_generate_data` on the device, oracle/datagen.py for the CPU arm), generated from ONE dictionary.
The only collective is the all-reduce of the packed sufficient statistics.
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# synthetic inputs and EA hyper-parameters per config
WORKLOADS = {
    # name: model, D, H, N (total), K, (selection, mutation, parents, children, generations), active latents
    "cfg1": dict(model="bsc", D=25, H=10, N=1000, K=20, ea=("fit", "randflip", 10, 1, 1), act=2.0, bars=True),
    "cfg2": dict(model="bsc", D=256, H=300, N=100000, K=64, ea=("fit", "randflip", 8, 8, 2), act=5.0),
    "cfg5": dict(model="bsc", D=1024, H=1024, N=1000000, K=128, ea=("fit", "randflip", 16, 8, 2), act=8.0),
    "cfg3": dict(model="es3c", D=144, H=512, N=250000, K=64, ea=("fit", "randflip", 8, 8, 2), act=4.0),
    "cfg4": dict(model="es3c", D=144, H=512, N=250000, K=64, ea=("fit", "randflip", 8, 8, 2), act=4.0, missing=0.3),
}
DESCR = {
    "cfg1": "examples/bars-test BSC, 5x5 bars (D=25, H=10), N=1000, K=20",
    "cfg2": "BSC feature learning, synthetic 16x16 patches (D=256, H=300), N=100k, K=64",
    "cfg5": "large-scale BSC scaling sweep (D=1024, H=1024), N=1M, K=128",
    "cfg3": "ES3C image denoising, synthetic 12x12 noisy patches (D=144, H=512), N=250k, K=64",
    "cfg4": "ES3C image inpainting, 30% missing (D=144, H=512), N=250k, K=64",
}
METRIC, UNIT = "log_joints_per_sec", "log-joints/s"
DATA_SEED, DICT_SEED = 1000, 7


# ---------------------------------------------------------------------------------------------
# ONE dataset per workload: generating parameters from a fixed legacy RandomState (identical on every
# rank and in the CPU arm), rows from Philox streams keyed by the GLOBAL row index.
# ---------------------------------------------------------------------------------------------
def gen_params(w):
    """Generating parameters of SURVEY 8(d): W_gen ~ N(0,1) (bars for cfg1), s ~ Bernoulli(act/H),
    ES3C z ~ N(1, I), noise N(0,1)."""
    D, H = w["D"], w["H"]
    if w.get("bars"):
        R = H // 2
        Wg = np.zeros((R, R, H))
        for i in range(R):
            Wg[i, :, i] = 10.0
            Wg[:, i, R + i] = 10.0
        Wg = Wg.reshape(R * R, H)
    else:
        Wg = np.random.RandomState(DICT_SEED).normal(size=(D, H))
    th = {"W": Wg, "pies": np.full(H, w["act"] / H)}
    if w["model"] == "es3c":
        th.update({"mus": np.ones(H), "Psi": np.eye(H), "sigma2": np.float64(1.0)})
    else:
        th["sigma"] = 1.0
    return th


def _hash_u32(x):
    """lowbias32-style integer hash on values held in 64-bit lanes (same arithmetic in NumPy and torch)."""
    x = x & 0xFFFFFFFF
    x = ((x ^ (x >> 16)) * 0x7FEB352D) & 0xFFFFFFFF
    x = ((x ^ (x >> 15)) * 0x846CA68B) & 0xFFFFFFFF
    return x ^ (x >> 16)


def observed_mask(rows, D, p_missing, xp):
    """Entry (n, d) is missing with probability p: a hash of the GLOBAL entry index (sharding independent)."""
    cols = xp.arange(D) if xp is np else xp.arange(D, device=rows.device)
    idx = rows[:, None] * D + cols[None, :]
    return _hash_u32(idx + 0x9E3779B9) >= int(p_missing * 4294967296.0)


def host_rows(w, rows):
    """Rows `rows` (global indices) of the workload's dataset on the host (the CPU arm's sample)."""
    from oracle import datagen as DG

    g = gen_params(w)
    if w["model"] == "es3c":
        out = DG.generate(g["W"], g["pies"], 1.0, rows, DATA_SEED, mus=g["mus"], Psi=g["Psi"], psi_diag=True)
    else:
        out = DG.generate(g["W"], g["pies"], g["sigma"], rows, DATA_SEED)
    X = observed_mask(np.asarray(rows, dtype=np.int64), w["D"], w["missing"], np) if w.get("missing") else None
    return out["y"], X


def device_rows(model, w, lo, hi, device):
    """Rows [lo, hi) of the same dataset, generated on the device by `evo_generate_data`."""
    import torch

    g = gen_params(w)
    Y = model.generate_data(g, hi - lo, seed=DATA_SEED, n_offset=lo, device_only=True)["y_dev"]
    X = None
    if w.get("missing"):
        X = observed_mask(torch.arange(lo, hi, dtype=torch.int64, device=device), w["D"], w["missing"], torch)
    return Y, X


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, "/tmp/evo_clocks_%d.csv" % os.getpid()

    def start(self):
        try:
            self.f = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def mark(self):
        """Start of the timed region: samples written before this point are dropped."""
        try:
            self.f.flush()
            self.n0 = sum(1 for _ in open(self.path))
        except Exception:
            self.n0 = 0

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        lines = open(self.path).read().splitlines()
        n0 = getattr(self, "n0", 0)
        lines = lines[n0:] if len(lines) > n0 else lines[-1:]
        for line in lines:
            p = [x.strip() for x in line.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for nm, v in zip(names, p[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "samples": len(sm)}
        try:
            os.remove(self.path)
        except OSError:
            pass
        return out


# ---------------------------------------------------------------------------------------------
# algorithmic HBM bytes per call -- what `roofline.achieved` is made of
# ---------------------------------------------------------------------------------------------
def algorithmic_bytes(w, N, U_mean):
    D, H, K = w["D"], w["H"], w["K"]
    Hw = (H + 31) // 32
    return {
        # (a') + (c): read y_n, read every packed state evaluated, write every lpj
        "evo_estep": N * (4 * D + ((D + 7) // 8 if w.get("missing") else 0) + (K + U_mean) * (4 * Hw + 4)),
        # tables of the Gram form: read Y, write B = Y W (f64)
        "evo_bsc_prepare": N * (4 * D + 8 * H),
        "evo_bsc_tables_tc": N * (4 * D + 8 * H),
        # (d): read K^n + lpj + y_n, one pass over the outputs
        "evo_bsc_suffstats": N * (K * (4 * Hw + 4) + 4 * D) + 8 * (H * D + H * H),
        "evo_bsc_suffstats_tc": N * (K * (4 * Hw + 4) + 4 * D) + 8 * (H * D + H * H),
        "evo_bsc_suffstats_tc_finish": N * (K * (4 * Hw + 4) + 4 * D) + 8 * (H * D + H * H),
        # transposed fp16 hi + lo planes of Y: read f32, write 2 x fp16
        "evo_bsc_wp_tc_set_data": N * D * 8,
        # ES3C (d): as above + mask bits + y_rec write, one pass over the four H x H + H x D outputs
        "evo_es3c_suffstats": N * (K * (4 * Hw + 4) + 4 * D + (D + 7) // 8 + 4 * D) + 8 * (H * D + 4 * H * H),
    }


def run_patches(args):
    """the overlapping-patch kernels either side of the ES3C path in the denoising driver
    (12 x 12 patches, shift 1), each call alone between CUDA events, against the HBM roofline: extract
    writes N*D*4 bytes, merge reads them.  One JSON line; not the headline metric."""
    import torch

    from evo_b200.patches import OverlappingPatches, mean_merger, median_merger

    torch.cuda.set_device(0)
    side, ph = (args.n if args.n else 2048), 12
    g = torch.Generator(device="cuda")
    g.manual_seed(0)
    img = torch.rand((side, side), generator=g, device="cuda") * 255
    ovp = OverlappingPatches(img, ph, ph, patch_shift=1)
    nbytes = ovp.N * ovp.D * 4
    peak, src = measured_peaks()

    def timed(fn):
        for _ in range(max(args.warmup, 3)):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / args.steps

    def extract():
        ovp._patches = None
        ovp.get()

    res = {}
    for name, fn in (("extract", extract), ("merge_mean", lambda: ovp.merge(mean_merger)),
                     ("merge_median", lambda: ovp.merge(median_merger))):
        ms = timed(fn)
        res[name] = {"ms": ms, "GB/s": nbytes / ms * 1e-6, "frac": nbytes / ms * 1e-6 / peak}
    print(json.dumps({"metric": "patch_pipeline", "image": [side, side], "patch": [ph, ph], "shift": 1, "patches": ovp.N,
                      "D": ovp.D, "algorithmic_bytes": nbytes, "peak_GB/s": peak, "peak_source": src,
                      "l2": "patch matrix larger than L2" if nbytes > 126e6 else "fits L2", "kernels": res}), flush=True)


def run_lpj_compare(args):
    """north_star (a): the dense tcgen05 log-joint (evo_bsc_lpj_gemm) timed against the sparse
    Gram/popc kernel (evo_bsc_lpj) on the SAME evolved K sets of a BSC workload, each call alone
    between CUDA events on the launching stream.  One JSON line; no EM step is timed here."""
    import torch

    from evo_b200 import _lib
    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    torch.cuda.set_device(0)
    device = torch.device("cuda", 0)
    w = WORKLOADS[args.workload]
    assert w["model"] == "bsc", "lpj comparison is defined for the BSC workloads"
    N = args.n if args.n else w["N"]
    D, H, K = w["D"], w["H"], w["K"]
    Hw = (H + 31) // 32
    model = BSC(D, H, K, device=device)
    Y, X = device_rows(model, w, 0, N, device)
    data = {"y": Y, "x_infr": None}
    np.random.seed(1)
    theta = model.check_params(model.standard_init(data))
    sel, mut, P, C, G = w["ea"]
    cont = init_states(N, K, H, sel, mut, P, C, G, device=device, seed=0x5EED)
    for _ in range(max(args.warmup, 3)):  # evolve the K sets so their sparsity is the workload's
        model.step(theta, cont, data)
    model.E_step_precompute(theta, cont, data)
    sd = cont["ss_dev"]
    dd = model.device_data(data)
    model.timing = []
    for _ in range(args.steps + 1):
        ws = model.dense_workspace(theta)
        dense = model.lpj_all_dense(theta, data, sd, workspace=ws)
        sparse = model.lpj_all(theta, data, sd)
    t = model.timing_summary()
    model.timing = None
    rel = float(((dense - sparse).abs() / sparse.abs().clamp_min(1e-300)).max().item())
    bits = sd.view(torch.int32).cpu().numpy().view(np.uint32)
    act = float(np.unpackbits(bits[: min(N, 2000)].view(np.uint8)).sum() / (min(N, 2000) * K))
    j = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak_tf, peak_bw = float(j.get("bf16_tflops", 1590.0)), float(j.get("hbm_gbs", 6650.0))
    src = "measured (MEASURED_PEAKS.json, burst)" if j else "fallback (B200_PROFILING.md)"
    lpj = N * K
    d_ms, s_ms = t["evo_bsc_lpj_gemm"][1], t["evo_bsc_lpj"][1]
    dp_ms, sp_ms = t["evo_bsc_lpj_gemm_prepare"][1], t["evo_bsc_prepare"][1]
    flop = 2.0 * H * D * lpj
    sbytes = N * (4 * D + K * (4 * Hw + 4))
    line = {
        "mode": "lpj_compare", "config": {"workload": "%s: %s" % (args.workload, DESCR[args.workload]), "N": N},
        "log_joints": lpj, "active_latents_per_state": act, "max_rel_diff_dense_vs_sparse": rel,
        "dense": {"kernel": "evo_bsc_lpj_gemm (tcgen05, fp16 hi+lo = 2 tensor passes)", "ms": d_ms, "prepare_ms": dp_ms,
                  "lpj_per_sec": lpj / (d_ms * 1e-3),
                  "roofline": {"bound": "tensor", "achieved": flop / (d_ms * 1e-3) / 1e12, "peak": peak_tf, "unit": "TFLOP/s",
                               "frac": flop / (d_ms * 1e-3) / 1e12 / peak_tf, "executed_frac": 2 * flop / (d_ms * 1e-3) / 1e12 / peak_tf,
                               "algorithmic_flop_per_launch": flop, "peak_source": src}},
        "sparse": {"kernel": "evo_bsc_lpj (Gram / popc form, f64)", "ms": s_ms, "prepare_ms": sp_ms,
                   "lpj_per_sec": lpj / (s_ms * 1e-3),
                   "roofline": {"bound": "hbm", "achieved": sbytes / (s_ms * 1e-3) / 1e9, "peak": peak_bw, "unit": "GB/s",
                                "frac": sbytes / (s_ms * 1e-3) / 1e9 / peak_bw, "algorithmic_bytes_per_launch": sbytes,
                                "peak_source": src}},
        "faster": "sparse" if s_ms + sp_ms < d_ms + dp_ms else "dense",
    }
    print(json.dumps(line), flush=True)


# dram__bytes_read.sum + dram__bytes_write.sum per call from the `ncu --set full` capture of the same command
# (profiles/, named per round); None where no capture at that size exists.  Filled by tools/ncu_traffic.py.
NCU_TRAFFIC = {}
try:
    NCU_TRAFFIC = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
except Exception:
    pass


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload_config(name, w, N_total, world, args):
    return {"workload": "%s: %s" % (name, DESCR[name]), "N_total": N_total,
            "ea": "%s/%s parents=%d children=%d generations=%d" % tuple(w["ea"]),
            "dataset": "one synthetic dataset, row n = f(seed %d, n) from one dictionary (seed %d)" % (DATA_SEED, DICT_SEED),
            "l2": "inputs larger than L2 (K sets + tables > 126 MB per GPU)" if (N_total // world) * w["K"] * 4 * ((w["H"] + 31) // 32) > 126e6
            else "working set fits L2 (small reference config)",
            "tables": ("B = Y W by tcgen05 fp16 hi+lo split GEMM (lpj within 1e-5 of f64)" if args.tables == "tc"
                       else "f64 SIMT GEMM") if w["model"] == "bsc" else "f64",
            "statistics": ("Wp = <S>'Y by split-K tcgen05 fp16 hi+lo GEMM (within 1e-5 of max|Wp|), Wq/pi/sigma f64"
                           if args.stats in ("tc", "auto") else "f64") if w["model"] == "bsc" else "f64",
            "parallelism": "dp%d over N (array_split shards of one dataset)" % world}


def dtype_string(w, args):
    if w["model"] == "bsc" and (args.tables == "tc" or args.stats in ("tc", "auto")):
        return "f64 + fp16x2-split tensor-core tables (fp32 accumulate)"
    return "f64"


class Job:
    """One workload on this rank's shard: model, resident data, parameters, K sets."""

    def __init__(self, name, N_total, args, rank, world, device, comm):
        import torch

        from evo_b200.comm import shard_bounds
        from evo_b200.variational import init_states

        self.name, self.w = name, WORKLOADS[name]
        w = self.w
        self.N_total = N_total
        self.lo, self.hi = shard_bounds(N_total, world, rank)
        self.N = self.hi - self.lo
        self.device, self.world, self.rank = device, world, rank
        if w["model"] == "bsc":
            from evo_b200.bsc import BSC
            self.model = BSC(w["D"], w["H"], w["K"], comm=comm, device=device)
            self.model.tables, self.model.stats = args.tables, args.stats
        else:
            from evo_b200.sssc import SSSC
            learn = ("W", "pies", "sigma2") if w.get("missing") else ("W", "pies", "mus", "sigma2", "Psi")
            self.model = SSSC(w["D"], w["H"], w["K"], to_learn=learn, comm=comm, device=device)
        Y, X = device_rows(self.model, w, self.lo, self.hi, device)
        self.data = {"y": Y, "x_infr": X}
        if X is not None:
            self.data["x"] = X
        np.random.seed(1)
        theta = self.model.check_params(self.model.standard_init(self.data))
        for k in list(theta):
            if isinstance(theta[k], np.ndarray) and theta[k].ndim >= 1:
                theta[k] = torch.as_tensor(theta[k], device=device)  # device-resident parameters
        self.theta = theta
        sel, mut, P, C, G = w["ea"]
        self.cont = init_states(self.N, w["K"], w["H"], sel, mut, P, C, G, device=device, seed=0x5EED, n_offset=self.lo)
        self.recon = bool(w.get("missing")) or w["model"] == "es3c"
        self.iters_done = 0

    def step(self):
        F, nu, nsub, self.theta = self.model.step(self.theta, self.cont, self.data, do_reconstruction=self.recon)
        self.iters_done += 1
        self.last_sub = nsub
        return F, nu, self.model.last_counts["lpj_evaluated"]

    def fused_rest_fraction(self):
        """share of this shard's datapoints whose statistics the fused E-step tail handed to the separate pass (last step)"""
        r = self.model.device_data(self.data).cache.get("fs_rest")
        return None if r is None else float(r[0].item()) / max(1, self.N)

    def mean_active(self, sample=2000):
        """mean |s|, |s|^2, |s|^3 over the K sets of the first datapoints of the shard"""
        from evo_b200.variational import unpack_states

        n = min(sample, self.N)
        a = unpack_states(self.cont["ss_dev"][:n], self.w["H"]).sum(-1).double()
        return float(a.mean().item()), float((a * a).mean().item()), float((a ** 3).mean().item())


def timed_window(job, steps, barrier, sampler=None):
    """`steps` EM iterations between CUDA events on the launching stream; per-call kernel times alongside."""
    import torch

    from evo_b200 import _lib

    lib = _lib.load()
    job.model.timing = []
    barrier()
    if sampler is not None:
        sampler.mark()
    l0 = lib.evo_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    lpj_total, U_sum, F = 0.0, 0.0, None
    for _ in range(steps):
        F, nu, cnt = job.step()
        lpj_total += cnt
        U_sum += nu
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=job.device)
    if job.world > 1:
        import torch.distributed as dist
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    kern = job.model.timing_summary()
    job.model.timing = None
    return {"ms": float(ms.item()), "lpj": lpj_total, "U_mean": U_sum / steps, "F": F, "kern": kern,
            "launches": int(lib.evo_launch_count() - l0), "sub": getattr(job, "last_sub", None),
            "rest": job.fused_rest_fraction()}


def graph_window(job, steps, barrier, warm=3):
    """The same EM iterations as ONE replayed CUDA graph per iteration (evo_b200/emgraph.py; BSC, complete data): `warm`
    iterations (1 eager, capture, replays), then `steps` replays between CUDA events, the run's single host read after."""
    import torch

    g = job.model.em_graph(job.theta, job.cont, job.data, max_iterations=warm + steps)
    g.run(warm)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    F, _, _ = g.run(steps)
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=job.device)
    if job.world > 1:
        import torch.distributed as dist
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    lpj = float(g.records[:, 6].sum())
    job.theta = g.theta_out()
    job.iters_done += warm + steps
    return {"steps": steps, "warmup": warm, "after_iterations": job.iters_done - steps, "ms_per_step": ms / steps,
            "value": lpj / (ms * 1e-3), "unit": UNIT, "em_iters_per_sec": steps / (ms * 1e-3), "free_energy": F[-1],
            "host_reads": 1, "note": "one CUDA graph replay per EM iteration, scalars and logging on the device"}


def roofline_of(job, win, peak, peak_src, fma=None):
    """`roofline` of the dominant call: algorithmic bytes / its mean CUDA-event duration / measured HBM
    peak; for the ES3C E-step also the algorithmic flops against the measured FMA peaks (bound = max of the two)."""
    w, kern = job.w, win["kern"]
    own = {k: v for k, v in kern.items() if k.startswith("evo_")}
    alg = algorithmic_bytes(w, job.N, win["U_mean"])
    dom = max(own, key=lambda k: own[k][0] * own[k][1])
    if dom not in alg:
        return None
    ach = alg[dom] / (kern[dom][1] * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "traffic": (NCU_TRAFFIC.get("%s:%d" % (job.name, job.N), {}) or {}).get(dom), "peak_source": peak_src,
            "launch_ms": kern[dom][1], "algorithmic_bytes_per_launch": alg[dom]}
    if dom == "evo_estep" and "evo_bsc_suffstats_tc_finish" in kern:
        roof["note"] = ("evo_estep also forms the M-step statistics of its datapoints in its tail (fs_*): its duration covers "
                        "rows a'+c AND most of row d of SURVEY 8(d); the bytes counted are rows a'+c only")
    wi = ((NCU_TRAFFIC.get("%s:%d" % (job.name, job.N), {}) or {}).get("warp_inst") or {}).get(dom)
    if wi:
        # the kernel is bound by instruction issue, not by HBM: warp instructions per launch (ncu, same command) against
        # 4 issue slots x SMs x clock -- the time the launch would take at 100 % issue-slot utilisation
        import torch
        sms = torch.cuda.get_device_properties(job.device).multi_processor_count
        ceil_ms = wi / (4.0 * sms * 1.965e9) * 1e3
        roof["issue"] = {"warp_instructions_per_launch": wi, "issue_slots_per_s": 4.0 * sms * 1.965e9,
                         "launch_ms_at_full_issue": ceil_ms, "frac_of_issue_ceiling": ceil_ms / kern[dom][1],
                         "hbm_frac_at_full_issue": alg[dom] / (ceil_ms * 1e-3) / 1e9 / peak}
    if w["model"] == "es3c" and fma:
        a1, a2, a3 = job.mean_active()
        d_obs = w["D"] * (1.0 - w.get("missing", 0.0))
        flop_lpj = d_obs * (a2 + 5 * a1) + 2.0 / 3.0 * a3 + 2 * a2
        t = kern["evo_estep"][1] * 1e-3
        lpj_launch = win["lpj"] / max(1, kern["evo_estep"][0]) / job.world
        tf = flop_lpj * lpj_launch / t / 1e12
        roof["simt"] = {"algorithmic_flop_per_lpj": flop_lpj, "mean_active": a1, "achieved_TFLOPs": tf,
                        "fp32_fma_peak_TFLOPs": fma[32], "fp64_fma_peak_TFLOPs": fma[64],
                        "frac_of_fp32_peak": tf / fma[32], "frac_of_fp64_peak": tf / fma[64],
                        "note": "kernel arithmetic is f64; SURVEY 8(d)(b) formula D_obs(a^2+5a)+2/3a^3+2a^2 at the measured moments of |s|"}
    return roof


def kernels_record(win):
    kern = win["kern"]
    tot = sum(c * t for c, t in kern.values())
    return {k: {"launches": c, "mean_ms": t, "share": c * t / tot} for k, (c, t) in kern.items()}


def comm_record(win, steps, world, nbytes):
    """all-reduce time per step and bus bandwidth of the packed statistics (the only collective), and the time of the
    replicated parameter solve -- the two terms that do not shrink with the rank count."""
    out = {}
    k = win["kern"].get("allreduce")
    if k:
        bus = (2.0 * (world - 1) / world) * nbytes / (k[1] * 1e-3) / 1e9 if world > 1 and k[1] > 0 else 0.0
        out.update({"allreduce_ms_per_step": k[1] * k[0] / steps, "bytes": nbytes, "bus_GBps": bus})
    if "solve" in win["kern"]:
        out["replicated_solve_ms_per_step"] = win["kern"]["solve"][1] * win["kern"]["solve"][0] / steps
    return out or None


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"  # the version banner goes to stdout, which carries ONE JSON line
        dist.init_process_group("nccl", device_id=device)
    from evo_b200 import _lib
    from evo_b200.comm import Comm

    lib = _lib.load()
    comm = Comm()
    name = args.workload
    w = WORKLOADS[name]
    N_total = args.n if args.n else w["N"]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    peak, peak_src = measured_peaks()
    fma = None
    if rank == 0:  # measured FP32 / FP64 FMA peaks (the SIMT denominators of the ES3C roofline)
        import ctypes
        fma = {}
        for bits in (32, 64):
            v = ctypes.c_double(0.0)
            _lib.check(lib.evo_fma_peak(bits, 4096, ctypes.byref(v)), "evo_fma_peak")
            fma[bits] = v.value
    job = Job(name, N_total, args, rank, world, device, comm)
    model, data, cont = job.model, job.data, job.cont
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # nvidia-smi needs ~0.1 s before its first sample: started ahead, marked below
    for _ in range(args.warmup):
        job.step()
    # snapshot of (theta, K sets, lpj, iteration counter): the end-to-end loop below restarts from it, so both
    # timed regions run the SAME EM iterations (the cost of an iteration grows as the states get denser)
    snap = {"theta": {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in job.theta.items()},
            "ss": cont["ss_dev"].clone(), "lpj": cont["lpj_dev"].clone(), "iteration": cont.get("iteration", 0)}
    # ---- timed region: inputs resident in HBM ------------------------------------------------
    win = timed_window(job, args.steps, barrier, sampler if rank == 0 else None)
    clocks = sampler.stop() if rank == 0 else None
    ms, lpj_total, F = win["ms"], win["lpj"], win["F"]
    act_timed = job.mean_active()[0]
    roof = roofline_of(job, win, peak, peak_src, fma) if rank == 0 else None
    comm_rec = comm_record(win, args.steps, world, int(model.last_counts.get("stats_bytes", 0)))

    # ---- end to end: batch arrives from pinned host memory, parameters and result go back ------
    Y = data["y"]
    Yh = torch.empty(Y.shape, dtype=Y.dtype, pin_memory=True)
    Yh.copy_(Y)
    hdata = dict(data)
    hdata["y"] = Yh
    model.rebind_data(data, hdata)  # same resident shard, now fed from the host buffer
    th_host = {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in job.theta.items()}
    h2d = Yh.numel() * Yh.element_size() + sum(v.nbytes for v in th_host.values() if isinstance(v, np.ndarray))
    d2h = sum(v.nbytes for v in th_host.values() if isinstance(v, np.ndarray)) + 8 * 8

    # every step consumes a batch that was copied host->device for THAT step; the copy of step
    # i+1 is issued on a side stream before step i's kernels (double-buffered shard), so it runs
    # under them.  One H2D copy of Y and one D2H read of theta/F per step, all inside the timed region.
    def e2e_step(th, last):
        model.adopt_staged(hdata)
        # parameters go up first (small), then the next batch starts streaming behind them
        thd = {k: (torch.as_tensor(v).to(device, non_blocking=True) if isinstance(v, np.ndarray) and v.ndim >= 1 else v)
               for k, v in th.items()}
        if not last:
            model.stage_data(hdata)
        F, nu, nsub, thd = model.step(thd, cont, hdata, do_reconstruction=job.recon)
        th = {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in thd.items()}
        return th, model.last_counts["lpj_evaluated"]

    model.stage_data(hdata)
    for _ in range(min(args.warmup, 2)):
        th_host, _ = e2e_step(th_host, False)
    # back to the state the resident timed region started from
    cont["ss_dev"].copy_(snap["ss"])
    cont["lpj_dev"].copy_(snap["lpj"])
    cont["iteration"] = snap["iteration"]
    th_host = {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in snap["theta"].items()}
    del snap
    torch.cuda.synchronize()
    barrier()
    t0 = time.perf_counter()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    e2e_lpj = 0.0
    model.stage_data(hdata)  # the staged copy left by the warm-up is overwritten: steps copies for steps steps
    for i in range(args.steps):
        th_host, cnt = e2e_step(th_host, i == args.steps - 1)
        e2e_lpj += cnt
    g1.record()
    barrier()
    e2e_ms = torch.tensor([max(g0.elapsed_time(g1), (time.perf_counter() - t0) * 1e3)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_ms.item())
    # back to device-resident parameters for the later windows
    model.rebind_data(hdata, data)
    job.theta = {k: (torch.as_tensor(v, device=device) if isinstance(v, np.ndarray) and v.ndim >= 1 else v)
                 for k, v in th_host.items()}
    del Yh, hdata

    # ---- the SAME iterations (fresh job, same warm-up) with one replayed CUDA graph per EM iteration --------------
    graph_rec = None
    if w["model"] == "bsc" and not w.get("missing") and args.stats in ("tc", "auto") and not args.no_graph:
        jg = Job(name, N_total, args, rank, world, device, comm)
        graph_rec = graph_window(jg, args.steps, barrier, warm=max(args.warmup, 3))
        del jg
        torch.cuda.empty_cache()

    # ---- converged window: the E-step gets dearer as the states densify ----------
    at_density = None
    if args.density_iters > 0:
        while job.iters_done < args.density_iters:
            job.step()
        k2 = max(3, min(args.steps, 5))
        win2 = timed_window(job, k2, barrier)
        at_density = {"after_iterations": job.iters_done - k2, "steps": k2, "ms_per_step": win2["ms"] / k2,
                      "value": win2["lpj"] / (win2["ms"] * 1e-3), "mean_active_latents_per_state": job.mean_active()[0],
                      "free_energy": win2["F"], "kernels": kernels_record(win2),
                      "substituted_per_datapoint": win2["sub"], "fused_statistics_rest_fraction": win2["rest"]}

    # ---- short sub-records of the other single-GPU workloads (N = 1 only) -------------------------------------
    also = {}
    if world == 1 and not args.no_also:
        for other in [x for x in ("cfg2", "cfg3", "cfg4", "cfg1") if x != name]:
            j2 = Job(other, WORKLOADS[other]["N"], args, rank, world, device, comm)
            for _ in range(3):
                j2.step()
            w2 = timed_window(j2, 3, barrier)
            also[other] = {"config": workload_config(other, j2.w, j2.N_total, world, args), "steps": 3, "warmup": 3,
                           "ms_per_step": w2["ms"] / 3, "value": w2["lpj"] / (w2["ms"] * 1e-3), "unit": UNIT,
                           "em_iters_per_sec": 3 / (w2["ms"] * 1e-3), "free_energy": w2["F"],
                           "dtype": dtype_string(j2.w, args), "kernels": kernels_record(w2),
                           "roofline": roofline_of(j2, w2, peak, peak_src, fma), "gpu_launches": w2["launches"]}
            del j2
            if WORKLOADS[other]["model"] == "bsc" and not WORKLOADS[other].get("missing") and args.stats in ("tc", "auto") \
                    and not args.no_graph:
                j3 = Job(other, WORKLOADS[other]["N"], args, rank, world, device, comm)  # fresh: the same iterations
                also[other]["graph"] = graph_window(j3, 200 if other == "cfg1" else 3, barrier)
                del j3

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": METRIC, "value": lpj_total / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": dtype_string(w, args), "data": "synthetic",
        "config": workload_config(name, w, N_total, world, args),
        "em_iters_per_sec": args.steps / (ms * 1e-3),
        "lpj_per_step": lpj_total / args.steps, "new_unique_per_datapoint": win["U_mean"], "free_energy": F,
        "mean_active_latents_per_state": act_timed,
        "substituted_per_datapoint": win["sub"], "fused_statistics_rest_fraction": win["rest"],
        "e2e": {"value": e2e_lpj / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": win["launches"],
        "kernels": kernels_record(win),
        "roofline": roof, "comm": comm_rec, "clocks": clocks, "at_density": at_density,
        "fma_peak_TFLOPs": fma, "graph": graph_rec, "also": also,
    }
    if not args.no_cpu:
        line["cpu_baseline"] = cpu_arm(name, steps=1, warmup=1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# CPU arm: the UNMODIFIED reference (baseline/_ref/evo, copied from /root/reference by tools/install_reference.py
# -- git-ignored, travels to the GPU box like the built .so) on the host cores, imported with the committed
# single-rank mpi4py stand-in (tests/golden/_standin), one process per core like the reference's MPI ranks, each
# on its own bounded slice of the SAME dataset rows the GPU arm holds (kind "reference").  If the copy is
# absent the pinned NumPy oracle port runs instead (kind "port").
# ---------------------------------------------------------------------------------------------
CPU_SAMPLE = {"cfg1": 1000, "cfg2": 1000, "cfg5": 64, "cfg3": 64, "cfg4": 64}  # datapoints per core per step
REF_DIR = os.path.join(ROOT, "baseline", "_ref")


def reference_available():
    return os.path.isdir(os.path.join(REF_DIR, "evo", "models"))


def cpu_worker(args):
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    import warnings

    warnings.simplefilter("ignore")
    w = WORKLOADS[args.workload]
    n = args.n
    D, H, K = w["D"], w["H"], w["K"]
    rows = np.arange(args.cpu_worker * n, (args.cpu_worker + 1) * n, dtype=np.int64)
    Y, X = host_rows(w, rows)
    obs = np.ones_like(Y, dtype=bool) if X is None else X
    sel, mut, P, C, G = w["ea"]
    np.random.seed(1 + args.cpu_worker)
    use_ref = reference_available() and not os.environ.get("EVO_BENCH_PORT")
    if use_ref:
        sys.path.insert(0, os.path.join(ROOT, "tests", "golden", "_standin"))
        sys.path.insert(0, REF_DIR)
        from evo.models import BSC as RBSC, SSSC as RSSSC
        from evo.variational.utils import init_states as ref_init_states

        Yn = np.where(obs, Y, np.nan)
        data = {"y": Yn, "x_infr": obs.copy()}
        if w["model"] == "bsc":
            model = RBSC(D, H, K)
            recon = False
        else:
            learn = ["W", "pies", "sigma2"] if X is not None else ["W", "pies", "mus", "sigma2", "Psi"]
            model = RSSSC(D, H, K, to_learn=learn)
            data["x"] = obs.copy() if X is not None else np.zeros_like(obs)
            recon = True
        theta = model.check_params(model.standard_init(data))
        cont = ref_init_states(n, K, H, sel, mut, P, C, G)
        step = lambda th: model.step(th, cont, data, do_reconstruction=recon)
    else:
        from oracle import bsc as OB
        from oracle import sssc as OS
        from oracle import variational as OV

        mean = (Y * obs).sum(0) / np.maximum(obs.sum(0), 1)
        sigma = float(np.sqrt((((Y - mean) * obs) ** 2).sum() / obs.sum()))
        W0 = mean[:, None] + np.random.normal(scale=sigma / 4.0, size=(D, H))
        cont = OV.init_states(n, K, H, sel, mut, P, C, G)
        if w["model"] == "bsc":
            theta = {"W": W0, "pi": 1.0 / H, "sigma": sigma}
            data = {"y": Y, "x_infr": obs}
            step = lambda th: OB.step(th, cont, data, D, H, K)
        else:
            learn = ("W", "pies", "sigma2") if X is not None else ("W", "pies", "mus", "sigma2", "Psi")
            theta = {"W": W0, "pies": np.random.uniform(0.1, 0.5, size=H),
                     "mus": np.ones(H) if X is not None else np.random.normal(0, 1, H),
                     "Psi": np.eye(H), "sigma2": np.float64(sigma ** 2)}
            Yn = np.where(obs, Y, np.nan)
            data = {"y": Yn, "x_infr": obs, "x": obs if X is not None else np.zeros_like(obs)}
            step = lambda th: OS.em_step(th, cont, data, D, H, K, to_learn=learn, do_reconstruction=True)
    for _ in range(args.warmup):
        F, nu, nsub, theta = step(theta)
    t0 = time.perf_counter()
    lpj = 0.0
    for _ in range(args.steps):
        F, nu, nsub, theta = step(theta)
        lpj += n * K + nu * n
    print(json.dumps({"lpj": lpj, "sec": time.perf_counter() - t0, "F": float(F), "kind": "reference" if use_ref else "port"}),
          flush=True)


def cpu_arm(workload, steps, warmup, n_sub=None, cores=None):
    cores = cores or len(os.sched_getaffinity(0))
    n_sub = n_sub or CPU_SAMPLE[workload]
    env = dict(os.environ, OMP_NUM_THREADS="1", OPENBLAS_NUM_THREADS="1", MKL_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    cmd = [sys.executable, os.path.abspath(__file__), "--workload", workload, "--n", str(n_sub), "--steps", str(steps),
           "--warmup", str(warmup)]
    t0 = time.perf_counter()
    procs = [subprocess.Popen(cmd + ["--cpu-worker", str(i)], stdout=subprocess.PIPE, env=env) for i in range(cores)]
    outs = [json.loads(p.communicate()[0].decode().strip().splitlines()[-1]) for p in procs]
    wall = time.perf_counter() - t0
    sec = max(o["sec"] for o in outs)
    lpj = sum(o["lpj"] for o in outs)
    kind = outs[0]["kind"]
    what = ("the unmodified reference (baseline/_ref/evo, mpi4py stand-in, one process per core)" if kind == "reference"
            else "NumPy oracle port (oracle/), one process per core")
    return {"value": lpj / sec, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "rows [0, %d) of the GPU arm's dataset: %d datapoints per core x %d cores, %d timed EM iteration(s) "
                      "after %d warm-up, %s, 1 BLAS thread per process; whole call %.1f s"
                      % (n_sub * cores, n_sub, cores, steps, warmup, what, wall),
            "ms_per_step": sec / steps * 1e3, "em_iters_per_sec_on_sample": steps / sec}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    total = max(1, args.steps + args.warmup)
    n_sub = max(8, int(CPU_SAMPLE[args.workload] * 2 / total))
    cb = cpu_arm(args.workload, steps=args.steps, warmup=args.warmup, n_sub=n_sub)
    cfg = workload_config(args.workload, w, args.n if args.n else w["N"], world, args)
    cfg.pop("tables"), cfg.pop("statistics")
    cfg["parallelism"] = "%d host processes over a bounded sample of N" % cb["cores"]
    cfg["N_per_step"] = n_sub * cb["cores"]
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=sorted(WORKLOADS))
    ap.add_argument("--n", type=int, default=0, help="datapoints IN TOTAL over all GPUs (default: the workload's N)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-also", action="store_true", help="skip the short sub-records of the other workloads")
    ap.add_argument("--no-graph", action="store_true", help="skip the CUDA-graph replay windows")
    ap.add_argument("--density-iters", type=int, default=40,
                    help="also time a window after this many EM iterations (0 = off): states are denser by then")
    ap.add_argument("--tables", default="tc", choices=["tc", "f64"],
                    help="BSC table B = Y W: tcgen05 split-precision GEMM (default) or the f64 SIMT GEMM")
    ap.add_argument("--stats", default="auto", choices=["auto", "tc", "f64"],
                    help="BSC M-step statistic Wp = <S>'Y: split-K tcgen05 GEMM (tc, = auto) or f64 kernels")
    ap.add_argument("--lpj-compare", action="store_true",
                    help="time the dense tcgen05 log-joint against the sparse Gram kernel (BSC workloads)")
    ap.add_argument("--patches", action="store_true",
                    help="time the overlapping-patch kernels (SURVEY 8f row 1) against the HBM roofline; --n = image side")
    ap.add_argument("--cpu-worker", type=int, default=-1, help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.cpu_worker >= 0:
        return cpu_worker(args)
    if args.impl == "reference":
        return run_reference(args)
    if args.lpj_compare:
        return run_lpj_compare(args)
    if args.patches:
        return run_patches(args)
    return run_ours(args)


if __name__ == "__main__":
    main()
