#!/usr/bin/env python
"""Benchmark of the evo hot path (truncated-variational E-step + M-step) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2] [--impl ours|reference]

A "step" is one EM iteration (`model.step`: E-step over every datapoint of the shard, sufficient
statistics, all-reduce, parameter update) on synthetic data of the shape BASELINE.json names.
Metric: log-joints evaluated per second (whole job, all ranks); EM iterations/s rides along.
Weak scaling: every rank owns N datapoints of the workload (no data-path collective; the only
collective is the all-reduce of the packed sufficient statistics).
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
    # name: model, D, H, N (per GPU), K, (selection, mutation, parents, children, generations), active latents
    "cfg1": dict(model="bsc", D=25, H=10, N=1000, K=20, ea=("fit", "randflip", 10, 1, 1), act=2.0, bars=True),
    "cfg2": dict(model="bsc", D=256, H=300, N=100000, K=64, ea=("fit", "randflip", 8, 8, 2), act=5.0),
    "cfg5": dict(model="bsc", D=1024, H=1024, N=1000000, K=128, ea=("fit", "randflip", 16, 8, 2), act=8.0),
    "cfg3": dict(model="es3c", D=144, H=512, N=250000, K=64, ea=("fit", "randflip", 8, 8, 2), act=4.0),
    "cfg4": dict(model="es3c", D=144, H=512, N=250000, K=64, ea=("fit", "randflip", 8, 8, 2), act=4.0, missing=0.3),
}
DESCR = {
    "cfg1": "examples/bars-test BSC, 5x5 bars (D=25, H=10), N=1000, K=20",
    "cfg2": "BSC feature learning, synthetic 16x16 patches (D=256, H=300), N=100k, K=64",
    "cfg5": "large-scale BSC (D=1024, H=1024), N=1M, K=128",
    "cfg3": "ES3C image denoising, synthetic 12x12 patches (D=144, H=512), N=250k, K=64",
    "cfg4": "ES3C image inpainting, 30% missing (D=144, H=512), N=250k, K=64",
}
METRIC, UNIT = "log_joints_per_sec", "log-joints/s"


def synth_host(w, N, seed):
    """Synthetic dataset with NumPy's legacy RandomState -- used by the CPU arm and,
    through the same recipe on the device, by the GPU arm."""
    rng = np.random.RandomState(seed)
    D, H = w["D"], w["H"]
    if w.get("bars"):
        R = H // 2
        Wg = np.zeros((R, R, H))
        for i in range(R):
            Wg[i, :, i] = 10.0
            Wg[:, i, R + i] = 10.0
        Wg = Wg.reshape(R * R, H)
    else:
        Wg = rng.normal(size=(D, H))
    s = rng.random_sample((N, H)) < w["act"] / H
    if w["model"] == "es3c":
        z = rng.normal(loc=1.0, size=(N, H))
        Y = (s * z) @ Wg.T + rng.normal(size=(N, D))
    else:
        Y = s.astype(np.float64) @ Wg.T + rng.normal(size=(N, D))
    Y = Y.astype(np.float32)
    X = None
    if w.get("missing"):
        X = rng.random_sample((N, D)) >= w["missing"]
    return Y, X


def synth_device(w, N, seed, device):
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(seed)
    D, H = w["D"], w["H"]
    if w.get("bars"):
        Y, X = synth_host(w, N, seed)
        return torch.as_tensor(Y, device=device), None
    Wg = torch.randn((D, H), generator=g, device=device)
    Y = torch.empty((N, D), dtype=torch.float32, device=device)
    X = None
    chunk = 65536
    for lo in range(0, N, chunk):
        hi = min(N, lo + chunk)
        s = (torch.rand((hi - lo, H), generator=g, device=device) < w["act"] / H).float()
        if w["model"] == "es3c":
            s = s * (1.0 + torch.randn((hi - lo, H), generator=g, device=device))
        Y[lo:hi] = s @ Wg.T + torch.randn((hi - lo, D), generator=g, device=device)
    if w.get("missing"):
        X = torch.rand((N, D), generator=g, device=device) >= w["missing"]
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
    Y, X = synth_device(w, N, 1000, device)
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


# dram__bytes_read.sum + dram__bytes_write.sum per call from the `ncu --set full` capture of exactly the
# default command (E-step: profiles/r01i_cfg2_full_estep_ncu_summary.txt, the other calls:
# profiles/r01g_cfg2_full_ncu_summary.txt -- unchanged kernels; a call = all its kernels); None where no
# capture at that size exists.
NCU_TRAFFIC = {("cfg2", 100000): {
    # list kernel + packed-bits pass over the overflow list
    "evo_estep": 497.574656e6 + 97.439744e6 + 0.063488e6,
    # statistics pass + split-K Wp GEMM (the memset of the <s> planes, 154 MB, is not a kernel)
    "evo_bsc_suffstats_tc": 413.756672e6 + 141.957888e6 + 260.738560e6 + 6.116864e6,
    # wmax + W' split + table GEMM
    "evo_bsc_tables_tc": 0.620288e6 + 0.629760e6 + 167.747072e6 + 188.397568e6,
    "evo_bsc_prepare": 0.621568e6}}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return float(j["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def run_ours(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    from evo_b200 import _lib
    from evo_b200.comm import Comm
    from evo_b200.variational import init_states

    w = WORKLOADS[args.workload]
    N = args.n if args.n else w["N"]
    comm = Comm()
    if w["model"] == "bsc":
        from evo_b200.bsc import BSC
        model = BSC(w["D"], w["H"], w["K"], comm=comm, device=device)
        model.tables = args.tables
        model.stats = args.stats
    else:
        from evo_b200.sssc import SSSC
        model = SSSC(w["D"], w["H"], w["K"], comm=comm, device=device)
    lib = _lib.load()

    Y, X = synth_device(w, N, 1000 + rank, device)
    data = {"y": Y, "x_infr": X}
    if X is not None:
        data["x"] = X
    np.random.seed(1)
    theta = model.check_params(model.standard_init(data))
    for k in list(theta):
        if isinstance(theta[k], np.ndarray) and theta[k].ndim >= 1:
            theta[k] = torch.as_tensor(theta[k], device=device)  # device-resident parameters
    sel, mut, P, C, G = w["ea"]
    cont = init_states(N, w["K"], w["H"], sel, mut, P, C, G, device=device, seed=0x5EED, n_offset=rank * N)
    recon = bool(w.get("missing")) or w["model"] == "es3c"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_step(th):
        F, nu, nsub, th = model.step(th, cont, data, do_reconstruction=recon)
        return F, nu, nsub, th, model.last_counts["lpj_evaluated"]

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # nvidia-smi needs ~0.1 s before its first sample: started ahead, marked below
    for _ in range(args.warmup):
        F, nu, nsub, theta, _ = one_step(theta)
    # snapshot of (theta, K sets, lpj, iteration counter): the end-to-end loop below restarts from it, so both
    # timed regions run the SAME EM iterations (the cost of an iteration grows as the states get denser)
    snap = {"theta": {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in theta.items()},
            "ss": cont["ss_dev"].clone(), "lpj": cont["lpj_dev"].clone(), "iteration": cont.get("iteration", 0)}
    # ---- timed region: inputs resident in HBM ------------------------------------------------
    model.timing = []
    barrier()
    if rank == 0:
        sampler.mark()
    l0 = lib.evo_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    lpj_total, U_sum = 0.0, 0.0
    for _ in range(args.steps):
        F, nu, nsub, theta, cnt = one_step(theta)
        lpj_total += cnt
        U_sum += nu
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = lib.evo_launch_count() - l0
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    kern = model.timing_summary()
    model.timing = None

    # ---- end to end: batch arrives from pinned host memory, parameters and result go back ------
    Yh = torch.empty(Y.shape, dtype=Y.dtype, pin_memory=True)
    Yh.copy_(Y)
    hdata = dict(data)
    hdata["y"] = Yh
    model._data = {id(Yh): model.device_data(data)}  # same resident shard, now fed from the host buffer
    th_host = {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in theta.items()}
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
            if os.environ.get("EVO_BENCH_E2E_SERIAL"):  # diagnostic: no copy/compute overlap
                torch.cuda.current_stream().wait_stream(model._copy_stream)
        F, nu, nsub, thd = model.step(thd, cont, hdata, do_reconstruction=recon)
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
    prof = None
    if os.environ.get("EVO_BENCH_PROFILE"):
        import cProfile
        prof = cProfile.Profile()
        prof.enable()
    if os.environ.get("EVO_BENCH_E2E_KERNELS"):  # diagnostic: per-call device time inside the e2e loop (stderr)
        model.timing = []
    for i in range(args.steps):
        th_host, cnt = e2e_step(th_host, i == args.steps - 1)
        e2e_lpj += cnt
    if prof is not None:
        import pstats
        prof.disable()
        pstats.Stats(prof, stream=sys.stderr).sort_stats("cumulative").print_stats(30)
    g1.record()
    barrier()
    e2e_ms = torch.tensor([max(g0.elapsed_time(g1), (time.perf_counter() - t0) * 1e3)], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_ms.item())
    if model.timing is not None:
        print("e2e kernels:", {k: round(v[1], 3) for k, v in model.timing_summary().items()}, file=sys.stderr)
        model.timing = None

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    U_mean = U_sum / args.steps
    alg = algorithmic_bytes(w, N, U_mean)
    peak, peak_src = measured_peaks()
    tot_t = sum(c * t for c, t in kern.values())
    dom = max(kern, key=lambda k: kern[k][0] * kern[k][1])
    roof = None
    if dom in alg:
        ach = alg[dom] / (kern[dom][1] * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": NCU_TRAFFIC.get((args.workload, N), {}).get(dom), "peak_source": peak_src, "launch_ms": kern[dom][1],
                "algorithmic_bytes_per_launch": alg[dom]}
    line = {
        "metric": METRIC, "value": lpj_total / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%s: %s" % (args.workload, DESCR[args.workload]), "N_per_gpu": N,
                   "ea": "%s/%s parents=%d children=%d generations=%d" % tuple(w["ea"]),
                   "l2": "inputs larger than L2 (K sets + tables > 126 MB)" if N * w["K"] * 4 * ((w["H"] + 31) // 32) > 126e6
                   else "working set fits L2 (small reference config)",
                   "tables": ("B = Y W by tcgen05 fp16 hi+lo split GEMM (lpj within 1e-5 of f64)" if args.tables == "tc"
                              else "f64 SIMT GEMM") if w["model"] == "bsc" else "f64",
                   "statistics": ("Wp = <S>'Y by split-K tcgen05 fp16 hi+lo GEMM (within 1e-5 of max|Wp|), Wq/pi/sigma f64"
                                  if args.stats in ("tc", "auto")
                                  else "f64") if w["model"] == "bsc" else "f64",
                   "parallelism": "dp%d over N" % world},
        "em_iters_per_sec": args.steps / (ms * 1e-3),
        "lpj_per_step": lpj_total / args.steps, "new_unique_per_datapoint": U_mean, "free_energy": F,
        "e2e": {"value": e2e_lpj / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": int(launches),
        "kernels": {k: {"launches": c, "mean_ms": t, "share": c * t / tot_t} for k, (c, t) in kern.items()},
        "roofline": roof, "clocks": clocks,
    }
    if not args.no_cpu:
        line["cpu_baseline"] = cpu_port(args.workload, steps=1, warmup=1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# CPU arm: the NumPy oracle (a restatement of the reference's per-datapoint NumPy path, pinned
# against the unmodified reference -- kind "port") on the host cores, one process per core like
# the reference's MPI ranks, each on its own bounded sample of the same workload.
# ---------------------------------------------------------------------------------------------
CPU_SAMPLE = {"cfg1": 1000, "cfg2": 1000, "cfg5": 64, "cfg3": 64, "cfg4": 64}  # datapoints per core per step


def cpu_worker(args):
    os.environ.setdefault("OMP_NUM_THREADS", "1")
    import warnings

    warnings.simplefilter("ignore")
    from oracle import bsc as OB
    from oracle import sssc as OS
    from oracle import variational as OV

    w = WORKLOADS[args.workload]
    n = args.n
    D, H, K = w["D"], w["H"], w["K"]
    Y, X = synth_host(w, n, 1000 + args.cpu_worker)
    Y = Y.astype(np.float64)
    np.random.seed(1 + args.cpu_worker)
    obs = np.ones_like(Y, dtype=bool) if X is None else X
    mean = (Y * obs).sum(0) / np.maximum(obs.sum(0), 1)
    sigma = float(np.sqrt((((Y - mean) * obs) ** 2).sum() / obs.sum()))
    W0 = mean[:, None] + np.random.normal(scale=sigma / 4.0, size=(D, H))
    sel, mut, P, C, G = w["ea"]
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
        step(theta)
    t0 = time.perf_counter()
    lpj = 0.0
    for _ in range(args.steps):
        F, nu, nsub, theta = step(theta)
        lpj += n * K + nu * n
    print(json.dumps({"lpj": lpj, "sec": time.perf_counter() - t0, "F": F}), flush=True)


def cpu_port(workload, steps, warmup, n_sub=None, cores=None):
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
    return {"value": lpj / sec, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d datapoints per core x %d cores, %d timed EM iteration(s) after %d warm-up, NumPy oracle "
                      "(oracle/bsc.py), 1 BLAS thread per process; whole call %.1f s" % (n_sub, cores, steps, warmup, wall),
            "ms_per_step": sec / steps * 1e3, "em_iters_per_sec_on_sample": steps / sec}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    total = max(1, args.steps + args.warmup)
    n_sub = max(16, int(CPU_SAMPLE[args.workload] * 2 / total))
    cb = cpu_port(args.workload, steps=args.steps, warmup=args.warmup, n_sub=n_sub)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": cb["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%s: %s" % (args.workload, DESCR[args.workload]), "N_per_step": n_sub * cb["cores"],
                   "ea": "%s/%s parents=%d children=%d generations=%d" % tuple(w["ea"]),
                   "parallelism": "%d host processes over N" % cb["cores"]},
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
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--n", type=int, default=0, help="datapoints per GPU (default: the workload's N)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
