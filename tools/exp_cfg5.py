"""Diagnostic: per-range state density / live states at cfg5 with large N."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import bench
from evo_b200.bsc import BSC
from evo_b200.variational import init_states

w = bench.WORKLOADS["cfg5"]
N = int(sys.argv[1]) if len(sys.argv) > 1 else 500000
dev = torch.device("cuda", 0)
Y, _ = bench.synth_device(w, N, 0, dev)
m = BSC(w["D"], w["H"], w["K"])
m.tables = "tc"; m.stats = "auto"
data = {"y": Y, "x_infr": None}
np.random.seed(1)
th = m.check_params(m.standard_init(data))
cont = init_states(N, w["K"], w["H"], *w["ea"], device=dev, seed=9)
lut = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int32, device=dev)
def report(tag):
    for lo in range(0, N, 100000):
        hi = min(N, lo + 100000)
        b = cont["ss_dev"][lo:hi].contiguous().view(torch.uint8).to(torch.int64)
        pc = lut[b].view(hi - lo, w["K"], -1).sum(-1).float()
        l = cont["lpj_dev"][lo:hi]
        live = ((l - l.max(1, keepdim=True).values) > np.log(1e-18)).float().sum(1)
        print(tag, lo, "mean|s| %.2f max %d" % (pc.mean().item(), int(pc.max().item())), "live/datapoint %.1f" % live.mean().item(),
              "lpj mean %.4g" % l.mean().item(), flush=True)
report("init")
for it in range(4):
    F, nu, ns, th = m.step(th, cont, data)
    print("iter", it, F, nu, ns, "ovf", int(m._ovf_list[0].item()))
report("after")

# ---- statistics call alone on prefixes / suffix windows of the same container
from evo_b200 import _lib
ss, lp = cont["ss_dev"], cont["lpj_dev"]
D, H, S = w["D"], w["H"], w["K"]
stats = torch.zeros(H * D + H * H + H + 1 + 16, dtype=torch.float64, device=dev)
def timed(lo, cnt):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stats.zero_()
    torch.cuda.synchronize()
    e0.record()
    m._call("evo_bsc_suffstats", _lib.ptr(ss[lo:]), _lib.ptr(lp[lo:]), _lib.ptr(Y[lo:]), cnt, D, H, S, 0, 0,
            float(th["pre1"]), float(th["pil_bar"]), _lib.ptr(stats), None, None, 0)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)
for lo, cnt in [(0, 100000), (0, 200000), (0, 300000), (0, 400000), (0, 500000), (300000, 100000), (400000, 100000), (0, 100000)]:
    if lo + cnt <= N:
        print("stats window", lo, cnt, "%.2f ms" % timed(lo, cnt), flush=True)
