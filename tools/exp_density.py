"""Diagnostic: active latents per state and list-overflow count of the E-step over EM iterations (cfg2)."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import bench
from evo_b200.bsc import BSC
from evo_b200.variational import init_states

w = bench.WORKLOADS["cfg2"] if hasattr(bench, "WORKLOADS") else None
name = [k for k, v in vars(bench).items() if isinstance(v, dict) and "cfg2" in v and isinstance(v["cfg2"], dict) and "ea" in v["cfg2"]][0]
w = getattr(bench, name)["cfg2"]
N = 20000
dev = torch.device("cuda", 0)
Y = bench.synth_device(w, N, 0, dev)
Y = Y[0] if isinstance(Y, tuple) else Y
m = BSC(w["D"], w["H"], w["K"])
m.tables = "tc"; m.stats = "auto"
data = {"y": Y, "x_infr": None}
np.random.seed(1)
th = m.check_params(m.standard_init(data))
cont = init_states(N, w["K"], w["H"], *w["ea"], device=dev, seed=9)
lut = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int32, device=dev)
for it in range(60):
    F, nu, ns, th = m.step(th, cont, data)
    if it % 5 == 4:
        b = cont["ss_dev"].view(torch.uint8).to(torch.int64)
        pc = lut[b].view(N, w["K"], -1).sum(-1).float()
        print(it + 1, "F %.3f" % F, "pi*H %.2f" % (th["pi"] * w["H"]), "mean|s| %.2f max|s| %d frac>15 %.4f" % (pc.mean().item(), int(pc.max().item()), (pc > 15).float().mean().item()),
              "ovf datapoints", int(m._ovf_list[0].item()), "of", N, flush=True)
