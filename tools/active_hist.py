"""Histogram of |s| over the K sets of a workload after some EM iterations (diagnostic for the ES3C evaluator paths).
usage: python tools/active_hist.py cfg3 20000 44"""
import argparse
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

name, N, iters = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
args = argparse.Namespace(tables="tc", stats="auto")
from evo_b200.comm import Comm  # noqa: E402
from evo_b200.variational import unpack_states  # noqa: E402

torch.cuda.set_device(0)
job = bench.Job(name, N, args, 0, 1, torch.device("cuda", 0), Comm())
for it in range(iters):
    job.step()
    if it in (5, iters - 1):
        a = unpack_states(job.cont["ss_dev"][:4000], job.w["H"]).sum(-1).cpu().numpy().ravel()
        h = np.bincount(a, minlength=40)
        cum = np.cumsum(h) / h.sum()
        print("iteration", it + 1, "mean |s| %.2f" % a.mean(), "share <=2 %.3f <=4 %.3f <=8 %.3f <=16 %.3f <=32 %.3f max %d"
              % (cum[2], cum[4], cum[8], cum[16], cum[min(32, len(cum) - 1)], a.max()))
        lpj = job.cont["lpj_dev"][:4000]
        q = torch.softmax(lpj, dim=1)[:, job.cont["S_perm"]:]
        live = (q > 1e-30).cpu().numpy().ravel()
        hl = np.bincount(a[live], minlength=40)
        cl = np.cumsum(hl) / max(1, hl.sum())
        print("   live states (q > 1e-30): %.1f of %d per datapoint; share <=2 %.3f <=8 %.3f <=16 %.3f"
              % (live.sum() / 4000, q.shape[1], cl[2], cl[8], cl[16]))
