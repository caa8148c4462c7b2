"""torchrun script (not a pytest): N sharded over WORLD_SIZE GPUs must give the same K sets and the
same F (up to reduction order) as one GPU holding all rows, in RNG mode -- the Philox streams are
keyed by the GLOBAL datapoint index.   torchrun --nproc-per-node 2 tests/run_multigpu_parity.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(model_name, comm, lo, hi, Y, theta0, N, D, H, K, iters, device):
    from evo_b200.bsc import BSC
    from evo_b200.sssc import SSSC
    from evo_b200.variational import init_states

    if model_name == "bsc":
        m = BSC(D, H, K, comm=comm, device=device)
    else:
        m = SSSC(D, H, K, comm=comm, device=device)
    data = {"y": Y[lo:hi].copy(), "x_infr": np.ones((hi - lo, D), bool), "x": np.zeros((hi - lo, D), bool)}
    cont = init_states(hi - lo, K, H, "fit", "randflip", 4, 3, 2, device=device, seed=7, n_offset=lo)
    th = {k: (np.array(v) if isinstance(v, np.ndarray) else v) for k, v in theta0.items()}
    Fs = []
    for _ in range(iters):
        F, a, b, th = m.step(th, cont, data, do_reconstruction=(model_name == "es3c"))
        Fs.append(F)
    return np.array(Fs), cont["ss"], th


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    device = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=device)
    from evo_b200.comm import Comm, shard_bounds

    ok = True
    for model_name in ("bsc", "es3c"):
        rng = np.random.RandomState(3)
        N, D, H, K, iters = 501, 20, 40, 16, 4
        Wg = rng.normal(size=(D, H)) * 2
        s = rng.rand(N, H) < 3.0 / H
        Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32).astype(np.float64)
        if model_name == "bsc":
            th0 = {"W": rng.normal(size=(D, H)), "pi": 0.1, "sigma": 1.5}
        else:
            th0 = {"W": rng.normal(size=(D, H)), "pies": np.full(H, 0.1), "mus": np.ones(H), "Psi": np.eye(H),
                   "sigma2": np.float64(2.0)}
        lo, hi = shard_bounds(N, world, rank)
        Fs, ss, th = run(model_name, Comm(), lo, hi, Y, th0, N, D, H, K, iters, device)
        parts = [None] * world
        dist.all_gather_object(parts, ss)
        if rank == 0:
            F1, ss1, th1 = run(model_name, Comm(single=True), 0, N, Y, th0, N, D, H, K, iters, device)
            same = np.array_equal(np.concatenate(parts, axis=0), ss1)
            ferr = float(np.max(np.abs(Fs - F1) / np.abs(F1)))
            werr = float(np.abs(np.asarray(th["W"]) - np.asarray(th1["W"])).max() / np.abs(np.asarray(th1["W"])).max())
            print("%s world=%d: K sets identical=%s  F rel err=%.2e  W rel err=%.2e" % (model_name, world, same, ferr, werr))
            ok &= same and ferr < 1e-10 and werr < 1e-8
        dist.barrier()
    # ---- CUDA-graph replay with the NCCL all-reduce captured inside (evo_b200/emgraph.py) against step-by-step ----
    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    rng = np.random.RandomState(5)
    N, D, H, K, iters = 2001, 24, 40, 16, 4
    Wg = rng.normal(size=(D, H)) * 2
    s = rng.rand(N, H) < 3.0 / H
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)
    lo, hi = shard_bounds(N, world, rank)
    res = []
    for use_graph in (False, True):
        m = BSC(D, H, K, comm=Comm(), device=device)
        m.tables = m.stats = "tc"
        data = {"y": torch.as_tensor(Y[lo:hi], device=device), "x_infr": None}
        cont = init_states(hi - lo, K, H, "fit", "randflip", 4, 3, 2, device=device, seed=7, n_offset=lo)
        th = {"W": np.random.RandomState(9).normal(size=(D, H)), "pi": 0.1, "sigma": 1.5}
        if use_graph:
            g = m.em_graph(th, cont, data)
            Fs, _, _ = g.run(iters)
            th = g.theta_out()
        else:
            Fs = []
            for _ in range(iters):
                F, _, _, th = m.step(th, cont, data)
                Fs.append(F)
        res.append((np.array(Fs), cont["ss_dev"].cpu().numpy(), np.asarray(th["W"])))
    same = np.array_equal(res[0][1], res[1][1])
    ferr = float(np.max(np.abs(res[0][0] - res[1][0]) / np.abs(res[0][0])))
    werr = float(np.abs(res[0][2] - res[1][2]).max() / np.abs(res[0][2]).max())
    flags = torch.tensor([int(same and ferr < 1e-8 and werr < 1e-6)], device=device)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("bsc graph replay world=%d: K sets identical=%s  F rel err=%.2e  W rel err=%.2e  all ranks ok=%d"
              % (world, same, ferr, werr, int(flags.item())))
        ok &= bool(flags.item())
    if rank == 0:
        print("MULTIGPU_PARITY", "OK" if ok else "FAIL")
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
