"""CPU tests of the N>1 path (world size 2, gloo): the mpi4py-shaped `Comm`, N sharding by the
np.array_split rule, and that per-shard sufficient statistics all-reduced through it equal the
single-rank statistics (computed with the NumPy oracle -- no GPU involved)."""
import os
import socket
import sys
import warnings

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out):
    warnings.simplefilter("ignore")
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from evo_b200.comm import Comm, shard_bounds
    from oracle import bsc as OB
    from oracle import variational as OV

    comm = Comm()
    ok = comm.rank == rank and comm.size == world
    # mpi4py-shaped calls
    ok &= comm.allreduce(rank + 1) == sum(range(1, world + 1))
    buf, res = np.full(3, float(rank + 1)), np.zeros(3)
    comm.Allreduce([buf, "f8"], [res, "f8"])
    ok &= np.allclose(res, sum(range(1, world + 1)))
    ok &= comm.bcast({"a": rank}, root=0) == {"a": 0}
    t = torch.full((4,), float(rank), dtype=torch.float64)
    comm.allreduce_(t)
    ok &= float(t[0]) == sum(range(world))
    comm.Barrier()
    # sharded statistics == single-rank statistics
    rng = np.random.RandomState(0)
    N, D, H, S = 37, 6, 8, 5
    W = rng.normal(size=(D, H))
    Y = rng.normal(size=(N, D))
    ss = rng.rand(N, S, H) < 0.3
    lpj = rng.normal(size=(N, S)) * 3
    perm = {"background": False, "allzero": False, "singletons": False}
    full = OB.m_step_stats({"W": W}, {"lpj": lpj, "ss": ss, "S_perm": 0, "permanent": perm},
                           {"y": Y, "x_infr": np.ones_like(Y, bool)})
    lo, hi = shard_bounds(N, world, rank)
    part = OB.m_step_stats({"W": W}, {"lpj": lpj[lo:hi], "ss": ss[lo:hi], "S_perm": 0, "permanent": perm},
                           {"y": Y[lo:hi], "x_infr": np.ones((hi - lo, D), bool)})
    packed = torch.from_numpy(np.concatenate([np.ravel(x) for x in part[:3]] + [[part[3]]]))
    comm.allreduce_(packed)
    want = np.concatenate([np.ravel(x) for x in full[:3]] + [[full[3]]])
    ok &= np.allclose(packed.numpy(), want, rtol=1e-12, atol=1e-12)
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_comm_world_size_2_gloo():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert all(out.get(r) for r in range(world)), dict(out)


def test_comm_single_process_is_identity():
    sys.path.insert(0, ROOT)
    from evo_b200.comm import Comm

    c = Comm()
    assert c.size == 1 and c.rank == 0
    assert c.allreduce(3.5) == 3.5
    t = torch.arange(3.0)
    assert torch.equal(c.allreduce_(t.clone()), t)
