"""Communicator with the mpi4py-shaped surface the reference's models use (`comm=` ctor argument,
evo/models/_models.py:20,40; call sites SURVEY.md section 2.3), backed by torch.distributed.

NCCL over NVLink for CUDA tensors, gloo for host tensors / CPU tests.  With world size 1 (or no
initialised process group) every call is the identity, like a single-rank MPI run.
"""
import numpy as np
import torch
import torch.distributed as dist


def split_rows(N, R):
    """Row counts per rank by the reference's np.array_split rule (evo/utils/parallel.py:88-114)."""
    base, extra = divmod(int(N), int(R))
    return [base + (1 if r < extra else 0) for r in range(R)]


def shard_bounds(N, R, r):
    counts = split_rows(N, R)
    lo = sum(counts[:r])
    return lo, lo + counts[r]


class Comm:
    def __init__(self, group=None, single=False):
        """`single=True` gives a world-size-1 communicator even inside an initialised process group."""
        self.group = group
        self._on = dist.is_available() and dist.is_initialized() and not single
        self.rank = dist.get_rank(group) if self._on else 0
        self.size = dist.get_world_size(group) if self._on else 1
        self._nccl = {}

    # ---- device / tensor path (one fused allreduce per EM iteration) -------------------------
    def _native(self, device):
        """NCCL communicator behind the library's own C-ABI (`evo_comm_*`, include/evo_b200.h), created lazily per
        device: rank 0 makes the unique id, torch.distributed ships the 128 bytes (host plumbing), every rank joins.
        EVO_COMM=torch keeps torch.distributed's all_reduce as the data plane."""
        import ctypes as C
        import os

        if os.environ.get("EVO_COMM", "") == "torch" or dist.get_backend(self.group) != "nccl":
            return None
        key = str(device)
        if key not in self._nccl:
            from . import _lib

            lib = _lib.load()
            buf = (C.c_ubyte * 128)()
            if self.rank == 0:
                _lib.check(lib.evo_comm_unique_id(buf), "evo_comm_unique_id")
            box = [bytes(buf)]
            dist.broadcast_object_list(box, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0,
                                       group=self.group)
            raw = (C.c_ubyte * 128).from_buffer_copy(box[0])
            h = C.c_void_p()
            with torch.cuda.device(device):
                _lib.check(lib.evo_comm_init(self.rank, self.size, raw, C.byref(h)), "evo_comm_init")
            self._nccl[key] = (lib, h)
        return self._nccl[key]

    def allreduce_(self, t):
        """In-place SUM all-reduce of a torch tensor: float64 CUDA tensors (the packed statistics) through
        `evo_comm_allreduce` (NCCL via the C-ABI) on the current stream; anything else through torch.distributed."""
        if self.size > 1:
            nat = self._native(t.device) if (t.is_cuda and t.dtype == torch.float64 and t.is_contiguous()) else None
            if nat is not None:
                from . import _lib

                lib, h = nat
                with torch.cuda.device(t.device):
                    _lib.check(lib.evo_comm_allreduce(h, _lib.ptr(t), t.numel(), _lib.stream()), "evo_comm_allreduce")
            else:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    # ---- mpi4py-shaped surface ---------------------------------------------------------------
    def allreduce(self, x, op=None):
        """Lower-case mpi4py call: python scalar / small array in, reduced value out."""
        if self.size == 1:
            return x
        arr = np.asarray(x, dtype=np.float64)
        t = torch.from_numpy(np.atleast_1d(arr).copy())
        if dist.get_backend(self.group) == "nccl":
            t = t.cuda()
        dist.all_reduce(t, group=self.group)
        out = t.cpu().numpy().reshape(arr.shape)
        return out if arr.shape else type(x)(out)

    def Allreduce(self, send, recv, op=None):
        sbuf = send[0] if isinstance(send, (list, tuple)) else send
        rbuf = recv[0] if isinstance(recv, (list, tuple)) else recv
        np.copyto(rbuf, self.allreduce(np.asarray(sbuf)))

    def bcast(self, x, root=0):
        if self.size == 1:
            return x
        box = [x]
        dist.broadcast_object_list(box, src=root, group=self.group)
        return box[0]

    def Bcast(self, buf, root=0):
        if self.size == 1:
            return
        arr = buf[0] if isinstance(buf, (list, tuple)) else buf
        t = torch.from_numpy(arr)
        if dist.get_backend(self.group) == "nccl":
            d = t.cuda()
            dist.broadcast(d, src=root, group=self.group)
            arr[...] = d.cpu().numpy()
        else:
            dist.broadcast(t, src=root, group=self.group)

    def Barrier(self):
        if self.size > 1:
            dist.barrier(group=self.group)


COMM_WORLD = None


def world():
    """Default communicator (the reference's MPI.COMM_WORLD default)."""
    global COMM_WORLD
    if COMM_WORLD is None or (COMM_WORLD.size == 1 and dist.is_available() and dist.is_initialized()):
        COMM_WORLD = Comm()
    return COMM_WORLD
