"""Sharding parity on real GPUs: N datapoints sharded over R ranks (NCCL all-reduce of the packed statistics)
give the same K sets, F and parameters as one GPU holding all rows (tests/run_multigpu_parity.py, RNG mode: the
Philox streams are keyed by the GLOBAL datapoint index).  Spawns R = 2 (and 4, 8 if visible) ranks with
torchrun; skipped on a box with a single GPU.  The world-size-2 host logic is covered on CPU by
tests/test_comm_gloo.py."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    import torch

    return torch.cuda.device_count()


@pytest.mark.parametrize("ranks", [2, 4, 8])
def test_sharded_run_equals_single_gpu_run(ranks):
    if _gpus() < ranks:
        pytest.skip("needs %d GPUs, %d visible" % (ranks, _gpus()))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(ranks),
           "--master-addr", "127.0.0.1", "--master-port", str(29610 + ranks), os.path.join(ROOT, "tests", "run_multigpu_parity.py")]
    r = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=900)
    out = r.stdout + r.stderr
    keep = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(keep):
        open(os.path.join(keep, "multigpu_parity_%dgpu.log" % ranks), "w").write(out)
    assert r.returncode == 0 and "MULTIGPU_PARITY OK" in out, out[-3000:]
