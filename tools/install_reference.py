"""Copy the UNMODIFIED reference package (pure Python, /root/reference/evo) to baseline/_ref/evo so that
`bench.py --impl reference` and the `cpu_baseline` leg can time it on the GPU box's host cores
(/root/reference itself does not exist there).  baseline/_ref/ is git-ignored, not gpurun-ignored: it travels
with the snapshot like the built .so.  The reference imports mpi4py at module import (evo/models/_models.py:7);
mpi4py is not installable here (no network), so it is imported with the committed single-rank stand-in
tests/golden/_standin/mpi4py.  Nothing is modified: the copy is verified file by file."""
import filecmp
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(os.environ.get("EVO_REFERENCE", "/root/reference"), "evo")
DST = os.path.join(ROOT, "baseline", "_ref", "evo")


def install(quiet=False):
    if not os.path.isdir(SRC):
        if not quiet:
            print("install_reference: %s not found (GPU box?) -- keeping whatever is in baseline/_ref" % SRC)
        return os.path.isdir(DST)
    if os.path.isdir(DST):
        shutil.rmtree(DST)
    os.makedirs(os.path.dirname(DST), exist_ok=True)
    shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    cmp = filecmp.dircmp(SRC, DST, ignore=["__pycache__"])
    bad = cmp.diff_files + cmp.left_only
    assert not bad, "copy differs from the reference: %s" % bad
    if not quiet:
        print("install_reference: copied %s -> %s (unmodified, %d top-level entries)" % (SRC, DST, len(os.listdir(DST))))
    return True


if __name__ == "__main__":
    sys.exit(0 if install() else 1)
