"""CPU tests: the C-ABI library loads and exports every symbol include/evo_b200.h declares; the
ctypes structs match the C layout; host-side sharding logic."""
import ctypes
import os
import re
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "evo_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(evo_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from evo_b200 import _lib

    lib = _lib.load()
    syms = declared_symbols()
    assert len(syms) >= 15
    for s in syms:
        assert hasattr(lib, s), s
    assert lib.evo_version() >= 100
    assert set(_lib.SIGNATURES) == set(syms)


def test_struct_layouts_match_c():
    from evo_b200 import _lib

    prog = r'''
#include <stdio.h>
#include <stddef.h>
#include "evo_b200.h"
int main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(evo_estep_args), offsetof(evo_estep_args, acc),
  offsetof(evo_estep_args, W), sizeof(evo_es3c_stats_args), offsetof(evo_es3c_stats_args, q_skip),
  sizeof(evo_evolve_args), offsetof(evo_evolve_args, pool0), offsetof(evo_evolve_args, ev_pool_n));}
'''
    path = "/tmp/evo_abi_check.c"
    open(path, "w").write(prog)
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), path, "-o", "/tmp/evo_abi_check"])
    a, b, c, d, e, f, g, h = map(int, subprocess.check_output(["/tmp/evo_abi_check"]).split())
    assert ctypes.sizeof(_lib.EvolveArgs) == f
    assert _lib.EvolveArgs.pool0.offset == g and _lib.EvolveArgs.ev_pool_n.offset == h
    assert ctypes.sizeof(_lib.EstepArgs) == a
    assert _lib.EstepArgs.acc.offset == b and _lib.EstepArgs.W.offset == c
    assert ctypes.sizeof(_lib.Es3cStatsArgs) == d and _lib.Es3cStatsArgs.q_skip.offset == e


def test_missing_library_fails_loudly(monkeypatch):
    from evo_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libevo_b200.so")
    try:
        _lib.load()
        raise AssertionError("load() must raise when the CUDA library is missing")
    except _lib.EvoError as e:
        assert "no CPU fallback" in str(e)


def test_split_rows_is_array_split_rule():
    from evo_b200.comm import shard_bounds, split_rows

    for N, R in [(10, 3), (1000, 8), (7, 7), (5, 8), (100000, 6)]:
        want = [len(x) for x in np.array_split(np.arange(N), R)]
        assert split_rows(N, R) == want
        lo = 0
        for r in range(R):
            assert shard_bounds(N, R, r) == (lo, lo + want[r])
            lo += want[r]
