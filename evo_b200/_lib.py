"""ctypes binding of libevo_b200.so (the C-ABI declared in include/evo_b200.h).

There is no CPU fallback: if the shared library is missing or a call fails this module raises.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libevo_b200.so")

SEL = {"fit": 0, "rand": 1, "all": 2}
MUT = {"randflip": 0, "sparseflip": 1, "cross": 2, "cross_randflip": 3, "cross_sparseflip": 4}
EVAL_BSC_GRAM, EVAL_BSC_DIRECT, EVAL_ES3C = 0, 1, 2


class EvoError(RuntimeError):
    pass


class EstepArgs(C.Structure):
    """Mirror of `evo_estep_args` (include/evo_b200.h)."""

    _fields_ = [
        ("N", C.c_int64), ("n_offset", C.c_int64),
        ("D", C.c_int32), ("H", C.c_int32), ("S", C.c_int32), ("S_perm", C.c_int32), ("Mprime", C.c_int32),
        ("n_parents", C.c_int32), ("n_children", C.c_int32), ("n_generations", C.c_int32),
        ("selection", C.c_int32), ("mutation", C.c_int32), ("background", C.c_int32),
        ("bitflip_prob", C.c_double), ("sparseness", C.c_double),
        ("seed", C.c_uint64), ("iteration", C.c_uint32), ("eval", C.c_int32),
        ("states", C.c_void_p), ("lpj", C.c_void_p),
        ("inj_children", C.c_void_p), ("inj_count", C.c_void_p), ("inj_cmax", C.c_int32), ("_pad0", C.c_int32),
        ("W", C.c_void_p), ("Y", C.c_void_p), ("mask", C.c_void_p),
        ("G", C.c_void_p), ("B", C.c_void_p), ("ynorm2", C.c_void_p),
        ("pre1", C.c_double), ("pil_bar", C.c_double),
        ("pil_bar_h", C.c_void_p), ("mus", C.c_void_p), ("Psi", C.c_void_p),
        ("sigma2", C.c_double),
        ("acc", C.c_void_p),
        ("overflow_list", C.c_void_p),
        ("fs_stats", C.c_void_p), ("fs_wp_workspace", C.c_void_p), ("fs_rest", C.c_void_p),
        ("theta_dev", C.c_void_p),
    ]


class EvolveArgs(C.Structure):
    """Mirror of `evo_evolve_args` (include/evo_b200.h)."""

    _fields_ = [
        ("e", EstepArgs),
        ("evolve_only", C.c_int32), ("lpj_given", C.c_int32), ("skip_eval", C.c_int32), ("gen0", C.c_int32),
        ("pool0", C.c_void_p), ("pool0_n", C.c_void_p),
        ("tr_parents", C.c_void_p), ("tr_npar", C.c_void_p), ("tr_children", C.c_void_p), ("tr_count", C.c_void_p),
        ("tr_pos", C.c_void_p),
        ("ev_states", C.c_void_p), ("ev_lpj", C.c_void_p), ("ev_count", C.c_void_p), ("ev_pool", C.c_void_p),
        ("ev_pool_n", C.c_void_p),
    ]


class Es3cStatsArgs(C.Structure):
    """Mirror of `evo_es3c_stats_args` (include/evo_b200.h)."""

    _fields_ = [
        ("N", C.c_int64),
        ("D", C.c_int32), ("H", C.c_int32), ("S", C.c_int32), ("S_perm", C.c_int32), ("learn_psi", C.c_int32),
        ("_pad", C.c_int32),
        ("states", C.c_void_p), ("lpj", C.c_void_p), ("Y", C.c_void_p), ("mask", C.c_void_p),
        ("recon_keep", C.c_void_p), ("G", C.c_void_p), ("B", C.c_void_p), ("ynorm2", C.c_void_p),
        ("Wt", C.c_void_p), ("Psi", C.c_void_p), ("mus", C.c_void_p),
        ("sigma2", C.c_double), ("stats", C.c_void_p), ("Yrec", C.c_void_p), ("q_skip", C.c_double),
    ]


class Es3cUpdateArgs(C.Structure):
    """Mirror of `evo_es3c_update_args` (include/evo_b200.h)."""

    _fields_ = [
        ("D", C.c_int32), ("H", C.c_int32),
        ("learn_W", C.c_int32), ("learn_pies", C.c_int32), ("learn_mus", C.c_int32), ("learn_Psi", C.c_int32),
        ("learn_sigma2", C.c_int32), ("background", C.c_int32), ("incomplete", C.c_int32), ("_pad", C.c_int32),
        ("N_total", C.c_double), ("n_obs_total", C.c_double), ("ysq", C.c_double), ("sigma2_old", C.c_double),
        ("eps_pies", C.c_double), ("eps_mus", C.c_double), ("eps_Psi", C.c_double), ("eps_sigma2", C.c_double),
        ("stats", C.c_void_p), ("W", C.c_void_p), ("pies", C.c_void_p), ("mus", C.c_void_p), ("Psi", C.c_void_p),
        ("scal", C.c_void_p), ("workspace", C.c_void_p),
    ]


_vp, _i, _i64, _d, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_uint64

SIGNATURES = {
    "evo_last_error": (C.c_char_p, []),
    "evo_version": (_i, []),
    "evo_device_sm_count": (_i, []),
    "evo_launch_count": (C.c_ulonglong, []),
    "evo_states_pack": (_i, [_vp, _vp, _i64, _i, _vp]),
    "evo_states_unpack": (_i, [_vp, _vp, _i64, _i, _vp]),
    "evo_states_init": (_i, [_vp, _i64, _i, _i, _i, _i, _d, _u64, _i64, _vp]),
    "evo_bsc_prepare": (_i, [_vp, _vp, _i64, _i, _i, _vp, _vp, _vp, _vp]),
    "evo_bsc_tables_tc_workspace_bytes": (_i64, [_i64, _i, _i]),
    "evo_bsc_tables_tc_set_data": (_i, [_vp, _i64, _i, _i, _vp, _vp]),
    "evo_bsc_tables_tc": (_i, [_vp, _i64, _i, _i, _vp, _vp, _vp]),
    "evo_bsc_lpj": (_i, [_i, _vp, _i, _i64, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _d, _d, _vp, _i64, _vp, _vp]),
    "evo_bsc_lpj_gemm_workspace_bytes": (_i64, [_i, _i]),
    "evo_bsc_lpj_gemm_prepare": (_i, [_vp, _i, _i, _vp, _vp]),
    "evo_bsc_lpj_gemm": (_i, [_vp, _i, _i64, _i, _i, _i, _vp, _vp, _vp, _d, _d, _vp, _i64, _vp, _vp]),
    "evo_estep": (_i, [C.POINTER(EstepArgs), _vp]),
    "evo_evolve": (_i, [C.POINTER(EvolveArgs), _vp]),
    "evo_evolve_dims": (_i, [C.POINTER(EstepArgs), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "evo_select_topk": (_i, [_vp, _vp, _i64, _i, _i, _i, _i, _vp, _vp, _vp, _i, _vp, _vp]),
    "evo_free_energy": (_i, [_vp, _i64, _i, _i64, _vp, _vp]),
    "evo_bsc_suffstats_workspace_bytes": (_i64, [_i64, _i, _i]),
    "evo_bsc_suffstats": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _i, _i, _d, _d, _vp, _vp, _vp, _i64, _vp]),
    "evo_bsc_wp_tc_workspace_bytes": (_i64, [_i64, _i, _i]),
    "evo_bsc_wp_tc_set_data": (_i, [_vp, _i64, _i, _i, _vp, _vp]),
    "evo_bsc_wp_tc_begin": (_i, [_i64, _i, _i, _vp, _vp]),
    "evo_bsc_suffstats_tc_finish": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _i, _d, _d, _vp, _vp, _vp, _vp, _vp]),
    "evo_bsc_suffstats_tc": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _i, _d, _d, _vp, _vp, _vp, _vp]),
    "evo_bsc_solve_workspace_bytes": (_i64, [_i, _i]),
    "evo_bsc_solve": (_i, [_vp, _d, _i, _i, _i, _i, _d, _d, _vp, _vp, _vp, _vp, _vp]),
    "evo_bsc_em_advance": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _d, _d, _d, _d, _vp]),
    "evo_bsc_solve_static": (_i, [_vp, _d, _i, _i, _i, _i, _d, _d, _vp, _vp, _vp, _vp]),
    "evo_es3c_update_workspace_bytes": (_i64, [_i, _i]),
    "evo_es3c_update": (_i, [C.POINTER(Es3cUpdateArgs), _vp]),
    "evo_es3c_set_scratch": (_i, [_vp, _i]),
    "evo_es3c_lpj": (_i, [_vp, _i, _i64, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _d, _vp, _i64, _vp, _vp]),
    "evo_es3c_suffstats": (_i, [C.POINTER(Es3cStatsArgs), _vp]),
    "evo_bsc_reconstruct": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _i, _vp, _vp]),
    "evo_generate_data": (_i, [_vp, _vp, _d, _vp, _vp, _i, _d, _i64, _i64, _i, _i, _u64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "evo_column_moments": (_i, [_vp, _vp, _i64, _i, _vp, _vp]),
    "evo_normal_fill": (_i, [_vp, _i64, _i, _vp, _d, _u64, _vp]),
    "evo_fma_peak": (_i, [_i, _i, C.POINTER(C.c_double)]),
    "evo_bsc_full_logsumexp": (_i, [_i, _i64, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _d, _d, _vp, _vp]),
    "evo_es3c_full_logsumexp": (_i, [_i64, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _d, _vp, _vp]),
    "evo_comm_load": (_i, [C.c_char_p]),
    "evo_comm_unique_id": (_i, [_vp]),
    "evo_comm_init": (_i, [_i, _i, _vp, C.POINTER(C.c_void_p)]),
    "evo_comm_allreduce": (_i, [_vp, _vp, _i64, _vp]),
    "evo_comm_destroy": (_i, [_vp]),
    "evo_patches_count": (_i64, [_i, _i, _i, _i, _i]),
    "evo_patches_extract": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "evo_patches_merge": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "evo_sum_sqdiff": (_i, [_vp, _vp, _i64, _vp, _vp]),
}

_lib = None


def load():
    """Load the shared library once; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise EvoError(
            "evo_b200: %s is missing -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)" % LIB_PATH
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().evo_last_error()
        raise EvoError("%s failed (%d): %s" % (what, rc, msg.decode() if msg else "?"))


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "evo_b200 kernels need contiguous CUDA tensors"
    return C.c_void_p(t.data_ptr())


def stream():
    import torch

    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
