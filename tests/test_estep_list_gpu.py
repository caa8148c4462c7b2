"""The E-step's index-list state representation (estep_flip.cuh; <= 31 active latents per state, 64 bytes
per explicit state in shared memory, children implicit) against the packed-bits kernel (EVO_ESTEP_BITS=1) in RNG mode: same Philox
streams and the same parameters, so K sets and lpj must be bit-identical (F to summation order) -- including datapoints whose states do not fit a list
and are redone by the packed-bits kernel inside the same call."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run_both(D, H, K, N, act, ea, iters=3):
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    old = os.environ.pop("EVO_ESTEP_BITS", None)
    try:
        rng = np.random.RandomState(D + H)
        Wg = rng.normal(size=(D, H)) * 2.0
        s = rng.random_sample((N, H)) < act / H
        Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)
        m = BSC(D, H, K)
        m.tables = "f64"
        data = {"y": torch.as_tensor(Y, device=m.device), "x_infr": None}
        np.random.seed(1)
        th = m.check_params(m.standard_init(data))
        cont = init_states(N, K, H, "fit", "randflip", *ea, p_init_Kn=act / H, device=m.device, seed=11)
        _, _, _, th = m.step(th, cont, data)  # one M-step so that the parameters carry structure
        # With H > N the M-step leaves latents without posterior mass an all-zero column of W (minimum-norm solve): states
        # that differ only in such latents have EXACTLY equal log-joints in exact arithmetic, and which of them survives
        # is decided by rounding (the reference's own order there is unspecified -- SURVEY App. A.6(i)).  A tiny
        # perturbation removes the exact ties, so that the two kernels can be compared state for state.
        th["W"] = np.asarray(th["W"]) + 1e-6 * rng.normal(size=(D, H))
        out = {}
        # E-steps under the SAME fixed parameters (the M-step's f64 atomics are unordered, so two
        # trajectories through M-steps differ in the last bits whichever kernel ran)
        for bits in (False, True):
            if bits:
                os.environ["EVO_ESTEP_BITS"] = "1"
            cont = init_states(N, K, H, "fit", "randflip", *ea, p_init_Kn=act / H, device=m.device, seed=11)
            Fs, ovf = [], 0
            for _ in range(iters):
                Fs.append(m.E_step(th, cont, data)[0])
                ovf = max(ovf, int(m._ovf_list[0].item()))  # datapoints redone by the packed-bits kernel
            out[bits] = (cont["ss_dev"].cpu().numpy(), cont["lpj_dev"].cpu().numpy(), np.array(Fs), ovf)
        return out
    finally:
        os.environ.pop("EVO_ESTEP_BITS", None)
        if old is not None:
            os.environ["EVO_ESTEP_BITS"] = old


@pytest.mark.parametrize(
    "D,H,K,N,act,ea,expect_overflow",
    [
        (64, 40, 32, 3000, 3.0, (6, 4, 2), False),     # sparse: every datapoint stays in list form
        (32, 96, 24, 2000, 24.0, (6, 4, 2), True),     # around the capacity: a mix of both kernels
        (32, 48, 16, 500, 34.0, (4, 3, 2), True),      # dense: (nearly) every datapoint overflows
        (256, 300, 64, 2000, 5.0, (8, 8, 2), None),    # cfg2 shape (a handful of initial states overflow)
        (40, 1030, 32, 300, 4.0, (8, 4, 2), False),    # H > 1024 (33 words per packed state)
        (16, 2100, 16, 150, 4.0, (4, 3, 2), False),    # more than 64 words per packed state: warp-scan load
    ],
)
def test_list_kernel_equals_bits_kernel(D, H, K, N, act, ea, expect_overflow):
    out = _run_both(D, H, K, N, act, ea)
    sl, ll, Fl, ovf = out[False]
    sb, lb, Fb, _ = out[True]
    assert np.array_equal(sl, sb)
    # children are evaluated incrementally from their parent's sums in the list kernel, in full in the packed-bits
    # kernel: same value up to rounding (datapoints the bits kernel redid are bit-identical)
    assert np.max(np.abs(ll - lb)) <= 1e-12 * np.max(np.abs(lb))
    assert np.max(np.abs(Fl - Fb) / np.abs(Fb)) < 1e-13  # F is summed with unordered f64 atomics
    if expect_overflow is not None:
        assert (ovf > 0) == expect_overflow, ovf
    assert ovf <= N
    if act == 24.0:
        assert 0 < ovf < N  # both kernels took part
