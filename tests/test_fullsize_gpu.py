"""Size-independent properties of the E-step at BASELINE.json's full sizes (the NumPy oracle cannot
reach them): after `E_step` in RNG mode on a config-sized shard

  * every K^n holds S pairwise distinct states (the de-duplication held for every datapoint),
  * the stored lpj row equals the stand-alone log-joint kernel re-evaluated on the stored states under
    the same parameters (the fused kernel's bookkeeping of which value belongs to which slot),
  * F equals ljc + mean_n logsumexp of those rows (`evo_free_energy`),
  * a second E-step under the same parameters never lowers F (top-K replacement is monotone),
  * the dense tensor-core log-joint agrees with the stored values (BSC).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _distinct_per_datapoint(ss_dev):
    """[N][S][Hw] int32 -> True if no datapoint holds two equal states (two independent 64-bit hashes)."""
    import torch

    x = ss_dev.to(torch.int64) & 0xFFFFFFFF
    Hw = x.shape[-1]
    g = torch.Generator(device="cpu").manual_seed(5)
    ok = True
    for _ in range(2):
        mult = torch.randint(1, 2 ** 62, (Hw,), generator=g, dtype=torch.int64).to(x.device) | 1
        h = (x * mult).sum(-1)  # wraps mod 2^64
        h = h ^ (h >> 29)
        hs, _ = torch.sort(h, dim=1)
        ok = ok and not bool((hs[:, 1:] == hs[:, :-1]).any().item())
    return ok


def _synth(D, H, N, act, es3c, device, seed=0):
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(seed)
    Wg = torch.randn((D, H), generator=g, device=device)
    s = (torch.rand((N, H), generator=g, device=device) < act / H).float()
    if es3c:
        s = s * (1.0 + torch.randn((N, H), generator=g, device=device))
    return s @ Wg.T + torch.randn((N, D), generator=g, device=device)


@pytest.mark.parametrize("D,H,K,N,ea", [(256, 300, 64, 100000, (8, 8, 2)), (1024, 1024, 128, 20000, (16, 8, 2))])
def test_bsc_estep_properties_at_config_size(D, H, K, N, ea):
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    m = BSC(D, H, K)
    Y = _synth(D, H, N, 5.0, False, m.device)
    data = {"y": Y, "x_infr": None}
    np.random.seed(1)
    th = m.check_params(m.standard_init(data))
    cont = init_states(N, K, H, "fit", "randflip", *ea, device=m.device, seed=9)
    for _ in range(2):  # move away from the random initial K sets
        _, _, _, th = m.step(th, cont, data)
    F1, nu, nsub = m.E_step(th, cont, data)
    assert _distinct_per_datapoint(cont["ss_dev"])
    lpj = cont["lpj_dev"]
    re = m.lpj_all(th, data, cont["ss_dev"])
    assert float(((lpj - re).abs() / re.abs()).max().item()) < 1e-12
    Fs = m._sum_logsumexp(lpj)
    assert abs(F1 - (th["ljc"] + float(Fs.item()) / N)) < 1e-9 * abs(F1)
    dense = m.lpj_all_dense(th, data, cont["ss_dev"])
    assert float(((dense - lpj).abs() / lpj.abs()).max().item()) < 1e-5
    F2, _, _ = m.E_step(th, cont, data)
    assert F2 >= F1 - 1e-9 * abs(F1)
    assert _distinct_per_datapoint(cont["ss_dev"])


@pytest.mark.parametrize("missing", [0.0, 0.3])
def test_es3c_estep_properties_at_config_shape(missing):
    import torch

    from evo_b200.sssc import SSSC
    from evo_b200.variational import init_states

    D, H, K, N = 144, 512, 64, 20000
    learn = ("W", "pies", "sigma2") if missing else ("W", "pies", "mus", "sigma2", "Psi")
    m = SSSC(D, H, K, to_learn=list(learn))
    Y = _synth(D, H, N, 4.0, True, m.device)
    X = None
    data = {"y": Y, "x_infr": None}
    if missing:
        g = torch.Generator(device=m.device)
        g.manual_seed(3)
        X = torch.rand((N, D), generator=g, device=m.device) >= missing
        data = {"y": Y, "x_infr": X, "x": X}
    np.random.seed(1)
    th = m.check_params(m.standard_init(data))
    cont = init_states(N, K, H, "fit", "randflip", 8, 8, 2, device=m.device, seed=9)
    F0, _, _, th = m.step(th, cont, data, do_reconstruction=True)
    F1, _, _, th = m.step(th, cont, data, do_reconstruction=True)
    assert np.isfinite(F0) and np.isfinite(F1)
    assert _distinct_per_datapoint(cont["ss_dev"])
    # stored lpj rows belong to the parameters BEFORE the last M-step; re-evaluate under the new ones and
    # check the packed (8-lane) and the warp-per-state evaluators through the stand-alone kernel
    re = m.lpj_all(th, data, cont["ss_dev"])
    assert bool(torch.isfinite(re).all().item())
    re2 = m.lpj_all(th, data, cont["ss_dev"][:, :3].contiguous())  # different batch split, same values
    assert float(((re[:, :3] - re2).abs() / re2.abs()).max().item()) < 1e-12
