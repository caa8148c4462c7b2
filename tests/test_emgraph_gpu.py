"""Whole EM iterations as one replayed CUDA graph (evo_b200/emgraph.py, `evo_bsc_em_advance`, `evo_estep_args.theta_dev`)
against the same number of `BSC.step` calls: same kernels and Philox streams, so the K sets must be bit-identical; F and
the parameters may differ by the last ulp of `log` (device vs NumPy) amplified through a few iterations: F 1e-10 over 4 iterations,
W 1e-6 (tensor-core Wp, stated 1e-5 per iteration)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _setup(D, H, K, N, act, ea, mutation="randflip", seed=3):
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.variational import init_states

    rng = np.random.RandomState(seed)
    Wg = rng.normal(size=(D, H)) * 2.0
    s = rng.random_sample((N, H)) < act / H
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)
    m = BSC(D, H, K)
    m.tables = m.stats = "tc"
    data = {"y": torch.as_tensor(Y, device=m.device), "x_infr": None}
    np.random.seed(1)
    th = m.check_params(m.standard_init(data))
    cont = init_states(N, K, H, "fit", mutation, *ea, bitflip_prob=1.0 / H if mutation == "sparseflip" else None,
                       p_init_Kn=act / H, device=m.device, seed=11)
    return m, th, cont, data


@pytest.mark.parametrize("D,H,K,N,act,ea,mutation", [
    (25, 10, 16, 1000, 2.0, (4, 3, 2), "randflip"),      # cfg1-like
    (64, 40, 32, 3000, 3.0, (6, 4, 2), "randflip"),
    (48, 300, 64, 4000, 5.0, (8, 8, 2), "randflip"),     # cfg2-like shape
    (32, 48, 24, 2000, 3.0, (5, 3, 2), "sparseflip"),    # generic kernel: sparseness from device memory
])
def test_graph_replay_equals_step_by_step(D, H, K, N, act, ea, mutation):
    n_it = 4  # the last-ulp differences grow ~ 40x per iteration (order of the f64 atomics, log on device vs NumPy)
    m, th, cont, data = _setup(D, H, K, N, act, ea, mutation)
    Fs = []
    for _ in range(n_it):
        F, nun, nsub, th = m.step(th, cont, data)
        Fs.append(F)
    ss_ref = cont["ss_dev"].cpu().numpy()

    m2, th2, cont2, data2 = _setup(D, H, K, N, act, ea, mutation)
    g = m2.em_graph(th2, cont2, data2)
    F1, _, _ = g.run(2)   # 1 eager + capture + 1 replay
    F2, nun2, nsub2 = g.run(n_it - 2)
    out = g.theta_out()
    assert int(cont2["iteration"]) == int(cont["iteration"]) == n_it
    assert np.array_equal(cont2["ss_dev"].cpu().numpy(), ss_ref)
    Fg = np.array(F1 + F2)
    relF = np.abs(Fg - np.array(Fs)) / np.abs(np.array(Fs))
    assert relF.max() < 1e-10
    assert abs(nun2[-1] - nun) < 1e-12 and abs(nsub2[-1] - nsub) < 1e-12
    # Wp goes through the fp16 hi + lo planes of <s> (2^-22 per entry): a last-ulp change of <s> can move an entry by that much
    assert np.abs(np.asarray(out["W"]) - np.asarray(th["W"])).max() < 1e-6 * np.abs(np.asarray(th["W"])).max()
    assert abs(out["pi"] - th["pi"]) < 1e-9 * th["pi"] and abs(out["sigma"] - th["sigma"]) < 1e-8 * th["sigma"]
    # and the step-by-step path continues from the graph's result
    F_next, _, _, _ = m2.step(out, cont2, data2)
    F_ref, _, _, _ = m.step(th, cont, data)
    assert abs(F_next - F_ref) < 1e-8 * abs(F_ref)
    assert np.array_equal(cont2["ss_dev"].cpu().numpy(), cont["ss_dev"].cpu().numpy())


def test_graph_rejects_what_it_does_not_cover():
    m, th, cont, data = _setup(25, 10, 16, 200, 2.0, (4, 3, 2))
    m.fuse_stats = False
    with pytest.raises(ValueError):
        m.em_graph(th, cont, data)
