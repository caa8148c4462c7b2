"""Checkpoint / resume (the reference cannot resume): a run interrupted after 2 of 4 EM
iterations, dumped, re-loaded into fresh objects and continued reproduces the uninterrupted run bit for bit
-- K sets, lpj, F and parameters -- because the Philox streams are keyed by (seed, global datapoint index,
iteration), all of which travel in the checkpoint."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model_name", ["bsc", "es3c"])
def test_resume_reproduces_uninterrupted_run(tmp_path, model_name):
    import torch

    from evo_b200.bsc import BSC
    from evo_b200.checkpoint import NpzLog, load_checkpoint, save_checkpoint
    from evo_b200.sssc import SSSC
    from evo_b200.variational import init_states

    D, H, K, N = 32, 24, 16, 1500
    rng = np.random.RandomState(2)
    Wg = rng.normal(size=(D, H)) * 2.0
    s = (rng.random_sample((N, H)) < 3.0 / H) * (1.0 + (model_name == "es3c") * rng.normal(size=(N, H)))
    Y = (s @ Wg.T + rng.normal(size=(N, D))).astype(np.float32)

    def fresh():
        m = BSC(D, H, K) if model_name == "bsc" else SSSC(D, H, K)
        return m, {"y": torch.as_tensor(Y, device=m.device), "x_infr": None}

    def run(m, data, th, cont, iters, log=None):
        Fs = []
        for _ in range(iters):
            F, nu, ns, th = m.step(th, cont, data, do_reconstruction=(model_name == "es3c"))
            Fs.append(F)
            if log is not None:
                log.append_all({"F": F, "S_nunique": nu, "S_sub": ns})
        return th, Fs

    # uninterrupted
    m, data = fresh()
    np.random.seed(1)
    th = m.check_params(m.standard_init(data))
    cont = init_states(N, K, H, "fit", "randflip", 5, 3, 2, device=m.device, seed=21)
    th_a, F_a = run(m, data, th, cont, 4)
    ss_a, lpj_a = cont["ss_dev"].cpu().numpy(), cont["lpj_dev"].cpu().numpy()

    # interrupted after 2, resumed in fresh objects
    m, data = fresh()
    np.random.seed(1)
    th = m.check_params(m.standard_init(data))
    cont = init_states(N, K, H, "fit", "randflip", 5, 3, 2, device=m.device, seed=21)
    log = NpzLog(str(tmp_path / "log.npz"))
    th, F_b = run(m, data, th, cont, 2, log)
    ck = str(tmp_path / "ck.npz")
    save_checkpoint(ck, th, cont, extra={"epoch": 2})
    log.flush()
    del m, cont, th
    m, data = fresh()
    th, cont, extra = load_checkpoint(ck, device=m.device)
    assert int(extra["epoch"]) == 2 and cont["iteration"] == 2
    th_b, F_c = run(m, data, m.check_params(th), cont, 2)

    assert np.array_equal(cont["ss_dev"].cpu().numpy(), ss_a)
    if model_name == "bsc":  # statistics are summed with unordered f64 atomics: F / theta to 1e-12, K sets exact
        assert np.allclose(cont["lpj_dev"].cpu().numpy(), lpj_a, rtol=1e-10)
    assert np.allclose(np.array(F_b + F_c), np.array(F_a), rtol=1e-10)
    Wa = th_a["W"].cpu().numpy() if isinstance(th_a["W"], torch.Tensor) else np.asarray(th_a["W"])
    Wb = th_b["W"].cpu().numpy() if isinstance(th_b["W"], torch.Tensor) else np.asarray(th_b["W"])
    assert np.abs(Wa - Wb).max() < 1e-8 * np.abs(Wa).max()
    with np.load(str(tmp_path / "log.npz")) as f:
        assert f["F"].shape == (2,) and np.allclose(f["F"], F_a[:2], rtol=1e-10)
