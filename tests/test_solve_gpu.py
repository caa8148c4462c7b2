"""The hand-written replicated parameter solve (csrc/solve.cu: block Gaussian elimination with 32 x 32 inverted
diagonal blocks, f64 SIMT GEMMs) against NumPy's LAPACK path the reference uses (bsc.py:233-235 lstsq,
sssc.py:693 inv): `evo_bsc_solve` and `evo_es3c_update` called directly through the C-ABI on synthetic
statistics.  Tolerance 1e-9 x cond-free scale for well-conditioned systems (f64 throughout)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _spd(rng, H, cond=1e3):
    Q, _ = np.linalg.qr(rng.normal(size=(H, H)))
    ev = np.exp(rng.uniform(0, np.log(cond), size=H))
    return (Q * ev) @ Q.T


@pytest.mark.parametrize("D,H", [(25, 10), (40, 32), (70, 33), (256, 300), (144, 512), (300, 1024)])
def test_bsc_solve_matches_lstsq(D, H):
    import torch

    from evo_b200 import _lib

    lib = _lib.load()
    dev = torch.device("cuda")
    rng = np.random.RandomState(H)
    N = 5000.0
    Wq = _spd(rng, H) * 50
    Wp = rng.normal(size=(H, D)) * 30
    pies = rng.uniform(1, 100, size=H)
    sig = 12345.678
    stats = np.concatenate([Wp.ravel(), np.tril(Wq).ravel(), pies, [sig]])
    sd = torch.as_tensor(stats, device=dev)
    ws = torch.empty(int(lib.evo_bsc_solve_workspace_bytes(D, H)), dtype=torch.uint8, device=dev)
    W = torch.zeros((D, H), dtype=torch.float64, device=dev)
    p = torch.zeros(H, dtype=torch.float64, device=dev)
    scal = torch.zeros(4, dtype=torch.float64, device=dev)
    for incomplete in (False, True):
        _lib.check(lib.evo_bsc_solve(_lib.ptr(sd), N, D, H, 1, 0, 777.0 if incomplete else -1.0, 0.9, _lib.ptr(W), _lib.ptr(p),
                                     _lib.ptr(scal), _lib.ptr(ws), _lib.stream()), "evo_bsc_solve")
        want = np.linalg.lstsq(Wq, Wp, rcond=-1)[0].T
        got, sc = W.cpu().numpy(), scal.cpu().numpy()
        assert sc[0] == 0 and sc[3] == 1
        assert np.abs(got - want).max() <= 1e-9 * np.abs(want).max()
        assert np.allclose(p.cpu().numpy(), pies / N, rtol=1e-14)
        assert abs(sc[1] - (pies / N).sum() / H) <= 1e-14
        ref_sigma = np.sqrt((sig + 777.0 * 0.81) / N / D) if incomplete else np.sqrt(sig / N / D)
        assert abs(sc[2] - ref_sigma) <= 1e-14 * ref_sigma
    # background unit + W not learned
    W.fill_(7.0)
    _lib.check(lib.evo_bsc_solve(_lib.ptr(sd), N, D, H, 0, 1, -1.0, 0.9, None, _lib.ptr(p), _lib.ptr(scal), _lib.ptr(ws),
                                 _lib.stream()), "evo_bsc_solve")
    assert abs(p.cpu().numpy()[-1] - (1.0 - 1.1e-5)) < 1e-16 and (W == 7.0).all()


def test_bsc_solve_flags_singular_statistics_and_the_model_falls_back_to_min_norm():
    """A latent without posterior mass makes Wq singular: `info` != 0, and BSC's M-step then returns the reference's
    minimum-norm lstsq solution (bsc.py:233-235) through the device pseudo-inverse."""
    import torch

    from evo_b200 import _lib
    from evo_b200.bsc import BSC
    from evo_b200.model import ACC_LEN

    lib = _lib.load()
    D, H = 20, 40
    rng = np.random.RandomState(1)
    Wq = _spd(rng, H)
    Wp = rng.normal(size=(H, D))
    dead = [3, 17]
    Wq[dead, :] = 0
    Wq[:, dead] = 0
    Wp[dead, :] = 0
    m = BSC(D, H, 8)
    stats = torch.as_tensor(np.concatenate([Wp.ravel(), np.tril(Wq).ravel(), np.ones(H), [10.0], np.zeros(ACC_LEN)]),
                            device=m.device)

    class DD:
        N_total, n_obs_total, incomplete = 100, 0, False
    th = {"W": np.zeros((D, H)), "pi": 0.1, "sigma": 1.0}
    pend = m._solve_enqueue(th, {"permanent": {"background": False}}, DD, stats)
    scal = pend["scal"].cpu().numpy()
    assert scal[0] != 0
    th = m._solve_finish(th, {}, DD, pend, scal)
    want = np.linalg.lstsq(Wq, Wp, rcond=-1)[0].T
    assert np.abs(th["W"] - want).max() <= 1e-8 * np.abs(want).max()
    assert (np.abs(th["W"][:, dead]) < 1e-12).all()


@pytest.mark.parametrize("D,H,incomplete,learn", [(12, 10, False, "W,pies,mus,Psi,sigma2"), (20, 45, True, "W,pies,sigma2"),
                                                  (144, 512, False, "W,pies,mus,Psi,sigma2"), (33, 64, False, "pies,mus")])
def test_es3c_update_matches_numpy(D, H, incomplete, learn):
    import torch

    from evo_b200 import _lib

    lib = _lib.load()
    dev = torch.device("cuda")
    rng = np.random.RandomState(D + H)
    N, H2 = 4000.0, H * H
    learn = learn.split(",")
    sum_s = rng.uniform(0.5, 200, size=H)
    sum_s[0] = 1e-3  # a nearly dead latent: clipped pies
    sum_sz = rng.normal(size=H) * 50
    sum_ss = _spd(rng, H) * 20
    sum_szsz = _spd(rng, H) * 40 + rng.normal(size=(H, H)) * 0.05  # not symmetric (the element-wise Psi update breaks symmetry)
    sum_s_sz = rng.normal(size=(H, H))
    sum_outer = rng.normal(size=(H, H))
    WpT = rng.normal(size=(H, D)) * 10
    trace = 4321.0
    stats = np.concatenate([sum_s, sum_sz, sum_ss.ravel(), sum_szsz.ravel(), sum_s_sz.ravel(), sum_outer.ravel(), WpT.ravel(), [trace]])
    old = {"W": rng.normal(size=(D, H)), "pies": rng.uniform(0.1, 0.3, size=H), "mus": rng.normal(size=H), "Psi": _spd(rng, H)}
    dv = {k: torch.as_tensor(v, device=dev).clone() for k, v in old.items()}
    scal = torch.zeros(4, dtype=torch.float64, device=dev)
    sd = torch.as_tensor(stats, device=dev)
    ws = torch.empty(int(lib.evo_es3c_update_workspace_bytes(D, H)), dtype=torch.uint8, device=dev)
    a = _lib.Es3cUpdateArgs()
    a.D, a.H = D, H
    a.learn_W, a.learn_pies, a.learn_mus = int("W" in learn), int("pies" in learn), int("mus" in learn)
    a.learn_Psi, a.learn_sigma2, a.background, a.incomplete = int("Psi" in learn), int("sigma2" in learn), 0, int(incomplete)
    ysq, s2old, nobs = 9.9e5, 1.7, 31000.0
    a.N_total, a.n_obs_total, a.ysq, a.sigma2_old = N, nobs, ysq, s2old
    a.eps_pies, a.eps_mus, a.eps_Psi, a.eps_sigma2 = 5e-5, np.finfo(np.float64).eps, 1e-5, 1e-5
    a.stats, a.scal, a.workspace = sd.data_ptr(), scal.data_ptr(), ws.data_ptr()
    a.W, a.pies, a.mus, a.Psi = dv["W"].data_ptr(), dv["pies"].data_ptr(), dv["mus"].data_ptr(), dv["Psi"].data_ptr()
    _lib.check(lib.evo_es3c_update(C.byref(a), _lib.stream()), "evo_es3c_update")
    got = {k: v.cpu().numpy() for k, v in dv.items()}
    sc = scal.cpu().numpy()
    assert sc[0] == 0 and sc[1] == 0 and sc[3] == 1
    # the reference's updates (sssc.py:688-770), NumPy
    W_new = WpT.T @ np.linalg.inv(sum_szsz) if "W" in learn else old["W"]
    pies = np.clip(sum_s / N, 5e-5, 1 - 5e-5) if "pies" in learn else old["pies"]
    mus = sum_sz / (sum_s + np.finfo(np.float64).eps) if "mus" in learn else old["mus"]
    Psi = old["Psi"]
    if "Psi" in learn:
        Psi = (np.outer(mus, mus) * sum_ss + sum_szsz - 2 * mus[:, None] * sum_s_sz) * np.linalg.inv(sum_ss + 1e-5 * np.eye(H))
    for k, want in (("W", W_new), ("pies", pies), ("mus", mus), ("Psi", Psi)):
        assert np.abs(got[k] - want).max() <= 1e-9 * np.abs(want).max(), k
    if "sigma2" in learn:
        if incomplete:
            s2 = ((ysq - trace) + nobs * s2old) / N / D + 1e-5
        else:
            s2 = (ysq - np.trace(sum_outer @ (W_new.T @ W_new))) / N / D + 1e-5
        assert abs(sc[2] - s2) <= 1e-9 * abs(s2)
    else:
        assert sc[2] == s2old


def test_es3c_singular_moments_take_the_reference_fallback():
    """A latent that is active in no state leaves a zero row / column in sum <sz sz'>: `evo_es3c_update` flags it and
    SSSC takes the reference's route (sssc.py:695-709: pseudo-inverse plus rank-one noise, drawn on rank 0 and
    broadcast) on the device -- finite parameters, the dead latent's column of W stays at the noise level."""
    from evo_b200.sssc import SSSC
    from evo_b200.variational import init_states

    rng = np.random.RandomState(0)
    D, H, N, S = 8, 6, 200, 4
    Y = rng.normal(size=(N, D)).astype(np.float32).astype(np.float64)
    m = SSSC(D, H, S)
    cont = init_states(N, S, H, "fit", "randflip", 2, 1, 0)  # no generations: the K sets stay as given
    ss = np.zeros((N, S, H), dtype=bool)
    for k in range(S):
        ss[:, k, k] = True  # states use latents 0..3 only; 4 and 5 never appear
    ss[:, 0, 1] = True
    cont["ss"] = ss
    th = {"W": rng.normal(size=(D, H)), "pies": np.full(H, 0.2), "mus": np.ones(H), "Psi": np.eye(H), "sigma2": np.float64(1.0)}
    np.random.seed(0)
    F, _, _, th = m.step(th, cont, {"y": Y, "x_infr": None, "x": np.zeros((N, D), bool)}, do_reconstruction=True)
    assert m.last_counts.get("solve_fallback", 0) == 1
    assert np.isfinite(F) and all(np.isfinite(np.asarray(th[k])).all() for k in ("W", "pies", "mus", "Psi", "sigma2"))
    assert np.abs(th["pies"][4:] - 5e-5).max() < 1e-12  # clipped at eps
