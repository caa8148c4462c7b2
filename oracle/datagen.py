"""CPU oracle -- synthetic data generation (TEST INFRASTRUCTURE, not product code).

NumPy restatement of the dataset DEFINITION that `evo_generate_data` (evo_b200/csrc/datagen.cu) implements:
row n of a dataset is a pure function of (seed, global row index) through Philox-4x32-10 counter streams.
It plays the role of the reference's `Model.generate_data` (evo/models/_models.py:73-99 with
`BSC.generate_from_hidden` bsc.py:27-57 / `SSSC.generate_from_hidden` sssc.py:66-102): same generative model
(s_h ~ Bernoulli(pi_h); BSC y = W s + sigma eps; ES3C z_A ~ N(mu_A, Psi_AA), y = W_A z_A + sigma eps), a
counter RNG instead of the reference's unseeded global MT19937 (SURVEY 8(b) "RNG"), so that the GPU arm and
the CPU arm of bench.py -- and every sharding of one dataset -- see the SAME rows.

Parity status: the generative model is pinned distributionally against the reference's generate_data by
tests/golden/make_golden_datagen.py (moments recorded from the unmodified reference); the stream layout has no
counterpart in the reference and is pinned by tests/test_datagen_gpu.py (device == this file: s bit-exact, y to
float32 rounding).
"""
import numpy as np

GEN_S, GEN_Z, GEN_NOISE = 6, 7, 8
GEN_CHOL = 32
_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32(seed, c0, c1, c2, c3):
    """Philox-4x32-10 (Salmon et al. 2011); key = 64-bit seed; counter words are uint32 arrays (broadcast)."""
    c0, c1, c2, c3 = [np.asarray(c, dtype=np.uint64) & _MASK for c in np.broadcast_arrays(c0, c1, c2, c3)]
    k0, k1 = int(seed) & 0xFFFFFFFF, (int(seed) >> 32) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = _M0 * c0, _M1 * c2
        h0, l0, h1, l1 = p0 >> np.uint64(32), p0 & _MASK, p1 >> np.uint64(32), p1 & _MASK
        c0, c1, c2, c3 = h1 ^ c1 ^ np.uint64(k0), l1, h0 ^ c3 ^ np.uint64(k1), l0
        k0, k1 = (k0 + _W0) & 0xFFFFFFFF, (k1 + _W1) & 0xFFFFFFFF
    return c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32)


def _draw(seed, n_global, purpose, sub, idx, iteration=0):
    n_global = np.asarray(n_global, dtype=np.uint64)
    n_lo = n_global & _MASK
    n_hi = ((n_global >> np.uint64(32)) & np.uint64(0xFFFF)) | np.uint64((purpose << 24) | ((sub & 0xFF) << 16))
    return philox4x32(seed, n_lo, n_hi, np.uint64(iteration), np.asarray(idx, dtype=np.uint64))


def _u01(a, b):
    v = ((a.astype(np.uint64) << np.uint64(32)) | b.astype(np.uint64)) >> np.uint64(11)
    return (v.astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)


def _box_muller(r):
    u1, u2 = _u01(r[0], r[1]), _u01(r[2], r[3])
    rad = np.sqrt(-2.0 * np.log(u1))
    return rad * np.cos(2.0 * np.pi * u2), rad * np.sin(2.0 * np.pi * u2)


def generate(W, pies, sigma, rows, seed, mus=None, Psi=None, psi_diag=True):
    """Rows `rows` (global indices) of the dataset defined by (seed, W (D,H), pies scalar or (H,), sigma[, mus,
    Psi]).  Returns {"y" float32-rounded f64 (n,D), "s" bool (n,H), "y_mean" f64, "z" f64 (ES3C)}."""
    W = np.asarray(W, dtype=np.float64)
    D, H = W.shape
    rows = np.asarray(rows, dtype=np.int64)
    n = rows.shape[0]
    p = np.broadcast_to(np.asarray(pies, dtype=np.float64), (H,))
    thr = np.minimum(4294967295.0, np.maximum(p, 0.0) * 4294967296.0).astype(np.uint32)
    q = np.arange((H + 3) // 4)
    r = _draw(seed, rows[:, None], GEN_S, 0, q[None, :])
    bits = np.stack(r, axis=-1).reshape(n, -1)[:, :H]  # component h % 4 of draw h // 4
    s = bits < thr[None, :]
    es3c = mus is not None
    z = s.astype(np.float64)
    if es3c:
        mus, Psi = np.asarray(mus, dtype=np.float64), np.asarray(Psi, dtype=np.float64)
        c, _ = _box_muller(_draw(seed, rows[:, None], GEN_Z, 0, np.arange(H)[None, :]))
        z = np.zeros((n, H))
        for i in range(n):
            A = np.flatnonzero(s[i])
            if A.size == 0:
                continue
            if psi_diag or A.size > GEN_CHOL:
                z[i, A] = mus[A] + np.sqrt(np.maximum(np.diag(Psi)[A], 0.0)) * c[i, A]
            else:
                sub = np.tril(Psi[np.ix_(A, A)])
                L = np.linalg.cholesky(sub + np.tril(sub, -1).T)
                z[i, A] = mus[A] + L @ c[i, A]
    y_mean = np.zeros((n, D))
    for i in range(n):
        acc = np.zeros(D)
        for h in np.flatnonzero(s[i]):  # ascending, one rounding per product and per sum like the kernel
            acc = acc + W[:, h] * z[i, h]
        y_mean[i] = acc
    d = np.arange(D)
    cs, sn = _box_muller(_draw(seed, rows[:, None], GEN_NOISE, 0, (d >> 1)[None, :]))
    eps = np.where((d & 1)[None, :] == 1, sn, cs)
    y = (y_mean + float(sigma) * eps).astype(np.float32).astype(np.float64)
    out = {"y": y, "s": s, "y_mean": y_mean}
    if es3c:
        out["z"] = z
    return out
