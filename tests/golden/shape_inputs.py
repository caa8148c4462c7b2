"""Synthetic inputs of the BASELINE-shape fixtures (tests/golden/shape_*.npz), regenerated from a legacy
`np.random.RandomState` seed (bit-stable across NumPy versions) by the generator (make_golden_shapes.py) AND by
the tests, so the large arrays (W0: 1024 x 1024 at cfg5) need not be committed; the fixtures pin them by SHA-1.
Test infrastructure; imports nothing but NumPy."""
import hashlib

import numpy as np


def f32(a):
    return np.asarray(a, dtype=np.float64).astype(np.float32).astype(np.float64)


def sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def shape_inputs(model, seed, D, H, N, act, missing=0.0):
    """Synthetic inputs of SURVEY 8(d): W_gen ~ N(0,1), s ~ Bernoulli(act/H), (z ~ N(1,1) for ES3C),
    y = W_gen (s o z) + N(0,1); theta_0 = W_gen + 0.5 N(0,1) so that the K sets meet well- and ill-fitting
    states; everything float32-representable.  Used by this generator AND by the tests (same seed -> same
    arrays; the fixture stores SHA-1 digests)."""
    rng = np.random.RandomState(seed)
    Wg = rng.normal(size=(D, H))
    s = rng.random_sample((N, H)) < act / H
    if model == "es3c":
        z = rng.normal(loc=1.0, size=(N, H))
        Y = (s * z) @ Wg.T + rng.normal(size=(N, D))
    else:
        Y = s.astype(np.float64) @ Wg.T + rng.normal(size=(N, D))
    Y = f32(Y)
    W0 = f32(Wg + 0.5 * rng.normal(size=(D, H)))
    X = np.ones((N, D), dtype=bool)
    if missing > 0:
        X = rng.random_sample((N, D)) >= missing
    out = {"Y": Y, "X": X, "W0": W0}
    if model == "es3c":
        A = rng.normal(size=(H, 8)) * 0.15
        out["Psi0"] = np.eye(H) + A @ A.T  # symmetric positive definite, dense
        out["mus0"] = rng.normal(size=H) * 0.3 + 1.0
        out["pies0"] = rng.uniform(0.5, 2.0, size=H) * act / H
    return out
