"""Overlapping-patch pipeline.  CPU: properties of the NumPy restatement
(oracle/patches.py; parity unpinned -- tvutil is not vendored with the reference).  GPU: the CUDA kernels
behind `evo_b200.patches` against it: extraction bit-exact, median bit-exact, mean within 1e-6 relative
(f64 accumulation, f32 result), PSNR within 1e-9, plus a full-size extract -> merge round trip."""
import numpy as np
import pytest

from oracle import patches as OP


def _img(shape, seed):
    rng = np.random.RandomState(seed)
    return (rng.random_sample(shape) * 255).astype(np.float32)


@pytest.mark.parametrize("shape,ph,pw,shift", [((1, 9, 11), 3, 4, 1), ((3, 8, 8), 2, 2, 1), ((1, 10, 10), 4, 4, 2), ((1, 11, 10), 4, 4, 3)])
def test_oracle_roundtrip_and_layout(shape, ph, pw, shift):
    img = _img(shape, 1)
    P = OP.extract(img, ph, pw, shift)
    ny, nx = (shape[1] - ph) // shift + 1, (shape[2] - pw) // shift + 1
    assert P.shape == (ny * nx, shape[0] * ph * pw)
    # patch (ky, kx), channel c, row r, column q
    assert P[1 * nx + 1, (shape[0] - 1) * ph * pw + 1 * pw + 0] == img[shape[0] - 1, shift + 1, shift]
    for method in ("mean", "median"):
        rec = OP.merge(P, shape, ph, pw, shift, method)
        cov = ~np.isnan(rec)
        assert np.allclose(rec[cov], img[cov], rtol=1e-6)  # merging untouched patches returns the image
        assert cov[:, : (ny - 1) * shift + ph, : (nx - 1) * shift + pw].all()
    assert OP.psnr(img, img + 1.0) == pytest.approx(10 * np.log10(255.0 ** 2), rel=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("shape,ph,pw,shift", [((1, 40, 37), 12, 12, 1), ((3, 21, 30), 5, 6, 1), ((1, 33, 33), 8, 8, 2),
                                               ((1, 30, 31), 7, 5, 3), ((1, 16, 16), 16, 16, 1)])
def test_patch_kernels_match_oracle(shape, ph, pw, shift):
    import torch

    from evo_b200.patches import MultiDimOverlappingPatches, OverlappingPatches, mean_merger, median_merger, psnr

    img = _img(shape, 7)
    if shape[0] == 1:
        ovp = OverlappingPatches(img[0], ph, pw, patch_shift=shift)
    else:
        ovp = MultiDimOverlappingPatches(np.ascontiguousarray(img.transpose(1, 2, 0)), ph, pw, patch_shift=shift)
    Y = ovp.get().T
    assert Y.is_contiguous()
    ref = OP.extract(img, ph, pw, shift)
    assert np.array_equal(Y.cpu().numpy(), ref)
    rng = np.random.RandomState(3)
    noisy = (ref + rng.normal(size=ref.shape) * 5).astype(np.float32)  # "reconstructed" patches disagree per pixel
    for merger, name in ((mean_merger, "mean"), (median_merger, "median")):
        got = ovp.set_and_merge(torch.as_tensor(noisy.T.copy()), merge_method=merger).cpu().numpy()
        if shape[0] > 1:
            got = got.transpose(2, 0, 1)
        else:
            got = got[None]
        want = OP.merge(noisy, shape, ph, pw, shift, name)
        assert np.array_equal(np.isnan(got), np.isnan(want))
        ok = ~np.isnan(want)
        if name == "median":
            assert np.array_equal(got[ok], want[ok])
        else:
            assert np.max(np.abs(got[ok] - want[ok]) / np.abs(want[ok])) < 1e-6
    assert psnr(img, img + 3.0) == pytest.approx(OP.psnr(img, img + 3.0), rel=1e-9)


@pytest.mark.gpu
def test_patch_roundtrip_at_denoising_size():
    """Size-independent property at the denoising driver's size (256 x 256 image, 12 x 12 patches, shift 1:
    60k patches of D = 144): extracting and merging untouched patches returns the image, for both mergers."""
    import torch

    from evo_b200.patches import OverlappingPatches, mean_merger, median_merger

    img = _img((256, 256), 11)
    ovp = OverlappingPatches(img, 12, 12, patch_shift=1)
    Y = ovp.get().T
    assert tuple(Y.shape) == (245 * 245, 144)
    for merger in (mean_merger, median_merger):
        rec = ovp.set_and_merge(Y.T, merge_method=merger).cpu().numpy()
        assert np.array_equal(rec, img) if merger is median_merger else np.allclose(rec, img, rtol=1e-6)
