"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py) -- CPU restatement of the overlapping-patch
pipeline the reference's image drivers call either side of the model:
`OverlappingPatches(image, ph, pw, patch_shift=1).get()`, `.set_and_merge(Y_rec_T, merge_method=mean_merger |
median_merger)` and the PSNR of `examples/image-denoising/utils.py:8-18`.

PARITY UNPINNED: the implementation is tvutil.prepost (github.com/tvlearn/tvutil, un-vendored and
unpinned, reference README.md:33) and scikit-image's `compare_psnr`; neither is installed and the
reference holds no test or golden vector for them.  This file restates the semantics visible at the call
sites (examples/image-denoising/main.py:85-88,130,169-177, examples/image-inpainting/main.py:86-88,174-182):
patches are the columns of `get()` (so `get().T` is (N, D)), one per top-left position in row-major
order, each flattened channel-slowest / row-major; merging reduces, per pixel, the estimates of all
covering patches with nan-mean or nan-median; PSNR = 10 log10(255^2 / mse)."""
import numpy as np


def _geom(shape, ph, pw, shift):
    C, Hi, Wi = shape
    return (Hi - ph) // shift + 1, (Wi - pw) // shift + 1


def extract(image, ph, pw, shift=1):
    """image (Hi, Wi) or (C, Hi, Wi) -> (N, D) float32, D = C*ph*pw."""
    img = np.asarray(image, dtype=np.float32)
    if img.ndim == 2:
        img = img[None]
    C, Hi, Wi = img.shape
    ny, nx = _geom(img.shape, ph, pw, shift)
    out = np.empty((ny * nx, C * ph * pw), dtype=np.float32)
    for ky in range(ny):
        for kx in range(nx):
            y, x = ky * shift, kx * shift
            out[ky * nx + kx] = img[:, y:y + ph, x:x + pw].reshape(-1)
    return out


def merge(patches, shape, ph, pw, shift=1, method="mean"):
    """(N, D) patches -> image (C, Hi, Wi) float32; pixels no patch covers are NaN."""
    C, Hi, Wi = shape
    ny, nx = _geom(shape, ph, pw, shift)
    P = np.asarray(patches, dtype=np.float32).reshape(ny, nx, C, ph, pw)
    cover = -(-ph // shift) * -(-pw // shift)
    est = np.full((C, Hi, Wi, cover), np.nan, dtype=np.float32)
    cnt = np.zeros((Hi, Wi), dtype=np.int64)
    for ky in range(ny):
        for kx in range(nx):
            y, x = ky * shift, kx * shift
            c = cnt[y:y + ph, x:x + pw]
            yy, xx = np.meshgrid(np.arange(y, y + ph), np.arange(x, x + pw), indexing="ij")
            est[:, yy, xx, c] = P[ky, kx]
            cnt[y:y + ph, x:x + pw] += 1
    import warnings

    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        if method == "mean":
            return np.nanmean(est.astype(np.float64), axis=-1).astype(np.float32)
        return np.nanmedian(est, axis=-1).astype(np.float32)


def psnr(target, reco, data_range=255.0):
    t, r = np.asarray(target, dtype=np.float64), np.asarray(reco, dtype=np.float64)
    return float(10.0 * np.log10(data_range ** 2 / np.mean((t - r) ** 2)))
