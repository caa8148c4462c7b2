"""Overlapping-patch pipeline of the image drivers on the device: the classes and
merger names the reference imports from `tvutil.prepost` (examples/image-denoising/main.py:17-22,85-88,
169-177; examples/image-inpainting/main.py:86-88,174-182), backed by `csrc/patches.cu`.

    ovp = OverlappingPatches(noisy, ph, pw, patch_shift=1)
    Y = ovp.get().T                                   # (N, D) float32 on the device, contiguous
    img = ovp.set_and_merge(Y_rec_T, merge_method=mean_merger)

tvutil is not vendored with the reference: the patch order and flattening are those stated in
csrc/patches.cu (parity unpinned)."""
import numpy as np
import torch

from . import _lib


class _Merger:
    def __init__(self, name, code):
        self.name, self.code = name, code

    def __repr__(self):
        return "<merger %s>" % self.name


mean_merger = _Merger("mean", 0)
median_merger = _Merger("median", 1)


class MultiDimOverlappingPatches:
    """image: (Hi, Wi, C) like the reference's RGB arrays (main.py:81,85), or (Hi, Wi)."""

    def __init__(self, image, patch_height, patch_width=None, patch_shift=1, device=None):
        self.lib = _lib.load()
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        img = torch.as_tensor(np.asarray(image) if not isinstance(image, torch.Tensor) else image)
        self.channels_last = img.ndim == 3
        if self.channels_last:
            img = img.permute(2, 0, 1)
        else:
            img = img[None]
        self.image = img.to(self.device, torch.float32).contiguous()
        self.C, self.Hi, self.Wi = (int(v) for v in self.image.shape)
        self.ph = int(patch_height)
        self.pw = int(patch_width) if patch_width is not None else self.ph
        self.shift = int(patch_shift)
        self.N = int(self.lib.evo_patches_count(self.Hi, self.Wi, self.ph, self.pw, self.shift))
        if self.N <= 0:
            raise ValueError("patch %dx%d does not fit image %dx%d" % (self.ph, self.pw, self.Hi, self.Wi))
        self.D = self.C * self.ph * self.pw
        self._patches = None

    def _geom(self):
        return self.C, self.Hi, self.Wi, self.ph, self.pw, self.shift

    def get(self):
        """(D, N) view of the (N, D) patch matrix: `get().T` is contiguous (N, D), as the drivers use it."""
        if self._patches is None:
            out = torch.empty((self.N, self.D), dtype=torch.float32, device=self.device)
            with torch.cuda.device(self.device):
                _lib.check(self.lib.evo_patches_extract(_lib.ptr(self.image), *self._geom(), _lib.ptr(out), _lib.stream()),
                           "evo_patches_extract")
            self._patches = out
        return self._patches.T

    def set(self, new_patches):
        p = new_patches if isinstance(new_patches, torch.Tensor) else torch.as_tensor(np.asarray(new_patches))
        if tuple(p.shape) != (self.D, self.N):
            raise ValueError("expected (D, N) = (%d, %d), got %s" % (self.D, self.N, tuple(p.shape)))
        self._patches = p.T.to(self.device, torch.float32).contiguous()

    def merge(self, merge_method=mean_merger):
        if self._patches is None:
            self.get()
        out = torch.empty((self.C, self.Hi, self.Wi), dtype=torch.float32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(self.lib.evo_patches_merge(_lib.ptr(self._patches), *self._geom(), int(merge_method.code),
                                                  _lib.ptr(out), _lib.stream()), "evo_patches_merge")
        return out.permute(1, 2, 0) if self.channels_last else out[0]

    def set_and_merge(self, new_patches, merge_method=mean_merger):
        self.set(new_patches)
        return self.merge(merge_method)


class OverlappingPatches(MultiDimOverlappingPatches):
    """Single-channel images (Hi, Wi)."""


def psnr(target, reco, data_range=255.0):
    """10 log10(R^2 / mse) (examples/image-denoising/utils.py:8-18) with the squared error summed on the device."""
    lib = _lib.load()
    r = reco if isinstance(reco, torch.Tensor) else torch.as_tensor(np.asarray(reco))
    dev = r.device if r.is_cuda else torch.device("cuda", torch.cuda.current_device())
    r = r.to(dev, torch.float32).contiguous()
    t = (target if isinstance(target, torch.Tensor) else torch.as_tensor(np.asarray(target))).to(dev, torch.float32).contiguous()
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.evo_sum_sqdiff(_lib.ptr(t), _lib.ptr(r), t.numel(), _lib.ptr(out), _lib.stream()), "evo_sum_sqdiff")
    mse = float(out.item()) / t.numel()
    return float(10.0 * np.log10(data_range ** 2 / mse))
