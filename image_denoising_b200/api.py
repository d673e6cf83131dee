"""Python host API over the C ABI: batches of BGR uint8 images in, corrupted +
denoised uint8 images out.

``apply``        - device tensors (N,H,W,3) uint8, asynchronous on the current stream;
``corrupt_host`` - NumPy / pinned host arrays through ``b2d_run_host`` (H2D, kernels,
                   D2H pipelined inside the library);
``corrupt``      - one HxWx3 image, the drop-in for ``im = add_<noise>_noise(spec)``
                   (lib/model/test.py:1678-1756).
PyTorch is used for device memory and streams only.
"""
import ctypes
import weakref

import numpy as np
import torch

from . import _lib
from .spec import Spec, parse, resolve, resolve_mix  # noqa: F401

DEFAULT_SEED = 3          # cfg.RNG_SEED, lib/model/config.py:255


def _as_specs(spec, n, overrides):
    if isinstance(spec, (str, Spec)):
        return [parse(spec, **overrides)]
    specs = [parse(s, **overrides) for s in spec]
    if len(specs) not in (1, n):
        raise ValueError("need 1 spec or one per image (%d), got %d" % (n, len(specs)))
    return specs


class Plan:
    """A (device, N, H, W, specs) plan: owns the library handle and its workspace."""

    def __init__(self, n, h, w, spec, device=None, **overrides):
        if not torch.cuda.is_available():
            raise RuntimeError("image_denoising_b200 needs a CUDA device (B200); there is no CPU fallback")
        L = _lib.lib()
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        self.n, self.h, self.w = int(n), int(h), int(w)
        self.specs = _as_specs(spec, self.n, overrides)
        arr = (_lib.B2DSpec * len(self.specs))(*[s.to_c() for s in self.specs])
        handle = ctypes.c_void_p()
        _lib.check(L.b2d_plan_create(ctypes.byref(handle), self.device.index, self.n, self.h, self.w, arr,
                                     len(self.specs)))
        self._h = handle
        self._finalizer = weakref.finalize(self, L.b2d_plan_destroy, handle)
        sz = ctypes.c_size_t()
        _lib.check(L.b2d_plan_workspace_bytes(self._h, ctypes.byref(sz)))
        self.workspace_bytes = sz.value
        self._ws = None

    def _workspace(self):
        if self._ws is None and self.workspace_bytes:
            self._ws = torch.empty(self.workspace_bytes, dtype=torch.uint8, device=self.device)
        return self._ws

    def _check(self, t, name):
        if t.dtype != torch.uint8 or t.device != self.device or not t.is_contiguous() or \
                tuple(t.shape) != (self.n, self.h, self.w, 3):
            raise ValueError("%s must be a contiguous uint8 tensor of shape %s on %s" %
                             (name, (self.n, self.h, self.w, 3), self.device))

    def run(self, batch, out=None, seed=DEFAULT_SEED, first_index=0, fields=None):
        """out[i] = closure(spec_i)(batch[i]); asynchronous on the current stream.
        fields: optional (f0, f1) float64 device tensors of injected noise."""
        self._check(batch, "batch")
        if out is None:
            out = torch.empty_like(batch)
        self._check(out, "out")
        ws = self._workspace()
        fptr = None
        if fields is not None:
            f0, f1 = (fields + (None,))[:2] if isinstance(fields, tuple) else (fields, None)
            for f in (f0, f1):
                if f is not None and (f.dtype != torch.float64 or f.device != self.device or not f.is_contiguous()):
                    raise ValueError("injected fields must be contiguous float64 tensors on %s" % self.device)
            fptr = ctypes.byref(_lib.B2DFields(f0.data_ptr(), f1.data_ptr() if f1 is not None else None))
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(_lib.lib().b2d_run(self._h, batch.data_ptr(), out.data_ptr(),
                                      ws.data_ptr() if ws is not None else None, fptr,
                                      int(seed) & (2 ** 64 - 1), int(first_index), stream))
        return out

    def run_host(self, h_in, h_out=None, seed=DEFAULT_SEED, first_index=0):
        """Host arrays in/out (np.ndarray or pinned CPU tensors); synchronous."""
        a = h_in.numpy() if isinstance(h_in, torch.Tensor) else np.ascontiguousarray(h_in)
        if a.dtype != np.uint8 or a.shape != (self.n, self.h, self.w, 3):
            raise ValueError("h_in must be uint8 of shape %s" % ((self.n, self.h, self.w, 3),))
        if h_out is None:
            h_out = np.empty_like(a)
        o = h_out.numpy() if isinstance(h_out, torch.Tensor) else h_out
        if o.dtype != np.uint8 or o.shape != a.shape or not o.flags["C_CONTIGUOUS"]:
            raise ValueError("h_out must be a C-contiguous uint8 array shaped like h_in")
        _lib.check(_lib.lib().b2d_run_host(self._h, a.ctypes.data, o.ctypes.data, int(seed) & (2 ** 64 - 1),
                                           int(first_index)))
        return h_out

    def noise_field(self, seed=DEFAULT_SEED, first_index=0):
        """Test hook: the float32 field(s) the generator draws (see b2d_noise_field)."""
        f0 = torch.zeros((self.n, self.h, self.w, 3), dtype=torch.float32, device=self.device)
        f1 = torch.zeros_like(f0) if any(s.noise == "sap" for s in self.specs) else None
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(_lib.lib().b2d_noise_field(self._h, f0.data_ptr(), f1.data_ptr() if f1 is not None else None,
                                              int(seed) & (2 ** 64 - 1), int(first_index), stream))
        return f0, f1

    def wavelet_stats(self, image=0):
        """Test hook (b2d_wavelet_stats): (3, 4) float32 array, per "YCbCr" channel {sigma, min, max, count}
        of image `image` (position in its spec group) as the LAST run left them."""
        out = np.zeros((3, 4), np.float32)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(_lib.lib().b2d_wavelet_stats(self._h, int(image), out.ctypes.data, stream))
        return out

    def stats(self, out, ref=None):
        """Per-image (checksum uint64 as int64 tensor, sse float64 or None)."""
        self._check(out, "out")
        cs = torch.zeros(self.n, dtype=torch.int64, device=self.device)
        sse = torch.zeros(self.n, dtype=torch.float64, device=self.device) if ref is not None else None
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(_lib.lib().b2d_stats(self._h, ref.data_ptr() if ref is not None else None, out.data_ptr(),
                                        cs.data_ptr(), sse.data_ptr() if sse is not None else None, stream))
        return cs, sse


_PLANS = {}


def get_plan(n, h, w, spec, device=None, **overrides):
    dev = torch.cuda.current_device() if device is None else device
    specs = tuple(_as_specs(spec, n, overrides))
    key = (dev, n, h, w, specs)
    p = _PLANS.get(key)
    if p is None:
        if len(_PLANS) > 64:
            _PLANS.clear()
        p = _PLANS[key] = Plan(n, h, w, list(specs), dev)
    return p


def apply(batch, spec, seed=DEFAULT_SEED, first_index=0, out=None, fields=None, **overrides):
    """Corrupt + denoise a device batch (N,H,W,3) uint8 BGR."""
    n, h, w, _ = batch.shape
    return get_plan(n, h, w, spec, batch.device.index, **overrides).run(batch, out, seed, first_index, fields)


def corrupt_host(images, spec, seed=DEFAULT_SEED, first_index=0, out=None, **overrides):
    """Host batch (N,H,W,3) uint8 -> host batch, through b2d_run_host."""
    n, h, w, _ = images.shape
    return get_plan(n, h, w, spec, None, **overrides).run_host(images, out, seed, first_index)


def corrupt(im, noise, seed=DEFAULT_SEED, index=0, side="test", rng=None, compat_flags=0, **overrides):
    """Drop-in for one pass of the reference's per-image block: HxWx3 uint8 BGR ndarray in, ndarray
    out (lib/model/test.py:1612-1831).  Concrete specs are one closure call.  `mix` specs are resolved
    with the reference's lists, and a denoiser named next to `mix` runs the way the reference runs it:
    as the post-denoise pass on the closure's uint8 result (``spec.resolve``).  ``compat_flags``:
    ``spec.COMPAT_*`` (Q3 is resolved here, the others travel in the C spec)."""
    concrete, post = resolve(noise, side, rng, compat_flags)
    if compat_flags & ~16:
        overrides = dict(overrides, compat_flags=compat_flags & ~16)
    im = np.ascontiguousarray(im)
    res = corrupt_host(im[None], concrete, seed, index, **overrides)
    if post:
        keep = {k: v for k, v in overrides.items() if k in ("ksize", "d", "sigma_color", "sigma_space", "wavelet", "levels")}
        res = corrupt_host(res, "original_" + post, seed, index, **keep)
    return res[0]


def lab_convert(pixels, direction):
    """Test hook (b2d_lab_convert): (..., 3) uint8 device tensor, 0 = BGR->Lab, 1 = Lab->BGR."""
    if pixels.dtype != torch.uint8 or not pixels.is_cuda or not pixels.is_contiguous() or pixels.shape[-1] != 3:
        raise ValueError("pixels must be a contiguous uint8 CUDA tensor of shape (..., 3)")
    out = torch.empty_like(pixels)
    stream = torch.cuda.current_stream(pixels.device).cuda_stream
    with torch.cuda.device(pixels.device):
        _lib.check(_lib.lib().b2d_lab_convert(pixels.data_ptr(), out.data_ptr(), pixels.numel() // 3, int(direction),
                                              stream))
    return out
