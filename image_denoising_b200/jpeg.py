"""Device JPEG decode in front of the path.

Mirrors the reference's ``img = cv2.imread(path)`` (lib/model/test.py:191,
lib/roi_data_layer/minibatch.py:85): same BGR uint8 H x W x 3 bytes, but for a whole batch and left on
the GPU, where ``api.apply`` / ``Plan.run`` consume it.  Entropy decoding runs on host threads inside
libb200denoise.so, everything else in CUDA kernels; there is no fallback decoder - an unsupported
stream (progressive, arithmetic, CMYK, 12-bit) raises ``B2DError``.
"""
import ctypes

import torch

from . import _lib


def info(data):
    """(h, w, components, hsamp, vsamp) of one JPEG stream (bytes)."""
    L = _lib.lib()
    buf = (ctypes.c_uint8 * len(data)).from_buffer_copy(data)
    v = [ctypes.c_int() for _ in range(5)]
    _lib.check(L.b2d_jpeg_info(ctypes.cast(buf, ctypes.c_void_p), len(data), *[ctypes.byref(x) for x in v]))
    return tuple(x.value for x in v)


def entropy_decode(data):
    """Host-only: the quantised coefficients of every component, [(blocks_y, blocks_x, 64) int16, ...] (natural order)."""
    import numpy as np
    L = _lib.lib()
    h, w = info(data)[:2]
    cap = 3 * ((h + 15) // 16 * 16) * ((w + 15) // 16 * 16)
    buf = (ctypes.c_uint8 * len(data)).from_buffer_copy(data)
    coef = np.zeros(cap, np.int16)
    bx, by = (ctypes.c_int * 3)(), (ctypes.c_int * 3)()
    _lib.check(L.b2d_jpeg_entropy(ctypes.cast(buf, ctypes.c_void_p), len(data), ctypes.c_void_p(coef.ctypes.data), cap, bx, by))
    out, off = [], 0
    for c in range(3):
        n = bx[c] * by[c] * 64
        if n:
            out.append(coef[off:off + n].reshape(by[c], bx[c], 64).copy())
        off += n
    return out


class Decoder:
    """Batches of up to ``max_n`` JPEG streams of one size -> uint8 CUDA tensor (n, h, w, 3), BGR."""

    def __init__(self, max_n, h, w, device=None, host_threads=0):
        self.L = _lib.lib()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.max_n, self.h, self.w = int(max_n), int(h), int(w)
        self.handle = ctypes.c_void_p()
        _lib.check(self.L.b2d_jpeg_create(ctypes.byref(self.handle), self.device.index or 0, self.max_n, self.h, self.w,
                                          int(host_threads)))

    def decode(self, streams, out=None):
        n = len(streams)
        if n > self.max_n:
            raise ValueError("%d streams, decoder holds %d" % (n, self.max_n))
        if out is None:
            out = torch.empty((n, self.h, self.w, 3), dtype=torch.uint8, device=self.device)
        elif out.dtype != torch.uint8 or tuple(out.shape) != (n, self.h, self.w, 3) or not out.is_contiguous() or out.device != self.device:
            raise ValueError("out must be a contiguous uint8 tensor of shape (%d, %d, %d, 3) on %s" % (n, self.h, self.w, self.device))
        keep = [bytes(s) for s in streams]                       # no copy for bytes objects; the pointers below borrow them
        ptrs = (ctypes.c_void_p * max(n, 1))(*[ctypes.cast(ctypes.c_char_p(b), ctypes.c_void_p) for b in keep])
        lens = (ctypes.c_size_t * max(n, 1))(*[len(s) for s in streams])
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(self.L.b2d_jpeg_decode(self.handle, n, ptrs, lens, ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(stream)))
        return out

    def close(self):
        if self.handle:
            self.L.b2d_jpeg_destroy(self.handle)
            self.handle = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_decoders = {}


def imdecode_batch(streams, device=None):
    """``[cv2.imdecode(s, IMREAD_COLOR) for s in streams]`` as one CUDA tensor; all streams must share one size."""
    if not streams:
        raise ValueError("no streams")
    h, w = info(streams[0])[:2]
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    key = (dev.index or 0, h, w)
    dec = _decoders.get(key)
    if dec is None or dec.max_n < len(streams):
        if dec is not None:
            dec.close()
        dec = _decoders[key] = Decoder(max(len(streams), 8), h, w, dev)
    return dec.decode(streams)


def imread_batch(paths, device=None):
    """``[cv2.imread(p) for p in paths]`` (lib/model/test.py:191) as one CUDA tensor."""
    streams = []
    for p in paths:
        with open(p, "rb") as f:
            streams.append(f.read())
    return imdecode_batch(streams, device)
