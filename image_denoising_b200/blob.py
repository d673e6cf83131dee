"""Blob helpers: the step right after the corrupt -> denoise path (lib/utils/blob.py:17-47;
SURVEY.md 8(f) f1).  Host-side mirror with the reference's names and semantics - mean
subtraction, bilinear resize so that the short side is `target_size` and the long side at most
`max_size`, NHWC float32 packing.  It is NOT on the timed hot path; the device version (fused into
the filter epilogue) is the next row of the scope table."""
import cv2
import numpy as np


def im_list_to_blob(ims):
    """List of HxWx3 float images -> (N, Hmax, Wmax, 3) float32, zero padded (blob.py:17-30)."""
    max_shape = np.array([im.shape for im in ims]).max(axis=0)
    blob = np.zeros((len(ims), max_shape[0], max_shape[1], 3), dtype=np.float32)
    for i, im in enumerate(ims):
        blob[i, 0:im.shape[0], 0:im.shape[1], :] = im
    return blob


def prep_im_for_blob(im, pixel_means, target_size, max_size):
    """Mean subtract and scale an image (blob.py:33-47).  Returns (im float32, im_scale)."""
    im = im.astype(np.float32, copy=True)
    im -= pixel_means
    im_size_min, im_size_max = np.min(im.shape[0:2]), np.max(im.shape[0:2])
    im_scale = float(target_size) / float(im_size_min)
    if np.round(im_scale * im_size_max) > max_size:          # keep the long side <= MAX_SIZE
        im_scale = float(max_size) / float(im_size_max)
    im = cv2.resize(im, None, None, fx=im_scale, fy=im_scale, interpolation=cv2.INTER_LINEAR)
    return im, im_scale


# ------------------------------------------------------------------ device version
def blob_shape(h, w, target_size, max_size):
    """(im_scale, out_h, out_w) of prep_im_for_blob for an h x w image, from the library
    (b2d_blob_shape: blob.py:37-45 + cv2.resize's dsize rounding)."""
    import ctypes
    from . import _lib
    s, oh, ow = ctypes.c_double(), ctypes.c_int(), ctypes.c_int()
    _lib.check(_lib.lib().b2d_blob_shape(int(h), int(w), int(target_size), int(max_size), ctypes.byref(s),
                                         ctypes.byref(oh), ctypes.byref(ow)))
    return s.value, oh.value, ow.value


BLOB_F32_COORDS = 1       # b2d_blob_prep flag: OpenCV's generic C++ path bit for bit (no IPP)


def prep_blob_device(batch, pixel_means, target_size, max_size, flipped=None, out=None, flags=0):
    """prep_im_for_blob + im_list_to_blob (+ the training-side horizontal flip) for a device batch.

    batch: (N, H, W, 3) uint8 CUDA tensor (the output of the corrupt -> denoise path);
    flipped: optional length-N booleans (roidb[i]['flipped'], minibatch.py:1676).
    flags: BLOB_F32_COORDS reproduces cv2.resize without IPP bit for bit; the default follows the
    IPP-backed wheels (double-precision source coordinates).
    Returns (blob (N, H', W', 3) float32 CUDA tensor, im_scale); asynchronous on the current stream."""
    import ctypes
    import torch
    from . import _lib
    if batch.dtype != torch.uint8 or not batch.is_cuda or not batch.is_contiguous() or batch.dim() != 4 or \
            batch.shape[3] != 3:
        raise ValueError("batch must be a contiguous (N, H, W, 3) uint8 CUDA tensor")
    n, h, w, _ = batch.shape
    scale, oh, ow = blob_shape(h, w, target_size, max_size)
    if out is None:
        out = torch.empty((n, oh, ow, 3), dtype=torch.float32, device=batch.device)
    elif out.dtype != torch.float32 or out.device != batch.device or not out.is_contiguous() or \
            out.dim() != 4 or out.shape[0] != n or out.shape[1] < oh or out.shape[2] < ow or out.shape[3] != 3:
        raise ValueError("out must be a contiguous (N, >=%d, >=%d, 3) float32 tensor on %s" % (oh, ow, batch.device))
    means = (ctypes.c_double * 3)(*[float(v) for v in np.asarray(pixel_means, np.float64).reshape(-1)[:3]])
    flip = None
    if flipped is not None:
        flip = torch.as_tensor(np.asarray(flipped, np.uint8).reshape(n)).to(batch.device)
    with torch.cuda.device(batch.device):
        stream = torch.cuda.current_stream(batch.device).cuda_stream
        _lib.check(_lib.lib().b2d_blob_prep(batch.data_ptr(), n, h, w, out.data_ptr(), out.shape[1], out.shape[2],
                                            scale, means, flip.data_ptr() if flip is not None else None,
                                            int(flags), stream))
    return out, scale
