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
