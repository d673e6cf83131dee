"""Drop-in for the per-image corruption block of `test_net` (lib/model/test.py:189-1831):

    for i in range(num_images):
        img = cv2.imread(imdb.image_path_at(i))        # test.py:191
        ... 1600 lines of closures + dispatch ...       # test.py:193-1831
        scores, boxes = im_detect(sess, net, im)        # test.py:1836

becomes `im = corrupt_for_test(img, noise, i)`; `get_test_blob` mirrors `_get_image_blob` of
test.py:49-83 (mean subtraction, TEST.SCALES / TEST.MAX_SIZE resize)."""
import numpy as np

from . import api
from .blob import im_list_to_blob
from .config import cfg


def corrupt_for_test(img, noise, index=0, rng=None, **overrides):
    """One image, test.py semantics: uint8 BGR in, uint8 out, `mix` lists of test.py:1613-1639."""
    return api.corrupt(img, noise, seed=cfg.RNG_SEED, index=index, side="test", rng=rng, **overrides)


def corrupt_batch_for_test(imgs, noise, first_index=0, **overrides):
    """Batched fast path: (N,H,W,3) uint8 host array, one spec, global indices first_index + i."""
    return api.corrupt_host(np.ascontiguousarray(imgs), noise, seed=cfg.RNG_SEED, first_index=first_index,
                            **overrides)


def get_test_blob(im):
    """test.py:49-83: image pyramid blob + scale factors."""
    import cv2
    im_orig = im.astype(np.float32, copy=True)
    im_orig -= cfg.PIXEL_MEANS
    im_size_min, im_size_max = np.min(im_orig.shape[0:2]), np.max(im_orig.shape[0:2])
    processed_ims, im_scale_factors = [], []
    for target_size in cfg.TEST.SCALES:
        im_scale = float(target_size) / float(im_size_min)
        if np.round(im_scale * im_size_max) > cfg.TEST.MAX_SIZE:
            im_scale = float(cfg.TEST.MAX_SIZE) / float(im_size_max)
        processed_ims.append(cv2.resize(im_orig, None, None, fx=im_scale, fy=im_scale,
                                        interpolation=cv2.INTER_LINEAR))
        im_scale_factors.append(im_scale)
    return im_list_to_blob(processed_ims), np.array(im_scale_factors)


def get_test_blob_device(batch):
    """test.py:49-83 for a device batch (N, H, W, 3) uint8 (the output of `corrupt_batch_for_test` kept
    on the GPU): blob (N * len(TEST.SCALES), Hmax, Wmax, 3) float32 CUDA tensor + the scale factors.
    Scale s of image i is blob[i * len(SCALES) + s], the order `_get_image_blob` gives for one image."""
    import torch
    from .blob import blob_shape, prep_blob_device
    n, h, w, _ = batch.shape
    shapes = [blob_shape(h, w, t, cfg.TEST.MAX_SIZE) for t in cfg.TEST.SCALES]
    hmax, wmax, ns = max(s[1] for s in shapes), max(s[2] for s in shapes), len(shapes)
    blob = torch.empty((n, ns, hmax, wmax, 3), dtype=torch.float32, device=batch.device)
    for k, target in enumerate(cfg.TEST.SCALES):
        if ns == 1:
            prep_blob_device(batch, cfg.PIXEL_MEANS, target, cfg.TEST.MAX_SIZE, out=blob.view(n, hmax, wmax, 3))
        else:
            tmp = torch.empty((n, hmax, wmax, 3), dtype=torch.float32, device=batch.device)
            prep_blob_device(batch, cfg.PIXEL_MEANS, target, cfg.TEST.MAX_SIZE, out=tmp)
            blob[:, k] = tmp
    return blob.view(n * ns, hmax, wmax, 3), np.array([s[0] for s in shapes])


def corrupt_paths_for_test_device(paths, noise, first_index=0, **overrides):
    """The first three lines of the test loop for a batch of same-sized JPEG files, everything on the GPU:
    `cv2.imread` (test.py:191) = `b2d_jpeg_decode` (bit-exact), closures (test.py:193-1831) = `b2d_run`.
    Returns the (N, H, W, 3) uint8 CUDA tensor `im_detect` would get image by image (feed `get_test_blob_device`)."""
    from . import jpeg
    from .spec import resolve
    batch = jpeg.imread_batch(paths)
    resolved = [resolve(noise, "test", None) for _ in paths]
    res = api.apply(batch, [r[0] for r in resolved], seed=cfg.RNG_SEED, first_index=first_index, **overrides)
    post = resolved[0][1] if resolved else None            # same top-level string: same post pass for every image
    if post:                                               # `noise_mix_*_wavelet`: the live post pass of test.py:1805-1812
        res = api.apply(res, "original_" + post, seed=cfg.RNG_SEED, first_index=first_index)
    return res


def nms_for_test(cls_dets, thresh=None):
    """`keep = nms(cls_dets, cfg.TEST.NMS)` of the per-class loop, test.py:1854-1862 (`b2d_nms`)."""
    from .nms import nms
    return nms(np.ascontiguousarray(cls_dets, dtype=np.float32), cfg.TEST.NMS if thresh is None else thresh)
