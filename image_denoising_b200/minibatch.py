"""Drop-in for lib/roi_data_layer/minibatch.py: `get_minibatch` / `_get_image_blob` with the
reference's names, roidb keys and blob layout; the 1600 lines of closures in between
(minibatch.py:87-1673) are ONE call into the CUDA library.

roidb entry keys read (set by lib/roi_data_layer/roidb.py:31,50 and lib/datasets/imdb.py:119-122):
  'image' (path, or an HxWx3 uint8 BGR ndarray for in-memory datasets), 'noise_type' (the spec
  string), 'flipped', 'boxes', 'gt_classes', 'gt_overlaps'.

Differences from the reference, all deliberate:
  * one denoise pass (the inner closure); `double_filter=True` re-applies the post pass of
    minibatch.py:1636-1663 (quirk Q1);
  * noise-only branches return uint8 like lib/model/test.py, not the unscaled float image (3.4);
  * randomness is Philox keyed by (cfg.RNG_SEED, image index) instead of the global np.random;
  * `_get_image_blob` accepts any number of same-shaped images and sends them as ONE batch.
"""
import cv2
import numpy as np
import numpy.random as npr

from . import api
from .blob import im_list_to_blob, prep_im_for_blob
from .config import cfg
from .spec import POST_DENOISE_ORDER, resolve


def _load(entry):
    im = entry['image']
    if isinstance(im, np.ndarray):
        return im
    img = cv2.imread(im)                                       # minibatch.py:85
    if img is None:
        raise IOError("cannot read image %r" % (im,))
    return img


class _JpegStream:
    """A not-yet-decoded JPEG file: bytes + the frame size from its header (decoded on the GPU)."""

    def __init__(self, data):
        from . import jpeg
        self.data = data
        h, w = jpeg.info(data)[:2]
        self.shape = (h, w, 3)


def _load_for_device(entry):
    """`_load`, except that JPEG files stay encoded: `cv2.imread` (minibatch.py:85) then happens on the GPU, bit-exact."""
    im = entry['image']
    if isinstance(im, np.ndarray):
        return im
    if str(im).lower().endswith((".jpg", ".jpeg")):
        with open(im, "rb") as f:
            return _JpegStream(f.read())
    return _load(entry)


def _to_device_batch(items, dev):
    """list of ndarrays / _JpegStream of one shape -> (n, h, w, 3) uint8 CUDA tensor."""
    import torch
    from . import jpeg
    enc = [k for k, it in enumerate(items) if isinstance(it, _JpegStream)]
    if len(enc) == len(items):
        return jpeg.imdecode_batch([it.data for it in items], dev)
    batch = torch.empty((len(items),) + tuple(items[0].shape), dtype=torch.uint8, device=dev)
    raw = [k for k in range(len(items)) if k not in enc]
    batch[raw] = torch.from_numpy(np.ascontiguousarray(np.stack([items[k] for k in raw]))).to(dev)
    if enc:
        batch[enc] = jpeg.imdecode_batch([items[k].data for k in enc], dev)
    return batch


def corrupt_images(imgs, specs, indices, side="train", double_filter=False, rng=None):
    """imgs: list of HxWx3 uint8 BGR arrays -> list of corrupted + denoised uint8 arrays.
    Images of equal shape go to the GPU as one batch (per-image specs)."""
    resolved = [resolve(s, side, rng) for s in specs]
    specs = [r[0] for r in resolved]
    posts = [r[1] for r in resolved]                     # mix + denoiser: the reference's post-denoise pass
    out = [None] * len(imgs)
    by_shape = {}
    for i, im in enumerate(imgs):
        by_shape.setdefault(im.shape, []).append(i)
    for shape, idx in by_shape.items():
        batch = np.ascontiguousarray(np.stack([imgs[i] for i in idx]))
        res = np.empty_like(batch)
        consecutive = all(indices[idx[j]] == indices[idx[0]] + j for j in range(len(idx)))
        if consecutive:      # one plan, per-image specs; the global image index keys the Philox counter
            api.corrupt_host(batch, [specs[i] for i in idx], seed=cfg.RNG_SEED, first_index=indices[idx[0]], out=res)
        else:
            for j, i in enumerate(idx):
                res[j] = api.corrupt_host(batch[j:j + 1], specs[i], seed=cfg.RNG_SEED, first_index=indices[i])[0]
        for j, i in enumerate(idx):
            post = posts[i]
            if post is None and double_filter and isinstance(specs[i], str):
                post = next((d for d in POST_DENOISE_ORDER if d in specs[i]), None)
            if post:
                res[j] = api.corrupt_host(res[j:j + 1], "original_" + post)[0]
        for j, i in enumerate(idx):
            out[i] = res[j]
    return out


def _get_image_blob(roidb, scale_inds, double_filter=False):
    """Builds an input blob from the images in the roidb at the specified scales
    (minibatch.py:77-86,1676-1690)."""
    imgs = [_load(r) for r in roidb]
    specs = [r.get('noise_type', 'original') for r in roidb]
    indices = [int(r.get('index', i)) for i, r in enumerate(roidb)]
    ims = corrupt_images(imgs, specs, indices, "train", double_filter)
    processed_ims, im_scales = [], []
    for i, im in enumerate(ims):
        if roidb[i].get('flipped', False):
            im = im[:, ::-1, :]                                # minibatch.py:1676-1677
        target_size = cfg.TRAIN.SCALES[scale_inds[i]]
        im, im_scale = prep_im_for_blob(im, cfg.PIXEL_MEANS, target_size, cfg.TRAIN.MAX_SIZE)
        im_scales.append(im_scale)
        processed_ims.append(im)
    return im_list_to_blob(processed_ims), im_scales


def _get_image_blob_device(roidb, scale_inds, double_filter=False, device_decode=False):
    """`_get_image_blob` with every step after the decode on the GPU: one H2D copy of
    the uint8 images, corrupt -> denoise (`b2d_run`), flip + mean subtraction + bilinear resize + NHWC
    packing (`b2d_blob_prep`).  Returns (blob: (N, Hmax, Wmax, 3) float32 CUDA tensor, im_scales).
    device_decode=True also moves `cv2.imread` of JPEG files to the GPU (8(f) f2: `b2d_jpeg_decode`; the H2D copy is
    then the Huffman-decoded coefficients); a stream that decoder refuses raises, nothing falls back."""
    import torch
    from .blob import blob_shape, prep_blob_device
    imgs = [_load_for_device(r) if device_decode else _load(r) for r in roidb]
    resolved = [resolve(r.get('noise_type', 'original'), "train", None) for r in roidb]
    specs = [r[0] for r in resolved]
    posts = [r[1] for r in resolved]
    indices = [int(r.get('index', i)) for i, r in enumerate(roidb)]
    targets = [cfg.TRAIN.SCALES[scale_inds[i]] for i in range(len(roidb))]
    shapes = [blob_shape(im.shape[0], im.shape[1], targets[i], cfg.TRAIN.MAX_SIZE) for i, im in enumerate(imgs)]
    hmax, wmax = max(s[1] for s in shapes), max(s[2] for s in shapes)
    dev = torch.device("cuda", torch.cuda.current_device())
    blob = torch.empty((len(imgs), hmax, wmax, 3), dtype=torch.float32, device=dev)
    groups = {}
    for i, im in enumerate(imgs):
        groups.setdefault((im.shape, targets[i]), []).append(i)
    for (shape, target), idx in groups.items():
        batch = _to_device_batch([imgs[i] for i in idx], dev)
        consecutive = all(indices[idx[j]] == indices[idx[0]] + j for j in range(len(idx)))
        if consecutive:
            res = api.apply(batch, [specs[i] for i in idx], seed=cfg.RNG_SEED, first_index=indices[idx[0]])
        else:
            res = torch.cat([api.apply(batch[j:j + 1], specs[i], seed=cfg.RNG_SEED, first_index=indices[i])
                             for j, i in enumerate(idx)])
        for j, i in enumerate(idx):
            post = posts[i]
            if post is None and double_filter and isinstance(specs[i], str):
                post = next((d for d in POST_DENOISE_ORDER if d in specs[i]), None)
            if post:
                res[j:j + 1] = api.apply(res[j:j + 1].contiguous(), "original_" + post)
        flips = [bool(roidb[i].get('flipped', False)) for i in idx]
        whole = idx == list(range(idx[0], idx[0] + len(idx)))
        if whole:                                              # the group is a contiguous slab of the blob
            prep_blob_device(res, cfg.PIXEL_MEANS, target, cfg.TRAIN.MAX_SIZE, flipped=flips,
                             out=blob[idx[0]:idx[0] + len(idx)])
        else:
            for j, i in enumerate(idx):
                prep_blob_device(res[j:j + 1], cfg.PIXEL_MEANS, target, cfg.TRAIN.MAX_SIZE, flipped=flips[j:j + 1],
                                 out=blob[i:i + 1])
    return blob, [s[0] for s in shapes]


def get_minibatch(roidb, num_classes, device_blob=False, device_decode=False):
    """Given a roidb, construct a minibatch sampled from it (minibatch.py:42-75).
    device_blob=True keeps blobs['data'] on the GPU (built by `_get_image_blob_device`); device_decode=True (with
    device_blob) also decodes JPEG files there."""
    num_images = len(roidb)
    random_scale_inds = npr.randint(0, high=len(cfg.TRAIN.SCALES), size=num_images)
    assert cfg.TRAIN.BATCH_SIZE % num_images == 0, \
        'num_images ({}) must divide BATCH_SIZE ({})'.format(num_images, cfg.TRAIN.BATCH_SIZE)
    if device_blob:
        im_blob, im_scales = _get_image_blob_device(roidb, random_scale_inds, device_decode=device_decode)
    else:
        im_blob, im_scales = _get_image_blob(roidb, random_scale_inds)
    blobs = {'data': im_blob}
    assert len(im_scales) == 1, "Single batch only"
    assert len(roidb) == 1, "Single batch only"
    if cfg.TRAIN.USE_ALL_GT:
        gt_inds = np.where(roidb[0]['gt_classes'] != 0)[0]
    else:
        gt_inds = np.where(roidb[0]['gt_classes'] != 0 &
                           np.all(roidb[0]['gt_overlaps'].toarray() > -1.0, axis=1))[0]
    gt_boxes = np.empty((len(gt_inds), 5), dtype=np.float32)
    gt_boxes[:, 0:4] = roidb[0]['boxes'][gt_inds, :] * im_scales[0]
    gt_boxes[:, 4] = roidb[0]['gt_classes'][gt_inds]
    blobs['gt_boxes'] = gt_boxes
    blobs['im_info'] = np.array([im_blob.shape[1], im_blob.shape[2], im_scales[0]], dtype=np.float32)
    return blobs
