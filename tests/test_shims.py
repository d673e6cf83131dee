"""The reference-named shims around the path: blob helpers (CPU) and, on the GPU, `get_minibatch` /
`_get_image_blob` (lib/roi_data_layer/minibatch.py:42-86,1676-1690) and the test_net hook
(lib/model/test.py:189-191,1836)."""
import numpy as np
import pytest

from conftest import synth_image
from image_denoising_b200 import blob
from image_denoising_b200.config import cfg


def test_prep_im_for_blob_matches_the_reference_arithmetic():
    import cv2
    im = synth_image(375, 500, 1, "smooth")
    out, scale = blob.prep_im_for_blob(im, cfg.PIXEL_MEANS, 600, 1000)
    assert scale == 600.0 / 375.0 and out.shape == (600, 800, 3) and out.dtype == np.float32
    want = cv2.resize(im.astype(np.float32) - cfg.PIXEL_MEANS.astype(np.float32), None, None, fx=scale, fy=scale,
                      interpolation=cv2.INTER_LINEAR)
    assert np.allclose(out, want, atol=1e-4)
    _, s2 = blob.prep_im_for_blob(synth_image(300, 900, 2), cfg.PIXEL_MEANS, 600, 1000)
    assert s2 == 1000.0 / 900.0                                   # long side capped at MAX_SIZE
    assert im.dtype == np.uint8                                   # input untouched


def test_im_list_to_blob_pads_to_the_largest():
    a, b = np.ones((4, 6, 3), np.float32), 2 * np.ones((5, 3, 3), np.float32)
    bl = blob.im_list_to_blob([a, b])
    assert bl.shape == (2, 5, 6, 3) and bl.dtype == np.float32
    assert bl[0, :4, :6].min() == 1 and bl[0, 4].max() == 0 and bl[1, :, :3].min() == 2 and bl[1, :, 3:].max() == 0


def _roidb(im, spec, flipped=False, index=0):
    return {"image": im, "noise_type": spec, "flipped": flipped, "index": index,
            "boxes": np.array([[10, 20, 110, 220], [5, 5, 50, 60]], np.float32),
            "gt_classes": np.array([3, 0], np.int32)}


@pytest.mark.gpu
def test_get_minibatch_blobs(cuda):
    from image_denoising_b200 import minibatch
    from oracle import closures as C
    im = synth_image(375, 500, 3, "smooth")
    for flipped in (False, True):
        blobs = minibatch.get_minibatch([_roidb(im, "periodic_median_var100", flipped)], 21)
        assert blobs["data"].shape == (1, 600, 800, 3) and blobs["data"].dtype == np.float32
        assert np.allclose(blobs["im_info"], [600, 800, 1.6])
        assert blobs["gt_boxes"].shape == (1, 5) and np.allclose(blobs["gt_boxes"][0], [16, 32, 176, 352, 3])
        want = C.corrupt(im, "periodic_median_var100")
        if flipped:
            want = want[:, ::-1, :]
        want, _ = __import__("image_denoising_b200.blob", fromlist=["x"]).prep_im_for_blob(want, cfg.PIXEL_MEANS, 600, 1000)
        assert np.allclose(blobs["data"][0], want, atol=1e-3)


@pytest.mark.gpu
def test_image_blob_batches_same_shape_images_and_double_filter(cuda):
    from image_denoising_b200 import minibatch
    from oracle import closures as C
    ims = [synth_image(64, 96, 10 + i) for i in range(3)]
    specs = ["shader", "original_mean", "periodic_gaus_blur_var3.14"]
    out = minibatch.corrupt_images(ims, specs, [5, 6, 7])
    for im, s, o in zip(ims, specs, out):
        want = C.corrupt(im, s) if "original" not in s else C.apply_denoise(im, "mean")
        assert np.array_equal(o, want), s
    twice = minibatch.corrupt_images(ims[2:], specs[2:], [7], double_filter=True)[0]      # quirk Q1
    assert np.array_equal(twice, C.apply_denoise(C.corrupt(ims[2], specs[2]), "gaus_blur"))
    # stochastic spec: independent of how the images are grouped (global index keys the counter)
    a = minibatch.corrupt_images(ims, ["gaussian_median_var1.0"] * 3, [5, 6, 7])
    b = minibatch.corrupt_images(ims[1:2], ["gaussian_median_var1.0"], [6])
    assert np.array_equal(a[1], b[0])


@pytest.mark.gpu
def test_test_hook(cuda):
    import random
    from image_denoising_b200 import test_hook, spec
    from oracle import closures as C
    im = synth_image(300, 400, 4, "smooth")          # H >= 300: every entry of the mix list is runnable
    assert np.array_equal(test_hook.corrupt_for_test(im, "shader"), C.corrupt(im, "shader"))
    assert np.array_equal(test_hook.corrupt_for_test(im, "original"), im)
    got = test_hook.corrupt_for_test(im, "noise_mix_var_low", index=3, rng=random.Random(1))
    chosen = spec.resolve_mix("noise_mix_var_low", "test", random.Random(1))
    assert got.shape == im.shape and got.dtype == np.uint8
    if chosen in ("original",):
        assert np.array_equal(got, im)
    bl, sc = test_hook.get_test_blob(got)
    assert bl.shape == (1, 600, 800, 3) and np.allclose(sc, [2.0])
    batch = np.stack([im, im])
    out = test_hook.corrupt_batch_for_test(batch, "sap_mean_var0.4", first_index=0)
    assert out.shape == batch.shape and not np.array_equal(out[0], out[1])                 # different counters
