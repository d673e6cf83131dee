"""GPU parity of the device blob stage against the reference's own calls:
prep_im_for_blob = astype(float32) - PIXEL_MEANS -> cv2.resize(INTER_LINEAR), im_list_to_blob,
and the training-side horizontal flip (lib/utils/blob.py:17-47, minibatch.py:1676-1688).
Two modes: BLOB_F32_COORDS is bit-exact with OpenCV's own C++ resize (IPP switched off in cv2);
the default follows the IPP-backed wheels, which compute source coordinates in higher precision,
within the north star's 1e-5 relative for fp32 outputs."""
import numpy as np
import pytest
import torch

from conftest import synth_batch

pytestmark = pytest.mark.gpu
MEANS = np.array([[[102.9801, 115.9465, 122.7717]]])


def check(got, want):
    assert got.shape == want.shape, (got.shape, want.shape)
    err = np.abs(got - want)
    assert err.max() <= 1e-5 * np.abs(want).max(), err.max()


@pytest.fixture
def cv2_without_ipp():
    import cv2
    was = cv2.ipp.useIPP() if hasattr(cv2, "ipp") else False
    if hasattr(cv2, "ipp"):
        cv2.ipp.setUseIPP(False)
    yield
    if hasattr(cv2, "ipp"):
        cv2.ipp.setUseIPP(was)


@pytest.mark.parametrize("shape", [(600, 1000), (375, 500), (480, 640), (1080, 1920), (301, 333)])
def test_blob_prep_bit_exact_with_opencv_cpp_path(cuda, cv2_without_ipp, shape):
    from image_denoising_b200 import blob
    h, w = shape
    batch = synth_batch(2, h, w, seed=h, kind="smooth")
    got, _ = blob.prep_blob_device(torch.from_numpy(batch).to(cuda), MEANS, 600, 1000, flipped=[True, False],
                                   flags=blob.BLOB_F32_COORDS)
    got = got.cpu().numpy()
    for i, fl in enumerate((True, False)):
        src = np.ascontiguousarray(batch[i][:, ::-1, :]) if fl else batch[i]
        want, _ = blob.prep_im_for_blob(src, MEANS, 600, 1000)
        assert np.array_equal(got[i], want), (shape, i, np.abs(got[i] - want).max())


@pytest.mark.parametrize("shape", [(600, 1000), (375, 500), (333, 500), (500, 375), (480, 640), (301, 333), (64, 96)])
@pytest.mark.parametrize("target,mx", [(600, 1000), (800, 1333)])
def test_blob_prep_matches_cv2(cuda, shape, target, mx):
    from image_denoising_b200 import blob
    h, w = shape
    batch = synth_batch(3, h, w, seed=h + w, kind="smooth")
    got, scale = blob.prep_blob_device(torch.from_numpy(batch).to(cuda), MEANS, target, mx)
    got = got.cpu().numpy()
    for i in range(3):
        want, s_ref = blob.prep_im_for_blob(batch[i], MEANS, target, mx)
        assert scale == s_ref
        check(got[i], want)


def test_blob_prep_downscale_flip_and_padding(cuda):
    from image_denoising_b200 import blob
    batch = synth_batch(2, 1080, 1920, seed=2, kind="smooth")          # scale 1000 / 1920: downscale
    dev = torch.from_numpy(batch).to(cuda)
    flipped = [False, True]
    got, scale = blob.prep_blob_device(dev, MEANS, 600, 1000, flipped=flipped)
    got = got.cpu().numpy()
    for i in range(2):
        src = batch[i][:, ::-1, :] if flipped[i] else batch[i]
        want, _ = blob.prep_im_for_blob(np.ascontiguousarray(src), MEANS, 600, 1000)
        check(got[i], want)
    # zero padding into a larger blob (im_list_to_blob with a bigger neighbour)
    oh, ow = got.shape[1:3]
    big = torch.full((2, oh + 7, ow + 13, 3), 5.0, dtype=torch.float32, device=cuda)
    blob.prep_blob_device(dev, MEANS, 600, 1000, out=big)
    big = big.cpu().numpy()
    want = blob.im_list_to_blob([blob.prep_im_for_blob(batch[0], MEANS, 600, 1000)[0], np.zeros((oh + 7, ow + 13, 3))])
    check(big[0], want[0])
    assert not big[:, oh:, :, :].any() and not big[:, :, ow:, :].any()


def test_blob_prep_scale_one_is_mean_subtraction(cuda):
    """600 x 1000 images keep their size: the blob is exactly float32(v - mean) (in-place numpy semantics)."""
    from image_denoising_b200 import blob
    batch = synth_batch(2, 600, 1000, seed=5)
    got, scale = blob.prep_blob_device(torch.from_numpy(batch).to(cuda), MEANS, 600, 1000)
    assert scale == 1.0
    want = batch.astype(np.float32)
    want -= MEANS
    assert np.array_equal(got.cpu().numpy(), want)


def test_minibatch_device_blob_equals_host_path(cuda):
    """get_minibatch(device_blob=True): same dict as the host path, 'data' on the GPU."""
    import numpy.random as npr
    import scipy.sparse
    from image_denoising_b200 import minibatch
    imgs = synth_batch(3, 375, 500, seed=8, kind="smooth")
    big = synth_batch(1, 480, 640, seed=9, kind="smooth")[0]

    def entry(im, i, spec, flipped):
        return {'image': im, 'noise_type': spec, 'flipped': flipped, 'index': i,
                'boxes': np.array([[10, 20, 100, 120]], np.float32), 'gt_classes': np.array([3], np.int32),
                'gt_overlaps': scipy.sparse.csr_matrix(np.ones((1, 21), np.float32))}

    roidb = [entry(imgs[0], 0, "sap_median_var0.4", False), entry(imgs[1], 1, "original_gaus_blur", True),
             entry(big, 2, "periodic_mean_var100", True), entry(imgs[2], 3, "shader", False)]
    host, s_host = minibatch._get_image_blob(roidb, [0, 0, 0, 0])
    dev, s_dev = minibatch._get_image_blob_device(roidb, [0, 0, 0, 0])
    assert s_host == s_dev and tuple(dev.shape) == host.shape
    check(dev.cpu().numpy(), host)
    npr.seed(1)
    a = minibatch.get_minibatch(roidb[:1], 21)
    npr.seed(1)
    b = minibatch.get_minibatch(roidb[:1], 21, device_blob=True)
    assert b['data'].is_cuda and np.array_equal(a['gt_boxes'], b['gt_boxes']) and np.array_equal(a['im_info'], b['im_info'])
    check(b['data'].cpu().numpy(), a['data'])


def test_test_side_blob_device(cuda):
    from image_denoising_b200 import test_hook
    batch = synth_batch(2, 375, 500, seed=11, kind="smooth")
    dev, scales = test_hook.get_test_blob_device(torch.from_numpy(batch).to(cuda))
    for i in range(2):
        want, s_ref = test_hook.get_test_blob(batch[i])
        assert np.array_equal(scales, s_ref)
        check(dev[i * len(scales):(i + 1) * len(scales)].cpu().numpy(), want)
