"""GPU: the device JPEG stage (SURVEY 8(f) f2, b2d_jpeg_decode through the C ABI) against cv2.imdecode - the call the
reference makes (cv2.imread, lib/model/test.py:191) - byte for byte, and against the pinned oracle."""
import cv2
import numpy as np
import pytest
import torch

from conftest import synth_image
from oracle import jpeg as J
from test_oracle_jpeg import CASES, S420, S422, S444, encode, streams

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("tag,data", list(streams()), ids=lambda v: str(v) if not isinstance(v, bytes) else "")
def test_single_streams_bit_exact(cuda, tag, data):
    from image_denoising_b200 import jpeg
    want = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
    got = jpeg.imdecode_batch([data], cuda).cpu().numpy()[0]
    assert got.shape == want.shape
    assert np.array_equal(got, want), (tag, int((got != want).sum()))
    assert np.array_equal(got, J.decode(data))


@pytest.mark.parametrize("shape", [(375, 500), (600, 1000), (601, 999)])
def test_mixed_batch_bit_exact(cuda, shape):
    """One batch, every image its own quality / sampling / restart interval (VOC files differ in all three)."""
    from image_denoising_b200 import jpeg
    h, w = shape
    datas = []
    for i, (q, samp, rst) in enumerate([(90, S420, 0), (75, S444, 0), (60, S422, 4), (95, S420, 11), (35, S420, 0), (85, None, 1)]):
        datas.append(encode(synth_image(h, w, 50 + i, "smooth" if i % 2 == 0 else "uniform"), q, samp, rst))
    datas.append(encode(synth_image(h, w, 9, "smooth")[..., 1], 80))                # greyscale file
    got = jpeg.imdecode_batch(datas, cuda).cpu().numpy()
    for i, d in enumerate(datas):
        want = cv2.imdecode(np.frombuffer(d, np.uint8), cv2.IMREAD_COLOR)
        assert np.array_equal(got[i], want), (shape, i, int((got[i] != want).sum()))
    again = jpeg.imdecode_batch(datas[:3], cuda).cpu().numpy()                      # decoder reuse, smaller batch
    assert np.array_equal(again, got[:3])


def test_natural_image_and_feeds_the_path(cuda):
    """A natural image (the crop of the reference's data/demo/000456.jpg kept under tests/golden) re-encoded, decoded on
    the device and handed straight to the corrupt -> denoise path: same bytes as cv2.imdecode + the same path."""
    import golden_cases as G
    from image_denoising_b200 import api, jpeg
    img = G.inputs()["natural_120x160"]
    datas = [encode(img, q, s) for q, s in ((92, S420), (70, S444))]
    dev = jpeg.imdecode_batch(datas, cuda)
    host = np.stack([cv2.imdecode(np.frombuffer(d, np.uint8), cv2.IMREAD_COLOR) for d in datas])
    assert np.array_equal(dev.cpu().numpy(), host)
    a = api.apply(dev, "periodic_median_var100")
    b = api.apply(torch.from_numpy(host).to(cuda), "periodic_median_var100")
    assert torch.equal(a, b)


def test_errors_are_loud(cuda):
    from image_denoising_b200 import _lib, jpeg
    img = synth_image(32, 48, 1, "smooth")
    good, other = encode(img, 90), encode(synth_image(40, 48, 1, "smooth"), 90)
    with pytest.raises(_lib.B2DError):
        jpeg.imdecode_batch([good, other], cuda)                                    # sizes differ inside one batch
    prog = encode(img, 90, None, 0, (cv2.IMWRITE_JPEG_PROGRESSIVE, 1))
    with pytest.raises(_lib.B2DError) as e:
        jpeg.imdecode_batch([prog], cuda)
    assert e.value.code == _lib.ERR_UNSUPPORTED
    ok = jpeg.imdecode_batch([good], cuda).cpu().numpy()[0]                         # decoder still usable afterwards
    assert np.array_equal(ok, cv2.imdecode(np.frombuffer(good, np.uint8), cv2.IMREAD_COLOR))


def test_minibatch_device_decode_equals_imread_path(cuda, tmp_path):
    """get_minibatch / _get_image_blob_device with device_decode=True: roidb entries name JPEG files, as in the
    reference (roidb[i]['image'], minibatch.py:85); the blob equals the one built from cv2.imread on the host."""
    import scipy.sparse
    from image_denoising_b200 import minibatch
    paths = []
    for i, (q, s) in enumerate([(90, S420), (80, S444), (70, S420)]):
        p = str(tmp_path / ("im%d.jpg" % i))
        with open(p, "wb") as f:
            f.write(encode(synth_image(375, 500, 20 + i, "smooth"), q, s))
        paths.append(p)
    raw = synth_image(375, 500, 30, "smooth")                                        # in-memory entry in the same shape group

    def entry(im, i, spec, flipped):
        return {'image': im, 'noise_type': spec, 'flipped': flipped, 'index': i,
                'boxes': np.array([[10, 20, 100, 120]], np.float32), 'gt_classes': np.array([3], np.int32),
                'gt_overlaps': scipy.sparse.csr_matrix(np.ones((1, 21), np.float32))}

    roidb = [entry(paths[0], 0, "sap_median_var0.4", False), entry(paths[1], 1, "original_gaus_blur", True),
             entry(raw, 2, "periodic_mean_var100", False), entry(paths[2], 3, "shader", True)]
    a, sa = minibatch._get_image_blob_device(roidb, [0, 0, 0, 0])
    b, sb = minibatch._get_image_blob_device(roidb, [0, 0, 0, 0], device_decode=True)
    assert sa == sb and a.shape == b.shape
    assert torch.equal(a, b), (int((a != b).sum()), [int((a[i] != b[i]).sum()) for i in range(len(roidb))])
    only_files = [roidb[0], roidb[1], roidb[3]]
    c, _ = minibatch._get_image_blob_device(only_files, [0, 0, 0], device_decode=True)
    assert torch.equal(c, a[[0, 1, 3]]), [int((c[j] != a[i]).sum()) for j, i in enumerate((0, 1, 3))]


@pytest.mark.parametrize("threads", [0, 1, 3])
def test_batches_larger_than_the_staging_ring(cuda, threads):
    """The decoder stages 128 images at a time in a ring (8 chunks of 16): batches beyond that wrap around while earlier
    chunks are still uploading; with 1 host thread the calling thread decodes inline."""
    from image_denoising_b200 import jpeg
    base = [encode(synth_image(48, 80, 70 + i, "smooth" if i % 3 else "uniform"), 60 + 3 * i, [S420, S444, S422][i % 3], i % 4)
            for i in range(11)]
    want = [cv2.imdecode(np.frombuffer(d, np.uint8), cv2.IMREAD_COLOR) for d in base]
    n = 437
    dec = jpeg.Decoder(n, 48, 80, cuda, host_threads=threads)
    for rep in range(2):
        got = dec.decode([base[(i * 7 + rep) % 11] for i in range(n)]).cpu().numpy()
        for i in range(n):
            assert np.array_equal(got[i], want[(i * 7 + rep) % 11]), (threads, rep, i)
    dec.close()


def test_test_side_chain_on_device(cuda, tmp_path):
    """test_hook.corrupt_paths_for_test_device = [corrupt_for_test(cv2.imread(p), noise, i) for p in paths] (test.py:191-1831)."""
    from image_denoising_b200 import test_hook
    paths = []
    for i in range(5):
        p = str(tmp_path / ("t%d.jpg" % i))
        with open(p, "wb") as f:
            f.write(encode(synth_image(96, 128, 40 + i, "smooth"), 88, S420))
        paths.append(p)
    for noise in ("periodic_median_var100", "shader", "sap_mean_var0.4"):
        dev = test_hook.corrupt_paths_for_test_device(paths, noise, first_index=7).cpu().numpy()
        for i, p in enumerate(paths):
            want = test_hook.corrupt_for_test(cv2.imread(p), noise, 7 + i)
            assert np.array_equal(dev[i], want), (noise, i)
    dets = np.array([[0, 0, 9, 9, 0.9], [0, 0, 9, 9, 0.8], [5, 0, 14, 9, 0.7], [100, 100, 109, 109, 0.6]], np.float32)
    assert [int(k) for k in test_hook.nms_for_test(dets)] == [0, 3]               # cfg.TEST.NMS = 0.3
