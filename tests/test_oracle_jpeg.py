"""CPU: (1) the JPEG oracle (oracle/jpeg.py) is PINNED against cv2.imdecode = libjpeg-turbo, byte for byte, on streams of
several sizes, qualities, chroma samplings and restart intervals; (2) the host Huffman stage of libb200denoise.so
(b2d_jpeg_entropy, no GPU needed) yields exactly the oracle's quantised coefficients; (3) refused streams fail loudly."""
import cv2
import numpy as np
import pytest

from conftest import synth_image
from oracle import jpeg as J

S444, S422, S420 = (cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_422,
                    cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420)


def encode(img, quality=90, sampling=None, rst=0, extra=()):
    params = [cv2.IMWRITE_JPEG_QUALITY, quality]
    if sampling is not None:
        params += [cv2.IMWRITE_JPEG_SAMPLING_FACTOR, sampling]
    if rst:
        params += [cv2.IMWRITE_JPEG_RST_INTERVAL, rst]
    ok, buf = cv2.imencode(".jpg", img, params + list(extra))
    assert ok
    return buf.tobytes()


CASES = [((64, 96), "smooth", 90, None, 0), ((75, 101), "smooth", 50, S444, 0), ((75, 101), "smooth", 75, S422, 0),
         ((37, 53), "uniform", 30, S420, 0), ((64, 64), "uniform", 95, S420, 3), ((120, 160), "smooth", 85, S420, 7),
         ((8, 8), "uniform", 90, S420, 0), ((1, 1), "uniform", 90, S420, 0), ((17, 3), "uniform", 100, S420, 0),
         ((16, 33), "smooth", 10, S422, 2), ((121, 47), "uniform", 60, S444, 5)]


def streams():
    for shape, kind, q, samp, rst in CASES:
        yield (shape, kind, q, samp, rst), encode(synth_image(shape[0], shape[1], q, kind), q, samp, rst)
    yield "grey", encode(synth_image(40, 56, 2, "smooth")[..., 0], 80)


@pytest.mark.parametrize("tag,data", list(streams()), ids=lambda v: str(v) if not isinstance(v, bytes) else "")
def test_oracle_equals_cv2_imdecode(tag, data):
    want = cv2.imdecode(np.frombuffer(data, np.uint8), cv2.IMREAD_COLOR)
    got = J.decode(data)
    assert got.shape == want.shape and np.array_equal(got, want), (tag, int((got != want).sum()))


@pytest.mark.parametrize("tag,data", list(streams()), ids=lambda v: str(v) if not isinstance(v, bytes) else "")
def test_host_huffman_stage_equals_oracle(tag, data):
    from image_denoising_b200 import jpeg
    _, _, want = J.entropy_decode(data)
    got = jpeg.entropy_decode(data)
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g.shape == w.shape and np.array_equal(g.astype(np.int32), w)
    h, w_, nc, hs, vs = jpeg.info(data)
    frame = J.parse(data)[0]
    assert (h, w_, nc) == (frame["h"], frame["w"], len(frame["comps"]))


def test_unsupported_streams_fail_loudly():
    from image_denoising_b200 import _lib, jpeg
    img = synth_image(32, 48, 1, "smooth")
    prog = encode(img, 90, None, 0, (cv2.IMWRITE_JPEG_PROGRESSIVE, 1))
    with pytest.raises(_lib.B2DError) as e:
        jpeg.entropy_decode(prog)
    assert e.value.code == _lib.ERR_UNSUPPORTED
    with pytest.raises(J.Unsupported):
        J.decode(prog)
    with pytest.raises(_lib.B2DError):
        jpeg.info(b"\x89PNG\r\n\x1a\n" + b"\0" * 32)
    good = encode(img, 90)
    with pytest.raises(_lib.B2DError):
        jpeg.entropy_decode(good[:len(good) // 3])           # truncated inside the headers / tables


def _insert_after_soi(data, marker, payload):
    import struct
    return data[:2] + bytes([0xFF, marker]) + struct.pack(">H", len(payload) + 2) + payload + data[2:]


def test_exif_orientation_and_rgb_streams_are_refused():
    """cv2.imread applies the EXIF orientation (checked here against cv2 itself) and libjpeg does not colour-convert
    RGB-coded streams: the device decoder, which does neither, must refuse such files instead of returning other bytes."""
    import struct
    from image_denoising_b200 import _lib, jpeg
    img = synth_image(24, 40, 4, "smooth")
    base = encode(img, 90)
    tiff = b"II*\x00" + struct.pack("<I", 8) + struct.pack("<H", 1) + struct.pack("<HHII", 0x0112, 3, 1, 6) + struct.pack("<I", 0)
    rotated = _insert_after_soi(base, 0xE1, b"Exif\x00\x00" + tiff)
    plain = cv2.imdecode(np.frombuffer(base, np.uint8), cv2.IMREAD_COLOR)
    turned = cv2.imdecode(np.frombuffer(rotated, np.uint8), cv2.IMREAD_COLOR)
    assert plain.shape == (24, 40, 3) and turned.shape == (40, 24, 3)           # orientation 6: cv2 rotates
    with pytest.raises(_lib.B2DError) as e:
        jpeg.entropy_decode(rotated)
    assert e.value.code == _lib.ERR_UNSUPPORTED
    upright = _insert_after_soi(base, 0xE1, b"Exif\x00\x00" + tiff.replace(struct.pack("<HHII", 0x0112, 3, 1, 6), struct.pack("<HHII", 0x0112, 3, 1, 1)))
    assert len(jpeg.entropy_decode(upright)) == 3                                # orientation 1 decodes
    # Adobe APP14 with transform 0 and no JFIF segment = RGB-coded
    assert base[2:4] == b"\xff\xe0"
    jfif_len = struct.unpack(">H", base[4:6])[0]
    no_jfif = base[:2] + base[4 + jfif_len:]
    adobe = _insert_after_soi(no_jfif, 0xEE, b"Adobe\x00\x64\x00\x00\x00\x00\x00")
    with pytest.raises(_lib.B2DError) as e:
        jpeg.entropy_decode(adobe)
    assert e.value.code == _lib.ERR_UNSUPPORTED
    ycc = _insert_after_soi(no_jfif, 0xEE, b"Adobe\x00\x64\x00\x00\x00\x00\x01")
    want = cv2.imdecode(np.frombuffer(ycc, np.uint8), cv2.IMREAD_COLOR)
    assert np.array_equal(J.decode(ycc), want) and len(jpeg.entropy_decode(ycc)) == 3
