"""Oracle: baseline JPEG decode as ``cv2.imread`` performs it (TEST INFRASTRUCTURE ONLY).

The reference decodes every image with ``cv2.imread`` (lib/model/test.py:191,
lib/roi_data_layer/minibatch.py:85 and again inside most closures); OpenCV delegates to
libjpeg-turbo with the library defaults: ``JDCT_ISLOW``, ``do_fancy_upsampling``,
YCbCr -> BGR.  This file restates that pipeline in NumPy:

  * marker parsing + sequential Huffman decoding            (ITU T.81 F.2.2, jdhuff.c)
  * dequantisation + the 8x8 "islow" integer IDCT           (jidctint.c, CONST_BITS 13, PASS1_BITS 2)
  * h2v2 / h2v1 "fancy" (triangle) chroma upsampling        (jdsample.c)
  * YCbCr -> RGB with the 16-bit fixed-point tables         (jdcolor.c)

PINNED: ``decode(data)`` is compared byte for byte with ``cv2.imdecode`` (libjpeg-turbo
3.1.2 inside cv2 4.13) in tests/test_oracle_jpeg.py on streams of several sizes, qualities,
samplings and restart intervals, and the CUDA decoder is held to both.
"""
import struct

import numpy as np

ZIGZAG = np.array([0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7,
                   14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39,
                   46, 53, 60, 61, 54, 47, 55, 62, 63])


class Unsupported(ValueError):
    pass


def parse(data):
    """Markers up to SOS.  Returns the frame description and the offset of the entropy-coded data."""
    if data[:2] != b"\xff\xd8":
        raise Unsupported("not a JPEG stream")
    qt, huff, frame, dri = {}, {}, None, 0
    i = 2
    while i < len(data):
        if data[i] != 0xFF:
            raise Unsupported("marker expected at %d" % i)
        while data[i + 1] == 0xFF:
            i += 1
        m = data[i + 1]
        if m in (0xD8, 0x01) or 0xD0 <= m <= 0xD7:
            i += 2
            continue
        L = struct.unpack(">H", data[i + 2:i + 4])[0]
        seg = data[i + 4:i + 2 + L]
        if m == 0xDB:
            k = 0
            while k < len(seg):
                pq, tq = seg[k] >> 4, seg[k] & 15
                if pq:
                    vals = np.frombuffer(seg[k + 1:k + 129], ">u2").astype(np.int32)
                    k += 129
                else:
                    vals = np.frombuffer(seg[k + 1:k + 65], np.uint8).astype(np.int32)
                    k += 65
                nat = np.zeros(64, np.int32)
                nat[ZIGZAG] = vals
                qt[tq] = nat
        elif m == 0xC4:
            k = 0
            while k < len(seg):
                tc, th = seg[k] >> 4, seg[k] & 15
                counts = list(seg[k + 1:k + 17])
                n = sum(counts)
                huff[(tc, th)] = (counts, list(seg[k + 17:k + 17 + n]))
                k += 17 + n
        elif m == 0xC0 or m == 0xC1:
            p, h, w, n = struct.unpack(">BHHB", seg[:6])
            if p != 8:
                raise Unsupported("sample precision %d" % p)
            comps = [(seg[6 + 3 * c], seg[7 + 3 * c] >> 4, seg[7 + 3 * c] & 15, seg[8 + 3 * c]) for c in range(n)]
            frame = {"h": h, "w": w, "comps": comps}
        elif m in (0xC2, 0xC3, 0xC5, 0xC6, 0xC7, 0xC9, 0xCA, 0xCB, 0xCD, 0xCE, 0xCF):
            raise Unsupported("SOF%d (only baseline sequential Huffman)" % (m - 0xC0))
        elif m == 0xDD:
            dri = struct.unpack(">H", seg[:2])[0]
        elif m == 0xDA:
            ns = seg[0]
            scan = [(seg[1 + 2 * c], seg[2 + 2 * c] >> 4, seg[2 + 2 * c] & 15) for c in range(ns)]
            if frame is None or ns != len(frame["comps"]):
                raise Unsupported("non-interleaved scan")
            return frame, qt, huff, dri, scan, i + 2 + L
        i += 2 + L
    raise Unsupported("no SOS")


def _huff_table(counts, symbols):
    """code -> symbol lookup per length (T.81 C.2)."""
    table, code, k = {}, 0, 0
    for length in range(1, 17):
        for _ in range(counts[length - 1]):
            table[(length, code)] = symbols[k]
            code += 1
            k += 1
        code <<= 1
    return table


class _Bits:
    def __init__(self, data, pos):
        self.d, self.p, self.acc, self.n = data, pos, 0, 0

    def bit(self):
        if self.n == 0:
            b = self.d[self.p]
            self.p += 1
            if b == 0xFF:
                if self.d[self.p] == 0:
                    self.p += 1
                else:                       # a marker: feed zeros (libjpeg does the same) and stay put
                    self.p -= 1
                    b = 0
            self.acc, self.n = b, 8
        self.n -= 1
        return (self.acc >> self.n) & 1

    def bits(self, k):
        v = 0
        for _ in range(k):
            v = (v << 1) | self.bit()
        return v

    def restart(self):
        self.n = 0
        while not (self.d[self.p] == 0xFF and 0xD0 <= self.d[self.p + 1] <= 0xD7):
            self.p += 1
        self.p += 2


def _decode_sym(br, table):
    code = 0
    for length in range(1, 17):
        code = (code << 1) | br.bit()
        s = table.get((length, code))
        if s is not None:
            return s
    raise Unsupported("bad Huffman code")


def _extend(v, t):
    return v - (1 << t) + 1 if t and v < (1 << (t - 1)) else v


def entropy_decode(data):
    """-> (frame, quantisation tables per component, [coefficients (by, bx, 64) int32 in natural order per component])."""
    frame, qt, huff, dri, scan, pos = parse(data)
    comps = frame["comps"]
    hmax = max(c[1] for c in comps)
    vmax = max(c[2] for c in comps)
    mcux = -(-frame["w"] // (8 * hmax))
    mcuy = -(-frame["h"] // (8 * vmax))
    coefs = [np.zeros((mcuy * c[2], mcux * c[1], 64), np.int32) for c in comps]
    tabs = {k: _huff_table(*v) for k, v in huff.items()}
    br = _Bits(data, pos)
    pred = [0] * len(comps)
    count = 0
    for my in range(mcuy):
        for mx in range(mcux):
            if dri and count and count % dri == 0:
                br.restart()
                pred = [0] * len(comps)
            count += 1
            for ci, (cid, hs, vs, tq) in enumerate(comps):
                _, td, ta = scan[ci]
                dc_t, ac_t = tabs[(0, td)], tabs[(1, ta)]
                for by in range(vs):
                    for bx in range(hs):
                        blk = coefs[ci][my * vs + by, mx * hs + bx]
                        t = _decode_sym(br, dc_t)
                        pred[ci] += _extend(br.bits(t), t) if t else 0
                        blk[0] = pred[ci]
                        k = 1
                        while k < 64:
                            rs = _decode_sym(br, ac_t)
                            r, s = rs >> 4, rs & 15
                            if s == 0:
                                if r != 15:
                                    break
                                k += 16
                                continue
                            k += r
                            blk[ZIGZAG[k]] = _extend(br.bits(s), s)
                            k += 1
    return frame, [qt[c[3]] for c in comps], coefs


# ------------------------------------------------------------------ jidctint.c
CONST_BITS, PASS1_BITS = 13, 2
F_0_298631336, F_0_390180644, F_0_541196100, F_0_765366865 = 2446, 3196, 4433, 6270
F_0_899976223, F_1_175875602, F_1_501321110, F_1_847759065 = 7373, 9633, 12299, 15137
F_1_961570560, F_2_053119869, F_2_562915447, F_3_072711026 = 16069, 16819, 20995, 25172


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def _idct_1d(v, shift, first):
    """One pass of jpeg_idct_islow over the LAST axis of v (..., 8), int64."""
    z2, z3 = v[..., 2], v[..., 6]
    z1 = (z2 + z3) * F_0_541196100
    tmp2 = z1 + z3 * (-F_1_847759065)
    tmp3 = z1 + z2 * F_0_765366865
    z2, z3 = v[..., 0], v[..., 4]
    tmp0 = (z2 + z3) << CONST_BITS
    tmp1 = (z2 - z3) << CONST_BITS
    tmp10, tmp13, tmp11, tmp12 = tmp0 + tmp3, tmp0 - tmp3, tmp1 + tmp2, tmp1 - tmp2
    tmp0, tmp1, tmp2, tmp3 = v[..., 7], v[..., 5], v[..., 3], v[..., 1]
    z1, z2, z3, z4 = tmp0 + tmp3, tmp1 + tmp2, tmp0 + tmp2, tmp1 + tmp3
    z5 = (z3 + z4) * F_1_175875602
    tmp0, tmp1, tmp2, tmp3 = tmp0 * F_0_298631336, tmp1 * F_2_053119869, tmp2 * F_3_072711026, tmp3 * F_1_501321110
    z1, z2, z3, z4 = z1 * (-F_0_899976223), z2 * (-F_2_562915447), z3 * (-F_1_961570560), z4 * (-F_0_390180644)
    z3, z4 = z3 + z5, z4 + z5
    tmp0, tmp1, tmp2, tmp3 = tmp0 + z1 + z3, tmp1 + z2 + z4, tmp2 + z2 + z3, tmp3 + z1 + z4
    out = np.stack([tmp10 + tmp3, tmp11 + tmp2, tmp12 + tmp1, tmp13 + tmp0,
                    tmp13 - tmp0, tmp12 - tmp1, tmp11 - tmp2, tmp10 - tmp3], -1)
    return _descale(out, shift)


def idct_islow(coef, quant):
    """coef (..., 64) natural order, quant (64,) -> samples (..., 8, 8) uint8."""
    w = (coef.astype(np.int64) * quant.astype(np.int64)).reshape(coef.shape[:-1] + (8, 8))
    # pass 1: columns (transform along axis -2), results scaled up by 2^PASS1_BITS
    ws = _idct_1d(np.swapaxes(w, -1, -2), CONST_BITS - PASS1_BITS, True)
    ws = np.swapaxes(ws, -1, -2)
    # pass 2: rows
    out = _idct_1d(ws, CONST_BITS + PASS1_BITS + 3, False)
    return np.clip(out + 128, 0, 255).astype(np.uint8)


def _plane(blocks):
    by, bx = blocks.shape[:2]
    return blocks.transpose(0, 2, 1, 3).reshape(by * 8, bx * 8)


# ------------------------------------------------------------------ jdsample.c
def upsample_h2v2(plane, dh, dw):
    """plane: uint8 with at least dh x dw valid samples -> (2 dh, 2 dw) uint8."""
    p = plane[:dh, :dw].astype(np.int32)
    if dw <= 2:                                    # jinit_upsampler: fancy only when downsampled_width > 2
        return np.repeat(np.repeat(p, 2, 0), 2, 1).astype(np.uint8)
    up = np.concatenate([p[:1], p[:-1]], 0)        # row above (first row duplicated)
    dn = np.concatenate([p[1:], p[-1:]], 0)        # row below (last real row duplicated)
    out = np.empty((2 * dh, 2 * dw), np.int32)
    for v, nb in ((0, up), (1, dn)):
        s = 3 * p + nb                             # column sums
        last = np.concatenate([s[:, :1], s[:, :-1]], 1)
        nxt = np.concatenate([s[:, 1:], s[:, -1:]], 1)
        even = (3 * s + last + 8) >> 4
        odd = (3 * s + nxt + 7) >> 4
        if dw == 1:
            even[:, 0] = (s[:, 0] * 4 + 8) >> 4
            odd[:, 0] = (s[:, 0] * 4 + 7) >> 4
        else:
            even[:, 0] = (s[:, 0] * 4 + 8) >> 4
            odd[:, -1] = (s[:, -1] * 4 + 7) >> 4
        out[v::2, 0::2] = even
        out[v::2, 1::2] = odd
    return out.astype(np.uint8)


def upsample_h2v1(plane, dh, dw):
    p = plane[:dh, :dw].astype(np.int32)
    if dw <= 2:
        return np.repeat(p, 2, 1).astype(np.uint8)
    last = np.concatenate([p[:, :1], p[:, :-1]], 1)
    nxt = np.concatenate([p[:, 1:], p[:, -1:]], 1)
    even = (3 * p + last + 1) >> 2
    odd = (3 * p + nxt + 2) >> 2
    even[:, 0] = p[:, 0]
    odd[:, -1] = p[:, -1]
    out = np.empty((dh, 2 * dw), np.int32)
    out[:, 0::2] = even
    out[:, 1::2] = odd
    return out.astype(np.uint8)


# ------------------------------------------------------------------- jdcolor.c
def ycc_to_bgr(y, cb, cr):
    fix = lambda v: int(v * 65536 + 0.5)  # noqa: E731
    x = np.arange(256, dtype=np.int64) - 128
    cr_r = (fix(1.40200) * x + 32768) >> 16
    cb_b = (fix(1.77200) * x + 32768) >> 16
    cr_g = -fix(0.71414) * x
    cb_g = -fix(0.34414) * x + 32768
    yy = y.astype(np.int64)
    r = yy + cr_r[cr]
    g = yy + ((cb_g[cb] + cr_g[cr]) >> 16)
    b = yy + cb_b[cb]
    return np.clip(np.stack([b, g, r], -1), 0, 255).astype(np.uint8)


def decode(data):
    """JPEG bytes -> H x W x 3 uint8 BGR (or H x W x 3 with equal channels for greyscale, as
    cv2.IMREAD_COLOR returns)."""
    frame, quants, coefs = entropy_decode(data)
    h, w, comps = frame["h"], frame["w"], frame["comps"]
    planes = [_plane(idct_islow(c, q)) for c, q in zip(coefs, quants)]
    if len(comps) == 1:
        g = planes[0][:h, :w]
        return np.stack([g, g, g], -1)
    if len(comps) != 3:
        raise Unsupported("%d components" % len(comps))
    hmax, vmax = comps[0][1], comps[0][2]
    if any(c[1] != 1 or c[2] != 1 for c in comps[1:]):
        raise Unsupported("subsampled luma / non-unit chroma factors")
    full = [planes[0]]
    for c in (1, 2):
        dh, dw = -(-h * 1 // vmax), -(-w * 1 // hmax)
        if (hmax, vmax) == (1, 1):
            full.append(planes[c])
        elif (hmax, vmax) == (2, 2):
            full.append(upsample_h2v2(planes[c], dh, dw))
        elif (hmax, vmax) == (2, 1):
            full.append(upsample_h2v1(planes[c], dh, dw))
        else:
            raise Unsupported("sampling %dx%d" % (hmax, vmax))
    return ycc_to_bgr(full[0][:h, :w], full[1][:h, :w], full[2][:h, :w])
