#!/usr/bin/env python
"""Generates image_denoising_b200/csrc/b2d_lab_tables.h: the lookup tables of OpenCV's 8-bit
(fixed point) BGR<->Lab conversion, the arithmetic `cv2.cvtColor(img, COLOR_BGR2LAB / COLOR_LAB2BGR)`
runs on uint8 images (reference call sites lib/roi_data_layer/minibatch.py:496,507).

The tables are DATA: float formulas of the published algorithm evaluated once here, then pinned
against the cv2 installed in this container (the oracle library) over ALL 2^24 inputs in each
direction.  Two places need the library's own rounding and are read off its observable behaviour:
  * cube-root table entry 324 (OpenCV's software cbrt lands on the other side of a .5);
  * the 4096-entry inverse sRGB gamma table (73 entries differ by one from libm's pow): the
    conversion output is a pure function of the table index, so the table is read back exactly.
Run:  python scripts/gen_lab_tables.py      (needs cv2; about a minute)
"""
import os
import sys

import cv2
import numpy as np

f32 = np.float32
BASE = 1 << 14
MIN_AB = -8145
M_FWD = np.array([0.412453, 0.357580, 0.180423, 0.212671, 0.715160, 0.072169, 0.019334, 0.119193, 0.950227]).reshape(3, 3)
M_INV = np.array([3.240479, -1.53715, -0.498535, -0.969256, 1.875991, 0.041556, 0.055648, -0.204043, 1.057311]).reshape(3, 3)
WHITE = np.array([0.950456, 1.0, 1.088754])


def ds(x, n):
    return (x + (1 << (n - 1))) >> n


def forward_tables():
    i = np.arange(256, dtype=np.float32)
    x = i / f32(255)
    g = np.where(x <= f32(0.04045), x / f32(12.92), np.power((x + f32(0.055)) / f32(1.055), f32(2.4))).astype(np.float32)
    gamma = np.rint(f32(255 * 8) * g).astype(np.int64)
    j = np.arange(256 * 3 // 2 * 8, dtype=np.float32)
    xx = j / f32(255 * 8)
    c = np.where(xx < f32(216 / 24389), xx * f32(841 / 108) + f32(16 / 116),
                 np.cbrt(xx.astype(np.float64)).astype(np.float32)).astype(np.float32)
    cbrt = np.rint(f32(32768) * c).astype(np.int64)
    cbrt[324] -= 1                      # pinned against cv2 (see module docstring)
    C = np.zeros((3, 3), np.int64)      # columns in B, G, R order of the source pixel
    for r in range(3):
        C[r, 2], C[r, 1], C[r, 0] = (int(np.rint(4096 * M_FWD[r, k] / WHITE[r])) for k in range(3))
    return gamma, cbrt, C


def bgr2lab(img, gamma, cbrt, C):
    s = gamma[img.astype(np.int64)]
    f = [cbrt[ds(s[..., 0] * C[r, 0] + s[..., 1] * C[r, 1] + s[..., 2] * C[r, 2], 12)] for r in range(3)]
    L = ds(296 * f[1] - 1336934, 15)
    a = ds(500 * (f[0] - f[1]) + 128 * 32768, 15)
    b = ds(200 * (f[1] - f[2]) + 128 * 32768, 15)
    return np.clip(np.stack([L, a, b], -1), 0, 255).astype(np.uint8)


def inverse_tables():
    i = np.arange(256, dtype=np.float32)
    li = i * f32(100) / f32(255)
    y_lo = li / f32(903.3)
    fy_lo = f32(7.787) * y_lo + f32(16) / f32(116)
    fy_hi = (li + f32(16)) / f32(116)
    y_hi = fy_hi * fy_hi * fy_hi
    y = np.where(li <= f32(8), y_lo, y_hi).astype(np.float32)
    fy = np.where(li <= f32(8), fy_lo, fy_hi).astype(np.float32)
    yf = np.stack([np.rint(f32(BASE) * y), np.rint(f32(BASE) * fy)], -1).astype(np.int64)   # LabToYF_b
    C = np.zeros((3, 3), np.int64)      # rows R, G, B
    for r in range(3):
        for k in range(3):
            C[r, k] = int(np.rint(4096 * M_INV[r, k] * WHITE[k]))
    return yf, C


def ab_to_xz(i):
    """abToXZ_b as a function (C integer division: truncation toward zero)."""
    def cdiv(a, b):
        return np.sign(a) * (np.abs(a) // b)
    lo = cdiv(i * 108, 841) - BASE * 16 // 116 * 108 // 841
    hi = cdiv(cdiv(i * i, BASE) * i, BASE)
    return np.where(i <= 3390, lo, hi)


def lab2rgb_index(lab, yf, C):
    """The three 12-bit indices (b, g, r order) into the inverse gamma table."""
    L, a, b = (lab[..., k].astype(np.int64) for k in range(3))
    y, ify = yf[L, 0], yf[L, 1]
    adiv = ((5 * a * 53687 + (1 << 7)) >> 13) - 128 * BASE // 500
    bdiv = ((b * 41943 + (1 << 4)) >> 9) - 128 * BASE // 200 + 1
    x, z = ab_to_xz(ify + adiv), ab_to_xz(ify - bdiv)
    rgb = [np.clip(ds(C[r, 0] * x + C[r, 1] * y + C[r, 2] * z, 14), 0, 4095) for r in range(3)]
    return np.stack([rgb[2], rgb[1], rgb[0]], -1)


def all_colours():
    a, b, c = np.meshgrid(np.arange(256), np.arange(256), np.arange(256), indexing="ij")
    return np.stack([a, b, c], -1).astype(np.uint8).reshape(4096, 4096, 3)


def fmt(name, ctype, arr, per=16):
    vals = [str(int(v)) for v in arr]
    lines = [", ".join(vals[i:i + per]) for i in range(0, len(vals), per)]
    return "B2D_LAB_TABLE %s %s[%d] = {\n  %s};\n" % (ctype, name, len(vals), ",\n  ".join(lines))


def main():
    cube = all_colours()
    gamma, cbrt, Cf = forward_tables()
    ref = cv2.cvtColor(cube, cv2.COLOR_BGR2LAB)
    bad = int((bgr2lab(cube, gamma, cbrt, Cf) != ref).any(-1).sum())
    print("BGR2LAB mismatching colours of 2^24:", bad)
    assert bad == 0
    yf, Ci = inverse_tables()
    idx = lab2rgb_index(cube, yf, Ci).reshape(-1, 3)
    ref = cv2.cvtColor(cube, cv2.COLOR_LAB2BGR).reshape(-1, 3)
    inv_gamma = np.full(4096, -1, np.int64)
    for ch in range(3):                                     # read the table back; must be a function of the index
        order = np.argsort(idx[:, ch], kind="stable")
        si, sr = idx[order, ch], ref[order, ch].astype(np.int64)
        starts = np.r_[0, np.nonzero(np.diff(si))[0] + 1]
        mn, mx = np.minimum.reduceat(sr, starts), np.maximum.reduceat(sr, starts)
        assert (mn == mx).all(), "inverse gamma is not a function of the index"
        cur = inv_gamma[si[starts]]
        assert ((cur == -1) | (cur == mn)).all()
        inv_gamma[si[starts]] = mn
    assert (inv_gamma >= 0).all() and (np.diff(inv_gamma) >= 0).all()
    bad = int((inv_gamma[idx] != ref).any(-1).sum())
    print("LAB2BGR mismatching colours of 2^24:", bad)
    assert bad == 0
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "image_denoising_b200", "csrc",
                       "b2d_lab_tables.h")
    with open(out, "w") as f:
        f.write("// b2d_lab_tables.h - GENERATED by scripts/gen_lab_tables.py; do not edit.\n"
                "// Lookup tables of OpenCV's 8-bit fixed-point BGR<->Lab conversion (cv2 %s), pinned\n"
                "// against cv2.cvtColor over all 2^24 inputs in both directions.\n"
                "#pragma once\n#include <stdint.h>\n#ifndef B2D_LAB_TABLE\n#define B2D_LAB_TABLE static const\n#endif\n\n"
                % cv2.__version__)
        f.write("// sRGBGammaTab_b: round(2040 * gamma(i / 255))\n" + fmt("kLabGamma", "uint16_t", gamma))
        f.write("// LabCbrtTab_b: round(32768 * f(i / 2040)), f = cube root with the linear toe\n" + fmt("kLabCbrt", "uint16_t", cbrt))
        f.write("// forward matrix, 12-bit fixed point, columns in source order B, G, R; rows X, Y, Z\n" + fmt("kLabFwd", "int32_t", Cf.ravel(), 9))
        f.write("// LabToYF_b: (2^14 * y, 2^14 * fy) per L byte\n" + fmt("kLabYF", "uint16_t", yf.ravel()))
        f.write("// inverse matrix, 12-bit fixed point, rows R, G, B; columns X, Y, Z\n" + fmt("kLabInv", "int32_t", Ci.ravel(), 9))
        f.write("// sRGBInvGammaTab_b: 12-bit linear -> uint8 sRGB\n" + fmt("kLabInvGamma", "uint8_t", inv_gamma, 32))
    print("wrote", out)
    return 0


if __name__ == "__main__":
    sys.exit(main())
