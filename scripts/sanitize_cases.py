#!/usr/bin/env python
"""Small-shape sweep of every kernel family through the C ABI, meant to run under compute-sanitizer
(scripts/sanitize.sh).  Shapes hit the partial-word stores at row ends (W % 4 != 0), rows of 4 (mod 8) bytes, the
marching kernels' double-buffered rings, apron regeneration across CTA borders (tiles wider / taller than one CTA),
images smaller than a tile, the multi-chunk b2d_run_host pipeline, and the static-scratch NMS entry point.
Results are also compared between b2d_run and b2d_run_host so that a race shows up as a mismatch too."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from conftest import synth_batch  # noqa: E402
from image_denoising_b200 import api, nms  # noqa: E402

SHAPES = [(2, 37, 53), (2, 50, 100), (1, 61, 99), (2, 64, 96), (1, 130, 1368), (1, 100, 500), (3, 24, 333)]
SPECS3 = ["sap_mean_var0.4", "gaussian_gaus_blur_var1.5", "speckle_median_var1.0", "uniform_mean_var0.6", "rayleigh_median_var0.2",
          "gamma_gaus_blur_var0.1", "periodic_mean_var100", "shader", "original_gaus_blur", "original_mean", "original_median",
          "sap_var0.4", "gaussian_var1.0", "poisson_median", "poisson", "brownian_mean_var0.09", "brownian_var0.9"]
HEAVY = [("original_median", {"ksize": 5}), ("original_median", {"ksize": 7}), ("gaussian_median_var1.0", {"ksize": 7}),
         ("original_bilateral", {}), ("speckle_bilateral_var1.0", {}), ("original_bilateral", {"d": 5}),
         ("gaussian_wavelet_var1.5", {}), ("original_wavelet", {"wavelet": "db1", "levels": 3}),
         ("gaussian_wavelet_var1.5", {"wavelet": "db1", "levels": 3}), ("poisson_wavelet", {}), ("quant_median_var7", {}),
         ("quant_wavelet_var3", {}), ("quant_var10", {})]


def main():
    dev = torch.device("cuda", 0)
    n_runs = 0
    for n, h, w in SHAPES:
        batch = torch.from_numpy(synth_batch(n, h, w, seed=h + w, kind="smooth")).to(dev)
        for spec in SPECS3:
            api.apply(batch, spec, seed=7)
            n_runs += 1
        for spec, kw in HEAVY:
            if "wavelet" in spec and kw.get("wavelet") == "db1" and False:
                continue
            api.apply(batch, spec, seed=7, **kw)
            n_runs += 1
    torch.cuda.synchronize()
    # bloom needs H >= 300
    b = torch.from_numpy(synth_batch(2, 304, 320, seed=3, kind="smooth")).to(dev)
    api.apply(b, "bloom", seed=5)
    # host pipeline with many chunks (B2D_HOST_CHUNK_MB=1 from sanitize.sh) vs the device call
    hb = synth_batch(12, 304, 320, seed=9, kind="smooth")
    for spec in ("sap_mean_var0.4", "poisson_median", "gaussian_bilateral_var1.0", "bloom", "brownian_mean_var0.09", "quant_median_var7",
                 "gaussian_wavelet_var1.0"):
        plan = api.Plan(12, 304, 320, spec)
        want = plan.run(torch.from_numpy(hb).to(dev), seed=11, first_index=3).cpu().numpy()
        got = plan.run_host(hb, seed=11, first_index=3)
        assert np.array_equal(got, want), spec
        n_runs += 2
    # NMS: host entry, device entry, static-scratch entry (twice)
    from nms_cases import make_dets
    dets = make_dets(700, seed=5, style="clustered")
    keep = nms.nms(dets, 0.5)
    order = dets[:, 4].argsort()[::-1]
    srt = torch.from_numpy(np.ascontiguousarray(dets[order])).to(dev)
    kd = nms.nms_device(srt, 0.5).cpu().numpy()
    assert [int(order[k]) for k in kd] == [int(k) for k in keep]
    import ctypes
    from image_denoising_b200 import _lib
    ko = torch.empty(700, dtype=torch.int32, device=dev)
    num = torch.zeros(1, dtype=torch.int32, device=dev)
    for _ in range(2):
        _lib.check(_lib.lib().b2d_nms_device_static(ctypes.c_void_p(srt.data_ptr()), 700, 5, 0.5, ctypes.c_void_p(ko.data_ptr()),
                                                    ctypes.c_void_p(num.data_ptr()), None))
        torch.cuda.synchronize()
        assert np.array_equal(ko[: int(num.item())].cpu().numpy(), kd)
    # blob stage
    from image_denoising_b200.blob import prep_blob_device
    prep_blob_device(torch.from_numpy(synth_batch(2, 75, 100, seed=2)).to(dev), np.array([[[102.9801, 115.9465, 122.7717]]]), 600, 1000)
    torch.cuda.synchronize()
    print("sanitize_cases: %d runs ok" % n_runs)


if __name__ == "__main__":
    main()
