"""Whole preprocessing chain of one training minibatch, files in -> network blob out (SURVEY 8(a) a1 with the 8(f) f2 and f1
stages either side of the path), against the reference's chain on one host core:

  reference (lib/roi_data_layer/minibatch.py:77-1690, one image per call):
      cv2.imread -> add_<noise>_noise(spec) closure -> prep_im_for_blob -> im_list_to_blob           (oracle transcription)
  here (one call per batch):
      file bytes -> b2d_jpeg_decode -> b2d_run(spec) -> b2d_blob_prep -> (N, H', W', 3) float32 on the GPU

  python tests/perf/bench_pipeline.py [--n 256 --batch 64 --spec gaussian_median_var1.0 ...]"""
import argparse
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cv2  # noqa: E402
import numpy as np  # noqa: E402
import torch  # noqa: E402

from conftest import synth_image  # noqa: E402
from image_denoising_b200 import minibatch  # noqa: E402
from image_denoising_b200.blob import im_list_to_blob, prep_im_for_blob  # noqa: E402
from image_denoising_b200.config import cfg  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=256)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--h", type=int, default=375)
    ap.add_argument("--w", type=int, default=500)
    ap.add_argument("--cpu-images", type=int, default=8)
    ap.add_argument("specs", nargs="*", default=["gaussian_median_var1.0", "sap_mean_var0.4", "speckle_bilateral_var1.0",
                                                 "gaussian_wavelet_var1.5"])
    a = ap.parse_args()
    tmp = tempfile.mkdtemp(prefix="b2d_pipe_")
    paths = []
    for i in range(a.n):
        p = os.path.join(tmp, "%06d.jpg" % i)
        if i < 8:
            cv2.imwrite(p, synth_image(a.h, a.w, 100 + i, "smooth"), [cv2.IMWRITE_JPEG_QUALITY, 90])
        else:
            with open(paths[i % 8], "rb") as f, open(p, "wb") as g:
                g.write(f.read())
        paths.append(p)
    for spec in a.specs:
        roidb = [{'image': p, 'noise_type': spec, 'flipped': bool(i & 1), 'index': i} for i, p in enumerate(paths)]
        scale = [0] * a.batch

        def run_all():
            for first in range(0, a.n, a.batch):
                blob, _ = minibatch._get_image_blob_device(roidb[first:first + a.batch], scale, device_decode=True)
            return blob
        run_all()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        blob = run_all()
        torch.cuda.synchronize()
        gpu_ms = (time.perf_counter() - t0) * 1e3 / a.n
        # the reference's chain, one image at a time on one core (cv2 threads = 1), oracle transcription of the closure
        from oracle import closures as C
        cv2.setNumThreads(1)
        np.random.seed(cfg.RNG_SEED)
        t0 = time.perf_counter()
        for i in range(a.cpu_images):
            img = cv2.imread(paths[i])
            im = C.corrupt(img, spec)
            if roidb[i]['flipped']:
                im = im[:, ::-1, :]
            im, s = prep_im_for_blob(im, cfg.PIXEL_MEANS, cfg.TRAIN.SCALES[0], cfg.TRAIN.MAX_SIZE)
            im_list_to_blob([im])
        cpu_ms = (time.perf_counter() - t0) * 1e3 / a.cpu_images
        cv2.setNumThreads(-1)
        print(json.dumps({"spec": spec, "images": a.n, "batch": a.batch, "shape": [a.h, a.w], "blob": list(blob.shape[1:]),
                          "b200_ms_per_image": round(gpu_ms, 4), "b200_images_s": round(1e3 / gpu_ms, 1),
                          "reference_chain_ms_per_image_1core": round(cpu_ms, 2), "cpu_images_timed": a.cpu_images,
                          "ratio_vs_1core": round(cpu_ms / gpu_ms, 1), "host_cores": os.cpu_count()}), flush=True)
    for p in paths:
        os.remove(p)
    os.rmdir(tmp)


if __name__ == "__main__":
    main()
