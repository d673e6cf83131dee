"""CPU baseline of the five BASELINE.json configurations: the reference's closures
(oracle transcription over the reference's own cv2 / NumPy / SciPy / sklearn calls) one image per call in ONE Python process,
as the reference runs them (cfg.TRAIN.IMS_PER_BATCH = 1, lib/model/config.py:69; loop lib/model/test.py:189), on the GPU
box's host cores.  C1 is timed in full, C2-C5 on a subsample and stated as such; each config once with cv2's own thread pool
and once with cv2.setNumThreads(1).  Image decode excluded (synthetic arrays in memory).

  python tests/perf/cpu_configs.py [--budget 25]     # seconds of CPU work per config and thread setting (upper bound)"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import cv2  # noqa: E402
import numpy as np  # noqa: E402

from conftest import synth_image  # noqa: E402
from oracle import closures as C  # noqa: E402

CONFIGS = [
    ("C1", 16, 600, 1000, ["gaussian_median_var1.0"], {}),
    ("C2", 512, 600, 1000, ["sap_mean_var0.4", "gaussian_gaus_blur_var1.5"], {}),
    ("C3", 1024, 1080, 1920, ["speckle_bilateral_var1.0"], {}),
    ("C4", 256, 2048, 2048, ["gaussian_wavelet_var1.5"], {"wavelet": "db1", "wavelet_levels": 3}),
    ("C5", 4096, 600, 1000, None, {"median_k": 7}),
]


def c5_specs(n, seed=3):          # the per-image draw of scripts/bench_configs.py::c5_specs
    levels = {"periodic": ["3.14", "100", "size"], "brownian": ["0.9", "0.09", "0.009"], "quant": ["3", "7", "10"]}
    out = []
    for i in range(n):
        hsh = (i * 2654435761 + seed * 40503) & 0xffffffff
        noise = ["poisson", "periodic", "brownian", "quant"][(hsh >> 8) % 4]
        out.append(noise + "_median" + ("" if noise == "poisson" else "_var" + levels[noise][(hsh >> 16) % 3]))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--budget", type=float, default=25.0)
    ap.add_argument("--max-images", type=int, default=16)
    a = ap.parse_args()
    import scipy
    import sklearn
    info = {"cpu_count": os.cpu_count(), "cv2": cv2.__version__, "numpy": np.__version__, "scipy": scipy.__version__,
            "sklearn": sklearn.__version__, "cv2_threads_default": cv2.getNumThreads()}
    print(json.dumps({"host": info}), flush=True)
    for name, n, h, w, specs, kw in CONFIGS:
        per_image_specs = c5_specs(n) if specs is None else None
        sample = min(n, a.max_images)
        imgs = [synth_image(h, w, 1000 + int(name[1]) * 100 + i, "smooth") for i in range(min(sample, 4))]
        for threads in (-1, 1):
            cv2.setNumThreads(threads)
            np.random.seed(3)
            t_all, done = 0.0, 0
            C.corrupt(imgs[0], (per_image_specs or specs)[0], **kw)          # warm-up (imports, cv2 pools)
            for i in range(sample):
                todo = [per_image_specs[i]] if per_image_specs else specs
                t0 = time.perf_counter()
                for s in todo:
                    C.corrupt(imgs[i % len(imgs)], s, **kw)
                t_all += time.perf_counter() - t0
                done += 1
                if t_all > a.budget:
                    break
            passes = 1 if per_image_specs else len(specs)
            mpix = done * h * w * passes / 1e6
            print(json.dumps({"config": name, "images_timed": done, "of": n, "shape": [h, w],
                              "specs": specs or "mixed poisson/periodic/brownian/quant -> median 7x7 (first %d of the 4096 draws)" % done,
                              "cv2_threads": cv2.getNumThreads(), "s_per_image": round(t_all / done, 4),
                              "mpix_s": round(mpix / t_all, 2), "extrapolated_full_config_s": round(t_all / done * n, 1),
                              "process": "single Python process, one image per call"}), flush=True)
    cv2.setNumThreads(-1)


if __name__ == "__main__":
    main()
