"""Writes tests/golden/manifest.json (+ natural_120x160.npz) - run HERE, in the container that has /root/reference:

  python tests/golden/make_manifest.py

For every case of tests/golden_cases.py: SHA-256 of the oracle's uint8 result.  The natural-image input is a
120 x 160 crop (rows 100.., cols 200..) of the reference's data/demo/000456.jpg decoded by cv2.imread, the call the
reference makes at lib/model/test.py:191."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import golden_cases as G  # noqa: E402


def main():
    import cv2
    src = "/root/reference/data/demo/000456.jpg"
    if os.path.exists(src):
        img = cv2.imread(src)
        np.savez_compressed(G.NATURAL, img=np.ascontiguousarray(img[100:220, 200:360]))
    out = {"versions": G.versions(), "rng": "numpy.random.RandomState(3) per case (cfg.RNG_SEED)", "cases": {}}
    for cid, img, spec, fields in G.cases():
        want = G.evaluate(img, spec, fields)
        out["cases"][cid] = {"sha256": G.digest(want), "class": G.tolerance_class(spec)}
    out["inputs"] = {k: G.digest(v) for k, v in G.inputs().items()}
    with open(G.MANIFEST, "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print(len(out["cases"]), "cases ->", G.MANIFEST)


if __name__ == "__main__":
    main()
