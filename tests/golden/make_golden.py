"""Generates the committed golden fixtures.  Run HERE (the container that has
/root/reference); the fixtures travel to the GPU box, this script's inputs do not.

  python tests/golden/make_golden.py

bloom_seed7_320x480.npz : output of the reference's own
    tools/Automold.py::add_sun_flare(img, flare_center=(100,100), angle=-pi/4)
    (the call at lib/roi_data_layer/minibatch.py:1487) with random.seed(7) on
    conftest.synth_image(320, 480, 7, 'smooth').
"""
import math
import os
import random
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, "/root/reference/tools")

from conftest import synth_image  # noqa: E402


def main():
    import Automold as am
    img = synth_image(320, 480, 7, "smooth")
    random.seed(7)
    out = am.add_sun_flare(img, flare_center=(100, 100), angle=-math.pi / 4)
    np.savez_compressed(os.path.join(HERE, "bloom_seed7_320x480.npz"), out=out)
    print("bloom fixture", out.shape, out.mean())


if __name__ == "__main__":
    main()
