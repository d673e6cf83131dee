import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def synth_image(h, w, seed=0, kind="uniform"):
    """Deterministic synthetic BGR uint8 image."""
    rng = np.random.RandomState(seed)
    if kind == "uniform":
        return rng.randint(0, 256, size=(h, w, 3)).astype(np.uint8)
    y, x = np.mgrid[0:h, 0:w]
    out = np.empty((h, w, 3), np.float64)
    for c in range(3):
        out[..., c] = (127.5 + 60 * np.cos(2 * np.pi * x / max(w, 1) * 3 + c)
                       + 60 * np.cos(2 * np.pi * y / max(h, 1) * 2 + 2 * c) + rng.normal(0, 8, (h, w)))
    return np.clip(out, 0, 255).astype(np.uint8)


def synth_batch(n, h, w, seed=0, kind="uniform"):
    return np.stack([synth_image(h, w, seed * 1000 + i, kind) for i in range(n)])


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from image_denoising_b200 import _lib
    _lib.lib()          # must load: the GPU tests never fall back to anything else
    return torch.device("cuda", 0)
