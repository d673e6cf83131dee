"""GPU: b2d_run_host (host buffers, chunked H2D / kernels / D2H on three streams) must give the bytes of b2d_run
for every kind of plan scratch - per-image parameters (poisson), uploaded tables (bloom, bilateral), pre-pass fields
(brownian), palette tables (quant), wavelet planes - when the batch is MANY chunks long.  Runs in a subprocess with
B2D_HOST_CHUNK_MB=1 (3 images per chunk here) so that consecutive chunks really overlap."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

SCRIPT = r'''
import sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tests")
import numpy as np, torch
from conftest import synth_batch
from image_denoising_b200 import api
N, H, W = 20, 304, 320
batch = synth_batch(N, H, W, seed=4, kind="smooth")
dev = torch.from_numpy(batch).cuda()
pin = torch.from_numpy(batch).pin_memory()
out = torch.empty_like(pin)
specs = ["poisson_median", "poisson_wavelet", "poisson_gaus_blur", "bloom", "gaussian_bilateral_var1.0",
         "brownian_mean_var0.09", "brownian_var0.9", "quant_median_var7", "quant_wavelet_var3", "gaussian_wavelet_var1.0",
         "sap_mean_var0.4", "gamma_median_var0.1", "periodic_bilateral_var100", "original_wavelet"]
mixed = [specs[i % len(specs)] for i in range(N)]
bad = []
for sp in specs + [mixed]:
    plan = api.Plan(N, H, W, sp)
    for rep in range(2):                       # second call reuses the plan's internal buffers / events
        want = plan.run(dev, seed=11, first_index=100).cpu().numpy()
        got = plan.run_host(pin, out, seed=11, first_index=100).numpy()
        if not np.array_equal(got, want):
            bad.append((sp if isinstance(sp, str) else "mixed", rep, int((got != want).any(axis=(1, 2, 3)).sum())))
print("BAD", bad)
sys.exit(1 if bad else 0)
'''


def test_run_host_many_chunks_equals_run(cuda):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, B2D_DEBUG_KNOBS="1", B2D_HOST_CHUNK_MB="1")
    res = subprocess.run([sys.executable, "-c", SCRIPT, root], env=env, timeout=900, stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True)
    assert res.returncode == 0, res.stdout[-3000:]
