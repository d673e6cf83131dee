import sys, numpy as np, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from conftest import synth_batch
from image_denoising_b200 import api
from oracle import closures as C, wavelet as W
dev = torch.device('cuda', 0)
np.set_printoptions(linewidth=250)
for shape in [(9, 12), (9, 16), (10, 12), (11, 16)]:
    b = synth_batch(1, shape[0], shape[1], seed=31, kind="smooth")
    got = api.apply(torch.from_numpy(b).to(dev), "original_wavelet", wavelet="db1", levels=1).cpu().numpy()[0]
    want, st = W.denoise_wavelet(b[0], "db1", 1, return_stats=True)
    want8 = (255 * want).astype(np.uint8)
    d = got.astype(int) - want8.astype(int)
    print(shape, "max", np.abs(d).max()); print(d[..., 0]); print([ (s['sigma'], s['min'], s['max']) for s in st])
