import sys, numpy as np, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from conftest import synth_batch
from image_denoising_b200 import api
from oracle import closures as C
dev = torch.device('cuda', 0)
for shape in [(64, 96), (375, 500), (376, 500), (375, 504), (37, 53), (128, 192)]:
    for wl, lv in [("bior1.5", 1), ("bior1.5", 2), ("db1", 1), ("db1", 2), ("db1", 3)]:
        b = synth_batch(1, shape[0], shape[1], seed=31, kind="smooth")
        got = api.apply(torch.from_numpy(b).to(dev), "original_wavelet", wavelet=wl, levels=lv).cpu().numpy()[0]
        want = C.apply_denoise(b[0], "wavelet", wavelet=wl, wavelet_levels=lv)
        d = np.abs(got.astype(int) - want.astype(int))
        ys, xs = np.nonzero(d.max(-1) > 1)
        print(shape, wl, lv, "max", d.max(), "n>1", (d > 1).sum(), "frac>0 %.4f" % (d > 0).mean(),
              "rows", (ys.min(), ys.max()) if len(ys) else None, "cols", (xs.min(), xs.max()) if len(xs) else None)
