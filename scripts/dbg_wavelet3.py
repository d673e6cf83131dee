import sys, numpy as np, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from conftest import synth_batch
from image_denoising_b200 import api
from oracle import closures as C
dev = torch.device('cuda', 0)
for shape in [(600, 1000), (300, 500), (160, 200)]:
    batch = synth_batch(1, shape[0], shape[1], seed=3, kind="smooth")
    for spec, kw in (("periodic_wavelet_var100", {}), ("original_wavelet", {}), ("periodic_var100", {}), ("original_wavelet", {"wavelet": "db1", "levels": 3})):
        got = api.apply(torch.from_numpy(batch).to(dev), spec, **kw).cpu().numpy()
        want = C.corrupt(batch[0], spec, wavelet=kw.get("wavelet", "bior1.5"), wavelet_levels=kw.get("levels") or None) if "original" not in spec else C.apply_denoise(batch[0], "wavelet", kw.get("wavelet", "bior1.5"), kw.get("levels") or None)
        d = np.abs(got[0].astype(int) - want.astype(int))
        print(shape, spec, kw, "max", d.max(), "frac>1", (d > 1).mean())
