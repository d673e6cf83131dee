"""Debug: GPU wavelet statistics vs the oracle's census for every golden wavelet case."""
import json, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, ROOT + "/tests")
import numpy as np, torch
import golden_cases as G
from oracle import closures as C, wavelet as W
from image_denoising_b200 import api
sys.path.insert(0, ROOT + "/tests")
from test_gpu_golden import _fields_to_dev
cuda = torch.device("cuda", 0)
rows = []
for cid, img, spec, fields in G.cases():
    if "wavelet" not in spec or "brownian" in spec:
        continue
    x = img if spec.startswith("original") else C.noise_stage(img, spec, fields)
    h, w, _ = img.shape
    plan = api.get_plan(1, h, w, spec, 0)
    got = plan.run(torch.from_numpy(img[None]).to(cuda), fields=_fields_to_dev(fields, cuda)).cpu().numpy()[0]
    st = plan.wavelet_stats(0)
    xf = x.astype(np.float64) / 255 if x.dtype == np.uint8 else x
    cen = W.sigma_census(xf, "bior1.5")
    ost = W.denoise_wavelet(xf, "bior1.5", None, True)[1]
    for c in range(3):
        s = float(st[c, 0]) / float(st[c, 2] - st[c, 1])
        rows.append(dict(cid=cid, c=c, gpu=s, gpu_n=int(st[c, 3]), smin=cen[c][0], smax=cen[c][1], n_sure=cen[c][2], n_amb=cen[c][3],
                         oracle=float(ost[c]["sigma"]), rel=(s - float(ost[c]["sigma"])) / float(ost[c]["sigma"])))
json.dump(rows, open(ROOT + "/gpurun_out/wv_census.json", "w"))
bad = [r for r in rows if not (r["smin"] * (1 - 5e-4) <= r["gpu"] <= r["smax"] * (1 + 5e-4))]
print(len(rows), "rows", len(bad), "outside")
for r in bad:
    print(r)
