import sys, numpy as np, torch, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from conftest import synth_batch
from image_denoising_b200 import api
from oracle import closures as C, denoise as D
dev = torch.device('cuda', 0)
print(torch.cuda.get_device_name(0))
b = synth_batch(2, 600, 1000, 1)
for den, f in [("gaus_blur", D.gaus_blur_cv2), ("mean", D.mean_cv2), ("median", lambda u: D.median_cv2(u,3))]:
    got = api.apply(torch.from_numpy(b).to(dev), "original_"+den).cpu().numpy()
    want = np.stack([f(x) for x in b]); d = np.abs(got.astype(int)-want)
    print(den, "maxdiff", d.max(), "nbad", (d>0).sum())
x = torch.from_numpy(synth_batch(64, 600, 1000, 2)).to(dev)
for spec in ["original_gaus_blur", "sap_mean_var0.4", "gaussian_gaus_blur_var1.5", "gaussian_median_var1.0", "gaussian_var1.0", "sap_var0.4","uniform_var0.6","rayleigh_var0.2","gamma_var0.1","poisson","speckle_var1.0","periodic_var100","brownian_var0.09","shader"]:
    try:
        p = api.get_plan(64, 600, 1000, spec, 0)
        out = torch.empty_like(x)
        for _ in range(3): p.run(x, out)
        torch.cuda.synchronize(); e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10): p.run(x, out)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)/10
        print("%-28s %.3f ms  %.1f Gpix/s  %.0f GB/s  mean=%.2f" % (spec, ms, 64*0.6/ms, 64*0.6*6/ms, out.float().mean().item()))
    except Exception as ex:
        print(spec, "FAILED", ex)
