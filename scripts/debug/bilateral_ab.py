import os, sys, subprocess, json
root = sys.argv[1]
code = r'''
import sys, numpy as np, torch
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[1] + "/tests")
from bench import smooth_batch_torch
from image_denoising_b200 import api
dev = torch.device("cuda", 0)
out = {}
for (n, h, w) in ((2, 61, 99), (2, 128, 200), (1, 375, 500)):
    b = smooth_batch_torch(n, h, w, dev, 5)
    out["%dx%d" % (h, w)] = api.apply(b, "speckle_bilateral_var1.0", seed=4).cpu().numpy()
b = smooth_batch_torch(128, 1080, 1920, dev, 1003); o = torch.empty_like(b)
plan = api.Plan(128, 1080, 1920, "original_bilateral", 0)
for _ in range(2): plan.run(b, o, 3, 0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): plan.run(b, o, 3, 0)
e1.record(); torch.cuda.synchronize()
print("ms per 128 x 1080p:", e0.elapsed_time(e1) / 5)
np.savez(sys.argv[2], **out)
'''
for tag, env in (("new", {}), ("old", {"B2D_BILATERAL": "1"})):
    subprocess.run([sys.executable, "-c", code, root, "/tmp/bil_%s.npz" % tag], check=True, env=dict(os.environ, B2D_DEBUG_KNOBS="1", **env))
import numpy as np
a, b = np.load("/tmp/bil_new.npz"), np.load("/tmp/bil_old.npz")
print("bit-identical:", all(np.array_equal(a[k], b[k]) for k in a.files))
