import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from conftest import synth_batch
from image_denoising_b200 import api
dev = torch.device('cuda', 0)
def bench(spec, n, h, w, reps=5, **kw):
    x = torch.from_numpy(synth_batch(min(n, 8), h, w, 2, "smooth")).to(dev)
    x = x.repeat((n + x.shape[0] - 1) // x.shape[0], 1, 1, 1)[:n].contiguous()
    p = api.Plan(n, h, w, spec, 0, **kw); out = torch.empty_like(x)
    for _ in range(2): p.run(x, out)
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): p.run(x, out)
    e1.record(); torch.cuda.synchronize(); ms = e0.elapsed_time(e1) / reps
    print("%-30s %s n=%d %dx%d  %.3f ms  %.1f Gpix/s" % (spec, kw, n, h, w, ms, n * h * w / 1e6 / ms))
for a in sys.argv[1:]:
    spec, n, h, w, *rest = a.split(":")
    kw = {k: int(v) for k, v in (r.split("=") for r in rest)}
    bench(spec, int(n), int(h), int(w), **kw)
