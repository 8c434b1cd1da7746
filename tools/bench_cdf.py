"""The materialised CDF path: mercury_cdf (exact serial-fp32 / fp64) + the inverse-CDF search, timed alone.
python tools/bench_cdf.py  -> JSON lines"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mercury_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)


def timed(fn, reps=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    z.record()
    torch.cuda.synchronize()
    return a.elapsed_time(z) / reps


for n in (50_000, 1_000_000, 1 << 24):
    g = torch.Generator(device=dev).manual_seed(1)
    w = torch.rand(n, generator=g, device=dev) + 0.01
    p = w / w.sum()
    u = torch.rand(256, dtype=torch.float64, generator=torch.Generator().manual_seed(2)).to(dev)
    for mode in ("exact", "f64"):
        ms = timed(lambda: ops.cdf(p, mode))
        print(json.dumps({"kernel": f"mercury_cdf ({mode})", "n": n, "ms": round(ms, 5), "bytes": 8 * n, "gbs": round(8 * n / ms / 1e6, 1)}))
    cdf, _ = ops.cdf(p, "exact")
    ms = timed(lambda: ops.sample_with_replacement(cdf, u))
    print(json.dumps({"kernel": "mercury_sample_with_replacement (256 draws, materialised CDF)", "n": n, "ms": round(ms, 5)}))
    ms = timed(lambda: torch.multinomial(p, 256, replacement=True), reps=20)
    print(json.dumps({"kernel": "stock torch.multinomial on the GPU (256 draws, its own RNG and arithmetic)", "n": n, "ms": round(ms, 5)}))
