"""Times mercury_score_head (classifier head folded into the scoring kernel) against its HBM roofline and against
stock ATen (F.linear + F.cross_entropy) on the same GPU.  python tools/bench_head.py [N] [D] [C] [f32|bf16]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from mercury_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 512
c = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dt = torch.bfloat16 if (len(sys.argv) > 4 and sys.argv[4] == "bf16") else torch.float32
g = torch.Generator(device=dev).manual_seed(0)
npools = 2 if n * d * dt.itemsize > (200 << 20) else 4          # rotate > 126 MB L2
pools = [torch.relu(torch.randn(n, d, generator=g, device=dev)).to(dt) for _ in range(npools)]
w = torch.randn(c, d, generator=g, device=dev) * 0.05
b = torch.randn(c, generator=g, device=dev) * 0.1
y = torch.randint(0, c, (n,), generator=g, device=dev)
out, st = torch.empty(n, device=dev), ops.new_status(dev)


def timed(fn, reps=20):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        fn(i)
    z.record()
    torch.cuda.synchronize()
    return a.elapsed_time(z) / reps


ms = timed(lambda i: ops.score_head(pools[i % npools], w, b, y, status=st, out=out))
ms_lib = timed(lambda i: F.cross_entropy(F.linear(pools[i % npools].float(), w, b), y, reduction="none"))
peak = 6555.8
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
bytes_ = n * (d * dt.itemsize + 8 + 4)
ops.check_status(st)
print(json.dumps({"kernel": "mercury_score_head", "n": n, "d": d, "c": c, "dtype": str(dt), "ms": ms, "candidates_per_s": n / (ms * 1e-3),
                  "algorithmic_bytes": bytes_, "achieved_gbs": bytes_ / (ms * 1e-3) / 1e9, "peak_gbs": peak,
                  "frac": bytes_ / (ms * 1e-3) / 1e9 / peak, "stock_aten_ms": ms_lib, "l2": f"{npools} feature pools rotated"}))
