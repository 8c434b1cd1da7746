"""K1 alone at the 1.28M x 1000 bf16 shape (or N C from the command line): time, GB/s, and the error against an fp64 log-sum-exp on a slice.
MERCURY_K1_POLY=0..3 python tools/bench_k1.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mercury_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
n, c = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1_280_000, 1000)
g = torch.Generator(device=dev).manual_seed(1)
x = torch.empty(n, c, dtype=torch.bfloat16, device=dev)
for i in range(0, n, 1 << 17):
    x[i:i + (1 << 17)] = (torch.randn(min(1 << 17, n - i), c, generator=g, device=dev) * 2).to(torch.bfloat16)
y = torch.randint(0, c, (n,), generator=g, device=dev)
out = torch.empty(n, device=dev)
for _ in range(5):
    ops.score(x, y, out=out)
torch.cuda.synchronize()
a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
reps = 30
for _ in range(reps):
    ops.score(x, y, out=out)
z.record()
torch.cuda.synchronize()
ms = a.elapsed_time(z) / reps
m = min(n, 65536 if c <= 1000 else 2048)
xs = x[:m].double()
ref = torch.logsumexp(xs, 1) - xs.gather(1, y[:m, None])[:, 0]
err = ((out[:m].double() - ref).abs() / ref.abs().clamp_min(1e-3)).max()
print(json.dumps({"poly": os.environ.get("MERCURY_K1_POLY", "default"), "ms": round(ms, 4), "GBs": round(n * (c * 2 + 12) / ms / 1e6, 1),
                  "max_rel_err_vs_fp64": float(err)}))
