"""A few eager whole-step calls on a cfg2-shaped pool, for ncu (see profiles/README.md):

    ncu --set full --import-source on --clock-control none -k regex:'summarize|spine|draw' -c 6 \
        -o gpurun_out/tail python tools/prof_step.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mercury_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
n, c, b = int(os.environ.get("N", "1000000")), 10, 128
g = torch.Generator(device=dev).manual_seed(100)
pools = []
for i in range(3):
    pools.append(((torch.randn(n, c, generator=g, device=dev) * 2), torch.randint(0, c, (n,), generator=g, device=dev)))
payload = torch.rand(n, 3, 32, 32, device=dev, generator=g)
index = torch.arange(n, device=dev)
u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(3)).to(dev)
ema = torch.zeros(1, device=dev)
out = {}
for i in range(int(os.environ.get("STEPS", "4"))):
    r = ops.update_samples(*pools[i % 3], u, ema, i == 0, data=payload, index=index, out=out)
    out.update({k: v for k, v in r.items() if k != "status"})
torch.cuda.synchronize()
ops.check_status(r.status)
print("ok", r.idx[:4].tolist())
