"""One GPU, world 1: the sharded step (exchange="peer", the rank stores into / polls its own box) against the plain step, each
as the chain and as the one-kernel step -- what the exchange code costs before any NVLink latency is added.
python tools/ab_peer.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mercury_b200 import ops  # noqa: E402
from mercury_b200.sharded import ShardedPool  # noqa: E402

dev = torch.device("cuda", 0)
n, c, b, NP = 1_000_000, 10, 128, 7
g = torch.Generator(device=dev).manual_seed(100)
pools = [((torch.randn(n, c, generator=g, device=dev) * 2), torch.randint(0, c, (n,), generator=g, device=dev)) for _ in range(NP)]
payload = torch.rand(n, 3, 32, 32, device=dev, generator=g)
index = torch.arange(n, device=dev)
u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(3)).to(dev)
ema = torch.zeros(1, device=dev)


def timed(step, label):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(3):
            step(i)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        for i in range(NP):
            step(i)
    for _ in range(5):
        gr.replay()
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(60):
        gr.replay()
    z.record()
    torch.cuda.synchronize()
    print(f"{label}: {a.elapsed_time(z) / (60 * NP) * 1e3:.2f} us", flush=True)


for fused in (-1, 1):
    ops.set_fused(fused)
    out, status = {}, ops.new_status(dev)

    def plain(i):
        r = ops.update_samples(*pools[i % NP], u, ema, False, data=payload, index=index, out=out, status=status)
        out.update({k: v for k, v in r.items() if k != "status" and v is not None})

    timed(plain, f"plain step, fused={fused}")
    sp = ShardedPool(n, dev, mode="totals", exchange="peer")
    sp.update_samples(*pools[0], u, data=payload, index=index)
    timed(lambda i: sp.update_samples(*pools[i % NP], u, data=payload, index=index), f"peer step (world 1), fused={fused}")
    if fused == 1:
        print("stamps:", ops.step_times(1) if hasattr(ops, "step_times") else None)
    sp.peer.close()
