"""Times the whole step and the step without its draws (CUDA-graph replay over rotating cfg2 pools), to attribute a
change to the chain (score/summarize/spine) or to the draw kernel.  python tools/ab_step.py [N]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

MODE = os.environ.get("CDF", "exact")   # or f64

from mercury_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
n, c, b = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000, 10, 128
NP = 7 if n * 48 >= (40 << 20) else min(64, -(-(320 << 20) // (n * 48)))   # >= 300 MB of pools in rotation
g = torch.Generator(device=dev).manual_seed(100)
pools = [((torch.randn(n, c, generator=g, device=dev) * 2), torch.randint(0, c, (n,), generator=g, device=dev)) for _ in range(NP)]
payload = torch.rand(n, 3, 32, 32, device=dev, generator=g)
index = torch.arange(n, device=dev)
u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(3)).to(dev)
u0 = torch.empty(0, dtype=torch.float64, device=dev)
ema = torch.zeros(1, device=dev)


def timed(uni, label):
    out, status = {}, ops.new_status(dev)

    def step(i):
        r = ops.update_samples(*pools[i % NP], uni, ema, False, data=payload, index=index, out=out, status=status, cdf_mode=MODE)
        out.update({k: v for k, v in r.items() if k != "status" and v is not None})

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for i in range(3):
            step(i)
    torch.cuda.synchronize()
    graphs = []
    for i in range(NP):
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=side):
            step(i)
        graphs.append(gr)
    for gr in graphs:
        gr.replay()
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(420):
        graphs[i % NP].replay()
    z.record()
    torch.cuda.synchronize()
    print(f"{label}: {a.elapsed_time(z) / 420 * 1e3:.2f} us")


def timed_one_graph(uni, label):
    """All seven pools' steps in ONE graph: what is left of the per-launch gap between consecutive steps."""
    out, status = {}, ops.new_status(dev)

    def step(i):
        r = ops.update_samples(*pools[i % NP], uni, ema, False, data=payload, index=index, out=out, status=status, cdf_mode=MODE)
        out.update({k: v for k, v in r.items() if k != "status" and v is not None})

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
    gr.replay()
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    reps = max(4, 420 // NP)
    for i in range(reps):
        gr.replay()
    z.record()
    torch.cuda.synchronize()
    print(f"{label}: {a.elapsed_time(z) / (reps * NP) * 1e3:.2f} us")


timed(u, "step")
timed(u0, "step without draws")
timed_one_graph(u, f"step, {NP} steps per graph")
