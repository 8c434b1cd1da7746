"""Times the exponential-race draws (north_star kernel 2, "Gumbel-top-k"): mercury_sample_topk (k without replacement over
N keys) and mercury_groupwise_draw (Groupwise_Sampler's group-masked draws: blocks of single draws, and one draw of k).
    python tools/bench_topk.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mercury_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    z.record()
    torch.cuda.synchronize()
    return a.elapsed_time(z) / reps


out = []
for n in (50_000, 1_000_000):
    g = torch.Generator(device=dev).manual_seed(n)
    w = torch.rand(n, generator=g, device=dev) + 1e-3
    p = w / w.sum()
    u = torch.rand(n, dtype=torch.float64, generator=g, device=dev)
    st = ops.new_status(dev)
    for k in (1, 256, 2048):
        ms = timed(lambda: ops.sample_topk(p, u, k, st))
        out.append({"kernel": "mercury_sample_topk", "n": n, "k": k, "ms": ms, "keys_per_s": n / (ms * 1e-3),
                    "bytes": n * (4 + 8 + 4), "gbs": n * 16 / (ms * 1e-3) / 1e9})
    ms_t = timed(lambda: torch.topk(p / torch.empty_like(p).exponential_(), 256))
    out.append({"kernel": "stock ATen exponential_ + div + topk(256) (its own RNG)", "n": n, "ms": ms_t})
    tags = torch.randint(1, 5, (n,), generator=g, device=dev).to(torch.int32)
    count = int((tags == 2).sum())
    for nd in (1, 64):
        ud = torch.rand(nd * count, dtype=torch.float64, generator=g, device=dev)
        ms = timed(lambda: ops.groupwise_draw(w, tags, 2, count, ud, ndraws=nd, status=st))
        out.append({"kernel": "mercury_groupwise_draw (single draws)", "n": n, "group_members": count, "draws_per_call": nd, "ms": ms,
                    "draws_per_s": nd / (ms * 1e-3)})
    ud = torch.rand(count, dtype=torch.float64, generator=g, device=dev)
    ms = timed(lambda: ops.groupwise_draw(w, tags, 2, count, ud, k=256, status=st))
    out.append({"kernel": "mercury_groupwise_draw (one draw of 256 without replacement)", "n": n, "group_members": count, "ms": ms})
    ops.check_status(st)
for o in out:
    print(json.dumps(o))
