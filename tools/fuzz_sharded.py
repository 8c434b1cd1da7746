"""Randomised check of the sharded step on real GPUs (launch with torchrun, 2+ ranks): random shard sizes, class counts and draw
counts; exchange over peer memory (kernel chain and one-kernel step) and over NCCL; every variant against the two-level oracle
on the job's own scores, with the assembled batch compared across ranks.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29520 tools/fuzz_sharded.py [iterations] [seed]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from mercury_b200 import ops  # noqa: E402
from mercury_b200.sharded import ShardedPool  # noqa: E402
from oracle import mercury_oracle as orc  # noqa: E402

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)     # (the same stream on every rank)
bad = 0
for it in range(iters):
    n = int(rng.choice([rng.integers(1, 3000), rng.integers(3000, 100_000), rng.integers(100_000, 700_000)]))
    c = int(rng.choice([3, 10, 10, 32]))
    b = int(rng.choice([1, 16, 128, 500]))
    seed = int(rng.integers(1 << 30))
    g = torch.Generator(device=dev).manual_seed(seed + rank)
    x = torch.randn(n, c, generator=g, device=dev) * 2
    y = torch.randint(0, c, (n,), generator=g, device=dev)
    data = torch.rand(n, 6, generator=g, device=dev)
    index = torch.arange(n, device=dev) + rank * n
    u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(seed)).to(dev)
    ref_idx = None
    for variant in ("nccl", "peer", "peer_one_kernel"):
        ops.set_fused(1 if variant == "peer_one_kernel" else -1)
        sp = ShardedPool(n, dev, mode="totals", exchange="nccl" if variant == "nccl" else "peer")
        for step in range(2):
            r = sp.update_samples(x + step, y, u, data=data, index=index, assemble=True)
            torch.cuda.synchronize()
            sp.check()
            scores = [torch.empty(n, device=dev) for _ in range(world)]
            dist.all_gather(scores, r["scores"].contiguous())
            hi, lo = r["idx"].clone(), r["idx"].clone()
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            mine = (r["idx"] >= sp.lo) & (r["idx"] < sp.lo + n)
            ok = bool(torch.equal(hi, lo)) and bool(torch.equal(r["data"][mine], data[r["idx"][mine] - sp.lo])) \
                and bool(torch.equal(r["index"][mine], index[r["idx"][mine] - sp.lo]))
            if rank == 0:
                st = r["stats"].cpu().numpy()
                sh = [((s.cpu().numpy() + np.float32(st[2])).astype(np.float32) / np.float32(st[3])).astype(np.float32) for s in scores]
                ok = ok and np.array_equal(r["idx"].cpu().numpy(), orc.sharded_draw_two_level(sh, u.cpu().numpy()))
            t = torch.tensor([int(ok)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            if int(t) == 0:
                bad += 1
                if rank == 0:
                    print(f"MISMATCH it={it} variant={variant} step={step} n={n} c={c} b={b}", flush=True)
        if sp.peer is not None:
            dist.barrier()
            torch.cuda.synchronize()
            sp.peer.close()
    if rank == 0 and it % 10 == 9:
        print(f"{it + 1} iterations, {bad} bad", flush=True)
ops.set_fused(-1)
if rank == 0:
    print("done:", iters, "iterations,", bad, "bad", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(1 if bad else 0)
