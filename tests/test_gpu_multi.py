"""Two real GPUs, one process each (NCCL): ShardedPool(mode="totals") with the exchange done by our kernels
over peer memory must give what the NCCL exchange gives, and both must match the two-level oracle.
Skipped on single-GPU boxes (the world-1 version of the peer path is in test_gpu_surface.py, the host logic
under gloo in test_sharded_gloo.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import mercury_oracle as orc

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _pool(n, c, b):
    g = torch.Generator().manual_seed(5)
    logits = torch.randn(n, c, generator=g) * 2
    labels = torch.randint(0, c, (n,), generator=g)
    data = torch.rand(n, 7, generator=g)
    u = torch.rand(b, dtype=torch.float64, generator=g)
    return logits, labels, data, u


def _worker(rank, world, port, n, c, b, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from mercury_b200 import ops
    from mercury_b200.sharded import ShardedPool, shard_bounds

    logits, labels, data, u = _pool(n, c, b)
    lo, nl = shard_bounds(n, world, rank)
    sl = slice(lo, lo + nl)
    pools = {"nccl": ShardedPool(nl, dev, mode="totals", exchange="nccl"),
             "peer": ShardedPool(nl, dev, mode="totals", exchange="peer"),                  # one C call per step
             "peer_staged": ShardedPool(nl, dev, mode="totals", exchange="peer", fused=False)}  # one call per stage
    pools["peer_one_kernel"] = pools["peer"]   # the same pool, the whole sharded step as ONE kernel (test hook: off by default)
    outs = {}
    for x, sp in pools.items():
        ops.set_fused(1 if x == "peer_one_kernel" else -1)
        steps = []
        for step in range(3):
            r = sp.update_samples((logits[sl] + step).to(dev), labels[sl].to(dev), u.to(dev), data=data[sl].to(dev), assemble=True)
            if r["xstatus"] is not None:
                ops.check_status(r["xstatus"])
            ops.check_status(r["status"])
            steps.append({k: r[k].cpu() for k in ("idx", "nprob", "seed", "data", "labels", "stats")})
            steps[-1]["scores"] = sp.local_scores.cpu()      # this rank's shard of the scores, as K1 wrote them
        outs[x] = steps
    ret[rank] = outs
    dist.barrier()
    torch.cuda.synchronize()
    pools["peer"].peer.close()
    pools["peer_staged"].peer.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_exchange_two_gpus_matches_nccl_and_oracle():
    n, c, b, world = 200000, 10, 128, 2
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), n, c, b, ret), nprocs=world, join=True)
    logits, labels, data, u = _pool(n, c, b)
    for step in range(3):
        for k in ("idx", "nprob", "seed", "data", "labels"):
            assert torch.equal(ret[0]["peer"][step][k], ret[0]["nccl"][step][k]), (step, k)   # peer == NCCL exchange
            assert torch.equal(ret[0]["peer_staged"][step][k], ret[0]["nccl"][step][k]), (step, k)
            assert torch.equal(ret[0]["peer"][step][k], ret[1]["peer"][step][k]), (step, k)   # assembled on both ranks
        st = ret[0]["peer"][step]["stats"]
        assert torch.equal(st[:4], ret[1]["peer"][step]["stats"][:4]) and torch.equal(st[:4], ret[0]["nccl"][step]["stats"][:4])
        ref_losses = orc.score(logits + step, labels)
        for x in ("peer", "peer_one_kernel"):
            dev_scores = torch.cat([ret[0][x][step]["scores"], ret[1][x][step]["scores"]])
            np.testing.assert_allclose(dev_scores.numpy(), ref_losses.numpy(), rtol=1e-5, atol=1e-6)
        # (the draws are checked on the device's own scores: one ulp in a score can move a draw that sits on a CDF step)
        losses = torch.cat([ret[0]["peer"][step]["scores"], ret[1]["peer"][step]["scores"]])
        p = ((losses.numpy() + np.float32(st[2])).astype(np.float32) / np.float32(st[3])).astype(np.float32)
        want = orc.sharded_draw_two_level([p[:n // 2], p[n // 2:]], u.numpy())
        assert np.array_equal(ret[0]["peer"][step]["idx"].numpy(), want)
        assert torch.equal(ret[0]["peer"][step]["data"], data[want])
        # the one-kernel sharded step sums its pool statistics per CTA instead of per tile (equal to 1e-6): checked against
        # the oracle on ITS OWN probabilities
        s1 = ret[0]["peer_one_kernel"][step]["stats"]
        np.testing.assert_allclose(s1[:4].numpy(), st[:4].numpy(), rtol=1e-6)
        losses = torch.cat([ret[0]["peer_one_kernel"][step]["scores"], ret[1]["peer_one_kernel"][step]["scores"]])
        p1 = ((losses.numpy() + np.float32(s1[2])).astype(np.float32) / np.float32(s1[3])).astype(np.float32)
        want1 = orc.sharded_draw_two_level([p1[:n // 2], p1[n // 2:]], u.numpy())
        assert np.array_equal(ret[0]["peer_one_kernel"][step]["idx"].numpy(), want1)
        assert torch.equal(ret[0]["peer_one_kernel"][step]["data"], data[want1])
        assert torch.equal(ret[0]["peer_one_kernel"][step]["idx"], ret[1]["peer_one_kernel"][step]["idx"])
