"""World-size-2 gloo test (CPU) of the sharded-pool host logic: shard arithmetic, the in-place
all-gather of scores, replicated global draw, ownership and batch assembly.  The compute steps
are injected from the oracle (tests may do that; the product backend is CUDA-only); the GPU
version of the same check lives in test_gpu_surface.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import mercury_oracle as orc


class OracleBackend:
    """CPU stand-in with the CudaBackend interface of mercury_b200.sharded (tests only)."""

    def score(self, logits, labels, out):
        out.copy_(orc.score(logits, labels))

    def draw(self, scores_all, uniforms, ema_state, ema_first, ema_alpha, smooth_alpha, out):
        ema = orc.EMAverage(ema_alpha)
        ema.first_update = bool(ema_first)
        ema.value = ema_state[0].clone()
        mean, w, p = orc.pool_stats(scores_all, ema, smooth_alpha)
        ema_state[0] = ema.value
        idx = torch.from_numpy(orc.draw_with_replacement(p.numpy(), uniforms.numpy()))
        nprob = p[idx] * p.numel()
        seed = (1.0 / idx.numel()) / nprob
        stats = torch.tensor([float(mean), float(ema.value), smooth_alpha * float(ema.value), float(w.sum()), 0, 0, 0, 0])
        return idx, nprob, seed, stats, torch.zeros(1, dtype=torch.int32)

    def gather_owned(self, idx, lo, n, data, labels, index, out):
        local = idx - lo
        mine = (local >= 0) & (local < n)
        safe = torch.where(mine, local, torch.zeros_like(local))

        def take(t):
            if t is None:
                return None
            rows = t[safe].clone()
            rows[~mine] = 0
            return rows

        return take(data), take(labels), take(index), mine.to(torch.int32)


    # ---- mode "totals"
    @staticmethod
    def _p(scores, stats):
        return ((scores.numpy() + np.float32(stats[2])).astype(np.float32) / np.float32(stats[3])).astype(np.float32)

    def sum_scores(self, scores, out):
        out.copy_(scores.double().sum().reshape(1))

    def local_chain(self, scores, sums_all, n_global, ema_state, ema_first, ema_alpha, smooth_alpha, out):
        tot = float(sums_all.sum())
        mean = np.float32(tot / n_global)
        ema = mean if ema_first else np.float32(np.float32(ema_alpha) * np.float32(ema_state[0]) + np.float32(1.0 - ema_alpha) * mean)
        ema_state[0] = float(ema)
        shift = np.float32(np.float32(smooth_alpha) * ema)
        total = np.float32(tot + n_global * float(shift))
        stats = torch.tensor([mean, ema, shift, total, 0, 0, 0, 0], dtype=torch.float32)
        stats[4] = float(orc.serial_chain_f32(self._p(scores, stats))[-1])
        return stats, torch.zeros(1, dtype=torch.int32)

    def shard_select(self, totals_all, rank, uniforms, out):
        c = np.cumsum(totals_all.numpy(), dtype=np.float32)
        cdf = (c / c[-1]).astype(np.float32).astype(np.float64)
        for j, uj in enumerate(uniforms.numpy()):
            g = 0
            while g < len(cdf) - 1 and cdf[g] < uj:
                g += 1
            lo, hi = (cdf[g - 1] if g else 0.0), cdf[g]
            out[j] = min(max((uj - lo) / (hi - lo), 0.0), 0.99999999999999989) if g == rank else -1.0
        return out

    def local_draw(self, scores, local_u, stats, sums_all, n_global, lo, ema_state, data, labels, index, out, status):
        p = self._p(scores, stats)
        mine = local_u >= 0
        idx_l = torch.zeros(local_u.numel(), dtype=torch.int64)
        if mine.any():
            idx_l[mine] = torch.from_numpy(orc.lower_bound(orc.serial_cdf_f32(p), local_u[mine].numpy()))
        nprob = torch.from_numpy(p)[idx_l] * n_global
        nprob[~mine] = 0
        seed = torch.where(mine, (1.0 / local_u.numel()) / nprob.clamp_min(1e-30), torch.zeros_like(nprob))

        def take(t):
            if t is None:
                return None
            rows = t[idx_l].clone()
            rows[~mine] = 0
            return rows

        idx = torch.where(mine, idx_l + lo, torch.zeros_like(idx_l))
        return dict(idx=idx, nprob=nprob, seed=seed, data=take(data), labels=take(labels), index=take(index))


def _pool(n, c, seed=0):
    g = torch.Generator().manual_seed(seed)
    logits = torch.randn(n, c, generator=g) * 2
    labels = torch.randint(0, c, (n,), generator=g)
    data = torch.rand(n, 3, 4, 4, generator=g)
    index = torch.randperm(n, generator=g)
    u = torch.rand(48, dtype=torch.float64, generator=g)
    return logits, labels, data, index, u


def _worker(rank, world, port, n, c, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mercury_b200.sharded import ShardedPool, shard_bounds

        logits, labels, data, index, u = _pool(n, c)
        lo, nl = shard_bounds(n, world, rank)
        sp = ShardedPool(nl, "cpu", backend=OracleBackend())
        outs = []
        for step in range(2):  # EMA carried across steps identically on every rank
            r = sp.update_samples(logits[lo:lo + nl] + step, labels[lo:lo + nl], u, data=data[lo:lo + nl],
                                  index=index[lo:lo + nl], assemble=True)
            outs.append({k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in r.items()})
        ret[rank] = outs
    finally:
        dist.destroy_process_group()


def _worker_totals(rank, world, port, n, c, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mercury_b200.sharded import ShardedPool, shard_bounds

        logits, labels, data, index, u = _pool(n, c)
        lo, nl = shard_bounds(n, world, rank)
        sp = ShardedPool(nl, "cpu", backend=OracleBackend(), mode="totals")
        outs = []
        for step in range(2):
            r = sp.update_samples(logits[lo:lo + nl] + step, labels[lo:lo + nl], u, data=data[lo:lo + nl], index=index[lo:lo + nl],
                                  assemble=True)
            outs.append({k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in r.items()})
        ret[rank] = outs
    finally:
        dist.destroy_process_group()


def test_sharded_pool_totals_mode_world2_matches_two_level_oracle():
    n, c, world = 4096, 10, 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker_totals, args=(world, _free_port(), n, c, ret), nprocs=world, join=True)
    logits, labels, data, index, u = _pool(n, c)
    for step in range(2):
        a, b = ret[0][step], ret[1][step]
        assert torch.equal(a["stats"][:4], b["stats"][:4])                       # whole-pool statistics agree
        assert torch.equal(a["idx"], b["idx"]) and torch.equal(a["data"], b["data"])   # assembled batch on both ranks
        assert int(a["owned"].sum() + b["owned"].sum()) == len(u)
        losses = orc.score(logits + step, labels)
        st = a["stats"]
        p = ((losses.numpy() + np.float32(st[2])).astype(np.float32) / np.float32(st[3])).astype(np.float32)
        want = orc.sharded_draw_two_level([p[:n // 2], p[n // 2:]], u.numpy())
        assert np.array_equal(a["idx"].numpy(), want)
        assert torch.equal(a["data"], data[want]) and torch.equal(a["labels"], labels[want]) and torch.equal(a["index"], index[want])
        np.testing.assert_allclose(float(st[0]), float(losses.mean()), rtol=1e-6)


class _Loader:
    def __init__(self, batches, batch_size):
        self.batches, self.batch_size = batches, batch_size

    def __iter__(self):
        return iter(self.batches)


def _worker_trainer_global(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mercury_b200.collab import Trainer
        from mercury_b200.sampler import EMAverage

        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Flatten(), torch.nn.Linear(48, 10))     # same weights on every worker
        g = torch.Generator().manual_seed(50 + rank)                                # each worker its own presample shard
        batches = [(torch.arange(i * 16, (i + 1) * 16) + 1000 * rank, torch.rand(16, 3, 4, 4, generator=g),
                    torch.randint(0, 10, (16,), generator=g)) for i in range(10)]
        tr = Trainer(net, None, _Loader([], 8), _Loader(batches, 16), None, "cpu", global_draw=True,
                     sharded_kwargs={"backend": OracleBackend()})
        ema = EMAverage(device="cpu")
        torch.manual_seed(77)        # (rank 0's stream supplies the uniforms)
        outs = []
        for step in range(2):
            probs, data, label, index, avg = tr.update_samples(ema)
            outs.append(dict(probs=probs.clone(), data=data.clone(), label=label.clone(), index=index.clone(), avg=float(avg),
                             idx=tr.last.idx.clone(), seed=tr.last.seed.clone(), scores=tr.last.scores.clone(), stats=tr.last.stats.clone(),
                             pool_data=tr._pool.tensors()[2].clone(), pool_index=tr._pool.tensors()[3].clone()))
        ret[rank] = outs
    finally:
        dist.destroy_process_group()


def test_trainer_global_draw_world2_draws_from_the_sharded_pool():
    """Trainer(global_draw=True): the two workers' presample pools are one sharded pool; one two-level draw of world * B
    candidates per step, worker r trains on rows [r*B, (r+1)*B) -- checked against the two-level oracle over both shards."""
    world, B, nl = 2, 8, 160
    ret = mp.Manager().dict()
    mp.spawn(_worker_trainer_global, args=(world, _free_port(), ret), nprocs=world, join=True)
    torch.manual_seed(77)
    for step in range(2):
        u = torch.rand(world * B, dtype=torch.float64).numpy()
        st = ret[0][step]["stats"].numpy()
        assert np.array_equal(st[:4], ret[1][step]["stats"].numpy()[:4])
        shards = [((ret[r][step]["scores"].numpy() + np.float32(st[2])).astype(np.float32) / np.float32(st[3])).astype(np.float32)
                  for r in range(world)]
        want = orc.sharded_draw_two_level(shards, u)
        p_all = np.concatenate(shards)
        for r in range(world):
            o = ret[r][step]
            mine = want[r * B:(r + 1) * B]
            assert np.array_equal(o["idx"].numpy(), mine)
            np.testing.assert_array_equal(o["probs"].numpy(), p_all[mine] * np.float32(world * nl))
            np.testing.assert_allclose(o["seed"].numpy(), (1.0 / B) / o["probs"].numpy(), rtol=1e-6)
            for j, gidx in enumerate(mine):          # rows come from the owning worker's pool
                owner, local = divmod(int(gidx), nl)
                assert torch.equal(o["data"][j], ret[owner][step]["pool_data"][local])
                assert int(o["index"][j]) == int(ret[owner][step]["pool_index"][local])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_sharded_pool_world2_matches_single_process():
    n, c, world = 4096, 10, 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), n, c, ret), nprocs=world, join=True)
    logits, labels, data, index, u = _pool(n, c)
    ema = orc.EMAverage()
    for step in range(2):
        ref = orc.update_samples_flat(logits + step, labels, len(u), ema, u=u.numpy(), data=data, index=index)
        for rank in range(world):
            got = ret[rank][step]
            assert np.array_equal(got["idx"].numpy(), ref["idx"].numpy())          # same global draw on every rank
            np.testing.assert_allclose(got["nprob"].numpy(), ref["nprob"].numpy(), rtol=1e-6)
            assert torch.equal(got["scores"], ref["scores"])                      # all-gathered score vector
            assert torch.equal(got["data"], ref["data"])                          # assembled batch
            assert torch.equal(got["labels"], ref["labels"]) and torch.equal(got["index"], ref["index"])
            lo = rank * (n // world)
            mine = (ref["idx"] >= lo) & (ref["idx"] < lo + n // world)
            assert torch.equal(got["owned"].bool(), mine)
        assert int(ret[0][step]["owned"].sum() + ret[1][step]["owned"].sum()) == len(u)  # every draw has one owner


def test_shard_bounds():
    from mercury_b200.sharded import shard_bounds

    assert shard_bounds(50000, 8, 3) == (18750, 6250)   # BASELINE cfg4: 8 shards of 6250
    assert [shard_bounds(16, 4, r)[0] for r in range(4)] == [0, 4, 8, 12]
    with pytest.raises(ValueError):
        shard_bounds(10, 4, 0)
