"""The BASELINE configs at their real sizes and the callers either side of the path, on the GPU:
  * configs[2]: 1.28 M candidates x 1000-class bf16 logits, B = 256 -- the whole step against the oracle;
  * Trainer(score_from_features=True): presample batches scored from the pooled features (split_head + score_head) give the
    same step as scoring the logits;
  * gradient averaging over NCCL on two GPUs (skipped on one).
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from oracle import mercury_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def test_whole_step_at_imagenet_shape_full_size():
    """BASELINE configs[2] at full size.  Scores against the upcast oracle in slices (1e-5), statistics, and the 256 indices
    bit-exact against the serial fp32 chain over the device's own probabilities (= torch.multinomial's arithmetic)."""
    from mercury_b200 import ops

    n, c, b = 1_280_000, 1000, 256
    g = torch.Generator(device=DEV).manual_seed(7)
    x = torch.empty(n, c, dtype=torch.bfloat16, device=DEV)
    step = 1 << 17
    for i in range(0, n, step):
        x[i:i + step] = (torch.randn(min(step, n - i), c, generator=g, device=DEV) * 2).to(torch.bfloat16)
    y = torch.randint(0, c, (n,), generator=g, device=DEV)
    u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(3))
    u[0], u[-1] = 0.0, 1.0 - 2.0 ** -53
    index = torch.arange(n, device=DEV) * 3
    r = ops.update_samples(x, y, u.to(DEV), torch.zeros(1, device=DEV), True, index=index)
    ops.check_status(r.status)
    scores = r.scores.cpu()
    for lo in (0, 640_000, n - step):           # three slices of 128 Ki rows through the CPU oracle
        ref = orc.score(x[lo:lo + step].cpu(), y[lo:lo + step].cpu())
        np.testing.assert_allclose(scores[lo:lo + step].numpy(), ref.numpy(), rtol=1e-5, atol=1e-5)
    mean, w, p_ref = orc.pool_stats(scores, orc.EMAverage(), 0.5)
    st = r.stats.cpu().numpy()
    np.testing.assert_allclose(st[0], float(mean), rtol=1e-6)
    np.testing.assert_allclose(st[3], float(w.sum()), rtol=1e-6)
    p = ops.normalize(r.scores, r.stats).cpu().numpy()
    np.testing.assert_allclose(p, p_ref.numpy(), rtol=1e-5)
    want = orc.draw_with_replacement(p, u.numpy())
    assert np.array_equal(r.idx.cpu().numpy(), want)
    assert np.array_equal(r.index.cpu().numpy(), want * 3) and torch.equal(r.labels.cpu(), y.cpu()[want])
    np.testing.assert_array_equal(r.nprob.cpu().numpy(), p[want] * np.float32(n))
    del x
    torch.cuda.empty_cache()


class TinyResNet(nn.Module):
    """The attribute layout of the reference's ResNet (pytorch_model.py:67-97: conv1, bn1, layer1..4, linear; forward =
    relu(bn1(conv1)) -> layers -> avg_pool2d(4) -> flatten -> linear), small enough for a test."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 8, 3, padding=1, bias=False)
        self.bn1 = nn.BatchNorm2d(8)
        self.layer1 = nn.Sequential(nn.Conv2d(8, 8, 3, padding=1), nn.ReLU())
        self.layer2 = nn.Sequential(nn.Conv2d(8, 16, 3, stride=2, padding=1), nn.ReLU())
        self.layer3 = nn.Sequential(nn.Conv2d(16, 32, 3, stride=2, padding=1), nn.ReLU())
        self.layer4 = nn.Sequential(nn.Conv2d(32, 128, 3, stride=2, padding=1), nn.ReLU())
        self.linear = nn.Linear(128, num_classes)

    def forward(self, x):
        out = F.relu(self.bn1(self.conv1(x)))
        out = self.layer4(self.layer3(self.layer2(self.layer1(out))))
        out = F.avg_pool2d(out, 4)
        return self.linear(out.view(out.size(0), -1))


class _Loader:
    def __init__(self, batches, batch_size):
        self.batches, self.batch_size = batches, batch_size

    def __iter__(self):
        return iter(self.batches)


def test_trainer_scoring_from_features_on_the_gpu():
    """Trainer(score_from_features=True): split_head + mercury_score_head fill a scores-only pool; the step must equal the
    logits path (net forward + the fused step): same draws, N*p to 1e-5 (the head accumulates in another order)."""
    from mercury_b200.collab import Trainer
    from mercury_b200.sampler import EMAverage

    torch.manual_seed(3)
    net = TinyResNet(10).to(DEV).eval()
    gen = torch.Generator().manual_seed(11)
    batches = []
    for bi in range(10):
        data = torch.rand(32, 3, 32, 32, generator=gen)
        label = torch.randint(0, 10, (32,), generator=gen)
        batches.append((torch.arange(bi * 32, (bi + 1) * 32) + 1000, data, label))
    out = {}
    for mode in (False, True):
        tr = Trainer(net, None, _Loader([], 64), _Loader(batches, 32), None, DEV, score_from_features=mode)
        ema = EMAverage(device=DEV)
        torch.manual_seed(5)
        out[mode] = [tuple(t.clone() for t in tr.update_samples(ema)) for _ in range(2)]
        assert (tr._pool.logits.dim() == 1) == mode
    for a, b in zip(out[False], out[True]):
        assert torch.equal(a[3], b[3]) and torch.equal(a[2], b[2]) and torch.equal(a[1], b[1])
        np.testing.assert_allclose(a[0].cpu().numpy(), b[0].cpu().numpy(), rtol=1e-5)
        np.testing.assert_allclose(float(a[4]), float(b[4]), rtol=1e-6)


def test_trainer_graphed_sampling_equals_eager():
    """graph_sampling=True: the fused sampling call replayed from a CUDA graph over the static pool gives the eager results."""
    from mercury_b200.collab import Trainer
    from mercury_b200.sampler import EMAverage

    torch.manual_seed(4)
    net = TinyResNet(10).to(DEV).eval()
    gen = torch.Generator().manual_seed(12)
    batches = [(torch.arange(bi * 32, (bi + 1) * 32), torch.rand(32, 3, 32, 32, generator=gen), torch.randint(0, 10, (32,), generator=gen))
               for bi in range(30)]
    res = {}
    for graphed in (False, True):
        tr = Trainer(net, None, _Loader([], 64), _Loader(batches, 32), None, DEV)
        tr.graph_sampling = graphed
        ema = EMAverage(device=DEV)
        torch.manual_seed(6)
        res[graphed] = []
        for _ in range(4):
            o = tr.update_samples(ema)
            torch.cuda.synchronize()
            res[graphed].append(tuple(t.clone() for t in o) + (ema.value.clone(),))
    for a, b in zip(res[False], res[True]):
        for x, y in zip(a, b):
            assert torch.equal(x, y)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _grad_worker(rank, world, port, ret):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from mercury_b200.collab import Trainer

    torch.manual_seed(0)
    net = TinyResNet(10).to(dev)
    tr = Trainer(net, torch.optim.SGD(net.parameters(), lr=0.1), None, None, None, dev)
    g = torch.Generator().manual_seed(100 + rank)
    ok = []
    for step in range(3):
        x, y = torch.rand(8, 3, 32, 32, generator=g).to(dev), torch.randint(0, 10, (8,), generator=g).to(dev)
        tr._zero_grads()
        F.cross_entropy(net(x), y).backward()
        flat = torch.cat([p.grad.reshape(-1).clone() for p in net.parameters()])   # the reference's procedure (pytorch_collab.py:236-249)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat /= float(world)
        tr.average_gradients()
        got = torch.cat([p.grad.reshape(-1) for p in net.parameters()])
        ok.append(bool(torch.allclose(got, flat, rtol=1e-6, atol=1e-9)))
    ret[rank] = ok
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_gradient_bucket_average_over_nccl_matches_the_reference_procedure():
    import torch.multiprocessing as mp

    ret = mp.Manager().dict()
    mp.spawn(_grad_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
    assert all(ret[0]) and all(ret[1])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_one_process_two_devices_cdf_and_static_cdf():
    """One process, cuda:0 then cuda:1: the >48 KB shared-memory opt-in of the spine and the static scratch address are
    per-device state -- mercury_cdf, mercury_cdf_static and the whole step must work on the second device as on the first."""
    from mercury_b200 import _lib, ops

    lib = _lib.load()
    g = torch.Generator().manual_seed(9)
    w_cpu = torch.rand(300_000, generator=g) + 0.01
    chain = np.cumsum(w_cpu.numpy(), dtype=np.float32)
    want = (chain / chain[-1]).astype(np.float32)
    logits = torch.randn(100_000, 10, generator=g) * 2
    labels = torch.randint(0, 10, (100_000,), generator=g)
    u = torch.rand(64, dtype=torch.float64, generator=g)
    results = []
    for d in (0, 1):
        dev = torch.device("cuda", d)
        with torch.cuda.device(dev):
            w = w_cpu.to(dev)
            cdf, total = ops.cdf(w, "exact")
            assert np.array_equal(cdf.cpu().numpy().view(np.uint32), want.view(np.uint32)), d
            out, tot, status = torch.empty_like(w), torch.empty(1, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)
            rc = lib.mercury_cdf_static(w.data_ptr(), w.numel(), 0, out.data_ptr(), tot.data_ptr(), status.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream)
            assert rc == 0, lib.mercury_last_error()
            torch.cuda.synchronize()
            assert np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32)), d
            for fused in (-1, 1):
                ops.set_fused(fused)
                r = ops.update_samples(logits.to(dev), labels.to(dev), u.to(dev), torch.zeros(1, device=dev), True)
                ops.check_status(r.status)
                results.append(r.idx.cpu())
            ops.set_fused(-1)
    assert all(torch.equal(results[0], x) for x in results[1:])


def test_chunk_ring_streamed_shard_equals_resident_step():
    """A shard streamed from pinned host memory through a 3-window ring (H2D on a copy stream behind K1): scores bit-equal to
    the resident call, and the step on those scores draws what the whole-step call draws."""
    from mercury_b200 import ops
    from mercury_b200.pool import ChunkRing

    g = torch.Generator().manual_seed(13)
    n, c, b, win = 300_000, 1000, 128, 1 << 15
    logits = (torch.randn(n, c, generator=g) * 2).to(torch.bfloat16).pin_memory()
    labels = torch.randint(0, c, (n,), generator=g).pin_memory()
    u = torch.rand(b, dtype=torch.float64, generator=g).to(DEV)
    ring = ChunkRing(win, c, dtype=torch.bfloat16, device=DEV, depth=3)
    scores = torch.empty(n, device=DEV)
    labs = torch.empty(n, dtype=torch.int64, device=DEV)
    for rep in range(2):          # (second pass: every window of the ring is reused behind its "free" event)
        scores.zero_()
        rows = ring.score(((logits[i:i + win], labels[i:i + win]) for i in range(0, n, win)), scores, labs)
        assert rows == n
        whole = ops.update_samples(logits.to(DEV), labels.to(DEV), u, torch.zeros(1, device=DEV), True)
        assert torch.equal(scores, whole.scores) and torch.equal(labs.cpu(), labels)
        r = ops.update_samples(None, labs, u, torch.zeros(1, device=DEV), True, scores_in=scores)
        ops.check_status(r.status)
        assert torch.equal(r.idx, whole.idx) and torch.equal(r.nprob, whole.nprob)


def _global_trainer_worker(rank, world, port, ret):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from mercury_b200.collab import Trainer
    from mercury_b200.sampler import EMAverage

    torch.manual_seed(0)
    net = nn.Sequential(nn.Flatten(), nn.Linear(3 * 32 * 32, 10)).to(dev)
    g = torch.Generator().manual_seed(50 + rank)
    batches = [(torch.arange(i * 625, (i + 1) * 625) + 100000 * rank, torch.rand(625, 3, 32, 32, generator=g),
                torch.randint(0, 10, (625,), generator=g)) for i in range(10)]      # 6250 candidates per worker (cfg4's shard)
    tr = Trainer(net, None, _Loader([], 16), _Loader(batches, 625), None, dev, global_draw=True, sharded_kwargs={"exchange": "peer"})
    ema = EMAverage(device=dev)
    torch.manual_seed(77)
    outs = []
    for step in range(2):
        probs, data, label, index, avg = tr.update_samples(ema)
        torch.cuda.synchronize()
        tr._sp.check()
        outs.append(dict(probs=probs.cpu(), data=data.cpu(), index=index.cpu(), idx=tr.last.idx.cpu(), scores=tr.last.scores.cpu(),
                         stats=tr.last.stats.cpu(), pool_data=tr._pool.tensors()[2].cpu(), pool_index=tr._pool.tensors()[3].cpu()))
    ret[rank] = outs
    dist.barrier()
    tr._sp.peer.close()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_trainer_global_draw_two_gpus_matches_the_two_level_oracle():
    """BASELINE configs[3] shape per worker (6250 candidates, 10 classes) on two real GPUs: Trainer(global_draw=True) over the
    sharded pool with the in-kernel peer exchange; each worker's B rows are its slice of the oracle's two-level draw."""
    import torch.multiprocessing as mp

    world, B, nl = 2, 16, 6250
    ret = mp.Manager().dict()
    mp.spawn(_global_trainer_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    torch.manual_seed(77)
    for step in range(2):
        u = torch.rand(world * B, dtype=torch.float64).numpy()
        st = ret[0][step]["stats"].numpy()
        shards = [((ret[r][step]["scores"].numpy() + np.float32(st[2])).astype(np.float32) / np.float32(st[3])).astype(np.float32)
                  for r in range(world)]
        want = orc.sharded_draw_two_level(shards, u)
        p_all = np.concatenate(shards)
        for r in range(world):
            o = ret[r][step]
            mine = want[r * B:(r + 1) * B]
            assert np.array_equal(o["idx"].numpy(), mine)
            np.testing.assert_array_equal(o["probs"].numpy(), p_all[mine] * np.float32(world * nl))
            for j, gidx in enumerate(mine):
                owner, local = divmod(int(gidx), nl)
                assert torch.equal(o["data"][j], ret[owner][step]["pool_data"][local])
                assert int(o["index"][j]) == int(ret[owner][step]["pool_index"][local])
