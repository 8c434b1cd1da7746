"""GPU tests of the reference-shaped Python surface (Trainer.update_samples, EMAverage,
Groupwise_Sampler, reweighted loss, ShardedPool) against the reference's own recorded run
(tests/golden/ref_*.npz) and the oracle."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from oracle import mercury_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


class ReplayNet(nn.Module):
    """Plays back the logits the reference's ResNet18 produced, batch by batch (the model is outside the path)."""

    def __init__(self, logits):
        super().__init__()
        self.logits = torch.from_numpy(np.asarray(logits)).to(DEV)
        self.pos = 0

    def forward(self, x):
        out = self.logits[self.pos:self.pos + x.shape[0]]
        self.pos += x.shape[0]
        return out


class _Loader:
    def __init__(self, batches, batch_size):
        self.batches, self.batch_size = batches, batch_size

    def __iter__(self):
        return iter(self.batches)


def _pool_batches(seed):
    gen = torch.Generator().manual_seed(seed)
    out = []
    for b in range(10):
        data = torch.rand(32, 3, 32, 32, generator=gen)
        label = torch.randint(0, 10, (32,), generator=gen)
        out.append((torch.arange(b * 32, (b + 1) * 32) + 1000, data, label))
    return out


@pytest.mark.parametrize("name", ["ref_update_samples_linear.npz", "ref_update_samples_resnet18.npz"])
def test_trainer_update_samples_equals_the_reference_run(golden_dir, name):
    from mercury_b200.collab import Trainer
    from mercury_b200.sampler import EMAverage

    g = np.load(os.path.join(golden_dir, name))
    batches = _pool_batches(int(g["pool_seed"]))
    net = ReplayNet(np.concatenate([g["call0_logits"], g["call1_logits"]]))
    tr = Trainer(net, None, _Loader([], int(g["draw_batch"])), _Loader(batches, 32), None, DEV)
    ema = EMAverage(device=DEV)
    torch.manual_seed(int(g["rng_seed"]))  # the generator torch.multinomial drew from in the reference run
    for call in range(2):
        probs, data, label, index, avg = tr.update_samples(ema)
        assert np.array_equal(index.cpu().numpy(), g[f"call{call}_index"])
        assert np.array_equal(label.cpu().numpy(), g[f"call{call}_label"])
        np.testing.assert_allclose(probs.cpu().numpy(), g[f"call{call}_probs"], rtol=1e-5)
        np.testing.assert_allclose(data.double().sum(dim=(1, 2, 3)).cpu().numpy(), g[f"call{call}_data_rowsum"], rtol=1e-12)
        np.testing.assert_allclose(float(avg), g[f"call{call}_avg"], rtol=1e-5)
        np.testing.assert_allclose(float(ema.value), g[f"call{call}_ema"], rtol=1e-5)


def test_emaverage_follows_util_py():
    from mercury_b200.sampler import EMAverage

    e, ref = EMAverage(device=DEV), orc.EMAverage()
    for v in [2.0, 4.0, 0.5, 7.25]:
        e.update(v)
        ref.update(torch.tensor(v))
        np.testing.assert_allclose(float(e.value), float(ref.value), rtol=1e-6)
    assert not e.first_update and "{:.6f}".format(float(ref.value)) == str(e)


def test_groupwise_sampler_equals_the_reference_run(golden_dir):
    from mercury_b200.sampler import Groupwise_Sampler

    g = np.load(os.path.join(golden_dir, "ref_groupwise.npz"))

    class DS:
        def __init__(self):
            gen = torch.Generator().manual_seed(701)
            self.data = torch.rand(192, 3, 32, 32, generator=gen)
            self.targets = torch.randint(0, 10, (192,), generator=gen)

        def __len__(self):
            return 192

        def get_slice(self, s, e):
            return self.data[s:e], self.targets[s:e]

    ds = DS()
    assert np.array_equal(ds.targets.numpy(), g["labels"])
    for tag, repl in (("repl", True), ("norepl", False)):
        net = ReplayNet(g["logits"][:192])
        s = Groupwise_Sampler(ds, replacement=repl, device=DEV)
        s.update_importance(0, 64, net)
        s.update_importance(0, 64, net)
        s.update_importance(1, 64, net)
        np.testing.assert_allclose(s.importance.cpu().numpy(), g[f"{tag}_importance"], rtol=1e-5, atol=1e-6)
        assert np.array_equal(s.group_indicator.cpu().numpy(), g[f"{tag}_group_indicator"].astype(np.int32))
        assert s.group_index == int(g[f"{tag}_group_index"])
        torch.manual_seed(int(g["rng_seed"]))
        it = iter(s)
        ys = [int(next(it)) for _ in range(len(g[f"{tag}_yield"]))]
        assert ys == list(g[f"{tag}_yield"])
    few = s.draw(8).cpu().numpy()
    assert len(set(few.tolist())) == 8 and few.min() >= 128  # one race, no replacement, current group only


def test_reweighted_mean_matches_autograd():
    from mercury_b200.sampler import reweighted_mean

    gen = torch.Generator().manual_seed(1)
    losses = (torch.rand(64, generator=gen) * 3).to(DEV).requires_grad_(True)
    nprob = (torch.rand(64, generator=gen) + 0.5).to(DEV)
    seed = (1.0 / 64) / nprob
    out = reweighted_mean(losses, seed)
    (out * 2.0).backward()
    ref_loss, ref_seed = orc.reweighted_loss(losses.detach().cpu(), nprob.cpu())
    np.testing.assert_allclose(float(out), float(ref_loss), rtol=1e-5)
    np.testing.assert_allclose(losses.grad.cpu().numpy(), 2.0 * ref_seed.numpy(), rtol=1e-5)


def test_trainer_runs_a_few_importance_sampled_steps():
    """cfg1-shaped smoke of the drop-in Trainer: synthetic CIFAR loaders, small CNN, 3 steps."""
    from mercury_b200.collab import Trainer
    from mercury_b200.exp_dataset import load_cifar10

    torch.manual_seed(0)
    train_loader, presam_loader, test_loader = load_cifar10(1, n_train=1280, n_test=64)
    net = nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.AdaptiveAvgPool2d(4), nn.Flatten(), nn.Linear(128, 10)).to(DEV)
    opt = torch.optim.Adam(net.parameters(), lr=1e-3)
    tr = Trainer(net, opt, train_loader, presam_loader, test_loader, DEV)
    before = [p.detach().clone() for p in net.parameters()]
    loss, _, ema = tr.train(max_steps=3)
    assert torch.isfinite(loss) and tr.step == 4 and float(ema.value) > 0
    assert any(not torch.equal(a, b) for a, b in zip(before, net.parameters()))
    r = tr.last
    assert r.idx.shape == (32,) and int(r.idx.max()) < 320 and r.data.shape == (32, 3, 32, 32)
    tl, acc = tr.evaluate()
    assert np.isfinite(tl) and 0.0 <= acc <= 1.0


def test_sharded_pool_single_rank_equals_flat_path():
    from mercury_b200 import ops
    from mercury_b200.sharded import ShardedPool

    gen = torch.Generator().manual_seed(5)
    n, c, b = 30000, 10, 96
    logits = (torch.randn(n, c, generator=gen) * 2).to(DEV)
    labels = torch.randint(0, c, (n,), generator=gen).to(DEV)
    data = torch.rand(n, 3, 8, 8, generator=gen).to(DEV)
    index = torch.randperm(n, generator=gen).to(DEV)
    u = torch.rand(b, dtype=torch.float64, generator=gen).to(DEV)
    sp = ShardedPool(n, DEV)
    flat_ema = torch.zeros(1, device=DEV)
    for step in range(2):
        r = sp.update_samples(logits, labels, u, data=data, index=index)
        f = ops.update_samples(logits, labels, u, flat_ema, step == 0, data=data, index=index)
        assert torch.equal(r["idx"], f.idx) and torch.equal(r["nprob"], f.nprob) and torch.equal(r["seed"], f.seed)
        assert torch.equal(r["data"], f.data) and torch.equal(r["labels"], f.labels) and torch.equal(r["index"], f.index)
        assert int(r["owned"].sum()) == b
        logits = logits + 0.25


def test_gather_owned_zero_fills_foreign_rows():
    from mercury_b200 import ops

    gen = torch.Generator().manual_seed(6)
    data = torch.rand(100, 7, generator=gen).to(DEV)
    labels = torch.randint(0, 10, (100,), generator=gen).to(DEV)
    idx = torch.tensor([5, 250, 199, 200, 299, 300, 0], device=DEV)
    d, l, ix, owned = ops.gather_owned(idx, 200, 100, data, labels, None)
    assert owned.tolist() == [0, 1, 0, 1, 1, 0, 0]
    assert torch.equal(d[1], data[50]) and torch.equal(d[3], data[0]) and torch.equal(d[4], data[99])
    assert float(d[0].abs().sum() + d[2].abs().sum() + d[5].abs().sum() + d[6].abs().sum()) == 0.0
    assert l.tolist() == [0, int(labels[50]), 0, int(labels[0]), int(labels[99]), 0, 0] and ix is None


def test_graphed_update_equals_eager_calls():
    from mercury_b200 import ops

    gen = torch.Generator().manual_seed(12)
    n, c, b = 40000, 10, 64
    logits = (torch.randn(n, c, generator=gen) * 2).to(DEV)
    labels = torch.randint(0, c, (n,), generator=gen).to(DEV)
    data = torch.rand(n, 16, generator=gen).to(DEV)
    u = torch.rand(b, dtype=torch.float64, generator=gen).to(DEV)
    ema_g, ema_e = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
    g = ops.GraphedUpdate(logits, labels, u, ema_g, data=data)   # one eager step (first EMA update), then capture
    ops.update_samples(logits, labels, u, ema_e, True, data=data)  # the same first step on the eager side
    assert torch.equal(ema_g, ema_e)
    for step in range(3):
        logits.add_(0.1 * (step + 1))                              # refill the static inputs in place
        u.copy_(torch.rand(b, dtype=torch.float64, generator=gen))
        r = g.replay()
        torch.cuda.synchronize()
        e = ops.update_samples(logits, labels, u, ema_e, False, data=data)
        assert torch.equal(r.idx, e.idx) and torch.equal(r.nprob, e.nprob) and torch.equal(r.seed, e.seed)
        assert torch.equal(r.data, e.data) and torch.equal(r.labels, e.labels) and torch.equal(ema_g, ema_e)
    ops.check_status(g.status)


@pytest.mark.parametrize("G,n_local,b", [(3, 40000, 200), (8, 6250, 128)])   # (8 x 6250: BASELINE configs[3], the 50k pool over 8 workers)
def test_two_level_sharded_draw_matches_its_oracle(G, n_local, b):
    """Mode "totals" (north-star literal): the shards emulated one after the other on one GPU --
    the kernels are the ones ShardedPool(mode="totals") calls, the exchanges are plain concatenations."""
    from mercury_b200 import _lib, ops

    gen = torch.Generator().manual_seed(21)
    c = 10
    N = G * n_local
    logits = [(torch.randn(n_local, c, generator=gen) * 2).to(DEV) for _ in range(G)]
    labels = [torch.randint(0, c, (n_local,), generator=gen).to(DEV) for _ in range(G)]
    data = [torch.rand(n_local, 6, generator=gen).to(DEV) for _ in range(G)]
    u = torch.rand(b, dtype=torch.float64, generator=gen).to(DEV)
    none = torch.empty(0, dtype=torch.float64, device=DEV)
    scores = [ops.score(logits[r], labels[r]) for r in range(G)]
    sums_all = torch.cat([ops.sum_scores(s) for s in scores])              # "all-gather" of the shard sums
    np.testing.assert_allclose(sums_all.cpu().numpy(), [float(s.double().sum()) for s in scores], rtol=1e-12)

    def chain(r):
        return ops.update_samples(None, None, none, torch.zeros(1, device=DEV), True, scores_in=scores[r],
                                  global_sums=sums_all, n_global=N)

    stats = [chain(r).stats.clone() for r in range(G)]
    for r in range(1, G):                                                   # whole-pool statistics: identical on all ranks
        assert torch.equal(stats[r][:4], stats[0][:4])
    allscores = torch.cat(scores).cpu()
    np.testing.assert_allclose(float(stats[0][0]), float(allscores.mean()), rtol=1e-6)
    totals_all = torch.stack([s[_lib.STAT_CHAIN_TOTAL] for s in stats])    # "all-gather" of the chain totals
    shift, total = np.float32(stats[0][2].item()), np.float32(stats[0][3].item())
    p_shards = [((s.cpu().numpy() + shift).astype(np.float32) / total).astype(np.float32) for s in scores]
    for r in range(G):
        assert totals_all[r].cpu().numpy().view(np.uint32) == orc.serial_chain_f32(p_shards[r])[-1:].view(np.uint32)
    want = orc.sharded_draw_two_level(p_shards, u.cpu().numpy())
    got_idx = torch.zeros(b, dtype=torch.int64, device=DEV)
    got_np = torch.zeros(b, device=DEV)
    got_data = torch.zeros(b, 6, device=DEV)
    owners = torch.zeros(b, dtype=torch.int64, device=DEV)
    for r in range(G):
        local_u = ops.shard_select(totals_all, r, u)
        st = chain(r)                                                       # (this rank's boundaries back into the workspace)
        d = ops.update_samples(None, labels[r], local_u, torch.zeros(1, device=DEV), False, scores_in=scores[r], data=data[r],
                               global_sums=sums_all, n_global=N, index_base=r * n_local, draw_only=True, out={"stats": st.stats})
        got_idx += d.idx
        got_np += d.nprob
        got_data += d.data
        owners += (local_u >= 0).long()
    assert torch.all(owners == 1)                                           # every draw has exactly one owner
    assert np.array_equal(got_idx.cpu().numpy(), want)
    p_all = np.concatenate(p_shards)
    np.testing.assert_allclose(got_np.cpu().numpy(), p_all[want] * np.float32(N), rtol=1e-6)
    assert torch.equal(got_data.cpu(), torch.cat(data).cpu()[want])
    # with a single shard the two-level draw is the flat draw
    one = ops.update_samples(None, None, u, torch.zeros(1, device=DEV), True, scores_in=scores[0])
    su = ops.sum_scores(scores[0])
    st1 = ops.update_samples(None, None, none, torch.zeros(1, device=DEV), True, scores_in=scores[0], global_sums=su, n_global=n_local)
    lu = ops.shard_select(st1.stats[_lib.STAT_CHAIN_TOTAL:_lib.STAT_CHAIN_TOTAL + 1].clone(), 0, u)
    two = ops.update_samples(None, None, lu, torch.zeros(1, device=DEV), False, scores_in=scores[0], global_sums=su,
                             n_global=n_local, draw_only=True, out={"stats": st1.stats})
    assert torch.equal(one.idx, two.idx) and torch.equal(one.nprob, two.nprob)


@pytest.mark.parametrize("fused", [True, False])
def test_peer_exchange_world1_equals_collective_path(fused):
    """exchange="peer" with a single rank: the kernels store into / poll their own box.  Three steps in a
    row (both parities of the double-buffered slots, EMA carried) must equal the exchange="nccl" path."""
    from mercury_b200 import ops
    from mercury_b200.sharded import ShardedPool

    gen = torch.Generator().manual_seed(33)
    n, c, b = 50000, 10, 96
    data = torch.rand(n, 5, generator=gen).to(DEV)
    index = torch.randperm(n, generator=gen).to(DEV)
    u = torch.rand(b, dtype=torch.float64, generator=gen).to(DEV)
    a = ShardedPool(n, DEV, mode="totals", exchange="nccl")
    p = ShardedPool(n, DEV, mode="totals", exchange="peer", fused=fused)
    try:
        for step in range(3):
            logits = (torch.randn(n, c, generator=gen) * 2).to(DEV)
            labels = torch.randint(0, c, (n,), generator=gen).to(DEV)
            ra = a.update_samples(logits, labels, u, data=data, index=index)
            rp = p.update_samples(logits, labels, u, data=data, index=index)
            for k in ("idx", "nprob", "seed", "data", "labels", "index", "stats"):
                assert torch.equal(ra[k], rp[k]), (step, k)
            assert torch.equal(a.sums_all, p.sums_all) and torch.equal(a.totals_all, p.totals_all)
            ops.check_status(rp["xstatus"])
            ops.check_status(rp["status"])
    finally:
        p.peer.close()


@pytest.mark.parametrize("path", ["staged", "chain", "one_kernel"])
def test_peer_timeout_poisons_the_step_and_raises_then_reset_recovers(path):
    """A peer that never shows up (a second, silent box on the same GPU stands in for rank 1 of 2): the wait gives up after
    the configured bound, the step's outputs are poisoned (idx -1, NaN nprob/seed -- they would survive the assembling SUM
    all-reduce), the next call raises, and after peer.reset() the exchange works again."""
    from mercury_b200 import _lib, ops
    from mercury_b200.sharded import PeerExchange, ShardedPool

    gen = torch.Generator().manual_seed(34)
    n, c, b = 30000, 10, 64
    u = torch.rand(b, dtype=torch.float64, generator=gen).to(DEV)
    logits = (torch.randn(n, c, generator=gen) * 2).to(DEV)
    labels = torch.randint(0, c, (n,), generator=gen).to(DEV)
    a = ShardedPool(n, DEV, mode="totals", exchange="nccl")
    p = ShardedPool(n, DEV, mode="totals", exchange="peer", fused=path != "staged")
    silent = PeerExchange(DEV)
    ops.set_fused(1 if path == "one_kernel" else -1)
    try:
        good = (p.peer.boxes, p.sums_all, p.totals_all)
        p.peer.set_wait(40)
        p.peer.boxes = torch.tensor([p.peer._box, silent._box], dtype=torch.int64, device=DEV)
        p.sums_all, p.totals_all = torch.zeros(2, dtype=torch.float64, device=DEV), torch.zeros(2, device=DEV)
        r = p.update_samples(logits, labels, u)
        torch.cuda.synchronize()
        assert int(r["xstatus"].item()) & _lib.ST_PEER_TIMEOUT
        assert bool((r["idx"] == -1).all()) and bool(torch.isnan(r["nprob"]).all()) and bool(torch.isnan(r["seed"]).all())
        with pytest.raises(RuntimeError, match="did not arrive"):
            p.update_samples(logits, labels, u)          # (the check on entry sees the finished, timed-out step)
        p.peer.boxes, p.sums_all, p.totals_all = good
        p.peer.reset()
        r["xstatus"].zero_()
        r["status"].zero_()
        p.ema_first = True
        ra, rp = a.update_samples(logits, labels, u), p.update_samples(logits, labels, u)
        p.check()
        assert torch.equal(ra["idx"], rp["idx"])
        if path != "one_kernel":
            assert torch.equal(ra["nprob"], rp["nprob"])
        ops.check_status(rp["status"])
    finally:
        ops.set_fused(-1)
        silent.close()
        p.peer.close()


@pytest.mark.parametrize("n,groups", [(192, 3), (50000, 4), (1000000, 7)])
def test_groupwise_draw_kernel_matches_the_restated_reference(n, groups):
    """mercury_groupwise_draw against the restatement of util.py:140-157 (oracle.groupwise_probabilities + the exponential
    race on the compacted group): probabilities to one fp32 rounding, single draws and a without-replacement draw equal
    index for index (ties in the race aside, which the inputs here do not produce)."""
    from mercury_b200 import ops

    g = torch.Generator().manual_seed(n)
    imp = (torch.rand(n, generator=g) * 3 + 1e-3).float()
    tags = torch.randint(1, groups + 1, (n,), generator=g).to(torch.int32)
    group = 2
    members = np.nonzero(tags.numpy() == group)[0]
    count = len(members)
    p_ref = orc.groupwise_probabilities(imp.numpy()[members].astype(np.float64)).numpy()
    ndraws = 5
    u = torch.rand(ndraws * count, dtype=torch.float64, generator=g)
    st = ops.new_status(DEV)
    idx, p = ops.groupwise_draw(imp.to(DEV), tags.to(DEV), group, count, u.to(DEV), ndraws=ndraws, want_p=True, status=st)
    ops.check_status(st)
    p = p.cpu().numpy()
    assert np.all(p[tags.numpy() != group] == 0)
    np.testing.assert_allclose(p[members], p_ref, rtol=2e-7)
    for j in range(ndraws):
        keys = orc.exp_race_keys(torch.from_numpy(p[members]), u[j * count:(j + 1) * count].numpy())
        assert int(idx[j]) == int(members[int(torch.argmax(keys))])
    k = 16
    idxk = ops.groupwise_draw(imp.to(DEV), tags.to(DEV), group, count, u[:count].to(DEV), k=k, status=st)
    ops.check_status(st)
    keys = orc.exp_race_keys(torch.from_numpy(p[members]), u[:count].numpy())
    want = members[torch.topk(keys, k).indices.numpy()]
    assert np.array_equal(idxk.cpu().numpy(), want)
    # a wrong member count is reported, not silently used
    ops.groupwise_draw(imp.to(DEV), tags.to(DEV), group, count - 1, u.to(DEV), status=st)
    with pytest.raises(RuntimeError, match="member count"):
        ops.check_status(st)
