"""Pins the oracle: against the reference's own run (tests/golden/ref_*.npz), against the ATen
known-answer vectors (tests/golden/aten_*.npz) and against the installed torch CPU ops live."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from oracle import c_oracle
from oracle import mercury_oracle as orc


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


class ReplayNet(nn.Module):
    """Returns the logits the reference's model produced, batch by batch."""

    def __init__(self, logits, batch=32):
        super().__init__()
        self.logits = torch.from_numpy(np.asarray(logits))
        self.batch = batch
        self.pos = 0

    def forward(self, x):
        out = self.logits[self.pos:self.pos + x.shape[0]]
        self.pos += x.shape[0]
        return out


def _pool_batches(g):
    """The loader of make_golden.synthetic_pool, rebuilt from its seed."""
    gen = torch.Generator().manual_seed(int(g["pool_seed"]))
    batches = []
    for b in range(10):
        data = torch.rand(32, 3, 32, 32, generator=gen)
        label = torch.randint(0, 10, (32,), generator=gen)
        idx = torch.arange(b * 32, (b + 1) * 32) + 1000
        batches.append((idx, data, label))
    return batches


@pytest.mark.parametrize("name", ["ref_update_samples_linear.npz", "ref_update_samples_resnet18.npz"])
def test_faithful_restatement_matches_reference_run(golden_dir, name):
    g = _load(golden_dir, name)
    batches = _pool_batches(g)
    assert np.array_equal(torch.cat([b[2] for b in batches]).numpy(), g["pool_labels"])
    ema = orc.EMAverage()
    torch.manual_seed(int(g["rng_seed"]))
    for call in range(2):
        net = ReplayNet(g[f"call{call}_logits"])
        probs, data, label, index, avg = orc.update_samples_faithful(net, iter(batches), int(g["draw_batch"]), ema)
        assert np.array_equal(index.numpy(), g[f"call{call}_index"])
        assert np.array_equal(label.numpy(), g[f"call{call}_label"])
        np.testing.assert_allclose(probs.numpy(), g[f"call{call}_probs"], rtol=1e-6)
        np.testing.assert_allclose(data.double().sum(dim=(1, 2, 3)).numpy(), g[f"call{call}_data_rowsum"], rtol=1e-12)
        np.testing.assert_allclose(float(avg), g[f"call{call}_avg"], rtol=1e-6)
        np.testing.assert_allclose(float(ema.value), g[f"call{call}_ema"], rtol=1e-6)


def test_explicit_uniform_draw_matches_reference_run(golden_dir):
    """The last of the 10 draws of each call, re-made from explicit uniforms."""
    g = _load(golden_dir, "ref_update_samples_resnet18.npz")
    B = int(g["draw_batch"])
    torch.manual_seed(int(g["rng_seed"]))
    u_all = torch.rand(20 * B, dtype=torch.float64).numpy()
    labels = torch.from_numpy(g["pool_labels"])
    ema = orc.EMAverage()
    for call in range(2):
        logits = torch.from_numpy(g[f"call{call}_logits"])
        for k in range(1, 11):  # growing pools, EMA updated each time (pytorch_collab.py:95-112)
            losses = orc.score(logits[:32 * k], labels[:32 * k])
            _, _, p = orc.pool_stats(losses, ema)
        u = u_all[(call * 10 + 9) * B:(call * 10 + 10) * B]
        idx = orc.draw_with_replacement(p.numpy(), u)
        assert np.array_equal(g["pool_index"][idx], g[f"call{call}_index"])
        nprob = orc.gather_reweight(p, torch.from_numpy(idx))[0]
        np.testing.assert_allclose(nprob.numpy(), g[f"call{call}_probs"], rtol=1e-6)


def test_groupwise_restatement_matches_reference_run(golden_dir):
    g = _load(golden_dir, "ref_groupwise.npz")
    for tag in ("repl", "norepl"):
        imp, ind, gi = g[f"{tag}_importance"], g[f"{tag}_group_indicator"], int(g[f"{tag}_group_index"])
        member = ind == gi
        loc = member.nonzero()[0]
        n_group = int(member.sum())
        torch.manual_seed(int(g["rng_seed"]))
        ys = []
        for _ in range(len(g[f"{tag}_yield"])):
            u = torch.rand(n_group, dtype=torch.float64).numpy()
            ys.append(int(loc[orc.groupwise_draw(imp[member], u)]))
        assert ys == list(g[f"{tag}_yield"])
    # the importance array holds the reference's CE losses of its three slices
    losses = orc.score(torch.from_numpy(g["logits"][:192]), torch.from_numpy(g["labels"]))
    np.testing.assert_allclose(g["repl_importance"], losses.numpy().astype(np.float64), rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("n", [17, 320, 5000, 50000])
def test_draws_match_aten_golden(golden_dir, n):
    g = _load(golden_dir, "aten_multinomial.npz")
    p, seed = g[f"n{n}_p"], int(g[f"n{n}_seed"])
    k = len(g[f"n{n}_with"])
    u = torch.rand(k, dtype=torch.float64, generator=torch.Generator().manual_seed(seed)).numpy()
    assert np.array_equal(orc.draw_with_replacement(p, u), g[f"n{n}_with"])
    st, chain, cdf = c_oracle.serial_cdf(p)
    assert st == 0 and np.array_equal(cdf, orc.serial_cdf_f32(p))
    assert np.array_equal(c_oracle.lower_bound(cdf, u), g[f"n{n}_with"])
    k2 = len(g[f"n{n}_without"])
    u2 = torch.rand(n, dtype=torch.float64, generator=torch.Generator().manual_seed(seed)).numpy()
    assert np.array_equal(orc.draw_without_replacement(p, u2, k2), g[f"n{n}_without"])
    if n <= 5000:
        keys, idx = c_oracle.exp_race_topk(p, u2, k2)
        assert np.array_equal(keys, orc.exp_race_keys(p, u2).numpy())
        assert np.array_equal(idx, g[f"n{n}_without"])


@pytest.mark.parametrize("n,k", [(1000, 64), (200000, 256)])
def test_draws_match_torch_live(n, k):
    w = torch.exp(torch.randn(n, generator=torch.Generator().manual_seed(5)) * 1.5)
    p = w / w.sum()
    g1, g2 = torch.Generator().manual_seed(77), torch.Generator().manual_seed(77)
    ref = torch.multinomial(p, k, True, generator=g1)
    u = orc.uniforms_like_multinomial(k, g2)
    assert np.array_equal(orc.draw_with_replacement(p.numpy(), u.numpy()), ref.numpy())
    assert torch.equal(g1.get_state(), g2.get_state())
    g1, g2 = torch.Generator().manual_seed(78), torch.Generator().manual_seed(78)
    ref = torch.multinomial(p, k, False, generator=g1)
    u = orc.uniforms_like_multinomial(n, g2)
    assert np.array_equal(orc.draw_without_replacement(p.numpy(), u.numpy(), k), ref.numpy())


@pytest.mark.parametrize("name,tol", [("c10_f32", 1e-5), ("c10_f32_wide", 1e-5), ("c1000_bf16", 1e-5),
                                      ("c1000_f32", 1e-5), ("c32768_bf16", 1e-5), ("c33_f32", 1e-5)])
def test_score_matches_aten_golden(golden_dir, name, tol):
    g = _load(golden_dir, "aten_cross_entropy.npz")
    raw, y, loss = g[f"{name}_logits"], g[f"{name}_labels"], g[f"{name}_loss"]
    bf16 = raw.dtype == np.int16
    x = torch.from_numpy(raw).view(torch.bfloat16) if bf16 else torch.from_numpy(raw)
    got = orc.score(x, torch.from_numpy(y)).numpy()
    np.testing.assert_allclose(got, loss, rtol=1e-6, atol=1e-6)
    assert got[0] == 0.0  # ignore_index row
    st, c_loss = c_oracle.score(raw, y, bf16=bf16)
    assert st == 0
    # tiny losses lose relative accuracy in fp32: relative on loss, absolute floor 1e-5
    np.testing.assert_allclose(c_loss, g[f"{name}_loss64"], rtol=tol, atol=1e-5)


def test_distribution_checks_have_aten_messages():
    for bad, msg in [([0.5, -0.1], "< 0"), ([0.5, float("inf")], "infinity or NaN"), ([0.0, 0.0], "<= 0")]:
        with pytest.raises(RuntimeError) as a:
            torch.multinomial(torch.tensor(bad), 2, True)
        with pytest.raises(RuntimeError) as b:
            orc.draw_with_replacement(np.asarray(bad, dtype=np.float32), np.asarray([0.3]))
        assert msg in str(b.value)
        assert msg in str(a.value)
    assert c_oracle.serial_cdf(np.asarray([0.5, -0.1], dtype=np.float32))[0] == 1
    assert c_oracle.serial_cdf(np.asarray([0.5, np.nan], dtype=np.float32))[0] == 1  # NaN: `val >= 0` fails first
    assert c_oracle.serial_cdf(np.asarray([0.5, np.inf], dtype=np.float32))[0] == 2
    assert c_oracle.serial_cdf(np.asarray([0.0, 0.0], dtype=np.float32))[0] == 3


def test_reweight_seed():
    g = torch.Generator().manual_seed(3)
    losses = torch.rand(64, generator=g) * 3
    nprob = torch.rand(64, generator=g) + 0.5
    loss, seed = orc.reweighted_loss(losses, nprob)
    np.testing.assert_allclose(seed.numpy(), (1.0 / (64 * nprob)).numpy(), rtol=1e-6)
    np.testing.assert_allclose(float(loss), float((losses / nprob).mean()), rtol=1e-6)
    p = (nprob / nprob.sum()).numpy()
    idx = np.arange(64, dtype=np.int64)
    np_, sd = c_oracle.gather_weights(p, idx)
    np.testing.assert_allclose(sd, 1.0 / (64 * p * 64), rtol=1e-6)


def test_ema_matches_reference_definition():
    e = orc.EMAverage()
    e.update(torch.tensor(2.0))
    assert float(e.value) == 2.0
    e.update(torch.tensor(4.0))
    np.testing.assert_allclose(float(e.value), 0.9 * 2.0 + 0.1 * 4.0, rtol=1e-6)


def test_score_head_is_linear_then_cross_entropy():
    """f4: the fused-head oracle is the reference's two calls, checked against fp64 and against its own ResNet head."""
    g = torch.Generator().manual_seed(8)
    x, w, b = torch.randn(64, 512, generator=g), torch.randn(10, 512, generator=g) * 0.05, torch.randn(10, generator=g) * 0.1
    y = torch.randint(0, 10, (64,), generator=g)
    got = orc.score_head(x, w, b, y)
    ref64 = F.cross_entropy(F.linear(x.double(), w.double(), b.double()), y, reduction="none")
    np.testing.assert_allclose(got.numpy(), ref64.numpy(), rtol=2e-5, atol=2e-5)
    lin = torch.nn.Linear(512, 10)
    with torch.no_grad():
        lin.weight.copy_(w)
        lin.bias.copy_(b)
    assert torch.equal(got, F.cross_entropy(lin(x), y, reduction="none"))
    y2 = y.clone()
    y2[3] = -100
    assert orc.score_head(x, w, None, y2)[3] == 0.0


@pytest.mark.skipif(not os.path.exists("/root/reference/pytorch_model.py"), reason="needs the reference checkout")
def test_split_head_reproduces_the_reference_resnet_forward():
    import importlib.util
    import sys
    spec = importlib.util.spec_from_file_location("ref_pytorch_model", "/root/reference/pytorch_model.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_pytorch_model"] = mod
    spec.loader.exec_module(mod)
    from mercury_b200.collab import split_head

    torch.manual_seed(0)
    net = mod.ResNet18(10).eval()
    x = torch.randn(6, 3, 32, 32)
    y = torch.randint(0, 10, (6,))
    features, lin = split_head(net)
    with torch.no_grad():
        f = features(x)
        assert f.shape == (6, 512)
        assert torch.equal(lin(f), net(x))
        assert torch.equal(orc.score_head(f, lin.weight, lin.bias, y), F.cross_entropy(net(x), y, reduction="none"))


def test_score_head_oracle_reproduces_the_reference_run(golden_dir):
    """tests/golden/ref_head.npz: the reference's ResNet18 forward, hooked at its linear layer."""
    z = np.load(os.path.join(golden_dir, "ref_head.npz"))
    f, w, b, y = (torch.from_numpy(z[k]) for k in ("features", "weight", "bias", "label"))
    assert f.shape == (64, 512) and w.shape == (10, 512)
    assert torch.equal(F.linear(f, w, b), torch.from_numpy(z["logits"]))          # same layer, same arithmetic
    assert torch.equal(orc.score_head(f, w, b, y), torch.from_numpy(z["scores"]))
    assert torch.equal(orc.score(torch.from_numpy(z["logits"]), y), torch.from_numpy(z["scores"]))


def test_c_oracle_head_and_two_level_draw_agree_with_the_python_oracle(golden_dir):
    z = np.load(os.path.join(golden_dir, "ref_head.npz"))
    st, got = c_oracle.score_head(z["features"], z["weight"], z["bias"], z["label"])
    assert st == 0
    np.testing.assert_allclose(got, z["scores"], rtol=2e-5, atol=2e-5)           # (index-order sums vs the BLAS)
    g = torch.Generator().manual_seed(77)
    shards = []
    for n in (5000, 1, 12345, 700):
        shards.append((torch.rand(n, generator=g) + 0.01).numpy().astype(np.float32))
    tot = np.float32(sum(float(s.astype(np.float64).sum()) for s in shards))
    shards = [(s / tot).astype(np.float32) for s in shards]
    u = torch.rand(500, dtype=torch.float64, generator=g).numpy()
    u[0], u[1] = 0.0, 1.0 - 2.0 ** -53
    st, idx = c_oracle.two_level_draw(shards, u)
    assert st == 0
    assert np.array_equal(idx, orc.sharded_draw_two_level(shards, u))


def test_fp64_draw_restatement_is_a_valid_inverse_cdf():
    """draw_with_replacement_f64: lower bound of an fp64 running total (the non-exact mode): agrees with the exact
    serial-fp32 draw on small pools and stays within a few candidates of it on a large one."""
    g = torch.Generator().manual_seed(2)
    for n in (17, 320, 200000):
        w = torch.rand(n, generator=g) + 0.01
        p = (w / w.sum()).numpy()
        u = torch.rand(256, dtype=torch.float64, generator=g).numpy()
        a, b = orc.draw_with_replacement(p, u), orc.draw_with_replacement_f64(p, u)
        assert np.abs(a - b).max() <= (0 if n <= 320 else 8)
