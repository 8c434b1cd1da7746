"""GPU parity tests: every kernel, called through the C ABI, against the CPU oracle.

Tolerances (BASELINE.json north_star): scores / N*p / seed within 1e-5 relative for fp32 logits
(absolute floor 1e-5 for near-zero losses), 1e-2 for bf16 logits against the CPU
bf16 op (1e-5 against the upcast oracle); sampled indices BIT-EXACT given the same importance
vector bits and the same uniform stream.
"""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import c_oracle
from oracle import mercury_oracle as orc

pytestmark = pytest.mark.gpu

RTOL_F32 = 1e-5
ATOL_LOSS = 1e-5


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mercury_b200 import ops as _ops

    return _ops


DEV = "cuda:0"


def _logits(n, c, dtype, sigma=2.0, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(n, c, generator=g) * sigma).to(dtype)
    y = torch.randint(0, c, (n,), generator=g)
    return x, y


# ------------------------------------------------------------------------------------------ K1
@pytest.mark.parametrize("n,c,dtype,sigma", [
    (1000, 10, torch.float32, 2.0),        # small rows, compile-time C, bulk-copy tiles
    (4096 + 37, 10, torch.float32, 8.0),   # ragged last tile, confident rows (tiny losses)
    (1, 10, torch.float32, 2.0),
    (3000, 10, torch.bfloat16, 2.0),
    (777, 33, torch.float32, 3.0),         # small rows, run-time C
    (513, 2, torch.float32, 1.0),
    (600, 100, torch.bfloat16, 2.0),       # 200-byte rows
    (512, 1000, torch.bfloat16, 2.0),      # warp per row, 4 vectors per lane
    (300, 1000, torch.float32, 4.0),       # warp per row, 8 vectors per lane
    (129, 136, torch.float32, 2.0),        # warp per row, 2 vectors per lane, idle lanes
    (40, 4096, torch.bfloat16, 2.0),       # 8 KB rows: 16 vectors per lane
    (64, 32768, torch.bfloat16, 2.0),      # wide rows: CTA per row, bulk ring, online softmax
    (10, 5000, torch.float32, 3.0),        # wide rows fp32 with a partial last chunk
    (50, 1001, torch.bfloat16, 2.0),       # unaligned rows -> scalar kernel
    (33, 70001, torch.float32, 2.0),       # unaligned and wide
])
def test_score_matches_oracle(ops, n, c, dtype, sigma):
    x, y = _logits(n, c, dtype, sigma)
    if n > 3:
        y[1] = -100  # ignore_index
    got = ops.score(x.to(DEV), y.to(DEV)).cpu()
    ref = orc.score(x, y)
    np.testing.assert_allclose(got.numpy(), ref.numpy(), rtol=RTOL_F32, atol=ATOL_LOSS)
    if n > 3:
        assert got[1] == 0.0
    if dtype == torch.bfloat16:  # the reference's literal CPU op returns bf16: 1e-2 budget
        lit = F.cross_entropy(x, y, reduction="none").float()
        np.testing.assert_allclose(got.numpy(), lit.numpy(), rtol=2e-2, atol=2e-2)
    st, cref = c_oracle.score(x.view(torch.int16).numpy() if dtype == torch.bfloat16 else x.numpy(), y.numpy(),
                              bf16=dtype == torch.bfloat16)
    assert st == 0
    np.testing.assert_allclose(got.numpy(), cref, rtol=RTOL_F32, atol=ATOL_LOSS)


def test_score_strided_rows_and_views(ops):
    x, y = _logits(500, 24, torch.float32)
    xd = x.to(DEV)
    wide = torch.zeros(500, 40, device=DEV)
    wide[:, :24] = xd
    view = wide[:, :24]  # row stride 40, not contiguous
    got = ops.score(view, y.to(DEV)).cpu()
    np.testing.assert_allclose(got.numpy(), orc.score(x, y).numpy(), rtol=RTOL_F32, atol=ATOL_LOSS)
    off = torch.zeros(500 * 24 + 1, device=DEV)[1:].view(500, 24)  # 4-byte aligned only
    off.copy_(xd)
    got = ops.score(off, y.to(DEV)).cpu()
    np.testing.assert_allclose(got.numpy(), orc.score(x, y).numpy(), rtol=RTOL_F32, atol=ATOL_LOSS)


def test_score_golden_vectors(ops, golden_dir):
    g = np.load(os.path.join(golden_dir, "aten_cross_entropy.npz"))
    for name in ["c10_f32", "c10_f32_wide", "c1000_bf16", "c1000_f32", "c32768_bf16", "c33_f32"]:
        raw, y, loss = g[f"{name}_logits"], g[f"{name}_labels"], g[f"{name}_loss"]
        x = torch.from_numpy(raw).view(torch.bfloat16) if raw.dtype == np.int16 else torch.from_numpy(raw)
        got = ops.score(x.to(DEV), torch.from_numpy(y).to(DEV)).cpu().numpy()
        np.testing.assert_allclose(got, loss, rtol=RTOL_F32, atol=ATOL_LOSS)
        np.testing.assert_allclose(got, g[f"{name}_loss64"], rtol=RTOL_F32, atol=ATOL_LOSS)


def test_score_bad_label_raises_like_torch(ops):
    x, y = _logits(100, 10, torch.float32)
    y[5] = 10
    with pytest.raises(IndexError):
        F.cross_entropy(x, y, reduction="none")
    st = ops.new_status(DEV)
    ops.score(x.to(DEV), y.to(DEV), status=st)
    with pytest.raises(IndexError):
        ops.check_status(st)


# ------------------------------------------------------------------------------------------ K2
def _weights(kind, n, seed=1):
    g = torch.Generator().manual_seed(seed)
    if kind == "uniform":
        w = torch.rand(n, generator=g) * 3 + 0.01
        return w / w.sum()
    if kind == "lognormal_raw":
        return torch.exp(torch.randn(n, generator=g) * 3)
    if kind == "ce_like":
        w = -torch.log(torch.rand(n, generator=g) + 1e-12) + 1.15
        return w / w.sum()
    if kind == "constant":
        return torch.full((n,), 1.0 / n)
    if kind == "ties":
        return torch.full((n,), 0.75)
    if kind == "pow2":
        return torch.pow(2.0, -torch.randint(0, 30, (n,), generator=g).float())
    if kind == "zeros_mixed":
        r = torch.rand(n, generator=g)
        w = torch.where(r < 0.2, r * 1e-42, torch.where(r < 0.6, r * 1e-30, r))
        w[: n // 8] = 0
        return w
    raise KeyError(kind)


@pytest.mark.parametrize("kind", ["uniform", "lognormal_raw", "ce_like", "constant", "ties", "pow2", "zeros_mixed"])
@pytest.mark.parametrize("n", [1, 7, 2048, 2049, 50000, 1000000])
def test_exact_cdf_is_bit_identical_to_serial_fp32_chain(ops, kind, n):
    w = _weights(kind, n)
    cdf, total = ops.cdf(w.to(DEV), "exact")
    chain = orc.serial_chain_f32(w.numpy())
    ref = (chain / chain[-1]).astype(np.float32)
    assert total.cpu().numpy().view(np.uint32)[0] == chain[-1:].view(np.uint32)[0]
    got = cdf.cpu().numpy()
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), \
        f"{(got.view(np.uint32) != ref.view(np.uint32)).sum()} of {n} cdf entries differ"
    st, cchain, ccdf = c_oracle.serial_cdf(w.numpy())
    assert st == 0 and np.array_equal(ccdf.view(np.uint32), got.view(np.uint32))


def test_f64_cdf_is_close_and_deterministic(ops):
    w = _weights("uniform", 300000)
    a, ta = ops.cdf(w.to(DEV), "f64")
    b, tb = ops.cdf(w.to(DEV), "f64")
    assert torch.equal(a, b) and torch.equal(ta, tb)
    ref = np.cumsum(w.numpy().astype(np.float64))
    np.testing.assert_allclose(a.cpu().numpy(), ref / ref[-1], rtol=2e-7, atol=1e-7)


@pytest.mark.parametrize("n,k", [(17, 8), (320, 32), (5000, 128), (50000, 128)])
def test_draws_match_aten_golden(ops, golden_dir, n, k):
    g = np.load(os.path.join(golden_dir, "aten_multinomial.npz"))
    p, seed = torch.from_numpy(g[f"n{n}_p"]), int(g[f"n{n}_seed"])
    got = ops.multinomial(p.to(DEV), len(g[f"n{n}_with"]), True, generator=torch.Generator().manual_seed(seed))
    assert np.array_equal(got.cpu().numpy(), g[f"n{n}_with"])
    got = ops.multinomial(p.to(DEV), len(g[f"n{n}_without"]), False, generator=torch.Generator().manual_seed(seed))
    assert np.array_equal(got.cpu().numpy(), g[f"n{n}_without"])


@pytest.mark.parametrize("n,k", [(1000, 64), (1000000, 256), (1 << 22, 128)])
def test_multinomial_is_bit_exact_vs_torch_cpu(ops, n, k):
    w = _weights("ce_like", n, seed=3)
    g1, g2 = torch.Generator().manual_seed(21), torch.Generator().manual_seed(21)
    ref = torch.multinomial(w, k, True, generator=g1)
    got = ops.multinomial(w.to(DEV), k, True, generator=g2)
    assert torch.equal(got.cpu(), ref)
    assert torch.equal(g1.get_state(), g2.get_state())
    g1, g2 = torch.Generator().manual_seed(22), torch.Generator().manual_seed(22)
    ref = torch.multinomial(w, k, False, generator=g1)
    got = ops.multinomial(w.to(DEV), k, False, generator=g2)
    assert torch.equal(got.cpu(), ref)
    # n_sample == 1 takes the race path whatever `replacement` says (util.py:150)
    g1, g2 = torch.Generator().manual_seed(23), torch.Generator().manual_seed(23)
    assert torch.equal(ops.multinomial(w.to(DEV), 1, True, generator=g2).cpu(), torch.multinomial(w, 1, True, generator=g1))


def test_search_edge_uniforms(ops):
    w = _weights("uniform", 10000)
    cdf, _ = ops.cdf(w.to(DEV), "exact")
    c = cdf.cpu().numpy()
    u = np.array([0.0, 1e-300, float(c[0]), np.nextafter(float(c[0]), 1), float(c[4999]), float(c[-2]),
                  np.nextafter(1.0, 0), 0.5, 0.25], dtype=np.float64)
    got = ops.sample_with_replacement(cdf, torch.from_numpy(u).to(DEV)).cpu().numpy()
    assert np.array_equal(got, orc.lower_bound(c, u))
    assert np.array_equal(got, c_oracle.lower_bound(c, u))


def test_distribution_errors_match_torch(ops):
    for bad, msg in [([0.5, -0.1, 0.2], "< 0"), ([0.5, float("inf"), 0.1], "infinity or NaN"), ([0.0, 0.0, 0.0], "<= 0"),
                     ([0.5, float("nan"), 0.2], "< 0")]:
        t = torch.tensor(bad)
        with pytest.raises(RuntimeError) as a:
            torch.multinomial(t, 2, True)
        with pytest.raises(RuntimeError) as b:
            ops.multinomial(t.to(DEV), 2, True)
        assert msg in str(a.value) and msg in str(b.value)
    with pytest.raises(RuntimeError, match="2\\^24"):
        ops.multinomial(torch.ones((1 << 24) + 1, device=DEV), 2, True)
    with pytest.raises(RuntimeError, match="without replacement"):
        ops.multinomial(torch.ones(4, device=DEV), 5, False)


def test_chi_squared_of_empirical_draws(ops):
    from scipy import stats as sps

    n, draws, bins = 4096, 1 << 20, 64
    w = _weights("ce_like", n, seed=9)
    u = torch.rand(draws, dtype=torch.float64, generator=torch.Generator().manual_seed(5))
    cdf, _ = ops.cdf(w.to(DEV), "exact")
    idx = ops.sample_with_replacement(cdf, u.to(DEV)).cpu().numpy()
    obs = np.bincount(idx // (n // bins), minlength=bins)
    exp = w.double().numpy().reshape(bins, -1).sum(1) * draws
    chi2 = ((obs - exp) ** 2 / exp).sum()
    assert sps.chi2.sf(chi2, bins - 1) > 1e-4, f"chi2={chi2}"


# ---------------------------------------------------------------------------------- stats + K3
def test_pool_stats_and_normalize(ops):
    x, y = _logits(70000, 10, torch.float32)
    losses = orc.score(x, y)
    ema_ref = orc.EMAverage()
    ema_dev = torch.zeros(1, device=DEV)
    first = True
    for _ in range(3):  # EMA carried across calls
        mean, w, p = orc.pool_stats(losses, ema_ref)
        st = ops.pool_stats(losses.to(DEV), ema_dev, first)
        first = False
        s = st.cpu().numpy()
        np.testing.assert_allclose(s[0], float(mean), rtol=1e-6)
        np.testing.assert_allclose(s[1], float(ema_ref.value), rtol=1e-6)
        np.testing.assert_allclose(s[2], 0.5 * float(ema_ref.value), rtol=1e-6)
        np.testing.assert_allclose(s[3], float(w.sum()), rtol=1e-6)
        np.testing.assert_allclose(float(ema_dev), float(ema_ref.value), rtol=1e-6)
        np.testing.assert_allclose(ops.normalize(losses.to(DEV), st).cpu().numpy(), p.numpy(), rtol=RTOL_F32)
        losses = losses * 0.9 + 0.01


def test_gather_reweight(ops):
    n, b = 5000, 77
    g = torch.Generator().manual_seed(4)
    p = torch.rand(n, generator=g)
    p = p / p.sum()
    idx = torch.randint(0, n, (b,), generator=g)
    data = torch.rand(n, 3, 32, 32, generator=g)
    odd = torch.rand(n, 5, generator=g).to(torch.bfloat16)  # 10-byte rows: unaligned copy path
    labels = torch.randint(0, 10, (n,), generator=g)
    index = torch.arange(n) * 3
    nprob, seed, d, l, ix = ops.gather_reweight(p.to(DEV), idx.to(DEV), data.to(DEV), labels.to(DEV), index.to(DEV))
    ref_nprob, rd, rl, ri = orc.gather_reweight(p, idx, data, labels, index)
    assert torch.equal(d.cpu(), rd) and torch.equal(l.cpu(), rl) and torch.equal(ix.cpu(), ri)
    np.testing.assert_allclose(nprob.cpu().numpy(), ref_nprob.numpy(), rtol=1e-6)
    losses = torch.rand(b, generator=g) * 3
    _, ref_seed = orc.reweighted_loss(losses, ref_nprob)
    np.testing.assert_allclose(seed.cpu().numpy(), ref_seed.numpy(), rtol=RTOL_F32)
    cn, cs = c_oracle.gather_weights(p.numpy(), idx.numpy())
    np.testing.assert_allclose(nprob.cpu().numpy(), cn, rtol=1e-6)
    np.testing.assert_allclose(seed.cpu().numpy(), cs, rtol=1e-6)
    _, _, d2, _, _ = ops.gather_reweight(p.to(DEV), idx.to(DEV), odd.to(DEV))
    assert torch.equal(d2.cpu(), odd[idx])


# ------------------------------------------------------------------------------ the whole step
@pytest.mark.parametrize("n,c,dtype,b", [(320, 10, torch.float32, 32), (50000, 10, torch.float32, 128),
                                         (1000000, 10, torch.float32, 128), (20000, 1000, torch.bfloat16, 256),
                                         (3000001, 2, torch.float32, 64)])  # multi-million pool: the chain drifts, guesses get contradicted
def test_update_samples_end_to_end(ops, n, c, dtype, b):
    x, y = _logits(n, c, dtype)
    g = torch.Generator().manual_seed(3)
    u = torch.rand(b, dtype=torch.float64, generator=g)
    data = torch.rand(n, 3, 4, 4, generator=g)
    index = torch.randperm(n, generator=g)
    ema = torch.zeros(1, device=DEV)
    r = ops.update_samples(x.to(DEV), y.to(DEV), u.to(DEV), ema, True, data=data.to(DEV), index=index.to(DEV))
    ops.check_status(r.status)
    ref = orc.update_samples_flat(x, y, b, orc.EMAverage(), u=u.numpy(), data=data, index=index)
    np.testing.assert_allclose(r.scores.cpu().numpy(), ref["scores"].numpy(), rtol=RTOL_F32, atol=ATOL_LOSS)
    np.testing.assert_allclose(r.stats.cpu().numpy()[0], float(ref["mean"]), rtol=1e-6)
    p_dev = ops.normalize(r.scores, r.stats).cpu()
    np.testing.assert_allclose(p_dev.numpy(), ref["p"].numpy(), rtol=RTOL_F32)
    # bit-exact indices for the same importance vector bits and the same uniforms
    want = orc.draw_with_replacement(p_dev.numpy(), u.numpy())
    assert np.array_equal(r.idx.cpu().numpy(), want)
    # ... which is what torch.multinomial itself returns for those bits and that generator
    tm = torch.multinomial(p_dev, b, True, generator=torch.Generator().manual_seed(3))
    assert np.array_equal(want, tm.numpy())
    assert torch.equal(r.data.cpu(), data[want]) and torch.equal(r.labels.cpu(), y[want]) and torch.equal(r.index.cpu(), index[want])
    np.testing.assert_allclose(r.nprob.cpu().numpy(), (p_dev[want] * n).numpy(), rtol=1e-6)
    np.testing.assert_allclose(r.seed.cpu().numpy(), (1.0 / b) / (p_dev[want] * n).numpy(), rtol=1e-6)
    # against the all-CPU result the draws agree except where a last-ulp difference of p moves a boundary
    assert (r.idx.cpu().numpy() == ref["idx"].numpy()).mean() > 0.5


def test_update_samples_reproduces_the_reference_run(ops, golden_dir):
    """The reference's own Trainer.update_samples output (ResNet18 logits, 10 growing pools)."""
    g = np.load(os.path.join(golden_dir, "ref_update_samples_resnet18.npz"))
    B = int(g["draw_batch"])
    torch.manual_seed(int(g["rng_seed"]))
    u_all = torch.rand(20 * B, dtype=torch.float64)
    labels = torch.from_numpy(g["pool_labels"]).to(DEV)
    index = torch.from_numpy(g["pool_index"]).to(DEV)
    ema = torch.zeros(1, device=DEV)
    bounds = [32 * k for k in range(1, 11)]
    for call in range(2):
        logits = torch.from_numpy(g[f"call{call}_logits"]).to(DEV)
        u = u_all[(call * 10 + 9) * B:(call * 10 + 10) * B].to(DEV)
        r = ops.update_samples(logits, labels, u, ema, call == 0, index=index, bounds=bounds)
        assert np.array_equal(r.index.cpu().numpy(), g[f"call{call}_index"])
        assert np.array_equal(r.labels.cpu().numpy(), g[f"call{call}_label"])
        np.testing.assert_allclose(r.nprob.cpu().numpy(), g[f"call{call}_probs"], rtol=RTOL_F32)
        np.testing.assert_allclose(r.stats.cpu().numpy()[0], g[f"call{call}_avg"], rtol=RTOL_F32)
        np.testing.assert_allclose(float(ema), g[f"call{call}_ema"], rtol=RTOL_F32)


def test_update_samples_from_precomputed_scores(ops):
    x, y = _logits(30000, 10, torch.float32)
    u = torch.rand(64, dtype=torch.float64, generator=torch.Generator().manual_seed(8)).to(DEV)
    a = ops.update_samples(x.to(DEV), y.to(DEV), u, torch.zeros(1, device=DEV), True)
    b = ops.update_samples(None, y.to(DEV), u, torch.zeros(1, device=DEV), True, scores_in=a.scores)
    assert torch.equal(a.idx, b.idx) and torch.equal(a.nprob, b.nprob) and torch.equal(a.labels, b.labels)


@pytest.mark.parametrize("bias", [0.0, 0.5, 0.997, 1.002, 4.0])
@pytest.mark.parametrize("kind,n", [("uniform", 1000000), ("lognormal_raw", 300000), ("constant", 1 << 21), ("zeros_mixed", 50000)])
def test_exact_cdf_is_exact_whatever_the_phase_guesses(ops, kind, n, bias):
    """Guesses only select which maps are precomputed: forcing them wrong must not change a bit."""
    from mercury_b200 import _lib

    w = _weights(kind, n, seed=11)
    chain = orc.serial_chain_f32(w.numpy())
    ref = (chain / chain[-1]).astype(np.float32)
    lib = _lib.load()
    lib.mercury_test_set_guess_bias(bias)
    try:
        cdf, total = ops.cdf(w.to(DEV), "exact")
        got = cdf.cpu().numpy()
    finally:
        lib.mercury_test_set_guess_bias(1.0)
    assert np.array_equal(got.view(np.uint32), ref.view(np.uint32))
    assert total.cpu().numpy().view(np.uint32)[0] == chain[-1:].view(np.uint32)[0]


# ------------------------------------------------------------------------------ maximum sizes
@pytest.mark.parametrize("kind", ["ce_like", "uniform", "lognormal_raw"])
def test_largest_pool_2_pow_24_is_bit_exact(ops, kind):
    """N = 2^24, the most categories torch.multinomial accepts (and BASELINE cfg5's pool size): the chain total,
    draws from the materialised CDF and the whole-step draw must equal the serial fp32 chain."""
    n, b = 1 << 24, 512
    w = _weights(kind, n, seed=9)
    chain = orc.serial_chain_f32(w.numpy())
    ref_cdf = (chain / chain[-1]).astype(np.float32)
    u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(4))
    u[0], u[1] = 0.0, 1.0 - 2.0 ** -53
    want = orc.lower_bound(ref_cdf, u.numpy())
    cdf, total = ops.cdf(w.to(DEV), "exact")
    assert total.cpu().numpy().view(np.uint32)[0] == chain[-1:].view(np.uint32)[0]
    assert np.array_equal(cdf.cpu().numpy().view(np.uint32), ref_cdf.view(np.uint32))
    got = ops.sample_with_replacement(cdf, u.to(DEV))
    assert np.array_equal(got.cpu().numpy(), want)
    # the whole-step entry never materialises the CDF: draw straight from the weights as precomputed scores
    # (smooth_alpha = 0: p = w / fp64-sum(w), then the chain over p)
    r = ops.update_samples(None, None, u.to(DEV), torch.zeros(1, device=DEV), True, smooth_alpha=0.0, scores_in=w.to(DEV))
    ops.check_status(r.status)
    p = (w.numpy() / np.float32(r.stats[3].item())).astype(np.float32)
    chain_p = orc.serial_chain_f32(p)
    want_p = orc.lower_bound((chain_p / chain_p[-1]).astype(np.float32), u.numpy())
    assert np.array_equal(r.idx.cpu().numpy(), want_p)


def test_score_rows_beyond_2_pow_31_elements(ops):
    """70 000 x 32 768 bf16 = 2.3e9 logits (4.6 GB): element offsets pass 2^31; rows on both sides are checked."""
    n, c = 70000, 32768
    g = torch.Generator(device=DEV).manual_seed(12)
    x = torch.empty(n, c, dtype=torch.bfloat16, device=DEV)
    for i in range(0, n, 4096):
        x[i:i + 4096] = (torch.randn(min(4096, n - i), c, generator=g, device=DEV) * 2).to(torch.bfloat16)
    y = torch.randint(0, c, (n,), generator=g, device=DEV)
    got = ops.score(x, y).cpu()
    for lo in (0, 65536 - 300, n - 600):   # 65536 rows * 32768 = 2^31 elements
        sl = slice(lo, lo + 600)
        ref = orc.score(x[sl].cpu(), y[sl].cpu())
        np.testing.assert_allclose(got[sl].numpy(), ref.numpy(), rtol=RTOL_F32, atol=ATOL_LOSS)
    del x
    torch.cuda.empty_cache()


@pytest.mark.parametrize("n", [100, 6250, 50000, 1000000, 3000001])
def test_whole_step_draws_cover_every_chunk(ops, n):
    """Many draws (incl. u = 0 and u -> 1) so that every kind of chunk is hit: the first chunks, where the running
    sum changes binade many times (sub-chunk boundaries left by the spine), and plain single-binade chunks."""
    b = 4096
    g = torch.Generator().manual_seed(n)
    scores = (-torch.log(torch.rand(n, generator=g) + 1e-9) * 0.7 + 0.01).float()
    u = torch.rand(b, dtype=torch.float64, generator=g)
    u[0], u[1], u[2] = 0.0, 1.0 - 2.0 ** -53, 1e-9
    u[3:1000] = u[3:1000] * (5000.0 / n if n > 5000 else 1.0)          # crowd the head of the pool
    u.clamp_(0.0, 1.0 - 2.0 ** -53)
    r = ops.update_samples(None, None, u.to(DEV), torch.zeros(1, device=DEV), True, scores_in=scores.to(DEV))
    ops.check_status(r.status)
    # (the pool total is an fp64 reduction on the device, ATen's is an fp32 cascade: last-bit differences of the total are
    # within the stated 1e-6; the draw is checked bit-exactly against the chain over the device's own probabilities)
    mean, w, p_ref = orc.pool_stats(scores, orc.EMAverage(), 0.5)
    np.testing.assert_allclose(float(r.stats[3]), float(w.sum()), rtol=1e-6)
    p = ((scores.numpy() + np.float32(r.stats[2].item())).astype(np.float32) / np.float32(r.stats[3].item())).astype(np.float32)
    assert np.array_equal(ops.normalize(scores.to(DEV), r.stats).cpu().numpy(), p)
    want = orc.draw_with_replacement(p, u.numpy())
    assert np.array_equal(r.idx.cpu().numpy(), want)
    np.testing.assert_array_equal(r.nprob.cpu().numpy(), p[want] * np.float32(n))


# ------------------------------------------------------------------ f4: classifier head folded into K1
@pytest.mark.parametrize("n,d,c,dtype,bias", [
    (1000, 512, 10, torch.float32, True),      # the reference's ResNet18 head
    (1003, 512, 10, torch.bfloat16, True),     # ragged tail (n % 4 != 0), bf16 features
    (1, 128, 2, torch.float32, False),
    (257, 2048, 16, torch.float32, True),      # Bottleneck head width, padded class count == c
    (300, 1024, 7, torch.bfloat16, False),     # c below its padded count
    (64, 256, 3, torch.float32, True),
])
def test_score_head_matches_linear_then_cross_entropy(ops, n, d, c, dtype, bias):
    g = torch.Generator().manual_seed(n + d + c)
    x = torch.relu(torch.randn(n, d, generator=g)).to(dtype)          # post-ReLU, pooled features
    w = torch.randn(c, d, generator=g) * (2.0 / d ** 0.5)
    b = torch.randn(c, generator=g) * 0.1 if bias else None
    y = torch.randint(0, c, (n,), generator=g)
    if n > 3:
        y[2] = -100
    got = ops.score_head(x.to(DEV), w.to(DEV), None if b is None else b.to(DEV), y.to(DEV)).cpu()
    ref = orc.score_head(x, w, b, y)
    # fp32 accumulation in an order that is not the BLAS's: 2e-5 (stated in include/mercury_b200.h)
    np.testing.assert_allclose(got.numpy(), ref.numpy(), rtol=2e-5, atol=2e-5)
    ref64 = F.cross_entropy(F.linear(x.double(), w.double(), None if b is None else b.double()), y, reduction="none")
    np.testing.assert_allclose(got.numpy(), ref64.numpy(), rtol=2e-5, atol=2e-5)
    if n > 3:
        assert got[2] == 0.0
    # same scores as K1 on the materialised logits (to the same tolerance), same bad-label behaviour
    logits = F.linear(x.float(), w, b)
    np.testing.assert_allclose(got.numpy(), ops.score(logits.to(DEV), y.to(DEV)).cpu().numpy(), rtol=2e-5, atol=2e-5)
    y_bad = y.clone()
    y_bad[0] = c
    st = ops.new_status(DEV)
    ops.score_head(x.to(DEV), w.to(DEV), None, y_bad.to(DEV), status=st)
    with pytest.raises(IndexError):
        ops.check_status(st)


def test_score_head_on_the_reference_resnet18_head(ops):
    """The reference's own model: ResNet18(10).linear applied to its pooled features (pytorch_model.py:94-96)."""
    import importlib.util
    import sys
    ref_model = os.path.join("/root/reference", "pytorch_model.py")
    torch.manual_seed(0)
    if os.path.exists(ref_model):       # (only in the build container; the GPU box has no /root/reference)
        spec = importlib.util.spec_from_file_location("ref_pytorch_model", ref_model)
        mod = importlib.util.module_from_spec(spec)
        sys.modules["ref_pytorch_model"] = mod
        spec.loader.exec_module(mod)
        lin = mod.ResNet18(10).linear
    else:
        lin = torch.nn.Linear(512, 10)  # the same layer: nn.Linear(512 * expansion, num_classes)
    g = torch.Generator().manual_seed(1)
    feats = torch.relu(torch.randn(5000, 512, generator=g))
    y = torch.randint(0, 10, (5000,), generator=g)
    with torch.no_grad():
        ref = F.cross_entropy(lin(feats), y, reduction="none")
    got = ops.score_head(feats.to(DEV), lin.weight.detach().to(DEV), lin.bias.detach().to(DEV), y.to(DEV)).cpu()
    np.testing.assert_allclose(got.numpy(), ref.numpy(), rtol=2e-5, atol=2e-5)


def test_score_head_matches_the_reference_run(ops, golden_dir):
    """Golden vector recorded from the reference's own ResNet18 forward (tests/golden/make_golden.py, ref_head)."""
    z = np.load(os.path.join(golden_dir, "ref_head.npz"))
    f, w, b, y = (torch.from_numpy(z[k]).to(DEV) for k in ("features", "weight", "bias", "label"))
    got = ops.score_head(f, w, b, y).cpu().numpy()
    np.testing.assert_allclose(got, z["scores"], rtol=2e-5, atol=2e-5)
    np.testing.assert_allclose(ops.score(torch.from_numpy(z["logits"]).to(DEV), y).cpu().numpy(), z["scores"], rtol=RTOL_F32, atol=ATOL_LOSS)
