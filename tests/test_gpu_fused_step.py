"""The one-kernel step (csrc/step_fused.cuh) against the four-kernel chain and against the oracle.

Both implement the arithmetic of Trainer.update_samples (reference pytorch_collab.py:102-117); for the same inputs they
must agree to the bit in every output (scores, statistics, indices, weights, seeds, gathered rows), and the indices must
equal the serial fp32 chain over the device's own probabilities (bit-exact, BASELINE.json north_star).
"""
import numpy as np
import pytest
import torch

from oracle import mercury_oracle as orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mercury_b200 import _lib
    from mercury_b200 import ops as _ops

    yield _ops
    _ops.set_fused(-1)
    _lib.load().mercury_test_set_guess_bias(1.0)


def _pool(n, c, dtype, seed, sigma=2.0, kind="randn"):
    g = torch.Generator().manual_seed(seed)
    if kind == "randn":
        x = (torch.randn(n, c, generator=g) * sigma).to(dtype)
    elif kind == "confident":   # most rows score ~0: long runs of tiny addends, exact ties, zero sub-chunks
        x = (torch.randn(n, c, generator=g) * 0.01).to(dtype)
        x[torch.arange(n), torch.randint(0, c, (n,), generator=g)] = 30.0
    else:                       # "flat": every row scores log(c): constant addends, ties everywhere
        x = torch.zeros(n, c, dtype=dtype)
    y = torch.randint(0, c, (n,), generator=g)
    if kind == "confident":
        y = x.float().argmax(1)
        y[::7] = (y[::7] + 1) % c   # some confidently wrong rows carry nearly all the mass
    return x, y


def _both(ops, x, y, u, ema0, first, **kw):
    res = []
    for mode in (0, 1):
        ops.set_fused(mode)
        ema = ema0.clone()
        r = ops.update_samples(x, y, u, ema, first, **kw)
        ops.check_status(r.status)
        torch.cuda.synchronize()
        res.append((r, ema))
    ops.set_fused(-1)
    return res


def _bits(t):
    return t.view(torch.uint8) if t.dtype.is_floating_point else t


def _assert_same(a, b):
    """Scores are the same kernel code: bit-equal.  The pool statistics are fp64 sums taken in different orders (per-tile
    sums in the chain, per-CTA sums in the one-kernel step): equal to fp32 rounding, 1e-6 relative at most."""
    (ra, ea), (rb, eb) = a, b
    assert torch.equal(_bits(ra.scores), _bits(rb.scores)), "scores differ between chain and one-kernel step"
    np.testing.assert_allclose(ra.stats.cpu().numpy()[:4], rb.stats.cpu().numpy()[:4], rtol=1e-6)
    np.testing.assert_allclose(ea.cpu().numpy(), eb.cpu().numpy(), rtol=1e-6)


def _assert_exact(ops, r, u, data, index, y):
    """Indices bit-exact for the device's own importance vector; gathered rows and weights belong to those indices."""
    p = ops.normalize(r.scores, r.stats).cpu().numpy()
    want = orc.draw_with_replacement(p, u.numpy())
    assert np.array_equal(r.idx.cpu().numpy(), want), "indices are not those of the serial fp32 chain"
    if data is not None:
        assert torch.equal(r.data.cpu(), data[want]) and torch.equal(r.index.cpu(), index[want]) and torch.equal(r.labels.cpu(), y[want])
    np.testing.assert_array_equal(r.nprob.cpu().numpy(), p[want] * np.float32(len(p)))
    chain = orc.serial_chain_f32(p)
    assert r.stats.cpu().numpy().view(np.uint32)[4] == chain[-1:].view(np.uint32)[0], "chain total"


@pytest.mark.parametrize("n,c,dtype,b,kind", [
    (1, 10, torch.float32, 4, "randn"), (33, 10, torch.float32, 16, "randn"), (320, 10, torch.float32, 32, "randn"),
    (2048, 10, torch.float32, 64, "randn"), (2049, 10, torch.float32, 64, "randn"), (6250, 10, torch.float32, 128, "randn"),
    (50000, 10, torch.float32, 128, "randn"), (1000000, 10, torch.float32, 128, "randn"),
    (1000000, 10, torch.float32, 1024, "confident"), (300000, 10, torch.float32, 256, "flat"),
    (2000000, 3, torch.float32, 256, "randn"),            # run-time class count, 12-byte rows (ragged bulk copies)
    (700001, 10, torch.bfloat16, 200, "randn"), (123457, 16, torch.bfloat16, 77, "confident"),
    (2097152, 8, torch.bfloat16, 512, "randn"),           # 1024 chunks: the largest pool the one-kernel step takes
])
def test_one_kernel_step_equals_chain_and_oracle(ops, n, c, dtype, b, kind):
    x, y = _pool(n, c, dtype, seed=n % 1000 + c, kind=kind)
    g = torch.Generator().manual_seed(3)
    u = torch.rand(b, dtype=torch.float64, generator=g)
    u[0] = 0.0
    u[-1] = 1.0 - 2.0 ** -53
    data = torch.rand(n, 6, generator=g)
    index = torch.arange(n) * 2 + 1
    xd, yd, ud = x.to(DEV), y.to(DEV), u.to(DEV)
    chain, fused = _both(ops, xd, yd, ud, torch.zeros(1, device=DEV), True, data=data.to(DEV), index=index.to(DEV))
    _assert_same(chain, fused)
    for r, _ in (chain, fused):
        _assert_exact(ops, r, u, data, index, y)
    ref = orc.score(x, y)
    np.testing.assert_allclose(fused[0].scores.cpu().numpy(), ref.numpy(), rtol=1e-5, atol=1e-5)


def test_one_kernel_step_carries_the_ema_and_reuses_its_workspace(ops):
    """Several steps in a row on one workspace (stale summaries, packs and draw words from the step before), EMA carried."""
    n, b = 200000, 96
    ema_c, ema_f = torch.zeros(1, device=DEV), torch.zeros(1, device=DEV)
    ema_ref = orc.EMAverage()
    for it in range(4):
        x, y = _pool(n - it * 777, 10, torch.float32, seed=40 + it)
        u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(it)).to(DEV)
        ops.set_fused(0)
        rc = ops.update_samples(x.to(DEV), y.to(DEV), u, ema_c, it == 0)
        ops.set_fused(1)
        rf = ops.update_samples(x.to(DEV), y.to(DEV), u, ema_f, it == 0)
        ops.set_fused(-1)
        ops.check_status(rf.status)
        np.testing.assert_allclose(rc.stats.cpu().numpy()[:4], rf.stats.cpu().numpy()[:4], rtol=1e-6)
        np.testing.assert_allclose(float(ema_c), float(ema_f), rtol=1e-6)
        _assert_exact(ops, rf, u.cpu(), None, None, None)
        orc.pool_stats(orc.score(x, y), ema_ref)
        np.testing.assert_allclose(float(ema_f), float(ema_ref.value), rtol=1e-6)


@pytest.mark.parametrize("bias", [0.0, 0.5, 0.997, 1.002, 4.0])
@pytest.mark.parametrize("n,kind", [(1000000, "randn"), (300000, "confident"), (50000, "flat")])
def test_one_kernel_step_is_exact_whatever_the_phase_guesses(ops, n, kind, bias):
    """Guesses only choose which maps are precomputed: forcing them wrong must not change a bit."""
    from mercury_b200 import _lib

    x, y = _pool(n, 10, torch.float32, seed=5, kind=kind)
    u = torch.rand(64, dtype=torch.float64, generator=torch.Generator().manual_seed(9))
    lib = _lib.load()
    ops.set_fused(1)
    lib.mercury_test_set_guess_bias(bias)
    try:
        r = ops.update_samples(x.to(DEV), y.to(DEV), u.to(DEV), torch.zeros(1, device=DEV), True)
        ops.check_status(r.status)
    finally:
        lib.mercury_test_set_guess_bias(1.0)
        ops.set_fused(-1)
    p = ops.normalize(r.scores, r.stats).cpu().numpy()
    assert np.array_equal(r.idx.cpu().numpy(), orc.draw_with_replacement(p, u.numpy()))
    chain = orc.serial_chain_f32(p)
    assert r.stats.cpu().numpy().view(np.uint32)[4] == chain[-1:].view(np.uint32)[0], "chain total"


def test_one_kernel_step_in_a_cuda_graph(ops):
    """The cooperative launch is capturable: replaying the graph gives the same draws as the eager call."""
    n, b = 400000, 128
    x, y = _pool(n, 10, torch.float32, seed=77)
    xd, yd = x.to(DEV), y.to(DEV)
    u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(1)).to(DEV)
    ops.set_fused(1)
    try:
        ema = torch.zeros(1, device=DEV)
        out = {}
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            r0 = ops.update_samples(xd, yd, u, ema, False, out=out)
            out.update({k: v for k, v in r0.items() if v is not None})
            torch.cuda.synchronize()
            want_idx, want_ema = r0.idx.clone(), ema.clone()
            ema.zero_()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                ops.update_samples(xd, yd, u, ema, False, out=out)
            out["idx"].zero_()
            ema.zero_()
            g.replay()
            torch.cuda.synchronize()
        assert torch.equal(out["idx"], want_idx) and torch.equal(ema, want_ema)
        ops.check_status(out["status"])
    finally:
        ops.set_fused(-1)


# ------------------------------------------------------------------------------------- the fp64 (non-exact) mode
@pytest.mark.parametrize("n,c,dtype,b", [(320, 10, torch.float32, 32), (50000, 10, torch.float32, 128), (1000000, 10, torch.float32, 256),
                                         (20000, 1000, torch.bfloat16, 256)])
def test_fp64_mode_matches_an_fp64_cdf_search(ops, n, c, dtype, b):
    """cdf_mode="f64": the indices are those of an fp64 running total searched as ATen searches.  The device adds in a
    tree, NumPy serially: the totals can differ in their last fp64 bits, which moves a draw only when its uniform sits
    within an fp32 rounding of a cdf value -- such a draw may land on the neighbouring candidate, nothing else may differ."""
    x, y = _pool(n, c, dtype, seed=21)
    g = torch.Generator().manual_seed(4)
    u = torch.rand(b, dtype=torch.float64, generator=g)
    u[0], u[-1] = 0.0, 1.0 - 2.0 ** -53
    data = torch.rand(n, 6, generator=g)
    r = ops.update_samples(x.to(DEV), y.to(DEV), u.to(DEV), torch.zeros(1, device=DEV), True, data=data.to(DEV), cdf_mode="f64")
    ops.check_status(r.status)
    p = ops.normalize(r.scores, r.stats).cpu().numpy()
    want = orc.draw_with_replacement_f64(p, u.numpy())
    got = r.idx.cpu().numpy()
    off = np.abs(got - want)
    assert off.max() <= 1 and (off == 0).mean() >= 0.99, (off.max(), (off == 0).mean())
    assert torch.equal(r.data.cpu(), data[got]) and torch.equal(r.labels.cpu(), y[got])
    np.testing.assert_array_equal(r.nprob.cpu().numpy(), p[got] * np.float32(n))
    # the same statistics as the exact mode (one code path), and mostly different indices for large pools
    e = ops.update_samples(x.to(DEV), y.to(DEV), u.to(DEV), torch.zeros(1, device=DEV), True, cdf_mode="exact")
    np.testing.assert_allclose(e.stats.cpu().numpy()[:4], r.stats.cpu().numpy()[:4], rtol=1e-6)
    np.testing.assert_allclose(float(r.stats[4]), 1.0, rtol=1e-5)


def test_fp64_mode_chi_squared(ops):
    from scipy import stats as sps

    n, draws, bins = 4096, 1 << 18, 64
    x, y = _pool(n, 10, torch.float32, seed=31)
    u = torch.rand(draws, dtype=torch.float64, generator=torch.Generator().manual_seed(5))
    r = ops.update_samples(x.to(DEV), y.to(DEV), u.to(DEV), torch.zeros(1, device=DEV), True, cdf_mode="f64")
    ops.check_status(r.status)
    p = ops.normalize(r.scores, r.stats).cpu().double().numpy()
    obs = np.bincount(r.idx.cpu().numpy() // (n // bins), minlength=bins)
    exp = p.reshape(bins, -1).sum(1) * draws
    chi2 = ((obs - exp) ** 2 / exp).sum()
    assert sps.chi2.sf(chi2, bins - 1) > 1e-4, f"chi2={chi2}"
