"""Randomised check of the one-kernel step (and the chain) against the serial fp32 chain on the device's own probabilities:
random pool sizes, class counts, draw counts, score distributions (heavy tails, exact ties, zeros, constant pools), EMA on/off,
repeated launches on one workspace.  python tools/fuzz_step.py [iterations] [seed]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from mercury_b200 import ops  # noqa: E402
from oracle import mercury_oracle as orc  # noqa: E402

dev = torch.device("cuda", 0)
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0
for it in range(iters):
    n = int(rng.choice([rng.integers(1, 5000), rng.integers(5000, 200_000), rng.integers(200_000, 1_570_000)]))
    c = int(rng.choice([2, 3, 10, 10, 10, 17, 32, 64]))
    b = int(rng.choice([1, 7, 128, 128, 256, 1000]))
    dt = torch.float32 if rng.random() < 0.7 or c % 8 else torch.bfloat16
    kind = int(rng.integers(0, 5))
    g = torch.Generator(device=dev).manual_seed(int(rng.integers(1 << 30)))
    x = torch.randn(n, c, generator=g, device=dev) * float(rng.choice([0.5, 2.0, 6.0]))
    if kind == 1:      # exact ties: few distinct rows
        x = x[torch.randint(0, min(4, n), (n,), generator=g, device=dev)]
    elif kind == 2:    # constant pool
        x = x[:1].expand(n, c).contiguous()
    elif kind == 3:    # many (near) zero losses
        y0 = torch.randint(0, c, (n,), generator=g, device=dev)
        x[torch.arange(n, device=dev), y0] += 40.0
    x = x.to(dt)
    y = torch.randint(0, c, (n,), generator=g, device=dev) if kind != 3 else y0
    if kind == 3:      # ... but not all of them
        flip = torch.rand(n, generator=g, device=dev) < 0.1
        y = torch.where(flip, (y + 1) % c, y)
    u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(int(rng.integers(1 << 30)))).to(dev)
    if b > 2:
        u[0], u[1] = 0.0, 1.0 - 2.0 ** -53
    for mode in (1, 0):
        ops.set_fused(mode)
        ema = torch.full((1,), float(rng.random() * 3), device=dev)
        first = bool(rng.random() < 0.5)
        for rep in range(2):
            r = ops.update_samples(x, y, u, ema, first and rep == 0)
            st = int(r.status.item())
            p = ops.normalize(r.scores, r.stats).cpu().numpy()
            want = orc.draw_with_replacement(p, u.cpu().numpy()) if st == 0 else None
            got = r.idx.cpu().numpy()
            ok = st != 0 or np.array_equal(got, want)
            # (a pool whose losses and EMA are all zero has total 0: torch.multinomial raises there too; 8 = zero sum, and the
            # 0/0 probabilities then also read as "not >= 0")
            if not ok or (st and not (st & 8)):
                bad += 1
                nb = int((got != want).sum()) if want is not None else -1
                print(f"MISMATCH it={it} mode={mode} rep={rep} n={n} c={c} b={b} dt={dt} kind={kind} status={st} differing={nb}", flush=True)
    if it % 4 == 0 and n >= 10:   # growing pools (pytorch_collab.py:95-112: one EMA update per presample batch) and scores handed in
        ops.set_fused(-1)
        nb = sorted(set(int(v) for v in rng.integers(1, n, size=int(rng.integers(1, 9)))) | {n})
        e0 = float(rng.random() * 3)
        ema = torch.full((1,), e0, device=dev)
        first = bool(rng.random() < 0.5)
        r = ops.update_samples(x, y, u, ema, first, bounds=nb)
        sc = r.scores.cpu()
        o = orc.EMAverage()
        o.first_update, o.value = first, torch.tensor(e0)
        for v in nb:
            mean, wsum, _ = orc.pool_stats(sc[:v], o, 0.5)
        st = r.stats.cpu().numpy()
        p = ops.normalize(r.scores, r.stats).cpu().numpy()
        ok = int(r.status.item()) == 0 and np.allclose(st[1], float(o.value), rtol=2e-6) and np.allclose(st[0], float(mean), rtol=2e-6) \
            and np.allclose(float(ema), float(o.value), rtol=2e-6) and np.array_equal(r.idx.cpu().numpy(), orc.draw_with_replacement(p, u.cpu().numpy()))
        r2 = ops.update_samples(None, y, u, torch.full((1,), e0, device=dev), first, scores_in=r.scores, bounds=nb)
        ok = ok and torch.equal(r2.idx, r.idx) and torch.equal(r2.nprob, r.nprob)
        if not ok:
            bad += 1
            print(f"MISMATCH (growing pools) it={it} n={n} c={c} b={b} bounds={nb} status={int(r.status.item())} ema={st[1]} want={float(o.value)}", flush=True)
    if it % 20 == 19:
        print(f"{it + 1} iterations, {bad} bad", flush=True)
ops.set_fused(-1)
print("done:", iters, "iterations,", bad, "bad")
sys.exit(1 if bad else 0)
