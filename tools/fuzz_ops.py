"""Randomised checks of the stand-alone operators against the oracle: K1 over all its layouts (row sizes around every
dispatch threshold, strided and unaligned inputs, ignored and bad labels), the materialised CDF (exact / static) and its
search, the exponential-race top-k, the group-masked draw, the tcgen05 head.  python tools/fuzz_ops.py [iterations] [seed]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from mercury_b200 import _lib, ops  # noqa: E402
from oracle import mercury_oracle as orc  # noqa: E402

dev = torch.device("cuda", 0)
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 100
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 0)
bad = 0


def report(what, **kw):
    global bad
    bad += 1
    print("MISMATCH", what, kw, flush=True)


def gen(seed):
    return torch.Generator(device=dev).manual_seed(int(seed))


for it in range(iters):
    # ---- K1
    c = int(rng.choice([1, 2, 3, 10, 22, 23, 44, 46, 47, 64, 65, 100, 128, 256, 257, 1000, 1024, 2048, 2049, 4096, 4100, 8192, 32768, 33000]))
    dt = torch.bfloat16 if rng.random() < 0.5 else torch.float32
    n = int(rng.integers(1, max(2, min(200_000, 40_000_000 // c))))
    g = gen(rng.integers(1 << 30))
    pad = int(rng.choice([0, 0, 0, 8, 3]))                      # row stride > c (3: unaligned rows)
    full = (torch.randn(n, c + pad, generator=g, device=dev) * float(rng.choice([1.0, 4.0, 12.0]))).to(dt)
    x = full[:, :c]
    y = torch.randint(0, c, (n,), generator=g, device=dev)
    if rng.random() < 0.3:
        y[torch.randint(0, n, (max(1, n // 50),), generator=g, device=dev)] = -100
    st = ops.new_status(dev)
    got = ops.score(x, y, status=st).cpu()
    ref = orc.score(x.cpu(), y.cpu())
    if int(st.item()) != 0 or not np.allclose(got.numpy(), ref.numpy(), rtol=1e-5, atol=2e-5):
        report("score", n=n, c=c, dt=str(dt), pad=pad, status=int(st.item()), maxerr=float((got - ref).abs().max()))
    # ---- materialised CDF + search, exact and static
    m = int(rng.choice([rng.integers(1, 3000), rng.integers(3000, 300_000), rng.integers(300_000, 3_000_000)]))
    g = gen(rng.integers(1 << 30))
    w = torch.rand(m, generator=g, device=dev) ** float(rng.choice([1.0, 4.0, 0.25])) + float(rng.choice([0.0, 1e-3]))
    if rng.random() < 0.3:
        w[torch.randint(0, m, (max(1, m // 10),), generator=g, device=dev)] = 0.0
    if float(w.sum()) > 0:
        p = w / w.sum()
        chain = np.cumsum(p.cpu().numpy(), dtype=np.float32)
        want = (chain / chain[-1]).astype(np.float32)
        cdf, tot = ops.cdf(p, "exact")
        if not np.array_equal(cdf.cpu().numpy().view(np.uint32), want.view(np.uint32)):
            report("cdf exact", n=m)
        if m <= _lib.STATIC_MAX_N:
            lib = _lib.load()
            out, t2, s2 = torch.empty_like(p), torch.empty(1, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)
            rc = lib.mercury_cdf_static(p.data_ptr(), m, 0, out.data_ptr(), t2.data_ptr(), s2.data_ptr(), torch.cuda.current_stream().cuda_stream)
            if rc != 0 or not np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32)):
                report("cdf static", n=m, rc=rc)
        u = torch.rand(int(rng.choice([1, 33, 256])), dtype=torch.float64, generator=torch.Generator().manual_seed(int(rng.integers(1 << 30))))
        idx = ops.sample_with_replacement(cdf, u.to(dev)).cpu().numpy()
        if not np.array_equal(idx, orc.lower_bound(want, u.numpy())):
            report("search", n=m, b=u.numel())
        # ---- exponential race top-k
        k = int(min(m, rng.choice([1, 2, 64, 2048])))
        ur = torch.rand(m, dtype=torch.float64, generator=torch.Generator().manual_seed(int(rng.integers(1 << 30))))
        st = ops.new_status(dev)
        tk = ops.sample_topk(p, ur.to(dev), k, status=st).cpu().numpy()
        keys = orc.exp_race_keys(p.cpu(), ur)
        if int(st.item()) & ~_lib.ST_TOPK_TIES:
            report("topk status", n=m, k=k, status=int(st.item()))
        elif not (int(st.item()) & _lib.ST_TOPK_TIES):
            wantk = torch.topk(keys, k).values.numpy()     # (equal keys may come in either order: compare the key values)
            if not np.array_equal(keys.numpy()[tk], wantk) or len(set(tk.tolist())) != k:
                report("topk", n=m, k=k)
    # ---- group-masked draw
    ng = int(rng.integers(1, 6))
    gn = int(rng.integers(ng, 100_000))
    g = gen(rng.integers(1 << 30))
    imp = torch.rand(gn, generator=g, device=dev) * 3 + 1e-3
    tags = torch.randint(0, ng, (gn,), generator=g, device=dev, dtype=torch.int32)
    grp = int(rng.integers(0, ng))
    members = torch.nonzero(tags == grp).flatten()
    cnt = int(members.numel())
    if cnt > 0:
        nd = int(rng.choice([1, 5]))
        ug = torch.rand(nd * cnt, dtype=torch.float64, generator=torch.Generator().manual_seed(int(rng.integers(1 << 30))))
        got = ops.groupwise_draw(imp, tags, grp, cnt, ug.to(dev), ndraws=nd).cpu().numpy()
        sub = imp[members].cpu().numpy()
        for j in range(nd):
            wantj = int(members[orc.groupwise_draw(sub, ug[j * cnt:(j + 1) * cnt])])
            if int(got[j]) != wantj:
                pg = orc.groupwise_probabilities(sub)
                kk = orc.exp_race_keys(pg, ug[j * cnt:(j + 1) * cnt])
                loc = {int(v): i for i, v in enumerate(members.tolist())}
                if float(kk[loc[int(got[j])]]) != float(kk.max()):      # (an exact tie of the two best keys is not a mismatch)
                    report("groupwise", n=gn, cnt=cnt, j=j)
    # ---- tcgen05 head (a lighter share of the iterations: it allocates n x d features)
    if it % 3 == 0:
        hn, hd, hc = int(rng.integers(1, 20_000)), int(rng.choice([8, 64, 72, 256, 512, 2048])), int(rng.choice([1, 10, 17, 64, 65, 128, 129, 300, 1000, 1030]))
        g = gen(rng.integers(1 << 30))
        f = torch.relu(torch.randn(hn, hd, generator=g, device=dev)).to(torch.bfloat16)
        wt = (torch.randn(hc, hd, generator=g, device=dev) * 0.05).to(torch.bfloat16)
        bs = torch.randn(hc, generator=g, device=dev) * 0.1 if rng.random() < 0.7 else None
        yy = torch.randint(0, hc, (hn,), generator=g, device=dev)
        st = ops.new_status(dev)
        got = ops.score_head_wide(f, wt, bs, yy, status=st).cpu()
        ref = orc.score_head(f.cpu(), wt.cpu(), None if bs is None else bs.cpu(), yy.cpu())
        if int(st.item()) != 0 or not np.allclose(got.numpy(), ref.numpy(), rtol=2e-5, atol=2e-5):
            report("head_wide", n=hn, d=hd, c=hc, status=int(st.item()), maxerr=float((got - ref).abs().max()))
    if it % 20 == 19:
        print(f"{it + 1} iterations, {bad} bad", flush=True)
print("done:", iters, "iterations,", bad, "bad")
sys.exit(1 if bad else 0)
