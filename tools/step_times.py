"""Phase timestamps of the one-kernel step (developer aid): [MERCURY_STEP_DBG=1] python tools/step_times.py [N] [B]
Prints, per pool, microseconds since the kernel's start at: CTA 0 scored | grid barrier passed | CTA 0 had the statistics |
CTA 0 published its summaries | last CTA arrived (walker starts) | walker loaded | walk done | draws published | CTA 0 done;
with MERCURY_STEP_DBG=1 also the cycles of each step of the walk."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from mercury_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
b = int(sys.argv[2]) if len(sys.argv) > 2 else 128
c = 10
g = torch.Generator(device=dev).manual_seed(100)
pools = [((torch.randn(n, c, generator=g, device=dev) * 2), torch.randint(0, c, (n,), generator=g, device=dev)) for _ in range(7)]
payload = torch.rand(n, 3, 32, 32, device=dev, generator=g)
index = torch.arange(n, device=dev)
u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(3)).to(dev)
ema = torch.zeros(1, device=dev)
out = {}
ops.set_fused(1)
order = [("scored", 1), ("all_scored", 11), ("barrier", 2), ("stats", 14), ("summarized", 15), ("head", 9), ("published", 3), ("summaries_seen", 4), ("packs_in", 10), ("loaded", 5),
         ("walked", 6), ("draws_out", 7), ("cta0_done", 8)]
for i in range(10):
    r = ops.update_samples(*pools[i % 7], u, ema, i == 0, data=payload, index=index, out=out)
    out.update({k: v for k, v in r.items() if v is not None})
    torch.cuda.synchronize()
    t = ops.step_times(n)
    if i >= 5:
        print("  ".join(f"{nm}={(t[k] - t[0]) / 1e3:6.2f}" for nm, k in order))
ops.check_status(r.status)
if int(os.environ.get("MERCURY_STEP_DBG", "0")) & 1:
    nev = t[12]
    ev = t[16:16 + nev + 1]
    print("walk steps (cycles : kind of the step, 0 add 1 map 3 pack 4 slow [general-loop scans / serial sub-chunks / tie maps]); head_k =", t[13])
    print(" ".join(f"{(ev[i] >> 16) - (ev[i - 1] >> 16)}:{ev[i] & 15}" + (f"[{(ev[i] >> 12) & 15}/{(ev[i] >> 8) & 15}/{(ev[i] >> 4) & 15}]" if ev[i] & 15 == 3 else "")
                   for i in range(1, len(ev))))
if int(os.environ.get("MERCURY_STEP_DBG", "0")) & 2:
    print("micro (cycles): 16 dependent SHFL loop iterations", t[16 + 104], "| 64 unrolled dependent FADD", t[16 + 105])
if int(os.environ.get("MERCURY_STEP_DBG", "0")) & 4:
    e = t[16 + 100:16 + 110]
    print("CTA 0, chunk 0 (us since start): p ready %.2f | thread maps %.2f | summary written %.2f | sub-chunk maps %.2f | pack written %.2f" % tuple((x - t[0]) / 1e3 for x in e[:5]))
    print("CTA 0 speculation (us since start): %.2f -> %.2f" % ((t[16 + 97] - t[0]) / 1e3, (t[16 + 98] - t[0]) / 1e3))
    print("speculative chunk preload of the draws: %d of %d hit (cumulative over the launches so far)" % (t[16 + 99] >> 32, t[16 + 99] & 0xffffffff))
    print("CTA 0 draw (us since start): word seen %.2f | chunk loaded %.2f | slot known %.2f | done %.2f" % tuple((x - t[0]) / 1e3 for x in (e[6], e[7], e[8], t[8])))
