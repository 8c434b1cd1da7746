"""Times mercury_score_head_wide (the classifier head on the tensor cores: TMA + tcgen05.mma + TMEM, cross-entropy in the
epilogue) against the tensor roofline, against cuBLAS bf16 `F.linear` alone (the GEMM it contains), and against the
unfused path it replaces: F.linear -> [N, C] logits in HBM -> mercury_score.
    python tools/bench_head_wide.py [N] [D] [C]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402

from mercury_b200 import ops  # noqa: E402

dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 262_144
d = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
c = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
g = torch.Generator(device=dev).manual_seed(0)
npools = 2
pools = [torch.relu(torch.randn(n, d, generator=g, device=dev)).to(torch.bfloat16) for _ in range(npools)]
w = (torch.randn(c, d, generator=g, device=dev) * 0.05).to(torch.bfloat16)
b = torch.randn(c, generator=g, device=dev) * 0.1
y = torch.randint(0, c, (n,), generator=g, device=dev)
out, st = torch.empty(n, device=dev), ops.new_status(dev)
logits = torch.empty(n, c, dtype=torch.bfloat16, device=dev)


def timed(fn, reps=10):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(reps):
        fn(i)
    z.record()
    torch.cuda.synchronize()
    return a.elapsed_time(z) / reps


def unfused(i):
    torch.addmm(b.to(torch.bfloat16), pools[i % npools], w.t(), out=logits)
    ops.score(logits, y, status=st, out=out)


ms = timed(lambda i: ops.score_head(pools[i % npools], w, b, y, status=st, out=out))
ms_gemm = timed(lambda i: torch.mm(pools[i % npools], w.t(), out=logits))
ms_unfused = timed(unfused)
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peak_tf, peak_sus = 1643.2, 1381.1
try:
    pk = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))
    peak_tf, peak_sus = float(pk["bf16_tflops"]), float(pk.get("bf16_tflops_sustained", pk["bf16_tflops"]))
except Exception:
    pass
flops = 2.0 * n * d * c
ops.check_status(st)
print(json.dumps({"kernel": "mercury_score_head_wide (tcgen05)", "n": n, "d": d, "c": c, "ms": ms, "candidates_per_s": n / (ms * 1e-3),
                  "flops": flops, "achieved_tflops": flops / (ms * 1e-3) / 1e12, "peak_tflops_burst": peak_tf, "peak_tflops_sustained": peak_sus,
                  "frac_of_burst": flops / (ms * 1e-3) / 1e12 / peak_tf, "frac_of_sustained": flops / (ms * 1e-3) / 1e12 / peak_sus,
                  "cublas_bf16_gemm_alone_ms": ms_gemm, "cublas_gemm_tflops": flops / (ms_gemm * 1e-3) / 1e12,
                  "unfused_linear_then_mercury_score_ms": ms_unfused, "speedup_vs_unfused": ms_unfused / ms,
                  "logits_bytes_not_written": n * c * 2}))
