#!/usr/bin/env python
"""Benchmark of the importance-sampling hot path (BASELINE.json metric: importance-scored +
sampled candidates/sec and % of the HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg3|cfg1|cfg5] [--impl reference]

A "step" is one pass of the hot path over one pool: score -> pool statistics -> exact CDF ->
B draws from caller-supplied uniforms -> gather + N*p weights + backward seed.  Prints ONE JSON
line (rank 0).  See DESIGN.md "Measurement" for what each field means.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOADS = {
    # name: candidates per GPU, classes, dtype, draws, payload shape per candidate (fp32)
    "cfg1": dict(n=50_000, c=10, dtype="f32", b=128, payload=(3, 32, 32), note="BASELINE configs[0] shape (pool 50k)"),
    "cfg2": dict(n=1_000_000, c=10, dtype="f32", b=128, payload=(3, 32, 32), note="BASELINE configs[1]: 1M x 10 fp32"),
    "cfg3": dict(n=1_280_000, c=1000, dtype="bf16", b=256, payload=None, note="BASELINE configs[2]: 1.28M x 1000 bf16"),
    "cfg5": dict(n=131_072, c=32768, dtype="bf16", b=256, payload=None,
                 note="BASELINE configs[4] row shape (32k-class bf16), 128k-candidate resident slice per GPU"),
    "cfg5shard": dict(n=1 << 21, c=32768, dtype="bf16", b=256, payload=None,
                      note="BASELINE configs[4] at full size: one GPU's shard of the 2^24 x 32768 bf16 pool split 8 ways (137 GB resident)"),
}
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md


def algorithmic_bytes(n, c, s, b, row_bytes):
    """logits + int64 label read once, fp32 score written once; per draw the payload
    row read+written, label and index read+written, nprob + seed + idx written."""
    return n * (c * s + 8 + 4) + b * (2 * row_bytes + 2 * 8 + 2 * 8 + 4 + 4 + 8)


def k1_traffic(workload):
    """DRAM bytes (read + write) of one K1 launch from the committed `ncu --set full` capture."""
    p = os.path.join(ROOT, "profiles", "r1_k1_traffic.json")
    try:
        t = json.load(open(p))[workload]
        return t["dram_bytes_read"] + t["dram_bytes_write"]
    except Exception:
        return None


def peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


def make_pool(wl, dev, seed, with_payload=True):
    """Synthetic pool of BASELINE.md's distributions: logits randn*2, labels randint, payload rand."""
    g = torch.Generator(device=dev).manual_seed(seed)
    n, c = wl["n"], wl["c"]
    dt = torch.float32 if wl["dtype"] == "f32" else torch.bfloat16
    logits = torch.empty(n, c, dtype=dt, device=dev)
    step = max(1, (1 << 27) // c)  # generate in slices to bound temporary memory
    for i in range(0, n, step):
        logits[i:i + step] = (torch.randn(min(step, n - i), c, generator=g, device=dev) * 2).to(dt)
    labels = torch.randint(0, c, (n,), generator=g, device=dev)
    return logits, labels


def cpu_baseline(wl, budget_s=12.0, max_rows=None, one_thread_s=4.0):
    """The reference's own CPU arithmetic (oracle restatement on torch CPU ops), all host threads."""
    from oracle import mercury_oracle as orc

    torch.set_num_threads(os.cpu_count() or 1)
    n, c, b = wl["n"], wl["c"], wl["b"]
    rows = n
    if max_rows:
        rows = min(n, max_rows)
    # bound the sample so that one pass stays within a few seconds
    while rows * c > 64_000_000 and rows > 4096:
        rows //= 2
    g = torch.Generator().manual_seed(0)
    dt = torch.float32 if wl["dtype"] == "f32" else torch.bfloat16
    x = (torch.randn(rows, c, generator=g) * 2).to(dt)
    y = torch.randint(0, c, (rows,), generator=g)
    gen = torch.Generator().manual_seed(3)
    t0 = time.perf_counter()
    reps = 0
    while True:
        ema = orc.EMAverage()
        r = orc.update_samples_flat(x, y, b, ema, generator=gen)
        _ = orc.reweighted_loss(torch.ones(b), r["nprob"])
        reps += 1
        el = time.perf_counter() - t0
        if el > budget_s or reps >= 5000:
            break
    one = None
    if one_thread_s:
        torch.set_num_threads(1)
        t1, r1 = time.perf_counter(), 0
        while True:
            r = orc.update_samples_flat(x, y, b, orc.EMAverage(), generator=gen)
            r1 += 1
            e1 = time.perf_counter() - t1
            if e1 > one_thread_s or r1 >= 50:
                break
        one = rows * r1 / e1
        torch.set_num_threads(os.cpu_count() or 1)
    return {"value": rows * reps / el, "unit": "candidates/s", "cores": torch.get_num_threads(), "kind": "port", "one_thread_value": one,
            "sample": f"{reps} passes over {rows} x {c} {wl['dtype']} candidates, B={b} "
                      f"({'full pool' if rows == n else 'row slice of the pool, rate per candidate'}), {el:.1f} s"}


def library_baseline(pools, payload, index, u, b, steps=20):
    """Stock ATen kernels on this GPU for the same step (SURVEY 8d "library" baseline): F.cross_entropy, mean,
    EMA, normalise, torch.multinomial (Philox: not reproducible against the CPU reference), index gathers."""
    import torch.nn.functional as F

    def one(lg, lb):
        loss = F.cross_entropy(lg.float(), lb, reduction="none")
        w = loss + 0.5 * loss.mean()
        p = w / w.sum()
        idx = torch.multinomial(p, b, replacement=True)
        nprob = p[idx] * p.numel()
        out = (idx, nprob, (1.0 / b) / nprob, lb[idx], index[idx])
        return out + ((payload[idx],) if payload is not None else ())

    for i in range(3):
        one(*pools[i % len(pools)])
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(steps):
        one(*pools[i % len(pools)])
    z.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(z) / steps
    return {"value": pools[0][0].shape[0] / (ms * 1e-3), "unit": "candidates/s", "ms_per_step": ms,
            "what": "stock ATen CUDA ops, eager, same pools (results not reproducible against the CPU reference)"}


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warm = args.steps, args.warmup
    base = cpu_baseline(wl, budget_s=max(5.0, min(60.0, 2.0 * steps)))
    line = {
        "impl": "reference", "metric": "importance-scored+sampled candidates/sec", "value": base["value"], "unit": "candidates/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": 1e3 * wl["n"] / base["value"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
        "config": {"workload": args.workload, "note": wl["note"], "candidates_per_gpu": wl["n"], "classes": wl["c"],
                   "draws": wl["b"], "payload_row_bytes": 4 * (wl["payload"][0] * wl["payload"][1] * wl["payload"][2]) if wl["payload"] else 0,
                   "cdf": "torch.multinomial on CPU (serial fp32)", "launch": "torch CPU ops, all host threads"},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


MULTI_GPU_NOTE = {
    "totals": "pool sharded by rank; every rank builds the exact chain of its own shard; one fp64 sum and one fp32 chain total "
              "per rank are exchanged; two-level draw (shard by totals, then position) by the owning rank",
    "scores": "pool sharded by rank; NCCL all-gather of scores; replicated exact CDF of the whole pool; rows gathered by their owner",
}


EXCHANGE_NOTE = {
    "peer": "in-kernel stores into the peers' exchange boxes (CUDA IPC over NVLink/NVSwitch), no collective in the step",
    "nccl": "NCCL all-gather",
}


def main():
    if os.environ.get("MERCURY_BENCH_FAULT_TIMEOUT"):  # developer aid: dump every thread's stack if the run stalls
        import faulthandler

        faulthandler.dump_traceback_later(float(os.environ["MERCURY_BENCH_FAULT_TIMEOUT"]), exit=True)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pools", type=int, default=0, help="distinct pools rotated to defeat L2 (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--shard-mode", choices=["totals", "scores"], default=os.environ.get("MERCURY_SHARD_MODE", "totals"),
                    help="N>1 only: 'totals' = shard-local exact chains, only per-shard sums/totals exchanged, two-level draw; "
                         "'scores' = all-gather the scores and replicate the flat pool's exact CDF on every rank")
    ap.add_argument("--exchange", choices=["peer", "nccl"], default=os.environ.get("MERCURY_EXCHANGE", "peer"),
                    help="N>1, --shard-mode totals: 'peer' = the sums/totals travel by stores into the peers' mapped memory from "
                         "inside our kernels (CUDA IPC over NVLink); 'nccl' = two small NCCL all-gathers")
    ap.add_argument("--no-graph", dest="graph", action="store_false", help="launch every step eagerly (no CUDA graph)")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])

    if args.impl == "reference":
        run_reference(args, wl)
        return

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (there is no CPU path); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)

    from mercury_b200 import ops

    n, c, b = wl["n"], wl["c"], wl["b"]
    s = 4 if wl["dtype"] == "f32" else 2
    row_bytes = 0
    payload = None
    if wl["payload"]:
        payload = torch.rand((n,) + wl["payload"], device=dev, generator=torch.Generator(device=dev).manual_seed(2))
        row_bytes = payload[0].numel() * 4
    index = torch.arange(n, device=dev) + rank * n
    pool_bytes = n * (c * s + 8)
    # >= 300 MB in rotation (> 126 MB L2); a pool of tens of GB is its own L2 flush
    npools = args.pools or (1 if pool_bytes >= (16 << 30) else max(2, min(8, -(-(300 << 20) // pool_bytes))))
    pools = [make_pool(wl, dev, seed=100 * rank + i) for i in range(npools)]
    gen = torch.Generator().manual_seed(3)
    u_host = torch.rand(b, dtype=torch.float64, generator=gen).pin_memory()
    u = u_host.to(dev)
    ema = torch.zeros(1, device=dev)
    out = {}  # result buffers reused across steps: no allocation inside the timed region

    if world > 1:
        from mercury_b200.sharded import ShardedPool

        exchange = args.exchange if args.shard_mode == "totals" else "nccl"
        if exchange == "peer":
            # peer mapping is a set-up step that a box may refuse (no IPC between these GPUs): agree on it across ranks
            try:
                sp = ShardedPool(n_local=n, device=dev, mode="totals", exchange="peer")
                ok = 1
            except Exception as e:  # noqa: BLE001
                print(f"[bench] rank {rank}: peer exchange unavailable ({e}); using NCCL", file=sys.stderr)
                ok = 0
            t = torch.tensor([ok], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MIN)
            if int(t) == 0:
                exchange = "nccl"
        if exchange == "nccl":
            sp = ShardedPool(n_local=n, device=dev, mode=args.shard_mode)

    status = ops.new_status(dev)

    def step(i, first=False):
        lg, lb = pools[i % npools]
        if world > 1:
            return sp.update_samples(lg, lb, u, data=payload, index=index, out=out)
        r = ops.update_samples(lg, lb, u, ema, first, data=payload, index=index, out=out, status=status)
        out.update({k: v for k, v in r.items() if k in ("scores", "stats", "idx", "nprob", "seed", "data", "labels", "index")})
        return r

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(max(3, args.warmup)):
        r = step(i, first=(i == 0))
    barrier()
    # One CUDA graph per pool (the pools differ only in their pointers): the step is replayed with
    # no host work between its kernels.  Multi-rank steps capture their NCCL all-gather in the graph
    # as well (MERCURY_BENCH_GRAPH_NCCL=0 keeps them eager).
    graphs = None
    if args.graph and (world == 1 or os.environ.get("MERCURY_BENCH_GRAPH_NCCL", "1") == "1"):
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            step(0)  # allocates this stream's workspace outside capture
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graphs = []
        for i in range(npools):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                step(i)
            graphs.append(g)
        for g in graphs:
            g.replay()
        torch.cuda.synchronize()

    def run(i):
        if graphs is not None:
            graphs[i % npools].replay()
        else:
            step(i)

    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clocks:
        barrier()
        e0.record()
        for i in range(args.steps):
            run(i)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1) / args.steps
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        # keep the clocks sampled over a region long enough for nvidia-smi to see it (~0.6 s; the
        # SAME number of steps on every rank: a step contains a collective)
        for i in range(max(1, min(20000, int(600.0 / max(ms, 1e-3))))):
            run(i)
        torch.cuda.synchronize()
    ops.check_status(r["status"])
    total_candidates = n * world
    value = total_candidates / (ms * 1e-3)

    # ---- dominant kernel (K1 score) timed alone on its launching stream.  The launches are
    # captured into one graph (one launch per rotating pool) so that the events bracket GPU
    # time, not the Python/ctypes launch rate.
    sc = torch.empty(n, dtype=torch.float32, device=dev)
    st = ops.new_status(dev)
    kstream = torch.cuda.Stream()
    kstream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(kstream):
        for i in range(3):
            ops.score(*pools[i % npools], status=st, out=sc)
    torch.cuda.synchronize()
    reps = max(2, -(-max(args.steps, 10) // npools))
    if world == 1 or os.environ.get("MERCURY_BENCH_GRAPH_NCCL", "1") == "1":
        kg = torch.cuda.CUDAGraph()
        with torch.cuda.graph(kg, stream=kstream):
            for i in range(npools):
                ops.score(*pools[i], status=st, out=sc)
        with torch.cuda.stream(kstream):
            kg.replay()
            a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(kstream)
            for _ in range(reps):
                kg.replay()
            z.record(kstream)
    else:
        # multi-rank: no graph capture beside NCCL's watchdog; stall the stream so that the launches queue up
        with torch.cuda.stream(kstream):
            torch.cuda._sleep(int(4e6))
            a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(kstream)
            for _ in range(reps):
                for i in range(npools):
                    ops.score(*pools[i], status=st, out=sc)
            z.record(kstream)
    torch.cuda.synchronize()
    k1_ms = a.elapsed_time(z) / (reps * npools)
    peak, peak_src = peak_hbm()
    k1_bytes = n * (c * s + 8 + 4)
    k1_gbs = k1_bytes / (k1_ms * 1e-3) / 1e9
    step_bytes = algorithmic_bytes(n, c, s, b, row_bytes)

    # ---- end to end through the public API with HOST buffers (pinned), copies inside the timed region
    e2e_rows = n
    if pool_bytes > (8 << 30) and world > 1:
        raise SystemExit("bench.py: workloads with pools above 8 GB per GPU are single-GPU runs (their host path is timed on a slice)")
    if pool_bytes > (8 << 30):
        # a pool that does not fit pinned host memory comfortably: time the host path on a 64 Ki-row slice and
        # report the rate per candidate (the copy dominates and is linear in the rows)
        e2e_rows = 1 << 16
    lg0, lb0 = pools[0][0][:e2e_rows], pools[0][1][:e2e_rows]
    e2e_payload = payload[:e2e_rows] if payload is not None else None
    e2e_index = index[:e2e_rows]
    h_logits = torch.empty(lg0.shape, dtype=lg0.dtype).pin_memory()
    h_labels = torch.empty(lb0.shape, dtype=lb0.dtype).pin_memory()
    h_logits.copy_(lg0)
    h_labels.copy_(lb0)
    d_logits, d_labels, d_u = torch.empty_like(lg0), torch.empty_like(lb0), torch.empty_like(u)
    h_idx = torch.empty(b, dtype=torch.int64).pin_memory()
    h_np = torch.empty(b, dtype=torch.float32).pin_memory()
    h_seed = torch.empty(b, dtype=torch.float32).pin_memory()
    h_stats = torch.empty(8, dtype=torch.float32).pin_memory()
    ema2 = torch.zeros(1, device=dev)
    e2e_out = out if e2e_rows == n else {}

    def e2e_step(first=False):
        d_logits.copy_(h_logits, non_blocking=True)
        d_labels.copy_(h_labels, non_blocking=True)
        d_u.copy_(u_host, non_blocking=True)
        if world > 1:
            rr = sp.update_samples(d_logits, d_labels, d_u, data=payload, index=index, out=out)
        else:
            rr = ops.update_samples(d_logits, d_labels, d_u, ema2, first, data=e2e_payload, index=e2e_index, out=e2e_out)
        h_idx.copy_(rr["idx"], non_blocking=True)
        h_np.copy_(rr["nprob"], non_blocking=True)
        h_seed.copy_(rr["seed"], non_blocking=True)
        h_stats.copy_(rr["stats"], non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the caller reads the result every step

    for i in range(3):
        e2e_step(first=(i == 0))
    barrier()
    e2e_steps = max(5, min(args.steps, 20))
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    if dist is not None:
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t)
    h2d = world * (e2e_rows * (c * s + 8) + b * 8)   # whole job: every rank copies its shard's logits, labels and the uniforms in
    d2h = world * (b * (8 + 4 + 4) + 32)      # ... and idx / N*p / seed / stats out

    if rank == 0:
        # our kernels per step and rank: score, summarize (statistics folded in), spine, draw; sharded: + chunk sums, gather_owned
        # totals: score, chunk sums, sum(+exchange), summarize, spine, select(+exchange), draw; scores: + gather_owned
        # totals over peer memory: score, summarize (+ exchange of the shard sums), spine (+ publishes the chain total),
        # draw (+ shard selection): the same four launches as on one GPU
        per_rank = 4 if (world == 1 or exchange == "peer") else 6 if args.shard_mode == "scores" else 7
        launches_per_step = per_rank * world
        line = {
            "metric": "importance-scored+sampled candidates/sec", "value": value, "unit": "candidates/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
            "config": {"workload": args.workload, "note": wl["note"], "candidates_per_gpu": n, "classes": c, "draws": b,
                       "payload_row_bytes": row_bytes, "l2": f"{npools} distinct pools rotated ({npools * pool_bytes >> 20} MiB > 126 MiB L2)",
                       "cdf": "exact serial-fp32 (bit-identical to torch.multinomial on CPU)",
                       "launch": "CUDA graph replay, one graph per pool" if graphs is not None else "eager launches",
                       "multi_gpu": None if world == 1 else MULTI_GPU_NOTE[args.shard_mode],
                       "exchange": None if world == 1 else EXCHANGE_NOTE[exchange]},
            "step_roofline": {"algorithmic_bytes": step_bytes, "achieved_gbs": step_bytes / (ms * 1e-3) / 1e9,
                              "frac": step_bytes / (ms * 1e-3) / 1e9 / peak},
            "roofline": {"kernel": "K1 score (mercury_score)", "bound": "hbm", "achieved": k1_gbs, "peak": peak, "unit": "GB/s",
                         "frac": k1_gbs / peak, "traffic": k1_traffic(args.workload) if world == 1 else None, "peak_source": peak_src, "ms_per_launch": k1_ms,
                         "algorithmic_bytes_per_launch": k1_bytes},
            "e2e": {"value": e2e_rows * world / (e2e_ms * 1e-3), "unit": "candidates/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "ms_per_step": e2e_ms, "steps": e2e_steps,
                    "rows_per_step": e2e_rows * world},
            "gpu_launches": launches_per_step * args.steps,
            "clocks": clocks.summary(),
        }
        if world == 1 and n * c * 4 <= (32 << 30):   # (ATen upcasts the logits: needs N*C*4 bytes of scratch)
            line["library_baseline"] = library_baseline(pools, payload, index, u, b)
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(wl)
        print(json.dumps(line))
    if dist is not None:
        # Graphs that captured NCCL work keep the communicator busy at teardown (destroy_process_group
        # was seen to block): agree that everybody is done, then leave without the collective teardown.
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
