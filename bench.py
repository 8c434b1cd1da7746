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
    "cfg5full": dict(n=None, c=32768, dtype="bf16", b=256, payload=None,
                     note="BASELINE configs[4] whole: the 2^24 x 32768 bf16 pool (1.1 TB) split over the ranks (--gpus >= 2); a shard "
                          "that does not fit in HBM (2 / 4 GPUs: 550 / 275 GB) streams from pinned host memory through pool.ChunkRing"),
}
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md


def algorithmic_bytes(n, c, s, b, row_bytes):
    """logits + int64 label read once, fp32 score written once; per draw the payload
    row read+written, label and index read+written, nprob + seed + idx written."""
    return n * (c * s + 8 + 4) + b * (2 * row_bytes + 2 * 8 + 2 * 8 + 4 + 4 + 8)


def k1_traffic(workload):
    """DRAM bytes (read + write) of one launch from the committed `ncu --set full` captures (profiles/r2_traffic.json):
    `<workload>` = one K1 launch, `<workload>_step` = one launch of the one-kernel step."""
    p = os.path.join(ROOT, "profiles", "r2_traffic.json")
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


def auto_pools(wl, pools_arg):
    """>= 300 MB of distinct pools in rotation (> 126 MB L2); a pool of tens of GB is its own L2 flush."""
    s = 4 if wl["dtype"] == "f32" else 2
    pool_bytes = wl["n"] * (wl["c"] * s + 8)
    return pools_arg or (1 if pool_bytes >= (16 << 30) else max(2, min(8, -(-(300 << 20) // pool_bytes))))


def make_config(args, wl, world, exchange=None):
    """The `config` object of the JSON line: the same for both arms (--impl ours / reference) of one command."""
    s = 4 if wl["dtype"] == "f32" else 2
    pool_bytes = wl["n"] * (wl["c"] * s + 8)
    npools = auto_pools(wl, args.pools)
    rotated = npools * pool_bytes
    l2 = (f"{npools} distinct pools rotated ({rotated >> 20} MiB > 126 MiB L2)" if rotated > (126 << 20) else
          f"{npools} distinct pools rotated ({rotated >> 20} MiB: they FIT the 126 MiB L2 -- a latency-bound workload, not an HBM number)")
    row_bytes = 4 * (wl["payload"][0] * wl["payload"][1] * wl["payload"][2]) if wl["payload"] else 0
    return {"workload": args.workload, "note": wl["note"], "candidates_per_gpu": wl["n"], "classes": wl["c"], "draws": wl["b"],
            "payload_row_bytes": row_bytes, "l2": l2,
            "cdf": "exact serial-fp32 (bit-identical to torch.multinomial on CPU)",
            "launch": (f"CUDA graph replay, one graph of {npools} steps (one per pool)" if args.graph else "eager launches"),
            "multi_gpu": None if world == 1 else MULTI_GPU_NOTE[args.shard_mode],
            "exchange": None if world == 1 else EXCHANGE_NOTE[exchange or (args.exchange if args.shard_mode == "totals" else "nccl")]}


def run_reference(args, wl):
    """The reference arm: the reference's own CPU arithmetic (oracle port on torch CPU ops: the reference is pure
    Python over torch, and its modules do not import under this torch) on all host threads.  W warm-up + K timed
    steps; a step is one pass over one pool (a row slice for the widest shapes); with N GPUs named, the N shards are
    processed one after the other by the one host (same rate)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import mercury_oracle as orc

    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.set_num_threads(os.cpu_count() or 1)
    n, c, b = wl["n"], wl["c"], wl["b"]
    rows = n
    while rows * c > 64_000_000 and rows > 4096:   # bound one pass to a few seconds
        rows //= 2
    g = torch.Generator().manual_seed(0)
    dt = torch.float32 if wl["dtype"] == "f32" else torch.bfloat16
    x = (torch.randn(rows, c, generator=g) * 2).to(dt)
    y = torch.randint(0, c, (rows,), generator=g)
    gen = torch.Generator().manual_seed(3)

    def one():
        r = orc.update_samples_flat(x, y, b, orc.EMAverage(), generator=gen)
        orc.reweighted_loss(torch.ones(b), r["nprob"])

    for _ in range(max(1, args.warmup)):
        one()
    steps = max(1, args.steps)
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    el = time.perf_counter() - t0
    value = rows * steps / el
    base = {"value": value, "unit": "candidates/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{steps} passes over {rows} x {c} {wl['dtype']} candidates, B={b} "
                      f"({'full pool' if rows == n else 'row slice of the pool, rate per candidate'}), {el:.1f} s; "
                      f"with {world} GPU(s) named the shards would be processed one after the other at this rate"}
    line = {
        "impl": "reference", "metric": "importance-scored+sampled candidates/sec", "value": value, "unit": "candidates/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * n / value,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
        "config": make_config(args, wl, world),
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": "candidates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


MULTI_GPU_NOTE = {
    "totals": "pool sharded by rank; every rank builds the exact chain of its own shard; one fp64 sum and one fp32 chain total "
              "per rank are exchanged; two-level draw (shard by totals, then position) by the owning rank; the drawn rows "
              "stay on their owner in the timed step (other ranks hold zero rows): the batch is NOT assembled",
    "scores": "pool sharded by rank; NCCL all-gather of scores; replicated exact CDF of the whole pool; rows gathered by their owner "
              "(other ranks hold zero rows: the batch is NOT assembled)",
}


EXCHANGE_NOTE = {
    "peer": "in-kernel stores into the peers' exchange boxes (CUDA IPC over NVLink/NVSwitch), no collective in the step",
    "nccl": "NCCL all-gather",
}


def bind_host_memory_near(dev_index):
    """Best effort, for the end-to-end leg at N > 1: prefer host memory (and cores) on the NUMA node the GPU hangs off, so that
    the pinned buffers of eight ranks do not all sit behind one socket's memory controller and the inter-socket link.  Returns
    a short description of what was done (goes into the JSON line)."""
    try:
        pr = torch.cuda.get_device_properties(dev_index)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return f"gpu {bdf}: no NUMA node reported"
        allowed = open("/proc/self/status").read().split("Mems_allowed_list:")[1].split()[0]
        nodes = set()
        for part in allowed.split(","):
            lo, _, hi = part.partition("-")
            nodes.update(range(int(lo), int(hi or lo) + 1))
        if node not in nodes:
            return f"gpu {bdf} on node {node}: not in this process's allowed memory nodes ({allowed})"
        import ctypes

        libc = ctypes.CDLL(None, use_errno=True)
        mask = (ctypes.c_ulong * 16)()
        mask[node // 64] = 1 << (node % 64)
        rc = libc.syscall(238, 1, ctypes.byref(mask), 16 * 64 + 1)          # set_mempolicy(MPOL_PREFERRED, node)  (x86_64)
        what = f"gpu {bdf} on node {node}: memory policy preferred({node})" + ("" if rc == 0 else f" failed (errno {ctypes.get_errno()})")
        try:
            cpus = set()
            for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
            cpus &= os.sched_getaffinity(0)
            if cpus:
                os.sched_setaffinity(0, cpus)
                what += f", {len(cpus)} cores of that node"
        except Exception:  # noqa: BLE001
            pass
        return what
    except Exception as e:  # noqa: BLE001
        return f"unavailable ({type(e).__name__})"


def timed_steps(step, npools, steps, warmup, use_graph, barrier, dist=None, dev=None, min_warm=3):
    """W warm-up steps, then EXACTLY `steps` timed steps between two barriers, CUDA events on the launching stream, max
    over ranks.  With graphs: one graph holding one step per pool, replayed steps // npools times, plus a graph with the
    remaining steps % npools steps.  Returns (ms per step, replay callable for extra untimed steps)."""
    for i in range(max(min_warm, warmup)):
        step(i)
    barrier()
    full = rest = None
    nrest = steps % npools
    if use_graph:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            step(0)  # allocates this stream's workspace outside capture
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        full = torch.cuda.CUDAGraph()
        with torch.cuda.graph(full, stream=side):
            for i in range(npools):
                step(i)
        if nrest:
            rest = torch.cuda.CUDAGraph()
            with torch.cuda.graph(rest, stream=side):
                for i in range(nrest):
                    step(i)
            rest.replay()
        full.replay()
        torch.cuda.synchronize()

    def run(nsteps):
        if full is None:
            for i in range(nsteps):
                step(i)
            return
        for _ in range(nsteps // npools):
            full.replay()
        if nsteps % npools:
            if nsteps % npools == nrest and rest is not None:
                rest.replay()
            else:
                for i in range(nsteps % npools):
                    step(i)

    if min_warm >= 3:
        run(max(npools, (max(3, warmup) // npools + 1) * npools))   # warm-up through the same path as the timed steps
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    run(steps)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / steps
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    return ms, run


def step_launches(ops, wl, world, exchange, shard_mode, fast=False):
    """Kernels of ours per step and rank.  One GPU, small rows: the whole step is ONE cooperative kernel (fused_step_kernel)
    up to 1.5 Mi candidates; else the chain score -> summarize -> spine -> draw (4).  fp64 mode outside the one-kernel
    step: score, chunk sums, statistics, chunk sums of p, scan, draw (6)."""
    s = 4 if wl["dtype"] == "f32" else 2
    one_kernel = wl["c"] * s <= 256 and wl["n"] <= 768 * 2048 and wl["b"] <= 1024
    if world == 1:
        return 1 if one_kernel else (6 if fast else 4)
    return 4 if exchange == "peer" else 6 if shard_mode == "scores" else 7


def bench_single(ops, wl, dev, args, npools, steps, warmup, payload_ok=True, rank=0):
    """One workload on one GPU: the exact step, the fp64 step, K1 alone.  Returns a dict of measurements and the live
    tensors the caller may reuse (pools, payload, index, u)."""
    n, c, b = wl["n"], wl["c"], wl["b"]
    s = 4 if wl["dtype"] == "f32" else 2
    row_bytes = 0
    payload = None
    if wl["payload"] and payload_ok:
        payload = torch.rand((n,) + wl["payload"], device=dev, generator=torch.Generator(device=dev).manual_seed(2))
        row_bytes = payload[0].numel() * 4
    index = torch.arange(n, device=dev) + rank * n
    pools = [make_pool(wl, dev, seed=100 * rank + i) for i in range(npools)]
    gen = torch.Generator().manual_seed(3)
    u_host = torch.rand(b, dtype=torch.float64, generator=gen).pin_memory()
    u = u_host.to(dev)
    ema = torch.zeros(1, device=dev)
    status = ops.new_status(dev)

    def barrier():
        torch.cuda.synchronize()

    def make_step(mode, out):
        def step(i):
            lg, lb = pools[i % npools]
            r = ops.update_samples(lg, lb, u, ema, False, data=payload, index=index, out=out, status=status, cdf_mode=mode)
            out.update({k: v for k, v in r.items() if k in ("scores", "stats", "idx", "nprob", "seed", "data", "labels", "index")})
            return r
        return step

    out = {}
    step = make_step("exact", out)
    ms, run = timed_steps(step, npools, steps, warmup, args.graph, barrier)
    ops.check_status(status)
    idx_exact = []
    for i in range(min(npools, 4)):   # indices of the exact mode on the first pools (for the fp64 mode's agreement figure)
        step(i)
        idx_exact.append(out["idx"].clone())
    out_f = {}
    step_f = make_step("f64", out_f)
    ms_f, _ = timed_steps(step_f, npools, steps, warmup, args.graph, barrier)
    ops.check_status(status)
    agree = []
    for i in range(min(npools, 4)):
        step_f(i)
        agree.append(float((out_f["idx"] == idx_exact[i]).float().mean()))
    # ---- K1 (score) alone, one launch per pool in a graph, timed on its launching stream
    sc = torch.empty(n, dtype=torch.float32, device=dev)
    st = ops.new_status(dev)
    kstream = torch.cuda.Stream()
    kstream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(kstream):
        for i in range(3):
            ops.score(*pools[i % npools], status=st, out=sc)
    torch.cuda.synchronize()
    reps = max(20, -(-max(steps, 10) // npools))
    kg = torch.cuda.CUDAGraph()
    with torch.cuda.graph(kg, stream=kstream):
        for i in range(npools):
            ops.score(*pools[i], status=st, out=sc)
    with torch.cuda.stream(kstream):
        for _ in range(10):   # (the GPU idled while the graph was captured: bring the clocks back up first)
            kg.replay()
        a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(kstream)
        for _ in range(reps):
            kg.replay()
        z.record(kstream)
    torch.cuda.synchronize()
    k1_ms = a.elapsed_time(z) / (reps * npools)
    peak, peak_src = peak_hbm()
    step_bytes = algorithmic_bytes(n, c, s, b, row_bytes)
    k1_bytes = n * (c * s + 8 + 4)
    res = {
        "ms_per_step": ms, "value": n / (ms * 1e-3),
        "step_bytes": step_bytes, "step_gbs": step_bytes / (ms * 1e-3) / 1e9, "step_frac": step_bytes / (ms * 1e-3) / 1e9 / peak,
        "fast": {"ms_per_step": ms_f, "value": n / (ms_f * 1e-3), "frac": step_bytes / (ms_f * 1e-3) / 1e9 / peak,
                 "index_agreement_vs_exact": sum(agree) / len(agree),
                 "what": "cdf_mode='f64': fp64 running totals, no serial walk; same distribution, not the same indices"},
        "k1_ms": k1_ms, "k1_bytes": k1_bytes, "k1_gbs": k1_bytes / (k1_ms * 1e-3) / 1e9, "k1_frac": k1_bytes / (k1_ms * 1e-3) / 1e9 / peak,
        "peak": peak, "peak_src": peak_src, "row_bytes": row_bytes,
    }
    live = {"pools": pools, "payload": payload, "index": index, "u": u, "u_host": u_host, "run": run, "status": status}
    return res, live


def e2e_single(ops, wl, dev, live, args, n, c, s, b):
    """The same step through the public API with HOST buffers (pinned): logits, labels and uniforms copied in, idx / weights /
    seed / statistics copied out, every step, all inside the timed region.  Pools above 8 GB: a 64 Ki-row slice (the copy
    dominates and is linear in the rows), rate per candidate."""
    pools, payload, index, u_host = live["pools"], live["payload"], live["index"], live["u_host"]
    pool_bytes = n * (c * s + 8)
    e2e_rows = n if pool_bytes <= (8 << 30) else 1 << 16
    lg0, lb0 = pools[0][0][:e2e_rows], pools[0][1][:e2e_rows]
    e2e_payload = payload[:e2e_rows] if payload is not None else None
    e2e_index = index[:e2e_rows]
    h_logits = torch.empty(lg0.shape, dtype=lg0.dtype).pin_memory()
    h_labels = torch.empty(lb0.shape, dtype=lb0.dtype).pin_memory()
    h_logits.copy_(lg0)
    h_labels.copy_(lb0)
    d_logits, d_labels, d_u = torch.empty_like(lg0), torch.empty_like(lb0), torch.empty_like(live["u"])
    h_idx = torch.empty(b, dtype=torch.int64).pin_memory()
    h_np = torch.empty(b, dtype=torch.float32).pin_memory()
    h_seed = torch.empty(b, dtype=torch.float32).pin_memory()
    h_stats = torch.empty(8, dtype=torch.float32).pin_memory()
    ema2 = torch.zeros(1, device=dev)
    e2e_out = {}

    def e2e_step(first=False):
        d_logits.copy_(h_logits, non_blocking=True)
        d_labels.copy_(h_labels, non_blocking=True)
        d_u.copy_(u_host, non_blocking=True)
        rr = ops.update_samples(d_logits, d_labels, d_u, ema2, first, data=e2e_payload, index=e2e_index, out=e2e_out)
        e2e_out.update({k: v for k, v in rr.items() if v is not None})
        h_idx.copy_(rr["idx"], non_blocking=True)
        h_np.copy_(rr["nprob"], non_blocking=True)
        h_seed.copy_(rr["seed"], non_blocking=True)
        h_stats.copy_(rr["stats"], non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the caller reads the result every step

    for i in range(3):
        e2e_step(first=(i == 0))
    torch.cuda.synchronize()
    e2e_steps = max(5, min(args.steps, 20))
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    return {"value": e2e_rows / (e2e_ms * 1e-3), "unit": "candidates/s", "h2d_bytes_per_step": e2e_rows * (c * s + 8) + b * 8,
            "d2h_bytes_per_step": b * (8 + 4 + 4) + 32, "ms_per_step": e2e_ms, "steps": e2e_steps, "rows_per_step": e2e_rows}


def main_single(args, wl, dev):
    from mercury_b200 import ops

    n, c, b = wl["n"], wl["c"], wl["b"]
    s = 4 if wl["dtype"] == "f32" else 2
    npools = auto_pools(wl, args.pools)
    with ClockSampler(dev.index or 0) as clocks:
        res, live = bench_single(ops, wl, dev, args, npools, args.steps, args.warmup)
        # keep the clocks sampled over a region long enough for nvidia-smi to see it (~0.6 s)
        live["run"](max(npools, min(20000, int(600.0 / max(res["ms_per_step"], 1e-3)))))
        torch.cuda.synchronize()
    ops.check_status(live["status"])
    e2e = e2e_single(ops, wl, dev, live, args, n, c, s, b)
    launches = step_launches(ops, wl, 1, None, None)
    one_kernel = launches == 1
    traffic = k1_traffic(args.workload)
    line = {
        "metric": "importance-scored+sampled candidates/sec", "value": res["value"], "unit": "candidates/s", "n_gpus": 1,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
        "config": make_config(args, wl, 1),
        # the headline fraction: the WHOLE step, which for small rows is one kernel launch
        "roofline": {"kernel": "whole step: " + ("fused_step_kernel (score + statistics + exact CDF + draws + gather, one cooperative launch)"
                                                 if one_kernel else "score -> summarize -> spine -> draw (4 launches; K1 dominates)"),
                     "bound": "hbm", "achieved": res["step_gbs"], "peak": res["peak"], "unit": "GB/s", "frac": res["step_frac"],
                     "traffic": k1_traffic(args.workload + "_step") if one_kernel else traffic, "peak_source": res["peak_src"], "ms_per_launch": res["ms_per_step"],
                     "algorithmic_bytes_per_launch": res["step_bytes"]},
        "k1_roofline": {"kernel": "K1 score (mercury_score) timed alone", "bound": "hbm", "achieved": res["k1_gbs"], "peak": res["peak"],
                        "unit": "GB/s", "frac": res["k1_frac"], "traffic": traffic, "ms_per_launch": res["k1_ms"],
                        "algorithmic_bytes_per_launch": res["k1_bytes"]},
        "fast_mode": res["fast"],
        "e2e": e2e,
        "gpu_launches": launches * args.steps,
        "clocks": clocks.summary(),
    }
    if n * c * 4 <= (32 << 30):   # (ATen upcasts the logits: needs N*C*4 bytes of scratch)
        line["library_baseline"] = library_baseline(live["pools"], live["payload"], live["index"], live["u"], b)
    del live
    torch.cuda.empty_cache()
    # ---- the bandwidth-bound workloads (BASELINE configs[2] and the configs[4] row shape), 2 pools each: a few seconds
    if args.workload == "cfg2" and not args.no_other_workloads:
        others = {}
        for name in ("cfg3", "cfg5"):
            w2 = dict(WORKLOADS[name])
            try:
                with ClockSampler(dev.index or 0) as ck:
                    r2, l2 = bench_single(ops, w2, dev, args, 2, 6, 3, payload_ok=False)
                    l2["run"](max(2, min(400, int(400.0 / max(r2["ms_per_step"], 1e-3)))))
                    torch.cuda.synchronize()
                others[name] = {"note": w2["note"], "ms_per_step": r2["ms_per_step"], "value": r2["value"],
                                "step_roofline": {"algorithmic_bytes": r2["step_bytes"], "achieved_gbs": r2["step_gbs"], "frac": r2["step_frac"]},
                                "k1_roofline": {"ms_per_launch": r2["k1_ms"], "frac": r2["k1_frac"]},
                                "fast_mode": {k: r2["fast"][k] for k in ("ms_per_step", "frac", "index_agreement_vs_exact")},
                                "pools": 2, "steps": 6, "clocks": ck.summary()}
                del l2
            except Exception as e:  # noqa: BLE001
                others[name] = {"error": str(e)[:200]}
            torch.cuda.empty_cache()
        line["other_workloads"] = others
    if not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(wl)
    print(json.dumps(line))


def parity_check(ops, sp, do_step, u, n, world, rank, dist, dev, mode):
    """Before the timed region: this step's draw against the oracle, on the job's own scores.  Every rank's shard of
    scores is gathered once; rank 0 recomputes the draw on the CPU (oracle.sharded_draw_two_level for the totals mode,
    the flat serial-fp32 draw for the scores mode) and compares the indices, which the ranks assemble by a SUM
    all-reduce (a draw is non-zero on its owner only)."""
    import numpy as np

    from oracle import mercury_oracle as orc

    r = do_step(0)
    torch.cuda.synchronize()
    idx = r["idx"].clone()
    if mode == "totals":
        dist.all_reduce(idx, op=dist.ReduceOp.SUM)   # (other ranks hold zeros)
    scores = [torch.empty(n, dtype=torch.float32, device=dev) for _ in range(world)]
    dist.all_gather(scores, r["scores"].contiguous())
    ok = None
    if rank == 0:
        st = r["stats"].cpu().numpy()
        sh = [((x.cpu().numpy() + np.float32(st[2])).astype(np.float32) / np.float32(st[3])).astype(np.float32) for x in scores]
        if mode == "totals":
            want = orc.sharded_draw_two_level(sh, u.cpu().numpy())
        else:
            want = orc.draw_with_replacement(np.concatenate(sh), u.cpu().numpy())
        ok = bool(np.array_equal(idx.cpu().numpy(), want))
    return {"mode": "two-level (oracle.sharded_draw_two_level)" if mode == "totals" else "flat (serial fp32 chain over the gathered scores)",
            "idx_equal": ok, "ranks": world, "draws": int(u.numel())}, idx


def main_multi(args, wl, dev, world, rank):
    import torch.distributed as dist

    numa = bind_host_memory_near(dev.index or 0)
    dist.init_process_group("nccl", device_id=dev)
    from mercury_b200 import ops
    from mercury_b200.sharded import ShardedPool

    n, c, b = wl["n"], wl["c"], wl["b"]
    s = 4 if wl["dtype"] == "f32" else 2
    pool_bytes = n * (c * s + 8)
    streamed = pool_bytes > (150 << 30)          # the shard does not fit beside its scores: stream it through the chunk ring
    npools = 1 if streamed else auto_pools(wl, args.pools)
    row_bytes = 0
    payload = None
    if wl["payload"]:
        payload = torch.rand((n,) + wl["payload"], device=dev, generator=torch.Generator(device=dev).manual_seed(2))
        row_bytes = payload[0].numel() * 4
    index = torch.arange(n, device=dev) + rank * n
    u = torch.rand(b, dtype=torch.float64, generator=torch.Generator().manual_seed(3)).to(dev)
    out = {}
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

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    stream_note = None
    if streamed:
        from mercury_b200.pool import ChunkRing

        win = 1 << 14                                 # 1 GiB of bf16 logits per window
        ring = ChunkRing(win, c, dtype=torch.bfloat16 if s == 2 else torch.float32, device=dev, depth=3)
        wlg, wlb = make_pool(dict(wl, n=win), dev, seed=100 * rank)
        h_logits, h_labels = torch.empty(wlg.shape, dtype=wlg.dtype).pin_memory(), torch.empty(wlb.shape, dtype=wlb.dtype).pin_memory()
        h_logits.copy_(wlg)
        h_labels.copy_(wlb)
        del wlg, wlb
        labels_dev = torch.empty(n, dtype=torch.int64, device=dev)
        nwin = n // win
        if sp.peer is not None:
            sp.peer.set_wait(300_000)   # (ranks that share a host link can be seconds apart after hundreds of GB of copies)
        stream_note = (f"shard of {pool_bytes >> 30} GiB streamed every step from pinned host memory through a ring of 3 x {win}-row windows "
                       f"({nwin} windows per step; the synthetic host window is reused, the copies are real): the step is bound by the "
                       f"host link, not by HBM")

        def step(i, assemble=False):
            ring.score(((h_logits, h_labels) for _ in range(nwin)), sp.local_scores, labels_dev)
            return sp.update_samples(None, labels_dev, u, data=None, index=index, assemble=assemble)
    else:
        pools = [make_pool(wl, dev, seed=100 * rank + i) for i in range(npools)]

        def step(i, assemble=False):
            lg, lb = pools[i % npools]
            return sp.update_samples(lg, lb, u, data=payload, index=index, out=out, assemble=assemble)

    pc, _ = parity_check(ops, sp, step, u, n, world, rank, dist, dev, args.shard_mode)
    use_graph = args.graph and not streamed and os.environ.get("MERCURY_BENCH_GRAPH_NCCL", "1") == "1"
    min_warm = 1 if streamed else 3
    with ClockSampler(dev.index or 0) as clocks:
        ms, run = timed_steps(step, npools, args.steps, args.warmup, use_graph, barrier, dist, dev, min_warm=min_warm)
        if not streamed:
            extra = max(npools, min(20000, int(600.0 / max(ms, 1e-3))))   # (the SAME number of steps on every rank)
            run(extra)
        torch.cuda.synchronize()
    r = step(0)
    torch.cuda.synchronize()
    ops.check_status(r["status"])
    if r.get("xstatus") is not None:
        ops.check_status(r["xstatus"])
    # ---- the same step with the batch ASSEMBLED on every rank: one all-reduce over the step's output arena
    assembled = None
    if args.shard_mode == "totals":
        ms_a, _ = timed_steps(lambda i: step(i, True), npools, args.steps, args.warmup, use_graph, barrier, dist, dev, min_warm=min_warm)
        ra = step(0, True)
        torch.cuda.synchronize()
        hi_, lo_ = ra["idx"].clone(), ra["idx"].clone()
        dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
        mine = (ra["idx"] >= sp.lo) & (ra["idx"] < sp.lo + n)
        same = bool(torch.equal(hi_, lo_)) and bool(torch.equal(mine, ra["owned"]))   # one batch everywhere; each draw has one owner
        rows_ok = bool(torch.equal(ra["index"][mine], index[ra["idx"][mine] - sp.lo]))
        if payload is not None:
            rows_ok = rows_ok and bool(torch.equal(ra["data"][mine], payload[ra["idx"][mine] - sp.lo]))
        t = torch.tensor([int(same and rows_ok)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        assembled = {"ms_per_step": ms_a, "value": n * world / (ms_a * 1e-3), "unit": "candidates/s",
                     "collective": "ONE all-reduce (SUM over the raw 32-bit words) of the step's output arena: idx, nprob, seed, labels, index "
                                   "and the payload rows in one buffer -- every row is zero on all ranks but its owner",
                     "arena_bytes_per_rank": int(sp._arena.numel()) if sp._arena is not None else None,
                     "batch_equal_on_all_ranks_and_to_the_owners_rows": bool(int(t))}
    peak, peak_src = peak_hbm()
    step_bytes = algorithmic_bytes(n, c, s, b, row_bytes)
    # ---- end to end with host buffers: every rank copies its shard's logits, labels and the uniforms in, results out
    e2e_rows = n if pool_bytes <= (8 << 30) else 1 << 16
    e2e = None
    if streamed:
        # the timed step above already starts from host memory; here the results are read back as well
        h_idx, h_np = torch.empty(b, dtype=torch.int64).pin_memory(), torch.empty(b, dtype=torch.float32).pin_memory()
        barrier()
        t0 = time.perf_counter()
        rr = step(0)
        h_idx.copy_(rr["idx"], non_blocking=True)
        h_np.copy_(rr["nprob"], non_blocking=True)
        barrier()
        e2e_ms = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t)
        e2e = {"value": n * world / (e2e_ms * 1e-3), "unit": "candidates/s", "h2d_bytes_per_step": world * pool_bytes,
               "d2h_bytes_per_step": world * b * 12, "ms_per_step": e2e_ms, "steps": 1, "rows_per_step": n * world}
    elif e2e_rows == n:
        lg0, lb0 = pools[0]
        h_logits = torch.empty(lg0.shape, dtype=lg0.dtype).pin_memory()
        h_labels = torch.empty(lb0.shape, dtype=lb0.dtype).pin_memory()
        h_logits.copy_(lg0)
        h_labels.copy_(lb0)
        u_host = u.cpu().pin_memory()
        d_logits, d_labels, d_u = torch.empty_like(lg0), torch.empty_like(lb0), torch.empty_like(u)
        h_idx = torch.empty(b, dtype=torch.int64).pin_memory()
        h_np = torch.empty(b, dtype=torch.float32).pin_memory()
        h_seed = torch.empty(b, dtype=torch.float32).pin_memory()
        h_stats = torch.empty(8, dtype=torch.float32).pin_memory()

        def e2e_step():
            d_logits.copy_(h_logits, non_blocking=True)
            d_labels.copy_(h_labels, non_blocking=True)
            d_u.copy_(u_host, non_blocking=True)
            rr = sp.update_samples(d_logits, d_labels, d_u, data=payload, index=index, out=out)
            h_idx.copy_(rr["idx"], non_blocking=True)
            h_np.copy_(rr["nprob"], non_blocking=True)
            h_seed.copy_(rr["seed"], non_blocking=True)
            h_stats.copy_(rr["stats"], non_blocking=True)
            torch.cuda.current_stream().synchronize()

        for i in range(3):
            e2e_step()
        barrier()
        e2e_steps = max(5, min(args.steps, 20))
        t0 = time.perf_counter()
        for i in range(e2e_steps):
            e2e_step()
        barrier()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
        t = torch.tensor([e2e_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t)
        e2e = {"value": n * world / (e2e_ms * 1e-3), "unit": "candidates/s", "h2d_bytes_per_step": world * (n * (c * s + 8) + b * 8),
               "d2h_bytes_per_step": world * (b * (8 + 4 + 4) + 32), "ms_per_step": e2e_ms, "steps": e2e_steps, "rows_per_step": n * world,
               "host_memory": numa}
    else:
        e2e = {"value": None, "unit": "candidates/s", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None,
               "note": "pool above 8 GB per GPU: the host path of this shape is timed on one GPU (row slice); not repeated per rank"}
    if rank == 0:
        value = n * world / (ms * 1e-3)
        line = {
            "metric": "importance-scored+sampled candidates/sec", "value": value, "unit": "candidates/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(min_warm, args.warmup), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
            "config": dict(make_config(args, wl, world, exchange), **({"launch": "eager launches (copy stream + compute stream)",
                                                                       "l2": "the shard streams through HBM once per step"} if streamed else {})),
            "roofline": {"kernel": "whole step per GPU: score -> summarize -> spine -> draw on the rank's shard, exchanges inside",
                         "bound": "hbm", "achieved": step_bytes / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": step_bytes / (ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src, "ms_per_launch": ms,
                         "algorithmic_bytes_per_launch": step_bytes},
            "parity_check": pc,
            "assembled": assembled,
            "streamed": stream_note,
            "e2e": e2e,
            "gpu_launches": (step_launches(ops, wl, world, exchange, args.shard_mode) + (n // (1 << 14) if streamed else 0)) * world * args.steps,
            "clocks": clocks.summary(),
        }
        print(json.dumps(line))
    # Graphs that captured NCCL work keep the communicator busy at teardown (destroy_process_group was seen to block):
    # agree that everybody is done, then leave without the collective teardown.
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


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
    ap.add_argument("--no-other-workloads", action="store_true", help="N=1, cfg2: skip the short cfg3 / cfg5 blocks")
    ap.add_argument("--shard-mode", choices=["totals", "scores"], default=os.environ.get("MERCURY_SHARD_MODE", "totals"),
                    help="N>1 only: 'totals' = shard-local exact chains, only per-shard sums/totals exchanged, two-level draw; "
                         "'scores' = all-gather the scores and replicate the flat pool's exact CDF on every rank")
    ap.add_argument("--exchange", choices=["peer", "nccl"], default=os.environ.get("MERCURY_EXCHANGE", "peer"),
                    help="N>1, --shard-mode totals: 'peer' = the sums/totals travel by stores into the peers' mapped memory from "
                         "inside our kernels (CUDA IPC over NVLink); 'nccl' = two small NCCL all-gathers")
    ap.add_argument("--no-graph", dest="graph", action="store_false", help="launch every step eagerly (no CUDA graph)")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    if wl["n"] is None:      # cfg5full: the 2^24 pool divided over the ranks
        w = int(os.environ.get("WORLD_SIZE", "1"))
        if w < 2 or (1 << 24) % w:
            raise SystemExit("cfg5full is the multi-GPU configuration: launch it under torchrun with 2, 4 or 8 ranks")
        wl["n"] = (1 << 24) // w

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
    if world > 1:
        main_multi(args, wl, dev, world, rank)
    else:
        main_single(args, wl, dev)


if __name__ == "__main__":
    main()
