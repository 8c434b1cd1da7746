"""f1: step time of the importance-sampled training loop (collab.Trainer.train = pytorch_collab.py:119-199) with and
without the sampling call overlapped with the gradient all-reduce, and with the sampling call replayed from a CUDA graph.

  python tools/bench_overlap.py                       # one GPU: no collective, shows what the side stream / graph buy alone
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/bench_overlap.py

Model: a CIFAR ResNet-18 (11.17 M parameters = the 44.7 MB gradient message of pytorch_collab.py:242), random weights,
synthetic CIFAR-shaped batches.  One JSON line on rank 0; times are CUDA-event times, max over ranks."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


class Block(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.c1, self.b1 = nn.Conv2d(cin, cout, 3, stride, 1, bias=False), nn.BatchNorm2d(cout)
        self.c2, self.b2 = nn.Conv2d(cout, cout, 3, 1, 1, bias=False), nn.BatchNorm2d(cout)
        self.short = None
        if stride != 1 or cin != cout:
            self.short = nn.Sequential(nn.Conv2d(cin, cout, 1, stride, bias=False), nn.BatchNorm2d(cout))

    def forward(self, x):
        y = self.b2(self.c2(F.relu(self.b1(self.c1(x)))))
        return F.relu(y + (x if self.short is None else self.short(x)))


class Net18(nn.Module):
    """Same attribute names as the reference's ResNet (conv1, bn1, layer1..4, linear) so split_head applies."""

    def __init__(self, classes=10):
        super().__init__()
        self.conv1, self.bn1 = nn.Conv2d(3, 64, 3, 1, 1, bias=False), nn.BatchNorm2d(64)
        cfg, cin, layers = [(64, 1), (128, 2), (256, 2), (512, 2)], 64, []
        for cout, s in cfg:
            layers.append(nn.Sequential(Block(cin, cout, s), Block(cout, cout, 1)))
            cin = cout
        self.layer1, self.layer2, self.layer3, self.layer4 = layers
        self.linear = nn.Linear(512, classes)

    def forward(self, x):
        x = F.relu(self.bn1(self.conv1(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.linear(F.avg_pool2d(x, 4).flatten(1))


class Loader:
    def __init__(self, batches, batch_size):
        self.batches, self.batch_size = batches, batch_size

    def __iter__(self):
        return iter(self.batches)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--presample", type=int, default=32)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from mercury_b200.collab import Trainer

    gen = torch.Generator().manual_seed(rank)
    presam = [(torch.arange(i * a.presample, (i + 1) * a.presample), torch.rand(a.presample, 3, 32, 32, generator=gen).pin_memory(),
               torch.randint(0, 10, (a.presample,), generator=gen).pin_memory()) for i in range(64)]
    total = a.steps + a.warmup
    out = {}
    for overlap in (False, True):
        for graphed in (False, True):
            torch.manual_seed(0)
            net = Net18().to(dev)
            tr = Trainer(net, torch.optim.Adam(net.parameters(), lr=1e-3 * world), Loader([None] * total, a.batch),
                         Loader(presam, a.presample), None, dev)
            tr.overlap_sampling, tr.graph_sampling = overlap, graphed
            tr.train(max_steps=a.warmup)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            tr.train(max_steps=a.steps)
            e1.record()
            torch.cuda.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            out[f"overlap={int(overlap)},graph={int(graphed)}"] = round(float(ms), 4)
    # the two things the overlap can hide behind each other, each timed alone
    def timed(fn, reps=20):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return round(float(t), 4)

    from mercury_b200.sampler import EMAverage
    ema = EMAverage(device=dev)
    tr.graph_sampling = False
    parts = {"presample_forward_plus_sampling_ms": timed(lambda: tr.update_samples(ema))}
    from mercury_b200 import ops
    logits, labels, datas, index = tr._pool.tensors()
    bounds = tr._pool.growing_bounds()
    parts["sampling_call_ms"] = timed(lambda: ops.update_samples(logits, labels, tr._u_dev, ema.state, False, data=datas, index=index,
                                                                bounds=bounds))     # the fused call alone, pool as it stands
    if world > 1:
        tr._zero_grads()
        parts["allreduce_avg_ms"] = timed(tr.average_gradients)
    if rank == 0:
        nparam = sum(p.numel() for p in Net18().parameters())
        print(json.dumps({"metric": "train_step_ms", "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "batch": a.batch,
                          "presample_batches": 10, "presample_batch": a.presample, "grad_message_bytes": 4 * nparam,
                          "collective": "ncclAllReduce(AVG) in place on the gradient bucket" if world > 1 else None,
                          "ms_per_step": out, "alone": parts}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
