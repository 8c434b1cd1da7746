#!/usr/bin/env python
"""Generate tests/golden/*.npz.  Run in the build container (needs /root/reference).

Two kinds of fixture are produced:

1. ``ref_*``: outputs of the REFERENCE'S OWN CODE executed here, unmodified, by importing
   ``pytorch_collab.Trainer`` / ``util.EMAverage`` / ``util.Groupwise_Sampler`` /
   ``pytorch_model.ResNet18`` from /root/reference.  Three things the reference needs do
   not exist under torch 2.11 / Python 3.12 and are supplied from outside, without
   touching the reference sources: stub modules for ``torch._six`` (util.py:4) and
   ``tensorboardX`` (pytorch_collab.py:10), and pool loaders whose iterators offer the
   Python-2 style ``.next()`` that pytorch_collab.py:78 calls.  A 1-process gloo group
   satisfies ``dist.get_rank()`` (pytorch_collab.py:45).
2. ``aten_*``: known-answer vectors from the installed torch CPU ops the reference
   delegates to (``torch.multinomial`` with and without replacement, ``F.cross_entropy``).

The fixtures are what travels to the GPU box; /root/reference does not.
"""
import os
import sys
import types

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def _install_stubs():
    six = types.ModuleType("torch._six")
    six.int_classes = int
    six.string_classes = (str,)
    sys.modules["torch._six"] = six
    tbx = types.ModuleType("tensorboardX")

    class SummaryWriter:  # never used: fit() is not called
        def __init__(self, *a, **k):
            pass

        def add_scalar(self, *a, **k):
            pass

    tbx.SummaryWriter = SummaryWriter
    sys.modules["tensorboardX"] = tbx
    if REF not in sys.path:
        sys.path.insert(0, REF)


class _Py2Iter:
    """Iterator with the ``.next()`` method pytorch_collab.py:78,81 calls."""

    def __init__(self, batches):
        self._it = iter(batches)

    def next(self):
        return next(self._it)

    __next__ = next

    def __iter__(self):
        return self


class _Loader:
    def __init__(self, batches, batch_size):
        self.batches = batches
        self.batch_size = batch_size

    def __iter__(self):
        return _Py2Iter(self.batches)


class _Recorder(nn.Module):
    """Wraps the model so the logits the reference computed can be stored."""

    def __init__(self, net):
        super().__init__()
        self.net = net
        self.logits = []

    def forward(self, x):
        out = self.net(x)
        self.logits.append(out.detach().clone())
        return out


def synthetic_pool(seed, n_batches=10, batch=32):
    g = torch.Generator().manual_seed(seed)
    batches = []
    for b in range(n_batches):
        data = torch.rand(batch, 3, 32, 32, generator=g)
        label = torch.randint(0, 10, (batch,), generator=g)
        idx = torch.arange(b * batch, (b + 1) * batch) + 1000  # dataset indices, offset to be visible
        batches.append((idx, data, label))
    return batches


def ref_update_samples(net_name, draw_batch, seed):
    import torch.distributed as dist
    from pytorch_collab import Trainer  # reference, unmodified
    from util import EMAverage  # reference, unmodified

    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29517")
        dist.init_process_group("gloo", rank=0, world_size=1)
    torch.manual_seed(seed)
    if net_name == "resnet18":
        from pytorch_model import ResNet18  # reference, unmodified

        net = ResNet18(num_classes=10)
    else:
        net = nn.Sequential(nn.Flatten(), nn.Linear(3072, 10))
    net.train()
    rec = _Recorder(net)
    batches = synthetic_pool(seed + 1)
    trainer = Trainer(rec, None, _Loader([], draw_batch), _Loader(batches, 32), None, "cpu")
    out = {}
    ema = EMAverage()
    torch.manual_seed(seed + 2)  # the generator torch.multinomial draws from
    for call in range(2):  # two consecutive calls share the EMA, as in Trainer.train (:125,:159)
        rec.logits.clear()
        # second call re-iterates the same 10 batches (loader exhausted -> restarted, :79-81)
        probs, data, label, index, avg = trainer.update_samples(ema)
        logits = torch.cat(rec.logits)
        out[f"call{call}_logits"] = logits.numpy()
        out[f"call{call}_probs"] = probs.numpy()
        out[f"call{call}_label"] = label.numpy()
        out[f"call{call}_index"] = index.numpy()
        out[f"call{call}_data_rowsum"] = data.double().sum(dim=(1, 2, 3)).numpy()
        out[f"call{call}_avg"] = np.float32(avg.item())
        out[f"call{call}_ema"] = np.float32(float(ema.value))
    out["pool_labels"] = torch.cat([b[2] for b in batches]).numpy()
    out["pool_index"] = torch.cat([b[0] for b in batches]).numpy()
    out["pool_data_rowsum"] = torch.cat([b[1] for b in batches]).double().sum(dim=(1, 2, 3)).numpy()
    out["pool_seed"] = np.int64(seed + 1)
    out["rng_seed"] = np.int64(seed + 2)
    out["draw_batch"] = np.int64(draw_batch)
    return out


class _SliceDataset:
    """Minimal dataset with the ``get_slice`` util.py:123 calls (util.py:264-273)."""

    def __init__(self, n, seed):
        g = torch.Generator().manual_seed(seed)
        self.data = torch.rand(n, 3, 32, 32, generator=g)
        self.targets = torch.randint(0, 10, (n,), generator=g)

    def __len__(self):
        return self.data.shape[0]

    def get_slice(self, s, e):
        return self.data[s:e], self.targets[s:e]


def ref_groupwise(seed, n=192, n_yield=48):
    from util import Groupwise_Sampler  # reference, unmodified

    torch.manual_seed(seed)
    net = nn.Sequential(nn.Flatten(), nn.Linear(3072, 10))
    rec = _Recorder(net)
    ds = _SliceDataset(n, seed + 1)
    out = {}
    for repl in (True, False):
        s = Groupwise_Sampler(ds, replacement=repl)
        rec.logits.clear()
        with torch.no_grad():
            s.update_importance(0, 64, rec, device="cpu")   # group 1: [0,64)
            s.update_importance(0, 64, rec, device="cpu")   # group 1: [64,128)
            s.update_importance(1, 64, rec, device="cpu")   # group 2: [128,192) -> current group
        torch.manual_seed(seed + 3)
        it = iter(s)
        ys = [int(next(it)) for _ in range(n_yield)]
        tag = "repl" if repl else "norepl"
        out[f"{tag}_yield"] = np.asarray(ys, dtype=np.int64)
        out[f"{tag}_importance"] = s.importance.copy()
        out[f"{tag}_group_indicator"] = s.group_indicator.copy()
        out[f"{tag}_group_index"] = np.int64(s.group_index)
    out["logits"] = torch.cat(rec.logits).numpy()
    out["labels"] = ds.targets.numpy()
    out["rng_seed"] = np.int64(seed + 3)
    return out


def aten_multinomial():
    out = {}
    for n, k, seed in [(17, 8, 1), (320, 32, 2), (5000, 128, 3), (50000, 128, 4)]:
        g = torch.Generator().manual_seed(100 + seed)
        if seed % 2:
            w = torch.rand(n, generator=g) * 3 + 0.01
        else:
            w = torch.exp(torch.randn(n, generator=g) * 2)  # heavy-tailed
        p = w / w.sum()
        g1 = torch.Generator().manual_seed(seed)
        idx_r = torch.multinomial(p, k, True, generator=g1)
        g2 = torch.Generator().manual_seed(seed)
        idx_n = torch.multinomial(p, min(k, n), False, generator=g2)
        out[f"n{n}_p"] = p.numpy()
        out[f"n{n}_seed"] = np.int64(seed)
        out[f"n{n}_with"] = idx_r.numpy()
        out[f"n{n}_without"] = idx_n.numpy()
    return out


def aten_cross_entropy():
    out = {}
    g = torch.Generator().manual_seed(11)
    for name, n, c, dt, sigma in [("c10_f32", 96, 10, torch.float32, 2.0), ("c10_f32_wide", 64, 10, torch.float32, 8.0),
                                  ("c1000_bf16", 24, 1000, torch.bfloat16, 2.0), ("c1000_f32", 8, 1000, torch.float32, 4.0),
                                  ("c32768_bf16", 2, 32768, torch.bfloat16, 2.0), ("c33_f32", 40, 33, torch.float32, 3.0)]:
        x = (torch.randn(n, c, generator=g) * sigma).to(dt)
        y = torch.randint(0, c, (n,), generator=g)
        y[0] = -100  # ignore_index row -> loss 0
        loss = F.cross_entropy(x.float(), y, reduction="none")
        loss64 = F.cross_entropy(x.double(), y, reduction="none")
        out[f"{name}_logits"] = x.view(torch.int16).numpy() if dt == torch.bfloat16 else x.numpy()
        out[f"{name}_labels"] = y.numpy()
        out[f"{name}_loss"] = loss.numpy()
        out[f"{name}_loss64"] = loss64.numpy()
    return out


def ref_head(seed, n=64):
    """The reference's ResNet18 run on synthetic images with its own forward (pytorch_model.py:88-97): the input of its
    ``linear`` layer (the pooled features), that layer's parameters, its output and the scores pytorch_collab.py:102
    computes from it -- the golden vector for the fused head (mercury_score_head)."""
    import torch.nn.functional as F
    from pytorch_model import ResNet18  # reference, unmodified

    torch.manual_seed(seed)
    net = ResNet18(num_classes=10)
    net.train()
    g = torch.Generator().manual_seed(seed + 1)
    data = torch.rand(n, 3, 32, 32, generator=g)
    label = torch.randint(0, 10, (n,), generator=g)
    seen = {}
    hook = net.linear.register_forward_hook(lambda m, i, o: seen.update(features=i[0].detach().clone(), logits=o.detach().clone()))
    with torch.no_grad():
        out = net(data)
    hook.remove()
    assert torch.equal(out, seen["logits"])
    scores = F.cross_entropy(out, label, reduction="none")  # pytorch_collab.py:102
    return dict(features=seen["features"].numpy(), weight=net.linear.weight.detach().numpy(), bias=net.linear.bias.detach().numpy(),
                label=label.numpy(), logits=out.numpy(), scores=scores.numpy(), seed=np.int64(seed))


def ref_partition(seed, n=6000, n_nets=4, alpha=0.5):
    """The reference's own Dirichlet ("hetero") partitioner, cifar10/data_loader.py:126-161, run on synthetic labels: its
    data loader (which would download CIFAR-10) is replaced from outside, the partitioning code runs unmodified."""
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from cifar10 import data_loader as dl

    rng = np.random.RandomState(seed)
    y = rng.randint(0, 10, size=n).astype(np.int64)
    dl.load_cifar10_data = lambda datadir: (np.zeros((n, 1), dtype=np.uint8), y, np.zeros((1, 1), dtype=np.uint8), y[:1])
    np.random.seed(seed)   # (the reference seeds the global generator, pytorch_collab.py:22, 280)
    _, _, _, _, net_map, counts = dl.partition_data("cifar10", "", "hetero", n_nets, alpha)
    out = dict(labels=y, seed=np.int64(seed), n_nets=np.int64(n_nets), alpha=np.float64(alpha))
    for j in range(n_nets):
        out[f"shard{j}"] = np.asarray(net_map[j], dtype=np.int64)
    return out


def main():
    _install_stubs()
    torch.set_num_threads(1)
    if len(sys.argv) > 1 and sys.argv[1] == "partition":   # only the fixture added in round 2
        np.savez_compressed(os.path.join(HERE, "ref_partition.npz"), **ref_partition(900))
        print("ref_partition.npz", os.path.getsize(os.path.join(HERE, "ref_partition.npz")))
        return
    if len(sys.argv) > 1 and sys.argv[1] == "head":   # only the fixture added after the others were committed
        np.savez_compressed(os.path.join(HERE, "ref_head.npz"), **ref_head(800))
        print("ref_head.npz", os.path.getsize(os.path.join(HERE, "ref_head.npz")))
        return
    np.savez_compressed(os.path.join(HERE, "aten_multinomial.npz"), **aten_multinomial())
    np.savez_compressed(os.path.join(HERE, "aten_cross_entropy.npz"), **aten_cross_entropy())
    np.savez_compressed(os.path.join(HERE, "ref_update_samples_linear.npz"), **ref_update_samples("linear", 32, 500))
    np.savez_compressed(os.path.join(HERE, "ref_update_samples_resnet18.npz"), **ref_update_samples("resnet18", 128, 600))
    np.savez_compressed(os.path.join(HERE, "ref_groupwise.npz"), **ref_groupwise(700))
    np.savez_compressed(os.path.join(HERE, "ref_head.npz"), **ref_head(800))
    np.savez_compressed(os.path.join(HERE, "ref_partition.npz"), **ref_partition(900))
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))


if __name__ == "__main__":
    main()
