"""Host logic of the drop-in Trainer (pool assembly, growing-pool bounds, the slice of the uniform stream that the
reference's tenth draw consumes) on CPU: the operator module is replaced by a stand-in built on the oracle -- tests
may do that, the product default is the CUDA ops.  The same check with the real kernels is in test_gpu_surface.py."""
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

from oracle import mercury_oracle as orc


class OracleOps:
    """``mercury_b200.ops`` as far as ``collab.Trainer.update_samples`` uses it, on the CPU oracle."""

    class Result(dict):
        __getattr__ = dict.__getitem__

    def score_head(self, features, weight, bias, labels):
        return orc.score_head(features, weight, bias, labels)

    def update_samples(self, logits, labels, uniforms, ema_state, ema_first, *, ema_alpha=0.9, smooth_alpha=0.5, data=None,
                       index=None, bounds=None, scores_in=None):
        scores = scores_in if scores_in is not None else orc.score(logits, labels)
        ema = orc.EMAverage(ema_alpha)
        ema.first_update = bool(ema_first)
        ema.value = ema_state[0].clone()
        for nb in (bounds or [scores.numel()]):               # the reference's growing pools (pytorch_collab.py:95-112)
            mean, w, p = orc.pool_stats(scores[:nb], ema, smooth_alpha)
        ema_state[0] = ema.value
        idx = torch.from_numpy(orc.draw_with_replacement(p.numpy(), uniforms.numpy()))
        nprob = p[idx] * p.numel()
        stats = torch.tensor([float(mean), float(ema.value), smooth_alpha * float(ema.value), float(w.sum()), 0, 0, 0, 0])
        return self.Result(scores=scores, stats=stats, idx=idx, nprob=nprob, seed=(1.0 / idx.numel()) / nprob,
                           data=None if data is None else data[idx], labels=labels[idx], index=None if index is None else index[idx],
                           status=torch.zeros(1, dtype=torch.int32))


class ReplayNet(nn.Module):
    """Plays back the logits the reference's net produced, batch by batch (the model is outside the path)."""

    def __init__(self, logits):
        super().__init__()
        self.logits = torch.from_numpy(np.asarray(logits))
        self.pos = 0

    def forward(self, x):
        out = self.logits[self.pos:self.pos + x.shape[0]]
        self.pos += x.shape[0]
        return out


class _Loader:
    def __init__(self, batches, batch_size):
        self.batches, self.batch_size = batches, batch_size

    def __iter__(self):
        return iter(self.batches)


def _pool_batches(seed):
    gen = torch.Generator().manual_seed(seed)
    out = []
    for b in range(10):
        data = torch.rand(32, 3, 32, 32, generator=gen)
        label = torch.randint(0, 10, (32,), generator=gen)
        out.append((torch.arange(b * 32, (b + 1) * 32) + 1000, data, label))
    return out


@pytest.mark.parametrize("name", ["ref_update_samples_linear.npz", "ref_update_samples_resnet18.npz"])
def test_trainer_host_logic_reproduces_the_reference_run(golden_dir, name):
    from mercury_b200.collab import Trainer
    from mercury_b200.sampler import EMAverage

    g = np.load(os.path.join(golden_dir, name))
    batches = _pool_batches(int(g["pool_seed"]))
    net = ReplayNet(np.concatenate([g["call0_logits"], g["call1_logits"]]))
    tr = Trainer(net, None, _Loader([], int(g["draw_batch"])), _Loader(batches, 32), None, "cpu", ops_module=OracleOps())
    ema = EMAverage(device="cpu")
    torch.manual_seed(int(g["rng_seed"]))  # the generator torch.multinomial drew from in the reference run
    for call in range(2):
        probs, data, label, index, avg = tr.update_samples(ema)
        assert np.array_equal(index.numpy(), g[f"call{call}_index"])
        assert np.array_equal(label.numpy(), g[f"call{call}_label"])
        np.testing.assert_allclose(probs.numpy(), g[f"call{call}_probs"], rtol=1e-6)
        np.testing.assert_allclose(data.double().sum(dim=(1, 2, 3)).numpy(), g[f"call{call}_data_rowsum"], rtol=1e-12)
        np.testing.assert_allclose(float(avg), g[f"call{call}_avg"], rtol=1e-6)
        np.testing.assert_allclose(float(ema.value), g[f"call{call}_ema"], rtol=1e-6)


@pytest.mark.skipif(not os.path.exists("/root/reference/pytorch_model.py"), reason="needs the reference checkout")
def test_trainer_scoring_from_features_equals_the_logits_path():
    """score_from_features=True: the pool holds scores from split_head + score_head instead of logits; same step."""
    import importlib.util
    import sys
    spec = importlib.util.spec_from_file_location("ref_pytorch_model", "/root/reference/pytorch_model.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules["ref_pytorch_model"] = mod
    spec.loader.exec_module(mod)
    from mercury_b200.collab import Trainer
    from mercury_b200.sampler import EMAverage

    torch.manual_seed(3)
    net = mod.ResNet18(10).eval()      # (eval: batch-norm statistics frozen, so both passes see the same network)
    batches = _pool_batches(11)
    out = {}
    for mode in (False, True):
        tr = Trainer(net, None, _Loader([], 64), _Loader(batches, 32), None, "cpu", score_from_features=mode, ops_module=OracleOps())
        ema = EMAverage(device="cpu")
        torch.manual_seed(5)
        out[mode] = [tuple(t.clone() for t in tr.update_samples(ema)) for _ in range(2)]
        assert (tr._pool.logits.dim() == 1) == mode                      # scores-only pool on the feature path
    for a, b in zip(out[False], out[True]):
        assert torch.equal(a[3], b[3]) and torch.equal(a[2], b[2])           # same draws (index, label)
        np.testing.assert_allclose(a[0].numpy(), b[0].numpy(), rtol=1e-5)    # N * p
        np.testing.assert_allclose(float(a[4]), float(b[4]), rtol=1e-6)      # pool mean loss
