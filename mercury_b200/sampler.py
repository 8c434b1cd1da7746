"""Host-side mirror of the reference's sampling interface, over the CUDA ops.

Names, argument meaning and error behaviour follow the reference:
  EMAverage            util.py:200-217
  ImportanceSampler    the arithmetic of Trainer.update_samples, pytorch_collab.py:102-117
  reweighted_mean      pytorch_collab.py:137,145,148 (loss / probs, mean, backward seed)
  Groupwise_Sampler    util.py:94-160
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np
import torch

from . import _lib, ops


class EMAverage:
    """util.py:200-217.  ``value`` lives on the device (no ``.item()`` on the hot path)."""

    def __init__(self, alpha: float = 0.9, device=None):
        self.first_update = True
        self.alpha = alpha
        self.device = torch.device(device if device is not None else "cuda")
        self.state = torch.zeros(1, dtype=torch.float32, device=self.device)

    @property
    def value(self) -> torch.Tensor:
        return self.state[0]

    def update(self, value) -> None:
        """value: number or 0-dim tensor.  Same recurrence as the reference, run by the stats kernel
        on a one-element pool (first update copies, then alpha*old + (1-alpha)*new)."""
        v = torch.as_tensor(value, dtype=torch.float32).reshape(1).to(self.device)
        ops.pool_stats(v, self.state, self.first_update, ema_alpha=self.alpha)
        self.first_update = False

    def __str__(self):
        return "{:.6f}".format(float(self.value))


class _ReweightedMean(torch.autograd.Function):
    """L = mean(loss_i / (N p_i)); dL/dloss_i = seed_i = 1 / (B N p_i), emitted by the gather kernel."""

    @staticmethod
    def forward(ctx, losses, seed):
        ctx.save_for_backward(seed)
        if losses.is_cuda:
            return ops.reweighted_mean(losses.detach(), seed)
        return (losses * seed).sum()      # (CPU tensors: host-side tests of the trainer logic only)

    @staticmethod
    def backward(ctx, grad_out):
        (seed,) = ctx.saved_tensors
        return grad_out * seed, None


def reweighted_mean(losses: torch.Tensor, seed: torch.Tensor) -> torch.Tensor:
    """``torch.mean(torch.div(losses, probs))`` of pytorch_collab.py:137,145 with the backward seed precomputed."""
    return _ReweightedMean.apply(losses, seed)


class ImportanceSampler:
    """Score a pool, smooth with the EMA of its mean loss, draw a batch, return it with its weights.

    ``update_samples`` returns the reference's 5-tuple ``(N*p_i, data, label, index, pool_mean_loss)``
    (pytorch_collab.py:116-117); ``last`` keeps the full result (scores, seed, idx, stats, status).
    """

    def __init__(self, device="cuda", ema_alpha: float = 0.9, smooth_alpha: float = 0.5,
                 generator: Optional[torch.Generator] = None):
        self.device = torch.device(device)
        self.ema = EMAverage(ema_alpha, self.device)
        self.smooth_alpha = smooth_alpha
        self.generator = generator
        self.last = None

    def reset_ema(self) -> None:
        """The reference creates a fresh EMAverage every epoch (pytorch_collab.py:121)."""
        self.ema = EMAverage(self.ema.alpha, self.device)

    def uniforms(self, batch_size: int, n_draws: int = 1) -> torch.Tensor:
        """The fp64 uniforms of the LAST of ``n_draws`` consecutive torch.multinomial calls."""
        u = torch.rand(n_draws * batch_size, dtype=torch.float64, generator=self.generator)
        return u[(n_draws - 1) * batch_size:].to(self.device, non_blocking=True)

    def update_samples(self, logits: torch.Tensor, labels: torch.Tensor, batch_size: int, data=None, index=None,
                       bounds: Optional[Sequence[int]] = None, uniforms: Optional[torch.Tensor] = None,
                       ema: Optional[EMAverage] = None, alpha: Optional[float] = None, check: bool = False):
        ema = ema or self.ema
        if uniforms is None:
            uniforms = self.uniforms(batch_size, len(bounds) if bounds else 1)
        r = ops.update_samples(logits, labels, uniforms, ema.state, ema.first_update, ema_alpha=ema.alpha,
                               smooth_alpha=self.smooth_alpha if alpha is None else alpha, data=data, index=index,
                               bounds=bounds)
        ema.first_update = False
        if check:
            ops.check_status(r.status)
        self.last = r
        return r.nprob, r.data, r.labels, r.index, r.stats[_lib.STAT_MEAN]


class Groupwise_Sampler(torch.utils.data.Sampler):
    """util.py:94-160 with the importance array, the group tags and the draws on the device.

    ``update_importance`` scores a slice of the dataset with the fused CE kernel and stores the losses (no device->host
    copy; util.py:132 syncs).  ``__iter__`` draws by the exponential race over the current group (util.py:144-151:
    ``imp + mean(imp)``, normalise in float64, ``torch.multinomial(.., 1, replacement)``) with ONE kernel sequence per
    block of draws and no compaction: mercury_groupwise_draw ranks the members in place.  Each single draw consumes
    ``n_group`` fp64 uniforms from the CPU generator, exactly as torch.multinomial does in the reference.

    ``prefetch`` single draws are raced per device round trip.  The reference re-reads the importance array before every
    yield; here a change made by ``update_importance`` takes effect at the next block, so ``prefetch=1`` (the default)
    is the reference's timing and larger values trade that for ``prefetch`` times fewer synchronisations.
    """

    def __init__(self, dataset, replacement: bool = True, device="cuda", generator: Optional[torch.Generator] = None,
                 prefetch: int = 1):
        self.dataset = dataset
        self.replacement = replacement
        self.device = torch.device(device)
        n = len(dataset)
        self.group_indicator = torch.zeros(n, dtype=torch.int32, device=self.device)
        self._tags_host = np.zeros(n, dtype=np.int32)   # mirror of the tags: the host wrote them, so it knows the group sizes
        self.importance = torch.ones(n, dtype=torch.float32, device=self.device)
        self.cur_sample_index = 0
        self.group_index = 0
        self.last_update_iteration = -1
        self.generator = generator
        self.prefetch = max(1, int(prefetch))
        self.status = ops.new_status(self.device)
        self.num_samples = n  # the reference's __len__ reads this (undefined there, util.py:159-160)

    def update_importance(self, iteration, update_batchsize, model, device=None):
        if iteration > self.last_update_iteration:
            self.group_index += 1
            self.last_update_iteration = iteration
        start = self.cur_sample_index
        end = min(start + update_batchsize, len(self.dataset))
        data, label = self.dataset.get_slice(start, end)
        data = data.to(self.device)
        label = label.to(self.device)
        with torch.no_grad():
            output = model(data)
        ops.score(output, label, out=self.importance[start:end])  # no device->host copy (util.py:132 syncs)
        self.group_indicator[start:end] = self.group_index
        self._tags_host[start:end] = self.group_index
        self.cur_sample_index = 0 if end == len(self.dataset) else end

    def group_size(self) -> int:
        return int((self._tags_host == self.group_index).sum())

    def group_probabilities(self):
        """p of the current group at the members' own positions (0 elsewhere): (imp + mean(imp)) / sum(...) (util.py:144-147)."""
        count = self.group_size()
        u = torch.full((count,), 0.5, dtype=torch.float64, device=self.device)
        _, p = ops.groupwise_draw(self.importance, self.group_indicator, self.group_index, count, u, want_p=True, status=self.status)
        return p

    def draw(self, k: int = 1, ndraws: int = 1) -> torch.Tensor:
        """Dataset positions from the current group: ``ndraws`` single draws (k == 1), or one race for k positions without
        replacement (k > 1).  Uniforms: ``count`` per draw from the CPU generator, like torch.multinomial."""
        count = self.group_size()
        if count == 0:
            raise RuntimeError("Groupwise_Sampler: the current group is empty (call update_importance first)")
        u = torch.rand(ndraws * count, dtype=torch.float64, generator=self.generator).to(self.device, non_blocking=True)
        return ops.groupwise_draw(self.importance, self.group_indicator, self.group_index, count, u, ndraws=ndraws, k=k,
                                  status=self.status)

    def __iter__(self):
        counter = 0
        total = len(self.dataset)
        while counter < total:
            block = min(self.prefetch, total - counter)
            for i in self.draw(1, block).tolist():   # one synchronisation per block of draws
                yield i
                counter += 1
        ops.check_status(self.status)

    def __len__(self):
        return self.num_samples
