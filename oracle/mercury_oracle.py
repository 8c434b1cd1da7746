"""CPU oracle for Mercury's importance-sampling hot path.  TEST INFRASTRUCTURE ONLY.

This module restates, on the CPU, the arithmetic the reference performs in
``pytorch_collab.py:89-117`` (score -> pool statistics -> draw -> gather),
``pytorch_collab.py:133-148`` (reweighted loss and its backward seed),
``util.py:200-217`` (``EMAverage``) and ``util.py:140-157`` (the group-wise
single draw).  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the
product package ``mercury_b200`` never does.

Where the arithmetic lives: the reference delegates every operation on the path
to PyTorch (``F.cross_entropy``, ``torch.mean``, ``torch.sum``,
``torch.multinomial``, fancy indexing, ``torch.div``).  PyTorch is a third-party
dependency that the reference does not pin (no requirements file; its use of
``torch._six`` dates it to torch <= 1.12).  The oracle therefore calls the
installed torch (2.11) CPU ops for the "verbatim" functions, and additionally
restates ``torch.multinomial``'s published CPU algorithm
(aten/src/ATen/native/cpu/MultinomialKernel.cpp, ``multinomial_with_replacement``
and the exponential-race fast path in aten/src/ATen/native/TensorRandom.cpp ...
``multinomial``) so that the draw can be driven by caller-supplied uniforms.

Parity pin: the reference ships no tests or golden vectors for this path, so the
reference itself leaves parity unpinned.  The explicit-uniform restatements below
are pinned against ``torch.multinomial`` run in this repository's test-suite
(``tests/test_oracle.py``) and against committed fixtures generated from
``torch.multinomial`` / ``F.cross_entropy`` (``tests/golden/make_golden.py``), and
against outputs of the reference's OWN code run in the build container
(``tests/golden/ref_update_samples_*.npz``, ``ref_groupwise.npz``: ``pytorch_collab.Trainer.update_samples``,
``util.EMAverage``, ``util.Groupwise_Sampler`` and ``pytorch_model.ResNet18`` imported unmodified by the
same script): pinned.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

MAX_CATEGORIES = 1 << 24  # torch.multinomial: "number of categories cannot exceed 2^24"


# --------------------------------------------------------------------------- a1
def score(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """Per-candidate importance score.  pytorch_collab.py:102, util.py:129.

    bf16 logits are upcast first: the CPU bf16 op returns bf16.
    """
    return F.cross_entropy(logits.float(), labels, reduction="none", ignore_index=ignore_index)


def score_head(features: torch.Tensor, weight: torch.Tensor, bias, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """Score from the penultimate features: the reference's classifier head followed by its
    scoring call -- ``self.linear(out)`` pytorch_model.py:96, then pytorch_collab.py:102.  bf16 features are upcast."""
    return score(F.linear(features.float(), weight.float(), None if bias is None else bias.float()), labels, ignore_index)


# --------------------------------------------------------------------------- a4
class EMAverage:
    """util.py:200-217, restated (first update copies, then alpha*old + (1-alpha)*new)."""

    def __init__(self, alpha: float = 0.9):
        self.first_update = True
        self.value = 0
        self.alpha = alpha

    def update(self, value):
        if self.first_update:
            self.value = value
            self.first_update = False
        else:
            self.value = self.alpha * self.value + (1 - self.alpha) * value

    def __str__(self):
        return "{:.6f}".format(float(self.value))


# ---------------------------------------------------------------------- a3,a5,a6
def pool_stats(losses: torch.Tensor, ema: EMAverage, alpha: float = 0.5):
    """pytorch_collab.py:109-112.  Returns (mean_loss, smoothed w, normalised p)."""
    mean = torch.mean(losses)
    ema.update(mean)
    w = losses + alpha * ema.value
    p = w / torch.sum(w)
    return mean, w, p


# --------------------------------------------------------------------------- a7
def serial_cdf_f32(p) -> np.ndarray:
    """ATen's CPU cumulative distribution: ``sum += val`` in fp32, then ``/= sum``.

    MultinomialKernel.cpp (multinomial_with_replacement_apply).  ``np.cumsum`` with
    ``dtype=float32`` performs the same left-to-right fp32 chain; ``torch.cumsum`` does
    not.
    """
    p = np.ascontiguousarray(np.asarray(p, dtype=np.float32))
    c = np.cumsum(p, dtype=np.float32)
    return (c / c[-1]).astype(np.float32)


def serial_chain_f32(p) -> np.ndarray:
    """The un-normalised serial fp32 running sum (what ``serial_cdf_f32`` divides)."""
    return np.cumsum(np.ascontiguousarray(np.asarray(p, dtype=np.float32)), dtype=np.float32)


def check_distribution(p) -> None:
    """The three checks ATen's CPU kernel performs, with its messages."""
    p = np.asarray(p, dtype=np.float32)
    if p.size > MAX_CATEGORIES:
        raise RuntimeError("number of categories cannot exceed 2^24")
    if np.any(~(p >= 0)):  # NaN fails `val >= 0` first, exactly as in ATen's loop
        raise RuntimeError("invalid multinomial distribution (encountering probability entry < 0)")
    if not np.all(np.isfinite(p)):
        raise RuntimeError(
            "invalid multinomial distribution (encountering probability entry = infinity or NaN)")
    if not serial_chain_f32(p)[-1] > 0:
        raise RuntimeError("invalid multinomial distribution (sum of probabilities <= 0)")


def lower_bound(cdf: np.ndarray, u: np.ndarray) -> np.ndarray:
    """ATen's binary search: first slot with ``double(cdf[slot]) >= u``."""
    return np.searchsorted(cdf.astype(np.float64), np.asarray(u, dtype=np.float64), side="left").astype(np.int64)


def draw_with_replacement(p, u) -> np.ndarray:
    """pytorch_collab.py:114 with the uniforms made explicit."""
    check_distribution(p)
    return lower_bound(serial_cdf_f32(p), u)


def draw_with_replacement_f64(p, u) -> np.ndarray:
    """The non-exact "fp64" mode of the draw (no counterpart in the reference, which always takes
    torch.multinomial's serial fp32 sum, pytorch_collab.py:114): running totals in fp64, each
    divided by the fp64 total and rounded to fp32, searched as ATen searches."""
    run = np.cumsum(np.asarray(p, dtype=np.float32).astype(np.float64))
    cdf = (run / run[-1]).astype(np.float32)
    return lower_bound(cdf, u)


def uniforms_like_multinomial(n: int, generator: torch.Generator) -> torch.Tensor:
    """The fp64 uniforms ``torch.multinomial`` consumes from a CPU generator."""
    return torch.rand(n, dtype=torch.float64, generator=generator)


# --------------------------------------------------------------------------- a8
def exp_race_keys(p, u) -> torch.Tensor:
    """Keys of the exponential race: ``p / float32(-log1p(-u))``."""
    p = torch.as_tensor(p, dtype=torch.float32)
    u = torch.as_tensor(u, dtype=torch.float64)
    q = (-torch.log1p(-u)).float()
    return p / q


def draw_without_replacement(p, u, k: int) -> np.ndarray:
    """util.py:150-151 / ``multinomial(replacement=False)``: top-k of the race keys."""
    keys = exp_race_keys(p, u)
    return torch.topk(keys, k).indices.numpy().astype(np.int64)


# ----------------------------------------------------------------------- a9,a10
def gather_reweight(p: torch.Tensor, idx: torch.Tensor, data=None, labels=None, index=None):
    """pytorch_collab.py:116-117: ``probs = p[idx] * N`` and the three gathers."""
    n = p.shape[0]
    nprob = p[idx] * n
    out = [nprob]
    for t in (data, labels, index):
        out.append(None if t is None else t[idx])
    return tuple(out)


def reweighted_loss(losses: torch.Tensor, nprob: torch.Tensor):
    """pytorch_collab.py:137,145: ``mean(loss / (N p))``; also the autograd seed."""
    losses = losses.detach().clone().requires_grad_(True)
    loss = torch.mean(torch.div(losses, nprob))
    loss.backward()
    return loss.detach(), losses.grad.detach()


# -------------------------------------------------------------------------- a11
def update_samples_flat(logits, labels, batch_size, ema: EMAverage, alpha=0.5, *,
                        generator=None, u=None, data=None, index=None):
    """One pass of pytorch_collab.py:102-117 over an already-assembled pool.

    Either ``generator`` (draw through ``torch.multinomial`` itself) or explicit
    uniforms ``u`` (draw through the restated inverse-CDF search).
    Returns dict(scores, mean, p, idx, nprob, data, labels, index).
    """
    losses = score(logits, labels)
    mean, w, p = pool_stats(losses, ema, alpha)
    if u is not None:
        idx = torch.from_numpy(draw_with_replacement(p.numpy(), np.asarray(u)))
    else:
        idx = torch.multinomial(p, batch_size, replacement=True, generator=generator)
    nprob, d, l, ix = gather_reweight(p, idx, data, labels, index)
    return dict(scores=losses, mean=mean, w=w, p=p, idx=idx, nprob=nprob, data=d, labels=l, index=ix)


def update_samples_faithful(net, batches, batch_size, ema: EMAverage, alpha=0.5, *, generator=None,
                            device="cpu"):
    """Verbatim pytorch_collab.py:89-117: 10 growing-pool iterations, EMA updated and a
    draw made every iteration, the last draw returned.

    ``batches`` is an iterator of ``(index, data, label)`` triples (util.py:242-262).
    """
    presam_losses_list, labels, datas, index = [], [], [], []
    cnt = 0
    while cnt < 10:
        cnt += 1
        with torch.no_grad():
            idx, data, label = next(batches)
            data = data.to(device)
            label = label.to(device)
            output = net(data)
            presam_losses = F.cross_entropy(output, label, reduction="none")
            presam_losses_list.append(presam_losses)
            labels.append(label)
            datas.append(data)
            index.append(idx)
        presam_losses = torch.cat(presam_losses_list)
        presam_avg_loss = torch.mean(presam_losses)
        ema.update(presam_avg_loss)
        presam_losses = presam_losses + alpha * ema.value
        losses_pros = presam_losses / torch.sum(presam_losses)
        important_idx = torch.multinomial(losses_pros, batch_size, replacement=True, generator=generator)
    return (losses_pros[important_idx] * presam_losses.size()[0], torch.cat(datas)[important_idx],
            torch.cat(labels)[important_idx], torch.cat(index)[important_idx], presam_avg_loss)


# -------------------------------------------------------------------- a8 / a12
def groupwise_probabilities(importance: np.ndarray) -> torch.Tensor:
    """util.py:146-150: ``imp + mean(imp)`` then normalise in float64, cast to fp32."""
    g = np.asarray(importance, dtype=np.float64)
    g = g + np.mean(g)
    g = g / np.sum(g)
    return torch.Tensor(g)


def groupwise_draw(importance: np.ndarray, u) -> int:
    """util.py:150-151 with n_sample=1: the exponential-race argmax (either ``replacement``)."""
    p = groupwise_probabilities(importance)
    return int(torch.argmax(exp_race_keys(p, u)))


# ---------------------------------------------------------------- sharded (8e)
def sharded_draw_flat(score_shards, u, ema: EMAverage, alpha=0.5):
    """North-star mode A: all-gather the scores, then the flat single-process draw."""
    losses = torch.cat(list(score_shards))
    mean, w, p = pool_stats(losses, ema, alpha)
    return p, torch.from_numpy(draw_with_replacement(p.numpy(), np.asarray(u)))


def sharded_draw_two_level(p_shards, u):
    """North-star mode B: per-shard totals are exchanged, the draw is two-level.

    p_shards: list of fp32 arrays, the shards of the globally normalised p.  First level: serial
    fp32 CDF over the shards' serial-chain totals, searched like ATen (first shard with
    double(cdf) >= u, the last shard catching the rest); second level: the uniform rescaled into
    the shard, u' = (u - lo) / (hi - lo), searched in the shard's own serial fp32 CDF.
    Returns global positions.  (Not index-identical to torch.multinomial on the flat vector.)
    Parity of this function is UNPINNED by the reference: the reference has no global draw (each worker draws from its own
    shard, pytorch_collab.py:282-287); the definition is this repository's, built from two pinned serial-fp32 draws.
    """
    totals = np.asarray([serial_chain_f32(p)[-1] for p in p_shards], dtype=np.float32)
    c = np.cumsum(totals, dtype=np.float32)
    cdf = (c / c[-1]).astype(np.float32).astype(np.float64)
    offsets = np.concatenate([[0], np.cumsum([len(p) for p in p_shards])])
    out = np.empty(len(u), dtype=np.int64)
    for j, uj in enumerate(np.asarray(u, dtype=np.float64)):
        g = 0
        while g < len(p_shards) - 1 and cdf[g] < uj:
            g += 1
        lo, hi = (cdf[g - 1] if g > 0 else 0.0), cdf[g]
        ul = (uj - lo) / (hi - lo) if hi > lo else 0.0
        ul = min(max(ul, 0.0), 0.99999999999999989)
        out[j] = offsets[g] + lower_bound(serial_cdf_f32(p_shards[g]), np.asarray([ul]))[0]
    return out
