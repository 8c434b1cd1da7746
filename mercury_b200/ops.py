"""Python operators over the C ABI (include/mercury_b200.h): one function per reference call site.

Every function takes CUDA tensors, launches hand-written sm_100a kernels on the current stream
through ctypes and returns CUDA tensors.  Nothing here computes on the CPU and nothing falls
back to stock PyTorch kernels; torch only owns the memory and the stream.

Reference call sites (AIoT-MLSys-Lab/Mercury):
  score                    F.cross_entropy(.., reduction='none')         pytorch_collab.py:102, util.py:129
  pool_stats / normalize   torch.mean, EMAverage.update, +alpha*ema, /sum pytorch_collab.py:109-112
  cdf + sample_with_replacement / multinomial   torch.multinomial(.., True)   pytorch_collab.py:114
  sample_topk / multinomial(replacement=False)  torch.multinomial(.., 1, r)   util.py:150-151
  gather_reweight          p[idx]*N, data[idx], labels[idx], index[idx]   pytorch_collab.py:116-117
  update_samples           the whole of Trainer.update_samples' arithmetic pytorch_collab.py:102-117
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch

from . import _lib
from ._lib import UpdateArgs

MAX_CATEGORIES = 1 << 24
MAX_TOPK = 2048

_workspaces: dict = {}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _cuda(t: torch.Tensor, name: str, dtype=None) -> torch.Tensor:
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"mercury_b200: `{name}` must be a CUDA tensor (there is no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"mercury_b200: `{name}` must have dtype {dtype}, got {t.dtype}")
    return t


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Per-(device, stream) scratch buffer, grown on demand (torch owns the allocation)."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), _stream())
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _workspaces[key] = buf
    return buf


def new_status(device) -> torch.Tensor:
    return torch.zeros(1, dtype=torch.int32, device=device)


def check_status(status: torch.Tensor) -> None:
    """Reads the device status word (synchronises) and raises what torch would have raised."""
    bits = int(status.item())
    if bits & _lib.ST_BAD_LABEL:
        raise IndexError("Target out of bounds.")
    if bits & _lib.ST_NEGATIVE:
        raise RuntimeError("invalid multinomial distribution (encountering probability entry < 0)")
    if bits & _lib.ST_NONFINITE:
        raise RuntimeError("invalid multinomial distribution (encountering probability entry = infinity or NaN)")
    if bits & _lib.ST_ZERO_SUM:
        raise RuntimeError("invalid multinomial distribution (sum of probabilities <= 0)")
    if bits & _lib.ST_PEER_TIMEOUT:
        raise RuntimeError("mercury_b200: a peer's exchange word did not arrive within the bound (ranks out of step? the step's outputs are poisoned)")
    if bits & _lib.ST_INTERNAL:
        raise RuntimeError("mercury_b200: a wait inside the one-kernel step hit its 2 s bound (internal error)")
    if bits & _lib.ST_GROUP_COUNT:
        raise RuntimeError("mercury_b200: the group does not have the member count the host passed (tags changed behind the sampler?)")
    if bits & _lib.ST_TOPK_TIES:
        raise RuntimeError("mercury_b200: too many exactly tied keys at the k-th place of the exponential race")


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return _lib.F32
    if t.dtype == torch.bfloat16:
        return _lib.BF16
    raise RuntimeError(f"mercury_b200: logits must be float32 or bfloat16, got {t.dtype}")


# ----------------------------------------------------------------------------------------- K1
def score(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100,
          status: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """F.cross_entropy(logits, labels, reduction='none') -> fp32 [N]  (pytorch_collab.py:102)."""
    _cuda(logits, "logits")
    _cuda(labels, "labels", torch.int64)
    if logits.dim() != 2 or labels.dim() != 1 or labels.shape[0] != logits.shape[0]:
        raise RuntimeError("mercury_b200.score: expected logits [N, C] and labels [N]")
    if logits.stride(1) != 1:
        logits = logits.contiguous()
    labels = labels.contiguous()
    n, c = logits.shape
    if status is None:
        status = new_status(logits.device)
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=logits.device)
    _lib.check(_lib.load().mercury_score(logits.data_ptr(), _dtype_code(logits), labels.data_ptr(), n, c,
                                         logits.stride(0) if n > 1 else c, ignore_index, out.data_ptr(),
                                         status.data_ptr(), _stream()))
    return out


def score_head(features: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], labels: torch.Tensor,
               ignore_index: int = -100, status: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Scores straight from the penultimate features: ``F.cross_entropy(F.linear(features, weight, bias), labels,
    reduction='none')`` without materialising the logits (pytorch_model.py:96 + pytorch_collab.py:102).
    features [N, D] fp32/bf16 with D a multiple of 128, weight fp32 [C, D] with C <= 16 (the reference's 512 -> 10 head)."""
    _cuda(features, "features")
    _cuda(weight, "weight")
    _cuda(labels, "labels", torch.int64)
    if features.dim() != 2 or weight.dim() != 2 or weight.shape[1] != features.shape[1] or labels.shape != features.shape[:1]:
        raise RuntimeError("mercury_b200.score_head: expected features [N, D], weight [C, D] and labels [N]")
    if weight.shape[0] > 16 or weight.dtype == torch.bfloat16:
        return score_head_wide(features, weight, bias, labels, ignore_index, status, out)
    _cuda(weight, "weight", torch.float32)
    if features.stride(1) != 1:
        features = features.contiguous()
    weight = weight.contiguous()
    if bias is not None:
        bias = _cuda(bias, "bias", torch.float32).contiguous()
    labels = labels.contiguous()
    n, d = features.shape
    if status is None:
        status = new_status(features.device)
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=features.device)
    _lib.check(_lib.load().mercury_score_head(features.data_ptr(), _dtype_code(features), n, d, features.stride(0) if n > 1 else d,
                                              weight.data_ptr(), _ptr(bias), weight.shape[0], labels.data_ptr(), ignore_index,
                                              out.data_ptr(), status.data_ptr(), _stream()))
    return out


def score_head_wide(features: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], labels: torch.Tensor,
                    ignore_index: int = -100, status: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """The wide head on the tensor cores (tcgen05): bf16 features [N, D] x bf16 weight [C, D]^T + bias, cross-entropy in the
    epilogue, logits never written (pytorch_model.py:94-96 for the Bottleneck ResNets' 2048 -> C head, then
    pytorch_collab.py:102).  fp32 inputs are cast to bf16 (a parameter / activation cast, as autocast would do)."""
    _cuda(features, "features")
    _cuda(weight, "weight")
    _cuda(labels, "labels", torch.int64)
    if features.dtype != torch.bfloat16:
        features = features.to(torch.bfloat16)
    if weight.dtype != torch.bfloat16:
        weight = weight.to(torch.bfloat16)
    if features.stride(1) != 1 or features.stride(0) % 8 or features.data_ptr() % 16:
        features = features.contiguous()
    if weight.stride(1) != 1 or weight.stride(0) % 8 or weight.data_ptr() % 16:
        weight = weight.contiguous()
    n, d = features.shape
    if d % 8:
        raise RuntimeError("mercury_b200.score_head_wide: the feature width must be a multiple of 8")
    if bias is not None:
        bias = _cuda(bias, "bias").float().contiguous()
    labels = labels.contiguous()
    if status is None:
        status = new_status(features.device)
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=features.device)
    _lib.check(_lib.load().mercury_score_head_wide(features.data_ptr(), n, d, features.stride(0) if n > 1 else d, weight.data_ptr(),
                                                   weight.stride(0) if weight.shape[0] > 1 else d, _ptr(bias), weight.shape[0],
                                                   labels.data_ptr(), ignore_index, out.data_ptr(), status.data_ptr(), _stream()))
    return out


# -------------------------------------------------------------------------------------- stats
def _bounds_array(bounds: Optional[Sequence[int]]):
    if not bounds:
        return None, 0
    arr = (ctypes.c_int64 * len(bounds))(*[int(b) for b in bounds])
    return arr, len(bounds)


def pool_stats(scores: torch.Tensor, ema_state: torch.Tensor, ema_first: bool, ema_alpha: float = 0.9,
               smooth_alpha: float = 0.5, bounds: Optional[Sequence[int]] = None) -> torch.Tensor:
    """mean / EMA / shift / total of a pool of scores (pytorch_collab.py:109-112).

    Returns the 8-float statistics block (see MERCURY_STAT_*) and updates ``ema_state`` in place.
    """
    _cuda(scores, "scores", torch.float32)
    _cuda(ema_state, "ema_state", torch.float32)
    n = scores.numel()
    lib = _lib.load()
    stats = torch.empty(_lib.STAT_COUNT, dtype=torch.float32, device=scores.device)
    ws = workspace(scores.device, lib.mercury_pool_stats_workspace_bytes(max(n, 1)))
    arr, nb = _bounds_array(bounds)
    _lib.check(lib.mercury_pool_stats(scores.data_ptr(), n, arr, nb, ema_state.data_ptr(), int(bool(ema_first)),
                                      float(ema_alpha), float(smooth_alpha), stats.data_ptr(), ws.data_ptr(),
                                      ws.numel(), _stream()))
    return stats


def normalize(scores: torch.Tensor, stats: torch.Tensor) -> torch.Tensor:
    """losses_pros = (scores + alpha*ema) / sum(...)  (pytorch_collab.py:111-112)."""
    _cuda(scores, "scores", torch.float32)
    _cuda(stats, "stats", torch.float32)
    p = torch.empty_like(scores)
    _lib.check(_lib.load().mercury_normalize(scores.data_ptr(), scores.numel(), stats.data_ptr(), p.data_ptr(), _stream()))
    return p


# ----------------------------------------------------------------------------------------- K2
def cdf(weights: torch.Tensor, mode: str = "exact", status: Optional[torch.Tensor] = None):
    """Normalised cumulative distribution of ``weights`` as ATen's CPU multinomial builds it.

    mode='exact': bit-identical to the serial fp32 chain; mode='f64': deterministic fp64 scan.
    Returns (cdf fp32 [N], total fp32 [1]).
    """
    _cuda(weights, "weights", torch.float32)
    weights = weights.contiguous()
    n = weights.numel()
    lib = _lib.load()
    if status is None:
        status = new_status(weights.device)
    out = torch.empty_like(weights)
    total = torch.empty(1, dtype=torch.float32, device=weights.device)
    ws = workspace(weights.device, lib.mercury_cdf_workspace_bytes(max(n, 1)))
    m = {"exact": _lib.CDF_EXACT_SERIAL_F32, "f64": _lib.CDF_F64}[mode]
    _lib.check(lib.mercury_cdf(weights.data_ptr(), n, m, out.data_ptr(), total.data_ptr(), status.data_ptr(),
                               ws.data_ptr(), ws.numel(), _stream()))
    return out, total


def sample_with_replacement(cdf_: torch.Tensor, uniforms: torch.Tensor) -> torch.Tensor:
    """For every fp64 uniform the first slot with double(cdf[slot]) >= u (pytorch_collab.py:114)."""
    _cuda(cdf_, "cdf", torch.float32)
    _cuda(uniforms, "uniforms", torch.float64)
    b = uniforms.numel()
    idx = torch.empty(b, dtype=torch.int64, device=cdf_.device)
    _lib.check(_lib.load().mercury_sample_with_replacement(cdf_.data_ptr(), cdf_.numel(), uniforms.contiguous().data_ptr(),
                                                           b, idx.data_ptr(), _stream()))
    return idx


def sample_topk(weights: torch.Tensor, uniforms: torch.Tensor, k: int, status: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Exponential-race top-k: indices of the k largest w_i / float(-log1p(-u_i)), descending (util.py:150)."""
    _cuda(weights, "weights", torch.float32)
    _cuda(uniforms, "uniforms", torch.float64)
    n = weights.numel()
    if uniforms.numel() != n:
        raise RuntimeError("mercury_b200.sample_topk: one uniform per candidate is required")
    lib = _lib.load()
    if status is None:
        status = new_status(weights.device)
    idx = torch.empty(k, dtype=torch.int64, device=weights.device)
    ws = workspace(weights.device, lib.mercury_sample_topk_workspace_bytes(max(n, 1), k))
    _lib.check(lib.mercury_sample_topk(weights.contiguous().data_ptr(), n, uniforms.contiguous().data_ptr(), k, idx.data_ptr(),
                                       status.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return idx


def multinomial(p: torch.Tensor, num_samples: int, replacement: bool = False, *, generator: Optional[torch.Generator] = None,
                uniforms: Optional[torch.Tensor] = None, check: bool = True) -> torch.Tensor:
    """Drop-in for ``torch.multinomial`` on a 1-D CUDA probability vector.

    The uniforms are the fp64 stream ``torch.multinomial`` would consume from the CPU
    ``generator`` (default generator if None), so for identical ``p`` bits the result equals
    ``torch.multinomial(p.cpu(), ...)`` index for index -- or pass device ``uniforms`` directly.
    """
    _cuda(p, "p", torch.float32)
    if p.dim() != 1:
        raise RuntimeError("mercury_b200.multinomial: 1-D distributions only")
    n = p.numel()
    if n > MAX_CATEGORIES:
        raise RuntimeError("number of categories cannot exceed 2^24")
    race = (not replacement) or num_samples == 1
    if not replacement and num_samples > n:
        raise RuntimeError("cannot sample n_sample > prob_dist.size(-1) samples without replacement")
    status = new_status(p.device)
    if uniforms is None:
        need = n if race else num_samples
        uniforms = torch.rand(need, dtype=torch.float64, generator=generator).to(p.device, non_blocking=True)
    if race:
        # ATen validates the distribution before the race as well
        c, _ = cdf(p, "f64", status) if check else (None, None)
        idx = sample_topk(p, uniforms, num_samples, status)
    else:
        c, _ = cdf(p, "exact", status)
        idx = sample_with_replacement(c, uniforms)
    if check:
        check_status(status)
    return idx


def groupwise_draw(importance: torch.Tensor, tags: torch.Tensor, group: int, count: int, uniforms: torch.Tensor, ndraws: int = 1,
                   k: int = 1, want_p: bool = False, status: Optional[torch.Tensor] = None):
    """Draws from the current group of a Groupwise_Sampler (util.py:140-157) without compacting it: ``ndraws`` single draws
    (k == 1; draw j races the group's members, in position order, with uniforms[j*count:(j+1)*count]) or one draw of k
    without replacement.  Returns dataset positions int64 [ndraws or k] (and p fp32 [n], 0 outside the group, if want_p)."""
    _cuda(importance, "importance", torch.float32)
    _cuda(tags, "tags", torch.int32)
    _cuda(uniforms, "uniforms", torch.float64)
    n = importance.numel()
    if uniforms.numel() < ndraws * count:
        raise RuntimeError("mercury_b200.groupwise_draw: one uniform per member and draw is required")
    lib = _lib.load()
    if status is None:
        status = new_status(importance.device)
    idx = torch.empty(ndraws if k == 1 else k, dtype=torch.int64, device=importance.device)
    p = torch.empty(n, dtype=torch.float32, device=importance.device) if want_p else None
    ws = workspace(importance.device, lib.mercury_groupwise_draw_workspace_bytes(n, ndraws))
    _lib.check(lib.mercury_groupwise_draw(importance.contiguous().data_ptr(), tags.contiguous().data_ptr(), int(group), n, int(count),
                                          uniforms.contiguous().data_ptr(), int(ndraws), int(k), idx.data_ptr(), _ptr(p),
                                          status.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return (idx, p) if want_p else idx


# ----------------------------------------------------------------------------------------- K3
def _row_bytes(data: torch.Tensor) -> int:
    n = data.shape[0]
    return (data.numel() // max(n, 1)) * data.element_size()


def gather_reweight(p: torch.Tensor, idx: torch.Tensor, data: Optional[torch.Tensor] = None,
                    labels: Optional[torch.Tensor] = None, index: Optional[torch.Tensor] = None):
    """(N*p[idx], seed, data[idx], labels[idx], index[idx])  (pytorch_collab.py:116-117,137,145)."""
    _cuda(p, "p", torch.float32)
    _cuda(idx, "idx", torch.int64)
    b, n = idx.numel(), p.numel()
    dev = p.device
    nprob = torch.empty(b, dtype=torch.float32, device=dev)
    seed = torch.empty(b, dtype=torch.float32, device=dev)
    out_data = out_labels = out_index = None
    row_bytes = 0
    if data is not None:
        data = _cuda(data, "data").contiguous()
        row_bytes = _row_bytes(data)
        out_data = torch.empty((b,) + tuple(data.shape[1:]), dtype=data.dtype, device=dev)
    if labels is not None:
        labels = _cuda(labels, "labels", torch.int64).contiguous()
        out_labels = torch.empty(b, dtype=torch.int64, device=dev)
    if index is not None:
        index = _cuda(index, "index", torch.int64).contiguous()
        out_index = torch.empty(b, dtype=torch.int64, device=dev)
    _lib.check(_lib.load().mercury_gather_reweight(p.data_ptr(), idx.data_ptr(), b, n, _ptr(data), row_bytes, _ptr(labels),
                                                   _ptr(index), _ptr(out_data), _ptr(out_labels), _ptr(out_index),
                                                   nprob.data_ptr(), seed.data_ptr(), _stream()))
    return nprob, seed, out_data, out_labels, out_index


def reweighted_mean(losses: torch.Tensor, seed: torch.Tensor) -> torch.Tensor:
    """sum_j losses_j * seed_j = mean(losses / (N p)) (pytorch_collab.py:137,145), 0-dim fp32."""
    _cuda(losses, "losses", torch.float32)
    _cuda(seed, "seed", torch.float32)
    if losses.numel() != seed.numel():
        raise ValueError("losses and seed must have the same number of elements")
    out = torch.empty(1, dtype=torch.float32, device=losses.device)
    _lib.check(_lib.load().mercury_reweighted_mean(losses.contiguous().data_ptr(), seed.contiguous().data_ptr(), losses.numel(),
                                                   out.data_ptr(), _stream()))
    return out[0]


def gather_owned(idx: torch.Tensor, shard_lo: int, shard_n: int, data: Optional[torch.Tensor] = None,
                 labels: Optional[torch.Tensor] = None, index: Optional[torch.Tensor] = None, out: Optional[dict] = None):
    """Rows of the draws this rank owns (global positions ``idx``), zero rows elsewhere; returns
    (data, labels, index, owned mask int32).  A SUM all-reduce over ranks then assembles the batch."""
    _cuda(idx, "idx", torch.int64)
    b, dev = idx.numel(), idx.device
    out = out or {}

    def buf(key, shape, dtype):
        t = out.get(key)
        return t if t is not None else torch.empty(shape, dtype=dtype, device=dev)

    out_data = out_labels = out_index = None
    row_bytes = 0
    if data is not None:
        data = _cuda(data, "data").contiguous()
        row_bytes = _row_bytes(data)
        out_data = buf("data", (b,) + tuple(data.shape[1:]), data.dtype)
    if labels is not None:
        labels = _cuda(labels, "labels", torch.int64).contiguous()
        out_labels = buf("labels", (b,), torch.int64)
    if index is not None:
        index = _cuda(index, "index", torch.int64).contiguous()
        out_index = buf("index", (b,), torch.int64)
    owned = buf("owned", (b,), torch.int32)
    _lib.check(_lib.load().mercury_gather_owned(idx.data_ptr(), b, shard_lo, shard_n, _ptr(data), row_bytes, _ptr(labels),
                                                _ptr(index), _ptr(out_data), _ptr(out_labels), _ptr(out_index),
                                                owned.data_ptr(), _stream()))
    return out_data, out_labels, out_index, owned


# ------------------------------------------------------------------------------ the whole step
class UpdateResult(dict):
    """scores, stats, idx, nprob, seed, data, labels, index, status (all CUDA tensors)."""

    __getattr__ = dict.__getitem__


def update_samples(logits: Optional[torch.Tensor], labels: Optional[torch.Tensor], uniforms: torch.Tensor,
                   ema_state: torch.Tensor, ema_first: bool, *, ema_alpha: float = 0.9, smooth_alpha: float = 0.5,
                   data: Optional[torch.Tensor] = None, index: Optional[torch.Tensor] = None,
                   scores_in: Optional[torch.Tensor] = None, bounds: Optional[Sequence[int]] = None,
                   ignore_index: int = -100, status: Optional[torch.Tensor] = None, out: Optional[dict] = None,
                   global_sums: Optional[torch.Tensor] = None, n_global: Optional[int] = None, index_base: int = 0,
                   draw_only: bool = False, peer: Optional[dict] = None, cdf_mode: str = "exact") -> UpdateResult:
    """Score -> statistics -> exact CDF -> B draws -> gather + weights + seed, in one C-ABI call.

    The arithmetic of Trainer.update_samples (pytorch_collab.py:102-117) on an assembled pool.
    ``uniforms`` are the fp64 numbers torch.multinomial would draw.  ``out`` may carry
    preallocated result tensors (keys as in the result) to keep the call allocation-free.
    ``cdf_mode="exact"`` (default): indices bit-identical to ``torch.multinomial`` on the CPU for the same p and
    uniforms; ``"f64"``: fp64 running totals, no serial walk -- the same distribution, not the same indices.
    """
    lib = _lib.load()
    _cuda(uniforms, "uniforms", torch.float64)
    _cuda(ema_state, "ema_state", torch.float32)
    dev = uniforms.device
    out = out or {}
    a = UpdateArgs()
    if scores_in is None:
        _cuda(logits, "logits")
        _cuda(labels, "labels", torch.int64)
        if logits.stride(1) != 1:
            logits = logits.contiguous()
        n, c = logits.shape
        a.logits, a.dtype, a.c = logits.data_ptr(), _dtype_code(logits), c
        a.row_stride = logits.stride(0) if n > 1 else c
        scores = out.get("scores")
        if scores is None:
            scores = torch.empty(n, dtype=torch.float32, device=dev)
        a.scores = scores.data_ptr()
    else:
        scores = _cuda(scores_in, "scores_in", torch.float32).contiguous()
        n = scores.numel()
        a.scores_in = scores.data_ptr()
    if labels is not None:
        labels = _cuda(labels, "labels", torch.int64).contiguous()
        a.labels = labels.data_ptr()
    b = uniforms.numel()
    a.n, a.ignore_index, a.b = n, ignore_index, b
    arr, nb = _bounds_array(bounds)
    a.h_bounds, a.n_bounds = arr, nb
    a.ema_state, a.ema_first, a.ema_alpha, a.smooth_alpha = ema_state.data_ptr(), int(bool(ema_first)), ema_alpha, smooth_alpha
    a.uniforms = uniforms.contiguous().data_ptr()
    if status is None:
        status = new_status(dev)
    a.status = status.data_ptr()

    def buf(key, shape, dtype):
        t = out.get(key)
        if t is None:
            t = torch.empty(shape, dtype=dtype, device=dev)
        return t

    stats = buf("stats", (_lib.STAT_COUNT,), torch.float32)
    idx = buf("idx", (b,), torch.int64)
    nprob = buf("nprob", (b,), torch.float32)
    seed = buf("seed", (b,), torch.float32)
    a.stats, a.idx, a.nprob, a.seed = stats.data_ptr(), idx.data_ptr(), nprob.data_ptr(), seed.data_ptr()
    out_data = out_labels = out_index = None
    if data is not None:
        data = _cuda(data, "data").contiguous()
        a.data, a.row_bytes = data.data_ptr(), _row_bytes(data)
        out_data = buf("data", (b,) + tuple(data.shape[1:]), data.dtype)
        a.out_data = out_data.data_ptr()
    if labels is not None:
        out_labels = buf("labels", (b,), torch.int64)
        a.out_labels = out_labels.data_ptr()
    if index is not None:
        index = _cuda(index, "index", torch.int64).contiguous()
        a.index = index.data_ptr()
        out_index = buf("index", (b,), torch.int64)
        a.out_index = out_index.data_ptr()
    if global_sums is not None:  # one shard of a larger pool (see sharded.ShardedPool, mode "totals")
        _cuda(global_sums, "global_sums", torch.float64)
        a.global_sums, a.n_global_sums = global_sums.data_ptr(), global_sums.numel()
        a.n_global, a.index_base = int(n_global), int(index_base)
    a.draw_only = int(bool(draw_only))
    a.cdf_mode = {"exact": _lib.CDF_EXACT_SERIAL_F32, "f64": _lib.CDF_F64}[cdf_mode]
    if peer is not None:  # the whole sharded step in this one call (sharded.ShardedPool, mode "totals", exchange "peer")
        boxes = _cuda(peer["boxes"], "peer boxes", torch.int64)
        a.peer_boxes, a.n_shards, a.shard_rank = boxes.data_ptr(), boxes.numel(), int(peer["rank"])
        a.sums_all = _cuda(peer["sums_all"], "sums_all", torch.float64).data_ptr()
        a.totals_all = _cuda(peer["totals_all"], "totals_all", torch.float32).data_ptr()
        a.local_uniforms = _cuda(peer["local_u"], "local_u", torch.float64).data_ptr()
        a.n_global, a.index_base = int(n_global), int(index_base)
    ws = workspace(dev, lib.mercury_update_samples_workspace_bytes(n, b))
    _lib.check(lib.mercury_update_samples(ctypes.byref(a), ws.data_ptr(), ws.numel(), _stream()))
    return UpdateResult(scores=scores, stats=stats, idx=idx, nprob=nprob, seed=seed, data=out_data, labels=out_labels,
                        index=out_index, status=status)


def set_fused(mode: int) -> None:
    """Test hook: 0 = always the four-kernel chain, 1 = the one-kernel step where it applies, -1 = environment."""
    _lib.load().mercury_test_set_fused(int(mode))


def step_times(n: int, device=None) -> list:
    """Developer aid: the stamps the last one-kernel step on the current stream left behind (mercury_debug_step_times)."""
    lib = _lib.load()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    ws = workspace(dev, lib.mercury_update_samples_workspace_bytes(n, 0))
    arr = (ctypes.c_ulonglong * 128)()
    _lib.check(lib.mercury_debug_step_times(ws.data_ptr(), n, arr))
    return list(arr)


def sum_scores(scores: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Deterministic fp64 sum of a shard's scores (one double), for the all-gather of shard sums."""
    _cuda(scores, "scores", torch.float32)
    lib = _lib.load()
    if out is None:
        out = torch.empty(1, dtype=torch.float64, device=scores.device)
    ws = workspace(scores.device, lib.mercury_update_samples_workspace_bytes(max(scores.numel(), 1), 0))
    _lib.check(lib.mercury_sum_scores(scores.contiguous().data_ptr(), scores.numel(), out.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return out


def shard_select(shard_totals: torch.Tensor, rank: int, uniforms: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """First level of the two-level draw: for every uniform, its position inside THIS rank's shard CDF
    (fp64 in [0,1)) if it falls in this shard, else -1."""
    _cuda(shard_totals, "shard_totals", torch.float32)
    _cuda(uniforms, "uniforms", torch.float64)
    if out is None:
        out = torch.empty_like(uniforms)
    _lib.check(_lib.load().mercury_shard_select(shard_totals.data_ptr(), shard_totals.numel(), int(rank), uniforms.data_ptr(),
                                                uniforms.numel(), out.data_ptr(), _stream()))
    return out


def sum_scores_exchange(scores: torch.Tensor, peer_boxes: torch.Tensor, rank: int, sums_all: torch.Tensor,
                        status: torch.Tensor) -> torch.Tensor:
    """``sum_scores`` plus the all-gather of the shard sums, done by the kernel through the peers' exchange
    boxes (``peer_boxes``: int64 device tensor of G box addresses, see sharded.PeerExchange)."""
    _cuda(scores, "scores", torch.float32)
    _cuda(sums_all, "sums_all", torch.float64)
    _cuda(peer_boxes, "peer_boxes", torch.int64)
    lib = _lib.load()
    ws = workspace(scores.device, lib.mercury_update_samples_workspace_bytes(max(scores.numel(), 1), 0))
    _lib.check(lib.mercury_sum_scores_exchange(scores.contiguous().data_ptr(), scores.numel(), peer_boxes.data_ptr(), peer_boxes.numel(),
                                               int(rank), sums_all.data_ptr(), status.data_ptr(), ws.data_ptr(), ws.numel(), _stream()))
    return sums_all


def shard_select_exchange(my_total: torch.Tensor, peer_boxes: torch.Tensor, rank: int, uniforms: torch.Tensor, totals_all: torch.Tensor,
                          status: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """All-gather of the shard chain totals through the exchange boxes, then ``shard_select``, in one kernel."""
    _cuda(my_total, "my_total", torch.float32)
    _cuda(totals_all, "totals_all", torch.float32)
    _cuda(uniforms, "uniforms", torch.float64)
    _cuda(peer_boxes, "peer_boxes", torch.int64)
    if out is None:
        out = torch.empty_like(uniforms)
    _lib.check(_lib.load().mercury_shard_select_exchange(my_total.data_ptr(), peer_boxes.data_ptr(), peer_boxes.numel(), int(rank),
                                                         uniforms.data_ptr(), uniforms.numel(), totals_all.data_ptr(), out.data_ptr(),
                                                         status.data_ptr(), _stream()))
    return out


class GraphedUpdate:
    """One importance-sampling step captured in a CUDA graph and replayed with no host work.

    The four kernels of ``update_samples`` take ~50 us at a million candidates while a Python
    call of ``update_samples`` costs about twice that on the host; a training loop that calls
    it every step should capture it once.  The tensors passed here are STATIC: refill them in
    place (``logits.copy_(...)``, ``uniforms.copy_(...)``) and call ``replay()``; the results
    are the tensors of ``self.result`` (overwritten by every replay).  ``ema_first`` is baked in
    at capture time: let one eager call initialise the EMA first, as ``__init__`` does.
    """

    def __init__(self, logits, labels, uniforms, ema_state, *, ema_alpha: float = 0.9, smooth_alpha: float = 0.5,
                 data=None, index=None, ema_first: bool = True, ignore_index: int = -100):
        self.status = new_status(uniforms.device)
        kw = dict(ema_alpha=ema_alpha, smooth_alpha=smooth_alpha, data=data, index=index, ignore_index=ignore_index,
                  status=self.status)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):  # warm-up on the capture stream: allocates its workspace, consumes ema_first
            r = update_samples(logits, labels, uniforms, ema_state, ema_first, **kw)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        out = {k: r[k] for k in ("scores", "stats", "idx", "nprob", "seed", "data", "labels", "index") if r[k] is not None}
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, stream=side):
            self.result = update_samples(logits, labels, uniforms, ema_state, False, out=out, **kw)

    def replay(self) -> UpdateResult:
        self.graph.replay()
        return self.result
