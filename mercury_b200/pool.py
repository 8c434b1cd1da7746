"""Device-resident candidate pool.

The reference assembles its pool per call by appending per-batch tensors to Python lists and
``torch.cat``-ing them once per loop iteration and three more times for the final gather
(pytorch_collab.py:103-108,116-117) -- O(pool^2) copies and a host round trip per batch.  Here
the pool is a set of preallocated tensors (logits, labels, payload, dataset indices) that
batches are written into in place; the sampling step reads them where they lie.  With
``ring=True`` the pool keeps the most recent ``capacity`` candidates (new batches overwrite the
oldest), otherwise ``reset()`` starts a new pool as the reference does on every call.
"""
from __future__ import annotations

from typing import List, Optional

import torch


class ResidentPool:
    def __init__(self, capacity: int, num_classes: Optional[int], payload_shape=(), device="cuda", logits_dtype=torch.float32,
                 payload_dtype=torch.float32, ring: bool = False):
        self.capacity, self.ring = int(capacity), ring
        dev = torch.device(device)
        # num_classes=None: the pool keeps per-candidate SCORES (fp32 [capacity]) instead of logits -- for callers that
        # score from the features (ops.score_head) and never hold logits
        self.logits = (torch.empty(capacity, dtype=torch.float32, device=dev) if num_classes is None
                       else torch.empty(capacity, num_classes, dtype=logits_dtype, device=dev))
        self.labels = torch.empty(capacity, dtype=torch.int64, device=dev)
        self.index = torch.empty(capacity, dtype=torch.int64, device=dev)
        self.data = torch.empty((capacity,) + tuple(payload_shape), dtype=payload_dtype, device=dev) if payload_shape else None
        self.reset()

    def reset(self) -> None:
        self.size = 0          # valid candidates
        self.cursor = 0        # next write position
        self.bounds: List[int] = []  # pool sizes after each append (the reference's growing pools, :95-110)

    def append(self, index, data, labels, logits) -> None:
        """Write one scored batch into the pool (no concatenation, no reallocation)."""
        m = int(logits.shape[0])
        if m > self.capacity:
            raise ValueError("batch larger than the pool")
        if self.cursor + m > self.capacity:
            if not self.ring:
                raise ValueError("pool is full (capacity %d); reset() it or create it with ring=True" % self.capacity)
            first = self.capacity - self.cursor  # split the batch at the end of the ring
            self._write(self.cursor, first, index[:first], None if data is None else data[:first], labels[:first], logits[:first])
            self._write(0, m - first, index[first:], None if data is None else data[first:], labels[first:], logits[first:])
        else:
            self._write(self.cursor, m, index, data, labels, logits)
        self.cursor = (self.cursor + m) % self.capacity if self.ring else self.cursor + m
        self.size = min(self.capacity, self.size + m)
        self.bounds.append(self.size)

    def _write(self, pos, m, index, data, labels, logits):
        if m == 0:
            return
        dev = self.logits.device
        self.logits[pos:pos + m].copy_(logits, non_blocking=True)
        self.labels[pos:pos + m].copy_(torch.as_tensor(labels).to(dev, non_blocking=True))
        self.index[pos:pos + m].copy_(torch.as_tensor(index).to(dev, non_blocking=True))
        if self.data is not None and data is not None:
            self.data[pos:pos + m].copy_(data, non_blocking=True)

    def tensors(self):
        """(logits, labels, data, index) of the valid candidates, as views of the resident buffers."""
        n = self.size
        return self.logits[:n], self.labels[:n], None if self.data is None else self.data[:n], self.index[:n]

    def growing_bounds(self) -> Optional[List[int]]:
        """Pool sizes for the reference's per-iteration EMA updates; None once the ring has wrapped."""
        if not self.bounds or (len(self.bounds) > 1 and self.bounds[-1] == self.bounds[-2]):
            return None
        return list(self.bounds)


class ChunkRing:
    """Scores a shard whose logits do not fit in HBM (the 2^24 x 32768 bf16 pool is 1.1 TB; at 2 / 4 GPUs a shard
    is 550 / 275 GB): the logits stream through a ring of ``depth`` resident windows -- host -> device copies on a copy
    stream, K1 on the caller's stream, one window of copy always in flight behind the window being scored -- and only the
    4-byte scores stay resident.  The step's CDF / draw / gather then run on the resident scores
    (``ops.update_samples(None, labels, ..., scores_in=scores)``).

    ``windows`` yields ``(logits_host, labels_host)`` pairs (pinned host tensors, at most ``window_rows`` rows each); labels
    are kept resident as well (8 B per candidate) when ``labels_out`` is given."""

    def __init__(self, window_rows: int, num_classes: int, dtype=torch.bfloat16, device="cuda", depth: int = 3, ops_module=None):
        if depth < 2:
            raise ValueError("ChunkRing: depth must be at least 2 (one window copying while one is scored)")
        if ops_module is None:
            from . import ops as ops_module
        self.ops = ops_module
        self.device = torch.device(device)
        self.window_rows, self.depth = int(window_rows), int(depth)
        self.logits = [torch.empty(window_rows, num_classes, dtype=dtype, device=self.device) for _ in range(depth)]
        self.labels = [torch.empty(window_rows, dtype=torch.int64, device=self.device) for _ in range(depth)]
        cuda = self.device.type == "cuda"
        self._copy_stream = torch.cuda.Stream(device=self.device) if cuda else None
        self._ready = [torch.cuda.Event() for _ in range(depth)] if cuda else None      # window copied in
        self._free = [torch.cuda.Event() for _ in range(depth)] if cuda else None       # window scored: may be overwritten
        self._used = [False] * depth
        self.status = None

    def score(self, windows, scores_out: torch.Tensor, labels_out: Optional[torch.Tensor] = None, status=None) -> int:
        """Fills ``scores_out[0:rows]`` window by window; returns the number of rows scored."""
        pos = 0
        cuda = self._copy_stream is not None
        main = torch.cuda.current_stream(self.device) if cuda else None
        for w, (h_logits, h_labels) in enumerate(windows):
            m = h_logits.shape[0]
            if m > self.window_rows:
                raise ValueError("ChunkRing: a window has more rows than the ring's windows")
            if pos + m > scores_out.numel():
                raise ValueError("ChunkRing: more rows than scores_out holds")
            k = w % self.depth
            lg, lb = self.logits[k][:m], self.labels[k][:m]
            if cuda:
                with torch.cuda.stream(self._copy_stream):
                    if self._used[k]:
                        self._copy_stream.wait_event(self._free[k])
                    lg.copy_(h_logits, non_blocking=True)
                    lb.copy_(h_labels, non_blocking=True)
                    self._ready[k].record(self._copy_stream)
                main.wait_event(self._ready[k])
            else:
                lg.copy_(h_logits)
                lb.copy_(h_labels)
            self.ops.score(lg, lb, status=status, out=scores_out[pos:pos + m])
            if labels_out is not None:
                labels_out[pos:pos + m].copy_(lb, non_blocking=True)
            if cuda:
                self._free[k].record(main)
                self._used[k] = True
            pos += m
        return pos
