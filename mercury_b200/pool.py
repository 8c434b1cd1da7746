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
