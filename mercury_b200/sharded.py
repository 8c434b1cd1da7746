"""Candidate pool sharded over the GPUs of one box (one process per GPU, torch.distributed/NCCL).

Reference behaviour: pytorch_collab.py:282-287 hands every worker its own presample shard and
``update_samples`` never communicates (each rank draws from its local pool only).  The
north-star mode implemented here makes the draw GLOBAL:

  1. every rank scores its contiguous shard with K1 straight into its slice of a pool-sized
     score vector (no communication);
  2. ``all_gather_into_tensor`` (NCCL over NVLink/NVSwitch) fills the other slices in place:
     N*4 bytes in total (4 MB at 1 M candidates);
  3. every rank runs the identical statistics + exact CDF + draw on the full vector with the
     same uniforms, so all ranks hold the same global indices / N*p weights / seeds, bit-equal
     to the single-process result on the concatenated pool, with no host round trip;
  4. rows are gathered by their owner (zero elsewhere); ``assemble=True`` SUM-all-reduces the
     B gathered rows so that every rank ends up with the whole batch.

That is ``mode="scores"`` (index-identical to the flat pool, K2 replicated).  ``mode="totals"`` is the north
star read literally and the one that scales: every rank builds the exact chain of its own shard only, the
ranks exchange one fp64 sum of scores and one fp32 chain total each, and the draw is two-level (shard by the
totals, then position inside the owner's exact CDF).  With ``exchange="peer"`` those two scalars travel by
stores into the peers' mapped memory from inside the step's own kernels (``PeerExchange``, CUDA IPC over
NVLink/NVSwitch) and the whole sharded step is one C call; ``exchange="nccl"`` uses two small all-gathers.
See DESIGN.md section 5.

The compute steps go through a small backend object so that the host-side logic (shard
arithmetic, collectives, ownership) can be exercised on CPU with gloo in the tests; the default
backend is the CUDA ops and has no CPU path.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

from . import _lib


class CudaBackend:
    """The product path: hand-written kernels through the C ABI."""

    def __init__(self):
        from . import ops

        self.ops = ops

    def score(self, logits, labels, out):
        self.ops.score(logits, labels, out=out)

    def draw(self, scores_all, uniforms, ema_state, ema_first, ema_alpha, smooth_alpha, out):
        r = self.ops.update_samples(None, None, uniforms, ema_state, ema_first, ema_alpha=ema_alpha,
                                    smooth_alpha=smooth_alpha, scores_in=scores_all, out=out)
        return r.idx, r.nprob, r.seed, r.stats, r.status

    def gather_owned(self, idx, lo, n, data, labels, index, out):
        return self.ops.gather_owned(idx, lo, n, data, labels, index, out=out)

    # ---- mode "totals": shard-local chains, only sums / totals are exchanged
    def sum_scores(self, scores, out):
        self.ops.sum_scores(scores, out=out)

    def local_chain(self, scores, sums_all, n_global, ema_state, ema_first, ema_alpha, smooth_alpha, out):
        empty_u = out.setdefault("_no_draws", torch.empty(0, dtype=torch.float64, device=scores.device))
        r = self.ops.update_samples(None, None, empty_u, ema_state, ema_first, ema_alpha=ema_alpha, smooth_alpha=smooth_alpha,
                                    scores_in=scores, global_sums=sums_all, n_global=n_global, out=out)
        out["stats"] = r.stats
        return r.stats, r.status

    def shard_select(self, totals_all, rank, uniforms, out):
        return self.ops.shard_select(totals_all, rank, uniforms, out=out)

    # ---- the two exchanges done by the kernels themselves over peer memory (PeerExchange)
    def sum_scores_exchange(self, scores, boxes, rank, sums_all, status):
        self.ops.sum_scores_exchange(scores, boxes, rank, sums_all, status)

    def shard_select_exchange(self, my_total, boxes, rank, uniforms, totals_all, status, out):
        return self.ops.shard_select_exchange(my_total, boxes, rank, uniforms, totals_all, status, out=out)

    def sharded_step(self, logits, labels, uniforms, scores_out, peer, n_global, lo, ema_state, ema_first, ema_alpha, smooth_alpha,
                     data, index, out):
        """score -> sum+exchange -> local chain -> total exchange + select -> draw: ONE C call, one PDL chain."""
        return self.ops.update_samples(logits, labels, uniforms, ema_state, ema_first, ema_alpha=ema_alpha, smooth_alpha=smooth_alpha,
                                       data=data, index=index, n_global=n_global, index_base=lo, peer=peer,
                                       out=dict(out, scores=scores_out))

    def local_draw(self, scores, local_u, stats, sums_all, n_global, lo, ema_state, data, labels, index, out, status):
        out = dict(out, stats=stats)
        r = self.ops.update_samples(None, labels, local_u, ema_state, False, scores_in=scores, data=data, index=index,
                                    global_sums=sums_all, n_global=n_global, index_base=lo, draw_only=True, out=out, status=status)
        return r


class PeerExchange:
    """This rank's exchange box plus every peer's box mapped over CUDA IPC (NVLink / NVSwitch peer memory).
    ``boxes`` is the int64 device tensor of G box addresses the exchange kernels take.  Set-up uses one
    all_gather_object of the 64-byte handles; after that the per-step exchange involves no collective."""

    def __init__(self, device, group=None):
        import ctypes

        lib = _lib.load()
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._lib, self._opened = lib, []
        self._box = None
        err = None
        handle_bytes = None
        with torch.cuda.device(device):
            try:
                box = ctypes.c_void_p()
                handle = (ctypes.c_ubyte * _lib.PEER_HANDLE_BYTES)()
                _lib.check(lib.mercury_peer_create(ctypes.byref(box), handle))
                self._box = box.value
                handle_bytes = bytes(handle)
            except Exception as e:  # noqa: BLE001 -- still take part in the collective below, then report
                err = e
            ptrs = [self._box]
            if self.world > 1:
                handles = [None] * self.world
                dist.all_gather_object(handles, handle_bytes, group=group)
                if err is None and any(h is None for h in handles):
                    err = RuntimeError("mercury_b200: a peer could not create its exchange box")
                ptrs = []
                for g, h in enumerate(handles):
                    if err is not None:
                        break
                    if g == self.rank:
                        ptrs.append(self._box)
                        continue
                    try:
                        p = ctypes.c_void_p()
                        _lib.check(lib.mercury_peer_open((ctypes.c_ubyte * _lib.PEER_HANDLE_BYTES).from_buffer_copy(h), ctypes.byref(p)))
                        self._opened.append(p.value)
                        ptrs.append(p.value)
                    except Exception as e:  # noqa: BLE001
                        err = e
            if err is not None:
                self.close()
                raise err
        self.boxes = torch.tensor(ptrs, dtype=torch.int64, device=device)

    def set_wait(self, wait_ms: int):
        """Bound of one wait for a peer's word (default 4 s, or MERCURY_PEER_WAIT_MS at creation)."""
        _lib.check(self._lib.mercury_peer_set_wait(self._box, int(wait_ms), torch.cuda.current_stream().cuda_stream))

    def reset(self, group=None):
        """Collective: after a MERCURY_ST_PEER_TIMEOUT, bring every rank's box back to its initial state (all ranks call
        this; barrier - clear - barrier, so no rank's new words meet a peer's clearing)."""
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=group)
        _lib.check(self._lib.mercury_peer_reset(self._box, torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier(group=group)

    def close(self):
        for p in self._opened:
            self._lib.mercury_peer_close(p)
        self._opened = []
        if self._box:
            self._lib.mercury_peer_destroy(self._box)
            self._box = None


def shard_bounds(n_global: int, world: int, rank: int):
    """Contiguous equal shards in rank order: shard g = [g*N/G, (g+1)*N/G)."""
    if n_global % world:
        raise ValueError("the pool must divide evenly across ranks (all_gather needs equal shards)")
    n_local = n_global // world
    return rank * n_local, n_local


class ShardedPool:
    """mode="scores": all-gather the scores, replicate the exact CDF (index-identical to the flat pool);
    mode="totals": every rank builds the exact CDF of its own shard only, the ranks exchange one sum and
    one total each, and the draw is two-level (shard, then position) -- constant work per rank."""

    def __init__(self, n_local: int, device, group=None, ema_alpha: float = 0.9, smooth_alpha: float = 0.5, backend=None,
                 mode: str = "scores", exchange: str = "nccl", fused: bool = True):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.n_local = n_local
        self.n_global = n_local * self.world
        if mode not in ("scores", "totals"):
            raise ValueError("mode must be 'scores' or 'totals'")
        self.mode = mode
        if (self.n_global if mode == "scores" else n_local) > (1 << 24):
            raise RuntimeError("number of categories cannot exceed 2^24")
        self.lo, _ = shard_bounds(self.n_global, self.world, self.rank)
        self.device = torch.device(device)
        self.backend = backend if backend is not None else CudaBackend()
        self.scores_all = torch.empty(self.n_global, dtype=torch.float32, device=self.device)
        self.ema_state = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.ema_first = True
        self.ema_alpha, self.smooth_alpha = ema_alpha, smooth_alpha
        self._draw_out: dict = {}
        self._gather_out: dict = {}
        self._arena = self._arena_key = None
        self.peer = None
        self._xstatus = None
        if mode == "totals":
            self.scores_all = torch.empty(n_local, dtype=torch.float32, device=self.device)  # (local shard only)
            self.sums_all = torch.zeros(self.world, dtype=torch.float64, device=self.device)
            self.totals_all = torch.zeros(self.world, dtype=torch.float32, device=self.device)
            self.local_u = None
            self._chain_out: dict = {}
            if exchange not in ("nccl", "peer"):
                raise ValueError("exchange must be 'nccl' or 'peer'")
            # "peer": the sums and totals travel by stores into the peers' mapped boxes from inside the kernels
            self.peer = PeerExchange(self.device, group) if exchange == "peer" else None
            self._xstatus = torch.zeros(1, dtype=torch.int32, device=self.device) if self.peer else None
        self.exchange = exchange if mode == "totals" else "nccl"
        # exchange="peer": the status word of every step is copied to pinned host memory behind the step; the NEXT call (or
        # check()) raises if a peer's word timed out -- a timed-out step is poisoned on the device (NaN nprob / seed, idx -1)
        self._host_status = torch.zeros(1, dtype=torch.int32).pin_memory() if self.peer is not None and self.device.type == "cuda" else None
        self._status_event = None
        self._last_xstatus = None
        self.fused = fused     # exchange="peer": the whole step as one C call / PDL chain (False: one call per stage)

    @property
    def local_scores(self) -> torch.Tensor:
        return self.scores_all if self.mode == "totals" else self.scores_all[self.lo:self.lo + self.n_local]

    def _use_arena(self, b, data, labels, index):
        """mode="totals", assemble=True: the step's outputs (idx, labels, index, nprob, seed, data rows) live in ONE buffer, so
        the batch is assembled by a single all-reduce instead of one per tensor.  Every row is zero on all ranks but its
        owner, so the integer SUM of the raw 32-bit words reproduces the owner's bits for every dtype."""
        key = (b, None if data is None else (tuple(data.shape[1:]), data.dtype), labels is not None, index is not None)
        if self._arena_key == key:
            return
        parts = [("idx", torch.int64, (b,)), ("nprob", torch.float32, (b,)), ("seed", torch.float32, (b,))]
        if labels is not None:
            parts.append(("labels", torch.int64, (b,)))
        if index is not None:
            parts.append(("index", torch.int64, (b,)))
        if data is not None:
            parts.append(("data", data.dtype, (b,) + tuple(data.shape[1:])))
        offs, total = [], 0
        for _, dt, shape in parts:
            offs.append(total)
            nbytes = torch.empty(0, dtype=dt).element_size() * int(torch.Size(shape).numel())
            total += (nbytes + 15) // 16 * 16
        self._arena = torch.zeros(max(total, 16), dtype=torch.uint8, device=self.device)
        for (name, dt, shape), off in zip(parts, offs):
            nbytes = torch.empty(0, dtype=dt).element_size() * int(torch.Size(shape).numel())
            self._draw_out[name] = self._arena[off:off + nbytes].view(dt).view(shape)
        self._arena_key = key

    def _assemble(self, tensors):
        lo, hi = (self._arena.data_ptr(), self._arena.data_ptr() + self._arena.numel()) if self._arena is not None else (0, 0)
        live = [t for t in tensors if t is not None]
        if live and all(lo <= t.data_ptr() < hi for t in live):
            dist.all_reduce(self._arena.view(torch.int32), op=dist.ReduceOp.SUM, group=self.group)   # one collective
            return
        for t in live:          # (a backend that allocates its own outputs)
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)

    def _all_gather_small(self, full: torch.Tensor):
        if self.world == 1:
            return
        mine = full[self.rank:self.rank + 1]
        try:
            dist.all_gather_into_tensor(full, mine, group=self.group)
        except (RuntimeError, NotImplementedError):
            parts = list(full.split(1))
            dist.all_gather(parts, mine.clone(), group=self.group)

    def _update_totals(self, logits_local, labels_local, uniforms, data, index, assemble):
        b = self.backend
        if assemble and self.world > 1 and self.device.type == "cuda":
            self._use_arena(uniforms.numel(), data, labels_local, index)
        if logits_local is not None and self.peer is not None and self.fused and hasattr(b, "sharded_step"):
            return self._update_totals_fused(logits_local, labels_local, uniforms, data, index, assemble)
        if logits_local is not None:     # (None: ``local_scores`` is filled already, e.g. by pool.ChunkRing for a streamed shard)
            b.score(logits_local, labels_local, self.local_scores)
        if self.peer is not None:
            b.sum_scores_exchange(self.local_scores, self.peer.boxes, self.rank, self.sums_all, self._xstatus)
        else:
            b.sum_scores(self.local_scores, self.sums_all[self.rank:self.rank + 1])
            self._all_gather_small(self.sums_all)                      # G doubles
        stats, status = b.local_chain(self.local_scores, self.sums_all, self.n_global, self.ema_state, self.ema_first,
                                      self.ema_alpha, self.smooth_alpha, self._chain_out)
        self.ema_first = False
        if self.local_u is None or self.local_u.shape != uniforms.shape:
            self.local_u = torch.empty_like(uniforms)
        if self.peer is not None:
            b.shard_select_exchange(stats[_lib.STAT_CHAIN_TOTAL:_lib.STAT_CHAIN_TOTAL + 1], self.peer.boxes, self.rank, uniforms,
                                    self.totals_all, self._xstatus, self.local_u)
        else:
            self.totals_all[self.rank:self.rank + 1].copy_(stats[_lib.STAT_CHAIN_TOTAL:_lib.STAT_CHAIN_TOTAL + 1])
            self._all_gather_small(self.totals_all)                    # G floats
            b.shard_select(self.totals_all, self.rank, uniforms, self.local_u)
        r = b.local_draw(self.local_scores, self.local_u, stats, self.sums_all, self.n_global, self.lo, self.ema_state, data,
                         labels_local, index, self._draw_out, status)
        self._draw_out.update({k: r[k] for k in ("idx", "nprob", "seed", "data", "labels", "index") if r[k] is not None})
        owned = self.local_u >= 0
        if assemble and self.world > 1:
            self._assemble((r["idx"], r["nprob"], r["seed"], r["data"], r["labels"], r["index"]))
        return dict(idx=r["idx"], nprob=r["nprob"], seed=r["seed"], stats=stats, status=status, data=r["data"], labels=r["labels"],
                    index=r["index"], owned=owned, scores=self.local_scores, xstatus=self._xstatus)

    def _update_totals_fused(self, logits_local, labels_local, uniforms, data, index, assemble):
        if self.local_u is None or self.local_u.shape != uniforms.shape:
            self.local_u = torch.empty_like(uniforms)
        peer = dict(boxes=self.peer.boxes, rank=self.rank, sums_all=self.sums_all, totals_all=self.totals_all, local_u=self.local_u)
        r = self.backend.sharded_step(logits_local, labels_local, uniforms, self.local_scores, peer, self.n_global, self.lo,
                                      self.ema_state, self.ema_first, self.ema_alpha, self.smooth_alpha, data, index, self._draw_out)
        self.ema_first = False
        self._draw_out.update({k: r[k] for k in ("idx", "nprob", "seed", "data", "labels", "index", "stats") if r[k] is not None})
        if assemble and self.world > 1:
            self._assemble((r.idx, r.nprob, r.seed, r.data, r.labels, r.index))
        return dict(idx=r.idx, nprob=r.nprob, seed=r.seed, stats=r.stats, status=r.status, data=r.data, labels=r.labels,
                    index=r.index, owned=self.local_u >= 0, scores=self.local_scores, xstatus=r.status)

    def check(self, wait: bool = True):
        """exchange="peer": raises if the last step's exchange timed out (``wait=False``: only if that step has finished --
        what update_samples does on entry, so a poisoned run stops at the next step instead of training on).  After the
        ranks have synchronised, ``pool.peer.reset()`` on every rank makes the exchange usable again."""
        ev = self._status_event
        if ev is None:
            # (steps replayed from a CUDA graph record no event: an explicit check reads the device word)
            bits = int(self._last_xstatus.item()) if wait and self._last_xstatus is not None and self.peer is not None else 0
        else:
            if wait:
                ev.synchronize()
            elif not ev.query():
                return
            self._status_event = None
            bits = int(self._host_status[0])
        if bits & _lib.ST_PEER_TIMEOUT:
            raise RuntimeError("mercury_b200: a peer's exchange word did not arrive within the bound (ranks more than the "
                               "wait apart? PeerExchange.set_wait / MERCURY_PEER_WAIT_MS; recover with peer.reset() on all ranks)")

    def _all_gather_scores(self):
        if self.world == 1:
            return
        try:
            dist.all_gather_into_tensor(self.scores_all, self.local_scores, group=self.group)
        except (RuntimeError, NotImplementedError):  # backends without the flat variant
            parts = list(self.scores_all.split(self.n_local))
            dist.all_gather(parts, self.local_scores.clone(), group=self.group)

    def update_samples(self, logits_local, labels_local, uniforms, data=None, index=None, assemble: bool = False,
                       out: Optional[dict] = None):
        """One global importance-sampling step (``logits_local=None``: the shard's scores are in ``local_scores``
        already -- a shard streamed through pool.ChunkRing).  Returns a dict with the replicated ``idx`` (global
        positions), ``nprob``, ``seed``, ``stats``, ``status`` and this rank's ``data`` / ``labels`` /
        ``index`` rows (zero where another rank owns the draw) plus the ``owned`` mask."""
        if self.mode == "totals":
            capturing = self.device.type == "cuda" and torch.cuda.is_current_stream_capturing()
            if not capturing:
                self.check(wait=False)
            r = self._update_totals(logits_local, labels_local, uniforms, data, index, assemble)
            self._last_xstatus = r["xstatus"]
            if self._host_status is not None and r["xstatus"] is not None and not capturing:
                self._host_status.copy_(r["xstatus"], non_blocking=True)
                self._status_event = torch.cuda.Event()
                self._status_event.record()
            return r
        b = self.backend
        if logits_local is not None:
            b.score(logits_local, labels_local, self.local_scores)
        self._all_gather_scores()
        idx, nprob, seed, stats, status = b.draw(self.scores_all, uniforms, self.ema_state, self.ema_first, self.ema_alpha,
                                                 self.smooth_alpha, self._draw_out)
        self.ema_first = False
        self._draw_out.update(idx=idx, nprob=nprob, seed=seed, stats=stats)
        d, l, ix, owned = b.gather_owned(idx, self.lo, self.n_local, data, labels_local, index, self._gather_out)
        self._gather_out.update({k: v for k, v in (("data", d), ("labels", l), ("index", ix), ("owned", owned)) if v is not None})
        if assemble and self.world > 1:
            for t in (d, l, ix):
                if t is not None:
                    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return dict(idx=idx, nprob=nprob, seed=seed, stats=stats, status=status, data=d, labels=l, index=ix, owned=owned,
                    scores=self.scores_all)
