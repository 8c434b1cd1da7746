"""Drop-in ``Trainer`` exposing the reference's hot-path surface (pytorch_collab.py:36-249).

Only the importance-sampling path is re-implemented: ``update_samples`` (pytorch_collab.py:89-117)
and the reweighted loss of ``train`` (:133-148) run on the CUDA ops; the model forward/backward,
the optimiser and the evaluation stay stock PyTorch and are kept deliberately thin (the trainer
loop, logging and TensorBoard output of the reference are out of scope).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
import torch.nn.functional as F

from . import _lib, ops
from .pool import ResidentPool
from .sampler import EMAverage, reweighted_mean


def split_head(net):
    """``(features, linear)`` for the reference's ResNets: ``features(x)`` is ResNet.forward up to the pooled activations
    (pytorch_model.py:89-95), ``linear`` the classifier head of :96, so that presample batches can be scored with
    ``ops.score_head(features(x), linear.weight, linear.bias, label)`` without materialising the logits."""
    def features(x):
        out = F.relu(net.bn1(net.conv1(x)))
        out = net.layer4(net.layer3(net.layer2(net.layer1(out))))
        out = F.avg_pool2d(out, 4)
        return out.view(out.size(0), -1)

    return features, net.linear


class _Last(dict):
    __getattr__ = dict.__getitem__


class Trainer(object):
    POOL_BATCHES = 10  # pytorch_collab.py:95

    def __init__(self, net, optimizer, train_loader, presam_loader, test_loader, device, score_from_features: bool = False,
                 ops_module=None, global_draw: bool = False, sharded_kwargs: Optional[dict] = None):
        # score_from_features: presample batches stop before the classifier head (``split_head``) and are scored by
        # ``ops.score_head`` -- the logits of the pool are never materialised; needs a ResNet-shaped net.
        # ops_module: the operator module (default: the CUDA ops; the CPU tests inject an oracle-backed stand-in).
        # global_draw (world > 1): the workers' presample pools form ONE sharded pool (sharded.ShardedPool, mode "totals"); every
        # step draws world * B candidates from it and worker r trains on rows [r*B, (r+1)*B) of the assembled batch.  The
        # reference's workers each draw from their own shard only (pytorch_collab.py:282-287); this is the north star's
        # "global draw" wired into the same loop (my_run / init_processes, :252-292).
        self.ops = ops_module if ops_module is not None else ops
        self.global_draw = global_draw
        self._sharded_kwargs = dict(sharded_kwargs or {})
        self._sp = None
        self.score_from_features = score_from_features
        self._head = split_head(net) if score_from_features else None
        self.net = net
        self.optimizer = optimizer
        self.train_loader = train_loader
        self.presam_loader = presam_loader
        self.test_loader = test_loader
        self.device = torch.device(device)
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world_size = dist.get_world_size() if dist.is_initialized() else 1
        self.epoch = 0
        self.step = 1
        self.next_batch_iter = None
        self.should_compute_importance = True
        self.last = None       # full result of the last update_samples (scores, seed, idx, ...)
        self._pool = None      # ResidentPool: preallocated, written in place (no torch.cat, pytorch_collab.py:108,116-117)
        self._grad_bucket = None
        self.overlap_sampling = True   # run update_samples on a side stream beside average_gradients
        self.graph_sampling = False    # replay the fused sampling call from a CUDA graph (static pool, static uniforms)
        self._graph = None
        self._u_dev = None
        self._side = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" and torch.cuda.is_available() else None

    # ---- pool feeder (pytorch_collab.py:74-82, with next(it) for Python 3)
    def get_next(self):
        if self.next_batch_iter is None:
            self.next_batch_iter = iter(self.presam_loader)
        try:
            b = next(self.next_batch_iter)
        except StopIteration:
            self.next_batch_iter = iter(self.presam_loader)
            b = next(self.next_batch_iter)
        return b

    def average_model(self):  # pytorch_collab.py:84-87
        if self.world_size > 1:
            for p in self.net.parameters():
                dist.all_reduce(p.data, op=dist.ReduceOp.SUM)
                p.data /= float(self.world_size)

    def update_samples(self, ema_loss: EMAverage, alpha=0.5):
        """pytorch_collab.py:89-117: same 5-tuple ``(N*p_i, data, label, index, pool_mean_loss)``.

        The ten presample batches are scored by the model (stock PyTorch, no_grad) straight into a
        device-resident pool (``ResidentPool``: no ``torch.cat``); one fused call then does what
        :102-117 do, including the ten EMA updates on the growing pools and the generator advancing
        by ten draws (the last draw is returned).
        """
        B = self.train_loader.batch_size
        pool = self._pool
        if pool is not None:
            pool.reset()
        for _ in range(self.POOL_BATCHES if self.should_compute_importance else 0):
            with torch.no_grad():
                idx, data, label = self.get_next()
                data = data.to(self.device, non_blocking=True)
                if self._head is not None:
                    features, linear = self._head
                    output = self.ops.score_head(features(data), linear.weight, linear.bias, label.to(self.device, non_blocking=True))
                else:
                    output = self.net(data)
                if pool is None:
                    pool = self._pool = ResidentPool(self.POOL_BATCHES * data.shape[0], None if self._head is not None else output.shape[1],
                                                     tuple(data.shape[1:]), device=self.device, logits_dtype=output.dtype,
                                                     payload_dtype=data.dtype)
                pool.append(idx, data, label, output)
        logits, labels, datas, index = pool.tensors()
        if self.global_draw and self.world_size > 1:
            return self._update_samples_global(ema_loss, alpha, B, logits, labels, datas, index)
        bounds = pool.growing_bounds()
        u = torch.rand(len(bounds) * B, dtype=torch.float64)[(len(bounds) - 1) * B:]
        if self._u_dev is None or self._u_dev.numel() != u.numel():
            self._u_dev = torch.empty(u.numel(), dtype=torch.float64, device=self.device)
            self._graph = None
        self._u_dev.copy_(u, non_blocking=True)
        u = self._u_dev

        def call(first, out=None):
            kw = {} if out is None else {"out": out}
            if self._head is not None:   # the pool holds scores
                return self.ops.update_samples(None, labels, u, ema_loss.state, first, ema_alpha=ema_loss.alpha, smooth_alpha=alpha,
                                               data=datas, index=index, bounds=bounds, scores_in=logits, **kw)
            return self.ops.update_samples(logits, labels, u, ema_loss.state, first, ema_alpha=ema_loss.alpha, smooth_alpha=alpha,
                                           data=datas, index=index, bounds=bounds, **kw)

        graphable = self.graph_sampling and self.device.type == "cuda" and not ema_loss.first_update
        if graphable and self._graph is not None:
            # the pool, the uniforms and the EMA slot are the captured tensors, refilled in place (the slot is the trainer's own,
            # so the graph survives the reference's per-epoch EMAverage, pytorch_collab.py:121)
            g, slot, r = self._graph
            slot.copy_(ema_loss.state)
            g.replay()
            ema_loss.state.copy_(slot)
        elif graphable and self.last is not None:
            # capture on the stream this call runs on (the pool tensors are static: ResidentPool writes them in place)
            out = {k: self.last[k] for k in ("scores", "stats", "idx", "nprob", "seed", "data", "labels", "index") if self.last[k] is not None}
            out["status"] = self.last["status"]
            cur = torch.cuda.current_stream()
            cap = torch.cuda.Stream(device=self.device)
            slot = ema_loss.state.clone()
            user_state, ema_loss.state = ema_loss.state, slot
            cap.wait_stream(cur)
            with torch.cuda.stream(cap):
                call(False, out)             # allocates this stream's workspace outside the capture
            cur.wait_stream(cap)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=cap):
                r = call(False, out)
            ema_loss.state = user_state
            slot.copy_(user_state)           # (the warm-up call above advanced the slot: reset it, then take this step's update)
            g.replay()
            user_state.copy_(slot)
            self._graph = (g, slot, r)
        else:
            r = call(ema_loss.first_update)
        ema_loss.first_update = False
        self.last = r
        return r.nprob, r.data, r.labels, r.index, r.stats[_lib.STAT_MEAN]

    def _update_samples_global(self, ema_loss, alpha, B, logits, labels, datas, index):
        """The global draw: one sharded step over all workers' pools (one EMA update per step on the whole pool's mean --
        the reference's ten updates on growing prefixes have no meaning across shards), world * B draws from rank 0's
        uniform stream, the batch assembled on every rank by one all-reduce, worker r keeps its B rows.  N in ``N * p`` is the
        global pool size; the backward seed is rescaled from 1 / (world * B) to 1 / B."""
        from .sharded import ShardedPool

        G = self.world_size
        if self._sp is None or self._sp.n_local != labels.numel():
            self._sp = ShardedPool(labels.numel(), self.device, ema_alpha=ema_loss.alpha, smooth_alpha=alpha,
                                   **dict({"mode": "totals"}, **self._sharded_kwargs))
        sp = self._sp
        sp.ema_state, sp.ema_first, sp.smooth_alpha = ema_loss.state, ema_loss.first_update, alpha
        u = torch.rand(G * B, dtype=torch.float64).to(self.device)
        dist.broadcast(u, src=0)
        if self._head is not None:          # the pool holds scores already
            sp.local_scores.copy_(logits)
            r = sp.update_samples(None, labels, u, data=datas, index=index, assemble=True)
        else:
            r = sp.update_samples(logits, labels, u, data=datas, index=index, assemble=True)
        ema_loss.first_update = False
        mine = slice(self.rank * B, (self.rank + 1) * B)
        self.last = _Last(scores=r["scores"], stats=r["stats"], idx=r["idx"][mine], nprob=r["nprob"][mine], seed=r["seed"][mine] * float(G),
                          data=r["data"][mine], labels=r["labels"][mine], index=r["index"][mine], status=r["status"])
        return self.last.nprob, self.last.data, self.last.labels, self.last.index, r["stats"][_lib.STAT_MEAN]

    def train(self, max_steps=None):
        """One epoch of pytorch_collab.py:119-199 (timers, printing and periodic evaluation dropped)."""
        presam_ema_loss = EMAverage(device=self.device)  # re-created every epoch, :121
        probs, i_data, i_label, _, _ = self.update_samples(presam_ema_loss)
        seed = self.last.seed
        losses_seen = []
        for n_step, _batch in enumerate(self.train_loader):  # the loader only counts steps (:127)
            output = self.net(i_data)
            losses = F.cross_entropy(output, i_label, reduction="none")      # :133 training side, stock
            loss = reweighted_mean(losses, seed)                             # :137,:145
            self._zero_grads()
            loss.backward()                                                  # :148
            self.should_compute_importance = True
            if self.overlap_sampling and self.device.type == "cuda":
                # the reference's intent (:154-156, commented out there): score + sample the next batch
                # while the gradients are averaged -- sampling on a side stream, the all-reduce on the main one
                main = torch.cuda.current_stream()
                self._side.wait_stream(main)
                with torch.cuda.stream(self._side):
                    probs, i_data, i_label, _, _ = self.update_samples(presam_ema_loss)   # :159, pre-step weights
                self.average_gradients()                                     # :160
                main.wait_stream(self._side)
                for t in (probs, i_data, i_label, self.last.seed):
                    t.record_stream(main)
            else:
                probs, i_data, i_label, _, _ = self.update_samples(presam_ema_loss)   # :159 next batch, pre-step weights
                self.average_gradients()                                     # :160
            seed = self.last.seed
            self.optimizer.step()                                            # :164
            losses_seen.append(loss.detach())
            self.step += 1
            if max_steps is not None and n_step + 1 >= max_steps:
                break
        return torch.stack(losses_seen).mean() if losses_seen else None, None, presam_ema_loss

    def fit(self, epochs, max_steps=None):
        self.average_model()
        for epoch in range(1, epochs + 1):
            self.epoch = epoch
            self.train(max_steps=max_steps)

    def _setup_grad_bucket(self):
        """One persistent flat buffer whose slices ARE the ``.grad`` tensors: the reference flattens the gradients into a
        fresh vector, all-reduces it and unflattens it every step (pytorch_collab.py:236-249, util.py:23-63); here backward
        writes into the bucket directly and the collective works on it in place -- no copy in, no copy out."""
        params = [p for p in self.net.parameters() if p.requires_grad]
        total = sum(p.numel() for p in params)
        self._grad_bucket = torch.zeros(total, dtype=params[0].dtype, device=params[0].device)
        off = 0
        for p in params:
            p.grad = self._grad_bucket[off:off + p.numel()].view_as(p)
            off += p.numel()

    def _zero_grads(self):
        if self.world_size > 1 and self._grad_bucket is None:
            self._setup_grad_bucket()
        if self._grad_bucket is not None:
            self._grad_bucket.zero_()            # (optimizer.zero_grad() would drop the views)
        else:
            self.optimizer.zero_grad()

    def average_gradients(self):
        """pytorch_collab.py:236-249 as ONE in-place collective on the persistent bucket: ncclAllReduce(AVG) (gloo has no AVG:
        SUM, then one division)."""
        if self.world_size == 1:
            return
        if self._grad_bucket is None:           # called outside train(): adopt the existing gradients once
            grads = [p.grad for p in self.net.parameters() if p.grad is not None]
            self._setup_grad_bucket()
            torch._foreach_copy_([p.grad for p in self.net.parameters() if p.requires_grad], grads)
        if dist.get_backend() == "nccl":
            dist.all_reduce(self._grad_bucket, op=dist.ReduceOp.AVG)
        else:
            dist.all_reduce(self._grad_bucket, op=dist.ReduceOp.SUM)
            self._grad_bucket /= self.world_size

    def evaluate(self):
        """pytorch_collab.py:201-234 reduced to the test split: (mean loss, accuracy)."""
        self.net.eval()
        tot, correct, cnt = 0.0, 0, 0
        with torch.no_grad():
            for idx, data, label in self.test_loader:
                data, label = data.to(self.device), label.to(self.device)
                output = self.net(data)
                tot += float(F.cross_entropy(output, label, reduction="sum"))
                correct += int((output.argmax(1) == label).sum())
                cnt += data.shape[0]
        self.net.train()
        return tot / max(cnt, 1), correct / max(cnt, 1)
