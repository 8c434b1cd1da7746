"""ResidentPool (device-resident pool feeder): host logic, runs on CPU tensors."""
import pytest
import torch

from mercury_b200.pool import ResidentPool


def _batch(k, m=4, c=3):
    g = torch.Generator().manual_seed(k)
    return torch.arange(m) + 100 * k, torch.rand(m, 2, generator=g), torch.randint(0, c, (m,), generator=g), torch.randn(m, c, generator=g)


def test_append_equals_concatenation():
    pool = ResidentPool(12, 3, (2,), device="cpu")
    bs = [_batch(k) for k in range(3)]
    for b in bs:
        pool.append(*b)
    logits, labels, data, index = pool.tensors()
    assert torch.equal(index, torch.cat([b[0] for b in bs])) and torch.equal(data, torch.cat([b[1] for b in bs]))
    assert torch.equal(labels, torch.cat([b[2] for b in bs])) and torch.equal(logits, torch.cat([b[3] for b in bs]))
    assert pool.growing_bounds() == [4, 8, 12]   # the reference's growing pools (pytorch_collab.py:95-110)
    with pytest.raises(ValueError):
        pool.append(*_batch(9))
    pool.reset()
    assert pool.size == 0 and pool.tensors()[0].shape[0] == 0


def test_ring_keeps_the_most_recent_candidates():
    pool = ResidentPool(10, 3, (2,), device="cpu", ring=True)
    bs = [_batch(k) for k in range(4)]  # 16 candidates through a ring of 10
    for b in bs:
        pool.append(*b)
    assert pool.size == 10 and pool.cursor == 6
    allidx = torch.cat([b[0] for b in bs])
    _, _, _, index = pool.tensors()
    assert sorted(index.tolist()) == sorted(allidx[-10:].tolist())
    assert index[:6].tolist() == allidx[10:16].tolist() and index[6:].tolist() == allidx[6:10].tolist()
    assert pool.growing_bounds() is None


def test_chunk_ring_scores_a_streamed_shard_like_one_resident_call():
    """pool.ChunkRing on CPU tensors (host logic: window rotation, ragged last window, bounds); the oracle stands in for K1."""
    import torch
    from oracle import mercury_oracle as orc
    from mercury_b200.pool import ChunkRing

    class Ops:
        def score(self, logits, labels, status=None, out=None):
            out.copy_(orc.score(logits, labels))
            return out

    g = torch.Generator().manual_seed(1)
    n, c, win = 1000, 7, 128
    logits, labels = torch.randn(n, c, generator=g), torch.randint(0, c, (n,), generator=g)
    ring = ChunkRing(win, c, dtype=torch.float32, device="cpu", depth=3, ops_module=Ops())
    scores, labs = torch.zeros(n), torch.zeros(n, dtype=torch.int64)
    rows = ring.score(((logits[i:i + win], labels[i:i + win]) for i in range(0, n, win)), scores, labs)
    assert rows == n and torch.equal(scores, orc.score(logits, labels)) and torch.equal(labs, labels)
    import pytest
    with pytest.raises(ValueError):
        ring.score([(logits[:win + 1], labels[:win + 1])], scores)
    with pytest.raises(ValueError):
        ring.score(((logits[i:i + win], labels[i:i + win]) for i in range(0, n, win)), scores[:500])
