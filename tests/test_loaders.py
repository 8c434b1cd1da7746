"""The data surface the hot path relies on: (index, image, target) triples, the loader factories of
exp_dataset.py:9-77 on synthetic data, and the Dirichlet partitioner against the reference's own run."""
import os

import numpy as np
import torch

from mercury_b200 import exp_dataset as ed


def test_partition_hetero_equals_the_reference_run(golden_dir):
    """tests/golden/ref_partition.npz: cifar10/data_loader.py:126-161 executed unmodified on synthetic labels
    (tests/golden/make_golden.py ref_partition); the restatement must give the same shards for the same seed."""
    g = np.load(os.path.join(golden_dir, "ref_partition.npz"))
    n_nets, alpha, seed = int(g["n_nets"]), float(g["alpha"]), int(g["seed"])
    np.random.seed(seed)
    got = ed.partition_hetero(g["labels"], n_nets, alpha)            # the global generator, as the reference uses it
    for j in range(n_nets):
        assert np.array_equal(np.asarray(got[j]), g[f"shard{j}"]), j
    again = ed.partition_hetero(g["labels"], n_nets, alpha, np.random.RandomState(seed))   # a private generator: same stream
    for j in range(n_nets):
        assert np.array_equal(np.asarray(again[j]), g[f"shard{j}"]), j
    allidx = np.sort(np.concatenate([g[f"shard{j}"] for j in range(n_nets)]))
    assert np.array_equal(allidx, np.arange(len(g["labels"])))      # a partition: every sample exactly once


def test_single_worker_loaders_yield_index_triples():
    train, presam, test = ed.load_cifar10(1, n_train=256, n_test=64)
    idx, img, tgt = next(iter(presam))
    assert idx.tolist() == list(range(32)) and img.shape == (32, 3, 32, 32) and tgt.shape == (32,)   # unshuffled (exp_dataset.py:56-61)
    assert torch.equal(img, presam.dataset.data[:32]) and train.batch_size == 32 and len(test.dataset) == 64
    idx2, img2, _ = next(iter(train))
    assert torch.equal(img2, train.dataset.data[idx2])              # the index really addresses the sample


def test_multi_worker_iid_split_covers_the_dataset_once():
    train, presams, _ = ed.load_cifar10(4, n_train=1000, n_test=10)
    assert len(presams) == 4
    seen = []
    for ld in presams:
        for idx, img, tgt in ld:
            assert torch.equal(img, train.dataset.data[idx])
            seen += idx.tolist()
    assert sorted(seen) == list(range(1000))


def test_noniid_loaders_are_skewed_partitions():
    presams, train, test = ed.load_cifar10_noniid(4, 0.5, n_train=4000, n_test=10, seed=3)
    seen, hists = [], []
    for ld in presams:
        ids = list(ld.dataset.indices)
        seen += ids
        assert len(ids) >= 10                                       # the reference's min_size loop
        hists.append(np.bincount(train.dataset.targets[ids].numpy(), minlength=10) / len(ids))
    assert sorted(seen) == list(range(4000))
    assert np.abs(np.asarray(hists) - 0.1).max() > 0.1              # class-skewed, unlike the IID split
