"""Synthetic stand-ins for the reference's loader factories (exp_dataset.py:9-77).

The reference downloads CIFAR-10 (exp_dataset.py:32,68) -- there is no network here and the
BASELINE configs say "synthetic", so the data are seeded CIFAR-shaped tensors.  What is kept is
the surface the hot path relies on: datasets yield ``(index, image, target)`` triples
(util.py:242-262), batch size 32 (exp_dataset.py:11,24), an unshuffled presample loader and a
shuffled train loader (:47-61), and ``split_num`` presample loaders for the multi-worker case.
"""
from __future__ import annotations

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset, Subset


class SyntheticCIFAR10(Dataset):
    """CIFAR-shaped (3x32x32, 10 classes) seeded data returning (index, img, target) like My_CIFAR10."""

    def __init__(self, n: int = 50000, seed: int = 0, num_classes: int = 10, shape=(3, 32, 32)):
        g = torch.Generator().manual_seed(seed)
        self.data = torch.rand((n,) + tuple(shape), generator=g)
        self.targets = torch.randint(0, num_classes, (n,), generator=g)

    def __len__(self):
        return self.data.shape[0]

    def __getitem__(self, index):
        return index, self.data[index], int(self.targets[index])

    def get_slice(self, start, end):  # util.py:264-273 without the Python loop
        return self.data[start:end], self.targets[start:end]


def load_cifar10(split_num: int = 1, n_train: int = 50000, n_test: int = 10000, seed: int = 0, batch_size: int = 32):
    """exp_dataset.py:23-77: (train_loader, presam_loader(s), test_loader)."""
    train = SyntheticCIFAR10(n_train, seed)
    test = SyntheticCIFAR10(n_test, seed + 1)
    test_loader = DataLoader(test, batch_size=batch_size, shuffle=False)
    if split_num > 1:  # :34-43 IID split into per-worker presample loaders
        parts = np.array_split(np.arange(n_train), split_num)
        presam = [DataLoader(Subset(train, p.tolist()), batch_size=batch_size, shuffle=True) for p in parts]
        return DataLoader(train, batch_size=batch_size, shuffle=True), presam, test_loader
    train_loader = DataLoader(train, batch_size=batch_size, shuffle=True)
    presam_loader = DataLoader(train, batch_size=batch_size, shuffle=False)  # :56-61 unshuffled
    return train_loader, presam_loader, test_loader


def partition_hetero(targets: np.ndarray, n_nets: int, alpha: float, rng=np.random, num_classes: int = 10, min_size_required: int = 10):
    """The reference's Dirichlet partitioner, cifar10/data_loader.py:138-161 ("hetero"), restated: per class, shuffle its
    indices, draw Dirichlet(alpha) proportions over the workers, zero the share of every worker that already holds its
    N / n_nets ("Balance", :150-151), split; repeat the whole assignment until the smallest shard has at least 10 samples
    (:144); finally shuffle every shard.  ``rng``: the module ``numpy.random`` (the reference uses the global generator it
    seeds at pytorch_collab.py:280) or a RandomState -- same calls in the same order, so the same seed gives the same map."""
    targets = np.asarray(targets)
    N = targets.shape[0]
    min_size = 0
    idx_batch = [[] for _ in range(n_nets)]
    while min_size < min_size_required:
        idx_batch = [[] for _ in range(n_nets)]
        for k in range(num_classes):
            idx_k = np.where(targets == k)[0]
            rng.shuffle(idx_k)
            proportions = rng.dirichlet(np.repeat(alpha, n_nets))
            proportions = np.array([p * (len(idx_j) < N / n_nets) for p, idx_j in zip(proportions, idx_batch)])
            proportions = proportions / proportions.sum()
            cuts = (np.cumsum(proportions) * len(idx_k)).astype(int)[:-1]
            idx_batch = [idx_j + idx.tolist() for idx_j, idx in zip(idx_batch, np.split(idx_k, cuts))]
            min_size = min(len(idx_j) for idx_j in idx_batch)
    for j in range(n_nets):
        rng.shuffle(idx_batch[j])
    return {j: idx_batch[j] for j in range(n_nets)}


def load_cifar10_noniid(split_num: int, alpha: float, n_train: int = 50000, n_test: int = 10000, seed: int = 0,
                        batch_size: int = 32):
    """exp_dataset.py:9-20: (presam_loaders[list], train_global, test_global) with the Dirichlet(alpha) class-skewed split
    across workers of cifar10/data_loader.py:138-161 (``partition_hetero``), on synthetic data."""
    assert split_num > 1
    train = SyntheticCIFAR10(n_train, seed)
    test = SyntheticCIFAR10(n_test, seed + 1)
    shards = partition_hetero(train.targets.numpy(), split_num, alpha, np.random.RandomState(seed))
    presam = [DataLoader(Subset(train, shards[j]), batch_size=batch_size, shuffle=True) for j in range(split_num)]
    return presam, DataLoader(train, batch_size=batch_size, shuffle=True), DataLoader(test, batch_size=batch_size)
