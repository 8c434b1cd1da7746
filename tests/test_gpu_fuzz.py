"""Randomised whole-step cases (tools/fuzz_step.py: random pool sizes, class counts, draw counts, ties, zeros, constant
pools, both the one-kernel step and the kernel chain, repeated launches on one workspace) against the serial fp32 chain on
the device's own probabilities; and the case that run found: rows of 185..256 bytes (64-row K1 tiles) in a large pool, whose
tile sums used to overrun their part of the workspace."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from oracle import mercury_oracle as orc

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fuzz_both_step_paths_against_the_serial_chain():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_step.py"), "150", "3"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "150 iterations, 0 bad" in r.stdout


def test_fuzz_standalone_operators_against_the_oracle():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_ops.py"), "60", "2"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "done: 60 iterations, 0 bad" in r.stdout


@pytest.mark.parametrize("fused", [0, 1])
@pytest.mark.parametrize("n,c", [(832_582, 64), (645_335, 64), (300_000, 48)])
def test_wide_small_rows_large_pool(fused, n, c):
    from mercury_b200 import ops

    dev = "cuda:0"
    g = torch.Generator(device=dev).manual_seed(n)
    x = torch.randn(n, c, generator=g, device=dev) * 2
    y = torch.randint(0, c, (n,), generator=g, device=dev)
    u = torch.rand(128, dtype=torch.float64, generator=torch.Generator().manual_seed(1)).to(dev)
    ops.set_fused(fused)
    try:
        for rep in range(3):
            r = ops.update_samples(x, y, u, torch.zeros(1, device=dev), True)
            assert int(r.status.item()) == 0
            p = ops.normalize(r.scores, r.stats).cpu().numpy()
            assert np.array_equal(r.idx.cpu().numpy(), orc.draw_with_replacement(p, u.cpu().numpy()))
    finally:
        ops.set_fused(-1)
    ref = orc.score(x[:4096].cpu(), y[:4096].cpu())
    np.testing.assert_allclose(r.scores[:4096].cpu().numpy(), ref.numpy(), rtol=1e-5, atol=1e-5)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_fuzz_sharded_step_on_two_gpus():
    """tools/fuzz_sharded.py under torchrun: random shard sizes / class counts / draw counts, exchange over peer memory (kernel
    chain and one-kernel step) and over NCCL, each against the two-level oracle, assembled batch compared across ranks."""
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tools", "fuzz_sharded.py"), "12", "4"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "done: 12 iterations, 0 bad" in r.stdout
