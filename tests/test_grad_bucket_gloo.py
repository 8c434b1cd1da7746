"""World-size-2 gloo test (CPU) of the gradient averaging of the drop-in Trainer: the gradients live
in ONE persistent flat bucket (``.grad`` tensors are views of it), backward accumulates into it and one in-place
all-reduce averages it -- against the reference's procedure, restated: flatten the gradients, all_reduce(SUM), divide by
the world size, unflatten (pytorch_collab.py:236-249, util.py:23-63)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _net():
    torch.manual_seed(0)
    return torch.nn.Sequential(torch.nn.Linear(12, 7), torch.nn.ReLU(), torch.nn.Linear(7, 3))


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mercury_b200.collab import Trainer

    net = _net()
    tr = Trainer(net, torch.optim.SGD(net.parameters(), lr=0.1), None, None, None, "cpu", ops_module=object())
    g = torch.Generator().manual_seed(100 + rank)
    outs = []
    for step in range(3):
        x, y = torch.randn(5, 12, generator=g), torch.randint(0, 3, (5,), generator=g)
        tr._zero_grads()
        torch.nn.functional.cross_entropy(net(x), y).backward()
        views_ok = all(p.grad.data_ptr() >= tr._grad_bucket.data_ptr() and
                       p.grad.data_ptr() < tr._grad_bucket.data_ptr() + tr._grad_bucket.numel() * 4 for p in net.parameters())
        # the reference's procedure on copies of the same gradients
        flat = torch.cat([p.grad.reshape(-1).clone() for p in net.parameters()])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM)
        flat /= float(world)
        tr.average_gradients()
        got = torch.cat([p.grad.reshape(-1) for p in net.parameters()])
        outs.append((views_ok, torch.equal(got, flat), got.clone()))
    ret[rank] = outs
    dist.destroy_process_group()


def test_bucket_views_average_like_flatten_allreduce_unflatten():
    world = 2
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    for step in range(3):
        for rank in range(world):
            views_ok, equal, _ = ret[rank][step]
            assert views_ok, "a .grad tensor is not a view of the bucket"
            assert equal, "averaged gradients differ from flatten -> all_reduce(SUM) -> / world -> unflatten"
        assert torch.equal(ret[0][step][2], ret[1][step][2])
