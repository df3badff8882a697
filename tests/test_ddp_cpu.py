"""CPU, gloo, world_size 2: the gradient arena's bucketing + allreduce logic (odvae_b200/ddp.py).
The kernels are not involved: gradients are written into the arena views by hand, the way the wgrad kernels do."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from odvae_b200.ddp import GradArena


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    params = [nn.Parameter(torch.zeros(s)) for s in [(7, 3), (1000,), (33, 5, 3, 3), (1,), (4096,), (13,)]]
    arena = GradArena(params, bucket_bytes=4096)  # several small buckets
    assert len(arena.buckets) >= 3
    assert all(p.grad.data_ptr() == p._odvae_grad_dst.data_ptr() for p in params)
    assert all(arena.offsets[i] % 4 == 0 for i in range(len(params)))
    arena.begin_step()
    g = torch.Generator().manual_seed(100 + rank)
    local = []
    for i, p in enumerate(params):  # "backward": gradients appear in arena order, buckets fire as they complete
        v = torch.randn(p.shape, generator=g)
        local.append(v)
        p._odvae_grad_dst.copy_(v)
        arena.ready_through(p)
    arena.finish()
    assert arena._next == len(arena.buckets)
    # expected: sum over ranks
    exp = []
    for i, p in enumerate(params):
        tot = torch.zeros(p.shape)
        for r in range(world):
            gr = torch.Generator().manual_seed(100 + r)
            for j in range(i + 1):
                v = torch.randn(params[j].shape, generator=gr)
            tot += v
        exp.append(tot)
    ok = all(torch.allclose(p.grad, e, atol=1e-6) for p, e in zip(params, exp))
    out[rank] = ok
    dist.destroy_process_group()


def test_arena_bucketed_allreduce_world2():
    world = 2
    port = _free_port()
    mgr = mp.get_context("spawn").Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert all(out[r] for r in range(world))


def test_arena_layout_single_process():
    params = [nn.Parameter(torch.zeros(5)), nn.Parameter(torch.zeros(3, 3)), nn.Parameter(torch.zeros(17))]
    arena = GradArena(params, bucket_bytes=64)
    assert arena.flat.numel() == 8 + 12 + 20
    assert arena.buckets[0][0] == 0 and arena.buckets[-1][1] == arena.flat.numel()
    for (s0, e0), (s1, e1) in zip(arena.buckets, arena.buckets[1:]):
        assert e0 == s1
    arena.release()
    assert params[0].grad is None and not hasattr(params[0], "_odvae_grad_dst")
