"""GPU, 2 ranks (skipped on a 1-GPU box): the data-parallel train step over NCCL (trainer.TrainStep + ddp.GradArena: bucketed
sum-allreduce overlapped with backward, 1/world folded into Adam) against ONE rank running the concatenated batch
(SURVEY 4 item 3 / VERDICT r1 weak #10). The reference's DDP semantics keep per-rank loss normalisers (sum(mask_bg)) and
a per-rank `_rescale`; the two shards are built with equal sum(mask_bg) and equal pixel min/max, for which the mean of the
per-rank gradients equals the gradient of the 2x batch -- so the parameters after the step must agree to fp32 rounding.
Run with: gpurun --gpus 2 -- python -m pytest tests/test_ddp_gpu.py -m gpu"""
import os
import tempfile

import pytest
import torch

import helpers as H
from odvae_b200 import synth

pytestmark = pytest.mark.gpu


def _shards():
    full = synth.synthetic_batch(4, 64, seed=91)
    full["class_id"] = torch.tensor([0, 1, 3, 1])  # one "class 1" (mask_bg = 0) sample per shard -> sum(mask_bg) = 1 each
    full["class_name"] = [synth.LABELS[int(c)] for c in full["class_id"]]
    for b0 in (0, 2):  # same pixel range in both shards (the reference rescales per rank, AE.py:434-436)
        full["patch"][b0, 0, 0, 0] = 0.0
        full["patch"][b0, 0, 0, 1] = 1.0
    cut = lambda lo, hi: {k: (v[lo:hi].clone() if torch.is_tensor(v) else list(v[lo:hi])) for k, v in full.items()}
    return full, cut(0, 2), cut(2, 4)


def _one_step(rank, world, batch, noise, out_path, port):
    import torch.distributed as dist
    from odvae_b200.autoencoder import PoseAutoencoder
    from odvae_b200.trainer import TrainStep
    torch.cuda.set_device(rank)
    if world > 1:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    params = synth.model_config(synth.DDCONFIG_SMALL, H.stats_path(23))
    model = PoseAutoencoder(**params, compute_dtype=torch.float32)
    synth.fill_state_dict_(model.state_dict(), seed=23)
    model = model.cuda()
    model.global_step = 80000
    model.inject_noise = noise
    step = TrainStep(model, lr=1e-3, max_grad_norm=1.0, bucket_bytes=1 << 20)
    b = {k: (v.cuda() if torch.is_tensor(v) else list(v)) for k, v in batch.items()}
    loss = step(b, 0)
    torch.cuda.synchronize()
    if rank == 0:
        torch.save({"loss": float(loss), "grads": step.arena.flat.detach().cpu() / world,
                    "params": {k: p.detach().cpu() for k, p in model.named_parameters()}}, out_path)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _worker(rank, world, shards, noises, out_path, port):
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.join(os.path.dirname(here), "generative-detection_b200"), os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    _one_step(rank, world, shards[rank], noises[rank], out_path, port)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_two_rank_step_equals_the_one_rank_step_on_the_concatenated_batch():
    import torch.multiprocessing as mp
    full, s0, s1 = _shards()
    nz = synth.draw_noise(4, 0.8, seed=17)
    cut = lambda lo, hi: tuple(t[lo:hi].clone() for t in nz)
    tmp = tempfile.mkdtemp()
    two, one = os.path.join(tmp, "two.pt"), os.path.join(tmp, "one.pt")
    mp.spawn(_worker, args=(2, [s0, s1], [cut(0, 2), cut(2, 4)], two, 29671), nprocs=2, join=True)
    mp.spawn(_worker, args=(1, [full], [nz], one, 0), nprocs=1, join=True)
    a, b = torch.load(two), torch.load(one)
    assert H.rel_l2(a["grads"], b["grads"]) < 1e-5       # mean of the two ranks' gradients == gradient of the 2x batch
    # ... and the same clip + Adam update on every parameter. Adam's first step is lr * g / (|g| + eps): an element whose
    # gradient is rounding noise can land anywhere in [-lr, lr], so the bound is lr-sized (measured 4e-5), not fp32-sized
    for k in a["params"]:
        assert H.rel_max(a["params"][k], b["params"][k]) < 2e-4, k
