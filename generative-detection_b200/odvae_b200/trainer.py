"""TrainStep: one optimizer_idx-0 training step of the reference loop, as Lightning would run it
(training_step -> backward -> gradient allreduce -> clip_grad_norm(1.0) -> Adam), on the B200 kernels.

Reference: PoseAutoencoder.training_step (src/models/autoencoder.py:295-319), configure_optimizers (:365-377),
`gradient_clip_val: 1.0` and `strategy: ddp` (configs/autoencoder/pose/autoencoder_kl_16x16x16.yaml:137,140),
LR scaling train.py:383. This is the public entry bench.py and the scaling runs call.
"""
import torch
import torch.distributed as dist

from . import modules as M
from . import ops
from .ddp import GradArena, bucket_order
from .optim import FusedAdam


class TrainStep:
    def __init__(self, model, lr=4.5e-6, max_grad_norm=1.0, bucket_bytes=32 << 20, process_group=None):
        self.model = model
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.arena = GradArena(bucket_order(model), bucket_bytes=bucket_bytes, process_group=process_group)
        # gradients are sum-allreduced; the 1/world of DDP's mean is folded into the Adam kernel
        self.opt = FusedAdam(model.ae_parameters(), lr=lr, betas=(0.5, 0.9), max_grad_norm=max_grad_norm,
                             grad_scale=1.0 / self.world)
        self.unused_in_pretrain = model.global_step < model.encoder_pretrain_steps

    def __call__(self, batch, batch_idx=0):
        model, arena = self.model, self.arena
        if model.global_step < model.encoder_pretrain_steps:
            arena.zero()  # decoder / pose-encoder gradients are not produced in the encoder-pretrain phase
        arena.begin_step()
        model.loss.logvar.grad = None  # not in the AE optimizer (AE.py:367-374); do not let it accumulate
        M._ARENA[0] = arena if self.world > 1 else None
        try:
            loss = model.training_step(batch, batch_idx, 0)
            loss.backward()
        finally:
            M._ARENA[0] = None
        arena.finish()
        self.opt.step()
        ops.refresh_param_caches()  # packed 16-bit weight copies for the next step, queued back to back
        return loss.detach()
