"""Fit: the training loop the reference gets from `train.py` + pytorch-lightning 1.9, without Lightning.

What it reproduces, with the reference line it comes from:
  * LR scaling           train.py:383  model.learning_rate = accumulate_grad_batches * ngpu * batch_size * base_lr
  * two optimizers       AE.py:365-377 [opt_ae, opt_disc]; per batch training_step(optimizer_idx=0) then (optimizer_idx=1)
                         (AE.py:295-330), each followed by its own optimizer step (PL 1.9 automatic optimization)
  * global_step          PL >= 1.6 counts optimizer steps across BOTH optimizers (two per batch); every phase threshold
                         of the model / loss (30000 / 37000 / 75000 / 120000) is in those units
  * gradient clipping    CFG:140 gradient_clip_val 1.0 (global 2-norm, per optimizer)
  * DDP                  CFG:137 strategy ddp: batch-sharded replicas, gradient mean over ranks (ddp.GradArena buckets)
  * checkpoints          PL layout subset: {"epoch", "global_step", "pytorch-lightning_version", "state_dict",
                         "optimizer_states"}; `state_dict` keys are the module's own (reference names/shapes), so
                         PoseAutoencoder.init_from_ckpt (AE.py:97, LD autoencoder.py:316-326) loads them unchanged,
                         and a reference `last.ckpt` loads here the same way.

The optimizer_idx-0 step runs through trainer.TrainStep (gradient arena, bucketed allreduce overlapped with backward,
fused clip + Adam). The discriminator step (optimizer_idx 1: hinge loss on detached real / fake crops, its own clip +
Adam, gradients all-reduced over ranks) is taken while the GAN is active (`disc_factor > 0` and `global_step >=
disc_start`); before that the reference's d_loss is identically 0 and its Adam update is a no-op on fresh moments, so
skipping it is exact; global_step still advances as in Lightning.
"""
import os

import torch
import torch.distributed as dist

from . import ops
from .trainer import TrainStep

PL_VERSION = "1.9.0"  # environment.yml:277 of the reference


class Fit:
    def __init__(self, model, batch_size, base_lr=4.5e-6, accumulate_grad_batches=1, scale_lr=True, max_grad_norm=1.0,
                 ckpt_dir=None, ckpt_every_n_batches=0, log_fn=None, process_group=None):
        if accumulate_grad_batches != 1:
            raise NotImplementedError("the reference config trains with accumulate_grad_batches = 1")
        self.model = model
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(process_group) if self.world > 1 else 0
        lr = accumulate_grad_batches * self.world * batch_size * base_lr if scale_lr else base_lr
        model.learning_rate = lr  # train.py:383 / :388
        self.ae_step = TrainStep(model, lr=lr, max_grad_norm=max_grad_norm, process_group=process_group)
        self.opt_ae = self.ae_step.opt
        self.opt_disc = model.configure_optimizers()[0][1]
        if hasattr(self.opt_disc, "max_grad_norm"):  # FusedAdam: clip folded into the multi-tensor kernel
            self.opt_disc.max_grad_norm = max_grad_norm
            self.opt_disc.grad_scale = 1.0 / self.world
        self.pg = process_group
        self.max_grad_norm = max_grad_norm
        self.ckpt_dir, self.ckpt_every = ckpt_dir, ckpt_every_n_batches
        self.log_fn = log_fn
        self.epoch = 0
        self.batches_seen = 0

    # ------------------------------------------------------------------ one batch = two optimizer steps
    def training_batch(self, batch, batch_idx):
        model = self.model
        aeloss = self.ae_step(batch, batch_idx)
        model.global_step += 1
        discloss = model.training_step(batch, batch_idx, 1)
        if model._gan_active() and torch.is_tensor(discloss) and discloss.requires_grad:
            self.disc_step(discloss)
        model.global_step += 1
        self.batches_seen += 1
        return aeloss, discloss.detach()

    def disc_step(self, discloss):
        """backward + gradient mean over ranks + clip + Adam for the discriminator (PL: second optimizer, CFG:137,140)."""
        params = [p for p in self.model.loss.discriminator.parameters() if p.requires_grad]
        self.opt_disc.zero_grad(set_to_none=True)
        discloss.backward()
        if self.world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.pg)
            o = 0
            for p in params:
                p.grad.copy_(flat[o:o + p.numel()].view_as(p))
                o += p.numel()
        if not hasattr(self.opt_disc, "max_grad_norm"):
            if self.world > 1:
                for p in params:
                    p.grad.div_(self.world)
            torch.nn.utils.clip_grad_norm_(params, self.max_grad_norm)
        self.opt_disc.step()
        ops.refresh_param_caches()  # the discriminator's packed weights, in one multi-tensor launch

    def fit(self, loader, max_steps=None, max_epochs=1):
        """Runs until `max_steps` optimizer steps (Lightning's `max_steps` unit) or `max_epochs` passes over `loader`."""
        model = self.model
        last = None
        for _ in range(max_epochs):
            for batch_idx, batch in enumerate(loader):
                if max_steps is not None and model.global_step >= max_steps:
                    return last
                last = self.training_batch(batch, batch_idx)
                if self.log_fn is not None:
                    self.log_fn(model.global_step, float(last[0]), float(last[1]))
                if self.ckpt_dir and self.ckpt_every and self.batches_seen % self.ckpt_every == 0:
                    self.save_checkpoint(os.path.join(self.ckpt_dir, "last.ckpt"))
            self.epoch += 1
        if self.ckpt_dir:
            self.save_checkpoint(os.path.join(self.ckpt_dir, "last.ckpt"))
        return last

    # ------------------------------------------------------------------ checkpoints (PL key layout)
    def save_checkpoint(self, path):
        if self.rank != 0:
            return
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        ckpt = {"epoch": self.epoch, "global_step": int(self.model.global_step),
                "pytorch-lightning_version": PL_VERSION,
                "state_dict": {k: v.detach().cpu() for k, v in self.model.state_dict().items()},
                "optimizer_states": [self.opt_ae.state_dict(), self.opt_disc.state_dict()]}
        tmp = path + ".tmp"
        torch.save(ckpt, tmp)
        os.replace(tmp, path)

    def load_checkpoint(self, path, strict=True):
        ckpt = torch.load(path, map_location="cpu", weights_only=False)
        self.model.load_state_dict(ckpt["state_dict"], strict=strict)
        self.model.global_step = int(ckpt.get("global_step", 0))
        self.epoch = int(ckpt.get("epoch", 0))
        states = ckpt.get("optimizer_states")
        if states:
            self.opt_ae.load_state_dict(states[0])
            if len(states) > 1:
                self.opt_disc.load_state_dict(states[1])
        return ckpt
