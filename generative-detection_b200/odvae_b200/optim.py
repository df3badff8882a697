"""FusedAdam: torch.optim.Adam semantics (AE.py:365-377: betas=(0.5, 0.9), eps=1e-8, no weight decay) with
optional global-norm clipping (Lightning `gradient_clip_val: 1.0`, CFG:140) in two launches for ALL tensors.

One multi-tensor kernel walks a device-resident chunk table (odvae_mt_adam_step) instead of the reference
stack's per-tensor loop (hundreds of launches). State keys (`step`, `exp_avg`, `exp_avg_sq`)
match torch.optim.Adam so optimizer checkpoints stay interchangeable.
"""
import numpy as np
import torch

from . import ops

CHUNK = 65536


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, max_grad_norm=None, grad_scale=1.0):
        defaults = dict(lr=lr, betas=betas, eps=eps)
        super().__init__(params, defaults)
        self.max_grad_norm = max_grad_norm
        self.grad_scale = grad_scale  # e.g. 1/world_size after a sum-allreduce
        self._plans = {}
        self.grad_norm_sq = None  # device scalar of the last step (unscaled gradients), when clipping is on

    def _plan(self, gi, group, params):
        """Static part of the chunk table for the parameters of this group that currently have gradients."""
        key = (gi, tuple(id(p) for p in params))
        plan = self._plans.get(key)
        if plan is not None:
            return plan
        dev = params[0].device
        counts, p_ptr, m_ptr, v_ptr, owner, offs = [], [], [], [], [], []
        for i, p in enumerate(params):
            st = self.state[p]
            if len(st) == 0:
                st["step"] = 0
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            if not p.is_contiguous():
                raise RuntimeError("FusedAdam needs contiguous parameters")
            n = p.numel()
            for o in range(0, n, CHUNK):
                owner.append(i)
                offs.append(o)
                counts.append(min(CHUNK, n - o))
                p_ptr.append(p.data_ptr() + 4 * o)
                m_ptr.append(st["exp_avg"].data_ptr() + 4 * o)
                v_ptr.append(st["exp_avg_sq"].data_ptr() + 4 * o)
        rows = len(counts)
        host = torch.empty(rows, 5, dtype=torch.int64, pin_memory=True)
        h = host.numpy()
        h[:, 0], h[:, 2], h[:, 3], h[:, 4] = p_ptr, m_ptr, v_ptr, counts
        plan = dict(rows=rows, host=host, owner=np.asarray(owner), offs=np.asarray(offs, dtype=np.int64) * 4,
                    table=torch.empty(rows, 5, dtype=torch.int64, device=dev), gkey=None,
                    ws=torch.empty(rows, dtype=torch.float32, device=dev),
                    norm=torch.empty(1, dtype=torch.float32, device=dev))
        self._plans[key] = plan
        return plan

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._plans = {}  # the chunk tables hold raw pointers into the replaced exp_avg / exp_avg_sq tensors

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            params = [p for p in group["params"] if p.grad is not None]
            if not params:
                continue
            if not params[0].is_cuda:
                raise RuntimeError("FusedAdam runs on the B200 only; use torch.optim.Adam on CPU")
            plan = self._plan(gi, group, params)
            gptrs = tuple(p.grad.data_ptr() for p in params)
            if gptrs != plan["gkey"]:
                for p in params:
                    if not p.grad.is_contiguous() or p.grad.dtype != torch.float32:
                        raise RuntimeError("FusedAdam needs contiguous fp32 gradients")
                plan["host"].numpy()[:, 1] = np.asarray(gptrs, dtype=np.int64)[plan["owner"]] + plan["offs"]
                plan["table"].copy_(plan["host"], non_blocking=True)
                plan["gkey"] = gptrs
            step = int(self.state[params[0]]["step"]) + 1  # a torch.optim.Adam checkpoint stores a tensor here
            for p in params:
                self.state[p]["step"] = step
            b1, b2 = group["betas"]
            norm = None
            if self.max_grad_norm is not None:
                norm = plan["norm"]
                ops.call("mt_sumsq", plan["table"], plan["rows"], norm, plan["ws"], plan["ws"].numel() * 4)
                self.grad_norm_sq = norm
            ops.call("mt_adam_step", plan["table"], plan["rows"], float(group["lr"]), float(b1), float(b2),
                     float(group["eps"]), 1.0 - b1 ** step, 1.0 - b2 ** step, norm,
                     float(self.max_grad_norm or 0.0), float(self.grad_scale))
        ops.bump_weight_epoch()
        return loss
