"""FusedAdam: torch.optim.Adam semantics (AE.py:365-377: betas=(0.5, 0.9), eps=1e-8, no weight decay) with
optional global-norm clipping (Lightning `gradient_clip_val: 1.0`, CFG:140) in three launches for ALL tensors.

One multi-tensor kernel walks a device-resident chunk table (odvae_mt_adam_step_dev) instead of the reference
stack's per-tensor loop (hundreds of launches). State keys (`step`, `exp_avg`, `exp_avg_sq`) match
torch.optim.Adam (its `capturable` flavour: `step` is a 0-dim float32 DEVICE tensor per parameter) so optimizer
checkpoints stay interchangeable.

Step counters live on the device, one per parameter: the kernel derives the bias corrections from them, so (a) a CUDA-graph
replay of the train step advances the optimizer without the host, and (b) a parameter whose gradient first appears at step
k starts its own count there. torch.optim.Adam skips parameters whose `.grad is None`; with the flat gradient arena every
parameter has a gradient buffer, so "has a gradient this step" is the arena's written-flag (`p._odvae_written`, set by the
kernels' wrappers through ops._new_grad and cleared by GradArena.begin_step): parameters not written in a step (decoder,
post_quant_conv and pose_encoder during the 30k-step encoder-pretrain phase) are neither updated nor aged.
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
        n = len(params)
        steps = torch.zeros(n, dtype=torch.float32, device=dev)
        active = torch.ones(n, dtype=torch.uint8, device=dev)
        init = np.zeros(n, dtype=np.float32)
        counts, p_ptr, m_ptr, v_ptr, owner, offs = [], [], [], [], [], []
        for i, p in enumerate(params):
            st = self.state[p]
            if len(st) == 0:
                st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
            else:  # a loaded checkpoint (python number, CPU or device tensor) or a previous plan's view
                init[i] = float(st.get("step", 0))
            if not p.is_contiguous():
                raise RuntimeError("FusedAdam needs contiguous parameters")
            numel = p.numel()
            for o in range(0, numel, CHUNK):
                owner.append(i)
                offs.append(o)
                counts.append(min(CHUNK, numel - o))
                p_ptr.append(p.data_ptr() + 4 * o)
                m_ptr.append(st["exp_avg"].data_ptr() + 4 * o)
                v_ptr.append(st["exp_avg_sq"].data_ptr() + 4 * o)
        steps.copy_(torch.from_numpy(init))
        for i, p in enumerate(params):
            self.state[p]["step"] = steps[i]  # 0-dim device view, like torch.optim.Adam(capturable=True)
        rows = len(counts)
        host = torch.empty(rows, 7, dtype=torch.int64, pin_memory=True)
        h = host.numpy()
        own = np.asarray(owner, dtype=np.int64)
        h[:, 0], h[:, 2], h[:, 3], h[:, 4] = p_ptr, m_ptr, v_ptr, counts
        h[:, 5] = steps.data_ptr() + 4 * own
        h[:, 6] = active.data_ptr() + own
        plan = dict(rows=rows, host=host, owner=own, offs=np.asarray(offs, dtype=np.int64) * 4,
                    table=torch.empty(rows, 7, dtype=torch.int64, device=dev), gkey=None,
                    ws=torch.empty(rows, dtype=torch.float32, device=dev),
                    norm=torch.empty(1, dtype=torch.float32, device=dev),
                    steps=steps, active=active, nparams=n,
                    active_host=torch.ones(n, dtype=torch.uint8, pin_memory=True), active_key=None)
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
            # which parameters received a gradient this step (arena mode; plain .grad tensors are always "written")
            akey = tuple(bool(getattr(p, "_odvae_written", True)) for p in params)
            if akey != plan["active_key"]:
                plan["active_host"].numpy()[:] = np.asarray(akey, dtype=np.uint8)
                plan["active"].copy_(plan["active_host"], non_blocking=True)
                plan["active_key"] = akey
            b1, b2 = group["betas"]
            ops.call("mt_step_incr", plan["steps"], plan["active"], plan["nparams"])
            norm = None
            if self.max_grad_norm is not None:
                norm = plan["norm"]
                ops.call("mt_sumsq", _table5(plan), plan["rows"], norm, plan["ws"], plan["ws"].numel() * 4)
                self.grad_norm_sq = norm
            ops.call("mt_adam_step_dev", plan["table"], plan["rows"], float(group["lr"]), float(b1), float(b2),
                     float(group["eps"]), norm, float(self.max_grad_norm or 0.0), float(self.grad_scale))
            ops.bump_param_epochs(params)
        return loss


def _table5(plan):
    """The 5-column {p, g, m, v, n} table odvae_mt_sumsq reads (kept next to the 7-column one, refreshed with it)."""
    t5 = plan.get("table5")
    if t5 is None or plan.get("table5_key") != plan["gkey"]:
        t5 = plan["table"][:, :5].contiguous() if t5 is None else t5.copy_(plan["table"][:, :5])
        plan["table5"], plan["table5_key"] = t5, plan["gkey"]
    return t5
