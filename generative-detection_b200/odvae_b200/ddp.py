"""Data parallelism for the train step: one process per GPU, flat gradient arena, bucketed NCCL allreduce
overlapped with backward (the reference relies on Lightning's `strategy: ddp`, CFG:137).

GradArena lays the AE parameter gradients out in ONE fp32 buffer in reverse execution order
(decoder.conv_out ... encoder.conv_in), so buckets complete front to back while backward runs. The wgrad /
GroupNorm-backward / Linear-backward kernels write their results STRAIGHT into the arena views (ops._new_grad):
no per-parameter gradient tensors, no gather copy into buckets, and FusedAdam's chunk table is static.
Each bucket's allreduce is enqueued on a side stream as soon as the activation-gradient hook at the bucket
boundary fires; the optimizer stream waits on the last bucket only.

Semantics kept from the reference's DDP run: gradients are averaged (sum-allreduce, 1/world folded into the Adam
kernel's grad_scale), loss normalisers stay per-rank, parameters unused in the encoder-pretrain phase get zero
gradients on every rank (the phase is a pure function of global_step, so ranks agree without communication).
"""
import torch
import torch.distributed as dist


class GradArena:
    def __init__(self, params, bucket_bytes=32 << 20, process_group=None):
        self.params = list(params)  # already in bucket (reverse execution) order
        dev = self.params[0].device
        offs, total = [], 0
        for p in self.params:
            offs.append(total)
            total += (p.numel() + 3) // 4 * 4  # keep every view 16-byte aligned
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        self.offsets = offs
        for p, o in zip(self.params, offs):
            view = self.flat[o:o + p.numel()].view(p.shape)
            p._odvae_grad_dst = view
            p.grad = view
        # bucket boundaries (element offsets)
        self.buckets, start = [], 0
        cap = bucket_bytes // 4
        for p, o in zip(self.params, offs):
            end = o + (p.numel() + 3) // 4 * 4
            if end - start >= cap:
                self.buckets.append((start, end))
                start = end
        if start < total:
            self.buckets.append((start, total))
        self.pg = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.comm_stream = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self._next = 0
        self._index = {id(p): i for i, p in enumerate(self.params)}

    def release(self):
        for p in self.params:
            if hasattr(p, "_odvae_grad_dst"):
                del p._odvae_grad_dst
            p.grad = None

    def zero(self):
        self.flat.zero_()

    def begin_step(self):
        self._next = 0
        for p in self.params:  # set again by every kernel wrapper that writes this parameter's gradient (ops._new_grad)
            p._odvae_written = False

    def _launch(self, b):
        s, e = self.buckets[b]
        if self.world > 1:
            if self.comm_stream is not None:
                self.comm_stream.wait_stream(torch.cuda.current_stream())
                from . import ops
                side = ops._SIDE.get(self.flat.device.index)
                if side is not None:  # weight gradients still running on the overlap stream (ops.side_release)
                    self.comm_stream.wait_stream(side)
                with torch.cuda.stream(self.comm_stream):
                    dist.all_reduce(self.flat[s:e], op=dist.ReduceOp.SUM, group=self.pg)
            else:
                dist.all_reduce(self.flat[s:e], op=dist.ReduceOp.SUM, group=self.pg)

    def ready_through(self, param):
        """All gradients of parameters up to and including `param` (arena order) have been enqueued."""
        i = self._index[id(param)]
        done = self.offsets[i] + (param.numel() + 3) // 4 * 4
        while self._next < len(self.buckets) and self.buckets[self._next][1] <= done:
            self._launch(self._next)
            self._next += 1

    def finish(self):
        """Flush the remaining buckets and make the compute stream wait for every allreduce."""
        while self._next < len(self.buckets):
            self._launch(self._next)
            self._next += 1
        if self.world > 1 and self.comm_stream is not None:
            torch.cuda.current_stream().wait_stream(self.comm_stream)

    def boundary_hook(self, param):
        """Tensor hook for the activation that enters the module owning `param` (its gradient is produced
        after every parameter gradient of that module and of all later modules)."""
        def hook(grad):
            from . import ops
            if ops._PROBE[0] == 0:  # not inside an autograd.grad probe (adaptive GAN weight)
                self.ready_through(param)
            return None
        return hook


def bucket_order(model):
    """AE parameters in reverse execution order of the forward pass (AE.py:208-257)."""
    enc, dec = model.encoder, model.decoder
    fwd = []

    def add(mod):
        fwd.extend(list(mod.parameters()))

    add(enc.conv_in)
    for lvl in enc.down:
        for i, blk in enumerate(lvl.block):
            add(blk)
            if len(lvl.attn) > 0:
                add(lvl.attn[i])
        if hasattr(lvl, "downsample"):
            add(lvl.downsample)
    add(enc.mid); add(enc.norm_out); add(enc.conv_out)
    add(model.quant_conv_obj); add(model.quant_conv_pose); add(model.pose_decoder); add(model.pose_encoder)
    add(model.post_quant_conv); add(dec.conv_in); add(dec.mid)
    for lvl in reversed(dec.up):
        for i, blk in enumerate(lvl.block):
            add(blk)
            if len(lvl.attn) > 0:
                add(lvl.attn[i])
        if hasattr(lvl, "upsample"):
            add(lvl.upsample)
    add(dec.norm_out); add(dec.conv_out)
    want = {id(p) for p in model.ae_parameters()}
    assert {id(p) for p in fwd} == want and len(fwd) == len(want)
    return list(reversed(fwd))
