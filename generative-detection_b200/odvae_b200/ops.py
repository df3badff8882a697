"""Autograd layer over the C ABI (include/odvae_b200.h).

Every Function below enqueues only kernels of libodvae_b200.so; torch supplies device memory
(`torch.empty`), streams and the autograd tape. Activations travel as NHWC tensors
[B,H,W,Cp] (Cp = channels padded to a multiple of 64) in the module's compute dtype:
bfloat16 (tcgen05 engine, the timed mode) or float32 (SIMT engine, the <=1e-4 validation mode).

Backward passes are hand-written per block (ResnetBlock, AttnBlock, ...) rather than per
primitive so that gradient accumulation is fused into kernel epilogues (`dx_add`) and the tape
stays short (~60 nodes for the whole autoencoder).

Reference call sites are cited per Function; LD/ = submodules/latent-diffusion/.
"""
import os
import weakref

import torch

from . import lib
from .lib import BF16, CONV_1x1, CONV_3x3_S1, CONV_3x3_S2, F16, F32, IMPL_SIMT, IMPL_TC, call, stream_handle

DS_SCALE_FP16 = 256.0  # extra factor on the attention score gradient while it is stored in fp16 (see AttnBlock.backward)
GROUPS = 32      # Normalize(): GroupNorm(num_groups=32, eps=1e-6)  LD model.py:38-39
GN_EPS = 1e-6

# bumped whenever parameters change behind autograd's back (FusedAdam kernels, load_state_dict)
_WEIGHT_EPOCH = [0]


def bump_weight_epoch():
    _WEIGHT_EPOCH[0] += 1


def bump_param_epochs(params):
    """An optimizer updated exactly these parameters in place (the fused Adam kernel does not touch `_version`): only the
    packed copies of THESE parameters go stale. With a global epoch the discriminator's optimizer step invalidated every
    packed autoencoder / LPIPS weight as well (216 single-weight re-pack launches per batch of the full objective)."""
    for p in params:
        p._odvae_epoch = getattr(p, "_odvae_epoch", 0) + 1


def act_code(dtype):
    if dtype == torch.bfloat16:
        return BF16
    if dtype == torch.float16:
        return F16
    if dtype == torch.float32:
        return F32
    raise TypeError("activations must be float32 or bfloat16, got %s" % dtype)


class Format:
    """Numeric formats of one compute mode (C-ABI dtype codes).

    fwd  : forward activations and packed forward weights of convs / GroupNorm outside attention blocks
    attn : everything inside an AttnBlock (normalised input, q|k|v, P, O and their packed weights)
    grad : every activation gradient, dgrad-packed weights and both operands of the weight-gradient GEMMs

    compute_dtype float32  -> fwd = attn = grad = f32  (SIMT engine; the <=1e-4 validation mode)
    compute_dtype bfloat16 -> fwd = attn = grad = bf16
    compute_dtype float16  -> fwd = attn = grad = IEEE fp16: the TIMED mode. 11-bit significands everywhere (bf16 storage
        measured 1.4-2.6e-2 on reconstructions and 1-3.4e-2 per-tensor on gradients through ~60 stacked layers; fp16 meets
        the 1e-2 tolerance on both). Range: forward tensors are O(1..1e3) (GroupNorm bounds them; stores saturate at
        65504 and raise the library's overflow flag); activation GRADIENTS travel multiplied by a static power-of-two
        scale per domain (GradScale: decoder side / encoder side, chosen from the batch size and the loss weights) that
        the parameter-gradient kernels divide out again (wgrad / GroupNorm-backward `*_scale` arguments).
    compute_dtype "mixed"  -> the round-1 mode, kept for A/B runs: fwd = fp16, attn = grad = bf16. tcgen05 needs both MMA
        operands in one format (a mix traps on B200), so its weight-gradient GEMMs read a bf16 copy of the forward operand
        (written by the GroupNorm apply kernel, or by `cast16` at the few other sites).

    Torch labelling: autograd casts a gradient to the dtype of the forward tensor it belongs to and ADDS the gradients of
    a tensor that has several consumers with torch's own kernels, so a buffer's torch dtype must say what its gradient
    holds: float32 / bfloat16 / float16 in the three uniform modes. Only in the "mixed" mode forward NHWC buffers are
    labelled torch.bfloat16 while holding fp16 bit patterns (their gradients are genuine bf16; only this library's kernels
    read the forward values; decode with `.view(torch.float16)`). The label is chosen where a tensor enters the NHWC domain
    (`Format.label`) and propagates from input to output tensors inside the kernels' wrappers.
    """

    def __init__(self, fwd, attn, grad):
        self.fwd, self.attn, self.grad = fwd, attn, grad
        self.mixed = fwd != grad
        self.buf = self.label(fwd)

    def label(self, code):
        """torch dtype that labels a buffer of format `code` in this mode (see the class docstring)."""
        if code == F32:
            return torch.float32
        return torch.float16 if (code == F16 and self.fwd == self.grad) else torch.bfloat16

    def __repr__(self):
        return "Format(fwd=%d, attn=%d, grad=%d)" % (self.fwd, self.attn, self.grad)


FORMAT_F32 = Format(F32, F32, F32)
FORMAT_BF16 = Format(BF16, BF16, BF16)
FORMAT_MIXED = Format(F16, BF16, BF16)
FORMAT_FP16 = Format(F16, F16, F16)


def format_for(compute_dtype):
    if isinstance(compute_dtype, Format):
        return compute_dtype
    if compute_dtype == "mixed":
        return FORMAT_MIXED
    return {torch.float32: FORMAT_F32, torch.bfloat16: FORMAT_BF16, torch.float16: FORMAT_FP16}[compute_dtype]


class GradScale:
    """Static scale S (a power of two) that the 16-bit activation gradients of one domain (decoder side / encoder side)
    are multiplied by while they travel through HBM in fp16; one shared object per domain, read at backward time."""

    def __init__(self, value=1.0):
        self.value = float(value)

    def __repr__(self):
        return "GradScale(%g)" % self.value


_UNIT_SCALE = GradScale(1.0)


def gs_of(holder):
    """Activation-gradient scale of the module / ConvSpec `holder` (1.0 unless assign_grad_scales stamped one)."""
    gs = getattr(holder, "gscale", None) if holder is not None else None
    return gs.value if gs is not None else 1.0


def assign_grad_scales(root, scale_obj):
    """Stamp every sub-module and ConvSpec under `root` with the shared GradScale object."""
    for m in root.modules():
        m.gscale = scale_obj
        for name in ("spec", "spec1", "spec2", "spec_nin", "spec_proj"):
            sp = getattr(m, name, None)
            if isinstance(sp, ConvSpec):
                sp.gscale = scale_obj


def pow2_near(v):
    import math
    return 2.0 ** round(math.log2(max(float(v), 1e-30)))


def disc_format(fmt):
    """Format of the PatchGAN discriminator inside a model of Format `fmt`: same forward format, gradients in bf16 whenever
    the model's gradient format is 16-bit (its logits gradient is ~1e-5 and unscaled: fp16 would flush it)."""
    if fmt.grad == F32:
        return fmt
    return Format(fmt.fwd, fmt.attn, BF16)


def format_of(x, holder=None):
    """Format carried by a module/spec (`holder.fmt`), else the one the tensor's torch dtype implies."""
    fmt = getattr(holder, "fmt", None) if holder is not None else None
    if fmt is not None:
        return fmt
    return FORMAT_F32 if x.dtype == torch.float32 else (FORMAT_FP16 if x.dtype == torch.float16 else FORMAT_BF16)


def assign_format(root, compute_dtype):
    """Stamp every sub-module and ConvSpec under `root` with the Format of `compute_dtype`."""
    fmt = format_for(compute_dtype)
    from .patchgan import NLayerDiscriminator
    skip = set()
    for m in root.modules():
        if isinstance(m, NLayerDiscriminator):
            m.fmt = disc_format(fmt)
            skip.update(id(c) for c in m.modules())
    for m in root.modules():
        if id(m) in skip:
            continue
        m.fmt = fmt
        spec = getattr(m, "spec", None)
        if isinstance(spec, ConvSpec):
            spec.fmt = fmt
    return fmt


def buf_dtype(code, like=None):
    """torch label of a buffer of format `code`: fp32 -> float32; 16-bit -> the label of the 16-bit tensor `like` it is derived
    from (so float16-labelled inputs give float16-labelled outputs), else bfloat16."""
    if code == F32:
        return torch.float32
    if like is not None and like in (torch.float16, torch.bfloat16):
        return like
    return torch.bfloat16


def cast16_raw(x, src, dst):
    """16-bit format conversion of a whole tensor (fp16 <-> bf16); identity when the codes agree."""
    if src == dst:
        return x
    x = _c(x)
    y = torch.empty_like(x)
    call("cast16", x, y, x.numel(), src, dst)
    return y


def cpad(c):
    """Channel padding of activation buffers (multiples of 64 feed TMA/tcgen05 tiles directly)."""
    return (c + 63) // 64 * 64


_WS = {}


def workspace(nbytes, device, tag):
    """Grow-only scratch per (device, stream, tag); never part of a state_dict."""
    key = (device.index, stream_handle(device.index), tag)
    buf = _WS.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 1 << 16), dtype=torch.uint8, device=device)
        _WS[key] = buf
    return buf


_OVERFLOW = {}


def overflow_flags(device):
    """int32[2] per device, set by the GroupNorm kernels: [0] an fp16 forward activation was stored saturated (+-65504),
    [1] an fp16 activation gradient was. Sticky until read_overflow() clears them."""
    t = _OVERFLOW.get(device.index)
    if t is None:
        t = torch.zeros(2, dtype=torch.int32, device=device)
        _OVERFLOW[device.index] = t
    return t


def read_overflow(device):
    """-> (forward_overflow, gradient_overflow) since the last call; synchronises with the device."""
    t = overflow_flags(device)
    f, g = (int(v) for v in t.tolist())
    if f or g:
        t.zero_()
    return bool(f), bool(g)


_TC_OK = {}


def conv_impl(code, B, H, W, CinP, CoutP, geom):
    """code: operand dtype code (or a torch dtype, for callers that predate Format)."""
    if isinstance(code, torch.dtype):
        code = act_code(code)
    if code == F32:
        return IMPL_SIMT
    key = (B, H, W, CinP, CoutP, geom)
    ok = _TC_OK.get(key)
    if ok is None:
        ok = bool(call("conv2d_tc_supported", B, H, W, CinP, CoutP, geom))
        _TC_OK[key] = ok
    return IMPL_TC if ok else IMPL_SIMT


def attn_fused(code, N, C):
    """The fused tcgen05 attention core needs 16-bit operands and TMA-tileable shapes."""
    return code != F32 and N % 256 == 0 and C % 64 == 0


def out_hw(geom, H, W):
    return (H // 2, W // 2) if geom == CONV_3x3_S2 else (H, W)


def geom_k(geom):
    return 1 if geom == CONV_1x1 else 3


# ------------------------------------------------------------------------------------------------
# parameter-side caches: packed weights / padded biases, refreshed when the parameter changes
# ------------------------------------------------------------------------------------------------
_CACHES = weakref.WeakSet()
_MT_PACK = {}  # signature of a set of stale packs -> device table (one per optimizer: autoencoder, discriminator)


class ParamCache:
    """Packed copies of one conv's parameters. Keyed by (param version, global weight epoch).
    `desc` = (w, geom, CoutP, CinP, which, code) marks an entry that is a plain odvae_pack_conv_weight of one
    parameter: refresh_param_caches() re-packs all of those in ONE multi-tensor launch."""

    def __init__(self):
        self.store = {}
        _CACHES.add(self)

    @staticmethod
    def _stamp(params):
        return (tuple(p._version for p in params) + tuple(p.data_ptr() for p in params) +
                tuple(getattr(p, "_odvae_epoch", 0) for p in params) + (_WEIGHT_EPOCH[0],))

    def get(self, key, params, build, desc=None):
        stamp = self._stamp(params)
        hit = self.store.get(key)
        if hit is not None and hit[0] == stamp:
            return hit[1]
        val = build(hit[1] if hit is not None else None)
        self.store[key] = (stamp, val, params, build, desc)
        return val

    def stale(self):
        for key, (stamp, val, params, build, desc) in list(self.store.items()):
            new = self._stamp(params)
            if new != stamp:
                yield key, new, val, params, build, desc


def refresh_param_caches():
    """Re-pack every cached weight copy whose parameter changed (called right after the optimizer step). Conv-weight
    packs go through one multi-tensor launch over a device table that is built once (parameter and packed-buffer
    addresses are stable across steps); anything else (padded biases, fused qkv) re-runs its own builder."""
    mt = []
    for c in list(_CACHES):
        for key, new, val, params, build, desc in c.stale():
            if desc is not None and val is not None:
                mt.append((c, key, new, val, params, build, desc))
            else:
                c.store[key] = (new, build(val), params, build, desc)
    if not mt:
        return
    sig = tuple((d[0].data_ptr(), val.data_ptr(), d[1], d[2], d[3], d[4], d[5]) for _, _, _, val, _, _, d in mt)
    entry = _MT_PACK.get(sig)
    if entry is None:
        rows, first = [], 0
        for _, _, _, val, _, _, (w, geom, coutp, cinp, which, code) in mt:
            k = geom_k(geom)
            flags = which | (2 if (geom == CONV_3x3_S2 and which == 1) else 0)
            rows.append([w.data_ptr(), val.data_ptr(), w.shape[0], w.shape[1], k, coutp, cinp, flags, code, first])
            first += call("pack_conv_weights_mt_blocks", coutp, cinp, geom)
        host = torch.tensor(rows, dtype=torch.int64)
        # (never evicted: a table is ~25 KB, and a CUDA-graph capture must find the table its warm-up steps built)
        entry = _MT_PACK[sig] = dict(table=host.to(mt[0][3].device), n=len(rows), blocks=first)
    call("pack_conv_weights_mt", entry["table"], entry["n"], entry["blocks"])
    for c, key, new, val, params, build, desc in mt:
        c.store[key] = (new, val, params, build, desc)


def pack_weight(w, geom, CoutP, CinP, which, code, out=None):
    """code: dtype code of the packed copy (a torch dtype is accepted too)."""
    if isinstance(code, torch.dtype):
        code = act_code(code)
    k = geom_k(geom)
    if out is None:
        out = torch.empty(CoutP * CinP * k * k, dtype=buf_dtype(code), device=w.device)
    call("pack_conv_weight", w, out, w.shape[0], w.shape[1], geom, CoutP, CinP, which, code)
    return out


def pad_vector(v, n, out=None):
    """fp32 [C] -> zero-padded fp32 [n] (bias rows of padded output channels)."""
    if v.numel() == n:
        return v
    if out is None:
        out = torch.empty(n, dtype=torch.float32, device=v.device)
    call("nchw_to_nhwc", v, out, 1, v.numel(), 1, 1, n, F32, 1.0, 0.0)
    return out


class ConvSpec:
    """Geometry + caches of one Conv2d (torch.nn.Conv2d call sites)."""

    fmt = None  # set by assign_format(); None = derive from the activation's torch dtype

    def __init__(self, cin, cout, geom, cout_pad=None):
        self.cin, self.cout, self.geom = cin, cout, geom
        self.cinp = cpad(cin)
        self.coutp = cout_pad if cout_pad is not None else cpad(cout)
        self.cache = ParamCache()

    def w_fwd(self, w, code, coutp=None):
        coutp = coutp or self.coutp
        code = act_code(code) if isinstance(code, torch.dtype) else code
        return self.cache.get(("f", code, coutp), (w,),
                              lambda old: pack_weight(w.detach(), self.geom, coutp, self.cinp, 0, code, old),
                              desc=(w, self.geom, coutp, self.cinp, 0, code))

    def w_dgrad(self, w, code, coutp=None):
        coutp = coutp or self.coutp
        code = act_code(code) if isinstance(code, torch.dtype) else code
        return self.cache.get(("d", code, coutp), (w,),
                              lambda old: pack_weight(w.detach(), self.geom, coutp, self.cinp, 1, code, old),
                              desc=(w, self.geom, coutp, self.cinp, 1, code))

    def bias(self, b, n=None):
        n = n or self.coutp
        if b is None:
            return None
        if b.numel() == n:
            return b.detach()
        return self.cache.get(("b", n), (b,), lambda old: pad_vector(b.detach(), n, old))


# ------------------------------------------------------------------------------------------------
# Backward overlap: the weight-gradient GEMMs of a block depend only on tensors that exist before the block's GroupNorm
# backward starts, and the two kinds of kernel bind on different resources (wgrad: tensor pipe / shared-memory bandwidth, ~1.8
# TB/s of DRAM; GroupNorm backward: HBM at 4.4-4.5 TB/s). The wgrad of a conv is therefore enqueued on a side stream right
# after the dgrad that precedes the GroupNorm backward, and co-runs with it (its persistent CTAs leave ~20 KB of shared
# memory and half the register file per SM: room for the GroupNorm CTAs). The side stream is joined before the block's
# backward returns, so nothing outside the block sees it (gradient-arena buckets, autograd, CUDA-graph capture).
# ------------------------------------------------------------------------------------------------
OVERLAP_WGRAD = [os.environ.get("ODVAE_OVERLAP_WGRAD", "1") != "0"]
_SIDE = {}


class side_stream:
    """`with side_stream(device):` -- work inside is enqueued on the device's side stream, ordered after everything enqueued so
    far on the current stream. join_side(device) makes the current stream wait for it. No-ops when overlap is disabled."""

    def __init__(self, device):
        self.device, self.ctx = device, None

    def __enter__(self):
        if not OVERLAP_WGRAD[0]:
            return self
        side = _SIDE.get(self.device.index)
        if side is None:
            # high priority: when a weight-gradient GEMM and a GroupNorm kernel become runnable together, the GEMM's large CTAs
            # (one per SM, ~200 KB of shared memory) must be placed FIRST -- the GroupNorm CTAs then fit beside them, while the
            # other order fills every SM with small CTAs and serialises the two kernels (CUDA-graph kernel nodes inherit it)
            prio = int(os.environ.get("ODVAE_SIDE_PRIORITY", "-1"))
            side = torch.cuda.Stream(self.device, priority=prio)
            _SIDE[self.device.index] = side
        side.wait_stream(torch.cuda.current_stream(self.device))
        self.ctx = torch.cuda.stream(side)
        self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


def join_side(device):
    side = _SIDE.get(device.index)
    if OVERLAP_WGRAD[0] and side is not None:
        torch.cuda.current_stream(device).wait_stream(side)


# Free-running variant (gradient-arena mode, i.e. trainer.TrainStep): the block does NOT wait for its weight-gradient GEMMs;
# the side stream is joined once, by a callback autograd runs at the end of the backward pass (and by the gradient arena
# before it all-reduces a bucket). The wgrad queue then always has work to co-run with whatever HBM-bound kernel the main
# chain is in. Tensors the side stream reads are handed to it with record_stream (they are freed by autograd as soon as the
# block's backward returns); the results live in the arena. Outside the arena the gradients are returned to autograd, which
# may touch them at once, so there the per-block join stays.
OVERLAP_FREE = [os.environ.get("ODVAE_OVERLAP_FREE", "0") != "0"]  # measured: no gain over the per-block join (135.9 ms both)
_JOIN_QUEUED = {}


def side_release(device, direct, *tensors):
    """End of a block's backward: free-running (-> True) when every parameter gradient of the block went straight into the
    arena, else join now (-> False)."""
    side = _SIDE.get(device.index)
    if not OVERLAP_WGRAD[0] or side is None:
        return False
    if not (OVERLAP_FREE[0] and direct and _PROBE[0] == 0):
        join_side(device)
        return False
    for t in tensors:
        if t is not None:
            t.record_stream(side)
    if not _JOIN_QUEUED.get(device.index):
        _JOIN_QUEUED[device.index] = True

        def _final_join():
            _JOIN_QUEUED[device.index] = False
            join_side(device)
        torch.autograd.Variable._execution_engine.queue_callback(_final_join)
    return True


def grad_dst(p):
    """Destination view for a parameter gradient when the flat gradient arena is enabled."""
    return getattr(p, "_odvae_grad_dst", None)


_PROBE = [0]  # > 0 while torch.autograd.grad probes run (PoseLoss.calculate_adaptive_weight)


class probe_gradients:
    """Context: parameter gradients produced inside are returned to autograd as fresh tensors instead of being written
    into the DDP gradient arena, and the arena's bucket hooks stay quiet (ddp.GradArena.boundary_hook)."""

    def __enter__(self):
        _PROBE[0] += 1

    def __exit__(self, *exc):
        _PROBE[0] -= 1
        return False


def _new_grad(p):
    dst = grad_dst(p) if _PROBE[0] == 0 else None
    if dst is not None:
        p._odvae_written = True  # FusedAdam: this parameter has a gradient in this step (ddp.GradArena.begin_step clears)
        return dst, True
    return torch.empty(p.shape, dtype=torch.float32, device=p.device), False


def _ret(g, direct):
    return None if direct else g


# ------------------------------------------------------------------------------------------------
# raw (non-autograd) kernel wrappers
# ------------------------------------------------------------------------------------------------
FUSE_GN_STATS = os.environ.get("ODVAE_FUSE_GN_STATS", "1") != "0"  # conv epilogues also emit the GroupNorm statistics of their output where the kernel serves it


def conv_fwd_raw(x, spec, w, b, residual=None, out_nchw=False, op=None, out=None, gn_stats=False):
    """op: format code of x and the packed weights; out: format code of the NHWC y / residual.
    Defaults: the spec's Format (fwd code for both), else what x's torch dtype implies.
    gn_stats: the caller knows y feeds a GroupNorm(32) next; where the tcgen05 swapped-role tiles serve the shape the
    statistics come out of the conv epilogue and ride on the result as `y._odvae_gn_stats` (see gn_stats_of)."""
    fmt = format_of(x, spec)
    op = fmt.fwd if op is None else op
    out = fmt.fwd if out is None else out
    B, H, W, CinP = x.shape
    assert CinP == spec.cinp, (CinP, spec.cinp)
    Ho, Wo = out_hw(spec.geom, H, W)
    if out_nchw:
        coutp = 16 if spec.cout <= 16 else spec.coutp
        impl = conv_impl(op, B, H, W, CinP, 64, spec.geom) if coutp == 16 else \
            conv_impl(op, B, H, W, CinP, coutp, spec.geom)
        if coutp == 16 and impl == IMPL_TC and spec.geom == CONV_3x3_S1 and Wo % 256 == 0:
            # N = 16 tiles are MMA-issue and L2->SM bound (1.7 ms for decoder.conv_out at batch 64); padded to one 128-row
            # weight tile the layer takes the swapped-role halo-slab tiles (0.9 ms), lanes >= Cout are simply not stored
            coutp = 128
        y = torch.empty(B, spec.cout, Ho, Wo, dtype=torch.float32, device=x.device)
        call("conv2d_fwd", x, spec.w_fwd(w, op, coutp), None if b is None else b.detach(), None, y, B, H, W,
             CinP, coutp, spec.cout, spec.geom, op, out, impl, 1)
        return y
    impl = conv_impl(op, B, H, W, CinP, spec.coutp, spec.geom)
    y = torch.empty(B, Ho, Wo, spec.coutp, dtype=buf_dtype(out, x.dtype), device=x.device)
    if gn_stats and FUSE_GN_STATS and impl == IMPL_TC and op != F32 and spec.coutp == spec.cout:
        nbytes = call("conv2d_fwd_gnstats_workspace_bytes", B, H, W, CinP, spec.coutp, spec.geom, GROUPS)
        if nbytes:
            ws = workspace(nbytes, x.device, "gnconv")
            stats = torch.empty(B, GROUPS, 2, dtype=torch.float32, device=x.device)
            call("conv2d_fwd_gnstats", x, spec.w_fwd(w, op), spec.bias(b), residual, y, stats, B, H, W, CinP,
                 spec.coutp, spec.geom, GROUPS, GN_EPS, op, out, ws, ws.numel())
            y._odvae_gn_stats = stats
            return y
    call("conv2d_fwd", x, spec.w_fwd(w, op), spec.bias(b), residual, y, B, H, W, CinP, spec.coutp, spec.cout,
         spec.geom, op, out, impl, 0)
    return y


def conv_dgrad_raw(dy, spec, w, in_hw, dx_add=None):
    """All operands in the gradient format."""
    g = format_of(dy, spec).grad
    B, Ho, Wo, CoutP = dy.shape
    H, W = in_hw
    impl = conv_impl(g, B, H, W, spec.cinp, CoutP, spec.geom)
    dx = torch.empty(B, H, W, spec.cinp, dtype=dy.dtype, device=dy.device)
    call("conv2d_dgrad", dy, spec.w_dgrad(w, g, CoutP), dx_add, dx, B, H, W, spec.cinp, CoutP, spec.geom, g, impl)
    return dx


def conv_dgrad_gnbwd_raw(dy, spec, w, in_hw, gn_x, gn_stats, gamma, beta):
    """dgrad whose epilogue also produces the GroupNorm+SiLU backward sums for the GroupNorm that made this conv's input
    (include/odvae_b200.h odvae_conv2d_dgrad_gnbwd). -> (dx, (part, nbox)) or None when the shape / mode is not served
    (uniform fp16 mode, tcgen05 swapped-role tiles of the 128-channel stride-1 convs)."""
    fmt = format_of(dy, spec)
    if fmt.grad != F16 or fmt.fwd != F16 or not _GN_BWD_FUSE[0] or gn_x.dtype != torch.float16:
        return None
    B, Ho, Wo, CoutP = dy.shape
    H, W = in_hw
    if conv_impl(F16, B, H, W, spec.cinp, CoutP, spec.geom) != IMPL_TC or gn_x.shape[-1] != spec.cinp:
        return None
    key = (B, H, W, spec.cinp, CoutP, spec.geom)
    nbytes = _GNBWD_BYTES.get(key)
    if nbytes is None:
        nbytes = _GNBWD_BYTES[key] = call("conv2d_dgrad_gnbwd_workspace_bytes", B, H, W, spec.cinp, CoutP, spec.geom, GROUPS)
    if nbytes == 0:
        return None
    dx = torch.empty(B, H, W, spec.cinp, dtype=dy.dtype, device=dy.device)
    part = torch.empty(nbytes // 4, dtype=torch.float32, device=dy.device)
    call("conv2d_dgrad_gnbwd", dy, spec.w_dgrad(w, F16, CoutP), dx, gn_x, gn_stats, gamma.detach(), beta.detach(), part, nbytes,
         B, H, W, spec.cinp, CoutP, spec.geom, GROUPS, F16)
    return dx, (part, nbytes // (8 * B * spec.cinp))


_GN_BWD_FUSE = [os.environ.get("ODVAE_GN_BWD_FUSE", "1") != "0"]
_GNBWD_BYTES = {}


def conv_wgrad_raw(x, dy, spec, w, b):
    """x must already be in the GRADIENT format (mixed mode: the bf16 copy of the fp16 forward operand).
    -> (dw, db) as autograd return values (None where written straight into the gradient arena)."""
    g = format_of(dy, spec).grad
    B, H, W, CinP = x.shape
    CoutP = dy.shape[-1]
    impl = conv_impl(g, B, H, W, CinP, CoutP, spec.geom)
    dw, dw_direct = _new_grad(w)
    db, db_direct = (None, True) if b is None else _new_grad(b)
    nbytes = call("conv2d_wgrad_workspace_bytes", B, H, W, CinP, CoutP, spec.geom, impl)
    ws = workspace(nbytes, x.device, "wgrad")
    call("conv2d_wgrad", x, dy, dw, db, B, H, W, CinP, CoutP, spec.cin, spec.cout, spec.geom, g, impl, 0,
         1.0 / gs_of(spec), ws, ws.numel())
    return _ret(dw, dw_direct), _ret(db, db_direct)


def grad_operand(x, fmt):
    """The copy of a forward activation that a weight-gradient GEMM may contract with dy."""
    return cast16_raw(x, fmt.fwd, fmt.grad) if fmt.mixed else x


def gn_stats_raw(x, C, code=None):
    B, H, W, Cp = x.shape
    assert Cp == C, "GroupNorm runs on unpadded channel counts (C % 64 == 0)"
    code = act_code(x.dtype) if code is None else code
    stats = torch.empty(B, GROUPS, 2, dtype=torch.float32, device=x.device)
    nbytes = call("groupnorm_workspace_bytes", B, H * W, C, GROUPS)
    ws = workspace(nbytes, x.device, "gn")
    call("groupnorm_stats", x, stats, B, H * W, C, GROUPS, GN_EPS, code, ws, ws.numel())
    return stats


def gn_stats_of(x, C, code=None):
    """GroupNorm statistics of x: the ones its producing conv attached (conv_fwd_raw gn_stats=True), else one pass."""
    stats = getattr(x, "_odvae_gn_stats", None)
    if stats is not None and stats.shape[0] == x.shape[0] and x.shape[-1] == C:
        return stats
    return gn_stats_raw(x, C, code)


def gn_apply_raw(x, stats, gamma, beta, silu, in_code=None, out_code=None, copy_code=None):
    """-> y, or (y, y2) when copy_code is given (second copy of the output in that format, same kernel)."""
    B, H, W, C = x.shape
    in_code = act_code(x.dtype) if in_code is None else in_code
    out_code = in_code if out_code is None else out_code
    y = torch.empty(B, H, W, C, dtype=buf_dtype(out_code, x.dtype), device=x.device)
    y2 = None if copy_code is None else torch.empty(B, H, W, C, dtype=buf_dtype(copy_code), device=x.device)
    call("groupnorm_apply", x, stats, gamma.detach(), beta.detach(), y, y2, B, H * W, C, GROUPS, 1 if silu else 0,
         in_code, out_code, copy_code if copy_code is not None else out_code,
         overflow_flags(x.device) if in_code == F16 else None)
    return y if copy_code is None else (y, y2)


def gn_small_ok(HW, C):
    """Shapes odvae_groupnorm_fwd_small serves (include/odvae_b200.h): small fp16 feature maps, one launch."""
    cpg = C // GROUPS
    return HW <= 1024 and C % 64 == 0 and C % GROUPS == 0 and cpg >= 2 and 64 % cpg == 0 and C <= 4096


def gn_forward(x, C, gamma, beta, silu, fmt, out_code=None):
    """-> (stats, y = [silu](GN(x))). Statistics: attached by the producing conv's epilogue if there, else ONE fused launch
    (statistics + apply) on the small fp16 feature maps, else the statistics pass followed by the apply pass."""
    out_code = fmt.fwd if out_code is None else out_code
    B, H, W, Cp = x.shape
    attached = getattr(x, "_odvae_gn_stats", None)
    if (attached is None and fmt.fwd == F16 and out_code == F16 and Cp == C and gn_small_ok(H * W, C)
            and os.environ.get("ODVAE_GN_SMALL", "1") != "0"):
        stats = torch.empty(B, GROUPS, 2, dtype=torch.float32, device=x.device)
        y = torch.empty(B, H, W, C, dtype=buf_dtype(F16, x.dtype), device=x.device)
        call("groupnorm_fwd_small", x, gamma.detach(), beta.detach(), y, stats, B, H * W, C, GROUPS, GN_EPS,
             1 if silu else 0, F16, overflow_flags(x.device))
        return stats, y
    stats = gn_stats_of(x, C, fmt.fwd)
    return stats, gn_apply_raw(x, stats, gamma, beta, silu, fmt.fwd, out_code)


def gn_apply_pair(x, stats, gamma, beta, silu, fmt):
    """(forward operand, weight-gradient operand) of the conv that consumes this GroupNorm output."""
    if fmt.mixed:
        return gn_apply_raw(x, stats, gamma, beta, silu, fmt.fwd, fmt.fwd, fmt.grad)
    y = gn_apply_raw(x, stats, gamma, beta, silu, fmt.fwd, fmt.fwd)
    return y, y


def gn_bwd_raw(x, stats, gamma, beta, dy, silu, dx_add=None, x_code=None, g_code=None, pscale=1.0, partials=None):
    """partials = (part, nbox): the two backward sums per pixel box from the dgrad that produced dy (conv_dgrad_gnbwd_raw):
    no reduce pass over x and dy."""
    B, H, W, C = x.shape
    x_code = act_code(x.dtype) if x_code is None else x_code
    g_code = act_code(dy.dtype) if g_code is None else g_code
    dx = torch.empty_like(dy)
    dgamma, g_direct = _new_grad(gamma)
    dbeta, b_direct = _new_grad(beta)
    nbytes = call("groupnorm_workspace_bytes", B, H * W, C, GROUPS)
    ws = workspace(nbytes, x.device, "gn")
    if partials is not None:
        assert silu and x_code == F16 and g_code == F16
        call("groupnorm_bwd_from_partials", x, stats, gamma.detach(), beta.detach(), dy, dx_add, dx, dgamma, dbeta, partials[0],
             partials[1], B, H * W, C, GROUPS, 0, float(pscale), overflow_flags(x.device), ws, ws.numel())
        return dx, _ret(dgamma, g_direct), _ret(dbeta, b_direct)
    call("groupnorm_bwd", x, stats, gamma.detach(), beta.detach(), dy, dx_add, dx, dgamma, dbeta, B, H * W, C,
         GROUPS, 1 if silu else 0, x_code, g_code, 0, float(pscale),
         overflow_flags(x.device) if g_code == F16 else None, ws, ws.numel())
    return dx, _ret(dgamma, g_direct), _ret(dbeta, b_direct)


def gemm_raw(A, Bm, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, a_mn, b_mn, alpha=1.0, bias=None, residual=None,
             code=None):
    """D[b] = alpha * op(A[b]) op(B[b])^T (+bias)(+residual); 16-bit operands (format `code`, default bf16) -> tcgen05
    with D in the same format (fp32 when D is a float tensor); fp32 operands -> the SIMT engine."""
    if A.dtype in (torch.bfloat16, torch.float16):
        code = act_code(A.dtype) if code is None else code
        call("gemm_16", A, Bm, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, int(a_mn), int(b_mn), code, code,
             F32 if D.dtype == torch.float32 else code, float(alpha), bias, residual)
    else:
        call("gemm_simt", A, Bm, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, int(a_mn), int(b_mn), F32, F32,
             float(alpha), bias, residual)
    return D


def nchw_to_nhwc_raw(x, Cpad, code, scale=1.0, shift=0.0, label=None):
    """code: dtype code of the NHWC result (a torch dtype is accepted too); label: its torch dtype (Format.label)."""
    if isinstance(code, torch.dtype):
        code = act_code(code)
    B, C, H, W = x.shape
    out = torch.empty(B, H, W, Cpad, dtype=buf_dtype(code, label), device=x.device)
    call("nchw_to_nhwc", x, out, B, C, H, W, Cpad, code, float(scale), float(shift))
    return out


def nhwc_to_nchw_raw(x, C, code=None):
    B, H, W, Cpad = x.shape
    out = torch.empty(B, C, H, W, dtype=torch.float32, device=x.device)
    call("nhwc_to_nchw", x, out, B, C, H, W, Cpad, act_code(x.dtype) if code is None else code)
    return out


def _c(t):
    return t if t.is_contiguous() else t.contiguous()


# ------------------------------------------------------------------------------------------------
# layout conversion at the module boundary (AE.py:223; LD/ldm/models/autoencoder.py:344-349)
# ------------------------------------------------------------------------------------------------
class ToNHWC(torch.autograd.Function):
    """gscale: GradScale of the domain the NHWC side belongs to (its gradient arrives scaled and leaves unscaled)."""

    @staticmethod
    def forward(ctx, x, Cpad, compute_dtype, gscale=None):
        fmt = format_for(compute_dtype)
        ctx.C, ctx.fmt, ctx.gscale = x.shape[1], fmt, gscale
        pre = getattr(x, "_odvae_nhwc", None)  # written by rescale_minmax in the same pass as x itself
        if pre is not None and pre[1] == Cpad and pre[2] == fmt.fwd and pre[0].shape[0] == x.shape[0]:
            return pre[0]
        return nchw_to_nhwc_raw(_c(x.float()), Cpad, fmt.fwd, label=fmt.label(fmt.fwd))

    @staticmethod
    def backward(ctx, dy):
        dx = nhwc_to_nchw_raw(_c(dy), ctx.C, ctx.fmt.grad)
        gs = ctx.gscale.value if ctx.gscale is not None else 1.0
        if gs != 1.0:
            dx.mul_(1.0 / gs)
        return dx, None, None, None


class ToNCHW(torch.autograd.Function):
    """gscale: GradScale of the NHWC side (the fp32 NCHW gradient is multiplied by it on the way in)."""

    @staticmethod
    def forward(ctx, x, C, compute_dtype=None, gscale=None):
        fmt = format_for(compute_dtype) if compute_dtype is not None else format_of(x)
        ctx.Cpad, ctx.fmt, ctx.gscale, ctx.label = x.shape[-1], fmt, gscale, x.dtype
        return nhwc_to_nchw_raw(x, C, fmt.fwd)

    @staticmethod
    def backward(ctx, dy):
        gs = ctx.gscale.value if ctx.gscale is not None else 1.0
        return nchw_to_nhwc_raw(_c(dy.float()), ctx.Cpad, ctx.fmt.grad, scale=gs, label=ctx.label), None, None, None


# ------------------------------------------------------------------------------------------------
# Conv2d  (LD model.py:92-106,383-387,428-432,487-491,529-533; Downsample :66-76; quant convs AE.py:88-90)
# ------------------------------------------------------------------------------------------------
class Conv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b, spec, out_nchw, need_dx):
        ctx.spec, ctx.out_nchw, ctx.need_dx = spec, out_nchw, need_dx
        ctx.fmt = format_of(x, spec)
        ctx.save_for_backward(x, w, b)
        B, H, W, _ = x.shape
        ctx.thin = out_nchw and thin_conv_ok(spec, ctx.fmt.grad, B, H, W, spec.cinp, spec.cout)
        return conv_fwd_raw(x, spec, w, b, out_nchw=out_nchw, gn_stats=not out_nchw)

    @staticmethod
    def backward(ctx, dy):
        x, w, b = ctx.saved_tensors
        spec, fmt = ctx.spec, ctx.fmt
        if ctx.thin:
            return Conv._thin_out_backward(ctx, dy, x, w, b)
        if ctx.out_nchw:  # dec_obj gradient arrives NCHW fp32 (masked-L1 backward): bring it to NHWC, times the domain scale
            dy = nchw_to_nhwc_raw(_c(dy.float()), cpad(spec.cout), fmt.grad, scale=gs_of(spec), label=x.dtype)
        else:
            dy = _c(dy)
        dx = None
        # probe mode (torch.autograd.grad of the adaptive GAN weight): only conv_out's weight gradient is wanted
        if ctx.need_dx and ctx.needs_input_grad[0] and not (_PROBE[0] > 0 and ctx.out_nchw):
            dx = conv_dgrad_raw(dy, spec, w, x.shape[1:3])
        dw, db = conv_wgrad_raw(grad_operand(x, fmt), dy, spec, w, b)
        return dx, dw, db, None, None, None

    @staticmethod
    def _thin_out_backward(ctx, dy, x, w, b):
        """decoder.conv_out (Cout <= 4, NCHW fp32 result): the dec_obj gradient [B,C,H,W] fp32 becomes ONE 64-column operand
        col[q, t*4 + co] = gs * dy[co, q - off_t]; dx = col Wd (1x1 forward), (dWd, dbd) = 1x1 weight gradient of (x, col):
        dw[co, i, t] = dWd[t*4 + co, i], db[co] = dbd[4*4 + co] (the centre tap's column sum)."""
        spec, fmt = ctx.spec, ctx.fmt
        g = fmt.grad
        C = spec.cout
        B, H, W, _ = x.shape
        col = thin_im2col_raw(_c(dy.float()), g, flip=True, scale=gs_of(spec), label=x.dtype)
        dx = None
        if ctx.need_dx and ctx.needs_input_grad[0] and _PROBE[0] == 0:
            wd = spec.cache.get(("thin_out_d", g), (w,), lambda old: _thin_weight(w, g, True, old))
            dx = torch.empty(B, H, W, spec.cinp, dtype=col.dtype, device=x.device)
            call("conv2d_fwd", col, wd, None, None, dx, B, H, W, THIN_K, spec.cinp, spec.cin, CONV_1x1, g, g, IMPL_TC, 0)
        dw, dw_direct = _new_grad(w)
        db, db_direct = (None, True) if b is None else _new_grad(b)
        dwp = torch.empty(THIN_K, spec.cin, dtype=torch.float32, device=w.device)
        dbp = torch.empty(THIN_K, dtype=torch.float32, device=w.device)
        _wgrad_1x1(grad_operand(x, fmt), col, spec.cin, THIN_K, g, 1.0 / gs_of(spec), dwp, dbp)
        dw.copy_(dwp.view(16, 4, spec.cin)[:9, :C].permute(1, 2, 0).reshape(dw.shape))
        if db is not None:
            db.copy_(dbp.view(16, 4)[4, :C])
        return dx, _ret(dw, dw_direct), _ret(db, db_direct), None, None, None


# ------------------------------------------------------------------------------------------------
# Thin edge convs: encoder.conv_in (3 -> 128, LD model.py:383-387) and the backward of decoder.conv_out (128 -> 3,
# LD model.py:529-533). Through the generic path the 3 channels are padded to a 64-channel tile and every one of the 9 taps
# costs a full K = 64 GEMM step (conv_in weight gradient 1.0 ms, conv_out dgrad + wgrad 2.5 ms at batch 64, 256x256: 21x padded
# FLOPs, and the weight gradients re-read dy once per filter row). Here the thin side is expanded once by
# odvae_im2col3x3_thin to 9 taps x 4 channels = 36 of 64 columns and the conv is ONE K = 64 step of the 1x1 path:
# y = col W1, dW1 = dy^T col, and for conv_out dx = col(dy) Wd, dWd = col(dy)^T x -- all HBM-bound (0.48 / 1.75 ms).
# ------------------------------------------------------------------------------------------------
THIN_CONV = os.environ.get("ODVAE_THIN_CONV", "1") != "0"
THIN_K = 64


def thin_conv_ok(spec, code, B, H, W, wide_p, thin_c):
    return (THIN_CONV and spec.geom == CONV_3x3_S1 and thin_c <= 4 and code != F32 and
            conv_impl(code, B, H, W, THIN_K, wide_p, CONV_1x1) == IMPL_TC)


def thin_im2col_raw(x, code, flip=False, scale=1.0, label=None):
    """NCHW fp32 [B,C<=4,H,W] -> [B,H,W,64] of format `code`, column t*4 + c = scale * x[c] shifted by tap t (flip: by -t)."""
    B, C, H, W = x.shape
    col = torch.empty(B, H, W, THIN_K, dtype=buf_dtype(code, label), device=x.device)
    call("im2col3x3_thin", x, col, B, C, H, W, 1 if flip else 0, float(scale), code)
    return col


def _thin_weight(w, code, transpose, old=None):
    """Packed 1x1 weight [wide][64] over the im2col columns. conv_in (transpose False): w [wide, C, 3, 3], column t*4 + c =
    w[o, c, t]. conv_out data gradient (transpose True): w [C, wide, 3, 3], column t*4 + co = w[co, i, t]. `old`: the previous
    packed copy, rewritten in place (captured CUDA graphs keep reading that buffer; ParamCache / refresh_param_caches)."""
    wd = w.detach().float()
    C = wd.shape[0] if transpose else wd.shape[1]
    wide = wd.shape[1] if transpose else wd.shape[0]
    w1 = torch.zeros(wide, 16, 4, dtype=torch.float32, device=w.device)
    w1[:, :9, :C] = (wd.permute(1, 2, 3, 0) if transpose else wd.permute(0, 2, 3, 1)).reshape(wide, 9, C)
    return pack_weight(w1.view(wide, THIN_K, 1, 1), CONV_1x1, cpad(wide), THIN_K, 0, code, old)


def _wgrad_1x1(x, dy, cin, cout, g, out_scale, dw, db):
    B, H, W, CinP = x.shape
    CoutP = dy.shape[-1]
    impl = conv_impl(g, B, H, W, CinP, CoutP, CONV_1x1)
    nbytes = call("conv2d_wgrad_workspace_bytes", B, H, W, CinP, CoutP, CONV_1x1, impl)
    ws = workspace(nbytes, x.device, "wgrad")
    call("conv2d_wgrad", x, dy, dw, db, B, H, W, CinP, CoutP, cin, cout, CONV_1x1, g, impl, 0, float(out_scale), ws,
         ws.numel())


class ThinConvIn(torch.autograd.Function):
    """x: NCHW fp32 image (no gradient wanted) -> NHWC [B,H,W,coutp] in the spec's forward format (+ GroupNorm statistics)."""

    @staticmethod
    def forward(ctx, x, w, b, spec):
        fmt = spec.fmt
        op = fmt.fwd
        B, C, H, W = x.shape
        col = thin_im2col_raw(_c(x.float()), op, label=fmt.label(op))
        w1 = spec.cache.get(("thin_in", op), (w,), lambda old: _thin_weight(w, op, False, old))
        y = torch.empty(B, H, W, spec.coutp, dtype=col.dtype, device=x.device)
        nbytes = 0
        if FUSE_GN_STATS and spec.coutp == spec.cout:
            nbytes = call("conv2d_fwd_gnstats_workspace_bytes", B, H, W, THIN_K, spec.coutp, CONV_1x1, GROUPS)
        if nbytes:
            ws = workspace(nbytes, x.device, "gnconv")
            stats = torch.empty(B, GROUPS, 2, dtype=torch.float32, device=x.device)
            call("conv2d_fwd_gnstats", col, w1, spec.bias(b), None, y, stats, B, H, W, THIN_K, spec.coutp, CONV_1x1,
                 GROUPS, GN_EPS, op, op, ws, ws.numel())
            y._odvae_gn_stats = stats
        else:
            call("conv2d_fwd", col, w1, spec.bias(b), None, y, B, H, W, THIN_K, spec.coutp, spec.cout, CONV_1x1, op, op,
                 IMPL_TC, 0)
        ctx.spec, ctx.C = spec, C
        ctx.save_for_backward(grad_operand(col, fmt), w, b)
        return y

    @staticmethod
    def backward(ctx, dy):
        col, w, b = ctx.saved_tensors
        spec, C = ctx.spec, ctx.C
        g = spec.fmt.grad
        dw, dw_direct = _new_grad(w)
        db, db_direct = (None, True) if b is None else _new_grad(b)
        dw1 = torch.empty(spec.cout, THIN_K, dtype=torch.float32, device=w.device)
        _wgrad_1x1(col, _c(dy), THIN_K, spec.cout, g, 1.0 / gs_of(spec), dw1, db)
        dw.copy_(dw1.view(spec.cout, 16, 4)[:, :9, :C].permute(0, 2, 1).reshape(dw.shape))
        return None, _ret(dw, dw_direct), _ret(db, db_direct), None


# ------------------------------------------------------------------------------------------------
# GroupNorm (+SiLU) as a standalone node (norm_out before conv_out: LD model.py:453-455,563-565)
# ------------------------------------------------------------------------------------------------
class GroupNormAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, silu, fmt=None, gscale=None):
        fmt = fmt if fmt is not None else format_of(x)
        stats, y = gn_forward(x, x.shape[-1], gamma, beta, silu, fmt)
        ctx.silu, ctx.fmt, ctx.gscale = silu, fmt, gscale
        ctx.save_for_backward(x, stats, gamma, beta)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, stats, gamma, beta = ctx.saved_tensors
        gs = ctx.gscale.value if ctx.gscale is not None else 1.0
        dx, dg, db = gn_bwd_raw(x, stats, gamma, beta, _c(dy), ctx.silu, x_code=ctx.fmt.fwd, g_code=ctx.fmt.grad,
                                pscale=1.0 / gs)
        return dx, dg, db, None, None, None


# ------------------------------------------------------------------------------------------------
# ResnetBlock  (LD model.py:121-141, temb=None, dropout 0):
#   out = shortcut(x) + conv2(silu(GN2(conv1(silu(GN1(x))))))
# Saved for backward: x, h1 (GroupNorm inputs, forward format) and a1, a2 in the GRADIENT format (they are only
# needed again as weight-gradient operands; the mixed mode's GN apply kernel writes that copy alongside the fp16 one).
# ------------------------------------------------------------------------------------------------
class ResBlock(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, g1, b1, w1, c1b, g2, b2, w2, c2b, wn, nb, blk):
        fmt = format_of(x, blk)
        if fmt.mixed:
            s1 = gn_stats_of(x, blk.cin, fmt.fwd)
            a1, a1g = gn_apply_pair(x, s1, g1, b1, True, fmt)
        else:
            s1, a1 = gn_forward(x, blk.cin, g1, b1, True, fmt)
            a1g = a1
        h1 = conv_fwd_raw(a1, blk.spec1, w1, c1b, op=fmt.fwd, out=fmt.fwd, gn_stats=True)
        del a1
        if fmt.mixed:
            s2 = gn_stats_of(h1, blk.cout, fmt.fwd)
            a2, a2g = gn_apply_pair(h1, s2, g2, b2, True, fmt)
        else:
            s2, a2 = gn_forward(h1, blk.cout, g2, b2, True, fmt)
            a2g = a2
        sc = x if wn is None else conv_fwd_raw(x, blk.spec_nin, wn, nb, op=fmt.fwd, out=fmt.fwd)
        out = conv_fwd_raw(a2, blk.spec2, w2, c2b, residual=sc, op=fmt.fwd, out=fmt.fwd, gn_stats=True)
        ctx.blk, ctx.fmt = blk, fmt
        ctx.has_nin = wn is not None
        ctx.save_for_backward(x, s1, a1g, h1, s2, a2g, g1, b1, w1, c1b, g2, b2, w2, c2b, wn, nb)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, s1, a1, h1, s2, a2, g1, b1, w1, c1b, g2, b2, w2, c2b, wn, nb = ctx.saved_tensors
        blk, fmt = ctx.blk, ctx.fmt
        dout = _c(dout)
        hw = x.shape[1:3]
        dwn = dnb = None
        if ctx.has_nin:
            dsc = conv_dgrad_raw(dout, blk.spec_nin, wn, hw)
            dwn, dnb = conv_wgrad_raw(grad_operand(x, fmt), dout, blk.spec_nin, wn, nb)
        else:
            dsc = dout
        dev = x.device
        # (128-channel layers, fp16 mode: the dgrad epilogue also produces the GroupNorm backward's two sums -> no reduce pass)
        fused = conv_dgrad_gnbwd_raw(dout, blk.spec2, w2, hw, h1, s2, g2, b2)
        da2, part2 = fused if fused is not None else (conv_dgrad_raw(dout, blk.spec2, w2, hw), None)
        with side_stream(dev):  # co-runs with the GroupNorm backward below
            dw2, dc2b = conv_wgrad_raw(a2, dout, blk.spec2, w2, c2b)
        ps = 1.0 / gs_of(blk)
        dh1, dg2, db2 = gn_bwd_raw(h1, s2, g2, b2, da2, True, x_code=fmt.fwd, g_code=fmt.grad, pscale=ps, partials=part2)
        fused = conv_dgrad_gnbwd_raw(dh1, blk.spec1, w1, hw, x, s1, g1, b1)
        da1, part1 = fused if fused is not None else (conv_dgrad_raw(dh1, blk.spec1, w1, hw), None)
        with side_stream(dev):
            dw1, dc1b = conv_wgrad_raw(a1, dh1, blk.spec1, w1, c1b)
        dx, dg1, db1 = gn_bwd_raw(x, s1, g1, b1, da1, True, dx_add=dsc, x_code=fmt.fwd, g_code=fmt.grad, pscale=ps,
                                  partials=part1)
        # the side stream read a1, a2, dout, dh1: wait for it here, or (arena mode) let it run on and join at the end of backward
        side_release(dev, dw1 is None and dw2 is None and (c1b is None or dc1b is None) and (c2b is None or dc2b is None),
                     a1, a2, dout, dh1)
        return dx, dg1, db1, dw1, dc1b, dg2, db2, dw2, dc2b, dwn, dnb, None


# ------------------------------------------------------------------------------------------------
# AttnBlock  (LD model.py:178-202): out = x + proj(softmax(q^T k / sqrt(C)) v), single head
# q, k, v projections run as ONE 1x1 conv with 3C output channels; the residual add is fused into
# the proj_out epilogue. Unfused core: S (fp32) and P (act dtype) are materialised per block.
# ------------------------------------------------------------------------------------------------
class AttnBlock(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, wq, bq, wk, bk, wv, bv, wp, bp, blk):
        B, H, W, C = x.shape
        N = H * W
        fmt = format_of(x, blk)
        ac = fmt.attn            # format code inside the block
        dt = buf_dtype(ac, x.dtype)
        stats, hn = gn_forward(x, C, gamma, beta, False, fmt, out_code=ac)
        wqkv, bqkv = blk.fused_qkv(wq, bq, wk, bk, wv, bv, ac)
        qkv = torch.empty(B, H, W, 3 * C, dtype=dt, device=x.device)
        impl = conv_impl(ac, B, H, W, C, 3 * C, CONV_1x1)
        call("conv2d_fwd", hn, wqkv, bqkv, None, qkv, B, H, W, C, 3 * C, 3 * C, CONV_1x1, ac, ac, impl, 0)
        q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
        P = torch.empty(B, N, N, dtype=dt, device=x.device)
        scale = float(int(C) ** (-0.5))
        if attn_fused(ac, N, C):
            # q k^T and the softmax in two tcgen05 GEMM passes; the fp32 scores never reach HBM
            nbytes = call("attn_softmax_workspace_bytes", B, N)
            ws = workspace(nbytes, x.device, "attn")
            call("attn_scores_softmax", q, k, P, B, N, C, 3 * C, 3 * C, N * 3 * C, N * 3 * C, scale, ac, ws,
                 ws.numel())
        else:
            S = torch.empty(B, N, N, dtype=torch.float32, device=x.device)
            gemm_raw(q, k, S, N, N, C, B, 3 * C, 3 * C, N, N * 3 * C, N * 3 * C, N * N, False, False, code=ac)
            call("softmax_rows", S, P, B * N, N, scale, ac)
            del S
        O = torch.empty(B, H, W, C, dtype=dt, device=x.device)
        gemm_raw(P, v, O, N, C, N, B, N, 3 * C, C, N * N, N * 3 * C, N * C, False, True, code=ac)
        out = conv_fwd_raw(O, blk.spec_proj, wp, bp, residual=x, op=ac, out=fmt.fwd)  # adds into the residual stream
        ctx.blk, ctx.fmt = blk, fmt
        ctx.save_for_backward(x, stats, hn, qkv, P, O, gamma, beta, wq, bq, wk, bk, wv, bv, wp, bp)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, stats, hn, qkv, P, O, gamma, beta, wq, bq, wk, bk, wv, bv, wp, bp = ctx.saved_tensors
        blk = ctx.blk
        B, H, W, C = x.shape
        N = H * W
        fmt = ctx.fmt
        ac = fmt.attn
        dt = buf_dtype(ac, x.dtype)
        dout = _c(dout)
        scale = float(int(C) ** (-0.5))
        q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
        dO = conv_dgrad_raw(dout, blk.spec_proj, wp, (H, W))
        dwp, dbp = conv_wgrad_raw(O, dout, blk.spec_proj, wp, bp)
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = dqkv[..., :C], dqkv[..., C:2 * C], dqkv[..., 2 * C:]
        # dV[j,c] = sum_i P[i,j] dO[i,c]
        gemm_raw(P, dO, dv, N, C, N, B, N, C, 3 * C, N * N, N * C, N * 3 * C, True, True, code=ac)
        # dP[i,j] = sum_c dO[i,c] v[j,c]
        dS = torch.empty(B, N, N, dtype=dt, device=x.device)
        # fp16: dS = P (dP - delta) / sqrt(C) is ~1/N of the (already scaled) gradient; it is stored times DS_SCALE so that it
        # stays in fp16's normal range, and the dQ / dK GEMMs divide the factor out again
        ds_scale = DS_SCALE_FP16 if ac == F16 else 1.0
        if attn_fused(ac, N, C):
            # delta_i = sum_j P_ij dP_ij = sum_c dO_ic O_ic; dS = P * (dO v^T - delta) * scale in the GEMM epilogue
            delta = torch.empty(B * N, dtype=torch.float32, device=x.device)
            call("rowdot", dO, O, delta, B * N, C, C, C, ac)
            call("attn_dsoftmax", dO, v, P, delta, dS, B, N, C, C, 3 * C, N * C, N * 3 * C, scale * ds_scale, ac)
        else:
            dP = torch.empty(B, N, N, dtype=torch.float32, device=x.device)
            gemm_raw(dO, v, dP, N, N, C, B, C, 3 * C, N, N * C, N * 3 * C, N * N, False, False, code=ac)
            call("softmax_rows_bwd", dP, P, dS, B * N, N, scale * ds_scale, ac)
            del dP
        # dQ[i,c] = sum_j dS[i,j] k[j,c] ; dK[j,c] = sum_i dS[i,j] q[i,c]
        gemm_raw(dS, k, dq, N, C, N, B, N, 3 * C, 3 * C, N * N, N * 3 * C, N * 3 * C, False, True, alpha=1.0 / ds_scale,
                 code=ac)
        gemm_raw(dS, q, dk, N, C, N, B, N, 3 * C, 3 * C, N * N, N * 3 * C, N * 3 * C, True, True, alpha=1.0 / ds_scale,
                 code=ac)
        del dS
        # fused qkv projection backward: dgrad as a GEMM against the forward packing ([3C][C], N-major view)
        wqkv, _ = blk.fused_qkv(wq, bq, wk, bk, wv, bv, ac)
        dhn = torch.empty_like(hn)
        M = B * N
        gemm_raw(dqkv, wqkv, dhn, M, C, 3 * C, 1, 3 * C, C, C, 0, 0, 0, False, True, code=ac)
        dwqkv, dbqkv, direct = blk.qkv_grad_buffers(wq, bq)
        impl = conv_impl(ac, B, H, W, C, 3 * C, CONV_1x1)
        nbytes = call("conv2d_wgrad_workspace_bytes", B, H, W, C, 3 * C, CONV_1x1, impl)
        ps = 1.0 / gs_of(blk)
        with side_stream(x.device):  # co-runs with the GroupNorm backward below
            ws = workspace(nbytes, x.device, "wgrad")
            call("conv2d_wgrad", hn, dqkv, dwqkv, dbqkv, B, H, W, C, 3 * C, C, 3 * C, CONV_1x1, ac, impl, 0, ps, ws,
                 ws.numel())
        dx, dg, db = gn_bwd_raw(x, stats, gamma, beta, dhn, False, dx_add=dout, x_code=fmt.fwd, g_code=fmt.grad,
                                pscale=ps)
        join_side(x.device)  # the q|k|v gradient is copied out of dwqkv right below
        if grad_dst(wq) is not None:  # arena mode: device-to-device copies of the three row blocks
            for i, (w_, b_) in enumerate(((wq, bq), (wk, bk), (wv, bv))):
                grad_dst(w_).copy_(dwqkv[i * C:(i + 1) * C].view(C, C, 1, 1))
                grad_dst(b_).copy_(dbqkv[i * C:(i + 1) * C])
                w_._odvae_written = b_._odvae_written = True
            gq = gk = gv = gbq = gbk = gbv = None
        else:
            gq, gk, gv = (dwqkv[i * C:(i + 1) * C].view(C, C, 1, 1) for i in range(3))
            gbq, gbk, gbv = (dbqkv[i * C:(i + 1) * C] for i in range(3))
        return dx, dg, db, gq, gbq, gk, gbk, gv, gbv, dwp, dbp, None


# ------------------------------------------------------------------------------------------------
# Upsample  (LD model.py:53-57): nearest 2x, then 3x3 conv.  Only the low-resolution input is kept.
# ------------------------------------------------------------------------------------------------
def upsample2x_raw(x):
    """Nearest 2x: a pure copy, so any 16-bit format goes through the 16-bit kernel."""
    B, H, W, C = x.shape
    y = torch.empty(B, 2 * H, 2 * W, C, dtype=x.dtype, device=x.device)
    call("upsample2x", x, y, B, H, W, C, act_code(x.dtype))
    return y


UPCONV_PHASES = os.environ.get("ODVAE_UPCONV_PHASES", "1") != "0"  # 0: the literal form (upsample2x kernel + 3x3 conv), for A/B
_UP_OK = {}


def upconv_impl(code, B, H, W, CinP, CoutP):
    if code == F32:
        return IMPL_SIMT
    key = (B, H, W, CinP, CoutP)
    ok = _UP_OK.get(key)
    if ok is None:
        ok = bool(call("upconv_tc_supported", B, H, W, CinP, CoutP))
        _UP_OK[key] = ok
    return IMPL_TC if ok else IMPL_SIMT


def upconv_weight(spec, w, code, which):
    """Pre-summed parity-phase weights of an Upsample conv (include/odvae_b200.h odvae_pack_upconv_weight)."""
    def build(old):
        out = old if old is not None else torch.empty(16 * spec.coutp * spec.cinp, dtype=buf_dtype(code), device=w.device)
        call("pack_upconv_weight", w.detach(), out, spec.cout, spec.cin, spec.coutp, spec.cinp, which, code)
        return out
    return spec.cache.get(("up", which, code), (w,), build)


class UpsampleConv(torch.autograd.Function):
    """nearest-2x upsample + 3x3 conv as four parity-phase 2x2 convs over the LOW-resolution input: no upsampled
    tensor in HBM, 16/36 of the FLOPs in forward, dgrad (one 16-tap launch) and wgrad (four 4-tap GEMMs folded onto 3x3)."""

    @staticmethod
    def forward(ctx, x, w, b, spec):
        fmt = format_of(x, spec)
        ctx.spec, ctx.fmt = spec, fmt
        ctx.save_for_backward(x, w, b)
        if not UPCONV_PHASES:
            return conv_fwd_raw(upsample2x_raw(x), spec, w, b, gn_stats=True)
        B, H, W, CinP = x.shape
        code = fmt.fwd
        impl = upconv_impl(code, B, H, W, CinP, spec.coutp)
        y = torch.empty(B, 2 * H, 2 * W, spec.coutp, dtype=buf_dtype(code, x.dtype), device=x.device)
        call("upconv_fwd", x, upconv_weight(spec, w, code, 0), spec.bias(b), y, B, H, W, CinP, spec.coutp, code, impl)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, b = ctx.saved_tensors
        spec, fmt = ctx.spec, ctx.fmt
        dy = _c(dy)
        B, H, W, C = x.shape
        if not UPCONV_PHASES:
            u = upsample2x_raw(grad_operand(x, fmt))  # re-materialised in the gradient format for the wgrad GEMM
            dw, db = conv_wgrad_raw(u, dy, spec, w, b)
            del u
            du = conv_dgrad_raw(dy, spec, w, (2 * H, 2 * W))
            dx = torch.empty(B, H, W, C, dtype=dy.dtype, device=dy.device)
            call("upsample2x_bwd", du, dx, B, H, W, C, fmt.grad)
            return dx, dw, db, None
        g = fmt.grad
        impl = upconv_impl(g, B, H, W, spec.cinp, spec.coutp)
        dx = torch.empty(B, H, W, spec.cinp, dtype=dy.dtype, device=dy.device)
        call("upconv_dgrad", dy, upconv_weight(spec, w, g, 1), dx, B, H, W, spec.cinp, spec.coutp, g, impl)
        dw, dw_direct = _new_grad(w)
        db, db_direct = (None, True) if b is None else _new_grad(b)
        nbytes = call("upconv_wgrad_workspace_bytes", B, H, W, spec.cinp, spec.coutp, spec.cin, spec.cout, impl)
        ws = workspace(nbytes, x.device, "upwgrad")
        call("upconv_wgrad", grad_operand(x, fmt), dy, dw, db, B, H, W, spec.cinp, spec.coutp, spec.cin, spec.cout, g, impl, 0,
             1.0 / gs_of(spec), ws, ws.numel())
        return dx, _ret(dw, dw_direct), _ret(db, db_direct), None


# ------------------------------------------------------------------------------------------------
# DiagonalGaussian sample + dropout + noise + pose-feature add + KL, one kernel each way
# (LD distributions.py:24-46; AE.py:226-241,253)
# ------------------------------------------------------------------------------------------------
class SampleKL(torch.autograd.Function):
    @staticmethod
    def forward(ctx, moments, enc_pose, eps, keep, noise, keep_scale, Cz, CzPad, want_z, fmt=None, gs_dec=None, gs_enc=None,
                keep_scale_dev=None):
        """gs_dec / gs_enc: GradScale of the decoder side (the scale dz arrives with) and of the encoder side (the scale
        dmoments leaves with). keep_scale_dev: device float [1] that overrides keep_scale (CUDA-graph replay)."""
        fmt = fmt if fmt is not None else format_of(moments)
        ctx.fmt, ctx.gs = fmt, (gs_dec, gs_enc)
        B, H, W, ld = moments.shape
        HW = H * W
        dev = moments.device
        mean = torch.empty(B, Cz, H, W, dtype=torch.float32, device=dev)
        logvar = torch.empty(B, Cz, H, W, dtype=torch.float32, device=dev)
        kl = torch.empty(B, dtype=torch.float32, device=dev)
        z = torch.empty(B, H, W, CzPad, dtype=moments.dtype, device=dev) if want_z else None
        call("sample_kl_fwd", moments, eps, keep, noise, enc_pose, float(keep_scale), mean, logvar, z, kl, B, HW, Cz,
             CzPad, ld, fmt.fwd, keep_scale_dev)
        ctx.cfg = (Cz, CzPad, ld, float(keep_scale), moments.dtype, enc_pose is not None)
        ctx.keep_scale_dev = keep_scale_dev
        ctx.save_for_backward(mean, logvar, eps, keep)
        if z is None:
            z = torch.empty(0, device=dev)
        return mean, logvar, z, kl

    @staticmethod
    def backward(ctx, dmean, dlogvar, dz, dkl):
        mean, logvar, eps, keep = ctx.saved_tensors
        Cz, CzPad, ld, keep_scale, dtype, has_pose = ctx.cfg
        B, _, H, W = mean.shape
        dev = mean.device
        if dz is not None and dz.numel() == 0:
            dz = None
        dz = None if dz is None else _c(dz)
        dkl = None if dkl is None else _c(dkl.float())
        dmean = None if dmean is None else _c(dmean.float())
        dlogvar = None if dlogvar is None else _c(dlogvar.float())
        dmom = torch.empty(B, H, W, ld, dtype=dtype, device=dev)
        dpose = torch.empty(B, Cz * H * W, dtype=torch.float32, device=dev) if (has_pose and dz is not None) else None
        gs_dec = ctx.gs[0].value if ctx.gs[0] is not None else 1.0
        gs_enc = ctx.gs[1].value if ctx.gs[1] is not None else 1.0
        call("sample_kl_bwd", dz, dkl, dmean, dlogvar, mean, logvar, eps, keep, keep_scale, dmom, dpose, B, H * W, Cz,
             CzPad, ld, ctx.fmt.grad, 1.0 / gs_dec, gs_enc, ctx.keep_scale_dev)
        if has_pose and dpose is None:
            dpose = torch.zeros(B, Cz * H * W, dtype=torch.float32, device=dev)
        return dmom, dpose, None, None, None, None, None, None, None, None, None, None, None


# ------------------------------------------------------------------------------------------------
# Masked L1 reconstruction sums (PoseLoss._get_rec_loss / mask multiplies, LOSS.py:134-145,250-257)
# ------------------------------------------------------------------------------------------------
class L1Masked(torch.autograd.Function):
    """sums[b] = sum_{c,p} |x*m - xhat*m|.  Gradient flows to xhat only (x is the ground truth)."""

    # Parity-test hook (tests/test_parity_gpu.py): a reconstruction whose sign pattern sign(x*m - ref*m) selects the
    # subgradient of |.| instead of our own xhat's. |x - xhat| has a kink at 0: two forwards that differ by 1e-3 flip the
    # sign of ~0.3 % of the pixels, which alone moves every upstream gradient by ~2e-2 whatever the backward's precision.
    sign_reference = None

    @staticmethod
    def forward(ctx, x, xhat, mask2d):
        B, C, H, W = xhat.shape
        HW = H * W
        if mask2d.numel() != B * HW:
            raise RuntimeError("mask_2d_bbox must hold one value per pixel ([B,1,H,W] or [B,H,W]), got %s for images %s"
                               % (tuple(mask2d.shape), tuple(xhat.shape)))
        sums = torch.empty(B, dtype=torch.float32, device=xhat.device)
        sgn = torch.empty(B, C, H, W, dtype=torch.int8, device=xhat.device)
        nbytes = call("l1_masked_workspace_bytes", B, C, HW)
        ws = workspace(nbytes, xhat.device, "l1")
        call("l1_masked_fwd", x, xhat, mask2d, sums, sgn, B, C, HW, ws, ws.numel())
        ref = L1Masked.sign_reference
        if ref is not None:
            scratch = torch.empty_like(sums)
            call("l1_masked_fwd", x, _c(ref.to(xhat)), mask2d, scratch, sgn, B, C, HW, ws, ws.numel())
        ctx.save_for_backward(sgn, mask2d)
        return sums

    @staticmethod
    def backward(ctx, dsums):
        sgn, mask2d = ctx.saved_tensors
        B, C, H, W = sgn.shape
        dxhat = torch.empty(B, C, H, W, dtype=torch.float32, device=sgn.device)
        call("l1_masked_bwd", sgn, mask2d, _c(dsums.float()), dxhat, B, C, H * W)
        return None, dxhat, None


# ------------------------------------------------------------------------------------------------
# nn.Linear / activations of the pose MLPs (pose_decoder.py:82-97, pose_encoder.py:82-131), fp32 SIMT GEMM
# ------------------------------------------------------------------------------------------------
class Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        M, K = x.shape
        N = w.shape[0]
        if w.shape[1] != K:  # F.linear's own error (e.g. 512x512 crops into the config's 16x16x16 pose MLP)
            raise RuntimeError("mat1 and mat2 shapes cannot be multiplied (%dx%d and %dx%d)" % (M, K, w.shape[1], N))
        y = torch.empty(M, N, dtype=torch.float32, device=x.device)
        gemm_raw(x, w.detach(), y, M, N, K, 1, K, K, N, 0, 0, 0, False, False, bias=None if b is None else b.detach())
        ctx.save_for_backward(x, w, b)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w, b = ctx.saved_tensors
        dy = _c(dy.float())
        M, K = x.shape
        N = w.shape[0]
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(M, K, dtype=torch.float32, device=x.device)
            gemm_raw(dy, w.detach(), dx, M, K, N, 1, N, K, K, 0, 0, 0, False, True)
        dw, dw_direct = _new_grad(w)
        gemm_raw(dy, x, dw, N, K, M, 1, N, K, K, 0, 0, 0, True, True)
        db = None
        if b is not None:
            db, db_direct = _new_grad(b)
            nbytes = call("colsum_workspace_bytes", N)
            ws = workspace(nbytes, x.device, "colsum")
            call("colsum", dy, M, N, N, F32, db, 0, ws, ws.numel())
            db = _ret(db, db_direct)
        return dx, _ret(dw, dw_direct), db


class Act(torch.autograd.Function):
    """act = 1 tanh, 2 silu on fp32 tensors."""

    @staticmethod
    def forward(ctx, x, act):
        x = _c(x)
        y = torch.empty_like(x)
        call("act_fwd", x, y, x.numel(), act)
        ctx.act = act
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        dx = torch.empty_like(x)
        call("act_bwd", x, _c(dy), dx, x.numel(), ctx.act)
        return dx, None


def rescale_minmax(x, nhwc=None):
    """AE.py:434-436 `_rescale`: 2*(x-min)/(max-min)-1 over the whole batch tensor; a two-stage min/max reduction and ONE
    pass that writes the rescaled NCHW fp32 tensor and, with nhwc = (Cpad, code, label), the NHWC 16-bit operand of
    encoder.conv_in as well; that copy rides on the result as `out._odvae_nhwc` and ToNHWC picks it up. No host
    sync."""
    x = _c(x.float())
    mm = torch.empty(2, dtype=torch.float32, device=x.device)
    ws = workspace(1 << 16, x.device, "minmax")
    call("minmax", x, x.numel(), mm, ws, ws.numel())
    out = torch.empty_like(x)
    if nhwc is not None and x.dim() == 4:
        Cpad, code, label = nhwc
        B, C, H, W = x.shape
        y = torch.empty(B, H, W, Cpad, dtype=buf_dtype(code, label), device=x.device)
        call("rescale_nhwc", x, mm, out, y, B, C, H, W, Cpad, code)
        out._odvae_nhwc = (y, Cpad, code)
    else:
        call("rescale", x, out, x.numel(), mm)
    return out
