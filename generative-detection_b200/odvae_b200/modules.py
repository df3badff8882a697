"""nn.Modules with the reference's parameter names and shapes over the kernels of ops.py.

Mirrors LD/ldm/modules/diffusionmodules/model.py (Encoder :368-459, Decoder :462-568, ResnetBlock
:82-141, AttnBlock :150-202, Upsample :42-57, Downsample :60-79) so that reference checkpoints load
unchanged (state_dict keys `encoder.down.{l}.block.{i}.norm1.weight`, ...). Parameters are fp32 in the
reference layouts; the packed bf16 copies the kernels consume are private caches, not state.

Activations between modules are NHWC tensors [B,H,W,Cp] in `compute_dtype` (see ops.py); the public
`Encoder.forward` / `Decoder.forward` take and return NCHW fp32 like the reference.
"""
import math

import torch
import torch.nn as nn

from . import ops
from .lib import CONV_1x1, CONV_3x3_S1, CONV_3x3_S2, require_device


def _check_cuda(x):
    if not x.is_cuda:
        raise RuntimeError("odvae_b200 modules run on a B200 (sm_100a) CUDA device only; there is no CPU fallback "
                           "(the CPU restatement lives under oracle/ and is test infrastructure)")
    require_device()


_ARENA = [None]  # set by odvae_b200.ddp users (TrainStep): bucket-boundary hooks are attached while it is set


def _run(mod, h, *args, **kw):
    """Call `mod(h)`; when a gradient arena is active, hook the input activation so that the bucket(s) holding
    the module's parameter gradients are all-reduced as soon as its backward has been enqueued."""
    arena = _ARENA[0]
    if arena is not None and h.requires_grad:
        first = next(mod.parameters(), None)
        if first is not None and id(first) in arena._index:
            h.register_hook(arena.boundary_hook(first))
    return mod(h, *args, **kw)


class Conv2d(nn.Module):
    """Parameter container + geometry of one torch.nn.Conv2d of the reference (same init as torch)."""

    def __init__(self, cin, cout, kernel_size, geom, cout_pad=None):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(cout, cin, kernel_size, kernel_size))
        self.bias = nn.Parameter(torch.empty(cout))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        bound = 1.0 / math.sqrt(cin * kernel_size * kernel_size)
        nn.init.uniform_(self.bias, -bound, bound)
        self.spec = ops.ConvSpec(cin, cout, geom, cout_pad)

    def forward(self, x, out_nchw=False, need_dx=True):
        return ops.Conv.apply(x, self.weight, self.bias, self.spec, out_nchw, need_dx)


class GroupNorm(nn.Module):
    """Normalize(): GroupNorm(32, C, eps=1e-6, affine=True)  (LD model.py:38-39)."""

    def __init__(self, channels):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(channels))
        self.bias = nn.Parameter(torch.zeros(channels))

    fmt = None  # set by ops.assign_format()

    def forward(self, x, silu=False):
        return ops.GroupNormAct.apply(x, self.weight, self.bias, silu, self.fmt, getattr(self, "gscale", None))


class ResnetBlock(nn.Module):
    def __init__(self, in_channels, out_channels=None):
        super().__init__()
        out_channels = in_channels if out_channels is None else out_channels
        self.cin, self.cout = in_channels, out_channels
        self.norm1 = GroupNorm(in_channels)
        self.conv1 = Conv2d(in_channels, out_channels, 3, CONV_3x3_S1)
        self.norm2 = GroupNorm(out_channels)
        self.conv2 = Conv2d(out_channels, out_channels, 3, CONV_3x3_S1)
        if in_channels != out_channels:
            self.nin_shortcut = Conv2d(in_channels, out_channels, 1, CONV_1x1)
        self.spec1, self.spec2 = self.conv1.spec, self.conv2.spec
        self.spec_nin = self.nin_shortcut.spec if in_channels != out_channels else None

    def forward(self, x):
        nin = getattr(self, "nin_shortcut", None)
        return ops.ResBlock.apply(x, self.norm1.weight, self.norm1.bias, self.conv1.weight, self.conv1.bias,
                                  self.norm2.weight, self.norm2.bias, self.conv2.weight, self.conv2.bias,
                                  None if nin is None else nin.weight, None if nin is None else nin.bias, self)


class AttnBlock(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.channels = channels
        self.norm = GroupNorm(channels)
        self.q = Conv2d(channels, channels, 1, CONV_1x1)
        self.k = Conv2d(channels, channels, 1, CONV_1x1)
        self.v = Conv2d(channels, channels, 1, CONV_1x1)
        self.proj_out = Conv2d(channels, channels, 1, CONV_1x1)
        self.spec_proj = self.proj_out.spec
        self._cache = ops.ParamCache()

    def fused_qkv(self, wq, bq, wk, bk, wv, bv, code):
        """-> (packed [3C][C] weight in the format `code`, fp32 [3C] bias) for the single q|k|v projection."""
        C = self.channels
        if isinstance(code, torch.dtype):
            code = ops.act_code(code)

        def build(old):
            wbuf, bbuf = old if old is not None else (
                torch.empty(3 * C * C, dtype=ops.buf_dtype(code), device=wq.device),
                torch.empty(3 * C, dtype=torch.float32, device=wq.device))
            for i, (w, b) in enumerate(((wq, bq), (wk, bk), (wv, bv))):
                ops.pack_weight(w.detach(), CONV_1x1, C, C, 0, code, out=wbuf[i * C * C:(i + 1) * C * C])
                bbuf[i * C:(i + 1) * C].copy_(b.detach())
            return wbuf, bbuf

        return self._cache.get(("qkv", code), (wq, bq, wk, bk, wv, bv), build)

    def qkv_grad_buffers(self, wq, bq):
        C = self.channels
        dw = torch.empty(3 * C, C, dtype=torch.float32, device=wq.device)
        db = torch.empty(3 * C, dtype=torch.float32, device=wq.device)
        return dw, db, False

    def forward(self, x):
        return ops.AttnBlock.apply(x, self.norm.weight, self.norm.bias, self.q.weight, self.q.bias, self.k.weight,
                                   self.k.bias, self.v.weight, self.v.bias, self.proj_out.weight,
                                   self.proj_out.bias, self)


class Downsample(nn.Module):
    """with_conv=True: F.pad(0,1,0,1) + 3x3 stride-2 conv  (LD model.py:66-76)."""

    def __init__(self, channels):
        super().__init__()
        self.conv = Conv2d(channels, channels, 3, CONV_3x3_S2)

    def forward(self, x):
        return self.conv(x)


class Upsample(nn.Module):
    """with_conv=True: nearest 2x + 3x3 conv  (LD model.py:53-57)."""

    def __init__(self, channels):
        super().__init__()
        self.conv = Conv2d(channels, channels, 3, CONV_3x3_S1)

    def forward(self, x):
        return ops.UpsampleConv.apply(x, self.conv.weight, self.conv.bias, self.conv.spec)


class _Level(nn.Module):
    def __init__(self):
        super().__init__()
        self.block = nn.ModuleList()
        self.attn = nn.ModuleList()


class _Mid(nn.Module):
    def __init__(self, channels):
        super().__init__()
        self.block_1 = ResnetBlock(channels)
        self.attn_1 = AttnBlock(channels)
        self.block_2 = ResnetBlock(channels)

    def forward(self, h):
        return _run(self.block_2, _run(self.attn_1, _run(self.block_1, h)))


class Encoder(nn.Module):
    """LD model.py:368-459 (attn_type 'vanilla', resamp_with_conv, temb_ch 0)."""

    def __init__(self, *, ch, out_ch, ch_mult=(1, 2, 4, 8), num_res_blocks, attn_resolutions, dropout=0.0,
                 resamp_with_conv=True, in_channels, resolution, z_channels, double_z=True,
                 compute_dtype=torch.bfloat16, **ignore_kwargs):
        super().__init__()
        assert resamp_with_conv and dropout == 0.0, "only the north-star configuration is implemented"
        self.ch, self.num_resolutions, self.num_res_blocks = ch, len(ch_mult), num_res_blocks
        self.resolution, self.in_channels = resolution, in_channels
        self.compute_dtype = compute_dtype
        self.conv_in = Conv2d(in_channels, ch, 3, CONV_3x3_S1)
        curr_res = resolution
        in_ch_mult = (1,) + tuple(ch_mult)
        self.down = nn.ModuleList()
        block_in = ch
        for i_level in range(self.num_resolutions):
            lvl = _Level()
            block_in = ch * in_ch_mult[i_level]
            block_out = ch * ch_mult[i_level]
            for _ in range(num_res_blocks):
                lvl.block.append(ResnetBlock(block_in, block_out))
                block_in = block_out
                if curr_res in attn_resolutions:
                    lvl.attn.append(AttnBlock(block_in))
            if i_level != self.num_resolutions - 1:
                lvl.downsample = Downsample(block_in)
                curr_res = curr_res // 2
            self.down.append(lvl)
        self.mid = _Mid(block_in)
        self.norm_out = GroupNorm(block_in)
        self.out_channels = 2 * z_channels if double_z else z_channels
        self.conv_out = Conv2d(block_in, self.out_channels, 3, CONV_3x3_S1)
        ops.assign_format(self, compute_dtype)

    def forward_nhwc(self, x):
        """x: NCHW fp32 image -> NHWC activations [B,h,w,pad64(out_channels)]."""
        _check_cuda(x)
        spec = self.conv_in.spec
        B, C, H, W = x.shape
        if not x.requires_grad and spec.fmt is not None and ops.thin_conv_ok(spec, spec.fmt.fwd, B, H, W, spec.coutp, C):
            h = ops.ThinConvIn.apply(x, self.conv_in.weight, self.conv_in.bias, spec)
        else:
            h = ops.ToNHWC.apply(x, ops.cpad(self.in_channels), self.compute_dtype, getattr(self, "gscale", None))
            h = self.conv_in(h, need_dx=x.requires_grad)
        for i_level in range(self.num_resolutions):
            lvl = self.down[i_level]
            for i_block in range(self.num_res_blocks):
                h = _run(lvl.block[i_block], h)
                if len(lvl.attn) > 0:
                    h = _run(lvl.attn[i_block], h)
            if i_level != self.num_resolutions - 1:
                h = _run(lvl.downsample, h)
        h = self.mid(h)
        h = _run(self.norm_out, h, silu=True)
        return _run(self.conv_out, h)

    def forward(self, x):
        return ops.ToNCHW.apply(self.forward_nhwc(x), self.out_channels, self.compute_dtype, getattr(self, "gscale", None))


class Decoder(nn.Module):
    """LD model.py:462-568 (give_pre_end False, tanh_out False)."""

    def __init__(self, *, ch, out_ch, ch_mult=(1, 2, 4, 8), num_res_blocks, attn_resolutions, dropout=0.0,
                 resamp_with_conv=True, in_channels, resolution, z_channels, give_pre_end=False, tanh_out=False,
                 compute_dtype=torch.bfloat16, **ignorekwargs):
        super().__init__()
        assert resamp_with_conv and dropout == 0.0 and not give_pre_end and not tanh_out
        self.ch, self.num_resolutions, self.num_res_blocks = ch, len(ch_mult), num_res_blocks
        self.resolution, self.in_channels, self.z_channels, self.out_ch = resolution, in_channels, z_channels, out_ch
        self.compute_dtype = compute_dtype
        block_in = ch * ch_mult[self.num_resolutions - 1]
        curr_res = resolution // 2 ** (self.num_resolutions - 1)
        self.z_shape = (1, z_channels, curr_res, curr_res)
        self.conv_in = Conv2d(z_channels, block_in, 3, CONV_3x3_S1)
        self.mid = _Mid(block_in)
        self.up = nn.ModuleList()
        for i_level in reversed(range(self.num_resolutions)):
            lvl = _Level()
            block_out = ch * ch_mult[i_level]
            for _ in range(num_res_blocks + 1):
                lvl.block.append(ResnetBlock(block_in, block_out))
                block_in = block_out
                if curr_res in attn_resolutions:
                    lvl.attn.append(AttnBlock(block_in))
            if i_level != 0:
                lvl.upsample = Upsample(block_in)
                curr_res = curr_res * 2
            self.up.insert(0, lvl)
        self.norm_out = GroupNorm(block_in)
        self.conv_out = Conv2d(block_in, out_ch, 3, CONV_3x3_S1)
        ops.assign_format(self, compute_dtype)

    def forward_nhwc(self, z):
        """z: NHWC [B,h,w,pad64(z_channels)] -> dec NCHW fp32 [B,out_ch,H,W]."""
        h = _run(self.conv_in, z)
        h = self.mid(h)
        for i_level in reversed(range(self.num_resolutions)):
            lvl = self.up[i_level]
            for i_block in range(self.num_res_blocks + 1):
                h = _run(lvl.block[i_block], h)
                if len(lvl.attn) > 0:
                    h = _run(lvl.attn[i_block], h)
            if i_level != 0:
                h = _run(lvl.upsample, h)
        h = _run(self.norm_out, h, silu=True)
        return _run(self.conv_out, h, out_nchw=True)

    def forward(self, z):
        _check_cuda(z)
        return self.forward_nhwc(ops.ToNHWC.apply(z, ops.cpad(self.z_channels), self.compute_dtype,
                                                  getattr(self, "gscale", None)))


class Linear(nn.Module):
    """nn.Linear with torch's default init; fp32 GEMM on the SIMT engine (M = batch is tiny)."""

    def __init__(self, in_features, out_features, bias=True):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if bias:
            self.bias = nn.Parameter(torch.empty(out_features))
            bound = 1.0 / math.sqrt(in_features)
            nn.init.uniform_(self.bias, -bound, bound)
        else:
            self.register_parameter("bias", None)

    def forward(self, x):
        _check_cuda(x)
        return ops.Linear.apply(ops._c(x.float()), self.weight, self.bias)


class Activation(nn.Module):
    CODES = {"tanh": 1, "silu": 2}

    def __init__(self, kind):
        super().__init__()
        self.code = self.CODES[kind]

    def forward(self, x):
        return ops.Act.apply(x, self.code)
