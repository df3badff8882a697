"""LPIPS perceptual distance on the B200 kernels.

Reference: `self.perceptual_loss = LPIPS().eval()` (LD/ldm/modules/losses/contperceptual.py:17) called as
`p_loss = self.perceptual_loss(inputs, reconstructions)` -> [B,1,1,1] (src/modules/losses/contperceptual.py:141-145).
The class itself lives in the un-vendored taming-transformers (pinned at 049ce3ce, `.SUBMODULES.json:44-50`); this
follows its public definition and keeps its state_dict names (`net.slice{1..5}.{idx}.weight|bias`,
`lin{0..4}.model.1.weight`, `scaling_layer.shift|scale`), so a real LPIPS checkpoint (`vgg.pth` + lin heads, not
downloadable here) loads with `load_state_dict`. PARITY UNPINNED: checked against oracle/lpips_oracle.py only.

Data path: both images -> scaling layer (torch, 3 channels) -> NHWC, 64-channel padded -> 13 x (3x3 conv on the conv
engine + in-place ReLU), 4 x 2x2 max pool -> per level one fused kernel (normalise, squared difference, lin weights,
spatial mean; fixed-order sums) accumulating into out[B]. All parameters are frozen: the backward pass only carries
the gradient of the reconstruction's tower back to the image (lpips_level_bwd -> relu_bwd -> conv dgrad -> maxpool_bwd).
16-bit modes run the tower in bf16 (as the attention blocks do), the fp32 validation mode in fp32.
"""
import torch
import torch.nn as nn

from . import ops
from .lib import BF16, CONV_3x3_S1, F32
from .ops import call

VGG_SLICES = [("slice1", [(0, 3, 64), (2, 64, 64)]),
              ("slice2", [(5, 64, 128), (7, 128, 128)]),
              ("slice3", [(10, 128, 256), (12, 256, 256), (14, 256, 256)]),
              ("slice4", [(17, 256, 512), (19, 512, 512), (21, 512, 512)]),
              ("slice5", [(24, 512, 512), (26, 512, 512), (28, 512, 512)])]
CHNS = [64, 128, 256, 512, 512]


class _Conv(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(cout, cin, 3, 3), requires_grad=False)
        self.bias = nn.Parameter(torch.zeros(cout), requires_grad=False)
        nn.init.kaiming_normal_(self.weight, nonlinearity="relu")
        self.spec = ops.ConvSpec(cin, cout, CONV_3x3_S1)


class _Lin(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.weight = nn.Parameter(torch.rand(1, c, 1, 1) / c, requires_grad=False)


class _Tower:
    """One pass of the VGG16 feature stack over an NHWC tensor; keeps every ReLU output (the backward needs them)."""

    def __init__(self, mod, x, code):
        self.acts = []     # ReLU outputs in execution order (13)
        self.levels = []   # indices into acts of the 5 level outputs
        for k, (name, convs) in enumerate(VGG_SLICES):
            if k > 0:
                B, H, W, C = x.shape
                y = torch.empty(B, H // 2, W // 2, C, dtype=x.dtype, device=x.device)
                call("maxpool2x2", x, y, B, H, W, C, code)
                x = y
            for i, _, _ in convs:
                cv = getattr(mod.net, name)[str(i)]
                x = ops.conv_fwd_raw(x, cv.spec, cv.weight, cv.bias, op=code, out=code)
                call("relu", x, x.numel(), code)
                self.acts.append(x)
            self.levels.append(len(self.acts) - 1)


class _LPIPSFn(torch.autograd.Function):
    """out[b] = LPIPS(x0[b], x1[b]); x0, x1: scaled NCHW fp32 images; gradient to x1 (the reconstruction) only."""

    @staticmethod
    def forward(ctx, x1, x0, mod, code):
        B, _, H, W = x1.shape
        if (H % 16) or (W % 16):
            raise RuntimeError("LPIPS: the four 2x2 max pools need H, W divisible by 16, got %dx%d" % (H, W))
        t0 = _Tower(mod, ops.nchw_to_nhwc_raw(ops._c(x0), 64, code), code)
        t1 = _Tower(mod, ops.nchw_to_nhwc_raw(ops._c(x1), 64, code), code)
        out = torch.empty(B, dtype=torch.float32, device=x1.device)
        ws = ops.workspace(call("lpips_level_workspace_bytes", B), x1.device, "lpips")
        lins = [getattr(mod, "lin%d" % k).model["1"].weight.detach().reshape(-1).contiguous() for k in range(5)]
        for k in range(5):
            f0, f1 = t0.acts[t0.levels[k]], t1.acts[t1.levels[k]]
            call("lpips_level_fwd", f0, f1, lins[k], out, B, f0.shape[1] * f0.shape[2], f0.shape[3], code,
                 1 if k else 0, ws, ws.numel())
        ctx.mod, ctx.code, ctx.hw = mod, code, (H, W)
        ctx.f0 = [t0.acts[i] for i in t0.levels]
        ctx.t1, ctx.lins = t1, lins
        return out

    @staticmethod
    def backward(ctx, dout):
        mod, code, t1 = ctx.mod, ctx.code, ctx.t1
        dout = ops._c(dout.float())
        g = None  # gradient w.r.t. the current activation of the reconstruction's tower
        for k in range(4, -1, -1):
            name, convs = VGG_SLICES[k]
            li = t1.levels[k]
            f1 = t1.acts[li]
            B, H, W, C = f1.shape
            if g is None:
                g = torch.empty_like(f1)
            call("lpips_level_bwd", ctx.f0[k], f1, ctx.lins[k], dout, g, B, H * W, C, code, 0 if k == 4 else 1)
            for j in range(len(convs) - 1, -1, -1):
                ai = li - (len(convs) - 1 - j)
                y = t1.acts[ai]
                call("relu_bwd", y, g, g, g.numel(), code)
                cv = getattr(mod.net, name)[str(convs[j][0])]
                g = ops.conv_dgrad_raw(g, cv.spec, cv.weight, (y.shape[1], y.shape[2]))
            if k > 0:  # g is the gradient of this slice's pooled input: route it back through the max pool
                xin = t1.acts[t1.levels[k - 1]]
                Bp, Hp, Wp, Cp = xin.shape
                gp = torch.empty_like(xin)
                call("maxpool2x2_bwd", xin, g, gp, Bp, Hp, Wp, Cp, code)
                g = gp
        dx1 = ops.nhwc_to_nchw_raw(g, 3, code)
        # the saved towers stay alive until autograd frees the node: PoseLoss.calculate_adaptive_weight runs this backward
        # once with retain_graph=True before the step's own backward
        return dx1, None, None, None


class LPIPS(nn.Module):
    def __init__(self, use_dropout=True, compute_dtype=torch.float16):
        super().__init__()
        self.scaling_layer = nn.Module()
        self.scaling_layer.register_buffer("shift", torch.tensor([-.030, -.088, -.188])[None, :, None, None])
        self.scaling_layer.register_buffer("scale", torch.tensor([.458, .448, .450])[None, :, None, None])
        self.net = nn.Module()
        for name, convs in VGG_SLICES:
            self.net.add_module(name, nn.ModuleDict({str(i): _Conv(ci, co) for i, ci, co in convs}))
        for k, c in enumerate(CHNS):
            lin = nn.Module()
            lin.model = nn.ModuleDict({"1": _Lin(c)})
            self.add_module("lin%d" % k, lin)
        self.compute_dtype = compute_dtype
        self.register_load_state_dict_post_hook(lambda *a: ops.bump_weight_epoch())

    def forward(self, input, target):
        if not target.is_cuda:
            raise RuntimeError("odvae_b200 LPIPS runs on the B200 only (no CPU fallback); the CPU restatement is oracle/")
        sl = self.scaling_layer
        code = F32 if self.compute_dtype == torch.float32 else BF16
        x0 = ((input.float() - sl.shift) / sl.scale).detach()
        x1 = (target.float() - sl.shift) / sl.scale
        return _LPIPSFn.apply(x1, x0, self, code).view(-1, 1, 1, 1)
