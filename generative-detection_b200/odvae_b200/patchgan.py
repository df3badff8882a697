"""PatchGAN discriminator on the B200 kernels.

Reference: taming.modules.discriminator.model.NLayerDiscriminator (un-vendored, pinned 049ce3ce; public definition in
SURVEY appendix B), constructed at LD/ldm/modules/losses/contperceptual.py:22-25 and called at
src/modules/losses/contperceptual.py:285 (generator side) and :355-356 (discriminator side). Same constructor kwargs
and `state_dict` keys (`main.{0,2,3,5,6,8,9,11}.*` incl. the BatchNorm running statistics), same init (`weights_init`).

Every Conv2d(k=4, s=2|1, p=1) runs as im2col + ONE GEMM on the tcgen05 engine (fp32 mode: the SIMT GEMM). The previous
layer's train-mode BatchNorm + LeakyReLU(0.2) is applied inside the im2col gather, so the normalised activations are
never written; the image mask (`reconstructions * mask_2d_bbox`, LOSS.py:250-257) is fused into the first gather.
Backward is hand-written: dgrad GEMM -> col2im (+ LeakyReLU') -> BatchNorm backward; weight gradients are GEMMs over
re-gathered patches in the gradient format with a fixed-order split over groups of images.
"""
import torch
import torch.nn as nn

from . import ops
from .lib import F32, call

SLOPE = 0.2
BN_EPS = 1e-5
BN_MOMENTUM = 0.1


def _out_hw(h, w, stride):
    return (h - 2) // stride + 1, (w - 2) // stride + 1


def _gemm(A, Bm, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, a_mn, b_mn, code, out_code, bias=None):
    """D[b] = op(A[b]) op(B[b])^T (+bias): tcgen05 engine for 16-bit operands, SIMT for fp32."""
    if code == F32:
        call("gemm_simt", A, Bm, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, int(a_mn), int(b_mn), F32, F32, 1.0,
             bias, None)
    else:
        call("gemm_16", A, Bm, D, M, N, K, batch, lda, ldb, ldd, sA, sB, sD, int(a_mn), int(b_mn), code, code,
             out_code, 1.0, bias, None)
    return D


class _Layer:
    """One Conv2d(k4) (+BatchNorm2d) of the Sequential: geometry, padded GEMM sizes, parameter caches."""

    def __init__(self, conv, bn, first, last):
        self.conv, self.bn = conv, bn
        self.cin, self.cout = conv.in_channels, conv.out_channels
        self.stride = conv.stride[0]
        self.first, self.last = first, last
        self.kp = 64 if first else 16 * self.cin     # K of the GEMM (first layer: 16 * 3 = 48 padded to one K block)
        self.coutp = ops.cpad(self.cout)             # last layer: Cout = 1 padded to one 64-wide N tile
        self.cache = ops.ParamCache()

    def w_packed(self, code):
        w = self.conv.weight

        def build(old):
            out = old if old is not None else torch.empty(self.coutp * self.kp, dtype=ops.buf_dtype(code), device=w.device)
            call("pack_k4_weight", w.detach(), out, self.cout, self.cin, self.coutp, self.kp, code)
            return out
        return self.cache.get(("w", code), (w,), build)

    def bias_padded(self):
        b = self.conv.bias
        if b is None:
            return None
        if b.numel() == self.coutp:
            return b.detach()
        return self.cache.get(("b", self.coutp), (b,), lambda old: ops.pad_vector(b.detach(), self.coutp, old))


def _split(B):
    """Weight-gradient GEMMs reduce over groups of images: the largest divisor of B that is <= 16."""
    for s in range(min(B, 16), 0, -1):
        if B % s == 0:
            return s
    return 1


class PatchGANFn(torch.autograd.Function):
    """logits[B,1,Ho,Wo] = D(x * mask).  tensors = (x, mask, w0, b0, w1, g1, be1, ..., wL, bL) in layer order."""

    @staticmethod
    def forward(ctx, disc, x, mask, *params):
        fmt = disc.fmt
        fc, gc = fmt.fwd, fmt.grad
        layers = disc.layers
        B, C, H, W = x.shape
        dev = x.device
        x = ops._c(x.float())
        mask = None if mask is None else ops._c(mask.float())
        saved = {"ys": [], "bn": [], "geo": []}
        act, act_hw, scale, shift = x, (H, W), None, None
        training = disc.training
        for li, L in enumerate(layers):
            h, w = act_hw
            ho, wo = _out_hw(h, w, L.stride)
            if ho <= 0 or wo <= 0:
                raise RuntimeError("PatchGAN input %dx%d is too small for %d layers" % (H, W, len(layers)))
            M = B * ho * wo
            P = torch.empty(M, L.kp, dtype=ops.buf_dtype(fc), device=dev)
            if L.first:
                call("patchgan_im2col", act, mask, None, None, 1.0, P, B, h, w, L.cin, L.stride, L.kp, 1, F32, fc)
            else:
                call("patchgan_im2col", act, None, scale, shift, SLOPE, P, B, h, w, L.cin, L.stride, L.kp, 0, fc, fc)
            out_code = F32 if L.last else fc
            Y = torch.empty(M, L.coutp, dtype=ops.buf_dtype(out_code), device=dev)
            _gemm(P, L.w_packed(fc), Y, M, L.coutp, L.kp, 1, L.kp, L.kp, L.coutp, 0, 0, 0, False, False, fc, out_code,
                  bias=L.bias_padded())
            del P
            saved["geo"].append((h, w, ho, wo, M))
            scale = shift = None
            if L.bn is not None:
                bn = L.bn
                C_ = L.cout
                scale = torch.empty(C_, dtype=torch.float32, device=dev)
                shift = torch.empty(C_, dtype=torch.float32, device=dev)
                mean = torch.empty(C_, dtype=torch.float32, device=dev)
                rstd = torch.empty(C_, dtype=torch.float32, device=dev)
                if training:
                    nbytes = call("bn_workspace_bytes", M, C_)
                    ws = ops.workspace(nbytes, dev, "bn")
                    call("bn_train_stats", Y, M, C_, fc, bn.weight.detach(), bn.bias.detach(), BN_EPS, BN_MOMENTUM,
                         bn.running_mean, bn.running_var, bn.num_batches_tracked, scale, shift, mean, rstd, ws,
                         ws.numel())
                else:  # eval(): running statistics (tiny per-channel vectors)
                    rstd = torch.rsqrt(bn.running_var + BN_EPS)
                    mean = bn.running_mean.clone()
                    scale = bn.weight.detach() * rstd
                    shift = bn.bias.detach() - mean * scale
                saved["bn"].append((scale, shift, mean, rstd))
            else:
                saved["bn"].append(None)
            saved["ys"].append(Y)
            act, act_hw = Y, (ho, wo)
        ctx.disc, ctx.saved, ctx.fmt = disc, saved, fmt
        ctx.x, ctx.mask, ctx.training = x, mask, training
        ctx.params = params
        _, _, ho, wo, M = saved["geo"][-1]
        return act.view(B, ho, wo, layers[-1].coutp)[..., :1].permute(0, 3, 1, 2)

    @staticmethod
    def backward(ctx, dlogits):
        disc, saved, fmt = ctx.disc, ctx.saved, ctx.fmt
        fc, gc = fmt.fwd, fmt.grad
        gdt = ops.buf_dtype(gc)
        layers = disc.layers
        x, mask = ctx.x, ctx.mask
        B = x.shape[0]
        dev = x.device
        need_x = ctx.needs_input_grad[1]
        # which parameters want gradients (positions 3.. of forward's inputs)
        need_p = ctx.needs_input_grad[3:]
        grads = [None] * len(ctx.params)
        # parameter positions per layer
        pos, p = [], 0
        for L in layers:
            ent = {"w": p}
            p += 1
            if L.conv.bias is not None:
                ent["b"] = p
                p += 1
            if L.bn is not None:
                ent["g"], ent["be"] = p, p + 1
                p += 2
            pos.append(ent)
        last = layers[-1]
        M = saved["geo"][-1][4]
        dY = torch.zeros(M, last.coutp, dtype=gdt, device=dev)
        dY[:, 0] = dlogits.reshape(-1).to(gdt)
        S = _split(B)
        for li in range(len(layers) - 1, -1, -1):
            L = layers[li]
            h, w, ho, wo, M = saved["geo"][li]
            ent = pos[li]
            want_w = need_p[ent["w"]]
            if want_w:
                # patches of this layer's input, re-gathered in the gradient format
                P = torch.empty(M, L.kp, dtype=gdt, device=dev)
                if L.first:
                    call("patchgan_im2col", x, mask, None, None, 1.0, P, B, h, w, L.cin, L.stride, L.kp, 1, F32, gc)
                else:
                    bnp = saved["bn"][li - 1]
                    sc, sh = (bnp[0], bnp[1]) if bnp is not None else (None, None)
                    call("patchgan_im2col", saved["ys"][li - 1], None, sc, sh, SLOPE, P, B, h, w, L.cin, L.stride, L.kp,
                         0, fc, gc)
                rows = M // S
                part = torch.empty(S, L.coutp, L.kp, dtype=torch.float32, device=dev)
                _gemm(dY, P, part, L.coutp, L.kp, rows, S, L.coutp, L.kp, L.kp, rows * L.coutp, rows * L.kp,
                      L.coutp * L.kp, True, True, gc, F32)
                del P
                dw, direct = ops._new_grad(L.conv.weight)
                call("unpack_k4_wgrad", part, S, dw, L.cout, L.cin, L.coutp, L.kp, 0)
                grads[ent["w"]] = ops._ret(dw, direct)
                del part
            if "b" in ent and need_p[ent["b"]]:
                db, direct = ops._new_grad(L.conv.bias)
                nbytes = call("colsum_workspace_bytes", L.cout)
                ws = ops.workspace(nbytes, dev, "colsum")
                call("colsum", dY, M, L.cout, L.coutp, gc, db, 0, ws, ws.numel())
                grads[ent["b"]] = ops._ret(db, direct)
            if li == 0 and not need_x:
                break
            # dP = dY W  (the packed weights in the gradient format, read MN-major)
            dP = torch.empty(M, L.kp, dtype=gdt, device=dev)
            _gemm(dY, L.w_packed(gc), dP, M, L.kp, L.coutp, 1, L.coutp, L.kp, L.kp, 0, 0, 0, False, True, gc, gc)
            del dY
            if li == 0:
                dx = torch.empty_like(x)
                call("patchgan_col2im", dP, None, None, None, 1.0, mask, dx, B, h, w, L.cin, L.stride, L.kp, 1, gc, fc,
                     F32)
                return (None, dx, None) + tuple(grads)
            Lp = layers[li - 1]
            bnp = saved["bn"][li - 1]
            Yp = saved["ys"][li - 1]
            Mp = saved["geo"][li - 1][4]
            dyp = torch.empty(Mp, Lp.coutp, dtype=gdt, device=dev)
            sc, sh = (bnp[0], bnp[1]) if bnp is not None else (None, None)
            call("patchgan_col2im", dP, Yp, sc, sh, SLOPE, None, dyp, B, h, w, L.cin, L.stride, L.kp, 0, gc, fc, gc)
            del dP
            if bnp is not None:
                entp = pos[li - 1]
                want_bn = need_p[entp["g"]] or need_p[entp["be"]]
                if ctx.training:
                    dg, dgd = ops._new_grad(Lp.bn.weight) if want_bn else (None, True)
                    db_, dbd = ops._new_grad(Lp.bn.bias) if want_bn else (None, True)
                    nbytes = call("bn_workspace_bytes", Mp, Lp.cout)
                    ws = ops.workspace(nbytes, dev, "bn")
                    call("bn_train_bwd", Yp, dyp, Mp, Lp.cout, fc, gc, Lp.bn.weight.detach(), bnp[2], bnp[3], dg, db_, 0,
                         ws, ws.numel())
                    if want_bn:
                        grads[entp["g"]], grads[entp["be"]] = ops._ret(dg, dgd), ops._ret(db_, dbd)
                else:  # eval(): y = x*scale + shift with constant statistics
                    if want_bn:
                        xhat = (Yp.view(ops.buf_dtype(fc) if fc == F32 else _view_dtype(fc)).float() - bnp[2]) * bnp[3]
                        grads[entp["g"]] = (dyp.float() * xhat).sum(0)
                        grads[entp["be"]] = dyp.float().sum(0)
                    dyp = (dyp.float() * sc).to(gdt)
            dY = dyp
        return (None, None, None) + tuple(grads)


def _view_dtype(code):
    from .lib import F16
    return torch.float16 if code == F16 else torch.bfloat16


def weights_init(m):  # taming.modules.discriminator.model.weights_init
    name = m.__class__.__name__
    if name.find("Conv") != -1:
        nn.init.normal_(m.weight.data, 0.0, 0.02)
    elif name.find("BatchNorm") != -1:
        nn.init.normal_(m.weight.data, 1.0, 0.02)
        nn.init.constant_(m.bias.data, 0)


class NLayerDiscriminator(nn.Module):
    """PatchGAN discriminator with taming's layout. forward(x[B,C,H,W] fp32 NCHW) -> logits [B,1,Ho,Wo] fp32.
    Extensions: `mask` ([B,1,H,W] or [B,H,W]; D(x * mask) with the multiply fused into the first gather) and
    `param_grads=False` (generator-side call: gradients flow to x only, like Lightning's toggle_optimizer)."""

    def __init__(self, input_nc=3, ndf=64, n_layers=3, use_actnorm=False):
        super().__init__()
        if use_actnorm:
            raise NotImplementedError("use_actnorm=True (ActNorm) is not used by the reference config")
        kw, padw = 4, 1
        seq = [nn.Conv2d(input_nc, ndf, kw, 2, padw), nn.LeakyReLU(SLOPE, True)]
        nf_mult = 1
        for n in range(1, n_layers):
            nf_prev, nf_mult = nf_mult, min(2 ** n, 8)
            seq += [nn.Conv2d(ndf * nf_prev, ndf * nf_mult, kw, 2, padw, bias=False),
                    nn.BatchNorm2d(ndf * nf_mult), nn.LeakyReLU(SLOPE, True)]
        nf_prev, nf_mult = nf_mult, min(2 ** n_layers, 8)
        seq += [nn.Conv2d(ndf * nf_prev, ndf * nf_mult, kw, 1, padw, bias=False),
                nn.BatchNorm2d(ndf * nf_mult), nn.LeakyReLU(SLOPE, True)]
        seq += [nn.Conv2d(ndf * nf_mult, 1, kw, 1, padw)]
        self.main = nn.Sequential(*seq)
        self.apply(weights_init)
        self.fmt = ops.FORMAT_BF16
        mods = list(self.main)
        convs = [i for i, m in enumerate(mods) if isinstance(m, nn.Conv2d)]
        self._layer_list = []
        for n, i in enumerate(convs):
            bn = mods[i + 1] if i + 1 < len(mods) and isinstance(mods[i + 1], nn.BatchNorm2d) else None
            self._layer_list.append(_Layer(mods[i], bn, n == 0, n == len(convs) - 1))
        if input_nc > 4:
            raise NotImplementedError("the first-layer gather takes the <= 4-channel image (disc_in_channels = 3)")

    @property
    def layers(self):
        return self._layer_list

    def _params(self, detach):
        out = []
        for L in self._layer_list:
            ps = [L.conv.weight] + ([L.conv.bias] if L.conv.bias is not None else [])
            if L.bn is not None:
                ps += [L.bn.weight, L.bn.bias]
            out += [p.detach() for p in ps] if detach else ps
        return out

    def forward(self, x, mask=None, param_grads=True):
        if not x.is_cuda:
            raise RuntimeError("odvae_b200 modules run on a B200 (sm_100a) CUDA device only; there is no CPU fallback")
        if mask is not None:
            mask = mask.reshape(x.shape[0], x.shape[2], x.shape[3])
        return PatchGANFn.apply(self, x, mask, *self._params(not param_grads))
