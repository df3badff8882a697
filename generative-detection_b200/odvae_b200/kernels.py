"""Functional wrappers over the C ABI: allocate outputs / workspaces with torch (device memory and
streams are the only things torch provides here) and call the hand-written kernels.

Activations are NHWC tensors [B,H,W,C] of dtype float32 (validation mode) or bfloat16 (timed mode).
"""
import torch

from . import lib
from .lib import BF16, CONV_1x1, CONV_3x3_S1, CONV_3x3_S2, F32, IMPL_SIMT, IMPL_TC, call


def act_code(dtype):
    if dtype == torch.bfloat16:
        return BF16
    if dtype == torch.float32:
        return F32
    if dtype == torch.float16:
        return lib.F16
    raise TypeError("activations must be float32, bfloat16 or float16, got %s" % dtype)


def pad_to(n, m):
    return (n + m - 1) // m * m


def chan_pad(c):
    """Channel padding used for activation buffers: multiples of 64 feed the tcgen05 engine directly;
    thin tensors (3, 16, 32 channels) are padded to 16."""
    return pad_to(c, 64) if c >= 64 else pad_to(c, 16)


_WS = {}


def workspace(nbytes, device, tag="default"):
    """Grow-only scratch buffer per (device, stream, tag); never part of any state_dict."""
    key = (device, torch.cuda.current_stream(device).cuda_stream, tag)
    buf = _WS.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _WS[key] = buf
    return buf


def geom_out_hw(geom, H, W):
    return (H // 2, W // 2) if geom == CONV_3x3_S2 else (H, W)


def geom_k(geom):
    return 1 if geom == CONV_1x1 else 3


def pack_conv_weight(w, geom, CoutP, CinP, which, dtype):
    Cout, Cin = w.shape[0], w.shape[1]
    k = geom_k(geom)
    out = torch.empty(CoutP * CinP * k * k, dtype=dtype, device=w.device)
    call("pack_conv_weight", w, out, Cout, Cin, geom, CoutP, CinP, which, act_code(dtype))
    return out


def conv_tc_supported(B, H, W, CinP, CoutP, geom):
    return bool(call("conv2d_tc_supported", B, H, W, CinP, CoutP, geom))


def conv_fwd(x, w_packed, bias, Cout, geom, impl, residual=None, out_nchw_f32=False):
    B, H, W, CinP = x.shape
    Ho, Wo = geom_out_hw(geom, H, W)
    k = geom_k(geom)
    CoutP = w_packed.numel() // (CinP * k * k)
    if out_nchw_f32:
        y = torch.empty(B, Cout, Ho, Wo, dtype=torch.float32, device=x.device)
    else:
        y = torch.empty(B, Ho, Wo, CoutP, dtype=x.dtype, device=x.device)
    call("conv2d_fwd", x, w_packed, bias, residual, y, B, H, W, CinP, CoutP, Cout, geom, act_code(x.dtype),
         act_code(x.dtype), impl,
         1 if out_nchw_f32 else 0)
    return y


def conv_dgrad(dy, w_packed_dgrad, CinP, geom, impl, in_hw, dx_add=None):
    B, Ho, Wo, CoutP = dy.shape
    H, W = in_hw
    dx = torch.empty(B, H, W, CinP, dtype=dy.dtype, device=dy.device)
    call("conv2d_dgrad", dy, w_packed_dgrad, dx_add, dx, B, H, W, CinP, CoutP, geom, act_code(dy.dtype), impl)
    return dx


def conv_wgrad(x, dy, Cin, Cout, geom, impl, want_bias=True, dw=None, db=None, accumulate=False, out_scale=1.0):
    B, H, W, CinP = x.shape
    CoutP = dy.shape[-1]
    k = geom_k(geom)
    if dw is None:
        dw = torch.empty(Cout, Cin, k, k, dtype=torch.float32, device=x.device)
    if want_bias and db is None:
        db = torch.empty(Cout, dtype=torch.float32, device=x.device)
    nbytes = call("conv2d_wgrad_workspace_bytes", B, H, W, CinP, CoutP, geom, impl)
    ws = workspace(nbytes, x.device, "wgrad")
    call("conv2d_wgrad", x, dy, dw, db if want_bias else None, B, H, W, CinP, CoutP, Cin, Cout, geom,
         act_code(x.dtype), impl, 1 if accumulate else 0, float(out_scale), ws, ws.numel())
    return dw, db


def gemm(A, B, M, N, K, batch=1, lda=None, ldb=None, ldd=None, strideA=0, strideB=0, strideD=0, a_mn=False,
         b_mn=False, out_dtype=torch.bfloat16, alpha=1.0, bias=None, residual=None, out=None, impl=IMPL_TC):
    """D[b] = alpha * op(A[b]) op(B[b])^T; see include/odvae_b200.h for operand conventions."""
    lda = lda if lda is not None else (M if a_mn else K)
    ldb = ldb if ldb is not None else (N if b_mn else K)
    ldd = ldd if ldd is not None else N
    if out is None:
        out = torch.empty(batch, M, ldd, dtype=out_dtype, device=A.device)
        strideD = M * ldd
    if impl == IMPL_TC:
        call("gemm_bf16", A, B, out, M, N, K, batch, lda, ldb, ldd, strideA, strideB, strideD, int(a_mn), int(b_mn),
             act_code(out.dtype), float(alpha), bias, residual)
    else:
        call("gemm_simt", A, B, out, M, N, K, batch, lda, ldb, ldd, strideA, strideB, strideD, int(a_mn), int(b_mn),
             act_code(A.dtype), act_code(out.dtype), float(alpha), bias, residual)
    return out


def nchw_to_nhwc(x, Cpad, dtype, scale=1.0, shift=0.0):
    B, C, H, W = x.shape
    out = torch.empty(B, H, W, Cpad, dtype=dtype, device=x.device)
    call("nchw_to_nhwc", x, out, B, C, H, W, Cpad, act_code(dtype), float(scale), float(shift))
    return out


def nhwc_to_nchw(x, C):
    B, H, W, Cpad = x.shape
    out = torch.empty(B, C, H, W, dtype=torch.float32, device=x.device)
    call("nhwc_to_nchw", x, out, B, C, H, W, Cpad, act_code(x.dtype))
    return out


def upsample2x(x):
    B, H, W, C = x.shape
    y = torch.empty(B, 2 * H, 2 * W, C, dtype=x.dtype, device=x.device)
    call("upsample2x", x, y, B, H, W, C, act_code(x.dtype))
    return y


def upsample2x_bwd(dy):
    B, H2, W2, C = dy.shape
    dx = torch.empty(B, H2 // 2, W2 // 2, C, dtype=dy.dtype, device=dy.device)
    call("upsample2x_bwd", dy, dx, B, H2 // 2, W2 // 2, C, act_code(dy.dtype))
    return dx
