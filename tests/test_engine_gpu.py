"""GPU parity of the GEMM / convolution engines (tcgen05 and SIMT) through the C ABI.

Reference here is torch fp32 math on the same (bf16-rounded) inputs: these are unit tests of single
kernels; whole-path parity against the oracle/goldens lives in test_parity_gpu.py.
Tolerances: bf16 tensor-core path <= 1e-2 relative (per-tensor max-abs / max-ref), fp32 SIMT <= 1e-4.
"""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from odvae_b200 import kernels as K  # noqa: E402
from odvae_b200.lib import CONV_1x1, CONV_3x3_S1, CONV_3x3_S2, IMPL_SIMT, IMPL_TC  # noqa: E402

DEV = "cuda"
# the torch reference must be true fp32 (cuDNN / cuBLAS default to TF32 on CUDA)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def rel_err(a, ref):
    a, ref = a.float(), ref.float()
    a, ref = a.detach(), ref.detach()
    return float((a - ref).abs().max() / ref.abs().max().clamp_min(1e-20))


def tol(dtype):
    return 1e-2 if dtype == torch.bfloat16 else 1e-4


def ref_conv(x_nhwc, w, b, geom):
    x = x_nhwc.float().permute(0, 3, 1, 2)
    if geom == CONV_3x3_S2:
        return F.conv2d(F.pad(x, (0, 1, 0, 1)), w, b, stride=2)
    if geom == CONV_1x1:
        return F.conv2d(x, w, b)
    return F.conv2d(x, w, b, padding=1)


CONV_CASES = [
    # B, H, W, Cin, Cout
    (2, 32, 32, 64, 128),
    (1, 16, 16, 128, 256),
    (2, 16, 16, 64, 64),
    (1, 128, 128, 64, 64),
    (2, 64, 64, 128, 128),   # N = 128 both ways: fwd and dgrad take the swapped-role tiles (M = channels, N = 256 pixels)
    (2, 256, 256, 64, 128),  # one full image row (256 pixels) per swapped tile, as in the 256x256 levels
    (3, 256, 256, 128, 128),  # halo-slab path with 2 channel chunks, fwd and dgrad (odd batch: rows cross images)
]


IMPLS3 = [pytest.param(IMPL_TC, torch.bfloat16, id="tc-bf16"), pytest.param(IMPL_SIMT, torch.float32, id="simt-f32"),
          pytest.param(IMPL_SIMT, torch.bfloat16, id="simt-bf16")]
IMPLS2 = IMPLS3[:2]


@pytest.mark.parametrize("impl,dtype", IMPLS3)
@pytest.mark.parametrize("geom", [CONV_3x3_S1, CONV_1x1, CONV_3x3_S2])
@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_fwd_dgrad_wgrad(impl, dtype, geom, case):
    B, H, W, Cin, Cout = case
    if impl == IMPL_TC and not K.conv_tc_supported(B, H, W, Cin, Cout, geom):
        pytest.skip("outside the tcgen05 engine's tiling domain (< 128 output pixels per image): SIMT engine serves it")
    torch.manual_seed(0)
    k = 1 if geom == CONV_1x1 else 3
    x = torch.randn(B, H, W, Cin, device=DEV).to(dtype)
    w = (torch.randn(Cout, Cin, k, k, device=DEV) / (Cin * k * k) ** 0.5)
    b = torch.randn(Cout, device=DEV)
    wq = w.to(dtype).float() if dtype == torch.bfloat16 else w
    Ho, Wo = K.geom_out_hw(geom, H, W)
    res = torch.randn(B, Ho, Wo, Cout, device=DEV).to(dtype)

    xr = x.float().requires_grad_(True)
    wr = wq.clone().requires_grad_(True)
    y_ref = ref_conv(xr, wr, b, geom) + res.float().permute(0, 3, 1, 2)
    dy = torch.randn(B, Ho, Wo, Cout, device=DEV).to(dtype)
    y_ref.backward(dy.float().permute(0, 3, 1, 2))

    wp = K.pack_conv_weight(w, geom, Cout, Cin, 0, dtype)
    y = K.conv_fwd(x, wp, b, Cout, geom, impl, residual=res)
    torch.cuda.synchronize()
    assert rel_err(y.permute(0, 3, 1, 2), y_ref) < tol(dtype)

    wd = K.pack_conv_weight(w, geom, Cout, Cin, 1, dtype)
    add = torch.randn(B, H, W, Cin, device=DEV).to(dtype)
    dx = K.conv_dgrad(dy, wd, Cin, geom, impl, (H, W), dx_add=add)
    torch.cuda.synchronize()
    assert rel_err(dx.float() - add.float(), xr.grad) < 2 * tol(dtype)

    dw, db = K.conv_wgrad(x, dy, Cin, Cout, geom, impl)
    torch.cuda.synchronize()
    assert rel_err(dw, wr.grad) < tol(dtype)
    assert rel_err(db, dy.float().sum(dim=(0, 1, 2))) < tol(dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,H,W", [(2, 256, 256), (1, 64, 512), (3, 32, 512), (3, 128, 128), (5, 6, 128)])
def test_conv_halo_slabs_match_tap_gather(dtype, B, H, W):
    """3x3 stride-1 at W = 256 (one tile per image row), W = 512 (two tiles per row, the configs[4] crops; the seam
    at x = 256 is an interior halo, not zero padding) and W = 128 (fp16 only: one tile = two pixel-interleaved image rows,
    engine_desc.h halo == 2): the shared-memory halo-slab path and the per-tap TMA gather accumulate the same
    products in the same order per output element only up to the K ordering (chunk-major vs tap-major), so they agree
    to fp32 summation noise; both are checked against torch fp32."""
    from odvae_b200 import lib
    C = 128
    torch.manual_seed(5)
    x = torch.randn(B, H, W, C, device=DEV).to(dtype)
    w = torch.randn(C, C, 3, 3, device=DEV) / (C * 9) ** 0.5
    b = torch.randn(C, device=DEV)
    res = torch.randn(B, H, W, C, device=DEV).to(dtype)
    wp = K.pack_conv_weight(w, CONV_3x3_S1, C, C, 0, dtype)
    wd = K.pack_conv_weight(w, CONV_3x3_S1, C, C, 1, dtype)
    outs = []
    try:
        for on in (1, 0):
            lib.call("set_conv_halo", on)
            y = K.conv_fwd(x, wp, b, C, CONV_3x3_S1, IMPL_TC, residual=res)
            dx = K.conv_dgrad(x, wd, C, CONV_3x3_S1, IMPL_TC, (H, W))  # dx in the operand format
            # wgrad: same products in the same K order -> bit-identical (gradients are bf16 in every 16-bit mode)
            dw, _ = K.conv_wgrad(x.to(torch.bfloat16), res.to(torch.bfloat16), C, C, CONV_3x3_S1, IMPL_TC)
            torch.cuda.synchronize()
            outs.append((y.float(), dx.float(), dw))
    finally:
        lib.call("set_conv_halo", 1)
    ref = ref_conv(x, w.to(dtype).float(), b, CONV_3x3_S1) + res.float().permute(0, 3, 1, 2)
    assert rel_err(outs[0][0].permute(0, 3, 1, 2), ref) < 1e-2
    ref_dx = F.conv_transpose2d(x.float().permute(0, 3, 1, 2), w.to(dtype).float(), padding=1)
    assert rel_err(outs[0][1].permute(0, 3, 1, 2), ref_dx) < 1e-2
    assert rel_err(outs[0][0], outs[1][0]) < 4e-3 and rel_err(outs[0][1], outs[1][1]) < 4e-3
    # image borders: halo columns / rows must be exact zeros, not neighbouring-row pixels
    edge = torch.cat([outs[0][0][:, :, 0], outs[0][0][:, :, -1], outs[0][0][:, 0], outs[0][0][:, -1]], 1)
    edge_ref = torch.cat([outs[1][0][:, :, 0], outs[1][0][:, :, -1], outs[1][0][:, 0], outs[1][0][:, -1]], 1)
    assert rel_err(edge, edge_ref) < 4e-3
    if W > 256:  # the columns either side of the tile seam
        assert rel_err(outs[0][0][:, :, 254:258], outs[1][0][:, :, 254:258]) < 4e-3
        assert rel_err(outs[0][1][:, :, 254:258], outs[1][1][:, :, 254:258]) < 4e-3
    assert rel_err(outs[0][2], outs[1][2]) < 1e-5


@pytest.mark.parametrize("B,H,W", [(2, 256, 256), (3, 128, 128), (1, 64, 512)])
def test_dgrad_with_fused_groupnorm_backward_sums(B, H, W):
    """odvae_conv2d_dgrad_gnbwd: the dgrad of a 128-channel 3x3 conv whose fp16 epilogue also produces the GroupNorm+SiLU
    backward sums per pixel box; odvae_groupnorm_bwd_from_partials then skips the reduce pass. dx of the conv is bit-identical
    to odvae_conv2d_dgrad; the GroupNorm backward agrees with the two-pass kernels and with torch autograd."""
    from odvae_b200 import lib, ops
    C, G = 128, 32
    torch.manual_seed(11)
    dy = torch.randn(B, H, W, C, device=DEV).half()
    gx = (torch.randn(B, H, W, C, device=DEV) * 1.5 + 0.3).half()  # the GroupNorm input
    add = torch.randn(B, H, W, C, device=DEV).half()
    gamma = 1 + 0.1 * torch.randn(C, device=DEV)
    beta = 0.1 * torch.randn(C, device=DEV)
    w = torch.randn(C, C, 3, 3, device=DEV) / (C * 9) ** 0.5
    wd = K.pack_conv_weight(w, CONV_3x3_S1, C, C, 1, torch.float16)
    stats = ops.gn_stats_raw(gx, C)
    nbytes = lib.call("conv2d_dgrad_gnbwd_workspace_bytes", B, H, W, C, C, CONV_3x3_S1, G)
    assert nbytes == B * (H * W // 128) * C * 8
    part = torch.empty(nbytes // 4, device=DEV)
    dx = torch.empty(B, H, W, C, dtype=torch.float16, device=DEV)
    lib.call("conv2d_dgrad_gnbwd", dy, wd, dx, gx, stats, gamma, beta, part, nbytes, B, H, W, C, C, CONV_3x3_S1, G, F16)
    dx_plain = K.conv_dgrad(dy, wd, C, CONV_3x3_S1, IMPL_TC, (H, W))
    torch.cuda.synchronize()
    assert torch.equal(dx, dx_plain)
    nbox = H * W // 128
    for dx_add in (None, add):
        a = ops.gn_bwd_raw(gx, stats, gamma, beta, dx, True, dx_add=dx_add, partials=(part, nbox))
        b = ops.gn_bwd_raw(gx, stats, gamma, beta, dx, True, dx_add=dx_add)
        torch.cuda.synchronize()
        for u, v in zip(a, b):
            assert rel_err(u, v) < 2e-3
    xr = gx[:1].float().permute(0, 3, 1, 2).requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    F.silu(F.group_norm(xr, 32, gr, br, eps=1e-6)).backward(dx[:1].float().permute(0, 3, 1, 2))
    assert rel_err(a[0][:1].float() - add[:1].float(), xr.grad.permute(0, 2, 3, 1)) < 8e-3


@pytest.mark.parametrize("impl,dtype", IMPLS2)
def test_conv_out_nchw_thin(impl, dtype):
    """decoder.conv_out: 128 -> 3 channels written as NCHW fp32 (weights padded to 16 rows)."""
    B, H, W, Cin, Cout = 2, 32, 32, 128, 3
    torch.manual_seed(1)
    x = torch.randn(B, H, W, Cin, device=DEV).to(dtype)
    w = torch.randn(Cout, Cin, 3, 3, device=DEV) / (Cin * 9) ** 0.5
    b = torch.randn(Cout, device=DEV)
    wq = w.to(dtype).float()
    wp = K.pack_conv_weight(w, CONV_3x3_S1, 16, Cin, 0, dtype)
    y = K.conv_fwd(x, wp, b, Cout, CONV_3x3_S1, impl, out_nchw_f32=True)
    torch.cuda.synchronize()
    assert y.shape == (B, Cout, H, W) and y.dtype == torch.float32
    assert rel_err(y, ref_conv(x, wq, b, CONV_3x3_S1)) < tol(dtype)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("B,H,W", [(3, 8, 256), (1, 4, 512), (2, 64, 64)])
def test_conv_out_nchw_swapped_tiles(dtype, B, H, W):
    """decoder.conv_out at full width: weights padded to ONE 128-row tile so the layer takes the swapped-role (halo-slab
    at W % 256 == 0) tiles; lanes >= Cout are not stored. Must equal the N = 16 tile result and torch."""
    Cin, Cout = 128, 3
    torch.manual_seed(2)
    x = torch.randn(B, H, W, Cin, device=DEV).to(dtype)
    w = torch.randn(Cout, Cin, 3, 3, device=DEV) / (Cin * 9) ** 0.5
    b = torch.randn(Cout, device=DEV)
    y128 = K.conv_fwd(x, K.pack_conv_weight(w, CONV_3x3_S1, 128, Cin, 0, dtype), b, Cout, CONV_3x3_S1, IMPL_TC, out_nchw_f32=True)
    y16 = K.conv_fwd(x, K.pack_conv_weight(w, CONV_3x3_S1, 16, Cin, 0, dtype), b, Cout, CONV_3x3_S1, IMPL_TC, out_nchw_f32=True)
    torch.cuda.synchronize()
    assert y128.shape == (B, Cout, H, W) and y128.dtype == torch.float32
    assert rel_err(y128, ref_conv(x, w.to(dtype).float(), b, CONV_3x3_S1)) < tol(dtype)
    assert rel_err(y128, y16) < 1e-5


GEMM_CASES = [
    # M, N, K, batch
    (256, 256, 128, 2),
    (128, 512, 64, 1),
    (200, 192, 72, 3),
    (1024, 1024, 256, 2),
]


@pytest.mark.parametrize("impl,dtype", IMPLS2)
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("case", GEMM_CASES)
@pytest.mark.parametrize("out_f32", [False, True])
def test_gemm(impl, dtype, a_mn, b_mn, case, out_f32):
    M, N, Kd, batch = case
    torch.manual_seed(2)
    A = torch.randn(batch, M, Kd, device=DEV).to(dtype)
    Bm = torch.randn(batch, N, Kd, device=DEV).to(dtype)
    ref = 0.5 * torch.bmm(A.float(), Bm.float().transpose(1, 2))
    As = A.transpose(1, 2).contiguous() if a_mn else A
    Bs = Bm.transpose(1, 2).contiguous() if b_mn else Bm
    out_dtype = torch.float32 if out_f32 else dtype
    D = K.gemm(As, Bs, M, N, Kd, batch=batch, strideA=M * Kd, strideB=N * Kd, a_mn=a_mn, b_mn=b_mn,
               out_dtype=out_dtype, alpha=0.5, impl=impl)
    torch.cuda.synchronize()
    assert rel_err(D, ref) < tol(dtype)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("case", [(512, 256, 1024, 3), (1000, 256, 200, 2), (1024, 256, 4096, 2), (256, 512, 128, 1)])
def test_gemm_m_tile_pairs(dtype, a_mn, b_mn, case, monkeypatch):
    """K-streaming GEMM with two m-tiles per CTA tile sharing every B tile (csrc/engine_desc.h `mpair`; the attention
    P v / P^T dO / dS k / dS^T q shape class): forced on for every shape that allows it, and the same call with the
    plain schedule must give the same bits (same MMAs per output, same K order)."""
    M, N, Kd, batch = case
    torch.manual_seed(4)
    A = torch.randn(batch, M, Kd, device=DEV).to(dtype)
    Bm = torch.randn(batch, N, Kd, device=DEV).to(dtype)
    ref = 0.25 * torch.bmm(A.float(), Bm.float().transpose(1, 2))
    As = A.transpose(1, 2).contiguous() if a_mn else A
    Bs = Bm.transpose(1, 2).contiguous() if b_mn else Bm
    from odvae_b200 import ops
    out = {}
    for mode in ("2", "0"):
        monkeypatch.setenv("ODVAE_GEMM_MPAIR", mode)
        D = torch.empty(batch, M, N, dtype=dtype, device=DEV)
        out[mode] = ops.gemm_raw(As, Bs, D, M, N, Kd, batch, M if a_mn else Kd, N if b_mn else Kd, N, M * Kd, N * Kd, M * N,
                                 a_mn, b_mn, alpha=0.25, code=ops.act_code(dtype))
    torch.cuda.synchronize()
    assert rel_err(out["2"], ref) < (2e-3 if dtype == torch.float16 else 1e-2)
    assert torch.equal(out["2"], out["0"])


def test_gemm_tc_strided_views():
    """q/k/v live interleaved in one [B,N,3C] tensor (row stride 3C); S = q k^T, O = P v."""
    Bn, N, C = 2, 256, 128
    torch.manual_seed(3)
    qkv = torch.randn(Bn, N, 3 * C, device=DEV).bfloat16()
    q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
    S = K.gemm(q, k, N, N, C, batch=Bn, lda=3 * C, ldb=3 * C, strideA=N * 3 * C, strideB=N * 3 * C,
               out_dtype=torch.float32, alpha=C ** -0.5)
    torch.cuda.synchronize()
    S_ref = torch.bmm(q.float(), k.float().transpose(1, 2)) * C ** -0.5
    assert rel_err(S, S_ref) < 1e-2
    P = torch.softmax(S_ref, dim=-1).bfloat16()
    O = K.gemm(P, v, N, C, N, batch=Bn, ldb=3 * C, strideA=N * N, strideB=N * 3 * C, b_mn=True)
    torch.cuda.synchronize()
    assert rel_err(O, torch.bmm(P.float(), v.float())) < 1e-2


# ------------------------------------------------------------------------------------------------
# fp16 / bf16 operand formats with independent output formats through odvae_gemm_16. A and B must share one format:
# on B200 a tcgen05.mma kind::f16 with different A/B formats traps (illegal instruction, measured in round 1), so the
# C ABI rejects that pair up front.
# ------------------------------------------------------------------------------------------------
from odvae_b200.lib import BF16, F16, F32, call  # noqa: E402

_CODE = {torch.bfloat16: BF16, torch.float16: F16, torch.float32: F32}


@pytest.mark.parametrize("a_dt,b_dt,o_dt", [
    (torch.float16, torch.float16, torch.float16),
    (torch.float16, torch.float16, torch.bfloat16),
    (torch.float16, torch.float16, torch.float32),
    (torch.bfloat16, torch.bfloat16, torch.float16),
])
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (1, 1), (0, 1)])
def test_gemm_mixed_16bit_formats(a_dt, b_dt, o_dt, a_mn, b_mn):
    torch.manual_seed(5)
    M, N, K, batch = 256, 192, 320, 2
    A = torch.randn(batch, K, M, device=DEV) if a_mn else torch.randn(batch, M, K, device=DEV)
    Bm = torch.randn(batch, K, N, device=DEV) if b_mn else torch.randn(batch, N, K, device=DEV)
    A, Bm = A.to(a_dt), Bm.to(b_dt)
    D = torch.empty(batch, M, N, dtype=o_dt, device=DEV)
    call("gemm_16", A, Bm, D, M, N, K, batch, M if a_mn else K, N if b_mn else K, N, M * K, N * K, M * N, a_mn, b_mn,
         _CODE[a_dt], _CODE[b_dt], _CODE[o_dt], 0.5, None, None)
    Af = A.float().transpose(1, 2) if a_mn else A.float()
    Bf = Bm.float() if b_mn else Bm.float().transpose(1, 2)
    ref = 0.5 * torch.bmm(Af, Bf)
    assert rel_err(D, ref) < (1e-2 if o_dt == torch.bfloat16 else (2e-3 if o_dt == torch.float16 else 1e-5))


def test_gemm_rejects_mixed_operand_formats():
    A = torch.zeros(128, 64, dtype=torch.float16, device=DEV)
    Bm = torch.zeros(128, 64, dtype=torch.bfloat16, device=DEV)
    D = torch.empty(128, 128, dtype=torch.float32, device=DEV)
    with pytest.raises(RuntimeError):
        call("gemm_16", A, Bm, D, 128, 128, 64, 1, 64, 64, 128, 0, 0, 0, 0, 0, F16, BF16, F32, 1.0, None, None)


@pytest.mark.parametrize("M,N,Kd", [(64, 500, 4096), (64, 27, 500), (5, 4096, 500), (64, 500, 500), (33, 19, 260)])
@pytest.mark.parametrize("b_mn", [False, True])
def test_skinny_fp32_linear(M, N, Kd, b_mn):
    """Pose-MLP Linears (M = batch rows, fp32): the K-split skinny kernels behind odvae_gemm_simt, forward layout
    (weights [N][K]) and the dx layout (weights [K][N]), against torch fp32 (<= 1e-4, fixed-order sums)."""
    torch.manual_seed(3)
    A = torch.randn(M, Kd, device=DEV)
    W = torch.randn(Kd, N, device=DEV) if b_mn else torch.randn(N, Kd, device=DEV)
    bias = torch.randn(N, device=DEV)
    ref = 0.5 * (A @ (W if b_mn else W.t())) + bias
    out = torch.empty(1, M, N, device=DEV)
    K.gemm(A, W, M, N, Kd, b_mn=b_mn, alpha=0.5, bias=bias, out=out, impl=IMPL_SIMT)
    out2 = torch.empty(1, M, N, device=DEV)
    K.gemm(A, W, M, N, Kd, b_mn=b_mn, alpha=0.5, bias=bias, out=out2, impl=IMPL_SIMT)
    torch.cuda.synchronize()
    assert rel_err(out[0], ref) < 1e-4
    assert torch.equal(out, out2)  # bit-stable run to run


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("H", [256, 64])
def test_conv_fwd_fused_groupnorm_stats(dtype, H):
    """odvae_conv2d_fwd_gnstats: GroupNorm(32) statistics of the conv output from the epilogue's fp32 accumulators
    (halo tiles at 256, per-tap gather tiles at 64) against the separate statistics pass over the stored output."""
    from odvae_b200 import lib
    B, W, C, G = 3, H, 128, 32
    torch.manual_seed(7)
    x = torch.randn(B, H, W, C, device=DEV).to(dtype)
    w = torch.randn(C, C, 3, 3, device=DEV) / (C * 9) ** 0.5
    b = torch.randn(C, device=DEV)
    res = torch.randn(B, H, W, C, device=DEV).to(dtype)
    wp = K.pack_conv_weight(w, CONV_3x3_S1, C, C, 0, dtype)
    code = K.act_code(dtype)
    nbytes = lib.call("conv2d_fwd_gnstats_workspace_bytes", B, H, W, C, C, CONV_3x3_S1, G)
    assert nbytes > 0
    ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
    y = torch.empty(B, H, W, C, dtype=dtype, device=DEV)
    stats = torch.empty(B, G, 2, device=DEV)
    lib.call("conv2d_fwd_gnstats", x, wp, b, res, y, stats, B, H, W, C, C, CONV_3x3_S1, G, 1e-6, code, code, ws, nbytes)
    y_plain = K.conv_fwd(x, wp, b, C, CONV_3x3_S1, IMPL_TC, residual=res)
    torch.cuda.synchronize()
    assert torch.equal(y, y_plain)  # the statistics ride along, the output is unchanged
    yf = y.float().reshape(B, H * W, G, C // G)
    mean = yf.mean(dim=(1, 3))
    rstd = (yf.var(dim=(1, 3), unbiased=False) + 1e-6).rsqrt()
    assert rel_err(stats[..., 0], mean) < 2e-3  # fp32 accumulators vs the 16-bit stored values
    assert rel_err(stats[..., 1], rstd) < 2e-3
    # shapes the swapped tiles do not serve report 0 and the caller keeps the separate pass
    assert lib.call("conv2d_fwd_gnstats_workspace_bytes", B, 16, 16, 512, 512, CONV_3x3_S1, G) == 0
