"""GPU unit tests of the memory-bound kernels through the C ABI against plain torch fp32 math on the same
(bf16-rounded) inputs. Tolerances: f32 kernels 1e-4 (2e-4 where a long fp32 reduction is involved), bf16 storage 1e-2,
fp16 storage 4e-3 (the fp16 GroupNorm backward computes in packed half2: a handful of 2^-11 roundings per element)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from odvae_b200 import ops  # noqa: E402

DEV = "cuda"


def rel(a, ref):
    a, ref = a.detach().float(), ref.detach().float()
    return float((a - ref).abs().max() / ref.abs().max().clamp_min(1e-20))


GN_TOL = {torch.float32: 2e-4, torch.bfloat16: 1e-2, torch.float16: 4e-3}


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(2, 16, 16, 128), (3, 32, 32, 64), (1, 64, 64, 256), (2, 8, 8, 512), (65, 16, 16, 128),
                                   (3, 25, 25, 64)])
@pytest.mark.parametrize("silu", [False, True])
def test_groupnorm_fwd_bwd(dtype, shape, silu):
    torch.manual_seed(0)
    B, H, W, C = shape
    x = (torch.randn(shape, device=DEV) * 1.5 + 0.3).to(dtype)
    gamma = 1 + 0.1 * torch.randn(C, device=DEV)
    beta = 0.1 * torch.randn(C, device=DEV)
    dy = torch.randn(shape, device=DEV).to(dtype)
    add = torch.randn(shape, device=DEV).to(dtype)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y_ref = F.group_norm(xr, 32, gr, br, eps=1e-6)
    if silu:
        y_ref = F.silu(y_ref)
    y_ref.backward(dy.float().permute(0, 3, 1, 2))
    stats = ops.gn_stats_raw(x, C)
    y = ops.gn_apply_raw(x, stats, gamma, beta, silu)
    dx, dg, db = ops.gn_bwd_raw(x, stats, gamma, beta, dy, silu, dx_add=add)
    dx0, dg0, db0 = ops.gn_bwd_raw(x, stats, gamma, beta, dy, silu)
    torch.cuda.synchronize()
    tol = GN_TOL[dtype]
    assert rel(y.permute(0, 3, 1, 2), y_ref) < tol
    assert rel(dx.float() - add.float(), xr.grad.permute(0, 2, 3, 1)) < 2 * tol
    assert rel(dx0, xr.grad.permute(0, 2, 3, 1)) < tol
    assert rel(dg, gr.grad) < tol and rel(db, br.grad) < tol
    assert torch.equal(dg, dg0) and torch.equal(db, db0)


@pytest.mark.parametrize("shape", [(2, 16, 16, 512), (65, 32, 32, 256), (3, 25, 25, 64), (1, 8, 8, 128)])
@pytest.mark.parametrize("silu", [False, True])
def test_groupnorm_small_maps_one_launch(shape, silu, monkeypatch):
    """fp16 feature maps of <= 1024 pixels: odvae_groupnorm_fwd_small (statistics + apply in one launch) and the one-launch
    backward odvae_groupnorm_bwd picks for them (ticketed dgamma / dbeta sums) against torch autograd, against the
    three-launch path, bit-stable over repeated calls."""
    torch.manual_seed(7)
    B, H, W, C = shape
    x = (torch.randn(shape, device=DEV) * 1.5 + 0.3).half()
    gamma = 1 + 0.1 * torch.randn(C, device=DEV)
    beta = 0.1 * torch.randn(C, device=DEV)
    dy = torch.randn(shape, device=DEV).half()
    add = torch.randn(shape, device=DEV).half()
    assert ops.gn_small_ok(H * W, C)
    stats = torch.empty(B, ops.GROUPS, 2, device=DEV)
    y = torch.empty_like(x)
    ops.call("groupnorm_fwd_small", x, gamma, beta, y, stats, B, H * W, C, ops.GROUPS, ops.GN_EPS, int(silu), ops.F16,
             ops.overflow_flags(x.device))
    stats3 = ops.gn_stats_raw(x, C)
    y3 = ops.gn_apply_raw(x, stats3, gamma, beta, silu)
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y_ref = F.group_norm(xr, 32, gr, br, eps=1e-6)
    if silu:
        y_ref = F.silu(y_ref)
    y_ref.backward(dy.float().permute(0, 3, 1, 2))
    tol = GN_TOL[torch.float16]
    assert rel(stats, stats3) < 1e-5
    assert rel(y.permute(0, 3, 1, 2), y_ref) < tol and rel(y, y3) < 1e-3
    outs = [ops.gn_bwd_raw(x, stats, gamma, beta, dy, silu, dx_add=add) for _ in range(3)]
    dx0, dg0, db0 = ops.gn_bwd_raw(x, stats, gamma, beta, dy, silu)
    torch.cuda.synchronize()
    for dx, dg, db in outs[1:]:
        assert torch.equal(dx, outs[0][0]) and torch.equal(dg, outs[0][1]) and torch.equal(db, outs[0][2])
    dx, dg, db = outs[0]
    assert rel(dx.float() - add.float(), xr.grad.permute(0, 2, 3, 1)) < 2 * tol
    assert rel(dx0, xr.grad.permute(0, 2, 3, 1)) < tol
    assert rel(dg, gr.grad) < tol and rel(db, br.grad) < tol
    assert torch.equal(dg, dg0) and torch.equal(db, db0)


@pytest.mark.parametrize("shape", [(2, 16, 16, 128), (2, 64, 64, 128)])  # one-launch small-map kernel / packed-half2 kernels
def test_groupnorm_bwd_flags_saturated_fp16_gradient(shape):
    """A dy value stored at +-65504 is a saturated store of its producer: both fp16 backward paths raise overflow flag [1]
    (and only then), and the outputs stay finite (saturating stores)."""
    torch.manual_seed(9)
    B, H, W, C = shape
    dev = torch.device("cuda", 0)
    x = torch.randn(shape, device=DEV).half()
    gamma, beta = torch.ones(C, device=DEV), torch.zeros(C, device=DEV)
    dy = torch.randn(shape, device=DEV).half()
    stats = ops.gn_stats_raw(x, C)
    ops.read_overflow(dev)
    ops.gn_bwd_raw(x, stats, gamma, beta, dy, True)
    assert ops.read_overflow(dev) == (False, False)
    dy[1, 3, 5, 7] = 65504.0
    dx, dg, db = ops.gn_bwd_raw(x, stats, gamma, beta, dy, True)
    assert ops.read_overflow(dev) == (False, True)
    assert bool(torch.isfinite(dx.float()).all()) and bool(torch.isfinite(dg).all())


@pytest.mark.parametrize("mean,std,gscale", [(8.0, 0.5, 1.0), (100.0, 3.0, 300.0), (-2.0, 40.0, 0.05)])
def test_groupnorm_bwd_fp16_offsets_and_scales(mean, std, gscale):
    """fp16 x / fp16 gradients (the timed mode; packed-half2 kernels of csrc/groupnorm.cu): a mean far from zero must not
    cost precision (x - fp16(mean) is exact where it matters), nor must large or small gradient magnitudes."""
    torch.manual_seed(3)
    shape, C = (2, 40, 40, 128), 128
    x = (torch.randn(shape, device=DEV) * std + mean).half()
    gamma = 1 + 0.1 * torch.randn(C, device=DEV)
    beta = 0.1 * torch.randn(C, device=DEV)
    dy = (torch.randn(shape, device=DEV) * gscale).half()
    xr = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    F.silu(F.group_norm(xr, 32, gr, br, eps=1e-6)).backward(dy.float().permute(0, 3, 1, 2))
    stats = ops.gn_stats_raw(x, C)
    dx, dg, db = ops.gn_bwd_raw(x, stats, gamma, beta, dy, True)
    torch.cuda.synchronize()
    tol = GN_TOL[torch.float16]
    assert rel(dx, xr.grad.permute(0, 2, 3, 1)) < tol
    assert rel(dg, gr.grad) < tol and rel(db, br.grad) < tol


@pytest.mark.parametrize("dtype,shape", [(torch.bfloat16, (13, 128, 128, 128)), (torch.bfloat16, (5, 256, 256, 128)),
                                         (torch.float32, (7, 128, 128, 128)), (torch.bfloat16, (50, 64, 64, 256)),
                                         (torch.float16, (13, 128, 128, 128)), (torch.float16, (5, 256, 256, 128))])
@pytest.mark.parametrize("with_add", [False, True])
def test_groupnorm_bwd_large(dtype, shape, with_add):
    """Tensors several times the L2 through the backward of csrc/groupnorm.cu (reduce -> finalize -> apply; fp16 pairs run
    the packed-half2 kernels): GroupNorm(32, C, 1e-6) + SiLU backward (LD model.py:33-39), checked against torch
    autograd; bit-stable run to run."""
    torch.manual_seed(5)
    B, H, W, C = shape
    x = (torch.randn(shape, device=DEV) * 1.5 + 0.3).to(dtype)
    gamma = 1 + 0.1 * torch.randn(C, device=DEV)
    beta = 0.1 * torch.randn(C, device=DEV)
    dy = torch.randn(shape, device=DEV).to(dtype)
    add = torch.randn(shape, device=DEV).to(dtype) if with_add else None
    stats = ops.gn_stats_raw(x, C)
    dx, dg, db = ops.gn_bwd_raw(x, stats, gamma, beta, dy, True, dx_add=add)
    dx2, dg2, db2 = ops.gn_bwd_raw(x, stats, gamma, beta, dy, True, dx_add=add)
    torch.cuda.synchronize()
    assert torch.equal(dx, dx2) and torch.equal(dg, dg2) and torch.equal(db, db2)
    tol = GN_TOL[dtype]
    dgr = torch.zeros(C, device=DEV)
    dbr = torch.zeros(C, device=DEV)
    for b in range(B):  # per-image autograd keeps the fp32 reference small
        xr = x[b:b + 1].float().permute(0, 3, 1, 2).requires_grad_(True)
        gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
        F.silu(F.group_norm(xr, 32, gr, br, eps=1e-6)).backward(dy[b:b + 1].float().permute(0, 3, 1, 2))
        got = dx[b:b + 1].float() - (add[b:b + 1].float() if with_add else 0)
        assert rel(got, xr.grad.permute(0, 2, 3, 1)) < 2 * tol, b
        dgr += gr.grad
        dbr += br.grad
    assert rel(dg, dgr) < tol and rel(db, dbr) < tol


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,cols,ld", [(4096, 128, 128), (1000, 3, 64), (77, 500, 500), (300000, 64, 64), (513, 512, 512)])
def test_colsum(dtype, rows, cols, ld):
    torch.manual_seed(1)
    m = torch.randn(rows, ld, device=DEV).to(dtype)
    out = torch.empty(cols, device=DEV)
    ws = ops.workspace(ops.call("colsum_workspace_bytes", cols), m.device, "colsum")
    ops.call("colsum", m, rows, cols, ld, ops.act_code(dtype), out, 0, ws, ws.numel())
    torch.cuda.synchronize()
    ref = m.float()[:, :cols].double().sum(0).float()
    assert float((out - ref).abs().max()) < 2e-4 * float(m.float().abs().double().sum(0).max())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [256, 4096])
def test_softmax_rows_fwd_bwd(dtype, n):
    torch.manual_seed(2)
    rows, scale = 300, 0.0625
    S = torch.randn(rows, n, device=DEV) * 8
    P = torch.empty(rows, n, dtype=dtype, device=DEV)
    ops.call("softmax_rows", S, P, rows, n, scale, ops.act_code(dtype))
    Sr = S.clone().requires_grad_(True)
    Pr = torch.softmax(Sr * scale, dim=-1)
    dP = torch.randn(rows, n, device=DEV)
    Pr.backward(dP)
    dS = torch.empty(rows, n, dtype=dtype, device=DEV)
    ops.call("softmax_rows_bwd", dP, P, dS, rows, n, scale, ops.act_code(dtype))
    torch.cuda.synchronize()
    tol = 1e-2 if dtype == torch.bfloat16 else 1e-4
    assert rel(P, Pr) < tol
    assert rel(dS, Sr.grad) < 2 * tol


def test_sample_kl_and_l1_against_torch():
    torch.manual_seed(3)
    B, HW, Cz = 3, 256, 16
    mom = torch.randn(B, 16, 16, 64, device=DEV) * 2
    mom[..., 16:32] *= 12  # exercise the clamp
    eps, noise = torch.randn(B, Cz, 16, 16, device=DEV), torch.randn(B, Cz, 16, 16, device=DEV)
    keep = (torch.rand(B, Cz, 16, 16, device=DEV) > 0.6).to(torch.uint8)
    pose = torch.randn(B, Cz * HW, device=DEV)
    momr = mom.clone().requires_grad_(True)
    poser = pose.clone().requires_grad_(True)
    mean, logvar, z, kl = ops.SampleKL.apply(momr, poser, eps, keep, noise, 2.5, Cz, 64, True)
    m = mom.permute(0, 3, 1, 2).clone().requires_grad_(True)
    mu, lv = m[:, :16], torch.clamp(m[:, 16:32], -30, 20)
    z_ref = (mu + torch.exp(0.5 * lv) * eps) * keep * 2.5 + noise + pose.view(B, Cz, 16, 16)
    kl_ref = 0.5 * (mu ** 2 + torch.exp(lv) - 1 - lv).sum(dim=[1, 2, 3])
    wz, wk = torch.randn_like(z_ref), torch.randn(B, device=DEV)
    ((z_ref * wz).sum() + (kl_ref * wk).sum()).backward()
    wz_nhwc = torch.zeros(B, 16, 16, 64, device=DEV)
    wz_nhwc[..., :16] = wz.permute(0, 2, 3, 1)
    ((z * wz_nhwc).sum() + (kl * wk).sum()).backward()
    torch.cuda.synchronize()
    assert rel(z[..., :16].permute(0, 3, 1, 2), z_ref) < 1e-5 and float(z[..., 16:].abs().max()) == 0
    assert rel(kl, kl_ref) < 1e-5 and rel(mean, mu) < 1e-6 and rel(logvar, lv) < 1e-6
    assert rel(momr.grad[..., :32].permute(0, 3, 1, 2), m.grad[:, :32]) < 1e-4
    assert rel(poser.grad.view(B, Cz, 16, 16), wz) < 1e-6
    # masked L1
    x, xh = torch.rand(B, 3, 32, 32, device=DEV), torch.randn(B, 3, 32, 32, device=DEV, requires_grad=True)
    mk = (torch.rand(B, 1, 32, 32, device=DEV) > 0.4).float()
    sums = ops.L1Masked.apply(x, xh, mk)
    w = torch.randn(B, device=DEV)
    (sums * w).sum().backward()
    xr = xh.detach().clone().requires_grad_(True)
    ref = (x * mk - xr * mk).abs().sum(dim=[1, 2, 3])
    (ref * w).sum().backward()
    torch.cuda.synchronize()
    assert rel(sums, ref) < 1e-5 and rel(xh.grad, xr.grad) < 1e-6
    # a SOFT mask (e.g. bilinearly resized): the backward multiplies the saved sign by the float mask
    soft = torch.rand(B, 1, 32, 32, device=DEV)
    xh2 = xh.detach().clone().requires_grad_(True)
    (ops.L1Masked.apply(x, xh2, soft) * w).sum().backward()
    xr2 = xh.detach().clone().requires_grad_(True)
    ((x * soft - xr2 * soft).abs().sum(dim=[1, 2, 3]) * w).sum().backward()
    torch.cuda.synchronize()
    assert rel(xh2.grad, xr2.grad) < 1e-6
    with pytest.raises(RuntimeError, match="one value per pixel"):
        ops.L1Masked.apply(x, xh2, soft[:, :, :16])


@pytest.mark.parametrize("B,N,C", [(2, 256, 512), (1, 1024, 256), (2, 4096, 256), (8, 4096, 256), (37, 1024, 128),
                                   (16, 4096, 256)])
def test_fused_attention_core_against_torch(B, N, C):
    """odvae_attn_scores_softmax / odvae_rowdot / odvae_attn_dsoftmax vs torch fp32 on the same bf16 inputs
    (AttnBlock core, LD model.py:187-197). Tolerance 1e-2 relative (bf16 outputs). In the last three shapes the softmax
    backward has enough (key tile, image) runs to take the engine's stationary-B-panel schedule (engine_desc.h bstat_mt:
    runs of 4 / 8 / 8 query tiles per resident v panel); the first three walk tile by tile."""
    torch.manual_seed(11)
    dt = torch.bfloat16
    qkv = torch.randn(B, N, 3 * C, device="cuda").to(dt)
    q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
    scale = float(C) ** -0.5
    P = torch.empty(B, N, N, dtype=dt, device="cuda")
    nbytes = ops.call("attn_softmax_workspace_bytes", B, N)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    ops.call("attn_scores_softmax", q, k, P, B, N, C, 3 * C, 3 * C, N * 3 * C, N * 3 * C, scale, ops.act_code(dt), ws,
             ws.numel())
    Pr = torch.softmax(torch.bmm(q.float(), k.float().transpose(1, 2)) * scale, dim=-1)
    assert float((P.float().sum(-1) - 1).abs().max()) < 2e-2
    assert float((P.float() - Pr).abs().max() / Pr.abs().max()) < 1e-2
    # backward: dS = P * (dP - rowsum(P*dP)) * scale with dP = dO v^T
    dO = torch.randn(B, N, C, device="cuda").to(dt)
    O = torch.bmm(P.float(), v.float()).to(dt)
    delta = torch.empty(B * N, dtype=torch.float32, device="cuda")
    ops.call("rowdot", dO, O, delta, B * N, C, C, C, ops.act_code(dt))
    dref = (dO.float() * O.float()).sum(-1).reshape(-1)
    assert float((delta - dref).abs().max() / dref.abs().max()) < 1e-5
    dS = torch.empty(B, N, N, dtype=dt, device="cuda")
    ops.call("attn_dsoftmax", dO, v, P, delta, dS, B, N, C, C, 3 * C, N * C, N * 3 * C, scale, ops.act_code(dt))
    dP = torch.bmm(dO.float(), v.float().transpose(1, 2))
    dSr = P.float() * (dP - delta.view(B, N, 1)) * scale
    assert float((dS.float() - dSr).abs().max() / dSr.abs().max()) < 1e-2


def test_multi_tensor_weight_pack_matches_single_packs():
    """refresh_param_caches(): one pack_conv_weights_mt launch over the device table must leave every packed copy
    bit-identical to odvae_pack_conv_weight of the same parameter (all layouts: forward, rotated dgrad, stride-2
    parity phases, 1x1; bf16 / fp16 / fp32 copies; padded channel counts)."""
    from odvae_b200.lib import BF16, CONV_1x1, CONV_3x3_S1, CONV_3x3_S2, F16, F32
    torch.manual_seed(11)
    specs, ws = [], []
    for cin, cout, geom in [(3, 128, CONV_3x3_S1), (128, 128, CONV_3x3_S1), (128, 3, CONV_3x3_S1), (128, 128, CONV_3x3_S2),
                            (256, 256, CONV_1x1), (128, 256, CONV_3x3_S1), (32, 16, CONV_1x1)]:
        k = 1 if geom == CONV_1x1 else 3
        w = torch.nn.Parameter(torch.randn(cout, cin, k, k, device=DEV))
        spec = ops.ConvSpec(cin, cout, geom)
        specs.append(spec)
        ws.append(w)
    packed = []
    for spec, w in zip(specs, ws):
        for code in (BF16, F16, F32):
            packed.append((spec, w, 0, code, spec.w_fwd(w, code)))
            packed.append((spec, w, 1, code, spec.w_dgrad(w, code)))
    with torch.no_grad():
        for w in ws:
            w.add_(torch.randn_like(w))  # bumps the parameter version: every cached copy is stale now
    ops.refresh_param_caches()
    torch.cuda.synchronize()
    for spec, w, which, code, buf in packed:
        again = spec.w_fwd(w, code) if which == 0 else spec.w_dgrad(w, code)
        assert again.data_ptr() == buf.data_ptr()  # cache hit: refreshed in place by the multi-tensor launch
        ref = ops.pack_weight(w.detach(), spec.geom, spec.coutp, spec.cinp, which, code)
        torch.cuda.synchronize()
        assert torch.equal(buf, ref), (spec.cin, spec.cout, spec.geom, which, code)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(2, 16, 16, 512), (3, 32, 64, 256), (1, 128, 128, 128), (5, 7, 9, 64)])
def test_upsample2x_matches_nearest_interpolate(dtype, shape):
    """Upsample.forward's F.interpolate(scale_factor=2.0, mode="nearest") (LD model.py:53-57) on NHWC: exact copy."""
    torch.manual_seed(9)
    x = torch.randn(shape, device=DEV).to(dtype)
    y = ops.upsample2x_raw(x)
    torch.cuda.synchronize()
    ref = F.interpolate(x.float().permute(0, 3, 1, 2), scale_factor=2.0, mode="nearest").permute(0, 2, 3, 1)
    assert torch.equal(y.float(), ref)


@pytest.mark.parametrize("C,Cpad,code", [(3, 64, "bf16"), (3, 64, "f16"), (16, 64, "f16"), (3, 8, "bf16"), (5, 24, "f16"), (3, 4, "f32"),
                                         (16, 64, "f32")])
def test_nchw_to_nhwc_pads_and_rescales(C, Cpad, code):
    """get_input/_rescale boundary (LD autoencoder.py:344-349, AE.py:434-436): NCHW fp32 -> padded NHWC with the affine
    `x * scale + shift` fused; padding channels are exact zeros."""
    from odvae_b200.lib import BF16, F16, F32
    torch.manual_seed(4)
    B, H, W = 3, 20, 12
    x = torch.rand(B, C, H, W, device=DEV)
    dc = {"bf16": BF16, "f16": F16, "f32": F32}[code]
    y = ops.nchw_to_nhwc_raw(x, Cpad, dc, 2.0, -1.0)
    torch.cuda.synchronize()
    tdt = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}[code]
    yv = y.view(tdt) if y.dtype != tdt else y
    assert tuple(yv.shape) == (B, H, W, Cpad)
    ref = (x * 2.0 - 1.0).permute(0, 2, 3, 1).to(tdt)
    # the kernel may contract x * scale + shift into one FMA: allow one rounding step of the output format
    assert float((yv[..., :C].float() - ref.float()).abs().max()) <= {"bf16": 2 ** -7, "f16": 2 ** -10, "f32": 2 ** -22}[code]
    if Cpad > C:
        assert float(yv[..., C:].float().abs().max()) == 0.0


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.float16, 3e-3), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("B,H,C,Cout", [(2, 16, 64, 64), (3, 32, 128, 128), (2, 64, 256, 128), (2, 128, 128, 128)])
def test_upsample_conv_phase_decomposition_against_torch(dtype, tol, B, H, C, Cout):
    """Upsample.forward (LD model.py:53-57): F.interpolate(2x, nearest) + Conv2d(3x3, p=1) -- run as four parity-phase 2x2
    convolutions over the low-resolution input: forward, input gradient (one 16-tap launch), weight and bias gradients
    (four 4-tap GEMMs folded back onto the 3x3 filter) against torch fp32 (TF32 off)."""
    import torch.nn.functional as F
    from odvae_b200 import modules as M
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator().manual_seed(B * 1000 + H)
    up = M.Upsample(C).cuda()
    if Cout != C:
        up = M.Upsample(C)
        up.conv = M.Conv2d(C, Cout, 3, up.conv.spec.geom)
        up = up.cuda()
    ops.assign_format(up, dtype)
    fmt = ops.format_for(dtype)
    xs = torch.randn(B, C, H, H, generator=g).cuda()
    x = ops.nchw_to_nhwc_raw(xs, C, fmt.fwd, label=fmt.label(fmt.fwd)).requires_grad_(True)
    y = up(x)
    assert y.shape == (B, 2 * H, 2 * H, Cout)
    wy = torch.randn(B, Cout, 2 * H, 2 * H, generator=g).cuda()
    dy = ops.nchw_to_nhwc_raw(wy, Cout, fmt.grad, label=fmt.label(fmt.grad))
    y.backward(dy)
    torch.cuda.synchronize()
    # reference on the values the kernels actually see (inputs / weights rounded to the 16-bit format)
    rd = {torch.float32: torch.float32, torch.float16: torch.float16, torch.bfloat16: torch.bfloat16}[dtype]
    xr = xs.to(rd).float().requires_grad_(True)
    wr = up.conv.weight.detach().to(rd).float().requires_grad_(True)
    br = up.conv.bias.detach().clone().requires_grad_(True)
    yr = F.conv2d(F.interpolate(xr, scale_factor=2.0, mode="nearest"), wr, br, padding=1)
    yr.backward(wy.to(rd).float())
    y_n = ops.nhwc_to_nchw_raw(y.detach(), Cout, fmt.fwd)
    dx_n = ops.nhwc_to_nchw_raw(x.grad, C, fmt.grad)
    rel = lambda a, b: float((a.float() - b.float()).norm() / b.float().norm())
    assert rel(y_n, yr) < tol
    assert rel(dx_n, xr.grad) < tol
    assert rel(up.conv.weight.grad, wr.grad) < tol
    assert rel(up.conv.bias.grad, br.grad) < tol


@pytest.mark.parametrize("dtype,tol", [(torch.float16, 3e-3), (torch.bfloat16, 2e-2), ("mixed", 2e-2)])
@pytest.mark.parametrize("B,H,W", [(2, 64, 64), (3, 32, 64), (1, 256, 256)])
def test_thin_edge_convs_against_torch(dtype, tol, B, H, W, monkeypatch):
    """encoder.conv_in (3 -> 128) forward / weight + bias gradient and decoder.conv_out (128 -> 3) input / weight / bias
    gradient through the thin path (odvae_im2col3x3_thin + ONE K = 64 step of the 1x1 tcgen05 path) against torch fp32 on
    the 16-bit-rounded operands, and against the generic 64-channel-padded path of the same layers."""
    import torch.nn.functional as F
    from odvae_b200 import modules as M
    from odvae_b200.lib import CONV_3x3_S1
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    g = torch.Generator().manual_seed(7 * B + H)
    fmt = ops.format_for(dtype)
    rel = lambda a, b: float((a.detach().float() - b.detach().float()).norm() / b.detach().float().norm())
    r16 = lambda t, code: (t.half() if code == ops.F16 else t.bfloat16()).float()
    # (mixed mode: the weight-gradient operand is the bf16 cast of the fp16 forward copy -- two roundings, 1.4e-3)
    wtol = 3e-3 if dtype == "mixed" else 1e-3

    # ---- conv_in ----
    conv = M.Conv2d(3, 128, 3, CONV_3x3_S1).cuda()
    ops.assign_format(conv, dtype)
    x = (torch.rand(B, 3, H, W, generator=g) * 2 - 1).cuda()
    wy = torch.randn(B, 128, H, W, generator=g).cuda()
    dy = ops.nchw_to_nhwc_raw(wy, 128, fmt.grad, label=fmt.label(fmt.grad))
    assert ops.thin_conv_ok(conv.spec, fmt.fwd, B, H, W, 128, 3)
    y = ops.ThinConvIn.apply(x, conv.weight, conv.bias, conv.spec)
    y.backward(dy)
    torch.cuda.synchronize()
    xr = r16(x, fmt.fwd)
    wr = r16(conv.weight.detach(), fmt.fwd).requires_grad_(True)
    br = conv.bias.detach().clone().requires_grad_(True)
    yr = F.conv2d(xr, wr, br, padding=1)
    assert rel(ops.nhwc_to_nchw_raw(y.detach(), 128, fmt.fwd), yr) < tol
    stats = getattr(y, "_odvae_gn_stats", None)
    if stats is not None:  # GroupNorm statistics from the epilogue: (mean, rstd) per (image, group)
        yv = ops.nhwc_to_nchw_raw(y.detach(), 128, fmt.fwd).view(B, 32, -1)
        assert float((stats[..., 0] - yv.mean(-1)).abs().max()) < 1e-3
    # weight gradient: contraction of the 16-bit col operand (gradient format) with the 16-bit dy
    xg = r16(x, fmt.grad)
    wg = conv.weight.detach().clone().requires_grad_(True)
    F.conv2d(xg, wg, None, padding=1).backward(r16(wy, fmt.grad))
    assert rel(conv.weight.grad, wg.grad) < wtol and rel(conv.bias.grad, r16(wy, fmt.grad).sum(dim=(0, 2, 3))) < 1e-3
    thin_dw = conv.weight.grad.clone()
    conv.weight.grad = conv.bias.grad = None
    h = ops.ToNHWC.apply(x, 64, dtype, None)
    conv(h, need_dx=False).backward(dy)
    assert rel(thin_dw, conv.weight.grad) < wtol

    # ---- conv_out ----
    conv = M.Conv2d(128, 3, 3, CONV_3x3_S1).cuda()
    ops.assign_format(conv, dtype)
    xs = torch.randn(B, 128, H, W, generator=g).cuda()
    dyo = torch.randn(B, 3, H, W, generator=g).cuda()
    res = {}
    for thin in (True, False):
        monkeypatch.setattr(ops, "THIN_CONV", thin)
        xh = ops.nchw_to_nhwc_raw(xs, 128, fmt.fwd, label=fmt.label(fmt.fwd)).requires_grad_(True)
        conv.weight.grad = conv.bias.grad = None
        yo = conv(xh, out_nchw=True)
        yo.backward(dyo)
        torch.cuda.synchronize()
        res[thin] = (yo.detach(), ops.nhwc_to_nchw_raw(xh.grad, 128, fmt.grad), conv.weight.grad.clone(), conv.bias.grad.clone())
    xr = r16(xs, fmt.fwd).requires_grad_(True)
    wr = r16(conv.weight.detach(), fmt.grad).requires_grad_(True)
    yr = F.conv2d(xr, wr, conv.bias.detach(), padding=1)
    yr.backward(r16(dyo, fmt.grad))
    wg = conv.weight.detach().clone().requires_grad_(True)
    F.conv2d(r16(xs, fmt.grad), wg, None, padding=1).backward(r16(dyo, fmt.grad))
    for thin in (True, False):
        yo, dx, dw, db = res[thin]
        assert rel(dx, xr.grad) < tol, thin
        assert rel(dw, wg.grad) < wtol and rel(db, r16(dyo, fmt.grad).sum(dim=(0, 2, 3))) < 1e-3, thin
    assert torch.equal(res[True][0], res[False][0])  # the forward is the same kernel either way
