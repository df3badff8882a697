"""GPU: the PatchGAN discriminator kernels (csrc/patchgan.cu + odvae_b200/patchgan.py, through the C ABI) against the
CPU oracle restatement of taming's NLayerDiscriminator (oracle/odvae_oracle.py `discriminator`, float64), and the whole
GAN path (generator step with the adaptive d_weight, discriminator step) against the goldens made from the reference's
loss code (tests/golden/make_golden.py GAN cases). The discriminator class itself is un-vendored (parity unpinned)."""
import pytest
import torch

import helpers as H
from odvae_b200 import ops, synth
from oracle import odvae_oracle

pytestmark = pytest.mark.gpu

def _fmt(mode):
    return ops.format_for({"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}[mode])


def _make_disc(mode, seed=23):
    from odvae_b200.patchgan import NLayerDiscriminator
    D = NLayerDiscriminator(3)
    sd = {"loss.discriminator." + k: v for k, v in D.state_dict().items()}
    synth.fill_state_dict_(sd, seed=seed)
    D = D.cuda()
    D.fmt = ops.disc_format(_fmt(mode))
    return D, sd


def _oracle_run(sd, x, mask, dlogits_fn, dtype=torch.float64):
    s = {}
    for k, v in sd.items():
        if torch.is_floating_point(v):
            t = v.detach().clone().to(dtype)
            s[k] = t if "running_" in k else t.requires_grad_(True)
        else:
            s[k] = v.clone()
    xx = x.detach().cpu().to(dtype).requires_grad_(True)
    m = mask.detach().cpu().to(dtype)
    logits = odvae_oracle.discriminator(s, xx * m)
    loss = dlogits_fn(logits)
    loss.backward()
    return logits.detach(), xx.grad, s


# Gradient tolerances: d logits / d input passes through four LeakyReLU(0.2) kinks; a pre-activation within the forward's
# rounding noise of 0 flips its slope and moves the gradient discretely. fp32 at (2, 64) has ~3e5 units and no flip (6e-5
# holds); the larger fp32 cases see a few flips (measured 2e-3), and the 16-bit forwards (1e-3 / 1e-2 noise) flip many:
# measured 2.7e-2 (fp16) and 1e-1 (bf16) on the input gradient. The same kink noise separates the fp32 CPU oracle from the
# fp32 reference (tests/test_oracle_cpu.py).
@pytest.mark.parametrize("mode,tol,gtol", [("fp32", 2e-5, 1e-2), ("fp16", 6e-3, 6e-2), ("bf16", 3e-2, 2e-1)])
@pytest.mark.parametrize("B,R", [(2, 64), (3, 96), (4, 256)])
def test_discriminator_forward_backward_against_oracle(mode, tol, gtol, B, R):
    D, sd = _make_disc(mode)
    g = torch.Generator().manual_seed(5 + B)
    x = (torch.rand(B, 3, R, R, generator=g) * 2 - 1).cuda().requires_grad_(True)
    mask = (torch.rand(B, 1, R, R, generator=g) > 0.3).float().cuda()
    w = torch.randn(B, 1, R // 8 - 2, R // 8 - 2, generator=g)  # a non-constant upstream gradient (not just the mean)
    fn = lambda lg: (lg * w.to(lg)).sum() - lg.mean()
    logits = D(x, mask=mask)
    assert logits.shape == (B, 1, R // 8 - 2, R // 8 - 2) and logits.dtype == torch.float32
    ((logits * w.cuda()).sum() - logits.mean()).backward()
    torch.cuda.synchronize()
    ref_logits, ref_dx, s = _oracle_run(sd, x, mask, fn)
    assert H.rel_l2(logits, ref_logits) < tol
    if mode == "fp32" and (B, R) == (2, 64):
        gtol = 6e-5  # no kink flip at this size: the backward kernels themselves are exact to fp32 rounding
    assert H.rel_l2(x.grad, ref_dx) < gtol
    for k, p in D.named_parameters():
        r = s["loss.discriminator." + k].grad
        assert H.rel_l2(p.grad, r) < gtol, k
    for k, b in D.named_buffers():  # running statistics / counters updated like torch.nn.BatchNorm2d
        assert H.rel_max(b.double(), s["loss.discriminator." + k].double()) < max(tol, 1e-5), k


def test_discriminator_generator_side_call_leaves_parameters_without_gradients():
    D, _ = _make_disc("fp16")
    x = torch.rand(2, 3, 64, 64, device="cuda").requires_grad_(True)
    (-D(x, param_grads=False).mean()).backward()
    assert x.grad is not None and float(x.grad.abs().max()) > 0
    assert all(p.grad is None for p in D.parameters())


def test_discriminator_is_bit_stable():
    outs = []
    for _ in range(2):
        D, _ = _make_disc("fp16")
        x = (torch.rand(3, 3, 64, 64, generator=torch.Generator().manual_seed(1)) * 2 - 1).cuda().requires_grad_(True)
        lg = D(x)
        lg.square().mean().backward()
        outs.append((lg.detach().clone(), x.grad.clone(), [p.grad.clone() for p in D.parameters()]))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    for a, b in zip(outs[0][2], outs[1][2]):
        assert torch.equal(a, b)


# ---------------------------------------------------------------------------------------------------------------------
# whole path with the GAN on, against the goldens (reference loss code + restated discriminator)
# ---------------------------------------------------------------------------------------------------------------------
def _run_gan(name, dtype, single_backward=True):
    from odvae_b200.autoencoder import PoseAutoencoder
    g = H.load_golden(name)
    params, batch, noise, step = H.case_inputs(g)
    model = PoseAutoencoder(**params, compute_dtype=dtype)
    synth.fill_state_dict_(model.state_dict(), seed=23)
    model = model.cuda()
    model.global_step = step
    model.inject_noise = noise
    model.loss.skip_inactive_gan = False
    model.loss.single_backward_gan = single_backward
    loss = model.training_step(H.clone_batch(batch), 0, 0)
    loss.backward()
    log0 = dict(model.logged)
    grads = {k: (None if p.grad is None else p.grad.clone()) for k, p in model.named_parameters()}
    model.zero_grad(set_to_none=True)
    d_loss = model.training_step(H.clone_batch(batch), 0, 1)
    d_loss.backward()
    torch.cuda.synchronize()
    return g, model, loss, log0, grads, d_loss, dict(model.logged)


# fp32: the LeakyReLU / L1 kink noise documented in tests/test_oracle_cpu.py bounds the gradients (the CPU fp32 oracle
# itself differs from the fp32 reference by 2.9e-3 on small_s80000_gan_b3); 16-bit: the timed mode's stated tolerance
@pytest.mark.parametrize("name", H.GAN_CASES)
@pytest.mark.parametrize("mode", ["fp32", "fp16", "fp32-literal"])
def test_gan_steps_match_reference(name, mode):
    """mode "...-literal": PoseLoss.single_backward_gan = False, the reference's own two autograd.grad calls + full backward;
    the default runs every expensive backward once (PoseLoss._generator_terms_single_backward)."""
    single = not mode.endswith("-literal")
    mode = mode.split("-")[0]
    dtype = {"fp32": torch.float32, "fp16": torch.float16}[mode]
    g, model, loss, log0, grads, d_loss, log1 = _run_gan(name, dtype, single)
    # 16-bit gradient bound: kink noise (L1 sign + LeakyReLU slope flips) times the adaptive weight's amplification; the
    # small config's 6x6-logit discriminator has so few units that single flips show (measured 7e-2; full config 3e-2)
    tol, gtol, dtol = (2e-3, 1e-2, 2e-3) if mode == "fp32" else (1e-2, 5e-2 if name.startswith("full") else 1e-1, 2e-2)
    assert H.rel_scalar(loss, g["loss"]) < tol
    for k, ref in g["log"].items():
        t = tol * (10 if k.endswith("d_weight") else 1)  # ||grad g|| in the denominator: see the kink note above
        assert H.rel_scalar(log0[k], ref, floor=1e-6) < t, (k, float(log0[k]), ref)
    errs = H.grad_diff_errors({k: v for k, v in grads.items()
                               if not k.startswith(("loss.discriminator", "loss.perceptual_loss"))}, g["grads"])
    worst = max(errs.items(), key=lambda kv: kv[1][0])
    print(name, mode, H.error_table(errs, 4))
    if mode == "fp32":
        assert worst[1][0] < gtol, (worst, H.error_table(errs))
    else:
        # 16-bit: 99 % of the tensors inside the bound and the tail within twice it. The tail is one near-cancelling tensor
        # (pose_encoder.latent_linear.weight: the decoder-input gradient summed over every latent pixel), whose error moves
        # between 4.5e-2 and 5.5e-2 with the summation order of the softmax row statistics -- which kinks flip, not how many
        l2 = sorted(v[0] for v in errs.values())
        assert l2[int(0.99 * len(l2))] < gtol and worst[1][0] < 2 * gtol, (worst, H.error_table(errs))
    ref = g["disc"]
    assert H.rel_scalar(d_loss, ref["loss"], floor=1e-6) < dtol
    for k, v in ref["log"].items():
        assert H.rel_scalar(log1[k], v, floor=1e-3) < dtol, (k, float(log1[k]), v)
    if ref["loss"] != 0.0:
        dg = {k: p.grad for k, p in model.named_parameters() if k.startswith("loss.discriminator")}
        errs = H.grad_diff_errors(dg, ref["grads"])
        worst = max(errs.items(), key=lambda kv: kv[1][0])
        # discriminator gradients: hinge relu(1 -+ logit) and LeakyReLU kinks; the small config's 6x6 logit map (72 values at
        # B = 2) makes single flips visible in the BatchNorm bias gradients (measured up to 0.16 in fp16; full config 0.03)
        dgtol = 5 * dtol if (mode == "fp32" or name.startswith("full")) else 0.3
        assert worst[1][0] < dgtol, (worst, H.error_table(errs))
    for k, v in ref["buffers"].items():
        assert H.rel_max(model.state_dict()[k].double(), v.double()) < dtol, k


def test_discriminator_step_is_skipped_but_differentiable_while_the_gan_is_off():
    """optimizer_idx 1 must return a tensor Lightning can call backward() on, also before disc_start."""
    from odvae_b200.autoencoder import PoseAutoencoder
    params = synth.model_config(synth.DDCONFIG_SMALL, H.stats_path(23))
    params["lossconfig"]["params"]["disc_factor"] = 1.0
    model = PoseAutoencoder(**params, compute_dtype=torch.float16).cuda()
    model.global_step = 100  # < disc_start = 30000
    batch = synth.synthetic_batch(2, 64, 3)
    d_loss = model.training_step(batch, 0, 1)
    assert d_loss.requires_grad and float(d_loss) == 0.0
    d_loss.backward()
    assert all(p.grad is None or float(p.grad.abs().max()) == 0.0 for p in model.loss.discriminator.parameters())
