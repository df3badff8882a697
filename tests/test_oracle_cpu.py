"""CPU: the oracle restatement (oracle/odvae_oracle.py) against the golden fixtures produced by the UNMODIFIED
reference (tests/golden/make_golden.py). This is what pins the oracle; tolerance 1e-4 relative (fp32 CPU vs
fp32 CPU, different op order only)."""
import pytest
import torch

import helpers as H
from odvae_b200 import synth
from oracle import odvae_oracle

CPU_CASES = [c for c in H.CASES if c.startswith("small")] + [pytest.param("full_s80000_b2", marks=pytest.mark.slow)]
CPU_GAN_CASES = [c for c in H.GAN_CASES if c.startswith("small")] + [pytest.param("full_s80000_gan_b2", marks=pytest.mark.slow)]


def oracle_state(params, dtype=torch.float32):
    """state_dict with the reference's names: names/shapes come from the product module (CPU construction), values from
    the seeded fill. Trainable tensors become leaves; BatchNorm running statistics / counters stay plain buffers."""
    from odvae_b200.autoencoder import PoseAutoencoder
    model = PoseAutoencoder(**params)
    synth.fill_state_dict_(model.state_dict(), seed=23)
    sd = {}
    for k, v in model.state_dict().items():
        if k.startswith("loss.perceptual_loss"):
            continue
        if not torch.is_floating_point(v) or "running_" in k:
            sd[k] = v.detach().clone().to(dtype) if torch.is_floating_point(v) else v.detach().clone()
        else:
            sd[k] = v.detach().clone().to(dtype).requires_grad_(True)
    return sd


def run_oracle(name, dtype=torch.float32):
    g = H.load_golden(name)
    params, batch, noise, step = H.case_inputs(g)
    sd = oracle_state(params, dtype)
    orc = odvae_oracle.Oracle(params, sd, synth.synthetic_dataset_stats(23), synth.LABELS)
    loss, log, out = orc.training_step(H.clone_batch(batch), step, tuple(t.to(dtype) for t in noise))
    loss.backward()
    return g, sd, loss, log, out


@pytest.mark.parametrize("name", CPU_CASES)
def test_oracle_matches_reference_golden(name):
    g, sd, loss, log, out = run_oracle(name)
    dec_obj, dec_pose, post, bbox_post, p = out
    assert abs(p - g["spec"]["dropout_p"]) < 1e-12
    assert H.rel_max(dec_obj, g["out"]["dec_obj"]) < 1e-4 or g["out"]["dec_obj"].abs().max() == 0
    assert H.rel_max(dec_pose, g["out"]["dec_pose"]) < 1e-4
    assert H.rel_max(post.mean, g["out"]["posterior_mean"]) < 1e-4
    assert H.rel_max(post.logvar, g["out"]["posterior_logvar"]) < 1e-4
    assert H.rel_max(bbox_post.mean, g["out"]["bbox_mean"]) < 1e-4
    assert H.rel_max(bbox_post.logvar, g["out"]["bbox_logvar"]) < 1e-4
    assert H.rel_scalar(loss, g["loss"]) < 1e-4
    for k, v in log.items():
        assert H.rel_scalar(v, g["log"][k], floor=1e-6) < 1e-4, k  # incl. g_loss: the discriminator runs on every step
    grads = {k: v.grad for k, v in sd.items() if not k.startswith("loss.discriminator")}
    for k in g["params_without_grad"]:
        assert grads[k] is None or float(grads[k].abs().max()) == 0.0, k
    # every gradient tensor against the reference's own (stored whole / as a fixed subset): relative L2 of the DIFFERENCE and
    # max-abs / inf-norm, both within the fp32 tolerance (measured 2e-6: same arithmetic, different op order)
    errs = H.grad_diff_errors(grads, g["grads"])
    worst = max(errs.items(), key=lambda kv: max(kv[1]))
    # full config (2 x 196608 L1 terms): two fp32 runs that differ by 1e-6 in dec_obj flip the sign of a few |x - x_hat| ~ 0
    # pixels, which moves every upstream gradient by ~1e-3 (the float64 oracle vs the fp32 reference: 3-5e-4, see the GAN test)
    assert max(worst[1]) < (1e-4 if name.startswith("small") else 3e-3), (worst, H.error_table(errs))


# GAN cases. The gradient of g_loss passes through LeakyReLU(0.2) and train-mode BatchNorm of a tiny discriminator input
# (small config: 64x64 crops -> 6x6 logits): a pre-activation within fp32 noise of 0 flips its slope between 1 and 0.2 and
# moves ||d g / d W_last|| -- hence the adaptive d_weight and every gradient -- by ~1e-4..3e-3. Measured here: the fp32
# oracle differs from the fp32 reference by up to 2.9e-3 on small_s80000_gan_b3 while the SAME oracle in float64 agrees with
# the fp32 reference to 2.3e-6. The fp64 run therefore carries the 1e-4 pin; the fp32 run is bounded by the kink noise.
GAN_FP32_KINK_TOL = 1e-2
FULL_FP32_NOISE_TOL = 2e-3  # measured 5.3e-4 (float64 oracle vs fp32 reference, full_s80000_gan_b2)


def _check_gan_case(g, sd, loss, log, out, tol, grad_tol):
    dec_obj, dec_pose, post, bbox_post, p = out
    assert H.rel_max(dec_obj, g["out"]["dec_obj"]) < 1e-4 or g["out"]["dec_obj"].abs().max() == 0
    assert H.rel_scalar(loss, g["loss"]) < tol
    for k, v in log.items():
        assert H.rel_scalar(v, g["log"][k], floor=1e-6) < tol, (k, v, g["log"][k])
    grads = {k: v.grad for k, v in sd.items() if not k.startswith("loss.discriminator")}
    errs = H.grad_diff_errors(grads, g["grads"])
    worst = max(errs.items(), key=lambda kv: max(kv[1]))
    assert max(worst[1]) < grad_tol, (worst, H.error_table(errs))


@pytest.mark.parametrize("name", CPU_GAN_CASES)
def test_oracle_gan_generator_step_matches_reference_golden(name):
    """optimizer_idx 0 with disc_factor = 1: g_loss, adaptive d_weight (two autograd.grad calls on decoder.conv_out.weight),
    total loss and every autoencoder gradient. float64 oracle vs fp32 reference: 1e-4; float32 oracle: kink-noise bound."""
    g, sd, loss, log, out = run_oracle(name, torch.float64)
    # full config: the fp32 REFERENCE run itself sits 3-5e-4 from this float64 run (its own L1-sign / LeakyReLU kink flips over
    # 2 x 196608 pixels), so that case is bounded by FULL_FP32_NOISE_TOL instead of 1e-4
    _check_gan_case(g, sd, loss, log, out, 1e-4, 1e-4 if name.startswith("small") else FULL_FP32_NOISE_TOL)
    if name.startswith("small"):
        g, sd, loss, log, out = run_oracle(name, torch.float32)
        _check_gan_case(g, sd, loss, log, out, 2e-3, GAN_FP32_KINK_TOL)


@pytest.mark.parametrize("name", [c for c in H.GAN_CASES if c.startswith("small")])
def test_oracle_gan_discriminator_step_matches_reference_golden(name):
    """optimizer_idx 1 (LOSS:352-375): d_loss, logged logits, discriminator gradients, and the BatchNorm running statistics
    after the batch's three discriminator forwards (one in the generator step, two here)."""
    g = H.load_golden(name)
    params, batch, noise, step = H.case_inputs(g)
    sd = oracle_state(params, torch.float64)  # float64: hinge / LeakyReLU kink flips put the fp32 oracle 2e-3 away
    noise = tuple(t.double() for t in noise)
    orc = odvae_oracle.Oracle(params, sd, synth.synthetic_dataset_stats(23), synth.LABELS)
    orc.training_step(H.clone_batch(batch), step, noise)  # generator-side forward: updates the running statistics once
    d_loss, log = orc.discriminator_step(H.clone_batch(batch), step, noise)
    d_loss.backward()
    ref = g["disc"]
    assert H.rel_scalar(d_loss, ref["loss"], floor=1e-6) < 1e-4
    for k, v in log.items():
        assert H.rel_scalar(v, ref["log"][k], floor=1e-6) < 1e-4, k
    grads = {k: v.grad for k, v in sd.items() if k.startswith("loss.discriminator") and v.requires_grad}
    if ref["loss"] != 0.0:
        errs = H.grad_diff_errors(grads, ref["grads"])
        worst = max(errs.items(), key=lambda kv: max(kv[1]))
        assert max(worst[1]) < 1e-4, (worst, H.error_table(errs))
    for k, v in ref["buffers"].items():
        assert H.rel_max(sd[k].double(), v.double()) < 1e-5, k
