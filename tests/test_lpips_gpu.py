"""LPIPS on the B200 kernels against the CPU restatement oracle/lpips_oracle.py. PARITY UNPINNED: taming's
LPIPS source and vgg.pth are not available offline, so both sides follow the public definition with the same seeded
stand-in weights; what is pinned here is that the CUDA path computes that definition (value and gradient)."""
import os
import sys

import pytest
import torch

import helpers as H
from odvae_b200 import synth

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle.lpips_oracle import LPIPSOracle, fill_lpips_state_  # noqa: E402

pytestmark = pytest.mark.gpu


def _pair(B, R, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, 3, R, R, generator=g) * 2 - 1
    y = (x + 0.3 * torch.randn(B, 3, R, R, generator=g)).clamp(-1, 1)
    return x, y


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-4), (torch.bfloat16, 3e-2)])
@pytest.mark.parametrize("B,R", [(2, 64), (1, 256)])
def test_lpips_value_and_gradient_match_oracle(dtype, tol, B, R):
    from odvae_b200.lpips import LPIPS
    ref = LPIPSOracle()
    fill_lpips_state_(ref.state_dict())
    mod = LPIPS(compute_dtype=dtype)
    mod.load_state_dict(ref.state_dict())  # same key layout as taming's LPIPS
    mod = mod.cuda()
    x, y = _pair(B, R, 3)
    yr = y.clone().requires_grad_(True)
    with torch.enable_grad():
        v_ref = ref(x, yr)
        (v_ref.sum() * 1.7).backward()
    yg = y.cuda().requires_grad_(True)
    v = mod(x.cuda(), yg)
    (v.sum() * 1.7).backward()
    torch.cuda.synchronize()
    assert v.shape == (B, 1, 1, 1)
    assert H.rel_max(v, v_ref) < tol
    # gradients cross 13 conv layers + max-pool routing: fp32 summation-order noise reaches ~6e-4 at 256x256
    assert H.rel_l2(yg.grad, yr.grad) < (5 * tol if dtype == torch.bfloat16 else 2e-3)
    # bit-stable run to run
    yg2 = y.cuda().requires_grad_(True)
    v2 = mod(x.cuda(), yg2)
    (v2.sum() * 1.7).backward()
    assert torch.equal(v, v2) and torch.equal(yg.grad, yg2.grad)


def test_pose_loss_with_perceptual_term_matches_oracle():
    """PoseLoss(perceptual_weight=1) inside the full train step (small config, fp32 mode): total loss, rec/nll log entries
    and the decoder gradients against the CPU oracle carrying the same LPIPS weights."""
    from oracle import odvae_oracle
    from odvae_b200.autoencoder import PoseAutoencoder
    g = H.load_golden("small_s80000_b3")
    params, batch, noise, step = H.case_inputs(g)
    params["lossconfig"]["params"]["perceptual_weight"] = 1.0
    model = PoseAutoencoder(**params, compute_dtype=torch.float32)
    synth.fill_state_dict_(model.state_dict(), seed=23)
    lp_ref = LPIPSOracle()
    fill_lpips_state_(lp_ref.state_dict())
    model.loss.perceptual_loss.load_state_dict(lp_ref.state_dict())
    model.loss.perceptual_loss.compute_dtype = torch.float32
    sd = {k: v.detach().clone() for k, v in model.state_dict().items() if not k.startswith("loss.perceptual_loss.")}
    model = model.cuda()
    model.global_step = step
    model.inject_noise = noise
    loss = model.training_step(H.clone_batch(batch), 0, 0)
    loss.backward()
    torch.cuda.synchronize()

    sd_ref = {k: v.float().cpu().requires_grad_(v.is_floating_point()) for k, v in sd.items()}
    orc = odvae_oracle.Oracle(params, sd_ref, synth.synthetic_dataset_stats(23), synth.LABELS)
    orc.lpips = lp_ref
    loss_ref, log_ref, _ = orc.training_step(H.clone_batch(batch), step, noise)
    loss_ref.backward()
    assert H.rel_scalar(loss, loss_ref) < 2e-4
    assert H.rel_scalar(model.logged["train/rec_loss"], log_ref["train/rec_loss"]) < 2e-4
    assert H.rel_scalar(model.logged["train/nll_loss"], log_ref["train/nll_loss"]) < 2e-4
    for k in ("decoder.conv_out.weight", "decoder.up.0.block.0.conv1.weight", "encoder.conv_in.weight"):
        gp = dict(model.named_parameters())[k].grad
        assert H.rel_l2(gp, sd_ref[k].grad) < 2e-3, k
