"""CPU: the oracle restatement (oracle/odvae_oracle.py) against the golden fixtures produced by the UNMODIFIED
reference (tests/golden/make_golden.py). This is what pins the oracle; tolerance 1e-4 relative (fp32 CPU vs
fp32 CPU, different op order only)."""
import pytest
import torch

import helpers as H
from odvae_b200 import synth
from oracle import odvae_oracle

CPU_CASES = [c for c in H.CASES if c.startswith("small")] + [pytest.param("full_s80000_b2", marks=pytest.mark.slow)]


def run_oracle(name):
    g = H.load_golden(name)
    params, batch, noise, step = H.case_inputs(g)
    # state_dict with the reference's names: take the names/shapes from the product module (CPU construction)
    from odvae_b200.autoencoder import PoseAutoencoder
    model = PoseAutoencoder(**params)
    synth.fill_state_dict_(model.state_dict(), seed=23)
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in model.state_dict().items()
          if torch.is_floating_point(v) and not k.startswith("loss.discriminator")}
    orc = odvae_oracle.Oracle(params, sd, synth.synthetic_dataset_stats(23), synth.LABELS)
    loss, log, out = orc.training_step(H.clone_batch(batch), step, noise)
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
        if k.endswith("g_loss"):
            continue
        assert H.rel_scalar(v, g["log"][k], floor=1e-6) < 1e-4, k
    grads = {k: v.grad for k, v in sd.items()}
    for k in g["params_without_grad"]:
        assert grads[k] is None or float(grads[k].abs().max()) == 0.0, k
    errs = H.grad_errors(grads, g["grads"])
    worst = max(errs.items(), key=lambda kv: max(kv[1]))
    assert max(worst[1]) < 2e-3, worst  # fp32 summation-order noise through ~60 layers with 1e6 loss weights
