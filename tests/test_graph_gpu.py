"""GPU: the whole-step CUDA graph (odvae_b200/graphs.py GraphedTrainStep) against the eager launch sequence
(trainer.TrainStep), and FusedAdam's device-resident per-parameter step counters against torch.optim.Adam."""
import pytest
import torch

import helpers as H
from odvae_b200 import synth

pytestmark = pytest.mark.gpu


def _model(step, dtype=torch.float16, disc_factor=0.0):
    from odvae_b200.autoencoder import PoseAutoencoder
    params = synth.model_config(synth.DDCONFIG_SMALL, H.stats_path(23))
    params["lossconfig"]["params"]["disc_factor"] = disc_factor
    model = PoseAutoencoder(**params, compute_dtype=dtype)
    synth.fill_state_dict_(model.state_dict(), seed=23)
    model = model.cuda()
    model.global_step = step
    return model


@pytest.mark.parametrize("start,disc", [(79998, 0.0), (100, 0.0), (74998, 1.0)])
def test_graphed_train_step_is_bit_identical_to_the_eager_sequence(start, disc):
    """Six optimizer steps with a changing batch: same loss every step and the same parameters at the end, bit for bit. start =
    79998: the dropout ramp (1/(1-p) changes every step and is read from device memory); 100: the encoder-pretrain phase (decoder
    parameters inactive: not updated, not aged); 74998 with the GAN on: the run crosses the 75000 threshold where dropout starts to
    ramp -> the phase key changes and a second graph is captured (generator term + adaptive weight inside the graph)."""
    from odvae_b200.graphs import GraphedTrainStep
    from odvae_b200.trainer import TrainStep
    B = 3
    batches = [synth.synthetic_batch(B, 64, seed=40 + i) for i in range(6)]
    noises = [synth.draw_noise(B, 0.8, seed=60 + i) for i in range(6)]
    out = []
    for graphed in (False, True):
        model = _model(start, disc_factor=disc)
        step = TrainStep(model, lr=1e-4, max_grad_norm=1.0)
        fn = GraphedTrainStep(step) if graphed else step
        losses = []
        for i in range(6):
            model.inject_noise = noises[i]
            b = {k: (v.cuda() if torch.is_tensor(v) else list(v)) for k, v in batches[i].items()}
            losses.append(float(fn(b, i)))
            model.global_step += 1
        torch.cuda.synchronize()
        out.append((losses, [p.detach().clone() for p in model.parameters()],
                    [b.detach().clone() for b in model.buffers()], fn))
    assert out[0][0] == out[1][0], (out[0][0], out[1][0])
    for a, b in zip(out[0][1], out[1][1]):
        assert torch.equal(a, b)
    for a, b in zip(out[0][2], out[1][2]):
        assert torch.equal(a, b)
    g = out[1][3]
    assert g.launches_per_replay > 100 and len(g.graphs) == (2 if start == 74998 else 1)


def test_graphed_train_step_draws_the_reference_noise_stream():
    """Without injected noise the graph path consumes the CPU generator exactly like the eager path (eps, z noise, bbox eps per
    step), so both runs see the same draws: same loss for every step at p = 1 (no device-side dropout draw)."""
    from odvae_b200.graphs import GraphedTrainStep
    from odvae_b200.trainer import TrainStep
    B = 2
    batch = synth.synthetic_batch(B, 64, seed=7)
    res = []
    for graphed in (False, True):
        model = _model(40000)
        step = TrainStep(model, lr=1e-4, max_grad_norm=1.0)
        fn = GraphedTrainStep(step) if graphed else step
        torch.manual_seed(1234)
        losses = []
        for i in range(4):
            b = {k: (v.cuda() if torch.is_tensor(v) else list(v)) for k, v in batch.items()}
            losses.append(float(fn(b, i)))
            model.global_step += 1
        res.append(losses)
    assert res[0] == res[1]


def test_fused_adam_starts_a_parameter_at_its_first_gradient_like_torch():
    """torch.optim.Adam skips parameters whose grad is None and starts their bias correction at their own first
    step. Half of the parameters get their first gradient at step 3."""
    from odvae_b200.optim import FusedAdam
    torch.manual_seed(0)
    shapes = [(64, 64, 3, 3), (1000,), (128, 64, 1, 1), (7,)]
    ps = [torch.randn(s, device="cuda") for s in shapes]
    a = [p.clone().requires_grad_(True) for p in ps]
    b = [p.clone().requires_grad_(True) for p in ps]
    oa = FusedAdam(a, lr=1e-2, betas=(0.5, 0.9))
    ob = torch.optim.Adam(b, lr=1e-2, betas=(0.5, 0.9))
    for it in range(6):
        for i, (x, y) in enumerate(zip(a, b)):
            if i % 2 == 1 and it < 2:
                x.grad = y.grad = None
            else:
                g = torch.randn_like(x)
                x.grad, y.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
    torch.cuda.synchronize()
    for x, y in zip(a, b):
        assert H.rel_max(x, y) < 1e-5
    assert float(oa.state[a[0]]["step"]) == 6 and float(oa.state[a[1]]["step"]) == 4


def test_arena_parameters_without_a_gradient_are_neither_updated_nor_aged():
    """Encoder-pretrain phase through TrainStep (gradient arena): decoder / post_quant_conv / pose_encoder are not on the path
    (AE.py:246-247) -> their values stay put and their step counters stay 0, while encoder parameters move."""
    from odvae_b200.trainer import TrainStep
    model = _model(100)
    before = {k: p.detach().clone() for k, p in model.named_parameters()}
    step = TrainStep(model, lr=1e-3, max_grad_norm=1.0)
    for i in range(2):
        b = {k: (v.cuda() if torch.is_tensor(v) else list(v)) for k, v in synth.synthetic_batch(2, 64, seed=i).items()}
        step(b, i)
        model.global_step += 1
    torch.cuda.synchronize()
    named = dict(model.named_parameters())
    assert torch.equal(named["decoder.conv_in.weight"], before["decoder.conv_in.weight"])
    assert torch.equal(named["post_quant_conv.weight"], before["post_quant_conv.weight"])
    assert not torch.equal(named["encoder.conv_in.weight"], before["encoder.conv_in.weight"])
    assert float(step.opt.state[named["decoder.conv_in.weight"]]["step"]) == 0
    assert float(step.opt.state[named["encoder.conv_in.weight"]]["step"]) == 2
