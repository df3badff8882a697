"""The non-Lightning training loop (odvae_b200/fit.py) on the B200: global_step accounting of the
two-optimizer loop, LR scaling (train.py:383), phase switch at encoder_pretrain_steps (AE.py:246-255), checkpoint in
the PL key layout, resume, and init_from_ckpt (AE.py:97) of a freshly built model from that checkpoint."""
import os

import pytest
import torch

import helpers as H
from odvae_b200 import synth

pytestmark = pytest.mark.gpu


def _model(step, dtype=torch.float32):
    from odvae_b200.autoencoder import PoseAutoencoder
    params = synth.model_config(synth.DDCONFIG_SMALL, H.stats_path(23))
    model = PoseAutoencoder(**params, compute_dtype=dtype)
    synth.fill_state_dict_(model.state_dict(), seed=23)
    model = model.cuda()
    model.global_step = step
    return model, params


def _loader(n, B=2, res=64):
    return [synth.synthetic_batch(B, res, 100 + i) for i in range(n)]


def test_fit_two_optimizer_loop_checkpoint_and_resume(tmp_path):
    from odvae_b200.autoencoder import PoseAutoencoder
    from odvae_b200.fit import Fit
    model, params = _model(79996)
    fit = Fit(model, batch_size=2, ckpt_dir=str(tmp_path))
    assert abs(model.learning_rate - 2 * 4.5e-6) < 1e-18  # accumulate(1) * ngpu(1) * bs(2) * base_lr
    w0 = model.decoder.conv_out.weight.detach().clone()
    seen = []
    fit.log_fn = lambda step, ae, disc: seen.append((step, ae, disc))
    fit.fit(_loader(3), max_steps=80000)  # 2 batches = 4 optimizer steps; the third batch is not started
    assert model.global_step == 80000 and [s for s, _, _ in seen] == [79998, 80000]
    assert all(torch.isfinite(torch.tensor(a)) for _, a, _ in seen) and all(d == 0.0 for _, _, d in seen)
    assert not torch.equal(w0, model.decoder.conv_out.weight)
    path = os.path.join(str(tmp_path), "last.ckpt")
    fit.save_checkpoint(path)
    ckpt = torch.load(path, map_location="cpu", weights_only=False)
    assert {"epoch", "global_step", "pytorch-lightning_version", "state_dict", "optimizer_states"} <= set(ckpt)
    assert ckpt["global_step"] == 80000 and len(ckpt["optimizer_states"]) == 2
    assert set(ckpt["state_dict"]) == set(model.state_dict())

    # init_from_ckpt on a fresh model (constructor path of AE.py:97): every tensor that exists at that point is restored
    fresh = PoseAutoencoder(**params, compute_dtype=torch.float32, ckpt_path=path).cuda()
    for k, v in model.state_dict().items():
        if not k.startswith(("pose_decoder.", "pose_encoder.")):  # created after init_from_ckpt in the reference too
            assert torch.equal(fresh.state_dict()[k].cpu(), v.cpu()), k

    # resume: same weights, same Adam moments, same next step as continuing the first run
    model2, _ = _model(0)
    fit2 = Fit(model2, batch_size=2)
    fit2.load_checkpoint(path)
    assert model2.global_step == 80000
    nxt = _loader(4)[3]
    noise = synth.draw_noise(2, model._get_dropout_prob(), seed=7)
    model.inject_noise = noise
    model2.inject_noise = tuple(t.clone() if torch.is_tensor(t) else t for t in noise)
    a1, _ = fit.training_batch(H.clone_batch(nxt), 0)
    a2, _ = fit2.training_batch(H.clone_batch(nxt), 0)
    torch.cuda.synchronize()
    assert float(a1) == float(a2)
    for (k, p1), (_, p2) in zip(model.named_parameters(), model2.named_parameters()):
        assert torch.equal(p1, p2), k


def test_fit_encoder_pretrain_phase_leaves_decoder_untouched():
    """global_step < encoder_pretrain_steps: the decoder is skipped (AE.py:246-247), so its parameters get no gradient
    and Adam must not move them; the encoder and the pose decoder train."""
    from odvae_b200.fit import Fit
    model, _ = _model(100)
    fit = Fit(model, batch_size=2)
    dec0 = {k: v.detach().clone() for k, v in model.decoder.state_dict().items()}
    enc0 = model.encoder.conv_in.weight.detach().clone()
    fit.fit(_loader(2), max_steps=104)
    assert model.global_step == 104
    for k, v in model.decoder.state_dict().items():
        assert torch.equal(v, dec0[k]), k
    assert not torch.equal(enc0, model.encoder.conv_in.weight)


def test_fit_with_the_gan_on_trains_both_networks_and_resumes_bit_exactly(tmp_path):
    """The yaml's own objective through the Lightning-free loop (SURVEY 8 f2 + f3): per batch the optimizer_idx 0 step
    (LPIPS off here: no weights offline; PatchGAN generator term with the adaptive weight) and the optimizer_idx 1 step
    (hinge loss on detached real / fake crops, clip, Adam on the discriminator). Both networks move, the discriminator's
    BatchNorm running statistics advance three forwards per batch, the checkpoint carries both optimizers, and a resumed run
    reproduces the next batch bit for bit (same injected noise)."""
    from odvae_b200.autoencoder import PoseAutoencoder
    from odvae_b200.fit import Fit
    params = synth.model_config(synth.DDCONFIG_SMALL, H.stats_path(23))
    params["lossconfig"]["params"]["disc_factor"] = 1.0

    def build(step):
        m = PoseAutoencoder(**params, compute_dtype=torch.float16)
        synth.fill_state_dict_(m.state_dict(), seed=23)
        m = m.cuda()
        m.global_step = step
        return m

    from odvae_b200 import ops
    ops.read_overflow(torch.device("cuda", 0))  # sticky per-device flags: drop what earlier tests left behind
    model = build(79996)
    fit = Fit(model, batch_size=2, ckpt_dir=str(tmp_path))
    d0 = {k: v.detach().clone() for k, v in model.loss.discriminator.state_dict().items()}
    g0 = model.decoder.conv_out.weight.detach().clone()
    seen = []
    fit.log_fn = lambda step, ae, disc: seen.append((step, ae, disc))
    fit.fit(_loader(2), max_steps=80000)
    assert model.global_step == 80000 and len(seen) == 2
    assert all(torch.isfinite(torch.tensor([a, d])).all() for _, a, d in seen) and all(d > 0.0 for _, _, d in seen)
    assert float(model.logged["train/d_weight"]) > 0.0 and float(model.logged["train/disc_factor"]) == 1.0
    d1 = model.loss.discriminator.state_dict()
    assert not torch.equal(d0["main.0.weight"], d1["main.0.weight"])           # discriminator trained
    assert not torch.equal(d0["main.3.running_mean"], d1["main.3.running_mean"])
    assert int(d1["main.3.num_batches_tracked"]) - int(d0["main.3.num_batches_tracked"]) == 6  # 3 passes per batch
    assert not torch.equal(g0, model.decoder.conv_out.weight)                   # generator trained
    model.check_overflow()
    path = os.path.join(str(tmp_path), "gan.ckpt")
    fit.save_checkpoint(path)

    model2 = build(0)
    fit2 = Fit(model2, batch_size=2)
    fit2.load_checkpoint(path)
    nxt = _loader(3)[2]
    noise = synth.draw_noise(2, model._get_dropout_prob(), seed=9)
    model.inject_noise = noise
    model2.inject_noise = tuple(t.clone() for t in noise)
    a1, dl1 = fit.training_batch(H.clone_batch(nxt), 0)
    a2, dl2 = fit2.training_batch(H.clone_batch(nxt), 0)
    torch.cuda.synchronize()
    assert float(a1) == float(a2) and float(dl1) == float(dl2)
    for (k, p1), (_, p2) in zip(model.state_dict().items(), model2.state_dict().items()):
        assert torch.equal(p1, p2), k
