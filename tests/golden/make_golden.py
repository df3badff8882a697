"""Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference
(`/root/reference`, via oracle/ref_shim.py) on CPU fp32 over seeded synthetic inputs.

Run in the build container only (the reference tree is not on the GPU box):

    python tests/golden/make_golden.py            # all cases
    python tests/golden/make_golden.py small_s80000_b3

Each case file holds: the case spec (seeds, geometry, global_step), the four noise draws the
reference consumed, the forward outputs, the loss + 26-entry log dict, and for every trainable tensor the
reference gradient itself (whole when it has <= 16384 elements, else a fixed RNG-free subset of 16384 elements)
plus summaries (l2 norm, sum, 32 sampled elements). Inputs and weights are
NOT stored: they are pure functions of the seeds (odvae_b200/synth.py).
"""
import math
import os
import pickle
import sys
import tempfile
import time
import zlib

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "generative-detection_b200"))
sys.path.insert(0, ROOT)

from odvae_b200 import synth  # noqa: E402
from oracle import ref_shim  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (ddconfig name, resolution, batch, global_step, batch seed, noise seed, class override)
    "full_s80000_b2": ("full", 256, 2, 80000, 23, 101, None),
    "small_s80000_b3": ("small", 64, 3, 80000, 24, 102, None),
    "small_s100_b2": ("small", 64, 2, 100, 25, 103, None),
    "small_s35000_b2": ("small", 64, 2, 35000, 26, 104, None),
    "small_s130000_b4": ("small", 64, 4, 130000, 27, 105, None),
    "small_s80000_allbg": ("small", 64, 2, 80000, 28, 106, 1),
    # GAN on (disc_factor = 1.0, the yaml's default; perceptual_weight stays 0: LPIPS weights are not available offline):
    # adaptive d_weight + generator term in the optimizer_idx 0 step, and the optimizer_idx 1 (discriminator) step
    "small_s80000_gan_b3": ("small", 64, 3, 80000, 29, 107, None, {"disc_factor": 1.0}),
    "small_s30000_gan_b2": ("small", 64, 2, 30000, 30, 108, None, {"disc_factor": 1.0}),
    "full_s80000_gan_b2": ("full", 256, 2, 80000, 31, 109, None, {"disc_factor": 1.0}),
}


def sample_indices(key, numel, n=32):
    g = torch.Generator().manual_seed(zlib.crc32(key.encode()) % (2 ** 31 - 1))
    return torch.randint(0, numel, (min(n, numel),), generator=g)


FULL_GRAD_CAP = 16384  # tensors up to this many elements are stored whole; larger ones as a fixed subset of this size


def subset_indices(numel, cap=FULL_GRAD_CAP):
    """Deterministic, RNG-free element subset of a large tensor: idx_i = (a * i + b) mod numel with a prime multiplier
    coprime to numel (distinct indices, no alignment with the tensor's tap / channel strides)."""
    if numel <= cap:
        return None
    a = next(q for q in (1000003, 1000033, 1000037, 1000039) if math.gcd(q, numel) == 1)
    return (torch.arange(cap, dtype=torch.int64) * a + 12345) % numel


def grad_summary(key, g):
    flat = g.detach().reshape(-1).double()
    idx = sample_indices(key, flat.numel())
    sub = subset_indices(flat.numel())
    ref = g.detach().reshape(-1).float()
    return {"l2": float(flat.norm()), "sum": float(flat.sum()), "abs_sum": float(flat.abs().sum()),
            "absmax": float(flat.abs().max()),
            "idx": idx, "val": g.detach().reshape(-1)[idx].clone(), "shape": tuple(g.shape),
            # the reference gradient itself (whole tensor, or the subset_indices() elements of a large one), so the
            # parity tests can bound ||g - g_ref|| / ||g_ref|| per tensor instead of comparing norms
            "ref": (ref if sub is None else ref[sub]).clone(), "ref_is_subset": sub is not None}


def tensor_summary(t):
    t = t.detach().float()
    return {"mean": float(t.mean()), "std": float(t.std()), "absmax": float(t.abs().max()),
            "shape": tuple(t.shape)}


def run_case(name):
    cfg_name, R, B, step, bseed, nseed, class_override = CASES[name][:7]
    loss_override = CASES[name][7] if len(CASES[name]) > 7 else {}
    ddconfig = synth.DDCONFIG if cfg_name == "full" else synth.DDCONFIG_SMALL
    tmp = tempfile.mkdtemp()
    stats_path = os.path.join(tmp, "all.pkl")
    with open(stats_path, "wb") as f:
        pickle.dump(synth.synthetic_dataset_stats(23), f)
    params = synth.model_config(ddconfig, stats_path)
    params["lossconfig"]["params"].update(loss_override)
    torch.manual_seed(23)
    model = ref_shim.build_reference_model(params, step)
    synth.fill_state_dict_(model.state_dict(), seed=23)
    model.train()

    batch = synth.synthetic_batch(B, R, bseed)
    if class_override is not None:
        batch["class_id"] = torch.full_like(batch["class_id"], class_override)
        batch["class_name"] = [synth.LABELS[class_override]] * B

    cap = {}
    acts = {}

    def fwd_hook(mod, inp, out):
        cap["forward_out"] = out

    def loss_hook(mod, inp, out):
        cap["loss_out"] = out
        cap["loss_in_rgb"] = inp[0]

    def pq_hook(mod, inp):
        cap["z_obj_pose"] = inp[0].detach().clone()

    def pe_hook(mod, inp, out):
        cap["enc_pose"] = out.detach().clone()

    def qp_hook(mod, inp, out):
        cap["pose_feat"] = out.detach().clone()

    orig_forward = model.forward

    def forward_capture(*a, **k):  # training_step calls self.forward(...) directly (no module hook)
        out = orig_forward(*a, **k)
        cap["forward_out"] = out
        return out

    model.forward = forward_capture
    hs = [model.loss.register_forward_hook(loss_hook),
          model.post_quant_conv.register_forward_pre_hook(pq_hook),
          model.pose_encoder.register_forward_hook(pe_hook),
          model.quant_conv_pose.register_forward_hook(qp_hook)]
    for mname, mod in model.named_modules():
        if mname.startswith(("encoder", "decoder")) and mname.count(".") <= 3 and mname.count(".") >= 1:
            def mk(n):
                def h(m, i, o):
                    if torch.is_tensor(o):
                        acts[n] = tensor_summary(o)
                return h
            hs.append(mod.register_forward_hook(mk(mname)))

    p_drop = None
    torch.manual_seed(nseed)
    t0 = time.time()
    b2 = {k: (v.clone() if torch.is_tensor(v) else list(v)) for k, v in batch.items()}
    loss = model.training_step(b2, 0, 0)
    loss.backward()
    dt = time.time() - t0
    p_drop = float(model.dropout_prob)
    for h in hs:
        h.remove()

    dec_obj, dec_pose, posterior_obj, bbox_posterior = cap["forward_out"]
    loss_val, log = cap["loss_out"]

    # replicate the reference's RNG consumption and check it (so products can inject the same draws)
    eps_obj, keep, noise, eps_bbox = synth.draw_noise(B, p_drop, seed=nseed)
    if step >= params["lossconfig"]["params"]["encoder_pretrain_steps"]:
        z_ref = cap["z_obj_pose"] - cap["enc_pose"].view(B, 16, 16, 16)
        z_mine = posterior_obj.mean + posterior_obj.std * eps_obj
        if p_drop > 0:
            z_mine = z_mine * keep / (1.0 - p_drop) if p_drop < 1.0 else z_mine * 0.0
        z_mine = z_mine + noise
        err = (z_ref - z_mine.detach()).abs().max().item()
        assert err < 1e-4, "noise replication mismatch: %g" % err
    bbox_pred = bbox_posterior.mean + bbox_posterior.std * eps_bbox
    err = (dec_pose[:, :8].detach() - bbox_pred.detach()).abs().max().item()
    assert err < 1e-5, "bbox eps replication mismatch: %g" % err

    grads = {}
    nograd = []
    for k, p in model.named_parameters():
        if k.startswith(("loss.discriminator", "loss.perceptual_loss")):
            continue
        if p.grad is None:
            nograd.append(k)
        else:
            grads[k] = grad_summary(k, p.grad)

    disc = None
    if loss_override.get("disc_factor", 0.0) > 0:
        # optimizer_idx 1 on the same batch with the same four draws (AE.py:321-330): d_loss, its log, the discriminator's
        # parameter gradients and the BatchNorm running statistics after the three discriminator forwards of this batch
        model.zero_grad(set_to_none=True)
        torch.manual_seed(nseed)
        b3 = {k: (v.clone() if torch.is_tensor(v) else list(v)) for k, v in batch.items()}
        hd = model.loss.register_forward_hook(loss_hook)
        d_loss = model.training_step(b3, 0, 1)
        hd.remove()
        d_loss.backward()
        d_log = cap["loss_out"][1]
        dgrads = {k: grad_summary(k, p.grad) for k, p in model.named_parameters()
                  if k.startswith("loss.discriminator") and p.grad is not None}
        disc = {"loss": float(d_loss.detach()),
                "log": {k: float(torch.as_tensor(v).detach().float().mean()) for k, v in d_log.items()},
                "grads": dgrads,
                "buffers": {k: v.detach().clone() for k, v in model.state_dict().items()
                            if k.startswith("loss.discriminator") and ("running" in k or "num_batches" in k)}}

    out = {
        "spec": {"name": name, "ddconfig": cfg_name, "resolution": R, "batch": B, "global_step": step,
                 "batch_seed": bseed, "noise_seed": nseed, "weight_seed": 23, "stats_seed": 23,
                 "class_override": class_override, "dropout_p": p_drop, "loss_override": loss_override,
                 "torch": torch.__version__, "reference_commit": "b09216ca0249d767eab18ebc8afb8bf69ded44e3",
                 "cpu_seconds_fwd_bwd": dt},
        "noise": {"eps_obj": eps_obj, "keep": keep.to(torch.uint8), "noise": noise, "eps_bbox": eps_bbox},
        "rgb_gt_summary": tensor_summary(cap["loss_in_rgb"]),
        "out": {"dec_obj": dec_obj.detach().clone(), "dec_pose": dec_pose.detach().clone(),
                "posterior_mean": posterior_obj.mean.detach().clone(),
                "posterior_logvar": posterior_obj.logvar.detach().clone(),
                "bbox_mean": bbox_posterior.mean.detach().clone(),
                "bbox_logvar": bbox_posterior.logvar.detach().clone(),
                "pose_feat": cap["pose_feat"]},
        "loss": float(loss_val.detach()),
        "log": {k: float(torch.as_tensor(v).detach().float().mean()) for k, v in log.items()},
        "grads": grads,
        "params_without_grad": nograd,
        "acts": acts,
        "disc": disc,
    }
    path = os.path.join(HERE, name + ".pt")
    torch.save(out, path)
    print("%s: loss=%.6g  %d grads, %d params without grad, %.1fs, %.2f MB" % (
        name, out["loss"], len(grads), len(nograd), dt, os.path.getsize(path) / 1e6))


if __name__ == "__main__":
    names = sys.argv[1:] or list(CASES)
    for n in names:
        run_case(n)
