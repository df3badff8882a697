"""Shared helpers for the parity tests: golden loading, seeded inputs, comparison metrics."""
import os
import pickle
import tempfile

import torch

from odvae_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
CASES = ["small_s80000_b3", "small_s100_b2", "small_s35000_b2", "small_s130000_b4", "small_s80000_allbg",
         "full_s80000_b2"]
# disc_factor = 1.0: generator term with the adaptive d_weight in the optimizer_idx 0 step + the optimizer_idx 1 step
GAN_CASES = ["small_s80000_gan_b3", "small_s30000_gan_b2", "full_s80000_gan_b2"]

_STATS = {}


def stats_path(seed=23):
    if seed not in _STATS:
        path = os.path.join(tempfile.mkdtemp(prefix="odvae_stats_"), "all.pkl")
        with open(path, "wb") as f:
            pickle.dump(synth.synthetic_dataset_stats(seed), f)
        _STATS[seed] = path
    return _STATS[seed]


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)


def case_inputs(g):
    """-> (model params, batch dict, noise tuple, global_step) of a golden case, rebuilt from its seeds."""
    spec = g["spec"]
    ddconfig = synth.DDCONFIG if spec["ddconfig"] == "full" else synth.DDCONFIG_SMALL
    params = synth.model_config(ddconfig, stats_path(spec["stats_seed"]))
    params["lossconfig"]["params"].update(spec.get("loss_override") or {})
    batch = synth.synthetic_batch(spec["batch"], spec["resolution"], spec["batch_seed"])
    if spec["class_override"] is not None:
        batch["class_id"] = torch.full_like(batch["class_id"], spec["class_override"])
        batch["class_name"] = [synth.LABELS[spec["class_override"]]] * spec["batch"]
    n = g["noise"]
    noise = (n["eps_obj"], n["keep"], n["noise"], n["eps_bbox"])
    return params, batch, noise, spec["global_step"]


def clone_batch(batch):
    return {k: (v.clone() if torch.is_tensor(v) else list(v)) for k, v in batch.items()}


def rel_l2(a, ref):
    a, ref = a.detach().double().cpu(), ref.detach().double().cpu()
    return float((a - ref).norm() / ref.norm().clamp_min(1e-30))


def rel_max(a, ref):
    a, ref = a.detach().double().cpu(), ref.detach().double().cpu()
    return float((a - ref).abs().max() / ref.abs().max().clamp_min(1e-30))


def rel_scalar(a, ref, floor=1e-12):
    return abs(float(a) - float(ref)) / max(abs(float(ref)), floor)


def grad_errors(named_grads, golden_grads, zero_grad_slack=1e-2):
    """Per-tensor errors against the golden gradient summaries (l2 norm + 32 sampled elements).
    -> {name: (|l2 - ref|/ref, max|sample - ref| / max|ref sample|)}; tensors with ~zero reference norm are
    compared absolutely against the largest gradient norm in the model."""
    out = {}
    scale = max(v["l2"] for v in golden_grads.values())
    for k, gs in golden_grads.items():
        g = named_grads[k]
        if g is None:
            # the reference reaches these tensors only through `0 * g_loss` (GAN term switched off): exact zeros
            assert gs["l2"] == 0.0, "no gradient for %s" % k
            out[k] = (0.0, 0.0)
            continue
        assert tuple(g.shape) == tuple(gs["shape"]), k
        flat = g.detach().double().cpu().reshape(-1)
        ref_l2 = gs["l2"]
        if k.endswith(".k.bias"):
            # analytically zero: a constant added to every key shifts each softmax row uniformly. The reference's
            # value is rounding noise; require ours to be noise as well, measured against the sibling q.bias gradient
            sib = golden_grads[k[:-len("k.bias")] + "q.bias"]["l2"]
            out[k] = (float(flat.norm()) / max(sib, 1e-30) * zero_grad_slack, 0.0)
            continue
        if ref_l2 < 1e-12 * scale:
            out[k] = (float(flat.norm()) / scale, 0.0)
            continue
        e_l2 = abs(float(flat.norm()) - ref_l2) / ref_l2
        ref_v = gs["val"].double()
        e_s = float((flat[gs["idx"]] - ref_v).abs().max() / ref_v.abs().max().clamp_min(1e-30 + 1e-6 * ref_l2 / flat.numel() ** 0.5))
        out[k] = (e_l2, e_s)
    return out


GRAD_SUBSET_CAP = 16384  # tests/golden/make_golden.py FULL_GRAD_CAP


def subset_indices(numel, cap=GRAD_SUBSET_CAP):
    """The RNG-free element subset make_golden.py stores for tensors above the cap (same formula)."""
    import math
    if numel <= cap:
        return None
    a = next(q for q in (1000003, 1000033, 1000037, 1000039) if math.gcd(q, numel) == 1)
    return (torch.arange(cap, dtype=torch.int64) * a + 12345) % numel


def grad_diff_errors(named_grads, golden_grads, noise_floor=1e-7):
    """Per-tensor errors of the DIFFERENCE against the stored reference gradients (per-tensor relative L2 +
    max-abs / ||ref||_inf): {name: (||g - ref||_2 / ||ref||_2, max|g - ref| / max|ref|)} over the stored elements (the whole
    tensor up to 16384 elements, a fixed subset of 16384 above). Tensors whose reference gradient is rounding noise
    (norm below `noise_floor` x the largest gradient norm of the model, e.g. attention k.bias, analytically zero) are
    measured against that largest norm instead of their own."""
    out = {}
    scale = max(v["l2"] for v in golden_grads.values())
    for k, gs in golden_grads.items():
        g = named_grads[k]
        ref = gs["ref"].double()
        if g is None:
            assert gs["l2"] == 0.0, "no gradient for %s" % k
            out[k] = (0.0, 0.0)
            continue
        assert tuple(g.shape) == tuple(gs["shape"]), k
        flat = g.detach().double().cpu().reshape(-1)
        sub = subset_indices(flat.numel())
        if sub is not None:
            flat = flat[sub]
        assert flat.numel() == ref.numel(), k
        diff = flat - ref
        if k.endswith(".k.bias"):
            sib = golden_grads[k[:-len("k.bias")] + "q.bias"]["l2"]
            out[k] = (float(diff.norm()) / max(sib, 1e-30), 0.0)
            continue
        if gs["l2"] < noise_floor * scale:
            out[k] = (float(diff.norm()) / scale, 0.0)
            continue
        out[k] = (float(diff.norm() / ref.norm().clamp_min(1e-300)),
                  float(diff.abs().max() / ref.abs().max().clamp_min(1e-300)))
    return out


def error_table(errs, top=8):
    """Text summary of grad_diff_errors(): quantiles of the per-tensor relative L2 and the worst tensors."""
    l2 = sorted(v[0] for v in errs.values())
    mx = sorted(v[1] for v in errs.values())
    q = lambda a, f: a[min(len(a) - 1, int(f * len(a)))]
    lines = ["tensors %d | rel-L2 of the difference: median %.3g  p90 %.3g  p99 %.3g  max %.3g | max-abs/inf-norm: median %.3g  max %.3g"
             % (len(l2), q(l2, 0.5), q(l2, 0.9), q(l2, 0.99), l2[-1], q(mx, 0.5), mx[-1])]
    for k, v in sorted(errs.items(), key=lambda kv: -kv[1][0])[:top]:
        lines.append("    %-58s rel-L2 %.3g  max-abs %.3g" % (k, v[0], v[1]))
    return "\n".join(lines)
