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
