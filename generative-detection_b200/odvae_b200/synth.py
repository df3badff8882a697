"""Synthetic inputs for the OD-VAE hot path: weights, object-crop batches, dataset stats.

Everything here is a pure function of integer seeds so the golden generator (which runs the
real reference in the build container), the oracle, the GPU parity tests and bench.py all see
bit-identical inputs without shipping large tensors.

Shapes follow the reference's batch dict (src/data/datasets/nuscenes.py:492-520,543-560) and the
dataset-stats pickle schema (compute_dataset_stats.py:18-25,134-135).
"""
import math
import zlib

import torch

LABELS = ['car', 'truck', 'trailer', 'bus', 'construction_vehicle', 'bicycle', 'motorcycle',
          'pedestrian', 'traffic_cone', 'barrier', 'background']

# configs/autoencoder/pose/autoencoder_kl_16x16x16.yaml:54-66
DDCONFIG = dict(double_z=True, z_channels=16, resolution=64, in_channels=3, out_ch=3, ch=128,
                ch_mult=[1, 1, 2, 2, 4], num_res_blocks=2, attn_resolutions=[16], dropout=0.0)

# a reduced geometry with the same structure (attention at level 1, 2 down/upsamples, mid block):
# used by fast unit tests; channel counts stay multiples of 64 so the tensor-core path is exercised.
DDCONFIG_SMALL = dict(double_z=True, z_channels=16, resolution=32, in_channels=3, out_ch=3, ch=64,
                      ch_mult=[1, 2, 2], num_res_blocks=1, attn_resolutions=[16], dropout=0.0)

# configs/autoencoder/pose/autoencoder_kl_16x16x16.yaml:17-32 (+ perceptual/disc off: SURVEY §8d config 2)
LOSS_PARAMS = dict(encoder_pretrain_steps=30000, disc_start=30000, kl_weight_obj=1.0,
                   kl_weight_bbox=0.000001, disc_weight=0.5, pose_weight=80000,
                   fill_factor_weight=500000, class_weight=1000000, bbox_weight=200000,
                   pose_loss_fn="l1", mask_weight=0, mask_loss_fn="l2", disc_in_channels=3,
                   num_classes=11, dataset_stats_path="dataset_stats/combined/all.pkl",
                   perceptual_weight=0.0, disc_factor=0.0)

POSE_DEC_PARAMS = dict(num_classes=11, num_channels=16, n=16, m=16, hidden_dim=500, num_layers=2,
                       activation="tanh", resid=False)
POSE_ENC_PARAMS = dict(num_classes=11, num_channels=16, n=16, m=16, hidden_dim=500, num_layers=2,
                       activation="swish")

MODEL_PARAMS = dict(monitor="val/rec_loss", embed_dim=16, euler_convention="XYZ", activation="relu",
                    dropout_prob_init=1.0, dropout_prob_final=0.7,
                    pose_conditioned_generation_steps=45000, dropout_warmup_steps=45000,
                    add_noise_to_z_obj=True, train_on_yaw=True)


def model_config(ddconfig=None, dataset_stats_path=None):
    """The `model.params` subtree of the reference yaml, as plain dicts."""
    loss_params = dict(LOSS_PARAMS)
    if dataset_stats_path is not None:
        loss_params["dataset_stats_path"] = dataset_stats_path
    params = dict(MODEL_PARAMS)
    params["ddconfig"] = dict(ddconfig or DDCONFIG)
    params["lossconfig"] = dict(target="src.modules.losses.PoseLoss", params=loss_params)
    params["pose_decoder_config"] = dict(
        target="src.modules.autoencodermodules.pose_decoder.PoseDecoderSpatialVAE",
        params=dict(POSE_DEC_PARAMS))
    params["pose_encoder_config"] = dict(
        target="src.modules.autoencodermodules.pose_encoder.PoseEncoderSpatialVAE",
        params=dict(POSE_ENC_PARAMS))
    return params


def _key_seed(key, seed):
    return (zlib.crc32(key.encode("utf-8")) + 7919 * seed) % (2 ** 31 - 1)


def fill_state_dict_(sd, seed=23):
    """Overwrite every floating tensor of `sd` in place with values that depend only on
    (key, shape, seed). Scales mimic torch's default init (bound 1/sqrt(fan_in)); GroupNorm
    affine parameters are perturbed away from (1, 0) so both are exercised."""
    for key in sorted(sd.keys()):
        t = sd[key]
        if not torch.is_floating_point(t):
            continue
        g = torch.Generator().manual_seed(_key_seed(key, seed))
        if t.ndim == 0:
            v = torch.full((), 0.1)
        elif t.ndim == 1:
            bn = key.startswith("loss.discriminator.") and t.numel() > 1 and not key.endswith(".bias")
            if key.endswith("running_var"):
                v = 1.0 + 0.5 * torch.rand(t.shape, generator=g)
            elif ("norm" in key and key.endswith("weight")) or (bn and key.endswith("weight")):
                v = 1.0 + 0.1 * torch.randn(t.shape, generator=g)
            else:
                v = 0.05 * torch.randn(t.shape, generator=g)
        else:
            fan_in = 1
            for d in t.shape[1:]:
                fan_in *= d
            bound = 1.0 / math.sqrt(fan_in)
            v = (torch.rand(t.shape, generator=g) * 2.0 - 1.0) * bound
        with torch.no_grad():
            t.copy_(v.to(t.dtype))
    return sd


def synthetic_batch(batch_size, resolution=256, seed=23):
    """Object-crop batch dict with the reference's keys."""
    g = torch.Generator().manual_seed(seed)
    B, R = batch_size, resolution
    patch = torch.rand(B, 3, R, R, generator=g)
    class_id = torch.randint(0, len(LABELS), (B,), generator=g)
    class_name = [LABELS[int(c)] for c in class_id]
    mask = torch.zeros(B, 1, R, R)
    frac = 0.3 + 0.7 * torch.rand(B, 2, generator=g)
    off = torch.rand(B, 2, generator=g)
    for b in range(B):
        if class_name[b] == "background":
            continue
        h = max(1, int(round(float(frac[b, 0]) ** 0.5 * R)))
        w = max(1, int(round(float(frac[b, 1]) ** 0.5 * R)))
        y0 = int(float(off[b, 0]) * (R - h))
        x0 = int(float(off[b, 1]) * (R - w))
        mask[b, :, y0:y0 + h, x0:x0 + w] = 1.0
    pose_6d = torch.randn(B, 4, generator=g)
    yaw = (torch.rand(B, generator=g, dtype=torch.float64) * 2.0 - 1.0) * math.pi
    bbox_sizes = 0.3 + 2.7 * torch.rand(B, 3, generator=g)
    fill_factor = 0.5 * torch.rand(B, generator=g, dtype=torch.float64)
    return {"patch": patch, "class_id": class_id, "class_name": class_name, "mask_2d_bbox": mask,
            "pose_6d": pose_6d, "yaw": yaw, "bbox_sizes": bbox_sizes, "fill_factor": fill_factor}


def synthetic_dataset_stats(seed=23):
    """{label: {key: tensor([mean, logvar])}} for the 10 foreground labels (the pickle schema the
    reference loads at src/modules/losses/contperceptual.py:75-80)."""
    g = torch.Generator().manual_seed(seed + 1000)
    stats = {}
    for label in LABELS[:-1]:
        d = {}
        for key in ["t1", "t2", "t3", "v3", "yaw", "l", "h", "w", "fill_factor"]:
            mean = float(torch.rand((), generator=g) * 2.0 - 1.0)
            logvar = float(-2.0 * torch.rand((), generator=g))
            d[key] = torch.tensor([mean, logvar])
        stats[label] = d
    return stats


def draw_noise(batch_size, dropout_p, latent_shape=(16, 16, 16), bbox_dim=8, seed=None):
    """The four random draws of one reference forward (src/models/autoencoder.py:226-241,153-154;
    distributions.py:35-37), taken from the CPU generator in the reference's order:
    eps_obj -> dropout keep-mask -> z noise -> eps_bbox. Returns CPU tensors; `keep` holds
    {0,1} (the reference scales by 1/(1-p) afterwards)."""
    if seed is not None:
        torch.manual_seed(seed)
    shape = (batch_size,) + tuple(latent_shape)
    eps_obj = torch.randn(shape)
    if dropout_p > 0:
        if dropout_p >= 1.0:
            keep = torch.zeros(shape)
        else:
            keep = torch.empty(shape).bernoulli_(1.0 - dropout_p)
    else:
        keep = torch.ones(shape)
    noise = torch.normal(torch.zeros(shape), torch.ones(shape))
    eps_bbox = torch.randn(batch_size, bbox_dim)
    return eps_obj, keep, noise, eps_bbox
