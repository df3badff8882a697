"""TEST INFRASTRUCTURE ONLY — CPU oracle for the OD-VAE AutoencoderKL hot path.

A plain-PyTorch, CPU, fp32 *restatement* of the reference algorithm (functional style over a
state_dict with the reference's parameter names). Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / `--impl reference` legs may import this file; the product package
(generative-detection_b200/) never does.

Parity status: PINNED. tests/test_oracle_cpu.py checks this restatement against the golden
fixtures under tests/golden/*.pt, which were produced by running the UNMODIFIED reference
(/root/reference @ b09216ca, via oracle/ref_shim.py) in the build container
(tests/golden/make_golden.py). Un-vendored third-party arithmetic (taming LPIPS /
NLayerDiscriminator, mmdet FocalLoss) is restated from the public definitions (SURVEY.md
Appendix B) and is therefore "parity unpinned" for those pieces only.

Citations: AE = src/models/autoencoder.py, LOSS = src/modules/losses/contperceptual.py,
MODEL = submodules/latent-diffusion/ldm/modules/diffusionmodules/model.py,
DIST = submodules/latent-diffusion/ldm/modules/distributions/distributions.py,
SDIST = src/util/distributions.py (all relative to the reference root).
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

BACKGROUND_CLASS_IDX = 1  # LOSS:17 (sic: index 1, not the dataset's background id)


# --------------------------------------------------------------------------------------------
# building blocks (MODEL)
# --------------------------------------------------------------------------------------------
def swish(x):  # MODEL:33-35
    return x * torch.sigmoid(x)


def group_norm(sd, pre, x):  # MODEL:38-39
    return F.group_norm(x, 32, sd[pre + ".weight"], sd[pre + ".bias"], eps=1e-6)


def conv(sd, pre, x, stride=1, padding=0):
    return F.conv2d(x, sd[pre + ".weight"], sd[pre + ".bias"], stride=stride, padding=padding)


def resnet_block(sd, pre, x):  # MODEL:121-141 (temb None, dropout 0)
    h = conv(sd, pre + ".conv1", swish(group_norm(sd, pre + ".norm1", x)), padding=1)
    h = conv(sd, pre + ".conv2", swish(group_norm(sd, pre + ".norm2", h)), padding=1)
    if (pre + ".nin_shortcut.weight") in sd:
        x = conv(sd, pre + ".nin_shortcut", x)
    return x + h


def attn_block(sd, pre, x):  # MODEL:178-202
    h = group_norm(sd, pre + ".norm", x)
    q, k, v = conv(sd, pre + ".q", h), conv(sd, pre + ".k", h), conv(sd, pre + ".v", h)
    b, c, hh, ww = q.shape
    q = q.reshape(b, c, hh * ww).permute(0, 2, 1)
    k = k.reshape(b, c, hh * ww)
    w = torch.softmax(torch.bmm(q, k) * (int(c) ** (-0.5)), dim=2)
    v = v.reshape(b, c, hh * ww)
    h = torch.bmm(v, w.permute(0, 2, 1)).reshape(b, c, hh, ww)
    return x + conv(sd, pre + ".proj_out", h)


def _levels_with_attn(dd):
    """Constructor bookkeeping of MODEL:389-411 / :481-524: which levels carry AttnBlocks."""
    res, out = dd["resolution"], []
    n = len(dd["ch_mult"])
    for lvl in range(n):
        out.append(res in dd["attn_resolutions"])
        if lvl != n - 1:
            res //= 2
    return out


def encoder(sd, dd, x, pre="encoder"):  # MODEL:434-459
    n = len(dd["ch_mult"])
    attn = _levels_with_attn(dd)
    h = conv(sd, pre + ".conv_in", x, padding=1)
    for lvl in range(n):
        for i in range(dd["num_res_blocks"]):
            h = resnet_block(sd, "%s.down.%d.block.%d" % (pre, lvl, i), h)
            if attn[lvl]:
                h = attn_block(sd, "%s.down.%d.attn.%d" % (pre, lvl, i), h)
        if lvl != n - 1:  # Downsample MODEL:72-76: pad right/bottom by one, stride-2 conv
            h = conv(sd, "%s.down.%d.downsample.conv" % (pre, lvl), F.pad(h, (0, 1, 0, 1)), stride=2)
    h = resnet_block(sd, pre + ".mid.block_1", h)
    h = attn_block(sd, pre + ".mid.attn_1", h)
    h = resnet_block(sd, pre + ".mid.block_2", h)
    return conv(sd, pre + ".conv_out", swish(group_norm(sd, pre + ".norm_out", h)), padding=1)


def decoder(sd, dd, z, pre="decoder"):  # MODEL:535-568
    n = len(dd["ch_mult"])
    attn = _levels_with_attn(dd)
    h = conv(sd, pre + ".conv_in", z, padding=1)
    h = resnet_block(sd, pre + ".mid.block_1", h)
    h = attn_block(sd, pre + ".mid.attn_1", h)
    h = resnet_block(sd, pre + ".mid.block_2", h)
    for lvl in reversed(range(n)):
        for i in range(dd["num_res_blocks"] + 1):
            h = resnet_block(sd, "%s.up.%d.block.%d" % (pre, lvl, i), h)
            if attn[lvl]:
                h = attn_block(sd, "%s.up.%d.attn.%d" % (pre, lvl, i), h)
        if lvl != 0:  # Upsample MODEL:53-57
            h = F.interpolate(h, scale_factor=2.0, mode="nearest")
            h = conv(sd, "%s.up.%d.upsample.conv" % (pre, lvl), h, padding=1)
    return conv(sd, pre + ".conv_out", swish(group_norm(sd, pre + ".norm_out", h)), padding=1)


def linear(sd, pre, x):
    return F.linear(x, sd[pre + ".weight"], sd.get(pre + ".bias"))


def pose_decoder(sd, x, pre="pose_decoder"):  # pose_decoder.py:82-97 (tanh, 2 hidden layers)
    h = torch.tanh(linear(sd, pre + ".layers.0", x))
    h = torch.tanh(linear(sd, pre + ".layers.2", h))
    return linear(sd, pre + ".layers.4", h)


def pose_encoder(sd, z, n=16, m=16, pre="pose_encoder"):  # pose_encoder.py:100-131
    xg, yg = np.meshgrid(np.linspace(-1, 1, m), np.linspace(1, -1, n))
    coords = torch.from_numpy(np.stack([xg.ravel(), yg.ravel()], 1)).float().reshape(1, -1)
    b = z.shape[0]
    h_x = linear(sd, pre + ".coord_linear", coords.expand(b, -1).to(z))
    h_z = F.linear(z, sd[pre + ".latent_linear.weight"])
    feat = h_z.shape[1]
    h_z = h_z.unsqueeze(1).expand(b, n * m, feat).reshape(b, n * m * feat)
    h = F.silu(h_x + h_z)
    h = F.silu(linear(sd, pre + ".layers.1", h))
    return linear(sd, pre + ".layers.3", h)


# --------------------------------------------------------------------------------------------
# PatchGAN discriminator: taming.modules.discriminator.model.NLayerDiscriminator(input_nc=3, ndf=64, n_layers=3)
# (un-vendored, PARITY UNPINNED; public definition in SURVEY appendix B). BatchNorm2d in training mode: batch statistics,
# running statistics updated in place in `sd` (momentum 0.1, unbiased variance) like torch.nn.BatchNorm2d.
# --------------------------------------------------------------------------------------------
def discriminator(sd, x, pre="loss.discriminator.main", n_layers=3, training=True):
    idx = 0
    h = F.leaky_relu(F.conv2d(x, sd["%s.%d.weight" % (pre, idx)], sd["%s.%d.bias" % (pre, idx)], stride=2, padding=1), 0.2)
    idx = 2
    for n in range(1, n_layers + 1):
        stride = 2 if n < n_layers else 1
        h = F.conv2d(h, sd["%s.%d.weight" % (pre, idx)], None, stride=stride, padding=1)
        bn = "%s.%d" % (pre, idx + 1)
        nbt = sd.get(bn + ".num_batches_tracked")
        if training and nbt is not None:
            nbt += 1
        h = F.batch_norm(h, sd[bn + ".running_mean"], sd[bn + ".running_var"], sd[bn + ".weight"], sd[bn + ".bias"],
                         training=training, momentum=0.1, eps=1e-5)
        h = F.leaky_relu(h, 0.2)
        idx += 3
    return F.conv2d(h, sd["%s.%d.weight" % (pre, idx)], sd["%s.%d.bias" % (pre, idx)], stride=1, padding=1)


def hinge_d_loss(logits_real, logits_fake):  # taming vqperceptual (same per-term form as LD vqperceptual.py:11-18)
    return 0.5 * (torch.mean(F.relu(1.0 - logits_real)) + torch.mean(F.relu(1.0 + logits_fake)))


def adopt_weight(weight, global_step, threshold=0, value=0.0):  # LD vqperceptual.py:20-23
    return value if global_step < threshold else weight


# --------------------------------------------------------------------------------------------
# distributions (DIST:24-62, SDIST:5-41)
# --------------------------------------------------------------------------------------------
class Gaussian:
    def __init__(self, parameters):
        self.parameters = parameters
        self.mean, logvar = torch.chunk(parameters, 2, dim=1)
        self.logvar = torch.clamp(logvar, -30.0, 20.0)
        self.std = torch.exp(0.5 * self.logvar)
        self.var = torch.exp(self.logvar)

    def kl(self, other=None):
        if other is None:
            return 0.5 * torch.sum(self.mean ** 2 + self.var - 1.0 - self.logvar, dim=[1, 2, 3])
        # SDIST:29-41: [n,1] vs [1,n] broadcast -> an [n,n] cross term summed over dim 1
        om = other.mean.squeeze().unsqueeze(0).to(self.mean)
        ov = other.var.squeeze().unsqueeze(0).to(self.mean)
        ol = other.logvar.squeeze().unsqueeze(0).to(self.mean)
        return 0.5 * torch.sum((self.mean - om) ** 2 / (ov + 1e-5) + self.var / (ov + 1e-5) - 1.0 - self.logvar + ol,
                               dim=[1])


def prior_table(dataset_stats, labels, train_on_yaw=True):
    """LOSS:82-109 -> {label: Gaussian over [8,1] means / logvars}."""
    out = {}
    for label, stats in dataset_stats.items():
        means, logvars = torch.zeros(8), torch.zeros(8)
        rot = "yaw" if train_on_yaw else "v3"
        for idx, key in enumerate(["t1", "t2", "t3", rot, "l", "h", "w", "fill_factor"]):
            if key == "yaw":
                mean, logvar = 0.0, 2 * torch.log(torch.tensor(math.pi))
            elif key in ("t1", "t2"):
                mean, logvar = 0.0, 2 * torch.log(torch.tensor(1.0))
            elif key == "fill_factor":
                mean, logvar = 0.5, 2 * torch.log(torch.tensor(math.sqrt(2)))
            else:
                mean, logvar = stats[key]
            means[idx], logvars[idx] = mean, logvar
        out[label] = Gaussian(torch.cat((means.unsqueeze(1), logvars.unsqueeze(1)), dim=1))
    return out


# --------------------------------------------------------------------------------------------
# schedules (AE:184-206)
# --------------------------------------------------------------------------------------------
def dropout_prob(step, mp, pretrain):
    init, final = mp["dropout_prob_init"], mp["dropout_prob_final"]
    gen, warm = mp["pose_conditioned_generation_steps"], mp["dropout_warmup_steps"]
    if step < pretrain or step < pretrain + gen:
        return init
    if step < warm + pretrain + gen:
        return init - (init - final) * (step - pretrain) / warm
    return final


def rescale(x):  # AE:434-436
    return 2.0 * (x - x.min()) / (x.max() - x.min()) - 1.0


def focal_loss(pred, target, gamma=2.0, alpha=0.25):  # mmdet 3.3.0 FocalLoss() defaults (Appendix B)
    C = pred.size(1)
    y = F.one_hot(target, C + 1)[:, :C].type_as(pred)
    p = pred.sigmoid()
    pt = (1 - p) * y + p * (1 - y)
    w = (alpha * y + (1 - alpha) * (1 - y)) * pt.pow(gamma)
    return (F.binary_cross_entropy_with_logits(pred, y, reduction="none") * w).mean()


# --------------------------------------------------------------------------------------------
# PoseAutoencoder.forward (AE:208-257) and PoseLoss.forward (LOSS:214-350), optimizer_idx = 0
# --------------------------------------------------------------------------------------------
class Oracle:
    """Functional CPU restatement of the hot path over a reference-layout state_dict.

    model_params: the `model.params` dict of the reference yaml (ddconfig, lossconfig, ...).
    sd: {reference parameter name: fp32 CPU tensor}; tensors with requires_grad receive .grad.
    """

    def __init__(self, model_params, sd, dataset_stats, labels):
        self.mp = model_params
        self.dd = model_params["ddconfig"]
        self.lp = model_params["lossconfig"]["params"]
        self.sd = sd
        self.labels = labels
        self.pretrain = self.lp["encoder_pretrain_steps"]
        self.priors = prior_table(dataset_stats, labels, model_params.get("train_on_yaw", True))
        self.has_disc = "loss.discriminator.main.0.weight" in sd
        # arithmetic type: float32 like the reference; float64 (sd given in double) is the "which side owns the error"
        # probe of tests/golden/make_golden.py -- the reference itself casts its inputs with .float() and cannot run in fp64
        self.dt = sd["encoder.conv_in.weight"].dtype

    def encode(self, x):  # AE:176-182
        h = encoder(self.sd, self.dd, x)
        return Gaussian(conv(self.sd, "quant_conv_obj", h)), conv(self.sd, "quant_conv_pose", h)

    def decode(self, z):  # ldm/models/autoencoder.py:330-333
        return decoder(self.sd, self.dd, conv(self.sd, "post_quant_conv", z))

    def forward(self, x, step, noise):
        """noise = (eps_obj, keep{0,1}, z_noise, eps_bbox): the reference's four random draws."""
        eps_obj, keep, z_noise, eps_bbox = noise
        x = x.contiguous().to(self.dt)
        post, pose_feat = self.encode(x)
        z = post.mean + post.std * eps_obj  # DIST:35-37
        p = dropout_prob(step, self.mp, self.pretrain)
        if p > 0:  # AE:233-235 (fresh nn.Dropout => always training mode)
            z = z * 0.0 if p >= 1.0 else z * keep.to(z) / (1.0 - p)
        if self.mp.get("add_noise_to_z_obj", True):
            z = z + z_noise
        zz = pose_decoder(self.sd, pose_feat.reshape(pose_feat.shape[0], -1))  # AE:137-160
        ncls = self.lp["num_classes"]
        bbox_post = Gaussian(torch.cat([zz[..., :8], zz[..., 8:16]], dim=-1))
        dec_pose = torch.cat([bbox_post.mean + bbox_post.std * eps_bbox, zz[..., -ncls:]], dim=-1)
        if step < self.pretrain:
            dec_obj = torch.zeros_like(x)
        else:
            enc_pose = pose_encoder(self.sd, dec_pose).view(z.shape)
            dec_obj = self.decode(z + enc_pose)
        return dec_obj, dec_pose, post, bbox_post, p

    def loss(self, rgb_gt, pose_gt, dec_obj, dec_pose, class_gt, class_names, bbox_gt, fill_gt, post, bbox_post,
             step, mask_2d_bbox, split="train"):
        lp, sd = self.lp, self.sd
        logvar = sd["loss.logvar"]
        gen_steps = lp.get("pose_conditioned_generation_steps", 7000)  # LOSS:31 default; CFG never passes it
        use_pixel = not (step < self.pretrain + gen_steps)
        mask_bg = torch.zeros_like(class_gt)
        mask_bg[class_gt != BACKGROUND_CLASS_IDX] = 1
        nbg = torch.sum(mask_bg)
        dev = rgb_gt.device  # CPU in the tests; bench.py's `gpu_eager_baseline` runs this same restatement on cuda
        zero = torch.zeros((), device=dev)
        inputs_rgb, rec_rgb = rgb_gt * mask_2d_bbox, dec_obj[:, :3] * mask_2d_bbox
        pose_rec, lhw_rec, fill_rec, logits = dec_pose[:, :4], dec_pose[:, 4:7], dec_pose[:, 7:8], dec_pose[:, 8:]

        class_loss = focal_loss(logits, class_gt)  # LOSS:176-181
        bbox_loss = (F.mse_loss(bbox_gt, lhw_rec, reduction="none") * mask_bg.unsqueeze(1)).sum() / nbg if nbg > 0 else zero
        # LOSS:269 passes (gt, rec) into (pred, gt) slots - symmetric losses
        t1 = (pose_gt[:, 0] - pose_rec[:, 0]).abs()
        t2 = (pose_gt[:, 1] - pose_rec[:, 1]).abs()
        t3 = (pose_gt[:, 2] - pose_rec[:, 2]).abs()
        v3 = F.smooth_l1_loss(torch.sin(pose_gt[:, 3]), torch.sin(pose_rec[:, 3]), reduction="none")
        pose_loss = ((t1 + t2 + t3 + v3) * mask_bg).sum() / nbg if nbg > 0 else zero
        fill_loss = (F.mse_loss(fill_gt, fill_rec.squeeze(), reduction="none") * mask_bg).sum() / nbg if nbg > 0 else zero

        rec = torch.abs(inputs_rgb - rec_rgb) if use_pixel else torch.zeros_like(inputs_rgb)  # LOSS:134-145
        if lp.get("perceptual_weight", 1.0) > 0:  # LOSS:141-145; taming's LPIPS restated in oracle/lpips_oracle.py
            lpips = getattr(self, "lpips", None)
            if lpips is None:
                raise NotImplementedError("perceptual_weight > 0 needs an LPIPS instance: set `.lpips` to an "
                                          "oracle.lpips_oracle.LPIPSOracle (vgg.pth is not downloadable offline)")
            rec = rec + lp["perceptual_weight"] * lpips(inputs_rgb.contiguous(), rec_rgb.contiguous())
        nll_el = rec / (torch.exp(logvar) + 1e-8) + logvar  # LOSS:147-158
        nll = (nll_el * mask_bg.view(-1, 1, 1, 1)).sum() / nbg if nbg > 0 else zero
        kl_obj = (post.kl() * mask_bg).sum() / nbg if nbg > 0 else zero  # LOSS:160-164
        kl_rows = torch.zeros(len(class_names), 8, device=dev, dtype=bbox_post.mean.dtype)  # LOSS:191-205
        for i, label in enumerate(class_names):
            if label == "background":
                continue
            cur = Gaussian(torch.cat((bbox_post.mean[i].unsqueeze(1), bbox_post.logvar[i].unsqueeze(1)), dim=1))
            kl_rows[i] = cur.kl(self.priors[label])
        kl_bbox = kl_rows.sum() / nbg if nbg > 0 else zero

        w_pose, w_class = lp["pose_weight"] * pose_loss, lp["class_weight"] * class_loss
        w_bbox, w_fill = lp["bbox_weight"] * bbox_loss, lp["fill_factor_weight"] * fill_loss
        # generator side of the GAN (LOSS:280-304); discriminator parameters are treated as constants here (Lightning's
        # toggle_optimizer freezes them during the optimizer_idx 0 step)
        g_loss, d_weight = torch.zeros((), device=dev), torch.zeros((), device=dev)
        disc_cfg = lp.get("disc_factor", 1.0)
        disc_factor = adopt_weight(disc_cfg, step, threshold=lp.get("disc_start", 0))
        if self.has_disc:
            dsd = {k: (v.detach() if "running" in k else v) for k, v in sd.items()
                   if k.startswith("loss.discriminator.")}
            logits_fake = discriminator(dsd, (dec_obj * mask_2d_bbox).contiguous()) * mask_bg.view(-1, 1, 1, 1)
            g_loss = -torch.mean(logits_fake)
            if disc_cfg > 0.0 and step > self.pretrain:  # LD contperceptual.py:32-43
                last = sd["decoder.conv_out.weight"]
                nll_g = torch.autograd.grad(nll, last, retain_graph=True)[0]
                g_g = torch.autograd.grad(g_loss, last, retain_graph=True)[0]
                d_weight = torch.clamp(torch.norm(nll_g) / (torch.norm(g_g) + 1e-4), 0.0, 1e4).detach() \
                    * lp.get("disc_weight", 1.0)
        if step > self.pretrain:  # LOSS:313-321
            loss = w_pose + 0.0 + nll + w_class + w_bbox + w_fill + lp["kl_weight_obj"] * kl_obj \
                + lp["kl_weight_bbox"] * kl_bbox + d_weight * disc_factor * g_loss
        else:
            loss = w_pose + w_class + w_bbox + w_fill + lp["kl_weight_bbox"] * kl_bbox
        d = lambda t: float(torch.as_tensor(t).detach().float().mean())
        log = {"total_loss": d(loss), "logvar": d(logvar), "kl_loss_obj": d(kl_obj), "nll_loss": d(nll),
               "weighted_nll_loss": d(nll), "rec_loss": d(rec), "d_weight": d(d_weight), "disc_factor": float(disc_factor), "g_loss": d(g_loss),
               "pose_loss": d(pose_loss), "weighted_pose_loss": d(w_pose), "mask_loss": 0.0,
               "weighted_mask_loss": 0.0, "class_loss": d(class_loss), "weighted_class_loss": d(w_class),
               "bbox_loss": d(bbox_loss), "weighted_bbox_loss": d(w_bbox), "t1_loss": d(t1), "t2_loss": d(t2),
               "t3_loss": d(t3), "v3_loss": d(v3), "kl_loss_bbox": d(kl_bbox),
               "weighted_kl_loss_bbox": lp["kl_weight_bbox"] * d(kl_bbox),
               "weighted_kl_loss_obj": lp["kl_weight_obj"] * d(kl_obj), "fill_factor_loss": d(fill_loss),
               "weighted_fill_factor_loss": d(w_fill)}
        return loss, {"%s/%s" % (split, k): v for k, v in log.items()}

    def discriminator_step(self, batch, step, noise, split="train"):
        """AE:321-330 + LOSS:352-375 (optimizer_idx = 1): disc_factor * hinge(D(x m) mask_bg, D(x_hat m) mask_bg) on detached
        images. Returns (d_loss, log)."""
        rgb_gt = rescale(batch["patch"].to(self.dt))
        with torch.no_grad():
            dec_obj = self.forward(rgb_gt, step, noise)[0]
        class_gt = batch["class_id"]
        mask_bg = torch.zeros_like(class_gt)
        mask_bg[class_gt != BACKGROUND_CLASS_IDX] = 1
        m2 = batch["mask_2d_bbox"]
        dsd = {k: v for k, v in self.sd.items() if k.startswith("loss.discriminator.")}
        logits_real = discriminator(dsd, (rgb_gt * m2).contiguous().detach()) * mask_bg.view(-1, 1, 1, 1)
        logits_fake = discriminator(dsd, (dec_obj * m2).contiguous().detach()) * mask_bg.view(-1, 1, 1, 1)
        disc_factor = adopt_weight(self.lp.get("disc_factor", 1.0), step, threshold=self.lp.get("disc_start", 0))
        d_loss = disc_factor * hinge_d_loss(logits_real, logits_fake)
        log = {"disc_loss": float(d_loss.detach()), "logits_real": float(logits_real.detach().mean()),
               "logits_fake": float(logits_fake.detach().mean())}
        return d_loss, {"%s/%s" % (split, k): v for k, v in log.items()}

    def training_step(self, batch, step, noise):
        """AE:295-319 with optimizer_idx = 0. Returns (loss, log, forward outputs)."""
        rgb_gt = rescale(batch["patch"].to(self.dt))  # get_input's permute round trip is an identity on values
        pose_gt = batch["pose_6d"].clone()
        pose_gt[:, 3] = batch["yaw"]  # AE:262-265
        pose_gt = pose_gt.to(self.dt)
        out = self.forward(rgb_gt, step, noise)
        dec_obj, dec_pose, post, bbox_post, _ = out
        loss, log = self.loss(rgb_gt, pose_gt, dec_obj, dec_pose, batch["class_id"], batch["class_name"],
                              batch["bbox_sizes"].to(self.dt), batch["fill_factor"].to(self.dt), post, bbox_post, step,
                              batch["mask_2d_bbox"].to(self.dt))
        return loss, log, out
