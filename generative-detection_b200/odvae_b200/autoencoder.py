"""PoseAutoencoder: the reference's LightningModule API over the B200 kernels.

Reference: src/models/autoencoder.py:34-443 (PoseAutoencoder) on LD/ldm/models/autoencoder.py:285-443
(AutoencoderKL: decode, get_input, get_last_layer, init_from_ckpt). Constructor kwargs, attribute names,
state_dict keys, `encode/decode/forward/training_step/validation_step/configure_optimizers/log_images`
signatures and return values follow the reference so that `instantiate_from_config` in the reference's
train.py builds this class unchanged. Extra (optional) kwargs: `compute_dtype` (torch.bfloat16 = tcgen05
engine, torch.float32 = SIMT validation mode) and `fused_optimizer`.

Random draws: like the reference, eps (posterior sample), the additive z noise and the bbox eps come from the
CPU generator in that order (DIST.py:36, AE.py:240, AE.py:154); they are drawn up front into pinned memory and
copied asynchronously, so no draw forces a device sync in the middle of the step. The dropout mask is drawn on
the device (nn.Dropout on a CUDA tensor in the reference). `inject_noise` overrides all four for parity tests.
"""
import importlib

import torch
import torch.nn as nn

from . import modules as M
from . import ops
from .distributions import DiagonalGaussianDistribution, FusedPosterior
from .lib import CONV_1x1

try:  # the reference harness is PyTorch-Lightning 1.9; absent in this image -> plain nn.Module stand-in
    import pytorch_lightning as pl
    _Base = pl.LightningModule
    _HAVE_PL = True
except Exception:  # pragma: no cover - depends on the environment
    _HAVE_PL = False

    class _Base(nn.Module):
        """Minimal LightningModule surface used by the hot path: global_step, device, log, log_dict."""

        def __init__(self, *a, **k):
            super().__init__()
            self._global_step = 0
            self.logged = {}

        @property
        def global_step(self):
            return self._global_step

        @global_step.setter
        def global_step(self, v):
            self._global_step = v

        @property
        def device(self):
            for p in self.parameters():
                return p.device
            return torch.device("cpu")

        def log(self, name, value, *a, **k):
            self.logged[name] = value

        def log_dict(self, d, *a, **k):
            self.logged.update(d)

POSE_6D_DIM = 4
FILL_FACTOR_DIM = 1
LHW_DIM = 3
BBOX_DIM = POSE_6D_DIM + LHW_DIM + FILL_FACTOR_DIM


def instantiate_from_config(config):
    """LD/ldm/util.py:78-93 (the reference's plugin loader), restated so this package does not import ldm."""
    if "target" not in config:
        raise KeyError("Expected key `target` to instantiate.")
    module, cls = config["target"].rsplit(".", 1)
    return getattr(importlib.import_module(module, package=None), cls)(**config.get("params", dict()))


class PoseHead(torch.autograd.Function):
    """zz [B,16+ncls] -> (dec_pose [B,8+ncls], bbox mean [B,8], clamped bbox logvar [B,8])  (AE.py:126-160)."""

    @staticmethod
    def forward(ctx, zz, eps, ncls):
        zz = ops._c(zz)
        B = zz.shape[0]
        dev = zz.device
        dec_pose = torch.empty(B, BBOX_DIM + ncls, dtype=torch.float32, device=dev)
        mean = torch.empty(B, BBOX_DIM, dtype=torch.float32, device=dev)
        logvar = torch.empty(B, BBOX_DIM, dtype=torch.float32, device=dev)
        ops.call("pose_head_fwd", zz, eps, dec_pose, mean, logvar, B, ncls)
        ctx.ncls = ncls
        ctx.save_for_backward(zz, eps)
        return dec_pose, mean, logvar

    @staticmethod
    def backward(ctx, ddec, dmean, dlogvar):
        zz, eps = ctx.saved_tensors
        dzz = torch.empty_like(zz)
        c = lambda t: None if t is None else ops._c(t.float())
        ops.call("pose_head_bwd", zz, eps, c(ddec), c(dmean), c(dlogvar), dzz, zz.shape[0], ctx.ncls)
        return dzz, None, None


class BBoxPosterior(DiagonalGaussianDistribution):
    """bbox posterior whose moments come from the pose-head kernel (same fields as the reference object)."""

    def __init__(self, mean, logvar, parameters):
        self.mean, self.logvar, self.parameters = mean, logvar, parameters
        self.deterministic = False

    @property
    def std(self):
        return torch.exp(0.5 * self.logvar)

    @property
    def var(self):
        return torch.exp(self.logvar)


class PoseAutoencoder(_Base):
    def __init__(self, ddconfig, lossconfig, embed_dim, euler_convention, ckpt_path=None, ignore_keys=[],
                 image_mask_key=None, image_rgb_key="patch", pose_key="pose_6d", fill_factor_key="fill_factor",
                 pose_perturbed_key="pose_6d_perturbed", class_key="class_id", bbox_key="bbox_sizes",
                 colorize_nlabels=None, monitor=None, activation="relu", feat_dims=[16, 16, 16],
                 pose_decoder_config=None, pose_encoder_config=None, dropout_prob_init=1.0, dropout_prob_final=0.7,
                 dropout_warmup_steps=5000, pose_conditioned_generation_steps=10000, add_noise_to_z_obj=True,
                 train_on_yaw=True, compute_dtype=torch.bfloat16, fused_optimizer=True):
        super().__init__()
        if isinstance(compute_dtype, str):
            compute_dtype = {"bf16": torch.bfloat16, "bfloat16": torch.bfloat16, "fp32": torch.float32,
                             "float32": torch.float32, "fp16": torch.float16, "float16": torch.float16,
                             "mixed": "mixed"}[compute_dtype]
        self.compute_dtype = compute_dtype
        self.fused_optimizer = fused_optimizer
        self.encoder_pretrain_steps = lossconfig["params"]["encoder_pretrain_steps"]
        self.train_on_yaw = train_on_yaw
        self.dropout_prob_final, self.dropout_prob_init = dropout_prob_final, dropout_prob_init
        self.dropout_prob = self.dropout_prob_init
        self.dropout_warmup_steps = dropout_warmup_steps
        self.pose_conditioned_generation_steps = pose_conditioned_generation_steps
        self.add_noise_to_z_obj = add_noise_to_z_obj
        self.feature_dims = feat_dims
        self.image_rgb_key, self.pose_key, self.pose_perturbed_key = image_rgb_key, pose_key, pose_perturbed_key
        self.class_key, self.bbox_key, self.fill_factor_key = class_key, bbox_key, fill_factor_key
        self.image_mask_key = image_mask_key
        self.encoder = M.Encoder(compute_dtype=compute_dtype, **ddconfig)
        self.decoder = M.Decoder(compute_dtype=compute_dtype, **ddconfig)
        lossconfig["params"]["train_on_yaw"] = self.train_on_yaw
        self.loss = instantiate_from_config(lossconfig)
        assert ddconfig["double_z"]
        zc = ddconfig["z_channels"]
        self.quant_conv_obj = M.Conv2d(2 * zc, 2 * embed_dim, 1, CONV_1x1)
        self.quant_conv_pose = M.Conv2d(2 * zc, embed_dim, 1, CONV_1x1)
        self.post_quant_conv = M.Conv2d(embed_dim, zc, 1, CONV_1x1)
        self.embed_dim = embed_dim
        if colorize_nlabels is not None:
            assert type(colorize_nlabels) == int
            self.register_buffer("colorize", torch.randn(3, colorize_nlabels, 1, 1))
        if monitor is not None:
            self.monitor = monitor
        if ckpt_path is not None:
            self.init_from_ckpt(ckpt_path, ignore_keys=ignore_keys)
        self.num_classes = lossconfig["params"]["num_classes"]
        self.pose_decoder = instantiate_from_config(pose_decoder_config)
        self.pose_encoder = instantiate_from_config(pose_encoder_config)
        self.z_channels = zc
        self.euler_convention = euler_convention
        self.feat_dims = feat_dims
        self.inject_noise = None  # (eps_obj, keep{0,1}, z_noise, eps_bbox) CPU/GPU tensors: parity tests only
        # optimizer_idx == 1 re-runs the whole autoencoder forward in the reference (AE.py:321-327) with fresh random draws.
        # True: reuse the reconstruction of the optimizer_idx == 0 call of the same batch_idx instead.
        self.reuse_forward_for_disc = False
        self._fwd_cache = None
        self._keep_scale_dev = None  # device float [1]: dropout 1/(1-p) read by the kernels (graphs.GraphedTrainStep)
        self.register_load_state_dict_post_hook(lambda *a: ops.bump_weight_epoch())
        self.fmt = ops.assign_format(self, compute_dtype)  # quant convs included (encoder/decoder did their own)
        # all-fp16 mode: static power-of-two scales the 16-bit activation gradients travel with (ops.GradScale); None = chosen
        # per batch size in _set_grad_scales, a number = fixed by the user
        self.grad_scale_dec = self.grad_scale_enc = None
        self.gs_dec, self.gs_enc = ops.GradScale(), ops.GradScale()
        for m in (self.decoder, self.post_quant_conv):
            ops.assign_grad_scales(m, self.gs_dec)
        for m in (self.encoder, self.quant_conv_obj, self.quant_conv_pose):
            ops.assign_grad_scales(m, self.gs_enc)

    # ------------------------------------------------------------------ step bookkeeping
    @property
    def global_step(self):
        """Lightning's optimizer-step counter. Under pytorch-lightning the base-class property is read-only and backed by
        the trainer; Lightning-free loops (fit.Fit, trainer.TrainStep users, bench.py, tests) assign to it, which installs
        an override; assigning None hands the counter back to the trainer."""
        ov = self.__dict__.get("_step_override")
        return ov if ov is not None else super().global_step

    @global_step.setter
    def global_step(self, v):
        self.__dict__["_step_override"] = None if v is None else int(v)

    # ------------------------------------------------------------------ checkpoint / misc API
    def init_from_ckpt(self, path, ignore_keys=list()):
        sd = torch.load(path, map_location="cpu")["state_dict"]
        for k in list(sd.keys()):
            for ik in ignore_keys:
                if k.startswith(ik):
                    print("Deleting key {} from state_dict.".format(k))
                    del sd[k]
        self.load_state_dict(sd, strict=False)
        print(f"Restored from {path}")

    def get_last_layer(self):
        return self.decoder.conv_out.weight

    def get_input(self, batch, k):  # LD autoencoder.py:344-349
        x = batch[k]
        if len(x.shape) == 3:
            x = x[..., None]
        return x.permute(0, 3, 1, 2).to(memory_format=torch.contiguous_format).float()

    def get_pose_input(self, batch, k):
        x = batch[k]
        if self.train_on_yaw:
            x[:, 3] = batch["yaw"]
        return x.to(memory_format=torch.contiguous_format).float()

    def get_mask_input(self, batch, k):
        return None if k is None else batch[k]

    def get_class_input(self, batch, k):
        return batch[k]

    def get_bbox_input(self, batch, k):
        return batch[k]

    def get_fill_factor_input(self, batch, k):
        return batch[k]

    def _get_perturbed_pose(self, batch, k):
        x = batch[k].squeeze(1)
        if self.train_on_yaw:
            x = torch.zeros_like(x)
            x[:, -1] = batch["yaw_perturbed"]
        return x

    def _rescale(self, x):  # AE.py:434-436, min/max over the whole batch tensor
        if x.is_cuda:
            f = self.fmt
            # (the NHWC copy for encoder.conv_in is only wanted where that conv takes the generic 64-channel-padded path)
            thin = ops.THIN_CONV and f.fwd != ops.F32 and x.dim() == 4 and x.shape[1] <= 4 and not x.requires_grad
            return ops.rescale_minmax(x, nhwc=(ops.cpad(x.shape[1]), f.fwd, f.label(f.fwd)) if x.dim() == 4 and not thin
                                      else None)
        return 2. * (x - x.min()) / (x.max() - x.min()) - 1.

    def check_overflow(self):
        """fp16 range check (synchronises): raises RuntimeError if, since the last call, a forward activation was stored
        saturated at +-65504 (pre-GroupNorm residual stream beyond fp16's range: run this model with compute_dtype=
        torch.bfloat16) or an activation gradient was (lower grad_scale_dec / grad_scale_enc)."""
        fwd, grad = ops.read_overflow(self.device)
        if fwd:
            raise RuntimeError("fp16 overflow: a forward activation exceeded 65504 and was stored saturated; this checkpoint "
                               "needs compute_dtype=torch.bfloat16 (8 exponent bits) -- results of this step are not valid")
        if grad:
            raise RuntimeError("fp16 overflow: an activation gradient exceeded 65504 after scaling (gs_dec=%g, gs_enc=%g); set "
                               "grad_scale_dec / grad_scale_enc lower -- results of this step are not valid"
                               % (self.gs_dec.value, self.gs_enc.value))

    def _get_dropout_prob(self):  # AE.py:184-206
        step = self.global_step
        if step < self.encoder_pretrain_steps:
            return self.dropout_prob_init
        if step < self.encoder_pretrain_steps + self.pose_conditioned_generation_steps:
            return self.dropout_prob_init
        if step < self.dropout_warmup_steps + self.encoder_pretrain_steps + self.pose_conditioned_generation_steps:
            return self.dropout_prob_init - (self.dropout_prob_init - self.dropout_prob_final) * \
                (step - self.encoder_pretrain_steps) / self.dropout_warmup_steps
        return self.dropout_prob_final

    # ------------------------------------------------------------------ hot path
    def _set_grad_scales(self, B):
        """fp16 gradients: the decoder-side gradient starts at |d loss / d dec_obj| <= 1 / (e^logvar * sum(mask_bg)) ~ 1/B per
        element and grows by <~ 30x towards the latent; the encoder-side one is dominated by the pose / class / bbox heads,
        ~ weight / B per element. Measured on the full config (tools/experiments notes in profiles/README.md, reference run
        with backward hooks): decoder rms 0.02-1.2 (max 32) and encoder rms 30-750 (max 9e3) at B = 2. S_dec = 64 B and
        S_enc = 16384 B / max loss weight put both into [2^-4, 2^12] with > 16x headroom below fp16's 65504."""
        if self.fmt.grad != ops.F16:
            self.gs_dec.value = self.gs_enc.value = 1.0
            return
        wmax = max(1.0, float(self.loss.pose_weight), float(self.loss.class_weight), float(self.loss.bbox_weight),
                   float(self.loss.fill_factor_weight))
        self.gs_dec.value = float(self.grad_scale_dec) if self.grad_scale_dec else ops.pow2_near(64.0 * B)
        self.gs_enc.value = float(self.grad_scale_enc) if self.grad_scale_enc else ops.pow2_near(16384.0 * B / wmax)

    def _draw_noise(self, B, lat_shape, p, sample_posterior, dev):
        """The reference's four draws (same generator, same order); returns device tensors or None."""
        shape = (B,) + tuple(lat_shape)
        if self.inject_noise is not None:
            eps, keep, noise, eps_bbox = self.inject_noise
            f = lambda t, dt: None if t is None else t.to(device=dev, dtype=dt).contiguous()
            keep = f(keep, torch.uint8) if p > 0 else None
            return (f(eps, torch.float32), keep, f(noise, torch.float32) if self.add_noise_to_z_obj else None,
                    f(eps_bbox, torch.float32))
        pin = dict(pin_memory=True) if dev.type == "cuda" else {}
        eps = torch.randn(shape, **pin).to(dev, non_blocking=True) if sample_posterior else None
        noise = None
        if self.add_noise_to_z_obj:  # Normal(0,1).sample(shape) == torch.normal(zeros, ones)
            noise = torch.normal(torch.zeros(shape), torch.ones(shape))
            noise = (noise.pin_memory() if pin else noise).to(dev, non_blocking=True)
        eps_bbox = torch.randn(B, BBOX_DIM, **pin).to(dev, non_blocking=True) if sample_posterior else None
        keep = None
        if p > 0:
            keep = torch.empty(shape, dtype=torch.float32, device=dev).bernoulli_(1.0 - p).to(torch.uint8) if p < 1.0 \
                else torch.zeros(shape, dtype=torch.uint8, device=dev)
        return eps, keep, noise, eps_bbox

    def _encode_nhwc(self, x):
        h = self.encoder.forward_nhwc(x)
        return self.quant_conv_obj(h), self.quant_conv_pose(h)

    def _posterior(self, moments, enc_pose=None, eps=None, keep=None, noise=None, keep_scale=1.0, want_z=False):
        zc = self.embed_dim
        if want_z and eps is None:
            eps = torch.zeros(moments.shape[0], zc, moments.shape[1], moments.shape[2], device=moments.device)
        mean, logvar, z, kl = ops.SampleKL.apply(moments, enc_pose, eps, keep, noise, keep_scale, zc, ops.cpad(zc),
                                                 want_z, self.fmt, self.gs_dec, self.gs_enc,
                                                 self._keep_scale_dev if keep is not None else None)
        post = FusedPosterior(mean, logvar, kl,
                              lambda: ops.ToNCHW.apply(moments, 2 * zc, self.compute_dtype, self.gs_enc))
        return post, (z if want_z else None)

    def encode(self, x):
        x = x.to(self.device)
        moments, pose_feat = self._encode_nhwc(x.contiguous().float())
        post, _ = self._posterior(moments)
        return post, ops.ToNCHW.apply(pose_feat, self.embed_dim, self.compute_dtype, self.gs_enc)

    def decode(self, z):
        h = ops.ToNHWC.apply(z.to(self.device), ops.cpad(self.embed_dim), self.compute_dtype, self.gs_dec)
        return self.decoder.forward_nhwc(self.post_quant_conv(h))

    def _decode_pose_to_distribution(self, z, sample_posterior=True):
        ncls = self.num_classes
        c_pred = z[..., -ncls:]
        bbox_moments = torch.cat([z[..., :BBOX_DIM], z[..., BBOX_DIM:2 * BBOX_DIM]], dim=-1).to(self.device)
        return DiagonalGaussianDistribution(bbox_moments), c_pred

    def _decode_pose(self, x, sample_posterior=True, eps_bbox=None):
        x = x.reshape(x.size(0), -1)
        zz = self.pose_decoder(x)
        if sample_posterior and eps_bbox is None:
            eps_bbox = torch.randn(x.size(0), BBOX_DIM).to(zz.device)
        dec_pose, mean, logvar = PoseHead.apply(zz, eps_bbox if sample_posterior else None, self.num_classes)
        return dec_pose, BBoxPosterior(mean, logvar, zz[..., :2 * BBOX_DIM])

    def _encode_pose(self, x):
        y = self.pose_encoder(x)
        return y.view(y.size(0), self.feature_dims[0], self.feature_dims[1], self.feature_dims[2])

    def forward(self, input_im, sample_posterior=True):
        input_im = input_im.to(self.device).to(memory_format=torch.contiguous_format).float()
        B = input_im.shape[0]
        self._set_grad_scales(B)
        self.dropout_prob = p = self._get_dropout_prob()
        eps, keep, noise, eps_bbox = self._draw_noise(B, self.feature_dims, p, sample_posterior, input_im.device)
        moments, pose_feat = self._encode_nhwc(input_im)
        dec_pose, bbox_posterior = self._decode_pose(
            ops.ToNCHW.apply(pose_feat, self.embed_dim, self.compute_dtype, self.gs_enc), sample_posterior, eps_bbox)
        keep_scale = 1.0 / (1.0 - p) if 0 < p < 1.0 else 1.0
        if self.global_step < self.encoder_pretrain_steps:  # AE.py:246-247: decoder skipped
            posterior_obj, _ = self._posterior(moments)
            dec_obj = torch.zeros_like(input_im)
        else:
            enc_pose = self._encode_pose(dec_pose)
            posterior_obj, z = self._posterior(moments, enc_pose.reshape(B, -1), eps, keep, noise, keep_scale, True)
            dec_obj = self.decoder.forward_nhwc(self.post_quant_conv(z))
        return dec_obj, dec_pose, posterior_obj, bbox_posterior

    def _unpack(self, batch):
        dev = self.device
        rgb_gt = batch[self.image_rgb_key]
        if rgb_gt.dim() == 3:
            rgb_gt = rgb_gt[..., None]
        # get_input's permute(0,3,1,2) followed by training_step's permute(0,2,3,1) is the identity on values
        rgb_gt = self._rescale(rgb_gt.to(dev, non_blocking=True).float())
        pose_gt = self.get_pose_input(batch, self.pose_key).to(dev, non_blocking=True)
        class_gt = self.get_class_input(batch, self.class_key).to(dev, non_blocking=True)
        bbox_gt = self.get_bbox_input(batch, self.bbox_key).to(dev, non_blocking=True)
        fill_gt = self.get_fill_factor_input(batch, self.fill_factor_key).to(dev, non_blocking=True).float()
        return rgb_gt, pose_gt, class_gt, batch["class_name"], bbox_gt, fill_gt, batch["mask_2d_bbox"]

    def _gan_active(self):
        from .losses import adopt_weight
        return adopt_weight(self.loss.disc_factor, self.global_step, threshold=self.loss.discriminator_iter_start) != 0

    def _discriminator_inputs(self, rgb_gt, batch_idx):
        """Forward outputs for the optimizer_idx == 1 call: only `dec_obj` is read there (detached, LOSS.py:355-356)."""
        B = rgb_gt.shape[0]
        if not self._gan_active() and self.loss.skip_inactive_gan:
            # d_loss = 0 * hinge: nothing of the forward is used; keep the CPU random stream where the reference leaves it
            self.dropout_prob = p = self._get_dropout_prob()
            if self.inject_noise is None:
                self._draw_noise(B, self.feature_dims, p, True, rgb_gt.device)
            return None, None, None, None
        cache, self._fwd_cache = self._fwd_cache, None
        if self.reuse_forward_for_disc and cache is not None and cache[0] == batch_idx and cache[1].shape[0] == B:
            return cache[1], None, None, None
        with torch.no_grad():
            return self.forward(rgb_gt)

    def training_step(self, batch, batch_idx, optimizer_idx=0):
        rgb_gt, pose_gt, class_gt, labels, bbox_gt, fill_gt, mask_2d_bbox = self._unpack(batch)
        mask_gt = self.get_mask_input(batch, self.image_mask_key)
        if optimizer_idx == 1:
            dec_obj, dec_pose, posterior_obj, bbox_posterior = self._discriminator_inputs(rgb_gt, batch_idx)
        else:
            dec_obj, dec_pose, posterior_obj, bbox_posterior = self.forward(rgb_gt)
            self._fwd_cache = (batch_idx, dec_obj.detach()) if self.reuse_forward_for_disc else None
        self.log("dropout_prob", self.dropout_prob, prog_bar=True, logger=True, on_step=True, on_epoch=True)
        loss, log_dict = self.loss(rgb_gt, mask_gt, pose_gt, dec_obj, dec_pose, class_gt, labels, bbox_gt, fill_gt,
                                   posterior_obj, bbox_posterior, optimizer_idx, self.global_step, mask_2d_bbox,
                                   last_layer=self.get_last_layer(), split="train")
        self.log("aeloss" if optimizer_idx == 0 else "discloss", loss, prog_bar=True, logger=True, on_step=True,
                 on_epoch=True)
        self.log_dict(log_dict, prog_bar=False, logger=True, on_step=True, on_epoch=False)
        return loss

    def validation_step(self, batch, batch_idx):
        rgb_gt, pose_gt, class_gt, labels, bbox_gt, fill_gt, mask_2d_bbox = self._unpack(batch)
        mask_gt = self.get_mask_input(batch, self.image_mask_key)
        dec_obj, dec_pose, posterior_obj, bbox_posterior = self.forward(rgb_gt)
        args = (rgb_gt, mask_gt, pose_gt, dec_obj, dec_pose, class_gt, labels, bbox_gt, fill_gt, posterior_obj,
                bbox_posterior)
        _, log_dict_ae = self.loss(*args, 0, self.global_step, mask_2d_bbox, last_layer=self.get_last_layer(),
                                   split="val")
        _, log_dict_disc = self.loss(*args, 1, self.global_step, mask_2d_bbox, last_layer=self.get_last_layer(),
                                     split="val")
        self.log("val/rec_loss", log_dict_ae["val/rec_loss"], sync_dist=True)
        del log_dict_ae["val/rec_loss"]
        self.log_dict(log_dict_ae)
        self.log_dict(log_dict_disc)
        return self.log_dict

    def ae_parameters(self):
        return (list(self.encoder.parameters()) + list(self.decoder.parameters()) +
                list(self.quant_conv_obj.parameters()) + list(self.quant_conv_pose.parameters()) +
                list(self.post_quant_conv.parameters()) + list(self.pose_encoder.parameters()) +
                list(self.pose_decoder.parameters()))

    def configure_optimizers(self):
        lr = self.learning_rate
        if self.fused_optimizer and next(self.parameters()).is_cuda:
            from .optim import FusedAdam
            opt_ae = FusedAdam(self.ae_parameters(), lr=lr, betas=(0.5, 0.9))
            opt_disc = FusedAdam(self.loss.discriminator.parameters(), lr=lr, betas=(0.5, 0.9))
        else:
            opt_ae = torch.optim.Adam(self.ae_parameters(), lr=lr, betas=(0.5, 0.9))
            opt_disc = torch.optim.Adam(self.loss.discriminator.parameters(), lr=lr, betas=(0.5, 0.9))
        return [opt_ae, opt_disc], []

    # ------------------------------------------------------------------ image logging (AE.py:379-432)
    def _perturb_poses(self, batch, dec_pose):
        v3 = self._get_perturbed_pose(batch, self.pose_perturbed_key).squeeze()[:, -1]
        out = dec_pose.clone()
        out[:, 3] = v3.to(out)
        return out.to(self.device)

    def _perturbed_pose_forward(self, posterior_obj, dec_pose, batch, sample_posterior=True):
        z_obj = posterior_obj.sample() if sample_posterior else posterior_obj.mode()
        enc = self._encode_pose(self._perturb_poses(batch, dec_pose))
        return self.decode(z_obj + enc)

    @torch.no_grad()
    def log_images(self, batch, only_inputs=False, **kwargs):
        log = dict()
        x_rgb = batch[self.image_rgb_key].to(self.device).float()
        x_rgb = self._rescale(x_rgb)
        if not only_inputs:
            xrec, poserec, posterior_obj, _ = self.forward(x_rgb)
            xrec_pp = self._perturbed_pose_forward(posterior_obj, poserec, batch)
            log["reconstructions_rgb"] = xrec[:, :3].clone().detach()
            log["perturbed_pose_reconstruction_rgb"] = xrec_pp[:, :3].clone().detach()
        log["inputs_rgb"] = x_rgb.clone().detach()
        return log


class Autoencoder(PoseAutoencoder):
    """Name kept for import parity with src/models/autoencoder.py:29-32."""
