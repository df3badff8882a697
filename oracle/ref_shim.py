"""TEST INFRASTRUCTURE ONLY (oracle/): import shim that lets the UNMODIFIED reference modules under
/root/reference run in the build container, where three of their dependencies are absent.

Used by tests/golden/make_golden.py (fixture generation) and by nothing on the product path.
/root/reference does not exist on the GPU box, so nothing that runs there may import this file.

What is shimmed:
  * pytorch_lightning.LightningModule  -> nn.Module + global_step/device/log/log_dict
  * taming.modules.{vqvae.quantize, discriminator.model, losses.lpips, losses.vqperceptual}
      restated from the public taming-transformers definitions;
      the source is NOT in /root/reference (pinned 049ce3ce, status "missing") -> parity unpinned
  * mmdet.models.losses.focal_loss.FocalLoss  (mmdet 3.3.0 defaults; Appendix B) -> parity unpinned
"""
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

REFERENCE_ROOT = os.environ.get("ODVAE_REFERENCE_ROOT", "/root/reference")


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "models"))


class _LightningModule(nn.Module):
    def __init__(self, *a, **k):
        super().__init__()
        self._shim_global_step = 0
        self._shim_logged = {}

    @property
    def global_step(self):
        return self._shim_global_step

    @global_step.setter
    def global_step(self, v):
        self._shim_global_step = v

    @property
    def device(self):
        for p in self.parameters():
            return p.device
        return torch.device("cpu")

    def log(self, name, value, *a, **k):
        self._shim_logged[name] = value

    def log_dict(self, d, *a, **k):
        self._shim_logged.update(d)


class NLayerDiscriminator(nn.Module):
    """PatchGAN discriminator (public taming definition)."""

    def __init__(self, input_nc=3, ndf=64, n_layers=3, use_actnorm=False):
        super().__init__()
        assert not use_actnorm
        kw, padw = 4, 1
        seq = [nn.Conv2d(input_nc, ndf, kw, 2, padw), nn.LeakyReLU(0.2, True)]
        nf_mult = 1
        for n in range(1, n_layers):
            nf_prev, nf_mult = nf_mult, min(2 ** n, 8)
            seq += [nn.Conv2d(ndf * nf_prev, ndf * nf_mult, kw, 2, padw, bias=False),
                    nn.BatchNorm2d(ndf * nf_mult), nn.LeakyReLU(0.2, True)]
        nf_prev, nf_mult = nf_mult, min(2 ** n_layers, 8)
        seq += [nn.Conv2d(ndf * nf_prev, ndf * nf_mult, kw, 1, padw, bias=False),
                nn.BatchNorm2d(ndf * nf_mult), nn.LeakyReLU(0.2, True)]
        seq += [nn.Conv2d(ndf * nf_mult, 1, kw, 1, padw)]
        self.main = nn.Sequential(*seq)

    def forward(self, x):
        return self.main(x)


def weights_init(m):
    name = m.__class__.__name__
    if name.find("Conv") != -1:
        nn.init.normal_(m.weight.data, 0.0, 0.02)
    elif name.find("BatchNorm") != -1:
        nn.init.normal_(m.weight.data, 1.0, 0.02)
        nn.init.constant_(m.bias.data, 0)


class LPIPS(nn.Module):
    """Placeholder: VGG weights cannot be fetched offline; never called with perceptual_weight=0."""

    def __init__(self, *a, **k):
        super().__init__()

    def forward(self, x, y):
        raise RuntimeError("LPIPS is unavailable offline (needs vgg.pth); use perceptual_weight=0")


def hinge_d_loss(logits_real, logits_fake):
    return 0.5 * (torch.mean(F.relu(1.0 - logits_real)) + torch.mean(F.relu(1.0 + logits_fake)))


def vanilla_d_loss(logits_real, logits_fake):
    return 0.5 * (torch.mean(F.softplus(-logits_real)) + torch.mean(F.softplus(logits_fake)))


def adopt_weight(weight, global_step, threshold=0, value=0.0):
    # same semantics as the in-tree copy at ldm/modules/losses/vqperceptual.py:20-23
    return value if global_step < threshold else weight


class FocalLoss(nn.Module):
    """mmdet 3.3.0 FocalLoss() defaults: sigmoid, gamma 2, alpha .25, mean, weight 1 (Appendix B)."""

    def __init__(self, gamma=2.0, alpha=0.25, loss_weight=1.0):
        super().__init__()
        self.gamma, self.alpha, self.loss_weight = gamma, alpha, loss_weight

    def forward(self, pred, target):
        C = pred.size(1)
        y = F.one_hot(target, C + 1)[:, :C].type_as(pred)
        p = pred.sigmoid()
        pt = (1 - p) * y + p * (1 - y)
        w = (self.alpha * y + (1 - self.alpha) * (1 - y)) * pt.pow(self.gamma)
        loss = F.binary_cross_entropy_with_logits(pred, y, reduction="none") * w
        return self.loss_weight * loss.mean()


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install():
    """Seed sys.modules / sys.path so `import src.models.autoencoder` resolves to the reference."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    if "pytorch_lightning" not in sys.modules:
        _module("pytorch_lightning", LightningModule=_LightningModule, __version__="1.9.0")
    _module("taming")
    _module("taming.modules")
    _module("taming.modules.vqvae")
    _module("taming.modules.vqvae.quantize", VectorQuantizer2=type("VectorQuantizer2", (nn.Module,), {}))
    _module("taming.modules.discriminator")
    _module("taming.modules.discriminator.model", NLayerDiscriminator=NLayerDiscriminator,
            weights_init=weights_init)
    _module("taming.modules.losses")
    _module("taming.modules.losses.lpips", LPIPS=LPIPS)
    _module("taming.modules.losses.vqperceptual", LPIPS=LPIPS, NLayerDiscriminator=NLayerDiscriminator,
            weights_init=weights_init, hinge_d_loss=hinge_d_loss, vanilla_d_loss=vanilla_d_loss,
            adopt_weight=adopt_weight, torch=torch, nn=nn, F=F)
    _module("mmdet")
    _module("mmdet.models")
    _module("mmdet.models.losses")
    _module("mmdet.models.losses.focal_loss", FocalLoss=FocalLoss)
    # the reference package is literally named `src`; drop any other `src` already imported
    for k in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
        del sys.modules[k]
    for p in (os.path.join(REFERENCE_ROOT, "submodules", "latent-diffusion"), REFERENCE_ROOT):
        if p in sys.path:
            sys.path.remove(p)
        sys.path.insert(0, p)


def build_reference_model(model_params, global_step):
    """Instantiate the reference PoseAutoencoder (CPU, fp32) from a `model.params` dict."""
    install()
    from src.models.autoencoder import PoseAutoencoder  # noqa: the REFERENCE's class
    model = PoseAutoencoder(**model_params)
    model.global_step = global_step
    return model
