"""DiagonalGaussianDistribution with the reference's API.

Reference: LD/ldm/modules/distributions/distributions.py:24-62 and the override at
src/util/distributions.py:5-41 (`kl(other)` with eps=1e-5 and the [n,1] vs [1,n] broadcast that
yields an [n,n] cross term summed over dim 1 -- reproduced as is).

Two flavours:
  * DiagonalGaussianDistribution(parameters): generic, for small/irregular parameter tensors (the
    [B,16] bbox posterior, the [8,2] per-class priors). Elementwise torch ops on a few hundred floats.
  * FusedPosterior: the object posterior [B,16,16,16]; mean / clamped logvar / per-sample KL come out
    of the fused sample+KL kernel (ops.SampleKL) and stay connected to autograd.
"""
import numpy as np
import torch


class DiagonalGaussianDistribution(object):
    def __init__(self, parameters, deterministic=False):
        self.parameters = parameters
        self.mean, self.logvar = torch.chunk(parameters, 2, dim=1)
        self.logvar = torch.clamp(self.logvar, -30.0, 20.0)
        self.deterministic = deterministic
        self.std = torch.exp(0.5 * self.logvar)
        self.var = torch.exp(self.logvar)
        if self.deterministic:
            self.var = self.std = torch.zeros_like(self.mean)

    def sample(self):
        # the reference draws on the CPU generator and moves the draw to the device (distributions.py:36)
        eps = torch.randn(self.mean.shape).to(device=self.parameters.device)
        return self.mean + self.std * eps

    def mode(self):
        return self.mean

    def kl(self, other=None):
        if self.deterministic:
            return torch.Tensor([0.])
        if other is None:
            return 0.5 * torch.sum(torch.pow(self.mean, 2) + self.var - 1.0 - self.logvar, dim=[1, 2, 3])
        om = other.mean.squeeze().unsqueeze(0).to(self.mean)
        ov = other.var.squeeze().unsqueeze(0).to(self.mean)
        ol = other.logvar.squeeze().unsqueeze(0).to(self.mean)
        dims = list(range(1, om.dim()))
        return 0.5 * torch.sum(torch.pow(self.mean - om, 2) / (ov + 1e-5) + self.var / (ov + 1e-5)
                               - 1.0 - self.logvar + ol, dim=dims)

    def nll(self, sample, dims=[1, 2, 3]):
        if self.deterministic:
            return torch.Tensor([0.])
        logtwopi = np.log(2.0 * np.pi)
        return 0.5 * torch.sum(logtwopi + self.logvar + torch.pow(sample - self.mean, 2) / self.var, dim=dims)


class FusedPosterior(DiagonalGaussianDistribution):
    """Object posterior whose moments were produced by the fused kernel (NCHW fp32 at the boundary)."""

    def __init__(self, mean, logvar, kl, parameters_fn):
        self.mean, self.logvar = mean, logvar
        self._kl = kl
        self._parameters_fn = parameters_fn
        self._parameters = None
        self.deterministic = False

    @property
    def parameters(self):
        if self._parameters is None:
            self._parameters = self._parameters_fn()
        return self._parameters

    @property
    def std(self):
        return torch.exp(0.5 * self.logvar)

    @property
    def var(self):
        return torch.exp(self.logvar)

    def sample(self):
        eps = torch.randn(self.mean.shape).to(device=self.mean.device)
        return self.mean + self.std * eps

    def kl(self, other=None):
        if other is None:
            return self._kl
        return super().kl(other)
