"""Pose MLPs (src/modules/autoencodermodules/pose_decoder.py:60-97, pose_encoder.py:59-131).

Same constructor kwargs, parameter names (`layers.{0,2,4}`, `coord_linear`, `latent_linear`,
`layers.{1,3}`) and maths as the reference; the Linear layers and activations run on the library's
fp32 GEMM / activation kernels (M = batch, launch-latency bound: SURVEY 8 a14).
"""
import numpy as np
import torch
import torch.nn as nn

from . import modules as M
from . import ops

POSE_DIM = 4
LHW_DIM = 3
FILL_FACTOR_DIM = 1


class PoseDecoderSpatialVAE(nn.Module):
    def __init__(self, num_classes=2, num_channels=16, n=16, m=16, activation="tanh", **kwargs):
        super().__init__()
        if activation != "tanh":
            raise NotImplementedError("only the tanh pose decoder of the north-star config is implemented")
        n_in = num_channels * n * m
        self.latent_dim = (POSE_DIM + LHW_DIM + FILL_FACTOR_DIM) * 2 + num_classes
        self.n = n_in
        hidden_dim = kwargs.get("hidden_dim", 500)
        num_layers = kwargs.get("num_layers", 2)
        if kwargs.get("resid", False):
            raise NotImplementedError("resid=True is not used by the north-star config")
        layers = [M.Linear(n_in, hidden_dim), M.Activation("tanh")]
        for _ in range(1, num_layers):
            layers += [M.Linear(hidden_dim, hidden_dim), M.Activation("tanh")]
        layers.append(M.Linear(hidden_dim, self.latent_dim))
        self.layers = nn.Sequential(*layers)

    def forward(self, x):
        return self.layers(x)


class TileAdd(torch.autograd.Function):
    """h[b, p*F + f] = h_x[p*F + f] + h_z[b, f]   (pose_encoder.py:121-125, h_x is batch-constant)."""

    @staticmethod
    def forward(ctx, h_x, h_z, P):
        B, F = h_z.shape
        out = torch.empty(B, P * F, dtype=torch.float32, device=h_z.device)
        ops.call("tile_add_fwd", h_x, h_z, out, B, P, F)
        ctx.P = P
        return out

    @staticmethod
    def backward(ctx, dh):
        dh = ops._c(dh.float())
        B = dh.shape[0]
        P = ctx.P
        F = dh.shape[1] // P
        dhx = torch.empty(1, P * F, dtype=torch.float32, device=dh.device)
        dhz = torch.empty(B, F, dtype=torch.float32, device=dh.device)
        ops.call("tile_add_bwd", dh, dhx, dhz, B, P, F)
        return dhx, dhz, None


class PoseEncoderSpatialVAE(nn.Module):
    def __init__(self, num_classes=2, num_channels=16, n=16, m=16, activation="swish", hidden_dim=500, num_layers=2):
        super().__init__()
        if activation != "swish":
            raise NotImplementedError("only the swish pose encoder of the north-star config is implemented")
        latent_dim = POSE_DIM + LHW_DIM + FILL_FACTOR_DIM + num_classes
        n_out = num_channels * n * m
        self.num_channels, self.n, self.m = num_channels, n, m
        self.in_dim = 2
        self.num_coords = n * m
        self.feat_size = 4
        self.h_dim = self.num_coords * self.feat_size
        self.x_dim = self.in_dim * self.num_coords
        self.z_dim = latent_dim
        self.coord_linear = M.Linear(self.x_dim, self.h_dim)
        self.latent_linear = M.Linear(self.z_dim, self.feat_size, bias=False)
        layers = [M.Activation("silu")]
        for layer_id in range(1, num_layers):
            layers.append(M.Linear(self.h_dim if layer_id == 1 else hidden_dim, hidden_dim))
            layers.append(M.Activation("silu"))
        layers.append(M.Linear(hidden_dim, n_out))
        self.layers = nn.Sequential(*layers)
        xgrid = np.linspace(-1, 1, m)
        ygrid = np.linspace(1, -1, n)
        x0, x1 = np.meshgrid(xgrid, ygrid)
        self.x = torch.from_numpy(np.stack([x0.ravel(), x1.ravel()], 1)).float()  # (n*m, 2), not a buffer (as in the reference)

    def forward(self, z):
        if z.dim() < 2:
            z = z.unsqueeze(0)
        if self.x.device != z.device:
            self.x = self.x.to(z.device)
        # the coordinate grid is the same for every sample: one [1,512] x [512,1024] product per call
        h_x = self.coord_linear(self.x.reshape(1, self.x_dim))
        h_z = self.latent_linear(z)
        h = TileAdd.apply(h_x, h_z, self.num_coords)
        return self.layers(h)
