"""TEST INFRASTRUCTURE ONLY (imported by tests/): CPU restatement of taming.modules.losses.lpips.LPIPS.

PARITY UNPINNED: taming-transformers @ 049ce3ce (the reference's `.SUBMODULES.json:44-50` pin) is not in /root/reference
and its `vgg.pth` cannot be downloaded here, so this follows the public definition and is checked only
against itself: ScalingLayer -> torchvision VGG16 `features` sliced at relu1_2 / relu2_2 / relu3_3 / relu4_3 / relu5_3 ->
normalize_tensor (x / (sqrt(sum_c x^2) + 1e-10)) -> squared difference -> NetLinLayer (Dropout in eval = identity, 1x1
conv without bias) -> spatial mean -> sum over levels -> [B,1,1,1]. Call site in the reference:
src/modules/losses/contperceptual.py:141-145. Parameter names match taming's state_dict (`net.slice1.0.weight`,
`lin0.model.1.weight`, `scaling_layer.shift`), so a real LPIPS checkpoint loads into it.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

# (slice name, [(features index, Cin, Cout), ...]); slices 2-5 start with MaxPool2d(2, 2)
VGG_SLICES = [("slice1", [(0, 3, 64), (2, 64, 64)]),
              ("slice2", [(5, 64, 128), (7, 128, 128)]),
              ("slice3", [(10, 128, 256), (12, 256, 256), (14, 256, 256)]),
              ("slice4", [(17, 256, 512), (19, 512, 512), (21, 512, 512)]),
              ("slice5", [(24, 512, 512), (26, 512, 512), (28, 512, 512)])]
CHNS = [64, 128, 256, 512, 512]
SHIFT = [-.030, -.088, -.188]
SCALE = [.458, .448, .450]


class LPIPSOracle(nn.Module):
    def __init__(self):
        super().__init__()
        self.scaling_layer = nn.Module()
        self.scaling_layer.register_buffer("shift", torch.tensor(SHIFT)[None, :, None, None])
        self.scaling_layer.register_buffer("scale", torch.tensor(SCALE)[None, :, None, None])
        self.net = nn.Module()
        for name, convs in VGG_SLICES:
            self.net.add_module(name, nn.ModuleDict({str(i): nn.Conv2d(ci, co, 3, padding=1) for i, ci, co in convs}))
        for k, c in enumerate(CHNS):
            lin = nn.Module()
            lin.model = nn.ModuleDict({"1": nn.Conv2d(c, 1, 1, bias=False)})
            self.add_module("lin%d" % k, lin)
        for p in self.parameters():
            p.requires_grad = False

    def features(self, x):
        outs = []
        for k, (name, convs) in enumerate(VGG_SLICES):
            if k > 0:
                x = F.max_pool2d(x, 2, 2)
            for i, _, _ in convs:
                x = F.relu(getattr(self.net, name)[str(i)](x))
            outs.append(x)
        return outs

    def forward(self, input, target):
        sl = self.scaling_layer
        f0 = self.features((input - sl.shift) / sl.scale)
        f1 = self.features((target - sl.shift) / sl.scale)
        val = 0
        for k in range(len(CHNS)):
            n0 = f0[k] / (torch.sqrt(torch.sum(f0[k] ** 2, dim=1, keepdim=True)) + 1e-10)
            n1 = f1[k] / (torch.sqrt(torch.sum(f1[k] ** 2, dim=1, keepdim=True)) + 1e-10)
            val = val + getattr(self, "lin%d" % k).model["1"]((n0 - n1) ** 2).mean([2, 3], keepdim=True)
        return val


def fill_lpips_state_(sd, seed=31):
    """Deterministic stand-in weights (no vgg.pth offline): He-scaled conv weights so activations keep their scale through
    the 13 ReLU layers, small biases, non-negative lin weights as in the trained LPIPS heads."""
    g = torch.Generator().manual_seed(seed)
    for k, v in sd.items():
        if k.startswith("scaling_layer"):
            continue
        if k.startswith("lin"):
            v.copy_(torch.rand(v.shape, generator=g) / v.shape[1])
        elif k.endswith(".weight"):
            fan_in = v.shape[1] * 9
            v.copy_(torch.randn(v.shape, generator=g) * (2.0 / fan_in) ** 0.5)
        else:
            v.copy_(0.05 * torch.randn(v.shape, generator=g))
    return sd
