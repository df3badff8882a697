"""The attention blocks' 1x1 convs alone (fused q/k/v 256->768 and proj_out 256->256 + residual at 64x64, batch 64):
CUDA-event timing and a small `ncu --set full` target (two engine launches per iteration)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "generative-detection_b200"), ROOT]
import torch
from odvae_b200 import kernels as K
from odvae_b200.lib import CONV_1x1, IMPL_TC

B, H, C = int(os.environ.get("PB", 64)), int(os.environ.get("PH", 64)), int(os.environ.get("PC", 256))
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 4
dt = torch.bfloat16
torch.manual_seed(0)
x = torch.randn(B, H, H, C, device="cuda").to(dt)
res = torch.randn(B, H, H, C, device="cuda").to(dt)
w3 = torch.randn(3 * C, C, 1, 1, device="cuda") / C ** 0.5
w1 = torch.randn(C, C, 1, 1, device="cuda") / C ** 0.5
b3, b1 = torch.randn(3 * C, device="cuda"), torch.randn(C, device="cuda")
wp3 = K.pack_conv_weight(w3, CONV_1x1, 3 * C, C, 0, dt)
wp1 = K.pack_conv_weight(w1, CONV_1x1, C, C, 0, dt)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
px = B * H * H
for it in range(iters):
    ev[0].record()
    y3 = K.conv_fwd(x, wp3, b3, 3 * C, CONV_1x1, IMPL_TC)
    ev[1].record()
    y1 = K.conv_fwd(x, wp1, b1, C, CONV_1x1, IMPL_TC, residual=res)
    ev[2].record()
    torch.cuda.synchronize()
    t = [ev[i].elapsed_time(ev[i + 1]) for i in range(2)]
    gb3, gb1 = px * C * 2 * 4 / 1e9, px * C * 2 * 3 / 1e9
    print("iter %d: qkv %.3f ms (%.0f TF/s, %.0f GB/s algorithmic)  proj+res %.3f ms (%.0f TF/s, %.0f GB/s)" % (
        it, t[0], 2.0 * px * C * 3 * C / t[0] / 1e9, gb3 / t[0] * 1e3, t[1], 2.0 * px * C * C / t[1] / 1e9, gb1 / t[1] * 1e3), flush=True)
ref = torch.nn.functional.conv2d(x.permute(0, 3, 1, 2).float(), w3.to(dt).float(), b3).permute(0, 2, 3, 1)
print("qkv max err", (y3.float() - ref).abs().max().item())
