"""Runs fwd / dgrad / wgrad of one conv shape alone (default: the 128->128 3x3 at 256x256, batch 64): CUDA-event timing
and a small `ncu --set full` target (engine launches per iteration: fwd, dgrad, wgrad)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "generative-detection_b200"), ROOT]
import torch
from odvae_b200 import kernels as K
from odvae_b200.lib import CONV_3x3_S1, IMPL_TC

B, H, C = int(os.environ.get("PB", 64)), int(os.environ.get("PH", 256)), int(os.environ.get("PC", 128))
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 4
dt = torch.float16 if os.environ.get("PDT", "fp16") == "fp16" else torch.bfloat16  # fp16 = the timed mode
torch.manual_seed(0)
x = torch.randn(B, H, H, C, device="cuda").to(dt)
dy = torch.randn(B, H, H, C, device="cuda").to(dt)
res = torch.randn(B, H, H, C, device="cuda").to(dt)
w = torch.randn(C, C, 3, 3, device="cuda") / (C * 9) ** 0.5
b = torch.randn(C, device="cuda")
wp = K.pack_conv_weight(w, CONV_3x3_S1, C, C, 0, dt)
wd = K.pack_conv_weight(w, CONV_3x3_S1, C, C, 1, dt)
fl = 2.0 * B * H * H * C * C * 9
from odvae_b200 import lib
G = 32
gn_bytes = lib.call("conv2d_fwd_gnstats_workspace_bytes", B, H, H, C, C, CONV_3x3_S1, G)
gn_ws = torch.empty(max(gn_bytes, 1), dtype=torch.uint8, device="cuda")
y_gn = torch.empty(B, H, H, C, dtype=dt, device="cuda")
stats = torch.empty(B, G, 2, device="cuda")
code = K.act_code(dt)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
for it in range(iters):
    ev[0].record()
    y = K.conv_fwd(x, wp, b, C, CONV_3x3_S1, IMPL_TC)
    ev[1].record()
    y2 = K.conv_fwd(x, wp, b, C, CONV_3x3_S1, IMPL_TC, residual=res)
    ev[2].record()
    dx = K.conv_dgrad(dy, wd, C, CONV_3x3_S1, IMPL_TC, (H, H))
    ev[3].record()
    dw, db = K.conv_wgrad(x, dy, C, C, CONV_3x3_S1, IMPL_TC)
    ev[4].record()
    if gn_bytes:
        lib.call("conv2d_fwd_gnstats", x, wp, b, None, y_gn, stats, B, H, H, C, C, CONV_3x3_S1, G, 1e-6, code, code, gn_ws, gn_bytes)
    ev[5].record()
    torch.cuda.synchronize()
    t = [ev[i].elapsed_time(ev[i + 1]) for i in range(5)]
    print("        fwd+gnstats(+finalize) %.3f ms (%.0f TF/s)" % (t[4], fl / t[4] / 1e9))
    print("iter %d: fwd %.3f ms (%.0f TF/s)  fwd+res %.3f ms (%.0f)  dgrad %.3f ms (%.0f)  wgrad(+reduce,colsum) %.3f ms (%.0f)" % (
        it, t[0], fl / t[0] / 1e9, t[1], fl / t[1] / 1e9, t[2], fl / t[2] / 1e9, t[3], fl / t[3] / 1e9), flush=True)
