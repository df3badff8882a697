"""Times encoder.conv_in (3 -> 128) forward + backward and decoder.conv_out (128 -> 3) backward at the bench shape through the
generic 64-channel-padded path and through the thin path (im2col + one K = 64 step):  python tools/thin_probe.py [B]"""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "generative-detection_b200"))
from odvae_b200 import modules as M, ops  # noqa: E402
from odvae_b200.lib import CONV_3x3_S1  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
H = W = 256
dtype = torch.float16
fmt = ops.format_for(dtype)
ops.OVERLAP_WGRAD[0] = False
g = torch.Generator().manual_seed(0)


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


cin = M.Conv2d(3, 128, 3, CONV_3x3_S1).cuda()
cout = M.Conv2d(128, 3, 3, CONV_3x3_S1).cuda()
ops.assign_format(cin, dtype)
ops.assign_format(cout, dtype)
x = (torch.rand(B, 3, H, W, generator=g) * 2 - 1).cuda()
dy = torch.randn(B, H, W, 128, generator=g).cuda().half()
xs = torch.randn(B, H, W, 128, generator=g).cuda().half()
dyo = torch.randn(B, 3, H, W, generator=g).cuda()

for thin in (False, True, False, True):
    ops.THIN_CONV = thin
    if thin:
        f_in = lambda: ops.ThinConvIn.apply(x, cin.weight, cin.bias, cin.spec)
    else:
        f_in = lambda: cin(ops.ToNHWC.apply(x, 64, dtype, None), need_dx=False)
    t_f = timed(f_in)
    y = f_in()
    t_b = timed(lambda: torch.autograd.backward(y, dy, retain_graph=True))
    xh = xs.clone().requires_grad_(True)
    yo = cout(xh, out_nchw=True)
    t_fo = timed(lambda: cout(xh, out_nchw=True))
    t_bo = timed(lambda: torch.autograd.backward(yo, dyo, retain_graph=True))
    print("thin=%d  conv_in fwd %.3f ms  bwd %.3f ms | conv_out fwd %.3f ms  bwd (dgrad + wgrad) %.3f ms" % (thin, t_f, t_b, t_fo, t_bo),
          flush=True)
