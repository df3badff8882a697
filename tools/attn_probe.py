"""Runs the fused attention core kernels alone at the bench shape (B=64, N=4096, C=256): timing with CUDA events, and
a small target for `ncu --set full` (3 engine launches per iteration: row-stats pass, softmax pass, dsoftmax)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "generative-detection_b200"), ROOT]
import torch
from odvae_b200 import ops

B, N, C = int(os.environ.get("PB", 64)), 4096, 256
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5
dt = torch.bfloat16
torch.manual_seed(0)
qkv = torch.randn(B, N, 3 * C, device="cuda").to(dt)
q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
P = torch.empty(B, N, N, dtype=dt, device="cuda")
dS = torch.empty(B, N, N, dtype=dt, device="cuda")
dO = torch.randn(B, N, C, device="cuda").to(dt)
O = torch.randn(B, N, C, device="cuda").to(dt)
delta = torch.empty(B * N, dtype=torch.float32, device="cuda")
ws = torch.empty(ops.call("attn_softmax_workspace_bytes", B, N), dtype=torch.uint8, device="cuda")
scale = C ** -0.5
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
for it in range(iters):
    ev[0].record()
    ops.call("attn_scores_softmax", q, k, P, B, N, C, 3 * C, 3 * C, N * 3 * C, N * 3 * C, scale, ops.act_code(dt), ws, ws.numel())
    ev[1].record()
    ops.call("rowdot", dO, O, delta, B * N, C, C, C, ops.act_code(dt))
    ev[2].record()
    ops.call("attn_dsoftmax", dO, v, P, delta, dS, B, N, C, C, 3 * C, N * C, N * 3 * C, scale, ops.act_code(dt))
    ev[3].record()
    torch.cuda.synchronize()
    print("iter %d: scores+softmax %.3f ms  rowdot %.3f ms  dsoftmax %.3f ms" % (
        it, ev[0].elapsed_time(ev[1]), ev[1].elapsed_time(ev[2]), ev[2].elapsed_time(ev[3])), flush=True)
