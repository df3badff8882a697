"""Times the attention score kernels of one AttnBlock at the bench shape (B x 4096 x 4096, C = 256) on the B200:
    python tools/attn_probe.py [B] [fp16|bf16]
q k^T + softmax (2 passes), dO v^T + softmax backward, and the plain P v GEMM, CUDA-event timed over 20 calls after 3."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "generative-detection_b200"))
from odvae_b200 import ops  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dt = torch.float16 if (len(sys.argv) < 3 or sys.argv[2] == "fp16") else torch.bfloat16
N, C = 4096, 256
torch.manual_seed(0)
qkv = (torch.randn(B, N, 3 * C, device="cuda") * 0.5).to(dt)
q, k, v = qkv[..., :C], qkv[..., C:2 * C], qkv[..., 2 * C:]
P = torch.empty(B, N, N, dtype=dt, device="cuda")
dS = torch.empty(B, N, N, dtype=dt, device="cuda")
dO = torch.randn(B, N, C, device="cuda").to(dt)
delta = torch.zeros(B * N, dtype=torch.float32, device="cuda")
ws = torch.empty(ops.call("attn_softmax_workspace_bytes", B, N), dtype=torch.uint8, device="cuda")
code = ops.act_code(dt)
scale = C ** -0.5


def timed(fn, n=20):
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


t1 = timed(lambda: ops.call("attn_scores_softmax", q, k, P, B, N, C, 3 * C, 3 * C, N * 3 * C, N * 3 * C, scale, code, ws,
                            ws.numel()))
t2 = timed(lambda: ops.call("attn_dsoftmax", dO, v, P, delta, dS, B, N, C, C, 3 * C, N * C, N * 3 * C, scale, code))
fl = 2.0 * B * N * N * C
print("B %d %s bstat=%s: qk^T+softmax (2 passes) %.3f ms = %.0f TF/s ; dO v^T+dsoftmax %.3f ms = %.0f TF/s (P+dS traffic %.0f GB/s)"
      % (B, dt, os.environ.get("ODVAE_GEMM_BSTAT", "1"), t1, 2 * fl / t1 * 1e-9, t2, fl / t2 * 1e-9,
         2.0 * B * N * N * 2 / t2 * 1e-6))

# ---- the plain GEMMs that stream P / dS from HBM: O = P v, dv = P^T dO, dq = dS k, dk = dS^T q ----
O = torch.empty(B, N, C, dtype=dt, device="cuda")
dqkv = torch.empty(B, N, 3 * C, dtype=dt, device="cuda")
t_pv = timed(lambda: ops.gemm_raw(P, v, O, N, C, N, B, N, 3 * C, C, N * N, N * 3 * C, N * C, False, True, code=code))
t_dv = timed(lambda: ops.gemm_raw(P, dO, dqkv[..., 2 * C:], N, C, N, B, N, C, 3 * C, N * N, N * C, N * 3 * C, True, True, code=code))
t_dq = timed(lambda: ops.gemm_raw(dS, k, dqkv[..., :C], N, C, N, B, N, 3 * C, 3 * C, N * N, N * 3 * C, N * 3 * C, False, True, code=code))
t_dk = timed(lambda: ops.gemm_raw(dS, q, dqkv[..., C:2 * C], N, C, N, B, N, 3 * C, 3 * C, N * N, N * 3 * C, N * 3 * C, True, True, code=code))
gb = B * N * N * 2 * 1e-9
print("prefetch=%s: P v %.3f ms (%.0f TF/s, %.2f TB/s) | P^T dO %.3f | dS k %.3f | dS^T q %.3f ms"
      % (os.environ.get("ODVAE_GEMM_PREFETCH", "8"), t_pv, fl / t_pv * 1e-9, gb / t_pv, t_dv, t_dq, t_dk))
