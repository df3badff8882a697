"""GroupNorm kernels alone, fp16 activations / fp16 gradients (the timed mode): forward statistics + apply, backward (reduce +
finalize + apply) with and without dx_add, per shape, CUDA-event timed with an L2 flush between calls; achieved ALGORITHMIC
GB/s (2 E s forward, 3 E s [+ E s] backward). ODVAE_LIB_PATH selects the library build (tools/gpu_gn_variants.sh)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "generative-detection_b200"), ROOT]
import torch
from odvae_b200 import ops

shapes = [(64, 256, 128), (64, 128, 128), (64, 64, 256), (64, 16, 512)]
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
tot = {"fwd": 0.0, "bwd": 0.0}
for B, H, C in shapes:
    torch.manual_seed(0)
    x = torch.randn(B, H, H, C, device="cuda").half()
    dy = torch.randn(B, H, H, C, device="cuda").half()
    add = torch.randn(B, H, H, C, device="cuda").half()
    gamma = 1 + 0.1 * torch.randn(C, device="cuda")
    beta = 0.1 * torch.randn(C, device="cuda")
    E = B * H * H * C * 2
    def timed(fn):
        best = 1e9
        for _ in range(iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best
    stats = ops.gn_stats_raw(x, C, code=ops.F16)
    t_stats = timed(lambda: ops.gn_stats_raw(x, C, code=ops.F16))
    t_apply = timed(lambda: ops.gn_apply_raw(x, stats, gamma, beta, True, ops.F16, ops.F16))
    t_bwd = timed(lambda: ops.gn_bwd_raw(x, stats, gamma, beta, dy, True, x_code=ops.F16, g_code=ops.F16))
    t_bwd_add = timed(lambda: ops.gn_bwd_raw(x, stats, gamma, beta, dy, True, dx_add=add, x_code=ops.F16, g_code=ops.F16))
    print("B%d %dx%d C%d: stats %.3f ms (%.0f GB/s) | apply %.3f ms (%.0f GB/s alg) | bwd %.3f ms (%.0f GB/s alg) | bwd+add %.3f ms (%.0f GB/s alg)" % (
        B, H, H, C, t_stats, E / t_stats / 1e6, t_apply, 2 * E / t_apply / 1e6, t_bwd, 3 * E / t_bwd / 1e6, t_bwd_add, 4 * E / t_bwd_add / 1e6), flush=True)
