"""GroupNorm backward of one shape alone (default 128 ch @256x256, batch 64, fp16 x / bf16 dy as in the timed mode):
CUDA-event timing per call and achieved algorithmic GB/s (3 E s [+ E s with dx_add]); target for an ncu DRAM-byte capture.
Env: PB/PH/PC shape, PADD=1 adds dx_add; the ODVAE_GN_* switches of csrc/groupnorm.cu select the path."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "generative-detection_b200"), ROOT]
import torch
from odvae_b200 import ops

B, H, C = int(os.environ.get("PB", 64)), int(os.environ.get("PH", 256)), int(os.environ.get("PC", 128))
with_add = int(os.environ.get("PADD", 0))
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 5
torch.manual_seed(0)
x = torch.randn(B, H, H, C, device="cuda").half()
dy = torch.randn(B, H, H, C, device="cuda").bfloat16()
add = torch.randn(B, H, H, C, device="cuda").bfloat16() if with_add else None
gamma = 1 + 0.1 * torch.randn(C, device="cuda")
beta = 0.1 * torch.randn(C, device="cuda")
stats = ops.gn_stats_raw(x, C, code=ops.F16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
E_bytes = B * H * H * C * 2
alg = (3 + with_add) * E_bytes
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for it in range(iters):
    flush.zero_()
    ev[0].record()
    dx, dg, db = ops.gn_bwd_raw(x, stats, gamma, beta, dy, True, dx_add=add, x_code=ops.F16)
    ev[1].record()
    torch.cuda.synchronize()
    t = ev[0].elapsed_time(ev[1])
    print("iter %d: gn_bwd B=%d %dx%d C=%d add=%d  %.3f ms  %.0f GB/s algorithmic (%.2f of 6457)" % (
        it, B, H, H, C, with_add, t, alg / t / 1e6, alg / t / 1e6 / 6457), flush=True)

if os.environ.get("PTIMING"):
    import ctypes
    from odvae_b200 import lib as L
    dll = ctypes.CDLL(L.LIB_PATH)
    buf = (ctypes.c_ulonglong * (2 * 16 * 8))()
    if hasattr(dll, "odvae_debug_gn_team_timing") and dll.odvae_debug_gn_team_timing(buf) == 0:
        names = ["p1 loop", "reduce+store", "arrive A", "duty", "wait B", "p2 loop"]
        for slot, who in enumerate(["cta 0 (duty)", "cta 100"]):
            print(who)
            for it in range(1, 9):
                st = [buf[(slot * 16 + it) * 8 + k] for k in range(7)]
                prev_end = buf[(slot * 16 + it - 1) * 8 + 6]
                print("  item %2d: gap %5.1f us | " % (it, (st[0] - prev_end) / 1e3) + "  ".join("%s %5.1f" % (names[k], (st[k + 1] - st[k]) / 1e3) for k in range(6)))
