"""Per-tensor gradient errors of the product against the reference gradients stored in the goldens (diagnostic; GPU box).
    python tests/report_grad_errors.py [fp32|fp16|mixed|bf16 ...] > profiles/rNN_grad_errors.txt
For every golden case and mode: outputs (relative L2), loss, and the distribution over parameter tensors of
||g - g_ref||_2 / ||g_ref||_2 and max|g - g_ref| / max|g_ref| (tests/helpers.py grad_diff_errors)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa: F401,E402
import torch  # noqa: E402

import helpers as H  # noqa: E402
from test_parity_gpu import run_product  # noqa: E402

MODES = {"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16, "mixed": "mixed"}
names = sys.argv[1:] or ["fp32", "fp16", "mixed", "bf16"]
for mode in names:
    for name in H.CASES:
        g, model, loss, out = run_product(name, MODES[mode])
        dec_obj, dec_pose, post, bbox = out
        head = "%s %s: loss rel %.3g  dec_pose %.3g  post_mean l2 %.3g" % (
            mode, name, H.rel_scalar(loss, g["loss"]), H.rel_max(dec_pose, g["out"]["dec_pose"]),
            H.rel_l2(post.mean, g["out"]["posterior_mean"]))
        if g["out"]["dec_obj"].abs().max() > 0:
            head += "  dec_obj l2 %.3g max %.3g" % (H.rel_l2(dec_obj, g["out"]["dec_obj"]),
                                                    H.rel_max(dec_obj, g["out"]["dec_obj"]))
        errs = H.grad_diff_errors({k: p.grad for k, p in model.named_parameters()}, g["grads"])
        print(head)
        print("  " + H.error_table(errs, 6), flush=True)
        del model, out, loss
        torch.cuda.empty_cache()
