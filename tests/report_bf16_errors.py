"""Prints the 16-bit-mode error table against the goldens (diagnostic; run on the GPU box).
    python tests/report_bf16_errors.py [bf16|mixed]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import conftest  # noqa
import torch
import helpers as H
from test_parity_gpu import run_product

MODE = torch.float16 if (len(sys.argv) > 1 and sys.argv[1] == "mixed") else torch.bfloat16
for name in H.CASES:
    g, model, loss, out = run_product(name, MODE)
    dec_obj, dec_pose, post, bbox = out
    row = {"loss": H.rel_scalar(loss, g["loss"]), "dec_pose": H.rel_max(dec_pose, g["out"]["dec_pose"]),
           "post_mean": H.rel_max(post.mean, g["out"]["posterior_mean"]), "post_mean_l2": H.rel_l2(post.mean, g["out"]["posterior_mean"]),
           "post_logvar": H.rel_max(post.logvar, g["out"]["posterior_logvar"])}
    if g["out"]["dec_obj"].abs().max() > 0:
        row["dec_obj_l2"] = H.rel_l2(dec_obj, g["out"]["dec_obj"]); row["dec_obj_max"] = H.rel_max(dec_obj, g["out"]["dec_obj"])
    logs = {k: H.rel_scalar(model.logged[k], v, 1e-6) for k, v in g["log"].items() if not k.endswith("g_loss")}
    row["worst_log"] = max(logs.items(), key=lambda kv: kv[1])
    errs = H.grad_errors({k: p.grad for k, p in model.named_parameters()}, g["grads"], 1e-1)
    l2s = sorted(((v[0], k) for k, v in errs.items()), reverse=True)
    row["grad_l2_p97"] = l2s[int(0.03 * len(l2s))][0]
    ss = sorted(((v[1], k) for k, v in errs.items()), reverse=True)
    import statistics
    row["grad_l2_worst"] = l2s[:3]; row["grad_l2_median"] = statistics.median(v for v, _ in l2s)
    row["grad_sample_worst"] = ss[:3]; row["grad_sample_median"] = statistics.median(v for v, _ in ss)
    print(name, row, flush=True)
