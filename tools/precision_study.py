"""Scratch study (CPU, oracle-based; NOT product code): where does 16-bit rounding error enter the forward pass?
Monkeypatches the oracle's building blocks to round at the points the CUDA path rounds and reports relative L2
error of latents / reconstruction against the fp32 oracle.   python tools/precision_study.py [small|full]
"""
import os, sys, pickle, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "generative-detection_b200"), ROOT]
import torch
import torch.nn.functional as F
from odvae_b200 import synth
from odvae_b200.autoencoder import PoseAutoencoder
from oracle import odvae_oracle as O

which = sys.argv[1] if len(sys.argv) > 1 else "small"
dd = synth.DDCONFIG_SMALL if which == "small" else synth.DDCONFIG
R, B = (64, 2) if which == "small" else (256, 1)
path = os.path.join(tempfile.mkdtemp(), "all.pkl")
pickle.dump(synth.synthetic_dataset_stats(23), open(path, "wb"))
params = synth.model_config(dd, path)
shell = PoseAutoencoder(**params)
synth.fill_state_dict_(shell.state_dict(), seed=23)
sd = {k: v.detach().clone() for k, v in shell.state_dict().items() if torch.is_floating_point(v)}
orc = O.Oracle(params, sd, synth.synthetic_dataset_stats(23), synth.LABELS)
batch = synth.synthetic_batch(B, R, seed=31)
step = 80000
p = O.dropout_prob(step, params, params["lossconfig"]["params"]["encoder_pretrain_steps"])
noise = synth.draw_noise(B, p, seed=77)
x = O.rescale(batch["patch"].float())

orig = dict(conv=O.conv, group_norm=O.group_norm, resnet_block=O.resnet_block, attn_block=O.attn_block, swish=O.swish)

def rnd(t, dt):
    return t if dt is None else t.to(dt).float()

def run(act=None, wt=None, out=None, resid=None, gn=None, attn=None, attn_all=None):
    """act: dtype conv inputs are rounded to; wt: weights; out: conv outputs; resid: block outputs; gn: GN+SiLU output"""
    def conv(sd_, pre, x_, stride=1, padding=0):
        w = rnd(sd_[pre + ".weight"], wt)
        y = F.conv2d(rnd(x_, act), w, sd_[pre + ".bias"], stride=stride, padding=padding)
        return rnd(y, out)
    def resnet_block(sd_, pre, x_):
        h = conv(sd_, pre + ".conv1", rnd(O.swish(O.group_norm(sd_, pre + ".norm1", x_)), gn), padding=1)
        h = F.conv2d(rnd(rnd(O.swish(O.group_norm(sd_, pre + ".norm2", h)), gn), act), rnd(sd_[pre + ".conv2.weight"], wt),
                     sd_[pre + ".conv2.bias"], padding=1)  # conv2 output stays fp32 until the residual add
        if (pre + ".nin_shortcut.weight") in sd_:
            x_ = F.conv2d(rnd(x_, act), rnd(sd_[pre + ".nin_shortcut.weight"], wt), sd_[pre + ".nin_shortcut.bias"])
        return rnd(x_ + h, resid)
    def attn_block(sd_, pre, x_):
        if attn_all is not None:  # the whole block in one format (operands, weights, P, O); residual add in `resid`
            a = attn_all
            h = rnd(O.group_norm(sd_, pre + ".norm", x_), a)
            cv = lambda n, t: rnd(F.conv2d(rnd(t, a), rnd(sd_[pre + n + ".weight"], a), sd_[pre + n + ".bias"]), a)
            q, k, v = cv(".q", h), cv(".k", h), cv(".v", h)
            b, c, hh, ww = q.shape
            q = q.reshape(b, c, hh * ww).permute(0, 2, 1)
            k = k.reshape(b, c, hh * ww)
            w = rnd(torch.softmax(torch.bmm(q, k) * (int(c) ** (-0.5)), dim=2), a)
            v = v.reshape(b, c, hh * ww)
            h = rnd(torch.bmm(v, w.permute(0, 2, 1)).reshape(b, c, hh, ww), a)
            y = F.conv2d(h, rnd(sd_[pre + ".proj_out.weight"], a), sd_[pre + ".proj_out.bias"])
            return rnd(x_ + y, resid)
        h = rnd(O.group_norm(sd_, pre + ".norm", x_), gn)
        q, k, v = conv(sd_, pre + ".q", h), conv(sd_, pre + ".k", h), conv(sd_, pre + ".v", h)
        b, c, hh, ww = q.shape
        q = q.reshape(b, c, hh * ww).permute(0, 2, 1)
        k = k.reshape(b, c, hh * ww)
        w = rnd(torch.softmax(torch.bmm(q, k) * (int(c) ** (-0.5)), dim=2), attn)
        v = v.reshape(b, c, hh * ww)
        h = rnd(torch.bmm(v, w.permute(0, 2, 1)).reshape(b, c, hh, ww), out)
        y = F.conv2d(rnd(h, act), rnd(sd_[pre + ".proj_out.weight"], wt), sd_[pre + ".proj_out.bias"])
        return rnd(x_ + y, resid)
    O.conv, O.resnet_block, O.attn_block = conv, resnet_block, attn_block
    try:
        with torch.no_grad():
            dec_obj, dec_pose, post, bbox_post, _ = orc.forward(x, step, noise)
    finally:
        O.conv, O.resnet_block, O.attn_block = orig["conv"], orig["resnet_block"], orig["attn_block"]
    return dec_obj, post.mean, post.logvar

def rel(a, b):
    return float((a - b).norm() / b.norm())

ref = run()
bf, hf = torch.bfloat16, torch.float16
cases = {
    "all bf16 (current)": dict(act=bf, wt=bf, out=bf, resid=bf, gn=bf, attn=bf),
    "all fp16": dict(act=hf, wt=hf, out=hf, resid=hf, gn=hf, attn=hf),
    "bf16, fp32 residual stream": dict(act=bf, wt=bf, out=bf, resid=None, gn=bf, attn=bf),
    "bf16 but weights exact": dict(act=bf, wt=None, out=bf, resid=bf, gn=bf, attn=bf),
    "only weights bf16": dict(wt=bf),
    "only residual bf16": dict(resid=bf),
    "only gn-out/act bf16": dict(act=bf, gn=bf),
    "only conv-out bf16": dict(out=bf),
    "all fp16, attention blocks bf16": dict(act=hf, wt=hf, out=hf, resid=hf, gn=hf, attn=hf, attn_all=bf),
    "all fp16, only P bf16": dict(act=hf, wt=hf, out=hf, resid=hf, gn=hf, attn=bf),
    "fp16 acts, bf16 weights": dict(act=hf, wt=bf, out=hf, resid=hf, gn=hf, attn=hf),
    "fp16 weights, bf16 acts": dict(act=bf, wt=hf, out=bf, resid=bf, gn=bf, attn=bf),
}
for name, kw in cases.items():
    d, m, lv = run(**kw)
    print("%-32s recon %.3e  mean %.3e  logvar %.3e" % (name, rel(d, ref[0]), rel(m, ref[1]), rel(lv, ref[2])))
