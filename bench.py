#!/usr/bin/env python
"""bench.py -- AutoencoderKL (OD-VAE PoseAutoencoder) train-step throughput on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 64] [--impl ours|reference]

One "step" = the reference's optimizer_idx-0 training step (PoseAutoencoder.training_step -> PoseLoss ->
backward -> gradient allreduce -> clip 1.0 -> Adam) on one synthetic object-crop batch of the
autoencoder_kl_16x16x16 pose config at 256x256, per-GPU batch 64 (BASELINE.json configs[1]), global_step 80000
(decoder, pixel loss and dropout schedule active), perceptual_weight = 0 and disc_factor = 0 (SURVEY 8d config 2:
a REDUCED objective -- the yaml's own defaults are perceptual_weight = 1 and disc_factor = 1; that step, with both
optimizer steps per batch, is measured next to it as `extra.full_objective`, or alone with `--objective full`).
Prints ONE JSON line (rank 0). At N = 1 the line also carries `extra`: the full-objective step, the encode-only sweep
(configs[2]), the 512x512 reconstruction step (configs[4]) and `gpu_eager_baseline` (the oracle restatement of the
reference's PyTorch code run in eager mode on the same B200: cuDNN / cuBLAS with TF32 and with bf16 autocast).
`--impl reference` times the CPU restatement of the reference's own PyTorch path (oracle/) on the host cores instead
(the reference tree itself is Python under /root/reference and does not travel to the GPU box).
"""
import argparse
import json
import os
import pickle
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "generative-detection_b200"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

STEP = 80000
RES = 256
FLOPS_TRAIN_PER_IMG = 1449.4e9  # fwd 483.30 GF (conv3x3 383.64 + conv1x1 13.49 + attn 86.17), x3 - 0.45


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=64, help="per-GPU batch (weak scaling)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-baseline", type=int, default=1, help="time the CPU oracle on rank 0 at N=1 (0 = skip)")
    ap.add_argument("--no-roofline", action="store_true")
    ap.add_argument("--mode", default="fp16", choices=["fp16", "mixed", "bf16"],
                    help="16-bit compute mode: fp16 = every 16-bit tensor IEEE fp16, activation gradients with a static "
                         "scale (meets the 1e-2 parity tolerance on outputs AND gradients; default); mixed = round 1's fp16 "
                         "forward + bf16 attention/gradients; bf16 = everything bf16")
    ap.add_argument("--objective", default="configs1", choices=["configs1", "full"],
                    help="configs1 = BASELINE.json configs[1] (perceptual_weight 0, disc_factor 0; the bench line); full = the "
                         "yaml's own objective: LPIPS + PatchGAN generator term with the adaptive weight in the optimizer_idx 0 "
                         "step AND the optimizer_idx 1 discriminator step, i.e. one whole batch of the reference loop")
    ap.add_argument("--step-graph", default="auto", choices=["auto", "on", "off"],
                    help="train workload: replay the whole step (fwd + loss + bwd + clip + Adam + weight re-pack, ~1000 "
                         "launches) as ONE CUDA graph (odvae_b200/graphs.py GraphedTrainStep; bit-identical to the eager launch "
                         "sequence; host enqueue 3 ms instead of ~125 ms per step). auto = off: measured on one box (profiles/"
                         "README.md r02 A/B) the eager two-stream launch sequence overlaps the weight-gradient GEMMs with the "
                         "GroupNorm backward better than the replayed graph does (139.4-140.3 vs 142.6 ms); the graph step is "
                         "reported next to the bench line as extra.cuda_graph_step")
    ap.add_argument("--extras", type=int, default=1,
                    help="N = 1 only: also measure the full objective, the encode sweep, the 512x512 step and the eager "
                         "cuDNN/cuBLAS baseline, reported under `extra` (0 = skip)")
    ap.add_argument("--ncu-step", action="store_true",
                    help="profiling aid: after warm-up run ONE step between cudaProfilerStart/Stop and exit "
                         "(use with `ncu --profile-from-start off`); prints no bench line")
    ap.add_argument("--shapes", default="", help="write the per-shape engine timing table (JSON) to this path")
    ap.add_argument("--perceptual-weight", type=float, default=0.0,
                    help="train workload: PoseLoss perceptual_weight (the yaml's default is 1.0 = LPIPS on; the bench line "
                         "keeps 0 as BASELINE.json configs[1] states, this flag measures the step with the LPIPS towers)")
    ap.add_argument("--graph", default="auto", choices=["auto", "on", "off"],
                    help="encode workload: replay the launch sequence as a CUDA graph (auto = batch <= 32)")
    ap.add_argument("--workload", default="train", choices=["train", "encode", "recon"],
                    help="train = BASELINE.json configs[1] (default, the bench line); encode = configs[2], encode-only "
                         "inference (Encoder + posterior.mode(), no_grad) at --batch; recon = configs[4]: encode -> mode "
                         "-> decode -> L1 -> backward through decoder and encoder (no pose heads: the config's pose MLPs "
                         "are sized for 16x16 latents, i.e. 256x256 crops only, in the reference too), use with --res 512")
    ap.add_argument("--res", type=int, default=RES, help="crop resolution (configs[4]: 512)")
    return ap.parse_args()


def stats_file():
    from odvae_b200 import synth
    path = os.path.join(tempfile.mkdtemp(prefix="odvae_stats_"), "all.pkl")
    with open(path, "wb") as f:
        pickle.dump(synth.synthetic_dataset_stats(23), f)
    return path


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


# --------------------------------------------------------------------------------------------------
# clocks sampling during the timed region
# --------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------------
# Reference arms: the oracle restatement of the reference's PyTorch path (oracle/odvae_oracle.py, pinned to goldens made
# from the unmodified reference) -- on the host cores (`cpu_baseline`, `--impl reference`) and, as the "existing Blackwell
# kernels on the same box" comparator of SURVEY 8d, in eager mode on the GPU (`extra.gpu_eager_baseline`: ATen / cuDNN /
# cuBLAS). Same step definition as the product arm: fwd + loss + bwd + clip_grad_norm_(1.0) + Adam(0.5, 0.9).
# --------------------------------------------------------------------------------------------------
def workload_string(R, perceptual_weight, disc_factor):
    return ("autoencoder_kl_16x16x16 pose config, %dx%d crops, train step (fwd + KL/L1/pose loss + bwd + allreduce + "
            "clip 1.0 + Adam), global_step 80000, perceptual_weight=%g, disc_factor=%g (BASELINE.json configs[%d])"
            % (R, R, perceptual_weight, disc_factor, 1 if R == RES else 4))


def oracle_step_fn(batch_size, device="cpu", autocast=None):
    from odvae_b200 import synth
    from odvae_b200.autoencoder import PoseAutoencoder
    from oracle import odvae_oracle
    dev = torch.device(device)
    if dev.type == "cpu":
        torch.set_num_threads(os.cpu_count() or 1)
    params = synth.model_config(synth.DDCONFIG, stats_file())
    shell = PoseAutoencoder(**params)  # CPU construction only: names/shapes of the reference state_dict
    synth.fill_state_dict_(shell.state_dict(), seed=23)
    sd = {}
    for k, v in shell.state_dict().items():
        if k.startswith("loss.perceptual_loss"):
            continue
        t = v.detach().clone().to(dev)
        frozen = (not torch.is_floating_point(t)) or "running_" in k or k.startswith("loss.discriminator")
        sd[k] = t if frozen else t.requires_grad_(True)
    del shell
    train = [v for k, v in sd.items() if v.requires_grad and k != "loss.logvar"]  # AE.py:367-374
    opt = torch.optim.Adam(train, lr=4.5e-6 * batch_size, betas=(0.5, 0.9))
    orc = odvae_oracle.Oracle(params, sd, synth.synthetic_dataset_stats(23), synth.LABELS)
    batch = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in synth.synthetic_batch(batch_size, RES, seed=23).items()}
    p = odvae_oracle.dropout_prob(STEP, params, params["lossconfig"]["params"]["encoder_pretrain_steps"])
    noise = tuple(t.to(dev) for t in synth.draw_noise(batch_size, p, seed=101))

    def step():
        opt.zero_grad(set_to_none=True)
        sd["loss.logvar"].grad = None
        if autocast is not None:
            with torch.autocast(dev.type, dtype=autocast):
                loss, _, _ = orc.training_step(batch, STEP, noise)
        else:
            loss, _, _ = orc.training_step(batch, STEP, noise)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(train, 1.0)
        opt.step()
        return loss.detach()

    return step


def cpu_baseline(batch_size=4, steps=3):
    step = oracle_step_fn(batch_size)
    step()  # warm-up (thread pools, mkldnn primitives)
    t0 = time.perf_counter()
    for _ in range(steps):
        float(step())
    dt = time.perf_counter() - t0
    return {"value": batch_size * steps / dt, "unit": "images/sec", "cores": os.cpu_count(), "kind": "port",
            "sample": "oracle/ (CPU fp32 restatement of the reference PyTorch path, pinned to the reference's goldens): 1 warm-up + "
                      "%d timed step(s) of fwd+loss+bwd+clip+Adam at batch %d, 256x256, all host threads, %.1f s"
                      % (steps, batch_size, dt)}


def gpu_eager_baseline(dev, steps=3):
    """The reference arithmetic (oracle restatement, plain torch ops) in eager mode on this B200: cuDNN 9 convs, cuBLAS
    bmm, ATen group_norm / softmax. (a) fp32 storage with TF32 tensor-core math (torch's default for convs; the reference
    sets matmul precision "medium", train.py:521), (b) bf16 autocast. Batch 64 if it fits, else the largest power of two."""
    out = {"what": "oracle/odvae_oracle.py on cuda (eager ATen/cuDNN/cuBLAS), same step: fwd+loss+bwd+clip+Adam, "
                   "perceptual_weight=0, disc_factor=0, 256x256, CUDA-event timed"}
    torch.backends.cudnn.allow_tf32 = True
    torch.backends.cuda.matmul.allow_tf32 = True
    for tag, ac in (("tf32", None), ("bf16_autocast", torch.bfloat16)):
        res = None
        for B in (64, 32, 16):
            try:
                step = oracle_step_fn(B, device=str(dev), autocast=ac)
                step()
                step()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(steps):
                    step()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / steps
                res = {"images_per_sec": B / (ms * 1e-3), "ms_per_step": ms, "batch": B, "steps": steps,
                       "peak_mem_GB": torch.cuda.max_memory_allocated() / 1e9}
                break
            except torch.cuda.OutOfMemoryError:
                res = {"error": "out of memory at batch %d" % B}
            finally:
                step = None
                import gc
                gc.collect()
                torch.cuda.empty_cache()
                torch.cuda.reset_peak_memory_stats()
        out[tag] = res
    return out


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B = 4  # the reference's own CPU-runnable case (configs[0]) -- a bounded sample of the batch-64 workload
    step = oracle_step_fn(B)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        float(step())
    dt = time.perf_counter() - t0
    v = B * args.steps / dt
    world = args.gpus
    line = {"impl": "reference", "metric": "AutoencoderKL train-step images/sec", "value": v, "unit": "images/sec",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_string(RES, 0.0, 0.0), "per_gpu_batch": args.batch,
                       "global_batch": args.batch * world, "resolution": RES, "parallelism": "dp%d" % world,
                       "l2": "per-step working set (activations ~40 GB at batch 64) far exceeds the 126 MB L2"},
            "per_step_sample": "each timed step = the same train step on a batch of %d crops (a bounded sample of the "
                               "batch-%d workload) on the host cores" % (B, args.batch),
            "cpu_baseline": {"value": v, "unit": "images/sec", "cores": os.cpu_count(), "kind": "port",
                             "sample": "%d timed steps of batch %d (oracle/ restatement of the reference, fwd+loss+bwd+clip+Adam)"
                                       % (args.steps, B)},
            "e2e": {"value": v, "unit": "images/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# roofline pass: per-call CUDA-event timing of the dominant kernel (tc_engine_kernel: every conv / GEMM)
# --------------------------------------------------------------------------------------------------
class EngineProfiler:
    """Wraps ops.call for ONE extra (untimed) step: brackets each conv/GEMM C-ABI call with CUDA events on the
    launching stream and adds up the algorithmic FLOPs (2*M*N*K of the real, unpadded problem)."""
    NAMES = {"conv2d_fwd", "conv2d_fwd_gnstats", "conv2d_dgrad", "conv2d_dgrad_gnbwd", "conv2d_wgrad", "gemm_bf16", "gemm_16", "attn_scores_softmax",
             "attn_dsoftmax", "upconv_fwd", "upconv_dgrad", "upconv_wgrad"}
    # memory-bound GroupNorm(+SiLU) calls, judged against the HBM roofline: algorithmic bytes at 2 B per
    # element are 2*E forward (stats + apply together: read x once, write y once) and 3*E backward (x, dy -> dx)
    GN_NAMES = {"groupnorm_stats", "groupnorm_apply", "groupnorm_fwd_small", "groupnorm_bwd", "groupnorm_bwd_from_partials"}

    def __init__(self):
        self.records = []
        self.gn = []

    @staticmethod
    def gn_bytes(name, a):
        if name == "groupnorm_stats":
            return 0.0  # its time counts, its bytes are the forward's 2*E attributed to the apply call
        if name == "groupnorm_apply":
            return 2.0 * a[6] * a[7] * a[8] * 2
        if name == "groupnorm_fwd_small":  # (x, gamma, beta, y, stats, B, HW, C, ...): statistics + apply in one launch
            return 2.0 * a[5] * a[6] * a[7] * 2
        if name == "groupnorm_bwd_from_partials":  # (x, stats, gamma, beta, dy, dx_add, dx, dgamma, dbeta, part, nbox, B, HW, C, ..)
            return 3.0 * a[11] * a[12] * a[13] * 2  # its reduce pass ran inside the dgrad epilogue (counted with the engine calls)
        return 3.0 * a[9] * a[10] * a[11] * 2

    def flops(self, name, a):
        if name in ("gemm_bf16", "gemm_16"):
            M, N, K, batch = a[3], a[4], a[5], a[6]
            return 2.0 * M * N * K * batch
        if name == "attn_scores_softmax":  # algorithmic: ONE q k^T (the second pass is a recompute, not counted)
            return 2.0 * a[3] * a[4] * a[4] * a[5]
        if name == "attn_dsoftmax":
            return 2.0 * a[5] * a[6] * a[6] * a[7]
        if name.startswith("upconv_"):
            # EXECUTED flops of the parity-phase form: 16 taps over the low-resolution grid (the literal upsample + 3x3 conv the
            # SURVEY's algorithmic figure counts is 36); args: fwd (x, w, bias, y, B, H, W, CinP, CoutP, ..), dgrad (dy, w, dx,
            # B, H, W, ..), wgrad (x, dy, dw, db, B, H, W, CinP, CoutP, ..)
            o = {"upconv_fwd": 4, "upconv_dgrad": 3, "upconv_wgrad": 4}[name]
            B, H, W, CinP, CoutP = a[o], a[o + 1], a[o + 2], a[o + 3], a[o + 4]
            return 2.0 * B * H * W * CoutP * CinP * 16
        if name == "conv2d_fwd":
            B, H, W, CinP, CoutP, Cout, geom = a[5], a[6], a[7], a[8], a[9], a[10], a[11]
            cin, cout = self._real(CinP), Cout
        elif name == "conv2d_fwd_gnstats":  # (x, w, bias, residual, y, stats, B, H, W, CinP, CoutP, geom, ...)
            B, H, W, CinP, CoutP, geom = a[6], a[7], a[8], a[9], a[10], a[11]
            cin, cout = CinP, CoutP
        elif name == "conv2d_dgrad":
            B, H, W, CinP, CoutP, geom = a[4], a[5], a[6], a[7], a[8], a[9]
            cin, cout = self._real(CinP), self._real(CoutP)
        elif name == "conv2d_dgrad_gnbwd":  # (dy, w, dx, gn_x, stats, gamma, beta, part, bytes, B, H, W, CinP, CoutP, geom, ..)
            B, H, W, CinP, CoutP, geom = a[9], a[10], a[11], a[12], a[13], a[14]
            cin, cout = CinP, CoutP
        else:
            B, H, W, Cin, Cout, geom = a[4], a[5], a[6], a[9], a[10], a[11]
            cin, cout = Cin, Cout
        taps = 1 if geom == 1 else 9
        Ho, Wo = (H // 2, W // 2) if geom == 2 else (H, W)
        return 2.0 * B * Ho * Wo * cout * cin * taps

    @staticmethod
    def shape_key(name, a):
        if name in ("gemm_bf16", "gemm_16"):
            return "gemm M%d N%d K%d b%d a_mn%d b_mn%d" % (a[3], a[4], a[5], a[6], a[13], a[14])
        if name == "attn_scores_softmax":
            return "attn qk^T+softmax (2 passes) B%d N%d C%d" % (a[3], a[4], a[5])
        if name == "attn_dsoftmax":
            return "attn dP+dsoftmax B%d N%d C%d" % (a[5], a[6], a[7])
        if name.startswith("upconv_"):
            o = {"upconv_fwd": 4, "upconv_dgrad": 3, "upconv_wgrad": 4}[name]
            return "%s (2x2 phases) B%d %dx%d->%dx%d %d->%d" % (name, a[o], a[o + 1], a[o + 2], 2 * a[o + 1], 2 * a[o + 2],
                                                               a[o + 3], a[o + 4])
        if name == "conv2d_fwd":
            return "fwd B%d %dx%d %d->%d g%d" % (a[5], a[6], a[7], a[8], a[9], a[11])
        if name == "conv2d_fwd_gnstats":
            return "fwd+gnstats B%d %dx%d %d->%d g%d" % (a[6], a[7], a[8], a[9], a[10], a[11])
        if name == "conv2d_dgrad":
            return "dgrad B%d %dx%d %d<-%d g%d" % (a[4], a[5], a[6], a[7], a[8], a[9])
        if name == "conv2d_dgrad_gnbwd":
            return "dgrad+gnbwd sums B%d %dx%d %d<-%d g%d" % (a[9], a[10], a[11], a[12], a[13], a[14])
        return "wgrad B%d %dx%d %dx%d g%d" % (a[4], a[5], a[6], a[7], a[8], a[11])

    @staticmethod
    def _real(cp):
        return cp  # activations are padded only at the 3/16/32-channel edges; counted via Cin/Cout where known

    def install(self):
        from odvae_b200 import ops, lib
        self.ops, self.orig = ops, ops.call

        def wrapped(name, *a):
            if name in self.NAMES:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                out = self.orig(name, *a)
                e1.record()
                self.records.append((name, e0, e1, self.flops(name, a), self.shape_key(name, a)))
                return out
            if name in self.GN_NAMES:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                out = self.orig(name, *a)
                e1.record()
                self.gn.append((name, e0, e1, self.gn_bytes(name, a)))
                return out
            return self.orig(name, *a)

        ops.call = wrapped

    def remove(self):
        self.ops.call = self.orig

    def summary(self):
        torch.cuda.synchronize()
        tot_ms = sum(e0.elapsed_time(e1) for _, e0, e1, _, _ in self.records)
        tot_fl = sum(r[3] for r in self.records)
        by = {}
        for n, e0, e1, f, _ in self.records:
            t, ff, c = by.get(n, (0.0, 0.0, 0))
            by[n] = (t + e0.elapsed_time(e1), ff + f, c + 1)
        return tot_ms, tot_fl, len(self.records), by

    def gn_summary(self, hbm_peak):
        torch.cuda.synchronize()
        out = {}
        for tag, names in (("fwd", ("groupnorm_stats", "groupnorm_apply", "groupnorm_fwd_small")), ("bwd", ("groupnorm_bwd", "groupnorm_bwd_from_partials"))):
            ms = sum(e0.elapsed_time(e1) for n, e0, e1, _ in self.gn if n in names)
            by = sum(b for n, _, _, b in self.gn if n in names)
            if ms > 0:
                gbs = by / (ms * 1e-3) / 1e9
                out[tag] = {"ms": ms, "algorithmic_GB": by / 1e9, "achieved_GBps": gbs, "frac_of_hbm_peak": gbs / hbm_peak}
        return out

    def shapes(self):
        tab = {}
        for n, e0, e1, f, key in self.records:
            t, ff, c = tab.get(key, (0.0, 0.0, 0))
            tab[key] = (t + e0.elapsed_time(e1), ff + f, c + 1)
        rows = [{"shape": k, "calls": c, "ms": t, "tflops": ff / (t * 1e-3) / 1e12 if t > 0 else None}
                for k, (t, ff, c) in tab.items()]
        rows.sort(key=lambda r: -r["ms"])
        return rows


# --------------------------------------------------------------------------------------------------
# encode-only inference (BASELINE.json configs[2]): PoseAutoencoder.encode + posterior.mode(), no_grad
# --------------------------------------------------------------------------------------------------
FLOPS_ENCODE_PER_IMG = 175.18e9


def run_encode(args, model, dev, B, R, rank, world):
    import torch.distributed as dist
    from odvae_b200 import lib, synth
    host = synth.synthetic_batch(B, R, seed=23 + rank)
    x_host = host["patch"].float().contiguous()  # [B,3,R,R] in [0,1), as the data loader delivers it
    x_pin = x_host.pin_memory()
    x_dev = x_host.to(dev)

    recon = args.workload == "recon"

    def step(x):
        if recon:
            for prm in model.parameters():
                prm.grad = None
            xr = model._rescale(x)
            post, _ = model.encode(xr)
            rec = model.decode(post.mode())
            loss = (rec - xr).abs().mean()
            loss.backward()
            return loss.detach().reshape(1)
        with torch.no_grad():
            post, _ = model.encode(model._rescale(x))
            return post.mode()

    def timed(src, steps, e2e):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            z = step(src.to(dev, non_blocking=True) if e2e else src)
            if e2e:
                z = z.cpu()  # device -> host read of the latents
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms, z

    # small batches are bound by the host's launch rate: replay the static launch sequence as one CUDA graph
    use_graph = (not recon) and (args.graph == "on" or (args.graph == "auto" and B <= 32))
    if use_graph:
        from odvae_b200.graphs import GraphedEncode
        genc = GraphedEncode(model, B, R)
        eager_step = step
        step = lambda x: genc(x)
        torch.cuda.synchronize()
        assert torch.equal(step(x_dev), eager_step(x_dev)), "graph replay differs from the eager launch sequence"
    for _ in range(max(args.warmup, 3)):
        step(x_dev)
    l0 = lib.call("launch_count")
    ms, z = timed(x_dev, args.steps, False)
    launches = genc.launches_per_replay * args.steps if use_graph else lib.call("launch_count") - l0
    ms_e2e, z = timed(x_pin, args.steps, True)
    if rank == 0:
        n_img = B * world * args.steps
        pk, how = peaks()
        line = {"metric": "AutoencoderKL encode+decode fwd+bwd images/sec" if recon else "AutoencoderKL encode-only images/sec", "value": n_img / (ms * 1e-3), "unit": "images/sec",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": DTYPE_STR[args.mode],
                "data": "synthetic",
                "config": {"workload": ("encode -> mode -> decode -> L1 -> backward (Encoder, Decoder, quant convs; no pose "
                                        "heads, no optimizer), %dx%d crops (BASELINE.json configs[4])" if recon else
                                        "encode-only inference: Encoder + DiagonalGaussian mode(), %dx%d crops, no_grad "
                                        "(BASELINE.json configs[2])") % (R, R), "per_gpu_batch": B, "resolution": R,
                           "cuda_graph": bool(use_graph)},
                "e2e": {"value": n_img / (ms_e2e * 1e-3), "unit": "images/sec",
                        "h2d_bytes_per_step": x_host.numel() * 4, "d2h_bytes_per_step": z.numel() * 4,
                        "ms_per_step": ms_e2e / args.steps},
                "gpu_launches": int(launches),
                # encode: conv/attention flops scale ~ (R/256)^2 for the convs; the recon figure uses SURVEY's 512^2 count
                "pct_of_tensor_roofline_whole_step": ((3 * 2967e9 if R == 512 else 3 * 483.3e9 * (R / 256.0) ** 2) if recon
                                                      else FLOPS_ENCODE_PER_IMG * (R / 256.0) ** 2) * B /
                (ms / args.steps * 1e-3) / 1e12 / pk["bf16_tflops_sustained"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------
MODES = {"fp16": torch.float16, "mixed": "mixed", "bf16": torch.bfloat16}
DTYPE_STR = {"fp16": "fp16 (tcgen05 kind::f16, fp32 accumulate; fp16 activation gradients with a static power-of-two scale)",
             "mixed": "fp16 fwd + bf16 attn/grad (tcgen05 kind::f16, fp32 accumulate)", "bf16": "bf16"}


def build_model(args, dev, world, perceptual_weight, disc_factor):
    from odvae_b200 import synth
    from odvae_b200.autoencoder import PoseAutoencoder
    params = synth.model_config(synth.DDCONFIG, stats_file())
    params["lossconfig"]["params"]["perceptual_weight"] = perceptual_weight
    params["lossconfig"]["params"]["disc_factor"] = disc_factor
    torch.manual_seed(23)
    model = PoseAutoencoder(**params, compute_dtype=MODES[args.mode])
    synth.fill_state_dict_({k: v for k, v in model.state_dict().items() if not k.startswith("loss.perceptual_loss.")},
                           seed=23)  # identical replicas on every rank (LPIPS keeps its seeded constructor weights)
    model = model.to(dev)
    model.global_step = STEP
    model.learning_rate = 1 * world * args.batch * 4.5e-6  # train.py:383: accumulate_grad_batches * ngpu * batch * base_lr
    return model


def replicas_identical(model, world):
    """max |param(rank r) - param(rank 0)| over all parameters and ranks after the timed steps (DDP keeps replicas in lockstep:
    identical initial weights + identical all-reduced gradients + deterministic kernels => exactly 0)."""
    import torch.distributed as dist
    flat = torch.cat([p.detach().reshape(-1).float() for p in model.parameters()])
    ref = flat.clone()
    dist.broadcast(ref, 0)
    d = (flat - ref).abs().max().reshape(1)
    dist.all_reduce(d, op=dist.ReduceOp.MAX)
    return float(d)


def time_graph_step(args, dev, B, R, steps):
    """The configs[1] step replayed as ONE CUDA graph (GraphedTrainStep): same launch sequence, host enqueue ~3 ms."""
    from odvae_b200 import synth
    from odvae_b200.graphs import GraphedTrainStep
    from odvae_b200.trainer import TrainStep
    model = build_model(args, dev, 1, 0.0, 0.0)
    fn = GraphedTrainStep(TrainStep(model, lr=model.learning_rate, max_grad_norm=1.0))
    res = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in synth.synthetic_batch(B, R, seed=23).items()}
    for i in range(3):
        fn(res, i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for i in range(steps):
        loss = fn(res, i)
    e1.record()
    host_ms = (time.perf_counter() - t0) * 1e3 / steps
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    out = {"images_per_sec": B / (ms * 1e-3), "ms_per_step": ms, "steps": steps, "host_enqueue_ms_per_step": host_ms,
           "launches_per_replay": fn.launches_per_replay, "last_loss": float(loss)}
    del fn, model
    return out


def time_full_objective(args, dev, B, R, steps, warmup):
    """One whole batch of the reference loop with the yaml's own objective (perceptual_weight 1, disc_factor 1): the
    optimizer_idx 0 step (LPIPS, PatchGAN generator term, adaptive d_weight, Adam) + the optimizer_idx 1 step (two
    discriminator passes, hinge loss, Adam). `reference_forward_twice` = AE.py:321-327 semantics (the AE forward runs again
    for the discriminator step); `reuse_forward` = SURVEY 8 f2's cached variant."""
    from odvae_b200 import lib, synth
    from odvae_b200.fit import Fit
    out = {}
    for tag, reuse in (("reference_forward_twice", False), ("reuse_forward", True)):
        model = build_model(args, dev, 1, 1.0, 1.0)
        model.reuse_forward_for_disc = reuse
        fit = Fit(model, batch_size=B, base_lr=4.5e-6)
        res = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in synth.synthetic_batch(B, R, seed=23).items()}
        clone = lambda b: {k: (v.clone() if torch.is_tensor(v) else list(v)) for k, v in b.items()}
        for i in range(warmup):
            fit.training_batch(clone(res), i)
        torch.cuda.synchronize()
        l0 = lib.call("launch_count")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            ae, dl = fit.training_batch(clone(res), i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out[tag] = {"images_per_sec": B / (ms * 1e-3), "ms_per_batch": ms, "steps": steps,
                    "gpu_launches_per_batch": int((lib.call("launch_count") - l0) / steps),
                    "aeloss": float(ae), "discloss": float(dl),
                    "d_weight": float(model.logged.get("train/d_weight", 0.0)), "g_loss": float(model.logged.get("train/g_loss", 0.0))}
        del fit, model
        import gc
        gc.collect()
        torch.cuda.empty_cache()
    out["what"] = ("one batch = optimizer_idx 0 step + optimizer_idx 1 step, perceptual_weight=1 (LPIPS with seeded stand-in "
                   "weights: vgg.pth is not downloadable), disc_factor=1, batch %d, %dx%d, global_step 80000+" % (B, R, R))
    return out


def run_ours(args):
    import torch.distributed as dist
    from odvae_b200 import lib, synth
    from odvae_b200.fit import Fit
    from odvae_b200.trainer import TrainStep

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    lib.require_device()
    B = args.batch
    full = args.objective == "full"
    pw = 1.0 if full else args.perceptual_weight
    df = 1.0 if full else 0.0
    model = build_model(args, dev, world, pw, df)
    if full:
        fit = Fit(model, batch_size=B, base_lr=4.5e-6)
        step_fn = lambda b, i: fit.training_batch(b, i)[0]
    else:
        step_fn = TrainStep(model, lr=model.learning_rate, max_grad_norm=1.0)
    eager_fn = step_fn
    use_graph = (not full) and args.workload == "train" and \
        args.step_graph == "on" and not args.ncu_step
    if use_graph:
        from odvae_b200.graphs import GraphedTrainStep
        step_fn = GraphedTrainStep(eager_fn)

    R = args.res
    if args.workload in ("encode", "recon"):
        return run_encode(args, model, dev, B, R, rank, world)
    host = synth.synthetic_batch(B, R, seed=23 + rank)  # each rank its own shard of the synthetic stream
    pinned = {k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in host.items()}
    resident = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in host.items()}
    h2d = sum(v.numel() * v.element_size() for v in host.values() if torch.is_tensor(v))
    h2d += B * 16 * 16 * 16 * 4 * 2 + B * 8 * 4  # eps, z noise, bbox eps drawn on the host (reference semantics)

    def clone(b):
        return {k: (v.clone() if torch.is_tensor(v) else list(v)) for k, v in b.items()}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_ms = [0.0]

    def timed(batch_src, steps, read_loss):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        t_host0 = time.perf_counter()
        for i in range(steps):
            if batch_src is pinned:
                b = {k: (v.to(dev, non_blocking=True) if torch.is_tensor(v) else list(v)) for k, v in pinned.items()}
            else:
                b = clone(resident)
            loss = step_fn(b, i)
            if read_loss:
                last = float(loss)  # device -> host read of the step's result
        e1.record()
        host_ms[0] = (time.perf_counter() - t_host0) * 1e3 / steps  # host time to ENQUEUE a step (no sync inside)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        return ms, last

    for i in range(max(args.warmup, 3)):
        step_fn(clone(resident), i)
    if args.ncu_step:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step_fn(clone(resident), 0)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    launches0 = lib.call("launch_count")
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms, _ = timed(resident, args.steps, False)
    host_enqueue_ms = host_ms[0]
    launches = step_fn.launches_per_replay * args.steps if use_graph else lib.call("launch_count") - launches0
    ms_e2e, last_loss = timed(pinned, args.steps, True)
    clocks = sampler.stop() if rank == 0 else None
    identical = replicas_identical(model, world) if world > 1 else None

    roof = None
    flops_img = FLOPS_TRAIN_PER_IMG
    if not args.no_roofline and rank != 0:
        eager_fn(clone(resident), 0)  # every rank runs the extra (untimed) step: its gradient all-reduces are collective
        torch.cuda.synchronize()
    if rank == 0 and not args.no_roofline:
        prof = EngineProfiler()
        prof.install()
        from odvae_b200 import ops as _ops
        overlap, _ops.OVERLAP_WGRAD[0] = _ops.OVERLAP_WGRAD[0], False  # per-call times: every kernel alone on the GPU
        eager_fn(clone(resident), 0)  # the SAME launch sequence eagerly, each conv / GEMM / GroupNorm call between CUDA events
        _ops.OVERLAP_WGRAD[0] = overlap
        prof.remove()
        eng_ms, counted, ncalls, by = prof.summary()
        if args.shapes:
            with open(args.shapes, "w") as f:
                json.dump(prof.shapes(), f, indent=1)
        pk, how = peaks()
        peak = pk["bf16_tflops_sustained"]
        if R != RES or full:  # other workloads: count the flops of the launched problems (no SURVEY figure)
            flops_img = counted / B
        achieved = flops_img * B / (eng_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "kernel": "tc_engine_kernel (tcgen05 implicit-GEMM conv fwd/dgrad/wgrad + attention GEMMs)",
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "peak_source": how + " (bf16_tflops_sustained: kernel timed inside a long step)",
                # DRAM bytes are not measurable from inside this process (no ncu under the timed run): null here; the per-launch
                # dram__bytes of this same command are in the committed ncu launch list (profiles/r02_launches*.txt)
                "traffic": None,
                "engine_ms_per_step": eng_ms, "engine_calls_per_step": ncalls,
                "engine_share_of_step": eng_ms / (ms / args.steps),
                "algorithmic_flops_per_step": flops_img * B, "counted_padded_flops_per_step": counted,
                "by_call": {k: {"ms": v[0], "tflops_padded": v[1] / (v[0] * 1e-3) / 1e12 if v[0] > 0 else None, "calls": v[2]}
                            for k, v in by.items()},
                # second roofline (HBM-bound kernels): GroupNorm(+SiLU) forward (stats + apply) and backward
                # (reduce + finalize + apply) C-ABI calls, algorithmic bytes over CUDA-event time, peak = hbm_gbs
                "hbm_kernels": {"peak_GBps": pk["hbm_gbs"], "groupnorm": prof.gn_summary(pk["hbm_gbs"])}}

    if rank == 0:
        n_img = B * world * args.steps
        line = {"metric": "AutoencoderKL train-step images/sec", "value": n_img / (ms * 1e-3), "unit": "images/sec",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": DTYPE_STR[args.mode], "data": "synthetic",
                "config": {"workload": workload_string(R, pw, df) if not full else
                           "autoencoder_kl_16x16x16 pose config, %dx%d crops, ONE BATCH of the reference loop with the yaml's own "
                           "objective: optimizer_idx 0 (LPIPS + GAN generator term, adaptive weight) + optimizer_idx 1 "
                           "(discriminator) steps, global_step 80000+" % (R, R),
                           "per_gpu_batch": B, "global_batch": B * world, "resolution": R,
                           "parallelism": "dp%d" % world,
                           "l2": "per-step working set (activations ~40 GB at batch 64) far exceeds the 126 MB L2"},
                "e2e": {"value": n_img / (ms_e2e * 1e-3), "unit": "images/sec", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e / args.steps, "last_loss": last_loss},
                "gpu_launches": int(launches), "host_enqueue_ms_per_step": host_enqueue_ms,
                "cuda_graph": bool(use_graph),
                "clocks": clocks, "roofline": roof, "cpu_baseline": None,
                "pct_of_tensor_roofline_whole_step": flops_img * B / (ms / args.steps * 1e-3) / 1e12 /
                peaks()[0]["bf16_tflops_sustained"]}
        if identical is not None:
            line["replicas_identical"] = {"max_abs_param_diff_vs_rank0": identical, "identical": identical == 0.0}
    # ---- N = 1 extras: the other BASELINE.json configs and comparators, each a short measurement after the main one
    if world == 1 and rank == 0:
        import gc
        del step_fn, eager_fn, model
        if full:
            del fit
        gc.collect()
        torch.cuda.empty_cache()
        if args.extras and not full and R == RES:
            extra = {}
            try:
                extra["cuda_graph_step"] = time_graph_step(args, dev, B, R, steps=args.steps)
            except Exception as e:
                extra["cuda_graph_step"] = {"error": repr(e)[:300]}
            gc.collect()
            torch.cuda.empty_cache()
            try:
                extra["full_objective"] = time_full_objective(args, dev, B, R, steps=3, warmup=2)
            except Exception as e:  # an extra must never take the bench line down
                extra["full_objective"] = {"error": repr(e)[:300]}
            try:
                extra["encode_sweep"] = encode_sweep(args, dev)
            except Exception as e:
                extra["encode_sweep"] = {"error": repr(e)[:300]}
            try:
                extra["recon_512"] = recon_512(args, dev)
            except Exception as e:
                extra["recon_512"] = {"error": repr(e)[:300]}
            gc.collect()
            torch.cuda.empty_cache()
            try:
                extra["gpu_eager_baseline"] = gpu_eager_baseline(dev)
            except Exception as e:
                extra["gpu_eager_baseline"] = {"error": repr(e)[:300]}
            line["extra"] = extra
        if args.cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(batch_size=4, steps=3)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def _timed_loop(fn, steps):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, out


def encode_sweep(args, dev, batches=(1, 16, 128), steps=10):
    """BASELINE.json configs[2]: encode-only inference (Encoder + posterior.mode(), no_grad) at 256x256; CUDA-graph replay for
    batch <= 32 (odvae_b200/graphs.py), eager launches above."""
    from odvae_b200 import synth
    pk, _ = peaks()
    model = build_model(args, dev, 1, 0.0, 0.0)
    rows = []
    for B in batches:
        x = synth.synthetic_batch(B, RES, seed=23)["patch"].float().to(dev)
        if B <= 32:
            from odvae_b200.graphs import GraphedEncode
            genc = GraphedEncode(model, B, RES)
            fn = lambda: genc(x)
        else:
            def fn():
                with torch.no_grad():
                    return model.encode(model._rescale(x))[0].mode()
        for _ in range(3):
            fn()
        ms, _ = _timed_loop(fn, steps)
        rows.append({"batch": B, "images_per_sec": B / (ms * 1e-3), "ms_per_step": ms, "cuda_graph": B <= 32,
                     "pct_of_tensor_roofline": FLOPS_ENCODE_PER_IMG * B / (ms * 1e-3) / 1e12 / pk["bf16_tflops_sustained"]})
    del model
    return rows


def recon_512(args, dev, steps=3):
    """BASELINE.json configs[4]: 512x512 crops, encode -> mode -> decode -> L1 -> backward through decoder and encoder (no pose
    heads: the config's pose MLPs are sized for 16x16 latents, i.e. 256x256 crops only, in the reference too)."""
    from odvae_b200 import synth
    pk, _ = peaks()
    out = None
    for B in (32, 16, 8):
        model = build_model(args, dev, 1, 0.0, 0.0)
        try:
            x = synth.synthetic_batch(B, 512, seed=23)["patch"].float().to(dev)

            def fn():
                for prm in model.parameters():
                    prm.grad = None
                xr = model._rescale(x)
                post, _ = model.encode(xr)
                rec = model.decode(post.mode())
                loss = (rec - xr).abs().mean()
                loss.backward()
                return loss.detach()
            fn()
            fn()
            ms, _ = _timed_loop(fn, steps)
            out = {"batch": B, "images_per_sec": B / (ms * 1e-3), "ms_per_step": ms,
                   "pct_of_tensor_roofline": 3 * 2967e9 * B / (ms * 1e-3) / 1e12 / pk["bf16_tflops_sustained"],
                   "peak_mem_GB": torch.cuda.max_memory_allocated() / 1e9}
            break
        except torch.cuda.OutOfMemoryError:
            out = {"error": "out of memory at batch %d" % B}
        finally:
            del model
            import gc
            gc.collect()
            torch.cuda.empty_cache()
    return out


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
