"""CUDA-graph replay of the encode-only inference path (BASELINE.json configs[2]).

`PoseAutoencoder.encode(x)` + `posterior.mode()` (AE.py:176-182, LD distributions.py:24-62) is ~75 kernel launches; at
batch 1-8 the kernels are a few microseconds each and the step is bound by the host's launch rate (3.4 ms floor through
ctypes + torch allocations). For a fixed (batch, resolution) the whole launch sequence is static - shapes, TMA
descriptors (passed by value as kernel parameters), workspaces and packed weights do not move - so it is captured once
into a CUDA graph and replayed with one driver call. Inputs are copied into the graph's static input buffer; the
returned latents live in the graph's static output buffer (clone them to keep them across calls).

After an optimizer step call `ops.refresh_param_caches()` before the next replay: the packed 16-bit weight copies are
refreshed IN PLACE (same addresses), so the captured graph picks the new weights up without re-capture.
"""
import torch

from . import lib


class GraphedEncode:
    def __init__(self, model, batch, resolution, rescale=True, warmup=2):
        self.model, self.rescale = model, rescale
        dev = next(model.parameters()).device
        self.x = torch.zeros(batch, 3, resolution, resolution, dtype=torch.float32, device=dev)
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side), torch.no_grad():  # packs weights, sizes workspaces, sets kernel attributes
            for _ in range(max(warmup, 1)):
                self._run()
        cur.wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        n0 = lib.call("launch_count")
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.z = self._run()
        self.launches_per_replay = int(lib.call("launch_count") - n0)

    def _run(self):
        x = self.model._rescale(self.x) if self.rescale else self.x
        post, _ = self.model.encode(x)
        return post.mode()

    def __call__(self, x):
        """x: [B,3,R,R] fp32 (host-pinned or device). Returns the graph's static latent buffer [B,16,R/16,R/16]."""
        self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        return self.z


class GraphedTrainStep:
    """CUDA-graph replay of trainer.TrainStep (training_step(optimizer_idx=0) -> backward -> gradient all-reduce -> clip ->
    Adam -> weight re-pack): ~1000 launches per step that the host otherwise spends 120-135 ms enqueuing through ctypes for a
    ~146 ms GPU step. Everything the host decides per step is either staged into static device buffers
    BEFORE the replay (the batch, the reference's four random draws, the dropout keep-mask and its 1/(1-p), the per-sample
    prior rows) or lives on the device (Adam's per-parameter step counters, optim.FusedAdam); what is baked into the graph
    is constant within a PHASE of the schedule, so one graph is captured per (phase, batch size) key and
    re-captured when `global_step` crosses a threshold.

    The replayed launch sequence is the eager one, kernel for kernel: losses and parameters stay bit-identical to
    TrainStep (tests/test_graph_gpu.py).
    """

    def __init__(self, step, warmup=2):
        self.step = step
        self.model = step.model
        self.graphs = {}
        self.warmup = warmup
        self.pool = None
        self.launches_per_replay = 0

    # ---- what the captured launch sequence depends on besides tensor contents
    def _key(self, B, R):
        m, ls = self.model, self.model.loss
        s = m.global_step
        p = m._get_dropout_prob()
        from .losses import adopt_weight
        return (B, R, s < m.encoder_pretrain_steps, s > ls.encoder_pretrain_steps,
                s < ls.encoder_pretrain_steps + ls.pose_conditioned_generation_steps,
                0 if p <= 0 else (2 if p >= 1.0 else 1),
                adopt_weight(ls.disc_factor, s, threshold=ls.discriminator_iter_start) != 0)

    def _stage(self, st, batch):
        """Host -> static device buffers (async copies from pinned memory where the source is pinned)."""
        m = self.model
        for k, buf in st["batch"].items():
            buf.copy_(batch[k], non_blocking=True)
        B = st["B"]
        p = m._get_dropout_prob()
        m.dropout_prob = p
        if m.inject_noise is not None:
            eps, keep, noise, eps_bbox = m.inject_noise
        else:  # the reference's draws, same generator and order as PoseAutoencoder._draw_noise
            shape = (B,) + tuple(m.feature_dims)
            eps = torch.randn(shape, pin_memory=True)
            noise = torch.normal(torch.zeros(shape), torch.ones(shape)).pin_memory() if m.add_noise_to_z_obj else None
            eps_bbox = torch.randn(B, 8, pin_memory=True)
            keep = None
        st["eps"].copy_(eps, non_blocking=True)
        if noise is not None and st["noise"] is not None:
            st["noise"].copy_(noise, non_blocking=True)
        st["eps_bbox"].copy_(eps_bbox, non_blocking=True)
        if st["keep"] is not None:
            if keep is not None:
                st["keep"].copy_(keep.to(torch.uint8), non_blocking=True)
            elif p >= 1.0:
                st["keep"].zero_()
            else:
                st["keep"].copy_(torch.empty(st["keep"].shape, dtype=torch.float32, device=st["keep"].device)
                                 .bernoulli_(1.0 - p).to(torch.uint8))
        # fresh pinned sources every step: the copies run asynchronously (the GPU is up to a step behind the host), so a
        # reused pinned buffer would be overwritten before its copy executes; the caching host allocator recycles them
        ks = torch.empty(1, pin_memory=True)
        ks[0] = 1.0 / (1.0 - p) if 0 < p < 1.0 else 1.0
        st["keep_scale"].copy_(ks, non_blocking=True)
        st["prior_rows"].copy_(m.loss._prior_rows_host(batch["class_name"]).pin_memory(), non_blocking=True)
        st["labels"] = list(batch["class_name"])

    def _static_batch(self, st):
        b = dict(st["batch"])
        b["class_name"] = st["labels"]
        return b

    def _run(self, st, batch_idx):
        m = self.model
        saved = (m.inject_noise, m._keep_scale_dev, m.loss._static_prior_rows)
        m.inject_noise = (st["eps"], st["keep"], st["noise"], st["eps_bbox"])
        m._keep_scale_dev = st["keep_scale"]
        m.loss._static_prior_rows = st["prior_rows"]
        try:
            return self.step(self._static_batch(st), batch_idx)
        finally:
            m.inject_noise, m._keep_scale_dev, m.loss._static_prior_rows = saved

    def _capture(self, key, batch, batch_idx):
        m = self.model
        dev = next(m.parameters()).device
        B, R = key[0], key[1]
        f32 = dict(dtype=torch.float32, device=dev)
        lat = (B,) + tuple(m.feature_dims)
        p = m._get_dropout_prob()
        st = {"B": B,
              "batch": {k: torch.empty(v.shape, dtype=v.dtype, device=dev) for k, v in batch.items() if torch.is_tensor(v)},
              "eps": torch.empty(lat, **f32), "noise": torch.empty(lat, **f32) if m.add_noise_to_z_obj else None,
              "eps_bbox": torch.empty(B, 8, **f32),
              "keep": torch.empty(lat, dtype=torch.uint8, device=dev) if p > 0 else None,
              "keep_scale": torch.ones(1, **f32), "prior_rows": torch.empty(B, dtype=torch.int32, device=dev)}
        # eager steps on a side stream first: pack weights, size workspaces, build the optimizer's tables, set kernel
        # attributes. They must leave no trace: parameters, buffers, optimizer state and the RNG streams are snapshotted and
        # restored, so that one call of this object == one optimizer step, like TrainStep.
        snap = self._snapshot()
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(self.warmup):
                self._stage(st, batch)
                self._run(st, batch_idx)
        cur.wait_stream(side)
        torch.cuda.synchronize(dev)
        self._restore(snap)
        del snap
        torch.cuda.synchronize(dev)
        torch.cuda.empty_cache()  # the eager steps' activations: the graph gets its own pool
        st["graph"] = torch.cuda.CUDAGraph()
        rng = (torch.get_rng_state(), torch.cuda.get_rng_state(dev))
        self._stage(st, batch)
        torch.set_rng_state(rng[0])
        torch.cuda.set_rng_state(rng[1], dev)
        torch.cuda.synchronize(dev)
        n0 = lib.call("launch_count")
        with torch.cuda.graph(st["graph"], pool=self.pool):
            st["loss"] = self._run(st, batch_idx)  # recorded, not executed
        if self.pool is None:
            self.pool = st["graph"].pool()
        self.launches_per_replay = int(lib.call("launch_count") - n0)
        self.graphs[key] = st
        return st

    def _snapshot(self):
        m, opt = self.model, self.step.opt
        dev = next(m.parameters()).device
        state = {}
        for group in opt.param_groups:
            for p in group["params"]:
                s = opt.state.get(p)
                if s:
                    state[p] = {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in s.items()}
        return {"tensors": [(t, t.detach().clone()) for t in list(m.parameters()) + list(m.buffers())],
                "opt": state, "rng": (torch.get_rng_state(), torch.cuda.get_rng_state(dev)), "dev": dev}

    def _restore(self, snap):
        from . import ops
        opt = self.step.opt
        with torch.no_grad():
            for t, c in snap["tensors"]:
                t.copy_(c)
            for group in opt.param_groups:
                for p in group["params"]:
                    s = opt.state.get(p)
                    if not s:
                        continue
                    old = snap["opt"].get(p)
                    for k, v in s.items():
                        if torch.is_tensor(v):
                            if old is not None and torch.is_tensor(old.get(k)):
                                v.copy_(old[k])
                            else:
                                v.zero_()  # state created by the warm-up steps: back to "never stepped"
        torch.set_rng_state(snap["rng"][0])
        torch.cuda.set_rng_state(snap["rng"][1], snap["dev"])
        ops.bump_weight_epoch()
        ops.refresh_param_caches()

    def __call__(self, batch, batch_idx=0):
        """One optimizer step, like TrainStep.__call__. -> loss (the graph's static output: clone it to keep it across calls).
        The first call for a (phase, batch-size) key also runs `warmup` eager steps (rolled back) and the capture pass."""
        patch = batch[self.model.image_rgb_key]
        key = self._key(patch.shape[0], patch.shape[-1])
        st = self.graphs.get(key)
        if st is None:
            st = self._capture(key, batch, batch_idx)
        self._stage(st, batch)
        st["graph"].replay()
        return st["loss"]
