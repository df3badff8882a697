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
