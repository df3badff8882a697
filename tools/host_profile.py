"""Host-side cost of ENQUEUEING one train step (bench configuration): cProfile over a few steps without synchronising.
The step is GPU-bound only while the host stays ahead: this shows where the Python / ctypes time goes."""
import cProfile, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "generative-detection_b200"), ROOT]
import torch
import bench
from odvae_b200 import synth
from odvae_b200.autoencoder import PoseAutoencoder
from odvae_b200.trainer import TrainStep

B = int(os.environ.get("PB", 64))
dev = torch.device("cuda:0")
params = synth.model_config(synth.DDCONFIG, bench.stats_file())
model = PoseAutoencoder(**params, compute_dtype=torch.float16)
synth.fill_state_dict_(model.state_dict(), seed=23)
model = model.to(dev)
model.global_step = 80000
step = TrainStep(model, lr=B * 4.5e-6, max_grad_norm=1.0)
host = synth.synthetic_batch(B, 256, seed=23)
res = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in host.items()}
clone = lambda b: {k: (v.clone() if torch.is_tensor(v) else list(v)) for k, v in b.items()}
for i in range(3):
    step(clone(res), i)
torch.cuda.synchronize()
n = 3
t0 = time.perf_counter()
pr = cProfile.Profile()
pr.enable()
for i in range(n):
    step(clone(res), i)
pr.disable()
t1 = time.perf_counter()
torch.cuda.synchronize()
print("host enqueue per step (under cProfile): %.1f ms" % ((t1 - t0) * 1e3 / n))
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(28)
