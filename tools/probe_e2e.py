# per-epoch durations of train_model() at the bench shape (where does e2e lose time vs the device-resident step?)
import sys, os, time, torch
sys.path.insert(0, "tests"); sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from helpers import build_gpt, scratch_cwd
scratch_cwd()
from penroz import shards
shards.write_synthetic_shards("synth50257", 50257, 1, 2 ** 22)
m = build_gpt("g124", 12, 12, 768, 1024, 50257); m.to("cuda")
import neural_net_model as nnm
orig = nnm._FusedStepper.run_epoch
marks = []
def timed(self, *a, **k):
    t0 = time.time(); r = orig(self, *a, **k); marks.append(time.time() - t0); return r
nnm._FusedStepper.run_epoch = timed
t0 = time.time()
m.train_model("synth50257", 0, 12, 32, 1024, 32)
print("total train_model wall", round(time.time() - t0, 2), "s")
print("epoch durationInSecs (ms):", [round(p["durationInSecs"] * 1e3, 1) for p in m.progress])
print("run_epoch host time (ms):  ", [round(x * 1e3, 1) for x in marks])
# cost of the end-of-training diagnostics alone (neural_net_model.py:724-777 equivalent), GPU summaries vs host path
acts = [(torch.randn(32, 1024, 768, device="cuda"), torch.randn(32, 1024, 768, device="cuda") * 1e-3) for _ in range(3)]
torch.cuda.synchronize(); t0 = time.time()
for a, g in acts:
    from penroz import ops
    ops.tensor_stats(a); ops.tensor_stats(g)
torch.cuda.synchronize(); t_gpu = (time.time() - t0) / 6
t0 = time.time()
m._density_histogram(acts[0][0]); acts[0][0].float().mean().item(); acts[0][0].float().std().item()
t_host = time.time() - t0
print(f"diagnostics of one [32,1024,768] fp32 tensor: GPU {t_gpu*1e3:.2f} ms, host path {t_host*1e3:.1f} ms")
