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
