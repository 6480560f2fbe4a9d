# quick perf probe of the fused step at the headline shape (not the bench; just to see where time goes)
import sys, os, time, torch
sys.path.insert(0, "tests"); sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
import penroz._lib as _L
if os.environ.get("PENROZ_AB_LIB"): _L.LIB_PATH = os.path.abspath(os.environ["PENROZ_AB_LIB"])   # A/B against another build
from helpers import build_gpt, scratch_cwd
root=os.getcwd(); scratch_cwd()
B=int(os.environ.get("PB","32")); T=1024
m=build_gpt("g124",12,12,768,1024,50257); m.to("cuda")
eng=m._engine(); eng.setup_training(use_amp=True)
x=torch.randint(0,50257,(B,T),device="cuda"); y=torch.randint(0,50257,(B,T),device="cuda")
def step():
    eng.zero_grad(); l=eng.train_micro_step(x,y,1.0); eng.optimizer_step(); return l
for _ in range(3): l=step()
torch.cuda.synchronize(); print("loss", l.item(), "mem GB", torch.cuda.max_memory_allocated()/2**30)
s=torch.cuda.Event(enable_timing=True); e=torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(5): step()
e.record(); torch.cuda.synchronize()
ms=s.elapsed_time(e)/5
print(f"ms/step {ms:.2f} tok/s {B*T/ms*1e3:.0f}  MFU(2.25PF) {B*T/ms*1e3*797.8e6/2.25e15:.3f}")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))
