# Where does the multi-GPU step lose time?  torchrun --nproc-per-node N tests/probe_scale.py
#   A: step with the overlapped bucketed all-reduce (the product path)   B: same step, no gradient exchange
#   C: the all-reduce alone (all buckets back to back, nothing else running)
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, "tests"); sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
import penroz._lib as _L
if os.environ.get("PENROZ_AB_LIB"): _L.LIB_PATH = os.path.abspath(os.environ["PENROZ_AB_LIB"])
from helpers import build_gpt, scratch_cwd
from penroz.grad_sync import GradSync, gpt2_segments
os.environ.setdefault("NCCL_DEBUG", "WARN")
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
scratch_cwd()
B, T = 32, 1024
m = build_gpt(f"g124_{rank}", 12, 12, 768, 1024, 50257); m.to("cuda")
eng = m._engine(); eng.setup_training(use_amp=True)
eng.grad_sync = GradSync(eng.flat.g, gpt2_segments(eng.flat, eng.plan), bucket_bytes=int(os.environ.get("PB_BUCKET_MB", "64")) << 20)
eng.grad_sync.broadcast_(eng.flat.p); eng.flat.refresh_shadow()
g = torch.Generator().manual_seed(rank)
x = torch.randint(0, 50257, (B, T), generator=g).cuda(); y = torch.randint(0, 50257, (B, T), generator=g).cuda()
def step(sync):
    eng.zero_grad(); l = eng.train_micro_step(x, y, 1.0, sync_grads=sync); eng.optimizer_step(); return l
def reduce_only():
    gs = eng.grad_sync
    for name in dict.fromkeys(n for _, _, n in gs.buckets): gs.ready(name)
    gs.wait_all()
def timed(fn, n=10):
    for _ in range(3): fn()
    dist.barrier(); torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / n], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()
a = timed(lambda: step(True)); b = timed(lambda: step(False)); c = timed(reduce_only); a2 = timed(lambda: step(True))
if rank == 0:
    print(f"lib {os.path.basename(os.path.dirname(_L.LIB_PATH))}/{os.path.basename(_L.LIB_PATH)} world {world} buckets {len(eng.grad_sync.buckets)}: "
          f"with sync {a:.2f} / {a2:.2f} ms   no sync {b:.2f} ms   all-reduce alone {c:.2f} ms   tok/s {world*B*T/a*1e3:.0f}")
dist.destroy_process_group()
