# /generate/-like request latency: NeuralNetworkModel.deserialize(id) + generate_tokens, first request vs later ones
import os, sys, time, torch
sys.path.insert(0, "tests"); sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from helpers import build_gpt, scratch_cwd
scratch_cwd()
import neural_net_model as nnm
m = build_gpt("g124", 12, 12, 768, 1024, 50257)
m.serialize(); del m
time.sleep(3)   # the disk flush runs in a child process
prompt = torch.randint(0, 50257, (1, 512), generator=torch.Generator().manual_seed(7)).tolist()
def request():
    t0 = time.perf_counter()
    model = nnm.NeuralNetworkModel.deserialize("g124")
    t1 = time.perf_counter()
    toks = model.generate_tokens(prompt, 1024, 64, temperature=0.0)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    return (t1 - t0) * 1e3, (t2 - t1) * 1e3
for mode in ("0", "1"):
    os.environ["PENROZ_MODEL_CACHE"] = mode
    nnm._RESIDENT.clear()
    for i in range(3):
        d, g = request()
        print(f"PENROZ_MODEL_CACHE={mode} request {i}: deserialize {d:8.1f} ms   generate 512+64 {g:8.1f} ms")
