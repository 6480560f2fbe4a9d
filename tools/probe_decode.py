# where does a B=1 decode token go?  wall per token vs GPU kernel time per token (torch profiler), 124M bf16
import os, sys, time, torch
sys.path.insert(0, "tests"); sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from helpers import build_gpt, scratch_cwd
scratch_cwd()
B = int(os.environ.get("PB", "1"))
m = build_gpt("dec", 12, 12, 768, 1024, 50257); m.to(dtype=torch.bfloat16); m.to("cuda")
g = torch.Generator().manual_seed(7)
prompt = torch.randint(0, 50257, (B, 512), generator=g).tolist()
m.generate_tokens([p[:16] for p in prompt], 1024, 4, temperature=0.0)
torch.cuda.synchronize()
for n in (1, 65):
    t0 = time.perf_counter(); m.generate_tokens(prompt, 1024, n, temperature=0.0); torch.cuda.synchronize()
    print(f"B={B} prompt 512 + {n} new: {1e3*(time.perf_counter()-t0):.2f} ms")
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    m.generate_tokens(prompt, 1024, 33, temperature=0.0); torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=60))
