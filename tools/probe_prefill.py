# host-side profile of generate_tokens(prompt 512, 1 new token): where do the ~8 ms go?
import os, sys, time, torch, cProfile, pstats
sys.path.insert(0, "tests"); sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from helpers import build_gpt, scratch_cwd
scratch_cwd()
m = build_gpt("dec", 12, 12, 768, 1024, 50257); m.to(dtype=torch.bfloat16); m.to("cuda")
g = torch.Generator().manual_seed(7)
prompt = torch.randint(0, 50257, (1, 512), generator=g).tolist()
for _ in range(3): m.generate_tokens(prompt, 1024, 1, temperature=0.0)
torch.cuda.synchronize()
pr = cProfile.Profile(); pr.enable()
for _ in range(5): m.generate_tokens(prompt, 1024, 1, temperature=0.0)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
