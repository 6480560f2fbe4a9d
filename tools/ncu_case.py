# One launch of each hot kernel at the headline shapes, for ncu (and runnable plain).
import os, sys, torch
sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from penroz import ops
M, C, H = 32768, 768, 12
dev = "cuda"
def rnd(*s): return (torch.randn(*s, device=dev) * 0.05).bfloat16()
x768, x3072 = rnd(M, 768), rnd(M, 3072)
wqkv, wproj, wfc, wproj2 = rnd(2304, 768), rnd(768, 768), rnd(3072, 768), rnd(768, 3072)
b768, b2304, b3072 = rnd(768), rnd(2304), rnd(3072)
res = torch.randn(M, 768, device=dev)
o768f = torch.empty(M, 768, device=dev)
o2304b = torch.empty(M, 2304, device=dev, dtype=torch.bfloat16)
o3072b, o3072b2 = torch.empty(M, 3072, device=dev, dtype=torch.bfloat16), torch.empty(M, 3072, device=dev, dtype=torch.bfloat16)
g_fc = torch.zeros(3072, 768, device=dev)
for rep in range(int(os.environ.get("NCU_REPS", "2"))):
    ops.gemm(x768, wqkv, bias=b2304, out=o2304b)                                               # plain
    ops.gemm(x768, wfc, epilogue=ops.EPI_GELU, bias=b3072, out=o3072b, out2=o3072b2)           # gelu
    ops.gemm(x768, wproj2, b_layout=1, epilogue=ops.EPI_DGELU, aux=o3072b, out=o3072b2)        # dgelu dgrad
    ops.gemm(x768, wproj, epilogue=ops.EPI_RESID_F32, bias=b768, aux=res, out=o768f)           # resid, K=768
    ops.gemm(x3072, x768, a_layout=1, b_layout=1, epilogue=ops.EPI_ACCUM_F32, out=g_fc, accumulate=True)  # wgrad split-K
qkv = torch.randn(32, 1024, 3 * C, device=dev).bfloat16()
dout = torch.randn(32, 1024, C, device=dev).bfloat16()
for rep in range(int(os.environ.get("NCU_REPS", "2"))):
    out, lse = ops.attn_fwd(qkv, H)
    dq = ops.attn_bwd(qkv, out, dout, lse, H)
x = torch.randn(M, C, device=dev); g = torch.randn(C, device=dev); b = torch.randn(C, device=dev)
y, mean, rstd = ops.layernorm_fwd(x, g, b, out_dtype=torch.bfloat16)
dg, db, dc = torch.zeros(C, device=dev), torch.zeros(C, device=dev), torch.zeros(C, device=dev)
for rep in range(int(os.environ.get("NCU_REPS", "2"))):
    ops.layernorm_bwd(y, x, g, mean, rstd, add=res, dgamma=dg, dbeta=db, dcol=dc, want_bf16=True)
torch.cuda.synchronize()
# one decode token step of the 124M bf16 model after a 512-token prefill (the fused small-batch kernels)
import os
sys.path.insert(0, "tests")
from helpers import build_gpt, scratch_cwd
root = os.getcwd(); scratch_cwd()
m = build_gpt("ncu_dec", 12, 12, 768, 1024, 50257); m.to(dtype=torch.bfloat16); m.to("cuda")
prompt = torch.randint(0, 50257, (1, 512), generator=torch.Generator().manual_seed(7)).tolist()
m.generate_tokens(prompt, 1024, 2, temperature=0.0)
torch.cuda.synchronize()
print("ok")
