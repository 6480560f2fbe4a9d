# attention backward with the QKV bias-gradient side output, B=32 T=1024 H=12: total time of the three kernels
import sys, torch
sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from penroz import ops
B, T, H, D = 32, 1024, 12, 64
C = H * D
qkv = torch.randn(B, T, 3 * C, device="cuda").bfloat16()
out, lse = ops.attn_fwd(qkv, H)
dout = torch.randn(B, T, C, device="cuda").bfloat16()
db = torch.zeros(3 * C, device="cuda")
scratch = torch.empty(ops.attn_bwd_scratch_bytes(B, T, H, D, qkv.dtype), device="cuda", dtype=torch.uint8)
dqkv = torch.empty_like(qkv)
for with_bias in (False, True):
    for _ in range(3): ops.attn_bwd(qkv, out, dout, lse, H, scratch=scratch, dqkv=dqkv, dbias=db if with_bias else None)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): ops.attn_bwd(qkv, out, dout, lse, H, scratch=scratch, dqkv=dqkv, dbias=db if with_bias else None)
    e1.record(); torch.cuda.synchronize()
    print(f"attn_bwd dbias={with_bias}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
