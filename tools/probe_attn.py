# timing probe of the tcgen05 attention kernels at the headline shape
import sys, torch
sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from penroz import ops
B, T, H = 32, 1024, 12
C = H * 64
qkv = torch.randn(B, T, 3 * C, device="cuda").bfloat16()
dout = torch.randn(B, T, C, device="cuda").bfloat16()
out, lse = ops.attn_fwd(qkv, H)
scratch = torch.empty(ops.attn_bwd_scratch_bytes(B, T, H, 64, qkv.dtype), device="cuda", dtype=torch.uint8)
dq = ops.attn_bwd(qkv, out, dout, lse, H, scratch=scratch)
torch.cuda.synchronize()
def t(fn, n=10):
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
fl = 4.0 * B * H * T * T * 64 / 2
f = t(lambda: ops.attn_fwd(qkv, H, out=out, lse=lse))
b = t(lambda: ops.attn_bwd(qkv, out, dout, lse, H, scratch=scratch, dqkv=dq))
print(f"fwd {f:.3f} ms {fl/f/1e9:.0f} TF/s | bwd {b:.3f} ms {2.5*fl/b/1e9:.0f} TF/s (causal-exact flops)")
