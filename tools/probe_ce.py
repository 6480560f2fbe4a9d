# Cross-entropy rows kernel at the LM-head chunk shape (5461 x 50257 bf16, row pitch 50304): time and GB/s of the
# algorithmic bytes (row read once + gradient written once). PENROZ_CE_GENERIC=1 for the two-pass generic kernel.
import sys, torch
sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from penroz import ops
dev = "cuda"
M, V, ld = 5461, 50257, 50304
src = (torch.randn(M, ld, device=dev) * 3).bfloat16()
tgt = torch.randint(0, V, (M,), device=dev)
nv = torch.zeros(1, device=dev, dtype=torch.int32); rows = torch.empty(M, device=dev)
ops.ce_count_valid(tgt, nv)
bufs = [src.clone() for _ in range(6)]   # 3.3 GB: every launch reads cold rows
for b in bufs[:2]: ops.ce_rows(b, tgt, nv, rows, V=V, grad_scale=1.0, write_grad=True)
for b in bufs: b.copy_(src)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for b in bufs: ops.ce_rows(b, tgt, nv, rows, V=V, grad_scale=1.0, write_grad=True)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / len(bufs) * 1e3
print(f"ce_rows {M}x{V}: {us:.1f} us  {4.0 * M * V / us / 1e3:.0f} GB/s")
