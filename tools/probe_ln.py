# LayerNorm backward at the headline shape (32768 x 768, bf16 dy, fp32 x / residual gradient): time and GB/s of the
# op's algorithmic bytes (dy 2 + x 4 + add 4 + dx 4 + dx_bf16 2 = 16 B per element). PENROZ_LN_BWD=two-kernel for A/B.
import os, sys, torch
sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from penroz import ops
dev = "cuda"
for M, C in ((32768, 768), (32768, 1024), (16384, 1600), (1024, 768)):
    x = torch.randn(M, C, device=dev); g = torch.randn(C, device=dev); b = torch.randn(C, device=dev)
    res = torch.randn(M, C, device=dev)
    y, mean, rstd = ops.layernorm_fwd(x, g, b, out_dtype=torch.bfloat16)
    dy = torch.randn(M, C, device=dev).bfloat16()
    dg, db, dc = (torch.zeros(C, device=dev) for _ in range(3))
    for _ in range(3): ops.layernorm_bwd(dy, x, g, mean, rstd, add=res, dgamma=dg, dbeta=db, dcol=dc, want_bf16=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n): ops.layernorm_bwd(dy, x, g, mean, rstd, add=res, dgamma=dg, dbeta=db, dcol=dc, want_bf16=True)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    print(f"ln_bwd {M}x{C}: {us:.1f} us  {16.0 * M * C / us / 1e3:.0f} GB/s")
