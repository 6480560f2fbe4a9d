# per-shape timing of the tcgen05 GEMM at the model's shapes (124M, B=32, T=1024)
import sys, torch
sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
from penroz import ops
M, C, V = 32768, 768, 50257
dev = "cuda"
def rnd(*s): return (torch.randn(*s, device=dev) * 0.05).bfloat16()
def t(fn, n=10):
    fn(); torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
x768, x2304, x3072 = rnd(M, 768), rnd(M, 2304), rnd(M, 3072)
wqkv, wproj, wfc, wproj2 = rnd(2304, 768), rnd(768, 768), rnd(3072, 768), rnd(768, 3072)
b768, b2304, b3072 = rnd(768), rnd(2304), rnd(3072)
res = torch.randn(M, 768, device=dev)
o768b, o2304b, o3072b, o3072b2 = [torch.empty(M, n, device=dev, dtype=torch.bfloat16) for n in (768, 2304, 3072, 3072)]
o768f = torch.empty(M, 768, device=dev)
g_fc, g_qkv, g_proj, g_p2 = torch.zeros(3072, 768, device=dev), torch.zeros(2304, 768, device=dev), torch.zeros(768, 768, device=dev), torch.zeros(768, 3072, device=dev)
cases = [
 ("qkv fwd  bias      ", 2*M*2304*768, lambda: ops.gemm(x768, wqkv, bias=b2304, out=o2304b)),
 ("proj fwd resid     ", 2*M*768*768, lambda: ops.gemm(x768, wproj, epilogue=ops.EPI_RESID_F32, bias=b768, aux=res, out=o768f)),
 ("fc fwd gelu        ", 2*M*3072*768, lambda: ops.gemm(x768, wfc, epilogue=ops.EPI_GELU, bias=b3072, out=o3072b, out2=o3072b2)),
 ("proj2 fwd resid    ", 2*M*768*3072, lambda: ops.gemm(x3072, wproj2, epilogue=ops.EPI_RESID_F32, bias=b768, aux=res, out=o768f)),
 ("dgelu dgrad (proj2)", 2*M*3072*768, lambda: ops.gemm(x768, wproj2, b_layout=1, epilogue=ops.EPI_DGELU, aux=o3072b, out=o3072b2)),
 ("fc dgrad           ", 2*M*768*3072, lambda: ops.gemm(x3072, wfc, b_layout=1, out=o768b)),
 ("proj dgrad         ", 2*M*768*768, lambda: ops.gemm(x768, wproj, b_layout=1, out=o768b)),
 ("qkv dgrad          ", 2*M*768*2304, lambda: ops.gemm(x2304, wqkv, b_layout=1, out=o768b)),
 ("proj2 wgrad        ", 2*M*768*3072, lambda: ops.gemm(x768, x3072, a_layout=1, b_layout=1, epilogue=ops.EPI_ACCUM_F32, out=g_p2, accumulate=True)),
 ("fc wgrad           ", 2*M*768*3072, lambda: ops.gemm(x3072, x768, a_layout=1, b_layout=1, epilogue=ops.EPI_ACCUM_F32, out=g_fc, accumulate=True)),
 ("proj wgrad         ", 2*M*768*768, lambda: ops.gemm(x768, x768, a_layout=1, b_layout=1, epilogue=ops.EPI_ACCUM_F32, out=g_proj, accumulate=True)),
 ("qkv wgrad          ", 2*M*768*2304, lambda: ops.gemm(x2304, x768, a_layout=1, b_layout=1, epilogue=ops.EPI_ACCUM_F32, out=g_qkv, accumulate=True)),
]
tot = 0
for name, fl, fn in cases:
    ms = t(fn); tot += ms
    print(f"{name}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.0f} TFLOP/s")
# cuBLAS reference points for the same contraction shapes (library baseline, not used by the product)
for name, a, b in (("torch qkv fwd", x768, wqkv), ("torch fc fwd", x768, wfc), ("torch proj2 fwd", x3072, wproj2)):
    ms = t(lambda: torch.matmul(a, b.t()))
    print(f"{name}: {ms*1e3:8.1f} us  {2*a.shape[0]*b.shape[0]*a.shape[1]/ms/1e9:7.0f} TFLOP/s")
print("sum of 12 GEMMs per layer:", tot, "ms")
