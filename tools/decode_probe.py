"""GPU time vs host time of the native decode token step (bf16 GPT-2 124M), batch 1 and 64, context ~1000.

    python tools/decode_probe.py [B ...]          # run under ncu for the per-kernel launch list
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "penr-oz-neural-network-v3-torch-ddp_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402
from helpers import build_gpt, scratch_cwd  # noqa: E402
from kv_cache import KVCache  # noqa: E402

scratch_cwd()
Bs = [int(a) for a in sys.argv[1:]] or [1, 64]
steps = int(os.environ.get("PROBE_STEPS", "32"))
m = build_gpt("probe", 12, 12, 768, 1024, 50257)
m.to(dtype=torch.bfloat16)
m.to("cuda")
eng, pl = m._engine(), m._plan
for B in Bs:
    kv = KVCache(pl.n_layer)
    kv.paged(B, pl.n_head, pl.head_dim, torch.bfloat16, "cuda", capacity_hint=1024)
    prompt = torch.randint(0, 50257, (B, 960), device="cuda")
    with torch.inference_mode():
        _, ids = eng.prefill(prompt, kv)
        for i in range(4):
            _, ids = eng.decode_step(ids, 960 + i, kv)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            _, ids = eng.decode_step(ids, 964 + i, kv)
        e1.record()
        t_host = time.perf_counter() - t0
        torch.cuda.synchronize()
    print(f"B={B}: GPU {e0.elapsed_time(e1) / steps * 1e3:.1f} us/step, host enqueue {t_host / steps * 1e6:.1f} us/step", flush=True)
    if os.environ.get("PROBE_GRAPH"):
        # the same step replayed from a CUDA graph (fixed position: timing only): what is left when the host and the
        # launch path are out of the picture
        side = torch.cuda.Stream()
        with torch.inference_mode(), torch.cuda.stream(side):
            kv.commit(-1)
            eng.decode_step(ids, 1000, kv)
            kv.commit(-1)
            side.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=side):
                eng.decode_step(ids, 1000, kv)
            kv.commit(-1)
            side.synchronize()
            gr.replay()
            side.synchronize()
            e0.record(side)
            for i in range(steps):
                gr.replay()
            e1.record(side)
            side.synchronize()
        print(f"B={B}: graph replay {e0.elapsed_time(e1) / steps * 1e3:.1f} us/step", flush=True)
