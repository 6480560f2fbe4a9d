"""Benchmark of the hot path: GPT-2 124M (12L/768d, T=1024, untied 50257 head = 163 M params) bf16-autocast DDP
training step on synthetic tokens — BASELINE.json's metric, configs[1].

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config 124m|350m-accum|xl|decode]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (N > 1)

One "step" = one optimizer step of the reference's train loop ("epoch", neural_net_model.py:614-677): zero_grad,
num_steps x [forward, loss, backward], bucketed gradient all-reduce (N > 1), AdamW — per rank micro-batch B x 1024
tokens (weak scaling).

  value  tokens/s with the step's token batches already resident in HBM (CUDA events, max over ranks)
  e2e    the same metric through the public API a user calls — NeuralNetworkModel.train_model() reading uint16
         shards on the host (H2D of inputs + D2H of the loss inside every step), timed per epoch by the wall
         clock after torch.cuda.synchronize(), which is how the reference times itself (neural_net_model.py:678-685)
  roofline      the kernel family with the largest share of the step, CUDA events around each of its launches
  cpu_baseline  the reference's own CPU train_model (oracle/_ref, staged unmodified) on the host cores, bounded sample

--config (default 124m = BASELINE.json configs[1], the one `metric` is quoted on) also measures the other configurations
BASELINE.json names, same JSON line: 350m-accum = configs[2] (24L/1024d, 4 accumulation micro-steps per optimizer
step, PENROZ_ACCUM_SYNC=reference|correct), xl = configs[4] (48L/1600d, 8 x 1024 tokens per rank), decode = configs[3]
(KV-cache generation grid: prompts 1/512/960/1023 out to the 1024-token context, greedy and top-k 50, batch 1 and 64).
"""
import argparse
import copy
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "penr-oz-neural-network-v3-torch-ddp_b200")
REF = os.path.join(ROOT, "oracle", "_ref")

CONFIGS = {
    "124m": dict(L=12, H=12, C=768, T=1024, V=50257, B=32, accum=1, metric="gpt2_124m_train_tokens_per_sec",
                 name="GPT-2 124M (12L/768d/12h, T=1024, V=50257 untied head, 163,037,184 params)"),
    "350m-accum": dict(L=24, H=16, C=1024, T=1024, V=50257, B=32, accum=4, metric="gpt2_350m_train_tokens_per_sec",
                       name="GPT-2 medium 350M (24L/1024d/16h, T=1024, V=50257 untied head, 406,286,336 params), "
                            "4 accumulation micro-steps per optimizer step"),
    "xl": dict(L=48, H=25, C=1600, T=1024, V=50257, B=8, accum=1, metric="gpt2_xl_train_tokens_per_sec",
               name="GPT-2 XL 1.5B (48L/1600d/25h, T=1024, V=50257 untied head, 1,638,022,400 params)"),
    "decode": dict(L=12, H=12, C=768, T=1024, V=50257, B=64, accum=1, metric="gpt2_124m_decode_tokens_per_sec",
                   name="GPT-2 124M KV-cache generation via generate_tokens, batch 1 and 64, 1024-token context"),
}
UNIT = "tokens/s"
TRAFFIC_JSON = os.path.join(ROOT, "profiles", "r02p_ncu_launch_list_step.json")
TRAFFIC_JSON_OLD = os.path.join(ROOT, "profiles", "r02_ncu_launch_list_step.json")


def _our_paths():
    for p in (PKG, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)


def f_tok(L, C, V, T):
    """Algorithmic training flops per token: 6(12 L C^2 + V C) + 6 L C T, causal-exact attention."""
    return 6 * (12 * L * C * C + V * C) + 6 * L * C * T


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def ncu_traffic_per_launch(family: str):
    """Mean DRAM bytes per launch of a kernel family from the committed ncu launch list of this command. The list is a
    capture of a past commit (named in its `commit` field); it is re-folded when the kernels change."""
    for path in (TRAFFIC_JSON, TRAFFIC_JSON_OLD):
        try:
            doc = json.load(open(path))
            fams = doc["families"]
        except Exception:
            continue
        key = "gemm_bf16" if family.startswith("gemm") else "attn_"
        sel = [f for f in fams if f["kernel"].startswith(key) and "delta" not in f["kernel"] and "finish" not in f["kernel"]]
        n = sum(f["launches"] for f in sel)
        if n:
            src = f"{os.path.relpath(path, ROOT)} (dram__bytes_read+write per launch, family mean; captured at commit " \
                  f"{doc.get('commit', 'end of round 1')})"
            return sum(f["dram_read_MB"] + f["dram_write_MB"] for f in sel) * 1e6 / n, src
    return None, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 6:
                continue
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except ValueError:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def build_model(model_id, cfg, seed=0):
    import torch
    import neural_net_model as nnm
    from penroz import dsl
    torch.manual_seed(seed)
    return nnm.NeuralNetworkModel(model_id, dsl.Mapper(dsl.gpt2_layers(cfg["L"], cfg["H"], cfg["C"], cfg["T"], cfg["V"]),
                                                       copy.deepcopy(dsl.DEFAULT_ADAMW)))


# ---------------------------------------------------------------------------------------------------- reference arm
def host_threads() -> int:
    """All the host threads this process may use (torch.distributed.run exports OMP_NUM_THREADS=1: do not inherit it)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


def reference_cpu_train(cfg, steps: int, warmup: int, sample_B: int = 1):
    """The UNMODIFIED reference (oracle/_ref, staged by oracle/stage_ref.py) training on the host cores through its own
    public API: NeuralNetworkModel.train_model (neural_net_model.py:552-722), fp32, world 1, `warmup + steps` epochs of
    ONE (sample_B x T) micro-batch of the full model each — a bounded sample of the workload (the bench's per-rank
    batch is B x T). Per-epoch time is the reference's own clock (progress[i].durationInSecs).
    Returns dict(rate, ms, cores, threads, sample, kind)."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_cuda
    work = tempfile.mkdtemp(prefix="penroz_bench_refcpu_")
    ref_nnm, Mapper = ref_cuda._import_reference(work)
    import torch
    n_thr = host_threads()
    torch.set_num_threads(n_thr)
    T, V = cfg["T"], cfg["V"]
    epochs = warmup + steps
    ref_cuda.write_shards(f"synth{V}", V, (epochs + 2) * sample_B * T + 2)
    torch.manual_seed(0)
    model = ref_nnm.NeuralNetworkModel("refcpu", Mapper(copy.deepcopy(ref_cuda.gpt_dsl(cfg["L"], cfg["H"], cfg["C"], T, V)),
                                                        copy.deepcopy(ref_cuda.OPTIM)))
    model.train_model(f"synth{V}", 0, epochs, sample_B, T, sample_B)
    secs = [p["durationInSecs"] for p in model.progress][warmup:]
    # the last epoch also keeps gradients of every activation for the end-of-training statistics; it is one of the K
    mean = sum(secs) / len(secs)
    sample = (f"{len(secs)} optimizer step(s) (after {warmup} warm-up) of the full {cfg['name'].split(' (')[0]} model on "
              f"{sample_B}x{T} tokens each: unmodified reference train_model on the CPU, fp32, torch {torch.__version__}, "
              f"{torch.get_num_threads()} threads")
    return {"rate": sample_B * T / mean, "ms": mean * 1e3, "cores": os.cpu_count(), "threads": torch.get_num_threads(),
            "sample": sample, "kind": "reference", "losses": [p["cost"] for p in model.progress]}


def port_cpu_train(cfg, steps: int, warmup: int, sample_B: int = 1):
    """Fallback when oracle/_ref is not staged: the numpy oracle port of the step."""
    _our_paths()
    import numpy as np
    from oracle.gpt2_oracle import Gpt2Oracle
    L, H, C, T, V = cfg["L"], cfg["H"], cfg["C"], cfg["T"], cfg["V"]
    rng = np.random.default_rng(0)
    state = {"layers.0.0.weight": rng.normal(0, 0.02, (V, C)), "layers.0.1.weight": rng.normal(0, 0.02, (T, C))}
    for l in range(L):
        pre = f"layers.{2 + l}."
        state.update({pre + "0.0.weight": np.ones(C), pre + "0.0.bias": np.zeros(C),
                      pre + "0.1.weight": rng.normal(0, 0.02, (3 * C, C)), pre + "0.1.bias": np.zeros(3 * C),
                      pre + "0.3.weight": rng.normal(0, 0.004, (C, C)), pre + "0.3.bias": np.zeros(C),
                      pre + "1.0.weight": np.ones(C), pre + "1.0.bias": np.zeros(C),
                      pre + "1.1.weight": rng.normal(0, 0.02, (4 * C, C)), pre + "1.1.bias": np.zeros(4 * C),
                      pre + "1.3.weight": rng.normal(0, 0.004, (C, 4 * C)), pre + "1.3.bias": np.zeros(C)})
    state.update({f"layers.{2 + L}.weight": np.ones(C), f"layers.{2 + L}.bias": np.zeros(C),
                  f"layers.{3 + L}.weight": rng.normal(0, 0.03, (V, C))})
    model = Gpt2Oracle(state, L, H, dtype=np.float32)
    toks = np.random.default_rng(1234).integers(0, V, size=sample_B * T + 1)
    x, y = toks[:-1].reshape(sample_B, T), toks[1:].reshape(sample_B, T)

    def step():
        _, loss = model.forward(x, y, keep=True)
        model.adamw_step(model.backward())
        return loss

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    secs = (time.perf_counter() - t0) / steps
    return {"rate": sample_B * T / secs, "ms": secs * 1e3, "cores": os.cpu_count(), "threads": host_threads(),
            "sample": f"{steps} optimizer step(s) of the full model on {sample_B}x{T} tokens (numpy fp32 port of the torch "
                      f"calls, BLAS threads = all cores)", "kind": "port"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config if args.config != "decode" else "124m"]
    # bounded so the whole run ends within a few minutes: the 124M step on 1x1024 tokens takes ~2-3 s on 8-16 cores
    steps = max(1, args.steps)
    warmup = max(0, args.warmup)
    if args.config in ("350m-accum", "xl"):
        steps, warmup = min(steps, 3), min(warmup, 1)
    staged = os.path.exists(os.path.join(REF, "neural_net_model.py"))
    res = reference_cpu_train(cfg, steps, warmup) if staged else port_cpu_train(cfg, steps, warmup)
    line = {"impl": "reference", "metric": cfg["metric"], "value": res["rate"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": res["ms"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{cfg['name']} training step, the reference's own CPU path on the host cores "
                                   f"({'unmodified reference staged in oracle/_ref' if staged else 'numpy port: oracle/_ref not staged'})",
                       "sample": res["sample"], "threads": res["threads"]},
            "cpu_baseline": {"value": res["rate"], "unit": UNIT, "cores": res["cores"], "threads": res["threads"],
                             "kind": res["kind"], "sample": res["sample"]},
            "e2e": {"value": res["rate"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def cpu_baseline_subprocess(config: str):
    """The reference arm in a process of its own (its modules carry the same names as the drop-ins loaded here):
    1 warm-up + 2 timed steps, about 10-30 s of CPU work."""
    try:
        res = subprocess.run([sys.executable, os.path.abspath(__file__), "--impl", "reference", "--steps", "2",
                              "--warmup", "1", "--config", config], capture_output=True, text=True, timeout=900,
                             env={k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK",
                                                                                   "OMP_NUM_THREADS")})
        for ln in reversed(res.stdout.strip().splitlines()):
            if ln.startswith("{"):
                return json.loads(ln)["cpu_baseline"]
        return {"error": (res.stderr or "no output")[-300:]}
    except Exception as exc:
        return {"error": repr(exc)}


# ---------------------------------------------------------------------------------------------------- decode
def _decode_bytes(cfg, B, ctx, es):
    """HBM bytes one token step must move: every weight once + the cached K/V of every sequence."""
    w = (12 * cfg["L"] * cfg["C"] ** 2 + cfg["V"] * cfg["C"]) * es
    kv = B * ctx * 2 * cfg["L"] * cfg["C"] * es
    return w + kv


def decode_case(torch, model, cfg, B, prompt_len, new, es, greedy=True, seed=7):
    """generate_tokens(prompt_len + new) through the public API; returns whole-call tokens/s and the incremental token
    step alone ((prefill + new) - (prefill + 1)) with the HBM fraction at its mean context."""
    V = cfg["V"]
    gen = torch.Generator().manual_seed(seed)
    prompt = torch.randint(0, V, (B, prompt_len), generator=gen).tolist()
    kw = dict(temperature=0.0) if greedy else dict(temperature=1.0, top_k=50)
    torch.manual_seed(42)
    model.generate_tokens(prompt, cfg["T"], min(new, 4), **kw)      # warm-up at the timed shape (first-use costs)
    torch.cuda.synchronize()
    torch.manual_seed(42)
    t0 = time.perf_counter()
    toks = model.generate_tokens(prompt, cfg["T"], new, **kw)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    assert len(toks) == prompt_len + new
    out = {"B": B, "prompt": prompt_len, "new": new, "mode": "greedy" if greedy else "top_k=50 T=1.0 seed 42",
           "tokens_per_s": B * new / dt, "seconds": dt}
    if new > 1:
        torch.manual_seed(42)
        t0 = time.perf_counter()
        model.generate_tokens(prompt, cfg["T"], 1, **kw)
        torch.cuda.synchronize()
        dt1 = time.perf_counter() - t0
        step_s = max(dt - dt1, 1e-9) / (new - 1)
        ctx = prompt_len + (new + 1) / 2.0
        byts = _decode_bytes(cfg, B, ctx, es)
        peak = measured_peaks()[0]["hbm_gbs"]
        out.update({"prefill_ms": dt1 * 1e3, "step_ms": step_s * 1e3, "step_tokens_per_s": B / step_s,
                    "mean_context": ctx, "hbm_gbps": byts / step_s / 1e9, "hbm_frac": byts / step_s / 1e9 / peak})
    return out


def decode_bench(torch, nnm, cfg, full_grid=False):
    """BASELINE config 4: KV-cache generation through generate_tokens(). Default: the 1024-token context (960-token
    prompt + 64 new tokens, mean context 992), batch 1 and 64, bf16 (HF-import style) and fp32 parameters.
    full_grid: SURVEY 8d config 4 — prompts 1 / 512 / 960 / 1023, greedy and top-k 50, and one run crossing block_size."""
    out = {}
    for tag, dtype in (("bf16", torch.bfloat16), ("f32", torch.float32)):
        model = build_model(f"dec_{tag}", cfg)
        es = 2 if dtype == torch.bfloat16 else 4
        if dtype != torch.float32:
            model.to(dtype=dtype)
        model.to("cuda")
        for B in (1, 64):
            out[f"{tag}_b{B}"] = decode_case(torch, model, cfg, B, 960, 64, es)
            if full_grid:
                out[f"{tag}_b{B}_p512"] = decode_case(torch, model, cfg, B, 512, 512, es)
                out[f"{tag}_b{B}_p1"] = decode_case(torch, model, cfg, B, 1, 1023 if B == 1 else 255, es)
                out[f"{tag}_b{B}_p1023"] = decode_case(torch, model, cfg, B, 1023, 1, es)
                out[f"{tag}_b{B}_topk50"] = decode_case(torch, model, cfg, B, 960, 64, es, greedy=False)
                if tag == "bf16":   # crosses block_size: clear + re-prefill of the cropped context on every later token
                    out[f"{tag}_b{B}_cross_block"] = decode_case(torch, model, cfg, B, 1016, 16, es)
                if tag == "f32":    # TURBO_QUANT_KV_CACHE=1: int8 codes + per-token scales, de-quantised inside attention
                    import kv_cache
                    kv_cache.TURBO_QUANT_ENABLED = True
                    try:
                        out[f"{tag}_b{B}_int8kv"] = decode_case(torch, model, cfg, B, 960, 64, es)
                    finally:
                        kv_cache.TURBO_QUANT_ENABLED = False
        del model
        torch.cuda.empty_cache()
    out["workload"] = ("GPT-2 124M generate_tokens, greedy, 960-token prompt + 64 new tokens (prefill included in "
                       "tokens_per_s; step_ms = the incremental token step alone at mean context 992 of 1024)")
    return out


# ---------------------------------------------------------------------------------------------------- our arm
class KernelTimer:
    """CUDA-event pairs around every launch of the C-ABI families that matter, for the roofline object."""

    def __init__(self, ops_mod, torch):
        self.ops, self.torch, self.records, self.saved = ops_mod, torch, [], {}

    def __enter__(self):
        torch, ops = self.torch, self.ops

        def wrap(name, flops_fn):
            fn = getattr(ops, name)
            self.saved[name] = fn

            def timed(*a, **k):
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                out = fn(*a, **k)
                e.record()
                self.records.append((name, flops_fn(*a, **k), s, e))
                return out
            setattr(ops, name, timed)

        def gemm_flops(a, b, **k):
            al, bl = k.get("a_layout", 0), k.get("b_layout", 0)
            M, K = (a.shape[0], a.shape[1]) if al == 0 else (a.shape[1], a.shape[0])
            N = b.shape[0] if bl == 0 else b.shape[1]
            return 2.0 * M * N * K

        def attn_fwd_flops(qkv, H, **k):
            B, T, C3 = qkv.shape
            return 4.0 * B * T * T * (C3 // 3) / 2      # causal-exact QK^T + PV

        def attn_bwd_flops(qkv, out, dout, lse, H, **k):
            B, T, C3 = qkv.shape
            return 2.5 * 4.0 * B * T * T * (C3 // 3) / 2

        wrap("gemm", gemm_flops)
        wrap("attn_fwd", attn_fwd_flops)
        wrap("attn_bwd", attn_bwd_flops)
        return self

    def __exit__(self, *exc):
        for name, fn in self.saved.items():
            setattr(self.ops, name, fn)

    def summary(self):
        self.torch.cuda.synchronize()
        fam = {}
        for name, flops, s, e in self.records:
            key = "gemm_bf16_tc" if name == "gemm" else "causal_attention"
            d = fam.setdefault(key, {"ms": 0.0, "flops": 0.0, "launches": 0})
            d["ms"] += s.elapsed_time(e)
            d["flops"] += flops
            d["launches"] += 1
        return fam


def run_decode_config(args):
    _our_paths()
    import torch
    import neural_net_model as nnm
    from penroz import _lib
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    _lib.require_device()
    cfg = CONFIGS["decode"]
    work = tempfile.mkdtemp(prefix=f"penroz_bench_dec_r{rank}_")
    os.chdir(work)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = _lib.lib().launch_count()
    grid = decode_bench(torch, nnm, cfg, full_grid=True)     # replicas only: every rank runs the same grid
    launches = _lib.lib().launch_count() - n0
    if rank != 0:
        return
    clocks = sampler.stop()
    head = grid["bf16_b64"]
    peaks, which = measured_peaks()
    line = {"metric": cfg["metric"], "value": world * head["step_tokens_per_s"], "unit": UNIT, "n_gpus": world,
            "steps": 64, "warmup": 4, "ms_per_step": head["step_ms"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": cfg["name"] + "; value = bf16 batch-64 token step at mean context 992 "
                                                 "(replicas only: decode does not shard, one replica per GPU)",
                       "l2_policy": "inputs larger than L2: each step streams 247 MB of weights + 2.4 GB of K/V"},
            "e2e": {"value": world * head["tokens_per_s"], "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0,
                    "api": "NeuralNetworkModel.generate_tokens (host token lists in, host token list out; prompt "
                           "upload + prefill + 64 token steps + final read-back inside the timed call)"},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"kernel": "decode token step (all launches)", "bound": "hbm", "achieved": head["hbm_gbps"],
                         "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": head["hbm_frac"], "traffic": None,
                         "peak_source": which},
            "decode": grid}
    print(json.dumps(line))


def run_ours(args):
    _our_paths()
    import numpy as np
    import torch
    import torch.distributed as dist
    import ddp
    import neural_net_model as nnm
    from penroz import _lib, ops, shards
    from penroz.grad_sync import GradSync, gpt2_segments

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group(backend="nccl", device_id=device)
    _lib.require_device()
    cfg = CONFIGS[args.config]
    B, T, V, accum = cfg["B"], cfg["T"], cfg["V"], cfg["accum"]
    sync_mode = os.environ.get("PENROZ_ACCUM_SYNC", "reference")
    work = tempfile.mkdtemp(prefix=f"penroz_bench_r{rank}_")
    os.chdir(work)
    nnm.NeuralNetworkModel.SHM_PATH = os.path.join(work, "shm")
    os.makedirs(nnm.NeuralNetworkModel.SHM_PATH, exist_ok=True)
    model = build_model("bench" + args.config.replace("-", ""), cfg)
    n_params = model.num_params
    model.to("cuda")
    eng = model._engine()
    eng.setup_training(use_amp=True)
    if world > 1:
        eng.grad_sync = GradSync(eng.flat.g, gpt2_segments(eng.flat, eng.plan))
        eng.grad_sync.broadcast_(eng.flat.p)
        eng.flat.refresh_shadow()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---------------- value: inputs resident in HBM; a different token batch every micro-step (nothing cacheable
    # matters: one 124M step streams ~25 GB of activations + 3 GB of optimizer state through a 126 MB L2)
    n_batches = 4
    gen = torch.Generator(device="cpu").manual_seed(1234 + rank)
    toks = [torch.randint(0, V, (B * T + 1,), generator=gen) for _ in range(n_batches)]
    xs = [t[:-1].view(B, T).to(device) for t in toks]
    ys = [t[1:].view(B, T).to(device) for t in toks]

    def step(i):
        # the body of _FusedStepper.run_epoch + apply_update (neural_net_model.py), inputs already on the device
        eng.zero_grad()
        loss = None
        for s in range(accum):
            sync_here = world > 1 and ((s == 0) if sync_mode == "reference" else (s == accum - 1))
            j = (i * accum + s) % n_batches
            part = eng.train_micro_step(xs[j], ys[j], 1.0 / accum, sync_grads=sync_here)
            if sync_here and s != accum - 1:
                eng.grad_sync.wait_all()
            loss = part / accum if loss is None else loss + part / accum
        # as _FusedStepper.apply_update does on an ordinary (non-statistics) epoch: rank 0 records the progress entry
        # (weight_upd_ratio measured inside the AdamW pass), the pass clears the gradients it consumed
        eng.optimizer_step(track_update=(rank == 0), clear_grads=True)
        return loss

    for i in range(args.warmup):
        step(i)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = _lib.lib().launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(args.steps):
        loss = step(i)
    ev1.record()
    barrier()
    launches = _lib.lib().launch_count() - n0
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = ms.item() / args.steps
    tokens_per_step = world * accum * B * T
    value = tokens_per_step / (ms_per_step / 1e3)
    final_loss = loss.item()

    parts = os.environ.get("PENROZ_BENCH_PARTS", "all")   # "value": only the timed loop (short runs under ncu)
    if parts == "value":
        if rank == 0:
            print(json.dumps({"metric": cfg["metric"], "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                              "warmup": args.warmup, "ms_per_step": ms_per_step, "gpu_launches": int(launches),
                              "note": "PENROZ_BENCH_PARTS=value: timed loop only"}))
        if world > 1:
            dist.destroy_process_group()
        return
    # ---------------- roofline: event pairs around each launch of the big kernel families, one extra step
    roofline = None
    if rank == 0:
        with KernelTimer(ops, torch) as kt:
            torch.cuda.synchronize()
            s_all, e_all = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s_all.record()
            step(0)                       # every rank runs this step (it contains the gradient all-reduce)
            e_all.record()
            fam = kt.summary()
        step_ms = s_all.elapsed_time(e_all)
        peaks, which = measured_peaks()
        top = max(fam.items(), key=lambda kv: kv[1]["ms"])
        name, d = top
        achieved = d["flops"] / (d["ms"] / 1e3) / 1e12
        peak = peaks["bf16_tflops_sustained"]
        traffic, traffic_src = ncu_traffic_per_launch(name) if args.config == "124m" else (None, None)
        roofline = {"kernel": name, "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                    "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": f"{which} (sustained: kernel timed inside a long step)",
                    "share_of_step": d["ms"] / step_ms, "launches_per_step": d["launches"],
                    "families_ms": {k: round(v["ms"], 3) for k, v in fam.items()},
                    "families_tflops": {k: round(v["flops"] / (v["ms"] / 1e3) / 1e12, 1) for k, v in fam.items()},
                    "instrumented_step_ms": step_ms}
    else:
        step(0)
    barrier()

    # ---------------- e2e: the public API, host shards in, loss out, every step
    e2e = None
    eng.release_workspace()
    e2e_epochs = args.warmup + args.steps + 1
    if rank == 0:
        need = (e2e_epochs * accum * world + 2) * B * T + 2
        shards.write_synthetic_shards("synth50257", V, n_shards=1, tokens_per_shard=max(need, 2 ** 22), folder="data")
    barrier()
    if world > 1:
        # every rank reads the same shard files: share rank 0's folder
        obj = [work if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        os.chdir(obj[0])
    # W warm-up epochs + K timed epochs + 1 final epoch that is NOT timed: on its last epoch train_model (like the
    # reference, neural_net_model.py:642-646,721) keeps every activation and gradient on rank 0 for the stats
    os.environ["PENROZ_TRAIN_PRECISION"] = "amp"
    step_size = max(1, B // (accum * world))
    model.train_model("synth50257", 0, e2e_epochs, B, T, step_size)
    barrier()
    if rank == 0:
        durs = [p["durationInSecs"] for p in model.progress][args.warmup:args.warmup + args.steps]
        e2e_secs = torch.tensor([sum(durs)], device=device)
    else:
        e2e_secs = torch.tensor([0.0], device=device)
    if world > 1:
        dist.all_reduce(e2e_secs, op=dist.ReduceOp.MAX)
    if rank == 0:
        n = len(durs)
        e2e = {"value": tokens_per_step * n / e2e_secs.item(), "unit": UNIT,
               "h2d_bytes_per_step": accum * 2 * B * T * 4, "d2h_bytes_per_step": 4,
               "api": "NeuralNetworkModel.train_model (uint16 shards on the host -> pinned int32 window -> async H2D -> widened on the device -> "
                      "fused step -> cost.item()); per-epoch wall clock after torch.cuda.synchronize(), as the reference "
                      "logs it; progress (incl. weight_upd_ratio) recorded every epoch; the final stats-collecting "
                      "epoch is run but not timed",
               "steps": n}

    decode = None
    if rank == 0 and world == 1 and args.config == "124m":
        del model, eng
        torch.cuda.empty_cache()
        try:
            decode = decode_bench(torch, nnm, cfg)
        except Exception as exc:  # the train metric is the headline; report, do not lose the line
            decode = {"error": repr(exc)}
    if rank == 0:
        ft = f_tok(cfg["L"], cfg["C"], V, T)
        peaks, which = measured_peaks()
        line = {"metric": cfg["metric"], "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": f"{cfg['name']} bf16-autocast DDP training step, AdamW lr 6e-4 wd 0.01, "
                                       f"num_steps={accum}" + (f", accumulation sync mode '{sync_mode}'" if accum > 1 else ""),
                           "per_gpu_batch": B, "seq_len": T, "micro_steps": accum, "params": n_params,
                           "global_batch_tokens": tokens_per_step, "parallelism": f"dp{world}",
                           "l2_policy": "inputs larger than L2: each step streams tens of GB of activations and the "
                                        "optimizer state through the 126 MB L2; token batches rotate every micro-step"},
                "mfu_vs_2.25PF": value * ft / (world * 2.25e15),
                "mfu_vs_measured_sustained": value * ft / (world * peaks["bf16_tflops_sustained"] * 1e12),
                "flops_per_token": ft, "final_loss": final_loss,
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
                "decode": decode}
        if world == 1 and args.config == "124m":
            line["cpu_baseline"] = cpu_baseline_subprocess(args.config)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _reserve_stdout():
    """The contract is ONE JSON line on stdout. Libraries (NCCL's version banner and NCCL_DEBUG=INFO log, torch-elastic
    notices) also write to file descriptor 1, so keep a private duplicate of it for the result line and point fd 1 at
    stderr."""
    global print
    sys.stdout.flush()
    keep = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    builtin_print = print

    def print(*a, **kw):        # noqa: A001 - module-local: every print() in this file is a result line
        kw.setdefault("file", keep)
        kw.setdefault("flush", True)
        builtin_print(*a, **kw)


def main():
    _reserve_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="124m", choices=list(CONFIGS))
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.config == "decode":
        run_decode_config(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
