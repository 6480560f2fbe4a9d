"""Benchmark of the hot path: GPT-2 124M (12L/768d, T=1024, untied 50257 head = 163 M params) bf16-autocast DDP
training step on synthetic tokens — BASELINE.json's metric, configs[1].

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (N > 1)

One "step" = one optimizer step of the reference's train loop with num_steps = 1: zero_grad, forward, loss,
backward, bucketed gradient all-reduce (N > 1), AdamW — per rank batch 32 x 1024 tokens (weak scaling).

  value  tokens/s with the step's token batch already resident in HBM (CUDA events, max over ranks)
  e2e    the same metric through the public API a user calls — NeuralNetworkModel.train_model() reading uint16
         shards on the host (H2D of inputs + D2H of the loss inside every step), timed per epoch by the wall
         clock after torch.cuda.synchronize(), which is how the reference times itself (neural_net_model.py:678-685)
  roofline      the kernel family with the largest share of the step, CUDA events around each of its launches
  cpu_baseline  the numpy oracle port of the same step on the host cores, bounded sample (rank 0, N = 1)
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
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

CFG = dict(L=12, H=12, C=768, T=1024, V=50257, B=32)
METRIC = "gpt2_124m_train_tokens_per_sec"
UNIT = "tokens/s"


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
    """Mean DRAM bytes per launch of a kernel family, from the committed ncu launch list of this same command."""
    path = os.path.join(ROOT, "profiles", "r01f_ncu_launch_list_step.json")
    try:
        fams = json.load(open(path))["families"]
    except Exception:
        return None
    key = "gemm_bf16" if family.startswith("gemm") else "attn_"
    sel = [f for f in fams if f["kernel"].startswith(key) and "delta" not in f["kernel"] and "finish" not in f["kernel"]]
    n = sum(f["launches"] for f in sel)
    return (sum(f["dram_read_MB"] + f["dram_write_MB"] for f in sel) * 1e6 / n) if n else None


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


def build_model(model_id="bench124m", cfg=CFG, seed=0):
    import torch
    import neural_net_model as nnm
    from penroz import dsl
    torch.manual_seed(seed)
    return nnm.NeuralNetworkModel(model_id, dsl.Mapper(dsl.gpt2_layers(cfg["L"], cfg["H"], cfg["C"], cfg["T"], cfg["V"]),
                                                       copy.deepcopy(dsl.DEFAULT_ADAMW)))


# ---------------------------------------------------------------------------------------------------- CPU baseline
def cpu_reference_step_rate(steps: int, warmup: int, cfg=CFG, sample_B=1, sample_T=None):
    """The oracle port of the step (numpy fp32, BLAS on every host core) on a bounded sample of the workload:
    one optimizer step of the full 124M model on a (sample_B x T) micro-batch. Returns (tokens/s, cores, sample)."""
    import numpy as np
    from oracle.gpt2_oracle import Gpt2Oracle
    L, H, C, T, V = cfg["L"], cfg["H"], cfg["C"], sample_T or cfg["T"], cfg["V"]
    rng = np.random.default_rng(0)
    state = {"layers.0.0.weight": rng.normal(0, 0.02, (V, C)), "layers.0.1.weight": rng.normal(0, 0.02, (cfg["T"], C))}
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
    return sample_B * T / secs, os.cpu_count(), f"{steps} optimizer step(s) of the full 124M model on {sample_B}x{T} tokens (numpy fp32 port, BLAS threads = all cores)", secs * 1e3


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 3))
    rate, cores, sample, ms = cpu_reference_step_rate(steps, 1)
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
            "warmup": 1, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "GPT-2 124M (12L/768d/T=1024/V=50257 untied) training step, reference CPU path "
                                   "(numpy port of the torch calls; the Python reference itself cannot travel to the GPU box)",
                       "sample": sample},
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def decode_bench(torch, nnm, device):
    """BASELINE config 4: greedy KV-cache generation through the public generate_tokens() API (prefill of a 512-token
    prompt + 64 new tokens), batch 1 and 64, bf16 (HF-import style) and fp32 parameters. tokens/s = B * new / wall."""
    out = {}
    V = CFG["V"]
    for tag, dtype in (("bf16", torch.bfloat16), ("f32", torch.float32)):
        model = build_model(f"dec_{tag}")
        if dtype != torch.float32:
            model.to(dtype=dtype)
        model.to("cuda")
        for B in (1, 64):
            gen = torch.Generator().manual_seed(7)
            prompt = torch.randint(0, V, (B, 512), generator=gen).tolist()
            model.generate_tokens(prompt, 1024, 4, temperature=0.0)     # warm-up at the timed shape (first-use costs)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            toks = model.generate_tokens(prompt, 1024, 64, temperature=0.0)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            assert len(toks) == 512 + 64
            # the incremental step alone: (prefill + 64 tokens) - (prefill + 1 token), and what it streams from HBM:
            # every weight once per step ((12 L C^2 + V C) elements) + the cached K/V of every sequence
            t0 = time.perf_counter()
            model.generate_tokens(prompt, 1024, 1, temperature=0.0)
            torch.cuda.synchronize()
            dt1 = time.perf_counter() - t0
            step_s = max(dt - dt1, 1e-9) / 63
            es = 2 if dtype == torch.bfloat16 else 4
            w_bytes = (12 * CFG["L"] * CFG["C"] ** 2 + V * CFG["C"]) * es
            kv_bytes = B * (512 + 32) * 2 * CFG["L"] * CFG["C"] * es          # mean context over the 64 steps
            peak = measured_peaks()[0]["hbm_gbs"]
            out[f"{tag}_b{B}"] = {"tokens_per_s": B * 64 / dt, "seconds": dt, "step_ms": step_s * 1e3,
                                  "step_tokens_per_s": B / step_s,
                                  "hbm_gbps": (w_bytes + kv_bytes) / step_s / 1e9,
                                  "hbm_frac": (w_bytes + kv_bytes) / step_s / 1e9 / peak}
        del model
        torch.cuda.empty_cache()
    out["workload"] = "GPT-2 124M greedy generate_tokens: 512-token prompt + 64 new tokens (prefill included), context 1024"
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


def run_ours(args):
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
    # rank 0 must print exactly ONE line: keep NCCL's version banner (NCCL_DEBUG=VERSION/INFO) off stdout
    os.environ["NCCL_DEBUG"] = os.environ.get("PENROZ_NCCL_DEBUG", "WARN")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group(backend="nccl", device_id=device)
    _lib.require_device()
    cfg = CFG
    B, T, V = cfg["B"], cfg["T"], cfg["V"]
    work = tempfile.mkdtemp(prefix=f"penroz_bench_r{rank}_")
    os.chdir(work)
    nnm.NeuralNetworkModel.SHM_PATH = os.path.join(work, "shm")
    os.makedirs(nnm.NeuralNetworkModel.SHM_PATH, exist_ok=True)
    model = build_model()
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

    # ---------------- value: inputs resident in HBM; a different token batch every step (each larger than nothing
    # cacheable matters here: one step streams ~25 GB of activations + 3 GB of optimizer state through a 126 MB L2)
    n_batches = 4
    gen = torch.Generator(device="cpu").manual_seed(1234 + rank)
    toks = [torch.randint(0, V, (B * T + 1,), generator=gen) for _ in range(n_batches)]
    xs = [t[:-1].view(B, T).to(device) for t in toks]
    ys = [t[1:].view(B, T).to(device) for t in toks]

    def step(i):
        eng.zero_grad()
        loss = eng.train_micro_step(xs[i % n_batches], ys[i % n_batches], 1.0, sync_grads=(world > 1))
        eng.optimizer_step()
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
    value = world * B * T / (ms_per_step / 1e3)
    final_loss = loss.item()

    parts = os.environ.get("PENROZ_BENCH_PARTS", "all")   # "value": only the timed loop (short runs under ncu)
    if parts == "value":
        if rank == 0:
            print(json.dumps({"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
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
        roofline = {"kernel": name, "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                    "frac": achieved / peak, "traffic": ncu_traffic_per_launch(name),
                    "traffic_source": "profiles/r01f_ncu_launch_list_step.json (dram__bytes_read+write per launch, family mean)",
                    "peak_source": f"{which} (sustained: kernel timed inside a long step)",
                    "share_of_step": d["ms"] / step_ms, "launches_per_step": d["launches"],
                    "families_ms": {k: round(v["ms"], 3) for k, v in fam.items()}, "instrumented_step_ms": step_ms}
    else:
        step(0)
    barrier()

    # ---------------- e2e: the public API, host shards in, loss out, every step
    e2e = None
    eng.release_workspace()
    if rank == 0:
        shards.write_synthetic_shards("synth50257", V, n_shards=1, tokens_per_shard=2 ** 23, folder="data")
    barrier()
    if world > 1:
        # every rank reads the same shard files: share rank 0's folder
        obj = [work if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        os.chdir(obj[0])
    # W warm-up epochs + K timed epochs + 1 final epoch that is NOT timed: on its last epoch train_model (like the
    # reference, neural_net_model.py:642-646,721) keeps every activation and gradient on rank 0 for the stats
    e2e_epochs = args.warmup + args.steps + 1
    os.environ["PENROZ_TRAIN_PRECISION"] = "amp"
    model.train_model("synth50257", 0, e2e_epochs, B, T, max(1, B // world))
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
        e2e = {"value": world * B * T * n / e2e_secs.item(), "unit": UNIT,
               "h2d_bytes_per_step": 2 * B * T * 8, "d2h_bytes_per_step": 4,
               "api": "NeuralNetworkModel.train_model (uint16 shards on the host -> pinned int64 batch -> async H2D -> fused "
                      "step -> cost.item()); per-epoch wall clock after torch.cuda.synchronize(), as the reference logs it; "
                      "the final stats-collecting epoch is run but not timed",
               "steps": n}

    decode = None
    if rank == 0 and world == 1:
        del model, eng
        torch.cuda.empty_cache()
        try:
            decode = decode_bench(torch, nnm, device)
        except Exception as exc:  # the train metric is the headline; report, do not lose the line
            decode = {"error": repr(exc)}
    if rank == 0:
        cpu_rate, cores, sample, _ = cpu_reference_step_rate(1, 1) if world == 1 else (None, None, None, None)
        ft = f_tok(cfg["L"], cfg["C"], V, T)
        peaks, which = measured_peaks()
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": {"workload": "GPT-2 124M (12L/768d/12h, T=1024, V=50257 untied head, 163,037,184 params) "
                                       "bf16-autocast DDP training step, AdamW lr 6e-4 wd 0.01, num_steps=1",
                           "per_gpu_batch": B, "seq_len": T, "global_batch_tokens": world * B * T,
                           "parallelism": f"dp{world}",
                           "l2_policy": "inputs larger than L2: each step streams >25 GB of activations and 3.3 GB of "
                                        "optimizer state through the 126 MB L2; token batches rotate every step"},
                "mfu_vs_2.25PF": value * ft / (world * 2.25e15),
                "mfu_vs_measured_sustained": value * ft / (world * peaks["bf16_tflops_sustained"] * 1e12),
                "flops_per_token": ft, "final_loss": final_loss,
                "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
                "decode": decode}
        if cpu_rate is not None:
            line["cpu_baseline"] = {"value": cpu_rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def _reserve_stdout():
    """The contract is ONE JSON line on stdout. Libraries (NCCL's version banner, torch-elastic notices) also write to
    file descriptor 1, so keep a private duplicate of it for the result line and point fd 1 at stderr."""
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
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
