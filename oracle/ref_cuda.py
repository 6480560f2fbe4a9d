"""Run the UNMODIFIED reference (staged in oracle/_ref by oracle/stage_ref.py) on the B200 box through its own public
API and stock code path: the bf16-autocast torch-CUDA training step (SURVEY 8c tier 2, BASELINE.md section 3 "B-GPU")
and its KV-cache generation. TEST INFRASTRUCTURE: the product never imports this; it produces
  * the bf16 golden the `-m gpu` parity tests compare the fused path with (tests/golden/golden_bf16_124m.npz is made from
    this script's output by tests/golden/make_golden_bf16.py), and
  * the performance bar "what the reference itself does on this hardware" (profiles/r02_ref_torch_cuda_*.json).

    python oracle/ref_cuda.py train   --config 124m --epochs 60 --out gpurun_out/ref_train_n1.json
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
           oracle/ref_cuda.py train --config 124m --epochs 30 --out gpurun_out/ref_train_n8.json
    python oracle/ref_cuda.py profile --config 124m --out gpurun_out/ref_profile.json
    python oracle/ref_cuda.py decode  --out gpurun_out/ref_decode.json

The only things done to the reference from outside (harness guards, as BASELINE.md section 3 plans):
  * NeuralNetworkModel._record_training_overall_progress is replaced by a no-op for the training runs: on CUDA under
    bf16 autocast it crashes in torch.histogram (neural_net_model.py:737-738) AFTER the last timed epoch,
    before the final serialize() that stores the progress list this harness reads. Per-epoch losses and durations are
    recorded before that point and are unaffected.
  * SHM_PATH and the working directory point into a scratch folder.
Everything else — Mapper, Loader, DistributedDataParallel wrap, autocast, AdamW, the epoch clock — is the reference's.
"""
import argparse
import copy
import json
import math
import os
import sys
import tempfile
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")

CONFIGS = {
    "tiny": dict(L=2, H=4, C=128, T=64, V=65, B=16),
    "124m": dict(L=12, H=12, C=768, T=1024, V=50257, B=32),
    "350m": dict(L=24, H=16, C=1024, T=1024, V=50257, B=32),
    "xl": dict(L=48, H=25, C=1600, T=1024, V=50257, B=8),
}
OPTIM = {"adamw": {"lr": 6e-4, "betas": [0.9, 0.95], "eps": 1e-8}}


def gpt_dsl(L, H, C, T, V):
    """The /model/ request body of main.py:57-83: normal(0.02) weights, 0.02/sqrt(2L) on the two residual projections,
    zero biases, exact GELU, dropout 0, untied default-initialised LM head."""
    pstd = 0.02 / math.sqrt(2 * L)

    def lin(i, o, std):
        return {"linear": {"in_features": i, "out_features": o}, "normal": {"mean": 0.0, "std": std}, "zeros": {}}

    layers = [{"summation": [
        {"embedding": {"num_embeddings": V, "embedding_dim": C}, "normal": {"mean": 0.0, "std": 0.02}},
        {"position": {"num_embeddings": T, "embedding_dim": C}, "normal": {"mean": 0.0, "std": 0.02}}]},
        {"dropout": {"p": 0.0}}]
    for _ in range(L):
        layers.append({"residual": [
            {"sequential": [{"layernorm": {"normalized_shape": C}}, lin(C, 3 * C, 0.02),
                            {"attention": {"num_heads": H, "dropout": 0.0}}, lin(C, C, pstd), {"dropout": {"p": 0.0}}]},
            {"sequential": [{"layernorm": {"normalized_shape": C}}, lin(C, 4 * C, 0.02), {"gelu": {}},
                            lin(4 * C, C, pstd), {"dropout": {"p": 0.0}}]}]})
    layers += [{"layernorm": {"normalized_shape": C}},
               {"linear": {"in_features": C, "out_features": V, "bias": False}},
               {"softmaxlast": {"dim": -1}}]
    return layers


def write_shards(dataset_id, vocab, tokens, folder="data"):
    """uint16 .npy shard in Downloader._save's format (loaders.py:22-25), tokens from default_rng(1234)."""
    import numpy as np
    os.makedirs(folder, exist_ok=True)
    arr = np.random.default_rng(1234).integers(0, vocab, size=tokens).astype(np.uint16)
    np.save(os.path.join(folder, f"{dataset_id}_000000"), arr)


def _env_int(name, default):
    return int(os.environ.get(name, default))


def _import_reference(work):
    """chdir into the scratch folder FIRST (loaders.py creates ./data at import), then import the staged modules."""
    if not os.path.exists(os.path.join(REF, "neural_net_model.py")):
        raise SystemExit("oracle/_ref is empty: run `python oracle/stage_ref.py` where /root/reference exists")
    os.chdir(work)
    if not os.path.exists("log_config.json"):
        import shutil
        shutil.copyfile(os.path.join(REF, "log_config.json"), "log_config.json")
    sys.path.insert(0, REF)
    import neural_net_model as ref_nnm
    from mappers import Mapper
    ref_nnm.NeuralNetworkModel.SHM_PATH = os.path.join(work, "shm")
    os.makedirs(os.path.join(work, "shm"), exist_ok=True)
    return ref_nnm, Mapper


def _shared_work_dir(tag):
    """One scratch folder shared by all ranks of a torchrun launch (rank 0's model file is what the others load)."""
    port = os.environ.get("MASTER_PORT", "0")
    work = os.path.join(tempfile.gettempdir(), f"penroz_ref_{tag}_{port}_{os.environ.get('TORCHELASTIC_RUN_ID', 'solo')}")
    os.makedirs(work, exist_ok=True)
    return work


def cmd_train(a):
    import torch
    cfg = CONFIGS[a.config]
    rank, world = _env_int("RANK", 0), _env_int("WORLD_SIZE", 1)
    work = _shared_work_dir("train")
    ref_nnm, Mapper = _import_reference(work)
    if a.threads:
        torch.set_num_threads(a.threads)
    B = a.batch or cfg["B"]
    T, V = cfg["T"], cfg["V"]
    step_size = a.step_size or max(1, B // world)
    num_steps = max(1, B // (step_size * world))
    dataset = f"synth{V}"
    ready = os.path.join(work, "READY")
    if rank == 0:
        need = (a.epochs * num_steps * world + 2) * B * T + 2
        write_shards(dataset, V, max(need, 2 ** 20))
        torch.manual_seed(0)
        model = ref_nnm.NeuralNetworkModel("ref", Mapper(copy.deepcopy(gpt_dsl(cfg["L"], cfg["H"], cfg["C"], T, V)),
                                                         copy.deepcopy(OPTIM)))
        n_params = model.num_params
        model.serialize()
        del model
        open(ready, "w").write(str(n_params))
    else:
        while not os.path.exists(ready):
            time.sleep(0.2)
    n_params = int(open(ready).read())
    # harness guard (see module docstring): the end-of-training diagnostics crash on bf16 activations
    ref_nnm.NeuralNetworkModel._record_training_overall_progress = lambda self, activations: None
    t0 = time.time()
    ref_nnm.NeuralNetworkModel.train_model_on_device("ref", a.device, dataset, 0, a.epochs, B, T, step_size)
    wall = time.time() - t0
    if rank != 0:
        return
    model = ref_nnm.NeuralNetworkModel.deserialize("ref")
    prog = model.progress
    losses = [p["cost"] for p in prog]
    secs = [p["durationInSecs"] for p in prog]
    steady = secs[a.skip:] if len(secs) > a.skip + 2 else secs
    steady_mean = sum(steady) / len(steady)
    steady_sorted = sorted(steady)
    out = {
        "what": "unmodified reference, NeuralNetworkModel.train_model_on_device, its own per-epoch clock "
                "(time.time() around the epoch after torch.cuda.synchronize(), neural_net_model.py:615,678-685)",
        "config": a.config, "cfg": cfg, "device": a.device, "world": world, "batch_size": B, "block_size": T,
        "step_size": step_size, "num_steps": num_steps, "epochs": a.epochs, "num_params": n_params,
        "torch": torch.__version__, "threads": torch.get_num_threads(), "cpu_count": os.cpu_count(),
        "gpu": torch.cuda.get_device_name(0) if torch.cuda.is_available() else None,
        "losses": losses, "epoch_secs": secs, "wall_secs_total": wall,
        "steady_epochs": len(steady), "steady_epoch_secs_mean": steady_mean,
        "steady_epoch_secs_median": steady_sorted[len(steady_sorted) // 2],
        # true tokens/s: every rank runs num_steps micro-batches of B*T tokens per epoch
        "tokens_per_sec": world * num_steps * B * T / steady_mean,
        "tokens_per_sec_median": world * num_steps * B * T / steady_sorted[len(steady_sorted) // 2],
        "guards": ["_record_training_overall_progress -> no-op (bf16 torch.histogram crash, SURVEY F5a)"],
    }
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(out, f)
    print(json.dumps({k: out[k] for k in ("config", "world", "tokens_per_sec", "steady_epoch_secs_mean")}
                     | {"loss_first": losses[0], "loss_last": losses[-1]}))


def cmd_profile(a):
    """torch.profiler over a few of the reference's own training micro-steps: which GEMM / SDPA / optimizer kernels
    torch picks on sm_100 and their share of the step (the 'kernel to beat' list)."""
    import torch
    from torch.profiler import ProfilerActivity, profile
    cfg = CONFIGS[a.config]
    work = tempfile.mkdtemp(prefix="penroz_ref_prof_")
    ref_nnm, Mapper = _import_reference(work)
    B, T, V = a.batch or cfg["B"], cfg["T"], cfg["V"]
    torch.manual_seed(0)
    model = ref_nnm.NeuralNetworkModel("refp", Mapper(copy.deepcopy(gpt_dsl(cfg["L"], cfg["H"], cfg["C"], T, V)),
                                                      copy.deepcopy(OPTIM)))
    model.to("cuda")
    model.train()
    gen = torch.Generator().manual_seed(1)
    toks = torch.randint(0, V, (B * T + 1,), generator=gen)
    x, y = toks[:-1].view(B, T).cuda(), toks[1:].view(B, T).cuda()

    def step():     # the body of the reference's epoch (neural_net_model.py:622-677) with num_steps = 1
        model.optimizer.zero_grad()
        with torch.amp.autocast("cuda", dtype=torch.bfloat16):
            _, cost = model(x, y, skip_softmax=True)
        cost.backward()
        model.optimizer.step()

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        step()
    e1.record()
    torch.cuda.synchronize()
    bare_ms = e0.elapsed_time(e1) / 5
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(2):
            step()
        torch.cuda.synchronize()
    rows = []
    for ev in prof.key_averages():
        t = getattr(ev, "device_time_total", None)
        if t is None:
            t = getattr(ev, "cuda_time_total", 0.0)
        if t and t > 0:
            rows.append({"kernel": ev.key[:160], "calls_per_step": ev.count / 2, "ms_per_step": t / 2 / 1e3})
    rows.sort(key=lambda r: -r["ms_per_step"])
    total = sum(r["ms_per_step"] for r in rows)
    out = {"what": "torch.profiler (CUDA activities) over 2 optimizer steps of the unmodified reference model: "
                   "zero_grad, autocast(bf16) forward + cross entropy, backward, torch AdamW; no rank-0 diagnostics",
           "config": a.config, "batch_size": B, "block_size": T, "torch": torch.__version__,
           "gpu": torch.cuda.get_device_name(0), "bare_step_ms_cuda_events": bare_ms,
           "tokens_per_sec_bare_step": B * T / (bare_ms / 1e3),
           "kernel_ms_per_step_total": total, "kernels": rows[:60]}
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps({"bare_step_ms": bare_ms, "tok_s": out["tokens_per_sec_bare_step"],
                      "top": [(r["kernel"][:60], round(r["ms_per_step"], 2)) for r in rows[:8]]}))


def cmd_decode(a):
    """The reference's own generate_tokens on CUDA for a bf16 (HF-import style) and an fp32 124M model: greedy ids, the
    top-2 logit gap at every generated position (from one teacher-forced forward of the reference over the generated
    sequence; near-ties are where two bf16 implementations may legitimately pick different ids), seeded top-k ids, and
    tokens/s at batch 1 and 64."""
    import numpy as np
    import torch
    cfg = CONFIGS[a.config]
    work = tempfile.mkdtemp(prefix="penroz_ref_dec_")
    ref_nnm, Mapper = _import_reference(work)
    L, H, C, T, V = cfg["L"], cfg["H"], cfg["C"], cfg["T"], cfg["V"]
    out = {"what": "unmodified reference generate_tokens on CUDA (the reference service itself decodes on the CPU, "
                   "SURVEY F5b; this is its same-box GPU twin)",
           "config": a.config, "torch": torch.__version__, "gpu": torch.cuda.get_device_name(0), "cases": {}}
    arrays = {}
    prompt_rng = np.random.default_rng(7)
    for tag, dtype in (("bf16", torch.bfloat16), ("f32", torch.float32)):
        torch.manual_seed(0)
        model = ref_nnm.NeuralNetworkModel(f"d{tag}", Mapper(copy.deepcopy(gpt_dsl(L, H, C, T, V)), copy.deepcopy(OPTIM)))
        if dtype != torch.float32:
            model.to(dtype=dtype)
        model.to("cuda")
        prompt = [[int(t) for t in np.random.default_rng(11).integers(0, V, size=a.prompt)]]
        ids = model.generate_tokens(prompt, T, a.new, temperature=0.0)
        arrays[f"{tag}/prompt"] = np.array(prompt[0], dtype=np.int64)
        arrays[f"{tag}/greedy_ids"] = np.array(ids, dtype=np.int64)
        with torch.no_grad():
            seq = torch.tensor([ids[:-1]], dtype=torch.long, device="cuda")
            acts, _ = model(seq, None, skip_softmax=True)
            logits = acts[-1][0].float()                       # [S, V]
            top2 = logits.topk(2, dim=-1)
        pos = slice(a.prompt - 1, a.prompt - 1 + a.new)
        arrays[f"{tag}/top2_gap"] = (top2.values[pos, 0] - top2.values[pos, 1]).cpu().numpy()
        arrays[f"{tag}/top1_value"] = top2.values[pos, 0].cpu().numpy()
        arrays[f"{tag}/teacher_forced_argmax"] = top2.indices[pos, 0].cpu().numpy()
        torch.manual_seed(42)
        arrays[f"{tag}/topk50_seed42_ids"] = np.array(
            model.generate_tokens(prompt, T, min(a.new, 32), temperature=1.0, top_k=50), dtype=np.int64)
        if a.time:
            for Bn in (1, 64):
                pr = [[int(t) for t in prompt_rng.integers(0, V, size=512)] for _ in range(Bn)]
                model.generate_tokens(pr, T, 2, temperature=0.0)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                model.generate_tokens(pr, T, 64, temperature=0.0)
                torch.cuda.synchronize()
                dt = time.perf_counter() - t0
                out["cases"][f"{tag}_b{Bn}"] = {"tokens_per_s": Bn * 64 / dt, "seconds": dt,
                                                "workload": "512-token prompt + 64 new tokens, greedy, prefill included"}
        del model
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    np.savez_compressed(os.path.splitext(a.out)[0] + ".npz", **arrays)
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out["cases"]))


def main():
    ap = argparse.ArgumentParser()
    sub = ap.add_subparsers(dest="cmd", required=True)
    t = sub.add_parser("train")
    t.add_argument("--config", default="124m", choices=list(CONFIGS))
    t.add_argument("--device", default="cuda")
    t.add_argument("--epochs", type=int, default=60)
    t.add_argument("--batch", type=int, default=0)
    t.add_argument("--step-size", type=int, default=0)
    t.add_argument("--skip", type=int, default=10, help="epochs left out of the steady-state mean (warm-up)")
    t.add_argument("--threads", type=int, default=0)
    t.add_argument("--out", required=True)
    p = sub.add_parser("profile")
    p.add_argument("--config", default="124m", choices=list(CONFIGS))
    p.add_argument("--batch", type=int, default=0)
    p.add_argument("--out", required=True)
    d = sub.add_parser("decode")
    d.add_argument("--config", default="124m", choices=list(CONFIGS))
    d.add_argument("--prompt", type=int, default=16)
    d.add_argument("--new", type=int, default=48)
    d.add_argument("--time", type=int, default=1)
    d.add_argument("--out", required=True)
    a = ap.parse_args()
    a.out = os.path.abspath(a.out)      # the commands chdir into a scratch folder before they write
    {"train": cmd_train, "profile": cmd_profile, "decode": cmd_decode}[a.cmd](a)


if __name__ == "__main__":
    main()
