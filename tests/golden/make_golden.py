"""Generate golden vectors by running the UNMODIFIED reference on CPU (fp32).

Run in the build container only (needs /root/reference, which is absent on the GPU box):

    python tests/golden/make_golden.py            # writes tests/golden/*.npz

What is pinned (reference call sites in parentheses):
  * tiny GPT (L=2,H=4,C=128,T=64,V=65) built through the reference Mapper
    (mappers.py:72-99) from torch.manual_seed(0): initial state dict, 50 train
    steps of NeuralNetworkModel.train_model (neural_net_model.py:552-722) on a
    synthetic uint16 shard -> per-step loss, final state dict, final logits,
    greedy and seeded top-k generations (neural_net_model.py:457-479).
  * Loader index arithmetic (loaders.py:44-87) for several (world, rank) and a
    shard roll-over.
  * GPT-2-small-shaped slice (L=2 of the 124M layout, V=50257 untied head) loss.

The script never copies reference source; it imports it from /root/reference.
"""
import json
import os
import sys
import tempfile
import hashlib

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def gpt_dsl(L, H, C, T, V, gelu=None, ln_bias=True):
    """The /model/ DSL of main.py:57-83 (normal 0.02, proj 0.02/sqrt(2L), zeros bias, exact GELU)."""
    import math
    gelu = gelu if gelu is not None else {"gelu": {}}
    ln = {"normalized_shape": C} if ln_bias else {"normalized_shape": C, "bias": False}
    pstd = 0.02 / math.sqrt(2 * L)
    layers = [{"summation": [
        {"embedding": {"num_embeddings": V, "embedding_dim": C}, "normal": {"mean": 0.0, "std": 0.02}},
        {"position": {"num_embeddings": T, "embedding_dim": C}, "normal": {"mean": 0.0, "std": 0.02}}]},
        {"dropout": {"p": 0.0}}]
    for _ in range(L):
        layers.append({"residual": [
            {"sequential": [{"layernorm": dict(ln)},
                            {"linear": {"in_features": C, "out_features": 3 * C},
                             "normal": {"mean": 0.0, "std": 0.02}, "zeros": {}},
                            {"attention": {"num_heads": H, "dropout": 0.0}},
                            {"linear": {"in_features": C, "out_features": C},
                             "normal": {"mean": 0.0, "std": pstd}, "zeros": {}},
                            {"dropout": {"p": 0.0}}]},
            {"sequential": [{"layernorm": dict(ln)},
                            {"linear": {"in_features": C, "out_features": 4 * C},
                             "normal": {"mean": 0.0, "std": 0.02}, "zeros": {}},
                            dict(gelu),
                            {"linear": {"in_features": 4 * C, "out_features": C},
                             "normal": {"mean": 0.0, "std": pstd}, "zeros": {}},
                            {"dropout": {"p": 0.0}}]}]})
    layers += [{"layernorm": dict(ln)},
               {"linear": {"in_features": C, "out_features": V, "bias": False}},
               {"softmaxlast": {"dim": -1}}]
    return layers


OPTIM = {"adamw": {"lr": 6e-4, "betas": [0.9, 0.95], "eps": 1e-8}}


def synth_shard(vocab, k, size):
    return np.random.default_rng(1234 + k).integers(0, vocab, size=size).astype(np.uint16)


def sd_to_np(sd):
    return {k: v.detach().cpu().float().numpy().copy() for k, v in sd.items()}


def main():
    work = tempfile.mkdtemp(prefix="penroz_golden_")
    os.chdir(work)
    sys.path.insert(0, REF)
    import copy
    import torch
    torch.set_num_threads(8)
    import neural_net_model as ref_nnm
    import loaders as ref_loaders
    from mappers import Mapper

    ref_nnm.NeuralNetworkModel.SHM_PATH = os.path.join(work, "shm")
    os.makedirs(ref_nnm.NeuralNetworkModel.SHM_PATH, exist_ok=True)

    out = {}
    meta = {"torch": torch.__version__, "numpy": np.__version__}

    # ---------------- loader indices (loaders.py:44-87) ----------------
    os.makedirs("data", exist_ok=True)
    shard_tokens = 5000
    for k in range(3):
        np.save(os.path.join("data", f"ldr_{k:06d}"), synth_shard(997, k, shard_tokens + 37 * k))
    loader_cases = []
    for world, B, T in [(1, 4, 16), (2, 4, 16), (4, 3, 32), (2, 16, 64)]:
        for rank in range(world):
            buf = B * T
            ld = ref_loaders.Loader("ldr", 0, buf * rank, buf, buf * world)
            rec = []
            for _ in range(90 // world + 5):
                x, y = ld.next_batch()
                rec.append([int(x[0]), int(x[-1]), int(y[0]), int(y[-1]), int(x.astype(np.int64).sum()),
                            int(y.astype(np.int64).sum()), int(len(x)), int(ld.shard_idx), int(ld.token_idx)])
            loader_cases.append({"world": world, "rank": rank, "B": B, "T": T, "rec": rec})
    # target_offset=0 variant used by evaluate_model (neural_net_model.py:342-343)
    ld = ref_loaders.Loader("ldr", 1, 0, 64, 64)
    rec0 = []
    for _ in range(100):
        x = ld.next_batch(target_offset=0)[0]
        rec0.append([int(x[0]), int(x[-1]), int(x.astype(np.int64).sum()), int(len(x))])
    meta["loader"] = {"shard_tokens": shard_tokens, "vocab": 997, "cases": loader_cases, "offset0": rec0}

    # ---------------- tiny GPT training (config 1) ----------------
    L, H, C, T, V = 2, 4, 128, 64, 65
    B, EPOCHS = 16, 50
    np.save(os.path.join("data", "synth65_000000"), synth_shard(V, 0, 2 ** 17))
    torch.manual_seed(0)
    mapper = Mapper(copy.deepcopy(gpt_dsl(L, H, C, T, V)), copy.deepcopy(OPTIM))
    model = ref_nnm.NeuralNetworkModel("tiny", mapper)
    assert model.num_params == 421632, model.num_params
    for k, v in sd_to_np(model.state_dict()).items():
        out["tiny/init/" + k] = v
    # fixed probe batch
    probe = synth_shard(V, 7, B * T + 1).astype(np.int64)
    px = torch.tensor(probe[:-1]).view(B, T)
    py = torch.tensor(probe[1:]).view(B, T)
    with torch.no_grad():
        acts, cost = model(px, py, skip_softmax=True)
    out["tiny/probe_x"] = px.numpy()
    out["tiny/probe_y"] = py.numpy()
    out["tiny/init_logits"] = acts[-1].numpy().copy()
    out["tiny/init_loss"] = np.array(cost.item(), dtype=np.float64)
    # one explicit fwd+bwd for gradient parity
    model.optimizer.zero_grad()
    _, cost = model(px, py, skip_softmax=True)
    cost.backward()
    for k, p in model.named_parameters():
        out["tiny/init_grad/" + k] = p.grad.detach().numpy().copy()
    model.optimizer.zero_grad()
    # generation before training (weights == init)
    prompt = [[int(t) for t in probe[:5]]]
    out["tiny/gen_greedy_b1"] = np.array(model.generate_tokens(prompt, T, 80, temperature=0.0))
    prompt3 = [[int(t) for t in probe[i * 7:i * 7 + 6]] for i in range(3)]
    out["tiny/gen_greedy_b3_row0"] = np.array(model.generate_tokens(prompt3, T, 20, temperature=0.0))
    torch.manual_seed(42)
    out["tiny/gen_topk5_seed42"] = np.array(model.generate_tokens(prompt, T, 30, temperature=1.0, top_k=5))
    torch.manual_seed(42)
    out["tiny/gen_sample_seed42"] = np.array(model.generate_tokens(prompt, T, 30, temperature=0.8))
    out["tiny/gen_prompt"] = np.array(prompt)
    out["tiny/gen_prompt3"] = np.array(prompt3)
    # compute_output (/output/) on init weights: probs of last position + cost
    probs, c = model.compute_output(px[:2].tolist(), py[:2].tolist())
    out["tiny/output_probs"] = np.array(probs, dtype=np.float32)
    out["tiny/output_cost"] = np.array(c, dtype=np.float64)

    # train 50 epochs (num_steps = 1), world 1
    model.train_model("synth65", 0, EPOCHS, B, T, B)
    out["tiny/train_loss"] = np.array([p["cost"] for p in model.progress], dtype=np.float64)
    assert len(model.progress) == EPOCHS
    for k, v in sd_to_np(model.state_dict()).items():
        out["tiny/final/" + k] = v
    with torch.no_grad():
        acts, cost = model(px, py, skip_softmax=True)
    out["tiny/final_logits"] = acts[-1].numpy().copy()
    out["tiny/final_loss"] = np.array(cost.item(), dtype=np.float64)
    out["tiny/gen_greedy_trained"] = np.array(model.generate_tokens(prompt, T, 80, temperature=0.0))
    st = model.optimizer.state_dict()
    out["tiny/final_adam_step"] = np.array(float(st["state"][0]["step"]))
    out["tiny/final_exp_avg0"] = st["state"][0]["exp_avg"].numpy().copy()
    out["tiny/final_exp_avg_sq0"] = st["state"][0]["exp_avg_sq"].numpy().copy()

    # grad accumulation variant: num_steps = 4 (step_size 4), 6 epochs, world 1
    torch.manual_seed(0)
    model2 = ref_nnm.NeuralNetworkModel("tiny2", Mapper(copy.deepcopy(gpt_dsl(L, H, C, T, V)), copy.deepcopy(OPTIM)))
    model2.train_model("synth65", 0, 6, B, T, 4)
    out["tiny/accum4_loss"] = np.array([p["cost"] for p in model2.progress], dtype=np.float64)

    # tanh-GELU + LayerNorm(bias=False) variant (HF gelu_new path, mappers.py:138-139; test_neural_net_model.py:58)
    torch.manual_seed(1)
    model3 = ref_nnm.NeuralNetworkModel("tiny3", Mapper(copy.deepcopy(
        gpt_dsl(L, H, C, T, V, gelu={"gelu": {"approximate": "tanh"}}, ln_bias=False)), copy.deepcopy(OPTIM)))
    for k, v in sd_to_np(model3.state_dict()).items():
        out["tanh/init/" + k] = v
    model3.train_model("synth65", 0, 5, B, T, B)
    out["tanh/train_loss"] = np.array([p["cost"] for p in model3.progress], dtype=np.float64)

    # ---------------- GPT-2 124M-shaped slice (V=50257, C=768, H=12, L=2, T=128) ----------------
    L2, H2, C2, T2, V2 = 2, 12, 768, 128, 50257
    torch.manual_seed(0)
    big = ref_nnm.NeuralNetworkModel("slice", Mapper(copy.deepcopy(gpt_dsl(L2, H2, C2, T2, V2)), copy.deepcopy(OPTIM)))
    bprobe = synth_shard(V2, 3, 4 * T2 + 1).astype(np.int64)
    bx = torch.tensor(bprobe[:-1]).view(4, T2)
    by = torch.tensor(bprobe[1:]).view(4, T2)
    losses = []
    for _ in range(3):
        big.optimizer.zero_grad()
        _, cost = big(bx, by, skip_softmax=True)
        cost.backward()
        big.optimizer.step()
        losses.append(cost.item())
    out["slice/x"] = bx.numpy()
    out["slice/y"] = by.numpy()
    out["slice/loss3"] = np.array(losses, dtype=np.float64)
    sdb = big.state_dict()
    meta["slice"] = {"L": L2, "H": H2, "C": C2, "T": T2, "V": V2, "seed": 0,
                     "final_param_sum": {k: float(v.double().sum()) for k, v in sdb.items()},
                     "num_params": big.num_params}
    with torch.no_grad():
        acts, _ = big(bx[:1], None, skip_softmax=True)
    out["slice/final_logits_row"] = acts[-1][0, -1].numpy().copy()

    meta["tiny"] = {"L": L, "H": H, "C": C, "T": T, "V": V, "B": B, "epochs": EPOCHS, "seed": 0,
                    "optim": OPTIM, "shard": "default_rng(1234).integers(0,65,2**17) uint16"}
    np.savez_compressed(os.path.join(HERE, "golden.npz"), **out)
    with open(os.path.join(HERE, "golden_meta.json"), "w") as f:
        json.dump(meta, f)
    h = hashlib.sha256(open(os.path.join(HERE, "golden.npz"), "rb").read()).hexdigest()
    print("wrote golden.npz", h, "keys", len(out))


if __name__ == "__main__":
    main()
