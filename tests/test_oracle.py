"""Pins the numpy oracle (oracle/gpt2_oracle.py) to golden vectors recorded from the unmodified reference
(tests/golden/make_golden.py). CPU only."""
import numpy as np
import pytest

from oracle.gpt2_oracle import Gpt2Oracle, LoaderOracle


def _state(g, prefix):
    return {k[len(prefix):]: g[k] for k in g.files if k.startswith(prefix)}


def test_forward_loss_and_logits(golden):
    g, _ = golden
    o = Gpt2Oracle(_state(g, "tiny/init/"), 2, 4)
    logits, loss = o.forward(g["tiny/probe_x"], g["tiny/probe_y"])
    np.testing.assert_allclose(loss, float(g["tiny/init_loss"]), rtol=2e-6)
    np.testing.assert_allclose(logits, g["tiny/init_logits"], rtol=1e-4, atol=2e-6)


def test_every_parameter_gradient(golden):
    g, _ = golden
    o = Gpt2Oracle(_state(g, "tiny/init/"), 2, 4)
    o.forward(g["tiny/probe_x"], g["tiny/probe_y"], keep=True)
    grads = o.backward()
    ref = _state(g, "tiny/init_grad/layers")
    assert len(ref) == len(grads)
    for k, v in ref.items():
        got = grads["layers" + k]
        assert np.abs(got - v).max() <= 2e-5 * np.abs(v).max() + 1e-9, k


def test_fifty_training_steps_and_final_weights(golden):
    g, meta = golden
    shard = np.random.default_rng(1234).integers(0, 65, size=2 ** 17).astype(np.uint16)
    o = Gpt2Oracle(_state(g, "tiny/init/"), 2, 4)
    ld = LoaderOracle([shard], 0, 0, 16 * 64, 16 * 64)
    losses = [o.train_epoch(ld, 1, 16, 64) for _ in range(50)]
    np.testing.assert_allclose(losses, g["tiny/train_loss"], rtol=2e-5)
    for k, v in _state(g, "tiny/final/").items():
        # AdamW moves each weight by ~lr per step whatever the gradient's size, so elements whose gradient is ~0
        # amplify fp32-vs-fp64 rounding: 1e-4 absolute is 0.3 % of the 50 * lr = 0.03 a weight can travel
        assert np.abs(o.p[k] - v).max() < 1e-4, k
    np.testing.assert_allclose(o.adam["m"]["layers.0.0.weight"], g["tiny/final_exp_avg0"], rtol=1e-3, atol=1e-8)
    assert o.generate_greedy(g["tiny/gen_prompt"], 64, 80) == g["tiny/gen_greedy_trained"].tolist()


def test_grad_accumulation_and_tanh_gelu(golden):
    g, _ = golden
    shard = np.random.default_rng(1234).integers(0, 65, size=2 ** 17).astype(np.uint16)
    o = Gpt2Oracle(_state(g, "tiny/init/"), 2, 4)
    ld = LoaderOracle([shard], 0, 0, 16 * 64, 16 * 64)
    losses = [o.train_epoch(ld, 4, 16, 64) for _ in range(6)]
    np.testing.assert_allclose(losses, g["tiny/accum4_loss"], rtol=2e-5)


def test_greedy_decode_with_cache_and_reprefill(golden):
    g, _ = golden
    o = Gpt2Oracle(_state(g, "tiny/init/"), 2, 4)
    assert o.generate_greedy(g["tiny/gen_prompt"], 64, 80) == g["tiny/gen_greedy_b1"].tolist()
    assert o.generate_greedy(g["tiny/gen_prompt3"], 64, 20) == g["tiny/gen_greedy_b3_row0"].tolist()


def test_loader_windows(golden):
    g, meta = golden
    ld = meta["loader"]
    shards = [np.random.default_rng(1234 + k).integers(0, ld["vocab"], size=ld["shard_tokens"] + 37 * k).astype(np.uint16)
              for k in range(3)]
    for case in ld["cases"]:
        buf = case["B"] * case["T"]
        lo = LoaderOracle(shards, 0, buf * case["rank"], buf, buf * case["world"])
        for rec in case["rec"]:
            x, y = lo.next_batch()
            got = [int(x[0]), int(x[-1]), int(y[0]), int(y[-1]), int(x.astype(np.int64).sum()),
                   int(y.astype(np.int64).sum()), len(x), lo.shard_idx, lo.token_idx]
            assert got == rec, (case["world"], case["rank"])
    lo = LoaderOracle(shards, 1, 0, 64, 64)
    for rec in ld["offset0"]:
        x, y = lo.next_batch(target_offset=0)
        assert y is None and [int(x[0]), int(x[-1]), int(x.astype(np.int64).sum()), len(x)] == rec


@pytest.mark.parametrize("tag,num_steps,epochs", [("ns1", 1, 10), ("ns2", 2, 6)])
def test_two_rank_ddp_epochs(golden, tag, num_steps, epochs):
    """The oracle's in-process DDP emulation against the reference run under 2-rank gloo DDP
    (tests/golden/make_golden_ddp.py): plain DDP, and accumulation where only micro-step 0 exchanges gradients
    and the two ranks' weights drift apart."""
    import os
    g, _ = golden
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_ddp.npz"))
    shard = np.random.default_rng(1234).integers(0, 65, size=2 ** 17).astype(np.uint16)
    micro = 16          # every micro-step reads a full batch_size x block_size buffer (neural_net_model.py:629-631);
    buf = micro * 64    # step_size only sets num_steps (:581)
    reps = [Gpt2Oracle(_state(g, "tiny/init/"), 2, 4) for _ in range(2)]
    lds = [LoaderOracle([shard], 0, buf * r, buf, buf * 2) for r in range(2)]
    costs = [Gpt2Oracle.ddp_train_epoch(reps, lds, num_steps, micro, 64) for _ in range(epochs)]
    np.testing.assert_allclose(costs, d[f"{tag}/r0/cost"], rtol=2e-5)
    for r in range(2):
        pre = f"{tag}/r{r}/final/"
        for k in [k for k in d.files if k.startswith(pre)]:
            assert np.abs(reps[r].p[k[len(pre):]] - d[k]).max() < 1e-4, k
    w0, w1 = d[f"{tag}/r0/final/layers.5.weight"], d[f"{tag}/r1/final/layers.5.weight"]
    assert np.array_equal(w0, w1) == (num_steps == 1)
    same = np.array_equal(reps[0].p["layers.5.weight"], reps[1].p["layers.5.weight"])
    assert same == (num_steps == 1)
