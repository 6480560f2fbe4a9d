"""CPU tier: the C-ABI library loads and exports every declared symbol; host logic (layer DSL RNG order, shard
windows, GPT-2 tree matcher, flat parameter state, bucketed gradient sync over Gloo with world_size 2, KV cache API)
behaves as pinned by the reference's golden vectors; the reference's own tests pass against the drop-in modules."""
import copy
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from helpers import build_gpt, scratch_cwd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "penr-oz-neural-network-v3-torch-ddp_b200")


def test_c_abi_library_loads_and_exports_every_declared_symbol():
    from penroz import _lib
    protos = _lib.parse_header()
    assert len(protos) >= 30
    lib = _lib.lib()                        # dlopen works without a GPU
    for name in protos:
        assert hasattr(lib._cdll, name), f"{name} declared in include/penroz_b200.h but not exported"
    assert lib.abi_version() == 4 and lib.kv_page_size() == 64
    assert lib.layernorm_bwd_scratch_bytes(32768, 768) > 0
    if not torch.cuda.is_available():
        with pytest.raises(_lib.PenrozError):
            _lib.require_device()           # the product path fails loudly without a device: no CPU fallback


def test_argument_validation_without_a_gpu():
    from penroz import _lib
    lib = _lib.lib()
    with pytest.raises(_lib.PenrozError, match="empty problem"):
        lib.gemm_bf16(None, 8, 0, None, 8, 0, 0, 8, 8, 0, None, None, 8, None, 0, None, 0, 0, None)
    with pytest.raises(_lib.PenrozError, match="head_dim"):
        lib.attn_fwd_tc(None, None, None, 1, 8, 1, 48, None)
    with pytest.raises(_lib.PenrozError, match="8192"):
        lib.colsum_static(1, 9000, 4, 9000, 1, 0, 0, None)


def test_dsl_builds_the_reference_weights_from_the_same_seed(golden):
    g, _ = golden
    m = build_gpt("t", 2, 4, 128, 64, 65)
    assert m.num_params == 421632 and m._plan is not None
    for k, v in m.state_dict().items():
        assert np.array_equal(v.numpy(), g["tiny/init/" + k]), k
    assert m._plan.n_layer == 2 and m._plan.n_head == 4 and m._plan.vocab == 65 and m._plan.block_size == 64


def test_plan_matcher_rejects_other_trees():
    import neural_net_model as nnm
    from penroz import dsl
    from penroz.plan import match_gpt2
    mlp = nnm.NeuralNetworkModel("mlp", dsl.Mapper([{"linear": {"in_features": 4, "out_features": 8}}, {"tanh": {}},
                                                    {"linear": {"in_features": 8, "out_features": 2}}, {"softmax": {"dim": -1}}],
                                                   {"sgd": {"lr": 0.1}}))
    assert mlp._plan is None
    spec = dsl.gpt2_layers(1, 2, 32, 16, 11)
    spec[2]["residual"][0]["sequential"][2] = {"attention": {"num_heads": 2, "num_kv_heads": 1, "rope_theta": 1e4, "head_dim": 16}}
    spec[2]["residual"][0]["sequential"][1] = {"linear": {"in_features": 32, "out_features": 64}}
    gq = nnm.NeuralNetworkModel("gqa", dsl.Mapper(spec, copy.deepcopy(dsl.DEFAULT_ADAMW)))
    assert gq._plan is None                 # RoPE / grouped-query attention is the Gemma family: not this path
    ok = build_gpt("ok", 1, 2, 32, 16, 11, dropout=0.1)
    assert ok._plan is not None and ok._plan.active_dropout(True) and not ok._plan.active_dropout(False)


def test_shard_loader_windows_match_the_reference(golden, tmp_path, monkeypatch):
    g, meta = golden
    ld = meta["loader"]
    monkeypatch.chdir(tmp_path)
    os.makedirs("data")
    for k in range(3):
        np.save(os.path.join("data", f"ldr_{k:06d}"),
                np.random.default_rng(1234 + k).integers(0, ld["vocab"], size=ld["shard_tokens"] + 37 * k).astype(np.uint16))
    from penroz.shards import Loader
    for case in ld["cases"]:
        buf = case["B"] * case["T"]
        lo = Loader("ldr", 0, buf * case["rank"], buf, buf * case["world"])
        assert lo.list() == ["ldr_000000.npy", "ldr_000001.npy", "ldr_000002.npy"]
        for rec in case["rec"]:
            x, y = lo.next_batch()
            assert x.dtype == np.int32
            got = [int(x[0]), int(x[-1]), int(y[0]), int(y[-1]), int(x.astype(np.int64).sum()),
                   int(y.astype(np.int64).sum()), len(x), lo.shard_idx, lo.token_idx]
            assert got == rec
    lo = Loader("ldr", 1, 0, 64, 64)
    for rec in ld["offset0"]:
        x, y = lo.next_batch(target_offset=0)
        assert y is None and [int(x[0]), int(x[-1]), int(x.astype(np.int64).sum()), len(x)] == rec


def test_dropin_on_cpu_follows_the_reference(golden, monkeypatch):
    """Host-side module composition (the route CPU-resident models take) against the golden run."""
    g, _ = golden
    monkeypatch.setenv("PENROZ_INFERENCE_DEVICE", "cpu")
    scratch_cwd()
    from penroz import shards
    shards.write_synthetic_shards("synth65", 65, 1, 2 ** 17)
    m = build_gpt("cpu", 2, 4, 128, 64, 65)
    assert m.generate_tokens(g["tiny/gen_prompt"].tolist(), 64, 80, temperature=0.0) == g["tiny/gen_greedy_b1"].tolist()
    torch.manual_seed(42)
    assert m.generate_tokens(g["tiny/gen_prompt"].tolist(), 64, 30, temperature=1.0, top_k=5) == g["tiny/gen_topk5_seed42"].tolist()
    m.train_model("synth65", 0, 5, 16, 64, 16)
    np.testing.assert_allclose([p["cost"] for p in m.progress], g["tiny/train_loss"][:5], rtol=1e-5)
    assert m.status["code"] == "Trained" and set(m.stats) == {"layers", "weights"}
    assert set(m.progress[0]) == {"dt", "epoch", "durationInSecs", "speedPerSec", "cost", "weight_upd_ratio"}


def test_flat_state_keeps_parameters_keys_and_optimizer_format():
    from penroz.engine import FlatState
    m = build_gpt("flat", 1, 2, 32, 16, 11)
    keys = list(m.state_dict().keys())
    before = {k: v.clone() for k, v in m.state_dict().items()}
    params = list(m.parameters())
    fl = FlatState(params, need_shadow=False)
    assert list(m.state_dict().keys()) == keys
    for k, v in m.state_dict().items():
        assert torch.equal(v, before[k])
    assert all(off % 64 == 0 for off in fl.offsets)
    assert all(p.data_ptr() == fl.p.data_ptr() + 4 * off for p, off in zip(params, fl.offsets))
    fl.g.fill_(2.0)
    assert all(float(p.grad.mean()) == 2.0 for p in params)
    params[0].grad = None
    fl.rebind_grads()
    assert params[0].grad.data_ptr() == fl.g.data_ptr()
    step = fl.adopt_optimizer_state(m.optimizer)
    assert step == 0
    sd = m.optimizer.state_dict()
    assert set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"} and len(sd["state"]) == len(params)
    fresh = build_gpt("flat2", 1, 2, 32, 16, 11)
    fresh.optimizer.load_state_dict(sd)     # stock torch.optim.AdamW accepts the state we expose


def test_checkpoint_saved_after_fused_steps_resumes_in_stock_adamw():
    """The on-disk optimizer state must resume correctly in the reference (stock torch.optim.AdamW): after loading a
    checkpoint written between fused steps, ONE optimizer.step() advances every parameter's `step` by exactly one.
    (A `step` tensor object shared by all states survives torch.save/torch.load as one object and would be incremented
    once per parameter.)"""
    import io
    from penroz.engine import FlatState
    m = build_gpt("resume", 1, 2, 32, 16, 11)
    params = list(m.parameters())
    fl = FlatState(params, need_shadow=False)
    fl.adopt_optimizer_state(m.optimizer)
    fl.step_vec.fill_(3.0)                                   # what three fused AdamW updates leave behind
    fl.m.normal_(); fl.v.uniform_(0.1, 1.0)
    buf = io.BytesIO()
    torch.save({"state": m.state_dict(), "optim_state": m.optimizer.state_dict()}, buf, pickle_protocol=5)
    buf.seek(0)
    data = torch.load(buf, weights_only=False)
    steps = [st["step"] for st in data["optim_state"]["state"].values()]
    assert len({id(t) for t in steps}) == len(steps) == len(params)
    fresh = build_gpt("resume2", 1, 2, 32, 16, 11)
    fresh.load_state_dict(data["state"])
    fresh.optimizer.load_state_dict(data["optim_state"])
    for p in fresh.parameters():
        p.grad = torch.ones_like(p)
    fresh.optimizer.step()
    assert [float(st["step"]) for st in fresh.optimizer.state.values()] == [4.0] * len(params)
    # and the fused side picks the count up again when a caller's stock optimizer stepped in between
    for p in params:
        m.optimizer.state[p]["step"] += 1
    assert float(fl.step_vec.max()) == 4.0 and float(fl.step_vec.min()) == 4.0


def _sync_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, PKG)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from helpers import build_gpt as bg
    from penroz.engine import FlatState
    from penroz.grad_sync import GradSync, gpt2_segments
    dist.init_process_group("gloo", rank=rank, world_size=world)
    m = bg("sync", 2, 2, 32, 16, 11, seed=rank)           # ranks start with DIFFERENT weights
    fl = FlatState(list(m.parameters()), need_shadow=False)
    segs = gpt2_segments(fl, m._plan)
    gs = GradSync(fl.g, segs, bucket_bytes=40_000)
    gs.broadcast_(fl.p)                                    # rank 0's weights everywhere, as DDP's constructor does
    fl.g.copy_(torch.arange(fl.total, dtype=torch.float32) * (rank + 1))
    order = []
    for name, _, _ in segs:                                # the order our backward finishes the stages
        gs.ready(name)
        order = list(gs.launched)
    gs.wait_all()
    g_all = fl.g.clone()
    # second round, drained bucket by bucket: a bucket is final when drain() yields it, whatever the later ones do
    fl.g.copy_(torch.arange(fl.total, dtype=torch.float32) * (rank + 1))
    for name, _, _ in segs:
        gs.ready(name)
    in_flight = gs.all_in_flight()
    drained = []
    for lo, hi in gs.drain():
        drained.append((lo, hi, bool(torch.equal(fl.g[lo:hi], torch.arange(lo, hi, dtype=torch.float32) * 1.5))))
    torch.save({"p": fl.p.clone(), "g": g_all, "order": order, "buckets": gs.buckets, "covers_all": gs.covers_all,
                "in_flight": in_flight, "after": gs.all_in_flight(), "drained": drained,
                "segs": segs, "total": fl.total}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


def test_bucketed_gradient_sync_world_size_2_gloo(tmp_path):
    import socket
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_sync_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = (torch.load(os.path.join(tmp_path, f"r{r}.pt"), weights_only=False) for r in (0, 1))
    assert torch.equal(r0["p"], r1["p"])                   # broadcast from rank 0
    want = torch.arange(r0["total"], dtype=torch.float32) * 1.5
    covered = torch.zeros(r0["total"], dtype=torch.bool)
    for lo, hi, _ in r0["buckets"]:
        covered[lo:hi] = True
    assert covered.all()                                   # every gradient element is in exactly one bucket
    assert sum(hi - lo for lo, hi, _ in r0["buckets"]) == r0["total"]
    assert torch.equal(r0["g"], want) and torch.equal(r1["g"], want)   # mean over the two ranks
    assert r0["order"] == list(range(len(r0["buckets"])))  # buckets go out in backward-completion order
    assert r0["segs"][0][0] == "head" and r0["segs"][-1][0] == "embed"
    assert 1 < len(r0["buckets"]) < len(r0["segs"]) + 1    # small adjacent stages were merged
    assert r0["covers_all"] and r0["in_flight"] and not r0["after"]
    assert [(lo, hi) for lo, hi, _ in r0["drained"]] == [(lo, hi) for lo, hi, _ in r0["buckets"]]
    assert all(ok for _, _, ok in r0["drained"]) and all(ok for _, _, ok in r1["drained"])


def test_kv_cache_api_dense_store():
    from kv_cache import KVCache, TurboQuantKVCache, KVCacheMetrics, create_kv_cache
    c = KVCache(num_layers=2)
    assert c.get(0) == (None, None) and c.seq_len() == 0
    k1, v1 = torch.randn(2, 3, 5, 8), torch.randn(2, 3, 5, 8)
    fk, fv = c.append(0, k1, v1)
    assert torch.equal(fk, k1) and torch.equal(fv, v1)
    for t in range(40):                                    # grows past the initial capacity
        kn, vn = torch.randn(2, 3, 1, 8), torch.randn(2, 3, 1, 8)
        fk, fv = c.append(0, kn, vn)
        assert fk.shape == (2, 3, 6 + t, 8) and torch.equal(fk[:, :, -1:], kn) and torch.equal(fk[:, :, :5], k1)
    assert c.seq_len(0) == 45 and c.seq_len(1) == 0
    assert c.metrics.num_appends == 41 and c.metrics.total_entries == 45 and c.metrics.compression_ratio == 1.0
    c.clear()
    assert c.seq_len() == 0 and c.metrics == KVCacheMetrics()
    q = TurboQuantKVCache(1)
    fk, _ = q.append(0, k1, v1)
    assert fk.dtype == torch.float32 and (fk - k1).abs().max() < 0.05 and q.metrics.compression_ratio > 2.5
    assert type(create_kv_cache(1)) is KVCache


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="needs the reference checkout (build container only)")
def test_reference_test_suite_passes_against_the_dropin_modules(tmp_path):
    """The reference's own unittest files for the four replaced modules, run with our modules shadowing theirs
    (mappers.py / loaders.py still come from the reference). `parameterized` is not installed: tests/shims has a
    stand-in for parameterized.expand."""
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([PKG, os.path.join(ROOT, "tests", "shims"), "/root/reference"])
    env["PENROZ_INFERENCE_DEVICE"] = "cpu"
    import shutil
    shutil.copy("/root/reference/log_config.json", tmp_path / "log_config.json")
    files = [f"/root/reference/{n}.py" for n in ("test_kv_cache", "test_ddp", "test_neural_net_layers", "test_neural_net_model")]
    res = subprocess.run([sys.executable, "-m", "pytest", "-q", "--import-mode=importlib", "-p", "no:cacheprovider", *files],
                         cwd=tmp_path, env=env, capture_output=True, text=True, timeout=900)
    tail = res.stdout[-1500:]
    assert res.returncode == 0, tail
    assert "151 passed" in tail, tail


def test_model_residency_cache(monkeypatch):
    """deserialize() keeps the last models resident while their cached file is unchanged
    (the service deserializes on every request, main.py:405-489); a rewritten file, delete() and the size limit evict."""
    import time
    import neural_net_model as nnm
    scratch_cwd()
    monkeypatch.setenv("PENROZ_MODEL_CACHE", "1")
    nnm._RESIDENT.clear()
    a = build_gpt("res_a", 1, 2, 32, 16, 11)
    a.serialize()
    m1 = nnm.NeuralNetworkModel.deserialize("res_a")
    m2 = nnm.NeuralNetworkModel.deserialize("res_a")
    assert m1 is m2 and m1 is not a                      # second request: no torch.load, the same resident object
    for k, v in a.state_dict().items():
        assert torch.equal(v, m1.state_dict()[k])
    # the resident object runs under its entry's lock; an unrelated object takes none
    assert nnm._Exclusive(m1).lock is not None and nnm._Exclusive(a).lock is None
    assert m1.generate_tokens([[1, 2, 3]], 16, 4, temperature=0.0) == m2.generate_tokens([[1, 2, 3]], 16, 4, temperature=0.0)
    # another writer (a training process) rewrites the file: the entry is stale, the next request reloads
    time.sleep(0.01)
    with torch.no_grad():
        a.layers[0][0].weight.add_(1.0)
    a.serialize()
    m3 = nnm.NeuralNetworkModel.deserialize("res_a")
    assert m3 is not m1 and torch.equal(m3.state_dict()["layers.0.0.weight"], a.state_dict()["layers.0.0.weight"])
    # the resident object serialising itself stays resident
    m3.progress.append({"cost": 1.0})
    m3.serialize()
    assert nnm.NeuralNetworkModel.deserialize("res_a") is m3
    # size limit (2) and delete()
    for name in ("res_b", "res_c"):
        build_gpt(name, 1, 2, 32, 16, 11).serialize()
        nnm.NeuralNetworkModel.deserialize(name)
    assert len(nnm._RESIDENT) == 2 and not any("res_a" in k for k in nnm._RESIDENT)
    nnm.NeuralNetworkModel.delete("res_c")
    assert len(nnm._RESIDENT) == 1
    with pytest.raises(KeyError):
        nnm.NeuralNetworkModel.deserialize("res_c")
    # switched off: every call loads afresh, as the reference does
    monkeypatch.setenv("PENROZ_MODEL_CACHE", "0")
    assert nnm.NeuralNetworkModel.deserialize("res_b") is not nnm.NeuralNetworkModel.deserialize("res_b")
    nnm._RESIDENT.clear()


def test_resident_model_serialises_concurrent_requests(monkeypatch):
    """Two request threads that deserialize the same resident model and generate at once (the service's threadpool,
    main.py:421-436) must each get the tokens a lone request gets: generation attaches its KV cache to the shared
    layers, so the residency entry's lock has to keep the calls apart."""
    import threading
    import neural_net_model as nnm
    scratch_cwd()
    monkeypatch.setenv("PENROZ_MODEL_CACHE", "1")
    nnm._RESIDENT.clear()
    build_gpt("res_t", 2, 2, 32, 32, 17).serialize()
    prompts = [[[1, 2, 3, 4]], [[9, 8, 7]], [[5, 5, 5, 5, 5]], [[2, 11, 3]]]
    alone = [nnm.NeuralNetworkModel.deserialize("res_t").generate_tokens(p, 32, 20, temperature=0.0) for p in prompts]
    got, errors = [None] * len(prompts), []

    def request(i):
        try:
            model = nnm.NeuralNetworkModel.deserialize("res_t")
            got[i] = list(model.generate_tokens_stream(prompts[i], 32, 20, temperature=0.0))
        except Exception as e:          # noqa: BLE001 - reported below
            errors.append(repr(e))
    for _ in range(3):
        threads = [threading.Thread(target=request, args=(i,)) for i in range(len(prompts))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        assert not errors, errors
        for i, p in enumerate(prompts):
            assert got[i] == alone[i][len(p[0]):], i
    assert len(nnm._RESIDENT) == 1
    nnm._RESIDENT.clear()


def test_stream_advanced_and_closed_from_different_threads(monkeypatch):
    """Starlette drives a sync generator with one threadpool dispatch per next() (the service wraps
    generate_tokens_stream in a StreamingResponse, main.py:427-431): the stream of a resident model is started on one
    thread, advanced on a second and closed on a third. The run lock must be released by the closing thread, a second
    request must wait while the stream is open (also when it runs on the owner's thread), and the model must still be
    locked for callers that hold it after the cache dropped its entry."""
    import threading
    import neural_net_model as nnm
    scratch_cwd()
    monkeypatch.setenv("PENROZ_MODEL_CACHE", "1")
    nnm._RESIDENT.clear()
    build_gpt("res_s", 2, 2, 32, 32, 17).serialize()
    model = nnm.NeuralNetworkModel.deserialize("res_s")
    alone = model.generate_tokens([[1, 2, 3, 4]], 32, 12, temperature=0.0)[4:]
    stream = model.generate_tokens_stream([[1, 2, 3, 4]], 32, 12, temperature=0.0)
    got, errors = [], []

    def on_thread(fn):
        def run():
            try:
                fn()
            except Exception as e:      # noqa: BLE001 - reported below
                errors.append(repr(e))
        t = threading.Thread(target=run)
        t.start()
        t.join()

    on_thread(lambda: got.append(next(stream)))
    lock = model.__dict__["_run_lock"]
    assert lock.locked()
    # a second request on this (any) thread cannot enter while the stream is open
    assert not lock.acquire(blocking=False)
    nnm._RESIDENT.clear()                                   # evicted while a caller still streams from the object
    assert nnm._Exclusive(model).lock is lock
    on_thread(lambda: got.extend(next(stream) for _ in range(5)))
    on_thread(stream.close)
    assert not errors, errors
    assert got == alone[:6]
    assert not lock.locked()
    assert model.generate_tokens([[1, 2, 3, 4]], 32, 12, temperature=0.0)[4:] == alone   # and the model is usable again
    nnm._RESIDENT.clear()
