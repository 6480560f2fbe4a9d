"""CPU oracle for the GPT-2 training step and KV-cache decode step — TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module,
and only as the checker or the timed CPU baseline. The product path (penr-oz-neural-network-v3-torch-ddp_b200/)
never imports it and has no CPU fallback.

What it restates. The reference (derinworks/penr-oz-neural-network-v3-torch-ddp) is pure Python whose arithmetic
lives in PyTorch (requirements.txt:11 pins torch>=2.0,<=2.7.0; this image has torch 2.11.0), a third-party
dependency that is not part of /root/reference. Each function below restates, in plain numpy, the published
algorithm of the torch call the reference makes, anchored on the reference's own call sites:

    embedding + position + sum      neural_net_layers.py:98-127 (Summation(Embedding, PositionEmbedding))
    LayerNorm (eps 1e-5, biased var) mappers.py:31,155,162,171 -> torch.nn.LayerNorm
    Linear                           mappers.py:21,156,158,163,165,172 -> y = x W^T + b
    causal attention, 1/sqrt(D)      neural_net_layers.py:59-95 -> F.scaled_dot_product_attention(is_causal=True)
    GELU erf / tanh                  mappers.py:25,139,164 -> torch.nn.GELU(approximate=...)
    residual add                     neural_net_layers.py:130-135
    mean cross entropy               neural_net_model.py:264-267 -> F.cross_entropy (ignore_index=-100)
    AdamW (wd 0.01 on everything)    mappers.py:267-274 -> torch.optim.AdamW defaults (optim/adamw.py)
    train epoch loop                 neural_net_model.py:614-677 (cost += step_cost/num_steps; one step per epoch)
    shard windows                    loaders.py:65-87
    KV-cache decode control flow     neural_net_model.py:360-406, kv_cache.py:41-68

PARITY PINNED: tests/test_oracle.py checks every function here against tests/golden/golden.npz, which
tests/golden/make_golden.py produced by running the UNMODIFIED reference (imported from /root/reference) on CPU in
fp32: initial logits and loss, every parameter gradient, 50 training-step losses, final weights, greedy and seeded
generations, loader windows. The reference's own tests pin no numeric results on this path.
"""
import math

import numpy as np

try:  # vectorised erf; falls back to math.erf element-wise
    from scipy.special import erf as _erf
except Exception:  # pragma: no cover
    _erf = np.vectorize(math.erf)

SQRT_2_OVER_PI = math.sqrt(2.0 / math.pi)


# ------------------------------------------------------------------ shard windows (loaders.py:65-87)
class LoaderOracle:
    """Strided window over a list of in-memory shards; mirrors loaders.Loader's cursor arithmetic."""

    def __init__(self, shards, begin_shard=0, begin_idx=0, buffer_size=0, idx_offset=0):
        self.shards = shards
        self.shard_idx = begin_shard
        self.token_idx = begin_idx
        self.buffer_size = buffer_size
        self.idx_offset = idx_offset
        self.tokens = np.empty((0,), dtype=np.int32)

    def _load(self):
        return np.asarray(self.shards[self.shard_idx]).astype(np.int32)

    def next_batch(self, target_offset=1):
        if len(self.tokens) == 0:
            self.tokens = self._load()
        for _ in range(len(self.shards)):
            if len(self.tokens) < self.token_idx + self.idx_offset + target_offset:
                self.shard_idx = (self.shard_idx + 1) % len(self.shards)
                self.tokens = np.concatenate((self.tokens[self.token_idx:], self._load()))
                self.token_idx = 0
            else:
                break
        i = self.token_idx
        x = self.tokens[i:i + self.buffer_size]
        y = self.tokens[i + target_offset:i + self.buffer_size + target_offset] if target_offset > 0 else None
        self.token_idx += self.idx_offset
        return x, y


# ------------------------------------------------------------------ primitive ops, forward + backward
def layer_norm_fwd(x, w, b, eps=1e-5):
    mu = x.mean(-1, keepdims=True)
    var = ((x - mu) ** 2).mean(-1, keepdims=True)
    rstd = 1.0 / np.sqrt(var + eps)
    xhat = (x - mu) * rstd
    y = xhat * w + (b if b is not None else 0.0)
    return y, (xhat, rstd)


def layer_norm_bwd(dy, w, cache):
    xhat, rstd = cache
    dw = (dy * xhat).reshape(-1, xhat.shape[-1]).sum(0)
    db = dy.reshape(-1, xhat.shape[-1]).sum(0)
    dyw = dy * w
    dx = rstd * (dyw - dyw.mean(-1, keepdims=True) - xhat * (dyw * xhat).mean(-1, keepdims=True))
    return dx, dw, db


def gelu_fwd(x, tanh=False):
    if tanh:
        return 0.5 * x * (1.0 + np.tanh(SQRT_2_OVER_PI * (x + 0.044715 * x ** 3)))
    return 0.5 * x * (1.0 + _erf(x / math.sqrt(2.0)))


def gelu_grad(x, tanh=False):
    if tanh:
        t = np.tanh(SQRT_2_OVER_PI * (x + 0.044715 * x ** 3))
        return 0.5 * (1.0 + t) + 0.5 * x * (1.0 - t * t) * SQRT_2_OVER_PI * (1.0 + 3 * 0.044715 * x * x)
    cdf = 0.5 * (1.0 + _erf(x / math.sqrt(2.0)))
    pdf = np.exp(-0.5 * x * x) / math.sqrt(2.0 * math.pi)
    return cdf + x * pdf


def attention_fwd(qkv, n_head, k_hist=None, v_hist=None, causal=True):
    """qkv [B,S,3C] -> out [B,S,C]. With k_hist/v_hist [B,H,t,D] (a KV cache) the new keys/values are appended
    first (kv_cache.py:53-60) and, for S == 1, no mask is applied (neural_net_layers.py:86)."""
    B, S, C3 = qkv.shape
    C = C3 // 3
    D = C // n_head
    q, k, v = [t.reshape(B, S, n_head, D).transpose(0, 2, 1, 3) for t in np.split(qkv, 3, axis=-1)]
    if k_hist is not None:
        k = np.concatenate([k_hist, k], axis=2)
        v = np.concatenate([v_hist, v], axis=2)
    s = q @ k.transpose(0, 1, 3, 2) / math.sqrt(D)
    if causal:
        tk = k.shape[2]
        mask = np.tril(np.ones((S, tk), dtype=bool), k=tk - S)
        s = np.where(mask, s, -np.inf)
    s = s - s.max(-1, keepdims=True)
    p = np.exp(s)
    p = p / p.sum(-1, keepdims=True)
    o = (p @ v).transpose(0, 2, 1, 3).reshape(B, S, C)
    return o, (q, k, v, p)


def attention_bwd(dout, n_head, cache):
    q, k, v, p = cache
    B, H, S, D = q.shape
    do = dout.reshape(B, S, H, D).transpose(0, 2, 1, 3)
    dv = p.transpose(0, 1, 3, 2) @ do
    dp = do @ v.transpose(0, 1, 3, 2)
    ds = p * (dp - (dp * p).sum(-1, keepdims=True)) / math.sqrt(D)
    dq = ds @ k
    dk = ds.transpose(0, 1, 3, 2) @ q
    back = lambda t: t.transpose(0, 2, 1, 3).reshape(B, S, H * D)
    return np.concatenate([back(dq), back(dk), back(dv)], axis=-1)


def cross_entropy(logits2d, targets, ignore_index=-100):
    """mean over rows with target != ignore_index; returns (loss, dlogits)."""
    valid = targets != ignore_index
    z = logits2d - logits2d.max(-1, keepdims=True)
    lse = np.log(np.exp(z).sum(-1, keepdims=True))
    logp = z - lse
    n = max(1, int(valid.sum()))
    rows = np.arange(len(targets))
    tsafe = np.where(valid, targets, 0)
    loss = -(logp[rows, tsafe] * valid).sum() / n
    d = np.exp(logp)
    d[rows, tsafe] -= 1.0
    d = d * valid[:, None] / n
    return loss, d


# ------------------------------------------------------------------ the model (state-dict keys of the reference)
class Gpt2Oracle:
    def __init__(self, state: dict, n_layer: int, n_head: int, tanh_gelu=False, dtype=np.float64):
        self.p = {k: np.asarray(v, dtype=dtype).copy() for k, v in state.items()}
        self.L, self.H, self.tanh, self.dtype = n_layer, n_head, tanh_gelu, dtype
        self.adam = None

    def _get(self, key):
        return self.p.get(key)

    def forward(self, idx, targets=None, pos_offset=0, kv=None, keep=False):
        """idx [B,S] ints -> (logits [B,S,V], loss|None). keep=True stores what backward() needs.
        kv: list of per-layer [k,v] arrays ([B,H,t,D]) updated in place (decode)."""
        p, L, H = self.p, self.L, self.H
        B, S = idx.shape
        x = p["layers.0.0.weight"][idx] + p["layers.0.1.weight"][pos_offset:pos_offset + S][None]
        tape = {"idx": idx, "blocks": []}
        for l in range(L):
            pre = f"layers.{2 + l}."
            h1, c1 = layer_norm_fwd(x, p[pre + "0.0.weight"], self._get(pre + "0.0.bias"))
            qkv = h1 @ p[pre + "0.1.weight"].T + p[pre + "0.1.bias"]
            if kv is not None:
                kh, vh = kv[l]
                att, ca = attention_fwd(qkv, H, kh, vh, causal=(S > 1))
                kv[l] = [ca[1], ca[2]]
            else:
                att, ca = attention_fwd(qkv, H)
            x_mid = x + (att @ p[pre + "0.3.weight"].T + p[pre + "0.3.bias"])
            h2, c2 = layer_norm_fwd(x_mid, p[pre + "1.0.weight"], self._get(pre + "1.0.bias"))
            hpre = h2 @ p[pre + "1.1.weight"].T + p[pre + "1.1.bias"]
            g = gelu_fwd(hpre, self.tanh)
            x_out = x_mid + (g @ p[pre + "1.3.weight"].T + p[pre + "1.3.bias"])
            if keep:
                tape["blocks"].append((h1, c1, ca, att, h2, c2, hpre, g))
            x = x_out
        hf, cf = layer_norm_fwd(x, p[f"layers.{2 + L}.weight"], self._get(f"layers.{2 + L}.bias"))
        logits = hf @ p[f"layers.{3 + L}.weight"].T
        loss = None
        if targets is not None:
            loss, dlogits = cross_entropy(logits.reshape(B * S, -1), targets.reshape(-1))
            tape["dlogits"] = dlogits.reshape(logits.shape)
        if keep:
            tape["hf"], tape["cf"] = hf, cf
            self.tape = tape
        return logits, loss

    def backward(self, scale=1.0):
        """Gradients of scale * loss for the last forward(keep=True); returns {key: grad}."""
        p, L, H, tp = self.p, self.L, self.H, self.tape
        g = {}
        C = p["layers.0.0.weight"].shape[1]
        flat = lambda t: t.reshape(-1, t.shape[-1])
        dlog = tp["dlogits"] * scale
        g[f"layers.{3 + L}.weight"] = flat(dlog).T @ flat(tp["hf"])
        dhf = dlog @ p[f"layers.{3 + L}.weight"]
        dx, dw, db = layer_norm_bwd(dhf, p[f"layers.{2 + L}.weight"], tp["cf"])
        g[f"layers.{2 + L}.weight"] = dw
        if f"layers.{2 + L}.bias" in p:
            g[f"layers.{2 + L}.bias"] = db
        for l in range(L - 1, -1, -1):
            pre = f"layers.{2 + l}."
            h1, c1, ca, att, h2, c2, hpre, gact = tp["blocks"][l]
            g[pre + "1.3.weight"] = flat(dx).T @ flat(gact)
            g[pre + "1.3.bias"] = flat(dx).sum(0)
            dh = (dx @ p[pre + "1.3.weight"]) * gelu_grad(hpre, self.tanh)
            g[pre + "1.1.weight"] = flat(dh).T @ flat(h2)
            g[pre + "1.1.bias"] = flat(dh).sum(0)
            d2, dw, db = layer_norm_bwd(dh @ p[pre + "1.1.weight"], p[pre + "1.0.weight"], c2)
            g[pre + "1.0.weight"] = dw
            if pre + "1.0.bias" in p:
                g[pre + "1.0.bias"] = db
            dx = dx + d2
            g[pre + "0.3.weight"] = flat(dx).T @ flat(att)
            g[pre + "0.3.bias"] = flat(dx).sum(0)
            dqkv = attention_bwd(dx @ p[pre + "0.3.weight"], H, ca)
            g[pre + "0.1.weight"] = flat(dqkv).T @ flat(h1)
            g[pre + "0.1.bias"] = flat(dqkv).sum(0)
            d1, dw, db = layer_norm_bwd(dqkv @ p[pre + "0.1.weight"], p[pre + "0.0.weight"], c1)
            g[pre + "0.0.weight"] = dw
            if pre + "0.0.bias" in p:
                g[pre + "0.0.bias"] = db
            dx = dx + d1
        idx = tp["idx"]
        gw = np.zeros_like(p["layers.0.0.weight"])
        np.add.at(gw, idx.reshape(-1), flat(dx))
        g["layers.0.0.weight"] = gw
        gp = np.zeros_like(p["layers.0.1.weight"])
        gp[:idx.shape[1]] = dx.sum(0)
        g["layers.0.1.weight"] = gp
        return g

    def adamw_step(self, grads, lr=6e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01):
        """torch.optim.AdamW, one param group, decay on every parameter."""
        if self.adam is None:
            self.adam = {"t": 0, "m": {k: np.zeros_like(v) for k, v in self.p.items()},
                         "v": {k: np.zeros_like(v) for k, v in self.p.items()}}
        a = self.adam
        a["t"] += 1
        b1, b2 = betas
        bc1, bc2 = 1 - b1 ** a["t"], 1 - b2 ** a["t"]
        for k in self.p:
            gk = grads[k]
            self.p[k] *= (1 - lr * weight_decay)
            a["m"][k] += (gk - a["m"][k]) * (1 - b1)
            a["v"][k] = a["v"][k] * b2 + (1 - b2) * gk * gk
            denom = np.sqrt(a["v"][k]) / math.sqrt(bc2) + eps
            self.p[k] -= (lr / bc1) * a["m"][k] / denom

    def train_epoch(self, loader, num_steps, batch_size, block_size, **optim):
        """One reference 'epoch' = one optimizer step over num_steps micro-batches (neural_net_model.py:614-677)."""
        total, cost = None, 0.0
        for _ in range(num_steps):
            x, y = loader.next_batch()
            x = np.asarray(x, dtype=np.int64).reshape(batch_size, block_size)
            y = np.asarray(y, dtype=np.int64).reshape(batch_size, block_size)
            _, loss = self.forward(x, y, keep=True)
            g = self.backward(1.0 / num_steps)
            total = g if total is None else {k: total[k] + g[k] for k in g}
            cost += loss / num_steps
        self.adamw_step(total, **optim)
        return cost

    @staticmethod
    def ddp_train_epoch(replicas, loaders, num_steps, batch_size, block_size, **optim):
        """One epoch of the reference under DistributedDataParallel, all ranks emulated in this process.

        replicas[r] / loaders[r] are rank r's model copy and its loader window (neural_net_model.py:581-594).
        DDP averages gradients over ranks during backward only while require_backward_grad_sync is set; the reference
        assigns that flag AFTER forward (neural_net_model.py:647-648), so the value seen by micro-step i's backward
        is the one written in micro-step i-1: step 0 sees the constructor default True, later steps see False unless
        num_steps == 1. Each rank then applies AdamW to its own gradient buffer (:664), and the logged
        cost is the all-reduced mean of the ranks' epoch costs (:667-668, ddp.py:80-85)."""
        world = len(replicas)
        totals, costs = [None] * world, [0.0] * world
        for i in range(num_steps):
            grads = []
            for r in range(world):
                x, y = loaders[r].next_batch()
                x = np.asarray(x, dtype=np.int64).reshape(batch_size, block_size)
                y = np.asarray(y, dtype=np.int64).reshape(batch_size, block_size)
                _, loss = replicas[r].forward(x, y, keep=True)
                grads.append(replicas[r].backward(1.0 / num_steps))
                costs[r] += loss / num_steps
            if i == 0:      # the only micro-step whose backward exchanges gradients
                mean = {k: sum(g[k] for g in grads) / world for k in grads[0]}
                grads = [{k: v.copy() for k, v in mean.items()} for _ in range(world)]
            for r in range(world):
                totals[r] = grads[r] if totals[r] is None else {k: totals[r][k] + grads[r][k] for k in grads[r]}
        for r in range(world):
            replicas[r].adamw_step(totals[r], **optim)
        return sum(costs) / world

    def generate_greedy(self, context, block_size, max_new_tokens, stop_token=None):
        """generate_tokens(temperature=0.0) with the KV cache: prefill, incremental steps, and clear + re-prefill
        once the cache holds block_size tokens (neural_net_model.py:376-389,457-479). Returns row 0."""
        ctx = np.asarray(context, dtype=np.int64)
        kv = None
        for _ in range(max_new_tokens):
            cached = 0 if kv is None else (kv[0][0].shape[2] if kv[0][0] is not None else 0)
            if cached > 0 and cached < block_size:
                inp, off = ctx[:, -1:], cached
            else:
                kv = [[None, None] for _ in range(self.L)]
                inp, off = ctx[:, -block_size:], 0
            logits, _ = self.forward(inp, None, pos_offset=off, kv=kv)
            nxt = logits[:, -1, :].argmax(-1)[:, None]
            ctx = np.concatenate([ctx, nxt], axis=1)
            if stop_token is not None and int(nxt[0, 0]) == stop_token:
                break
        return ctx[0].tolist()
