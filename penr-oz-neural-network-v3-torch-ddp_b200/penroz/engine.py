"""Fused GPT-2 step engine: runs the matched module tree (plan.py) through the sm_100a kernels.

What it replaces (reference): NeuralNetworkModel.forward (neural_net_model.py:250-271) + autograd .backward()
(:650-653) + optimizer.zero_grad()/step() (:622,677) for training, and the forward used by compute_output /
evaluate_model / _generate_next_token (:273-406) for inference.

Precision modes (chosen from the parameter dtype and whether the caller trains under autocast):
  "f32"  fp32 parameters, everything fp32 (FP32-pipe GEMM, exact attention)       -> CPU-oracle parity, rtol 1e-4
  "amp"  fp32 master parameters + bf16 shadow copies; residual stream and LayerNorm statistics fp32, every Linear /
         GELU / attention output rounded once to bf16 — the rounding points of torch.amp.autocast(bf16)
         (neural_net_model.py:567-578,633)
  "bf16" bf16 parameters end to end (HF-imported checkpoints, neural_net_model.py:221-222), inference only

Training keeps the reference's semantics that change numbers: untied LM head, AdamW weight decay on every parameter,
no gradient clipping unless asked.
"""
import os
from typing import Optional

import torch

from . import ops
from ._lib import (EPI_ACCUM_F32, EPI_DGELU, EPI_GELU, EPI_RESID_BF16, EPI_RESID_F32, EPI_STORE, PenrozError,
                   require_device)
from .plan import Gpt2Plan

_ALIGN = 64  # elements: keeps every parameter 256-byte aligned in the fp32 buffers and 128-byte aligned in bf16


def _round_up(n: int, a: int) -> int:
    return (n + a - 1) // a * a


class FlatState:
    """All parameters of the model re-homed (as views) into one fp32 buffer, with matching flat gradient, AdamW
    moment and bf16 shadow buffers. Parameter objects, their order and their state-dict keys are unchanged."""

    def __init__(self, params: list, need_shadow: bool):
        self.params = params
        dev = params[0].device
        self.offsets = []
        total = 0
        for p in params:
            if p.dtype != torch.float32:
                raise PenrozError(f"fused training needs float32 master parameters, found {p.dtype}")
            self.offsets.append(total)
            total += _round_up(p.numel(), _ALIGN)
        self.total = total
        self.p = torch.zeros(total, device=dev, dtype=torch.float32)
        self.g = torch.zeros(total, device=dev, dtype=torch.float32)
        self.m = torch.zeros(total, device=dev, dtype=torch.float32)
        self.v = torch.zeros(total, device=dev, dtype=torch.float32)
        self.shadow = torch.empty(total, device=dev, dtype=torch.bfloat16) if need_shadow else None
        with torch.no_grad():
            for p, off in zip(params, self.offsets):
                view = self.p[off:off + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p.grad = self.g[off:off + p.numel()].view(p.shape)
        self._index = {id(p): i for i, p in enumerate(params)}
        # which tracked tensor (= parameter index, 2-D weights only) each 64-element group of the flat buffer belongs to
        seg = torch.full((total // _ALIGN,), -1, dtype=torch.int32)
        for i, (p, off) in enumerate(zip(params, self.offsets)):
            if p.ndim == 2:
                seg[off // _ALIGN:(off + _round_up(p.numel(), _ALIGN)) // _ALIGN] = i
        self.seg_of_group = seg.to(dev)
        self.update_stats = torch.zeros((len(params), 4), device=dev, dtype=torch.float64)
        if need_shadow:
            self.refresh_shadow()

    def refresh_shadow(self):
        if self.shadow is not None:
            ops.cast(self.p, self.shadow)

    def _view(self, buf, p):
        off = self.offsets[self._index[id(p)]]
        return buf[off:off + p.numel()].view(p.shape)

    def grad(self, p):
        return self._view(self.g, p)

    def shadow_of(self, p):
        return self._view(self.shadow, p)

    def rebind_grads(self):
        """zero_grad(set_to_none=True) or foreign code may have dropped .grad: point it back at the flat buffer."""
        for p, off in zip(self.params, self.offsets):
            g = p.grad
            want = self.g[off:off + p.numel()]
            if g is None or g.data_ptr() != want.data_ptr():
                p.grad = want.view(p.shape)

    def adopt_optimizer_state(self, optimizer) -> int:
        """Move any existing AdamW moments into the flat buffers and make optimizer.state[p] views of them.
        Returns the step count found (0 for a fresh optimizer). Every state's `step` is its OWN 0-dim element of one
        host vector: the fused update sets all of them with a single fill, while a stock torch.optim.AdamW that later
        loads the saved state (the reference resuming from this checkpoint) increments each parameter's count exactly
        once per step -- a tensor object shared between the states would be bumped once per parameter."""
        step = 0
        for p in self.params:
            st = optimizer.state.get(p)
            if st:
                if "exp_avg" in st:
                    self._view(self.m, p).copy_(st["exp_avg"].to(self.p.device, torch.float32))
                    self._view(self.v, p).copy_(st["exp_avg_sq"].to(self.p.device, torch.float32))
                s = st.get("step", 0)
                step = max(step, int(s.item()) if torch.is_tensor(s) else int(s))
        self.step_vec = torch.full((len(self.params),), float(step), dtype=torch.float32)
        self.step_views = [self.step_vec[i] for i in range(len(self.params))]
        for i, p in enumerate(self.params):
            optimizer.state[p] = {"step": self.step_views[i], "exp_avg": self._view(self.m, p),
                                  "exp_avg_sq": self._view(self.v, p)}
        return step


class Gpt2Engine:
    def __init__(self, plan: Gpt2Plan, parameters: list, optimizer=None):
        require_device()
        self.plan = plan
        self.parameters = parameters
        self.optimizer = optimizer
        self.device = plan.wte.weight.device
        self.flat: Optional[FlatState] = None
        self.mode = None
        self._ws = {}
        self.grad_sync = None          # penroz.grad_sync.GradSync when world_size > 1
        # rows of logits alive at once: 6144 = 24 CTA-pair tiles of 256 rows, so the chunk's dX GEMM (3 column tiles) is
        # 72 units = one wave of the 74 CTA pairs; the [rows, 50304] bf16 chunk is 618 MB
        self.head_chunk_rows = int(os.environ.get("PENROZ_HEAD_CHUNK_ROWS", "6144"))
        self.clip_grad_norm = float(os.environ.get("PENROZ_CLIP_GRAD_NORM", "0") or 0.0)
        self._err_flag = torch.zeros(1, device=self.device, dtype=torch.int32)
        self.keep_for_stats = None     # dict filled on stats epochs (rank 0)

    # ------------------------------------------------------------------ parameter access
    def _wb(self, p):
        """weight/bias operand for a Linear in the current mode"""
        if p is None:
            return None
        return self.flat.shadow_of(p) if self.mode == "amp" else p.data

    def param_mode(self) -> str:
        dt = self.plan.wte.weight.dtype
        if dt == torch.bfloat16:
            return "bf16"
        if dt == torch.float32:
            return "f32"
        raise PenrozError(f"unsupported parameter dtype {dt}")

    # ------------------------------------------------------------------ inference
    def _linear_infer(self, x2d, lin, epilogue=EPI_STORE, aux=None, tanh=False):
        M = x2d.shape[0]
        w, b = lin.weight.data, (lin.bias.data if lin.bias is not None else None)
        skinny_rows = 64 if w.dtype == torch.float32 else 16
        if M <= skinny_rows:
            epi = {EPI_STORE: 0, EPI_GELU: 1, EPI_RESID_F32: 2, EPI_RESID_BF16: 2}[epilogue]
            return ops.skinny_linear(x2d, w, b, epilogue=epi, aux=aux, tanh=tanh)
        if epilogue == EPI_GELU:
            return ops.gemm(x2d, w, epilogue=EPI_GELU, bias=b, tanh=tanh)[1]
        return ops.gemm(x2d, w, epilogue=epilogue, bias=b, aux=aux, out_dtype=w.dtype)

    # ------------------------------------------------------------------ native single-token decode step
    def _native_layers(self, kv):
        """ctypes array of penroz_gpt2_layer_t for this model and this cache's page pools (built once per cache)."""
        import ctypes
        store = kv._paged
        cached = getattr(store, "native_layers", None)
        if cached is not None:
            return cached

        class Layer(ctypes.Structure):
            _fields_ = [(n, ctypes.c_void_p) for n in ("ln1_w", "ln1_b", "qkv_w", "qkv_b", "proj_w", "proj_b", "ln2_w",
                                                        "ln2_b", "fc_w", "fc_b", "proj2_w", "proj2_b", "kv_pool")] + \
                       [("ln1_eps", ctypes.c_float), ("ln2_eps", ctypes.c_float), ("gelu_tanh", ctypes.c_int),
                        ("ffn", ctypes.c_int), ("kv_scales", ctypes.c_void_p), ("kv_quant", ctypes.c_int),
                        ("kv_recip", ctypes.c_int)]

        def ptr(p):
            return None if p is None else p.data_ptr()

        arr = (Layer * self.plan.n_layer)()
        for i, blk in enumerate(self.plan.blocks):
            L = arr[i]
            L.ln1_w, L.ln1_b = ptr(blk.ln1.weight), ptr(blk.ln1.bias)
            L.qkv_w, L.qkv_b = ptr(blk.qkv.weight), ptr(blk.qkv.bias)
            L.proj_w, L.proj_b = ptr(blk.proj.weight), ptr(blk.proj.bias)
            L.ln2_w, L.ln2_b = ptr(blk.ln2.weight), ptr(blk.ln2.bias)
            L.fc_w, L.fc_b = ptr(blk.fc.weight), ptr(blk.fc.bias)
            L.proj2_w, L.proj2_b = ptr(blk.proj2.weight), ptr(blk.proj2.bias)
            L.kv_pool = store.pools[i].data_ptr()
            L.ln1_eps, L.ln2_eps = blk.ln1.eps, blk.ln2.eps
            L.gelu_tanh, L.ffn = int(blk.gelu_tanh), blk.fc.out_features
            L.kv_scales = store.scale_pools[i].data_ptr() if store.quant else None
            L.kv_quant, L.kv_recip = int(store.quant), 0      # true division: the arithmetic of the reference's CPU decode
        store.native_layers = arr
        return arr

    @torch.no_grad()
    def decode_step(self, tokens, pos, kv):
        """One incremental decode step for tokens [B,1] at position `pos` through penroz_gpt2_decode_step (one host
        call enqueues every kernel of the step). Returns (logits [B,1,V], greedy ids [B,1])."""
        pl = self.plan
        B = tokens.shape[0]
        C, H, V = pl.n_embd, pl.n_head, pl.vocab
        if pos + 1 > pl.block_size:
            raise IndexError(f"position {pos + 1} exceeds the position table ({pl.block_size})")
        store = kv._paged
        store.ensure_pages(pos + 1)
        layers = self._native_layers(kv)
        F = max(b.fc.out_features for b in pl.blocks)
        need = ops.lib().gpt2_decode_workspace_bytes(B, C, H, F, pl.block_size)
        ws = getattr(store, "native_ws", None)
        if ws is None or ws.numel() < need:
            ws = store.native_ws = torch.zeros(need, device=self.device, dtype=torch.uint8)   # zeroed once: holds the step's barrier words
        dt = pl.wte.weight.dtype
        logits = torch.empty((B, 1, V), device=self.device, dtype=dt)
        ids = torch.empty((B, 1), device=self.device, dtype=torch.int64)
        self._err_flag.zero_()
        lnf_b = None if pl.lnf.bias is None else pl.lnf.bias.data_ptr()
        head_b = None if pl.head.bias is None else pl.head.bias.data_ptr()
        ops.lib().gpt2_decode_step(tokens.contiguous().data_ptr(), pl.wte.weight.data_ptr(), pl.wpe.weight.data_ptr(),
                                   layers, pl.n_layer, pl.lnf.weight.data_ptr(), lnf_b, pl.lnf.eps,
                                   pl.head.weight.data_ptr(), head_b, store.table_dev.data_ptr(), store.max_pages, B, C,
                                   H, V, pos, ops.dt(pl.wte.weight), ws.data_ptr(), ws.numel(), logits.data_ptr(),
                                   ids.data_ptr(), self._err_flag.data_ptr(), ops.stream())
        kv.account_tokens(1, 2 * B * C * logits.element_size(), pl.n_layer)
        kv.commit(1)
        return logits, ids

    @torch.no_grad()
    def prefill(self, tokens, kv):
        """The prompt pass of a generation ([B,S] tokens, empty paged cache) through penroz_gpt2_prefill: one host call
        enqueues every kernel, K/V of all S positions land in the pages, LayerNorm + LM head run on the last position
        only. Returns (logits [B,1,V], greedy ids [B,1])."""
        pl = self.plan
        B, S = tokens.shape
        C, H, V = pl.n_embd, pl.n_head, pl.vocab
        if S > pl.block_size:
            raise IndexError(f"position {S} exceeds the position table ({pl.block_size})")
        store = kv._paged
        if store.length != 0:
            raise PenrozError("multi-token append onto a non-empty KV cache is not part of the decode path")
        store.ensure_pages(S)
        layers = self._native_layers(kv)
        F = max(b.fc.out_features for b in pl.blocks)
        need = ops.lib().gpt2_prefill_workspace_bytes(B, S, C, F)
        ws = getattr(store, "prefill_ws", None)
        if ws is None or ws.numel() < need:
            ws = store.prefill_ws = torch.empty(need, device=self.device, dtype=torch.uint8)
        dt = pl.wte.weight.dtype
        logits = torch.empty((B, 1, V), device=self.device, dtype=dt)
        ids = torch.empty((B, 1), device=self.device, dtype=torch.int64)
        self._err_flag.zero_()
        lnf_b = None if pl.lnf.bias is None else pl.lnf.bias.data_ptr()
        head_b = None if pl.head.bias is None else pl.head.bias.data_ptr()
        ops.lib().gpt2_prefill(tokens.contiguous().data_ptr(), pl.wte.weight.data_ptr(), pl.wpe.weight.data_ptr(),
                               layers, pl.n_layer, pl.lnf.weight.data_ptr(), lnf_b, pl.lnf.eps,
                               pl.head.weight.data_ptr(), head_b, store.table_dev.data_ptr(), store.max_pages, B, S, C,
                               H, V, ops.dt(pl.wte.weight), ws.data_ptr(), ws.numel(), logits.data_ptr(),
                               ids.data_ptr(), self._err_flag.data_ptr(), ops.stream())
        kv.account_tokens(S, 2 * B * S * C * logits.element_size(), pl.n_layer)
        kv.commit(S)
        return logits, ids

    @torch.no_grad()
    def forward_infer(self, idx, targets=None, pos_offset=0, kv=None, last_only=False, collect=False):
        """Forward in the parameter dtype, no autograd. Returns (logits [B,S|1,V], loss|None, activations|None).

        kv: a penroz PagedKV (kv_cache.py) — S == 1 appends the token and runs split-KV decode attention; S > 1
        (prefill, cache empty) runs causal attention and bulk-appends K/V."""
        pl = self.plan
        mode = self.param_mode()
        B, S = idx.shape
        C, H = pl.n_embd, pl.n_head
        if pos_offset + S > pl.block_size:
            raise IndexError(f"position {pos_offset + S} exceeds the position table ({pl.block_size})")
        resid = EPI_RESID_BF16 if mode == "bf16" else EPI_RESID_F32
        self._err_flag.zero_()
        x = ops.embed_fwd(idx, pl.wte.weight.data, pl.wpe.weight.data, pos_offset, err_flag=self._err_flag)
        acts = [x, x] if collect else None
        x2 = x.view(B * S, C)
        t0 = kv.seq_len() if kv is not None else 0
        for li, blk in enumerate(pl.blocks):
            ln1, _, _ = ops.layernorm_fwd(x2, blk.ln1.weight.data, None if blk.ln1.bias is None else blk.ln1.bias.data,
                                          blk.ln1.eps, save_stats=False)
            qkv = self._linear_infer(ln1, blk.qkv)
            if kv is None:
                att, _ = ops.attn_fwd(qkv.view(B, S, 3 * C), H, want_lse=False)
                att = att.view(B * S, C)
            else:
                kv.write_packed(li, qkv, B, S, t0)
                if S == 1 and kv.scale_pool(li) is not None:
                    att = ops.attn_decode_q8(qkv, kv.pool(li), kv.scale_pool(li), kv.block_table, B, H, C // H, t0 + 1,
                                             scratch=kv.scratch)
                elif S == 1:
                    att = ops.attn_decode(qkv, kv.pool(li), kv.block_table, B, H, C // H, t0 + 1, scratch=kv.scratch)
                else:
                    if t0 != 0:
                        raise PenrozError("multi-token append onto a non-empty KV cache is not part of the decode path")
                    att, _ = ops.attn_fwd(qkv.view(B, S, 3 * C), H, want_lse=False)
                    att = att.view(B * S, C)
            x2 = self._linear_infer(att, blk.proj, resid, aux=x2)
            ln2, _, _ = ops.layernorm_fwd(x2, blk.ln2.weight.data, None if blk.ln2.bias is None else blk.ln2.bias.data,
                                          blk.ln2.eps, save_stats=False)
            g = self._linear_infer(ln2, blk.fc, EPI_GELU, tanh=blk.gelu_tanh)
            x2 = self._linear_infer(g, blk.proj2, resid, aux=x2)
            if collect:
                acts.append(x2.view(B, S, C))
        if kv is not None:
            kv.commit(S)
        lnf, _, _ = ops.layernorm_fwd(x2, pl.lnf.weight.data, None if pl.lnf.bias is None else pl.lnf.bias.data,
                                      pl.lnf.eps, save_stats=False)
        if collect:
            acts.append(lnf.view(B, S, C))
        if last_only and targets is None:
            rows = lnf.view(B, S, C)[:, -1, :].contiguous()
            logits = self._linear_infer(rows, pl.head).view(B, 1, pl.vocab)
        else:
            logits = self._linear_infer(lnf, pl.head).view(B, S, pl.vocab)
        loss = None
        if targets is not None:
            loss = self._ce_forward_only(logits.view(B * S, pl.vocab), targets.reshape(-1))
        if collect:
            acts.append(logits)
        self._check_index_errors()
        return logits, loss, acts

    def _ce_forward_only(self, logits2d, targets):
        M = logits2d.shape[0]
        nv = torch.zeros(1, device=self.device, dtype=torch.int32)
        rows = torch.empty(M, device=self.device, dtype=torch.float32)
        loss = torch.zeros((), device=self.device, dtype=torch.float32)
        targets = targets.contiguous()
        ops.ce_count_valid(targets, nv)
        ops.ce_rows(logits2d, targets, nv, rows, write_grad=False, err_flag=self._err_flag)
        ops.ce_reduce(rows, nv, loss)
        return loss

    def _check_index_errors(self):
        # deferred: only looked at when someone already synchronises (cost.item(), tolist(), ...)
        pass

    def raise_on_index_errors(self):
        code = int(self._err_flag.item())
        if code == 1:
            raise IndexError("index out of range in self")  # torch's embedding message
        if code == 2:
            raise IndexError("Target is out of bounds")

    # ------------------------------------------------------------------ training set-up
    def setup_training(self, use_amp: bool):
        if self.plan.active_dropout(True):
            raise PenrozError("dropout p > 0 is not on the fused training path (SURVEY H6): "
                              "build the model with p = 0.0 or train it through torch")
        if self.param_mode() != "f32":
            raise PenrozError("fused training needs float32 master parameters")
        mode = "amp" if use_amp else "f32"
        if self.flat is None or any(p.data_ptr() != self.flat._view(self.flat.p, p).data_ptr() for p in self.parameters):
            self.flat = FlatState(self.parameters, need_shadow=(mode == "amp"))
            self.step_count = self.flat.adopt_optimizer_state(self.optimizer) if self.optimizer is not None else 0
        elif mode == "amp":
            if self.flat.shadow is None:
                self.flat.shadow = torch.empty(self.flat.total, device=self.device, dtype=torch.bfloat16)
            self.flat.refresh_shadow()
        self.mode = mode
        self.flat.rebind_grads()

    def zero_grad(self):
        # the previous fused AdamW pass may already have cleared the gradients it consumed (one 4-byte store per element
        # it was touching anyway, instead of a separate 4 P-byte fill)
        if not getattr(self, "_grads_clean", False):
            self.flat.g.zero_()
        self.flat.rebind_grads()

    def _buf(self, name, shape, dtype):
        key = (name, tuple(shape), dtype)
        t = self._ws.get(key)
        if t is None:
            t = torch.empty(shape, device=self.device, dtype=dtype)
            self._ws[key] = t
        return t

    def release_workspace(self):
        self._ws.clear()

    # ------------------------------------------------------------------ one micro-step: forward + backward
    def train_micro_step(self, idx, targets, grad_scale: float, sync_grads: bool = False, keep_stats: bool = False,
                         cost_out=None, cost_first: bool = True):
        """Forward + backward of one (B,T) micro-batch; gradients are ACCUMULATED (+=) into the flat buffer scaled by
        grad_scale (= 1/num_steps, neural_net_model.py:636). Returns the micro-batch mean loss (device scalar).
        sync_grads: launch the bucketed all-reduce of each bucket as soon as its gradients are final.
        cost_out: 0-dim fp32 device tensor receiving the epoch cost `cost += step_cost / num_steps` (:636-638) from the
        loss reduction itself (set on the first micro-step, added to on later ones)."""
        pl, fl = self.plan, self.flat
        amp = self.mode == "amp"
        act = torch.bfloat16 if amp else torch.float32
        B, T = idx.shape
        C, H, V, L = pl.n_embd, pl.n_head, pl.vocab, pl.n_layer
        M = B * T
        if T > pl.block_size:
            raise IndexError(f"sequence length {T} exceeds the position table ({pl.block_size})")
        idx = idx.contiguous()
        targets = targets.reshape(-1).contiguous()
        buf = self._buf
        self._grads_clean = False
        gs = self.grad_sync if sync_grads else None
        self._err_flag.zero_()

        # ---------------- forward
        x = [buf(f"x{l}", (M, C), torch.float32) for l in range(L + 1)]
        ops.embed_fwd(idx, pl.wte.weight.data, pl.wpe.weight.data, 0, err_flag=self._err_flag, out=x[0].view(B, T, C))
        saved = []
        for l, blk in enumerate(pl.blocks):
            F = blk.fc.out_features
            s = dict(
                ln1=buf(f"ln1_{l}", (M, C), act), m1=buf(f"m1_{l}", (M,), torch.float32), r1=buf(f"r1_{l}", (M,), torch.float32),
                qkv=buf(f"qkv_{l}", (M, 3 * C), act), att=buf(f"att_{l}", (M, C), act),
                lse=buf(f"lse_{l}", (B, H, T), torch.float32), xmid=buf(f"xmid_{l}", (M, C), torch.float32),
                ln2=buf(f"ln2_{l}", (M, C), act), m2=buf(f"m2_{l}", (M,), torch.float32), r2=buf(f"r2_{l}", (M,), torch.float32),
                h=buf(f"h_{l}", (M, F), act), g=buf(f"g_{l}", (M, F), act))
            saved.append(s)
            ops.layernorm_fwd(x[l], blk.ln1.weight.data, None if blk.ln1.bias is None else blk.ln1.bias.data,
                              blk.ln1.eps, out=s["ln1"], mean=s["m1"], rstd=s["r1"])
            ops.gemm(s["ln1"], self._wb(blk.qkv.weight), bias=self._wb(blk.qkv.bias), out=s["qkv"])
            ops.attn_fwd(s["qkv"].view(B, T, 3 * C), H, out=s["att"].view(B, T, C), lse=s["lse"])
            ops.gemm(s["att"], self._wb(blk.proj.weight), epilogue=EPI_RESID_F32, bias=self._wb(blk.proj.bias),
                     aux=x[l], out=s["xmid"])
            ops.layernorm_fwd(s["xmid"], blk.ln2.weight.data, None if blk.ln2.bias is None else blk.ln2.bias.data,
                              blk.ln2.eps, out=s["ln2"], mean=s["m2"], rstd=s["r2"])
            ops.gemm(s["ln2"], self._wb(blk.fc.weight), epilogue=EPI_GELU, bias=self._wb(blk.fc.bias), out=s["h"],
                     out2=s["g"], tanh=blk.gelu_tanh)
            ops.gemm(s["g"], self._wb(blk.proj2.weight), epilogue=EPI_RESID_F32, bias=self._wb(blk.proj2.bias),
                     aux=s["xmid"], out=x[l + 1])
        lnf = buf("lnf", (M, C), act)
        mf, rf = buf("mf", (M,), torch.float32), buf("rf", (M,), torch.float32)
        ops.layernorm_fwd(x[L], pl.lnf.weight.data, None if pl.lnf.bias is None else pl.lnf.bias.data, pl.lnf.eps,
                          out=lnf, mean=mf, rstd=rf)

        # ---------------- LM head + cross entropy + head backward, chunked over rows: the [M,V] logits never exist
        n_valid = buf("n_valid", (1,), torch.int32)
        loss_rows = buf("loss_rows", (M,), torch.float32)
        loss = torch.zeros((), device=self.device, dtype=torch.float32)
        ops.ce_count_valid(targets, n_valid)
        Vp = _round_up(V, 64)
        # equal chunks (a short tail chunk leaves the K = V data-gradient GEMM with a third of the SMs: 0.5 ms)
        n_chunks = max(1, -(-M // self.head_chunk_rows))
        rows_per = min(M, _round_up(-(-M // n_chunks), 128))
        if keep_stats:
            rows_per = M
        logits = buf("logits", (rows_per, Vp), act)
        d_lnf = buf("d_lnf", (M, C), act)
        w_head = self._wb(pl.head.weight)
        b_head = self._wb(pl.head.bias)
        g_head = fl.grad(pl.head.weight)
        for r0 in range(0, M, rows_per):
            r1 = min(M, r0 + rows_per)
            n = r1 - r0
            lg = logits[:n]
            ops.gemm(lnf[r0:r1], w_head, bias=b_head, out=lg[:, :V])
            if keep_stats:
                self.keep_for_stats["logits"] = lg[:, :V].clone().view(B, T, V)
            ops.ce_rows(lg, targets[r0:r1], n_valid, loss_rows[r0:r1], V=V, grad_scale=grad_scale, write_grad=True,
                        err_flag=self._err_flag)
            if keep_stats:
                self.keep_for_stats["logits_grad"] = lg[:, :V].clone().view(B, T, V)
            dl = lg[:, :V]
            # d lnf = dlogits . W   (B operand MN-major: W is [V, C])
            ops.gemm(dl, w_head, b_layout=1, out=d_lnf[r0:r1])
            # dW += dlogits^T . lnf (both operands MN-major; TMA zero-fills past column V of the padded chunk)
            ops.gemm(dl, lnf[r0:r1], a_layout=1, b_layout=1, epilogue=EPI_ACCUM_F32, out=g_head, accumulate=True)
            if pl.head.bias is not None:
                ops.colsum(dl, fl.grad(pl.head.bias), accumulate=True, scratch=self._scratch(M, self._widest(), B, T))
        ops.ce_reduce(loss_rows, n_valid, loss)
        if cost_out is not None:
            ops.ce_reduce(loss_rows, n_valid, cost_out, out_scale=grad_scale, accumulate=not cost_first)

        # ---------------- backward through the blocks
        scratch = self._scratch(M, self._widest(), B, T)
        dx_a, dx_b = buf("dx_a", (M, C), torch.float32), buf("dx_b", (M, C), torch.float32)
        dxb_a = buf("dxb_a", (M, C), torch.bfloat16) if amp else None
        dxb_b = buf("dxb_b", (M, C), torch.bfloat16) if amp else None
        last = pl.blocks[-1]
        ops.layernorm_bwd(d_lnf, x[L], pl.lnf.weight.data, mf, rf, add=None, dgamma=fl.grad(pl.lnf.weight),
                          dbeta=None if pl.lnf.bias is None else fl.grad(pl.lnf.bias),
                          dcol=None if last.proj2.bias is None else fl.grad(last.proj2.bias), want_bf16=amp,
                          scratch=scratch, dx=dx_a, dxb=dxb_a)
        if keep_stats:
            self.keep_for_stats["lnf"] = lnf.clone().view(B, T, C)
            self.keep_for_stats["lnf_grad"] = d_lnf.clone().view(B, T, C)
            self.keep_for_stats["x"] = [t.clone().view(B, T, C) for t in x]
            self.keep_for_stats["x_grad"] = [None] * (L + 1)
            self.keep_for_stats["x_grad"][L] = dx_a.clone().view(B, T, C)
        if gs is not None:
            gs.ready("head")
        dx, dxb = dx_a, dxb_a          # gradient wrt x[l+1] (fp32) and its bf16 copy
        dx_o, dxb_o = dx_b, dxb_b
        for l in range(L - 1, -1, -1):
            blk, s = pl.blocks[l], saved[l]
            F = blk.fc.out_features
            dy = dxb if amp else dx
            # ---- MLP branch
            dh = buf("dh", (M, F), act)
            ops.gemm(dy, self._wb(blk.proj2.weight), b_layout=1, epilogue=EPI_DGELU, aux=s["h"], out=dh,
                     tanh=blk.gelu_tanh, colsum_out=None if blk.fc.bias is None else fl.grad(blk.fc.bias),
                     colsum_scratch=scratch)
            ops.gemm(dy, s["g"], a_layout=1, b_layout=1, epilogue=EPI_ACCUM_F32, out=fl.grad(blk.proj2.weight),
                     accumulate=True)
            d_ln2 = buf("d_ln", (M, C), act)
            ops.gemm(dh, self._wb(blk.fc.weight), b_layout=1, out=d_ln2)
            ops.gemm(dh, s["ln2"], a_layout=1, b_layout=1, epilogue=EPI_ACCUM_F32, out=fl.grad(blk.fc.weight),
                     accumulate=True)
            ops.layernorm_bwd(d_ln2, s["xmid"], blk.ln2.weight.data, s["m2"], s["r2"], add=dx,
                              dgamma=fl.grad(blk.ln2.weight), dbeta=None if blk.ln2.bias is None else fl.grad(blk.ln2.bias),
                              dcol=None if blk.proj.bias is None else fl.grad(blk.proj.bias), want_bf16=amp,
                              scratch=scratch, dx=dx_o, dxb=dxb_o)
            dx, dx_o = dx_o, dx
            dxb, dxb_o = dxb_o, dxb
            dy = dxb if amp else dx
            # ---- attention branch
            d_att = buf("d_att", (M, C), act)
            # the GEMM that produces dO also leaves delta = rowsum(dO * O) per head at the head of the scratch buffer,
            # where the attention backward looks for it (nothing else touches scratch in between)
            fuse_delta = (ops.rowdot_fusable(M, C, C // H, act) and ops._use_tc_attention(s["qkv"], C // H)
                          and os.environ.get("PENROZ_FUSE_DELTA", "1") != "0")
            if fuse_delta:
                ops.gemm(dy, self._wb(blk.proj.weight), b_layout=1, out=d_att, epilogue=ops.EPI_ROWDOT, aux=s["att"],
                         rowdot_out=scratch, seq_len=T)
            else:
                ops.gemm(dy, self._wb(blk.proj.weight), b_layout=1, out=d_att)
            ops.gemm(dy, s["att"], a_layout=1, b_layout=1, epilogue=EPI_ACCUM_F32, out=fl.grad(blk.proj.weight),
                     accumulate=True)
            d_qkv = buf("d_qkv", (M, 3 * C), act)
            ops.attn_bwd(s["qkv"].view(B, T, 3 * C), s["att"].view(B, T, C), d_att.view(B, T, C), s["lse"], H,
                         scratch=scratch, dqkv=d_qkv.view(B, T, 3 * C),
                         dbias=None if blk.qkv.bias is None else fl.grad(blk.qkv.bias), delta_ready=fuse_delta)
            d_ln1 = buf("d_ln", (M, C), act)
            ops.gemm(d_qkv, self._wb(blk.qkv.weight), b_layout=1, out=d_ln1)
            ops.gemm(d_qkv, s["ln1"], a_layout=1, b_layout=1, epilogue=EPI_ACCUM_F32, out=fl.grad(blk.qkv.weight),
                     accumulate=True)
            prev = pl.blocks[l - 1] if l > 0 else None
            ops.layernorm_bwd(d_ln1, x[l], blk.ln1.weight.data, s["m1"], s["r1"], add=dx,
                              dgamma=fl.grad(blk.ln1.weight), dbeta=None if blk.ln1.bias is None else fl.grad(blk.ln1.bias),
                              dcol=None if (prev is None or prev.proj2.bias is None) else fl.grad(prev.proj2.bias),
                              want_bf16=(amp and l > 0), scratch=scratch, dx=dx_o, dxb=dxb_o)
            dx, dx_o = dx_o, dx
            dxb, dxb_o = dxb_o, dxb
            if keep_stats:
                self.keep_for_stats["x_grad"][l] = dx.clone().view(B, T, C)
            if gs is not None:
                gs.ready(f"block{l}")
        ops.embed_bwd(idx, dx, fl.grad(pl.wte.weight), fl.grad(pl.wpe.weight), 0)
        if gs is not None:
            gs.ready("embed")
        return loss

    # ------------------------------------------------------------------ forward() under autograd
    def forward_with_pending_grads(self, idx, targets, use_amp: bool):
        """forward() called with gradients enabled (a caller's own training loop: cost.backward(); optimizer.step()).
        The hand-written backward cannot be replayed later by autograd, so it runs now, into a PENDING gradient buffer
        of the flat layout; backward_pending() then adds it, scaled by the cost's incoming gradient, to the real
        .grad buffer. Returns (loss, kept activations / activation gradients)."""
        were_none = [p.grad is None for p in self.parameters]
        self.setup_training(use_amp)
        fl = self.flat
        if getattr(fl, "g_pending", None) is None:
            fl.g_pending = torch.zeros_like(fl.g)
        else:
            fl.g_pending.zero_()
        real, fl.g = fl.g, fl.g_pending
        try:
            self.keep_for_stats = {}
            loss = self.train_micro_step(idx, targets, 1.0, sync_grads=False, keep_stats=True)
            kept = self.keep_for_stats
        finally:
            fl.g = real
            self.keep_for_stats = None
        for p, none in zip(self.parameters, were_none):     # zero_grad(set_to_none=True) stays visible until backward
            if none:
                p.grad = None
        return loss, kept

    def _sync_step_count(self):
        """A caller's own torch optimizer.step() between fused steps advances the per-parameter `step` entries (each
        state owns one element of flat.step_vec); the fused update continues from the largest count. States whose
        `step` object was replaced by foreign code (optimizer.load_state_dict) are read and pointed back."""
        if self.optimizer is None:
            return
        fl = self.flat
        count = max(self.step_count, int(fl.step_vec.max().item()))
        for i, p in enumerate(self.parameters):
            st = self.optimizer.state.get(p)
            if st is None:
                continue
            s = st.get("step")
            if s is not fl.step_views[i]:
                if s is not None:
                    count = max(count, int(s.item()) if torch.is_tensor(s) else int(s))
                st["step"] = fl.step_views[i]
        self.step_count = count

    def backward_pending(self, grad_out):
        fl = self.flat
        dropped = [p for p in self.parameters if p.grad is None]
        self._grads_clean = False
        if len(dropped) == len(self.parameters):
            fl.g.zero_()
        else:
            for p in dropped:
                fl.grad(p).zero_()
        fl.rebind_grads()
        ops.axpy_(fl.g, fl.g_pending, grad_out.detach().to(torch.float32).reshape(()).contiguous())

    def _widest(self):
        """widest matrix whose column sums (bias gradients) go through the scratch: logits, any block's FFN, packed QKV"""
        pl = self.plan
        return max(pl.vocab, max(b.fc.out_features for b in pl.blocks), 3 * pl.n_embd)

    def _scratch(self, M, widest, B, T):
        C, H = self.plan.n_embd, self.plan.n_head
        act = torch.bfloat16 if self.mode == "amp" else torch.float32
        need = max(ops.lib().layernorm_bwd_scratch_bytes(M, C), ops.lib().colsum_scratch_bytes(M, widest),
                   ops.attn_bwd_scratch_bytes(B, T, H, C // H, act) + 64)
        return self._buf("scratch", (int(need),), torch.uint8)

    # ------------------------------------------------------------------ optimizer
    def optimizer_step(self, track_update: bool = False, clear_grads: bool = False):
        """One AdamW update over the flat buffers with the hyper-parameters of the (stock) torch optimizer's single
        param group; also refreshes the bf16 shadow weights in the same pass. track_update: the pass also measures the
        update per 2-D weight (update_ratios() reads it). clear_grads: the pass zeroes the gradients it consumed (the
        caller does not look at .grad afterwards), so the next zero_grad() costs nothing."""
        fl = self.flat
        gs = self.grad_sync
        # With every bucket still in flight (the usual one-micro-step epoch) the update runs bucket by bucket in
        # arrival order, so AdamW on the early buckets hides the all-reduce of the last ones (the embeddings, whose
        # gradient only exists once backward has finished). AdamW is elementwise, so the result is unchanged.
        # Clipping needs the norm of the whole reduced gradient first.
        ranges = None
        if gs is not None:
            if self.clip_grad_norm <= 0.0 and gs.covers_all and gs.all_in_flight():
                ranges = gs.drain()
            else:
                gs.wait_all()
        grp = self.optimizer.param_groups[0]
        if len(self.optimizer.param_groups) != 1:
            raise PenrozError("fused AdamW expects the reference's single parameter group (mappers.py:267-274)")
        if grp.get("amsgrad", False) or grp.get("maximize", False):
            raise PenrozError("fused AdamW does not implement amsgrad / maximize")
        self._sync_step_count()
        self.step_count += 1
        mult = None
        if self.clip_grad_norm > 0.0:
            norm = self._buf("gnorm", (1,), torch.float32)
            mult = self._buf("gmult", (1,), torch.float32)
            ops.grad_norm(fl.g, self.clip_grad_norm, norm, mult)
        shadow = fl.shadow if self.mode == "amp" else None
        if track_update:
            fl.update_stats.zero_()
        for lo, hi in (ranges if ranges is not None else [(0, fl.p.numel())]):
            ops.adamw_flat(fl.p[lo:hi], fl.g[lo:hi], fl.m[lo:hi], fl.v[lo:hi], None if shadow is None else shadow[lo:hi],
                           lr=grp["lr"], betas=grp["betas"], eps=grp["eps"], weight_decay=grp["weight_decay"],
                           step=self.step_count, grad_mult=mult, zero_grad=clear_grads,
                           seg_of_group=fl.seg_of_group[lo // _ALIGN:hi // _ALIGN] if track_update else None,
                           seg_stats=fl.update_stats if track_update else None)
        self._grads_clean = bool(clear_grads)
        fl.step_vec.fill_(float(self.step_count))

    def update_ratios(self) -> list:
        """std(w - w_prev) / (std(w) + 1e-8) per parameter for the last optimizer_step(track_update=True), None for
        parameters that are not 2-D weights: the reference's `weight_upd_ratio` (neural_net_model.py:696-699; torch's
        std is the unbiased one) from the sums the AdamW pass accumulated -- one small device->host read."""
        vals = self.flat.update_stats.tolist()
        out = []
        for p, (sd, sdd, sw, sww) in zip(self.parameters, vals):
            n = p.numel()
            if p.ndim != 2 or n < 2:
                out.append(None)
                continue
            var_d = max(0.0, (sdd - sd * sd / n) / (n - 1))
            var_w = max(0.0, (sww - sw * sw / n) / (n - 1))
            out.append(var_d ** 0.5 / (var_w ** 0.5 + 1e-8))
        return out
