"""Drop-in for the reference's kv_cache.py (same classes, methods, metrics and env flag; reference: kv_cache.py:14-206).

Two stores sit behind the one KVCache interface:

  * dense  — what generic callers (CausalSelfAttention.forward's cache branch, tests) use through
             append()/get(): per-layer [B,H,capacity,D] buffers that grow geometrically, so an append copies only
             the new tokens (the reference re-copies the whole cache with torch.cat on every token) and still hands
             back contiguous-in-S `[B,H,t,D]` views.
  * paged  — what the fused decode path (penroz/engine.py) uses: a page pool [n_pages][2][H][PAGE][D] per layer and
             one int32 block table [B][max_pages] shared by all layers; pages are handed out on demand from a free
             list (PAGE = 64 tokens), written by the kv_append kernel and read by the split-KV decode-attention
             kernel through the table. get() gathers a paged layer back to `[B,H,t,D]` for introspection.
"""
import logging
import os
import time
from dataclasses import dataclass

import torch
from torch import Tensor

log = logging.getLogger(__name__)

# Environment flag to enable the int8-compressed KV cache (default OFF)
TURBO_QUANT_ENABLED = os.environ.get("TURBO_QUANT_KV_CACHE", "0") == "1"


@dataclass
class KVCacheMetrics:
    """Lightweight metrics for KV cache usage."""
    num_appends: int = 0
    total_entries: int = 0
    memory_bytes: int = 0
    compressed_memory_bytes: int = 0
    compression_ratio: float = 1.0
    last_append_latency_ms: float = 0.0


class _GrowingSlab:
    """[B,H,capacity,D] buffer with amortised O(new tokens) append along dim 2."""

    def __init__(self):
        self.buf = None
        self.used = 0

    def push(self, piece: Tensor) -> Tensor:
        n = piece.shape[2]
        if self.buf is None:
            cap = max(16, 2 * n)
            self.buf = piece.new_empty(piece.shape[0], piece.shape[1], cap, piece.shape[3])
        elif (self.buf.shape[:2] != piece.shape[:2] or self.buf.shape[3] != piece.shape[3]
              or self.buf.dtype != piece.dtype or self.buf.device != piece.device):
            raise ValueError("KV cache append: shape/dtype/device differs from what is already cached")
        if self.used + n > self.buf.shape[2]:
            cap = max(2 * self.buf.shape[2], self.used + n)
            bigger = self.buf.new_empty(self.buf.shape[0], self.buf.shape[1], cap, self.buf.shape[3])
            bigger[:, :, :self.used] = self.buf[:, :, :self.used]
            self.buf = bigger
        self.buf[:, :, self.used:self.used + n] = piece
        self.used += n
        return self.view()

    def view(self):
        return None if self.buf is None else self.buf[:, :, :self.used]


class KVCache:
    """Per-layer key/value storage across generation steps: append, retrieve, clear."""

    def __init__(self, num_layers: int = 0):
        self._num_layers = num_layers
        self._k = [_GrowingSlab() for _ in range(num_layers)]
        self._v = [_GrowingSlab() for _ in range(num_layers)]
        self._metrics = KVCacheMetrics()
        self._paged = None

    @property
    def metrics(self) -> KVCacheMetrics:
        return self._metrics

    def _account(self, n_tokens: int, new_bytes: int, t0: float):
        m = self._metrics
        m.num_appends += 1
        m.total_entries += n_tokens
        m.memory_bytes += new_bytes
        m.compressed_memory_bytes = m.memory_bytes
        m.compression_ratio = 1.0
        m.last_append_latency_ms = (time.monotonic() - t0) * 1000

    def append(self, layer_idx: int, key: Tensor, value: Tensor) -> tuple[Tensor, Tensor]:
        """Append `[B,H,S_new,D]` keys/values for a layer; returns everything cached for it, `[B,H,t,D]`."""
        t0 = time.monotonic()
        full_k = self._k[layer_idx].push(key)
        full_v = self._v[layer_idx].push(value)
        self._account(key.shape[2], key.nelement() * key.element_size() + value.nelement() * value.element_size(), t0)
        return full_k, full_v

    def get(self, layer_idx: int) -> tuple[Tensor | None, Tensor | None]:
        if self._paged is not None and self._paged.length > 0:
            return self._paged.gather(layer_idx)
        return self._k[layer_idx].view(), self._v[layer_idx].view()

    def clear(self):
        for slab in self._k + self._v:
            slab.buf, slab.used = None, 0
        if self._paged is not None:
            self._paged.reset()
        self._metrics = KVCacheMetrics()

    def seq_len(self, layer_idx: int = 0) -> int:
        if self._paged is not None and self._paged.length > 0:
            return self._paged.length
        return self._k[layer_idx].used if self._num_layers else 0

    def log_metrics(self):
        m = self._metrics
        log.info(f"KVCache metrics: entries={m.total_entries}, memory={m.memory_bytes / 1024:.1f}KB, "
                 f"compression_ratio={m.compression_ratio:.2f}, last_append={m.last_append_latency_ms:.3f}ms")

    # ---- paged store, used by the fused decode path -------------------------------------------------------------
    _quantised = False       # TurboQuantKVCache: the paged store keeps int8 codes + scales

    def paged(self, B: int, H: int, D: int, dtype, device, capacity_hint: int = 0) -> "_PagedStore":
        if self._paged is None:
            self._paged = _PagedStore(self._num_layers, B, H, D, dtype, device, capacity_hint, quant=self._quantised)
        elif not self._paged.compatible(B, H, D, dtype, device):
            if self._paged.length > 0:
                raise ValueError("paged KV cache: batch/head shape changed while tokens are cached")
            self._paged = _PagedStore(self._num_layers, B, H, D, dtype, device, capacity_hint, quant=self._quantised)
        return self._paged

    # engine-facing façade (valid after paged(...) was called)
    def write_packed(self, layer_idx: int, qkv2d: Tensor, B: int, S: int, t0: int):
        t = time.monotonic()
        self._paged.write(layer_idx, qkv2d, B, S, t0)
        self._account(S, 2 * B * S * self._paged.H * self._paged.D * qkv2d.element_size(), t)

    def pool(self, layer_idx: int) -> Tensor:
        return self._paged.pools[layer_idx]

    def scale_pool(self, layer_idx: int):
        return self._paged.scale_pools[layer_idx] if self._paged.quant else None

    @property
    def block_table(self) -> Tensor:
        return self._paged.table_dev

    @property
    def scratch(self) -> Tensor:
        return self._paged.scratch

    def commit(self, n_tokens: int):
        self._paged.length += n_tokens

    def account_tokens(self, n_tokens: int, bytes_per_layer: int, n_layers: int):
        """Metrics for appends done inside the native decode step (one append per layer, as append() counts them)."""
        m = self._metrics
        m.num_appends += n_layers
        m.total_entries += n_tokens * n_layers
        m.memory_bytes += bytes_per_layer * n_layers
        m.compressed_memory_bytes = m.memory_bytes
        m.compression_ratio = 1.0


class _PagedStore:
    def __init__(self, num_layers, B, H, D, dtype, device, capacity_hint, quant=False):
        from penroz import _lib
        self.PAGE = _lib.lib().kv_page_size()
        self.L, self.B, self.H, self.D, self.dtype, self.device = num_layers, B, H, D, dtype, device
        self.quant = quant           # int8 codes + one scale per cached token and head (TurboQuantKVCache)
        self.max_pages = max(1, (max(capacity_hint, 1) + self.PAGE - 1) // self.PAGE)
        self.n_pages = B * self.max_pages
        self.pools = [torch.empty(self.n_pages, 2, H, self.PAGE, D, device=device,
                                  dtype=torch.int8 if quant else dtype) for _ in range(num_layers)]
        self.scale_pools = ([torch.empty(self.n_pages, 2, H, self.PAGE, device=device, dtype=dtype)
                             for _ in range(num_layers)] if quant else None)
        self.table_host = torch.full((B, self.max_pages), 0, dtype=torch.int32)
        self.table_dev = self.table_host.to(device)
        self.free = list(range(self.n_pages - 1, -1, -1))
        self.pages_of = [0] * B            # pages handed to each sequence so far
        self.length = 0
        self.scratch = torch.empty(_lib.lib().attn_decode_scratch_bytes(B, H, D, 1), device=device, dtype=torch.uint8)

    def compatible(self, B, H, D, dtype, device):
        return (B, H, D, dtype) == (self.B, self.H, self.D, self.dtype) and torch.device(device) == torch.device(self.device)

    def reset(self):
        self.free = list(range(self.n_pages - 1, -1, -1))
        self.pages_of = [0] * self.B
        self.length = 0

    def _grow(self, need_pages_per_seq: int):
        new_max = max(2 * self.max_pages, need_pages_per_seq)
        add = self.B * (new_max - self.max_pages)
        for i, p in enumerate(self.pools):
            extra = torch.empty(add, 2, self.H, self.PAGE, self.D, device=self.device, dtype=p.dtype)
            self.pools[i] = torch.cat([p, extra], dim=0)
            if self.quant:
                more = torch.empty(add, 2, self.H, self.PAGE, device=self.device, dtype=self.dtype)
                self.scale_pools[i] = torch.cat([self.scale_pools[i], more], dim=0)
        self.native_layers = None      # page pools moved: the native decode step's pointer table is stale
        self.free = list(range(self.n_pages + add - 1, self.n_pages - 1, -1)) + self.free
        self.n_pages += add
        table = torch.zeros((self.B, new_max), dtype=torch.int32)
        table[:, :self.max_pages] = self.table_host
        self.table_host = table
        self.max_pages = new_max

    def ensure_pages(self, upto_tokens: int):
        self._ensure(upto_tokens)

    def _ensure(self, upto_tokens: int):
        need = (upto_tokens + self.PAGE - 1) // self.PAGE
        if need > self.max_pages:
            self._grow(need)
        changed = False
        for b in range(self.B):
            while self.pages_of[b] < need:
                self.table_host[b, self.pages_of[b]] = self.free.pop()
                self.pages_of[b] += 1
                changed = True
        if changed or self.table_dev.shape != self.table_host.shape:
            self.table_dev = self.table_host.to(self.device)

    def write(self, layer_idx, qkv2d, B, S, t0):
        from penroz import ops
        if layer_idx == 0:
            self._ensure(t0 + S)
        if self.quant:
            # a multi-token append is the prompt pass: its own attention reads the de-quantised K/V (kv_cache.py:165-168)
            ops.kv_append_q8(qkv2d, self.pools[layer_idx], self.scale_pools[layer_idx], self.table_dev, B, S, self.H,
                             self.D, t0, writeback=(S > 1))
        else:
            ops.kv_append(qkv2d, self.pools[layer_idx], self.table_dev, B, S, self.H, self.D, t0)

    def gather(self, layer_idx):
        n = (self.length + self.PAGE - 1) // self.PAGE
        idx = self.table_dev[:, :n].long()                       # [B, n]
        pages = self.pools[layer_idx][idx]                       # [B, n, 2, H, PAGE, D]
        if self.quant:                                           # codes.float() * scale, as _dequantize returns them
            pages = pages.float() * self.scale_pools[layer_idx][idx].unsqueeze(-1)
        kv = pages.permute(2, 0, 3, 1, 4, 5).reshape(2, self.B, self.H, n * self.PAGE, self.D)[:, :, :, :self.length]
        return kv[0], kv[1]


class TurboQuantKVCache(KVCache):
    """KV cache that stores int8 codes with one scale per cached token (absmax / 127) and hands back de-quantised
    float tensors (reference: kv_cache.py:101-195). Enabled by TURBO_QUANT_KV_CACHE=1. Generic callers go through
    append()/get() (torch ops on growing slabs, the reference's arithmetic); the fused decode path keeps the codes in a
    paged int8 pool written by penroz_kv_append_q8 and read by penroz_attn_decode_q8."""
    _quantised = True

    def account_tokens(self, n_tokens: int, bytes_per_layer: int, n_layers: int):
        """Metrics for appends done inside the native decode step: what append() would have counted per layer."""
        st = self._paged
        es = bytes_per_layer // max(1, 2 * st.B * n_tokens * st.H * st.D)          # bytes per uncompressed element
        rows = 2 * st.B * n_tokens * st.H                                            # (k|v, sequence, token, head) rows
        m = self._metrics
        m.num_appends += n_layers
        m.total_entries += n_tokens * n_layers
        m.memory_bytes += bytes_per_layer * n_layers
        m.compressed_memory_bytes += rows * (st.D + es) * n_layers                  # int8 codes + one scale per row
        m.compression_ratio = m.memory_bytes / m.compressed_memory_bytes if m.compressed_memory_bytes > 0 else 1.0

    def write_packed(self, layer_idx: int, qkv2d: Tensor, B: int, S: int, t0: int):
        self._paged.write(layer_idx, qkv2d, B, S, t0)
        if layer_idx == self._num_layers - 1:
            self.account_tokens(S, 2 * B * S * self._paged.H * self._paged.D * qkv2d.element_size(), self._num_layers)

    def __init__(self, num_layers: int = 0):
        super().__init__(num_layers)
        self._sk = [_GrowingSlab() for _ in range(num_layers)]
        self._sv = [_GrowingSlab() for _ in range(num_layers)]

    @staticmethod
    def _quantize(tensor: Tensor) -> tuple[Tensor, Tensor]:
        scale = tensor.abs().amax(dim=-1, keepdim=True) / 127.0
        scale = torch.where(scale == 0, torch.ones_like(scale), scale)
        codes = torch.clamp(torch.round(tensor / scale), -128, 127).to(torch.int8)
        return codes, scale

    @staticmethod
    def _dequantize(quantized: Tensor, scale: Tensor) -> Tensor:
        return quantized.float() * scale

    def append(self, layer_idx: int, key: Tensor, value: Tensor) -> tuple[Tensor, Tensor]:
        t0 = time.monotonic()
        qk, sk = self._quantize(key)
        qv, sv = self._quantize(value)
        all_qk, all_sk = self._k[layer_idx].push(qk), self._sk[layer_idx].push(sk)
        all_qv, all_sv = self._v[layer_idx].push(qv), self._sv[layer_idx].push(sv)
        m = self._metrics
        m.num_appends += 1
        m.total_entries += key.shape[2]
        m.compressed_memory_bytes += sum(t.nelement() * t.element_size() for t in (qk, qv, sk, sv))
        m.memory_bytes += key.nelement() * key.element_size() + value.nelement() * value.element_size()
        m.compression_ratio = m.memory_bytes / m.compressed_memory_bytes if m.compressed_memory_bytes > 0 else 1.0
        m.last_append_latency_ms = (time.monotonic() - t0) * 1000
        return self._dequantize(all_qk, all_sk), self._dequantize(all_qv, all_sv)

    def clear(self):
        super().clear()
        for slab in self._sk + self._sv:
            slab.buf, slab.used = None, 0

    @property
    def _scales_k(self):
        return [s.view() for s in self._sk]

    @property
    def _scales_v(self):
        return [s.view() for s in self._sv]


def create_kv_cache(num_layers: int) -> KVCache:
    """KVCache, or TurboQuantKVCache when TURBO_QUANT_KV_CACHE=1."""
    if TURBO_QUANT_ENABLED:
        log.info("TurboQuant KV cache enabled (TURBO_QUANT_KV_CACHE=1)")
        return TurboQuantKVCache(num_layers)
    return KVCache(num_layers)
