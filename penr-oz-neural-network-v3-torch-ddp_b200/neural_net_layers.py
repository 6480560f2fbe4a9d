"""Drop-in for the reference's neural_net_layers.py (same class names, constructor kwargs, attributes and state-dict
keys — mappers.py:32-40 constructs these by name from JSON kwargs).

On the GPT-2 hot path these modules are only *structure*: NeuralNetworkModel recognises the tree they form and runs
it through the fused sm_100a engine (penroz/engine.py), never through the forward() methods below, whenever the
parameters live on a CUDA device. The forward() methods are the module-composition semantics of the reference
(neural_net_layers.py:59-225) for trees the engine does not cover — Gemma-family blocks (RMSNorm, GatedMLP, RoPE, GQA:
out of scope) and host-side (CPU) tensors. The one kernel-backed forward here is plain causal
attention on CUDA tensors, which goes through the penroz attention kernels with a hand-written backward.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F
from torch import Tensor


class _PackedCausalAttention(torch.autograd.Function):
    """softmax(QK^T/sqrt(D), causal) V on the packed [B,T,3C] buffer via the C-ABI kernels."""

    @staticmethod
    def forward(ctx, qkv: Tensor, num_heads: int):
        from penroz import ops
        qkv = qkv.contiguous()
        out, lse = ops.attn_fwd(qkv, num_heads)
        ctx.save_for_backward(qkv, out, lse)
        ctx.num_heads = num_heads
        return out

    @staticmethod
    def backward(ctx, dout: Tensor):
        from penroz import ops
        qkv, out, lse = ctx.saved_tensors
        return ops.attn_bwd(qkv, out, dout.contiguous().to(qkv.dtype), lse, ctx.num_heads), None


class CausalSelfAttention(nn.Module):
    """Multi-head causal attention over a fused QKV projection (reference: neural_net_layers.py:7-95)."""

    def __init__(self, num_heads: int, dropout: float = 0.0, num_kv_heads: int = None, rope_theta: float = None,
                 head_dim: int = None):
        super().__init__()
        self.num_heads = num_heads
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        self.dropout = dropout
        self.rope_theta = rope_theta
        self._kv_cache = None
        self._layer_idx = 0
        if rope_theta is not None and head_dim is not None:
            exponents = torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim
            self.register_buffer("inv_freq", 1.0 / (rope_theta ** exponents), persistent=False)

    def set_kv_cache(self, kv_cache, layer_idx: int):
        """Attach (or, with None, detach) a KV cache used for incremental decoding."""
        self._kv_cache, self._layer_idx = kv_cache, layer_idx

    @staticmethod
    def _rotate_half(x: Tensor) -> Tensor:
        half = x.shape[-1] // 2
        return torch.cat((-x[..., half:], x[..., :half]), dim=-1)

    def _apply_rope(self, q: Tensor, k: Tensor, head_dim: int, offset: int = 0) -> tuple[Tensor, Tensor]:
        inv_freq = getattr(self, "inv_freq", None)
        if inv_freq is None:
            steps = torch.arange(0, head_dim, 2, device=q.device, dtype=torch.float32) / head_dim
            inv_freq = 1.0 / (self.rope_theta ** steps)
        positions = torch.arange(offset, offset + q.shape[2], device=q.device, dtype=torch.float32)
        angles = torch.outer(positions, inv_freq.to(q.device)).repeat(1, 2)
        cos, sin = angles.cos()[None, None].to(q.dtype), angles.sin()[None, None].to(q.dtype)
        return q * cos + self._rotate_half(q) * sin, k * cos + self._rotate_half(k) * sin

    def _kernel_path(self, x: Tensor) -> bool:
        return (x.is_cuda and self.rope_theta is None and self.num_kv_heads == self.num_heads
                and self._kv_cache is None and not (self.training and self.dropout > 0.0)
                and x.dtype in (torch.float32, torch.bfloat16))

    def forward(self, query_key_value: Tensor) -> Tensor:
        if self._kernel_path(query_key_value):
            return _PackedCausalAttention.apply(query_key_value, self.num_heads)
        bsz, seq, width = query_key_value.shape
        hd = width // (self.num_heads + 2 * self.num_kv_heads)
        nq, nkv = self.num_heads * hd, self.num_kv_heads * hd
        q, k, v = torch.split(query_key_value, [nq, nkv, nkv], dim=-1)
        q = q.reshape(bsz, seq, self.num_heads, hd).permute(0, 2, 1, 3)
        k = k.reshape(bsz, seq, self.num_kv_heads, hd).permute(0, 2, 1, 3)
        v = v.reshape(bsz, seq, self.num_kv_heads, hd).permute(0, 2, 1, 3)
        if self.rope_theta is not None:
            start = self._kv_cache.seq_len() if self._kv_cache is not None else 0
            q, k = self._apply_rope(q, k, hd, start)
        if self.num_kv_heads < self.num_heads:
            groups = self.num_heads // self.num_kv_heads
            k = k.repeat_interleave(groups, dim=1)
            v = v.repeat_interleave(groups, dim=1)
        causal = True
        if self._kv_cache is not None:
            k, v = self._kv_cache.append(self._layer_idx, k, v)
            causal = seq > 1
        p_drop = self.dropout if self.training else 0.0
        ctx = F.scaled_dot_product_attention(q, k, v, attn_mask=None, dropout_p=p_drop, is_causal=causal)
        return ctx.permute(0, 2, 1, 3).reshape(bsz, seq, nq)


class PositionEmbedding(nn.Embedding):
    """Learned absolute positions; `position_offset` shifts the window during cached decoding."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._position_offset = 0

    @property
    def position_offset(self) -> int:
        return self._position_offset

    @position_offset.setter
    def position_offset(self, value: int):
        self._position_offset = value

    def forward(self, input_data: Tensor) -> Tensor:
        count = input_data.shape[1]
        first = self._position_offset
        where = torch.arange(first, first + count, dtype=torch.long, device=input_data.device)
        return super().forward(where)


class Summation(nn.Sequential):
    """Sum of every child applied to the same input (token + position embedding)."""

    def forward(self, input_data: Tensor) -> Tensor:
        total = None
        for child in self:
            term = child(input_data)
            total = term if total is None else total + term
        return total


class ResidualConnection(nn.Sequential):
    """x <- x + child(x) for each child in order."""

    def forward(self, forwarded: Tensor) -> Tensor:
        for child in self:
            forwarded = forwarded + child(forwarded)
        return forwarded


class SoftmaxOnLast(nn.Softmax):
    """Softmax over the last sequence position only: [B,S,V] -> [B,V]."""

    def forward(self, logits: Tensor) -> Tensor:
        return super().forward(logits[:, -1, :])


# ---- Gemma-family blocks: kept for mappers.py (constructed by name); no kernels here, out of scope ----

class RMSNorm(nn.Module):
    def __init__(self, normalized_shape: int, eps: float = 1e-6):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(normalized_shape))
        self.eps = eps

    def forward(self, x: Tensor) -> Tensor:
        xf = x.float()
        inv = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.eps)
        return (xf * inv).to(x.dtype) * self.weight


class GatedMLP(nn.Module):
    def __init__(self, in_features: int, intermediate_size: int, bias: bool = False,
                 activation: str = "gelu_pytorch_tanh"):
        super().__init__()
        self.gate_proj = nn.Linear(in_features, intermediate_size, bias=bias)
        self.up_proj = nn.Linear(in_features, intermediate_size, bias=bias)
        self.down_proj = nn.Linear(intermediate_size, in_features, bias=bias)
        if activation in ("silu", "swish"):
            self.act = nn.SiLU()
        else:
            self.act = nn.GELU(approximate="tanh") if activation == "gelu_pytorch_tanh" else nn.GELU()

    def forward(self, x: Tensor) -> Tensor:
        return self.down_proj(self.act(self.gate_proj(x)) * self.up_proj(x))


class ScaledEmbedding(nn.Embedding):
    def __init__(self, num_embeddings: int, embedding_dim: int, scale: float = 1.0, **kwargs):
        super().__init__(num_embeddings, embedding_dim, **kwargs)
        self.scale = scale

    def forward(self, input_data: Tensor) -> Tensor:
        return super().forward(input_data) * self.scale


class TransformerBlock(nn.Module):
    """attention branch + MLP branch with optional post-norms, applied to the branch (Gemma 2) or to the sum
    (Gemma 3+, `post_norm_on_residual`)."""

    def __init__(self, attn_block: nn.Module, mlp_block: nn.Module, post_attn_norm: nn.Module = None,
                 post_mlp_norm: nn.Module = None, post_norm_on_residual: bool = True):
        super().__init__()
        self.attn_block = attn_block
        self.mlp_block = mlp_block
        self.post_attn_norm = post_attn_norm
        self.post_mlp_norm = post_mlp_norm
        self.post_norm_on_residual = post_norm_on_residual

    def _branch(self, x: Tensor, block: nn.Module, norm) -> Tensor:
        y = block(x)
        if norm is None:
            return x + y
        return norm(x + y) if self.post_norm_on_residual else x + norm(y)

    def forward(self, x: Tensor) -> Tensor:
        h = self._branch(x, self.attn_block, self.post_attn_norm)
        return self._branch(h, self.mlp_block, self.post_mlp_norm)
