"""Torch-tensor front ends of the C-ABI entry points (include/penroz_b200.h).

torch is only the allocator and stream owner here: every function takes CUDA tensors, passes raw device pointers
and the current stream through ctypes, and returns tensors it allocated. No function falls back to a torch op.
"""
import math
import os

import torch

from . import _lib
from ._lib import (BF16, EPI_ACCUM_F32, EPI_DGELU, EPI_F32, EPI_GELU, EPI_RESID_BF16, EPI_RESID_F32, EPI_ROWDOT, EPI_STORE, F32,
                   FLAG_ACCUM, FLAG_TANH, dt, lib, ptr, stream)


def _chk(t: torch.Tensor, name: str):
    if not t.is_cuda:
        raise _lib.PenrozError(f"{name} must be a CUDA tensor (penroz_b200 has no CPU path)")
    if not t.is_contiguous() and t.dim() != 2:
        raise _lib.PenrozError(f"{name} must be contiguous")


def _mat(t: torch.Tensor, name: str):
    """2-D view with unit inner stride -> (tensor, ld)."""
    _chk(t, name)
    if t.dim() != 2 or t.stride(1) != 1:
        raise _lib.PenrozError(f"{name} must be a row-major 2-D tensor, got shape {tuple(t.shape)} strides {t.stride()}")
    return t, t.stride(0)


def gemm(a, b, *, a_layout=0, b_layout=0, epilogue=EPI_STORE, bias=None, out=None, out2=None, aux=None,
         tanh=False, accumulate=False, out_dtype=None, flags_extra=0, colsum_out=None, colsum_scratch=None,
         rowdot_out=None, seq_len=0):
    """D[M,N] = sum_k A(m,k) B(n,k) with a fused epilogue. a: [M,K] (layout 0) or [K,M] (layout 1); b likewise.

    bf16 operands -> tcgen05 kernel; fp32 operands -> FP32-pipe kernel. Returns out (and out2 for the GELU epilogue).
    colsum_out (EPI_DGELU only): fp32 [N] vector that receives += the column sums of `out` (the bias gradient of the
    Linear in front of the activation), produced by the same kernel on the bf16 path.
    rowdot_out / seq_len (EPI_ROWDOT only, bf16 shapes of the CTA-pair kernel): fp32 buffer that receives, per row and
    64-column group, the dot product of the output row with the aux row, laid out [M / seq_len][N / 64][seq_len].
    """
    a, lda = _mat(a, "a")
    b, ldb = _mat(b, "b")
    M, K = (a.shape[0], a.shape[1]) if a_layout == 0 else (a.shape[1], a.shape[0])
    N, Kb = (b.shape[0], b.shape[1]) if b_layout == 0 else (b.shape[1], b.shape[0])
    if K != Kb:
        raise _lib.PenrozError(f"gemm: K mismatch {K} vs {Kb}")
    if a.dtype != b.dtype:
        raise _lib.PenrozError(f"gemm: operand dtypes differ ({a.dtype} vs {b.dtype})")
    is_bf16 = a.dtype == torch.bfloat16
    if out is None:
        if out_dtype is None:
            out_dtype = torch.float32 if (epilogue in (EPI_RESID_F32, EPI_ACCUM_F32, EPI_F32) or not is_bf16) else torch.bfloat16
        out = torch.empty((M, N), device=a.device, dtype=out_dtype)
    if epilogue == EPI_GELU and out2 is None:
        out2 = torch.empty((M, N), device=a.device, dtype=out.dtype)
    flags = (FLAG_TANH if tanh else 0) | (FLAG_ACCUM if accumulate else 0) | flags_extra
    ldo = out.stride(0)
    ldo2 = out2.stride(0) if out2 is not None else 0
    if epilogue == EPI_ROWDOT:
        if rowdot_out is None or seq_len <= 0 or not is_bf16:
            raise _lib.PenrozError("gemm: EPI_ROWDOT needs bf16 operands, rowdot_out and seq_len")
        out2, ldo2 = rowdot_out, int(seq_len)
    ldaux = aux.stride(0) if aux is not None else 0
    if colsum_out is not None and epilogue != EPI_DGELU:
        raise _lib.PenrozError("gemm: colsum_out goes with EPI_DGELU")
    if is_bf16:
        lib().gemm_bf16(ptr(a), lda, a_layout, ptr(b), ldb, b_layout, M, N, K, epilogue, ptr(bias), ptr(out), ldo,
                        ptr(colsum_out) if colsum_out is not None else ptr(out2), ldo2, ptr(aux), ldaux, flags, stream())
    else:
        epi = EPI_STORE if epilogue == EPI_F32 else epilogue
        lib().gemm_f32(ptr(a), lda, a_layout, ptr(b), ldb, b_layout, M, N, K, epi, ptr(bias), ptr(out), ldo,
                       ptr(out2), ldo2, ptr(aux), ldaux, flags, stream())
        if colsum_out is not None:
            colsum(out, colsum_out, accumulate=True, scratch=colsum_scratch)
    return (out, out2) if epilogue == EPI_GELU else out


def kv_append_q8(qkv2d, codes, scales, block_table, B, S, H, D, t0, writeback=False, recip_scale=False):
    """Quantise the K and V columns of S new rows per sequence into the int8 page pool (+ per-token scales)."""
    lib().kv_append_q8(ptr(qkv2d), ptr(codes), ptr(scales), ptr(block_table), block_table.shape[1], B, S, H, D, t0,
                       dt(qkv2d), int(writeback), int(recip_scale), stream())


def attn_decode_q8(qkv2d, codes, scales, block_table, B, H, D, t_len, scratch=None, out=None):
    """Single-query attention over the int8 cache, de-quantising inside the kernel. Returns [B, H*D]."""
    if out is None:
        out = torch.empty((B, H * D), device=qkv2d.device, dtype=qkv2d.dtype)
    need = lib().attn_decode_q8_scratch_bytes(B, H, D)
    if scratch is None or scratch.numel() < need:
        scratch = torch.empty(need, device=qkv2d.device, dtype=torch.uint8)
    lib().attn_decode_q8(ptr(qkv2d), ptr(codes), ptr(scales), ptr(block_table), block_table.shape[1], ptr(out), B, H, D,
                         t_len, dt(qkv2d), ptr(scratch), scratch.numel(), stream())
    return out


def wstream_linear(x, W, bias=None, epilogue=0, aux=None, tanh=False, ln_w=None, ln_b=None, ln_eps=1e-5):
    """Weight-streaming Linear (decode, bf16, M <= 64 rows): epi(LN?(x) W^T + bias). epilogue 0 store, 1 GELU, 2 residual."""
    M, K = x.shape
    N = W.shape[0]
    out = torch.empty((M, N), device=x.device, dtype=x.dtype)
    lib().wstream_linear(ptr(x), x.stride(0), ptr(W), W.stride(0), ptr(bias), ptr(out), N, ptr(aux),
                         aux.stride(0) if aux is not None else 0, ptr(ln_w), ptr(ln_b), float(ln_eps), M, N, K,
                         int(epilogue), int(bool(tanh)), stream())
    return out


def widen_tokens(src_i32, dst_i64):
    """int32 token ids -> int64, on the device (dst may be any int64 tensor with as many elements)."""
    lib().widen_tokens(ptr(src_i32), ptr(dst_i64), src_i32.numel(), stream())
    return dst_i64


def embed_fwd(idx, wte, wpe, pos_offset=0, out_dtype=None, err_flag=None, out=None):
    idx = idx.contiguous()
    B, T = idx.shape
    V, C = wte.shape
    if out is None:
        out_dtype = out_dtype or wte.dtype
        out = torch.empty((B, T, C), device=idx.device, dtype=out_dtype)
    out_dtype = out.dtype
    lib().embed_fwd(ptr(idx), ptr(wte), ptr(wpe), ptr(out), B, T, C, V, pos_offset, dt(wte),
                    F32 if out_dtype == torch.float32 else BF16, ptr(err_flag), stream())
    return out


def embed_bwd(idx, dx, dwte, dwpe, pos_offset=0):
    idx = idx.contiguous()
    B, T = idx.shape
    V, C = dwte.shape
    lib().embed_bwd(ptr(idx), ptr(dx), ptr(dwte), ptr(dwpe), B, T, C, V, pos_offset, stream())


def layernorm_fwd(x, gamma, beta, eps=1e-5, out_dtype=None, save_stats=True, out=None, mean=None, rstd=None):
    C = x.shape[-1]
    rows = x.numel() // C
    if out is None:
        out_dtype = out_dtype or x.dtype
        out = torch.empty(x.shape, device=x.device, dtype=out_dtype)
    out_dtype = out.dtype
    if save_stats and mean is None:
        mean = torch.empty(rows, device=x.device, dtype=torch.float32)
        rstd = torch.empty(rows, device=x.device, dtype=torch.float32)
    lib().layernorm_fwd(ptr(x), ptr(gamma), ptr(beta), ptr(out), ptr(mean), ptr(rstd), rows, C, eps, dt(x), dt(gamma),
                        F32 if out_dtype == torch.float32 else BF16, stream())
    return out, mean, rstd


def layernorm_bwd_scratch(rows, C, device):
    n = lib().layernorm_bwd_scratch_bytes(rows, C)
    return torch.empty(n, device=device, dtype=torch.uint8)


def layernorm_bwd(dy, x, gamma, mean, rstd, *, add=None, dgamma=None, dbeta=None, dcol=None, want_bf16=False,
                  scratch=None, dx=None, dxb=None):
    """Returns (dx, dx_bf16|None); accumulates (+=) into dgamma/dbeta/dcol."""
    C = x.shape[-1]
    rows = x.numel() // C
    if dx is None:
        dx = torch.empty_like(x)
    if want_bf16 and dxb is None:
        dxb = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16)
    if not want_bf16:
        dxb = None
    if scratch is None:
        scratch = layernorm_bwd_scratch(rows, C, x.device)
    lib().layernorm_bwd(ptr(dy), ptr(x), ptr(gamma), ptr(mean), ptr(rstd), ptr(add), ptr(dx), ptr(dxb), ptr(dgamma),
                        ptr(dbeta), ptr(dcol), rows, C, dt(x), dt(gamma), dt(dy), ptr(scratch), scratch.numel(),
                        stream())
    return dx, dxb


def gelu_fwd(x, tanh=False):
    y = torch.empty_like(x)
    lib().gelu_fwd(ptr(x), ptr(y), x.numel(), int(tanh), dt(x), stream())
    return y


def gelu_bwd(dy, x, tanh=False):
    dx = torch.empty_like(x)
    lib().gelu_bwd(ptr(dy), ptr(x), ptr(dx), x.numel(), int(tanh), dt(x), stream())
    return dx


def colsum(x2d, out, accumulate=True, scratch=None):
    x2d, ld = _mat(x2d, "x")
    M, N = x2d.shape
    if scratch is None:
        scratch = torch.empty(lib().colsum_scratch_bytes(M, N), device=x2d.device, dtype=torch.uint8)
    lib().colsum(ptr(x2d), ld, M, N, ptr(out), int(accumulate), dt(x2d), ptr(scratch), scratch.numel(), stream())
    return out


def cast(src, dst):
    lib().cast(ptr(src), ptr(dst), src.numel(), dt(src), dt(dst), stream())
    return dst


def axpy_(y, x, alpha):
    """y += alpha * x (flat fp32 CUDA buffers; alpha a 0-dim fp32 CUDA tensor)."""
    lib().axpy_f32(ptr(y), ptr(x), ptr(alpha), y.numel(), stream())
    return y


def attn_fwd(qkv, H, want_lse=True, out=None, lse=None):
    """qkv [B,T,3C] -> (out [B,T,C], lse [B,H,T])."""
    B, T, C3 = qkv.shape
    C = C3 // 3
    D = C // H
    if out is None:
        out = torch.empty((B, T, C), device=qkv.device, dtype=qkv.dtype)
    if want_lse and lse is None:
        lse = torch.empty((B, H, T), device=qkv.device, dtype=torch.float32)
    if _use_tc_attention(qkv, D):
        lib().attn_fwd_tc(ptr(qkv), ptr(out), ptr(lse), B, T, H, D, stream())
    else:
        lib().attn_fwd_simt(ptr(qkv), ptr(out), ptr(lse), B, T, H, D, dt(qkv), stream())
    return out, lse


def _use_tc_attention(qkv, D) -> bool:
    """bf16 + head_dim 64 -> tcgen05 kernels; fp32 models and other head sizes -> exact FP32-pipe kernels."""
    return qkv.dtype == torch.bfloat16 and D == 64 and os.environ.get("PENROZ_ATTENTION", "tc") != "simt"


def attn_bwd_scratch_bytes(B, T, H, D, dtype) -> int:
    if dtype == torch.bfloat16 and D == 64:
        return max(lib().attn_bwd_tc_scratch_bytes(B, T, H, D), lib().attn_bwd_scratch_bytes(B, T, H))
    return lib().attn_bwd_scratch_bytes(B, T, H)


def rowdot_fusable(M, N, D, dtype) -> bool:
    """Can the GEMM that produces dO also produce the attention backward's row dot-products (EPI_ROWDOT)?"""
    return dtype == torch.bfloat16 and D == 64 and M >= 256 and N > 128 and N % 64 == 0


def attn_bwd(qkv, out, dout, lse, H, scratch=None, dqkv=None, dbias=None, delta_ready=False):
    """dbias: fp32 [3C] vector that receives += the column sums of dqkv (the QKV bias gradient); the tensor-core kernels
    produce it themselves, the fp32 path adds a column-sum pass."""
    B, T, C3 = qkv.shape
    D = (C3 // 3) // H
    if dqkv is None:
        dqkv = torch.empty_like(qkv)
    if scratch is None:
        scratch = torch.empty(attn_bwd_scratch_bytes(B, T, H, D, qkv.dtype), device=qkv.device, dtype=torch.uint8)
    if _use_tc_attention(qkv, D):
        lib().attn_bwd_tc(ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(dqkv), ptr(dbias), 1 if delta_ready else 0,
                          B, T, H, D, ptr(scratch), scratch.numel(), stream())
    else:
        lib().attn_bwd_simt(ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(dqkv), B, T, H, D, dt(qkv), ptr(scratch),
                            scratch.numel(), stream())
        if dbias is not None:
            ok = scratch.numel() >= lib().colsum_scratch_bytes(B * T, C3)
            colsum(dqkv.view(B * T, C3), dbias, accumulate=True, scratch=scratch if ok else None)
    return dqkv


def ce_count_valid(targets, n_valid, ignore_index=-100):
    lib().ce_count_valid(ptr(targets), targets.numel(), ignore_index, ptr(n_valid), stream())


def ce_rows(logits2d, targets, n_valid, loss_rows, *, V=None, grad_scale=1.0, write_grad=True, ignore_index=-100,
            err_flag=None):
    logits2d, ld = _mat(logits2d, "logits")
    M = logits2d.shape[0]
    V = V if V is not None else logits2d.shape[1]
    lib().ce_rows(ptr(logits2d), ld, ptr(targets), M, V, ignore_index, ptr(n_valid), float(grad_scale),
                  ptr(loss_rows), int(write_grad), dt(logits2d), ptr(err_flag), stream())


def ce_reduce(loss_rows, n_valid, loss_out, out_scale=1.0, accumulate=False):
    lib().ce_reduce(ptr(loss_rows), loss_rows.numel(), ptr(n_valid), float(out_scale), ptr(loss_out), int(accumulate),
                    stream())


def adamw_flat(p, g, m, v, shadow, *, lr, betas, eps, weight_decay, step, grad_mult=None, zero_grad=False,
               seg_of_group=None, seg_stats=None):
    """One AdamW pass over flat buffers. With seg_of_group / seg_stats (see penroz_adamw_flat_stats) the pass also
    accumulates sum(d), sum(d^2), sum(w), sum(w^2) of the update d = w - w_prev per tracked tensor."""
    if seg_of_group is not None:
        lib().adamw_flat_stats(ptr(p), ptr(g), ptr(m), ptr(v), ptr(shadow), p.numel(), float(lr), float(betas[0]),
                               float(betas[1]), float(eps), float(weight_decay), int(step), ptr(grad_mult),
                               int(zero_grad), ptr(seg_of_group), ptr(seg_stats), stream())
        return
    lib().adamw_flat(ptr(p), ptr(g), ptr(m), ptr(v), ptr(shadow), p.numel(), float(lr), float(betas[0]),
                     float(betas[1]), float(eps), float(weight_decay), int(step), ptr(grad_mult), int(zero_grad),
                     stream())


def grad_norm(g, max_norm, norm_out, mult_out, scratch=None):
    if scratch is None:
        scratch = torch.empty(lib().grad_norm_scratch_bytes(g.numel()), device=g.device, dtype=torch.uint8)
    lib().grad_norm(ptr(g), g.numel(), float(max_norm), ptr(norm_out), ptr(mult_out), ptr(scratch), scratch.numel(),
                    stream())


def tensor_stats(t, bins: int = 100, abs_threshold: float = 5.0):
    """mean / std / min / max / saturated fraction and histogram counts of a CUDA tensor with one host read.
    Returns (mean, std, min, max, saturated, lo, hi, counts[list of float])."""
    t = t.detach()
    if not t.is_contiguous():
        t = t.contiguous()
    out = torch.empty(7 + bins, device=t.device, dtype=torch.float64)
    scratch = torch.empty(lib().tensor_stats_scratch_bytes(bins), device=t.device, dtype=torch.uint8)
    lib().tensor_stats(ptr(t), t.numel(), dt(t), int(bins), float(abs_threshold), ptr(out), ptr(scratch),
                       scratch.numel(), stream())
    vals = out.tolist()
    return vals[0], vals[1], vals[2], vals[3], vals[4], vals[5], vals[6], vals[7:]


def kv_append(qkv2d, pool, block_table, B, S, H, D, t0):
    lib().kv_append(ptr(qkv2d), ptr(pool), ptr(block_table), block_table.shape[1], B, S, H, D, t0, dt(qkv2d), stream())


def attn_decode(qkv2d, pool, block_table, B, H, D, t_len, scratch=None):
    out = torch.empty((B, H * D), device=qkv2d.device, dtype=qkv2d.dtype)
    if scratch is None:
        scratch = torch.empty(lib().attn_decode_scratch_bytes(B, H, D, t_len), device=qkv2d.device, dtype=torch.uint8)
    lib().attn_decode(ptr(qkv2d), ptr(pool), ptr(block_table), block_table.shape[1], ptr(out), B, H, D, t_len,
                      dt(qkv2d), ptr(scratch), scratch.numel(), stream())
    return out


def skinny_linear(x, W, bias=None, *, epilogue=0, aux=None, tanh=False, out=None, out_f32=None):
    x, ldx = _mat(x, "x")
    W, ldw = _mat(W, "W")
    M, K = x.shape
    N = W.shape[0]
    if out is None and out_f32 is None:
        out = torch.empty((M, N), device=x.device, dtype=x.dtype)
    lib().skinny_linear(ptr(x), ldx, ptr(W), ldw, ptr(bias), ptr(out), out.stride(0) if out is not None else 0,
                        ptr(aux), aux.stride(0) if aux is not None else 0, ptr(out_f32),
                        out_f32.stride(0) if out_f32 is not None else 0, M, N, K, epilogue, int(tanh), dt(x), stream())
    return out if out is not None else out_f32


def argmax_rows(x2d, V=None):
    x2d, ld = _mat(x2d, "x")
    rows = x2d.shape[0]
    out = torch.empty(rows, device=x2d.device, dtype=torch.int64)
    lib().argmax_rows(ptr(x2d), ld, rows, V if V is not None else x2d.shape[1], ptr(out), dt(x2d), stream())
    return out
