"""Per-kernel parity on the GPU: every C-ABI entry point against a plain PyTorch fp32 statement of the same op.

Tolerances are written beside each check. bf16 kernels are compared with an fp32 reference evaluated on the same
bf16-rounded inputs; the bound is a few bf16 ulps of the output scale (2^-8 relative) unless stated.
"""
import math

import numpy as np

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from penroz import _lib
    _lib.require_device()
    return torch.device("cuda:0")


def rel_err(a, b):
    a = a.float()
    b = b.float()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


# ------------------------------------------------------------------ GEMM (tcgen05)
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 768), (300, 200, 136), (1024, 2304, 768),
                                    (77, 50257, 768), (4096, 768, 3072)])
@pytest.mark.parametrize("al,bl", [(0, 0), (0, 1), (1, 1), (1, 0)])
@pytest.mark.parametrize("single_cta", [0, 16])
def test_gemm_bf16_layouts(dev, M, N, K, al, bl, single_cta):
    """single_cta = 16 forces the one-CTA kernel; 0 lets shapes with M, N >= 256 take the CTA-pair (cta_group::2) kernel."""
    from penroz import ops
    if N == 50257 and bl == 1:
        pytest.skip("MN-major operand needs ld % 8 == 0")
    if al == 1 and M % 8:
        pytest.skip("MN-major operand needs ld % 8 == 0")
    torch.manual_seed(M + N + K)
    A = torch.randn(M, K, device=dev).bfloat16()
    B = torch.randn(N, K, device=dev).bfloat16()
    ref = A.float() @ B.float().t()
    a_in = A if al == 0 else A.t().contiguous()
    b_in = B if bl == 0 else B.t().contiguous()
    out = ops.gemm(a_in, b_in, a_layout=al, b_layout=bl, epilogue=ops.EPI_F32, flags_extra=single_cta)
    torch.cuda.synchronize()
    # fp32 accumulation of exact bf16 products: only summation-order noise
    assert rel_err(out, ref) < 2e-5, f"fp32-out rel err {rel_err(out, ref)}"
    out_b = ops.gemm(a_in, b_in, a_layout=al, b_layout=bl, epilogue=ops.EPI_STORE, flags_extra=single_cta)
    assert torch.equal(out_b, out.bfloat16()) or rel_err(out_b, ref) < 4e-3


@pytest.mark.parametrize("bn_flag", [4, 8])
def test_gemm_bf16_epilogues(dev, bn_flag):
    from penroz import ops
    torch.manual_seed(0)
    M, N, K = 512, 768, 256
    A = torch.randn(M, K, device=dev).bfloat16()
    W = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    bias = torch.randn(N, device=dev).bfloat16()
    acc = A.float() @ W.float().t() + bias.float()
    pre_ref = acc.bfloat16()
    out = ops.gemm(A, W, epilogue=ops.EPI_STORE, bias=bias, flags_extra=bn_flag)
    assert rel_err(out, acc) < 4e-3
    for tanh in (False, True):
        pre, post = ops.gemm(A, W, epilogue=ops.EPI_GELU, bias=bias, tanh=tanh, flags_extra=bn_flag)
        g_ref = F.gelu(pre.float(), approximate="tanh" if tanh else "none")
        assert rel_err(pre, acc) < 4e-3
        assert rel_err(post, g_ref) < 4e-3
        # DGELU: out = bf16(acc) * gelu'(h)
        h = pre
        hh = h.float().requires_grad_(True)
        F.gelu(hh, approximate="tanh" if tanh else "none").sum().backward()
        d = ops.gemm(A, W, epilogue=ops.EPI_DGELU, aux=h, tanh=tanh, flags_extra=bn_flag)
        d_ref = (A.float() @ W.float().t()).bfloat16().float() * hh.grad
        assert rel_err(d, d_ref) < 6e-3
        # the same with the bias-gradient side output: += column sums of the output as written
        cs = torch.full((N,), 3.0, device=dev)
        d2 = ops.gemm(A, W, epilogue=ops.EPI_DGELU, aux=h, tanh=tanh, flags_extra=bn_flag, colsum_out=cs)
        assert torch.equal(d2, d)
        assert rel_err(cs - 3.0, d.float().sum(0)) < 1e-5
    resid = torch.randn(M, N, device=dev)
    o = ops.gemm(A, W, epilogue=ops.EPI_RESID_F32, bias=bias, aux=resid, flags_extra=bn_flag)
    assert (o - (resid + pre_ref.float())).abs().max().item() < 0.05
    assert rel_err(o, resid + acc) < 4e-3
    rb = resid.bfloat16()
    ob = ops.gemm(A, W, epilogue=ops.EPI_RESID_BF16, bias=bias, aux=rb, flags_extra=bn_flag)
    assert rel_err(ob, rb.float() + acc) < 8e-3
    # accumulate (weight-gradient form, split-K engaged: few tiles, long K)
    dY = torch.randn(4096, 256, device=dev).bfloat16()
    X = torch.randn(4096, 384, device=dev).bfloat16()
    ref = dY.float().t() @ X.float()
    g0 = torch.zeros(256, 384, device=dev)
    ops.gemm(dY, X, a_layout=1, b_layout=1, epilogue=ops.EPI_ACCUM_F32, out=g0, flags_extra=bn_flag)
    assert rel_err(g0, ref) < 2e-5
    ops.gemm(dY, X, a_layout=1, b_layout=1, epilogue=ops.EPI_ACCUM_F32, out=g0, accumulate=True, flags_extra=bn_flag)
    assert rel_err(g0, 2 * ref) < 2e-5


def test_gemm_rowdot_epilogue_feeds_attention_backward(dev):
    """EPI_ROWDOT: the GEMM that produces dO also leaves delta[b,h,t] = sum_d dO*O where the attention backward reads it."""
    from penroz import ops
    torch.manual_seed(11)
    B, T, H, D = 2, 256, 4, 64
    C, M = H * D, B * T
    dy = torch.randn(M, C, device=dev).bfloat16()
    W = (torch.randn(C, C, device=dev) * 0.05).bfloat16()          # [N_out, K_in] as stored: dO = dy . W (MN-major B)
    qkv = torch.randn(B, T, 3 * C, device=dev).bfloat16()
    O, lse = ops.attn_fwd(qkv, H)
    scratch = torch.zeros(ops.attn_bwd_scratch_bytes(B, T, H, D, qkv.dtype), device=dev, dtype=torch.uint8)
    dO_plain = ops.gemm(dy, W, b_layout=1)
    dO = ops.gemm(dy, W, b_layout=1, epilogue=ops.EPI_ROWDOT, aux=O.view(M, C), rowdot_out=scratch, seq_len=T)
    assert torch.equal(dO, dO_plain)
    delta = scratch[: B * H * T * 4].view(torch.float32).view(B, H, T)
    ref = (dO.float() * O.view(M, C).float()).view(B, T, H, D).sum(-1).permute(0, 2, 1)
    assert rel_err(delta, ref) < 1e-5
    d1 = ops.attn_bwd(qkv, O, dO.view(B, T, C), lse, H, scratch=scratch, delta_ready=True)
    d2 = ops.attn_bwd(qkv, O, dO.view(B, T, C), lse, H)
    assert rel_err(d1, d2) < 2e-3    # dQ is reduced with floating-point atomics; delta differs by summation order only


def test_gemm_f32(dev):
    from penroz import ops
    torch.manual_seed(1)
    for (M, N, K, al, bl) in [(100, 70, 33, 0, 0), (256, 128, 64, 1, 0), (65, 193, 128, 0, 1), (128, 65, 128, 1, 1)]:
        A = torch.randn(M, K, device=dev)
        B = torch.randn(N, K, device=dev)
        bias = torch.randn(N, device=dev)
        ref = (A.double() @ B.double().t() + bias.double()).float()
        a_in = A if al == 0 else A.t().contiguous()
        b_in = B if bl == 0 else B.t().contiguous()
        out = ops.gemm(a_in, b_in, a_layout=al, b_layout=bl, bias=bias)
        assert rel_err(out, ref) < 1e-5
    A = torch.randn(64, 128, device=dev); W = torch.randn(96, 128, device=dev) * 0.1; b = torch.randn(96, device=dev)
    pre, post = ops.gemm(A, W, epilogue=ops.EPI_GELU, bias=b)
    assert rel_err(post, F.gelu(A @ W.t() + b)) < 1e-5
    r = torch.randn(64, 96, device=dev)
    o = ops.gemm(A, W, epilogue=ops.EPI_RESID_F32, bias=b, aux=r)
    assert rel_err(o, r + A @ W.t() + b) < 1e-5


# ------------------------------------------------------------------ row-wise kernels
@pytest.mark.parametrize("C", [128, 768, 1024, 1600])
def test_layernorm_fwd_bwd(dev, C):
    from penroz import ops
    torch.manual_seed(C)
    rows = 1000
    x = torch.randn(rows, C, device=dev) * 2 + 0.5
    g = torch.randn(C, device=dev)
    b = torch.randn(C, device=dev)
    ref = F.layer_norm(x, (C,), g, b, 1e-5)
    out, mean, rstd = ops.layernorm_fwd(x, g, b, out_dtype=torch.float32)
    assert rel_err(out, ref) < 1e-5   # fp32 in/out: summation order only
    outb, _, _ = ops.layernorm_fwd(x, g, b, out_dtype=torch.bfloat16)
    assert rel_err(outb, ref) < 4e-3  # one bf16 rounding of the output
    out_nb, _, _ = ops.layernorm_fwd(x, g, None, out_dtype=torch.float32)
    assert rel_err(out_nb, F.layer_norm(x, (C,), g, None, 1e-5)) < 1e-5
    # backward, fp32 dy, with residual add and fused column sums
    xr = x.clone().requires_grad_(True); gr = g.clone().requires_grad_(True); br = b.clone().requires_grad_(True)
    dy = torch.randn(rows, C, device=dev)
    add = torch.randn(rows, C, device=dev)
    F.layer_norm(xr, (C,), gr, br, 1e-5).backward(dy)
    dgam = torch.zeros(C, device=dev); dbet = torch.zeros(C, device=dev); dcol = torch.zeros(C, device=dev)
    dx, dxb = ops.layernorm_bwd(dy, x, g, mean, rstd, add=add, dgamma=dgam, dbeta=dbet, dcol=dcol, want_bf16=False)
    assert rel_err(dx, xr.grad + add) < 2e-5
    assert rel_err(dgam, gr.grad) < 2e-5
    assert rel_err(dbet, br.grad) < 2e-5
    assert rel_err(dcol, (xr.grad + add).sum(0)) < 1e-4
    # bf16 dy (autocast), bf16 copy of dx, accumulate semantics (+=)
    dyb = dy.bfloat16()
    xr.grad = None; gr.grad = None; br.grad = None
    F.layer_norm(xr, (C,), gr, br, 1e-5).backward(dyb.float())
    dx2, dxb2 = ops.layernorm_bwd(dyb, x, g, mean, rstd, add=None, dgamma=dgam, dbeta=dbet, dcol=None, want_bf16=True)
    assert rel_err(dx2, xr.grad) < 2e-5
    assert torch.equal(dxb2, dx2.bfloat16())
    assert rel_err(dgam, gr.grad * 2) < 1e-2  # second accumulation used the bf16-rounded dy


@pytest.mark.parametrize("rows,C", [(1001, 768), (7, 768), (1, 1024), (513, 100), (300, 2048), (4097, 1280)])
@pytest.mark.parametrize("dy_dtype", [torch.float32, torch.bfloat16])
def test_layernorm_bwd_shapes(dev, rows, C, dy_dtype):
    """Odd row counts (a last stage with one row), one- and two-vector rows per thread (C <= 1024 / > 1024), a row
    length the bulk copies cannot take (C % 8 != 0 -> the two-kernel form), with and without the residual gradient."""
    from penroz import ops
    torch.manual_seed(rows + C)
    x = torch.randn(rows, C, device=dev) * 1.5 - 0.3
    g = torch.randn(C, device=dev); b = torch.randn(C, device=dev)
    _, mean, rstd = ops.layernorm_fwd(x, g, b, out_dtype=torch.float32)
    dy = torch.randn(rows, C, device=dev).to(dy_dtype)
    add = torch.randn(rows, C, device=dev)
    xr = x.clone().requires_grad_(True); gr = g.clone().requires_grad_(True); br = b.clone().requires_grad_(True)
    F.layer_norm(xr, (C,), gr, br, 1e-5).backward(dy.float())
    for use_add in (True, False):
        dgam = torch.zeros(C, device=dev); dbet = torch.zeros(C, device=dev); dcol = torch.zeros(C, device=dev)
        want_bf16 = dy_dtype == torch.bfloat16
        dx, dxb = ops.layernorm_bwd(dy, x, g, mean, rstd, add=add if use_add else None, dgamma=dgam, dbeta=dbet,
                                    dcol=dcol, want_bf16=want_bf16)
        ref_dx = xr.grad + (add if use_add else 0)
        assert rel_err(dx, ref_dx) < 3e-5
        assert rel_err(dgam, gr.grad) < 3e-5 and rel_err(dbet, br.grad) < 3e-5
        seen = dx.bfloat16().float() if want_bf16 else dx
        assert rel_err(dcol, seen.sum(0)) < 1e-4
        if want_bf16:
            assert torch.equal(dxb, dx.bfloat16())


def test_embedding(dev):
    from penroz import ops
    torch.manual_seed(3)
    V, T, C, B = 65, 64, 128, 16
    wte = torch.randn(V, C, device=dev); wpe = torch.randn(T, C, device=dev)
    idx = torch.randint(0, V, (B, T), device=dev)
    out = ops.embed_fwd(idx, wte, wpe)
    assert torch.equal(out, wte[idx] + wpe[None, :, :])
    out5 = ops.embed_fwd(idx[:, :7], wte, wpe, pos_offset=5)
    assert torch.equal(out5, wte[idx[:, :7]] + wpe[None, 5:12, :])
    ob = ops.embed_fwd(idx, wte.bfloat16(), wpe.bfloat16())
    assert torch.equal(ob, (wte.bfloat16()[idx] + wpe.bfloat16()[None]))
    dx = torch.randn(B, T, C, device=dev)
    dwte = torch.zeros_like(wte); dwpe = torch.zeros_like(wpe)
    ops.embed_bwd(idx, dx, dwte, dwpe)
    ref_wte = torch.zeros_like(wte).index_add_(0, idx.view(-1), dx.view(-1, C))
    assert rel_err(dwte, ref_wte) < 1e-5  # atomics: order noise only
    assert rel_err(dwpe, dx.sum(0)) < 1e-6
    flag = torch.zeros(1, device=dev, dtype=torch.int32)
    bad = idx.clone(); bad[0, 0] = V + 3
    ops.embed_fwd(bad, wte, wpe, err_flag=flag)
    assert flag.item() == 1


@pytest.mark.parametrize("tanh", [False, True])
def test_gelu_and_colsum_and_cast(dev, tanh):
    from penroz import ops
    torch.manual_seed(4)
    x = torch.randn(1000, 515, device=dev) * 3
    appr = "tanh" if tanh else "none"
    assert rel_err(ops.gelu_fwd(x, tanh), F.gelu(x, approximate=appr)) < 2e-6
    xr = x.clone().requires_grad_(True)
    dy = torch.randn_like(x)
    F.gelu(xr, approximate=appr).backward(dy)
    assert rel_err(ops.gelu_bwd(dy, x, tanh), xr.grad) < 5e-6
    xb = x.bfloat16()
    assert rel_err(ops.gelu_fwd(xb, tanh), F.gelu(xb.float(), approximate=appr)) < 4e-3
    y = torch.randn(4099, 772, device=dev)
    out = torch.ones(772, device=dev)
    ops.colsum(y, out, accumulate=True)
    assert rel_err(out, 1 + y.sum(0)) < 1e-5
    outb = torch.zeros(772, device=dev)
    ops.colsum(y.bfloat16(), outb, accumulate=False)
    assert rel_err(outb, y.bfloat16().float().sum(0)) < 1e-5
    dst = torch.empty(x.numel(), device=dev, dtype=torch.bfloat16)
    ops.cast(x.view(-1), dst)
    assert torch.equal(dst, x.view(-1).bfloat16())


# ------------------------------------------------------------------ attention (exact path)
@pytest.mark.parametrize("B,T,H,D,dtype", [(2, 64, 4, 32, torch.float32), (2, 200, 3, 64, torch.float32),
                                            (1, 1024, 2, 64, torch.bfloat16), (2, 96, 2, 128, torch.float32)])
def test_attention_simt(dev, B, T, H, D, dtype, monkeypatch):
    from penroz import ops
    monkeypatch.setenv("PENROZ_ATTENTION", "simt")   # the exact FP32-pipe kernels, also for bf16 storage
    torch.manual_seed(T)
    C = H * D
    qkv = (torch.randn(B, T, 3 * C, device=dev)).to(dtype)
    q, k, v = [t.view(B, T, H, D).transpose(1, 2).float().requires_grad_(True) for t in qkv.split(C, dim=2)]
    ref = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    out, lse = ops.attn_fwd(qkv, H)
    ref_out = ref.transpose(1, 2).reshape(B, T, C)
    tol = 1e-5 if dtype == torch.float32 else 8e-3
    assert rel_err(out, ref_out) < tol
    s = (q @ k.transpose(-1, -2)) / math.sqrt(D)
    mask = torch.ones(T, T, device=dev, dtype=torch.bool).tril()
    lse_ref = torch.logsumexp(s.masked_fill(~mask, float("-inf")), dim=-1)
    assert rel_err(lse, lse_ref) < 1e-5
    dout = torch.randn(B, T, C, device=dev).to(dtype)
    ref.backward(dout.float().view(B, T, H, D).transpose(1, 2))
    db = torch.zeros(3 * C, device=dev)
    dqkv = ops.attn_bwd(qkv, out, dout, lse, H, dbias=db)
    ref_d = torch.cat([t.grad.transpose(1, 2).reshape(B, T, C) for t in (q, k, v)], dim=2)
    btol = 2e-5 if dtype == torch.float32 else 1.5e-2
    assert rel_err(dqkv, ref_d) < btol
    assert rel_err(db, dqkv.float().sum((0, 1))) < 1e-5


@pytest.mark.parametrize("B,T,H", [(1, 128, 1), (2, 256, 3), (1, 200, 2), (2, 1024, 4), (1, 1000, 2)])
def test_attention_tensor_core(dev, B, T, H, monkeypatch):
    """tcgen05 forward/backward (bf16, head_dim 64) against SDPA evaluated in fp32 on the same bf16 inputs.
    Tolerance: P, dS and the outputs are rounded to bf16 inside the kernel -> a few bf16 ulps (2^-8) of the scale."""
    from penroz import ops
    torch.manual_seed(T * 7 + H)
    D, C = 64, H * 64
    qkv = torch.randn(B, T, 3 * C, device=dev).bfloat16()
    q, k, v = [t.view(B, T, H, D).transpose(1, 2).float().requires_grad_(True) for t in qkv.split(C, dim=2)]
    ref = F.scaled_dot_product_attention(q, k, v, is_causal=True)
    out, lse = ops.attn_fwd(qkv, H)
    assert rel_err(out, ref.transpose(1, 2).reshape(B, T, C)) < 1e-2
    monkeypatch.setenv("PENROZ_ATTENTION", "simt")
    out_s, lse_s = ops.attn_fwd(qkv, H)
    monkeypatch.delenv("PENROZ_ATTENTION")
    assert (lse - lse_s).abs().max().item() < 5e-3
    dout = torch.randn(B, T, C, device=dev).bfloat16()
    ref.backward(dout.float().view(B, T, H, D).transpose(1, 2))
    dqkv = ops.attn_bwd(qkv, out, dout, lse, H)
    ref_d = torch.cat([t.grad.transpose(1, 2).reshape(B, T, C) for t in (q, k, v)], dim=2)
    for name, sl in (("dq", slice(0, C)), ("dk", slice(C, 2 * C)), ("dv", slice(2 * C, 3 * C))):
        assert rel_err(dqkv[..., sl], ref_d[..., sl]) < 2e-2, name
    # the QKV bias gradient as a side output: += column sums of dqkv as written
    db = torch.full((3 * C,), -2.0, device=dev)
    dqkv2 = ops.attn_bwd(qkv, out, dout, lse, H, dbias=db)
    assert rel_err(dqkv2, dqkv) < 1e-6   # dQ is reduced with floating-point atomics: not bit-stable run to run
    assert rel_err(db + 2.0, dqkv2.float().sum((0, 1))) < 1e-5


@pytest.mark.parametrize("mode", ["lazy", "classic"])
def test_attention_forward_with_growing_scores(dev, mode, monkeypatch):
    """Scores that keep growing along the sequence, so the running row maximum jumps by far more than the 2^8 the
    forward kernel tolerates before it rescales its TMEM accumulator -- and a plain random case where it never does.
    Both schedules (output accumulated in TMEM with lazy rescale / read back every tile) against SDPA in fp32."""
    import subprocess, sys, os, textwrap
    code = textwrap.dedent("""
        import sys, torch, torch.nn.functional as F
        sys.path.insert(0, "penr-oz-neural-network-v3-torch-ddp_b200")
        from penroz import ops
        torch.manual_seed(3)
        B, T, H, D = 2, 1024, 3, 64
        C = H * D
        for ramp in (0.0, 1.0):
            qkv = torch.randn(B, T, 3 * C, device="cuda")
            grow = 1.0 + ramp * 6.0 * torch.arange(T, device="cuda").float().view(1, T, 1) / T       # keys get up to 7x larger
            qkv[:, :, C:2 * C] *= grow
            qkv = qkv.bfloat16()
            q, k, v = [t.view(B, T, H, D).transpose(1, 2).float() for t in qkv.split(C, dim=2)]
            ref = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, C)
            out, lse = ops.attn_fwd(qkv, H)
            err = ((out.float() - ref).norm() / ref.norm()).item()
            ref_lse = torch.logsumexp((q @ k.transpose(-1, -2) / 8.0).masked_fill(
                torch.ones(T, T, device="cuda").triu(1).bool(), float("-inf")), dim=-1)
            lerr = (lse - ref_lse).abs().max().item()
            print("ramp", ramp, "rel err", err, "lse err", lerr)
            assert err < 1e-2 and lerr < 2e-2, (ramp, err, lerr)
        print("OK")
    """)
    env = dict(os.environ)
    if mode == "classic":
        env["PENROZ_ATTN_FWD"] = "classic"
    else:
        env.pop("PENROZ_ATTN_FWD", None)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "OK" in r.stdout, r.stdout[-1500:] + r.stderr[-3000:]


# ------------------------------------------------------------------ cross entropy, AdamW, grad norm
@pytest.mark.parametrize("V,ld,dtype", [(65, 65, torch.float32), (50257, 50304, torch.bfloat16), (50257, 50304, torch.float32),
                                         (1000, 1000, torch.bfloat16)])
def test_cross_entropy(dev, V, ld, dtype):
    from penroz import ops
    torch.manual_seed(V)
    M = 257
    buf = torch.zeros(M, ld, device=dev, dtype=dtype)
    buf[:, :V] = (torch.randn(M, V, device=dev) * 3).to(dtype)
    tgt = torch.randint(0, V, (M,), device=dev)
    tgt[5] = -100; tgt[77] = -100
    lg = buf[:, :V].float().requires_grad_(True)
    ref = F.cross_entropy(lg, tgt)
    ref.backward()
    nv = torch.zeros(1, device=dev, dtype=torch.int32)
    rows = torch.empty(M, device=dev)
    loss = torch.zeros(1, device=dev)
    ops.ce_count_valid(tgt, nv)
    assert nv.item() == M - 2
    work = buf.clone()
    ops.ce_rows(work, tgt, nv, rows, V=V, grad_scale=1.0, write_grad=True)
    ops.ce_reduce(rows, nv, loss)
    assert abs(loss.item() - ref.item()) < 2e-5 * max(1.0, abs(ref.item()))
    g = work[:, :V].float()
    tol = 1e-5 if dtype == torch.float32 else 8e-3
    assert rel_err(g, lg.grad) < tol
    assert g[5].abs().max().item() == 0.0
    # forward only leaves the logits untouched
    work2 = buf.clone()
    ops.ce_rows(work2, tgt, nv, rows, V=V, write_grad=False)
    assert torch.equal(work2, buf)


def test_adamw_and_grad_norm(dev):
    from penroz import ops
    torch.manual_seed(5)
    n = 1_000_003
    p = torch.randn(n, device=dev); g = torch.randn(n, device=dev) * 0.1
    ref_p = p.clone().requires_grad_(True)
    opt = torch.optim.AdamW([ref_p], lr=6e-4, betas=(0.9, 0.95), eps=1e-8, foreach=False, fused=False)
    m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
    shadow = torch.empty(n, device=dev, dtype=torch.bfloat16)
    for step in range(1, 4):
        gg = g * step
        ref_p.grad = gg.clone()
        opt.step()
        gbuf = gg.clone()
        ops.adamw_flat(p, gbuf, m, v, shadow, lr=6e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01, step=step,
                       zero_grad=True)
        assert gbuf.abs().max().item() == 0.0
    st = opt.state[ref_p]
    # same operation order as torch's single-tensor AdamW: agreement to a few fp32 ulps
    assert (p - ref_p.detach()).abs().max().item() < 2e-6
    assert rel_err(m, st["exp_avg"]) < 1e-6
    assert rel_err(v, st["exp_avg_sq"]) < 1e-6
    assert torch.equal(shadow, p.bfloat16())
    norm = torch.zeros(1, device=dev); mult = torch.zeros(1, device=dev)
    ops.grad_norm(g, 1.0, norm, mult)
    assert abs(norm.item() - g.double().norm().item()) < 1e-3 * g.norm().item()
    assert abs(mult.item() - min(1.0, 1.0 / (g.norm().item() + 1e-6))) < 1e-5


def test_adamw_pass_measures_its_own_update(dev):
    """penroz_adamw_flat_stats: the same parameters / moments / shadow as penroz_adamw_flat bit for bit, plus, per tracked
    tensor, the sums behind the reference's weight_upd_ratio = std(w - w_prev) / (std(w) + 1e-8)
    (neural_net_model.py:696-699), checked against torch on cloned weights -- also when the pass is split into
    ranges (the bucket-wise update under DDP) and when a tracked tensor is smaller than a sweep tile."""
    from penroz import ops
    torch.manual_seed(9)
    sizes = [(65, 128), (128,), (384, 128), (384,), (3, 64), (50257, 16), (7,)]      # 2-D ones are tracked
    offs, total = [], 0
    for shp in sizes:
        offs.append(total)
        total += (int(np.prod(shp)) + 63) // 64 * 64
    p = torch.zeros(total, device=dev); g = torch.zeros(total, device=dev)
    for shp, off in zip(sizes, offs):
        n = int(np.prod(shp))
        p[off:off + n] = torch.randn(n, device=dev) * 0.02
        g[off:off + n] = torch.randn(n, device=dev) * 1e-3
    seg = torch.full((total // 64,), -1, dtype=torch.int32)
    for i, (shp, off) in enumerate(zip(sizes, offs)):
        if len(shp) == 2:
            seg[off // 64:(off + (int(np.prod(shp)) + 63) // 64 * 64) // 64] = i
    seg = seg.to(dev)
    kw = dict(lr=6e-4, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.01)
    for split in (None, [0, offs[2], offs[5], total]):
        p1, p2 = p.clone(), p.clone()
        m1, v1, m2, v2 = (torch.zeros(total, device=dev) for _ in range(4))
        s1 = torch.empty(total, device=dev, dtype=torch.bfloat16); s2 = torch.empty_like(s1)
        g1, g2 = g.clone(), g.clone()
        stats = torch.zeros(len(sizes), 4, device=dev, dtype=torch.float64)
        for step in (1, 2):
            before = p2.clone()
            ops.adamw_flat(p1, g1, m1, v1, s1, step=step, **kw)
            stats.zero_()
            for lo, hi in (zip(split[:-1], split[1:]) if split else [(0, total)]):
                ops.adamw_flat(p2[lo:hi], g2[lo:hi], m2[lo:hi], v2[lo:hi], s2[lo:hi], step=step, zero_grad=False,
                               seg_of_group=seg[lo // 64:hi // 64], seg_stats=stats, **kw)
            assert torch.equal(p1, p2) and torch.equal(m1, m2) and torch.equal(v1, v2) and torch.equal(s1, s2)
            vals = stats.tolist()
            for i, (shp, off) in enumerate(zip(sizes, offs)):
                n = int(np.prod(shp))
                if len(shp) != 2:
                    assert vals[i] == [0.0] * 4
                    continue
                w, pw = p2[off:off + n], before[off:off + n]
                want = ((w - pw).std() / (w.std() + 1e-8)).item()
                sd, sdd, sw, sww = vals[i]
                got = (max(0.0, (sdd - sd * sd / n) / (n - 1)) ** 0.5) / (max(0.0, (sww - sw * sw / n) / (n - 1)) ** 0.5 + 1e-8)
                assert abs(got - want) <= 1e-4 * abs(want), (i, got, want)


# ------------------------------------------------------------------ decode kernels
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_decode_kernels(dev, dtype):
    from penroz import ops, _lib
    torch.manual_seed(6)
    B, H, D, Tctx = 3, 4, 64, 200
    C = H * D
    PAGE = _lib.lib().kv_page_size()
    max_pages = (Tctx + 1 + PAGE - 1) // PAGE
    pool = torch.zeros(B * max_pages, 2, H, PAGE, D, device=dev, dtype=dtype)
    # scrambled page assignment to exercise the indirection
    perm = torch.randperm(B * max_pages, device=dev).view(B, max_pages).int().contiguous()
    hist = torch.randn(B, Tctx, 3 * C, device=dev).to(dtype)
    ops.kv_append(hist.view(B * Tctx, 3 * C), pool, perm, B, Tctx, H, D, 0)
    new = torch.randn(B, 3 * C, device=dev).to(dtype)
    ops.kv_append(new, pool, perm, B, 1, H, D, Tctx)
    out = ops.attn_decode(new, pool, perm, B, H, D, Tctx + 1)
    allk = torch.cat([hist[:, :, C:2 * C], new[:, None, C:2 * C]], 1).view(B, Tctx + 1, H, D).transpose(1, 2).float()
    allv = torch.cat([hist[:, :, 2 * C:], new[:, None, 2 * C:]], 1).view(B, Tctx + 1, H, D).transpose(1, 2).float()
    q = new[:, :C].view(B, 1, H, D).transpose(1, 2).float()
    ref = F.scaled_dot_product_attention(q, allk, allv).transpose(1, 2).reshape(B, C)
    assert rel_err(out, ref) < (1e-5 if dtype == torch.float32 else 8e-3)
    # skinny linear
    for M in (1, 3, 11):
        x = torch.randn(M, 768, device=dev).to(dtype)
        W = (torch.randn(1000, 768, device=dev) * 0.05).to(dtype)
        bias = torch.randn(1000, device=dev).to(dtype)
        ref = x.float() @ W.float().t() + bias.float()
        o = ops.skinny_linear(x, W, bias)
        tol = 1e-5 if dtype == torch.float32 else 4e-3
        assert rel_err(o, ref) < tol
        og = ops.skinny_linear(x, W, bias, epilogue=1)
        assert rel_err(og, F.gelu(ref.to(dtype).float())) < tol * 2
        aux = torch.randn(M, 1000, device=dev).to(dtype)
        orr = ops.skinny_linear(x, W, bias, epilogue=2, aux=aux)
        assert rel_err(orr, aux.float() + ref) < tol * 2
    lg = torch.randn(5, 50304, device=dev).to(dtype)
    lg[2, 100] = 50.0; lg[2, 40000] = 50.0  # tie -> lowest index
    am = ops.argmax_rows(lg, V=50257)
    assert torch.equal(am, lg[:, :50257].float().argmax(dim=-1))
    assert am[2].item() == 100


# ------------------------------------------------------------------ rank-0 diagnostics
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_tensor_stats_match_torch_histogram(dev, dtype):
    from penroz import ops
    torch.manual_seed(11)
    cases = [torch.randn(1000003, device=dev) * 3.0 + 0.5, torch.randn(257, 129, device=dev) ** 3,
             torch.full((4097,), 1.25, device=dev), torch.rand(65536, device=dev) * 1e-3 - 5e-4,
             torch.tensor([-2.0, 7.0], device=dev)]
    for c in cases:
        t = c.to(dtype)
        mean, std, mn, mx, sat, lo, hi, counts = ops.tensor_stats(t, 100, 5.0)
        ref = t.float().cpu()
        h = torch.histogram(ref, bins=100)
        assert counts == h.hist.tolist(), (dtype, tuple(c.shape))           # bin by bin, exactly
        assert lo == h.bin_edges[0].item() and hi == h.bin_edges[-1].item()
        assert mn == ref.min().item() and mx == ref.max().item()
        assert abs(mean - ref.double().mean().item()) <= 1e-6 * max(1.0, abs(mean))
        assert abs(std - ref.double().std().item()) <= 1e-6 * max(1.0, std)
        assert abs(sat - (ref.abs() > 5.0).double().mean().item()) < 1e-12


def test_training_stats_gpu_path_equals_host_path(dev):
    """_record_training_overall_progress on CUDA tensors (penroz_tensor_stats) against the same method on host copies
    (the reference's torch calls, neural_net_model.py:743-777)."""
    from helpers import build_gpt, scratch_cwd
    scratch_cwd()
    m = build_gpt("stats", 2, 4, 128, 64, 65)
    torch.manual_seed(5)
    acts = [(torch.randn(4, 64, 128) * (1 + i), torch.randn(4, 64, 128) * 1e-3) for i in range(5)]
    acts.append((torch.randn(4, 64, 65) * 4, torch.randn(4, 64, 65) * 1e-4))
    for p in m.parameters():
        p.grad = torch.randn_like(p) * 1e-2
    m.progress = [{"cost": 4.0}, {"cost": 3.5}]
    m._record_training_overall_progress(acts)
    host = m.stats
    m.to("cuda")
    for p in m.parameters():
        assert p.grad is not None and p.grad.is_cuda
    m._record_training_overall_progress([(a.cuda(), g.cuda()) for a, g in acts])
    gpu = m.stats

    def close(a, b, path):
        if isinstance(a, dict):
            assert a.keys() == b.keys(), path
            for k in a:
                if k == "mean" and "std" in a:      # a mean near zero: fp32 (host) vs fp64 (GPU) summation, scale = std
                    assert abs(a[k] - b[k]) <= 1e-5 * max(a["std"], 1e-30), (path, a[k], b[k])
                else:
                    close(a[k], b[k], path + "/" + k)
        elif isinstance(a, list):
            assert len(a) == len(b), path
            for i, (x, y) in enumerate(zip(a, b)):
                close(x, y, f"{path}[{i}]")
        elif isinstance(a, float):
            assert abs(a - b) <= 2e-5 * max(abs(a), abs(b), 1e-30) + 1e-12, (path, a, b)
        else:
            assert a == b, (path, a, b)
    close(host, gpu, "stats")


# ------------------------------------------------------------------ int8 KV cache kernels
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_int8_kv_cache_kernels_match_the_torch_class(dev, dtype):
    """penroz_kv_append_q8 / penroz_attn_decode_q8 against TurboQuantKVCache's own torch arithmetic
    (kv_cache.py:119-125,135): codes and scales torch.equal to _quantize run on CUDA tensors (recip_scale=1) and on CPU
    tensors (recip_scale=0, what the reference service computes), the written-back K/V equal to _dequantize, and
    decode attention over the int8 pages equal to SDPA over the de-quantised cache."""
    from kv_cache import TurboQuantKVCache
    from penroz import ops
    torch.manual_seed(3)
    B, S, H, D = 3, 70, 4, 64
    C = H * D
    qkv = (torch.randn(B * S, 3 * C, device=dev) * 1.7).to(dtype)
    qkv[5, C:C + D] = 0                                     # an all-zero key row: scale falls back to 1
    max_pages = 2
    table = torch.arange(B * max_pages, dtype=torch.int32, device=dev).view(B, max_pages)

    def rows(t):        # [B*S, C] column block -> [B, H, S, D]
        return t.view(B, S, H, D).permute(0, 2, 1, 3)

    for recip in (1, 0):
        codes = torch.zeros(B * max_pages, 2, H, 64, D, device=dev, dtype=torch.int8)
        scales = torch.zeros(B * max_pages, 2, H, 64, device=dev, dtype=dtype)
        work = qkv.clone()
        ops.kv_append_q8(work, codes, scales, table, B, S, H, D, 0, writeback=True, recip_scale=bool(recip))
        for kv, lo in ((0, C), (1, 2 * C)):
            src = rows(qkv[:, lo:lo + C])
            want_q, want_s = TurboQuantKVCache._quantize(src if recip else src.cpu())
            want_q, want_s = want_q.to(dev), want_s.to(dev)
            got_q = codes.view(B, max_pages, 2, H, 64, D)[:, :, kv].permute(0, 2, 1, 3, 4).reshape(B, H, 128, D)[:, :, :S]
            got_s = scales.view(B, max_pages, 2, H, 64)[:, :, kv].permute(0, 2, 1, 3).reshape(B, H, 128)[:, :, :S]
            assert torch.equal(got_q, want_q), (recip, kv)
            assert torch.equal(got_s.unsqueeze(-1), want_s), (recip, kv)
            deq = TurboQuantKVCache._dequantize(want_q, want_s)               # fp32
            back = rows(work[:, lo:lo + C])
            assert torch.equal(back.float(), deq.to(dtype).float())
        assert torch.equal(work[:, :C], qkv[:, :C])                             # queries untouched
    # decode attention: one new token per sequence at position S (second page), q8 cache vs SDPA on the de-quantised cache
    new = (torch.randn(B, 3 * C, device=dev) * 1.7).to(dtype)
    ops.kv_append_q8(new, codes, scales, table, B, 1, H, D, S, writeback=False, recip_scale=False)
    out = ops.attn_decode_q8(new, codes, scales, table, B, H, D, S + 1)
    k_all = (codes.view(B, max_pages, 2, H, 64, D)[:, :, 0].permute(0, 2, 1, 3, 4).reshape(B, H, 128, D)[:, :, :S + 1].float()
             * scales.view(B, max_pages, 2, H, 64)[:, :, 0].permute(0, 2, 1, 3).reshape(B, H, 128)[:, :, :S + 1].float().unsqueeze(-1))
    v_all = (codes.view(B, max_pages, 2, H, 64, D)[:, :, 1].permute(0, 2, 1, 3, 4).reshape(B, H, 128, D)[:, :, :S + 1].float()
             * scales.view(B, max_pages, 2, H, 64)[:, :, 1].permute(0, 2, 1, 3).reshape(B, H, 128)[:, :, :S + 1].float().unsqueeze(-1))
    q = new[:, :C].float().view(B, H, 1, D)
    want = F.scaled_dot_product_attention(q, k_all, v_all).reshape(B, C)
    tol = 1e-5 if dtype == torch.float32 else 2 ** -7
    assert (out.float() - want).abs().max().item() <= tol * max(1.0, want.abs().max().item())


# ------------------------------------------------------------------ weight-streaming Linear of the decode step
@pytest.mark.parametrize("M", [1, 5, 8, 16, 33, 64])
def test_wstream_linear(dev, M):
    """penroz_wstream_linear (mma.sync, 16-row units, LayerNorm / GELU / residual fused) against torch on the same
    bf16 inputs with fp32 accumulation: plain, LN + bias, GELU (erf and tanh), residual with a long reduction, and a row
    count that is not a multiple of 16 (the LM head's 50257)."""
    from penroz import ops
    torch.manual_seed(M)
    bf = torch.bfloat16

    def rnd(*shape, s=1.0):
        return (torch.randn(*shape, device=dev) * s).to(bf)
    cases = [(768, 2304, 0, True), (768, 3072, 1, True), (3072, 768, 2, False), (768, 50257, 0, True), (1600, 4800, 1, True),
             (768, 768, 2, False)]
    for K, N, epi, ln in cases:
        x, W, b = rnd(M, K), rnd(N, K, s=0.03), rnd(N, s=0.1)
        aux = rnd(M, N) if epi == 2 else None
        g, be = (1 + rnd(K, s=0.1)), rnd(K, s=0.1)
        xin = x.float()
        if ln:
            xin = F.layer_norm(xin, (K,), g.float(), be.float(), 1e-5).to(bf).float()
        want = (xin @ W.float().t() + b.float()).to(bf).float()
        for tanh in ((False, True) if epi == 1 else (False,)):
            w2 = want
            if epi == 1:
                w2 = F.gelu(want, approximate="tanh" if tanh else "none").to(bf).float()
            elif epi == 2:
                w2 = (aux.float() + want).to(bf).float()
            got = ops.wstream_linear(x, W, b, epilogue=epi, aux=aux, tanh=tanh, ln_w=g if ln else None,
                                     ln_b=be if ln else None).float()
            scale = max(1.0, w2.abs().max().item())
            assert (got - w2).abs().max().item() <= 2 ** -7 * scale, (K, N, epi, ln, tanh, (got - w2).abs().max().item())
