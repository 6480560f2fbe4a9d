// Softmax cross-entropy over a chunk of LM-head logits (forward + in-place gradient), the fused flat AdamW
// step, and the global gradient-norm reduction.
//
// Reference call sites:
//   F.cross_entropy(logits.view(-1,V), target.view(-1))   neural_net_model.py:264-267  (mean, ignore_index=-100)
//   torch.optim.AdamW(...).step()/zero_grad()             mappers.py:267-274, neural_net_model.py:622,677
// The reference has no gradient clipping (grep: none); the norm/clip path is opt-in (PENROZ_CLIP_GRAD_NORM).
#include "common.cuh"
#include <stdlib.h>

namespace penroz {

// ------------------------------------------------------------------ cross entropy
__global__ void ce_count_valid_kernel(const long long* __restrict__ tgt, int M, long long ignore_index,
                                      int* __restrict__ n_valid) {
    __shared__ float red[32];
    float c = 0.f;
    for (int i = threadIdx.x; i < M; i += blockDim.x) c += (tgt[i] != ignore_index) ? 1.f : 0.f;
    c = block_sum(c, red);
    if (threadIdx.x == 0) *n_valid = (int)(c + 0.5f);
}

// One CTA per row. Pass 1: online (max, sum-exp) over the row, 16-byte loads. Pass 2 (optional): overwrite the
// row with d loss / d logits = (softmax - onehot) * grad_scale / n_valid, rounded to the logits dtype (this is
// what autograd hands the LM-head matmuls under autocast). The second read is served from L2.
template <typename T>
__global__ void __launch_bounds__(512)
ce_rows_kernel(T* __restrict__ logits, long long ld, const long long* __restrict__ tgt, int V,
               long long ignore_index, const int* __restrict__ n_valid, float grad_scale,
               float* __restrict__ loss_rows, int write_grad, int* __restrict__ err_flag) {
    __shared__ float red[32];
    const int row = blockIdx.x;
    T* x = logits + (size_t)row * ld;
    const long long t = tgt[row];
    const bool ignored = (t == ignore_index);
    if (!ignored && (t < 0 || t >= V)) {
        if (threadIdx.x == 0 && err_flag) atomicExch(err_flag, 2);
    }
    constexpr int VN = Vec16<T>::N;
    const bool vec = (ld % VN == 0) && (((uintptr_t)logits) % 16 == 0);
    const int Vv = vec ? (V / VN) * VN : 0;
    float m = -INFINITY, s = 0.f;
    if (!ignored || write_grad) {
        for (int i = threadIdx.x * VN; i < Vv; i += blockDim.x * VN) {
            Vec16<T> v;
            v.load(x + i);
            float lm = v.get(0);
#pragma unroll
            for (int k = 1; k < VN; ++k) lm = fmaxf(lm, v.get(k));
            const float nm = fmaxf(m, lm);
            float acc = 0.f;
#pragma unroll
            for (int k = 0; k < VN; ++k) acc += __expf(v.get(k) - nm);
            s = s * __expf(m - nm) + acc;
            m = nm;
        }
        for (int i = Vv + threadIdx.x; i < V; i += blockDim.x) {
            const float v = to_f32(x[i]);
            const float nm = fmaxf(m, v);
            s = s * __expf(m - nm) + __expf(v - nm);
            m = nm;
        }
    }
    const float gm = block_max(m, red);
    const float part = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
    const float gs = block_sum(part, red);
    const float lse = gm + logf(gs);
    if (threadIdx.x == 0) {
        float l = 0.f;
        if (!ignored && t >= 0 && t < V) l = lse - to_f32(x[t]);
        loss_rows[row] = l;
    }
    if (!write_grad) return;
    __syncthreads();  // x[t] read above before the row is overwritten
    const float scale = ignored ? 0.f : grad_scale / (float)max(1, *n_valid);
    for (int i = threadIdx.x * VN; i < Vv; i += blockDim.x * VN) {
        Vec16<T> v;
        v.load(x + i);
#pragma unroll
        for (int k = 0; k < VN; ++k) {
            float p = __expf(v.get(k) - lse);
            if ((long long)(i + k) == t) p -= 1.f;
            v.set(k, p * scale);
        }
        v.store(x + i);
    }
    for (int i = Vv + threadIdx.x; i < V; i += blockDim.x) {
        float p = __expf(to_f32(x[i]) - lse);
        if ((long long)i == t) p -= 1.f;
        x[i] = from_f32<T>(p * scale);
    }
}

// The training form for bf16 logits (16-byte aligned rows, gradient written). Same two passes, built for the two
// limits the generic kernel runs into at V = 50257 (1.65 G elements per step): the SFU (one ex2 per element and pass,
// 0.39 ms each on 148 SMs) and issue slots (15 per element there). Here: the running maximum is rescaled only when it
// moves (a branch taken a handful of times per thread instead of an ex2 per vector), the exponent's argument is one
// packed FMA (x log2e - m log2e), sums and the gradient scale are packed too, the gradient is packed with cvt.bf16x2,
// and the target column is patched by one thread after the pass instead of being tested for on every element.
__global__ void __launch_bounds__(512)
ce_rows_bf16_kernel(__nv_bfloat16* __restrict__ logits, long long ld, const long long* __restrict__ tgt, int V,
                    long long ignore_index, const int* __restrict__ n_valid, float grad_scale,
                    float* __restrict__ loss_rows, int* __restrict__ err_flag) {
    pdl_trigger();   // a dependent-launch successor (the next GEMM) may set itself up while this grid drains
    __shared__ float red[32];
    constexpr float kLog2e = 1.44269504088896340736f;
    const int row = blockIdx.x;
    __nv_bfloat16* x = logits + (size_t)row * ld;
    const long long t = tgt[row];
    const bool ignored = (t == ignore_index);
    const bool t_ok = !ignored && t >= 0 && t < V;
    if (!ignored && !t_ok) {
        if (threadIdx.x == 0 && err_flag) atomicExch(err_flag, 2);
    }
    const int nvec = V >> 3;
    uint4* x4 = reinterpret_cast<uint4*>(x);
    float m = -INFINITY;
    f32x2 s2 = f2_all(0.f);
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
        const uint4 raw = x4[i];
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
        float lo[4], hi[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) { lo[k] = __uint_as_float(w[k] << 16); hi[k] = __uint_as_float(w[k] & 0xffff0000u); }
        const float lm = fmaxf(fmaxf(fmaxf(lo[0], hi[0]), fmaxf(lo[1], hi[1])), fmaxf(fmaxf(lo[2], hi[2]), fmaxf(lo[3], hi[3])));
        if (lm > m) {   // rare after the first vectors
            const float r = ex2_fast((m - lm) * kLog2e);   // m = -inf: 0
            s2 = f2_mul(s2, f2_all(r));
            m = lm;
        }
        const f32x2 off = f2_all(-m * kLog2e);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float a0, a1;
            f2_unpack(f2_fma(f2_pack(lo[k], hi[k]), f2_all(kLog2e), off), a0, a1);
            s2 = f2_add(s2, f2_pack(ex2_fast(a0), ex2_fast(a1)));
        }
    }
    float s;
    { float s0, s1; f2_unpack(s2, s0, s1); s = s0 + s1; }
    for (int i = nvec * 8 + threadIdx.x; i < V; i += blockDim.x) {   // < 8 tail elements
        const float v = __bfloat162float(x[i]);
        const float nm = fmaxf(m, v);
        s = s * __expf(m - nm) + __expf(v - nm);
        m = nm;
    }
    const float gm = block_max(m, red);
    const float part = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
    const float gs = block_sum(part, red);
    const float lse = gm + logf(gs);
    float xt = 0.f;
    if (threadIdx.x == 0) {
        if (t_ok) xt = __bfloat162float(x[t]);
        loss_rows[row] = t_ok ? lse - xt : 0.f;
    }
    const float scale = ignored ? 0.f : grad_scale / (float)max(1, *n_valid);
    const f32x2 off = f2_all(-lse * kLog2e), sc2 = f2_all(scale);
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
        const uint4 raw = x4[i];   // served from L2
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float a0, a1, p0, p1;
            f2_unpack(f2_fma(f2_pack(__uint_as_float(w[k] << 16), __uint_as_float(w[k] & 0xffff0000u)), f2_all(kLog2e), off), a0, a1);
            f2_unpack(f2_mul(f2_pack(ex2_fast(a0), ex2_fast(a1)), sc2), p0, p1);
            o[k] = pack_bf16x2(p0, p1);
        }
        x4[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
    for (int i = nvec * 8 + threadIdx.x; i < V; i += blockDim.x)
        x[i] = __float2bfloat16_rn(__expf(__bfloat162float(x[i]) - lse) * scale);
    __syncthreads();   // the pass has written column t
    if (threadIdx.x == 0 && t_ok) x[t] = __float2bfloat16_rn((__expf(xt - lse) - 1.f) * scale);
}

// loss = sum(loss_rows) / n_valid, single CTA, fixed order (deterministic); *loss_out (+)= loss * out_scale
__global__ void ce_reduce_kernel(const float* __restrict__ loss_rows, int M, const int* __restrict__ n_valid,
                                 float out_scale, float* __restrict__ loss_out, int accumulate) {
    __shared__ float red[32];
    float s = 0.f;
    for (int i = threadIdx.x; i < M; i += blockDim.x) s += loss_rows[i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
        const int nv = *n_valid;
        const float l = (nv > 0) ? s / (float)nv : __int_as_float(0x7fc00000);  // torch: mean over zero rows = nan
        *loss_out = accumulate ? *loss_out + l * out_scale : l * out_scale;
    }
}

// ------------------------------------------------------------------ AdamW over flat fp32 buffers
// Same operation order as torch's _single_tensor_adamw (param decay, lerp, addcmul, sqrt/bias-corr, addcdiv).
// Written with explicit rounding intrinsics so that every kernel using it produces the same bits (the compiler may
// contract a*b+c differently from one kernel to the next).
struct AdamwCoef { float lr_wd_keep, one_minus_beta1, beta2, one_minus_beta2, eps, step_size_neg, bc2_sqrt; };
__device__ __forceinline__ void adamw_update(float& p, float g, float& m, float& v, const AdamwCoef& c) {
    const float pk = __fmul_rn(p, c.lr_wd_keep);                                   // param.mul_(1 - lr * wd)
    m = __fmaf_rn(__fsub_rn(g, m), c.one_minus_beta1, m);                              // exp_avg.lerp_(grad, 1 - beta1)
    v = __fmaf_rn(__fmul_rn(c.one_minus_beta2, g), g, __fmul_rn(v, c.beta2));          // .mul_(beta2).addcmul_(g, g, 1 - beta2)
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), c.bc2_sqrt), c.eps);   // sqrt / bias_correction2_sqrt + eps
    p = __fmaf_rn(c.step_size_neg, __fdiv_rn(m, denom), pk);                       // param.addcdiv_(m, denom, -step_size)
}

__global__ void adamw_flat_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                  float* __restrict__ v, __nv_bfloat16* __restrict__ shadow, size_t n, float lr_wd_keep,
                                  float omb1, float beta2, float omb2, float eps, float step_size_neg, float bc2_sqrt,
                                  const float* __restrict__ grad_mult, int zero_grad) {
    const float gm = grad_mult ? *grad_mult : 1.f;
    const AdamwCoef coef{lr_wd_keep, omb1, beta2, omb2, eps, step_size_neg, bc2_sqrt};
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (; i < n; i += stride) {
        if (i + 3 < n) {
            float4 pv = *reinterpret_cast<float4*>(p + i);
            float4 gv = *reinterpret_cast<float4*>(g + i);
            float4 mv = *reinterpret_cast<float4*>(m + i);
            float4 vv = *reinterpret_cast<float4*>(v + i);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                adamw_update((&pv.x)[k], __fmul_rn((&gv.x)[k], gm), (&mv.x)[k], (&vv.x)[k], coef);
            *reinterpret_cast<float4*>(p + i) = pv;
            *reinterpret_cast<float4*>(m + i) = mv;
            *reinterpret_cast<float4*>(v + i) = vv;
            if (zero_grad) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
            if (shadow) {
                uint2 w;
                w.x = pack_bf16x2(pv.x, pv.y);
                w.y = pack_bf16x2(pv.z, pv.w);
                *reinterpret_cast<uint2*>(shadow + i) = w;
            }
        } else {
            for (size_t k = i; k < n; ++k) {
                float pk = p[k], mk = m[k], vk = v[k];
                adamw_update(pk, __fmul_rn(g[k], gm), mk, vk, coef);
                p[k] = pk; m[k] = mk; v[k] = vk;
                if (zero_grad) g[k] = 0.f;
                if (shadow) shadow[k] = __float2bfloat16_rn(pk);
            }
        }
    }
}

// The same update, additionally folding what the reference's progress entry needs per 2-D weight -- std(w_new - w_old)
// and std(w_new) (neural_net_model.py:619-620,696-699: a clone of every weight before the step and ~3 reductions per
// weight after it) -- into the pass that already holds w_old and w_new in registers. The flat buffer is cut into
// 64-element groups (every parameter starts on a group boundary); seg_of_group[group] names the tracked tensor the
// group belongs to (-1: not tracked). A CTA sweeps a CONTIGUOUS range of the buffer, a half-warp owns one group per
// sweep step and keeps fp64 running sums {sum d, sum d^2, sum w, sum w^2} for its current tensor; they are flushed with
// four fp64 atomics when the tensor changes and at the end (a few thousand atomics per step in total).
__global__ void __launch_bounds__(256)
adamw_flat_stats_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                        __nv_bfloat16* __restrict__ shadow, size_t n, float lr_wd_keep, float omb1, float beta2,
                        float omb2, float eps, float step_size_neg, float bc2_sqrt, const float* __restrict__ grad_mult,
                        int zero_grad, const int* __restrict__ seg_of_group, double* __restrict__ seg_stats) {
    const float gm = grad_mult ? *grad_mult : 1.f;
    const AdamwCoef coef{lr_wd_keep, omb1, beta2, omb2, eps, step_size_neg, bc2_sqrt};
    constexpr size_t TILE = 256 * 4;                       // elements per sweep step of a CTA
    const size_t tiles = (n + TILE - 1) / TILE;
    const size_t t0 = tiles * blockIdx.x / gridDim.x, t1 = tiles * (blockIdx.x + 1) / gridDim.x;
    int cur = -1;
    double sd = 0.0, sdd = 0.0, sw = 0.0, sww = 0.0;
    const unsigned lane = threadIdx.x & 31;
    auto flush = [&](bool mine) {                          // called by whole warps; `mine`: uniform per half-warp
        double a = sd, b = sdd, c = sw, d = sww;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
            c += __shfl_xor_sync(0xffffffffu, c, o);
            d += __shfl_xor_sync(0xffffffffu, d, o);
        }
        if (mine) {
            if ((lane & 15) == 0 && cur >= 0) {
                double* out = seg_stats + (size_t)cur * 4;
                atomicAdd(out + 0, a); atomicAdd(out + 1, b); atomicAdd(out + 2, c); atomicAdd(out + 3, d);
            }
            sd = sdd = sw = sww = 0.0;
        }
    };
    for (size_t t = t0; t < t1; ++t) {
        const size_t i = t * TILE + (size_t)threadIdx.x * 4;
        const bool live = i < n;                           // n is a multiple of 64: a float4 is all in or all out
        const int seg = live ? seg_of_group[i >> 6] : cur;
        const bool changed = seg != cur;
        if (__any_sync(0xffffffffu, changed)) {
            flush(changed);
            if (changed) cur = seg;
        }
        if (!live) continue;
        float4 pv = *reinterpret_cast<float4*>(p + i);
        float4 gv = *reinterpret_cast<float4*>(g + i);
        float4 mv = *reinterpret_cast<float4*>(m + i);
        float4 vv = *reinterpret_cast<float4*>(v + i);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float p_old = (&pv.x)[k];
            adamw_update((&pv.x)[k], __fmul_rn((&gv.x)[k], gm), (&mv.x)[k], (&vv.x)[k], coef);
            const float pk = (&pv.x)[k];
            const double dk = (double)__fsub_rn(pk, p_old), wk = (double)pk;   // w - w_prev as the reference forms it: in fp32
            sd += dk; sdd = fma(dk, dk, sdd); sw += wk; sww = fma(wk, wk, sww);
        }
        *reinterpret_cast<float4*>(p + i) = pv;
        *reinterpret_cast<float4*>(m + i) = mv;
        *reinterpret_cast<float4*>(v + i) = vv;
        if (zero_grad) *reinterpret_cast<float4*>(g + i) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (shadow) {
            uint2 w;
            w.x = pack_bf16x2(pv.x, pv.y);
            w.y = pack_bf16x2(pv.z, pv.w);
            *reinterpret_cast<uint2*>(shadow + i) = w;
        }
    }
    flush(true);
}

// sum of squares -> per-block partials (scratch) -> norm, clip multiplier
__global__ void sumsq_partial_kernel(const float* __restrict__ g, size_t n, float* __restrict__ part) {
    __shared__ float red[32];
    float s = 0.f;
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (; i < n; i += stride) {
        if (i + 3 < n) {
            float4 v = *reinterpret_cast<const float4*>(g + i);
            s += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
        } else {
            for (size_t k = i; k < n; ++k) s += g[k] * g[k];
        }
    }
    s = block_sum(s, red);
    if (threadIdx.x == 0) part[blockIdx.x] = s;
}
__global__ void norm_finish_kernel(const float* __restrict__ part, int nparts, float max_norm,
                                   float* __restrict__ norm_out, float* __restrict__ mult_out) {
    __shared__ float red[32];
    float s = 0.f;
    for (int i = threadIdx.x; i < nparts; i += blockDim.x) s += part[i];
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
        const float nrm = sqrtf(s);
        if (norm_out) *norm_out = nrm;
        if (mult_out) {
            // torch.nn.utils.clip_grad_norm_: clip_coef = max_norm / (norm + 1e-6), clamped to 1
            const float c = max_norm / (nrm + 1e-6f);
            *mult_out = (max_norm > 0.f && c < 1.f) ? c : 1.f;
        }
    }
}

static int flat_grid(size_t n) {
    size_t g = (n / 4 + 255) / 256;
    size_t cap = (size_t)num_sms() * 8;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace penroz

using namespace penroz;

extern "C" int penroz_ce_count_valid(const long long* targets, int M, long long ignore_index, int* n_valid,
                                     void* stream) {
    PENROZ_REQUIRE(M > 0 && targets && n_valid, "ce_count_valid: bad arguments");
    ce_count_valid_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(targets, M, ignore_index, n_valid);
    count_launch();
    return check_launch("ce_count_valid");
}

extern "C" int penroz_ce_rows(void* logits, long long ld, const long long* targets, int M, int V,
                              long long ignore_index, const int* n_valid, float grad_scale, float* loss_rows,
                              int write_grad, int dtype, int* err_flag, void* stream) {
    PENROZ_REQUIRE(M > 0 && V > 0 && ld >= V, "ce_rows: bad shape M=%d V=%d ld=%lld", M, V, ld);
    PENROZ_REQUIRE(logits && targets && loss_rows && n_valid, "ce_rows: null argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int threads = V >= 8192 ? 512 : (V >= 1024 ? 256 : 64);
    static const bool generic = getenv("PENROZ_CE_GENERIC") != nullptr;   // A/B switch
    if (!generic && dtype == BF16 && write_grad && V >= 4096 && (ld % 8 == 0) && ((uintptr_t)logits % 16 == 0)) {
        ce_rows_bf16_kernel<<<M, 512, 0, st>>>((__nv_bfloat16*)logits, ld, targets, V, ignore_index, n_valid, grad_scale, loss_rows, err_flag);
    } else if (dtype == BF16)
        ce_rows_kernel<__nv_bfloat16><<<M, threads, 0, st>>>((__nv_bfloat16*)logits, ld, targets, V, ignore_index, n_valid, grad_scale, loss_rows, write_grad, err_flag);
    else if (dtype == F32)
        ce_rows_kernel<float><<<M, threads, 0, st>>>((float*)logits, ld, targets, V, ignore_index, n_valid, grad_scale, loss_rows, write_grad, err_flag);
    else { set_error("ce_rows: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("ce_rows");
}

extern "C" int penroz_ce_reduce(const float* loss_rows, int M, const int* n_valid, float out_scale, float* loss_out,
                                int accumulate, void* stream) {
    PENROZ_REQUIRE(M > 0 && loss_rows && n_valid && loss_out, "ce_reduce: bad arguments");
    ce_reduce_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(loss_rows, M, n_valid, out_scale, loss_out, accumulate);
    count_launch();
    return check_launch("ce_reduce");
}

extern "C" int penroz_adamw_flat(float* p, float* g, float* m, float* v, void* shadow_bf16, long long n, double lr,
                                 double beta1, double beta2, double eps, double weight_decay, long long step,
                                 const float* grad_mult, int zero_grad, void* stream) {
    PENROZ_REQUIRE(n > 0 && p && g && m && v, "adamw_flat: bad arguments");
    PENROZ_REQUIRE(step >= 1, "adamw_flat: step must be >= 1");
    PENROZ_REQUIRE(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) % 16 == 0, "adamw_flat: buffers must be 16-byte aligned");
    const double bc1 = 1.0 - pow(beta1, (double)step);
    const double bc2 = 1.0 - pow(beta2, (double)step);
    const double step_size = lr / bc1;
    const double bc2_sqrt = sqrt(bc2);
    adamw_flat_kernel<<<flat_grid((size_t)n), 256, 0, (cudaStream_t)stream>>>(
        p, g, m, v, (__nv_bfloat16*)shadow_bf16, (size_t)n, (float)(1.0 - lr * weight_decay), (float)(1.0 - beta1), (float)beta2,
        (float)(1.0 - beta2), (float)eps, (float)(-step_size), (float)bc2_sqrt, grad_mult, zero_grad);
    count_launch();
    return check_launch("adamw_flat");
}

extern "C" int penroz_adamw_flat_stats(float* p, float* g, float* m, float* v, void* shadow_bf16, long long n, double lr,
                                       double beta1, double beta2, double eps, double weight_decay, long long step,
                                       const float* grad_mult, int zero_grad, const int* seg_of_group,
                                       double* seg_stats, void* stream) {
    PENROZ_REQUIRE(n > 0 && p && g && m && v && seg_of_group && seg_stats, "adamw_flat_stats: bad arguments");
    PENROZ_REQUIRE(n % 64 == 0, "adamw_flat_stats: n=%lld must be a multiple of the 64-element group", n);
    PENROZ_REQUIRE(step >= 1, "adamw_flat_stats: step must be >= 1");
    PENROZ_REQUIRE(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) % 16 == 0, "adamw_flat_stats: buffers must be 16-byte aligned");
    const double bc1 = 1.0 - pow(beta1, (double)step);
    const double bc2 = 1.0 - pow(beta2, (double)step);
    const double step_size = lr / bc1;
    const double bc2_sqrt = sqrt(bc2);
    size_t tiles = ((size_t)n + 1023) / 1024;
    size_t grid = (size_t)num_sms() * 8;
    if (grid > tiles) grid = tiles;
    adamw_flat_stats_kernel<<<(int)grid, 256, 0, (cudaStream_t)stream>>>(
        p, g, m, v, (__nv_bfloat16*)shadow_bf16, (size_t)n, (float)(1.0 - lr * weight_decay), (float)(1.0 - beta1), (float)beta2,
        (float)(1.0 - beta2), (float)eps, (float)(-step_size), (float)bc2_sqrt, grad_mult, zero_grad, seg_of_group, seg_stats);
    count_launch();
    return check_launch("adamw_flat_stats");
}

extern "C" size_t penroz_grad_norm_scratch_bytes(long long n) { return (size_t)flat_grid((size_t)n) * sizeof(float); }

extern "C" int penroz_grad_norm(const float* g, long long n, float max_norm, float* norm_out, float* mult_out,
                                void* scratch, size_t scratch_bytes, void* stream) {
    PENROZ_REQUIRE(n > 0 && g, "grad_norm: bad arguments");
    PENROZ_REQUIRE(scratch && scratch_bytes >= penroz_grad_norm_scratch_bytes(n), "grad_norm: scratch too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = flat_grid((size_t)n);
    sumsq_partial_kernel<<<grid, 256, 0, st>>>(g, (size_t)n, (float*)scratch);
    norm_finish_kernel<<<1, 256, 0, st>>>((const float*)scratch, grid, max_norm, norm_out, mult_out);
    count_launch(2);
    return check_launch("grad_norm");
}
