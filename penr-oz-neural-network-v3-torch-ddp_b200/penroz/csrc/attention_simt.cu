// Causal self-attention forward/backward on the packed QKV buffer, fp32 arithmetic on the FP32 pipes.
//
// Replaces CausalSelfAttention.forward's split / view / transpose / F.scaled_dot_product_attention(is_causal=True)
// / transpose+contiguous sequence (neural_net_layers.py:59-95, GPT-2 branch: no RoPE, no GQA) by kernels that read
// q,k,v straight out of the [B,T,3C] Linear output and write [B,T,C]: no transposes are materialised.
// This file is the exact-arithmetic path (fp32 models, and the cross-check for the tensor-core kernel):
// online-softmax flash attention, one query (or key) row per D/32 threads, K/V (or Q/dO) tiles staged in smem.
// scale = 1/sqrt(D) (torch default). lse[b,h,t] = log sum_j exp(scale * q_t.k_j) is saved for the backward.
#include "common.cuh"

namespace penroz {

constexpr int ATT_THREADS = 128;
constexpr int ATT_TILE = 32;

template <int TPR> __device__ __forceinline__ float row_sum(float v) {
    if (TPR >= 2) v += __shfl_xor_sync(0xffffffffu, v, 1);
    if (TPR >= 4) v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}

// smem tile [TPR][ATT_TILE][32] with +4 float skew between parts (keeps the TPR parts on different banks)
template <int TPR> struct TileS {
    static constexpr int kPart = ATT_TILE * 32 + 4;
    static constexpr int kFloats = TPR * kPart;
};

// load rows [r0, r0+32) x D of one head (column offset col0 inside the packed row) into a tile
template <typename T, int TPR>
__device__ __forceinline__ void load_tile(float* dst, const T* __restrict__ base, long long row_stride, int r0,
                                          int nrows, float mul) {
    constexpr int D = TPR * 32;
    for (int e = threadIdx.x; e < ATT_TILE * D; e += ATT_THREADS) {
        const int r = e / D, d = e % D;
        float v = 0.f;
        if (r0 + r < nrows) v = to_f32(base[(long long)(r0 + r) * row_stride + d]) * mul;
        dst[(d >> 5) * TileS<TPR>::kPart + r * 32 + (d & 31)] = v;
    }
}

template <typename T, int TPR>
__global__ void __launch_bounds__(ATT_THREADS)
attn_fwd_simt_kernel(const T* __restrict__ qkv, T* __restrict__ out, float* __restrict__ lse, int Tn, int H,
                     float scale) {
    constexpr int D = TPR * 32, RPB = ATT_THREADS / TPR;
    __shared__ float Ks[TileS<TPR>::kFloats];
    __shared__ float Vs[TileS<TPR>::kFloats];
    const int b = blockIdx.z, h = blockIdx.y;
    const int C = H * D;
    const long long rs = 3LL * C;
    const T* base = qkv + (long long)b * Tn * rs + (long long)h * D;
    const int part = threadIdx.x % TPR;
    const int i = blockIdx.x * RPB + threadIdx.x / TPR;
    const bool valid = i < Tn;
    const int ic = valid ? i : Tn - 1;
    float q[32], o[32];
#pragma unroll
    for (int d = 0; d < 32; ++d) {
        q[d] = to_f32(base[(long long)ic * rs + part * 32 + d]) * scale;
        o[d] = 0.f;
    }
    float m = -INFINITY, l = 0.f;
    const int last = min(Tn, (int)(blockIdx.x + 1) * RPB) - 1;
    const int ntiles = last / ATT_TILE + 1;
    for (int t = 0; t < ntiles; ++t) {
        __syncthreads();
        load_tile<T, TPR>(Ks, base + C, rs, t * ATT_TILE, Tn, 1.f);
        load_tile<T, TPR>(Vs, base + 2 * C, rs, t * ATT_TILE, Tn, 1.f);
        __syncthreads();
        const float* kp = Ks + part * TileS<TPR>::kPart;
        const float* vp = Vs + part * TileS<TPR>::kPart;
        float s[ATT_TILE];
        float mt = -INFINITY;
#pragma unroll
        for (int j = 0; j < ATT_TILE; ++j) {
            float a = 0.f;
#pragma unroll
            for (int d = 0; d < 32; d += 4) {
                const float4 kv = *reinterpret_cast<const float4*>(kp + j * 32 + d);
                a = fmaf(q[d], kv.x, a); a = fmaf(q[d + 1], kv.y, a);
                a = fmaf(q[d + 2], kv.z, a); a = fmaf(q[d + 3], kv.w, a);
            }
            a = row_sum<TPR>(a);
            const int kj = t * ATT_TILE + j;
            s[j] = (kj <= ic) ? a : -INFINITY;
            mt = fmaxf(mt, s[j]);
        }
        const float mn = fmaxf(m, mt);
        if (mn != -INFINITY) {
            const float corr = __expf(m - mn);
            l *= corr;
#pragma unroll
            for (int d = 0; d < 32; ++d) o[d] *= corr;
#pragma unroll
            for (int j = 0; j < ATT_TILE; ++j) {
                const float pj = __expf(s[j] - mn);
                l += pj;
#pragma unroll
                for (int d = 0; d < 32; d += 4) {
                    const float4 vv = *reinterpret_cast<const float4*>(vp + j * 32 + d);
                    o[d] = fmaf(pj, vv.x, o[d]); o[d + 1] = fmaf(pj, vv.y, o[d + 1]);
                    o[d + 2] = fmaf(pj, vv.z, o[d + 2]); o[d + 3] = fmaf(pj, vv.w, o[d + 3]);
                }
            }
            m = mn;
        }
    }
    if (valid) {
        const float inv = 1.f / l;
        T* op = out + ((long long)b * Tn + i) * C + h * D + part * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) op[d] = from_f32<T>(o[d] * inv);
        if (part == 0 && lse) lse[((long long)b * H + h) * Tn + i] = m + logf(l);
    }
}

// dQ (and delta = rowsum(dO * O), saved for the dK/dV kernel): one query row per TPR threads
template <typename T, int TPR>
__global__ void __launch_bounds__(ATT_THREADS)
attn_bwd_dq_simt_kernel(const T* __restrict__ qkv, const T* __restrict__ out, const T* __restrict__ dout,
                        const float* __restrict__ lse, T* __restrict__ dqkv, float* __restrict__ delta, int Tn, int H,
                        float scale) {
    constexpr int D = TPR * 32, RPB = ATT_THREADS / TPR;
    __shared__ float Ks[TileS<TPR>::kFloats];
    __shared__ float Vs[TileS<TPR>::kFloats];
    const int b = blockIdx.z, h = blockIdx.y;
    const int C = H * D;
    const long long rs = 3LL * C;
    const T* base = qkv + (long long)b * Tn * rs + (long long)h * D;
    const int part = threadIdx.x % TPR;
    const int i = blockIdx.x * RPB + threadIdx.x / TPR;
    const bool valid = i < Tn;
    const int ic = valid ? i : Tn - 1;
    float q[32], go[32], dq[32];
    float dl = 0.f;
    {
        const T* op = out + ((long long)b * Tn + ic) * C + h * D + part * 32;
        const T* gp = dout + ((long long)b * Tn + ic) * C + h * D + part * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) {
            q[d] = to_f32(base[(long long)ic * rs + part * 32 + d]) * scale;
            go[d] = to_f32(gp[d]);
            dl = fmaf(go[d], to_f32(op[d]), dl);
            dq[d] = 0.f;
        }
    }
    dl = row_sum<TPR>(dl);
    const float L = lse[((long long)b * H + h) * Tn + ic];
    if (valid && part == 0) delta[((long long)b * H + h) * Tn + i] = dl;
    const int last = min(Tn, (int)(blockIdx.x + 1) * RPB) - 1;
    const int ntiles = last / ATT_TILE + 1;
    for (int t = 0; t < ntiles; ++t) {
        __syncthreads();
        load_tile<T, TPR>(Ks, base + C, rs, t * ATT_TILE, Tn, 1.f);
        load_tile<T, TPR>(Vs, base + 2 * C, rs, t * ATT_TILE, Tn, 1.f);
        __syncthreads();
        const float* kp = Ks + part * TileS<TPR>::kPart;
        const float* vp = Vs + part * TileS<TPR>::kPart;
#pragma unroll 4
        for (int j = 0; j < ATT_TILE; ++j) {
            float a = 0.f, dp = 0.f;
#pragma unroll
            for (int d = 0; d < 32; d += 4) {
                const float4 kv = *reinterpret_cast<const float4*>(kp + j * 32 + d);
                const float4 vv = *reinterpret_cast<const float4*>(vp + j * 32 + d);
                a = fmaf(q[d], kv.x, a); a = fmaf(q[d + 1], kv.y, a);
                a = fmaf(q[d + 2], kv.z, a); a = fmaf(q[d + 3], kv.w, a);
                dp = fmaf(go[d], vv.x, dp); dp = fmaf(go[d + 1], vv.y, dp);
                dp = fmaf(go[d + 2], vv.z, dp); dp = fmaf(go[d + 3], vv.w, dp);
            }
            a = row_sum<TPR>(a);
            dp = row_sum<TPR>(dp);
            const int kj = t * ATT_TILE + j;
            const float pj = (kj <= ic) ? __expf(a - L) : 0.f;
            const float ds = pj * (dp - dl);
#pragma unroll
            for (int d = 0; d < 32; d += 4) {
                const float4 kv = *reinterpret_cast<const float4*>(kp + j * 32 + d);
                dq[d] = fmaf(ds, kv.x, dq[d]); dq[d + 1] = fmaf(ds, kv.y, dq[d + 1]);
                dq[d + 2] = fmaf(ds, kv.z, dq[d + 2]); dq[d + 3] = fmaf(ds, kv.w, dq[d + 3]);
            }
        }
    }
    if (valid) {
        T* gq = dqkv + ((long long)b * Tn + i) * rs + h * D + part * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) gq[d] = from_f32<T>(dq[d] * scale);
    }
}

// dK, dV: one key row per TPR threads; Q and dO tiles staged in smem
template <typename T, int TPR>
__global__ void __launch_bounds__(ATT_THREADS)
attn_bwd_dkv_simt_kernel(const T* __restrict__ qkv, const T* __restrict__ dout, const float* __restrict__ lse,
                         const float* __restrict__ delta, T* __restrict__ dqkv, int Tn, int H, float scale) {
    constexpr int D = TPR * 32, RPB = ATT_THREADS / TPR;
    __shared__ float Qs[TileS<TPR>::kFloats];
    __shared__ float Gs[TileS<TPR>::kFloats];
    __shared__ float Ls[ATT_TILE], Ds[ATT_TILE];
    const int b = blockIdx.z, h = blockIdx.y;
    const int C = H * D;
    const long long rs = 3LL * C;
    const T* base = qkv + (long long)b * Tn * rs + (long long)h * D;
    const T* gbase = dout + (long long)b * Tn * C + (long long)h * D;
    const int part = threadIdx.x % TPR;
    const int j = blockIdx.x * RPB + threadIdx.x / TPR;
    const bool valid = j < Tn;
    const int jc = valid ? j : Tn - 1;
    float k[32], v[32], dk[32], dv[32];
#pragma unroll
    for (int d = 0; d < 32; ++d) {
        k[d] = to_f32(base[(long long)jc * rs + C + part * 32 + d]);
        v[d] = to_f32(base[(long long)jc * rs + 2 * C + part * 32 + d]);
        dk[d] = 0.f;
        dv[d] = 0.f;
    }
    const int first_tile = (blockIdx.x * RPB) / ATT_TILE;
    const int ntiles = (Tn + ATT_TILE - 1) / ATT_TILE;
    for (int t = first_tile; t < ntiles; ++t) {
        __syncthreads();
        load_tile<T, TPR>(Qs, base, rs, t * ATT_TILE, Tn, scale);
        load_tile<T, TPR>(Gs, gbase, (long long)C, t * ATT_TILE, Tn, 1.f);
        if (threadIdx.x < ATT_TILE) {
            const int i = t * ATT_TILE + threadIdx.x;
            Ls[threadIdx.x] = (i < Tn) ? lse[((long long)b * H + h) * Tn + i] : 0.f;
            Ds[threadIdx.x] = (i < Tn) ? delta[((long long)b * H + h) * Tn + i] : 0.f;
        }
        __syncthreads();
        const float* qp = Qs + part * TileS<TPR>::kPart;
        const float* gp = Gs + part * TileS<TPR>::kPart;
#pragma unroll 2
        for (int r = 0; r < ATT_TILE; ++r) {
            const int i = t * ATT_TILE + r;
            float a = 0.f, dp = 0.f;
#pragma unroll
            for (int d = 0; d < 32; d += 4) {
                const float4 qv = *reinterpret_cast<const float4*>(qp + r * 32 + d);
                const float4 gv = *reinterpret_cast<const float4*>(gp + r * 32 + d);
                a = fmaf(qv.x, k[d], a); a = fmaf(qv.y, k[d + 1], a);
                a = fmaf(qv.z, k[d + 2], a); a = fmaf(qv.w, k[d + 3], a);
                dp = fmaf(gv.x, v[d], dp); dp = fmaf(gv.y, v[d + 1], dp);
                dp = fmaf(gv.z, v[d + 2], dp); dp = fmaf(gv.w, v[d + 3], dp);
            }
            a = row_sum<TPR>(a);
            dp = row_sum<TPR>(dp);
            const float pj = (i >= jc && i < Tn) ? __expf(a - Ls[r]) : 0.f;
            const float ds = pj * (dp - Ds[r]);
#pragma unroll
            for (int d = 0; d < 32; d += 4) {
                const float4 qv = *reinterpret_cast<const float4*>(qp + r * 32 + d);
                const float4 gv = *reinterpret_cast<const float4*>(gp + r * 32 + d);
                dv[d] = fmaf(pj, gv.x, dv[d]); dv[d + 1] = fmaf(pj, gv.y, dv[d + 1]);
                dv[d + 2] = fmaf(pj, gv.z, dv[d + 2]); dv[d + 3] = fmaf(pj, gv.w, dv[d + 3]);
                dk[d] = fmaf(ds, qv.x, dk[d]); dk[d + 1] = fmaf(ds, qv.y, dk[d + 1]);   // Qs already carries `scale`
                dk[d + 2] = fmaf(ds, qv.z, dk[d + 2]); dk[d + 3] = fmaf(ds, qv.w, dk[d + 3]);
            }
        }
    }
    if (valid) {
        T* gk = dqkv + ((long long)b * Tn + j) * rs + C + h * D + part * 32;
        T* gv = dqkv + ((long long)b * Tn + j) * rs + 2 * C + h * D + part * 32;
#pragma unroll
        for (int d = 0; d < 32; ++d) {
            gk[d] = from_f32<T>(dk[d]);
            gv[d] = from_f32<T>(dv[d]);
        }
    }
}

template <typename T, int TPR>
static void launch_attn_fwd(const void* qkv, void* out, float* lse, int B, int Tn, int H, cudaStream_t st) {
    constexpr int RPB = ATT_THREADS / TPR;
    dim3 grid((Tn + RPB - 1) / RPB, H, B);
    attn_fwd_simt_kernel<T, TPR><<<grid, ATT_THREADS, 0, st>>>((const T*)qkv, (T*)out, lse, Tn, H, rsqrtf((float)(TPR * 32)));
}
template <typename T, int TPR>
static void launch_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                            float* delta, int B, int Tn, int H, cudaStream_t st) {
    constexpr int RPB = ATT_THREADS / TPR;
    dim3 grid((Tn + RPB - 1) / RPB, H, B);
    const float scale = rsqrtf((float)(TPR * 32));
    attn_bwd_dq_simt_kernel<T, TPR><<<grid, ATT_THREADS, 0, st>>>((const T*)qkv, (const T*)out, (const T*)dout, lse, (T*)dqkv, delta, Tn, H, scale);
    attn_bwd_dkv_simt_kernel<T, TPR><<<grid, ATT_THREADS, 0, st>>>((const T*)qkv, (const T*)dout, lse, delta, (T*)dqkv, Tn, H, scale);
}

}  // namespace penroz

using namespace penroz;

extern "C" int penroz_attn_fwd_simt(const void* qkv, void* out, float* lse, int B, int T, int H, int D, int dtype,
                                    void* stream) {
    PENROZ_REQUIRE(B > 0 && T > 0 && H > 0, "attn_fwd: bad shape");
    PENROZ_REQUIRE(D == 32 || D == 64 || D == 128, "attn_fwd: head_dim %d unsupported (32, 64, 128)", D);
    PENROZ_REQUIRE(B <= 65535 && H <= 65535, "attn_fwd: B/H too large for the launch grid");
    cudaStream_t st = (cudaStream_t)stream;
#define ATT_FWD(T_) \
    if (D == 32) launch_attn_fwd<T_, 1>(qkv, out, lse, B, T, H, st); \
    else if (D == 64) launch_attn_fwd<T_, 2>(qkv, out, lse, B, T, H, st); \
    else launch_attn_fwd<T_, 4>(qkv, out, lse, B, T, H, st)
    if (dtype == F32) { ATT_FWD(float); }
    else if (dtype == BF16) { ATT_FWD(__nv_bfloat16); }
    else { set_error("attn_fwd: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
#undef ATT_FWD
    count_launch();
    return check_launch("attn_fwd_simt");
}

extern "C" size_t penroz_attn_bwd_scratch_bytes(int B, int T, int H) { return (size_t)B * H * T * sizeof(float); }

extern "C" int penroz_attn_bwd_simt(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                    int B, int T, int H, int D, int dtype, void* scratch, size_t scratch_bytes,
                                    void* stream) {
    PENROZ_REQUIRE(B > 0 && T > 0 && H > 0, "attn_bwd: bad shape");
    PENROZ_REQUIRE(D == 32 || D == 64 || D == 128, "attn_bwd: head_dim %d unsupported (32, 64, 128)", D);
    PENROZ_REQUIRE(scratch && scratch_bytes >= penroz_attn_bwd_scratch_bytes(B, T, H), "attn_bwd: scratch too small");
    cudaStream_t st = (cudaStream_t)stream;
    float* delta = (float*)scratch;
#define ATT_BWD(T_) \
    if (D == 32) launch_attn_bwd<T_, 1>(qkv, out, dout, lse, dqkv, delta, B, T, H, st); \
    else if (D == 64) launch_attn_bwd<T_, 2>(qkv, out, dout, lse, dqkv, delta, B, T, H, st); \
    else launch_attn_bwd<T_, 4>(qkv, out, dout, lse, dqkv, delta, B, T, H, st)
    if (dtype == F32) { ATT_BWD(float); }
    else if (dtype == BF16) { ATT_BWD(__nv_bfloat16); }
    else { set_error("attn_bwd: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
#undef ATT_BWD
    count_launch(2);
    return check_launch("attn_bwd_simt");
}
