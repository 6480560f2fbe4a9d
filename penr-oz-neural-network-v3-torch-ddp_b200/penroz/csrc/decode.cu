// KV-cache decode step: paged K/V append, split-KV single-query attention, skinny (few-row) Linear, argmax.
//
// Reference behaviour being replaced:
//   KVCache.append: torch.cat([old,new], dim=2) per layer per token  -> O(t) recopy  (kv_cache.py:41-68)
//   CausalSelfAttention cache branch: SDPA with S_q = 1, is_causal=False            (neural_net_layers.py:83-92)
//   logits[:, -1, :].argmax(dim=-1)                                                 (neural_net_model.py:393-394)
// Here K/V live in a page pool [n_pages][2][H][PAGE][D]; a per-sequence block table maps position -> page, so an
// append writes only the new token (O(1)) and attention streams each cached byte exactly once. All HBM-bound.
#include "common.cuh"
#include <stdlib.h>
#include <string.h>

namespace penroz {

// decode_fused.cu: split-KV attention with the cluster merge; returns 1 when the shape is not covered
int attn_decode_cluster(const void* qkv, const void* pool, const int* block_table, int max_pages, void* out, int B,
                        int H, int D, int t_len, int dtype, cudaStream_t st);

constexpr int KV_PAGE = 64;

template <typename T>
__device__ __forceinline__ const T* page_row(const T* pool, const int* bt, int max_pages, int b, int kv, int H, int h,
                                             int D, int pos) {
    const int page = bt[(size_t)b * max_pages + pos / KV_PAGE];
    return pool + ((((size_t)page * 2 + kv) * H + h) * KV_PAGE + (pos % KV_PAGE)) * D;
}

// qkv rows [B*S, 3C] -> pool at positions t0 .. t0+S-1
template <typename T>
__global__ void kv_append_kernel(const T* __restrict__ qkv, T* __restrict__ pool, const int* __restrict__ bt,
                                 int max_pages, int B, int S, int H, int D, int t0) {
    const int C = H * D;
    const size_t total = (size_t)B * S * 2 * C;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(e % C);
        const int kv = (int)((e / C) % 2);
        const int s = (int)((e / (2 * (size_t)C)) % S);
        const int b = (int)(e / (2 * (size_t)C * S));
        const int h = c / D, d = c % D;
        const T v = qkv[((size_t)b * S + s) * 3 * C + (size_t)(1 + kv) * C + c];
        T* dst = const_cast<T*>(page_row(pool, bt, max_pages, b, kv, H, h, D, t0 + s)) + d;
        *dst = v;
    }
}

// same, 16 bytes per thread: a thread moves VN consecutive head-dimension elements of one (row, k|v, head)
template <typename T>
__global__ void kv_append_vec_kernel(const T* __restrict__ qkv, T* __restrict__ pool, const int* __restrict__ bt,
                                     int max_pages, int B, int S, int H, int D, int t0) {
    constexpr int VN = Vec16<T>::N;
    const int C = H * D, dv = D / VN, cv = C / VN;
    const size_t total = (size_t)B * S * 2 * cv;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x) {
        const int c = (int)(e % cv);
        const int kv = (int)((e / cv) % 2);
        const size_t row = e / (2 * (size_t)cv);          // b * S + s
        const int s = (int)(row % S), b = (int)(row / S);
        const int h = c / dv, d = (c - h * dv) * VN;
        Vec16<T> v;
        v.load(qkv + row * 3 * C + (size_t)(1 + kv) * C + (size_t)c * VN);
        v.store(const_cast<T*>(page_row(pool, bt, max_pages, b, kv, H, h, D, t0 + s)) + d);
    }
}

// partial attention of one query row against keys [k0, k1) of its sequence; part layout [B][H][splits][D + 2]
template <typename T>
__global__ void __launch_bounds__(128)
attn_decode_split_kernel(const T* __restrict__ qkv, const T* __restrict__ pool, const int* __restrict__ bt,
                         int max_pages, float* __restrict__ part, int H, int D, int t_len, int chunk, float scale) {
    extern __shared__ float sm[];  // q[D] | s[chunk]
    __shared__ float red[32];
    float* qs = sm;
    float* ss = sm + D;
    const int split = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int C = H * D;
    const int k0 = split * chunk, k1 = min(t_len, k0 + chunk);
    for (int d = threadIdx.x; d < D; d += blockDim.x) qs[d] = to_f32(qkv[(size_t)b * 3 * C + h * D + d]) * scale;
    __syncthreads();
    float lm = -INFINITY;
    for (int j = k0 + threadIdx.x; j < k1; j += blockDim.x) {
        const T* kr = page_row(pool, bt, max_pages, b, 0, H, h, D, j);
        float a = 0.f;
        for (int d = 0; d < D; d += Vec16<T>::N) {
            Vec16<T> v;
            v.load(kr + d);
#pragma unroll
            for (int k = 0; k < Vec16<T>::N; ++k) a = fmaf(qs[d + k], v.get(k), a);
        }
        ss[j - k0] = a;
        lm = fmaxf(lm, a);
    }
    const float m = block_max(lm, red);
    float ls = 0.f;
    for (int j = k0 + threadIdx.x; j < k1; j += blockDim.x) {
        const float p = __expf(ss[j - k0] - m);
        ss[j - k0] = p;
        ls += p;
    }
    const float l = block_sum(ls, red);  // also orders the ss[] writes before the reads below
    float* out = part + (((size_t)b * H + h) * gridDim.x + split) * (D + 2);
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float o = 0.f;
        for (int j = k0; j < k1; ++j) {
            const T* vr = page_row(pool, bt, max_pages, b, 1, H, h, D, j);
            o = fmaf(ss[j - k0], to_f32(vr[d]), o);
        }
        out[d] = o;
    }
    if (threadIdx.x == 0) {
        out[D] = (k1 > k0) ? m : -INFINITY;
        out[D + 1] = (k1 > k0) ? l : 0.f;
    }
}

template <typename T>
__global__ void attn_decode_combine_kernel(const float* __restrict__ part, T* __restrict__ out, int H, int D,
                                           int splits) {
    const int h = blockIdx.x, b = blockIdx.y;
    const float* p = part + ((size_t)b * H + h) * splits * (D + 2);
    float m = -INFINITY;
    for (int s = 0; s < splits; ++s) m = fmaxf(m, p[(size_t)s * (D + 2) + D]);
    float l = 0.f;
    for (int s = 0; s < splits; ++s) {
        const float ms = p[(size_t)s * (D + 2) + D];
        if (ms != -INFINITY) l += p[(size_t)s * (D + 2) + D + 1] * __expf(ms - m);
    }
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float o = 0.f;
        for (int s = 0; s < splits; ++s) {
            const float ms = p[(size_t)s * (D + 2) + D];
            if (ms != -INFINITY) o += p[(size_t)s * (D + 2) + d] * __expf(ms - m);
        }
        out[((size_t)b * H + h) * D + d] = from_f32<T>(o / l);
    }
}

// Skinny Linear for decode: out[m,n] = epi(sum_k x[m,k] W[n,k] + bias[n]), M <= 8 rows per pass.
// One warp per output column, W row streamed with 16-byte loads, x rows staged in smem as fp32.
template <typename T, int MB>
__global__ void __launch_bounds__(256)
skinny_linear_kernel(const T* __restrict__ x, long long ldx, const T* __restrict__ W, long long ldw,
                     const T* __restrict__ bias, T* __restrict__ out, long long ldo, const T* __restrict__ aux,
                     long long ldaux, float* __restrict__ out_f32, long long ldo32, int M, int N, int K, int epi,
                     int gelu_tanh_) {
    extern __shared__ float xs[];  // [MB][K]
    const int m0 = blockIdx.y * MB;
    const int mrows = min(MB, M - m0);
    for (int e = threadIdx.x; e < MB * K; e += blockDim.x) {
        const int r = e / K, k = e % K;
        xs[e] = (r < mrows) ? to_f32(x[(size_t)(m0 + r) * ldx + k]) : 0.f;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n = blockIdx.x * (blockDim.x >> 5) + warp;
    if (n >= N) return;
    constexpr int VN = Vec16<T>::N;
    float acc[MB];
#pragma unroll
    for (int r = 0; r < MB; ++r) acc[r] = 0.f;
    const T* wr = W + (size_t)n * ldw;
    const bool vec = (K % VN == 0) && (ldw % VN == 0) && (((uintptr_t)W) % 16 == 0);
    if (vec) {
        for (int k = lane * VN; k < K; k += 32 * VN) {
            Vec16<T> w;
            w.load(wr + k);
#pragma unroll
            for (int i = 0; i < VN; i += 4) {
                const float w0 = w.get(i), w1 = w.get(i + 1), w2 = w.get(i + 2), w3 = w.get(i + 3);
#pragma unroll
                for (int r = 0; r < MB; ++r) {
                    const float4 xv = *reinterpret_cast<const float4*>(xs + r * K + k + i);
                    acc[r] = fmaf(xv.x, w0, acc[r]); acc[r] = fmaf(xv.y, w1, acc[r]);
                    acc[r] = fmaf(xv.z, w2, acc[r]); acc[r] = fmaf(xv.w, w3, acc[r]);
                }
            }
        }
    } else {
        for (int k = lane; k < K; k += 32) {
            const float wv = to_f32(wr[k]);
#pragma unroll
            for (int r = 0; r < MB; ++r) acc[r] = fmaf(xs[r * K + k], wv, acc[r]);
        }
    }
#pragma unroll
    for (int r = 0; r < MB; ++r) acc[r] = warp_sum(acc[r]);
    if (lane == 0) {
        const float bv = bias ? to_f32(bias[n]) : 0.f;
        for (int r = 0; r < mrows; ++r) {
            const int m = m0 + r;
            // rounding points follow torch: Linear output rounded to T, then GELU / residual add rounded again
            float v = to_f32(from_f32<T>(acc[r] + bv));
            if (epi == 1) v = gelu_tanh_ ? gelu_tanh(v) : gelu_erf(v);
            else if (epi == 2) v = to_f32(aux[(size_t)m * ldaux + n]) + v;
            if (out) out[(size_t)m * ldo + n] = from_f32<T>(v);
            if (out_f32) out_f32[(size_t)m * ldo32 + n] = to_f32(from_f32<T>(v));
        }
    }
}

// first index of the row maximum (torch.argmax tie rule), one CTA per row
template <typename T>
__global__ void argmax_rows_kernel(const T* __restrict__ x, long long ld, int V, long long* __restrict__ out) {
    __shared__ float sv[32];
    __shared__ int si[32];
    const T* r = x + (size_t)blockIdx.x * ld;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
        const float v = to_f32(r[i]);
        if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { sv[w] = best; si[w] = bi; }
    __syncthreads();
    if (w == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        best = lane < nw ? sv[lane] : -INFINITY;
        bi = lane < nw ? si[lane] : 0x7fffffff;
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        if (lane == 0) out[blockIdx.x] = (bi == 0x7fffffff) ? 0 : bi;
    }
}

}  // namespace penroz

using namespace penroz;

extern "C" int penroz_kv_page_size(void) { return KV_PAGE; }

extern "C" int penroz_kv_append(const void* qkv, void* pool, const int* block_table, int max_pages, int B, int S,
                                int H, int D, int t0, int dtype, void* stream) {
    PENROZ_REQUIRE(B > 0 && S > 0 && H > 0 && D > 0 && t0 >= 0, "kv_append: bad shape");
    PENROZ_REQUIRE((t0 + S + KV_PAGE - 1) / KV_PAGE <= max_pages, "kv_append: position %d exceeds the block table (%d pages)", t0 + S, max_pages);
    cudaStream_t st = (cudaStream_t)stream;
    const int vn = dtype == F32 ? 4 : 8;
    const bool vec = (D % vn == 0) && (((uintptr_t)qkv | (uintptr_t)pool) % 16 == 0);
    const size_t total = (size_t)B * S * 2 * H * D / (vec ? vn : 1);
    size_t g = (total + 255) / 256;
    const size_t cap = (size_t)num_sms() * 16;
    const int grid = (int)(g > cap ? cap : g);
    if (vec && dtype == F32) kv_append_vec_kernel<float><<<grid, 256, 0, st>>>((const float*)qkv, (float*)pool, block_table, max_pages, B, S, H, D, t0);
    else if (vec && dtype == BF16) kv_append_vec_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)qkv, (__nv_bfloat16*)pool, block_table, max_pages, B, S, H, D, t0);
    else if (dtype == F32) kv_append_kernel<float><<<grid, 256, 0, st>>>((const float*)qkv, (float*)pool, block_table, max_pages, B, S, H, D, t0);
    else if (dtype == BF16) kv_append_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)qkv, (__nv_bfloat16*)pool, block_table, max_pages, B, S, H, D, t0);
    else { set_error("kv_append: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("kv_append");
}

static int decode_splits(int B, int H, int t_len) {
    int want = (2 * num_sms() + B * H - 1) / (B * H);
    int maxs = (t_len + 63) / 64;
    int s = want < 1 ? 1 : want;
    if (s > maxs) s = maxs;
    if (s > 32) s = 32;
    return s < 1 ? 1 : s;
}
extern "C" size_t penroz_attn_decode_scratch_bytes(int B, int H, int D, int t_len) {
    return (size_t)B * H * 32 * (size_t)(D + 2) * sizeof(float);
    (void)t_len;
}

extern "C" int penroz_attn_decode(const void* qkv, const void* pool, const int* block_table, int max_pages, void* out,
                                  int B, int H, int D, int t_len, int dtype, void* scratch, size_t scratch_bytes,
                                  void* stream) {
    PENROZ_REQUIRE(B > 0 && H > 0 && D > 0 && t_len > 0, "attn_decode: bad shape");
    PENROZ_REQUIRE(D % 8 == 0 && D <= 256, "attn_decode: head_dim %d unsupported", D);
    PENROZ_REQUIRE(B <= 65535 && H <= 65535, "attn_decode: B/H too large for the launch grid");
    PENROZ_REQUIRE(scratch && scratch_bytes >= penroz_attn_decode_scratch_bytes(B, H, D, t_len), "attn_decode: scratch too small");
    cudaStream_t st = (cudaStream_t)stream;
    if (!getenv("PENROZ_DECODE_ATTN") || strcmp(getenv("PENROZ_DECODE_ATTN"), "split")) {
        // one thread-block cluster per head, partials merged through distributed shared memory (decode_fused.cu)
        const int rc = attn_decode_cluster(qkv, pool, block_table, max_pages, out, B, H, D, t_len, dtype, st);
        if (rc <= 0) return rc;     // done or failed; 1 = shape not covered, use the two-kernel form below
    }
    const int splits = decode_splits(B, H, t_len);
    const int chunk = (t_len + splits - 1) / splits;
    const size_t smem = (size_t)(D + chunk) * sizeof(float);
    PENROZ_REQUIRE(smem <= 48 * 1024, "attn_decode: context chunk %d too long for shared memory", chunk);
    const float scale = 1.0f / sqrtf((float)D);
    dim3 g(splits, H, B), gc(H, B);
    if (dtype == F32) {
        attn_decode_split_kernel<float><<<g, 128, smem, st>>>((const float*)qkv, (const float*)pool, block_table, max_pages, (float*)scratch, H, D, t_len, chunk, scale);
        attn_decode_combine_kernel<float><<<gc, 64, 0, st>>>((const float*)scratch, (float*)out, H, D, splits);
    } else if (dtype == BF16) {
        attn_decode_split_kernel<__nv_bfloat16><<<g, 128, smem, st>>>((const __nv_bfloat16*)qkv, (const __nv_bfloat16*)pool, block_table, max_pages, (float*)scratch, H, D, t_len, chunk, scale);
        attn_decode_combine_kernel<__nv_bfloat16><<<gc, 64, 0, st>>>((const float*)scratch, (__nv_bfloat16*)out, H, D, splits);
    } else { set_error("attn_decode: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch(2);
    return check_launch("attn_decode");
}

extern "C" int penroz_skinny_linear(const void* x, long long ldx, const void* W, long long ldw, const void* bias,
                                    void* out, long long ldo, const void* aux, long long ldaux, float* out_f32,
                                    long long ldo32, int M, int N, int K, int epilogue, int flags, int dtype,
                                    void* stream) {
    PENROZ_REQUIRE(M > 0 && N > 0 && K > 0, "skinny_linear: bad shape");
    PENROZ_REQUIRE(epilogue >= 0 && epilogue <= 2, "skinny_linear: bad epilogue %d", epilogue);
    PENROZ_REQUIRE(out || out_f32, "skinny_linear: no output");
    if (epilogue == 2) PENROZ_REQUIRE(aux != nullptr, "skinny_linear: residual epilogue needs aux");
    constexpr int MB = 8;
    const size_t smem = (size_t)MB * K * sizeof(float);
    PENROZ_REQUIRE(smem <= 200 * 1024, "skinny_linear: K=%d too large", K);
    cudaStream_t st = (cudaStream_t)stream;
    dim3 grid((N + 7) / 8, (M + MB - 1) / MB);
    PENROZ_REQUIRE(grid.y <= 65535, "skinny_linear: too many rows");
    if (dtype == F32) {
        auto k = skinny_linear_kernel<float, MB>;
        if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, 256, smem, st>>>((const float*)x, ldx, (const float*)W, ldw, (const float*)bias, (float*)out, ldo, (const float*)aux, ldaux, out_f32, ldo32, M, N, K, epilogue, flags & 1);
    } else if (dtype == BF16) {
        auto k = skinny_linear_kernel<__nv_bfloat16, MB>;
        if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k<<<grid, 256, smem, st>>>((const __nv_bfloat16*)x, ldx, (const __nv_bfloat16*)W, ldw, (const __nv_bfloat16*)bias, (__nv_bfloat16*)out, ldo, (const __nv_bfloat16*)aux, ldaux, out_f32, ldo32, M, N, K, epilogue, flags & 1);
    } else { set_error("skinny_linear: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("skinny_linear");
}

extern "C" int penroz_argmax_rows(const void* x, long long ld, int rows, int V, long long* out, int dtype,
                                  void* stream) {
    PENROZ_REQUIRE(rows > 0 && V > 0 && x && out, "argmax_rows: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == F32) argmax_rows_kernel<float><<<rows, 512, 0, st>>>((const float*)x, ld, V, out);
    else if (dtype == BF16) argmax_rows_kernel<__nv_bfloat16><<<rows, 512, 0, st>>>((const __nv_bfloat16*)x, ld, V, out);
    else { set_error("argmax_rows: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("argmax_rows");
}

// ------------------------------------------------------------------ whole single-token step, enqueued from C++
// The decode step of a GPT-2 tree is ~9 kernels per block; driving them one ctypes call at a time costs more host
// time than the GPU needs for the token (B = 1). This entry point enqueues the whole step — embedding, every
// block (LN, QKV, paged K/V append, split-KV attention, proj + residual, LN, fc + GELU, proj2 + residual), final LN,
// LM head, argmax — with one host call, using the same kernels as the individual entry points.
// Replaces one iteration of the loop at neural_net_model.py:465-469 / 499-503 (incremental branch :382-386).
extern "C" {
int penroz_embed_fwd(const long long*, const void*, const void*, void*, int, int, int, int, int, int, int, int*, void*);
int penroz_layernorm_fwd(const void*, const void*, const void*, void*, float*, float*, int, int, float, int, int, int, void*);
int penroz_gemm_bf16(const void*, long long, int, const void*, long long, int, int, int, int, int, const void*, void*,
                     long long, void*, long long, const void*, long long, int, void*);
int penroz_gemm_f32(const float*, long long, int, const float*, long long, int, int, int, int, int, const float*, float*,
                    long long, float*, long long, const float*, long long, int, void*);
}

struct penroz_gpt2_layer_t {
    const void *ln1_w, *ln1_b, *qkv_w, *qkv_b, *proj_w, *proj_b, *ln2_w, *ln2_b, *fc_w, *fc_b, *proj2_w, *proj2_b;
    void* kv_pool;
    float ln1_eps, ln2_eps;
    int gelu_tanh, ffn;
    void* kv_scales;             // int8 cache: per-token scales beside the int8 kv_pool (NULL: exact cache)
    int kv_quant, kv_recip;
};
extern "C" {
int penroz_kv_append_q8(void*, void*, void*, const int*, int, int, int, int, int, int, int, int, int, void*);
int penroz_attn_decode_q8(const void*, const void*, const void*, const int*, int, void*, int, int, int, int, int, void*,
                          size_t, void*);
}

namespace penroz {
// y = epi(x W^T + b): skinny kernel for few rows, tensor-core / FP32-pipe GEMM otherwise. epi: 0 store, 1 gelu, 2 resid
static int decode_linear(const void* x, const void* W, const void* b, void* out, const void* aux, int M, int N, int K,
                         int epi, int tanh_, int dtype, void* scratch_pre, void* stream) {
    static const int bf16_rows = getenv("PENROZ_DECODE_SKINNY_ROWS") ? atoi(getenv("PENROZ_DECODE_SKINNY_ROWS")) : 8;
    const int skinny_rows = (dtype == F32) ? 64 : bf16_rows;
    if (M <= skinny_rows)
        return penroz_skinny_linear(x, K, W, K, b, out, N, aux, N, nullptr, 0, M, N, K, epi, tanh_, dtype, stream);
    if (dtype == BF16) {
        if (epi == 1) return penroz_gemm_bf16(x, K, 0, W, K, 0, M, N, K, 1, b, scratch_pre, N, out, N, nullptr, 0, tanh_, stream);
        return penroz_gemm_bf16(x, K, 0, W, K, 0, M, N, K, epi == 2 ? 3 : 0, b, out, N, nullptr, 0, aux, N, 0, stream);
    }
    if (epi == 1) return penroz_gemm_f32((const float*)x, K, 0, (const float*)W, K, 0, M, N, K, 1, (const float*)b, (float*)scratch_pre, N, (float*)out, N, nullptr, 0, tanh_, stream);
    return penroz_gemm_f32((const float*)x, K, 0, (const float*)W, K, 0, M, N, K, epi == 2 ? 2 : 0, (const float*)b, (float*)out, N, nullptr, 0, (const float*)aux, N, 0, stream);
}
}  // namespace penroz

namespace penroz {   // decode_fused.cu: the five-launches-per-block step for small batches
bool decode_small_supported(int B, int C, int H, int dtype);
size_t decode_small_extra_bytes(int B);
int decode_step_small(const long long* tokens, const void* wte, const void* wpe, const penroz_gpt2_layer_t* layers,
                      int n_layer, const void* lnf_w, const void* lnf_b, float lnf_eps, const void* head_w,
                      const void* head_b, const int* block_table, int max_pages, int B, int C, int H, int V, int pos,
                      int dtype, void* xa, void* xb, void* qkv, void* att, void* g, float* part, void* extra,
                      void* logits_out, long long* argmax_out, int* err_flag, cudaStream_t st);
}

namespace penroz {   // decode_mega.cu: the whole step as one persistent cooperative kernel (bf16, batch <= 64)
bool decode_mega_supported(int B, int C, int H, int V, int F, int n_layer, int dtype);
size_t decode_mega_extra_bytes(int B, int H);
int decode_step_mega(const long long* tokens, const void* wte, const void* wpe, const penroz_gpt2_layer_t* layers,
                     int n_layer, const void* lnf_w, const void* lnf_b, float lnf_eps, const void* head_w,
                     const void* head_b, const int* block_table, int max_pages, int B, int C, int H, int V, int pos,
                     void* xa, void* xb, void* qkv, void* att, void* g, float* part, float* amax_val, int* amax_idx,
                     unsigned* sync, void* logits_out, long long* argmax_out, int* err_flag, cudaStream_t st);
}

extern "C" size_t penroz_gpt2_decode_workspace_bytes(int B, int C, int H, int F, int t_len_max) {
    const size_t es = 4;  // sized for fp32; bf16 uses half of the activation part
    const size_t act = (size_t)B * ((size_t)2 * C + C + 3 * C + C + 2 * (size_t)F + C) * es + 8 * 256;
    return decode_mega_extra_bytes(B, H) + act + penroz_attn_decode_scratch_bytes(B, H, C / H, t_len_max) + 256 +
           decode_small_extra_bytes(B);
}

extern "C" int penroz_gpt2_decode_step(const long long* tokens, const void* wte, const void* wpe,
                                       const penroz_gpt2_layer_t* layers, int n_layer, const void* lnf_w,
                                       const void* lnf_b, float lnf_eps, const void* head_w, const void* head_b,
                                       const int* block_table, int max_pages, int B, int C, int H, int V, int pos,
                                       int dtype, void* workspace, size_t workspace_bytes, void* logits_out,
                                       long long* argmax_out, int* err_flag, void* stream) {
    PENROZ_REQUIRE(B > 0 && C > 0 && H > 0 && V > 0 && n_layer > 0 && pos >= 0, "gpt2_decode_step: bad shape");
    PENROZ_REQUIRE(C % H == 0, "gpt2_decode_step: n_embd %d not divisible by %d heads", C, H);
    PENROZ_REQUIRE(dtype == F32 || dtype == BF16, "gpt2_decode_step: bad dtype %d", dtype);
    int F = 0;
    for (int l = 0; l < n_layer; ++l) F = layers[l].ffn > F ? layers[l].ffn : F;
    PENROZ_REQUIRE(workspace && workspace_bytes >= penroz_gpt2_decode_workspace_bytes(B, C, H, F, pos + 1),
                   "gpt2_decode_step: workspace too small");
    const int D = C / H;
    const size_t es = dtype == F32 ? 4 : 2;
    uint8_t* w = (uint8_t*)workspace;
    unsigned* sync_words = (unsigned*)w;          // zero when the workspace is handed over, left zero by every step
    w += decode_mega_extra_bytes(B, H);
    auto carve = [&](size_t elems) { void* p = w; w += ((elems * es + 255) / 256) * 256; return p; };
    void* xa = carve((size_t)B * C);
    void* xb = carve((size_t)B * C);
    void* ln = carve((size_t)B * C);
    void* qkv = carve((size_t)B * 3 * C);
    void* att = carve((size_t)B * C);
    void* g = carve((size_t)B * F);
    void* hpre = carve((size_t)B * F);
    void* attn_scratch = w;
    const size_t attn_scratch_bytes = penroz_attn_decode_scratch_bytes(B, H, D, pos + 1);
    const bool q8 = layers[0].kv_quant != 0;
    if (!q8 && !getenv("PENROZ_DECODE_UNFUSED") && decode_mega_supported(B, C, H, V, F, n_layer, dtype)) {
        uint8_t* extra = w + ((attn_scratch_bytes + 255) / 256) * 256;
        float* amax_val = (float*)extra;
        int* amax_idx = (int*)(amax_val + (size_t)B * 16384);
        return decode_step_mega(tokens, wte, wpe, layers, n_layer, lnf_w, lnf_b, lnf_eps, head_w, head_b, block_table,
                                max_pages, B, C, H, V, pos, xa, xb, qkv, att, g, (float*)attn_scratch, amax_val,
                                amax_idx, sync_words, logits_out, argmax_out, err_flag, (cudaStream_t)stream);
    }      // int8 cache: one kernel per op, quantise-on-append + dequantising attention
    if (!q8 && decode_small_supported(B, C, H, dtype) && !getenv("PENROZ_DECODE_UNFUSED")) {
        void* extra = w + ((attn_scratch_bytes + 255) / 256) * 256;
        return decode_step_small(tokens, wte, wpe, layers, n_layer, lnf_w, lnf_b, lnf_eps, head_w, head_b, block_table,
                                 max_pages, B, C, H, V, pos, dtype, xa, xb, qkv, att, g, (float*)attn_scratch, extra,
                                 logits_out, argmax_out, err_flag, (cudaStream_t)stream);
    }
    int rc = penroz_embed_fwd(tokens, wte, wpe, xa, B, 1, C, V, pos, dtype, dtype, err_flag, stream);
    if (rc) return rc;
    void *x = xa, *xo = xb;
    for (int l = 0; l < n_layer; ++l) {
        const penroz_gpt2_layer_t& L = layers[l];
        if ((rc = penroz_layernorm_fwd(x, L.ln1_w, L.ln1_b, ln, nullptr, nullptr, B, C, L.ln1_eps, dtype, dtype, dtype, stream))) return rc;
        if ((rc = decode_linear(ln, L.qkv_w, L.qkv_b, qkv, nullptr, B, 3 * C, C, 0, 0, dtype, nullptr, stream))) return rc;
        if (q8) {
            if ((rc = penroz_kv_append_q8(qkv, L.kv_pool, L.kv_scales, block_table, max_pages, B, 1, H, D, pos, dtype, 0, L.kv_recip, stream))) return rc;
            if ((rc = penroz_attn_decode_q8(qkv, L.kv_pool, L.kv_scales, block_table, max_pages, att, B, H, D, pos + 1, dtype, attn_scratch, attn_scratch_bytes, stream))) return rc;
        } else {
            if ((rc = penroz_kv_append(qkv, L.kv_pool, block_table, max_pages, B, 1, H, D, pos, dtype, stream))) return rc;
            if ((rc = penroz_attn_decode(qkv, L.kv_pool, block_table, max_pages, att, B, H, D, pos + 1, dtype, attn_scratch, attn_scratch_bytes, stream))) return rc;
        }
        if ((rc = decode_linear(att, L.proj_w, L.proj_b, xo, x, B, C, C, 2, 0, dtype, nullptr, stream))) return rc;
        if ((rc = penroz_layernorm_fwd(xo, L.ln2_w, L.ln2_b, ln, nullptr, nullptr, B, C, L.ln2_eps, dtype, dtype, dtype, stream))) return rc;
        if ((rc = decode_linear(ln, L.fc_w, L.fc_b, g, nullptr, B, L.ffn, C, 1, L.gelu_tanh, dtype, hpre, stream))) return rc;
        if ((rc = decode_linear(g, L.proj2_w, L.proj2_b, x, xo, B, C, L.ffn, 2, 0, dtype, nullptr, stream))) return rc;
    }
    if ((rc = penroz_layernorm_fwd(x, lnf_w, lnf_b, ln, nullptr, nullptr, B, C, lnf_eps, dtype, dtype, dtype, stream))) return rc;
    if ((rc = decode_linear(ln, head_w, head_b, logits_out, nullptr, B, V, C, 0, 0, dtype, nullptr, stream))) return rc;
    if (argmax_out) rc = penroz_argmax_rows(logits_out, V, B, V, argmax_out, dtype, stream);
    return rc;
}

// ------------------------------------------------------------------ whole prompt pass (prefill), enqueued from C++
// The first decode iteration runs the full prompt through the model with an empty cache (neural_net_model.py:387-390,
// 250-258 with the cache branch of neural_net_layers.py:83-92 appending S tokens) and keeps only the last position's
// logits. Same kernels as the per-op entry points -- tensor-core GEMMs, tcgen05 flash attention, paged K/V append for
// all S positions -- enqueued with ONE host call; LayerNorm + LM head run on the B last rows only.
extern "C" {
int penroz_attn_fwd_simt(const void*, void*, float*, int, int, int, int, int, void*);
int penroz_attn_fwd_tc(const void*, void*, float*, int, int, int, int, void*);
}

extern "C" size_t penroz_gpt2_prefill_workspace_bytes(int B, int S, int C, int F) {
    const size_t M = (size_t)B * S;
    // x, x', LN out, qkv, attention out, GELU out, pre-activation, last rows (fp32-sized; bf16 uses half)
    return (M * ((size_t)3 * C + 3 * C + C + 2 * (size_t)F) + (size_t)B * C) * 4 + 10 * 256;
}

extern "C" int penroz_gpt2_prefill(const long long* tokens, const void* wte, const void* wpe,
                                   const penroz_gpt2_layer_t* layers, int n_layer, const void* lnf_w, const void* lnf_b,
                                   float lnf_eps, const void* head_w, const void* head_b, const int* block_table,
                                   int max_pages, int B, int S, int C, int H, int V, int dtype, void* workspace,
                                   size_t workspace_bytes, void* logits_out, long long* argmax_out, int* err_flag,
                                   void* stream) {
    PENROZ_REQUIRE(B > 0 && S > 0 && C > 0 && H > 0 && V > 0 && n_layer > 0, "gpt2_prefill: bad shape");
    PENROZ_REQUIRE(C % H == 0, "gpt2_prefill: n_embd %d not divisible by %d heads", C, H);
    PENROZ_REQUIRE(dtype == F32 || dtype == BF16, "gpt2_prefill: bad dtype %d", dtype);
    PENROZ_REQUIRE((long long)B * S < (1LL << 31), "gpt2_prefill: too many rows");
    int F = 0;
    for (int l = 0; l < n_layer; ++l) F = layers[l].ffn > F ? layers[l].ffn : F;
    PENROZ_REQUIRE(workspace && workspace_bytes >= penroz_gpt2_prefill_workspace_bytes(B, S, C, F),
                   "gpt2_prefill: workspace too small");
    const int D = C / H, M = B * S;
    const size_t es = dtype == F32 ? 4 : 2;
    uint8_t* w = (uint8_t*)workspace;
    auto carve = [&](size_t elems) { void* p = w; w += ((elems * es + 255) / 256) * 256; return p; };
    void* xa = carve((size_t)M * C);
    void* xb = carve((size_t)M * C);
    void* ln = carve((size_t)M * C);
    void* qkv = carve((size_t)M * 3 * C);
    void* att = carve((size_t)M * C);
    void* g = carve((size_t)M * F);
    void* hpre = carve((size_t)M * F);
    void* last = carve((size_t)B * C);
    const bool tc_attn = (dtype == BF16) && D == 64 && !(getenv("PENROZ_ATTENTION") && !strcmp(getenv("PENROZ_ATTENTION"), "simt"));
    int rc = penroz_embed_fwd(tokens, wte, wpe, xa, B, S, C, V, 0, dtype, dtype, err_flag, stream);
    if (rc) return rc;
    void *x = xa, *xo = xb;
    for (int l = 0; l < n_layer; ++l) {
        const penroz_gpt2_layer_t& L = layers[l];
        if ((rc = penroz_layernorm_fwd(x, L.ln1_w, L.ln1_b, ln, nullptr, nullptr, M, C, L.ln1_eps, dtype, dtype, dtype, stream))) return rc;
        if ((rc = decode_linear(ln, L.qkv_w, L.qkv_b, qkv, nullptr, M, 3 * C, C, 0, 0, dtype, nullptr, stream))) return rc;
        // int8 cache: the prompt's own attention already sees the de-quantised K/V (KVCache.append returns them,
        // kv_cache.py:165-168), so the append writes codes.float() * scale back over the K and V columns of qkv
        if (L.kv_quant) rc = penroz_kv_append_q8(qkv, L.kv_pool, L.kv_scales, block_table, max_pages, B, S, H, D, 0, dtype, 1, L.kv_recip, stream);
        else rc = penroz_kv_append(qkv, L.kv_pool, block_table, max_pages, B, S, H, D, 0, dtype, stream);
        if (rc) return rc;
        if (tc_attn) rc = penroz_attn_fwd_tc(qkv, att, nullptr, B, S, H, D, stream);
        else rc = penroz_attn_fwd_simt(qkv, att, nullptr, B, S, H, D, dtype, stream);
        if (rc) return rc;
        if ((rc = decode_linear(att, L.proj_w, L.proj_b, xo, x, M, C, C, 2, 0, dtype, nullptr, stream))) return rc;
        if ((rc = penroz_layernorm_fwd(xo, L.ln2_w, L.ln2_b, ln, nullptr, nullptr, M, C, L.ln2_eps, dtype, dtype, dtype, stream))) return rc;
        if ((rc = decode_linear(ln, L.fc_w, L.fc_b, g, nullptr, M, L.ffn, C, 1, L.gelu_tanh, dtype, hpre, stream))) return rc;
        if ((rc = decode_linear(g, L.proj2_w, L.proj2_b, x, xo, M, C, L.ffn, 2, 0, dtype, nullptr, stream))) return rc;
    }
    // final LayerNorm and LM head on the last position of every sequence only
    for (int b = 0; b < B; ++b) {
        const uint8_t* src = (const uint8_t*)x + ((size_t)b * S + (S - 1)) * C * es;
        if ((rc = penroz_layernorm_fwd(src, lnf_w, lnf_b, (uint8_t*)last + (size_t)b * C * es, nullptr, nullptr, 1, C, lnf_eps, dtype, dtype, dtype, stream))) return rc;
    }
    if ((rc = decode_linear(last, head_w, head_b, logits_out, nullptr, B, V, C, 0, 0, dtype, nullptr, stream))) return rc;
    if (argmax_out) rc = penroz_argmax_rows(logits_out, V, B, V, argmax_out, dtype, stream);
    return rc;
}
