// int8 KV cache (TURBO_QUANT_KV_CACHE=1): quantise-on-append and dequantise-inside-attention.
//
// Reference behaviour being replaced (kv_cache.py:101-195):
//   _quantize:   scale = amax(|x|, dim=-1) / 127 per cached token and head, 1 where that is 0;
//                codes = (x / scale).round().clamp(-128, 127).to(int8)                          (:119-125)
//   append:      torch.cat of codes and scales per layer per token, then _dequantize of the WHOLE cache,
//                codes.float() * scale, handed to SDPA                                           (:137-168, :135)
// Here codes live in a page pool int8 [n_pages][2][H][PAGE][D] with scales [n_pages][2][H][PAGE] beside it (the block
// table is the one the exact cache uses); an append writes only the new tokens and decode attention reads each cached
// byte once, multiplying by the token's scale on the fly -- nothing is ever dequantised back to HBM.
//
// Arithmetic follows torch's kernels on CUDA tensors step by step so codes and scales are bit-equal to the torch class
// (checked with torch.equal): `amax / 127.0` on a CUDA tensor is a multiplication by the fp32 reciprocal of 127
// (ATen's division-by-a-scalar path), `x / scale` is a true division, round() is to-nearest-even; for bf16 tensors every
// step is computed in fp32 and rounded to bf16, as ATen does. recip_scale = 0 selects the true division torch performs on
// CPU tensors instead (the reference service decodes on the CPU).
#include "common.cuh"

namespace penroz {

constexpr int KVQ_PAGE = 64;

template <typename T> __device__ __forceinline__ float round_to(float x);
template <> __device__ __forceinline__ float round_to<float>(float x) { return x; }
template <> __device__ __forceinline__ float round_to<__nv_bfloat16>(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// One warp per (row = b*S + s, k|v, head): D <= 256 elements, lane l holds elements l, l+32, ...
template <typename T>
__global__ void __launch_bounds__(256)
kv_append_q8_kernel(T* __restrict__ qkv, int8_t* __restrict__ codes, T* __restrict__ scales,
                    const int* __restrict__ bt, int max_pages, int B, int S, int H, int D, int t0, int writeback,
                    int recip_scale) {
    const int lane = threadIdx.x & 31;
    const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const size_t total = (size_t)B * S * 2 * H;
    if (warp >= total) return;
    const int h = (int)(warp % H);
    const int kv = (int)((warp / H) % 2);
    const size_t row = warp / (2 * (size_t)H);
    const int s = (int)(row % S), b = (int)(row / S);
    const int C = H * D;
    T* src = qkv + row * 3 * C + (size_t)(1 + kv) * C + (size_t)h * D;
    float x[8];
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int d = lane + 32 * i;
        x[i] = d < D ? to_f32(src[d]) : 0.f;
        amax = fmaxf(amax, fabsf(x[i]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    float scale = recip_scale ? round_to<T>(__fmul_rn(amax, 1.0f / 127.0f)) : round_to<T>(__fdiv_rn(amax, 127.0f));
    if (scale == 0.f) scale = 1.f;
    const int pos = t0 + s;
    const int page = bt[(size_t)b * max_pages + pos / KVQ_PAGE];
    const size_t slot = (((size_t)page * 2 + kv) * H + h) * KVQ_PAGE + (pos % KVQ_PAGE);
    int8_t* dst = codes + slot * D;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int d = lane + 32 * i;
        if (d < D) {
            const float q = fminf(fmaxf(rintf(round_to<T>(__fdiv_rn(x[i], scale))), -128.f), 127.f);
            dst[d] = (int8_t)(int)q;
            if (writeback) src[d] = from_f32<T>(__fmul_rn(q, scale));     // codes.float() * scale
        }
    }
    if (lane == 0) scales[slot] = from_f32<T>(scale);
}

// partial attention of one query row against keys [k0, k1) of its sequence, K/V dequantised on the fly;
// part layout [B][H][splits][D + 2]
template <typename T>
__global__ void __launch_bounds__(128)
attn_decode_q8_split_kernel(const T* __restrict__ qkv, const int8_t* __restrict__ codes, const T* __restrict__ scales,
                            const int* __restrict__ bt, int max_pages, float* __restrict__ part, int H, int D,
                            int t_len, int chunk, float scale) {
    extern __shared__ float sm[];  // q[D] | s[chunk] | vscale[chunk]
    __shared__ float red[32];
    float* qs = sm;
    float* ss = sm + D;
    float* vs = ss + chunk;
    const int split = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int C = H * D;
    const int k0 = split * chunk, k1 = min(t_len, k0 + chunk);
    for (int d = threadIdx.x; d < D; d += blockDim.x) qs[d] = to_f32(qkv[(size_t)b * 3 * C + h * D + d]) * scale;
    __syncthreads();
    float lm = -INFINITY;
    for (int j = k0 + threadIdx.x; j < k1; j += blockDim.x) {
        const int page = bt[(size_t)b * max_pages + j / KVQ_PAGE];
        const size_t slot_k = (((size_t)page * 2 + 0) * H + h) * KVQ_PAGE + (j % KVQ_PAGE);
        const size_t slot_v = (((size_t)page * 2 + 1) * H + h) * KVQ_PAGE + (j % KVQ_PAGE);
        const float sk = to_f32(scales[slot_k]);
        vs[j - k0] = to_f32(scales[slot_v]);
        const int8_t* kr = codes + slot_k * D;
        float a = 0.f;
        for (int d = 0; d < D; d += 16) {
            const int4 v = *reinterpret_cast<const int4*>(kr + d);
            const int w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float kd = __fmul_rn((float)(int8_t)((w[u] >> (8 * k)) & 0xff), sk);   // codes.float() * scale
                    a = fmaf(qs[d + 4 * u + k], kd, a);
                }
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
    const float l = block_sum(ls, red);  // also orders the ss[] / vs[] writes before the reads below
    float* out = part + (((size_t)b * H + h) * gridDim.x + split) * (D + 2);
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
        float o = 0.f;
        for (int j = k0; j < k1; ++j) {
            const int page = bt[(size_t)b * max_pages + j / KVQ_PAGE];
            const size_t slot_v = (((size_t)page * 2 + 1) * H + h) * KVQ_PAGE + (j % KVQ_PAGE);
            o = fmaf(ss[j - k0], __fmul_rn((float)codes[slot_v * D + d], vs[j - k0]), o);
        }
        out[d] = o;
    }
    if (threadIdx.x == 0) {
        out[D] = (k1 > k0) ? m : -INFINITY;
        out[D + 1] = (k1 > k0) ? l : 0.f;
    }
}

template <typename T>
__global__ void attn_decode_q8_combine_kernel(const float* __restrict__ part, T* __restrict__ out, int H, int D,
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

static int q8_splits(int B, int H, int t_len) {
    int want = (2 * num_sms() + B * H - 1) / (B * H);
    int maxs = (t_len + 63) / 64;
    int s = want < 1 ? 1 : want;
    if (s > maxs) s = maxs;
    if (s > 32) s = 32;
    return s < 1 ? 1 : s;
}

}  // namespace penroz

using namespace penroz;

extern "C" int penroz_kv_append_q8(void* qkv, void* codes, void* scales, const int* block_table, int max_pages, int B,
                                   int S, int H, int D, int t0, int dtype, int writeback, int recip_scale,
                                   void* stream) {
    PENROZ_REQUIRE(B > 0 && S > 0 && H > 0 && D > 0 && t0 >= 0, "kv_append_q8: bad shape");
    PENROZ_REQUIRE(D <= 256 && D % 16 == 0, "kv_append_q8: head_dim %d unsupported (multiple of 16, at most 256)", D);
    PENROZ_REQUIRE(qkv && codes && scales && block_table, "kv_append_q8: null argument");
    PENROZ_REQUIRE((t0 + S + KVQ_PAGE - 1) / KVQ_PAGE <= max_pages, "kv_append_q8: position %d exceeds the block table (%d pages)", t0 + S, max_pages);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t warps = (size_t)B * S * 2 * H;
    const size_t grid = (warps + 7) / 8;
    PENROZ_REQUIRE(grid < (1ull << 31), "kv_append_q8: too many rows");
    if (dtype == F32) kv_append_q8_kernel<float><<<(unsigned)grid, 256, 0, st>>>((float*)qkv, (int8_t*)codes, (float*)scales, block_table, max_pages, B, S, H, D, t0, writeback, recip_scale);
    else if (dtype == BF16) kv_append_q8_kernel<__nv_bfloat16><<<(unsigned)grid, 256, 0, st>>>((__nv_bfloat16*)qkv, (int8_t*)codes, (__nv_bfloat16*)scales, block_table, max_pages, B, S, H, D, t0, writeback, recip_scale);
    else { set_error("kv_append_q8: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("kv_append_q8");
}

extern "C" size_t penroz_attn_decode_q8_scratch_bytes(int B, int H, int D) {
    return (size_t)B * H * 32 * (size_t)(D + 2) * sizeof(float);
}

extern "C" int penroz_attn_decode_q8(const void* qkv, const void* codes, const void* scales, const int* block_table,
                                     int max_pages, void* out, int B, int H, int D, int t_len, int dtype,
                                     void* scratch, size_t scratch_bytes, void* stream) {
    PENROZ_REQUIRE(B > 0 && H > 0 && D > 0 && t_len > 0, "attn_decode_q8: bad shape");
    PENROZ_REQUIRE(D % 16 == 0 && D <= 256, "attn_decode_q8: head_dim %d unsupported", D);
    PENROZ_REQUIRE(B <= 65535 && H <= 65535, "attn_decode_q8: B/H too large for the launch grid");
    PENROZ_REQUIRE(scratch && scratch_bytes >= penroz_attn_decode_q8_scratch_bytes(B, H, D), "attn_decode_q8: scratch too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int splits = q8_splits(B, H, t_len);
    const int chunk = (t_len + splits - 1) / splits;
    const size_t smem = (size_t)(D + 2 * chunk) * sizeof(float);
    PENROZ_REQUIRE(smem <= 48 * 1024, "attn_decode_q8: context chunk %d too long for shared memory", chunk);
    const float scale = 1.0f / sqrtf((float)D);
    dim3 g(splits, H, B), gc(H, B);
    if (dtype == F32) {
        attn_decode_q8_split_kernel<float><<<g, 128, smem, st>>>((const float*)qkv, (const int8_t*)codes, (const float*)scales, block_table, max_pages, (float*)scratch, H, D, t_len, chunk, scale);
        attn_decode_q8_combine_kernel<float><<<gc, 64, 0, st>>>((const float*)scratch, (float*)out, H, D, splits);
    } else if (dtype == BF16) {
        attn_decode_q8_split_kernel<__nv_bfloat16><<<g, 128, smem, st>>>((const __nv_bfloat16*)qkv, (const int8_t*)codes, (const __nv_bfloat16*)scales, block_table, max_pages, (float*)scratch, H, D, t_len, chunk, scale);
        attn_decode_q8_combine_kernel<__nv_bfloat16><<<gc, 64, 0, st>>>((const float*)scratch, (__nv_bfloat16*)out, H, D, splits);
    } else { set_error("attn_decode_q8: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch(2);
    return check_launch("attn_decode_q8");
}
