// Causal flash attention on the 5th-gen tensor cores (bf16, head_dim 64), forward and backward, reading q/k/v
// straight from the packed [B*T, 3C] QKV buffer through ONE 2-D TMA tensor map and writing [B*T, C].
//
// Replaces F.scaled_dot_product_attention(q,k,v,is_causal=True) plus the view/transpose/contiguous copies around it
// (neural_net_layers.py:59-95, GPT-2 branch) for the bf16-autocast training step.
//
// Forward, one CTA per (128-query tile, head, batch), two CTAs resident per SM so one tile's softmax overlaps the
// other's MMAs:
//     warp 4, lane 0 : TMA loads (Q once; K_j, V_j double buffered) and tcgen05.mma issue
//                      S   = Q K_j^T      (M128 N128 K64, both operands K-major)                -> TMEM cols [0,128)
//                      O_j = P V_j        (M128 N64 K128, A = P from smem, B = V MN-major)      -> TMEM cols [128,192)
//     warps 0-3      : one thread per query row (TMEM lane): tcgen05.ld the row of S, online softmax in registers
//                      (no shuffles: a thread owns its whole row), P -> bf16 -> smem in the 128B-swizzled K-major
//                      layout the MMA reads, then O <- O*corr + O_j in registers.
// Backward, one CTA per (128-key tile, head, batch) looping over query tiles i >= j, five MMAs per step
// (S, dP, dV += P^T dO, dK += dS^T Q, dQ_i = dS K) with dV/dK accumulating in TMEM and dQ reduced across key tiles
// with fp32 red.add into a scratch buffer, converted to bf16 by a small finishing kernel.
#include "common.cuh"

namespace penroz {

constexpr int AT_BM = 128;   // query rows per tile
constexpr int AT_BN = 128;   // keys per tile
constexpr int AT_D = 64;

__device__ __forceinline__ uint64_t at_desc(uint32_t saddr, uint32_t lbo_enc, uint32_t sbo_enc) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)lbo_enc << 16) | ((uint64_t)sbo_enc << 32) |
           (1ull << 46) | (2ull << 61);
}
// idesc: fp32 accum, bf16 x bf16, majors, N, M=128
__device__ __forceinline__ uint32_t at_idesc(int a_mn, int b_mn, int N) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// byte offset of element (row, col) in a [128 rows][64 bf16] tile with the TMA/UMMA 128-byte swizzle
__device__ __forceinline__ uint32_t sw128_off(int row, int col) {
    return (uint32_t)(row * 128 + ((((col >> 3) ^ (row & 7)) << 4)) + ((col & 7) << 1));
}

// ---------------------------------------------------------------------------------------------- forward
struct AttnFwdSmem {
    static constexpr int kQ = 0;                     // 16 KB
    static constexpr int kK0 = 16384;                // 2 stages x 16 KB
    static constexpr int kV0 = kK0 + 2 * 16384;      // 2 stages x 16 KB
    static constexpr int kP = kV0 + 2 * 16384;       // 32 KB: [2 key blocks][128 rows][64]
    static constexpr int kBar = kP + 32768;
    static constexpr int kBytes = kBar + 128 + 1024;  // + alignment slack
};

__global__ void __launch_bounds__(160, 2)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, __nv_bfloat16* __restrict__ out,
                   float* __restrict__ lse, int Tn, int H, float scale_log2e) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
    uint8_t* smem = smem_raw + pad;
    const uint32_t sbase = raw + pad;
    const uint32_t bar = sbase + AttnFwdSmem::kBar;
    const uint32_t q_full = bar, kv_full0 = bar + 8, kv_empty0 = bar + 24, s_full = bar + 40, p_ready = bar + 48,
                   o_full = bar + 56;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + AttnFwdSmem::kBar + 64);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_qt = (Tn + AT_BM - 1) / AT_BM;
    const int qt = n_qt - 1 - (int)blockIdx.x;       // heaviest tiles first
    const int h = blockIdx.y, b = blockIdx.z;
    const int C = H * AT_D;
    const int row0 = b * Tn + qt * AT_BM;            // first row of this query tile in the [B*T, 3C] matrix
    const int n_kv = qt + 1;

    if (threadIdx.x == 128) {
        tma_prefetch_desc(&tmQKV);
        mbar_init(q_full, 1);
        mbar_init(kv_full0, 1); mbar_init(kv_full0 + 8, 1);
        mbar_init(kv_empty0, 1); mbar_init(kv_empty0 + 8, 1);
        mbar_init(s_full, 1);
        mbar_init(p_ready, 4);
        mbar_init(o_full, 1);
        mbar_fence_init();
    }
    if (warp == 0) {
        tmem_alloc(smem_u32(tmem_slot), 256);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t tS = tmem, tO = tmem + 128;

    if (warp == 4) {
        if (lane == 0) {
            const uint32_t sQ = sbase + AttnFwdSmem::kQ;
            const uint32_t sP = sbase + AttnFwdSmem::kP;
            mbar_expect_tx(q_full, 16384);
            tma_load_2d(sQ, &tmQKV, q_full, h * AT_D, row0);
            mbar_expect_tx(kv_full0, 32768);
            tma_load_2d(sbase + AttnFwdSmem::kK0, &tmQKV, kv_full0, C + h * AT_D, b * Tn);
            tma_load_2d(sbase + AttnFwdSmem::kV0, &tmQKV, kv_full0, 2 * C + h * AT_D, b * Tn);
            const uint32_t idS = at_idesc(0, 0, AT_BN);
            const uint32_t idO = at_idesc(0, 1, AT_D);
            mbar_wait(q_full, 0);
            for (int j = 0; j < n_kv; ++j) {
                const int s = j & 1;
                const uint32_t sK = sbase + AttnFwdSmem::kK0 + s * 16384;
                const uint32_t sV = sbase + AttnFwdSmem::kV0 + s * 16384;
                mbar_wait(kv_full0 + 8 * s, (uint32_t)((j >> 1) & 1));
                tc_fence_after();
#pragma unroll
                for (int k4 = 0; k4 < AT_D / 16; ++k4)
                    umma_bf16(tS, at_desc(sQ + k4 * 32, 1, 64), at_desc(sK + k4 * 32, 1, 64), idS, (uint32_t)(k4 != 0));
                umma_commit(s_full);
                if (j + 1 < n_kv) {
                    const int s2 = s ^ 1;
                    if (j >= 1) mbar_wait(kv_empty0 + 8 * s2, (uint32_t)(((j - 1) >> 1) & 1));
                    mbar_expect_tx(kv_full0 + 8 * s2, 32768);
                    tma_load_2d(sbase + AttnFwdSmem::kK0 + s2 * 16384, &tmQKV, kv_full0 + 8 * s2, C + h * AT_D,
                                b * Tn + (j + 1) * AT_BN);
                    tma_load_2d(sbase + AttnFwdSmem::kV0 + s2 * 16384, &tmQKV, kv_full0 + 8 * s2, 2 * C + h * AT_D,
                                b * Tn + (j + 1) * AT_BN);
                }
                mbar_wait(p_ready, (uint32_t)(j & 1));
                tc_fence_after();
                // O_j = P V_j : A = P K-major (two 64-key blocks 16 KB apart), B = V MN-major (16 keys = 2048 B per step)
#pragma unroll
                for (int k8 = 0; k8 < AT_BN / 16; ++k8) {
                    const uint32_t pa = sP + (k8 >> 2) * 16384 + (k8 & 3) * 32;
                    umma_bf16(tO, at_desc(pa, 1, 64), at_desc(sV + k8 * 2048, 512, 64), idO, (uint32_t)(k8 != 0));
                }
                umma_commit(o_full);
                umma_commit(kv_empty0 + 8 * s);
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------ softmax / output threads: one query row each
        const int r = warp * 32 + lane;
        const int i = qt * AT_BM + r;                      // query position inside the sequence
        const uint32_t lane_addr = (uint32_t)(warp * 32) << 16;
        float o[AT_D];
#pragma unroll
        for (int d = 0; d < AT_D; ++d) o[d] = 0.f;
        float m = -INFINITY, l = 0.f;
        uint8_t* sPp = smem + AttnFwdSmem::kP;
        for (int j = 0; j < n_kv; ++j) {
            mbar_wait(s_full, (uint32_t)(j & 1));
            tc_fence_after();
            const bool diag = (j == qt);
            const int key0 = j * AT_BN;
            // pass 1: row maximum
            float mt = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tS + lane_addr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int k = 0; k < 32; ++k) {
                    float sv = __uint_as_float(v[k]);
                    if (diag && key0 + c * 32 + k > i) sv = -INFINITY;
                    mt = fmaxf(mt, sv);
                }
            }
            const float mn = fmaxf(m, mt * scale_log2e);   // scale > 0: max commutes with the scaling
            const float corr = exp2f(m - mn);              // m = -inf on the first tile -> 0
            float ls = 0.f;
            // pass 2: p = 2^(s*scale*log2e - mn), row sum, P -> smem (bf16, swizzled K-major)
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tS + lane_addr + c * 32, v);
                tmem_ld_wait();
                uint32_t packed[16];
#pragma unroll
                for (int k = 0; k < 32; k += 2) {
                    float a = __uint_as_float(v[k]) * scale_log2e - mn;
                    float bb = __uint_as_float(v[k + 1]) * scale_log2e - mn;
                    float pa = (diag && key0 + c * 32 + k > i) ? 0.f : exp2f(a);
                    float pb = (diag && key0 + c * 32 + k + 1 > i) ? 0.f : exp2f(bb);
                    // the row sum uses the bf16-rounded probabilities the PV product will see
                    __nv_bfloat162 hb = __floats2bfloat162_rn(pa, pb);
                    float2 back = __bfloat1622float2(hb);
                    ls += back.x + back.y;
                    packed[k >> 1] = *reinterpret_cast<uint32_t*>(&hb);
                }
                // 32 keys = 4 chunks of 16 B inside key block (c >> 1), columns (c & 1) * 32 ..
                uint8_t* blk = sPp + (c >> 1) * 16384;
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                    const int col = (c & 1) * 32 + q4 * 8;
                    *reinterpret_cast<uint4*>(blk + sw128_off(r, col)) =
                        make_uint4(packed[4 * q4], packed[4 * q4 + 1], packed[4 * q4 + 2], packed[4 * q4 + 3]);
                }
            }
            tc_fence_before();
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_ready);
            l = l * corr + ls;
            m = mn;
            mbar_wait(o_full, (uint32_t)(j & 1));
            tc_fence_after();
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tO + lane_addr + c * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int k = 0; k < 32; ++k) o[c * 32 + k] = o[c * 32 + k] * corr + __uint_as_float(v[k]);
            }
            tc_fence_before();
        }
        if (i < Tn) {
            const float inv = 1.f / l;
            __nv_bfloat16* op = out + ((size_t)(b * Tn + i)) * C + h * AT_D;
#pragma unroll
            for (int d = 0; d < AT_D; d += 8) {
                uint4 w;
                w.x = pack_bf16x2(o[d] * inv, o[d + 1] * inv);
                w.y = pack_bf16x2(o[d + 2] * inv, o[d + 3] * inv);
                w.z = pack_bf16x2(o[d + 4] * inv, o[d + 5] * inv);
                w.w = pack_bf16x2(o[d + 6] * inv, o[d + 7] * inv);
                *reinterpret_cast<uint4*>(op + d) = w;
            }
            if (lse) lse[((size_t)b * H + h) * Tn + i] = (m + log2f(l)) * 0.69314718055994530942f;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFnA)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static int encode_rows_bf16(CUtensorMap* tm, const void* ptr, uint64_t cols, uint64_t rows, uint32_t box_rows) {
    static std::atomic<void*> cached{nullptr};
    void* sym = cached.load(std::memory_order_acquire);
    if (!sym) {
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
            qres != cudaDriverEntryPointSuccess || !sym) {
            set_error("cuTensorMapEncodeTiled entry point unavailable");
            return PENROZ_ERR_CUDA;
        }
        cached.store(sym, std::memory_order_release);
    }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 2};
    cuuint32_t box[2] = {64, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = ((EncodeTiledFnA)sym)(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides,
                                       box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                       CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("attention: cuTensorMapEncodeTiled failed (%d)", (int)r); return PENROZ_ERR_CUDA; }
    return PENROZ_OK;
}

}  // namespace penroz

using namespace penroz;

extern "C" int penroz_attn_fwd_tc(const void* qkv, void* out, float* lse, int B, int T, int H, int D, void* stream) {
    PENROZ_REQUIRE(B > 0 && T > 0 && H > 0, "attn_fwd_tc: bad shape");
    PENROZ_REQUIRE(D == 64, "attn_fwd_tc: head_dim %d unsupported (tensor-core path is head_dim 64)", D);
    PENROZ_REQUIRE(B <= 65535 && H <= 65535, "attn_fwd_tc: B/H too large for the launch grid");
    PENROZ_REQUIRE(((uintptr_t)qkv % 16 == 0) && ((uintptr_t)out % 16 == 0), "attn_fwd_tc: buffers must be 16-byte aligned");
    CUtensorMap tm;
    int rc = encode_rows_bf16(&tm, qkv, (uint64_t)3 * H * D, (uint64_t)B * T, 128);
    if (rc) return rc;
    static std::atomic<int> configured{0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!((configured.load() >> dev) & 1)) {
        cudaError_t e = cudaFuncSetAttribute(attn_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnFwdSmem::kBytes);
        if (e != cudaSuccess) { set_error("attn_fwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
        configured.fetch_or(1 << dev);
    }
    dim3 grid((T + AT_BM - 1) / AT_BM, H, B);
    const float scale_log2e = (1.0f / sqrtf((float)D)) * 1.44269504088896340736f;
    attn_fwd_tc_kernel<<<grid, 160, AttnFwdSmem::kBytes, (cudaStream_t)stream>>>(tm, (__nv_bfloat16*)out, lse, T, H, scale_log2e);
    count_launch();
    return check_launch("attn_fwd_tc");
}
