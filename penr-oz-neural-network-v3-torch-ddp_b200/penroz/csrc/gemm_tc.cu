// bf16 GEMM on the 5th-gen tensor cores: D[M,N] = sum_k A(m,k) * B(n,k), fp32 accumulate in TMEM.
//
// Replaces the cuBLASLt calls torch makes for nn.Linear forward/backward on the reference's
// hot path (reference: mappers.py:21,156,158,163,165,172 build the nn.Linear modules;
// neural_net_model.py:250-258 runs them; .backward() at neural_net_model.py:650-653).
//
// Design (B200 / sm_100a):
//   * persistent, warp-specialised CTA (384 threads, one CTA per SM):
//       warp 0  TMA producer   (cp.async.bulk.tensor 2D, 128B swizzle, mbarrier complete_tx)
//       warp 1  MMA issuer     (one elected lane issues tcgen05.mma cta_group::1 kind::f16, M=128 N=BN K=16)
//       warp 2  TMEM allocator
//       warps 4-11 epilogue    (tcgen05.ld 32x32b -> registers -> fused epilogue -> global; lane quarter = warp % 4,
//                               column half = (warp - 4) / 4)
//   * smem ring of kStages x {A 128x64, B BNx64} bf16 tiles, full/empty mbarriers
//   * TMEM double buffered (2 x BN fp32 columns) so the epilogue of tile i overlaps the MMAs of tile i+1
//   * both operands may be K-major (row-major [rows,K]) or MN-major (row-major [K,rows]); MN-major is what
//     the weight-gradient (dW = dY^T X) and data-gradient (dX = dY W) contractions need, so no transposed
//     copies of activations or weights are ever materialised.
//   * split-K (fp32 red.add epilogue) for the weight-gradient shapes whose output has too few tiles for 148 SMs.
//   * work units are not dealt round robin: the producer draws them from a global counter (see "dynamic tile feed"
//     below), so a grid that loses SMs to a concurrent kernel (the gradient all-reduce) slows down in proportion
//     instead of waiting for its unlucky CTAs.
#include "common.cuh"
#include <stdlib.h>
#include <string.h>

namespace penroz {

constexpr int BM = 128;
constexpr int BK = 64;   // one 128-byte swizzle atom of bf16 along K
constexpr int UK = 16;   // K per tcgen05.mma for 16-bit inputs

enum Epi {
    EPI_BF16 = 0,        // out = bf16(acc + bias)
    EPI_GELU = 1,        // out = pre = bf16(acc + bias); out2 = bf16(gelu(pre))
    EPI_RESID_F32 = 2,   // out(f32) = aux(f32) + float(bf16(acc + bias))
    EPI_RESID_BF16 = 3,  // out(bf16) = bf16(float(aux bf16) + float(bf16(acc + bias)))
    EPI_ACCUM_F32 = 4,   // out(f32) (+)= acc        (beta flag; split-K uses red.add)
    EPI_DGELU = 5,       // out = bf16(float(bf16(acc)) * gelu'(aux bf16))
    EPI_F32 = 6,         // out(f32) = acc + bias
    EPI_ROWDOT = 7,      // out = bf16(acc + bias); out2(f32)[(r / T) * (N/64) + g][r % T] = sum_c out[r,64g+c] * aux[r,64g+c]
};

struct GemmParams {
    int M, N, K;
    int a_mn, b_mn;
    int tiles_m, tiles_n, splits, kb_per_split, kb_total;
    const __nv_bfloat16* bias;
    void* out;
    long long ldo;
    void* out2;
    long long ldo2;
    const void* aux;
    long long ldaux;
    int gelu_tanh;
    int beta;
    int feed_slot;   // which g_feed counter pair this launch draws its work units from
    int seq;         // EPI_ROWDOT: rows per sequence (T)
};

// ------------------------------------------------------------------ dynamic tile feed
// A persistent grid with a fixed round-robin schedule assumes it owns every SM for its whole duration. It does not:
// the gradient all-reduce (NCCL, on a side stream) takes SMs away while backward GEMMs run, and a CTA that cannot
// become resident leaves its whole share of tiles waiting until another CTA has finished ITS share -- the kernel
// then takes up to twice as long. Here every work unit, the first included, is drawn from a global counter by the
// CTA's (or CTA pair's) TMA producer, one unit ahead of use so the atomic's latency hides behind a tile's loads;
// a late CTA finds the counter exhausted and exits. The producer hands unit ids to the MMA and epilogue warps (and,
// in the pair kernel, to the odd CTA through distributed shared memory) over a small ring guarded by mbarriers.
// Counters live in a file-scope device array, one pair per launch in flight (the host deals them out round robin);
// the last CTA to leave resets its pair, so no launch ever needs a memset.
constexpr int kFeedSlots = 6;        // ring depth: ids are published a tile ahead of a producer that itself runs
                                     // up to ~2 tiles ahead of the epilogue
constexpr int kFeedCounters = 1024;
__device__ unsigned int g_feed[kFeedCounters][2];   // {next unit, CTAs done}

// ring layout at `base` (120 bytes): full[6] mbarriers | empty[6] mbarriers | unit id[6]
constexpr uint32_t kFeedOffset = 136u;   // behind the pipeline barriers (<= 128 B) and the TMEM pointer, inside 256 B
static_assert(kFeedOffset + 20u * kFeedSlots <= 256u, "tile-feed ring must fit the 256-byte barrier area");
__device__ __forceinline__ uint32_t feed_full(uint32_t base, int s) { return base + 8u * s; }
__device__ __forceinline__ uint32_t feed_empty(uint32_t base, int s) { return base + 8u * kFeedSlots + 8u * s; }
__device__ __forceinline__ uint32_t feed_unit(uint32_t base, int s) { return base + 16u * kFeedSlots + 4u * s; }

// Producer side, one lane: publish `u` (-1 = no more work) in ring slot `s`. PAIR: also into the odd CTA's ring.
template <bool PAIR>
__device__ __forceinline__ void feed_publish(uint32_t base, int s, uint32_t ph, int u) {
    mbar_wait(feed_empty(base, s), ph ^ 1u);
    st_shared_s32(feed_unit(base, s), u);
    mbar_arrive(feed_full(base, s));
    if (PAIR) st_async_s32(mapa_rank(feed_unit(base, s), 1), u, mapa_rank(feed_full(base, s), 1));
}

// Consumer side: a warp-uniform cursor over the ring. next() returns the unit id, or -1 when the feed is exhausted.
// In the odd CTA of a pair (`odd`) the slot's "empty" arrival goes to the even CTA's barrier, which the publisher
// waits on; the shuffle between the read and that arrival keeps the read ahead of any overwrite. One reader per odd
// CTA (`rearm`, its producer warp) re-arms each full barrier for the next st.async as soon as it has seen the
// current phase complete.
template <bool PAIR>
struct FeedCursor {
    uint32_t base;
    int s;
    uint32_t ph;
    bool odd, rearm;
    __device__ __forceinline__ explicit FeedCursor(uint32_t b, bool odd_cta = false, bool rearm_full = false)
        : base(b), s(0), ph(0), odd(odd_cta), rearm(rearm_full) {}
    __device__ __forceinline__ int next(int lane) {
        int u = 0;
        if (lane == 0) {
            mbar_wait(feed_full(base, s), ph);
            u = ld_shared_s32(feed_unit(base, s));
            if (PAIR && rearm) mbar_expect_tx(feed_full(base, s), 4);
        }
        u = __shfl_sync(0xffffffffu, u, 0);
        if (lane == 0) {
            if (PAIR && odd) mbar_arrive_on_leader(feed_empty(base, s)); else mbar_arrive(feed_empty(base, s));
        }
        if (++s == kFeedSlots) { s = 0; ph ^= 1u; }
        return u;
    }
};

// Called by a drawing thread right after the draw that came back exhausted (its last): the last of the launch's
// `drawers` to get here resets the counters for the slot's next user. Nobody draws after its own exhausted draw, so
// the reset races with nothing, and it sits behind the CTA's remaining tiles instead of at the kernel's tail.
__device__ __forceinline__ void feed_retire(int slot, unsigned int drawers) {
    const unsigned int done = atomicAdd(&g_feed[slot][1], 1u);
    if (done == drawers - 1) {
        g_feed[slot][0] = 0u;
        g_feed[slot][1] = 0u;
    }
}



// ------------------------------------------------------------------ epilogue helpers (one warp, one 32x32 chunk)
// stg: per-warp staging tile, 32 rows x 33 floats. "own row" access = stg[lane*33 + i]; "coalesced" access walks
// rows with 4 (bf16, 8 values) or 8 (fp32, 4 values) lanes per row.
template <typename T>
__device__ __forceinline__ void stage_load_rows(float* stg, const T* g, long long ld, int lane, int row0, int col0, int M);
template <>
__device__ __forceinline__ void stage_load_rows<float>(float* stg, const float* g, long long ld, int lane, int row0,
                                                       int col0, int M) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int r = 4 * k + (lane >> 3), c = (lane & 7) * 4;
        float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row0 + r < M) w = *reinterpret_cast<const float4*>(g + (long long)(row0 + r) * ld + col0 + c);
        float* d = stg + r * 33 + c;
        d[0] = w.x; d[1] = w.y; d[2] = w.z; d[3] = w.w;
    }
}
template <>
__device__ __forceinline__ void stage_load_rows<__nv_bfloat16>(float* stg, const __nv_bfloat16* g, long long ld,
                                                               int lane, int row0, int col0, int M) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int r = 8 * k + (lane >> 2), c = (lane & 3) * 8;
        uint4 w = make_uint4(0u, 0u, 0u, 0u);
        if (row0 + r < M) w = *reinterpret_cast<const uint4*>(g + (long long)(row0 + r) * ld + col0 + c);
        const float2 a = unpack_bf16x2(w.x), b = unpack_bf16x2(w.y), cc = unpack_bf16x2(w.z), d2 = unpack_bf16x2(w.w);
        float* d = stg + r * 33 + c;
        d[0] = a.x; d[1] = a.y; d[2] = b.x; d[3] = b.y; d[4] = cc.x; d[5] = cc.y; d[6] = d2.x; d[7] = d2.y;
    }
}
__device__ __forceinline__ void stage_store_rows_bf16(const float* stg, __nv_bfloat16* g, long long ld, int lane,
                                                      int row0, int col0, int M) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int r = 8 * k + (lane >> 2), c = (lane & 3) * 8;
        const float* d = stg + r * 33 + c;
        uint4 w;
        w.x = pack_bf16x2(d[0], d[1]); w.y = pack_bf16x2(d[2], d[3]);
        w.z = pack_bf16x2(d[4], d[5]); w.w = pack_bf16x2(d[6], d[7]);
        if (row0 + r < M) *reinterpret_cast<uint4*>(g + (long long)(row0 + r) * ld + col0 + c) = w;
    }
}
// mode 0 store, 1 load-add-store, 2 red.add
__device__ __forceinline__ void stage_store_rows_f32(const float* stg, float* g, long long ld, int lane, int row0,
                                                     int col0, int M, int mode) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int r = 4 * k + (lane >> 3), c = (lane & 7) * 4;
        const float* d = stg + r * 33 + c;
        if (row0 + r < M) {
            float* o = g + (long long)(row0 + r) * ld + col0 + c;
            if (mode == 2) {
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o), "f"(d[0]), "f"(d[1]), "f"(d[2]),
                             "f"(d[3])
                             : "memory");
            } else {
                float4 w = make_float4(d[0], d[1], d[2], d[3]);
                if (mode == 1) {
                    const float4 old = *reinterpret_cast<const float4*>(o);
                    w.x += old.x; w.y += old.y; w.z += old.z; w.w += old.w;
                }
                *reinterpret_cast<float4*>(o) = w;
            }
        }
    }
}

// aux operand of a chunk, fetched one chunk ahead into registers (row-coalesced mapping) and later dropped into the
// staging tile: fp32 aux uses 8 x 16 B per lane, bf16 aux 4 x 16 B.
template <int EPI>
__device__ __forceinline__ void aux_issue(const GemmParams& p, uint4 (&pre)[8], int lane, int row0, int col0) {
    if (EPI == EPI_RESID_F32) {
        const float* g = reinterpret_cast<const float*>(p.aux);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int r = 4 * k + (lane >> 3), c = (lane & 7) * 4;
            pre[k] = make_uint4(0u, 0u, 0u, 0u);
            if (row0 + r < p.M) pre[k] = *reinterpret_cast<const uint4*>(g + (long long)(row0 + r) * p.ldaux + col0 + c);
        }
    } else if (EPI == EPI_DGELU || EPI == EPI_RESID_BF16) {
        const __nv_bfloat16* g = reinterpret_cast<const __nv_bfloat16*>(p.aux);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = 8 * k + (lane >> 2), c = (lane & 3) * 8;
            pre[k] = make_uint4(0u, 0u, 0u, 0u);
            if (row0 + r < p.M) pre[k] = *reinterpret_cast<const uint4*>(g + (long long)(row0 + r) * p.ldaux + col0 + c);
        }
    }
}
template <int EPI>
__device__ __forceinline__ void aux_deposit(float* stg, const uint4 (&pre)[8], int lane) {
    if (EPI == EPI_RESID_F32) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int r = 4 * k + (lane >> 3), c = (lane & 7) * 4;
            float* d = stg + r * 33 + c;
            d[0] = __uint_as_float(pre[k].x); d[1] = __uint_as_float(pre[k].y);
            d[2] = __uint_as_float(pre[k].z); d[3] = __uint_as_float(pre[k].w);
        }
    } else if (EPI == EPI_DGELU || EPI == EPI_RESID_BF16) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int r = 8 * k + (lane >> 2), c = (lane & 3) * 8;
            const float2 a = unpack_bf16x2(pre[k].x), b = unpack_bf16x2(pre[k].y), cc = unpack_bf16x2(pre[k].z), d2 = unpack_bf16x2(pre[k].w);
            float* d = stg + r * 33 + c;
            d[0] = a.x; d[1] = a.y; d[2] = b.x; d[3] = b.y; d[4] = cc.x; d[5] = cc.y; d[6] = d2.x; d[7] = d2.y;
        }
    }
}
// can this chunk take the vector path? (full 32 columns and 16-byte aligned rows in every buffer involved)
template <int EPI>
__device__ __forceinline__ bool chunk_is_fast(const GemmParams& p, int col0) {
    bool fast = (col0 + 32 <= p.N);
    if (EPI == EPI_BF16 || EPI == EPI_GELU || EPI == EPI_DGELU || EPI == EPI_RESID_BF16) fast = fast && (p.ldo % 8 == 0);
    if (EPI == EPI_GELU) fast = fast && (p.ldo2 % 8 == 0);
    if (EPI == EPI_DGELU || EPI == EPI_RESID_BF16) fast = fast && (p.ldaux % 8 == 0);
    if (EPI == EPI_RESID_F32 || EPI == EPI_F32 || EPI == EPI_ACCUM_F32) fast = fast && (p.ldo % 4 == 0);
    if (EPI == EPI_RESID_F32) fast = fast && (p.ldaux % 4 == 0);
    return fast;
}

template <int EPI>
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, float (&acc)[32], float* stg, int lane, int row0,
                                               int col0, bool full, uint4 (&pre)[8], bool pre_valid, int next_col0) {
    const int row = row0 + lane;
    const bool row_ok = row < p.M;
    float* own = stg + lane * 33;
    // ---- bias (same 64 B for every lane: broadcast loads)
    if (EPI != EPI_ACCUM_F32 && EPI != EPI_DGELU && p.bias != nullptr) {
        if (full && ((col0 & 7) == 0) && ((reinterpret_cast<uintptr_t>(p.bias) & 15) == 0)) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 w = *reinterpret_cast<const uint4*>(p.bias + col0 + 8 * j);
                const float2 b0 = unpack_bf16x2(w.x), b1 = unpack_bf16x2(w.y), b2 = unpack_bf16x2(w.z), b3 = unpack_bf16x2(w.w);
                acc[8 * j + 0] += b0.x; acc[8 * j + 1] += b0.y; acc[8 * j + 2] += b1.x; acc[8 * j + 3] += b1.y;
                acc[8 * j + 4] += b2.x; acc[8 * j + 5] += b2.y; acc[8 * j + 6] += b3.x; acc[8 * j + 7] += b3.y;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (full || col0 + i < p.N) acc[i] += __bfloat162float(p.bias[col0 + i]);
        }
    }
    const bool fast = chunk_is_fast<EPI>(p, col0);
    constexpr bool kHasAux = (EPI == EPI_RESID_F32 || EPI == EPI_DGELU || EPI == EPI_RESID_BF16);

    if (fast) {
        float aux[32];
        if (kHasAux) {
            if (!pre_valid) aux_issue<EPI>(p, pre, lane, row0, col0);   // first chunk of a tile whose prefetch was skipped
            aux_deposit<EPI>(stg, pre, lane);
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 32; ++i) aux[i] = own[i];
            __syncwarp();
            // next chunk's aux goes in flight now and lands while this chunk is computed and stored
            if (next_col0 >= 0) aux_issue<EPI>(p, pre, lane, row0, next_col0);
        }
        if (EPI == EPI_GELU) {
            // pre-activation first, then the activation, through the same staging tile
#pragma unroll
            for (int i = 0; i < 32; ++i) own[i] = acc[i];
            __syncwarp();
            stage_store_rows_bf16(stg, reinterpret_cast<__nv_bfloat16*>(p.out), p.ldo, lane, row0, col0, p.M);
            __syncwarp();
            if (p.gelu_tanh) {   // warp-uniform choice outside the element loop: straight-line code per element
#pragma unroll
                for (int i = 0; i < 32; ++i) own[i] = gelu_tanh_fast(bf16_round(acc[i]));
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) own[i] = gelu_erf_fast(bf16_round(acc[i]));
            }
            __syncwarp();
            stage_store_rows_bf16(stg, reinterpret_cast<__nv_bfloat16*>(p.out2), p.ldo2, lane, row0, col0, p.M);
            __syncwarp();
            return;
        }
        if (EPI == EPI_DGELU) {
            if (p.gelu_tanh) {
#pragma unroll
                for (int i = 0; i < 32; ++i) own[i] = bf16_round(acc[i]) * gelu_tanh_grad_fast(aux[i]);
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) own[i] = bf16_round(acc[i]) * gelu_erf_grad_fast(aux[i]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                float r = acc[i];
                if (EPI == EPI_RESID_F32 || EPI == EPI_RESID_BF16) r = aux[i] + bf16_round(acc[i]);
                own[i] = r;
            }
        }
        __syncwarp();
        if (EPI == EPI_BF16 || EPI == EPI_DGELU || EPI == EPI_RESID_BF16)
            stage_store_rows_bf16(stg, reinterpret_cast<__nv_bfloat16*>(p.out), p.ldo, lane, row0, col0, p.M);
        else
            stage_store_rows_f32(stg, reinterpret_cast<float*>(p.out), p.ldo, lane, row0, col0, p.M,
                                 (EPI == EPI_ACCUM_F32) ? ((p.splits > 1) ? 2 : (p.beta ? 1 : 0)) : 0);
        __syncwarp();
        return;
    }
    // ---- edge chunks / unaligned leading dimensions: one thread per row, element by element
    if (!row_ok) return;
    if (EPI == EPI_BF16 || EPI == EPI_GELU || EPI == EPI_DGELU || EPI == EPI_RESID_BF16) {
        __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + (long long)row * p.ldo + col0;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            if (col0 + i >= p.N) continue;   // fully unrolled + predicated: acc[] stays in registers
            float r = acc[i];
            if (EPI == EPI_DGELU) {
                const float hv = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.aux)[(long long)row * p.ldaux + col0 + i]);
                r = bf16_round(r) * (p.gelu_tanh ? gelu_tanh_grad_fast(hv) : gelu_erf_grad_fast(hv));
            }
            if (EPI == EPI_RESID_BF16)
                r = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.aux)[(long long)row * p.ldaux + col0 + i]) + bf16_round(r);
            o[i] = __float2bfloat16_rn(r);
            if (EPI == EPI_GELU) {
                const float pre = bf16_round(acc[i]);
                reinterpret_cast<__nv_bfloat16*>(p.out2)[(long long)row * p.ldo2 + col0 + i] =
                    __float2bfloat16_rn(p.gelu_tanh ? gelu_tanh_fast(pre) : gelu_erf_fast(pre));
            }
        }
    } else {
        float* o = reinterpret_cast<float*>(p.out) + (long long)row * p.ldo + col0;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            if (col0 + i >= p.N) continue;
            if (EPI == EPI_RESID_F32) o[i] = reinterpret_cast<const float*>(p.aux)[(long long)row * p.ldaux + col0 + i] + bf16_round(acc[i]);
            else if (EPI == EPI_F32) o[i] = acc[i];
            else if (p.splits > 1) atomicAdd(o + i, acc[i]);
            else o[i] = p.beta ? o[i] + acc[i] : acc[i];
        }
    }
}

// ------------------------------------------------------------------ epilogue of the pair kernel: registers ->
// swizzled shared memory -> bulk tensor store. A thread owns one accumulator row (tcgen05.ld hands TMEM out that way):
// it packs its 32 columns of the chunk, writes them as 16-byte pieces into a [32 rows][32 columns] box laid out with
// the TMA swizzle (conflict-free), and one lane hands the box to the TMA unit (cp.async.bulk.tensor store, or
// cp.reduce ... .add for the accumulating weight-gradient form, which replaces a load-add-store round trip or 1024
// red.global per chunk). Rows and columns past M / N are clipped by the tensor map. The round-1 epilogue moved every
// element through an fp32 staging tile twice (33-float stride transpose) to get row-contiguous global accesses: 11
// instructions per element for a plain bf16 store and 38 with GELU, which made the GELU / GELU' GEMMs epilogue-bound
// (ncu: 120.8 M warp instructions against 26.5 M for the plain kernel, 201 us against ~125 us of mainloop).
template <int EPI> struct EpiTraits {
    static constexpr bool kF32Out = (EPI == EPI_RESID_F32 || EPI == EPI_ACCUM_F32 || EPI == EPI_F32);
    static constexpr bool kHasAux = (EPI == EPI_RESID_F32 || EPI == EPI_DGELU || EPI == EPI_RESID_BF16 || EPI == EPI_ROWDOT);
    static constexpr bool kAuxF32 = (EPI == EPI_RESID_F32);
    static constexpr int kBoxBytes = kF32Out ? 4096 : 2048;
};
constexpr int kStgPerWarp = 4096;   // one fp32 box, or two bf16 boxes (alternating; out / out2 for GELU)

__device__ __forceinline__ void tma_store_2d(const void* tmap, uint32_t smem_src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(tmap), "r"(smem_src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const void* tmap, uint32_t smem_src, int c0, int c1) {
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(tmap), "r"(smem_src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// this thread's row of a bf16 box (32 rows x 64 B, SWIZZLE_64B: 16-byte piece j of row r sits at j ^ ((r >> 1) & 3))
__device__ __forceinline__ void box_put_bf16(uint32_t box, int lane, const uint32_t (&w)[16]) {
    const uint32_t row = box + (uint32_t)lane * 64u, x = ((uint32_t)lane >> 1) & 3u;
#pragma unroll
    for (uint32_t j = 0; j < 4; ++j) sts128(row + ((j ^ x) << 4), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
}
// fp32 box (32 rows x 128 B, SWIZZLE_128B: piece j of row r at j ^ (r & 7))
__device__ __forceinline__ void box_put_f32(uint32_t box, int lane, const float (&a)[32]) {
    const uint32_t row = box + (uint32_t)lane * 128u, x = (uint32_t)lane & 7u;
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j)
        sts128(row + ((j ^ x) << 4), __float_as_uint(a[4 * j]), __float_as_uint(a[4 * j + 1]), __float_as_uint(a[4 * j + 2]),
               __float_as_uint(a[4 * j + 3]));
}
// hand a filled box to the TMA unit. PENDING = bulk groups of this lane that may still be reading shared memory when
// the box is written (the other box of an alternating pair).
template <int PENDING>
__device__ __forceinline__ void box_begin(int lane) {
    if (lane == 0) bulk_wait_read<PENDING>();
    __syncwarp();
}
__device__ __forceinline__ void box_send(const void* tmap, uint32_t box, int lane, int col0, int row0, bool reduce) {
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
        if (reduce) tma_reduce_add_2d(tmap, box, col0, row0); else tma_store_2d(tmap, box, col0, row0);
        bulk_commit();
    }
}

// Epilogues with an aux operand: the aux chunk sits in a swizzled box (bulk tensor load), is replaced in place by the
// result (a thread touches only its own row's bytes) and leaves by a bulk tensor store from the same box.
//   EPI_RESID_F32  fp32 box (SWIZZLE_128B): box += bf16(acc + bias)
//   EPI_RESID_BF16 bf16 box (SWIZZLE_64B):  box = bf16(box + bf16(acc + bias))
//   EPI_DGELU      bf16 box:                box = bf16(bf16(acc) * gelu'(box))
//   EPI_ROWDOT     bf16 box:                returns sum_c bf16(acc + bias)[c] * box[c], then box = bf16(acc + bias)
template <int EPI>
__device__ __forceinline__ float epilogue_chunk_box(const GemmParams& p, const uint32_t (&v)[32], uint32_t box, int lane,
                                                    int col0) {
    float dot = 0.f;
    float acc[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] = __uint_as_float(v[i]);
    if (EPI != EPI_DGELU && p.bias != nullptr) {
        if (col0 + 32 <= p.N && ((reinterpret_cast<uintptr_t>(p.bias + col0) & 15) == 0)) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 w = *reinterpret_cast<const uint4*>(p.bias + col0 + 8 * j);
                const uint32_t bw[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 b = unpack_bf16x2(bw[k]);
                    acc[8 * j + 2 * k] += b.x; acc[8 * j + 2 * k + 1] += b.y;
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (col0 + i < p.N) acc[i] += __bfloat162float(p.bias[col0 + i]);
        }
    }
    if (EPI == EPI_RESID_F32) {
        const uint32_t row = box + (uint32_t)lane * 128u, x = (uint32_t)lane & 7u;
#pragma unroll
        for (uint32_t j = 0; j < 8; ++j) {
            const uint32_t addr = row + ((j ^ x) << 4);
            float4 a;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "r"(addr) : "memory");
            sts128(addr, __float_as_uint(a.x + bf16_round(acc[4 * j])), __float_as_uint(a.y + bf16_round(acc[4 * j + 1])),
                   __float_as_uint(a.z + bf16_round(acc[4 * j + 2])), __float_as_uint(a.w + bf16_round(acc[4 * j + 3])));
        }
    } else {
        const uint32_t row = box + (uint32_t)lane * 64u, x = ((uint32_t)lane >> 1) & 3u;
        uint32_t h[16];
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j)
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(h[4 * j]), "=r"(h[4 * j + 1]), "=r"(h[4 * j + 2]), "=r"(h[4 * j + 3])
                         : "r"(row + ((j ^ x) << 4)) : "memory");
        uint32_t w[16];
        if (EPI == EPI_RESID_BF16) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 a = unpack_bf16x2(h[i]);
                w[i] = pack_bf16x2(a.x + bf16_round(acc[2 * i]), a.y + bf16_round(acc[2 * i + 1]));
            }
        } else if (EPI == EPI_ROWDOT) {
            float d0 = 0.f, d1 = 0.f;
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                w[i] = pack_bf16x2(acc[2 * i], acc[2 * i + 1]);
                const float2 a = unpack_bf16x2(h[i]), r = unpack_bf16x2(w[i]);
                d0 = fmaf(r.x, a.x, d0);
                d1 = fmaf(r.y, a.y, d1);
            }
            dot = d0 + d1;
        } else if (p.gelu_tanh) {   // warp-uniform choice outside the element loop: straight-line code per element
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 a = unpack_bf16x2(h[i]);
                w[i] = pack_bf16x2(bf16_round(acc[2 * i]) * gelu_tanh_grad_fast(a.x), bf16_round(acc[2 * i + 1]) * gelu_tanh_grad_fast(a.y));
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 a = unpack_bf16x2(h[i]);
                float lo, hi;
                gelu_erf_grad_times_fast2(a.x, a.y, bf16_round(acc[2 * i]), bf16_round(acc[2 * i + 1]), lo, hi);
                w[i] = pack_bf16x2(lo, hi);
            }
        }
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j) sts128(row + ((j ^ x) << 4), w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
        if (EPI == EPI_DGELU && p.out2 != nullptr) {
            // column sums of the chunk as its consumers see it (bf16), for the bias gradient of the Linear in front of
            // the activation: lane = column, down the 32 rows of the box (rows past M hold zeros: their A rows and
            // their aux rows were zero-filled by the loads), then one red.add per column
            __syncwarp();
            const uint32_t cb = box + (((uint32_t)lane & 7u) << 1), piece = (uint32_t)lane >> 3;
            float cs = 0.f;
#pragma unroll
            for (uint32_t r = 0; r < 32; ++r) {
                unsigned short hv;
                asm volatile("ld.shared.u16 %0, [%1];" : "=h"(hv) : "r"(cb + r * 64u + ((piece ^ ((r >> 1) & 3u)) << 4)) : "memory");
                cs += __uint_as_float((uint32_t)hv << 16);
            }
            if (col0 + lane < p.N) atomicAdd(reinterpret_cast<float*>(p.out2) + col0 + lane, cs);
        }
    }
    return dot;
}

// Epilogues without an aux operand (EPI_BF16, EPI_GELU, EPI_F32, EPI_ACCUM_F32): registers -> box -> store
template <int EPI>
__device__ __forceinline__ void epilogue_chunk_ts(const GemmParams& p, const void* tmO, const void* tmO2,
                                                  const uint32_t (&v)[32], uint32_t stg, int& nbox, int lane, int row0,
                                                  int col0) {
    using T = EpiTraits<EPI>;
    float acc[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] = __uint_as_float(v[i]);
    // ---- bias (same 64 B for every lane: broadcast loads)
    if (EPI != EPI_ACCUM_F32 && p.bias != nullptr) {
        if (col0 + 32 <= p.N && ((reinterpret_cast<uintptr_t>(p.bias + col0) & 15) == 0)) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint4 w = *reinterpret_cast<const uint4*>(p.bias + col0 + 8 * j);
                const uint32_t bw[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float2 b = unpack_bf16x2(bw[k]);
                    f2_unpack(f2_add(f2_pack(acc[8 * j + 2 * k], acc[8 * j + 2 * k + 1]), f2_pack(b.x, b.y)), acc[8 * j + 2 * k],
                              acc[8 * j + 2 * k + 1]);
                }
            }
        } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (col0 + i < p.N) acc[i] += __bfloat162float(p.bias[col0 + i]);
        }
    }
    if constexpr (T::kF32Out) {
        box_begin<0>(lane);
        box_put_f32(stg, lane, acc);
        box_send(tmO, stg, lane, col0, row0, EPI == EPI_ACCUM_F32 && (p.splits > 1 || p.beta));
    } else {
    uint32_t w[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) w[i] = pack_bf16x2(acc[2 * i], acc[2 * i + 1]);
    const uint32_t box = stg + (uint32_t)(nbox & 1) * 2048u;
    box_begin<1>(lane);
    box_put_bf16(box, lane, w);
    box_send(tmO, box, lane, col0, row0, false);
    ++nbox;
    if (EPI == EPI_GELU) {
        // the activation of the rounded pre-activation, through the other box
        if (p.gelu_tanh) {   // warp-uniform choice outside the element loop: straight-line code per element
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 r = unpack_bf16x2(w[i]);
                w[i] = pack_bf16x2(gelu_tanh_fast(r.x), gelu_tanh_fast(r.y));
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                const float2 r = unpack_bf16x2(w[i]);
                float lo, hi;
                gelu_erf_fast2(r.x, r.y, lo, hi);
                w[i] = pack_bf16x2(lo, hi);
            }
        }
        const uint32_t box2 = stg + (uint32_t)(nbox & 1) * 2048u;
        box_begin<1>(lane);
        box_put_bf16(box2, lane, w);
        box_send(tmO2, box2, lane, col0, row0, false);
        ++nbox;
    }
    }
}

template <int BN> struct GemmCfg {
    static constexpr int kABytes = BM * BK * 2;
    static constexpr int kBBytes = BN * BK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = (BN == 256) ? 4 : 6;
    static constexpr int kTmemCols = 2 * BN;  // 512 or 256: power of two >= 32
    static constexpr int kStagingBytes = 8 * 32 * 33 * 4;  // epilogue staging tiles, one per epilogue warp
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/ + kStagingBytes;
};

template <int BN, int EPI>
__global__ void __launch_bounds__(384, 1)
gemm_bf16_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const GemmParams p) {
    using Cfg = GemmCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    const uint32_t pad = (1024u - (raw_addr & 1023u)) & 1023u;
    uint8_t* smem = smem_raw + pad;
    const uint32_t smem_base = raw_addr + pad;
    const uint32_t bar_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
    // barrier slots (8 bytes each): full[kStages], empty[kStages], tmem_full[2], tmem_empty[2], then tmem ptr
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + 2 + s); };
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + Cfg::kStages * Cfg::kStageBytes + 8 * (2 * Cfg::kStages + 4));
    const uint32_t feed = bar_base + kFeedOffset;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    unsigned int first_unit = 0;
    if (warp == 0 && lane == 0) {
        first_unit = atomicAdd(&g_feed[p.feed_slot][0], 1u);   // in flight while the CTA sets itself up
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < Cfg::kStages; ++s) {
            mbar_init(full_bar(s), 1);
            mbar_init(empty_bar(s), 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(tfull_bar(s), 1);
            mbar_init(tempty_bar(s), 8);
        }
        for (int s = 0; s < kFeedSlots; ++s) {
            mbar_init(feed_full(feed, s), 1);
            mbar_init(feed_empty(feed, s), 1 + 8);   // MMA warp + 8 epilogue warps
        }
        mbar_fence_init();
    }
    if (warp == 2) {
        tmem_alloc(smem_u32(tmem_ptr_smem), Cfg::kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;

    const int units = p.tiles_m * p.tiles_n * p.splits;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        int s = 0, fs = 0;
        uint32_t ph = 0, fph = 0;
        // Unit ids are published one tile ahead: the next unit is drawn when a tile's loads start and handed to the
        // other warps after its second k-block has been issued, by which time the atomic has long returned.
        int u = __shfl_sync(0xffffffffu, (int)first_unit, 0);
        bool live = (unsigned)u < (unsigned)units;
        if (lane == 0) {
            feed_publish<false>(feed, fs, fph, live ? u : -1);
            if (!live) feed_retire(p.feed_slot, gridDim.x);
        }
        if (++fs == kFeedSlots) { fs = 0; fph ^= 1u; }
        while (live) {
            int drawn = 0;
            if (lane == 0) drawn = (int)atomicAdd(&g_feed[p.feed_slot][0], 1u);
            const int tile = u % (p.tiles_m * p.tiles_n);
            const int split = u / (p.tiles_m * p.tiles_n);
            const int m0 = (tile / p.tiles_n) * BM;
            const int n0 = (tile % p.tiles_n) * BN;
            const int kb0 = split * p.kb_per_split;
            const int nkb = min(p.kb_per_split, p.kb_total - kb0);
            for (int kb = 0; kb < nkb; ++kb) {
                if (lane == 0) {
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    mbar_expect_tx(full_bar(s), Cfg::kStageBytes);
                    const uint32_t sA = smem_base + s * Cfg::kStageBytes;
                    const uint32_t sB = sA + Cfg::kABytes;
                    const int k0 = (kb0 + kb) * BK;
                    if (!p.a_mn) {
                        tma_load_2d(sA, &tmA, full_bar(s), k0, m0);  // box {64 k, 128 rows}
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j)  // box {64 m, 64 k}
                            tma_load_2d(sA + j * 8192, &tmA, full_bar(s), m0 + 64 * j, k0);
                    }
                    if (!p.b_mn) {
                        tma_load_2d(sB, &tmB, full_bar(s), k0, n0);  // box {64 k, BN rows}
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_2d(sB + j * 8192, &tmB, full_bar(s), n0 + 64 * j, k0);
                    }
                    if (kb == min(1, nkb - 1)) {
                        const bool more = (unsigned)drawn < (unsigned)units;
                        feed_publish<false>(feed, fs, fph, more ? drawn : -1);
                        if (!more) feed_retire(p.feed_slot, gridDim.x);
                    }
                }
                __syncwarp();
                if (++s == Cfg::kStages) { s = 0; ph ^= 1u; }
            }
            if (++fs == kFeedSlots) { fs = 0; fph ^= 1u; }
            u = __shfl_sync(0xffffffffu, drawn, 0);
            live = (unsigned)u < (unsigned)units;
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        // cute::UMMA::InstrDescriptor: c_format F32 (1<<4), a/b format BF16 (1<<7, 1<<10),
        // a_major bit 15, b_major bit 16, N>>3 at [17,23), M>>4 at [24,29)
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                               ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
        // K-major SW128: 8-row groups 1024 B apart (SBO), LBO unused (=1). k-step of 16 elems = +32 B.
        // MN-major SW128: 64-wide MN blocks 8192 B apart (LBO), 8-k groups 1024 B apart (SBO). k-step = 16 rows = +2048 B.
        const uint32_t a_lbo = p.a_mn ? 512u : 1u, b_lbo = p.b_mn ? 512u : 1u;
        const uint32_t a_step = p.a_mn ? 2048u : 32u, b_step = p.b_mn ? 2048u : 32u;
        constexpr uint32_t kDescHi = desc_hi(64u);   // SBO = 1024 B for both majors
        int s = 0, as = 0;
        uint32_t ph = 0, aph = 0;
        FeedCursor<false> cur(feed);
        for (int u = cur.next(lane); u >= 0; u = cur.next(lane)) {
            const int split = u / (p.tiles_m * p.tiles_n);
            const int kb0 = split * p.kb_per_split;
            const int nkb = min(p.kb_per_split, p.kb_total - kb0);
            if (lane == 0) mbar_wait(tempty_bar(as), aph ^ 1u);
            __syncwarp();
            tc_fence_after();
            const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
            for (int kb = 0; kb < nkb; ++kb) {
                if (lane == 0) {
                    mbar_wait(full_bar(s), ph);
                    tc_fence_after();
                    const uint32_t sA = smem_base + s * Cfg::kStageBytes;
                    const uint32_t a_lo = desc_lo(sA, a_lbo), b_lo = desc_lo(sA + Cfg::kABytes, b_lbo);
#pragma unroll
                    for (int k4 = 0; k4 < BK / UK; ++k4)
                        umma_bf16(d_tmem, desc_make(a_lo + k4 * (a_step >> 4), kDescHi), desc_make(b_lo + k4 * (b_step >> 4), kDescHi),
                                  idesc, (uint32_t)((kb | k4) != 0));
                    umma_commit(empty_bar(s));               // frees the smem slot once these MMAs retire
                    if (kb == nkb - 1) umma_commit(tfull_bar(as));  // accumulator ready for the epilogue
                }
                __syncwarp();
                if (++s == Cfg::kStages) { s = 0; ph ^= 1u; }
            }
            as ^= 1;
            if (as == 0) aph ^= 1u;
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue: 8 warps. TMEM lane quarter = warp % 4
        // (hardware rule), column half = (warp - 4) / 4: two warps per SM sub-partition hide each other's latency.
        // A thread owns one accumulator row (that is how tcgen05.ld hands TMEM out), but global memory wants a warp to
        // touch whole rows: every 32x32 chunk goes through a per-warp smem staging tile (row stride 33 floats, conflict
        // free both ways) so that global loads of aux and global stores are row-contiguous 16-byte accesses.
        const int q = warp & 3;
        const int half = (warp - 4) >> 2;
        constexpr int kChunks = BN / 64;   // 32-column chunks per warp
        float* stg = reinterpret_cast<float*>(smem + Cfg::kStages * Cfg::kStageBytes + 256) + (warp - 4) * (32 * 33);
        int as = 0;
        uint32_t aph = 0;
        FeedCursor<false> cur(feed);
        for (int u = cur.next(lane); u >= 0; u = cur.next(lane)) {
            const int tile = u % (p.tiles_m * p.tiles_n);
            const int m0 = (tile / p.tiles_n) * BM;
            const int n0 = (tile % p.tiles_n) * BN;
            const int row0 = m0 + q * 32;      // first of this warp's 32 rows
            constexpr bool kHasAux = (EPI == EPI_RESID_F32 || EPI == EPI_DGELU || EPI == EPI_RESID_BF16);
            uint4 pre[8];
            bool pre_valid = false;
            {   // the aux operand does not depend on the accumulator: fetch the first chunk's before waiting for the MMAs
                const int c0 = n0 + half * kChunks * 32;
                if (kHasAux && row0 < p.M && c0 < p.N && chunk_is_fast<EPI>(p, c0)) {
                    aux_issue<EPI>(p, pre, lane, row0, c0);
                    pre_valid = true;
                }
            }
            mbar_wait(tfull_bar(as), aph);
            tc_fence_after();
#pragma unroll 1
            for (int cc = 0; cc < kChunks; ++cc) {
                const int c = half * kChunks + cc;
                uint32_t v[32];
                __syncwarp();  // tcgen05.ld is .sync.aligned
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + c * 32), v);
                tmem_ld_wait();
                const int col0 = n0 + c * 32;
                if (col0 >= p.N || row0 >= p.M) continue;  // warp-uniform
                const bool full = (col0 + 32 <= p.N);
                float acc[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) acc[i] = __uint_as_float(v[i]);
                int next_col0 = -1;
                if (kHasAux && cc + 1 < kChunks && col0 + 32 < p.N && chunk_is_fast<EPI>(p, col0 + 32)) next_col0 = col0 + 32;
                epilogue_chunk<EPI>(p, acc, stg, lane, row0, col0, full, pre, pre_valid, next_col0);
                pre_valid = (next_col0 >= 0);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(as));
            as ^= 1;
            if (as == 0) aph ^= 1u;
        }
    }


    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

// ------------------------------------------------------------------ CTA-pair kernel (cta_group::2)
// Output tile 256 x 256 per cluster of two CTAs (same TPC): CTA r owns accumulator rows [128 r, 128 r + 128) in its
// TMEM and loads only its 128 rows of A and its 128-row half of B per k-block (32 KB instead of 48 KB), so the operand
// traffic per flop through L2 drops by a third and the smem ring gets 6 stages deep. The even CTA issues every
// tcgen05.mma.cta_group::2 (M = 256) and owns the full / tmem-empty barriers; completion of MMAs is multicast to both
// CTAs' empty / tmem-full barriers. Epilogue code is shared with the single-CTA kernel.
// EW = epilogue warps per CTA. Measured with the register -> TMA-store epilogue: 16 warps (5 stages) are no faster
// than 8 (6 stages) for any epilogue, so 8 it is.
// The fp32 residual epilogue keeps two 4 KB boxes per warp: the residual chunk arrives in one by a bulk tensor load,
// is updated in place and leaves from the same box by a bulk tensor store, while the other box's load is in flight
// (row-owner register loads of a [32][32] fp32 chunk are 32 cache lines per instruction, and went through the LSU tag
// stage at 6 us per tile). That takes one stage of the operand ring.
template <int EW, int EPI> struct GemmCfg2 {
    static constexpr int BN = 256;
    static constexpr int kABytes = BM * BK * 2;            // 128 rows of A
    static constexpr int kBBytes = (BN / 2) * BK * 2;      // this CTA's half of B
    static constexpr int kStageBytes = kABytes + kBBytes;  // 32 KB
    static constexpr bool kAuxBoxes = EpiTraits<EPI>::kHasAux;
    static constexpr int kBoxBytes = EpiTraits<EPI>::kBoxBytes;
    static constexpr int kStgWarp = (EPI == EPI_RESID_F32) ? 2 * kStgPerWarp : kStgPerWarp;   // two boxes either way
    static constexpr int kStages = (EPI == EPI_RESID_F32) ? 5 : 6;
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kStagingBytes = EW * kStgWarp;    // TMA boxes, 1024-byte aligned behind the barriers
    static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 1024 /*barriers*/ + kStagingBytes;
    static constexpr int kThreads = (4 + EW) * 32;
    static_assert(kSmemBytes <= 232448, "pair kernel exceeds the shared memory of an SM");
};

template <int EPI, int EW>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__((4 + EW) * 32, 1)
gemm_bf16_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmO, const __grid_constant__ CUtensorMap tmO2,
                     const GemmParams p) {
    using Cfg = GemmCfg2<EW, EPI>;
    constexpr int BN = Cfg::BN;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw_addr = smem_u32(smem_raw);
    const uint32_t pad = (1024u - (raw_addr & 1023u)) & 1023u;
    uint8_t* smem = smem_raw + pad;
    const uint32_t smem_base = raw_addr + pad;
    const uint32_t bar_base = smem_base + Cfg::kStages * Cfg::kStageBytes;
    auto full_bar = [&](int s) { return bar_base + 8u * s; };
    auto empty_bar = [&](int s) { return bar_base + 8u * (Cfg::kStages + s); };
    auto tfull_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + s); };
    auto tempty_bar = [&](int s) { return bar_base + 8u * (2 * Cfg::kStages + 2 + s); };
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + Cfg::kStages * Cfg::kStageBytes + 8 * (2 * Cfg::kStages + 4));

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = (rank == 0);
    const uint32_t feed = bar_base + kFeedOffset;   // tile-feed ring; the leader's producer fills both CTAs' copies

    pdl_trigger();   // the next kernel in the stream may begin its own setup as SMs free up
    unsigned int first_unit = 0;
    if (warp == 0 && lane == 0) {
        if (leader) first_unit = atomicAdd(&g_feed[p.feed_slot][0], 1u);
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmO);
        if (EPI == EPI_GELU) tma_prefetch_desc(&tmO2);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < Cfg::kStages; ++s) {
            mbar_init(full_bar(s), 1);     // the leader's producer arrives (with the pair's byte count)
            mbar_init(empty_bar(s), 1);    // one multicast commit
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(tfull_bar(s), 1);    // one multicast commit
            mbar_init(tempty_bar(s), 2 * EW);  // every epilogue warp of both CTAs (only the leader's copy is used)
        }
        if (Cfg::kAuxBoxes)
            for (int w = 0; w < 2 * EW; ++w) mbar_init(bar_base + 512u + 8u * w, 1);
        for (int s = 0; s < kFeedSlots; ++s) {
            mbar_init(feed_full(feed, s), 1);
            // readers: the leader's MMA warp, the odd CTA's producer, every epilogue warp of both CTAs
            // (only the leader's copy of the empty barriers is used)
            mbar_init(feed_empty(feed, s), 2 + 2 * EW);
            if (!leader) mbar_expect_tx(feed_full(feed, s), 4);   // armed for the leader's first st.async
        }
        mbar_fence_init();
    }
    if (warp == 2) {
        tmem_alloc_pair(smem_u32(tmem_ptr_smem), Cfg::kTmemCols);
        tmem_relinquish_pair();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr_smem;
    // Programmatic dependent launch: this grid may have been started while its predecessor in the stream was still
    // draining (its CTAs take SMs as the predecessor's leave them), so everything above -- barrier setup, TMEM
    // allocation, descriptor prefetch -- overlapped that tail. Nothing of the predecessor's output is touched before
    // this point; from here on its memory is complete and visible.
    pdl_wait();

    const int units = p.tiles_m * p.tiles_n * p.splits;   // tiles_m counts 256-row tiles here

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer (both CTAs); the leader's also
        // draws the pair's work units and publishes them to both CTAs
        int s = 0, fs = 0;
        uint32_t ph = 0, fph = 0;
        // (published one tile ahead, so the odd CTA's producer never waits for the hop through DSMEM)
        FeedCursor<true> cur(feed, !leader, !leader);
        int u;
        if (leader) {
            u = __shfl_sync(0xffffffffu, (int)first_unit, 0);
            if ((unsigned)u >= (unsigned)units) u = -1;
            if (lane == 0) {
                feed_publish<true>(feed, fs, fph, u);
                if (u < 0) feed_retire(p.feed_slot, gridDim.x >> 1);
            }
            if (++fs == kFeedSlots) { fs = 0; fph ^= 1u; }
        } else {
            u = cur.next(lane);
        }
        while (u >= 0) {
            int drawn = 0;
            if (leader && lane == 0) drawn = (int)atomicAdd(&g_feed[p.feed_slot][0], 1u);
            const int tile = u % (p.tiles_m * p.tiles_n);
            const int split = u / (p.tiles_m * p.tiles_n);
            const int m0 = (tile / p.tiles_n) * (2 * BM) + (int)rank * BM;
            const int n0 = (tile % p.tiles_n) * BN + (int)rank * (BN / 2);
            const int kb0 = split * p.kb_per_split;
            const int nkb = min(p.kb_per_split, p.kb_total - kb0);
            for (int kb = 0; kb < nkb; ++kb) {
                if (lane == 0) {
                    mbar_wait(empty_bar(s), ph ^ 1u);
                    if (leader) mbar_expect_tx(full_bar(s), 2 * Cfg::kStageBytes);
                    const uint32_t sA = smem_base + s * Cfg::kStageBytes;
                    const uint32_t sB = sA + Cfg::kABytes;
                    const int k0 = (kb0 + kb) * BK;
                    if (!p.a_mn) {
                        tma_load_2d_pair(sA, &tmA, full_bar(s), k0, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BM / 64; ++j) tma_load_2d_pair(sA + j * 8192, &tmA, full_bar(s), m0 + 64 * j, k0);
                    }
                    if (!p.b_mn) {
                        tma_load_2d_pair(sB, &tmB, full_bar(s), k0, n0);   // box {64 k, 128 rows}
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 128; ++j) tma_load_2d_pair(sB + j * 8192, &tmB, full_bar(s), n0 + 64 * j, k0);
                    }
                    if (leader && kb == min(1, nkb - 1)) {
                        const bool more = (unsigned)drawn < (unsigned)units;
                        feed_publish<true>(feed, fs, fph, more ? drawn : -1);
                        if (!more) feed_retire(p.feed_slot, gridDim.x >> 1);
                    }
                }
                __syncwarp();
                if (++s == Cfg::kStages) { s = 0; ph ^= 1u; }
            }
            if (leader) {
                if (++fs == kFeedSlots) { fs = 0; fph ^= 1u; }
                u = __shfl_sync(0xffffffffu, drawn, 0);
                if ((unsigned)u >= (unsigned)units) u = -1;
            } else {
                u = cur.next(lane);
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (even CTA only)
        if (leader) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)p.a_mn << 15) | ((uint32_t)p.b_mn << 16) |
                                   ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
            const uint32_t a_lbo = p.a_mn ? 512u : 1u, b_lbo = p.b_mn ? 512u : 1u;
            const uint32_t a_step = p.a_mn ? 2048u : 32u, b_step = p.b_mn ? 2048u : 32u;
            constexpr uint32_t kDescHi = desc_hi(64u);
            int s = 0, as = 0;
            uint32_t ph = 0, aph = 0;
            FeedCursor<true> cur(feed);
            for (int u = cur.next(lane); u >= 0; u = cur.next(lane)) {
                const int split = u / (p.tiles_m * p.tiles_n);
                const int kb0 = split * p.kb_per_split;
                const int nkb = min(p.kb_per_split, p.kb_total - kb0);
                if (lane == 0) mbar_wait(tempty_bar(as), aph ^ 1u);
                __syncwarp();
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
                for (int kb = 0; kb < nkb; ++kb) {
                    if (lane == 0) {
                        mbar_wait(full_bar(s), ph);
                        tc_fence_after();
                        const uint32_t sA = smem_base + s * Cfg::kStageBytes;
                        const uint32_t a_lo = desc_lo(sA, a_lbo), b_lo = desc_lo(sA + Cfg::kABytes, b_lbo);
#pragma unroll
                        for (int k4 = 0; k4 < BK / UK; ++k4)
                            umma_bf16_pair(d_tmem, desc_make(a_lo + k4 * (a_step >> 4), kDescHi),
                                           desc_make(b_lo + k4 * (b_step >> 4), kDescHi), idesc, (uint32_t)((kb | k4) != 0));
                        umma_commit_pair(empty_bar(s));
                        if (kb == nkb - 1) umma_commit_pair(tfull_bar(as));
                    }
                    __syncwarp();
                    if (++s == Cfg::kStages) { s = 0; ph ^= 1u; }
                }
                as ^= 1;
                if (as == 0) aph ^= 1u;
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------ epilogue (both CTAs, own 128 accumulator rows)
        // TMEM lane quarter = warp % 4 (hardware rule), column slice = (warp - 4) / 4. The accumulator buffer is handed
        // back to the MMA warp as soon as this warp's last chunk sits in registers, before its math and stores.
        const int q = warp & 3;
        const int slice = (warp - 4) >> 2;
        constexpr int kChunks = BN / 32 / (EW / 4);
        const uint32_t stg = smem_base + Cfg::kStages * Cfg::kStageBytes + 1024 + (uint32_t)(warp - 4) * Cfg::kStgWarp;
        const uint32_t aux_bar = bar_base + 512u + 16u * (uint32_t)(warp - 4);   // two mbarriers per epilogue warp
        int as = 0, nbox = 0;
        uint32_t aph = 0, aux_ph = 0;
        FeedCursor<true> cur(feed, !leader);
        for (int u = cur.next(lane); u >= 0; u = cur.next(lane)) {
            const int tile = u % (p.tiles_m * p.tiles_n);
            const int m0 = (tile / p.tiles_n) * (2 * BM) + (int)rank * BM;
            const int n0 = (tile % p.tiles_n) * BN;
            const int row0 = m0 + q * 32;
            const int c_first = n0 + slice * kChunks * 32;
            if constexpr (Cfg::kAuxBoxes) {
                // residual chunks cc and cc + 1 go into boxes 0 and 1 now; chunk cc + 2 follows into box cc & 1 once the
                // store of chunk cc has read it
                auto load_box = [&](int cc) {
                    const uint32_t bar = aux_bar + 8u * (cc & 1);
                    mbar_expect_tx(bar, (uint32_t)Cfg::kBoxBytes);
                    tma_load_2d(stg + (uint32_t)((cc & 1) * Cfg::kBoxBytes), &tmO2, bar, c_first + cc * 32, row0);
                };
                const bool live = row0 < p.M;
                float rowdot = 0.f;
                if (lane == 0 && live) {
                    if (c_first < p.N) load_box(0);
                    if (c_first + 32 < p.N) load_box(1);
                }
#pragma unroll 1
                for (int cc = 0; cc < kChunks; ++cc) {
                    const int col0 = c_first + cc * 32;
                    uint32_t v[32];
                    if (cc == 0) { mbar_wait(tfull_bar(as), aph); tc_fence_after(); }
                    __syncwarp();
                    tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + (col0 - n0)), v);
                    tmem_ld_wait();
                    if (cc == kChunks - 1) {
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_on_leader(tempty_bar(as));
                    }
                    if (col0 >= p.N || !live) continue;   // warp-uniform
                    const uint32_t box = stg + (uint32_t)((cc & 1) * Cfg::kBoxBytes);
                    mbar_wait(aux_bar + 8u * (cc & 1), (aux_ph >> (cc & 1)) & 1u);
                    aux_ph ^= 1u << (cc & 1);
                    const float dot = epilogue_chunk_box<EPI>(p, v, box, lane, col0);
                    if (EPI == EPI_ROWDOT) {
                        // two chunks make a 64-column group (a head): the row's dot product goes to [sequence][group][t]
                        if ((cc & 1) == 0) {
                            rowdot = dot;
                        } else if (row0 + lane < p.M) {
                            const int row = row0 + lane, g = col0 >> 6;
                            reinterpret_cast<float*>(p.out2)[((size_t)(row / p.seq) * (p.N >> 6) + g) * p.seq + row % p.seq] =
                                rowdot + dot;
                        }
                    }
                    fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        tma_store_2d(&tmO, box, col0, row0);
                        bulk_commit();
                        if (cc + 2 < kChunks && col0 + 64 < p.N) {
                            bulk_wait_read<0>();
                            load_box(cc + 2);
                        }
                    }
                }
                if (lane == 0) bulk_wait_read<0>();   // both boxes are free for the next tile's loads
                __syncwarp();
            } else {
            mbar_wait(tfull_bar(as), aph);
            tc_fence_after();
#pragma unroll 1
            for (int cc = 0; cc < kChunks; ++cc) {
                const int col0 = c_first + cc * 32;
                uint32_t v[32];
                __syncwarp();
                tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN + (col0 - n0)), v);
                tmem_ld_wait();
                if (cc == kChunks - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_on_leader(tempty_bar(as));
                }
                if (col0 >= p.N || row0 >= p.M) continue;   // warp-uniform
                epilogue_chunk_ts<EPI>(p, &tmO, &tmO2, v, stg, nbox, lane, row0, col0);
            }
            }
            as ^= 1;
            if (as == 0) aph ^= 1u;
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // stores have left shared memory
    }

    tc_fence_before();
    cluster_sync_all();   // the peer may still be reading this CTA's smem (B halves) or signalling its barriers
    if (warp == 2) tmem_dealloc_pair(tmem_base, Cfg::kTmemCols);
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static std::atomic<void*> cached{nullptr};
    void* fn = cached.load(std::memory_order_acquire);
    if (fn) return (EncodeTiledFn)fn;
    cudaDriverEntryPointQueryResult qres;
    void* sym = nullptr;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !sym) return nullptr;
    cached.store(sym, std::memory_order_release);
    return (EncodeTiledFn)sym;
}

// rows x inner bf16 matrix, row stride ld elements; box {box_inner, box_rows}, 128B swizzle, zero OOB fill
static int encode_bf16_2d(CUtensorMap* tm, const void* ptr, uint64_t inner, uint64_t rows, uint64_t ld,
                          uint32_t box_inner, uint32_t box_rows) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) { set_error("cuTensorMapEncodeTiled entry point unavailable"); return PENROZ_ERR_CUDA; }
    cuuint64_t dims[2] = {inner, rows};
    cuuint64_t strides[1] = {ld * 2};
    cuuint32_t box[2] = {box_inner, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d): ptr=%p inner=%llu rows=%llu ld=%llu box=%ux%u", (int)r, ptr,
                  (unsigned long long)inner, (unsigned long long)rows, (unsigned long long)ld, box_inner, box_rows);
        return PENROZ_ERR_CUDA;
    }
    return PENROZ_OK;
}

// output matrix of the pair kernel: box {32 columns, 32 rows}; bf16 rows of 64 B (SWIZZLE_64B), fp32 rows of 128 B
// (SWIZZLE_128B) -- the layouts box_put_bf16 / box_put_f32 write
static int encode_out_2d(CUtensorMap* tm, const void* ptr, bool f32, uint64_t cols, uint64_t rows, uint64_t ld) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) { set_error("cuTensorMapEncodeTiled entry point unavailable"); return PENROZ_ERR_CUDA; }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {ld * (f32 ? 4u : 2u)};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr),
                    dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    f32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled (output) failed (%d): ptr=%p cols=%llu rows=%llu ld=%llu", (int)r, ptr,
                  (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)ld);
        return PENROZ_ERR_CUDA;
    }
    return PENROZ_OK;
}

// out[c] += sum_r x[r, c] for a bf16 matrix: the bias-gradient side output of EPI_DGELU for the shapes that do not
// take the pair kernel (small problems: one thread per column pair and row slab, red.add)
__global__ void colsum_atomic_bf16_kernel(const __nv_bfloat16* __restrict__ x, long long ld, int rows, int N,
                                          float* __restrict__ out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= N) return;
    const int per = (rows + gridDim.y - 1) / gridDim.y;
    const int r0 = blockIdx.y * per, r1 = min(rows, r0 + per);
    float s = 0.f;
    for (int r = r0; r < r1; ++r) s += __bfloat162float(x[(long long)r * ld + c]);
    if (r1 > r0) atomicAdd(out + c, s);
}

template <int BN, int EPI>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int grid, cudaStream_t st) {
    using Cfg = GemmCfg<BN>;
    auto kern = gemm_bf16_tc_kernel<BN, EPI>;
    static std::atomic<int> configured{0};  // attribute is per-function, per-device; set every time a new device shows up
    int dev = 0;
    cudaGetDevice(&dev);
    if (!((configured.load() >> dev) & 1)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
        configured.fetch_or(1 << dev);
    }
    kern<<<grid, 384, Cfg::kSmemBytes, st>>>(tmA, tmB, p);
    count_launch();
    return check_launch("gemm_bf16_tc");
}

struct OutMaps { CUtensorMap o, o2; };

template <int EPI, int EW>
static int launch_gemm2_ew(const CUtensorMap& tmA, const CUtensorMap& tmB, const OutMaps& om, const GemmParams& p, int grid,
                           cudaStream_t st) {
    using Cfg2 = GemmCfg2<EW, EPI>;
    auto kern = gemm_bf16_tc2_kernel<EPI, EW>;
    static std::atomic<int> configured{0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!((configured.load() >> dev) & 1)) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg2::kSmemBytes);
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
        configured.fetch_or(1 << dev);
    }
    static const bool pdl = !(getenv("PENROZ_GEMM_PDL") && !strcmp(getenv("PENROZ_GEMM_PDL"), "0"));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(Cfg2::kThreads); cfg.dynamicSmemBytes = Cfg2::kSmemBytes; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    cudaError_t le = cudaLaunchKernelEx(&cfg, kern, tmA, tmB, om.o, om.o2, p);
    if (le != cudaSuccess) { set_error("gemm_bf16_tc2 launch: %s", cudaGetErrorString(le)); return PENROZ_ERR_CUDA; }
    count_launch();
    return check_launch("gemm_bf16_tc2");
}
template <int EPI>
static int launch_gemm2(const CUtensorMap& tmA, const CUtensorMap& tmB, const OutMaps& om, const GemmParams& p, int grid,
                        cudaStream_t st) {
    return launch_gemm2_ew<EPI, 8>(tmA, tmB, om, p, grid, st);
}
static int dispatch_epi2(int epi, const CUtensorMap& a, const CUtensorMap& b, const OutMaps& om, const GemmParams& p, int grid,
                         cudaStream_t st) {
    switch (epi) {
        case EPI_BF16: return launch_gemm2<EPI_BF16>(a, b, om, p, grid, st);
        case EPI_GELU: return launch_gemm2<EPI_GELU>(a, b, om, p, grid, st);
        case EPI_RESID_F32: return launch_gemm2<EPI_RESID_F32>(a, b, om, p, grid, st);
        case EPI_RESID_BF16: return launch_gemm2<EPI_RESID_BF16>(a, b, om, p, grid, st);
        case EPI_ACCUM_F32: return launch_gemm2<EPI_ACCUM_F32>(a, b, om, p, grid, st);
        case EPI_DGELU: return launch_gemm2<EPI_DGELU>(a, b, om, p, grid, st);
        case EPI_F32: return launch_gemm2<EPI_F32>(a, b, om, p, grid, st);
        case EPI_ROWDOT: return launch_gemm2<EPI_ROWDOT>(a, b, om, p, grid, st);
    }
    set_error("gemm_bf16: unknown epilogue %d", epi);
    return PENROZ_ERR_ARG;
}

template <int BN>
static int dispatch_epi(int epi, const CUtensorMap& a, const CUtensorMap& b, const GemmParams& p, int grid, cudaStream_t st) {
    switch (epi) {
        case EPI_BF16: return launch_gemm<BN, EPI_BF16>(a, b, p, grid, st);
        case EPI_GELU: return launch_gemm<BN, EPI_GELU>(a, b, p, grid, st);
        case EPI_RESID_F32: return launch_gemm<BN, EPI_RESID_F32>(a, b, p, grid, st);
        case EPI_RESID_BF16: return launch_gemm<BN, EPI_RESID_BF16>(a, b, p, grid, st);
        case EPI_ACCUM_F32: return launch_gemm<BN, EPI_ACCUM_F32>(a, b, p, grid, st);
        case EPI_DGELU: return launch_gemm<BN, EPI_DGELU>(a, b, p, grid, st);
        case EPI_F32: return launch_gemm<BN, EPI_F32>(a, b, p, grid, st);
    }
    set_error("gemm_bf16: unknown epilogue %d", epi);
    return PENROZ_ERR_ARG;
}

}  // namespace penroz

using namespace penroz;

extern "C" int penroz_gemm_bf16(const void* A, long long lda, int a_layout, const void* B, long long ldb, int b_layout,
                                int M, int N, int K, int epilogue, const void* bias, void* out, long long ldo,
                                void* out2, long long ldo2, const void* aux, long long ldaux, int flags,
                                void* stream) {
    PENROZ_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_bf16: empty problem M=%d N=%d K=%d", M, N, K);
    PENROZ_REQUIRE(A && B && out, "gemm_bf16: null operand");
    PENROZ_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm_bf16: lda/ldb must be multiples of 8 elements (16 B) for TMA");
    PENROZ_REQUIRE(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0), "gemm_bf16: operands must be 16-byte aligned");
    PENROZ_REQUIRE(epilogue >= 0 && epilogue <= EPI_ROWDOT, "gemm_bf16: bad epilogue %d", epilogue);
    if (epilogue == EPI_ROWDOT)
        PENROZ_REQUIRE(out2 != nullptr && ldo2 > 0 && N % 64 == 0 && M % ldo2 == 0,
                       "gemm_bf16: row-dot epilogue needs out2, rows per sequence in ldo2 (dividing M=%d) and N=%d %% 64 == 0", M, N);
    if (epilogue == EPI_GELU) PENROZ_REQUIRE(out2 != nullptr, "gemm_bf16: GELU epilogue needs out2");
    if (epilogue == EPI_RESID_F32 || epilogue == EPI_RESID_BF16 || epilogue == EPI_DGELU || epilogue == EPI_ROWDOT)
        PENROZ_REQUIRE(aux != nullptr, "gemm_bf16: epilogue %d needs aux", epilogue);
    cudaStream_t st = (cudaStream_t)stream;

    GemmParams p;
    p.M = M; p.N = N; p.K = K;
    p.a_mn = a_layout ? 1 : 0;
    p.b_mn = b_layout ? 1 : 0;
    p.bias = reinterpret_cast<const __nv_bfloat16*>(bias);
    p.out = out; p.ldo = ldo; p.out2 = out2; p.ldo2 = ldo2; p.aux = aux; p.ldaux = ldaux;
    p.gelu_tanh = (flags & 1) ? 1 : 0;
    p.beta = (flags & 2) ? 1 : 0;
    p.seq = (epilogue == EPI_ROWDOT) ? (int)ldo2 : 1;
    static std::atomic<unsigned int> feed_seq{0};
    p.feed_slot = (int)(feed_seq.fetch_add(1, std::memory_order_relaxed) % kFeedCounters);

    const int sms = num_sms();
    const int tiles_m = (M + BM - 1) / BM;
    // Tile shape: estimate the number of waves of each configuration, weighted by the work one wave does per SM and
    // by the measured efficiency of that kernel (pair 256x256 > single 128x256 > single 128x128), and take the
    // cheapest. The weight-gradient form (split-K supplies parallelism) always takes the widest tile that fits.
    auto ceil_div = [](long long a, long long b) { return (a + b - 1) / b; };
    const long long u_pair = ceil_div(M, 2 * BM) * ceil_div(N, 256);
    const long long u_256 = (long long)tiles_m * ceil_div(N, 256);
    const long long u_128 = (long long)tiles_m * ceil_div(N, 128);
    // the pair kernel writes its output with bulk tensor stores: 16-byte aligned base and row pitch
    const bool out_f32 = (epilogue == EPI_RESID_F32 || epilogue == EPI_ACCUM_F32 || epilogue == EPI_F32);
    const size_t oesz = out_f32 ? 4 : 2;
    const bool tma_out_ok = ((uintptr_t)out % 16 == 0) && ((size_t)ldo * oesz % 16 == 0) &&
                            (epilogue != EPI_GELU || (((uintptr_t)out2 % 16 == 0) && ((size_t)ldo2 * 2 % 16 == 0))) &&
                            (epilogue != EPI_RESID_F32 || (((uintptr_t)aux % 16 == 0) && ((size_t)ldaux * 4 % 16 == 0))) &&
                            ((epilogue != EPI_RESID_BF16 && epilogue != EPI_DGELU && epilogue != EPI_ROWDOT) ||
                             (((uintptr_t)aux % 16 == 0) && ((size_t)ldaux * 2 % 16 == 0)));
    const bool pair_ok = (M >= 256) && (N > 128) && (sms >= 2) && !(flags & 16) && tma_out_ok;
    const double c_pair = pair_ok ? (double)ceil_div(u_pair, sms / 2) * 1.0 : 1e30;
    const double c_256 = (N > 128) ? (double)ceil_div(u_256, sms) * 1.12 : 1e30;
    const double c_128 = (double)ceil_div(u_128, sms) * 0.5 * 1.35;
    int bn = 256;
    bool pair = false;
    if (epilogue == EPI_ACCUM_F32) {
        pair = pair_ok;
        bn = (N > 128) ? 256 : 128;
    } else if (c_pair <= c_256 && c_pair <= c_128) {
        pair = true;
    } else if (c_128 < c_256) {
        bn = 128;
    }
    if (flags & 4) { bn = 128; pair = false; }
    if (flags & 8) { bn = 256; pair = pair_ok; }
    if (epilogue == EPI_ROWDOT) {   // exists in the CTA-pair kernel only
        PENROZ_REQUIRE(pair_ok, "gemm_bf16: the row-dot epilogue needs M >= 256, N > 128 and 16-byte aligned out / aux rows");
        bn = 256; pair = true;
    }
    // a single 128-row tile of activations (decode at batch <= 128): the kernel is a weight stream and every CTA pulls
    // BN rows of W, so narrower tiles = more CTAs pulling in parallel
    if (!(flags & 12) && M <= 128 && epilogue != EPI_ACCUM_F32 && N > 64 && ceil_div(N, 128) * 2 <= sms) { bn = 64; pair = false; }
    p.tiles_m = pair ? (M + 2 * BM - 1) / (2 * BM) : tiles_m;
    p.tiles_n = (N + bn - 1) / bn;
    p.kb_total = (K + BK - 1) / BK;
    p.splits = 1;
    const int workers = pair ? sms / 2 : sms;
    if (epilogue == EPI_ACCUM_F32) {
        // split K so that the units fill whole waves of the workers: the fewest splits that leave the last wave at
        // least 95 % full (27 tiles of the 2304 x 768 QKV gradient: 8 splits = 216 units on 74 CTA pairs, not 2 = 54),
        // each split at least 8 k-blocks deep
        const int tiles = p.tiles_m * p.tiles_n;
        if (tiles < workers) {
            const int smax = max(1, min(p.kb_total / 8, 4 * workers / tiles));
            int best = 1;
            double best_eff = 0.0;
            for (int sp = 1; sp <= smax; ++sp) {
                const int units = tiles * sp;
                const double eff = (double)units / (double)(((units + workers - 1) / workers) * workers);
                if (eff > best_eff + 1e-9) { best_eff = eff; best = sp; }
                if (eff >= 0.95) { best = sp; break; }
            }
            p.splits = best;
        }
        // red.add accumulates into `out`: with beta == 0 the caller must have zeroed it
        if (p.splits > 1 && !p.beta) {
            cudaError_t e = cudaMemset2DAsync(out, (size_t)ldo * 4, 0, (size_t)N * 4, (size_t)M, st);
            if (e != cudaSuccess) { set_error("gemm_bf16: memset: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
        }
    }
    p.kb_per_split = (p.kb_total + p.splits - 1) / p.splits;
    p.splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;  // no empty splits

    CUtensorMap tmA, tmB;
    int rc;
    if (!p.a_mn) rc = encode_bf16_2d(&tmA, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, 64, BM);
    else         rc = encode_bf16_2d(&tmA, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, 64, 64);
    if (rc) return rc;
    if (!p.b_mn) rc = encode_bf16_2d(&tmB, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, 64, (uint32_t)(pair ? bn / 2 : bn));
    else         rc = encode_bf16_2d(&tmB, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, 64, 64);
    if (rc) return rc;

    if (pair) {
        OutMaps om;
        rc = encode_out_2d(&om.o, out, out_f32, (uint64_t)N, (uint64_t)M, (uint64_t)ldo);
        if (rc) return rc;
        if (epilogue == EPI_GELU) {
            rc = encode_out_2d(&om.o2, out2, false, (uint64_t)N, (uint64_t)M, (uint64_t)ldo2);
            if (rc) return rc;
        } else if (epilogue == EPI_RESID_F32 || epilogue == EPI_RESID_BF16 || epilogue == EPI_DGELU || epilogue == EPI_ROWDOT) {
            // the aux operand, loaded box by box
            rc = encode_out_2d(&om.o2, aux, epilogue == EPI_RESID_F32, (uint64_t)N, (uint64_t)M, (uint64_t)ldaux);
            if (rc) return rc;
        } else {
            om.o2 = om.o;
        }
        const int units2 = p.tiles_m * p.tiles_n * p.splits;
        return dispatch_epi2(epilogue, tmA, tmB, om, p, 2 * min(units2, workers), st);
    }
    const int units = p.tiles_m * p.tiles_n * p.splits;
    const int grid = min(units, sms);
    if (bn == 256) rc = dispatch_epi<256>(epilogue, tmA, tmB, p, grid, st);
    else if (bn == 64) rc = dispatch_epi<64>(epilogue, tmA, tmB, p, grid, st);
    else rc = dispatch_epi<128>(epilogue, tmA, tmB, p, grid, st);
    if (rc == PENROZ_OK && epilogue == EPI_DGELU && out2 != nullptr) {
        dim3 g((N + 127) / 128, (unsigned)min(64, (M + 63) / 64));
        colsum_atomic_bf16_kernel<<<g, 128, 0, st>>>((const __nv_bfloat16*)out, ldo, M, N, (float*)out2);
        count_launch();
        rc = check_launch("gemm_bf16/colsum");
    }
    return rc;
}
