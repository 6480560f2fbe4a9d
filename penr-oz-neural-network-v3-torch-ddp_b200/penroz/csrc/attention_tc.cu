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
#include <stdlib.h>
#include <string.h>

namespace penroz {

constexpr int AT_BM = 128;   // query rows per tile
constexpr int AT_BN = 128;   // keys per tile
constexpr int AT_D = 64;

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
// Where the forward stands (clock64 timeline of one CTA, B=32 T=1024 H=12, round 2): a key tile takes ~3,080 cycles per
// CTA, two CTAs per SM. 2,080 of them are the exponential pass -- 64 MUFU.EX2 per thread on 16 SFU lanes per SM, which
// both resident CTAs want at once -- and ~1,000 are the hand-off chain (P ready -> S_{j+1} issued, 8 tcgen05.mma at ~75
// cycles of issue each -> S complete -> softmax wakes). The SFU is busy 66 % of the time: the two CTAs fall into
// lock-step (they share the SFU fairly, finish their passes together and idle through the chain together). Packing two
// exponentials per SFU operation (ex2.approx.f16x2) measured SLOWER (171 vs 153 us: the packed form is not double
// rate and adds conversions), and moving a quarter of the exponentials to the FMA pipe (degree-3 polynomial on packed
// fp32 pairs, exponent added as an integer) left the kernel at 151 us: the pass has a second bound of the same size,
// the 64 B/clk at which tensor memory is read (a 128x128 fp32 S tile is 64 KB = 1,024 cycles per CTA, 2,048 for the two
// resident CTAs). What is left is taking the ~1,000-cycle hand-off chain off the critical path: S_{j+1} computed while
// softmax_j runs needs a second S buffer, which two 256-column CTAs per SM do not have (S 128 + P 64 + O 64) -- i.e. one
// 512-column CTA per SM working on two query tiles with explicitly alternating softmax groups.
struct AttnFwdSmem {
    // sized so that TWO CTAs fit in one SM's 227 KB (one CTA's softmax overlaps the other's MMAs): K is single
    // buffered (its only reader, the S MMA, retires long before the next K is needed), V double buffered
    static constexpr int kQ = 0;                     // 16 KB
    static constexpr int kK0 = 16384;                // 16 KB
    static constexpr int kV0 = kK0 + 16384;          // 2 stages x 16 KB
    static constexpr int kBar = kV0 + 2 * 16384;     // 128 B of barriers + tmem slot (P lives in tensor memory)
    static constexpr int kMax = kBar + 128;          // [2 parity][2 halves][128 rows] floats: row-max exchange
    static constexpr int kSum = kMax + 2048;         // [2 halves][128 rows] floats: final row-sum exchange
    static constexpr int kBytes = kSum + 1024 + 1024;  // + alignment slack
};

__device__ __forceinline__ void softmax_pair_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// LAZY: the output accumulates in TMEM across key tiles (the PV MMA adds), and a row is rescaled only when its running
// maximum grows by more than 2^8 -- rare after the first tile -- so the softmax warps neither wait for PV_j nor read
// O back every tile, and the issuing thread queues S_{j+1} ahead of PV_j (S is free as soon as P_j is written).
constexpr float kLazyTau = 8.0f;   // log2 units; probabilities stay <= 2^8, far inside bf16 / fp32 range

template <bool LAZY>
__global__ void __launch_bounds__(320, 2)
attn_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, __nv_bfloat16* __restrict__ out,
                   float* __restrict__ lse, int Tn, int H, float scale_log2e) {
    pdl_trigger();   // a dependent-launch successor (the next GEMM) may set itself up while this grid drains
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
    uint8_t* smem = smem_raw + pad;
    const uint32_t sbase = raw + pad;
    const uint32_t bar = sbase + AttnFwdSmem::kBar;
    const uint32_t q_full = bar, v_full0 = bar + 8, v_empty0 = bar + 24, s_full = bar + 40, p_ready = bar + 48,
                   o_full = bar + 56, k_full = bar + 64, k_empty = bar + 72;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + AttnFwdSmem::kBar + 96);
    float* mx_s = reinterpret_cast<float*>(smem + AttnFwdSmem::kMax);
    float* sum_s = reinterpret_cast<float*>(smem + AttnFwdSmem::kSum);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_qt = (Tn + AT_BM - 1) / AT_BM;
    const int qt = n_qt - 1 - (int)blockIdx.x;       // heaviest tiles first
    const int h = blockIdx.y, b = blockIdx.z;
    const int C = H * AT_D;
    const int row0 = b * Tn + qt * AT_BM;            // first row of this query tile in the [B*T, 3C] matrix
    const int n_kv = qt + 1;

    if (threadIdx.x == 256) {
        tma_prefetch_desc(&tmQKV);
        mbar_init(q_full, 1);
        mbar_init(v_full0, 1); mbar_init(v_full0 + 8, 1);
        mbar_init(v_empty0, 1); mbar_init(v_empty0 + 8, 1);
        mbar_init(k_full, 1); mbar_init(k_empty, 1);
        mbar_init(s_full, 1);
        mbar_init(p_ready, 8);
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
    // S [0,128) fp32 | P [128,192): the probabilities as bf16 pairs, the A operand of the PV product read straight from
    // tensor memory (no shared-memory round trip: the kernel was bound by shared-memory bandwidth, which the MMAs'
    // operand fetch alone saturates -- 8 KB per 64-cycle S step, 6 KB per 32-cycle PV step) | O [192,256)
    const uint32_t tS = tmem, tP = tmem + 128, tO = tmem + 192;

    if (warp >= 8) {
        if (warp == 8 && lane == 0) {
            const uint32_t sQ = sbase + AttnFwdSmem::kQ;
            mbar_expect_tx(q_full, 16384);
            tma_load_2d(sQ, &tmQKV, q_full, h * AT_D, row0);
            const uint32_t sK = sbase + AttnFwdSmem::kK0;
            mbar_expect_tx(k_full, 16384);
            tma_load_2d(sK, &tmQKV, k_full, C + h * AT_D, b * Tn);
            mbar_expect_tx(v_full0, 16384);
            tma_load_2d(sbase + AttnFwdSmem::kV0, &tmQKV, v_full0, 2 * C + h * AT_D, b * Tn);
            const uint32_t idS = at_idesc(0, 0, AT_BN);
            const uint32_t idO = at_idesc(0, 1, AT_D);
            constexpr uint32_t kHi = desc_hi(64u);
            mbar_wait(q_full, 0);
            auto issue_s = [&]() {
                const uint32_t q_lo = desc_lo(sQ, 1), k_lo = desc_lo(sK, 1);
#pragma unroll
                for (int k4 = 0; k4 < AT_D / 16; ++k4)
                    umma_bf16(tS, desc_make(q_lo + k4 * 2, kHi), desc_make(k_lo + k4 * 2, kHi), idS, (uint32_t)(k4 != 0));
                umma_commit(s_full);
                umma_commit(k_empty);
            };
            if (LAZY) {
                mbar_wait(k_full, 0);
                tc_fence_after();
                issue_s();                                               // S_0
                if (n_kv > 1) {
                    mbar_wait(k_empty, 0);
                    mbar_expect_tx(k_full, 16384);
                    tma_load_2d(sK, &tmQKV, k_full, C + h * AT_D, b * Tn + AT_BN);
                    mbar_expect_tx(v_full0 + 8, 16384);
                    tma_load_2d(sbase + AttnFwdSmem::kV0 + 16384, &tmQKV, v_full0 + 8, 2 * C + h * AT_D, b * Tn + AT_BN);
                }
                for (int j = 0; j < n_kv; ++j) {
                    const int s = j & 1;
                    const uint32_t sV = sbase + AttnFwdSmem::kV0 + s * 16384;
                    mbar_wait(v_full0 + 8 * s, (uint32_t)((j >> 1) & 1));
                    mbar_wait(p_ready, (uint32_t)(j & 1));              // P_j in smem, S free, O rescaled if it had to be
                    tc_fence_after();
                    if (j + 1 < n_kv) {
                        mbar_wait(k_full, (uint32_t)((j + 1) & 1));
                        tc_fence_after();
                        issue_s();                                       // S_{j+1} ahead of PV_j
                    }
                    {
                        const uint32_t v_lo = desc_lo(sV, 512);
#pragma unroll
                        for (int k8 = 0; k8 < AT_BN / 16; ++k8)   // 16 keys = 8 columns of P per step
                            umma_bf16_ts(tO, tP + k8 * 8, desc_make(v_lo + k8 * 128, kHi), idO,
                                         (uint32_t)((j | k8) != 0));     // accumulates across key tiles
                    }
                    umma_commit(o_full);
                    umma_commit(v_empty0 + 8 * s);
                    if (j + 2 < n_kv) {
                        mbar_wait(k_empty, (uint32_t)((j + 1) & 1));    // S_{j+1} has consumed K_{j+1}
                        mbar_expect_tx(k_full, 16384);
                        tma_load_2d(sK, &tmQKV, k_full, C + h * AT_D, b * Tn + (j + 2) * AT_BN);
                        mbar_wait(v_empty0 + 8 * s, (uint32_t)((j >> 1) & 1));   // PV_j has consumed V_j
                        mbar_expect_tx(v_full0 + 8 * s, 16384);
                        tma_load_2d(sV, &tmQKV, v_full0 + 8 * s, 2 * C + h * AT_D, b * Tn + (j + 2) * AT_BN);
                    }
                }
            }
            for (int j = 0; !LAZY && j < n_kv; ++j) {
                const int s = j & 1;
                const uint32_t sV = sbase + AttnFwdSmem::kV0 + s * 16384;
                mbar_wait(k_full, (uint32_t)(j & 1));
                tc_fence_after();
                {
                    const uint32_t q_lo = desc_lo(sQ, 1), k_lo = desc_lo(sK, 1);
#pragma unroll
                    for (int k4 = 0; k4 < AT_D / 16; ++k4)
                        umma_bf16(tS, desc_make(q_lo + k4 * 2, kHi), desc_make(k_lo + k4 * 2, kHi), idS, (uint32_t)(k4 != 0));
                }
                umma_commit(s_full);
                umma_commit(k_empty);
                if (j + 1 < n_kv) {
                    const int s2 = s ^ 1;
                    mbar_wait(k_empty, (uint32_t)(j & 1));          // S_j has consumed K_j: refill the single K buffer
                    mbar_expect_tx(k_full, 16384);
                    tma_load_2d(sK, &tmQKV, k_full, C + h * AT_D, b * Tn + (j + 1) * AT_BN);
                    if (j >= 1) mbar_wait(v_empty0 + 8 * s2, (uint32_t)(((j - 1) >> 1) & 1));
                    mbar_expect_tx(v_full0 + 8 * s2, 16384);
                    tma_load_2d(sbase + AttnFwdSmem::kV0 + s2 * 16384, &tmQKV, v_full0 + 8 * s2, 2 * C + h * AT_D,
                                b * Tn + (j + 1) * AT_BN);
                }
                mbar_wait(v_full0 + 8 * s, (uint32_t)((j >> 1) & 1));
                mbar_wait(p_ready, (uint32_t)(j & 1));
                tc_fence_after();
                // O_j = P V_j : A = P K-major (two 64-key blocks 16 KB apart), B = V MN-major (16 keys = 2048 B per step)
                {
                    const uint32_t v_lo = desc_lo(sV, 512);
#pragma unroll
                    for (int k8 = 0; k8 < AT_BN / 16; ++k8)
                        umma_bf16_ts(tO, tP + k8 * 8, desc_make(v_lo + k8 * 128, kHi), idO, (uint32_t)(k8 != 0));
                }
                umma_commit(o_full);
                umma_commit(v_empty0 + 8 * s);
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------ softmax / output threads: 8 warps, two threads per query row.
        // TMEM lane quarter = warp % 4 (hardware rule); warps 0-3 own key columns 0..63 of every S tile and output
        // columns 0..31, warps 4-7 the other halves. The pair shares its row maximum through smem each tile; the row
        // sums stay private until the end.
        const int qd = warp & 3, half = warp >> 2;
        const int r = qd * 32 + lane;
        const int i = qt * AT_BM + r;                      // query position inside the sequence
        const uint32_t lane_addr = (uint32_t)(qd * 32) << 16;
        float o[32];
#pragma unroll
        for (int d = 0; d < 32; ++d) o[d] = 0.f;
        float m = -INFINITY, l = 0.f;
        const uint32_t tPmine = tP + lane_addr + half * 32;      // this thread's 64 probabilities = 32 packed columns
        int* redo_flag = reinterpret_cast<int*>(sum_s + 256);   // 3 rotating CTA-wide flags (slack after the row sums)
        if (threadIdx.x < 3) redo_flag[threadIdx.x] = 0;
        softmax_pair_sync();
        for (int j = 0; j < n_kv; ++j) {
            mbar_wait(s_full, (uint32_t)(j & 1));
            tc_fence_after();
            const bool diag = (j == qt);
            const int key0 = j * AT_BN + half * 64;
            if (LAZY && j > 0) {
                // ---- ONE pass over S for every tile after the first: the exponent uses the running maximum as it stands
                // (it may lag the tile's own maximum by up to 2^tau, which bf16 / fp32 hold easily), the tile's maximum
                // is tracked on the side, and only if some row of the CTA outgrew its maximum by more than tau is the
                // tile redone by the two-pass code below. S is read from tensor memory once instead of twice -- the
                // kernel sat on TMEM read bandwidth (two 64 KB passes per 128x128 tile), not on the exponentials.
                if (threadIdx.x == 0) redo_flag[(j + 1) % 3] = 0;
                const float neg_m = -m;
                float mt = -INFINITY, ls = 0.f;
#pragma unroll 1
                for (int c = 0; c < 2; ++c) {
                    uint32_t v[32];
                    tmem_ld_32x32(tS + lane_addr + half * 64 + c * 32, v);
                    tmem_ld_wait();
                    uint32_t packed[16];
                    if (!diag) {
                        float ls0 = 0.f, ls1 = 0.f, m0 = -INFINITY, m1 = -INFINITY;
#pragma unroll
                        for (int k = 0; k < 32; k += 2) {
                            const float sa = __uint_as_float(v[k]), sb = __uint_as_float(v[k + 1]);
                            m0 = fmaxf(m0, sa); m1 = fmaxf(m1, sb);
                            const float pa = ex2_fast(fmaf(sa, scale_log2e, neg_m));
                            const float pb = ex2_fast(fmaf(sb, scale_log2e, neg_m));
                            ls0 += pa; ls1 += pb;
                            packed[k >> 1] = pack_bf16x2(pa, pb);
                        }
                        ls += ls0 + ls1;
                        mt = fmaxf(mt, fmaxf(m0, m1));
                    } else {
#pragma unroll
                        for (int k = 0; k < 32; k += 2) {
                            const bool oka = key0 + c * 32 + k <= i, okb = key0 + c * 32 + k + 1 <= i;
                            const float sa = __uint_as_float(v[k]), sb = __uint_as_float(v[k + 1]);
                            if (oka) mt = fmaxf(mt, sa);
                            if (okb) mt = fmaxf(mt, sb);
                            const float pa = oka ? ex2_fast(fmaf(sa, scale_log2e, neg_m)) : 0.f;
                            const float pb = okb ? ex2_fast(fmaf(sb, scale_log2e, neg_m)) : 0.f;
                            ls += pa + pb;
                            packed[k >> 1] = pack_bf16x2(pa, pb);
                        }
                    }
                    if (c == 0) {
                        // PV_{j-1} reads the P buffer about to be overwritten (and it is what a rescale would scale): it
                        // was issued a whole S MMA ago and has normally retired while the first half was computed
                        mbar_wait(o_full, (uint32_t)((j - 1) & 1));
                        tc_fence_after();
                    }
                    tmem_st_32x16(tPmine + c * 16, packed);
                }
                tmem_st_wait();
                if (fmaf(mt, scale_log2e, neg_m) > kLazyTau) redo_flag[j % 3] = 1;
                softmax_pair_sync();
                if (redo_flag[j % 3] == 0) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(p_ready);
                    l += ls;
                    if (j == n_kv - 1) {
                        mbar_wait(o_full, (uint32_t)(j & 1));
                        tc_fence_after();
                        uint32_t v[32];
                        tmem_ld_32x32(tO + lane_addr + half * 32, v);
                        tmem_ld_wait();
#pragma unroll
                        for (int k = 0; k < 32; ++k) o[k] = __uint_as_float(v[k]);
                    }
                    tc_fence_before();
                    continue;
                }
                // some row jumped: fall through to the two-pass code, which rescales O and rewrites P for this tile
            }
            // pass 1: maximum over this thread's 64 columns, then over the pair (mask only on the diagonal tile)
            float mt = -INFINITY;
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tS + lane_addr + half * 64 + c * 32, v);
                tmem_ld_wait();
                if (diag) {
#pragma unroll
                    for (int k = 0; k < 32; ++k)
                        mt = fmaxf(mt, (key0 + c * 32 + k > i) ? -INFINITY : __uint_as_float(v[k]));
                } else {
#pragma unroll
                    for (int k = 0; k < 32; ++k) mt = fmaxf(mt, __uint_as_float(v[k]));
                }
            }
            float* mx = mx_s + (j & 1) * 256;
            mx[half * 128 + r] = mt;
            softmax_pair_sync();
            mt = fmaxf(mt, mx[(half ^ 1) * 128 + r]);
            float mn = fmaxf(m, mt * scale_log2e);   // scale > 0: max commutes with the scaling
            float corr = ex2_fast(m - mn);           // m = -inf on the first tile -> 0
            if (LAZY) {
                // keep the stale maximum unless it grew by more than tau; the first tile only sets it (O is overwritten)
                const bool jump = (j > 0) && (mn - m > kLazyTau);
                if (j > 0 && !jump) { mn = m; corr = 1.f; }
                if (j > 0) {
                    // PV_{j-1} must have finished: it reads the P buffer pass 2 is about to overwrite, and it is what a
                    // rescale would scale. By now (a whole pass over S later) it normally has; nothing waits here.
                    mbar_wait(o_full, (uint32_t)((j - 1) & 1));
                    tc_fence_after();
                }
                if (__any_sync(0xffffffffu, jump)) {
                    uint32_t v[32];
                    tmem_ld_32x32(tO + lane_addr + half * 32, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int k = 0; k < 32; ++k) v[k] = __float_as_uint(__uint_as_float(v[k]) * corr);
                    tmem_st_32x32(tO + lane_addr + half * 32, v);
                    tmem_st_wait();
                }
            }
            float ls = 0.f;
            // pass 2: p = 2^(s*scale*log2e - mn), partial row sum, P -> smem (bf16, swizzled K-major). Off the
            // diagonal tile (all but one tile per row of tiles) there is no mask: per pair of scores 2 FFMA, 2 MUFU.EX2,
            // one pack and 2 FADD -- the row sum adds the fp32 probabilities (as torch's softmax does before its own
            // rounding); the kernel was issue-bound at 9 instructions per score with the mask selects and the bf16
            // round trip of the sum in every tile.
            const float neg_mn = -mn;
#pragma unroll 1
            for (int c = 0; c < 2; ++c) {
                uint32_t v[32];
                tmem_ld_32x32(tS + lane_addr + half * 64 + c * 32, v);
                tmem_ld_wait();
                uint32_t packed[16];
                if (!diag) {
                    float ls0 = 0.f, ls1 = 0.f;
#pragma unroll
                    for (int k = 0; k < 32; k += 2) {
                        const float pa = ex2_fast(fmaf(__uint_as_float(v[k]), scale_log2e, neg_mn));
                        const float pb = ex2_fast(fmaf(__uint_as_float(v[k + 1]), scale_log2e, neg_mn));
                        ls0 += pa; ls1 += pb;
                        packed[k >> 1] = pack_bf16x2(pa, pb);
                    }
                    ls += ls0 + ls1;
                } else {
#pragma unroll
                    for (int k = 0; k < 32; k += 2) {
                        float pa = ex2_fast(fmaf(__uint_as_float(v[k]), scale_log2e, neg_mn));
                        float pb = ex2_fast(fmaf(__uint_as_float(v[k + 1]), scale_log2e, neg_mn));
                        if (key0 + c * 32 + k > i) pa = 0.f;
                        if (key0 + c * 32 + k + 1 > i) pb = 0.f;
                        ls += pa + pb;
                        packed[k >> 1] = pack_bf16x2(pa, pb);
                    }
                }
                tmem_st_32x16(tPmine + c * 16, packed);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_ready);
            l = (LAZY && j == 0) ? ls : l * corr + ls;
            m = mn;
            if (!LAZY || j == n_kv - 1) {
                mbar_wait(o_full, (uint32_t)(j & 1));
                tc_fence_after();
                uint32_t v[32];
                tmem_ld_32x32(tO + lane_addr + half * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int k = 0; k < 32; ++k) o[k] = LAZY ? __uint_as_float(v[k]) : o[k] * corr + __uint_as_float(v[k]);
            }
            tc_fence_before();
        }
        sum_s[half * 128 + r] = l;
        softmax_pair_sync();
        l += sum_s[(half ^ 1) * 128 + r];
        if (i < Tn) {
            const float inv = 1.f / l;
            __nv_bfloat16* op = out + ((size_t)(b * Tn + i)) * C + h * AT_D + half * 32;
#pragma unroll
            for (int d = 0; d < 32; d += 8) {
                uint4 w;
                w.x = pack_bf16x2(o[d] * inv, o[d + 1] * inv);
                w.y = pack_bf16x2(o[d + 2] * inv, o[d + 3] * inv);
                w.z = pack_bf16x2(o[d + 4] * inv, o[d + 5] * inv);
                w.w = pack_bf16x2(o[d + 6] * inv, o[d + 7] * inv);
                *reinterpret_cast<uint4*>(op + d) = w;
            }
            if (lse && half == 0) lse[((size_t)b * H + h) * Tn + i] = (m + log2f(l)) * 0.69314718055994530942f;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ---------------------------------------------------------------------------------------------- backward
// delta[b,h,t] = sum_d dO[b,t,h,d] * O[b,t,h,d]: 8 lanes per (row, head), 16-byte loads, 3-step shuffle reduce
__global__ void attn_bwd_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ dout,
                                      float* __restrict__ delta, int rows, int Tn, int H) {
    const long long g = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3;   // (row, head) index
    const int sub = threadIdx.x & 7;
    float s = 0.f;
    const bool ok = g < (long long)rows * H;
    if (ok) {
        const size_t off = (size_t)g * AT_D + sub * 8;            // [rows, H*64] is contiguous in (row, head)
        const uint4 a = *reinterpret_cast<const uint4*>(o + off);
        const uint4 b = *reinterpret_cast<const uint4*>(dout + off);
        const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), a2 = unpack_bf16x2(a.z), a3 = unpack_bf16x2(a.w);
        const float2 b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y), b2 = unpack_bf16x2(b.z), b3 = unpack_bf16x2(b.w);
        s = (a0.x * b0.x + a0.y * b0.y) + (a1.x * b1.x + a1.y * b1.y) + (a2.x * b2.x + a2.y * b2.y) + (a3.x * b3.x + a3.y * b3.y);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (ok && sub == 0) {
        const int row = (int)(g / H), h = (int)(g % H);
        delta[((size_t)(row / Tn) * H + h) * Tn + (row % Tn)] = s;
    }
}

// dq_accum (fp32 [rows, C]) -> bf16 into the q third of dqkv [rows, 3C]
__global__ void attn_bwd_dq_finish_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dqkv, size_t rows,
                                          int C) {
    const size_t n4 = rows * (size_t)C / 4;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = reinterpret_cast<const float4*>(acc)[i];
        const size_t e = i * 4, r = e / C, c = e % C;
        uint2 w;
        w.x = pack_bf16x2(v.x, v.y);
        w.y = pack_bf16x2(v.z, v.w);
        *reinterpret_cast<uint2*>(dqkv + r * 3 * (size_t)C + c) = w;
    }
}

// The same conversion fused with the QKV bias gradient: one pass over the three thirds of dqkv that converts the Q
// third from the fp32 accumulator on the way and leaves per-slab column sums of the bf16 values (Q as just rounded,
// K and V as the main kernel wrote them) in part[gridDim.y][3C]; attn_bwd_bias_reduce_kernel folds the slabs in fixed
// order. Replaces the conversion pass plus a column-sum pass over dqkv (250 MB instead of 300 MB, two launches fewer).
// A thread owns 8 columns (32 lanes x 8 = 256 columns per CTA), 8 row lanes, four rows in flight.
__global__ void __launch_bounds__(256)
attn_bwd_finish_colsum_kernel(const float* __restrict__ acc, __nv_bfloat16* __restrict__ dqkv, int rows, int C,
                              float* __restrict__ part) {
    pdl_trigger();   // a dependent-launch successor (the next GEMM) may set itself up while this grid drains
    __shared__ float sm[8][32][9];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = (blockIdx.x * 32 + tx) * 8;
    const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
    const int r0 = blockIdx.y * rows_per, r1 = min(rows, r0 + rows_per);
    float s[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) s[k] = 0.f;
    if (col < 3 * C) {
        const bool is_q = col < C;
        for (int r = r0 + ty; r < r1; r += 32) {
            uint4 w[4];
            if (is_q) {
                float4 a[4][2];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int rr = r + 8 * u;
                    if (rr < r1) {
                        a[u][0] = *reinterpret_cast<const float4*>(acc + (size_t)rr * C + col);
                        a[u][1] = *reinterpret_cast<const float4*>(acc + (size_t)rr * C + col + 4);
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int rr = r + 8 * u;
                    if (rr < r1) {
                        w[u] = make_uint4(pack_bf16x2(a[u][0].x, a[u][0].y), pack_bf16x2(a[u][0].z, a[u][0].w),
                                          pack_bf16x2(a[u][1].x, a[u][1].y), pack_bf16x2(a[u][1].z, a[u][1].w));
                        *reinterpret_cast<uint4*>(dqkv + (size_t)rr * 3 * C + col) = w[u];
                    }
                }
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int rr = r + 8 * u;
                    if (rr < r1) w[u] = *reinterpret_cast<const uint4*>(dqkv + (size_t)rr * 3 * C + col);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (r + 8 * u < r1) {
                    const float2 p0 = unpack_bf16x2(w[u].x), p1 = unpack_bf16x2(w[u].y), p2 = unpack_bf16x2(w[u].z), p3 = unpack_bf16x2(w[u].w);
                    s[0] += p0.x; s[1] += p0.y; s[2] += p1.x; s[3] += p1.y;
                    s[4] += p2.x; s[5] += p2.y; s[6] += p3.x; s[7] += p3.y;
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) sm[ty][tx][k] = s[k];
    __syncthreads();
    if (ty == 0 && col < 3 * C) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float t = 0.f;
#pragma unroll
            for (int w8 = 0; w8 < 8; ++w8) t += sm[w8][tx][k];
            part[(size_t)blockIdx.y * 3 * C + col + k] = t;
        }
    }
}
__global__ void attn_bwd_bias_reduce_kernel(const float* __restrict__ part, int nparts, int N, float* __restrict__ out) {
    pdl_trigger();   // a dependent-launch successor (the next GEMM) may set itself up while this grid drains
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= N) return;
    float t = 0.f;
    for (int q = 0; q < nparts; ++q) t += part[(size_t)q * N + c];
    out[c] += t;
}

struct AttnBwdSmem {
    static constexpr int kK = 0;                       // 16 KB   K_j  [128 keys][64]
    static constexpr int kV = 16384;                   // 16 KB   V_j
    static constexpr int kQ0 = 32768;                  // 2 x 16 KB  Q_i
    static constexpr int kG0 = kQ0 + 2 * 16384;        // 2 x 16 KB  dO_i
    static constexpr int kP = kG0 + 2 * 16384;         // 32 KB  P   [2 key blocks][128 q][64]
    static constexpr int kDS = kP + 32768;             // 32 KB  dS  same layout
    static constexpr int kBar = kDS + 32768;           // barriers + tmem slot (128 B)
    static constexpr int kStage = kBar + 128;          // 16 warps x [32 rows][17 floats]: dQ staging for coalesced red.add
    static constexpr int kBytes = kStage + 16 * 32 * 17 * 4 + 1024;
};

// 18 warps: 0-15 softmax-gradient (4 threads per query row: TMEM lane quarter = warp % 4, 32-key column slice =
// warp / 4), 16 = MMA issuer, 17 = TMA producer. TMEM: S [0,128) dP [128,256) dV [256,320) dK [320,384) dQ x2 [384,512).
// Per query tile `it`:   MMA: S,dP(it) ........ dV,dK,dQ(it) -> S,dP(it+1) ...
//                        SM :        softmax-grad(it) -> P,dS | dQ(it-1) red.add | softmax-grad(it+1) ...
// dQ is double buffered so the reduction of tile it-1 runs under the MMAs of tile it.
__global__ void __launch_bounds__(576, 1)
attn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                   const __grid_constant__ CUtensorMap tmDQ,
                   const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_accum,
                   __nv_bfloat16* __restrict__ dqkv, int Tn, int H, float scale, float scale_log2e) {
    pdl_trigger();   // a dependent-launch successor (the next GEMM) may set itself up while this grid drains
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t pad = (1024u - (raw & 1023u)) & 1023u;
    uint8_t* smem = smem_raw + pad;
    const uint32_t sbase = raw + pad;
    const uint32_t bar = sbase + AttnBwdSmem::kBar;
    const uint32_t kv_full = bar, qg_full0 = bar + 8, qg_empty0 = bar + 24, sdp_full = bar + 40, pds_ready = bar + 48,
                   dq_full = bar + 56;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + AttnBwdSmem::kBar + 64);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_qt = (Tn + AT_BM - 1) / AT_BM;
    const int jt = blockIdx.x;                         // key tile; tile 0 has the most query tiles -> scheduled first
    const int h = blockIdx.y, b = blockIdx.z;
    const int C = H * AT_D;
    const int n_it = n_qt - jt;                        // query tiles jt .. n_qt-1

    if (threadIdx.x == 512) {
        tma_prefetch_desc(&tmQKV);
        tma_prefetch_desc(&tmDO);
        tma_prefetch_desc(&tmDQ);
        mbar_init(kv_full, 1);
        mbar_init(qg_full0, 1); mbar_init(qg_full0 + 8, 1);
        mbar_init(qg_empty0, 1); mbar_init(qg_empty0 + 8, 1);
        mbar_init(sdp_full, 1);
        mbar_init(pds_ready, 16);
        mbar_init(dq_full, 1);
        mbar_fence_init();
    }
    if (warp == 0) {
        tmem_alloc(smem_u32(tmem_slot), 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t tS = tmem, tDP = tmem + 128, tDV = tmem + 256, tDK = tmem + 320, tDQ = tmem + 384;

    if (warp == 17) {
        // ------------------------------------------------ TMA producer
        if (lane == 0) {
            mbar_expect_tx(kv_full, 32768);
            tma_load_2d(sbase + AttnBwdSmem::kK, &tmQKV, kv_full, C + h * AT_D, b * Tn + jt * AT_BN);
            tma_load_2d(sbase + AttnBwdSmem::kV, &tmQKV, kv_full, 2 * C + h * AT_D, b * Tn + jt * AT_BN);
            for (int it = 0; it < n_it; ++it) {
                const int s = it & 1;
                if (it >= 2) mbar_wait(qg_empty0 + 8 * s, (uint32_t)(((it - 2) >> 1) & 1));
                const int qrow = b * Tn + (jt + it) * AT_BM;
                mbar_expect_tx(qg_full0 + 8 * s, 32768);
                tma_load_2d(sbase + AttnBwdSmem::kQ0 + s * 16384, &tmQKV, qg_full0 + 8 * s, h * AT_D, qrow);
                tma_load_2d(sbase + AttnBwdSmem::kG0 + s * 16384, &tmDO, qg_full0 + 8 * s, h * AT_D, qrow);
            }
        }
        __syncwarp();
    } else if (warp == 16) {
        // ------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t sK = sbase + AttnBwdSmem::kK, sV = sbase + AttnBwdSmem::kV;
            const uint32_t sP = sbase + AttnBwdSmem::kP, sDS = sbase + AttnBwdSmem::kDS;
            const uint32_t id_kk128 = at_idesc(0, 0, 128);   // A K-major, B K-major, N=128  (S, dP)
            const uint32_t id_mm64 = at_idesc(1, 1, 64);     // A MN-major, B MN-major, N=64 (dV, dK)
            const uint32_t id_km64 = at_idesc(0, 1, 64);     // A K-major, B MN-major, N=64  (dQ)
            constexpr uint32_t kHi = desc_hi(64u);
            // loop-invariant descriptor words: K/V as K-major B (S, dP), K as MN-major B (dQ),
            // P/dS as MN-major A (dV, dK; 64-key blocks 16 KB apart) and dS as K-major A (dQ)
            const uint32_t k_lo = desc_lo(sK, 1), v_lo = desc_lo(sV, 1), km_lo = desc_lo(sK, 512);
            const uint32_t pm_lo = desc_lo(sP, 1024), dsm_lo = desc_lo(sDS, 1024), dsk_lo = desc_lo(sDS, 1);
            auto issue_s_dp = [&](int it) {
                const int s = it & 1;
                const uint32_t q_lo = desc_lo(sbase + AttnBwdSmem::kQ0 + s * 16384, 1);
                const uint32_t g_lo = desc_lo(sbase + AttnBwdSmem::kG0 + s * 16384, 1);
                mbar_wait(qg_full0 + 8 * s, (uint32_t)((it >> 1) & 1));
                tc_fence_after();
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4)   // S = Q K^T
                    umma_bf16(tS, desc_make(q_lo + k4 * 2, kHi), desc_make(k_lo + k4 * 2, kHi), id_kk128, (uint32_t)(k4 != 0));
#pragma unroll
                for (int k4 = 0; k4 < 4; ++k4)   // dP = dO V^T
                    umma_bf16(tDP, desc_make(g_lo + k4 * 2, kHi), desc_make(v_lo + k4 * 2, kHi), id_kk128, (uint32_t)(k4 != 0));
                umma_commit(sdp_full);
            };
            mbar_wait(kv_full, 0);
            issue_s_dp(0);
            for (int it = 0; it < n_it; ++it) {
                const int s = it & 1;
                const uint32_t qm_lo = desc_lo(sbase + AttnBwdSmem::kQ0 + s * 16384, 512);
                const uint32_t gm_lo = desc_lo(sbase + AttnBwdSmem::kG0 + s * 16384, 512);
                const uint32_t tDQb = tDQ + (uint32_t)(s * 64);
                mbar_wait(pds_ready, (uint32_t)(it & 1));     // P, dS of this tile are in smem; S/dP TMEM is free again
                tc_fence_after();
                // S, dP of the NEXT query tile go first: they only need the S/dP columns the softmax-gradient threads have
                // just finished reading, and they are what those threads wait for. Issued after dV/dK/dQ (as this loop
                // first did) the 24 + 8 MMA issues (~65 cycles each) sat between two softmax-gradient passes: 4,600 cycles
                // per tile, of which the tensor pipe was busy 1,280 (clock64 timeline, round 2). Now the pass over tile
                // it+1 (~1,900 cycles of exponentials) runs under dV, dK, dQ of tile it.
                if (it + 1 < n_it) issue_s_dp(it + 1);
#pragma unroll
                for (int k8 = 0; k8 < 8; ++k8)   // dV += P^T dO   (M = keys, K = queries; both operands MN-major)
                    umma_bf16(tDV, desc_make(pm_lo + k8 * 128, kHi), desc_make(gm_lo + k8 * 128, kHi), id_mm64,
                              (uint32_t)((it | k8) != 0));
#pragma unroll
                for (int k8 = 0; k8 < 8; ++k8)   // dK += dS^T Q
                    umma_bf16(tDK, desc_make(dsm_lo + k8 * 128, kHi), desc_make(qm_lo + k8 * 128, kHi), id_mm64,
                              (uint32_t)((it | k8) != 0));
#pragma unroll
                for (int k8 = 0; k8 < 8; ++k8)   // dQ_i = dS K   (M = queries, K = keys)
                    umma_bf16(tDQb, desc_make(dsk_lo + (k8 >> 2) * 1024 + (k8 & 3) * 2, kHi), desc_make(km_lo + k8 * 128, kHi),
                              id_km64, (uint32_t)(k8 != 0));
                umma_commit(dq_full);
                umma_commit(qg_empty0 + 8 * s);
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------ softmax-gradient threads
        const int qd = warp & 3, part = warp >> 2;          // lane quarter, 32-key column slice
        const int r = qd * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(qd * 32) << 16;
        uint8_t* pb = smem + AttnBwdSmem::kP + (part >> 1) * 16384;
        uint8_t* db = smem + AttnBwdSmem::kDS + (part >> 1) * 16384;
        float* stg = reinterpret_cast<float*>(smem + AttnBwdSmem::kStage) + warp * (32 * 17);
        const float log2e = 1.44269504088896340736f;
        const size_t stat0 = ((size_t)b * H + h) * Tn;
        float lse_next = 0.f, dl_next = 0.f;
        if (jt * AT_BM + r < Tn) { lse_next = lse[stat0 + jt * AT_BM + r]; dl_next = delta[stat0 + jt * AT_BM + r]; }

        // dQ of query tile `t` (TMEM buffer t & 1): this warp's [32 rows][16 columns] piece goes to shared memory and from
        // there into the fp32 accumulator with ONE bulk tensor reduce-add (cp.reduce.async.bulk.tensor ... .add): the TMA
        // unit performs the 512 atomic adds. Issued per lane as red.global.add.v4 they cost every softmax-gradient warp
        // ~2,300 cycles per tile (clock64 timeline) -- half of the tile's critical path once S, dP were issued early.
        // Rows past the sequence end carry dS = 0, so a piece that straddles into the next sequence adds zeros there.
        auto reduce_dq = [&](int t) {
            uint32_t v[16];
            __syncwarp();
            tmem_ld_32x16(tDQ + (uint32_t)((t & 1) * 64) + lane_addr + part * 16, v);
            tmem_ld_wait();
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // staging is free again
            __syncwarp();
            float4* row = reinterpret_cast<float4*>(stg + lane * 16);
#pragma unroll
            for (int k = 0; k < 4; ++k)
                row[k] = make_float4(__uint_as_float(v[4 * k]), __uint_as_float(v[4 * k + 1]), __uint_as_float(v[4 * k + 2]),
                                     __uint_as_float(v[4 * k + 3]));
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
                const int c0 = h * AT_D + part * 16;
                const int r0 = b * Tn + (jt + t) * AT_BM + qd * 32;
                asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2}], [%3];"
                             ::"l"(&tmDQ), "r"(c0), "r"(r0), "r"(smem_u32(stg)) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            __syncwarp();
            tc_fence_before();
        };

        for (int it = 0; it < n_it; ++it) {
            const int qt = jt + it;
            const int i = qt * AT_BM + r;               // query position
            const bool row_ok = i < Tn;
            const float L2 = lse_next * log2e;          // fetched one iteration ahead: no exposed global latency
            const float dl = dl_next;
            if (it + 1 < n_it && i + AT_BM < Tn) { lse_next = lse[stat0 + i + AT_BM]; dl_next = delta[stat0 + i + AT_BM]; }
            else { lse_next = 0.f; dl_next = 0.f; }
            const bool diag = (it == 0);
            const bool edge = diag || ((qt + 1) * AT_BM > Tn);
            const int key0 = jt * AT_BN + part * 32;
            mbar_wait(sdp_full, (uint32_t)(it & 1));
            tc_fence_after();
            {
                uint32_t sv[32], dv[32];
                tmem_ld_32x32(tS + lane_addr + part * 32, sv);
                tmem_ld_32x32(tDP + lane_addr + part * 32, dv);
                tmem_ld_wait();
                uint32_t pp[16], dd[16];
                // dS = P (dP - delta) scale with the scale folded into one FFMA: (dP - delta) scale = dP scale - delta scale.
                // Interior tiles (no mask, all rows live) run a loop without selects: 5 instructions per score.
                const float negL2 = -L2, ndl = -dl * scale;
                if (!edge) {
#pragma unroll
                    for (int k = 0; k < 32; k += 2) {
                        const float p0 = ex2_fast(fmaf(__uint_as_float(sv[k]), scale_log2e, negL2));
                        const float p1 = ex2_fast(fmaf(__uint_as_float(sv[k + 1]), scale_log2e, negL2));
                        const float d0 = p0 * fmaf(__uint_as_float(dv[k]), scale, ndl);
                        const float d1 = p1 * fmaf(__uint_as_float(dv[k + 1]), scale, ndl);
                        pp[k >> 1] = pack_bf16x2(p0, p1);
                        dd[k >> 1] = pack_bf16x2(d0, d1);
                    }
                } else {      // diagonal tile or a tile with rows past the sequence end
#pragma unroll
                    for (int k = 0; k < 32; k += 2) {
                        float p0 = ex2_fast(fmaf(__uint_as_float(sv[k]), scale_log2e, negL2));
                        float p1 = ex2_fast(fmaf(__uint_as_float(sv[k + 1]), scale_log2e, negL2));
                        if (!row_ok || (diag && key0 + k > i)) p0 = 0.f;
                        if (!row_ok || (diag && key0 + k + 1 > i)) p1 = 0.f;
                        const float d0 = p0 * fmaf(__uint_as_float(dv[k]), scale, ndl);
                        const float d1 = p1 * fmaf(__uint_as_float(dv[k + 1]), scale, ndl);
                        pp[k >> 1] = pack_bf16x2(p0, p1);
                        dd[k >> 1] = pack_bf16x2(d0, d1);
                    }
                }
                // dV, dK, dQ of the previous tile (issued after this tile's S, dP) read the P / dS buffers: they must have
                // retired before the buffers are overwritten -- normally they have, a whole pass of exponentials later
                if (it > 0) { mbar_wait(dq_full, (uint32_t)((it - 1) & 1)); tc_fence_after(); }
#pragma unroll
                for (int q4 = 0; q4 < 4; ++q4) {
                    const uint32_t off = sw128_off(r, (part & 1) * 32 + q4 * 8);
                    *reinterpret_cast<uint4*>(pb + off) = make_uint4(pp[4 * q4], pp[4 * q4 + 1], pp[4 * q4 + 2], pp[4 * q4 + 3]);
                    *reinterpret_cast<uint4*>(db + off) = make_uint4(dd[4 * q4], dd[4 * q4 + 1], dd[4 * q4 + 2], dd[4 * q4 + 3]);
                }
            }
            tc_fence_before();
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(pds_ready);
            if (it > 0) { tc_fence_after(); reduce_dq(it - 1); }   // runs under this tile's dV/dK/dQ MMAs
        }
        mbar_wait(dq_full, (uint32_t)((n_it - 1) & 1));
        tc_fence_after();
        reduce_dq(n_it - 1);
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // the adds have landed before the CTA exits
        // dV, dK of this key tile: lane = key row, 16 columns per thread
        const int key = jt * AT_BN + r;
        const bool key_ok = key < Tn;
        __nv_bfloat16* gk = dqkv + ((size_t)(b * Tn) + (key_ok ? key : 0)) * 3 * C + C + h * AT_D + part * 16;
        __nv_bfloat16* gv = gk + C;
#pragma unroll 1
        for (int which = 0; which < 2; ++which) {
            uint32_t v[16];
            __syncwarp();
            tmem_ld_32x16((which == 0 ? tDV : tDK) + lane_addr + part * 16, v);
            tmem_ld_wait();
            if (key_ok) {
                __nv_bfloat16* dst = which == 0 ? gv : gk;
#pragma unroll
                for (int k = 0; k < 16; k += 8) {
                    uint4 w;
                    w.x = pack_bf16x2(__uint_as_float(v[k]), __uint_as_float(v[k + 1]));
                    w.y = pack_bf16x2(__uint_as_float(v[k + 2]), __uint_as_float(v[k + 3]));
                    w.z = pack_bf16x2(__uint_as_float(v[k + 4]), __uint_as_float(v[k + 5]));
                    w.w = pack_bf16x2(__uint_as_float(v[k + 6]), __uint_as_float(v[k + 7]));
                    *reinterpret_cast<uint4*>(dst + k) = w;
                }
            }
        }
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
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

// fp32 [rows, cols] accumulator, box {16 columns, 32 rows}, no swizzle: the target of the dQ bulk reduce-add
static int encode_dq_f32(CUtensorMap* tm, const void* ptr, uint64_t cols, uint64_t rows) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess || !sym) {
        set_error("cuTensorMapEncodeTiled entry point unavailable");
        return PENROZ_ERR_CUDA;
    }
    cuuint64_t dims[2] = {cols, rows};
    cuuint64_t strides[1] = {cols * 4};
    cuuint32_t box[2] = {16, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = ((EncodeTiledFnA)sym)(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(ptr), dims, strides, box,
                                       estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("attention: cuTensorMapEncodeTiled (dQ) failed (%d)", (int)r); return PENROZ_ERR_CUDA; }
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
        cudaError_t e = cudaFuncSetAttribute(attn_fwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnFwdSmem::kBytes);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(attn_fwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnFwdSmem::kBytes);
        if (e != cudaSuccess) { set_error("attn_fwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
        configured.fetch_or(1 << dev);
    }
    dim3 grid((T + AT_BM - 1) / AT_BM, H, B);
    const float scale_log2e = (1.0f / sqrtf((float)D)) * 1.44269504088896340736f;
    // PENROZ_ATTN_FWD=classic: rescale and read the output back every key tile (the first version of this kernel)
    static const bool lazy = !(getenv("PENROZ_ATTN_FWD") && !strcmp(getenv("PENROZ_ATTN_FWD"), "classic"));
    if (lazy)
        attn_fwd_tc_kernel<true><<<grid, 320, AttnFwdSmem::kBytes, (cudaStream_t)stream>>>(tm, (__nv_bfloat16*)out, lse, T, H, scale_log2e);
    else
        attn_fwd_tc_kernel<false><<<grid, 320, AttnFwdSmem::kBytes, (cudaStream_t)stream>>>(tm, (__nv_bfloat16*)out, lse, T, H, scale_log2e);
    count_launch();
    return check_launch("attn_fwd_tc");
}

extern "C" size_t penroz_colsum_scratch_bytes(int M, int N);

extern "C" size_t penroz_attn_bwd_tc_scratch_bytes(int B, int T, int H, int D) {
    // row dot-products | fp32 dQ accumulator | partial column sums of the optional bias-gradient pass
    return ((((size_t)B * H * T + 3) / 4) * 4 + (size_t)B * T * H * D) * sizeof(float) +
           penroz_colsum_scratch_bytes(B * T, 3 * H * D);
}

// dqkv [B,T,3C] <- gradient of causal attention wrt the packed QKV buffer. scratch: delta [B,H,T] + dq accumulator
// [B*T, C] fp32 (zeroed here with cudaMemsetAsync on `stream`).
extern "C" int penroz_attn_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                  float* dbias, int flags, int B, int T, int H, int D, void* scratch,
                                  size_t scratch_bytes, void* stream) {
    PENROZ_REQUIRE(B > 0 && T > 0 && H > 0, "attn_bwd_tc: bad shape");
    PENROZ_REQUIRE(D == 64, "attn_bwd_tc: head_dim %d unsupported (tensor-core path is head_dim 64)", D);
    PENROZ_REQUIRE(B <= 65535 && H <= 65535, "attn_bwd_tc: B/H too large for the launch grid");
    PENROZ_REQUIRE(scratch && scratch_bytes >= penroz_attn_bwd_tc_scratch_bytes(B, T, H, D), "attn_bwd_tc: scratch too small");
    PENROZ_REQUIRE(((uintptr_t)qkv % 16 == 0) && ((uintptr_t)dout % 16 == 0) && ((uintptr_t)dqkv % 16 == 0) &&
                   ((uintptr_t)scratch % 16 == 0), "attn_bwd_tc: buffers must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const int C = H * D;
    const size_t rows = (size_t)B * T;
    float* delta = (float*)scratch;
    float* dq_acc = delta + (((size_t)B * H * T + 3) / 4) * 4;
    CUtensorMap tmQ, tmG;
    int rc = encode_rows_bf16(&tmQ, qkv, (uint64_t)3 * C, rows, 128);
    if (rc) return rc;
    rc = encode_rows_bf16(&tmG, dout, (uint64_t)C, rows, 128);
    if (rc) return rc;
    CUtensorMap tmDQ;
    rc = encode_dq_f32(&tmDQ, dq_acc, (uint64_t)C, rows);
    if (rc) return rc;
    static std::atomic<int> configured{0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!((configured.load() >> dev) & 1)) {
        cudaError_t e = cudaFuncSetAttribute(attn_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnBwdSmem::kBytes);
        if (e != cudaSuccess) { set_error("attn_bwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
        configured.fetch_or(1 << dev);
    }
    cudaError_t e = cudaMemsetAsync(dq_acc, 0, rows * C * sizeof(float), st);
    if (e != cudaSuccess) { set_error("attn_bwd_tc: memset: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
    const long long pairs = (long long)rows * H;
    if (!(flags & 1))   // bit 0: the caller's projection-gradient GEMM already left delta at the head of scratch
        attn_bwd_delta_kernel<<<(unsigned)((pairs * 8 + 255) / 256), 256, 0, st>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, delta, (int)rows, T, H);
    dim3 grid((T + AT_BN - 1) / AT_BN, H, B);
    const float scale = 1.0f / sqrtf((float)D);
    attn_bwd_tc_kernel<<<grid, 576, AttnBwdSmem::kBytes, st>>>(tmQ, tmG, tmDQ, lse, delta, dq_acc, (__nv_bfloat16*)dqkv, T, H, scale, scale * 1.44269504088896340736f);
    if (dbias != nullptr) {
        // dQ conversion and the bias gradient of the QKV Linear (column sums of dqkv as written) in one pass
        constexpr int kParts = 32;
        float* part = dq_acc + rows * C;
        dim3 g((3 * C + 255) / 256, kParts);
        attn_bwd_finish_colsum_kernel<<<g, 256, 0, st>>>(dq_acc, (__nv_bfloat16*)dqkv, (int)rows, C, part);
        attn_bwd_bias_reduce_kernel<<<(3 * C + 255) / 256, 256, 0, st>>>(part, kParts, 3 * C, dbias);
        count_launch((flags & 1) ? 3 : 4);
    } else {
        size_t n4 = rows * C / 4;
        size_t g = (n4 + 255) / 256;
        size_t cap = (size_t)num_sms() * 16;
        attn_bwd_dq_finish_kernel<<<(unsigned)(g > cap ? cap : g), 256, 0, st>>>(dq_acc, (__nv_bfloat16*)dqkv, rows, C);
        count_launch((flags & 1) ? 2 : 3);
    }
    rc = check_launch("attn_bwd_tc");
    return rc;
}
