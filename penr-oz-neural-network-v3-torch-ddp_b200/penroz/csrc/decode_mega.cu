// The whole single-token decode step of a GPT-2 tree as ONE persistent kernel (bf16, batch <= 64).
//
// Why: at batch 1 a 124M token step streams 247 MB of weights (38 us at the measured 6.4 TB/s) through 61 dependent
// stages (5 per block + the LM head) of 1-5 MB each. As 61 launches -- even with programmatic dependent launch -- every
// stage pays kernel start-up, a cold instruction cache, the drain of the previous grid and a fresh round trip for its
// first weight bytes: ~5.5 us per stage, 0.34 ms per token, 10 % of the HBM roofline. Here one cooperative grid (one
// CTA per SM) walks the stages itself:
//   * stages are separated by a grid-wide barrier (one atomic arrive + one polling thread per CTA, ~1 us) instead of a
//     launch boundary;
//   * the barrier is split: a CTA arrives, THEN issues the weight loads of its first unit of the next stage into
//     registers, THEN waits -- the weights (which no stage writes) are in flight while the slowest CTA finishes;
//   * every Linear is the weight-streaming unit of wstream.cuh (16 weight rows per unit, mma.sync with the batch on the
//     N side, LayerNorm / GELU / residual / K-V page write fused); the embedding is folded into the first stage's
//     activation staging and the arg-max into the LM head;
//   * attention runs as (sequence, head, key-split) units dealt over the CTAs; partials meet in L2 and the last CTA to
//     finish a head merges them (a counter per head), so the stage needs no second barrier.
// Everything a stage reads that another CTA wrote in this launch goes through L2 (ld.global.cg / cp.async.cg); L1 could
// hold a line of the same buffer from two stages ago.
// Reference behaviour replaced: one iteration of the generation loop, neural_net_model.py:465-469 (incremental branch
// :382-386) = forward :250-258 over mappers.py:145-172's tree with the cache branch of neural_net_layers.py:83-92 and
// KVCache.append (kv_cache.py:41-68), then logits[:, -1, :].argmax (:393-394).
#include "common.cuh"
#include "wstream.cuh"
#include <mutex>
#include <stdlib.h>
#include <string.h>

struct penroz_gpt2_layer_t {   // same layout as include/penroz_b200.h
    const void *ln1_w, *ln1_b, *qkv_w, *qkv_b, *proj_w, *proj_b, *ln2_w, *ln2_b, *fc_w, *fc_b, *proj2_w, *proj2_b;
    void* kv_pool;
    float ln1_eps, ln2_eps;
    int gelu_tanh, ffn;
    void* kv_scales;
    int kv_quant, kv_recip;
};

namespace penroz {

constexpr int MEGA_MAX_LAYERS = 64;
constexpr int MEGA_SYNC_BYTES = 16 * 1024;     // barrier words + one counter per (sequence, head)

struct MegaParams {
    const long long* tokens;
    const __nv_bfloat16 *wte, *wpe, *lnf_w, *lnf_b, *head_w, *head_b;
    float lnf_eps;
    const int* bt;
    int max_pages, B, C, H, D, V, pos, n_layer, kchunk, splits, chunk;
    __nv_bfloat16 *xa, *xb, *qkv, *att, *g;
    float* part;
    float* amax_val; int* amax_idx;
    unsigned* sync;              // [0] barrier arrivals, [1] exits, [2] abort flag, [64 + b*H + h] head counters
    __nv_bfloat16* logits;
    long long* argmax_out;
    int* err_flag;
    penroz_gpt2_layer_t layers[MEGA_MAX_LAYERS];
};

__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// split grid barrier: arrive() after a stage's last store, wait() before the next stage's first dependent load
struct GridBar {
    unsigned* word;
    unsigned* abort_word;
    unsigned expect;             // arrivals that complete the next wait
    unsigned nctas;
    bool dead;
    __device__ __forceinline__ void arrive() {
        __syncthreads();
        // one release-reduction: orders this CTA's stores (made visible to thread 0 by the barrier above) before the
        // arrival without a separate fence round trip
        if (threadIdx.x == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(word) : "memory");
    }
    __device__ __forceinline__ void wait() {
        __shared__ int s_dead;
        expect += nctas;
        if (threadIdx.x == 0) {
            int d = 0;
            if (!dead) {
                const unsigned long long t0 = global_ns();
                unsigned spins = 0;
                while (ld_acquire_gpu(word) < expect) {
                    if ((++spins & 1023u) == 0) {
                        // a CTA that never arrives (a fault elsewhere, a stale counter) must not hang the device
                        if (ld_acquire_gpu(abort_word) != 0u || global_ns() - t0 > 2000000000ull) {
                            atomicExch(abort_word, 1u);
                            d = 1;
                            break;
                        }
                    }
                }
                __threadfence();
            } else {
                d = 1;
            }
            s_dead = d;
        }
        __syncthreads();
        dead = s_dead != 0;
        __syncthreads();
    }
};

__device__ __forceinline__ uint4 ldcg16(const void* p) {
    uint4 r;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float ldcg_f32(const float* p) {
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

// ---- attention stage: one (sequence, head, key split) unit per call, whole CTA (8 warps)
// Lanes are (key slot, 16-byte slice of the head dimension): a K or V row is one coalesced read by D/8 lanes.
// sm: s[chunk] | o[8][D] | red[32] | page ids
__device__ __forceinline__ void mega_attn_unit(const MegaParams& P, const __nv_bfloat16* pool, int b, int h, int split,
                                               float* sm, float scale) {
    constexpr int NW = WS_WARPS, NP = 8, VN = 8;
    const int D = P.D, H = P.H, C = P.C, t_len = P.pos + 1, chunk = P.chunk, S = P.splits;
    float* ss = sm;
    float* osm = sm + chunk;
    float* red = osm + NW * D;
    int* pg = reinterpret_cast<int*>(red + 32);
    const int k0 = split * chunk, k1 = min(t_len, k0 + chunk);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int lpk = D / VN, kpw = 32 / lpk;
    const int sub = lane / lpk, dl = (lane - sub * lpk) * VN;
    const int pg0 = k0 / WS_KVP;
    __syncthreads();             // the previous unit's readers of sm are done
    for (int i = tid; k1 > k0 && i <= (k1 - 1) / WS_KVP - pg0; i += WS_THREADS) pg[i] = P.bt[(size_t)b * P.max_pages + pg0 + i];
    float qv[VN];
    {
        const uint4 q = ldcg16(P.qkv + (size_t)b * 3 * C + h * D + dl);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) { const float2 f = unpack_bf16x2(w[j]); qv[2 * j] = f.x * scale; qv[2 * j + 1] = f.y * scale; }
    }
    __syncthreads();
    auto row = [&](int kv, int j) {
        return pool + ((((size_t)pg[j / WS_KVP - pg0] * 2 + kv) * H + h) * WS_KVP + (j % WS_KVP)) * D + dl;
    };
    const int stride = NW * kpw;
    float lm = -INFINITY;
    for (int base = k0 + warp * kpw; base < k1; base += NP * stride) {
        uint4 kx[NP];
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const int j = base + u * stride + sub;
            if (j < k1) kx[u] = ldcg16(row(0, j));
        }
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const int j = base + u * stride + sub;
            float a = 0.f;
            if (j < k1) {
                const uint32_t w[4] = {kx[u].x, kx[u].y, kx[u].z, kx[u].w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 f = unpack_bf16x2(w[i]);
                    a = fmaf(qv[2 * i], f.x, a);
                    a = fmaf(qv[2 * i + 1], f.y, a);
                }
            }
            for (int o = lpk >> 1; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
            if (j < k1) {
                if (dl == 0) ss[j - k0] = a;
                lm = fmaxf(lm, a);
            }
        }
    }
    const float m = block_max(lm, red);   // its barriers also publish ss[]
    float o[VN];
#pragma unroll
    for (int i = 0; i < VN; ++i) o[i] = 0.f;
    float ls = 0.f;
    for (int base = k0 + warp * kpw; base < k1; base += NP * stride) {
        uint4 vx[NP];
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const int j = base + u * stride + sub;
            if (j < k1) vx[u] = ldcg16(row(1, j));
        }
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const int j = base + u * stride + sub;
            if (j < k1) {
                const float pj = __expf(ss[j - k0] - m);
                if (dl == 0) ls += pj;
                const uint32_t w[4] = {vx[u].x, vx[u].y, vx[u].z, vx[u].w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const float2 f = unpack_bf16x2(w[i]);
                    o[2 * i] = fmaf(pj, f.x, o[2 * i]);
                    o[2 * i + 1] = fmaf(pj, f.y, o[2 * i + 1]);
                }
            }
        }
    }
    for (int off = lpk; off < 32; off <<= 1)   // fold the warp's key slots
#pragma unroll
        for (int i = 0; i < VN; ++i) o[i] += __shfl_xor_sync(0xffffffffu, o[i], off);
    if (sub == 0)
#pragma unroll
        for (int i = 0; i < VN; ++i) osm[warp * D + dl + i] = o[i];
    const float l = block_sum(ls, red);   // barriers publish osm[]
    auto folded = [&](int d) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < NW; ++w) t += osm[w * D + d];
        return t;
    };
    __nv_bfloat16* out = P.att + ((size_t)b * H + h) * D;
    if (S == 1) {
        const float inv = 1.f / l;
        for (int d = tid; d < D; d += WS_THREADS) out[d] = __float2bfloat16_rn(folded(d) * inv);
        return;
    }
    // partial {o[D], max, sum} to L2; the CTA that completes the head merges
    float* mine = P.part + (((size_t)b * H + h) * S + split) * (D + 2);
    for (int d = tid; d < D; d += WS_THREADS) mine[d] = folded(d);
    if (tid == 0) {
        mine[D] = (k1 > k0) ? m : -INFINITY;
        mine[D + 1] = (k1 > k0) ? l : 0.f;
    }
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        unsigned* cnt = P.sync + 64 + b * H + h;
        const unsigned old = atomicAdd(cnt, 1u);
        s_last = (old == (unsigned)S - 1u);
        if (s_last) { *cnt = 0u; __threadfence(); }    // every split has arrived: ready for the next block's stage
    }
    __syncthreads();
    if (!s_last) return;
    const float* all = P.part + ((size_t)b * H + h) * S * (D + 2);
    float* wsm = ss;             // S <= 32 merge weights
    if (warp == 0) {
        const float ms = lane < S ? ldcg_f32(all + (size_t)lane * (D + 2) + D) : -INFINITY;
        const float lsv = lane < S ? ldcg_f32(all + (size_t)lane * (D + 2) + D + 1) : 0.f;
        const float mm = warp_max(ms);
        const float w = (ms == -INFINITY) ? 0.f : __expf(ms - mm);
        const float ll = warp_sum(lsv * w);
        wsm[lane] = w / ll;
    }
    __syncthreads();
    for (int d = tid; d < D; d += WS_THREADS) {
        float acc = 0.f;
        for (int r = 0; r < S; ++r) acc = fmaf(ldcg_f32(all + (size_t)r * (D + 2) + d), wsm[r], acc);
        out[d] = __float2bfloat16_rn(acc);
    }
}

__device__ __forceinline__ WsParams mega_linear(const MegaParams& P, const __nv_bfloat16* x, const void* W, const void* bias,
                                                __nv_bfloat16* out, int N, int K, int epi, const __nv_bfloat16* aux) {
    WsParams q{};
    q.x = x; q.ldx = K; q.W = (const __nv_bfloat16*)W; q.ldw = K; q.bias = (const __nv_bfloat16*)bias;
    q.out = out; q.ldo = N; q.aux = aux; q.ldaux = N; q.M = P.B; q.N = N; q.K = K; q.epi = epi;
    return q;
}

// Dynamic shared memory: xs[8 NT][kchunk + WS_XPAD] bf16 | red[WS_WARPS][8 NT][16] fp32 (the attention stage reuses it)
template <int NT>
__global__ void __launch_bounds__(WS_THREADS, 1)
decode_mega_kernel(const __grid_constant__ MegaParams P) {
    extern __shared__ __align__(16) unsigned char mega_smem[];
    constexpr int MB = 8 * NT;
    const int C = P.C, B = P.B;
    const int kchunk = P.kchunk;
    __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(mega_smem);
    float* red = reinterpret_cast<float*>(mega_smem + (size_t)MB * (min(kchunk, max(C, P.layers[0].ffn)) + WS_XPAD) * sizeof(__nv_bfloat16));
    GridBar bar{P.sync, P.sync + 2, 0u, gridDim.x, false};
    uint4 fa[WS_MAXBLK][2];
    const float scale = rsqrtf((float)P.D);
    const int cta = blockIdx.x, ncta = gridDim.x;

    // Stage s of block l = s / 5: 0 LN1+QKV, 1 attention, 2 proj+residual, 3 LN2+fc+GELU, 4 proj2+residual; stage 5 L is the
    // LM head. ONE loop body (one copy of the Linear code in the instruction cache) serves them all.
    const int nstage = 5 * P.n_layer + 1;
    auto build = [&](int s) {
        const int l = s / 5, k = s - 5 * l;
        __nv_bfloat16 *x = P.xa, *xo = P.xb;
        WsParams q{};
        if (s == nstage - 1) {
            q = mega_linear(P, x, P.head_w, P.head_b, P.logits, P.V, C, 0, nullptr);
            q.ln = 1; q.ln_w = P.lnf_w; q.ln_b = P.lnf_b; q.ln_eps = P.lnf_eps;
            if (P.argmax_out) { q.amax_val = P.amax_val; q.amax_idx = P.amax_idx; }
            return q;
        }
        const penroz_gpt2_layer_t& L = P.layers[l];
        if (k == 0) {
            q = mega_linear(P, x, L.qkv_w, L.qkv_b, P.qkv, 3 * C, C, 0, nullptr);
            q.ln = 1; q.ln_w = (const __nv_bfloat16*)L.ln1_w; q.ln_b = (const __nv_bfloat16*)L.ln1_b; q.ln_eps = L.ln1_eps;
            q.pool = (__nv_bfloat16*)L.kv_pool; q.bt = P.bt; q.max_pages = P.max_pages; q.pos = P.pos; q.C = C; q.H = P.H; q.D = P.D;
            if (l == 0) { q.tok = P.tokens; q.wte = P.wte; q.wpe = P.wpe; q.emb_out = x; q.V = P.V; q.err_flag = P.err_flag; }
        } else if (k == 2) {
            q = mega_linear(P, P.att, L.proj_w, L.proj_b, xo, C, C, 2, x);
        } else if (k == 3) {
            q = mega_linear(P, xo, L.fc_w, L.fc_b, P.g, L.ffn, C, 1, nullptr);
            q.gelu_tanh = L.gelu_tanh;
            q.ln = 1; q.ln_w = (const __nv_bfloat16*)L.ln2_w; q.ln_b = (const __nv_bfloat16*)L.ln2_b; q.ln_eps = L.ln2_eps;
        } else {
            q = mega_linear(P, P.g, L.proj2_w, L.proj2_b, x, C, L.ffn, 2, xo);
        }
        return q;
    };
    WsParams q = build(0);
    ws_prefetch_first(q, kchunk, cta, fa);
#pragma unroll 1
    for (int s = 0; s < nstage; ++s) {
        const bool attn = (s != nstage - 1) && (s % 5 == 1);
        if (!bar.dead) {
            if (attn) {
                const penroz_gpt2_layer_t& L = P.layers[s / 5];
                const int units = B * P.H * P.splits;
                for (int u = cta; u < units; u += ncta) {
                    const int split = u % P.splits, bh = u / P.splits;
                    mega_attn_unit(P, (const __nv_bfloat16*)L.kv_pool, bh / P.H, bh % P.H, split, reinterpret_cast<float*>(mega_smem), scale);
                }
            } else {
                const int kc = min(kchunk, q.K);
                ws_stage_rows<NT>(q, xs, kc + WS_XPAD, 0, kc);
                ws_linear<NT>(q, xs, red, kchunk, cta, ncta, cta, ncta, fa);
            }
        }
        if (s == nstage - 1) break;
        bar.arrive();
        const bool next_attn = (s + 1 != nstage - 1) && ((s + 1) % 5 == 1);
        if (!next_attn) {           // the next Linear's first weights: in flight while the slowest CTA finishes this stage
            q = build(s + 1);
            ws_prefetch_first(q, kchunk, cta, fa);
        }
        bar.wait();
    }
    // ---- the arg-max of each row from the LM head's per-CTA candidates
    if (P.argmax_out) {
        bar.arrive();
        bar.wait();
        if (!bar.dead) {
            float* sv = reinterpret_cast<float*>(mega_smem);
            int* si = reinterpret_cast<int*>(sv + 32);
            for (int r = cta; r < B; r += ncta) {
                float best = -INFINITY;
                int bi = 0x7fffffff;
                for (int i = threadIdx.x; i < ncta; i += WS_THREADS) {
                    const float v = ldcg_f32(P.amax_val + (size_t)r * ncta + i);
                    const int ix = __ldcg(P.amax_idx + (size_t)r * ncta + i);
                    if (v > best || (v == best && ix < bi)) { best = v; bi = ix; }
                }
                for (int o = 16; o > 0; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
                }
                const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
                __syncthreads();
                if (lane == 0) { sv[w] = best; si[w] = bi; }
                __syncthreads();
                if (w == 0) {
                    best = lane < WS_WARPS ? sv[lane] : -INFINITY;
                    bi = lane < WS_WARPS ? si[lane] : 0x7fffffff;
                    for (int o = 16; o > 0; o >>= 1) {
                        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
                        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
                    }
                    if (lane == 0) P.argmax_out[r] = (bi == 0x7fffffff) ? 0 : bi;
                }
            }
        }
    }
    // ---- leave the counters as they were found: the last CTA out clears them
    __syncthreads();
    if (threadIdx.x == 0) {
        if (bar.dead && P.err_flag) *P.err_flag = 3;
        __threadfence();
        const unsigned old = atomicAdd(P.sync + 1, 1u);
        if (old == gridDim.x - 1) {
            P.sync[0] = 0u; P.sync[1] = 0u; P.sync[2] = 0u;
            if (bar.dead) for (int i = 0; i < B * P.H; ++i) P.sync[64 + i] = 0u;
            __threadfence();
        }
    }
}

// splits of a head's keys: fill the grid in as few rounds as possible, prefer shorter chunks among equals
static void mega_attn_plan(int B, int H, int t_len, int grid, int* splits, int* chunk) {
    int best_s = 1;
    double best_eff = -1.0;
    for (int s = 1; s <= 32; ++s) {
        if (s > 1 && (t_len + s - 1) / s < 32) break;
        const long long units = (long long)B * H * s;
        const long long rounds = (units + grid - 1) / grid;
        const double eff = (double)units / (double)(rounds * grid) / (double)rounds;   // fewer rounds first
        if (eff > best_eff + 1e-9) { best_eff = eff; best_s = s; }
    }
    *splits = best_s;
    *chunk = (t_len + best_s - 1) / best_s;
}

// Opt-in (PENROZ_DECODE_MEGA=1): measured on B200 at 124M bf16 the persistent form is correct (same parity tests) but
// SLOWER than the launch-per-stage step -- 0.40-0.50 ms against 0.32 ms per token at batch 1, 2.6-3.4 ms against 1.39 ms
// at batch 64: a stage still costs 6-8 us (barrier round trips, one 256-thread CTA per SM keeps too few loads in flight
// for the attention stage, 255 registers with spills), so programmatic dependent launch stays the default.
static bool mega_enabled() {
    static const bool on = getenv("PENROZ_DECODE_MEGA") && !strcmp(getenv("PENROZ_DECODE_MEGA"), "1");
    return on;
}

size_t decode_mega_extra_bytes(int B, int H) { (void)B; (void)H; return MEGA_SYNC_BYTES + 256; }

bool decode_mega_supported(int B, int C, int H, int V, int F, int n_layer, int dtype) {
    if (!mega_enabled() || dtype != BF16 || B < 1 || B > 16 || C % H || n_layer > MEGA_MAX_LAYERS) return false;
    const int D = C / H;
    if (C % 32 || F % 32 || D % 8) return false;
    const int lpk = D / 8;
    if (lpk < 1 || lpk > 32 || (lpk & (lpk - 1))) return false;
    if (C > 1024) return false;                       // LayerNorm rows are normalised out of registers (<= 1024 columns)
    if (B * H > (MEGA_SYNC_BYTES / 4) - 64) return false;
    const int nt = ws_nt(B);
    const int kmax = C > F ? C : F;
    const int kchunk = ws_kchunk(nt, kmax, 200 * 1024);
    if (ws_smem_bytes(nt, kchunk < kmax ? kchunk : kmax) > 200 * 1024) return false;
    if (kchunk < C) return false;                     // LayerNorm / embedding rows are staged whole
    // a Linear whose reduction is staged in chunks restages per unit: allow it only with one unit per CTA
    if (kchunk < F && (C + 15) / 16 > num_sms()) return false;
    (void)V;
    return true;
}

int decode_step_mega(const long long* tokens, const void* wte, const void* wpe, const penroz_gpt2_layer_t* layers,
                     int n_layer, const void* lnf_w, const void* lnf_b, float lnf_eps, const void* head_w,
                     const void* head_b, const int* block_table, int max_pages, int B, int C, int H, int V, int pos,
                     void* xa, void* xb, void* qkv, void* att, void* g, float* part, float* amax_val, int* amax_idx,
                     unsigned* sync, void* logits_out, long long* argmax_out, int* err_flag, cudaStream_t st) {
    PENROZ_REQUIRE(pos / WS_KVP < max_pages, "gpt2_decode_step: position %d exceeds the block table (%d pages)", pos, max_pages);
    static MegaParams P;         // 9 KB: filled per call under the lock below, passed by value
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    int F = 0;
    for (int l = 0; l < n_layer; ++l) {
        F = layers[l].ffn > F ? layers[l].ffn : F;
        PENROZ_REQUIRE(layers[l].ffn == layers[0].ffn, "gpt2_decode_step: blocks with different FFN widths");
    }
    const int nt = ws_nt(B), kmax = C > F ? C : F;
    const int grid = num_sms();
    P.tokens = tokens; P.wte = (const __nv_bfloat16*)wte; P.wpe = (const __nv_bfloat16*)wpe;
    P.lnf_w = (const __nv_bfloat16*)lnf_w; P.lnf_b = (const __nv_bfloat16*)lnf_b; P.lnf_eps = lnf_eps;
    P.head_w = (const __nv_bfloat16*)head_w; P.head_b = (const __nv_bfloat16*)head_b;
    P.bt = block_table; P.max_pages = max_pages; P.B = B; P.C = C; P.H = H; P.D = C / H; P.V = V; P.pos = pos;
    P.n_layer = n_layer;
    P.kchunk = ws_kchunk(nt, kmax, 200 * 1024);
    mega_attn_plan(B, H, pos + 1, grid, &P.splits, &P.chunk);
    P.xa = (__nv_bfloat16*)xa; P.xb = (__nv_bfloat16*)xb; P.qkv = (__nv_bfloat16*)qkv; P.att = (__nv_bfloat16*)att;
    P.g = (__nv_bfloat16*)g; P.part = part; P.amax_val = amax_val; P.amax_idx = amax_idx; P.sync = sync;
    P.logits = (__nv_bfloat16*)logits_out; P.argmax_out = argmax_out; P.err_flag = err_flag;
    for (int l = 0; l < n_layer; ++l) P.layers[l] = layers[l];
    const size_t lin_smem = ws_smem_bytes(nt, P.kchunk < kmax ? P.kchunk : kmax);
    const size_t attn_smem = (size_t)(P.chunk + WS_WARPS * P.D + 32 + P.chunk / WS_KVP + 4) * sizeof(float);
    const size_t smem = lin_smem > attn_smem ? lin_smem : attn_smem;
    void* args[] = {(void*)&P};
    cudaError_t e = cudaSuccess;
    static std::atomic<int> granted[4][64];
    int dev = 0;
    cudaGetDevice(&dev);
#define PENROZ_MEGA(NTV, SLOT)                                                                                       \
    {                                                                                                                \
        auto k = decode_mega_kernel<NTV>;                                                                            \
        if (smem > 48 * 1024 && (int)smem > granted[SLOT][dev & 63].load(std::memory_order_relaxed)) {               \
            e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                     \
            if (e == cudaSuccess) granted[SLOT][dev & 63].store((int)smem, std::memory_order_relaxed);               \
        }                                                                                                            \
        if (e == cudaSuccess) e = cudaLaunchCooperativeKernel((const void*)k, dim3(grid), dim3(WS_THREADS), args, smem, st); \
    }
    if (nt == 1) PENROZ_MEGA(1, 0) else PENROZ_MEGA(2, 1)
#undef PENROZ_MEGA
    if (e != cudaSuccess) { set_error("gpt2_decode_step (persistent): %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
    count_launch();
    return check_launch("decode_mega");
}

}  // namespace penroz
