// Weight-streaming Linear for the decode step: parameters, device code and host-side launch, shared by wstream.cu
// (one launch per Linear), decode_fused.cu (the step built from those launches) and decode_mega.cu (the whole token
// step as one persistent kernel). See wstream.cu for the design notes.
#pragma once
#include "common.cuh"

namespace penroz {

constexpr int WS_KVP = 64;        // KV page size (decode.cu)
constexpr int WS_WARPS = 8;
constexpr int WS_THREADS = WS_WARPS * 32;
constexpr int WS_MAXBLK = 4;      // k32 blocks a warp owns per pass
constexpr int WS_KPASS = WS_WARPS * WS_MAXBLK * 32;   // 1024 reduction elements per pass
constexpr int WS_XPAD = 32;       // bf16 elements of padding per staged activation row (64 bytes)

struct WsParams {
    const __nv_bfloat16* x; long long ldx;      // activation rows [M, K]
    const __nv_bfloat16* W; long long ldw;      // [N, K]
    const __nv_bfloat16* bias;
    __nv_bfloat16* out; long long ldo;
    const __nv_bfloat16* aux; long long ldaux;  // residual rows (epi 2)
    int M, N, K, epi, gelu_tanh;                // epi: 0 store, 1 GELU, 2 residual
    const __nv_bfloat16* ln_w; const __nv_bfloat16* ln_b; float ln_eps; int ln;
    __nv_bfloat16* pool; const int* bt; int max_pages, pos, C, H, D;   // QKV: columns >= C also go to the page pool
    float* amax_val; int* amax_idx;             // LM head: per-CTA arg-max [M][gridDim.x]
    // the activations are the embedding of `tok` at position `pos` (first Linear of the persistent step):
    // x[m,:] = wte[tok[m]] + wpe[pos]; CTA 0 also stores them to emb_out (the residual stream)
    const long long* tok; const __nv_bfloat16* wte; const __nv_bfloat16* wpe; __nv_bfloat16* emb_out; int V; int* err_flag;
};

bool wstream_supported(int M, int N, int K, int dtype);
int wstream_grid(int M, int N);                 // CTAs (= arg-max candidates per row) a launch uses
int launch_wstream(const WsParams& p, cudaStream_t st);
int ws_nt(int M);
size_t ws_smem_bytes(int nt, int kchunk);
int ws_kchunk(int nt, int K, size_t budget);    // columns of the activations staged at once (K, or a multiple of 1024)

// ------------------------------------------------------------------------------------------------ device code
__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint4 ldg_stream16(const void* p) {      // weights are read once: do not keep them in L1
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ uint4 ldg_act16(const void* p) {         // activations: written by another CTA / launch, L2
    uint4 r;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {   // .cg: L2 only, never a stale L1 line
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
    asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
}

// weight fragments of one 16-row unit for this warp: k32 blocks warp, warp + 8, ... of [kb, kb + 32 nblk) of rows
// n0 + g and n0 + g + 8
__device__ __forceinline__ void ws_load_unit(const WsParams& p, int n0, int kb, int nblk, uint4 (&fa)[WS_MAXBLK][2]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    const int r0 = min(n0 + g, p.N - 1), r1 = min(n0 + g + 8, p.N - 1);     // tail rows: clamped, computed, dropped
    const __nv_bfloat16* w0 = p.W + (size_t)r0 * p.ldw + kb + 8 * t;
    const __nv_bfloat16* w1 = p.W + (size_t)r1 * p.ldw + kb + 8 * t;
#pragma unroll
    for (int i = 0; i < WS_MAXBLK; ++i) {
        const int blk = warp + WS_WARPS * i;
        if (blk < nblk) {
            fa[i][0] = ldg_stream16(w0 + 32 * blk);
            fa[i][1] = ldg_stream16(w1 + 32 * blk);
        }
    }
}

// first fetch of a Linear for this CTA's first unit (issued before the activations are waited for)
__device__ __forceinline__ void ws_prefetch_first(const WsParams& p, int kchunk, int unit, uint4 (&fa)[WS_MAXBLK][2]) {
    const int nunits = (p.N + 15) / 16;
    if (unit < nunits) ws_load_unit(p, unit * 16, 0, min(min(kchunk, p.K), WS_KPASS) / 32, fa);
}

// Stage columns [k0, k0 + kc) of the MB activation rows into xs (bf16, row stride xstride), all threads, every 16-byte
// load independent of the others; rows >= M are zero. With p.ln (needs k0 == 0 and kc == K) LayerNorm is applied in
// place afterwards, one warp per row out of shared memory. Ends with a barrier.
template <int NT>
__device__ __forceinline__ void ws_stage_rows_general_impl(const WsParams& p, __nv_bfloat16* xs, int xstride, int k0, int kc) {
    constexpr int MB = 8 * NT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int kv8 = kc >> 3, total = MB * kv8;
    if (p.tok) {   // embedding rows: wte[tok] + wpe[pos], rounded to bf16 like the embedding kernel
#pragma unroll 2
        for (int e = tid; e < total; e += WS_THREADS) {
            const int r = e / kv8, c = e - r * kv8;
            uint4 o = make_uint4(0, 0, 0, 0);
            if (r < p.M) {
                long long tk = p.tok[r];
                if (tk < 0 || tk >= p.V) { if (p.err_flag) *p.err_flag = 1; tk = 0; }
                const uint4 a = *reinterpret_cast<const uint4*>(p.wte + (size_t)tk * p.K + k0 + c * 8);
                const uint4 b = *reinterpret_cast<const uint4*>(p.wpe + (size_t)p.pos * p.K + k0 + c * 8);
                const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
                uint32_t ow[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 fa2 = unpack_bf16x2(aw[j]), fb2 = unpack_bf16x2(bw[j]);
                    ow[j] = pack_bf16x2(fa2.x + fb2.x, fa2.y + fb2.y);
                }
                o = make_uint4(ow[0], ow[1], ow[2], ow[3]);
                if (p.emb_out && blockIdx.x == 0) *reinterpret_cast<uint4*>(p.emb_out + (size_t)r * p.K + k0 + c * 8) = o;
            }
            *reinterpret_cast<uint4*>(xs + (size_t)r * xstride + c * 8) = o;
        }
    } else if (p.ln && kc < p.K) {
        // LayerNorm over a row longer than what is staged at once: row statistics straight from global memory first
        // (one warp per row, the row in registers), then every staged slice is normalised on its way in
        __shared__ float ws_mean[64], ws_rstd[64];
        if (k0 == 0) {
            const int K = p.K;
            for (int r = warp; r < p.M; r += WS_WARPS) {
                const __nv_bfloat16* src = p.x + (size_t)r * p.ldx;
                uint4 v[8];
                float s = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k = lane * 8 + 256 * i;
                    if (k < K) {
                        v[i] = ldg_act16(src + k);
                        const uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) { const float2 f = unpack_bf16x2(w[j]); s += f.x + f.y; }
                    }
                }
                const float mean = warp_sum(s) / (float)K;
                float q = 0.f;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int k = lane * 8 + 256 * i;
                    if (k < K) {
                        const uint32_t w[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 f = unpack_bf16x2(w[j]);
                            q += (f.x - mean) * (f.x - mean) + (f.y - mean) * (f.y - mean);
                        }
                    }
                }
                const float rstd = rsqrtf(warp_sum(q) / (float)K + p.ln_eps);
                if (lane == 0) { ws_mean[r] = mean; ws_rstd[r] = rstd; }
            }
            __syncthreads();
        }
#pragma unroll 2
        for (int e = tid; e < total; e += WS_THREADS) {
            const int r = e / kv8, c = e - r * kv8;
            uint4 o4 = make_uint4(0, 0, 0, 0);
            if (r < p.M) {
                const int k = k0 + c * 8;
                const uint4 v = ldg_act16(p.x + (size_t)r * p.ldx + k);
                const uint4 gw = *reinterpret_cast<const uint4*>(p.ln_w + k);
                uint4 gb = make_uint4(0, 0, 0, 0);
                if (p.ln_b) gb = *reinterpret_cast<const uint4*>(p.ln_b + k);
                const float mean = ws_mean[r], rstd = ws_rstd[r];
                const uint32_t w[4] = {v.x, v.y, v.z, v.w}, gws[4] = {gw.x, gw.y, gw.z, gw.w}, gbs[4] = {gb.x, gb.y, gb.z, gb.w};
                uint32_t o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_bf16x2(w[j]), a = unpack_bf16x2(gws[j]), b = unpack_bf16x2(gbs[j]);
                    o[j] = pack_bf16x2((f.x - mean) * rstd * a.x + b.x, (f.y - mean) * rstd * a.y + b.y);
                }
                o4 = make_uint4(o[0], o[1], o[2], o[3]);
            }
            *reinterpret_cast<uint4*>(xs + (size_t)r * xstride + c * 8) = o4;
        }
        __syncthreads();
        return;
    } else {
        // asynchronous 16-byte copies global (L2) -> shared: no register staging, so a thread has all of its slices in
        // flight at once (64 rows x 768 columns = 24 copies per thread) instead of a few loads per round trip
        for (int e = tid; e < total; e += WS_THREADS) {
            const int r = e / kv8, c = e - r * kv8;
            __nv_bfloat16* dst = xs + (size_t)r * xstride + c * 8;
            if (r < p.M) cp_async16(dst, p.x + (size_t)r * p.ldx + k0 + c * 8);
            else *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
        }
        cp_async_wait_all();
    }
    __syncthreads();
    if (p.ln) {
        const int K = p.K;
        if (K <= 1024) {
            // A lane owns the same <= 4 16-byte column slices in every row: gamma and beta are unpacked once, a row's
            // values stay in registers from the first statistics pass to the store (one shared-memory read and one
            // write per element, ~6 instructions per element)
            uint4 gp[4], bp[4];      // kept packed: 8 registers per slice instead of 16
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int k = lane * 8 + 256 * i;
                gp[i] = make_uint4(0, 0, 0, 0); bp[i] = make_uint4(0, 0, 0, 0);
                if (k < K) {
                    gp[i] = *reinterpret_cast<const uint4*>(p.ln_w + k);
                    if (p.ln_b) bp[i] = *reinterpret_cast<const uint4*>(p.ln_b + k);
                }
            }
            const float inv_k = 1.f / (float)K;
            for (int r = warp; r < p.M; r += WS_WARPS) {
                __nv_bfloat16* row = xs + (size_t)r * xstride;
                float xv[4][8];
                float s = 0.f;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int k = lane * 8 + 256 * i;
                    uint4 v = make_uint4(0, 0, 0, 0);
                    if (k < K) v = *reinterpret_cast<const uint4*>(row + k);
                    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 f = unpack_bf16x2(w[j]);
                        xv[i][2 * j] = f.x; xv[i][2 * j + 1] = f.y;
                        s += f.x + f.y;
                    }
                }
                const float mean = warp_sum(s) * inv_k;
                float q = 0.f;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const bool on = lane * 8 + 256 * i < K;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        xv[i][j] -= mean;
                        if (on) q = fmaf(xv[i][j], xv[i][j], q);
                    }
                }
                const float rstd = rsqrtf(warp_sum(q) * inv_k + p.ln_eps);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int k = lane * 8 + 256 * i;
                    if (k < K) {
                        const uint32_t gws[4] = {gp[i].x, gp[i].y, gp[i].z, gp[i].w};
                        const uint32_t gbs[4] = {bp[i].x, bp[i].y, bp[i].z, bp[i].w};
                        uint32_t o[4];
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 a2 = unpack_bf16x2(gws[j]), b2 = unpack_bf16x2(gbs[j]);
                            o[j] = pack_bf16x2(fmaf(xv[i][2 * j] * rstd, a2.x, b2.x), fmaf(xv[i][2 * j + 1] * rstd, a2.y, b2.y));
                        }
                        *reinterpret_cast<uint4*>(row + k) = make_uint4(o[0], o[1], o[2], o[3]);
                    }
                }
            }
            __syncthreads();
            return;
        }
        for (int r = warp; r < p.M; r += WS_WARPS) {
            __nv_bfloat16* row = xs + (size_t)r * xstride;
            float s = 0.f;
            for (int k = lane * 8; k < K; k += 256) {
                const uint4 v = *reinterpret_cast<const uint4*>(row + k);
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) { const float2 f = unpack_bf16x2(w[j]); s += f.x + f.y; }
            }
            const float mean = warp_sum(s) / (float)K;
            float q = 0.f;
            for (int k = lane * 8; k < K; k += 256) {
                const uint4 v = *reinterpret_cast<const uint4*>(row + k);
                const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_bf16x2(w[j]);
                    q += (f.x - mean) * (f.x - mean) + (f.y - mean) * (f.y - mean);
                }
            }
            const float rstd = rsqrtf(warp_sum(q) / (float)K + p.ln_eps);
            for (int k = lane * 8; k < K; k += 256) {
                const uint4 v = *reinterpret_cast<const uint4*>(row + k);
                const uint4 gw = *reinterpret_cast<const uint4*>(p.ln_w + k);
                uint4 gb = make_uint4(0, 0, 0, 0);
                if (p.ln_b) gb = *reinterpret_cast<const uint4*>(p.ln_b + k);
                const uint32_t w[4] = {v.x, v.y, v.z, v.w}, gws[4] = {gw.x, gw.y, gw.z, gw.w}, gbs[4] = {gb.x, gb.y, gb.z, gb.w};
                uint32_t o[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float2 f = unpack_bf16x2(w[j]), a = unpack_bf16x2(gws[j]), b = unpack_bf16x2(gbs[j]);
                    o[j] = pack_bf16x2((f.x - mean) * rstd * a.x + b.x, (f.y - mean) * rstd * a.y + b.y);
                }
                *reinterpret_cast<uint4*>(row + k) = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
        __syncthreads();
    }
}

template <int NT>
__device__ __forceinline__ void ws_stage_rows(const WsParams& p, __nv_bfloat16* xs, int xstride, int k0, int kc) {
    ws_stage_rows_general_impl<NT>(p, xs, xstride, k0, kc);
}

// One Linear for this CTA: units first_unit, first_unit + unit_stride, ... of 16 weight rows each.
//   xs      staged activations (ws_stage_rows); when kchunk < K the columns are staged chunk by chunk here
//   red     [WS_WARPS][MB][16] fp32
//   fa      weight fragments of (first_unit, columns [0, min(kchunk, K, 1024))), already in flight (ws_prefetch_first)
//   staged  the caller has staged chunk 0 (always the case when kchunk >= K)
// Writes the outputs (and the K/V page slots, and the per-CTA arg-max candidates) of its units.
template <int NT>
__device__ __forceinline__ void ws_linear(const WsParams& p, __nv_bfloat16* xs, float* red, int kchunk, int first_unit,
                                          int unit_stride, int cand_slot, int cand_count, uint4 (&fa)[WS_MAXBLK][2]) {
    constexpr int MB = 8 * NT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
    const int K = p.K;
    const int kc_full = min(kchunk, K);
    const int xstride = kc_full + WS_XPAD;
    const int nunits = (p.N + 15) / 16;
    const int nchunks = (K + kc_full - 1) / kc_full;
    float best_v = -INFINITY;          // LM head: running arg-max of this CTA's columns (threads tid < M)
    int best_i = 0x7fffffff;
    bool have_first = true;            // fa holds (unit, chunk 0, pass 0)
    for (int unit = first_unit; unit < nunits; unit += unit_stride) {
        const int n0 = unit * 16;
        float acc[NT][4];
#pragma unroll
        for (int j = 0; j < NT; ++j) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.f; }
        // this thread's epilogue operands (bias, residual): requested now, used after the reduction
        constexpr int EPT = (MB * 16 + WS_THREADS - 1) / WS_THREADS;
        float e_bias[EPT], e_aux[EPT];
#pragma unroll
        for (int i = 0; i < EPT; ++i) {
            const int e = tid + WS_THREADS * i, m = e >> 4, n = n0 + (e & 15);
            e_bias[i] = 0.f; e_aux[i] = 0.f;
            if (e < MB * 16 && m < p.M && n < p.N) {
                if (p.bias) e_bias[i] = __bfloat162float(p.bias[n]);
                if (p.epi == 2)
                    e_aux[i] = __bfloat162float(__ushort_as_bfloat16(
                        __ldcg(reinterpret_cast<const unsigned short*>(p.aux + (size_t)m * p.ldaux + n))));
            }
        }
        for (int ch = 0; ch < nchunks; ++ch) {
            const int c0 = ch * kc_full, kc = min(kc_full, K - c0);
            if (nchunks > 1 && !(ch == 0 && unit == first_unit)) {   // chunk 0 of the first unit was staged by the caller
                __syncthreads();
                ws_stage_rows<NT>(p, xs, xstride, c0, kc);
            }
            for (int kb = 0; kb < kc; kb += WS_KPASS) {
                const int nblk = min(WS_KPASS, kc - kb) / 32;
                if (!(have_first && ch == 0 && kb == 0)) ws_load_unit(p, n0, c0 + kb, nblk, fa);
#pragma unroll
                for (int i = 0; i < WS_MAXBLK; ++i) {
                    const int blk = warp + WS_WARPS * i;
                    if (blk < nblk) {
                        const uint4 a = fa[i][0], b8 = fa[i][1];
                        const __nv_bfloat16* xk = xs + kb + 32 * blk + 8 * t;
#pragma unroll
                        for (int j = 0; j < NT; ++j) {
                            const uint4 xb = *reinterpret_cast<const uint4*>(xk + (size_t)(8 * j + g) * xstride);
                            mma_bf16_16816(acc[j], a.x, b8.x, a.y, b8.y, xb.x, xb.y);
                            mma_bf16_16816(acc[j], a.z, b8.z, a.w, b8.w, xb.z, xb.w);
                        }
                    }
                }
            }
        }
        // next unit's first weights: in flight during the reduction and the epilogue
        have_first = false;
        if (unit + unit_stride < nunits) {
            ws_load_unit(p, (unit + unit_stride) * 16, 0, min(kc_full, WS_KPASS) / 32, fa);
            have_first = true;
        }
        // ---- the warps' partial sums meet: red[warp][batch row][weight row]
        float* mine = red + (size_t)warp * MB * 16;
#pragma unroll
        for (int j = 0; j < NT; ++j) {
            mine[(8 * j + 2 * t) * 16 + g] = acc[j][0];
            mine[(8 * j + 2 * t + 1) * 16 + g] = acc[j][1];
            mine[(8 * j + 2 * t) * 16 + g + 8] = acc[j][2];
            mine[(8 * j + 2 * t + 1) * 16 + g + 8] = acc[j][3];
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < EPT; ++i) {
            const int e = tid + WS_THREADS * i;
            if (e >= MB * 16) break;
            const int m = e >> 4, c = e & 15, n = n0 + c;
            float total = 0.f;
#pragma unroll
            for (int w = 0; w < WS_WARPS; ++w) total += red[(size_t)w * MB * 16 + e];
            float v = -INFINITY;
            if (m < p.M && n < p.N) {
                v = bf16_round(total + e_bias[i]);
                if (p.epi == 1) v = bf16_round(p.gelu_tanh ? gelu_tanh(v) : gelu_erf(v));
                else if (p.epi == 2) v = bf16_round(e_aux[i] + v);
                const __nv_bfloat16 tv = __float2bfloat16_rn(v);
                p.out[(size_t)m * p.ldo + n] = tv;
                if (p.pool && n >= p.C) {   // K or V column of the QKV projection: straight into this token's page slot
                    const int cc0 = n - p.C, kv = cc0 / p.C, cc = cc0 - kv * p.C, h = cc / p.D, d = cc - h * p.D;
                    const int page = p.bt[(size_t)m * p.max_pages + p.pos / WS_KVP];
                    p.pool[((((size_t)page * 2 + kv) * p.H + h) * WS_KVP + (p.pos % WS_KVP)) * p.D + d] = tv;
                }
            }
            if (p.amax_val) red[e] = v;          // warp 0's slice doubles as the candidate table (read after the barrier)
        }
        __syncthreads();
        if (p.amax_val) {
            if (tid < p.M) {                     // first index wins ties, like torch.argmax (columns visited in order)
                for (int c = 0; c < 16; ++c) {
                    const float cv = red[tid * 16 + c];
                    if (cv > best_v) { best_v = cv; best_i = n0 + c; }
                }
            }
            __syncthreads();                     // the table is overwritten by the next unit's partial sums
        }
    }
    if (p.amax_val && tid < p.M) {
        p.amax_val[(size_t)tid * cand_count + cand_slot] = best_v;
        p.amax_idx[(size_t)tid * cand_count + cand_slot] = best_i;
    }
}

}  // namespace penroz
