// Small-batch (B <= 8) single-token step: five launches per transformer block instead of nine, every one a
// weight-streaming kernel sized for HBM latency rather than for FLOPs.
//
// At B = 1 a GPT-2 124M token reads 326 MB of weights (51 us at the measured 6.4 TB/s) but each Linear is only
// 1-5 MB: a launch lives or dies by how many 16-byte loads it has in flight in its first microsecond, and by how
// many launches there are. So:
//   * gemv_kernel: one warp owns R weight rows at a time (R x K/256 independent 16-byte loads per lane in flight),
//     the activation rows (<= 8) sit in shared memory as fp32;
//   * what used to be separate launches rides in its prologue or epilogue:
//       prologue  LN      LayerNorm of the activation rows (every CTA recomputes it: K values, nothing next to W)
//                 COMBINE the split-KV attention partials are merged while being loaded (no combine launch)
//       epilogue  bias, GELU, residual; K/V columns of the QKV projection are written straight into the page pool
//                 (no kv_append launch); the LM head leaves a per-CTA arg-max so the final arg-max reads 3k
//                 candidates instead of 50k logits.
// Rounding points are those of the unfused path (and of torch): LN output, Linear output, GELU, residual sum and the
// attention output are each rounded to the storage type.
// Reference behaviour replaced: one iteration of neural_net_model.py:465-469 with the cache branch of
// neural_net_layers.py:83-92 and kv_cache.py:41-68.
#include "common.cuh"
#include "wstream.cuh"
#include <string.h>
#include <stdlib.h>

namespace penroz {

constexpr int KVP = 64;   // page size, same as decode.cu (penroz_kv_page_size)

enum { PRO_PLAIN = 0, PRO_LN = 1, PRO_COMBINE = 2 };

struct GemvParams {
    const void* x; long long ldx;           // activation rows [M, K] (PLAIN, LN)
    const void* W; long long ldw;           // [N, K]
    const void* bias;
    void* out; long long ldo;
    const void* aux; long long ldaux;       // residual rows (epi 2)
    int M, N, K, epi, gelu_tanh;            // epi: 0 store, 1 GELU, 2 residual
    const void* ln_w; const void* ln_b; float ln_eps;         // LN prologue
    const float* part; int splits, H, D;                      // COMBINE prologue: [M][H][splits][D + 2]
    void* pool; const int* bt; int max_pages, pos, C;         // QKV epilogue: columns >= C also go to the page pool
    float* amax_val; int* amax_idx;                           // LM-head epilogue: per-CTA arg-max [M][gridDim.x]
};

template <typename T> struct GemvRows { static constexpr int R = 4; };   // weight rows a warp loads per pass (bf16)
template <> struct GemvRows<float> { static constexpr int R = 2; };

// One CTA (4 warps) owns `rpb` consecutive output columns. The reduction dimension is dealt to the warps in chunks
// of 32 lanes x 16 bytes, so even a 768-column Linear has N/rpb >= 190 CTAs and every lane R x (K / chunk / 4)
// independent 16-byte loads in flight; per-warp partial sums meet in shared memory once, at the end.
// Dynamic shared memory: xs[MB][K] | psum[4][rpb * MB] | (COMBINE) cw, cl [MB][H * splits].
template <typename T, int MB, int PRO>
__global__ void __launch_bounds__(128)
gemv_kernel(const GemvParams p, int rpb) {
    extern __shared__ float xs[];            // activation rows as fp32, already rounded to T where torch rounds
    const T* x = reinterpret_cast<const T*>(p.x);
    const T* W = reinterpret_cast<const T*>(p.W);
    const int K = p.K;
    float* psum = xs + MB * K;
    const int m0 = blockIdx.y * MB;
    const int mrows = min(MB, p.M - m0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- before the producer of the activations has even finished (programmatic dependent launch): pull this CTA's
    // weight rows towards L2. Weights are the one thing this kernel reads that no kernel of the step writes.
    pdl_trigger();
    {
        const int nbp = blockIdx.x * rpb;
        const int rows = min(rpb, p.N - nbp);
        const size_t row_bytes = (size_t)K * sizeof(T);
        const int lines = (int)((row_bytes + 127) / 128);
        if (blockIdx.y == 0)
            for (int e = tid; e < rows * lines; e += 128) {
                const int r = e / lines, c = e - r * lines;
                prefetch_l2(reinterpret_cast<const char*>(W + (size_t)(nbp + r) * p.ldw) + (size_t)c * 128);
            }
    }
    // LayerNorm's gamma / beta are constants too: one 16-byte slice per thread, fetched while the producer still runs
    constexpr int VNP = Vec16<T>::N;
    Vec16<T> g_pre, b_pre;
    const bool ln_pre = (PRO == PRO_LN) && (K % VNP == 0) && (K / VNP <= 128) &&
                        ((reinterpret_cast<uintptr_t>(p.ln_w) & 15) == 0) &&
                        (!p.ln_b || (reinterpret_cast<uintptr_t>(p.ln_b) & 15) == 0);
    if (ln_pre && tid < K / VNP) {
        g_pre.load(reinterpret_cast<const T*>(p.ln_w) + tid * VNP);
        if (p.ln_b) b_pre.load(reinterpret_cast<const T*>(p.ln_b) + tid * VNP);
    }
    pdl_wait();
    // ---- prologue: activation rows -> xs. All global loads of all rows are issued before anything waits on them.
    {
        constexpr int VNX = Vec16<T>::N;
        const bool xvec = (K % VNX == 0) && (p.ldx % VNX == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
        if (PRO != PRO_COMBINE) {
            if (xvec) {
                const int kv = K / VNX;
#pragma unroll 4
                for (int e = tid; e < MB * kv; e += 128) {
                    const int r = e / kv, c = e - r * kv;
                    float* d = xs + r * K + c * VNX;
                    if (r < mrows) {
                        Vec16<T> v;
                        v.load(x + (size_t)(m0 + r) * p.ldx + c * VNX);
#pragma unroll
                        for (int i = 0; i < VNX; ++i) d[i] = v.get(i);
                    } else {
#pragma unroll
                        for (int i = 0; i < VNX; ++i) d[i] = 0.f;
                    }
                }
            } else {
                for (int e = tid; e < MB * K; e += 128) {
                    const int r = e / K, k = e - r * K;
                    xs[e] = (r < mrows) ? to_f32(x[(size_t)(m0 + r) * p.ldx + k]) : 0.f;
                }
            }
        }
    }
    if (PRO == PRO_LN) {   // statistics: one warp per row, from shared memory; normalisation: all threads, 16-byte
        const T* gw = reinterpret_cast<const T*>(p.ln_w);    // loads of gamma / beta
        const T* gb = reinterpret_cast<const T*>(p.ln_b);
        __shared__ float st_mean[MB], st_rstd[MB];
        __syncthreads();
        for (int r = warp; r < mrows; r += 4) {
            const float* xr = xs + r * K;
            float s = 0.f;
            for (int k = lane; k < K; k += 32) s += xr[k];
            const float mean = warp_sum(s) / (float)K;
            float q = 0.f;
            for (int k = lane; k < K; k += 32) { const float a = xr[k] - mean; q += a * a; }
            const float rstd = rsqrtf(warp_sum(q) / (float)K + p.ln_eps);
            if (lane == 0) { st_mean[r] = mean; st_rstd[r] = rstd; }
        }
        __syncthreads();
        constexpr int VNX = Vec16<T>::N;
        const bool gvec = (K % VNX == 0) && ((reinterpret_cast<uintptr_t>(gw) & 15) == 0) &&
                          (!gb || (reinterpret_cast<uintptr_t>(gb) & 15) == 0);
        if (gvec) {
            const int kv = K / VNX;
#pragma unroll 2
            for (int c = tid; c < kv; c += 128) {
                Vec16<T> gv, bv;
                if (ln_pre) {
                    gv = g_pre; bv = b_pre;
                } else {
                    gv.load(gw + c * VNX);
                    if (gb) bv.load(gb + c * VNX);
                }
                for (int r = 0; r < mrows; ++r) {
                    float* d = xs + r * K + c * VNX;
                    const float mean = st_mean[r], rstd = st_rstd[r];
#pragma unroll
                    for (int i = 0; i < VNX; ++i)
                        d[i] = to_f32(from_f32<T>((d[i] - mean) * rstd * gv.get(i) + (gb ? bv.get(i) : 0.f)));
                }
            }
        } else {
            for (int e = tid; e < mrows * K; e += 128) {
                const int r = e / K, k = e - r * K;
                xs[e] = to_f32(from_f32<T>((xs[e] - st_mean[r]) * st_rstd[r] * to_f32(gw[k]) + (gb ? to_f32(gb[k]) : 0.f)));
            }
        }
    } else if (PRO == PRO_COMBINE) {   // K == H * D: softmax weight of every (row, head, split), then the merge
        const int D = p.D, S = p.splits, HS = p.H * p.splits;
        float* cw = psum + 4 * rpb * MB;     // [MB][H][S] split max -> weight
        float* cl = cw + MB * HS;            // [MB][H][S] split sum
        for (int e = tid; e < MB * HS; e += 128) {
            const int r = e / HS;
            const float* pp = p.part + ((size_t)(m0 + r) * HS + (e - r * HS)) * (D + 2) + D;
            cw[e] = (r < mrows) ? pp[0] : -INFINITY;
            cl[e] = (r < mrows) ? pp[1] : 0.f;
        }
        __syncthreads();
        for (int e = tid; e < mrows * p.H; e += 128) {
            float* c = cw + (size_t)e * S;
            const float* ls = cl + (size_t)e * S;
            float m = -INFINITY;
            for (int s = 0; s < S; ++s) m = fmaxf(m, c[s]);
            float l = 0.f;
            for (int s = 0; s < S; ++s) {
                const float w = (c[s] == -INFINITY) ? 0.f : __expf(c[s] - m);
                l += ls[s] * w;
                c[s] = w;
            }
            const float inv = 1.f / l;
            for (int s = 0; s < S; ++s) c[s] *= inv;
        }
        __syncthreads();
        for (int e = tid; e < MB * K; e += 128) {
            const int r = e / K, c = e - r * K;
            float v = 0.f;
            if (r < mrows) {
                const int h = c / D, d = c - h * D;
                const float* pp = p.part + ((size_t)(m0 + r) * p.H + h) * S * (D + 2) + d;
                const float* w = cw + ((size_t)r * p.H + h) * S;
                float o = 0.f;
#pragma unroll 8
                for (int s = 0; s < S; ++s) o = fmaf(pp[(size_t)s * (D + 2)], w[s], o);
                v = to_f32(from_f32<T>(o));
            }
            xs[e] = v;
        }
    }
    __syncthreads();

    constexpr int R = GemvRows<T>::R;
    constexpr int VN = Vec16<T>::N;
    const int nb = blockIdx.x * rpb;         // this CTA's first output column
    const bool vec = (K % VN == 0) && (p.ldw % VN == 0) && ((reinterpret_cast<uintptr_t>(W) & 15) == 0);
    const int span = vec ? 32 * VN : 32;     // reduction elements one warp covers per chunk
    const int nchunks = (K + span - 1) / span;
#pragma unroll 2
    for (int g = 0; g < rpb; g += R) {
        float acc[R][MB];
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int m = 0; m < MB; ++m) acc[r][m] = 0.f;
        if (nb + g < p.N) {
            if (vec) {
                for (int c = warp; c < nchunks; c += 4) {
                    const int k = c * span + lane * VN;
                    if (k < K) {
                        Vec16<T> w[R];
#pragma unroll
                        for (int r = 0; r < R; ++r)     // tail rows are clamped, computed and dropped
                            w[r].load(W + (size_t)min(nb + g + r, p.N - 1) * p.ldw + k);
#pragma unroll
                        for (int i = 0; i < VN; i += 4) {
#pragma unroll
                            for (int m = 0; m < MB; ++m) {
                                const float4 xv = *reinterpret_cast<const float4*>(xs + m * K + k + i);
#pragma unroll
                                for (int r = 0; r < R; ++r) {
                                    acc[r][m] = fmaf(xv.x, w[r].get(i), acc[r][m]);
                                    acc[r][m] = fmaf(xv.y, w[r].get(i + 1), acc[r][m]);
                                    acc[r][m] = fmaf(xv.z, w[r].get(i + 2), acc[r][m]);
                                    acc[r][m] = fmaf(xv.w, w[r].get(i + 3), acc[r][m]);
                                }
                            }
                        }
                    }
                }
            } else {
                for (int c = warp; c < nchunks; c += 4) {
                    const int k = c * 32 + lane;
                    if (k < K) {
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            const float wv = to_f32(W[(size_t)min(nb + g + r, p.N - 1) * p.ldw + k]);
#pragma unroll
                            for (int m = 0; m < MB; ++m) acc[r][m] = fmaf(xs[m * K + k], wv, acc[r][m]);
                        }
                    }
                }
            }
        }
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int m = 0; m < MB; ++m) acc[r][m] = warp_sum(acc[r][m]);
        float my = 0.f;      // lane j keeps partial (r, m) = (j / MB, j % MB); R * MB <= 32
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int m = 0; m < MB; ++m)
                if (lane == r * MB + m) my = acc[r][m];
        if (lane < R * MB) psum[(warp * rpb + g) * MB + lane] = my;
    }
    __syncthreads();

    // thread t finishes output (column nb + t / MB, row m0 + t % MB)
    float* cand = psum;      // reused below for the arg-max candidates (after a barrier)
    float v = -INFINITY;
    int n = 0x7fffffff;
    const bool owner = tid < rpb * MB;
    bool live = false;
    if (owner) {
        const int col = tid / MB, m_out = tid - col * MB;
        n = nb + col;
        live = (n < p.N) && (m_out < mrows);
        const float total = (psum[tid] + psum[rpb * MB + tid]) + (psum[2 * rpb * MB + tid] + psum[3 * rpb * MB + tid]);
        if (live) {
            const int m = m0 + m_out;
            const T* bias = reinterpret_cast<const T*>(p.bias);
            v = to_f32(from_f32<T>(total + (bias ? to_f32(bias[n]) : 0.f)));
            if (p.epi == 1) v = to_f32(from_f32<T>(p.gelu_tanh ? gelu_tanh(v) : gelu_erf(v)));
            else if (p.epi == 2) v = to_f32(from_f32<T>(to_f32(reinterpret_cast<const T*>(p.aux)[(size_t)m * p.ldaux + n]) + v));
            const T tv = from_f32<T>(v);
            reinterpret_cast<T*>(p.out)[(size_t)m * p.ldo + n] = tv;
            if (p.pool && n >= p.C) {   // K or V column of the QKV projection: straight into this token's page slot
                const int c = n - p.C, kv = c / p.C, cc = c - kv * p.C, h = cc / p.D, d = cc - h * p.D;
                const int page = p.bt[(size_t)m * p.max_pages + p.pos / KVP];
                reinterpret_cast<T*>(p.pool)[((((size_t)page * 2 + kv) * p.H + h) * KVP + (p.pos % KVP)) * p.D + d] = tv;
            }
        } else {
            v = -INFINITY;
        }
    }
    if (p.amax_val) {   // per-CTA arg-max per activation row; first index wins ties, like torch.argmax
        __syncthreads();
        if (owner) cand[tid] = v;
        __syncthreads();
        if (tid < mrows) {
            float bv = -INFINITY;
            int bi = 0x7fffffff;
            for (int col = 0; col < rpb; ++col) {
                const float cv = cand[col * MB + tid];
                if (cv > bv) { bv = cv; bi = nb + col; }
            }
            p.amax_val[(size_t)(m0 + tid) * gridDim.x + blockIdx.x] = bv;
            p.amax_idx[(size_t)(m0 + tid) * gridDim.x + blockIdx.x] = bi;
        }
    }
}

// Partial attention of one query row against keys [k0, k1): lanes are laid out as (key slot, 16-byte slice of the
// head dimension), so a K or V row is one coalesced 128-byte read by D/VN lanes and every thread works in both the
// score and the value pass. part layout [B][H][splits][D + 2] = {unnormalised output, row max, row sum}.
// CLUSTER: the `splits` CTAs of one head form a thread-block cluster; each leaves {output, max, sum} in its own shared
// memory, and after one cluster barrier rank 0 merges them through distributed shared memory and writes the head's
// attention output -- the partials never touch global memory and no combine pass exists anywhere. Otherwise the
// partials go to `part` and the consumer (gemv_kernel's COMBINE prologue) merges them.
constexpr int kAttnWarps = 8;   // warps per CTA of the decode-attention kernel
// key rows whose loads one lane has in flight together: 4, or 8 when the launch has few CTAs (measured at 124M bf16:
// batch 1 token step 360 -> 346 us with 8, batch 16 / 64 654 -> 797 / 1267 -> 1347 us)

template <typename T, bool CLUSTER, int NP = 4>
__global__ void __launch_bounds__(kAttnWarps * 32)
attn_decode_split2_kernel(const T* __restrict__ qkv, const T* __restrict__ pool, const int* __restrict__ bt,
                          int max_pages, float* __restrict__ part, int H, int D, int t_len, int chunk, float scale,
                          T* __restrict__ out_final) {
    extern __shared__ float sm[];   // s[chunk] | o[NW][D] | res[D + 2] | page ids
    __shared__ float red[32];
    constexpr int VN = Vec16<T>::N, NW = kAttnWarps;
    float* ss = sm;
    float* osm = sm + chunk;
    int* pg = reinterpret_cast<int*>(osm + NW * D + D + 2);
    const int split = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int C = H * D;
    const int k0 = split * chunk, k1 = min(t_len, k0 + chunk);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int lpk = D / VN, kpw = 32 / lpk;          // lanes per key row, key rows per warp pass
    const int sub = lane / lpk, dl = (lane - sub * lpk) * VN;
    pdl_trigger();
    // the block table does not change during a token step: read this CTA's page ids before waiting for the producer
    const int pg0 = k0 / KVP;
    for (int i = threadIdx.x; i <= (k1 - 1) / KVP - pg0 && k1 > k0; i += blockDim.x) pg[i] = bt[(size_t)b * max_pages + pg0 + i];
    pdl_wait();
    __syncthreads();
    float qv[VN];
    {
        Vec16<T> q;
        q.load(qkv + (size_t)b * 3 * C + h * D + dl);
#pragma unroll
        for (int i = 0; i < VN; ++i) qv[i] = q.get(i) * scale;
    }
    auto row = [&](int kv, int j) {
        return pool + ((((size_t)pg[j / KVP - pg0] * 2 + kv) * H + h) * KVP + (j % KVP)) * D + dl;
    };
    const int stride = NW * kpw;
    float lm = -INFINITY;
    for (int base = k0 + warp * kpw; base < k1; base += NP * stride) {   // warp-uniform trip count
        Vec16<T> kx[NP];
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const int j = base + u * stride + sub;
            if (j < k1) kx[u].load(row(0, j));
        }
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const int j = base + u * stride + sub;
            float a = 0.f;
            if (j < k1) {
#pragma unroll
                for (int i = 0; i < VN; ++i) a = fmaf(qv[i], kx[u].get(i), a);
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
        Vec16<T> vx[NP];
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const int j = base + u * stride + sub;
            if (j < k1) vx[u].load(row(1, j));
        }
#pragma unroll
        for (int u = 0; u < NP; ++u) {
            const int j = base + u * stride + sub;
            if (j < k1) {
                const float pj = __expf(ss[j - k0] - m);
                if (dl == 0) ls += pj;
#pragma unroll
                for (int i = 0; i < VN; ++i) o[i] = fmaf(pj, vx[u].get(i), o[i]);
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
    if (!CLUSTER) {
        float* out = part + (((size_t)b * H + h) * gridDim.x + split) * (D + 2);
        for (int d = threadIdx.x; d < D; d += blockDim.x) out[d] = folded(d);
        if (threadIdx.x == 0) {
            out[D] = (k1 > k0) ? m : -INFINITY;
            out[D + 1] = (k1 > k0) ? l : 0.f;
        }
        return;
    }
    float* res = osm + NW * D;   // this CTA's {o[D], m, l}
    for (int d = threadIdx.x; d < D; d += blockDim.x) res[d] = folded(d);
    if (threadIdx.x == 0) {
        res[D] = (k1 > k0) ? m : -INFINITY;
        res[D + 1] = (k1 > k0) ? l : 0.f;
    }
    cluster_sync_all();
    if (split == 0) {   // cluster rank 0 (the cluster spans gridDim.x)
        const int S = gridDim.x;   // <= 8
        float* wsm = osm;          // free again: every thread passed the barrier after reading it
        const uint32_t res_addr = smem_u32(res);
        if (warp == 0) {
            const float ms = lane < S ? ld_cluster_f32(mapa_rank(res_addr + 4u * D, lane)) : -INFINITY;
            const float lsv = lane < S ? ld_cluster_f32(mapa_rank(res_addr + 4u * (D + 1), lane)) : 0.f;
            const float mm = warp_max(ms);
            const float w = (ms == -INFINITY) ? 0.f : __expf(ms - mm);
            const float ll = warp_sum(lsv * w);
            wsm[lane] = w / ll;
        }
        __syncthreads();
        for (int d = threadIdx.x; d < D; d += blockDim.x) {
            float acc = 0.f;
            for (int r = 0; r < S; ++r) acc = fmaf(ld_cluster_f32(mapa_rank(res_addr + 4u * d, r)), wsm[r], acc);
            out_final[((size_t)b * H + h) * D + d] = from_f32<T>(acc);
        }
    }
    cluster_sync_all();   // the other CTAs' shared memory must outlive rank 0's reads
}

// arg-max over the per-CTA candidates the LM-head kernel left: one CTA per activation row
__global__ void argmax_candidates_kernel(const float* __restrict__ val, const int* __restrict__ idx, int n,
                                         long long* __restrict__ out) {
    __shared__ float sv[32];
    __shared__ int si[32];
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float v = val[(size_t)blockIdx.x * n + i];
        const int ix = idx[(size_t)blockIdx.x * n + i];
        if (v > best || (v == best && ix < bi)) { best = v; bi = ix; }
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

// ------------------------------------------------------------------ host side
using namespace penroz;

extern "C" int penroz_embed_fwd(const long long*, const void*, const void*, void*, int, int, int, int, int, int, int, int*, void*);

struct penroz_gpt2_layer_t {   // same layout as include/penroz_b200.h
    const void *ln1_w, *ln1_b, *qkv_w, *qkv_b, *proj_w, *proj_b, *ln2_w, *ln2_b, *fc_w, *fc_b, *proj2_w, *proj2_b;
    void* kv_pool;
    float ln1_eps, ln2_eps;
    int gelu_tanh, ffn;
    void* kv_scales;
    int kv_quant, kv_recip;
};

namespace penroz {

constexpr int kAmaxSlots = 16384;   // per-row arg-max candidates the workspace reserves room for

// PENROZ_DECODE_PDL=0 launches the step's kernels strictly one after the other
static bool use_pdl() {
    static const bool on = !(getenv("PENROZ_DECODE_PDL") && !strcmp(getenv("PENROZ_DECODE_PDL"), "0"));
    return on;
}

// CTAs per SM a GEMV grid may reach before a CTA takes more than one pass of columns (PENROZ_GEMV_GRID_CAP)
static int gemv_grid_cap() {
    static const int v = getenv("PENROZ_GEMV_GRID_CAP") ? atoi(getenv("PENROZ_GEMV_GRID_CAP")) : 24;
    return v < 1 ? 1 : v;
}

template <typename T, int MB>
static int launch_gemv_mb(const GemvParams& p, int pro, cudaStream_t st) {
    constexpr int R = GemvRows<T>::R;
    // columns per CTA: as few as keeps the grid at a few CTAs per SM (every CTA redoes the prologue), at most 128 / MB
    // (one finishing thread per output)
    int rpb = R;
    const int cap = 128 / MB;
    while (rpb * 2 <= cap && (p.N + rpb - 1) / rpb > gemv_grid_cap() * num_sms()) rpb *= 2;
    const size_t smem = ((size_t)MB * p.K + (size_t)4 * rpb * MB + (pro == PRO_COMBINE ? (size_t)2 * MB * p.H * p.splits : 0)) * sizeof(float);
    PENROZ_REQUIRE(smem <= 200 * 1024, "gemv: K=%d needs %zu bytes of shared memory", p.K, smem);
    dim3 grid((p.N + rpb - 1) / rpb, (p.M + MB - 1) / MB);
#define PENROZ_GEMV(PRO)                                                                                          \
    {                                                                                                             \
        auto k = gemv_kernel<T, MB, PRO>;                                                                         \
        if (smem > 48 * 1024) {                                                                                   \
            cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);      \
            if (e != cudaSuccess) { set_error("gemv: shared memory %zu: %s", smem, cudaGetErrorString(e)); return PENROZ_ERR_CUDA; } \
        }                                                                                                         \
        cudaLaunchConfig_t cfg = {};                                                                              \
        cfg.gridDim = grid; cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = st;              \
        cudaLaunchAttribute at[1];                                                                                \
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                            \
        at[0].val.programmaticStreamSerializationAllowed = 1;                                                     \
        cfg.attrs = at; cfg.numAttrs = use_pdl() ? 1 : 0;                                                         \
        cudaError_t le = cudaLaunchKernelEx(&cfg, k, p, rpb);                                                     \
        if (le != cudaSuccess) { set_error("gemv launch: %s", cudaGetErrorString(le)); return PENROZ_ERR_CUDA; }  \
    }
    if (pro == PRO_LN) PENROZ_GEMV(PRO_LN)
    else if (pro == PRO_COMBINE) PENROZ_GEMV(PRO_COMBINE)
    else PENROZ_GEMV(PRO_PLAIN)
#undef PENROZ_GEMV
    count_launch();
    return check_launch("gemv");
}

// number of CTAs (= arg-max candidates per row) launch_gemv_mb will use for N columns
template <typename T, int MB>
static int gemv_grid_x(int N) {
    constexpr int R = GemvRows<T>::R;
    int rpb = R;
    const int cap = 128 / MB;
    while (rpb * 2 <= cap && (N + rpb - 1) / rpb > gemv_grid_cap() * num_sms()) rpb *= 2;
    return (N + rpb - 1) / rpb;
}
template <typename T>
static int gemv_candidates(int M, int N) {
    if (M <= 1) return gemv_grid_x<T, 1>(N);
    if (M <= 2) return gemv_grid_x<T, 2>(N);
    if (M <= 4) return gemv_grid_x<T, 4>(N);
    return gemv_grid_x<T, 8>(N);
}

template <typename T>
static int launch_gemv(const GemvParams& p, int pro, cudaStream_t st) {
    if (p.M <= 1) return launch_gemv_mb<T, 1>(p, pro, st);
    if (p.M <= 2) return launch_gemv_mb<T, 2>(p, pro, st);
    if (p.M <= 4) return launch_gemv_mb<T, 4>(p, pro, st);
    return launch_gemv_mb<T, 8>(p, pro, st);
}

// bf16 Linears of the step go through the weight-streaming kernel (wstream.cu: 16-row units over every SM, mma.sync,
// any batch up to 64); fp32 models and PENROZ_DECODE_WSTREAM=0 keep the FP32-pipe GEMV above (batch <= 8).
static bool use_wstream() {
    static const bool on = !(getenv("PENROZ_DECODE_WSTREAM") && !strcmp(getenv("PENROZ_DECODE_WSTREAM"), "0"));
    return on;
}
template <typename T> struct IsBf16 { static constexpr bool value = false; };
template <> struct IsBf16<__nv_bfloat16> { static constexpr bool value = true; };

template <typename T>
static int launch_linear(const GemvParams& p, int pro, cudaStream_t st) {
    if (IsBf16<T>::value && use_wstream() && pro != PRO_COMBINE && wstream_supported(p.M, p.N, p.K, BF16) &&
        (pro != PRO_LN || p.K <= 2048)) {
        WsParams w{};
        w.x = (const __nv_bfloat16*)p.x; w.ldx = p.ldx; w.W = (const __nv_bfloat16*)p.W; w.ldw = p.ldw;
        w.bias = (const __nv_bfloat16*)p.bias; w.out = (__nv_bfloat16*)p.out; w.ldo = p.ldo;
        w.aux = (const __nv_bfloat16*)p.aux; w.ldaux = p.ldaux;
        w.M = p.M; w.N = p.N; w.K = p.K; w.epi = p.epi; w.gelu_tanh = p.gelu_tanh;
        w.ln = pro == PRO_LN; w.ln_w = (const __nv_bfloat16*)p.ln_w; w.ln_b = (const __nv_bfloat16*)p.ln_b; w.ln_eps = p.ln_eps;
        w.pool = (__nv_bfloat16*)p.pool; w.bt = p.bt; w.max_pages = p.max_pages; w.pos = p.pos; w.C = p.C; w.H = p.H; w.D = p.D;
        w.amax_val = p.amax_val; w.amax_idx = p.amax_idx;
        return launch_wstream(w, st);
    }
    PENROZ_REQUIRE(p.M <= 8, "decode step: %d activation rows need the weight-streaming kernel (bf16, K %% 32 == 0)", p.M);
    return launch_gemv<T>(p, pro, st);
}
template <typename T>
static int linear_candidates(int M, int N, int K) {
    if (IsBf16<T>::value && use_wstream() && wstream_supported(M, N, K, BF16) && K <= 2048) return wstream_grid(M, N);
    return gemv_candidates<T>(M, N);
}

static int fused_splits(int B, int H, int t_len) {
    int want = (2 * num_sms() + B * H - 1) / (B * H);
    const int maxs = (t_len + 31) / 32;
    if (want > maxs) want = maxs;
    if (want > 32) want = 32;
    return want < 1 ? 1 : want;
}

// small batch (measured at 124M bf16: faster than the tensor-core path at B = 8, slower at B = 16) and a head dimension the lane layout of attn_decode_split2_kernel covers
static int small_batch_limit() {   // PENROZ_DECODE_SMALL_MAX overrides the largest batch the fused step takes
    static const int v = getenv("PENROZ_DECODE_SMALL_MAX") ? atoi(getenv("PENROZ_DECODE_SMALL_MAX")) : 8;
    return v;
}
bool decode_small_supported(int B, int C, int H, int dtype) {
    const bool stream = dtype == BF16 && use_wstream() && C % 32 == 0 && !getenv("PENROZ_DECODE_SMALL_MAX") &&
                        !(getenv("PENROZ_DECODE_ATTN") && !strcmp(getenv("PENROZ_DECODE_ATTN"), "split"));
    if (B > (stream ? 64 : small_batch_limit()) || C % H) return false;
    const int D = C / H, vn = (dtype == F32) ? 4 : 8;
    if (D % vn) return false;
    const int lpk = D / vn;
    return lpk >= 1 && lpk <= 32 && (lpk & (lpk - 1)) == 0;
}

// Split-KV attention for ANY batch with the cluster merge (what penroz_attn_decode runs when the head dimension fits
// the lane layout): out [B, H*D]. Returns PENROZ_OK, or 1 when the shape is not covered (the caller falls back to the
// two-kernel form).
int attn_decode_cluster(const void* qkv, const void* pool, const int* block_table, int max_pages, void* out, int B,
                        int H, int D, int t_len, int dtype, cudaStream_t st) {
    const int vn = (dtype == F32) ? 4 : 8;
    if (D % vn) return 1;
    const int lpk = D / vn;
    if (lpk < 1 || lpk > 32 || (lpk & (lpk - 1)) || D < 8) return 1;
    // as many CTAs per head as keeps ~2 waves of 256-thread CTAs in flight (8 per SM), at least 32 keys each
    int splits = 1;
    const long long room = 16LL * num_sms();
    while (splits * 2 <= 8 && splits * 2 * 32 <= t_len + 31 && (long long)B * H * splits * 2 <= room) splits *= 2;
    const int chunk = (t_len + splits - 1) / splits;
    const size_t smem = (size_t)(chunk + (kAttnWarps + 1) * D + 2 + chunk / KVP + 2) * sizeof(float);
    if (smem > 48 * 1024) return 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(splits, H, B);
    cfg.blockDim = dim3(kAttnWarps * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = splits; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    const float scale = 1.0f / sqrtf((float)D);
    cudaError_t e;
    if (dtype == F32)
        e = cudaLaunchKernelEx(&cfg, attn_decode_split2_kernel<float, true>, (const float*)qkv, (const float*)pool,
                               block_table, max_pages, (float*)nullptr, H, D, t_len, chunk, scale, (float*)out);
    else
        e = cudaLaunchKernelEx(&cfg, attn_decode_split2_kernel<__nv_bfloat16, true>, (const __nv_bfloat16*)qkv,
                               (const __nv_bfloat16*)pool, block_table, max_pages, (float*)nullptr, H, D, t_len, chunk,
                               scale, (__nv_bfloat16*)out);
    if (e != cudaSuccess) { set_error("attn_decode cluster launch: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
    count_launch();
    return check_launch("attn_decode_cluster");
}

size_t decode_small_extra_bytes(int B) { return (size_t)B * kAmaxSlots * 8 + 256; }

template <typename T>
static int decode_step_small_t(const long long* tokens, const void* wte, const void* wpe,
                               const penroz_gpt2_layer_t* layers, int n_layer, const void* lnf_w, const void* lnf_b,
                               float lnf_eps, const void* head_w, const void* head_b, const int* block_table,
                               int max_pages, int B, int C, int H, int V, int pos, int dtype, void* xa, void* xb,
                               void* qkv, void* att, void* g, float* part, float* amax_val, int* amax_idx, void* logits_out,
                               long long* argmax_out, int* err_flag, cudaStream_t st) {
    const int D = C / H;
    PENROZ_REQUIRE(pos / KVP < max_pages, "gpt2_decode_step: position %d exceeds the block table (%d pages)", pos, max_pages);
    int rc = penroz_embed_fwd(tokens, wte, wpe, xa, B, 1, C, V, pos, dtype, dtype, err_flag, st);
    if (rc) return rc;
    const int t_len = pos + 1;
    // one cluster of up to 8 CTAs per head (32+ keys each); PENROZ_DECODE_ATTN=split keeps the partials in global memory
    static const bool use_cluster = !(getenv("PENROZ_DECODE_ATTN") && !strcmp(getenv("PENROZ_DECODE_ATTN"), "split"));
    int splits = fused_splits(B, H, t_len);
    if (use_cluster) {
        int s2 = 1;
        const long long room = 16LL * num_sms();
        while (s2 * 2 <= 8 && s2 * 2 * 32 <= t_len + 31 && (long long)B * H * s2 * 2 <= room) s2 *= 2;
        splits = s2;
    }
    const int chunk = (t_len + splits - 1) / splits;
    const size_t attn_smem = (size_t)(chunk + (kAttnWarps + 1) * D + 2 + chunk / KVP + 2) * sizeof(float);
    PENROZ_REQUIRE(attn_smem <= 48 * 1024, "gpt2_decode_step: context chunk %d too long for shared memory", chunk);
    const float scale = 1.0f / sqrtf((float)D);
    void *x = xa, *xo = xb;
    for (int l = 0; l < n_layer; ++l) {
        const penroz_gpt2_layer_t& L = layers[l];
        GemvParams p{};
        // LN1 + QKV projection; K and V columns land in the page pool
        p.x = x; p.ldx = C; p.W = L.qkv_w; p.ldw = C; p.bias = L.qkv_b; p.out = qkv; p.ldo = 3 * C;
        p.M = B; p.N = 3 * C; p.K = C; p.epi = 0;
        p.ln_w = L.ln1_w; p.ln_b = L.ln1_b; p.ln_eps = L.ln1_eps;
        p.pool = L.kv_pool; p.bt = block_table; p.max_pages = max_pages; p.pos = pos; p.C = C; p.H = H; p.D = D;
        if ((rc = launch_linear<T>(p, PRO_LN, st))) return rc;
        // split-KV attention over the pages (current token included)
        {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = dim3(splits, H, B);
            cfg.blockDim = dim3(kAttnWarps * 32);
            cfg.dynamicSmemBytes = attn_smem;
            cfg.stream = st;
            cudaLaunchAttribute at[2];
            int na = 0;
            if (use_pdl()) {
                at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                at[na].val.programmaticStreamSerializationAllowed = 1;
                ++na;
            }
            if (use_cluster) {
                at[na].id = cudaLaunchAttributeClusterDimension;
                at[na].val.clusterDim.x = splits; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1;
                ++na;
            }
            cfg.attrs = at; cfg.numAttrs = na;
            cudaError_t e;
            if (use_cluster && (long long)B * H * splits <= 2LL * num_sms())
                e = cudaLaunchKernelEx(&cfg, attn_decode_split2_kernel<T, true, 8>, (const T*)qkv, (const T*)L.kv_pool,
                                       block_table, max_pages, part, H, D, t_len, chunk, scale, (T*)att);
            else if (use_cluster)
                e = cudaLaunchKernelEx(&cfg, attn_decode_split2_kernel<T, true>, (const T*)qkv, (const T*)L.kv_pool,
                                       block_table, max_pages, part, H, D, t_len, chunk, scale, (T*)att);
            else
                e = cudaLaunchKernelEx(&cfg, attn_decode_split2_kernel<T, false>, (const T*)qkv, (const T*)L.kv_pool,
                                       block_table, max_pages, part, H, D, t_len, chunk, scale, (T*)nullptr);
            if (e != cudaSuccess) { set_error("attn_decode launch: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
        }
        count_launch();
        if ((rc = check_launch("attn_decode_split2"))) return rc;
        // output projection + residual (merging the partials on the way in if the attention kernel did not)
        p = GemvParams{};
        p.x = att; p.ldx = C;
        p.W = L.proj_w; p.ldw = C; p.bias = L.proj_b; p.out = xo; p.ldo = C; p.aux = x; p.ldaux = C;
        p.M = B; p.N = C; p.K = C; p.epi = 2; p.part = part; p.splits = splits; p.H = H; p.D = D;
        if ((rc = launch_linear<T>(p, use_cluster ? PRO_PLAIN : PRO_COMBINE, st))) return rc;
        // LN2 + fc + GELU
        p = GemvParams{};
        p.x = xo; p.ldx = C; p.W = L.fc_w; p.ldw = C; p.bias = L.fc_b; p.out = g; p.ldo = L.ffn;
        p.M = B; p.N = L.ffn; p.K = C; p.epi = 1; p.gelu_tanh = L.gelu_tanh;
        p.ln_w = L.ln2_w; p.ln_b = L.ln2_b; p.ln_eps = L.ln2_eps;
        if ((rc = launch_linear<T>(p, PRO_LN, st))) return rc;
        // proj2 + residual
        p = GemvParams{};
        p.x = g; p.ldx = L.ffn; p.W = L.proj2_w; p.ldw = L.ffn; p.bias = L.proj2_b; p.out = x; p.ldo = C;
        p.aux = xo; p.ldaux = C; p.M = B; p.N = C; p.K = L.ffn; p.epi = 2;
        if ((rc = launch_linear<T>(p, PRO_PLAIN, st))) return rc;
    }
    // final LN + LM head (+ per-CTA arg-max candidates)
    GemvParams p{};
    p.x = x; p.ldx = C; p.W = head_w; p.ldw = C; p.bias = head_b; p.out = logits_out; p.ldo = V;
    p.M = B; p.N = V; p.K = C; p.epi = 0; p.ln_w = lnf_w; p.ln_b = lnf_b; p.ln_eps = lnf_eps;
    const int nblk = linear_candidates<T>(B, V, C);
    const bool fused_amax = argmax_out && nblk <= kAmaxSlots;
    if (fused_amax) { p.amax_val = amax_val; p.amax_idx = amax_idx; }
    if ((rc = launch_linear<T>(p, PRO_LN, st))) return rc;
    if (fused_amax) {
        argmax_candidates_kernel<<<B, 256, 0, st>>>(amax_val, amax_idx, nblk, argmax_out);
        count_launch();
        return check_launch("argmax_candidates");
    }
    return PENROZ_OK;
}

int decode_step_small(const long long* tokens, const void* wte, const void* wpe, const penroz_gpt2_layer_t* layers,
                      int n_layer, const void* lnf_w, const void* lnf_b, float lnf_eps, const void* head_w,
                      const void* head_b, const int* block_table, int max_pages, int B, int C, int H, int V, int pos,
                      int dtype, void* xa, void* xb, void* qkv, void* att, void* g, float* part, void* extra,
                      void* logits_out, long long* argmax_out, int* err_flag, cudaStream_t st) {
    float* amax_val = reinterpret_cast<float*>(extra);
    int* amax_idx = reinterpret_cast<int*>(amax_val + (size_t)B * kAmaxSlots);
    if (dtype == F32)
        return decode_step_small_t<float>(tokens, wte, wpe, layers, n_layer, lnf_w, lnf_b, lnf_eps, head_w, head_b,
                                          block_table, max_pages, B, C, H, V, pos, dtype, xa, xb, qkv, att, g, part,
                                          amax_val, amax_idx, logits_out, argmax_out, err_flag, st);
    return decode_step_small_t<__nv_bfloat16>(tokens, wte, wpe, layers, n_layer, lnf_w, lnf_b, lnf_eps, head_w, head_b,
                                              block_table, max_pages, B, C, H, V, pos, dtype, xa, xb, qkv, att, g, part,
                                              amax_val, amax_idx, logits_out, argmax_out, err_flag, st);
}

}  // namespace penroz
