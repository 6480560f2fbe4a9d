// Memory-bound row-wise kernels: embedding gather+add, LayerNorm fwd/bwd (+residual-gradient add, +fused
// bias-gradient column sums), GELU fwd/bwd, column sums, casts.
//
// Reference call sites these replace (all are torch library calls in the reference):
//   Summation(Embedding, PositionEmbedding)   neural_net_layers.py:98-127
//   nn.LayerNorm (eps 1e-5)                   mappers.py:31,155,162,171
//   nn.GELU (erf | tanh)                      mappers.py:25,139,164
//   ResidualConnection                        neural_net_layers.py:130-135
// All of them are HBM-bound: one warp per row, 16-byte vector accesses, warp-shuffle reductions.
#include "common.cuh"
#include <stdlib.h>
#include <string.h>

namespace penroz {

// 4 consecutive elements as float4, whatever the storage type
template <typename T> __device__ __forceinline__ float4 load4(const T* p);
template <> __device__ __forceinline__ float4 load4<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
template <> __device__ __forceinline__ float4 load4<__nv_bfloat16>(const __nv_bfloat16* p) {
    uint2 w = *reinterpret_cast<const uint2*>(p);
    float2 a = unpack_bf16x2(w.x), b = unpack_bf16x2(w.y);
    return make_float4(a.x, a.y, b.x, b.y);
}
template <typename T> __device__ __forceinline__ void store4(T* p, float4 v);
template <> __device__ __forceinline__ void store4<float>(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
template <> __device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, float4 v) {
    uint2 w;
    w.x = pack_bf16x2(v.x, v.y);
    w.y = pack_bf16x2(v.z, v.w);
    *reinterpret_cast<uint2*>(p) = w;
}

// ------------------------------------------------------------------ embedding
// out[b,t,:] = wte[idx[b,t],:] + wpe[t + pos_offset,:]   (sum formed in fp32, rounded once to TO)
template <typename TW, typename TO>
__global__ void embed_fwd_kernel(const long long* __restrict__ idx, const TW* __restrict__ wte,
                                 const TW* __restrict__ wpe, TO* __restrict__ out, int rows, int T, int C, int V,
                                 int pos_offset, int* __restrict__ err_flag) {
    const int warps_per_block = blockDim.x >> 5;
    const int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    long long tok = idx[row];
    if (tok < 0 || tok >= V) {  // torch raises an index error here; we flag it and clamp
        if (lane == 0 && err_flag) atomicExch(err_flag, 1);
        tok = tok < 0 ? 0 : V - 1;
    }
    const int t = row % T + pos_offset;
    const TW* a = wte + (size_t)tok * C;
    const TW* b = wpe + (size_t)t * C;
    TO* o = out + (size_t)row * C;
    if ((C & 3) == 0) {
        for (int i = lane * 4; i < C; i += 128) {
            float4 x = load4(a + i), y = load4(b + i);
            store4(o + i, make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w));
        }
    } else {
        for (int i = lane; i < C; i += 32) o[i] = from_f32<TO>(to_f32(a[i]) + to_f32(b[i]));
    }
}

// dwte[idx[b,t],:] += dx[b,t,:] (fp32 red.add: rows collide when a token repeats)
__global__ void embed_bwd_wte_kernel(const long long* __restrict__ idx, const float* __restrict__ dx,
                                     float* __restrict__ dwte, int rows, int C, int V) {
    const int warps_per_block = blockDim.x >> 5;
    const int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    long long tok = idx[row];
    if (tok < 0 || tok >= V) return;
    const float* g = dx + (size_t)row * C;
    float* o = dwte + (size_t)tok * C;
    for (int i = lane; i < C; i += 32) atomicAdd(o + i, g[i]);
}

// dwpe[t + off,:] += sum_b dx[b,t,:]  (deterministic: one thread per (t, c4), loop over the batch)
__global__ void embed_bwd_wpe_kernel(const float* __restrict__ dx, float* __restrict__ dwpe, int B, int T, int C,
                                     int pos_offset) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= T * C) return;
    const int t = i / C, c = i % C;
    float s = 0.f;
    for (int b = 0; b < B; ++b) s += dx[((size_t)b * T + t) * C + c];
    dwpe[(size_t)(t + pos_offset) * C + c] += s;
}

// ------------------------------------------------------------------ LayerNorm forward
// One warp per row. Statistics in fp32 (two-pass: mean, then centred sum of squares), like ATen.
template <typename TX, typename TW, typename TO>
__global__ void layernorm_fwd_kernel(const TX* __restrict__ x, const TW* __restrict__ gamma,
                                     const TW* __restrict__ beta, TO* __restrict__ out, float* __restrict__ mean_out,
                                     float* __restrict__ rstd_out, int rows, int C, float eps) {
    const int warps_per_block = blockDim.x >> 5;
    const int row = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const TX* xr = x + (size_t)row * C;
    TO* o = out + (size_t)row * C;
    const bool vec = (C & 3) == 0;
    float s = 0.f;
    if (vec) {
        for (int i = lane * 4; i < C; i += 128) { float4 v = load4(xr + i); s += (v.x + v.y) + (v.z + v.w); }
    } else {
        for (int i = lane; i < C; i += 32) s += to_f32(xr[i]);
    }
    const float mean = warp_sum(s) / (float)C;
    float q = 0.f;
    if (vec) {
        for (int i = lane * 4; i < C; i += 128) {
            float4 v = load4(xr + i);
            float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
            q += (a * a + b * b) + (c * c + d * d);
        }
    } else {
        for (int i = lane; i < C; i += 32) { float a = to_f32(xr[i]) - mean; q += a * a; }
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)C + eps);
    if (lane == 0) {
        if (mean_out) mean_out[row] = mean;
        if (rstd_out) rstd_out[row] = rstd;
    }
    if (vec) {
        for (int i = lane * 4; i < C; i += 128) {
            float4 v = load4(xr + i), g = load4(gamma + i);
            float4 b = beta ? load4(beta + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            store4(o + i, make_float4((v.x - mean) * rstd * g.x + b.x, (v.y - mean) * rstd * g.y + b.y,
                                      (v.z - mean) * rstd * g.z + b.z, (v.w - mean) * rstd * g.w + b.w));
        }
    } else {
        for (int i = lane; i < C; i += 32)
            o[i] = from_f32<TO>((to_f32(xr[i]) - mean) * rstd * to_f32(gamma[i]) + (beta ? to_f32(beta[i]) : 0.f));
    }
}

// Same arithmetic with the row held in registers (J = ceil(C / 128) float4 per lane): one pass over global memory and
// every load of the row in flight at once, instead of three dependent passes.
template <typename TX, typename TW, typename TO, int J>
__global__ void __launch_bounds__(256)
layernorm_fwd_reg_kernel(const TX* __restrict__ x, const TW* __restrict__ gamma, const TW* __restrict__ beta,
                         TO* __restrict__ out, float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows,
                         int C, float eps) {
    pdl_trigger();   // a dependent-launch successor (the next GEMM) may set itself up while this grid drains
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const TX* xr = x + (size_t)row * C;
    float4 v[J];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int i = (lane + 32 * j) * 4;
        v[j] = (i < C) ? load4(xr + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
    }
    const float mean = warp_sum(s) / (float)C;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int i = (lane + 32 * j) * 4;
        if (i < C) {
            const float a = v[j].x - mean, b = v[j].y - mean, c = v[j].z - mean, d = v[j].w - mean;
            q += (a * a + b * b) + (c * c + d * d);
        }
    }
    const float rstd = rsqrtf(warp_sum(q) / (float)C + eps);
    if (lane == 0) {
        if (mean_out) mean_out[row] = mean;
        if (rstd_out) rstd_out[row] = rstd;
    }
    TO* o = out + (size_t)row * C;
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int i = (lane + 32 * j) * 4;
        if (i < C) {
            const float4 g = load4(gamma + i);
            const float4 b = beta ? load4(beta + i) : make_float4(0.f, 0.f, 0.f, 0.f);
            store4(o + i, make_float4((v[j].x - mean) * rstd * g.x + b.x, (v[j].y - mean) * rstd * g.y + b.y,
                                      (v[j].z - mean) * rstd * g.z + b.z, (v[j].w - mean) * rstd * g.w + b.w));
        }
    }
}

// ------------------------------------------------------------------ LayerNorm backward
// dx_total = add + rstd * (dy*g - mean(dy*g) - xhat * mean(dy*g*xhat)); written as TX and (optionally) a bf16 copy.
// Column reductions (dgamma = sum dy*xhat, dbeta = sum dy, dcol = sum bf16-or-f32(dx_total)) are accumulated in
// registers by each warp over its rows, combined across the block in shared memory and written to
// scratch[block][3][C]; layernorm_bwd_reduce_kernel then folds the per-block partials into the gradients.
// J = ceil(C / 128): float4 chunks per lane.
template <typename TX, typename TW, typename TG, int J, bool DCOL>
__global__ void __launch_bounds__(256)
layernorm_bwd_kernel(const TG* __restrict__ dy, const TX* __restrict__ x, const TW* __restrict__ gamma,
                     const float* __restrict__ mean, const float* __restrict__ rstd, const TX* __restrict__ add,
                     TX* __restrict__ dx, __nv_bfloat16* __restrict__ dx_bf16, float* __restrict__ scratch, int rows,
                     int C) {
    extern __shared__ float sred[];  // [8 warps][C]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4 ag[J], ab[J], ac[DCOL ? J : 1];
#pragma unroll
    for (int j = 0; j < J; ++j) {
        ag[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        ab[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (DCOL) ac[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float4 gam[J];
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int i = (lane + 32 * j) * 4;
        gam[j] = (i < C) ? load4(gamma + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int row = blockIdx.x * 8 + warp; row < rows; row += gridDim.x * 8) {
        const TG* dyr = dy + (size_t)row * C;
        const TX* xr = x + (size_t)row * C;
        const float mu = mean[row], rs = rstd[row];
        float4 dyv[J], xh[J];
        float s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int i = (lane + 32 * j) * 4;
            if (i < C) {
                dyv[j] = load4(dyr + i);
                float4 xv = load4(xr + i);
                xh[j] = make_float4((xv.x - mu) * rs, (xv.y - mu) * rs, (xv.z - mu) * rs, (xv.w - mu) * rs);
                const float4 g = gam[j];
                const float a = dyv[j].x * g.x, b = dyv[j].y * g.y, c = dyv[j].z * g.z, d = dyv[j].w * g.w;
                s1 += (a + b) + (c + d);
                s2 += (a * xh[j].x + b * xh[j].y) + (c * xh[j].z + d * xh[j].w);
                ag[j].x += dyv[j].x * xh[j].x; ag[j].y += dyv[j].y * xh[j].y;
                ag[j].z += dyv[j].z * xh[j].z; ag[j].w += dyv[j].w * xh[j].w;
                ab[j].x += dyv[j].x; ab[j].y += dyv[j].y; ab[j].z += dyv[j].z; ab[j].w += dyv[j].w;
            }
        }
        s1 = warp_sum(s1) / (float)C;
        s2 = warp_sum(s2) / (float)C;
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int i = (lane + 32 * j) * 4;
            if (i < C) {
                const float4 g = gam[j];
                float4 r;
                r.x = rs * (dyv[j].x * g.x - s1 - xh[j].x * s2);
                r.y = rs * (dyv[j].y * g.y - s1 - xh[j].y * s2);
                r.z = rs * (dyv[j].z * g.z - s1 - xh[j].z * s2);
                r.w = rs * (dyv[j].w * g.w - s1 - xh[j].w * s2);
                if (add) {
                    float4 a = load4(add + (size_t)row * C + i);
                    r.x += a.x; r.y += a.y; r.z += a.z; r.w += a.w;
                }
                store4(dx + (size_t)row * C + i, r);
                if (dx_bf16) store4(dx_bf16 + (size_t)row * C + i, r);
                if (DCOL) {
                    // the bias gradient of the Linear that produced this residual branch sums the gradient
                    // as that Linear sees it: rounded to bf16 when a bf16 copy is requested
                    if (dx_bf16) { r.x = bf16_round(r.x); r.y = bf16_round(r.y); r.z = bf16_round(r.z); r.w = bf16_round(r.w); }
                    ac[j].x += r.x; ac[j].y += r.y; ac[j].z += r.z; ac[j].w += r.w;
                }
            }
        }
    }
    // block combine: quantity by quantity through sred[8][C]
    const int nq = DCOL ? 3 : 2;
    for (int qn = 0; qn < nq; ++qn) {
        __syncthreads();
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int i = (lane + 32 * j) * 4;
            if (i < C) {
                const float4 v = (qn == 0) ? ag[j] : (qn == 1) ? ab[j] : ac[DCOL ? j : 0];
                *reinterpret_cast<float4*>(sred + (size_t)warp * C + i) = v;
            }
        }
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) s += sred[(size_t)w * C + c];
            scratch[((size_t)blockIdx.x * 3 + qn) * C + c] = s;
        }
    }
}

// Two-kernel form used for wide launches (enough rows to fill the GPU): (1) a row kernel that only produces dx
// (few registers -> high occupancy -> HBM-bound), (2) a column kernel that re-reads dy, x and the written gradient
// and accumulates dgamma/dbeta/dcol for 4 columns per thread over a slice of the rows.
template <typename TX, typename TW, typename TG, int J>
__global__ void __launch_bounds__(256, (J <= 6) ? 3 : 1)
layernorm_bwd_rows_kernel(const TG* __restrict__ dy, const TX* __restrict__ x, const TW* __restrict__ gamma,
                          const float* __restrict__ mean, const float* __restrict__ rstd, const TX* __restrict__ add,
                          TX* __restrict__ dx, __nv_bfloat16* __restrict__ dx_bf16, int rows, int C) {
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (row >= rows) return;
    const TG* dyr = dy + (size_t)row * C;
    const TX* xr = x + (size_t)row * C;
    const float mu = mean[row], rs = rstd[row];
    float4 dyg[J], xh[J], ad[J];
    float s1 = 0.f, s2 = 0.f;
    // the residual gradient does not depend on the row statistics: fetch it with dy and x, not after the reductions
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int i = (lane + 32 * j) * 4;
        ad[j] = (add && i < C) ? load4(add + (size_t)row * C + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int i = (lane + 32 * j) * 4;
        if (i < C) {
            const float4 d = load4(dyr + i), xv = load4(xr + i), g = load4(gamma + i);
            xh[j] = make_float4((xv.x - mu) * rs, (xv.y - mu) * rs, (xv.z - mu) * rs, (xv.w - mu) * rs);
            dyg[j] = make_float4(d.x * g.x, d.y * g.y, d.z * g.z, d.w * g.w);
            s1 += (dyg[j].x + dyg[j].y) + (dyg[j].z + dyg[j].w);
            s2 += (dyg[j].x * xh[j].x + dyg[j].y * xh[j].y) + (dyg[j].z * xh[j].z + dyg[j].w * xh[j].w);
        }
    }
    s1 = warp_sum(s1) / (float)C;
    s2 = warp_sum(s2) / (float)C;
#pragma unroll
    for (int j = 0; j < J; ++j) {
        const int i = (lane + 32 * j) * 4;
        if (i < C) {
            float4 r;
            r.x = rs * (dyg[j].x - s1 - xh[j].x * s2);
            r.y = rs * (dyg[j].y - s1 - xh[j].y * s2);
            r.z = rs * (dyg[j].z - s1 - xh[j].z * s2);
            r.w = rs * (dyg[j].w - s1 - xh[j].w * s2);
            r.x += ad[j].x; r.y += ad[j].y; r.z += ad[j].z; r.w += ad[j].w;
            store4(dx + (size_t)row * C + i, r);
            if (dx_bf16) store4(dx_bf16 + (size_t)row * C + i, r);
        }
    }
}

// scratch[blockIdx.y][q][c]: q = 0 sum dy*xhat, 1 sum dy, 2 sum of the written gradient (bf16 copy when present)
template <typename TX, typename TG>
__global__ void __launch_bounds__(256)
layernorm_bwd_cols_kernel(const TG* __restrict__ dy, const TX* __restrict__ x, const float* __restrict__ mean,
                          const float* __restrict__ rstd, const TX* __restrict__ dx,
                          const __nv_bfloat16* __restrict__ dx_bf16, float* __restrict__ scratch, int rows, int C,
                          int want_dcol) {
    __shared__ float4 sm[3][8][32];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = (blockIdx.x * 32 + tx) * 4;
    const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
    const int r0 = blockIdx.y * rows_per, r1 = min(rows, r0 + rows_per);
    float4 ag = make_float4(0.f, 0.f, 0.f, 0.f), ab = ag, ac = ag;
    if (col < C) {
        for (int r = r0 + ty; r < r1; r += 8) {
            const float mu = mean[r], rs = rstd[r];
            const float4 d = load4(dy + (size_t)r * C + col), xv = load4(x + (size_t)r * C + col);
            ag.x += d.x * (xv.x - mu) * rs; ag.y += d.y * (xv.y - mu) * rs;
            ag.z += d.z * (xv.z - mu) * rs; ag.w += d.w * (xv.w - mu) * rs;
            ab.x += d.x; ab.y += d.y; ab.z += d.z; ab.w += d.w;
            if (want_dcol) {
                const float4 g = dx_bf16 ? load4(dx_bf16 + (size_t)r * C + col) : load4(dx + (size_t)r * C + col);
                ac.x += g.x; ac.y += g.y; ac.z += g.z; ac.w += g.w;
            }
        }
    }
    sm[0][ty][tx] = ag; sm[1][ty][tx] = ab; sm[2][ty][tx] = ac;
    __syncthreads();
    if (ty < 3 && col < C) {
        float4 s = sm[ty][0][tx];
        for (int k = 1; k < 8; ++k) { const float4 v = sm[ty][k][tx]; s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
        *reinterpret_cast<float4*>(scratch + ((size_t)blockIdx.y * 3 + ty) * C + col) = s;
    }
}

// One-pass form. The CTA spans the row (thread t owns float4 columns t, t + blockDim, ...), so a thread carries
// 12 VPT column accumulators instead of the 72 of the row-per-warp kernel above, and dy, x and the residual gradient
// are read once (the two-kernel form reads dy and x twice and the written gradient back: 24 bytes per element against
// the 16 the op needs). What keeps HBM busy is not registers but shared memory: R consecutive rows of dy, x and add
// are three contiguous ranges, so one thread fetches them with bulk copies into a ring of S stages (about 100 KB per
// CTA, two CTAs per SM) while the others work on an earlier stage. The two row sums cross the warps through a
// double-buffered shared array: one __syncthreads per R rows, after which the stage consumed before this one is
// refilled. scratch[block][q][c] as above.
template <typename TG, int VPT, int R>
__global__ void __launch_bounds__(256)
layernorm_bwd_staged_kernel(const TG* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ gamma,
                            const float* __restrict__ mean, const float* __restrict__ rstd,
                            const float* __restrict__ add, float* __restrict__ dx, __nv_bfloat16* __restrict__ dx_bf16,
                            float* __restrict__ scratch, int rows, int C, int S, int want_dcol) {
    pdl_trigger();   // a dependent-launch successor (the next GEMM) may set itself up while this grid drains
    extern __shared__ __align__(128) unsigned char ln_smem[];
    __shared__ __align__(16) float red[2][R][2][8];
    __shared__ __align__(8) unsigned long long full_bar[8];
    const int nv = C >> 2;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t dy_bytes = (uint32_t)R * C * sizeof(TG), x_bytes = (uint32_t)R * C * sizeof(float);
    const uint32_t x_off = (dy_bytes + 127u) & ~127u, add_off = x_off + ((x_bytes + 127u) & ~127u);
    const uint32_t stage_bytes = add_off + ((x_bytes + 127u) & ~127u);
    const int nchunks = (rows + R - 1) / R;
    const int mine = ((int)blockIdx.x < nchunks) ? (nchunks - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    auto fill = [&](int i) {     // chunk i of this CTA into stage i % S (one thread)
        const int row0 = ((int)blockIdx.x + i * (int)gridDim.x) * R;
        const int nr = min(R, rows - row0);
        const uint32_t bar = smem_u32(&full_bar[i % S]);
        const uint32_t base = smem_u32(ln_smem) + (uint32_t)(i % S) * stage_bytes;
        const uint32_t nb_dy = (uint32_t)nr * C * sizeof(TG), nb_x = (uint32_t)nr * C * sizeof(float);
        mbar_expect_tx(bar, nb_dy + nb_x + (add ? nb_x : 0u));
        bulk_load_1d(base, dy + (size_t)row0 * C, nb_dy, bar);
        bulk_load_1d(base + x_off, x + (size_t)row0 * C, nb_x, bar);
        if (add) bulk_load_1d(base + add_off, add + (size_t)row0 * C, nb_x, bar);
    };
    if (tid == 0) {
        for (int s = 0; s < S; ++s) mbar_init(smem_u32(&full_bar[s]), 1);
        mbar_fence_init();
        for (int i = 0; i < S && i < mine; ++i) fill(i);
    }
    for (int i = tid; i < 2 * R * 2 * 8; i += blockDim.x) (&red[0][0][0][0])[i] = 0.f;   // absent warps add zero
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 gam[VPT], ag[VPT], ab[VPT], ac[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = tid + k * blockDim.x;
        gam[k] = (v < nv) ? load4(gamma + 4 * v) : zero4;
        ag[k] = zero4; ab[k] = zero4; ac[k] = zero4;
    }
    const float invC = 1.f / (float)C;
    float mu_n[R], rs_n[R];   // row statistics of the next chunk, fetched one iteration ahead
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int row = (int)blockIdx.x * R + r;
        mu_n[r] = (row < rows) ? mean[row] : 0.f;
        rs_n[r] = (row < rows) ? rstd[row] : 0.f;
    }
    __syncthreads();   // barriers initialised, red zeroed
    for (int i = 0; i < mine; ++i) {
        const int row0 = ((int)blockIdx.x + i * (int)gridDim.x) * R;
        const int nr = min(R, rows - row0);
        const int buf = i & 1;
        float mu[R], rs[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            mu[r] = mu_n[r]; rs[r] = rs_n[r];
            const int row = row0 + (int)gridDim.x * R + r;
            mu_n[r] = (row < rows) ? mean[row] : 0.f;
            rs_n[r] = (row < rows) ? rstd[row] : 0.f;
        }
        const unsigned char* st = ln_smem + (size_t)(i % S) * stage_bytes;
        const TG* sdy = reinterpret_cast<const TG*>(st);
        const float* sx = reinterpret_cast<const float*>(st + x_off);
        const float* sadd = reinterpret_cast<const float*>(st + add_off);
        mbar_wait(smem_u32(&full_bar[i % S]), (uint32_t)((i / S) & 1));
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float s1 = 0.f, s2 = 0.f;
            if (r < nr) {
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    const int v = tid + k * blockDim.x;
                    if (v < nv) {
                        const float4 d = load4(sdy + (size_t)r * C + 4 * v), xv = load4(sx + (size_t)r * C + 4 * v);
                        const float4 g = gam[k];
                        const float4 h = make_float4((xv.x - mu[r]) * rs[r], (xv.y - mu[r]) * rs[r],
                                                     (xv.z - mu[r]) * rs[r], (xv.w - mu[r]) * rs[r]);
                        const float a = d.x * g.x, b = d.y * g.y, c = d.z * g.z, e = d.w * g.w;
                        s1 += (a + b) + (c + e);
                        s2 += (a * h.x + b * h.y) + (c * h.z + e * h.w);
                        ag[k].x += d.x * h.x; ag[k].y += d.y * h.y; ag[k].z += d.z * h.z; ag[k].w += d.w * h.w;
                        ab[k].x += d.x; ab[k].y += d.y; ab[k].z += d.z; ab[k].w += d.w;
                    }
                }
            }
            s1 = warp_sum(s1);
            s2 = warp_sum(s2);
            if (lane == 0) { red[buf][r][0][warp] = s1; red[buf][r][1][warp] = s2; }
        }
        __syncthreads();
        // every thread is past its reads of the stage consumed in the previous iteration: refill it
        if (tid == 0 && i >= 1 && i - 1 + S < mine) fill(i - 1 + S);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (r < nr) {
                const float4 a0 = *reinterpret_cast<const float4*>(&red[buf][r][0][0]);
                const float4 a1 = *reinterpret_cast<const float4*>(&red[buf][r][0][4]);
                const float4 b0 = *reinterpret_cast<const float4*>(&red[buf][r][1][0]);
                const float4 b1 = *reinterpret_cast<const float4*>(&red[buf][r][1][4]);
                const float s1 = (((a0.x + a0.y) + (a0.z + a0.w)) + ((a1.x + a1.y) + (a1.z + a1.w))) * invC;
                const float s2 = (((b0.x + b0.y) + (b0.z + b0.w)) + ((b1.x + b1.y) + (b1.z + b1.w))) * invC;
#pragma unroll
                for (int k = 0; k < VPT; ++k) {
                    const int v = tid + k * blockDim.x;
                    if (v < nv) {
                        const float4 d = load4(sdy + (size_t)r * C + 4 * v), xv = load4(sx + (size_t)r * C + 4 * v);
                        const float4 ad = add ? load4(sadd + (size_t)r * C + 4 * v) : zero4;
                        const float4 g = gam[k];
                        float4 o;
                        o.x = rs[r] * (d.x * g.x - s1 - (xv.x - mu[r]) * rs[r] * s2) + ad.x;
                        o.y = rs[r] * (d.y * g.y - s1 - (xv.y - mu[r]) * rs[r] * s2) + ad.y;
                        o.z = rs[r] * (d.z * g.z - s1 - (xv.z - mu[r]) * rs[r] * s2) + ad.z;
                        o.w = rs[r] * (d.w * g.w - s1 - (xv.w - mu[r]) * rs[r] * s2) + ad.w;
                        const size_t off = (size_t)(row0 + r) * C + 4 * v;
                        store4(dx + off, o);
                        if (dx_bf16) {
                            store4(dx_bf16 + off, o);
                            // the bias gradient of the Linear that produced this residual branch sums the gradient
                            // as that Linear sees it: rounded to bf16 when a bf16 copy is requested
                            o.x = bf16_round(o.x); o.y = bf16_round(o.y); o.z = bf16_round(o.z); o.w = bf16_round(o.w);
                        }
                        ac[k].x += o.x; ac[k].y += o.y; ac[k].z += o.z; ac[k].w += o.w;
                    }
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int v = tid + k * blockDim.x;
        if (v < nv) {
            float* s = scratch + (size_t)blockIdx.x * 3 * C + 4 * v;
            *reinterpret_cast<float4*>(s) = ag[k];
            *reinterpret_cast<float4*>(s + C) = ab[k];
            if (want_dcol) *reinterpret_cast<float4*>(s + 2 * (size_t)C) = ac[k];
        }
    }
}

// out_q[c] += sum_blocks scratch[b][q][c]   (fixed summation order: deterministic). 32 columns (8 float4 lanes) x 128
// partial-sum lanes per CTA, so the several hundred per-block partials of a column are four or five independent 16-byte
// loads per thread -- the kernel is a chain of load latencies, not bandwidth.
__global__ void __launch_bounds__(1024)
layernorm_bwd_reduce_kernel(const float* __restrict__ scratch, int nblocks, int C, float* __restrict__ dgamma,
                            float* __restrict__ dbeta, float* __restrict__ dcol) {
    pdl_trigger();   // a dependent-launch successor (the next GEMM) may set itself up while this grid drains
    __shared__ float4 sm[128][9];
    const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
    const int c = blockIdx.x * 32 + tx * 4;
    const int qn = blockIdx.y;
    float* dst = (qn == 0) ? dgamma : (qn == 1) ? dbeta : dcol;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (dst && c < C) {
#pragma unroll 8
        for (int b = ty; b < nblocks; b += 128) {
            const float4 v = *reinterpret_cast<const float4*>(scratch + ((size_t)b * 3 + qn) * C + c);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
    }
    sm[ty][tx] = s;
    __syncthreads();
    if (ty < 32) {
        float4 t = sm[ty][tx];
#pragma unroll
        for (int k = 1; k < 4; ++k) { const float4 v = sm[ty + 32 * k][tx]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
        sm[ty][tx] = t;
    }
    __syncthreads();
    if (ty == 0 && dst && c < C) {
        float4 t = sm[0][tx];
#pragma unroll
        for (int k = 1; k < 32; ++k) { const float4 v = sm[k][tx]; t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w; }
        if ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
            float4 o = *reinterpret_cast<float4*>(dst + c);
            o.x += t.x; o.y += t.y; o.z += t.z; o.w += t.w;
            *reinterpret_cast<float4*>(dst + c) = o;
        } else {
            dst[c] += t.x; dst[c + 1] += t.y; dst[c + 2] += t.z; dst[c + 3] += t.w;
        }
    }
}

// ------------------------------------------------------------------ GELU
template <typename T>
__global__ void gelu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, size_t n, int tanh_approx) {
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (; i + 3 < n; i += stride) {
        float4 v = load4(x + i);
        if (tanh_approx) v = make_float4(gelu_tanh(v.x), gelu_tanh(v.y), gelu_tanh(v.z), gelu_tanh(v.w));
        else v = make_float4(gelu_erf(v.x), gelu_erf(v.y), gelu_erf(v.z), gelu_erf(v.w));
        store4(y + i, v);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (size_t k = n & ~(size_t)3; k < n; ++k) {
            float v = to_f32(x[k]);
            y[k] = from_f32<T>(tanh_approx ? gelu_tanh(v) : gelu_erf(v));
        }
}
template <typename T>
__global__ void gelu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, size_t n,
                                int tanh_approx) {
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (; i + 3 < n; i += stride) {
        float4 v = load4(x + i), g = load4(dy + i);
        if (tanh_approx) v = make_float4(gelu_tanh_grad(v.x), gelu_tanh_grad(v.y), gelu_tanh_grad(v.z), gelu_tanh_grad(v.w));
        else v = make_float4(gelu_erf_grad(v.x), gelu_erf_grad(v.y), gelu_erf_grad(v.z), gelu_erf_grad(v.w));
        store4(dx + i, make_float4(g.x * v.x, g.y * v.y, g.z * v.z, g.w * v.w));
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (size_t k = n & ~(size_t)3; k < n; ++k) {
            float v = to_f32(x[k]);
            dx[k] = from_f32<T>(to_f32(dy[k]) * (tanh_approx ? gelu_tanh_grad(v) : gelu_erf_grad(v)));
        }
}

// ------------------------------------------------------------------ column sums (bias gradients)
// part[y][n] = sum over this block's row slice of X[m,n]; 32 x 8 threads, one 16-byte vector of columns per thread
// (4 fp32 or 8 bf16), four rows in flight per thread.
template <typename T>
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const T* __restrict__ X, long long ld, int M, int N, float* __restrict__ part) {
    constexpr int VN = Vec16<T>::N;
    __shared__ float sm[8][32][VN + 1];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = (blockIdx.x * 32 + tx) * VN;
    const int rows_per = (M + gridDim.y - 1) / gridDim.y;
    const int r0 = blockIdx.y * rows_per, r1 = min(M, r0 + rows_per);
    float s[VN];
#pragma unroll
    for (int k = 0; k < VN; ++k) s[k] = 0.f;
    const bool vec_ok = (col + VN <= N) && (ld % VN == 0) && ((reinterpret_cast<uintptr_t>(X) & 15) == 0);
    if (vec_ok) {
        int r = r0 + ty;
        for (; r + 24 < r1; r += 32) {
            Vec16<T> v0, v1, v2, v3;
            v0.load(X + (size_t)r * ld + col);
            v1.load(X + (size_t)(r + 8) * ld + col);
            v2.load(X + (size_t)(r + 16) * ld + col);
            v3.load(X + (size_t)(r + 24) * ld + col);
#pragma unroll
            for (int k = 0; k < VN; ++k) s[k] += (v0.get(k) + v1.get(k)) + (v2.get(k) + v3.get(k));
        }
        for (; r < r1; r += 8) {
            Vec16<T> v;
            v.load(X + (size_t)r * ld + col);
#pragma unroll
            for (int k = 0; k < VN; ++k) s[k] += v.get(k);
        }
    } else if (col < N) {
        for (int r = r0 + ty; r < r1; r += 8)
            for (int k = 0; k < VN && col + k < N; ++k) s[k] += to_f32(X[(size_t)r * ld + col + k]);
    }
#pragma unroll
    for (int k = 0; k < VN; ++k) sm[ty][tx][k] = s[k];
    __syncthreads();
    if (ty == 0 && col < N) {
        for (int k = 0; k < VN && col + k < N; ++k) {
            float t = 0.f;
#pragma unroll
            for (int w = 0; w < 8; ++w) t += sm[w][tx][k];
            part[(size_t)blockIdx.y * N + col + k] = t;
        }
    }
}
__global__ void colsum_reduce_kernel(const float* __restrict__ part, int nparts, int N, float* __restrict__ out,
                                     int accumulate) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= N) return;
    float s = 0.f;
    for (int p = 0; p < nparts; ++p) s += part[(size_t)p * N + c];
    out[c] = accumulate ? out[c] + s : s;
}

template <typename TI, typename TO>
__global__ void cast_kernel(const TI* __restrict__ in, TO* __restrict__ out, size_t n) {
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (; i + 3 < n; i += stride) store4(out + i, load4(in + i));
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (size_t k = n & ~(size_t)3; k < n; ++k) out[k] = from_f32<TO>(to_f32(in[k]));
}

// y += alpha * x over a flat fp32 buffer, alpha read from device memory (no host sync)
__global__ void axpy_f32_kernel(float* __restrict__ y, const float* __restrict__ x, const float* __restrict__ alpha, size_t n) {
    const float a = *alpha;
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (; i + 3 < n; i += stride) {
        float4 yv = *reinterpret_cast<const float4*>(y + i);
        const float4 xv = *reinterpret_cast<const float4*>(x + i);
        yv.x = fmaf(a, xv.x, yv.x); yv.y = fmaf(a, xv.y, yv.y); yv.z = fmaf(a, xv.z, yv.z); yv.w = fmaf(a, xv.w, yv.w);
        *reinterpret_cast<float4*>(y + i) = yv;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (size_t k = n & ~(size_t)3; k < n; ++k) y[k] = fmaf(a, x[k], y[k]);
}

template <int J, bool DCOL, typename TX, typename TW, typename TG>
static void launch_ln_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                          const void* add, void* dx, void* dx_bf16, float* scratch, int rows, int C, int grid,
                          cudaStream_t st) {
    auto k = layernorm_bwd_kernel<TX, TW, TG, J, DCOL>;
    const size_t smem = (size_t)8 * C * sizeof(float);
    if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<grid, 256, smem, st>>>((const TG*)dy, (const TX*)x, (const TW*)gamma, mean, rstd, (const TX*)add, (TX*)dx,
                              (__nv_bfloat16*)dx_bf16, scratch, rows, C);
}

template <bool DCOL, typename TX, typename TW, typename TG>
static int dispatch_ln_bwd_j(int J, const void* dy, const void* x, const void* gamma, const float* mean,
                             const float* rstd, const void* add, void* dx, void* dx_bf16, float* scratch, int rows,
                             int C, int grid, cudaStream_t st) {
#define PENROZ_LN_CASE(j) \
    if (J <= j) { launch_ln_bwd<j, DCOL, TX, TW, TG>(dy, x, gamma, mean, rstd, add, dx, dx_bf16, scratch, rows, C, grid, st); return 0; }
    PENROZ_LN_CASE(1) PENROZ_LN_CASE(2) PENROZ_LN_CASE(4) PENROZ_LN_CASE(6) PENROZ_LN_CASE(8)
    if (!DCOL) { PENROZ_LN_CASE(13) PENROZ_LN_CASE(16) }
#undef PENROZ_LN_CASE
    return -1;
}

}  // namespace penroz

using namespace penroz;

static inline int grid_for(size_t n_vec4, int threads) {
    size_t g = (n_vec4 + threads - 1) / threads;
    size_t cap = (size_t)num_sms() * 16;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

// int32 token ids (what the shard loader yields, loaders.py:62-63) -> the int64 ids the embedding and cross-entropy
// kernels index with. The reference widens on the host (torch.tensor(arr, dtype=torch.long), neural_net_model.py:630);
// here the int32 window is what crosses PCIe (half the bytes) and the GPU widens it.
namespace penroz {
__global__ void widen_tokens_kernel(const int* __restrict__ src, long long* __restrict__ dst, size_t n) {
    size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    const size_t stride = (size_t)gridDim.x * blockDim.x * 4;
    for (; i < n; i += stride) {
        if (i + 3 < n) {
            const int4 v = *reinterpret_cast<const int4*>(src + i);
            longlong2 a, b;
            a.x = v.x; a.y = v.y; b.x = v.z; b.y = v.w;
            *reinterpret_cast<longlong2*>(dst + i) = a;
            *reinterpret_cast<longlong2*>(dst + i + 2) = b;
        } else {
            for (size_t k = i; k < n; ++k) dst[k] = src[k];
        }
    }
}
}  // namespace penroz

extern "C" int penroz_widen_tokens(const int* src, long long* dst, long long n, void* stream) {
    PENROZ_REQUIRE(src && dst && n > 0, "widen_tokens: bad arguments");
    PENROZ_REQUIRE(((uintptr_t)src | (uintptr_t)dst) % 16 == 0, "widen_tokens: buffers must be 16-byte aligned");
    const int grid = grid_for((size_t)(n + 3) / 4, 256);
    penroz::widen_tokens_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(src, dst, (size_t)n);
    count_launch();
    return check_launch("widen_tokens");
}

extern "C" int penroz_embed_fwd(const long long* idx, const void* wte, const void* wpe, void* out, int B, int T,
                                int C, int V, int pos_offset, int w_dtype, int out_dtype, int* err_flag,
                                void* stream) {
    PENROZ_REQUIRE(B > 0 && T > 0 && C > 0 && V > 0, "embed_fwd: bad shape");
    PENROZ_REQUIRE(pos_offset >= 0, "embed_fwd: negative position offset");
    cudaStream_t st = (cudaStream_t)stream;
    const int rows = B * T, wpb = 8;
    const int grid = (rows + wpb - 1) / wpb;
    if (w_dtype == F32 && out_dtype == F32)
        embed_fwd_kernel<float, float><<<grid, wpb * 32, 0, st>>>(idx, (const float*)wte, (const float*)wpe, (float*)out, rows, T, C, V, pos_offset, err_flag);
    else if (w_dtype == BF16 && out_dtype == BF16)
        embed_fwd_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, wpb * 32, 0, st>>>(idx, (const __nv_bfloat16*)wte, (const __nv_bfloat16*)wpe, (__nv_bfloat16*)out, rows, T, C, V, pos_offset, err_flag);
    else { set_error("embed_fwd: unsupported dtype combination w=%d out=%d", w_dtype, out_dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("embed_fwd");
}

extern "C" int penroz_embed_bwd(const long long* idx, const float* dx, float* dwte, float* dwpe, int B, int T, int C,
                                int V, int pos_offset, void* stream) {
    PENROZ_REQUIRE(B > 0 && T > 0 && C > 0 && V > 0, "embed_bwd: bad shape");
    cudaStream_t st = (cudaStream_t)stream;
    const int rows = B * T;
    embed_bwd_wte_kernel<<<(rows + 7) / 8, 256, 0, st>>>(idx, dx, dwte, rows, C, V);
    embed_bwd_wpe_kernel<<<(T * C + 255) / 256, 256, 0, st>>>(dx, dwpe, B, T, C, pos_offset);
    count_launch(2);
    return check_launch("embed_bwd");
}

extern "C" int penroz_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* out, float* mean,
                                    float* rstd, int rows, int C, float eps, int x_dtype, int w_dtype, int out_dtype,
                                    void* stream) {
    PENROZ_REQUIRE(rows > 0 && C > 0, "layernorm_fwd: bad shape rows=%d C=%d", rows, C);
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = (rows + 7) / 8;
    const int J = (C + 127) / 128;
    const bool reg = (C % 4 == 0) && J <= 16;
#define LN_FWD_J(TX, TW, TO, j) layernorm_fwd_reg_kernel<TX, TW, TO, j><<<grid, 256, 0, st>>>((const TX*)x, (const TW*)gamma, (const TW*)beta, (TO*)out, mean, rstd, rows, C, eps)
#define LN_FWD(TX, TW, TO)                                                                                    \
    if (!reg) layernorm_fwd_kernel<TX, TW, TO><<<grid, 256, 0, st>>>((const TX*)x, (const TW*)gamma, (const TW*)beta, (TO*)out, mean, rstd, rows, C, eps); \
    else if (J <= 1) LN_FWD_J(TX, TW, TO, 1); else if (J <= 2) LN_FWD_J(TX, TW, TO, 2);                        \
    else if (J <= 4) LN_FWD_J(TX, TW, TO, 4); else if (J <= 6) LN_FWD_J(TX, TW, TO, 6);                        \
    else if (J <= 8) LN_FWD_J(TX, TW, TO, 8); else if (J <= 13) LN_FWD_J(TX, TW, TO, 13);                      \
    else LN_FWD_J(TX, TW, TO, 16)
    if (x_dtype == F32 && w_dtype == F32 && out_dtype == BF16) LN_FWD(float, float, __nv_bfloat16);
    else if (x_dtype == F32 && w_dtype == F32 && out_dtype == F32) LN_FWD(float, float, float);
    else if (x_dtype == BF16 && w_dtype == BF16 && out_dtype == BF16) LN_FWD(__nv_bfloat16, __nv_bfloat16, __nv_bfloat16);
    else { set_error("layernorm_fwd: unsupported dtypes x=%d w=%d out=%d", x_dtype, w_dtype, out_dtype); return PENROZ_ERR_ARG; }
#undef LN_FWD
#undef LN_FWD_J
    count_launch();
    return check_launch("layernorm_fwd");
}

// scratch must hold penroz_layernorm_bwd_scratch_bytes(rows, C) bytes
static int ln_bwd_grid_old(int rows) {
    int g = (rows + 7) / 8;
    int cap = 2 * num_sms();
    return g < cap ? g : cap;
}
// upper bound of the one-pass kernel's CTAs = per-block partial-sum slots in scratch (at most 4 resident per SM)
static int ln_bwd_grid(int rows) {
    int g = (rows + 1) / 2;
    int cap = 4 * num_sms();
    return g < 1 ? 1 : (g < cap ? g : cap);
}
static bool ln_bwd_two_kernel() {
    static const bool on = getenv("PENROZ_LN_BWD") && !strcmp(getenv("PENROZ_LN_BWD"), "two-kernel");
    return on;
}
extern "C" size_t penroz_layernorm_bwd_scratch_bytes(int rows, int C) {
    return (size_t)ln_bwd_grid(rows) * 3 * (size_t)C * sizeof(float);
}

namespace penroz {
int layernorm_bwd_impl(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                       const void* add, void* dx, void* dx_bf16, float* dgamma, float* dbeta, float* dcol,
                       int rows, int C, int x_dtype, int w_dtype, int dy_dtype, float* scratch, cudaStream_t st) {
    PENROZ_REQUIRE(rows > 0 && C > 0, "layernorm_bwd: bad shape rows=%d C=%d", rows, C);
    PENROZ_REQUIRE((C & 3) == 0, "layernorm_bwd: C=%d must be a multiple of 4", C);
    PENROZ_REQUIRE(C <= 2048, "layernorm_bwd: C=%d exceeds the supported 2048", C);
    PENROZ_REQUIRE(scratch != nullptr, "layernorm_bwd: scratch is required");
    const int J = (C + 127) / 128;
    int grid = ln_bwd_grid(rows);
    // one-pass staged kernel: fp32 x / gamma, rows of 16-byte multiples in every operand, at least two stages
    {
        const int nv = C / 4, vpt = nv <= 256 ? 1 : 2;
        const int threads = (((nv + vpt - 1) / vpt) + 31) / 32 * 32;
        const size_t esz = dy_dtype == BF16 ? 2 : 4;
        auto stage_bytes = [&](int R) {
            const size_t a = ((size_t)R * C * esz + 127) & ~(size_t)127, b = ((size_t)R * C * 4 + 127) & ~(size_t)127;
            return a + 2 * b;
        };
        // measured (32768 x 768 / x 1024, 16384 x 1600): resident warps matter more than ring depth -- 2 rows per
        // stage, 3 stages, as many CTAs per SM as 200 KB of shared memory hold (4 / 3 / 2): 74 / 91 / 76 us, 5.4-5.9
        // TB/s of the op's bytes; 4-row stages or deeper rings with fewer CTAs are 10-40 % slower
        const int R = 2;
        int S = 3;
        if ((size_t)S * stage_bytes(R) > 200 * 1024) S = 2;
        int per_sm = (int)((size_t)200 * 1024 / ((size_t)S * stage_bytes(R) + 1024));
        per_sm = per_sm > 4 ? 4 : per_sm;
        const bool ok = !ln_bwd_two_kernel() && x_dtype == F32 && w_dtype == F32 && (dy_dtype == BF16 || dy_dtype == F32) &&
                        C % 8 == 0 && per_sm >= 1 &&
                        (((uintptr_t)dy | (uintptr_t)x | (uintptr_t)add) & 15) == 0;
        if (ok) {
            const int nchunks = (rows + R - 1) / R;
            grid = per_sm * num_sms();            // <= ln_bwd_grid's cap, which sizes scratch
            if (grid > nchunks) grid = nchunks;
            const size_t smem = (size_t)S * stage_bytes(R);
            cudaError_t e = cudaSuccess;
            int dev = 0;
            cudaGetDevice(&dev);
#define LN_STAGED(TG, V, RR)                                                                                            \
            {                                                                                                           \
                auto k = layernorm_bwd_staged_kernel<TG, V, RR>;                                                        \
                static std::atomic<int> granted[64];                                                                    \
                if (!granted[dev & 63].load(std::memory_order_relaxed)) {                                               \
                    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);               \
                    if (e == cudaSuccess) granted[dev & 63].store(1, std::memory_order_relaxed);                        \
                }                                                                                                       \
                if (e == cudaSuccess)                                                                                   \
                    k<<<grid, threads, smem, st>>>((const TG*)dy, (const float*)x, (const float*)gamma, mean, rstd,      \
                                                   (const float*)add, (float*)dx, (__nv_bfloat16*)dx_bf16, scratch, rows, \
                                                   C, S, dcol != nullptr);                                              \
            }
            if (dy_dtype == BF16) { if (vpt == 1) LN_STAGED(__nv_bfloat16, 1, 2) else LN_STAGED(__nv_bfloat16, 2, 2) }
            else { if (vpt == 1) LN_STAGED(float, 1, 2) else LN_STAGED(float, 2, 2) }
#undef LN_STAGED
            if (e != cudaSuccess) { set_error("layernorm_bwd: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
            dim3 rg((C + 31) / 32, dcol ? 3 : 2);
            layernorm_bwd_reduce_kernel<<<rg, 1024, 0, st>>>(scratch, grid, C, dgamma, dbeta, dcol);
            count_launch(2);
            return check_launch("layernorm_bwd");
        }
    }
    // ---- two-kernel form of round 1: other operand types and alignments, and PENROZ_LN_BWD=two-kernel for A/B
    grid = ln_bwd_grid_old(rows);
    bool fuse_dcol = dcol != nullptr && J <= 8;
    int rc = -1;
    const bool wide = rows >= 4096;   // enough rows to fill the GPU with the one-row-per-warp kernel
    if (wide) {
        const int rgrid = (rows + 7) / 8;
#define LN_ROWS_J(TX, TW, TG, j) layernorm_bwd_rows_kernel<TX, TW, TG, j><<<rgrid, 256, 0, st>>>((const TG*)dy, (const TX*)x, (const TW*)gamma, mean, rstd, (const TX*)add, (TX*)dx, (__nv_bfloat16*)dx_bf16, rows, C)
#define LN_ROWS(TX, TW, TG)                                                      \
        if (J <= 1) LN_ROWS_J(TX, TW, TG, 1); else if (J <= 2) LN_ROWS_J(TX, TW, TG, 2); \
        else if (J <= 4) LN_ROWS_J(TX, TW, TG, 4); else if (J <= 6) LN_ROWS_J(TX, TW, TG, 6); \
        else if (J <= 8) LN_ROWS_J(TX, TW, TG, 8); else if (J <= 13) LN_ROWS_J(TX, TW, TG, 13); \
        else LN_ROWS_J(TX, TW, TG, 16)
        grid = ln_bwd_grid_old(rows);    // number of row slices == partial-sum slots in scratch
        dim3 cg((C + 127) / 128, grid);
        if (x_dtype == F32 && w_dtype == F32 && dy_dtype == BF16) {
            LN_ROWS(float, float, __nv_bfloat16);
            layernorm_bwd_cols_kernel<float, __nv_bfloat16><<<cg, 256, 0, st>>>((const __nv_bfloat16*)dy, (const float*)x, mean, rstd, (const float*)dx, (const __nv_bfloat16*)dx_bf16, scratch, rows, C, dcol != nullptr);
        } else if (x_dtype == F32 && w_dtype == F32 && dy_dtype == F32) {
            LN_ROWS(float, float, float);
            layernorm_bwd_cols_kernel<float, float><<<cg, 256, 0, st>>>((const float*)dy, (const float*)x, mean, rstd, (const float*)dx, (const __nv_bfloat16*)dx_bf16, scratch, rows, C, dcol != nullptr);
        } else { set_error("layernorm_bwd: unsupported dtypes x=%d w=%d dy=%d", x_dtype, w_dtype, dy_dtype); return PENROZ_ERR_ARG; }
#undef LN_ROWS
#undef LN_ROWS_J
        fuse_dcol = dcol != nullptr;
        rc = 0;
        count_launch(1);
    } else {
#define LN_BWD(TX, TW, TG)                                                                                              \
    rc = fuse_dcol ? dispatch_ln_bwd_j<true, TX, TW, TG>(J, dy, x, gamma, mean, rstd, add, dx, dx_bf16, scratch, rows, C, grid, st) \
                   : dispatch_ln_bwd_j<false, TX, TW, TG>(J, dy, x, gamma, mean, rstd, add, dx, dx_bf16, scratch, rows, C, grid, st)
    if (x_dtype == F32 && w_dtype == F32 && dy_dtype == BF16) { LN_BWD(float, float, __nv_bfloat16); }
    else if (x_dtype == F32 && w_dtype == F32 && dy_dtype == F32) { LN_BWD(float, float, float); }
    else { set_error("layernorm_bwd: unsupported dtypes x=%d w=%d dy=%d", x_dtype, w_dtype, dy_dtype); return PENROZ_ERR_ARG; }
#undef LN_BWD
    }
    if (rc != 0) { set_error("layernorm_bwd: no kernel for C=%d", C); return PENROZ_ERR_ARG; }
    dim3 rg((C + 31) / 32, fuse_dcol ? 3 : 2);
    layernorm_bwd_reduce_kernel<<<rg, 1024, 0, st>>>(scratch, grid, C, dgamma, dbeta, fuse_dcol ? dcol : nullptr);
    count_launch(2);
    int e = check_launch("layernorm_bwd");
    if (e) return e;
    if (dcol != nullptr && !fuse_dcol) {
        // wide rows: column sums of the written gradient as a separate pass (re-uses the same scratch)
        const void* src = dx_bf16 ? dx_bf16 : dx;
        const int parts = 32;
        dim3 g((C + 255) / 256, parts), g4((C + 127) / 128, parts);
        if (dx_bf16) colsum_partial_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)src, C, rows, C, scratch);
        else colsum_partial_kernel<float><<<g4, 256, 0, st>>>((const float*)src, C, rows, C, scratch);
        colsum_reduce_kernel<<<(C + 255) / 256, 256, 0, st>>>(scratch, parts, C, dcol, 1);
        count_launch(2);
        return check_launch("layernorm_bwd/colsum");
    }
    return PENROZ_OK;
}
}  // namespace penroz

extern "C" int penroz_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean,
                                    const float* rstd, const void* add, void* dx, void* dx_bf16, float* dgamma,
                                    float* dbeta, float* dcol, int rows, int C, int x_dtype, int w_dtype,
                                    int dy_dtype, void* scratch, size_t scratch_bytes, void* stream) {
    PENROZ_REQUIRE(scratch_bytes >= penroz_layernorm_bwd_scratch_bytes(rows, C),
                   "layernorm_bwd: scratch too small (%zu < %zu)", scratch_bytes, penroz_layernorm_bwd_scratch_bytes(rows, C));
    return layernorm_bwd_impl(dy, x, gamma, mean, rstd, add, dx, dx_bf16, dgamma, dbeta, dcol, rows, C, x_dtype,
                              w_dtype, dy_dtype, (float*)scratch, (cudaStream_t)stream);
}

extern "C" int penroz_gelu_fwd(const void* x, void* y, long long n, int tanh_approx, int dtype, void* stream) {
    PENROZ_REQUIRE(n > 0, "gelu_fwd: empty");
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = grid_for((size_t)(n + 3) / 4, 256);
    if (dtype == F32) gelu_fwd_kernel<float><<<grid, 256, 0, st>>>((const float*)x, (float*)y, (size_t)n, tanh_approx);
    else if (dtype == BF16) gelu_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, (size_t)n, tanh_approx);
    else { set_error("gelu_fwd: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("gelu_fwd");
}

extern "C" int penroz_gelu_bwd(const void* dy, const void* x, void* dx, long long n, int tanh_approx, int dtype,
                               void* stream) {
    PENROZ_REQUIRE(n > 0, "gelu_bwd: empty");
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = grid_for((size_t)(n + 3) / 4, 256);
    if (dtype == F32) gelu_bwd_kernel<float><<<grid, 256, 0, st>>>((const float*)dy, (const float*)x, (float*)dx, (size_t)n, tanh_approx);
    else if (dtype == BF16) gelu_bwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)dy, (const __nv_bfloat16*)x, (__nv_bfloat16*)dx, (size_t)n, tanh_approx);
    else { set_error("gelu_bwd: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("gelu_bwd");
}

extern "C" size_t penroz_colsum_scratch_bytes(int M, int N) { (void)M; return (size_t)32 * (size_t)N * sizeof(float); }

// out[n] (+)= sum_m X[m,n]   (bias gradient of a Linear whose output gradient is X)
extern "C" int penroz_colsum(const void* X, long long ld, int M, int N, float* out, int accumulate, int dtype,
                             void* scratch, size_t scratch_bytes, void* stream) {
    PENROZ_REQUIRE(M > 0 && N > 0, "colsum: bad shape");
    PENROZ_REQUIRE(ld >= N, "colsum: leading dimension %lld is smaller than the row length %d (rows would alias)", ld, N);
    PENROZ_REQUIRE(scratch && scratch_bytes >= penroz_colsum_scratch_bytes(M, N), "colsum: scratch too small");
    cudaStream_t st = (cudaStream_t)stream;
    const int parts = 32;
    dim3 g4((N + 127) / 128, parts), g8((N + 255) / 256, parts);
    if (dtype == F32) colsum_partial_kernel<float><<<g4, 256, 0, st>>>((const float*)X, ld, M, N, (float*)scratch);
    else if (dtype == BF16) colsum_partial_kernel<__nv_bfloat16><<<g8, 256, 0, st>>>((const __nv_bfloat16*)X, ld, M, N, (float*)scratch);
    else { set_error("colsum: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    colsum_reduce_kernel<<<(N + 255) / 256, 256, 0, st>>>((const float*)scratch, parts, N, out, accumulate);
    count_launch(2);
    return check_launch("colsum");
}

extern "C" int penroz_cast(const void* in, void* out, long long n, int in_dtype, int out_dtype, void* stream) {
    PENROZ_REQUIRE(n > 0, "cast: empty");
    cudaStream_t st = (cudaStream_t)stream;
    const int grid = grid_for((size_t)(n + 3) / 4, 256);
    if (in_dtype == F32 && out_dtype == BF16) cast_kernel<float, __nv_bfloat16><<<grid, 256, 0, st>>>((const float*)in, (__nv_bfloat16*)out, (size_t)n);
    else if (in_dtype == BF16 && out_dtype == F32) cast_kernel<__nv_bfloat16, float><<<grid, 256, 0, st>>>((const __nv_bfloat16*)in, (float*)out, (size_t)n);
    else { set_error("cast: unsupported %d -> %d", in_dtype, out_dtype); return PENROZ_ERR_ARG; }
    count_launch();
    return check_launch("cast");
}

extern "C" int penroz_axpy_f32(float* y, const float* x, const float* alpha, long long n, void* stream) {
    PENROZ_REQUIRE(y && x && alpha && n > 0, "axpy_f32: bad arguments");
    PENROZ_REQUIRE((((uintptr_t)y | (uintptr_t)x) & 15) == 0, "axpy_f32: buffers must be 16-byte aligned");
    const int grid = grid_for((size_t)(n + 3) / 4, 256);
    axpy_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(y, x, alpha, (size_t)n);
    count_launch();
    return check_launch("axpy_f32");
}

// ------------------------------------------------------------------ column sums without caller scratch
// Same two kernels as penroz_colsum, but the [32][N] partial-sum buffer lives in a fixed array inside the library
// image instead of being passed in. It exists for hosts that hand this library nothing but input and output
// pointers and cannot obtain device memory themselves (see the header comment of penroz_colsum_static).
namespace penroz {
constexpr int kStaticColsumMaxCols = 8192;   // == PENROZ_COLSUM_STATIC_MAX_COLS in include/penroz_b200.h
constexpr int kStaticColsumParts = 32;
// 32 x 8192 floats = 1 MiB of global memory on every GPU that loads the library, held until the process exits.
static __device__ __align__(256) float g_colsum_static_scratch[kStaticColsumParts * kStaticColsumMaxCols];

// Device address of the array above for the CURRENT device, resolved once per device. In host code the symbol
// name is only a handle (its host-side value is not a device address), so kernels are always given the pointer
// that cudaGetSymbolAddress returns, never the name.
static int static_colsum_scratch(float** out) {
    static std::atomic<float*> resolved[64];
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_error("colsum_static: cudaGetDevice: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
    if (dev < 0 || dev >= 64) { set_error("colsum_static: device ordinal %d not supported", dev); return PENROZ_ERR_LIMIT; }
    float* p = resolved[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        void* addr = nullptr;
        e = cudaGetSymbolAddress(&addr, g_colsum_static_scratch);
        if (e != cudaSuccess || addr == nullptr) {
            set_error("colsum_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return PENROZ_ERR_CUDA;
        }
        p = static_cast<float*>(addr);
        resolved[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return PENROZ_OK;
}
}  // namespace penroz

extern "C" int penroz_colsum_static(const void* X, long long ld, int M, int N, float* out, int accumulate, int dtype,
                                    void* stream) {
    PENROZ_REQUIRE(M > 0 && N > 0, "colsum_static: bad shape M=%d N=%d", M, N);
    PENROZ_REQUIRE(X != nullptr && out != nullptr, "colsum_static: null argument");
    PENROZ_REQUIRE(ld >= N, "colsum_static: leading dimension %lld is smaller than the row length %d (rows would alias)",
                   ld, N);
    if (N > kStaticColsumMaxCols) {
        set_error("colsum_static: N=%d exceeds the built-in scratch (at most %d columns); use penroz_colsum with "
                  "caller-provided scratch for wider matrices", N, kStaticColsumMaxCols);
        return PENROZ_ERR_LIMIT;
    }
    float* scratch = nullptr;
    int rc = static_colsum_scratch(&scratch);
    if (rc != PENROZ_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 g4((N + 127) / 128, kStaticColsumParts), g8((N + 255) / 256, kStaticColsumParts);
    // every [part][column < N] slot is written by the first kernel before the second reads it, so nothing left over
    // from an earlier call is ever consumed and no clearing is needed
    if (dtype == F32) colsum_partial_kernel<float><<<g4, 256, 0, st>>>((const float*)X, ld, M, N, scratch);
    else if (dtype == BF16) colsum_partial_kernel<__nv_bfloat16><<<g8, 256, 0, st>>>((const __nv_bfloat16*)X, ld, M, N, scratch);
    else { set_error("colsum_static: bad dtype %d", dtype); return PENROZ_ERR_ARG; }
    colsum_reduce_kernel<<<(N + 255) / 256, 256, 0, st>>>(scratch, kStaticColsumParts, N, out, accumulate);
    count_launch(2);
    return check_launch("colsum_static");
}
