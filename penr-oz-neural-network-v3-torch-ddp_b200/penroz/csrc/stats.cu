// Rank-0 training diagnostics on the GPU: mean / std / min / max / saturated fraction and the 100-bin density
// histogram of an activation, gradient or weight tensor, with ONE small device->host read per tensor.
//
// Replaces, for CUDA tensors, what _record_training_overall_progress does per layer and per weight
// (neural_net_model.py:743-777): t.mean().item(), t.std().item(), a saturation test and
// torch.histogram(t.cpu(), density=True) -- i.e. a full device->host copy of every activation (6.6 GB for the 124M
// logits alone) and a single-threaded CPU histogram. Here each tensor is read twice from HBM (statistics, then
// centred sum of squares + histogram) and 105 doubles come back.
//
// Bin assignment follows ATen's CPU histogram for uniform bins step by step, in the tensor's float arithmetic
// (aten/src/ATen/native/cpu/HistogramKernel.cpp, LINEAR_INTERPOLATION_WITH_LOCAL_SEARCH): position estimate
// (x - lo) * bins / (hi - lo), a local search against the torch.linspace bin edges, the right edge closed; a
// constant tensor widens its range by 0.5 each side. Counts are therefore the ones torch.histogram returns.
#include "common.cuh"

namespace penroz {

constexpr int kStatBlocks = 592;      // 4 CTAs per SM on a 148-SM part; partials are folded in fixed order
constexpr int kMaxBins = 1024;

struct StatPartial { double sum; float mn, mx; unsigned long long hit; };

template <typename T>
__global__ void __launch_bounds__(256)
stats_pass1_kernel(const T* __restrict__ x, long long n, float thr, StatPartial* __restrict__ part) {
    __shared__ double s_sum[8];
    __shared__ float s_mn[8], s_mx[8];
    __shared__ unsigned long long s_hit[8];
    double sum = 0.0;
    float mn = INFINITY, mx = -INFINITY;
    unsigned long long hit = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = to_f32(x[i]);
        sum += (double)v;
        mn = fminf(mn, v);
        mx = fmaxf(mx, v);
        hit += (fabsf(v) > thr) ? 1ull : 0ull;
    }
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        hit += __shfl_xor_sync(0xffffffffu, hit, o);
    }
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { s_sum[w] = sum; s_mn[w] = mn; s_mx[w] = mx; s_hit[w] = hit; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) { sum += s_sum[k]; mn = fminf(mn, s_mn[k]); mx = fmaxf(mx, s_mx[k]); hit += s_hit[k]; }
        part[blockIdx.x] = StatPartial{sum, mn, mx, hit};
    }
}

// head[0..4] = mean, (std later), min, max, saturated fraction; range[0..1] = histogram range lo, hi (float)
__global__ void stats_fold1_kernel(const StatPartial* __restrict__ part, int nparts, long long n, double* __restrict__ head,
                                   float* __restrict__ range, unsigned long long* __restrict__ counts, int bins,
                                   double* __restrict__ sq_part) {
    if (threadIdx.x == 0) {
        double sum = 0.0;
        float mn = INFINITY, mx = -INFINITY;
        unsigned long long hit = 0;
        for (int k = 0; k < nparts; ++k) { sum += part[k].sum; mn = fminf(mn, part[k].mn); mx = fmaxf(mx, part[k].mx); hit += part[k].hit; }
        head[0] = sum / (double)n;
        head[2] = (double)mn;
        head[3] = (double)mx;
        head[4] = (double)hit / (double)n;
        float lo = mn, hi = mx;
        if (lo == hi) { lo -= 0.5f; hi += 0.5f; }    // ATen select_outer_bin_edges
        range[0] = lo;
        range[1] = hi;
    }
    for (int b = threadIdx.x; b < bins; b += blockDim.x) counts[b] = 0ull;
    for (int b = threadIdx.x; b < kStatBlocks; b += blockDim.x) sq_part[b] = 0.0;
}

template <typename T>
__global__ void __launch_bounds__(256)
stats_pass2_kernel(const T* __restrict__ x, long long n, const double* __restrict__ head, const float* __restrict__ range,
                   int bins, double* __restrict__ sq_part, unsigned long long* __restrict__ counts) {
    __shared__ float edges[kMaxBins + 1];
    __shared__ unsigned int hist[kMaxBins];
    __shared__ double s_sq[8];
    const float lo = range[0], hi = range[1];
    const int n_edges = bins + 1;
    {   // torch.linspace(lo, hi, bins + 1) in float: filled from both ends (aten RangeFactories linspace kernel)
        const float step = (hi - lo) / (float)(n_edges - 1);
        const int halfway = n_edges / 2;
        for (int i = threadIdx.x; i < n_edges; i += blockDim.x)
            edges[i] = (i < halfway) ? (lo + step * (float)i) : (hi - step * (float)(n_edges - i - 1));
        for (int i = threadIdx.x; i < bins; i += blockDim.x) hist[i] = 0u;
    }
    __syncthreads();
    const double mean = head[0];
    const float width = hi - lo;
    double sq = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = to_f32(x[i]);
        const double d = (double)v - mean;
        sq += d * d;
        if (v >= lo && v <= hi) {
            long long pos = (long long)((v - lo) * (float)(n_edges - 1) / width);
            long long p0 = pos - 1 < 0 ? 0 : pos - 1;
            long long p1 = pos + 2 < n_edges ? pos + 2 : n_edges;
            // upper_bound(edges + p0, edges + p1, v) - edges - 1
            long long ub = p0;
            while (ub < p1 && !(v < edges[ub])) ++ub;
            pos = ub - 1;
            if (pos == n_edges - 1) pos -= 1;
            if (pos >= 0 && pos < bins) atomicAdd(&hist[pos], 1u);
        }
    }
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) s_sq[w] = sq;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 1; k < 8; ++k) sq += s_sq[k];
        sq_part[blockIdx.x] = sq;
    }
    for (int b = threadIdx.x; b < bins; b += blockDim.x)
        if (hist[b]) atomicAdd(&counts[b], (unsigned long long)hist[b]);
}

// out[0..4] = mean, std (unbiased), min, max, saturated fraction; out[5] = lo, out[6] = hi; out[7 .. 7+bins) = counts
__global__ void stats_fold2_kernel(const double* __restrict__ head, const float* __restrict__ range,
                                   const double* __restrict__ sq_part, int nparts, long long n,
                                   const unsigned long long* __restrict__ counts, int bins, double* __restrict__ out) {
    if (threadIdx.x == 0) {
        double sq = 0.0;
        for (int k = 0; k < nparts; ++k) sq += sq_part[k];
        out[0] = head[0];
        out[1] = n > 1 ? sqrt(sq / (double)(n - 1)) : NAN;
        out[2] = head[2];
        out[3] = head[3];
        out[4] = head[4];
        out[5] = (double)range[0];
        out[6] = (double)range[1];
    }
    for (int b = threadIdx.x; b < bins; b += blockDim.x) out[7 + b] = (double)counts[b];
}

}  // namespace penroz

using namespace penroz;

extern "C" size_t penroz_tensor_stats_scratch_bytes(int bins) {
    return sizeof(StatPartial) * kStatBlocks + sizeof(double) * (8 + kStatBlocks) + sizeof(float) * 2 + 8 +
           sizeof(unsigned long long) * (size_t)(bins > 0 ? bins : 1) + 256;
}

extern "C" int penroz_tensor_stats(const void* x, long long n, int dtype, int bins, float abs_threshold, double* out,
                                   void* scratch, size_t scratch_bytes, void* stream) {
    PENROZ_REQUIRE(x && out && n > 0, "tensor_stats: bad arguments");
    PENROZ_REQUIRE(bins >= 1 && bins <= kMaxBins, "tensor_stats: bins=%d outside [1, %d]", bins, kMaxBins);
    PENROZ_REQUIRE(dtype == F32 || dtype == BF16, "tensor_stats: bad dtype %d", dtype);
    PENROZ_REQUIRE(scratch && scratch_bytes >= penroz_tensor_stats_scratch_bytes(bins), "tensor_stats: scratch too small");
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t* w = (uint8_t*)scratch;
    StatPartial* part = (StatPartial*)w;            w += sizeof(StatPartial) * kStatBlocks;
    double* head = (double*)w;                      w += sizeof(double) * 8;
    double* sq_part = (double*)w;                   w += sizeof(double) * kStatBlocks;
    unsigned long long* counts = (unsigned long long*)w;  w += sizeof(unsigned long long) * bins;
    float* range = (float*)w;
    long long want = (n + 255) / 256;
    const int grid = (int)(want < kStatBlocks ? want : kStatBlocks);
    if (dtype == F32) stats_pass1_kernel<float><<<grid, 256, 0, st>>>((const float*)x, n, abs_threshold, part);
    else stats_pass1_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, n, abs_threshold, part);
    stats_fold1_kernel<<<1, 256, 0, st>>>(part, grid, n, head, range, counts, bins, sq_part);
    if (dtype == F32) stats_pass2_kernel<float><<<grid, 256, 0, st>>>((const float*)x, n, head, range, bins, sq_part, counts);
    else stats_pass2_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, n, head, range, bins, sq_part, counts);
    stats_fold2_kernel<<<1, 256, 0, st>>>(head, range, sq_part, grid, n, counts, bins, out);
    count_launch(4);
    return check_launch("tensor_stats");
}
