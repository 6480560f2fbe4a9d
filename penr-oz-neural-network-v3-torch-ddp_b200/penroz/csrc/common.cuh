// Shared device/host helpers for the penroz B200 (sm_100a) kernels.
// Everything here is private to the library; the public surface is include/penroz_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>

#define PENROZ_OK 0
#define PENROZ_ERR_ARG (-1)
#define PENROZ_ERR_CUDA (-2)
#define PENROZ_ERR_ARCH (-3)
#define PENROZ_ERR_LIMIT (-4)

namespace penroz {

// thread-local last error text (re-entrant: /generate/ runs on uvicorn worker threads)
void set_error(const char* fmt, ...);
extern std::atomic<unsigned long long> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
int check_launch(const char* what);  // cudaGetLastError -> PENROZ_ERR_CUDA with text
int num_sms();                        // cached per device

enum DType { F32 = 0, BF16 = 1 };

#define PENROZ_REQUIRE(cond, ...)                 \
    do {                                          \
        if (!(cond)) {                            \
            penroz::set_error(__VA_ARGS__);       \
            return PENROZ_ERR_ARG;                \
        }                                         \
    } while (0)

// ---------------------------------------------------------------- scalar conversions
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t v) {
    __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&v);
    return __bfloat1622float2(h);
}

// 16-byte vector of T (4 floats or 8 bf16)
template <typename T> struct Vec16;
template <> struct Vec16<float> {
    static constexpr int N = 4;
    float4 raw;
    __device__ __forceinline__ void load(const float* p) { raw = *reinterpret_cast<const float4*>(p); }
    __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = raw; }
    __device__ __forceinline__ float get(int i) const { return (&raw.x)[i]; }
    __device__ __forceinline__ void set(int i, float v) { (&raw.x)[i] = v; }
};
template <> struct Vec16<__nv_bfloat16> {
    static constexpr int N = 8;
    uint4 raw;
    __device__ __forceinline__ void load(const __nv_bfloat16* p) { raw = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
    __device__ __forceinline__ float get(int i) const {
        uint32_t w = (&raw.x)[i >> 1];
        return __uint_as_float((i & 1) ? (w & 0xffff0000u) : (w << 16));
    }
    __device__ __forceinline__ void set(int i, float v) {
        uint32_t b = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(v));
        uint32_t& w = (&raw.x)[i >> 1];
        w = (i & 1) ? ((w & 0x0000ffffu) | (b << 16)) : ((w & 0xffff0000u) | b);
    }
};

// ---------------------------------------------------------------- warp / block reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// block-wide sum; `red` is a __shared__ float[32]; every thread gets the result
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = (lane < nw) ? red[lane] : 0.f;
    return warp_sum(t);
}
__device__ __forceinline__ float block_max(float v, float* red) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[w] = v;
    __syncthreads();
    float t = (lane < nw) ? red[lane] : -INFINITY;
    return warp_max(t);
}

// exact (erf) and tanh GELU with their derivatives; fp32 math like ATen's GeluCUDAKernelImpl
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.f + erff(x * 0.70710678118654752440f)); }
__device__ __forceinline__ float gelu_erf_grad(float x) {
    const float cdf = 0.5f * (1.f + erff(x * 0.70710678118654752440f));
    const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
    return cdf + x * pdf;
}
__device__ __forceinline__ float gelu_tanh(float x) {
    const float k = 0.79788456080286535588f, c = 0.044715f;
    return 0.5f * x * (1.f + tanhf(k * (x + c * x * x * x)));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
    const float k = 0.79788456080286535588f, c = 0.044715f;
    const float x2 = x * x;
    const float t = tanhf(k * (x + c * x * x2));
    const float dt = (1.f - t * t) * k * (1.f + 3.f * c * x2);
    return 0.5f * (1.f + t) + 0.5f * x * dt;
}

// Branch-free variants for the tensor-core epilogues, whose outputs are rounded to bf16 (2^-9 relative): erf by
// Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7), exp/reciprocal/tanh on the SFU. No branches, so the 32 independent
// elements a thread holds pipeline instead of serialising on erff's control flow.
__device__ __forceinline__ float rcp_fast(float x) {   // MUFU.RCP, no slow-path branch (x is in [1, inf) here)
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float ex2_fast(float x) {   // MUFU.EX2; -inf -> 0, large negative -> 0 (ftz)
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// hc = 0.5 erfc(|x| / sqrt 2) = Phi(-|x|) and g = exp(-x^2 / 2); the A&S coefficients carry the factor 0.5
__device__ __forceinline__ float half_erfc_fast(float x, float& g) {
    const float t = rcp_fast(fmaf(0.3275911f * 0.70710678118654752440f, fabsf(x), 1.f));
    g = ex2_fast((x * x) * (-0.5f * 1.44269504088896340736f));
    float poly = fmaf(0.5f * 1.061405429f, t, 0.5f * -1.453152027f);
    poly = fmaf(poly, t, 0.5f * 1.421413741f);
    poly = fmaf(poly, t, 0.5f * -0.284496736f);
    poly = fmaf(poly, t, 0.5f * 0.254829592f);
    return (poly * t) * g;
}
// x Phi(x) = relu(x) - |x| Phi(-|x|): no cancellation in the negative tail
__device__ __forceinline__ float gelu_erf_fast(float x) {
    float g;
    const float hc = half_erfc_fast(x, g);
    return fmaf(-fabsf(x), hc, fmaxf(x, 0.f));
}
// Phi(x) + x phi(x), Phi(x) = 0.5 + copysign(0.5 - hc, x)
__device__ __forceinline__ float gelu_erf_grad_fast(float x) {
    float g;
    const float hc = half_erfc_fast(x, g);
    const float cdf = 0.5f + copysignf(0.5f - hc, x);
    return fmaf(x * 0.39894228040143267794f, g, cdf);
}
// ---- the same two functions on pairs, with the packed fp32 instructions of sm_100 (FFMA2 / FMUL2 / FADD2: two
// IEEE fp32 results per issue slot). The GEMM epilogues that apply them are bound by issue slots, not by any one pipe
// (ncu: no pipe above 30 %, 45 % of issue slots used by 2 warps per scheduler), so halving the instruction count of the
// polynomial is what shortens them.
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 f2_pack(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void f2_unpack(f32x2 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ f32x2 f2_fma(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 f2_mul(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 f2_add(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 f2_all(float c) { return f2_pack(c, c); }
// q = 0.5 - Phi(-|x|) (so Phi(x) = 0.5 + copysign(q, x)), g = exp(-x^2 / 2), for the pair x
__device__ __forceinline__ f32x2 half_erf2_fast(f32x2 x, f32x2 ax, f32x2& g) {
    float a0, a1, d0, d1;
    f2_unpack(f2_mul(f2_mul(x, x), f2_all(-0.5f * 1.44269504088896340736f)), a0, a1);
    g = f2_pack(ex2_fast(a0), ex2_fast(a1));
    f2_unpack(f2_fma(ax, f2_all(0.3275911f * 0.70710678118654752440f), f2_all(1.f)), d0, d1);
    const f32x2 t = f2_pack(rcp_fast(d0), rcp_fast(d1));
    f32x2 poly = f2_fma(f2_all(-0.5f * 1.061405429f), t, f2_all(0.5f * 1.453152027f));   // minus the A&S polynomial / 2
    poly = f2_fma(poly, t, f2_all(-0.5f * 1.421413741f));
    poly = f2_fma(poly, t, f2_all(0.5f * 0.284496736f));
    poly = f2_fma(poly, t, f2_all(-0.5f * 0.254829592f));
    return f2_fma(f2_mul(poly, t), g, f2_all(0.5f));
}
// gelu(x) = x / 2 + |x| q
__device__ __forceinline__ void gelu_erf_fast2(float x0, float x1, float& y0, float& y1) {
    const f32x2 x = f2_pack(x0, x1), ax = f2_pack(fabsf(x0), fabsf(x1));
    f32x2 g;
    const f32x2 q = half_erf2_fast(x, ax, g);
    f2_unpack(f2_fma(ax, q, f2_mul(x, f2_all(0.5f))), y0, y1);
}
// r gelu'(x), gelu'(x) = 0.5 + copysign(q, x) + x g / sqrt(2 pi)
__device__ __forceinline__ void gelu_erf_grad_times_fast2(float x0, float x1, float r0, float r1, float& y0, float& y1) {
    const f32x2 x = f2_pack(x0, x1), ax = f2_pack(fabsf(x0), fabsf(x1));
    f32x2 g;
    const f32x2 q = half_erf2_fast(x, ax, g);
    float q0, q1;
    f2_unpack(q, q0, q1);
    const f32x2 cdf = f2_add(f2_pack(copysignf(q0, x0), copysignf(q1, x1)), f2_all(0.5f));
    const f32x2 grad = f2_fma(f2_mul(x, g), f2_all(0.39894228040143267794f), cdf);
    f2_unpack(f2_mul(f2_pack(r0, r1), grad), y0, y1);
}
__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float gelu_tanh_fast(float x) {
    const float k = 0.79788456080286535588f, c = 0.044715f;
    return 0.5f * x * (1.f + tanh_fast(k * (x + c * x * x * x)));
}
__device__ __forceinline__ float gelu_tanh_grad_fast(float x) {
    const float k = 0.79788456080286535588f, c = 0.044715f;
    const float x2 = x * x;
    const float t = tanh_fast(k * (x + c * x * x2));
    const float dt = (1.f - t * t) * k * (1.f + 3.f * c * x2);
    return 0.5f * (1.f + t) + 0.5f * x * dt;
}

// ---------------------------------------------------------------- sm_100a PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug traps (context error) instead of hanging the GPU box.
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    const unsigned long long t0 = global_ns();
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if ((++spins & 255u) == 0 && global_ns() - t0 > 4000000000ull) {  // 4 s: far beyond any legitimate wait
            printf("penroz: mbarrier wait timeout (block %d thread %d bar %u parity %u)\n", (int)blockIdx.x,
                   (int)threadIdx.x, bar, parity);
            __trap();
        }
    }
}

__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const void* tmap, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}

// contiguous global -> shared bulk copy (16-byte aligned, size a multiple of 16), completing on an mbarrier
__device__ __forceinline__ void bulk_load_1d(uint32_t smem_dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t smem_result, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result), "r"(ncols)
                 : "memory");
}
__device__ __forceinline__ void tmem_relinquish() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// UMMA shared-memory descriptor split into words so the issuing thread advances it with one 32-bit add:
//   lo = start address >> 4 [0,14) | LBO >> 4 [16,30);  hi = SBO >> 4 [0,14) | version 1 [14,16) | SWIZZLE_128B (2) [29,32)
// (cute::UMMA::SmemDescriptor). Offsets added to `lo` are byte offsets >> 4 and never carry out of the 14-bit
// address field (shared memory is < 256 KB).
__device__ __forceinline__ uint32_t desc_lo(uint32_t saddr, uint32_t lbo_enc) {
    return ((saddr & 0x3FFFFu) >> 4) | (lbo_enc << 16);
}
__device__ __forceinline__ constexpr uint32_t desc_hi(uint32_t sbo_enc) { return sbo_enc | (1u << 14) | (2u << 29); }
__device__ __forceinline__ uint64_t desc_make(uint32_t lo, uint32_t hi) {
    uint64_t d;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "r"(lo), "r"(hi));
    return d;
}
// D[tmem] (+)= A[smem desc] * B[smem desc], bf16 inputs, fp32 accumulate
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem desc]: the A operand (M rows on the lanes, K along the columns, two bf16 per 32-bit
// column) is read from tensor memory instead of shared memory
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// mbarrier arrives when all previously issued tcgen05.mma of this thread have completed
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// ---- CTA-pair (cta_group::2) forms: two SMs of one TPC cooperate on a 256-row MMA, each feeding its own A rows
// and half of B from its shared memory. Only the even CTA (cluster rank 0) issues MMAs and owns the barriers the
// pair synchronises on; `& kPeerMask` turns a barrier's address into the same offset in the even CTA.
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t smem_dst, const void* tmap, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_dst), "l"(tmap), "r"(bar & kPeerMask), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void mbar_arrive_on_leader(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & kPeerMask) : "memory");
}
// ---- handing a word to the peer CTA of a pair: st.async writes it into the peer's shared memory and completes 4
// transaction bytes on the peer's mbarrier, so the reader needs nothing stronger than its usual CTA-scope wait.
// (A release.cluster arrive would cost a MEMBAR.GPU per hand-off, an acquire.cluster wait an L1 invalidate.)
__device__ __forceinline__ uint32_t mapa_rank(uint32_t saddr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
// Programmatic dependent launch: a kernel launched with cudaLaunchAttributeProgrammaticStreamSerialization may start
// while its predecessor in the stream is still running. pdl_trigger() lets the NEXT kernel start early; pdl_wait()
// returns once the PREVIOUS kernel has completed and its writes are visible. Both are no-ops in an ordinary launch.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
__device__ __forceinline__ float ld_cluster_f32(uint32_t cluster_addr) {
    float v;
    asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(cluster_addr) : "memory");
    return v;
}
__device__ __forceinline__ void st_async_s32(uint32_t cluster_addr, int v, uint32_t cluster_bar) {
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];"
                 ::"r"(cluster_addr), "r"(v), "r"(cluster_bar) : "memory");
}
__device__ __forceinline__ int ld_shared_s32(uint32_t saddr) {
    int v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(saddr) : "memory");
    return v;
}
__device__ __forceinline__ void st_shared_s32(uint32_t saddr, int v) {
    asm volatile("st.shared.s32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}
__device__ __forceinline__ void tmem_alloc_pair(uint32_t smem_result, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_result), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tmem_relinquish_pair() {
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrives on the barrier at this offset in BOTH CTAs of the pair once all earlier MMAs have completed
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
        ::"r"(bar), "h"((uint16_t)3)
        : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp gets row (lane base + i)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
          "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]),
          "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]),
          "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
          "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

}  // namespace penroz
