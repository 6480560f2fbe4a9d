// Weight-streaming Linear for the decode step, 1 <= M <= 64 activation rows (a CTA takes up to 16 of them), bf16.
//
// A decode-step Linear reads its whole weight matrix once for a handful of activation rows: 247 MB of weights per GPT-2
// 124M token step against 64 x 768 activations. It is bound by how evenly and how early all 148 SMs pull weight bytes,
// not by arithmetic. The 128-row tcgen05 tiles of the training GEMM are the wrong shape for that: at M = 64 the QKV
// projection (N = 2304) has 18 tiles, so 130 SMs stream nothing (measured: 1.2 ms of a 1.67 ms batch-64 step is in the
// Linears, whose bytes take 40 us). Here the unit of work is 16 weight rows (N / 16 units: 144 for QKV, 192 for fc,
// 3142 for the LM head), the K dimension is dealt to the 8 warps of a CTA, and the products run on the warp-level
// tensor-core path (mma.sync m16n8k16, bf16 x bf16 -> fp32: weight rows on the M side, the batch on the N side, "swap
// AB"), which has exactly that 16-row granularity and leaves the FP32 pipes free. The arithmetic is a by-product;
// what is engineered is the memory access:
//   * every lane fetches weights with 16-byte loads straight into the registers the MMA consumes. The reduction index
//     inside a 32-wide block is permuted identically for both operands (lane (g,t) takes k0 + 8t .. 8t+7 of rows g and
//     g+8 and of activation row g), which a dot product does not notice, so no shuffle or ldmatrix is needed;
//   * all weight loads of a unit are issued before anything waits (and, under programmatic dependent launch, before the
//     producer of the activations has finished); in the persistent form the next unit's weights are in flight while
//     the current one is reduced;
//   * the activation rows are staged once per CTA in shared memory as bf16 (LayerNorm applied on the way in when the
//     Linear follows one), padded so the 16-byte operand reads are conflict-free.
// Epilogue (after the 8 warps' partial sums meet in shared memory): bias, GELU or residual, rounding at torch's points,
// K/V columns of the QKV projection written straight into the page pool, per-CTA arg-max candidates for the LM head.
// Reference behaviour replaced: nn.Linear (+ LayerNorm / GELU / residual add around it) inside one iteration of the
// generation loop, neural_net_model.py:465-469 through neural_net_layers.py:83-92,130-135 and mappers.py:155-172.
#include "common.cuh"
#include "wstream.cuh"
#include <stdlib.h>
#include <string.h>

namespace penroz {

// Dynamic shared memory: xs[8 NT][kchunk + WS_XPAD] bf16 | red[WS_WARPS][8 NT][16] fp32
template <int NT>
__global__ void __launch_bounds__(WS_THREADS, 1)
wstream_linear_kernel(const WsParams p0, int kchunk) {
    extern __shared__ __align__(16) unsigned char ws_smem[];
    // blockIdx.y = group of 8 NT activation rows: everything indexed by row moves with it
    WsParams p = p0;
    if (gridDim.y > 1) {
        const int r0 = blockIdx.y * 8 * NT;
        p.M = min(8 * NT, p0.M - r0);
        p.x = p0.x + (size_t)r0 * p0.ldx;
        p.out = p0.out + (size_t)r0 * p0.ldo;
        if (p0.aux) p.aux = p0.aux + (size_t)r0 * p0.ldaux;
        if (p0.bt) p.bt = p0.bt + (size_t)r0 * p0.max_pages;
        if (p0.amax_val) { p.amax_val = p0.amax_val + (size_t)r0 * gridDim.x; p.amax_idx = p0.amax_idx + (size_t)r0 * gridDim.x; }
        if (p0.tok) { p.tok = p0.tok + r0; if (p0.emb_out) p.emb_out = p0.emb_out + (size_t)r0 * p0.K; }
    }
    constexpr int MB = 8 * NT;
    const int kc = min(kchunk, p.K);
    const int xstride = kc + WS_XPAD;
    __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(ws_smem);
    float* red = reinterpret_cast<float*>(ws_smem + (size_t)MB * xstride * sizeof(__nv_bfloat16));
    pdl_trigger();
    // ---- weights of the first unit: in flight before the activations exist
    uint4 fa[WS_MAXBLK][2];
    ws_prefetch_first(p, kchunk, blockIdx.x, fa);
    pdl_wait();
    ws_stage_rows<NT>(p, xs, xstride, 0, kc);
    ws_linear<NT>(p, xs, red, kchunk, blockIdx.x, gridDim.x, blockIdx.x, gridDim.x, fa);
}

static bool ws_use_pdl() {
    static const bool on = !(getenv("PENROZ_DECODE_PDL") && !strcmp(getenv("PENROZ_DECODE_PDL"), "0"));
    return on;
}

int ws_nt(int M) { return M <= 8 ? 1 : 2; }   // rows per CTA / 8 (see ws_groups)

// CTAs the launch uses for N output columns (= arg-max candidates per row)
// more than 16 activation rows run as groups of 16 (blockIdx.y): a CTA then stages and normalises 16 rows instead of
// all of them (measured at 64 rows: every one of 144 CTAs spent most of its 18 us on LayerNorm and staging of the same
// 64 x 768 activations) and streams 4 units of weights instead of one; the groups' reads of a weight row meet in L2
int ws_groups(int M) { return M > 16 ? (M + 15) / 16 : 1; }

int wstream_grid(int M, int N) {
    const int nunits = (N + 15) / 16;
    if (M > 16) {
        const int cap = max(1, num_sms() / ws_groups(M));
        return nunits < cap ? nunits : cap;
    }
    // one resident CTA per SM looping over its units, the next unit's weights in flight while the current one is reduced.
    // Measured at 124M bf16 (the kernel holds 255 registers, so extra CTAs only run as later waves): token step at batch 1
    // 347 us with 1 CTA per SM's worth of grid, 371 with 2, 375 with 4; batch 8: 494 / 518 / 522 (PENROZ_WS_PER_SM for A/B)
    static const int small_per_sm = getenv("PENROZ_WS_PER_SM") ? atoi(getenv("PENROZ_WS_PER_SM")) : 1;
    const int cap = num_sms() * small_per_sm;
    return nunits < cap ? nunits : cap;
}

bool wstream_supported(int M, int N, int K, int dtype) {
    return dtype == BF16 && M >= 1 && M <= 64 && K % 32 == 0 && K >= 32 && K <= 8192 && N >= 1;
}

size_t ws_smem_bytes(int nt, int kchunk) {
    return (size_t)(8 * nt) * (kchunk + WS_XPAD) * 2 + (size_t)WS_WARPS * (8 * nt) * 16 * 4;
}

// Columns of the activation rows staged in shared memory at once: all K when they fit, else a multiple of 1024 (the
// rest of a long reduction -- proj2's K = 4C at batch 64 -- is staged chunk by chunk inside the kernel).
int ws_kchunk(int nt, int K, size_t budget) {
    if (ws_smem_bytes(nt, K) <= budget) return K;
    int kc = (K / WS_KPASS) * WS_KPASS;
    while (kc > WS_KPASS && ws_smem_bytes(nt, kc) > budget) kc -= WS_KPASS;
    return kc;
}

int launch_wstream(const WsParams& p, cudaStream_t st) {
    PENROZ_REQUIRE(wstream_supported(p.M, p.N, p.K, BF16), "wstream_linear: unsupported shape M=%d N=%d K=%d", p.M, p.N, p.K);
    PENROZ_REQUIRE((((uintptr_t)p.x | (uintptr_t)p.W) & 15) == 0 && p.ldx % 8 == 0 && p.ldw % 8 == 0,
                   "wstream_linear: operands must be 16-byte aligned");
    const int nt = ws_nt(p.M);
    const size_t budget = 216 * 1024;
    const int kchunk = ws_kchunk(nt, p.K, budget);
    const size_t smem = ws_smem_bytes(nt, kchunk);
    PENROZ_REQUIRE(smem <= budget, "wstream_linear: M=%d K=%d needs %zu bytes of shared memory", p.M, p.K, smem);
    PENROZ_REQUIRE(!p.ln || p.K <= 2048, "wstream_linear: LayerNorm prologue needs K <= 2048 (got %d)", p.K);
    PENROZ_REQUIRE(!p.tok || kchunk >= p.K, "wstream_linear: embedding prologue needs the whole row staged");
    const int grid = wstream_grid(p.M, p.N);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid, ws_groups(p.M)); cfg.blockDim = dim3(WS_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = ws_use_pdl() ? 1 : 0;
    cudaError_t e = cudaSuccess;
    // cudaFuncSetAttribute costs microseconds of host time: raise a kernel's limit only when a launch needs more
    static std::atomic<int> granted[2][64];
    int dev = 0;
    cudaGetDevice(&dev);
#define PENROZ_WS(NTV, SLOT)                                                                                         \
    {                                                                                                                \
        auto k = wstream_linear_kernel<NTV>;                                                                         \
        if (smem > 48 * 1024 && (int)smem > granted[SLOT][dev & 63].load(std::memory_order_relaxed)) {               \
            e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);                     \
            if (e == cudaSuccess) granted[SLOT][dev & 63].store((int)smem, std::memory_order_relaxed);               \
        }                                                                                                            \
        if (e == cudaSuccess) e = cudaLaunchKernelEx(&cfg, k, p, kchunk);                                            \
    }
    if (nt == 1) PENROZ_WS(1, 0) else PENROZ_WS(2, 1)
#undef PENROZ_WS
    if (e != cudaSuccess) { set_error("wstream_linear launch: %s", cudaGetErrorString(e)); return PENROZ_ERR_CUDA; }
    count_launch();
    return check_launch("wstream_linear");
}

}  // namespace penroz

using namespace penroz;

// Stand-alone entry point: out[M,N] = epi(LN?(x)[M,K] W[N,K]^T + bias), bf16, M <= 64.
extern "C" int penroz_wstream_linear(const void* x, long long ldx, const void* W, long long ldw, const void* bias,
                                     void* out, long long ldo, const void* aux, long long ldaux, const void* ln_w,
                                     const void* ln_b, float ln_eps, int M, int N, int K, int epilogue, int flags,
                                     void* stream) {
    PENROZ_REQUIRE(x && W && out, "wstream_linear: null argument");
    PENROZ_REQUIRE(epilogue >= 0 && epilogue <= 2, "wstream_linear: bad epilogue %d", epilogue);
    if (epilogue == 2) PENROZ_REQUIRE(aux != nullptr, "wstream_linear: residual epilogue needs aux");
    WsParams p{};
    p.x = (const __nv_bfloat16*)x; p.ldx = ldx; p.W = (const __nv_bfloat16*)W; p.ldw = ldw;
    p.bias = (const __nv_bfloat16*)bias; p.out = (__nv_bfloat16*)out; p.ldo = ldo;
    p.aux = (const __nv_bfloat16*)aux; p.ldaux = ldaux; p.M = M; p.N = N; p.K = K; p.epi = epilogue;
    p.gelu_tanh = flags & 1;
    p.ln = ln_w != nullptr; p.ln_w = (const __nv_bfloat16*)ln_w; p.ln_b = (const __nv_bfloat16*)ln_b; p.ln_eps = ln_eps;
    return launch_wstream(p, (cudaStream_t)stream);
}
