// fp32 GEMM on the FP32 pipes (FFMA), for fp32-parameter models: the reference's parity anchor (tiny GPT, CPU fp32,
// rtol 1e-4) and /output/ + /generate/ on fp32 checkpoints run without autocast (neural_net_model.py:273-298,
// 360-406), where tensor-core tf32 (10-bit mantissa) cannot hold 1e-4 on the logits.
// D[M,N] = sum_k A(m,k) * B(n,k); operands addressed through (row stride, col stride) so K-major and MN-major
// storage both work without copies. 64x64x16 tiles, 4x4 outputs per thread, k ascending (one FFMA chain per output).
#include "common.cuh"

namespace penroz {

struct GemmF32Params {
    const float* A; long long a_rs, a_cs;
    const float* B; long long b_rs, b_cs;
    int M, N, K;
    int epi;
    const float* bias;
    float* out; long long ldo;
    float* out2; long long ldo2;
    const float* aux; long long ldaux;
    int gelu_tanh, beta;
};

__global__ void __launch_bounds__(256)
gemm_f32_kernel(const GemmF32Params p) {
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    // loader mapping: contiguous global dimension varies fastest across threads
    const bool a_kfast = (p.a_cs == 1), b_kfast = (p.b_cs == 1);
    for (int k0 = 0; k0 < p.K; k0 += 16) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int e = tid + r * 256;
            int mm, kk;
            if (a_kfast) { mm = e >> 4; kk = e & 15; } else { kk = e >> 6; mm = e & 63; }
            const int gm = m0 + mm, gk = k0 + kk;
            As[kk][mm] = (gm < p.M && gk < p.K) ? p.A[(long long)gm * p.a_rs + (long long)gk * p.a_cs] : 0.f;
            int nn, kb;
            if (b_kfast) { nn = e >> 4; kb = e & 15; } else { kb = e >> 6; nn = e & 63; }
            const int gn = n0 + nn, gkb = k0 + kb;
            Bs[kb][nn] = (gn < p.N && gkb < p.K) ? p.B[(long long)gn * p.b_rs + (long long)gkb * p.b_cs] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= p.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n >= p.N) continue;
            float v = acc[i][j];
            float* o = p.out + (long long)m * p.ldo + n;
            switch (p.epi) {
                case 0: case 6: *o = v + (p.bias ? p.bias[n] : 0.f); break;
                case 1: {
                    const float pre = v + (p.bias ? p.bias[n] : 0.f);
                    *o = pre;
                    p.out2[(long long)m * p.ldo2 + n] = p.gelu_tanh ? gelu_tanh(pre) : gelu_erf(pre);
                } break;
                case 2: *o = p.aux[(long long)m * p.ldaux + n] + (v + (p.bias ? p.bias[n] : 0.f)); break;
                case 4: *o = p.beta ? *o + v : v; break;
                case 5: {
                    const float h = p.aux[(long long)m * p.ldaux + n];
                    *o = v * (p.gelu_tanh ? gelu_tanh_grad(h) : gelu_erf_grad(h));
                } break;
            }
        }
    }
}

}  // namespace penroz

using namespace penroz;

extern "C" int penroz_gemm_f32(const float* A, long long lda, int a_layout, const float* B, long long ldb,
                               int b_layout, int M, int N, int K, int epilogue, const float* bias, float* out,
                               long long ldo, float* out2, long long ldo2, const float* aux, long long ldaux,
                               int flags, void* stream) {
    PENROZ_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_f32: empty problem M=%d N=%d K=%d", M, N, K);
    PENROZ_REQUIRE(A && B && out, "gemm_f32: null operand");
    PENROZ_REQUIRE(epilogue == 0 || epilogue == 1 || epilogue == 2 || epilogue == 4 || epilogue == 5 || epilogue == 6,
                   "gemm_f32: unsupported epilogue %d", epilogue);
    if (epilogue == 1) PENROZ_REQUIRE(out2 != nullptr, "gemm_f32: GELU epilogue needs out2");
    if (epilogue == 2 || epilogue == 5) PENROZ_REQUIRE(aux != nullptr, "gemm_f32: epilogue %d needs aux", epilogue);
    GemmF32Params p;
    p.A = A; p.B = B;
    if (a_layout == 0) { p.a_rs = lda; p.a_cs = 1; } else { p.a_rs = 1; p.a_cs = lda; }
    if (b_layout == 0) { p.b_rs = ldb; p.b_cs = 1; } else { p.b_rs = 1; p.b_cs = ldb; }
    p.M = M; p.N = N; p.K = K; p.epi = epilogue; p.bias = bias;
    p.out = out; p.ldo = ldo; p.out2 = out2; p.ldo2 = ldo2; p.aux = aux; p.ldaux = ldaux;
    p.gelu_tanh = (flags & 1) ? 1 : 0;
    p.beta = (flags & 2) ? 1 : 0;
    dim3 grid((N + 63) / 64, (M + 63) / 64);
    PENROZ_REQUIRE(grid.y <= 65535, "gemm_f32: M=%d too large for this launch shape", M);
    gemm_f32_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    count_launch();
    return check_launch("gemm_f32");
}
