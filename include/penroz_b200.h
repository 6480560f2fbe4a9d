/* penroz_b200 — C ABI of the B200 (sm_100a) kernels behind the GPT-2 DDP training step and the KV-cache decode
 * step of derinworks/penr-oz-neural-network-v3-torch-ddp.
 *
 * The reference has no native interface: its hot path is Python calling PyTorch. The entry points below are what a
 * binding for that path needs — one per PyTorch library call the reference makes on the path (file:line cited per
 * function, relative to the reference repo). They are called through ctypes from the drop-in Python modules
 * (neural_net_model / neural_net_layers / ddp / kv_cache) — see INTEGRATION.md.
 *
 * Conventions
 *   - every function returns 0 on success or a negative PENROZ_ERR_* code; penroz_last_error() gives the text
 *     (thread-local);
 *   - all pointers are DEVICE pointers unless stated; the caller owns every buffer, including scratch;
 *   - no entry point allocates, frees or synchronises unless its comment says so; work is enqueued on `stream`
 *     (a cudaStream_t passed as void*);
 *   - dtype codes: 0 = float32, 1 = bfloat16; matrices are row-major with a leading dimension in ELEMENTS;
 *   - the library is re-entrant: calls on any number of host threads and streams are independent. Its mutable globals
 *     are a launch counter, per-thread error text, and two pieces of device-resident state named where they are used:
 *     the tile-counter pairs of penroz_gemm_bf16 (1024 pairs, one dealt per call, reset by the call's last CTA; a
 *     launch that dies mid-kernel can leave its pair non-zero -- after a CUDA error the process should not go on
 *     using the library) and the fixed scratch array of penroz_colsum_static (calls sharing a device must not overlap).
 */
#ifndef PENROZ_B200_H
#define PENROZ_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PENROZ_OK 0
#define PENROZ_ERR_ARG (-1)
#define PENROZ_ERR_CUDA (-2)
#define PENROZ_ERR_ARCH (-3)
#define PENROZ_ERR_LIMIT (-4)

#define PENROZ_F32 0
#define PENROZ_BF16 1

/* GEMM epilogues (penroz_gemm_bf16 / penroz_gemm_f32) */
#define PENROZ_EPI_STORE 0      /* out = T(acc + bias)                                     */
#define PENROZ_EPI_GELU 1       /* out = pre = T(acc + bias); out2 = T(gelu(pre))          */
#define PENROZ_EPI_RESID_F32 2  /* out(f32) = aux(f32) + float(T(acc + bias))              */
#define PENROZ_EPI_RESID_BF16 3 /* out(bf16) = bf16(float(aux) + float(bf16(acc + bias)))  */
#define PENROZ_EPI_ACCUM_F32 4  /* out(f32) (+)= acc   (flags bit 1 = accumulate)          */
#define PENROZ_EPI_DGELU 5      /* out = T(T(acc) * gelu'(aux)); penroz_gemm_bf16: a non-null out2 is an fp32 [N]
                                   vector that receives += the column sums of out (bias gradient of the Linear in
                                   front of the activation)                                 */
#define PENROZ_EPI_F32 6        /* out(f32) = acc + bias                                   */
#define PENROZ_EPI_STORE_ROWDOT 7 /* penroz_gemm_bf16, shapes of the CTA-pair kernel (M >= 256, N > 128, N % 64 == 0):
                                   out = bf16(acc + bias) and, for every 64-column group g of row r,
                                   out2(f32)[((r / T)·(N/64) + g)·T + r % T] = sum_c out[r,64g+c]·aux[r,64g+c]
                                   with T = ldo2 rows per sequence: the attention backward's delta, produced by the
                                   GEMM that computes dO (aux = the attention output)          */
/* flags: bit0 tanh-GELU (else erf), bit1 accumulate into out, bit2 force 128-wide tile, bit3 force 256-wide tile */

/* ---- library ---- */
const char* penroz_last_error(void);
int penroz_abi_version(void);
/* number of kernel launches this process has issued through the library (for bench.py's gpu_launches) */
unsigned long long penroz_launch_count(void);
/* fails with PENROZ_ERR_ARCH unless the current device is compute capability 10.x */
int penroz_query(int* cc_major, int* cc_minor, int* sm_count);

/* ---- dense contractions: D[M,N] = sum_k A(m,k) B(n,k) ----
 * Replaces torch nn.Linear forward and its autograd backward (modules built at mappers.py:21,156,158,163,165,172;
 * run at neural_net_model.py:250-258; backward at neural_net_model.py:650-653).
 * a_layout/b_layout: 0 = operand stored [rows, K] (K contiguous), 1 = stored [K, rows] (rows contiguous).
 * penroz_gemm_bf16: tcgen05.mma (TMEM accumulators) fed by TMA; bias is bf16. penroz_gemm_f32: FP32 FMA pipes.
 * penroz_gemm_bf16 draws its output tiles from a counter pair inside the library (1024 pairs, one dealt per call,
 * reset by the call's last CTA): calls on any number of streams or threads are independent, but ONE captured launch
 * replayed concurrently with itself would share its pair -- replay such a graph on one stream at a time. */
int penroz_gemm_bf16(const void* A, long long lda, int a_layout, const void* B, long long ldb, int b_layout, int M,
                     int N, int K, int epilogue, const void* bias, void* out, long long ldo, void* out2,
                     long long ldo2, const void* aux, long long ldaux, int flags, void* stream);
int penroz_gemm_f32(const float* A, long long lda, int a_layout, const float* B, long long ldb, int b_layout, int M,
                    int N, int K, int epilogue, const float* bias, float* out, long long ldo, float* out2,
                    long long ldo2, const float* aux, long long ldaux, int flags, void* stream);

/* ---- embedding: out[b,t,:] = wte[idx[b,t]] + wpe[t + pos_offset] ----
 * Replaces Summation(nn.Embedding, PositionEmbedding) (neural_net_layers.py:98-127) and its backward.
 * err_flag (device int, may be NULL) is set to 1 when an index is outside [0, V). */
int penroz_embed_fwd(const long long* idx, const void* wte, const void* wpe, void* out, int B, int T, int C, int V,
                     int pos_offset, int w_dtype, int out_dtype, int* err_flag, void* stream);
int penroz_embed_bwd(const long long* idx, const float* dx, float* dwte, float* dwpe, int B, int T, int C, int V,
                     int pos_offset, void* stream);

/* int32 token ids -> int64 on the device. Replaces the host-side widening of every batch,
 * torch.tensor(arr, dtype=torch.long).view(B, T).to(device) (neural_net_model.py:630-631): the
 * loader's int32 window is uploaded as it is and widened here. */
int penroz_widen_tokens(const int* src, long long* dst, long long n, void* stream);

/* ---- LayerNorm (eps as given; statistics in fp32) ----
 * Replaces nn.LayerNorm forward/backward (mappers.py:31,155,162,171). beta may be NULL (bias=False).
 * Backward also folds in the residual-stream gradient (`add`, neural_net_layers.py:130-135), can emit a bf16 copy
 * of dx for the next matmul, and accumulates (+=) dgamma, dbeta and — optionally — dcol = column sums of dx
 * (the bias gradient of the Linear feeding the residual add). Needs caller scratch. */
int penroz_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* out, float* mean, float* rstd,
                         int rows, int C, float eps, int x_dtype, int w_dtype, int out_dtype, void* stream);
size_t penroz_layernorm_bwd_scratch_bytes(int rows, int C);
int penroz_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                         const void* add, void* dx, void* dx_bf16, float* dgamma, float* dbeta, float* dcol,
                         int rows, int C, int x_dtype, int w_dtype, int dy_dtype, void* scratch,
                         size_t scratch_bytes, void* stream);

/* ---- GELU (erf or tanh), standalone forms; the bf16 GEMM fuses both into its epilogues ----
 * Replaces nn.GELU (mappers.py:25,139,164). */
int penroz_gelu_fwd(const void* x, void* y, long long n, int tanh_approx, int dtype, void* stream);
int penroz_gelu_bwd(const void* dy, const void* x, void* dx, long long n, int tanh_approx, int dtype, void* stream);

/* ---- column sums: out[n] (+)= sum_m X[m,n] (bias gradients of nn.Linear) ---- */
size_t penroz_colsum_scratch_bytes(int M, int N);
int penroz_colsum(const void* X, long long ld, int M, int N, float* out, int accumulate, int dtype, void* scratch,
                  size_t scratch_bytes, void* stream);

/* ---- column sums with no caller scratch and no allocation: out[n] (+)= sum_m X[m,n], N <= 8192 ----
 * No counterpart in the reference (it is a variant of penroz_colsum added for this library's own callers).
 * Who needs it: a host that binds this C ABI from outside PyTorch — for example the reference service's worker
 * calling through a ctypes/cffi stub — receives device pointers for inputs and outputs from its framework but owns
 * no device allocator; and inside a training worker PyTorch's caching allocator has usually reserved almost all of
 * HBM, so a raw cudaMalloc from such a host can fail (or stall the device while it synchronises) even though plenty
 * of memory is "free" inside torch. For those callers the partial-sum buffer is a fixed 1 MiB array (32 x 8192
 * floats) that is part of the library image: every GPU that loads the library holds it until the process exits.
 *   - N > PENROZ_COLSUM_STATIC_MAX_COLS is rejected with PENROZ_ERR_LIMIT; use penroz_colsum for wider matrices.
 *   - The first call on each device looks up the array's device address (one cudaGetSymbolAddress, that device
 *     current) and remembers it; make that first call before starting any stream capture. Later calls only launch.
 *   - The call never allocates or frees: no cudaMalloc/cudaFree, no library routine that allocates.
 *   - All calls on a device share the one array and it keeps whatever the previous call left in it. Each call
 *     overwrites every slot it later reads, so stale contents are harmless, but two calls whose kernels overlap on
 *     the GPU (different streams, or different host threads on different streams) race on it: issue them on one
 *     stream, or order them with events. */
#define PENROZ_COLSUM_STATIC_MAX_COLS 8192
int penroz_colsum_static(const void* X, long long ld, int M, int N, float* out, int accumulate, int dtype,
                         void* stream);

/* ---- dtype cast of a flat buffer (autocast's per-step weight cast, done once per optimizer step here) ---- */
int penroz_cast(const void* in, void* out, long long n, int in_dtype, int out_dtype, void* stream);

/* ---- y += alpha * x over flat fp32 buffers, alpha in device memory ----
 * What autograd's AccumulateGrad does per parameter when `cost.backward()` is called on the cost forward() returned
 * (neural_net_model.py:250-271, :650-653): here ONE pass over the flat gradient buffer, scaled by the incoming
 * gradient of the cost without reading it on the host. */
int penroz_axpy_f32(float* y, const float* x, const float* alpha, long long n, void* stream);

/* ---- causal self-attention on the packed [B,T,3C] QKV buffer (exact fp32 arithmetic path) ----
 * Replaces CausalSelfAttention.forward, GPT-2 branch (neural_net_layers.py:59-95) and its backward.
 * out is [B,T,C]; lse is [B,H,T] fp32. D in {32, 64, 128}. */
int penroz_attn_fwd_simt(const void* qkv, void* out, float* lse, int B, int T, int H, int D, int dtype, void* stream);
size_t penroz_attn_bwd_scratch_bytes(int B, int T, int H);
int penroz_attn_bwd_simt(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int B,
                         int T, int H, int D, int dtype, void* scratch, size_t scratch_bytes, void* stream);

/* ---- the same attention on the tensor cores (bf16, head_dim 64): tcgen05.mma with S / dP / dK / dV / dQ in TMEM,
 * q, k, v and dO tiles fetched from the packed buffers by TMA. Replaces the same reference lines as above for the
 * bf16-autocast training step. The backward needs scratch (row dot-products + an fp32 dQ accumulator it zeroes).
 * dbias (optional, fp32 [3C]) receives += the column sums of dqkv as written (bf16): the bias gradient of the QKV Linear
 * (nn.Linear backward, mappers.py:156). flags bit 0: the row dot-products delta[b,h,t] = sum_d dO·O are already at the
 * head of scratch (B·H·T floats) -- written by the GEMM that produced dout (PENROZ_EPI_STORE_ROWDOT) -- and the pass that
 * would compute them is skipped. */
int penroz_attn_fwd_tc(const void* qkv, void* out, float* lse, int B, int T, int H, int D, void* stream);
size_t penroz_attn_bwd_tc_scratch_bytes(int B, int T, int H, int D);
int penroz_attn_bwd_tc(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* dbias,
                       int flags, int B, int T, int H, int D, void* scratch, size_t scratch_bytes, void* stream);

/* ---- softmax cross-entropy over a chunk of logits rows ----
 * Replaces F.cross_entropy(logits.view(-1,V), target.view(-1)) (neural_net_model.py:264-267): mean over rows whose
 * target != ignore_index. ce_rows writes per-row losses and (write_grad) overwrites the logits rows in place with
 * (softmax - onehot) * grad_scale / n_valid. err_flag is set to 2 for a target outside [0, V). */
int penroz_ce_count_valid(const long long* targets, int M, long long ignore_index, int* n_valid, void* stream);
int penroz_ce_rows(void* logits, long long ld, const long long* targets, int M, int V, long long ignore_index,
                   const int* n_valid, float grad_scale, float* loss_rows, int write_grad, int dtype, int* err_flag,
                   void* stream);
int penroz_ce_reduce(const float* loss_rows, int M, const int* n_valid, float out_scale, float* loss_out,
                     int accumulate, void* stream);

/* ---- AdamW over flat fp32 buffers + global gradient norm ----
 * Replaces torch.optim.AdamW.step()/zero_grad() (mappers.py:267-274; neural_net_model.py:622,677), same operation
 * order as torch's single-tensor AdamW. Optionally writes the bf16 shadow weights and zeroes the gradients in the
 * same pass. grad_mult (device float, may be NULL) scales gradients (clipping; the reference has none). */
int penroz_adamw_flat(float* p, float* g, float* m, float* v, void* shadow_bf16, long long n, double lr, double beta1,
                      double beta2, double eps, double weight_decay, long long step, const float* grad_mult,
                      int zero_grad, void* stream);
/* The same AdamW pass that also measures the update it applies. Replaces, for the progress entry the
 * reference records every epoch, the clone of every 2-D weight before the step (neural_net_model.py:619-620) and
 * std(w - w_prev) / (std(w) + 1e-8) per weight after it (:696-699): the kernel holds w_prev and w in registers anyway.
 * n must be a multiple of 64; seg_of_group[n / 64] (device int32) gives, per 64-element group of the flat buffer, the
 * row of seg_stats the group's tensor accumulates into, or -1; seg_stats[row][4] (device doubles, zeroed by the caller)
 * receives sum(d), sum(d^2), sum(w), sum(w^2) with d = w - w_prev formed in fp32 as the reference forms it. */
int penroz_adamw_flat_stats(float* p, float* g, float* m, float* v, void* shadow_bf16, long long n, double lr,
                            double beta1, double beta2, double eps, double weight_decay, long long step,
                            const float* grad_mult, int zero_grad, const int* seg_of_group, double* seg_stats,
                            void* stream);
size_t penroz_grad_norm_scratch_bytes(long long n);
int penroz_grad_norm(const float* g, long long n, float max_norm, float* norm_out, float* mult_out, void* scratch,
                     size_t scratch_bytes, void* stream);

/* ---- rank-0 training diagnostics ----
 * Replaces, for CUDA tensors, the per-layer / per-weight summaries of _record_training_overall_progress
 * (neural_net_model.py:743-777): t.mean(), t.std() (unbiased), a saturation fraction |t| > abs_threshold and
 * torch.histogram(t.cpu(), density=True) -- which copies every activation to the host and bins it on one CPU thread.
 * out (device, 7 + bins doubles) = {mean, std, min, max, saturated fraction, range lo, range hi, counts[bins]}.
 * Counts follow ATen's CPU histogram for uniform bins step by step (float position estimate, local search against the
 * torch.linspace edges, closed right edge, a constant tensor widens its range by 0.5 each side). One tensor = two
 * passes over HBM; the caller reads `out` back with a single copy. */
size_t penroz_tensor_stats_scratch_bytes(int bins);
int penroz_tensor_stats(const void* x, long long n, int dtype, int bins, float abs_threshold, double* out,
                        void* scratch, size_t scratch_bytes, void* stream);

/* ---- KV-cache decode step ----
 * Replaces KVCache.append's torch.cat (kv_cache.py:41-68), the cache branch of CausalSelfAttention.forward
 * (neural_net_layers.py:83-92) and logits[:, -1, :].argmax (neural_net_model.py:393-394).
 * pool is [n_pages][2][H][PAGE][D]; block_table is int32 [B][max_pages]. */
int penroz_kv_page_size(void);
int penroz_kv_append(const void* qkv, void* pool, const int* block_table, int max_pages, int B, int S, int H, int D,
                     int t0, int dtype, void* stream);
size_t penroz_attn_decode_scratch_bytes(int B, int H, int D, int t_len);
int penroz_attn_decode(const void* qkv, const void* pool, const int* block_table, int max_pages, void* out, int B,
                       int H, int D, int t_len, int dtype, void* scratch, size_t scratch_bytes, void* stream);
/* ---- int8 KV cache (TURBO_QUANT_KV_CACHE=1) ----
 * Replaces TurboQuantKVCache._quantize / append / _dequantize (kv_cache.py:119-125, 137-168, 135): one scale per cached
 * token and head (amax|x| / 127, 1 where that is 0), codes = round(x / scale) clamped to [-128, 127]. codes is an int8
 * pool [n_pages][2][H][PAGE][D], scales a pool [n_pages][2][H][PAGE] of the model dtype, both addressed through the
 * same block table as the exact cache. kv_append_q8 quantises the K and V columns of S new rows of qkv [B*S, 3C];
 * writeback != 0 also overwrites those columns with codes.float() * scale (what the reference's append hands to the
 * prompt's own attention). recip_scale != 0 computes the scale as amax * (1/127) in fp32 -- what ATen does for a CUDA
 * tensor divided by a Python scalar, so codes and scales equal the torch class's bit for bit on the GPU; 0 is the true
 * division of CPU tensors. attn_decode_q8 is penroz_attn_decode reading that cache, de-quantising inside the kernel. */
int penroz_kv_append_q8(void* qkv, void* codes, void* scales, const int* block_table, int max_pages, int B, int S,
                        int H, int D, int t0, int dtype, int writeback, int recip_scale, void* stream);
size_t penroz_attn_decode_q8_scratch_bytes(int B, int H, int D);
int penroz_attn_decode_q8(const void* qkv, const void* codes, const void* scales, const int* block_table,
                          int max_pages, void* out, int B, int H, int D, int t_len, int dtype, void* scratch,
                          size_t scratch_bytes, void* stream);
/* few-row Linear (M small): epilogue 0 store, 1 GELU, 2 residual add; out_f32 (optional) receives an fp32 copy */
int penroz_skinny_linear(const void* x, long long ldx, const void* W, long long ldw, const void* bias, void* out,
                         long long ldo, const void* aux, long long ldaux, float* out_f32, long long ldo32, int M,
                         int N, int K, int epilogue, int flags, int dtype, void* stream);
int penroz_argmax_rows(const void* x, long long ld, int rows, int V, long long* out, int dtype, void* stream);
/* Weight-streaming Linear for 1 <= M <= 64 activation rows, bf16: out[M,N] = epi(LN?(x)[M,K] W[N,K]^T + bias).
 * Replaces nn.Linear inside the generation loop (mappers.py:156-172 run at neural_net_model.py:465-469) together with
 * the LayerNorm in front of it (ln_w != NULL: nn.LayerNorm with ln_eps, mappers.py:155,162,171), the GELU behind it
 * (epilogue 1; flags bit0 = tanh form) or the residual add (epilogue 2, aux; neural_net_layers.py:130-135). Units of 16
 * weight rows are dealt to every SM, the products run on mma.sync m16n8k16 (weights on the M side), every weight byte
 * is read once with 16-byte loads issued before the activations are waited for. K % 32 == 0; LayerNorm needs K <= 2048. */
int penroz_wstream_linear(const void* x, long long ldx, const void* W, long long ldw, const void* bias, void* out,
                          long long ldo, const void* aux, long long ldaux, const void* ln_w, const void* ln_b,
                          float ln_eps, int M, int N, int K, int epilogue, int flags, void* stream);

/* ---- one whole single-token decode step of a GPT-2 tree, enqueued with a single host call ----
 * Replaces one iteration of the generation loop (neural_net_model.py:465-469, incremental branch :382-386): embedding
 * at position `pos`, every block (LN, QKV, paged K/V append at `pos`, split-KV attention over pos+1 keys, proj+residual,
 * LN, fc+GELU, proj2+residual), final LN, LM head into logits_out [B,V] (model dtype) and, when argmax_out is not NULL,
 * the greedy token ids. `layers` is a HOST array; every pointer inside it is a device pointer.
 * With PENROZ_DECODE_MEGA=1 (bf16, B <= 64, n_embd <= 1024) the whole step is ONE persistent cooperative kernel
 * (decode_mega.cu; measured slower than the default on B200, kept as an experiment). By default:
 * The step runs as five launches per block (decode_fused.cu): LayerNorm, bias, GELU, residual and the K/V page write
 * ride in the prologue / epilogue of weight-streaming Linear kernels (bf16, B <= 64: penroz_wstream_linear's kernel;
 * fp32, B <= 8: an FP32-pipe GEMV), the split-KV attention merges its partials inside a thread-block cluster, and
 * consecutive launches overlap through programmatic dependent launch; larger batches, and the int8 cache, take one
 * kernel per op. Same results up to summation order. */
typedef struct penroz_gpt2_layer_t {
    const void *ln1_w, *ln1_b, *qkv_w, *qkv_b, *proj_w, *proj_b, *ln2_w, *ln2_b, *fc_w, *fc_b, *proj2_w, *proj2_b;
    void* kv_pool;
    float ln1_eps, ln2_eps;
    int gelu_tanh, ffn;
    void* kv_scales;        /* int8 cache (kv_quant != 0): kv_pool holds int8 codes, this the per-token scales; else NULL */
    int kv_quant, kv_recip; /* kv_recip: see penroz_kv_append_q8's recip_scale */
} penroz_gpt2_layer_t;
/* The workspace must be ZERO when it is first handed to penroz_gpt2_decode_step (its head holds the persistent
 * kernel's barrier and per-head counters); every step leaves those words zero again, so one cudaMemset at allocation
 * is enough. A workspace is used by one step at a time (steps on one stream; one workspace per concurrent generation). */
size_t penroz_gpt2_decode_workspace_bytes(int B, int C, int H, int F, int t_len_max);
/* The prompt pass that starts a generation (neural_net_model.py:387-390 with an empty cache): all S positions through
 * every block with K/V appended at positions 0..S-1, final LayerNorm + LM head on the last position of each sequence
 * only; logits_out [B,V]. Same kernels as the per-op entry points, one host call. */
size_t penroz_gpt2_prefill_workspace_bytes(int B, int S, int C, int F);
int penroz_gpt2_prefill(const long long* tokens, const void* wte, const void* wpe, const penroz_gpt2_layer_t* layers,
                        int n_layer, const void* lnf_w, const void* lnf_b, float lnf_eps, const void* head_w,
                        const void* head_b, const int* block_table, int max_pages, int B, int S, int C, int H, int V,
                        int dtype, void* workspace, size_t workspace_bytes, void* logits_out, long long* argmax_out,
                        int* err_flag, void* stream);
int penroz_gpt2_decode_step(const long long* tokens, const void* wte, const void* wpe,
                            const penroz_gpt2_layer_t* layers, int n_layer, const void* lnf_w, const void* lnf_b,
                            float lnf_eps, const void* head_w, const void* head_b, const int* block_table,
                            int max_pages, int B, int C, int H, int V, int pos, int dtype, void* workspace,
                            size_t workspace_bytes, void* logits_out, long long* argmax_out, int* err_flag,
                            void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PENROZ_B200_H */
