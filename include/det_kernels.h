/* det_kernels.h — C ABI of libdet_kernels.so (hand-written sm_100a kernels).
 *
 * Every function returns 0 on success; on failure det_last_error() holds a message (thread-local).
 * Pointers are device pointers unless a parameter is documented as a host array. Kernels are enqueued on
 * `stream` and never synchronise. bf16 tensors are row-major with the given leading dimension (elements).
 * Reference citations say which library call of the reference's stack (PyTorch 1.4 + Horovod/NCCL + apex;
 * the reference repository itself contains no GPU code) each entry point replaces.
 */
#ifndef DET_KERNELS_H_
#define DET_KERNELS_H_

#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* det_last_error(void);
/* Cap the number of SMs the persistent kernels size their grids from (0 = whole device). Used while a collective runs
 * on a side stream so its CTAs find free SMs. Host-side state, takes effect for subsequent launches. */
int det_set_sm_limit(int n);
int det_kernels_abi_version(void);

/* ---- tensor-core GEMM (tcgen05 + TMEM + TMA) -------------------------------------------------------------
 * C[M,N] = epilogue(alpha * A[M,K] B[N,K]^T). a_mn/b_mn: 0 = K-major ([M,K] / [N,K] storage), 1 = MN-major
 * ([K,M] / [K,N] storage). out_mode: 0 bf16, 1 fp32, 2 fp32 atomic accumulate (split-K allowed).
 * epi: 0 none, 1 +bias, 2 +bias then GELU (pre-activation to aux_out), 3 *gelu'(aux_in), 4 dropout(+bias)+aux_in.
 * Replaces torch.addmm / cuBLAS behind nn.Linear (reference: examples/nlp/bert_glue_pytorch/model_def.py:60-68). */
int det_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int a_mn, int b_mn, long long lda,
                  long long ldb, long long ldc, int out_mode, int epi, float alpha, const float* bias,
                  const void* aux_in, void* aux_out, long long ld_aux, float dropout_p, unsigned int dropout_seed,
                  const unsigned int* dropout_seed_dev, int splitk, int force_bn, cudaStream_t stream);

/* Strided CUDA-core GEMM for tiny heads (pooler / classifier): C[m,n] (+)= act(alpha*sum_k A[m,k]B[n,k]+bias[n]). */
int det_gemm_small(const void* A, const void* B, void* C, int M, int N, int K, long long sam, long long sak,
                   long long sbn, long long sbk, long long scm, long long scn, int a_f32, int b_f32, int c_f32,
                   const float* bias, int act, float alpha, int accumulate, cudaStream_t stream);

/* ---- fused self-attention, seq <= 128, head_dim 64 (reference: HF BertSelfAttention via model_def.py:60) --- */
int det_attention_fwd(const void* qkv, const float* mask, void* ctx, float* lse, int B, int S, int H,
                      float dropout_p, unsigned int seed, const unsigned int* seed_dev, cudaStream_t stream);
int det_attention_bwd(const void* qkv, const void* dctx, const float* mask, const float* lse, void* dqkv, int B,
                      int S, int H, float dropout_p, unsigned int seed, const unsigned int* seed_dev,
                      cudaStream_t stream);

/* ---- tiled (online-softmax) self-attention, any sequence length, head_dim 64; same tensors as above.
 *      bwd also takes the forward output `ctx`, an fp32 scratch dq_accum[B*S, H*64] and an fp32 scratch delta[B, H, S]. */
int det_attention_flash_fwd(const void* qkv, const float* mask, void* ctx, float* lse, int B, int S, int H,
                            float dropout_p, unsigned int seed, const unsigned int* seed_dev, cudaStream_t stream);
int det_attention_flash_bwd(const void* qkv, const void* dctx, const void* ctx, const float* mask, const float* lse,
                            void* dqkv, float* dq_accum, float* delta, int B, int S, int H, float dropout_p,
                            unsigned int seed, const unsigned int* seed_dev, cudaStream_t stream);

/* ---- depthwise convolution (groups == channels), NHWC bf16, stride / padding / dilation; w and dw are fp32 tap-major
 *      [R*S, C]; dgrad takes the INPUT sizes H, W; wgrad accumulates (+=). Replaces torch.nn.Conv2d(groups=C) → cuDNN of the
 *      reference's NAS examples (examples/hp_search_benchmarks/darts_cifar10_pytorch/operations.py). */
int det_dwconv2d_fwd(const void* x, const float* w, const float* bias, void* y, int N, int H, int W, int C, int R, int S,
                     int stride, int pad, int dil, int relu, cudaStream_t stream);
int det_dwconv2d_dgrad(const void* dy, const float* w, void* dx, int N, int H, int W, int C, int R, int S, int stride, int pad,
                       int dil, cudaStream_t stream);
int det_dwconv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int R, int S, int stride, int pad,
                       int dil, cudaStream_t stream);

/* windowed average pooling (torch.nn.AvgPool2d); bwd takes the INPUT sizes H, W */
int det_avgpool2d_fwd(const void* x, void* y, int N, int H, int W, int C, int k, int stride, int pad, int count_include_pad,
                      cudaStream_t stream);
int det_avgpool2d_bwd(const void* dy, void* dx, int N, int H, int W, int C, int k, int stride, int pad, int count_include_pad,
                      cudaStream_t stream);

/* ---- recurrent-cell gate: s_j = sp + sigmoid(c_j) * (f_j(h_j) - sp) for the k nodes sharing predecessor sp; ch is
 * [B, k*2*nh] with node j's [c | h] at columns j*2*nh; acts: 0 tanh, 1 relu, 2 sigmoid, 3 identity; mask [B, nh] or null
 * (then the masked copies s_j * mask are written too). Pointer arrays are host arrays of k device pointers. */
int det_highway_gate_fwd(const void* ch, const void* sp, const void* mask, void* const* s_out, void* const* sm_out,
                         const int* acts, int k, int B, int nh, cudaStream_t stream);
int det_highway_gate_bwd(const void* ch, const void* sp, const void* mask, const void* const* ds, const void* const* dsm,
                         void* dch, void* dsp, const int* acts, int k, int B, int nh, cudaStream_t stream);

/* ---- LayerNorm / reductions / embeddings / loss (reference: torch.nn.LayerNorm, Embedding, CrossEntropyLoss) */
int det_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                      int T, int H, float eps, float dropout_p, unsigned int seed, const unsigned int* seed_dev,
                      cudaStream_t stream);
int det_layernorm_bwd(const void* dy, const void* x, const float* gamma, const float* mean, const float* rstd,
                      void* dx, void* dx_masked, float* dgamma, float* dbeta, float* dxsum, int T, int H, float in_dropout_p,
                      unsigned int in_seed, float out_dropout_p, unsigned int out_seed,
                      const unsigned int* seed_dev, cudaStream_t stream);
int det_colsum_bf16(const void* x, float* out, int T, int N, long long ld, cudaStream_t stream);
int det_colsum_select_bf16(const void* x, const long long* sel, long long sel_value, float* out, int T, int N,
                           long long ld, cudaStream_t stream);
int det_embed_gather(const long long* ids, const long long* type_ids, const float* word, const float* pos,
                     const float* type, void* out, int T, int S, int H, cudaStream_t stream);
int det_embed_scatter(const long long* ids, const long long* type_ids, const void* de, float* dword, float* dpos,
                      float* dtype, int T, int S, int H, cudaStream_t stream);
int det_softmax_ce(const float* logits, const long long* labels, float* loss, float* dlogits, int B, int C,
                   float gscale, cudaStream_t stream);
int det_cast_f32_bf16(const float* in, void* out, long long n, cudaStream_t stream);

/* ---- fused optimizers on flat buckets (reference: optimizer.step()/zero_grad()/_average_gradients/clip_grads,
 *      harness/determined/pytorch/_pytorch_context.py:470-545; apex.amp loss scaling :444-462) -------------
 * variant: 0 Adam, 1 AdamW, 2 SGD(momentum), 3 RMSprop, 4 Adadelta. hyper_dev: 12 fp32 (see fused_optim.cu). */
int det_fused_optimizer_step(int variant, float* param, float* grad, float* state1, float* state2,
                             void* shadow_bf16, long long n, const float* hyper_dev, const float* grad_scale_dev,
                             const int* found_inf_dev, int zero_grad, cudaStream_t stream);
/* sumsq_dev (+=) and found_inf_dev (set) are accumulators the CALLER allocates and clears. */
int det_grad_sumsq_finite(const float* grad, long long n, const float* grad_scale_dev, float* sumsq_dev,
                          int* found_inf_dev, cudaStream_t stream);
int det_clip_coef(const float* sumsq_dev, float max_norm, const float* base_scale_dev, float* out2_dev,
                  cudaStream_t stream);
/* Gradient L2 norm + non-finite flag for callers that cannot allocate device memory (a hook inside a trainer
 * whose memory is all held by the framework's allocator; a C plugin called inside the step loop). Same math as
 * det_grad_sumsq_finite, but the scratch is a fixed ~256 KB array inside the library (held by every GPU that
 * loads it), so n is limited to det_grad_norm_noalloc_max_elems() = 2^30 and larger inputs are refused.
 * out2_dev[0] = norm, out2_dev[1] = norm^2; found_inf_out_dev (nullable) = 0/1. The first call on a device
 * resolves the array's address: make it before any stream capture. Calls on one device share the array and
 * must not overlap on the GPU. Allocates nothing. No counterpart in the reference. */
long long det_grad_norm_noalloc_max_elems(void);
int det_grad_norm_noalloc(const float* grad, long long n, const float* grad_scale_dev, float* out2_dev,
                          int* found_inf_out_dev, cudaStream_t stream);
int det_loss_scale_update(float* state3_dev, int* found_inf_dev, float growth, float backoff, int growth_interval,
                          float divisor, cudaStream_t stream);

/* ---- in-kernel collectives over NVLink peer memory (reference: Horovod/NCCL allreduce + broadcast,
 *      _pytorch_context.py:192-198, _pytorch_trial.py:71-73; ZMQ metric gather _pytorch_trial.py:292-335) -----
 * peer_bufs / peer_flags are HOST arrays of `world` device pointers (this rank's mapping of every rank). */
long long det_comm_flag_bytes(void);
long long det_comm_epoch_bytes(void);
int det_fused_allreduce(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr, void* epoch_dev,
                        long long n_f32, int rank, int world, float scale, const float* scale_dev,
                        float* sumsq_dev, int* found_inf_dev, int max_blocks, cudaStream_t stream);
/* Same collective with an explicit CTA size and algorithm. `threads` = 128 gives CTAs small enough to be co-resident with
 * a GEMM CTA on the same SM; 0 = the default 1024. `algo`: 0 = automatic (also DET_AR_ALGO), 1 = multimem.ld_reduce /
 * multimem.st through the NVLS multicast address, 2 = 16-byte peer loads / stores, 3 = TMA bulk copies over peer memory
 * (cp.async.bulk, 3 x 64 KB stages per CTA). On return sumsq_dev / found_inf_dev describe the WHOLE bucket on every rank
 * (the per-rank partials are exchanged inside the kernel). */
int det_fused_allreduce_cta(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr, void* epoch_dev,
                            long long n_f32, int rank, int world, float scale, const float* scale_dev,
                            float* sumsq_dev, int* found_inf_dev, int max_blocks, int threads, int algo,
                            cudaStream_t stream);
/* Pipelined all-reduce (all-gather under the optimizer). The bucket is `chunks` (<= 8) chunks of det_pipe_chunk_floats()
 * floats, each chunk `world` equal shares, rank r owning share r of every chunk. 1: reduce-scatter + scale + whole-bucket
 * statistics, the reduced shares stay in the owner's bucket. 2 (every rank, c = 0 .. chunks-1 in order): push the own share of
 * chunk c into every rank's bucket and raise its flag there. 3 (every rank): hold the stream until every OTHER rank's share of
 * chunk c has landed here — one warp spins; the working kernels never wait. */
long long det_pipe_chunk_floats(long long n_f32, int world, int chunks);
int det_reduce_scatter_stats(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr, void* epoch_dev, long long n_f32,
                             int rank, int world, int chunks, float scale, const float* scale_dev, float* sumsq_dev,
                             int* found_inf_dev, int max_blocks, cudaStream_t stream);
int det_allgather_send(void* const* peer_bufs, void* const* peer_flags, void* mc_ptr, void* epoch_dev, long long n_f32, int rank,
                       int world, int chunks, int c, int max_blocks, cudaStream_t stream);
int det_wait_chunk(void* local_flags, int c, int rank, int world, cudaStream_t stream);
/* bf16 on the wire (optimizations.gradient_compression): pack `grad` (local fp32, n % 8 == 0) into this rank's bf16 staging
 * buffer, all-reduce the staging buffers (multimem bf16x2 with fp32 accumulation when mc_wire is given, 16-byte peer loads
 * summed in fp32 otherwise; scale, non-finite test and grad-norm partial as above), unpack into `grad`. */
int det_fused_allreduce_bf16wire(float* grad, void* const* peer_wire, void* const* peer_flags, void* mc_wire, void* epoch_dev,
                                 long long n, int rank, int world, float scale, const float* scale_dev, float* sumsq_dev,
                                 int* found_inf_dev, int max_blocks, cudaStream_t stream);
int det_p2p_broadcast(void* const* peer_bufs, void* const* peer_flags, void* epoch_dev, long long nbytes, int rank,
                      int world, int root, cudaStream_t stream);
int det_metric_allreduce(void* const* peer_bufs, void* const* peer_flags, void* epoch_dev, float* out_dev,
                         const float* weights_dev, int n, int rank, int world, int op, cudaStream_t stream);
int det_arena_alloc(long long bytes, void** out);
int det_arena_free(void* p);
int det_ipc_export(void* p, unsigned char* handle64);
int det_ipc_import(const unsigned char* handle64, void** out);
int det_ipc_close(void* p);

/* ---- convolutional stack, NHWC bf16 (reference: torch.nn.Conv2d / BatchNorm2d / pooling → cuDNN in user model code,
 *      e.g. examples/computer_vision/cifar10_pytorch/model_def.py:49-74) -------------------------
 * det_conv2d_implicit: unit-stride implicit GEMM on tcgen05; mode 0 fprop (src = x, other = weight [Cout,R,S,Cin],
 * out = y bf16), 1 dgrad (src = dy, other = weight, out = dx bf16), 2 wgrad (src = dy, other = x, out = dW fp32,
 * accumulated). Cin, Cout multiples of 8. */
int det_conv2d_implicit(int mode, const void* src, const void* other, void* out, int N, int H, int W, int Cin,
                        int Cout, int R, int S, int pad_h, int pad_w, const float* bias, int relu, int splitk,
                        int force_bn, cudaStream_t stream);
/* Forward variants that also accumulate per-output-channel sum and sum of squares of the bf16 output into
 * colstats [2*Cout] (caller clears it): BatchNorm batch statistics fused into the producing kernel's epilogue. */
int det_conv2d_fprop_stats(const void* x, const void* weight, void* y, int N, int H, int W, int Cin, int Cout, int R, int S,
                           int pad_h, int pad_w, const float* bias, float* colstats, cudaStream_t stream);
int det_gemm_bf16_stats(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                        long long ldc, int epi, const float* bias, float* colstats, cudaStream_t stream);
int det_bn_finalize(const float* sums, float* mean, float* rstd, float* running_mean, float* running_var, long long M, int C,
                    float eps, float momentum, cudaStream_t stream);
/* gather path for strided / few-channel convolutions: col [N*Ho*Wo, Kp] with column (r, s, c); Kp % 8 == 0 */
int det_im2col(const void* x, void* col, int N, int H, int W, int C, int R, int S, int stride, int pad_h, int pad_w,
               int Kp, cudaStream_t stream);
int det_col2im(const void* col, void* dx, int N, int H, int W, int C, int R, int S, int stride, int pad_h, int pad_w,
               int Kp, cudaStream_t stream);
/* BatchNorm over the rows of x [M, C]; sums_scratch is 2*C floats; act: 0 none, 1 relu, 2 leaky relu (slope) */
int det_bn_stats(const void* x, float* sums_scratch, float* mean, float* rstd, float* running_mean, float* running_var,
                 long long M, int C, float eps, float momentum, cudaStream_t stream);
int det_bn_apply(const void* x, const void* resid, void* y, const float* mean, const float* rstd, const float* gamma,
                 const float* beta, long long M, int C, int act, float slope, cudaStream_t stream);
int det_bn_bwd(const void* dy, const void* x, const void* a, const float* mean, const float* rstd, const float* gamma,
               float* sums_scratch, void* dx, void* dresid, float* dgamma, float* dbeta, long long M, int C, int act,
               float slope, cudaStream_t stream);
int det_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride, int pad,
                    cudaStream_t stream);
int det_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k, int stride, int pad,
                    cudaStream_t stream);
int det_avgpool_fwd(const void* x, void* y, int N, int HW, int C, cudaStream_t stream);
int det_avgpool_bwd(const void* dy, void* dx, int N, int HW, int C, cudaStream_t stream);
/* act: 1 relu, 2 leaky relu, 3 tanh, 4 sigmoid; backward reads the saved OUTPUT y */
int det_act_fwd(const void* x, void* y, long long n, int act, float slope, cudaStream_t stream);
int det_act_bwd(const void* dy, const void* y, void* dx, long long n, int act, float slope, cudaStream_t stream);
int det_dropout(const void* x, void* y, long long n, float p, unsigned int seed, const unsigned int* seed_dev,
                cudaStream_t stream);
int det_nchw_to_nhwc_bf16(const float* in, void* out, int N, int C, int HW, cudaStream_t stream);
/* decoded image bytes (uint8, ...HWC, C <= 8) -> bf16 with (x/255 - mean[c]) / std[c]; mean/std are HOST arrays */
int det_u8_to_bf16_norm(const void* in, void* out, long long n, int C, const float* mean_host, const float* std_host,
                        cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DET_KERNELS_H_ */
