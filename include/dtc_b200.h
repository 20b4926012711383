/* C ABI of libdtc_b200.so — the hand-written sm_100a (B200) kernel library of this framework.
 *
 * Conventions
 *  - every function returns 0 on success, a positive cudaError_t on a CUDA failure, a negative value on an
 *    argument error; dtc_last_error() returns a human-readable message for the calling thread.
 *  - all pointers are device pointers unless stated otherwise; "bf16" buffers are passed as void*.
 *  - kernels are enqueued on `stream` and never synchronise, allocate or free device memory, so every
 *    call can be captured into a CUDA graph.
 *  - citations "ref: path:line" point at the reference implementation (KT19/distributed-training-compare-jax)
 *    whose behaviour the op reproduces; SURVEY.md §2.5 ids (K*, X*) name the op site.
 */
#ifndef DTC_B200_H_
#define DTC_B200_H_

#include <cuda_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime --------------------------------------------------------------------------------------- */
const char* dtc_last_error(void);
int dtc_version(void);
int dtc_set_pdl(int enabled); /* programmatic dependent launch for all kernels: 0 (default) / 1; env DTC_PDL=1 */
int dtc_device_arch(void); /* compute capability major*10+minor of the current device, <0 on error */
int dtc_cast_f32_to_bf16(const float* in, void* out, long long n, cudaStream_t stream);
int dtc_cast_f32_to_bf16_2d_clear(float* in, long long ld_in, void* out, long long ld_out, int rows, int cols,
                                  cudaStream_t stream); /* same, and `in` is zeroed behind the read */
int dtc_cast_f32_to_bf16_2d(const float* in, long long ld_in, void* out, long long ld_out, int rows, int cols,
                            cudaStream_t stream);

/* ---- tcgen05 GEMM (K5 K7 K8 K9 K10 K11 K16; ref: model/CausalSelfAttention.py:18-20,46, model/MLP.py:13-24,
 *      model/GPTModel.py:72) --------------------------------------------------------------------------
 * D[M,N] = A[M,K] * B[N,K]^T, bf16 operands, fp32 accumulation in tensor memory.
 *   a_mn_major == 0: A is [M,K] row-major (lda);  == 1: A is [K,M] row-major.
 *   b_mn_major == 0: B is [N,K] row-major (ldb);  == 1: B is [K,N] row-major.
 * epi_mode: 0 out_bf16 = alpha*acc + bias          1 out = acc+bias, out2 = gelu_tanh(out)
 *           2 out_bf16 = acc + bias + aux          3 out_bf16 = acc * gelu_tanh'(aux)
 *           4 out_f32 += alpha*acc (split-K ok)    5 out_f32 = alpha*acc + bias
 * bias: bf16 [N] or NULL.  splits/stages/block_n: 0 = library default.  N % 8 == 0, ld* % 8 == 0. */
int dtc_gemm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major, int M,
                  int N, int K, int epi_mode, float alpha, const void* bias, const void* aux, long long ld_aux,
                  void* out, long long ldo, void* out2, long long ldo2, int splits, int stages, int block_n,
                  cudaStream_t stream);
/* same, kernel variant chosen per call instead of through dtc_gemm_set_impl (0 default policy, 1 one tile per CTA,
 * 2 / 3 CTA pairs 256x128 / 256x256, 4 persistent 1-CTA) */
int dtc_gemm_bf16_ex(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major, int M,
                  int N, int K, int epi_mode, float alpha, const void* bias, const void* aux, long long ld_aux,
                  void* out, long long ldo, void* out2, long long ldo2, int splits, int stages, int block_n, int impl,
                     cudaStream_t stream);

int dtc_gemm_set_impl(int impl); /* process-wide default variant of dtc_gemm_bf16 (same codes as `impl` above; tools) */

/* ---- causal flash attention (K3 K6; ref: model/CausalSelfAttention.py:34-42, model/GPTModel.py:50-51) ---- */
int dtc_flash_attn_fwd(const void* q, const void* k, const void* v, long long ld, void* out, long long ldo,
                       float* lse2, int B, int H, int T, int head_dim, cudaStream_t stream);
/* dq_accum (fp32) must be zero on entry; delta [B,H,T] fp32 is scratch. */
int dtc_flash_attn_bwd(const void* q, const void* k, const void* v, long long ld, const void* o, long long ldo,
                       const void* d_o, long long lddo, const float* lse2, float* delta, float* dq_accum,
                       long long lddq, void* dk, void* dv, long long lddkv, int B, int H, int T, int head_dim,
                       cudaStream_t stream);

/* No counterpart in the reference.  Allocation-free variant of dtc_flash_attn_bwd for callers that may not
 * allocate device memory (plugins inside a runtime that owns all allocations, loops that cannot take the
 * device-wide sync of cudaMalloc/cudaFree, jobs with a device-memory budget fixed up front):
 * the [B,H,T] delta scratch lives in a 1 MiB static device array inside the library (held by every GPU that loads
 * it, until exit).  Limit B*H*T <= 262144, larger sizes are rejected.  The first call on a device resolves the
 * array address — make it before any stream capture.  Calls share the array: overlapping calls on one GPU race. */
int dtc_flash_attn_bwd_static(const void* q, const void* k, const void* v, long long ld, const void* o, long long ldo,
                              const void* d_o, long long lddo, const float* lse2, float* dq_accum, long long lddq,
                              void* dk, void* dv, long long lddkv, int B, int H, int T, int head_dim,
                              cudaStream_t stream);

/* ---- LayerNorm (K4; ref: model/TransformerBlock.py:16,22, model/GPTModel.py:71) -------------------- */
int dtc_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd, int M,
                      int D, float eps, cudaStream_t stream);
/* dx = (dres ? dres : 0) + LN'(dy); dgamma/dbeta fp32 [D] are accumulated into. */
int dtc_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd, const void* gamma,
                      const void* dres, void* dx, float* dgamma, float* dbeta, int M, int D, cudaStream_t stream);

/* ---- embedding + positional + dropout (K1 K2; ref: model/GPTModel.py:30-38) ------------------------ */
int dtc_embed_fwd(const int* ids, const void* wte, const void* wpe, void* out, int M, int T, int D, int V,
                  float dropout_p, unsigned long long seed, const int* step_ptr, unsigned int stream_id,
                  long long row_offset, cudaStream_t stream);
int dtc_embed_bwd(const int* ids, const void* dh, float* dwte, float* dwpe, int M, int T, int D, int V,
                  float dropout_p, unsigned long long seed, const int* step_ptr, unsigned int stream_id,
                  long long row_offset, cudaStream_t stream);

/* ---- softmax cross-entropy (K12 X7; ref: train/create_train_step.py:32-34,148) --------------------- */
int dtc_ce_fwd_bwd(void* logits, long long ld, const int* labels, float* loss_rows, int M, int V, float gscale,
                   cudaStream_t stream);
/* loss + in-place gradient from the {max, sum exp} parts written by the lm_head GEMM (dtc_gemm_bf16_ex, epi_mode 8:
 * out2 = float2 parts [M, 2*ceil(N/block_n)], ldo2 = number of valid columns, block_n 128 or 256 given explicitly) —
 * one pass over the logits instead of two (K11 + K12 fused at the statistics level; ref: train/create_train_step.py:31-34). */
int dtc_ce_bwd_from_parts(void* logits, long long ld, const int* labels, const void* parts, int n_parts, float* loss_rows,
                          int M, int V, float gscale, cudaStream_t stream);
int dtc_ce_stats(const void* logits, long long ld, const int* labels, float* stats, int M, int V_local,
                 int vocab_start, cudaStream_t stream);
int dtc_ce_combine(const float* stats, int n_shards, int M, float* lse, float* loss_rows, cudaStream_t stream);
int dtc_ce_grad(void* logits, long long ld, const int* labels, const float* lse, int M, int V_local, int vocab_start,
                float gscale, cudaStream_t stream);
int dtc_reduce_sum(const float* in, int n, float scale, float* out, int accumulate, cudaStream_t stream);

/* ---- optimizer (K13 K14; ref: train/create_optimizer.py:9-12) -------------------------------------- */
int dtc_sumsq_num_partials(long long n);
int dtc_sumsq_partials(const float* g, long long n, float* partials, cudaStream_t stream);
/* scalars[0] = ||g||*grad_scale, scalars[1] = clip factor; *step += 1 when step != NULL */
int dtc_clip_finalize(const float* partials, int n_partials, float grad_scale, float max_norm, float* scalars,
                      int* step, cudaStream_t stream);
int dtc_adamw(float* p, float* m, float* v, float* g, void* p_bf16, long long n, const float* scalars,
              const int* step_ptr, float lr, float b1, float b2, float eps, float wd, float grad_scale, int zero_grad,
              cudaStream_t stream);
int dtc_colsum(const void* x, long long ld, int M, int N, float* out, cudaStream_t stream);

/* ---- symmetric-memory data-parallel step (R10 R11 R13, X1 X2; ref: the gradient all-reduce GSPMD inserts for
 *      parallel/sharding.py:25-27 + train/create_train_step.py:34-40, then train/create_optimizer.py:9-12) ----
 * *_ptrs: HOST arrays of `world` device addresses (every rank's copy of a symmetric buffer); mc_*: NVSwitch
 * multicast address of the same buffer (0 = none: peer loads / stores are used instead). */
int dtc_symm_barrier(const unsigned long long* flag_ptrs, int rank, int world, int slot, cudaStream_t stream);
/* Protocol stress aid (the reference has no race tooling): stalls the stream for
 * hash(seed, *step) % max_ns nanoseconds, max_ns <= 1e6; `step` = device int that changes between graph replays (or NULL). */
int dtc_jitter(unsigned seed, unsigned max_ns, const int* step, cudaStream_t stream);
/* small all-gather / all-reduce through symmetric memory (vocab-parallel CE statistics X7, partial LayerNorm / row-bias
 * gradients of the tensor-parallel step; ref: the max / sum exchanges GSPMD derives for train/create_train_step.py:32
 * over parallel/sharding.py:33-39): push a local fp32 range into slot `dst_off` of EVERY rank's buffer, then (after a
 * dtc_symm_barrier) sum the `world` slots locally. */
int dtc_symm_push_f32(const float* src, unsigned long long mc_dst, const unsigned long long* dst_ptrs, int world,
                      long long dst_off, long long n, cudaStream_t stream);
int dtc_sum_slots_f32(const float* slots, long long stride, int world, float* out, long long n, int accumulate,
                      cudaStream_t stream);
int dtc_dp_num_partials(long long n);
int dtc_dp_reduce_scatter(float* local_grad, unsigned long long mc_grad, const unsigned long long* grad_ptrs, int rank,
                          int world, long long off, long long n, float* partials, cudaStream_t stream);
int dtc_dp_reduce_scatter_ranges(float* local_grad, unsigned long long mc_grad, const unsigned long long* grad_ptrs,
                                 int rank, int world, const long long* offs, const long long* lens, int count,
                                 int blocks, float* partials, cudaStream_t stream);
int dtc_dp_norm_exchange(const float* partials, int n_partials, const unsigned long long* norm_slot_ptrs,
                         const unsigned long long* flag_ptrs, int rank, int world, int slot, float max_norm,
                         float* scalars, int* step, cudaStream_t stream);
int dtc_dp_adamw_allgather(float* p, float* m, float* v, float* g, unsigned long long mc_compute,
                           const unsigned long long* compute_ptrs, int world, long long off, long long n,
                           long long total_n, const float* scalars, const int* step_ptr, float lr, float b1, float b2,
                           float eps, float wd, cudaStream_t stream);
int dtc_dp_adamw_allgather_ranges(float* p, float* m, float* v, float* g, unsigned long long mc_compute,
                                  const unsigned long long* compute_ptrs, int world, const long long* offs,
                                  const long long* prefix, int n_ranges, int blocks, const float* scalars,
                                  const int* step_ptr, float lr, float b1, float b2, float eps, float wd,
                                  cudaStream_t stream);

/* ---- sequence-parallel tensor parallelism (rows R3 / R4 of SURVEY.md section 8; the TP all-reduces X3 / X4 / X5 / X8 the
 *      reference gets from GSPMD, model/CausalSelfAttention.py:49-50, model/MLP.py:23-24, are split into a reduce-scatter
 *      that is the GEMM epilogue and an all-gather that is the LayerNorm kernel's store) ----
 * dtc_gemm_bf16_scatter: D[M,N] = A[M,K] · Bᵀ; the row block owned by rank o (M/n_ranks rows each) is stored into slot
 *   `rank` of peer_out[o], a [M, ldo] bf16 buffer viewed as [n_ranks, M/n_ranks, ldo] (P2P stores from the epilogue).
 * dtc_sp_ln_fwd: owned rows x = res + bias + sum_p part[p]  -> hsum (local); LayerNorm(x) -> rows
 *   [out_row0, out_row0+Ms) of the destination on EVERY rank (mc_out multicast address, else the n_out peer_out bases).
 * dtc_sp_ln_bwd: dy = sum_p part[p]; dx = LayerNorm'(dy) + dres -> pushed like the forward; dgamma/dbeta += owned rows. */
int dtc_gemm_bf16_scatter(const void* A, long long lda, const void* B, long long ldb, int b_mn_major, int M, int N, int K,
                          const unsigned long long* peer_out, int n_ranks, int rank, long long ldo, int stages,
                          int block_n, cudaStream_t stream);
int dtc_sp_ln_fwd(const void* res, const void* bias, const void* part, int n_part, const void* gamma, const void* beta,
                  void* hsum, float* mean, float* rstd, unsigned long long mc_out, const unsigned long long* peer_out,
                  int n_out, long long out_row0, int Ms, int D, float eps, cudaStream_t stream);
int dtc_sp_ln_bwd(const void* part, int n_part, const void* x, const float* mean, const float* rstd, const void* gamma,
                  const void* dres, float* dgamma, float* dbeta, unsigned long long mc_out,
                  const unsigned long long* peer_out, int n_out, long long out_row0, int Ms, int D, cudaStream_t stream);

/* ---- grouped parameter-gradient launches (experimental, `DTC_WGRAD_GROUPED=1`): the four weight gradients of a layer
 *      (K16 of SURVEY.md section 2.5) as ONE persistent tcgen05 launch without split-K, and its four bias gradients as
 *      one column-sum launch.  All arrays are HOST arrays of `count` (<= 4) entries. ----
 * out[p][M_p, N_p] (fp32) += alpha * A[p]^T B[p],  A[p] = [K, M_p], B[p] = [K, N_p] bf16 row-major. */
int dtc_gemm_wgrad_grouped(int count, const unsigned long long* A, const long long* lda, const unsigned long long* B,
                           const long long* ldb, const int* Ms, const int* Ns, int K, const unsigned long long* out,
                           const long long* ldo, float alpha, int splits, int stages, cudaStream_t stream);
/* same contract on CTA pairs (cta_group::2), 256 x 256 output tiles; `splits` slices the token reduction */
int dtc_gemm_wgrad_grouped_pair(int count, const unsigned long long* A, const long long* lda, const unsigned long long* B,
                           const long long* ldb, const int* Ms, const int* Ns, int K, const unsigned long long* out,
                           const long long* ldo, float alpha, int splits, int stages, cudaStream_t stream);
int dtc_colsum_grouped(int count, const unsigned long long* x, const long long* ld, const int* M, const int* N,
                       const unsigned long long* out, cudaStream_t stream);
/* folded-barrier variants (experimental, `DTC_SP_FOLD=1`): same operations with the cross-rank synchronisation inside the
 * kernels.  sig_ptrs: HOST array of `world` device addresses of every rank's int[world] flag array (symmetric memory, zeroed);
 * counters: LOCAL zeroed int[4].  wait_in: wait in the prologue for the signal that publishes the partial-sum slots;
 * signal_out: the last CTA publishes this kernel's remote stores; dtc_sp_wait: one-CTA wait in front of a stock GEMM. */
int dtc_gemm_bf16_scatter_signal(const void* A, long long lda, const void* B, long long ldb, int b_mn_major, int M, int N,
                                 int K, const unsigned long long* peer_out, int n_ranks, int rank, long long ldo,
                                 int stages, int block_n, const unsigned long long* sig_ptrs, int* counters,
                                 cudaStream_t stream);
int dtc_sp_ln_fwd_sync(const void* res, const void* bias, const void* part, int n_part, const void* gamma, const void* beta,
                       void* hsum, float* mean, float* rstd, unsigned long long mc_out, const unsigned long long* peer_out,
                       int n_out, long long out_row0, int Ms, int D, float eps, const unsigned long long* sig_ptrs,
                       int* counters, int rank, int world, int wait_in, int signal_out, cudaStream_t stream);
int dtc_sp_ln_bwd_sync(const void* part, int n_part, const void* x, const float* mean, const float* rstd, const void* gamma,
                       const void* dres, float* dgamma, float* dbeta, unsigned long long mc_out,
                       const unsigned long long* peer_out, int n_out, long long out_row0, int Ms, int D,
                       const unsigned long long* sig_ptrs, int* counters, int rank, int world, int wait_in, int signal_out,
                       cudaStream_t stream);
int dtc_sp_wait(const unsigned long long* sig_ptrs, int* counters, int rank, int world, cudaStream_t stream);
int dtc_gemm_set_dual(int bits);   /* tuning: bit0 two-CTAs-per-SM GEMM, bit1 64-wide tiles for small grids, bit2 coop stores */

#ifdef __cplusplus
}
#endif
#endif /* DTC_B200_H_ */
