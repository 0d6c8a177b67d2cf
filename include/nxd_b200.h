/* nxd_b200 -- C ABI of the sm_100a kernel library (libnxd_b200.so).
 *
 * Conventions: every function returns 0 on success, a cudaError_t value or -1 otherwise
 * (nxd_last_error() gives the message of the last failure on the calling thread).  Pointers
 * are device pointers unless stated otherwise; `stream` is a cudaStream_t; bf16 tensors are
 * row-major with the stated shape.  No entry point allocates device memory: scratch buffers
 * are passed by the caller (their sizes come from the *_num_* helpers).
 * The "reference" citations name the file:line of the upstream NeuronxDistributed code whose
 * behaviour the kernel provides on B200.
 */
#ifndef NXD_B200_H
#define NXD_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ---------------------------------------------------------------------------- */
const char* nxd_last_error(void);
int nxd_version(void);
int nxd_device_sm_count(void);
/* Kernel launches made by this library in the calling process so far (bench.py's "gpu_launches"). */
unsigned long long nxd_kernel_launches(void);
/* Overwrite `bytes` of `buf` (> L2 size) so the next timed iteration starts with a cold L2. */
int nxd_l2_flush(void* buf, size_t bytes, void* stream);

/* ---- GEMM: D[M,N] (+)= A_op[M,K] * B_op[N,K]^T on tcgen05/TMEM/TMA -------------------------
 * reference: parallel_layers/layers.py:325 (fwd), :348 (dgrad), :409 (wgrad) -- torch.matmul.
 * layout 0 "nt": a = [M,K], b = [N,K];  1 "nn": a = [M,K], b = [K,N];  2 "tn": a = [K,M], b = [K,N].
 * lda/ldb/ldd: row pitch in elements of the tensors as stored.  out_mode 0 bf16, 1 fp32,
 * 2 fp32 accumulate (d += result).  block_n 0 = auto, else 128 or 256. */
int nxd_gemm_bf16(int layout, const void* a, const void* b, void* d, int M, int N, int K, long long lda,
                  long long ldb, long long ldd, int out_mode, int block_n, void* stream);

/* ---- RMSNorm (reference: examples/training/llama/modeling_llama_nxd.py:142-149;
 *      examples/inference/modules/custom_calls.py:5-16) ---------------------------------------- */
int nxd_rmsnorm_fwd(const void* x, const void* residual, const void* w, void* y, void* residual_out,
                    float* rstd, int rows, int H, float eps, void* stream);
int nxd_rmsnorm_bwd_num_partials(int rows);
/* scratch: nxd_rmsnorm_bwd_num_partials(rows) * H floats */
int nxd_rmsnorm_bwd(const void* dy, const void* x, const void* w, const float* rstd, const void* dresidual,
                    void* dx, void* dw_bf16, float* dw_f32, int accumulate, float* scratch, int rows, int H,
                    void* stream);

/* No counterpart in the reference.  Same computation as nxd_rmsnorm_bwd for callers that cannot supply (or allocate)
 * a scratch buffer: the dW partial sums go to a 4 MiB __device__ array inside the library.  H <= 8192 (see
 * nxd_rmsnorm_bwd_static_max_h), larger H returns an error.  No allocation happens in the call; the first call on a
 * device looks up the array's address, so issue it before any stream capture.  Calls share the array: do not let two
 * of them overlap on the GPU. */
int nxd_rmsnorm_bwd_static_max_h(void);
int nxd_rmsnorm_bwd_static(const void* dy, const void* x, const void* w, const float* rstd, const void* dresidual,
                           void* dx, void* dw_bf16, float* dw_f32, int accumulate, int rows, int H, void* stream);

/* ---- SwiGLU / RoPE (reference: modeling_llama_nxd.py:197-208, :398-399;
 *      examples/inference/modules/attention/utils.py:27-41) ------------------------------------- */
int nxd_swiglu_fwd(const void* gate_up, void* out, long long rows, int I, void* stream);
int nxd_swiglu_bwd(const void* gate_up, const void* dy, void* dgate_up, long long rows, int I, void* stream);
int nxd_rope_inplace(void* x, const float* cos_t, const float* sin_t, const int* positions, long long T,
                     int heads_total, int n_rot_heads, int D, int pos_div, int pos_mod, int inverse,
                     void* stream);

/* ---- vocab-parallel cross entropy (reference: parallel_layers/loss_functions.py:11-130) ------- */
int nxd_ce_stats(const void* logits, const long long* target, float* stats, long long T, int V, int vocab_start,
                 void* stream);
int nxd_ce_bwd(const void* logits, const long long* target, const float* lse, const float* grad_out,
               void* grad_logits, long long T, int V, int vocab_start, float smoothing, int vocab_global,
               void* stream);

/* ---- optimizer / grad-norm (reference: optimizer/zero_redundancy_optimizer.py:29-156,
 *      utils/adamw_fp32_optim_params.py:88-153, parallel_layers/grads.py:33-240) ----------------- */
int nxd_sumsq_num_partials(void);
int nxd_sumsq(const void* x, long long n, int dtype, float* scratch, float* out, int accumulate, void* stream);
int nxd_clip_coef(const float* total_sumsq, float max_norm, float* coef, float* norm_out, void* stream);
int nxd_adamw_step(float* master, float* exp_avg, float* exp_avg_sq, const void* grad, int grad_dtype,
                   void* param_out, long long n, float lr, float beta1, float beta2, float eps,
                   float weight_decay, int step, float grad_scale, const float* clip_coef, void* stream);
int nxd_scale_inplace(void* x, long long n, int dtype, float scale, const float* dev_scale, void* stream);

/* ---- sampling (reference: utils/sampling.py:27-77) -------------------------------------------- */
int nxd_argmax(const void* logits, int dtype, int B, int V, long long row_stride, long long* out, void* stream);
int nxd_topk_sample(const void* logits, int dtype, int B, int V, long long row_stride, int top_k, float top_p,
                    float temperature, const float* uniform, long long* out, void* stream);

/* ---- decode attention + KV cache (reference: examples/inference/modules/attention/
 *      attention_base.py:141-170; modules/model_base.py:388-422) -------------------------------- */
int nxd_decode_attn_num_splits(int B, int n_kv, int max_len);
/* scratch: B * n_q * n_splits * (D + 2) floats */
int nxd_decode_attn(const void* q, const void* cache_k, const void* cache_v, const int* lengths,
                    const int* block_table, long long stride_b, long long stride_h, int page_size,
                    int max_pages, void* out, float* scratch, int B, int n_q, int n_kv, int D, int n_splits,
                    float scale, void* stream);
/* Same with a cache-line table: seq_ids[b] = cache line (contiguous cache) or block-table row (paged cache) of batch
 * row b (continuous batching reads the lines in place instead of gathering them first). */
int nxd_decode_attn_lines(const void* q, const void* cache_k, const void* cache_v, const int* lengths,
                          const int* block_table, const int* seq_ids, long long stride_b, long long stride_h,
                          int page_size, int max_pages, void* out, float* scratch, int B, int n_q, int n_kv, int D,
                          int n_splits, float scale, void* stream);
int nxd_kv_cache_update(const void* new_k, const void* new_v, void* cache_k, void* cache_v, const int* positions,
                        const int* seq_ids, const int* block_table, long long stride_b, long long stride_h,
                        int page_size, int max_pages, int B, int n_kv, int T_new, int D, void* stream);

/* ---- embedding (reference: parallel_layers/layers.py:215-238) ---------------------------------- */
int nxd_embedding_fwd(const long long* ids, const void* weight, void* out, long long T, int H, int vocab_start,
                      int vocab_local, int B, int S, int seq_first_out, void* stream);
int nxd_embedding_bwd(const long long* ids, const void* dy, float* dw, long long T, int H, int vocab_start,
                      int vocab_local, int B, int S, int seq_first_in, void* stream);

/* ---- flash attention (reference: kernels/flash_attn.py nki_flash_attn_func; attention_base.py prefill) ----------
 * bf16 tensors addressed as (batch, head, row) element strides with unit stride on the head dimension:
 * strides = {q_b,q_h,q_s, k_b,k_h,k_s, v_b,v_h,v_s, o_b,o_h,o_s} (forward, 12 entries) followed in the backward
 * by the same triple for dout, dq, dk, dv (24 entries).  lse: [B, Hq, Sq] floats (natural log).  GQA: Hq % Hkv == 0.
 * *_tc = tcgen05 / TMEM / TMA kernels (forward D = 64 or 128; backward D = 128), the others are the portable
 * mma.sync kernels (D = 64 or 128). */
int nxd_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, const long long* strides,
                       int B, int Hq, int Hkv, int Sq, int Sk, int D, float scale, int causal, void* stream);
/* delta: [B, Hq, Sq] floats of scratch */
int nxd_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout,
                       const float* lse, float* delta, void* dq, void* dk, void* dv, const long long* strides, int B,
                       int Hq, int Hkv, int Sq, int Sk, int D, float scale, int causal, void* stream);
int nxd_flash_attn_fwd_tc(const void* q, const void* k, const void* v, void* o, float* lse,
                          const long long* strides, int B, int Hq, int Hkv, int Sq, int Sk, int D, float scale,
                          int causal, void* stream);
/* scratch: lse_pad, delta_pad = B * Hq * roundup(Sq, 64) floats each; dq_accum = B * Hq * D * roundup(Sq, 64) floats (d-major), ZEROED
 * by the caller before the call (dQ is accumulated with TMA reduce-add).  dkv_accum: NULL, or 2 * B * Hkv * Sk * D
 * ZEROED floats: then nxd_flash_attn_bwd_tc_num_splits() CTAs share each K/V block and reduce-add their partial
 * dK / dV (few heads per GPU under tensor parallelism would otherwise leave most SMs idle). */
int nxd_flash_attn_bwd_tc_num_splits(int B, int Hq, int Hkv, int Sq, int Sk);
int nxd_flash_attn_bwd_tc(const void* q, const void* k, const void* v, const void* o, const void* dout,
                          const float* lse, float* lse_pad, float* delta_pad, float* dq_accum, float* dkv_accum,
                          void* dq, void* dk, void* dv, const long long* strides, int B, int Hq, int Hkv, int Sq,
                          int Sk, int D, float scale, int causal, void* stream);

/* ---- symmetric (peer-mapped) memory over CUDA IPC + fused compute/collective kernels ---------------------------
 * No counterpart in the reference (its collectives are XLA ops).  One process per GPU: every rank allocates a
 * region, exchanges the 64-byte handles out of band (torch.distributed), opens the peers' handles and passes the
 * resulting pointer tables to the kernels.  A signal pad (uint32 flags, zero-initialised) is one more such region. */
int nxd_symm_handle_bytes(void);
int nxd_symm_alloc(size_t bytes, void** ptr, unsigned char* handle_out);   /* cudaMalloc + memset 0 + export */
int nxd_symm_open(const unsigned char* handle, void** ptr);                /* map a peer's region */
int nxd_symm_close(void* ptr);
int nxd_symm_free(void* ptr);
/* all ranks rendezvous on the device: flag `slot` of every pad, `epoch` strictly increasing per slot */
int nxd_symm_barrier(void* const* sig_peers, void* sig_local, int world, int rank, int slot, unsigned epoch,
                     void* stream);
/* Y[tp*R, N] = all_gather_r(X_r[R, K]) * op(W)  (reference: layers.py:312-325 all_gather then matmul).
 * layout 0: W is [N, K] (nt); 1: W is [K, N] (nn).  xt_ptrs: tp pointers to every rank's gathered buffer
 * [tp*R, K] (symmetric memory; on return the local one holds the gathered input), sig_ptrs: tp signal pads,
 * slot_chunks: first of tp * R/256 flags, epoch: call counter of this (buffer, flags) pair starting at 1,
 * comm_ctas: even number of CTAs that drive the copy engine, block_n: 0 = auto, trace: NULL or 16 x u64. */
int nxd_ag_gemm_bf16(int layout, const void* x_local, const void* w, void* y, int R, int N, int K, long long ldw,
                     long long ldy, void* const* xt_ptrs, void* const* sig_ptrs, int tp, int rank, int slot_chunks,
                     unsigned epoch, int comm_ctas, int block_n, void* trace, void* stream);
/* EXPERIMENTAL (compiled, not yet run on hardware): same, but the copy CTAs store the local shard ONCE to the NVLS
 * multicast address of the gathered buffer (multimem.st) instead of once per peer. */
int nxd_ag_gemm_bf16_nvls(int layout, const void* x_local, const void* w, void* y, int R, int N, int K, long long ldw,
                          long long ldy, void* const* xt_ptrs, void* const* sig_ptrs, int tp, int rank, int slot_chunks,
                          unsigned epoch, int comm_ctas, int block_n, void* mc_total, void* stream);
/* out[R, N] = reduce_scatter_r(X[tp*R, K] * op(W))  (reference: layers.py:452-461 matmul then reduce_scatter).
 * stage_ptrs: tp pointers to every rank's staging region [tp][R][N] bf16 (double-buffered by the caller through
 * epoch parity), slot_cnt: first of tp * R/256 flags. */
int nxd_gemm_rs_bf16(int layout, const void* x, const void* w, void* out, int R, int N, int K, long long ldx,
                     long long ldw, long long ldo, void* const* stage_ptrs, void* const* sig_ptrs, int tp, int rank,
                     int slot_cnt, unsigned epoch, int block_n, void* trace, void* stream);
/* Expert parallelism (reference: modules/moe/experts.py:121-152, enter / exit_expert_parallel_region all-to-all around
 * the expert matmuls).  Dispatch: x_local is [ep][E_loc, C, K] (slab d = tokens routed to the experts of rank d); every
 * rank receives X_total = [ep(source)][E_loc, C, K] in its symmetric buffer and y[(s, e, c), :] = X_total[s, e, c, :] * W_e.
 * Combine: x is [ep(source)][E_loc, C, K]; result tiles are stored straight into the token owner's output
 * [E, C, N] (out_ptrs: every rank's symmetric output buffer of this call), nothing is reduced.
 * W: stacked expert weights, layout 0 = [E_loc * N, K], 1 = [E_loc * K, N].  R = E_loc * C, C % 256 == 0. */
int nxd_a2a_gemm_bf16(int layout, const void* x_local, const void* w, void* y, int R, int C, int N, int K,
                      long long ldw, long long ldy, void* const* xt_ptrs, void* const* sig_ptrs, int ep, int rank,
                      int slot_chunks, unsigned epoch, int comm_ctas, int block_n, void* trace, void* stream);
int nxd_gemm_a2a_bf16(int layout, const void* x, const void* w, int R, int C, int N, int K, long long ldx,
                      long long ldw, void* const* out_ptrs, void* const* sig_ptrs, int ep, int rank, int slot_cnt,
                      unsigned epoch, int block_n, void* trace, void* stream);
/* ZeRO-1 step over peer memory (reference: the ZeRO-1 optimizer configured in trainer/trainer.py:244-269: reduce-scatter
 * of the gradient bucket, inner optimizer on the shard, all-gather of the parameters).  grad_ptrs / param_ptrs: dp
 * pointers to every rank's flat gradient / parameter buffer (symmetric memory).  reduce_shard: out[0..n) (fp32) = sum
 * over ranks of grad_r[lo..lo+n) in rank order (grad_dtype 0 = fp32, 1 = bf16).  adamw_allgather: AdamW on the shard,
 * the bf16 result is stored into param_r[lo..lo+n) of EVERY rank.  lo, n in elements, multiples of 8.  The caller
 * brackets the pair with nxd_symm_barrier. */
int nxd_zero1_reduce_shard(void* const* grad_ptrs, int grad_dtype, long long lo, long long n, float* out, int dp, int rank,
                           void* stream);
/* The same reduction with the local part of the gradient norm folded in: sum_sharded[0] += sum of out[i]^2
 * for i < split_at, sum_replicated[0] += the rest; partials = 2 * nxd_zero1_reduce_blocks(n, dtype) floats of scratch. */
int nxd_zero1_reduce_blocks(long long n, int grad_dtype);
int nxd_zero1_reduce_shard_sumsq(void* const* grad_ptrs, int grad_dtype, long long lo, long long n, float* out, int dp,
                                 int rank, long long split_at, float* partials, float* sum_sharded, float* sum_replicated,
                                 void* stream);
int nxd_zero1_adamw_allgather(float* master, float* exp_avg, float* exp_avg_sq, const float* grad,
                              void* const* param_ptrs, long long lo, long long n, float lr, float beta1, float beta2,
                              float eps, float weight_decay, int step, float grad_scale, const float* clip_coef, int dp,
                              int rank, void* stream);
/* Token-generation path (reference: RowParallelLinear matmul + all-reduce, layers.py:452-468 / mappings.py:176-190):
 * out[B, N] (+ residual) = sum over ranks of x[B, K] * w[N, K]^T in ONE launch, B <= 16; the fp32 partial blocks are
 * pushed into every peer's staging slot and added in rank order (bit-identical on all ranks).  stage_ptrs: tp regions
 * of 2 * tp * slot_elems floats; slot: first of nxd_oneshot_num_flags() flags; state: 2 zeroed uint32 of local device
 * memory per staging region (call counter, kept on the device so that CUDA-graph replays stay correct).
 * tp == 1 is a plain weight-streaming skinny GEMM (no symmetric memory needed). */
int nxd_oneshot_num_flags(void);
int nxd_skinny_gemm_ar_bf16(const void* x, const void* w, const void* residual, void* out, int B, int N, int K,
                            long long ldx, long long ldw, long long ldo, void* const* stage_ptrs,
                            void* const* sig_ptrs, size_t slot_elems, int tp, int rank, int slot, void* state,
                            void* stream);
/* Weight-streaming tcgen05 GEMM for 1..128 activation rows (batched token generation): out[B, N] = x[B, K] @ w[N, K]^T
 * (+ residual[B, N]); the weight rows are the UMMA M dimension, the padded row count (16/32/64/128) the UMMA N dimension;
 * split-K over a thread-block cluster with a distributed-shared-memory reduction.  K, ldx, ldw % 8 == 0;
 * splits: 0 = choose, else 1 / 2 / 4 / 8. */
int nxd_gemm_smallm_bf16(const void* x, const void* w, const void* residual, void* out, int B, int N, int K,
                         long long ldx, long long ldw, long long ldo, long long ldr, int splits, void* stream);
/* The same kernel with the decoder layer's element-wise neighbours in its epilogue (batched counterparts of
 * nxd_decode_qkv_rope_cache / nxd_decode_gate_up_swiglu, without the norm: x is the normalised activation).
 * qkv: B <= 128, D in {64, 128}; gate_up: B <= 64, I % 8 == 0 (a CTA streams gate slab t and up slab t). */
int nxd_smallm_qkv_rope_cache(const void* x, long long ldx, const void* w, long long ldw, const float* cos_t,
                              const float* sin_t, const int* positions, void* q_out, void* cache_k, void* cache_v,
                              const int* seq_ids, const int* block_table, long long stride_b, long long stride_h,
                              int page_size, int max_pages, int B, int K, int n_q, int n_kv, int D, int splits,
                              void* stream);
int nxd_smallm_gate_up_swiglu(const void* x, long long ldx, const void* w, long long ldw, void* out, long long ldo, int B,
                              int K, int I, int splits, void* stream);
/* Weight-only int8 skinny GEMM (reference: quantization_layers.py:444-478 dequantise + matmul): out[B, N] (+ residual) =
 * (x[B, K] bf16 * w[N, K]^T int8) * scale, fp32 accumulation; scale_n == 1 (per tensor) or N (per output channel).
 * B <= 8, K % 16 == 0, ldw (bytes) % 16 == 0.  Streams half the bytes of the bf16 skinny GEMM. */
int nxd_skinny_gemm_i8(const void* x, const void* w, const void* scale, int scale_n, const void* residual, void* out,
                       int B, int N, int K, long long ldx, long long ldw, long long ldo, void* stream);
/* out[n] = sum over ranks of in[n] (bf16, n % 8 == 0, n <= slot_elems): one-shot all-reduce for small messages */
int nxd_allreduce_oneshot(const void* in, void* out, long long n, void* const* stage_ptrs, void* const* sig_ptrs,
                          size_t slot_elems, int tp, int rank, int slot, void* state, void* stream);
/* Sequence-parallel embedding forward without the reduce-scatter of a mostly-zero tensor (reference: layers.py:215-238
 * + mappings.py reduce_scatter): the rank that owns a token's vocab row writes it into the output shard
 * [S/tp, B, H] of the rank that owns the sequence position (symmetric memory), then the ranks exchange flags.
 * ids: all B*S ids, identical on every rank; done: one zeroed uint32 of local device memory. */
int nxd_embedding_fwd_sp_push(const long long* ids, const void* weight, void* const* out_ptrs, void* const* sig_ptrs,
                              void* done, long long T, int H, int vocab_start, int vocab_local, int B, int S, int tp,
                              int rank, int slot, unsigned epoch, void* stream);
/* tooling: NVLink push-bandwidth probe (tools/bench_p2p.py) */
int nxd_p2p_push_probe(int mode, const void* src, const void* const* dst_ptrs, int n_dst, size_t bytes, int ctas,
                       int chunk, int bufs, int read_lag, int sig_mode, int sig_every, size_t flag_off, void* stream);

/* ---- MoE: grouped expert GEMM and token permutation (reference: modules/moe/moe_parallel_layers.py:78,94,108 einsum
 *      over e; modules/moe/expert_mlps.py:183-265 gather by token_assignments / weighted scatter-combine) -------------
 * E problems of one shape in one launch; stride_* = elements between consecutive problems; M needs no alignment. */
int nxd_gemm_grouped_bf16(int layout, const void* a, const void* b, void* d, int E, int M, int N, int K, long long lda,
                          long long ldb, long long ldd, long long stride_a, long long stride_b, long long stride_d,
                          int out_mode, void* stream);
/* out[r] = slot_token[r] >= 0 ? scale[r] * x[slot_token[r]] : 0   (scale NULL = 1) */
int nxd_moe_gather(const void* x, const int* slot_token, const float* scale, void* out, long long slots, int H,
                   void* stream);
/* out[t] = sum_j w[t, j] * y[slot_of[t, j]]   (pairs with w == 0 are skipped) */
int nxd_moe_combine(const void* y, const int* slot_of, const float* w, void* out, long long T, int k, int H, void* stream);
/* dw[t, j] = kept[t, j] ? <dout[t], y[slot_of[t, j]]> : 0 */
int nxd_moe_combine_dw(const void* y, const void* dout, const int* slot_of, const unsigned char* kept, float* dw,
                       long long T, int k, int H, void* stream);

/* Attention dropout (reference: nki_flash_attn_func(q, k, v, dropout_p, ..., seed), kernels/flash_attn.py:122-127,151-160).
 * Same arguments as the calls above plus dropout_p in [0, 1) and a seed; the keep mask is a hash of
 * (seed, batch, head, query, key), so the backward call regenerates it from the same two values. */
int nxd_flash_attn_fwd_tc_dropout(const void* q, const void* k, const void* v, void* o, float* lse,
                                  const long long* strides, int B, int Hq, int Hkv, int Sq, int Sk, int D, float scale,
                                  int causal, float dropout_p, unsigned long long seed, void* stream);
int nxd_flash_attn_bwd_tc_dropout(const void* q, const void* k, const void* v, const void* o, const void* dout,
                                  const float* lse, float* lse_pad, float* delta_pad, float* dq_accum, float* dkv_accum,
                                  void* dq, void* dk, void* dv, const long long* strides, int B, int Hq, int Hkv, int Sq,
                                  int Sk, int D, float scale, int causal, float dropout_p, unsigned long long seed,
                                  void* stream);

/* ---- token-generation projections with fused neighbours (reference: the token-generation graph,
 *      examples/inference/modules/model_base.py:459-518, attention_base.py:172-200) ----------------------------------
 * q, k, v = W_qkv rmsnorm(x); RoPE(q, k) at positions[b]; q -> q_out [B, n_q, D]; k, v -> the KV cache row
 * (contiguous [lines, n_kv, S, D] with strides in elements, or paged with block_table [lines, max_pages]).
 * norm_w NULL = no norm.  B <= 8, D in {64, 128}. */
int nxd_decode_qkv_rope_cache(const void* x, long long ldx, const void* norm_w, float eps, const void* w, long long ldw,
                              const float* cos_t, const float* sin_t, const int* positions, void* q_out, void* cache_k,
                              void* cache_v, const int* seq_ids, const int* block_table, long long stride_b,
                              long long stride_h, int page_size, int max_pages, int B, int K, int n_q, int n_kv, int D,
                              void* stream);
/* out[B, I] = silu(W_gate rmsnorm(x)) * (W_up rmsnorm(x)); w = [2 I, K] with the gate rows first.  B <= 8. */
int nxd_decode_gate_up_swiglu(const void* x, long long ldx, const void* norm_w, float eps, const void* w, long long ldw,
                              void* out, long long ldo, int B, int K, int I, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NXD_B200_H */
