/* C ABI of libllmctl_b200.so — hand-written sm_100a kernels.
 *
 * Conventions
 *   - every function returns a cudaError_t value as int (0 = success, 1 = invalid argument);
 *   - all device pointers are raw addresses, bf16 unless stated; `stream` is a cudaStream_t;
 *   - nothing here allocates device memory: outputs and scratch are passed in by the caller;
 *   - launches are asynchronous on `stream`.
 *
 * The reference repository ships no native code; the "Reference:" notes name the
 * library path the reference reaches for the same work.
 */
#ifndef LLMCTL_B200_H_
#define LLMCTL_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

/* ---- K3 norms.  Reference: ATen rms_norm/layer_norm inside HF modules (engine.py:284). ---- */
/* h = x (+ residual); y = h * rsqrt(mean(h^2) + eps) * w.  h_out/residual/rstd may be NULL. */
int llmctl_rmsnorm_fwd(const void* x, const void* residual, const void* w, void* h_out, void* y,
                       float* rstd, int rows, int cols, float eps, void* stream);
/* dx = d(norm)/dh (+ dres); dw (fp32) (+)= sum_rows dy * xhat.
 * scratch: fp32 [scratch_rows * cols]; scratch_rows CTAs are used (1..rows). */
int llmctl_rmsnorm_bwd(const void* dy, const void* h, const void* w, const float* rstd,
                       const void* dres, void* dx, float* dw, float* scratch, int scratch_rows,
                       int rows, int cols, int accumulate, void* stream);
/* RMSNorm backward for callers that cannot allocate device memory (a host runtime whose arena already owns the
 * whole device budget, or a latency-bound loop that must stay away from cudaMalloc's driver lock).  Same math
 * as llmctl_rmsnorm_bwd; the dw partials live in a 2 MiB array built into the library (held on every GPU that
 * loads it).  cols <= 8192 and a multiple of 8, else cudaErrorInvalidValue + llmctl_noalloc_last_error().  The
 * first call on a device resolves the array's address (do it before any stream capture).  Calls overlapping on
 * the GPU race on the shared array.  This entry point has no counterpart in the reference. */
int llmctl_rmsnorm_bwd_noalloc(const void* dy, const void* h, const void* w, const float* rstd,
                               const void* dres, void* dx, float* dw, int rows, int cols, int accumulate,
                               void* stream);
const char* llmctl_noalloc_last_error(void);
int llmctl_layernorm_fwd(const void* x, const void* residual, const void* w, const void* b,
                         void* h_out, void* y, float* mean, float* rstd, int rows, int cols,
                         float eps, void* stream);
/* scratch: fp32 [2 * scratch_rows * cols] */
int llmctl_layernorm_bwd(const void* dy, const void* h, const void* w, const float* mean,
                         const float* rstd, const void* dres, void* dx, float* dw, float* db,
                         float* scratch, int scratch_rows, int rows, int cols, int accumulate,
                         void* stream);

/* ---- K4 RoPE (rotate-half), in place; sign=+1 forward, -1 backward.  Reference: HF rotary. ---- */
int llmctl_rope_inplace(void* x, long long row_stride, int n_tokens, int n_heads, int head_dim,
                        const float* cos_t, const float* sin_t, const int* positions, int seq_len,
                        float sign, void* stream);

/* ---- K5 SwiGLU on packed [M, 2F] gate|up.  Reference: ATen silu + mul. ---- */
int llmctl_swiglu_fwd(const void* gate_up, void* out, int M, int F, long long ld_gu, void* stream);
int llmctl_swiglu_bwd(const void* dout, const void* gate_up, void* d_gate_up, int M, int F,
                      long long ld_gu, void* stream);

/* ---- K9 fused AdamW + grad-norm pieces.  Reference: torch.optim.AdamW (engine.py:230-235),
 *      clip_grad_norm_ (engine.py:295-299). ---- */
int llmctl_adamw_step(void* master, void* m, void* v, const void* grad, int grad_is_fp32,
                      void* param_out, long long n, float lr, float beta1, float beta2, float eps,
                      float weight_decay, float bias_corr1, float bias_corr2, float host_scale,
                      const float* scale_dev, void* stream);
/* out (+)= sum(x^2); scratch: fp32 [scratch_elems] partial sums (deterministic two-stage). */
int llmctl_sumsq(const void* x, int is_fp32, long long n, float* scratch, int scratch_elems,
                 float* out, int accumulate, void* stream);

/* ---- K15 cross entropy (fused fwd+bwd, in-place gradient).  Reference: HF CrossEntropyLoss. ---- */
int llmctl_cross_entropy_fwd_bwd(void* logits, long long ld, const long long* targets,
                                 float* loss_per_row, float* lse_per_row, int rows, int V,
                                 long long ignore_index, float host_scale, const float* scale_dev,
                                 int write_grad, void* stream);
int llmctl_cross_entropy_partial(const void* logits, long long ld, const long long* targets,
                                 float* stats, int rows, int V_local, long long vocab_start,
                                 void* stream);
int llmctl_cross_entropy_finalize(void* logits, long long ld, const long long* targets,
                                  const float* lse, int rows, int V_local, long long vocab_start,
                                  long long ignore_index, float host_scale, const float* scale_dev,
                                  void* stream);

/* ---- K0 tcgen05/TMA GEMM.  Reference: cuBLAS via torch.matmul / nn.Linear. ----
 * D[M,N] = op(A) op(B) (+ D if beta).  a_mn=0: A is [M,K] row-major; a_mn=1: A stored [K,M].
 * b_mn=0: B is [N,K] row-major (Linear weight); b_mn=1: B stored [K,N].
 * out_fp32: D fp32 else bf16.  tile_n: 0 auto | 128 | 192 | 256 (single-CTA tiles) | 2 = 256x256 and 3 = 256x128 tile on a CTA pair
 * (cta_group::2); auto picks the 256x256 pair tile whenever M, N > 128.  max_ctas: 0 = all SMs. */
int llmctl_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                     long long ldb, long long ldc, int a_mn, int b_mn, int out_fp32, int beta,
                     int tile_n, int max_ctas, void* stream);

/* ---- K13 / K12 paged KV cache.  Reference: none (server.py:199-204 recomputes the whole sequence). ----
 * caches: [num_blocks, block_size, heads, head_dim] bf16; slot = block * block_size + offset. */
int llmctl_kv_append(const void* qkv, long long row_stride, int k_off, int v_off, void* k_cache,
                     void* v_cache, const long long* slot_mapping, int num_tokens, int heads,
                     int head_dim, void* stream);
long long llmctl_paged_attn_scratch_elems(int batch, int heads, int head_dim, int num_splits);
/* scratch: fp32 [batch * heads * num_splits * (head_dim + 2)] split-K partials. */
int llmctl_paged_attn_decode(const void* q, long long q_stride, const void* k_cache,
                             const void* v_cache, const int* block_tables, int max_blocks_per_seq,
                             const int* context_lens, void* out, long long out_stride, float* scratch,
                             int batch, int heads, int head_dim, int block_size, int max_context,
                             int num_splits, float scale, void* stream);

/* ---- K14 sampling.  Reference: per-request Python loop with .item() syncs (server.py:209-235). ---- */
int llmctl_sample(const void* logits, int logits_is_fp32, long long ld, int rows, int V,
                  const float* temperature, const int* top_k, const float* top_p, const float* uniform,
                  long long* out_token, float* out_logprob, void* stream);

/* ---- K0s skinny (decode) GEMM: C[Mb,N] = X[Mb,K] W[N,K]^T, 1 <= Mb <= 32.  Swap-AB (UMMA M = 128 weight rows,
 * N = 16/32 token rows), split-K over CTAs with a deterministic in-kernel fix-up.  workspace: fp32 scratch;
 * counters: >= ceil(N/128) zeroed uint32 tickets (left zero on return); one call at a time per pair.
 * Reference: cuBLAS inside the HF model (server.py:199-204). */
int llmctl_gemm_skinny_bf16(const void* X, const void* W, void* C, int Mb, int N, int K, long long ldx,
                            long long ldw, long long ldc, float* workspace, long long workspace_elems,
                            unsigned int* counters, int n_counters, int splits, void* stream);

/* ---- K6 causal flash attention, head_dim 128, bf16.  q/k/v/o: [B,S,H,128] views with token strides in
 * elements; lse: fp32 [B,H,S] (natural log).  S % 128 == 0.  Reference: SDPA/eager inside the HF model
 * (engine.py:284); baseline cuDNN SDPA.
 * Forward: 256 queries per CTA as two tiles sharing K/V, S/P/O in TMEM, P never leaves TMEM.
 * Backward: two deterministic kernels (dK/dV on transposed scores, then dQ); scratch >= 2*B*H*S floats
 * (the single-kernel atomic variant selected by LLMCTL_FLASH_BWD_V1 needs B*S*H*129). */
int llmctl_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int S, int H,
                          int D, long long q_tok_stride, long long k_tok_stride, long long v_tok_stride,
                          long long o_tok_stride, float scale, int causal, void* stream);
int llmctl_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                          const float* lse, void* dq, void* dk, void* dv, float* scratch, int B, int S, int H,
                          int D, long long q_tok_stride, long long k_tok_stride, long long v_tok_stride,
                          long long o_tok_stride, long long dqkv_tok_stride, float scale, int causal, void* stream);
/* In-kernel timeline of the forward's CTA 0 (only recorded when LLMCTL_FLASH_TRACE is set in the environment):
 * copies the 2048-word (clock64, event) log to host_out (synchronous); read with tools/trace_flash.py.  Returns a
 * cudaError_t value: 0 = ok, cudaErrorNotReady when tracing is off. */
int llmctl_flash_trace_read(long long* host_out);

/* ---- Peer-memory (NVLink 5 / NVSwitch) kernels.  No counterpart in the reference (llmctl/comms is empty).
 * llmctl_peer_table: base pointers of one symmetric allocation on every rank of the group (unused = NULL);
 * `pads` is the table of per-rank signal pads (uint32 flags); `epoch` must increase by one per call on every
 * rank for a given slot range.  A wait that exceeds the watchdog deadline gives up and is reported by
 * llmctl_comm_watchdog_poll instead of hanging the GPU. */
typedef struct { void* ptr[8]; } llmctl_peer_table;

/* stand-alone collectives on a symmetric bf16 workspace (pad slots [slot, slot + world)) */
int llmctl_allgather_pull(llmctl_peer_table buf, long long chunk_bytes, llmctl_peer_table pads, int slot, int rank,
                          int world, unsigned int epoch, int max_ctas, void* stream);
int llmctl_reduce_scatter_pull(llmctl_peer_table in, long long n_per_rank, void* out, llmctl_peer_table pads,
                               int slot, int rank, int world, unsigned int epoch, int max_ctas, void* stream);
int llmctl_allreduce_oneshot(llmctl_peer_table in, long long n, void* out, llmctl_peer_table pads, int slot,
                             int rank, int world, unsigned int epoch, int max_ctas, void* stream);
int llmctl_allreduce_multimem(void* mc, long long n, llmctl_peer_table pads, int slot, int rank, int world,
                              unsigned int epoch, int max_ctas, void* stream);   /* NVLS multimem.ld_reduce/st */

/* K1: sequence-parallel all-gather -> GEMM.  The owner pushes its rows into every rank's symmetric buffer
 * (llmctl_ag_push: NVLS multimem.st when mc_xg != NULL, else P2P stores) slab by slab and raises per-slab epoch flags
 * at pads[*][flag_base + row / rows_per_flag]; the consumer is the tcgen05 GEMM whose TMA producer waits per tile for
 * the flags of the rows it reads (llmctl_gemm_bf16_gated; flags = this rank's pad + flag_base). counters: >= 32
 * zeroed uint32 in local memory. */
/* K8: GEMM whose B operand (a ZeRO-3 parameter matrix inside a gather buffer that its shard owners are still filling
 * with llmctl_ag_push) is consumed as it arrives: before a tile's weight rows are loaded, the flags
 * b_flags[(b_off + element) / b_slab] must have reached `epoch`.  llmctl_signal_all / llmctl_signal_wait hand a gather
 * buffer back and forth between its readers and its next writers. */
int llmctl_gemm_bf16_bgated(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                            long long ldc, int a_mn, int b_mn, int out_fp32, int beta, int tile_n, int max_ctas,
                            const unsigned int* b_flags, unsigned int epoch, long long b_off, long long b_slab,
                            void* stream);
int llmctl_signal_all(llmctl_peer_table pads, int rank, int world, int slot, unsigned int epoch, void* stream);
int llmctl_ag_gemm(llmctl_peer_table xg, void* mc_xg, const void* W, void* Y, int M, int N, int K, long long ldw,
                   long long ldy, int w_mn, int rows_per_flag, llmctl_peer_table pads, int flag_base, int rank,
                   int world, unsigned int epoch, unsigned int* counters, int push_ctas, int tile_n,
                   void* push_stream, void* stream);   /* push (side stream) + gated GEMM in one call */
int llmctl_ag_push(const void* src, llmctl_peer_table xg, void* mc_xg, int rows, int cols, int rows_per_flag,
                   llmctl_peer_table pads, int flag_base, int rank, int world, unsigned int epoch,
                   unsigned int* counters, int src_in_xg, int max_ctas, void* stream);
int llmctl_gemm_bf16_gated(const void* A, const void* B, void* C, int M, int N, int K, long long lda, long long ldb,
                           long long ldc, int b_mn, int out_fp32, int beta, int tile_n, int max_ctas,
                           const unsigned int* flags, unsigned int epoch, int rows_per_flag, int rank, int world,
                           void* stream);
/* K2: row-parallel GEMM whose epilogue pushes bf16 tiles into the owner's staging slot; finished by
 * llmctl_rs_reduce_rmsnorm (sum of the world slots + residual (+ RMSNorm)) on the sequence shard. */
int llmctl_gemm_rs(const void* X, const void* W, llmctl_peer_table stage, int M, int N, int K, long long ldx,
                   long long ldw, int w_mn /* W stored [K, N] */, llmctl_peer_table pads, unsigned int* local_flags,
                   int rank, int world, unsigned int epoch, void* stream);
int llmctl_rs_reduce_rmsnorm(const void* stage_local, llmctl_peer_table pads, int rank, int world,
                             const void* residual, const void* w, void* h_out, void* y, float* rstd, int rows,
                             int cols, float eps, unsigned int epoch, void* stream);
/* K7: ZeRO gradient reduce-scatter of this rank's slice (in-switch multimem.ld_reduce when mc_grads != NULL,
 * else rank-ordered P2P pull) + squared-norm partial; then AdamW on the shard + publication of the updated
 * parameters (NVLS multimem.st when mc_params != NULL, else P2P stores; push = 0 writes local_out only).
 * done_slot >= 0 announces "finished reading this bucket" to every rank; the bucket's next writer waits for it
 * with llmctl_signal_wait (transient ZeRO-2/3 buckets). */
int llmctl_rs_sumsq(llmctl_peer_table grads, long long bucket_off, const void* mc_grads, llmctl_peer_table pads,
                    int rank, int world, long long shard_lo, long long shard_n, float* reduced, int accumulate_out,
                    float* scratch, int scratch_elems, float* sumsq_out, int accumulate_sumsq, unsigned int epoch,
                    int done_slot, unsigned int done_epoch, int max_ctas, void* stream);
int llmctl_signal_wait(llmctl_peer_table pads, int rank, int world, int slot, unsigned int epoch, void* stream);
int llmctl_adam_allgather(llmctl_peer_table params, long long bucket_off, void* mc_params, llmctl_peer_table pads,
                          llmctl_peer_table scalars, const float* total_sq_dev, int rank, int world,
                          long long shard_lo, long long shard_n, float* master, float* m, float* v,
                          const float* reduced, void* local_out, int push, float lr, float beta1, float beta2,
                          float eps, float weight_decay, float bias_corr1, float bias_corr2, float grad_scale,
                          float max_norm, float* norm_out, unsigned int epoch_begin, int begin_barrier,
                          unsigned int epoch_end, int end_barrier, int max_ctas, void* stream);
/* K11: decode all-reduce of [rows, cols] partial sums + residual add + RMSNorm in one launch. */
int llmctl_ar_residual_rmsnorm(llmctl_peer_table x, long long x_off, const void* mc_x, llmctl_peer_table pads,
                               int slot_base, int rank, int world, const void* residual, void* h_out,
                               const void* w, void* y_out, int rows, int cols, float eps, unsigned int epoch,
                               const unsigned int* step_counter /* optional: epoch += *step_counter * calls_per_step */,
                               unsigned int calls_per_step, void* stream);
int llmctl_counter_add(unsigned int* counter, unsigned int value, void* stream);
/* watchdog: out[0] timed-out waits since the last clearing poll, out[1..6] = expected epoch, value seen,
 * blockIdx.x, flag address, ns waited, module index of the first one (out must hold 8 values). */
int llmctl_comm_watchdog_poll(unsigned long long* out, char* module_name, int name_cap, int clear);
int llmctl_comm_watchdog_set_timeout_ms(unsigned long long ms);
int llmctl_comm_watchdog_modules(void);

#ifdef __cplusplus
}
#endif
#endif /* LLMCTL_B200_H_ */
