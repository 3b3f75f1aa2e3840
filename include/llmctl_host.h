/* C ABI of libllmctl_host.so — native host-side runtime for serving (no CUDA dependency).
 *
 *   KV block manager  : paged KV-cache bookkeeping (block tables, ref-counted blocks, watermark, fork/COW)
 *   Batch scheduler   : continuous-batching admission / preemption under the reference's two budgets
 *                       (max_batch_size, max_batch_tokens)
 *
 * Reference: the reference's "KV cache" is an LRU dict of per-request blobs that is never filled
 * (reference llmctl/serve/server.py:57-87) and its scheduler is a FIFO admit loop with head-of-line
 * blocking (server.py:102-125).  Thread-safety: every handle carries its own mutex.
 */
#ifndef LLMCTL_HOST_H_
#define LLMCTL_HOST_H_

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------- KV block manager ---------------- */
void* llmctl_kv_create(int num_blocks, int block_size, float watermark);
void llmctl_kv_destroy(void* kv);
/* Make sure `seq_id` owns enough blocks for `num_tokens` tokens in total. 0 = ok, -1 = out of blocks
 * (nothing is changed in that case). */
int llmctl_kv_allocate(void* kv, long long seq_id, int num_tokens);
int llmctl_kv_free(void* kv, long long seq_id);            /* returns number of blocks released */
int llmctl_kv_num_tokens(void* kv, long long seq_id);      /* -1 if unknown */
int llmctl_kv_block_table(void* kv, long long seq_id, int* out, int max_out); /* returns #blocks */
int llmctl_kv_free_blocks(void* kv);
int llmctl_kv_total_blocks(void* kv);
int llmctl_kv_block_size(void* kv);
/* 1 if `num_tokens` more tokens can be allocated while keeping `watermark` of the pool free. */
int llmctl_kv_can_allocate(void* kv, int num_tokens);
/* Share parent's blocks with child (prefix sharing).  If the last block is partially filled it must be
 * copied before the child appends: returns 1 and sets *cow_src / *cow_dst, else 0; -1 on error. */
int llmctl_kv_fork(void* kv, long long parent, long long child, int* cow_src, int* cow_dst);

/* ---------------- batch scheduler ---------------- */
enum { LLMCTL_SCHED_DYNAMIC = 0, LLMCTL_SCHED_SIMPLE = 1, LLMCTL_SCHED_VLLM_LIKE = 2 };
enum { LLMCTL_PHASE_PREFILL = 0, LLMCTL_PHASE_DECODE = 1 };
void* llmctl_sched_create(int max_batch_size, int max_batch_tokens, int policy, void* kv /* may be NULL */);
void llmctl_sched_destroy(void* s);
/* 0 = queued, -2 = can never fit (prompt_len + max_tokens > max_batch_tokens, or more KV blocks than exist). */
int llmctl_sched_add(void* s, long long id, int prompt_len, int max_tokens);
/* Build the next step.  Fills parallel arrays; returns the number of entries (0 = nothing to run). */
int llmctl_sched_step(void* s, long long* ids, int* phases, int* num_tokens, int max_out);
/* Record that `id` produced one token in the last step (grows its KV allocation). 0 ok, -1 preempted (by this
 * call, or earlier in the step: the token is still counted so that re-admission re-prefills prompt + generated). */
int llmctl_sched_on_token(void* s, long long id);
void llmctl_sched_finish(void* s, long long id);
int llmctl_sched_num_waiting(void* s);
int llmctl_sched_num_running(void* s);
/* Sequences evicted since the last call (their KV was freed; they will be re-prefilled). */
int llmctl_sched_pop_preempted(void* s, long long* ids, int max_out);

#ifdef __cplusplus
}
#endif
#endif /* LLMCTL_HOST_H_ */
