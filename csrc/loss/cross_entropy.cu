// K15 (single-rank part): fused softmax cross-entropy forward + backward over bf16 logits.
// One pass over HBM: the row is held in registers, the per-row loss (fp32) is written out and the
// logits buffer is overwritten in place with d(loss)/d(logits) = (softmax - onehot) * grad_scale.
//
// Reference: HF `CrossEntropyLoss` on fp32-upcast logits inside `self.model(**batch)`
// (reference llmctl/runtime/engine.py:284-285).  Baseline to beat: ATen log_softmax + nll_loss
// (+ their backward), which materialises [M, V] fp32 twice.
//
// The vocab-parallel variant (logits sharded over TP ranks) splits this into "partial max/sumexp",
// an all-reduce, and "finalize"; both halves are exposed below.
#include "../common/util.cuh"

namespace llmctl {

constexpr int kCEThreads = 512;

// kVec: 16-byte groups cached in registers per thread (V <= kVec * 512 * 8).
template <int kVec>
__global__ void __launch_bounds__(kCEThreads)
ce_fwd_bwd_kernel(__nv_bfloat16* __restrict__ logits, long long ld, const long long* __restrict__ targets,
                  float* __restrict__ loss_out, float* __restrict__ lse_out, int V,
                  long long ignore_index, float host_scale, const float* __restrict__ scale_dev,
                  int write_grad) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  __nv_bfloat16* p = logits + static_cast<long long>(row) * ld;
  const int nvec = V >> 3;
  const long long tgt = targets[row];
  bf16x8 data[kVec];
  float mx = -INFINITY;
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int v = i * kCEThreads + threadIdx.x;
    if (v < nvec) {
      data[i] = load8(p + v * 8);
      float a[8];
      unpack8(data[i], a);
#pragma unroll
      for (int j = 0; j < 8; ++j) mx = fmaxf(mx, a[j]);
    }
  }
  mx = block_max(mx, red);
  float se = 0.f;
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int v = i * kCEThreads + threadIdx.x;
    if (v < nvec) {
      float a[8];
      unpack8(data[i], a);
#pragma unroll
      for (int j = 0; j < 8; ++j) se += __expf(a[j] - mx);
    }
  }
  se = block_sum(se, red);
  const float lse = mx + __logf(se);
  const bool ignored = (tgt == ignore_index) || tgt < 0 || tgt >= V;
  if (threadIdx.x == 0) {
    if (lse_out) lse_out[row] = lse;
    float tl = 0.f;
    if (!ignored) tl = lse - __bfloat162float(p[tgt]);
    loss_out[row] = tl;
  }
  if (!write_grad) return;
  float scale = host_scale;
  if (scale_dev != nullptr) scale *= *scale_dev;
  if (ignored) scale = 0.f;
  __syncthreads();  // thread 0 has read p[tgt] before anyone overwrites the row
  const float inv_se = 1.f / se;
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int v = i * kCEThreads + threadIdx.x;
    if (v < nvec) {
      float a[8], g[8];
      unpack8(data[i], a);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float sm = __expf(a[j] - mx) * inv_se;
        if (v * 8 + j == tgt) sm -= 1.f;
        g[j] = sm * scale;
      }
      store8(p + v * 8, pack8(g));
    }
  }
}

// Large-vocab fallback: three sweeps over the row (the 2nd and 3rd hit L2).
__global__ void __launch_bounds__(kCEThreads)
ce_fwd_bwd_stream_kernel(__nv_bfloat16* __restrict__ logits, long long ld,
                         const long long* __restrict__ targets, float* __restrict__ loss_out,
                         float* __restrict__ lse_out, int V, long long ignore_index,
                         float host_scale, const float* __restrict__ scale_dev, int write_grad) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  __nv_bfloat16* p = logits + static_cast<long long>(row) * ld;
  const int nvec = V >> 3;
  const long long tgt = targets[row];
  float mx = -INFINITY;
  for (int v = threadIdx.x; v < nvec; v += kCEThreads) {
    float a[8];
    unpack8(load8(p + v * 8), a);
#pragma unroll
    for (int j = 0; j < 8; ++j) mx = fmaxf(mx, a[j]);
  }
  mx = block_max(mx, red);
  float se = 0.f;
  for (int v = threadIdx.x; v < nvec; v += kCEThreads) {
    float a[8];
    unpack8(load8(p + v * 8), a);
#pragma unroll
    for (int j = 0; j < 8; ++j) se += __expf(a[j] - mx);
  }
  se = block_sum(se, red);
  const float lse = mx + __logf(se);
  const bool ignored = (tgt == ignore_index) || tgt < 0 || tgt >= V;
  if (threadIdx.x == 0) {
    if (lse_out) lse_out[row] = lse;
    loss_out[row] = ignored ? 0.f : lse - __bfloat162float(p[tgt]);
  }
  if (!write_grad) return;
  float scale = host_scale;
  if (scale_dev != nullptr) scale *= *scale_dev;
  if (ignored) scale = 0.f;
  __syncthreads();
  const float inv_se = 1.f / se;
  for (int v = threadIdx.x; v < nvec; v += kCEThreads) {
    float a[8], g[8];
    unpack8(load8(p + v * 8), a);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float sm = __expf(a[j] - mx) * inv_se;
      if (v * 8 + j == tgt) sm -= 1.f;
      g[j] = sm * scale;
    }
    store8(p + v * 8, pack8(g));
  }
}

// Vocab-parallel pieces.  Stage A: per-row local max and sum(exp(x - local_max)) plus the target
// logit if this shard owns it (else 0).  stats[row] = {max, sumexp, target_logit}.
__global__ void __launch_bounds__(kCEThreads)
ce_partial_kernel(const __nv_bfloat16* __restrict__ logits, long long ld,
                  const long long* __restrict__ targets, float* __restrict__ stats, int V_local,
                  long long vocab_start) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  const __nv_bfloat16* p = logits + static_cast<long long>(row) * ld;
  const int nvec = V_local >> 3;
  float mx = -INFINITY;
  for (int v = threadIdx.x; v < nvec; v += kCEThreads) {
    float a[8];
    unpack8(load8(p + v * 8), a);
#pragma unroll
    for (int j = 0; j < 8; ++j) mx = fmaxf(mx, a[j]);
  }
  mx = block_max(mx, red);
  float se = 0.f;
  for (int v = threadIdx.x; v < nvec; v += kCEThreads) {
    float a[8];
    unpack8(load8(p + v * 8), a);
#pragma unroll
    for (int j = 0; j < 8; ++j) se += __expf(a[j] - mx);
  }
  se = block_sum(se, red);
  if (threadIdx.x == 0) {
    const long long t = targets[row] - vocab_start;
    stats[row * 3 + 0] = mx;
    stats[row * 3 + 1] = se;
    stats[row * 3 + 2] = (t >= 0 && t < V_local) ? __bfloat162float(p[t]) : 0.f;
  }
}
// Stage B: given the global log-sum-exp per row, overwrite the local shard with the gradient.
__global__ void __launch_bounds__(kCEThreads)
ce_finalize_kernel(__nv_bfloat16* __restrict__ logits, long long ld,
                   const long long* __restrict__ targets, const float* __restrict__ lse, int V_local,
                   long long vocab_start, long long ignore_index, float host_scale,
                   const float* __restrict__ scale_dev) {
  const int row = blockIdx.x;
  __nv_bfloat16* p = logits + static_cast<long long>(row) * ld;
  const int nvec = V_local >> 3;
  const long long tg = targets[row];
  const long long t = tg - vocab_start;
  float scale = host_scale;
  if (scale_dev != nullptr) scale *= *scale_dev;
  if (tg == ignore_index || tg < 0) scale = 0.f;
  const float l = lse[row];
  for (int v = threadIdx.x; v < nvec; v += kCEThreads) {
    float a[8], g[8];
    unpack8(load8(p + v * 8), a);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float sm = __expf(a[j] - l);
      if (v * 8 + j == t) sm -= 1.f;
      g[j] = sm * scale;
    }
    store8(p + v * 8, pack8(g));
  }
}

}  // namespace llmctl

using namespace llmctl;

LLMCTL_API int llmctl_cross_entropy_fwd_bwd(void* logits, long long ld, const long long* targets,
                                            float* loss_per_row, float* lse_per_row, int rows,
                                            int V, long long ignore_index, float host_scale,
                                            const float* scale_dev, int write_grad, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(V % 8 == 0 && ld % 8 == 0 && ld >= V);
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const int nvec = V / 8;
  const int kv = (nvec + kCEThreads - 1) / kCEThreads;
#define LAUNCH(KV)                                                                             \
  ce_fwd_bwd_kernel<KV><<<rows, kCEThreads, 0, s>>>((__nv_bfloat16*)logits, ld, targets,        \
                                                    loss_per_row, lse_per_row, V, ignore_index, \
                                                    host_scale, scale_dev, write_grad)
  if (kv <= 2) LAUNCH(2);
  else if (kv <= 4) LAUNCH(4);
  else if (kv <= 8) LAUNCH(8);
  else
    ce_fwd_bwd_stream_kernel<<<rows, kCEThreads, 0, s>>>((__nv_bfloat16*)logits, ld, targets,
                                                         loss_per_row, lse_per_row, V, ignore_index,
                                                         host_scale, scale_dev, write_grad);
#undef LAUNCH
  LLMCTL_RETURN_LAST();
}

LLMCTL_API int llmctl_cross_entropy_partial(const void* logits, long long ld,
                                            const long long* targets, float* stats, int rows,
                                            int V_local, long long vocab_start, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(V_local % 8 == 0 && ld % 8 == 0);
  if (rows == 0) return 0;
  ce_partial_kernel<<<rows, kCEThreads, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)logits, ld, targets, stats, V_local, vocab_start);
  LLMCTL_RETURN_LAST();
}

LLMCTL_API int llmctl_cross_entropy_finalize(void* logits, long long ld, const long long* targets,
                                             const float* lse, int rows, int V_local,
                                             long long vocab_start, long long ignore_index,
                                             float host_scale, const float* scale_dev,
                                             void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(V_local % 8 == 0 && ld % 8 == 0);
  if (rows == 0) return 0;
  ce_finalize_kernel<<<rows, kCEThreads, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)logits, ld, targets, lse, V_local, vocab_start, ignore_index, host_scale,
      scale_dev);
  LLMCTL_RETURN_LAST();
}
