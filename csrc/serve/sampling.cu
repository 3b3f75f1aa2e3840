// K14: batched temperature / top-k / top-p / multinomial sampling with no host synchronisation.
//
// Semantics are the reference's (reference llmctl/serve/server.py:209-235): logits / temperature (if != 1)
// -> keep the k largest (if k > 0) -> nucleus: sort descending, softmax, cumulative sum, drop tokens whose
// cumulative probability *before* them already exceeds p, so the first token crossing p is kept (if p < 1)
// -> softmax -> multinomial(1).  The reference does this in a Python loop per request with a `.item()`
// host sync each; here one CTA per row does it on the device.  temperature == 0 means greedy (argmax).
//
// Instead of sorting, thresholds are found by bisection on the order-preserving integer image of the
// float: 32 passes for the k-th largest value, 32 for the nucleus cut (each pass is a register sweep and a
// block reduction).  Ties at a threshold are all kept, like a stable sort would keep the first of them.
#include "../common/util.cuh"

namespace llmctl {

constexpr int kSampleThreads = 512;
// values per thread: 64 (V <= 32768, all in registers), 128 or 256 (larger vocabularies; partly in local memory)

__device__ __forceinline__ uint32_t float_to_ordered(float f) {
  uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ordered_to_float(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

__device__ __forceinline__ float block_sum_1024(float v, float* red) { return block_sum(v, red); }
__device__ __forceinline__ int block_sum_int(int v, int* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  int t = (lane < nwarp) ? red[lane] : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  return t;
}

template <int kMaxPerThread>
__global__ void __launch_bounds__(kSampleThreads)
sample_kernel(const __nv_bfloat16* __restrict__ logits_bf16, const float* __restrict__ logits_f32,
              long long ld, int V, const float* __restrict__ temperature, const int* __restrict__ top_k,
              const float* __restrict__ top_p, const float* __restrict__ uniform,
              long long* __restrict__ out_token, float* __restrict__ out_logprob) {
  __shared__ float redf[32];
  __shared__ int redi[32];
  __shared__ float s_prefix[kSampleThreads / 32];
  __shared__ int s_choice;
  const int row = blockIdx.x;
  const float T = temperature ? temperature[row] : 1.f;
  const int k = top_k ? top_k[row] : -1;
  const float p = top_p ? top_p[row] : 1.f;

  float x[kMaxPerThread];
  const int per = (V + kSampleThreads - 1) / kSampleThreads;   // contiguous chunk per thread (index order)
  const int base = threadIdx.x * per;
  const bool greedy = (T <= 0.f) || (k == 1);
  const float invT = greedy ? 1.f : 1.f / T;
  float mx = -INFINITY;
  int amax = 0;
#pragma unroll
  for (int i = 0; i < kMaxPerThread; ++i) {
    const int idx = base + i;
    float v = -INFINITY;
    if (i < per && idx < V) {
      v = logits_bf16 ? __bfloat162float(logits_bf16[row * ld + idx]) : logits_f32[row * ld + idx];
      v *= invT;
    }
    x[i] = v;
    if (v > mx) {
      mx = v;
      amax = idx;
    }
  }
  // block argmax (value, lowest index on ties)
  {
    float bm = block_max(mx, redf);
    int cand = (mx == bm) ? amax : 0x7fffffff;
    __syncthreads();
    // min-reduce candidate index
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) cand = min(cand, __shfl_xor_sync(0xffffffffu, cand, o));
    if (lane == 0) redi[warp] = cand;
    __syncthreads();
    int t = (lane < (kSampleThreads / 32)) ? redi[lane] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t = min(t, __shfl_xor_sync(0xffffffffu, t, o));
    mx = bm;
    amax = t;
  }
  if (greedy) {
    // log-prob of the argmax under the (untempered) softmax
    float se = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxPerThread; ++i) se += (x[i] == -INFINITY) ? 0.f : __expf(x[i] - mx);
    se = block_sum_1024(se, redf);
    if (threadIdx.x == 0) {
      out_token[row] = amax;
      if (out_logprob) out_logprob[row] = -__logf(se);
    }
    return;
  }

  // ---- top-k: bisection for the k-th largest value --------------------------------------------
  uint32_t thr_k = 0;  // keep x with ordered(x) >= thr_k ; 0 keeps everything
  if (k > 0 && k < V) {
    uint32_t lo = 0, hi = 0xffffffffu;   // invariant: count(>= lo) >= k
    for (int it = 0; it < 32; ++it) {
      const uint32_t mid = lo + ((hi - lo) >> 1) + ((hi - lo) & 1u);
      int c = 0;
#pragma unroll
      for (int i = 0; i < kMaxPerThread; ++i) c += (x[i] != -INFINITY && float_to_ordered(x[i]) >= mid) ? 1 : 0;
      c = block_sum_int(c, redi);
      if (c >= k) lo = mid; else hi = mid - 1;
      if (lo >= hi) break;
    }
    thr_k = lo;
  }
  // unnormalised probabilities of the survivors
  float* e = x;   // the logits are not needed past this point: reuse their registers
  float z = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxPerThread; ++i) {
    const bool keep = x[i] != -INFINITY && float_to_ordered(x[i]) >= thr_k;
    const float ev = keep ? __expf(x[i] - mx) : 0.f;
    e[i] = ev;
    z += ev;
  }
  z = block_sum_1024(z, redf);

  // ---- top-p: largest threshold whose strictly-greater mass is < p * z (first crossing token kept) ----
  if (p < 1.f) {
    const float target = p * z;
    // find the smallest value v such that mass(e > v) < target; keep e >= v
    uint32_t lo = 0, hi = float_to_ordered(1.0f);   // e in [0, 1]
    for (int it = 0; it < 32; ++it) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      const float fv = ordered_to_float(mid);
      float mass = 0.f;
#pragma unroll
      for (int i = 0; i < kMaxPerThread; ++i) mass += (e[i] > fv) ? e[i] : 0.f;
      mass = block_sum_1024(mass, redf);
      if (mass < target) hi = mid; else lo = mid + 1;
      if (lo >= hi) break;
    }
    const float cut = ordered_to_float(hi);
    z = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxPerThread; ++i) {
      if (e[i] < cut) e[i] = 0.f;
      z += e[i];
    }
    z = block_sum_1024(z, redf);
  }

  // ---- multinomial: inverse CDF in index order ----------------------------------------------------
  float local = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxPerThread; ++i) local += e[i];
  // exclusive prefix over threads: warp scan then scan of warp totals
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += n;
  }
  if (lane == 31) s_prefix[warp] = incl;
  if (threadIdx.x == 0) s_choice = -1;
  __syncthreads();
  if (warp == 0) {
    float w = (lane < kSampleThreads / 32) ? s_prefix[lane] : 0.f;
    float wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float n = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += n;
    }
    if (lane < kSampleThreads / 32) s_prefix[lane] = wi - w;   // exclusive prefix of warp totals
  }
  __syncthreads();
  const float before = s_prefix[warp] + incl - local;
  const float u = fminf(fmaxf(uniform[row], 0.f), 0.99999994f) * z;
  if (u >= before && u < before + local) {
    float c = before;
    int pick = -1;
#pragma unroll
    for (int i = 0; i < kMaxPerThread; ++i) {
      if (pick < 0 && e[i] > 0.f) {
        c += e[i];
        if (u < c) pick = base + i;
      }
    }
    if (pick < 0) {   // numerical edge: take this thread's last surviving token
#pragma unroll
      for (int i = 0; i < kMaxPerThread; ++i) if (e[i] > 0.f) pick = base + i;
    }
    if (pick >= 0) atomicMax(&s_choice, pick);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int c = s_choice;
    if (c < 0) c = amax;  // z rounding: fall back to the mode
    out_token[row] = c;
  }
  if (out_logprob) {
    __syncthreads();
    const int c = s_choice < 0 ? amax : s_choice;
    if (c >= base && c < base + per) {
#pragma unroll
      for (int i = 0; i < kMaxPerThread; ++i)
        if (base + i == c) out_logprob[row] = __logf(fmaxf(e[i], 1e-30f) / z);
    }
  }
}

}  // namespace llmctl

using namespace llmctl;

// logits: [rows, V] bf16 (logits_is_fp32 = 0) or fp32.  temperature/top_k/top_p: per-row arrays (NULL =
// 1.0 / disabled / 1.0).  uniform: per-row U[0,1) numbers.  out_token: int64 [rows].
LLMCTL_API int llmctl_sample(const void* logits, int logits_is_fp32, long long ld, int rows, int V,
                             const float* temperature, const int* top_k, const float* top_p,
                             const float* uniform, long long* out_token, float* out_logprob,
                             void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(V >= 1 && V <= 256 * kSampleThreads && uniform != nullptr);
  if (rows == 0) return 0;
  const __nv_bfloat16* lb = logits_is_fp32 ? nullptr : (const __nv_bfloat16*)logits;
  const float* lf = logits_is_fp32 ? (const float*)logits : nullptr;
  cudaStream_t s = (cudaStream_t)stream;
  if (V <= 64 * kSampleThreads)
    sample_kernel<64><<<rows, kSampleThreads, 0, s>>>(lb, lf, ld, V, temperature, top_k, top_p, uniform, out_token, out_logprob);
  else if (V <= 128 * kSampleThreads)
    sample_kernel<128><<<rows, kSampleThreads, 0, s>>>(lb, lf, ld, V, temperature, top_k, top_p, uniform, out_token, out_logprob);
  else
    sample_kernel<256><<<rows, kSampleThreads, 0, s>>>(lb, lf, ld, V, temperature, top_k, top_p, uniform, out_token, out_logprob);
  LLMCTL_RETURN_LAST();
}
