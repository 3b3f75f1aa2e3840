// K9: fused AdamW over flat buffers (fp32 master weights + moments, bf16 or fp32 gradients, bf16
// parameter write-back) and the gradient sum-of-squares reduction used for global-norm clipping.
//
// Reference: torch.optim.AdamW(betas=(0.9,0.95), eps=1e-8) at reference
// llmctl/runtime/engine.py:230-235 and clip_grad_norm_ at engine.py:295-299.  Unlike the reference
// (which steps AdamW directly on fp16 weights, engine.py:132) the update runs on fp32 master weights.
//
// The gradient scale (1/world * clip coefficient) is read from device memory so the host never
// synchronises; a non-finite scale (NaN/Inf gradient norm) turns the whole step into a no-op.
#include "../common/util.cuh"

namespace llmctl {

template <typename GradT>
__device__ __forceinline__ void load_grad8(const GradT* g, float* out);
template <>
__device__ __forceinline__ void load_grad8<__nv_bfloat16>(const __nv_bfloat16* g, float* out) {
  unpack8(load8(g), out);
}
template <>
__device__ __forceinline__ void load_grad8<float>(const float* g, float* out) {
  *reinterpret_cast<float4*>(out) = *reinterpret_cast<const float4*>(g);
  *reinterpret_cast<float4*>(out + 4) = *reinterpret_cast<const float4*>(g + 4);
}

template <typename GradT>
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
             const GradT* __restrict__ grad, __nv_bfloat16* __restrict__ param_out, long long n,
             float lr, float beta1, float beta2, float eps, float wd, float bc1, float bc2,
             float host_scale, const float* __restrict__ scale_dev) {
  float scale = host_scale;
  if (scale_dev != nullptr) scale *= *scale_dev;
  if (!isfinite(scale)) return;  // overflow guard: skip the step entirely
  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const float decay = 1.f - lr * wd;
  const long long nvec = n >> 3;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long o = i << 3;
    float g[8], p[8], mm[8], vv[8];
    load_grad8<GradT>(grad + o, g);
    load_grad8<float>(master + o, p);
    load_grad8<float>(m + o, mm);
    load_grad8<float>(v + o, vv);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = g[j] * scale;
      mm[j] = beta1 * mm[j] + (1.f - beta1) * gj;
      vv[j] = beta2 * vv[j] + (1.f - beta2) * gj * gj;
      const float denom = sqrtf(vv[j]) * inv_sqrt_bc2 + eps;
      p[j] = p[j] * decay - step_size * (mm[j] / denom);
    }
    *reinterpret_cast<float4*>(master + o) = *reinterpret_cast<float4*>(p);
    *reinterpret_cast<float4*>(master + o + 4) = *reinterpret_cast<float4*>(p + 4);
    *reinterpret_cast<float4*>(m + o) = *reinterpret_cast<float4*>(mm);
    *reinterpret_cast<float4*>(m + o + 4) = *reinterpret_cast<float4*>(mm + 4);
    *reinterpret_cast<float4*>(v + o) = *reinterpret_cast<float4*>(vv);
    *reinterpret_cast<float4*>(v + o + 4) = *reinterpret_cast<float4*>(vv + 4);
    if (param_out != nullptr) store8(param_out + o, pack8(p));
  }
  // scalar tail (n not a multiple of 8)
  const long long tail0 = nvec << 3;
  for (long long i = tail0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float gj = static_cast<float>(grad[i]) * scale;
    const float mj = beta1 * m[i] + (1.f - beta1) * gj;
    const float vj = beta2 * v[i] + (1.f - beta2) * gj * gj;
    const float denom = sqrtf(vj) * inv_sqrt_bc2 + eps;
    const float pj = master[i] * decay - step_size * (mj / denom);
    master[i] = pj;
    m[i] = mj;
    v[i] = vj;
    if (param_out != nullptr) param_out[i] = __float2bfloat16(pj);
  }
}

// Stage 1 of the deterministic sum of squares: one fp32 partial per CTA.
template <typename T>
__global__ void __launch_bounds__(512)
sumsq_partial_kernel(const T* __restrict__ x, long long n, float* __restrict__ partials) {
  __shared__ float red[32];
  float acc = 0.f;
  const long long nvec = n >> 3;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float g[8];
    load_grad8<T>(x + (i << 3), g);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += g[j] * g[j];
  }
  for (long long i = (nvec << 3) + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
       i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float g = static_cast<float>(x[i]);
    acc += g * g;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}
// Stage 2: a single CTA folds the partials in a fixed order.
__global__ void __launch_bounds__(256)
sumsq_final_kernel(const float* __restrict__ partials, int nparts, float* __restrict__ out,
                   int accumulate) {
  __shared__ float red[32];
  float acc = 0.f;
  for (int i = threadIdx.x; i < nparts; i += blockDim.x) acc += partials[i];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) *out = accumulate ? *out + acc : acc;
}

}  // namespace llmctl

using namespace llmctl;

LLMCTL_API int llmctl_adamw_step(void* master, void* m, void* v, const void* grad, int grad_is_fp32,
                                 void* param_out, long long n, float lr, float beta1, float beta2,
                                 float eps, float weight_decay, float bias_corr1, float bias_corr2,
                                 float host_scale, const float* scale_dev, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(n >= 0);
  if (n == 0) return 0;
  long long blocks = ((n >> 3) + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  if (grad_is_fp32) {
    adamw_kernel<float><<<static_cast<int>(blocks), 256, 0, (cudaStream_t)stream>>>(
        (float*)master, (float*)m, (float*)v, (const float*)grad, (__nv_bfloat16*)param_out, n, lr,
        beta1, beta2, eps, weight_decay, bias_corr1, bias_corr2, host_scale, scale_dev);
  } else {
    adamw_kernel<__nv_bfloat16><<<static_cast<int>(blocks), 256, 0, (cudaStream_t)stream>>>(
        (float*)master, (float*)m, (float*)v, (const __nv_bfloat16*)grad, (__nv_bfloat16*)param_out,
        n, lr, beta1, beta2, eps, weight_decay, bias_corr1, bias_corr2, host_scale, scale_dev);
  }
  LLMCTL_RETURN_LAST();
}

// out (+)= sum(x^2).  `scratch` must hold `scratch_elems` floats (>= 1); the number of stage-1 CTAs
// is min(scratch_elems, 8 * SMs).
LLMCTL_API int llmctl_sumsq(const void* x, int is_fp32, long long n, float* scratch,
                            int scratch_elems, float* out, int accumulate, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(n >= 0 && scratch_elems >= 1);
  long long blocks = ((n >> 3) + 511) / 512;
  if (blocks < 1) blocks = 1;
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  if (blocks > scratch_elems) blocks = scratch_elems;
  if (is_fp32)
    sumsq_partial_kernel<float><<<static_cast<int>(blocks), 512, 0, (cudaStream_t)stream>>>(
        (const float*)x, n, scratch);
  else
    sumsq_partial_kernel<__nv_bfloat16><<<static_cast<int>(blocks), 512, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, n, scratch);
  sumsq_final_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(scratch, static_cast<int>(blocks), out,
                                                          accumulate);
  LLMCTL_RETURN_LAST();
}
