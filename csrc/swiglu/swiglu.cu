// K5: SwiGLU forward / backward on a packed [M, 2F] gate|up activation (gate = cols [0,F),
// up = cols [F,2F)).  out = silu(gate) * up.
//
// Reference: none — HF MLP runs ATen silu + mul as separate kernels behind
// reference llmctl/runtime/engine.py:284; `activation = "silu"` comes from
// reference configs/models/llama-7b.json.  Baseline to beat: ATen silu + mul (3 passes over HBM).
#include "../common/util.cuh"

namespace llmctl {

__global__ void __launch_bounds__(256)
swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ out, int M,
                  int F, long long ld_gu) {
  const int vec_per_row = F >> 3;
  const long long total = static_cast<long long>(M) * vec_per_row;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(idx % vec_per_row);
    const long long r = idx / vec_per_row;
    float g[8], u[8], o[8];
    unpack8(load8(gu + r * ld_gu + v * 8), g);
    unpack8(load8(gu + r * ld_gu + F + v * 8), u);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sig = 1.f / (1.f + __expf(-g[j]));
      o[j] = g[j] * sig * u[j];
    }
    store8(out + r * F + v * 8, pack8(o));
  }
}

__global__ void __launch_bounds__(256)
swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ dout, const __nv_bfloat16* __restrict__ gu,
                  __nv_bfloat16* __restrict__ dgu, int M, int F, long long ld_gu) {
  const int vec_per_row = F >> 3;
  const long long total = static_cast<long long>(M) * vec_per_row;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(idx % vec_per_row);
    const long long r = idx / vec_per_row;
    float g[8], u[8], d[8], dg[8], du[8];
    unpack8(load8(gu + r * ld_gu + v * 8), g);
    unpack8(load8(gu + r * ld_gu + F + v * 8), u);
    unpack8(load8(dout + r * F + v * 8), d);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sig = 1.f / (1.f + __expf(-g[j]));
      const float silu = g[j] * sig;
      du[j] = d[j] * silu;
      dg[j] = d[j] * u[j] * (sig + silu * (1.f - sig));
    }
    store8(dgu + r * ld_gu + v * 8, pack8(dg));
    store8(dgu + r * ld_gu + F + v * 8, pack8(du));
  }
}

}  // namespace llmctl

using namespace llmctl;

LLMCTL_API int llmctl_swiglu_fwd(const void* gate_up, void* out, int M, int F, long long ld_gu,
                                 void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(F % 8 == 0 && ld_gu % 8 == 0 && ld_gu >= 2LL * F);
  const long long total = static_cast<long long>(M) * (F / 8);
  if (total == 0) return 0;
  long long blocks = (total + 255) / 256;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  swiglu_fwd_kernel<<<static_cast<int>(blocks), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)gate_up, (__nv_bfloat16*)out, M, F, ld_gu);
  LLMCTL_RETURN_LAST();
}

LLMCTL_API int llmctl_swiglu_bwd(const void* dout, const void* gate_up, void* d_gate_up, int M,
                                 int F, long long ld_gu, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(F % 8 == 0 && ld_gu % 8 == 0 && ld_gu >= 2LL * F);
  const long long total = static_cast<long long>(M) * (F / 8);
  if (total == 0) return 0;
  long long blocks = (total + 255) / 256;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  swiglu_bwd_kernel<<<static_cast<int>(blocks), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dout, (const __nv_bfloat16*)gate_up, (__nv_bfloat16*)d_gate_up, M, F,
      ld_gu);
  LLMCTL_RETURN_LAST();
}
