// K4: rotary position embedding (rotate-half / NeoX pairing, as used by LLaMA-style models),
// applied in place to the q and k heads of a packed [tokens, (q|k|v)] activation.  `sign = -1`
// applies the inverse rotation, which is exactly the backward pass.
//
// Reference: none — the reference reaches HF rotary through `self.model(**batch)`
// (reference llmctl/runtime/engine.py:284); config keys rope.base / rope.scaling come from
// reference configs/models/llama-7b.json:11-14.  The cos/sin table is built by the caller so
// "linear"/"su" scaling only changes the table, not the kernel.
#include "../common/util.cuh"

namespace llmctl {

// One thread rotates 8 pairs: elements [d0, d0+8) of the first half with the matching second half.
__global__ void __launch_bounds__(256)
rope_kernel(__nv_bfloat16* __restrict__ x, long long row_stride, int n_tokens, int n_heads,
            int head_dim, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
            const int* __restrict__ positions, int seq_len, float sign) {
  const int half = head_dim >> 1;
  const int vec_per_head = half >> 3;
  const long long total = static_cast<long long>(n_tokens) * n_heads * vec_per_head;
  for (long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; idx < total;
       idx += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int vh = static_cast<int>(idx % vec_per_head);
    const long long th = idx / vec_per_head;
    const int head = static_cast<int>(th % n_heads);
    const int tok = static_cast<int>(th / n_heads);
    const int pos = positions ? positions[tok] : (tok % seq_len);
    __nv_bfloat16* p = x + tok * row_stride + static_cast<long long>(head) * head_dim + vh * 8;
    float a[8], b[8], c[8], s[8], oa[8], ob[8];
    unpack8(load8(p), a);
    unpack8(load8(p + half), b);
    const float* cp = cos_t + static_cast<long long>(pos) * half + vh * 8;
    const float* sp = sin_t + static_cast<long long>(pos) * half + vh * 8;
    *reinterpret_cast<float4*>(c) = *reinterpret_cast<const float4*>(cp);
    *reinterpret_cast<float4*>(c + 4) = *reinterpret_cast<const float4*>(cp + 4);
    *reinterpret_cast<float4*>(s) = *reinterpret_cast<const float4*>(sp);
    *reinterpret_cast<float4*>(s + 4) = *reinterpret_cast<const float4*>(sp + 4);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float sj = s[j] * sign;
      oa[j] = a[j] * c[j] - b[j] * sj;
      ob[j] = b[j] * c[j] + a[j] * sj;
    }
    store8(p, pack8(oa));
    store8(p + half, pack8(ob));
  }
}

}  // namespace llmctl

using namespace llmctl;

LLMCTL_API int llmctl_rope_inplace(void* x, long long row_stride, int n_tokens, int n_heads,
                                   int head_dim, const float* cos_t, const float* sin_t,
                                   const int* positions, int seq_len, float sign, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(head_dim % 16 == 0 && row_stride % 8 == 0 && seq_len > 0);
  const long long total = static_cast<long long>(n_tokens) * n_heads * (head_dim / 16);
  if (total == 0) return 0;
  long long blocks = (total + 255) / 256;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  rope_kernel<<<static_cast<int>(blocks), 256, 0, (cudaStream_t)stream>>>(
      (__nv_bfloat16*)x, row_stride, n_tokens, n_heads, head_dim, cos_t, sin_t, positions, seq_len,
      sign);
  LLMCTL_RETURN_LAST();
}
