// K12 + K13: paged KV cache writes and single-query (decode) attention over block tables.
//
// Cache layout per layer: k_cache / v_cache = [num_blocks, block_size, heads, head_dim] bf16.
// A sequence's tokens live in the blocks listed by its block table (host side: csrc/host/
// kv_block_manager.cpp); token t sits in slot block_table[t / block_size] * block_size + t % block_size.
//
// Decode attention is HBM-bound (every K and V byte of the context is read once), so the kernel is built
// for bytes in flight, not FLOPs: the context is split over CTAs (split-K) so B x H x splits covers the
// 148 SMs even at batch 1, each half-warp streams one token (16-byte loads), keeps an online-softmax
// state in registers, and a tiny second kernel merges the per-split (max, sum, acc) triples.
//
// Reference: the reference has no paged attention — it re-runs the HF model over the whole right-padded
// sequence every step and throws the cache away (reference llmctl/serve/server.py:199-204; README.md:12,67
// only claims the feature).
#include "../common/util.cuh"

namespace llmctl {

// ---------------------------------------------------------------------------------------------
// K13: scatter new K/V rows (post-RoPE) from the packed [tokens, 3*H*D] activation into the cache.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
kv_append_kernel(const __nv_bfloat16* __restrict__ qkv, long long row_stride, int k_off, int v_off,
                 __nv_bfloat16* __restrict__ k_cache, __nv_bfloat16* __restrict__ v_cache,
                 const long long* __restrict__ slot_mapping, int row_elems) {
  const int tok = blockIdx.x;
  const long long slot = slot_mapping[tok];
  if (slot < 0) return;  // padded token
  const __nv_bfloat16* ksrc = qkv + tok * row_stride + k_off;
  const __nv_bfloat16* vsrc = qkv + tok * row_stride + v_off;
  __nv_bfloat16* kdst = k_cache + slot * row_elems;
  __nv_bfloat16* vdst = v_cache + slot * row_elems;
  for (int i = threadIdx.x * 8; i < row_elems; i += blockDim.x * 8) {
    *reinterpret_cast<uint4*>(kdst + i) = *reinterpret_cast<const uint4*>(ksrc + i);
    *reinterpret_cast<uint4*>(vdst + i) = *reinterpret_cast<const uint4*>(vsrc + i);
  }
}

// ---------------------------------------------------------------------------------------------
// K12: decode attention.  grid = (heads, splits, batch), block = 128 threads.
// head_dim = 128: a half-warp (16 lanes x 8 elements) covers one K/V row.
// partial layout: [batch, heads, splits, head_dim + 2] fp32 = {acc[head_dim], max, sum}.
// ---------------------------------------------------------------------------------------------
template <int kHeadDim>
__global__ void __launch_bounds__(128)
paged_attn_decode_kernel(const __nv_bfloat16* __restrict__ q, long long q_stride,
                         const __nv_bfloat16* __restrict__ k_cache,
                         const __nv_bfloat16* __restrict__ v_cache,
                         const int* __restrict__ block_tables, int max_blocks_per_seq,
                         const int* __restrict__ context_lens, float* __restrict__ partial,
                         int heads, int block_size, int tokens_per_split, float scale_log2) {
  constexpr int kLanesPerRow = kHeadDim / 8;          // 16 for D=128, 8 for D=64
  constexpr int kRowsPerWarp = 32 / kLanesPerRow;
  // heads vary fastest in launch order: the CTAs of all heads of one (sequence, split) are co-resident and walk
  // the same token rows together, so each [heads x head_dim] cache row is fetched from DRAM as one burst
  const int head = blockIdx.x, split = blockIdx.y, b = blockIdx.z;
  const int nsplit = gridDim.y;
  const int ctx = context_lens[b];
  const int t0 = split * tokens_per_split;
  const int t1 = min(ctx, t0 + tokens_per_split);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int sub = lane / kLanesPerRow;                // which row of the warp's group
  const int li = lane % kLanesPerRow;                 // 8-element slice of the row
  const int* bt = block_tables + static_cast<long long>(b) * max_blocks_per_seq;
  const long long row_elems = static_cast<long long>(heads) * kHeadDim;

  float qf[8];
  unpack8(load8(q + b * q_stride + head * kHeadDim + li * 8), qf);
#pragma unroll
  for (int j = 0; j < 8; ++j) qf[j] *= scale_log2;

  float m = -INFINITY, l = 0.f, acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;

  const int rows_per_iter = 4 * kRowsPerWarp;
  constexpr int kUnroll = 4;     // tokens per group per iteration: 8 independent 16-byte loads in flight per lane
  // The trip count must be uniform across the warp: the shuffles below use the full mask, so a lane whose token
  // is out of range still executes them (with a masked score) instead of leaving the loop early.
  for (int tb = t0 + warp * kRowsPerWarp; tb < t1; tb += rows_per_iter * kUnroll) {
    bf16x8 kraw[kUnroll], vraw[kUnroll];
    bool valid[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int t = tb + u * rows_per_iter + sub;
      valid[u] = t < t1;
      if (valid[u]) {
        const long long slot = static_cast<long long>(bt[t / block_size]) * block_size + (t % block_size);
        kraw[u] = load8(k_cache + slot * row_elems + head * kHeadDim + li * 8);
        vraw[u] = load8(v_cache + slot * row_elems + head * kHeadDim + li * 8);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      float kf[8], vf[8];
      if (valid[u]) {
        unpack8(kraw[u], kf);
        unpack8(vraw[u], vf);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) kf[j] = vf[j] = 0.f;
      }
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) s += qf[j] * kf[j];
#pragma unroll
      for (int o = kLanesPerRow / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (valid[u]) {
        const float m_new = fmaxf(m, s);
        const float corr = exp2f(m - m_new);
        const float pr = exp2f(s - m_new);
        l = l * corr + pr;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = acc[j] * corr + pr * vf[j];
        m = m_new;
      }
    }
  }

  // merge the (4 warps x kRowsPerWarp) partial states through shared memory
  __shared__ float s_m[16], s_l[16];
  __shared__ float s_acc[16][kHeadDim];
  const int g = warp * kRowsPerWarp + sub;            // group id, 0 .. 4*kRowsPerWarp-1
  if (li == 0) {
    s_m[g] = m;
    s_l[g] = l;
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) s_acc[g][li * 8 + j] = acc[j];
  __syncthreads();
  const int ngroups = 4 * kRowsPerWarp;
  if (threadIdx.x < kHeadDim) {
    float gm = -INFINITY;
    for (int i = 0; i < ngroups; ++i) gm = fmaxf(gm, s_m[i]);
    float gl = 0.f, ga = 0.f;
    for (int i = 0; i < ngroups; ++i) {
      const float w = (s_m[i] == -INFINITY) ? 0.f : exp2f(s_m[i] - gm);
      gl += s_l[i] * w;
      ga += s_acc[i][threadIdx.x] * w;
    }
    float* out = partial + ((static_cast<long long>(b) * heads + head) * nsplit + split) * (kHeadDim + 2);
    out[threadIdx.x] = ga;
    if (threadIdx.x == 0) {
      out[kHeadDim] = gm;
      out[kHeadDim + 1] = gl;
    }
  }
}

template <int kHeadDim>
__global__ void __launch_bounds__(kHeadDim)
paged_attn_merge_kernel(const float* __restrict__ partial, __nv_bfloat16* __restrict__ out,
                        long long out_stride, int heads, int nsplit) {
  const int head = blockIdx.x, b = blockIdx.y;
  const float* p = partial + (static_cast<long long>(b) * heads + head) * nsplit * (kHeadDim + 2);
  float gm = -INFINITY;
  for (int s = 0; s < nsplit; ++s) gm = fmaxf(gm, p[s * (kHeadDim + 2) + kHeadDim]);
  float gl = 0.f, ga = 0.f;
  for (int s = 0; s < nsplit; ++s) {
    const float ms = p[s * (kHeadDim + 2) + kHeadDim];
    const float w = (ms == -INFINITY) ? 0.f : exp2f(ms - gm);
    gl += p[s * (kHeadDim + 2) + kHeadDim + 1] * w;
    ga += p[s * (kHeadDim + 2) + threadIdx.x] * w;
  }
  out[b * out_stride + head * kHeadDim + threadIdx.x] = __float2bfloat16(gl > 0.f ? ga / gl : 0.f);
}

}  // namespace llmctl

using namespace llmctl;

LLMCTL_API int llmctl_kv_append(const void* qkv, long long row_stride, int k_off, int v_off,
                                void* k_cache, void* v_cache, const long long* slot_mapping,
                                int num_tokens, int heads, int head_dim, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE((heads * head_dim) % 8 == 0 && row_stride % 8 == 0 && k_off % 8 == 0 && v_off % 8 == 0);
  if (num_tokens == 0) return 0;
  kv_append_kernel<<<num_tokens, 128, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)qkv, row_stride, k_off, v_off, (__nv_bfloat16*)k_cache,
      (__nv_bfloat16*)v_cache, slot_mapping, heads * head_dim);
  LLMCTL_RETURN_LAST();
}

// Number of fp32 scratch elements llmctl_paged_attn_decode needs for a given launch.
LLMCTL_API long long llmctl_paged_attn_scratch_elems(int batch, int heads, int head_dim, int num_splits) {
  return static_cast<long long>(batch) * heads * num_splits * (head_dim + 2);
}

// out[b, h, :] = softmax(q[b, h, :] . K[b]^T * scale) V[b] over the paged cache.
// scratch: fp32 [batch * heads * num_splits * (head_dim + 2)].  num_splits >= 1.
LLMCTL_API int llmctl_paged_attn_decode(const void* q, long long q_stride, const void* k_cache,
                                        const void* v_cache, const int* block_tables,
                                        int max_blocks_per_seq, const int* context_lens, void* out,
                                        long long out_stride, float* scratch, int batch, int heads,
                                        int head_dim, int block_size, int max_context, int num_splits,
                                        float scale, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(head_dim == 128 || head_dim == 64);
  LLMCTL_REQUIRE(num_splits >= 1 && block_size >= 1 && q_stride % 8 == 0);
  if (batch == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  int per = (max_context + num_splits - 1) / num_splits;
  per = ((per + block_size - 1) / block_size) * block_size;  // whole blocks per split
  if (per < block_size) per = block_size;
  const float scale_log2 = scale * 1.4426950408889634f;
  dim3 grid(heads, num_splits, batch);
  if (head_dim == 128) {
    paged_attn_decode_kernel<128><<<grid, 128, 0, s>>>(
        (const __nv_bfloat16*)q, q_stride, (const __nv_bfloat16*)k_cache, (const __nv_bfloat16*)v_cache,
        block_tables, max_blocks_per_seq, context_lens, scratch, heads, block_size, per, scale_log2);
    paged_attn_merge_kernel<128><<<dim3(heads, batch), 128, 0, s>>>(scratch, (__nv_bfloat16*)out,
                                                                    out_stride, heads, num_splits);
  } else {
    paged_attn_decode_kernel<64><<<grid, 128, 0, s>>>(
        (const __nv_bfloat16*)q, q_stride, (const __nv_bfloat16*)k_cache, (const __nv_bfloat16*)v_cache,
        block_tables, max_blocks_per_seq, context_lens, scratch, heads, block_size, per, scale_log2);
    paged_attn_merge_kernel<64><<<dim3(heads, batch), 64, 0, s>>>(scratch, (__nv_bfloat16*)out,
                                                                  out_stride, heads, num_splits);
  }
  LLMCTL_RETURN_LAST();
}
