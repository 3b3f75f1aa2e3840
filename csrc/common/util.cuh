// Small device helpers shared by the memory-bound kernels: 16-byte vector access, bf16 <-> fp32
// packing, warp/block reductions, and the host-side launch-check macro used by every C-ABI entry.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define LLMCTL_API extern "C" __attribute__((visibility("default")))

// Clear any stale (non-sticky) error left behind by another library on this thread, so that the status
// we return describes *our* launch only.
#define LLMCTL_BEGIN() (void)cudaGetLastError()

// Every entry point returns a cudaError_t cast to int (0 == success).
// Failures are described on stderr when LLMCTL_DEBUG is set in the environment.
static inline bool llmctl_debug() {
  static int v = -1;
  if (v < 0) v = getenv("LLMCTL_DEBUG") ? 1 : 0;
  return v == 1;
}
#define LLMCTL_FAIL(code, what)                                                          \
  do {                                                                                   \
    if (llmctl_debug())                                                                  \
      fprintf(stderr, "[llmctl] %s:%d: %s (%d: %s)\n", __FILE__, __LINE__, what, (int)(code), \
              cudaGetErrorString((cudaError_t)(code)));                                  \
    return static_cast<int>(code);                                                       \
  } while (0)

#define LLMCTL_RETURN_LAST()                              \
  do {                                                    \
    cudaError_t e__ = cudaGetLastError();                 \
    if (e__ != cudaSuccess) LLMCTL_FAIL(e__, "launch");   \
    return 0;                                             \
  } while (0)

#define LLMCTL_CHECK(expr)                               \
  do {                                                   \
    cudaError_t e__ = (expr);                            \
    if (e__ != cudaSuccess) LLMCTL_FAIL(e__, #expr);     \
  } while (0)

// Argument errors are reported as cudaErrorInvalidValue (1) so python raises uniformly.
#define LLMCTL_REQUIRE(cond)                                          \
  do {                                                                \
    if (!(cond)) LLMCTL_FAIL(cudaErrorInvalidValue, "require " #cond); \
  } while (0)

namespace llmctl {

constexpr int kNumSMs = 148;

struct __align__(16) bf16x8 {
  __nv_bfloat162 v[4];
};

__device__ __forceinline__ bf16x8 load8(const __nv_bfloat16* p) {
  bf16x8 r;
  *reinterpret_cast<uint4*>(&r) = *reinterpret_cast<const uint4*>(p);
  return r;
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const bf16x8& v) {
  *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(&v);
}
__device__ __forceinline__ void unpack8(const bf16x8& in, float* f) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 t = __bfloat1622float2(in.v[i]);
    f[2 * i] = t.x;
    f[2 * i + 1] = t.y;
  }
}
__device__ __forceinline__ bf16x8 pack8(const float* f) {
  bf16x8 r;
#pragma unroll
  for (int i = 0; i < 4; ++i) r.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  return r;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// Block-wide sum; `red` must hold >= 32 floats of shared memory.  All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` against the previous use
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (lane < nwarp) ? red[lane] : 0.f;
  return warp_sum(t);
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float t = (lane < nwarp) ? red[lane] : -INFINITY;
  return warp_max(t);
}

}  // namespace llmctl
