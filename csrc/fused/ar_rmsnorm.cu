// K11: tensor-parallel all-reduce fused with the residual add and the following RMSNorm, for the decode step.
//
// A row-parallel projection leaves every TP rank with a partial [rows, hidden] bf16 sum in its symmetric
// buffer.  Each rank then needs  h = residual + sum_r partial_r  and  y = rmsnorm(h) * w.  Decode messages are
// tiny (rows <= batch, <= 512 KiB) so this is latency-bound: one kernel does a one-shot all-reduce — signal
// "my partial is written", wait for all peers, pull every peer's row (16-byte P2P loads, or a single NVLS
// multimem.ld_reduce when a multicast mapping is given) — and finishes the norm in the same registers.
// 64 of these run per generated token for a 32-layer model (two per layer).
//
// Reference: NCCL all-reduce + ATen add + rms_norm (three launches and a ~10 us collective); the reference
// itself has no tensor parallelism at all (SURVEY.md §2.4 TP row; §2.6.b K11).
#include "../comm/symm.cuh"
#include "../common/util.cuh"

namespace llmctl {
using namespace symm;

// blockDim = 256, one CTA per row, kVec 16-byte groups per thread (cols <= 256 * 8 * kVec).
template <int kVec>
__global__ void __launch_bounds__(256)
ar_residual_rmsnorm_kernel(PeerTable x, long long x_off, const void* mc_x, PeerTable pads, int slot_base,
                           int rank, int world, const __nv_bfloat16* __restrict__ residual,
                           __nv_bfloat16* __restrict__ h_out, const __nv_bfloat16* __restrict__ w,
                           __nv_bfloat16* __restrict__ y_out, int cols, float eps, uint32_t epoch,
                           const uint32_t* __restrict__ step_counter, uint32_t calls_per_step) {
  __shared__ float red[32];
  // CUDA-graph replay: launch arguments are frozen, so the epoch is derived from a device-side step counter that a
  // one-thread kernel bumps at the start of every decode step (epoch = call index within the step in that case)
  if (step_counter != nullptr) epoch += *step_counter * calls_per_step;
  const int row = blockIdx.x;
  const int nvec = cols >> 3;
  // 1. publish / wait (every CTA waits; CTA 0 publishes after all local producers are done — the partial
  //    was written by the previous kernel on this stream, so it is complete when this kernel starts)
  if (blockIdx.x == 0 && threadIdx.x < world)
    signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + slot_base + rank, epoch);
  if (threadIdx.x < world)
    signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + slot_base + threadIdx.x, epoch);
  __syncthreads();

  const long long base = x_off + static_cast<long long>(row) * cols;
  float acc[kVec][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int v = i * 256 + threadIdx.x;
    if (v < nvec) {
      if (mc_x != nullptr) {
        const uint4 r = ptx::multimem_ld_reduce_bf16x8(reinterpret_cast<const __nv_bfloat16*>(mc_x) + base + v * 8);
        unpack8(*reinterpret_cast<const bf16x8*>(&r), acc[i]);
      } else {
        uint4 raw[kMaxRanks];
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < world) raw[r] = ptx::ld_relaxed_sys_v4(reinterpret_cast<const __nv_bfloat16*>(x.ptr[r]) + base + v * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r) {
          if (r < world) {
            float f[8];
            unpack8(*reinterpret_cast<const bf16x8*>(&raw[r]), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] += f[j];
          }
        }
      }
      if (residual != nullptr) {
        float f[8];
        unpack8(load8(residual + static_cast<long long>(row) * cols + v * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] += f[j];
      }
      const bf16x8 hq = pack8(acc[i]);          // the norm sees the bf16-rounded residual stream
      if (h_out != nullptr) store8(h_out + static_cast<long long>(row) * cols + v * 8, hq);
      unpack8(hq, acc[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += acc[i][j] * acc[i][j];
    }
  }
  ss = block_sum(ss, red);
  const float rstd = rsqrtf(ss / cols + eps);
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int v = i * 256 + threadIdx.x;
    if (v < nvec) {
      float g[8], o[8];
      unpack8(load8(w + v * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = acc[i][j] * rstd * g[j];
      store8(y_out + static_cast<long long>(row) * cols + v * 8, pack8(o));
    }
  }
}

__global__ void counter_add_kernel(uint32_t* p, uint32_t v) { *p += v; }

}  // namespace llmctl

using namespace llmctl;
LLMCTL_SYMM_WATCHDOG_MODULE(ar_rmsnorm)

// *counter += value on the stream (one thread): the per-step epoch source of graph-replayed peer-memory kernels.
LLMCTL_API int llmctl_counter_add(unsigned int* counter, unsigned int value, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(counter != nullptr);
  counter_add_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(counter, value);
  LLMCTL_RETURN_LAST();
}

// x: per-rank base pointers of the symmetric partial-sum buffer (bf16), x_off in elements, [rows, cols] each.
// mc_x: NVLS multicast address of the same buffer or NULL.  slot_base: first of `world` pad slots to use.
// residual / h_out may be NULL.  epoch: launch counter for this slot range (identical on all ranks).
// step_counter (optional, device uint32): the effective epoch becomes *step_counter * calls_per_step + epoch, with
// `epoch` the 1-based index of this call inside the step — the form used when the decode step is replayed as a CUDA
// graph (llmctl_counter_add bumps the counter once per step, inside the graph).
LLMCTL_API int llmctl_ar_residual_rmsnorm(llmctl::symm::PeerTable x, long long x_off, const void* mc_x,
                                          llmctl::symm::PeerTable pads, int slot_base, int rank, int world,
                                          const void* residual, void* h_out, const void* w, void* y_out,
                                          int rows, int cols, float eps, unsigned int epoch,
                                          const unsigned int* step_counter, unsigned int calls_per_step,
                                          void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && cols % 8 == 0 && cols <= 8192 && x_off % 8 == 0);
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const int kv = ((cols >> 3) + 255) / 256;
#define LAUNCH(KV)                                                                                        \
  ar_residual_rmsnorm_kernel<KV><<<rows, 256, 0, s>>>(x, x_off, mc_x, pads, slot_base, rank, world,        \
                                                      (const __nv_bfloat16*)residual, (__nv_bfloat16*)h_out, \
                                                      (const __nv_bfloat16*)w, (__nv_bfloat16*)y_out, cols, eps, epoch, \
                                                      step_counter, calls_per_step)
  if (kv <= 1) LAUNCH(1);
  else if (kv <= 2) LAUNCH(2);
  else LAUNCH(4);
#undef LAUNCH
  LLMCTL_RETURN_LAST();
}
