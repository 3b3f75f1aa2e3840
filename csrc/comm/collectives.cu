// Stand-alone collectives over symmetric memory (NVLink 5 / NVSwitch), written as plain kernels so they can be
// measured against NCCL (`llmctl bench comms`) and reused by the runtime where no GEMM is available to fuse
// with (ZeRO-3 parameter prefetch, evaluation-time gathers):
//   all-gather      : every rank PULLS the other ranks' chunks with 16-byte P2P loads (rank-staggered order)
//   reduce-scatter  : every rank pulls its slice from all ranks and sums in fp32 in rank order (deterministic)
//   all-reduce      : one-shot pull (latency-optimal for small messages) or two-shot NVLS
//                     (multimem.ld_reduce of the owned slice + multimem.st broadcast: the switch reduces)
// Synchronisation: epoch flags in the signal pads (csrc/comm/symm.cuh).  Each kernel starts with an all-rank
// "inputs are ready" exchange and ends with an "I am done reading/writing" exchange, so buffers may be
// reused immediately after the kernel by stream-ordered work on any rank.
//
// Reference: none — the reference contains no collective call at all; its "communication
// tuner" multiplies 1 ms by random noise (reference llmctl/plugins/autotuning.py:222-245).
#include "symm.cuh"
#include "../common/util.cuh"

namespace llmctl {
using namespace symm;

// pad slots: [slot, slot+8) start barrier, [slot+8, slot+16) end barrier, slot+16 local CTA counter
__device__ __forceinline__ void start_sync(const PeerTable& pads, int slot, int rank, int world, uint32_t epoch) {
  if (blockIdx.x == 0 && threadIdx.x < world)
    signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + slot + rank, epoch);
  if (threadIdx.x < world)
    signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + slot + threadIdx.x, epoch);
  __syncthreads();
}
// The last CTA to finish tells every peer "done" and waits for everyone (so nobody's buffer is reused early).
__device__ __forceinline__ void end_sync(const PeerTable& pads, int slot, int rank, int world, uint32_t epoch) {
  __shared__ unsigned int s_last;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* counter = reinterpret_cast<unsigned int*>(pads.ptr[rank]) + slot + 16;
    const unsigned int done = atomicAdd(counter, 1u) + 1u;
    s_last = (done == gridDim.x) ? 1u : 0u;
    if (s_last) atomicExch(counter, 0u);
  }
  __syncthreads();
  if (s_last) {
    if (threadIdx.x < world) {
      signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + slot + 8 + rank, epoch);
      signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + slot + 8 + threadIdx.x, epoch);
    }
  }
}

__global__ void __launch_bounds__(512)
allgather_pull_kernel(PeerTable buf, long long chunk_vec, PeerTable pads, int slot, int rank, int world,
                      uint32_t epoch) {
  start_sync(pads, slot, rank, world, epoch);
  uint4* local = reinterpret_cast<uint4*>(buf.ptr[rank]);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (int i = 1; i < world; ++i) {
    const int src = (rank + i) % world;
    const uint4* remote = reinterpret_cast<const uint4*>(buf.ptr[src]) + src * chunk_vec;
    uint4* dst = local + src * chunk_vec;
    long long v = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    for (; v + 3 * stride < chunk_vec; v += 4 * stride) {       // 4 independent 16-byte loads in flight per thread
      const uint4 a = ptx::ld_nc_v4(remote + v), b = ptx::ld_nc_v4(remote + v + stride),
                  c = ptx::ld_nc_v4(remote + v + 2 * stride), d = ptx::ld_nc_v4(remote + v + 3 * stride);
      dst[v] = a; dst[v + stride] = b; dst[v + 2 * stride] = c; dst[v + 3 * stride] = d;
    }
    for (; v < chunk_vec; v += stride) dst[v] = ptx::ld_nc_v4(remote + v);
  }
  end_sync(pads, slot, rank, world, epoch);
}

__global__ void __launch_bounds__(512)
reduce_scatter_pull_kernel(PeerTable in, long long n_vec_per_rank, __nv_bfloat16* __restrict__ out, PeerTable pads,
                           int slot, int rank, int world, uint32_t epoch) {
  start_sync(pads, slot, rank, world, epoch);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long v0 = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; v0 < n_vec_per_rank; v0 += 2 * stride) {
    const long long v1 = v0 + stride;
    const bool has1 = v1 < n_vec_per_rank;
    uint4 raw0[kMaxRanks], raw1[kMaxRanks];
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r) {
      if (r < world) {
        const uint4* base = reinterpret_cast<const uint4*>(in.ptr[r]) + rank * n_vec_per_rank;
        raw0[r] = ptx::ld_nc_v4(base + v0);
        if (has1) raw1[r] = ptx::ld_nc_v4(base + v1);
      }
    }
    float s0[8], s1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s0[j] = s1[j] = 0.f;
    // fixed summation order (rank 0 .. world-1) -> deterministic
#pragma unroll
    for (int rr = 0; rr < kMaxRanks; ++rr) {
      if (rr < world) {
        float f[8];
        unpack8(*reinterpret_cast<const bf16x8*>(&raw0[rr]), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) s0[j] += f[j];
        if (has1) {
          unpack8(*reinterpret_cast<const bf16x8*>(&raw1[rr]), f);
#pragma unroll
          for (int j = 0; j < 8; ++j) s1[j] += f[j];
        }
      }
    }
    store8(out + v0 * 8, pack8(s0));
    if (has1) store8(out + v1 * 8, pack8(s1));
  }
  end_sync(pads, slot, rank, world, epoch);
}

// one-shot: every rank reads everything from everyone (small messages)
__global__ void __launch_bounds__(512)
allreduce_oneshot_kernel(PeerTable in, long long n_vec, __nv_bfloat16* __restrict__ out, PeerTable pads, int slot,
                         int rank, int world, uint32_t epoch) {
  start_sync(pads, slot, rank, world, epoch);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long v = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; v < n_vec; v += stride) {
    float s[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = 0.f;
    for (int r = 0; r < world; ++r) {
      const uint4 raw = ptx::ld_relaxed_sys_v4(reinterpret_cast<const uint4*>(in.ptr[r]) + v);
      float f[8];
      unpack8(*reinterpret_cast<const bf16x8*>(&raw), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += f[j];
    }
    store8(out + v * 8, pack8(s));
  }
  end_sync(pads, slot, rank, world, epoch);
}

// two-shot NVLS, in place on the symmetric buffer: rank r reduces vectors [r*per, (r+1)*per) through the switch
// and broadcasts the result to every rank with one multicast store.
__global__ void __launch_bounds__(512)
allreduce_multimem_kernel(void* mc, long long n_vec, PeerTable pads, int slot, int rank, int world, uint32_t epoch) {
  start_sync(pads, slot, rank, world, epoch);
  const long long per = (n_vec + world - 1) / world;
  const long long lo = rank * per, hi = min(n_vec, lo + per);
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  uint4* p = reinterpret_cast<uint4*>(mc);
  long long v = lo + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  for (; v + 3 * stride < hi; v += 4 * stride) {
    const uint4 a = ptx::multimem_ld_reduce_bf16x8(p + v), b = ptx::multimem_ld_reduce_bf16x8(p + v + stride),
                c = ptx::multimem_ld_reduce_bf16x8(p + v + 2 * stride), d = ptx::multimem_ld_reduce_bf16x8(p + v + 3 * stride);
    ptx::multimem_st_v4(p + v, a); ptx::multimem_st_v4(p + v + stride, b);
    ptx::multimem_st_v4(p + v + 2 * stride, c); ptx::multimem_st_v4(p + v + 3 * stride, d);
  }
  for (; v < hi; v += stride) ptx::multimem_st_v4(p + v, ptx::multimem_ld_reduce_bf16x8(p + v));
  end_sync(pads, slot, rank, world, epoch);
}

static int comm_grid(long long n_vec, int max_ctas) {
  long long g = (n_vec + 511) / 512;
  if (g < 1) g = 1;
  if (g > max_ctas) g = max_ctas;
  return static_cast<int>(g);
}

}  // namespace llmctl

using namespace llmctl;
LLMCTL_SYMM_WATCHDOG_MODULE(collectives)

// buf: symmetric [world * chunk_bytes]; rank r's chunk is valid on rank r at entry, all chunks valid everywhere at exit.
LLMCTL_API int llmctl_allgather_pull(llmctl::symm::PeerTable buf, long long chunk_bytes, llmctl::symm::PeerTable pads,
                                     int slot, int rank, int world, unsigned int epoch, int max_ctas, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(chunk_bytes % 16 == 0 && world >= 1 && world <= symm::kMaxRanks);
  allgather_pull_kernel<<<comm_grid(chunk_bytes / 16, max_ctas > 0 ? max_ctas : 148), 512, 0, (cudaStream_t)stream>>>(
      buf, chunk_bytes / 16, pads, slot, rank, world, epoch);
  LLMCTL_RETURN_LAST();
}
// in: symmetric bf16 [world * n_per_rank]; out (local bf16 [n_per_rank]) = sum over ranks of slice `rank`.
LLMCTL_API int llmctl_reduce_scatter_pull(llmctl::symm::PeerTable in, long long n_per_rank, void* out,
                                          llmctl::symm::PeerTable pads, int slot, int rank, int world,
                                          unsigned int epoch, int max_ctas, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(n_per_rank % 8 == 0 && world >= 1 && world <= symm::kMaxRanks);
  reduce_scatter_pull_kernel<<<comm_grid(n_per_rank / 8, max_ctas > 0 ? max_ctas : 148), 512, 0, (cudaStream_t)stream>>>(
      in, n_per_rank / 8, (__nv_bfloat16*)out, pads, slot, rank, world, epoch);
  LLMCTL_RETURN_LAST();
}
// in: symmetric bf16 [n]; out: local bf16 [n] = sum over ranks.
LLMCTL_API int llmctl_allreduce_oneshot(llmctl::symm::PeerTable in, long long n, void* out,
                                        llmctl::symm::PeerTable pads, int slot, int rank, int world,
                                        unsigned int epoch, int max_ctas, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(n % 8 == 0 && world >= 1 && world <= symm::kMaxRanks);
  allreduce_oneshot_kernel<<<comm_grid(n / 8, max_ctas > 0 ? max_ctas : 32), 512, 0, (cudaStream_t)stream>>>(
      in, n / 8, (__nv_bfloat16*)out, pads, slot, rank, world, epoch);
  LLMCTL_RETURN_LAST();
}
// mc: NVLS multicast address of a symmetric bf16 [n] buffer; reduced in place on every rank.
LLMCTL_API int llmctl_allreduce_multimem(void* mc, long long n, llmctl::symm::PeerTable pads, int slot, int rank,
                                         int world, unsigned int epoch, int max_ctas, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(mc != nullptr && n % 8 == 0 && world >= 1 && world <= symm::kMaxRanks);
  const long long per = ((n / 8) + world - 1) / world;
  allreduce_multimem_kernel<<<comm_grid(per, max_ctas > 0 ? max_ctas : 148), 512, 0, (cudaStream_t)stream>>>(
      mc, n / 8, pads, slot, rank, world, epoch);
  LLMCTL_RETURN_LAST();
}
