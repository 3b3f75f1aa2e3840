// Device-side primitives for kernels that talk to peer GPUs through symmetric (peer-mapped) memory over
// NVLink 5 / NVSwitch.  A "symmetric buffer" is the same allocation made on every rank and mapped into every
// rank's address space; kernels receive the table of per-rank base pointers (`peers[r]`) plus a table of
// per-rank signal pads (arrays of 32-bit flags).
//
// Signalling protocol (epoch numbered, so flags are never reset):
//   producer:  all data stores  ->  __threadfence_system()  ->  st.release.sys flag = epoch
//   consumer:  spin ld.acquire.sys flag >= epoch  ->  (fence.proxy.async if the data is then read by TMA)
// The host passes a fresh, monotonically increasing `epoch` to every launch that uses a given pad slot.
//
// Reference: `llmctl.comms` is an empty package (reference llmctl/comms/__init__.py); SURVEY.md §5.8.
#pragma once
#include "../common/ptx.cuh"

namespace llmctl {
namespace symm {

constexpr int kMaxRanks = 8;

struct PeerTable {
  void* ptr[kMaxRanks];
};

__device__ __forceinline__ void signal_set(uint32_t* flag, uint32_t epoch) {
  __threadfence_system();
  ptx::st_release_sys(flag, epoch);
}
__device__ __forceinline__ void signal_wait_ge(const uint32_t* flag, uint32_t epoch) {
  while (static_cast<int32_t>(ptx::ld_acquire_sys(flag) - epoch) < 0) {
    __nanosleep(40);
  }
}

// All-rank barrier executed by ONE CTA (callers gate on blockIdx): thread r < world signals peer r's pad slot
// [slot_base + my_rank], then waits for peer r's arrival in the local pad slot [slot_base + r].
__device__ __forceinline__ void barrier_all(const PeerTable& pads, int rank, int world, int slot_base,
                                            uint32_t epoch) {
  __syncthreads();
  if (threadIdx.x < world) {
    uint32_t* remote = reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + slot_base + rank;
    signal_set(remote, epoch);
    const uint32_t* local = reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + slot_base + threadIdx.x;
    signal_wait_ge(local, epoch);
  }
  __syncthreads();
}

}  // namespace symm
}  // namespace llmctl
