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

// ---- watchdog -------------------------------------------------------------------------------------------
// A peer that never signals (crashed rank, mismatched epoch, a launch that was skipped on one rank) would leave
// the waiting kernel spinning forever and wedge the GPU.  Every wait therefore carries a deadline: when it
// expires the waiter records what it was waiting for (flag address, expected epoch, last value seen, CTA) in a
// per-module record and *returns*, so the kernel drains and the host can read the record
// (`llmctl_comm_watchdog_poll`) and raise instead of hanging.  Data produced after a trip is garbage; the
// engine polls before it logs or checkpoints a step.
struct WatchdogRecord {
  unsigned int tripped;            // number of waits that timed out since the last poll
  unsigned int expected;           // epoch the first timed-out wait was waiting for
  unsigned int seen;               // last flag value it saw
  unsigned int block;              // blockIdx.x of that waiter
  unsigned long long flag_addr;    // address of the flag
  unsigned long long waited_ns;
};
static __device__ WatchdogRecord g_wd_record;
static __device__ unsigned long long g_wd_timeout_ns = 30ull * 1000ull * 1000ull * 1000ull;

__device__ __forceinline__ unsigned long long wd_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
static __device__ __noinline__ void wd_trip(const uint32_t* flag, uint32_t epoch, uint32_t seen, unsigned long long waited) {
  if (atomicAdd(&g_wd_record.tripped, 1u) == 0u) {
    g_wd_record.expected = epoch;
    g_wd_record.seen = seen;
    g_wd_record.block = blockIdx.x;
    g_wd_record.flag_addr = reinterpret_cast<unsigned long long>(flag);
    g_wd_record.waited_ns = waited;
  }
}

__device__ __forceinline__ void signal_wait_ge(const uint32_t* flag, uint32_t epoch) {
  uint32_t v = ptx::ld_acquire_sys(flag);
  if (static_cast<int32_t>(v - epoch) >= 0) return;
  const unsigned long long t0 = wd_now_ns();
  unsigned int spins = 0;
  while (static_cast<int32_t>((v = ptx::ld_acquire_sys(flag)) - epoch) < 0) {
    __nanosleep(40);
    if ((++spins & 1023u) == 0u) {
      const unsigned long long waited = wd_now_ns() - t0;
      if (waited > g_wd_timeout_ns) {
        wd_trip(flag, epoch, v, waited);
        return;
      }
    }
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

// Host side: every .cu that includes this header owns a private copy of the record (no relocatable device
// code), so each registers a reader / timeout setter with the registry in comm/watchdog.cu.
using WatchdogReadFn = int (*)(symm::WatchdogRecord* out, int clear);
using WatchdogTimeoutFn = int (*)(unsigned long long ns);
void register_watchdog_module(const char* name, WatchdogReadFn rd, WatchdogTimeoutFn st);

#define LLMCTL_SYMM_WATCHDOG_MODULE(modname)                                                            \
  namespace {                                                                                           \
  int wd_read_##modname(llmctl::symm::WatchdogRecord* out, int clear) {                                 \
    cudaError_t e = cudaMemcpyFromSymbol(out, llmctl::symm::g_wd_record, sizeof(*out));                 \
    if (e == cudaSuccess && clear && out->tripped) {                                                    \
      llmctl::symm::WatchdogRecord z{};                                                                 \
      e = cudaMemcpyToSymbol(llmctl::symm::g_wd_record, &z, sizeof(z));                                 \
    }                                                                                                   \
    return static_cast<int>(e);                                                                         \
  }                                                                                                     \
  int wd_timeout_##modname(unsigned long long ns) {                                                     \
    return static_cast<int>(cudaMemcpyToSymbol(llmctl::symm::g_wd_timeout_ns, &ns, sizeof(ns)));        \
  }                                                                                                     \
  struct WdReg_##modname {                                                                              \
    WdReg_##modname() { llmctl::register_watchdog_module(#modname, wd_read_##modname, wd_timeout_##modname); } \
  } wd_reg_instance_##modname;                                                                          \
  }
}  // namespace llmctl
