// K1 producer side: sequence-parallel all-gather by PUSH.
//
// Every rank owns rows [r*Ms, (r+1)*Ms) of the activation X[M, K].  Instead of the consumer pulling seven remote
// shards (NVLink read round trips in front of the tensor core), the owner pushes its rows into the same offset of
// every rank's symmetric buffer Xg — one NVLS `multimem.st` per 16 bytes, replicated by the switch (egress = one
// shard, not seven), or P2P stores when no multicast mapping exists — slab by slab, and raises the slab's epoch
// flag on every rank once the slab is complete.  The consumer is the tcgen05 GEMM itself
// (llmctl_gemm_bf16_gated, csrc/gemm/gemm_tcgen05.cu): all 148 SMs do MMAs, its TMA producer waits per tile for
// the flags of the rows it needs, so the transfer of later slabs overlaps the math on earlier ones.
//
// Flag f = row / rows_per_flag lives at pads[dst][flag_base + f]; epochs increase per call, flags are never reset.
// Buffer reuse: Xg alternates between two buffers; a rank can only push call n after its own GEMM n-1 finished,
// which waited for every peer's push n-1, which those peers issued after their GEMM n-2 — so nobody still reads the
// buffer that call n overwrites.
//
// Reference: none (SP is a preset flag only: reference configs/presets/llama-7b-a100x8.toml:36).  SURVEY.md §2.6.b K1.
#include "../comm/symm.cuh"
#include "../common/util.cuh"

namespace llmctl {
using namespace symm;

constexpr int kPushThreads = 256;   // small enough to stay co-resident with a persistent GEMM CTA on the same SM

__global__ void __launch_bounds__(kPushThreads)
ag_push_kernel(const uint4* __restrict__ src, uint4* __restrict__ mc_dst, PeerTable peers, long long dst_off_vec,
               long long slab_vec, int n_slabs, PeerTable pads, int flag_base, int first_flag, int rank, int world,
               uint32_t epoch, uint32_t* __restrict__ counters, int skip_self) {
  __shared__ unsigned int s_last;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  // P2P destinations in staggered order (rank, rank+1, ...): resolved once, kept in registers
  uint4* dptr[kMaxRanks];
#pragma unroll
  for (int i = 0; i < kMaxRanks; ++i) {
    int d = rank + i;
    if (d >= world) d -= world;
    dptr[i] = (i < world && !(i == 0 && skip_self)) ? reinterpret_cast<uint4*>(peers.ptr[d]) : nullptr;
  }
  for (int sl = 0; sl < n_slabs; ++sl) {
    const uint4* s_src = src + sl * slab_vec;
    const long long base = dst_off_vec + sl * slab_vec;
    long long v = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (mc_dst != nullptr) {
      for (; v + 7 * stride < slab_vec; v += 8 * stride) {       // 8 independent 16-byte loads in flight per thread
        uint4 a[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) a[u] = s_src[v + u * stride];
#pragma unroll
        for (int u = 0; u < 8; ++u) ptx::multimem_st_v4(mc_dst + base + v + u * stride, a[u]);
      }
      for (; v < slab_vec; v += stride) ptx::multimem_st_v4(mc_dst + base + v, s_src[v]);
    } else {
      auto put = [&](long long at, const uint4& a) {
#pragma unroll
        for (int i = 0; i < kMaxRanks; ++i)
          if (dptr[i] != nullptr) ptx::st_v4(dptr[i] + base + at, a);      // staggered: no egress hot-spot
      };
      for (; v + 3 * stride < slab_vec; v += 4 * stride) {
        uint4 a[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) a[u] = s_src[v + u * stride];
#pragma unroll
        for (int u = 0; u < 4; ++u) put(v + u * stride, a[u]);
      }
      for (; v < slab_vec; v += stride) put(v, s_src[v]);
    }
    // slab complete on this CTA -> the last CTA to get here raises the slab's flag on every rank
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
      const unsigned int done = atomicAdd(&counters[sl], 1u) + 1u;
      s_last = (done == gridDim.x) ? 1u : 0u;
      if (s_last) atomicExch(&counters[sl], 0u);
    }
    __syncthreads();
    // (the local copy needs a flag too when this kernel wrote it, i.e. when the source was not already in place)
    if (s_last && threadIdx.x < world && (static_cast<int>(threadIdx.x) != rank || !skip_self))
      signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + flag_base + first_flag + sl, epoch);
    __syncthreads();
  }
}

}  // namespace llmctl

using namespace llmctl;
LLMCTL_SYMM_WATCHDOG_MODULE(ag_push)

// Push this rank's [rows, cols] bf16 block (contiguous, `src`) to rows [rank*rows, (rank+1)*rows) of every rank's
// symmetric [world*rows, cols] buffer.
//   xg / mc_xg : per-rank base pointers / NVLS multicast address (or NULL) of the symmetric buffer
//   rows_per_flag : slab size; must divide rows.  Flags used: pads[*][flag_base + rank*rows/rows_per_flag + slab]
//   counters   : >= rows/rows_per_flag zero-initialised uint32 in local memory (left zero on exit)
//   src_in_xg  : 1 when `src` already IS this rank's block of the local buffer (no local write needed on the P2P path)
LLMCTL_API int llmctl_ag_push(const void* src, llmctl::symm::PeerTable xg, void* mc_xg, int rows, int cols,
                              int rows_per_flag, llmctl::symm::PeerTable pads, int flag_base, int rank, int world,
                              unsigned int epoch, unsigned int* counters, int src_in_xg, int max_ctas,
                              void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && cols % 8 == 0 && rows > 0 && rows_per_flag > 0 &&
                 rows % rows_per_flag == 0 && rows / rows_per_flag <= 32 &&
                 flag_base + world * (rows / rows_per_flag) <= 480);
  const int n_slabs = rows / rows_per_flag;
  const long long slab_vec = static_cast<long long>(rows_per_flag) * cols / 8;
  const long long dst_off_vec = static_cast<long long>(rank) * rows * cols / 8;
  long long grid = (slab_vec + kPushThreads - 1) / kPushThreads;
  if (max_ctas <= 0) max_ctas = kNumSMs;
  if (grid > max_ctas) grid = max_ctas;
  if (grid < 1) grid = 1;
  ag_push_kernel<<<static_cast<int>(grid), kPushThreads, 0, (cudaStream_t)stream>>>(
      (const uint4*)src, (uint4*)mc_xg, xg, dst_off_vec, slab_vec, n_slabs, pads, flag_base, rank * n_slabs, rank,
      world, epoch, counters, src_in_xg);
  LLMCTL_RETURN_LAST();
}

namespace llmctl {
__global__ void __launch_bounds__(32)
signal_all_kernel(symm::PeerTable pads, int rank, int world, int slot, uint32_t epoch) {
  if (threadIdx.x < world)
    symm::signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + slot + rank, epoch);
}
}  // namespace llmctl

// Stream-ordered "I am done" announcement: pads[r][slot + rank] = epoch on every rank r (the counterpart of
// llmctl_signal_wait; used to hand a symmetric buffer back before its next writer may touch it).
LLMCTL_API int llmctl_signal_all(llmctl::symm::PeerTable pads, int rank, int world, int slot, unsigned int epoch,
                                 void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && slot >= 0 && slot < 480);
  signal_all_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pads, rank, world, slot, epoch);
  LLMCTL_RETURN_LAST();
}

extern "C" int llmctl_gemm_bf16_gated(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                                      long long ldb, long long ldc, int b_mn, int out_fp32, int beta, int tile_n,
                                      int max_ctas, const unsigned int* flags, unsigned int epoch,
                                      int rows_per_flag, int rank, int world, void* stream);

namespace {
struct ForkJoinEvents {
  cudaEvent_t ready = nullptr, pushed = nullptr;
};
ForkJoinEvents g_fj[64];
}  // namespace

// K1 in one call: Y[M, N] = all_gather(X rows)[M, K] * W^T.  This rank's rows must already sit in its block of the
// local symmetric buffer (xg.ptr[rank] + rank*rows*K).  The push runs on `push_stream`, forked from / joined to
// `stream` with events, so it overlaps the GEMM's work on the local rows; the GEMM consumes remote slabs as their
// flags arrive.  W: [N, K] (ldw) or, with w_mn, stored [K, N].  tile_n as in llmctl_gemm_bf16 (the time between the
// last slab's arrival and the end of the kernel is one tile, so smaller tiles shorten the un-overlapped tail).
// The first call on a device creates two events.
LLMCTL_API int llmctl_ag_gemm(llmctl::symm::PeerTable xg, void* mc_xg, const void* W, void* Y, int M, int N, int K,
                              long long ldw, long long ldy, int w_mn, int rows_per_flag,
                              llmctl::symm::PeerTable pads, int flag_base, int rank, int world, unsigned int epoch,
                              unsigned int* counters, int push_ctas, int tile_n, void* push_stream, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && M % world == 0);
  int dev = 0;
  LLMCTL_CHECK(cudaGetDevice(&dev));
  LLMCTL_REQUIRE(dev >= 0 && dev < 64);
  ForkJoinEvents& fj = g_fj[dev];
  if (fj.ready == nullptr) {
    LLMCTL_CHECK(cudaEventCreateWithFlags(&fj.ready, cudaEventDisableTiming));
    LLMCTL_CHECK(cudaEventCreateWithFlags(&fj.pushed, cudaEventDisableTiming));
  }
  cudaStream_t main_s = (cudaStream_t)stream, side = (cudaStream_t)push_stream;
  const int rows = M / world;
  const __nv_bfloat16* mine = reinterpret_cast<const __nv_bfloat16*>(xg.ptr[rank]) + static_cast<long long>(rank) * rows * K;
  if (world > 1) {
    LLMCTL_CHECK(cudaEventRecord(fj.ready, main_s));
    LLMCTL_CHECK(cudaStreamWaitEvent(side, fj.ready, 0));
    int rc = llmctl_ag_push(mine, xg, mc_xg, rows, K, rows_per_flag, pads, flag_base, rank, world, epoch, counters, 1,
                            push_ctas, side);
    if (rc) return rc;
    LLMCTL_CHECK(cudaEventRecord(fj.pushed, side));
  }
  int rc = llmctl_gemm_bf16_gated(xg.ptr[rank], W, Y, M, N, K, K, ldw, ldy, w_mn, 0, 0, tile_n, 0,
                                  reinterpret_cast<const unsigned int*>(pads.ptr[rank]) + flag_base, epoch,
                                  rows_per_flag, rank, world, stream);
  if (rc) return rc;
  if (world > 1) LLMCTL_CHECK(cudaStreamWaitEvent(main_s, fj.pushed, 0));   // buffer reuse: our own push is over
  return 0;
}
