// K0s: "skinny" bf16 GEMM for decode, C[Mb, N] = X[Mb, K] . W[N, K]^T with Mb <= 32 rows.
//
// A decode step multiplies a handful of token rows by every weight matrix of the model: the work is the
// weight stream (HBM-bound), not the FLOPs.  The general kernel tiles C as 128 x BN with X as the 128-row A
// operand, which for Mb = 8 (a) launches only N/128 CTAs — 32 for a 4096-wide layer, so 116 of the 148 SMs
// pull nothing — and (b) spends half of every pipeline stage on a mostly zero-filled X tile.  Here the
// operands are swapped and K is split:
//   * UMMA M = 128 weight rows (A operand = W tile, 16 KB per 64-wide k block), UMMA N = MB = 16 or 32 token
//     rows (B operand = X tile, 2-4 KB): the stage is almost all weights, 10-11 stages = ~170 KB of weight
//     bytes in flight per SM;
//   * stream-K: the (128-row tile, 64-wide k block) space is flattened and cut into one equal range per SM, so
//     every SM streams the same number of weight bytes through ONE pipeline that never drains at a tile
//     boundary (persistent CTA, barriers / TMEM set up once).  A CTA accumulates each tile piece in TMEM
//     (128 lanes x MB fp32 columns, double buffered so the epilogue overlaps the next piece); tiles cut by a
//     range boundary go through an fp32 workspace and the last piece to arrive (atomic ticket) sums the
//     pieces in slot order — deterministic — and writes bf16 C.  The ticket counter is reset by that CTA, so
//     the workspace needs no clearing between calls.
// Measured alternative (B200, 8 x 4096 -> 12288): two 5-stage CTAs per SM with four ranges per SM (hardware hands the
// second wave to whichever SM drains first) — 38.5 us against 28.9 us for one 11-stage CTA per SM; kept as is.
// Warp roles as in gemm_tcgen05.cu: warp 0 TMA, warp 1 MMA, warps 2-5 epilogue (thread = weight row).
//
// Reference: decode in the reference is the HF model's cuBLAS GEMMs over the whole padded batch (reference
// llmctl/serve/server.py:199-204).  Baseline: cuBLAS (gemv-like kernels) at the same shapes.
#include "../common/ptx.cuh"
#include "../common/tma_host.h"
#include "../common/util.cuh"

namespace llmctl {
using namespace ptx;

namespace skinny {
constexpr int kBM = 128;                 // weight rows per CTA (UMMA M)
constexpr int kBK = 64;                  // k elements per stage (128 bytes: one swizzle row)
constexpr int kWBytes = kBM * kBK * 2;   // 16 KB
constexpr int kThreads = 192;
template <int MB>
struct Cfg {
  static constexpr int kXBytes = MB * kBK * 2;                 // 2 KB (MB = 16) or 4 KB (MB = 32)
  static constexpr int kStageBytes = kWBytes + kXBytes;
  static constexpr int kStages = (MB == 16) ? 11 : 10;
  static constexpr int kSmem = kStages * kStageBytes + 1024;
};
}  // namespace skinny

template <int MB>
__global__ void __launch_bounds__(skinny::kThreads, 1)
gemm_skinny_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX,
                   __nv_bfloat16* __restrict__ C, long long ldc, int Mb, int N, int num_kb, int n_tiles, int L,
                   int max_slots, float* __restrict__ workspace, unsigned int* __restrict__ counters) {
  using namespace skinny;
  using C_ = Cfg<MB>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar_full[C_::kStages], bar_empty[C_::kStages], bar_acc_full[2], bar_acc_empty[2];
  __shared__ uint32_t tmem_base_smem;
  __shared__ int s_last;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // stream-K: the (tile, k block) space is flattened and cut into equal ranges, one per CTA
  const long long total = static_cast<long long>(n_tiles) * num_kb;
  const long long begin = static_cast<long long>(blockIdx.x) * L;
  const long long end = (begin + L < total) ? begin + L : total;

  if (threadIdx.x == 0) {
    for (int i = 0; i < C_::kStages; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_acc_full[i], 1);
      mbar_init(&bar_acc_empty[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmX);
  }
  if (warp == 1) tmem_alloc<1>(&tmem_base_smem, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int tile = static_cast<int>(begin / num_kb), kb = static_cast<int>(begin % num_kb);
      for (long long pos = begin; pos < end; ++pos) {
        mbar_wait(&bar_empty[stage], phase ^ 1u);
        uint8_t* sW = smem + stage * C_::kStageBytes;
        uint8_t* sX = sW + kWBytes;
        mbar_expect_tx(&bar_full[stage], C_::kStageBytes);
        tma_load_2d(sW, &tmW, &bar_full[stage], kb * kBK, tile * kBM);   // rows past N read as zero
        tma_load_2d(sX, &tmX, &bar_full[stage], kb * kBK, 0);            // rows past Mb read as zero
        if (++stage == C_::kStages) {
          stage = 0;
          phase ^= 1u;
        }
        if (++kb == num_kb) {
          kb = 0;
          ++tile;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kBM, MB, 0u, 0u);
      int stage = 0;
      uint32_t phase = 0;
      int seg = 0;
      for (long long pos = begin; pos < end; ++seg) {
        const int kb_s = static_cast<int>(pos % num_kb);
        const long long left = end - pos;
        const int seg_len = (num_kb - kb_s < left) ? num_kb - kb_s : static_cast<int>(left);
        const int buf = seg & 1;
        mbar_wait(&bar_acc_empty[buf], ((seg >> 1) & 1u) ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 32;
        for (int i = 0; i < seg_len; ++i) {
          mbar_wait(&bar_full[stage], phase);
          tc_fence_after();
          const uint32_t w_addr = smem_base + stage * C_::kStageBytes;
          const uint32_t x_addr = w_addr + kWBytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k)
            umma_ss<1>(d_tmem, make_smem_desc(w_addr + k * 32, 16, 1024, 2), make_smem_desc(x_addr + k * 32, 16, 1024, 2),
                       idesc, (i | k) != 0 ? 1u : 0u);
          umma_commit<1>(&bar_empty[stage]);
          if (++stage == C_::kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit<1>(&bar_acc_full[buf]);
        pos += seg_len;
      }
    }
  } else {
    // ------------------------------ epilogue (4 warps, thread = weight row) ------------------------------
    const int q = warp & 3;
    const int row = q * 32 + lane;
    int seg = 0;
    for (long long pos = begin; pos < end; ++seg) {
      const int tile = static_cast<int>(pos / num_kb);
      const int kb_s = static_cast<int>(pos % num_kb);
      const long long left = end - pos;
      const int seg_len = (num_kb - kb_s < left) ? num_kb - kb_s : static_cast<int>(left);
      const int buf = seg & 1;
      const int n = tile * kBM + row;
      uint32_t r[MB];
      mbar_wait(&bar_acc_full[buf], (seg >> 1) & 1u);
      tc_fence_after();
      const uint32_t taddr = tmem_base + buf * 32 + (static_cast<uint32_t>(q * 32) << 16);
      if constexpr (MB == 16) {
        tmem_ld_x16(taddr, r);
      } else {
        tmem_ld_x32(taddr, r);
      }
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_acc_empty[buf]);       // the MMA warp may start the segment after next
      // CTAs that own a piece of this tile: ranges are L long, so it is pure arithmetic
      const long long t_begin = static_cast<long long>(tile) * num_kb;
      const int first = static_cast<int>(t_begin / L);
      const int last = static_cast<int>((t_begin + num_kb - 1) / L);
      const int count = last - first + 1;
      if (count == 1) {
        if (n < N) {
#pragma unroll
          for (int m = 0; m < MB; ++m)
            if (m < Mb) C[m * ldc + n] = __float2bfloat16(__uint_as_float(r[m]));
        }
      } else {
        float* wt = workspace + static_cast<size_t>(tile) * max_slots * MB * kBM;
        float* ws = wt + static_cast<size_t>(static_cast<int>(blockIdx.x) - first) * MB * kBM;
#pragma unroll
        for (int m = 0; m < MB; ++m)
          if (m < Mb) ws[m * kBM + row] = __uint_as_float(r[m]);
        __threadfence();
        named_bar_sync(1, 128);
        if (threadIdx.x == 64) {
          const unsigned int ticket = atomicAdd(&counters[tile], 1u);
          s_last = (ticket == static_cast<unsigned int>(count - 1)) ? 1 : 0;
        }
        named_bar_sync(1, 128);
        if (s_last) {                        // last piece to arrive: sum the pieces in slot order (deterministic)
          __threadfence();
          if (n < N) {
#pragma unroll 1
            for (int m = 0; m < Mb; ++m) {
              float sum = 0.f;
              for (int sp = 0; sp < count; ++sp) sum += __ldcg(wt + (static_cast<size_t>(sp) * MB + m) * kBM + row);
              C[m * ldc + n] = __float2bfloat16(sum);
            }
          }
          if (threadIdx.x == 64) counters[tile] = 0u;     // ready for the next call on this stream
        }
      }
      pos += seg_len;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 64);
  }
}

}  // namespace llmctl

using namespace llmctl;

// C[Mb, N] (bf16, row stride ldc) = X[Mb, K] (row stride ldx) . W[N, K]^T (row stride ldw), 1 <= Mb <= 32.
// workspace: fp32 scratch for split-K partials (workspace_elems floats); counters: n_counters zero-initialised
// uint32 tickets (>= ceil(N/128)), left zero again on return.  One call at a time per (workspace, counters)
// pair: calls on the same stream are fine, concurrent streams need their own pair.  splits <= 0 picks it.
LLMCTL_API int llmctl_gemm_skinny_bf16(const void* X, const void* W, void* C, int Mb, int N, int K, long long ldx,
                                       long long ldw, long long ldc, float* workspace, long long workspace_elems,
                                       unsigned int* counters, int n_counters, int splits, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(Mb >= 1 && Mb <= 32 && N >= 1 && K >= 8 && K % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0);
  const int MB = Mb <= 16 ? 16 : 32;
  const int n_tiles = (N + skinny::kBM - 1) / skinny::kBM;
  const int num_kb = (K + skinny::kBK - 1) / skinny::kBK;
  const long long total = static_cast<long long>(n_tiles) * num_kb;
  const bool have_ws = workspace != nullptr && counters != nullptr && n_counters >= n_tiles;
  // k blocks per CTA: `splits` > 0 forces num_kb / splits (tests); otherwise an even share of all SMs, but
  // never less than 8 blocks (128 KB of weights) so small layers are not shredded
  long long L;
  if (splits > 0) {
    L = (num_kb + splits - 1) / splits;
  } else {
    L = (total + kNumSMs - 1) / kNumSMs;
    if (L < 8) L = 8;
  }
  auto slots_for = [&](long long l) { return (num_kb + l - 1) / l + 1; };
  if (!have_ws || static_cast<long long>(n_tiles) * slots_for(L) * MB * skinny::kBM > workspace_elems) {
    // no (or too little) scratch: whole tiles per CTA, no fix-up needed
    const long long tiles_per_cta = (n_tiles + kNumSMs - 1) / kNumSMs;
    L = tiles_per_cta * num_kb;
  }
  const int max_slots = static_cast<int>(slots_for(L));
  const unsigned grid = static_cast<unsigned>((total + L - 1) / L);
  CUtensorMap tw, tx;
  int rc = get_tmap_2d_bf16(&tw, W, N, K, ldw, skinny::kBM, skinny::kBK, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map W");
  rc = get_tmap_2d_bf16(&tx, X, Mb, K, ldx, MB, skinny::kBK, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map X");
  cudaStream_t s = (cudaStream_t)stream;
  if (MB == 16) {
    LLMCTL_CHECK(cudaFuncSetAttribute(gemm_skinny_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, skinny::Cfg<16>::kSmem));
    gemm_skinny_kernel<16><<<grid, skinny::kThreads, skinny::Cfg<16>::kSmem, s>>>(
        tw, tx, (__nv_bfloat16*)C, ldc, Mb, N, num_kb, n_tiles, static_cast<int>(L), max_slots, workspace, counters);
  } else {
    LLMCTL_CHECK(cudaFuncSetAttribute(gemm_skinny_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, skinny::Cfg<32>::kSmem));
    gemm_skinny_kernel<32><<<grid, skinny::kThreads, skinny::Cfg<32>::kSmem, s>>>(
        tw, tx, (__nv_bfloat16*)C, ldc, Mb, N, num_kb, n_tiles, static_cast<int>(L), max_slots, workspace, counters);
  }
  LLMCTL_RETURN_LAST();
}
