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
//   * grid = (N/128) x splits with splits chosen so that every SM has a CTA; each CTA accumulates its K range
//     in TMEM (128 lanes x MB fp32 columns), writes the partial to an fp32 workspace, and the last CTA of a
//     tile to arrive (atomic ticket) sums the partials in split order — deterministic — and writes bf16 C.
//     The ticket counter is reset by that CTA, so the workspace needs no clearing between calls.
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
                   __nv_bfloat16* __restrict__ C, long long ldc, int Mb, int N, int num_kb, int splits,
                   float* __restrict__ workspace, unsigned int* __restrict__ counters) {
  using namespace skinny;
  using C_ = Cfg<MB>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar_full[C_::kStages], bar_empty[C_::kStages], bar_acc;
  __shared__ uint32_t tmem_base_smem;
  __shared__ int s_last;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x / splits, split = blockIdx.x % splits;
  const int n0 = tile * kBM;
  // this CTA's k blocks: [kb0, kb1)
  const int per = (num_kb + splits - 1) / splits;
  const int kb0 = split * per;
  const int kb1 = min(num_kb, kb0 + per);
  const int nkb = max(kb1 - kb0, 0);

  if (threadIdx.x == 0) {
    for (int i = 0; i < C_::kStages; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], 1);
    }
    mbar_init(&bar_acc, 1);
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmX);
  }
  if (warp == 1) tmem_alloc<1>(&tmem_base_smem, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&bar_empty[stage], phase ^ 1u);
        uint8_t* sW = smem + stage * C_::kStageBytes;
        uint8_t* sX = sW + kWBytes;
        mbar_expect_tx(&bar_full[stage], C_::kStageBytes);
        tma_load_2d(sW, &tmW, &bar_full[stage], kb * kBK, n0);     // rows past N read as zero
        tma_load_2d(sX, &tmX, &bar_full[stage], kb * kBK, 0);      // rows past Mb read as zero
        if (++stage == C_::kStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kBM, MB, 0u, 0u);
      int stage = 0;
      uint32_t phase = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait(&bar_full[stage], phase);
        tc_fence_after();
        const uint32_t w_addr = smem_base + stage * C_::kStageBytes;
        const uint32_t x_addr = w_addr + kWBytes;
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k)
          umma_ss<1>(tmem_base, make_smem_desc(w_addr + k * 32, 16, 1024, 2), make_smem_desc(x_addr + k * 32, 16, 1024, 2),
                     idesc, (i | k) != 0 ? 1u : 0u);
        umma_commit<1>(&bar_empty[stage]);
        if (++stage == C_::kStages) {
          stage = 0;
          phase ^= 1u;
        }
      }
      umma_commit<1>(&bar_acc);
    }
  } else {
    // ------------------------------ epilogue (4 warps, thread = weight row) ------------------------------
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int n = n0 + row;
    uint32_t r[MB];
    if (nkb > 0) {
      mbar_wait(&bar_acc, 0);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
      if constexpr (MB == 16) {
        tmem_ld_x16(taddr, r);
      } else {
        tmem_ld_x32(taddr, r);
      }
      tmem_ld_wait();
    } else {
#pragma unroll
      for (int m = 0; m < MB; ++m) r[m] = 0u;
    }
    if (splits == 1) {
      if (n < N) {
#pragma unroll
        for (int m = 0; m < MB; ++m)
          if (m < Mb) C[m * ldc + n] = __float2bfloat16(__uint_as_float(r[m]));
      }
    } else {
      float* ws = workspace + (static_cast<size_t>(tile) * splits + split) * MB * kBM;
#pragma unroll
      for (int m = 0; m < MB; ++m)
        if (m < Mb) ws[m * kBM + row] = __uint_as_float(r[m]);
      __threadfence();
      named_bar_sync(1, 128);
      if (threadIdx.x == 64) {
        const unsigned int ticket = atomicAdd(&counters[tile], 1u);
        s_last = (ticket == static_cast<unsigned int>(splits - 1)) ? 1 : 0;
      }
      named_bar_sync(1, 128);
      if (s_last) {
        __threadfence();
        if (n < N) {
          const float* wt = workspace + static_cast<size_t>(tile) * splits * MB * kBM;
#pragma unroll 1
          for (int m = 0; m < Mb; ++m) {
            float sum = 0.f;
            for (int sp = 0; sp < splits; ++sp) sum += __ldcg(wt + (static_cast<size_t>(sp) * MB + m) * kBM + row);
            C[m * ldc + n] = __float2bfloat16(sum);
          }
        }
        if (threadIdx.x == 64) counters[tile] = 0u;     // ready for the next call on this stream
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 32);
  }
}

// cost model for the split count: waves over 148 SMs x (k blocks per CTA + pipeline fill, ~6 k blocks)
static int choose_splits(int n_tiles, int num_kb, long long max_by_workspace) {
  int best = 1;
  long long best_cost = -1;
  for (int s = 1; s <= 8 && s <= num_kb; ++s) {
    if (s > 1 && s > max_by_workspace) break;
    const int per = (num_kb + s - 1) / s;
    const int eff = (num_kb + per - 1) / per;              // splits that actually get k blocks
    if (eff != s) continue;
    const long long waves = (static_cast<long long>(n_tiles) * s + kNumSMs - 1) / kNumSMs;
    const long long cost = waves * (per + 6);
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best = s;
    }
  }
  return best;
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
  const long long per_split_elems = static_cast<long long>(n_tiles) * MB * skinny::kBM;
  long long max_ws = (workspace != nullptr && counters != nullptr && n_counters >= n_tiles) ? workspace_elems / per_split_elems : 1;
  if (splits <= 0) splits = choose_splits(n_tiles, num_kb, max_ws);
  if (splits > max_ws) splits = static_cast<int>(max_ws < 1 ? 1 : max_ws);
  if (splits > num_kb) splits = num_kb;
  {
    const int per = (num_kb + splits - 1) / splits;      // drop splits that would get no k block
    splits = (num_kb + per - 1) / per;
  }
  CUtensorMap tw, tx;
  int rc = get_tmap_2d_bf16(&tw, W, N, K, ldw, skinny::kBM, skinny::kBK, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map W");
  rc = get_tmap_2d_bf16(&tx, X, Mb, K, ldx, MB, skinny::kBK, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map X");
  cudaStream_t s = (cudaStream_t)stream;
  const unsigned grid = static_cast<unsigned>(n_tiles) * splits;
  if (MB == 16) {
    LLMCTL_CHECK(cudaFuncSetAttribute(gemm_skinny_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, skinny::Cfg<16>::kSmem));
    gemm_skinny_kernel<16><<<grid, skinny::kThreads, skinny::Cfg<16>::kSmem, s>>>(tw, tx, (__nv_bfloat16*)C, ldc, Mb, N, num_kb,
                                                                                  splits, workspace, counters);
  } else {
    LLMCTL_CHECK(cudaFuncSetAttribute(gemm_skinny_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, skinny::Cfg<32>::kSmem));
    gemm_skinny_kernel<32><<<grid, skinny::kThreads, skinny::Cfg<32>::kSmem, s>>>(tw, tx, (__nv_bfloat16*)C, ldc, Mb, N, num_kb,
                                                                                  splits, workspace, counters);
  }
  LLMCTL_RETURN_LAST();
}
