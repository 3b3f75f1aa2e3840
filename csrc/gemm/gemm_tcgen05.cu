// K0: bf16 GEMM on the 5th-generation tensor cores.  D[M,N] (+)= op(A) * op(B), fp32 accumulation in
// tensor memory.
//
//   * operands staged by TMA (cp.async.bulk.tensor, 128-byte swizzle) into a multi-stage smem ring
//   * one elected thread issues tcgen05.mma (UMMA 128 x BN x 16, kind::f16) per k-slice
//   * accumulators double-buffered in TMEM (2 x BN columns) so the epilogue of tile i overlaps the
//     main loop of tile i+1
//   * persistent CTAs (one per SM) walk the tile list; warp roles: 0 = TMA producer, 1 = MMA issuer
//     (+ TMEM alloc), 2..5 = epilogue (tcgen05.ld -> convert -> global)
//
// Operand layouts (both operands independently):
//   K-major  : A is [M,K] row-major (K contiguous) / B is [N,K] row-major   (torch Linear weight)
//   MN-major : A is stored [K,M] row-major (M contiguous) / B is stored [K,N] row-major
// which covers forward (x Wt), dgrad (dy W) and wgrad (dyT x) without any transposes in HBM.
//
// Reference: none — every GEMM in the reference is a cuBLAS call made by HF modules behind
// `self.model(**batch)` (reference llmctl/runtime/engine.py:284) and `torch.matmul`
// (reference llmctl/cli/commands/hw.py:317-337, llmctl/plugins/autotuning.py:89-114).
// Baseline to beat: cuBLASLt bf16 (MEASURED_PEAKS.json).
#include "../comm/symm.cuh"
#include "../common/tma_host.h"
#include "../common/util.cuh"

namespace llmctl {
using namespace ptx;

// K1 (sequence-parallel all-gather -> GEMM) runs on this kernel: the rows of A live in a symmetric buffer that the
// peers' producers fill over NVLink (csrc/fused/ag_push.cu); every `rows_per_flag` rows carry an epoch flag in the
// local signal pad.  The TMA producer waits for the flags of the rows a tile needs, so the MMAs of what has arrived
// overlap the transfer of what has not; tile rows are visited in arrival order (own chunk first, then the first
// slab of every peer, the second slab of every peer, ...).  flags == NULL: plain GEMM.
struct GemmGate {
  const uint32_t* flags;
  uint32_t epoch;
  int rows_per_flag;
  int own_lo, own_hi;        // rows that are local and never waited for
  int perm;                  // 1: arrival-order permutation of the tile-row blocks
  int rank, world;
  int bpc, bps;              // tile-row blocks per rank chunk / per flag slab
  // K8 (ZeRO-3 parameter gather consumed by the GEMM): the B operand lives in a symmetric buffer that the owners of
  // its shards fill over NVLink; flag f covers flat elements [f*b_slab, (f+1)*b_slab) of that buffer.
  const uint32_t* b_flags;
  uint32_t b_epoch;
  long long b_off;           // element offset of B inside the gathered buffer
  long long b_slab;          // elements per flag
  long long b_row_stride;    // K-major B: elements between consecutive N rows
  long long b_span;          // MN-major B: elements of the whole matrix (a tile touches every K row)
};
__device__ __forceinline__ int gate_perm(const GemmGate& g, int i) {
  if (!g.perm) return i;
  if (i < g.bpc) return g.rank * g.bpc + i;
  const int j = i - g.bpc, per_s = (g.world - 1) * g.bps;
  const int s = j / per_s, w = j - s * per_s, p = w / g.bps, b = w - p * g.bps;
  return ((g.rank + 1 + p) % g.world) * g.bpc + s * g.bps + b;
}
__device__ __forceinline__ void gate_wait(const GemmGate& g, int m0, int m_end) {
  if (g.flags == nullptr) return;
  const int R = g.rows_per_flag;
  for (int f = m0 / R; f <= (m_end - 1) / R; ++f) {
    const int fr = f * R;
    if (fr >= g.own_lo && fr < g.own_hi) continue;
    symm::signal_wait_ge(g.flags + f, g.epoch);
  }
  fence_proxy_async_all();      // generic-proxy flag reads -> async-proxy (TMA) reads of the rows
}
template <bool B_MN>
__device__ __forceinline__ void gate_wait_b(const GemmGate& g, int n0, int n_end) {
  if (g.b_flags == nullptr || n0 >= n_end) return;
  long long lo, hi;
  if (!B_MN) {
    lo = g.b_off + static_cast<long long>(n0) * g.b_row_stride;
    hi = g.b_off + static_cast<long long>(n_end) * g.b_row_stride;
  } else {
    lo = g.b_off;
    hi = g.b_off + g.b_span;
  }
  for (long long f = lo / g.b_slab; f <= (hi - 1) / g.b_slab; ++f) symm::signal_wait_ge(g.b_flags + f, g.b_epoch);
  fence_proxy_async_all();
}

constexpr int kBM = 128;
constexpr int kBK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int kGemmThreads = 192;

template <int BN>
struct GemmCfg {
  static constexpr int kStages = (BN == 256) ? 4 : (BN == 192 ? 5 : 6);
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024;  // + alignment slack
  static constexpr int kTmemCols = (2 * BN <= 256) ? 256 : 512;   // power of two >= 2 accumulators of BN columns
};

// Epilogue modes
enum : int { kEpiStore = 0, kEpiAccumulate = 1 };

// Grouped rasterisation: consecutive tile ids walk kGroupM row-blocks before moving to the next column block, so
// the tiles that are in flight at any moment share a small set of A and B panels (both stay L2-resident even
// when an operand is larger than L2; measured on the 7B wgrad shape: DRAM reads 1.06 GB -> ~0.2 GB).
constexpr int kGroupM = 8;
__device__ __forceinline__ void tile_to_mn(int tile, int num_m, int num_n, int& m_blk, int& n_blk) {
  const int per_group = kGroupM * num_n;
  const int group = tile / per_group;
  const int first_m = group * kGroupM;
  const int gsize = min(num_m - first_m, kGroupM);
  const int in_group = tile - group * per_group;
  m_blk = first_m + in_group % gsize;
  n_blk = in_group / gsize;
}

template <int BN, bool A_MN, bool B_MN, typename OutT>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    OutT* __restrict__ C, long long ldc, int M, int N, int K, int beta, const GemmGate gate) {
  using Cfg = GemmCfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar_full[Cfg::kStages];
  __shared__ uint64_t bar_empty[Cfg::kStages];
  __shared__ uint64_t bar_tfull[2];
  __shared__ uint64_t bar_tempty[2];
  __shared__ uint32_t tmem_base_smem;

  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int num_m = (M + kBM - 1) / kBM;
  const int num_n = (N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_kb = (K + kBK - 1) / kBK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::kStages; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_tfull[i], 1);
      mbar_init(&bar_tempty[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
  }
  if (warp == 1) tmem_alloc<1>(&tmem_base_smem, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ------------------------------ TMA producer ------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int mb_, nb_;
        tile_to_mn(tile, num_m, num_n, mb_, nb_);
        mb_ = gate_perm(gate, mb_);
        const int m0 = mb_ * kBM;
        const int n0 = nb_ * BN;
        gate_wait(gate, m0, min(m0 + kBM, M));
        gate_wait_b<B_MN>(gate, n0, min(n0 + BN, N));
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&bar_empty[stage], phase ^ 1u);
          uint8_t* sA = smem + stage * Cfg::kStageBytes;
          uint8_t* sB = sA + Cfg::kABytes;
          mbar_expect_tx(&bar_full[stage], Cfg::kStageBytes);
          const int k0 = kb * kBK;
          if (!A_MN) {
            tma_load_2d(sA, &tmA, &bar_full[stage], k0, m0);
          } else {
#pragma unroll
            for (int g = 0; g < kBM / 64; ++g)
              tma_load_2d(sA + g * 8192, &tmA, &bar_full[stage], m0 + g * 64, k0);
          }
          if (!B_MN) {
            tma_load_2d(sB, &tmB, &bar_full[stage], k0, n0);
          } else {
#pragma unroll
            for (int g = 0; g < BN / 64; ++g)
              tma_load_2d(sB + g * 8192, &tmB, &bar_full[stage], n0 + g * 64, k0);
          }
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer ------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kBM, BN, A_MN ? 1u : 0u, B_MN ? 1u : 0u);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t aphase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&bar_tempty[acc], aphase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&bar_full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_base + stage * Cfg::kStageBytes;
          const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = A_MN ? make_smem_desc(a_addr + k * 2048, 8192, 1024, 2)
                                     : make_smem_desc(a_addr + k * 32, 16, 1024, 2);
            const uint64_t db = B_MN ? make_smem_desc(b_addr + k * 2048, 8192, 1024, 2)
                                     : make_smem_desc(b_addr + k * 32, 16, 1024, 2);
            umma_ss<1>(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit<1>(&bar_empty[stage]);  // frees the smem slot once these MMAs retire
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit<1>(&bar_tfull[acc]);  // accumulator complete -> epilogue
        acc ^= 1;
        if (acc == 0) aphase ^= 1u;
      }
    }
  } else {
    // ------------------------------ epilogue (4 warps) ------------------------------
    const int q = warp & 3;  // TMEM lane quadrant this warp may access
    int acc = 0;
    uint32_t aphase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int mb_, nb_;
      tile_to_mn(tile, num_m, num_n, mb_, nb_);
      mb_ = gate_perm(gate, mb_);
      const int m0 = mb_ * kBM;
      const int n0 = nb_ * BN;
      mbar_wait(&bar_tfull[acc], aphase);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                             static_cast<uint32_t>(acc * BN);
      OutT* crow = C + static_cast<long long>(row) * ldc + n0;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_x32(t_row + c * 32, r);
        tmem_ld_wait();
        if (row < M) {
          if constexpr (sizeof(OutT) == 2) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int col = n0 + c * 32 + j * 8;
              if (col < N) {
                float f[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[j * 8 + e]);
                __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(crow) + c * 32 + j * 8;
                if (beta) {
                  float old[8];
                  unpack8(load8(dst), old);
#pragma unroll
                  for (int e = 0; e < 8; ++e) f[e] += old[e];
                }
                store8(dst, pack8(f));
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int col = n0 + c * 32 + j * 4;
              if (col < N) {
                float4 v = make_float4(__uint_as_float(r[j * 4]), __uint_as_float(r[j * 4 + 1]),
                                       __uint_as_float(r[j * 4 + 2]), __uint_as_float(r[j * 4 + 3]));
                float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(crow) + c * 32 + j * 4);
                if (beta) {
                  const float4 o = *dst;
                  v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                }
                *dst = v;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_tempty[acc]);
      acc ^= 1;
      if (acc == 0) aphase ^= 1u;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, Cfg::kTmemCols);
  }
}


// ================================================================================================
// 2-CTA variant (cta_group::2): a CTA pair (cluster of 2, same TPC) computes one 256 x 256 tile.
// CTA r stages A rows [m0 + 128 r, +128) and B rows (N) [n0 + 128 r, +128); the leader CTA issues
// UMMA 256 x 256 x 16 which reads both CTAs' shared memory, so each SM moves only half of B through
// its own shared memory (smem traffic per SM: 64 KB/stage-pair instead of 96 KB) — this is what lets
// the tensor pipe stay fed.  Each CTA's TMEM holds its own 128 accumulator rows.
//   full[s]   (leader only): 1 arrive (leader producer expect_tx of BOTH CTAs' bytes) + tx bytes
//   empty[s]  (both CTAs)  : tcgen05.commit multicast from the leader
//   tfull[a]  (both CTAs)  : tcgen05.commit multicast from the leader
//   tempty[a] (leader only): 8 arrives = 4 epilogue warps x 2 CTAs (peer arrives remotely)
// ================================================================================================
// PBN = N of the pair tile: 256 (the throughput tile) or 128 (half the work per tile: fills the last wave of the
// 1-3 wave problems of a TP=8 shard, where a 256 x 256 tile is 26 us of tensor-core time).
template <int PBN>
struct Gemm2Cfg {
  static constexpr int kBN = PBN;
  static constexpr int kStages = (PBN == 256) ? 6 : 8;
  static constexpr int kABytes = kBM * kBK * 2;          // 128 x 64 per CTA
  static constexpr int kBBytes = (kBN / 2) * kBK * 2;    // (PBN/2) x 64 per CTA
  static constexpr int kStageBytes = kABytes + kBBytes;  // 32 / 24 KB per CTA
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024;
  static constexpr int kTmemCols = 2 * PBN;              // 512 / 256
};

template <int PBN, bool A_MN, bool B_MN, typename OutT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kGemmThreads, 1)
gemm2_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     OutT* __restrict__ C, long long ldc, int M, int N, int K, int beta, const GemmGate gate) {
  using Cfg = Gemm2Cfg<PBN>;
  constexpr int BN = Cfg::kBN;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar_full[Cfg::kStages];
  __shared__ uint64_t bar_empty[Cfg::kStages];
  __shared__ uint64_t bar_tfull[2];
  __shared__ uint64_t bar_tempty[2];
  __shared__ uint32_t tmem_base_smem;

  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();   // 0 = leader
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;

  const int num_m = (M + 2 * kBM - 1) / (2 * kBM);
  const int num_n = (N + BN - 1) / BN;
  const int num_tiles = num_m * num_n;
  const int num_kb = (K + kBK - 1) / kBK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < Cfg::kStages; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_tfull[i], 1);
      mbar_init(&bar_tempty[i], 8);
    }
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmB);
  }
  cluster_sync();   // barrier inits visible to the peer before any remote arrive / multicast commit
  if (warp == 1) tmem_alloc<2>(&tmem_base_smem, Cfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  if (warp == 0) {
    // ------------------------------ TMA producer (both CTAs) ------------------------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        int mb_, nb_;
        tile_to_mn(tile, num_m, num_n, mb_, nb_);
        mb_ = gate_perm(gate, mb_);
        const int m0 = mb_ * (2 * kBM) + static_cast<int>(cta_rank) * kBM;
        const int n0 = nb_ * BN + static_cast<int>(cta_rank) * (BN / 2);
        if (m0 < M) gate_wait(gate, m0, min(m0 + kBM, M));
        gate_wait_b<B_MN>(gate, n0, min(n0 + BN / 2, N));
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&bar_empty[stage], phase ^ 1u);
          uint8_t* sA = smem + stage * Cfg::kStageBytes;
          uint8_t* sB = sA + Cfg::kABytes;
          // all bytes of the pair are accounted on the leader's barrier
          const uint32_t full_leader = mapa_u32(smem_u32(&bar_full[stage]), 0);
          if (cta_rank == 0) mbar_expect_tx(&bar_full[stage], 2 * Cfg::kStageBytes);
          const int k0 = kb * kBK;
          if (!A_MN) {
            tma_load_2d_cg2(sA, &tmA, full_leader, k0, m0);
          } else {
#pragma unroll
            for (int g = 0; g < kBM / 64; ++g)
              tma_load_2d_cg2(sA + g * 8192, &tmA, full_leader, m0 + g * 64, k0);
          }
          if (!B_MN) {
            tma_load_2d_cg2(sB, &tmB, full_leader, k0, n0);
          } else {
#pragma unroll
            for (int g = 0; g < (BN / 2) / 64; ++g)
              tma_load_2d_cg2(sB + g * 8192, &tmB, full_leader, n0 + g * 64, k0);
          }
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------ MMA issuer (leader CTA only) ------------------------------
    if (cta_rank == 0 && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * kBM, BN, A_MN ? 1u : 0u, B_MN ? 1u : 0u);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t aphase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        mbar_wait(&bar_tempty[acc], aphase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&bar_full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_base + stage * Cfg::kStageBytes;
          const uint32_t b_addr = a_addr + Cfg::kABytes;
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = A_MN ? make_smem_desc(a_addr + k * 2048, 8192, 1024, 2)
                                     : make_smem_desc(a_addr + k * 32, 16, 1024, 2);
            const uint64_t db = B_MN ? make_smem_desc(b_addr + k * 2048, 8192, 1024, 2)
                                     : make_smem_desc(b_addr + k * 32, 16, 1024, 2);
            umma_ss<2>(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit_mc2(&bar_empty[stage], 0x3);   // both CTAs' slots are free once these retire
          if (++stage == Cfg::kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit_mc2(&bar_tfull[acc], 0x3);
        acc ^= 1;
        if (acc == 0) aphase ^= 1u;
      }
    }
  } else {
    // ------------------------------ epilogue (4 warps, both CTAs) ------------------------------
    const int q = warp & 3;
    int acc = 0;
    uint32_t aphase = 0;
    for (int tile = pair; tile < num_tiles; tile += num_pairs) {
      int mb_, nb_;
      tile_to_mn(tile, num_m, num_n, mb_, nb_);
      mb_ = gate_perm(gate, mb_);
      const int m0 = mb_ * (2 * kBM) + static_cast<int>(cta_rank) * kBM;
      const int n0 = nb_ * BN;
      mbar_wait(&bar_tfull[acc], aphase);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                             static_cast<uint32_t>(acc * BN);
      OutT* crow = C + static_cast<long long>(row) * ldc + n0;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_x32(t_row + c * 32, r);
        tmem_ld_wait();
        if (row < M) {
          if constexpr (sizeof(OutT) == 2) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const int col = n0 + c * 32 + j * 8;
              if (col < N) {
                float f[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[j * 8 + e]);
                __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(crow) + c * 32 + j * 8;
                if (beta) {
                  float old[8];
                  unpack8(load8(dst), old);
#pragma unroll
                  for (int e = 0; e < 8; ++e) f[e] += old[e];
                }
                store8(dst, pack8(f));
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int col = n0 + c * 32 + j * 4;
              if (col < N) {
                float4 v = make_float4(__uint_as_float(r[j * 4]), __uint_as_float(r[j * 4 + 1]),
                                       __uint_as_float(r[j * 4 + 2]), __uint_as_float(r[j * 4 + 3]));
                float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(crow) + c * 32 + j * 4);
                if (beta) {
                  const float4 o = *dst;
                  v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
                }
                *dst = v;
              }
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&bar_tempty[acc], 0);   // always the leader's barrier
      acc ^= 1;
      if (acc == 0) aphase ^= 1u;
    }
  }

  tc_fence_before();
  cluster_sync();   // the leader's MMAs read the peer's smem: nobody leaves before everything retired
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<2>(tmem_base, Cfg::kTmemCols);
  }
}

static int g_num_sms = 0;
static int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      g_num_sms = n;
    else
      g_num_sms = kNumSMs;
  }
  return g_num_sms;
}

template <int BN, bool A_MN, bool B_MN, typename OutT>
static int launch_gemm(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                       long long ldb, long long ldc, int beta, int max_ctas, cudaStream_t s, GemmGate gate) {
  if (gate.flags) {   // tile-row permutation needs whole tile blocks per chunk and per slab
    gate.bpc = (M / gate.world) / kBM;
    gate.bps = gate.rows_per_flag / kBM;
    gate.perm = (gate.perm && (M / gate.world) % kBM == 0 && gate.rows_per_flag % kBM == 0 && M % gate.world == 0) ? 1 : 0;
  }
  using Cfg = GemmCfg<BN>;
  CUtensorMap tmA, tmB;
  int rc;
  if (!A_MN) rc = get_tmap_2d_bf16(&tmA, A, M, K, lda, kBM, kBK, 1);
  else rc = get_tmap_2d_bf16(&tmA, A, K, M, lda, kBK, 64, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map A");
  if (!B_MN) rc = get_tmap_2d_bf16(&tmB, B, N, K, ldb, BN, kBK, 1);
  else rc = get_tmap_2d_bf16(&tmB, B, K, N, ldb, kBK, 64, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map B");
  auto kern = gemm_tcgen05_kernel<BN, A_MN, B_MN, OutT>;
  LLMCTL_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::kSmemBytes));
  const int tiles = ((M + kBM - 1) / kBM) * ((N + BN - 1) / BN);
  int grid = num_sms();
  if (max_ctas > 0 && max_ctas < grid) grid = max_ctas;
  if (tiles < grid) grid = tiles;
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, s>>>(tmA, tmB, (OutT*)C, ldc, M, N, K, beta, gate);
  LLMCTL_RETURN_LAST();
}

template <int BN, typename OutT>
static int dispatch_layout(int a_mn, int b_mn, const void* A, const void* B, void* C, int M, int N,
                           int K, long long lda, long long ldb, long long ldc, int beta,
                           int max_ctas, cudaStream_t s, const GemmGate& g) {
  if (!a_mn && !b_mn) return launch_gemm<BN, false, false, OutT>(A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  if (!a_mn && b_mn) return launch_gemm<BN, false, true, OutT>(A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  if (a_mn && !b_mn) return launch_gemm<BN, true, false, OutT>(A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  return launch_gemm<BN, true, true, OutT>(A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
}


template <int PBN, bool A_MN, bool B_MN, typename OutT>
static int launch_gemm2(const void* A, const void* B, void* C, int M, int N, int K, long long lda,
                        long long ldb, long long ldc, int beta, int max_ctas, cudaStream_t s, GemmGate gate) {
  if (gate.flags) {
    gate.bpc = (M / gate.world) / (2 * kBM);
    gate.bps = gate.rows_per_flag / (2 * kBM);
    gate.perm = (gate.perm && (M / gate.world) % (2 * kBM) == 0 && gate.rows_per_flag % (2 * kBM) == 0 &&
                 M % gate.world == 0) ? 1 : 0;
  }
  using Cfg = Gemm2Cfg<PBN>;
  CUtensorMap tmA, tmB;
  int rc;
  if (!A_MN) rc = get_tmap_2d_bf16(&tmA, A, M, K, lda, kBM, kBK, 1);
  else rc = get_tmap_2d_bf16(&tmA, A, K, M, lda, kBK, 64, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map A");
  if (!B_MN) rc = get_tmap_2d_bf16(&tmB, B, N, K, ldb, Cfg::kBN / 2, kBK, 1);
  else rc = get_tmap_2d_bf16(&tmB, B, K, N, ldb, kBK, 64, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map B");
  auto kern = gemm2_tcgen05_kernel<PBN, A_MN, B_MN, OutT>;
  LLMCTL_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    Cfg::kSmemBytes));
  const int tiles = ((M + 2 * kBM - 1) / (2 * kBM)) * ((N + Cfg::kBN - 1) / Cfg::kBN);
  int pairs = num_sms() / 2;
  if (max_ctas > 0 && max_ctas / 2 < pairs) pairs = max_ctas / 2;
  if (pairs < 1) pairs = 1;
  if (tiles < pairs) pairs = tiles;
  kern<<<2 * pairs, kGemmThreads, Cfg::kSmemBytes, s>>>(tmA, tmB, (OutT*)C, ldc, M, N, K, beta, gate);
  LLMCTL_RETURN_LAST();
}

template <int PBN, typename OutT>
static int dispatch_layout2(int a_mn, int b_mn, const void* A, const void* B, void* C, int M, int N,
                            int K, long long lda, long long ldb, long long ldc, int beta,
                            int max_ctas, cudaStream_t s, const GemmGate& g) {
  if (!a_mn && !b_mn) return launch_gemm2<PBN, false, false, OutT>(A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  if (!a_mn && b_mn) return launch_gemm2<PBN, false, true, OutT>(A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  if (a_mn && !b_mn) return launch_gemm2<PBN, true, false, OutT>(A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  return launch_gemm2<PBN, true, true, OutT>(A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
}

}  // namespace llmctl

using namespace llmctl;

// D[M,N] = op(A) op(B) (+ D if beta).  a_mn / b_mn select the operand storage (see file header).
// out_fp32: D is fp32 (else bf16).  tile_n: 0 = auto, 128 / 192 / 256 (1-CTA tiles), 2 = 256x256 CTA-pair tile, 3 = 256x128 CTA-pair tile.  max_ctas: 0 = all SMs
// (a smaller value leaves SMs free for a concurrent communication kernel).
static int gemm_bf16_impl(const void* A, const void* B, void* C, int M, int N, int K,
                          long long lda, long long ldb, long long ldc, int a_mn, int b_mn,
                          int out_fp32, int beta, int tile_n, int max_ctas, void* stream, const GemmGate& g) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(M > 0 && N > 0 && K > 0);
  LLMCTL_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && N % 8 == 0);
  LLMCTL_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(C) & 15) == 0);
  LLMCTL_REQUIRE(out_fp32 ? (ldc % 4 == 0) : (ldc % 8 == 0));
  cudaStream_t s = (cudaStream_t)stream;
  // auto: the CTA-pair kernel wins whenever the problem has at least one full pair tile in each dimension
  // (measured on B200: 1.44-1.61 PFLOP/s vs 1.24-1.45 for the single-CTA 128x256 tile on the GPT-7B shapes)
  // auto: the 256 x 256 CTA-pair tile wins whenever the problem has at least one full pair tile in each dimension
  // (measured on B200: 1.44-1.61 PFLOP/s vs 1.24-1.45 for the single-CTA 128x256 tile on the GPT-7B shapes).  The
  // narrower tiles (192, pair 256 x 128) were added for the 1-3 wave problems of a TP=8 shard and measured SLOWER
  // there too (4096 x 1536 x 4096: 55 us pair-256, 53 us 192, 85 us pair-128, cuBLAS 42 us): the main loop costs about
  // the same per 64-deep k-block whatever the tile width (operand traffic from L2, not the tensor pipe, sets the pace),
  // so less work per tile only lowers the rate.  They stay selectable (tile_n = 192 / 3), never chosen automatically.
  if (tile_n == 0 && M > 128 && N > 128) tile_n = 2;
  if (tile_n == 2) {   // 2-CTA pair tile 256 x 256 (cta_group::2)
    if (out_fp32) return dispatch_layout2<256, float>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
    return dispatch_layout2<256, __nv_bfloat16>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  }
  if (tile_n == 3) {   // 2-CTA pair tile 256 x 128
    if (out_fp32) return dispatch_layout2<128, float>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
    return dispatch_layout2<128, __nv_bfloat16>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  }
  int bn = tile_n;
  if (bn != 128 && bn != 192 && bn != 256) {
    const int sms = num_sms();
    const long long t256 = static_cast<long long>((M + 127) / 128) * ((N + 255) / 256);
    const long long t128 = static_cast<long long>((M + 127) / 128) * ((N + 127) / 128);
    const double c256 = static_cast<double>((t256 + sms - 1) / sms) * 2.0;
    const double c128 = static_cast<double>((t128 + sms - 1) / sms) * 1.45;  // measured: a 128-wide tile runs at ~0.7x the rate
    bn = (c128 < c256) ? 128 : 256;
  }
  if (bn == 192) {
    if (out_fp32) return dispatch_layout<192, float>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
    return dispatch_layout<192, __nv_bfloat16>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  }
  if (bn == 256) {
    if (out_fp32) return dispatch_layout<256, float>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
    return dispatch_layout<256, __nv_bfloat16>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  }
  if (out_fp32) return dispatch_layout<128, float>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
  return dispatch_layout<128, __nv_bfloat16>(a_mn, b_mn, A, B, C, M, N, K, lda, ldb, ldc, beta, max_ctas, s, g);
}

LLMCTL_API int llmctl_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                                long long lda, long long ldb, long long ldc, int a_mn, int b_mn,
                                int out_fp32, int beta, int tile_n, int max_ctas, void* stream) {
  return gemm_bf16_impl(A, B, C, M, N, K, lda, ldb, ldc, a_mn, b_mn, out_fp32, beta, tile_n, max_ctas, stream,
                        GemmGate{});
}

LLMCTL_SYMM_WATCHDOG_MODULE(gemm)

// K1 consumer: the same GEMM, but rows of A (K-major, [M, K] in a symmetric buffer) outside [own_lo, own_hi) are only
// read once their flag  flags[row / rows_per_flag]  (local signal pad, written by the peers' llmctl_ag_push) reached
// `epoch`.  M = world * rows_per_rank; tile rows are visited in arrival order when the shapes allow it.
LLMCTL_API int llmctl_gemm_bf16_gated(const void* A, const void* B, void* C, int M, int N, int K,
                                      long long lda, long long ldb, long long ldc, int b_mn, int out_fp32,
                                      int beta, int tile_n, int max_ctas, const unsigned int* flags,
                                      unsigned int epoch, int rows_per_flag, int rank, int world,
                                      void* stream) {
  LLMCTL_REQUIRE(flags != nullptr && world >= 1 && world <= symm::kMaxRanks && M % world == 0 && rows_per_flag > 0 &&
                 (M / world) % rows_per_flag == 0 && M / rows_per_flag <= 256);
  GemmGate g{};
  g.flags = flags;
  g.epoch = epoch;
  g.rows_per_flag = rows_per_flag;
  g.own_lo = rank * (M / world);
  g.own_hi = g.own_lo + M / world;
  g.perm = 1;
  g.rank = rank;
  g.world = world;
  return gemm_bf16_impl(A, B, C, M, N, K, lda, ldb, ldc, 0, b_mn, out_fp32, beta, tile_n, max_ctas, stream, g);
}

// K8 consumer: the same GEMM, with the B operand (a ZeRO-3 parameter matrix inside a gathered symmetric buffer) read
// tile by tile as its shards arrive: flag  b_flags[(b_off + element) / b_slab]  (local signal pad, raised by the shard
// owners' llmctl_ag_push) must have reached `epoch` before the tile's rows are loaded.  ldb must equal the row pitch of
// B inside the buffer.
LLMCTL_API int llmctl_gemm_bf16_bgated(const void* A, const void* B, void* C, int M, int N, int K,
                                       long long lda, long long ldb, long long ldc, int a_mn, int b_mn,
                                       int out_fp32, int beta, int tile_n, int max_ctas,
                                       const unsigned int* b_flags, unsigned int epoch, long long b_off,
                                       long long b_slab, void* stream) {
  LLMCTL_REQUIRE(b_flags != nullptr && b_slab > 0 && b_off >= 0);
  GemmGate g{};
  g.b_flags = b_flags;
  g.b_epoch = epoch;
  g.b_off = b_off;
  g.b_slab = b_slab;
  g.b_row_stride = ldb;
  g.b_span = static_cast<long long>(K) * ldb;
  return gemm_bf16_impl(A, B, C, M, N, K, lda, ldb, ldc, a_mn, b_mn, out_fp32, beta, tile_n, max_ctas, stream, g);
}
