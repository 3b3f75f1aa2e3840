// K2: tensor/sequence-parallel row-parallel GEMM fused with its reduce-scatter over NVLink peer memory.
// (K1, the all-gather -> GEMM direction, is csrc/fused/ag_push.cu + the gated GEMM in csrc/gemm/gemm_tcgen05.cu.)
//
// K2  gemm_rs   : row-parallel GEMM -> sequence-parallel REDUCE-SCATTER in ONE kernel.
//     Partial sums D[M,N] = X[M,Kl] Wl^T are produced tile by tile; the epilogue converts each 128-row slab to
//     bf16 and stores it straight into the OWNER rank's staging slot  stage[src_rank][Ms, N]  with 16-byte
//     P2P stores while the tensor core is already working on the next tile (TMEM double buffering).  Tiles
//     are visited destination by destination (rank+1, ..., rank last); when the last tile for a destination has
//     been pushed, that destination's flag is raised.  The owner's follow-up kernel
//     (llmctl_rs_reduce_rmsnorm, below) sums the `world` slots in fp32 in rank order — deterministic, no
//     atomics — fused with the residual add and the RMSNorm that sits on the sequence shard (K3).
//     The weight may be K-major ([N, Kl], forward) or MN-major ([Kl, N] = the forward weight of a column-parallel
//     layer used untransposed for its dgrad).
//
// Reference: no TP/SP exists in the reference (planner integers only: reference
// llmctl/cli/commands/plan.py:79-80,110-113; `sequence_parallel` flag: configs/presets/llama-7b-a100x8.toml:36).
// Baseline to beat: cuBLAS + NCCL reduce-scatter, run back to back.
#include "../comm/symm.cuh"
#include "../common/tma_host.h"
#include "../common/util.cuh"

namespace llmctl {
using namespace ptx;
using namespace symm;

constexpr int kFBM = 128, kFBN = 256, kFBK = 64;
constexpr int kFStages = 4;
constexpr int kFABytes = kFBM * kFBK * 2;
constexpr int kFBBytes = kFBN * kFBK * 2;
constexpr int kFStageBytes = kFABytes + kFBBytes;
constexpr int kFSmem = kFStages * kFStageBytes + 1024;
constexpr int kFThreads = 192;
constexpr int kEpiRowBytes = 128 + 16;       // K2 epilogue staging: 64 bf16 per row + pad (conflict-free 16 B access)
constexpr int kEpiWarpBytes = 32 * kEpiRowBytes;
constexpr int kFSmemRS = kFSmem + 4 * kEpiWarpBytes;

struct FusedArgs {
  int M, N, K;
  int rank, world;
  int rows_per_rank;          // Ms = M / world
  uint32_t epoch;
};

// --------------------------------------------------------------------------------------------------
// GEMM core (1-CTA, 128 x 256 tile) whose epilogue pushes bf16 tiles to the owner's staging slot.
// Tiles are enumerated chunk-major: chunk i covers the M rows owned by rank c_i = (rank + i + 1) % world.
// --------------------------------------------------------------------------------------------------
template <bool B_MN>
__global__ void __launch_bounds__(kFThreads, 1)
gemm_rs_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, FusedArgs a,
               PeerTable xg /* staging buffers */, PeerTable pads, uint32_t* __restrict__ local_flags /* counters */) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar_full[kFStages];
  __shared__ uint64_t bar_empty[kFStages];
  __shared__ uint64_t bar_tfull[2];
  __shared__ uint64_t bar_tempty[2];
  __shared__ uint32_t tmem_base_smem;

  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int Ms = a.rows_per_rank;

  // ================================ GEMM CTAs ================================
  const int cta = static_cast<int>(blockIdx.x);
  const int n_cta = static_cast<int>(gridDim.x);
  const int m_per_chunk = (Ms + kFBM - 1) / kFBM;
  const int num_n = (a.N + kFBN - 1) / kFBN;
  const int tiles_per_chunk = m_per_chunk * num_n;
  const int num_tiles = tiles_per_chunk * a.world;
  const int num_kb = (a.K + kFBK - 1) / kFBK;
  // Each CTA takes a CONTIGUOUS range of the chunk-major tile list: its tiles share a destination (and a B panel), so
  // an epilogue warp publishes once per destination instead of once per tile — the system-scope fence in front of
  // the publication waits for the warp's outstanding NVLink stores and used to stall the epilogue after every tile —
  // and at any moment the CTAs together feed all destinations, not one.
  const int t_begin = static_cast<int>(static_cast<long long>(cta) * num_tiles / n_cta);
  const int t_end = static_cast<int>(static_cast<long long>(cta + 1) * num_tiles / n_cta);

  if (threadIdx.x == 0) {
    for (int i = 0; i < kFStages; ++i) {
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
  if (warp == 1) tmem_alloc<1>(&tmem_base_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  // tile -> (chunk rank, m0 in global rows, n0)
  auto decode = [&](int tile, int& crank, int& m0, int& n0, int& m_end) {
    const int ci = tile / tiles_per_chunk;
    const int t = tile % tiles_per_chunk;
    crank = (a.rank + ci + 1) % a.world;
    m0 = crank * Ms + (t % m_per_chunk) * kFBM;
    n0 = (t / m_per_chunk) * kFBN;
    m_end = (crank + 1) * Ms;
  };

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = t_begin; tile < t_end; ++tile) {
        int crank, m0, n0, m_end;
        decode(tile, crank, m0, n0, m_end);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&bar_empty[stage], phase ^ 1u);
          uint8_t* sA = smem + stage * kFStageBytes;
          uint8_t* sB = sA + kFABytes;
          mbar_expect_tx(&bar_full[stage], kFStageBytes);
          tma_load_2d(sA, &tmA, &bar_full[stage], kb * kFBK, m0);
          if (!B_MN) {
            tma_load_2d(sB, &tmB, &bar_full[stage], kb * kFBK, n0);
          } else {
#pragma unroll
            for (int g = 0; g < kFBN / 64; ++g)
              tma_load_2d(sB + g * 8192, &tmB, &bar_full[stage], n0 + g * 64, kb * kFBK);
          }
          if (++stage == kFStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kFBM, kFBN, 0u, B_MN ? 1u : 0u);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t aphase = 0;
      for (int tile = t_begin; tile < t_end; ++tile) {
        mbar_wait(&bar_tempty[acc], aphase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * kFBN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&bar_full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_base + stage * kFStageBytes;
          const uint32_t b_addr = a_addr + kFABytes;
#pragma unroll
          for (int k = 0; k < kFBK / 16; ++k)
            umma_ss<1>(d_tmem, make_smem_desc(a_addr + k * 32, 16, 1024, 2),
                       B_MN ? make_smem_desc(b_addr + k * 2048, 8192, 1024, 2)
                            : make_smem_desc(b_addr + k * 32, 16, 1024, 2),
                       idesc, (kb | k) != 0 ? 1u : 0u);
          umma_commit<1>(&bar_empty[stage]);
          if (++stage == kFStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit<1>(&bar_tfull[acc]);
        acc ^= 1;
        if (acc == 0) aphase ^= 1u;
      }
    }
  } else {
    const int q = warp & 3;
    int acc = 0;
    uint32_t aphase = 0;
    int cur_dest = -1;
    uint32_t unpublished = 0;       // tiles of cur_dest whose stores this warp has issued but not yet announced
    // all of this warp's remote stores -> system fence -> warp converges -> lane 0 counts them in; the last
    // (tile, warp) pair of a destination raises that destination's flag
    auto publish = [&]() {
      if (unpublished == 0) return;
      __threadfence_system();
      __syncwarp();
      if (lane == 0) {
        const uint32_t done = atomicAdd(&local_flags[cur_dest], unpublished) + unpublished;
        if (done == static_cast<uint32_t>(tiles_per_chunk * 4)) {
          atomicExch(&local_flags[cur_dest], 0u);
          signal_set(reinterpret_cast<uint32_t*>(pads.ptr[cur_dest]) + 8 + a.rank, a.epoch);
        }
      }
      unpublished = 0;
    };
    for (int tile = t_begin; tile < t_end; ++tile) {
      int crank, m0, n0, m_end;
      decode(tile, crank, m0, n0, m_end);
      if (crank != cur_dest) {
        publish();
        cur_dest = crank;
      }
      mbar_wait(&bar_tfull[acc], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * kFBN);
      {
        // Remote stores must be wide: stage 64 columns of the warp's 32 rows in shared memory (row = lane), then
        // let 8 consecutive lanes write one row's 128 contiguous bytes -> full-size NVLink write packets.
        uint8_t* st = smem + kFStages * kFStageBytes + q * kEpiWarpBytes;
        __nv_bfloat16* owner_base = reinterpret_cast<__nv_bfloat16*>(xg.ptr[crank]) +
                                    static_cast<long long>(a.rank) * Ms * a.N;
#pragma unroll 1
        for (int hc = 0; hc < kFBN / 64; ++hc) {
#pragma unroll
          for (int c2 = 0; c2 < 2; ++c2) {
            uint32_t r[32];
            tmem_ld_x32(t_row + hc * 64 + c2 * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(r[j * 8 + e]);
              const bf16x8 pk = pack8(f);
              *reinterpret_cast<uint4*>(st + lane * kEpiRowBytes + (c2 * 4 + j) * 16) = *reinterpret_cast<const uint4*>(&pk);
            }
          }
          __syncwarp();
          const int seg = lane & 7;
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + (lane >> 3);
            const int grow = m0 + q * 32 + rr;
            const int col = n0 + hc * 64 + seg * 8;
            if (grow < m_end && grow < a.M && col < a.N) {
              const uint4 v = *reinterpret_cast<const uint4*>(st + rr * kEpiRowBytes + seg * 16);
              st_v4(owner_base + static_cast<long long>(grow - crank * Ms) * a.N + col, v);
            }
          }
          __syncwarp();
        }
      }
      tc_fence_before();
      __syncwarp();
      ++unpublished;
      if (lane == 0) mbar_arrive(&bar_tempty[acc]);
      acc ^= 1;
      if (acc == 0) aphase ^= 1u;
    }
    publish();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

// --------------------------------------------------------------------------------------------------
// K2 consumer / K3 on the sequence shard: out rows = sum over source ranks of stage[s] (+ residual), RMSNorm.
// One warp per row (cols <= 8192), fp32 accumulation in rank order.
// --------------------------------------------------------------------------------------------------
template <int kMaxVec>
__global__ void __launch_bounds__(128)
rs_reduce_rmsnorm_kernel(const __nv_bfloat16* __restrict__ stage, PeerTable pads, int rank, int world,
                         const __nv_bfloat16* __restrict__ residual, const __nv_bfloat16* __restrict__ w,
                         __nv_bfloat16* __restrict__ h_out, __nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out,
                         int rows, int cols, float eps, uint32_t epoch) {
  // wait until every source rank has pushed all its tiles for us
  if (threadIdx.x < world)
    signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + 8 + threadIdx.x, epoch);
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nvec = cols >> 3;
  const long long slot = static_cast<long long>(rows) * cols;
  float acc[kMaxVec][8];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxVec; ++i) {
    const int v = i * 32 + lane;
    if (v < nvec) {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
      for (int s = 0; s < world; ++s) {
        float f[8];
        unpack8(load8(stage + s * slot + static_cast<long long>(row) * cols + v * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] += f[j];
      }
      if (residual != nullptr) {
        float f[8];
        unpack8(load8(residual + static_cast<long long>(row) * cols + v * 8), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] += f[j];
      }
      const bf16x8 hq = pack8(acc[i]);
      if (h_out != nullptr) store8(h_out + static_cast<long long>(row) * cols + v * 8, hq);
      unpack8(hq, acc[i]);
#pragma unroll
      for (int j = 0; j < 8; ++j) ss += acc[i][j] * acc[i][j];
    }
  }
  ss = warp_sum(ss);
  const float rstd = rsqrtf(ss / cols + eps);
  if (lane == 0 && rstd_out) rstd_out[row] = rstd;
  if (y == nullptr) return;
#pragma unroll
  for (int i = 0; i < kMaxVec; ++i) {
    const int v = i * 32 + lane;
    if (v < nvec) {
      float g[8], o[8];
      unpack8(load8(w + v * 8), g);
#pragma unroll
      for (int j = 0; j < 8; ++j) o[j] = acc[i][j] * rstd * g[j];
      store8(y + static_cast<long long>(row) * cols + v * 8, pack8(o));
    }
  }
}

static int fused_num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = kNumSMs;
  }
  return n;
}

}  // namespace llmctl

using namespace llmctl;
LLMCTL_SYMM_WATCHDOG_MODULE(tp_fused_gemm)

// K2.  X: [M, K] local activation (K = local shard of the contraction dim).  W: local weight shard, [N, K] (ldw = row
// stride) or, with w_mn = 1, stored [K, N] (MN-major: a column-parallel forward weight reused for its dgrad).
// stage: symmetric staging buffer [world, Ms, N] bf16 on every rank (slot s receives rank s's partial rows).
// pads slots [8, 8 + world) are used.  local_flags: 32 zero-initialised uint32 (tile counters).
LLMCTL_API int llmctl_gemm_rs(const void* X, const void* W, llmctl::symm::PeerTable stage, int M, int N, int K,
                              long long ldx, long long ldw, int w_mn, llmctl::symm::PeerTable pads,
                              unsigned int* local_flags, int rank, int world, unsigned int epoch,
                              void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && M % world == 0 && K % 8 == 0 && N % 8 == 0);
  CUtensorMap tmA, tmB;
  int rc = get_tmap_2d_bf16(&tmA, X, M, K, ldx, kFBM, kFBK, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map A");
  if (!w_mn) rc = get_tmap_2d_bf16(&tmB, W, N, K, ldw, kFBN, kFBK, 1);
  else rc = get_tmap_2d_bf16(&tmB, W, K, N, ldw, kFBK, 64, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map B");
  FusedArgs a{M, N, K, rank, world, M / world, epoch};
  const int Ms = M / world;
  const int tiles = ((Ms + kFBM - 1) / kFBM) * ((N + kFBN - 1) / kFBN) * world;
  int grid = fused_num_sms();
  if (tiles < grid) grid = tiles;
  if (w_mn) {
    auto kern = gemm_rs_kernel<true>;
    LLMCTL_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFSmemRS));
    kern<<<grid, kFThreads, kFSmemRS, (cudaStream_t)stream>>>(tmA, tmB, a, stage, pads, local_flags);
  } else {
    auto kern = gemm_rs_kernel<false>;
    LLMCTL_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFSmemRS));
    kern<<<grid, kFThreads, kFSmemRS, (cudaStream_t)stream>>>(tmA, tmB, a, stage, pads, local_flags);
  }
  LLMCTL_RETURN_LAST();
}

// out[rows, cols] = sum_s stage[s] (+ residual); y = rmsnorm(out) * w.  rows = Ms.  Waits for the K2 flags of
// `epoch`.  h_out / residual / y / rstd may be NULL (y == NULL: reduce only).
LLMCTL_API int llmctl_rs_reduce_rmsnorm(const void* stage_local, llmctl::symm::PeerTable pads, int rank,
                                        int world, const void* residual, const void* w, void* h_out, void* y,
                                        float* rstd, int rows, int cols, float eps, unsigned int epoch,
                                        void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(cols % 8 == 0 && cols <= 8192 && world >= 1 && world <= symm::kMaxRanks);
  if (rows == 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  const int per_lane = ((cols >> 3) + 31) / 32;
  dim3 grid((rows + 3) / 4);
#define LAUNCH(MV)                                                                                          \
  rs_reduce_rmsnorm_kernel<MV><<<grid, 128, 0, s>>>((const __nv_bfloat16*)stage_local, pads, rank, world,   \
                                                    (const __nv_bfloat16*)residual, (const __nv_bfloat16*)w, \
                                                    (__nv_bfloat16*)h_out, (__nv_bfloat16*)y, rstd, rows, cols, eps, epoch)
  if (per_lane <= 4) LAUNCH(4);
  else if (per_lane <= 8) LAUNCH(8);
  else if (per_lane <= 16) LAUNCH(16);
  else LAUNCH(32);
#undef LAUNCH
  LLMCTL_RETURN_LAST();
}
