// K6 (backward), second generation: two deterministic tcgen05 kernels instead of one kernel with atomics.
//
//   dKV kernel  — one CTA per (batch, head, 128-key block): K_j, V_j resident in shared memory, query sub-blocks
//                 of 64 rows streamed through a 4-stage TMA ring.  It works on the TRANSPOSED scores,
//                     S^T = K_j Q_i^T      dP^T = V_j dO_i^T        (M = 128 keys in TMEM lanes, N = 64 queries)
//                 so that P^T and dS^T — written back by the softmax threads as bf16 over the first half of the
//                 very TMEM columns they were read from — are directly the A operands of
//                     dV += P^T dO_i       dK += dS^T Q_i
//                 Nothing but Q/dO tiles and 512 bytes of per-query statistics ever goes through shared memory.
//   dQ kernel   — one CTA per (batch, head, 128-query block): Q_i, dO_i resident, key sub-blocks of 64 streamed;
//                     S = Q_i K_j^T    dP = dO_i V_j^T    dS -> TMEM    dQ += dS K_j
// Both kernels double-buffer S / dP in TMEM and alternate two softmax warpgroups over the sub-blocks, so the
// tensor pipe computes sub-block n+1/n+2 while sub-block n is in the softmax.  Recomputing S and dP in the
// second kernel costs 7 GEMMs instead of 5, but removes the fp32 atomic accumulation of dQ — which ran at the
// L2 atomic rate, ~3x slower than the tensor work — together with the accumulator clear and cast passes, and
// makes the gradient bit-reproducible.  The 1/sqrt(d) factor of dS is applied once in the dK / dQ epilogues.
//
// Reference: autograd through HF attention (reference llmctl/runtime/engine.py:288-291).  Baseline: cuDNN SDPA bwd.
#include "../common/ptx.cuh"
#include "../common/tma_host.h"
#include "../common/util.cuh"

namespace llmctl {
using namespace ptx;

namespace bwd2 {
constexpr int kD = 128;
constexpr int kRes = 128;                    // resident rows (keys for dKV, queries for dQ)
constexpr int kSub = 64;                     // streamed rows per sub-block
constexpr int kResPanel = kRes * 128;        // bytes of a [128 x 64 bf16] swizzled panel
constexpr int kResTile = 2 * kResPanel;      // 32 KB
constexpr int kSubPanel = kSub * 128;        // 8 KB
constexpr int kSubTile = 2 * kSubPanel;      // 16 KB
constexpr int kStages = 4;
constexpr int kStatBytes = kSub * 8;         // (−lse·log2e, −delta) per streamed query
constexpr int kStageBytes = 2 * kSubTile + 1024;
constexpr int kSmem = 2 * kResTile + kStages * kStageBytes + 1024;
constexpr int kThreads = 320;
}  // namespace bwd2

// stats[b,h,s] = (−lse·log2(e), −sum_d dO·O).  One warp per (token, head).
__global__ void __launch_bounds__(256)
flash_bwd2_prep_kernel(const __nv_bfloat16* __restrict__ O, long long o_tok, const __nv_bfloat16* __restrict__ dO,
                       long long do_tok, const float* __restrict__ lse, float2* __restrict__ stats, int B, int S, int H) {
  const int lane = threadIdx.x & 31;
  const long long w = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const long long total = static_cast<long long>(B) * S * H;
  if (w >= total) return;
  const int h = static_cast<int>(w % H);
  const long long tok = w / H;
  const uint2 ou = *reinterpret_cast<const uint2*>(O + tok * o_tok + h * bwd2::kD + lane * 4);
  const uint2 du = *reinterpret_cast<const uint2*>(dO + tok * do_tok + h * bwd2::kD + lane * 4);
  const float2 o0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ou.x));
  const float2 o1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ou.y));
  const float2 d0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&du.x));
  const float2 d1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&du.y));
  float s = o0.x * d0.x + o0.y * d0.y + o1.x * d1.x + o1.y * d1.y;
  s = warp_sum(s);
  if (lane == 0) {
    const int b = static_cast<int>(tok / S), sidx = static_cast<int>(tok % S);
    const long long i = (static_cast<long long>(b) * H + h) * S + sidx;
    stats[i] = make_float2(-lse[i] * 1.4426950408889634f, -s);
  }
}

// kDQ = false: dKV kernel (resident K,V; streamed Q,dO).   kDQ = true: dQ kernel (resident Q,dO; streamed K,V).
// tmRes0/1: 128-row boxes of the resident operands; tmSub0/1: 64-row boxes of the streamed operands, where
// operand 0 pairs up in the S-like product and operand 1 in the dP-like product.
// tmOut0/1: outputs staged through the resident tiles (dKV: dV from acc0, dK from acc1;  dQ: dQ from acc0).
template <bool kDQ>
__global__ void __launch_bounds__(bwd2::kThreads, 1)
flash_bwd2_kernel(const __grid_constant__ CUtensorMap tmRes0, const __grid_constant__ CUtensorMap tmRes1,
                  const __grid_constant__ CUtensorMap tmSub0, const __grid_constant__ CUtensorMap tmSub1,
                  const __grid_constant__ CUtensorMap tmOut0, const __grid_constant__ CUtensorMap tmOut1,
                  const float2* __restrict__ stats, int S, int H, float scale_log2, float scale, int causal) {
  using namespace bwd2;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar_res, bar_full[kStages], bar_empty[kStages], bar_s[2], bar_p[2], bar_done, bar_tres;
  __shared__ uint32_t tmem_base_smem;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  uint8_t* sRes0 = smem;
  uint8_t* sRes1 = smem + kResTile;
  uint8_t* sRing = smem + 2 * kResTile;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int head = blockIdx.y, b = blockIdx.z;
  const int tok0 = b * S;
  // dKV: key block j, streams queries [sub_begin, S).   dQ: query block i (heavy blocks first), streams keys [0, sub_end)
  const int rblk = kDQ ? static_cast<int>(gridDim.x - 1 - blockIdx.x) : static_cast<int>(blockIdx.x);
  const int r0 = rblk * kRes;
  const int sub_begin = kDQ ? 0 : (causal ? r0 : 0);
  const int sub_end = kDQ ? (causal ? r0 + kRes : S) : S;
  const int nsub = (sub_end - sub_begin) / kSub;

  if (threadIdx.x == 0) {
    mbar_init(&bar_res, 1);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&bar_full[i], 1);
      mbar_init(&bar_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_s[i], 1);
      mbar_init(&bar_p[i], 4);
    }
    mbar_init(&bar_done, 1);
    mbar_init(&bar_tres, 8);
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmRes0);
    prefetch_tmap(&tmRes1);
    prefetch_tmap(&tmSub0);
    prefetch_tmap(&tmSub1);
    prefetch_tmap(&tmOut0);
    if (!kDQ) prefetch_tmap(&tmOut1);
  }
  if (warp == 1) tmem_alloc<1>(&tmem_base_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  // TMEM: S buffers [0,64) [64,128);  dP buffers [128,192) [192,256);  acc0 [256,384);  acc1 [384,512)
  const uint32_t tS = tmem_base, tdP = tmem_base + 128, tAcc0 = tmem_base + 256, tAcc1 = tmem_base + 384;
  // dQ kernel only (acc1 is unused there): the resident Q and dO tiles are copied once into TMEM as bf16x2
  // A operands ([384,448) and [448,512)), so S = Q K^T and dP = dO V^T read only their 2 KB B operand from
  // shared memory per MMA (64 B/clk) instead of 6 KB (192 B/clk, the operand-bandwidth limit of N = 64 tiles)
  const uint32_t tRes0 = tmem_base + 384, tRes1 = tmem_base + 448;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(&bar_res, 2 * kResTile);
      tma_load_3d(sRes0, &tmRes0, &bar_res, 0, head, tok0 + r0);
      tma_load_3d(sRes0 + kResPanel, &tmRes0, &bar_res, 64, head, tok0 + r0);
      tma_load_3d(sRes1, &tmRes1, &bar_res, 0, head, tok0 + r0);
      tma_load_3d(sRes1 + kResPanel, &tmRes1, &bar_res, 64, head, tok0 + r0);
      for (int n = 0; n < nsub; ++n) {
        const int st = n % kStages;
        const uint32_t ph = (n / kStages) & 1u;
        uint8_t* stage = sRing + st * kStageBytes;
        const int row = sub_begin + n * kSub;
        mbar_wait(&bar_empty[st], ph ^ 1u);
        mbar_expect_tx(&bar_full[st], 2 * kSubTile + (kDQ ? 0 : kStatBytes));
        tma_load_3d(stage, &tmSub0, &bar_full[st], 0, head, tok0 + row);
        tma_load_3d(stage + kSubPanel, &tmSub0, &bar_full[st], 64, head, tok0 + row);
        tma_load_3d(stage + kSubTile, &tmSub1, &bar_full[st], 0, head, tok0 + row);
        tma_load_3d(stage + kSubTile + kSubPanel, &tmSub1, &bar_full[st], 64, head, tok0 + row);
        if (!kDQ)
          bulk_load_1d(stage + 2 * kSubTile, stats + (static_cast<long long>(b) * H + head) * S + row, kStatBytes,
                       &bar_full[st]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t id_s = make_idesc_bf16(kRes, kSub, 0u, 0u);     // A, B K-major, N = 64
      constexpr uint32_t id_acc = make_idesc_bf16(kRes, kD, 0u, 1u);     // A from TMEM, B MN-major, N = 128
      const uint32_t aRes0 = smem_base, aRes1 = smem_base + kResTile, aRing = smem_base + 2 * kResTile;
      auto kdesc_res = [](uint32_t base, int ks) {
        return make_smem_desc(base + (ks >> 2) * kResPanel + (ks & 3) * 32, 16, 1024, 2);
      };
      auto kdesc_sub = [](uint32_t base, int ks) {
        return make_smem_desc(base + (ks >> 2) * kSubPanel + (ks & 3) * 32, 16, 1024, 2);
      };
      auto mdesc_sub = [](uint32_t base, int ks) {       // [64 k-rows x 128] MN-major: 16 k-rows = 2048 B, panels 8 KB apart
        return make_smem_desc(base + ks * 2048, kSubPanel, 1024, 2);
      };
      auto issue_s = [&](int n) {
        const int st = n % kStages, buf = n & 1;
        mbar_wait(&bar_full[st], (n / kStages) & 1u);
        tc_fence_after();
        const uint32_t a0 = aRing + st * kStageBytes, a1 = a0 + kSubTile;
        if (kDQ) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_ts<1>(tS + buf * 64, tRes0 + ks * 8, kdesc_sub(a0, ks), id_s, ks != 0 ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_ts<1>(tdP + buf * 64, tRes1 + ks * 8, kdesc_sub(a1, ks), id_s, ks != 0 ? 1u : 0u);
        } else {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_ss<1>(tS + buf * 64, kdesc_res(aRes0, ks), kdesc_sub(a0, ks), id_s, ks != 0 ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)
            umma_ss<1>(tdP + buf * 64, kdesc_res(aRes1, ks), kdesc_sub(a1, ks), id_s, ks != 0 ? 1u : 0u);
        }
        umma_commit<1>(&bar_s[buf]);
      };
      mbar_wait(&bar_res, 0);
      if (kDQ) {
        mbar_wait(&bar_tres, 0);         // Q and dO have been copied into TMEM by the softmax warps
        tc_fence_after();
      }
      issue_s(0);
      if (nsub > 1) issue_s(1);
      for (int n = 0; n < nsub; ++n) {
        const int st = n % kStages, buf = n & 1;
        mbar_wait(&bar_p[buf], (n >> 1) & 1u);
        tc_fence_after();
        const uint32_t a0 = aRing + st * kStageBytes, a1 = a0 + kSubTile;
        if (!kDQ) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)       // dV += P^T dO
            umma_ts<1>(tAcc0, tS + buf * 64 + ks * 8, mdesc_sub(a1, ks), id_acc, (n | ks) != 0 ? 1u : 0u);
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)       // dK += dS^T Q
            umma_ts<1>(tAcc1, tdP + buf * 64 + ks * 8, mdesc_sub(a0, ks), id_acc, (n | ks) != 0 ? 1u : 0u);
        } else {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)       // dQ += dS K
            umma_ts<1>(tAcc0, tdP + buf * 64 + ks * 8, mdesc_sub(a0, ks), id_acc, (n | ks) != 0 ? 1u : 0u);
        }
        umma_commit<1>(&bar_empty[st]);
        if (n + 2 < nsub) issue_s(n + 2);
      }
      umma_commit<1>(&bar_done);
    }
  } else {
    // ------------------------- softmax / dS warpgroups (sub-blocks of parity wg) -------------------------
    const int wg = (warp - 2) >> 2;
    const int q4 = warp & 3;
    const int row = q4 * 32 + lane;                       // resident row == TMEM lane
    const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
    float nl2 = 0.f, ndl = 0.f;                           // dQ kernel: this query row's statistics
    if (kDQ) {
      const float2 sv = stats[(static_cast<long long>(b) * H + head) * S + r0 + row];
      nl2 = sv.x;
      ndl = sv.y;
      // warpgroup 0 moves Q, warpgroup 1 moves dO: this thread's row, 16 swizzled 16-byte chunks -> 64 TMEM columns
      mbar_wait(&bar_res, 0);
      const uint8_t* src = (wg == 0 ? sRes0 : sRes1) + row * 128;
      const uint32_t dst = (wg == 0 ? tRes0 : tRes1) + lane_off;
#pragma unroll
      for (int pnl = 0; pnl < 2; ++pnl) {
#pragma unroll
        for (int c2 = 0; c2 < 4; ++c2) {
          uint32_t w[8];
          const uint4 lo = *reinterpret_cast<const uint4*>(src + pnl * kResPanel + (((2 * c2) ^ (row & 7)) << 4));
          const uint4 hi = *reinterpret_cast<const uint4*>(src + pnl * kResPanel + (((2 * c2 + 1) ^ (row & 7)) << 4));
          w[0] = lo.x; w[1] = lo.y; w[2] = lo.z; w[3] = lo.w;
          w[4] = hi.x; w[5] = hi.y; w[6] = hi.z; w[7] = hi.w;
          tmem_st_x8(dst + pnl * 32 + c2 * 8, w);
        }
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_tres);
    }
    for (int n = wg; n < nsub; n += 2) {
      const int st = n % kStages;
      const int c0 = sub_begin + n * kSub;                // first streamed row of this sub-block
      // dKV: keys r0+row vs queries c0+col, masked where key > query.  dQ: query r0+row vs keys c0+col.
      const bool diag = causal && (kDQ ? (c0 + kSub - 1 > r0) : (c0 < r0 + kRes - 1));
      const float4* st4 = reinterpret_cast<const float4*>(sRing + st * kStageBytes + 2 * kSubTile);
      if (!kDQ) mbar_wait(&bar_full[st], (n / kStages) & 1u);      // statistics of the streamed queries
      mbar_wait(&bar_s[wg], (n >> 1) & 1u);
      tc_fence_after();
      const uint32_t tSb = tS + wg * 64 + lane_off, tPb = tdP + wg * 64 + lane_off;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t rs[32], rp[32];
        tmem_ld_x32(tSb + half * 32, rs);
        tmem_ld_x32(tPb + half * 32, rp);
        tmem_ld_wait();
        uint32_t pk[16], dk[16];
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          float a0, a1, t0, t1, d0, d1;
          if (kDQ) {
            fma2(a0, a1, __uint_as_float(rs[e]), __uint_as_float(rs[e + 1]), scale_log2, nl2);
            add2v(t0, t1, __uint_as_float(rp[e]), __uint_as_float(rp[e + 1]), ndl, ndl);
          } else {
            const float4 sv = st4[(half * 32 + e) >> 1];          // (-lse2, -delta) of two queries, warp-uniform
            fma2v(a0, a1, __uint_as_float(rs[e]), __uint_as_float(rs[e + 1]), scale_log2, sv.x, sv.z);
            add2v(t0, t1, __uint_as_float(rp[e]), __uint_as_float(rp[e + 1]), sv.y, sv.w);
          }
          float p0 = ex2_approx(a0), p1 = ex2_approx(a1);
          if (diag) {
            const int col = c0 + half * 32 + e;
            const int rr = r0 + row;
            if (kDQ ? (col > rr) : (rr > col)) p0 = 0.f;
            if (kDQ ? (col + 1 > rr) : (rr > col + 1)) p1 = 0.f;
          }
          mul2v(d0, d1, p0, p1, t0, t1);
          if (!kDQ) pk[e >> 1] = pack_bf16x2(p0, p1);
          dk[e >> 1] = pack_bf16x2(d0, d1);
        }
        if (!kDQ) tmem_st_x16(tSb + half * 16, pk);
        tmem_st_x16(tPb + half * 16, dk);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_p[wg]);
    }
    // ---- epilogue: accumulators -> bf16 -> swizzled staging (the idle resident tiles) -> TMA store ----
    // dKV: warpgroup 0 writes dV (acc0), warpgroup 1 writes scale * dK (acc1).  dQ: both halves of acc0 by wg 0/1.
    mbar_wait(&bar_done, 0);
    tc_fence_after();
    {
      const uint32_t tA = (kDQ ? tAcc0 : (wg == 0 ? tAcc0 : tAcc1)) + lane_off;
      const float mul = kDQ ? scale : (wg == 0 ? 1.f : scale);
      uint8_t* stage = kDQ ? sRes0 : (wg == 0 ? sRes1 : sRes0);
      const CUtensorMap* tm = kDQ ? &tmOut0 : (wg == 0 ? &tmOut0 : &tmOut1);
      const int c_begin = kDQ ? wg * 2 : 0, c_end = kDQ ? wg * 2 + 2 : 4;      // 32-column chunks
#pragma unroll 1
      for (int c = c_begin; c < c_end; ++c) {
        uint32_t r[32];
        tmem_ld_x32(tA + c * 32, r);
        tmem_ld_wait();
        uint8_t* prow = stage + (c >> 1) * kResPanel + row * 128;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          uint32_t w[4];
#pragma unroll
          for (int e = 0; e < 4; ++e)
            w[e] = pack_bf16x2(__uint_as_float(r[ch * 8 + 2 * e]) * mul, __uint_as_float(r[ch * 8 + 2 * e + 1]) * mul);
          const int chunk = (c & 1) * 4 + ch;
          *reinterpret_cast<uint4*>(prow + ((chunk ^ (row & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (kDQ) {
          tma_store_3d(tm, stage + wg * kResPanel + q4 * 32 * 128, wg * 64, head, tok0 + r0 + q4 * 32);
        } else {
          tma_store_3d(tm, stage + q4 * 32 * 128, 0, head, tok0 + r0 + q4 * 32);
          tma_store_3d(tm, stage + kResPanel + q4 * 32 * 128, 64, head, tok0 + r0 + q4 * 32);
        }
        tma_store_commit();
        tma_store_wait_read<0>();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

// Launch both kernels.  scratch: fp32, at least 2*B*H*S elements (the per-row statistics).
int flash_bwd_v2_launch(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                        void* dq, void* dk, void* dv, float* scratch, int B, int S, int H, long long q_tok,
                        long long k_tok, long long v_tok, long long o_tok, long long dqkv_tok, float scale, int causal,
                        cudaStream_t s) {
  using namespace bwd2;
  const long long n_tok = static_cast<long long>(B) * S;
  float2* stats = reinterpret_cast<float2*>(scratch);
  {
    const long long blocks = (n_tok * H * 32 + 255) / 256;
    flash_bwd2_prep_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(
        (const __nv_bfloat16*)o, o_tok, (const __nv_bfloat16*)d_o, o_tok, lse, stats, B, S, H);
  }
  const uint64_t rows = static_cast<uint64_t>(n_tok);
  CUtensorMap q128, k128, v128, do128, q64, k64, v64, do64, tdq, tdk, tdv;
  int rc = 0;
  auto mk = [&](CUtensorMap* m, const void* base, long long tok, uint32_t box_rows) {
    if (!rc) rc = get_tmap_3d_bf16(m, base, kD, H, rows, kD, tok, 64, 1, box_rows, 1);
  };
  mk(&q128, q, q_tok, 128);
  mk(&k128, k, k_tok, 128);
  mk(&v128, v, v_tok, 128);
  mk(&do128, d_o, o_tok, 128);
  mk(&q64, q, q_tok, 64);
  mk(&k64, k, k_tok, 64);
  mk(&v64, v, v_tok, 64);
  mk(&do64, d_o, o_tok, 64);
  mk(&tdq, dq, dqkv_tok, 32);
  mk(&tdk, dk, dqkv_tok, 32);
  mk(&tdv, dv, dqkv_tok, 32);
  if (rc) return rc;
  cudaError_t e = cudaFuncSetAttribute(flash_bwd2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  if (e != cudaSuccess) return static_cast<int>(e);
  e = cudaFuncSetAttribute(flash_bwd2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmem);
  if (e != cudaSuccess) return static_cast<int>(e);
  const float sl2 = scale * 1.4426950408889634f;
  dim3 grid(S / kRes, H, B);
  // dKV: S^T = K Q^T (res0 = K, sub0 = Q), dP^T = V dO^T (res1 = V, sub1 = dO); outputs dV (acc0), dK (acc1)
  flash_bwd2_kernel<false><<<grid, kThreads, kSmem, s>>>(k128, v128, q64, do64, tdv, tdk, stats, S, H, sl2, scale, causal);
  // dQ: S = Q K^T (res0 = Q, sub0 = K), dP = dO V^T (res1 = dO, sub1 = V); output dQ (acc0)
  flash_bwd2_kernel<true><<<grid, kThreads, kSmem, s>>>(q128, do128, k64, v64, tdq, tdq, stats, S, H, sl2, scale, causal);
  return static_cast<int>(cudaGetLastError());
}

}  // namespace llmctl

using namespace llmctl;

// q,k,v,o,do: [B,S,H,128] bf16 views (token strides given in elements; do and o share o_tok_stride);
// dq,dk,dv: outputs with token stride dqkv_tok_stride; lse: fp32 [B,H,S] from the forward;
// scratch: fp32 [B*S*H*128 + B*H*S] (only the first 2*B*H*S floats are used: per-row log-sum-exp and delta).
LLMCTL_API int llmctl_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                                     const float* lse, void* dq, void* dk, void* dv, float* scratch, int B, int S,
                                     int H, int D, long long q_tok_stride, long long k_tok_stride,
                                     long long v_tok_stride, long long o_tok_stride, long long dqkv_tok_stride,
                                     float scale, int causal, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(D == 128 && S % 128 == 0 && B > 0 && H > 0);
  LLMCTL_REQUIRE(q_tok_stride % 8 == 0 && k_tok_stride % 8 == 0 && v_tok_stride % 8 == 0 && o_tok_stride % 8 == 0 &&
                 dqkv_tok_stride % 8 == 0);
  // two deterministic kernels (dK/dV, then dQ): no atomics, no fp32 dQ accumulator
  const int rc2 = flash_bwd_v2_launch(q, k, v, o, d_o, lse, dq, dk, dv, scratch, B, S, H, q_tok_stride, k_tok_stride,
                                      v_tok_stride, o_tok_stride, dqkv_tok_stride, scale, causal, (cudaStream_t)stream);
  if (rc2) LLMCTL_FAIL(rc2, "flash backward");
  return 0;
}
