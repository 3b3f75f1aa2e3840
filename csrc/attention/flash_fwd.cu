// K6 (forward): causal flash attention on the 5th-generation tensor cores, head_dim = 128, bf16.
//
// One CTA = one (batch, head, 128-query block).  Warp roles:
//   warp 0     TMA producer  — Q once, then K and V tiles of 128 keys into 2-deep smem rings (3-D tensor maps
//                              over the packed [tokens, heads, 128] activations, 128-byte swizzle)
//   warp 1     MMA issuer    — S = Q K^T  (UMMA 128x128x16, both operands K-major, accumulator S[j%2] in TMEM)
//                              O += P V   (UMMA 128x128x16, A = P from smem (K-major), B = V (MN-major), O in TMEM)
//   warps 2-5  softmax       — thread = query row: tcgen05.ld S, scale / causal mask / running max, exp2,
//                              P -> bf16 -> swizzled smem, lazy O rescale in TMEM, final normalisation + store
// S is double buffered in TMEM so Q K^T of block j+1 runs on the tensor core while the softmax warps work on
// block j.  The running max is only advanced when it grows by more than 8 (log2 units): between rescales the
// reference stays put and P may reach 2^8, which fp32 accumulation and bf16 P hold exactly as well — this
// removes almost all TMEM read-modify-write passes over O.
// Outputs: O [B,S,H,128] bf16 and the row log-sum-exp (natural log, scale included) for the backward pass.
//
// Reference: HF modules choose SDPA/eager attention behind `self.model(**batch)` (reference
// llmctl/runtime/engine.py:284); the reference's tuner times a naive softmax(QK^T)V (reference
// llmctl/plugins/autotuning.py:149-188).  Baseline to beat: cuDNN / flash-attn SDPA.
#include "../common/ptx.cuh"
#include "../common/tma_host.h"
#include "../common/util.cuh"

namespace llmctl {
using namespace ptx;

constexpr int kAD = 128;                 // head dim
constexpr int kABM = 128;                // queries per CTA
constexpr int kABN = 128;                // keys per tile
constexpr int kPanel = 128 * 128;        // bytes of one [128 rows x 64 bf16] swizzled panel
constexpr int kTile = 2 * kPanel;        // 128 x 128 bf16
constexpr int kAttnSmem = kTile /*Q*/ + 2 * kTile /*K*/ + 2 * kTile /*V*/ + kTile /*P*/ + 1024;
constexpr int kAttnThreads = 192;
constexpr float kRescaleThreshold = 8.f;

// =====================================================================================================
// Tile structure: one work item = one (batch, head, 256-query block) handled as two 128-query tiles A and B that share
// every K/V tile (half the K/V smem traffic) and ping-pong on the tensor core: while the softmax warpgroup of
// tile A works on S_A, the MMA warp runs P_B V / Q_B K^T and vice versa.
//   warps 0-3   softmax warpgroup of tile A      warps 4-7  softmax warpgroup of tile B
//   warp 8      TMA producer (Q_A, Q_B once; K, V rings of 2 x 128 keys)
//   warp 9      MMA issuer                       warps 10-11 idle (they complete the third warpgroup so that
//               setmaxnreg can move its registers to the softmax warpgroups: 216 per softmax thread, enough to
//               hold a whole 128-column S row plus the packed P chunk without spilling)
// TMEM (all 512 columns): S_A [0,128) S_B [128,256) O_A [256,384) O_B [384,512).  P never touches shared
// memory: the softmax thread overwrites the first 64 columns of its own S row with bf16x2 probabilities
// (tcgen05.st) and P V reads its A operand straight from TMEM.  The tensor pipe executes in issue order, so
// Q K^T of block j+1 (which overwrites S/P) is simply issued after P V of block j; tcgen05.commit covers every
// earlier MMA, so "S(j+1) ready" also means "P V(j) has retired" and the O rescale needs no extra barrier.
// The epilogue stages O through the (now idle) Q tile and writes it with TMA.
constexpr int kF2Threads = 384;
// kPolyCount of every kPolyPeriod score pairs can take the FMA-pipe exponential (ptx.cuh ex2_poly2) instead of
// MUFU.EX2: per 256x128 tile pair the SFU needs 32768 / 16 = 2048 clk, the same as the tensor core.  Measured on
// B200 (B1 S4096 H32): 0 -> 170 us, 2 of 4 -> 180 us — the softmax warps are bound by dependent-issue latency,
// not by SFU throughput, so the extra FMA instructions cost more than the SFU time they free.  Default off.
#ifndef LLMCTL_FLASH_POLY_COUNT
#define LLMCTL_FLASH_POLY_COUNT 0
#endif
constexpr int kPolyPeriod = 4;
constexpr int kPolyCount = LLMCTL_FLASH_POLY_COUNT;
constexpr int kF2Smem = 6 * kTile + 1024;

// On the diagonal block the mask is resolved per 32-column chunk, which is warp-uniform (warp q owns rows
// 32q .. 32q+31): chunks left of the diagonal chunk take the plain path, chunks right of it are all masked and
// cost nothing (no compare, no exponential, P = 0), only chunk q compares its 32 columns against the lane.
template <bool kDiag>
__device__ __forceinline__ void softmax_block_v2(uint32_t tS, uint32_t tO, int row, bool first, float scale_log2,
                                                 float& m_ref, float& l0, float& l1) {
  const int dq = row >> 5;            // diagonal chunk of this warp
  const int dl = row & 31;            // lane = row within the chunk
  // The whole S row (128 fp32) is read from TMEM once and stays in registers for both passes: every
  // tcgen05.ld -> wait round trip removed here is latency the tensor pipe would otherwise sit idle for.
  uint32_t r[128];
  tmem_ld_x32(tS, r);
  tmem_ld_x32(tS + 32, r + 32);
  tmem_ld_x32(tS + 64, r + 64);
  tmem_ld_x32(tS + 96, r + 96);
  tmem_ld_wait();
  // pass 1: raw row maximum (the scale is positive, so it commutes with max)
  float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    if (kDiag && c > dq) continue;                       // fully masked chunk
    const bool edge = kDiag && c == dq;
#pragma unroll
    for (int e = c * 32; e < c * 32 + 32; e += 4) {
      float a = __uint_as_float(r[e]), b2 = __uint_as_float(r[e + 1]), c2 = __uint_as_float(r[e + 2]),
            d2 = __uint_as_float(r[e + 3]);
      if (edge) {
        if ((e & 31) > dl) a = -INFINITY;
        if ((e & 31) + 1 > dl) b2 = -INFINITY;
        if ((e & 31) + 2 > dl) c2 = -INFINITY;
        if ((e & 31) + 3 > dl) d2 = -INFINITY;
      }
      m0 = fmaxf(m0, a);
      m1 = fmaxf(m1, b2);
      m2 = fmaxf(m2, c2);
      m3 = fmaxf(m3, d2);
    }
  }
  const float bm = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)) * scale_log2;
  // lazy reference update: only move the reference when the block maximum outgrows it by more than 2^8
  const bool need = bm > m_ref + kRescaleThreshold;
  float alpha = 1.f;
  if (need) {
    alpha = (m_ref == -INFINITY) ? 0.f : ex2_approx(m_ref - bm);
    m_ref = bm;
    l0 *= alpha;
    l1 *= alpha;
  }
  const bool rescale = !first && __any_sync(0xffffffffu, need);
  // pass 2: p = 2^(s*scale - m_ref) -> bf16x2 -> back into the first 64 columns of this S row
  const float neg_m = -m_ref;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    uint32_t pk[16];
    if (kDiag && c > dq) {                               // fully masked chunk: P = 0, nothing to compute
#pragma unroll
      for (int e = 0; e < 16; ++e) pk[e] = 0u;
      tmem_st_x16(tS + c * 16, pk);
      continue;
    }
    const bool edge = kDiag && c == dq;
#pragma unroll
    for (int e = 0; e < 32; e += 2) {
      float a0, a1;
      fma2(a0, a1, __uint_as_float(r[c * 32 + e]), __uint_as_float(r[c * 32 + e + 1]), scale_log2, neg_m);
      if (edge) {
        if (e > dl) a0 = -INFINITY;
        if (e + 1 > dl) a1 = -INFINITY;
      }
      float p0, p1;
      if (((e >> 1) % kPolyPeriod) < kPolyCount) {
        ex2_poly2(p0, p1, a0, a1);         // FMA-pipe exponential for this pair (see ptx.cuh)
      } else {
        p0 = ex2_approx(a0);
        p1 = ex2_approx(a1);
      }
      add2(l0, l1, p0, p1);
      pk[e >> 1] = pack_bf16x2(p0, p1);
    }
    tmem_st_x16(tS + c * 16, pk);
  }
  if (rescale) {
    // rare (reference moved by more than 2^8): scale the running O row; P stores above are already in flight
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t o[32];
      tmem_ld_x32(tO + c * 32, o);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; e += 2) {
        float x0, x1;
        mul2(x0, x1, __uint_as_float(o[e]), __uint_as_float(o[e + 1]), alpha);
        o[e] = __float_as_uint(x0);
        o[e + 1] = __float_as_uint(x1);
      }
      tmem_st_x16(tO + c * 32, o);
      tmem_st_x16(tO + c * 32 + 16, o + 16);
    }
  }
  tmem_st_wait();
}

// =====================================================================================================
// The kernel is PERSISTENT.  The timeline of a one-CTA-per-item launch (tools/trace_flash.py) showed ~5.5 k clocks
// between kernel start and the first usable S (barrier / TMEM set-up, tensor-map fetch, Q and K TMA latency,
// first Q K^T) and ~3 k for the epilogue, against ~3.5 k per 256x128 iteration: at the training shape
// (S = 512, 2-4 iterations per CTA) more than half of every CTA's life was set-up and drain, with nothing else
// resident on the SM (197 KB of shared memory) to hide it.  Here one CTA per SM walks a heavy-first list of
// (batch, head, 256-query block) items: TMEM, barriers and tensor maps are set up once, the K/V ring, the
// S/P ping-pong and all barrier phases run straight through item boundaries, the TMA warp prefetches the next
// item's Q tiles as soon as the last Q K^T of the current one has retired, and the MMA warp starts the next
// item's Q K^T while the softmax warpgroups are still normalising and storing O of the current one (O goes
// from registers straight to global memory, so no staging buffer blocks the Q prefetch).
__global__ void __launch_bounds__(kF2Threads, 1)
flash_fwd3_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, __nv_bfloat16* __restrict__ O, long long o_tok_stride,
                  float* __restrict__ lse, int B, int S, int H, float scale_log2, int causal,
                  long long* __restrict__ trace) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar_q[2], bar_qfree[2], bar_kfull[2], bar_vfull[2], bar_kempty[2], bar_vempty[2];
  __shared__ uint64_t bar_s[2], bar_p[2], bar_o[2], bar_ofree[2];
  __shared__ uint32_t tmem_base_smem;

  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));      // layout: Q_A, Q_B, K[2], V[2]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nqb = (S + 2 * kABM - 1) / (2 * kABM);
  const int per_qb = H * B;
  const int n_items = nqb * per_qb;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bar_q[i], 1);
      mbar_init(&bar_qfree[i], 1);
      mbar_init(&bar_kfull[i], 1);
      mbar_init(&bar_vfull[i], 1);
      mbar_init(&bar_kempty[i], 1);
      mbar_init(&bar_vempty[i], 1);
      mbar_init(&bar_s[i], 1);
      mbar_init(&bar_p[i], 4);
      mbar_init(&bar_o[i], 1);
      mbar_init(&bar_ofree[i], 4);
    }
    fence_mbar_init();
  }
  if (warp == 8 && lane == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmK);
    prefetch_tmap(&tmV);
  }
  if (warp == 9) tmem_alloc<1>(&tmem_base_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;

  // optional event log of CTA 0 (LLMCTL_FLASH_TRACE): per role 127 x (code, clock) pairs after a count word.
  // roles: 0/1 softmax A/B (codes 1 = S ready, 2 = P written, 3 = O stored), 2 / 4 = MMA issuer of tile A / B (10+t = P seen,
  // 20+t = issued, 30 = item start), 3 = TMA (40 = item start)
  const bool tracing = trace != nullptr && blockIdx.x == 0;
  const long long t_start = tracing ? clock64() : 0;
  int n_ev = 0;
  auto ev = [&](int role, int code) {
    if (tracing && n_ev < 127) {
      trace[role * 256 + 2 + 2 * n_ev] = code;
      trace[role * 256 + 3 + 2 * n_ev] = clock64() - t_start;
      ++n_ev;
      trace[role * 256] = n_ev;
    }
  };
  // item -> (query block pair, head, batch); heavy (late) query blocks first so the tail is made of light items
  auto decode = [&](int item, int& qb2, int& head, int& b) {
    qb2 = nqb - 1 - item / per_qb;
    const int rem = item % per_qb;
    head = rem % H;
    b = rem / H;
  };

  if (warp >= 8) {
   // register budget: the CTA owns 384 x 168; what warps 8-11 give back ((168-104) x 128) is exactly what the two
   // softmax warpgroups take ((200-168) x 256) — asking for more than was released blocks forever.  
   setmaxnreg_dec<104>();
   if (warp == 8) {
    // ------------------------------------ TMA producer ------------------------------------
    if (lane == 0) {
      uint32_t g = 0;                 // K/V blocks loaded so far (ring position and phase)
      uint32_t nA = 0, nB = 0;        // Q_A / Q_B loads so far
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int qb2, head, b;
        decode(item, qb2, head, b);
        const int q0 = qb2 * 2 * kABM;
        const bool validB = q0 + kABM < S;
        const int nblk = causal ? (validB ? 2 * qb2 + 2 : 2 * qb2 + 1) : (S / kABN);
        const int tok0 = b * S;
        ev(3, 40);
        mbar_wait(&bar_qfree[0], (nA & 1u) ^ 1u);        // the previous item's last Q_A K^T has retired
        mbar_expect_tx(&bar_q[0], kTile);
        tma_load_3d(smem, &tmQ, &bar_q[0], 0, head, tok0 + q0);
        tma_load_3d(smem + kPanel, &tmQ, &bar_q[0], 64, head, tok0 + q0);
        ++nA;
        for (int j = 0; j < nblk; ++j, ++g) {
          const int st = g & 1u;
          const uint32_t ph = (g >> 1) & 1u;
          uint8_t* sK = smem + 2 * kTile + st * kTile;
          uint8_t* sV = smem + 4 * kTile + st * kTile;
          mbar_wait(&bar_kempty[st], ph ^ 1u);
          mbar_expect_tx(&bar_kfull[st], kTile);
          tma_load_3d(sK, &tmK, &bar_kfull[st], 0, head, tok0 + j * kABN);
          tma_load_3d(sK + kPanel, &tmK, &bar_kfull[st], 64, head, tok0 + j * kABN);
          if (j == 0 && validB) {
            mbar_wait(&bar_qfree[1], (nB & 1u) ^ 1u);
            mbar_expect_tx(&bar_q[1], kTile);
            tma_load_3d(smem + kTile, &tmQ, &bar_q[1], 0, head, tok0 + q0 + kABM);
            tma_load_3d(smem + kTile + kPanel, &tmQ, &bar_q[1], 64, head, tok0 + q0 + kABM);
            ++nB;
          }
          mbar_wait(&bar_vempty[st], ph ^ 1u);
          mbar_expect_tx(&bar_vfull[st], kTile);
          tma_load_3d(sV, &tmV, &bar_vfull[st], 0, head, tok0 + j * kABN);
          tma_load_3d(sV + kPanel, &tmV, &bar_vfull[st], 64, head, tok0 + j * kABN);
        }
      }
    }
   } else if (warp == 9) {
    // ------------------------------------ MMA issuer ------------------------------------
    // One thread issues for both tiles in strict A, B, A, B order: the issue order is what keeps the two tiles in
    // anti-phase (one in its softmax while the other owns the tensor pipe).  Measured alternatives (B1 S8192 H32):
    // two issuing warps, one per tile — the tiles drift into lock-step and the exponential unit and the tensor
    // core alternate instead of overlapping, 1010 -> 750 TFLOP/s; issuing the next item's first Q K^T early from
    // inside this loop — the extra live state spills into the issue sequences (or, with more registers here,
    // starves the softmax warps), 1010 -> 720-830 TFLOP/s.  This thread's issue latency is on the critical path.
    if (lane == 0) {
      constexpr uint32_t idesc_qk = make_idesc_bf16(kABM, kABN, 0u, 0u);   // A, B K-major
      constexpr uint32_t idesc_pv = make_idesc_bf16(kABM, kAD, 0u, 1u);    // A from TMEM, B MN-major (V)
      uint32_t g0 = 0;                 // K/V blocks consumed by earlier items
      uint32_t nQ[2] = {0, 0};         // items done per tile (Q / O barrier phases)
      uint32_t nP[2] = {0, 0};         // P V products issued per tile (bar_p phase)
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int qb2, head, b;
        decode(item, qb2, head, b);
        const int q0 = qb2 * 2 * kABM;
        const bool validB = q0 + kABM < S;
        const int nblk_t0 = causal ? (2 * qb2 + 1) : (S / kABN);
        const int nblk_t1 = validB ? (causal ? (2 * qb2 + 2) : (S / kABN)) : 0;
        const int nblk = nblk_t1 > nblk_t0 ? nblk_t1 : nblk_t0;
        auto nblk_of = [&](int t) { return t == 0 ? nblk_t0 : nblk_t1; };
        auto issue_qk = [&](int t, int j, bool release_k) {
          const uint32_t g = g0 + j;
          const int st = g & 1u;
          mbar_wait(&bar_kfull[st], (g >> 1) & 1u);
          tc_fence_after();
          const uint32_t q_addr = smem_base + t * kTile;
          const uint32_t k_addr = smem_base + 2 * kTile + st * kTile;
#pragma unroll
          for (int ks = 0; ks < kAD / 16; ++ks) {
            const uint32_t off = (ks >> 2) * kPanel + (ks & 3) * 32;
            umma_ss<1>(tmem_base + t * 128, make_smem_desc(q_addr + off, 16, 1024, 2),
                       make_smem_desc(k_addr + off, 16, 1024, 2), idesc_qk, ks != 0 ? 1u : 0u);
          }
          umma_commit<1>(&bar_s[t]);
          if (release_k) umma_commit<1>(&bar_kempty[st]);
          if (j + 1 == nblk_of(t)) umma_commit<1>(&bar_qfree[t]);       // last use of this item's Q_t
        };
        mbar_wait(&bar_q[0], nQ[0] & 1u);
        issue_qk(0, 0, !validB);
        if (validB) {
          mbar_wait(&bar_q[1], nQ[1] & 1u);
          issue_qk(1, 0, true);
        }
        for (int j = 0; j < nblk; ++j) {
          const uint32_t g = g0 + j;
          const int st = g & 1u;
          for (int t = 0; t < 2; ++t) {
            if (j >= nblk_of(t)) continue;
            mbar_wait(&bar_vfull[st], (g >> 1) & 1u);
            mbar_wait(&bar_p[t], nP[t] & 1u);                  // P_t(j) is in TMEM, O_t rescaled if needed
            ++nP[t];
            if (j == 0 && nQ[t] > 0) mbar_wait(&bar_ofree[t], (nQ[t] - 1) & 1u);   // previous item's O_t was read out
            tc_fence_after();
            const uint32_t v_addr = smem_base + 4 * kTile + st * kTile;
            const uint32_t tP = tmem_base + t * 128;
            const uint32_t tOt = tmem_base + 256 + t * 128;
#pragma unroll
            for (int ks = 0; ks < kABN / 16; ++ks)
              umma_ts<1>(tOt, tP + ks * 8, make_smem_desc(v_addr + ks * 2048, kPanel, 1024, 2), idesc_pv,
                         (j | ks) != 0 ? 1u : 0u);
            if (t == 1 || j >= nblk_t1) umma_commit<1>(&bar_vempty[st]);    // last reader of V(j)
            if (j + 1 < nblk_of(t))
              issue_qk(t, j + 1, t == 1 || j + 1 >= nblk_t1);
            else
              umma_commit<1>(&bar_o[t]);
          }
        }
        g0 += nblk;
        ++nQ[0];
        if (validB) ++nQ[1];
      }
    }
   }
  } else {
    setmaxnreg_inc<200>();
    // ---------------------------- softmax warpgroups ----------------------------
    const int t = warp >> 2;
    const int q = warp & 3;                        // TMEM lane quadrant this warp may touch
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t tS = tmem_base + t * 128 + lane_off;
    const uint32_t tOt = tmem_base + 256 + t * 128 + lane_off;
    uint32_t nS = 0;                               // S tiles consumed so far (bar_s phase)
    uint32_t nO = 0;                               // items finished by this tile (bar_o phase)
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      int qb2, head, b;
      decode(item, qb2, head, b);
      const int q0 = qb2 * 2 * kABM;
      const int qt0 = q0 + t * kABM;
      if (qt0 >= S) continue;                      // tile B past the end of the sequence: nobody touches its barriers
      const int nb = causal ? (2 * qb2 + t + 1) : (S / kABN);
      const int tok0 = b * S;
      float m_ref = -INFINITY, l0 = 0.f, l1 = 0.f;
      for (int j = 0; j < nb; ++j, ++nS) {
        mbar_wait(&bar_s[t], nS & 1u);
        tc_fence_after();
        if (q == 0 && lane == 0) ev(t, 1);
        if (causal && j == nb - 1)
          softmax_block_v2<true>(tS, tOt, row, j == 0, scale_log2, m_ref, l0, l1);
        else
          softmax_block_v2<false>(tS, tOt, row, j == 0, scale_log2, m_ref, l0, l1);
        tc_fence_before();
        __syncwarp();
        if (q == 0 && lane == 0) ev(t, 2);
        if (lane == 0) mbar_arrive(&bar_p[t]);
      }
      // ---- epilogue: O / l -> bf16 -> global (row per thread), LSE ----
      mbar_wait(&bar_o[t], nO & 1u);
      ++nO;
      tc_fence_after();
      const float l = l0 + l1;
      const float inv_l = 1.f / l;
      __nv_bfloat16* orow = O + (static_cast<long long>(tok0 + qt0 + row)) * o_tok_stride + head * kAD;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        tmem_ld_x32(tOt + c * 32, r);
        tmem_ld_wait();
        if (c == 3) {                              // O_t is in registers: the next item's P V may overwrite it
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bar_ofree[t]);
        }
#pragma unroll
        for (int gq = 0; gq < 4; ++gq) {
          uint4 w;
          w.x = pack_bf16x2(__uint_as_float(r[gq * 8 + 0]) * inv_l, __uint_as_float(r[gq * 8 + 1]) * inv_l);
          w.y = pack_bf16x2(__uint_as_float(r[gq * 8 + 2]) * inv_l, __uint_as_float(r[gq * 8 + 3]) * inv_l);
          w.z = pack_bf16x2(__uint_as_float(r[gq * 8 + 4]) * inv_l, __uint_as_float(r[gq * 8 + 5]) * inv_l);
          w.w = pack_bf16x2(__uint_as_float(r[gq * 8 + 6]) * inv_l, __uint_as_float(r[gq * 8 + 7]) * inv_l);
          *reinterpret_cast<uint4*>(orow + c * 32 + gq * 8) = w;
        }
      }
      if (lse != nullptr)
        lse[(static_cast<long long>(b) * H + head) * S + qt0 + row] = (m_ref + log2f(l)) * 0.6931471805599453f;
      if (q == 0 && lane == 0) ev(t, 3);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

// Debug timeline buffer: allocated once when LLMCTL_FLASH_TRACE is set in the environment, NULL otherwise.
static long long* flash_trace_buffer() {
  static long long* buf = nullptr;
  static int init = 0;
  if (!init) {
    init = 1;
    if (getenv("LLMCTL_FLASH_TRACE") != nullptr && cudaMalloc(&buf, 2048 * sizeof(long long)) == cudaSuccess)
      cudaMemset(buf, 0, 2048 * sizeof(long long));
    else
      buf = nullptr;
  }
  return buf;
}

}  // namespace llmctl

using namespace llmctl;

// Copies the forward kernel's debug event log (2048 int64: per role a count, then (code, clock) pairs; see the
// kernel for roles and codes) to the host.
LLMCTL_API int llmctl_flash_trace_read(long long* host_out) {
  long long* b = flash_trace_buffer();
  if (b == nullptr) return static_cast<int>(cudaErrorNotReady);
  return static_cast<int>(cudaMemcpy(host_out, b, 2048 * sizeof(long long), cudaMemcpyDeviceToHost));
}

// q,k,v: [B, S, H, 128] bf16 views with a common layout: element (b,s,h,d) at ((b*S+s)*tok_stride + h*128 + d).
// o: same indexing with o_tok_stride.  lse: fp32 [B, H, S] (may be NULL).  S must be a multiple of 128.
LLMCTL_API int llmctl_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B,
                                     int S, int H, int D, long long q_tok_stride, long long k_tok_stride,
                                     long long v_tok_stride, long long o_tok_stride, float scale, int causal,
                                     void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(D == kAD && S % kABM == 0 && B > 0 && H > 0);
  LLMCTL_REQUIRE(q_tok_stride % 8 == 0 && k_tok_stride % 8 == 0 && v_tok_stride % 8 == 0 && o_tok_stride % 8 == 0);
  CUtensorMap tq, tk, tv;
  const uint64_t rows = static_cast<uint64_t>(B) * S;
  int rc = get_tmap_3d_bf16(&tq, q, kAD, H, rows, kAD, q_tok_stride, 64, 1, kABM, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map Q");
  rc = get_tmap_3d_bf16(&tk, k, kAD, H, rows, kAD, k_tok_stride, 64, 1, kABN, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map K");
  rc = get_tmap_3d_bf16(&tv, v, kAD, H, rows, kAD, v_tok_stride, 64, 1, kABN, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map V");
  LLMCTL_REQUIRE(scale > 0.f);
  // persistent kernel: one CTA per SM over a heavy-first item list
  LLMCTL_CHECK(cudaFuncSetAttribute(flash_fwd3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kF2Smem));
  const int n_items = ((S + 2 * kABM - 1) / (2 * kABM)) * H * B;
  const int grid3 = n_items < kNumSMs ? n_items : kNumSMs;
  flash_fwd3_kernel<<<grid3, kF2Threads, kF2Smem, (cudaStream_t)stream>>>(
      tq, tk, tv, (__nv_bfloat16*)o, o_tok_stride, lse, B, S, H, scale * 1.4426950408889634f, causal,
      flash_trace_buffer());
  LLMCTL_RETURN_LAST();
}
