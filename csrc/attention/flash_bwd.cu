// K6 (backward): causal flash attention backward on the 5th-generation tensor cores, head_dim = 128, bf16.
//
// One CTA = one (batch, head, 128-key block j); it keeps K_j and V_j in shared memory and walks the query
// blocks i >= j.  Five tensor-core GEMMs per (i, j) pair, all accumulating in TMEM (512 columns):
//     S  = Q_i K_j^T                (cols   0..127)      dP = dO_i V_j^T              (cols 128..255)
//     dV += P^T dO_i                (cols 256..383)      dK += dS^T Q_i               (cols 384..511)
//     dQ_i(partial) = dS K_j        (reuses cols 0..127 once S has been consumed)
// with  P = exp(scale S - LSE_i),  dS = scale * P o (dP - delta_i),  delta_i = rowsum(dO_i o O_i).
// P and dS are written once to shared memory in the 128-byte-swizzled K-major layout and serve both as the
// K-major A operand of dQ = dS K and — the same bytes — as the MN-major A operand of P^T dO and dS^T Q, so
// no transposed copy is ever made.  dQ partials are added to an fp32 accumulator with vector reductions
// (red.global.add.v4.f32); a small epilogue kernel casts it to bf16.
// Warp roles: 0 = TMA producer, 1 = MMA issuer, 2..5 = softmax/dS (thread = query row) and epilogue.
//
// scratch (fp32): [B*S*H*128] dQ accumulator followed by [B*H*S] delta.
//
// Reference: autograd through HF attention (reference llmctl/runtime/engine.py:288-291).  Baseline: cuDNN SDPA bwd.
#include "../common/ptx.cuh"
#include "../common/tma_host.h"
#include "../common/util.cuh"

namespace llmctl {
using namespace ptx;

constexpr int kBD = 128, kBBM = 128, kBBN = 128;
constexpr int kBPanel = 128 * 128;
constexpr int kBTile = 2 * kBPanel;
constexpr int kBwdSmem = 6 * kBTile + 1024;          // K, V, Q, dO, P, dS
constexpr int kBwdThreads = 192;

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// delta[b,h,s] = sum_d dO*O ; also clears the dQ accumulator.  One warp per (token, head).
__global__ void __launch_bounds__(256)
flash_bwd_prep_kernel(const __nv_bfloat16* __restrict__ O, long long o_tok, const __nv_bfloat16* __restrict__ dO,
                      long long do_tok, float* __restrict__ dq_acc, float* __restrict__ delta, int B, int S, int H) {
  const int lane = threadIdx.x & 31;
  const long long w = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5;
  const long long total = static_cast<long long>(B) * S * H;
  if (w >= total) return;
  const int h = static_cast<int>(w % H);
  const long long tok = w / H;
  const __nv_bfloat16* o = O + tok * o_tok + h * kBD + lane * 4;
  const __nv_bfloat16* d = dO + tok * do_tok + h * kBD + lane * 4;
  const uint2 ou = *reinterpret_cast<const uint2*>(o), du = *reinterpret_cast<const uint2*>(d);
  const float2 o0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ou.x));
  const float2 o1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&ou.y));
  const float2 d0 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&du.x));
  const float2 d1 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&du.y));
  float s = o0.x * d0.x + o0.y * d0.y + o1.x * d1.x + o1.y * d1.y;
  s = warp_sum(s);
  *reinterpret_cast<float4*>(dq_acc + (tok * H + h) * kBD + lane * 4) = make_float4(0.f, 0.f, 0.f, 0.f);
  if (lane == 0) {
    const int b = static_cast<int>(tok / S), sidx = static_cast<int>(tok % S);
    delta[(static_cast<long long>(b) * H + h) * S + sidx] = s;
  }
}

__global__ void __launch_bounds__(kBwdThreads, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                 const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                 const float* __restrict__ lse, const float* __restrict__ delta, float* __restrict__ dq_acc,
                 __nv_bfloat16* __restrict__ dK, __nv_bfloat16* __restrict__ dV, long long dkv_tok, int S, int H,
                 float scale_log2, float scale, int causal) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bar_kv, bar_qdo_full, bar_qdo_empty, bar_sfull, bar_pready, bar_dqfull, bar_sfree;
  __shared__ uint32_t tmem_base_smem;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (smem_base - smem_u32(smem_raw));
  uint8_t *sK = smem, *sV = smem + kBTile, *sQ = smem + 2 * kBTile, *sdO = smem + 3 * kBTile, *sP = smem + 4 * kBTile,
          *sdS = smem + 5 * kBTile;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int jblk = blockIdx.x, head = blockIdx.y, b = blockIdx.z;
  const int nq = S / kBBM;
  const int i0 = causal ? jblk : 0;
  const int niter = nq - i0;
  const int tok0 = b * S;
  const int k0 = jblk * kBBN;

  if (threadIdx.x == 0) {
    mbar_init(&bar_kv, 1);
    mbar_init(&bar_qdo_full, 1);
    mbar_init(&bar_qdo_empty, 1);
    mbar_init(&bar_sfull, 1);
    mbar_init(&bar_pready, 4);
    mbar_init(&bar_dqfull, 1);
    mbar_init(&bar_sfree, 4);
    fence_mbar_init();
  }
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmQ);
    prefetch_tmap(&tmK);
    prefetch_tmap(&tmV);
    prefetch_tmap(&tmdO);
  }
  if (warp == 1) tmem_alloc<1>(&tmem_base_smem, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_smem;
  const uint32_t tS = tmem_base, tdP = tmem_base + 128, tdV = tmem_base + 256, tdK = tmem_base + 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(&bar_kv, 2 * kBTile);
      tma_load_3d(sK, &tmK, &bar_kv, 0, head, tok0 + k0);
      tma_load_3d(sK + kBPanel, &tmK, &bar_kv, 64, head, tok0 + k0);
      tma_load_3d(sV, &tmV, &bar_kv, 0, head, tok0 + k0);
      tma_load_3d(sV + kBPanel, &tmV, &bar_kv, 64, head, tok0 + k0);
      for (int it = 0; it < niter; ++it) {
        const int q0 = (i0 + it) * kBBM;
        mbar_wait(&bar_qdo_empty, (it & 1u) ^ 1u);
        mbar_expect_tx(&bar_qdo_full, 2 * kBTile);
        tma_load_3d(sQ, &tmQ, &bar_qdo_full, 0, head, tok0 + q0);
        tma_load_3d(sQ + kBPanel, &tmQ, &bar_qdo_full, 64, head, tok0 + q0);
        tma_load_3d(sdO, &tmdO, &bar_qdo_full, 0, head, tok0 + q0);
        tma_load_3d(sdO + kBPanel, &tmdO, &bar_qdo_full, 64, head, tok0 + q0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t id_kk = make_idesc_bf16(128, 128, 0u, 0u);   // A K-major, B K-major
      constexpr uint32_t id_mm = make_idesc_bf16(128, 128, 1u, 1u);   // A MN-major, B MN-major
      constexpr uint32_t id_km = make_idesc_bf16(128, 128, 0u, 1u);   // A K-major, B MN-major
      const uint32_t aK = smem_base, aV = smem_base + kBTile, aQ = smem_base + 2 * kBTile, adO = smem_base + 3 * kBTile,
                     aP = smem_base + 4 * kBTile, adS = smem_base + 5 * kBTile;
      auto kdesc = [](uint32_t base, int ks) {      // K-major operand, k-step ks (16 elements of the 128-wide K dim)
        return make_smem_desc(base + (ks >> 2) * kBPanel + (ks & 3) * 32, 16, 1024, 2);
      };
      auto mdesc = [](uint32_t base, int ks) {      // MN-major operand: 64-wide MN groups are panels, k rows 128 B apart
        return make_smem_desc(base + ks * 2048, kBPanel, 1024, 2);
      };
      mbar_wait(&bar_kv, 0);
      for (int it = 0; it < niter; ++it) {
        mbar_wait(&bar_qdo_full, it & 1u);
        mbar_wait(&bar_sfree, (it & 1u) ^ 1u);        // dQ of the previous block has been drained from cols 0..127
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) umma_ss<1>(tS, kdesc(aQ, ks), kdesc(aK, ks), id_kk, ks != 0 ? 1u : 0u);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) umma_ss<1>(tdP, kdesc(adO, ks), kdesc(aV, ks), id_kk, ks != 0 ? 1u : 0u);
        umma_commit<1>(&bar_sfull);
        mbar_wait(&bar_pready, it & 1u);              // P and dS are in shared memory; S / dP have been consumed
        tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) umma_ss<1>(tdV, mdesc(aP, ks), mdesc(adO, ks), id_mm, (it | ks) != 0 ? 1u : 0u);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) umma_ss<1>(tdK, mdesc(adS, ks), mdesc(aQ, ks), id_mm, (it | ks) != 0 ? 1u : 0u);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) umma_ss<1>(tS, kdesc(adS, ks), mdesc(aK, ks), id_km, ks != 0 ? 1u : 0u);
        umma_commit<1>(&bar_dqfull);
        umma_commit<1>(&bar_qdo_empty);
      }
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const float log2e = 1.4426950408889634f;
    for (int it = 0; it < niter; ++it) {
      const int iblk = i0 + it;
      const int qi = iblk * kBBM + row;
      const bool diag = causal && (iblk == jblk);
      const long long stat = (static_cast<long long>(b) * H + head) * S + qi;
      const float lse2 = lse[stat] * log2e;
      const float dlt = delta[stat];
      mbar_wait(&bar_sfull, it & 1u);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t rs[32], rp[32];
        tmem_ld_x32(tS + lane_off + c * 32, rs);
        tmem_ld_x32(tdP + lane_off + c * 32, rp);
        tmem_ld_wait();
        uint32_t pk[16], dk[16];
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          float p0 = exp2f(__uint_as_float(rs[e]) * scale_log2 - lse2);
          float p1 = exp2f(__uint_as_float(rs[e + 1]) * scale_log2 - lse2);
          if (diag && (k0 + c * 32 + e > qi)) p0 = 0.f;
          if (diag && (k0 + c * 32 + e + 1 > qi)) p1 = 0.f;
          const float d0 = p0 * (__uint_as_float(rp[e]) - dlt) * scale;
          const float d1 = p1 * (__uint_as_float(rp[e + 1]) - dlt) * scale;
          pk[e >> 1] = pack_bf16x2(p0, p1);
          dk[e >> 1] = pack_bf16x2(d0, d1);
        }
        uint8_t* prow = sP + (c >> 1) * kBPanel + row * 128;
        uint8_t* drow = sdS + (c >> 1) * kBPanel + row * 128;
#pragma unroll
        for (int ch = 0; ch < 4; ++ch) {
          const int off = (((c & 1) * 4 + ch) ^ (row & 7)) << 4;
          *reinterpret_cast<uint4*>(prow + off) = make_uint4(pk[ch * 4], pk[ch * 4 + 1], pk[ch * 4 + 2], pk[ch * 4 + 3]);
          *reinterpret_cast<uint4*>(drow + off) = make_uint4(dk[ch * 4], dk[ch * 4 + 1], dk[ch * 4 + 2], dk[ch * 4 + 3]);
        }
      }
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_pready);
      // drain the dQ partial of this (i, j) pair into the fp32 accumulator
      mbar_wait(&bar_dqfull, it & 1u);
      tc_fence_after();
      float* dq_row = dq_acc + ((static_cast<long long>(tok0) + qi) * H + head) * kBD;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t r[32];
        tmem_ld_x32(tS + lane_off + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 8; ++g)
          red_add_v4(dq_row + c * 32 + g * 4, __uint_as_float(r[g * 4]), __uint_as_float(r[g * 4 + 1]),
                     __uint_as_float(r[g * 4 + 2]), __uint_as_float(r[g * 4 + 3]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_sfree);
    }
    // ---- epilogue: dK, dV (row = key index) ----
    // the last bar_dqfull wait above also guarantees that the dV / dK accumulations have retired
    tc_fence_after();
    const int ki = k0 + row;
    __nv_bfloat16* dkrow = dK + (static_cast<long long>(tok0) + ki) * dkv_tok + head * kBD;
    __nv_bfloat16* dvrow = dV + (static_cast<long long>(tok0) + ki) * dkv_tok + head * kBD;
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      uint32_t r1[32], r2[32];
      tmem_ld_x32(tdK + lane_off + c * 32, r1);
      tmem_ld_x32(tdV + lane_off + c * 32, r2);
      tmem_ld_wait();
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        float f[8], h2[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          f[e] = __uint_as_float(r1[g * 8 + e]);
          h2[e] = __uint_as_float(r2[g * 8 + e]);
        }
        store8(dkrow + c * 32 + g * 8, pack8(f));
        store8(dvrow + c * 32 + g * 8, pack8(h2));
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

// dq (bf16, packed layout) = dq_acc (fp32 [tokens, H, 128])
__global__ void __launch_bounds__(256)
flash_bwd_dq_cast_kernel(const float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dQ, long long dq_tok, long long n_tok,
                         int H) {
  const long long vecs = n_tok * H * (kBD / 8);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < vecs;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(i % (kBD / 8));
    const long long th = i / (kBD / 8);
    const int h = static_cast<int>(th % H);
    const long long tok = th / H;
    const float* src = dq_acc + (tok * H + h) * kBD + v * 8;
    float f[8];
    *reinterpret_cast<float4*>(f) = *reinterpret_cast<const float4*>(src);
    *reinterpret_cast<float4*>(f + 4) = *reinterpret_cast<const float4*>(src + 4);
    store8(dQ + tok * dq_tok + h * kBD + v * 8, pack8(f));
  }
}

int flash_bwd_v2_launch(const void* q, const void* k, const void* v, const void* o, const void* d_o, const float* lse,
                        void* dq, void* dk, void* dv, float* scratch, int B, int S, int H, long long q_tok,
                        long long k_tok, long long v_tok, long long o_tok, long long dqkv_tok, float scale, int causal,
                        cudaStream_t s);

}  // namespace llmctl

using namespace llmctl;

// q,k,v,o,do: [B,S,H,128] bf16 views (token strides given in elements; do and o share o_tok_stride);
// dq,dk,dv: outputs with token stride dqkv_tok_stride; lse: fp32 [B,H,S] from the forward;
// scratch: fp32 [B*S*H*128 + B*H*S].
LLMCTL_API int llmctl_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* d_o,
                                     const float* lse, void* dq, void* dk, void* dv, float* scratch, int B, int S,
                                     int H, int D, long long q_tok_stride, long long k_tok_stride,
                                     long long v_tok_stride, long long o_tok_stride, long long dqkv_tok_stride,
                                     float scale, int causal, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(D == kBD && S % kBBM == 0 && B > 0 && H > 0);
  LLMCTL_REQUIRE(q_tok_stride % 8 == 0 && k_tok_stride % 8 == 0 && v_tok_stride % 8 == 0 && o_tok_stride % 8 == 0 &&
                 dqkv_tok_stride % 8 == 0);
  cudaStream_t s = (cudaStream_t)stream;
  static const int use_v1 = getenv("LLMCTL_FLASH_BWD_V1") ? 1 : 0;
  if (!use_v1) {
    // two deterministic kernels (dK/dV, then dQ), see flash_bwd2.cu; the single-kernel version below accumulates
    // dQ with fp32 atomics and is kept for comparison
    const int rc2 = flash_bwd_v2_launch(q, k, v, o, d_o, lse, dq, dk, dv, scratch, B, S, H, q_tok_stride, k_tok_stride,
                                        v_tok_stride, o_tok_stride, dqkv_tok_stride, scale, causal, s);
    if (rc2) LLMCTL_FAIL(rc2, "flash backward v2");
    return 0;
  }
  const long long n_tok = static_cast<long long>(B) * S;
  float* dq_acc = scratch;
  float* delta = scratch + n_tok * H * kBD;
  {
    const long long warps = n_tok * H;
    const long long blocks = (warps * 32 + 255) / 256;
    flash_bwd_prep_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>((const __nv_bfloat16*)o, o_tok_stride,
                                                                         (const __nv_bfloat16*)d_o, o_tok_stride, dq_acc,
                                                                         delta, B, S, H);
  }
  CUtensorMap tq, tk, tv, tdo;
  const uint64_t rows = static_cast<uint64_t>(n_tok);
  int rc = get_tmap_3d_bf16(&tq, q, kBD, H, rows, kBD, q_tok_stride, 64, 1, kBBM, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map Q");
  rc = get_tmap_3d_bf16(&tk, k, kBD, H, rows, kBD, k_tok_stride, 64, 1, kBBN, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map K");
  rc = get_tmap_3d_bf16(&tv, v, kBD, H, rows, kBD, v_tok_stride, 64, 1, kBBN, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map V");
  rc = get_tmap_3d_bf16(&tdo, d_o, kBD, H, rows, kBD, o_tok_stride, 64, 1, kBBM, 1);
  if (rc) LLMCTL_FAIL(rc, "tensor map dO");
  LLMCTL_CHECK(cudaFuncSetAttribute(flash_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem));
  dim3 grid(S / kBBN, H, B);
  flash_bwd_kernel<<<grid, kBwdThreads, kBwdSmem, s>>>(tq, tk, tv, tdo, lse, delta, dq_acc, (__nv_bfloat16*)dk,
                                                       (__nv_bfloat16*)dv, dqkv_tok_stride, S, H,
                                                       scale * 1.4426950408889634f, scale, causal);
  {
    long long blocks = (n_tok * H * (kBD / 8) + 255) / 256;
    if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
    flash_bwd_dq_cast_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(dq_acc, (__nv_bfloat16*)dq, dqkv_tok_stride,
                                                                           n_tok, H);
  }
  LLMCTL_RETURN_LAST();
}
