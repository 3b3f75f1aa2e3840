// K7: ZeRO gradient reduce-scatter + AdamW + parameter all-gather over NVLink peer memory, without NCCL.
//
// Every data-parallel rank keeps its bf16 gradient buckets and (ZeRO-1/2) its bf16 parameter buffer in symmetric
// memory.
//   kernel 1  rs_sumsq      : each rank reduces ITS 1/dp slice of a bucket across all ranks — either in the
//                             NVSwitch (one `multimem.ld_reduce` per 16 bytes: the switch reads the eight peers'
//                             copies and returns the fp32-accumulated sum, so the SM issues 1/8 of the loads and
//                             few CTAs saturate the links) or with 16-byte P2P loads summed in a fixed rank order
//                             (deterministic mode / no multicast mapping) — stores the reduced slice locally in
//                             fp32 (optionally accumulating over micro-batches) and produces sum-of-squares
//                             partials for the global gradient norm.  Launched per unit from backward as soon as
//                             the unit's last gradient landed, so the transfer overlaps the rest of backward.
//   kernel 2  adam_allgather: clip coefficient from the global norm (peers' partial sums read over NVLink, or a
//                             device scalar when TP/PP ranks contribute), AdamW on the local slice (fp32 master +
//                             moments), then the updated bf16 parameters are PUSHED into every rank's parameter
//                             buffer (one NVLS `multimem.st` or staggered P2P stores); ZeRO-3 writes only the
//                             local shard.  One launch per unit, in next-forward order on a side stream, each
//                             ending with its own all-rank "unit complete" exchange so the next step's forward
//                             can start on unit 0 while later units are still in flight.
// Cross-GPU ordering uses epoch-numbered flags in the symmetric signal pads (csrc/comm/symm.cuh).
//
// Reference: NCCL all-reduce inside DDP + torch.optim.AdamW + clip_grad_norm_
// (reference llmctl/runtime/engine.py:265-267,289,295-303).  SURVEY.md §2.6.b K7.
#include "../comm/symm.cuh"
#include "../common/util.cuh"

namespace llmctl {
using namespace symm;

// slot layout inside the signal pads (uint32 indices).  rs_sumsq uses the gradient buffer's pads, adam_allgather
// the parameter buffer's (or, for ZeRO-3, the scalar buffer's) — the two never share a pad.
constexpr int kSlotGradReady = 0;    // [0, 8)   peer r finished writing its bucket for launch #epoch
constexpr int kSlotPullDone = 8;     // [8, 16)  peer r finished all pulls of this step
constexpr int kSlotParamsDone = 16;  // [16, 24) peer r finished pushing this unit
constexpr int kSlotCounter = 24;     // local CTA counter (adam)
constexpr int kSlotRsCounter = 25;   // local CTA counter (rs, only with done_slot)
// [32, 32 + 8 * n) : "every pull of the bucket in transient slot i is finished" (host chooses done_slot)

template <bool kNvls>
__global__ void __launch_bounds__(256)
rs_sumsq_kernel(PeerTable grads, const __nv_bfloat16* __restrict__ mc, long long bucket_off, PeerTable pads,
                int rank, int world, long long shard_lo, long long shard_n, float* __restrict__ reduced,
                int accumulate_out, float* __restrict__ partials, uint32_t epoch, int done_slot,
                uint32_t done_epoch) {
  __shared__ float red[32];
  // announce "my bucket is complete" (block 0), then everybody waits for all peers
  if (blockIdx.x == 0 && threadIdx.x < world)
    signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + kSlotGradReady + rank, epoch);
  if (threadIdx.x < world)
    signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + kSlotGradReady + threadIdx.x, epoch);
  __syncthreads();

  float acc_sq = 0.f;
  const long long nvec = shard_n >> 3;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  auto finish = [&](long long i, const float* s_in) {
    float s[8];
    float* dst = reduced + (i << 3);
    if (accumulate_out) {
      const float4 a = *reinterpret_cast<const float4*>(dst), b = *reinterpret_cast<const float4*>(dst + 4);
      s[0] = s_in[0] + a.x; s[1] = s_in[1] + a.y; s[2] = s_in[2] + a.z; s[3] = s_in[3] + a.w;
      s[4] = s_in[4] + b.x; s[5] = s_in[5] + b.y; s[6] = s_in[6] + b.z; s[7] = s_in[7] + b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] = s_in[j];
    }
    *reinterpret_cast<float4*>(dst) = make_float4(s[0], s[1], s[2], s[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(s[4], s[5], s[6], s[7]);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_sq += s[j] * s[j];
  };

  if constexpr (kNvls) {
    const __nv_bfloat16* src = mc + bucket_off + shard_lo;
    long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    for (; i + 3 * stride < nvec; i += 4 * stride) {          // 4 independent switch reductions in flight
      uint4 raw[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) raw[u] = ptx::multimem_ld_reduce_bf16x8(src + ((i + u * stride) << 3));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        unpack8(*reinterpret_cast<const bf16x8*>(&raw[u]), f);
        finish(i + u * stride, f);
      }
    }
    for (; i < nvec; i += stride) {
      const uint4 raw = ptx::multimem_ld_reduce_bf16x8(src + (i << 3));
      float f[8];
      unpack8(*reinterpret_cast<const bf16x8*>(&raw), f);
      finish(i, f);
    }
  } else {
    const __nv_bfloat16* src[kMaxRanks];
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      src[r] = (r < world) ? reinterpret_cast<const __nv_bfloat16*>(grads.ptr[r]) + bucket_off + shard_lo : nullptr;
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec; i += stride) {
      float s[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] = 0.f;
      uint4 raw[kMaxRanks];
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r)
        if (r < world) raw[r] = ptx::ld_nc_v4(src[r] + (i << 3));     // all loads in flight first
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r) {
        if (r < world) {
          float f[8];
          unpack8(*reinterpret_cast<const bf16x8*>(&raw[r]), f);
#pragma unroll
          for (int j = 0; j < 8; ++j) s[j] += f[j];                    // fixed order r = 0..world-1
        }
      }
      finish(i, s);
    }
  }
  acc_sq = block_sum(acc_sq, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc_sq;

  if (done_slot >= 0) {
    // transient buckets: tell every peer that this rank no longer reads the bucket (last CTA only)
    __shared__ unsigned int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned int* counter = reinterpret_cast<unsigned int*>(pads.ptr[rank]) + kSlotRsCounter;
      const unsigned int done = atomicAdd(counter, 1u) + 1u;
      s_last = (done == gridDim.x) ? 1u : 0u;
      if (s_last) atomicExch(counter, 0u);
    }
    __syncthreads();
    if (s_last && threadIdx.x < world)
      signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + done_slot + rank, done_epoch);
  }
}

__global__ void __launch_bounds__(256)
sumsq_fold_kernel(const float* __restrict__ partials, int n, float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  float a = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) a += partials[i];
  a = block_sum(a, red);
  if (threadIdx.x == 0) *out = accumulate ? *out + a : a;
}

// One CTA: wait until every peer's flag [slot + r] in the LOCAL pad reached `epoch`.
__global__ void __launch_bounds__(32)
signal_wait_kernel(PeerTable pads, int rank, int world, int slot, uint32_t epoch) {
  if (threadIdx.x < world)
    signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + slot + threadIdx.x, epoch);
}

// scalars: symmetric float buffer; scalars[0] of each rank = that rank's total sum of squares.
__global__ void __launch_bounds__(256)
adam_allgather_kernel(PeerTable params, long long bucket_off, void* mc_params, PeerTable pads,
                      PeerTable scalars, const float* __restrict__ total_sq_dev, int rank, int world,
                      long long shard_lo, long long shard_n, float* __restrict__ master,
                      float* __restrict__ m, float* __restrict__ v, const float* __restrict__ reduced,
                      __nv_bfloat16* __restrict__ local_out, int push, float lr, float beta1, float beta2,
                      float eps, float wd, float bc1, float bc2, float grad_scale, float max_norm,
                      float* __restrict__ norm_out, uint32_t epoch_begin, int begin_barrier,
                      uint32_t epoch_end, int end_barrier) {
  if (begin_barrier) {
    if (blockIdx.x == 0 && threadIdx.x < world)
      signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + kSlotPullDone + rank, epoch_begin);
    if (threadIdx.x < world)
      signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + kSlotPullDone + threadIdx.x, epoch_begin);
    __syncthreads();
  }
  // global gradient norm: every rank reduced a disjoint slice, so the squares simply add up
  float total = 0.f;
  if (total_sq_dev != nullptr) {
    total = *total_sq_dev;
  } else {
    for (int r = 0; r < world; ++r)
      total += __uint_as_float(ptx::ld_relaxed_sys(reinterpret_cast<const uint32_t*>(scalars.ptr[r])));
  }
  const float norm = sqrtf(total) * grad_scale;
  float coef = 1.f;
  if (max_norm > 0.f) coef = fminf(1.f, max_norm / (norm + 1e-6f));
  if (blockIdx.x == 0 && threadIdx.x == 0 && norm_out) *norm_out = norm;
  const float scale = grad_scale * coef;
  const bool skip = !isfinite(norm);

  __nv_bfloat16* dst[kMaxRanks];
#pragma unroll
  for (int r = 0; r < kMaxRanks; ++r)
    dst[r] = (push && r < world) ? reinterpret_cast<__nv_bfloat16*>(params.ptr[r]) + bucket_off + shard_lo : nullptr;
  __nv_bfloat16* mc = (push && mc_params) ? reinterpret_cast<__nv_bfloat16*>(mc_params) + bucket_off + shard_lo : nullptr;

  const float step_size = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const float decay = 1.f - lr * wd;
  const long long nvec = shard_n >> 3;
  if (!skip) {
    for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
      const long long o = i << 3;
      float g[8], p[8], mm[8], vv[8];
      *reinterpret_cast<float4*>(g) = *reinterpret_cast<const float4*>(reduced + o);
      *reinterpret_cast<float4*>(g + 4) = *reinterpret_cast<const float4*>(reduced + o + 4);
      *reinterpret_cast<float4*>(p) = *reinterpret_cast<const float4*>(master + o);
      *reinterpret_cast<float4*>(p + 4) = *reinterpret_cast<const float4*>(master + o + 4);
      *reinterpret_cast<float4*>(mm) = *reinterpret_cast<const float4*>(m + o);
      *reinterpret_cast<float4*>(mm + 4) = *reinterpret_cast<const float4*>(m + o + 4);
      *reinterpret_cast<float4*>(vv) = *reinterpret_cast<const float4*>(v + o);
      *reinterpret_cast<float4*>(vv + 4) = *reinterpret_cast<const float4*>(v + o + 4);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float gj = g[j] * scale;
        mm[j] = beta1 * mm[j] + (1.f - beta1) * gj;
        vv[j] = beta2 * vv[j] + (1.f - beta2) * gj * gj;
        const float denom = sqrtf(vv[j]) * inv_sqrt_bc2 + eps;
        p[j] = p[j] * decay - step_size * (mm[j] / denom);
      }
      *reinterpret_cast<float4*>(master + o) = *reinterpret_cast<float4*>(p);
      *reinterpret_cast<float4*>(master + o + 4) = *reinterpret_cast<float4*>(p + 4);
      *reinterpret_cast<float4*>(m + o) = *reinterpret_cast<float4*>(mm);
      *reinterpret_cast<float4*>(m + o + 4) = *reinterpret_cast<float4*>(mm + 4);
      *reinterpret_cast<float4*>(v + o) = *reinterpret_cast<float4*>(vv);
      *reinterpret_cast<float4*>(v + o + 4) = *reinterpret_cast<float4*>(vv + 4);
      const bf16x8 packed = pack8(p);
      const uint4 bits = *reinterpret_cast<const uint4*>(&packed);
      if (local_out != nullptr) *reinterpret_cast<uint4*>(local_out + o) = bits;
      if (mc != nullptr) {
        ptx::multimem_st_v4(mc + o, bits);                           // one store, the switch replicates it
      } else if (push) {
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r)
          if (r < world) ptx::st_v4(dst[(rank + r) % world] + o, bits);   // staggered start: no egress hot-spot
      }
    }
  }
  if (end_barrier) {
    // every block must have finished its stores before this rank announces completion
    __threadfence_system();
    __shared__ unsigned int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned int* counter = reinterpret_cast<unsigned int*>(pads.ptr[rank]) + kSlotCounter;   // local CTA counter
      const unsigned int done = atomicAdd(counter, 1u) + 1u;
      s_last = (done == gridDim.x) ? 1u : 0u;
      if (s_last) atomicExch(counter, 0u);        // leave the counter clean for the next launch
    }
    __syncthreads();
    if (s_last) {
      if (threadIdx.x < world)
        signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + kSlotParamsDone + rank, epoch_end);
      if (threadIdx.x < world)
        signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + kSlotParamsDone + threadIdx.x, epoch_end);
    }
  }
}

}  // namespace llmctl

using namespace llmctl;
LLMCTL_SYMM_WATCHDOG_MODULE(rs_adam)

// Reduce this rank's [shard_lo, shard_lo + shard_n) slice of a bf16 bucket over all ranks.
//   grads      : per-rank base pointers of the symmetric gradient buffer; bucket_off = unit offset (elements)
//   mc_grads   : NVLS multicast address of the same buffer (in-switch reduction) or NULL (P2P pull, rank order)
//   reduced    : local fp32 [shard_n] output (+= when accumulate_out);  scratch: fp32 [scratch_elems] partials
//   sumsq_out  : local float, (+)= sum of squares of `reduced` after this call, or NULL
//   epoch      : launch counter of this entry point (same sequence on every rank)
//   done_slot  : >= 0: pad slot base where every peer is told "rank finished reading this bucket" (= done_epoch);
//                the bucket's next writer waits for it with llmctl_signal_wait.  -1: no announcement.
LLMCTL_API int llmctl_rs_sumsq(llmctl::symm::PeerTable grads, long long bucket_off, const void* mc_grads,
                               llmctl::symm::PeerTable pads, int rank, int world, long long shard_lo,
                               long long shard_n, float* reduced, int accumulate_out, float* scratch,
                               int scratch_elems, float* sumsq_out, int accumulate_sumsq, unsigned int epoch,
                               int done_slot, unsigned int done_epoch, int max_ctas, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && shard_n % 8 == 0 && shard_lo % 8 == 0 &&
                 bucket_off % 8 == 0 && scratch_elems >= 1 && done_slot < 480);
  cudaStream_t s = (cudaStream_t)stream;
  long long blocks = ((shard_n >> 3) + 255) / 256;
  if (max_ctas <= 0) max_ctas = mc_grads ? 128 : kNumSMs * 2;
  if (blocks < 1) blocks = 1;
  if (blocks > max_ctas) blocks = max_ctas;
  if (blocks > scratch_elems) blocks = scratch_elems;
  if (mc_grads)
    rs_sumsq_kernel<true><<<static_cast<int>(blocks), 256, 0, s>>>(
        grads, (const __nv_bfloat16*)mc_grads, bucket_off, pads, rank, world, shard_lo, shard_n, reduced,
        accumulate_out, scratch, epoch, done_slot, done_epoch);
  else
    rs_sumsq_kernel<false><<<static_cast<int>(blocks), 256, 0, s>>>(
        grads, nullptr, bucket_off, pads, rank, world, shard_lo, shard_n, reduced, accumulate_out, scratch,
        epoch, done_slot, done_epoch);
  if (sumsq_out) sumsq_fold_kernel<<<1, 256, 0, s>>>(scratch, static_cast<int>(blocks), sumsq_out, accumulate_sumsq);
  LLMCTL_RETURN_LAST();
}

// Stream-ordered wait until all `world` flags [slot, slot + world) of this rank's pad reached `epoch`.
LLMCTL_API int llmctl_signal_wait(llmctl::symm::PeerTable pads, int rank, int world, int slot,
                                  unsigned int epoch, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && slot >= 0 && slot < 480);
  signal_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(pads, rank, world, slot, epoch);
  LLMCTL_RETURN_LAST();
}

// AdamW on the local slice + publication of the updated bf16 parameters.
//   scalars      : per-rank pointers to each rank's total sum of squares (written by llmctl_rs_sumsq); ignored
//                  when total_sq_dev (device float: global sum of squares, already reduced over every group) is set
//   push         : 1 = write the slice into every rank's parameter buffer (mc_params: NVLS multicast address of
//                  that buffer or NULL for P2P stores); 0 = no remote write
//   local_out    : optional local bf16 [shard_n] copy of the updated slice (ZeRO-3 parameter shard)
//   begin_barrier: wait until every rank finished all its pulls of this step (first unit of a step; epoch_begin)
//   end_barrier  : all-rank "this unit is complete everywhere" exchange (epoch_end, fresh per launch)
LLMCTL_API int llmctl_adam_allgather(llmctl::symm::PeerTable params, long long bucket_off, void* mc_params,
                                     llmctl::symm::PeerTable pads, llmctl::symm::PeerTable scalars,
                                     const float* total_sq_dev, int rank, int world, long long shard_lo,
                                     long long shard_n, float* master, float* m, float* v,
                                     const float* reduced, void* local_out, int push, float lr, float beta1,
                                     float beta2, float eps, float weight_decay, float bias_corr1,
                                     float bias_corr2, float grad_scale, float max_norm, float* norm_out,
                                     unsigned int epoch_begin, int begin_barrier, unsigned int epoch_end,
                                     int end_barrier, int max_ctas, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && shard_n % 8 == 0 && shard_lo % 8 == 0 &&
                 bucket_off % 8 == 0);
  long long blocks = ((shard_n >> 3) + 255) / 256;
  if (max_ctas <= 0) max_ctas = kNumSMs * 2;
  if (blocks < 1) blocks = 1;
  if (blocks > max_ctas) blocks = max_ctas;
  adam_allgather_kernel<<<static_cast<int>(blocks), 256, 0, (cudaStream_t)stream>>>(
      params, bucket_off, mc_params, pads, scalars, total_sq_dev, rank, world, shard_lo, shard_n, master, m, v,
      reduced, (__nv_bfloat16*)local_out, push, lr, beta1, beta2, eps, weight_decay, bias_corr1, bias_corr2,
      grad_scale, max_norm, norm_out, epoch_begin, begin_barrier, epoch_end, end_barrier);
  LLMCTL_RETURN_LAST();
}
