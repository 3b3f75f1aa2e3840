// K7: ZeRO gradient reduce-scatter + AdamW + parameter all-gather over NVLink peer memory, without NCCL.
//
// Every data-parallel rank keeps its bf16 gradient bucket and its bf16 parameter buffer in symmetric memory.
//   kernel 1  rs_sumsq      : each rank PULLS its 1/dp slice of the bucket from every peer (16-byte P2P loads),
//                             sums in fp32 in a fixed rank order (deterministic), stores the reduced slice
//                             locally (fp32) and produces sum-of-squares partials for the global grad norm.
//                             Launched per unit from backward as soon as the unit's last gradient landed, so the
//                             transfer overlaps the rest of backward.
//   kernel 2  adam_allgather: global-norm clip coefficient from every rank's sum of squares (read from peers),
//                             AdamW on the local slice (fp32 master + moments), then the updated bf16 parameters
//                             are PUSHED into every rank's parameter buffer (P2P stores, or one NVLS
//                             multimem.st when a multicast mapping is available).
// Cross-GPU ordering uses epoch-numbered flags in the symmetric signal pads (csrc/comm/symm.cuh): kernel 1
// waits until every peer's backward for that unit has finished; kernel 2 waits until every peer has finished
// all its pulls (so buckets may be overwritten by the next backward) and ends with an all-rank barrier after
// which every parameter buffer is complete.
//
// Reference: NCCL all-reduce inside DDP + torch.optim.AdamW + clip_grad_norm_
// (reference llmctl/runtime/engine.py:265-267,289,295-303).  SURVEY.md §2.6.b K7.
#include "../comm/symm.cuh"
#include "../common/util.cuh"

namespace llmctl {
using namespace symm;

// slot layout inside each rank's signal pad (uint32 indices)
constexpr int kSlotGradReady = 0;    // [0, 8)   peer r finished writing its bucket for launch #epoch
constexpr int kSlotPullDone = 8;     // [8, 16)  peer r finished all pulls of this step
constexpr int kSlotParamsDone = 16;  // [16, 24) peer r finished pushing parameters

__global__ void __launch_bounds__(256)
rs_sumsq_kernel(PeerTable grads, long long bucket_off, PeerTable pads, int rank, int world,
                long long shard_lo, long long shard_n, float* __restrict__ reduced,
                float* __restrict__ partials, uint32_t epoch) {
  __shared__ float red[32];
  // announce "my bucket is complete" (block 0), then everybody waits for all peers
  if (blockIdx.x == 0 && threadIdx.x < world)
    signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + kSlotGradReady + rank, epoch);
  if (threadIdx.x < world)
    signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + kSlotGradReady + threadIdx.x, epoch);
  __syncthreads();

  const __nv_bfloat16* src[kMaxRanks];
#pragma unroll
  for (int r = 0; r < kMaxRanks; ++r)
    src[r] = (r < world) ? reinterpret_cast<const __nv_bfloat16*>(grads.ptr[r]) + bucket_off + shard_lo : nullptr;

  float acc_sq = 0.f;
  const long long nvec = shard_n >> 3;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
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
    float* dst = reduced + (i << 3);
    *reinterpret_cast<float4*>(dst) = make_float4(s[0], s[1], s[2], s[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(s[4], s[5], s[6], s[7]);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc_sq += s[j] * s[j];
  }
  acc_sq = block_sum(acc_sq, red);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc_sq;
}

__global__ void __launch_bounds__(256)
sumsq_fold_kernel(const float* __restrict__ partials, int n, float* __restrict__ out, int accumulate) {
  __shared__ float red[32];
  float a = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) a += partials[i];
  a = block_sum(a, red);
  if (threadIdx.x == 0) *out = accumulate ? *out + a : a;
}

// scalars: symmetric float buffer; scalars[rank_slot] of each rank = that rank's total sum of squares.
__global__ void __launch_bounds__(256)
adam_allgather_kernel(PeerTable params, long long bucket_off, void* mc_params, PeerTable pads,
                      PeerTable scalars, int rank, int world, long long shard_lo, long long shard_n,
                      float* __restrict__ master, float* __restrict__ m, float* __restrict__ v,
                      const float* __restrict__ reduced, float lr, float beta1, float beta2, float eps,
                      float wd, float bc1, float bc2, float grad_scale, float max_norm,
                      float* __restrict__ norm_out, uint32_t epoch, int begin_barrier, int end_barrier) {
  if (begin_barrier) {
    if (blockIdx.x == 0 && threadIdx.x < world)
      signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + kSlotPullDone + rank, epoch);
    if (threadIdx.x < world)
      signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + kSlotPullDone + threadIdx.x, epoch);
    __syncthreads();
  }
  // global gradient norm: every rank reduced a disjoint slice, so the squares simply add up
  float total = 0.f;
  for (int r = 0; r < world; ++r)
    total += __uint_as_float(ptx::ld_relaxed_sys(reinterpret_cast<const uint32_t*>(scalars.ptr[r])));
  const float norm = sqrtf(total) * grad_scale;
  float coef = 1.f;
  if (max_norm > 0.f) coef = fminf(1.f, max_norm / (norm + 1e-6f));
  if (blockIdx.x == 0 && threadIdx.x == 0 && norm_out) *norm_out = norm;
  const float scale = grad_scale * coef;
  const bool skip = !isfinite(norm);

  __nv_bfloat16* dst[kMaxRanks];
#pragma unroll
  for (int r = 0; r < kMaxRanks; ++r)
    dst[r] = (r < world) ? reinterpret_cast<__nv_bfloat16*>(params.ptr[r]) + bucket_off + shard_lo : nullptr;
  __nv_bfloat16* mc = mc_params ? reinterpret_cast<__nv_bfloat16*>(mc_params) + bucket_off + shard_lo : nullptr;

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
      if (mc != nullptr) {
        ptx::multimem_st_v4(mc + o, bits);                           // one store, the switch replicates it
      } else {
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
      unsigned int* counter = reinterpret_cast<unsigned int*>(pads.ptr[rank]) + 24;   // local CTA counter slot
      const unsigned int done = atomicAdd(counter, 1u) + 1u;
      s_last = (done == gridDim.x) ? 1u : 0u;
      if (s_last) atomicExch(counter, 0u);        // leave the counter clean for the next launch
    }
    __syncthreads();
    if (s_last) {
      if (threadIdx.x < world)
        signal_set(reinterpret_cast<uint32_t*>(pads.ptr[threadIdx.x]) + kSlotParamsDone + rank, epoch);
      if (threadIdx.x < world)
        signal_wait_ge(reinterpret_cast<const uint32_t*>(pads.ptr[rank]) + kSlotParamsDone + threadIdx.x, epoch);
    }
  }
}

}  // namespace llmctl

using namespace llmctl;
LLMCTL_SYMM_WATCHDOG_MODULE(rs_adam)

// Pull-reduce this rank's [shard_lo, shard_lo + shard_n) slice of a bf16 bucket from every rank.
//   grads      : per-rank base pointers of the symmetric gradient buffer; bucket_off = unit offset (elements)
//   reduced    : local fp32 [shard_n] output;  scratch: fp32 [scratch_elems] partials
//   sumsq_out  : local float, (+)= sum of squares of the reduced slice (lives in the symmetric scalar buffer)
//   epoch      : launch counter of this entry point (same sequence on every rank)
LLMCTL_API int llmctl_rs_sumsq(llmctl::symm::PeerTable grads, long long bucket_off,
                               llmctl::symm::PeerTable pads, int rank, int world, long long shard_lo,
                               long long shard_n, float* reduced, float* scratch, int scratch_elems,
                               float* sumsq_out, int accumulate, unsigned int epoch, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && shard_n % 8 == 0 && shard_lo % 8 == 0 &&
                 bucket_off % 8 == 0 && scratch_elems >= 1);
  cudaStream_t s = (cudaStream_t)stream;
  long long blocks = ((shard_n >> 3) + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > kNumSMs * 2) blocks = kNumSMs * 2;
  if (blocks > scratch_elems) blocks = scratch_elems;
  rs_sumsq_kernel<<<static_cast<int>(blocks), 256, 0, s>>>(grads, bucket_off, pads, rank, world, shard_lo,
                                                           shard_n, reduced, scratch, epoch);
  sumsq_fold_kernel<<<1, 256, 0, s>>>(scratch, static_cast<int>(blocks), sumsq_out, accumulate);
  LLMCTL_RETURN_LAST();
}

// AdamW on the local slice + push of the updated bf16 parameters into every rank's parameter buffer.
//   scalars    : per-rank pointers to each rank's total sum of squares (written by llmctl_rs_sumsq)
//   mc_params  : NVLS multicast address of the parameter buffer, or NULL for P2P stores
//   begin_barrier / end_barrier : set on the first / last unit of a step (see file header)
LLMCTL_API int llmctl_adam_allgather(llmctl::symm::PeerTable params, long long bucket_off, void* mc_params,
                                     llmctl::symm::PeerTable pads, llmctl::symm::PeerTable scalars,
                                     int rank, int world, long long shard_lo, long long shard_n,
                                     float* master, float* m, float* v, const float* reduced, float lr,
                                     float beta1, float beta2, float eps, float weight_decay,
                                     float bias_corr1, float bias_corr2, float grad_scale, float max_norm,
                                     float* norm_out, unsigned int epoch, int begin_barrier,
                                     int end_barrier, void* stream) {
  LLMCTL_BEGIN();
  LLMCTL_REQUIRE(world >= 1 && world <= symm::kMaxRanks && shard_n % 8 == 0 && shard_lo % 8 == 0 &&
                 bucket_off % 8 == 0);
  long long blocks = ((shard_n >> 3) + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > kNumSMs * 2) blocks = kNumSMs * 2;
  adam_allgather_kernel<<<static_cast<int>(blocks), 256, 0, (cudaStream_t)stream>>>(
      params, bucket_off, mc_params, pads, scalars, rank, world, shard_lo, shard_n, master, m, v, reduced,
      lr, beta1, beta2, eps, weight_decay, bias_corr1, bias_corr2, grad_scale, max_norm, norm_out, epoch,
      begin_barrier, end_barrier);
  LLMCTL_RETURN_LAST();
}
