// K3: RMSNorm / LayerNorm forward + backward for bf16 activations with fp32 statistics, with the
// residual add fused into the forward and the residual-gradient add fused into the backward.
//
// The reference has no norm kernel: HF modules call ATen (SURVEY.md §2.6.a "Train fwd+loss",
// reference llmctl/runtime/engine.py:284).  Baseline to beat: ATen rms_norm / layer_norm.
//
// Forward : one warp per row, the whole row lives in registers (<= 32 x 16 B per lane), one pass over
//           HBM.  h = x (+ residual) ; y = h * rsqrt(mean(h^2)+eps) * w.
// Backward: persistent CTAs loop over rows; dw/db partial sums stay in registers and are flushed once
//           per CTA into a scratch [grid, cols] fp32 buffer which a second kernel column-reduces.
#include "../common/util.cuh"

namespace llmctl {

template <int kMaxVec, bool kLayerNorm>
__global__ void __launch_bounds__(128)
norm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ residual,
                const __nv_bfloat16* __restrict__ w, const __nv_bfloat16* __restrict__ b,
                __nv_bfloat16* __restrict__ h_out, __nv_bfloat16* __restrict__ y,
                float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int cols,
                float eps) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const int nvec = cols >> 3;
  const size_t base = static_cast<size_t>(row) * cols;

  bf16x8 data[kMaxVec];
#pragma unroll
  for (int i = 0; i < kMaxVec; ++i) {
    const int v = i * 32 + lane;
    if (v < nvec) data[i] = load8(x + base + v * 8);
  }
  if (residual != nullptr) {
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i) {
      const int v = i * 32 + lane;
      if (v < nvec) {
        bf16x8 r = load8(residual + base + v * 8);
        float a[8], c[8];
        unpack8(data[i], a);
        unpack8(r, c);
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] += c[j];
        data[i] = pack8(a);  // the normalised value is computed from the bf16-rounded sum
      }
    }
  }
  if (h_out != nullptr) {
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i) {
      const int v = i * 32 + lane;
      if (v < nvec) store8(h_out + base + v * 8, data[i]);
    }
  }
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxVec; ++i) {
    const int v = i * 32 + lane;
    if (v < nvec) {
      float a[8];
      unpack8(data[i], a);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        s1 += a[j];
        s2 += a[j] * a[j];
      }
    }
  }
  float mean = 0.f, rstd;
  if (kLayerNorm) {
    s1 = warp_sum(s1);
    mean = s1 / cols;
    // second pass over registers for a numerically safer variance
    float var = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i) {
      const int v = i * 32 + lane;
      if (v < nvec) {
        float a[8];
        unpack8(data[i], a);
#pragma unroll
        for (int j = 0; j < 8; ++j) var += (a[j] - mean) * (a[j] - mean);
      }
    }
    var = warp_sum(var);
    rstd = rsqrtf(var / cols + eps);
  } else {
    s2 = warp_sum(s2);
    rstd = rsqrtf(s2 / cols + eps);
  }
  if (lane == 0) {
    if (rstd_out) rstd_out[row] = rstd;
    if (kLayerNorm && mean_out) mean_out[row] = mean;
  }
#pragma unroll
  for (int i = 0; i < kMaxVec; ++i) {
    const int v = i * 32 + lane;
    if (v < nvec) {
      float a[8], g[8], o[8];
      unpack8(data[i], a);
      unpack8(load8(w + v * 8), g);
      if (kLayerNorm && b != nullptr) {
        float bb[8];
        unpack8(load8(b + v * 8), bb);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (a[j] - mean) * rstd * g[j] + bb[j];
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (a[j] - mean) * rstd * g[j];
      }
      store8(y + base + v * 8, pack8(o));
    }
  }
}

// Backward.  blockDim = 256; each thread owns kVec 16-byte column groups: v = i*256 + tid.
template <int kVec, bool kLayerNorm>
__global__ void __launch_bounds__(256)
norm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h,
                const __nv_bfloat16* __restrict__ w, const float* __restrict__ mean_in,
                const float* __restrict__ rstd_in, const __nv_bfloat16* __restrict__ dres,
                __nv_bfloat16* __restrict__ dx, float* __restrict__ dw_partial,
                float* __restrict__ db_partial, int rows, int cols) {
  __shared__ float red[32];
  const int nvec = cols >> 3;
  float wreg[kVec][8];
  float dwacc[kVec][8];
  float dbacc[kLayerNorm ? kVec : 1][8];
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int v = i * 256 + threadIdx.x;
    if (v < nvec) unpack8(load8(w + v * 8), wreg[i]);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      dwacc[i][j] = 0.f;
      if (kLayerNorm) dbacc[i][j] = 0.f;
    }
  }
  // Software pipeline over this CTA's rows: the loads of row r+1 (dy, h) and the residual-gradient load of row r
  // are issued before the block reduction of row r, so two rows of traffic are in flight per CTA instead of one
  // load -> reduce -> load -> store chain per row (the kernel ran at 36% of HBM bandwidth when it was serial).
  bf16x8 ndy[kVec], nh[kVec];
  {
    const int row0 = blockIdx.x;
    if (row0 < rows) {
      const size_t base0 = static_cast<size_t>(row0) * cols;
#pragma unroll
      for (int i = 0; i < kVec; ++i) {
        const int v = i * 256 + threadIdx.x;
        if (v < nvec) {
          ndy[i] = load8(dy + base0 + v * 8);
          nh[i] = load8(h + base0 + v * 8);
        }
      }
    }
  }
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = static_cast<size_t>(row) * cols;
    const float rstd = rstd_in[row];
    const float mean = kLayerNorm ? mean_in[row] : 0.f;
    float g[kVec][8], xh[kVec][8];
    bf16x8 rres[kVec];
    float c1 = 0.f, c2 = 0.f;
    if (dres != nullptr) {
#pragma unroll
      for (int i = 0; i < kVec; ++i) {
        const int v = i * 256 + threadIdx.x;
        if (v < nvec) rres[i] = load8(dres + base + v * 8);
      }
    }
#pragma unroll
    for (int i = 0; i < kVec; ++i) {
      const int v = i * 256 + threadIdx.x;
      if (v < nvec) {
        float d[8], a[8];
        unpack8(ndy[i], d);
        unpack8(nh[i], a);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[i][j] = (a[j] - mean) * rstd;
          g[i][j] = d[j] * wreg[i][j];
          c1 += g[i][j] * xh[i][j];
          c2 += g[i][j];
          dwacc[i][j] += d[j] * xh[i][j];
          if (kLayerNorm) dbacc[i][j] += d[j];
        }
      }
    }
    {
      const int nrow = row + gridDim.x;
      if (nrow < rows) {
        const size_t nbase = static_cast<size_t>(nrow) * cols;
#pragma unroll
        for (int i = 0; i < kVec; ++i) {
          const int v = i * 256 + threadIdx.x;
          if (v < nvec) {
            ndy[i] = load8(dy + nbase + v * 8);
            nh[i] = load8(h + nbase + v * 8);
          }
        }
      }
    }
    c1 = block_sum(c1, red) / cols;
    if (kLayerNorm) c2 = block_sum(c2, red) / cols; else c2 = 0.f;
#pragma unroll
    for (int i = 0; i < kVec; ++i) {
      const int v = i * 256 + threadIdx.x;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (g[i][j] - xh[i][j] * c1 - c2);
        if (dres != nullptr) {
          float r[8];
          unpack8(rres[i], r);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] += r[j];
        }
        store8(dx + base + v * 8, pack8(o));
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kVec; ++i) {
    const int v = i * 256 + threadIdx.x;
    if (v < nvec) {
      float* p = dw_partial + static_cast<size_t>(blockIdx.x) * cols + v * 8;
      *reinterpret_cast<float4*>(p) = make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
      *reinterpret_cast<float4*>(p + 4) = make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
      if (kLayerNorm && db_partial != nullptr) {
        float* q = db_partial + static_cast<size_t>(blockIdx.x) * cols + v * 8;
        *reinterpret_cast<float4*>(q) = make_float4(dbacc[i][0], dbacc[i][1], dbacc[i][2], dbacc[i][3]);
        *reinterpret_cast<float4*>(q + 4) = make_float4(dbacc[i][4], dbacc[i][5], dbacc[i][6], dbacc[i][7]);
      }
    }
  }
}

// Column-reduce the [nparts, cols] fp32 partials; writes (or accumulates into) an fp32 gradient.
__global__ void colsum_kernel(const float* __restrict__ partial, float* __restrict__ out, int nparts,
                              int cols, int accumulate) {
  // blockDim = (32, 8): 32 consecutive columns x 8 row-slices
  __shared__ float tile[8][33];
  const int col = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (col < cols)
    for (int r = threadIdx.y; r < nparts; r += 8) s += partial[static_cast<size_t>(r) * cols + col];
  tile[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && col < cols) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) t += tile[r][threadIdx.x];
    out[col] = accumulate ? out[col] + t : t;
  }
}

template <bool kLN>
static int launch_fwd(const void* x, const void* residual, const void* w, const void* b, void* h_out,
                      void* y, float* mean, float* rstd, int rows, int cols, float eps,
                      cudaStream_t s) {
  LLMCTL_REQUIRE(cols % 8 == 0 && cols <= 8192 && rows >= 0);
  if (rows == 0) return 0;
  const int nvec = cols / 8;
  const int per_lane = (nvec + 31) / 32;
  dim3 block(128), grid((rows + 3) / 4);
#define LAUNCH(MV)                                                                              \
  norm_fwd_kernel<MV, kLN><<<grid, block, 0, s>>>(                                             \
      (const __nv_bfloat16*)x, (const __nv_bfloat16*)residual, (const __nv_bfloat16*)w,        \
      (const __nv_bfloat16*)b, (__nv_bfloat16*)h_out, (__nv_bfloat16*)y, mean, rstd, rows, cols, \
      eps)
  if (per_lane <= 4) LAUNCH(4);
  else if (per_lane <= 8) LAUNCH(8);
  else if (per_lane <= 16) LAUNCH(16);
  else if (per_lane <= 20) LAUNCH(20);
  else LAUNCH(32);
#undef LAUNCH
  LLMCTL_RETURN_LAST();
}

template <bool kLN>
static int launch_bwd(const void* dy, const void* h, const void* w, const float* mean,
                      const float* rstd, const void* dres, void* dx, float* dw, float* db,
                      float* scratch, int scratch_rows, int rows, int cols, int accumulate,
                      cudaStream_t s) {
  LLMCTL_REQUIRE(cols % 8 == 0 && cols <= 8192 && scratch_rows >= 1);
  if (rows == 0) return 0;
  const int nvec = cols / 8;
  const int kv = (nvec + 255) / 256;
  int grid = scratch_rows < rows ? scratch_rows : rows;
  float* dwp = scratch;
  float* dbp = kLN ? scratch + static_cast<size_t>(scratch_rows) * cols : nullptr;
#define LAUNCH(KV)                                                                            \
  norm_bwd_kernel<KV, kLN><<<grid, 256, 0, s>>>(                                               \
      (const __nv_bfloat16*)dy, (const __nv_bfloat16*)h, (const __nv_bfloat16*)w, mean, rstd, \
      (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, dwp, dbp, rows, cols)
  if (kv <= 1) LAUNCH(1);
  else if (kv <= 2) LAUNCH(2);
  else if (kv <= 3) LAUNCH(3);
  else LAUNCH(4);
#undef LAUNCH
  dim3 b2(32, 8), g2((cols + 31) / 32);
  colsum_kernel<<<g2, b2, 0, s>>>(dwp, dw, grid, cols, accumulate);
  if (kLN && db != nullptr) colsum_kernel<<<g2, b2, 0, s>>>(dbp, db, grid, cols, accumulate);
  LLMCTL_RETURN_LAST();
}

}  // namespace llmctl

using namespace llmctl;

LLMCTL_API int llmctl_rmsnorm_fwd(const void* x, const void* residual, const void* w, void* h_out,
                                  void* y, float* rstd, int rows, int cols, float eps,
                                  void* stream) {
  LLMCTL_BEGIN();
  return launch_fwd<false>(x, residual, w, nullptr, h_out, y, nullptr, rstd, rows, cols, eps,
                           (cudaStream_t)stream);
}
LLMCTL_API int llmctl_rmsnorm_bwd(const void* dy, const void* h, const void* w, const float* rstd,
                                  const void* dres, void* dx, float* dw, float* scratch,
                                  int scratch_rows, int rows, int cols, int accumulate,
                                  void* stream) {
  LLMCTL_BEGIN();
  return launch_bwd<false>(dy, h, w, nullptr, rstd, dres, dx, dw, nullptr, scratch, scratch_rows,
                           rows, cols, accumulate, (cudaStream_t)stream);
}
LLMCTL_API int llmctl_layernorm_fwd(const void* x, const void* residual, const void* w,
                                    const void* b, void* h_out, void* y, float* mean, float* rstd,
                                    int rows, int cols, float eps, void* stream) {
  LLMCTL_BEGIN();
  return launch_fwd<true>(x, residual, w, b, h_out, y, mean, rstd, rows, cols, eps,
                          (cudaStream_t)stream);
}
LLMCTL_API int llmctl_layernorm_bwd(const void* dy, const void* h, const void* w, const float* mean,
                                    const float* rstd, const void* dres, void* dx, float* dw,
                                    float* db, float* scratch, int scratch_rows, int rows, int cols,
                                    int accumulate, void* stream) {
  LLMCTL_BEGIN();
  return launch_bwd<true>(dy, h, w, mean, rstd, dres, dx, dw, db, scratch, scratch_rows, rows, cols,
                          accumulate, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------
// Allocation-free RMSNorm backward.
//
// Who this is for: programs that reach this library through the C ABI and cannot obtain device memory at
// call time.  Two real cases: (1) a host runtime that hands its whole device budget to its own arena at
// start-up (an inference server, a simulator embedding) so that a later cudaMalloc would either fail or
// fragment the arena; (2) a control loop with a latency bound, where cudaMalloc/cudaFree are off limits
// because they take the driver's memory lock and may synchronise the device.  llmctl_rmsnorm_bwd needs an
// fp32 scratch area for the per-CTA partial sums of dw and leaves its provision to the caller; this variant
// carries the scratch inside the library image instead.
//
// The scratch is the file-scope array below: kNoAllocCtas x kNoAllocMaxCols floats = 2 MiB of global memory
// that every GPU which loads this library keeps reserved until the process exits.
//
// Set-up: on the first call made on a given device (with that device current) the entry point asks the
// runtime for the array's device address with cudaGetSymbolAddress, checks the result and caches it per
// device.  In host code the array's name is only a stand-in for the symbol, not a device address, so it is
// never handed to the kernel.  Make that first call outside any stream capture.
//
// Limits and sharing: cols <= 8192 (and a multiple of 8); larger rows are rejected with an error message
// (llmctl_noalloc_last_error) and cudaErrorInvalidValue.  rows is unlimited (CTAs loop over rows).  All calls
// on a device share the one array and it keeps whatever the last call left in it.  Nothing in it has to start
// at zero: each CTA overwrites its own partial row in full before the column reduction reads exactly those
// rows.  Two calls whose kernels overlap on the GPU (different streams, no dependency) race on the array and
// produce wrong dw; serialise them.
//
// No cudaMalloc / cudaFree / allocating library call is made here.
// ------------------------------------------------------------------------------------------------
#include <mutex>

namespace llmctl {
constexpr int kNoAllocCtas = 64;
constexpr int kNoAllocMaxCols = 8192;
constexpr int kNoAllocMaxDevices = 64;
static __device__ __align__(256) float g_rmsnorm_bwd_scratch[kNoAllocCtas * kNoAllocMaxCols];
static float* g_noalloc_scratch_addr[kNoAllocMaxDevices] = {nullptr};
static std::mutex g_noalloc_mu;
static thread_local char g_noalloc_err[256] = "";
}  // namespace llmctl

// Text of the last error raised by an allocation-free entry point on this thread ("" if none).
LLMCTL_API const char* llmctl_noalloc_last_error() { return llmctl::g_noalloc_err; }

LLMCTL_API int llmctl_rmsnorm_bwd_noalloc(const void* dy, const void* h, const void* w, const float* rstd,
                                          const void* dres, void* dx, float* dw, int rows, int cols,
                                          int accumulate, void* stream) {
  LLMCTL_BEGIN();
  g_noalloc_err[0] = '\0';
  if (cols <= 0 || cols % 8 != 0 || cols > kNoAllocMaxCols || rows < 0) {
    snprintf(g_noalloc_err, sizeof(g_noalloc_err),
             "llmctl_rmsnorm_bwd_noalloc: cols=%d rows=%d not supported: cols must be a positive multiple of 8 "
             "and at most %d (the built-in scratch holds %d x %d floats)",
             cols, rows, kNoAllocMaxCols, kNoAllocCtas, kNoAllocMaxCols);
    fprintf(stderr, "[llmctl] %s\n", g_noalloc_err);
    return static_cast<int>(cudaErrorInvalidValue);
  }
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kNoAllocMaxDevices) {
    snprintf(g_noalloc_err, sizeof(g_noalloc_err), "llmctl_rmsnorm_bwd_noalloc: cannot identify the current device (%d)", (int)e);
    return static_cast<int>(e != cudaSuccess ? e : cudaErrorInvalidDevice);
  }
  float* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (g_noalloc_scratch_addr[dev] == nullptr) {
      // once per device, with that device current: resolve the symbol to its address on THIS device
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_rmsnorm_bwd_scratch);
      if (e != cudaSuccess || p == nullptr) {
        snprintf(g_noalloc_err, sizeof(g_noalloc_err),
                 "llmctl_rmsnorm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        fprintf(stderr, "[llmctl] %s\n", g_noalloc_err);
        return static_cast<int>(e != cudaSuccess ? e : cudaErrorInvalidSymbol);
      }
      g_noalloc_scratch_addr[dev] = static_cast<float*>(p);
    }
    scratch = g_noalloc_scratch_addr[dev];
  }
  return launch_bwd<false>(dy, h, w, nullptr, rstd, dres, dx, dw, nullptr, scratch, kNoAllocCtas, rows, cols,
                           accumulate, (cudaStream_t)stream);
}
