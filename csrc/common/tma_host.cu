#include "tma_host.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <unordered_map>
#include <string>

namespace llmctl {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn resolve_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, []() {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct Key {
  uint64_t v[11];
  bool operator==(const Key& o) const { return memcmp(v, o.v, sizeof(v)) == 0; }
};
struct KeyHash {
  size_t operator()(const Key& k) const {
    uint64_t h = 1469598103934665603ull;
    for (int i = 0; i < 11; ++i) {
      h ^= k.v[i];
      h *= 1099511628211ull;
    }
    return static_cast<size_t>(h);
  }
};
static std::mutex g_mu;
static std::unordered_map<Key, CUtensorMap, KeyHash> g_cache;

static int encode_cached(CUtensorMap* out, const Key& key, int rank, const void* base,
                         const cuuint64_t* dims, const cuuint64_t* strides_bytes,
                         const cuuint32_t* box, int swizzle128) {
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_cache.find(key);
    if (it != g_cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  EncodeTiledFn fn = resolve_encode();
  if (!fn) return static_cast<int>(cudaErrorNotSupported);
  // Driver entry points need a current context on *this* thread (autograd worker threads may not have
  // touched the runtime yet): a no-op runtime call binds the primary context.
  // Once per thread is enough, and keeps this path free of runtime calls while a CUDA graph is being captured.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    (void)cudaFree(nullptr);
    ctx_bound = true;
  }
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), dims,
                  strides_bytes, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    if (getenv("LLMCTL_DEBUG"))
      fprintf(stderr, "[llmctl] cuTensorMapEncodeTiled failed: CUresult %d rank %d base %p dims %llu %llu box %u %u stride0 %llu\n",
              (int)r, rank, base, (unsigned long long)dims[0], (unsigned long long)dims[1], box[0], box[1],
              (unsigned long long)strides_bytes[0]);
    return static_cast<int>(cudaErrorInvalidValue);
  }
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_cache.size() > 4096) g_cache.clear();
  g_cache.emplace(key, *out);
  return 0;
}

int get_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                     uint32_t box_rows, uint32_t box_cols, int swizzle128) {
  Key key;
  memset(&key, 0, sizeof(key));
  key.v[0] = reinterpret_cast<uint64_t>(base);
  key.v[1] = rows;
  key.v[2] = cols;
  key.v[3] = ld;
  key.v[4] = box_rows;
  key.v[5] = box_cols;
  key.v[6] = swizzle128;
  key.v[7] = 2;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {box_cols, box_rows};
  return encode_cached(out, key, 2, base, dims, strides, box, swizzle128);
}

int get_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                     uint64_t s1, uint64_t s2, uint32_t b0, uint32_t b1, uint32_t b2,
                     int swizzle128) {
  Key key;
  memset(&key, 0, sizeof(key));
  key.v[0] = reinterpret_cast<uint64_t>(base);
  key.v[1] = d0;
  key.v[2] = d1;
  key.v[3] = d2;
  key.v[4] = s1;
  key.v[5] = s2;
  key.v[6] = b0;
  key.v[7] = b1;
  key.v[8] = b2;
  key.v[9] = swizzle128;
  key.v[10] = 3;
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {s1 * 2, s2 * 2};
  cuuint32_t box[3] = {b0, b1, b2};
  return encode_cached(out, key, 3, base, dims, strides, box, swizzle128);
}

}  // namespace llmctl
