// Host-side helper: build (and cache) CUtensorMap descriptors without linking against libcuda.
// cuTensorMapEncodeTiled is resolved through cudaGetDriverEntryPoint at first use so that the
// library also loads on machines that have no driver (CPU build check).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace llmctl {

// 2D bf16 tensor map over a row-major [rows, cols] matrix with leading dimension `ld` (elements).
// Box = [box_rows, box_cols]; swizzle 128B when box_cols * 2 == 128, none otherwise.
// Returns 0 on success.  Thread-safe; descriptors are cached by value of all arguments.
int get_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                     uint32_t box_rows, uint32_t box_cols, int swizzle128);

// 3D bf16 tensor map: dims (inner -> outer) = {d0, d1, d2} with strides (elements) {1, s1, s2}.
int get_tmap_3d_bf16(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                     uint64_t s1, uint64_t s2, uint32_t b0, uint32_t b1, uint32_t b2,
                     int swizzle128);

}  // namespace llmctl
