// Host-side TMA descriptor construction. The driver entry point is resolved at run time through the
// CUDA runtime, so the library has no link-time dependency on libcuda (it must load on CPU-only boxes).
#pragma once
#include <cstdint>
#include <cuda.h>

namespace b2 {
// Row-major bf16 matrix [outer][inner], 128B swizzle, zero fill for out-of-bounds boxes.
// Returns 0 on success, a CUresult/cudaError value otherwise.
int make_tmap_2d_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint32_t box_inner,
                      uint32_t box_outer);
// The same matrix as `panels` side-by-side 64-column panels in ONE box: a 3-D view (64 | rows | inner/64) whose box
// [64][box_outer][panels] lands in shared memory as `panels` consecutive [box_outer x 128 B] swizzled blocks - the
// MN-major operand layout of tcgen05.mma, fetched with a single TMA instead of one per 64 columns.
// inner must be a multiple of 64. Coordinates: (0, row, column / 64).
int make_tmap_panels_bf16(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer, uint32_t box_outer,
                          uint32_t panels);
// General tiled bf16 map: dims/box innermost first, strides_bytes[i] is the byte stride of dimension i+1 (any multiple
// of 16, overlapping windows allowed), swizzle_bytes in {0, 32, 64, 128}, zero fill out of bounds.
int make_tmap_tiled_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);
// Tiled map over 8-byte elements, no swizzle (raw row copies; inner box <= 256 elements = 2 KB).
int make_tmap_tiled_u64(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                        const uint64_t* strides_bytes, const uint32_t* box);
// im2col view of an NHWC bf16 tensor [N][H][W][C] (128B swizzle, zero fill): boxes of `channels` x `pixels`,
// bounding box corners as in cuTensorMapEncodeIm2col (lower = -pad, upper = pad - (K-1)*dilation), traversal stride.
int make_tmap_im2col_bf16(CUtensorMap* out, const void* base, int N, int H, int W, int C, int lower_w, int lower_h,
                          int upper_w, int upper_h, int channels, int pixels, int stride);
}  // namespace b2
