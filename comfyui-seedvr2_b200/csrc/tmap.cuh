// Host-side TMA tensor-map builders shared by the attention and GEMM launchers (cuTensorMapEncodeTiled through the runtime's
// driver entry point: the library links only libcudart).
#pragma once
#include <cuda.h>
#include <stdint.h>

namespace svr2 {

// [P, heads, 128] bf16 viewed as a 3-D tensor (d, head, row); box = 64 d x 1 head x box_rows rows, 128-byte swizzle.
int make_tmap_box(CUtensorMap* m, const void* ptr, int64_t P, int heads, int box_rows);
inline int make_tmap(CUtensorMap* m, const void* ptr, int64_t P, int heads) { return make_tmap_box(m, ptr, P, heads, 128); }
// Row-major [rows, cols] bf16 matrix with a row pitch of ld elements; box = 64 cols x box_rows rows, 128-byte swizzle (the
// canonical K-major UMMA operand layout).  Out-of-range rows / cols are zero-filled by the TMA unit.
int make_tmap_2d(CUtensorMap* m, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows);
// Channels-last [T, H, W, C] tensor of 2-byte elements viewed as (c, w, h, t); box = 64 c x box_w x box_h x 1 frame, 128-byte
// swizzle: rows of the box (w fastest) land as consecutive 128-byte rows — a K-major UMMA operand of box_w * box_h positions.
// Coordinates may be negative / past the edge: the TMA unit zero-fills on load (the convolution's zero padding), clips on store.
// stride > 1: the box samples every stride-th row / column (TMA element strides): box_w x box_h positions are still what lands.
int make_tmap_thwc(CUtensorMap* m, const void* ptr, int64_t T, int64_t H, int64_t W, int64_t C, int box_w, int box_h, int stride = 1);

}  // namespace svr2
