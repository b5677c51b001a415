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

}  // namespace svr2
