// Launchers of the window-attention kernels (one translation unit per kernel) and the state they share.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "tmap.cuh"

namespace svr2 {

extern long long* g_trace;      // svr2_debug_attn_trace: device buffer for timeline / wait-profile builds, or nullptr
extern int g_trace_ctas;

int launch_txt_pool(const void* txt_part, void* out_txt, int n_windows, int txt_len, int heads, cudaStream_t st);

// ---- product kernels (bit-identical outputs)
// generation 4 (window_attn_v2.cu): persistent, two query tiles per CTA share every K/V tile — long sequences.
// halves = 2: generation 10, the same kernel with two threads per query row (64 key columns each), 24 warps
int launch_window_attn_tc2(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int n_windows, int H, int W, int txt_len, int heads, int64_t P, void* out_vid,
                           void* txt_part, int halves, cudaStream_t st);
// generation 8 (window_attn_v6.cu): CTA pairs, tcgen05.mma.cta_group::2 — short sequences
int launch_window_attn_tc6(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int n_windows, int H, int W, int txt_len, int heads, int64_t P, void* out_vid,
                           void* txt_part, cudaStream_t st);
// ---- test comparator (window_attn_mma.cu): legacy mma.sync kernel, never dispatched by svr2_window_attn_fwd
int launch_window_attn_mma(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int H, int W, int txt_len, int heads, void* out_vid, void* txt_part,
                           cudaStream_t st);
#ifdef SVR2_EXPERIMENTS
// ---- csrc/experiments/: earlier generations, built only with `make EXPERIMENTS=1` for A/B runs
int launch_window_attn_tc(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                          int n_work, int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                          cudaStream_t st);
int launch_window_attn_tc3(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                           cudaStream_t st);
int launch_window_attn_tc4(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                           cudaStream_t st);
int launch_window_attn_tc5(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int n_windows, int H, int W, int txt_len, int heads, int64_t P, void* out_vid,
                           void* txt_part, cudaStream_t st);
int launch_window_attn_tc7(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work4, int n_work4,
                           int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part, cudaStream_t st);
#endif

}  // namespace svr2
