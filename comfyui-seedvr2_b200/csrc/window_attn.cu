// svr2_window_attn_fwd — host side: kernel selection, work lists, debug hook.  The kernels live in one file each:
//   window_attn_v6.cu   generation 8: CTA pairs (tcgen05.mma.cta_group::2), one query tile per CTA — short sequences
//   window_attn_v2.cu   generation 4: persistent CTA per SM, two query tiles share every K/V tile — long sequences
//   window_attn_mma.cu  legacy mma.sync kernel: a TEST COMPARATOR only (svr2_window_attn_fwd_impl with SVR2_ATTN_MMA)
//   experiments/        generations 1, 5, 6, 7, 9: compiled only with `make EXPERIMENTS=1` (A/B history)
//
// Replaces FlashAttentionVarlen / pytorch_varlen_attention (reference src/models/dit_3b/attention.py:24-61,138-157: a
// Python loop of one SDPA call per window), unconcat_win (na.py:396-417) and window_reverse (na.py:638).
// The kernel is chosen from the launch arguments alone — no environment variable is read on the product path.
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "attn_launch.cuh"
#include "common.cuh"

namespace svr2 {
long long* g_trace = nullptr;      // svr2_debug_attn_trace
int g_trace_ctas = 0;
}  // namespace svr2

extern "C" int svr2_debug_attn_trace(void* trace_buf, int n_ctas) {
    svr2::g_trace = (long long*)trace_buf;
    svr2::g_trace_ctas = trace_buf ? n_ctas : 0;
    return 64;      // slots per CTA (kTraceSlots of the kernels)
}

extern "C" int svr2_window_attn_worklist(const int32_t* win_table_host, int n_windows, int txt_len, int32_t* work,
                                         int max_items, int64_t* n_items) {
    return svr2_window_attn_worklist_tiles(win_table_host, n_windows, txt_len, 2, work, max_items, n_items);
}

extern "C" int svr2_window_attn_worklist_tiles(const int32_t* win_table_host, int n_windows, int txt_len, int tiles_per_item,
                                               int32_t* work, int max_items, int64_t* n_items) {
    SVR2_CHECK_ARG(win_table_host && n_items && n_windows > 0 && txt_len >= 0 && tiles_per_item > 0 && tiles_per_item <= 64,
                   "svr2_window_attn_worklist: bad arguments");
    const int rows_per_item = 128 * tiles_per_item;
    struct Item { int seq, win, qp; };
    std::vector<Item> items;
    for (int i = 0; i < n_windows; ++i) {
        const int seq = win_table_host[i * SVR2_WIN_FIELDS + 7] + txt_len;
        for (int qp = 0; qp * rows_per_item < seq; ++qp) items.push_back({seq, i, qp});
    }
    // longest sequences first (LPT): the ragged shifted-window remnants fill the tail of the last wave
    std::stable_sort(items.begin(), items.end(), [](const Item& a, const Item& b) { return a.seq > b.seq; });
    *n_items = (int64_t)items.size();
    if (work == nullptr) return SVR2_OK;
    SVR2_CHECK_ARG((int64_t)items.size() <= max_items, "svr2_window_attn_worklist: %zu items do not fit %d", items.size(), max_items);
    for (size_t i = 0; i < items.size(); ++i) {
        work[2 * i] = items[i].win;
        work[2 * i + 1] = items[i].qp;
    }
    return SVR2_OK;
}

extern "C" int svr2_window_attn_fwd_impl(const void* q_pack, const void* k_pack, const void* v_pack, const int32_t* win_table,
                                         const int32_t* work, int n_work, const int32_t* work4, int n_work4, int n_windows,
                                         int T, int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                                         void* out_txt, int impl, void* stream) {
    SVR2_CHECK_ARG(q_pack && k_pack && v_pack && win_table && work && out_vid, "svr2_window_attn_fwd: null pointer");
    SVR2_CHECK_ARG(txt_len == 0 || txt_part, "svr2_window_attn_fwd: text workspace missing");
    SVR2_CHECK_ARG(n_work > 0 && n_windows > 0 && heads > 0 && heads <= 65535 && P > 0 && P < ((int64_t)1 << 31),
                   "svr2_window_attn_fwd: bad sizes");
    (void)T; (void)work4; (void)n_work4;
    cudaStream_t st = (cudaStream_t)stream;
    // Product path, two kernels with identical output bits (tests/test_gpu_kernels.py::test_window_attn_generations_bit_identical):
    //  * sequences of up to ~4 key tiles (stills, small windows: mean packed length <= 576 rows): generation 8 — in the
    //    power-capped C1 step 151.5 us against 163.7 us per launch (round 1, run 42);
    //  * longer sequences: generation 4 — 2 % ahead at C2, equal at the C3 per-rank size.
    if (impl == SVR2_ATTN_AUTO) impl = (P <= (int64_t)576 * n_windows) ? SVR2_ATTN_GEN8 : SVR2_ATTN_GEN4;
    int rc;
    switch (impl) {
    case SVR2_ATTN_GEN8:
        rc = svr2::launch_window_attn_tc6(q_pack, k_pack, v_pack, win_table, work, n_work, n_windows, H, W, txt_len, heads, P, out_vid, txt_part, st);
        break;
    case SVR2_ATTN_GEN4:
        rc = svr2::launch_window_attn_tc2(q_pack, k_pack, v_pack, win_table, work, n_work, n_windows, H, W, txt_len, heads, P, out_vid, txt_part, 1, st);
        break;
    case SVR2_ATTN_GEN10:
        rc = svr2::launch_window_attn_tc2(q_pack, k_pack, v_pack, win_table, work, n_work, n_windows, H, W, txt_len, heads, P, out_vid, txt_part, 2, st);
        break;
    case SVR2_ATTN_MMA:
        rc = svr2::launch_window_attn_mma(q_pack, k_pack, v_pack, win_table, work, n_work, H, W, txt_len, heads, out_vid, txt_part, st);
        break;
#ifdef SVR2_EXPERIMENTS
    case 1: rc = svr2::launch_window_attn_tc(q_pack, k_pack, v_pack, win_table, work, n_work, H, W, txt_len, heads, P, out_vid, txt_part, st); break;
    case 5: rc = svr2::launch_window_attn_tc3(q_pack, k_pack, v_pack, win_table, work, n_work, H, W, txt_len, heads, P, out_vid, txt_part, st); break;
    case 6: rc = svr2::launch_window_attn_tc4(q_pack, k_pack, v_pack, win_table, work, n_work, H, W, txt_len, heads, P, out_vid, txt_part, st); break;
    case 7: rc = svr2::launch_window_attn_tc5(q_pack, k_pack, v_pack, win_table, work, n_work, n_windows, H, W, txt_len, heads, P, out_vid, txt_part, st); break;
    case 9:
        SVR2_CHECK_ARG(work4 && n_work4 > 0, "svr2_window_attn_fwd_impl: generation 9 needs the 4-tile work list");
        rc = svr2::launch_window_attn_tc7(q_pack, k_pack, v_pack, win_table, work4, n_work4, H, W, txt_len, heads, P, out_vid, txt_part, st);
        break;
#endif
    default:
        svr2::set_error("svr2_window_attn_fwd_impl: kernel %d is not part of this build (experiments need `make EXPERIMENTS=1`)", impl);
        return SVR2_E_UNSUPPORTED;
    }
    if (rc != SVR2_OK) return rc;
    if (out_txt == nullptr) return SVR2_OK;      // caller pools the text rows itself with svr2_txt_pool_fwd
    return svr2::launch_txt_pool(txt_part, out_txt, n_windows, txt_len, heads, st);
}

extern "C" int svr2_window_attn_fwd(const void* q_pack, const void* k_pack, const void* v_pack, const int32_t* win_table,
                                    const int32_t* work, int n_work, const int32_t* work4, int n_work4, int n_windows, int T,
                                    int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                                    void* out_txt, void* stream) {
    return svr2_window_attn_fwd_impl(q_pack, k_pack, v_pack, win_table, work, n_work, work4, n_work4, n_windows, T, H, W, txt_len,
                                     heads, P, out_vid, txt_part, out_txt, SVR2_ATTN_AUTO, stream);
}

extern "C" int svr2_txt_pool_fwd(const void* txt_part, void* out_txt, int n_windows, int txt_len, int heads, void* stream) {
    SVR2_CHECK_ARG(txt_part && out_txt && n_windows > 0 && txt_len >= 0 && heads > 0, "svr2_txt_pool_fwd: bad arguments");
    return svr2::launch_txt_pool(txt_part, out_txt, n_windows, txt_len, heads, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------------------------------
// svr2_window_attn_fwd_noalloc — svr2_window_attn_fwd for callers that cannot allocate (CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The per-window text rows (the `txt_part` workspace of svr2_window_attn_fwd) live in this file-scope device array instead of
// caller memory.  Every GPU that loads the library holds the array until the process exits.  All calls on a device share it:
// calls that overlap on the GPU (different streams) race.  Nothing in it has to start at zero — the attention kernel writes
// every (window, text row, head) slot it later pools — so the previous call's contents are simply overwritten.
static __device__ __align__(256) unsigned char g_noalloc_txt_part[SVR2_NOALLOC_SCRATCH_BYTES];

namespace svr2 {
namespace {
// The array's name is only a host-side stand-in in host code; the device address is looked up once per device, with that
// device current, and cached.
int noalloc_scratch(void** out) {
    static void* ptr[kMaxDevices] = {};
    const int dev = cur_device();
    if (ptr[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_txt_part);
        if (e != cudaSuccess || p == nullptr) {
            set_error("svr2_window_attn_fwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return SVR2_E_CUDA;
        }
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return SVR2_OK;
}
}  // namespace
}  // namespace svr2

extern "C" int svr2_window_attn_fwd_noalloc(const void* q_pack, const void* k_pack, const void* v_pack, const int32_t* win_table,
                                            const int32_t* work, int n_work, int n_windows, int T, int H, int W, int txt_len,
                                            int heads, int64_t P, void* out_vid, void* out_txt, void* stream) {
    SVR2_CHECK_ARG(n_windows > 0 && heads > 0 && txt_len >= 0, "svr2_window_attn_fwd_noalloc: bad sizes");
    SVR2_CHECK_ARG(txt_len == 0 || out_txt, "svr2_window_attn_fwd_noalloc: out_txt is required (the pooling runs inside the call)");
    const int64_t need = (int64_t)n_windows * txt_len * heads * SVR2_HEAD_DIM * 2;
    SVR2_CHECK_ARG(need <= (int64_t)SVR2_NOALLOC_SCRATCH_BYTES,
                   "svr2_window_attn_fwd_noalloc: %d windows x %d text rows x %d heads need %lld B of scratch, the static array "
                   "holds %lld B; use svr2_window_attn_fwd with a caller workspace", n_windows, txt_len, heads, (long long)need,
                   (long long)SVR2_NOALLOC_SCRATCH_BYTES);
    void* scratch = nullptr;
    const int rc = svr2::noalloc_scratch(&scratch);
    if (rc != SVR2_OK) return rc;
    return svr2_window_attn_fwd_impl(q_pack, k_pack, v_pack, win_table, work, n_work, nullptr, 0, n_windows, T, H, W, txt_len, heads,
                                     P, out_vid, scratch, txt_len ? out_txt : nullptr, SVR2_ATTN_AUTO, stream);
}
