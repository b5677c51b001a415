// Error plumbing + host-side window geometry (pure double/integer arithmetic, no device work).
#include <math.h>
#include <stdarg.h>
#include <string.h>

#include <vector>

#include "common.cuh"

namespace svr2 {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
}  // namespace svr2

extern "C" const char* svr2_last_error(void) { return svr2::g_err; }
extern "C" int svr2_abi_version(void) { return 14; }   // 14: svr2_row_softmax_fwd; 13: svr2_causal_conv3d_down_fwd; 12: svr2_pixel_shuffle3d_fwd; 11: svr2_transpose_2b; 10: GroupNorm statistics from the convolution's epilogue; 9: svr2_groupnorm_silu_fwd; 8: svr2_causal_conv3d_fwd; 7: LAB conversions; 6: post-processing (svr2_wavelet_level_fwd, svr2_colorfix_finish_fwd); 5: svr2_linear_qkv_pack_fwd + svr2_rope_plan; 4: svr2_linear_*_fwd; 3: svr2_window_attn_fwd_impl

namespace {

struct Span {
    int lo, hi;
};

// Python's round(): ties to even. nearbyint() under the default FE_TONEAREST mode does exactly that.
inline long py_round(double x) { return (long)nearbyint(x); }

// Regular windows along one axis: [i*w, min((i+1)*w, size)), empty ones dropped.   window.py:37-49
void axis_regular(int size, int win, std::vector<Span>& out) {
    const int n = (int)ceil((double)size / (double)win);
    for (int i = 0; i < n; ++i) {
        const int lo = i * win, hi = ((i + 1) * win < size) ? (i + 1) * win : size;
        if (hi > lo) out.push_back({lo, hi});
    }
}

// Shifted windows along one axis: half-window shift when the window is smaller than the axis.   window.py:60-83
void axis_shifted(int size, int win, std::vector<Span>& out) {
    const double s = (win < size) ? 0.5 : 0.0;
    int n = (int)ceil(((double)size - s) / (double)win);
    n = (s > 0.0) ? n + 1 : 1;
    for (int i = 0; i < n; ++i) {
        // int() in Python truncates toward zero; so does the C cast.
        int lo = (int)(((double)i - s) * (double)win);
        int hi = (int)(((double)i - s + 1.0) * (double)win);
        if (lo < 0) lo = 0;
        if (hi > size) hi = size;
        if (hi > lo) out.push_back({lo, hi});
    }
}

}  // namespace

extern "C" int svr2_window_plan(int T, int H, int W, int nt, int nh, int nw, int shifted, int txt_len,
                                int32_t* win_table, int max_windows, int64_t* out_counts) {
    SVR2_CHECK_ARG(T > 0 && H > 0 && W > 0, "svr2_window_plan: grid must be positive (got %dx%dx%d)", T, H, W);
    SVR2_CHECK_ARG(nt > 0 && nh > 0 && nw > 0, "svr2_window_plan: window triple must be positive");
    SVR2_CHECK_ARG(txt_len >= 0 && out_counts != nullptr, "svr2_window_plan: bad txt_len / out_counts");
    SVR2_CHECK_ARG((int64_t)T * H * W < (int64_t)1 << 30, "svr2_window_plan: grid too large");

    // window.py:32-36 — all in double precision exactly like CPython floats
    const double scale = sqrt((45.0 * 80.0) / ((double)H * (double)W));
    const long rh = py_round((double)H * scale), rw = py_round((double)W * scale);
    const int wh = (int)ceil((double)rh / (double)nh);
    const int ww = (int)ceil((double)rw / (double)nw);
    const int wt = (int)ceil((double)(T < 30 ? T : 30) / (double)nt);
    SVR2_CHECK_ARG(wh > 0 && ww > 0 && wt > 0, "svr2_window_plan: degenerate window size");

    std::vector<Span> ts, hs, ws;
    if (shifted) {
        axis_shifted(T, wt, ts); axis_shifted(H, wh, hs); axis_shifted(W, ww, ws);
    } else {
        axis_regular(T, wt, ts); axis_regular(H, wh, hs); axis_regular(W, ww, ws);
    }
    const int64_t n = (int64_t)ts.size() * hs.size() * ws.size();
    int64_t sum = 0, mx = 0, mn = INT64_MAX, qtiles = 0, row = 0;
    int k = 0;
    const bool fill = (win_table != nullptr) && (n <= max_windows);
    for (const Span& sw : ws)          // iw outer
        for (const Span& sh : hs)      // ih middle
            for (const Span& st : ts)  // it inner
            {
                const int64_t lw = (int64_t)(st.hi - st.lo) * (sh.hi - sh.lo) * (sw.hi - sw.lo);
                if (fill) {
                    int32_t* e = win_table + (int64_t)k * SVR2_WIN_FIELDS;
                    e[0] = st.lo; e[1] = sh.lo; e[2] = sw.lo;
                    e[3] = st.hi - st.lo; e[4] = sh.hi - sh.lo; e[5] = sw.hi - sw.lo;
                    e[6] = (int32_t)row; e[7] = (int32_t)lw;
                }
                sum += lw;
                if (lw > mx) mx = lw;
                if (lw < mn) mn = lw;
                qtiles += (lw + txt_len + 127) / 128;
                row += lw + txt_len;
                ++k;
            }
    out_counts[0] = n; out_counts[1] = sum; out_counts[2] = mx; out_counts[3] = row; out_counts[4] = qtiles;
    out_counts[5] = (n > 0) ? mn : 0;
    if (win_table != nullptr && n > max_windows) {
        svr2::set_error("svr2_window_plan: %lld windows do not fit max_windows=%d", (long long)n, max_windows);
        return SVR2_E_INVALID;
    }
    return SVR2_OK;
}
