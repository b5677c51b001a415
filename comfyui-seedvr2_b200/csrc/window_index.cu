// svr2_window_index — NaSwin window partition indices, bit-exact integers.
// Replaces na.window_idx / na.window (reference src/models/dit_3b/na.py:583-641): the reference slices an arange tensor
// per window on the host and argsorts it; here every packed position computes its own global token id and writes the
// inverse permutation in the same pass.  HBM-bound integer writes: 16 B per token (tgt + src), O(L).
#include "common.cuh"

namespace svr2 {

// grid.y = window, grid.x CTAs stride over the window's positions.
__global__ void __launch_bounds__(256) window_index_kernel(const int32_t* __restrict__ tab, int n_windows, int H, int W,
                                                           int txt_len, int64_t* __restrict__ tgt,
                                                           int64_t* __restrict__ src, int64_t* __restrict__ wshape,
                                                           int32_t* __restrict__ cu) {
    const int wi = blockIdx.y;
    const Win w = load_win(tab, wi);
    // packed-vid offset of this window = row_start minus the text rows of the windows before it
    const int64_t vid_off = (int64_t)w.row_start - (int64_t)wi * txt_len;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        wshape[3 * wi + 0] = w.nt;
        wshape[3 * wi + 1] = w.nh;
        wshape[3 * wi + 2] = w.nw;
        cu[wi + 1] = w.row_start + w.n_vid + txt_len;
        if (wi == 0) cu[0] = 0;
    }
    for (int j = blockIdx.x * 256 + threadIdx.x; j < w.n_vid; j += gridDim.x * 256) {
        int lt, lh, lw;
        const int64_t tok = win_token(w, j, H, W, &lt, &lh, &lw);
        tgt[vid_off + j] = tok;          // coalesced
        src[tok] = vid_off + j;          // runs of nw consecutive tokens -> coalesced per window row
    }
}

}  // namespace svr2

extern "C" int svr2_window_index(const int32_t* win_table, int n_windows, int T, int H, int W, int txt_len,
                                 int64_t* tgt_idx, int64_t* src_idx, int64_t* window_shape, int32_t* cu_seqlens,
                                 void* stream) {
    SVR2_CHECK_ARG(win_table && tgt_idx && src_idx && window_shape && cu_seqlens, "svr2_window_index: null pointer");
    SVR2_CHECK_ARG(n_windows > 0 && n_windows <= 65535 && T > 0 && H > 0 && W > 0 && txt_len >= 0,
                   "svr2_window_index: bad sizes (n_windows=%d)", n_windows);
    // a window never exceeds ceil(30/4)*ceil(45/3)*ceil(80/3) = 8*15*27 = 3240 tokens (window.py:32-36); 4 CTAs of 256
    // threads stride over it, and the loop keeps any size correct.
    dim3 grid(4, n_windows);
    svr2::window_index_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(win_table, n_windows, H, W, txt_len, tgt_idx,
                                                                       src_idx, window_shape, cu_seqlens);
    SVR2_CHECK_LAUNCH("svr2_window_index");
    return SVR2_OK;
}
