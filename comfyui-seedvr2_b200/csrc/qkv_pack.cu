// svr2_qkv_pack_fwd — q/k RMSNorm (affine, per head) + 3D RoPE + window partition + text interleave in ONE pass.
//
// Replaces, per attention call (reference src/models/dit_3b/nablocks/attention/mmattn.py:199-248 and the 7B twin
// dit_7b/nablocks/mmsr_block.py:106-128):
//   window_partition index_select (na.py:637)              -> the gather is the read address of this kernel
//   norm_q / norm_k CustomRMSNorm (normalization.py:88-109) -> bf16 rounding points reproduced with bf16r()
//   text q/k repeat per window (mmattn.py:223-236)          -> text rows are written once per window
//   RoPE (dit_3b/rope.py:92-127, dit_7b/rope.py:73-91, third-party apply_rotary_emb) -> fp32, no FMA contraction
//   concat_win (na.py:320-424)                              -> output is already the packed varlen layout
// Output: q,k,v [P, heads, 128] bf16, sequence k = [video rows of window k (row-major t,h,w) ; Lt text rows].
//
// HBM-bound: reads 3*heads*256 B per packed row, writes the same.  One CTA per packed row; a half-warp owns one
// (q|k|v, head) slice of 128 values as 16-byte vectors, so every access is a fully coalesced 256-B segment.
#include "common.cuh"

namespace svr2 {

__device__ __forceinline__ float torch_linspace_m1_p1(int idx, int n) {
    // torch.linspace(-1, 1, steps=n)[idx] as the CUDA kernel evaluates it (RangeFactories.cu): symmetric halves.
    if (n <= 1) return -1.0f;
    const float step = __fdiv_rn(2.0f, (float)(n - 1));
    return (idx < n / 2) ? __fadd_rn(-1.0f, __fmul_rn(step, (float)idx))
                         : __fsub_rn(1.0f, __fmul_rn(step, (float)(n - idx - 1)));
}

__global__ void __launch_bounds__(128)
qkv_pack_kernel(const __nv_bfloat16* __restrict__ qkv_vid, const __nv_bfloat16* __restrict__ qkv_txt,
                const __nv_bfloat16* __restrict__ wq_vid, const __nv_bfloat16* __restrict__ wk_vid,
                const __nv_bfloat16* __restrict__ wq_txt, const __nv_bfloat16* __restrict__ wk_txt,
                const float* __restrict__ freqs, int n_freqs, int rope_kind, const int32_t* __restrict__ tab,
                int n_windows, int H, int W, int txt_len, int heads, float eps, __nv_bfloat16* __restrict__ qp,
                __nv_bfloat16* __restrict__ kp, __nv_bfloat16* __restrict__ vp) {
    __shared__ float s_cos[64], s_sin[64];
    const int row = blockIdx.x;
    // window of this packed row: last window whose row_start <= row (uniform across the CTA, table is L1/L2 resident)
    int lo = 0, hi = n_windows - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(tab + mid * SVR2_WIN_FIELDS + 6) <= row) lo = mid; else hi = mid - 1;
    }
    const Win w = load_win(tab, lo);
    const int j = row - w.row_start;
    const bool is_txt = j >= w.n_vid;
    int lt = 0, lh = 0, lw = 0;
    const __nv_bfloat16* src;
    if (is_txt) {
        src = qkv_txt + (int64_t)(j - w.n_vid) * 3 * heads * SVR2_HEAD_DIM;
    } else {
        const int64_t tok = win_token(w, j, H, W, &lt, &lh, &lw);
        src = qkv_vid + tok * 3 * heads * SVR2_HEAD_DIM;
    }
    const int n_rot_pairs = (rope_kind == SVR2_ROPE_3B_MM || (rope_kind == SVR2_ROPE_7B_PIX && !is_txt)) ? 3 * n_freqs : 0;
    if (threadIdx.x < 64) {
        const int p = threadIdx.x;
        float c = 1.f, s = 0.f;
        if (p < n_rot_pairs) {
            const int axis = p / n_freqs, fi = p - axis * n_freqs;
            const float f = __ldg(freqs + fi);
            if (rope_kind == SVR2_ROPE_3B_MM) {
                // video: positions (Lt + t, h, w) inside the window (rope.py:172); text row r: (r, r, r) (rope.py:173)
                const int pos = is_txt ? (j - w.n_vid) : (axis == 0 ? txt_len + lt : (axis == 1 ? lh : lw));
                const float ang = __fmul_rn((float)pos, f);
                c = cosf(ang);
                s = sinf(ang);
            } else {
                // 7B: linspace(-1,1,n) over the window's own extent; the angle AND cos/sin are bf16 tensors
                // (freqs.to(q.dtype), dit_7b/rope.py:85; freqs.cos() of a bf16 tensor is bf16)
                const int n = axis == 0 ? w.nt : (axis == 1 ? w.nh : w.nw);
                const int idx = axis == 0 ? lt : (axis == 1 ? lh : lw);
                const float ang = bf16r(__fmul_rn(torch_linspace_m1_p1(idx, n), f));
                c = bf16r(cosf(ang));
                s = bf16r(sinf(ang));
            }
        }
        s_cos[p] = c;
        s_sin[p] = s;
    }
    __syncthreads();

    const int hl = threadIdx.x & 15;              // lane inside the half-warp: owns elements [8*hl, 8*hl+8)
    const int hw_id = threadIdx.x >> 4;           // 8 half-warps per CTA
    const int col = hl * 8;
    const unsigned half_mask = 0xFFFFu << (threadIdx.x & 16);
    float wq[8], wk[8];
    unpack8(ldg8((is_txt ? wq_txt : wq_vid) + col), wq);
    unpack8(ldg8((is_txt ? wk_txt : wk_vid) + col), wk);
    const int64_t drow = (int64_t)row * heads * SVR2_HEAD_DIM;
    const int n_items = 3 * heads;                // item = kind * heads + head; source offset = item * 128
    for (int it = hw_id; it < n_items; it += 8) {
        const int kind = it / heads, head = it - kind * heads;
        const bf16x8 raw = ldg8_stream(src + it * SVR2_HEAD_DIM + col);
        __nv_bfloat16* dst = (kind == 0 ? qp : (kind == 1 ? kp : vp)) + drow + head * SVR2_HEAD_DIM + col;
        if (kind == 2) {                          // v: gather only
            stg8(dst, raw);
            continue;
        }
        float x[8];
        unpack8(raw, x);
        float ss = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += bf16r(x[i] * x[i]);
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) ss += __shfl_xor_sync(half_mask, ss, o);     // the two halves may diverge
        const float var = bf16r(ss * (1.0f / SVR2_HEAD_DIM));
        const float rms = bf16r(sqrtf(bf16r(var + eps)));
        const float* wgt = kind == 0 ? wq : wk;
#pragma unroll
        for (int i = 0; i < 8; ++i) x[i] = bf16r(bf16r(__fdiv_rn(x[i], rms)) * wgt[i]);
        // interleaved-pair rotation in fp32, mul and add rounded separately like the two ATen ops
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int p = hl * 4 + i;
            if (p < n_rot_pairs) {
                const float c = s_cos[p], s = s_sin[p];
                const float x0 = x[2 * i], x1 = x[2 * i + 1];
                x[2 * i] = __fadd_rn(__fmul_rn(x0, c), __fmul_rn(-x1, s));
                x[2 * i + 1] = __fadd_rn(__fmul_rn(x1, c), __fmul_rn(x0, s));
            }
        }
        stg8(dst, pack8(x));
    }
}

}  // namespace svr2

extern "C" int svr2_qkv_pack_fwd(const void* qkv_vid, const void* qkv_txt, const void* wq_vid, const void* wk_vid,
                                 const void* wq_txt, const void* wk_txt, const float* freqs, int n_freqs, int rope_kind,
                                 const int32_t* win_table, int n_windows, int T, int H, int W, int txt_len, int heads,
                                 float eps, void* q_pack, void* k_pack, void* v_pack, int64_t P, void* stream) {
    SVR2_CHECK_ARG(qkv_vid && wq_vid && wk_vid && win_table && q_pack && k_pack && v_pack, "svr2_qkv_pack_fwd: null pointer");
    SVR2_CHECK_ARG(txt_len == 0 || (qkv_txt && wq_txt && wk_txt), "svr2_qkv_pack_fwd: text inputs missing");
    SVR2_CHECK_ARG(heads > 0 && n_windows > 0 && P > 0 && P < ((int64_t)1 << 31), "svr2_qkv_pack_fwd: bad sizes");
    SVR2_CHECK_ARG(rope_kind == SVR2_ROPE_NONE || rope_kind == SVR2_ROPE_3B_MM || rope_kind == SVR2_ROPE_7B_PIX,
                   "svr2_qkv_pack_fwd: unknown rope_kind %d", rope_kind);
    SVR2_CHECK_ARG(rope_kind == SVR2_ROPE_NONE || (freqs && n_freqs > 0 && 3 * n_freqs <= 64),
                   "svr2_qkv_pack_fwd: need 0 < 3*n_freqs <= 64 rotary pairs (got n_freqs=%d)", n_freqs);
    (void)T;
    svr2::qkv_pack_kernel<<<(unsigned)P, 128, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)qkv_vid, (const __nv_bfloat16*)qkv_txt, (const __nv_bfloat16*)wq_vid,
        (const __nv_bfloat16*)wk_vid, (const __nv_bfloat16*)wq_txt, (const __nv_bfloat16*)wk_txt, freqs, n_freqs,
        rope_kind, win_table, n_windows, H, W, txt_len, heads, eps, (__nv_bfloat16*)q_pack, (__nv_bfloat16*)k_pack,
        (__nv_bfloat16*)v_pack);
    SVR2_CHECK_LAUNCH("svr2_qkv_pack_fwd");
    return SVR2_OK;
}
