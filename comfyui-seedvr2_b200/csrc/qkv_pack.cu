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
// HBM-bound: reads 3*heads*256 B per packed row, writes the same.  One CTA per packed row; a half-warp owns up to five
// (q|k|v, head) slices of 128 values as 16-byte vectors (every access a fully coalesced 256-B segment) and issues all of
// its loads before any arithmetic; products that the reference rounds to bf16 use HMUL2.BF16 (exact), the division is
// one reciprocal + a Newton step.  Where the row comes from is read from the plan's packed-row map (svr2_pack_rowmap).
// Measured (profiles/r01_run9_*): the kernel is latency bound, not issue bound — one row per CTA at 48 registers
// (5 CTAs = 40 warps per SM) runs at 3.6 TB/s, two rows per CTA at 74 registers (3 CTAs) at 3.1 TB/s, and a version with
// 30 % fewer instructions but 96-120 registers (2 CTAs) at 2.7 TB/s.  A persistent variant that stages whole rows through
// shared memory with 1-D bulk copies (3 stages x 4 rows, 512 threads) was correct but slower (149.5 us against 136.2 us at
// C1): with the loads off the critical path the kernel is issue bound — ~5,200 warp instructions per row, 63 % issue-slot
// use with 4 warps per scheduler (profiles/r01_qkv_pack_staged_experiment.txt) — so the exact-rounding arithmetic, not the
// memory system, sets the floor (~85 us of issue at IPC 1 against 84 us of HBM time).
#include "common.cuh"

namespace svr2 {

__device__ __forceinline__ float torch_linspace_m1_p1(int idx, int n) {
    // torch.linspace(-1, 1, steps=n)[idx] as the CUDA kernel evaluates it (RangeFactories.cu): symmetric halves.
    if (n <= 1) return -1.0f;
    const float step = __fdiv_rn(2.0f, (float)(n - 1));
    return (idx < n / 2) ? __fadd_rn(-1.0f, __fmul_rn(step, (float)idx))
                         : __fsub_rn(1.0f, __fmul_rn(step, (float)(n - idx - 1)));
}

// cos / sin of rotary pair p of one packed row; pairs at or beyond n_rot rotate by (1, 0) = identity
__device__ __forceinline__ void rope_pair(int rope_kind, int p, int n_rot, int n_freqs, const float* __restrict__ freqs,
                                          bool is_txt, int txt_row, int txt_len, int lt, int lh, int lw, int wnt, int wnh,
                                          int wnw, float* c_out, float* s_out) {
    float c = 1.f, sn = 0.f;
    if (p < n_rot) {
        const int axis = p / n_freqs, fi = p - axis * n_freqs;
        const float f = __ldg(freqs + fi);
        if (rope_kind == SVR2_ROPE_3B_MM) {
            // video: positions (Lt + t, h, w) inside the window (rope.py:172); text row r: (r, r, r) (rope.py:173)
            const int pos = is_txt ? txt_row : sel_i32(axis == 0, txt_len + lt, sel_i32(axis == 1, lh, lw));
            const float ang = __fmul_rn((float)pos, f);
            c = cosf(ang);
            sn = sinf(ang);
        } else {
            // 7B: linspace(-1,1,n) over the window's own extent; the angle AND cos/sin are bf16 tensors
            // (freqs.to(q.dtype), dit_7b/rope.py:85; freqs.cos() of a bf16 tensor is bf16)
            const int n = sel_i32(axis == 0, wnt, sel_i32(axis == 1, wnh, wnw));
            const int idx = sel_i32(axis == 0, lt, sel_i32(axis == 1, lh, lw));
            const float ang = bf16r(__fmul_rn(torch_linspace_m1_p1(idx, n), f));
            c = bf16r(cosf(ang));
            sn = bf16r(sinf(ang));
        }
    }
    *c_out = c;
    *s_out = sn;
}

// One (q|k, head) slice of 128 values held by a half-warp, 8 per thread: RMSNorm over the slice with the reference's bf16
// rounding points (normalization.py:88-109), affine weight, interleaved-pair rotation (apply_rotary_emb).
__device__ __forceinline__ bf16x8 norm_rope_slice(const bf16x8& raw, bool is_q, const uint4& wq, const uint4& wk,
                                                  const float* cs, const float* sn, float eps) {
    float x[8];
    unpack8(raw, x);
    // x.pow(2) is a bf16 tensor; its fp32 sum over the half-warp's 128 values: one HMMA per quad + two butterfly steps
    // (both half-warps of a warp are always on this path together: q|k slices end at an even item index)
    float ss = quad_sum_bf16x8(mul8(raw, raw));
    ss += __shfl_xor_sync(0xffffffffu, ss, 4);
    ss += __shfl_xor_sync(0xffffffffu, ss, 8);
    const float var = bf16r(ss * (1.0f / SVR2_HEAD_DIM));
    const float rms = bf16r(sqrt_approx(bf16r(var + eps)));
    const float rcp = rcp_approx(rms);
#pragma unroll
    for (int i = 0; i < 8; ++i) x[i] = x[i] * rcp;       // rounds to the same bf16 as x / rms (common.cuh)
    // field-wise selects (selp): a struct-valued ?: would go through local memory
    bf16x8 wgt;
    *reinterpret_cast<uint4*>(&wgt) = make_uint4(sel_u32(is_q, wq.x, wk.x), sel_u32(is_q, wq.y, wk.y),
                                                 sel_u32(is_q, wq.z, wk.z), sel_u32(is_q, wq.w, wk.w));
    unpack8(mul8(pack8(x), wgt), x);                    // (x / rms) -> bf16, * weight -> bf16
    // interleaved-pair rotation in fp32, mul and add rounded separately like the two ATen ops
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float x0 = x[2 * i], x1 = x[2 * i + 1];
        x[2 * i] = __fadd_rn(__fmul_rn(x0, cs[i]), __fmul_rn(-x1, sn[i]));
        x[2 * i + 1] = __fadd_rn(__fmul_rn(x1, cs[i]), __fmul_rn(x0, sn[i]));
    }
    return pack8(x);
}

constexpr int kRowsPerCta = 1;     // packed rows per CTA (the code below is written for 1 or 2)
constexpr int kMaxItems = 5;       // (q|k|v, head) slices per half-warp: 16 half-warps x 5 >= 3 * 24 heads
constexpr int kPackThreads = 256;

// Per-row bookkeeping every thread derives redundantly (uniform loads of the tiny L1-resident window table).
struct PackRow {
    const __nv_bfloat16* src;   // qkv row of the source token (video) or text token
    int valid, is_txt, n_rot;
};

#ifndef SVR2_PACK_MINB
#define SVR2_PACK_MINB 5      // resident CTAs per SM the register budget is sized for
#endif
__global__ void __launch_bounds__(kPackThreads, SVR2_PACK_MINB)
qkv_pack_kernel(const __nv_bfloat16* __restrict__ qkv_vid, const __nv_bfloat16* __restrict__ qkv_txt,
                const __nv_bfloat16* __restrict__ wq_vid, const __nv_bfloat16* __restrict__ wk_vid,
                const __nv_bfloat16* __restrict__ wq_txt, const __nv_bfloat16* __restrict__ wk_txt,
                const float* __restrict__ freqs, int n_freqs, int rope_kind, const int32_t* __restrict__ tab,
                const int4* __restrict__ row_map, int n_windows, int H, int W, int txt_len, int heads, float eps,
                __nv_bfloat16* __restrict__ qp,
                __nv_bfloat16* __restrict__ kp, __nv_bfloat16* __restrict__ vp, int64_t P) {
    __shared__ float s_cos[kRowsPerCta][64], s_sin[kRowsPerCta][64];
    const int hl = threadIdx.x & 15;              // lane inside the half-warp: owns elements [8*hl, 8*hl+8)
    const int hw_id = threadIdx.x >> 4;           // 16 half-warps per CTA
    const int col = hl * 8;
    const int n_items = 3 * heads;                // item = kind * heads + head; source offset = item * 128
    const int64_t qkv_stride = (int64_t)n_items * SVR2_HEAD_DIM;

    PackRow pr[kRowsPerCta];
    bf16x8 raw[kRowsPerCta][kMaxItems];
#pragma unroll
    for (int r = 0; r < kRowsPerCta; ++r) {
        const int64_t row = (int64_t)blockIdx.x * kRowsPerCta + r;
        pr[r].valid = row < P;
        pr[r].is_txt = 0; pr[r].n_rot = 0; pr[r].src = qkv_vid;
        if (!pr[r].valid) continue;
        bool is_txt;
        int lt = 0, lh = 0, lw = 0, txt_row = 0, wnt = 1, wnh = 1, wnw = 1;
        if (row_map != nullptr) {
            // the plan's row map (svr2_pack_rowmap): one 16-B load instead of a search of the window table
            const int4 m = __ldg(row_map + row);
            is_txt = m.x < 0;
            txt_row = -(m.x + 1);
            lt = m.y & 0xffff; lh = (m.y >> 16) & 0xffff; lw = m.z;
            pr[r].src = is_txt ? qkv_txt + (int64_t)txt_row * qkv_stride : qkv_vid + (int64_t)m.x * qkv_stride;
            if (rope_kind == SVR2_ROPE_7B_PIX && !is_txt) {
                const int32_t* t = tab + m.w * SVR2_WIN_FIELDS;
                wnt = __ldg(t + 3); wnh = __ldg(t + 4); wnw = __ldg(t + 5);
            }
        } else {
            // window of this packed row: last window whose row_start <= row
            int lo = 0, hi = n_windows - 1;
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (__ldg(tab + mid * SVR2_WIN_FIELDS + 6) <= row) lo = mid; else hi = mid - 1;
            }
            const Win w = load_win(tab, lo);
            const int j = (int)row - w.row_start;
            is_txt = j >= w.n_vid;
            txt_row = j - w.n_vid;
            wnt = w.nt; wnh = w.nh; wnw = w.nw;
            if (is_txt) {
                pr[r].src = qkv_txt + (int64_t)txt_row * qkv_stride;
            } else {
                pr[r].src = qkv_vid + win_token(w, j, H, W, &lt, &lh, &lw) * qkv_stride;
            }
        }
        pr[r].is_txt = is_txt;
        pr[r].n_rot = (rope_kind == SVR2_ROPE_3B_MM || (rope_kind == SVR2_ROPE_7B_PIX && !is_txt)) ? 3 * n_freqs : 0;
        // all of this thread's loads for the row go out before any arithmetic
#pragma unroll
        for (int it = 0; it < kMaxItems; ++it) {
            const int item = hw_id + it * 16;
            if (item < n_items) raw[r][it] = ldg8_stream(pr[r].src + item * SVR2_HEAD_DIM + col);
        }
        // cos/sin of the row's rotary pairs: threads [64r, 64r+64) — overlaps the loads above
        const int p = (int)threadIdx.x - 64 * r;
        if (p >= 0 && p < 64) {
            float c, sn;
            rope_pair(rope_kind, p, pr[r].n_rot, n_freqs, freqs, is_txt, txt_row, txt_len, lt, lh, lw, wnt, wnh, wnw, &c, &sn);
            s_cos[r][p] = c;        // pairs beyond n_rot rotate by (1, 0): identity, branch-free below
            s_sin[r][p] = sn;
        }
    }
    __syncthreads();

#pragma unroll
    for (int r = 0; r < kRowsPerCta; ++r) {
        if (!pr[r].valid) continue;              // uniform across the CTA
        const int64_t drow = ((int64_t)blockIdx.x * kRowsPerCta + r) * heads * SVR2_HEAD_DIM;
        // the row's norm weights (video or text set): two L1-resident 16-B loads, selected by pointer
        const uint4 wq = __ldg(reinterpret_cast<const uint4*>((pr[r].is_txt ? wq_txt : wq_vid) + col));
        const uint4 wk = __ldg(reinterpret_cast<const uint4*>((pr[r].is_txt ? wk_txt : wk_vid) + col));
        float cs[4], sn[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { cs[i] = s_cos[r][hl * 4 + i]; sn[i] = s_sin[r][hl * 4 + i]; }
#pragma unroll
        for (int it = 0; it < kMaxItems; ++it) {
            const int item = hw_id + it * 16;
            // q | k | v without an integer division; kind 3 = no slice for this half-warp
            const int kind = item >= n_items ? 3 : (item >= heads) + (item >= 2 * heads);
            // the q|k -> v boundary (item 2 * heads) is even and the two half-warps of a warp hold items (2m, 2m+1): a warp is
            // never split between the norm path and the others, so the row-sum HMMA below always sees a converged warp
            if (kind == 3) continue;             // uniform across the half-warp
            const int head = item - kind * heads;
            __nv_bfloat16* dst = (kind == 0 ? qp : (kind == 1 ? kp : vp)) + drow + head * SVR2_HEAD_DIM + col;
            if (kind == 2) {                     // v: gather only
                stg8(dst, raw[r][it]);
                continue;
            }
            stg8(dst, norm_rope_slice(raw[r][it], kind == 0, wq, wk, cs, sn, eps));
        }
    }
}


// one thread per packed row: the search every qkv_pack CTA used to repeat, done once per plan
__global__ void pack_rowmap_kernel(const int32_t* __restrict__ tab, int n_windows, int H, int W, int4* __restrict__ row_map,
                                   int64_t P) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= P) return;
    int lo = 0, hi = n_windows - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(tab + mid * SVR2_WIN_FIELDS + 6) <= row) lo = mid; else hi = mid - 1;
    }
    const Win w = load_win(tab, lo);
    const int j = (int)row - w.row_start;
    int4 m = make_int4(0, 0, 0, lo);
    if (j >= w.n_vid) {
        m.x = -(1 + (j - w.n_vid));
    } else {
        int lt, lh, lw;
        m.x = (int)win_token(w, j, H, W, &lt, &lh, &lw);
        m.y = lt | (lh << 16);
        m.z = lw;
    }
    row_map[row] = m;
}

// ---- per-plan tables of the fused projection (svr2_linear_qkv_pack_fwd, gemm.cu) ---------------------------------------------
// Rope table: one row of n_freqs (cos, sin) pairs per distinct POSITION VALUE of the plan, computed by the very code path of
// rope_pair() above so both qkv paths rotate by identical bits.  3B: row r = position r (video: Lt + t, h, w; text row j: j for
// all three axes).  7B: row base_of_n[n] + idx = linspace(-1, 1, n)[idx] for every window extent n that occurs in the plan.
__global__ void rope_table_kernel(int rope_kind, const float* __restrict__ freqs, int n_freqs, int rows,
                                  const int32_t* __restrict__ base_of_n, float2* __restrict__ out) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= rows * n_freqs) return;
    const int r = e / n_freqs, fi = e - r * n_freqs;
    const float f = __ldg(freqs + fi);
    float c, sn;
    if (rope_kind == SVR2_ROPE_3B_MM) {
        const float ang = __fmul_rn((float)r, f);
        c = cosf(ang);
        sn = sinf(ang);
    } else {
        int n = 1, idx = 0;
        for (int k = 1; k <= 64; ++k) {
            const int b = __ldg(base_of_n + k);
            if (b >= 0 && r >= b && r < b + k) { n = k; idx = r - b; }
        }
        const float ang = bf16r(__fmul_rn(torch_linspace_m1_p1(idx, n), f));
        c = bf16r(cosf(ang));
        sn = bf16r(sinf(ang));
    }
    out[e] = make_float2(c, sn);
}

// Token map: the inverse of the packed-row map.  tok_map[token] = (packed row, table row of axis t, h, w); the text rows follow
// the L video tokens: tok_map[L + j] = (-(1 + j), j, j, j) (3B) or (-(1 + j), -1, -1, -1) (7B / no rope: text is not rotated).
__global__ void tok_map_kernel(const int32_t* __restrict__ tab, const int4* __restrict__ row_map, int64_t P, int64_t L,
                               int txt_len, int rope_kind, const int32_t* __restrict__ base_of_n, int4* __restrict__ tok_map) {
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= P) return;
    const int4 m = __ldg(row_map + row);
    if (m.x < 0) {
        if (m.w != 0) return;           // every window repeats the text rows; window 0 writes their entries
        const int j = -(m.x + 1);
        const int r = rope_kind == SVR2_ROPE_3B_MM ? j : -1;
        tok_map[L + j] = make_int4(-(1 + j), r, r, r);
        return;
    }
    const int lt = m.y & 0xffff, lh = (m.y >> 16) & 0xffff, lw = m.z;
    int4 o = make_int4((int)row, -1, -1, -1);
    if (rope_kind == SVR2_ROPE_3B_MM) {
        o.y = txt_len + lt; o.z = lh; o.w = lw;
    } else if (rope_kind == SVR2_ROPE_7B_PIX) {
        const int32_t* t = tab + m.w * SVR2_WIN_FIELDS;
        o.y = __ldg(base_of_n + __ldg(t + 3)) + lt;
        o.z = __ldg(base_of_n + __ldg(t + 4)) + lh;
        o.w = __ldg(base_of_n + __ldg(t + 5)) + lw;
    }
    tok_map[m.x] = o;
}

// Text rows of q, k, v (normalised + rotated once by the projection's epilogue) copied into every window's sequence
// (mmattn.py:223-236 repeats them per window; concat_win puts them after the window's video rows).
__global__ void txt_replicate_kernel(const uint4* __restrict__ txt_qkv, const int32_t* __restrict__ tab, int txt_len, int heads,
                                     uint4* __restrict__ qp, uint4* __restrict__ kp, uint4* __restrict__ vp) {
    const int w = blockIdx.x / txt_len, j = blockIdx.x - w * txt_len;
    const int32_t* t = tab + w * SVR2_WIN_FIELDS;
    const int64_t drow = (int64_t)__ldg(t + 6) + __ldg(t + 7) + j;       // row_start + n_vid + j
    const int per_kind = heads * (SVR2_HEAD_DIM * 2 / 16);                  // uint4 per (row, kind)
    for (int i = threadIdx.x; i < 3 * per_kind; i += blockDim.x) {
        const int kind = i / per_kind, c = i - kind * per_kind;
        uint4* dst = kind == 0 ? qp : (kind == 1 ? kp : vp);
        dst[drow * per_kind + c] = __ldg(txt_qkv + ((int64_t)j * 3 + kind) * per_kind + c);
    }
}

}  // namespace svr2

extern "C" int svr2_rope_plan(const int32_t* win_table, const int32_t* row_map, int n_windows, int T, int H, int W, int txt_len,
                              int64_t P, int rope_kind, const float* freqs, int n_freqs, const int32_t* base_of_n, int rope_rows,
                              int32_t* tok_map, float* rope_cs, void* stream) {
    SVR2_CHECK_ARG(win_table && row_map && tok_map && n_windows > 0 && P > 0 && P < ((int64_t)1 << 31), "svr2_rope_plan: bad arguments");
    SVR2_CHECK_ARG(rope_kind == SVR2_ROPE_NONE || rope_kind == SVR2_ROPE_3B_MM || rope_kind == SVR2_ROPE_7B_PIX,
                   "svr2_rope_plan: unknown rope_kind %d", rope_kind);
    SVR2_CHECK_ARG(rope_kind == SVR2_ROPE_NONE || (freqs && rope_cs && n_freqs > 0 && rope_rows > 0), "svr2_rope_plan: rope inputs missing");
    SVR2_CHECK_ARG(rope_kind != SVR2_ROPE_7B_PIX || base_of_n, "svr2_rope_plan: the 7B table needs base_of_n[65]");
    const int64_t L = (int64_t)T * H * W;
    cudaStream_t st = (cudaStream_t)stream;
    if (rope_kind != SVR2_ROPE_NONE) {
        const int n = rope_rows * n_freqs;
        svr2::rope_table_kernel<<<(n + 255) / 256, 256, 0, st>>>(rope_kind, freqs, n_freqs, rope_rows, base_of_n, (float2*)rope_cs);
        SVR2_CHECK_LAUNCH("svr2_rope_plan(table)");
    }
    svr2::tok_map_kernel<<<(unsigned)((P + 255) / 256), 256, 0, st>>>(win_table, (const int4*)row_map, P, L, txt_len, rope_kind,
                                                                      base_of_n, (int4*)tok_map);
    SVR2_CHECK_LAUNCH("svr2_rope_plan(token map)");
    return SVR2_OK;
}

extern "C" int svr2_txt_replicate_fwd(const void* txt_qkv, const int32_t* win_table, int n_windows, int txt_len, int heads,
                                      void* q_pack, void* k_pack, void* v_pack, void* stream) {
    SVR2_CHECK_ARG(win_table && q_pack && k_pack && v_pack && n_windows > 0 && heads > 0 && txt_len >= 0, "svr2_txt_replicate_fwd: bad arguments");
    if (txt_len == 0) return SVR2_OK;
    SVR2_CHECK_ARG(txt_qkv, "svr2_txt_replicate_fwd: null pointer");
    svr2::txt_replicate_kernel<<<(unsigned)(n_windows * txt_len), 256, 0, (cudaStream_t)stream>>>(
        (const uint4*)txt_qkv, win_table, txt_len, heads, (uint4*)q_pack, (uint4*)k_pack, (uint4*)v_pack);
    SVR2_CHECK_LAUNCH("svr2_txt_replicate_fwd");
    return SVR2_OK;
}

extern "C" int svr2_pack_rowmap(const int32_t* win_table, int n_windows, int T, int H, int W, int txt_len, int32_t* row_map,
                                int64_t P, void* stream) {
    SVR2_CHECK_ARG(win_table && row_map && n_windows > 0 && P > 0 && P < ((int64_t)1 << 31), "svr2_pack_rowmap: bad arguments");
    SVR2_CHECK_ARG(T > 0 && H > 0 && W > 0 && T < 65536 && H < 65536 && (int64_t)T * H * W < ((int64_t)1 << 31) && txt_len >= 0,
                   "svr2_pack_rowmap: grid %dx%dx%d out of range", T, H, W);
    svr2::pack_rowmap_kernel<<<(unsigned)((P + 255) / 256), 256, 0, (cudaStream_t)stream>>>(win_table, n_windows, H, W,
                                                                                             (int4*)row_map, P);
    SVR2_CHECK_LAUNCH("svr2_pack_rowmap");
    return SVR2_OK;
}

extern "C" int svr2_qkv_pack_fwd(const void* qkv_vid, const void* qkv_txt, const void* wq_vid, const void* wk_vid,
                                 const void* wq_txt, const void* wk_txt, const float* freqs, int n_freqs, int rope_kind,
                                 const int32_t* win_table, const int32_t* row_map, int n_windows, int T, int H, int W,
                                 int txt_len, int heads, float eps, void* q_pack, void* k_pack, void* v_pack, int64_t P,
                                 void* stream) {
    SVR2_CHECK_ARG(qkv_vid && wq_vid && wk_vid && win_table && q_pack && k_pack && v_pack, "svr2_qkv_pack_fwd: null pointer");
    SVR2_CHECK_ARG(txt_len == 0 || (qkv_txt && wq_txt && wk_txt), "svr2_qkv_pack_fwd: text inputs missing");
    SVR2_CHECK_ARG(heads > 0 && n_windows > 0 && P > 0 && P < ((int64_t)1 << 31), "svr2_qkv_pack_fwd: bad sizes");
    SVR2_CHECK_ARG(rope_kind == SVR2_ROPE_NONE || rope_kind == SVR2_ROPE_3B_MM || rope_kind == SVR2_ROPE_7B_PIX,
                   "svr2_qkv_pack_fwd: unknown rope_kind %d", rope_kind);
    SVR2_CHECK_ARG(rope_kind == SVR2_ROPE_NONE || (freqs && n_freqs > 0 && 3 * n_freqs <= 64),
                   "svr2_qkv_pack_fwd: need 0 < 3*n_freqs <= 64 rotary pairs (got n_freqs=%d)", n_freqs);
    (void)T;
    SVR2_CHECK_ARG(3 * heads <= 16 * svr2::kMaxItems, "svr2_qkv_pack_fwd: at most %d heads", 16 * svr2::kMaxItems / 3);
    const unsigned grid = (unsigned)((P + svr2::kRowsPerCta - 1) / svr2::kRowsPerCta);
    svr2::qkv_pack_kernel<<<grid, svr2::kPackThreads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)qkv_vid, (const __nv_bfloat16*)qkv_txt, (const __nv_bfloat16*)wq_vid,
        (const __nv_bfloat16*)wk_vid, (const __nv_bfloat16*)wq_txt, (const __nv_bfloat16*)wk_txt, freqs, n_freqs,
        rope_kind, win_table, (const int4*)row_map, n_windows, H, W, txt_len, heads, eps, (__nv_bfloat16*)q_pack, (__nv_bfloat16*)k_pack,
        (__nv_bfloat16*)v_pack, P);
    SVR2_CHECK_LAUNCH("svr2_qkv_pack_fwd");
    return SVR2_OK;
}
