// Legacy-tensor-path (mma.sync / HMMA) varlen window attention.  NOT the product path: it is the bring-up comparator for
// the tcgen05 kernel in window_attn.cu (select with SVR2_ATTN_IMPL=mma for A/B debugging only) and shares its epilogue
// contract: window reverse + per-window text rows fused into the store.
//
// CTA = 128 query rows of one (window, head); 8 warps x 16 rows; K/V tiles of 64 rows double-buffered with cp.async.
#include "common.cuh"
#include "attn_launch.cuh"

namespace svr2 {
namespace mma_attn {

constexpr int BM = 128, BN = 64, D = SVR2_HEAD_DIM, NTHREADS = 256;
constexpr int Q_BYTES = BM * D * 2, KV_BYTES = BN * D * 2;
constexpr int SMEM_BYTES = Q_BYTES + 4 * KV_BYTES;   // Q + 2 x (K, V)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
// element offset (in bytes) of (row, 16-byte chunk) in a [rows][128] bf16 tile with an 8-chunk XOR swizzle
__device__ __forceinline__ uint32_t sw(int row, int chunk) { return (uint32_t)(row * 256 + (((chunk & 8) | ((chunk ^ row) & 7)) << 4)); }

__device__ __forceinline__ void cp16(uint32_t dst, const void* src, bool valid) {
    const int n = valid ? 16 : 0;   // src-size 0 -> zero fill
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(n));
}
__device__ __forceinline__ void ldsm4(uint32_t a, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(a));
}
__device__ __forceinline__ void ldsm4t(uint32_t a, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(a));
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

// load `rows` x 128 tile starting at packed row `row0` (rows >= row_end are zero-filled) for one head
__device__ __forceinline__ void load_tile(uint32_t sbase, const __nv_bfloat16* g, int64_t row0, int64_t row_end, int rows,
                                          int64_t row_stride, int tid) {
    for (int i = tid; i < rows * 16; i += NTHREADS) {
        const int r = i >> 4, c = i & 15;
        const bool ok = row0 + r < row_end;
        cp16(sbase + sw(r, c), g + (ok ? (row0 + r) * row_stride + c * 8 : 0), ok);
    }
}

__global__ void __launch_bounds__(NTHREADS, 1)
window_attn_mma_kernel(const __nv_bfloat16* __restrict__ qp, const __nv_bfloat16* __restrict__ kp,
                       const __nv_bfloat16* __restrict__ vp, const int32_t* __restrict__ tab,
                       const int32_t* __restrict__ work, int H, int W, int txt_len, int heads,
                       __nv_bfloat16* __restrict__ out_vid, __nv_bfloat16* __restrict__ txt_part) {
    extern __shared__ __align__(128) unsigned char smem[];
    const uint32_t sQ = smem_u32(smem), sK = sQ + Q_BYTES, sV = sK + 2 * KV_BYTES;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
    const int wi = work[2 * blockIdx.x], qt = 2 * work[2 * blockIdx.x + 1] + blockIdx.z, head = blockIdx.y;
    const Win w = load_win(tab, wi);
    const int seq = w.n_vid + txt_len;
    if (qt * BM >= seq) return;      // work items are PAIRS of 128-row query tiles; the second may not exist
    const int64_t row_stride = (int64_t)heads * D;
    const __nv_bfloat16* qh = qp + head * D;
    const __nv_bfloat16* kh = kp + head * D;
    const __nv_bfloat16* vh = vp + head * D;
    const int64_t s0 = w.row_start, s_end = s0 + seq;
    const int n_kv = (seq + BN - 1) / BN;

    load_tile(sQ, qh, s0 + qt * BM, s_end, BM, row_stride, tid);
    load_tile(sK, kh, s0, s_end, BN, row_stride, tid);
    load_tile(sV, vh, s0, s_end, BN, row_stride, tid);
    asm volatile("cp.async.commit_group;");

    uint32_t qf[8][4];
    float o[16][4];
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
    const float sl2 = 0.08838834764831845f * 1.4426950408889634f;   // 1/sqrt(128) * log2(e)

    for (int kt = 0; kt < n_kv; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < n_kv) {
            load_tile(sK + (buf ^ 1) * KV_BYTES, kh, s0 + (kt + 1) * BN, s_end, BN, row_stride, tid);
            load_tile(sV + (buf ^ 1) * KV_BYTES, vh, s0 + (kt + 1) * BN, s_end, BN, row_stride, tid);
            asm volatile("cp.async.commit_group;");
            asm volatile("cp.async.wait_group 1;");
        } else {
            asm volatile("cp.async.wait_group 0;");
        }
        __syncthreads();
        if (kt == 0) {
#pragma unroll
            for (int ks = 0; ks < 8; ++ks)
                ldsm4(sQ + sw(warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, ks * 2 + (lane >> 4)), qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3]);
        }
        const uint32_t kb = sK + buf * KV_BYTES, vb = sV + buf * KV_BYTES;
        float s[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
#pragma unroll
            for (int np = 0; np < 4; ++np) {   // pairs of 8-row kv tiles
                uint32_t b0, b1, b2, b3;
                ldsm4(kb + sw(np * 16 + (lane & 7) + (lane >> 4) * 8, ks * 2 + ((lane >> 3) & 1)), b0, b1, b2, b3);
                mma16816(s[2 * np], qf[ks], b0, b1);
                mma16816(s[2 * np + 1], qf[ks], b2, b3);
            }
        }
        // mask + online softmax (rows g and g+8 of this warp's 16)
        const int col0 = kt * BN + 2 * t;
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const bool ok = col0 + nt * 8 + e < seq;
                s[nt][e] = ok ? s[nt][e] : -INFINITY;
                s[nt][2 + e] = ok ? s[nt][2 + e] : -INFINITY;
                mx0 = fmaxf(mx0, s[nt][e]);
                mx1 = fmaxf(mx1, s[nt][2 + e]);
            }
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float nm0 = fmaxf(m0, mx0), nm1 = fmaxf(m1, mx1);   // finite: every kv tile has >= 1 valid column
        const float a0 = exp2f((m0 - nm0) * sl2), a1 = exp2f((m1 - nm1) * sl2);
        m0 = nm0; m1 = nm1;
        float rs0 = 0.f, rs1 = 0.f;
        uint32_t pf[4][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
            const float p0 = exp2f((s[nt][0] - m0) * sl2), p1 = exp2f((s[nt][1] - m0) * sl2);
            const float p2 = exp2f((s[nt][2] - m1) * sl2), p3 = exp2f((s[nt][3] - m1) * sl2);
            rs0 += p0 + p1; rs1 += p2 + p3;
            pf[nt >> 1][(nt & 1) * 2] = pack_bf16(p0, p1);
            pf[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
        }
        l0 = l0 * a0 + rs0; l1 = l1 * a1 + rs1;
#pragma unroll
        for (int dt = 0; dt < 16; ++dt) { o[dt][0] *= a0; o[dt][1] *= a0; o[dt][2] *= a1; o[dt][3] *= a1; }
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
            for (int dp = 0; dp < 8; ++dp) {   // pairs of 8-col d tiles
                uint32_t b0, b1, b2, b3;
                ldsm4t(vb + sw(kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, dp * 2 + (lane >> 4)), b0, b1, b2, b3);
                mma16816(o[2 * dp], pf[kk], b0, b1);
                mma16816(o[2 * dp + 1], pf[kk], b2, b3);
            }
        }
        __syncthreads();   // all warps done with buf before it is refilled two iterations later
    }
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float inv0 = 1.f / l0, inv1 = 1.f / l1;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        const int j = qt * BM + warp * 16 + g + half * 8;   // row inside the sequence
        if (j >= seq) continue;
        __nv_bfloat16* dst;
        if (j < w.n_vid) {
            int lt, lh, lw;
            dst = out_vid + win_token(w, j, H, W, &lt, &lh, &lw) * row_stride + head * D;
        } else {
            dst = txt_part + ((int64_t)wi * txt_len + (j - w.n_vid)) * row_stride + head * D;
        }
        const float inv = half ? inv1 : inv0;
#pragma unroll
        for (int dt = 0; dt < 16; ++dt)
            *reinterpret_cast<uint32_t*>(dst + dt * 8 + 2 * t) = pack_bf16(o[dt][half * 2] * inv, o[dt][half * 2 + 1] * inv);
    }
}

}  // namespace mma_attn

// out_txt[j, c] = mean over windows of txt_part[w, j, c]  (reference na.py:414: bf16 rows, fp32 accumulate, bf16 out)
__global__ void txt_pool_kernel(const __nv_bfloat16* __restrict__ part, __nv_bfloat16* __restrict__ out, int n_windows,
                                int64_t n8 /* Lt*heads*128/8 */) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n8) return;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int wv = 0; wv < n_windows; ++wv) {
        float f[8];
        unpack8(ldg8(part + ((int64_t)wv * n8 + i) * 8), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += f[e];
    }
    const float nwf = (float)n_windows;
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[e] = __fdiv_rn(acc[e], nwf);
    stg8(out + i * 8, pack8(acc));
}

int launch_txt_pool(const void* txt_part, void* out_txt, int n_windows, int txt_len, int heads, cudaStream_t st) {
    if (txt_len == 0) return SVR2_OK;
    const int64_t n8 = (int64_t)txt_len * heads * SVR2_HEAD_DIM / 8;
    txt_pool_kernel<<<(unsigned)((n8 + 127) / 128), 128, 0, st>>>((const __nv_bfloat16*)txt_part, (__nv_bfloat16*)out_txt, n_windows, n8);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(txt_pool)");
    return SVR2_OK;
}

int launch_window_attn_mma(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int H, int W, int txt_len, int heads, void* out_vid, void* txt_part,
                           cudaStream_t st) {
    int rc;
    static PerDeviceFlag smem_ok[1];
    if ((rc = ensure_dyn_smem(mma_attn::window_attn_mma_kernel, mma_attn::SMEM_BYTES, smem_ok[0], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    dim3 grid(n_work, heads, 2);
    mma_attn::window_attn_mma_kernel<<<grid, mma_attn::NTHREADS, mma_attn::SMEM_BYTES, st>>>(
        (const __nv_bfloat16*)q, (const __nv_bfloat16*)k, (const __nv_bfloat16*)v, tab, work, H, W, txt_len, heads,
        (__nv_bfloat16*)out_vid, (__nv_bfloat16*)txt_part);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(mma)");
    return SVR2_OK;
}

}  // namespace svr2
