// Fused adaLN kernels: RMSNorm + modulate ("in"), gate + residual ("out"), and the pair fused.
// Replaces CustomRMSNorm.forward (reference src/models/dit_3b/normalization.py:88-109) and AdaSingle.forward
// (modulation.py:65-118) plus the residual adds of mmsr_block.py:114,126.
//
// HBM-bound.  Algorithmic bytes (bf16): modnorm 4*L*D, gate_residual 6*L*D, fused pair 8*L*D.
// A thread owns 8 fixed columns, rows live in registers as packed bf16 (lossless: every value is bf16 already), all
// global accesses are 16-byte vectors, consecutive lanes touch consecutive 16 B.  The (embA + paramB) sums are formed
// once per CTA in registers; CTAs are persistent (grid = resident CTAs x #SM) and stride over groups of kRows rows.
//
// Numerics: the reference runs op by op in bf16 — ATen computes each op in fp32 and rounds the result to bf16.  Every
// such rounding point is reproduced with bf16r(), so results match the reference's bf16 path to the
// last bit except for the fp32 summation order inside mean().
#include <cstdlib>

#include "common.cuh"

namespace svr2 {

// Thread t of a CTA owns columns [8t, 8t+8) of EVERY row the CTA touches (D/8 threads, rounded up to a warp), so the
// (embA + paramB) vectors live in registers for the CTA's lifetime and a row costs each thread exactly one 16-byte load
// (two when the gate/residual is fused) and one (two) 16-byte store.  A CTA walks kRows rows per iteration — kRows
// independent loads in flight per thread — and the row sums are reduced with warp shuffles + one __syncthreads per
// iteration (double-buffered partials).  Few registers -> 3+ CTAs per SM -> >= 60 KB of loads in flight per SM.
constexpr int kMaxWarps = 16;     // D <= 4096

template <bool FUSE_GATE, int kRows, bool AFFINE = false, bool PREFETCH = true>
__global__ void __launch_bounds__(kMaxWarps * 32, 2)
modnorm_kernel(const __nv_bfloat16* __restrict__ x,       // FUSE_GATE: h (branch output), else the row to normalise
               const __nv_bfloat16* __restrict__ resid,   // FUSE_GATE only
               const __nv_bfloat16* __restrict__ emb, const __nv_bfloat16* __restrict__ gateB,
               const __nv_bfloat16* __restrict__ shiftB, const __nv_bfloat16* __restrict__ scaleB,
               __nv_bfloat16* __restrict__ out,           // FUSE_GATE only: gated residual sum
               __nv_bfloat16* __restrict__ y, int64_t L, int D, int gate_layer, int mod_layer, float eps, int has_mod,
               const __nv_bfloat16* __restrict__ normW = nullptr) {   // AFFINE only: RMSNorm weight (vid_out_norm)
    __shared__ float s_part[2][kRows][kMaxWarps];
    __shared__ __align__(16) float s_tot[2][4];      // row totals of the current group (kRows <= 4)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    const int col = threadIdx.x * 8;
    const bool act = col < D;
    bf16x8 sc_p, sh_p, g_p;      // (embA + paramB), rounded to bf16 like the reference's tensor add; kept packed
    {
        float sc[8], sh[8], g[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { sc[i] = 1.f; sh[i] = 0.f; g[i] = 1.f; }
        if (act) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int d = col + i;
                if (has_mod) {
                    sh[i] = __bfloat162float(emb[d * 6 + mod_layer * 3 + 0]) + __bfloat162float(shiftB[d]);
                    sc[i] = __bfloat162float(emb[d * 6 + mod_layer * 3 + 1]) + __bfloat162float(scaleB[d]);
                }
                if (FUSE_GATE) g[i] = __bfloat162float(emb[d * 6 + gate_layer * 3 + 2]) + __bfloat162float(gateB[d]);
            }
        }
        sc_p = pack8(sc); sh_p = pack8(sh); g_p = pack8(g);
    }
    bf16x8 w_p;
    if (AFFINE) {
        if (act) w_p = ldg8_stream(normW + col);
    }
    const float inv_d = 1.0f / (float)D;
    int buf = 0;
    // Software pipeline: the loads of the NEXT row group are issued before this group's reduction, barrier and stores, so a
    // CTA always has kRows (x2 when fused) 16-byte loads per thread in flight instead of alternating load / store phases.
    bf16x8 nv[kRows], nr[kRows];
    auto fetch = [&](int64_t row0) {
#pragma unroll
        for (int k = 0; k < kRows; ++k) {
            if (act && (row0 + k < L)) {
                nv[k] = ldg8_stream(x + (row0 + k) * D + col);
                if (FUSE_GATE) nr[k] = ldg8_stream(resid + (row0 + k) * D + col);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) nv[k].v[i] = __floats2bfloat162_rn(0.f, 0.f);
            }
        }
    };
    const int64_t stride = (int64_t)gridDim.x * kRows;
    fetch((int64_t)blockIdx.x * kRows);
    for (int64_t row0 = (int64_t)blockIdx.x * kRows; row0 < L; row0 += stride, buf ^= 1) {
        bf16x8 v[kRows], r[kRows];
#pragma unroll
        for (int k = 0; k < kRows; ++k) {
            v[k] = nv[k];
            if (FUSE_GATE) r[k] = nr[k];
        }
        if (PREFETCH) {
            if (row0 + stride < L) fetch(row0 + stride);
        }
        float ss[kRows];
#pragma unroll
        for (int k = 0; k < kRows; ++k) {
            if (FUSE_GATE) {
                if (act && (row0 + k < L)) {
                    v[k] = add8_rn(mul8_rn(v[k], g_p), r[k]);                         // mul_ rounds to bf16, then the add rounds
                    stg8(out + (row0 + k) * D + col, v[k]);
                }
            }
            float a = quad_sum_bf16x8(mul8(v[k], v[k]));                             // input.pow(2) is a bf16 tensor; fp32 sum
            a += __shfl_xor_sync(0xffffffffu, a, 4);
            a += __shfl_xor_sync(0xffffffffu, a, 8);
            a += __shfl_xor_sync(0xffffffffu, a, 16);
            ss[k] = a;
        }
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < kRows; ++k) s_part[buf][k][warp] = ss[k];
        }
        __syncthreads();
        // second level: warp k adds the n_warps partials of row k (one lane per partial, a 4-step butterfly) and publishes the
        // total; everybody then reads the kRows totals with one 16-byte load.  Letting every thread add all partials itself
        // cost ~10 FADD + 10 uniform branches per row and thread — a quarter of this issue-bound kernel's instructions.
        for (int k = warp; k < kRows; k += n_warps) {       // (a CTA of a narrow D has fewer warps than rows)
            float t = lane < n_warps ? s_part[buf][k][lane] : 0.f;
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
            if (lane == 0) s_tot[buf][k] = t;
        }
        __syncthreads();
        const float4 tot4 = *reinterpret_cast<const float4*>(s_tot[buf]);
#pragma unroll
        for (int k = 0; k < kRows; ++k) {
            if (!(act && (row0 + k < L))) continue;
            const float tot = k == 0 ? tot4.x : (k == 1 ? tot4.y : (k == 2 ? tot4.z : tot4.w));
            const float var = bf16r(tot * inv_d);                                     // .mean() -> bf16
            const float rms = bf16r(sqrt_approx(bf16r(var + eps)));
            const float rcp = rcp_approx(rms);
            float f[8];
            unpack8(v[k], f);
#pragma unroll
            for (int i = 0; i < 8; ++i) f[i] = f[i] * rcp;       // rounds to the same bf16 as x / rms (common.cuh)
            bf16x8 yv = pack8(f);                                                     // (x / rms) -> bf16
            if (AFFINE) yv = mul8_rn(yv, w_p);                                        // * weight -> bf16 (normalization.py:108)
            if (has_mod) yv = add8_rn(mul8_rn(yv, sc_p), sh_p);                       // * scale -> bf16, + shift -> bf16
            stg8(y + (row0 + k) * D + col, yv);
        }
        if (!PREFETCH) {
            if (row0 + stride < L) fetch(row0 + stride);      // A/B variant: load / compute / store phases alternate
        }
    }
}

// out = h * gate + resid, pure streaming.  A thread owns 8 fixed columns (gate values live in registers) and walks rows.
__global__ void __launch_bounds__(1024)
gate_residual_kernel(const __nv_bfloat16* __restrict__ h, const __nv_bfloat16* __restrict__ resid,
                     const __nv_bfloat16* __restrict__ emb, const __nv_bfloat16* __restrict__ gateB,
                     __nv_bfloat16* __restrict__ out, int64_t L, int D, int layer, int has_gate) {
    const int col = threadIdx.x * 8;
    if (col >= D) return;
    bf16x8 g_p;
    {
        float g[8];
#pragma unroll
        for (int i = 0; i < 8; ++i)
            g[i] = has_gate ? __bfloat162float(emb[(col + i) * 6 + layer * 3 + 2]) + __bfloat162float(gateB[col + i]) : 1.f;
        g_p = pack8(g);       // gateA + gateB is a bf16 tensor in the reference
    }
    int64_t row = blockIdx.x;
    const int64_t step = gridDim.x;
    for (; row + 3 * step < L; row += 4 * step) {     // 8 independent 16-B loads in flight per thread
        bf16x8 hv[4], rv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            hv[u] = ldg8_stream(h + (row + u * step) * D + col);
            rv[u] = ldg8_stream(resid + (row + u * step) * D + col);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            stg8_stream(out + (row + u * step) * D + col, add8_rn(has_gate ? mul8_rn(hv[u], g_p) : hv[u], rv[u]));
        }
    }
    for (; row < L; row += step) {
        const bf16x8 hv = ldg8_stream(h + row * D + col);
        stg8_stream(out + row * D + col, add8_rn(has_gate ? mul8_rn(hv, g_p) : hv, ldg8_stream(resid + row * D + col)));
    }
}

// out = silu(gate) * up  — the SwiGLU activation between proj_in_gate/proj_in and proj_out (reference mlp.py:60-62:
// `F.silu(self.proj_in_gate(x)) * self.proj_in(x)`, two ATen kernels = 5 [L,hidden] passes; here 3).  silu is rounded to
// bf16 before the product like the reference's intermediate tensor.  HBM-bound: 6 B per element.
__global__ void __launch_bounds__(256)
swiglu_kernel(const __nv_bfloat16* __restrict__ gate, const __nv_bfloat16* __restrict__ up,
              __nv_bfloat16* __restrict__ out, int64_t n8) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n8; i += 4 * stride) {       // 8 independent 16-B loads in flight per thread
        bf16x8 g[4], u[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            g[k] = ldg8_stream(gate + (i + k * stride) * 8);
            u[k] = ldg8_stream(up + (i + k * stride) * 8);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float f[8];
            unpack8(g[k], f);
#pragma unroll
            for (int e = 0; e < 8; ++e) f[e] = __fdividef(f[e], 1.0f + __expf(-f[e]));
            stg8(out + (i + k * stride) * 8, mul8(pack8(f), u[k]));
        }
    }
    for (; i < n8; i += stride) {
        float f[8];
        unpack8(ldg8_stream(gate + i * 8), f);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = __fdividef(f[e], 1.0f + __expf(-f[e]));
        stg8(out + i * 8, mul8(pack8(f), ldg8_stream(up + i * 8)));
    }
}

// The same on the output of ONE GEMM with the weights [W_gate ; W_up] stacked: row r holds gate at [0, hidden) and up at
// [hidden, 2 hidden).  (Experiment of run 72: does cuBLAS run the MLP in-projection better as one N = 2 hidden GEMM?)
__global__ void __launch_bounds__(256)
swiglu_rows_kernel(const __nv_bfloat16* __restrict__ gate_up, __nv_bfloat16* __restrict__ out, int64_t rows, int h8) {
    const int64_t n8 = rows * h8;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / h8;
        const int c = (int)(i - r * h8);
        const __nv_bfloat16* g = gate_up + (r * 2 * h8 + c) * 8;
        float f[8];
        unpack8(ldg8_stream(g), f);
        const bf16x8 u = ldg8_stream(g + (int64_t)h8 * 8);
#pragma unroll
        for (int e = 0; e < 8; ++e) f[e] = __fdividef(f[e], 1.0f + __expf(-f[e]));
        stg8(out + i * 8, mul8(pack8(f), u));
    }
}

template <bool FUSE>
static int launch_modnorm(const void* x, const void* resid, const void* emb, const void* gateB, const void* shiftB,
                          const void* scaleB, void* out, void* y, int64_t L, int D, int gate_layer, int mod_layer,
                          float eps, int has_mod, cudaStream_t st, const char* name) {
    if (L == 0) return SVR2_OK;
    const int threads = ((D / 8) + 31) / 32 * 32;
    constexpr int kRows = FUSE ? 2 : 4;          // same bytes in flight per thread either way (the fused kernel loads h and resid)
    static int occ_dev[kMaxDevices][kMaxWarps + 1] = {};   // resident CTAs per SM for this block size (queried once per device)
    int* occ = occ_dev[cur_device()];
    const int wi = threads / 32;
    if (occ[wi] == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, modnorm_kernel<FUSE, kRows>, threads, 0) != cudaSuccess || n <= 0) n = 2;
        occ[wi] = n;
    }
    const int64_t want = (L + kRows - 1) / kRows;
    const int64_t cap = (int64_t)sm_count() * occ[wi];
    const int grid = (int)(want < cap ? want : cap);
    constexpr bool prefetch = true;      // next row group's loads issued before this group's reduction
    auto kern = prefetch ? modnorm_kernel<FUSE, kRows, false, true> : modnorm_kernel<FUSE, kRows, false, false>;
    kern<<<grid, threads, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)resid, (const __nv_bfloat16*)emb,
                                   (const __nv_bfloat16*)gateB, (const __nv_bfloat16*)shiftB, (const __nv_bfloat16*)scaleB,
                                   (__nv_bfloat16*)out, (__nv_bfloat16*)y, L, D, gate_layer, mod_layer, eps, has_mod, nullptr);
    SVR2_CHECK_LAUNCH(name);
    return SVR2_OK;
}

}  // namespace svr2

extern "C" int svr2_modnorm_fwd(const void* x, const void* emb, const void* shiftB, const void* scaleB, void* y,
                                int64_t L, int D, int layer, float eps, void* stream) {
    SVR2_CHECK_ARG(x && y, "svr2_modnorm_fwd: null x/y");
    SVR2_CHECK_ARG(L >= 0 && D > 0 && D % 8 == 0 && D <= 4096, "svr2_modnorm_fwd: D=%d must be a multiple of 8, <= 4096", D);
    SVR2_CHECK_ARG(layer == 0 || layer == 1, "svr2_modnorm_fwd: layer must be 0 (attn) or 1 (mlp)");
    const int has_mod = (shiftB != nullptr || scaleB != nullptr || emb != nullptr) ? 1 : 0;
    SVR2_CHECK_ARG(!has_mod || (shiftB && scaleB && emb), "svr2_modnorm_fwd: emb, shiftB and scaleB go together");
    return svr2::launch_modnorm<false>(x, nullptr, emb, nullptr, shiftB, scaleB, nullptr, y, L, D, 0, layer, eps,
                                       has_mod, (cudaStream_t)stream, "svr2_modnorm_fwd");
}

// RMSNorm WITH elementwise-affine weight, then the "in" modulation: the model's output norm + vid_out_ada
// (reference dit_3b/nadit.py:233-244; CustomRMSNorm normalization.py:88-109 with weight; quirk Q1 — the caller passes
// layer 0, the attn-layer (shift, scale) pair the reference's cache hands to vid_out_ada).
extern "C" int svr2_modnorm_affine_fwd(const void* x, const void* weight, const void* emb, const void* shiftB,
                                       const void* scaleB, void* y, int64_t L, int D, int layer, float eps, void* stream) {
    SVR2_CHECK_ARG(x && y && weight, "svr2_modnorm_affine_fwd: null x/y/weight");
    SVR2_CHECK_ARG(L >= 0 && D > 0 && D % 8 == 0 && D <= 4096, "svr2_modnorm_affine_fwd: D=%d must be a multiple of 8, <= 4096", D);
    SVR2_CHECK_ARG(layer == 0 || layer == 1, "svr2_modnorm_affine_fwd: layer must be 0 (attn) or 1 (mlp)");
    const int has_mod = (shiftB != nullptr || scaleB != nullptr || emb != nullptr) ? 1 : 0;
    SVR2_CHECK_ARG(!has_mod || (shiftB && scaleB && emb), "svr2_modnorm_affine_fwd: emb, shiftB and scaleB go together");
    if (L == 0) return SVR2_OK;
    const int threads = ((D / 8) + 31) / 32 * 32;
    const int64_t want = (L + 3) / 4, cap = (int64_t)svr2::sm_count() * 2;
    svr2::modnorm_kernel<false, 4, true><<<(int)(want < cap ? want : cap), threads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)x, nullptr, (const __nv_bfloat16*)emb, nullptr, (const __nv_bfloat16*)shiftB,
        (const __nv_bfloat16*)scaleB, nullptr, (__nv_bfloat16*)y, L, D, 0, layer, eps, has_mod, (const __nv_bfloat16*)weight);
    SVR2_CHECK_LAUNCH("svr2_modnorm_affine_fwd");
    return SVR2_OK;
}

extern "C" int svr2_gate_residual_modnorm_fwd(const void* h, const void* resid, const void* emb, const void* gateB,
                                              const void* shiftB2, const void* scaleB2, void* out, void* y, int64_t L,
                                              int D, int layer, float eps, void* stream) {
    SVR2_CHECK_ARG(h && resid && emb && gateB && shiftB2 && scaleB2 && out && y, "svr2_gate_residual_modnorm_fwd: null pointer");
    SVR2_CHECK_ARG(L >= 0 && D > 0 && D % 8 == 0 && D <= 4096, "svr2_gate_residual_modnorm_fwd: bad D=%d", D);
    SVR2_CHECK_ARG(layer == 0, "svr2_gate_residual_modnorm_fwd: gate layer must be 0 (attn); modulation uses layer 1");
    return svr2::launch_modnorm<true>(h, resid, emb, gateB, shiftB2, scaleB2, out, y, L, D, layer, layer + 1, eps, 1,
                                      (cudaStream_t)stream, "svr2_gate_residual_modnorm_fwd");
}

extern "C" int svr2_gate_residual_fwd(const void* h, const void* resid, const void* emb, const void* gateB, void* out,
                                      int64_t L, int D, int layer, void* stream) {
    SVR2_CHECK_ARG(h && resid && out, "svr2_gate_residual_fwd: null pointer");
    SVR2_CHECK_ARG(L >= 0 && D > 0 && D % 8 == 0 && D <= 4096, "svr2_gate_residual_fwd: bad D=%d", D);
    SVR2_CHECK_ARG(layer == 0 || layer == 1, "svr2_gate_residual_fwd: layer must be 0 or 1");
    const int has_gate = (emb != nullptr || gateB != nullptr) ? 1 : 0;
    SVR2_CHECK_ARG(!has_gate || (emb && gateB), "svr2_gate_residual_fwd: emb and gateB go together");
    if (L == 0) return SVR2_OK;
    const int threads = ((D / 8) + 31) / 32 * 32;
    const int64_t cap = (int64_t)svr2::sm_count() * (2048 / threads);
    const int grid = (int)(L < cap ? L : cap);
    svr2::gate_residual_kernel<<<grid, threads, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)h, (const __nv_bfloat16*)resid, (const __nv_bfloat16*)emb, (const __nv_bfloat16*)gateB,
        (__nv_bfloat16*)out, L, D, layer, has_gate);
    SVR2_CHECK_LAUNCH("svr2_gate_residual_fwd");
    return SVR2_OK;
}

extern "C" int svr2_swiglu_fwd(const void* gate, const void* up, void* out, int64_t n, void* stream) {
    SVR2_CHECK_ARG(gate && up && out, "svr2_swiglu_fwd: null pointer");
    SVR2_CHECK_ARG(n >= 0 && n % 8 == 0, "svr2_swiglu_fwd: element count %lld must be a multiple of 8", (long long)n);
    if (n == 0) return SVR2_OK;
    const int64_t n8 = n / 8;
    int64_t grid = (n8 + 256 * 4 - 1) / (256 * 4);
    const int64_t cap = (int64_t)svr2::sm_count() * 8;
    if (grid > cap) grid = cap;
    svr2::swiglu_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)gate, (const __nv_bfloat16*)up,
                                                                           (__nv_bfloat16*)out, n8);
    SVR2_CHECK_LAUNCH("svr2_swiglu_fwd");
    return SVR2_OK;
}

extern "C" int svr2_swiglu_rows_fwd(const void* gate_up, void* out, int64_t rows, int hidden, void* stream) {
    SVR2_CHECK_ARG(gate_up && out, "svr2_swiglu_rows_fwd: null pointer");
    SVR2_CHECK_ARG(rows >= 0 && hidden > 0 && hidden % 8 == 0, "svr2_swiglu_rows_fwd: hidden %d must be a positive multiple of 8", hidden);
    if (rows == 0) return SVR2_OK;
    const int64_t n8 = rows * (hidden / 8);
    int64_t grid = (n8 + 255) / 256;
    const int64_t cap = (int64_t)svr2::sm_count() * 8;
    if (grid > cap) grid = cap;
    svr2::swiglu_rows_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)gate_up, (__nv_bfloat16*)out, rows, hidden / 8);
    SVR2_CHECK_LAUNCH("svr2_swiglu_rows_fwd");
    return SVR2_OK;
}
