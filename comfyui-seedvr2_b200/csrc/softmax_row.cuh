// Row-softmax building blocks of the tcgen05 window-attention kernel: one thread owns one query row, a 32-column chunk of
// the S row lives in 32 registers.  Shared with scripts/microbench/softmax_step.cu, which times them in isolation.
#pragma once
#include <cuda_bf16.h>
#include <cstdint>

namespace svr2 {
namespace rowsm {

constexpr float kScaleLog2 = 0.08838834764831845f * 1.4426950408889634f;   // 1/sqrt(128) * log2(e)

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

// packed fp32 pairs (Blackwell FFMA2 / FADD2): two lanes of arithmetic per issue slot
__device__ __forceinline__ uint64_t pack_f2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack_f2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t add_f2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// row max of one 32-column chunk (masked columns were set to -inf beforehand)
__device__ __forceinline__ float chunk_max(const uint32_t* v) {
    float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
    for (int c = 0; c < 32; c += 8) {
        m0 = max3(m0, __uint_as_float(v[c]), __uint_as_float(v[c + 1]));
        m1 = max3(m1, __uint_as_float(v[c + 2]), __uint_as_float(v[c + 3]));
        m2 = max3(m2, __uint_as_float(v[c + 4]), __uint_as_float(v[c + 5]));
        m3 = max3(m3, __uint_as_float(v[c + 6]), __uint_as_float(v[c + 7]));
    }
    return fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
}
// P = exp2(s k - mb) for one 32-column chunk, packed to bf16 pairs; returns the chunk's row sum.  Branch free: FFMA2 (1/2
// issue slot per element), MUFU.EX2, F2FP (1/2), FADD2 (1/2).  -inf (masked) columns give exactly 0.  The exponent never
// exceeds kRescaleThreshold (lazy rescale), so no clamp is needed.
__device__ __forceinline__ float chunk_exp(const uint32_t* v, uint32_t* pk, float mb) {
    const uint64_t kk = pack_f2(kScaleLog2, kScaleLog2), nmb = pack_f2(-mb, -mb);
    uint64_t acc0 = pack_f2(0.f, 0.f), acc1 = acc0;
#pragma unroll
    for (int c = 0; c < 16; c += 2) {
        float x0, x1, x2, x3;
        unpack_f2(fma_f2(pack_f2(__uint_as_float(v[2 * c]), __uint_as_float(v[2 * c + 1])), kk, nmb), x0, x1);
        unpack_f2(fma_f2(pack_f2(__uint_as_float(v[2 * c + 2]), __uint_as_float(v[2 * c + 3])), kk, nmb), x2, x3);
        const float p0 = ex2(x0), p1 = ex2(x1), p2 = ex2(x2), p3 = ex2(x3);
        pk[c] = pack2(p0, p1);
        pk[c + 1] = pack2(p2, p3);
        acc0 = add_f2(acc0, pack_f2(p0, p1));
        acc1 = add_f2(acc1, pack_f2(p2, p3));
    }
    float s0, s1;
    unpack_f2(add_f2(acc0, acc1), s0, s1);
    return s0 + s1;
}
__device__ __forceinline__ void mask_chunk(uint32_t* v, int n_valid) {      // columns >= n_valid -> -inf
#pragma unroll
    for (int c = 0; c < 32; ++c) v[c] = (c < n_valid) ? v[c] : 0xff800000u;
}


// running row max over one 32-column chunk kept in 8 independent chains (a lone warp issuing dependent half-rate FMNMX3s
// is latency bound: 4 chains took ~8 cycles per instruction)
__device__ __forceinline__ void chunk_max8(float* m, const uint32_t* v) {
#pragma unroll
    for (int q = 0; q < 8; ++q) m[q] = max3(m[q], __uint_as_float(v[2 * q]), __uint_as_float(v[2 * q + 1]));
#pragma unroll
    for (int q = 0; q < 8; ++q) m[q] = max3(m[q], __uint_as_float(v[16 + 2 * q]), __uint_as_float(v[16 + 2 * q + 1]));
}
// row max of a whole 128-column row with 8 independent FMNMX3 chains (the per-chunk version has 4 chains of dependent
// half-rate instructions and was latency bound: ~8 cycles per instruction for a lone warp)
__device__ __forceinline__ float row_max128(const uint32_t* v) {
    float m[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) m[q] = fmaxf(__uint_as_float(v[2 * q]), __uint_as_float(v[2 * q + 1]));
#pragma unroll
    for (int c = 16; c < 128; c += 16)
#pragma unroll
        for (int q = 0; q < 8; ++q) m[q] = max3(m[q], __uint_as_float(v[c + 2 * q]), __uint_as_float(v[c + 2 * q + 1]));
    return fmaxf(max3(m[0], m[1], m[2]), max3(max3(m[3], m[4], m[5]), m[6], m[7]));
}

// exp2 of a pair on the FMA pipe instead of MUFU (the exp2 unit, 16 results/clk/SM, is as loaded as the tensor core in
// this kernel): Cody-Waite split x = i + f, f in [-0.5, 0.5], 2^f by a degree-3 polynomial (relative error 1.1e-4, far below
// bf16's 2^-9), 2^i by an add into the exponent field.  x is clamped at -126 (result ~1e-38, i.e. nothing).
__device__ __forceinline__ void ex2_fma_pair(uint64_t xx, float& p0, float& p1) {
    float x0, x1;
    unpack_f2(xx, x0, x1);
    xx = pack_f2(fmaxf(x0, -126.f), fmaxf(x1, -126.f));
    const float magic = 12582912.f;                                  // 1.5 * 2^23: the sum's low mantissa bits hold round(x)
    const uint64_t t = add_f2(xx, pack_f2(magic, magic));
    const uint64_t xi = add_f2(t, pack_f2(-magic, -magic));
    const uint64_t f = fma_f2(xi, pack_f2(-1.f, -1.f), xx);
    uint64_t p = fma_f2(pack_f2(0.05459282f, 0.05459282f), f, pack_f2(0.24221784f, 0.24221784f));
    p = fma_f2(p, f, pack_f2(0.6933686f, 0.6933686f));
    p = fma_f2(p, f, pack_f2(1.f, 1.f));
    float q0, q1, t0, t1;
    unpack_f2(p, q0, q1);
    unpack_f2(t, t0, t1);
    p0 = __int_as_float(__float_as_int(q0) + (__float_as_int(t0) << 23));
    p1 = __int_as_float(__float_as_int(q1) + (__float_as_int(t1) << 23));
}
// pair q of a chunk goes to the FMA pipe when this is true: E of every 8 pairs, evenly spread
template <int E> __host__ __device__ constexpr bool on_fma_pipe(int q) { return ((q % 8 + 1) * E) / 8 != ((q % 8) * E) / 8; }

// chunk_exp with E/8 of the exponentials evaluated on the FMA pipe
template <int E>
__device__ __forceinline__ float chunk_exp_mix(const uint32_t* v, uint32_t* pk, float mb) {
    const uint64_t kk = pack_f2(kScaleLog2, kScaleLog2), nmb = pack_f2(-mb, -mb);
    uint64_t acc0 = pack_f2(0.f, 0.f), acc1 = acc0;
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        const uint64_t xx = fma_f2(pack_f2(__uint_as_float(v[2 * q]), __uint_as_float(v[2 * q + 1])), kk, nmb);
        float p0, p1;
        if (on_fma_pipe<E>(q)) {
            ex2_fma_pair(xx, p0, p1);
        } else {
            float x0, x1;
            unpack_f2(xx, x0, x1);
            p0 = ex2(x0);
            p1 = ex2(x1);
        }
        pk[q] = pack2(p0, p1);
        if (q & 1) acc1 = add_f2(acc1, pack_f2(p0, p1));
        else acc0 = add_f2(acc0, pack_f2(p0, p1));
    }
    float s0, s1;
    unpack_f2(add_f2(acc0, acc1), s0, s1);
    return s0 + s1;
}
}  // namespace rowsm
}  // namespace svr2
