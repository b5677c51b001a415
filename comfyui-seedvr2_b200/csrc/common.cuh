// Shared helpers for the seedvr2_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/seedvr2_b200.h"

namespace svr2 {

void set_error(const char* fmt, ...);

#define SVR2_CHECK_ARG(cond, ...)                 \
    do {                                          \
        if (!(cond)) {                            \
            ::svr2::set_error(__VA_ARGS__);       \
            return SVR2_E_INVALID;                \
        }                                         \
    } while (0)

#define SVR2_CHECK_LAUNCH(name)                                                          \
    do {                                                                                 \
        cudaError_t e__ = cudaGetLastError();                                            \
        if (e__ != cudaSuccess) {                                                        \
            ::svr2::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));   \
            return SVR2_E_CUDA;                                                          \
        }                                                                                \
    } while (0)

struct __align__(16) bf16x8 {
    __nv_bfloat162 v[4];
};

// Round an fp32 value to bf16 precision and return it as fp32: the rounding point of every elementwise ATen op of the
// reference's bf16 path (ATen computes in fp32 and stores bf16).
__device__ __forceinline__ float bf16r(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }

// register selects the compiler cannot turn into an indexed local-memory array
__device__ __forceinline__ uint32_t sel_u32(bool c, uint32_t a, uint32_t b) {
    uint32_t r;
    asm("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %1, 0;\n\tselp.u32 %0, %2, %3, p;\n\t}" : "=r"(r) : "r"((uint32_t)c), "r"(a), "r"(b));
    return r;
}
__device__ __forceinline__ int sel_i32(bool c, int a, int b) { return (int)sel_u32(c, (uint32_t)a, (uint32_t)b); }

__device__ __forceinline__ bf16x8 ldg8(const __nv_bfloat16* p) {
    bf16x8 r;
    *reinterpret_cast<uint4*>(&r) = __ldg(reinterpret_cast<const uint4*>(p));
    return r;
}
// streaming variants: data touched once, keep it out of L1
__device__ __forceinline__ bf16x8 ldg8_stream(const __nv_bfloat16* p) {
    bf16x8 r;
    uint4 u;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w) : "l"(p));
    *reinterpret_cast<uint4*>(&r) = u;
    return r;
}
__device__ __forceinline__ void stg8(__nv_bfloat16* p, const bf16x8& v) {
    *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(&v);
}
__device__ __forceinline__ void stg8_stream(__nv_bfloat16* p, const bf16x8& v) {
    const uint4 u = *reinterpret_cast<const uint4*>(&v);
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(u.x), "r"(u.y), "r"(u.z), "r"(u.w) : "memory");
}
__device__ __forceinline__ void unpack8(const bf16x8& v, float* f) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 t = __bfloat1622float2(v.v[i]);
        f[2 * i] = t.x;
        f[2 * i + 1] = t.y;
    }
}
__device__ __forceinline__ bf16x8 pack8(const float* f) {
    bf16x8 r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return r;
}

// Element-wise bf16 product, rounded once to bf16.  The exact product of two bf16 values fits in fp32 (16 significant
// bits), so HMUL2.BF16 equals ATen's "compute in fp32, round the result to bf16" bit for bit — at half the instructions
// of unpack / FMUL / round.  (Additions are NOT done this way: fp32-add-then-round double-rounds, HADD2.BF16 may not.)
__device__ __forceinline__ bf16x8 mul8(const bf16x8& a, const bf16x8& b) {
    bf16x8 r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.v[i] = __hmul2(a.v[i], b.v[i]);
    return r;
}
// a * b and a + b as the reference's bf16 tensor ops compute them (fp32 result rounded once to bf16) on packed pairs.
// The product of two bf16 values is exact in fp32; their sum is exact in fp32 unless the exponents are > 16 bits apart,
// where the small operand cannot reach the bf16 rounding boundary either way — so HMUL2.BF16 / HADD2.BF16 give the same
// bits as "fp32 op, then round".  The _rn intrinsics keep the compiler from contracting the pair into one HFMA2 (which
// would round once instead of twice).
__device__ __forceinline__ bf16x8 mul8_rn(const bf16x8& a, const bf16x8& b) {
    bf16x8 r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.v[i] = __hmul2_rn(a.v[i], b.v[i]);
    return r;
}
__device__ __forceinline__ bf16x8 add8_rn(const bf16x8& a, const bf16x8& b) {
    bf16x8 r;
#pragma unroll
    for (int i = 0; i < 4; ++i) r.v[i] = __hadd2_rn(a.v[i], b.v[i]);
    return r;
}
// bf16(fp32(x / d)) for bf16-valued x and d, without the division.  The quotient of two 8-bit significands is never a
// bf16 rounding midpoint (its odd part would have to exceed 8 bits) and lies at least 2^-17 (relative) away from one, so
// x * rcp with rcp within a few fp32 ulps of 1/d rounds to the same bf16 value as the reference's correctly rounded fp32
// division followed by its bf16 store.  The same holds for sqrt of a bf16 value (distance >= 2^-19).  Both checked
// exhaustively over all significand pairs with the approximation perturbed by +-2 / +-4 ulps (tests/test_oracle_golden.py).
__device__ __forceinline__ float rcp_approx(float d) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return r;
}
__device__ __forceinline__ float sqrt_approx(float v) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}

// Sum of the 8 bf16 values held by each of the 4 lanes of a quad (lanes 4q .. 4q+3), in fp32, in ONE instruction: the four
// registers are the A fragment of mma.sync.m16n8k16 (a0, a2 -> row q, a1, a3 -> row q + 8) and B is all ones, so D[q][*] +
// D[q+8][*] is the quad's total.  Replaces 8 conversions + 8 FADD + 2 butterfly steps of a row-sum — the RMSNorm kernels are
// issue bound, not memory bound (profiles/r01_run25_ncu_elementwise_c1.txt).  Products with 1.0 are exact and the accumulation
// is fp32; like any change of summation order it can move a sum by an fp32 ulp.  The whole warp must be converged.
__device__ __forceinline__ float quad_sum_bf16x8(const bf16x8& v) {
    const uint4 a = *reinterpret_cast<const uint4*>(&v);
    float c0, c1, c2, c3;
    const uint32_t ones = 0x3F803F80u;
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%8}, {%9,%9,%9,%9};"
                 : "=f"(c0), "=f"(c1), "=f"(c2), "=f"(c3)
                 : "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(ones), "f"(0.f));
    (void)c1; (void)c3;
    return c0 + c2;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// device window table entry (SVR2_WIN_FIELDS int32)
struct Win {
    int t0, h0, w0, nt, nh, nw, row_start, n_vid;
};
__device__ __forceinline__ Win load_win(const int32_t* tab, int i) {
    const int4 a = __ldg(reinterpret_cast<const int4*>(tab) + 2 * i);
    const int4 b = __ldg(reinterpret_cast<const int4*>(tab) + 2 * i + 1);
    Win w;
    w.t0 = a.x; w.h0 = a.y; w.w0 = a.z; w.nt = a.w;
    w.nh = b.x; w.nw = b.y; w.row_start = b.z; w.n_vid = b.w;
    return w;
}
// global token id of the j-th video row of a window (row-major t,h,w inside the window)
__device__ __forceinline__ int64_t win_token(const Win& w, int j, int H, int W, int* lt, int* lh, int* lw) {
    const int hw = w.nh * w.nw;
    const int t = j / hw, r = j - t * hw;
    const int h = r / w.nw, x = r - h * w.nw;
    *lt = t; *lh = h; *lw = x;
    return ((int64_t)(w.t0 + t) * H + (w.h0 + h)) * W + (w.w0 + x);
}

// ---- per-device host-side caches (one process may drive several GPUs: ops._call switches the current device per call)
constexpr int kMaxDevices = 64;
inline int cur_device() {
    int dev = 0;
    cudaGetDevice(&dev);
    return (dev < 0 || dev >= kMaxDevices) ? 0 : dev;
}
inline int sm_count() {
    static int n[kMaxDevices] = {0};
    const int dev = cur_device();
    if (n[dev] == 0) {
        cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
        if (n[dev] <= 0) n[dev] = 148;
    }
    return n[dev];
}
// Opt a kernel in to `bytes` of dynamic shared memory on the CURRENT device, once per device (the attribute is per device).
struct PerDeviceFlag {
    bool set[kMaxDevices] = {};
};
template <class Kernel>
inline int ensure_dyn_smem(Kernel kernel, int bytes, PerDeviceFlag& done, const char* what) {
    const int dev = cur_device();
    if (done.set[dev]) return SVR2_OK;
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e != cudaSuccess) {
        set_error("%s: cannot reserve %d B shared memory: %s", what, bytes, cudaGetErrorString(e));
        return SVR2_E_CUDA;
    }
    done.set[dev] = true;
    return SVR2_OK;
}

}  // namespace svr2
