// svr2_wavelet_level_fwd / svr2_colorfix_finish_fwd — the pipeline's default colour correction ("wavelet") and the conversion of the
// corrected frames to the output format, on the GPU.
//
// Replaces, per decoded batch (reference src/core/generation_phases.py:1286-1337, src/utils/color_fix.py:122-246):
//   wavelet_blur            3x3 binomial kernel [1 2 1]^T [1 2 1] / 16, dilation = radius, replicate padding — a grouped conv2d
//   wavelet_decomposition   5 levels, radius 1, 2, 4, 8, 16 (each capped at min(H, W) / 8): low_{i+1} = blur(low_i),
//                           high += low_i; high -= low_{i+1}   (two separately rounded in-place ops per level)
//   wavelet_reconstruction  out = clamp(high(content) + low_5(style), -1, 1)
//   normalisation           clamp(-1, 1) * 0.5 + 0.5  -> [T, H, W, C] in [0, 1]   (generation_phases.py:1337)
//   uint8 frames            (x * 255).astype(uint8), RGB -> BGR for OpenCV          (inference_cli.py:691-695,734-738)
// One launch per level and image (HBM-bound stencil: the nine taps of a pixel sit 2r+1 rows apart, which L2 holds, so the
// algorithmic traffic is one read + one write of the plane set per level: 8 B per element in fp32, 4 B in bf16), one finishing
// launch.  Arithmetic follows ATen's depthwise convolution (nine FMAs in row-major tap order from 0, fp32 accumulator) and rounds
// to bf16 where the reference's bf16 tensors do (the pipeline post-processes in its compute dtype, generation_phases.py:1229-1237).
#include "common.cuh"

namespace svr2 {
namespace post {

template <bool BF16> __device__ __forceinline__ float ld(const void* p, int64_t i) {
    if constexpr (BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
    else return reinterpret_cast<const float*>(p)[i];
}
template <bool BF16> __device__ __forceinline__ void st(void* p, int64_t i, float v) {
    if constexpr (BF16) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16_rn(v);
    else reinterpret_cast<float*>(p)[i] = v;
}
template <bool BF16> __device__ __forceinline__ float rnd(float v) { return BF16 ? bf16r(v) : v; }

struct Strides {
    int64_t t, c, y, x;      // element strides of a [T, C, H, W]-indexed tensor (any memory layout)
};

// low[t, c, y, x] = blur_r(src)[t, c, y, x];  high (optional, planar like low) = (high + src) - low, `first`: high starts at zero
template <bool BF16>
__global__ void __launch_bounds__(256) wavelet_level_kernel(const void* __restrict__ src, Strides ss, void* __restrict__ low,
                                                            void* __restrict__ high, int first, int C, int H, int W, int radius) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    const int plane = blockIdx.z;
    if (x >= W) return;
    const int t = plane / C, c = plane - t * C;
    const int64_t sbase = (int64_t)t * ss.t + (int64_t)c * ss.c;
    const int ys[3] = {max(y - radius, 0), y, min(y + radius, H - 1)};
    const int xs[3] = {max(x - radius, 0), x, min(x + radius, W - 1)};
    const float wgt[3] = {0.25f, 0.5f, 0.25f};
    float acc = 0.f;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
        for (int kx = 0; kx < 3; ++kx)
            acc = fmaf(ld<BF16>(src, sbase + ys[ky] * ss.y + xs[kx] * ss.x), wgt[ky] * wgt[kx], acc);
    const float lo = rnd<BF16>(acc);
    const int64_t o = ((int64_t)plane * H + y) * W + x;
    st<BF16>(low, o, lo);
    if (high != nullptr) {
        const float centre = ld<BF16>(src, sbase + (int64_t)y * ss.y + (int64_t)x * ss.x);
        const float h0 = first ? 0.f : ld<BF16>(high, o);
        st<BF16>(high, o, rnd<BF16>(__fsub_rn(rnd<BF16>(__fadd_rn(h0, centre)), lo)));
    }
}

// out[t, y, x, c] = clamp(clamp(high + low_style, -1, 1), -1, 1) * 0.5 + 0.5  (and / or its uint8 form)
template <bool BF16>
__global__ void __launch_bounds__(256) colorfix_finish_kernel(const void* __restrict__ high, const void* __restrict__ low_style,
                                                              void* __restrict__ out, uint8_t* __restrict__ out_u8, int bgr, int C,
                                                              int H, int W, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;       // index into [T, H, W, C]
    if (i >= n) return;
    const int c = (int)(i % C);
    const int64_t pix = i / C;
    const int x = (int)(pix % W);
    const int64_t ty = pix / W;
    const int y = (int)(ty % H);
    const int64_t t = ty / H;
    const int64_t p = ((t * C + c) * H + y) * W + x;
    float v = rnd<BF16>(__fadd_rn(ld<BF16>(high, p), ld<BF16>(low_style, p)));       // content_high_freq.add_(style_low_freq)
    v = fminf(fmaxf(v, -1.f), 1.f);                                                   // .clamp_(-1, 1)  (twice in the reference: idempotent)
    v = rnd<BF16>(__fmul_rn(v, 0.5f));                                                // .mul_(0.5)
    v = rnd<BF16>(__fadd_rn(v, 0.5f));                                                // .add_(0.5)
    if (out != nullptr) st<BF16>(out, i, v);
    if (out_u8 != nullptr) {
        // numpy: (float32 frame * 255.0).astype(uint8) truncates; values are in [0, 1] here
        const int cc = (bgr && C >= 3 && c < 3) ? 2 - c : c;
        out_u8[pix * C + cc] = (uint8_t)(int)__fmul_rn(v, 255.0f);
    }
}

// ---- LAB colour transfer (the CLI's default colour correction, color_fix.py:249-470): the two colour-space conversions.
// The histogram matching between them (color_fix.py:473-513) is a global sort + scatter per channel and stays on torch's
// sort (comfyui-seedvr2_b200/postprocess.py).  Arithmetic follows the reference's fp32 tensor ops one by one (separately
// rounded, powf as ATen's pow kernel calls it, x*x*x for the cube as ATen special-cases exponent 3); the 3x3 colour matrix is a
// library matmul in the reference, whose summation order is unspecified — the products are summed in index order here.
__device__ __forceinline__ float srgb_to_linear(float v) {
    return v > 0.04045f ? powf(__fdiv_rn(__fadd_rn(v, 0.055f), 1.055f), 2.4f) : __fdiv_rn(v, 12.92f);
}
__device__ __forceinline__ float dot3(float a0, float a1, float a2, float m0, float m1, float m2) {
    return fmaf(a2, m2, fmaf(a1, m1, __fmul_rn(a0, m0)));
}
__device__ __forceinline__ float lab_f(float v) {
    const float eps3 = (float)((6.0 / 29.0) * (6.0 / 29.0) * (6.0 / 29.0));
    const float kappa = (float)((29.0 / 3.0) * (29.0 / 3.0) * (29.0 / 3.0));
    return v > eps3 ? powf(v, (float)(1.0 / 3.0)) : __fdiv_rn(__fadd_rn(__fmul_rn(v, kappa), 16.0f), 116.0f);
}

// lab[t, {L,a,b}, y, x] (fp32) from rgb[t, {r,g,b}, y, x] in [-1, 1]; with `add` != NULL the input is clamp(rnd(rgb + add), -1, 1)
// (wavelet_reconstruction's combine, so that the LAB path does not need the reconstructed image in memory)
template <bool BF16>
__global__ void __launch_bounds__(256) rgb_to_lab_kernel(const void* __restrict__ rgb, const void* __restrict__ add,
                                                         float* __restrict__ lab, int64_t plane, int64_t n_pix) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;       // pixel index over [T, H, W]
    if (i >= n_pix) return;
    const int64_t t = i / plane, p = i - t * plane, base = t * 3 * plane + p;
    float c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float v = ld<BF16>(rgb, base + k * plane);
        if (add != nullptr) v = fminf(fmaxf(rnd<BF16>(__fadd_rn(v, ld<BF16>(add, base + k * plane))), -1.f), 1.f);
        v = fminf(fmaxf(__fmul_rn(__fadd_rn(v, 1.0f), 0.5f), 0.f), 1.f);       // .add_(1).mul_(0.5).clamp_(0, 1) in fp32
        c[k] = srgb_to_linear(v);
    }
    float X = dot3(c[0], c[1], c[2], 0.4124564f, 0.3575761f, 0.1804375f);
    const float Y = dot3(c[0], c[1], c[2], 0.2126729f, 0.7151522f, 0.0721750f);
    float Z = dot3(c[0], c[1], c[2], 0.0193339f, 0.1191920f, 0.9503041f);
    X = __fdiv_rn(X, 0.95047f);
    Z = __fdiv_rn(Z, 1.08883f);
    const float fx = lab_f(X), fy = lab_f(Y), fz = lab_f(Z);
    lab[base] = __fsub_rn(__fmul_rn(fy, 116.0f), 16.0f);
    lab[base + plane] = __fmul_rn(__fsub_rn(fx, fy), 500.0f);
    lab[base + 2 * plane] = __fmul_rn(__fsub_rn(fy, fz), 200.0f);
}

__device__ __forceinline__ float lab_finv(float f) {
    const float eps = (float)(6.0 / 29.0);
    const float kappa = (float)((29.0 / 3.0) * (29.0 / 3.0) * (29.0 / 3.0));
    return f > eps ? __fmul_rn(__fmul_rn(f, f), f) : __fdiv_rn(__fsub_rn(__fmul_rn(f, 116.0f), 16.0f), kappa);
}
__device__ __forceinline__ float linear_to_srgb(float v) {
    return v > 0.0031308f ? __fsub_rn(__fmul_rn(powf(fmaxf(v, 0.f), (float)(1.0 / 2.4)), 1.055f), 0.055f) : __fmul_rn(v, 12.92f);
}

// lab (fp32, planar) -> rgb in [-1, 1] rounded to the pipeline dtype (`result.to(original_dtype)`), then the batch normalisation
// clamp(-1, 1) * 0.5 + 0.5 and the writer's conversion, as colorfix_finish_kernel
template <bool BF16>
__global__ void __launch_bounds__(256) lab_to_rgb_finish_kernel(const float* __restrict__ lab, void* __restrict__ out,
                                                                uint8_t* __restrict__ out_u8, int bgr, int64_t plane, int64_t n_pix) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pix) return;
    const int64_t t = i / plane, p = i - t * plane, base = t * 3 * plane + p;
    const float L = lab[base], a = lab[base + plane], b = lab[base + 2 * plane];
    const float fy = __fdiv_rn(__fadd_rn(L, 16.0f), 116.0f);
    const float fx = __fadd_rn(__fdiv_rn(a, 500.0f), fy);
    const float fz = __fsub_rn(fy, __fdiv_rn(b, 200.0f));
    const float X = __fmul_rn(lab_finv(fx), 0.95047f), Y = lab_finv(fy), Z = __fmul_rn(lab_finv(fz), 1.08883f);
    float c[3];
    c[0] = dot3(X, Y, Z, 3.2404542f, -1.5371385f, -0.4985314f);
    c[1] = dot3(X, Y, Z, -0.9692660f, 1.8760108f, 0.0415560f);
    c[2] = dot3(X, Y, Z, 0.0556434f, -0.2040259f, 1.0572252f);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float v = fminf(fmaxf(linear_to_srgb(c[k]), 0.f), 1.f);
        v = rnd<BF16>(__fsub_rn(__fmul_rn(v, 2.0f), 1.0f));                  // .mul_(2).sub_(1) in fp32, then back to the pipeline dtype
        v = fminf(fmaxf(v, -1.f), 1.f);
        v = rnd<BF16>(__fmul_rn(v, 0.5f));
        v = rnd<BF16>(__fadd_rn(v, 0.5f));
        const int64_t o = i * 3 + k;                                          // [T, H, W, C]
        if (out != nullptr) st<BF16>(out, o, v);
        if (out_u8 != nullptr) out_u8[i * 3 + (bgr ? 2 - k : k)] = (uint8_t)(int)__fmul_rn(v, 255.0f);
    }
}

}  // namespace post
}  // namespace svr2

extern "C" int svr2_wavelet_level_fwd(const void* src, int64_t stride_t, int64_t stride_c, int64_t stride_y, int64_t stride_x,
                                      void* low, void* high, int first, int T, int C, int H, int W, int radius, int dtype,
                                      void* stream) {
    SVR2_CHECK_ARG(src && low, "svr2_wavelet_level_fwd: null pointer");
    SVR2_CHECK_ARG(T > 0 && C > 0 && H > 0 && W > 0 && (int64_t)T * C <= 65535 && H <= 65535, "svr2_wavelet_level_fwd: bad sizes %d x %d x %d x %d", T, C, H, W);
    SVR2_CHECK_ARG(radius > 0, "svr2_wavelet_level_fwd: radius must be positive");
    SVR2_CHECK_ARG(dtype == SVR2_DTYPE_F32 || dtype == SVR2_DTYPE_BF16, "svr2_wavelet_level_fwd: dtype must be SVR2_DTYPE_F32 or SVR2_DTYPE_BF16");
    const svr2::post::Strides ss{stride_t, stride_c, stride_y, stride_x};
    dim3 grid((W + 255) / 256, H, T * C);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SVR2_DTYPE_BF16) svr2::post::wavelet_level_kernel<true><<<grid, 256, 0, st>>>(src, ss, low, high, first, C, H, W, radius);
    else svr2::post::wavelet_level_kernel<false><<<grid, 256, 0, st>>>(src, ss, low, high, first, C, H, W, radius);
    SVR2_CHECK_LAUNCH("svr2_wavelet_level_fwd");
    return SVR2_OK;
}

extern "C" int svr2_colorfix_finish_fwd(const void* high, const void* low_style, void* out, void* out_u8, int bgr, int T, int C,
                                        int H, int W, int dtype, void* stream) {
    SVR2_CHECK_ARG(high && low_style && (out || out_u8), "svr2_colorfix_finish_fwd: null pointer");
    SVR2_CHECK_ARG(T > 0 && C > 0 && H > 0 && W > 0, "svr2_colorfix_finish_fwd: bad sizes");
    SVR2_CHECK_ARG(dtype == SVR2_DTYPE_F32 || dtype == SVR2_DTYPE_BF16, "svr2_colorfix_finish_fwd: dtype must be SVR2_DTYPE_F32 or SVR2_DTYPE_BF16");
    const int64_t n = (int64_t)T * C * H * W;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SVR2_DTYPE_BF16)
        svr2::post::colorfix_finish_kernel<true><<<blocks, 256, 0, st>>>(high, low_style, out, (uint8_t*)out_u8, bgr, C, H, W, n);
    else
        svr2::post::colorfix_finish_kernel<false><<<blocks, 256, 0, st>>>(high, low_style, out, (uint8_t*)out_u8, bgr, C, H, W, n);
    SVR2_CHECK_LAUNCH("svr2_colorfix_finish_fwd");
    return SVR2_OK;
}

extern "C" int svr2_rgb_to_lab_fwd(const void* rgb, const void* add, float* lab, int T, int H, int W, int dtype, void* stream) {
    SVR2_CHECK_ARG(rgb && lab && T > 0 && H > 0 && W > 0, "svr2_rgb_to_lab_fwd: bad arguments");
    SVR2_CHECK_ARG(dtype == SVR2_DTYPE_F32 || dtype == SVR2_DTYPE_BF16, "svr2_rgb_to_lab_fwd: dtype must be SVR2_DTYPE_F32 or SVR2_DTYPE_BF16");
    const int64_t plane = (int64_t)H * W, n = plane * T;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SVR2_DTYPE_BF16) svr2::post::rgb_to_lab_kernel<true><<<blocks, 256, 0, st>>>(rgb, add, lab, plane, n);
    else svr2::post::rgb_to_lab_kernel<false><<<blocks, 256, 0, st>>>(rgb, add, lab, plane, n);
    SVR2_CHECK_LAUNCH("svr2_rgb_to_lab_fwd");
    return SVR2_OK;
}

extern "C" int svr2_lab_to_rgb_finish_fwd(const float* lab, void* out, void* out_u8, int bgr, int T, int H, int W, int dtype,
                                          void* stream) {
    SVR2_CHECK_ARG(lab && (out || out_u8) && T > 0 && H > 0 && W > 0, "svr2_lab_to_rgb_finish_fwd: bad arguments");
    SVR2_CHECK_ARG(dtype == SVR2_DTYPE_F32 || dtype == SVR2_DTYPE_BF16, "svr2_lab_to_rgb_finish_fwd: dtype must be SVR2_DTYPE_F32 or SVR2_DTYPE_BF16");
    const int64_t plane = (int64_t)H * W, n = plane * T;
    const unsigned blocks = (unsigned)((n + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SVR2_DTYPE_BF16)
        svr2::post::lab_to_rgb_finish_kernel<true><<<blocks, 256, 0, st>>>(lab, out, (uint8_t*)out_u8, bgr, plane, n);
    else
        svr2::post::lab_to_rgb_finish_kernel<false><<<blocks, 256, 0, st>>>(lab, out, (uint8_t*)out_u8, bgr, plane, n);
    SVR2_CHECK_LAUNCH("svr2_lab_to_rgb_finish_fwd");
    return SVR2_OK;
}
