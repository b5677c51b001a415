// svr2_row_softmax_fwd — in-place row softmax of a padded 2-byte matrix, for the VAE mid block's attention.
//
// The mid block of the VAE's encoder / decoder holds one attention layer (diffusers `Attention`, built by UNetMidBlock3D,
// attn_video_vae.py:563-668): one head of 512 channels over the H x W positions of a latent frame (32,400 tokens at 1080p), which
// the reference hands to torch's scaled_dot_product_attention (head dim 512: only the memory-efficient backend takes it,
// 12.2 ms per frame on a B200).  Here S = Q K^T and O = P V are two calls of the convolution kernel as a GEMM (conv3d.cu,
// 1 x 1 x 1) and this kernel is the softmax between them: one CTA per row, the row (<= 32 values per thread) stays in registers.
// Columns n_valid .. pitch - 1 (the GEMM's channel padding) are written as zeros so that P can be the next GEMM's K operand.
// HBM-bound: reads and writes the matrix once.
#include <cuda_fp16.h>

#include "common.cuh"

namespace svr2 {
namespace softmax {

constexpr int kThreads = 1024;
constexpr int kVecs = 4;            // 16-byte vectors per thread: rows of up to 1024 * 4 * 8 = 32,768 columns (32 values in registers)

template <bool F16>
__device__ __forceinline__ void unpack8(const uint4& u, float* v) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        float2 f;
        if (F16) f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
        else f = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w[i]));
        v[2 * i] = f.x;
        v[2 * i + 1] = f.y;
    }
}
template <bool F16>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
    if (F16) {
        const __half2 h = __floats2half2_rn(a, b);
        return *reinterpret_cast<const uint32_t*>(&h);
    }
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<const uint32_t*>(&h);
}

template <bool MAX>
__device__ __forceinline__ float block_reduce(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float w = __shfl_xor_sync(0xffffffffu, v, o);
        v = MAX ? fmaxf(v, w) : v + w;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();                     // `red` may still be read from the previous reduction
    if (lane == 0) red[warp] = v;
    __syncthreads();
    v = lane < kThreads / 32 ? red[lane] : (MAX ? -INFINITY : 0.f);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float w = __shfl_xor_sync(0xffffffffu, v, o);
        v = MAX ? fmaxf(v, w) : v + w;
    }
    return v;
}

template <bool F16>
__global__ void __launch_bounds__(kThreads) row_softmax_kernel(uint4* __restrict__ x, int n_valid, int pitch8) {
    __shared__ float red[kThreads / 32];
    uint4* row = x + (int64_t)blockIdx.x * pitch8;
    float v[kVecs][8];
    float m = -INFINITY;
#pragma unroll
    for (int k = 0; k < kVecs; ++k) {
        const int c8 = threadIdx.x + k * kThreads;
        if (c8 < pitch8) {
            unpack8<F16>(row[c8], v[k]);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                if (c8 * 8 + i >= n_valid) v[k][i] = -INFINITY;
                m = fmaxf(m, v[k][i]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) v[k][i] = -INFINITY;
        }
    }
    m = block_reduce<true>(m, red);
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kVecs; ++k)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            v[k][i] = __expf(v[k][i] - m);          // exp(-inf) = 0 for the padding
            s += v[k][i];
        }
    s = block_reduce<false>(s, red);
    const float inv = 1.0f / s;
#pragma unroll
    for (int k = 0; k < kVecs; ++k) {
        const int c8 = threadIdx.x + k * kThreads;
        if (c8 < pitch8) {
            uint4 o;
            o.x = pack2<F16>(v[k][0] * inv, v[k][1] * inv); o.y = pack2<F16>(v[k][2] * inv, v[k][3] * inv);
            o.z = pack2<F16>(v[k][4] * inv, v[k][5] * inv); o.w = pack2<F16>(v[k][6] * inv, v[k][7] * inv);
            row[c8] = o;
        }
    }
}

}  // namespace softmax
}  // namespace svr2

extern "C" int svr2_row_softmax_fwd(void* x, int64_t rows, int n_valid, int pitch, int dtype, void* stream) {
    using namespace svr2::softmax;
    const char* name = "svr2_row_softmax_fwd";
    SVR2_CHECK_ARG(x, "%s: null pointer", name);
    SVR2_CHECK_ARG(dtype == SVR2_DTYPE_F16 || dtype == SVR2_DTYPE_BF16, "%s: dtype %d (fp16 or bf16)", name, dtype);
    SVR2_CHECK_ARG(rows >= 0 && rows < ((int64_t)1 << 31) && n_valid > 0 && pitch >= n_valid && pitch % 8 == 0, "%s: bad extents", name);
    SVR2_CHECK_ARG(pitch <= kThreads * kVecs * 8, "%s: rows of at most %d columns (got %d)", name, kThreads * kVecs * 8, pitch);
    if (rows == 0) return SVR2_OK;
    if (dtype == SVR2_DTYPE_F16)
        row_softmax_kernel<true><<<(unsigned)rows, kThreads, 0, (cudaStream_t)stream>>>((uint4*)x, n_valid, pitch / 8);
    else
        row_softmax_kernel<false><<<(unsigned)rows, kThreads, 0, (cudaStream_t)stream>>>((uint4*)x, n_valid, pitch / 8);
    SVR2_CHECK_LAUNCH(name);
    return SVR2_OK;
}
