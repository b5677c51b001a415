// svr2_row_softmax_fwd — in-place row softmax of a padded 2-byte matrix, for the VAE mid block's attention.
//
// The mid block of the VAE's encoder / decoder holds one attention layer (diffusers `Attention`, built by UNetMidBlock3D,
// attn_video_vae.py:563-668): one head of 512 channels over the H x W positions of a latent frame (32,400 tokens at 1080p), which
// the reference hands to torch's scaled_dot_product_attention (head dim 512: only the memory-efficient backend takes it,
// 12.2 ms per frame on a B200).  Here S = Q K^T and O = P V are two calls of the convolution kernel as a GEMM (conv3d.cu,
// 1 x 1 x 1) and this kernel is the softmax between them: one CTA per row.
// Columns n_valid .. pitch - 1 (the GEMM's channel padding) are written as zeros so that P can be the next GEMM's K operand.
// HBM-bound: reads and writes the matrix once.
#include <cuda_fp16.h>

#include "common.cuh"

namespace svr2 {
namespace softmax {

constexpr int kThreads = 512;
constexpr int kMaxPitch = 1 << 20;   // any row length a 32-bit column index covers comfortably; S itself is the practical limit

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

// Three passes over the row (max, sum, normalise) straight from memory: the first read comes from HBM, the other two hit L2 (a
// row is 64 KB and its CTA lives for microseconds), so DRAM still sees one read and one write of the matrix.  Holding the row in
// registers instead (first version: 1024 threads, 51 registers) left ONE CTA per SM with nothing to overlap its two block
// reductions with — 2.9 TB/s; packed-register variants spilled.  Here a thread keeps a handful of registers and four CTAs share an SM.
template <bool F16>
__global__ void __launch_bounds__(kThreads) row_softmax_kernel(uint4* __restrict__ x, int n_valid, int pitch8) {
    __shared__ float red[kThreads / 32];
    uint4* row = x + (int64_t)blockIdx.x * pitch8;
    const int nv8 = (n_valid + 7) >> 3;                 // vectors that hold at least one valid column
    float m = -INFINITY;
#pragma unroll 4
    for (int c8 = threadIdx.x; c8 < nv8; c8 += kThreads) {
        float v[8];
        unpack8<F16>(row[c8], v);
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (c8 * 8 + i < n_valid) m = fmaxf(m, v[i]);
    }
    m = block_reduce<true>(m, red);
    float s = 0.f;
#pragma unroll 4
    for (int c8 = threadIdx.x; c8 < nv8; c8 += kThreads) {
        float v[8];
        unpack8<F16>(row[c8], v);
#pragma unroll
        for (int i = 0; i < 8; ++i)
            if (c8 * 8 + i < n_valid) s += __expf(v[i] - m);
    }
    s = block_reduce<false>(s, red);
    const float inv = 1.0f / s;
#pragma unroll 4
    for (int c8 = threadIdx.x; c8 < pitch8; c8 += kThreads) {
        float v[8];
        if (c8 < nv8) unpack8<F16>(row[c8], v);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = (c8 * 8 + i < n_valid) ? __expf(v[i] - m) * inv : 0.f;      // zeros in the padding
        uint4 o;
        o.x = pack2<F16>(v[0], v[1]); o.y = pack2<F16>(v[2], v[3]); o.z = pack2<F16>(v[4], v[5]); o.w = pack2<F16>(v[6], v[7]);
        row[c8] = o;
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
    SVR2_CHECK_ARG(pitch <= kMaxPitch, "%s: rows of at most %d columns (got %d)", name, kMaxPitch, pitch);
    if (rows == 0) return SVR2_OK;
    if (dtype == SVR2_DTYPE_F16)
        row_softmax_kernel<true><<<(unsigned)rows, kThreads, 0, (cudaStream_t)stream>>>((uint4*)x, n_valid, pitch / 8);
    else
        row_softmax_kernel<false><<<(unsigned)rows, kThreads, 0, (cudaStream_t)stream>>>((uint4*)x, n_valid, pitch / 8);
    SVR2_CHECK_LAUNCH(name);
    return SVR2_OK;
}
