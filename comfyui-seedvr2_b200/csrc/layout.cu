// svr2_transpose_2b — layout changes at the boundary between this library's channels-last VAE modules and the reference's NCDHW
// modules: [C, T*H*W] <-> [T*H*W, C] for 2-byte elements, batched, with row pitches (so that a channel slice of
// a wider tensor can be the source: conv_out's 3 real channels of its 128-wide padded result).
//
// The reference never does this explicitly — it stays NCDHW and pays for it inside cuDNN — but every hand-over between a swapped
// module and an unswapped one needs it, and ATen's generic strided copy (`.contiguous()` of a permuted view) moves such a tensor at
// ~400 GB/s (measured: 10.5 ms per 4.2-GB copy, 52 % of a decoder pass, run 38).  HBM-bound: 64 x 64 tiles through shared memory, 16-byte
// global accesses on both sides; a scalar path covers unaligned shapes.
#include "common.cuh"

namespace svr2 {
namespace layout {

constexpr int TILE = 64;
constexpr int PITCH = 66;      // halfwords per transposed tile row: 33 words -> column gathers spread over the banks

// dst[b][j][i] = src[b][i][j], i < rows, j < cols.  VEC: rows % 8 == 0, cols % 8 == 0, pitches % 8 == 0, 16-byte aligned bases.
template <bool VEC>
__global__ void __launch_bounds__(256) transpose_kernel(const uint16_t* __restrict__ src, uint16_t* __restrict__ dst, int64_t rows, int64_t cols,
                                                        int64_t ld_src, int64_t ld_dst, int64_t bs_src, int64_t bs_dst, int64_t tiles_j) {
    __shared__ uint16_t t[TILE * PITCH];      // t[col][row]
    const int64_t tile = blockIdx.x;
    const int64_t i0 = (tile / tiles_j) * TILE, j0 = (tile % tiles_j) * TILE;
    src += (int64_t)blockIdx.y * bs_src;
    dst += (int64_t)blockIdx.y * bs_dst;
    if (VEC) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int idx = threadIdx.x + k * 256;
            const int r = idx >> 3, v = idx & 7;
            const int64_t i = i0 + r, j = j0 + 8 * v;
            if (i < rows && j < cols) {
                const uint4 q = __ldg(reinterpret_cast<const uint4*>(src + i * ld_src + j));
                const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    t[(8 * v + 2 * e) * PITCH + r] = (uint16_t)(w[e] & 0xffffu);
                    t[(8 * v + 2 * e + 1) * PITCH + r] = (uint16_t)(w[e] >> 16);
                }
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int idx = threadIdx.x + k * 256;
            const int c = idx >> 3, v = idx & 7;
            const int64_t j = j0 + c, i = i0 + 8 * v;
            if (j < cols && i < rows) {
                const uint32_t* p = reinterpret_cast<const uint32_t*>(t + c * PITCH + 8 * v);      // 4-byte aligned: PITCH and 8 v even
                *reinterpret_cast<uint4*>(dst + j * ld_dst + i) = make_uint4(p[0], p[1], p[2], p[3]);
            }
        }
    } else {
        for (int idx = threadIdx.x; idx < TILE * TILE; idx += 256) {
            const int r = idx >> 6, c = idx & 63;
            if (i0 + r < rows && j0 + c < cols) t[c * PITCH + r] = src[(i0 + r) * ld_src + j0 + c];
        }
        __syncthreads();
        for (int idx = threadIdx.x; idx < TILE * TILE; idx += 256) {
            const int c = idx >> 6, r = idx & 63;
            if (i0 + r < rows && j0 + c < cols) dst[(j0 + c) * ld_dst + i0 + r] = t[c * PITCH + r];
        }
    }
}

// Pixel (and time) shuffle of the up-samplers, channels-last: out[fo, h*S + x, w*S + y, c] = in[f, h, w, ((x*S + y)*Z + z)*C + c]
// with fo = f*Z + z, minus the duplicated second frame of the first temporal slice when `drop` is set (remove_head).
// One block per output row (fo, ho): everything but the column is a per-block constant, a thread keeps its 16-byte channel vector
// and walks the row's columns — no division in the loop (the first version, one flat index per vector with 64-bit div / mod, ran at
// 30 % of the HBM roofline).  Both sides move whole C-channel runs (>= 128 B).
__global__ void __launch_bounds__(256) pixel_shuffle_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, int H, int W, int C8,
                                                            int S, int Z, int drop) {
    const int ho = blockIdx.x, Wo = W * S;
    int fo = blockIdx.y;
    const int64_t out_row = ((int64_t)blockIdx.y * gridDim.x + ho) * Wo * C8;
    if (drop && fo >= 1) fo += 1;                          // output frames 1.. come from shuffled frames 2..
    const int f = fo / Z, z = fo - f * Z;
    const int h = ho / S, x = ho - h * S;
    const int R = S * S * Z;                               // channel planes of the expanded tensor
    const int64_t in_row = ((int64_t)f * H + h) * W * R * C8 + (int64_t)(x * S * Z + z) * C8;
    const int per = blockDim.x / C8 > 0 ? blockDim.x / C8 : 1;      // columns in flight (C8 <= 256 divides 256 for the VAE's widths)
    if (C8 <= (int)blockDim.x && (int)blockDim.x % C8 == 0) {
        const int c = threadIdx.x % C8;
        for (int wo = threadIdx.x / C8; wo < Wo; wo += per) {
            const int w = wo / S, y = wo - w * S;          // S is 1 or 2: a shift
            dst[out_row + (int64_t)wo * C8 + c] = __ldg(src + in_row + ((int64_t)w * R + y * Z) * C8 + c);
        }
    } else {
        for (int i = threadIdx.x; i < Wo * C8; i += blockDim.x) {
            const int wo = i / C8, c = i - wo * C8;
            const int w = wo / S, y = wo - w * S;
            dst[out_row + i] = __ldg(src + in_row + ((int64_t)w * R + y * Z) * C8 + c);
        }
    }
}

}  // namespace layout
}  // namespace svr2

extern "C" int svr2_pixel_shuffle3d_fwd(const void* src, void* dst, int T, int H, int W, int C, int spatial, int temporal,
                                        int drop_second_frame, void* stream) {
    SVR2_CHECK_ARG(src && dst, "svr2_pixel_shuffle3d_fwd: null pointer");
    SVR2_CHECK_ARG(T >= 0 && H > 0 && W > 0 && C > 0 && C % 8 == 0, "svr2_pixel_shuffle3d_fwd: bad shape (C %% 8 == 0)");
    SVR2_CHECK_ARG((spatial == 1 || spatial == 2) && (temporal == 1 || temporal == 2), "svr2_pixel_shuffle3d_fwd: ratios are 1 or 2");
    SVR2_CHECK_ARG(drop_second_frame == 0 || (drop_second_frame == 1 && temporal == 2 && T >= 1), "svr2_pixel_shuffle3d_fwd: drop needs temporal 2");
    const int64_t To = (int64_t)T * temporal - drop_second_frame;
    if (To <= 0) return SVR2_OK;
    SVR2_CHECK_ARG(To < 65536, "svr2_pixel_shuffle3d_fwd: %lld output frames in one call", (long long)To);
    const dim3 grid((unsigned)(H * spatial), (unsigned)To);
    svr2::layout::pixel_shuffle_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const uint4*)src, (uint4*)dst, H, W, C / 8, spatial, temporal,
                                                                              drop_second_frame);
    SVR2_CHECK_LAUNCH("svr2_pixel_shuffle3d_fwd");
    return SVR2_OK;
}

extern "C" int svr2_transpose_2b(const void* src, void* dst, int64_t rows, int64_t cols, int64_t ld_src, int64_t ld_dst, int batch,
                                 int64_t batch_stride_src, int64_t batch_stride_dst, void* stream) {
    using namespace svr2::layout;
    SVR2_CHECK_ARG(src && dst, "svr2_transpose_2b: null pointer");
    SVR2_CHECK_ARG(rows >= 0 && cols >= 0 && ld_src >= cols && ld_dst >= rows && batch >= 0, "svr2_transpose_2b: bad extents");
    if (rows == 0 || cols == 0 || batch == 0) return SVR2_OK;
    const int64_t ti = (rows + TILE - 1) / TILE, tj = (cols + TILE - 1) / TILE;
    SVR2_CHECK_ARG(ti * tj < ((int64_t)1 << 31) && batch < 65536, "svr2_transpose_2b: too many tiles");
    const bool vec = rows % 8 == 0 && cols % 8 == 0 && ld_src % 8 == 0 && ld_dst % 8 == 0 && batch_stride_src % 8 == 0 &&
                     batch_stride_dst % 8 == 0 && ((uintptr_t)src & 15) == 0 && ((uintptr_t)dst & 15) == 0;
    const dim3 grid((unsigned)(ti * tj), (unsigned)batch);
    if (vec)
        transpose_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>((const uint16_t*)src, (uint16_t*)dst, rows, cols, ld_src, ld_dst,
                                                                        batch_stride_src, batch_stride_dst, tj);
    else
        transpose_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>((const uint16_t*)src, (uint16_t*)dst, rows, cols, ld_src, ld_dst,
                                                                         batch_stride_src, batch_stride_dst, tj);
    SVR2_CHECK_LAUNCH("svr2_transpose_2b");
    return SVR2_OK;
}
