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
// One thread per 16-byte vector of the output; both sides move whole C-channel runs (>= 128 B).
__global__ void __launch_bounds__(256) pixel_shuffle_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, int T, int H, int W,
                                                            int C8, int S, int Z, int drop, int64_t n_out) {
    const int To = T * Z - drop, Ho = H * S, Wo = W * S;
    (void)To;
    for (int64_t o = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; o < n_out; o += (int64_t)gridDim.x * blockDim.x) {
        const int c = (int)(o % C8);
        int64_t r = o / C8;
        const int wo = (int)(r % Wo); r /= Wo;
        const int ho = (int)(r % Ho);
        int fo = (int)(r / Ho);
        if (drop && fo >= 1) fo += 1;                      // output frame 1.. come from shuffled frames 2..
        const int f = fo / Z, z = fo - f * Z;
        const int h = ho / S, x = ho - h * S, w = wo / S, y = wo - w * S;
        const int64_t s_idx = ((((int64_t)f * H + h) * W + w) * (S * S * Z) + ((x * S + y) * Z + z)) * C8 + c;
        dst[o] = __ldg(src + s_idx);
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
    const int64_t n_out = To * H * spatial * W * spatial * (C / 8);
    if (n_out <= 0) return SVR2_OK;
    int64_t blocks = (n_out + 255) / 256;
    const int64_t cap = (int64_t)svr2::sm_count() * 32;
    if (blocks > cap) blocks = cap;
    svr2::layout::pixel_shuffle_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const uint4*)src, (uint4*)dst, T, H, W, C / 8,
                                                                                          spatial, temporal, drop_second_frame, n_out);
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
