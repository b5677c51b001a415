// svr2_halo_blend — crossfade of the `overlap` frames shared by two neighbouring temporal segments.
// Replaces blend_overlapping_frames (reference src/core/generation_utils.py:284-312) as applied by the multi-GPU parent
// process on CPU tensors (inference_cli.py:1163-1198).  fp32, HBM-bound: 12 B per element (two reads, one write).
#include "common.cuh"

namespace svr2 {

__device__ __forceinline__ float prev_weight(int f, int overlap) {
    if (overlap >= 3) {
        // t = linspace(0,1,overlap)[f]; u = clamp((t - 1/3) / (2/3 - 1/3), 0, 1); w = 0.5 + 0.5 cos(pi u)
        const float step = 1.0f / (float)(overlap - 1);
        const float t = (f < overlap / 2) ? step * (float)f : 1.0f - step * (float)(overlap - 1 - f);
        const float bs = 1.0f / 3.0f, be = 2.0f / 3.0f;
        float u = (t - bs) / (be - bs);
        u = fminf(fmaxf(u, 0.f), 1.f);
        return 0.5f + 0.5f * cosf(3.14159265358979323846f * u);
    }
    // linspace(1, 0, overlap)
    if (overlap <= 1) return 1.0f;
    const float step = -1.0f / (float)(overlap - 1);
    return (f < overlap / 2) ? 1.0f + step * (float)f : 0.0f - step * (float)(overlap - 1 - f);
}

// one expression for every kernel that blends, so that they round identically
__device__ __forceinline__ float4 blend4(const float4 a, const float4 b, float wp, float wc) {
    float4 r;
    // separately rounded mul, mul, add like the reference's three fp32 tensor ops (no FMA contraction)
    r.x = __fadd_rn(__fmul_rn(a.x, wp), __fmul_rn(b.x, wc));
    r.y = __fadd_rn(__fmul_rn(a.y, wp), __fmul_rn(b.y, wc));
    r.z = __fadd_rn(__fmul_rn(a.z, wp), __fmul_rn(b.z, wc));
    r.w = __fadd_rn(__fmul_rn(a.w, wp), __fmul_rn(b.w, wc));
    return r;
}

// Halo exchange + crossfade + gather in ONE kernel over peer memory.  `mine` is this rank's segment, `prev_tail` the previous
// rank's last `head` frames (a peer mapping: NVLink loads), `out` this rank's slot in the ROOT's stitched video (a peer
// mapping for every rank but the root: NVLink stores).  Output frame f < head is the crossfade of prev_tail[f] and mine[f];
// output frame f >= head is mine[first_body + f - head].  blockIdx.y = output frame.
__global__ void __launch_bounds__(256) segment_stitch_peer_kernel(const float4* __restrict__ mine, const float4* __restrict__ prev_tail,
                                                                 float4* __restrict__ out, int head, int first_body,
                                                                 int64_t n4_per_frame) {
    const int f = blockIdx.y;
    float4* dst = out + (int64_t)f * n4_per_frame;
    if (f < head) {
        const float wp = prev_weight(f, head), wc = 1.0f - wp;
        const float4* a = prev_tail + (int64_t)f * n4_per_frame;
        const float4* b = mine + (int64_t)f * n4_per_frame;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4_per_frame; i += (int64_t)gridDim.x * blockDim.x)
            dst[i] = blend4(a[i], b[i], wp, wc);
    } else {
        const float4* b = mine + (int64_t)(first_body + f - head) * n4_per_frame;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4_per_frame; i += (int64_t)gridDim.x * blockDim.x)
            dst[i] = b[i];
    }
}

__global__ void __launch_bounds__(256) halo_blend_kernel(const float4* __restrict__ prev, const float4* __restrict__ cur,
                                                         float4* __restrict__ out, int overlap, int64_t n4_per_frame) {
    const int f = blockIdx.y;
    const float wp = prev_weight(f, overlap), wc = 1.0f - wp;
    const int64_t base = (int64_t)f * n4_per_frame;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4_per_frame; i += (int64_t)gridDim.x * blockDim.x) {
        out[base + i] = blend4(prev[base + i], cur[base + i], wp, wc);
    }
}

}  // namespace svr2

extern "C" int svr2_halo_blend(const float* prev_tail, const float* cur_head, float* out, int overlap,
                               int64_t frame_elems, void* stream) {
    SVR2_CHECK_ARG(prev_tail && cur_head && out, "svr2_halo_blend: null pointer");
    SVR2_CHECK_ARG(overlap > 0 && overlap <= 65535 && frame_elems > 0 && frame_elems % 4 == 0,
                   "svr2_halo_blend: need overlap > 0 and frame_elems %% 4 == 0 (got %d, %lld)", overlap, (long long)frame_elems);
    const int64_t n4 = frame_elems / 4;
    int64_t gx = (n4 + 255) / 256;
    const int64_t cap = (int64_t)svr2::sm_count() * 8;
    if (gx > cap) gx = cap;
    dim3 grid((unsigned)gx, overlap);
    svr2::halo_blend_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const float4*)prev_tail, (const float4*)cur_head,
                                                                    (float4*)out, overlap, n4);
    SVR2_CHECK_LAUNCH("svr2_halo_blend");
    return SVR2_OK;
}

extern "C" int svr2_segment_stitch_peer(const float* mine, const float* prev_tail, float* out, int head, int first_body,
                                        int n_body, int64_t frame_elems, void* stream) {
    SVR2_CHECK_ARG(mine && out, "svr2_segment_stitch_peer: null pointer");
    SVR2_CHECK_ARG(head >= 0 && first_body >= 0 && n_body >= 0 && head + n_body <= 65535 && frame_elems > 0 && frame_elems % 4 == 0,
                   "svr2_segment_stitch_peer: bad frame counts (%d head, %d body from %d) or frame_elems %% 4 != 0 (%lld)", head, n_body,
                   first_body, (long long)frame_elems);
    SVR2_CHECK_ARG(head == 0 || prev_tail, "svr2_segment_stitch_peer: head frames need the previous segment's tail");
    if (head + n_body == 0) return SVR2_OK;
    const int64_t n4 = frame_elems / 4;
    int64_t gx = (n4 + 255) / 256;
    const int64_t cap = (int64_t)svr2::sm_count() * 8;
    if (gx > cap) gx = cap;
    dim3 grid((unsigned)gx, head + n_body);
    svr2::segment_stitch_peer_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const float4*)mine, (const float4*)prev_tail, (float4*)out,
                                                                             head, first_body, n4);
    SVR2_CHECK_LAUNCH("svr2_segment_stitch_peer");
    return SVR2_OK;
}
