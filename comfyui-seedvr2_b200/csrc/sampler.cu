// The two elementwise steps either side of the DiT forward in the reference's one-step super-resolution call.
//
// svr2_sr_input_fwd  — builds the DiT input [N, 2C+1] = [ x_t | blurred latent | 1 ] in one pass.  Replaces get_condition
//   (task "sr", reference src/core/infer.py:54-78: zeros + two slice assignments), the optional latent-noise augmentation
//   schedule.forward(latent, aug_noise, t) (src/core/generation_phases.py:685-693, schedules/base.py:82-87) and
//   torch.cat([x_t, latents_cond], -1) (src/core/infer.py:341).  With augmentation the reference's timestep is fp32 after
//   timestep_transform, so A x_0 + B x_T is evaluated in fp32 (each op rounded) and rounded to bf16 once, by the assignment
//   into the bf16 condition — reproduced here with _rn intrinsics (no FMA contraction).
// svr2_v_lerp_endpoint_fwd — x_0 = (x_t - B pred) / (A + B), the Euler sampler's endpoint for prediction type v_lerp
//   (samplers/base.py:88-98, schedules/base.py:107-109) with the bf16 sampler state of the GPU path: every op rounds to bf16.
//   For the one-step schedule t = T: A = 0, B = 1 and the result is bf16(x_t - pred).
// Both are HBM-trivial (tens of MB against ~100 GB per forward); they exist so the whole call is stream-ordered device work.
#include "common.cuh"

namespace svr2 {

template <int MODE>   // 0: blur = latent; 1: blur = bf16(fl(fl(A latent) + fl(B aug)))
__global__ void __launch_bounds__(256) sr_input_kernel(const __nv_bfloat16* __restrict__ noise, const __nv_bfloat16* __restrict__ latent,
                                                       const __nv_bfloat16* __restrict__ aug, float A, float B,
                                                       __nv_bfloat16* __restrict__ out, int64_t total, int C) {
    const int W = 2 * C + 1;
    const __nv_bfloat16 one = __float2bfloat16_rn(1.0f);
    for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i0 < total; i0 += (int64_t)gridDim.x * blockDim.x * 8) {
        int64_t n = i0 / W;
        int c = (int)(i0 - n * W);
        __align__(16) __nv_bfloat16 v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            __nv_bfloat16 r = one;
            if (i0 + e < total) {
                if (c < C) {
                    r = noise[n * C + c];
                } else if (c < 2 * C) {
                    const int64_t s = n * C + (c - C);
                    if (MODE == 0) {
                        r = latent[s];
                    } else {
                        const float a = __fmul_rn(A, __bfloat162float(latent[s]));
                        const float b = __fmul_rn(B, __bfloat162float(aug[s]));
                        r = __float2bfloat16_rn(__fadd_rn(a, b));
                    }
                }
            }
            v[e] = r;
            if (++c == W) { c = 0; ++n; }
        }
        if (i0 + 8 <= total) {
            *reinterpret_cast<uint4*>(out + i0) = *reinterpret_cast<const uint4*>(v);
        } else {
            for (int e = 0; i0 + e < total; ++e) out[i0 + e] = v[e];
        }
    }
}

__global__ void __launch_bounds__(256) v_lerp_endpoint_kernel(const __nv_bfloat16* __restrict__ x_t, const __nv_bfloat16* __restrict__ pred,
                                                              float B, float denom, __nv_bfloat16* __restrict__ out, int64_t n) {
    for (int64_t i0 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i0 < n; i0 += (int64_t)gridDim.x * blockDim.x * 8) {
        if (i0 + 8 <= n) {
            float x[8], p[8], r[8];
            unpack8(ldg8_stream(x_t + i0), x);
            unpack8(ldg8_stream(pred + i0), p);
#pragma unroll
            for (int e = 0; e < 8; ++e) r[e] = bf16r(__fdiv_rn(bf16r(__fsub_rn(x[e], bf16r(__fmul_rn(B, p[e])))), denom));
            stg8_stream(out + i0, pack8(r));
        } else {
            for (int64_t i = i0; i < n; ++i) {
                const float x = __bfloat162float(x_t[i]), p = __bfloat162float(pred[i]);
                out[i] = __float2bfloat16_rn(__fdiv_rn(bf16r(__fsub_rn(x, bf16r(__fmul_rn(B, p)))), denom));
            }
        }
    }
}

static inline unsigned grid_for(int64_t elems) {
    int64_t g = (elems / 8 + 255) / 256 + 1;
    const int64_t cap = (int64_t)sm_count() * 8;
    return (unsigned)(g > cap ? cap : g);
}

}  // namespace svr2

extern "C" int svr2_sr_input_fwd(const void* noise, const void* latent, const void* aug_noise, float A, float B, int augment,
                                 void* out, int64_t n_rows, int channels, void* stream) {
    SVR2_CHECK_ARG(noise && latent && out, "svr2_sr_input_fwd: null pointer");
    SVR2_CHECK_ARG(n_rows > 0 && channels > 0 && channels <= 4096, "svr2_sr_input_fwd: bad shape (%lld rows, %d channels)",
                   (long long)n_rows, channels);
    SVR2_CHECK_ARG(!augment || aug_noise, "svr2_sr_input_fwd: augmentation asked for without aug_noise");
    const int64_t total = n_rows * (2 * (int64_t)channels + 1);
    const unsigned g = svr2::grid_for(total);
    using bf = __nv_bfloat16;
    if (augment)
        svr2::sr_input_kernel<1><<<g, 256, 0, (cudaStream_t)stream>>>((const bf*)noise, (const bf*)latent, (const bf*)aug_noise, A, B,
                                                                      (bf*)out, total, channels);
    else
        svr2::sr_input_kernel<0><<<g, 256, 0, (cudaStream_t)stream>>>((const bf*)noise, (const bf*)latent, nullptr, 1.f, 0.f,
                                                                      (bf*)out, total, channels);
    SVR2_CHECK_LAUNCH("svr2_sr_input_fwd");
    return SVR2_OK;
}

extern "C" int svr2_v_lerp_endpoint_fwd(const void* x_t, const void* pred, float A, float B, void* out, int64_t n, void* stream) {
    SVR2_CHECK_ARG(x_t && pred && out && n > 0, "svr2_v_lerp_endpoint_fwd: null pointer or empty input");
    // A and B arrive as the bf16 values of the sampler state; their sum rounds to bf16 like the reference's (A_t + B_t)
    const float denom = __bfloat162float(__float2bfloat16_rn(A + B));
    SVR2_CHECK_ARG(denom != 0.f, "svr2_v_lerp_endpoint_fwd: A + B == 0");
    using bf = __nv_bfloat16;
    svr2::v_lerp_endpoint_kernel<<<svr2::grid_for(n), 256, 0, (cudaStream_t)stream>>>((const bf*)x_t, (const bf*)pred, B, denom,
                                                                                     (bf*)out, n);
    SVR2_CHECK_LAUNCH("svr2_v_lerp_endpoint_fwd");
    return SVR2_OK;
}
