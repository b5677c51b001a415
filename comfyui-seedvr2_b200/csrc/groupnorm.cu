// svr2_groupnorm_silu_fwd — the video VAE's per-frame GroupNorm followed by SiLU, channels-last.
//
// Reference: causal_norm_wrapper (src/models/video_vae_v3/modules/causal_inflation_lib.py:354-409) folds time into the batch
// ("b c t h w -> (b t) c h w", :368), so every frame is normalised on its own statistics — one (mean, rstd) per (frame, group)
// over (C / G) x H x W values — then ResnetBlock3D applies `nonlinearity` = SiLU (attn_video_vae.py:316-322, 348-356).  The
// channel-chunked path the reference takes above its memory limit (:371-395) computes the same function.
//
// Rounding points (fp16 / bf16 tensors): ATen's group_norm computes the moments in fp32 but returns mean and rstd in the tensor
// dtype, and normalises with those ROUNDED statistics (measured: its output is a full ulp from the fp32 result, not half):
// y = a * x + b with a = rstd * gamma, b = -a * mean + beta in fp32, rounded to the tensor dtype; SiLU then reads that tensor,
// computes x / (1 + exp(-x)) in fp32 and rounds again.  All three roundings are reproduced; the moments are Chan-merged
// (count, mean, M2) triples — the quantity ATen's Welford pass computes, in a different but fixed order (no atomics).
//
// HBM-bound: the statistics pass reads the tensor once, the apply pass reads it and writes the result: 3 x T x H x W x C x 2 B.
// Layout [T, H, W, C]: a thread owns one 16-byte vector (8 channels) of a position, so its groups are fixed for the whole pass.
#include <cuda_fp16.h>

#include "common.cuh"

namespace svr2 {
namespace gn {

constexpr int kThreads = 256;
constexpr int kChunks = SVR2_GN_CHUNKS;      // statistics partials per frame
constexpr int kUnroll = 4;                   // independent 16-byte loads in flight per thread
constexpr int64_t kLutMinElements = 32ll << 20;      // tensors from this size on take the table-driven activation

struct Moments {
    float n, mean, m2;
};
// Chan et al. pairwise merge of two (count, mean, M2) triples
__device__ __forceinline__ Moments merge(const Moments& a, const Moments& b) {
    if (b.n == 0.f) return a;
    if (a.n == 0.f) return b;
    Moments r;
    r.n = a.n + b.n;
    const float d = b.mean - a.mean;
    const float f = b.n / r.n;
    r.mean = a.mean + d * f;
    r.m2 = a.m2 + b.m2 + d * d * a.n * f;
    return r;
}

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
__device__ __forceinline__ float round_t(float x) {
    return F16 ? __half2float(__float2half_rn(x)) : __bfloat162float(__float2bfloat16_rn(x));
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

// SiLU whose ROUNDED result equals that of ATen's fp32 formula x / (1 + exp(-x)) (expf + IEEE division, ~25 instructions per
// value: the apply pass was issue-bound, not HBM-bound, with it).  Fast path: ex2.approx + rcp.approx (relative error well
// below 2^-19); its rounding to the tensor dtype can only differ from the exact value's when it lies within that error of a
// rounding midpoint — then, and for results in the dtype's subnormal range, the exact formula is evaluated (~1 % of values).
template <bool F16>
__device__ __forceinline__ float silu_rounding_safe(float x) {
    float e;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * -1.4426950408889634f));
    float r = __fdividef(x, 1.0f + e);
    constexpr int DROP = F16 ? 13 : 16;                              // fp32 mantissa bits the rounding drops
    const uint32_t u = __float_as_uint(r);
    const int dist = abs((int)(u & ((1u << DROP) - 1u)) - (1 << (DROP - 1)));
    const bool subnormal = (u & 0x7FFFFFFFu) < (F16 ? 0x38800000u : 0x00800000u);
    if (dist <= 32 || subnormal) r = x / (1.0f + expf(-x));
    return r;
}

// Pass 1.  grid (kChunks, T): block (chunk, t) reduces positions [chunk * per, (chunk + 1) * per) of frame t.
// CPG = channels per group; a thread's 8 channels hold NG = max(1, 8 / CPG) groups (CPG = 16: half a group).
template <bool F16, int CPG>
__global__ void __launch_bounds__(kThreads) stats_kernel(const uint4* __restrict__ x, float* __restrict__ partial, int64_t HW, int C, int G) {
    constexpr int NG = CPG >= 8 ? 1 : 8 / CPG;
    constexpr int PER = CPG >= 8 ? 8 : CPG;            // values of one vector that fall in one of its groups
    const int vpp = C >> 3;                            // vectors per position
    const int lanes = kThreads / vpp;                  // positions in flight per block iteration
    const int vec = threadIdx.x % vpp, plane = threadIdx.x / vpp;
    const int t = blockIdx.y;
    const int64_t per = (HW + kChunks - 1) / kChunks;
    const int64_t p0 = (int64_t)blockIdx.x * per, p1 = min(HW, p0 + per);
    const uint4* src = x + ((int64_t)t * HW) * vpp + vec;

    float s[NG], ss[NG], pivot[NG], cnt = 0.f;
#pragma unroll
    for (int g = 0; g < NG; ++g) s[g] = ss[g] = pivot[g] = 0.f;
    bool have_pivot = false;
    if (plane < lanes) {
        for (int64_t pos = p0 + plane; pos < p1; pos += (int64_t)lanes * kUnroll) {
            uint4 raw[kUnroll];
#pragma unroll
            for (int u = 0; u < kUnroll; ++u)                      // independent loads first: the pass is latency x bandwidth bound
                if (pos + (int64_t)u * lanes < p1) raw[u] = __ldg(src + (pos + (int64_t)u * lanes) * vpp);
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                if (pos + (int64_t)u * lanes >= p1) break;
                float v[8];
                unpack8<F16>(raw[u], v);
                if (!have_pivot) {                     // shifted sums: no cancellation when |mean| >> std
#pragma unroll
                    for (int g = 0; g < NG; ++g) pivot[g] = v[g * PER];
                    have_pivot = true;
                }
#pragma unroll
                for (int g = 0; g < NG; ++g) {
#pragma unroll
                    for (int i = 0; i < PER; ++i) {
                        const float d = v[g * PER + i] - pivot[g];
                        s[g] += d;
                        ss[g] = fmaf(d, d, ss[g]);
                    }
                }
                cnt += (float)PER;
            }
        }
    }
    // thread-local triples -> shared memory -> one thread per group merges its contributors in a fixed order
    __shared__ Moments sm[kThreads * NG];
#pragma unroll
    for (int g = 0; g < NG; ++g) {
        Moments m;
        m.n = cnt;
        const float ms = cnt > 0.f ? s[g] / cnt : 0.f;
        m.mean = pivot[g] + ms;
        m.m2 = fmaxf(ss[g] - s[g] * ms, 0.f);
        sm[threadIdx.x * NG + g] = m;
    }
    __syncthreads();
    if (threadIdx.x < G) {
        const int grp = threadIdx.x;
        Moments acc = {0.f, 0.f, 0.f};
        // contributors of group grp: vectors [grp * CPG / 8, (grp + 1) * CPG / 8) (CPG >= 8) or slot grp % NG of vector grp / NG
        const int v0 = CPG >= 8 ? grp * (CPG / 8) : grp / NG;
        const int nv = CPG >= 8 ? CPG / 8 : 1;
        const int slot = CPG >= 8 ? 0 : grp % NG;
        for (int pl = 0; pl < lanes; ++pl)
            for (int j = 0; j < nv; ++j) acc = merge(acc, sm[(pl * vpp + v0 + j) * NG + slot]);
        float* dst = partial + (((int64_t)t * kChunks + blockIdx.x) * G + grp) * 3;
        dst[0] = acc.n; dst[1] = acc.mean; dst[2] = acc.m2;
    }
}

// Pass 2.  grid (blocks, T).  y = act(round(a * x + b)), a = rstd * gamma, b = -a * mean + beta.
template <bool F16, int CPG>
__global__ void __launch_bounds__(kThreads) apply_kernel(const uint4* __restrict__ x, const float* __restrict__ partial,
                                                         const void* __restrict__ gamma, const void* __restrict__ beta,
                                                         uint4* __restrict__ y, int64_t HW, int C, int G, float eps, int silu, int n_partials) {
    __shared__ float s_mean[64], s_rstd[64];
    __shared__ Moments s_red[kThreads];
    const int t = blockIdx.y;
    {
        // the frame's kChunks partials -> (mean, rstd) per group: kThreads / G threads per group, then one; fixed order
        const int grp = threadIdx.x % G, sub = threadIdx.x / G, nsub = kThreads / G;
        Moments acc = {0.f, 0.f, 0.f};
        if (sub < nsub) {
            const float* src = partial + ((int64_t)t * n_partials * G + grp) * 3;
            for (int c = sub; c < n_partials; c += nsub) {
                const Moments m = {src[(int64_t)c * G * 3], src[(int64_t)c * G * 3 + 1], src[(int64_t)c * G * 3 + 2]};
                acc = merge(acc, m);
            }
            s_red[sub * G + grp] = acc;
        }
        __syncthreads();
        if (threadIdx.x < G) {
            acc = s_red[grp];
            for (int k = 1; k < nsub; ++k) acc = merge(acc, s_red[k * G + grp]);
            // nn.GroupNorm: biased variance; ATen returns mean and rstd as tensors of the INPUT dtype and normalises with those
            // rounded values (group_norm_kernel.cu: mean / rstd are `T`, only the arithmetic is fp32) — so does this kernel
            s_mean[grp] = round_t<F16>(acc.mean);
            s_rstd[grp] = round_t<F16>(rsqrtf(acc.m2 / acc.n + eps));
        }
    }
    __syncthreads();
    const int vpp = C >> 3;
    const int lanes = kThreads / vpp;
    const int vec = threadIdx.x % vpp, plane = threadIdx.x / vpp;
    if (plane >= lanes) return;
    const bool affine = gamma != nullptr || beta != nullptr;
    float a[8], b[8];
    {
        float gm[8], bt[8];
        if (gamma != nullptr) unpack8<F16>(__ldg(reinterpret_cast<const uint4*>(gamma) + vec), gm);
        if (beta != nullptr) unpack8<F16>(__ldg(reinterpret_cast<const uint4*>(beta) + vec), bt);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int grp = (vec * 8 + i) / CPG;
            a[i] = gamma != nullptr ? s_rstd[grp] * gm[i] : s_rstd[grp];
            b[i] = affine ? fmaf(-a[i], s_mean[grp], beta != nullptr ? bt[i] : 0.f) : s_mean[grp];
        }
    }
    const int64_t base = ((int64_t)t * HW) * vpp + vec;
    const int64_t step = (int64_t)gridDim.x * lanes;
    for (int64_t pos = (int64_t)blockIdx.x * lanes + plane; pos < HW; pos += step * kUnroll) {
        uint4 raw[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
            if (pos + u * step < HW) raw[u] = __ldg(x + base + (pos + u * step) * vpp);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            if (pos + u * step >= HW) break;
            float v[8];
            unpack8<F16>(raw[u], v);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                // ATen: a * x + b with the fused per-channel a, b when there is an affine, (x - mean) * rstd when there is none
                float r = round_t<F16>(affine ? fmaf(a[i], v[i], b[i]) : (v[i] - b[i]) * a[i]);
                if (silu) r = silu_rounding_safe<F16>(r);                       // ATen silu: fp32 opmath, rounded by the pack below
                v[i] = r;
            }
            uint4 o;
            o.x = pack2<F16>(v[0], v[1]); o.y = pack2<F16>(v[2], v[3]); o.z = pack2<F16>(v[4], v[5]); o.w = pack2<F16>(v[6], v[7]);
            y[base + (pos + u * step) * vpp] = o;
        }
    }
}

// ---- SiLU by table.  The activation's input is the normalisation's ROUNDED output, i.e. one of 65,536 bit patterns, so
// silu -> round is a pure function of 16 bits: a 128-KB table (per dtype, filled once per device with ATen's exact fp32 formula)
// held in shared memory turns ~25 issue slots per value into one conflict-prone but cheap LDS.  The arithmetic version above
// stays for small tensors, where loading 128 KB per CTA would cost more than it saves.  Measured (run 33): apply pass at
// 1080p x 128 channels 1.61 ms (issue-bound, 29.5 instructions per value) -> see DESIGN.md.
static __device__ uint16_t g_silu_table[2][65536];      // [0] fp16, [1] bf16

template <bool F16>
__global__ void silu_table_kernel(uint16_t* __restrict__ table) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 65536u) return;
    float x;
    if (F16) x = __half2float(__ushort_as_half((unsigned short)i));
    else x = __bfloat162float(__ushort_as_bfloat16((unsigned short)i));
    const float r = x / (1.0f + expf(-x));
    table[i] = F16 ? __half_as_ushort(__float2half_rn(r)) : __bfloat16_as_ushort(__float2bfloat16_rn(r));
}

constexpr int kLutThreads = 1024;
constexpr int kLutSmem = 65536 * 2;

template <bool F16, int CPG>
__global__ void __launch_bounds__(kLutThreads, 1) apply_lut_kernel(const uint4* __restrict__ x, const float* __restrict__ partial,
                                                                   const void* __restrict__ gamma, const void* __restrict__ beta,
                                                                   uint4* __restrict__ y, const uint4* __restrict__ table, int64_t HW,
                                                                   int C, int G, float eps, int n_partials) {
    extern __shared__ __align__(16) unsigned char lut_raw[];
    uint16_t* lut = reinterpret_cast<uint16_t*>(lut_raw);
    __shared__ float s_mean[64], s_rstd[64];
    __shared__ Moments s_red[kLutThreads];
    for (int i = threadIdx.x; i < kLutSmem / 16; i += kLutThreads) reinterpret_cast<uint4*>(lut_raw)[i] = __ldg(table + i);
    const int t = blockIdx.y;
    {
        const int grp = threadIdx.x % G, sub = threadIdx.x / G, nsub = kLutThreads / G;
        Moments acc = {0.f, 0.f, 0.f};
        if (sub < nsub) {
            const float* src = partial + ((int64_t)t * n_partials * G + grp) * 3;
            for (int c = sub; c < n_partials; c += nsub) {
                const Moments m = {src[(int64_t)c * G * 3], src[(int64_t)c * G * 3 + 1], src[(int64_t)c * G * 3 + 2]};
                acc = merge(acc, m);
            }
            s_red[sub * G + grp] = acc;
        }
        __syncthreads();
        if (threadIdx.x < G) {
            acc = s_red[grp];
            for (int k = 1; k < nsub; ++k) acc = merge(acc, s_red[k * G + grp]);
            s_mean[grp] = round_t<F16>(acc.mean);                                  // see apply_kernel
            s_rstd[grp] = round_t<F16>(rsqrtf(acc.m2 / acc.n + eps));
        }
    }
    __syncthreads();
    const int vpp = C >> 3;
    const int lanes = kLutThreads / vpp;
    const int vec = threadIdx.x % vpp, plane = threadIdx.x / vpp;
    if (plane >= lanes) return;
    const bool affine = gamma != nullptr || beta != nullptr;
    float a[8], b[8];
    {
        float gm[8], bt[8];
        if (gamma != nullptr) unpack8<F16>(__ldg(reinterpret_cast<const uint4*>(gamma) + vec), gm);
        if (beta != nullptr) unpack8<F16>(__ldg(reinterpret_cast<const uint4*>(beta) + vec), bt);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int grp = (vec * 8 + i) / CPG;
            a[i] = gamma != nullptr ? s_rstd[grp] * gm[i] : s_rstd[grp];
            b[i] = affine ? fmaf(-a[i], s_mean[grp], beta != nullptr ? bt[i] : 0.f) : s_mean[grp];
        }
    }
    const int64_t base = ((int64_t)t * HW) * vpp + vec;
    const int64_t step = (int64_t)gridDim.x * lanes;
    for (int64_t pos = (int64_t)blockIdx.x * lanes + plane; pos < HW; pos += step * kUnroll) {
        uint4 raw[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u)
            if (pos + u * step < HW) raw[u] = __ldg(x + base + (pos + u * step) * vpp);
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            if (pos + u * step >= HW) break;
            float v[8];
            unpack8<F16>(raw[u], v);
            uint32_t o[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float r = affine ? fmaf(a[i], v[i], b[i]) : (v[i] - b[i]) * a[i];
                const uint32_t bits = F16 ? (uint32_t)__half_as_ushort(__float2half_rn(r)) : (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(r));
                o[i] = lut[bits];
            }
            uint4 w;
            w.x = o[0] | (o[1] << 16); w.y = o[2] | (o[3] << 16); w.z = o[4] | (o[5] << 16); w.w = o[6] | (o[7] << 16);
            y[base + (pos + u * step) * vpp] = w;
        }
    }
}

// device address of the dtype's table, filled on first use on this device (on `st`; inside a stream capture the fill is
// recorded with the capture every time, because a captured launch has not run yet when the next eager call looks at the flag)
template <bool F16>
int silu_table(const uint4** out, cudaStream_t st) {
    static bool filled[kMaxDevices] = {};
    static uint16_t* base[kMaxDevices] = {};
    const int dev = cur_device();
    if (base[dev] == nullptr) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_silu_table);
        if (e != cudaSuccess || p == nullptr) {
            set_error("svr2_groupnorm_silu_fwd: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
            return SVR2_E_CUDA;
        }
        base[dev] = (uint16_t*)p;
    }
    uint16_t* tab = base[dev] + (F16 ? 0 : 65536);
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cap);
    if (!filled[dev] || cap != cudaStreamCaptureStatusNone) {
        silu_table_kernel<F16><<<65536 / 256, 256, 0, st>>>(tab);
        SVR2_CHECK_LAUNCH("svr2_groupnorm_silu_fwd(table)");
        if (cap == cudaStreamCaptureStatusNone) {
            // other streams of this device may use the table right after this call returns
            if (cudaStreamSynchronize(st) != cudaSuccess) {
                set_error("svr2_groupnorm_silu_fwd: table fill failed");
                return SVR2_E_CUDA;
            }
            filled[dev] = true;
        }
    }
    *out = reinterpret_cast<const uint4*>(tab);
    return SVR2_OK;
}

template <bool F16, int CPG>
int launch(const void* x, const void* gamma, const void* beta, void* y, float* ws, int T, int64_t HW, int C, int G, float eps, int silu,
           int have_partials, cudaStream_t st) {
    const int n_partials = have_partials > 0 ? have_partials : kChunks;
    if (have_partials <= 0) {
        stats_kernel<F16, CPG><<<dim3(kChunks, T), kThreads, 0, st>>>((const uint4*)x, ws, HW, C, G);
        SVR2_CHECK_LAUNCH("svr2_groupnorm_silu_fwd(stats)");
    }
    if (silu && (int64_t)T * HW * C >= kLutMinElements) {
        const uint4* table = nullptr;
        int rc = silu_table<F16>(&table, st);
        if (rc != SVR2_OK) return rc;
        static PerDeviceFlag smem_ok;
        if ((rc = ensure_dyn_smem(apply_lut_kernel<F16, CPG>, kLutSmem, smem_ok, "svr2_groupnorm_silu_fwd")) != SVR2_OK) return rc;
        int per_frame = sm_count() / T;
        if (per_frame < 1) per_frame = 1;
        apply_lut_kernel<F16, CPG><<<dim3((unsigned)per_frame, T), kLutThreads, kLutSmem, st>>>((const uint4*)x, ws, gamma, beta, (uint4*)y,
                                                                                              table, HW, C, G, eps, n_partials);
        SVR2_CHECK_LAUNCH("svr2_groupnorm_silu_fwd(apply, table)");
        return SVR2_OK;
    }
    const int lanes = kThreads / (C >> 3);
    int64_t blocks = (HW + lanes * 8 - 1) / (lanes * 8);             // ~8 positions per thread
    const int cap = sm_count() * 16 / (T > 0 ? T : 1) + 1;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    apply_kernel<F16, CPG><<<dim3((unsigned)blocks, T), kThreads, 0, st>>>((const uint4*)x, ws, gamma, beta, (uint4*)y, HW, C, G, eps, silu, n_partials);
    SVR2_CHECK_LAUNCH("svr2_groupnorm_silu_fwd(apply)");
    return SVR2_OK;
}

template <bool F16>
int dispatch(int cpg, const void* x, const void* gamma, const void* beta, void* y, float* ws, int T, int64_t HW, int C, int G, float eps,
             int silu, int have_partials, cudaStream_t st) {
    switch (cpg) {
        case 2: return launch<F16, 2>(x, gamma, beta, y, ws, T, HW, C, G, eps, silu, have_partials, st);
        case 4: return launch<F16, 4>(x, gamma, beta, y, ws, T, HW, C, G, eps, silu, have_partials, st);
        case 8: return launch<F16, 8>(x, gamma, beta, y, ws, T, HW, C, G, eps, silu, have_partials, st);
        case 16: return launch<F16, 16>(x, gamma, beta, y, ws, T, HW, C, G, eps, silu, have_partials, st);
        default: return SVR2_E_INVALID;
    }
}

}  // namespace gn
}  // namespace svr2

extern "C" int64_t svr2_groupnorm_workspace_bytes(int T, int groups) {
    return (int64_t)(T > 0 ? T : 0) * SVR2_GN_CHUNKS * (groups > 0 ? groups : 0) * 3 * (int64_t)sizeof(float);
}

extern "C" int svr2_groupnorm_silu_fwd(const void* x, const void* weight, const void* bias, void* y, void* workspace, int have_partials,
                                       int T, int64_t HW, int C, int groups, float eps, int silu, int dtype, void* stream) {
    const char* name = "svr2_groupnorm_silu_fwd";
    SVR2_CHECK_ARG(x && y && workspace, "%s: null pointer", name);
    SVR2_CHECK_ARG(dtype == SVR2_DTYPE_F16 || dtype == SVR2_DTYPE_BF16, "%s: dtype %d (fp16 or bf16)", name, dtype);
    SVR2_CHECK_ARG(T >= 0 && HW > 0, "%s: bad extent T=%d HW=%lld", name, T, (long long)HW);
    SVR2_CHECK_ARG(groups > 0 && groups <= 64 && C > 0 && C % groups == 0, "%s: C=%d groups=%d", name, C, groups);
    SVR2_CHECK_ARG(have_partials >= 0 && (have_partials == 0 || groups == 32), "%s: statistics partials of a convolution are for 32 groups", name);
    const int cpg = C / groups;
    SVR2_CHECK_ARG(cpg == 2 || cpg == 4 || cpg == 8 || cpg == 16, "%s: %d channels per group (2, 4, 8 or 16)", name, cpg);
    SVR2_CHECK_ARG(C % 8 == 0 && C / 8 <= svr2::gn::kThreads && svr2::gn::kThreads % (C / 8) == 0,
                   "%s: C=%d must be 8 x a power of two <= 2048", name, C);
    if (T == 0) return SVR2_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (dtype == SVR2_DTYPE_F16)
        return svr2::gn::dispatch<true>(cpg, x, weight, bias, y, (float*)workspace, T, HW, C, groups, eps, silu, have_partials, st);
    return svr2::gn::dispatch<false>(cpg, x, weight, bias, y, (float*)workspace, T, HW, C, groups, eps, silu, have_partials, st);
}
