// svr2_causal_conv3d_fwd / svr2_causal_conv3d_down_fwd — the video VAE's causal 3-D convolutions as an implicit GEMM on tcgen05
// (SURVEY §8f row N3): the stride-1 convolutions of the residual blocks, up-samplers and model boundaries (~99 % of the VAE's
// FLOPs) and the encoder's stride-2 down-samplers.
//
// Reference: InflatedCausalConv3d (src/models/video_vae_v3/modules/causal_inflation_lib.py:61-262).  The temporal padding is
// removed from the module's padding (:74) and replaced by a causal head: 2 * temporal_padding copies of the first frame
// (extend_head, :422-437, basic_forward :228), or the last kernel_t - 1 frames of the previous temporal slice kept in
// `self.memory` (:226-227, slicing_forward :254-272).  Spatially the convolution zero-pads (Conv3d padding (0, 1, 1)).
//
//   out[t, h, w, co] = bias[co] + sum_{dt, dh, dw, ci} W[co, dt, dh, dw, ci] * X[t + dt - (kt - 1), h + dh - ks/2, w + dw - ks/2, ci]
//   X[t' < 0] = head[t' + (kt - 1)]   (memory of the previous slice)   or   X[0]   (first slice: replicated first frame)
//   (+ resid[t, h, w, co], rounded after the convolution's own rounding: ResnetBlock3D's input_tensor + hidden_states)
//
// Layout: activations channels-last [T, H, W, C] (torch.channels_last_3d of the reference's [1, C, T, H, W] — same bytes), weights
// pre-packed once to [C_out, kt, ks, ks, C_in] (svr2_conv3d_pack_weight), so that the GEMM's K axis is (tap, channel).
//
// The A operand is never materialised.  An output tile of one CTA is 16 rows x 8 columns of one frame.  For each (temporal tap,
// 64-channel chunk) ONE 4-D TMA box brings the tile's receptive field — 18 rows x 16 column slots x 64 channels (36 KB; the
// box starts one column left of the tile and the TMA unit zero-fills whatever lies outside the frame: the spatial zero padding)
// — into shared memory as consecutive 128-byte rows, 128-byte swizzled.  That is already a K-major UMMA operand for EVERY one
// of the 9 spatial taps: tap (dh, dw) is the same patch read from row slot (dh * 16 + dw) on, eight consecutive slots (the
// tile's 8 columns) per 8-row core group and a group stride of one patch row (2048 B).  The descriptor's start address is then
// only 128-byte aligned; that is fine because tcgen05.mma, like the TMA unit that wrote the patch, derives the 128-byte-swizzle
// XOR from the absolute shared-memory address (bits 7-9 into bits 4-6), not from the row's index in the tile — verified on the
// B200 (run 26: base-offset field left 0 gives exact results, setting it to the phase dw does not).  The nine taps therefore re-read
// shared memory, not L2: 36 KB of input per 9 K steps instead of 9 x 16 KB.  1 x 1 kernels use a 16 x 8 box (the tile itself).
// Stride-2 down-samplers (Downsample3D, attn_video_vae.py:176-246) run on the 1 x 1 instantiation: each of the 9 spatial taps is
// its own K-step group fed by a 16 x 8 box with TMA element strides 2 (every second row / column from the tap's offset on), the
// zero row / column the reference appends is the unit's zero fill, temporal stride 2 is a frame coordinate 2 t' + dt - head.
//
// The epilogue can also gather the (count, mean, M2) partials of the output per frame and GroupNorm group (gn_partials), so that
// the per-frame GroupNorm that follows every convolution of a residual block needs no statistics pass of its own.
//
// Kernel shape: the GEMM of gemm.cu.  CTA pair, tcgen05.mma.cta_group::2 with M = 256 output positions (16 x 16 of one frame;
// each CTA stages the patch of its 8 columns and half of the weight tile), N = 256 or 128 output channels, fp32 accumulators
// double-buffered in TMEM, persistent over (frame, tile, channel tile); warp 0 = TMA producer (a patch ring and a weight ring),
// warp 1 = MMA issuer, warps 4-7 = epilogue (bias, residual, rounding, swizzled slab, 4-D TMA store clipped at the frame's
// edge).  fp16 (the VAE's dtype) or bf16, fp32 accumulation.
#include <cuda.h>
#include <cuda_fp16.h>

#include "common.cuh"
#include "sm100_pair.cuh"
#include "sm100_ptx.cuh"
#include "tmap.cuh"

namespace svr2 {
namespace conv {

using namespace ptx;
using namespace pairx;

constexpr int BM = 128;                  // output positions per CTA: 16 rows x 8 columns of one frame
constexpr int TILE_H = 16, TILE_W = 8;   // per CTA; the pair covers 16 x 16 (CTA r: columns 8r .. 8r + 7)
constexpr int BK = 64;                   // channels per K step (one 128-byte swizzle atom)
constexpr int PATCH_BYTES = 18 * 16 * 128;   // 3 x 3 taps: 18 rows x 16 column slots (10 used) x 64 channels; 1 x 1: 16 x 8 of it
constexpr int SLAB_BYTES = 32 * 128;     // 4 tile rows x 8 columns x 64 channels: one TMA store box
constexpr uint32_t TMEM_COLS = 512;

template <int BN>
struct Cfg {
    static constexpr int B_BYTES = (BN / 2) * BK * 2;
    static constexpr int PSTAGES = 2;                     // patch ring: one stage feeds ks * ks K steps (1 x 1: 4 stages of 16 KB)
    static constexpr int PSTAGES_MAX = 4;
    // weight ring: one stage holds the weight boxes of TPB consecutive taps (a row of the 3 x 3 stencil when N = 128, where a K
    // step is only 256 tensor-pipe cycles: one barrier round trip per row instead of per tap, 1211 -> 1294 TFLOP/s, run 29)
    static constexpr int TPB = BN == 256 ? 1 : 3;
    static constexpr int BSTAGE_BYTES = TPB * B_BYTES;
    static constexpr int BSTAGES = BN == 256 ? 7 : 5;
    static constexpr int NTHREADS = 256;
    static constexpr int OFF_B = PSTAGES * PATCH_BYTES;
    static constexpr int OFF_SLAB = OFF_B + BSTAGES * BSTAGE_BYTES;
    static constexpr int OFF_BAR = OFF_SLAB + 4 * 2 * SLAB_BYTES;
    static constexpr int SMEM_BYTES = OFF_BAR + 512;
    static constexpr int BAR_PFULL = 0, BAR_PEMPTY = PSTAGES_MAX, BAR_BFULL = 2 * PSTAGES_MAX, BAR_BEMPTY = 2 * PSTAGES_MAX + BSTAGES,
                         BAR_TFULL = 2 * (PSTAGES_MAX + BSTAGES), BAR_TEMPTY = BAR_TFULL + 2, BAR_COUNT = BAR_TFULL + 4;
    static_assert(BAR_COUNT * 8 + 8 <= 512, "barrier block too small");
    static_assert(SMEM_BYTES <= 232448, "shared memory");
    static_assert(OFF_B % 1024 == 0 && OFF_SLAB % 1024 == 0 && PATCH_BYTES % 1024 == 0 && B_BYTES % 1024 == 0,
                  "swizzled tiles need 1024-byte alignment");
};

struct Params {
    int T, H, W, Cin, Cout;
    int kt, ks;                 // temporal taps (1 | 3), spatial taps per axis (1 | 3)
    int h_tiles, w_tiles, n_tiles;
    int has_head;               // frames t' < 0 come from the head tensor (head_len frames); otherwise frame 0 is replicated
    int head_len;               // frames in front of the slice: kt - 1 (stride 1), 2 or the memory's length (temporal stride 2)
    int st;                     // temporal stride (1 | 2)
    int s2;                     // spatial stride 2 (down-sampler): one strided TMA box per tap, KS = 1 instantiation, 9 x the K steps
    int Ti, Hi, Wi;             // input extents (host: tensor maps); T, H, W above are the OUTPUT's
    const void* bias;           // [Cout] or nullptr
    const void* resid;          // [T, H, W, Cout] or nullptr
    float* gn_partials;         // [T, gn_slots, 32, 3] (count, mean, M2) of the OUTPUT per frame and group of Cout / 32 channels, or nullptr
    int gn_slots;
};

__device__ __forceinline__ void tma_load_4d_pair(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(dst), "l"(m), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, uint32_t src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(m), "r"(src), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
// kind::f16 instruction descriptor, fp16 or bf16 operands (both K-major), fp32 accumulator
__host__ __device__ constexpr uint32_t idesc_16(bool f16, int M, int N) {
    return (1u << 4) | ((f16 ? 0u : 1u) << 7) | ((f16 ? 0u : 1u) << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

template <bool F16> struct Num;
template <> struct Num<true> {
    using T2 = __half2;
    static __device__ __forceinline__ uint32_t pack(float a, float b) { const __half2 v = __floats2half2_rn(a, b); return *reinterpret_cast<const uint32_t*>(&v); }
    static __device__ __forceinline__ float2 unpack(uint32_t u) { return __half22float2(*reinterpret_cast<const __half2*>(&u)); }
    static __device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) {
        const __half2 v = __hadd2_rn(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
        return *reinterpret_cast<const uint32_t*>(&v);
    }
};
template <> struct Num<false> {
    static __device__ __forceinline__ uint32_t pack(float a, float b) { const __nv_bfloat162 v = __floats2bfloat162_rn(a, b); return *reinterpret_cast<const uint32_t*>(&v); }
    static __device__ __forceinline__ float2 unpack(uint32_t u) { return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&u)); }
    static __device__ __forceinline__ uint32_t add(uint32_t a, uint32_t b) {
        const __nv_bfloat162 v = __hadd2_rn(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
        return *reinterpret_cast<const uint32_t*>(&v);
    }
};

struct Tile {
    int t, h0, w0, nt;
};
__device__ __forceinline__ Tile decode_tile(const Params& p, int id) {
    Tile r;
    r.nt = id % p.n_tiles;
    int sp = id / p.n_tiles;
    r.w0 = (sp % p.w_tiles) * (2 * TILE_W);
    sp /= p.w_tiles;
    r.h0 = (sp % p.h_tiles) * TILE_H;
    r.t = sp / p.h_tiles;
    return r;
}

// ---- GroupNorm statistics of the output, gathered by the epilogue (the next op of the VAE's residual blocks is a per-frame
// GroupNorm over 32 groups: its statistics pass — one full read of the tensor — disappears).  Per 64-channel slab a thread holds
// its position's rounded outputs; sums are taken against a per-group pivot (lane 0's first channel: no cancellation when
// |mean| >> std), reduced over the warp's 32 positions by a butterfly, and the lane whose index is the group's number merges the
// tile's (count, mean, M2) into its running triple (Chan).  One partial per (frame, epilogue warp): deterministic, no atomics.
struct Running {
    float n, mean, m2;
};
__device__ __forceinline__ void chan_merge(Running& a, float n, float mean, float m2) {
    if (n == 0.f) return;
    if (a.n == 0.f) { a.n = n; a.mean = mean; a.m2 = m2; return; }
    const float tot = a.n + n, d = mean - a.mean, f = n / tot;
    a.m2 = a.m2 + m2 + d * d * a.n * f;
    a.mean = a.mean + d * f;
    a.n = tot;
}
// QPG = channel quads per group (Cout / 32 / 4 = 1, 2, 4); o = the slab's 64 outputs of this thread's position, packed pairs
template <bool F16, int QPG>
__device__ __forceinline__ void slab_stats(const uint32_t* o, bool valid, int lane, int gbase, Running& run) {
    using N = Num<F16>;
    constexpr int GPS = 16 / QPG;                       // groups in the slab
    float S[GPS], Q[GPS], K[GPS];
#pragma unroll
    for (int g = 0; g < GPS; ++g) {
        K[g] = __shfl_sync(0xffffffffu, N::unpack(o[2 * g * QPG]).x, 0);
        float s = 0.f, q = 0.f;
#pragma unroll
        for (int j = 0; j < 2 * QPG; ++j) {
            const float2 v = N::unpack(o[2 * g * QPG + j]);
            const float d0 = v.x - K[g], d1 = v.y - K[g];
            s += d0 + d1;
            q = fmaf(d0, d0, fmaf(d1, d1, q));
        }
        S[g] = valid ? s : 0.f;
        Q[g] = valid ? q : 0.f;
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
        for (int g = 0; g < GPS; ++g) {
            S[g] += __shfl_xor_sync(0xffffffffu, S[g], off);
            Q[g] += __shfl_xor_sync(0xffffffffu, Q[g], off);
        }
    }
    const float n = (float)(__popc(__ballot_sync(0xffffffffu, valid)) * 4 * QPG);
    float s = 0.f, q = 0.f, k = 0.f;
    bool mine = false;
#pragma unroll
    for (int g = 0; g < GPS; ++g)
        if (lane == gbase + g) { s = S[g]; q = Q[g]; k = K[g]; mine = true; }
    if (mine && n > 0.f) {
        const float ms = s / n;
        chan_merge(run, n, k + ms, fmaxf(q - s * ms, 0.f));
    }
}

template <int BN, bool F16, int KS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(Cfg<BN>::NTHREADS, 1)
conv3d_tc_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_head,
                 const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_y, const Params p) {
    using C = Cfg<BN>;
    using N = Num<F16>;
    constexpr int BSTAGES = C::BSTAGES, B_BYTES = C::B_BYTES, BSTAGE_BYTES = C::BSTAGE_BYTES;
    constexpr int BAR_PFULL = C::BAR_PFULL, BAR_PEMPTY = C::BAR_PEMPTY, BAR_BFULL = C::BAR_BFULL, BAR_BEMPTY = C::BAR_BEMPTY,
                  BAR_TFULL = C::BAR_TFULL, BAR_TEMPTY = C::BAR_TEMPTY;
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t base = smem_u32(smem);
    const uint32_t sBar = base + C::OFF_BAR;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + C::OFF_BAR + C::BAR_COUNT * 8);
    auto bar = [&](int slot) { return sBar + 8u * (uint32_t)slot; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();
    const bool leader = rank == 0;
    const int first = (int)(blockIdx.x >> 1), stride = (int)(gridDim.x >> 1);
    const int n_tiles_total = p.T * p.h_tiles * p.w_tiles * p.n_tiles;
    const int c_chunks = p.Cin / BK;
    constexpr int half = KS >> 1;
    constexpr int taps_sp = KS * KS;                           // K steps fed by one patch
    constexpr int slot_w = KS == 3 ? 16 : 8;                   // column slots per patch row
    constexpr uint32_t patch_bytes = (uint32_t)((TILE_H + 2 * half) * slot_w * 128);
    const int n_patches = p.kt * c_chunks * ((KS == 1 && p.s2) ? 9 : 1);      // per tile
    constexpr int tpb = KS == 3 ? C::TPB : 1;                  // taps per weight stage (divides KS * KS)
    constexpr uint32_t pstages = KS == 3 ? C::PSTAGES : C::PSTAGES_MAX;
    constexpr uint32_t pslot = KS == 3 ? (uint32_t)PATCH_BYTES : 16384u;      // patch ring slot

    if (threadIdx.x == 0) {
        if (base & 1023u) __trap();
        for (int s = 0; s < C::PSTAGES_MAX; ++s) {
            mbar_init(bar(BAR_PFULL + s), 1);
            mbar_init(bar(BAR_PEMPTY + s), 1);
        }
        for (int s = 0; s < BSTAGES; ++s) {
            mbar_init(bar(BAR_BFULL + s), 1);
            mbar_init(bar(BAR_BEMPTY + s), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(bar(BAR_TFULL + b), 1);
            mbar_init(bar(BAR_TEMPTY + b), 8);          // one arrival per epilogue warp of the pair
        }
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc_pair(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_x);
        prefetch_tmap(&tm_head);
        prefetch_tmap(&tm_w);
        prefetch_tmap(&tm_y);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer: one patch per (dt, chunk), one weight box per tap
        if (lane == 0) {
            uint32_t pit = 0, bit = 0;
            for (int id = first; id < n_tiles_total; id += stride) {
                const Tile tl = decode_tile(p, id);
                const int wtile = tl.w0 + (int)rank * TILE_W;
                const int b_row = tl.nt * BN + (int)rank * (BN / 2);
                // strided (down-sampler) form: the 9 spatial taps are separate K-step groups, each a TMA box that samples every
                // second row / column (element strides 2) starting at the tap's offset — the rows past the frame's bottom / right
                // edge that the reference pads with zeros (attn_video_vae.py:239-241) are the TMA unit's zero fill
                const int n_sp = (KS == 1 && p.s2) ? 9 : 1;
                for (int dt = 0; dt < p.kt; ++dt) {
                    // the causal head: frames before the slice come from the previous slice's memory, or repeat frame 0
                    int ts = p.st * tl.t + dt - p.head_len;
                    const CUtensorMap* ta = &tm_x;
                    if (ts < 0) {
                        if (p.has_head) { ta = &tm_head; ts += p.head_len; }
                        else ts = 0;
                    }
                    for (int sp = 0; sp < n_sp; ++sp)
                    for (int cc = 0; cc < c_chunks; ++cc, ++pit) {
                        const uint32_t ps = pit % pstages, puse = pit / pstages;
                        mbar_wait(bar(BAR_PEMPTY + ps), (puse & 1) ^ 1);
                        if (leader) mbar_expect_tx(bar(BAR_PFULL + ps), 2 * patch_bytes);
                        const int sdh = sp / 3, sdw = sp - sdh * 3;
                        const int ww = (KS == 1 && p.s2) ? 2 * wtile + sdw : wtile - half;
                        const int hh = (KS == 1 && p.s2) ? 2 * tl.h0 + sdh : tl.h0 - half;
                        tma_load_4d_pair(base + ps * pslot, ta, bar(BAR_PFULL + ps), cc * BK, ww, hh, ts);
                        const int kcol0 = (dt * n_sp + sp) * taps_sp * p.Cin + cc * BK;
                        for (int tap = 0; tap < taps_sp; tap += tpb, ++bit) {
                            const uint32_t bs = bit % BSTAGES, buse = bit / BSTAGES;
                            mbar_wait(bar(BAR_BEMPTY + bs), (buse & 1) ^ 1);
                            if (leader) mbar_expect_tx(bar(BAR_BFULL + bs), 2 * tpb * B_BYTES);
                            for (int j = 0; j < tpb; ++j)
                                tma_load_2d_pair(base + C::OFF_B + bs * BSTAGE_BYTES + j * B_BYTES, &tm_w, bar(BAR_BFULL + bs),
                                                 kcol0 + (tap + j) * p.Cin, b_row);
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ tcgen05.mma issuer of the pair (leader CTA)
        // The whole warp walks the loop and waits on the barriers; one elected lane issues.  (A single lane looping while the
        // other 31 sit at the final barrier measured 30 % slower at N = 128, 10 % at N = 256 — run 29.)
        if (leader) {
            constexpr uint32_t idesc = idesc_16(F16, 2 * BM, BN);
            constexpr uint32_t hi_b = smem_desc_hi_sw128(1024);
            constexpr uint32_t hi_a = smem_desc_hi_sw128((uint32_t)slot_w * 128u);      // 8-row core groups one patch row apart
            uint32_t pit = 0, bit = 0, tc = 0;
            for (int id = first; id < n_tiles_total; id += stride, ++tc) {
                const uint32_t b = tc & 1;
                mbar_wait(bar(BAR_TEMPTY + b), ((tc >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem + b * BN;
                for (int pp = 0; pp < n_patches; ++pp, ++pit) {
                    const uint32_t ps = pit % pstages, puse = pit / pstages;
                    mbar_wait(bar(BAR_PFULL + ps), puse & 1);
                    const uint32_t patch = base + ps * pslot;
#pragma unroll
                    for (int tap0 = 0; tap0 < taps_sp; tap0 += tpb, ++bit) {
                        const uint32_t bs = bit % BSTAGES, buse = bit / BSTAGES;
                        mbar_wait(bar(BAR_BFULL + bs), buse & 1);
                        tc_fence_after();
                        if (elect_one()) {
                            const uint32_t bstage = base + C::OFF_B + bs * BSTAGE_BYTES;
#pragma unroll
                            for (int j = 0; j < tpb; ++j) {
                                // tap (dh, dw): the patch from row slot dh * slot_w + dw on.  The start is 128-byte, not
                                // 1024-byte aligned; the swizzle XOR is taken from the absolute shared-memory address bits
                                // (measured, run 26: base-offset field 0 is right, base offset = dw is wrong), which is also
                                // what the TMA write used
                                const int dh = (tap0 + j) / KS, dw = (tap0 + j) % KS;
                                const uint32_t alo = smem_desc_lo(patch + (uint32_t)(dh * slot_w + dw) * 128u, 16);
                                const uint32_t blo = smem_desc_lo(bstage + j * B_BYTES, 16);
#pragma unroll
                                for (int kk = 0; kk < BK / 16; ++kk)
                                    mma2_ss(d_tmem, make_desc(alo + kk * 2, hi_a), make_desc(blo + kk * 2, hi_b), idesc,
                                            (uint32_t)((pp | (tap0 + j) | kk) != 0));
                            }
                            tc_commit_pair(bar(BAR_BEMPTY + bs));
                            if (tap0 + tpb == taps_sp) {
                                tc_commit_pair(bar(BAR_PEMPTY + ps));
                                if (pp == n_patches - 1) tc_commit_pair(bar(BAR_TFULL + b));
                            }
                        }
                        __syncwarp();
                    }
                }
            }
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue: one thread per output position
        const int q = warp & 3;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        unsigned char* slab0 = smem + C::OFF_SLAB + q * 2 * SLAB_BYTES;
        const uint32_t tempty_g0 = mapa(bar(BAR_TEMPTY + 0), 0), tempty_g1 = mapa(bar(BAR_TEMPTY + 1), 0);
        const __half* bias = reinterpret_cast<const __half*>(p.bias);          // 2-byte elements either way
        const __half* resid = reinterpret_cast<const __half*>(p.resid);
        uint32_t tc = 0, n_store = 0;
        Running run = {0.f, 0.f, 0.f};            // lane L: group L of the current frame, over this warp's positions
        int run_t = -1;
        const int qpg = p.Cout >> 7;              // channel quads per group (32 groups)
        auto flush_stats = [&]() {
            if (run_t >= 0 && run.n > 0.f) {
                float* dst = p.gn_partials + (((int64_t)run_t * p.gn_slots + (int)blockIdx.x * 4 + q) * 32 + lane) * 3;
                dst[0] = run.n; dst[1] = run.mean; dst[2] = run.m2;
            }
            run.n = run.mean = run.m2 = 0.f;
        };
        for (int id = first; id < n_tiles_total; id += stride, ++tc) {
            const Tile tl = decode_tile(p, id);
            const uint32_t b = tc & 1;
            if (p.gn_partials != nullptr && tl.t != run_t) {
                flush_stats();
                run_t = tl.t;
            }
            const int hs = tl.h0 + q * 4, ws = tl.w0 + (int)rank * TILE_W;     // this warp's 4 rows x 8 columns of the tile
            const int h = hs + (lane >> 3), w = ws + (lane & 7);
            const bool valid = h < p.H && w < p.W;
            const int64_t pos = ((int64_t)tl.t * p.H + h) * p.W + w;
            const int col_tile = tl.nt * BN;
            mbar_wait(bar(BAR_TFULL + b), (tc >> 1) & 1);
            tc_fence_after();
            const uint32_t tacc = tmem + lane_off + b * BN;
#pragma unroll 1
            for (int sl = 0; sl < BN / 64; ++sl) {
                const int col0 = col_tile + sl * 64;
                uint32_t o[32];
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    const int c = col0 + hf * 32;
                    uint32_t v[32];
                    tmem_ld32(tacc + sl * 64 + hf * 32, v);
                    uint4 bq[4], rq[4];
                    if (bias != nullptr) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) bq[j] = __ldg(reinterpret_cast<const uint4*>(bias + c) + j);
                    }
                    if (resid != nullptr) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            rq[j] = valid ? __ldg(reinterpret_cast<const uint4*>(resid + pos * p.Cout + c) + j) : make_uint4(0u, 0u, 0u, 0u);
                    }
                    tc_wait_ld();
                    const uint32_t* bw = reinterpret_cast<const uint32_t*>(bq);
                    const uint32_t* rw = reinterpret_cast<const uint32_t*>(rq);
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        float a0 = __uint_as_float(v[2 * i]), a1 = __uint_as_float(v[2 * i + 1]);
                        if (bias != nullptr) {
                            const float2 bf = N::unpack(bw[i]);
                            a0 += bf.x; a1 += bf.y;
                        }
                        uint32_t h2 = N::pack(a0, a1);                        // the convolution's output in the tensor dtype
                        if (resid != nullptr) h2 = N::add(h2, rw[i]);         // + input_tensor, rounded again
                        o[hf * 16 + i] = h2;
                    }
                }
                if (sl == BN / 64 - 1) {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(b ? tempty_g1 : tempty_g0);
                }
                if (p.gn_partials != nullptr) {
                    const int gbase = col0 / (4 * qpg);                          // the slab's first group
                    if (qpg == 1) slab_stats<F16, 1>(o, valid, lane, gbase, run);
                    else if (qpg == 2) slab_stats<F16, 2>(o, valid, lane, gbase, run);
                    else slab_stats<F16, 4>(o, valid, lane, gbase, run);
                }
                unsigned char* slab = slab0 + (n_store & 1) * SLAB_BYTES;
                if (lane == 0) bulk_wait_read<1>();
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<uint4*>(slab + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                        make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0 && hs < p.H && ws < p.W) {
                    tma_store_4d(&tm_y, smem_u32(slab), col0, ws, hs, tl.t);      // box 64 c x 8 w x 4 h x 1 t, clipped at the edges
                    bulk_commit();
                }
                ++n_store;
            }
        }
        if (p.gn_partials != nullptr) flush_stats();
        if (lane == 0) bulk_wait<0>();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 2) {
        __syncwarp();
        tmem_dealloc_pair(tmem, TMEM_COLS);
    }
}

template <int BN, bool F16, int KS>
int launch(const void* x, const void* head, const void* w, void* y, Params p, cudaStream_t st) {
    const char* name = "svr2_causal_conv3d_fwd";
    p.h_tiles = (p.H + TILE_H - 1) / TILE_H;
    p.w_tiles = (p.W + 2 * TILE_W - 1) / (2 * TILE_W);
    p.n_tiles = p.Cout / BN;
    constexpr int slot_w = KS == 3 ? 16 : 8, box_h = TILE_H + 2 * (KS / 2);
    CUtensorMap tx, th, tw, ty;
    int rc;
    const int es = p.s2 ? 2 : 1;                 // element stride of the input boxes along W and H
    if ((rc = make_tmap_thwc(&tx, x, p.Ti, p.Hi, p.Wi, p.Cin, slot_w, box_h, es)) != SVR2_OK) return rc;
    if (p.has_head) {
        if ((rc = make_tmap_thwc(&th, head, p.head_len, p.Hi, p.Wi, p.Cin, slot_w, box_h, es)) != SVR2_OK) return rc;
    } else {
        th = tx;
    }
    const int64_t ktot = (int64_t)p.kt * p.ks * p.ks * p.Cin;
    if ((rc = make_tmap_2d(&tw, w, p.Cout, ktot, ktot, BN / 2)) != SVR2_OK) return rc;
    if ((rc = make_tmap_thwc(&ty, y, p.T, p.H, p.W, p.Cout, TILE_W, 4)) != SVR2_OK) return rc;
    static PerDeviceFlag smem_ok;
    if ((rc = ensure_dyn_smem(conv3d_tc_kernel<BN, F16, KS>, Cfg<BN>::SMEM_BYTES, smem_ok, name)) != SVR2_OK) return rc;
    const int pairs = sm_count() / 2;
    const int64_t tiles = (int64_t)p.T * p.h_tiles * p.w_tiles * p.n_tiles;
    const int grid = 2 * (int)(tiles < pairs ? tiles : pairs);
    conv3d_tc_kernel<BN, F16, KS><<<grid, Cfg<BN>::NTHREADS, Cfg<BN>::SMEM_BYTES, st>>>(tx, th, tw, ty, p);
    SVR2_CHECK_LAUNCH(name);
    return SVR2_OK;
}

// [Cout, Cin, kt, kh, kw] (nn.Conv3d) -> [Cout, kt, kh, kw, Cin]: the implicit GEMM's K axis is (tap, channel)
__global__ void pack_weight_kernel(const uint16_t* __restrict__ w, uint16_t* __restrict__ out, int Cout, int Cin, int taps) {
    const int64_t n = (int64_t)Cout * Cin * taps;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int ci = (int)(i % Cin);
        const int tap = (int)((i / Cin) % taps);
        const int64_t co = i / ((int64_t)Cin * taps);
        out[i] = w[(co * Cin + ci) * taps + tap];
    }
}

}  // namespace conv
}  // namespace svr2

extern "C" int svr2_conv3d_pack_weight(const void* weight, void* packed, int c_out, int c_in, int kt, int kh, int kw, void* stream) {
    SVR2_CHECK_ARG(weight && packed, "svr2_conv3d_pack_weight: null pointer");
    SVR2_CHECK_ARG(c_out > 0 && c_in > 0 && kt > 0 && kh > 0 && kw > 0, "svr2_conv3d_pack_weight: bad shape");
    const int64_t n = (int64_t)c_out * c_in * kt * kh * kw;
    const int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
    svr2::conv::pack_weight_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>((const uint16_t*)weight, (uint16_t*)packed, c_out, c_in,
                                                                            kt * kh * kw);
    SVR2_CHECK_LAUNCH("svr2_conv3d_pack_weight");
    return SVR2_OK;
}

extern "C" int svr2_conv3d_gn_slots(void) { return svr2::sm_count() * 4; }

extern "C" int svr2_causal_conv3d_fwd(const void* x, const void* head, const void* weight_packed, const void* bias, const void* resid,
                                      void* y, void* gn_partials, int T, int H, int W, int c_in, int c_out, int kt, int ks, int dtype,
                                      void* stream) {
    const char* name = "svr2_causal_conv3d_fwd";
    SVR2_CHECK_ARG(x && weight_packed && y, "%s: null pointer", name);
    SVR2_CHECK_ARG(dtype == SVR2_DTYPE_F16 || dtype == SVR2_DTYPE_BF16, "%s: dtype %d (fp16 or bf16)", name, dtype);
    SVR2_CHECK_ARG(T >= 0 && H > 0 && W > 0, "%s: bad extent T=%d H=%d W=%d", name, T, H, W);
    SVR2_CHECK_ARG((kt == 1 || kt == 3) && (ks == 1 || ks == 3), "%s: kernel %dx%dx%d (temporal 1|3, spatial 1|3, stride 1)", name, kt, ks, ks);
    SVR2_CHECK_ARG(c_in > 0 && c_in % 64 == 0, "%s: c_in=%d must be a positive multiple of 64", name, c_in);
    SVR2_CHECK_ARG(c_out > 0 && c_out % 128 == 0, "%s: c_out=%d must be a positive multiple of 128", name, c_out);
    SVR2_CHECK_ARG((int64_t)T * H * W < ((int64_t)1 << 31), "%s: too many positions", name);
    if (T == 0) return SVR2_OK;
    svr2::conv::Params p{};
    p.T = T; p.H = H; p.W = W; p.Cin = c_in; p.Cout = c_out; p.kt = kt; p.ks = ks;
    p.Ti = T; p.Hi = H; p.Wi = W; p.st = 1; p.s2 = 0; p.head_len = kt - 1;
    p.has_head = (head != nullptr && kt > 1) ? 1 : 0;
    p.bias = bias; p.resid = resid;
    SVR2_CHECK_ARG(gn_partials == nullptr || c_out == 128 || c_out == 256 || c_out == 512,
                   "%s: output statistics are gathered for 32 groups of 4, 8 or 16 channels (c_out=%d)", name, c_out);
    p.gn_partials = (float*)gn_partials;
    p.gn_slots = svr2::sm_count() * 4;
    if (gn_partials != nullptr) {
        // a warp writes its slot of a frame only if it saw positions of that frame: the rest must read as count 0
        const cudaError_t e = cudaMemsetAsync(gn_partials, 0, (size_t)T * p.gn_slots * 32 * 3 * sizeof(float), (cudaStream_t)stream);
        if (e != cudaSuccess) { svr2::set_error("%s: cudaMemsetAsync: %s", name, cudaGetErrorString(e)); return SVR2_E_CUDA; }
    }
    const bool f16 = dtype == SVR2_DTYPE_F16;
    cudaStream_t st = (cudaStream_t)stream;
    using namespace svr2::conv;
#define SVR2_CONV_GO(BN, F, K) return launch<BN, F, K>(x, head, weight_packed, y, p, st)
    if (c_out % 256 == 0) {
        if (ks == 3) { if (f16) SVR2_CONV_GO(256, true, 3); SVR2_CONV_GO(256, false, 3); }
        if (f16) SVR2_CONV_GO(256, true, 1);
        SVR2_CONV_GO(256, false, 1);
    }
    if (ks == 3) { if (f16) SVR2_CONV_GO(128, true, 3); SVR2_CONV_GO(128, false, 3); }
    if (f16) SVR2_CONV_GO(128, true, 1);
    SVR2_CONV_GO(128, false, 1);
#undef SVR2_CONV_GO
}

extern "C" int svr2_causal_conv3d_down_fwd(const void* x, const void* head, int head_frames, const void* weight_packed, const void* bias,
                                           void* y, int T, int H, int W, int c_in, int c_out, int kt, int pad_hw, int dtype, void* stream) {
    const char* name = "svr2_causal_conv3d_down_fwd";
    SVR2_CHECK_ARG(x && weight_packed && y, "%s: null pointer", name);
    SVR2_CHECK_ARG(dtype == SVR2_DTYPE_F16 || dtype == SVR2_DTYPE_BF16, "%s: dtype %d (fp16 or bf16)", name, dtype);
    SVR2_CHECK_ARG(kt == 1 || kt == 3, "%s: temporal kernel %d (1: spatial down-sampling, 3: spatial + temporal)", name, kt);
    SVR2_CHECK_ARG(pad_hw == 0 || pad_hw == 1, "%s: pad_hw is the number of zero rows / columns after the last one (0 | 1)", name);
    SVR2_CHECK_ARG(c_in > 0 && c_in % 64 == 0 && c_out > 0 && c_out % 128 == 0, "%s: c_in %% 64, c_out %% 128 (got %d, %d)", name, c_in, c_out);
    SVR2_CHECK_ARG(T >= 0 && H + pad_hw >= 3 && W + pad_hw >= 3, "%s: bad extent T=%d H=%d W=%d", name, T, H, W);
    SVR2_CHECK_ARG(head == nullptr || (kt == 3 && head_frames >= 1 && head_frames <= 2), "%s: head of %d frames", name, head_frames);
    svr2::conv::Params p{};
    p.kt = kt; p.ks = 3; p.Cin = c_in; p.Cout = c_out;
    p.Ti = T; p.Hi = H; p.Wi = W;
    p.s2 = 1;
    p.st = kt == 3 ? 2 : 1;
    p.has_head = head != nullptr ? 1 : 0;
    p.head_len = kt == 3 ? (head != nullptr ? head_frames : 2) : 0;      // first slice: the first frame twice (extend_head, times = 2)
    p.T = kt == 3 ? (T + p.head_len - 3) / 2 + 1 : T;
    p.H = (H + pad_hw - 3) / 2 + 1;
    p.W = (W + pad_hw - 3) / 2 + 1;
    if (T == 0 || p.T <= 0) return SVR2_OK;
    p.bias = bias; p.resid = nullptr; p.gn_partials = nullptr; p.gn_slots = 0;
    const bool f16 = dtype == SVR2_DTYPE_F16;
    cudaStream_t st = (cudaStream_t)stream;
    using namespace svr2::conv;
    if (c_out % 256 == 0) return f16 ? launch<256, true, 1>(x, head, weight_packed, y, p, st) : launch<256, false, 1>(x, head, weight_packed, y, p, st);
    return f16 ? launch<128, true, 1>(x, head, weight_packed, y, p, st) : launch<128, false, 1>(x, head, weight_packed, y, p, st);
}
