// svr2_window_attn_fwd — PERSISTENT variable-length NaSwin window attention for sm_100a (TMA + tcgen05 + TMEM).
//
// Replaces FlashAttentionVarlen / pytorch_varlen_attention (reference src/models/dit_3b/attention.py:24-61,138-157: a
// Python loop of one SDPA call per window), unconcat_win (na.py:396-417) and window_reverse (na.py:638).
//
// Design history (measurements under profiles/):
//  gen 1  one CTA per work item: with NaSwin's short sequences (4 key tiles per window at 2048^2) 40 % of a CTA's life
//         was exposed prologue / epilogue / launch gap                                   -> 285 TFLOP/s at C1
//  gen 2  persistent CTA per SM, roles run ahead across items                             -> 476 TFLOP/s at C1
//         but the chain  S(j) -> softmax -> P(j) -> [PV(j), QK(j+1)] -> S(j+1)  is serial per query tile: a lone warp
//         per SM sub-partition needs ~590 cycles per 32 columns (issue latency, not a pipe limit — the microbenchmarks
//         show tcgen05.ld at ~1 KB/clk/SM and no MMA/LDTM interference), so one key tile cost 3200 + 1300 cycles.
//  gen 3  (this file) the key dimension is processed in HALF tiles of 64 columns with TWO S buffers per query tile:
//         QK(h+1) and QK(h+2) are queued before P(h) exists, so the softmax warps never wait for the tensor pipe and the
//         tensor pipe never waits for more than the P it is about to consume.
//
// Work item = (window, pair of 128-row query tiles, head); items are sorted longest sequence first and CTA b takes
// items b, b+G, b+2G, ...   Roles (15 warps, 480 threads):
//   warps 0-3   softmax group 0 (query tile 0): one thread per query row, S read from TMEM once per half tile, exp2
//               against a lagging reference max (softmax is shift invariant), bf16 P written back over S
//   warps 4-7   softmax group 1 (query tile 1)
//   warps 8-11  epilogue: O/l -> bf16 -> swizzled staging tile -> coalesced 256-B row stores to the token's ORIGINAL
//               position (window reverse) or the per-window text buffer; frees O for the next item as soon as it is read
//   warp  12    TMA producer (one thread): Q tiles, then K/V through a 4-stage ring of 32-KB tiles, across items
//   warps 13,14 tcgen05.mma issuers, one per query tile: S_i = Q_i K^T (SS), O_i += P_i V (A = P from TMEM, B = V MN-major);
//               one elected lane issues, the whole warp runs the (warp-uniform) control flow
// The key length is trimmed to a multiple of 32 columns (MMA N / K steps and the softmax chunks shrink with it).
//
// TMEM map (columns): S_i[b] at 128 i + 64 b (i = query tile, b = half-tile parity), O_0 [256,384), O_1 [384,512);
//                     P aliases the first 32 columns of its S buffer.
// SMEM map: Q0 | Q1 | ring[4] | staging (32 KB each) | row sums l[2][2][128] | mbarriers.
#include <cuda.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace svr2 {
namespace tc2 {

using namespace ptx;

constexpr int D = SVR2_HEAD_DIM;          // 128
constexpr int BM = 128, BN = 128, HN = 64;
constexpr int TILE_BYTES = BM * D * 2;    // 32768
constexpr int BOX_BYTES = TILE_BYTES / 2; // 128 rows x 64 cols bf16
constexpr int STAGES = 4;
constexpr int NTHREADS = 480;    // 15 warps
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_STAGE = OFF_RING + STAGES * TILE_BYTES;
constexpr int OFF_L = OFF_STAGE + TILE_BYTES;          // float l[2 parities][2 tiles][128 rows]
constexpr int OFF_BAR = OFF_L + 2 * 2 * 128 * 4;
constexpr int SMEM_BYTES = OFF_BAR + 256;              // 231,680 B of the 232,448 B a CTA may opt in to
constexpr uint32_t TMEM_COLS = 512;
constexpr float kScaleLog2 = 0.08838834764831845f * 1.4426950408889634f;   // 1/sqrt(128) * log2(e)
constexpr float kRescaleThreshold = 8.0f;   // lazy rescale: the reference lags the running max by up to 2^8
constexpr float kMaxExponent = 100.0f;      // P <= 2^100: a row can never overflow bf16 / the fp32 sums (128 * 2^100 << 2^127)

// barrier slots (8 bytes each); SFULL / PREADY are per (query tile i, S buffer b): slot + 2 i + b
enum { BAR_QFULL = 0, BAR_QEMPTY = 2, BAR_SFULL = 4, BAR_PREADY = 8, BAR_OFINAL = 12, BAR_OEMPTY = 14, BAR_LREADY = 16,
       BAR_PVDONE = 18, BAR_KVFULL = 20, BAR_KVEMPTY = 20 + STAGES, BAR_COUNT = 20 + 2 * STAGES };
static_assert(BAR_COUNT * 8 + 8 <= 256, "barrier block too small");

struct Params {
    const int32_t* tab;
    const int32_t* work;
    int n_items;           // work-list entries x heads
    int H, W, txt_len, heads;
    __nv_bfloat16* out_vid;
    __nv_bfloat16* txt_part;
    long long* trace;      // debug timeline (svr2_debug_attn_trace) or nullptr
    int trace_ctas;
};
// 0 entry, 1 setup done, 5 end, 6 smid, 7 items done; for the CTA's first kTraceItems items n, at 8 + 9n:
// +0 MMA thread starts the item (Q and K0 landed)   +1 softmax-0 sees S(0)   +2 softmax-0 published P(0)
// +3 softmax-0 sees S(last)   +4 softmax-0 published P(last)   +5 epilogue sees O_0 final   +6 tile-0 stores issued
// +7 tile-1 stores issued   +8 TMA thread issued the item's last K/V load
constexpr int kTraceSlots = 64, kTraceItems = 4;
// fine stamps (item 2, half tile 3) of softmax group 0 at 44.. and group 1 at 52..: S seen, S landed in registers, max done,
// MUFU token acquired, exp2 done, wait::st done, P published, next S seen
#define SVR2_FINE(k) do { if (p.trace != nullptr && fine) trace_at(p, 44 + 8 * i + (k), clock64()); } while (0)

// Everything a role needs to know about one work item; every role decodes it identically.
struct Item {
    Win w;
    int wi, head, seq, q0, nq, n_kv, n_half, cols_total;
};
__device__ __forceinline__ Item decode_item(const Params& p, int item) {
    Item it;
    const int e = item / p.heads;
    it.head = item - e * p.heads;
    it.wi = __ldg(p.work + 2 * e);
    const int qpair = __ldg(p.work + 2 * e + 1);
    it.w = load_win(p.tab, it.wi);
    it.seq = it.w.n_vid + p.txt_len;
    it.q0 = qpair * 2 * BM;
    it.nq = (it.seq - it.q0 > BM) ? 2 : 1;
    it.n_kv = (it.seq + BN - 1) / BN;                  // 128-row K/V tiles moved by TMA
    it.cols_total = (it.seq + 31) & ~31;               // key columns the MMAs / softmax touch (multiple of 32)
    it.n_half = (it.cols_total + HN - 1) / HN;         // 64-column half tiles
    return it;
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(fmaxf(a, b), c); }

// packed fp32 pairs (Blackwell FFMA2 / FADD2): two lanes of arithmetic per issue slot
__device__ __forceinline__ uint64_t pack_f2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack_f2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma_f2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t add_f2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// row max of one 32-column chunk (masked columns were set to -inf beforehand)
__device__ __forceinline__ float chunk_max(const uint32_t* v) {
    float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
    for (int c = 0; c < 32; c += 8) {
        m0 = max3(m0, __uint_as_float(v[c]), __uint_as_float(v[c + 1]));
        m1 = max3(m1, __uint_as_float(v[c + 2]), __uint_as_float(v[c + 3]));
        m2 = max3(m2, __uint_as_float(v[c + 4]), __uint_as_float(v[c + 5]));
        m3 = max3(m3, __uint_as_float(v[c + 6]), __uint_as_float(v[c + 7]));
    }
    return fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
}
// P = exp2(min(s k - mb, kMaxExponent)) for one chunk, packed to bf16 pairs; returns the chunk's row sum.  Branch free:
// FFMA2 (1/2 per element), FMNMX clamp, MUFU.EX2, F2FP (1/2), FADD2 (1/2).  -inf (masked) columns give exactly 0.
__device__ __forceinline__ float chunk_exp(const uint32_t* v, uint32_t* pk, float mb) {
    const uint64_t kk = pack_f2(kScaleLog2, kScaleLog2), nmb = pack_f2(-mb, -mb);
    uint64_t acc0 = pack_f2(0.f, 0.f), acc1 = acc0;
#pragma unroll
    for (int c = 0; c < 16; c += 2) {
        float x0, x1, x2, x3;
        unpack_f2(fma_f2(pack_f2(__uint_as_float(v[2 * c]), __uint_as_float(v[2 * c + 1])), kk, nmb), x0, x1);
        unpack_f2(fma_f2(pack_f2(__uint_as_float(v[2 * c + 2]), __uint_as_float(v[2 * c + 3])), kk, nmb), x2, x3);
        const float p0 = ex2(fminf(x0, kMaxExponent)), p1 = ex2(fminf(x1, kMaxExponent));
        const float p2 = ex2(fminf(x2, kMaxExponent)), p3 = ex2(fminf(x3, kMaxExponent));
        pk[c] = pack2(p0, p1);
        pk[c + 1] = pack2(p2, p3);
        acc0 = add_f2(acc0, pack_f2(p0, p1));
        acc1 = add_f2(acc1, pack_f2(p2, p3));
    }
    float s0, s1;
    unpack_f2(add_f2(acc0, acc1), s0, s1);
    return s0 + s1;
}
__device__ __forceinline__ void mask_chunk(uint32_t* v, int n_valid) {      // columns >= n_valid -> -inf
#pragma unroll
    for (int c = 0; c < 32; ++c) v[c] = (c < n_valid) ? v[c] : 0xff800000u;
}

__device__ __forceinline__ void trace_at(const Params& p, int slot, long long v) {
    if (p.trace != nullptr && (int)blockIdx.x < p.trace_ctas && slot < kTraceSlots)
        p.trace[(long long)blockIdx.x * kTraceSlots + slot] = v;
}
__device__ __forceinline__ void trace_item(const Params& p, int n, int what) {
    if (p.trace != nullptr && n < kTraceItems) trace_at(p, 8 + 9 * n + what, clock64());
}

// 15 warps = 4+4+4+3 per SM sub-partition; a sub-partition's 16 K registers / 4 warps caps a thread at 128 registers
__global__ void __launch_bounds__(NTHREADS, 1)
window_attn_tc2_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t base = smem_u32(smem);                          // SWIZZLE_128B tiles need 1024-B alignment
    const uint32_t sRing = base + OFF_RING;
    const uint32_t sBar = base + OFF_BAR;
    float* l_buf = reinterpret_cast<float*>(smem + OFF_L);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + OFF_BAR + BAR_COUNT * 8);
    auto bar = [&](int slot) { return sBar + 8u * (uint32_t)slot; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        if (base & 1023u) {
            printf("svr2: dynamic shared memory base 0x%x is not 1024-B aligned\n", base);
            __trap();
        }
        trace_at(p, 0, clock64());
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(BAR_QFULL + i), 1);
            mbar_init(bar(BAR_QEMPTY + i), 1);
            mbar_init(bar(BAR_OFINAL + i), 1);
            mbar_init(bar(BAR_OEMPTY + i), 128);
            mbar_init(bar(BAR_LREADY + i), 128);
            mbar_init(bar(BAR_PVDONE + i), 1);
        }
        for (int s = 0; s < 4; ++s) {
            mbar_init(bar(BAR_SFULL + s), 1);
            mbar_init(bar(BAR_PREADY + s), 128);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_KVFULL + s), 1);
            mbar_init(bar(BAR_KVEMPTY + s), 2);       // one commit per MMA warp (two from warp 13 when tile 1 is absent)
        }
        fence_mbar_init();
    }
    if (warp == 13) tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
    if (warp == 12 && lane == 0) {
        prefetch_tmap(&tm_q);
        prefetch_tmap(&tm_k);
        prefetch_tmap(&tm_v);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (threadIdx.x == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        trace_at(p, 1, clock64());
        trace_at(p, 6, smid);
    }

    if (warp == 12) {
        // ------------------------------------------------------------------ TMA producer (runs ahead across items)
        if (lane == 0) {
            uint32_t kv_n = 0, c0 = 0, c1 = 0;
            int n_it = 0;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++n_it) {
                const Item it = decode_item(p, item);
                for (int i = 0; i < it.nq; ++i) {
                    const int row = it.w.row_start + it.q0 + i * BM;
                    const uint32_t c = i ? c1 : c0;
                    mbar_wait(bar(BAR_QEMPTY + i), (c & 1) ^ 1);              // all QK MMAs of the previous user are done
                    mbar_expect_tx(bar(BAR_QFULL + i), TILE_BYTES);
                    tma_load_3d(base + i * TILE_BYTES, &tm_q, bar(BAR_QFULL + i), 0, it.head, row);
                    tma_load_3d(base + i * TILE_BYTES + BOX_BYTES, &tm_q, bar(BAR_QFULL + i), 64, it.head, row);
                    if (i) ++c1; else ++c0;
                }
                for (int n = 0; n < 2 * it.n_kv; ++n, ++kv_n) {               // ring order K0 V0 K1 V1 ...
                    const uint32_t st = kv_n % STAGES, use = kv_n / STAGES;
                    const CUtensorMap* map = (n & 1) ? &tm_v : &tm_k;
                    const int row = it.w.row_start + (n >> 1) * BN;
                    mbar_wait(bar(BAR_KVEMPTY + st), (use & 1) ^ 1);
                    mbar_expect_tx(bar(BAR_KVFULL + st), TILE_BYTES);
                    tma_load_3d(sRing + st * TILE_BYTES, map, bar(BAR_KVFULL + st), 0, it.head, row);
                    tma_load_3d(sRing + st * TILE_BYTES + BOX_BYTES, map, bar(BAR_KVFULL + st), 64, it.head, row);
                }
                trace_item(p, n_it, 8);
            }
        }
    } else if (warp >= 13) {
        // ------------------------------------------------------------------ tcgen05.mma issuer
        // The WHOLE warp runs this loop (warp-uniform control flow keeps the descriptors in uniform registers: a
        // lane-0-only branch made the compiler serialise every UTCHMMA operand through a per-lane waterfall loop and the
        // issue rate, ~90 cycles per MMA, was the kernel's bottleneck); one elected lane issues the MMAs and commits.
        {
            constexpr uint32_t idesc_pv = idesc_bf16(BM, D, 1);
            constexpr uint32_t hi_k = smem_desc_hi_sw128(1024);       // K-major operands (Q, K): 8-row groups 1024 B apart
            constexpr uint32_t hi_v = smem_desc_hi_sw128(1024);       // MN-major V: 8-key groups 1024 B apart, LBO = next 64 d
            // S_i[h&1] = Q_i K(half h)^T : 8 x (128 x ncols x 16); the half's 64 key rows start 64 * 128 B into each box
            auto issue_qk = [&](int i, int h, uint32_t kbase, int ncols) {
                const uint32_t idesc_qk = idesc_bf16(BM, ncols, 0);
                const uint32_t d_tmem = tmem + i * 128 + (h & 1) * HN;
                const uint32_t qlo = smem_desc_lo(base + i * TILE_BYTES, 16);
                const uint32_t klo = smem_desc_lo(kbase + (h & 1) * (HN * 128), 16);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t off = ((uint32_t)(kk >> 2) * BOX_BYTES + (uint32_t)(kk & 3) * 32u) >> 4;
                    mma_ss(d_tmem, make_desc(qlo + off, hi_k), make_desc(klo + off, hi_k), idesc_qk, kk > 0);
                }
            };
            // O_i += P_i(h) V(half h) : A = P from TMEM (16 keys = 8 columns per step), B = V MN-major (16 keys = 2048 B)
            auto issue_pv = [&](int i, int h, uint32_t vbase, int ksteps) {
                const uint32_t a_tmem = tmem + i * 128 + (h & 1) * HN;
                const uint32_t vlo = smem_desc_lo(vbase + (h & 1) * 4 * 2048, BOX_BYTES);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                    if (kk < ksteps)
                        mma_ts(tmem + 256 + i * 128, a_tmem + kk * 8, make_desc(vlo + kk * (2048 >> 4), hi_v), idesc_pv,
                               !(h == 0 && kk == 0));
            };
            const int i = warp - 13;     // this warp's query tile
            uint32_t kv0 = 0;            // ring index of the current item's K tile 0 (K_t at 2t, V_t at 2t+1)
            uint32_t cq = 0;             // items in which this query tile took part
            uint32_t pr_par = 0;         // parity bits of PREADY(i, b) at bit b
            int n_done = 0;
            auto ring_slot = [&](uint32_t idx) { return (kv0 + idx) % STAGES; };
            auto ring_par = [&](uint32_t idx) { return ((kv0 + idx) / STAGES) & 1u; };
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++n_done) {
                const Item it = decode_item(p, item);
                if (i >= it.nq) {                       // tile 1 absent: warp 13 releases the K/V tiles for both warps
                    kv0 += 2 * it.n_kv;
                    continue;
                }
                const bool both = it.nq == 1;           // commit the ring barriers twice
                auto cols = [&](int h) { const int r = it.cols_total - h * HN; return r < HN ? r : HN; };
                auto release = [&](uint32_t idx) {
                    tc_commit(bar(BAR_KVEMPTY + ring_slot(idx)));
                    if (both) tc_commit(bar(BAR_KVEMPTY + ring_slot(idx)));
                };
                mbar_wait(bar(BAR_QFULL + i), cq & 1);
                mbar_wait(bar(BAR_KVFULL + ring_slot(0)), ring_par(0));
                tc_fence_after();
                // two half tiles of S are queued before any P exists
                if (elect_one()) {
                    if (i == 0) trace_item(p, n_done, 0);
                    const uint32_t k0 = sRing + ring_slot(0) * TILE_BYTES;
                    issue_qk(i, 0, k0, cols(0));
                    tc_commit(bar(BAR_SFULL + 2 * i));
                    if (it.n_half > 1) {
                        issue_qk(i, 1, k0, cols(1));
                        tc_commit(bar(BAR_SFULL + 2 * i + 1));
                    }
                    if (it.n_half <= 2) tc_commit(bar(BAR_QEMPTY + i));                  // that was the last read of Q_i
                    release(0);                                                          // K tile 0 is consumed
                }
                __syncwarp();
                for (int h = 0; h < it.n_half; ++h) {
                    const int t = h >> 1, hn = h + 2, b = h & 1;
                    const bool last = h == it.n_half - 1;
                    if (b == 0) mbar_wait(bar(BAR_KVFULL + ring_slot(2 * t + 1)), ring_par(2 * t + 1));                 // V_t
                    if (hn < it.n_half && b == 0) mbar_wait(bar(BAR_KVFULL + ring_slot(hn)), ring_par(hn));              // K_{hn/2}
                    mbar_wait(bar(BAR_PREADY + 2 * i + b), (pr_par >> b) & 1u);
                    pr_par ^= 1u << b;
                    const bool mfine = p.trace != nullptr && i == 0 && lane == 0 && n_done == 2 && h >= 2 && h <= 4;
                    if (mfine) trace_at(p, 52 + 2 * (h - 2), clock64());
                    if (h == 0) mbar_wait(bar(BAR_OEMPTY + i), (cq & 1) ^ 1);            // epilogue drained O_i of the previous item
                    tc_fence_after();
                    if (elect_one()) {
                        issue_pv(i, h, sRing + ring_slot(2 * t + 1) * TILE_BYTES, cols(h) >> 4);
                        tc_commit(bar(BAR_PVDONE + i));
                        if (hn < it.n_half) {
                            issue_qk(i, hn, sRing + ring_slot(2 * (hn >> 1)) * TILE_BYTES, cols(hn));
                            tc_commit(bar(BAR_SFULL + 2 * i + b));
                            if (hn == it.n_half - 1) tc_commit(bar(BAR_QEMPTY + i));
                        } else if (last) {
                            tc_commit(bar(BAR_OFINAL + i));
                        }
                        if (b == 1 || last) release(2 * t + 1);
                        if (hn < it.n_half && ((hn & 1) == 1 || hn == it.n_half - 1)) release(2 * (hn >> 1));
                    }
                    __syncwarp();
                    if (mfine) trace_at(p, 53 + 2 * (h - 2), clock64());
                }
                kv0 += 2 * it.n_kv;
                ++cq;
            }
            if (lane == 0 && i == 0) trace_at(p, 7, n_done);
        }
    } else if (warp < 8) {
        // ------------------------------------------------------------------ softmax group i
        const int i = warp >> 2, quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        const uint32_t tO = tmem + lane_off + 256 + i * 128;
        uint32_t sf_par = 0;        // parity bits of SFULL(i, b) at bit b
        uint32_t g = 0;             // half tiles this group has processed = PV_i MMAs that were or will be issued before the next one
        uint32_t c_cnt = 0;
        int n_it = -1;
        // MUFU token: the two groups share each sub-partition's exp2 pipe (16 lanes/clk/SM, the kernel's tightest pipe after
        // the tensor core).  Left alone they run in lockstep — both in their exp2 burst, then both in their TMEM / barrier
        // overhead — and the pipe idles half the time.  Named barriers 1 (group 1 -> group 0) and 2 (group 0 -> group 1)
        // make the bursts alternate, so one group's overhead hides under the other's exp2.
        if (i == 1) named_bar_arrive(1, 256);
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const Item it = decode_item(p, item);
            ++n_it;
            if (i >= it.nq) continue;
            const bool paired = it.nq == 2;
            // The exponent reference `m` is NOT the exact running max: it is the max of the first 32 columns, raised lazily
            // (with O and l rescaled) once the true running max `m_seen` leads it by more than 2^8.  Softmax is shift
            // invariant, bf16 / fp32 precision is relative and the exponent is clamped far below overflow, so the result
            // is the same — and no half tile ever has to wait for a full-row max before its first exp2.
            float m = -INFINITY, m_seen = -INFINITY, l = 0.f;
            for (int h = 0; h < it.n_half; ++h, ++g) {
                const int b = h & 1;
                const bool last = h == it.n_half - 1;
                const int rem = it.cols_total - h * HN;
                const bool two = rem > 32;                               // 64 or 32 key columns in this half tile
                const int valid = it.seq - h * HN;                       // key columns >= valid are beyond the sequence
                const uint32_t tS = tmem + lane_off + i * 128 + b * HN;
                if (threadIdx.x == 0 && n_it == 2 && h == 4 && p.trace != nullptr) {
                    trace_at(p, 60, clock64());
                    trace_at(p, 61, mbar_try_wait(bar(BAR_SFULL + 2 * i + b), (sf_par >> b) & 1u) ? 1 : 0);
                }
                mbar_wait(bar(BAR_SFULL + 2 * i + b), (sf_par >> b) & 1u);
                sf_par ^= 1u << b;
                tc_fence_after();
                if (threadIdx.x == 0) {
                    if (h == 0) trace_item(p, n_it, 1);
                    if (last) trace_item(p, n_it, 3);
                }
                const bool fine = threadIdx.x == 0 && n_it == 2 && h == 3;
                if (threadIdx.x == 0 && n_it == 2 && h == 4 && p.trace != nullptr) trace_at(p, 44 + 7, clock64());
                SVR2_FINE(0);
                uint32_t a[32], a2[32];
                tmem_ld32(tS, a);
                if (two) tmem_ld32(tS + 32, a2);
                if (h > 0) {
                    const bool need = (m_seen - m) * kScaleLog2 > kRescaleThreshold;
                    if (__any_sync(0xffffffffu, need)) {                  // warp-uniform: tcgen05.ld/st are warp collectives
                        // rare.  O_i still receives P(h-1) V(h-1): wait for that MMA (the g-th PV of this query tile)
                        mbar_wait(bar(BAR_PVDONE + i), (g - 1) & 1u);
                        tc_fence_after();
                        const float m_use = need ? m_seen : m;
                        const float alpha = ex2((m - m_use) * kScaleLog2);
                        l *= alpha;
                        m = m_use;
                        tc_wait_ld();
                        uint32_t ov[32];
#pragma unroll 1
                        for (int c4 = 0; c4 < 4; ++c4) {
                            tmem_ld32(tO + c4 * 32, ov);
                            tc_wait_ld();
#pragma unroll
                            for (int c = 0; c < 32; ++c) ov[c] = __float_as_uint(__uint_as_float(ov[c]) * alpha);
                            tmem_st32(tO + c4 * 32, ov);
                        }
                    }
                }
                tc_wait_ld();
                SVR2_FINE(1);
                if (valid < rem) {                                       // only the last half tile of a sequence
                    if (valid < 32) mask_chunk(a, valid);
                    if (two) mask_chunk(a2, valid - 32);
                }
                float cm = chunk_max(a);
                if (h == 0) m = cm;
                if (two) cm = fmaxf(cm, chunk_max(a2));
                const float mb = m * kScaleLog2;
                uint32_t pk[16];
                SVR2_FINE(2);
                if (paired) named_bar_sync(i == 0 ? 1 : 2, 256);
                SVR2_FINE(3);
                float rs = chunk_exp(a, pk, mb);
                tmem_st16(tS, pk);
                if (two) {
                    uint32_t pk2[16];
                    rs += chunk_exp(a2, pk2, mb);
                    tmem_st16(tS + 16, pk2);
                }
                if (paired) named_bar_arrive(i == 0 ? 2 : 1, 256);
                SVR2_FINE(4);
                l += rs;
                m_seen = fmaxf(m_seen, cm);
                tc_wait_st();
                SVR2_FINE(5);
                tc_fence_before();
                mbar_arrive(bar(BAR_PREADY + 2 * i + b));
                SVR2_FINE(6);
                if (threadIdx.x == 0) {
                    if (h == 0) trace_item(p, n_it, 2);
                    if (last) trace_item(p, n_it, 4);
                }
            }
            // publish the row sum for the epilogue warps (double-buffered by item parity)
            l_buf[((c_cnt & 1) * 2 + i) * 128 + row] = l;
            mbar_arrive(bar(BAR_LREADY + i));
            ++c_cnt;
        }
    } else if (warp < 12) {
        // ------------------------------------------------------------------ epilogue warps (both query tiles, in turn)
        const int quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        unsigned char* stage = smem + OFF_STAGE;
        const int64_t row_stride = (int64_t)p.heads * D;
        uint32_t c0 = 0, c1 = 0;
        int n_it = 0;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++n_it) {
            const Item it = decode_item(p, item);
            for (int i = 0; i < it.nq; ++i) {
                const uint32_t cc = i ? c1 : c0;
                const uint32_t tO = tmem + lane_off + 256 + i * 128;
                mbar_wait(bar(BAR_LREADY + i), cc & 1);
                mbar_wait(bar(BAR_OFINAL + i), cc & 1);
                tc_fence_after();
                if (threadIdx.x == 256 && i == 0) trace_item(p, n_it, 5);
                const float inv = 1.0f / l_buf[((cc & 1) * 2 + i) * 128 + row];
#pragma unroll
                for (int c4 = 0; c4 < 4; ++c4) {
                    uint32_t ov[32];
                    tmem_ld32(tO + c4 * 32, ov);
                    tc_wait_ld();
                    if (c4 == 3) {                                   // O_i is in registers: the next item may overwrite it
                        tc_fence_before();
                        mbar_arrive(bar(BAR_OEMPTY + i));
                    }
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        uint4 u;
                        u.x = pack2(__uint_as_float(ov[8 * q + 0]) * inv, __uint_as_float(ov[8 * q + 1]) * inv);
                        u.y = pack2(__uint_as_float(ov[8 * q + 2]) * inv, __uint_as_float(ov[8 * q + 3]) * inv);
                        u.z = pack2(__uint_as_float(ov[8 * q + 4]) * inv, __uint_as_float(ov[8 * q + 5]) * inv);
                        u.w = pack2(__uint_as_float(ov[8 * q + 6]) * inv, __uint_as_float(ov[8 * q + 7]) * inv);
                        const int ch = c4 * 4 + q;
                        *reinterpret_cast<uint4*>(stage + row * 256 + (((ch & 8) | ((ch ^ row) & 7)) << 4)) = u;
                    }
                }
                // destination of my row: video rows go back to their original token (window reverse), text rows to the
                // per-window text buffer that svr2_txt_pool_fwd averages
                const int jrow = it.q0 + i * BM + row;
                unsigned long long dst = 0ull;
                if (jrow < it.w.n_vid) {
                    int lt, lh, lw;
                    dst = (unsigned long long)(p.out_vid + win_token(it.w, jrow, p.H, p.W, &lt, &lh, &lw) * row_stride + it.head * D);
                } else if (jrow < it.seq) {
                    dst = (unsigned long long)(p.txt_part + ((int64_t)it.wi * p.txt_len + (jrow - it.w.n_vid)) * row_stride + it.head * D);
                }
                __syncwarp();
                const int ch = lane & 15;
#pragma unroll 4
                for (int t = 0; t < 16; ++t) {
                    const int rl = 2 * t + (lane >> 4);                // row (inside this warp's 32) served by my half-warp
                    const unsigned long long d = __shfl_sync(0xffffffffu, dst, rl);
                    const int r = quad * 32 + rl;
                    const uint4 u = *reinterpret_cast<const uint4*>(stage + r * 256 + (((ch & 8) | ((ch ^ r) & 7)) << 4));
                    if (d != 0ull) *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(d) + ch * 8) = u;
                }
                __syncwarp();                                           // staging rows of this warp are free again
                if (threadIdx.x == 256) trace_item(p, n_it, 6 + i);
                if (i) ++c1; else ++c0;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) trace_at(p, 5, clock64());
    if (warp == 13) {
        __syncwarp();
        tmem_dealloc(tmem, TMEM_COLS);
    }
}

}  // namespace tc2
}  // namespace svr2

// ------------------------------------------------------------------------------------------------ host side
namespace svr2 {
extern long long* g_trace;
extern int g_trace_ctas;
int make_tmap(CUtensorMap* m, const void* ptr, int64_t P, int heads);

int launch_window_attn_tc2(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                           cudaStream_t st) {
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap(&tq, q, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tk, k, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tv, v, P, heads)) != SVR2_OK) return rc;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(tc2::window_attn_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             tc2::SMEM_BYTES);
        if (e != cudaSuccess) {
            set_error("svr2_window_attn_fwd: cannot reserve %d B shared memory: %s", tc2::SMEM_BYTES, cudaGetErrorString(e));
            return SVR2_E_CUDA;
        }
        attr_set = true;
    }
    tc2::Params p;
    p.tab = tab; p.work = work; p.n_items = n_work * heads;
    p.H = H; p.W = W; p.txt_len = txt_len; p.heads = heads;
    p.out_vid = (__nv_bfloat16*)out_vid; p.txt_part = (__nv_bfloat16*)txt_part;
    p.trace = g_trace; p.trace_ctas = g_trace_ctas;
    const int sms = sm_count();
    const int grid = p.n_items < sms ? p.n_items : sms;       // one persistent CTA per SM
    tc2::window_attn_tc2_kernel<<<grid, tc2::NTHREADS, tc2::SMEM_BYTES, st>>>(tq, tk, tv, p);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(tcgen05, persistent)");
    return SVR2_OK;
}

}  // namespace svr2
