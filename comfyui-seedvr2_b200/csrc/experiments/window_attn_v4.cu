// svr2_window_attn_fwd — PERSISTENT variable-length NaSwin window attention for sm_100a (TMA + tcgen05 + TMEM), generation 6:
// 64-column half tiles with DOUBLE-BUFFERED S.
//
// Generation 4 (window_attn_v2.cu) keeps S_i single-buffered, so a query tile's next Q K^T can only start after its P V has
// read P (aliased on S): softmax and MMAs of one tile strictly alternate, and with only two tiles in TMEM the tensor core
// idles half the time.  Q K^T of the NEXT key columns does not depend on the softmax of the current ones — it only needs a
// free S buffer.  512 TMEM columns hold O_0, O_1 (2 x 128) and FOUR 64-column S buffers (2 per query tile): the MMA warp keeps
// two half tiles of S queued per query tile, the softmax warps consume one while the tensor core fills the other.
// (This structure was first tried as "generation 3" and lost; that version still carried a printf in every barrier wait,
// whose local-memory traffic cost 600-900 cycles per softmax step, an inexact lagging row max, and a ring-parity race.)
//
// Replaces FlashAttentionVarlen / pytorch_varlen_attention (reference src/models/dit_3b/attention.py:24-61,138-157),
// unconcat_win (na.py:396-417) and window_reverse (na.py:638).
//
// Roles (15 warps): warps 0-3 softmax of query tile 0, 4-7 softmax of tile 1 (one thread per row, 64 columns per step,
// exact running max with lazy rescale, 3/8 of the exp2 on the FMA pipe), 8-11 epilogue, 12 TMA, 13 / 14 MMA issuers of
// tile 0 / 1 (warp-uniform control flow, one elected lane issues).
// TMEM map (columns): S_i[b] at 128 i + 64 b, O_i at 256 + 128 i;  P aliases the first 32 columns of its S buffer.
// SMEM map: Q0 | Q1 | ring[4] | staging | l_buf[2][2][128] | barriers.
#include <cuda.h>

#include "../common.cuh"
#include "../attn_launch.cuh"
#include "../sm100_ptx.cuh"
#include "../softmax_row.cuh"

namespace svr2 {
namespace tc4 {

using namespace ptx;
using namespace rowsm;

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
// Exponent reference of a row, in the log2 domain.  Default: the exact scaled row max (P <= 1, the rounding pattern of the
// reference's fp32 softmax).  SVR2_INT_REF=1 rounds it UP to an integer, which makes every lazy rescale an exact power of two;
// measured: it does NOT make the result independent of the tiling (the FFMA that forms s k - c rounds differently for a
// different c), and it moved three block outputs by one bf16 ulp against the golden vectors — kept off.
#ifndef SVR2_INT_REF
#define SVR2_INT_REF 0
#endif
__device__ __forceinline__ float ref_of(float row_max) {
    return SVR2_INT_REF ? ceilf(row_max * kScaleLog2) : row_max * kScaleLog2;
}
constexpr float kRescaleThreshold = 8.0f;   // lazy rescale: tolerate P up to 2^8 before touching O
constexpr int kFmaEighths = 3;              // of every 8 exp2 pairs, this many run on the FMA pipe (softmax_row.cuh)

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

// The timeline hooks exist only in the kTrace instantiation (launched while svr2_debug_attn_trace is armed).
template <bool kTrace>
__device__ __forceinline__ void trace_at(const Params& p, int slot, long long v) {
    if constexpr (kTrace) {
        if ((int)blockIdx.x < p.trace_ctas && slot < kTraceSlots) p.trace[(long long)blockIdx.x * kTraceSlots + slot] = v;
    }
}
template <bool kTrace>
__device__ __forceinline__ void trace_item(const Params& p, int n, int what) {
    if constexpr (kTrace) {
        if (n < kTraceItems) trace_at<kTrace>(p, 8 + 9 * n + what, clock64());
    }
}

// 15 warps = 4+4+4+3 per SM sub-partition; a sub-partition's 16 K registers / 4 warps caps a thread at 128 registers
template <bool kTrace>
__global__ void __launch_bounds__(NTHREADS, 1)
window_attn_tc4_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
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
        trace_at<kTrace>(p, 0, clock64());
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
            mbar_init(bar(BAR_KVEMPTY + s), 2);       // one commit per MMA warp
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
        if constexpr (kTrace) {
            unsigned smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            trace_at<kTrace>(p, 1, clock64());
            trace_at<kTrace>(p, 6, smid);
        }
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
                trace_item<kTrace>(p, n_it, 8);
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
                if (i >= it.nq) {
                    // Query tile 1 absent.  This warp still walks the item's K/V fills — waits for each one and releases it
                    // with an (empty) commit.  Skipping ahead instead would let it test a ring barrier two phases early,
                    // where the parity aliases and an unfinished fill reads as complete.
                    for (uint32_t n = 0; n < 2u * it.n_kv; ++n) {
                        mbar_wait(bar(BAR_KVFULL + ring_slot(n)), ring_par(n));
                        if (elect_one()) tc_commit(bar(BAR_KVEMPTY + ring_slot(n)));
                        __syncwarp();
                    }
                    kv0 += 2 * it.n_kv;
                    continue;
                }
                auto cols = [&](int h) { const int r = it.cols_total - h * HN; return r < HN ? r : HN; };
                auto release = [&](uint32_t idx) { tc_commit(bar(BAR_KVEMPTY + ring_slot(idx))); };
                mbar_wait(bar(BAR_QFULL + i), cq & 1);
                mbar_wait(bar(BAR_KVFULL + ring_slot(0)), ring_par(0));
                tc_fence_after();
                // two half tiles of S are queued before any P exists
                if (elect_one()) {
                    if (i == 0) trace_item<kTrace>(p, n_done, 0);
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
                }
                kv0 += 2 * it.n_kv;
                ++cq;
            }
            if (lane == 0 && i == 0) trace_at<kTrace>(p, 7, n_done);
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
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const Item it = decode_item(p, item);
            ++n_it;
            if (i >= it.nq) continue;
            // Bit-compatibility with generation 4 (window_attn_v2.cu), which works on whole 128-column key tiles: the exponent
            // reference of the first tile is the max over all of ITS columns (both halves), and a tile's row sum is formed as
            // ((c0 + c1) + c2) + c3 before it is added to l.  With the same reference every P, every P V accumulation and l
            // are the same numbers in the same order, so both generations return identical bits unless a lazy rescale
            // fires (they test for it at different granularity) — tests/test_gpu_kernels.py checks the identity.
            float m = -INFINITY, l = 0.f;       // exponent reference (log2 domain, raised lazily) and row sum
            float rs_even = 0.f;                // row sum of the even half of the current key tile, not yet in l
            for (int h = 0; h < it.n_half; ++h, ++g) {
                const int b = h & 1;
                const bool last = h == it.n_half - 1;
                const int rem = it.cols_total - h * HN;
                const bool two = rem > 32;                               // 64 or 32 key columns in this half tile
                const int valid = it.seq - h * HN;                       // key columns >= valid are beyond the sequence
                const uint32_t tS = tmem + lane_off + i * 128 + b * HN;
                if (h != 1) mbar_wait(bar(BAR_SFULL + 2 * i + b), (sf_par >> b) & 1u);      // S(1) was awaited at h == 0
                sf_par ^= 1u << b;
                tc_fence_after();
                if (kTrace && threadIdx.x == 0) {
                    if (h == 0) trace_item<kTrace>(p, n_it, 1);
                    if (last) trace_item<kTrace>(p, n_it, 3);
                }
                float m8[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) m8[q] = -INFINITY;
                if (h == 0 && it.n_half > 1) {
                    // first key tile: fold the row max of its second half (already queued behind the first) into the reference
                    mbar_wait(bar(BAR_SFULL + 2 * i + 1), (sf_par >> 1) & 1u);
                    tc_fence_after();
                    const int rem1 = it.cols_total - HN, valid1 = it.seq - HN;
                    uint32_t t[32];
                    tmem_ld32(tS + HN, t);
                    tc_wait_ld();
                    if (valid1 < 32) mask_chunk(t, valid1);
                    chunk_max8(m8, t);
                    if (rem1 > 32) {
                        tmem_ld32(tS + HN + 32, t);
                        tc_wait_ld();
                        if (valid1 < rem1) mask_chunk(t, valid1 - 32);
                        chunk_max8(m8, t);
                    }
                }
                uint32_t a[32], a2[32];
                tmem_ld32(tS, a);
                if (two) tmem_ld32(tS + 32, a2);
                tc_wait_ld();
                if (valid < rem) {                                       // only the last half tile of a sequence
                    if (valid < 32) mask_chunk(a, valid);
                    if (two) mask_chunk(a2, valid - 32);
                }
                chunk_max8(m8, a);
                if (two) chunk_max8(m8, a2);
                const float mx = fmaxf(max3(m8[0], m8[1], m8[2]), max3(max3(m8[3], m8[4], m8[5]), m8[6], m8[7]));
                if (h == 0) {
                    m = ref_of(mx);
                } else {
                    // lazy rescale: O_i and l keep the old reference until the row max has moved by more than 2^8
                    const bool need = mx * kScaleLog2 - m > kRescaleThreshold;
                    if (__any_sync(0xffffffffu, need)) {                  // warp-uniform: tcgen05.ld/st are warp collectives
                        // rare.  O_i may still be receiving P(h-1) V(h-1) — S(h) was queued before it: wait for that MMA
                        // (the g-th P V of this query tile; it cannot be more than one phase away)
                        mbar_wait(bar(BAR_PVDONE + i), (g - 1) & 1u);
                        tc_fence_after();
                        const float m_new = need ? ref_of(mx) : m;
                        const float alpha = ex2(m - m_new);
                        l *= alpha;
                        rs_even *= alpha;
                        m = m_new;
#pragma unroll 1
                        for (int c8 = 0; c8 < 8; ++c8) {
                            uint32_t ov[16];
                            tmem_ld16(tO + c8 * 16, ov);
                            tc_wait_ld();
#pragma unroll
                            for (int c = 0; c < 16; ++c) ov[c] = __float_as_uint(__uint_as_float(ov[c]) * alpha);
                            tmem_st16(tO + c8 * 16, ov);
                        }
                    }
                }
                const float mb = m;
                float rs = b ? rs_even : 0.f;                            // ((c0 + c1) + c2) + c3 per key tile
                {
                    uint32_t pk[16];
                    rs += chunk_exp_mix<kFmaEighths>(a, pk, mb);
                    tmem_st16(tS, pk);
                }
                if (two) {
                    uint32_t pk[16];
                    rs += chunk_exp_mix<kFmaEighths>(a2, pk, mb);
                    tmem_st16(tS + 16, pk);
                }
                if (b == 0 && !last) {
                    rs_even = rs;
                } else {
                    l += rs;
                    rs_even = 0.f;
                }
                tc_wait_st();
                tc_fence_before();
                mbar_arrive(bar(BAR_PREADY + 2 * i + b));
                if (kTrace && threadIdx.x == 0) {
                    if (h == 0) trace_item<kTrace>(p, n_it, 2);
                    if (last) trace_item<kTrace>(p, n_it, 4);
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
                if (threadIdx.x == 256 && i == 0) trace_item<kTrace>(p, n_it, 5);
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
                if (threadIdx.x == 256) trace_item<kTrace>(p, n_it, 6 + i);
                if (i) ++c1; else ++c0;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) trace_at<kTrace>(p, 5, clock64());
    if (warp == 13) {
        __syncwarp();
        tmem_dealloc(tmem, TMEM_COLS);
    }
}

}  // namespace tc4
}  // namespace svr2

// ------------------------------------------------------------------------------------------------ host side
namespace svr2 {

int launch_window_attn_tc4(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                           cudaStream_t st) {
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap(&tq, q, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tk, k, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tv, v, P, heads)) != SVR2_OK) return rc;
    static PerDeviceFlag smem_ok[2];
    if ((rc = ensure_dyn_smem(tc4::window_attn_tc4_kernel<false>, tc4::SMEM_BYTES, smem_ok[0], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    if ((rc = ensure_dyn_smem(tc4::window_attn_tc4_kernel<true>, tc4::SMEM_BYTES, smem_ok[1], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    tc4::Params p;
    p.tab = tab; p.work = work; p.n_items = n_work * heads;
    p.H = H; p.W = W; p.txt_len = txt_len; p.heads = heads;
    p.out_vid = (__nv_bfloat16*)out_vid; p.txt_part = (__nv_bfloat16*)txt_part;
    p.trace = g_trace; p.trace_ctas = g_trace_ctas;
    const int sms = sm_count();
    const int grid = p.n_items < sms ? p.n_items : sms;       // one persistent CTA per SM
    if (p.trace != nullptr) tc4::window_attn_tc4_kernel<true><<<grid, tc4::NTHREADS, tc4::SMEM_BYTES, st>>>(tq, tk, tv, p);
    else tc4::window_attn_tc4_kernel<false><<<grid, tc4::NTHREADS, tc4::SMEM_BYTES, st>>>(tq, tk, tv, p);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(tcgen05, persistent)");
    return SVR2_OK;
}

}  // namespace svr2
