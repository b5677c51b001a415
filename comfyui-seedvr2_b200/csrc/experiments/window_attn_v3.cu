// svr2_window_attn_fwd — PERSISTENT variable-length NaSwin window attention for sm_100a (TMA + tcgen05 + TMEM), generation 5.
//
// Replaces FlashAttentionVarlen / pytorch_varlen_attention (reference src/models/dit_3b/attention.py:24-61,138-157: a
// Python loop of one SDPA call per window), unconcat_win (na.py:396-417) and window_reverse (na.py:638).
//
// Generation 4 (window_attn_v2.cu, kept as SVR2_ATTN_IMPL=v2) gives each query tile its own softmax warpgroup.  Measured
// there (profiles/r01_attn_g4_*): a softmax step of one tile takes ~2000 cycles for a lone warp per sub-partition (exp2
// unit + single-warp issue latency), the tile's MMAs another ~1100 + ~400 of hand-over latency, and with only two tiles in
// TMEM the tensor core idles half the time whatever the scheduling between the two groups.
// Generation 5 halves the latency of a softmax step instead: BOTH softmax warpgroups work on the SAME tile, each on one
// 64-column half of the S row (two warps per sub-partition share a row's exp2 work and hide each other's latencies), and
// they alternate between the two query tiles.  The two halves agree on the row max through shared memory (one 64-thread
// named barrier per row quadrant), keep partial row sums that the epilogue adds, and split the rare rescale of O by column.
//
// Work item = (window, pair of 128-row query tiles, head); items are sorted longest sequence first and CTA b takes
// items b, b+G, b+2G, ...   Roles (16 warps, 512 threads, 128 registers each):
//   warp  0      TMA producer (one thread): Q tiles, then K/V through a 4-stage ring of 32-KB tiles, across items
//   warps 1,2    tcgen05.mma issuers, one per query tile: S_i = Q_i K^T (SS), O_i += P_i V (A = P from TMEM, B = V
//                MN-major); warp-uniform control flow, one elected lane issues; warp 3 idle
//   warps 4-7    epilogue: O/(l_a + l_b) -> bf16 -> swizzled 16-KB staging half tile -> coalesced 128-B row segments to the
//                token's ORIGINAL position (window reverse) or the per-window text buffer
//   warps 8-11   softmax half A: key columns [0,64) of the current tile's S row, one thread per query row
//   warps 12-15  softmax half B: key columns [64,128)
// Per key tile j the softmax warps process tile 0 then tile 1; while they are on one tile the tensor core runs the other
// tile's P V and next Q K^T.  exp2: 3 of every 8 on the FMA pipe (softmax_row.cuh).  The last key tile of a sequence is
// trimmed to a multiple of 32 columns (MMA N / K steps and the softmax shrink with it).
//
// TMEM map (columns): S0 [0,128)  S1 [128,256)  O0 [256,384)  O1 [384,512);  P_i aliases the first 64 columns of S_i.
// SMEM map: Q0 | Q1 | ring[4] (32 KB each) | staging (16 KB) | row sums l[2][2][2][128] | row max exchange [2][2][128] | mbarriers.
#include <cuda.h>

#include "../common.cuh"
#include "../attn_launch.cuh"
#include "../sm100_ptx.cuh"
#include "../softmax_row.cuh"

namespace svr2 {
namespace tc3 {

using namespace ptx;
using namespace rowsm;

constexpr int D = SVR2_HEAD_DIM;          // 128
constexpr int BM = 128, BN = 128;
constexpr int TILE_BYTES = BM * D * 2;    // 32768
constexpr int BOX_BYTES = TILE_BYTES / 2; // 128 rows x 64 cols bf16
constexpr int STAGES = 4;
constexpr int NTHREADS = 512;
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_STAGE = OFF_RING + STAGES * TILE_BYTES;
constexpr int STAGE_BYTES = TILE_BYTES / 2;            // 128 rows x 64 columns bf16
constexpr int OFF_L = OFF_STAGE + STAGE_BYTES;         // float l[2 item parities][2 tiles][2 halves][128 rows]
constexpr int OFF_XM = OFF_L + 2 * 2 * 2 * 128 * 4;    // float row-max exchange [2 step parities][2 halves][128 rows]
constexpr int OFF_BAR = OFF_XM + 2 * 2 * 128 * 4;
constexpr int SMEM_BYTES = OFF_BAR + 256;
constexpr uint32_t TMEM_COLS = 512;
#ifndef SVR2_FMA8
#define SVR2_FMA8 3
#endif
constexpr int kFmaEighths = SVR2_FMA8;              // of every 8 exp2 pairs, this many are evaluated on the FMA pipe (softmax_row.cuh)
constexpr float kRescaleThreshold = 8.0f;   // lazy rescale: tolerate P up to 2^8 before touching O

// barrier slots (8 bytes each), per query tile i: slot + i
enum { BAR_QFULL = 0, BAR_QEMPTY = 2, BAR_SFULL = 4, BAR_PREADY = 6, BAR_OFINAL = 8, BAR_OEMPTY = 10, BAR_LREADY = 12,
       BAR_KVFULL = 14, BAR_KVEMPTY = 14 + STAGES, BAR_COUNT = 14 + 2 * STAGES };
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
// +0 MMA warp 0 starts the item (Q and K0 landed)   +1 softmax-0 sees S(0)   +2 softmax-0 published P(0)
// +3 softmax-0 sees S(last)   +4 softmax-0 published P(last)   +5 epilogue sees O_0 final   +6 tile-0 stores issued
// +7 tile-1 stores issued   +8 TMA thread issued the item's last K/V load
constexpr int kTraceSlots = 64, kTraceItems = 4;
// step stamps for the CTA's third item, key tiles j < 4: softmax group i saw S(j) at 44 + 8i + 2j, published P(j) at 45 + 8i + 2j;
// MMA warp i finished issuing the step-j MMAs at 60 + 2i (j = 1 only: slots are scarce)

// Everything a role needs to know about one work item; every role decodes it identically.
struct Item {
    Win w;
    int wi, head, seq, q0, nq, n_kv, last_cols;
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
    it.n_kv = (it.seq + BN - 1) / BN;
    const int rem = it.seq - (it.n_kv - 1) * BN;       // 1..128 valid key columns in the last tile
    it.last_cols = (rem + 31) & ~31;                   // trimmed to a multiple of 32 (MMA N, softmax chunks, PV K steps)
    return it;
}

#ifndef SVR2_EXP
#define SVR2_EXP 0
#endif
// one lane polls the barrier, the rest of the warp parks at the warp barrier (no issue slots burnt by 31 spinning lanes)
__device__ __forceinline__ void warp_wait(uint32_t bar, uint32_t parity) {
    if ((threadIdx.x & 31) == 0) mbar_wait(bar, parity);
    __syncwarp();
}
#define SOFTMAX_WAIT(b, par) mbar_wait(b, par)
#define OTHER_WAIT(b, par) mbar_wait(b, par)


// The timeline hooks exist only in the kTrace instantiation (launched while svr2_debug_attn_trace is armed): the
// production kernel carries none of their code — the roles' instruction streams have to share a 32-KB instruction cache.
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

template <bool kTrace>
__global__ void __launch_bounds__(NTHREADS, 1)
window_attn_tc3_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
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
            mbar_init(bar(BAR_SFULL + i), 1);
            mbar_init(bar(BAR_PREADY + i), 256);      // both softmax halves
            mbar_init(bar(BAR_OFINAL + i), 1);
            mbar_init(bar(BAR_OEMPTY + i), 128);
            mbar_init(bar(BAR_LREADY + i), 256);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_KVFULL + s), 1);
            mbar_init(bar(BAR_KVEMPTY + s), 2);       // one commit per MMA warp
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
    if (warp == 0 && lane == 0) {
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

    if (warp < 4) {
        if (warp == 0) {
            // -------------------------------------------------------------- TMA producer (runs ahead across items)
            if (lane == 0) {
                uint32_t kv_n = 0, c0 = 0, c1 = 0;
                int n_it = 0;
                for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++n_it) {
                    const Item it = decode_item(p, item);
                    for (int i = 0; i < it.nq; ++i) {
                        const int row = it.w.row_start + it.q0 + i * BM;
                        const uint32_t c = i ? c1 : c0;
                        mbar_wait(bar(BAR_QEMPTY + i), (c & 1) ^ 1);          // all QK MMAs of the previous user are done
                        mbar_expect_tx(bar(BAR_QFULL + i), TILE_BYTES);
                        tma_load_3d(base + i * TILE_BYTES, &tm_q, bar(BAR_QFULL + i), 0, it.head, row);
                        tma_load_3d(base + i * TILE_BYTES + BOX_BYTES, &tm_q, bar(BAR_QFULL + i), 64, it.head, row);
                        if (i) ++c1; else ++c0;
                    }
                    for (int n = 0; n < 2 * it.n_kv; ++n, ++kv_n) {           // ring order K0 V0 K1 V1 ...
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
        } else if (warp <= 2) {
            // -------------------------------------------------------------- tcgen05.mma issuer for query tile i
            // The WHOLE warp runs this loop: warp-uniform control flow keeps the descriptors in uniform registers (a
            // lane-0-only branch made the compiler serialise every UTCHMMA operand through a per-lane waterfall loop,
            // ~90 cycles per MMA); one elected lane issues the MMAs and commits.
            constexpr uint32_t idesc_pv = idesc_bf16(BM, D, 1);
            constexpr uint32_t hi_sw = smem_desc_hi_sw128(1024);   // 8-row (K-major Q, K) / 8-key (MN-major V) groups 1024 B apart
            const int i = warp - 1;
            // S_i = Q_i K^T : 8 x (128 x ncols x 16)
            auto issue_qk = [&](uint32_t kbase, int ncols) {
                const uint32_t idesc_qk = idesc_bf16(BM, ncols, 0);
                const uint32_t qlo = smem_desc_lo(base + i * TILE_BYTES, 16);
                const uint32_t klo = smem_desc_lo(kbase, 16);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t off = ((uint32_t)(kk >> 2) * BOX_BYTES + (uint32_t)(kk & 3) * 32u) >> 4;
                    mma_ss(tmem + i * 128, make_desc(qlo + off, hi_sw), make_desc(klo + off, hi_sw), idesc_qk, kk > 0);
                }
            };
            // O_i += P_i V : A = P from TMEM (16 keys = 8 columns per step), B = V MN-major (16 keys = 2048 B; LBO = next 64 d)
            auto issue_pv = [&](uint32_t vbase, int ksteps, bool first) {
                const uint32_t vlo = smem_desc_lo(vbase, BOX_BYTES);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk)
                    if (kk < ksteps)
                        mma_ts(tmem + 256 + i * 128, tmem + i * 128 + kk * 8, make_desc(vlo + kk * (2048 >> 4), hi_sw),
                               idesc_pv, !(first && kk == 0));
            };
            uint32_t kv0 = 0;            // ring index of the current item's K tile 0 (K_t at 2t, V_t at 2t+1)
            uint32_t cq = 0;             // items in which this query tile took part
            uint32_t pr = 0;             // PREADY_i phases consumed
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
                        OTHER_WAIT(bar(BAR_KVFULL + ring_slot(n)), ring_par(n));
                        if (elect_one()) tc_commit(bar(BAR_KVEMPTY + ring_slot(n)));
                        __syncwarp();
                    }
                    kv0 += 2 * it.n_kv;
                    continue;
                }
                auto cols = [&](int j) { return j == it.n_kv - 1 ? it.last_cols : BN; };
                auto release = [&](uint32_t idx) { tc_commit(bar(BAR_KVEMPTY + ring_slot(idx))); };
                OTHER_WAIT(bar(BAR_QFULL + i), cq & 1);
                OTHER_WAIT(bar(BAR_KVFULL + ring_slot(0)), ring_par(0));
                tc_fence_after();
                if (elect_one()) {
                    if (i == 0) trace_item<kTrace>(p, n_done, 0);
                    issue_qk(sRing + ring_slot(0) * TILE_BYTES, cols(0));
                    tc_commit(bar(BAR_SFULL + i));
                    if (it.n_kv == 1) tc_commit(bar(BAR_QEMPTY + i));                    // that was the last read of Q_i
                    release(0);
                }
                __syncwarp();
                for (int j = 0; j < it.n_kv; ++j) {
                    const bool more = j + 1 < it.n_kv;
                    OTHER_WAIT(bar(BAR_KVFULL + ring_slot(2 * j + 1)), ring_par(2 * j + 1));              // V_j
                    if (more) OTHER_WAIT(bar(BAR_KVFULL + ring_slot(2 * j + 2)), ring_par(2 * j + 2));    // K_{j+1}
                    OTHER_WAIT(bar(BAR_PREADY + i), pr & 1u);
                    ++pr;
                    if (j == 0) OTHER_WAIT(bar(BAR_OEMPTY + i), (cq & 1) ^ 1);            // epilogue drained O_i of the previous item
                    tc_fence_after();
                    if (elect_one()) {
                        issue_pv(sRing + ring_slot(2 * j + 1) * TILE_BYTES, cols(j) >> 4, j == 0);
                        if (more) {
                            // the tensor core runs MMAs in issue order: S_i (and the P_i aliased on it) is overwritten
                            // only after P_i V has read it
                            issue_qk(sRing + ring_slot(2 * j + 2) * TILE_BYTES, cols(j + 1));
                            tc_commit(bar(BAR_SFULL + i));
                            if (j + 2 == it.n_kv) tc_commit(bar(BAR_QEMPTY + i));
                        } else {
                            tc_commit(bar(BAR_OFINAL + i));
                        }
                        release(2 * j + 1);
                        if (more) release(2 * j + 2);
                    }
                    __syncwarp();
                }
                kv0 += 2 * it.n_kv;
                ++cq;
            }
            if (lane == 0 && i == 0) trace_at<kTrace>(p, 7, n_done);
        }
    } else if (warp < 8) {
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
                OTHER_WAIT(bar(BAR_LREADY + i), cc & 1);
                OTHER_WAIT(bar(BAR_OFINAL + i), cc & 1);
                tc_fence_after();
                if (threadIdx.x == 128 && i == 0) trace_item<kTrace>(p, n_it, 5);
                const float* lp = l_buf + (((cc & 1) * 2 + i) * 2) * 128 + row;
                const float inv = 1.0f / (lp[0] + lp[128]);                // the two column halves' partial row sums
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
#pragma unroll 1
                for (int half = 0; half < 2; ++half) {                      // 64 output columns = 128-B row segments
#pragma unroll
                    for (int c2 = 0; c2 < 2; ++c2) {
                        uint32_t ov[32];
                        tmem_ld32(tO + half * 64 + c2 * 32, ov);
                        tc_wait_ld();
                        if (half == 1 && c2 == 1) {                         // O_i is in registers: the next item may overwrite it
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
                            const int ch = c2 * 4 + q;                      // 16-B chunk of the 128-B row segment, XOR-swizzled
                            *reinterpret_cast<uint4*>(stage + row * 128 + (((ch ^ row) & 7) << 4)) = u;
                        }
                    }
                    __syncwarp();
                    const int ch = lane & 7;
#pragma unroll 4
                    for (int t = 0; t < 8; ++t) {
                        const int rl = 4 * t + (lane >> 3);                 // row (inside this warp's 32) served by my quarter-warp
                        const unsigned long long d = __shfl_sync(0xffffffffu, dst, rl);
                        const int r = quad * 32 + rl;
                        const uint4 u = *reinterpret_cast<const uint4*>(stage + r * 128 + (((ch ^ r) & 7) << 4));
                        if (d != 0ull) *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(d) + half * 64 + ch * 8) = u;
                    }
                    __syncwarp();                                           // staging rows of this warp are free again
                }
                if (threadIdx.x == 128) trace_item<kTrace>(p, n_it, 6 + i);
                if (i) ++c1; else ++c0;
            }
        }
    } else {
        // ------------------------------------------------------------------ softmax: half g of the key columns, one thread per row
        const int g = (warp - 8) >> 2, quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        float* xm = reinterpret_cast<float*>(smem + OFF_XM);
        const bool tracer = kTrace && (threadIdx.x & 127) == 0;
        uint32_t sf0 = 0, sf1 = 0;  // SFULL_i phases consumed
        uint32_t step = 0;          // softmax steps done (parity selects the row-max exchange buffer)
        uint32_t c0 = 0, c1 = 0;    // items in which tile i took part
        int n_it = -1;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const Item it = decode_item(p, item);
            ++n_it;
            float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;      // per query tile: exponent reference, partial row sum
            for (int j = 0; j < it.n_kv; ++j) {
                const bool last = j == it.n_kv - 1;
                const int cols = last ? it.last_cols : BN;                  // key columns of this tile (multiple of 32)
                int nch = (cols >> 5) - 2 * g;                              // my 32-column chunks: half A has the first two
                nch = nch < 0 ? 0 : (nch > 2 ? 2 : nch);
                const int valid = it.seq - j * BN - 64 * g;                 // my columns >= valid are beyond the sequence
#pragma unroll 1
                for (int i = 0; i < it.nq; ++i, ++step) {
                    const uint32_t tS = tmem + lane_off + i * 128 + 64 * g;  // my half of S_i
                    const uint32_t tP = tmem + lane_off + i * 128 + 32 * g;  // my half of P_i (bf16 pairs)
                    const uint32_t tO = tmem + lane_off + 256 + i * 128 + 64 * g;   // my half of O_i's columns (rescale only)
                    float m = i ? m1 : m0, l = i ? l1 : l0;
                    const bool fine = kTrace && (threadIdx.x & 127) == 0 && n_it == 2 && j < 4;
                    mbar_wait(bar(BAR_SFULL + i), (i ? sf1 : sf0) & 1u);
                    if (i) ++sf1; else ++sf0;
                    tc_fence_after();
                    if (fine) trace_at<kTrace>(p, 44 + 8 * i + 2 * j, clock64());
                    if (tracer && g == 0 && i == 0) {
                        if (j == 0) trace_item<kTrace>(p, n_it, 1);
                        if (last) trace_item<kTrace>(p, n_it, 3);
                    }
                    // Register block 1 holds my LAST chunk (the one that may contain the end of the sequence), block 0 the
                    // one before it, so the masking code exists once.
                    uint32_t s[64];
                    if (nch > 1) tmem_ld32(tS, s);
                    if (nch > 0) tmem_ld32(tS + 32 * (nch - 1), s + 32);
                    tc_wait_ld();
                    float mx = -INFINITY;
                    if (nch > 0) {
                        if (valid < nch * 32) mask_chunk(s + 32, valid - 32 * (nch - 1));      // only in the last key tile
                        float m8[8];
#pragma unroll
                        for (int q = 0; q < 8; ++q) m8[q] = -INFINITY;
                        if (nch > 1) chunk_max8(m8, s);
                        chunk_max8(m8, s + 32);
                        mx = fmaxf(max3(m8[0], m8[1], m8[2]), max3(max3(m8[3], m8[4], m8[5]), m8[6], m8[7]));
                    }
                    // the two halves of the row agree on its max through shared memory
                    float* xs = xm + (step & 1u) * 256;
                    xs[g * 128 + row] = mx;
                    named_bar_sync(1 + quad, 64);
                    mx = fmaxf(mx, xs[(g ^ 1) * 128 + row]);
                    if (j == 0) {
                        m = mx;
                    } else {
                        // lazy rescale: O_i and l keep the old reference until the row max has moved by more than 2^8
                        const bool need = (mx - m) * kScaleLog2 > kRescaleThreshold;
                        if (__any_sync(0xffffffffu, need)) {                // same decision in both halves (same m, same mx)
                            // S_i(j) was committed after P_i V(j-1): O_i is quiescent until P(j) is published
                            const float m_new = need ? mx : m;
                            const float alpha = ex2((m - m_new) * kScaleLog2);
                            l *= alpha;
                            m = m_new;
#pragma unroll 1
                            for (int c8 = 0; c8 < 4; ++c8) {                 // my half of O_i's 128 columns
                                uint32_t ov[16];
                                tmem_ld16(tO + c8 * 16, ov);
                                tc_wait_ld();
#pragma unroll
                                for (int c = 0; c < 16; ++c) ov[c] = __float_as_uint(__uint_as_float(ov[c]) * alpha);
                                tmem_st16(tO + c8 * 16, ov);
                            }
                        }
                    }
                    const float mb = m * kScaleLog2;
                    // bf16 P overwrites the front of S_i.  P columns [0,32) alias S columns [0,32) of half A, [32,64) alias
                    // half A's second chunk: half B may store only after half A has loaded its S columns.  Half A's loads
                    // precede its arrival at the named barrier above, half B's stores follow it.
                    float rs = 0.f;
                    if (nch > 1) {
                        uint32_t pk[16];
                        rs = chunk_exp_mix<kFmaEighths>(s, pk, mb);
                        tmem_st16(tP, pk);
                    }
                    if (nch > 0) {
                        uint32_t pk[16];
                        rs += chunk_exp_mix<kFmaEighths>(s + 32, pk, mb);
                        tmem_st16(tP + 16 * (nch - 1), pk);
                    }
                    l += rs;
                    tc_wait_st();
                    tc_fence_before();
                    mbar_arrive(bar(BAR_PREADY + i));
                    if (fine) trace_at<kTrace>(p, 45 + 8 * i + 2 * j, clock64());
                    if (tracer && g == 0 && i == 0) {
                        if (j == 0) trace_item<kTrace>(p, n_it, 2);
                        if (last) trace_item<kTrace>(p, n_it, 4);
                    }
                    if (i) { m1 = m; l1 = l; } else { m0 = m; l0 = l; }
                }
            }
            // publish the partial row sums for the epilogue warps (double-buffered by the tile's item parity)
            l_buf[(((c0 & 1) * 2 + 0) * 2 + g) * 128 + row] = l0;
            mbar_arrive(bar(BAR_LREADY + 0));
            ++c0;
            if (it.nq == 2) {
                l_buf[(((c1 & 1) * 2 + 1) * 2 + g) * 128 + row] = l1;
                mbar_arrive(bar(BAR_LREADY + 1));
                ++c1;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) trace_at<kTrace>(p, 5, clock64());
    if (warp == 1) {
        __syncwarp();
        tmem_dealloc(tmem, TMEM_COLS);
    }
}

}  // namespace tc3
}  // namespace svr2

// ------------------------------------------------------------------------------------------------ host side
namespace svr2 {

int launch_window_attn_tc3(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                           cudaStream_t st) {
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap(&tq, q, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tk, k, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tv, v, P, heads)) != SVR2_OK) return rc;
    static PerDeviceFlag smem_ok[2];
    if ((rc = ensure_dyn_smem(tc3::window_attn_tc3_kernel<false>, tc3::SMEM_BYTES, smem_ok[0], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    if ((rc = ensure_dyn_smem(tc3::window_attn_tc3_kernel<true>, tc3::SMEM_BYTES, smem_ok[1], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    tc3::Params p;
    p.tab = tab; p.work = work; p.n_items = n_work * heads;
    p.H = H; p.W = W; p.txt_len = txt_len; p.heads = heads;
    p.out_vid = (__nv_bfloat16*)out_vid; p.txt_part = (__nv_bfloat16*)txt_part;
    p.trace = g_trace; p.trace_ctas = g_trace_ctas;
    const int sms = sm_count();
    const int grid = p.n_items < sms ? p.n_items : sms;       // one persistent CTA per SM
    if (p.trace != nullptr) tc3::window_attn_tc3_kernel<true><<<grid, tc3::NTHREADS, tc3::SMEM_BYTES, st>>>(tq, tk, tv, p);
    else tc3::window_attn_tc3_kernel<false><<<grid, tc3::NTHREADS, tc3::SMEM_BYTES, st>>>(tq, tk, tv, p);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(tcgen05, persistent)");
    return SVR2_OK;
}

}  // namespace svr2
