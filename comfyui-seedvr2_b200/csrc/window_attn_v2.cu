// svr2_window_attn_fwd — PERSISTENT variable-length NaSwin window attention for sm_100a (TMA + tcgen05 + TMEM).
//
// Replaces FlashAttentionVarlen / pytorch_varlen_attention (reference src/models/dit_3b/attention.py:24-61,138-157: a
// Python loop of one SDPA call per window), unconcat_win (na.py:396-417) and window_reverse (na.py:638).
//
// Design history (measurements under profiles/):
//  gen 1  one CTA per work item: with NaSwin's short sequences (4 key tiles per window at 2048^2) 40 % of a CTA's life
//         was exposed prologue / epilogue / launch gap                                          -> 285 TFLOP/s at C1
//  gen 2  persistent CTA per SM, roles run ahead across items                                    -> 476 TFLOP/s at C1
//  gen 3  experiments (profiles/r01_microbench_*.txt, r01_attn_*timeline*.txt) that shaped this version:
//         * tcgen05.ld moves ~1 KB/clk/SM and neither it nor MUFU interferes with running MMAs — TMEM is not a limit;
//         * a lane-0-only MMA branch makes the compiler waterfall every UTCHMMA operand (~90 cycles per MMA issued, the
//           real bottleneck of gen 2): the whole MMA warp now runs warp-uniform control flow and one ELECTed lane
//           issues, with descriptors advanced by adding to their low word; one MMA warp per query tile;
//         * a softmax warp is latency bound, not pipe bound (issue slots 32 % used): every TMEM round trip, barrier and
//           fence costs 100-250 cycles and nothing else runs in that warp meanwhile.  Half tiles (64 columns) doubled
//           those fixed costs per column and were slower; the fix is the opposite — ONE round trip per 128 columns
//           with the whole S row in registers (setmaxnreg gives the softmax warps 200 registers).
//
// Work item = (window, pair of 128-row query tiles, head); items are sorted longest sequence first and CTA b takes
// items b, b+G, b+2G, ...   Roles (16 warps, 512 threads, registers re-balanced with setmaxnreg):
//   warp  0      TMA producer (one thread): Q tiles, then K/V through a 4-stage ring of 32-KB tiles, across items
//   warps 1,2    tcgen05.mma issuers, one per query tile: S_i = Q_i K^T (SS), O_i += P_i V (A = P from TMEM, B = V
//                MN-major); warp 3 idle                                                           (56 registers)
//   warps 4-7    epilogue: O/l -> bf16 -> swizzled staging tile -> coalesced 256-B row stores to the token's ORIGINAL
//                position (window reverse) or the per-window text buffer                           (56 registers)
//   warps 8-11   softmax group 0 (query tile 0): one thread per query row, whole 128-column S row in registers, exact
//                row max, lazy rescale of O (only when the max moved by more than 2^8), bf16 P written back over S
//   warps 12-15  softmax group 1 (query tile 1)                                                   (200 registers)
// The two query tiles ping-pong: while one group runs its exp2 burst the tensor core works for the other tile.
// The last key tile of a sequence is trimmed to a multiple of 32 columns (MMA N / K steps and the softmax shrink with it).
//
// Generation 10 = the same kernel instantiated with HALVES = 2 (24 warps): TWO threads per query row, each owning 64 of the key
// tile's 128 columns (softmax warps 8-15 take columns [0,64), warps 16-23 columns [64,128) of the same rows).  Every generation
// up to 9 ran two softmax warps per scheduler, each a long dependent chain (~770 warp-instructions per step at IPC 0.25 —
// latency bound, run 54), and a third warp could not fit in registers next to two 128-column rows.  A half row needs ~100
// registers, so four softmax warps share a scheduler and hide each other's latencies.  The two halves of a row exchange their
// partial row max through shared memory (one 64-thread named barrier per step); their partial row sums are only added by the
// epilogue.  P(j) written by the upper half aliases S columns the lower half has already loaded (the exchange barrier orders it).
//
// TMEM map (columns): S0 [0,128)  S1 [128,256)  O0 [256,384)  O1 [384,512);  P_i aliases the first 64 columns of S_i.
// SMEM map: Q0 | Q1 | ring[4] | staging (32 KB each) | row sums l[2][2][128] | mbarriers.
#include <cuda.h>
#include <cstdlib>

#include "common.cuh"
#include "attn_launch.cuh"
#include "sm100_ptx.cuh"
#include "softmax_row.cuh"

namespace svr2 {
namespace tc2 {

using namespace ptx;
using namespace rowsm;

constexpr int D = SVR2_HEAD_DIM;          // 128
constexpr int BM = 128, BN = 128;
constexpr int TILE_BYTES = BM * D * 2;    // 32768
constexpr int BOX_BYTES = TILE_BYTES / 2; // 128 rows x 64 cols bf16
constexpr int STAGES = 4;
constexpr int NTHREADS = 512;             // generation 4; generation 10 adds 8 softmax warps (768)
#ifndef SVR2_REGS_C      // experiment builds: -DSVR2_REGS_C=.. -DSVR2_REGS_E=.. -DSVR2_REGS_S=..  (nvcc splits -D values at commas)
#define SVR2_REGS_C 56     // run 60/61: 56/56/200 (no spills) is 2.6 % faster at C2 than 72/72/184; 48/64/200 and 40/72/200 spill and lose
#define SVR2_REGS_E 56
#define SVR2_REGS_S 200
#endif
constexpr int kRegs[3] = {SVR2_REGS_C, SVR2_REGS_E, SVR2_REGS_S};
// per sub-partition: 32 * (C + E + 2 S) = 16384
constexpr int kRegs10[3] = {48, 48, 96};     // generation 10: the CTA is launched with 80 x 768 registers; 128 (C + E) + 512 S must fit in them
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_STAGE = OFF_RING + STAGES * TILE_BYTES;
constexpr int STAGE_BYTES = BM * 128;                  // epilogue staging: 128 rows x 64 columns, two passes per tile
constexpr int OFF_L = OFF_STAGE + STAGE_BYTES;         // float l[2 parities][2 tiles][2 halves][128 rows]
constexpr int OFF_XCH = OFF_L + 2 * 2 * 2 * 128 * 4;   // float partial row max [2 tiles][2 halves][128 rows] (generation 10)
constexpr int OFF_BAR = OFF_XCH + 2 * 2 * 128 * 4;
constexpr int SMEM_BYTES = OFF_BAR + 256;              // 219,392 B of the 232,448 B a CTA may opt in to
constexpr uint32_t TMEM_COLS = 512;
#ifndef SVR2_FMA8
#define SVR2_FMA8 3
#endif
constexpr int kFmaEighths = SVR2_FMA8;              // of every 8 exp2 pairs, this many are evaluated on the FMA pipe (softmax_row.cuh)
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

// barrier slots (8 bytes each), per query tile i: slot + i
enum { BAR_QFULL = 0, BAR_QEMPTY = 2, BAR_SFULL = 4, BAR_PREADY = 6, BAR_OFINAL = 8, BAR_OEMPTY = 10, BAR_LREADY = 12,
       BAR_SLOADED = 14, BAR_KVFULL = 16, BAR_KVEMPTY = 16 + STAGES, BAR_COUNT = 16 + 2 * STAGES };
// SVR2_V2_SPLIT_QK=1: Q K^T(j+1) is issued in two halves.  Key columns 64..127 of S_i do not alias P_i(j) (columns 0..63), so
// their MMAs go out as soon as softmax(j) has the S row in registers (BAR_SLOADED) and run under its exp phase; only the lower
// half has to follow P_i V(j).  The chain between "P is ready" and "S is full" shrinks from 16 to 12 MMAs.
#ifndef SVR2_V2_SPLIT_QK
#define SVR2_V2_SPLIT_QK 0
#endif
#ifndef SVR2_TURNS
#define SVR2_TURNS 0      // measured: the bounded counter version gains nothing (see the comment in the softmax role); kept for experiments
#endif
#ifndef SVR2_TURN_BOUND
#define SVR2_TURN_BOUND 4096      // cycles a group waits for its turn at most
#endif
#ifndef SVR2_TURN_SLEEP
#define SVR2_TURN_SLEEP 20        // ns between polls of the turn counter
#endif
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
    int turns;             // softmax groups take turns (short sequences) or run freely (long ones)
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
#ifdef SVR2_DEBUG_WAIT
#define DBG_PROGRESS(role, val) do { if ((threadIdx.x & 31) == 0) reinterpret_cast<volatile uint32_t*>(smem + OFF_BAR + 200)[role] = (uint32_t)(val); } while (0)
#else
#define DBG_PROGRESS(role, val) do { } while (0)
#endif
#define SOFTMAX_WAIT(b, par) mbar_wait(b, par)
#define OTHER_WAIT(b, par) mbar_wait(b, par)

template <int N> __device__ __forceinline__ void reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

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

template <bool kTrace, int HALVES>
__global__ void __launch_bounds__(NTHREADS + 256 * (HALVES - 1), 1)
window_attn_tc2_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t base = smem_u32(smem);                          // SWIZZLE_128B tiles need 1024-B alignment
    const uint32_t sRing = base + OFF_RING;
    const uint32_t sBar = base + OFF_BAR;
    constexpr int REGS_CTRL = HALVES == 2 ? kRegs10[0] : kRegs[0], REGS_EPI = HALVES == 2 ? kRegs10[1] : kRegs[1],
                  REGS_SOFTMAX = HALVES == 2 ? kRegs10[2] : kRegs[2];
    float* l_buf = reinterpret_cast<float*>(smem + OFF_L);
    float* xch = reinterpret_cast<float*>(smem + OFF_XCH);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + OFF_BAR + BAR_COUNT * 8);
    volatile uint32_t* turn_cnt = reinterpret_cast<volatile uint32_t*>(smem + OFF_BAR + 248);   // softmax steps finished, x4 warps, per group
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
            mbar_init(bar(BAR_PREADY + i), 128 * HALVES);
            mbar_init(bar(BAR_OFINAL + i), 1);
            mbar_init(bar(BAR_OEMPTY + i), 128);
            mbar_init(bar(BAR_LREADY + i), 128 * HALVES);
            mbar_init(bar(BAR_SLOADED + i), 128 * HALVES);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_KVFULL + s), 1);
            mbar_init(bar(BAR_KVEMPTY + s), 2);       // one commit per MMA warp
        }
        turn_cnt[0] = 0;
        turn_cnt[1] = 0;
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
        reg_dealloc<REGS_CTRL>();
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
                        DBG_PROGRESS(0, (n_it << 12) | (it.nq << 8) | n);
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
            // key columns [c0, c0 + ncols) of the tile at kbase (key row r of a K-major SW128 box sits 128 r bytes in)
            auto issue_qk_cols = [&](uint32_t kbase, int c0, int ncols) {
                const uint32_t idesc_qk = idesc_bf16(BM, ncols, 0);
                const uint32_t qlo = smem_desc_lo(base + i * TILE_BYTES, 16);
                const uint32_t klo = smem_desc_lo(kbase + (uint32_t)c0 * 128u, 16);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t off = ((uint32_t)(kk >> 2) * BOX_BYTES + (uint32_t)(kk & 3) * 32u) >> 4;
                    mma_ss(tmem + i * 128 + c0, make_desc(qlo + off, hi_sw), make_desc(klo + off, hi_sw), idesc_qk, kk > 0);
                }
            };
            auto issue_qk = [&](uint32_t kbase, int ncols) { issue_qk_cols(kbase, 0, ncols); };
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
            uint32_t pr = 0;             // PREADY_i phases consumed (= softmax steps of this tile: also the SLOADED_i phase)
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
                        DBG_PROGRESS(1 + i, 0x80000000u | (n_done << 12) | n);
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
                    DBG_PROGRESS(1 + i, (n_done << 12) | (it.nq << 8) | j);
#if SVR2_V2_SPLIT_QK
                    const int hi_cols = more ? cols(j + 1) - 64 : 0;       // > 0: the next S tile has an upper half
                    if (hi_cols > 0) {
                        OTHER_WAIT(bar(BAR_KVFULL + ring_slot(2 * j + 2)), ring_par(2 * j + 2));          // K_{j+1}
                        OTHER_WAIT(bar(BAR_SLOADED + i), pr & 1u);                                        // S_i(j) is in registers
                        tc_fence_after();
                        if (elect_one()) issue_qk_cols(sRing + ring_slot(2 * j + 2) * TILE_BYTES, 64, hi_cols);
                        __syncwarp();
                    }
#endif
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
#if SVR2_V2_SPLIT_QK
                            issue_qk_cols(sRing + ring_slot(2 * j + 2) * TILE_BYTES, 0, cols(j + 1) < 64 ? cols(j + 1) : 64);
#else
                            issue_qk(sRing + ring_slot(2 * j + 2) * TILE_BYTES, cols(j + 1));
#endif
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
            DBG_PROGRESS(1 + i, 0xffff0000u | n_done);
            if (lane == 0 && i == 0) trace_at<kTrace>(p, 7, n_done);
        }
    } else if (warp < 8) {
        reg_dealloc<REGS_EPI>();
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
                if (quad == 0) DBG_PROGRESS(3, (n_it << 12) | (it.nq << 8) | i);
                OTHER_WAIT(bar(BAR_LREADY + i), cc & 1);
                OTHER_WAIT(bar(BAR_OFINAL + i), cc & 1);
                tc_fence_after();
                if (threadIdx.x == 128 && i == 0) trace_item<kTrace>(p, n_it, 5);
                const float* lrow = l_buf + (((cc & 1) * 2 + i) * 2) * 128 + row;
                const float inv = 1.0f / (HALVES == 2 ? lrow[0] + lrow[128] : lrow[0]);
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
                // two passes of 64 columns: registers -> swizzled 128-byte staging rows -> 128-byte row segments in HBM
#pragma unroll 1
                for (int hf = 0; hf < 2; ++hf) {
#pragma unroll 1
                    for (int c2 = 0; c2 < 2; ++c2) {
                        uint32_t ov[32];
                        tmem_ld32(tO + (hf * 2 + c2) * 32, ov);
                        tc_wait_ld();
                        if (hf == 1 && c2 == 1) {                        // O_i is in registers: the next item may overwrite it
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
                            const int ch = c2 * 4 + q;
                            *reinterpret_cast<uint4*>(stage + row * 128 + (((ch ^ row) & 7) << 4)) = u;
                        }
                    }
                    __syncwarp();
                    const int ch = lane & 7;
#pragma unroll 4
                    for (int t = 0; t < 8; ++t) {
                        const int rl = 4 * t + (lane >> 3);              // row (inside this warp's 32) served by my quarter-warp
                        const unsigned long long d = __shfl_sync(0xffffffffu, dst, rl);
                        const int r = quad * 32 + rl;
                        const uint4 u = *reinterpret_cast<const uint4*>(stage + r * 128 + (((ch ^ r) & 7) << 4));
                        if (d != 0ull) *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(d) + hf * 64 + ch * 8) = u;
                    }
                    __syncwarp();                                       // staging rows of this warp are free again
                }
                if (threadIdx.x == 128) trace_item<kTrace>(p, n_it, 6 + i);
                if (i) ++c1; else ++c0;
            }
        }
    } else if constexpr (HALVES == 2) {
        reg_alloc<REGS_SOFTMAX>();
        // ------------------------------------------------------------------ softmax group i, column half h: TWO threads per query row
        // warps 8-15: columns [0, 64) of the key tile (tiles 0, 1); warps 16-23: columns [64, 128).  The arithmetic per element
        // is generation 4's; what changes is who holds which columns, the exchange of the partial row max and the split row sum.
        const int half = (warp - 8) >> 3, i = ((warp - 8) >> 2) & 1, quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        const uint32_t tS = tmem + lane_off + i * 128;
        const uint32_t tO = tmem + lane_off + 256 + i * 128 + half * 64;
        float* x_mine = xch + (i * 2 + half) * 128 + row;
        const float* x_other = xch + (i * 2 + (half ^ 1)) * 128 + row;
        const int bar_id = 1 + i * 4 + quad;        // named barrier of the two warps that hold the same 32 rows
        uint32_t sf = 0, c_cnt = 0;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const Item it = decode_item(p, item);
            if (i >= it.nq) continue;
            float m = -INFINITY, l = 0.f;
            for (int j = 0; j < it.n_kv; ++j, ++sf) {
                const bool last = j == it.n_kv - 1;
                const int nch = (last ? it.last_cols : BN) >> 5;            // 32-column chunks in this key tile (warp-uniform)
                const int my_n = min(max(nch - 2 * half, 0), 2);            // ... of which this half owns chunks 2h, 2h + 1
                const int v0 = it.seq - j * BN - 64 * half;                 // my columns >= v0 are beyond the sequence
                SOFTMAX_WAIT(bar(BAR_SFULL + i), sf & 1u);
                tc_fence_after();
                // Pass 1: this half's partial row max straight from TMEM, 32 columns at a time.  The S row is NOT kept in registers:
                // a half row next to the exp temporaries needs ~110 registers, the role has 96, and a spill here is an L2 round trip
                // (this kernel leaves ~8 KB of L1) — run 13: 20 % slower than generation 4 with a dozen spilled scalars per step.
                float m8[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) m8[q] = -INFINITY;
                if (my_n > 0) {
                    uint32_t s[32];
                    tmem_ld32(tS + 64 * half, s);
                    tc_wait_ld();
                    if (v0 < 32) mask_chunk(s, v0);
                    chunk_max8(m8, s);
                }
                if (my_n > 1) {
                    uint32_t s[32];
                    tmem_ld32(tS + 64 * half + 32, s);
                    tc_wait_ld();
                    if (v0 < 64) mask_chunk(s, v0 - 32);
                    chunk_max8(m8, s);
                }
                float mx = fmaxf(max3(m8[0], m8[1], m8[2]), max3(max3(m8[3], m8[4], m8[5]), m8[6], m8[7]));
                // the row max over both halves (every later write of x_mine follows the partner's read: its P-ready arrival, which
                // the next S depends on, comes after it)
                *x_mine = mx;
                named_bar_sync(bar_id, 64);
                mx = fmaxf(mx, *x_other);
                if (j == 0) {
                    m = ref_of(mx);
                } else {
                    // lazy rescale, decided identically by both halves (same rows, same values): each rescales its 64 columns of O_i
                    const bool need = mx * kScaleLog2 - m > kRescaleThreshold;
                    if (__any_sync(0xffffffffu, need)) {
                        const float m_new = need ? ref_of(mx) : m;
                        const float alpha = ex2(m - m_new);
                        l *= alpha;
                        m = m_new;
#pragma unroll 1
                        for (int c8 = 0; c8 < 4; ++c8) {
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
                // Pass 2: S again from TMEM, chunk by chunk, P written over the front of S_i.  P chunk c (32 keys = 16 packed
                // columns) sits at tS + 16 c: chunks 0, 1 (lower half) land on S columns [0, 32), which the lower half itself has
                // consumed by then; chunks 2, 3 (upper half) land on S columns [32, 64) = the lower half's SECOND chunk, so the upper
                // half stores only after the lower half has that chunk in registers (arrive / sync on the pair's named barrier).
                const uint32_t tP = tS + 32 * half;
                float rs = 0.f;
                if (half == 0) {
                    if (my_n > 0) {
                        uint32_t s[32], pk[16];
                        tmem_ld32(tS, s);
                        tc_wait_ld();
                        if (v0 < 32) mask_chunk(s, v0);
                        rs = chunk_exp_mix<kFmaEighths>(s, pk, mb);
                        tmem_st16(tP, pk);
                    }
                    if (my_n > 1) {
                        uint32_t s[32], pk[16];
                        tmem_ld32(tS + 32, s);
                        tc_wait_ld();
                        named_bar_arrive(bar_id, 64);
                        if (v0 < 64) mask_chunk(s, v0 - 32);
                        rs += chunk_exp_mix<kFmaEighths>(s, pk, mb);
                        tmem_st16(tP + 16, pk);
                    } else {
                        named_bar_arrive(bar_id, 64);
                    }
                } else {
                    if (my_n > 0) {
                        uint32_t s[32], pk[16];
                        tmem_ld32(tS + 64, s);
                        tc_wait_ld();
                        if (v0 < 32) mask_chunk(s, v0);
                        rs = chunk_exp_mix<kFmaEighths>(s, pk, mb);
                        named_bar_sync(bar_id, 64);
                        tmem_st16(tP, pk);
                    } else {
                        named_bar_sync(bar_id, 64);
                    }
                    if (my_n > 1) {
                        uint32_t s[32], pk[16];
                        tmem_ld32(tS + 96, s);
                        tc_wait_ld();
                        if (v0 < 64) mask_chunk(s, v0 - 32);
                        rs += chunk_exp_mix<kFmaEighths>(s, pk, mb);
                        tmem_st16(tP + 16, pk);
                    }
                }
                l += rs;
                tc_wait_st();
                tc_fence_before();
                mbar_arrive(bar(BAR_PREADY + i));
            }
            // publish this half's partial row sum for the epilogue warps (double-buffered by item parity).  A one-key-tile item
            // can finish before the epilogue has picked up the PREVIOUS item's row sums (it may still be storing the item before
            // that): LREADY would then complete two phases and its parity alias.  The epilogue's "O is drained" arrival of the
            // previous item follows its LREADY wait, so waiting for it here keeps LREADY at most one phase ahead.
            if (c_cnt > 0) SOFTMAX_WAIT(bar(BAR_OEMPTY + i), (c_cnt - 1) & 1u);
            l_buf[(((c_cnt & 1) * 2 + i) * 2 + half) * 128 + row] = l;
            mbar_arrive(bar(BAR_LREADY + i));
            ++c_cnt;
        }
    } else {
        reg_alloc<REGS_SOFTMAX>();
        // ------------------------------------------------------------------ softmax group i: one thread per query row
        const int i = (warp - 8) >> 2, quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        const uint32_t tS = tmem + lane_off + i * 128;
        const uint32_t tO = tmem + lane_off + 256 + i * 128;
        const bool tracer = (threadIdx.x & 127) == 0 && i == 0;
        uint32_t sf = 0;            // SFULL_i phases consumed
        uint32_t c_cnt = 0;
        // Taking turns.  Left alone the two groups fall into lockstep: both run their softmax step at once, fighting for the
        // exp2 unit and the issue slots (~2600 cycles each instead of ~1300), then both wait while their MMAs run one after
        // the other.  turn_cnt[g] counts the softmax steps group g's warps have finished (bumped after the last exp2 is
        // issued); a group starts a step only after the other group finished the step that precedes it in the order
        // (item, key tile, group).  That staggers the groups by one softmax step, so one tile's MMAs run under the other
        // tile's softmax.  The turn is a scheduling hint and must never be able to block: it is a monotonic counter (an
        // mbarrier phase can be missed by a slow warp — that deadlocked an earlier version) and the wait is bounded.
        uint32_t other_steps = 0;   // steps of the other group in the items before this one
        int n_it = -1;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const Item it = decode_item(p, item);
            ++n_it;
            if (i >= it.nq) {
                other_steps += it.n_kv;
                continue;
            }
            float m = -INFINITY, l = 0.f;
            for (int j = 0; j < it.n_kv; ++j, ++sf) {
                if (SVR2_TURNS && p.turns) {
                    // group 1 follows group 0's step j of this item; group 0 follows group 1's previous step
                    const uint32_t need = 4u * (i ? other_steps + j + 1 : other_steps + (it.nq == 2 ? j : 0));
                    uint32_t t0;
                    asm volatile("mov.u32 %0, %%clock;" : "=r"(t0));
                    while (turn_cnt[i ^ 1] < need) {
                        uint32_t t1;
                        asm volatile("mov.u32 %0, %%clock;" : "=r"(t1));
                        if (t1 - t0 > (uint32_t)SVR2_TURN_BOUND) break;
                        if (SVR2_TURN_SLEEP) __nanosleep(SVR2_TURN_SLEEP);
                    }
                }
                const bool last = j == it.n_kv - 1;
                const int nch = (last ? it.last_cols : BN) >> 5;            // 32-column chunks in this key tile (warp-uniform)
                const int valid = it.seq - j * BN;                          // key columns >= valid are beyond the sequence
                const bool fine = kTrace && (threadIdx.x & 127) == 0 && n_it == 2 && j < 4;
                DBG_PROGRESS(4 + 4 * i + quad, (n_it << 12) | (it.nq << 8) | j);
                SOFTMAX_WAIT(bar(BAR_SFULL + i), sf & 1u);
                tc_fence_after();
                if (fine) trace_at<kTrace>(p, 44 + 8 * i + 2 * j, clock64());
                if (tracer) {
                    if (j == 0) trace_item<kTrace>(p, n_it, 1);
                    if (last) trace_item<kTrace>(p, n_it, 3);
                }
                // The whole S row in registers: ONE TMEM round trip per key tile.  Register block b holds chunk nch - 4 + b,
                // so the chunk with the end of the sequence is always block 3 and the masking code exists once.
                uint32_t s[128];
                const uint32_t tS3 = tS + 32 * (nch - 1);
                if (nch > 3) tmem_ld32(tS3 - 96, s);
                if (nch > 2) tmem_ld32(tS3 - 64, s + 32);
                if (nch > 1) tmem_ld32(tS3 - 32, s + 64);
                tmem_ld32(tS3, s + 96);
                tc_wait_ld();
#if SVR2_V2_SPLIT_QK
                tc_fence_before();
                mbar_arrive(bar(BAR_SLOADED + i));                                     // the upper half of S_i may be overwritten
#endif
                if (valid < nch * 32) mask_chunk(s + 96, valid - 32 * (nch - 1));      // only in the last key tile of a sequence
                float m8[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) m8[q] = -INFINITY;
                if (nch > 3) chunk_max8(m8, s);
                if (nch > 2) chunk_max8(m8, s + 32);
                if (nch > 1) chunk_max8(m8, s + 64);
                chunk_max8(m8, s + 96);
                const float mx = fmaxf(max3(m8[0], m8[1], m8[2]), max3(max3(m8[3], m8[4], m8[5]), m8[6], m8[7]));
                if (j == 0) {
                    m = ref_of(mx);
                } else {
                    // lazy rescale: O_i and l keep the old reference until the row max has moved by more than 2^8
                    const bool need = mx * kScaleLog2 - m > kRescaleThreshold;
                    if (__any_sync(0xffffffffu, need)) {                    // warp-uniform: tcgen05.ld/st are warp collectives
                        // S_i(j) was committed after P_i V(j-1): O_i is quiescent until this thread publishes P(j)
                        const float m_new = need ? ref_of(mx) : m;
                        const float alpha = ex2(m - m_new);                    // an exact power of two
                        l *= alpha;
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
                // bf16 P overwrites the front of S_i chunk by chunk (the whole row is already in registers); kFmaEighths/8 of the
                // exponentials run on the FMA pipe, the rest on MUFU
                const uint32_t tP3 = tS + 16 * (nch - 1);
                float rs = 0.f;
                if (nch > 3) {
                    uint32_t pk[16];
                    rs = chunk_exp_mix<kFmaEighths>(s, pk, mb);
                    tmem_st16(tP3 - 48, pk);
                }
                if (nch > 2) {
                    uint32_t pk[16];
                    rs += chunk_exp_mix<kFmaEighths>(s + 32, pk, mb);
                    tmem_st16(tP3 - 32, pk);
                }
                if (nch > 1) {
                    uint32_t pk[16];
                    rs += chunk_exp_mix<kFmaEighths>(s + 64, pk, mb);
                    tmem_st16(tP3 - 16, pk);
                }
                {
                    uint32_t pk[16];
                    rs += chunk_exp_mix<kFmaEighths>(s + 96, pk, mb);
                    tmem_st16(tP3, pk);
                }
                l += rs;
                if (SVR2_TURNS && p.turns && lane == 0) atomicAdd(const_cast<uint32_t*>(&turn_cnt[i]), 1u);
                tc_wait_st();
                tc_fence_before();
                mbar_arrive(bar(BAR_PREADY + i));
                if (fine) trace_at<kTrace>(p, 45 + 8 * i + 2 * j, clock64());
                if (tracer) {
                    if (j == 0) trace_item<kTrace>(p, n_it, 2);
                    if (last) trace_item<kTrace>(p, n_it, 4);
                }
            }
            other_steps += (i == 1 || it.nq == 2) ? it.n_kv : 0;
            // publish the row sum for the epilogue warps (double-buffered by item parity)
            if (c_cnt > 0) SOFTMAX_WAIT(bar(BAR_OEMPTY + i), (c_cnt - 1) & 1u);      // see generation 10 above: LREADY stays one phase ahead at most
            l_buf[(((c_cnt & 1) * 2 + i) * 2) * 128 + row] = l;
            mbar_arrive(bar(BAR_LREADY + i));
            ++c_cnt;
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

}  // namespace tc2
}  // namespace svr2

// ------------------------------------------------------------------------------------------------ host side
namespace svr2 {

int launch_window_attn_tc2(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int n_windows, int H, int W, int txt_len, int heads, int64_t P, void* out_vid,
                           void* txt_part, int halves, cudaStream_t st) {
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap(&tq, q, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tk, k, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tv, v, P, heads)) != SVR2_OK) return rc;
    static PerDeviceFlag smem_ok[3];
    if ((rc = ensure_dyn_smem(tc2::window_attn_tc2_kernel<false, 1>, tc2::SMEM_BYTES, smem_ok[0], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    if ((rc = ensure_dyn_smem(tc2::window_attn_tc2_kernel<true, 1>, tc2::SMEM_BYTES, smem_ok[1], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
#ifdef SVR2_EXPERIMENTS
    if ((rc = ensure_dyn_smem(tc2::window_attn_tc2_kernel<false, 2>, tc2::SMEM_BYTES, smem_ok[2], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
#else
    if (halves != 1) {
        set_error("svr2_window_attn_fwd_impl: generation 10 (two threads per query row) is an experiment — measured 4-8 %% slower than "
                  "generation 4 (run 14) — and is only compiled with `make EXPERIMENTS=1`");
        return SVR2_E_UNSUPPORTED;
    }
#endif
    tc2::Params p;
    p.tab = tab; p.work = work; p.n_items = n_work * heads;
    p.H = H; p.W = W; p.txt_len = txt_len; p.heads = heads;
    p.out_vid = (__nv_bfloat16*)out_vid; p.txt_part = (__nv_bfloat16*)txt_part;
    p.trace = g_trace; p.trace_ctas = g_trace_ctas;
    // Turn-taking between the two softmax groups is compiled out (SVR2_TURNS 0: the bounded-counter version measured no
    // gain in round 1); builds with -DSVR2_TURNS=1 enable it for short sequences only.
    p.turns = SVR2_TURNS && (P / n_windows <= 8 * tc2::BN);
    const int sms = sm_count();
    const int grid = p.n_items < sms ? p.n_items : sms;       // one persistent CTA per SM
#ifdef SVR2_EXPERIMENTS
    if (halves == 2) tc2::window_attn_tc2_kernel<false, 2><<<grid, tc2::NTHREADS + 256, tc2::SMEM_BYTES, st>>>(tq, tk, tv, p);
    else
#endif
    if (p.trace != nullptr) tc2::window_attn_tc2_kernel<true, 1><<<grid, tc2::NTHREADS, tc2::SMEM_BYTES, st>>>(tq, tk, tv, p);
    else tc2::window_attn_tc2_kernel<false, 1><<<grid, tc2::NTHREADS, tc2::SMEM_BYTES, st>>>(tq, tk, tv, p);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(tcgen05, persistent)");
    return SVR2_OK;
}

}  // namespace svr2
