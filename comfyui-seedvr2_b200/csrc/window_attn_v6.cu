// svr2_window_attn_fwd, generation 8 — generation 7's pipeline on CTA PAIRS: tcgen05.mma.cta_group::2 (M = 256 over two SMs).
//
// Why: generation 7's wait profile (profiles/r01_run28_attn_gen7_wait_profile.txt) showed that the 16 MMAs of a tile step
// occupy the tensor pipe for ~1,730 cycles instead of 1,086 because the SS / TS operands and the TMA fills move 128-160 KB
// of shared memory per tile step at 128 B/clk.  In a CTA pair each SM supplies HALF of the B operand:
//   S = Q K^T : A = this CTA's 128 query rows (smem), B = K tile, this CTA holds N/2 of its key rows
//   O += P V  : A = this CTA's P rows (TMEM),          B = V tile, this CTA holds 64 of its 128 d columns
// so a CTA reads 32 + 16 + 16 KB and its TMA writes 16 + 16 KB per tile step: 96 KB.
//
// The two query tiles of a work-list entry (window, pair of query tiles, head) go to the two CTAs of a cluster; each CTA
// keeps generation 7's layout — S0 S1 | P0 P1 | O in TMEM, two softmax warpgroups on alternate key tiles, m and l handed
// over through shared memory, arithmetic and therefore output bits identical to generation 4.  What becomes pair-wide:
//   * one thread of the LEADER CTA (cluster rank 0) issues every MMA; its commits are multicast to both CTAs' barriers;
//   * both CTAs' TMA loads complete on the leader's "full" barriers (cp.async.bulk.tensor ... .cta_group::2);
//   * "S is in registers", "P is ready" and "O is drained" are counted on the leader's barriers from both CTAs (256 arrivals).
// A pair whose second tile does not exist (sequence <= 128 rows of the pair) still runs it on CTA 1 and stores nothing.
//
//   TMEM columns: S[b] = 128 b, P[b] = 256 + 64 b, O = 384                       (b = g & 1, g = step counter of the pair)
//   SMEM per CTA: Q[2] (32 KB each) | ring[8] of 16-KB half tiles | staging (32 KB) | m[128] l[128] l_final[2][128] | mbarriers
// Ring order = issue order: K(0), K(1), K(2), then K(g+3), V(g).
#include <cuda.h>
#include <cstdlib>

#include "common.cuh"
#include "attn_launch.cuh"
#include "sm100_pair.cuh"
#include "sm100_ptx.cuh"
#include "softmax_row.cuh"

namespace svr2 {
namespace tc6 {

using namespace ptx;
using namespace pairx;
using namespace rowsm;

constexpr int D = SVR2_HEAD_DIM;          // 128
constexpr int BM = 128, BN = 128;
constexpr int TILE_BYTES = BM * D * 2;    // 32768: a query tile
constexpr int BOX_BYTES = TILE_BYTES / 2; // 128 rows x 64 cols bf16
constexpr int HALF_BYTES = TILE_BYTES / 2;   // this CTA's half of a K tile (64 keys x 128 d) or V tile (128 keys x 64 d)
constexpr int KBOX_BYTES = HALF_BYTES / 2;   // 64 keys x 64 d
constexpr int STAGES = 8;
constexpr int NTHREADS = 512;
#ifndef SVR2_REGS_C      // experiment builds: -DSVR2_REGS_C=.. -DSVR2_REGS_E=.. -DSVR2_REGS_S=..
#define SVR2_REGS_C 48     // run 61: 48/64/200 removes the softmax role's spills: +3.7 % at C1 over 72/72/184
#define SVR2_REGS_E 64
#define SVR2_REGS_S 200
#endif
constexpr int REGS_CTRL = SVR2_REGS_C, REGS_EPI = SVR2_REGS_E, REGS_SOFTMAX = SVR2_REGS_S;   // 4 x (C + E + 2 S) x 32 = 65,536
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_STAGE = OFF_RING + STAGES * HALF_BYTES;
constexpr int OFF_M = OFF_STAGE + TILE_BYTES;          // float m[128]
constexpr int OFF_L = OFF_M + 128 * 4;                 // float l[128]
constexpr int OFF_LFIN = OFF_L + 128 * 4;              // float l_final[2][128]
constexpr int OFF_BAR = OFF_LFIN + 2 * 128 * 4;
constexpr int SMEM_BYTES = OFF_BAR + 512;
constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t COL_P = 256, COL_O = 384;
constexpr int kFmaEighths = 3;
// Q K^T runs kAhead steps ahead of P V in the tensor-core program: QK(0..kAhead-1), then QK(g+kAhead), PV(g).  QK(g+3) needs
// S[(g+1) & 1] in the registers of step g+1's softmax — always earlier than P(g) — so with 3 no softmax waits for its S while the
// issuer sits in front of PV(g) waiting for P(g) (2 left ~500 cycles of "S full" waits per step in the wait profile).
#ifndef SVR2_V6_AHEAD
#define SVR2_V6_AHEAD 3
#endif
constexpr int kAhead = SVR2_V6_AHEAD;
constexpr float kRescaleThreshold = 8.0f;
__device__ __forceinline__ float ref_of(float row_max) { return row_max * kScaleLog2; }

// [L] = lives in each CTA and is signalled locally or by a multicast commit; [G] = only the leader's copy is used (256 arrivals
// or the transaction bytes of both CTAs)
enum { BAR_QFULL = 0 /*G*/, BAR_QEMPTY = 2 /*L*/, BAR_SFULL = 4 /*L*/, BAR_SLOADED = 6 /*G*/, BAR_PREADY = 8 /*G*/,
       BAR_PVDONE = 10 /*L*/, BAR_MPUB = 12 /*L*/, BAR_LPUB = 14 /*L*/, BAR_OFINAL = 16 /*L*/, BAR_OEMPTY_G = 17 /*G*/,
       BAR_OEMPTY_L = 18 /*L*/, BAR_LFINAL = 19 /*L*/, BAR_KVFULL = 20 /*G*/, BAR_KVEMPTY = 20 + STAGES /*L*/,
       BAR_COUNT = 20 + 2 * STAGES };
static_assert(BAR_COUNT * 8 + 8 <= 512, "barrier block too small");
static_assert(SMEM_BYTES <= 232448, "shared memory");

struct Params {
    const int32_t* tab;
    const int32_t* work;
    int n_items;           // work-list entries x heads
    int H, W, txt_len, heads;
    __nv_bfloat16* out_vid;
    __nv_bfloat16* txt_part;
    long long* prof;       // kProf instantiation: wait / work cycle counters of cluster 0 (layout of window_attn_v5.cu)
};

struct Item {
    Win w;
    int wi, head, seq, q0, n_kv, last_cols;
};
__device__ __forceinline__ void decode_item(const Params& p, int item, int cta_rank, Item& it) {
    const int e = item / p.heads;
    it.head = item - e * p.heads;
    it.wi = __ldg(p.work + 2 * e);
    const int qpair = __ldg(p.work + 2 * e + 1);
    it.w = load_win(p.tab, it.wi);
    it.seq = it.w.n_vid + p.txt_len;
    it.q0 = (2 * qpair + cta_rank) * BM;               // may lie beyond the sequence on CTA 1: computed, never stored
    it.n_kv = (it.seq + BN - 1) / BN;
    const int rem = it.seq - (it.n_kv - 1) * BN;
    it.last_cols = (rem + 31) & ~31;
}

// Walks the (item, key tile) steps of this CTA pair in order; every role of both CTAs runs the same walk.
struct Cursor {
    int item, n, j;
    bool valid;
    Item it;
};
struct Walk {
    int first, stride, rank;
};
__device__ __forceinline__ void cur_load(const Params& p, const Walk& w, Cursor& c) {
    c.valid = c.item < p.n_items;
    if (c.valid) decode_item(p, c.item, w.rank, c.it);
}
__device__ __forceinline__ void cur_init(const Params& p, const Walk& w, Cursor& c) {
    c.item = w.first; c.n = 0; c.j = 0;
    cur_load(p, w, c);
}
__device__ __forceinline__ void cur_next(const Params& p, const Walk& w, Cursor& c) {
    if (++c.j == c.it.n_kv) {
        c.j = 0; ++c.n; c.item += w.stride;
        cur_load(p, w, c);
    }
}
__device__ __forceinline__ void cur_next_item(const Params& p, const Walk& w, Cursor& c) {
    c.j = 0; ++c.n; c.item += w.stride;
    cur_load(p, w, c);
}

template <int N> __device__ __forceinline__ void reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

#ifndef SVR2_V6_FLAT_WALK
#define SVR2_V6_FLAT_WALK 0       // 1: the softmax warpgroups walk the generic (item, key tile) cursor like the other roles (A/B)
#endif
#define PROF_WAIT(slot, stmt) do { if constexpr (kProf) { const long long t0_ = clock64(); stmt; prof_acc[slot] += clock64() - t0_; } else { stmt; } } while (0)
template <bool kProf>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
window_attn_tc6_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t base = smem_u32(smem);
    const uint32_t sRing = base + OFF_RING;
    const uint32_t sBar = base + OFF_BAR;
    float* m_sh = reinterpret_cast<float*>(smem + OFF_M);
    float* l_sh = reinterpret_cast<float*>(smem + OFF_L);
    float* l_fin = reinterpret_cast<float*>(smem + OFF_LFIN);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + OFF_BAR + BAR_COUNT * 8);
    auto bar = [&](int slot) { return sBar + 8u * (uint32_t)slot; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();
    const bool leader = rank == 0;
    Walk walk;
    walk.first = (int)(blockIdx.x >> 1);
    walk.stride = (int)(gridDim.x >> 1);
    walk.rank = (int)rank;

    if (threadIdx.x == 0) {
        if (base & 1023u) __trap();
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(BAR_QFULL + i), 1);
            mbar_init(bar(BAR_QEMPTY + i), 1);
            mbar_init(bar(BAR_SFULL + i), 1);
            mbar_init(bar(BAR_SLOADED + i), 8);       // one arrival per softmax warp of the pair: a remote arrival is a
            mbar_init(bar(BAR_PREADY + i), 8);        // transaction on the leader's shared memory, 256 of them serialise
            mbar_init(bar(BAR_PVDONE + i), 1);
            mbar_init(bar(BAR_MPUB + i), 128);        // local barriers: one arrival per thread (per-warp arrivals behind a
            mbar_init(bar(BAR_LPUB + i), 128);        // __syncwarp measured 1 % slower, run 41)
        }
        mbar_init(bar(BAR_OFINAL), 1);
        mbar_init(bar(BAR_OEMPTY_G), 8);
        mbar_init(bar(BAR_OEMPTY_L), 128);
        mbar_init(bar(BAR_LFINAL), 128);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_KVFULL + s), 1);
            mbar_init(bar(BAR_KVEMPTY + s), 1);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_pair(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_q);
        prefetch_tmap(&tm_k);
        prefetch_tmap(&tm_v);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();              // the peer's barriers are initialised and its TMEM allocated before anything crosses over
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp < 4) {
        reg_dealloc<REGS_CTRL>();
        if (warp == 0) {
            // -------------------------------------------------------------- TMA producer: this CTA's halves
            if (lane == 0) {
                uint32_t kv_n = 0;
                auto next_stage = [&](uint32_t& st) {
                    st = kv_n % STAGES;
                    const uint32_t use = kv_n / STAGES;
                    mbar_wait(bar(BAR_KVEMPTY + st), (use & 1) ^ 1);                      // local copy of the multicast commit
                    if (leader) mbar_expect_tx(bar(BAR_KVFULL + st), 2 * HALF_BYTES);     // both CTAs' halves
                    ++kv_n;
                };
                auto load_k = [&](const Cursor& c) {
                    if (c.j == 0) {                                     // the item's query tile first
                        const uint32_t qb = c.n & 1;
                        if (c.n >= 2) mbar_wait(bar(BAR_QEMPTY + qb), ((c.n >> 1) - 1) & 1);
                        if (leader) mbar_expect_tx(bar(BAR_QFULL + qb), 2 * TILE_BYTES);
                        const int row = c.it.w.row_start + c.it.q0;
                        tma_load_3d_pair(base + qb * TILE_BYTES, &tm_q, bar(BAR_QFULL + qb), 0, c.it.head, row);
                        tma_load_3d_pair(base + qb * TILE_BYTES + BOX_BYTES, &tm_q, bar(BAR_QFULL + qb), 64, c.it.head, row);
                    }
                    // this CTA's half of the key tile's N columns: key rows [rank * N/2, (rank + 1) * N/2); the box is 64 rows,
                    // the MMA reads only the first N/2 of them
                    const int ncols = (c.j == c.it.n_kv - 1) ? c.it.last_cols : BN;
                    const int row = c.it.w.row_start + c.j * BN + (int)rank * (ncols >> 1);
                    uint32_t st;
                    next_stage(st);
                    tma_load_3d_pair(sRing + st * HALF_BYTES, &tm_k, bar(BAR_KVFULL + st), 0, c.it.head, row);
                    tma_load_3d_pair(sRing + st * HALF_BYTES + KBOX_BYTES, &tm_k, bar(BAR_KVFULL + st), 64, c.it.head, row);
                };
                auto load_v = [&](const Cursor& c) {
                    // this CTA's half of the value tile's N = d columns: d [64 rank, 64 rank + 64), all 128 keys
                    uint32_t st;
                    next_stage(st);
                    tma_load_3d_pair(sRing + st * HALF_BYTES, &tm_v, bar(BAR_KVFULL + st), 64 * (int)rank, c.it.head,
                                     c.it.w.row_start + c.j * BN);
                };
                Cursor kc, vc;
                cur_init(p, walk, kc);
                vc = kc;
                for (int a = 0; a < kAhead && kc.valid; ++a) {
                    load_k(kc);
                    cur_next(p, walk, kc);
                }
                while (vc.valid) {
                    if (kc.valid) {
                        load_k(kc);
                        cur_next(p, walk, kc);
                    }
                    load_v(vc);
                    cur_next(p, walk, vc);
                }
            }
        } else if (warp == 1 && leader) {
            // -------------------------------------------------------------- tcgen05.mma issuer of the pair
            constexpr uint32_t idesc_pv = idesc_bf16(2 * BM, D, 1);
            constexpr uint32_t hi_sw = smem_desc_hi_sw128(1024);
            uint32_t kv_n = 0, g_qk = 0, g_pv = 0;
            long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const long long prof_t0 = kProf ? clock64() : 0;
            auto do_qk = [&](const Cursor& c) {
                const uint32_t b = g_qk & 1, qb = c.n & 1;
                const int ncols = (c.j == c.it.n_kv - 1) ? c.it.last_cols : BN;
                const uint32_t st = kv_n % STAGES, par = (kv_n / STAGES) & 1;
                if (c.j == 0) PROF_WAIT(7, mbar_wait_cluster(bar(BAR_QFULL + qb), (c.n >> 1) & 1));
                if (g_qk >= 2) PROF_WAIT(2, mbar_wait_cluster(bar(BAR_SLOADED + b), ((g_qk >> 1) - 1) & 1));   // softmax(g-2) has S[b] in registers
                PROF_WAIT(3, mbar_wait_cluster(bar(BAR_KVFULL + st), par));
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t idesc_qk = idesc_bf16(2 * BM, ncols, 0);
                    const uint32_t qlo = smem_desc_lo(base + qb * TILE_BYTES, 16);
                    const uint32_t klo = smem_desc_lo(sRing + st * HALF_BYTES, 16);
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        const uint32_t qoff = ((uint32_t)(kk >> 2) * BOX_BYTES + (uint32_t)(kk & 3) * 32u) >> 4;
                        const uint32_t koff = ((uint32_t)(kk >> 2) * KBOX_BYTES + (uint32_t)(kk & 3) * 32u) >> 4;
                        mma2_ss(tmem + b * 128, make_desc(qlo + qoff, hi_sw), make_desc(klo + koff, hi_sw), idesc_qk, kk > 0);
                    }
                    tc_commit_pair(bar(BAR_SFULL + b));
                    tc_commit_pair(bar(BAR_KVEMPTY + st));
                    if (c.j == c.it.n_kv - 1) tc_commit_pair(bar(BAR_QEMPTY + qb));
                }
                __syncwarp();
                ++kv_n; ++g_qk;
            };
            auto do_pv = [&](const Cursor& c) {
                const uint32_t b = g_pv & 1;
                const bool last = c.j == c.it.n_kv - 1;
                const int ksteps = (last ? c.it.last_cols : BN) >> 4;
                const uint32_t st = kv_n % STAGES, par = (kv_n / STAGES) & 1;
                PROF_WAIT(4, mbar_wait_cluster(bar(BAR_KVFULL + st), par));
                PROF_WAIT(5, mbar_wait_cluster(bar(BAR_PREADY + b), (g_pv >> 1) & 1));
                if (c.j == 0 && c.n > 0) PROF_WAIT(6, mbar_wait_cluster(bar(BAR_OEMPTY_G), (c.n - 1) & 1));
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t vlo = smem_desc_lo(sRing + st * HALF_BYTES, BOX_BYTES);
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)
                        if (kk < ksteps)
                            mma2_ts(tmem + COL_O, tmem + COL_P + b * 64 + kk * 8, make_desc(vlo + kk * (2048 >> 4), hi_sw),
                                    idesc_pv, !(c.j == 0 && kk == 0));
                    tc_commit_pair(bar(BAR_PVDONE + b));
                    tc_commit_pair(bar(BAR_KVEMPTY + st));
                    if (last) tc_commit_pair(bar(BAR_OFINAL));
                }
                __syncwarp();
                ++kv_n; ++g_pv;
            };
            Cursor kc, vc;
            cur_init(p, walk, kc);
            vc = kc;
            for (int a = 0; a < kAhead && kc.valid; ++a) {
                do_qk(kc);
                cur_next(p, walk, kc);
            }
            while (vc.valid) {
                if (kc.valid) {
                    do_qk(kc);
                    cur_next(p, walk, kc);
                }
                do_pv(vc);
                cur_next(p, walk, vc);
            }
            if constexpr (kProf) {
                if (blockIdx.x == 0 && lane == 0 && p.prof != nullptr) {
                    prof_acc[0] = g_pv; prof_acc[1] = clock64() - prof_t0;
                    for (int q = 0; q < 8; ++q) p.prof[16 + q] = prof_acc[q];
                }
            }
        }
    } else if (warp < 8) {
        reg_dealloc<REGS_EPI>();
        // ------------------------------------------------------------------ epilogue warps (this CTA's query tile)
        const int quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        unsigned char* stage = smem + OFF_STAGE;
        const int64_t row_stride = (int64_t)p.heads * D;
        const uint32_t tO = tmem + lane_off + COL_O;
        const uint32_t oempty_g = mapa(bar(BAR_OEMPTY_G), 0);
        Cursor c;
        cur_init(p, walk, c);
        while (c.valid) {
            const Item it = c.it;
            const uint32_t n = c.n;
            mbar_wait(bar(BAR_LFINAL), n & 1);
            mbar_wait(bar(BAR_OFINAL), n & 1);
            tc_fence_after();
            const float inv = 1.0f / l_fin[(n & 1) * 128 + row];
#pragma unroll 1
            for (int c4 = 0; c4 < 4; ++c4) {
                uint32_t ov[32];
                tmem_ld32(tO + c4 * 32, ov);
                tc_wait_ld();
                if (c4 == 3) {                                   // O is in registers: the next item may overwrite it
                    tc_fence_before();
                    mbar_arrive(bar(BAR_OEMPTY_L));
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(oempty_g);
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
            const int jrow = it.q0 + row;
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
                const int rl = 2 * t + (lane >> 4);
                const unsigned long long d = __shfl_sync(0xffffffffu, dst, rl);
                const int r = quad * 32 + rl;
                const uint4 u = *reinterpret_cast<const uint4*>(stage + r * 256 + (((ch & 8) | ((ch ^ r) & 7)) << 4));
                if (d != 0ull) *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(d) + ch * 8) = u;
            }
            __syncwarp();
            cur_next_item(p, walk, c);
        }
    } else {
        reg_alloc<REGS_SOFTMAX>();
        // ------------------------------------------------------------------ softmax warpgroup wg: steps g with (g & 1) == wg
        const int wg = (warp - 8) >> 2, quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        const uint32_t tS = tmem + lane_off + wg * 128;
        const uint32_t tP = tmem + lane_off + COL_P + wg * 64;
        const uint32_t tO = tmem + lane_off + COL_O;
        const uint32_t b = wg, ob = wg ^ 1;
        const uint32_t sloaded_g = mapa(bar(BAR_SLOADED + b), 0), pready_g = mapa(bar(BAR_PREADY + b), 0);
        long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        long long ph_acc[4] = {0, 0, 0, 0};          // kProf: cycles in ld S | mask + max | exp2 + pack + st issue | wait::st
        long long ph_t = 0;
        long long x_acc[6] = {0, 0, 0, 0, 0, 0};     // kProf: glue between the phases (see scripts/attn_v5_prof.py)
        long long x_t = 0;
        const long long prof_t0 = kProf ? clock64() : 0;
#if SVR2_V6_FLAT_WALK
        Cursor c;
        cur_init(p, walk, c);
        for (uint32_t g = 0; c.valid; ++g, cur_next(p, walk, c)) {
            if ((g & 1u) != (uint32_t)wg) continue;
            const int j = c.j, cn = c.n, it_n_kv = c.it.n_kv, it_last_cols = c.it.last_cols, it_seq = c.it.seq;
#else
        // The softmax warpgroups need only the packed length of an item.  They walk items in an outer loop and their own key
        // tiles (every other one) in an inner loop; the next item's length is loaded two items ahead (window index) and one item
        // ahead (length), so an item boundary exposes no global-load latency.  (The generic cursor below costs ~170 more
        // instructions per step at IPC 0.25: profiles/r01_run54_attn_ncu_stall_sampling.txt.)
        auto item_wi = [&](int item) { return item < p.n_items ? __ldg(p.work + 2 * (item / p.heads)) : -1; };
        auto wi_seq = [&](int wi) { return wi >= 0 ? __ldg(p.tab + SVR2_WIN_FIELDS * wi + 7) + p.txt_len : 0; };
        int next_seq = wi_seq(item_wi(walk.first));
        int next_wi = item_wi(walk.first + walk.stride);
        uint32_t g0 = 0;
        for (int item = walk.first, cn = 0; item < p.n_items; item += walk.stride, ++cn) {
            const int it_seq = next_seq;
            next_seq = wi_seq(next_wi);
            next_wi = item_wi(item + 2 * walk.stride);
            const int it_n_kv = (it_seq + BN - 1) / BN;
            const int it_last_cols = (it_seq - (it_n_kv - 1) * BN + 31) & ~31;
          for (int j = (int)((g0 ^ (uint32_t)wg) & 1u); j < it_n_kv; j += 2) {
            const uint32_t g = g0 + (uint32_t)j;
#endif
            const bool last = j == it_n_kv - 1;
            const int nch = (last ? it_last_cols : BN) >> 5;
            const int valid = it_seq - j * BN;
            const uint32_t par = (g >> 1) & 1u;
            const uint32_t ppar = ((g - 1) >> 1) & 1u;
            if constexpr (kProf) { if (x_t != 0) x_acc[5] += clock64() - x_t; }
            PROF_WAIT(2, mbar_wait(bar(BAR_SFULL + b), par));
            if constexpr (kProf) prof_acc[0] += 1;
            tc_fence_after();
            if constexpr (kProf) ph_t = clock64();
            uint32_t s[128];
            const uint32_t tS3 = tS + 32 * (nch - 1);
            if (nch > 3) tmem_ld32(tS3 - 96, s);
            if (nch > 2) tmem_ld32(tS3 - 64, s + 32);
            if (nch > 1) tmem_ld32(tS3 - 32, s + 64);
            tmem_ld32(tS3, s + 96);
            tc_wait_ld();
            if constexpr (kProf) { const long long t = clock64(); ph_acc[0] += t - ph_t; ph_t = t; }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(sloaded_g);              // Q K^T(g+2) may overwrite S[b] (counted over both CTAs)
            if constexpr (kProf) { const long long t = clock64(); x_acc[1] += t - ph_t; ph_t = t; }
            if (valid < nch * 32) mask_chunk(s + 96, valid - 32 * (nch - 1));
            float m8[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) m8[q] = -INFINITY;
            if (nch > 3) chunk_max8(m8, s);
            if (nch > 2) chunk_max8(m8, s + 32);
            if (nch > 1) chunk_max8(m8, s + 64);
            chunk_max8(m8, s + 96);
            const float mx = fmaxf(max3(m8[0], m8[1], m8[2]), max3(max3(m8[3], m8[4], m8[5]), m8[6], m8[7]));
            if constexpr (kProf) { if (mx > -1e30f) ph_acc[1] += clock64() - ph_t; }
            if (g > 0) PROF_WAIT(3, mbar_wait(bar(BAR_MPUB + ob), ppar));
            if constexpr (kProf) x_t = clock64();
            float m, alpha = 1.0f;
            if (j == 0) {
                m = ref_of(mx);
            } else {
                m = m_sh[row];
                const bool need = mx * kScaleLog2 - m > kRescaleThreshold;
                if (__any_sync(0xffffffffu, need)) {
                    mbar_wait(bar(BAR_PVDONE + ob), ppar);
                    tc_fence_after();
                    const float m_new = need ? ref_of(mx) : m;
                    alpha = ex2(m - m_new);
                    m = m_new;
#pragma unroll 1
                    for (int c8 = 0; c8 < 8; ++c8) {
                        uint32_t ov[16];
                        tmem_ld16(tO + c8 * 16, ov);
                        tc_wait_ld();
#pragma unroll
                        for (int cc = 0; cc < 16; ++cc) ov[cc] = __float_as_uint(__uint_as_float(ov[cc]) * alpha);
                        tmem_st16(tO + c8 * 16, ov);
                    }
                    tc_wait_st();
                }
            }
            m_sh[row] = m;
            mbar_arrive(bar(BAR_MPUB + b));
            if constexpr (kProf) x_acc[2] += clock64() - x_t;
            // P[b] is free once P V(g-2) has read it.  (Waiting later, in front of the first store of P, costs registers the
            // step does not have: the spills made the step 450 cycles longer, run 38.)
            if (g >= 2) PROF_WAIT(4, mbar_wait(bar(BAR_PVDONE + b), (par ^ 1u)));
            const float mb = m;
            const uint32_t tP3 = tP + 16 * (nch - 1);
            if constexpr (kProf) ph_t = clock64();
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
            if constexpr (kProf) { if (rs >= 0.f) { const long long t = clock64(); ph_acc[2] += t - ph_t; ph_t = t; } }
            tc_wait_st();
            if constexpr (kProf) { const long long t = clock64(); ph_acc[3] += t - ph_t; ph_t = t; }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(pready_g);
            if constexpr (kProf) x_acc[3] += clock64() - ph_t;
            if (g > 0) PROF_WAIT(5, mbar_wait(bar(BAR_LPUB + ob), ppar));
            if constexpr (kProf) x_t = clock64();
            float l = 0.f;
            if (j > 0) l = l_sh[row];
            l *= alpha;
            l += rs;
            if (last) {
                if (cn > 0) PROF_WAIT(6, mbar_wait(bar(BAR_OEMPTY_L), (cn - 1) & 1));
                l_fin[(cn & 1) * 128 + row] = l;
                mbar_arrive(bar(BAR_LFINAL));
            }
            l_sh[row] = l;
            mbar_arrive(bar(BAR_LPUB + b));
            if constexpr (kProf) { const long long t = clock64(); x_acc[4] += t - x_t; x_t = t; }
#if !SVR2_V6_FLAT_WALK
          }
          g0 += (uint32_t)it_n_kv;
#endif
        }
        if constexpr (kProf) {
            if (blockIdx.x < 2 && (threadIdx.x & 127) == 0 && p.prof != nullptr) {
                prof_acc[1] = clock64() - prof_t0;
                for (int q = 0; q < 8; ++q) p.prof[(blockIdx.x == 0 ? 0 : 24) + 8 * wg + q] = prof_acc[q];
                if (blockIdx.x == 0) for (int q = 0; q < 4; ++q) p.prof[40 + 4 * wg + q] = ph_acc[q];
                if (blockIdx.x == 0) for (int q = 0; q < 6; ++q) p.prof[48 + 8 * wg + q] = x_acc[q];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();              // nothing of this CTA may disappear while the peer can still signal or read it
    if (warp == 1) {
        __syncwarp();
        tmem_dealloc_pair(tmem, TMEM_COLS);
    }
}

}  // namespace tc6


int launch_window_attn_tc6(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int n_windows, int H, int W, int txt_len, int heads, int64_t P, void* out_vid,
                           void* txt_part, cudaStream_t st) {
    (void)n_windows;
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap_box(&tq, q, P, heads, 128)) != SVR2_OK) return rc;
    if ((rc = make_tmap_box(&tk, k, P, heads, 64)) != SVR2_OK) return rc;      // half key tiles
    if ((rc = make_tmap_box(&tv, v, P, heads, 128)) != SVR2_OK) return rc;     // one 64-column box = half of the d columns
    static PerDeviceFlag smem_ok[2];
    if ((rc = ensure_dyn_smem(tc6::window_attn_tc6_kernel<false>, tc6::SMEM_BYTES, smem_ok[0], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    if ((rc = ensure_dyn_smem(tc6::window_attn_tc6_kernel<true>, tc6::SMEM_BYTES, smem_ok[1], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    tc6::Params p;
    p.tab = tab; p.work = work; p.n_items = n_work * heads;
    p.H = H; p.W = W; p.txt_len = txt_len; p.heads = heads;
    p.out_vid = (__nv_bfloat16*)out_vid; p.txt_part = (__nv_bfloat16*)txt_part;
    const int pairs = sm_count() / 2;
    const int grid = 2 * (p.n_items < pairs ? p.n_items : pairs);       // one persistent CTA pair per TPC
    p.prof = g_trace;
    if (p.prof != nullptr) tc6::window_attn_tc6_kernel<true><<<grid, tc6::NTHREADS, tc6::SMEM_BYTES, st>>>(tq, tk, tv, p);
    else tc6::window_attn_tc6_kernel<false><<<grid, tc6::NTHREADS, tc6::SMEM_BYTES, st>>>(tq, tk, tv, p);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(tcgen05 cta_group::2, generation 8)");
    return SVR2_OK;
}

}  // namespace svr2
