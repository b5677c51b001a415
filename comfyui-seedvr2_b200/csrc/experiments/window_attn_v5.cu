// svr2_window_attn_fwd, generation 7 — ONE query tile per work item, S and P double-buffered in TMEM, two softmax
// warpgroups taking alternate key tiles of the same rows.
//
// Why (measurements of generation 4, window_attn_v2.cu, profiles/r01_attn_v2_fine_timeline_c1.txt): with two query tiles
// per CTA all 512 TMEM columns hold S0 S1 O0 O1, P must alias S, and each tile is a serial chain
//     softmax(j) -> [P V(j), Q K^T(j+1)] -> softmax(j+1)          (3,270 + 1,340 + ~100 cycles per key tile)
// two chains share one tensor core and one set of schedulers: 4,700 cycles per pair of key tiles against 2,172 of tensor
// work.  Here a CTA keeps ONE 128-row query tile: S0 S1 | P0 P1 | O = 128+128+64+64+128 columns.  P no longer aliases S,
// so Q K^T(g+2) is issued as soon as the softmax of step g has pulled S(g) into registers, and the two warpgroups always
// find their next S waiting: the softmax runs back to back and the tensor core works underneath it.
//
// The same rows are handled by both warpgroups (step g -> warpgroup g & 1), so the running reference exponent m and the
// row sum l travel through shared memory: step g reads m(g-1) right after its own row max (published early in step
// g-1), and l(g-1) at its very end.  The arithmetic — exponent reference, lazy rescale, summation order — is generation
// 4's, so the output bits are identical (tests/test_gpu_kernels.py::test_window_attn_generations_bit_identical).
//
// Steps are numbered per CTA across work items (g = 0, 1, 2, ...): buffers, warpgroups and barrier phases are functions
// of g alone, and the tensor-core program simply runs  QK(g+2), PV(g)  for g = 0, 1, ... across item boundaries.
//   TMEM columns: S[b] = 128 b, P[b] = 256 + 64 b, O = 384            (b = g & 1)
//   SMEM: Q[2] | K/V ring[4] | staging   (32 KB each)  | m[128] l[128] l_final[2][128] | mbarriers
// Ring order = issue order: K(0), K(1), then K(g+2), V(g) for g = 0, 1, ...
#include <cuda.h>
#include <cstdlib>

#include "../common.cuh"
#include "../attn_launch.cuh"
#include "../sm100_ptx.cuh"
#include "../softmax_row.cuh"

namespace svr2 {
namespace tc5 {

using namespace ptx;
using namespace rowsm;

constexpr int D = SVR2_HEAD_DIM;          // 128
constexpr int BM = 128, BN = 128;
constexpr int TILE_BYTES = BM * D * 2;    // 32768
constexpr int BOX_BYTES = TILE_BYTES / 2; // 128 rows x 64 cols bf16
constexpr int STAGES = 4;
constexpr int NTHREADS = 512;
constexpr int REGS_CTRL = 72, REGS_EPI = 72, REGS_SOFTMAX = 184;      // 32 * (4 * 72 + 4 * 72 + 8 * 184) = 65536
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_STAGE = OFF_RING + STAGES * TILE_BYTES;
// m and l need one slot per row each: reads and writes of the two warpgroups strictly alternate (step g+1 reads m(g), then
// writes m(g+1) to the same word; step g keeps its m in a register)
constexpr int OFF_M = OFF_STAGE + TILE_BYTES;          // float m[128]
constexpr int OFF_L = OFF_M + 128 * 4;                 // float l[128]
constexpr int OFF_LFIN = OFF_L + 128 * 4;              // float l_final[2][128]  (by item parity, for the epilogue)
constexpr int OFF_BAR = OFF_LFIN + 2 * 128 * 4;
constexpr int SMEM_BYTES = OFF_BAR + 256;
constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t COL_P = 256, COL_O = 384;
constexpr int kFmaEighths = 3;            // of every 8 exp2 pairs, this many on the FMA pipe (softmax_row.cuh)
constexpr float kRescaleThreshold = 8.0f;
__device__ __forceinline__ float ref_of(float row_max) { return row_max * kScaleLog2; }

enum { BAR_QFULL = 0, BAR_QEMPTY = 2, BAR_SFULL = 4, BAR_SLOADED = 6, BAR_PREADY = 8, BAR_PVDONE = 10, BAR_MPUB = 12,
       BAR_LPUB = 14, BAR_OFINAL = 16, BAR_OEMPTY = 17, BAR_LFINAL = 18, BAR_KVFULL = 19, BAR_KVEMPTY = 19 + STAGES,
       BAR_COUNT = 19 + 2 * STAGES };
static_assert(BAR_COUNT * 8 + 8 <= 256, "barrier block too small");
static_assert(SMEM_BYTES <= 232448, "shared memory");

struct Params {
    const int32_t* tab;
    const int32_t* work;
    int n_items;           // work-list entries x heads x 2 (query tiles of a pair)
    int H, W, txt_len, heads;
    __nv_bfloat16* out_vid;
    __nv_bfloat16* txt_part;
    long long* prof;       // kProf build: per-role wait / work cycle counters of CTA 0 (svr2_debug_attn_trace buffer)
};

struct Item {
    Win w;
    int wi, head, seq, q0, n_kv, last_cols;
};
// item = ((entry * heads + head) << 1) | tile of the pair; false when the pair has no second tile
__device__ __forceinline__ bool decode_item(const Params& p, int item, Item& it) {
    const int half = item & 1, eh = item >> 1;
    const int e = eh / p.heads;
    it.head = eh - e * p.heads;
    it.wi = __ldg(p.work + 2 * e);
    const int qpair = __ldg(p.work + 2 * e + 1);
    it.w = load_win(p.tab, it.wi);
    it.seq = it.w.n_vid + p.txt_len;
    it.q0 = (2 * qpair + half) * BM;
    it.n_kv = (it.seq + BN - 1) / BN;
    const int rem = it.seq - (it.n_kv - 1) * BN;       // 1..128 valid key columns in the last tile
    it.last_cols = (rem + 31) & ~31;                   // trimmed to a multiple of 32 (MMA N, softmax chunks, PV K steps)
    return it.q0 < it.seq;
}

// Walks the (item, key tile) steps of this CTA in order; every role runs the same walk.
struct Cursor {
    int item;       // global item index
    int n;          // ordinal of the item among this CTA's existing items
    int j;          // key tile inside the item
    bool valid;
    Item it;
};
__device__ __forceinline__ void cur_seek(const Params& p, Cursor& c) {
    c.valid = false;
    while (c.item < p.n_items) {
        if (decode_item(p, c.item, c.it)) {
            c.valid = true;
            return;
        }
        c.item += gridDim.x;
    }
}
__device__ __forceinline__ void cur_init(const Params& p, Cursor& c) {
    c.item = blockIdx.x; c.n = 0; c.j = 0;
    cur_seek(p, c);
}
__device__ __forceinline__ void cur_next(const Params& p, Cursor& c) {
    if (++c.j == c.it.n_kv) {
        c.j = 0; ++c.n; c.item += gridDim.x;
        cur_seek(p, c);
    }
}

template <int N> __device__ __forceinline__ void reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

// kProf: CTA 0 accumulates the cycles each role spends in each of its waits (clock64 around the wait) and writes them to
// p.prof at the end — [wg][0..7] = steps, total, SFULL, MPUB, PVDONE(P free), LPUB, OEMPTY, rescales; [2][..] = MMA warp:
// steps, total, SLOADED, KVFULL(K), KVFULL(V), PREADY, OEMPTY, QFULL.  The production instantiation carries none of it.
#define PROF_WAIT(slot, stmt) do { if constexpr (kProf) { const long long t0_ = clock64(); stmt; prof_acc[slot] += clock64() - t0_; } else { stmt; } } while (0)
template <bool kProf>
__global__ void __launch_bounds__(NTHREADS, 1)
window_attn_tc5_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t base = smem_u32(smem);                          // SWIZZLE_128B tiles need 1024-B alignment
    const uint32_t sRing = base + OFF_RING;
    const uint32_t sBar = base + OFF_BAR;
    float* m_sh = reinterpret_cast<float*>(smem + OFF_M);
    float* l_sh = reinterpret_cast<float*>(smem + OFF_L);
    float* l_fin = reinterpret_cast<float*>(smem + OFF_LFIN);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + OFF_BAR + BAR_COUNT * 8);
    auto bar = [&](int slot) { return sBar + 8u * (uint32_t)slot; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        if (base & 1023u) __trap();
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(BAR_QFULL + i), 1);
            mbar_init(bar(BAR_QEMPTY + i), 1);
            mbar_init(bar(BAR_SFULL + i), 1);
            mbar_init(bar(BAR_SLOADED + i), 128);
            mbar_init(bar(BAR_PREADY + i), 128);
            mbar_init(bar(BAR_PVDONE + i), 1);
            mbar_init(bar(BAR_MPUB + i), 128);
            mbar_init(bar(BAR_LPUB + i), 128);
        }
        mbar_init(bar(BAR_OFINAL), 1);
        mbar_init(bar(BAR_OEMPTY), 128);
        mbar_init(bar(BAR_LFINAL), 128);
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_KVFULL + s), 1);
            mbar_init(bar(BAR_KVEMPTY + s), 1);
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

    if (warp < 4) {
        reg_dealloc<REGS_CTRL>();
        if (warp == 0) {
            // -------------------------------------------------------------- TMA producer
            if (lane == 0) {
                uint32_t kv_n = 0;
                auto load_tile = [&](const CUtensorMap* map, int head, int row) {
                    const uint32_t st = kv_n % STAGES, use = kv_n / STAGES;
                    mbar_wait(bar(BAR_KVEMPTY + st), (use & 1) ^ 1);
                    mbar_expect_tx(bar(BAR_KVFULL + st), TILE_BYTES);
                    tma_load_3d(sRing + st * TILE_BYTES, map, bar(BAR_KVFULL + st), 0, head, row);
                    tma_load_3d(sRing + st * TILE_BYTES + BOX_BYTES, map, bar(BAR_KVFULL + st), 64, head, row);
                    ++kv_n;
                };
                auto load_k = [&](const Cursor& c) {
                    if (c.j == 0) {                                     // the item's query tile first
                        const uint32_t qb = c.n & 1;
                        if (c.n >= 2) mbar_wait(bar(BAR_QEMPTY + qb), ((c.n >> 1) - 1) & 1);      // item n-2 is done with Q[qb]
                        mbar_expect_tx(bar(BAR_QFULL + qb), TILE_BYTES);
                        const int row = c.it.w.row_start + c.it.q0;
                        tma_load_3d(base + qb * TILE_BYTES, &tm_q, bar(BAR_QFULL + qb), 0, c.it.head, row);
                        tma_load_3d(base + qb * TILE_BYTES + BOX_BYTES, &tm_q, bar(BAR_QFULL + qb), 64, c.it.head, row);
                    }
                    load_tile(&tm_k, c.it.head, c.it.w.row_start + c.j * BN);
                };
                Cursor kc, vc;
                cur_init(p, kc);
                vc = kc;
                for (int a = 0; a < 2 && kc.valid; ++a) {       // K runs two steps ahead of V (see the MMA issuer)
                    load_k(kc);
                    cur_next(p, kc);
                }
                while (vc.valid) {
                    if (kc.valid) {
                        load_k(kc);
                        cur_next(p, kc);
                    }
                    load_tile(&tm_v, vc.it.head, vc.it.w.row_start + vc.j * BN);
                    cur_next(p, vc);
                }
            }
        } else if (warp == 1) {
            // -------------------------------------------------------------- tcgen05.mma issuer (whole warp, one elected lane)
            constexpr uint32_t idesc_pv = idesc_bf16(BM, D, 1);
            constexpr uint32_t hi_sw = smem_desc_hi_sw128(1024);
            uint32_t kv_n = 0;       // ring tiles consumed
            uint32_t g_qk = 0, g_pv = 0;
            long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const long long prof_t0 = kProf ? clock64() : 0;
            // S[b] = Q K^T : 8 x (128 x ncols x 16), SS
            auto do_qk = [&](const Cursor& c) {
                const uint32_t b = g_qk & 1, qb = c.n & 1;
                const int ncols = (c.j == c.it.n_kv - 1) ? c.it.last_cols : BN;
                const uint32_t st = kv_n % STAGES, par = (kv_n / STAGES) & 1;
                if (c.j == 0) PROF_WAIT(7, mbar_wait(bar(BAR_QFULL + qb), (c.n >> 1) & 1));
                if (g_qk >= 2) PROF_WAIT(2, mbar_wait(bar(BAR_SLOADED + b), ((g_qk >> 1) - 1) & 1));      // softmax(g-2) has S[b] in registers
                PROF_WAIT(3, mbar_wait(bar(BAR_KVFULL + st), par));
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t idesc_qk = idesc_bf16(BM, ncols, 0);
                    const uint32_t qlo = smem_desc_lo(base + qb * TILE_BYTES, 16);
                    const uint32_t klo = smem_desc_lo(sRing + st * TILE_BYTES, 16);
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        const uint32_t off = ((uint32_t)(kk >> 2) * BOX_BYTES + (uint32_t)(kk & 3) * 32u) >> 4;
                        mma_ss(tmem + b * 128, make_desc(qlo + off, hi_sw), make_desc(klo + off, hi_sw), idesc_qk, kk > 0);
                    }
                    tc_commit(bar(BAR_SFULL + b));
                    tc_commit(bar(BAR_KVEMPTY + st));
                    if (c.j == c.it.n_kv - 1) tc_commit(bar(BAR_QEMPTY + qb));           // last read of this item's Q
                }
                __syncwarp();
                ++kv_n; ++g_qk;
            };
            // O += P[b] V : A = P from TMEM (16 keys = 8 columns per step), B = V MN-major
            auto do_pv = [&](const Cursor& c) {
                const uint32_t b = g_pv & 1;
                const bool last = c.j == c.it.n_kv - 1;
                const int ksteps = (last ? c.it.last_cols : BN) >> 4;
                const uint32_t st = kv_n % STAGES, par = (kv_n / STAGES) & 1;
                PROF_WAIT(4, mbar_wait(bar(BAR_KVFULL + st), par));
                PROF_WAIT(5, mbar_wait(bar(BAR_PREADY + b), (g_pv >> 1) & 1));
                if (c.j == 0 && c.n > 0) PROF_WAIT(6, mbar_wait(bar(BAR_OEMPTY), (c.n - 1) & 1));       // epilogue drained the previous item's O
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t vlo = smem_desc_lo(sRing + st * TILE_BYTES, BOX_BYTES);
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk)
                        if (kk < ksteps)
                            mma_ts(tmem + COL_O, tmem + COL_P + b * 64 + kk * 8, make_desc(vlo + kk * (2048 >> 4), hi_sw),
                                   idesc_pv, !(c.j == 0 && kk == 0));
                    tc_commit(bar(BAR_PVDONE + b));
                    tc_commit(bar(BAR_KVEMPTY + st));
                    if (last) tc_commit(bar(BAR_OFINAL));
                }
                __syncwarp();
                ++kv_n; ++g_pv;
            };
            // Program: QK(0), QK(1), then  QK(g+2), PV(g)  for g = 0, 1, ...  QK(g+2) only needs S[g & 1] to be in the
            // registers of step g's softmax (early in that step), so it runs UNDER that softmax and S(g+2) is waiting when the
            // warpgroup comes back for it; PV(g) follows when P(g) is published.
            Cursor kc, vc;
            cur_init(p, kc);
            vc = kc;
            for (int a = 0; a < 2 && kc.valid; ++a) {
                do_qk(kc);
                cur_next(p, kc);
            }
            while (vc.valid) {
                if (kc.valid) {
                    do_qk(kc);
                    cur_next(p, kc);
                }
                do_pv(vc);
                cur_next(p, vc);
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
        // ------------------------------------------------------------------ epilogue warps
        const int quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        unsigned char* stage = smem + OFF_STAGE;
        const int64_t row_stride = (int64_t)p.heads * D;
        const uint32_t tO = tmem + lane_off + COL_O;
        Cursor c;
        cur_init(p, c);
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
                    mbar_arrive(bar(BAR_OEMPTY));
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
            __syncwarp();                                           // staging rows of this warp are free again
            // skip to the next item
            c.j = 0; ++c.n; c.item += gridDim.x;
            cur_seek(p, c);
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
        long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const long long prof_t0 = kProf ? clock64() : 0;
        Cursor c;
        cur_init(p, c);
        for (uint32_t g = 0; c.valid; ++g, cur_next(p, c)) {
            if ((g & 1u) != (uint32_t)wg) continue;
            const Item& it = c.it;
            const int j = c.j;
            const bool last = j == it.n_kv - 1;
            const int nch = (last ? it.last_cols : BN) >> 5;            // 32-column chunks in this key tile (warp-uniform)
            const int valid = it.seq - j * BN;                          // key columns >= valid are beyond the sequence
            const uint32_t par = (g >> 1) & 1u;                         // phase parity of this step on the [b] barriers
            const uint32_t ppar = ((g - 1) >> 1) & 1u;                  // phase parity of step g-1 on the [ob] barriers
            PROF_WAIT(2, mbar_wait(bar(BAR_SFULL + b), par));
            if constexpr (kProf) prof_acc[0] += 1;
            tc_fence_after();
            // the whole S row in registers; register block q holds chunk nch - 4 + q (the sequence end is always in block 3)
            uint32_t s[128];
            const uint32_t tS3 = tS + 32 * (nch - 1);
            if (nch > 3) tmem_ld32(tS3 - 96, s);
            if (nch > 2) tmem_ld32(tS3 - 64, s + 32);
            if (nch > 1) tmem_ld32(tS3 - 32, s + 64);
            tmem_ld32(tS3, s + 96);
            tc_wait_ld();
            tc_fence_before();
            mbar_arrive(bar(BAR_SLOADED + b));                          // Q K^T(g+2) may overwrite S[b]
            if (valid < nch * 32) mask_chunk(s + 96, valid - 32 * (nch - 1));
            float m8[8];
#pragma unroll
            for (int q = 0; q < 8; ++q) m8[q] = -INFINITY;
            if (nch > 3) chunk_max8(m8, s);
            if (nch > 2) chunk_max8(m8, s + 32);
            if (nch > 1) chunk_max8(m8, s + 64);
            chunk_max8(m8, s + 96);
            const float mx = fmaxf(max3(m8[0], m8[1], m8[2]), max3(max3(m8[3], m8[4], m8[5]), m8[6], m8[7]));
            // the reference exponent of step g-1 (the other warpgroup published it right after ITS row max).  Waited for at
            // every step, first tiles of an item included, so that the two warpgroups' arrivals strictly alternate and a
            // barrier can never run two phases ahead of a waiter.
            if (g > 0) PROF_WAIT(3, mbar_wait(bar(BAR_MPUB + ob), ppar));
            float m, alpha = 1.0f;
            if (j == 0) {
                m = ref_of(mx);
            } else {
                m = m_sh[row];
                const bool need = mx * kScaleLog2 - m > kRescaleThreshold;
                if (__any_sync(0xffffffffu, need)) {                    // warp-uniform: tcgen05.ld/st are warp collectives
                    // O must hold P V(0..j-1) and nothing newer: P V(g-1) done (P V(g) needs this step's P)
                    mbar_wait(bar(BAR_PVDONE + ob), ppar);
                    if constexpr (kProf) prof_acc[7] += 1;
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
            // P[b] is free once P V(g-2) has read it
            if (g >= 2) PROF_WAIT(4, mbar_wait(bar(BAR_PVDONE + b), (par ^ 1u)));
            const float mb = m;
            const uint32_t tP3 = tP + 16 * (nch - 1);
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
            tc_wait_st();
            tc_fence_before();
            mbar_arrive(bar(BAR_PREADY + b));
            // row sum: l(g) = l(g-1) * alpha + rs, handed from warpgroup to warpgroup (waited for at every step, see above)
            if (g > 0) PROF_WAIT(5, mbar_wait(bar(BAR_LPUB + ob), ppar));
            float l = 0.f;
            if (j > 0) l = l_sh[row];
            l *= alpha;
            l += rs;
            if (last) {
                // the epilogue must be done with item n-1 (its l_final slot is two items old, and LFINAL must not run a
                // phase ahead of it): O of item n-1 drained
                if (c.n > 0) PROF_WAIT(6, mbar_wait(bar(BAR_OEMPTY), (c.n - 1) & 1));
                l_fin[(c.n & 1) * 128 + row] = l;
                mbar_arrive(bar(BAR_LFINAL));
            }
            l_sh[row] = l;
            mbar_arrive(bar(BAR_LPUB + b));
        }
        if constexpr (kProf) {
            if (blockIdx.x == 0 && (threadIdx.x & 127) == 0 && p.prof != nullptr) {
                prof_acc[1] = clock64() - prof_t0;
                for (int q = 0; q < 8; ++q) p.prof[8 * wg + q] = prof_acc[q];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tmem_dealloc(tmem, TMEM_COLS);
    }
}

}  // namespace tc5


int launch_window_attn_tc5(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                           int n_work, int n_windows, int H, int W, int txt_len, int heads, int64_t P, void* out_vid,
                           void* txt_part, cudaStream_t st) {
    (void)n_windows;
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap(&tq, q, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tk, k, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tv, v, P, heads)) != SVR2_OK) return rc;
    static PerDeviceFlag smem_ok[2];
    if ((rc = ensure_dyn_smem(tc5::window_attn_tc5_kernel<false>, tc5::SMEM_BYTES, smem_ok[0], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    if ((rc = ensure_dyn_smem(tc5::window_attn_tc5_kernel<true>, tc5::SMEM_BYTES, smem_ok[1], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    tc5::Params p;
    p.tab = tab; p.work = work; p.n_items = n_work * heads * 2;
    p.H = H; p.W = W; p.txt_len = txt_len; p.heads = heads;
    p.out_vid = (__nv_bfloat16*)out_vid; p.txt_part = (__nv_bfloat16*)txt_part;
    const int sms = sm_count();
    const int grid = p.n_items < sms ? p.n_items : sms;       // one persistent CTA per SM
    p.prof = g_trace;
    if (p.prof != nullptr) tc5::window_attn_tc5_kernel<true><<<grid, tc5::NTHREADS, tc5::SMEM_BYTES, st>>>(tq, tk, tv, p);
    else tc5::window_attn_tc5_kernel<false><<<grid, tc5::NTHREADS, tc5::SMEM_BYTES, st>>>(tq, tk, tv, p);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(tcgen05, generation 7)");
    return SVR2_OK;
}

}  // namespace svr2
