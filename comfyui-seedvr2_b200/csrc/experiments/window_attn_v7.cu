// svr2_window_attn_fwd, generation 9 — generation 4's two-query-tiles-per-CTA layout on CTA PAIRS (tcgen05.mma.cta_group::2).
//
// Why (runs 46-52): generation 4 is limited by the shared-memory operand traffic of its MMAs, generation 8 (CTA
// pairs, one query tile per CTA shared by two softmax warpgroups) by the fixed-latency glue of handing m and l from one
// warpgroup to the other.  Here every softmax warpgroup OWNS a query tile (no hand-over, generation 4's softmax loop unchanged)
// and every MMA spans the pair (M = 256: tile i of CTA 0 and tile i of CTA 1), so each CTA supplies half of every B operand:
//   S_i = Q_i K^T : A = this CTA's query tile i (smem), B = K tile, this CTA holds N/2 of its key rows
//   O_i += P_i V  : A = this CTA's P_i (TMEM),           B = V tile, this CTA holds 64 of its 128 d columns
// Work item = (window, FOUR 128-row query tiles, head); CTA r of the cluster owns tiles 2 i + r (i = 0, 1).  Tiles beyond the
// sequence are computed and never stored.  Pair-wide protocol as in generation 8: the leader CTA's two MMA warps issue, their
// commits are multicast to both CTAs, both CTAs' TMA bytes complete on the leader's "full" barriers, "P is ready" and "O is
// drained" are counted on the leader's barriers with one remote arrival per warp.
//
// TMEM map (columns, each CTA): S0 [0,128)  S1 [128,256)  O0 [256,384)  O1 [384,512);  P_i aliases the first 64 columns of S_i.
// SMEM map: Q0 | Q1 (32 KB each) | ring[8] of 16-KB half tiles | staging (32 KB) | row sums l[2][2][128] | mbarriers.
#include <cuda.h>
#include <cstdlib>

#include "../common.cuh"
#include "../attn_launch.cuh"
#include "../sm100_pair.cuh"
#include "../sm100_ptx.cuh"
#include "../softmax_row.cuh"

namespace svr2 {
namespace tc7 {

using namespace ptx;
using namespace pairx;
using namespace rowsm;

constexpr int D = SVR2_HEAD_DIM;          // 128
constexpr int BM = 128, BN = 128;
constexpr int TILE_BYTES = BM * D * 2;    // 32768
constexpr int BOX_BYTES = TILE_BYTES / 2; // 128 rows x 64 cols bf16
constexpr int HALF_BYTES = TILE_BYTES / 2;   // this CTA's half of a K tile (64 keys x 128 d) or V tile (128 keys x 64 d)
constexpr int KBOX_BYTES = HALF_BYTES / 2;   // 64 keys x 64 d
constexpr int STAGES = 8;
constexpr int NTHREADS = 512;
#ifndef SVR2_REGS_C      // experiment builds: -DSVR2_REGS_C=.. -DSVR2_REGS_E=.. -DSVR2_REGS_S=..
#define SVR2_REGS_C 56     // run 61: +3-4 % over 72/72/184
#define SVR2_REGS_E 56
#define SVR2_REGS_S 200
#endif
constexpr int REGS_CTRL = SVR2_REGS_C, REGS_EPI = SVR2_REGS_E, REGS_SOFTMAX = SVR2_REGS_S;   // 4 x (C + E + 2 S) x 32 = 65,536
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_STAGE = OFF_RING + STAGES * HALF_BYTES;
constexpr int OFF_L = OFF_STAGE + TILE_BYTES;          // float l[2 parities][2 tiles][128 rows]
constexpr int OFF_BAR = OFF_L + 2 * 2 * 128 * 4;
constexpr int SMEM_BYTES = OFF_BAR + 256;
constexpr uint32_t TMEM_COLS = 512;
constexpr int kFmaEighths = 3;
__device__ __forceinline__ float ref_of(float row_max) { return row_max * kScaleLog2; }
constexpr float kRescaleThreshold = 8.0f;

// per query tile i: slot + i.  [L] lives in each CTA (local arrivals or multicast commits); [G] only the leader's copy is used
enum { BAR_QFULL = 0 /*G*/, BAR_QEMPTY = 2 /*L*/, BAR_SFULL = 4 /*L*/, BAR_PREADY = 6 /*G*/, BAR_OFINAL = 8 /*L*/, BAR_OEMPTY = 10 /*G*/,
       BAR_LREADY = 12 /*L*/, BAR_KVFULL = 14 /*G*/, BAR_KVEMPTY = 14 + STAGES /*L*/, BAR_COUNT = 14 + 2 * STAGES };
static_assert(BAR_COUNT * 8 + 8 <= 256, "barrier block too small");
static_assert(SMEM_BYTES <= 232448, "shared memory");

struct Params {
    const int32_t* tab;
    const int32_t* work;   // (window, quad of query tiles) entries
    int n_items;           // work-list entries x heads
    int H, W, txt_len, heads;
    __nv_bfloat16* out_vid;
    __nv_bfloat16* txt_part;
};

// Everything a role needs to know about one work item; every role of both CTAs decodes it identically (q0 differs by rank).
struct Item {
    Win w;
    int wi, head, seq, q0, nq, n_kv, last_cols;
};
__device__ __forceinline__ Item decode_item(const Params& p, int item, int rank) {
    Item it;
    const int e = item / p.heads;
    it.head = item - e * p.heads;
    it.wi = __ldg(p.work + 2 * e);
    const int quad = __ldg(p.work + 2 * e + 1);
    it.w = load_win(p.tab, it.wi);
    it.seq = it.w.n_vid + p.txt_len;
    const int base = quad * 4 * BM;
    it.q0 = base + rank * BM;                          // local tile i starts at q0 + 2 i BM
    it.nq = (it.seq - base > 2 * BM) ? 2 : 1;          // MMA i spans global tiles 2 i and 2 i + 1
    it.n_kv = (it.seq + BN - 1) / BN;
    const int rem = it.seq - (it.n_kv - 1) * BN;
    it.last_cols = (rem + 31) & ~31;
    return it;
}

#define SOFTMAX_WAIT(b, par) mbar_wait(b, par)
#define OTHER_WAIT(b, par) mbar_wait(b, par)

template <int N> __device__ __forceinline__ void reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
window_attn_tc7_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const Params p) {
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t base = smem_u32(smem);
    const uint32_t sRing = base + OFF_RING;
    const uint32_t sBar = base + OFF_BAR;
    float* l_buf = reinterpret_cast<float*>(smem + OFF_L);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + OFF_BAR + BAR_COUNT * 8);
    auto bar = [&](int slot) { return sBar + 8u * (uint32_t)slot; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_rank();
    const bool leader = rank == 0;
    const int first = (int)(blockIdx.x >> 1), stride = (int)(gridDim.x >> 1);

    if (threadIdx.x == 0) {
        if (base & 1023u) __trap();
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(BAR_QFULL + i), 1);
            mbar_init(bar(BAR_QEMPTY + i), 1);
            mbar_init(bar(BAR_SFULL + i), 1);
            mbar_init(bar(BAR_PREADY + i), 8);        // one remote arrival per softmax warp of the pair
            mbar_init(bar(BAR_OFINAL + i), 1);
            mbar_init(bar(BAR_OEMPTY + i), 8);        // one remote arrival per epilogue warp of the pair
            mbar_init(bar(BAR_LREADY + i), 128);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_KVFULL + s), 1);
            mbar_init(bar(BAR_KVEMPTY + s), 2);       // one multicast commit per MMA warp of the leader
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
            // -------------------------------------------------------------- TMA producer: this CTA's query tiles and halves
            if (lane == 0) {
                uint32_t kv_n = 0, c0 = 0, c1 = 0;
                for (int item = first; item < p.n_items; item += stride) {
                    const Item it = decode_item(p, item, rank);
                    for (int i = 0; i < it.nq; ++i) {
                        const int row = it.w.row_start + it.q0 + 2 * i * BM;
                        const uint32_t c = i ? c1 : c0;
                        mbar_wait(bar(BAR_QEMPTY + i), (c & 1) ^ 1);          // local copy of the multicast commit
                        if (leader) mbar_expect_tx(bar(BAR_QFULL + i), 2 * TILE_BYTES);
                        tma_load_3d_pair(base + i * TILE_BYTES, &tm_q, bar(BAR_QFULL + i), 0, it.head, row);
                        tma_load_3d_pair(base + i * TILE_BYTES + BOX_BYTES, &tm_q, bar(BAR_QFULL + i), 64, it.head, row);
                        if (i) ++c1; else ++c0;
                    }
                    for (int n = 0; n < 2 * it.n_kv; ++n, ++kv_n) {           // ring order K0 V0 K1 V1 ...
                        const uint32_t st = kv_n % STAGES, use = kv_n / STAGES;
                        const int j = n >> 1;
                        mbar_wait(bar(BAR_KVEMPTY + st), (use & 1) ^ 1);
                        if (leader) mbar_expect_tx(bar(BAR_KVFULL + st), 2 * HALF_BYTES);
                        if (n & 1) {
                            // this CTA's half of the value tile's d columns: d [64 rank, 64 rank + 64), all 128 keys
                            tma_load_3d_pair(sRing + st * HALF_BYTES, &tm_v, bar(BAR_KVFULL + st), 64 * rank, it.head,
                                             it.w.row_start + j * BN);
                        } else {
                            // this CTA's half of the key tile's N columns: key rows [rank N/2, (rank + 1) N/2); the box is 64
                            // rows, the MMA reads only the first N/2 of them
                            const int ncols = (j == it.n_kv - 1) ? it.last_cols : BN;
                            const int row = it.w.row_start + j * BN + rank * (ncols >> 1);
                            tma_load_3d_pair(sRing + st * HALF_BYTES, &tm_k, bar(BAR_KVFULL + st), 0, it.head, row);
                            tma_load_3d_pair(sRing + st * HALF_BYTES + KBOX_BYTES, &tm_k, bar(BAR_KVFULL + st), 64, it.head, row);
                        }
                    }
                }
            }
        } else if (warp <= 2 && leader) {
            // -------------------------------------------------------------- tcgen05.mma issuer of the pair for query tiles i
            // (warp-uniform control flow, one elected lane issues: see window_attn_v2.cu)
            constexpr uint32_t idesc_pv = idesc_bf16(2 * BM, D, 1);
            constexpr uint32_t hi_sw = smem_desc_hi_sw128(1024);
            const int i = warp - 1;
            auto issue_qk = [&](uint32_t kbase, int ncols) {
                const uint32_t idesc_qk = idesc_bf16(2 * BM, ncols, 0);
                const uint32_t qlo = smem_desc_lo(base + i * TILE_BYTES, 16);
                const uint32_t klo = smem_desc_lo(kbase, 16);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t qoff = ((uint32_t)(kk >> 2) * BOX_BYTES + (uint32_t)(kk & 3) * 32u) >> 4;
                    const uint32_t koff = ((uint32_t)(kk >> 2) * KBOX_BYTES + (uint32_t)(kk & 3) * 32u) >> 4;
                    mma2_ss(tmem + i * 128, make_desc(qlo + qoff, hi_sw), make_desc(klo + koff, hi_sw), idesc_qk, kk > 0);
                }
            };
            auto issue_pv = [&](uint32_t vbase, int ksteps, bool first_tile) {
                const uint32_t vlo = smem_desc_lo(vbase, BOX_BYTES);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk)
                    if (kk < ksteps)
                        mma2_ts(tmem + 256 + i * 128, tmem + i * 128 + kk * 8, make_desc(vlo + kk * (2048 >> 4), hi_sw),
                                idesc_pv, !(first_tile && kk == 0));
            };
            uint32_t kv0 = 0;            // ring index of the current item's K tile 0 (K_t at 2t, V_t at 2t+1)
            uint32_t cq = 0;             // items in which this query tile took part
            uint32_t pr = 0;             // PREADY_i phases consumed
            auto ring_slot = [&](uint32_t idx) { return (kv0 + idx) % STAGES; };
            auto ring_par = [&](uint32_t idx) { return ((kv0 + idx) / STAGES) & 1u; };
            for (int item = first; item < p.n_items; item += stride) {
                const Item it = decode_item(p, item, rank);
                if (i >= it.nq) {
                    // tiles 2 and 3 absent: walk the item's K/V fills and release them with empty commits (see window_attn_v2.cu)
                    for (uint32_t n = 0; n < 2u * it.n_kv; ++n) {
                        OTHER_WAIT(bar(BAR_KVFULL + ring_slot(n)), ring_par(n));
                        if (elect_one()) tc_commit_pair(bar(BAR_KVEMPTY + ring_slot(n)));
                        __syncwarp();
                    }
                    kv0 += 2 * it.n_kv;
                    continue;
                }
                auto cols = [&](int j) { return j == it.n_kv - 1 ? it.last_cols : BN; };
                auto release = [&](uint32_t idx) { tc_commit_pair(bar(BAR_KVEMPTY + ring_slot(idx))); };
                OTHER_WAIT(bar(BAR_QFULL + i), cq & 1);
                OTHER_WAIT(bar(BAR_KVFULL + ring_slot(0)), ring_par(0));
                tc_fence_after();
                if (elect_one()) {
                    issue_qk(sRing + ring_slot(0) * HALF_BYTES, cols(0));
                    tc_commit_pair(bar(BAR_SFULL + i));
                    if (it.n_kv == 1) tc_commit_pair(bar(BAR_QEMPTY + i));
                    release(0);
                }
                __syncwarp();
                for (int j = 0; j < it.n_kv; ++j) {
                    const bool more = j + 1 < it.n_kv;
                    OTHER_WAIT(bar(BAR_KVFULL + ring_slot(2 * j + 1)), ring_par(2 * j + 1));              // V_j
                    if (more) OTHER_WAIT(bar(BAR_KVFULL + ring_slot(2 * j + 2)), ring_par(2 * j + 2));    // K_{j+1}
                    OTHER_WAIT(bar(BAR_PREADY + i), pr & 1u);
                    ++pr;
                    if (j == 0) OTHER_WAIT(bar(BAR_OEMPTY + i), (cq & 1) ^ 1);            // both epilogues drained O_i of the previous item
                    tc_fence_after();
                    if (elect_one()) {
                        issue_pv(sRing + ring_slot(2 * j + 1) * HALF_BYTES, cols(j) >> 4, j == 0);
                        if (more) {
                            issue_qk(sRing + ring_slot(2 * j + 2) * HALF_BYTES, cols(j + 1));
                            tc_commit_pair(bar(BAR_SFULL + i));
                            if (j + 2 == it.n_kv) tc_commit_pair(bar(BAR_QEMPTY + i));
                        } else {
                            tc_commit_pair(bar(BAR_OFINAL + i));
                        }
                        release(2 * j + 1);
                        if (more) release(2 * j + 2);
                    }
                    __syncwarp();
                }
                kv0 += 2 * it.n_kv;
                ++cq;
            }
        }
    } else if (warp < 8) {
        reg_dealloc<REGS_EPI>();
        // ------------------------------------------------------------------ epilogue warps (both query tiles, in turn)
        const int quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        unsigned char* stage = smem + OFF_STAGE;
        const int64_t row_stride = (int64_t)p.heads * D;
        uint32_t c0 = 0, c1 = 0;
        const uint32_t oempty_g[2] = {mapa(bar(BAR_OEMPTY + 0), 0), mapa(bar(BAR_OEMPTY + 1), 0)};
        for (int item = first; item < p.n_items; item += stride) {
            const Item it = decode_item(p, item, rank);
            for (int i = 0; i < it.nq; ++i) {
                const uint32_t cc = i ? c1 : c0;
                const uint32_t tO = tmem + lane_off + 256 + i * 128;
                OTHER_WAIT(bar(BAR_LREADY + i), cc & 1);
                OTHER_WAIT(bar(BAR_OFINAL + i), cc & 1);
                tc_fence_after();
                const float inv = 1.0f / l_buf[((cc & 1) * 2 + i) * 128 + row];
#pragma unroll 1
                for (int c4 = 0; c4 < 4; ++c4) {
                    uint32_t ov[32];
                    tmem_ld32(tO + c4 * 32, ov);
                    tc_wait_ld();
                    if (c4 == 3) {                                   // O_i is in registers: the next item may overwrite it
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive_cluster(oempty_g[i]);
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
                const int jrow = it.q0 + 2 * i * BM + row;
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
                if (i) ++c1; else ++c0;
            }
        }
    } else {
        reg_alloc<REGS_SOFTMAX>();
        // ------------------------------------------------------------------ softmax group i: one thread per query row
        const int i = (warp - 8) >> 2, quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        const uint32_t tS = tmem + lane_off + i * 128;
        const uint32_t tO = tmem + lane_off + 256 + i * 128;
        uint32_t sf = 0;            // SFULL_i phases consumed
        uint32_t c_cnt = 0;
        const uint32_t pready_g = mapa(bar(BAR_PREADY + i), 0);
        for (int item = first; item < p.n_items; item += stride) {
            const Item it = decode_item(p, item, rank);
            if (i >= it.nq) continue;
            float m = -INFINITY, l = 0.f;
            for (int j = 0; j < it.n_kv; ++j, ++sf) {
                const bool last = j == it.n_kv - 1;
                const int nch = (last ? it.last_cols : BN) >> 5;            // 32-column chunks in this key tile (warp-uniform)
                const int valid = it.seq - j * BN;                          // key columns >= valid are beyond the sequence
                SOFTMAX_WAIT(bar(BAR_SFULL + i), sf & 1u);
                tc_fence_after();
                // The whole S row in registers: ONE TMEM round trip per key tile.  Register block b holds chunk nch - 4 + b,
                // so the chunk with the end of the sequence is always block 3 and the masking code exists once.
                uint32_t s[128];
                const uint32_t tS3 = tS + 32 * (nch - 1);
                if (nch > 3) tmem_ld32(tS3 - 96, s);
                if (nch > 2) tmem_ld32(tS3 - 64, s + 32);
                if (nch > 1) tmem_ld32(tS3 - 32, s + 64);
                tmem_ld32(tS3, s + 96);
                tc_wait_ld();
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
                tc_wait_st();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(pready_g);        // counted over both CTAs on the leader's barrier
            }
            // publish the row sum for the epilogue warps (double-buffered by item parity)
            l_buf[((c_cnt & 1) * 2 + i) * 128 + row] = l;
            mbar_arrive(bar(BAR_LREADY + i));
            ++c_cnt;
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

}  // namespace tc7


int launch_window_attn_tc7(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work4,
                           int n_work4, int H, int W, int txt_len, int heads, int64_t P, void* out_vid,
                           void* txt_part, cudaStream_t st) {
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap_box(&tq, q, P, heads, 128)) != SVR2_OK) return rc;
    if ((rc = make_tmap_box(&tk, k, P, heads, 64)) != SVR2_OK) return rc;      // half key tiles
    if ((rc = make_tmap_box(&tv, v, P, heads, 128)) != SVR2_OK) return rc;     // one 64-column box = half of the d columns
    static PerDeviceFlag smem_ok[1];
    if ((rc = ensure_dyn_smem(tc7::window_attn_tc7_kernel, tc7::SMEM_BYTES, smem_ok[0], "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    tc7::Params p;
    p.tab = tab; p.work = work4; p.n_items = n_work4 * heads;
    p.H = H; p.W = W; p.txt_len = txt_len; p.heads = heads;
    p.out_vid = (__nv_bfloat16*)out_vid; p.txt_part = (__nv_bfloat16*)txt_part;
    const int pairs = sm_count() / 2;
    const int grid = 2 * (p.n_items < pairs ? p.n_items : pairs);       // one persistent CTA pair per TPC
    tc7::window_attn_tc7_kernel<<<grid, tc7::NTHREADS, tc7::SMEM_BYTES, st>>>(tq, tk, tv, p);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(tcgen05 cta_group::2, generation 9)");
    return SVR2_OK;
}

}  // namespace svr2
