// svr2_window_attn_fwd, generation 2 — PERSISTENT variable-length NaSwin window attention for sm_100a.
//
// Replaces FlashAttentionVarlen / pytorch_varlen_attention (reference src/models/dit_3b/attention.py:24-61,138-157: a
// Python loop of one SDPA call per window), unconcat_win (na.py:396-417) and window_reverse (na.py:638).
//
// Why a second generation: the per-CTA timeline of generation 1 (profiles/r01_run4_attn_v1_timeline_c1.txt) showed that
// with NaSwin's short sequences (4 key tiles per window at 2048^2) only 60 % of a CTA's life is the main loop — the
// first Q/K loads (3300 cycles), the epilogue (4600) and the launch gap between CTAs (2200) were all exposed because one
// CTA owned an SM at a time.  Here ONE CTA PER SM stays resident and walks a static slice of the work list; every role
// runs ahead across item boundaries, so loads of item n+1 and the epilogue of item n-1 overlap the MMAs of item n.
//
// Work item = (window, pair of 128-row query tiles, head); items are sorted longest sequence first and CTA b takes
// items b, b+G, b+2G, ...   Roles (14 warps, 448 threads):
//   warps 0-3   softmax group 0 (query tile 0): one thread per query row; S is read from TMEM ONCE in 32-column chunks
//               (exp2 against a lagging reference max + bf16 P written back over S) so no thread ever holds a whole row
//   warps 4-7   softmax group 1 (query tile 1)
//   warps 8-11  epilogue: O/l -> bf16 -> swizzled staging tile -> coalesced 256-B row stores to the token's ORIGINAL
//               position (window reverse) or the per-window text buffer; frees O for the next item as soon as it is read
//   warp  12    TMA producer (one thread): Q tiles, then K/V through a 4-stage ring of 32-KB tiles, across items
//   warp  13    tcgen05.mma issuer (one thread): S_i = Q_i K^T (SS), O_i += P_i V (A = P from TMEM, B = V MN-major)
// The last key tile of a sequence is trimmed to a multiple of 32 columns (MMA N / K steps and the softmax chunks shrink
// with it), which removes most of the tile-quantisation waste of the 458-token windows.
//
// TMEM map (columns): S0 [0,128)  S1 [128,256)  O0 [256,384)  O1 [384,512);  P_i aliases the first 64 columns of S_i.
// SMEM map: Q0 | Q1 | ring[4] | staging (32 KB each) | row sums l[2][2][128] | mbarriers.
#include <cuda.h>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace svr2 {
namespace tc2 {

using namespace ptx;

constexpr int D = SVR2_HEAD_DIM;          // 128
constexpr int BM = 128, BN = 128;
constexpr int TILE_BYTES = BM * D * 2;    // 32768
constexpr int BOX_BYTES = TILE_BYTES / 2; // 128 rows x 64 cols bf16
constexpr int STAGES = 4;
constexpr int NTHREADS = 448;
constexpr int OFF_RING = 2 * TILE_BYTES;
constexpr int OFF_STAGE = OFF_RING + STAGES * TILE_BYTES;
constexpr int OFF_L = OFF_STAGE + TILE_BYTES;          // float l[2 parities][2 tiles][128 rows]
constexpr int OFF_BAR = OFF_L + 2 * 2 * 128 * 4;
constexpr int SMEM_BYTES = OFF_BAR + 256;              // 231,680 B of the 232,448 B a CTA may opt in to
constexpr uint32_t TMEM_COLS = 512;
constexpr float kScaleLog2 = 0.08838834764831845f * 1.4426950408889634f;   // 1/sqrt(128) * log2(e)
constexpr float kRescaleThreshold = 8.0f;   // lazy rescale: the reference lags the running max by up to 2^8 between tiles
constexpr float kMaxExponent = 100.0f;      // P <= 2^100: a row can never overflow bf16 / the fp32 sums (128 * 2^100 << 2^127)

// barrier slots (8 bytes each)
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
// +0 MMA thread starts the item (Q and K0 landed)   +1 softmax-0 sees S(0)   +2 softmax-0 published P(0)
// +3 softmax-0 sees S(last)   +4 softmax-0 published P(last)   +5 epilogue sees O_0 final   +6 tile-0 stores issued
// +7 tile-1 stores issued   +8 TMA thread issued the item's last K/V load
constexpr int kTraceSlots = 64, kTraceItems = 4;
// fine stamps of item 2, key tile 1 (steady state): 44-51 softmax-0 thread 0 (S seen, chunk-0 landed, after chunks 0..3, after
// wait::st, after arrive), 52-55 MMA thread (P0 ready, tile-0 MMAs issued, P1 ready, tile-1 MMAs issued), 56-58 epilogue tile 0
// (O final seen, staged, stored)
#define SVR2_FINE(slot) do { if (p.trace != nullptr && fine) trace_at(p, slot, clock64()); } while (0)

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

// row max of one 32-column chunk; columns >= n_valid (only possible in the last chunk of a sequence) are ignored
__device__ __forceinline__ float chunk_max(const uint32_t* v, int n_valid) {
    float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
    if (n_valid >= 32) {
#pragma unroll
        for (int c = 0; c < 32; c += 8) {
            m0 = max3(m0, __uint_as_float(v[c]), __uint_as_float(v[c + 1]));
            m1 = max3(m1, __uint_as_float(v[c + 2]), __uint_as_float(v[c + 3]));
            m2 = max3(m2, __uint_as_float(v[c + 4]), __uint_as_float(v[c + 5]));
            m3 = max3(m3, __uint_as_float(v[c + 6]), __uint_as_float(v[c + 7]));
        }
    } else {
#pragma unroll
        for (int c = 0; c < 32; ++c) m0 = fmaxf(m0, c < n_valid ? __uint_as_float(v[c]) : -INFINITY);
    }
    return fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
}
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

// P = exp2(s * k - mb) for one full chunk, packed to bf16 pairs; returns the chunk's row sum.  3 issue slots per element:
// FFMA2 (half), MUFU.EX2, F2FP (half), FADD2 (half) + the FMNMX3 (half) of chunk_max.  The caller guarantees
// s * k - mb <= kMaxExponent for every element (warp vote on the chunk max), so there is no per-element clamp.
__device__ __forceinline__ float chunk_exp(const uint32_t* v, uint32_t* pk, float mb) {
    const uint64_t kk = pack_f2(kScaleLog2, kScaleLog2), nmb = pack_f2(-mb, -mb);
    uint64_t acc0 = pack_f2(0.f, 0.f), acc1 = acc0;
#pragma unroll
    for (int c = 0; c < 16; c += 2) {
        float x0, x1, x2, x3;
        unpack_f2(fma_f2(pack_f2(__uint_as_float(v[2 * c]), __uint_as_float(v[2 * c + 1])), kk, nmb), x0, x1);
        unpack_f2(fma_f2(pack_f2(__uint_as_float(v[2 * c + 2]), __uint_as_float(v[2 * c + 3])), kk, nmb), x2, x3);
        const float p0 = ex2(x0), p1 = ex2(x1), p2 = ex2(x2), p3 = ex2(x3);
        pk[c] = pack2(p0, p1);
        pk[c + 1] = pack2(p2, p3);
        acc0 = add_f2(acc0, pack_f2(p0, p1));
        acc1 = add_f2(acc1, pack_f2(p2, p3));
    }
    float s0, s1;
    unpack_f2(add_f2(acc0, acc1), s0, s1);
    return s0 + s1;
}
// The rare general form: columns >= n_valid masked to zero (last chunk of a sequence) and the exponent clamped.
__device__ __forceinline__ float chunk_exp_slow(const uint32_t* v, uint32_t* pk, float mb, int n_valid) {
    float r0 = 0.f, r1 = 0.f;
#pragma unroll
    for (int c = 0; c < 16; ++c) {
        float p0 = ex2(fminf(fmaf(__uint_as_float(v[2 * c]), kScaleLog2, -mb), kMaxExponent));
        float p1 = ex2(fminf(fmaf(__uint_as_float(v[2 * c + 1]), kScaleLog2, -mb), kMaxExponent));
        p0 = (2 * c < n_valid) ? p0 : 0.f;
        p1 = (2 * c + 1 < n_valid) ? p1 : 0.f;
        r0 += p0;
        r1 += p1;
        pk[c] = pack2(p0, p1);
    }
    return r0 + r1;
}
// one chunk: running max bookkeeping + P; `cm` = this chunk's row max (already masked)
__device__ __forceinline__ float chunk_softmax(const uint32_t* v, uint32_t* pk, float mb, float cm, int n_valid) {
    const bool danger = fmaf(cm, kScaleLog2, -mb) > kMaxExponent;       // reference lags by more than 2^100: never in practice
    if (n_valid >= 32 && !__any_sync(0xffffffffu, danger)) return chunk_exp(v, pk, mb);
    return chunk_exp_slow(v, pk, mb, n_valid);
}

__device__ __forceinline__ void trace_at(const Params& p, int slot, long long v) {
    if (p.trace != nullptr && (int)blockIdx.x < p.trace_ctas && slot < kTraceSlots)
        p.trace[(long long)blockIdx.x * kTraceSlots + slot] = v;
}
__device__ __forceinline__ void trace_item(const Params& p, int n, int what) {
    if (p.trace != nullptr && n < kTraceItems) trace_at(p, 8 + 9 * n + what, clock64());
}

// 14 warps = 4+4+3+3 per SM sub-partition; a sub-partition's 16 K registers / 4 warps caps a thread at 128 registers
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
            mbar_init(bar(BAR_SFULL + i), 1);
            mbar_init(bar(BAR_PREADY + i), 128);
            mbar_init(bar(BAR_OFINAL + i), 1);
            mbar_init(bar(BAR_OEMPTY + i), 128);
            mbar_init(bar(BAR_LREADY + i), 128);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_KVFULL + s), 1);
            mbar_init(bar(BAR_KVEMPTY + s), 1);
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
            uint32_t kv_n = 0, c[2] = {0, 0};
            int n_it = 0;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++n_it) {
                const Item it = decode_item(p, item);
                for (int i = 0; i < it.nq; ++i) {
                    const int row = it.w.row_start + it.q0 + i * BM;
                    mbar_wait(bar(BAR_QEMPTY + i), (c[i] & 1) ^ 1);           // all QK MMAs of the previous user are done
                    mbar_expect_tx(bar(BAR_QFULL + i), TILE_BYTES);
                    tma_load_3d(base + i * TILE_BYTES, &tm_q, bar(BAR_QFULL + i), 0, it.head, row);
                    tma_load_3d(base + i * TILE_BYTES + BOX_BYTES, &tm_q, bar(BAR_QFULL + i), 64, it.head, row);
                    ++c[i];
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
    } else if (warp == 13) {
        // ------------------------------------------------------------------ tcgen05.mma issuer (single thread)
        if (lane == 0) {
            constexpr uint32_t idesc_pv = idesc_bf16(BM, D, 1);
            auto issue_qk = [&](int i, uint32_t kbase, int ncols) {   // S_i = Q_i K^T : 8 x (128 x ncols x 16), both K-major
                const uint32_t idesc_qk = idesc_bf16(BM, ncols, 0);
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t off = (uint32_t)(kk >> 2) * BOX_BYTES + (uint32_t)(kk & 3) * 32u;
                    mma_ss(tmem + i * 128, smem_desc_sw128(base + i * TILE_BYTES + off, 16, 1024),
                           smem_desc_sw128(kbase + off, 16, 1024), idesc_qk, kk > 0);
                }
            };
            auto issue_pv = [&](int i, uint32_t vbase, bool first, int ksteps) {   // O_i += P_i V : A = P (TMEM), B = V MN-major
                for (int kk = 0; kk < ksteps; ++kk)
                    mma_ts(tmem + 256 + i * 128, tmem + i * 128 + kk * 8,
                           smem_desc_sw128(vbase + kk * 2048, BOX_BYTES, 1024), idesc_pv, !(first && kk == 0));
            };
            uint32_t kv_c = 0, c[2] = {0, 0}, s[2] = {0, 0};
            int n_done = 0;
            for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++n_done) {
                const Item it = decode_item(p, item);
                auto cols = [&](int j) { return j == it.n_kv - 1 ? it.last_cols : BN; };
                for (int i = 0; i < it.nq; ++i) mbar_wait(bar(BAR_QFULL + i), c[i] & 1);
                {
                    const uint32_t st = kv_c % STAGES;
                    mbar_wait(bar(BAR_KVFULL + st), (kv_c / STAGES) & 1);
                    tc_fence_after();
                    trace_item(p, n_done, 0);
                    for (int i = 0; i < it.nq; ++i) {
                        issue_qk(i, sRing + st * TILE_BYTES, cols(0));
                        tc_commit(bar(BAR_SFULL + i));
                        if (it.n_kv == 1) tc_commit(bar(BAR_QEMPTY + i));
                    }
                    tc_commit(bar(BAR_KVEMPTY + st));
                    ++kv_c;
                }
                for (int j = 0; j < it.n_kv; ++j) {
                    const bool has_next = j + 1 < it.n_kv;
                    const uint32_t stV = kv_c % STAGES, stK = (kv_c + 1) % STAGES;
                    mbar_wait(bar(BAR_KVFULL + stV), (kv_c / STAGES) & 1);
                    if (has_next) mbar_wait(bar(BAR_KVFULL + stK), ((kv_c + 1) / STAGES) & 1);
                    const bool fine = n_done == 2 && j == 1;
                    for (int i = 0; i < it.nq; ++i) {
                        mbar_wait(bar(BAR_PREADY + i), s[i] & 1);
                        SVR2_FINE(52 + 2 * i);
                        ++s[i];
                        if (j == 0) mbar_wait(bar(BAR_OEMPTY + i), (c[i] & 1) ^ 1);   // epilogue drained O_i of the previous item
                        tc_fence_after();
                        issue_pv(i, sRing + stV * TILE_BYTES, j == 0, cols(j) >> 4);
                        if (has_next) {
                            issue_qk(i, sRing + stK * TILE_BYTES, cols(j + 1));
                            tc_commit(bar(BAR_SFULL + i));
                            if (j + 2 == it.n_kv) tc_commit(bar(BAR_QEMPTY + i));      // that was the last read of Q_i
                        } else {
                            tc_commit(bar(BAR_OFINAL + i));
                        }
                        SVR2_FINE(53 + 2 * i);
                    }
                    tc_commit(bar(BAR_KVEMPTY + stV));
                    ++kv_c;
                    if (has_next) {
                        tc_commit(bar(BAR_KVEMPTY + stK));
                        ++kv_c;
                    }
                }
                for (int i = 0; i < it.nq; ++i) ++c[i];
            }
            trace_at(p, 7, n_done);
        }
    } else if (warp < 8) {
        // ------------------------------------------------------------------ softmax group i
        const int i = warp >> 2, quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        const uint32_t tS = tmem + lane_off + i * 128, tO = tmem + lane_off + 256 + i * 128;
        uint32_t s_cnt = 0, c_cnt = 0;
        int n_it = -1;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x) {
            const Item it = decode_item(p, item);
            ++n_it;
            if (i >= it.nq) continue;
            float m = -INFINITY, m_seen = -INFINITY, l = 0.f;
            for (int j = 0; j < it.n_kv; ++j) {
                const bool last = j == it.n_kv - 1;
                const int n_ch = (last ? it.last_cols : BN) >> 5;         // 32-column chunks of this key tile
                const int valid = it.seq - j * BN;                        // key columns >= valid are beyond the sequence
                mbar_wait(bar(BAR_SFULL + i), s_cnt & 1);
                ++s_cnt;
                tc_fence_after();
                if (threadIdx.x == 0) {
                    if (j == 0) trace_item(p, n_it, 1);
                    if (last) trace_item(p, n_it, 3);
                }
                // S is read from TMEM exactly ONCE (TMEM -> registers moves 64 B/clk/SM: one 128x128 fp32 tile costs as many
                // cycles as its two MMAs), chunk by chunk with the next chunk in flight.  The exponent reference `m` is
                // therefore not the exact row max of this tile: it is the running max of the previous tiles (first tile:
                // the max of its first 32 columns).  Softmax is shift invariant, bf16/fp32 precision is relative, and the
                // exponent is clamped far below overflow, so the result is the same; the true running max `m_seen` is
                // tracked alongside and `m` catches up lazily (rescaling O and l) once it lags by more than 2^8.
                const bool fine = threadIdx.x == 0 && n_it == 2 && j == 1;
                SVR2_FINE(44);
                uint32_t a[32], b[32];
                tmem_ld32(tS, a);
                if (j > 0) {
                    const bool need = (m_seen - m) * kScaleLog2 > kRescaleThreshold;
                    if (__any_sync(0xffffffffu, need)) {                  // warp-uniform: tcgen05.ld/st are warp collectives
                        const float m_use = need ? m_seen : m;            // SFULL(j) implies O += P(j-1) V(j-1) has completed
                        const float alpha = ex2((m - m_use) * kScaleLog2);
                        l *= alpha;
                        m = m_use;
                        tc_wait_ld();
#pragma unroll 1
                        for (int c4 = 0; c4 < 4; ++c4) {
                            tmem_ld32(tO + c4 * 32, b);
                            tc_wait_ld();
#pragma unroll
                            for (int c = 0; c < 32; ++c) b[c] = __float_as_uint(__uint_as_float(b[c]) * alpha);
                            tmem_st32(tO + c4 * 32, b);
                        }
                    }
                }
                uint32_t pk[16];
                float rs, cm;
                tc_wait_ld();
                SVR2_FINE(45);
                if (n_ch > 1) tmem_ld32(tS + 32, b);
                cm = chunk_max(a, valid);
                if (j == 0) m = cm;
                const float mb = m * kScaleLog2;
                rs = chunk_softmax(a, pk, mb, cm, valid);
                tmem_st16(tS, pk);
                SVR2_FINE(46);
                if (n_ch > 1) {
                    tc_wait_ld();
                    if (n_ch > 2) tmem_ld32(tS + 64, a);
                    const float c1 = chunk_max(b, valid - 32);
                    cm = fmaxf(cm, c1);
                    rs += chunk_softmax(b, pk, mb, c1, valid - 32);
                    tmem_st16(tS + 16, pk);
                    SVR2_FINE(47);
                }
                if (n_ch > 2) {
                    tc_wait_ld();
                    if (n_ch > 3) tmem_ld32(tS + 96, b);
                    const float c2 = chunk_max(a, valid - 64);
                    cm = fmaxf(cm, c2);
                    rs += chunk_softmax(a, pk, mb, c2, valid - 64);
                    tmem_st16(tS + 32, pk);
                    SVR2_FINE(48);
                }
                if (n_ch > 3) {
                    tc_wait_ld();
                    const float c3 = chunk_max(b, valid - 96);
                    cm = fmaxf(cm, c3);
                    rs += chunk_softmax(b, pk, mb, c3, valid - 96);
                    tmem_st16(tS + 48, pk);
                    SVR2_FINE(49);
                }
                m_seen = fmaxf(m_seen, cm);
                l += rs;
                tc_wait_st();
                SVR2_FINE(50);
                tc_fence_before();
                mbar_arrive(bar(BAR_PREADY + i));
                SVR2_FINE(51);
                if (threadIdx.x == 0) {
                    if (j == 0) trace_item(p, n_it, 2);
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
        uint32_t c[2] = {0, 0};
        int n_it = 0;
        for (int item = blockIdx.x; item < p.n_items; item += gridDim.x, ++n_it) {
            const Item it = decode_item(p, item);
            for (int i = 0; i < it.nq; ++i) {
                const uint32_t tO = tmem + lane_off + 256 + i * 128;
                mbar_wait(bar(BAR_LREADY + i), c[i] & 1);
                mbar_wait(bar(BAR_OFINAL + i), c[i] & 1);
                tc_fence_after();
                if (threadIdx.x == 256 && i == 0) trace_item(p, n_it, 5);
                const bool fine = threadIdx.x == 256 && i == 0 && n_it == 2;
                SVR2_FINE(56);
                const float inv = 1.0f / l_buf[((c[i] & 1) * 2 + i) * 128 + row];
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
                SVR2_FINE(57);
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
                SVR2_FINE(58);
                if (threadIdx.x == 256) trace_item(p, n_it, 6 + i);
                ++c[i];
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
