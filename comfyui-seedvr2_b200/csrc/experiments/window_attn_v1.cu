// EXPERIMENT (generation 1; built only with `make EXPERIMENTS=1`, svr2_window_attn_fwd_impl code 1)
// svr2_window_attn_fwd — variable-length NaSwin window attention for sm_100a: TMA-staged Q/K/V tiles, tcgen05 MMA with
// TMEM accumulators, online softmax, window reverse + text pooling fused into the epilogue.
//
// Replaces FlashAttentionVarlen / pytorch_varlen_attention (reference src/models/dit_3b/attention.py:24-61,138-157: a
// Python loop of one SDPA call per window), unconcat_win (na.py:396-417) and window_reverse (na.py:638).
//
// Work item = (window, pair of 128-row query tiles, head).  One CTA per item, one CTA per SM (all 512 TMEM columns):
//   warps 0-3  softmax group 0 (query tile 0): one thread per query row, tcgen05.ld S -> exp2 -> tcgen05.st P (bf16)
//   warps 4-7  softmax group 1 (query tile 1)
//   warp  8    TMA producer: Q tiles once, then K(j), V(j) through a 4-stage ring of 32-KB tiles
//   warp  9    tcgen05.mma issuer (one thread): S_i = Q_i K^T (SS), O_i += P_i V (A = P from TMEM, B = V MN-major)
// The two query tiles ping-pong: while group 0 runs softmax on S_0(j) the tensor core works on tile 1, and vice versa
// (the FlashAttention-4 schedule).  P aliases the first 64 columns of its S buffer; tensor-core ops execute in issue
// order, so S_i(j+1) = Q_i K(j+1)^T issued after O_i += P_i(j) V(j) cannot clobber P_i(j) early.
//
// TMEM map (columns): S0 [0,128)  S1 [128,256)  O0 [256,384)  O1 [384,512).
// SMEM map: Q0 | Q1 (32 KB each, reused as the bf16 output staging tile) | ring[4] x 32 KB | barriers.
// Every 128x128 bf16 tile is two TMA boxes of 128 rows x 64 cols with the 128-byte swizzle (canonical UMMA layout).
#include <cuda.h>

#include "../common.cuh"
#include "../attn_launch.cuh"
#include "../sm100_ptx.cuh"

namespace svr2 {


namespace tc_attn {

using namespace ptx;

constexpr int D = SVR2_HEAD_DIM;          // 128
constexpr int BM = 128, BN = 128;
constexpr int TILE_BYTES = BM * D * 2;    // 32768
constexpr int BOX_BYTES = TILE_BYTES / 2; // 128 rows x 64 cols bf16
constexpr int STAGES = 4;
constexpr int NTHREADS = 320;
constexpr int SMEM_BYTES = 1024 /*align slack*/ + (2 + STAGES) * TILE_BYTES + 256 /*barriers*/;
constexpr uint32_t TMEM_COLS = 512;
constexpr float kScaleLog2 = 0.08838834764831845f * 1.4426950408889634f;   // 1/sqrt(128) * log2(e)
constexpr float kRescaleThreshold = 8.0f;   // lazy rescale: tolerate P up to 2^8 before touching O

// barrier slots (8 bytes each) inside the barrier block
enum { BAR_QFULL = 0, BAR_SFULL = 2, BAR_PREADY = 4, BAR_OFINAL = 6, BAR_KVFULL = 8, BAR_KVEMPTY = 8 + STAGES,
       BAR_COUNT = 8 + 2 * STAGES };

struct Params {
    const int32_t* tab;
    const int32_t* work;
    int H, W, txt_len, heads;
    __nv_bfloat16* out_vid;
    __nv_bfloat16* txt_part;
    long long* trace;      // debug timeline (svr2_debug_attn_trace): kTraceSlots clock64 stamps per CTA, or nullptr
    int trace_ctas;
};

constexpr int kTraceSlots = 64;
// slot map: 0 entry, 1 setup done, 2 Q+K0 landed (MMA warp), 3 O final seen (group 0), 4 epilogue stores issued (group 0),
// 5 CTA end, 6 smid, 7 n_kv, 8+j S(j) ready seen by group 0, 24+j P(j) published by group 0, 40+j group-1 S(j) ready
__device__ __forceinline__ void trace_stamp(const Params& p, int slot) {
    if (p.trace != nullptr) {
        const int cta = blockIdx.y * gridDim.x + blockIdx.x;
        if (cta < p.trace_ctas && slot < kTraceSlots) p.trace[(long long)cta * kTraceSlots + slot] = clock64();
    }
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

__global__ void __launch_bounds__(NTHREADS, 1)
window_attn_tc_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                      const __grid_constant__ CUtensorMap tm_v, const Params p) {
    extern __shared__ unsigned char smem_raw[];
    const uint32_t raw_u32 = smem_u32(smem_raw);
    const uint32_t base = (raw_u32 + 1023u) & ~1023u;            // SWIZZLE_128B tiles need 1024-B alignment
    unsigned char* gbase = smem_raw + (base - raw_u32);
    const uint32_t sRing = base + 2 * TILE_BYTES;
    const uint32_t sBar = sRing + STAGES * TILE_BYTES;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(gbase + (2 + STAGES) * TILE_BYTES + BAR_COUNT * 8);
    auto bar = [&](int slot) { return sBar + 8u * (uint32_t)slot; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wi = __ldg(p.work + 2 * blockIdx.x), qpair = __ldg(p.work + 2 * blockIdx.x + 1), head = blockIdx.y;
    const Win w = load_win(p.tab, wi);
    const int seq = w.n_vid + p.txt_len;
    const int q0 = qpair * 2 * BM;                                // first query row (inside the sequence) of this item
    const int nq = (seq - q0 > BM) ? 2 : 1;
    const int n_kv = (seq + BN - 1) / BN;

    if (threadIdx.x == 0) {
        trace_stamp(p, 0);
        for (int i = 0; i < 2; ++i) {
            mbar_init(bar(BAR_QFULL + i), 1);
            mbar_init(bar(BAR_SFULL + i), 1);
            mbar_init(bar(BAR_PREADY + i), 128);
            mbar_init(bar(BAR_OFINAL + i), 1);
        }
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_KVFULL + s), 1);
            mbar_init(bar(BAR_KVEMPTY + s), 1);
        }
        fence_mbar_init();
    }
    if (warp == 9) tmem_alloc(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
    if (warp == 8 && lane == 0) {
        prefetch_tmap(&tm_q);
        prefetch_tmap(&tm_k);
        prefetch_tmap(&tm_v);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    if (threadIdx.x == 0) {
        trace_stamp(p, 1);
        if (p.trace != nullptr) {
            const int cta = blockIdx.y * gridDim.x + blockIdx.x;
            if (cta < p.trace_ctas) {
                unsigned smid;
                asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
                p.trace[(long long)cta * kTraceSlots + 6] = smid;
                p.trace[(long long)cta * kTraceSlots + 7] = n_kv;
            }
        }
    }

    if (warp == 8) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            for (int i = 0; i < nq; ++i) {
                const int row = w.row_start + q0 + i * BM;
                mbar_expect_tx(bar(BAR_QFULL + i), TILE_BYTES);
                tma_load_3d(base + i * TILE_BYTES, &tm_q, bar(BAR_QFULL + i), 0, head, row);
                tma_load_3d(base + i * TILE_BYTES + BOX_BYTES, &tm_q, bar(BAR_QFULL + i), 64, head, row);
            }
            for (int n = 0; n < 2 * n_kv; ++n) {                  // ring order K0 V0 K1 V1 ...
                const int st = n % STAGES, use = n / STAGES;
                const CUtensorMap* map = (n & 1) ? &tm_v : &tm_k;
                const int row = w.row_start + (n >> 1) * BN;
                mbar_wait(bar(BAR_KVEMPTY + st), (use & 1) ^ 1);
                mbar_expect_tx(bar(BAR_KVFULL + st), TILE_BYTES);
                tma_load_3d(sRing + st * TILE_BYTES, map, bar(BAR_KVFULL + st), 0, head, row);
                tma_load_3d(sRing + st * TILE_BYTES + BOX_BYTES, map, bar(BAR_KVFULL + st), 64, head, row);
            }
        }
    } else if (warp == 9) {
        // ------------------------------------------------------------------ tcgen05.mma issuer (single thread)
        if (lane == 0) {
            constexpr uint32_t idesc_qk = idesc_bf16(BM, BN, 0);
            constexpr uint32_t idesc_pv = idesc_bf16(BM, D, 1);
            auto issue_qk = [&](int i, uint32_t kbase) {          // S_i = Q_i K^T : 8 x (128x128x16), both K-major
#pragma unroll
                for (int kk = 0; kk < 8; ++kk) {
                    const uint32_t off = (uint32_t)(kk >> 2) * BOX_BYTES + (uint32_t)(kk & 3) * 32u;
                    mma_ss(tmem + i * 128, smem_desc_sw128(base + i * TILE_BYTES + off, 16, 1024),
                           smem_desc_sw128(kbase + off, 16, 1024), idesc_qk, kk > 0);
                }
            };
            auto issue_pv = [&](int i, uint32_t vbase, bool first) {   // O_i += P_i V : A = P (TMEM), B = V MN-major
#pragma unroll
                for (int kk = 0; kk < 8; ++kk)
                    mma_ts(tmem + 256 + i * 128, tmem + i * 128 + kk * 8,
                           smem_desc_sw128(vbase + kk * 2048, BOX_BYTES, 1024), idesc_pv, !(first && kk == 0));
            };
            for (int i = 0; i < nq; ++i) mbar_wait(bar(BAR_QFULL + i), 0);
            mbar_wait(bar(BAR_KVFULL + 0), 0);
            tc_fence_after();
            trace_stamp(p, 2);
            for (int i = 0; i < nq; ++i) {
                issue_qk(i, sRing);
                tc_commit(bar(BAR_SFULL + i));
            }
            tc_commit(bar(BAR_KVEMPTY + 0));
            for (int j = 0; j < n_kv; ++j) {
                const int nV = 2 * j + 1, stV = nV % STAGES;
                const bool has_next = j + 1 < n_kv;
                const int nK = 2 * j + 2, stK = nK % STAGES;
                mbar_wait(bar(BAR_KVFULL + stV), (nV / STAGES) & 1);
                if (has_next) mbar_wait(bar(BAR_KVFULL + stK), (nK / STAGES) & 1);
                for (int i = 0; i < nq; ++i) {
                    mbar_wait(bar(BAR_PREADY + i), j & 1);
                    tc_fence_after();
                    issue_pv(i, sRing + stV * TILE_BYTES, j == 0);
                    if (has_next) {
                        issue_qk(i, sRing + stK * TILE_BYTES);
                        tc_commit(bar(BAR_SFULL + i));
                    } else {
                        tc_commit(bar(BAR_OFINAL + i));
                    }
                }
                tc_commit(bar(BAR_KVEMPTY + stV));
                if (has_next) tc_commit(bar(BAR_KVEMPTY + stK));
            }
        }
    } else if ((warp >> 2) < nq) {
        // ------------------------------------------------------------------ softmax + epilogue group i
        const int i = warp >> 2, quad = warp & 3, row = quad * 32 + lane;
        const uint32_t lane_off = (uint32_t)(quad * 32) << 16;
        const uint32_t tS = tmem + lane_off + i * 128, tO = tmem + lane_off + 256 + i * 128;
        float m = -INFINITY, l = 0.f;
        for (int j = 0; j < n_kv; ++j) {
            mbar_wait(bar(BAR_SFULL + i), j & 1);
            tc_fence_after();
            if ((threadIdx.x & 127) == 0 && j < 16) trace_stamp(p, (i == 0 ? 8 : 40) + j);
            uint32_t sv[128];
            tmem_ld32(tS, sv);
            tmem_ld32(tS + 32, sv + 32);
            tmem_ld32(tS + 64, sv + 64);
            tmem_ld32(tS + 96, sv + 96);
            tc_wait_ld();
            const int valid = seq - j * BN;                        // columns >= valid are beyond the sequence
            float mx = -INFINITY;
            if (valid < BN) {
#pragma unroll
                for (int c = 0; c < 128; ++c) {
                    const float v = (c < valid) ? __uint_as_float(sv[c]) : -INFINITY;
                    sv[c] = __float_as_uint(v);
                    mx = fmaxf(mx, v);
                }
            } else {
#pragma unroll
                for (int c = 0; c < 128; ++c) mx = fmaxf(mx, __uint_as_float(sv[c]));
            }
            const float m_new = fmaxf(m, mx);
            if (j == 0) {
                m = m_new;
            } else {
                const bool need = (m_new - m) * kScaleLog2 > kRescaleThreshold;
                if (__any_sync(0xffffffffu, need)) {               // warp-uniform: tcgen05.ld/st are warp collectives
                    const float m_use = need ? m_new : m;
                    const float alpha = ex2((m - m_use) * kScaleLog2);
                    l *= alpha;
                    m = m_use;
#pragma unroll
                    for (int c4 = 0; c4 < 4; ++c4) {
                        uint32_t ov[32];
                        tmem_ld32(tO + c4 * 32, ov);
                        tc_wait_ld();
#pragma unroll
                        for (int c = 0; c < 32; ++c) ov[c] = __float_as_uint(__uint_as_float(ov[c]) * alpha);
                        tmem_st32(tO + c4 * 32, ov);
                    }
                }
            }
            const float mb = m * kScaleLog2;
            uint32_t pk[64];
            float rs = 0.f;
#pragma unroll
            for (int c = 0; c < 64; ++c) {
                const float p0 = ex2(fmaf(__uint_as_float(sv[2 * c]), kScaleLog2, -mb));
                const float p1 = ex2(fmaf(__uint_as_float(sv[2 * c + 1]), kScaleLog2, -mb));
                rs += p0 + p1;
                pk[c] = pack2(p0, p1);
            }
            l += rs;
            tmem_st32(tS, pk);
            tmem_st32(tS + 32, pk + 32);
            tc_wait_st();
            tc_fence_before();
            mbar_arrive(bar(BAR_PREADY + i));
            if (threadIdx.x == 0 && j < 16) trace_stamp(p, 24 + j);
        }
        // ---- epilogue: O_i / l -> bf16 -> swizzled staging tile (the Q_i buffer) -> coalesced 256-B row stores
        mbar_wait(bar(BAR_OFINAL + i), 0);
        tc_fence_after();
        if (threadIdx.x == 0) trace_stamp(p, 3);
        const float inv = 1.0f / l;
        unsigned char* stage = gbase + i * TILE_BYTES;
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4) {
            uint32_t ov[32];
            tmem_ld32(tO + c4 * 32, ov);
            tc_wait_ld();
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
        // per-window text buffer that svr2's text pooling averages
        const int jrow = q0 + i * BM + row;
        const int64_t row_stride = (int64_t)p.heads * D;
        unsigned long long dst = 0ull;
        if (jrow < w.n_vid) {
            int lt, lh, lw;
            dst = (unsigned long long)(p.out_vid + win_token(w, jrow, p.H, p.W, &lt, &lh, &lw) * row_stride + head * D);
        } else if (jrow < seq) {
            dst = (unsigned long long)(p.txt_part + ((int64_t)wi * p.txt_len + (jrow - w.n_vid)) * row_stride + head * D);
        }
        __syncwarp();
        const int ch = lane & 15;
#pragma unroll 4
        for (int it = 0; it < 16; ++it) {
            const int rl = 2 * it + (lane >> 4);                   // row (inside this warp's 32) served by my half-warp
            const unsigned long long d = __shfl_sync(0xffffffffu, dst, rl);
            const int r = quad * 32 + rl;
            const uint4 u = *reinterpret_cast<const uint4*>(stage + r * 256 + (((ch & 8) | ((ch ^ r) & 7)) << 4));
            if (d != 0ull) *reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(d) + ch * 8) = u;
        }
        if (threadIdx.x == 0) trace_stamp(p, 4);
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) trace_stamp(p, 5);
    if (warp == 9) {
        __syncwarp();
        tmem_dealloc(tmem, TMEM_COLS);
    }
}

}  // namespace tc_attn
}  // namespace svr2

namespace svr2 {

int launch_window_attn_tc(const void* q, const void* k, const void* v, const int32_t* tab, const int32_t* work,
                          int n_work, int H, int W, int txt_len, int heads, int64_t P, void* out_vid, void* txt_part,
                          cudaStream_t st) {
    CUtensorMap tq, tk, tv;
    int rc;
    if ((rc = make_tmap(&tq, q, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tk, k, P, heads)) != SVR2_OK) return rc;
    if ((rc = make_tmap(&tv, v, P, heads)) != SVR2_OK) return rc;
    static PerDeviceFlag smem_ok;
    if ((rc = ensure_dyn_smem(tc_attn::window_attn_tc_kernel, tc_attn::SMEM_BYTES, smem_ok, "svr2_window_attn_fwd")) != SVR2_OK) return rc;
    tc_attn::Params p;
    p.tab = tab; p.work = work; p.H = H; p.W = W; p.txt_len = txt_len; p.heads = heads;
    p.out_vid = (__nv_bfloat16*)out_vid; p.txt_part = (__nv_bfloat16*)txt_part;
    p.trace = g_trace; p.trace_ctas = g_trace_ctas;
    dim3 grid(n_work, heads);
    tc_attn::window_attn_tc_kernel<<<grid, tc_attn::NTHREADS, tc_attn::SMEM_BYTES, st>>>(tq, tk, tv, p);
    SVR2_CHECK_LAUNCH("svr2_window_attn_fwd(tcgen05)");
    return SVR2_OK;
}


}  // namespace svr2
