// svr2_linear_*_fwd — the block's linear layers as ONE persistent tcgen05 GEMM kernel with the row-wise work that follows
// each of them fused into its epilogue.
//
//   y = x W^T (+ bias)                                   proj_qkv                      mmattn.py:56-59,173
//   y = silu(x Wg^T) * (x Wu^T)                          SwiGLU in-projections         mlp.py:46-62   (two B operands, one tile)
//   y = gelu_tanh(x W^T + bias)                          7B MLP proj_in                mlp.py:28-43
//   y = (x W^T + bias) * (gateA + gateB) + resid         proj_out / MLP proj_out       mmattn.py:269, mlp.py:43,62 + AdaSingle
//                                                        "out" (modulation.py:111-116) + residual (mmsr_block.py:114,126)
//   q,k,v packed = scatter(rope(rmsnorm(x Wqkv^T)))      proj_qkv + norm_q / norm_k + RoPE + window partition + text split
//                                                        mmattn.py:173,199-248 (7B: mmsr_block.py:106-128) — the epilogue of
//                                                        svr2_linear_qkv_pack_fwd; see "QKV epilogue" below
// x [M, K] bf16 row-major, W [N, K] bf16 row-major (nn.Linear layout) — both operands are K-major, the canonical UMMA case.
//
// Shape of the kernel (sm_100a):
//   * CTA pair (cluster of 2, one TPC): tcgen05.mma.cta_group::2, M = 256 (128 rows per CTA), N = 256, K = 16 per instruction.
//     Each CTA stages its own 128 A rows and HALF of the B tile (128 of the 256 weight rows): 32 KB per 64-wide K step per CTA
//     for 256 tensor-pipe cycles.
//   * warp 0: TMA producer (5-stage ring, 128-byte swizzle boxes of 64 K x 128 rows; both CTAs' bytes complete on the leader's
//     "full" barrier); warp 1 of the leader CTA: the pair's MMA issuer (one elected lane), commits multicast to both CTAs;
//     warps 4-7: epilogue, one thread per accumulator row (TMEM lane).
//   * TMEM: two 256-column fp32 accumulators; the epilogue of tile i drains buffer i & 1 while the tensor pipe fills the other.
//   * persistent: grid = one pair per TPC; tiles are walked in groups of 8 row tiles x all column tiles so that the ~74 tiles in
//     flight share A rows and W rows in L2.
//   * epilogue: tcgen05.ld (32 lanes x 32 columns) -> bias / activation / gate / residual in the reference's bf16 rounding
//     order -> packed bf16 -> 128-byte-swizzled 4-KB slab per warp -> TMA store (rows past M are clipped by the tensor map).
//
// QKV epilogue (EPI_QKV*): a 256-column tile is two whole heads of q, k or v, and an epilogue thread owns one token row, so the
// per-head RMSNorm is a thread-local reduction over the 128 accumulator columns it has just read from TMEM, and the interleaved-
// pair rotation needs no exchange either.  cos / sin come from a small per-plan table in shared memory (svr2_rope_plan: one row
// per distinct position value, the three axes of a token index it through the token map); the token map also holds the row's
// destination in the packed window layout, so the window partition is the store address.  Each thread stages its 256-byte head
// row in a private shared-memory line and sends it with ONE bulk copy (cp.async.bulk, 256 B contiguous in HBM) — no per-lane
// scattered stores.  Eight epilogue warps (two per TMEM lane quarter, one head each).  The qkv tensor, the qkv_pack pass that
// re-read it and the packed q,k,v it re-wrote (0 algorithmic bytes for partition / interleave) are gone.
//
// Rounding points: the linear's output is a bf16 tensor in the reference, so the accumulator (+ bias, in fp32
// like cuBLASLt's epilogue) is rounded to bf16 FIRST; every following elementwise op rounds again, exactly as
// svr2_swiglu_fwd / svr2_gate_residual_fwd do on the materialised tensors.
#include <cuda.h>

#include "common.cuh"
#include "sm100_pair.cuh"
#include "sm100_ptx.cuh"
#include "tmap.cuh"

namespace svr2 {
namespace gemm {

using namespace ptx;
using namespace pairx;

constexpr int BM = 128;                  // accumulator rows per CTA (pair: 256)
constexpr int BN = 256;                  // accumulator columns = weight rows per tile (this CTA stages BN / 2 of them)
constexpr int BK = 64;                   // one 128-byte swizzle atom of bf16
#ifndef SVR2_GEMM_STAGES
#define SVR2_GEMM_STAGES 5
#endif
constexpr int A_BYTES = BM * BK * 2;             // 16384
constexpr int B_BYTES = (BN / 2) * BK * 2;       // 16384
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int SLAB_BYTES = 32 * 128;             // 32 rows x 64 bf16 columns, one TMA store box
constexpr int kMaxGateN = 4096;
constexpr uint32_t TMEM_COLS = 512;
constexpr int GROUP_M = 8;                       // row tiles per rasterisation group
constexpr int QKV_ROW_STRIDE = 272;              // staged head row: 256 B + 16 B pad (quarter-warp STS.128 hit 32 distinct banks)
constexpr int kMaxRopeEntries = 2048;            // float2 (cos, sin) entries of the rope table held in shared memory (16 KB)

enum Epilogue { EPI_BIAS = 0, EPI_GELU = 1, EPI_GATE_RESID = 2, EPI_SWIGLU = 3,
                EPI_QKV0 = 4 /*no rope*/, EPI_QKV21 = 5 /*3B: 21 frequencies per axis*/, EPI_QKV10 = 6 /*7B: 10*/ };

// Per-epilogue shape of the CTA: ring depth, epilogue warps, shared-memory map, barrier slots.
template <int EPI>
struct Cfg {
    static constexpr bool QKV = EPI >= EPI_QKV0;
    static constexpr int NF = EPI == EPI_QKV21 ? 21 : (EPI == EPI_QKV10 ? 10 : 0);     // rotary frequencies per axis
    static constexpr int STAGES = QKV ? 4 : SVR2_GEMM_STAGES;
    static constexpr int EPI_WARPS = QKV ? 8 : 4;
    static constexpr int NTHREADS = 128 + 32 * EPI_WARPS;
    static constexpr int OFF_SLAB = STAGES * STAGE_BYTES;      // [4 warps][2] swizzled slabs | QKV: [8 warps][32] staged head rows
    static constexpr int SLAB_TOTAL = QKV ? EPI_WARPS * 32 * QKV_ROW_STRIDE : 4 * 2 * SLAB_BYTES;
    static constexpr int OFF_AUX = OFF_SLAB + SLAB_TOTAL;      // bf16 gate[kMaxGateN] | QKV: rope table, then 4 x 128 norm weights
    static constexpr int AUX_BYTES = QKV ? kMaxRopeEntries * 8 + 4 * 128 * 2 : kMaxGateN * 2;
    static constexpr int OFF_BAR = OFF_AUX + AUX_BYTES;
    static constexpr int SMEM_BYTES = OFF_BAR + 256;
    static constexpr int BAR_FULL = 0;                  // leader's copy counts both CTAs' bytes
    static constexpr int BAR_EMPTY = STAGES;            // local, multicast commit
    static constexpr int BAR_TFULL = 2 * STAGES;        // local, multicast commit
    static constexpr int BAR_TEMPTY = 2 * STAGES + 2;   // leader's copy, one arrival per epilogue warp of the pair
    static constexpr int BAR_COUNT = 2 * STAGES + 4;
    static_assert(BAR_COUNT * 8 + 8 <= 256, "barrier block too small");
    static_assert(SMEM_BYTES <= 232448, "shared memory");
    static_assert(OFF_SLAB % 1024 == 0, "swizzled slabs need 1024-byte alignment");
    static_assert(OFF_AUX % 16 == 0 && OFF_BAR % 8 == 0, "alignment");
};

struct Params {
    int64_t M;
    int N;                 // weight rows of ONE B operand (= output columns)
    int K;
    int m_tiles, n_tiles;  // pair tiles: 256 rows x (256 | 128 for SwiGLU) output columns
    const __nv_bfloat16* bias;      // [N] or nullptr
    const __nv_bfloat16* emb;       // EPI_GATE_RESID: [6 N] (d, layer, {shift, scale, gate})
    const __nv_bfloat16* gateB;     // EPI_GATE_RESID: [N]
    const __nv_bfloat16* resid;     // EPI_GATE_RESID: [M, N]
    int layer;
    // EPI_QKV*: N = 3 * heads * 128
    const int4* tok_map;            // [M] (packed destination row | -(1 + text row), rope-table row of axis t, h, w | -1)
    const float2* rope_cs;          // [rope_entries] (cos, sin), row-major [table row][frequency]
    int rope_entries;
    const __nv_bfloat16 *wq_vid, *wk_vid, *wq_txt, *wk_txt;     // norm_q / norm_k weights [128]
    float eps;
    int heads;
    __nv_bfloat16 *q_pack, *k_pack, *v_pack;                     // [P, heads, 128]
    __nv_bfloat16* txt_qkv;                                      // [Lt, 3, heads, 128]: text rows, replicated per window afterwards
};

__device__ __forceinline__ void decode_tile(const Params& p, int t, int& mt, int& nt) {
    const int per_group = GROUP_M * p.n_tiles;
    const int g = t / per_group;
    const int first = g * GROUP_M;
    const int gsz = min(GROUP_M, p.m_tiles - first);
    const int r = t - g * per_group;
    mt = first + r % gsz;
    nt = r / gsz;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ __nv_bfloat162 as_bf162(const uint32_t& u) { return *reinterpret_cast<const __nv_bfloat162*>(&u); }
__device__ __forceinline__ uint32_t as_u32(const __nv_bfloat162& v) { return *reinterpret_cast<const uint32_t*>(&v); }

// GELU, approximate="tanh" (ATen: 0.5 x (1 + tanh(beta (x + kappa x^3))), fp32 opmath).  0.5 (1 + tanh u) == sigmoid(2u), so
// gelu(x) = x / (1 + exp(-2u)) — two MUFU ops instead of a tanhf expansion, relative error ~1e-7, no cancellation in the tail.
__device__ __forceinline__ float gelu_tanh_f(float x) {
    constexpr float kBeta = 0.7978845608028654f;             // sqrt(2 / pi)
    constexpr float kKappa = 0.044715f;
    const float u = kBeta * (x + kKappa * x * x * x);
    return __fdividef(x, 1.0f + __expf(-2.0f * u));
}
__device__ __forceinline__ float silu_f(float g) { return __fdividef(g, 1.0f + __expf(-g)); }      // as svr2_swiglu_fwd

// 256-byte bulk copy shared -> global of one staged head row (completion tracked by the issuing thread's bulk group)
__device__ __forceinline__ void bulk_store_256(void* gdst, uint32_t ssrc) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 256;" ::"l"(gdst), "r"(ssrc) : "memory");
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(Cfg<EPI>::NTHREADS, 1)
linear_tc_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b0,
                 const __grid_constant__ CUtensorMap tm_b1, const __grid_constant__ CUtensorMap tm_c, const Params p) {
    using C = Cfg<EPI>;
    constexpr int STAGES = C::STAGES, NTHREADS = C::NTHREADS, OFF_SLAB = C::OFF_SLAB, OFF_GATE = C::OFF_AUX, OFF_BAR = C::OFF_BAR;
    constexpr int BAR_FULL = C::BAR_FULL, BAR_EMPTY = C::BAR_EMPTY, BAR_TFULL = C::BAR_TFULL, BAR_TEMPTY = C::BAR_TEMPTY,
                  BAR_COUNT = C::BAR_COUNT;
    extern __shared__ __align__(1024) unsigned char smem[];
    const uint32_t base = smem_u32(smem);
    const uint32_t sBar = base + OFF_BAR;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + OFF_BAR + BAR_COUNT * 8);
    auto bar = [&](int slot) { return sBar + 8u * (uint32_t)slot; };

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();
    const bool leader = rank == 0;
    const int first = (int)(blockIdx.x >> 1), stride = (int)(gridDim.x >> 1);
    const int n_tiles_total = p.m_tiles * p.n_tiles;
    const int kb_count = (p.K + BK - 1) / BK;
    constexpr int OUT_N = EPI == EPI_SWIGLU ? BN / 2 : BN;        // output columns of a tile

    if (threadIdx.x == 0) {
        if (base & 1023u) __trap();
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(bar(BAR_FULL + s), 1);
            mbar_init(bar(BAR_EMPTY + s), 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(bar(BAR_TFULL + b), 1);
            mbar_init(bar(BAR_TEMPTY + b), 2 * C::EPI_WARPS);      // one arrival per epilogue warp of the pair
        }
        fence_mbar_init();
    }
    if (warp == 2) tmem_alloc_pair(smem_u32(const_cast<uint32_t*>(tmem_slot)), TMEM_COLS);
    if (warp == 0 && lane == 0) {
        prefetch_tmap(&tm_a);
        prefetch_tmap(&tm_b0);
        prefetch_tmap(&tm_b1);
        prefetch_tmap(&tm_c);
    }
    if constexpr (EPI == EPI_GATE_RESID) {
        // gateA + gateB is a bf16 tensor in the reference (modulation.py:111-116); formed once per CTA
        __nv_bfloat16* gate_s = reinterpret_cast<__nv_bfloat16*>(smem + OFF_GATE);
        for (int d = threadIdx.x; d < p.N; d += NTHREADS)
            gate_s[d] = __float2bfloat16_rn(__bfloat162float(p.emb[d * 6 + p.layer * 3 + 2]) + __bfloat162float(p.gateB[d]));
    }
    if constexpr (C::QKV) {
        // the plan's rope table and the four norm weight vectors, once per CTA
        float2* cs_s = reinterpret_cast<float2*>(smem + OFF_GATE);
        for (int i = threadIdx.x; i < p.rope_entries; i += NTHREADS) cs_s[i] = __ldg(p.rope_cs + i);
        uint32_t* w_s = reinterpret_cast<uint32_t*>(smem + OFF_GATE + kMaxRopeEntries * 8);
        if (threadIdx.x < 256) {
            const int which = threadIdx.x >> 6, i = threadIdx.x & 63;
            const __nv_bfloat16* src = which == 0 ? p.wq_vid : (which == 1 ? p.wk_vid : (which == 2 ? p.wq_txt : p.wk_txt));
            w_s[which * 64 + i] = src != nullptr ? __ldg(reinterpret_cast<const uint32_t*>(src) + i) : 0x3F803F80u;
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();              // the peer's barriers are initialised and its TMEM allocated before anything crosses over
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer: this CTA's A rows and B half
        if (lane == 0) {
            // SwiGLU: CTA 0 stages 128 rows of W_gate, CTA 1 the same 128 rows of W_up -> accumulator columns [0,128) | [128,256)
            const CUtensorMap* tb = (EPI == EPI_SWIGLU && rank == 1) ? &tm_b1 : &tm_b0;
            const int b_row_off = EPI == EPI_SWIGLU ? 0 : (int)rank * (BN / 2);
            uint32_t it = 0;
            for (int t = first; t < n_tiles_total; t += stride) {
                int mt, nt;
                decode_tile(p, t, mt, nt);
                const int a_row = mt * (2 * BM) + (int)rank * BM;
                const int b_row = nt * OUT_N + b_row_off;
                for (int kb = 0; kb < kb_count; ++kb, ++it) {
                    const uint32_t s = it % STAGES, use = it / STAGES;
                    mbar_wait(bar(BAR_EMPTY + s), (use & 1) ^ 1);
                    if (leader) mbar_expect_tx(bar(BAR_FULL + s), 2 * STAGE_BYTES);
                    const uint32_t dst = base + s * STAGE_BYTES;
                    tma_load_2d_pair(dst, &tm_a, bar(BAR_FULL + s), kb * BK, a_row);
                    tma_load_2d_pair(dst + A_BYTES, tb, bar(BAR_FULL + s), kb * BK, b_row);
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ tcgen05.mma issuer of the pair (leader CTA)
        if (leader) {
            constexpr uint32_t idesc = idesc_bf16(2 * BM, BN, 0);
            constexpr uint32_t hi_sw = smem_desc_hi_sw128(1024);
            uint32_t it = 0, tc = 0;
            for (int t = first; t < n_tiles_total; t += stride, ++tc) {
                const uint32_t b = tc & 1;
                mbar_wait(bar(BAR_TEMPTY + b), ((tc >> 1) & 1) ^ 1);       // the epilogues of both CTAs drained buffer b
                tc_fence_after();
                const uint32_t d_tmem = tmem + b * BN;
                for (int kb = 0; kb < kb_count; ++kb, ++it) {
                    const uint32_t s = it % STAGES, use = it / STAGES;
                    mbar_wait(bar(BAR_FULL + s), use & 1);
                    tc_fence_after();
                    if (elect_one()) {
                        const uint32_t alo = smem_desc_lo(base + s * STAGE_BYTES, 16);
                        const uint32_t blo = smem_desc_lo(base + s * STAGE_BYTES + A_BYTES, 16);
#pragma unroll
                        for (int kk = 0; kk < BK / 16; ++kk)
                            mma2_ss(d_tmem, make_desc(alo + kk * 2, hi_sw), make_desc(blo + kk * 2, hi_sw), idesc,
                                    (uint32_t)((kb | kk) != 0));
                        tc_commit_pair(bar(BAR_EMPTY + s));
                        if (kb == kb_count - 1) tc_commit_pair(bar(BAR_TFULL + b));
                    }
                    __syncwarp();
                }
            }
        }
    } else if (C::QKV && warp >= 4) {
        // ------------------------------------------------------------------ QKV epilogue: one thread per token row, one head per warp
        if constexpr (C::QKV) {
            constexpr int NF = C::NF, NROT = 3 * NF;
            const int q = warp & 3, hsel = (warp - 4) >> 2;
            const uint32_t lane_off = (uint32_t)(q * 32) << 16;
            unsigned char* my_row = smem + OFF_SLAB + ((warp - 4) * 32 + lane) * QKV_ROW_STRIDE;
            const uint32_t my_row_s = smem_u32(my_row);
            const uint32_t tempty_g0 = mapa(bar(BAR_TEMPTY + 0), 0), tempty_g1 = mapa(bar(BAR_TEMPTY + 1), 0);
            const float2* cs_s = reinterpret_cast<const float2*>(smem + OFF_GATE);
            const uint32_t* w_s = reinterpret_cast<const uint32_t*>(smem + OFF_GATE + kMaxRopeEntries * 8);
            const int kind_cols = p.heads * 128;
            uint32_t tc = 0;
            for (int t = first; t < n_tiles_total; t += stride, ++tc) {
                int mt, nt;
                decode_tile(p, t, mt, nt);
                const uint32_t b = tc & 1;
                const int64_t row = (int64_t)mt * (2 * BM) + (int)rank * BM + q * 32 + lane;
                const int col_tile = nt * BN;
                const int kind = col_tile / kind_cols;                      // 0 q, 1 k, 2 v (a tile never straddles: heads is even)
                const int head = (col_tile - kind * kind_cols) / 128 + hsel;
                const bool valid = row < p.M;
                int4 tk = make_int4(0, -1, -1, -1);
                if (valid) tk = __ldg(p.tok_map + row);
                mbar_wait(bar(BAR_TFULL + b), (tc >> 1) & 1);
                tc_fence_after();
                const uint32_t tacc = tmem + lane_off + b * BN + hsel * 128;
                uint32_t xb[64];                    // the head's 128 linear outputs of this row, packed bf16 (the Linear's output dtype)
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    uint32_t v0[32], v1[32];
                    tmem_ld32(tacc + c * 64, v0);
                    tmem_ld32(tacc + c * 64 + 32, v1);
                    tc_wait_ld();
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        xb[c * 32 + i] = pack_bf16x2(__uint_as_float(v0[2 * i]), __uint_as_float(v0[2 * i + 1]));
                        xb[c * 32 + 16 + i] = pack_bf16x2(__uint_as_float(v1[2 * i]), __uint_as_float(v1[2 * i + 1]));
                    }
                }
                // the accumulator is in registers: hand the TMEM buffer back to the MMA issuer
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive_cluster(b ? tempty_g1 : tempty_g0);
                if (valid) {
                if (kind < 2) {
                    // CustomRMSNorm over the head (normalization.py:88-109) with its bf16 rounding points: x.pow(2) is a bf16 tensor,
                    // its mean an fp32 sum rounded to bf16, (var + eps) and sqrt round again, x / rms and * weight round each
                    float ss0 = 0.f, ss1 = 0.f, ss2 = 0.f, ss3 = 0.f;
#pragma unroll
                    for (int i = 0; i < 64; i += 4) {
                        const float2 a = __bfloat1622float2(__hmul2(as_bf162(xb[i]), as_bf162(xb[i])));
                        const float2 bb = __bfloat1622float2(__hmul2(as_bf162(xb[i + 1]), as_bf162(xb[i + 1])));
                        const float2 cc = __bfloat1622float2(__hmul2(as_bf162(xb[i + 2]), as_bf162(xb[i + 2])));
                        const float2 dd = __bfloat1622float2(__hmul2(as_bf162(xb[i + 3]), as_bf162(xb[i + 3])));
                        ss0 += a.x + a.y; ss1 += bb.x + bb.y; ss2 += cc.x + cc.y; ss3 += dd.x + dd.y;
                    }
                    const float ss = (ss0 + ss1) + (ss2 + ss3);
                    const float var = bf16r(ss * (1.0f / 128.0f));
                    const float rms = bf16r(sqrt_approx(bf16r(var + p.eps)));
                    const float rcp = rcp_approx(rms);
                    const bool is_txt = tk.x < 0;
                    const uint32_t* w2 = w_s + ((is_txt ? 2 : 0) + kind) * 64;
                    const bool rot = tk.y >= 0;
                    const int r0 = rot ? tk.y * NF : 0, r1 = rot ? tk.z * NF : 0, r2 = rot ? tk.w * NF : 0;
#pragma unroll
                    for (int i = 0; i < 64; ++i) {
                        const float2 xf = __bfloat1622float2(as_bf162(xb[i]));
                        const __nv_bfloat162 nb = __floats2bfloat162_rn(xf.x * rcp, xf.y * rcp);       // x / rms -> bf16
                        __nv_bfloat162 y = __hmul2(nb, as_bf162(w2[i]));                               // * weight -> bf16
                        if (i < NROT) {
                            // interleaved-pair rotation in fp32, mul and add rounded separately like the two ATen ops
                            const int axis = i / NF, fi = i - axis * NF;
                            float2 cs = make_float2(1.f, 0.f);
                            if (rot) cs = cs_s[(axis == 0 ? r0 : (axis == 1 ? r1 : r2)) + fi];
                            const float2 yf = __bfloat1622float2(y);
                            const float o0 = __fadd_rn(__fmul_rn(yf.x, cs.x), __fmul_rn(-yf.y, cs.y));
                            const float o1 = __fadd_rn(__fmul_rn(yf.y, cs.x), __fmul_rn(yf.x, cs.y));
                            y = __floats2bfloat162_rn(o0, o1);
                        }
                        xb[i] = as_u32(y);
                    }
                }
                // stage the 256-byte head row in this thread's private line, then one bulk copy to its packed position
                bulk_wait_read<0>();                 // the previous copy out of this line has finished reading it
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    *reinterpret_cast<uint4*>(my_row + j * 16) = make_uint4(xb[4 * j], xb[4 * j + 1], xb[4 * j + 2], xb[4 * j + 3]);
                fence_proxy_async_smem();
                __nv_bfloat16* dst;
                if (tk.x >= 0) {
                    __nv_bfloat16* pk = kind == 0 ? p.q_pack : (kind == 1 ? p.k_pack : p.v_pack);
                    dst = pk + ((int64_t)tk.x * p.heads + head) * 128;
                } else {
                    dst = p.txt_qkv + (((int64_t)(-(tk.x + 1)) * 3 + kind) * p.heads + head) * 128;
                }
                bulk_store_256(dst, my_row_s);
                bulk_commit();
                }
                __syncwarp();
            }
            bulk_wait<0>();
        }
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue: one thread per accumulator row
        const int q = warp & 3;
        const uint32_t lane_off = (uint32_t)(q * 32) << 16;
        unsigned char* slab0 = smem + OFF_SLAB + q * 2 * SLAB_BYTES;
        const uint32_t tempty_g0 = mapa(bar(BAR_TEMPTY + 0), 0), tempty_g1 = mapa(bar(BAR_TEMPTY + 1), 0);
        const __nv_bfloat16* gate_s = reinterpret_cast<const __nv_bfloat16*>(smem + OFF_GATE);
        uint32_t tc = 0, n_store = 0;
        for (int t = first; t < n_tiles_total; t += stride, ++tc) {
            int mt, nt;
            decode_tile(p, t, mt, nt);
            const uint32_t b = tc & 1;
            const int row0 = mt * (2 * BM) + (int)rank * BM + q * 32;      // first row of this warp's slabs
            const int64_t row = row0 + lane;
            const int col_tile = nt * OUT_N;
            const int n_slabs = min(OUT_N / 64, (p.N - col_tile) / 64);    // >= 1; N % 64 == 0
            mbar_wait(bar(BAR_TFULL + b), (tc >> 1) & 1);
            tc_fence_after();
            const uint32_t tacc = tmem + lane_off + b * BN;
#pragma unroll 1
            for (int sl = 0; sl < n_slabs; ++sl) {
                const int col0 = col_tile + sl * 64;
                uint32_t o[32];                     // 64 output columns of this row, packed bf16
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    const int c = col0 + hf * 32;   // first output column of this 32-column half
                    uint32_t v[32];
                    if constexpr (EPI == EPI_SWIGLU) {
                        uint32_t u[32];
                        tmem_ld32(tacc + sl * 64 + hf * 32, v);               // gate projection
                        tmem_ld32(tacc + BN / 2 + sl * 64 + hf * 32, u);      // up projection
                        tc_wait_ld();
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            // g, u -> bf16 (the two linears' outputs); silu in fp32 -> bf16; product -> bf16
                            const float g0 = bf16r(__uint_as_float(v[2 * i])), g1 = bf16r(__uint_as_float(v[2 * i + 1]));
                            const uint32_t s2 = pack_bf16x2(silu_f(g0), silu_f(g1));
                            const uint32_t u2 = pack_bf16x2(__uint_as_float(u[2 * i]), __uint_as_float(u[2 * i + 1]));
                            o[hf * 16 + i] = as_u32(__hmul2(as_bf162(s2), as_bf162(u2)));
                        }
                    } else {
                        tmem_ld32(tacc + sl * 64 + hf * 32, v);
                        uint4 bq[4], rq[4];
                        if (p.bias != nullptr) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) bq[j] = __ldg(reinterpret_cast<const uint4*>(p.bias + c) + j);
                        }
                        if constexpr (EPI == EPI_GATE_RESID) {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                rq[j] = row < p.M ? __ldg(reinterpret_cast<const uint4*>(p.resid + row * p.N + c) + j)
                                                  : make_uint4(0u, 0u, 0u, 0u);
                        }
                        tc_wait_ld();
                        const uint32_t* bw = reinterpret_cast<const uint32_t*>(bq);
                        const uint32_t* rw = reinterpret_cast<const uint32_t*>(rq);
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            float a0 = __uint_as_float(v[2 * i]), a1 = __uint_as_float(v[2 * i + 1]);
                            if (p.bias != nullptr) {
                                const float2 bf = __bfloat1622float2(as_bf162(bw[i]));
                                a0 += bf.x; a1 += bf.y;
                            }
                            uint32_t h2 = pack_bf16x2(a0, a1);              // the linear's bf16 output
                            if constexpr (EPI == EPI_GELU) {
                                const float2 hf2 = __bfloat1622float2(as_bf162(h2));
                                h2 = pack_bf16x2(gelu_tanh_f(hf2.x), gelu_tanh_f(hf2.y));
                            }
                            if constexpr (EPI == EPI_GATE_RESID) {
                                const __nv_bfloat162 g2 = *reinterpret_cast<const __nv_bfloat162*>(gate_s + c + 2 * i);
                                const __nv_bfloat162 t2 = __hmul2_rn(as_bf162(h2), g2);            // mul_ -> bf16
                                h2 = as_u32(__hadd2_rn(t2, as_bf162(rw[i])));                      // + resid -> bf16
                            }
                            o[hf * 16 + i] = h2;
                        }
                    }
                }
                if (sl == n_slabs - 1) {
                    // the accumulator is in registers: hand the TMEM buffer back to the MMA issuer
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive_cluster(b ? tempty_g1 : tempty_g0);
                }
                // stage the 32 x 64 slab (128-byte swizzle: 16-byte chunk j of row r at r * 128 + ((j ^ (r & 7)) << 4))
                unsigned char* slab = slab0 + (n_store & 1) * SLAB_BYTES;
                if (lane == 0) bulk_wait_read<1>();          // the store that last read this slab has finished reading it
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<uint4*>(slab + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                        make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0 && row0 < p.M) {
                    tma_store_2d(&tm_c, smem_u32(slab), col0, row0);
                    bulk_commit();
                }
                ++n_store;
            }
        }
        if (lane == 0) bulk_wait<0>();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();              // nothing of this CTA may disappear while the peer can still signal or read it
    if (warp == 2) {
        __syncwarp();
        tmem_dealloc_pair(tmem, TMEM_COLS);
    }
}

template <int EPI>
int launch(const void* x, const void* w0, const void* w1, void* y, const Params& p0, cudaStream_t st, const char* name) {
    Params p = p0;
    constexpr int OUT_N = EPI == EPI_SWIGLU ? BN / 2 : BN;
    p.m_tiles = (int)((p.M + 2 * BM - 1) / (2 * BM));
    p.n_tiles = (p.N + OUT_N - 1) / OUT_N;
    CUtensorMap ta, tb0, tb1, tc;
    int rc;
    if ((rc = make_tmap_2d(&ta, x, p.M, p.K, p.K, BM)) != SVR2_OK) return rc;
    if ((rc = make_tmap_2d(&tb0, w0, p.N, p.K, p.K, BN / 2)) != SVR2_OK) return rc;
    if ((rc = make_tmap_2d(&tb1, w1 ? w1 : w0, p.N, p.K, p.K, BN / 2)) != SVR2_OK) return rc;
    if (Cfg<EPI>::QKV) tc = ta;        // no dense output: the epilogue scatters head rows with bulk copies
    else if ((rc = make_tmap_2d(&tc, y, p.M, p.N, p.N, 32)) != SVR2_OK) return rc;
    static PerDeviceFlag smem_ok;
    if ((rc = ensure_dyn_smem(linear_tc_kernel<EPI>, Cfg<EPI>::SMEM_BYTES, smem_ok, name)) != SVR2_OK) return rc;
    const int pairs = sm_count() / 2;
    const int tiles = p.m_tiles * p.n_tiles;
    const int grid = 2 * (tiles < pairs ? tiles : pairs);
    linear_tc_kernel<EPI><<<grid, Cfg<EPI>::NTHREADS, Cfg<EPI>::SMEM_BYTES, st>>>(ta, tb0, tb1, tc, p);
    SVR2_CHECK_LAUNCH(name);
    return SVR2_OK;
}

int check_shapes(const char* name, int64_t M, int N, int K) {
    SVR2_CHECK_ARG(M > 0 && M < ((int64_t)1 << 31) - 512, "%s: M=%lld out of range", name, (long long)M);
    SVR2_CHECK_ARG(N > 0 && N % 64 == 0, "%s: N=%d must be a positive multiple of 64", name, N);
    SVR2_CHECK_ARG(K > 0 && K % 8 == 0, "%s: K=%d must be a positive multiple of 8 (16-byte rows)", name, K);
    return SVR2_OK;
}

}  // namespace gemm
}  // namespace svr2

using svr2::gemm::Params;

extern "C" int svr2_linear_fwd(const void* x, const void* weight, const void* bias, void* y, int64_t M, int N, int K,
                               int activation, void* stream) {
    SVR2_CHECK_ARG(x && weight && y, "svr2_linear_fwd: null pointer");
    SVR2_CHECK_ARG(activation == SVR2_ACT_NONE || activation == SVR2_ACT_GELU_TANH, "svr2_linear_fwd: unknown activation %d", activation);
    if (M == 0) return SVR2_OK;
    int rc = svr2::gemm::check_shapes("svr2_linear_fwd", M, N, K);
    if (rc != SVR2_OK) return rc;
    Params p{};
    p.M = M; p.N = N; p.K = K; p.bias = (const __nv_bfloat16*)bias;
    if (activation == SVR2_ACT_GELU_TANH)
        return svr2::gemm::launch<svr2::gemm::EPI_GELU>(x, weight, nullptr, y, p, (cudaStream_t)stream, "svr2_linear_fwd(gelu)");
    return svr2::gemm::launch<svr2::gemm::EPI_BIAS>(x, weight, nullptr, y, p, (cudaStream_t)stream, "svr2_linear_fwd");
}

extern "C" int svr2_linear_swiglu_fwd(const void* x, const void* w_gate, const void* w_up, void* y, int64_t M, int N, int K,
                                      void* stream) {
    SVR2_CHECK_ARG(x && w_gate && w_up && y, "svr2_linear_swiglu_fwd: null pointer");
    if (M == 0) return SVR2_OK;
    int rc = svr2::gemm::check_shapes("svr2_linear_swiglu_fwd", M, N, K);
    if (rc != SVR2_OK) return rc;
    SVR2_CHECK_ARG(N % 128 == 0, "svr2_linear_swiglu_fwd: hidden size %d must be a multiple of 128", N);
    Params p{};
    p.M = M; p.N = N; p.K = K;
    return svr2::gemm::launch<svr2::gemm::EPI_SWIGLU>(x, w_gate, w_up, y, p, (cudaStream_t)stream, "svr2_linear_swiglu_fwd");
}

extern "C" int svr2_linear_gate_residual_fwd(const void* x, const void* weight, const void* bias, const void* emb,
                                             const void* gateB, int layer, const void* resid, void* out, int64_t M, int N,
                                             int K, void* stream) {
    SVR2_CHECK_ARG(x && weight && emb && gateB && resid && out, "svr2_linear_gate_residual_fwd: null pointer");
    SVR2_CHECK_ARG(layer == 0 || layer == 1, "svr2_linear_gate_residual_fwd: layer must be 0 (attn) or 1 (mlp)");
    if (M == 0) return SVR2_OK;
    int rc = svr2::gemm::check_shapes("svr2_linear_gate_residual_fwd", M, N, K);
    if (rc != SVR2_OK) return rc;
    SVR2_CHECK_ARG(N <= svr2::gemm::kMaxGateN, "svr2_linear_gate_residual_fwd: N=%d exceeds %d", N, svr2::gemm::kMaxGateN);
    Params p{};
    p.M = M; p.N = N; p.K = K; p.bias = (const __nv_bfloat16*)bias;
    p.emb = (const __nv_bfloat16*)emb; p.gateB = (const __nv_bfloat16*)gateB; p.resid = (const __nv_bfloat16*)resid; p.layer = layer;
    return svr2::gemm::launch<svr2::gemm::EPI_GATE_RESID>(x, weight, nullptr, out, p, (cudaStream_t)stream,
                                                          "svr2_linear_gate_residual_fwd");
}

extern "C" int svr2_linear_qkv_pack_fwd(const void* x, const void* weight, int64_t M, int K, int heads, const int32_t* tok_map,
                                        const float* rope_cs, int rope_rows, int n_freqs, const void* wq_vid, const void* wk_vid,
                                        const void* wq_txt, const void* wk_txt, float eps, void* q_pack, void* k_pack,
                                        void* v_pack, void* txt_qkv, void* stream) {
    SVR2_CHECK_ARG(x && weight && tok_map && wq_vid && wk_vid && q_pack && k_pack && v_pack, "svr2_linear_qkv_pack_fwd: null pointer");
    SVR2_CHECK_ARG(heads > 0 && heads % 2 == 0, "svr2_linear_qkv_pack_fwd: heads=%d must be even (a 256-column tile holds two whole heads)", heads);
    SVR2_CHECK_ARG(n_freqs == 0 || n_freqs == 21 || n_freqs == 10,
                   "svr2_linear_qkv_pack_fwd: n_freqs=%d — built for 21 (3B), 10 (7B) or 0 (no rope) frequencies per axis", n_freqs);
    SVR2_CHECK_ARG(n_freqs == 0 || (rope_cs && rope_rows > 0 && (int64_t)rope_rows * n_freqs <= svr2::gemm::kMaxRopeEntries),
                   "svr2_linear_qkv_pack_fwd: rope table of %d rows x %d frequencies exceeds the %d entries held in shared memory",
                   rope_rows, n_freqs, svr2::gemm::kMaxRopeEntries);
    if (M == 0) return SVR2_OK;
    const int N = 3 * heads * 128;
    int rc = svr2::gemm::check_shapes("svr2_linear_qkv_pack_fwd", M, N, K);
    if (rc != SVR2_OK) return rc;
    Params p{};
    p.M = M; p.N = N; p.K = K;
    p.tok_map = (const int4*)tok_map; p.rope_cs = (const float2*)rope_cs; p.rope_entries = n_freqs ? rope_rows * n_freqs : 0;
    p.wq_vid = (const __nv_bfloat16*)wq_vid; p.wk_vid = (const __nv_bfloat16*)wk_vid;
    p.wq_txt = (const __nv_bfloat16*)wq_txt; p.wk_txt = (const __nv_bfloat16*)wk_txt;
    p.eps = eps; p.heads = heads;
    p.q_pack = (__nv_bfloat16*)q_pack; p.k_pack = (__nv_bfloat16*)k_pack; p.v_pack = (__nv_bfloat16*)v_pack;
    p.txt_qkv = (__nv_bfloat16*)txt_qkv;
    const char* name = "svr2_linear_qkv_pack_fwd";
    if (n_freqs == 21) return svr2::gemm::launch<svr2::gemm::EPI_QKV21>(x, weight, nullptr, nullptr, p, (cudaStream_t)stream, name);
    if (n_freqs == 10) return svr2::gemm::launch<svr2::gemm::EPI_QKV10>(x, weight, nullptr, nullptr, p, (cudaStream_t)stream, name);
    return svr2::gemm::launch<svr2::gemm::EPI_QKV0>(x, weight, nullptr, nullptr, p, (cudaStream_t)stream, name);
}
