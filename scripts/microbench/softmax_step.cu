// One softmax step of the window-attention kernel in isolation: 4 x tcgen05.ld of a 128-column S row, row max, exp2 + bf16
// pack + tcgen05.st of P, per-phase cycle counts, with 1 or 2 warps per SM sub-partition (= one or both query-tile groups).
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -I../../comfyui-seedvr2_b200/csrc -o softmax_step softmax_step.cu
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "sm100_ptx.cuh"
#include "softmax_row.cuh"
using namespace svr2::ptx;
using namespace svr2::rowsm;

__device__ __forceinline__ uint32_t clk() { uint32_t c; asm volatile("mov.u32 %0, %%clock;" : "=r"(c)); return c; }

// MAXV: 0 = FMNMX3, 4 chains per chunk (kernel as of gen 4); 1 = 2-input FMNMX, 8 chains over the row; 2 = FMNMX3, 8 chains over the row
// EXPV: 0 = all MUFU; 1 = every 4th element on the FMA pipe; 2 = every 2nd pair ... see chunk_exp_mix
template <int MAXV, int EXPV, int UNROLL, int MMA>
__global__ void __launch_bounds__(288, 1) k(long long* res, int n_it, int nwarps, volatile int* stop) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar_mem[2];
    __shared__ int done_warps;
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar_mem[0]), 1); mbar_init(smem_u32(&bar_mem[1]), 1); fence_mbar_init(); done_warps = 0; }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(smem_u32(&slot), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = slot;
    if (warp == 8) {
        if (MMA != 0) {
            const uint32_t base = smem_u32(smem), bar = smem_u32(&bar_mem[0]);
            const uint32_t idesc = idesc_bf16(128, 128, 0);
            uint32_t nb = 0;
            long long t0 = clock64();
            while (*(volatile int*)&done_warps < nwarps) {
                if (nb >= 2) mbar_wait(bar + 8 * (nb & 1), ((nb >> 1) - 1) & 1);
                if (elect_one()) {
#pragma unroll
                    for (int kk = 0; kk < 16; ++kk) {
                        const uint32_t off = (uint32_t)((kk & 7) >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
                        const bool ts = MMA == 2 || (MMA == 3 && kk < 8);
                        if (!ts) mma_ss(tmem + 256, smem_desc_sw128(base + off, 16, 1024), smem_desc_sw128(base + 32768 + off, 16, 1024), idesc, 1);
                        else mma_ts(tmem + 384, tmem + 256 + 64 + (kk & 7) * 8, smem_desc_sw128(base + 32768 + off, 16, 1024), idesc, 1);
                    }
                    tc_commit(bar + 8 * (nb & 1));
                }
                __syncwarp();
                ++nb;
            }
            for (uint32_t b = (nb >= 2 ? nb - 2 : 0); b < nb; ++b) mbar_wait(bar + 8 * (b & 1), (b >> 1) & 1);
            if (lane == 0) { res[200] = clock64() - t0; res[201] = nb * 16; }
        }
    } else if (warp < nwarps) {
        const uint32_t tS = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;
        uint32_t s[128];
#pragma unroll
        for (int c = 0; c < 128; ++c) s[c] = __float_as_uint(-0.37f * (float)((c * 7 + lane * 3 + warp) % 41));
        tmem_st32(tS, s); tmem_st32(tS + 32, s + 32); tmem_st32(tS + 64, s + 64); tmem_st32(tS + 96, s + 96);
        tc_wait_st();
        uint32_t t_ld = 0, t_max = 0, t_exp = 0, t_st = 0;
        float l = 0.f;
        const uint32_t t_begin = clk();
#pragma unroll UNROLL
        for (int it = 0; it < n_it; ++it) {
            const uint32_t c0 = clk();
            tmem_ld32(tS, s); tmem_ld32(tS + 32, s + 32); tmem_ld32(tS + 64, s + 64); tmem_ld32(tS + 96, s + 96);
            tc_wait_ld();
            { uint32_t d; asm volatile("mov.u32 %0, %1;" : "=r"(d) : "r"(s[127])); }
            const uint32_t c1 = clk();
            float mx;
            if (MAXV == 0) {
                mx = fmaxf(fmaxf(chunk_max(s), chunk_max(s + 32)), fmaxf(chunk_max(s + 64), chunk_max(s + 96)));
            } else if (MAXV == 1) {
                float m[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) m[q] = __uint_as_float(s[q]);
#pragma unroll
                for (int c = 8; c < 128; c += 8)
#pragma unroll
                    for (int q = 0; q < 8; ++q) m[q] = fmaxf(m[q], __uint_as_float(s[c + q]));
                mx = fmaxf(fmaxf(fmaxf(m[0], m[1]), fmaxf(m[2], m[3])), fmaxf(fmaxf(m[4], m[5]), fmaxf(m[6], m[7])));
            } else {
                mx = row_max128(s);
            }
            { uint32_t d; asm volatile("mov.u32 %0, %1;" : "=r"(d) : "r"(__float_as_uint(mx))); }
            const uint32_t c2 = clk();
            const float mb = mx * kScaleLog2;
            float rs = 0.f;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t pk[16];
                if (EXPV == 0) rs += chunk_exp(s + 32 * c, pk, mb);
                else rs += chunk_exp_mix<EXPV>(s + 32 * c, pk, mb);
                tmem_st16(tS + 16 * c, pk);
            }
            l += rs;
            const uint32_t c3 = clk();
            tc_wait_st();
            const uint32_t c4 = clk();
            t_ld += c1 - c0; t_max += c2 - c1; t_exp += c3 - c2; t_st += c4 - c3;
        }
        const uint32_t t_end = clk();
        if (lane == 0) {
            long long* r = res + warp * 8;
            r[0] = t_ld; r[1] = t_max; r[2] = t_exp; r[3] = t_st; r[4] = t_end - t_begin;
        }
        if (l == 123.456f) res[255] = 1;
        if (lane == 0) atomicAdd(&done_warps, 1);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int M, int E, int U = 1, int MMA = 0>
void run(const char* name, int nwarps) {
    long long* res; cudaMalloc(&res, 2048); cudaMemset(res, 0, 2048);
    const int n = 512;
    auto kern = k<M, E, U, MMA>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    kern<<<1, 288, 65536>>>(res, n, nwarps, nullptr);
    kern<<<1, 288, 65536>>>(res, n, nwarps, nullptr);
    cudaDeviceSynchronize();
    long long h[256]; cudaMemcpy(h, res, 2048, cudaMemcpyDeviceToHost);
    if (MMA) printf("   [MMA stream: %.1f cycles per 128x128x16] ", (double)h[200] / (double)h[201]);
    printf("%-46s warps/sub-partition %d | ld %6.1f | max %6.1f | exp+pack+st %7.1f | wait::st %5.1f | step %7.1f cycles   (%s)\n", name,
           nwarps / 4, (double)h[0] / n, (double)h[1] / n, (double)h[2] / n, (double)h[3] / n, (double)h[4] / n,
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(res);
}

int main() {
    for (int nw = 4; nw <= 8; nw += 4) {
        run<0, 0>("max: FMNMX3 4 chains/chunk, exp: MUFU", nw);
        run<1, 0>("max: FMNMX 8 chains,        exp: MUFU", nw);
        run<2, 0>("max: FMNMX3 8 chains,       exp: MUFU", nw);
        run<2, 1>("max: FMNMX3 8 chains, exp: 1/8 on FMA pipe", nw);
        run<2, 2>("max: FMNMX3 8 chains, exp: 1/4 on FMA pipe", nw);
        run<2, 3>("max: FMNMX3 8 chains, exp: 3/8 on FMA pipe", nw);
        run<2, 4>("max: FMNMX3 8 chains, exp: 1/2 on FMA pipe", nw);
        run<2, 3, 8>("3/8 FMA, loop body unrolled x8 (~80 KB)", nw);
        run<2, 3, 1, 1>("3/8 FMA + SS MMA stream", nw);
        run<2, 3, 1, 2>("3/8 FMA + TS MMA stream", nw);
        run<2, 3, 1, 3>("3/8 FMA + 8 TS / 8 SS MMA stream", nw);
        run<2, 0, 1, 3>("all MUFU + 8 TS / 8 SS MMA stream", nw);
    }
    return 0;
}
