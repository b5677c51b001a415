// Interaction of tcgen05.mma (accumulating into TMEM) with concurrent tcgen05.ld / MUFU work, one CTA on one SM.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -I../../comfyui-seedvr2_b200/csrc -o tmem3 tmem3.cu
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "sm100_ptx.cuh"
using namespace svr2::ptx;

// warps 0-3: LD_MODE 0 idle, 1 = ld32 loop on cols [0,128), 2 = ld32 + 32 MUFU per load, 3 = MUFU only
// warp 4 lane 0: MMA_MODE 0 none, 1 = SS MMAs (A,B smem) into cols [256,384), 2 = TS MMAs (A from TMEM cols [128,192))
template <int LD_MODE, int MMA_MODE>
__global__ void __launch_bounds__(160, 1) k(long long* res, int n_ld, int n_mma) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) unsigned long long bar_mem;
    const uint32_t base = smem_u32(smem), bar = smem_u32(&bar_mem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
    if (warp == 4) tmem_alloc(smem_u32(&slot), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = slot;
    if (warp < 4) {
        const uint32_t t = tmem + ((uint32_t)(warp * 32) << 16);
        uint32_t a[32]; float acc = 0.f;
        long long t0 = clock64();
        if (LD_MODE != 0) {
#pragma unroll 1
            for (int it = 0; it < n_ld; ++it) {
                if (LD_MODE == 1 || LD_MODE == 2) { tmem_ld32(t + (it & 3) * 32, a); tc_wait_ld(); acc += __uint_as_float(a[0]) + __uint_as_float(a[31]); }
                if (LD_MODE == 2 || LD_MODE == 3) {
#pragma unroll
                    for (int c = 0; c < 32; ++c) { float x = (LD_MODE == 2) ? __uint_as_float(a[c]) : acc + c; float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); acc += y; }
                }
            }
        }
        long long t1 = clock64();
        if (lane == 0) res[warp] = t1 - t0;
        if (acc == 123.456f) res[10] = 1;
    } else if (lane == 0) {
        long long t0 = clock64();
        if (MMA_MODE != 0) {
            const uint32_t idesc = idesc_bf16(128, 128, 0);
            uint32_t phase = 0;
            for (int it = 0; it < n_mma; it += 16) {
                for (int kk = 0; kk < 16; ++kk) {
                    const uint32_t off = (uint32_t)((kk & 7) >> 2) * 16384u + (uint32_t)(kk & 3) * 32u;
                    if (MMA_MODE == 1) mma_ss(tmem + 256, smem_desc_sw128(base + off, 16, 1024), smem_desc_sw128(base + 32768 + off, 16, 1024), idesc, 1);
                    else mma_ts(tmem + 256, tmem + 128 + (kk & 7) * 8, smem_desc_sw128(base + 32768 + off, 16, 1024), idesc, 1);
                }
                tc_commit(bar);
                mbar_wait(bar, phase); phase ^= 1;      // wait for the batch of 16: one 128x128x128 block pair
            }
        }
        long long t1 = clock64();
        res[4] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem, 512);
}

template <int L, int M>
void run(const char* name) {
    long long* res; cudaMalloc(&res, 128); cudaMemset(res, 0, 128);
    auto kern = k<L, M>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    const int n_ld = 2048, n_mma = 4096;
    kern<<<1, 160, 65536>>>(res, n_ld, n_mma);
    kern<<<1, 160, 65536>>>(res, n_ld, n_mma);
    cudaDeviceSynchronize();
    long long h[8]; cudaMemcpy(h, res, 64, cudaMemcpyDeviceToHost);
    printf("%-44s ld-warp: %7.1f cycles per iteration | MMA: %6.1f cycles per 128x128x16 MMA   (%s)\n", name,
           L ? (double)h[0] / n_ld : 0.0, M ? (double)h[4] / n_mma : 0.0, cudaGetErrorString(cudaGetLastError()));
    cudaFree(res);
}

int main() {
    run<1, 0>("ld32 alone");
    run<0, 1>("SS MMA alone");
    run<0, 2>("TS MMA alone");
    run<1, 1>("ld32 + SS MMA");
    run<1, 2>("ld32 + TS MMA");
    run<3, 0>("32 MUFU alone");
    run<2, 0>("ld32 + 32 MUFU, no MMA");
    run<2, 1>("ld32 + 32 MUFU + SS MMA");
    run<2, 2>("ld32 + 32 MUFU + TS MMA");
    run<3, 1>("32 MUFU + SS MMA");
    return 0;
}
