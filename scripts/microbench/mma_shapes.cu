// tcgen05.mma issue-rate microbenchmark for sm_100a: cycles per instruction for kind::f16 shapes, operands in shared memory
// (SS) or A in tensor memory (TS), cta_group::1 and ::2.  One CTA (pair) per launch, ITER back-to-back MMAs, one commit.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../comfyui-seedvr2_b200/csrc -o mma_shapes mma_shapes.cu
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>

#include "sm100_pair.cuh"
#include "sm100_ptx.cuh"

using namespace svr2::ptx;
using namespace svr2::pairx;

#define ITER 4096

__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// CG: cta_group; M: instruction M (per pair for CG=2); N; TS: A operand from TMEM; DISTINCT: rotate A / B through 4 smem tiles
template <int CG, int M, int N, bool TS, bool DISTINCT>
__global__ void __launch_bounds__(128, 1) k(long long* cyc) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t slot;
    __shared__ __align__(8) uint64_t barmem;
    const uint32_t base = smem_u32(smem);
    const uint32_t bar = smem_u32(&barmem);
    const int warp = threadIdx.x >> 5;
    const uint32_t rank = CG == 2 ? cluster_rank() : 0;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        fence_mbar_init();
    }
    if (warp == 1) {
        if (CG == 2) tmem_alloc_pair(smem_u32(&slot), 512);
        else tmem_alloc(smem_u32(&slot), 512);
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = slot;
    constexpr uint32_t idesc = idesc_f16(M, N);
    constexpr uint32_t hi = smem_desc_hi_sw128(1024);
    long long t0 = 0, t1 = 0;
    if (warp == 0 && rank == 0) {
        if (elect_one()) {
            t0 = clock64();
            for (int i = 0; i < ITER; ++i) {
                const uint32_t a_off = DISTINCT ? (uint32_t)(i & 3) * 16384u : 0u;
                const uint32_t b_off = 65536u + (DISTINCT ? (uint32_t)(i & 3) * 32768u : 0u);
                const uint64_t bd = make_desc(smem_desc_lo(base + b_off, 16) + (i & 3) * 2 * (DISTINCT ? 0 : 1), hi);
                if (TS) {
                    if (CG == 2) mma2_ts(tmem, tmem + 256, bd, idesc, 1);
                    else mma_ts(tmem, tmem + 256, bd, idesc, 1);
                } else {
                    const uint64_t ad = make_desc(smem_desc_lo(base + a_off, 16) + (i & 3) * 2 * (DISTINCT ? 0 : 1), hi);
                    if (CG == 2) mma2_ss(tmem, ad, bd, idesc, 1);
                    else mma_ss(tmem, ad, bd, idesc, 1);
                }
            }
            if (CG == 2) tc_commit_pair(bar);
            else tc_commit(bar);
        }
        __syncwarp();
        mbar_wait(bar, 0);
        t1 = clock64();
        if (threadIdx.x == 0) cyc[0] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();
    if (warp == 1) {
        if (CG == 2) tmem_dealloc_pair(tmem, 512);
        else tmem_dealloc(tmem, 512);
    }
}

template <int CG, int M, int N, bool TS, bool DISTINCT>
void run(long long* d) {
    auto fn = k<CG, M, N, TS, DISTINCT>;
    const int smem = 200 * 1024;
    cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(CG, 1, 1);
    cfg.blockDim = dim3(128, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = CG;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    for (int rep = 0; rep < 2; ++rep) {
        cudaError_t e = cudaLaunchKernelEx(&cfg, fn, d);
        if (e != cudaSuccess) { printf("launch failed: %s\n", cudaGetErrorString(e)); return; }
        e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(e)); return; }
    }
    long long h = 0;
    cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    const double per = (double)h / ITER;
    const double flop_per_sm = 2.0 * M * N * 16 / CG;
    printf("cta_group::%d M=%3d N=%3d %s %-8s : %7.1f cycles / MMA  -> %6.0f FLOP/clk/SM\n", CG, M, N, TS ? "TS" : "SS",
           DISTINCT ? "4 tiles" : "1 tile", per, flop_per_sm / per);
}

int main() {
    long long* d;
    cudaMalloc(&d, 8);
    run<1, 128, 256, false, false>(d);
    run<1, 128, 128, false, false>(d);
    run<1, 128, 64, false, false>(d);
    run<1, 64, 256, false, false>(d);
    run<2, 256, 256, false, false>(d);
    run<2, 256, 128, false, false>(d);
    run<2, 256, 64, false, false>(d);
    run<2, 128, 256, false, false>(d);
    run<2, 128, 128, false, false>(d);
    run<2, 256, 256, false, true>(d);
    run<2, 256, 128, false, true>(d);
    run<1, 128, 128, false, true>(d);
    run<1, 128, 128, true, false>(d);
    run<2, 256, 128, true, false>(d);
    run<2, 256, 256, true, false>(d);
    return 0;
}
