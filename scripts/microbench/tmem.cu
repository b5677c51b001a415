// TMEM -> register (tcgen05.ld) throughput microbenchmark for sm_100a, one CTA on one SM, no MMA running.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tmem tmem.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITER 2048

__device__ __forceinline__ void ld32(uint32_t taddr, uint32_t* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void ld16(uint32_t taddr, uint32_t* v) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void st16(uint32_t taddr, const uint32_t* v) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
          "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// MODE 0: x32, 1 in flight; 1: x32, 2 in flight; 2: x32, 3 in flight; 3: x16, 4 in flight; 4: x16 stores, 2 in flight
template <int MODE>
__global__ void k(uint32_t* out, long long* cyc) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 128) % 512;
    uint32_t a[32], b[32], c[32], acc = 0;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
        if (MODE == 0) { ld32(tmem, a); wait_ld(); acc ^= a[it & 31]; }
        if (MODE == 1) { ld32(tmem, a); ld32(tmem + 32, b); wait_ld(); acc ^= a[it & 31] ^ b[(it + 1) & 31]; }
        if (MODE == 2) { ld32(tmem, a); ld32(tmem + 32, b); ld32(tmem + 64, c); wait_ld(); acc ^= a[it & 31] ^ b[(it + 1) & 31] ^ c[(it + 2) & 31]; }
        if (MODE == 3) { ld16(tmem, a); ld16(tmem + 16, a + 16); ld16(tmem + 32, b); ld16(tmem + 48, b + 16); wait_ld(); acc ^= a[it & 31] ^ b[(it + 1) & 31]; }
        if (MODE == 4) { a[0] = it; st16(tmem, a); st16(tmem + 16, a + 16); wait_st(); }
    }
    long long t1 = clock64();
    out[threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(slot) : "memory");
}

template <int MODE>
void run(const char* name, int bytes_per_warp_iter) {
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, 4096); cudaMalloc(&cyc, 8);
    for (int warps : {4, 8, 12, 16}) {
        k<MODE><<<1, warps * 32>>>(out, cyc);
        k<MODE><<<1, warps * 32>>>(out, cyc);
        cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        printf("%-34s warps %2d: %7.1f cycles/iter  -> %.1f B/clk/SM\n", name, warps, (double)c / ITER, (double)bytes_per_warp_iter * warps * ITER / c);
    }
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("ld 32x32b.x32, 1 in flight", 4096);
    run<1>("ld 32x32b.x32, 2 in flight", 8192);
    run<2>("ld 32x32b.x32, 3 in flight", 12288);
    run<3>("ld 32x32b.x16, 4 in flight", 8192);
    run<4>("st 32x32b.x16, 2 in flight", 4096);
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
