// Pipe-rate microbenchmark for sm_100a: cycles per warp-instruction per SM sub-partition for the instructions the
// attention softmax is made of.  One CTA on one SM, W warps per sub-partition, ILP 8 independent chains.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes pipes.cu ; run: ./pipes
#include <cstdio>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#define ITER 4096
#define ILP 8

template <int OP>
__global__ void k(float* out, long long* cyc, float seed) {
    float x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; ++i) x[i] = seed + i * 0.001f + threadIdx.x * 1e-6f;
    unsigned acc = 0;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 1) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(x[i]) : "f"(seed));
            if (OP == 2) asm volatile("max.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(seed));
            if (OP == 3) asm volatile("add.f32 %0, %0, %1;" : "+f"(x[i]) : "f"(seed));
            if (OP == 4) { unsigned r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(x[i]), "f"(seed)); acc ^= r; }
            if (OP == 5) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[i]));
            if (OP == 6) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[i])); asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(x[(i + 1) % ILP]) : "f"(seed)); }
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < ILP; ++i) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + acc;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

template <int OP>
void run(const char* name) {
    float* out; long long* cyc;
    cudaMalloc(&out, 1024 * 4); cudaMalloc(&cyc, 8);
    for (int warps : {4, 8, 16, 32}) {
        k<OP><<<1, warps * 32>>>(out, cyc, 0.5f);
        k<OP><<<1, warps * 32>>>(out, cyc, 0.5f);
        cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
        double per = (double)c / (ITER * ILP) / (warps / 4.0);
        printf("%-28s warps/SM %2d: %.2f cycles per warp-instruction per sub-partition -> %.1f lanes/clk/SM\n", name, warps, per, 32.0 * 4 / per);
    }
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("ex2.approx (MUFU.EX2)");
    run<5>("rcp.approx (MUFU.RCP)");
    run<1>("fma.rn.f32 (FFMA)");
    run<2>("max.f32 (FMNMX)");
    run<3>("add.f32 (FADD)");
    run<4>("cvt.rn.bf16x2.f32 (F2FP)");
    run<6>("ex2 + fma interleaved (pair)");
    cudaError_t e = cudaGetLastError();
    printf("status: %s\n", cudaGetErrorString(e));
    return 0;
}
