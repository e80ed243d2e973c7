// ubench_icache.cu — how fast does straight-line ALU code run as a function of the loop body size?
// One loop whose body is N LOP3 (8 independent chains), N = 256 .. 32768 instructions (4 KB .. 512 KB of SASS),
// run with 1 or 2 warps per SM sub-partition.  Prints warp-instructions per clock per SM.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_icache ubench_icache.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int N>
__device__ __forceinline__ void body(unsigned (&a)[8], unsigned m)
{
#pragma unroll
    for (int i = 0; i < N / 8; i++) {
#pragma unroll
        for (int c = 0; c < 8; c++)
            asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a[c]) : "r"(a[(c + 1) & 7]), "r"(m));
    }
}

template <int N>
__global__ void k(unsigned *out, unsigned seed, int iters)
{
    unsigned a[8];
#pragma unroll
    for (int c = 0; c < 8; c++)
        a[c] = threadIdx.x * (2 * c + 3) + seed;
    unsigned m = seed | 0x0f0f0f0f;
#pragma unroll 1
    for (int it = 0; it < iters; it++)
        body<N>(a, m);
    unsigned r = 0;
#pragma unroll
    for (int c = 0; c < 8; c++)
        r ^= a[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int N>
void run(unsigned *d, int sms, double ghz)
{
    for (int wps = 1; wps <= 2; wps++) { // warps per sub-partition
        const int threads = 128 * wps;
        const int iters = (1 << 22) / N;
        k<N><<<sms, threads>>>(d, 1, 2);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0);
        cudaEventCreate(&e1);
        cudaEventRecord(e0);
        k<N><<<sms, threads>>>(d, 1, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        const double winstr = (double)iters * N * (threads / 32);
        printf("body %6d instr (%4d KB)  %d warp/SMSP  %.3f ms  %.2f warp-instr/clk/SM\n", N, N * 16 / 1024, wps, ms,
               winstr / (ms * 1e-3) / (ghz * 1e9));
    }
}

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int khz;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double ghz = khz * 1e-6;
    unsigned *d;
    cudaMalloc(&d, 1 << 24);
    run<256>(d, p.multiProcessorCount, ghz);
    run<512>(d, p.multiProcessorCount, ghz);
    run<1024>(d, p.multiProcessorCount, ghz);
    run<2048>(d, p.multiProcessorCount, ghz);
    run<4096>(d, p.multiProcessorCount, ghz);
    run<6144>(d, p.multiProcessorCount, ghz);
    run<8192>(d, p.multiProcessorCount, ghz);
    run<12288>(d, p.multiProcessorCount, ghz);
    run<16384>(d, p.multiProcessorCount, ghz);
    run<32768>(d, p.multiProcessorCount, ghz);
    return 0;
}
