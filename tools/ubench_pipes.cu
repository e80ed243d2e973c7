// ubench_pipes.cu — throughput of the integer pipes the bitsliced kernels live on (B200, sm_100a).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench_pipes ubench_pipes.cu
// Prints warp-instructions per clock per SM for LOP3, IMAD, LOP3+IMAD mix, SHF, PRMT, SHFL, LOP3+SHFL mix, IMAD.HI
// (a right shift on the FMA pipe) and LOP3+IMAD.HI; with an argument, also writes them as JSON to that file
// (profiles/r02_ubench_pipes.json: the ALU-roofline denominator of bench.py).
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 4096

template <int MODE>
__global__ void k(unsigned *out, unsigned seed)
{
    unsigned a0 = threadIdx.x ^ seed, a1 = a0 * 3 + 1, a2 = a0 * 5 + 2, a3 = a0 * 7 + 3;
    unsigned a4 = a0 * 11 + 4, a5 = a0 * 13 + 5, a6 = a0 * 17 + 6, a7 = a0 * 19 + 7;
    unsigned m = seed | 0x0f0f0f0f, s = (seed & 3) + 1;
#pragma unroll 1
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
            if (MODE == 0) { // LOP3 (3-input xor), 8 independent chains
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a0) : "r"(a1), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(a2), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a2) : "r"(a3), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a3) : "r"(a4), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a4) : "r"(a5), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a5) : "r"(a6), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a6) : "r"(a7), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a7) : "r"(a0), "r"(m));
            } else if (MODE == 1) { // IMAD
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a0) : "r"(a1), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a1) : "r"(a2), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a2) : "r"(a3), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a3) : "r"(a4), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a4) : "r"(a5), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a5) : "r"(a6), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a6) : "r"(a7), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a7) : "r"(a0), "r"(m));
            } else if (MODE == 2) { // 4 LOP3 + 4 IMAD interleaved
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a0) : "r"(a1), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a4) : "r"(a5), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(a2), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a5) : "r"(a6), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a2) : "r"(a3), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a6) : "r"(a7), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a3) : "r"(a0), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a7) : "r"(a4), "r"(m));
            } else if (MODE == 3) { // SHF (funnel shift)
                asm volatile("shf.l.wrap.b32 %0,%0,%1,%2;" : "+r"(a0) : "r"(a1), "r"(s));
                asm volatile("shf.l.wrap.b32 %0,%0,%1,%2;" : "+r"(a1) : "r"(a2), "r"(s));
                asm volatile("shf.l.wrap.b32 %0,%0,%1,%2;" : "+r"(a2) : "r"(a3), "r"(s));
                asm volatile("shf.l.wrap.b32 %0,%0,%1,%2;" : "+r"(a3) : "r"(a4), "r"(s));
                asm volatile("shf.l.wrap.b32 %0,%0,%1,%2;" : "+r"(a4) : "r"(a5), "r"(s));
                asm volatile("shf.l.wrap.b32 %0,%0,%1,%2;" : "+r"(a5) : "r"(a6), "r"(s));
                asm volatile("shf.l.wrap.b32 %0,%0,%1,%2;" : "+r"(a6) : "r"(a7), "r"(s));
                asm volatile("shf.l.wrap.b32 %0,%0,%1,%2;" : "+r"(a7) : "r"(a0), "r"(s));
            } else if (MODE == 4) { // PRMT
                asm volatile("prmt.b32 %0,%0,%1,%2;" : "+r"(a0) : "r"(a1), "r"(m));
                asm volatile("prmt.b32 %0,%0,%1,%2;" : "+r"(a1) : "r"(a2), "r"(m));
                asm volatile("prmt.b32 %0,%0,%1,%2;" : "+r"(a2) : "r"(a3), "r"(m));
                asm volatile("prmt.b32 %0,%0,%1,%2;" : "+r"(a3) : "r"(a4), "r"(m));
                asm volatile("prmt.b32 %0,%0,%1,%2;" : "+r"(a4) : "r"(a5), "r"(m));
                asm volatile("prmt.b32 %0,%0,%1,%2;" : "+r"(a5) : "r"(a6), "r"(m));
                asm volatile("prmt.b32 %0,%0,%1,%2;" : "+r"(a6) : "r"(a7), "r"(m));
                asm volatile("prmt.b32 %0,%0,%1,%2;" : "+r"(a7) : "r"(a0), "r"(m));
            } else if (MODE == 5) { // SHFL
                a0 = __shfl_xor_sync(0xffffffffu, a0, 1); a1 = __shfl_xor_sync(0xffffffffu, a1, 1);
                a2 = __shfl_xor_sync(0xffffffffu, a2, 1); a3 = __shfl_xor_sync(0xffffffffu, a3, 1);
                a4 = __shfl_xor_sync(0xffffffffu, a4, 1); a5 = __shfl_xor_sync(0xffffffffu, a5, 1);
                a6 = __shfl_xor_sync(0xffffffffu, a6, 1); a7 = __shfl_xor_sync(0xffffffffu, a7, 1);
            } else if (MODE == 6) { // 7 LOP3 + 1 SHFL
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a0) : "r"(a1), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(a2), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a2) : "r"(a3), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a3) : "r"(a4), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a4) : "r"(a5), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a5) : "r"(a6), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a6) : "r"(a0), "r"(m));
                a7 = __shfl_xor_sync(0xffffffffu, a7, 1);
            } else if (MODE == 7) { // IADD3
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a0) : "r"(a1));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a1) : "r"(a2));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a2) : "r"(a3));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a3) : "r"(a4));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a4) : "r"(a5));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a5) : "r"(a6));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a6) : "r"(a7));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a7) : "r"(a0));
            } else if (MODE == 8) { // 4 LOP3 + 4 imul-by-16 (shift on the fma pipe?)
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a0) : "r"(a1), "r"(m));
                asm volatile("mul.lo.u32 %0,%0,%1;" : "+r"(a4) : "r"(s));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(a2), "r"(m));
                asm volatile("mul.lo.u32 %0,%0,%1;" : "+r"(a5) : "r"(s));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a2) : "r"(a3), "r"(m));
                asm volatile("mul.lo.u32 %0,%0,%1;" : "+r"(a6) : "r"(s));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a3) : "r"(a0), "r"(m));
                asm volatile("mul.lo.u32 %0,%0,%1;" : "+r"(a7) : "r"(s));
            } else if (MODE == 9) { // mul.hi (right shift by a constant on the fma pipe?)
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a0) : "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a1) : "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a2) : "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a3) : "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a4) : "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a5) : "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a6) : "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a7) : "r"(m));
            } else if (MODE == 10) { // 4 LOP3 + 4 mul.hi
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a0) : "r"(a1), "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a4) : "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(a2), "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a5) : "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a2) : "r"(a3), "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a6) : "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a3) : "r"(a0), "r"(m));
                asm volatile("mul.hi.u32 %0,%0,%1;" : "+r"(a7) : "r"(m));
            } else if (MODE == 11) { // 6 LOP3 + 2 IMAD (3:1)
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a0) : "r"(a1), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(a2), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a2) : "r"(a3), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a6) : "r"(a7), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a3) : "r"(a4), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a4) : "r"(a5), "r"(m));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a5) : "r"(a0), "r"(m));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a7) : "r"(a6), "r"(m));
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
}

// The bitsliced GF(2^8) product as the kernels issue it: acc[s + t] ^= a[s] & b[t], 64 LOP3 with three DIFFERENT
// register operands each (8 + 8 inputs, 15 accumulators), 16 products per loop trip with the inputs rotating.  The
// solver runs 8 warps per SM at 255 registers, the transform kernels 16 at 128: the rate is measured at those warp
// counts, not only at 32.
__global__ void kmac(unsigned *out, unsigned seed)
{
    unsigned a[8], b[8], acc[15];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        a[i] = threadIdx.x * (2 * i + 3) ^ seed;
        b[i] = threadIdx.x * (2 * i + 5) + seed;
    }
#pragma unroll
    for (int i = 0; i < 15; i++)
        acc[i] = i;
#pragma unroll 1
    for (int it = 0; it < ITERS / 16; it++) {
#pragma unroll
        for (int rep = 0; rep < 16; rep++) {
#pragma unroll
            for (int q = 0; q < 8; q++)
#pragma unroll
                for (int t = 0; t < 8; t++)
                    asm volatile("lop3.b32 %0,%1,%2,%0,0x6a;" : "+r"(acc[q + t]) : "r"(a[(q + rep) & 7]), "r"(b[t]));
        }
    }
    unsigned r = 0;
#pragma unroll
    for (int i = 0; i < 15; i++)
        r ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

// the same product through the eight held multiples m[s][t] of one factor: v[t] ^= x[s] & m[s][t] (64 + 8 + 8 registers)
__global__ void __launch_bounds__(256, 1) ktab(unsigned *out, unsigned seed)
{
    unsigned m[8][8], x[8], v[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        x[i] = threadIdx.x * (2 * i + 3) ^ seed;
        v[i] = i;
#pragma unroll
        for (int j = 0; j < 8; j++)
            m[i][j] = threadIdx.x * (8 * i + j + 7) + seed;
    }
#pragma unroll 1
    for (int it = 0; it < ITERS / 16; it++) {
#pragma unroll
        for (int rep = 0; rep < 16; rep++) {
#pragma unroll
            for (int q = 0; q < 8; q++)
#pragma unroll
                for (int t = 0; t < 8; t++)
                    asm volatile("lop3.b32 %0,%1,%2,%0,0x6a;" : "+r"(v[t]) : "r"(x[(q + rep) & 7]), "r"(m[q][t]));
        }
    }
    unsigned r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++)
        r ^= v[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

static FILE *g_json = nullptr;
static int g_first = 1;

template <int MODE>
void run(const char *name, int sms, double ghz_guess)
{
    unsigned *out;
    const int threads = 256, blocks = sms * 4; // 32 warps per SM
    cudaMalloc(&out, sizeof(unsigned) * threads * blocks);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    k<MODE><<<blocks, threads>>>(out, 1);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k<MODE><<<blocks, threads>>>(out, 2);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double winstr = (double)blocks * (threads / 32) * ITERS * 32.0; // warp instructions in the hot loop
    double per_s = winstr / (ms * 1e-3);
    printf("%-28s %8.3f ms  %8.2f Gwarp-instr/s  = %6.2f warp-instr/clk/SM @%.3f GHz (lanes/clk/SM %.1f)\n", name, ms,
           per_s / 1e9, per_s / sms / (ghz_guess * 1e9), ghz_guess, 32 * per_s / sms / (ghz_guess * 1e9));
    if (g_json) {
        fprintf(g_json, "%s\n  \"%s\": {\"warp_instr_per_clk_per_sm\": %.4f, \"lanes_per_clk_per_sm\": %.2f, \"gwarp_instr_per_s\": %.2f}",
                g_first ? "" : ",", name, per_s / sms / (ghz_guess * 1e9), 32 * per_s / sms / (ghz_guess * 1e9), per_s / 1e9);
        g_first = 0;
    }
    cudaFree(out);
}

template <class K>
void run_product(const char *name, K kernel, int warps_per_sm, int sms, double ghz_guess)
{
    unsigned *out;
    const int threads = 256, blocks = sms * warps_per_sm / 8;
    cudaMalloc(&out, sizeof(unsigned) * threads * blocks);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    kernel<<<blocks, threads>>>(out, 1);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    kernel<<<blocks, threads>>>(out, 2);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double winstr = (double)blocks * (threads / 32) * (ITERS / 16) * 16.0 * 64.0;
    double per_s = winstr / (ms * 1e-3);
    char full[96];
    snprintf(full, sizeof full, "%s, %d warps/SM", name, warps_per_sm);
    printf("%-40s %8.3f ms  %8.2f Gwarp-instr/s  = %6.2f warp-instr/clk/SM (lanes/clk/SM %.1f)\n", full, ms, per_s / 1e9,
           per_s / sms / (ghz_guess * 1e9), 32 * per_s / sms / (ghz_guess * 1e9));
    if (g_json) {
        fprintf(g_json, "%s\n  \"%s\": {\"warp_instr_per_clk_per_sm\": %.4f, \"lanes_per_clk_per_sm\": %.2f, \"gwarp_instr_per_s\": %.2f}",
                g_first ? "" : ",", full, per_s / sms / (ghz_guess * 1e9), 32 * per_s / sms / (ghz_guess * 1e9), per_s / 1e9);
        g_first = 0;
    }
    cudaFree(out);
}

int main(int argc, char **argv)
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double ghz = khz / 1e6;
    printf("%s, %d SMs, clock attr %.3f GHz\n", p.name, sms, ghz);
    if (argc > 1 && (g_json = fopen(argv[1], "w")))
        fprintf(g_json, "{\"device\": \"%s\", \"sms\": %d, \"clock_ghz\": %.3f, \"warps_per_sm\": 32, \"note\": \"8 independent chains per thread, 4096 x 4 x 8 instructions per thread; rates are per clock of the attribute clock\", \"pipes\": {", p.name, sms, ghz);
    run<0>("LOP3 xor3", sms, ghz);
    run<1>("IMAD", sms, ghz);
    run<2>("LOP3+IMAD 1:1", sms, ghz);
    run<3>("SHF", sms, ghz);
    run<4>("PRMT", sms, ghz);
    run<5>("SHFL.xor", sms, ghz);
    run<6>("LOP3x7 + SHFL", sms, ghz);
    run<7>("IADD", sms, ghz);
    run<8>("LOP3+IMUL 1:1", sms, ghz);
    run<9>("IMAD.HI (mul.hi)", sms, ghz);
    run<10>("LOP3+IMAD.HI 1:1", sms, ghz);
    run<11>("LOP3+IMAD 3:1", sms, ghz);
    for (int w : {8, 16, 32})
        run_product("GF product acc[s+t] ^= a[s] & b[t]", kmac, w, sms, ghz);
    for (int w : {8, 16})
        run_product("GF product on held multiples", ktab, w, sms, ghz);
    if (g_json) {
        fprintf(g_json, "\n}}\n");
        fclose(g_json);
    }
    return 0;
}
