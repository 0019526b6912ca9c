// Which pipe do the packed min / max flavours run on?  Time N independent chains of one flavour alone and two flavours side by
// side: when the pair costs what the slower one costs alone, they sit on different pipes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scripts/scratch/pipes scripts/scratch/pipes.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>

__device__ __forceinline__ unsigned umin2(unsigned a, unsigned b) { unsigned d; asm("min.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ unsigned umax2(unsigned a, unsigned b) { unsigned d; asm("max.u16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ unsigned hmin2(unsigned a, unsigned b) { unsigned d; asm("min.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ unsigned hmax2(unsigned a, unsigned b) { unsigned d; asm("max.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ unsigned hrelu(unsigned a, unsigned b, unsigned c) { unsigned d; asm("fma.rn.relu.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ unsigned hadd2u(unsigned a, unsigned b) { unsigned d; asm("add.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ float fmn(float a, float b) { return fminf(a, b); }

template <int MODE>
__global__ void __launch_bounds__(256) k(unsigned seed, int iters, unsigned* sink)
{
    const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned a[8], b[8], h[8], g[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        a[i] = seed * (2654435761u + i) + t * (40503u + 2 * i); b[i] = a[i] * 3u + 7u;
        h[i] = 0x64006400u | (a[i] & 0x00ff00ffu); g[i] = 0x64006400u | (b[i] & 0x00ff00ffu);
    }
    const unsigned one = 0x3c003c00u;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0 || MODE == 2 || MODE == 4) { a[i] = umin2(a[i], b[(i + 1) & 7]); b[i] = umax2(b[i], a[(i + 3) & 7]); }
            if (MODE == 1 || MODE == 2) { h[i] = hmin2(h[i], g[(i + 1) & 7]); g[i] = hmax2(g[i], h[(i + 3) & 7]); }
            if (MODE == 3 || MODE == 4) { h[i] = hrelu(h[i], one, g[(i + 1) & 7]); g[i] = hadd2u(g[i], h[(i + 3) & 7]); }
            if (MODE == 5) { a[i] = umin2(umin2(a[i], b[(i + 1) & 7]), b[(i + 2) & 7]); b[i] = umax2(umax2(b[i], a[(i + 3) & 7]), a[(i + 5) & 7]); }
            if (MODE == 7 || MODE == 11) { a[i] = a[i] * b[(i + 1) & 7] + b[(i + 2) & 7]; b[i] = b[i] * a[(i + 3) & 7] + a[(i + 5) & 7]; }
            if (MODE == 8 || MODE == 10) { a[i] = __umulhi(a[i], b[(i + 1) & 7]); b[i] = __umulhi(b[i], a[(i + 3) & 7]); }
            if (MODE == 9) { a[i] = __dp2a_lo(a[i], b[(i + 1) & 7], b[(i + 2) & 7]); b[i] = __dp2a_hi(b[i], a[(i + 3) & 7], a[(i + 5) & 7]); }
            if (MODE == 10) { h[i] = umin2(h[i], g[(i + 1) & 7]); g[i] = umax2(g[i], h[(i + 3) & 7]); }
            if (MODE == 11) { h[i] = __float_as_uint(__fmaf_rn(__uint_as_float(h[i]), 1.0001f, __uint_as_float(g[(i + 1) & 7]))); g[i] = __float_as_uint(__fmaf_rn(__uint_as_float(g[i]), 0.9999f, __uint_as_float(h[(i + 3) & 7]))); }
            if (MODE == 12) { a[i] = __byte_perm(a[i], b[(i + 1) & 7], b[(i + 2) & 7]); b[i] = __byte_perm(b[i], a[(i + 3) & 7], a[(i + 5) & 7]); }
            if (MODE == 15) { a[i] = __byte_perm(a[i], b[(i + 1) & 7], 0x5140); b[i] = __byte_perm(b[i], a[(i + 3) & 7], 0x7250); }
            if (MODE == 16) { a[i] = __byte_perm(a[i], b[(i + 1) & 7], 0x5140); b[i] = __byte_perm(b[i], a[(i + 3) & 7], 0x7250); h[i] = umin2(h[i], g[(i + 1) & 7]); g[i] = umax2(g[i], h[(i + 3) & 7]); }
            if (MODE == 17) { a[i] = (a[i] & b[(i + 1) & 7]) | (b[(i + 2) & 7] & ~a[i]); b[i] = b[i] ^ a[(i + 3) & 7] ^ a[(i + 5) & 7]; }
            if (MODE == 18) { a[i] = a[i] + b[(i + 1) & 7] + b[(i + 2) & 7]; b[i] = b[i] + a[(i + 3) & 7] - a[(i + 5) & 7]; }
            if (MODE == 13) { a[i] = __funnelshift_r(a[i], b[(i + 1) & 7], 7); b[i] = __funnelshift_r(b[i], a[(i + 3) & 7], 9); }
            if (MODE == 14) { h[i] = __float_as_uint(__fmaf_rn(__uint_as_float(h[i]), 1.0001f, __uint_as_float(g[(i + 1) & 7]))); g[i] = __float_as_uint(__fmaf_rn(__uint_as_float(g[i]), 0.9999f, __uint_as_float(h[(i + 3) & 7]))); }
            if (MODE == 6) { a[i] = __float_as_uint(fmn(__uint_as_float(a[i]), __uint_as_float(b[(i + 1) & 7]))); b[i] = __float_as_uint(fmaxf(__uint_as_float(b[i]), __uint_as_float(a[(i + 3) & 7]))); }
        }
    }
    unsigned s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i] + b[i] + h[i] + g[i];
    if (s == 0x12345678u) sink[0] = s;
}

template <int MODE> void run(const char* name, double opsPerIter)
{
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int res = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&res, k<MODE>, 256, 0);
    const int blocks = sms * res, iters = 8192;
    unsigned* sink; cudaMalloc(&sink, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 30; w++) k<MODE><<<blocks, 256>>>(1u, iters, sink);
    cudaEventRecord(e0);
    for (int w = 0; w < 5; w++) k<MODE><<<blocks, 256>>>(2u, iters, sink);
    cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
    const double inst = (double)blocks * 256 * iters * opsPerIter;
    printf("%-34s %8.3f ms  %7.2f T thread-instr/s  (%d blocks/SM)  = %.1f lanes/clk/SM at 1.965 GHz\n", name, ms, inst / (ms * 1e-3) / 1e12, res,
           inst / (ms * 1e-3) / sms / 1.965e9);
    cudaFree(sink);
}

int main()
{
    run<0>("min/max.u16x2 (VIMNMX.U16x2)", 16);
    run<1>("min/max.f16x2 (HMNMX2)", 16);
    run<2>("u16x2 + f16x2 side by side", 32);
    run<3>("fma.relu.f16x2 + add.f16x2", 16);
    run<4>("u16x2 + (fma.relu, add).f16x2", 32);
    run<5>("3-input u16x2 (VIMNMX3)", 16);
    run<6>("fminf/fmaxf (FMNMX)", 16);
    run<7>("IMAD (32-bit mad.lo)", 16);
    run<8>("IMAD.HI (umulhi)", 16);
    run<9>("IDP.2A (dp2a)", 16);
    run<10>("IMAD.HI + VIMNMX.U16x2", 32);
    run<14>("FFMA", 16);
    run<11>("IMAD + FFMA side by side", 32);
    run<12>("PRMT", 16);
    run<13>("SHF (funnel shift)", 16);
    run<15>("PRMT, immediate selector", 16);
    run<16>("PRMT imm + VIMNMX.U16x2", 32);
    run<17>("LOP3", 16);
    run<18>("IADD3", 16);
    return 0;
}
