// Can byte-window extraction be moved off the ALU pipe?  SHF vs IMAD.HI/IMAD throughput and their overlap with VIMNMX3.
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 512
__device__ __forceinline__ unsigned madhi(unsigned a, unsigned b, unsigned c) { unsigned r; asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
__device__ __forceinline__ unsigned mulhi(unsigned a, unsigned b) { unsigned r; asm("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ unsigned madlo(unsigned a, unsigned b, unsigned c) { unsigned r; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
template <int MODE>
__global__ void k(unsigned* out, unsigned seed, long long* cyc, unsigned k1, unsigned k2)
{
    unsigned a[8], b[8], c[8];
    for (int i = 0; i < 8; i++) { a[i] = seed * (threadIdx.x + 1) + i * 77; b[i] = a[i] ^ 0x5a5a5a5a; c[i] = a[i] * 3 + i; }
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]);
            if (MODE == 1) c[i] = __funnelshift_r(c[i], c[(i + 1) & 7], 8);
            if (MODE == 2) c[i] = madhi(c[i], k1, c[(i + 1) & 7]);
            if (MODE == 3) c[i] = madlo(c[i], k2, c[(i + 1) & 7]);
            if (MODE == 4) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); c[i] = madhi(c[i], k1, c[(i + 1) & 7]); }
            if (MODE == 5) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); c[i] = madhi(c[i], k1, c[(i + 1) & 7]); b[i] = madlo(b[i], k2, c[(i + 2) & 7]); }
            if (MODE == 6) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); c[i] = __funnelshift_r(c[i], c[(i + 1) & 7], 8); }
            if (MODE == 7) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); b[i] = __vimax3_u16x2(b[i], a[(i + 2) & 7], a[(i + 5) & 7]);
                             a[i] = __vimin3_u16x2(a[i], b[(i + 2) & 7], b[(i + 4) & 7]); c[i] = madhi(c[i], k1, c[(i + 1) & 7]); }
            if (MODE == 9) c[i] = mulhi(c[i], k1) ^ 0;
            if (MODE == 10) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); c[i] = madlo(c[i], k2, c[(i + 1) & 7]); }
            if (MODE == 11) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); c[i] = mulhi(c[(i + 1) & 7], k1); }
            if (MODE == 12) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); b[i] = __vimax3_u16x2(b[i], a[(i + 2) & 7], a[(i + 5) & 7]);
                              a[i] = __vimin3_u16x2(a[i], b[(i + 2) & 7], b[(i + 4) & 7]); c[i] = madlo(c[(i + 3) & 7], k2, mulhi(c[(i + 1) & 7], k1)); }
            if (MODE == 13) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); b[i] = __vimax3_u16x2(b[i], a[(i + 2) & 7], a[(i + 5) & 7]);
                              a[i] = __vimin3_u16x2(a[i], b[(i + 2) & 7], b[(i + 4) & 7]); b[i] = __vimax3_u16x2(b[i], a[(i + 1) & 7], a[(i + 6) & 7]);
                              a[i] = __vimin3_u16x2(a[i], b[(i + 3) & 7], b[(i + 5) & 7]); b[i] = __vimax3_u16x2(b[i], a[(i + 3) & 7], a[(i + 7) & 7]);
                              c[i] = madlo(c[(i + 3) & 7], k2, mulhi(c[(i + 1) & 7], k1)); }
            if (MODE == 8) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); c[i] = __byte_perm(c[i], c[(i + 1) & 7], 0x4321); }
        }
    }
    long long t1 = clock64();
    unsigned r = 0;
    for (int i = 0; i < 8; i++) r ^= a[i] ^ b[i] ^ c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE> void run(const char* name, int nops, unsigned* out, long long* cyc)
{
    k<MODE><<<1, 512>>>(out, 12345, cyc, 1u << 24, 1u << 8); cudaDeviceSynchronize();
    k<MODE><<<1, 512>>>(out, 12345, cyc, 1u << 24, 1u << 8); cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s cycles %8lld  -> %.2f cycles per SMSP per iteration-slot (%d instr)\n", name, h, (double)h / (4.0 * ITER * 8), nops);
}
int main()
{
    unsigned* out; long long* cyc; cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
    run<0>("VIMNMX3", 1, out, cyc);
    run<1>("SHF (funnelshift)", 1, out, cyc);
    run<2>("IMAD.HI", 1, out, cyc);
    run<3>("IMAD", 1, out, cyc);
    run<4>("VIMNMX3 + IMAD.HI", 2, out, cyc);
    run<5>("VIMNMX3 + IMAD.HI + IMAD", 3, out, cyc);
    run<6>("VIMNMX3 + SHF", 2, out, cyc);
    run<7>("VIMNMX3 x3 + IMAD.HI", 4, out, cyc);
    run<8>("VIMNMX3 + PRMT", 2, out, cyc);
    run<9>("IMAD.HI (no accumulate)", 1, out, cyc);
    run<10>("VIMNMX3 + IMAD", 2, out, cyc);
    run<11>("VIMNMX3 + IMAD.HI(RZ)", 2, out, cyc);
    run<12>("VIMNMX3 x3 + IMAD.HI + IMAD", 5, out, cyc);
    run<13>("VIMNMX3 x6 + IMAD.HI + IMAD", 8, out, cyc);
    return 0;
}
