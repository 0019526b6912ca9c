// which pipe runs half2 min/max, and does it overlap with the integer min/max pipe?  (cycles per SMSP for 4 warps)
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#define ITER 512
template <int MODE>
__global__ void k(unsigned* out, unsigned seed, long long* cyc)
{
    unsigned a[8], b[8];
    __half2 h[8], g[8];
    for (int i = 0; i < 8; i++) {
        a[i] = seed * (threadIdx.x + 1) + i * 77; b[i] = a[i] ^ 0x5a5a5a5a;
        h[i] = __halves2half2(__ushort_as_half(0x6400 | ((a[i] >> 3) & 0xFF)), __ushort_as_half(0x6400 | ((a[i] >> 13) & 0xFF)));
        g[i] = __halves2half2(__ushort_as_half(0x6400 | ((b[i] >> 3) & 0xFF)), __ushort_as_half(0x6400 | ((b[i] >> 13) & 0xFF)));
    }
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0 || MODE == 2) a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]);
            if (MODE == 1 || MODE == 2) h[i] = __hmin2(h[i], g[(i + 1) & 7]);
            if (MODE == 3) { h[i] = __hmin2(h[i], g[(i + 1) & 7]); g[i] = __hmax2(g[i], h[(i + 3) & 7]); }
            if (MODE == 4) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); b[i] = __vimax3_u16x2(b[i], a[(i + 2) & 7], a[(i + 5) & 7]); }
            if (MODE == 5) { a[i] = __vimin3_u16x2(a[i], b[(i + 1) & 7], b[(i + 3) & 7]); h[i] = __hmin2(h[i], g[(i + 1) & 7]); g[i] = __hmax2(g[i], h[(i + 3) & 7]); }
        }
    }
    long long t1 = clock64();
    unsigned r = 0;
    for (int i = 0; i < 8; i++) r ^= a[i] ^ b[i] ^ (unsigned)__half2float(__low2half(h[i])) ^ (unsigned)__half2float(__high2half(g[i]));
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE> void run(const char* name, int nops, unsigned* out, long long* cyc)
{
    k<MODE><<<1, 512>>>(out, 12345, cyc); cudaDeviceSynchronize();
    k<MODE><<<1, 512>>>(out, 12345, cyc); cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-44s cycles %8lld  -> %.2f cycles per warp-instruction per SMSP (%d instr / iteration-slot)\n", name, h, (double)h / (4.0 * ITER * 8 * nops), nops);
}
int main()
{
    unsigned* out; long long* cyc; cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
    run<0>("VIMNMX3.U16x2", 1, out, cyc);
    run<1>("HMNMX2 (hmin2)", 1, out, cyc);
    run<2>("VIMNMX3.U16x2 + HMNMX2 interleaved", 2, out, cyc);
    run<3>("HMNMX2 x2", 2, out, cyc);
    run<4>("VIMNMX3 x2", 2, out, cyc);
    run<5>("VIMNMX3 + HMNMX2 x2", 3, out, cyc);
    return 0;
}
