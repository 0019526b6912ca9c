// instruction-throughput micro-benchmark for the integer min/max family on sm_100a (cycles per warp-instruction per SM sub-partition)
#include <cstdio>
#include <cuda_runtime.h>
#define ITER 256
template <int OP>
__global__ void k(unsigned* out, unsigned seed, long long* cyc)
{
    unsigned a[8];
    for (int i = 0; i < 8; i++) a[i] = seed * (threadIdx.x + 1) + i * 77;
    unsigned b = seed ^ 0x1234567, c = seed + threadIdx.x;
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITER; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if (OP == 0) a[i] = __vimin3_u16x2(a[i], b, c);
            if (OP == 1) a[i] = __vimin3_s32((int)a[i], (int)b, (int)c);
            if (OP == 2) a[i] = __vminu2(a[i], b);
            if (OP == 3) a[i] = min(a[i], b);
            if (OP == 4) a[i] = __byte_perm(a[i], b, 0x5140);
            if (OP == 5) a[i] = (a[i] & b) | c;
            if (OP == 6) a[i] = a[i] + b;
            if (OP == 7) a[i] = __vimax3_u16x2(a[i], b, c);
            if (OP == 8) a[i] = a[i] * b + c;
            if (OP == 9) a[i] = __vimin3_u16x2(a[i], a[(i + 1) & 7], a[(i + 3) & 7]);
            b += 1;   // keep operands varying a little (1 extra IADD per op, measured separately by OP 6)
        }
    }
    long long t1 = clock64();
    unsigned r = 0;
    for (int i = 0; i < 8; i++) r ^= a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r + b;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int OP> void run(const char* name, unsigned* out, long long* cyc)
{
    // 4 warps per SMSP-quadrant... use 1 block of 512 threads (16 warps = 4 per SMSP) on one SM
    k<OP><<<1, 512>>>(out, 12345, cyc); cudaDeviceSynchronize();
    k<OP><<<1, 512>>>(out, 12345, cyc); cudaDeviceSynchronize();
    long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    // per SMSP: 4 warps x ITER x 8 ops (+8 IADD)
    printf("%-28s cycles %8lld  -> %.2f cycles per (op + 1 IADD) per SMSP-warp-instr-pair\n", name, h, (double)h / (4.0 * ITER * 8));
}
int main()
{
    unsigned* out; long long* cyc; cudaMalloc(&out, 1 << 20); cudaMalloc(&cyc, 8);
    run<6>("IADD only (x2)", out, cyc);
    run<0>("VIMNMX3.U16x2 min", out, cyc);
    run<7>("VIMNMX3.U16x2 max", out, cyc);
    run<1>("VIMNMX3 s32", out, cyc);
    run<2>("VIMNMX.U16x2 (vminu2)", out, cyc);
    run<3>("VIMNMX u32", out, cyc);
    run<4>("PRMT", out, cyc);
    run<5>("LOP3", out, cyc);
    run<8>("IMAD", out, cyc);
    run<9>("VIMNMX3.U16x2 reg-reg-reg", out, cyc);
    return 0;
}
