#include <cstdio>
#include <cuda_runtime.h>
__global__ void spin(long long cycles, unsigned long long* out, int heavy)
{
    unsigned long long t0, t1; long long c0 = clock64();
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned x = threadIdx.x, y = blockIdx.x;
    while (clock64() - c0 < cycles) { if (heavy) { for (int i = 0; i < 64; i++) { x = __popc(x ^ y) + (x << 3); y = (y & x) | (y ^ 0x5555u); } } }
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = clock64() - c0; out[2] = x + y; }
}
int main()
{
    unsigned long long* d; cudaMalloc(&d, 64); unsigned long long h[3];
    for (int heavy = 0; heavy < 2; heavy++)
        for (int blocks : { 1, 148 * 8 }) {
            for (int rep = 0; rep < 3; rep++) {
                spin<<<blocks, 256>>>(200000000LL, d, heavy); cudaDeviceSynchronize();
                cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
            }
            printf("heavy=%d blocks=%d: %llu ns for %llu cycles -> %.1f MHz\n", heavy, blocks, h[0], h[1], h[1] * 1e3 / h[0]);
        }
    return 0;
}
