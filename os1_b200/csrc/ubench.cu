// Issue-rate micro-benchmarks behind the Hamming roofline ("measure popc/clk/SM on the box with a
// micro-kernel; do not assume").  Built into its own small library (os1_b200/libos1_ubench.so) that only bench.py loads:
// measurement aid, not part of the product ABI.
//
//   popc   : independent POPC chains                         -> POPC thread-ops per clock per SM
//   lop3   : independent three-input logic chains            -> LOP3 thread-ops per clock per SM
//   dist   : os1_dist256 of a register row against a resident shared-memory tile (8 XOR + 4 carry-save adders + 4 POPC + adds)
//   scan   : dist + the key / best / second update — the matchers' inner loop exactly, with everything around it removed
//            (no tile refills, no barriers, no tails, no merge, 64 resident warps per SM)
// The `scan` rate is the ceiling the matcher kernels are held to.
#include <cuda_runtime.h>
#include <stdint.h>
#include "hamming_dist.cuh"

namespace {

template <int MODE>
__global__ void __launch_bounds__(256) ubench_kernel(unsigned seed, int iters, unsigned* sink, long long* cycles)
{
    const unsigned t = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned r[8];
#pragma unroll
    for (int i = 0; i < 8; i++) r[i] = seed * (2654435761u + i) + t * (40503u + 2 * i);
    const long long c0 = clock64();
    if (MODE == 0) {            // 8 independent popc per iteration
        unsigned acc[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
#pragma unroll 1
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) { acc[i] += __popc(r[i]); r[i] = acc[i] ^ r[i]; }   // (the xor keeps the popc input changing)
        }
#pragma unroll
        for (int i = 0; i < 8; i++) r[0] += acc[i];
    } else if (MODE == 1) {     // 16 independent-ish lop3 per iteration
#pragma unroll 1
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                r[i] = (r[i] & r[(i + 1) & 7]) | (r[(i + 3) & 7] & (r[i] ^ r[(i + 1) & 7]));
                r[i] = r[i] ^ r[(i + 2) & 7] ^ r[(i + 5) & 7];
            }
        }
    } else {
        // the matchers' inner loop as it is (hamming.cu): row of A in registers, candidates read from a shared-memory tile
        // as broadcast 16-byte loads — but the tile is never refilled, there are no barriers, no tails and no merge
        __shared__ uint4 tile[128];
        for (int i = threadIdx.x; i < 128; i += blockDim.x) tile[i] = make_uint4(r[0] * (i + 1), r[1] ^ (i * 0x9E3779B9u), r[2] + i, r[3] * (2 * i + 1));
        __syncthreads();
        const uint4 a0 = make_uint4(r[0], r[1], r[2], r[3]), a1 = make_uint4(r[4], r[5], r[6], r[7]);
        unsigned best = 0xFFFFFFFFu, second = 0xFFFFFFFFu, sum = 0;
#pragma unroll 1
        for (int it = 0; it < iters; it += 16) {
#pragma unroll 4
            for (int j = 0; j < 64; j++) {
                const unsigned d = MODE == 4 ? (unsigned)os1_dist256<1>(a0, a1, tile[2 * j], tile[2 * j + 1])
                                   : MODE == 5 ? (unsigned)os1_dist256<2>(a0, a1, tile[2 * j], tile[2 * j + 1])
                                               : (unsigned)os1_dist256<OS1_DIST_VARIANT>(a0, a1, tile[2 * j], tile[2 * j + 1]);
                if (MODE == 2) sum += d;
                else {
                    const unsigned k = MODE == 3 ? os1_dist256_key(a0, a1, tile[2 * j], tile[2 * j + 1], (unsigned)(4 * it + j)) : ((d << 20) | (unsigned)(4 * it + j));
                    second = min(second, max(k, best));
                    best = min(best, k);
                }
            }
        }
        r[0] = sum + best + second;
    }
    const long long c1 = clock64();
    if (r[0] == 0x12345678u) sink[0] = r[0];
    if (threadIdx.x == 0) cycles[blockIdx.x] = c1 - c0;
}

template <int MODE>
int run(int device, int iters, double* ops_per_clk_per_sm, double* ops_per_s, double* eff_mhz = nullptr)
{
    int sms = 0;
    if (cudaSetDevice(device) != cudaSuccess) return -1;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    int resident = 0, clockKHz = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, ubench_kernel<MODE>, 256, 0);
    cudaDeviceGetAttribute(&clockKHz, cudaDevAttrClockRate, device);
    if (resident < 1) resident = 1;
    const int blocks = sms * resident;              // exactly one full wave: every SM holds `resident` blocks from start to end
    unsigned* sink = nullptr; long long* cyc = nullptr;
    if (cudaMalloc(&sink, 4) != cudaSuccess || cudaMalloc(&cyc, blocks * sizeof(long long)) != cudaSuccess) return -1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    // warm-up long enough for the clocks to settle (an idle GPU starts far below its running clock), then the timed launch
    {
        cudaEvent_t w0, w1;
        cudaEventCreate(&w0); cudaEventCreate(&w1);
        cudaEventRecord(w0);
        float wms = 0;
        for (int k = 0; k < 400 && wms < 250.f; k++) {
            ubench_kernel<MODE><<<blocks, 256>>>(1u, iters, sink, cyc);
            cudaEventRecord(w1);
            cudaEventSynchronize(w1);
            cudaEventElapsedTime(&wms, w0, w1);
        }
        cudaEventDestroy(w0); cudaEventDestroy(w1);
    }
    cudaEventRecord(e0);
    ubench_kernel<MODE><<<blocks, 256>>>(2u, iters, sink, cyc);
    cudaEventRecord(e1);
    if (cudaDeviceSynchronize() != cudaSuccess) return -1;
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    long long* h = new long long[blocks];
    cudaMemcpy(h, cyc, blocks * sizeof(long long), cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < blocks; i++) mean += (double)h[i];
    mean /= blocks;
    delete[] h;
    const double per_iter = MODE == 0 ? 8.0 : (MODE == 1 ? 16.0 : 4.0);      // popc / lop3 / distances per thread per iteration
    const double thread_ops = (double)blocks * 256.0 * iters * per_iter;
    // every SM runs its resident blocks side by side for ~`mean` SM cycles (clock64 counts at the clock the SM really runs at)
    *ops_per_clk_per_sm = (double)resident * 256.0 * iters * per_iter / mean;
    (void)clockKHz;
    *ops_per_s = thread_ops / (ms * 1e-3);
    if (eff_mhz) *eff_mhz = mean / (ms * 1e-3) / 1e6;      // SM cycles a block counted over the kernel's wall time: the clock the SMs ran at
    cudaFree(sink); cudaFree(cyc); cudaEventDestroy(e0); cudaEventDestroy(e1);
    return 0;
}

} // namespace

// out[0..1] popc, out[2..3] lop3, out[4..5] dist256, out[6..7] dist256 + best/second update: (per clock per SM, per second);
// out[9], out[10] = scan pairs per second with distance variants 1 and 2;
// out[8] = SM clock (MHz) while the last kernel ran (these loops keep every integer pipe busy: the chip may run them below its
// nominal clock, which is why the per-second and per-clock figures are both given)
extern "C" int os1_ubench_issue_rates(int device, double* out8)
{
    int prev = -1;
    cudaGetDevice(&prev);
    int rc = 0;
    rc |= run<0>(device, 4096, out8 + 0, out8 + 1);
    rc |= run<1>(device, 4096, out8 + 2, out8 + 3);
    rc |= run<2>(device, 2048, out8 + 4, out8 + 5);
    rc |= run<3>(device, 2048, out8 + 6, out8 + 7, out8 + 8);
    // the same scan with fewer carry-save adders and more POPCs (hamming_dist.cuh variants 1 and 2), for the record of why
    // the shipped variant was chosen
    double t = 0;
    rc |= run<4>(device, 2048, &t, out8 + 9);
    rc |= run<5>(device, 2048, &t, out8 + 10);
    if (prev >= 0) cudaSetDevice(prev);
    return rc;
}
