// MapPoint::ComputeDistinctiveDescriptors for many map points in one launch (SURVEY.md 8f.3;
// replaces src/MapPoint.cc:226-292: N x N DescriptorDistance matrix, median of every row, first row with the
// least median).
//
// One warp per map point.  For row i the lanes compute the distances to columns lane, lane+32, ... (at most eight per
// lane, kept in registers; up to 32 observations the descriptors themselves stay in registers and travel by shuffles) and the
// median — element int(0.5*(N-1)) of the sorted row, MapPoint.cc:279 — is found without sorting: distances are
// integers in [0, 256], so it is the smallest v with |{j : d_ij <= v}| >= k+1, located by a nine-step binary search
// whose counts are ballots.  Rows are visited in ascending order and a row replaces the best only when its median is
// strictly smaller, which is the reference's tie rule (:281).
#include "os1_internal.cuh"
#include "../../include/os1_b200.h"
#include <climits>
#include <vector>

#define DD_MAXN 256
#define DD_WARPS 4

namespace {

__device__ __forceinline__ void csa3(unsigned a, unsigned b, unsigned c, unsigned& sum, unsigned& carry)
{
    sum = a ^ b ^ c;
    carry = (a & b) | (c & (a ^ b));
}
// 256-bit Hamming distance (ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1605-1621) with four POPC
__device__ __forceinline__ int dist256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    unsigned sa, ca, sb, cb, sc, cc, sd, cd;
    csa3(x0, x1, x2, sa, ca);
    csa3(x3, x4, x5, sb, cb);
    csa3(x6, x7, sa, sc, cc);
    csa3(ca, cb, cc, sd, cd);
    return __popc(sb) + __popc(sc) + 2 * __popc(sd) + 4 * __popc(cd);
}

__device__ __forceinline__ uint4 shfl4(const uint4& v, int src)
{
    return make_uint4(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src), __shfl_sync(0xffffffffu, v.z, src),
                      __shfl_sync(0xffffffffu, v.w, src));
}

// smallest v in [0, 256] with |{d <= v}| >= need, d[0..T) per lane: nine ballot-counted steps
template <int TMAX>
__device__ __forceinline__ int median_of_row(const int (&d)[TMAX], int T, int need)
{
    int lo = 0, hi = 256;
#pragma unroll 1
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        int c = 0;
#pragma unroll
        for (int t = 0; t < TMAX; t++)
            if (t < T) c += __popc(__ballot_sync(0xffffffffu, d[t] <= mid));
        if (c >= need) hi = mid; else lo = mid + 1;
    }
    return lo;
}

__global__ void __launch_bounds__(DD_WARPS * 32) distinctive_kernel(const uint4* __restrict__ desc, const int* __restrict__ ptr, int npoints,
                                                                      int* __restrict__ bestIdx, int* __restrict__ bestMedian, int* status)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int m = blockIdx.x * DD_WARPS + warp;
    if (m >= npoints) return;
    const int p0 = ptr[m], N = ptr[m + 1] - p0;
    if (N <= 0 || N > DD_MAXN) {
        if (lane == 0) {
            bestIdx[m] = -1;
            if (bestMedian) bestMedian[m] = INT_MAX;
            if (N > DD_MAXN || N < 0) atomicOr(status, 1);
        }
        return;
    }
    const uint4* D = desc + 2 * (long long)p0;
    const int need = ((N - 1) >> 1) + 1;    // int(0.5*(N-1)) + 1 values must be <= the median
    int best = INT_MAX, bi = 0;
    if (N <= 32) {
        // the usual case (a map point is seen from a handful of keyframes): lane j keeps descriptor j in registers, row i
        // reaches the others by shuffles — no shared memory, no reloads
        uint4 c0 = make_uint4(0, 0, 0, 0), c1 = c0;
        if (lane < N) { c0 = __ldg(D + 2 * lane); c1 = __ldg(D + 2 * lane + 1); }
#pragma unroll 1
        for (int i = 0; i < N; i++) {
            const uint4 a0 = shfl4(c0, i), a1 = shfl4(c1, i);
            int d[1];
            d[0] = lane < N ? dist256(a0, a1, c0, c1) : 0x7FFF;
            const int med = median_of_row<1>(d, 1, need);
            if (med < best) { best = med; bi = i; }
        }
    } else {
        // up to 256 observations: columns lane, lane+32, ... straight from global memory (8 KB per point: L1 hits after the
        // first row), at most eight distances per lane in registers
        const int T = (N + 31) >> 5;
#pragma unroll 1
        for (int i = 0; i < N; i++) {
            const uint4 a0 = __ldg(D + 2 * i), a1 = __ldg(D + 2 * i + 1);
            int d[8];
#pragma unroll
            for (int t = 0; t < 8; t++) {
                const int j = lane + 32 * t;
                d[t] = (t < T && j < N) ? dist256(a0, a1, __ldg(D + 2 * j), __ldg(D + 2 * j + 1)) : 0x7FFF;
            }
            const int med = median_of_row<8>(d, T, need);
            if (med < best) { best = med; bi = i; }
        }
    }
    if (lane == 0) {
        bestIdx[m] = bi;
        if (bestMedian) bestMedian[m] = best;
    }
}

int set_err(const char* msg, int rc) { os1_internal_set_error(msg); return rc; }

// status words of the calls in flight (a call takes the next slot): static, so that the device-pointer variant allocates
// nothing — a stream-ordered allocation is returned to the driver at every synchronisation and would cost more than the kernel
__device__ int d_ddStatus[64];

} // namespace

extern "C" int os1_distinctive_descriptors(const uint8_t* desc, const int32_t* ptr, int npoints, int32_t* best_idx, int32_t* best_median,
                                           int on_device, int device, void* cuda_stream)
{
    if (npoints < 0 || (npoints > 0 && (!ptr || !best_idx))) return set_err("os1_distinctive_descriptors: bad argument", OS1_E_INVALID);
    if (npoints == 0) return OS1_OK;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    DeviceScope _ds(device); cudaError_t e = _ds.err;
    if (e != cudaSuccess) { cudaGetLastError(); return set_err("os1_distinctive_descriptors: cudaSetDevice failed (no CPU fallback)", OS1_E_CUDA); }
    const int grid = (npoints + DD_WARPS - 1) / DD_WARPS;
    int* dStatus = nullptr;
    {
        static unsigned next = 0;
        void* base = nullptr;
        e = cudaGetSymbolAddress(&base, d_ddStatus);
        if (e == cudaSuccess) dStatus = (int*)base + (__atomic_fetch_add(&next, 1u, __ATOMIC_RELAXED) & 63u);
    }
    if (e == cudaSuccess) e = cudaMemsetAsync(dStatus, 0, 4, s);
    int hostStatus = 0;
    if (on_device) {
        if (!desc) { return set_err("os1_distinctive_descriptors: NULL descriptors", OS1_E_INVALID); }
        if (e == cudaSuccess) {
            distinctive_kernel<<<grid, DD_WARPS * 32, 0, s>>>((const uint4*)desc, ptr, npoints, best_idx, best_median, dStatus);
            e = cudaGetLastError();
        }
        // the size limit can only be seen on the device here: one word comes back
        if (e == cudaSuccess) e = cudaMemcpyAsync(&hostStatus, dStatus, 4, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) { cudaGetLastError(); return set_err("os1_distinctive_descriptors: CUDA failure", OS1_E_CUDA); }
        if (hostStatus) return set_err("os1_distinctive_descriptors: a map point has more than 256 observations", OS1_E_UNSUPPORTED);
        return OS1_OK;
    }
    // host buffers: validate, stage, run, fetch
    if (ptr[0] < 0) { return set_err("os1_distinctive_descriptors: ptr is not a prefix sum", OS1_E_INVALID); }
    for (int m = 0; m < npoints; m++) {
        const int n = ptr[m + 1] - ptr[m];
        if (n < 0) { return set_err("os1_distinctive_descriptors: ptr is not monotone", OS1_E_INVALID); }
        if (n > DD_MAXN) { return set_err("os1_distinctive_descriptors: a map point has more than 256 observations", OS1_E_UNSUPPORTED); }
    }
    const long long total = ptr[npoints];
    if (total > 0 && !desc) { return set_err("os1_distinctive_descriptors: NULL descriptors", OS1_E_INVALID); }
    uint8_t* dDesc = nullptr; int *dPtr = nullptr, *dIdx = nullptr, *dMed = nullptr;
    const long long first = ptr[0];
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dDesc, (size_t)std::max(total, 1LL) * 32, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dPtr, (size_t)(npoints + 1) * 4, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dIdx, (size_t)npoints * 4, s);
    if (e == cudaSuccess) e = cudaMallocAsync((void**)&dMed, (size_t)npoints * 4, s);
    if (e == cudaSuccess && total > first) e = cudaMemcpyAsync(dDesc + first * 32, desc + first * 32, (size_t)(total - first) * 32, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dPtr, ptr, (size_t)(npoints + 1) * 4, cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) {
        distinctive_kernel<<<grid, DD_WARPS * 32, 0, s>>>((const uint4*)dDesc, dPtr, npoints, dIdx, dMed, dStatus);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(best_idx, dIdx, (size_t)npoints * 4, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess && best_median) e = cudaMemcpyAsync(best_median, dMed, (size_t)npoints * 4, cudaMemcpyDeviceToHost, s);
    void* frees[] = { dDesc, dPtr, dIdx, dMed };
    for (void* p : frees) if (p) cudaFreeAsync(p, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) { cudaGetLastError(); return set_err("os1_distinctive_descriptors: CUDA failure", OS1_E_CUDA); }
    return OS1_OK;
}
