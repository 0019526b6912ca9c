// K6 — batched Hamming consumer (replaces the scalar ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1605-1621,
// inside the scan loops of SearchForInitialization :431-452 and SearchByProjection :1360-1376, :79-107).
//
// 256-bit descriptors are 8 x u32; a distance is 8 XOR + 8 POPC (not a dense contraction: no tensor cores).
// The scan loops keep (best, second best, index of best) with strict '<' while walking candidates in list
// order, i.e. first-seen wins ties and "second" is the second-smallest VALUE of the multiset.  Both are
// order statistics of the keys  k = dist << 20 | position :   best = min k,  second = min over max(k, best-so-far),
// which turns the branchy scan into three integer min/max per pair and merges exactly across partitions:
//      best' = min(b1, b2),   second' = min(max(b1, b2), s1, s2).
//
// knn2 (all pairs, config 4 of BASELINE.json): thread = one row of A held in registers, B streamed through
// shared memory in 256-descriptor tiles read as broadcast uint4; B is split over blockIdx.y so that a 2000x2000
// problem fills the GPU; the per-split partials go through a scratch array and a small merge kernel.
// window (CSR candidate lists, what Frame::GetFeaturesInArea returns, src/Frame.cc:209-262): one warp per row.
#include "os1_internal.cuh"
#include "hamming_dist.cuh"
#include "../../include/os1_b200.h"
#include <climits>
#include <mutex>

#define HM_ROWS 64
#define HM_TILE 64
#define HM_POSBITS 20
#define HM_POSMASK ((1u << HM_POSBITS) - 1)
#define HM_EMPTY 0xFFFFFFFFu

namespace {

__device__ __forceinline__ int dist256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) { return os1_dist256(a0, a1, b0, b1); }

// merge of the per-split partials of one row (ascending index ranges: the earlier split wins ties)
__device__ __forceinline__ void merge_row(const uint2* partial, int row, int nA, int nSplit, int splitLen, int* bestIdx, int* bestDist, int* secondDist)
{
    int bd = INT_MAX, bi = -1, sd = INT_MAX;
    for (int s0 = 0; s0 < nSplit; s0 += 8) {
        // eight loads in flight at once (written by other CTAs of this launch: read through L2), then the ordered merge
        uint2 p[8];
#pragma unroll
        for (int k = 0; k < 8; k++) p[k] = (s0 + k < nSplit) ? __ldcg(partial + (long long)(s0 + k) * nA + row) : make_uint2(HM_EMPTY, HM_EMPTY);
#pragma unroll
        for (int k = 0; k < 8; k++) {
            if (p[k].x == HM_EMPTY) continue;
            const int d = (int)(p[k].x >> HM_POSBITS), i = (s0 + k) * splitLen + (int)(p[k].x & HM_POSMASK);
            const int d2 = p[k].y == HM_EMPTY ? INT_MAX : (int)(p[k].y >> HM_POSBITS);
            if (d < bd) { sd = min(bd, d2); bd = d; bi = i; }
            else sd = min(sd, d);
        }
    }
    bestIdx[row] = bi; bestDist[row] = bd; secondDist[row] = sd;
}

// partial[split * nA + row] = { bestKey, secondKey } with positions relative to the split start.  With `counters` the
// launch also merges: every CTA publishes its partials, then takes a ticket of its row block; the CTA that draws the last
// ticket sees the partials of all splits (fence + atomic, the "last block" pattern), merges them and puts the counter
// back to zero, so the counters are clean again when the launch ends — one launch instead of partial + merge.
__global__ void __launch_bounds__(HM_ROWS) hamming_knn2_partial(const uint4* __restrict__ A, int nA, const uint4* __restrict__ B, int nB,
                                                                 int splitLen, uint2* partial, unsigned* counters, int* bestIdx,
                                                                 int* bestDist, int* secondDist)
{
    __shared__ uint4 tile[HM_TILE * 2];
    __shared__ unsigned ticket;
    const int row = blockIdx.x * HM_ROWS + threadIdx.x;
    const int jb = blockIdx.y * splitLen, je = min(nB, jb + splitLen);
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (row < nA) { a0 = __ldg(A + 2 * row); a1 = __ldg(A + 2 * row + 1); }
    unsigned best = HM_EMPTY, second = HM_EMPTY;
    for (int t0 = jb; t0 < je; t0 += HM_TILE) {
        const int cnt = min(HM_TILE, je - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < cnt * 2; i += HM_ROWS) tile[i] = __ldg(B + 2 * (long long)t0 + i);
        __syncthreads();
        const unsigned base = (unsigned)(t0 - jb);
#pragma unroll 4
        for (int j = 0; j < cnt; j++) {
            const unsigned k = os1_dist256_key(a0, a1, tile[2 * j], tile[2 * j + 1], base + j);
            second = min(second, max(k, best));
            best = min(best, k);
        }
    }
    if (counters && gridDim.y == 1) {      // a single split: nothing to merge
        if (row < nA) {
            bestIdx[row] = best == HM_EMPTY ? -1 : (int)(best & HM_POSMASK);
            bestDist[row] = best == HM_EMPTY ? INT_MAX : (int)(best >> HM_POSBITS);
            secondDist[row] = second == HM_EMPTY ? INT_MAX : (int)(second >> HM_POSBITS);
        }
        return;
    }
    if (row < nA) partial[(long long)blockIdx.y * nA + row] = make_uint2(best, second);
    if (!counters) return;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) ticket = atomicAdd(counters + blockIdx.x, 1u);
    __syncthreads();
    if (ticket != gridDim.y - 1) return;
    __threadfence();
    if (row < nA) merge_row(partial, row, nA, gridDim.y, splitLen, bestIdx, bestDist, secondDist);
    if (threadIdx.x == 0) counters[blockIdx.x] = 0;
}

__global__ void hamming_knn2_merge(const uint2* __restrict__ partial, int nA, int nSplit, int splitLen, int* bestIdx, int* bestDist,
                                   int* secondDist)
{
    const int row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nA) return;
    merge_row(partial, row, nA, nSplit, splitLen, bestIdx, bestDist, secondDist);
}

// All pairs for SMALL problems (a single frame pair: BASELINE config 4, 2000 x 2000) — one launch, no partials in global memory.
// The tiled kernel above gives a thread a whole row and needs many CTAs per SM (hence a second grid dimension over B, partial
// results and a merge) before a 2000 x 2000 problem fills the GPU.  Here the columns of B are dealt out inside the CTA instead:
// a warp owns TWO rows of A, and its 16 lane pairs, together with one or three more warps of the same rows, split the columns 32
// or 64 ways (column j goes to slice j mod 32 or 64).  B is staged once per CTA into shared memory by ONE bulk copy (cp.async.bulk + mbarrier: the
// SM's copy engine streams the 64 KB of a 2000-column B while the warps fetch their rows); a warp's load then touches 16 columns,
// each address shared by the warp's two rows, in two conflict-free wavefronts — a quarter of what one row per warp costs the
// shared-memory pipe (that version was bound by it: l1tex 80 %, ncu).  The (best, second) keys carry global column positions, so the butterfly over a warp's slices and
// the two-way merge across the warp pair are exact in any order.  The launcher picks the CTA height so that a problem of this size
// is ONE CTA per SM: 2000 rows on 148 SMs -> 14 rows, 14 warps, 143 CTAs.
#define HL_MAX_COLS 3072            // columns of B per shared-memory pass (96 KB): larger B takes several passes
#define HL_SG_MAX 4                 // warps that can share a pair of rows
#define HL_STAGES 4                 // bulk copies (and barriers) a pass of B is brought in with
__global__ void __launch_bounds__(1024) hamming_knn2_lanes(const uint4* __restrict__ A, int nA, const uint4* __restrict__ B, int nB, int cols, int SG,
                                                           int* bestIdx, int* bestDist, int* secondDist)
{
    extern __shared__ __align__(128) uint4 tileB[];         // [cols][2], as in global memory
    __shared__ __align__(8) unsigned long long bar[HL_STAGES];
    __shared__ uint2 part[HL_SG_MAX - 1][32];               // (best, second) of the other warps of every row
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int RG = nwarps / SG, rg = warp % RG, sg = warp / RG;
    const int rowInCta = rg * 2 + (lane >> 4);
    const int row = blockIdx.x * (2 * RG) + rowInCta;
    const int slice = sg * 16 + (lane & 15);
    // Bank layout: a warp's load touches 16 columns = 512 consecutive bytes, of which every lane wants 16.  If all lanes took the
    // same half of their column, the 16 addresses would fall into 4 bank groups (4-way conflict); with slices 0-3 and 8-11 taking
    // the first half first and slices 4-7 and 12-15 the second half first, every bank group is hit exactly twice: two wavefronts,
    // the minimum for 256 bytes.  The row's halves are swapped for those lanes once, up front, so the loop does not care.
    const int h = (lane >> 2) & 1;
    const unsigned barAddr = (unsigned)__cvta_generic_to_shared(bar), tileAddr = (unsigned)__cvta_generic_to_shared(tileB);
    if (threadIdx.x == 0) {
        for (int i = 0; i < HL_STAGES; i++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(barAddr + 8 * i));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (row < nA) { a0 = __ldg(A + 2 * row + h); a1 = __ldg(A + 2 * row + (h ^ 1)); }
    unsigned best = HM_EMPTY, second = HM_EMPTY;
    unsigned parity = 0;
    for (int c0 = 0; c0 < nB; c0 += cols) {
        const int cnt = min(cols, nB - c0);
        // the pass arrives in HL_STAGES bulk copies (the SM's copy engine streams them: no registers, no per-thread loop), each with
        // its own barrier, so the distances of the first columns are under way while the later ones are still travelling
        const int per = ((cnt + HL_STAGES - 1) / HL_STAGES + 63) & ~63;
        if (c0) __syncthreads();                            // everybody is done with the previous pass's columns
        if (threadIdx.x == 0) {
            for (int i = 0; i < HL_STAGES; i++) {
                const int b0 = min(i * per, cnt), b1 = min(b0 + per, cnt);
                const unsigned bytes = (unsigned)(b1 - b0) * 32u;
                if (bytes) {
                    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(barAddr + 8 * i), "r"(bytes) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(tileAddr + 32u * b0),
                                 "l"(B + 2 * (long long)(c0 + b0)), "r"(bytes), "r"(barAddr + 8 * i)
                                 : "memory");
                } else {
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(barAddr + 8 * i) : "memory");
                }
            }
        }
#pragma unroll 1
        for (int i = 0; i < HL_STAGES; i++) {
            const int b0 = min(i * per, cnt), b1 = min(b0 + per, cnt);
            unsigned done;
            do {
                asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(done) : "r"(barAddr + 8 * i), "r"(parity) : "memory");
            } while (!done);
            if (row < nA) {
#pragma unroll 4
                for (int j = b0 + slice; j < b1; j += 16 * SG) {   // (b0 is a multiple of 64: column j still belongs to slice j mod 16*SG)
                    const unsigned k = os1_dist256_key(a0, a1, tileB[2 * j + h], tileB[2 * j + (h ^ 1)], (unsigned)(c0 + j));
                    second = min(second, max(k, best));
                    best = min(best, k);
                }
            }
        }
        parity ^= 1u;
    }
    // the 16 slices of this warp's two rows
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
        const unsigned ob = __shfl_xor_sync(0xffffffffu, best, o), os = __shfl_xor_sync(0xffffffffu, second, o);
        second = min(min(second, os), max(best, ob));
        best = min(best, ob);
    }
    if (sg > 0 && (lane & 15) == 0) part[sg - 1][rowInCta] = make_uint2(best, second);
    __syncthreads();
    if (sg == 0 && (lane & 15) == 0 && row < nA) {
        for (int g = 0; g < SG - 1; g++) {
            const uint2 p = part[g][rowInCta];
            second = min(min(second, p.y), max(best, p.x));
            best = min(best, p.x);
        }
        bestIdx[row] = best == HM_EMPTY ? -1 : (int)(best & HM_POSMASK);
        bestDist[row] = best == HM_EMPTY ? INT_MAX : (int)(best >> HM_POSBITS);
        secondDist[row] = second == HM_EMPTY ? INT_MAX : (int)(second >> HM_POSBITS);
    }
}

__global__ void __launch_bounds__(128) hamming_window_kernel(const uint4* __restrict__ A, int nA, const uint4* __restrict__ B,
                                                             const int* __restrict__ ptr, const int* __restrict__ cand, int* bestIdx,
                                                             int* bestDist, int* secondDist)
{
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= nA) return;
    const uint4 a0 = __ldg(A + 2 * row), a1 = __ldg(A + 2 * row + 1);
    const int p0 = ptr[row], p1 = ptr[row + 1];
    unsigned best = HM_EMPTY, second = HM_EMPTY;
    for (int k = p0 + lane; k < p1; k += 32) {
        const int j = __ldg(cand + k);
        const unsigned key = ((unsigned)dist256(a0, a1, __ldg(B + 2 * (long long)j), __ldg(B + 2 * (long long)j + 1)) << HM_POSBITS) |
                             (unsigned)(k - p0);
        second = min(second, max(key, best));
        best = min(best, key);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned ob = __shfl_xor_sync(0xffffffffu, best, o), os = __shfl_xor_sync(0xffffffffu, second, o);
        second = min(min(second, os), max(best, ob));
        best = min(best, ob);
    }
    if (lane == 0) {
        if (best == HM_EMPTY) { bestIdx[row] = -1; bestDist[row] = INT_MAX; secondDist[row] = INT_MAX; }
        else {
            bestIdx[row] = cand[p0 + (int)(best & HM_POSMASK)];
            bestDist[row] = (int)(best >> HM_POSBITS);
            secondDist[row] = second == HM_EMPTY ? INT_MAX : (int)(second >> HM_POSBITS);
        }
    }
}

// Many frame pairs in one launch (the cross-frame sweep of BASELINE.json configs[4]): blockIdx.y selects the pair
// (i, j) of a slab of per-frame descriptor sets; every thread scans the whole of frame j for its row of frame i, so
// no partials or merge are needed.  `accepted[pair]` counts rows passing ORBmatcher's test
// best <= th_low && best < nnratio * second  (src/ORBmatcher.cc:454-456).
__global__ void __launch_bounds__(HM_ROWS) hamming_pairs_kernel(const uint4* __restrict__ desc, const int* __restrict__ counts, int cap,
                                                                const int2* __restrict__ pairs, int* bestIdx, int* bestDist, int* secondDist,
                                                                int* accepted, float nnratio, int thLow)
{
    __shared__ uint4 tile[HM_TILE * 2];
    const int2 pr = pairs[blockIdx.y];
    const int nA = min(counts[pr.x], cap), nB = min(counts[pr.y], cap);      // a count above the slab's capacity is clamped, never followed out of bounds
    if ((int)(blockIdx.x * HM_ROWS) >= nA) return;
    const uint4* A = desc + (long long)pr.x * cap * 2;
    const uint4* B = desc + (long long)pr.y * cap * 2;
    const int row = blockIdx.x * HM_ROWS + threadIdx.x;
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (row < nA) { a0 = __ldg(A + 2 * row); a1 = __ldg(A + 2 * row + 1); }
    unsigned best = HM_EMPTY, second = HM_EMPTY;
    for (int t0 = 0; t0 < nB; t0 += HM_TILE) {
        const int cnt = min(HM_TILE, nB - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < cnt * 2; i += HM_ROWS) tile[i] = __ldg(B + 2 * (long long)t0 + i);
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < cnt; j++) {
            const unsigned k = os1_dist256_key(a0, a1, tile[2 * j], tile[2 * j + 1], (unsigned)(t0 + j));
            second = min(second, max(k, best));
            best = min(best, k);
        }
    }
    int ok = 0;
    if (row < nA) {
        const int bd = best == HM_EMPTY ? INT_MAX : (int)(best >> HM_POSBITS);
        const int sd = second == HM_EMPTY ? INT_MAX : (int)(second >> HM_POSBITS);
        const long long o = (long long)blockIdx.y * cap + row;
        if (bestIdx) bestIdx[o] = best == HM_EMPTY ? -1 : (int)(best & HM_POSMASK);
        if (bestDist) bestDist[o] = bd;
        if (secondDist) secondDist[o] = sd;
        ok = bd <= thLow && (float)bd < __fmul_rn((float)sd, nnratio);
    }
    if (accepted) {
        const unsigned m = __ballot_sync(0xffffffffu, ok);
        if ((threadIdx.x & 31) == 0 && m) atomicAdd(&accepted[blockIdx.y], __popc(m));
    }
}

// dist[k] = distance(A[row of k], B[cand[k]]) for every CSR entry; one warp per row
// (lists that end beyond `cap` entries were not written by their producer: they are left alone)
__global__ void __launch_bounds__(128) hamming_csr_kernel(const uint4* __restrict__ A, int nA, const uint4* __restrict__ B,
                                                          const int* __restrict__ ptr, const int* __restrict__ cand, uint16_t* __restrict__ dist, int cap)
{
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= nA || ptr[row + 1] > cap) return;
    const uint4 a0 = __ldg(A + 2 * row), a1 = __ldg(A + 2 * row + 1);
    const int p0 = ptr[row], p1 = ptr[row + 1];
    for (int k = p0 + lane; k < p1; k += 32) {
        const int j = __ldg(cand + k);
        dist[k] = (uint16_t)dist256(a0, a1, __ldg(B + 2 * (long long)j), __ldg(B + 2 * (long long)j + 1));
    }
}

} // namespace

cudaError_t os1_launch_hamming_pairs(const uint8_t* desc, const int* counts, int cap, const int* pairs, int npairs, int* bestIdx, int* bestDist,
                                     int* secondDist, int* accepted, float nnratio, int thLow, cudaStream_t s)
{
    if (npairs <= 0 || cap <= 0) return cudaSuccess;
    if (cap > (int)HM_POSMASK) return cudaErrorInvalidValue;
    if (accepted) { cudaError_t e = cudaMemsetAsync(accepted, 0, (size_t)npairs * 4, s); if (e != cudaSuccess) return e; }
    for (int p0 = 0; p0 < npairs; p0 += 65535) {
        const int n = min(65535, npairs - p0);
        dim3 grid((cap + HM_ROWS - 1) / HM_ROWS, n);
        const long long o = (long long)p0 * cap;
        hamming_pairs_kernel<<<grid, HM_ROWS, 0, s>>>((const uint4*)desc, counts, cap, (const int2*)pairs + p0, bestIdx ? bestIdx + o : nullptr,
                                                     bestDist ? bestDist + o : nullptr, secondDist ? secondDist + o : nullptr,
                                                     accepted ? accepted + p0 : nullptr, nnratio, thLow);
    }
    return cudaGetLastError();
}

cudaError_t os1_launch_hamming_csr(const uint8_t* A, int nA, const uint8_t* B, const int* ptr, const int* cand, uint16_t* dist, cudaStream_t s, int cap)
{
    if (nA <= 0) return cudaSuccess;
    hamming_csr_kernel<<<(nA + 3) / 4, 128, 0, s>>>((const uint4*)A, nA, (const uint4*)B, ptr, cand, dist, cap);
    return cudaGetLastError();
}

int os1_hamming_row_blocks(int nA) { return (nA + HM_ROWS - 1) / HM_ROWS; }

// number of B partitions used for an nA x nB problem on a GPU with `sms` SMs
int os1_hamming_splits(int nA, int nB, int sms)
{
    const int rowBlocks = (nA + HM_ROWS - 1) / HM_ROWS;
    int s = (8 * sms + rowBlocks - 1) / rowBlocks;            // aim at ~8 two-warp CTAs per SM: the scan is latency bound
    const int maxS = (nB + HM_TILE - 1) / HM_TILE;             // at least one tile per split
    if (s > maxS) s = maxS;
    if (s < 1) s = 1;
    return s;
}

// Small problems — fewer row blocks than it takes to fill the GPU even with every split the tiled kernel could use — go to the
// lane-split kernel: one CTA per SM, nothing to merge.  Returns false when the problem is not of that kind.
static bool launch_lanes(const uint8_t* A, int nA, const uint8_t* B, int nB, int* bestIdx, int* bestDist, int* secondDist, int sms, cudaStream_t s,
                         cudaError_t* err)
{
    if (nB < 1 || nB > (int)HM_POSMASK || nA > 16 * sms) return false;
    int rows = (nA + sms - 1) / sms;                     // rows per CTA: one CTA per SM where that is possible
    rows = (rows + 1) & ~1;
    rows = rows < 4 ? 4 : (rows > 16 ? 16 : rows);
    const int SG = 2;         // warps that share a pair of rows (4 was measured too: 9.0 instead of 8.9 us for 2000 x 2000)
    const int warps = (rows / 2) * SG;
    const int cols = nB < HL_MAX_COLS ? ((nB + 31) & ~31) : HL_MAX_COLS;
    const size_t smem = (size_t)cols * 32;
    if (smem > 48 * 1024) { *err = os1_optin_smem((const void*)hamming_knn2_lanes); if (*err != cudaSuccess) return true; }
    hamming_knn2_lanes<<<(nA + rows - 1) / rows, warps * 32, smem, s>>>((const uint4*)A, nA, (const uint4*)B, nB, cols, SG, bestIdx, bestDist, secondDist);
    *err = cudaGetLastError();
    return true;
}

// `counters` (one zero-initialised unsigned per block of HM_ROWS rows, left at zero again by the launch) selects the
// single-launch form; without it the partials are merged by a second kernel.
cudaError_t os1_launch_hamming_knn2_splits(const uint8_t* A, int nA, const uint8_t* B, int nB, int* bestIdx, int* bestDist,
                                           int* secondDist, uint64_t* scratch, int nSplit, cudaStream_t s, unsigned* counters)
{
    if (nA <= 0) return cudaSuccess;
    if (counters) {
        int dev = 0, sms = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaSuccess;
        if (launch_lanes(A, nA, B, nB, bestIdx, bestDist, secondDist, sms, s, &e)) return e;
    }
    int splitLen = (nB + nSplit - 1) / nSplit;
    splitLen = ((splitLen + HM_TILE - 1) / HM_TILE) * HM_TILE;
    if (splitLen < HM_TILE) splitLen = HM_TILE;
    if (splitLen > (int)HM_POSMASK) return cudaErrorInvalidValue;
    dim3 grid((nA + HM_ROWS - 1) / HM_ROWS, nSplit);
    hamming_knn2_partial<<<grid, HM_ROWS, 0, s>>>((const uint4*)A, nA, (const uint4*)B, nB, splitLen, (uint2*)scratch, counters, bestIdx, bestDist,
                                                  secondDist);
    if (!counters) hamming_knn2_merge<<<(nA + 127) / 128, 128, 0, s>>>((const uint2*)scratch, nA, nSplit, splitLen, bestIdx, bestDist, secondDist);
    return cudaGetLastError();
}

cudaError_t os1_launch_hamming_window(const uint8_t* A, int nA, const uint8_t* B, const int* ptr, const int* cand, int* bestIdx,
                                      int* bestDist, int* secondDist, cudaStream_t s)
{
    if (nA <= 0) return cudaSuccess;
    hamming_window_kernel<<<(nA + 3) / 4, 128, 0, s>>>((const uint4*)A, nA, (const uint4*)B, ptr, cand, bestIdx, bestDist, secondDist);
    return cudaGetLastError();
}

// ---- allocation-free all-pairs matcher --------------------------------------------------------------------------
// os1_hamming_knn2_noalloc (include/os1_b200.h): same kernels as os1_hamming_knn2, but the per-split partial results
// live in a fixed array that is part of this library's device image instead of a stream-ordered allocation.
// Limits: nA <= 4096 rows, nB <= 65536 rows, at most 32 partitions of B -> 32 * 4096 * 8 B = 1 MiB (+ 256 B of merge
// tickets), resident on every GPU that loads the library until the process exits.
#define OS1_NOALLOC_MAX_A 4096
#define OS1_NOALLOC_MAX_B 65536
#define OS1_NOALLOC_MAX_SPLITS 32

static __device__ __align__(256) uint2 g_noalloc_partials[OS1_NOALLOC_MAX_SPLITS * OS1_NOALLOC_MAX_A];
// tickets of the single-launch merge, one per block of HM_ROWS rows: zero when the library is loaded, and put back to
// zero by every launch before it ends (the CTA that draws a row block's last ticket resets it)
static __device__ __align__(256) unsigned g_noalloc_tickets[OS1_NOALLOC_MAX_A / HM_ROWS];

namespace {
struct NoAllocDevice { uint2* partials; unsigned* tickets; int sms; bool ready; };
NoAllocDevice g_noalloc_dev[64];
std::mutex g_noalloc_mu;
}

void os1_internal_set_error(const char* msg);      // api.cu

extern "C" int os1_hamming_knn2_noalloc(const uint8_t* A, int nA, const uint8_t* B, int nB, int32_t* best_idx, int32_t* best_dist,
                                        int32_t* second_dist, int device, void* cuda_stream)
{
    if (nA < 0 || nB < 0 || (nA > 0 && (!A || !best_idx || !best_dist || !second_dist)) || (nB > 0 && !B) || device < 0 || device >= 64) {
        os1_internal_set_error("os1_hamming_knn2_noalloc: bad argument");
        return OS1_E_INVALID;
    }
    if (nA > OS1_NOALLOC_MAX_A || nB > OS1_NOALLOC_MAX_B) {
        os1_internal_set_error("os1_hamming_knn2_noalloc: problem larger than the built-in scratch (limits: nA <= 4096, nB <= 65536); "
                               "use os1_hamming_knn2, which sizes its scratch per call");
        return OS1_E_CAPACITY;
    }
    if (nA == 0) return OS1_OK;
    DeviceScope _ds(device); cudaError_t e = _ds.err;
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_hamming_knn2_noalloc: cudaSetDevice failed (no CPU fallback)"); return OS1_E_CUDA; }
    NoAllocDevice dev;
    {
        // One-time set-up per device, with that device current: in host code `g_noalloc_partials` is only a handle
        // for the symbol, not a device address, so the address the kernels need is asked from the runtime.
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        NoAllocDevice& d = g_noalloc_dev[device];
        if (!d.ready) {
            uint2* p = nullptr;
            e = cudaGetSymbolAddress((void**)&p, g_noalloc_partials);
            if (e != cudaSuccess || !p) {
                cudaGetLastError();
                os1_internal_set_error("os1_hamming_knn2_noalloc: cudaGetSymbolAddress(scratch) failed");
                return OS1_E_CUDA;
            }
            unsigned* t = nullptr;
            e = cudaGetSymbolAddress((void**)&t, g_noalloc_tickets);
            if (e != cudaSuccess || !t) {
                cudaGetLastError();
                os1_internal_set_error("os1_hamming_knn2_noalloc: cudaGetSymbolAddress(tickets) failed");
                return OS1_E_CUDA;
            }
            int sms = 0;
            if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || sms <= 0) { cudaGetLastError(); sms = 148; }
            d.partials = p; d.tickets = t; d.sms = sms; d.ready = true;
        }
        dev = d;
    }
    int nSplit = os1_hamming_splits(nA, nB, dev.sms);
    if (nSplit > OS1_NOALLOC_MAX_SPLITS) nSplit = OS1_NOALLOC_MAX_SPLITS;
    // The partials need no clearing: the merge reads exactly the nSplit x nA entries that THIS launch has just written, and
    // whatever an earlier call left behind is never read.  The tickets must start at zero and do: each launch hands them
    // back at zero.
    e = os1_launch_hamming_knn2_splits(A, nA, B, nB, best_idx, best_dist, second_dist, (uint64_t*)dev.partials, nSplit,
                                       (cudaStream_t)cuda_stream, dev.tickets);
    if (e != cudaSuccess) { cudaGetLastError(); os1_internal_set_error("os1_hamming_knn2_noalloc: kernel launch failed"); return OS1_E_CUDA; }
    return OS1_OK;
}
