// K3 — keypoint distribution (replaces ORBextractor::DistributeOctTree and ExtractorNode::DivideNode,
// src/ORBextractor.cc:571-795, :513-569, plus the fix-up of :879-889 which the describe kernel applies).
//
// One CTA per (frame, level) runs the whole culling in rounds, without the reference's std::list:
//   * prologue: exclusive scan of the per-cell counts written by K2 and an ordered gather of the staged
//     candidates -> the reference's vToDistributeKeys order (cell-row-major, row-major inside a cell);
//   * every candidate carries a PATH CODE: root index plus, for each depth, the (x>=mid, y>=mid) bits of
//     DivideNode's split.  DivideNode splits x and y independently (halfX from the x range only, :515-516),
//     so the bits come from two small per-level tables indexed by x and by y (built on the host with the
//     reference's float arithmetic); the node a candidate belongs to at depth d is a prefix of its code;
//   * a round = count the four children of every live (multi-key) node with warp-aggregated shared-memory
//     atomics, decide which nodes split (all of them in a full pass, :626-697; largest-first until the count
//     reaches N in the final phase, :708-769, by a shared-memory bitonic sort on (size, creation rank) and a
//     prefix sum), give the children their creation ranks by prefix sums, and re-label the candidates;
//   * single-key children and unsplit nodes become leaves keyed by (round, rank); the best response per leaf
//     is an atomicMax on (response<<24 | ~index) so that the first maximum wins as in :784;
//   * the leaves are sorted into std::list front-to-back order (:776): newest round first, newest rank first,
//     roots in push_back order.
// Tie rule for equal-size nodes: later-created first (DESIGN.md "canonical tie rule").
#include "os1_internal.cuh"
#include <cstdlib>

#define QT_DONE 0xFFFFu
// batches: four 512-thread CTAs per SM (32 registers per thread, no spills) — the kernel is bound by barriers and latency,
// and at the compiler's own 56 registers only two CTAs were resident (0.44 -> 0.35 ms per 256 frames)
#ifndef QT_MIN_CTAS_512
#define QT_MIN_CTAS_512 4
#endif
#define QT_LEAF 0x80000000u

namespace {

struct QtParams {
    Geometry g;
    DevPtrs d;
    int maxNodes, maxLeaves, maxCells, sortCap;
};

template <int QT_THREADS>
__device__ __forceinline__ int block_scan_excl(int v, int* warpSums, int& total)
{
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    __syncthreads();                 // protect warpSums from the previous call
    if (lane == 31) warpSums[w] = inc;
    __syncthreads();
    if (w == 0) {
        int s = lane < QT_THREADS / 32 ? warpSums[lane] : 0;
        int si = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, si, o);
            if (lane >= o) si += t;
        }
        warpSums[lane] = si - s;     // exclusive warp offsets
        if (lane == 31) warpSums[32] = si;
    }
    __syncthreads();
    total = warpSums[32];
    return warpSums[w] + inc - v;
}

// bitonic sort of n (a power of two) 64-bit keys in a[0..n), descending when desc != 0; `a` has room for 2n keys.
// Up to one key per thread the keys stay in registers: partners less than 32 apart are exchanged by shuffles, the others
// through shared memory, alternating between the two halves of `a` so that such a step costs one barrier — 10 barriers
// instead of 45 for 512 keys, which is most of what a latency-bound CTA spends in its sorts.
template <int QT_THREADS>
__device__ void bitonic_sort(unsigned long long* a, int n, bool desc)
{
    const int tid = threadIdx.x;
    if (n <= QT_THREADS) {
        if ((tid & ~31) >= n) {
            // a warp without keys only keeps the barrier count of the exchange steps below
#pragma unroll 1
            for (int k = 64; k <= n; k <<= 1)
#pragma unroll 1
                for (int j = k >> 1; j >= 32; j >>= 1) __syncthreads();
            __syncthreads();
            __syncthreads();
            return;
        }
        unsigned long long x = tid < n ? a[tid] : 0ull;
        int cur = 0;                        // which half of `a` the next shared-memory exchange writes
#pragma unroll 1
        for (int k = 2; k <= n; k <<= 1) {
#pragma unroll 1
            for (int j = k >> 1; j > 0; j >>= 1) {
                unsigned long long y;
                if (j >= 32) {
                    unsigned long long* buf = a + cur * n;
                    if (tid < n) buf[tid] = x;
                    __syncthreads();
                    y = tid < n ? buf[tid ^ j] : 0ull;
                    cur ^= 1;               // the next exchange writes the other half: no second barrier needed
                } else {
                    y = __shfl_xor_sync(0xffffffffu, x, j);
                }
                const bool lower = (tid & j) == 0;
                const bool up = ((tid & k) == 0) != desc;
                const bool takeMin = lower == up;
                x = takeMin ? (x < y ? x : y) : (x < y ? y : x);
            }
        }
        __syncthreads();                    // every exchange buffer has been read
        if (tid < n) a[tid] = x;
        __syncthreads();
        return;
    }
#pragma unroll 1
    for (int k = 2; k <= n; k <<= 1)
#pragma unroll 1
        for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll 1
            for (int i = threadIdx.x; i < n; i += QT_THREADS) {
                int ixj = i ^ j;
                if (ixj > i) {
                    unsigned long long x = a[i], y = a[ixj];
                    bool up = ((i & k) == 0) != desc;
                    if ((x > y) == up) { a[i] = y; a[ixj] = x; }
                }
            }
            __syncthreads();
        }
}

template <int QT_THREADS>
__global__ void __launch_bounds__(QT_THREADS, QT_THREADS == 512 ? QT_MIN_CTAS_512 : 1) quadtree_kernel(const __grid_constant__ QtParams P)
{
    extern __shared__ __align__(16) uint8_t smraw[];
    const Geometry& g = P.g;
    const int level = blockIdx.y, frame = blockIdx.x;      // level-major launch order: the big levels first, the small ones fill the tail
    const LevelGeom& L = g.lv[level];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = L.quota;
    const int D = OS1_PATH_BITS;

    // ---- shared memory carve-up ---------------------------------------------------------------------------
    unsigned long long* sortBuf = (unsigned long long*)smraw;                 // max(sortCap, 2 * min(sortCap, 1024)): the register sort ping-pongs
    uint32_t* cellOffs = (uint32_t*)(sortBuf + max(P.sortCap, 2 * min(P.sortCap, 1024)));       // maxCells + 1
    uint32_t* slotCnt = cellOffs + P.maxCells + 1;                             // 4 * maxNodes
    uint32_t* slotInfo = slotCnt + 4 * P.maxNodes;                             // 4 * maxNodes
    uint32_t* liveSize = slotInfo + 4 * P.maxNodes;                            // maxNodes  (x2 buffers)
    uint32_t* liveKey = liveSize + 2 * P.maxNodes;                             // maxNodes  (x2 buffers)
    uint32_t* procPos = liveKey + 2 * P.maxNodes;                              // maxNodes
    uint32_t* order = procPos + P.maxNodes;                                    // maxNodes
    uint32_t* nzCum = order + P.maxNodes;                                      // maxNodes + 1
    uint32_t* mzCum = nzCum + P.maxNodes + 1;                                  // maxNodes + 1
    uint32_t* unsplitLeaf = mzCum + P.maxNodes + 1;                            // maxNodes
    uint32_t* leafKey = unsplitLeaf + P.maxNodes;                              // maxLeaves
    uint32_t* leafBest = leafKey + P.maxLeaves;                                // maxLeaves
    int* warpSums = (int*)(leafBest + P.maxLeaves);                            // 33
    __shared__ int s_leafCount, s_cut, s_err;

    const long long fbase = (long long)frame * g.candPerFrame + L.candOff;
    const uint32_t* stage = P.d.stage + fbase;
    uint32_t* cand = P.d.cand + fbase;
    uint32_t* code = P.d.code + fbase;
    uint16_t* node = P.d.node + fbase;
    const uint32_t* cellCnt = P.d.cellCnt + (long long)frame * g.cellsPerFrame + L.cellOff;
    const uint32_t* xs = P.d.ptab + L.xsOff;
    const uint32_t* ys = P.d.ptab + L.ysOff;
    uint32_t* outKp = P.d.lvKp + (long long)frame * g.kpCapPerFrame + L.kpOff;

    if (tid == 0) { s_leafCount = 0; s_err = 0; }

    // ---- 0. scan of cell counts -------------------------------------------------------------------------------
    int n;
    {
        int carry = 0;
#pragma unroll 1
        for (int base = 0; base < L.nCells; base += QT_THREADS) {
            int c = base + tid;
            int v = (c < L.nCells) ? (int)cellCnt[c] : 0;
            int tot;
            int ex = block_scan_excl<QT_THREADS>(v, warpSums, tot);
            if (c < L.nCells) cellOffs[c] = carry + ex;
            carry += tot;
        }
        n = carry;
        if (tid == 0) { cellOffs[L.nCells] = n; P.d.candCount[frame * g.nlevels + level] = n; }
    }
    __syncthreads();

    // ---- 1. ordered gather + path codes (eight lanes per cell: a cell rarely holds more candidates, and four cells per
    // warp trip means four times fewer dependent-load round trips) ---------------------------------------------------
    {
        const int sub = lane >> 3, l8 = lane & 7;
#pragma unroll 2
        for (int c = warp * 4 + sub; c < L.nCells; c += QT_THREADS / 8) {
            const int off = cellOffs[c], cnt = cellOffs[c + 1] - off;
            const uint32_t* s = stage + (long long)c * L.cellCap;
#pragma unroll 1
            for (int r = l8; r < cnt; r += 8) {
                uint32_t v = s[r];
                cand[off + r] = v;
                code[off + r] = __ldg(xs + OS1_X(v)) | __ldg(ys + OS1_Y(v));
            }
        }
    }
    __syncthreads();

    // ---- 2. roots (:584-617) ----------------------------------------------------------------------------------------
    const int nIni = L.nIni;
#pragma unroll 1
    for (int i = tid; i < 4 * P.maxNodes; i += QT_THREADS) slotCnt[i] = 0;
    __syncthreads();
#pragma unroll 1
    for (int base = 0; base < n; base += 4 * QT_THREADS) {
        uint32_t cd[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const int i = base + u * QT_THREADS + tid;
            cd[u] = i < n ? code[i] : 0xFFFFFFFFu;
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const bool valid = cd[u] != 0xFFFFFFFFu;
            const unsigned m = __ballot_sync(0xffffffffu, valid);
            if (valid) {
                const int r = cd[u] >> (2 * D);
                const unsigned grp = __match_any_sync(m, r);
                if (lane == __ffs(grp) - 1) atomicAdd(&slotCnt[r], __popc(grp));
            }
        }
    }
    __syncthreads();
    int E = 0, Lcount = 0;          // live multi-key nodes, lNodes.size()
    {
        // nIni is tiny: every thread walks the roots redundantly, thread 0 writes
        int e = 0, lc = 0;
#pragma unroll 1
        for (int r = 0; r < nIni; r++) {
            const uint32_t c = slotCnt[r];
            if (c == 0) continue;
            lc++;
            if (c == 1) {
                if (tid == 0) { int id = s_leafCount++; leafKey[id] = (0xFFFFu << 16) | r; leafBest[id] = 0; slotInfo[r] = QT_LEAF | id; }
            } else {
                if (tid == 0) { liveSize[e] = c; liveKey[e] = r; slotInfo[r] = e; }
                e++;
            }
        }
        E = e; Lcount = lc;
    }
    __syncthreads();
#pragma unroll 1
    for (int i = tid; i < n; i += QT_THREADS) {
        const uint32_t info = slotInfo[code[i] >> (2 * D)];
        if (info & QT_LEAF) { atomicMax(&leafBest[info & 0xFFFF], (cand[i] & 0xFF000000u) | (0xFFFFFFu - i)); node[i] = QT_DONE; }
        else node[i] = (uint16_t)info;
    }
    __syncthreads();

    // ---- 3. rounds ------------------------------------------------------------------------------------------------------
    int phase = 1, depth = 0, round = 0, buf = 0;
    bool haveCounts = false;                      // this round's children sizes were counted by the previous round's sweep B
    uint32_t* nextCnt = (uint32_t*)sortBuf;       // ... into the sort buffer, which is idle during a sweep
    const int nextCap = 2 * max(P.sortCap, 2 * min(P.sortCap, 1024));      // its size in 32-bit words
    while (E > 0) {
        round++;
        uint32_t* curSize = liveSize + buf * P.maxNodes;
        uint32_t* curKey = liveKey + buf * P.maxNodes;
        uint32_t* newSize = liveSize + (buf ^ 1) * P.maxNodes;
        uint32_t* newKey = liveKey + (buf ^ 1) * P.maxNodes;
        if (depth + 1 > D) { if (tid == 0) atomicOr(P.d.status, OS1_ERR_DEPTH); break; }
        const int shift = 2 * (D - (depth + 1));

        // children sizes counted by the previous round's re-labelling sweep wait in the sort buffer: move them out before the
        // final phase sorts in it
        if (haveCounts) {
#pragma unroll 1
            for (int i = tid; i < 4 * E; i += QT_THREADS) slotCnt[i] = nextCnt[i];
            __syncthreads();
        }

        // processing order: list order in a full pass (roots ascending, afterwards newest first);
        // (size, creation rank) descending in the final phase (:716-717)
        if (phase == 1) {
#pragma unroll 1
            for (int j = tid; j < E; j += QT_THREADS) {
                int p = (round == 1) ? j : E - 1 - j;
                procPos[j] = p; order[p] = j;
            }
        } else {
            int n2 = 1; while (n2 < E) n2 <<= 1;
#pragma unroll 1
            for (int j = tid; j < n2; j += QT_THREADS)
                sortBuf[j] = (j < E) ? (((unsigned long long)curSize[j] << 32) | ((unsigned long long)(curKey[j] & 0xFFFFu) << 16) | (unsigned)j) : 0ull;
            __syncthreads();
            bitonic_sort<QT_THREADS>(sortBuf, n2, true);
#pragma unroll 1
            for (int p = tid; p < E; p += QT_THREADS) {
                int j = (int)(sortBuf[p] & 0xFFFFu);
                order[p] = j; procPos[j] = p;
            }
        }
        if (!haveCounts) {
#pragma unroll 1
        for (int i = tid; i < 4 * E; i += QT_THREADS) slotCnt[i] = 0;
        __syncthreads();

        // sweep A: children sizes (first round only: from the second round on the re-labelling sweep of the previous round has
        // already counted them)
        // (four candidates per thread per trip: the eight loads are in flight together — the sweeps are latency bound)
#pragma unroll 1
        for (int base = 0; base < n; base += 4 * QT_THREADS) {
            int nd[4]; uint32_t cd[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int i = base + u * QT_THREADS + tid;
                nd[u] = i < n ? (int)node[i] : (int)QT_DONE;
                cd[u] = i < n ? code[i] : 0u;
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const bool valid = nd[u] != QT_DONE;
                const unsigned m = __ballot_sync(0xffffffffu, valid);
                if (valid) {
                    const int slot = nd[u] * 4 + ((cd[u] >> shift) & 3);
                    const unsigned grp = __match_any_sync(m, slot);
                    if (lane == __ffs(grp) - 1) atomicAdd(&slotCnt[slot], __popc(grp));
                }
            }
        }
        }
        __syncthreads();

        // non-empty / multi-key children per node, cumulated in processing order
        {
            int carryN = 0, carryM = 0;
#pragma unroll 1
            for (int base = 0; base < E; base += QT_THREADS) {
                const int p = base + tid;
                int nz = 0, mz = 0;
                if (p < E) {
                    const int j = order[p];
#pragma unroll
                    for (int q = 0; q < 4; q++) { uint32_t c = slotCnt[4 * j + q]; nz += (c > 0); mz += (c > 1); }
                }
                // one scan for both counts: each stays below 2^16 per trip (at most 4 children x QT_THREADS nodes)
                int tot;
                const int ex = block_scan_excl<QT_THREADS>(nz | (mz << 16), warpSums, tot);
                if (p < E) { nzCum[p] = carryN + (ex & 0xFFFF); mzCum[p] = carryM + (ex >> 16); }
                carryN += tot & 0xFFFF; carryM += tot >> 16;
            }
            if (tid == 0) { nzCum[E] = carryN; mzCum[E] = carryM; s_cut = E - 1; }
        }
        __syncthreads();
        if (phase == 2) {
            // first processing position at which the node count reaches N (:762)
#pragma unroll 1
            for (int p = tid; p < E; p += QT_THREADS) {
                const bool here = Lcount + (int)nzCum[p + 1] - (p + 1) >= N;
                const bool before = p > 0 && (Lcount + (int)nzCum[p] - p >= N);
                if (here && !before) s_cut = p;
            }
            __syncthreads();
        }
        const int nSplit = s_cut + 1;
        const int Lnew = Lcount + (int)nzCum[nSplit] - nSplit;
        const int newE = (int)mzCum[nSplit];
        bool fin = (Lnew >= N) || (Lnew == Lcount) || (newE == 0);
        int nextPhase = phase;
        if (phase == 1 && !fin && Lnew + 3 * newE > N) nextPhase = 2;
        if (!fin && newE > P.maxNodes) { fin = true; if (tid == 0) atomicOr(P.d.status, OS1_ERR_NODE_OVERFLOW); }

        // children of split nodes; unsplit nodes turn into leaves as they are
#pragma unroll 1
        for (int p = tid; p < E; p += QT_THREADS) {
            const int j = order[p];
            if (p < nSplit) {
                int rank = nzCum[p], mi = mzCum[p];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const uint32_t c = slotCnt[4 * j + q];
                    if (c == 0) continue;
                    const uint32_t key = ((uint32_t)round << 16) | (uint32_t)rank;
                    rank++;
                    if (c == 1 || fin) {
                        int id = atomicAdd(&s_leafCount, 1);
                        if (id < P.maxLeaves) {
                            leafKey[id] = ((0xFFFFu - (key >> 16)) << 16) | (0xFFFFu - (key & 0xFFFFu));
                            leafBest[id] = 0;
                        } else { id = 0; s_err = 1; }
                        slotInfo[4 * j + q] = QT_LEAF | id;
                        if (c > 1) mi++;
                    } else {
                        newSize[mi] = c; newKey[mi] = key;
                        slotInfo[4 * j + q] = mi;
                        mi++;
                    }
                }
            } else {
                int id = atomicAdd(&s_leafCount, 1);
                const uint32_t key = curKey[j];
                if (id < P.maxLeaves) {
                    // roots keep push_back order and sort after everything else
                    leafKey[id] = (key >> 16) == 0 ? ((0xFFFFu << 16) | key) : (((0xFFFFu - (key >> 16)) << 16) | (0xFFFFu - (key & 0xFFFFu)));
                    leafBest[id] = 0;
                } else { id = 0; s_err = 1; }
                unsplitLeaf[j] = id;
            }
        }
        __syncthreads();

        // sweep B: re-label candidates / vote for the best response of finished leaves — and, for the candidates that stay live,
        // count the children of their NEW node right away (what sweep A of the next round would do with a second pass over the
        // same candidates)
        const bool countNext = !fin && depth + 2 <= D && 4 * newE <= nextCap;
        const int shiftN = countNext ? shift - 2 : 0;
        if (countNext) {
#pragma unroll 1
            for (int i = tid; i < 4 * newE; i += QT_THREADS) nextCnt[i] = 0;
            __syncthreads();
        }
#pragma unroll 1
        for (int base = 0; base < n; base += 4 * QT_THREADS) {
            int nd[4]; uint32_t cd[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int i = base + u * QT_THREADS + tid;
                nd[u] = i < n ? (int)node[i] : (int)QT_DONE;
                cd[u] = i < n ? code[i] : 0u;
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                int slotN = -1;
                if (nd[u] != QT_DONE) {
                    const int i = base + u * QT_THREADS + tid;
                    if ((int)procPos[nd[u]] >= nSplit) {
                        atomicMax(&leafBest[unsplitLeaf[nd[u]]], (cand[i] & 0xFF000000u) | (0xFFFFFFu - i));
                        node[i] = QT_DONE;
                    } else {
                        const uint32_t info = slotInfo[nd[u] * 4 + ((cd[u] >> shift) & 3)];
                        if (info & QT_LEAF) {
                            atomicMax(&leafBest[info & 0xFFFF], (cand[i] & 0xFF000000u) | (0xFFFFFFu - i));
                            node[i] = QT_DONE;
                        } else {
                            node[i] = (uint16_t)info;
                            if (countNext) slotN = (int)info * 4 + (int)((cd[u] >> shiftN) & 3);
                        }
                    }
                }
                if (countNext) {
                    const unsigned m = __ballot_sync(0xffffffffu, slotN >= 0);
                    if (slotN >= 0) {
                        const unsigned grp = __match_any_sync(m, slotN);
                        if (lane == __ffs(grp) - 1) atomicAdd(&nextCnt[slotN], __popc(grp));
                    }
                }
            }
        }
        __syncthreads();
        Lcount = Lnew;
        if (fin) break;
        E = newE; phase = nextPhase; depth++; buf ^= 1;
        haveCounts = countNext;
    }
    __syncthreads();

    // ---- 4. leaves -> list order -> kept keypoints ------------------------------------------------------------------
    int nLeaves = min(s_leafCount, P.maxLeaves);
    if (s_err && tid == 0) atomicOr(P.d.status, OS1_ERR_NODE_OVERFLOW);
    {
        int n2 = 1; while (n2 < nLeaves) n2 <<= 1;
#pragma unroll 1
        for (int j = tid; j < n2; j += QT_THREADS)
            sortBuf[j] = (j < nLeaves) ? (((unsigned long long)leafKey[j] << 32) | leafBest[j]) : ~0ull;
        __syncthreads();
        bitonic_sort<QT_THREADS>(sortBuf, n2, false);
        if (nLeaves > L.kpCap) { if (tid == 0) atomicOr(P.d.status, OS1_ERR_KP_OVERFLOW); nLeaves = L.kpCap; }
#pragma unroll 1
        for (int j = tid; j < nLeaves; j += QT_THREADS) {
            const uint32_t best = (uint32_t)sortBuf[j];
            const int idx = 0xFFFFFF - (best & 0xFFFFFF);
            outKp[j] = cand[idx];
        }
        if (tid == 0) P.d.lvCount[frame * g.nlevels + level] = nLeaves;
    }
}

} // namespace

// shared memory of one quadtree CTA, and the sizes it is derived from (the launcher and the geometry check agree on it)
static size_t quadtree_smem(int maxQuota, int maxCells, int maxIni, QtParams* P)
{
    const int maxNodes = max(maxQuota, maxIni) + 8;
    const int maxLeaves = max(maxQuota + 2, 4 * maxIni) + 8;
    int sc = 1; while (sc < max(maxNodes, maxLeaves)) sc <<= 1;
    if (P) { P->maxNodes = maxNodes; P->maxLeaves = maxLeaves; P->maxCells = maxCells; P->sortCap = sc; }
    // the sort keeps its keys in registers and ping-pongs between two halves only up to one key per thread (<= 1024)
    const size_t sortKeys = (size_t)max(sc, 2 * min(sc, 1024));
    return sortKeys * 8 + (size_t)(maxCells + 1) * 4 + (size_t)maxNodes * 4 * (4 + 4 + 2 + 2 + 1 + 1 + 1 + 1 + 1) + 8 * 4 + (size_t)maxLeaves * 8 + 34 * 4;
}

bool os1_quadtree_fits(int maxQuota, int maxCells, int maxIni) { return quadtree_smem(maxQuota, maxCells, maxIni, nullptr) <= 220 * 1024; }

cudaError_t os1_launch_quadtree(const Geometry& g, const DevPtrs& d, int nframes, int maxQuota, int maxCells, cudaStream_t s)
{
    static_assert(sizeof(QtParams) < 4000, "kernel parameter block too large");
    QtParams P;
    P.g = g; P.d = d;
    int maxIni = 1;
    for (int l = 0; l < g.nlevels; l++) maxIni = max(maxIni, g.lv[l].nIni);
    const size_t smem = quadtree_smem(maxQuota, maxCells, maxIni, &P);
    if (smem > 220 * 1024 || P.maxNodes > 16000) return cudaErrorNotSupported;
    dim3 grid(nframes, g.nlevels);
    // few frames: the level-0 CTA is the critical path, give it 1024 threads; batches: 512 threads, two CTAs per SM
    if (nframes <= 4) {
        if (smem > 48 * 1024) {
            cudaError_t e = os1_optin_smem((const void*)quadtree_kernel<1024>);
            if (e != cudaSuccess) return e;
        }
        quadtree_kernel<1024><<<grid, 1024, smem, s>>>(P);
    } else {
        if (smem > 48 * 1024) {
            cudaError_t e = os1_optin_smem((const void*)quadtree_kernel<512>);
            if (e != cudaSuccess) return e;
        }
        quadtree_kernel<512><<<grid, 512, smem, s>>>(P);
    }
    return cudaGetLastError();
}
