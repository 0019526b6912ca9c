// K2 — FAST-9/16 score, cell-local non-max suppression and the per-cell 20 -> 7 retry
// (replaces the cell loop of ORBextractor::ComputeKeyPointsOctTree, src/ORBextractor.cc:801-870, and the
// cv::FAST calls at :848,:854).
//
// Formulation (verified identical to the per-cell cv::FAST calls): the reference's
// overlapping (wCell+6)-wide sub-images have TESTED regions that tile the level exactly, so the true FAST
// score s(p) = max over the 16 arcs of 9 ring pixels of min|ring - centre| - 1 is computed once per pixel;
// a pixel survives when s(p) is strictly greater than its 8 neighbours *inside the same cell's tested
// region*; a cell emits its survivors with s >= iniTh, or, if there are none, those with s >= minTh.
// Survivors are written to the cell's staging slots in row-major order (= cv::FAST's output order), with the
// in-cell rank obtained from warp ballots + popc; the quadtree kernel concatenates cells in cell-row-major
// order, reproducing vToDistributeKeys.
//
// A CTA handles FAST_G horizontally adjacent cells of one cell row: tile -> shared memory, score map,
// keep map, then one warp per cell for the threshold vote and the ordered emission.
#include "os1_internal.cuh"

#define FAST_THREADS 256

namespace {

struct FastParams {
    Geometry g;
    DevPtrs d;
    int tilePitch;      // shared-memory pitch (bytes)
    int tileRows;
};

// strength = max over arcs of 9 contiguous ring pixels of min(c - ring) and of min(ring - c).
// Both polarities are computed as MIN chains over d = c - ring and e = ring - c.  Do NOT rewrite the dark side
// as max(best, -max(d...)): nvcc 12.9 / ptxas for sm_100a miscompiles an integer negation feeding min/max
// (scripts/scratch/var.cu reproduces it: variant A wrong on a B200, variants B/C right; DESIGN.md "toolchain").
__device__ __forceinline__ int fast_strength(const uint8_t* p, int pitch)
{
    const int c = p[0];
    int r[16];
    r[0] = p[3 * pitch];
    r[1] = p[3 * pitch + 1];
    r[2] = p[2 * pitch + 2];
    r[3] = p[pitch + 3];
    r[4] = p[3];
    r[5] = p[-pitch + 3];
    r[6] = p[-2 * pitch + 2];
    r[7] = p[-3 * pitch + 1];
    r[8] = p[-3 * pitch];
    r[9] = p[-3 * pitch - 1];
    r[10] = p[-2 * pitch - 2];
    r[11] = p[-pitch - 3];
    r[12] = p[-3];
    r[13] = p[pitch - 3];
    r[14] = p[2 * pitch - 2];
    r[15] = p[3 * pitch - 1];
    // min over 3, then min over 9 = min3 of three min3; same with max for the ring itself
    int lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        lo3[k] = __vimin3_s32(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
        hi3[k] = __vimax3_s32(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
    }
    int darkest = 0, brightest = 255;      // max over arcs of (min ring), min over arcs of (max ring)
#pragma unroll
    for (int k = 0; k < 16; k++) {
        darkest = max(darkest, __vimin3_s32(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]));
        brightest = min(brightest, __vimax3_s32(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]));
    }
    // ring all brighter than c by (darkest - c); ring all darker than c by (c - brightest)
    return max(max(darkest - c, c - brightest), 0);
}

__global__ void __launch_bounds__(FAST_THREADS) fast_kernel(const __grid_constant__ FastParams P)
{
    extern __shared__ __align__(16) uint8_t sm[];
    const Geometry& g = P.g;
    const int frame = blockIdx.y;
    // locate (level, cell row, group)
    int level = 0;
    while (level + 1 < g.nlevels && (int)blockIdx.x >= g.lv[level + 1].groupOff) level++;
    const LevelGeom& L = g.lv[level];
    const int local = blockIdx.x - L.groupOff;
    const int ci = local / L.groupsPerRow;
    const int j0 = (local % L.groupsPerRow) * FAST_G;
    const int ncell = min(FAST_G, L.nCols - j0);

    const int x0 = OS1_BORDER + j0 * L.wCell;
    const int x1 = min(OS1_BORDER + (j0 + ncell) * L.wCell + 6, L.maxBX);
    const int y0 = OS1_BORDER + ci * L.hCell;
    const int y1 = min(y0 + L.hCell + 6, L.maxBY);
    const int tw = x1 - x0, th = y1 - y0;
    const int pitch = P.tilePitch;
    uint8_t* img = sm;                              // tile pixels, origin (xa, y0)
    uint8_t* sc = sm + pitch * P.tileRows;          // scores, origin (xa, y0)
    uint8_t* kp = sc + pitch * P.tileRows;          // kept scores
    const int xa = x0 & ~3;                         // aligned tile origin
    const int xo = x0 - xa;

    const uint8_t* src;
    int spitch;
    if (level == 0) { src = P.d.lvl0 + (long long)frame * P.d.lvl0FrameStride; spitch = P.d.lvl0Pitch; }
    else { src = P.d.slab + (long long)frame * g.slabBytes + L.off; spitch = L.pitch; }

    uint32_t* cellCnt = P.d.cellCnt + (long long)frame * g.cellsPerFrame + L.cellOff + ci * L.nCols + j0;
    if (tw < 7 || th < 7) {       // nothing can be tested (ORBextractor.cc:831,:841 skip these cells)
        if (threadIdx.x < ncell) cellCnt[threadIdx.x] = 0;
        return;
    }

    // ---- phase 1: tile -> shared (aligned 32-bit words; rows of the slab are 16-byte aligned) ------------
    {
        const int words = (xo + tw + 3) >> 2;
        for (int i = threadIdx.x; i < words * th; i += FAST_THREADS) {
            int r = i / words, w = i - r * words;
            uint32_t v = *(const uint32_t*)(src + (long long)(y0 + r) * spitch + xa + 4 * w);
            *(uint32_t*)(img + r * pitch + 4 * w) = v;
        }
    }
    __syncthreads();

    // ---- phase 2: score of every tested pixel ---------------------------------------------------------------
    {
        const int iw = tw - 6, ih = th - 6;
        const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
        for (int r = wrp; r < ih; r += FAST_THREADS / 32)
            for (int c = lane; c < iw; c += 32) {
                const uint8_t* p = img + (r + 3) * pitch + xo + c + 3;
                int s = fast_strength(p, pitch) - 1;
                sc[(r + 3) * pitch + xo + c + 3] = (uint8_t)max(s, 0);
            }
    }
    __syncthreads();

    // ---- phase 3 + 4: one warp per cell ---------------------------------------------------------------------
    const int lane = threadIdx.x & 31, wrp = threadIdx.x >> 5;
    for (int cc = wrp; cc < ncell; cc += FAST_THREADS / 32) {
        const int cx0 = OS1_BORDER + (j0 + cc) * L.wCell;
        const int cx1 = min(cx0 + L.wCell + 6, L.maxBX);
        // tested region of this cell in tile coordinates (relative to xa, y0)
        const int ax0 = cx0 + 3 - xa, ax1 = cx1 - 3 - xa;
        const int ay0 = 3, ay1 = th - 3;
        int any = 0;
        // keep map + vote
        for (int r = ay0; r < ay1; r++)
            for (int cb = ax0; cb < ax1; cb += 32) {
                const int c = cb + lane;
                int keep = 0;
                if (c < ax1) {
                    const uint8_t* q = sc + r * pitch + c;
                    const int s = q[0];
                    if (s > 0) {
                        const bool L_ = c > ax0, R_ = c + 1 < ax1, U_ = r > ay0, D_ = r + 1 < ay1;
                        int m = 0;
                        if (L_) m = max(m, (int)q[-1]);
                        if (R_) m = max(m, (int)q[1]);
                        if (U_) {
                            m = max(m, (int)q[-pitch]);
                            if (L_) m = max(m, (int)q[-pitch - 1]);
                            if (R_) m = max(m, (int)q[-pitch + 1]);
                        }
                        if (D_) {
                            m = max(m, (int)q[pitch]);
                            if (L_) m = max(m, (int)q[pitch - 1]);
                            if (R_) m = max(m, (int)q[pitch + 1]);
                        }
                        if (s > m) keep = s;
                    }
                    kp[r * pitch + c] = (uint8_t)keep;
                }
                any |= (keep >= g.iniTh);
            }
        any = __any_sync(0xffffffffu, any);
        const int thr = any ? g.iniTh : g.minTh;
        __syncwarp();
        // ordered emission
        uint32_t* out = P.d.stage + (long long)frame * g.candPerFrame + L.candOff + (long long)(ci * L.nCols + j0 + cc) * L.cellCap;
        int base = 0;
        for (int r = ay0; r < ay1; r++)
            for (int cb = ax0; cb < ax1; cb += 32) {
                const int c = cb + lane;
                int s = (c < ax1) ? kp[r * pitch + c] : 0;
                const bool hit = s >= thr;
                const uint32_t bal = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    int rank = base + __popc(bal & ((1u << lane) - 1));
                    if (rank < L.cellCap) out[rank] = OS1_PACK_XYS(c + xa - OS1_BORDER, r + y0 - OS1_BORDER, s);
                    else atomicOr(P.d.status, OS1_ERR_CELL_OVERFLOW);
                }
                base += __popc(bal);
            }
        if (lane == 0) cellCnt[cc] = min(base, L.cellCap);
    }
}

} // namespace

cudaError_t os1_launch_fast(const Geometry& g, const DevPtrs& d, int nframes, cudaStream_t s)
{
    static_assert(sizeof(FastParams) < 4000, "kernel parameter block too large");
    FastParams P;
    P.g = g; P.d = d;
    int maxW = 0, maxH = 0;
    for (int l = 0; l < g.nlevels; l++) {
        maxW = max(maxW, FAST_G * g.lv[l].wCell + 6);
        maxH = max(maxH, g.lv[l].hCell + 6);
    }
    P.tilePitch = (maxW + 3 + 3) & ~3;
    P.tileRows = maxH;
    size_t smem = (size_t)3 * P.tilePitch * P.tileRows;
    if (smem > 200 * 1024) return cudaErrorNotSupported;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    dim3 grid(g.groupsPerFrame, nframes);
    fast_kernel<<<grid, FAST_THREADS, smem, s>>>(P);
    return cudaGetLastError();
}
