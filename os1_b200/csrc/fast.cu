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
// A CTA handles FAST_G horizontally adjacent cells of one cell row.  The work is integer-issue bound, so the
// kernel is organised around instruction count, not bytes:
//   phase 1  tile -> shared memory as aligned 32-bit words;
//   phase 2  a thread scores FOUR adjacent pixels from 21 word loads: each ring value of a pixel pair is lifted
//            into a u16x2 register by ONE byte-permute (value*257 per half-word, order preserving) and the
//            min-of-9 / max-of-9 arcs are min3(min3) chains on the packed VIMNMX3.U16x2 path (2 pixels/instr);
//            the response is finished in packed form too (saturating differences as max(a,b)-b, no cross-lane
//            borrow) and stored as max(strength - tmin, 0): zero below min(iniTh, minTh), which leaves NMS
//            unchanged and makes the maps sparse;
//   phase 3  NMS on the packed maps (native VIMNMX3.U16x2), skipping empty groups; cell boundaries are lane masks;
//            the few survivors are appended to a per-cell list in shared memory;
//   phase 4  per cell: vote "any survivor >= iniTh", rank every survivor by (row, column) by counting, emit.
#include "os1_internal.cuh"

#define FAST_THREADS 256

namespace {

struct FastParams {
    Geometry g;
    DevPtrs d;
    int pitchW;         // shared-memory row pitch in 32-bit words
    int tileRows;
    int listCap;        // survivor slots per cell in shared memory
};

// ---- packed ring access -------------------------------------------------------------------------------------
// The three words (wm, w0, wp) of a tile row hold bytes -4..7 relative to the first of the thread's four pixels.
// ring_pair<DX, ODD> returns, for the pixel pair (ODD ? {1,3} : {0,2}), the bytes at horizontal offset DX as
// u16x2 with each half-word = byte * 257.
template <int DX, int ODD>
__device__ __forceinline__ unsigned ring_pair(unsigned wm, unsigned w0, unsigned wp)
{
    constexpr int pa = DX + ODD, pb = pa + 2;
    if (DX <= 0) {
        constexpr unsigned ia = pa + 4, ib = pb + 4;
        return __byte_perm(wm, w0, ia | (ia << 4) | (ib << 8) | (ib << 12));
    } else {
        constexpr unsigned ia = pa, ib = pb;
        return __byte_perm(w0, wp, ia | (ia << 4) | (ib << 8) | (ib << 12));
    }
}

// max over the 16 arcs of 9 of (min ring), min over the arcs of (max ring), for one pixel pair
template <int ODD>
__device__ __forceinline__ void arc_extremes(const unsigned (&W)[7][3], unsigned& darkest, unsigned& brightest)
{
    unsigned r[16];
    // ring order of cv::FAST 9_16: (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
    r[0] = ring_pair<0, ODD>(W[6][0], W[6][1], W[6][2]);
    r[1] = ring_pair<1, ODD>(W[6][0], W[6][1], W[6][2]);
    r[2] = ring_pair<2, ODD>(W[5][0], W[5][1], W[5][2]);
    r[3] = ring_pair<3, ODD>(W[4][0], W[4][1], W[4][2]);
    r[4] = ring_pair<3, ODD>(W[3][0], W[3][1], W[3][2]);
    r[5] = ring_pair<3, ODD>(W[2][0], W[2][1], W[2][2]);
    r[6] = ring_pair<2, ODD>(W[1][0], W[1][1], W[1][2]);
    r[7] = ring_pair<1, ODD>(W[0][0], W[0][1], W[0][2]);
    r[8] = ring_pair<0, ODD>(W[0][0], W[0][1], W[0][2]);
    r[9] = ring_pair<-1, ODD>(W[0][0], W[0][1], W[0][2]);
    r[10] = ring_pair<-2, ODD>(W[1][0], W[1][1], W[1][2]);
    r[11] = ring_pair<-3, ODD>(W[2][0], W[2][1], W[2][2]);
    r[12] = ring_pair<-3, ODD>(W[3][0], W[3][1], W[3][2]);
    r[13] = ring_pair<-3, ODD>(W[4][0], W[4][1], W[4][2]);
    r[14] = ring_pair<-2, ODD>(W[5][0], W[5][1], W[5][2]);
    r[15] = ring_pair<-1, ODD>(W[6][0], W[6][1], W[6][2]);
    unsigned lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        lo3[k] = __vimin3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
        hi3[k] = __vimax3_u16x2(r[k], r[(k + 1) & 15], r[(k + 2) & 15]);
    }
    unsigned lo9[16], hi9[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        lo9[k] = __vimin3_u16x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);      // min of ring[k..k+8]
        hi9[k] = __vimax3_u16x2(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
    }
    unsigned d = __vimax3_u16x2(lo9[0], lo9[1], lo9[2]), b = __vimin3_u16x2(hi9[0], hi9[1], hi9[2]);
#pragma unroll
    for (int k = 3; k < 15; k += 2) {
        d = __vimax3_u16x2(d, lo9[k], lo9[k + 1]);
        b = __vimin3_u16x2(b, hi9[k], hi9[k + 1]);
    }
    darkest = __vimax3_u16x2(d, lo9[15], lo9[15]);
    brightest = __vimin3_u16x2(b, hi9[15], hi9[15]);
}

// u16x2 lanes holding value*257: per-lane saturating difference without any cross-lane borrow
__device__ __forceinline__ unsigned usubsat2(unsigned a, unsigned b) { return __vmaxu2(a, b) - b; }     // max(a-b, 0)

template <int PW>
__global__ void __launch_bounds__(FAST_THREADS) fast_kernel(const __grid_constant__ FastParams P)
{
    extern __shared__ __align__(16) unsigned smw[];
    const Geometry& g = P.g;
    const int frame = blockIdx.y;
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
    constexpr int pw = PW;                 // compile-time pitch: every shared-memory access gets an immediate offset
    // shared memory: tile words; two score maps holding the even (px 0,2) and odd (px 1,3) pixels of every group
    // of four as u16x2; six mask rows; per-cell survivor lists
    unsigned* img = smw;
    unsigned* scE = img + pw * P.tileRows;
    unsigned* scO = scE + pw * P.tileRows;
    unsigned* vmE = scO + pw * P.tileRows;      // lanes inside the tested columns
    unsigned* vmO = vmE + pw;
    unsigned* lmE = vmO + pw;                    // 0 in the lane whose pixel is a cell's FIRST tested column
    unsigned* lmO = lmE + pw;
    unsigned* rmE = lmO + pw;                    // 0 in the lane whose pixel is a cell's LAST tested column
    unsigned* rmO = rmE + pw;
    unsigned* lists = rmO + pw;                  // [FAST_G][listCap]: survivors >= iniTh grow from the front of a cell's
                                                 // slots, the weaker ones from the back (together <= the NMS bound)
    unsigned char* cellOf = (unsigned char*)(lists + FAST_G * P.listCap);   // cell index of every tile byte column
    __shared__ int s_n[2][FAST_G];
    const int listCap = P.listCap;
    const int xorg = (x0 & ~3) - 4;             // global x of tile byte 0 (one spare word on the left)
    const int words = ((x1 + 3 - xorg) >> 2) + 1;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const uint8_t* src;
    int spitch;
    if (level == 0) { src = P.d.lvl0 + (long long)frame * P.d.lvl0FrameStride; spitch = P.d.lvl0Pitch; }
    else { src = P.d.slab + (long long)frame * g.slabBytes + L.off; spitch = L.pitch; }

    uint32_t* cellCnt = P.d.cellCnt + (long long)frame * g.cellsPerFrame + L.cellOff + ci * L.nCols + j0;
    if (tw < 7 || th < 7) {       // nothing can be tested (ORBextractor.cc:831,:841 skip these cells)
        if (tid < ncell) cellCnt[tid] = 0;
        return;
    }
    const int tmin = min(g.iniTh, g.minTh);
    const int tc0 = x0 + 3 - xorg, tc1 = x1 - 3 - xorg;      // tested byte columns of the tile
    const int tr0 = 3, tr1 = th - 3;                          // tested rows
    const int wCell = L.wCell;

    // ---- phase 0/1: lane masks, tile -> shared, cleared score maps --------------------------------------------------
    if (tid < 2 * FAST_G) s_n[tid / FAST_G][tid % FAST_G] = 0;
    for (int w = tid; w < pw; w += FAST_THREADS) {
        unsigned vm[4], lm[4], rm[4];
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int col = 4 * w + b;
            const bool in = col >= tc0 && col < tc1;
            const int rel = col - tc0;
            const int q = in ? rel % wCell : 1;
            cellOf[col] = (unsigned char)(in ? rel / wCell : 0);
            vm[b] = in ? 0xFFFFu : 0u;
            lm[b] = (in && q == 0) ? 0u : 0xFFFFu;
            rm[b] = (in && (q == wCell - 1 || col == tc1 - 1)) ? 0u : 0xFFFFu;
        }
        vmE[w] = vm[0] | (vm[2] << 16); vmO[w] = vm[1] | (vm[3] << 16);
        lmE[w] = lm[0] | (lm[2] << 16); lmO[w] = lm[1] | (lm[3] << 16);
        rmE[w] = rm[0] | (rm[2] << 16); rmO[w] = rm[1] | (rm[3] << 16);
    }
    const int g0 = tc0 >> 2, g1 = (tc1 + 3) >> 2;
    // tile: asynchronous 4-byte copies straight into shared memory (LDGSTS), all in flight at once
    for (int r = warp; r < th; r += FAST_THREADS / 32) {
        const uint32_t* srow = (const uint32_t*)(src + (long long)(y0 + r) * spitch + xorg);
        for (int w = lane; w < words; w += 32)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(img + r * pw + w)), "l"(srow + w) : "memory");
    }
    // the NMS reads one word / one row beyond the scored area: only that frame has to be zero
    for (int w = g0 - 1 + tid; w <= g1; w += FAST_THREADS) {
        scE[(tr0 - 1) * pw + w] = 0; scO[(tr0 - 1) * pw + w] = 0;
        scE[tr1 * pw + w] = 0; scO[tr1 * pw + w] = 0;
    }
    for (int r = tr0 + tid; r < tr1; r += FAST_THREADS) {
        scE[r * pw + g0 - 1] = 0; scO[r * pw + g0 - 1] = 0;
        scE[r * pw + g1] = 0; scO[r * pw + g1] = 0;
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();

    const int ng = g1 - g0, nItems = ng * (tr1 - tr0);
    const unsigned rcpNg = (1u << 24) / (unsigned)ng + 1u;          // exact quotient for idx < 2^16 (nItems <= 136 * 64)
    // stored s8 = strength - tmin; response = strength - 1 = s8 + tmin - 1
    const int iniS8 = g.iniTh - tmin + 1, minS8 = g.minTh - tmin + 1;
    // ---- phase 2: scores, four pixels per thread, kept packed ---------------------------------------------------------
    {
        const unsigned T = (unsigned)tmin * 0x01010101u;       // tmin * 257 in both lanes
        // (rows x groups flattened so that every lane has work: 75 groups per row would leave a third of the lanes
        //  idle in the last of three 32-wide trips; the division by the group count is a multiply by its reciprocal)
        for (int idx = tid; idx < nItems; idx += FAST_THREADS) {
                const int rr = (int)(((unsigned)idx * rcpNg) >> 24);
                const int r = tr0 + rr, gq = g0 + idx - rr * ng;
                unsigned W[7][3];
                const unsigned* row = img + (r - 3) * pw + gq;
#pragma unroll
                for (int dy = 0; dy < 7; dy++) {
                    W[dy][0] = row[dy * pw - 1]; W[dy][1] = row[dy * pw]; W[dy][2] = row[dy * pw + 1];
                }
                unsigned dE, bE, dO, bO;
                arc_extremes<0>(W, dE, bE);
                arc_extremes<1>(W, dO, bO);
                const unsigned cE = __byte_perm(W[3][1], 0, 0x2200), cO = __byte_perm(W[3][1], 0, 0x3311);
                // strength*257 = max(darkest - c, c - brightest) per lane; stored value = max(strength - tmin, 0)*257,
                // i.e. non-zero exactly when the cv::FAST response (strength-1) reaches tmin — order preserving for NMS
                const unsigned sE = __vmaxu2(usubsat2(dE, cE), usubsat2(cE, bE));
                const unsigned sO = __vmaxu2(usubsat2(dO, cO), usubsat2(cO, bO));
                scE[r * pw + gq] = usubsat2(sE, T) & vmE[gq];
                scO[r * pw + gq] = usubsat2(sO, T) & vmO[gq];
            }
    }
    __syncthreads();

    // ---- phase 3: NMS on the packed maps; survivors go to their cell's list ----------------------------------------------
    for (int idx = tid; idx < nItems; idx += FAST_THREADS) {
            const int rr = (int)(((unsigned)idx * rcpNg) >> 24);
            const int r = tr0 + rr, gq = g0 + idx - rr * ng;
            const int o = r * pw + gq;
            const unsigned vE = scE[o], vO = scO[o];
            if ((vE | vO) == 0) continue;
            unsigned keepE = 0, keepO = 0;          // lanes that beat all eight neighbours
            {
                // even pixels (0,2): left neighbours = (odd[-1].hi, odd.lo), right neighbours = odd
                const unsigned lm = lmE[gq], rm = rmE[gq];
                unsigned m = __vimax3_u16x2(__byte_perm(scO[o - pw - 1], scO[o - pw], 0x5432) & lm, scE[o - pw], scO[o - pw] & rm);
                m = __vimax3_u16x2(m, __byte_perm(scO[o + pw - 1], scO[o + pw], 0x5432) & lm, scE[o + pw]);
                m = __vimax3_u16x2(m, scO[o + pw] & rm, __byte_perm(scO[o - 1], vO, 0x5432) & lm);
                m = __vmaxu2(m, vO & rm);
                if ((vE & 0xFFFFu) > (m & 0xFFFFu)) keepE |= 1;
                if ((vE >> 16) > (m >> 16)) keepE |= 2;
            }
            {
                // odd pixels (1,3): left neighbours = even, right neighbours = (even.hi, even[+1].lo)
                const unsigned lm = lmO[gq], rm = rmO[gq];
                unsigned m = __vimax3_u16x2(scE[o - pw] & lm, scO[o - pw], __byte_perm(scE[o - pw], scE[o - pw + 1], 0x5432) & rm);
                m = __vimax3_u16x2(m, scE[o + pw] & lm, scO[o + pw]);
                m = __vimax3_u16x2(m, __byte_perm(scE[o + pw], scE[o + pw + 1], 0x5432) & rm, vE & lm);
                m = __vmaxu2(m, __byte_perm(vE, scE[o + 1], 0x5432) & rm);
                if ((vO & 0xFFFFu) > (m & 0xFFFFu)) keepO |= 1;
                if ((vO >> 16) > (m >> 16)) keepO |= 2;
            }
            // bit b of `kept` <-> pixel b of the group; survivors are few, so this loop rarely runs
            unsigned kept = (keepE & 1) | ((keepO & 1) << 1) | ((keepE & 2) << 1) | ((keepO & 2) << 2);
            while (kept) {
                const int b = __ffs(kept) - 1;
                kept &= kept - 1;
                const unsigned v = (b & 1) ? vO : vE;
                const int s8 = ((b >> 1) ? (v >> 24) : (v >> 8)) & 0xFF;       // strength - tmin  (>= 1)
                const int col = 4 * gq + b;
                const int cc = cellOf[col];
                const int which = s8 >= iniS8 ? 0 : 1;
                const int slot = atomicAdd(&s_n[which][cc], 1);
                if (slot < listCap) lists[cc * listCap + (which ? listCap - 1 - slot : slot)] = ((unsigned)r << 22) | ((unsigned)col << 8) | (unsigned)s8;
            }
        }
    __syncthreads();

    // ---- phase 4: per cell — the >= iniTh list if it is not empty, else the weaker one (the reference's retry with
    // minThFAST, ORBextractor.cc:852-856); rank by (row, column) by counting; emit in cv::FAST's order -----------------
    for (int cc = warp; cc < ncell; cc += FAST_THREADS / 32) {          // one warp per cell, lanes = survivors
        const int nA = s_n[0][cc], nB = s_n[1][cc];
        if (nA + nB > listCap && lane == 0) atomicOr(P.d.status, OS1_ERR_CELL_OVERFLOW);
        const int which = nA > 0 ? 0 : 1;
        // with iniTh <= minTh the retry adds nothing: the weaker list is then empty (tmin = iniTh), so the rule above
        // still reproduces the reference
        const int n = min(which ? nB : nA, listCap);
        const unsigned* lst = lists + cc * listCap + (which ? listCap - n : 0);
        uint32_t* out = P.d.stage + (long long)frame * g.candPerFrame + L.candOff + (long long)(ci * L.nCols + j0 + cc) * L.cellCap;
        const int lim = which ? minS8 : 0;                                 // every entry of the strong list passes
        int emitted = 0;
        for (int i0 = 0; i0 < n; i0 += 32) {
            const int i = i0 + lane;
            const unsigned e = i < n ? lst[i] : 0u;
            const bool pass = i < n && (int)(e & 0xFF) >= lim;
            int rank = 0;
            for (int j = 0; j < n; j++) {
                const unsigned f = lst[j];                                  // broadcast read
                rank += (f < e) && ((int)(f & 0xFF) >= lim);
            }
            if (pass) {
                const int r = e >> 22, col = (e >> 8) & 0x3FFF;
                if (rank < L.cellCap)
                    out[rank] = OS1_PACK_XYS(col + xorg - OS1_BORDER, r + y0 - OS1_BORDER, (int)(e & 0xFF) + tmin - 1);
            }
            emitted += __popc(__ballot_sync(0xffffffffu, pass));
        }
        if (lane == 0) cellCnt[cc] = min(emitted, L.cellCap);
    }
}

} // namespace

template <int PW>
static cudaError_t launch_fast_pw(const FastParams& P, size_t smem, dim3 grid, cudaStream_t s)
{
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(fast_kernel<PW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    fast_kernel<PW><<<grid, FAST_THREADS, smem, s>>>(P);
    return cudaGetLastError();
}

cudaError_t os1_launch_fast(const Geometry& g, const DevPtrs& d, int nframes, cudaStream_t s)
{
    static_assert(sizeof(FastParams) < 4000, "kernel parameter block too large");
    if (g.groupsPerFrame == 0) return cudaSuccess;
    FastParams P;
    P.g = g; P.d = d;
    int maxW = 0, maxH = 0;
    P.listCap = 1;
    for (int l = 0; l < g.nlevels; l++) {
        if (g.lv[l].nCells == 0) continue;
        maxW = max(maxW, FAST_G * g.lv[l].wCell + 6);
        maxH = max(maxH, g.lv[l].hCell + 6);
        P.listCap = max(P.listCap, g.lv[l].cellCap);
    }
    const int need = ((maxW + 3 + 3) >> 2) + 2;        // + alignment slack + one spare word on each side
    const int PW = need <= 81 ? 81 : 135;              // odd pitches: rows start in different banks
    if (need > 135) return cudaErrorNotSupported;
    P.pitchW = PW;
    P.tileRows = maxH;
    size_t smem = (size_t)4 * (PW * (3 * P.tileRows + 6) + FAST_G * P.listCap) + 4 * PW + 16;
    if (smem > 200 * 1024) return cudaErrorNotSupported;
    dim3 grid(g.groupsPerFrame, nframes);
    return PW == 81 ? launch_fast_pw<81>(P, smem, grid, s) : launch_fast_pw<135>(P, smem, grid, s);
}
