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
//   phase 1  tile -> shared memory with 16-byte asynchronous copies from a 16-byte aligned origin;
//   phase 2  (after a cheap compass-point test that queues, per warp, the groups that can hold a corner at all)
//            a thread scores FOUR adjacent pixels from 21 word loads: each ring value of a pixel pair sits in the high
//            bytes of a u16x2 register (at most one byte-permute or shift, see ring_pair) and the min-of-9 / max-of-9
//            arcs are min/max trees on the packed VIMNMX.U16x2 path (2 pixels/instr, 36 per polarity);
//            the response is finished in packed form too (saturating differences as max(a,b)-b, no cross-lane
//            borrow) and stored as max(strength - tmin, 0): zero below min(iniTh, minTh), which leaves NMS
//            unchanged and makes the maps sparse;
//   phase 3  NMS as a column sweep: a thread owns one group of four columns and walks down a band of rows with the
//            horizontal neighbour maxima of three rows in registers (4 word loads + ~25 ALU instructions per row,
//            no divergence); cell boundaries are lane masks; survivors become bits of a per-row bitmap of the tile;
//   phase 4  one warp per cell: lanes = rows cut the cell's columns out of the row bitmaps and write the survivors in
//            row-major order (cv::FAST's output order) to a list; lanes = survivors vote "any survivor >= iniTh" and
//            emit at their ballot rank, so nothing is ever sorted.
#include "os1_internal.cuh"

#define FAST_THREADS 256
#ifndef FAST_MIN_CTAS
#define FAST_MIN_CTAS 4
#endif

namespace {

// n / d == (n * c_rcp.v[d]) >> 24 for d <= 256 while n * d < 2^24 (and n * v[d] < 2^32): every division of the kernel
// is by a small CTA-uniform number, and an integer division costs ~25 instructions per thread
struct RcpTab {
    unsigned v[257];
    constexpr RcpTab() : v() { for (int i = 1; i < 257; i++) v[i] = (1u << 24) / (unsigned)i + 1u; }
};
__constant__ RcpTab c_rcp = RcpTab();

struct FastParams {
    Geometry g;
    DevPtrs d;
    int pitchW;         // shared-memory row pitch in 32-bit words
    int tileRows;
    int bmWords;        // words per row of the survivor bitmap (one bit per tile byte column, + 2 spare)
    int listStride;     // u16 slots per warp for the survivor list of a cell (aliases the image tile)
    int qStride;        // u16 slots per warp for the queue of groups that pass the compass test
};

// ---- packed ring access -------------------------------------------------------------------------------------
// The three words (wm, w0, wp) of a tile row hold bytes -4..7 relative to the first of the thread's four pixels.
// ring_pair<DX, ODD> returns, for the pixel pair (ODD ? {1,3} : {0,2}), the bytes at horizontal offset DX in the HIGH
// bytes of the two u16 lanes; the low bytes hold whatever neighbours them in memory.  min/max of such lanes order by
// the high byte first, so a chain of them ends with the exact min/max of the pixel values in the high bytes — and the
// value needed is simply the 32-bit window at byte offset o = 3 + DX + ODD of the row: a register as it is (o = 0, 4),
// one shift (o = 3, 7; IMAD.SHL, not on the saturated ALU pipe) or one byte-permute (the rest), 18 permutes per four
// pixels instead of 32.
template <int DX, int ODD>
__device__ __forceinline__ unsigned ring_pair(unsigned wm, unsigned w0, unsigned wp)
{
    constexpr int o = 3 + DX + ODD;
    static_assert(o >= 0 && o <= 7, "ring offset outside the three words");
    if (o == 0) return wm;
    if (o == 4) return w0;
    if (o == 3) return w0 << 8;
    if (o == 7) return wp << 8;
    if (o < 3) return __byte_perm(wm, w0, ((o + 1) << 4) | ((o + 3) << 12));             // bytes o+1, o+3 of (wm, w0)
    return __byte_perm(w0, wp, ((o - 3) << 4) | ((o - 1) << 12));                          // bytes o-3, o-1 of (w0, wp)
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
    // Arcs 2i and 2i+1 share the eight ring values 2i+1 .. 2i+8 (four adjacent pairs), and
    //   max(min(W, r[2i]), min(W, r[2i+9])) == min(W, max(r[2i], r[2i+9])),
    // so each polarity takes 8 pair ops + 8 pair-of-pairs + 8 end ops + 8 min3/max3 + 4 to fold: 36 instead of the 40 of
    // a sliding window-3 / window-9 scheme.
    unsigned pl[8], ph[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        pl[j] = __vminu2(r[2 * j + 1], r[(2 * j + 2) & 15]);
        ph[j] = __vmaxu2(r[2 * j + 1], r[(2 * j + 2) & 15]);
    }
    unsigned ql[8], qh[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        ql[j] = __vminu2(pl[j], pl[(j + 1) & 7]);
        qh[j] = __vmaxu2(ph[j], ph[(j + 1) & 7]);
    }
    unsigned vd[8], vb[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const unsigned ehi = __vmaxu2(r[2 * i], r[(2 * i + 9) & 15]), elo = __vminu2(r[2 * i], r[(2 * i + 9) & 15]);
        vd[i] = __vimin3_u16x2(ql[i], ql[(i + 2) & 7], ehi);       // the better of arcs 2i, 2i+1: min over the arc
        vb[i] = __vimax3_u16x2(qh[i], qh[(i + 2) & 7], elo);       // ... max over the arc
    }
    unsigned d = __vimax3_u16x2(vd[0], vd[1], vd[2]), bb = __vimin3_u16x2(vb[0], vb[1], vb[2]);
    d = __vimax3_u16x2(d, vd[3], vd[4]); bb = __vimin3_u16x2(bb, vb[3], vb[4]);
    d = __vimax3_u16x2(d, vd[5], vd[6]); bb = __vimin3_u16x2(bb, vb[5], vb[6]);
    darkest = __vmaxu2(d, vd[7]);
    brightest = __vminu2(bb, vb[7]);
}

// per-lane saturating difference of u16x2 lanes without any cross-lane borrow
__device__ __forceinline__ unsigned usubsat2(unsigned a, unsigned b) { return __vmaxu2(a, b) - b; }     // max(a-b, 0)

template <int PW>
__global__ void __launch_bounds__(FAST_THREADS, FAST_MIN_CTAS) fast_kernel(const __grid_constant__ FastParams P)
{
    extern __shared__ __align__(16) unsigned smw[];
    const Geometry& g = P.g;
    const int frame = blockIdx.y;
    const int2 gt = __ldg(P.d.gtab + blockIdx.x);
    const int level = gt.x & 0xFF, ci = gt.x >> 8, j0 = gt.y;
    const LevelGeom& L = g.lv[level];
    const int ncell = min(FAST_G, L.nCols - j0);

    const int x0 = OS1_BORDER + j0 * L.wCell;
    const int x1 = min(OS1_BORDER + (j0 + ncell) * L.wCell + 6, L.maxBX);
    const int y0 = OS1_BORDER + ci * L.hCell;
    const int y1 = min(y0 + L.hCell + 6, L.maxBY);
    const int tw = x1 - x0, th = y1 - y0;
    constexpr int pw = PW;                 // compile-time pitch: every shared-memory access gets an immediate offset
    // shared memory: tile words; two score maps holding the even (px 0,2) and odd (px 1,3) pixels of every group
    // of four as u16x2; six mask rows; the survivor bitmap; the per-warp queues of phase 2 (the per-cell survivor lists of
    // phase 4 reuse the image tile)
    unsigned* img = smw;
    unsigned* scE = img + pw * P.tileRows;
    unsigned* scO = scE + pw * P.tileRows;
    unsigned* vmE = scO + pw * P.tileRows;      // lanes inside the tested columns
    unsigned* vmO = vmE + pw;
    unsigned* lmE = vmO + pw;                    // 0 in the lane whose pixel is a cell's FIRST tested column
    unsigned* lmO = lmE + pw;
    unsigned* rmE = lmO + pw;                    // 0 in the lane whose pixel is a cell's LAST tested column
    unsigned* rmO = rmE + pw;
    unsigned* bm = smw + ((pw * (3 * P.tileRows + 6) + 3) & ~3);      // [tileRows][bmWords] (16-byte aligned): bit c of a row <-> tile byte column c survived the NMS
    const int bw = P.bmWords;
    const int xorg = (x0 - 1) & ~15;            // global x of tile byte 0: 16-byte aligned, at least one spare word on the left
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

    // tile: asynchronous 16-byte copies straight into shared memory (LDGSTS), all in flight at once — issued first, so
    // the masks and the clearing below run while the bytes travel.  (Level rows start 16-byte aligned and have a pitch
    // that is a multiple of 16; a chunk that would end beyond the level's last column is fetched word by word, so that
    // nothing is ever read past the end of a frame the caller owns.)
    {
        const int nch = (words + 3) >> 2;                   // 16-byte chunks per tile row
        const unsigned rcpCh = c_rcp.v[nch];
        const int lw = L.w;
#pragma unroll 1
        for (int i = tid; i < th * nch; i += FAST_THREADS) {
            const int r = (int)(((unsigned)i * rcpCh) >> 24), c = i - r * nch;
            const uint8_t* gsrc = src + (long long)(y0 + r) * spitch + xorg + 16 * c;
            const unsigned sdst = (unsigned)__cvta_generic_to_shared(img + r * pw + 4 * c);
            if (xorg + 16 * c + 16 <= lw) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sdst), "l"(gsrc) : "memory");
            } else {
                for (int k = 0; k < 4; k++)
                    if (4 * c + k < words)
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sdst + 4 * k), "l"(gsrc + 4 * k) : "memory");
            }
        }
    }
    // ---- phase 0/1: lane masks, tile -> shared, cleared score-map frame and bitmap ------------------------------------
    const unsigned rcpW = c_rcp.v[wCell];
#pragma unroll 1
    for (int w = tid; w < pw; w += FAST_THREADS) {
        unsigned vm[4], lm[4], rm[4];
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int col = 4 * w + b;
            const bool in = col >= tc0 && col < tc1;
            const int rel = in ? col - tc0 : 1;
            const int q = rel - (int)(((unsigned)rel * rcpW) >> 24) * wCell;       // rel % wCell
            vm[b] = in ? 0xFF00u : 0u;
            lm[b] = (in && q == 0) ? 0u : 0xFFFFu;
            rm[b] = (in && (q == wCell - 1 || col == tc1 - 1)) ? 0u : 0xFFFFu;
        }
        vmE[w] = vm[0] | (vm[2] << 16); vmO[w] = vm[1] | (vm[3] << 16);
        lmE[w] = lm[0] | (lm[2] << 16); lmO[w] = lm[1] | (lm[3] << 16);
        rmE[w] = rm[0] | (rm[2] << 16); rmO[w] = rm[1] | (rm[3] << 16);
    }
#pragma unroll 1
    for (int i = tid; 4 * i < th * bw; i += FAST_THREADS) ((uint4*)bm)[i] = make_uint4(0, 0, 0, 0);
    const int g0 = tc0 >> 2, g1 = (tc1 + 3) >> 2;
    // both score maps start as zeros (16-byte stores): the groups the compass test rejects are never touched again, and the
    // NMS may read one word / one row beyond the scored area
    {
        uint4* z = (uint4*)scE;                       // scE and scO are adjacent; pw is a multiple of 4
        const int nz = (pw * P.tileRows + pw * th) >> 2;
#pragma unroll 1
        for (int i = tid; i < nz; i += FAST_THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncthreads();

    const int ng = g1 - g0, nItems = ng * (tr1 - tr0);
    const unsigned rcpNg = c_rcp.v[ng];                             // exact quotient for idx < 2^16 (nItems <= 136 * 64)
    // stored s8 = strength - tmin; response = strength - 1 = s8 + tmin - 1
    const int iniS8 = g.iniTh - tmin + 1, minS8 = g.minTh - tmin + 1;
    // ---- phase 2: scores, four pixels per thread, kept packed ---------------------------------------------------------
    // (rows x groups flattened so that every lane has work: 75 groups per row would leave a third of the lanes idle in
    //  the last of three 32-wide trips; the division by the group count is a multiply by its reciprocal)
    {
        const unsigned T = (unsigned)tmin * 0x01000100u;       // tmin << 8 in both lanes
        // 2a: a necessary condition from the four compass points of the ring.  An arc of 9 contains two ADJACENT compass
        // points, so  max over adjacent pairs of min(pair) >= darkest  and  min over pairs of max(pair) <= brightest:
        // if that bound does not exceed tmin, the group's four scores are zero.  Groups that pass are queued (per warp,
        // ballot-compacted) and scored in full by the same warp in 2b, all lanes busy.
        const unsigned T1 = (unsigned)((tmin + 1) * 256 - 1) * 0x00010001u;
        unsigned short* queue = (unsigned short*)(bm + ((th * bw + 3) & ~3)) + warp * P.qStride;
        int qn = 0;
        const unsigned ltmask = (1u << lane) - 1u;
#pragma unroll 1
        for (int idx0 = warp * 32; idx0 < nItems; idx0 += FAST_THREADS) {
            const int idx = idx0 + lane;
            bool pass = false;
            if (idx < nItems) {
                const int rr = (int)(((unsigned)idx * rcpNg) >> 24);
                const int r = tr0 + rr, gq = g0 + idx - rr * ng;
                const unsigned* row = img + r * pw + gq;
                const unsigned tm = row[-3 * pw - 1], t0 = row[-3 * pw], tp = row[-3 * pw + 1];      // ring 8: (0,-3)
                const unsigned cm = row[-1], c0 = row[0], cp = row[1];                                 // ring 4: (3,0), 12: (-3,0)
                const unsigned bmw_ = row[3 * pw - 1], b0 = row[3 * pw], bp = row[3 * pw + 1];       // ring 0: (0,3)
                const unsigned cE = __byte_perm(c0, 0, 0x2404), cO = __byte_perm(c0, 0, 0x3414);
                unsigned u[2];
#pragma unroll
                for (int odd = 0; odd < 2; odd++) {
                    const unsigned r0 = odd ? ring_pair<0, 1>(bmw_, b0, bp) : ring_pair<0, 0>(bmw_, b0, bp);
                    const unsigned r4 = odd ? ring_pair<3, 1>(cm, c0, cp) : ring_pair<3, 0>(cm, c0, cp);
                    const unsigned r8 = odd ? ring_pair<0, 1>(tm, t0, tp) : ring_pair<0, 0>(tm, t0, tp);
                    const unsigned r12 = odd ? ring_pair<-3, 1>(cm, c0, cp) : ring_pair<-3, 0>(cm, c0, cp);
                    const unsigned c = odd ? cO : cE;
                    // the four adjacent pairs are exactly {r0, r8} x {r4, r12}, so
                    //   max over the pairs of min(pair) == min(max(r0, r8), max(r4, r12))   (and dually for the other bound):
                    // three operations per bound instead of six
                    const unsigned lo = __vminu2(__vmaxu2(r0, r8), __vmaxu2(r4, r12));
                    const unsigned hi = __vmaxu2(__vminu2(r0, r8), __vminu2(r4, r12));
                    u[odd] = __vmaxu2(usubsat2(lo, c), usubsat2(c | 0x00FF00FFu, hi));
                }
                pass = __vmaxu2(__vmaxu2(u[0], u[1]), T1) != T1;          // some lane's bound exceeds tmin
            }
            const unsigned bal = __ballot_sync(0xffffffffu, pass);
            if (pass) queue[qn + __popc(bal & ltmask)] = (unsigned short)idx;
            qn += __popc(bal);
        }
        __syncwarp();
        // 2b: the full score of the queued groups
#pragma unroll 1
        for (int i = lane; i < qn; i += 32) {
                const int idx = queue[i];
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
                // centre pixels in the high bytes; low bytes 0x00 for `ring - centre`, 0xFF for `centre - ring`, so the
                // unknown low bytes of the ring extremes can neither borrow from nor carry into the high byte
                const unsigned cE = __byte_perm(W[3][1], 0, 0x2404), cO = __byte_perm(W[3][1], 0, 0x3414);
                const unsigned cEf = cE | 0x00FF00FFu, cOf = cO | 0x00FF00FFu;
                // high byte of a lane = max(darkest - c, c - brightest) = strength; stored = max(strength - tmin, 0) << 8,
                // i.e. non-zero exactly when the cv::FAST response (strength-1) reaches tmin — order preserving for NMS
                const unsigned sE = __vmaxu2(usubsat2(dE, cE), usubsat2(cEf, bE));
                const unsigned sO = __vmaxu2(usubsat2(dO, cO), usubsat2(cOf, bO));
                scE[r * pw + gq] = usubsat2(sE, T) & vmE[gq];        // (the masks also clear the low bytes)
                scO[r * pw + gq] = usubsat2(sO, T) & vmO[gq];
        }
    }
    __syncthreads();

    // ---- phase 3: NMS sweep.  Thread = (group of four columns, band of rows).  For a row, H3 = max(left, centre, right)
    // and H2 = max(left, right) of the two pixel pairs (E = pixels 0,2; O = pixels 1,3), with the neighbours that lie in
    // another cell masked out; a pixel survives when it beats H3 of the rows above and below and H2 of its own row. -----
    {
        const int nr = tr1 - tr0;
        const int nb = max(1, min(ng > 1 ? (int)(((unsigned)FAST_THREADS * rcpNg) >> 24) : FAST_THREADS, nr));     // bands of rows
        const int rpb = (int)(((unsigned)(nr + nb - 1) * c_rcp.v[nb]) >> 24);
        const int band = (int)(((unsigned)tid * rcpNg) >> 24);
        const int gq = g0 + tid - band * ng;
        const int rb = tr0 + band * rpb, re = min(tr1, rb + rpb);
        if (band < nb && rb < re) {
            const unsigned lE = lmE[gq], rE = rmE[gq], lO = lmO[gq], rO = rmO[gq];
            const unsigned ONES = 0x00010001u;
            const int sh = 4 * (gq & 7);
            unsigned* bmw = bm + (gq >> 3);
            unsigned h3Ep, h3Op, h2Ec, h3Ec, h2Oc, h3Oc, ec, oc;
            // horizontal maxima of one row
            auto hrow = [&](int r, unsigned& e, unsigned& o, unsigned& h2E, unsigned& h3E, unsigned& h2O, unsigned& h3O) {
                const unsigned* q = scE + r * pw + gq;
                const unsigned* t = scO + r * pw + gq;
                e = q[0]; o = t[0];
                const unsigned leftE = __byte_perm(t[-1], o, 0x5432) & lE;       // pixels -1, 1
                const unsigned rightE = o & rE;                                  // pixels  1, 3
                const unsigned leftO = e & lO;                                   // pixels  0, 2
                const unsigned rightO = __byte_perm(e, q[1], 0x5432) & rO;       // pixels  2, 4
                h2E = __vmaxu2(leftE, rightE); h3E = __vimax3_u16x2(leftE, e, rightE);
                h2O = __vmaxu2(leftO, rightO); h3O = __vimax3_u16x2(leftO, o, rightO);
            };
            {
                unsigned e, o, h2E, h2O;
                hrow(rb - 1, e, o, h2E, h3Ep, h2O, h3Op);
            }
            hrow(rb, ec, oc, h2Ec, h3Ec, h2Oc, h3Oc);
#pragma unroll 2
            for (int r = rb; r < re; r++) {
                unsigned en, on, h2En, h3En, h2On, h3On;
                hrow(r + 1, en, on, h2En, h3En, h2On, h3On);
                const unsigned mE = __vimax3_u16x2(h3Ep, h2Ec, h3En), mO = __vimax3_u16x2(h3Op, h2Oc, h3On);
                // lanes that beat all eight neighbours -> 1
                const unsigned nE = __vminu2(__vmaxu2(ec, mE) - mE, ONES), nO = __vminu2(__vmaxu2(oc, mO) - mO, ONES);
                const unsigned x = 2u * nO + nE;                                  // bits 0,1 <-> pixels 0,1; bits 16,17 <-> pixels 2,3
                const unsigned nib = (x | (x >> 14)) & 0xFu;
                if (nib) atomicOr(bmw + r * bw, nib << sh);
                h3Ep = h3Ec; h3Op = h3Oc;
                h2Ec = h2En; h3Ec = h3En; h2Oc = h2On; h3Oc = h3On; ec = en; oc = on;
            }
        }
    }
    __syncthreads();

    // ---- phase 4: per cell (one warp).  Pass A, lanes = rows: cut the cell's columns out of the row bitmaps and write
    // the survivors, in row-major order (cv::FAST's), to a list in the dead image tile.  Pass B, lanes = survivors: vote
    // "any survivor >= iniTh", keep those or else the ones >= minTh (the reference's retry with minThFAST,
    // ORBextractor.cc:852-856) and emit them at their rank among the kept ones. -------------------------------------
#pragma unroll 1
    for (int cc = warp; cc < ncell; cc += FAST_THREADS / 32) {
        const int cx0 = tc0 + cc * wCell, cx1 = min(cx0 + wCell, tc1);      // tile byte columns of the cell (at most 64)
        const int wsel = cx0 >> 5, bsh = cx0 & 31, width = cx1 - cx0;
        const unsigned keepLo = width >= 32 ? 0xFFFFFFFFu : ((1u << width) - 1u);
        const unsigned keepHi = width >= 64 ? 0xFFFFFFFFu : (width > 32 ? ((1u << (width - 32)) - 1u) : 0u);
        uint32_t* out = P.d.stage + (long long)frame * g.candPerFrame + L.candOff + (long long)(ci * L.nCols + j0 + cc) * L.cellCap;
        unsigned short* list = (unsigned short*)img + warp * P.listStride;      // entries: row << 6 | column inside the cell
        int n = 0;
#pragma unroll 1
        for (int r0 = tr0; r0 < tr1; r0 += 32) {                                // (cells taller than 32 tested rows: two trips)
            const int r = r0 + lane;
            unsigned lo = 0, hi = 0;
            if (r < tr1) {
                const unsigned* w = bm + r * bw + wsel;
                const unsigned w0 = w[0], w1 = w[1], w2 = w[2];
                lo = __funnelshift_r(w0, w1, bsh) & keepLo;
                hi = __funnelshift_r(w1, w2, bsh) & keepHi;
            }
            const int cnt = __popc(lo) + __popc(hi);
            int incl = cnt;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += t;
            }
            unsigned short* dst = list + n + incl - cnt;
            const unsigned tag = (unsigned)r << 6;
            while (lo) { *dst++ = (unsigned short)(tag | (unsigned)(__ffs(lo) - 1)); lo &= lo - 1; }
            while (hi) { *dst++ = (unsigned short)(tag | (unsigned)(__ffs(hi) + 31)); hi &= hi - 1; }
            n += __shfl_sync(0xffffffffu, incl, 31);
        }
        __syncwarp();
        // score byte (strength - tmin) of list entry e
        auto score8 = [&](unsigned e) -> int {
            const int r = e >> 6, c = cx0 + (int)(e & 63u);
            const unsigned v = ((c & 1) ? scO : scE)[r * pw + (c >> 2)];
            return (int)(((c & 2) ? (v >> 24) : (v >> 8)) & 0xFFu);
        };
        unsigned e = 0; int s8 = 0;
        bool st = false;
#pragma unroll 1
        for (int i = lane; i < n; i += 32) { e = list[i]; s8 = score8(e); st = st || s8 >= iniS8; }
        const int lim = __any_sync(0xffffffffu, st) ? iniS8 : minS8;
        int base = 0;
#pragma unroll 1
        for (int i0 = 0; i0 < n; i0 += 32) {
            const int i = i0 + lane;
            if (n > 32 && i < n) { e = list[i]; s8 = score8(e); }       // up to 32 survivors: the vote's registers are this trip's
            const bool pass = i < n && s8 >= lim;
            const unsigned bal = __ballot_sync(0xffffffffu, pass);
            const int rank = base + __popc(bal & ((1u << lane) - 1u));
            if (pass && rank < L.cellCap)
                out[rank] = OS1_PACK_XYS(cx0 + (int)(e & 63u) + xorg - OS1_BORDER, (int)(e >> 6) + y0 - OS1_BORDER, s8 + tmin - 1);
            base += __popc(bal);
        }
        if (lane == 0) cellCnt[cc] = min(base, L.cellCap);
    }
}

} // namespace

template <int PW>
static cudaError_t launch_fast_pw(const FastParams& P, size_t smem, dim3 grid, cudaStream_t s)
{
    if (smem > 48 * 1024) {
        cudaError_t e = os1_optin_smem((const void*)fast_kernel<PW>);
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
    int maxW = 0, maxH = 0, maxCap = 1;
    for (int l = 0; l < g.nlevels; l++) {
        if (g.lv[l].nCells == 0) continue;
        maxW = max(maxW, FAST_G * g.lv[l].wCell + 6);
        maxH = max(maxH, g.lv[l].hCell + 6);
        maxCap = max(maxCap, g.lv[l].cellCap);
        if (g.lv[l].wCell > 64) return cudaErrorNotSupported;       // a cell's columns must fit the 64-bit row masks of phase 4
    }
    // widest tile: x0 - xorg <= 16 (tile byte 0 sits on a 16-byte boundary), one more word for the last group's right
    // neighbour, rounded up to whole 16-byte chunks
    const int need = (((((maxW + 16 + 3) >> 2) + 1) + 3) >> 2) << 2;
    const int PW = need <= 72 ? 72 : (need <= 88 ? 88 : 136);      // multiples of 4: tile rows are 16-byte aligned in shared memory
    if (need > 136) return cudaErrorNotSupported;
    P.pitchW = PW;
    P.tileRows = maxH;
    P.bmWords = (PW + 7) / 8 + 2;
    P.listStride = (maxCap + 1) & ~1;
    {
        int maxItems = 1;
        for (int l = 0; l < g.nlevels; l++)
            if (g.lv[l].nCells) maxItems = max(maxItems, (((FAST_G * g.lv[l].wCell + 6 + 16) >> 2) + 1) * g.lv[l].hCell);
        P.qStride = ((maxItems + FAST_THREADS - 1) / FAST_THREADS) * 32;       // a warp sees at most every 8th trip of 32 items
    }
    if ((size_t)2 * P.listStride * (FAST_THREADS / 32) > (size_t)4 * PW * P.tileRows) return cudaErrorNotSupported;
    size_t smem = (size_t)4 * (((PW * (3 * P.tileRows + 6) + 3) & ~3) + ((P.tileRows * P.bmWords + 3) & ~3)) + (size_t)2 * P.qStride * (FAST_THREADS / 32) + 16;
    if (smem > 200 * 1024) return cudaErrorNotSupported;
    dim3 grid(g.groupsPerFrame, nframes);
    return PW == 72 ? launch_fast_pw<72>(P, smem, grid, s) : (PW == 88 ? launch_fast_pw<88>(P, smem, grid, s) : launch_fast_pw<136>(P, smem, grid, s));
}
