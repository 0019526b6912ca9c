// K1 — fused scale-pyramid build (replaces ORBextractor::ComputePyramid, src/ORBextractor.cc:971-996).
//
// One launch builds levels 1..L-1 of every frame of the batch.  A CTA owns one tile of the COARSEST level and
// the chain of source footprints below it: it pulls its level-0 footprint into shared memory with one TMA
// tile load (cp.async.bulk.tensor, SASS UTMALDG), then resizes level by level inside shared memory with
// cv::resize's INTER_LINEAR fixed-point arithmetic (11-bit coefficients) and stores
// only the pixels it owns.  Level 0 is read once from HBM and every other level is written once; no level is
// ever re-read from global memory.  The 19-px REFLECT_101 border of the reference (:986,:991) is not produced:
// no later stage reads it.
//
// Ownership: tile column tx owns [b_l[tx], b_l[tx+1]) at level l, where b_{L-1}[i] = i*TW and
// b_{l-1}[i] = first source tap of b_l[i].  The computed (shared-memory) range of a level is the owned range
// plus the source footprint of the next level's computed range, i.e. a halo of at most ~2 px per level.
#include "os1_internal.cuh"
#include <vector>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <cstdlib>

namespace {

struct PyrParams {
    int nlevels;
    int ntx, nty;
    int boxBytes;
    int barOff;
    const uint8_t* dbgLvl0; long long dbgStride; int dbgPitch, dbgW, dbgH;
    const int4* xt;              // [level][ntx]  {c0, c1, own0, own1}
    const int4* yt;              // [level][nty]
    const int4* xg;              // [level][ntx]  thread-mapping constants {ngroups, rcp, nbands, rcp}
    const int4* qtab;            // per level, per group of four destination columns: 2 x int4 (see QuadTaps)
    const int4* ptab;            // per level, per group: 1 x int4 (see PairTaps), first entry of a level = qOff / 2
    int qOff[OS1_MAX_LEVELS];    // first entry (in int4 units) of a level in qtab
    const int4* ytab;            // per level, per destination row: {source row 0, source row 1, c0, c1}
    int yOff[OS1_MAX_LEVELS];
    uint8_t* slab;
    long long slabBytes;
    int smemOff[OS1_MAX_LEVELS];
    int smemPitch[OS1_MAX_LEVELS];
    long long lvOff[OS1_MAX_LEVELS];
    int lvPitch[OS1_MAX_LEVELS];
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t done;
    do {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int x, int y, int z)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_u32(dst)),
        "l"(map), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
        : "memory");
}

// One resized pixel: cv::resize's fixed-point bilinear from the two horizontally
// interpolated values of the source rows (already >> 4).
__device__ __forceinline__ unsigned vresize(int h0, int h1, int b0, int b1)
{
    return (unsigned)((((b0 * h0) >> 16) + ((b1 * h1) >> 16) + 2) >> 2);
}

template <bool USE_TMA>
__global__ void __launch_bounds__(256) pyramid_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ PyrParams p)
{
    extern __shared__ __align__(128) uint8_t sm_raw[];
    // TMA wants a 128-byte aligned destination: align by hand (the launcher allocates 128 spare bytes) and keep
    // the mbarrier in the same dynamic block
    // (pointer arithmetic, not an integer round trip, so the compiler keeps addressing it as shared memory)
    uint8_t* sm = sm_raw + ((128u - ((unsigned)__cvta_generic_to_shared(sm_raw) & 127u)) & 127u);
    uint64_t& bar = *(uint64_t*)(sm + p.barOff);
    const int tx = blockIdx.x, ty = blockIdx.y, frame = blockIdx.z;
    const int tid = threadIdx.y * 32 + threadIdx.x;

    int4 PX = p.xt[tx], PY = p.yt[ty];      // level-0 ranges
    int srcX0 = PX.x, srcY0 = PY.x;         // global coordinates of byte (0,0) of the source level's tile
    if (USE_TMA) {
        if (tid == 0) {
            mbar_init(&bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            mbar_expect_tx(&bar, (uint32_t)p.boxBytes);
            tma_load_3d(sm + p.smemOff[0], &tmap, &bar, PX.x, PY.x, frame);
        }
        __syncthreads();
        mbar_wait(&bar, 0);
    } else {
        // debug path (OS1_DEBUG_NOTMA=1): plain loads of the same box
        const uint8_t* g0 = p.dbgLvl0 + (long long)frame * p.dbgStride;
        const int bw = p.smemPitch[0], bh = p.boxBytes / p.smemPitch[0];
        for (int i = tid; i < bw * bh; i += 256) {
            int r = i / bw, c = i - r * bw;
            int gx = PX.x + c, gy = PY.x + r;
            sm[p.smemOff[0] + i] = (gx < p.dbgW && gy < p.dbgH) ? g0[(long long)gy * p.dbgPitch + gx] : 0;
        }
        __syncthreads();
    }

    uint8_t* slab = p.slab + (long long)frame * p.slabBytes;
    for (int l = 1; l < p.nlevels; l++) {
        const int4 X = p.xt[l * p.ntx + tx], Y = p.yt[l * p.nty + ty];
        const int gx0 = X.x & ~3;                                   // tile byte 0 of this level <-> global x = gx0
        const int ngroups = ((X.y + 3) >> 2) - (gx0 >> 2);           // groups of four destination columns
        const int rows = Y.y - Y.x;
        const uint8_t* src = sm + p.smemOff[l - 1];
        uint32_t* dst = (uint32_t*)(sm + p.smemOff[l]);
        const int sp = p.smemPitch[l - 1], dpw = p.smemPitch[l] >> 2;
        const int4* ry = p.ytab + p.yOff[l];
        uint8_t* g = slab + p.lvOff[l];
        const int gp = p.lvPitch[l];
        if (ngroups > 0 && rows > 0) {
            // a thread owns four destination columns and walks a band of rows, so the horizontal taps live in
            // registers and the horizontally interpolated source row is reused when two destination rows share it
            // (divisions by per-tile constants through host-made reciprocals: exact for operands < 2^16)
            const int4 M = p.xg[l * p.ntx + tx];                    // {ngroups, 2^24/ngroups+1, nbands, 2^24/nbands+1}
            const int nbands = M.z;
            const int band = (int)(((unsigned)tid * (unsigned)M.y) >> 24), grp = tid - band * ngroups;
            const int rpb = (int)(((unsigned)(rows + nbands - 1) * (unsigned)M.w) >> 24);
            const int yb = band * rpb, ye = min(rows, yb + rpb);
            if (band < nbands && yb < ye) {
                // horizontal taps of the four columns, prepared on the host: the four tap bytes of a pixel pair lie in two
                // consecutive source words (PairTaps), one byte-permute lines them up as (p0,p1) byte pairs and one dp2a per
                // pixel does p0*c0 + p1*c1 (QuadTaps' coefficients)
                const int gq = (gx0 >> 2) + grp;                 // global group index
                const int4 t1 = __ldg(p.qtab + p.qOff[l] + 2 * gq + 1), pt = __ldg(p.ptab + (p.qOff[l] >> 1) + gq);
                // first source words of the two pairs, relative to the tile (columns left of the computed range may point
                // before it: their results are discarded, so any in-bounds word will do)
                const int w01 = max(pt.x - (srcX0 >> 2), 0), w23 = max(pt.y - (srcX0 >> 2), 0);
                const unsigned sel01 = (unsigned)pt.z & 0xFFFFu, sel23 = (unsigned)pt.z >> 16;
                const unsigned A0 = t1.x, A1 = t1.y, A2 = t1.z, A3 = t1.w;       // c0 | c1 << 16 per column
                const int gxb = gx0 + 4 * grp;
                const bool fullX = gxb >= X.z && gxb + 4 <= X.w;
                const int spw = sp >> 2;
                const uint32_t* srcw = (const uint32_t*)src - srcY0 * spw;           // word 0 of global source row 0
                auto hrow = [&](int srow, unsigned (&h)[4]) {
                    const uint32_t* r = srcw + srow * spw;
                    const unsigned B01 = __byte_perm(r[w01], r[w01 + 1], sel01);
                    const unsigned B23 = __byte_perm(r[w23], r[w23 + 1], sel23);
                    // kept as H >> 4, what cv::resize's vertical pass multiplies
                    h[0] = __dp2a_lo(A0, B01, 0u) >> 4;
                    h[1] = __dp2a_hi(A1, B01, 0u) >> 4;
                    h[2] = __dp2a_lo(A2, B23, 0u) >> 4;
                    h[3] = __dp2a_hi(A3, B23, 0u) >> 4;
                };
                unsigned h0[4], h1[4] = { 0, 0, 0, 0 };
                int prev1 = -1;
                const int4* ryp = ry + Y.x + yb;
                uint32_t* dp = dst + yb * dpw + grp;
                uint8_t* grow = g + (long long)(Y.x + yb) * gp + gxb;
                // rows of the band this tile owns (the others are halo for the next level): [o0, o1)
                const int o0 = max(Y.z - Y.x, yb), o1 = min(Y.w - Y.x, ye);
                for (int ly = yb; ly < ye; ly++, ryp++, dp += dpw, grow += gp) {
                    const int4 ey = __ldg(ryp);
                    if (ey.x == prev1) {
#pragma unroll
                        for (int i = 0; i < 4; i++) h0[i] = h1[i];
                    } else hrow(ey.x, h0);
                    if (ey.y != ey.x) hrow(ey.y, h1);
                    else {
#pragma unroll
                        for (int i = 0; i < 4; i++) h1[i] = h0[i];
                    }
                    prev1 = ey.y;
                    const unsigned b0 = (unsigned)ey.z, b1 = (unsigned)ey.w;
                    // ((b0*(H0>>4))>>16) + ((b1*(H1>>4))>>16) + 2) >> 2: full 32-bit products (< 2^26; a multiply-high runs at half
                    // the rate of a multiply here), their upper halves picked out two pixels at a time by a byte-permute, one
                    // add per two pixels in 16-bit lanes (sums <= 1022: no carry between the lanes), rounding included
                    const unsigned ev = __byte_perm(b0 * h0[0], b0 * h0[2], 0x7632) + __byte_perm(b1 * h1[0], b1 * h1[2], 0x7632) + 0x00020002u;
                    const unsigned od = __byte_perm(b0 * h0[1], b0 * h0[3], 0x7632) + __byte_perm(b1 * h1[1], b1 * h1[3], 0x7632) + 0x00020002u;
                    const unsigned word = ((ev >> 2) & 0x00FF00FFu) | ((od << 6) & 0xFF00FF00u);
                    *dp = word;
                    if (ly >= o0 && ly < o1) {
                        if (fullX) *(uint32_t*)grow = word;
                        else {
#pragma unroll
                            for (int i = 0; i < 4; i++)
                                if (gxb + i >= X.z && gxb + i < X.w) grow[i] = (uint8_t)(word >> (8 * i));
                        }
                    }
                }
            }
        }
        srcX0 = gx0; srcY0 = Y.x;
        __syncthreads();
    }
}

// ---- batches: one launch per level, HBM-streaming ---------------------------------------------------------------------
// The fused chain above reads level 0 once and keeps everything else on chip, which is what a single frame wants (one launch, no
// round trips).  For a batch the bytes were never the problem — a frame's pyramid is 10.6 MB of traffic even unfused, 1.6 us at
// HBM speed — the instructions were: halo pixels recomputed at every level, upper levels of a tile too small for 256 threads, lanes
// of one warp sitting in different row bands and diverging on the row-reuse logic (33 thread-instructions per output pixel).
// Here level l is one launch over all frames: a CTA owns a 128 x PL_TH tile of the DESTINATION level in TWO consecutive frames and
// pulls the source footprints of both in with one TMA tile load (box depth 2).  A warp is a band of 16 rows and a lane a group of
// four adjacent columns IN BOTH FRAMES — the two frames have the same geometry, so a lane's taps, selectors and coefficients
// serve eight pixels, every lane of a warp walks the same rows (the horizontally filtered source row is reused without any
// divergence), the per-row bookkeeping is paid once per eight pixels, and every output row of a frame's tile is one 128-byte
// store.  No pixel is computed twice; the source halo re-read is 4-5 %.
#define PL_TW 128
#define PL_TH 64
#define PL_WARPS 4
struct LvlParams {
    int level;
    int W, H;                    // destination level
    int nframes;
    int boxW, boxH;              // TMA box (source footprint of a tile, inner extent a multiple of 16 bytes)
    const int4* qtab;            // this level's QuadTaps (2 x int4 per group of four destination columns): the coefficients
    const int4* ptab;            // this level's PairTaps (1 x int4 per group): source words and byte selectors
    const int4* ytab;            // this level's vertical taps per destination row
    uint8_t* dst; long long dstFrameStride; int dstPitch;
    const uint8_t* dbgSrc; long long dbgStride; int dbgPitch, dbgW, dbgH;
};

template <bool USE_TMA>
__global__ void __launch_bounds__(32 * PL_WARPS, 7) pyramid_level_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ LvlParams p)
{
    extern __shared__ __align__(128) uint8_t sm_raw[];
    uint8_t* sm = sm_raw + ((128u - ((unsigned)__cvta_generic_to_shared(sm_raw) & 127u)) & 127u);
    __shared__ __align__(8) uint64_t bar;
    const int lane = threadIdx.x, warp = threadIdx.y, tid = warp * 32 + lane;
    const int frame0 = 2 * blockIdx.z;
    const int X0 = blockIdx.x * PL_TW, Y0 = blockIdx.y * PL_TH;
    const int Y1 = min(Y0 + PL_TH, p.H);
    // source origin of the tile: the first tap of its first column, aligned down to 16 bytes (TMA), and of its first row
    const int4 tq0 = __ldg(p.qtab + 2 * (X0 >> 2));
    const int srcXw = tq0.x & ~3;                          // word index
    const int srcY0 = __ldg(p.ytab + Y0).x;
    const int plane = p.boxW * p.boxH;
    if (USE_TMA) {
        if (tid == 0) {
            mbar_init(&bar, 1);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            mbar_expect_tx(&bar, (uint32_t)(2 * plane));                 // (a frame beyond the batch arrives as zeros, and counts)
            tma_load_3d(sm, &tmap, &bar, srcXw * 4, srcY0, frame0);
        }
    } else {
        for (int i = tid; i < 2 * plane; i += 32 * PL_WARPS) {
            const int f = i / plane, j = i - f * plane;
            const int r = j / p.boxW, c = j - r * p.boxW;
            const int gx = srcXw * 4 + c, gy = srcY0 + r;
            sm[i] = (frame0 + f < p.nframes && gx < p.dbgW && gy < p.dbgH) ? p.dbgSrc[(long long)(frame0 + f) * p.dbgStride + (long long)gy * p.dbgPitch + gx] : 0;
        }
    }
    // this thread's four columns (two pixel pairs) in BOTH frames: coefficients and tap positions from the host-made tables
    const int gq = (X0 >> 2) + lane;
    const bool active = 4 * gq < p.W;
    const int4 ta1 = active ? __ldg(p.qtab + 2 * gq + 1) : make_int4(0, 0, 0, 0);
    const int4 pa = active ? __ldg(p.ptab + gq) : make_int4(srcXw, srcXw, 0, 0);
    const int spw = p.boxW >> 2, planeW = plane >> 2;
    // row band of this warp
    constexpr int rpw = PL_TH / PL_WARPS;
    const int yb = Y0 + warp * rpw, ye = min(yb + rpw, Y1);
    __syncthreads();
    if (USE_TMA) mbar_wait(&bar, 0);
    if (!active || yb >= ye) return;
    const unsigned sel01 = (unsigned)pa.z & 0xFFFFu, sel23 = (unsigned)pa.z >> 16;
    const int4* ryp = p.ytab + yb;
    int4 ey = __ldg(ryp), eyNext = __ldg(ryp + 1);       // (the same address in every lane; the table has two spare entries at its end)
    int srow = ey.x, left = ye - yb;
    // the two pairs' first source words in the current source row of the first frame (the second frame's plane follows the
    // first's; the pointers step down a row after each filtered row)
    const uint32_t* rowp = (const uint32_t*)sm + (srow - srcY0) * spw - srcXw;
    const uint32_t *p01 = rowp + pa.x, *p23 = rowp + pa.y;
    auto hrow = [&](unsigned (&h)[8]) {
        const unsigned a0 = p01[0], a1 = p01[1], a2 = p23[0], a3 = p23[1];
        const unsigned b0 = p01[planeW], b1 = p01[planeW + 1], b2 = p23[planeW], b3 = p23[planeW + 1];
        p01 += spw; p23 += spw;
        const unsigned A01 = __byte_perm(a0, a1, sel01), A23 = __byte_perm(a2, a3, sel23);
        const unsigned B01 = __byte_perm(b0, b1, sel01), B23 = __byte_perm(b2, b3, sel23);
        // kept as H >> 4, what cv::resize's vertical pass multiplies
        h[0] = __dp2a_lo((unsigned)ta1.x, A01, 0u) >> 4;
        h[1] = __dp2a_hi((unsigned)ta1.y, A01, 0u) >> 4;
        h[2] = __dp2a_lo((unsigned)ta1.z, A23, 0u) >> 4;
        h[3] = __dp2a_hi((unsigned)ta1.w, A23, 0u) >> 4;
        h[4] = __dp2a_lo((unsigned)ta1.x, B01, 0u) >> 4;
        h[5] = __dp2a_hi((unsigned)ta1.y, B01, 0u) >> 4;
        h[6] = __dp2a_lo((unsigned)ta1.z, B23, 0u) >> 4;
        h[7] = __dp2a_hi((unsigned)ta1.w, B23, 0u) >> 4;
    };
    const int ncols = min(4, p.W - 4 * gq);
    const bool second = frame0 + 1 < p.nframes;           // (a batch with an odd number of frames: the last CTA layer has one)
    uint8_t* grow = p.dst + (long long)frame0 * p.dstFrameStride + (long long)yb * p.dstPitch + 4 * gq;
    // one destination row (of both frames) from two filtered source rows: ((c0*(H0>>4))>>16) + ((c1*(H1>>4))>>16) + 2) >> 2.
    // A multiply-high (IMAD.HI) runs at a quarter of the lane rate on this part, a 32-bit multiply at half of it: the products
    // (< 2^26) are taken in full and their upper halves, which are what is wanted, picked out two pixels at a time by a
    // byte-permute.  The four sums of a word are then shifted and merged in their 16-bit lanes
    auto emit = [&](const unsigned (&top)[8], const unsigned (&bot)[8], unsigned c0, unsigned c1) {
        unsigned word[2];
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const unsigned* t = top + 4 * k;
            const unsigned* b = bot + 4 * k;
            // upper halves of the products of pixels (0, 2) and of pixels (1, 3) side by side in 16-bit lanes: one add per
            // two pixels gives their sums (<= 1022: no carry between the lanes), rounding included
            const unsigned ev = __byte_perm(c0 * t[0], c0 * t[2], 0x7632) + __byte_perm(c1 * b[0], c1 * b[2], 0x7632) + 0x00020002u;
            const unsigned od = __byte_perm(c0 * t[1], c0 * t[3], 0x7632) + __byte_perm(c1 * b[1], c1 * b[3], 0x7632) + 0x00020002u;
            word[k] = ((ev >> 2) & 0x00FF00FFu) | ((od << 6) & 0xFF00FF00u);
        }
        uint8_t* grow1 = grow + p.dstFrameStride;
        if (ncols == 4) {
            *(uint32_t*)grow = word[0];
            if (second) *(uint32_t*)grow1 = word[1];
        } else {
#pragma unroll
            for (int i = 0; i < 3; i++)
                if (i < ncols) {
                    grow[i] = (uint8_t)(word[0] >> (8 * i));
                    if (second) grow1[i] = (uint8_t)(word[1] >> (8 * i));
                }
        }
        grow += p.dstPitch;
    };
    // The loop steps through the SOURCE rows of the band, filtering each once into one of two register sets used alternately
    // (the loop is unrolled by two, so which set holds which row is static and nothing is ever copied); a destination row is
    // emitted at the step that filters its second source row — its first is the row of the step before (the plan checks
    // that the second source rows of a level increase strictly), or the same row where cv::resize clamps at the last row.
    unsigned hA[8], hB[8];
    const int srowEnd = srcY0 + p.boxH;                  // (never reached: the band's last row is emitted inside the box; a bound all the same)
#pragma unroll 1
    while (srow < srowEnd) {
        hrow(hA);
        if (srow == ey.y) {
            if (ey.x == ey.y) emit(hA, hA, (unsigned)ey.z, (unsigned)ey.w);
            else emit(hB, hA, (unsigned)ey.z, (unsigned)ey.w);
            if (--left == 0) break;
            ey = eyNext; ryp++; eyNext = __ldg(ryp + 1);
        }
        srow++;
        hrow(hB);
        if (srow == ey.y) {
            if (ey.x == ey.y) emit(hB, hB, (unsigned)ey.z, (unsigned)ey.w);
            else emit(hA, hB, (unsigned)ey.z, (unsigned)ey.w);
            if (--left == 0) break;
            ey = eyNext; ryp++; eyNext = __ldg(ryp + 1);
        }
        srow++;
    }
}

} // namespace

struct PyramidPlan {
    int TW, TH, ntx, nty;
    int4* d_xt = nullptr;
    int4* d_yt = nullptr;
    int4* d_xg = nullptr;
    int4* d_qtab = nullptr;
    int4* d_ptab = nullptr;    // pair taps of the per-level kernel: one int4 per group of four columns
    bool pairsFit = true;      // the four tap bytes of every pixel pair lie in two consecutive source words
    int4* d_ytab = nullptr;
    bool tapsFit = true;       // every group of four columns finds its eight taps in three consecutive source words
    cudaError_t cudaErr = cudaSuccess;      // first failure of a table allocation / upload
    PyrParams prm;
    int boxW, boxH;
    size_t smemBytes;
    // per-level launches (batches): TMA box of a destination tile's source footprint, per level
    bool lvOk = false;
    int lvBoxW[OS1_MAX_LEVELS], lvBoxH[OS1_MAX_LEVELS];
    CUresult (*encode)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                       const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
};

// host copy of the taps of one axis for level l as destination (mirrors what api.cu uploads into rtab)
void os1_host_resize_taps(int ssize, int dsize, bool vertical, std::vector<short>& s0, std::vector<short>& s1, std::vector<short>& c0,
                          std::vector<short>& c1);

static void build_axis(const Geometry& g, bool vertical, int T, int& nt, std::vector<int4>& tab, std::vector<int>& maxC)
{
    const int L = g.nlevels, top = L - 1;
    auto size = [&](int l) { return vertical ? g.lv[l].h : g.lv[l].w; };
    nt = (size(top) + T - 1) / T;
    std::vector<std::vector<short> > S0(L), S1(L);
    for (int l = 1; l < L; l++) {
        std::vector<short> c0, c1;
        os1_host_resize_taps(size(l - 1), size(l), vertical, S0[l], S1[l], c0, c1);
    }
    std::vector<std::vector<int> > b(L, std::vector<int>(nt + 1));
    for (int i = 0; i <= nt; i++) b[top][i] = std::min(i * T, size(top));
    for (int l = top; l >= 1; l--)
        for (int i = 0; i <= nt; i++) {
            int v = (i == 0) ? 0 : (i == nt || b[l][i] >= size(l)) ? size(l - 1) : (int)S0[l][b[l][i]];
            if (i > 0) v = std::max(v, b[l - 1][i - 1]);
            b[l - 1][i] = v;
        }
    tab.assign((size_t)L * nt, make_int4(0, 0, 0, 0));
    maxC.assign(L, 0);
    for (int t = 0; t < nt; t++) {
        int c0 = b[top][t], c1 = b[top][t + 1];
        for (int l = top; l >= 0; l--) {
            // TMA tile loads fault ("illegal instruction") unless the innermost start coordinate is a multiple
            // of 16 bytes (measured on B200, CUDA 12.9): the level-0 footprint starts on a 16-px boundary
            if (l == 0 && !vertical) c0 &= ~15;
            tab[(size_t)l * nt + t] = make_int4(c0, c1, b[l][t], b[l][t + 1]);
            maxC[l] = std::max(maxC[l], c1 - c0);
            if (l == 0) break;
            int n0 = b[l - 1][t], n1 = b[l - 1][t + 1];
            if (c1 > c0) {
                // (x axis: threads work on groups of four columns, and a group's source words are anchored at its
                //  FIRST column's tap, so the footprint starts at the tap of the group that contains c0)
                n0 = std::min(n0, (int)S0[l][vertical ? c0 : (c0 & ~3)]);
                n1 = std::max(n1, (int)S1[l][c1 - 1] + 1);
            }
            c0 = n0; c1 = n1;
        }
    }
}

PyramidPlan* os1_pyramid_plan_create(const Geometry& g, int tileW, int tileH)
{
    PyramidPlan* P = new PyramidPlan();
    P->TW = tileW; P->TH = tileH;
    std::vector<int4> xt, yt;
    std::vector<int> mw, mh;
    build_axis(g, false, tileW, P->ntx, xt, mw);
    build_axis(g, true, tileH, P->nty, yt, mh);
    auto upload = [&](int4*& dptr, const std::vector<int4>& v) {
        cudaError_t e = cudaMalloc(&dptr, v.size() * sizeof(int4));
        if (e == cudaSuccess) e = cudaMemcpy(dptr, v.data(), v.size() * sizeof(int4), cudaMemcpyHostToDevice);
        if (e != cudaSuccess && P->cudaErr == cudaSuccess) { P->cudaErr = e; cudaGetLastError(); }
    };
    upload(P->d_xt, xt);
    upload(P->d_yt, yt);
    std::vector<int4> xg(xt.size());
    for (size_t i = 0; i < xt.size(); i++) {
        const int ng = std::max(1, ((xt[i].y + 3) >> 2) - ((xt[i].x & ~3) >> 2));
        const int nb = std::max(1, 256 / ng);
        xg[i] = make_int4(ng, (1 << 24) / ng + 1, nb, (1 << 24) / nb + 1);
    }
    upload(P->d_xg, xg);
    PyrParams& p = P->prm;
    memset(&p, 0, sizeof(p));
    p.nlevels = g.nlevels; p.ntx = P->ntx; p.nty = P->nty;
    p.xt = P->d_xt; p.yt = P->d_yt; p.xg = P->d_xg;
    {
        // QuadTaps: for destination columns 4q..4q+3 of level l — int4 #0 = {first source word (global index),
        // selLo01 | selHi01 << 16, selLo23 | selHi23 << 16, 0}, int4 #1 = {c0 | c1 << 16 for the four columns}.
        // Tap byte o (0..11, relative to the first word) is routed by two byte-permutes: the first picks from words
        // 0-1, the second substitutes bytes that live in word 2.
        std::vector<int4> q, pt;
        for (int l = 1; l < g.nlevels; l++) {
            p.qOff[l] = (int)q.size();
            std::vector<short> s0, s1, c0, c1;
            os1_host_resize_taps(g.lv[l - 1].w, g.lv[l].w, false, s0, s1, c0, c1);
            const int W = g.lv[l].w, nq = (W + 3) / 4;
            for (int qi = 0; qi < nq; qi++) {
                int a[8]; unsigned A[4];
                for (int i = 0; i < 4; i++) {
                    const int x = std::min(4 * qi + i, W - 1);
                    a[2 * i] = s0[x]; a[2 * i + 1] = s1[x];
                    A[i] = (unsigned)(unsigned short)c0[x] | ((unsigned)(unsigned short)c1[x] << 16);
                }
                const int wbase = a[0] >> 2;
                unsigned sel[2][2] = { { 0, 0 }, { 0, 0 } };
                for (int k = 0; k < 8; k++) {
                    const int o = a[k] - 4 * wbase;
                    if (o < 0 || o > 11) { P->tapsFit = false; continue; }
                    const int word = k >> 2, byte = k & 3;
                    sel[word][0] |= (unsigned)(o < 8 ? o : 0) << (4 * byte);
                    sel[word][1] |= (unsigned)(o < 8 ? byte : 4 + (o - 8)) << (4 * byte);
                }
                q.push_back(make_int4(wbase, (int)(sel[0][0] | (sel[0][1] << 16)), (int)(sel[1][0] | (sel[1][1] << 16)), 0));
                q.push_back(make_int4((int)A[0], (int)A[1], (int)A[2], (int)A[3]));
                // PairTaps (per-level kernel): the four tap bytes of a pixel pair span at most four source bytes, i.e. two
                // consecutive words: {first word of pair 0-1, first word of pair 2-3, selector 01 | selector 23 << 16, 0} — one
                // byte-permute per pair instead of two
                int pw[2]; unsigned ps[2] = { 0, 0 };
                for (int j = 0; j < 2; j++) {
                    pw[j] = std::min(std::min(a[4 * j], a[4 * j + 1]), std::min(a[4 * j + 2], a[4 * j + 3])) >> 2;
                    for (int k = 0; k < 4; k++) {
                        const int o = a[4 * j + k] - 4 * pw[j];
                        if (o < 0 || o > 7) { P->pairsFit = false; continue; }
                        ps[j] |= (unsigned)o << (4 * k);
                    }
                }
                pt.push_back(make_int4(pw[0], pw[1], (int)(ps[0] | (ps[1] << 16)), 0));
            }
        }
        // vertical taps per destination row
        std::vector<int4> yt4;
        for (int l = 1; l < g.nlevels; l++) {
            p.yOff[l] = (int)yt4.size();
            std::vector<short> s0, s1, c0, c1;
            os1_host_resize_taps(g.lv[l - 1].h, g.lv[l].h, true, s0, s1, c0, c1);
            for (int y = 0; y < g.lv[l].h; y++) yt4.push_back(make_int4(s0[y], s1[y], (int)c0[y], (int)c1[y]));
        }
        yt4.push_back(make_int4(0, 0, 0, 0));           // spare entries: pyramid_level_kernel reads two rows ahead
        yt4.push_back(make_int4(0, 0, 0, 0));
        upload(P->d_ytab, yt4);
        p.ytab = P->d_ytab;
        if (q.empty()) q.push_back(make_int4(0, 0, 0, 0));
        upload(P->d_qtab, q);
        p.qtab = P->d_qtab;
        if (pt.empty()) pt.push_back(make_int4(0, 0, 0, 0));
        upload(P->d_ptab, pt);
        p.ptab = P->d_ptab;
    }
    P->boxW = (mw[0] + 15) & ~15;
    P->boxH = mh[0];
    p.boxBytes = P->boxW * P->boxH;
    // two shared-memory buffers, used alternately: level l+1 is written over level l-1, which nobody reads any more
    // (the block-wide barrier after level l separates the two) — the tile chain costs size(L0) + size(L1), not the sum of
    // all levels, and more CTAs fit an SM
    size_t bufSize[2] = { 0, 0 };
    for (int l = 0; l < g.nlevels; l++) {
        p.smemPitch[l] = (l == 0) ? P->boxW : ((mw[l] + 3 + 3) & ~3);     // + up to 3 bytes: column 0 sits on a 4-px boundary
        bufSize[l & 1] = std::max(bufSize[l & 1], ((size_t)p.smemPitch[l] * mh[l] + 127) & ~(size_t)127);
        p.lvOff[l] = g.lv[l].off; p.lvPitch[l] = g.lv[l].pitch;
    }
    for (int l = 0; l < g.nlevels; l++) p.smemOff[l] = (l & 1) ? (int)bufSize[0] : 0;
    size_t off = bufSize[0] + bufSize[1];
    p.slabBytes = g.slabBytes;
    p.barOff = (int)off;
    P->smemBytes = off + 16 + 128;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess || !fn) {
        // leave encode null; the launcher reports the failure
        fn = nullptr;
    }
    P->encode = (decltype(P->encode))fn;
    // per-level launches: the largest source footprint of a PL_TW x PL_TH destination tile, per level
    P->lvOk = g.nlevels >= 2;
    for (int l = 1; l < g.nlevels; l++) {
        std::vector<short> s0, s1, c0, c1, t0, t1;
        os1_host_resize_taps(g.lv[l - 1].w, g.lv[l].w, false, s0, s1, c0, c1);
        os1_host_resize_taps(g.lv[l - 1].h, g.lv[l].h, true, t0, t1, c0, c1);
        int bw = 16, bh = 1;
        for (int X0 = 0; X0 < g.lv[l].w; X0 += PL_TW) {
            const int X1 = std::min(X0 + PL_TW, g.lv[l].w);
            const int first = ((int)s0[X0] >> 2 & ~3) * 4;                  // as the kernel aligns it
            int last = 0;
            for (int x = X0; x < X1; x++) last = std::max(last, (int)s1[x]);
            // a group's three source words start at its first tap's word: the last group may read up to 11 bytes past its first tap
            const int lastGroupFirst = ((int)s0[(X1 - 1) & ~3] >> 2) * 4;
            last = std::max(last, lastGroupFirst + 11);
            bw = std::max(bw, last + 1 - first);
        }
        for (int y = 1; y < g.lv[l].h; y++)
            if (t1[y] <= t1[y - 1]) P->lvOk = false;      // (the kernel emits at most one destination row per source row)
        for (int Y0 = 0; Y0 < g.lv[l].h; Y0 += PL_TH) {
            const int Y1 = std::min(Y0 + PL_TH, g.lv[l].h);
            int last = 0;
            for (int y = Y0; y < Y1; y++) last = std::max(last, (int)t1[y]);
            bh = std::max(bh, last + 1 - (int)t0[Y0]);
        }
        bw = (bw + 15) & ~15;
        P->lvBoxW[l] = bw; P->lvBoxH[l] = bh;
        if (bw > 256 || bh > 256 || 2 * (size_t)bw * bh + 256 > 200 * 1024) P->lvOk = false;
    }
    if (!P->tapsFit || !P->pairsFit || !P->encode || P->cudaErr != cudaSuccess) P->lvOk = false;
    return P;
}

bool os1_pyramid_plan_ok(const PyramidPlan* P)
{
    return P && P->cudaErr == cudaSuccess && P->encode && P->tapsFit && P->pairsFit && P->boxW <= 256 && P->boxH <= 256 && P->smemBytes <= 200 * 1024;
}

void os1_pyramid_plan_destroy(PyramidPlan* P)
{
    if (!P) return;
    cudaFree(P->d_xt); cudaFree(P->d_yt); cudaFree(P->d_xg); cudaFree(P->d_qtab); cudaFree(P->d_ptab); cudaFree(P->d_ytab);
    delete P;
}

// batches: levels 1..L-1 one launch each (see pyramid_level_kernel)
static cudaError_t launch_levels(PyramidPlan* P, const Geometry& g, const DevPtrs& d, int nframes, cudaStream_t s)
{
    static const bool noTma = getenv("OS1_DEBUG_NOTMA") != nullptr;
    for (int l = 1; l < g.nlevels; l++) {
        const LevelGeom& S = g.lv[l - 1];
        const LevelGeom& D = g.lv[l];
        const uint8_t* src = l == 1 ? d.lvl0 : d.slab + S.off;
        const long long srcStride = l == 1 ? d.lvl0FrameStride : g.slabBytes;
        const int srcPitch = l == 1 ? d.lvl0Pitch : S.pitch;
        CUtensorMap tmap;
        cuuint64_t dims[3] = { (cuuint64_t)S.w, (cuuint64_t)S.h, (cuuint64_t)nframes };
        cuuint64_t strides[2] = { (cuuint64_t)srcPitch, (cuuint64_t)srcStride };
        cuuint32_t box[3] = { (cuuint32_t)P->lvBoxW[l], (cuuint32_t)P->lvBoxH[l], 2 };           // two frames per CTA
        cuuint32_t estr[3] = { 1, 1, 1 };
        CUresult r = P->encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void*)src, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
        LvlParams p;
        p.level = l; p.W = D.w; p.H = D.h; p.nframes = nframes; p.boxW = P->lvBoxW[l]; p.boxH = P->lvBoxH[l];
        p.qtab = P->d_qtab + P->prm.qOff[l]; p.ptab = P->d_ptab + P->prm.qOff[l] / 2; p.ytab = P->d_ytab + P->prm.yOff[l];
        p.dst = d.slab + D.off; p.dstFrameStride = g.slabBytes; p.dstPitch = D.pitch;
        p.dbgSrc = src; p.dbgStride = srcStride; p.dbgPitch = srcPitch; p.dbgW = S.w; p.dbgH = S.h;
        const size_t smem = 2 * (size_t)p.boxW * p.boxH + 128;
        if (smem > 48 * 1024) {
            cudaError_t e = os1_optin_smem(noTma ? (const void*)pyramid_level_kernel<false> : (const void*)pyramid_level_kernel<true>);
            if (e != cudaSuccess) return e;
        }
        dim3 grid((D.w + PL_TW - 1) / PL_TW, (D.h + PL_TH - 1) / PL_TH, (nframes + 1) / 2), block(32, PL_WARPS);
        if (noTma) pyramid_level_kernel<false><<<grid, block, smem, s>>>(tmap, p);
        else pyramid_level_kernel<true><<<grid, block, smem, s>>>(tmap, p);
    }
    return cudaGetLastError();
}

bool os1_pyramid_plan_levels_ok(const PyramidPlan* P) { return P && P->lvOk; }

// Frames from which a batch takes the per-level launches.  Seven launches cost about 60 us whatever the batch (launch latency,
// the tail of each small grid), the fused chain's single launch nothing: measured cross-over at about 32 M pixels per call (16 frames
// of 1920x1080, 35 of 1280x720, 100 of 640x480; scripts/pyr_crossover.py).  OS1_PYR_LEVELS_MIN=<frames> overrides (tests, experiments).
static int levels_min_frames(const Geometry& g)
{
    const char* env = getenv("OS1_PYR_LEVELS_MIN");          // (read at every call: the tests switch paths inside one process)
    const int envMin = env ? atoi(env) : 0;
    if (envMin > 0) return envMin;
    return (int)((32000000LL + (long long)g.rows * g.cols - 1) / ((long long)g.rows * g.cols));
}

int os1_pyramid_launch_count(const PyramidPlan* P, const Geometry& g, int nframes)
{
    if (g.nlevels < 2) return 0;
    return (P && P->lvOk && nframes >= levels_min_frames(g)) ? g.nlevels - 1 : 1;
}

cudaError_t os1_launch_pyramid(PyramidPlan* P, const Geometry& g, const DevPtrs& d, int nframes, cudaStream_t s)
{
    if (g.nlevels < 2) return cudaSuccess;
    if (P && P->lvOk && nframes >= levels_min_frames(g)) return launch_levels(P, g, d, nframes, s);
    if (!os1_pyramid_plan_ok(P)) return cudaErrorNotSupported;
    CUtensorMap tmap;
    cuuint64_t dims[3] = { (cuuint64_t)g.cols, (cuuint64_t)g.rows, (cuuint64_t)nframes };
    cuuint64_t strides[2] = { (cuuint64_t)d.lvl0Pitch, (cuuint64_t)d.lvl0FrameStride };
    cuuint32_t box[3] = { (cuuint32_t)P->boxW, (cuuint32_t)P->boxH, 1 };
    cuuint32_t estr[3] = { 1, 1, 1 };
    CUresult r = P->encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (void*)d.lvl0, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
    PyrParams p = P->prm;
    p.slab = d.slab;
    dim3 grid(P->ntx, P->nty, nframes), block(32, 8);
    p.dbgLvl0 = d.lvl0; p.dbgStride = d.lvl0FrameStride; p.dbgPitch = d.lvl0Pitch; p.dbgW = g.cols; p.dbgH = g.rows;
    static const bool noTma = getenv("OS1_DEBUG_NOTMA") != nullptr;
    if (P->smemBytes > 48 * 1024) {
        cudaError_t e = os1_optin_smem(noTma ? (const void*)pyramid_kernel<false> : (const void*)pyramid_kernel<true>);
        if (e != cudaSuccess) return e;
    }
    if (noTma) pyramid_kernel<false><<<grid, block, P->smemBytes, s>>>(tmap, p);
    else pyramid_kernel<true><<<grid, block, P->smemBytes, s>>>(tmap, p);
    return cudaGetLastError();
}
