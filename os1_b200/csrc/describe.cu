// K4/K5 — orientation + blurred rBRIEF descriptor + final keypoint assembly, one warp per kept keypoint.
// Replaces, for every kept keypoint:
//   IC_Angle / computeOrientation          src/ORBextractor.cc:86-113, :504-511   (un-blurred level image)
//   cv::fastAtan2                          :112                                    
//   GaussianBlur(7x7, sigma 2, REFLECT_101) :949-950                               (Appendix A.2; only the 37x37
//                                          window the descriptor can touch is blurred — the fixed-point blur is
//                                          exact per pixel, so blurring a window == cropping the blurred level)
//   computeOrbDescriptor / bit_pattern_31_ :132-171, :182-440
//   border fix-up, octave, size, scaling   :879-889, :959-967   and the level-major concatenation of :936-967
// The warp stages the 43x43 source window in shared memory once; orientation, both blur passes and the 512
// rotated taps all read it from there.  No FMA contraction anywhere in the float path (x86 reference build has
// none); cos/sin are evaluated in double and rounded, which is what glibc's cosf/sinf deliver.
#include "os1_internal.cuh"
#include <cmath>
#include <cfloat>

#define DESC_WARPS 4
#define WIN 43            // 2*(18+3)+1 : rotated pattern reach 18 (|p| <= 18.38) + blur radius 3
#define WIN_R 21
#define RAW_PW 13         // window row pitch in words (52 bytes >= 3 + 43, odd)
#define HB_W 37           // blurred window side
#define HB_Q 10           // groups of four blurred columns
#define BL_PITCH 40

namespace {

__device__ __align__(16) const int8_t d_pattern[1024] = {
#include "orb_pattern.inc"
};

struct DescParams {
    Geometry g;
    DevPtrs d;
    os1_kp_dev* kp;
    uint8_t* desc;
    int* counts;
    int capPerFrame;
    int umax[16];
    float p1, p3, p5, p7, eps, factorPI;
};

struct WarpSmem {
    unsigned raw[WIN * RAW_PW];         // source window as fetched: window column c sits at byte c + shift;
                                        // after the horizontal pass the same words hold the blurred 37x37 window (bytes)
    uint2 hb[WIN * HB_Q];               // horizontal pass, per row and group of four: {(H0,H2), (H1,H3)} as u16x2
};

__device__ __forceinline__ int reflect101(int p, int len)
{
    if (p < 0) p = -p;
    if (p >= len) p = 2 * len - 2 - p;
    return p;
}

__device__ __forceinline__ float fast_atan2_deg(float y, float x, const DescParams& P)
{
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, P.eps));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(P.p7, c2), P.p5), c2), P.p3), c2), P.p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, P.eps));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(P.p7, c2), P.p5), c2), P.p3), c2), P.p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

// one output of the vertical 7-tap pass: (sum k_j * H_j + 32768) >> 16 with k = {18,34,48,56,48,34,18}
__device__ __forceinline__ unsigned vtap(int r0, int r1, int r2, int r3, int r4, int r5, int r6)
{
    return (unsigned)((18 * (r0 + r6) + 34 * (r1 + r5) + 48 * (r2 + r4) + 56 * r3 + 32768) >> 16);
}

__global__ void __launch_bounds__(DESC_WARPS * 32, 9) describe_kernel(const __grid_constant__ DescParams P)
{
    __shared__ WarpSmem sm[DESC_WARPS];
    const Geometry& g = P.g;
    const int frame = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot = blockIdx.x * DESC_WARPS + warp;

    // level of this slot from the per-level kept counts (:936-967 concatenates levels in order)
    const uint32_t* lvCount = P.d.lvCount + frame * g.nlevels;
    int level = -1, j = slot, total = 0;
    for (int l = 0; l < g.nlevels; l++) {
        const int c = (int)lvCount[l];
        if (level < 0 && slot - total < c) { level = l; j = slot - total; }
        total += c;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        P.counts[frame] = min(total, P.capPerFrame);
        if (total > P.capPerFrame) atomicOr(P.d.status, OS1_ERR_KP_OVERFLOW);
    }
    if (level < 0 || slot >= P.capPerFrame) return;

    const LevelGeom& L = g.lv[level];
    const uint32_t v = P.d.lvKp[(long long)frame * g.kpCapPerFrame + L.kpOff + j];
    const int x = OS1_X(v) + OS1_BORDER, y = OS1_Y(v) + OS1_BORDER, score = OS1_S(v);
    const uint8_t* img;
    int pitch;
    if (level == 0) { img = P.d.lvl0 + (long long)frame * P.d.lvl0FrameStride; pitch = P.d.lvl0Pitch; }
    else { img = P.d.slab + (long long)frame * g.slabBytes + L.off; pitch = L.pitch; }

    WarpSmem& S = sm[warp];
    // ---- 43x43 window -> shared.  Interior keypoints: aligned 32-bit loads of the rows (12 words cover the 43 bytes
    // wherever they start); the window keeps the byte phase `sh` of its first column.  Keypoints within 21 px of the
    // level's edge take the byte path with REFLECT_101 indices, as the blur of the borderless clone sees them. ----------
    const int xs = x - WIN_R, ys = y - WIN_R;
    int sh;
    if (xs >= 0 && ys >= 0 && x + WIN_R < L.w && y + WIN_R < L.h) {
        sh = xs & 3;
        const uint8_t* base = img + (long long)ys * pitch + (xs - sh);
        for (int it = lane; it < WIN * 12; it += 32) {
            const int r = (it * 5462) >> 16;            // it / 12
            const int w = it - r * 12;
            if (4 * w <= sh + WIN - 1)                  // never touch a word that holds no window byte (row end)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(S.raw + r * RAW_PW + w)),
                             "l"(base + (long long)r * pitch + 4 * w) : "memory");      // all 17 copies of a lane in flight at once
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
    } else {
        sh = 0;
        uint8_t* rb = (uint8_t*)S.raw;
        for (int i = lane; i < WIN * WIN; i += 32) {
            const int r = i / WIN, c = i - r * WIN;
            const int yy = reflect101(ys + r, L.h), xx = reflect101(xs + c, L.w);
            rb[r * (RAW_PW * 4) + c] = img[(long long)yy * pitch + xx];
        }
    }
    __syncwarp();

    // ---- intensity centroid (lanes = columns u, loop over rows v) ----------------------------------------------
    int m10 = 0, m01 = 0;
    {
        const int u = lane - 15;
        if (lane < 31) {
            const int au = u < 0 ? -u : u;
            const uint8_t* col = (const uint8_t*)S.raw + WIN_R + u + sh;
#pragma unroll 1
            for (int vv = -15; vv <= 15; vv++) {
                const int av = vv < 0 ? -vv : vv;
                if (au <= P.umax[av]) {
                    const int val = col[(WIN_R + vv) * (RAW_PW * 4)];
                    m10 += u * val;
                    m01 += vv * val;
                }
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, o);
            m01 += __shfl_xor_sync(0xffffffffu, m01, o);
        }
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10, P);

    // ---- 7x7 Gaussian, Q8 taps {18,34,48,56,48,34,18}, exact two-pass fixed point ----------------------------------
    // Horizontal pass, four outputs per item on packed u16x2 arithmetic (an output is <= 65280, so two of them share
    // a 32-bit multiply-add without carries): pair E = outputs (0,2) of the group, pair O = outputs (1,3);
    // tap i of E is the byte pair (i, i+2) of the aligned 12-byte block, tap i of O is byte pair (i+1, i+3).
    {
        const int s8 = 8 * sh;
        for (int it = lane; it < WIN * HB_Q; it += 32) {
            const int r = (it * 6554) >> 16;            // it / 10
            const int q = it - r * HB_Q;
            const unsigned* w = S.raw + r * RAW_PW + q;
            const unsigned W0 = w[0], W1 = w[1], W2 = w[2], W3 = w[3];
            const unsigned A0 = __funnelshift_r(W0, W1, s8), A1 = __funnelshift_r(W1, W2, s8), A2 = __funnelshift_r(W2, W3, s8);
            const unsigned M = 0x00FF00FFu;
            const unsigned p0 = A0 & M, p1 = (A0 >> 8) & M, p2 = __funnelshift_r(A0, A1, 16) & M, p3 = __funnelshift_r(A0, A1, 24) & M;
            const unsigned p4 = A1 & M, p5 = (A1 >> 8) & M, p6 = __funnelshift_r(A1, A2, 16) & M, p7 = __funnelshift_r(A1, A2, 24) & M;
            uint2 o;
            o.x = 18u * (p0 + p6) + 34u * (p1 + p5) + 48u * (p2 + p4) + 56u * p3;
            o.y = 18u * (p1 + p7) + 34u * (p2 + p6) + 48u * (p3 + p5) + 56u * p4;
            S.hb[r * HB_Q + q] = o;
        }
    }
    __syncwarp();
    unsigned* bl = S.raw;               // the raw window is dead from here on (37 * 10 words <= 43 * 13)
    // Vertical pass: lane = (group of four columns, third of the rows); a 7-row window of unpacked sums slides down.
    if (lane < 3 * HB_Q) {
        const int seg = lane / HB_Q, q = lane - seg * HB_Q;
        const int rb = seg == 0 ? 0 : (seg == 1 ? 13 : 25), re = seg == 0 ? 13 : (seg == 1 ? 25 : 37);
        int wnd[7][4];                                  // circular: step t keeps rows rb+t .. rb+t+6 in slots (t+j) % 7
#pragma unroll
        for (int k = 0; k < 6; k++) {
            const uint2 hh = S.hb[(rb + k) * HB_Q + q];
            wnd[k][0] = hh.x & 0xFFFF; wnd[k][2] = hh.x >> 16; wnd[k][1] = hh.y & 0xFFFF; wnd[k][3] = hh.y >> 16;
        }
#pragma unroll
        for (int t = 0; t < 13; t++) {
            const int r = rb + t;
            if (r < re) {
                const uint2 hh = S.hb[(r + 6) * HB_Q + q];
                const int n = (t + 6) % 7;
                wnd[n][0] = hh.x & 0xFFFF; wnd[n][2] = hh.x >> 16; wnd[n][1] = hh.y & 0xFFFF; wnd[n][3] = hh.y >> 16;
                unsigned o = 0;
#pragma unroll
                for (int c = 0; c < 4; c++)
                    o |= vtap(wnd[t % 7][c], wnd[(t + 1) % 7][c], wnd[(t + 2) % 7][c], wnd[(t + 3) % 7][c], wnd[(t + 4) % 7][c],
                              wnd[(t + 5) % 7][c], wnd[(t + 6) % 7][c]) << (8 * c);
                bl[r * (BL_PITCH / 4) + q] = o;
            }
        }
    }
    __syncwarp();

    // ---- rotated BRIEF: lane = descriptor byte ---------------------------------------------------------------------------
    const float ang = __fmul_rn(angle, P.factorPI);
    const float a = (float)cos((double)ang), b = (float)sin((double)ang);
    const int4* pat = (const int4*)d_pattern;
    int4 w0 = __ldg(pat + 2 * lane), w1 = __ldg(pat + 2 * lane + 1);
    const int words[8] = { w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w };
    const uint8_t* center = (const uint8_t*)bl + 18 * BL_PITCH + 18;
    int val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int w = words[k];
        const float x0 = (float)(int)(int8_t)(w & 0xFF), y0 = (float)(int)(int8_t)((w >> 8) & 0xFF);
        const float x1 = (float)(int)(int8_t)((w >> 16) & 0xFF), y1 = (float)(int)(int8_t)((w >> 24) & 0xFF);
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, b), __fmul_rn(y0, a)));
        const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, a), __fmul_rn(y0, b)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, b), __fmul_rn(y1, a)));
        const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, a), __fmul_rn(y1, b)));
        const int t0 = center[r0 * BL_PITCH + c0], t1 = center[r1 * BL_PITCH + c1];
        val |= (t0 < t1) << k;
    }
    const long long o = (long long)frame * P.capPerFrame + slot;
    P.desc[o * 32 + lane] = (uint8_t)val;
    if (lane == 0) {
        os1_kp_dev k;
        k.x = (float)x; k.y = (float)y;
        if (level != 0) { k.x = __fmul_rn(k.x, L.scale); k.y = __fmul_rn(k.y, L.scale); }
        k.size = L.kpSize; k.angle = angle; k.response = (float)score; k.octave = level; k.class_id = -1;
        P.kp[o] = k;
    }
}

} // namespace

cudaError_t os1_launch_describe(const Geometry& g, const DevPtrs& d, int nframes, os1_kp_dev* kp, uint8_t* desc, int capPerFrame,
                                int* counts, cudaStream_t s)
{
    static_assert(sizeof(DescParams) < 4000, "kernel parameter block too large");
    DescParams P;
    P.g = g; P.d = d; P.kp = kp; P.desc = desc; P.counts = counts; P.capPerFrame = capPerFrame;
    // umax exactly as the reference builds it (ORBextractor.cc:486-501)
    {
        const int HP = 15;
        int v, v0, vmax = (int)floor(HP * sqrt(2.f) / 2 + 1);
        int vmin = (int)ceil(HP * sqrt(2.f) / 2);
        const double hp2 = HP * HP;
        for (v = 0; v <= vmax; ++v) P.umax[v] = (int)lrint(sqrt(hp2 - v * v));
        for (v = HP, v0 = 0; v >= vmin; --v) {
            while (P.umax[v0] == P.umax[v0 + 1]) ++v0;
            P.umax[v] = v0;
            ++v0;
        }
    }
    volatile float R = (float)(180.0 / 3.1415926535897932384626433832795);
    volatile float t;
    t = 0.9997878412794807f * R; P.p1 = t;
    t = -0.3258083974640975f * R; P.p3 = t;
    t = 0.1555786518463281f * R; P.p5 = t;
    t = -0.04432655554792128f * R; P.p7 = t;
    P.eps = (float)DBL_EPSILON;
    P.factorPI = (float)(3.1415926535897932384626433832795 / 180.f);
    dim3 grid((min(capPerFrame, g.kpCapPerFrame) + DESC_WARPS - 1) / DESC_WARPS, nframes);
    describe_kernel<<<grid, DESC_WARPS * 32, 0, s>>>(P);
    return cudaGetLastError();
}
