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
#define RAW_PITCH 48
#define HB_W 37           // blurred window side
#define HB_PITCH 38
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
    uint8_t raw[WIN * RAW_PITCH];
    uint16_t hb[WIN * HB_PITCH];
    uint8_t bl[HB_W * BL_PITCH];
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

__global__ void __launch_bounds__(DESC_WARPS * 32) describe_kernel(const __grid_constant__ DescParams P)
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
    // ---- window -> shared (REFLECT_101 at the level's edges, as the blur of the borderless clone sees it) ----
    for (int i = lane; i < WIN * WIN; i += 32) {
        const int r = i / WIN, c = i - r * WIN;
        const int yy = reflect101(y - WIN_R + r, L.h), xx = reflect101(x - WIN_R + c, L.w);
        S.raw[r * RAW_PITCH + c] = img[(long long)yy * pitch + xx];
    }
    __syncwarp();

    // ---- intensity centroid (lanes = columns u, loop over rows v) ----------------------------------------------
    int m10 = 0, m01 = 0;
    {
        const int u = lane - 15;
        if (lane < 31) {
            const int au = u < 0 ? -u : u;
#pragma unroll 1
            for (int vv = -15; vv <= 15; vv++) {
                const int av = vv < 0 ? -vv : vv;
                if (au <= P.umax[av]) {
                    const int val = S.raw[(WIN_R + vv) * RAW_PITCH + WIN_R + u];
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
    for (int i = lane; i < WIN * HB_W; i += 32) {
        const int r = i / HB_W, c = i - r * HB_W;
        const uint8_t* q = S.raw + r * RAW_PITCH + c;       // output column c <-> raw columns c..c+6
        S.hb[r * HB_PITCH + c] = (uint16_t)(18 * (q[0] + q[6]) + 34 * (q[1] + q[5]) + 48 * (q[2] + q[4]) + 56 * q[3]);
    }
    __syncwarp();
    for (int i = lane; i < HB_W * HB_W; i += 32) {
        const int r = i / HB_W, c = i - r * HB_W;
        const uint16_t* q = S.hb + r * HB_PITCH + c;
        const int s = 18 * (q[0] + q[6 * HB_PITCH]) + 34 * (q[HB_PITCH] + q[5 * HB_PITCH]) + 48 * (q[2 * HB_PITCH] + q[4 * HB_PITCH]) +
                      56 * q[3 * HB_PITCH];
        S.bl[r * BL_PITCH + c] = (uint8_t)((s + 32768) >> 16);
    }
    __syncwarp();

    // ---- rotated BRIEF: lane = descriptor byte ---------------------------------------------------------------------------
    const float ang = __fmul_rn(angle, P.factorPI);
    const float a = (float)cos((double)ang), b = (float)sin((double)ang);
    const int4* pat = (const int4*)d_pattern;
    int4 w0 = __ldg(pat + 2 * lane), w1 = __ldg(pat + 2 * lane + 1);
    const int words[8] = { w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w };
    const uint8_t* center = S.bl + 18 * BL_PITCH + 18;
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
