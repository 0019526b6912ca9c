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
// none); cos/sin follow glibc's cosf/sinf algorithm (see libm_sincosf).
#include "os1_internal.cuh"
#include <mutex>
#include <cmath>
#include <cfloat>

#define DESC_WARPS 4
#define WIN 43            // 2*(18+3)+1 : rotated pattern reach 18 (|p| <= 18.38) + blur radius 3
#define WIN_R 21
#define RAW_PW 20         // window row pitch in words: four 16-byte chunks hold the 43 bytes wherever they start, and a fifth
                          // of padding spreads the rows over the banks (pitch 16 would put every row pair on the same ones)
#define HB_W 37           // blurred window side
#define HB_Q 10           // groups of four blurred columns
#define HB_PAIRS 22       // pairs of window rows (43 rows; the last pair's second row is padding)
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
    uint4 raw4[HB_PAIRS * HB_Q];        // source window as fetched (43 rows of 80 bytes: window column c sits at byte c + shift,
                                        // shift < 16), overwritten IN PLACE by the horizontal pass: a pair of rows (2e, 2e+1) is 160
                                        // bytes, and so are its ten groups of four u16x2 words {H[2e][j] | H[2e+1][j] << 16}
    unsigned bl[HB_W * (BL_PITCH / 4)]; // the blurred 37x37 window (bytes, 40-byte rows)
};
static_assert(HB_PAIRS * HB_Q * 16 >= WIN * RAW_PW * 4 && 2 * RAW_PW * 4 == HB_Q * 16, "a row pair and its horizontal-pass output must overlay exactly");

// Intensity-centroid coefficients, one row of the 31x31 patch per lane: for |v| = 0..15 and the nine aligned words that
// hold patch columns 4..39 of the window, the signed column offsets u (0 outside the circle) and the 0/1 membership
// mask (ORBextractor.cc:86-113 with umax of :486-501).  Filled by the launcher once per device.
struct IcTab { int cu[16][9]; unsigned cm[16][9]; };
__device__ IcTab d_ic;

__device__ __forceinline__ int dp4a_us(unsigned a, int b, int c)
{
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

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

// cosf / sinf exactly as the reference's C library evaluates them (ORBextractor.cc:136-137 calls cos/sin on a float,
// i.e. cosf/sinf).  glibc >= 2.28 (sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, sincosf.h — the ARM optimized-routines
// implementation) is accurate to 0.56 ULP, NOT correctly rounded: it differs from round(cos((double)x)) for about one
// float in a thousand, and a 1-ULP difference in a or b flips one of the 512 rounded tap coordinates of a keypoint about
// once in 10^5 keypoints.  So the algorithm itself is restated: reduction by pi/2 with n = (int)(x * 2/pi * 2^24 + 2^23)
// >> 24, then the degree-7 sine or degree-8 cosine polynomial in double, no fused multiply-adds.  Verified on the host
// (same constants, same order of operations) against libm's cosf / sinf on 155 million floats of [0, 6.5): no difference.
// Valid for 0 <= x < 120; the angles here are in [0, 2 pi].
__device__ __forceinline__ float libm_sincosf_poly(double x, double x2, bool flip, int n)
{
    const double C0 = 0x1p0, C1 = -0x1.ffffffd0c621cp-2, C2 = 0x1.55553e1068f19p-5, C3 = -0x1.6c087e89a359dp-10, C4 = 0x1.99343027bf8c3p-16;
    const double S1 = -0x1.555545995a603p-3, S2 = 0x1.1107605230bc4p-7, S3 = -0x1.994eb3774cf24p-13;
    if ((n & 1) == 0) {
        const double x3 = __dmul_rn(x, x2);
        const double s1 = __dadd_rn(S2, __dmul_rn(x2, S3));
        const double x7 = __dmul_rn(x3, x2);
        const double s = __dadd_rn(x, __dmul_rn(x3, S1));
        return (float)__dadd_rn(s, __dmul_rn(x7, s1));
    }
    // (the second table of the reference has c0..c4 negated: every term, and so the sum, changes sign exactly)
    const double x4 = __dmul_rn(x2, x2);
    const double c2 = __dadd_rn(C3, __dmul_rn(x2, C4));
    const double c1 = __dadd_rn(C0, __dmul_rn(x2, C1));
    const double x6 = __dmul_rn(x4, x2);
    const double c = __dadd_rn(c1, __dmul_rn(x4, C2));
    const double r = __dadd_rn(c, __dmul_rn(x6, c2));
    return (float)(flip ? -r : r);
}
__device__ __forceinline__ void libm_sincosf(float y, float& sn, float& cs)
{
    const unsigned top = (__float_as_uint(y) >> 20) & 0x7ffu;
    double x = (double)y;
    if (top < ((__float_as_uint(0x1.921FB6p-1f) >> 20) & 0x7ffu)) {           // |y| < pi/4
        const double x2 = __dmul_rn(x, x);
        if (top < ((__float_as_uint(0x1p-12f) >> 20) & 0x7ffu)) { sn = y; cs = 1.0f; return; }
        sn = libm_sincosf_poly(x, x2, false, 0);
        cs = libm_sincosf_poly(x, x2, false, 1);
        return;
    }
    const double r = __dmul_rn(x, 0x1.45F306DC9C883p+23);                      // x * 2/pi * 2^24
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    x = __dsub_rn(x, __dmul_rn((double)n, 0x1.921FB54442D18p0));
    const double sgn = ((n & 3) == 1 || (n & 3) == 2) ? -1.0 : 1.0;             // sign[n & 3] = {1, -1, -1, 1}
    const bool flip = (n & 2) != 0;
    const double xs = __dmul_rn(x, sgn), x2 = __dmul_rn(x, x);
    sn = libm_sincosf_poly(xs, x2, flip, n);
    cs = libm_sincosf_poly(xs, x2, flip, n ^ 1);
}

__global__ void __launch_bounds__(DESC_WARPS * 32, 10) describe_kernel(const __grid_constant__ DescParams P)
{
    __shared__ WarpSmem sm[DESC_WARPS];
    __shared__ IcTab ic;
    for (int i = threadIdx.x; i < (int)(sizeof(IcTab) / 4); i += DESC_WARPS * 32) ((unsigned*)&ic)[i] = __ldg((const unsigned*)&d_ic + i);
    __syncthreads();
    const Geometry& g = P.g;
    const int frame = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int slot = blockIdx.x * DESC_WARPS + warp;

    // level of this slot from the per-level kept counts (:936-967 concatenates levels in order)
    // (lane l holds level l's count; an inclusive scan over the first 16 lanes gives every level's end slot)
    const int cnt = lane < g.nlevels ? (int)P.d.lvCount[frame * g.nlevels + lane] : 0;
    int incl = cnt;
#pragma unroll
    for (int d = 1; d < 16; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += t;
    }
    const int total = __shfl_sync(0xffffffffu, incl, OS1_MAX_LEVELS - 1);
    const int below = __popc(__ballot_sync(0xffffffffu, lane < g.nlevels && incl <= slot));     // levels that end at or before the slot
    const int level = below < g.nlevels ? below : -1;
    const int j = slot - __shfl_sync(0xffffffffu, incl - cnt, below & 31);
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
    // ---- 43x43 window -> shared.  Interior keypoints: 16-byte asynchronous copies of the rows, starting at the 16-byte
    // boundary at or before the window's first column (level rows start aligned and have a pitch that is a multiple of
    // 16); the window keeps the byte phase `sh` (< 16) of its first column.  Keypoints within 21 px of the level's edge
    // take the byte path with REFLECT_101 indices, as the blur of the borderless clone sees them — and so do the few
    // whose last chunk would end beyond the bytes the level owns. ----------------------------------------------------
    const int xs = x - WIN_R, ys = y - WIN_R;
    int sh = xs & 15;
    const int nch = (sh + WIN + 15) >> 4;                       // chunks that hold window bytes: 3 or 4
    if (xs >= 0 && ys >= 0 && x + WIN_R < L.w && y + WIN_R < L.h && xs - sh + 16 * nch <= (level == 0 ? L.w : pitch)) {
        const uint8_t* base = img + (long long)ys * pitch + (xs - sh);
        for (int it = lane; it < WIN * 4; it += 32) {
            const int r = it >> 2, c = it & 3;
            if (c < nch)
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(S.raw4 + r * (RAW_PW / 4) + c)),
                             "l"(base + (long long)r * pitch + 16 * c) : "memory");      // all 6 copies of a lane in flight at once
        }
        asm volatile("cp.async.wait_all;" ::: "memory");
    } else {
        sh = 0;
        uint8_t* rb = (uint8_t*)S.raw4;
        for (int i = lane; i < WIN * WIN; i += 32) {
            const int r = i / WIN, c = i - r * WIN;
            const int yy = reflect101(ys + r, L.h), xx = reflect101(xs + c, L.w);
            rb[r * (RAW_PW * 4) + c] = img[(long long)yy * pitch + xx];
        }
    }
    __syncwarp();
    const unsigned* raw = (const unsigned*)S.raw4 + (sh >> 2);   // word that holds window column 0
    const int s8 = 8 * (sh & 3);

    // ---- intensity centroid: lane = patch row v.  The row's 31 pixels sit in nine aligned words; one dp4a per word
    // against the signed column offsets gives sum(u * I), one against the membership mask gives sum(I). -----------------
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int vv = lane - 15, av = vv < 0 ? -vv : vv;
        const unsigned* w = raw + (WIN_R + vv) * RAW_PW + 1;
        // (rows 8 apart start in the same bank: each lane walks its nine words from a different one)
        const int rot = lane >> 3;
        int rowSum = 0;
#pragma unroll
        for (int t = 0; t < 9; t++) {
            int k = t + rot;
            if (k >= 9) k -= 9;
            const unsigned a = __funnelshift_r(w[k], w[k + 1], s8);     // window columns 4(k+1) .. 4(k+1)+3
            m10 = dp4a_us(a, ic.cu[av][k], m10);
            rowSum = (int)__dp4a(a, ic.cm[av][k], (unsigned)rowSum);
        }
        m01 = vv * rowSum;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = fast_atan2_deg((float)m01, (float)m10, P);

    // ---- 7x7 Gaussian, Q8 taps {18,34,48,56,48,34,18}, exact two-pass fixed point ----------------------------------
    // Horizontal pass: item = (pair of rows, group of four columns).  After aligning the row to the window's byte phase,
    // output x of the group is dp4a(bytes x..x+3, {18,34,48,56}) + dp4a(bytes x+4..x+7, {48,34,18,0}); the results of the
    // two rows (<= 65280 each) are packed as u16x2 so that the vertical pass can use dp2a on them.
    {
        const unsigned K1 = 18u | (34u << 8) | (48u << 16) | (56u << 24), K2 = 48u | (34u << 8) | (18u << 16);
        // (three pairs per trip, lanes 30 and 31 idle: every lane of a pair has read its rows before any of them writes)
        const int sub = lane / HB_Q, q = lane - sub * HB_Q;
        for (int e0 = 0; e0 < HB_PAIRS; e0 += 3) {
            const int e = e0 + sub;
            const bool on = sub < 3 && e < HB_PAIRS;
            unsigned h[2][4];
            if (on) {
#pragma unroll
                for (int rr = 0; rr < 2; rr++) {
                    const int r = min(2 * e + rr, WIN - 1);
                    const unsigned* w = raw + r * RAW_PW + q;
                    const unsigned W0 = w[0], W1 = w[1], W2 = w[2], W3 = w[3];
                    const unsigned A0 = __funnelshift_r(W0, W1, s8), A1 = __funnelshift_r(W1, W2, s8), A2 = __funnelshift_r(W2, W3, s8);
                    h[rr][0] = __dp4a(A0, K1, __dp4a(A1, K2, 0u));
                    h[rr][1] = __dp4a(__funnelshift_r(A0, A1, 8), K1, __dp4a(__funnelshift_r(A1, A2, 8), K2, 0u));
                    h[rr][2] = __dp4a(__funnelshift_r(A0, A1, 16), K1, __dp4a(__funnelshift_r(A1, A2, 16), K2, 0u));
                    h[rr][3] = __dp4a(__funnelshift_r(A0, A1, 24), K1, __dp4a(__funnelshift_r(A1, A2, 24), K2, 0u));
                }
            }
            __syncwarp();
            if (on) S.raw4[e * HB_Q + q] = make_uint4(h[0][0] | (h[1][0] << 16), h[0][1] | (h[1][1] << 16), h[0][2] | (h[1][2] << 16), h[0][3] | (h[1][3] << 16));
        }
    }
    __syncwarp();
    const uint4* hp = S.raw4;
    unsigned* bl = S.bl;
    // Vertical pass: lane = (group of four columns, third of the rows).  Output rows 2m and 2m+1 both read the row pairs
    // m .. m+3 — with taps {k0,k1 | k2,k3 | k4,k5 | k6,0} and {0,k0 | k1,k2 | k3,k4 | k5,k6} — so a four-pair window
    // slides down one pair per step, and an output is four dp2a starting from the rounding constant.
    if (lane < 3 * HB_Q) {
        const int seg = lane / HB_Q, q = lane - seg * HB_Q;
        const int mb = seg == 0 ? 0 : (seg == 1 ? 7 : 13), me = seg == 0 ? 7 : (seg == 1 ? 13 : 19);
        const unsigned CE = 18u | (34u << 8) | (48u << 16) | (56u << 24), CE2 = 48u | (34u << 8) | (18u << 16);
        const unsigned CO = (18u << 8) | (34u << 16) | (48u << 24), CO2 = 56u | (48u << 8) | (34u << 16) | (18u << 24);
        uint4 wnd[4];                                   // circular: step t keeps pairs mb+t .. mb+t+3 in slots (t+j) % 4
#pragma unroll
        for (int k = 0; k < 3; k++) wnd[k] = hp[(mb + k) * HB_Q + q];
#pragma unroll
        for (int t = 0; t < 7; t++) {
            const int m = mb + t;
            if (m < me) {
                wnd[(t + 3) % 4] = hp[(m + 3) * HB_Q + q];
                const uint4 p0 = wnd[t % 4], p1 = wnd[(t + 1) % 4], p2 = wnd[(t + 2) % 4], p3 = wnd[(t + 3) % 4];
                const unsigned a0[4] = { p0.x, p0.y, p0.z, p0.w }, a1[4] = { p1.x, p1.y, p1.z, p1.w };
                const unsigned a2[4] = { p2.x, p2.y, p2.z, p2.w }, a3[4] = { p3.x, p3.y, p3.z, p3.w };
                unsigned ev[4], od[4];
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    ev[c] = __dp2a_hi(a3[c], CE2, __dp2a_lo(a2[c], CE2, __dp2a_hi(a1[c], CE, __dp2a_lo(a0[c], CE, 32768u))));
                    od[c] = __dp2a_hi(a3[c], CO2, __dp2a_lo(a2[c], CO2, __dp2a_hi(a1[c], CO, __dp2a_lo(a0[c], CO, 32768u))));
                }
                // (sum + 32768) >> 16 is byte 2 of each accumulator (< 2^24)
                bl[(2 * m) * (BL_PITCH / 4) + q] = __byte_perm(__byte_perm(ev[0], ev[1], 0x0062), __byte_perm(ev[2], ev[3], 0x0062), 0x5410);
                if (2 * m + 1 < HB_W)
                    bl[(2 * m + 1) * (BL_PITCH / 4) + q] = __byte_perm(__byte_perm(od[0], od[1], 0x0062), __byte_perm(od[2], od[3], 0x0062), 0x5410);
            }
        }
    }
    __syncwarp();

    // ---- rotated BRIEF: lane = descriptor byte ---------------------------------------------------------------------------
    const float ang = __fmul_rn(angle, P.factorPI);
    float a, b;
    libm_sincosf(ang, b, a);                   // a = cosf(ang), b = sinf(ang) as the reference's libm computes them
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

// umax exactly as the reference builds it (ORBextractor.cc:486-501)
static void build_umax(int* umax)
{
    const int HP = 15;
    int v, v0, vmax = (int)floor(HP * sqrt(2.f) / 2 + 1);
    int vmin = (int)ceil(HP * sqrt(2.f) / 2);
    const double hp2 = HP * HP;
    for (v = 0; v <= vmax; ++v) umax[v] = (int)lrint(sqrt(hp2 - v * v));
    for (v = HP, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
}

// Per-device constant table of the intensity-centroid pass.  Called from handle set-up (never from inside a launch
// sequence, which may be under stream capture), once per device, under a lock; synchronous.
cudaError_t os1_describe_init()
{
    static std::mutex mu;
    static bool filled[64] = { false };
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(mu);
    if (filled[dev]) return cudaSuccess;
    int umax[16];
    build_umax(umax);
    IcTab t;
    for (int av = 0; av < 16; av++)
        for (int k = 0; k < 9; k++) {
            int cu = 0; unsigned cm = 0;
            for (int b = 0; b < 4; b++) {
                const int u = 4 * (k + 1) + b - WIN_R;               // window column 4(k+1)+b  <->  patch offset u
                if (u >= -umax[av] && u <= umax[av]) { cu |= (u & 0xFF) << (8 * b); cm |= 1u << (8 * b); }
            }
            t.cu[av][k] = cu; t.cm[av][k] = cm;
        }
    e = cudaMemcpyToSymbol(d_ic, &t, sizeof(t), 0, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaStreamSynchronize(0);
    if (e == cudaSuccess) filled[dev] = true;
    return e;
}

cudaError_t os1_launch_describe(const Geometry& g, const DevPtrs& d, int nframes, os1_kp_dev* kp, uint8_t* desc, int capPerFrame,
                                int* counts, cudaStream_t s)
{
    static_assert(sizeof(DescParams) < 4000, "kernel parameter block too large");
    DescParams P;
    P.g = g; P.d = d; P.kp = kp; P.desc = desc; P.counts = counts; P.capPerFrame = capPerFrame;
    build_umax(P.umax);
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
