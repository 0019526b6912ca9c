// 256-bit Hamming distance shared by the matcher kernels (hamming.cu) and the issue-rate micro-benchmarks (ubench.cu).
#pragma once
#include <cuda_runtime.h>

// POPC issues at 16 per clock per SM, a quarter of the logic rate, so the eight XOR words are first squeezed through carry-save
// adders (sum = a^b^c, carry = majority: one LOP3 each).  Three adders leave two weight-1 and three weight-2 words: 5 POPC and 14
// logic ops, the measured optimum of the inner loop (os1_b200/csrc/ubench.cu: 786 / 839 / 709 G pairs/s with four / three / two
// adders on a B200; no adders at all — 8 POPC — was 539 in the 48-frame sweep where four gave 760).
__device__ __forceinline__ void os1_csa(unsigned a, unsigned b, unsigned c, unsigned& sum, unsigned& carry)
{
    sum = a ^ b ^ c;
    carry = (a & b) | (c & (a ^ b));
}
// V selects how far the words are compressed before the population counts (all exact):
//   0: four carry-save adders, 4 POPC, 16 logic ops      1: three adders, 5 POPC, 14 logic ops      2: two adders, 6 POPC, 12 logic ops
#ifndef OS1_DIST_VARIANT
#define OS1_DIST_VARIANT 1
#endif
template <int V = OS1_DIST_VARIANT>
__device__ __forceinline__ int os1_dist256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    unsigned sa, ca, sb, cb, sc, cc, sd, cd;
    os1_csa(x0, x1, x2, sa, ca);
    os1_csa(x3, x4, x5, sb, cb);
    if (V == 2) return __popc(sa) + __popc(sb) + __popc(x6) + __popc(x7) + 2 * (__popc(ca) + __popc(cb));
    os1_csa(x6, x7, sa, sc, cc);
    if (V == 1) return __popc(sb) + __popc(sc) + 2 * (__popc(ca) + __popc(cb) + __popc(cc));
    os1_csa(ca, cb, cc, sd, cd);
    return __popc(sb) + __popc(sc) + 2 * __popc(sd) + 4 * __popc(cd);
}

// The matchers' key  dist << 20 | position  built straight from the population counts by multiply-adds: on sm_100 the integer
// multiply-add issues on the FMA pipe, which the distance leaves idle, while shifts, ORs and three-input adds compete with the XORs
// and carry-save adders for the logic pipe that bounds the loop.  key = pos + 2^20 * (w1a + w1b) + 2^21 * (w2a + w2b + w2c).
__device__ __forceinline__ unsigned os1_mad(unsigned a, unsigned b, unsigned c)
{
    unsigned r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ unsigned os1_dist256_key(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1, unsigned pos)
{
    const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    unsigned sa, ca, sb, cb, sc, cc;
    os1_csa(x0, x1, x2, sa, ca);
    os1_csa(x3, x4, x5, sb, cb);
    os1_csa(x6, x7, sa, sc, cc);
    unsigned k = os1_mad((unsigned)__popc(sb), 1u << 20, pos);
    k = os1_mad((unsigned)__popc(sc), 1u << 20, k);
    k = os1_mad((unsigned)__popc(ca), 1u << 21, k);
    k = os1_mad((unsigned)__popc(cb), 1u << 21, k);
    return os1_mad((unsigned)__popc(cc), 1u << 21, k);
}
