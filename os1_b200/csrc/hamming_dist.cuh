// 256-bit Hamming distance shared by the matcher kernels (hamming.cu) and the issue-rate micro-benchmarks (ubench.cu).
#pragma once
#include <cuda_runtime.h>

// POPC issues at a quarter of the LOP3 rate, so the eight XOR words are first squeezed through four carry-save adders
// (sum = a^b^c, carry = majority: one LOP3 each) into two weight-1 words, one weight-2 word and one weight-4 word:
// 4 POPC instead of 8 (48-frame sweep: 539 -> 760 G descriptor pairs/s on a B200).
__device__ __forceinline__ void os1_csa(unsigned a, unsigned b, unsigned c, unsigned& sum, unsigned& carry)
{
    sum = a ^ b ^ c;
    carry = (a & b) | (c & (a ^ b));
}
__device__ __forceinline__ int os1_dist256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1)
{
    const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
    const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
    unsigned sa, ca, sb, cb, sc, cc, sd, cd;
    os1_csa(x0, x1, x2, sa, ca);
    os1_csa(x3, x4, x5, sb, cb);
    os1_csa(x6, x7, sa, sc, cc);
    os1_csa(ca, cb, cc, sd, cd);
    return __popc(sb) + __popc(sc) + 2 * __popc(sd) + 4 * __popc(cd);
}
