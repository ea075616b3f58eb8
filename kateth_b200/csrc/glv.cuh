// glv.cuh -- split of a scalar along the G1 endomorphism of BLS12-381.
//
// phi(x, y) = (beta x, y) acts on G1 as multiplication by -x^2 (x = -0xd201000000010000 the curve parameter; see
// g1_in_subgroup in g1.cuh), so with N = x^2 = 0xac45a4010001a4020000000100000000 (r = N^2 - N + 1)
//     k = k1 + k2 * N,  0 <= k1 < N,  0 <= k2 < N      =>      [k]P = [k1]P + [k2]Q,   Q = -phi(P) = (beta x, -y).
// Used by the fixed-base MSM in GLV form (window_bits 16 or 100 + c: two 128-bit halves per scalar share one table of
// ceil(128 / c) windows, and the second half-sum is mapped through Q once per thread) and by the RLC scalar
// multiplications of batch verification (verify.cuh).
// Plain C++ (no PTX): compiled for the host by tests/test_glv.py and checked against Python integers.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GLV_HD __host__ __device__ __forceinline__
#else
#define GLV_HD static inline
#endif

namespace kb200 {

// k: 8 little-endian limbs, k < r.  k1, k2: 4 limbs each.
GLV_HD void glv_split(uint32_t (&k1)[4], uint32_t (&k2)[4], const uint32_t (&k)[8]) {
    const uint32_t Nl[4] = {0x00000000u, 0x00000001u, 0x0001a402u, 0xac45a401u};
    const uint32_t Ml[5] = {0xf6cfee2eu, 0x63f6e522u, 0xe01faaddu, 0x7c6becf1u, 0x00000001u};     // floor(2^256 / N)
    // q = (k * M) >> 256  (k2 - 1 <= q <= k2)
    uint32_t prod[13];
#pragma unroll
    for (int i = 0; i < 13; i++) prod[i] = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        uint64_t c = 0;
#pragma unroll
        for (int j = 0; j < 5; j++) {
            uint64_t t = (uint64_t)k[i] * Ml[j] + prod[i + j] + c;
            prod[i + j] = (uint32_t)t; c = t >> 32;
        }
        prod[i + 5] = (uint32_t)c;
    }
    uint32_t q[5] = {prod[8], prod[9], prod[10], prod[11], prod[12]};
    // rem = k - q * N  (fits in 5 limbs: rem < 2N)
    uint32_t qn[5] = {0, 0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 5; i++) {
        uint64_t c = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (i + j < 5) {
                uint64_t t = (uint64_t)q[i] * Nl[j] + qn[i + j] + c;
                qn[i + j] = (uint32_t)t; c = t >> 32;
            }
        }
        if (i == 0) qn[4] = (uint32_t)c;          // the only row whose carry-out still lands inside the five limbs kept
    }
    uint32_t rem[5], bw = 0;
#pragma unroll
    for (int i = 0; i < 5; i++) { uint64_t t = (uint64_t)k[i] - qn[i] - bw; rem[i] = (uint32_t)t; bw = (uint32_t)(t >> 63); }
    // at most two corrections (one observed): rem >= N  ->  rem -= N, q += 1
#pragma unroll
    for (int rep = 0; rep < 2; rep++) {
        uint32_t d[5], b2 = 0;
#pragma unroll
        for (int i = 0; i < 5; i++) { uint64_t t = (uint64_t)rem[i] - (i < 4 ? Nl[i] : 0u) - b2; d[i] = (uint32_t)t; b2 = (uint32_t)(t >> 63); }
        const uint32_t ge = b2 ^ 1u;
        uint32_t cy = ge;
#pragma unroll
        for (int i = 0; i < 5; i++) {
            rem[i] = ge ? d[i] : rem[i];
            uint32_t s = q[i] + cy; cy = (s < cy) ? 1u : 0u; q[i] = s;
        }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) { k1[i] = rem[i]; k2[i] = q[i]; }
}

}  // namespace kb200
