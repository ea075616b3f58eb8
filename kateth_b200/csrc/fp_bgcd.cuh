// fp_bgcd.cuh -- modular inversion in Fp (BLS12-381 base field) by a branchless binary GCD with word-sized
// approximations (the published "optimized binary GCD" construction: 30 inner steps on 62-bit approximations of
// (a, b) build a 2x2 transition matrix with |f| + |g| <= 2^30, which is then applied to the full 384-bit values;
// 26 outer iterations cover the 2*381 - 1 steps the worst case needs).
//
// Why it exists: kateth divides per element (Fr::div, /root/reference/src/bls.rs:297-312) and normalises per point
// (blst_p1_to_affine behind P1::compress, src/bls.rs:491-503).  On the GPU the batch-affine MSM (msm_ba.cuh) shares one
// inversion among a few dozen additions, so the inversion itself has to be cheap on the 32x32 multiplier: this one
// costs ~3.1k MAC32 (about 10 Fp products) plus ~20k ALU instructions that issue on the otherwise idle integer-ALU
// pipe, against ~460 Fp products for a^(p-2).  Every lane executes the same instruction stream (no data-dependent
// branch), so a warp inverts 32 values in lock step.
//
// Plain C++ on purpose (no PTX): the same code compiles for the host, where tests/test_bgcd.py checks it against
// pow(x, -1, p) without a GPU.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define KB_HD __host__ __device__ __forceinline__
#else
#define KB_HD static inline
#endif

namespace kb200 {

#define BG_K 30
#define BG_ITERS 26
#define BG_MASK ((1u << BG_K) - 1u)
#define BG_MINV 0x3ffcfffdu   /* -p^-1 mod 2^30 */
#define BG_P_LIMBS {0xffffaaabu, 0xb9feffffu, 0xb153ffffu, 0x1eabfffeu, 0xf6b0f624u, 0x6730d2a0u, 0xf38512bfu, 0x64774b84u, 0x434bacd7u, 0x4b1ba7b6u, 0x397fe69au, 0x1a0111eau}

KB_HD int bg_clz(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int)x);
#else
    return x ? __builtin_clz(x) : 32;
#endif
}

// t (13 limbs, two's complement) = a*f + b*g;  a, b unsigned 12 limbs, |f| + |g| <= 2^30
KB_HD void bg_lincomb(uint32_t (&t)[13], const uint32_t (&a)[12], const uint32_t (&b)[12], int32_t f, int32_t g) {
    int64_t acc = 0;
#pragma unroll
    for (int i = 0; i < 12; i++) {
        acc += (int64_t)(uint64_t)a[i] * (int64_t)f + (int64_t)(uint64_t)b[i] * (int64_t)g;
        t[i] = (uint32_t)acc;
        acc >>= 32;
    }
    t[12] = (uint32_t)acc;
}

// r = |(a*f + b*g) / 2^30| (the division is exact by construction); returns 1 when the quotient was negative
KB_HD uint32_t bg_update_ab(uint32_t (&r)[12], const uint32_t (&a)[12], const uint32_t (&b)[12], int32_t f, int32_t g) {
    uint32_t t[13];
    bg_lincomb(t, a, b, f, g);
    const uint32_t sign = t[12] >> 31, mask = 0u - sign;
    uint32_t carry = sign;
#pragma unroll
    for (int i = 0; i < 12; i++) {
        uint32_t w = ((t[i] >> BG_K) | (t[i + 1] << (32 - BG_K))) ^ mask;
        uint32_t s = w + carry;
        carry = (s < w) ? 1u : 0u;
        r[i] = s;
    }
    return sign;
}

// r = (u*f + v*g) / 2^30 mod p, in [0, p);  u, v in [0, p)
KB_HD void bg_update_uv(uint32_t (&r)[12], const uint32_t (&u)[12], const uint32_t (&v)[12], int32_t f, int32_t g) {
    const uint32_t p[12] = BG_P_LIMBS;
    uint32_t t[13];
    bg_lincomb(t, u, v, f, g);
    const uint32_t q = (t[0] * BG_MINV) & BG_MASK;
    uint64_t c = 0;
#pragma unroll
    for (int i = 0; i < 12; i++) {
        uint64_t s = (uint64_t)t[i] + (uint64_t)q * p[i] + c;
        t[i] = (uint32_t)s;
        c = s >> 32;
    }
    t[12] += (uint32_t)c;
    // (t >> 30) lies in (-p, 2p): 12 limbs and a sign
    uint32_t w[12];
#pragma unroll
    for (int i = 0; i < 12; i++) w[i] = (t[i] >> BG_K) | (t[i + 1] << (32 - BG_K));
    const uint32_t neg = t[12] >> 31;
    uint32_t plus[12], minus[12];
    uint32_t cy = 0, bw = 0;
#pragma unroll
    for (int i = 0; i < 12; i++) {
        uint64_t s = (uint64_t)w[i] + p[i] + cy;
        plus[i] = (uint32_t)s; cy = (uint32_t)(s >> 32);
        uint64_t d = (uint64_t)w[i] - p[i] - bw;
        minus[i] = (uint32_t)d; bw = (uint32_t)(d >> 63);
    }
#pragma unroll
    for (int i = 0; i < 12; i++) r[i] = neg ? plus[i] : (bw ? w[i] : minus[i]);
}

// out = x^-1 mod p as a plain integer for 0 < x < p.  Returns 1 when the gcd came out as 1 (always, for x != 0).
KB_HD int bg_inverse(uint32_t (&out)[12], const uint32_t (&x)[12]) {
    const uint32_t p[12] = BG_P_LIMBS;
    uint32_t a[12], b[12], u[12], v[12];
#pragma unroll
    for (int i = 0; i < 12; i++) { a[i] = x[i]; b[i] = p[i]; u[i] = 0; v[i] = 0; }
    u[0] = 1;
#pragma unroll 1
    for (int it = 0; it < BG_ITERS; it++) {
        // 62-bit approximations: the 32 bits below the common top bit of (a | b) and the low 30 bits (exact when both
        // values fit in 64 bits)
        uint32_t ahi = 0, alo = 0, bhi = 0, blo = 0, found = 0;
#pragma unroll
        for (int i = 11; i >= 2; i--) {
            uint32_t w = a[i] | b[i];
            uint32_t take = (w != 0u && !found) ? 1u : 0u;
            ahi = take ? a[i] : ahi; alo = take ? a[i - 1] : alo;
            bhi = take ? b[i] : bhi; blo = take ? b[i - 1] : blo;
            found |= (w != 0u) ? 1u : 0u;
        }
        uint64_t abar, bbar;
        {
            int sh = bg_clz(ahi | bhi) & 31;        // found => (ahi | bhi) != 0 => sh in [0, 31]
            uint32_t at = sh ? ((ahi << sh) | (alo >> (32 - sh))) : ahi;
            uint32_t bt = sh ? ((bhi << sh) | (blo >> (32 - sh))) : bhi;
            uint64_t aa = ((uint64_t)at << BG_K) | (a[0] & BG_MASK), ba = ((uint64_t)bt << BG_K) | (b[0] & BG_MASK);
            uint64_t ae = ((uint64_t)a[1] << 32) | a[0], be = ((uint64_t)b[1] << 32) | b[0];
            abar = found ? aa : ae; bbar = found ? ba : be;
        }
        int32_t f0 = 1, g0 = 0, f1 = 0, g1 = 1;
#pragma unroll 2
        for (int s = 0; s < BG_K; s++) {
            const uint32_t odd = (uint32_t)abar & 1u;
            const bool sw = odd && (abar < bbar);
            const uint64_t ta = sw ? bbar : abar, tb = sw ? abar : bbar;
            const int32_t tf0 = sw ? f1 : f0, tf1 = sw ? f0 : f1, tg0 = sw ? g1 : g0, tg1 = sw ? g0 : g1;
            const uint64_t m64 = 0ull - (uint64_t)odd;
            const int32_t m32 = -(int32_t)odd;
            abar = (ta - (tb & m64)) >> 1; bbar = tb;
            f0 = tf0 - (tf1 & m32); g0 = tg0 - (tg1 & m32);
            f1 = tf1 * 2; g1 = tg1 * 2;
        }
        uint32_t na[12], nb[12];
        const uint32_t sa = bg_update_ab(na, a, b, f0, g0);
        const uint32_t sb = bg_update_ab(nb, a, b, f1, g1);
        f0 = sa ? -f0 : f0; g0 = sa ? -g0 : g0;
        f1 = sb ? -f1 : f1; g1 = sb ? -g1 : g1;
        uint32_t nu[12], nv[12];
        bg_update_uv(nu, u, v, f0, g0);
        bg_update_uv(nv, u, v, f1, g1);
        uint32_t anz = 0;
#pragma unroll
        for (int i = 0; i < 12; i++) { a[i] = na[i]; b[i] = nb[i]; u[i] = nu[i]; v[i] = nv[i]; anz |= na[i]; }
        // Once a == 0 every further iteration is the identity on (b, v) (the matrix is ((1, 0), (0, 2^30)) and the update divides
        // by 2^30), so the loop may stop as soon as every lane of the warp is there: 17-20 iterations for random inputs
        // instead of the worst-case 26.
#if defined(__CUDA_ARCH__)
        if (!__any_sync(__activemask(), anz != 0u)) break;
#else
        if (anz == 0u) break;
#endif
    }
    uint32_t rest = b[0] ^ 1u;
#pragma unroll
    for (int i = 1; i < 12; i++) rest |= b[i];
#pragma unroll
    for (int i = 0; i < 12; i++) { rest |= a[i]; out[i] = v[i]; }
    return rest == 0u;
}

#if defined(__CUDACC__)
// r = a^-1 in Montgomery form (a = x*R  ->  r = x^-1 * R):  (x*R)^-1 * R^3 * R^-1.  0 -> 0 like fp_inv.
static __device__ __noinline__ void fp_inv_bgcd(Fp& r, const Fp& a) {
    const uint32_t r3[12] = {0xd94ca1e0u, 0xed48ac6bu, 0x03a7adf8u, 0x315f831eu, 0x615e29ddu, 0x9a53352au, 0x921e1761u, 0x34c04e5eu, 0x65724728u, 0x2512d435u, 0x91755d4du, 0x0aa63460u};
    uint32_t t[12];
    if (bg_inverse(t, a.l)) { fp_mul(r.l, t, r3); return; }
    fp_inv(r, a);     // never taken for a != 0; keeps the result defined (and identical to fp_inv) whatever happens
}
#endif

}  // namespace kb200
