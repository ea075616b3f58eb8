// pairing.cuh -- the two-pairing check ON THE DEVICE, one item per thread: per-item verdicts of verify_kzg_proof at batch scale.
//
// Replaces, for many items at once, bls::verify_pairings (/root/reference/src/bls.rs:572-598) as called by
// Setup::verify_proof_inner (/root/reference/src/kzg/setup.rs:84-94).  kateth checks
//     e(pi, [tau]_2 - [z]G2) == e(C - [y]G1, G2);
// moving [z]G2 to the other side gives the same verdict with both G2 arguments FIXED:
//     e(-pi, [tau]_2) * e(C - [y]G1 + [z]pi, G2) == 1,
// so the Miller-loop line coefficients of [tau]_2 and G2 are the ones the host module precomputes at setup load (host_bls.cpp,
// g2_prepare) and every thread only evaluates them at its own two G1 points.  A batch verdict needs one pairing on the host (the
// RLC path, north_star); 65,536 PER-ITEM verdicts at ~1.5 ms of host pairing each would take minutes -- here they take a fraction
// of a second.
//
// Tower and formulas follow host_bls.cpp line for line (Fp2 = Fp[u]/(u^2+1), Fp6 = Fp2[v]/(v^3-(1+u)), Fp12 = Fp6[w]/(w^2-v); final
// exponentiation: easy part by conjugate / inverse / Frobenius, hard part (x-1)^2 (x+p)(x^2+p^2-1) + 3), on the device Fp.
// An Fp12 is 144 words: values live in local memory (L1) and every tower routine is out of line; this kernel is bound by that
// traffic and by the dependent chain of one pairing per thread, not by the multiplier -- it is off the north_star hot path.
#pragma once
#include "g1.cuh"

namespace kb200 {

struct Fp2 { Fp c0, c1; };
struct Fp6 { Fp2 c0, c1, c2; };
struct Fp12 { Fp6 c0, c1; };
struct PairLine { Fp2 c0, c1; };                  // host::Line, same bytes (6 x u64 Montgomery limbs == 12 x u32)
struct PairConsts {                               // filled at setup load from the host module
    PairLine tau[68], g2[68];
    Fp2 frob1[6];                                 // gamma^k, gamma = xi^((p-1)/6)
    Fp frob2[6];                                  // xi^((p^2-1)/6 * k)
    int nlines;
};

// ------------------------------------------------------------------ Fp2
__device__ __forceinline__ void f2add(Fp2& r, const Fp2& a, const Fp2& b) { fp_add(r.c0, a.c0, b.c0); fp_add(r.c1, a.c1, b.c1); }
__device__ __forceinline__ void f2sub(Fp2& r, const Fp2& a, const Fp2& b) { fp_sub(r.c0, a.c0, b.c0); fp_sub(r.c1, a.c1, b.c1); }
__device__ __forceinline__ void f2neg(Fp2& r, const Fp2& a) { fp_neg(r.c0, a.c0); fp_neg(r.c1, a.c1); }
__device__ __forceinline__ void f2conj(Fp2& r, const Fp2& a) { r.c0 = a.c0; fp_neg(r.c1, a.c1); }
__device__ __forceinline__ bool f2zero(const Fp2& a) { return fp_is_zero(a.c0) && fp_is_zero(a.c1); }
__device__ __forceinline__ void f2mulxi(Fp2& r, const Fp2& a) { Fp t0, t1; fp_sub(t0, a.c0, a.c1); fp_add(t1, a.c0, a.c1); r.c0 = t0; r.c1 = t1; }
static __device__ __noinline__ void f2mul(Fp2& r, const Fp2& a, const Fp2& b) {          // Karatsuba, 3 products
    Fp t0, t1, s0, s1, m;
    fp_mul(t0, a.c0, b.c0); fp_mul(t1, a.c1, b.c1);
    fp_add(s0, a.c0, a.c1); fp_add(s1, b.c0, b.c1); fp_mul(m, s0, s1);
    fp_sub(r.c0, t0, t1); fp_sub(m, m, t0); fp_sub(r.c1, m, t1);
}
static __device__ __noinline__ void f2sqr(Fp2& r, const Fp2& a) {                        // (a0+a1)(a0-a1), 2 a0 a1
    Fp s, d, m;
    fp_add(s, a.c0, a.c1); fp_sub(d, a.c0, a.c1); fp_mul(m, a.c0, a.c1);
    fp_mul(r.c0, s, d); fp_add(r.c1, m, m);
}
static __device__ __noinline__ void f2mulfp(Fp2& r, const Fp2& a, const Fp& b) { fp_mul(r.c0, a.c0, b); fp_mul(r.c1, a.c1, b); }
static __device__ __noinline__ void f2inv(Fp2& r, const Fp2& a) {
    Fp n, t; fp_sqr(n, a.c0); fp_sqr(t, a.c1); fp_add(n, n, t); fp_inv_bgcd(n, n);
    fp_mul(r.c0, a.c0, n); fp_mul(t, a.c1, n); fp_neg(r.c1, t);
}

// ------------------------------------------------------------------ Fp6 / Fp12
__device__ __forceinline__ void f6add(Fp6& r, const Fp6& a, const Fp6& b) { f2add(r.c0, a.c0, b.c0); f2add(r.c1, a.c1, b.c1); f2add(r.c2, a.c2, b.c2); }
__device__ __forceinline__ void f6sub(Fp6& r, const Fp6& a, const Fp6& b) { f2sub(r.c0, a.c0, b.c0); f2sub(r.c1, a.c1, b.c1); f2sub(r.c2, a.c2, b.c2); }
__device__ __forceinline__ void f6neg(Fp6& r, const Fp6& a) { f2neg(r.c0, a.c0); f2neg(r.c1, a.c1); f2neg(r.c2, a.c2); }
static __device__ __noinline__ void f6mul(Fp6& r, const Fp6& a, const Fp6& b) {
    Fp2 v0, v1, v2, t0, t1, t2, s, u;
    f2mul(v0, a.c0, b.c0); f2mul(v1, a.c1, b.c1); f2mul(v2, a.c2, b.c2);
    f2add(s, a.c1, a.c2); f2add(u, b.c1, b.c2); f2mul(t0, s, u); f2sub(t0, t0, v1); f2sub(t0, t0, v2); f2mulxi(t0, t0); f2add(t0, t0, v0);
    f2add(s, a.c0, a.c1); f2add(u, b.c0, b.c1); f2mul(t1, s, u); f2sub(t1, t1, v0); f2sub(t1, t1, v1); f2mulxi(s, v2); f2add(t1, t1, s);
    f2add(s, a.c0, a.c2); f2add(u, b.c0, b.c2); f2mul(t2, s, u); f2sub(t2, t2, v0); f2sub(t2, t2, v2); f2add(t2, t2, v1);
    r.c0 = t0; r.c1 = t1; r.c2 = t2;
}
__device__ __forceinline__ void f6mulv(Fp6& r, const Fp6& a) { Fp2 t; f2mulxi(t, a.c2); Fp2 a0 = a.c0, a1 = a.c1; r.c0 = t; r.c1 = a0; r.c2 = a1; }
static __device__ __noinline__ void f6inv(Fp6& r, const Fp6& a) {
    Fp2 t0, t1, t2, s, d;
    f2sqr(t0, a.c0); f2mul(s, a.c1, a.c2); f2mulxi(s, s); f2sub(t0, t0, s);
    f2sqr(t1, a.c2); f2mulxi(t1, t1); f2mul(s, a.c0, a.c1); f2sub(t1, t1, s);
    f2sqr(t2, a.c1); f2mul(s, a.c0, a.c2); f2sub(t2, t2, s);
    f2mul(d, a.c2, t1); f2mul(s, a.c1, t2); f2add(d, d, s); f2mulxi(d, d); f2mul(s, a.c0, t0); f2add(d, d, s);
    f2inv(d, d);
    f2mul(r.c0, t0, d); f2mul(r.c1, t1, d); f2mul(r.c2, t2, d);
}
static __device__ __noinline__ void f12mul(Fp12& r, const Fp12& a, const Fp12& b) {
    Fp6 t0, t1, s, u, m;
    f6mul(t0, a.c0, b.c0); f6mul(t1, a.c1, b.c1);
    f6add(s, a.c0, a.c1); f6add(u, b.c0, b.c1); f6mul(m, s, u); f6sub(m, m, t0); f6sub(m, m, t1);
    f6mulv(s, t1); f6add(r.c0, t0, s); r.c1 = m;
}
static __device__ __noinline__ void f12sqr(Fp12& r, const Fp12& a) {
    Fp6 ab, s, t, va1;
    f6mul(ab, a.c0, a.c1);
    f6add(s, a.c0, a.c1); f6mulv(va1, a.c1); f6add(t, a.c0, va1);
    f6mul(s, s, t); f6sub(s, s, ab); f6mulv(t, ab); f6sub(s, s, t);
    r.c0 = s; f6add(r.c1, ab, ab);
}
__device__ __forceinline__ void f12conj(Fp12& r, const Fp12& a) { r.c0 = a.c0; f6neg(r.c1, a.c1); }
static __device__ __noinline__ void f12inv(Fp12& r, const Fp12& a) {
    Fp6 t0, t1; f6mul(t0, a.c0, a.c0); f6mul(t1, a.c1, a.c1); f6mulv(t1, t1); f6sub(t0, t0, t1); f6inv(t0, t0);
    f6mul(r.c0, a.c0, t0); f6mul(t1, a.c1, t0); f6neg(r.c1, t1);
}
__device__ __forceinline__ void f12one(Fp12& r) {
    uint32_t* w = reinterpret_cast<uint32_t*>(&r);
    for (int k = 0; k < 144; k++) w[k] = 0;
    fp_set(r.c0.c0.c0, c_k.fp_one);
}
__device__ __forceinline__ bool f12isone(const Fp12& a) {
    Fp one; fp_set(one, c_k.fp_one);
    return fp_eq(a.c0.c0.c0, one) && fp_is_zero(a.c0.c0.c1) && f2zero(a.c0.c1) && f2zero(a.c0.c2) && f2zero(a.c1.c0) && f2zero(a.c1.c1) && f2zero(a.c1.c2);
}
// coefficient of w^k (k = 0..5) in the flat basis: w^0->c0.c0, w^1->c1.c0, w^2->c0.c1, w^3->c1.c1, w^4->c0.c2, w^5->c1.c2
__device__ __forceinline__ Fp2& f12wcoef(Fp12& a, int k) { Fp6& h = (k & 1) ? a.c1 : a.c0; return (k >> 1) == 0 ? h.c0 : ((k >> 1) == 1 ? h.c1 : h.c2); }
static __device__ __noinline__ void f12frob(Fp12& r, const Fp12& a, const PairConsts* pc) {       // a^p
    Fp12 t = a;
    for (int k = 0; k < 6; k++) { Fp2 c; f2conj(c, f12wcoef(t, k)); Fp2 g = pc->frob1[k]; f2mul(f12wcoef(t, k), c, g); }
    r = t;
}
static __device__ __noinline__ void f12frob2(Fp12& r, const Fp12& a, const PairConsts* pc) {      // a^(p^2)
    Fp12 t = a;
    for (int k = 0; k < 6; k++) { Fp2 c = f12wcoef(t, k); Fp g = pc->frob2[k]; f2mulfp(f12wcoef(t, k), c, g); }
    r = t;
}
// f *= l,  l = c0 + c2 w^2 + c3 w^3 (c3 in Fp): in tower form (c0 + c2 v) + (c3 v) w
static __device__ __noinline__ void f12mul_line(Fp12& f, const Fp2& c0, const Fp2& c2, const Fp& c3) {
    Fp12 l;
    uint32_t* w = reinterpret_cast<uint32_t*>(&l);
    for (int k = 0; k < 144; k++) w[k] = 0;
    l.c0.c0 = c0; l.c0.c1 = c2; l.c1.c1.c0 = c3;
    f12mul(f, f, l);
}
#define KB_BLS_X_ABS 0xd201000000010000ull
static __device__ __noinline__ void f12pow_x(Fp12& r, const Fp12& a) {                 // a^|x|
    Fp12 acc = a;
#pragma unroll 1
    for (int i = 62; i >= 0; i--) { f12sqr(acc, acc); if ((KB_BLS_X_ABS >> i) & 1) f12mul(acc, acc, a); }
    r = acc;
}
__device__ __forceinline__ void f12pow_negx(Fp12& r, const Fp12& a) { f12pow_x(r, a); f12conj(r, r); }   // unitary a, x < 0

// e(-a1, [tau]_2) * e(b1, G2) == 1 ?   a1, b1 affine Montgomery, (0, 0) = infinity  (host::verify_pairings)
static __device__ __noinline__ bool pairing_check(const G1Affine& a1, const G1Affine& b1, const PairConsts* pc) {
    const bool use_a = !g1a_is_inf(a1), use_b = !g1a_is_inf(b1);
    if (!use_a && !use_b) return true;
    Fp nay; fp_neg(nay, a1.y);
    Fp12 f; f12one(f);
    int li = 0; bool started = false;
#pragma unroll 1
    for (int i = 63; i >= 0; i--) {
        const bool bit = (KB_BLS_X_ABS >> i) & 1;
        if (!started) { started = bit; continue; }
        f12sqr(f, f);
#pragma unroll 1
        for (int rep = 0; rep < (bit ? 2 : 1); rep++) {
            if (use_a) { PairLine l = pc->tau[li]; Fp2 c2; f2mulfp(c2, l.c1, a1.x); f12mul_line(f, l.c0, c2, nay); }
            if (use_b) { PairLine l = pc->g2[li]; Fp2 c2; f2mulfp(c2, l.c1, b1.x); f12mul_line(f, l.c0, c2, b1.y); }
            li++;
        }
    }
    // easy part: f^((p^6-1)(p^2+1))
    Fp12 t, u; f12conj(t, f); f12inv(u, f); f12mul(f, t, u);
    f12frob2(t, f, pc); f12mul(f, t, f);
    // hard part: f^((x-1)^2 (x+p)(x^2+p^2-1) + 3)
    Fp12 a, b, c, fc;
    f12conj(fc, f);
    f12pow_negx(a, f); f12mul(a, a, fc);
    f12pow_negx(t, a); f12conj(u, a); f12mul(a, t, u);
    f12pow_negx(t, a); f12frob(u, a, pc); f12mul(b, t, u);
    f12pow_negx(t, b); f12pow_negx(t, t); f12frob2(u, b, pc); f12mul(c, t, u); f12conj(u, b); f12mul(c, c, u);
    f12sqr(t, f); f12mul(t, t, f); f12mul(c, c, t);
    return f12isone(c);
}

// neg_g_bits[k] = 2^k * (-G1), affine, k = 0..254 (built at setup): [y](-G1) is at most 255 mixed additions, no doublings
__global__ void k_neg_g_bits(G1Affine* __restrict__ out) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= 255) return;
    G1Affine ng; fp_set(ng.x, c_k.g1_gen_x); fp_set(ng.y, c_k.g1_neg_gen_y);
    G1 t; g1_from_affine(t, ng);
#pragma unroll 1
    for (int j = 0; j < k; j++) g1_dbl(t);
    G1Affine a; g1_to_affine(a, t); out[k] = a;
}

// Per item: R = C - [y]G1 + [z]pi (affine), then the two-pairing check against the fixed G2 points.  One thread per item.
//   ok[i] = 1 / 0; items whose status is non-zero (a decoding error) are skipped with ok[i] = 0.
#define PAIR_THREADS 64
__global__ void __launch_bounds__(PAIR_THREADS) k_verify_proof_many(const G1Affine* __restrict__ proofs, const G1Affine* __restrict__ comms,
                                                                     const Fr* __restrict__ z_canon, const Fr* __restrict__ y_canon, const int32_t* __restrict__ status,
                                                                     const G1Affine* __restrict__ neg_g_bits, const PairConsts* __restrict__ pc, int n,
                                                                     int32_t* __restrict__ ok) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (status != nullptr && status[i] != 0) { ok[i] = 0; return; }
    G1Affine pi = proofs[i];
    G1 acc;
    { Fr z = z_canon[i]; g1_mul(acc, pi, z); }                       // [z]pi  (infinity stays infinity)
    { G1Affine c = comms[i]; g1_madd_signed(acc, c, false); }        // + C
    {
        Fr y = y_canon[i];
#pragma unroll 1
        for (int k = 0; k < 255; k++)
            if ((y.l[k >> 5] >> (k & 31)) & 1) { G1Affine b = neg_g_bits[k]; g1_madd(acc, b.x, b.y); }     // + [y](-G1)
    }
    G1Affine r; g1_to_affine(r, acc);
    ok[i] = pairing_check(pi, r, pc) ? 1 : 0;
}

}  // namespace kb200
