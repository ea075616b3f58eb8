// g1.cuh -- the group G1 of BLS12-381 (y^2 = x^3 + 4 over Fp) on the device.
//
// Replaces blst's P1 arithmetic behind kateth's `P1` wrapper (/root/reference/src/bls.rs:362-552):
//   blst_p1_add (complete addition), blst_p1_cneg, blst_p1_mult, blst_p1_to_affine / from_affine,
//   blst_p1_compress (src/bls.rs:491-503), blst_p1_uncompress + blst_p1_affine_in_g1 (src/bls.rs:505-531).
// Accumulators use extended Jacobian "XYZZ" coordinates (x = X/ZZ, y = Y/ZZZ, ZZ^3 = ZZZ^2; ZZ == 0 is the point at
// infinity); table / input points are affine with (0,0) standing for infinity ((0,0) is not on the curve).
// Every addition handles P == Q, P == -Q and infinity, so the result is the unique group element blst returns.
#pragma once
#include "field.cuh"
#include "fp_bgcd.cuh"
#include "glv.cuh"

namespace kb200 {

struct G1Affine { Fp x, y; };             // 96 bytes, Montgomery form; (0,0) = infinity
struct G1 { Fp x, y, zz, zzz; };          // XYZZ

__device__ __forceinline__ bool g1a_is_inf(const G1Affine& p) { return fp_is_zero(p.x) && fp_is_zero(p.y); }
__device__ __forceinline__ bool g1_is_inf(const G1& p) { return fp_is_zero(p.zz); }
__device__ __forceinline__ void g1_set_inf(G1& p) { fp_zero(p.x); fp_zero(p.y); fp_zero(p.zz); fp_zero(p.zzz); }
__device__ __forceinline__ void g1_from_affine(G1& r, const G1Affine& p) {
    if (g1a_is_inf(p)) { g1_set_inf(r); return; }
    r.x = p.x; r.y = p.y; fp_set(r.zz, c_k.fp_one); fp_set(r.zzz, c_k.fp_one);
}

// r = 2*(x, y) for an affine, finite point
static __device__ __noinline__ void g1_dbl_affine(G1& r, const Fp& x, const Fp& y) {
    Fp u, v, w, s, m, t;
    fp_dbl(u, y); FPS(v, u); FPM(w, u, v); FPM(s, x, v);
    FPS(m, x); fp_dbl(t, m); fp_add(m, m, t);
    FPS(r.x, m); fp_sub(r.x, r.x, s); fp_sub(r.x, r.x, s);
    fp_sub(t, s, r.x); FPM(t, m, t); FPM(u, w, y); fp_sub(r.y, t, u);
    r.zz = v; r.zzz = w;
}
// p = 2*p
static __device__ __noinline__ void g1_dbl(G1& p) {
    if (g1_is_inf(p)) return;
    Fp u, v, w, s, m, t;
    fp_dbl(u, p.y); FPS(v, u); FPM(w, u, v); FPM(s, p.x, v);
    FPS(m, p.x); fp_dbl(t, m); fp_add(m, m, t);
    FPM(t, w, p.y);                      // W*Y1 before Y is overwritten
    FPS(p.x, m); fp_sub(p.x, p.x, s); fp_sub(p.x, p.x, s);
    fp_sub(s, s, p.x); FPM(s, m, s); fp_sub(p.y, s, t);
    FPM(p.zz, v, p.zz); FPM(p.zzz, w, p.zzz);
}
// p = 2^k * p.  Runs of three or more doublings leave XYZZ for Jacobian coordinates, where a doubling on this curve (a = 0) is
// 2M + 5S (dbl-2009-l) against 9 products in XYZZ: (X, Y, ZZ, ZZZ) -> (X*ZZ, Y*ZZZ, Z = ZZ), back with ZZ = Z^2, ZZZ = Z^3.  E(Fp) has
// odd order, so Y = 0 never occurs on a finite point; infinity (ZZ = 0) stays Z = 0.
static __device__ __noinline__ void g1_dbl_k(G1& p, int k) {
    if (k < 3) { for (int i = 0; i < k; i++) g1_dbl(p); return; }
    if (g1_is_inf(p)) return;
    Fp X, Y, Z, A, B, C, D, E, t;
    FPM(X, p.x, p.zz); FPM(Y, p.y, p.zzz); Z = p.zz;
#pragma unroll 1
    for (int i = 0; i < k; i++) {
        FPS(A, X); FPS(B, Y); FPS(C, B);
        fp_add(t, X, B); FPS(t, t); fp_sub(t, t, A); fp_sub(t, t, C); fp_dbl(D, t);
        fp_dbl(E, A); fp_add(E, E, A);
        FPM(Z, Y, Z); fp_dbl(Z, Z);                     // Z3 = 2 Y1 Z1, before Y is overwritten
        FPS(t, E); fp_sub(t, t, D); fp_sub(X, t, D);    // X3 = E^2 - 2D
        fp_sub(t, D, X); FPM(t, E, t);
        fp_dbl(C, C); fp_dbl(C, C); fp_dbl(C, C);
        fp_sub(Y, t, C);                                   // Y3 = E (D - X3) - 8C
    }
    p.x = X; p.y = Y; FPS(p.zz, Z); FPM(p.zzz, p.zz, Z);
}
// rare paths of the mixed addition, kept out of line so the hot loop stays small
static __device__ __noinline__ void g1_madd_rare(G1* acc, const Fp* x2, const Fp* y2, bool same_y) {
    if (same_y) { G1 t; g1_dbl_affine(t, *x2, *y2); *acc = t; } else g1_set_inf(*acc);
}
// acc += (x2, y2)   [affine, finite].  8M + 2S on the common path.
__device__ __forceinline__ void g1_madd(G1& acc, const Fp& x2, const Fp& y2) {
    if (g1_is_inf(acc)) { acc.x = x2; acc.y = y2; fp_set(acc.zz, c_k.fp_one); fp_set(acc.zzz, c_k.fp_one); return; }
    Fp p, r, pp, ppp, q, t;
    fp_mul(p, x2, acc.zz); fp_sub(p, p, acc.x);          // P = U2 - X1
    fp_mul(r, y2, acc.zzz); fp_sub(r, r, acc.y);         // R = S2 - Y1
    if (fp_is_zero(p)) { G1 tmp = acc; g1_madd_rare(&tmp, &x2, &y2, fp_is_zero(r)); acc = tmp; return; }
    fp_sqr(pp, p); fp_mul(ppp, p, pp); fp_mul(q, acc.x, pp);
    fp_sqr(t, r); fp_sub(t, t, ppp); fp_sub(t, t, q); fp_sub(acc.x, t, q);     // X3 = R^2 - PPP - 2Q
    fp_sub(q, q, acc.x); fp_mul(q, r, q); fp_mul(t, acc.y, ppp); fp_sub(acc.y, q, t);
    fp_mul(acc.zz, acc.zz, pp); fp_mul(acc.zzz, acc.zzz, ppp);
}
// the same mixed addition out of line on the shared product (callers outside the MSM hot loop)
static __device__ __noinline__ void g1_madd_c(G1& acc, const Fp& x2, const Fp& y2) {
    if (g1_is_inf(acc)) { acc.x = x2; acc.y = y2; fp_set(acc.zz, c_k.fp_one); fp_set(acc.zzz, c_k.fp_one); return; }
    Fp p, r, pp, ppp, q, t;
    FPM(p, x2, acc.zz); fp_sub(p, p, acc.x);
    FPM(r, y2, acc.zzz); fp_sub(r, r, acc.y);
    if (fp_is_zero(p)) { G1 tmp = acc; g1_madd_rare(&tmp, &x2, &y2, fp_is_zero(r)); acc = tmp; return; }
    FPS(pp, p); FPM(ppp, p, pp); FPM(q, acc.x, pp);
    FPS(t, r); fp_sub(t, t, ppp); fp_sub(t, t, q); fp_sub(acc.x, t, q);
    fp_sub(q, q, acc.x); FPM(q, r, q); FPM(t, acc.y, ppp); fp_sub(acc.y, q, t);
    FPM(acc.zz, acc.zz, pp); FPM(acc.zzz, acc.zzz, ppp);
}
// acc += pt (affine, possibly the (0,0) infinity sentinel), optionally negated
__device__ __forceinline__ void g1_madd_signed(G1& acc, const G1Affine& pt, bool neg) {
    if (g1a_is_inf(pt)) return;
    Fp y; fp_cneg(y, pt.y, neg);
    g1_madd(acc, pt.x, y);
}
// acc += b   (both XYZZ).  12M + 2S on the common path.
static __device__ __noinline__ void g1_add(G1& acc, const G1& b) {
    if (g1_is_inf(b)) return;
    if (g1_is_inf(acc)) { acc = b; return; }
    Fp u1, u2, s1, s2, p, r, pp, ppp, q, t;
    FPM(u1, acc.x, b.zz); FPM(u2, b.x, acc.zz);
    FPM(s1, acc.y, b.zzz); FPM(s2, b.y, acc.zzz);
    fp_sub(p, u2, u1); fp_sub(r, s2, s1);
    if (fp_is_zero(p)) {
        if (fp_is_zero(r)) g1_dbl(acc); else g1_set_inf(acc);
        return;
    }
    FPS(pp, p); FPM(ppp, p, pp); FPM(q, u1, pp);
    FPS(t, r); fp_sub(t, t, ppp); fp_sub(t, t, q); fp_sub(acc.x, t, q);
    fp_sub(q, q, acc.x); FPM(q, r, q); FPM(t, s1, ppp); fp_sub(acc.y, q, t);
    FPM(acc.zz, acc.zz, b.zz); FPM(acc.zz, acc.zz, pp);
    FPM(acc.zzz, acc.zzz, b.zzz); FPM(acc.zzz, acc.zzz, ppp);
}
// XYZZ -> affine (one field inversion)
static __device__ __noinline__ void g1_to_affine(G1Affine& r, const G1& p) {
    if (g1_is_inf(p)) { fp_zero(r.x); fp_zero(r.y); return; }
    Fp i3, zi, i2;
    fp_inv_bgcd(i3, p.zzz);                  // Z^-3 (branchless binary GCD, fp_bgcd.cuh; same value as a^(p-2))
    FPM(zi, p.zz, i3); FPS(i2, zi);    // Z^-1, Z^-2
    FPM(r.x, p.x, i2); FPM(r.y, p.y, i3);
}
// r = [k]P for a plain 255-bit scalar (little-endian limbs); MSB-first double-and-add  (blst_p1_mult(..., 255))
static __device__ __noinline__ void g1_mul(G1& r, const G1Affine& p, const Fr& k, int nbits = 255) {
    G1 acc; g1_set_inf(acc);
    if (!g1a_is_inf(p)) {
#pragma unroll 1
        for (int i = nbits - 1; i >= 0; i--) {
            g1_dbl(acc);
            if ((k.l[i >> 5] >> (i & 31)) & 1) g1_madd(acc, p.x, p.y);
        }
    }
    r = acc;
}

// GLV split of a canonical scalar k < r (glv.cuh): [k]P = [k1]P + [k2]Q, Q = -phi(P) = (beta x, -y), two 128-bit scalar
// multiplications that can run on two threads.
__device__ __forceinline__ void fr_glv_split(Fr& k1, Fr& k2, const Fr& k) {
    uint32_t a[4], b[4];
    glv_split(a, b, k.l);
#pragma unroll
    for (int j = 0; j < 8; j++) { k1.l[j] = j < 4 ? a[j] : 0u; k2.l[j] = j < 4 ? b[j] : 0u; }
}

// ---- ZCash 48-byte compressed form ----
// out: 12 big-endian words.  p affine.
__device__ __forceinline__ void g1_compress(uint32_t* out12, const G1Affine& p) {
    if (g1a_is_inf(p)) {
#pragma unroll
        for (int i = 0; i < 12; i++) out12[i] = 0;
        out12[0] = __byte_perm(0xc0000000u, 0, 0x0123);
        return;
    }
    Fp x, y; fp_from_mont(x, p.x); fp_from_mont(y, p.y);
    uint32_t top = x.l[11] | 0x80000000u; if (fp_plain_gt_half(y)) top |= 0x20000000u;
    out12[0] = __byte_perm(top, 0, 0x0123);
#pragma unroll
    for (int i = 1; i < 12; i++) out12[i] = __byte_perm(x.l[11 - i], 0, 0x0123);
}
// status codes (include/kateth_b200.h)
#define KB_OK 0
#define KB_INVALID_LEN 1
#define KB_INVALID_FIELD_ELEMENT 2
#define KB_G1_INVALID_ENCODING 3
#define KB_G1_NOT_ON_CURVE 4
#define KB_G1_NOT_IN_GROUP 5
#define KB_SCALAR_INVALID_ENCODING 6
#define KB_SCALAR_NOT_IN_FIELD 7

// in12: the 48 bytes as 12 words (memory order).  Curve check only (blst_p1_uncompress).
static __device__ __noinline__ int g1_uncompress(G1Affine& r, const uint32_t* in12) {
    uint32_t w0 = __byte_perm(in12[0], 0, 0x0123);
    if (!(w0 & 0x80000000u)) return KB_G1_INVALID_ENCODING;
    if (w0 & 0x40000000u) {
        uint32_t rest = w0 & 0x3fffffffu;
        for (int i = 1; i < 12; i++) rest |= in12[i];
        if (rest) return KB_G1_INVALID_ENCODING;
        fp_zero(r.x); fp_zero(r.y); return KB_OK;
    }
    Fp x;
    x.l[11] = w0 & 0x1fffffffu;
#pragma unroll
    for (int i = 1; i < 12; i++) x.l[11 - i] = __byte_perm(in12[i], 0, 0x0123);
    if (fp_plain_geq_p(x)) return KB_G1_INVALID_ENCODING;
    Fp xm, y2, y, t, b;
    fp_to_mont(xm, x);
    FPS(y2, xm); FPM(y2, y2, xm); fp_set(b, c_k.fp_b4); fp_add(y2, y2, b);
    fp_pow_c(y, y2, c_k.fp_sqrt_exp);
    FPS(t, y);
    if (!fp_eq(t, y2)) return KB_G1_NOT_ON_CURVE;
    fp_from_mont(t, y);
    bool want = (w0 & 0x20000000u) != 0;
    fp_cneg(y, y, fp_plain_gt_half(t) != want);
    r.x = xm; r.y = y; return KB_OK;
}
// acc = [|x|] base, |x| = 0xd201000000010000 (the BLS parameter; bits 63, 62, 60, 57, 48, 16), MSB-first double-and-add.
// AFFINE = true: base is the affine point (bx, by); false: base is the XYZZ point *bz.
template <bool AFFINE>
static __device__ __noinline__ void g1_mul_bls_x(G1& acc, const Fp& bx, const Fp& by, const G1* bz) {
    const uint64_t X = 0xd201000000010000ull;
    if (AFFINE) { acc.x = bx; acc.y = by; fp_set(acc.zz, c_k.fp_one); fp_set(acc.zzz, c_k.fp_one); } else acc = *bz;
    // doublings between the set bits below bit 63: 1, 2, 3, 9, 32, and 16 after the last addition
#pragma unroll 1
    for (int i = 62, run = 0; i >= 0; i--) {
        run++;
        if (((X >> i) & 1) || i == 0) {
            g1_dbl_k(acc, run); run = 0;
            if ((X >> i) & 1) { if (AFFINE) g1_madd_c(acc, bx, by); else g1_add(acc, *bz); }
        }
    }
}
// Subgroup membership (blst_p1_affine_in_g1).  With beta the cube root of unity below and phi(x, y) = (beta x, y), G1 is the
// eigenspace phi(P) = [-x^2]P (r = x^4 - x^2 + 1, so lambda = -x^2 satisfies lambda^2 + lambda + 1 = 0 mod r); a point of the
// curve is in G1 iff that holds (Scott, "A note on group membership tests for G1, G2 and GT on BLS pairing-friendly curves",
// 2021 -- the test blst uses).  126 doublings + 10 additions instead of the 255 + 128 of [r]P == infinity; checked against the
// definition on random curve points, cofactor-cleared points and small-order torsion points (tests/test_gpu_parity.py).
static __device__ __noinline__ bool g1_in_subgroup(const G1Affine& p) {
    if (g1a_is_inf(p)) return true;
    G1 t1, t2;
    g1_mul_bls_x<true>(t1, p.x, p.y, nullptr);
    if (g1_is_inf(t1)) return false;
    g1_mul_bls_x<false>(t2, p.x, p.y, &t1);          // t2 = [x^2]P
    if (g1_is_inf(t2)) return false;
    // phi(P) == -t2  <=>  beta*x*ZZ == X  and  y*ZZZ == -Y
    Fp bx, l, ry;
    fp_set(bx, c_k.fp_beta); FPM(bx, bx, p.x); FPM(l, bx, t2.zz);
    if (!fp_eq(l, t2.x)) return false;
    FPM(l, p.y, t2.zzz); fp_neg(ry, t2.y);
    return fp_eq(l, ry);
}
// the definition, kept for the unit test that pins the fast test: [r]P == infinity
static __device__ __noinline__ bool g1_in_subgroup_by_order(const G1Affine& p) {
    if (g1a_is_inf(p)) return true;
    Fr k; fr_set(k, c_k.fr_mod);
    G1 t; g1_mul(t, p, k);
    return g1_is_inf(t);
}

}  // namespace kb200
