/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY.  Not part of the shipped engine: only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may build, load or call anything under oracle/.
 *
 * kzo_field.h: prime-field arithmetic for BLS12-381 on the CPU, 64-bit limbs, Montgomery form.
 *   Fp: 381-bit base field, 6 limbs, R = 2^384.   Fr: 255-bit scalar field, 4 limbs, R = 2^256.
 *
 * Restates the semantics kateth gets from the (absent, unpinned) third-party crate blst 0.3.x through
 *   /root/reference/src/bls.rs:110-139  (blst_scalar_from_bendian, blst_scalar_fr_check, blst_fr_from_scalar)
 *   /root/reference/src/bls.rs:141-159  (blst_scalar_from_fr, blst_bendian/lendian_from_scalar)
 *   /root/reference/src/bls.rs:237-360  (blst_fr_add/sub/mul/cneg/eucl_inverse, lshift/rshift)
 * from the published definitions of the two primes.  blst's limb layout is the same
 * (little-endian u64 limbs, Montgomery with R = 2^256 / 2^384), so Fr::ONE / Fr::MAX in
 * /root/reference/src/bls.rs:87-106 are directly comparable (tests/test_oracle.py checks them).
 */
#ifndef KZO_FIELD_H
#define KZO_FIELD_H
#include <stdint.h>
#include <string.h>
#include "constants.h"

typedef unsigned __int128 u128;
typedef uint64_t u64;

/* ---------- generic little helpers on n-limb integers ---------- */
static inline int bn_is_zero(const u64 *a, int n) { u64 t = 0; for (int i = 0; i < n; i++) t |= a[i]; return t == 0; }
static inline int bn_eq(const u64 *a, const u64 *b, int n) { u64 t = 0; for (int i = 0; i < n; i++) t |= a[i] ^ b[i]; return t == 0; }
/* a >= b ? */
static inline int bn_geq(const u64 *a, const u64 *b, int n) {
    for (int i = n - 1; i >= 0; i--) { if (a[i] > b[i]) return 1; if (a[i] < b[i]) return 0; }
    return 1;
}
static inline u64 bn_add(u64 *r, const u64 *a, const u64 *b, int n) {
    u128 c = 0; for (int i = 0; i < n; i++) { c += (u128)a[i] + b[i]; r[i] = (u64)c; c >>= 64; } return (u64)c;
}
static inline u64 bn_sub(u64 *r, const u64 *a, const u64 *b, int n) {
    u64 br = 0; for (int i = 0; i < n; i++) { u128 t = (u128)a[i] - b[i] - br; r[i] = (u64)t; br = (u64)(t >> 64) & 1; } return br;
}
static inline void bn_shr1(u64 *a, int n, u64 top) {
    for (int i = 0; i < n - 1; i++) a[i] = (a[i] >> 1) | (a[i + 1] << 63);
    a[n - 1] = (a[n - 1] >> 1) | (top << 63);
}

/* ---------- Montgomery core, specialised by constant n through inlining ---------- */
static inline void mont_add(u64 *r, const u64 *a, const u64 *b, const u64 *m, int n) {
    u64 t[6], c = bn_add(t, a, b, n);
    if (c || bn_geq(t, m, n)) bn_sub(t, t, m, n);
    memcpy(r, t, 8 * n);
}
static inline void mont_sub(u64 *r, const u64 *a, const u64 *b, const u64 *m, int n) {
    u64 t[6]; if (bn_sub(t, a, b, n)) bn_add(t, t, m, n);
    memcpy(r, t, 8 * n);
}
static inline void mont_neg(u64 *r, const u64 *a, const u64 *m, int n) {
    if (bn_is_zero(a, n)) { memset(r, 0, 8 * n); return; }
    u64 t[6]; bn_sub(t, m, a, n); memcpy(r, t, 8 * n);
}
/* CIOS Montgomery product: r = a*b/2^(64n) mod m */
static inline void mont_mul(u64 *r, const u64 *a, const u64 *b, const u64 *m, u64 inv, int n) {
    u64 t[8] = {0};
    for (int i = 0; i < n; i++) {
        u128 c = 0;
        for (int j = 0; j < n; j++) { c += (u128)a[j] * b[i] + t[j]; t[j] = (u64)c; c >>= 64; }
        c += t[n]; t[n] = (u64)c; t[n + 1] = (u64)(c >> 64);
        u64 q = t[0] * inv;
        c = (u128)q * m[0] + t[0]; c >>= 64;
        for (int j = 1; j < n; j++) { c += (u128)q * m[j] + t[j]; t[j - 1] = (u64)c; c >>= 64; }
        c += t[n]; t[n - 1] = (u64)c; t[n] = t[n + 1] + (u64)(c >> 64);
    }
    if (t[n] || bn_geq(t, m, n)) bn_sub(t, t, m, n);
    memcpy(r, t, 8 * n);
}

/* ======================= Fp ======================= */
typedef struct { u64 l[6]; } fp_t;
static const fp_t FP_ZERO_C = {{0, 0, 0, 0, 0, 0}};
static inline void fp_add(fp_t *r, const fp_t *a, const fp_t *b) { mont_add(r->l, a->l, b->l, FP_P, 6); }
static inline void fp_sub(fp_t *r, const fp_t *a, const fp_t *b) { mont_sub(r->l, a->l, b->l, FP_P, 6); }
static inline void fp_neg(fp_t *r, const fp_t *a) { mont_neg(r->l, a->l, FP_P, 6); }
static inline void fp_mul(fp_t *r, const fp_t *a, const fp_t *b) { mont_mul(r->l, a->l, b->l, FP_P, FP_INV, 6); }
static inline void fp_sqr(fp_t *r, const fp_t *a) { mont_mul(r->l, a->l, a->l, FP_P, FP_INV, 6); }
static inline int fp_is_zero(const fp_t *a) { return bn_is_zero(a->l, 6); }
static inline int fp_eq(const fp_t *a, const fp_t *b) { return bn_eq(a->l, b->l, 6); }
static inline void fp_one(fp_t *r) { memcpy(r->l, FP_R, 48); }
static inline void fp_dbl(fp_t *r, const fp_t *a) { fp_add(r, a, a); }
/* plain integer (< p) -> Montgomery */
static inline void fp_to_mont(fp_t *r, const u64 *plain) { fp_t t; memcpy(t.l, plain, 48); fp_t r2; memcpy(r2.l, FP_R2, 48); fp_mul(r, &t, &r2); }
static inline void fp_from_mont(u64 *plain, const fp_t *a) { fp_t one = {{1, 0, 0, 0, 0, 0}}, t; fp_mul(&t, a, &one); memcpy(plain, t.l, 48); }
static inline void fp_set_u64(fp_t *r, u64 v) { u64 t[6] = {v, 0, 0, 0, 0, 0}; fp_to_mont(r, t); }
/* r = a^e, e an nl-limb plain integer */
static void fp_pow(fp_t *r, const fp_t *a, const u64 *e, int nl) {
    fp_t acc; fp_one(&acc);
    for (int i = nl * 64 - 1; i >= 0; i--) {
        fp_sqr(&acc, &acc);
        if ((e[i / 64] >> (i % 64)) & 1) fp_mul(&acc, &acc, a);
    }
    *r = acc;
}
static inline void fp_inv(fp_t *r, const fp_t *a) { fp_pow(r, a, FP_PM2, 6); }
/* square root for p = 3 mod 4; returns 1 iff a is a square (r*r == a) */
static int fp_sqrt(fp_t *r, const fp_t *a) {
    fp_t s, c; fp_pow(&s, a, FP_SQRT_EXP, 6); fp_sqr(&c, &s);
    *r = s; return fp_eq(&c, a);
}
/* 48-byte big-endian -> plain limbs; returns 0 if value >= p */
static int fp_from_be48(fp_t *r, const uint8_t *in) {
    u64 t[6];
    for (int i = 0; i < 6; i++) { u64 w = 0; for (int k = 0; k < 8; k++) w = (w << 8) | in[(5 - i) * 8 + k]; t[i] = w; }
    if (bn_geq(t, FP_P, 6)) return 0;
    fp_to_mont(r, t); return 1;
}
static void fp_to_be48(uint8_t *out, const fp_t *a) {
    u64 t[6]; fp_from_mont(t, a);
    for (int i = 0; i < 6; i++) for (int k = 0; k < 8; k++) out[(5 - i) * 8 + k] = (uint8_t)(t[i] >> (56 - 8 * k));
}
/* "lexicographically largest": plain(a) > (p-1)/2 */
static int fp_is_lex_largest(const fp_t *a) {
    u64 t[6]; fp_from_mont(t, a);
    return !bn_geq(FP_HALF, t, 6);
}

/* ======================= Fr ======================= */
typedef struct { u64 l[4]; } fr_t;
static inline void fr_add(fr_t *r, const fr_t *a, const fr_t *b) { mont_add(r->l, a->l, b->l, FR_R, 4); }
static inline void fr_sub(fr_t *r, const fr_t *a, const fr_t *b) { mont_sub(r->l, a->l, b->l, FR_R, 4); }
static inline void fr_neg(fr_t *r, const fr_t *a) { mont_neg(r->l, a->l, FR_R, 4); }
static inline void fr_mul(fr_t *r, const fr_t *a, const fr_t *b) { mont_mul(r->l, a->l, b->l, FR_R, FR_INV, 4); }
static inline int fr_is_zero(const fr_t *a) { return bn_is_zero(a->l, 4); }
static inline int fr_eq(const fr_t *a, const fr_t *b) { return bn_eq(a->l, b->l, 4); }
static inline void fr_one(fr_t *r) { memcpy(r->l, FR_ONE, 32); }
static inline void fr_zero(fr_t *r) { memset(r->l, 0, 32); }
/* plain integer (any 256-bit value; reduced by the Montgomery product as blst_fr_from_scalar does) -> Montgomery */
static inline void fr_to_mont(fr_t *r, const u64 *plain) { fr_t t, r2; memcpy(t.l, plain, 32); memcpy(r2.l, FR_R2, 32); fr_mul(r, &t, &r2); }
static inline void fr_from_mont(u64 *plain, const fr_t *a) { fr_t one = {{1, 0, 0, 0}}, t; fr_mul(&t, a, &one); memcpy(plain, t.l, 32); }
static inline void fr_set_u64(fr_t *r, u64 v) { u64 t[4] = {v, 0, 0, 0}; fr_to_mont(r, t); }
static inline void be32_to_limbs(u64 *t, const uint8_t *in) {
    for (int i = 0; i < 4; i++) { u64 w = 0; for (int k = 0; k < 8; k++) w = (w << 8) | in[(3 - i) * 8 + k]; t[i] = w; }
}
/* Fr::from_be_slice (src/bls.rs:130-139): 0 if value >= r (NotInFiniteField) */
static int fr_from_be32(fr_t *r, const uint8_t *in) {
    u64 t[4]; be32_to_limbs(t, in);
    if (bn_geq(t, FR_R, 4)) return 0;
    fr_to_mont(r, t); return 1;
}
/* Fr::hash_to's tail (src/bls.rs:198-203): BE bytes -> integer -> Montgomery WITHOUT the range check.
 * A 256-bit value v enters the Montgomery product as-is: result = v * R^2 / R mod r = v*R mod r, i.e. v mod r. */
static void fr_from_be32_reduce(fr_t *r, const uint8_t *in) { u64 t[4]; be32_to_limbs(t, in); fr_to_mont(r, t); }
static void fr_to_be32(uint8_t *out, const fr_t *a) {
    u64 t[4]; fr_from_mont(t, a);
    for (int i = 0; i < 4; i++) for (int k = 0; k < 8; k++) out[(3 - i) * 8 + k] = (uint8_t)(t[i] >> (56 - 8 * k));
}
static void fr_to_le32(uint8_t *out, const fr_t *a) {
    u64 t[4]; fr_from_mont(t, a);
    for (int i = 0; i < 4; i++) for (int k = 0; k < 8; k++) out[i * 8 + k] = (uint8_t)(t[i] >> (8 * k));
}
/* Field inverse by binary extended Euclid on plain integers (stands in for blst_fr_eucl_inverse,
 * src/bls.rs:297-312 -- one full inversion per division, as the reference does).  a != 0. */
static void fr_inv(fr_t *r, const fr_t *a) {
    u64 u[4], v[4], x1[4] = {1, 0, 0, 0}, x2[4] = {0, 0, 0, 0};
    fr_from_mont(u, a); memcpy(v, FR_R, 32);
    static const u64 ONE[4] = {1, 0, 0, 0};
    while (!bn_eq(u, ONE, 4) && !bn_eq(v, ONE, 4)) {
        while (!(u[0] & 1)) {
            bn_shr1(u, 4, 0);
            if (x1[0] & 1) { u64 c = bn_add(x1, x1, FR_R, 4); bn_shr1(x1, 4, c); } else bn_shr1(x1, 4, 0);
        }
        while (!(v[0] & 1)) {
            bn_shr1(v, 4, 0);
            if (x2[0] & 1) { u64 c = bn_add(x2, x2, FR_R, 4); bn_shr1(x2, 4, c); } else bn_shr1(x2, 4, 0);
        }
        if (bn_geq(u, v, 4)) { bn_sub(u, u, v, 4); mont_sub(x1, x1, x2, FR_R, 4); }
        else { bn_sub(v, v, u, 4); mont_sub(x2, x2, x1, FR_R, 4); }
    }
    fr_to_mont(r, bn_eq(u, ONE, 4) ? x1 : x2);
}
/* Fr Div (src/bls.rs:297-312) */
static inline void fr_div(fr_t *r, const fr_t *a, const fr_t *b) { fr_t i; fr_inv(&i, b); fr_mul(r, a, &i); }
static inline int fr_is_odd(const fr_t *a) { u64 t[4]; fr_from_mont(t, a); return (int)(t[0] & 1); }
static inline void fr_shr1(fr_t *a) { u64 t[4]; fr_from_mont(t, a); bn_shr1(t, 4, 0); fr_to_mont(a, t); }
/* Fr::pow (src/bls.rs:169-187), INCLUDING its quirk: pow(x, 0) == x (the loop is skipped and out*tmp = x*1). */
static void fr_pow_kateth(fr_t *r, const fr_t *base, const fr_t *power_in) {
    fr_t power = *power_in, out = *base, tmp, one, zero;
    fr_one(&one); fr_zero(&zero); tmp = one;
    while (!fr_eq(&power, &one) && !fr_eq(&power, &zero)) {
        if (fr_is_odd(&power)) { fr_mul(&tmp, &out, &tmp); fr_sub(&power, &power, &one); }
        fr_mul(&out, &out, &out);
        fr_shr1(&power);
    }
    fr_mul(r, &out, &tmp);
}
#endif
