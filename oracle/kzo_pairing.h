/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY (see kzo_field.h).
 *
 * kzo_pairing.h: Fp2, the group G2 (on the twist y^2 = x^3 + 4(1+u)), Fp12 and the pairing-product test
 *   e(-a1, a2) * e(b1, b2) == 1      (/root/reference/src/bls.rs:572-598, bls::verify_pairings)
 * plus what P2 needs at load time and in verify_proof_inner (/root/reference/src/bls.rs:554-570,
 * /root/reference/src/kzg/setup.rs:84-94): decompress (96-byte ZCash form), subgroup test, add, neg, mul.
 *
 * Deliberately the *textbook* construction, chosen to be easy to audit rather than fast:
 *   - Fp12 is Fp2[w]/(w^6 - (1+u)) with schoolbook products (no tower, no Frobenius, no inversion);
 *   - the Miller loop walks |x| = 0xd201000000010000 with the G2 point kept in AFFINE twist coordinates;
 *     a line through T' with twist-slope L evaluated at P = (xP, yP) in G1 is, after scaling by w^3
 *     (an element of the subfield Fp4, wiped out by the final exponentiation),
 *         (L*xT' - yT')  +  (-L*xP) * w^2  +  yP * w^3;
 *   - the final exponentiation is a plain square-and-multiply by (p^12 - 1)/r  (constants.h: FINAL_EXP);
 *   - x is negative, so every f is the inverse of the true Miller value; a product of pairings is 1 iff
 *     the product of inverses is 1, so no conjugation is needed for an equality-to-one test;
 *   - a pair with an infinite member contributes 1.
 * Verdicts are booleans, so agreement with blst's optimised pairing is exact by bilinearity/non-degeneracy.
 */
#ifndef KZO_PAIRING_H
#define KZO_PAIRING_H
#include "kzo_g1.h"

/* ---------------- Fp2 = Fp[u]/(u^2+1) ---------------- */
typedef struct { fp_t c0, c1; } fp2_t;
static inline void fp2_add(fp2_t *r, const fp2_t *a, const fp2_t *b) { fp_add(&r->c0, &a->c0, &b->c0); fp_add(&r->c1, &a->c1, &b->c1); }
static inline void fp2_sub(fp2_t *r, const fp2_t *a, const fp2_t *b) { fp_sub(&r->c0, &a->c0, &b->c0); fp_sub(&r->c1, &a->c1, &b->c1); }
static inline void fp2_neg(fp2_t *r, const fp2_t *a) { fp_neg(&r->c0, &a->c0); fp_neg(&r->c1, &a->c1); }
static inline void fp2_dbl(fp2_t *r, const fp2_t *a) { fp2_add(r, a, a); }
static inline int fp2_is_zero(const fp2_t *a) { return fp_is_zero(&a->c0) && fp_is_zero(&a->c1); }
static inline int fp2_eq(const fp2_t *a, const fp2_t *b) { return fp_eq(&a->c0, &b->c0) && fp_eq(&a->c1, &b->c1); }
static inline void fp2_zero(fp2_t *r) { memset(r, 0, sizeof *r); }
static inline void fp2_one(fp2_t *r) { fp_one(&r->c0); memset(&r->c1, 0, sizeof(fp_t)); }
static void fp2_mul(fp2_t *r, const fp2_t *a, const fp2_t *b) {
    fp_t t0, t1, t2, t3;
    fp_mul(&t0, &a->c0, &b->c0); fp_mul(&t1, &a->c1, &b->c1);
    fp_mul(&t2, &a->c0, &b->c1); fp_mul(&t3, &a->c1, &b->c0);
    fp_sub(&r->c0, &t0, &t1); fp_add(&r->c1, &t2, &t3);
}
static inline void fp2_sqr(fp2_t *r, const fp2_t *a) { fp2_mul(r, a, a); }
static inline void fp2_mul_fp(fp2_t *r, const fp2_t *a, const fp_t *b) { fp_mul(&r->c0, &a->c0, b); fp_mul(&r->c1, &a->c1, b); }
/* multiply by xi = 1 + u */
static inline void fp2_mul_xi(fp2_t *r, const fp2_t *a) { fp_t t0, t1; fp_sub(&t0, &a->c0, &a->c1); fp_add(&t1, &a->c0, &a->c1); r->c0 = t0; r->c1 = t1; }
static void fp2_inv(fp2_t *r, const fp2_t *a) {
    fp_t n, t; fp_sqr(&n, &a->c0); fp_sqr(&t, &a->c1); fp_add(&n, &n, &t); fp_inv(&n, &n);
    fp_mul(&r->c0, &a->c0, &n); fp_mul(&t, &a->c1, &n); fp_neg(&r->c1, &t);
}
static void fp2_pow(fp2_t *r, const fp2_t *a, const u64 *e, int nl) {
    fp2_t acc; fp2_one(&acc);
    for (int i = nl * 64 - 1; i >= 0; i--) { fp2_sqr(&acc, &acc); if ((e[i / 64] >> (i % 64)) & 1) fp2_mul(&acc, &acc, a); }
    *r = acc;
}
/* square root in Fp2 for p = 3 mod 4 (Adj--Rodriguez-Henriquez, alg. 9); returns 1 iff a is a square */
static int fp2_sqrt(fp2_t *r, const fp2_t *a) {
    if (fp2_is_zero(a)) { fp2_zero(r); return 1; }
    fp2_t a1, alpha, x0, m1, t;
    fp2_pow(&a1, a, FP_PM3D4, 6);
    fp2_sqr(&alpha, &a1); fp2_mul(&alpha, &alpha, a);
    fp2_mul(&x0, &a1, a);
    fp2_one(&m1); fp2_neg(&m1, &m1);
    if (fp2_eq(&alpha, &m1)) { /* x = u * x0 */
        fp_neg(&t.c0, &x0.c1); t.c1 = x0.c0; *r = t;
    } else {
        fp2_t b; fp2_one(&b); fp2_add(&b, &b, &alpha); fp2_pow(&b, &b, FP_HALF, 6); fp2_mul(r, &b, &x0);
    }
    fp2_sqr(&t, r); return fp2_eq(&t, a);
}
/* lexicographic order on Fp2 as in the ZCash format: compare c1 first, then c0 */
static int fp2_is_lex_largest(const fp2_t *a) {
    if (!fp_is_zero(&a->c1)) return fp_is_lex_largest(&a->c1);
    return fp_is_lex_largest(&a->c0);
}

/* ---------------- G2 (Jacobian over Fp2) ---------------- */
typedef struct { fp2_t x, y, z; } g2_t;
typedef struct { fp2_t x, y; int inf; } g2a_t;
static inline void g2_set_inf(g2_t *r) { memset(r, 0, sizeof *r); }
static inline int g2_is_inf(const g2_t *a) { return fp2_is_zero(&a->z); }
static inline void g2_from_affine(g2_t *r, const g2a_t *a) { if (a->inf) { g2_set_inf(r); return; } r->x = a->x; r->y = a->y; fp2_one(&r->z); }
static void g2_generator(g2_t *r) {
    fp_to_mont(&r->x.c0, G2_GEN_X0); fp_to_mont(&r->x.c1, G2_GEN_X1); fp_to_mont(&r->y.c0, G2_GEN_Y0); fp_to_mont(&r->y.c1, G2_GEN_Y1); fp2_one(&r->z);
}
static inline void g2_neg(g2_t *r, const g2_t *a) { r->x = a->x; fp2_neg(&r->y, &a->y); r->z = a->z; }
static void g2_dbl(g2_t *r, const g2_t *p) {
    if (g2_is_inf(p)) { g2_set_inf(r); return; }
    fp2_t a, b, c, d, e, f, t, x3, y3, z3;
    fp2_sqr(&a, &p->x); fp2_sqr(&b, &p->y); fp2_sqr(&c, &b);
    fp2_add(&t, &p->x, &b); fp2_sqr(&t, &t); fp2_sub(&t, &t, &a); fp2_sub(&t, &t, &c); fp2_dbl(&d, &t);
    fp2_dbl(&e, &a); fp2_add(&e, &e, &a); fp2_sqr(&f, &e);
    fp2_dbl(&t, &d); fp2_sub(&x3, &f, &t);
    fp2_sub(&t, &d, &x3); fp2_mul(&y3, &e, &t);
    fp2_dbl(&t, &c); fp2_dbl(&t, &t); fp2_dbl(&t, &t); fp2_sub(&y3, &y3, &t);
    fp2_mul(&z3, &p->y, &p->z); fp2_dbl(&z3, &z3);
    r->x = x3; r->y = y3; r->z = z3;
}
static void g2_add(g2_t *r, const g2_t *p, const g2_t *q) {
    if (g2_is_inf(p)) { *r = *q; return; }
    if (g2_is_inf(q)) { *r = *p; return; }
    fp2_t z1z1, z2z2, u1, u2, s1, s2, h, i, j, rr, v, t, x3, y3, z3;
    fp2_sqr(&z1z1, &p->z); fp2_sqr(&z2z2, &q->z);
    fp2_mul(&u1, &p->x, &z2z2); fp2_mul(&u2, &q->x, &z1z1);
    fp2_mul(&s1, &p->y, &q->z); fp2_mul(&s1, &s1, &z2z2);
    fp2_mul(&s2, &q->y, &p->z); fp2_mul(&s2, &s2, &z1z1);
    if (fp2_eq(&u1, &u2)) { if (fp2_eq(&s1, &s2)) g2_dbl(r, p); else g2_set_inf(r); return; }
    fp2_sub(&h, &u2, &u1); fp2_dbl(&i, &h); fp2_sqr(&i, &i); fp2_mul(&j, &h, &i);
    fp2_sub(&rr, &s2, &s1); fp2_dbl(&rr, &rr); fp2_mul(&v, &u1, &i);
    fp2_sqr(&x3, &rr); fp2_sub(&x3, &x3, &j); fp2_dbl(&t, &v); fp2_sub(&x3, &x3, &t);
    fp2_sub(&t, &v, &x3); fp2_mul(&y3, &rr, &t); fp2_mul(&t, &s1, &j); fp2_dbl(&t, &t); fp2_sub(&y3, &y3, &t);
    fp2_add(&z3, &p->z, &q->z); fp2_sqr(&z3, &z3); fp2_sub(&z3, &z3, &z1z1); fp2_sub(&z3, &z3, &z2z2); fp2_mul(&z3, &z3, &h);
    r->x = x3; r->y = y3; r->z = z3;
}
static void g2_mul_limbs(g2_t *r, const g2_t *p, const u64 *k, int nbits) {
    g2_t acc; g2_set_inf(&acc);
    for (int i = nbits - 1; i >= 0; i--) { g2_dbl(&acc, &acc); if ((k[i / 64] >> (i % 64)) & 1) g2_add(&acc, &acc, p); }
    *r = acc;
}
static void g2_mul_fr(g2_t *r, const g2_t *p, const fr_t *s) { u64 k[4]; fr_from_mont(k, s); g2_mul_limbs(r, p, k, 255); }
static void g2_to_affine(g2a_t *r, const g2_t *p) {
    if (g2_is_inf(p)) { memset(r, 0, sizeof *r); r->inf = 1; return; }
    fp2_t zi, zi2, zi3; fp2_inv(&zi, &p->z); fp2_sqr(&zi2, &zi); fp2_mul(&zi3, &zi2, &zi);
    fp2_mul(&r->x, &p->x, &zi2); fp2_mul(&r->y, &p->y, &zi3); r->inf = 0;
}
/* P2::decompress (src/bls.rs:505-531 instantiated at :554-570): 96 bytes = x.c1 (48 BE, flags on top) || x.c0 */
static int g2_decompress(g2_t *r, const uint8_t in[96]) {
    uint8_t b0 = in[0];
    if (!(b0 & 0x80)) return KZ_G1_INVALID_ENCODING;
    if (b0 & 0x40) {
        if (b0 & 0x3f) return KZ_G1_INVALID_ENCODING;
        for (int i = 1; i < 96; i++) if (in[i]) return KZ_G1_INVALID_ENCODING;
        g2_set_inf(r); return KZ_OK;
    }
    uint8_t tmp[48]; memcpy(tmp, in, 48); tmp[0] &= 0x1f;
    g2a_t a; a.inf = 0;
    if (!fp_from_be48(&a.x.c1, tmp)) return KZ_G1_INVALID_ENCODING;
    if (!fp_from_be48(&a.x.c0, in + 48)) return KZ_G1_INVALID_ENCODING;
    fp2_t y2, b; fp2_sqr(&y2, &a.x); fp2_mul(&y2, &y2, &a.x);
    fp_set_u64(&b.c0, 4); fp_set_u64(&b.c1, 4); fp2_add(&y2, &y2, &b);        /* b' = 4(1+u) */
    if (!fp2_sqrt(&a.y, &y2)) return KZ_G1_NOT_ON_CURVE;
    if (fp2_is_lex_largest(&a.y) != ((b0 >> 5) & 1)) fp2_neg(&a.y, &a.y);
    g2_t p, t; g2_from_affine(&p, &a);
    g2_mul_limbs(&t, &p, FR_R, 255);
    if (!g2_is_inf(&t)) return KZ_G1_NOT_IN_GROUP;
    *r = p; return KZ_OK;
}

/* ---------------- Fp12 = Fp2[w]/(w^6 - xi) ---------------- */
typedef struct { fp2_t c[6]; } fp12_t;
static void fp12_one(fp12_t *r) { memset(r, 0, sizeof *r); fp2_one(&r->c[0]); }
static int fp12_is_one(const fp12_t *a) {
    fp2_t one; fp2_one(&one); if (!fp2_eq(&a->c[0], &one)) return 0;
    for (int i = 1; i < 6; i++) if (!fp2_is_zero(&a->c[i])) return 0;
    return 1;
}
static void fp12_mul(fp12_t *r, const fp12_t *a, const fp12_t *b) {
    fp2_t acc[11]; memset(acc, 0, sizeof acc);
    for (int i = 0; i < 6; i++) {
        if (fp2_is_zero(&a->c[i])) continue;
        for (int j = 0; j < 6; j++) { fp2_t t; fp2_mul(&t, &a->c[i], &b->c[j]); fp2_add(&acc[i + j], &acc[i + j], &t); }
    }
    for (int k = 6; k < 11; k++) { fp2_t t; fp2_mul_xi(&t, &acc[k]); fp2_add(&acc[k - 6], &acc[k - 6], &t); }
    memcpy(r->c, acc, sizeof r->c);
}

/* One Miller loop f_{|x|,Q}(P) with P in G1 (affine, over Fp) and Q in G2 (affine, on the twist). */
static void miller_loop(fp12_t *f, const g1a_t *P, const g2a_t *Q) {
    fp12_one(f);
    if (P->inf || Q->inf) return;
    fp2_t tx = Q->x, ty = Q->y;
    fp_t nxp; fp_neg(&nxp, &P->x);
    int started = 0;
    for (int i = 63; i >= 0; i--) {
        int bit = (int)((BLS_X_ABS >> i) & 1);
        if (!started) { if (bit) started = 1; continue; }
        /* doubling step: slope L = 3 x^2 / (2 y) */
        fp2_t L, t, n, line0; fp12_t l;
        fp2_sqr(&n, &tx); fp2_dbl(&t, &n); fp2_add(&n, &n, &t);
        fp2_dbl(&t, &ty); fp2_inv(&t, &t); fp2_mul(&L, &n, &t);
        fp2_mul(&line0, &L, &tx); fp2_sub(&line0, &line0, &ty);
        memset(&l, 0, sizeof l); l.c[0] = line0; fp2_mul_fp(&l.c[2], &L, &nxp); l.c[3].c0 = P->y;
        fp12_mul(f, f, f); fp12_mul(f, &l, f);
        fp2_t x3, y3; fp2_sqr(&x3, &L); fp2_sub(&x3, &x3, &tx); fp2_sub(&x3, &x3, &tx);
        fp2_sub(&t, &tx, &x3); fp2_mul(&y3, &L, &t); fp2_sub(&y3, &y3, &ty);
        tx = x3; ty = y3;
        if (bit) {
            /* addition step T + Q: slope L = (yT - yQ)/(xT - xQ); T != +-Q inside the loop for points of order r */
            fp2_sub(&n, &ty, &Q->y); fp2_sub(&t, &tx, &Q->x); fp2_inv(&t, &t); fp2_mul(&L, &n, &t);
            fp2_mul(&line0, &L, &tx); fp2_sub(&line0, &line0, &ty);
            memset(&l, 0, sizeof l); l.c[0] = line0; fp2_mul_fp(&l.c[2], &L, &nxp); l.c[3].c0 = P->y;
            fp12_mul(f, &l, f);
            fp2_sqr(&x3, &L); fp2_sub(&x3, &x3, &tx); fp2_sub(&x3, &x3, &Q->x);
            fp2_sub(&t, &tx, &x3); fp2_mul(&y3, &L, &t); fp2_sub(&y3, &y3, &ty);
            tx = x3; ty = y3;
        }
    }
}
static void final_exp(fp12_t *r, const fp12_t *a) {
    fp12_t acc; fp12_one(&acc); int started = 0;
    for (int i = FINAL_EXP_LIMBS * 64 - 1; i >= 0; i--) {
        int bit = (int)((FINAL_EXP[i / 64] >> (i % 64)) & 1);
        if (started) fp12_mul(&acc, &acc, &acc);
        if (bit) { if (started) fp12_mul(&acc, &acc, a); else { acc = *a; started = 1; } }
    }
    *r = acc;
}
/* bls::verify_pairings (src/bls.rs:572-598): e(-a1, a2) * e(b1, b2) == 1 */
static int verify_pairings(const g1_t *a1, const g2_t *a2, const g1_t *b1, const g2_t *b2) {
    g1_t na1; g1_neg(&na1, a1);
    g1a_t pa, pb; g2a_t qa, qb;
    g1_to_affine(&pa, &na1); g1_to_affine(&pb, b1); g2_to_affine(&qa, a2); g2_to_affine(&qb, b2);
    fp12_t e1, e2, prod, ex;
    miller_loop(&e1, &pa, &qa); miller_loop(&e2, &pb, &qb);
    fp12_mul(&prod, &e1, &e2); final_exp(&ex, &prod);
    return fp12_is_one(&ex);
}
#endif
