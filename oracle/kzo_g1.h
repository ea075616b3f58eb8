/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY (see kzo_field.h).
 *
 * kzo_g1.h: the group G1 of BLS12-381 (y^2 = x^3 + 4 over Fp) in Jacobian coordinates, plus the ZCash
 * 48-byte compressed encoding and SHA-256.
 *
 * Restates what kateth's P1 wrapper obtains from blst:
 *   /root/reference/src/bls.rs:452-489  Add (complete), Neg, Mul<Fr> (255-bit scalar)
 *   /root/reference/src/bls.rs:406-413  lincomb (naive: n scalar multiplications)
 *   /root/reference/src/bls.rs:416-437  lincomb_pippenger (bases to affine on every call, LE scalars, 255 bits)
 *   /root/reference/src/bls.rs:491-503  Compress   (blst_p1_compress)
 *   /root/reference/src/bls.rs:505-531  Decompress (blst_p1_uncompress: BAD_ENCODING / NOT_ON_CURVE, then
 *                                        blst_p1_affine_in_g1 -> NotInGroup)
 *   /root/reference/src/bls.rs:189-205  blst_sha256
 * Group elements are unique, so any correct addition law yields the same compressed bytes as blst.
 */
#ifndef KZO_G1_H
#define KZO_G1_H
#include <stdlib.h>
#include "kzo_field.h"

/* status codes shared with include/kateth_b200.h (SURVEY.md 8b "Errors") */
enum { KZ_OK = 0, KZ_INVALID_LEN = 1, KZ_INVALID_FIELD_ELEMENT = 2, KZ_G1_INVALID_ENCODING = 3, KZ_G1_NOT_ON_CURVE = 4,
       KZ_G1_NOT_IN_GROUP = 5, KZ_SCALAR_INVALID_ENCODING = 6, KZ_SCALAR_NOT_IN_FIELD = 7 };

typedef struct { fp_t x, y, z; } g1_t;          /* Jacobian; z == 0 <=> infinity */
typedef struct { fp_t x, y; int inf; } g1a_t;   /* affine */

static inline void g1_set_inf(g1_t *r) { memset(r, 0, sizeof *r); }
static inline int g1_is_inf(const g1_t *a) { return fp_is_zero(&a->z); }
static inline void g1_from_affine(g1_t *r, const g1a_t *a) { if (a->inf) { g1_set_inf(r); return; } r->x = a->x; r->y = a->y; fp_one(&r->z); }
static void g1_generator(g1_t *r) { fp_to_mont(&r->x, G1_GEN_X); fp_to_mont(&r->y, G1_GEN_Y); fp_one(&r->z); }
static inline void g1_neg(g1_t *r, const g1_t *a) { r->x = a->x; fp_neg(&r->y, &a->y); r->z = a->z; }

static void g1_dbl(g1_t *r, const g1_t *p) {
    if (g1_is_inf(p)) { g1_set_inf(r); return; }
    fp_t a, b, c, d, e, f, t, x3, y3, z3;
    fp_sqr(&a, &p->x); fp_sqr(&b, &p->y); fp_sqr(&c, &b);
    fp_add(&t, &p->x, &b); fp_sqr(&t, &t); fp_sub(&t, &t, &a); fp_sub(&t, &t, &c); fp_dbl(&d, &t);
    fp_dbl(&e, &a); fp_add(&e, &e, &a);
    fp_sqr(&f, &e);
    fp_dbl(&t, &d); fp_sub(&x3, &f, &t);
    fp_sub(&t, &d, &x3); fp_mul(&y3, &e, &t);
    fp_dbl(&t, &c); fp_dbl(&t, &t); fp_dbl(&t, &t); fp_sub(&y3, &y3, &t);
    fp_mul(&z3, &p->y, &p->z); fp_dbl(&z3, &z3);
    r->x = x3; r->y = y3; r->z = z3;
}
/* complete addition (branches to doubling / infinity like blst_p1_add's contract) */
static void g1_add(g1_t *r, const g1_t *p, const g1_t *q) {
    if (g1_is_inf(p)) { *r = *q; return; }
    if (g1_is_inf(q)) { *r = *p; return; }
    fp_t z1z1, z2z2, u1, u2, s1, s2, h, i, j, rr, v, t, x3, y3, z3;
    fp_sqr(&z1z1, &p->z); fp_sqr(&z2z2, &q->z);
    fp_mul(&u1, &p->x, &z2z2); fp_mul(&u2, &q->x, &z1z1);
    fp_mul(&s1, &p->y, &q->z); fp_mul(&s1, &s1, &z2z2);
    fp_mul(&s2, &q->y, &p->z); fp_mul(&s2, &s2, &z1z1);
    if (fp_eq(&u1, &u2)) { if (fp_eq(&s1, &s2)) g1_dbl(r, p); else g1_set_inf(r); return; }
    fp_sub(&h, &u2, &u1);
    fp_dbl(&i, &h); fp_sqr(&i, &i);
    fp_mul(&j, &h, &i);
    fp_sub(&rr, &s2, &s1); fp_dbl(&rr, &rr);
    fp_mul(&v, &u1, &i);
    fp_sqr(&x3, &rr); fp_sub(&x3, &x3, &j); fp_dbl(&t, &v); fp_sub(&x3, &x3, &t);
    fp_sub(&t, &v, &x3); fp_mul(&y3, &rr, &t); fp_mul(&t, &s1, &j); fp_dbl(&t, &t); fp_sub(&y3, &y3, &t);
    fp_add(&z3, &p->z, &q->z); fp_sqr(&z3, &z3); fp_sub(&z3, &z3, &z1z1); fp_sub(&z3, &z3, &z2z2); fp_mul(&z3, &z3, &h);
    r->x = x3; r->y = y3; r->z = z3;
}
static void g1_add_affine(g1_t *r, const g1_t *p, const g1a_t *q) { g1_t t; g1_from_affine(&t, q); g1_add(r, p, &t); }

/* scalar given as 4 plain little-endian limbs, nbits significant bits (blst_p1_mult(..., 255)) */
static void g1_mul_limbs(g1_t *r, const g1_t *p, const u64 *k, int nbits) {
    g1_t acc; g1_set_inf(&acc);
    for (int i = nbits - 1; i >= 0; i--) {
        g1_dbl(&acc, &acc);
        if ((k[i / 64] >> (i % 64)) & 1) g1_add(&acc, &acc, p);
    }
    *r = acc;
}
/* P1 * Fr (src/bls.rs:477-489) */
static void g1_mul_fr(g1_t *r, const g1_t *p, const fr_t *s) { u64 k[4]; fr_from_mont(k, s); g1_mul_limbs(r, p, k, 255); }

static void g1_to_affine(g1a_t *r, const g1_t *p) {
    if (g1_is_inf(p)) { memset(r, 0, sizeof *r); r->inf = 1; return; }
    fp_t zi, zi2, zi3; fp_inv(&zi, &p->z); fp_sqr(&zi2, &zi); fp_mul(&zi3, &zi2, &zi);
    fp_mul(&r->x, &p->x, &zi2); fp_mul(&r->y, &p->y, &zi3); r->inf = 0;
}
/* batch Jacobian -> affine with one inversion (what p1_affines::from does on every lincomb_pippenger call) */
static void g1_batch_to_affine(g1a_t *out, const g1_t *in, size_t n) {
    fp_t *pre = (fp_t *)malloc(sizeof(fp_t) * (n + 1)); fp_t acc; fp_one(&acc);
    for (size_t i = 0; i < n; i++) { pre[i] = acc; if (!g1_is_inf(&in[i])) fp_mul(&acc, &acc, &in[i].z); }
    fp_t inv; fp_inv(&inv, &acc);
    for (size_t i = n; i-- > 0;) {
        if (g1_is_inf(&in[i])) { memset(&out[i], 0, sizeof out[i]); out[i].inf = 1; continue; }
        fp_t zi, zi2, zi3; fp_mul(&zi, &inv, &pre[i]); fp_mul(&inv, &inv, &in[i].z);
        fp_sqr(&zi2, &zi); fp_mul(&zi3, &zi2, &zi);
        fp_mul(&out[i].x, &in[i].x, &zi2); fp_mul(&out[i].y, &in[i].y, &zi3); out[i].inf = 0;
    }
    free(pre);
}
static int g1_eq(const g1_t *a, const g1_t *b) {
    g1a_t x, y; g1_to_affine(&x, a); g1_to_affine(&y, b);
    if (x.inf || y.inf) return x.inf == y.inf;
    return fp_eq(&x.x, &y.x) && fp_eq(&x.y, &y.y);
}

/* ZCash compressed encoding: bit7 = compressed, bit6 = infinity, bit5 = y is the lexicographically larger root */
static void g1_compress(uint8_t out[48], const g1_t *p) {
    g1a_t a; g1_to_affine(&a, p);
    if (a.inf) { memset(out, 0, 48); out[0] = 0xc0; return; }
    fp_to_be48(out, &a.x); out[0] |= 0x80; if (fp_is_lex_largest(&a.y)) out[0] |= 0x20;
}
/* curve membership only (blst_p1_uncompress); returns KZ_OK / KZ_G1_INVALID_ENCODING / KZ_G1_NOT_ON_CURVE */
static int g1_uncompress(g1a_t *r, const uint8_t in[48]) {
    uint8_t b0 = in[0];
    if (!(b0 & 0x80)) return KZ_G1_INVALID_ENCODING;                  /* uncompressed form not accepted here */
    if (b0 & 0x40) {                                                   /* infinity: every other bit must be 0 */
        if (b0 & 0x3f) return KZ_G1_INVALID_ENCODING;
        for (int i = 1; i < 48; i++) if (in[i]) return KZ_G1_INVALID_ENCODING;
        memset(r, 0, sizeof *r); r->inf = 1; return KZ_OK;
    }
    uint8_t tmp[48]; memcpy(tmp, in, 48); tmp[0] &= 0x1f;
    fp_t x; if (!fp_from_be48(&x, tmp)) return KZ_G1_INVALID_ENCODING; /* x >= p */
    fp_t y2, b, y; fp_sqr(&y2, &x); fp_mul(&y2, &y2, &x); fp_set_u64(&b, 4); fp_add(&y2, &y2, &b);
    if (!fp_sqrt(&y, &y2)) return KZ_G1_NOT_ON_CURVE;
    if (fp_is_lex_largest(&y) != ((b0 >> 5) & 1)) fp_neg(&y, &y);
    r->x = x; r->y = y; r->inf = 0; return KZ_OK;
}
/* subgroup test by definition: [r]P == infinity */
static int g1_in_subgroup(const g1a_t *a) {
    if (a->inf) return 1;
    g1_t p, t; g1_from_affine(&p, a); g1_mul_limbs(&t, &p, FR_R, 255); return g1_is_inf(&t);
}
/* P1::decompress (src/bls.rs:505-531) */
static int g1_decompress(g1_t *r, const uint8_t in[48]) {
    g1a_t a; int st = g1_uncompress(&a, in); if (st) return st;
    if (!g1_in_subgroup(&a)) return KZ_G1_NOT_IN_GROUP;
    g1_from_affine(r, &a); return KZ_OK;
}

/* P1::lincomb (src/bls.rs:406-413): sum of n independent scalar multiplications */
static void g1_lincomb_naive(g1_t *r, const g1_t *pts, const fr_t *sc, size_t n) {
    g1_t acc, t; g1_set_inf(&acc);
    for (size_t i = 0; i < n; i++) { g1_mul_fr(&t, &pts[i], &sc[i]); g1_add(&acc, &acc, &t); }
    *r = acc;
}
/* P1::lincomb_pippenger (src/bls.rs:416-437): bases converted to affine per call, scalars serialised LE,
 * then a bucket-method MSM over 255 bits (blst's p1_affines::mult).  Window c, signed digits. */
static void g1_lincomb_pippenger(g1_t *r, const g1_t *pts, const fr_t *sc, size_t n) {
    if (n == 0) { g1_set_inf(r); return; }
    g1a_t *aff = (g1a_t *)malloc(sizeof(g1a_t) * n);
    g1_batch_to_affine(aff, pts, n);
    uint8_t *le = (uint8_t *)malloc(32 * n);
    for (size_t i = 0; i < n; i++) fr_to_le32(le + 32 * i, &sc[i]);
    int c = n >= 2048 ? 10 : (n >= 256 ? 8 : (n >= 32 ? 5 : 3));
    int nwin = (255 + c) / c;                       /* one spare window absorbs the signed-digit carry */
    size_t nb = (size_t)1 << (c - 1);
    int32_t *dig = (int32_t *)malloc(sizeof(int32_t) * n * nwin);
    for (size_t i = 0; i < n; i++) {
        u64 k[5] = {0}; memcpy(k, le + 32 * i, 32);
        int carry = 0;
        for (int w = 0; w < nwin; w++) {
            int bit = w * c; u64 v = (k[bit / 64] >> (bit % 64));
            if (bit % 64 + c > 64) v |= k[bit / 64 + 1] << (64 - bit % 64);
            int d = (int)(v & (((u64)1 << c) - 1)) + carry;
            if (d > (int)nb) { d -= (1 << c); carry = 1; } else carry = 0;
            dig[i * nwin + w] = d;
        }
    }
    g1_t *bk = (g1_t *)malloc(sizeof(g1_t) * nb);
    g1_t total; g1_set_inf(&total);
    for (int w = nwin - 1; w >= 0; w--) {
        for (int k = 0; k < c; k++) g1_dbl(&total, &total);
        for (size_t b = 0; b < nb; b++) g1_set_inf(&bk[b]);
        for (size_t i = 0; i < n; i++) {
            int d = dig[i * nwin + w]; if (d == 0 || aff[i].inf) continue;
            g1a_t q = aff[i]; if (d < 0) { fp_neg(&q.y, &q.y); d = -d; }
            g1_add_affine(&bk[d - 1], &bk[d - 1], &q);
        }
        g1_t run, sum; g1_set_inf(&run); g1_set_inf(&sum);
        for (size_t b = nb; b-- > 0;) { g1_add(&run, &run, &bk[b]); g1_add(&sum, &sum, &run); }
        g1_add(&total, &total, &sum);
    }
    *r = total;
    free(bk); free(dig); free(le); free(aff);
}

/* ---------------- SHA-256 (FIPS 180-4) ---------------- */
static const uint32_t SHA_K[64] = {
    0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01, 0x243185be,
    0x550c7dc3, 0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa,
    0x5cb0a9dc, 0x76f988da, 0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85,
    0x2e1b2138, 0x4d2c6dfc, 0x53380d13, 0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3,
    0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070, 0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f,
    0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208, 0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
typedef struct { uint32_t h[8]; uint8_t buf[64]; size_t fill; u64 total; } sha256_t;
static inline uint32_t rotr32(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
static void sha256_block(uint32_t h[8], const uint8_t *p) {
    uint32_t w[64];
    for (int i = 0; i < 16; i++) w[i] = ((uint32_t)p[4 * i] << 24) | ((uint32_t)p[4 * i + 1] << 16) | ((uint32_t)p[4 * i + 2] << 8) | p[4 * i + 3];
    for (int i = 16; i < 64; i++) {
        uint32_t s0 = rotr32(w[i - 15], 7) ^ rotr32(w[i - 15], 18) ^ (w[i - 15] >> 3);
        uint32_t s1 = rotr32(w[i - 2], 17) ^ rotr32(w[i - 2], 19) ^ (w[i - 2] >> 10);
        w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    for (int i = 0; i < 64; i++) {
        uint32_t t1 = hh + (rotr32(e, 6) ^ rotr32(e, 11) ^ rotr32(e, 25)) + ((e & f) ^ (~e & g)) + SHA_K[i] + w[i];
        uint32_t t2 = (rotr32(a, 2) ^ rotr32(a, 13) ^ rotr32(a, 22)) + ((a & b) ^ (a & c) ^ (b & c));
        hh = g; g = f; f = e; e = d + t1; d = c; c = b; b = a; a = t1 + t2;
    }
    h[0] += a; h[1] += b; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
}
static void sha256_init(sha256_t *s) {
    static const uint32_t iv[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    memcpy(s->h, iv, 32); s->fill = 0; s->total = 0;
}
static void sha256_update(sha256_t *s, const uint8_t *d, size_t n) {
    s->total += n;
    while (n) {
        if (s->fill == 0 && n >= 64) { sha256_block(s->h, d); d += 64; n -= 64; continue; }
        size_t k = 64 - s->fill; if (k > n) k = n;
        memcpy(s->buf + s->fill, d, k); s->fill += k; d += k; n -= k;
        if (s->fill == 64) { sha256_block(s->h, s->buf); s->fill = 0; }
    }
}
static void sha256_final(sha256_t *s, uint8_t out[32]) {
    u64 bits = s->total * 8; uint8_t pad[72] = {0x80}; size_t padlen = (s->fill < 56) ? 56 - s->fill : 120 - s->fill;
    uint8_t len[8]; for (int i = 0; i < 8; i++) len[i] = (uint8_t)(bits >> (56 - 8 * i));
    sha256_update(s, pad, padlen); sha256_update(s, len, 8);
    for (int i = 0; i < 8; i++) { out[4 * i] = s->h[i] >> 24; out[4 * i + 1] = s->h[i] >> 16; out[4 * i + 2] = s->h[i] >> 8; out[4 * i + 3] = s->h[i]; }
}
/* Fr::hash_to (src/bls.rs:189-205) */
static void fr_hash_to(fr_t *r, const uint8_t *data, size_t n) {
    sha256_t s; uint8_t h[32]; sha256_init(&s); sha256_update(&s, data, n); sha256_final(&s, h); fr_from_be32_reduce(r, h);
}
#endif
