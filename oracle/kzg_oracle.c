/*
 * ORACLE -- TEST INFRASTRUCTURE ONLY.  CPU restatement of kateth's EIP-4844 path, function for function,
 * quirks included.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build, load or call this; the shipped engine (kateth_b200/csrc) never does.
 *
 * Parity is PINNED: tools/reconstruct_vectors.py rebuilds all 253 consensus-spec-tests KZG data.yaml files
 * from this oracle's outputs and matches each against the Git-LFS sha256 oid kept in
 * tests/golden/lfs_oids.json.
 *
 * The arithmetic kateth delegates to the third-party crate blst ("0.3.11", caret requirement, exact
 * version unpinned, source absent from /root/reference) is restated in kzo_field.h / kzo_g1.h /
 * kzo_pairing.h from the published BLS12-381 definitions.
 *
 * Build: make -C oracle   (gcc -O3 -march=native -shared -fPIC)
 */
#include <stdio.h>
#include <pthread.h>
#include "kzo_pairing.h"

#define KZ_N 4096
#define KZ_BLOB_BYTES (KZ_N * 32)

typedef struct {
    g1_t g1_lagrange_brp[KZ_N];   /* src/kzg/setup.rs:39 */
    g2_t g2_monomial[65];         /* src/kzg/setup.rs:40 */
    fr_t roots_of_unity_brp[KZ_N];/* src/kzg/setup.rs:41 */
} kzo_setup;

/* src/math.rs:72-74 */
static unsigned brp_index(unsigned i, unsigned log2n) { unsigned r = 0; for (unsigned k = 0; k < log2n; k++) r |= ((i >> k) & 1) << (log2n - 1 - k); return r; }

/* Setup::load_json minus the JSON/hex layer (src/kzg/setup.rs:46-82): the caller passes the decoded bytes. */
int kzo_setup_load(const uint8_t *g1_lagrange48, size_t n_g1, const uint8_t *g2_monomial96, size_t n_g2, kzo_setup **out) {
    if (n_g1 != KZ_N) return -10;   /* InvalidLenG1Lagrange */
    if (n_g2 != 65) return -11;     /* InvalidLenG2Monomial */
    kzo_setup *s = (kzo_setup *)malloc(sizeof *s);
    g1_t *tmp = (g1_t *)malloc(sizeof(g1_t) * KZ_N);
    for (int i = 0; i < KZ_N; i++) { int st = g1_decompress(&tmp[i], g1_lagrange48 + 48 * i); if (st) { free(tmp); free(s); return st; } }
    for (int i = 0; i < KZ_N; i++) s->g1_lagrange_brp[i] = tmp[brp_index(i, 12)];
    free(tmp);
    for (int i = 0; i < 65; i++) { int st = g2_decompress(&s->g2_monomial[i], g2_monomial96 + 96 * i); if (st) { free(s); return st; } }
    /* src/math.rs:7-29: primitive = 7^((r-1)/4096) via Fr::pow; roots[i] = primitive^i */
    fr_t w, cur; { fr_t t; memcpy(t.l, FR_OMEGA, 32); fr_to_mont(&w, t.l); }
    fr_t *roots = (fr_t *)malloc(sizeof(fr_t) * KZ_N);
    fr_one(&cur);
    for (int i = 0; i < KZ_N; i++) { roots[i] = cur; fr_mul(&cur, &cur, &w); }
    for (int i = 0; i < KZ_N; i++) s->roots_of_unity_brp[i] = roots[brp_index(i, 12)];
    free(roots);
    *out = s; return 0;
}
void kzo_setup_free(kzo_setup *s) { free(s); }

/* Blob::from_slice (src/blob.rs:26-37) */
static int blob_from_slice(fr_t *el, const uint8_t *bytes, size_t len) {
    if (len != KZ_BLOB_BYTES) return KZ_INVALID_LEN;
    for (int i = 0; i < KZ_N; i++) if (!fr_from_be32(&el[i], bytes + 32 * i)) return KZ_INVALID_FIELD_ELEMENT;
    return KZ_OK;
}
/* Blob::commitment (src/blob.rs:48-53) */
static void blob_commitment(g1_t *c, const fr_t *el, const kzo_setup *s) { g1_lincomb_pippenger(c, s->g1_lagrange_brp, el, KZ_N); }
/* Blob::challenge (src/blob.rs:78-97) */
static void blob_challenge(fr_t *z, const fr_t *el, const g1_t *commitment) {
    size_t n = 16 + 16 + KZ_BLOB_BYTES + 48; uint8_t *data = (uint8_t *)malloc(n);
    memcpy(data, "FSBLOBVERIFY_V1_", 16);
    memset(data + 16, 0, 16); data[30] = (KZ_N >> 8) & 0xff; data[31] = KZ_N & 0xff;   /* (N as u128).to_be_bytes() */
    for (int i = 0; i < KZ_N; i++) fr_to_be32(data + 32 + 32 * i, &el[i]);
    g1_compress(data + 32 + KZ_BLOB_BYTES, commitment);
    fr_hash_to(z, data, n); free(data);
}
/* Polynomial::evaluate (src/kzg/poly.rs:10-33): per-term division, as written */
static void poly_evaluate(fr_t *y, const fr_t *f, const fr_t *point, const kzo_setup *s) {
    const fr_t *roots = s->roots_of_unity_brp;
    for (int i = 0; i < KZ_N; i++) if (fr_eq(point, &roots[i])) { *y = f[i]; return; }
    fr_t eval; fr_zero(&eval);
    for (int i = 0; i < KZ_N; i++) {
        fr_t numer, denom, term; fr_mul(&numer, &f[i], &roots[i]); fr_sub(&denom, point, &roots[i]);
        fr_div(&term, &numer, &denom); fr_add(&eval, &eval, &term);
    }
    fr_t n, one, t; fr_set_u64(&n, KZ_N); fr_one(&one);
    fr_pow_kateth(&t, point, &n); fr_sub(&t, &t, &one); fr_div(&t, &t, &n);
    fr_mul(y, &eval, &t);
}
/* Polynomial::prove (src/kzg/poly.rs:36-71) */
static void poly_prove(fr_t *eval, g1_t *proof, const fr_t *f, const fr_t *point, const kzo_setup *s) {
    const fr_t *roots = s->roots_of_unity_brp;
    poly_evaluate(eval, f, point, s);
    fr_t *q = (fr_t *)malloc(sizeof(fr_t) * KZ_N);
    for (int i = 0; i < KZ_N; i++) {
        fr_t numer, denom; fr_sub(&numer, &f[i], eval); fr_sub(&denom, &roots[i], point);
        if (!fr_is_zero(&denom)) { fr_div(&q[i], &numer, &denom); continue; }
        fr_t quo; fr_zero(&quo);
        for (int j = 0; j < KZ_N; j++) {
            if (j == i) continue;
            fr_t coef, num, d1, d2, term; fr_sub(&coef, &f[j], eval); fr_mul(&num, &coef, &roots[j]);
            fr_mul(&d1, &roots[i], &roots[i]); fr_mul(&d2, &roots[i], &roots[j]); fr_sub(&d1, &d1, &d2);
            fr_div(&term, &num, &d1); fr_add(&quo, &quo, &term);
        }
        q[i] = quo;
    }
    g1_lincomb_pippenger(proof, s->g1_lagrange_brp, q, KZ_N);
    free(q);
}
/* Setup::verify_proof_inner (src/kzg/setup.rs:84-94) */
static int verify_proof_inner(const kzo_setup *s, const g1_t *proof, const g1_t *commitment, const fr_t *point, const fr_t *eval) {
    g2_t g2, ng2, t, a2; g2_generator(&g2); g2_neg(&ng2, &g2);
    g2_mul_fr(&t, &ng2, point); g2_add(&a2, &s->g2_monomial[1], &t);
    g1_t g1, ng1, u, b1; g1_generator(&g1); g1_neg(&ng1, &g1);
    g1_mul_fr(&u, &ng1, eval); g1_add(&b1, commitment, &u);
    return verify_pairings(proof, &a2, &b1, &g2);
}
/* Setup::verify_proof_batch (src/kzg/setup.rs:115-161) up to the pairing: the three lincombs.
 * rpowers[i] = Fr::pow(r, i) with the pow(x,0)=x quirk, i.e. [r, r, r^2, r^3, ...]. */
static void verify_batch_sums(g1_t *proof_lincomb, g1_t *rhs, const g1_t *proofs, const g1_t *comms, const fr_t *points, const fr_t *evals, size_t n) {
    uint8_t data[48]; memcpy(data, "RCKZGBATCH___V1_", 16);
    memset(data + 16, 0, 32); data[30] = (KZ_N >> 8) & 0xff; data[31] = KZ_N & 0xff;
    for (int k = 0; k < 8; k++) data[47 - k] = (uint8_t)((u64)n >> (8 * k));
    fr_t r; fr_hash_to(&r, data, 48);
    fr_t *rp = (fr_t *)malloc(sizeof(fr_t) * (n + 1)), *zr = (fr_t *)malloc(sizeof(fr_t) * (n + 1));
    g1_t *cme = (g1_t *)malloc(sizeof(g1_t) * (n + 1));
    g1_t g1, ng1; g1_generator(&g1); g1_neg(&ng1, &g1);
    for (size_t i = 0; i < n; i++) {
        fr_t e; fr_set_u64(&e, (u64)i); fr_pow_kateth(&rp[i], &r, &e);
        fr_mul(&zr[i], &points[i], &rp[i]);
        g1_t t; g1_mul_fr(&t, &ng1, &evals[i]); g1_add(&cme[i], &comms[i], &t);
    }
    g1_t pz, cl;
    g1_lincomb_naive(proof_lincomb, proofs, rp, n);
    g1_lincomb_naive(&pz, proofs, zr, n);
    g1_lincomb_naive(&cl, cme, rp, n);
    g1_add(rhs, &cl, &pz);
    free(rp); free(zr); free(cme);
}
static int verify_proof_batch(const kzo_setup *s, const g1_t *proofs, const g1_t *comms, const fr_t *points, const fr_t *evals, size_t n) {
    g1_t a, b; verify_batch_sums(&a, &b, proofs, comms, points, evals, n);
    g2_t g2; g2_generator(&g2);
    return verify_pairings(&a, &s->g2_monomial[1], &b, &g2);
}

/* ======================= public (ctypes) surface ======================= */
int kzo_blob_to_commitment(const kzo_setup *s, const uint8_t *blob, size_t len, uint8_t out48[48]) {
    fr_t *el = (fr_t *)malloc(sizeof(fr_t) * KZ_N); int st = blob_from_slice(el, blob, len);
    if (!st) { g1_t c; blob_commitment(&c, el, s); g1_compress(out48, &c); }
    free(el); return st;
}
/* Setup::blob_proof (src/kzg/setup.rs:177-183) */
int kzo_blob_proof(const kzo_setup *s, const uint8_t *blob, size_t len, const uint8_t *comm48, size_t clen, uint8_t out48[48]) {
    fr_t *el = (fr_t *)malloc(sizeof(fr_t) * KZ_N); int st = blob_from_slice(el, blob, len);
    g1_t c;
    if (!st && clen != 48) st = KZ_G1_INVALID_ENCODING;
    if (!st) st = g1_decompress(&c, comm48);
    if (!st) { fr_t z, y; g1_t p; blob_challenge(&z, el, &c); poly_prove(&y, &p, el, &z, s); g1_compress(out48, &p); }
    free(el); return st;
}
/* Setup::proof (src/kzg/setup.rs:185-194) */
int kzo_proof(const kzo_setup *s, const uint8_t *blob, size_t len, const uint8_t *z32, size_t zlen, uint8_t out48[48], uint8_t y32[32]) {
    fr_t *el = (fr_t *)malloc(sizeof(fr_t) * KZ_N); int st = blob_from_slice(el, blob, len);
    fr_t z;
    if (!st && zlen != 32) st = KZ_SCALAR_INVALID_ENCODING;
    if (!st && !fr_from_be32(&z, z32)) st = KZ_SCALAR_NOT_IN_FIELD;
    if (!st) { fr_t y; g1_t p; poly_prove(&y, &p, el, &z, s); g1_compress(out48, &p); fr_to_be32(y32, &y); }
    free(el); return st;
}
/* Setup::verify_proof (src/kzg/setup.rs:96-113); argument order proof, commitment, point, eval */
int kzo_verify_proof(const kzo_setup *s, const uint8_t *proof48, size_t plen, const uint8_t *comm48, size_t clen,
                     const uint8_t *z32, size_t zlen, const uint8_t *y32, size_t ylen, int *ok) {
    g1_t p, c; fr_t z, y; int st;
    if (plen != 48) return KZ_G1_INVALID_ENCODING;
    if ((st = g1_decompress(&p, proof48))) return st;
    if (clen != 48) return KZ_G1_INVALID_ENCODING;
    if ((st = g1_decompress(&c, comm48))) return st;
    if (zlen != 32) return KZ_SCALAR_INVALID_ENCODING;
    if (!fr_from_be32(&z, z32)) return KZ_SCALAR_NOT_IN_FIELD;
    if (ylen != 32) return KZ_SCALAR_INVALID_ENCODING;
    if (!fr_from_be32(&y, y32)) return KZ_SCALAR_NOT_IN_FIELD;
    *ok = verify_proof_inner(s, &p, &c, &z, &y); return KZ_OK;
}
/* Setup::verify_blob_proof (src/kzg/setup.rs:196-221) */
int kzo_verify_blob_proof(const kzo_setup *s, const uint8_t *blob, size_t len, const uint8_t *comm48, size_t clen,
                          const uint8_t *proof48, size_t plen, int *ok) {
    fr_t *el = (fr_t *)malloc(sizeof(fr_t) * KZ_N); int st = blob_from_slice(el, blob, len);
    g1_t c, p;
    if (!st && clen != 48) st = KZ_G1_INVALID_ENCODING;
    if (!st) st = g1_decompress(&c, comm48);
    if (!st && plen != 48) st = KZ_G1_INVALID_ENCODING;
    if (!st) st = g1_decompress(&p, proof48);
    if (!st) { fr_t z, y; blob_challenge(&z, el, &c); poly_evaluate(&y, el, &z, s); *ok = verify_proof_inner(s, &p, &c, &z, &y); }
    free(el); return st;
}
/* Setup::verify_blob_proof_batch (src/kzg/setup.rs:223-275): blobs first, then commitments, then proofs;
 * first error wins.  blobs is n contiguous 131072-byte blobs. */
static int batch_prepare(const kzo_setup *s, const uint8_t *blobs, const uint8_t *comms48, const uint8_t *proofs48, size_t n,
                         g1_t **pc, g1_t **pp, fr_t **pz, fr_t **py) {
    fr_t *el = (fr_t *)malloc(sizeof(fr_t) * KZ_N * (n ? n : 1));
    g1_t *c = (g1_t *)malloc(sizeof(g1_t) * (n + 1)), *p = (g1_t *)malloc(sizeof(g1_t) * (n + 1));
    fr_t *z = (fr_t *)malloc(sizeof(fr_t) * (n + 1)), *y = (fr_t *)malloc(sizeof(fr_t) * (n + 1));
    int st = 0;
    for (size_t i = 0; i < n && !st; i++) st = blob_from_slice(el + i * KZ_N, blobs + i * KZ_BLOB_BYTES, KZ_BLOB_BYTES);
    for (size_t i = 0; i < n && !st; i++) st = g1_decompress(&c[i], comms48 + 48 * i);
    for (size_t i = 0; i < n && !st; i++) st = g1_decompress(&p[i], proofs48 + 48 * i);
    if (!st) for (size_t i = 0; i < n; i++) { blob_challenge(&z[i], el + i * KZ_N, &c[i]); poly_evaluate(&y[i], el + i * KZ_N, &z[i], s); }
    free(el);
    if (st) { free(c); free(p); free(z); free(y); return st; }
    *pc = c; *pp = p; *pz = z; *py = y; return 0;
}
int kzo_verify_blob_proof_batch(const kzo_setup *s, const uint8_t *blobs, const uint8_t *comms48, const uint8_t *proofs48, size_t n, int *ok) {
    g1_t *c, *p; fr_t *z, *y; int st = batch_prepare(s, blobs, comms48, proofs48, n, &c, &p, &z, &y); if (st) return st;
    *ok = verify_proof_batch(s, p, c, z, y, n);
    free(c); free(p); free(z); free(y); return KZ_OK;
}
/* The two G1 operands of the final pairing check (A = sum r^i pi_i ; B = sum r^i (C_i - y_i G) + sum r^i z_i pi_i),
 * compressed, so the GPU path's intermediate sums can be diffed against the reference's. */
int kzo_verify_blob_proof_batch_sums(const kzo_setup *s, const uint8_t *blobs, const uint8_t *comms48, const uint8_t *proofs48, size_t n,
                                     uint8_t a48[48], uint8_t b48[48]) {
    g1_t *c, *p; fr_t *z, *y; int st = batch_prepare(s, blobs, comms48, proofs48, n, &c, &p, &z, &y); if (st) return st;
    g1_t a, b; verify_batch_sums(&a, &b, p, c, z, y, n); g1_compress(a48, &a); g1_compress(b48, &b);
    free(c); free(p); free(z); free(y); return KZ_OK;
}

/* ======================= low-level hooks for unit parity tests (K1/K5/K6/K8 on the GPU side) ======================= */
/* Montgomery products on raw little-endian limb arrays (blst layout): n items each */
void kzo_fp_mul_raw(const u64 *a, const u64 *b, u64 *out, size_t n) { for (size_t i = 0; i < n; i++) mont_mul(out + 6 * i, a + 6 * i, b + 6 * i, FP_P, FP_INV, 6); }
void kzo_fr_mul_raw(const u64 *a, const u64 *b, u64 *out, size_t n) { for (size_t i = 0; i < n; i++) mont_mul(out + 4 * i, a + 4 * i, b + 4 * i, FR_R, FR_INV, 4); }
void kzo_fr_consts(u64 one[4], u64 max[4]) { fr_t o; fr_one(&o); memcpy(one, o.l, 32); fr_t z, m; fr_zero(&z); fr_sub(&m, &z, &o); memcpy(max, m.l, 32); }
void kzo_sha256(const uint8_t *d, size_t n, uint8_t out[32]) { sha256_t s; sha256_init(&s); sha256_update(&s, d, n); sha256_final(&s, out); }
int kzo_g1_decompress_check(const uint8_t in[48]) { g1_t p; return g1_decompress(&p, in); }
/* out = compress(sum_i k_i * P_i), scalars 32-byte BE canonical, points compressed */
int kzo_g1_lincomb(const uint8_t *pts48, const uint8_t *sc32, size_t n, uint8_t out48[48]) {
    g1_t acc; g1_set_inf(&acc);
    for (size_t i = 0; i < n; i++) {
        g1_t p, t; fr_t k; int st = g1_decompress(&p, pts48 + 48 * i); if (st) return st;
        if (!fr_from_be32(&k, sc32 + 32 * i)) return KZ_SCALAR_NOT_IN_FIELD;
        g1_mul_fr(&t, &p, &k); g1_add(&acc, &acc, &t);
    }
    g1_compress(out48, &acc); return 0;
}
int kzo_g1_add(const uint8_t a48[48], const uint8_t b48[48], uint8_t out48[48]) {
    g1_t a, b, c; int st; if ((st = g1_decompress(&a, a48))) return st; if ((st = g1_decompress(&b, b48))) return st;
    g1_add(&c, &a, &b); g1_compress(out48, &c); return 0;
}
/* challenge and evaluation for one blob (Blob::challenge + Polynomial::evaluate) */
int kzo_blob_challenge_eval(const kzo_setup *s, const uint8_t *blob, const uint8_t comm48[48], uint8_t z32[32], uint8_t y32[32]) {
    fr_t *el = (fr_t *)malloc(sizeof(fr_t) * KZ_N); int st = blob_from_slice(el, blob, KZ_BLOB_BYTES); g1_t c;
    if (!st) st = g1_decompress(&c, comm48);
    if (!st) { fr_t z, y; blob_challenge(&z, el, &c); poly_evaluate(&y, el, &z, s); fr_to_be32(z32, &z); fr_to_be32(y32, &y); }
    free(el); return st;
}
/* e(a1,a2) == e(b1,b2) on compressed operands; exercises the pairing alone (bilinearity tests) */
int kzo_pairing_eq(const uint8_t a1[48], const uint8_t a2[96], const uint8_t b1[48], const uint8_t b2[96], int *ok) {
    g1_t p1, p2; g2_t q1, q2; int st;
    if ((st = g1_decompress(&p1, a1)) || (st = g1_decompress(&p2, b1)) || (st = g2_decompress(&q1, a2)) || (st = g2_decompress(&q2, b2))) return st;
    *ok = verify_pairings(&p1, &q1, &p2, &q2); return 0;
}
int kzo_g2_mul_gen(const uint8_t k32[32], uint8_t x96_out[96]) { /* returns affine x of [k]G2 as c1||c0 BE, for tests */
    fr_t k; if (!fr_from_be32(&k, k32)) return KZ_SCALAR_NOT_IN_FIELD; g2_t g, t; g2_generator(&g); g2_mul_fr(&t, &g, &k);
    g2a_t a; g2_to_affine(&a, &t); if (a.inf) { memset(x96_out, 0, 96); x96_out[0] = 0xc0; return 0; }
    fp_to_be48(x96_out, &a.x.c1); fp_to_be48(x96_out + 48, &a.x.c0); x96_out[0] |= 0x80; if (fp2_is_lex_largest(&a.y)) x96_out[0] |= 0x20; return 0;
}

/* ======================= CPU baseline: the same functions over many blobs on T host threads ======================= */
typedef struct { const kzo_setup *s; const uint8_t *blobs, *comms, *proofs; uint8_t *out_c, *out_p; size_t lo, hi; int mode, status; } job_t;
static void *job_run(void *arg) {
    job_t *j = (job_t *)arg;
    for (size_t i = j->lo; i < j->hi; i++) {
        const uint8_t *b = j->blobs + i * KZ_BLOB_BYTES; int st = 0;
        if (j->mode == 0 || j->mode == 2) st = kzo_blob_to_commitment(j->s, b, KZ_BLOB_BYTES, j->out_c + 48 * i);
        if (!st && (j->mode == 1 || j->mode == 2)) st = kzo_blob_proof(j->s, b, KZ_BLOB_BYTES, (j->mode == 2 ? j->out_c : j->comms) + 48 * i, 48, j->out_p + 48 * i);
        if (st) j->status = st;
    }
    return NULL;
}
/* mode 0: commit; 1: blob_proof with given commitments; 2: commit then blob_proof.  Blobs are split across nthreads. */
int kzo_commit_prove_many(const kzo_setup *s, const uint8_t *blobs, const uint8_t *comms, size_t n, int mode, int nthreads,
                          uint8_t *out_c, uint8_t *out_p) {
    if (nthreads < 1) nthreads = 1;
    pthread_t th[256]; job_t jb[256]; if (nthreads > 256) nthreads = 256;
    for (int t = 0; t < nthreads; t++) {
        jb[t] = (job_t){s, blobs, comms, NULL, out_c, out_p, n * t / nthreads, n * (t + 1) / nthreads, mode, 0};
        pthread_create(&th[t], NULL, job_run, &jb[t]);
    }
    int st = 0; for (int t = 0; t < nthreads; t++) { pthread_join(th[t], NULL); if (jb[t].status) st = jb[t].status; }
    return st;
}
/* Batch verification over n blobs as nthreads independent sub-batches (each a full kateth verify_blob_proof_batch
 * call on its contiguous share, its own r and pairing) -- the way a multi-core host would shard kateth. */
typedef struct { const kzo_setup *s; const uint8_t *blobs, *comms, *proofs; size_t lo, hi; int ok, status; } vjob_t;
static void *vjob_run(void *arg) {
    vjob_t *j = (vjob_t *)arg; j->ok = 0;
    j->status = kzo_verify_blob_proof_batch(j->s, j->blobs + j->lo * KZ_BLOB_BYTES, j->comms + 48 * j->lo, j->proofs + 48 * j->lo, j->hi - j->lo, &j->ok);
    return NULL;
}
int kzo_verify_many(const kzo_setup *s, const uint8_t *blobs, const uint8_t *comms, const uint8_t *proofs, size_t n, int nthreads, int *ok) {
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256]; vjob_t jb[256];
    for (int t = 0; t < nthreads; t++) { jb[t] = (vjob_t){s, blobs, comms, proofs, n * t / nthreads, n * (t + 1) / nthreads, 0, 0}; pthread_create(&th[t], NULL, vjob_run, &jb[t]); }
    int st = 0, all = 1; for (int t = 0; t < nthreads; t++) { pthread_join(th[t], NULL); if (jb[t].status) st = jb[t].status; if (!jb[t].ok) all = 0; }
    *ok = all; return st;
}
