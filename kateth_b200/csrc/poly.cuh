// poly.cuh -- per-blob scalar-field work of the KZG path: blob validation, the Fiat-Shamir challenge,
// barycentric evaluation with a batched inversion, and the quotient polynomial.
//
// Replaces (file:line in /root/reference):
//   Blob::from_slice            src/blob.rs:26-37      (k_blob_check: every element < r)
//   Blob::challenge             src/blob.rs:78-97      (k_challenge: SHA-256 over domain | degree | blob | commitment)
//   Polynomial::evaluate        src/kzg/poly.rs:10-33  (k_eval_prod / k_eval_inv / k_eval_finish)
//   Polynomial::prove (quotient) src/kzg/poly.rs:45-66 (k_eval_finish with WANT_Q)
// kateth divides per term (8,192 Euclidean inversions per proof).  Here all 4,096 denominators of a blob are
// inverted with ONE inversion: per-thread products, a product tree over the CTA in shared memory, the root
// inverted in a separate kernel where 32 blobs share a warp (one inversion per lane), then the tree is walked
// back down.  Field elements are unique, so the results are bit-identical to the per-term divisions.
//
// Domain mixing: blob elements f_i stay in canonical (non-Montgomery) form; multiplying a canonical value by a
// Montgomery-form value with the Montgomery product yields a canonical value, so y and the quotient come out
// canonical -- exactly what the output encoding and the MSM's digit recoding need -- with no conversions.
#pragma once
#include "field.cuh"
#include "sha256.cuh"
#include "msm.cuh"

namespace kb200 {

// ---- roots of unity in bit-reversed order (src/math.rs:7-29, 72-74), Montgomery form, built once ----
__device__ __forceinline__ uint32_t brp12(uint32_t i) { return __brev(i) >> 20; }
__global__ void k_roots(Fr* roots_brp) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= KB_N) return;
    Fr w; fr_set(w, c_k.fr_omega);
    Fr r; fr_pow_u32(r, w, brp12(i));
    roots_brp[i] = r;
}

// ---- monomial cross-check of a setup: blob k0 + b holds the evaluations of X^(k0+b) over the domain in kateth's blob order, i.e. element
// i = roots_brp[i]^(k0+b) = w^(brp(i) * (k0+b) mod 4096), as canonical big-endian bytes; its commitment must be [tau^(k0+b)]G1 ----
__global__ void k_vandermonde(const Fr* __restrict__ roots_brp, int k0, int count, uint32_t* __restrict__ blobs) {
    size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (size_t)count * KB_N) return;
    const uint32_t b = (uint32_t)(gid >> 12), i = (uint32_t)(gid & (KB_N - 1));
    const uint32_t e = (brp12(i) * (uint32_t)(k0 + b)) & (KB_N - 1);
    Fr w = roots_brp[brp12(e)], plain;
    fr_from_mont(plain, w);
    fr_store_be(blobs + gid * 8, plain);
}

// ---- Blob::from_slice: status[b] = 2 (InvalidFieldElement) if any element >= r ----
__global__ void k_blob_check(const uint32_t* __restrict__ blobs, int n, int32_t* __restrict__ status) {
    size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (size_t)n * KB_N) return;
    const uint4* p = reinterpret_cast<const uint4*>(blobs + gid * 8);
    uint4 a = __ldg(p), b = __ldg(p + 1);
    Fr v;
    v.l[7] = bswap32(a.x); v.l[6] = bswap32(a.y); v.l[5] = bswap32(a.z); v.l[4] = bswap32(a.w);
    v.l[3] = bswap32(b.x); v.l[2] = bswap32(b.y); v.l[1] = bswap32(b.z); v.l[0] = bswap32(b.w);
    if (fr_plain_geq_r(v)) status[gid >> 12] = KB_INVALID_FIELD_ELEMENT;
}
// 32-byte big-endian scalars (z or y): status 7 (NotInFiniteField) if >= r; out = canonical limbs
__global__ void k_scalar_check(const uint32_t* __restrict__ in, int n, Fr* __restrict__ out, int32_t* __restrict__ status) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr v; fr_load_be(v, in + (size_t)i * 8);
    if (fr_plain_geq_r(v)) { if (status[i] == 0) status[i] = KB_SCALAR_NOT_IN_FIELD; }
    out[i] = v;
}

// ---- Blob::challenge: z = SHA-256("FSBLOBVERIFY_V1_" | u128_be(4096) | blob | compress(C)) mod r, canonical ----
__global__ void k_challenge(const uint32_t* __restrict__ blobs, const uint32_t* __restrict__ comm48, int n, const int32_t* __restrict__ status,
                            Fr* __restrict__ z_out) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n) return;
    if (status[b] != 0) { Fr zz; fr_zero(zz); z_out[b] = zz; return; }
    const uint4* bp = reinterpret_cast<const uint4*>(blobs + (size_t)b * (KB_N * 8));
    uint32_t h[8], w[16];
    sha256_init(h);
    // block 0: domain tag, degree, first 32 bytes of the blob
    w[0] = 0x4653424c; w[1] = 0x4f425645; w[2] = 0x52494659; w[3] = 0x5f56315f;   // "FSBLOBVERIFY_V1_"
    w[4] = 0; w[5] = 0; w[6] = 0; w[7] = KB_N;
    uint4 c0 = __ldg(bp), c1 = __ldg(bp + 1);
    w[8] = bswap32(c0.x); w[9] = bswap32(c0.y); w[10] = bswap32(c0.z); w[11] = bswap32(c0.w);
    w[12] = bswap32(c1.x); w[13] = bswap32(c1.y); w[14] = bswap32(c1.z); w[15] = bswap32(c1.w);
    sha256_block(h, w);
    // blocks 1..2047: blob bytes [64k-32, 64k+32)
#pragma unroll 1
    for (int k = 1; k < 2048; k++) {
        const uint4* p = bp + (4 * k - 2);
        uint4 v0 = __ldg(p), v1 = __ldg(p + 1), v2 = __ldg(p + 2), v3 = __ldg(p + 3);
        w[0] = bswap32(v0.x); w[1] = bswap32(v0.y); w[2] = bswap32(v0.z); w[3] = bswap32(v0.w);
        w[4] = bswap32(v1.x); w[5] = bswap32(v1.y); w[6] = bswap32(v1.z); w[7] = bswap32(v1.w);
        w[8] = bswap32(v2.x); w[9] = bswap32(v2.y); w[10] = bswap32(v2.z); w[11] = bswap32(v2.w);
        w[12] = bswap32(v3.x); w[13] = bswap32(v3.y); w[14] = bswap32(v3.z); w[15] = bswap32(v3.w);
        sha256_block(h, w);
    }
    // block 2048: last 32 blob bytes + first 32 bytes of the commitment
    {
        const uint4* p = bp + (4 * 2048 - 2);
        uint4 v0 = __ldg(p), v1 = __ldg(p + 1);
        const uint32_t* cp = comm48 + (size_t)b * 12;
        w[0] = bswap32(v0.x); w[1] = bswap32(v0.y); w[2] = bswap32(v0.z); w[3] = bswap32(v0.w);
        w[4] = bswap32(v1.x); w[5] = bswap32(v1.y); w[6] = bswap32(v1.z); w[7] = bswap32(v1.w);
#pragma unroll
        for (int k = 0; k < 8; k++) w[8 + k] = bswap32(cp[k]);
        sha256_block(h, w);
        // block 2049: last 16 commitment bytes, padding, bit length (131152 * 8 = 1049216 = 0x100280)
#pragma unroll
        for (int k = 0; k < 4; k++) w[k] = bswap32(cp[8 + k]);
        w[4] = 0x80000000u;
#pragma unroll
        for (int k = 5; k < 15; k++) w[k] = 0;
        w[15] = 131152u * 8u;
        sha256_block(h, w);
    }
    Fr z;
#pragma unroll
    for (int k = 0; k < 8; k++) z.l[7 - k] = h[k];
    fr_reduce_256(z);
    z_out[b] = z;
}

// The same hash over a range [blk0, blk1) of the 2,050 blocks of the preimage, with the chaining value parked in `state`
// (8 words per blob) between launches.  The host-buffer entry points upload a batch of blobs in column pieces (bytes
// [j*W, (j+1)*W) of every blob) and hash each piece while the next one crosses PCIe: block k needs blob bytes up to 64k + 32.
// blk1 == 2050 finishes (both commitment blocks, reduction mod r); otherwise blk1 <= 2048.
__global__ void k_challenge_range(const uint32_t* __restrict__ blobs, const uint32_t* __restrict__ comm48, int n, uint32_t* __restrict__ state,
                                  Fr* __restrict__ z_out, int blk0, int blk1) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n) return;
    const uint4* bp = reinterpret_cast<const uint4*>(blobs + (size_t)b * (KB_N * 8));
    uint32_t h[8], w[16];
    int k = blk0;
    if (k == 0) {
        sha256_init(h);
        w[0] = 0x4653424c; w[1] = 0x4f425645; w[2] = 0x52494659; w[3] = 0x5f56315f;   // "FSBLOBVERIFY_V1_"
        w[4] = 0; w[5] = 0; w[6] = 0; w[7] = KB_N;
        uint4 c0 = __ldg(bp), c1 = __ldg(bp + 1);
        w[8] = bswap32(c0.x); w[9] = bswap32(c0.y); w[10] = bswap32(c0.z); w[11] = bswap32(c0.w);
        w[12] = bswap32(c1.x); w[13] = bswap32(c1.y); w[14] = bswap32(c1.z); w[15] = bswap32(c1.w);
        sha256_block(h, w);
        k = 1;
    } else {
#pragma unroll
        for (int q = 0; q < 8; q++) h[q] = state[(size_t)q * n + b];
    }
    const int kend = blk1 < 2048 ? blk1 : 2048;
#pragma unroll 1
    for (; k < kend; k++) {
        const uint4* p = bp + (4 * k - 2);
        uint4 v0 = __ldg(p), v1 = __ldg(p + 1), v2 = __ldg(p + 2), v3 = __ldg(p + 3);
        w[0] = bswap32(v0.x); w[1] = bswap32(v0.y); w[2] = bswap32(v0.z); w[3] = bswap32(v0.w);
        w[4] = bswap32(v1.x); w[5] = bswap32(v1.y); w[6] = bswap32(v1.z); w[7] = bswap32(v1.w);
        w[8] = bswap32(v2.x); w[9] = bswap32(v2.y); w[10] = bswap32(v2.z); w[11] = bswap32(v2.w);
        w[12] = bswap32(v3.x); w[13] = bswap32(v3.y); w[14] = bswap32(v3.z); w[15] = bswap32(v3.w);
        sha256_block(h, w);
    }
    if (blk1 <= 2048) {
#pragma unroll
        for (int q = 0; q < 8; q++) state[(size_t)q * n + b] = h[q];
        return;
    }
    {
        const uint4* p = bp + (4 * 2048 - 2);
        uint4 v0 = __ldg(p), v1 = __ldg(p + 1);
        const uint32_t* cp = comm48 + (size_t)b * 12;
        w[0] = bswap32(v0.x); w[1] = bswap32(v0.y); w[2] = bswap32(v0.z); w[3] = bswap32(v0.w);
        w[4] = bswap32(v1.x); w[5] = bswap32(v1.y); w[6] = bswap32(v1.z); w[7] = bswap32(v1.w);
#pragma unroll
        for (int q = 0; q < 8; q++) w[8 + q] = bswap32(cp[q]);
        sha256_block(h, w);
#pragma unroll
        for (int q = 0; q < 4; q++) w[q] = bswap32(cp[8 + q]);
        w[4] = 0x80000000u;
#pragma unroll
        for (int q = 5; q < 15; q++) w[q] = 0;
        w[15] = 131152u * 8u;
        sha256_block(h, w);
    }
    Fr z;
#pragma unroll
    for (int q = 0; q < 8; q++) z.l[7 - q] = h[q];
    fr_reduce_256(z);
    z_out[b] = z;
}

// ---- the same hash with the work of one blob spread over TWO warps (one CTA = 32 blobs, 64 threads) ----
// A SHA-256 stream is serial, and with one thread per blob the kernel's duration IS the latency of one thread's chain: ~3.9 k cycles per
// 64-byte block, of which only the 64 rounds (four dependent operations each) are inherently serial -- the message schedule (48 words of
// sigma0 / sigma1 expansion), the loads and the byte swaps are not.  So lane l of warp 1 prepares K[i] + w[i] for blob l's next block in
// shared memory (double-buffered, named barriers) while lane l of warp 0 runs the rounds of the current one.
// Covers both forms: all 2,050 blocks (state == nullptr) or a range [blk0, blk1) with the chaining value parked in `state`.
#define CH2_FULL 1      // named barriers: FULL + j = buffer j holds a block, EMPTY + j = buffer j has been read
#define CH2_EMPTY 3
__device__ __forceinline__ void ch2_bar_sync(int id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void ch2_bar_arrive(int id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }
// the sixteen big-endian message words of block k of  "FSBLOBVERIFY_V1_" | u128_be(4096) | blob | compress(C), in two steps so that
// the producer can have the loads of block k + 1 in flight while it expands block k: the raw 16-byte pieces, then the words
__device__ __forceinline__ void ch2_block_raw(uint4 (&v)[4], int k, const uint4* __restrict__ bp, const uint32_t* __restrict__ cp) {
    if (k == 0) { v[0] = __ldg(bp); v[1] = __ldg(bp + 1); }
    else if (k < 2048) { const uint4* p = bp + (4 * k - 2); v[0] = __ldg(p); v[1] = __ldg(p + 1); v[2] = __ldg(p + 2); v[3] = __ldg(p + 3); }
    else if (k == 2048) {
        const uint4* p = bp + (4 * 2048 - 2);
        v[0] = __ldg(p); v[1] = __ldg(p + 1);
        v[2] = make_uint4(cp[0], cp[1], cp[2], cp[3]); v[3] = make_uint4(cp[4], cp[5], cp[6], cp[7]);
    } else v[0] = make_uint4(cp[8], cp[9], cp[10], cp[11]);
}
__device__ __forceinline__ void ch2_words4(uint32_t* w, const uint4& v) { w[0] = bswap32(v.x); w[1] = bswap32(v.y); w[2] = bswap32(v.z); w[3] = bswap32(v.w); }
__device__ __forceinline__ void ch2_block_words(uint32_t (&w)[16], int k, const uint4 (&v)[4]) {
    if (k == 0) {
        w[0] = 0x4653424c; w[1] = 0x4f425645; w[2] = 0x52494659; w[3] = 0x5f56315f; w[4] = 0; w[5] = 0; w[6] = 0; w[7] = KB_N;
        ch2_words4(w + 8, v[0]); ch2_words4(w + 12, v[1]);
    } else if (k <= 2048) {
        ch2_words4(w, v[0]); ch2_words4(w + 4, v[1]); ch2_words4(w + 8, v[2]); ch2_words4(w + 12, v[3]);
    } else {
        ch2_words4(w, v[0]);
        w[4] = 0x80000000u;
#pragma unroll
        for (int q = 5; q < 15; q++) w[q] = 0;
        w[15] = 131152u * 8u;
    }
}
#define CH2_THREADS 128  // two producer/consumer pairs per CTA: 4096 blobs occupy 64 SMs, so that the point decompression of a batch check
                         // (64 CTAs of its own) finds SMs to itself -- sharing an SM stretches the SHA chains from 2.4 to 5 ms
__global__ void __launch_bounds__(CH2_THREADS) k_challenge_w2(const uint32_t* __restrict__ blobs, const uint32_t* __restrict__ comm48, int n, const int32_t* __restrict__ status,
                                                     uint32_t* __restrict__ state, Fr* __restrict__ z_out, int blk0, int blk1) {
    __shared__ uint32_t kw_all[CH2_THREADS / 64][2][64][32];
    const int lane = threadIdx.x & 31, role = (threadIdx.x >> 5) & 1, pair = threadIdx.x >> 6;
    uint32_t (*kw)[64][32] = kw_all[pair];
    const int bar0 = 4 * pair;                                               // named barriers of this pair
    const int b = (blockIdx.x * (CH2_THREADS / 64) + pair) * 32 + lane, bc = b < n ? b : n - 1;          // lanes past the end hash the last blob again and write nothing
    const uint4* bp = reinterpret_cast<const uint4*>(blobs + (size_t)bc * (KB_N * 8));
    const uint32_t* cp = comm48 + (size_t)bc * 12;
    if (role == 1) {
        // ---- producer: K[i] + w[i], i = 0..63, for every block of the range ----
        uint4 nxt[4] = {};
        ch2_block_raw(nxt, blk0, bp, cp);
#pragma unroll 1
        for (int k = blk0; k < blk1; k++) {
            const int j = (k - blk0) & 1;
            uint4 cur[4] = {nxt[0], nxt[1], nxt[2], nxt[3]};
            if (k + 1 < blk1) ch2_block_raw(nxt, k + 1, bp, cp);       // in flight during the 48 schedule steps below
            uint32_t w[16];
            ch2_block_words(w, k, cur);
            if (k - blk0 >= 2) ch2_bar_sync(bar0 + CH2_EMPTY + j);
#pragma unroll
            for (int i = 0; i < 64; i++) {
                if (i >= 16) {
                    const uint32_t w15 = w[(i + 1) & 15], w2 = w[(i + 14) & 15];
                    const uint32_t s0 = rotr(w15, 7) ^ rotr(w15, 18) ^ (w15 >> 3);
                    const uint32_t s1 = rotr(w2, 17) ^ rotr(w2, 19) ^ (w2 >> 10);
                    w[i & 15] = w[i & 15] + s0 + w[(i + 9) & 15] + s1;
                }
                kw[j][i][lane] = w[i & 15] + c_sha_k[i];
            }
            ch2_bar_arrive(bar0 + CH2_FULL + j);
        }
        return;
    }
    // ---- consumer: the 64 rounds of every block ----
    uint32_t h[8];
    if (blk0 == 0) sha256_init(h);
    else {
#pragma unroll
        for (int q = 0; q < 8; q++) h[q] = state[(size_t)q * n + bc];
    }
#pragma unroll 1
    for (int k = blk0; k < blk1; k++) {
        const int j = (k - blk0) & 1;
        ch2_bar_sync(bar0 + CH2_FULL + j);
        uint32_t x[64];
#pragma unroll
        for (int i = 0; i < 64; i++) x[i] = kw[j][i][lane];
        if (k + 2 < blk1) ch2_bar_arrive(bar0 + CH2_EMPTY + j);
        uint32_t a = h[0], bb = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
        // The chain of e is the critical path of a round (a is needed four rounds later): h + K + w and d + h + K + w are formed one
        // round ahead from values that are already final, so e_new = (d + h + kw) + Sigma1(e) + Ch(e, f, g) is three dependent
        // instructions (shift, xor, three-input add) instead of four.
        uint32_t hx = hh + x[0], pre = d + hx;
#pragma unroll
        for (int i = 0; i < 64; i++) {
            const uint32_t s1 = rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25), ch = (e & f) ^ (~e & g);
            const uint32_t en = pre + s1 + ch, t1 = hx + s1 + ch;
            const uint32_t an = t1 + (rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22)) + ((a & bb) ^ (a & c) ^ (bb & c));
            hh = g; g = f; f = e; e = en; d = c; c = bb; bb = a; a = an;
            if (i < 63) { hx = hh + x[i + 1]; pre = d + hx; }
        }
        h[0] += a; h[1] += bb; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
    }
    if (b >= n) return;
    if (blk1 < 2050) {
#pragma unroll
        for (int q = 0; q < 8; q++) state[(size_t)q * n + b] = h[q];
        return;
    }
    Fr z;
#pragma unroll
    for (int q = 0; q < 8; q++) z.l[7 - q] = h[q];
    fr_reduce_256(z);
    if (status != nullptr && status[b] != 0) fr_zero(z);
    z_out[b] = z;
}

// ---------------------------------------------------------------------------------------------------------
// Evaluation, phase 1.  CTA per blob, EV_THREADS threads, EV_PER = 4096/EV_THREADS denominators per thread
// (element index i = k*EV_THREADS + tid).  d_i = z - w_i; if some d_m == 0 (z in the domain, poly.rs:14-18) the
// index m is recorded and that factor is replaced by 1.  Writes the product of all 4096 factors.
// ---------------------------------------------------------------------------------------------------------
#define EV_THREADS 512
#define EV_PER 8
__device__ __forceinline__ void cta_tree_up(Fr* tree, const Fr& leaf, int tid) {
    // tree[EV_THREADS + tid] = leaf; internal node n has children 2n, 2n+1; root at 1
    tree[EV_THREADS + tid] = leaf;
    __syncthreads();
#pragma unroll 1
    for (int width = EV_THREADS / 2; width >= 1; width >>= 1) {
        if (tid < width) { Fr a = tree[2 * (width + tid)], b = tree[2 * (width + tid) + 1], r; fr_mul(r, a, b); tree[width + tid] = r; }
        __syncthreads();
    }
}
// half_out != nullptr: also the products of the factors of elements [0, 2048) and [2048, 4096) (k_eval_finish2 works in halves)
__global__ void __launch_bounds__(EV_THREADS) k_eval_prod(const Fr* __restrict__ z_canon, const Fr* __restrict__ roots, int n,
                                                           Fr* __restrict__ prod_out, int32_t* __restrict__ indomain_out, Fr* __restrict__ half_out) {
    __shared__ Fr tree[2 * EV_THREADS];
    __shared__ int s_m;
    const int b = blockIdx.x, tid = threadIdx.x;
    if (tid == 0) s_m = -1;
    __syncthreads();
    Fr z; { Fr zc = z_canon[b]; fr_to_mont(z, zc); }
    Fr local; fr_set(local, c_k.fr_one);
    Fr lo = local;
#pragma unroll 1
    for (int k = 0; k < EV_PER; k++) {
        int i = k * EV_THREADS + tid;
        if (k == EV_PER / 2) { lo = local; fr_set(local, c_k.fr_one); }       // elements below 2048 are done
        Fr d; fr_sub(d, z, roots[i]);
        if (fr_is_zero(d)) s_m = i; else fr_mul(local, local, d);
    }
    if (half_out == nullptr) {
        fr_mul(local, local, lo);
        cta_tree_up(tree, local, tid);
        if (tid == 0) { prod_out[b] = tree[1]; indomain_out[b] = s_m; }
        return;
    }
    cta_tree_up(tree, lo, tid);
    Fr p0 = tree[1];
    __syncthreads();
    cta_tree_up(tree, local, tid);
    if (tid == 0) { Fr p1 = tree[1], t; fr_mul(t, p0, p1); prod_out[b] = t; half_out[2 * b] = p0; half_out[2 * b + 1] = p1; indomain_out[b] = s_m; }
}
// phase 2: one thread per blob inverts the product (32 blobs per warp, one Fermat inversion per lane)
__global__ void k_eval_inv(Fr* __restrict__ prod, int n) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n) return;
    Fr p = prod[b], r; fr_inv(r, p); prod[b] = r;
}
// phase 3: walk the tree down, recover 1/d_i, evaluate; optionally emit the quotient.
//   y (canonical)  = (z^N - 1)/N * sum_i f_i * w_i / (z - w_i)                      [poly.rs:23-32]
//   q_i (canonical) = (f_i - y)/(w_i - z) = (y - f_i) * inv_i                        [poly.rs:45-49]
//   z == w_m:  y = f_m;  q_m = sum_{j!=m} (f_j - y) w_j / (w_m (w_m - w_j)) = -w_m^-1 * sum_{j!=m} q_j w_j   [poly.rs:50-63]
template <bool WANT_Q>
__global__ void __launch_bounds__(EV_THREADS, 1) k_eval_finish(const uint32_t* __restrict__ blobs, const Fr* __restrict__ z_canon,
                                                             const Fr* __restrict__ roots, const Fr* __restrict__ prod_inv,
                                                             const int32_t* __restrict__ indomain, const int32_t* __restrict__ status, int n,
                                                             Fr* __restrict__ y_out, uint32_t* __restrict__ q_out) {
    __shared__ Fr tree[2 * EV_THREADS];
    __shared__ Fr s_y;
    const int b = blockIdx.x, tid = threadIdx.x;
    if (status[b] != 0) return;
    const int m = indomain[b];
    Fr z; { Fr zc = z_canon[b]; fr_to_mont(z, zc); }
    Fr one; fr_set(one, c_k.fr_one);
    // recompute the factors and the per-thread prefix products
    Fr pre[EV_PER];            // pre[k] = product of this thread's factors 0..k-1
    Fr local = one;
#pragma unroll
    for (int k = 0; k < EV_PER; k++) {
        int i = k * EV_THREADS + tid;
        Fr d; fr_sub(d, z, roots[i]);
        pre[k] = local;
        if (i != m) fr_mul(local, local, d);
    }
    cta_tree_up(tree, local, tid);
    // down-sweep: tree[n] becomes the inverse of the product under node n
    if (tid == 0) tree[1] = prod_inv[b];
    __syncthreads();
#pragma unroll 1
    for (int width = 1; width < EV_THREADS; width <<= 1) {
        if (tid < width) {
            int nd = width + tid;
            Fr inv = tree[nd], l = tree[2 * nd], r = tree[2 * nd + 1], il, ir;
            fr_mul(il, inv, r); fr_mul(ir, inv, l);
            tree[2 * nd] = il; tree[2 * nd + 1] = ir;
        }
        __syncthreads();
    }
    Fr run = tree[EV_THREADS + tid];          // inverse of this thread's product
    __syncthreads();
    // back-substitute: inv[k] = run * pre[k]; run *= d_k   (k descending)
    Fr (&inv)[EV_PER] = pre;      // inv[k] overwrites pre[k] in place (register budget)
    Fr acc; fr_zero(acc);
#pragma unroll
    for (int k = EV_PER - 1; k >= 0; k--) {
        int i = k * EV_THREADS + tid;
        Fr w = roots[i];
        if (i == m) { fr_zero(inv[k]); continue; }
        Fr d; fr_sub(d, z, w);
        fr_mul(inv[k], run, inv[k]);
        fr_mul(run, run, d);
        // term = f_i * (w_i * inv_i): canonical f times Montgomery value -> canonical
        Fr f; fr_load_be(f, blobs + ((size_t)b * KB_N + i) * 8);
        Fr t; fr_mul(t, w, inv[k]); fr_mul(t, f, t);
        fr_add(acc, acc, t);
    }
    // CTA sum of acc
    tree[tid] = acc;
    __syncthreads();
#pragma unroll 1
    for (int stride = EV_THREADS / 2; stride >= 1; stride >>= 1) {
        if (tid < stride) { Fr a = tree[tid], c = tree[tid + stride]; fr_add(a, a, c); tree[tid] = a; }
        __syncthreads();
    }
    if (tid == 0) {
        Fr y;
        if (m >= 0) {
            fr_load_be(y, blobs + ((size_t)b * KB_N + m) * 8);
        } else {
            Fr zn = z;
#pragma unroll 1
            for (int k = 0; k < 12; k++) fr_sqr(zn, zn);       // z^4096
            fr_sub(zn, zn, one);
            Fr ninv; fr_set(ninv, c_k.fr_ninv);
            fr_mul(zn, zn, ninv);                               // (z^N - 1)/N, Montgomery
            Fr s = tree[0];
            fr_mul(y, s, zn);                                   // canonical
        }
        s_y = y; y_out[b] = y;
    }
    if (!WANT_Q) return;
    __syncthreads();
    Fr y = s_y;
    Fr qsum; fr_zero(qsum);
#pragma unroll
    for (int k = 0; k < EV_PER; k++) {
        int i = k * EV_THREADS + tid;
        if (i == m) continue;
        Fr f; fr_load_be(f, blobs + ((size_t)b * KB_N + i) * 8);
        Fr q; fr_sub(q, y, f); fr_mul(q, q, inv[k]);            // canonical
        uint4* dst = reinterpret_cast<uint4*>(q_out + ((size_t)b * KB_N + i) * 8);
        dst[0] = make_uint4(q.l[0], q.l[1], q.l[2], q.l[3]); dst[1] = make_uint4(q.l[4], q.l[5], q.l[6], q.l[7]);
        if (m >= 0) { Fr t; fr_mul(t, q, roots[i]); fr_add(qsum, qsum, t); }
    }
    if (m >= 0) {          // uniform across the CTA
        __syncthreads();
        tree[tid] = qsum;
        __syncthreads();
#pragma unroll 1
        for (int stride = EV_THREADS / 2; stride >= 1; stride >>= 1) {
            if (tid < stride) { Fr a = tree[tid], c = tree[tid + stride]; fr_add(a, a, c); tree[tid] = a; }
            __syncthreads();
        }
        if (tid == 0) {
            // w_m^-1 = w^(4096 - e) where w_m = w^e, e = brp(m)
            uint32_t e = brp12((uint32_t)m);
            Fr winv = roots[brp12((KB_N - e) & (KB_N - 1))];
            Fr q = tree[0], nq; fr_mul(q, q, winv); fr_neg(nq, q);
            uint4* dst = reinterpret_cast<uint4*>(q_out + ((size_t)b * KB_N + m) * 8);
            dst[0] = make_uint4(nq.l[0], nq.l[1], nq.l[2], nq.l[3]); dst[1] = make_uint4(nq.l[4], nq.l[5], nq.l[6], nq.l[7]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Evaluation only (batch verification), in two halves of 2048 elements on 256 threads: the same arithmetic as k_eval_finish<false>,
// but two CTAs fit on an SM (128 registers x 256 threads), so the tree sweeps of one -- nine barriers with ever fewer threads at work --
// are covered by the bulk loops of the other.  The inverse of a half's product is prod_inv * (the other half's product).
// ---------------------------------------------------------------------------------------------------------
#define EV2_THREADS 256
// WANT_Q: the inverses 1/(z - w_i) are parked in the quotient's own slots (Montgomery form; a thread reads back only what it wrote, the
// resident CTAs' 128 KiB each stay in L2) until y is known, then q_i = (y - f_i) * inv_i overwrites them (poly.rs:45-63 as in k_eval_finish).
template <bool WANT_Q>
__global__ void __launch_bounds__(EV2_THREADS, 2) k_eval_finish2(const uint32_t* __restrict__ blobs, const Fr* __restrict__ z_canon,
                                                                const Fr* __restrict__ roots, const Fr* __restrict__ prod_inv,
                                                                const Fr* __restrict__ half_prod, const int32_t* __restrict__ indomain,
                                                                const int32_t* __restrict__ status, int n, Fr* __restrict__ y_out,
                                                                uint32_t* __restrict__ q_out) {
    __shared__ Fr tree[2 * EV2_THREADS];
    __shared__ Fr s_y;
    const int b = blockIdx.x, tid = threadIdx.x;
    if (status[b] != 0) return;
    const int m = indomain[b];
    Fr z; { Fr zc = z_canon[b]; fr_to_mont(z, zc); }
    Fr one; fr_set(one, c_k.fr_one);
    Fr acc; fr_zero(acc);
#pragma unroll 1
    for (int h = 0; h < 2; h++) {
        Fr pre[EV_PER];
        Fr local = one;
#pragma unroll
        for (int k = 0; k < EV_PER; k++) {
            const int i = (h * EV_PER + k) * EV2_THREADS + tid;
            Fr d; fr_sub(d, z, roots[i]);
            pre[k] = local;
            if (i != m) fr_mul(local, local, d);
        }
        // up-sweep over 256 leaves
        tree[EV2_THREADS + tid] = local;
        __syncthreads();
#pragma unroll 1
        for (int width = EV2_THREADS / 2; width >= 1; width >>= 1) {
            if (tid < width) { Fr a = tree[2 * (width + tid)], c = tree[2 * (width + tid) + 1], r; fr_mul(r, a, c); tree[width + tid] = r; }
            __syncthreads();
        }
        if (tid == 0) { Fr pi = prod_inv[b], po = half_prod[2 * b + (1 - h)], ih; fr_mul(ih, pi, po); tree[1] = ih; }
        __syncthreads();
#pragma unroll 1
        for (int width = 1; width < EV2_THREADS; width <<= 1) {
            if (tid < width) {
                const int nd = width + tid;
                Fr inv = tree[nd], l = tree[2 * nd], r = tree[2 * nd + 1], il, ir;
                fr_mul(il, inv, r); fr_mul(ir, inv, l);
                tree[2 * nd] = il; tree[2 * nd + 1] = ir;
            }
            __syncthreads();
        }
        Fr run = tree[EV2_THREADS + tid];
        __syncthreads();
#pragma unroll
        for (int k = EV_PER - 1; k >= 0; k--) {
            const int i = (h * EV_PER + k) * EV2_THREADS + tid;
            if (i == m) continue;
            Fr w = roots[i];
            Fr d; fr_sub(d, z, w);
            Fr inv; fr_mul(inv, run, pre[k]);
            fr_mul(run, run, d);
            if (WANT_Q) {
                uint4* dst = reinterpret_cast<uint4*>(q_out + ((size_t)b * KB_N + i) * 8);
                dst[0] = make_uint4(inv.l[0], inv.l[1], inv.l[2], inv.l[3]); dst[1] = make_uint4(inv.l[4], inv.l[5], inv.l[6], inv.l[7]);
            }
            Fr f; fr_load_be(f, blobs + ((size_t)b * KB_N + i) * 8);
            Fr t; fr_mul(t, w, inv); fr_mul(t, f, t);
            fr_add(acc, acc, t);
        }
    }
    tree[tid] = acc;
    __syncthreads();
#pragma unroll 1
    for (int stride = EV2_THREADS / 2; stride >= 1; stride >>= 1) {
        if (tid < stride) { Fr a = tree[tid], c = tree[tid + stride]; fr_add(a, a, c); tree[tid] = a; }
        __syncthreads();
    }
    if (tid == 0) {
        Fr y;
        if (m >= 0) fr_load_be(y, blobs + ((size_t)b * KB_N + m) * 8);
        else {
            Fr zn = z;
#pragma unroll 1
            for (int k = 0; k < 12; k++) fr_sqr(zn, zn);
            fr_sub(zn, zn, one);
            Fr ninv; fr_set(ninv, c_k.fr_ninv);
            fr_mul(zn, zn, ninv);
            Fr s = tree[0];
            fr_mul(y, s, zn);
        }
        y_out[b] = y;
        if (WANT_Q) s_y = y;
    }
    if (!WANT_Q) return;
    __syncthreads();
    const Fr y = s_y;
    Fr qsum; fr_zero(qsum);
#pragma unroll 1
    for (int k = 0; k < 2 * EV_PER; k++) {
        const int i = k * EV2_THREADS + tid;
        if (i == m) continue;
        uint4* slot = reinterpret_cast<uint4*>(q_out + ((size_t)b * KB_N + i) * 8);
        const uint4 v0 = slot[0], v1 = slot[1];
        Fr inv; inv.l[0] = v0.x; inv.l[1] = v0.y; inv.l[2] = v0.z; inv.l[3] = v0.w; inv.l[4] = v1.x; inv.l[5] = v1.y; inv.l[6] = v1.z; inv.l[7] = v1.w;
        Fr f; fr_load_be(f, blobs + ((size_t)b * KB_N + i) * 8);
        Fr q; fr_sub(q, y, f); fr_mul(q, q, inv);            // canonical
        slot[0] = make_uint4(q.l[0], q.l[1], q.l[2], q.l[3]); slot[1] = make_uint4(q.l[4], q.l[5], q.l[6], q.l[7]);
        if (m >= 0) { Fr t; fr_mul(t, q, roots[i]); fr_add(qsum, qsum, t); }
    }
    if (m >= 0) {          // uniform across the CTA
        __syncthreads();
        tree[tid] = qsum;
        __syncthreads();
#pragma unroll 1
        for (int stride = EV2_THREADS / 2; stride >= 1; stride >>= 1) {
            if (tid < stride) { Fr a = tree[tid], c = tree[tid + stride]; fr_add(a, a, c); tree[tid] = a; }
            __syncthreads();
        }
        if (tid == 0) {
            const uint32_t e = brp12((uint32_t)m);
            Fr winv = roots[brp12((KB_N - e) & (KB_N - 1))];
            Fr q = tree[0], nq; fr_mul(q, q, winv); fr_neg(nq, q);
            uint4* dst = reinterpret_cast<uint4*>(q_out + ((size_t)b * KB_N + m) * 8);
            dst[0] = make_uint4(nq.l[0], nq.l[1], nq.l[2], nq.l[3]); dst[1] = make_uint4(nq.l[4], nq.l[5], nq.l[6], nq.l[7]);
        }
    }
}

}  // namespace kb200
