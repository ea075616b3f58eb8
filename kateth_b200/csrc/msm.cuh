// msm.cuh -- fixed-base multi-scalar multiplication over the 4096 Lagrange points of the trusted setup.
//
// Replaces P1::lincomb_pippenger (/root/reference/src/bls.rs:416-437: p1_affines::from on every call +
// blst_p1s_mult_pippenger over 255 bits) as used by Blob::commitment (/root/reference/src/blob.rs:48-53) and
// Polynomial::prove (/root/reference/src/kzg/poly.rs:68).
//
// B200 design: the bases never change, HBM is 180 GB, so the Pippenger *bucket* phase is precomputed away.
// For every base i, window j and digit magnitude d in [1, 2^(c-1)] the table holds the affine point
//     T[j][d][i] = d * 2^(c*j) * L_brp[i]                 (96 bytes, Montgomery x,y; (0,0) = infinity)
// so an MSM is: signed-digit recode each scalar (c-bit windows), gather one table point per (scalar, window),
// negate y for negative digits, and add it into a per-thread XYZZ accumulator (mixed addition, 8M+2S); then a
// tree reduction per CTA.  No buckets, no sorting, no atomics, no inter-thread conflicts; 4096*W gathers of 96 B
// per MSM (W = floor(255/c)+1; c = 13 -> 20 windows, 32 GB table).
//
// Table layout in HBM: entry (j, d, i) at index ((j << (c-1)) + (d-1)) * 4096 + i, i.e. i fastest: the build
// kernel writes coalesced rows; the MSM's gathers are random 96-byte reads (3 full 32-byte sectors each).
#pragma once
#include "g1.cuh"
#include "msm_table.cuh"

namespace kb200 {

__device__ __forceinline__ void load_affine(G1Affine& p, const uint4* src) {
    uint4 v[6];
#pragma unroll
    for (int k = 0; k < 6; k++) v[k] = __ldg(src + k);
#pragma unroll
    for (int k = 0; k < 3; k++) {
        p.x.l[4 * k] = v[k].x; p.x.l[4 * k + 1] = v[k].y; p.x.l[4 * k + 2] = v[k].z; p.x.l[4 * k + 3] = v[k].w;
        p.y.l[4 * k] = v[k + 3].x; p.y.l[4 * k + 1] = v[k + 3].y; p.y.l[4 * k + 2] = v[k + 3].z; p.y.l[4 * k + 3] = v[k + 3].w;
    }
}
__device__ __forceinline__ void store_affine(uint4* dst, const G1Affine& p) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
        dst[k] = make_uint4(p.x.l[4 * k], p.x.l[4 * k + 1], p.x.l[4 * k + 2], p.x.l[4 * k + 3]);
        dst[k + 3] = make_uint4(p.y.l[4 * k], p.y.l[4 * k + 1], p.y.l[4 * k + 2], p.y.l[4 * k + 3]);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Table build.  One thread per (base i, window j): B = 2^(c*j) * L_i (affine), then the chain d*B for d = 1..D
// in XYZZ with an explicit Z (ZZ = Z^2, ZZZ = Z^3), X,Y parked in the table slot and (Z, prefix product) in a
// same-shaped scratch array; one inversion per thread and a backward sweep normalise everything to affine.
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_table_build(uint4* table, uint4* scratch, const G1Affine* bases, int c, int nwin, int dlast, int j0, int jcount) {
    int gid = blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= KB_N * jcount) return;
    int i = gid & (KB_N - 1), j = j0 + (gid >> 12);
    scratch -= ((size_t)j0 << (c - 1)) * KB_N * 6;      // scratch covers windows [j0, j0 + jcount) only
    int D = 1 << (c - 1);
    int nd = (j == nwin - 1) ? dlast : D;
    size_t base_e = ((size_t)j << (c - 1)) * KB_N + i;     // entry index of d = 1
    G1Affine b = bases[i];
    if (g1a_is_inf(b)) {
        for (int d = 0; d < nd; d++) { uint4* dst = table + (base_e + (size_t)d * KB_N) * 6; for (int k = 0; k < 6; k++) dst[k] = make_uint4(0, 0, 0, 0); }
        return;
    }
    {   // B = 2^(c*j) * L_i
        G1 t; g1_from_affine(t, b);
        for (int k = 0; k < c * j; k++) g1_dbl(t);
        g1_to_affine(b, t);
    }
    G1 acc; Fp z, pref;
    fp_set(pref, c_k.fp_one);
    // forward
    for (int d = 1; d <= nd; d++) {
        if (d == 1) { g1_from_affine(acc, b); fp_set(z, c_k.fp_one); }
        else if (d == 2) { g1_dbl_affine(acc, b.x, b.y); fp_dbl(z, b.y); }
        else {
            // acc += b with Z tracked: Z3 = Z1 * P.  d*B != +-B and != infinity for 3 <= d <= 2^(c-1) < r.
            Fp p, r, pp, ppp, q, t;
            fp_mul(p, b.x, acc.zz); fp_sub(p, p, acc.x);
            fp_mul(r, b.y, acc.zzz); fp_sub(r, r, acc.y);
            fp_sqr(pp, p); fp_mul(ppp, p, pp); fp_mul(q, acc.x, pp);
            fp_sqr(t, r); fp_sub(t, t, ppp); fp_sub(t, t, q); fp_sub(acc.x, t, q);
            fp_sub(q, q, acc.x); fp_mul(q, r, q); fp_mul(t, acc.y, ppp); fp_sub(acc.y, q, t);
            fp_mul(acc.zz, acc.zz, pp); fp_mul(acc.zzz, acc.zzz, ppp);
            fp_mul(z, z, p);
        }
        size_t e = (base_e + (size_t)(d - 1) * KB_N) * 6;
        G1Affine xy; xy.x = acc.x; xy.y = acc.y; store_affine(table + e, xy);
        G1Affine zp; zp.x = z; zp.y = pref; store_affine(scratch + e, zp);
        fp_mul(pref, pref, z);
    }
    // backward
    Fp inv; fp_inv_bgcd(inv, pref);
    for (int d = nd; d >= 1; d--) {
        size_t e = (base_e + (size_t)(d - 1) * KB_N) * 6;
        G1Affine xy, zp; load_affine(xy, table + e); load_affine(zp, scratch + e);
        Fp zi, zi2, zi3;
        fp_mul(zi, inv, zp.y); fp_mul(inv, inv, zp.x);
        fp_sqr(zi2, zi); fp_mul(zi3, zi2, zi);
        fp_mul(xy.x, xy.x, zi2); fp_mul(xy.y, xy.y, zi3);
        store_affine(table + e, xy);
    }
}

// ---------------------------------------------------------------------------------------------------------
// MSM.  grid = n_items * S CTAs of 128 threads; CTA (item, slice) sums the points [slice*4096/S, (slice+1)*4096/S)
// of item's scalars and writes one XYZZ partial.  S = 1 is the throughput shape (one CTA per blob, 32 bases per
// thread); S = 32 spreads one blob over 32 CTAs (latency shape).
//   FMT 0: scalars are the blob itself (32-byte big-endian canonical field elements, kateth's wire format)
//   FMT 1: scalars are 8 little-endian 32-bit limbs each (canonical; the quotient kernel's output)
// Items whose status is non-zero are skipped (their partial is infinity).
// ---------------------------------------------------------------------------------------------------------
#define MSM_THREADS 128
template <int FMT>
__global__ void __launch_bounds__(MSM_THREADS) k_msm(MsmTable tab, const uint32_t* __restrict__ scalars, const int32_t* __restrict__ status,
                                                      G1* __restrict__ partials, int S) {
    __shared__ G1 sm[MSM_THREADS / 2];
    const int item = blockIdx.x / S, slice = blockIdx.x - item * S;
    const int tid = threadIdx.x;
    G1 acc; g1_set_inf(acc);
    if (status == nullptr || status[item] == 0) {
        const int per_slice = KB_N / S;
        const int c = tab.c, nwin = tab.nwin;
        const uint32_t mask = (1u << c) - 1, D = 1u << (c - 1);
        const uint32_t* sbase = scalars + (size_t)item * (KB_N * 8);
        // GLV tables (tab.glv): pass h = 1 sums the k2 halves, the sum is mapped through (X, Y) -> (beta X, -Y), then pass
        // h = 0 adds the k1 halves on top.  Plain tables: one pass over the whole scalar.
#pragma unroll 1
        for (int h = tab.glv; h >= 0; h--) {
#pragma unroll 1
        for (int i = slice * per_slice + tid; i < (slice + 1) * per_slice; i += MSM_THREADS) {
            uint32_t s[8];
            {
                const uint4* sp = reinterpret_cast<const uint4*>(sbase + (size_t)i * 8);
                uint4 a = __ldg(sp), b = __ldg(sp + 1);
                if (FMT == 0) {
                    s[7] = bswap32(a.x); s[6] = bswap32(a.y); s[5] = bswap32(a.z); s[4] = bswap32(a.w);
                    s[3] = bswap32(b.x); s[2] = bswap32(b.y); s[1] = bswap32(b.z); s[0] = bswap32(b.w);
                } else {
                    s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
                }
            }
            if (tab.glv) {
                uint32_t k1[4], k2[4];
                glv_split(k1, k2, s);
#pragma unroll
                for (int k = 0; k < 4; k++) { s[k] = h ? k2[k] : k1[k]; s[4 + k] = 0; }
            }
            uint32_t carry = 0;
            size_t e = (size_t)i;                                  // entry index of (window 0, d = 1, base i)
#pragma unroll 1
            for (int j = 0; j < nwin; j++) {
                uint32_t d = (s[0] & mask) + carry;
                // s >>= c
#pragma unroll
                for (int k = 0; k < 7; k++) s[k] = __funnelshift_r(s[k], s[k + 1], c);
                s[7] >>= c;
                bool neg = d > D && j != nwin - 1;                 // the top window is unsigned (its digits are bounded by dlast)
                carry = neg ? 1u : 0u;
                uint32_t mag = neg ? ((1u << c) - d) : d;
                if (mag != 0) {
                    G1Affine pt;
                    load_affine(pt, tab.data + (e + (size_t)(mag - 1) * KB_N) * 6);
                    g1_madd_signed(acc, pt, neg);
                }
                e += (size_t)D * KB_N;
            }
        }
        if (h == 1) { Fp b; fp_set(b, c_k.fp_beta); fp_mul(acc.x, acc.x, b); fp_neg(acc.y, acc.y); }
        }
    }
    // CTA tree reduction through shared memory
#pragma unroll 1
    for (int stride = MSM_THREADS / 2; stride >= 1; stride >>= 1) {
        if (tid >= stride && tid < 2 * stride) sm[tid - stride] = acc;
        __syncthreads();
        if (tid < stride) g1_add(acc, sm[tid]);
        __syncthreads();
    }
    if (tid == 0) partials[blockIdx.x] = acc;
}

// ---------------------------------------------------------------------------------------------------------
// Sum the S <= 32 partials of each item, normalise (one inversion) and emit the 48-byte compressed point
// (Compress::compress, /root/reference/src/bls.rs:491-503).  Optionally also keeps the affine point.
// One WARP per item: lane l holds partial l, five shuffle levels add them pairwise (the single-blob latency shape has
// S = 32 partials; summing them one after the other on one thread cost half of the call), lane 0 normalises.
// ---------------------------------------------------------------------------------------------------------
#define FIN_WARPS 4
__device__ __forceinline__ void g1_shfl_down(G1& r, const G1& v, int off) {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(&v);
    uint32_t* dst = reinterpret_cast<uint32_t*>(&r);
#pragma unroll
    for (int k = 0; k < 48; k++) dst[k] = __shfl_down_sync(0xffffffffu, src[k], off);
}
__global__ void __launch_bounds__(32 * FIN_WARPS) k_msm_finish(const G1* __restrict__ partials, int S, int n, const int32_t* __restrict__ status,
                                                                uint32_t* __restrict__ out48, G1Affine* __restrict__ out_affine) {
    const int item = blockIdx.x * FIN_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (item >= n) return;                                  // whole warp
    uint32_t* o = out48 + (size_t)item * 12;
    if (status != nullptr && status[item] != 0) {
        if (lane < 12) o[lane] = 0;
        return;
    }
    G1 acc;
    if (lane < S) acc = partials[(size_t)item * S + lane]; else g1_set_inf(acc);
    if (S > 1) {
#pragma unroll 1
        for (int off = 16; off >= 1; off >>= 1) {
            G1 other; g1_shfl_down(other, acc, off);
            if (lane < off && off < S) g1_add(acc, other);
        }
    }
    if (lane == 0) {
        G1Affine a; g1_to_affine(a, acc);
        if (out_affine) out_affine[item] = a;
        uint32_t w[12]; g1_compress(w, a);
        for (int k = 0; k < 12; k++) o[k] = w[k];
    }
}

}  // namespace kb200
