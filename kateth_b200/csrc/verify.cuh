// verify.cuh -- the verification side: G1 decompression + subgroup check, and the random-linear-combination sums
// of batch verification.
//
// Replaces (file:line in /root/reference):
//   P1::decompress                    src/bls.rs:505-531        (k_g1_decompress)
//   Setup::verify_proof_batch         src/kzg/setup.rs:115-161  (k_batch_r; k_rlc or k_rlc_scalars / k_rlc_bucket / k_rlc_shift; k_rlc_s,
//                                                               k_rlc_finish, k_rlc_combine) up to,
//                                                               but not including, the two-pairing check (host)
//   Setup::verify_proof_inner         src/kzg/setup.rs:84-94    (same kernels in "raw" mode, weights = 1)
// kateth's quirks are reproduced: r = SHA-256("RCKZGBATCH___V1_" | u128_be(4096) | u128_be(n)) mod r depends on n
// only, and rpowers = [r, r, r^2, r^3, ...] because Fr::pow(x, 0) returns x (src/bls.rs:169-187).
//
// Per GPU the result is one 336-byte record {A, B, C as affine Montgomery points, s, err[4]}:
//   A = sum w_i pi_i,  B = sum w_i z_i pi_i,  C = sum w_i C_i,  s = sum w_i y_i        (w_i = rpowers[i])
//   err = the shard's first error of each class (blob batches: blob, commitment, proof; point batches: proof, commitment, z, y;
//   0 = none): errors travel WITH the sums, so that every
//   rank that finishes a sharded batch returns the error kateth would (all blobs, then all commitments, then all proofs,
//   src/kzg/setup.rs:259-271) instead of a verdict computed from unvalidated terms.  Items with a non-zero status contribute
//   nothing to the sums.
// Records of several GPUs are exchanged with one all-gather, then k_rlc_combine forms the pairing operands
//   lhs = sum_g A_g            rhs = sum_g C_g - (sum_g s_g) * G1 + sum_g B_g
// and the host checks e(lhs, [tau]_2) == e(rhs, G2).
#pragma once
#include "poly.cuh"

namespace kb200 {

struct RlcRecord { G1Affine a, b, c; Fr s; int32_t err[4]; };   // 3*96 + 32 + 16 = 336 bytes
static_assert(sizeof(RlcRecord) == 336, "record layout");

// Two arrays of n compressed points each (the second may be null) in ONE launch, one thread per point: out = affine
// Montgomery; status[i] set to 3/4/5 on failure (kept if already non-zero).  by_order != 0 uses the [r]P definition of
// subgroup membership instead of the endomorphism test (unit test only).
__global__ void __launch_bounds__(128) k_g1_decompress(const uint32_t* __restrict__ in_a, const uint32_t* __restrict__ in_b, int n,
                                                      G1Affine* __restrict__ out_a, G1Affine* __restrict__ out_b, int32_t* __restrict__ st_a,
                                                      int32_t* __restrict__ st_b, int by_order) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (in_b ? 2 * n : n)) return;
    const uint32_t* in48 = in_a; G1Affine* out = out_a; int32_t* status = st_a;
    if (i >= n) { i -= n; in48 = in_b; out = out_b; status = st_b; }
    uint32_t w[12];
    for (int k = 0; k < 12; k++) w[k] = in48[(size_t)i * 12 + k];
    G1Affine p;
    int st = g1_uncompress(p, w);
    if (st == KB_OK && !(by_order ? g1_in_subgroup_by_order(p) : g1_in_subgroup(p))) st = KB_G1_NOT_IN_GROUP;
    if (st != KB_OK) { fp_zero(p.x); fp_zero(p.y); if (status[i] == 0) status[i] = st; }
    out[i] = p;
}

// r (Montgomery) for a batch of n_total proofs
__global__ void k_batch_r(uint32_t n_total, Fr* r_mont) {
    uint32_t h[8], w[16];
    sha256_init(h);
    w[0] = 0x52434b5a; w[1] = 0x47424154; w[2] = 0x43485f5f; w[3] = 0x5f56315f;   // "RCKZGBATCH___V1_"
    w[4] = 0; w[5] = 0; w[6] = 0; w[7] = KB_N;
    w[8] = 0; w[9] = 0; w[10] = 0; w[11] = n_total;
    w[12] = 0x80000000u; w[13] = 0; w[14] = 0; w[15] = 48 * 8;
    sha256_block(h, w);
    Fr v;
    for (int k = 0; k < 8; k++) v.l[7 - k] = h[k];
    fr_reduce_256(v);
    Fr m; fr_to_mont(m, v);
    *r_mont = m;
}

// Consensus-spec batch challenge (optional mode): the transcript of verify_kzg_proof_batch is
// domain | u64(4096) | u64(n) | (commitment | z | y | proof) per item; this kernel lays out the per-item part (160 bytes
// each) from the compressed points and the canonical scalars; the host hashes it (one sequential SHA-256).
__global__ void k_transcript(const uint32_t* __restrict__ c48, const uint32_t* __restrict__ p48, const Fr* __restrict__ z_canon,
                             const Fr* __restrict__ y_canon, int n, uint32_t* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t* o = out + (size_t)i * 40;
    for (int k = 0; k < 12; k++) o[k] = c48[(size_t)i * 12 + k];
    Fr z = z_canon[i], y = y_canon[i];
    fr_store_be(o + 12, z); fr_store_be(o + 20, y);
    for (int k = 0; k < 12; k++) o[28 + k] = p48[(size_t)i * 12 + k];
}
// r given as 8 plain little-endian limbs (already < 2^256; reduced here) -> Montgomery
__global__ void k_set_r(const uint32_t* plain, Fr* r_mont) {
    Fr v;
    for (int k = 0; k < 8; k++) v.l[k] = plain[k];
    fr_reduce_256(v);
    Fr m; fr_to_mont(m, v);
    *r_mont = m;
}

#define RLC_THREADS 64
// grid (ceil(n/RLC_THREADS), 3).  which = blockIdx.y: 0 -> w_i*pi_i, 1 -> (w_i z_i)*pi_i, 2 -> w_i*C_i.
// raw != 0: weights are 1 (single-proof check, src/kzg/setup.rs:84-94 rearranged to avoid the G2 scalar mul:
//   e(pi, [tau]_2) == e(C - y G1 + z pi, G2)).
__global__ void __launch_bounds__(RLC_THREADS) k_rlc(const G1Affine* __restrict__ proofs, const G1Affine* __restrict__ comms,
                                                      const Fr* __restrict__ z_canon, const Fr* __restrict__ y_canon, int n, uint32_t index_offset,
                                                      const Fr* __restrict__ r_mont, int raw, int spec, const int32_t* __restrict__ status,
                                                      G1* __restrict__ partials, Fr* __restrict__ s_partials) {
    __shared__ G1 sm[RLC_THREADS / 2];
    __shared__ Fr sms[RLC_THREADS];
    const int tid = threadIdx.x, which = blockIdx.y;
    // two threads per term: the scalar is split along the endomorphism (fr_glv_split) and each thread runs a 128-bit
    // double-and-add; the halves need no recombination because everything is summed anyway.
    const int t = blockIdx.x * RLC_THREADS + tid;
    const int i = t >> 1, half = t & 1;
    G1 acc; g1_set_inf(acc);
    Fr s; fr_zero(s);
    if (i < n && (status == nullptr || status[i] == 0)) {     // an item that failed validation contributes nothing
        Fr wgt;    // Montgomery
        if (raw) fr_set(wgt, c_k.fr_one);
        else {
            uint32_t gi = index_offset + (uint32_t)i;
            Fr r = *r_mont;
            fr_pow_u32(wgt, r, (gi == 0 && !spec) ? 1u : gi);     // Fr::pow(r, 0) == r quirk (kateth); r^0 = 1 in spec mode
        }
        Fr k;      // canonical scalar
        if (which == 1) { Fr z = z_canon[i]; fr_mul(k, z, wgt); }
        else fr_from_mont(k, wgt);
        G1Affine p = (which == 2) ? comms[i] : proofs[i];
        Fr k1, k2; fr_glv_split(k1, k2, k);
        if (half) {
            if (!g1a_is_inf(p)) { Fp b; fp_set(b, c_k.fp_beta); fp_mul(p.x, p.x, b); fp_neg(p.y, p.y); }      // Q = (beta x, -y)
            k1 = k2;
        } else if (which == 0 && s_partials != nullptr) { Fr y = y_canon[i]; fr_mul(s, y, wgt); }
        g1_mul(acc, p, k1, 128);          // ONE call site: the two halves of a warp must not serialise their scalar multiplications
    }
#pragma unroll 1
    for (int stride = RLC_THREADS / 2; stride >= 1; stride >>= 1) {
        if (tid >= stride && tid < 2 * stride) sm[tid - stride] = acc;
        __syncthreads();
        if (tid < stride) g1_add(acc, sm[tid]);
        __syncthreads();
    }
    if (tid == 0) partials[(size_t)which * gridDim.x + blockIdx.x] = acc;
    if (which == 0 && s_partials != nullptr) {
        sms[tid] = s;
        __syncthreads();
#pragma unroll 1
        for (int stride = RLC_THREADS / 2; stride >= 1; stride >>= 1) {
            if (tid < stride) { Fr a = sms[tid], c = sms[tid + stride]; fr_add(a, a, c); sms[tid] = a; }
            __syncthreads();
        }
        if (tid == 0) s_partials[blockIdx.x] = sms[0];
    }
}
// The scalar part of the record on its own: s_partials[b] = sum of w_i * y_i over the items of k_rlc's block b (RLC_THREADS / 2 items).
// Only this needs the evaluations y, so batch verification runs the barycentric evaluation BESIDE the scalar multiplications of
// k_rlc (which need the challenges z only) and joins here.
__global__ void __launch_bounds__(RLC_THREADS / 2) k_rlc_s(const Fr* __restrict__ y_canon, int n, uint32_t index_offset, const Fr* __restrict__ r_mont, int raw, int spec,
                                                          const int32_t* __restrict__ status, Fr* __restrict__ s_partials) {
    __shared__ Fr sms[RLC_THREADS / 2];
    const int tid = threadIdx.x, i = blockIdx.x * (RLC_THREADS / 2) + tid;
    Fr s; fr_zero(s);
    if (i < n && (status == nullptr || status[i] == 0)) {
        Fr wgt;
        if (raw) fr_set(wgt, c_k.fr_one);
        else { const uint32_t gi = index_offset + (uint32_t)i; Fr r = *r_mont; fr_pow_u32(wgt, r, (gi == 0 && !spec) ? 1u : gi); }
        Fr y = y_canon[i]; fr_mul(s, y, wgt);
    }
    sms[tid] = s;
    __syncthreads();
#pragma unroll 1
    for (int stride = RLC_THREADS / 4; stride >= 1; stride >>= 1) {
        if (tid < stride) { Fr a = sms[tid], c = sms[tid + stride]; fr_add(a, a, c); sms[tid] = a; }
        __syncthreads();
    }
    if (tid == 0) s_partials[blockIdx.x] = sms[0];
}
// ---------------------------------------------------------------------------------------------------------
// The same three sums for LARGE shards as a bucket (Pippenger) multi-scalar multiplication over the batch itself.
// k_rlc spends one 128-step double-and-add ladder per term: 3 x 2 x n ladders, ~760 M warp instructions at n = 4096 -- as much
// multiplier work as the barycentric evaluation it runs beside.  Here the 2n terms (P_i, k1_i), (Q_i, k2_i) of a sum are cut into 16
// byte-wide windows; per (sum, window, slice of 1024 items) one CTA counting-sorts its <= 2048 terms by digit in shared memory, thread d sums
// the points of bucket d (mixed additions), the CTA forms sum_d d * B_d with a suffix scan + tree (16 levels of full additions), and
// k_rlc_shift adds the slices of a window and doubles the window sum 8j times.  ~14x fewer group operations than the ladders; what is
// left on the critical path is the chain of <= 120 doublings no method avoids.  Same group elements, hence the same record bytes.
// ---------------------------------------------------------------------------------------------------------
#define RLCB_THREADS 256
#define RLCB_WINDOWS 16
#define RLCB_ITEMS 1024                           // items per slice: 2048 terms, 8 per thread
#define RLCB_PER_THREAD (2 * RLCB_ITEMS / RLCB_THREADS)

// dig[(sel * 16 + j) * n + i] = byte j of k1 | byte j of k2 << 8, (k1, k2) = GLV halves of the canonical scalar of item i:
// sel 0: w_i (sums A and C), sel 1: w_i z_i (sum B).  Items with a non-zero status get zero digits: they contribute nothing.
__global__ void __launch_bounds__(128) k_rlc_scalars(const Fr* __restrict__ z_canon, int n, uint32_t index_offset, const Fr* __restrict__ r_mont, int raw, int spec,
                                                     const int32_t* __restrict__ status, uint16_t* __restrict__ dig) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr ka, kb; fr_zero(ka); fr_zero(kb);
    if (status == nullptr || status[i] == 0) {
        Fr wgt;
        if (raw) fr_set(wgt, c_k.fr_one);
        else { const uint32_t gi = index_offset + (uint32_t)i; Fr r = *r_mont; fr_pow_u32(wgt, r, (gi == 0 && !spec) ? 1u : gi); }
        fr_from_mont(ka, wgt);
        Fr z = z_canon[i]; fr_mul(kb, z, wgt);
    }
    Fr a1, a2, b1, b2;
    fr_glv_split(a1, a2, ka); fr_glv_split(b1, b2, kb);
#pragma unroll
    for (int j = 0; j < RLCB_WINDOWS; j++) {
        const int sh = 8 * (j & 3), l = j >> 2;
        dig[(size_t)j * n + i] = (uint16_t)(((a1.l[l] >> sh) & 0xffu) | (((a2.l[l] >> sh) & 0xffu) << 8));
        dig[(size_t)(RLCB_WINDOWS + j) * n + i] = (uint16_t)(((b1.l[l] >> sh) & 0xffu) | (((b2.l[l] >> sh) & 0xffu) << 8));
    }
}

// grid (16 * nslices, 3): CTA (window j, slice, which) -> wsum[(which * 16 + j) * nslices + slice] = sum over the slice's terms of digit * point
__global__ void __launch_bounds__(RLCB_THREADS, 2) k_rlc_bucket(const G1Affine* __restrict__ proofs, const G1Affine* __restrict__ comms,
                                                             const uint16_t* __restrict__ dig, int n, int nslices, G1* __restrict__ wsum) {
    __shared__ __align__(16) unsigned char raw_sm[RLCB_THREADS * sizeof(G1)];       // 48 KiB: sort arrays first, bucket sums afterwards
    G1* bkt = reinterpret_cast<G1*>(raw_sm);
    uint32_t* cnt = reinterpret_cast<uint32_t*>(raw_sm);
    uint32_t* cur = cnt + RLCB_THREADS;
    uint16_t* sorted = reinterpret_cast<uint16_t*>(cur + RLCB_THREADS);
    const int tid = threadIdx.x, which = blockIdx.y, j = blockIdx.x % RLCB_WINDOWS, slice = blockIdx.x / RLCB_WINDOWS;
    const int i0 = slice * RLCB_ITEMS, i1 = (n < i0 + RLCB_ITEMS) ? n : i0 + RLCB_ITEMS, T = 2 * (i1 - i0);
    const uint16_t* dg = dig + (size_t)((which == 1 ? RLCB_WINDOWS : 0) + j) * n;
    const G1Affine* pts = (which == 2) ? comms : proofs;
    cnt[tid] = 0;
    __syncthreads();
    uint32_t myd[RLCB_PER_THREAD];
#pragma unroll
    for (int k = 0; k < RLCB_PER_THREAD; k++) {
        const int t = tid + RLCB_THREADS * k;
        uint32_t d = 0;
        if (t < T) d = ((uint32_t)dg[i0 + (t >> 1)] >> (8 * (t & 1))) & 0xffu;
        myd[k] = d;
        if (d) atomicAdd(&cnt[d], 1u);
    }
    __syncthreads();
    uint32_t off = 0;
    for (int e = 1; e < tid; e++) off += cnt[e];
    const uint32_t mycnt = cnt[tid];              // bucket 0 stays empty
    cur[tid] = off;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RLCB_PER_THREAD; k++)
        if (myd[k]) { const uint32_t pos = atomicAdd(&cur[myd[k]], 1u); sorted[pos] = (uint16_t)(tid + RLCB_THREADS * k); }
    __syncthreads();
    G1 acc; g1_set_inf(acc);
#pragma unroll 1
    for (uint32_t q = off; q < off + mycnt; q++) {
        const uint32_t t = sorted[q];
        G1Affine pt = pts[i0 + (int)(t >> 1)];
        if (g1a_is_inf(pt)) continue;
        if (t & 1) { Fp b; fp_set(b, c_k.fp_beta); FPM(pt.x, pt.x, b); fp_neg(pt.y, pt.y); }        // Q = (beta x, -y)
        g1_madd_c(acc, pt.x, pt.y);
    }
    __syncthreads();                               // the sort arrays are dead: their bytes become the bucket sums
    bkt[tid] = acc;
    __syncthreads();
    // suffix sums S_d = sum_{e >= d} B_e (Hillis-Steele), then sum_{d >= 1} S_d = sum_d d * B_d
#pragma unroll 1
    for (int st = 1; st < RLCB_THREADS; st <<= 1) {
        const bool has = tid + st < RLCB_THREADS;
        G1 other;
        if (has) other = bkt[tid + st];
        __syncthreads();
        if (has) { g1_add(acc, other); bkt[tid] = acc; }
        __syncthreads();
    }
    if (tid == 0) { g1_set_inf(acc); bkt[0] = acc; }
    __syncthreads();
#pragma unroll 1
    for (int stride = RLCB_THREADS / 2; stride >= 1; stride >>= 1) {
        if (tid < stride) { G1 other = bkt[tid + stride]; g1_add(acc, other); bkt[tid] = acc; }
        __syncthreads();
    }
    if (tid == 0) wsum[(size_t)(which * RLCB_WINDOWS + j) * nslices + slice] = acc;
}
// grid (16, 3), one warp: window sum = sum of its slices, shifted: partials[which * 16 + j] = 2^(8j) * W_j  (k_rlc_finish adds the 16)
__global__ void __launch_bounds__(32) k_rlc_shift(const G1* __restrict__ wsum, int nslices, G1* __restrict__ partials) {
    __shared__ G1 sm[16];
    const int lane = threadIdx.x, j = blockIdx.x, which = blockIdx.y;
    G1 acc; g1_set_inf(acc);
    for (int k = lane; k < nslices; k += 32) { G1 t = wsum[(size_t)(which * RLCB_WINDOWS + j) * nslices + k]; g1_add(acc, t); }
#pragma unroll 1
    for (int stride = 16; stride >= 1; stride >>= 1) {
        if (stride < nslices) {                    // uniform: lanes >= nslices hold infinity
            if (lane >= stride && lane < 2 * stride) sm[lane - stride] = acc;
            __syncwarp();
            if (lane < stride) g1_add(acc, sm[lane]);
            __syncwarp();
        }
    }
    if (lane == 0) {
        g1_dbl_k(acc, 8 * j);
        partials[which * RLCB_WINDOWS + j] = acc;
    }
}

// grid 3 CTAs of RLC_FIN_THREADS: CTA `which` reduces its row of partials, normalises, writes the record field.
#define RLC_FIN_THREADS 128
__global__ void __launch_bounds__(RLC_FIN_THREADS) k_rlc_finish(const G1* __restrict__ partials, int nparts, const Fr* __restrict__ s_partials, int nparts_s,
                                                                 RlcRecord* __restrict__ rec) {
    __shared__ G1 sm[RLC_FIN_THREADS / 2];
    __shared__ Fr sms[RLC_FIN_THREADS];
    const int tid = threadIdx.x, which = blockIdx.x;
    G1 acc; g1_set_inf(acc);
    for (int k = tid; k < nparts; k += RLC_FIN_THREADS) { G1 t = partials[(size_t)which * nparts + k]; g1_add(acc, t); }
#pragma unroll 1
    for (int stride = RLC_FIN_THREADS / 2; stride >= 1; stride >>= 1) {
        if (tid >= stride && tid < 2 * stride) sm[tid - stride] = acc;
        __syncthreads();
        if (tid < stride) g1_add(acc, sm[tid]);
        __syncthreads();
    }
    if (tid == 0) {
        G1Affine a; g1_to_affine(a, acc);
        if (which == 0) rec->a = a; else if (which == 1) rec->b = a; else rec->c = a;
    }
    if (which == 0) {
        Fr s; fr_zero(s);
        for (int k = tid; k < nparts_s; k += RLC_FIN_THREADS) { Fr t = s_partials[k]; fr_add(s, s, t); }
        sms[tid] = s;
        __syncthreads();
#pragma unroll 1
        for (int stride = RLC_FIN_THREADS / 2; stride >= 1; stride >>= 1) {
            if (tid < stride) { Fr a = sms[tid], c = sms[tid + stride]; fr_add(a, a, c); sms[tid] = a; }
            __syncthreads();
        }
        if (tid == 0) { rec->s = sms[0]; for (int k = 0; k < 4; k++) rec->err[k] = 0; }
    }
}
// Several records of ONE rank (the chunks of a host-buffer shard) -> one record: lane 0..2 sum A, B, C, lane 3 sums s and keeps
// the first error of each class in chunk order.
__global__ void __launch_bounds__(32) k_rlc_merge(const RlcRecord* __restrict__ recs, int nrec, RlcRecord* __restrict__ out) {
    const int lane = threadIdx.x;
    if (lane < 3) {
        G1 acc; g1_set_inf(acc);
        for (int g = 0; g < nrec; g++) { const G1Affine* p = lane == 0 ? &recs[g].a : (lane == 1 ? &recs[g].b : &recs[g].c); G1Affine t = *p; g1_madd_signed(acc, t, false); }
        G1Affine a; g1_to_affine(a, acc);
        if (lane == 0) out->a = a; else if (lane == 1) out->b = a; else out->c = a;
    } else if (lane == 3) {
        Fr s; fr_zero(s);
        int32_t e[4] = {0, 0, 0, 0};
        for (int g = 0; g < nrec; g++) {
            Fr t = recs[g].s; fr_add(s, s, t);
            for (int k = 0; k < 4; k++) if (e[k] == 0) e[k] = recs[g].err[k];
        }
        out->s = s;
        for (int k = 0; k < 4; k++) out->err[k] = e[k];
    }
}
// neg_g_tab[w * 256 + d] = d * 2^(8w) * (-G1), affine, w = 0..31, d = 0..255 (d = 0: the infinity sentinel): fixed-base table for the
// [s](-G1) term of the combine step, built at setup (768 KiB).  One thread per entry: 8w doublings, an 8-step ladder, one inversion.
__global__ void __launch_bounds__(256) k_neg_g_table(G1Affine* __restrict__ neg_g_tab) {
    const int w = blockIdx.x, d = threadIdx.x;
    G1Affine ng; fp_set(ng.x, c_k.g1_gen_x); fp_set(ng.y, c_k.g1_neg_gen_y);
    G1 t; g1_from_affine(t, ng);
    g1_dbl_k(t, 8 * w);
    G1 acc; g1_set_inf(acc);
#pragma unroll 1
    for (int i = 7; i >= 0; i--) { g1_dbl(acc); if ((d >> i) & 1) g1_add(acc, t); }
    G1Affine a; g1_to_affine(a, acc);
    neg_g_tab[w * 256 + d] = a;
}
// Combine the records of all ranks into the two G1 pairing operands, as plain (non-Montgomery) affine coordinates:
// out[0..11] = lhs.x, [12..23] = lhs.y, [24..35] = rhs.x, [36..47] = rhs.y, out[48] = lhs is infinity, out[49] = rhs is infinity,
// out[50..53] = first error of each of the four classes over the records in rank order (0 = none).
//   lhs = sum_g A_g        rhs = sum_g C_g + [sum_g s_g](-G1) + sum_g B_g
// One warp: lane w looks up [byte w of s] 2^(8w)(-G1) in the table, the warp tree-sums the 32 terms, lane 0 adds the records, lanes 0
// and 1 normalise one operand each.
__global__ void __launch_bounds__(32) k_rlc_combine(const RlcRecord* __restrict__ recs, int nrec, const G1Affine* __restrict__ neg_g_tab,
                                                    uint32_t* __restrict__ out) {
    __shared__ G1 sm[16];
    const int lane = threadIdx.x;
    Fr s; fr_zero(s);
    for (int g = 0; g < nrec; g++) { Fr t = recs[g].s; fr_add(s, s, t); }
    const uint32_t digit = (s.l[lane >> 2] >> (8 * (lane & 3))) & 0xffu;
    G1 acc;
    { G1Affine e = neg_g_tab[lane * 256 + digit]; g1_from_affine(acc, e); }
#pragma unroll 1
    for (int stride = 16; stride >= 1; stride >>= 1) {
        if (lane >= stride && lane < 2 * stride) sm[lane - stride] = acc;
        __syncwarp();
        if (lane < stride) g1_add(acc, sm[lane]);
        __syncwarp();
    }
    if (lane == 2) {
        int32_t e[4] = {0, 0, 0, 0};
        for (int g = 0; g < nrec; g++)
            for (int k = 0; k < 4; k++) if (e[k] == 0) e[k] = recs[g].err[k];
        for (int k = 0; k < 4; k++) out[50 + k] = (uint32_t)e[k];
    }
    if (lane == 0) {
        G1 A, B; g1_set_inf(A); g1_set_inf(B);
        for (int g = 0; g < nrec; g++) {
            RlcRecord r = recs[g];
            if (!g1a_is_inf(r.a)) g1_madd_c(A, r.a.x, r.a.y);
            if (!g1a_is_inf(r.b)) g1_madd_c(B, r.b.x, r.b.y);
            if (!g1a_is_inf(r.c)) g1_madd_c(acc, r.c.x, r.c.y);
        }
        g1_add(acc, B);
        sm[0] = A; sm[1] = acc;
    }
    __syncwarp();
    if (lane < 2) {
        G1 t = sm[lane];
        G1Affine a; g1_to_affine(a, t);
        Fp v;
        fp_from_mont(v, a.x); for (int k = 0; k < 12; k++) out[24 * lane + k] = v.l[k];
        fp_from_mont(v, a.y); for (int k = 0; k < 12; k++) out[24 * lane + 12 + k] = v.l[k];
        out[48 + lane] = g1a_is_inf(a) ? 1u : 0u;
    }
}

}  // namespace kb200
