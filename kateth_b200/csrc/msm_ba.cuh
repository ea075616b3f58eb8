// msm_ba.cuh -- the throughput shape of the fixed-base MSM: batch-affine summation of the gathered table points.
//
// Same job as k_msm (msm.cuh): replaces P1::lincomb_pippenger (/root/reference/src/bls.rs:416-437) behind
// Blob::commitment (src/blob.rs:48-53) and Polynomial::prove (src/kzg/poly.rs:68).  k_msm adds every gathered table
// point into an XYZZ accumulator (8M + 2S = 10 Fp products per point).  All 4096 * W points of one MSM only have to be
// SUMMED, so here they are first added pairwise in AFFINE coordinates, where an addition costs 1S + 2M once the
// inverse of (x2 - x1) is known, and the inverses of a whole batch come from one shared inversion (3 products per
// element for the prefix/suffix products + one binary-GCD inversion per batch, fp_bgcd.cuh: ~10 products' worth of
// multiplier work, ~44k integer-ALU instructions).  A thread owns 32 bases x W windows = 32W gathered points and halves
// them five times:
//     round 1: 16W pairs of table points -> 16W affine sums;  rounds 2..5: 8W, 4W, 2W, W pairs of the previous sums;
//     then W mixed additions into the thread's XYZZ accumulator,
// i.e. (6 * 31W + 10 * W) / 32W = 6.1 products per gathered point (+ 5 inversions per thread) instead of 10: 2.13e5
// instead of 3.48e5 Fp products per 4096-point MSM at W = 17.  The sums of a round live in
// a per-CTA scratch area in HBM, interleaved by thread so every access is a coalesced 512-byte warp transaction; the
// table points are simply gathered twice (once for the denominators, once for the additions).
// Exceptional pairs (an operand at infinity = zero digit, P == Q, P == -Q) are detected from the x coordinates in both
// passes and take a slow path with the complete formulas (doubling shares the batch inversion with denominator 2y),
// so the 48-byte result is the same group element whatever the scalars or the setup are.
//
// The scalars are recoded once per MSM by k_msm_recode into 32-bit table row references (signed digits: adding
// 2^(c-1) to every window below the top one turns the unsigned window value v into the signed digit v - 2^(c-1)).
#pragma once
#include "msm_table.cuh"
#include "fp_bgcd.cuh"

namespace kb200 {

#define BA_THREADS 128
#define BA_UB_LOG 5
#define BA_UB (1 << BA_UB_LOG)     // bases per unit per thread; a unit is reduced by BA_UB_LOG halving rounds
#define BA_UNITS (KB_N / (BA_THREADS * BA_UB))   // units per thread (one CTA = one blob)
#define BA_ZERO 0xffffffffu        // digit 0: nothing to add
// workspace shape (msm_ba.cu sizes the buffers from these): scratch points / prefix products per CTA, in units of K1 = pairs of round 1
// BA_STAGE 1 (default): the operands of the next pair are copied into shared memory with cp.async while the current pair is being
// added, and read from there (LDS) when they are needed -- nothing waits on HBM inside the loops, and the operands that an addition
// needs again late (x1, x2, y1) are re-read from shared memory instead of being held in registers across the products (the kernel
// sits at the 128-register cap: held, they spill).  Per thread: two buffers of {x1, x2, y1} (the next pair lands in the other
// one), single slots for {prefix product, y2} (consumed before the next pair is requested) and one slot for the digits of the
// pair after that: 25 x 16 bytes, 50 KiB per CTA; the forward pass uses the same bytes as a two-deep ring of (x1, x2).
// BA_STAGE 0: round 1's kernel (operands loaded into registers at the top of each pair; A/B reference).
#ifndef BA_STAGE
#define BA_STAGE 1
#endif
#ifndef BA_SPLIT_TAB
#define BA_SPLIT_TAB 0   // 1: two instantiations of a round's loops (table operands / scratch operands)
#endif
#ifndef BA_SMEM_BYTES
#define BA_STAGE_SLOTS 25
#define BA_SMEM_BYTES (BA_STAGE ? BA_STAGE_SLOTS * BA_THREADS * 16 : 0)            // dynamic shared memory of k_msm_ba
#endif
#define BA_PTS_SLOTS_NUM 1
#define BA_PTS_SLOTS_DEN 1
#define BA_PRE_SLOTS_NUM 1
#define BA_PRE_SLOTS_DEN 1

struct BaOffset { uint32_t l[8]; };     // sum_{j < nwin-1} 2^(c-1) * 2^(c*j)

// One out-of-line copy of the Montgomery product for everything in this kernel.  The batch-affine kernel has its warps
// in different phases (prefix products, inversion, additions, final accumulation) at the same time; with the 400-
// instruction product inlined at every call site the instruction working set is > 100 KB and the warps stall on
// instruction fetch (ncu: no_instructions 28 %).  Arguments and result travel in registers (by-value structs: no stack).
static __device__ __noinline__ Fp fp_mulc(Fp a, Fp b) { Fp r; fp_mul(r, a, b); return r; }
#define BA_MUL(r, a, b) r = fp_mulc(a, b)
#define BA_SQR(r, a) r = fp_mulc(a, a)
// BA_INLINE_MASK (experiment): bit i set = product i of the backward pass (0 dinv, 1 inv*d, 2 lambda, 3 lambda^2, 4 lambda*(x1-x3)) is inlined
#ifndef BA_INLINE_MASK
#define BA_INLINE_MASK 0
#endif
#define BA_MULX(i, r, a, b) do { if ((BA_INLINE_MASK >> (i)) & 1) fp_mul(r, a, b); else r = fp_mulc(a, b); } while (0)

// g1_dbl_affine / g1_dbl / g1_madd / g1_add (g1.cuh) on the shared product.  All inlined into the kernel: nothing in
// k_msm_ba passes a pointer to a local object to an out-of-line function, so every operand stays in registers.
__device__ __forceinline__ void ba_dbl_affine(G1& r, const Fp& x, const Fp& y) {
    Fp u, v, w, s, m, t;
    fp_dbl(u, y); BA_SQR(v, u); BA_MUL(w, u, v); BA_MUL(s, x, v);
    BA_SQR(m, x); fp_dbl(t, m); fp_add(m, m, t);
    BA_SQR(r.x, m); fp_sub(r.x, r.x, s); fp_sub(r.x, r.x, s);
    fp_sub(t, s, r.x); BA_MUL(t, m, t); BA_MUL(u, w, y); fp_sub(r.y, t, u);
    r.zz = v; r.zzz = w;
}
__device__ __forceinline__ void ba_dbl(G1& p) {
    if (g1_is_inf(p)) return;
    Fp u, v, w, s, m, t;
    fp_dbl(u, p.y); BA_SQR(v, u); BA_MUL(w, u, v); BA_MUL(s, p.x, v);
    BA_SQR(m, p.x); fp_dbl(t, m); fp_add(m, m, t);
    BA_MUL(t, w, p.y);
    BA_SQR(p.x, m); fp_sub(p.x, p.x, s); fp_sub(p.x, p.x, s);
    fp_sub(s, s, p.x); BA_MUL(s, m, s); fp_sub(p.y, s, t);
    BA_MUL(p.zz, v, p.zz); BA_MUL(p.zzz, w, p.zzz);
}
__device__ __forceinline__ void ba_madd(G1& acc, const Fp& x2, const Fp& y2) {
    if (fp_is_zero(x2)) return;                                     // (0,0): nothing to add
    if (g1_is_inf(acc)) { acc.x = x2; acc.y = y2; fp_set(acc.zz, c_k.fp_one); fp_set(acc.zzz, c_k.fp_one); return; }
    Fp p, r, pp, ppp, q, t;
    BA_MUL(p, x2, acc.zz); fp_sub(p, p, acc.x);
    BA_MUL(r, y2, acc.zzz); fp_sub(r, r, acc.y);
    if (fp_is_zero(p)) {
        if (fp_is_zero(r)) ba_dbl_affine(acc, x2, y2); else g1_set_inf(acc);
        return;
    }
    BA_SQR(pp, p); BA_MUL(ppp, p, pp); BA_MUL(q, acc.x, pp);
    BA_SQR(t, r); fp_sub(t, t, ppp); fp_sub(t, t, q); fp_sub(acc.x, t, q);
    fp_sub(q, q, acc.x); BA_MUL(q, r, q); BA_MUL(t, acc.y, ppp); fp_sub(acc.y, q, t);
    BA_MUL(acc.zz, acc.zz, pp); BA_MUL(acc.zzz, acc.zzz, ppp);
}
__device__ __forceinline__ void ba_add(G1& acc, const G1& b) {
    if (g1_is_inf(b)) return;
    if (g1_is_inf(acc)) { acc = b; return; }
    Fp u1, u2, s1, s2, p, r, pp, ppp, q, t;
    BA_MUL(u1, acc.x, b.zz); BA_MUL(u2, b.x, acc.zz);
    BA_MUL(s1, acc.y, b.zzz); BA_MUL(s2, b.y, acc.zzz);
    fp_sub(p, u2, u1); fp_sub(r, s2, s1);
    if (fp_is_zero(p)) {
        if (fp_is_zero(r)) ba_dbl(acc); else g1_set_inf(acc);
        return;
    }
    BA_SQR(pp, p); BA_MUL(ppp, p, pp); BA_MUL(q, u1, pp);
    BA_SQR(t, r); fp_sub(t, t, ppp); fp_sub(t, t, q); fp_sub(acc.x, t, q);
    fp_sub(q, q, acc.x); BA_MUL(q, r, q); BA_MUL(t, s1, ppp); fp_sub(acc.y, q, t);
    BA_MUL(acc.zz, acc.zz, b.zz); BA_MUL(acc.zz, acc.zz, pp);
    BA_MUL(acc.zzz, acc.zzz, b.zzz); BA_MUL(acc.zzz, acc.zzz, ppp);
}
// inversion on the shared product (fp_inv_bgcd, fp_bgcd.cuh)
static __device__ __noinline__ Fp ba_inv(Fp a) {
    Fp r3 = {{0xd94ca1e0u, 0xed48ac6bu, 0x03a7adf8u, 0x315f831eu, 0x615e29ddu, 0x9a53352au, 0x921e1761u, 0x34c04e5eu, 0x65724728u, 0x2512d435u, 0x91755d4du, 0x0aa63460u}};
    Fp t;
    if (bg_inverse(t.l, a.l)) return fp_mulc(t, r3);
    Fp r; fp_inv(r, a); return r;
}

// digits: dig[((item * BA_UNITS + u) * NP + p) * BA_THREADS + t],  NP = BA_UB * nwin * halves,  p = (h * nwin + j) * BA_UB + b,
// for base i = (u * BA_UB + b) * BA_THREADS + t, window j and GLV half h (halves = 1 for a plain table: the whole scalar).  Value: bit 31 = negative, low bits = table row
// (j << (c-1)) + |d| - 1, or BA_ZERO.
template <int FMT>
__global__ void __launch_bounds__(256) k_msm_recode(MsmTable tab, BaOffset off, const uint32_t* __restrict__ scalars, const int32_t* __restrict__ status,
                                                    uint32_t* __restrict__ dig, int n) {
    size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    int item = (int)(gid >> 12), i = (int)(gid & (KB_N - 1));
    if (item >= n) return;
    if (status != nullptr && status[item] != 0) return;
    uint32_t s[8];
    {
        const uint4* sp = reinterpret_cast<const uint4*>(scalars + gid * 8);
        uint4 a = __ldg(sp), b = __ldg(sp + 1);
        if (FMT == 0) {
            s[7] = bswap32(a.x); s[6] = bswap32(a.y); s[5] = bswap32(a.z); s[4] = bswap32(a.w);
            s[3] = bswap32(b.x); s[2] = bswap32(b.y); s[1] = bswap32(b.z); s[0] = bswap32(b.w);
        } else {
            s[0] = a.x; s[1] = a.y; s[2] = a.z; s[3] = a.w; s[4] = b.x; s[5] = b.y; s[6] = b.z; s[7] = b.w;
        }
    }
    const int c = tab.c, nwin = tab.nwin, halves = tab.glv + 1;
    const uint32_t mask = (1u << c) - 1, D = 1u << (c - 1);
    const int t = i & (BA_THREADS - 1), ub = i >> 7, u = ub >> BA_UB_LOG, b = ub & (BA_UB - 1);
    const int NP = BA_UB * nwin * halves;
    uint32_t* out = dig + (((size_t)item * BA_UNITS + u) * NP + b) * BA_THREADS + t;
    uint32_t k2[4] = {0, 0, 0, 0};
    if (tab.glv) {
        uint32_t k1[4];
        glv_split(k1, k2, s);
#pragma unroll
        for (int k = 0; k < 4; k++) { s[k] = k1[k]; s[4 + k] = 0; }
    }
#pragma unroll 1
    for (int h = 0; h < halves; h++) {
        if (h == 1) {
#pragma unroll
            for (int k = 0; k < 4; k++) { s[k] = k2[k]; s[4 + k] = 0; }
        }
        {   // s += off  (s < r < 2^255 and off < 2^255, or s < N < 2^128 and off < 2^128: no carry out of the eight limbs)
            uint32_t cy = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) { uint64_t w = (uint64_t)s[k] + off.l[k] + cy; s[k] = (uint32_t)w; cy = (uint32_t)(w >> 32); }
        }
#pragma unroll 1
        for (int j = 0; j < nwin; j++) {
            uint32_t v = s[0] & mask;
#pragma unroll
            for (int k = 0; k < 7; k++) s[k] = __funnelshift_r(s[k], s[k + 1], c);
            s[7] >>= c;
            uint32_t e;
            if (j == nwin - 1) {
                v = s[0] << c | v;       // the top window keeps everything that is left (it is unsigned)
                e = v ? (((uint32_t)j << (c - 1)) + v - 1) : BA_ZERO;
            } else {
                int d = (int)v - (int)D;
                uint32_t mag = d < 0 ? (uint32_t)(-d) : (uint32_t)d;
                e = d ? ((((uint32_t)j << (c - 1)) + mag - 1) | (d < 0 ? 0x80000000u : 0u)) : BA_ZERO;
            }
            out[(size_t)(h * nwin + j) * BA_UB * BA_THREADS] = e;
        }
    }
}

struct BaArgs {
    MsmTable tab;
    const uint32_t* dig;
    uint4* scr_pts;          // per CTA slot: K1 points, [(slot * K1 + s) * 6 + q] * BA_THREADS + t
    uint4* scr_pre;          // per CTA slot: K1 prefix products, [(slot * K1 + k) * 3 + q] * BA_THREADS + t
    G1* partials;            // one XYZZ sum per item
    const int32_t* status;
    unsigned int* counter;   // work counter (zeroed before the launch)
    int n;
};

// a reference to one operand of a round: table entry (stride 1) or scratch slot (stride BA_THREADS)
struct BaRef { const uint4* p; int stride; bool neg, zero; };
// what a thread needs to find the operands of its current unit
struct BaCtx { const uint4* tab; const uint32_t* dg; uint4* pts; int base0; };

// KB200_TAB_LD (A/B switch, default off): other ways of loading a table entry (stride 1) -- 1: ld.global.nc.L1::no_allocate,
// 2: ld.global.cs, 3: one 256-bit + one 128-bit load per coordinate.  All three measured slower than the plain loads: L1 is
// what merges the six 16-byte loads of an entry into three sector requests.
#if defined(KB200_TAB_LD) && KB200_TAB_LD > 0 && KB200_TAB_LD < 3
__device__ __forceinline__ uint4 ba_ld_tab(const uint4* p) {
    uint4 v;
#if KB200_TAB_LD == 1
    asm("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
#else
    asm("ld.global.cs.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
#endif
    return v;
}
#endif
__device__ __forceinline__ void ba_ld(Fp& r, const uint4* p, int stride) {
#if defined(KB200_TAB_LD) && KB200_TAB_LD == 3
    // 3: a coordinate (48 bytes) of a table entry as one 256-bit and one 128-bit load (x starts a 32-byte sector, y ends one)
    uint4 a, b, c;
    if (stride == 1) {
        if (((size_t)p & 31) == 0) {
            asm("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(a.x), "=r"(a.y), "=r"(a.z), "=r"(a.w), "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w) : "l"(p));
            c = p[2];
        } else {
            a = p[0];
            asm("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(b.x), "=r"(b.y), "=r"(b.z), "=r"(b.w), "=r"(c.x), "=r"(c.y), "=r"(c.z), "=r"(c.w) : "l"(p + 1));
        }
    } else { a = p[0]; b = p[stride]; c = p[2 * stride]; }
#elif defined(KB200_TAB_LD) && KB200_TAB_LD > 0
    uint4 a, b, c;
    if (stride == 1) { a = ba_ld_tab(p); b = ba_ld_tab(p + 1); c = ba_ld_tab(p + 2); }
    else { a = p[0]; b = p[stride]; c = p[2 * stride]; }
#else
    uint4 a = p[0], b = p[stride], c = p[2 * stride];
#endif
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w; r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    r.l[8] = c.x; r.l[9] = c.y; r.l[10] = c.z; r.l[11] = c.w;
}
__device__ __forceinline__ void ba_st(uint4* p, const Fp& r) {
    p[0] = make_uint4(r.l[0], r.l[1], r.l[2], r.l[3]);
    p[BA_THREADS] = make_uint4(r.l[4], r.l[5], r.l[6], r.l[7]);
    p[2 * BA_THREADS] = make_uint4(r.l[8], r.l[9], r.l[10], r.l[11]);
}
__device__ __forceinline__ uint32_t ba_digit(const BaCtx& c, int idx) { return c.dg[(size_t)idx * BA_THREADS]; }
// operand idx of a round: round 0 reads the table through the digit d, later rounds read scratch slot idx
__device__ __forceinline__ BaRef ba_ref(const BaCtx& c, int round, int idx, uint32_t d) {
    BaRef r;
    if (round == 0) {
        r.zero = d == BA_ZERO; r.neg = (d >> 31) != 0;
        size_t e = (size_t)(d & 0x7fffffffu) * KB_N + (size_t)(c.base0 + (idx & (BA_UB - 1)) * BA_THREADS);
        r.p = c.tab + (r.zero ? 0 : e * 6); r.stride = 1;
    } else {
        r.zero = false; r.neg = false; r.p = c.pts + (size_t)idx * 6 * BA_THREADS; r.stride = BA_THREADS;
    }
    return r;
}
__device__ __forceinline__ void ba_ldx(Fp& x, const BaRef& r) { if (r.zero) fp_zero(x); else ba_ld(x, r.p, r.stride); }
__device__ __forceinline__ void ba_ldy(Fp& y, const BaRef& r) {
    if (r.zero) { fp_zero(y); return; }
    Fp t; ba_ld(t, r.p + 3 * r.stride, r.stride);
    fp_cneg(y, t, r.neg);
}
// pull a table entry (96 bytes: at most two 128-byte lines) into L2 ahead of its use
__device__ __forceinline__ void ba_prefetch(const BaRef& r) {
    if (r.zero || r.stride != 1) return;
    asm volatile("prefetch.global.L2 [%0];" ::"l"(r.p));
    asm volatile("prefetch.global.L2 [%0];" ::"l"(r.p + 5));
}

// Pair k of a round adds operands i1 = ba_i1(k, Kh) and i1 + Kh and leaves the sum in slot k.  Kh = pairs per GLV half: a
// plain table has one half (i1 = k); with two halves the operands of half h sit at [2*h*Kh, 2*(h+1)*Kh) and its sums land in
// [h*Kh, (h+1)*Kh), so the halves never mix and every round halves both.  The sums are written in ascending k: slot k is
// never an operand of a later pair.
struct BaTabYes {}; struct BaTabNo {};     // compile-time operand source of a round (BA_SPLIT_TAB); a plain bool keeps it a run-time value
__device__ __forceinline__ constexpr bool ba_tab_value(BaTabYes) { return true; }
__device__ __forceinline__ constexpr bool ba_tab_value(BaTabNo) { return false; }
__device__ __forceinline__ constexpr bool ba_tab_value(bool b) { return b; }
__device__ __forceinline__ int ba_i1(int k, int Kh) { return k + (k >= Kh ? Kh : 0); }

// Exceptional pairs (an operand at infinity, or x1 == x2) leave the hot loops through these two out-of-line functions;
// they re-read their operands themselves so that nothing in the hot loops has its address taken.
// Denominator: 2*y1 for a doubling, else 1 (nothing to invert).
static __device__ __noinline__ Fp ba_rare_den(BaCtx c, int round, int k, int Kh) {
    const int i1 = ba_i1(k, Kh), i2 = i1 + Kh;
    BaRef r1 = ba_ref(c, round, i1, round == 0 ? ba_digit(c, i1) : 0u), r2 = ba_ref(c, round, i2, round == 0 ? ba_digit(c, i2) : 0u);
    Fp den, x1, x2;
    fp_set(den, c_k.fp_one);
    ba_ldx(x1, r1); ba_ldx(x2, r2);
    if (fp_is_zero(x1) || fp_is_zero(x2)) return den;
    Fp y1, y2; ba_ldy(y1, r1); ba_ldy(y2, r2);
    if (fp_eq(x1, x2) && fp_eq(y1, y2)) fp_dbl(den, y1);
    return den;
}
// Sum of an exceptional pair, written to scratch slot k; returns the running inverse advanced exactly as the forward pass
// advanced the running product.
static __device__ __noinline__ Fp ba_rare_add(BaCtx c, int round, int k, int Kh, Fp inv, Fp pk) {
    const int o1 = ba_i1(k, Kh), o2 = o1 + Kh;
    BaRef r1 = ba_ref(c, round, o1, round == 0 ? ba_digit(c, o1) : 0u), r2 = ba_ref(c, round, o2, round == 0 ? ba_digit(c, o2) : 0u);
    Fp x1, y1, x2, y2, x3, y3;
    ba_ldx(x1, r1); ba_ldy(y1, r1); ba_ldx(x2, r2); ba_ldy(y2, r2);
    const bool i1 = fp_is_zero(x1), i2 = fp_is_zero(x2);
    if (i1 && i2) { fp_zero(x3); fp_zero(y3); }
    else if (i1) { x3 = x2; y3 = y2; }
    else if (i2) { x3 = x1; y3 = y1; }
    else if (!fp_eq(y1, y2)) { fp_zero(x3); fp_zero(y3); }            // P + (-P)
    else {
        Fp den, dinv, lam, t, u;
        fp_dbl(den, y1);
        BA_MUL(dinv, inv, pk); BA_MUL(inv, inv, den);
        BA_SQR(t, x1); fp_dbl(u, t); fp_add(t, t, u);                   // 3 x1^2
        BA_MUL(lam, t, dinv);
        BA_SQR(t, lam); fp_sub(t, t, x1); fp_sub(t, t, x1);
        fp_sub(u, x1, t); BA_MUL(u, lam, u); fp_sub(y3, u, y1);
        x3 = t;
    }
    uint4* o = c.pts + (size_t)k * 6 * BA_THREADS;
    ba_st(o, x3); ba_st(o + 3 * BA_THREADS, y3);
    return inv;
}


// r = neg ? p - a : a.  For a != 0 this is -a in one borrow chain (fp_cneg goes through 0 - a + p: two chains); a == 0 only occurs
// for the (0, 0) infinity sentinel, whose pair leaves the hot loop through the rare path before y is used (no finite point of the
// curve has y == 0: its group order is odd).
__device__ __forceinline__ void ba_cneg_nz(Fp& r, const Fp& a, bool neg) {
    uint32_t t[12];
    asm volatile(
        "sub.cc.u32 %0, 0xffffaaab, %12;\n\t"
        "subc.cc.u32 %1, 0xb9feffff, %13;\n\t"
        "subc.cc.u32 %2, 0xb153ffff, %14;\n\t"
        "subc.cc.u32 %3, 0x1eabfffe, %15;\n\t"
        "subc.cc.u32 %4, 0xf6b0f624, %16;\n\t"
        "subc.cc.u32 %5, 0x6730d2a0, %17;\n\t"
        "subc.cc.u32 %6, 0xf38512bf, %18;\n\t"
        "subc.cc.u32 %7, 0x64774b84, %19;\n\t"
        "subc.cc.u32 %8, 0x434bacd7, %20;\n\t"
        "subc.cc.u32 %9, 0x4b1ba7b6, %21;\n\t"
        "subc.cc.u32 %10, 0x397fe69a, %22;\n\t"
        "subc.u32 %11, 0x1a0111ea, %23;\n\t"
        : "=&r"(t[0]), "=&r"(t[1]), "=&r"(t[2]), "=&r"(t[3]), "=&r"(t[4]), "=&r"(t[5]), "=&r"(t[6]), "=&r"(t[7]), "=&r"(t[8]), "=&r"(t[9]), "=&r"(t[10]), "=&r"(t[11])
        : "r"(a.l[0]), "r"(a.l[1]), "r"(a.l[2]), "r"(a.l[3]), "r"(a.l[4]), "r"(a.l[5]), "r"(a.l[6]), "r"(a.l[7]), "r"(a.l[8]), "r"(a.l[9]), "r"(a.l[10]), "r"(a.l[11]));
#pragma unroll
    for (int i = 0; i < 12; i++) r.l[i] = neg ? t[i] : a.l[i];
}

#if BA_STAGE
extern __shared__ uint4 ba_smem[];      // [15][BA_THREADS]: column `tid` belongs to thread tid alone (no barriers needed)
// 16 bytes global -> shared, asynchronously; bytes == 0 writes zeros without touching the source (an operand at infinity)
__device__ __forceinline__ void ba_cp16(uint4* sdst, const uint4* gsrc, int bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" ::"r"((uint32_t)__cvta_generic_to_shared(sdst)), "l"(gsrc), "r"(bytes) : "memory");
}
__device__ __forceinline__ void ba_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void ba_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// one coordinate of operand r (q0 = 0: x, 3: y) -> slots [c0, c0 + 3) of this thread's column
__device__ __forceinline__ void ba_stage_fp(uint4* st, int c0, const BaRef& r, int q0) {
    const uint4* p = r.p + (size_t)q0 * r.stride;
    const int bytes = r.zero ? 0 : 16;
#pragma unroll
    for (int q = 0; q < 3; q++) ba_cp16(st + (c0 + q) * BA_THREADS, p + (size_t)q * r.stride, bytes);
}
__device__ __forceinline__ void ba_lds(Fp& r, const uint4* st, int c0) {
    const uint4 a = st[c0 * BA_THREADS], b = st[(c0 + 1) * BA_THREADS], c = st[(c0 + 2) * BA_THREADS];
    r.l[0] = a.x; r.l[1] = a.y; r.l[2] = a.z; r.l[3] = a.w; r.l[4] = b.x; r.l[5] = b.y; r.l[6] = b.z; r.l[7] = b.w;
    r.l[8] = c.x; r.l[9] = c.y; r.l[10] = c.z; r.l[11] = c.w;
}
// slot map of the backward pass: buffer b in {0, 1}: x1 -> 9b, x2 -> 9b + 3, y1 -> 9b + 6;  prefix product -> 18, y2 -> 21;  digits -> 24
#define BA_SLOT_PK 18
#define BA_SLOT_Y2 21
#define BA_SLOT_DIG 24
__device__ __forceinline__ void ba_sts(uint4* st, int c0, const Fp& r) {
    st[c0 * BA_THREADS] = make_uint4(r.l[0], r.l[1], r.l[2], r.l[3]);
    st[(c0 + 1) * BA_THREADS] = make_uint4(r.l[4], r.l[5], r.l[6], r.l[7]);
    st[(c0 + 2) * BA_THREADS] = make_uint4(r.l[8], r.l[9], r.l[10], r.l[11]);
}
// the two digits of pair k (table round only) -> words w0, w0 + 1 of the digit slot
__device__ __forceinline__ void ba_stage_digits(uint4* st, int w0, const BaCtx& C, int k, int Kh) {
    const int i1 = ba_i1(k, Kh);
    uint32_t* dst = reinterpret_cast<uint32_t*>(st + BA_SLOT_DIG * BA_THREADS) + w0;
    const uint32_t* g = C.dg + (size_t)i1 * BA_THREADS;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(g) : "memory");
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + 1)), "l"(g + (size_t)Kh * BA_THREADS) : "memory");
}
// backward pass: operands of pair k -> buffer b and the single slots (one commit group, together with whatever digits were
// requested before); returns the sign bits of the two operands
__device__ __forceinline__ uint32_t ba_stage_bwd(uint4* st, int b, const BaCtx& C, const uint4* pre, int round, int k, int Kh, uint32_t d1, uint32_t d2) {
    const int i1 = ba_i1(k, Kh);
    const BaRef r1 = ba_ref(C, round, i1, d1), r2 = ba_ref(C, round, i1 + Kh, d2);
    ba_stage_fp(st, 9 * b, r1, 0); ba_stage_fp(st, 9 * b + 3, r2, 0);
    const uint4* pp = pre + (size_t)k * 3 * BA_THREADS;
#pragma unroll
    for (int q = 0; q < 3; q++) ba_cp16(st + (BA_SLOT_PK + q) * BA_THREADS, pp + (size_t)q * BA_THREADS, 16);
    ba_stage_fp(st, 9 * b + 6, r1, 3); ba_stage_fp(st, BA_SLOT_Y2, r2, 3);
    ba_commit();
    return (r1.neg ? 1u : 0u) | (r2.neg ? 2u : 0u);
}
// forward pass: x1, x2 of pair k -> slots [c0, c0 + 6)
__device__ __forceinline__ void ba_stage_fwd(uint4* st, int c0, const BaCtx& C, int round, int k, int Kh, uint32_t d1, uint32_t d2) {
    const int i1 = ba_i1(k, Kh);
    const BaRef r1 = ba_ref(C, round, i1, d1), r2 = ba_ref(C, round, i1 + Kh, d2);
    ba_stage_fp(st, c0, r1, 0); ba_stage_fp(st, c0 + 3, r2, 0);
    ba_commit();
}
#endif

#ifndef BA_MINB
#define BA_MINB 4          // CTAs per SM the register budget is cut for (128 registers)
#endif
__global__ void __launch_bounds__(BA_THREADS, BA_MINB) k_msm_ba(BaArgs A) {
    static_assert(BA_UNITS == 1, "the GLV map below is applied to the accumulator of one unit");
#if BA_STAGE
    G1* const sm = reinterpret_cast<G1*>(ba_smem);      // the tree reduction reuses the staging bytes (barriers on both sides)
    static_assert(sizeof(G1) * (BA_THREADS / 2) <= BA_SMEM_BYTES, "reduction buffer");
#else
    __shared__ G1 sm[BA_THREADS / 2];
#endif
    __shared__ unsigned int s_item;
    const int tid = threadIdx.x;
    const int nwin = A.tab.nwin, halves = A.tab.glv + 1;
    const int NP = BA_UB * nwin * halves, K1 = NP / 2;
    uint4* const pts = A.scr_pts + (size_t)blockIdx.x * K1 * 6 * BA_THREADS + tid;
    uint4* const pre = A.scr_pre + (size_t)blockIdx.x * K1 * 3 * BA_THREADS + tid;
    for (;;) {
        if (tid == 0) s_item = atomicAdd(A.counter, 1u);
        __syncthreads();
        const unsigned int item = s_item;
        __syncthreads();
        if (item >= (unsigned int)A.n) break;
        G1 acc; g1_set_inf(acc);
        if (A.status == nullptr || A.status[item] == 0) {
            BaCtx C;
            C.tab = A.tab.data; C.dg = A.dig + (size_t)item * NP * BA_THREADS + tid; C.pts = pts;
            C.base0 = tid;
#pragma unroll 1
            for (int round = 0, Kh = (BA_UB / 2) * nwin; round < BA_UB_LOG; round++, Kh >>= 1) {
                const bool tab = round == 0;
                const int K = Kh * halves;
#if BA_STAGE
                // The body of a round, instantiated twice when BA_SPLIT_TAB is set: round 0 reads its operands from the window table
                // (96 contiguous bytes per point, digits decide sign and row), later rounds from the scratch slots (rows 2 KiB apart);
                // with the source known at compile time the sixteen request addresses of a pair become one base + immediates.
                auto stage_round = [&](auto tabc) {
                const bool tab = ba_tab_value(tabc);
                const int round = tab ? 0 : 1;      // (shadows the loop variable) only "round == 0" matters to the helpers
                uint4* const st = ba_smem + tid;
                // ---- forward (k descending): prefix products of the denominators.  The x coordinates of pairs k-1 and k-2 are on
                // their way into the two halves of the staging column while pair k is multiplied in.
                Fp run; fp_set(run, c_k.fp_one);
                {
                    uint32_t dn1 = 0, dn2 = 0;          // digits of the pair that is staged next
                    for (int q = 0; q < 2 && q < K; q++) {
                        const int kk = K - 1 - q, i1 = ba_i1(kk, Kh);
                        ba_stage_fwd(st, 6 * q, C, round, kk, Kh, tab ? ba_digit(C, i1) : 0u, tab ? ba_digit(C, i1 + Kh) : 0u);
                    }
                    if (tab && K > 2) { const int i1 = ba_i1(K - 3, Kh); dn1 = ba_digit(C, i1); dn2 = ba_digit(C, i1 + Kh); }
#pragma unroll 1
                    for (int k = K - 1; k >= 0; k--) {
                        const int c0 = ((K - 1 - k) & 1) * 6;
                        if (k > 0) ba_wait<1>(); else ba_wait<0>();
                        Fp x1, x2, d;
                        ba_lds(x1, st, c0); ba_lds(x2, st, c0 + 3);
                        ba_st(pre + (size_t)k * 3 * BA_THREADS, run);
                        fp_sub(d, x2, x1);
                        if (fp_is_zero(x1) || fp_is_zero(x2) || fp_is_zero(d)) d = ba_rare_den(C, round, k, Kh);
                        if (k >= 2) {
                            ba_stage_fwd(st, c0, C, round, k - 2, Kh, dn1, dn2);
                            if (tab && k >= 3) { const int i1 = ba_i1(k - 3, Kh); dn1 = ba_digit(C, i1); dn2 = ba_digit(C, i1 + Kh); }
                        }
#if defined(BA_FWD_CALL) && BA_FWD_CALL
                        BA_MUL(run, run, d);
#else
                        fp_mul(run, run, d);
#endif
                    }
                }
                Fp inv = ba_inv(run);
                // ---- backward (k ascending): the additions.  Pair k+1 is requested as soon as the single slots of pair k are consumed
                // (after the first product); x1, x2 and y1 of pair k stay in their buffer and are read again where the formulas use them.
                {
                    uint32_t negs_next;
                    uint32_t* const sdig = reinterpret_cast<uint32_t*>(st + BA_SLOT_DIG * BA_THREADS);
                    {
                        const int i1 = ba_i1(0, Kh);
                        if (tab && K > 1) ba_stage_digits(st, 2, C, 1, Kh);       // digits of pair 1 -> words 2, 3 (same group as pair 0)
                        negs_next = ba_stage_bwd(st, 0, C, pre, round, 0, Kh, tab ? ba_digit(C, i1) : 0u, tab ? ba_digit(C, i1 + Kh) : 0u);
                    }
#pragma unroll 1
                    for (int k = 0; k < K; k++) {
                        const int b = k & 1;
                        ba_wait<0>();
                        const uint32_t negs = negs_next;
                        // Order of work: what is live across a product call must fit the registers the product leaves alone, so
                        // values are read from the staging column right where the formulas use them.
                        Fp d, pk;
                        bool rare;
                        {
                            Fp x1, x2;
                            ba_lds(x1, st, 9 * b); ba_lds(x2, st, 9 * b + 3);
                            fp_sub(d, x2, x1);
                            rare = fp_is_zero(x1) || fp_is_zero(x2) || fp_is_zero(d);
                            ba_lds(pk, st, BA_SLOT_PK);
                        }
                        const uint32_t dn1 = sdig[2 * (b ^ 1)], dn2 = sdig[2 * (b ^ 1) + 1];     // digits of pair k + 1 (requested one pair ago)
                        if (rare) {
                            inv = ba_rare_add(C, round, k, Kh, inv, pk);
                            if (k + 1 < K) {
                                if (tab && k + 2 < K) ba_stage_digits(st, 2 * b, C, k + 2, Kh);
                                negs_next = ba_stage_bwd(st, b ^ 1, C, pre, round, k + 1, Kh, dn1, dn2);
                            }
                        } else {
                            Fp dinv, t, lam, x3, y3;
                            BA_MULX(0, dinv, inv, pk);
                            BA_MULX(1, inv, inv, d);
                            {
                                Fp y1, y2;
                                { Fp r; ba_lds(r, st, 9 * b + 6); ba_cneg_nz(y1, r, (negs & 1u) != 0); }
                                { Fp r; ba_lds(r, st, BA_SLOT_Y2); ba_cneg_nz(y2, r, (negs & 2u) != 0); }
                                ba_sts(st, 9 * b + 6, y1);                               // the signed y1 is needed again for y3
                                fp_sub(t, y2, y1);
                            }
                            // the single slots (prefix product, y2) and the digit words are consumed: request pair k + 1 into the other
                            // buffer; three products remain to cover its way from HBM
                            if (k + 1 < K) {
                                if (tab && k + 2 < K) ba_stage_digits(st, 2 * b, C, k + 2, Kh);
                                negs_next = ba_stage_bwd(st, b ^ 1, C, pre, round, k + 1, Kh, dn1, dn2);
                            }
                            BA_MULX(2, lam, t, dinv);
                            BA_MULX(3, x3, lam, lam);
                            { Fp x1, x2; ba_lds(x1, st, 9 * b); ba_lds(x2, st, 9 * b + 3); fp_sub(x3, x3, x1); fp_sub(x3, x3, x2); fp_sub(t, x1, x3); }
                            BA_MULX(4, t, lam, t);
                            { Fp y1; ba_lds(y1, st, 9 * b + 6); fp_sub(y3, t, y1); }
                            uint4* o = pts + (size_t)k * 6 * BA_THREADS;
                            ba_st(o, x3); ba_st(o + 3 * BA_THREADS, y3);
                        }
                    }
                }
                };
#if BA_SPLIT_TAB
                if (tab) stage_round(BaTabYes{}); else stage_round(BaTabNo{});
#else
                stage_round(tab);
#endif
#else
                // ---- forward (k descending): prefix products of the denominators.  Operands of the next pair are loaded
                // before the product of this pair is issued, the digits of the pair after that before that.
                Fp run; fp_set(run, c_k.fp_one);
                {
                    int i1 = ba_i1(K - 1, Kh);
                    uint32_t d1 = tab ? ba_digit(C, i1) : 0u, d2 = tab ? ba_digit(C, i1 + Kh) : 0u;
                    BaRef r1 = ba_ref(C, round, i1, d1), r2 = ba_ref(C, round, i1 + Kh, d2);
                    Fp nx1, nx2; ba_ldx(nx1, r1); ba_ldx(nx2, r2);
                    if (tab && K > 1) { i1 = ba_i1(K - 2, Kh); d1 = ba_digit(C, i1); d2 = ba_digit(C, i1 + Kh); }
#pragma unroll 1
                    for (int k = K - 1; k >= 0; k--) {
                        Fp x1 = nx1, x2 = nx2, d;
                        if (k > 0) {
                            i1 = ba_i1(k - 1, Kh);
                            r1 = ba_ref(C, round, i1, d1); r2 = ba_ref(C, round, i1 + Kh, d2);
                            ba_ldx(nx1, r1); ba_ldx(nx2, r2);
                            if (tab && k > 1) { i1 = ba_i1(k - 2, Kh); d1 = ba_digit(C, i1); d2 = ba_digit(C, i1 + Kh); }
                        }
                        ba_st(pre + (size_t)k * 3 * BA_THREADS, run);
                        fp_sub(d, x2, x1);
                        if (fp_is_zero(x1) || fp_is_zero(x2) || fp_is_zero(d)) d = ba_rare_den(C, round, k, Kh);
                        fp_mul(run, run, d);      // inlined: the only product of this loop, so the loads of the next pair stay in flight across it
                    }
                }
                Fp inv = ba_inv(run);
                // ---- backward (k ascending): the additions.  The table entries of pair k+1 are pulled into L2 while pair k is added.
                {
                    uint32_t d1 = tab ? ba_digit(C, 0) : 0u, d2 = tab ? ba_digit(C, Kh) : 0u;
#pragma unroll 1
                    for (int k = 0; k < K; k++) {
                        const int i1 = ba_i1(k, Kh);
                        BaRef r1 = ba_ref(C, round, i1, d1), r2 = ba_ref(C, round, i1 + Kh, d2);
                        Fp x1, y1, x2, y2, pk, d;
                        ba_ldx(x1, r1); ba_ldx(x2, r2);
                        ba_ld(pk, pre + (size_t)k * 3 * BA_THREADS, BA_THREADS);
                        ba_ldy(y1, r1); ba_ldy(y2, r2);
                        if (tab && k + 1 < K) {
                            const int j1 = ba_i1(k + 1, Kh);
                            d1 = ba_digit(C, j1); d2 = ba_digit(C, j1 + Kh);
                            ba_prefetch(ba_ref(C, 0, j1, d1)); ba_prefetch(ba_ref(C, 0, j1 + Kh, d2));
                        }
                        fp_sub(d, x2, x1);
                        if (fp_is_zero(x1) || fp_is_zero(x2) || fp_is_zero(d)) {
                            inv = ba_rare_add(C, round, k, Kh, inv, pk);
                        } else {
                            Fp dinv, lam, t, x3, y3;
                            BA_MUL(dinv, inv, pk); BA_MUL(inv, inv, d);
                            fp_sub(t, y2, y1); BA_MUL(lam, t, dinv);
                            BA_SQR(x3, lam); fp_sub(x3, x3, x1); fp_sub(x3, x3, x2);
                            fp_sub(t, x1, x3); BA_MUL(t, lam, t); fp_sub(y3, t, y1);
                            uint4* o = pts + (size_t)k * 6 * BA_THREADS;
                            ba_st(o, x3); ba_st(o + 3 * BA_THREADS, y3);
                        }
                    }
                }
#endif
            }
            // the nwin sums of each half -> XYZZ accumulator; the k2 half first, mapped through Q = (beta x, -y) as one sum
#pragma unroll 1
            for (int h = halves - 1; h >= 0; h--) {
#pragma unroll 1
                for (int k = h * nwin; k < (h + 1) * nwin; k++) {
                    G1Affine pt;
                    ba_ld(pt.x, pts + (size_t)k * 6 * BA_THREADS, BA_THREADS);
                    ba_ld(pt.y, pts + (size_t)k * 6 * BA_THREADS + 3 * BA_THREADS, BA_THREADS);
                    ba_madd(acc, pt.x, pt.y);
                }
                if (h == 1) { Fp b; fp_set(b, c_k.fp_beta); BA_MUL(acc.x, acc.x, b); fp_neg(acc.y, acc.y); }
            }
        }
        // CTA tree reduction through shared memory
#if BA_STAGE
        __syncthreads();        // every warp is out of its loops: the staging columns are free
#endif
#pragma unroll 1
        for (int stride = BA_THREADS / 2; stride >= 1; stride >>= 1) {
            if (tid >= stride && tid < 2 * stride) sm[tid - stride] = acc;
            __syncthreads();
            if (tid < stride) { G1 other = sm[tid]; ba_add(acc, other); }
            __syncthreads();
        }
        if (tid == 0) A.partials[item] = acc;
    }
}

}  // namespace kb200
