// host_bls.hpp -- the HOST side of the engine's BLS12-381 needs (product code; independent of oracle/).
//
// north_star keeps "the final two-pairing check in the host bls module as a non-hot step".  kateth's host bls
// module is blst behind /root/reference/src/bls.rs:554-598 (P2, verify_pairings); neither Rust nor blst exists in
// this image, so this file is that module written from scratch in C++:
//   * Fp (6x64-bit Montgomery), the tower Fp2 = Fp[u]/(u^2+1), Fp6 = Fp2[v]/(v^3-(1+u)), Fp12 = Fp6[w]/(w^2-v)
//   * G2: 96-byte ZCash decompression + subgroup check (P2::decompress at setup load, src/kzg/setup.rs:68-72)
//   * optimal-ate pairing product check  e(-a1, a2) * e(b1, b2) == 1  (bls::verify_pairings, src/bls.rs:572-598)
//     with the G2 arguments FIXED at setup time ([tau]_2 and the generator): their Miller-loop line coefficients
//     are precomputed once, so a check costs 63 Fp12 squarings + 136 sparse products + one final exponentiation
//     (easy part by conjugate/inverse/Frobenius, hard part by the (x-1)^2 (x+p)(x^2+p^2-1)+3 chain).
// Unlike the oracle's textbook pairing (affine, no tower, plain 4314-bit exponent) this one is built for speed;
// the two share no code, and both must agree with the sha-pinned verdicts of the consensus-spec-tests vectors.
#pragma once
#include <stdint.h>
#include <stddef.h>

namespace kb200 { namespace host {

struct Fp { uint64_t l[6]; };
struct Fp2 { Fp c0, c1; };
struct G1A { Fp x, y; bool inf; };     // Montgomery form
struct G2A { Fp2 x, y; bool inf; };

// line coefficients of one Miller-loop step for a fixed Q: l(P) = c0 + (c1 * xP) w^2 + yP w^3  (c1 already negated)
struct Line { Fp2 c0, c1; };
struct G2Prepared { Line lines[68]; int n; bool inf; };

// plain little-endian limbs (12 x u32, as the device emits them) -> Montgomery Fp
void fp_from_plain_u32(Fp& r, const uint32_t* limbs12);

// 96-byte compressed G2 -> affine; returns 0 ok, 3 bad encoding, 4 not on curve, 5 not in subgroup
int g2_decompress(G2A& out, const uint8_t in[96]);
void g2_generator(G2A& out);
void g2_prepare(G2Prepared& out, const G2A& q);

// Frobenius constants of the tower (gamma^k in Fp2, (gamma conj(gamma))^k in Fp, k = 0..5) for the device copy of the pairing
void frobenius_consts(Fp2 frob1[6], Fp frob2[6]);

// e(-a1, q1) * e(b1, q2) == 1 ?
bool verify_pairings(const G1A& a1, const G2Prepared& q1, const G1A& b1, const G2Prepared& q2);

}}  // namespace
