#!/usr/bin/env python3
"""Emit kateth_b200/csrc/mont_gen.cuh: fully unrolled Montgomery products for BLS12-381 Fp (12x32-bit limbs,
R = 2^384) and Fr (8x32-bit limbs, R = 2^256) as inline PTX carry chains.

Shape of the algorithm (word-serial CIOS, one 32-bit word of b per round):
  T = E + (O << 32): two accumulator arrays.  E collects a_even * b_j (and p_even * m), O collects a_odd * b_j
  (and p_odd * m), so every `mad.lo.cc / madc.hi.cc` pair lands on an *aligned* register pair and ptxas fuses it
  into ONE `IMAD.WIDE.U32(.X)` (32x32+64 with predicate carry).  After the reduction step word 0 of E is zero;
  instead of shifting, the arrays swap roles for the next round, the shift of the other array by two words is
  folded into that round's first chain (dest = src+2), and the one misaligned word is added at the head of the
  chain so its carry enters the chain.  Cost: 2N^2 + N IMAD per product (300 for Fp, 136 for Fr) plus O(N) adds
  that issue on the ALU pipe.  The modulus limbs are immediates, so they occupy no registers.
Inputs and outputs are fully reduced (< modulus).
"""
import os

P = 0x1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffffb9feffffffffaaab
R = 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def limbs(v, n):
    return [(v >> (32 * i)) & 0xffffffff for i in range(n)]


class Asm:
    """Builds one asm volatile statement with named operands."""

    def __init__(self):
        self.lines, self.outs, self.ins, self.idx = [], [], [], {}

    def ref(self, expr, write=False):
        if expr not in self.idx:
            self.idx[expr] = None
            (self.outs if write else self.ins).append(expr)
        elif write and expr in self.ins:
            self.ins.remove(expr); self.outs.append(expr)
        return "@" + expr + "@"

    def op(self, text):
        self.lines.append(text)

    def render(self, indent="    "):
        order = self.outs + self.ins
        num = {e: i for i, e in enumerate(order)}
        body = []
        for l in self.lines:
            for e in order:
                l = l.replace("@" + e + "@", "%" + str(num[e]))
            body.append(l)
        s = indent + "asm volatile(\n"
        for l in body:
            s += indent + '    "%s;\\n\\t"\n' % l
        s += indent + "    : " + ", ".join('"+r"(%s)' % e for e in self.outs) + "\n"
        s += indent + "    : " + ", ".join('"r"(%s)' % e for e in self.ins) + ");\n"
        return s


def gen_mul(name, n, mod, inv, const_arr=None):
    p = limbs(mod, n)
    def pl(s, i):
        return ("0x%08x" % p[i]) if const_arr is None else s.ref("pm_[%d]" % i)
    out = "__device__ __forceinline__ void %s_mul(uint32_t (&r)[%d], const uint32_t (&a)[%d], const uint32_t (&b)[%d]) {\n" % (name, n, n, n)
    out += "    uint32_t e[%d], o[%d], m;\n" % (n, n)
    if const_arr == "opaque":
        # ptxas strength-reduces products by the special limbs of r (1, 0xffffffff) and then fails to fuse the
        # rest of the chain into IMAD.WIDE; hiding the limbs behind a runtime zero keeps the chains intact.
        out += "    const uint32_t z_ = opaque_zero();\n    const uint32_t pm_[%d] = {%s};\n" % (n, ", ".join("0x%08xu ^ z_" % x for x in p))
    # round 0: plain products
    for i in range(0, n, 2):
        out += "    { uint64_t t = (uint64_t)a[%d] * b[0]; e[%d] = (uint32_t)t; e[%d] = (uint32_t)(t >> 32); }\n" % (i, i, i + 1)
    for i in range(1, n, 2):
        out += "    { uint64_t t = (uint64_t)a[%d] * b[0]; o[%d] = (uint32_t)t; o[%d] = (uint32_t)(t >> 32); }\n" % (i, i - 1, i)
    for j in range(n):
        A, B = ("e", "o") if j % 2 == 0 else ("o", "e")
        if j > 0:
            s = Asm()
            # (a) misaligned word, (b) B chain with the two-word shift folded in
            s.op("add.cc.u32 %s, %s, %s" % (s.ref("%s[0]" % A, True), s.ref("%s[0]" % A, True), s.ref("%s[1]" % B, True)))
            k = 0
            for i in range(1, n, 2):
                src0 = s.ref("%s[%d]" % (B, k + 2), True) if k + 2 < n else "0"
                src1 = s.ref("%s[%d]" % (B, k + 3), True) if k + 3 < n else "0"
                last = (i == n - 1)
                s.op("madc.lo.cc.u32 %s, %s, %s, %s" % (s.ref("%s[%d]" % (B, k), True), s.ref("a[%d]" % i), s.ref("b[%d]" % j), src0))
                s.op("madc.hi%s.u32 %s, %s, %s, %s" % ("" if last else ".cc", s.ref("%s[%d]" % (B, k + 1), True), s.ref("a[%d]" % i), s.ref("b[%d]" % j), src1))
                k += 2
            out += s.render()
            # (c) A chain
            s = Asm()
            for i in range(0, n, 2):
                first = (i == 0)
                s.op("mad%s.lo.cc.u32 %s, %s, %s, %s" % ("" if first else "c", s.ref("%s[%d]" % (A, i), True), s.ref("a[%d]" % i), s.ref("b[%d]" % j), s.ref("%s[%d]" % (A, i), True)))
                s.op("madc.hi.cc.u32 %s, %s, %s, %s" % (s.ref("%s[%d]" % (A, i + 1), True), s.ref("a[%d]" % i), s.ref("b[%d]" % j), s.ref("%s[%d]" % (A, i + 1), True)))
            s.op("addc.u32 %s, %s, 0" % (s.ref("%s[%d]" % (B, n - 1), True), s.ref("%s[%d]" % (B, n - 1), True)))
            out += s.render()
        out += "    m = %s[0] * 0x%08xu;\n" % (A, inv)
        # (e) reduction: A += p_even * m (carry into top of B), B += p_odd * m
        s = Asm()
        for i in range(0, n, 2):
            first = (i == 0)
            s.op("mad%s.lo.cc.u32 %s, %s, %s, %s" % ("" if first else "c", s.ref("%s[%d]" % (A, i), True), s.ref("m"), pl(s, i), s.ref("%s[%d]" % (A, i), True)))
            s.op("madc.hi.cc.u32 %s, %s, %s, %s" % (s.ref("%s[%d]" % (A, i + 1), True), s.ref("m"), pl(s, i), s.ref("%s[%d]" % (A, i + 1), True)))
        s.op("addc.u32 %s, %s, 0" % (s.ref("%s[%d]" % (B, n - 1), True), s.ref("%s[%d]" % (B, n - 1), True)))
        k = 0
        for i in range(1, n, 2):
            first, last = (i == 1), (i == n - 1)
            s.op("mad%s.lo.cc.u32 %s, %s, %s, %s" % ("" if first else "c", s.ref("%s[%d]" % (B, k), True), s.ref("m"), pl(s, i), s.ref("%s[%d]" % (B, k), True)))
            s.op("madc.hi%s.u32 %s, %s, %s, %s" % ("" if last else ".cc", s.ref("%s[%d]" % (B, k + 1), True), s.ref("m"), pl(s, i), s.ref("%s[%d]" % (B, k + 1), True)))
            k += 2
        out += s.render()
    # merge: after the last round (j = n-1, odd) A = o, B = e:  r[k] = o[k+1] + e[k]
    A, B = ("e", "o") if (n - 1) % 2 == 0 else ("o", "e")
    s = Asm()
    for k in range(n):
        opn = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < n - 1 else "addc.u32")
        src = s.ref("%s[%d]" % (A, k + 1), True) if k + 1 < n else "0"
        s.op("%s %s, %s, %s" % (opn, s.ref("%s[%d]" % (B, k), True), s.ref("%s[%d]" % (B, k), True), src))
    out += s.render()
    out += "    %s_reduce_once(r, %s);\n}\n\n" % (name, B)
    return out


def gen_addsub(name, n, mod):
    p = limbs(mod, n)
    o = ""
    # reduce_once: r = (t >= p) ? t - p : t      (t < 2p, no carry word)
    o += "__device__ __forceinline__ void %s_reduce_once(uint32_t (&r)[%d], const uint32_t (&t)[%d]) {\n    uint32_t d[%d], bw;\n" % (name, n, n, n)
    s = Asm()
    for k in range(n):
        opn = "sub.cc.u32" if k == 0 else "subc.cc.u32"
        s.op("%s %s, %s, 0x%08x" % (opn, s.ref("d[%d]" % k, True), s.ref("t[%d]" % k), p[k]))
    s.op("subc.u32 %s, 0, 0" % s.ref("bw", True))
    o += s.render().replace('"+r"', '"=&r"')
    o += "    #pragma unroll\n    for (int i = 0; i < %d; i++) r[i] = bw ? t[i] : d[i];\n}\n\n" % n
    # add
    o += "__device__ __forceinline__ void %s_add(uint32_t (&r)[%d], const uint32_t (&a)[%d], const uint32_t (&b)[%d]) {\n    uint32_t t[%d];\n" % (name, n, n, n, n)
    s = Asm()
    for k in range(n):
        opn = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < n - 1 else "addc.u32")
        s.op("%s %s, %s, %s" % (opn, s.ref("t[%d]" % k, True), s.ref("a[%d]" % k), s.ref("b[%d]" % k)))
    o += s.render().replace('"+r"', '"=&r"')
    o += "    %s_reduce_once(r, t);\n}\n\n" % name
    # sub: t = a - b; if borrow t += p
    o += "__device__ __forceinline__ void %s_sub(uint32_t (&r)[%d], const uint32_t (&a)[%d], const uint32_t (&b)[%d]) {\n    uint32_t t[%d], bw;\n" % (name, n, n, n, n)
    s = Asm()
    for k in range(n):
        opn = "sub.cc.u32" if k == 0 else "subc.cc.u32"
        s.op("%s %s, %s, %s" % (opn, s.ref("t[%d]" % k, True), s.ref("a[%d]" % k), s.ref("b[%d]" % k)))
    s.op("subc.u32 %s, 0, 0" % s.ref("bw", True))
    o += s.render().replace('"+r"', '"=&r"')
    s = Asm()
    for k in range(n):
        opn = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < n - 1 else "addc.u32")
        s.op("%s %s, %s, %s" % (opn, s.ref("t[%d]" % k, True), s.ref("t[%d]" % k, True), s.ref("q[%d]" % k)))
    # two forms of the correction.  KB200_SUB_PAR: u = t + p as a second chain that follows t limb by limb (it does not wait for the final
    # borrow), then a select; default: t += (p & borrow mask), which can only start once the borrow is known.
    s2 = Asm()
    for k in range(n):
        opn = "add.cc.u32" if k == 0 else ("addc.cc.u32" if k < n - 1 else "addc.u32")
        s2.op("%s %s, %s, 0x%08x" % (opn, s2.ref("u[%d]" % k, True), s2.ref("t[%d]" % k), p[k]))
    o += "#if defined(KB200_SUB_PAR) && KB200_SUB_PAR\n    uint32_t u[%d];\n" % n
    o += s2.render().replace('"+r"', '"=&r"')
    o += "    #pragma unroll\n    for (int i = 0; i < %d; i++) r[i] = bw ? u[i] : t[i];\n#else\n" % n
    o += "    uint32_t q[%d] = {%s};\n" % (n, ", ".join("bw & 0x%08xu" % x for x in p))
    o += s.render()
    o += "    #pragma unroll\n    for (int i = 0; i < %d; i++) r[i] = t[i];\n#endif\n}\n\n" % n
    return o


def main():
    out = "// GENERATED by tools/gen_mont.py -- do not edit.\n#pragma once\n#include <stdint.h>\nnamespace kb200 {\n\n__device__ __forceinline__ uint32_t opaque_zero() { uint32_t l; asm volatile(\"mov.u32 %0, %%laneid;\" : \"=r\"(l)); return l >> 5; }\n\n"
    out += gen_addsub("fp", 12, P) + gen_mul("fp", 12, P, (-pow(P, -1, 1 << 32)) % (1 << 32))
    out += "static __device__ __constant__ uint32_t FR_MOD_C[8] = {%s};\n\n" % ", ".join("0x%08xu" % x for x in limbs(R, 8))
    out += gen_addsub("fr", 8, R) + gen_mul("fr", 8, R, (-pow(R, -1, 1 << 32)) % (1 << 32))
    out += "} // namespace kb200\n"
    with open(os.path.join(ROOT, "kateth_b200", "csrc", "mont_gen.cuh"), "w") as f:
        f.write(out)
    print("wrote kateth_b200/csrc/mont_gen.cuh (%d bytes)" % len(out))


if __name__ == "__main__":
    main()
