"""Host-side check of kateth_b200/csrc/fp_bgcd.cuh (the branchless binary-GCD inversion the batch-affine MSM uses):
the header is plain C++, so it is compiled here with g++ and compared with pow(x, -1, p).  No GPU needed."""
import ctypes as C
import os, random, subprocess, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = 0x1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffffb9feffffffffaaab
SRC = r'''
#include "fp_bgcd.cuh"
extern "C" int bgcd_inverse_many(const uint32_t* in, uint32_t* out, int n) {
    int ok = 1;
    for (int k = 0; k < n; k++) {
        uint32_t x[12], y[12];
        for (int i = 0; i < 12; i++) x[i] = in[12 * k + i];
        ok &= kb200::bg_inverse(y, x);
        for (int i = 0; i < 12; i++) out[12 * k + i] = y[i];
    }
    return ok;
}
'''


def _build(tmp):
    src = os.path.join(tmp, "bgcd_host.cpp")
    open(src, "w").write(SRC)
    so = os.path.join(tmp, "libbgcd_host.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I", os.path.join(ROOT, "kateth_b200", "csrc"), src, "-o", so])
    return C.CDLL(so)


def test_bgcd_inverse_matches_pow():
    rng = random.Random(0x4b5a)
    xs = [1, 2, 3, P - 1, P - 2, (P + 1) // 2, 1 << 380, P >> 1, 0x1234, (1 << 64) - 1, 1 << 64, (1 << 62) + 1]
    xs += [rng.randrange(1, P) for _ in range(4000)]
    xs += [rng.randrange(1, 1 << rng.randrange(1, 381)) for _ in range(2000)]
    xs += [P - rng.randrange(1, 1 << rng.randrange(1, 380)) for _ in range(1000)]
    n = len(xs)
    buf = (C.c_uint32 * (12 * n))(*[(x >> (32 * i)) & 0xffffffff for x in xs for i in range(12)])
    out = (C.c_uint32 * (12 * n))()
    with tempfile.TemporaryDirectory() as tmp:
        lib = _build(tmp)
        assert lib.bgcd_inverse_many(buf, out, n) == 1
    for k, x in enumerate(xs):
        y = sum(out[12 * k + i] << (32 * i) for i in range(12))
        assert y == pow(x, -1, P), hex(x)


def test_bgcd_zero_is_flagged():
    buf = (C.c_uint32 * 12)()
    out = (C.c_uint32 * 12)()
    with tempfile.TemporaryDirectory() as tmp:
        lib = _build(tmp)
        assert lib.bgcd_inverse_many(buf, out, 1) == 0      # gcd(0, p) = p: the caller falls back
