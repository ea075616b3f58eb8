"""Host-side check of kateth_b200/csrc/glv.cuh (scalar split along the G1 endomorphism): compiled with g++ and compared
with Python integers.  No GPU needed."""
import ctypes as C
import os, random, subprocess, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R = 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001
N = 0xd201000000010000 ** 2
SRC = r'''
#include "glv.cuh"
extern "C" void glv_split_many(const uint32_t* in, uint32_t* out, int n) {
    for (int t = 0; t < n; t++) {
        uint32_t k[8], k1[4], k2[4];
        for (int i = 0; i < 8; i++) k[i] = in[8 * t + i];
        kb200::glv_split(k1, k2, k);
        for (int i = 0; i < 4; i++) { out[8 * t + i] = k1[i]; out[8 * t + 4 + i] = k2[i]; }
    }
}
'''


def test_glv_split_matches_divmod():
    assert R == N * N - N + 1
    rng = random.Random(0x61)
    ks = [0, 1, N - 1, N, N + 1, R - 1, R - 2, 2 * N - 1, 2 * N, (N - 1) * N, (N - 1) * N + N - 1, R >> 1]
    ks += [rng.randrange(R) for _ in range(20000)]
    ks += [min(R - 1, q * N + d) for q in (1, 2, N - 2, N - 1) for d in (0, 1, N - 1)]
    n = len(ks)
    buf = (C.c_uint32 * (8 * n))(*[(k >> (32 * i)) & 0xffffffff for k in ks for i in range(8)])
    out = (C.c_uint32 * (8 * n))()
    with tempfile.TemporaryDirectory() as tmp:
        src = os.path.join(tmp, "glv_host.cpp"); open(src, "w").write(SRC)
        so = os.path.join(tmp, "libglv_host.so")
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I", os.path.join(ROOT, "kateth_b200", "csrc"), src, "-o", so])
        C.CDLL(so).glv_split_many(buf, out, n)
    for t, k in enumerate(ks):
        k1 = sum(out[8 * t + i] << (32 * i) for i in range(4)); k2 = sum(out[8 * t + 4 + i] << (32 * i) for i in range(4))
        assert (k2, k1) == divmod(k, N), hex(k)
