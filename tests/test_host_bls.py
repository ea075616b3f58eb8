"""CPU test of the engine's host pairing module (kateth_b200/csrc/host_bls.cpp, product code): built with g++ alone and run here.
The Granger-Scott cyclotomic squaring of the final exponentiation must equal the generic Fp12 squaring on a cyclotomic element, and the
two-pairing check must accept e(-G1, G2) e(G1, G2) and reject e(-G1, G2) e(2 G1, G2).  (The verdict vectors pin it on the GPU box.)"""
import os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_host_pairing_module(tmp_path):
    exe = str(tmp_path / "host_pairing_check")
    src = [os.path.join(ROOT, "tools", "host_pairing_check.cpp"), os.path.join(ROOT, "kateth_b200", "csrc", "host_bls.cpp")]
    subprocess.run(["g++", "-O2", "-std=c++17", "-DKB200_HOST_SELFTEST", "-I", os.path.join(ROOT, "kateth_b200", "csrc")] + src + ["-o", exe], check=True)
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    got = dict(line.split() for line in out.strip().splitlines())
    assert got["cyclotomic_equals_generic"] == "1" and got["pairing_same"] == "1" and got["pairing_different"] == "0", out
