#!/usr/bin/env python3
"""Debug probe: one blob of the degenerate-setup test through the batch-affine MSM, in its own process."""
import json, os, random, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import reconstruct_vectors as RV
import test_gpu_parity as T
import kateth_b200 as K
from oracle import oracle as O
which, wb = int(sys.argv[1]), int(sys.argv[2])
R = RV.R
real = json.load(open(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json")))
gx = int.from_bytes(bytes([RV.G1[0] & 0x1f]) + RV.G1[1:], "big")
gy = 0x08b3f481e3aaa0f1a09e30ed741d8ae4fcf5e095d5d00af600db18cb2c04b3edd03cc744a2888ae40caa232946c5e7e1
mult = [None] + [T._ec_mul(k, (gx, gy)) for k in range(1, 8)]
pts = ["0x" + T._compress(mult[1 + (i * 7 + i // 64) % 7]).hex() for i in range(4096)]
path = "/tmp/degenerate_setup.json"
open(path, "w").write(json.dumps({"g1_lagrange": pts, "g2_monomial": real["g2_monomial"]}))
rng = random.Random(77)
small = [0, 1, 2, 3, 4, R - 1, R - 2, R - 3, 255, 256, 257, R - 256]
blobs = [b"".join(RV.fe(rng.choice(small)) for _ in range(4096)),
         b"".join(RV.fe(1) for _ in range(4096)),
         b"".join(RV.fe(R - 1 if k & 1 else 1) for k in range(4096)),
         b"".join(RV.fe(rng.randrange(R)) for _ in range(4096)),
         b"".join(RV.fe(rng.randrange(1 << 18)) for _ in range(4096)),
         b"".join(RV.fe(0) for _ in range(4096)),
         b"".join(RV.fe(1 if k == 5 else 0) for k in range(4096))]
s = K.Setup.load_json(path, window_bits=wb)
s.set_msm_mode(2)
try:
    c = s.blob_to_commitment(blobs[which])
    o = O.Setup.load_json(path)
    print("blob", which, "wb", wb, "OK" if c == o.blob_to_commitment(blobs[which]) else "MISMATCH", flush=True)
except Exception as e:
    print("blob", which, "wb", wb, "FAIL", str(e)[:80], flush=True)
