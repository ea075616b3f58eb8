#!/usr/bin/env python3
"""Latency of small batch checks under the two forms of the sums (kb200_set_rlc_mode 1 = ladders, 2 = buckets): verify_proof (n = 1) and
verify_blob_proof_batch at n = 1, 8, 64, 256."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import kateth_b200 as K
import bench
s = K.Setup.load_json(bench.SETUP_JSON, window_bits=int(os.environ.get("KB200_WINDOW_BITS", "13")))
n = 256
cat = bytes(bench.synth_blobs(n, 5))
comm, _ = s.blob_to_commitment_many(cat, n)
pr, _ = s.blob_proof_many(cat, comm, n)
zs = cat[:32]
p1, y1, _ = s.proof_many(cat[:131072], zs, 1)
for mode in (1, 2):
    s.set_rlc_mode(mode)
    out = []
    for _ in range(3): s.verify_proof(p1, comm[:48], zs, y1)
    t0 = time.perf_counter()
    for _ in range(20): ok = s.verify_proof(p1, comm[:48], zs, y1)
    out.append("verify_proof %.2f ms" % ((time.perf_counter() - t0) / 20 * 1e3))
    for m in (1, 8, 64, 256):
        for _ in range(2): s.verify_blob_proof_batch_concat(cat[:m * 131072], comm[:48 * m], pr[:48 * m], m)
        t0 = time.perf_counter()
        for _ in range(10): ok = s.verify_blob_proof_batch_concat(cat[:m * 131072], comm[:48 * m], pr[:48 * m], m) and ok
        out.append("batch %d: %.2f ms" % (m, (time.perf_counter() - t0) / 10 * 1e3))
    print("rlc_mode", mode, ok, " | ".join(out), flush=True)
s.close()
