#!/usr/bin/env python3
"""First-contact probe on a GPU box: K0 integer-MAC peak, field parity, setup load time, golden commitment,
per-stage timings.  Writes progress line by line so a hang shows where."""
import os, sys, time, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
def log(*a):
    print(time.strftime("%H:%M:%S"), *a, flush=True)
import kateth_b200 as K
import reconstruct_vectors as RV
log("nvidia-smi", subprocess.run(["nvidia-smi", "--query-gpu=name,clocks.sm,clocks.max.sm,memory.used,memory.total", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip())
w, f = K.test_hooks.mac_peak(8192)
log("K0 mad.wide peak T MAC32/s:", round(w, 3), " fp_mul chain T MAC32/s:", round(f, 3))
w, f = K.test_hooks.mac_peak(32768)
log("K0 (longer) mad.wide:", round(w, 3), " fp_mul chain:", round(f, 3))
vb = RV.valid_blobs()
for wb in (9, 13):
    t = time.time()
    s = K.Setup.load_json(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json"), window_bits=wb)
    log("setup c=%d" % wb, s.info(), "wall", round(time.time() - t, 2))
    for i in (0, 1, 2, 5, 6):
        t = time.time(); c = s.blob_to_commitment(vb[i]); log(" commit", i, c.hex()[:24], round((time.time() - t) * 1e3, 2), "ms")
    c = s.blob_to_commitment(vb[2])
    t = time.time(); p = s.blob_proof(vb[2], c); log(" blob_proof r1", p.hex()[:24], round((time.time() - t) * 1e3, 2), "ms")
    t = time.time(); ok = s.verify_blob_proof(vb[2], c, p); log(" verify_blob_proof", ok, round((time.time() - t) * 1e3, 2), "ms")
    t = time.time(); ok = s.verify_blob_proof_batch(vb[1:5], [s.blob_to_commitment(b) for b in vb[1:5]], [s.blob_proof(b, s.blob_to_commitment(b)) for b in vb[1:5]]); log(" verify batch 4", ok, round((time.time() - t) * 1e3, 2), "ms")
    n = 512
    cat = b"".join(vb[2 + (i % 3)] for i in range(n))
    for rep in range(2):
        t = time.time(); cm, st = s.blob_to_commitment_many(cat, n); dt = time.time() - t
        log(" commit x%d (host buffers)" % n, round(dt * 1e3, 1), "ms ->", round(n / dt), "blobs/s", "launches", s.info()["kernel_launches"])
    t = time.time(); pr, st = s.blob_proof_many(cat, cm, n); dt = time.time() - t
    log(" blob_proof x%d" % n, round(dt * 1e3, 1), "ms ->", round(n / dt), "blobs/s")
    t = time.time(); ok = s.verify_blob_proof_batch_concat(cat, cm, pr, n); dt = time.time() - t
    log(" verify batch x%d" % n, ok, round(dt * 1e3, 1), "ms ->", round(n / dt), "blobs/s")
    s.close()
log("probe done")
