#!/usr/bin/env python3
"""Small end-to-end pass for compute-sanitizer (memcheck / racecheck): a narrow table (window_bits 6), 600 blobs through commit, blob proof,
batch verification in both forms of the sums, the allocation-free commit and the device pairing.
  compute-sanitizer --tool memcheck  python tools/sanitize_probe.py
  compute-sanitizer --tool racecheck python tools/sanitize_probe.py
(compute-sanitizer is closed on this build's GPU pool; without it the script is still a 6-second smoke pass over every pipeline.)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import kateth_b200 as K
import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 600
s = K.Setup.load_json(bench.SETUP_JSON, window_bits=6)
blobs = bench.synth_blobs(n, 11)
cat = bytes(blobs)
comm, st = s.blob_to_commitment_many(cat, n)
assert not any(st)
proofs, st = s.blob_proof_many(cat, comm, n)
assert not any(st)
for mode in (1, 2):
    s.set_rlc_mode(mode)
    assert s.verify_blob_proof_batch_concat(cat, comm, proofs, n) is True
s.set_rlc_mode(0)
d_b = torch.from_numpy(np.frombuffer(cat, dtype=np.uint8)[: 64 * 131072].copy()).cuda()
out = torch.zeros(64 * 48, dtype=torch.uint8, device="cuda"); dst = torch.zeros(64, dtype=torch.int32, device="cuda")
torch.cuda.synchronize()
assert s.blob_to_commitment_device_noalloc(d_b.data_ptr(), 64, out.data_ptr(), dst.data_ptr()) == 0
torch.cuda.synchronize()
assert bytes(out.cpu().numpy()) == comm[: 64 * 48]
zs = cat[: 32 * 8]
pr, ys, st = s.proof_many(cat[: 8 * 131072], zs, 8)
assert s.verify_proof_many(pr, comm[: 8 * 48], zs, ys, 8)[0] == [True] * 8
s.close()
print("sanitize_probe ok")
