#!/usr/bin/env python3
"""Where the batch-verification stage spends its time (4096 resident blobs): the shard call, the finish call (combine kernel + host pairing)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import kateth_b200 as K
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
s = K.Setup.load_json(bench.SETUP_JSON, window_bits=16)
dev = torch.device("cuda:0")
d_blobs = torch.from_numpy(bench.synth_blobs(n, 7)).to(dev)
d_c = torch.empty(n * 48, dtype=torch.uint8, device=dev); d_p = torch.empty(n * 48, dtype=torch.uint8, device=dev)
d_st = torch.zeros(n, dtype=torch.int32, device=dev); rec = torch.zeros(K.RECORD_BYTES, dtype=torch.uint8, device=dev)
assert s.blob_to_commitment_device(d_blobs.data_ptr(), n, d_c.data_ptr(), d_st.data_ptr()) == 0
assert s.blob_proof_device(d_blobs.data_ptr(), d_c.data_ptr(), n, d_p.data_ptr(), d_st.data_ptr()) == 0
for it in range(6):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    rc = s.verify_partial_device(d_blobs.data_ptr(), d_c.data_ptr(), d_p.data_ptr(), n, 0, n, rec.data_ptr(), d_st.data_ptr())
    t1 = time.perf_counter()
    ok = s.verify_finish_device(rec.data_ptr(), 1)
    t2 = time.perf_counter()
    print("partial %.2f ms  finish %.2f ms  rc %d ok %s" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, rc, ok), flush=True)
s.close()
