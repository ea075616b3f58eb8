#!/usr/bin/env python3
"""Side measurements for the BASELINE.json configs that are not the bench.py line, one JSON object per line:

  config 2  single-blob blob_to_commitment (4096-point MSM) on one B200: median latency through the host-buffer C ABI
            (131072 bytes in, 48 bytes out) and with the blob resident in HBM.
  config 5  compute_kzg_proof at random evaluation points + verify_kzg_proof (one RLC batch) over --blobs blobs:
            blobs/s of proof(z) -> (proof, y), then of verify_proof_batch, inputs resident.
  sweep     commit throughput against batch size (where the XYZZ shape hands over to the batch-affine shape).

Every output is cross-checked (proofs verify, a corrupted proof does not).  Needs a GPU; no CPU fallback.
"""
import argparse, json, os, statistics, sys, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--blobs", type=int, default=4096)
    ap.add_argument("--window-bits", type=int, default=16)
    ap.add_argument("--iters", type=int, default=300)
    args = ap.parse_args()
    import torch
    import kateth_b200 as K
    dev = torch.device("cuda", 0)
    s = K.Setup.load_json(B.SETUP_JSON, device=0, window_bits=args.window_bits)
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    n = args.blobs
    h_blobs = torch.from_numpy(B.synth_blobs(n, 0x4b5a4700 + 5)).pin_memory()
    d_blobs = h_blobs.to(dev)
    d_comm = torch.empty(n * 48, dtype=torch.uint8, device=dev)
    d_proof = torch.empty(n * 48, dtype=torch.uint8, device=dev)
    d_y = torch.empty(n * 32, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(n, dtype=torch.int32, device=dev)

    def ev_ms(fn, reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    # ---- config 2: single blob
    h_out = torch.empty(48, dtype=torch.uint8).pin_memory()
    lat = []
    for i in range(args.iters + 20):
        t = time.perf_counter()
        rc = s.blob_to_commitment_host(h_blobs.data_ptr() + (i % n) * B.BLOB, 1, h_out.data_ptr())
        dt = time.perf_counter() - t
        assert rc == 0
        if i >= 20:
            lat.append(dt * 1e3)
    res_ms = ev_ms(lambda: s.blob_to_commitment_device(d_blobs.data_ptr(), 1, d_comm.data_ptr(), d_status.data_ptr()), 50)
    print(json.dumps({"config": 2, "what": "single-blob blob_to_commitment", "host_buffer_latency_ms_median": statistics.median(lat),
                      "host_buffer_latency_ms_p95": sorted(lat)[int(0.95 * len(lat))], "resident_latency_ms": res_ms,
                      "blobs_per_s_one_at_a_time": 1e3 / statistics.median(lat), "msm_shape": "xyzz, 32 CTAs per blob"}), flush=True)

    # ---- sweep: commit throughput against batch size
    sweep = []
    for m in [1, 8, 64, 148, 296, 592, 1184, 2368, 4096]:
        if m > n:
            break
        s.blob_to_commitment_device(d_blobs.data_ptr(), m, d_comm.data_ptr(), d_status.data_ptr())      # first call at a new size grows the scratch buffers
        ms = ev_ms(lambda: s.blob_to_commitment_device(d_blobs.data_ptr(), m, d_comm.data_ptr(), d_status.data_ptr()), 3)
        sweep.append({"blobs": m, "ms": ms, "blobs_per_s": m / ms * 1e3, "batch_affine": bool(s.info()["last_msm_batch_affine"])})
    print(json.dumps({"config": "sweep", "what": "blob_to_commitment, resident, by batch size", "points": sweep}), flush=True)

    # ---- config 5: proofs at random points, then one RLC verification of all of them
    z = torch.from_numpy(B.synth_blobs((n + 4095) // 4096, 0x4b5a4700 + 55)[: n * 32].copy()).to(dev)      # canonical field elements
    assert s.blob_to_commitment_device(d_blobs.data_ptr(), n, d_comm.data_ptr(), d_status.data_ptr()) == 0
    assert s.proof_device(d_blobs.data_ptr(), z.data_ptr(), n, d_proof.data_ptr(), d_y.data_ptr(), d_status.data_ptr()) == 0   # first call at this size grows the workspace: not timed
    ms_proof = ev_ms(lambda: s.proof_device(d_blobs.data_ptr(), z.data_ptr(), n, d_proof.data_ptr(), d_y.data_ptr(), d_status.data_ptr()), 3)
    pr, cm, zz, yy = (bytes(t.cpu().numpy()) for t in (d_proof, d_comm, z, d_y))
    ok = s.verify_proof_batch(pr, cm, zz, yy, n)          # first call allocates
    t0 = time.perf_counter()
    for _ in range(3):
        ok = s.verify_proof_batch(pr, cm, zz, yy, n) and ok
    ms_verify = (time.perf_counter() - t0) * 1e3 / 3
    bad = bytearray(pr); bad[48:96] = pr[0:48]
    ok_bad = s.verify_proof_batch(bytes(bad), cm, zz, yy, n)
    one = s.verify_proof(pr[:48], cm[:48], zz[:32], yy[:32])
    assert ok is True and ok_bad is False and one is True
    print(json.dumps({"config": 5, "what": "compute_kzg_proof at random z + verify_kzg_proof batch", "blobs": n, "proof_ms": ms_proof,
                      "proof_blobs_per_s": n / ms_proof * 1e3, "verify_batch_ms_host_buffers": ms_verify, "verify_blobs_per_s": n / ms_verify * 1e3,
                      "verdict_valid": ok, "verdict_one_corrupted": ok_bad}), flush=True)
    s.close()


if __name__ == "__main__":
    main()
