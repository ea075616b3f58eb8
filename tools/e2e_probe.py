#!/usr/bin/env python3
"""Wall-clock time of each host-buffer C-ABI call of the bench step (pinned host buffers, 4096 blobs) next to the same call
on resident buffers: where the end-to-end step spends what the resident step does not.  KB200_H2D_SPLIT=0 for one-piece copies."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B


def main():
    import torch
    import kateth_b200 as K
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    dev = torch.device("cuda", 0)
    s = K.Setup.load_json(B.SETUP_JSON, device=0, window_bits=int(os.environ.get("KB200_WINDOW_BITS", "16")))
    s.set_stream(torch.cuda.current_stream().cuda_stream)
    h_blobs = torch.from_numpy(B.synth_blobs(n, 0x4b5a4700)).pin_memory()
    h_comm = torch.empty(n * 48, dtype=torch.uint8).pin_memory()
    h_proof = torch.empty(n * 48, dtype=torch.uint8).pin_memory()
    d_blobs = h_blobs.to(dev)
    d_comm = torch.empty(n * 48, dtype=torch.uint8, device=dev)
    d_proof = torch.empty(n * 48, dtype=torch.uint8, device=dev)
    d_status = torch.zeros(n, dtype=torch.int32, device=dev)
    d_rec = torch.empty(320, dtype=torch.uint8, device=dev)

    def wall(fn, reps=3):
        fn(); torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t) / reps * 1e3

    out = {"blobs": n, "split": os.environ.get("KB200_H2D_SPLIT", "1")}
    out["h2d_blobs_ms"] = wall(lambda: d_blobs.copy_(h_blobs, non_blocking=True))
    out["commit_host_ms"] = wall(lambda: s.blob_to_commitment_host(h_blobs.data_ptr(), n, h_comm.data_ptr()))
    out["commit_resident_ms"] = wall(lambda: s.blob_to_commitment_device(d_blobs.data_ptr(), n, d_comm.data_ptr(), d_status.data_ptr()))
    out["prove_host_ms"] = wall(lambda: s.blob_proof_host(h_blobs.data_ptr(), h_comm.data_ptr(), n, h_proof.data_ptr()))
    out["prove_resident_ms"] = wall(lambda: s.blob_proof_device(d_blobs.data_ptr(), d_comm.data_ptr(), n, d_proof.data_ptr(), d_status.data_ptr()))
    out["verify_host_ms"] = wall(lambda: s.verify_partial_host(h_blobs.data_ptr(), h_comm.data_ptr(), h_proof.data_ptr(), n, 0, n, d_rec.data_ptr()))
    out["verify_resident_ms"] = wall(lambda: s.verify_partial_device(d_blobs.data_ptr(), d_comm.data_ptr(), d_proof.data_ptr(), n, 0, n, d_rec.data_ptr(), d_status.data_ptr()))
    print(json.dumps(out), flush=True)
    s.close()


if __name__ == "__main__":
    main()
