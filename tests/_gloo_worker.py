"""Worker of tests/test_multirank_cpu.py (launched by torch.distributed.run, backend gloo, CPU only).

The multi-GPU algorithm of batch verification (kateth_b200.setup.BatchVerifier): rank g owns the contiguous
shard shard_bounds(n, world, g), reduces it to a record {A_g, B_g, C_g, s_g} with weights that use the GLOBAL index, the
records are exchanged with ONE all-gather, every rank combines them into the two pairing operands.  Here the per-shard
records are computed with the oracle's G1 arithmetic (the CUDA engine needs a GPU), exchanged over gloo in the engine's
336-byte wire size (sums + the shard's first error of each class), and the combined operands must equal the oracle's
single-call sums.  A second exchange plants errors in the shards: every rank must read the SAME batch error off the records,
in kateth's order (blob errors of any rank before commitment errors before proof errors)."""
import hashlib, os, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.distributed as dist
import reconstruct_vectors as RV
from oracle import oracle as O
from kateth_b200.setup import shard_bounds, RECORD_BYTES, batch_error_of_records

R = RV.R


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    s = O.Setup.load_json(os.path.join(ROOT, "tests", "golden", "trusted_setup_4096.json"))
    n = 5
    blobs = RV.valid_blobs()[1:1 + n]                      # every rank derives the same inputs
    cs = [s.blob_to_commitment(b) for b in blobs]
    ps = [s.blob_proof(b, c) for b, c in zip(blobs, cs)]
    # kateth's batch challenge and weights (src/kzg/setup.rs:127-136, Fr::pow(r, 0) == r)
    r = int.from_bytes(hashlib.sha256(b"RCKZGBATCH___V1_" + (4096).to_bytes(16, "big") + n.to_bytes(16, "big")).digest(), "big") % R
    w = [pow(r, max(i, 1), R) for i in range(n)]
    lo, hi = shard_bounds(n, world, rank)
    zy = [s.challenge_eval(blobs[i], cs[i]) for i in range(lo, hi)]
    be = lambda x: (x % R).to_bytes(32, "big")
    if hi > lo:
        A = O.g1_lincomb(ps[lo:hi], [be(w[i]) for i in range(lo, hi)])
        B = O.g1_lincomb(ps[lo:hi], [be(w[i] * int.from_bytes(zy[i - lo][0], "big")) for i in range(lo, hi)])
        Cc = O.g1_lincomb(cs[lo:hi], [be(w[i]) for i in range(lo, hi)])
        sg = sum(w[i] * int.from_bytes(zy[i - lo][1], "big") for i in range(lo, hi)) % R
    else:
        A = B = Cc = RV.INF; sg = 0
    rec = torch.zeros(RECORD_BYTES, dtype=torch.uint8)
    rec[:176] = torch.frombuffer(bytearray(A + B + Cc + be(sg)), dtype=torch.uint8)     # 3 x 48 + 32 bytes inside the record slot
    gathered = torch.zeros(world * RECORD_BYTES, dtype=torch.uint8)
    dist.all_gather_into_tensor(gathered, rec)                                         # the ONLY data-path collective
    lhs, rhs, ssum = RV.INF, RV.INF, 0
    for g in range(world):
        raw = bytes(gathered[g * RECORD_BYTES:g * RECORD_BYTES + 176].numpy())
        lhs = O.g1_add(lhs, raw[0:48]); rhs = O.g1_add(rhs, raw[96:144]); rhs = O.g1_add(rhs, raw[48:96])
        ssum = (ssum + int.from_bytes(raw[144:176], "big")) % R
    rhs = O.g1_add(rhs, O.g1_lincomb([RV.G1], [be(R - ssum)]))
    want = s.verify_blob_proof_batch_sums(blobs, cs, ps)
    assert (lhs, rhs) == want, "rank %d: sharded operands differ from the single-call sums" % rank
    assert batch_error_of_records(gathered.numpy().tobytes()) == 0
    # errors travel with the records: rank 0's shard has a bad proof (class 2 of the record), the last rank's a bad blob (class 0)
    import struct
    rec2 = rec.clone()
    if rank == 0:
        rec2[320:336] = torch.frombuffer(bytearray(struct.pack("<4i", 0, 0, 4, 0)), dtype=torch.uint8)
    if rank == world - 1:
        rec2[320:336] = torch.frombuffer(bytearray(struct.pack("<4i", 2, 0, 4 if rank == 0 else 0, 0)), dtype=torch.uint8)
    dist.all_gather_into_tensor(gathered, rec2)
    assert batch_error_of_records(gathered.numpy().tobytes()) == 2, "rank %d: the blob error of the last rank must win" % rank
    ok = torch.tensor([1], dtype=torch.int32)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    assert int(ok) == 1
    print("rank %d of %d: shard [%d, %d) ok" % (rank, world, lo, hi), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
