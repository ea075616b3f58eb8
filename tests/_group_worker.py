"""Worker of tests/test_gpu_group.py::test_group_one_process_per_gpu (torchrun, one rank per GPU).  Every rank loads its own
replica, joins the library's NCCL group (kb200_group_create_rank) and verifies ITS shard of one batch; all ranks must get the same
verdict / the same error."""
import os, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import torch.distributed as dist
import reconstruct_vectors as RV
import kateth_b200 as K
from kateth_b200.setup import shard_bounds
from oracle import oracle as O
from test_gpu_parity import rand_blob


def main():
    dist.init_process_group("gloo")
    rank, world, local = dist.get_rank(), dist.get_world_size(), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    path = os.path.join(ROOT, "tests", "golden", "trusted_setup_4096.json")
    s = K.Setup.load_json(path, device=local, window_bits=9)
    uid = [K.Group.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    g = K.Group.join(s, rank, world, uid[0])
    assert g.info()["world"] == world and g.info()["first_rank"] == rank
    o = O.Setup.load_json(path)
    n = 2 * world + 1
    base = [rand_blob(13000 + i) for i in range(2)]
    blobs = [base[i % 2] for i in range(n)]
    cs = [o.blob_to_commitment(b) for b in base]
    ps = [o.blob_proof(b, c) for b, c in zip(base, cs)]
    lo, hi = shard_bounds(n, world, rank)
    cat = b"".join(blobs[lo:hi]); comm = b"".join(cs[i % 2] for i in range(lo, hi)); proofs = b"".join(ps[i % 2] for i in range(lo, hi))
    # commit / prove on the shard: the engine's bytes equal the oracle's
    c, st = g.blob_to_commitment_many(cat, hi - lo)
    assert c == comm and st == [0] * (hi - lo)
    assert g.verify_blob_proof_batch(cat, comm, proofs, hi - lo, lo, n) is True
    want = o.verify_blob_proof_batch_sums(blobs, [cs[i % 2] for i in range(n)], [ps[i % 2] for i in range(n)])
    assert s.last_pairing_operands() == want, "rank %d: operands differ from the oracle's single-call sums" % rank
    # the last rank corrupts one proof: every rank sees false
    if rank == world - 1:
        bad = bytearray(proofs); bad[0:48] = ps[(lo + 1) % 2]
        proofs_bad = bytes(bad)
    else:
        proofs_bad = proofs
    assert g.verify_blob_proof_batch(cat, comm, proofs_bad, hi - lo, lo, n) is False
    # rank 0 holds a proof that is not a curve point, the last rank a non-canonical blob: every rank reports the blob error
    p2 = proofs
    c2 = cat
    if rank == 0:
        t = bytearray(proofs); t[0:48] = RV.INVALID_G1[3]; p2 = bytes(t)
    if rank == world - 1:
        t = bytearray(cat); t[64:96] = RV.be32(RV.R); c2 = bytes(t)
    try:
        g.verify_blob_proof_batch(c2, comm, p2, hi - lo, lo, n)
        raise AssertionError("no error")
    except K.KzgError as e:
        assert e.status == 2, e.status
    print("rank %d ok" % rank, flush=True)
    g.close(); s.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
