"""GPU tests of the multi-GPU front end inside the C ABI (kb200_group_*, kateth_b200/csrc/group.cpp): one replica per GPU,
contiguous shards, the NCCL all-gather of batch verification done by the library, one verdict per call
(kateth: Setup::verify_blob_proof_batch, /root/reference/src/kzg/setup.rs:247-275).  Runs on however many GPUs the box has:
a group of one exercises the same code without a communicator; the multi-GPU cases skip on a single-GPU box."""
import hashlib, os, subprocess, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import reconstruct_vectors as RV
from test_gpu_parity import rand_blob

pytestmark = pytest.mark.gpu
R = RV.R
SETUP_JSON = os.path.join(ROOT, "tests", "golden", "trusted_setup_4096.json")


def _check_group(g, oracle_setup, n):
    import kateth_b200 as K
    base = [rand_blob(12000 + i) for i in range(3)]
    want_c = [oracle_setup.blob_to_commitment(b) for b in base]
    want_p = [oracle_setup.blob_proof(b, c) for b, c in zip(base, want_c)]
    idx = [(5 * i + i // 3) % 3 for i in range(n)]
    cat = b"".join(base[k] for k in idx)
    comm, st = g.blob_to_commitment_many(cat, n)
    assert st == [0] * n and comm == b"".join(want_c[k] for k in idx)
    proofs, st = g.blob_proof_many(cat, comm, n)
    assert st == [0] * n and proofs == b"".join(want_p[k] for k in idx)
    zs = b"".join(RV.fe(int.from_bytes(hashlib.sha256(b"gz%d" % i).digest(), "big")) for i in range(n))
    pr, ys, st = g.proof_many(cat, zs, n)
    assert st == [0] * n
    for i in (0, n - 1):
        assert (pr[48 * i:48 * i + 48], ys[32 * i:32 * i + 32]) == oracle_setup.proof(base[idx[i]], zs[32 * i:32 * i + 32])
    assert g.verify_proof_batch(pr, comm, zs, ys, n) is True
    ok, st = g.verify_proof_many(pr, comm, zs, ys, n)
    assert ok == [True] * n and st == [0] * n
    bad_y = bytearray(ys); bad_y[32 * (n - 1):32 * n] = RV.fe(int.from_bytes(ys[32 * (n - 1):32 * n], "big") + 1)
    assert g.verify_proof_batch(pr, comm, zs, bytes(bad_y), n) is False
    ok, st = g.verify_proof_many(pr, comm, zs, bytes(bad_y), n)
    assert ok == [True] * (n - 1) + [False] and st == [0] * n
    assert g.verify_blob_proof_batch(cat, comm, proofs, n) is True
    # the operands of the one pairing check equal the oracle's single-call sums, however the batch was sharded
    blobs = [base[k] for k in idx]
    assert g.replica(0).last_pairing_operands() == oracle_setup.verify_blob_proof_batch_sums(
        blobs, [comm[48 * i:48 * i + 48] for i in range(n)], [proofs[48 * i:48 * i + 48] for i in range(n)])
    wrong = bytearray(proofs); wrong[48 * (n - 1):48 * n] = proofs[48 * (n - 2):48 * (n - 1)] if idx[n - 1] != idx[n - 2] else want_p[(idx[n - 1] + 1) % 3]
    assert g.verify_blob_proof_batch(cat, comm, bytes(wrong), n) is False
    # errors: a bad proof in the first shard, a bad blob in the last one -> the blob error (kateth's order), whichever GPU saw it
    badp = bytearray(proofs); badp[0:48] = RV.INVALID_G1[3]
    badb = bytearray(cat); badb[131072 * (n - 1) + 64:131072 * (n - 1) + 96] = RV.be32(R)
    with pytest.raises(K.KzgError) as e:
        g.verify_blob_proof_batch(bytes(badb), comm, bytes(badp), n)
    assert e.value.status == 2
    with pytest.raises(K.KzgError) as e:
        g.verify_blob_proof_batch(cat, comm, bytes(badp), n)
    assert e.value.status == 4
    c2, st2 = g.blob_to_commitment_many(bytes(badb), n)
    assert st2 == [0] * (n - 1) + [2] and c2[:48 * (n - 1)] == comm[:48 * (n - 1)]
    assert g.verify_blob_proof_batch(b"", b"", b"", 0) is True
    # the three pipelines over one upload per GPU
    oc, op, ok = g.commit_prove_verify(cat, n)
    assert ok is True and oc == comm and op == proofs
    with pytest.raises(K.KzgError) as e:
        g.commit_prove_verify(bytes(badb), n)
    assert e.value.status == 2


def test_group_of_one(oracle_setup, gpu_free):
    import kateth_b200 as K
    g = K.Group.load_json(SETUP_JSON, devices=[0], window_bits=9)
    try:
        assert g.info() == {"local_replicas": 1, "world": 1, "first_rank": 0, "nccl_version": 0}
        _check_group(g, oracle_setup, 7)
    finally:
        g.close()


def test_group_over_all_gpus_of_the_box(oracle_setup, gpu_free):
    """One process drives every GPU (ncclCommInitAll); 11 blobs do not divide evenly, so the shards are ragged."""
    import torch
    import kateth_b200 as K
    nd = torch.cuda.device_count()
    if nd < 2:
        pytest.skip("needs at least two GPUs")
    g = K.Group.load_json(SETUP_JSON, window_bits=9)
    try:
        info = g.info()
        assert info["local_replicas"] == nd and info["world"] == nd and info["nccl_version"] > 20000
        _check_group(g, oracle_setup, 11)
        _check_group(g, oracle_setup, nd - 1 if nd > 2 else 3)      # fewer items than GPUs: empty shards take part in the exchange
    finally:
        g.close()


def test_group_one_process_per_gpu(gpu_free):
    """torchrun, one rank per GPU: rank 0's kb200_group_unique_id travels over the (gloo) process group, every rank joins with its own
    handle, the all-gather of the records is the library's own NCCL call (tests/_group_worker.py)."""
    import torch
    nd = torch.cuda.device_count()
    if nd < 2:
        pytest.skip("needs at least two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(min(nd, 4)), "--master-addr", "127.0.0.1", "--master-port", "29641",
           os.path.join(ROOT, "tests", "_group_worker.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert p.returncode == 0, p.stdout[-3000:] + p.stderr[-3000:]
    assert "rank 0 ok" in p.stdout
