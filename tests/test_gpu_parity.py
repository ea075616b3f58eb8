"""GPU parity tests (run with -m gpu on a B200): the CUDA engine, called through the C ABI, against
  (1) the CPU oracle on the same seeded inputs,
  (2) the 253 sha-pinned consensus-spec-tests vectors (every handler kateth tests, src/kzg/setup.rs:319-463),
  (3) size-independent properties at larger batch sizes.
Bit-exact is the bar everywhere: 48-byte compressed G1, 32-byte scalars, booleans, error classes."""
import hashlib, os, random, sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import reconstruct_vectors as RV

pytestmark = pytest.mark.gpu
R = RV.R
P = 0x1a0111ea397fe69a4b1ba7b6434bacd764774b84f38512bf6730d2a0f6b0f6241eabfffeb153ffffb9feffffffffaaab


def rand_blob(seed):
    """Canonical field elements: SHA-256 counter stream, top bit cleared, rejected if >= r."""
    out = bytearray()
    ctr = 0
    while len(out) < 131072:
        h = bytearray(hashlib.sha256(b"kb200-blob" + seed.to_bytes(8, "little") + ctr.to_bytes(8, "little")).digest())
        ctr += 1
        h[0] &= 0x7f
        if int.from_bytes(h, "big") < R:
            out += h
    return bytes(out)


def limbs32(v, n):
    return [(v >> (32 * i)) & 0xffffffff for i in range(n)]


# ---------------------------------------------------------------- K1: field products vs the oracle
def test_fp_mul_bit_exact():
    import kateth_b200 as K
    from oracle import oracle as O
    rng = random.Random(1)
    edge = [0, 1, 2, P - 1, P - 2, (P - 1) // 2, 2 ** 380, 2 ** 381 - 1 - (2 ** 381 - 1 >= P) * (2 ** 381 - P), 0xffffffff, 2 ** 64 - 1]
    vals = [(a, b) for a in edge for b in edge] + [(rng.randrange(P), rng.randrange(P)) for _ in range(20000)]
    a = np.array([limbs32(x, 12) for x, _ in vals], dtype=np.uint32)
    b = np.array([limbs32(y, 12) for _, y in vals], dtype=np.uint32)
    got = K.test_hooks.fp_mul(a, b)
    exp = O.fp_mul_raw(a.view(np.uint64), b.view(np.uint64)).view(np.uint32)
    assert np.array_equal(got, exp)
    # and against plain integers: a*b/2^384 mod p
    rinv = pow(2 ** 384, -1, P)
    for k in range(0, len(vals), 997):
        x, y = vals[k]
        assert sum(int(got[k][i]) << (32 * i) for i in range(12)) == x * y * rinv % P


def test_fr_mul_bit_exact():
    import kateth_b200 as K
    from oracle import oracle as O
    rng = random.Random(2)
    edge = [0, 1, 2, R - 1, R - 2, (R - 1) // 2, 2 ** 254, 0xffffffff, 2 ** 64 - 1]
    vals = [(a, b) for a in edge for b in edge] + [(rng.randrange(R), rng.randrange(R)) for _ in range(20000)]
    a = np.array([limbs32(x, 8) for x, _ in vals], dtype=np.uint32)
    b = np.array([limbs32(y, 8) for _, y in vals], dtype=np.uint32)
    got = K.test_hooks.fr_mul(a, b)
    exp = O.fr_mul_raw(a.view(np.uint64), b.view(np.uint64)).view(np.uint32)
    assert np.array_equal(got, exp)


# ---------------------------------------------------------------- golden commitments / proofs
GOLD_COMM = {
    0: "c00000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000000",
    1: "a572cbea904d67468808c8eb50a9450c9721db309128012543902d0ac358a62ae28f75bb8f1c7c42c39a8c5529bf0f4e",
    2: "a421e229565952cfff4ef3517100a97da1d4fe57956fa50a442f92af03b1bf37adacc8ad4ed209b31287ea5bb94d9d06",
    5: "b7f1d3a73197d7942695638c4fa9ac0fc3688c4f9774b905a14e3a3f171bac586c55e83ff97a1aeffb3af00adb22c6bb",
    6: "93efc82d2017e9c57834a1246463e64774e56183bb247c8fc9dd98c56817e878d97b05f5c8d900acf1fbbbca6f146556",
}


def test_golden_commitments(engine_setup):
    vb = RV.valid_blobs()
    for i, hexv in GOLD_COMM.items():
        assert engine_setup.blob_to_commitment(vb[i]).hex() == hexv, i


def test_golden_blob_proof_and_in_domain_proofs(engine_setup):
    vb = RV.valid_blobs()
    c = engine_setup.blob_to_commitment(vb[2])
    assert engine_setup.blob_proof(vb[2], c).hex() == "a2aeea08a9cd37fb0b089b1938bbe7eedd4ea6120dc70f45d59ad077008d08be115b858350b1eff645148fe4470b65c8"
    # r1 at z = 1 (in the domain, brp index 0) and z = omega (brp index 2048)
    p, y = engine_setup.proof(vb[2], RV.FE_VALID[1])
    assert p.hex() == "b0c829a8d2d3405304fecbea193e6c67f7c3912a6adc7c3737ad3f8a3b750425c1531a7426f03033a3994bc82a10609f"
    assert y.hex() == "1824b159acc5056f998c4fefecbc4ff55884b7fa0003480200000001fffffffe"
    p, y = engine_setup.proof(vb[2], RV.FE_VALID[5])
    assert p.hex() == "a444d6bb5aadc3ceb615b50d6606bd54bfe529f59247987cd1ab848d19de599a9052f1835fb0d0d44cf70183e19a68c9"
    assert y.hex() == "6d928e13fe443e957d82e3e71d48cb65d51028eb4483e719bf8efcdf12f7c321"


# ---------------------------------------------------------------- all 253 vectors through the engine
def test_all_253_vectors_through_engine(engine_setup):
    engine_setup.set_msm_mode(0)
    counts, bad, missing = RV.check_all("engine", setup=engine_setup)
    assert sum(counts.values()) == 253
    assert not bad, bad[:5]
    assert not missing


# ---------------------------------------------------------------- random blobs vs the oracle, same seeded inputs
def test_random_blobs_match_oracle(engine_setup, oracle_setup):
    n = 6
    blobs = [rand_blob(100 + i) for i in range(n)]
    cat = b"".join(blobs)
    comm, st = engine_setup.blob_to_commitment_many(cat, n)
    assert st == [0] * n
    for i in range(n):
        assert comm[48 * i:48 * i + 48] == oracle_setup.blob_to_commitment(blobs[i])
    proofs, st = engine_setup.blob_proof_many(cat, comm, n)
    assert st == [0] * n
    for i in range(2):
        assert proofs[48 * i:48 * i + 48] == oracle_setup.blob_proof(blobs[i], comm[48 * i:48 * i + 48])
    # challenge/evaluation and the RLC operands of the pairing check are diffed too (intermediate sums)
    cs = [comm[48 * i:48 * i + 48] for i in range(n)]
    ps = [proofs[48 * i:48 * i + 48] for i in range(n)]
    assert engine_setup.verify_blob_proof_batch(blobs, cs, ps) is True
    lhs, rhs = engine_setup.last_pairing_operands()
    assert (lhs, rhs) == oracle_setup.verify_blob_proof_batch_sums(blobs, cs, ps)
    # proof at random points
    zs = [RV.fe(int.from_bytes(hashlib.sha256(b"z%d" % i).digest(), "big")) for i in range(n)]
    pr, ys, st = engine_setup.proof_many(cat, b"".join(zs), n)
    assert st == [0] * n
    for i in range(2):
        assert (pr[48 * i:48 * i + 48], ys[32 * i:32 * i + 32]) == oracle_setup.proof(blobs[i], zs[i])
    assert engine_setup.verify_proof_batch(pr, comm, b"".join(zs), ys, n) is True
    for i in range(n):
        assert engine_setup.verify_proof(pr[48 * i:48 * i + 48], cs[i], zs[i], ys[32 * i:32 * i + 32]) is True


# ---------------------------------------------------------------- properties at batch scale
def test_batch_properties(engine_setup, oracle_setup):
    import kateth_b200 as K
    n = 300                                    # > 148 SMs: exercises the one-CTA-per-blob shape and ragged tails
    base = [rand_blob(1000 + i) for i in range(3)]
    # linearity: commit(a) + commit(b) == commit(a + b mod r)
    from oracle import oracle as O
    a_int = [int.from_bytes(base[0][32 * k:32 * k + 32], "big") for k in range(4096)]
    b_int = [int.from_bytes(base[1][32 * k:32 * k + 32], "big") for k in range(4096)]
    sum_blob = b"".join(RV.fe(x + y) for x, y in zip(a_int, b_int))
    blobs = [base[i % 3] for i in range(n - 1)] + [sum_blob]
    cat = b"".join(blobs)
    comm, st = engine_setup.blob_to_commitment_many(cat, n)
    assert st == [0] * n
    cs = [comm[48 * i:48 * i + 48] for i in range(n)]
    for i in range(n - 1):
        assert cs[i] == cs[i % 3]            # same blob -> same commitment, whatever CTA computed it
    assert O.g1_add(cs[0], cs[1]) == cs[n - 1]
    proofs, st = engine_setup.blob_proof_many(cat, comm, n)
    assert st == [0] * n
    ps = [proofs[48 * i:48 * i + 48] for i in range(n)]
    assert ps[0] == oracle_setup.blob_proof(base[0], cs[0])
    assert engine_setup.verify_blob_proof_batch_concat(cat, comm, proofs, n) is True
    # one corrupted proof flips the verdict
    bad = bytearray(proofs); bad[48 * 137:48 * 138] = O.g1_add(ps[137], RV.G1)
    assert engine_setup.verify_blob_proof_batch_concat(cat, comm, bytes(bad), n) is False
    # one bad blob makes the whole call an error, with per-item status
    badcat = bytearray(cat); badcat[131072 * 5 + 32 * 77:131072 * 5 + 32 * 78] = RV.be32(R)
    c2, st2 = engine_setup.blob_to_commitment_many(bytes(badcat), 8)
    assert st2 == [0, 0, 0, 0, 0, 2, 0, 0] and c2[48 * 4:48 * 5] == cs[4]
    with pytest.raises(K.KzgError) as e:
        engine_setup.verify_blob_proof_batch_concat(bytes(badcat), comm, proofs, n)
    assert e.value.status == 2


def test_sharded_partials_equal_single_call(engine_setup):
    """The multi-GPU path emulated on one GPU: shards with index offsets + finish == single call."""
    import torch
    n = 10
    blobs = [rand_blob(2000 + i) for i in range(n)]
    cat = b"".join(blobs)
    comm, _ = engine_setup.blob_to_commitment_many(cat, n)
    proofs, _ = engine_setup.blob_proof_many(cat, comm, n)
    assert engine_setup.verify_blob_proof_batch_concat(cat, comm, proofs, n) is True
    single = engine_setup.last_pairing_operands()
    dev = torch.device("cuda:0")
    d_b = torch.frombuffer(bytearray(cat), dtype=torch.uint8).to(dev)
    d_c = torch.frombuffer(bytearray(comm), dtype=torch.uint8).to(dev)
    d_p = torch.frombuffer(bytearray(proofs), dtype=torch.uint8).to(dev)
    import kateth_b200 as K
    RB = K.RECORD_BYTES
    recs = torch.zeros(3 * RB, dtype=torch.uint8, device=dev)
    st = torch.zeros(n, dtype=torch.int32, device=dev)
    bounds = [(0, 3), (3, 4), (4, 10)]
    for g, (lo, hi) in enumerate(bounds):
        rc = engine_setup.verify_partial_device(d_b.data_ptr() + lo * 131072, d_c.data_ptr() + lo * 48, d_p.data_ptr() + lo * 48, hi - lo, lo, n,
                                                recs.data_ptr() + g * RB, st.data_ptr() + 4 * lo)
        assert rc == 0
    assert engine_setup.verify_finish_device(recs.data_ptr(), 3) is True
    assert engine_setup.last_pairing_operands() == single


def test_g1_decompress_classes(engine_setup):
    import kateth_b200 as K
    pts = [RV.G1, RV.INF, RV.INVALID_G1[2], RV.INVALID_G1[3], bytes([0x17]) + RV.G1[1:], b"\xe0" + bytes(47), b"\x9a" + b"\xff" * 47]
    st, re = K.test_hooks.g1_decompress(engine_setup, pts)
    assert st == [0, 0, 5, 4, 3, 3, 3]
    assert re[0] == RV.G1 and re[1] == RV.INF
    assert K.g1_validate(pts) == st          # the handle-free form behind the Rust wrapper's P1::decompress


GLV_N = 0xac45a4010001a4020000000100000000      # x^2; r = N^2 - N + 1 (kateth_b200/csrc/glv.cuh)


@pytest.mark.parametrize("wb,mode", [(9, 1), (15, 1), (9, 2), (13, 2), (15, 2), (109, 1), (109, 2), (111, 2), (16, 1), (16, 2)])
def test_other_window_width_gives_same_bytes(wb, mode, oracle_setup, gpu_free):
    """The window width is a tuning knob, not a semantic one.  15 is the widest plain table (17 windows, the top one unsigned
    and 15 bits wide, 107 GiB); 9 exercises a 3-bit top window; 16 and 100 + c are the GLV form (two 128-bit halves per scalar
    on ceil(128 / c) windows, the second half mapped through the endomorphism)."""
    import kateth_b200 as K
    s = K.Setup.load_json(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json"), window_bits=wb)
    try:
        s.set_msm_mode(mode)        # 1 = XYZZ accumulation (k_msm), 2 = batch-affine summation (k_msm_ba)
        vb = RV.valid_blobs()
        for i, hexv in GOLD_COMM.items():
            assert s.blob_to_commitment(vb[i]).hex() == hexv
        assert R == GLV_N * GLV_N - GLV_N + 1
        assert s.info()["window_bits"] == wb % 100 and s.info()["glv"] == (1 if wb >= 16 else 0)
        # r - 1 everywhere, then scalars whose top window is maximal with and without a carry coming in, then scalars whose GLV
        # halves are extreme (k1 = N - 1, k2 = N - 1, either half zero, top window of a half maximal)
        hi = [R - 1, R - 2 ** 240, (R >> 240 << 240) + 2 ** 239 + 5, (R >> 240 << 240) + 2 ** 239 - 5, 2 ** 254, 2 ** 240 - 1,
              GLV_N - 1 + 700, GLV_N + 700, (GLV_N - 2) * GLV_N + GLV_N - 1, (GLV_N - 1) * GLV_N, 5 * GLV_N + 700, 2 ** 128 + 700,
              (GLV_N >> 112 << 112) + 2 ** 111 + 705, ((GLV_N >> 112 << 112) + 2 ** 111 + 695) * GLV_N + GLV_N - 1]
        assert all(0 <= v - 700 and v < R for v in hi)
        blob = b"".join(RV.fe(hi[k % len(hi)] - (k // len(hi))) for k in range(4096))
        c = s.blob_to_commitment(blob)
        assert c == oracle_setup.blob_to_commitment(blob)
        blobs = [rand_blob(3000 + i) for i in range(3)] + [blob]
        cat = b"".join(blobs)
        comm, st = s.blob_to_commitment_many(cat, 4)
        proofs, st2 = s.blob_proof_many(cat, comm, 4)
        assert st == [0] * 4 and st2 == [0] * 4
        for i in range(4):
            assert comm[48 * i:48 * i + 48] == oracle_setup.blob_to_commitment(blobs[i])
        assert proofs[:48] == oracle_setup.blob_proof(blobs[0], comm[:48])
        assert s.verify_blob_proof_batch_concat(cat, comm, proofs, 4) is True
    finally:
        s.close()


# ---------------------------------------------------------------- subgroup test: endomorphism criterion == [r]P definition
def _ec_add(P, Q):
    if P is None: return Q
    if Q is None: return P
    x1, y1 = P; x2, y2 = Q
    if x1 == x2:
        if (y1 + y2) % P_MOD == 0: return None
        l = 3 * x1 * x1 * pow(2 * y1, -1, P_MOD) % P_MOD
    else:
        l = (y2 - y1) * pow(x2 - x1, -1, P_MOD) % P_MOD
    x3 = (l * l - x1 - x2) % P_MOD
    return x3, (l * (x1 - x3) - y1) % P_MOD


def _ec_mul(k, pt):
    acc = None
    while k:
        if k & 1: acc = _ec_add(acc, pt)
        pt = _ec_add(pt, pt); k >>= 1
    return acc


P_MOD = P


def _compress(pt):
    if pt is None:
        return RV.INF
    x, y = pt
    b = bytearray(x.to_bytes(48, "big")); b[0] |= 0x80
    if y > (P_MOD - 1) // 2: b[0] |= 0x20
    return bytes(b)


def test_subgroup_check_matches_definition(engine_setup):
    """k_g1_decompress uses phi(P) == [-x^2]P (g1.cuh); it must agree with [r]P == infinity on points of the curve that are
    outside G1 (random points, small-order torsion points) and inside it (cofactor-cleared points)."""
    import kateth_b200 as K
    H_COF = 0x396c8c005555e1568c00aaab0000aaab
    rng = random.Random(7)
    pts = []
    while len(pts) < 40:
        x = rng.randrange(P_MOD); a = (x ** 3 + 4) % P_MOD; y = pow(a, (P_MOD + 1) // 4, P_MOD)
        if y * y % P_MOD == a:
            pts.append((x, y))
    cleared = [_ec_mul(H_COF, q) for q in pts[:6]]                       # in G1
    torsion = [_ec_mul(R * (H_COF // q), pts[10 + k]) for k, q in enumerate((3, 11, 121, 10177))]   # order divides q: not in G1
    allp = pts + cleared + [t for t in torsion if t is not None]
    enc = [_compress(q) for q in allp]
    st_fast, re_fast = K.test_hooks.g1_decompress(engine_setup, enc)
    st_def, re_def = K.test_hooks.g1_decompress(engine_setup, enc, by_order=True)
    assert st_fast == st_def
    assert st_fast[:40] == [5] * 40 and st_fast[40:46] == [0] * 6 and all(s == 5 for s in st_fast[46:])
    assert re_fast[40:46] == enc[40:46]


# ---------------------------------------------------------------- batch-affine MSM: all vectors, and a degenerate setup
def test_all_253_vectors_through_batch_affine_msm(engine_setup):
    engine_setup.set_msm_mode(2)
    try:
        counts, bad, missing = RV.check_all("engine", setup=engine_setup)
        assert engine_setup.info()["last_msm_batch_affine"] == 1
    finally:
        engine_setup.set_msm_mode(0)
    assert sum(counts.values()) == 253
    assert not bad, bad[:5]
    assert not missing


@pytest.mark.parametrize("mode,wb", [(1, 9), (2, 9), (1, 109), (2, 109), (2, 13), (2, 116), (2, 16)])
def test_degenerate_setup_exercises_exceptional_additions(mode, wb, tmp_path, gpu_free):
    """A setup whose Lagrange points are small multiples of the generator makes gathered table points and partial sums
    collide all the time (P == Q, P == -Q, infinity): the complete-formula paths of both MSM shapes must still give the
    group element the oracle's Pippenger gives."""
    import json
    import kateth_b200 as K
    from oracle import oracle as O
    real = json.load(open(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json")))
    gx = int.from_bytes(bytes([RV.G1[0] & 0x1f]) + RV.G1[1:], "big")
    gy = 0x08b3f481e3aaa0f1a09e30ed741d8ae4fcf5e095d5d00af600db18cb2c04b3edd03cc744a2888ae40caa232946c5e7e1
    mult = [None] + [_ec_mul(k, (gx, gy)) for k in range(1, 8)]
    pts = ["0x" + _compress(mult[1 + (i * 7 + i // 64) % 7]).hex() for i in range(4096)]
    path = tmp_path / "degenerate_setup.json"
    path.write_text(json.dumps({"g1_lagrange": pts, "g2_monomial": real["g2_monomial"]}))
    s = K.Setup.load_json(str(path), window_bits=wb)
    o = O.Setup.load_json(str(path))
    try:
        s.set_msm_mode(mode)
        rng = random.Random(77)
        small = [0, 1, 2, 3, 4, R - 1, R - 2, R - 3, 255, 256, 257, R - 256]
        blobs = [b"".join(RV.fe(rng.choice(small)) for _ in range(4096)),
                 b"".join(RV.fe(1) for _ in range(4096)),
                 b"".join(RV.fe(R - 1 if k & 1 else 1) for k in range(4096)),
                 b"".join(RV.fe(rng.randrange(R)) for _ in range(4096)),
                 b"".join(RV.fe(rng.randrange(1 << 18)) for _ in range(4096))]
        comm, st = s.blob_to_commitment_many(b"".join(blobs), len(blobs))
        assert st == [0] * len(blobs)
        for i, b in enumerate(blobs):
            assert comm[48 * i:48 * i + 48] == o.blob_to_commitment(b), i
    finally:
        s.close()


# ---------------------------------------------------------------- optional consensus-spec batch challenge
def test_consensus_spec_batch_challenge(engine_setup, oracle_setup):
    """Mode 1 of kb200_set_batch_challenge: r comes from the verify_kzg_proof_batch transcript and the weights are
    [1, r, r^2, ...].  The left pairing operand sum r^i * proof_i is recomputed here from hashlib + the oracle's G1."""
    from oracle import oracle as O
    s = engine_setup
    n = 5
    blobs = [rand_blob(4200 + i) for i in range(n)]
    cat = b"".join(blobs)
    comm, _ = s.blob_to_commitment_many(cat, n)
    proofs, _ = s.blob_proof_many(cat, comm, n)
    cs = [comm[48 * i:48 * i + 48] for i in range(n)]
    ps = [proofs[48 * i:48 * i + 48] for i in range(n)]
    s.set_batch_challenge(1)
    try:
        assert s.verify_blob_proof_batch(blobs, cs, ps) is True
        lhs, _ = s.last_pairing_operands()
        zy = [oracle_setup.challenge_eval(blobs[i], cs[i]) for i in range(n)]
        data = b"RCKZGBATCH___V1_" + (4096).to_bytes(8, "big") + n.to_bytes(8, "big")
        data += b"".join(cs[i] + zy[i][0] + zy[i][1] + ps[i] for i in range(n))
        r = int.from_bytes(hashlib.sha256(data).digest(), "big") % R
        assert lhs == O.g1_lincomb(ps, [pow(r, i, R).to_bytes(32, "big") for i in range(n)])
        # a wrong proof is still rejected, the vectors' verdicts do not change, shards are refused
        assert s.verify_blob_proof_batch(blobs, cs, [ps[1]] + ps[1:]) is False
        assert s.verify_blob_proof_batch([], [], []) is True
        zs, ys = b"".join(z for z, _ in zy), b"".join(y for _, y in zy)
        assert s.verify_proof_batch(proofs, comm, zs, ys, n) is True
        import torch
        d = lambda b: torch.frombuffer(bytearray(b), dtype=torch.uint8).cuda()
        db, dc, dp = d(cat), d(comm), d(proofs)
        import kateth_b200 as K
        rec = torch.zeros(K.RECORD_BYTES, dtype=torch.uint8, device="cuda"); st = torch.zeros(n, dtype=torch.int32, device="cuda")
        assert s.verify_partial_device(db.data_ptr(), dc.data_ptr(), dp.data_ptr(), 2, 0, n, rec.data_ptr(), st.data_ptr()) == -20
    finally:
        s.set_batch_challenge(0)
    assert s.verify_blob_proof_batch(blobs, cs, ps) is True
    assert s.last_pairing_operands() == oracle_setup.verify_blob_proof_batch_sums(blobs, cs, ps)


# ---------------------------------------------------------------- the batch-affine shape at its own batch size
def test_batch_affine_at_scale_with_bad_items(engine_setup, oracle_setup):
    """n >= 592 selects k_msm_ba on its own (automatic mode): persistent CTAs pull several blobs each, items with a non-zero
    status are skipped by the recoder and by the MSM, every other output is unaffected (idempotence over copies + oracle)."""
    import kateth_b200 as K
    from oracle import oracle as O
    n = 700
    base = [rand_blob(5000 + i) for i in range(4)] + [bytes(131072), b"".join(RV.fe(R - 1) for _ in range(4096))]
    blobs = [base[i % len(base)] for i in range(n)]
    bad_at = [0, 17, 599, 699]
    for i in bad_at:
        b = bytearray(blobs[i]); b[32 * (i % 4096):32 * (i % 4096) + 32] = RV.be32(R + i); blobs[i] = bytes(b)
    cat = b"".join(blobs)
    engine_setup.set_msm_mode(0)
    comm, st = engine_setup.blob_to_commitment_many(cat, n)
    assert engine_setup.info()["last_msm_batch_affine"] == 1
    assert [i for i in range(n) if st[i] != 0] == bad_at and all(st[i] == 2 for i in bad_at)
    want = [oracle_setup.blob_to_commitment(b) for b in base]
    for i in range(n):
        if i not in bad_at:
            assert comm[48 * i:48 * i + 48] == want[i % len(base)], i
    # proofs over the good items (quotient scalars, FMT 1 recoding)
    good = [i for i in range(n) if i not in bad_at][:640]
    gcat = b"".join(blobs[i] for i in good); gcomm = b"".join(comm[48 * i:48 * i + 48] for i in good)
    proofs, st2 = engine_setup.blob_proof_many(gcat, gcomm, len(good))
    assert engine_setup.info()["last_msm_batch_affine"] == 1 and st2 == [0] * len(good)
    wantp = [oracle_setup.blob_proof(b, c) for b, c in zip(base, want)]
    for k, i in enumerate(good):
        assert proofs[48 * k:48 * k + 48] == wantp[i % len(base)], i
    assert engine_setup.verify_blob_proof_batch_concat(gcat, gcomm, proofs, len(good)) is True


def test_host_batches_copied_in_two_pieces(engine_setup, oracle_setup):
    """From 2 x 592 blobs on, the host-buffer commit call copies a head of one MSM wave and the rest separately and runs them as
    two device passes (engine.cu h2d_split): outputs and statuses are those of one pass; the proof call takes the same inputs
    in one piece."""
    n = 1300
    base = [rand_blob(6000 + i) for i in range(3)] + [b"".join(RV.fe(R - 1 - k) for k in range(4096))]
    blobs = [base[i % len(base)] for i in range(n)]
    bad_at = [5, 591, 592, 1299]                  # both sides of the split
    for i in bad_at:
        b = bytearray(blobs[i]); b[32 * (i % 4096):32 * (i % 4096) + 32] = RV.be32(R + i); blobs[i] = bytes(b)
    cat = b"".join(blobs)
    engine_setup.set_msm_mode(0)
    comm, st = engine_setup.blob_to_commitment_many(cat, n)
    assert [i for i in range(n) if st[i] != 0] == bad_at
    want = [oracle_setup.blob_to_commitment(b) for b in base]
    for i in range(n):
        if i not in bad_at:
            assert comm[48 * i:48 * i + 48] == want[i % len(base)], i
    good = [i for i in range(n) if i not in bad_at][:1250]
    gcat = b"".join(blobs[i] for i in good); gcomm = b"".join(comm[48 * i:48 * i + 48] for i in good)
    proofs, st2 = engine_setup.blob_proof_many(gcat, gcomm, len(good))
    assert st2 == [0] * len(good)
    wantp = [oracle_setup.blob_proof(b, c) for b, c in zip(base, want)]
    for k, i in enumerate(good):
        assert proofs[48 * k:48 * k + 48] == wantp[i % len(base)], i
    assert engine_setup.verify_blob_proof_batch_concat(gcat, gcomm, proofs, len(good)) is True
    # proof and verify calls of this size upload the blobs in column pieces and hash each piece as it lands (h2d_blobs_hashed):
    # a flipped low byte in any piece of any blob must change that blob's challenge, hence the verdict
    for item, byte in ((0, 31), (700, 32768 + 63), (1249, 65536 + 95), (3, 131071)):
        gb = bytearray(gcat); gb[131072 * item + byte] ^= 1
        assert engine_setup.verify_blob_proof_batch_concat(bytes(gb), gcomm, proofs, len(good)) is False, (item, byte)
    # a bad commitment in the tail is reported in the status of its own item only
    bc = bytearray(gcomm); bc[48 * 1000:48 * 1000 + 48] = bytes([0xff] * 48)
    _, st3 = engine_setup.blob_proof_many(gcat, bytes(bc), len(good))
    assert [k for k in range(len(good)) if st3[k] != 0] == [1000]


def test_txt_setup_gives_the_same_engine(tmp_path, gpu_free):
    import json
    import kateth_b200 as K
    d = json.load(open(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json")))
    p = tmp_path / "trusted_setup.txt"
    p.write_text("\n".join(["4096", "65"] + [x[2:] for x in d["g1_lagrange"]] + [x[2:] for x in d["g2_monomial"]]) + "\n")
    s = K.Setup.load_txt(str(p), window_bits=9)
    try:
        vb = RV.valid_blobs()
        c = s.blob_to_commitment(vb[2])
        assert c.hex() == GOLD_COMM[2]
        pr = s.blob_proof(vb[2], c)
        assert s.verify_blob_proof(vb[2], c, pr) is True
    finally:
        s.close()


def test_full_size_batch_checksum_of_checksums(engine_setup, oracle_setup):
    """BASELINE configs[2] at its full size, 4096 DISTINCT blobs (bench.py's generator), through the host-buffer C ABI, pinned by
    size-independent properties: (1) linearity -- the sum of the 4096 commitments is the commitment of the element-wise sum
    of the blobs mod r (a checksum of checksums: any wrong MSM in the batch breaks it); (2) commit -> prove -> batch-verify
    round trip is true, and false once one proof or one blob byte is changed; (3) a seeded sample equals the oracle."""
    from oracle import oracle as O
    sys.path.insert(0, ROOT)
    import bench
    n = 4096
    flat = bench.synth_blobs(n, 0x4b5a4700 + 77)
    cat = flat.tobytes()
    engine_setup.set_msm_mode(0)
    comm, st = engine_setup.blob_to_commitment_many(cat, n)
    assert st == [0] * n and engine_setup.info()["last_msm_batch_affine"] == 1
    # (1) element-wise sum of the blobs: per 32-bit big-endian limb in uint64 (4096 x 2^32 < 2^64), then carried in Python
    limbs = flat.view(">u4").reshape(n, 4096, 8).astype(np.uint64).sum(axis=0)
    sum_blob = b"".join((sum(int(limbs[e, k]) << (32 * (7 - k)) for k in range(8)) % R).to_bytes(32, "big") for e in range(4096))
    cs = [comm[48 * i:48 * i + 48] for i in range(n)]
    assert len(set(cs)) == n
    assert O.g1_lincomb(cs, [(1).to_bytes(32, "big")] * n) == engine_setup.blob_to_commitment(sum_blob) == oracle_setup.blob_to_commitment(sum_blob)
    # (2) round trip
    proofs, st2 = engine_setup.blob_proof_many(cat, comm, n)
    assert st2 == [0] * n
    assert engine_setup.verify_blob_proof_batch_concat(cat, comm, proofs, n) is True
    bad = bytearray(proofs); bad[48 * 4095:48 * 4096] = proofs[0:48]
    assert engine_setup.verify_blob_proof_batch_concat(cat, comm, bytes(bad), n) is False
    gb = bytearray(cat); gb[131072 * 2048 + 131071] ^= 1
    assert engine_setup.verify_blob_proof_batch_concat(bytes(gb), comm, proofs, n) is False
    # (3) sample against the oracle: every position of the persistent loop of k_msm_ba (592 resident CTAs pull items from a counter:
    # first wave, wave boundaries, later waves, the last item)
    assert engine_setup.info()["ba_grid"] == 592
    for i in (0, 1, 591, 592, 593, 1183, 1184, 1775, 1776, 2047, 2368, 2959, 2960, 3551, 3552, 4094, 4095):
        b = cat[131072 * i:131072 * (i + 1)]
        assert cs[i] == oracle_setup.blob_to_commitment(b), i
        assert proofs[48 * i:48 * i + 48] == oracle_setup.blob_proof(b, cs[i]), i


# ---------------------------------------------------------------- sharded batches: errors travel with the records, empty shards
def test_sharded_errors_reach_every_rank(engine_setup, oracle_setup):
    """SURVEY 8e + kateth's error order (src/kzg/setup.rs:259-271: all blobs, then all commitments, then all proofs).  A shard with
    an invalid item returns that error itself AND writes it into its record; whoever finishes the gathered records gets the
    error kateth would return for the whole batch, never a verdict computed from unvalidated terms."""
    import torch
    import kateth_b200 as K
    RB = K.RECORD_BYTES
    n = 7
    blobs = [rand_blob(7000 + i) for i in range(n)]
    cat = b"".join(blobs)
    comm, _ = engine_setup.blob_to_commitment_many(cat, n)
    proofs, _ = engine_setup.blob_proof_many(cat, comm, n)
    dev = torch.device("cuda:0")
    up = lambda b: torch.frombuffer(bytearray(b), dtype=torch.uint8).to(dev)
    bounds = [(0, 3), (3, 5), (5, 7)]

    def run(cat_, comm_, proofs_, bounds_=bounds, total=n):
        d_b, d_c, d_p = up(cat_), up(comm_), up(proofs_)
        recs = torch.zeros(len(bounds_) * RB, dtype=torch.uint8, device=dev)
        st = torch.zeros(max(total, 1), dtype=torch.int32, device=dev)
        rcs = []
        for g, (lo, hi) in enumerate(bounds_):
            rcs.append(engine_setup.verify_partial_device(d_b.data_ptr() + lo * 131072, d_c.data_ptr() + lo * 48, d_p.data_ptr() + lo * 48, hi - lo, lo, total,
                                                          recs.data_ptr() + g * RB, st.data_ptr() + 4 * lo))
        ok = __import__("ctypes").c_int(-1)
        rc = K.lib().kb200_verify_batch_finish_device(engine_setup._h, __import__("ctypes").c_void_p(recs.data_ptr()), len(bounds_), __import__("ctypes").byref(ok))
        return rcs, rc, ok.value, st.cpu().tolist()

    assert run(cat, comm, proofs) == ([0, 0, 0], 0, 1, [0] * n)
    # a proof that is not a curve point in shard 1, a non-canonical blob in shard 2: every finisher reports the BLOB error
    badp = bytearray(proofs); badp[48 * 4:48 * 5] = RV.INVALID_G1[3]
    badb = bytearray(cat); badb[131072 * 6 + 32 * 9:131072 * 6 + 32 * 10] = RV.be32(R)
    rcs, rc, ok, st = run(bytes(badb), comm, bytes(badp))
    assert rcs == [0, 4, 2] and rc == 2 and ok == 0 and st == [0, 0, 0, 0, 4, 0, 2]
    # only the bad proof: its class; and the single-call entry points agree
    rcs, rc, ok, st = run(cat, comm, bytes(badp))
    assert rcs == [0, 4, 0] and rc == 4 and ok == 0
    with pytest.raises(K.KzgError) as e:
        engine_setup.verify_blob_proof_batch_concat(cat, comm, bytes(badp), n)
    assert e.value.status == 4
    # a commitment outside G1 in shard 0 and the bad proof in shard 1: commitments come before proofs
    badc = bytearray(comm); badc[0:48] = RV.INVALID_G1[2]
    rcs, rc, ok, st = run(cat, bytes(badc), bytes(badp))
    assert rcs == [5, 4, 0] and rc == 5
    # the failing item is left out of the sums: dropping it from the batch by hand gives the same operands as ... nothing to
    # compare (an error has no verdict), but a WRONG proof (valid point) still only flips the verdict
    wrong = bytearray(proofs); wrong[48 * 4:48 * 5] = proofs[0:48]
    assert run(cat, comm, bytes(wrong))[1:3] == (0, 0)
    # empty shards: a batch of one over three ranks
    rcs, rc, ok, st = run(cat[:131072], comm[:48], proofs[:48], [(0, 0), (0, 1), (1, 1)], 1)
    assert rcs == [0, 0, 0] and rc == 0 and ok == 1
    assert engine_setup.last_pairing_operands() == oracle_setup.verify_blob_proof_batch_sums(blobs[:1], [comm[:48]], [proofs[:48]])


def test_batches_larger_than_a_chunk(engine_setup, oracle_setup):
    """The host-buffer entry points work in chunks of 4096 blobs: outputs, statuses, error order and the batch verdict must not
    depend on where the chunk boundary falls (4100 blobs: the second chunk holds 4), in either batch-challenge mode."""
    import torch
    import kateth_b200 as K
    n = 4100
    base = [rand_blob(8000 + i) for i in range(3)] + [b"".join(RV.fe(R - 2 - k) for k in range(4096))]
    want_c = [oracle_setup.blob_to_commitment(b) for b in base]
    want_p = [oracle_setup.blob_proof(b, c) for b, c in zip(base, want_c)]
    idx = [(i * 7 + i // 5) % len(base) for i in range(n)]
    bad_at = 4098                                      # second chunk
    blobs = [base[k] for k in idx]
    b = bytearray(blobs[bad_at]); b[32 * 17:32 * 18] = RV.be32(R + 1); blobs[bad_at] = bytes(b)
    cat = b"".join(blobs)
    comm, st = engine_setup.blob_to_commitment_many(cat, n)
    assert [i for i in range(n) if st[i]] == [bad_at] and st[bad_at] == 2
    assert all(comm[48 * i:48 * i + 48] == want_c[idx[i]] for i in range(n) if i != bad_at)
    good = [i for i in range(n) if i != bad_at] + [0]                # 4100 again, all valid; the chunk boundary moved by one
    gcat = b"".join(blobs[i] for i in good); gcomm = b"".join(want_c[idx[i]] for i in good)
    proofs, st2 = engine_setup.blob_proof_many(gcat, gcomm, n)
    assert st2 == [0] * n and all(proofs[48 * k:48 * k + 48] == want_p[idx[i]] for k, i in enumerate(good))
    zs = b"".join(RV.fe(int.from_bytes(hashlib.sha256(b"zc%d" % (i % 5)).digest(), "big")) for i in range(n))
    pr, ys, st3 = engine_setup.proof_many(gcat, zs, n)
    assert st3 == [0] * n
    for k in (0, 4095, 4096, 4099):
        assert (pr[48 * k:48 * k + 48], ys[32 * k:32 * k + 32]) == oracle_setup.proof(blobs[good[k]], zs[32 * k:32 * k + 32]), k
    for mode in (0, 1):
        engine_setup.set_batch_challenge(mode)
        try:
            assert engine_setup.verify_blob_proof_batch_concat(gcat, gcomm, proofs, n) is True
            if mode == 0:
                single = engine_setup.last_pairing_operands()
            wrong = bytearray(proofs); wrong[48 * 4097:48 * 4098] = proofs[48 * 4098:48 * 4099] if proofs[48 * 4097:48 * 4098] != proofs[48 * 4098:48 * 4099] else proofs[48 * 4099:48 * 4100]
            assert engine_setup.verify_blob_proof_batch_concat(gcat, gcomm, bytes(wrong), n) is False
            # error order across chunks: a bad proof in chunk 0 and a bad blob in chunk 1 -> the blob error
            badp = bytearray(proofs); badp[0:48] = RV.INVALID_G1[3]
            bb = bytearray(gcat); bb[131072 * 4097 + 64:131072 * 4097 + 96] = RV.be32(R)
            with pytest.raises(K.KzgError) as e:
                engine_setup.verify_blob_proof_batch_concat(bytes(bb), gcomm, bytes(badp), n)
            assert e.value.status == 2
            with pytest.raises(K.KzgError) as e:
                engine_setup.verify_blob_proof_batch_concat(gcat, gcomm, bytes(badp), n)
            assert e.value.status == 4
        finally:
            engine_setup.set_batch_challenge(0)
    # commit + prove + verify over ONE upload (kb200_group_commit_prove_verify, a group of one around this handle): 4100 blobs are two
    # chunks, taken last-first inside the library -- same bytes, same verdict, same operands; a bad blob in the second chunk is the error
    g = K.Group.join(engine_setup, 0, 1, None)
    try:
        oc, op, ok = g.commit_prove_verify(gcat, n)
        assert ok is True and oc == gcomm and op == proofs and engine_setup.last_pairing_operands() == single
        with pytest.raises(K.KzgError) as e:
            g.commit_prove_verify(bytes(bb), n)
        assert e.value.status == 2
    finally:
        g.close()
    # the host-buffer shard call has no size limit either: one "rank" holding all 4100 blobs, finished from its record
    rec = torch.zeros(K.RECORD_BYTES, dtype=torch.uint8, device="cuda")
    assert engine_setup.verify_partial_host(gcat, gcomm, proofs, n, 0, n, rec.data_ptr()) == 0
    assert engine_setup.verify_finish_device(rec.data_ptr(), 1) is True
    assert engine_setup.last_pairing_operands() == single


def test_concurrent_callers_share_one_handle(engine_setup, oracle_setup):
    """kateth's Setup is Send + Sync and its tests share it through an Arc (src/kzg/setup.rs:323): four host threads hammer ONE
    handle with commit / blob-proof / proof-at-z / verify calls on their own blobs; every output must equal the oracle's."""
    import threading
    blobs = [rand_blob(9000 + i) for i in range(4)]
    want = []
    for b in blobs:
        c = oracle_setup.blob_to_commitment(b)
        want.append((c, oracle_setup.blob_proof(b, c)))
    errors = []

    def worker(t):
        try:
            b, (c, p) = blobs[t], want[t]
            for it in range(6):
                assert engine_setup.blob_to_commitment(b) == c
                assert engine_setup.blob_proof(b, c) == p
                assert engine_setup.verify_blob_proof(b, c, p) is True
                assert engine_setup.verify_blob_proof(b, c, want[(t + 1) % 4][1]) is False
                k = 2 + (t + it) % 3
                cat = b"".join(blobs[(t + j) % 4] for j in range(k))
                comm, st = engine_setup.blob_to_commitment_many(cat, k)
                assert st == [0] * k and comm == b"".join(want[(t + j) % 4][0] for j in range(k))
                assert engine_setup.verify_blob_proof_batch_concat(cat, comm, b"".join(want[(t + j) % 4][1] for j in range(k)), k) is True
        except BaseException as ex:                      # noqa: B036 -- reported by the main thread
            errors.append((t, repr(ex)))

    th = [threading.Thread(target=worker, args=(t,)) for t in range(4)]
    for x in th: x.start()
    for x in th: x.join()
    assert not errors, errors


def test_first_error_follows_blob_order(engine_setup):
    """kateth decodes blob by blob (src/kzg/setup.rs:259-262): [blob with a non-canonical element, blob one byte short] is
    InvalidFieldElement, the reverse order is InvalidLen."""
    import kateth_b200 as K
    good = rand_blob(9100)
    bad_el = bytearray(good); bad_el[32 * 5:32 * 6] = RV.be32(R); bad_el = bytes(bad_el)
    short = good[:-1]
    c = engine_setup.blob_to_commitment(good); p = engine_setup.blob_proof(good, c)
    for blobs, want in (([good, bad_el, short], 2), ([good, short, bad_el], 1), ([short], 1)):
        with pytest.raises(K.KzgError) as e:
            engine_setup.verify_blob_proof_batch(blobs, [c] * len(blobs), [p] * len(blobs))
        assert e.value.status == want


# ---------------------------------------------------------------- per-item verdicts from the device pairing
def test_device_pairing_matches_host_pairing_on_the_verify_vectors(engine_setup):
    """kb200_verify_proof_many runs the two-pairing check on the GPU, one item per thread (pairing.cuh).  Its verdicts and statuses must
    equal Setup::verify_proof's -- the host pairing, itself pinned by the 122 sha-checked verify_kzg_proof vectors -- on exactly
    those cases: correct and incorrect proofs for 7 blobs x 6 points (3 of them in the domain), proofs at infinity for the zero and
    the constant polynomial, invalid commitments / proofs / z / y."""
    import kateth_b200 as K
    from oracle import oracle as O
    s = engine_setup
    VB = RV.valid_blobs()
    comm = [s.blob_to_commitment(b) for b in VB]
    items = []          # (proof, commitment, z, y)
    for bi, b in enumerate(VB):
        for z in RV.FE_VALID:
            pr, y = s.proof(b, z)
            items.append((pr, comm[bi], z, y))
            items.append((O.g1_add(pr, RV.G1), comm[bi], z, y))
    for z in RV.FE_VALID:
        _, y = s.proof(VB[2], z)
        items.append((RV.INF, comm[2], z, y))                 # incorrect: proof at infinity
        items.append((RV.INF, RV.INF, z, RV.fe(0)))           # correct: zero polynomial
        items.append((RV.INF, comm[1], z, RV.fe(2)))          # correct: constant polynomial
    pr1, y1 = s.proof(VB[2], RV.FE_VALID[1])
    for g in RV.INVALID_G1[2:]:                               # full-length encodings (the shorter ones never reach the ABI)
        items.append((pr1, g, RV.FE_VALID[1], y1)); items.append((g, comm[2], RV.FE_VALID[1], y1))
    for f in RV.FE_INVALID[:4]:
        items.append((pr1, comm[2], f, y1)); items.append((pr1, comm[2], RV.FE_VALID[1], f))
    n = len(items)
    ok, st = s.verify_proof_many(b"".join(i[0] for i in items), b"".join(i[1] for i in items), b"".join(i[2] for i in items), b"".join(i[3] for i in items), n)
    seen = {True: 0, False: 0, "err": 0}
    for k, (p, c, z, y) in enumerate(items):
        try:
            want = s.verify_proof(p, c, z, y)
            assert st[k] == 0 and ok[k] is want, (k, ok[k], want)
            seen[want] += 1
        except K.KzgError as e:
            assert st[k] == e.status and ok[k] is False, (k, st[k], e.status)
            seen["err"] += 1
    assert seen[True] >= 42 + 12 and seen[False] >= 42 + 6 and seen["err"] == 12


def test_device_pairing_at_scale_with_planted_failures(engine_setup):
    """4,000 (proof, commitment, z, y) tuples built from real proofs with failures planted at seeded positions (a wrong y, a proof of
    another item, a swapped commitment): every per-item verdict must be the planted one, and the RLC batch check over the good items
    (one host pairing) must agree."""
    s = engine_setup
    rng = random.Random(99)
    nb, nz = 4, 25
    blobs = [rand_blob(14000 + i) for i in range(nb)]
    cat = b"".join(blobs)
    comm, _ = s.blob_to_commitment_many(cat, nb)
    base = []
    for j in range(nz):
        zs = b"".join(RV.fe(rng.randrange(R)) for _ in range(nb))
        pr, ys, st = s.proof_many(cat, zs, nb)
        assert st == [0] * nb
        for i in range(nb):
            base.append((pr[48 * i:48 * i + 48], comm[48 * i:48 * i + 48], zs[32 * i:32 * i + 32], ys[32 * i:32 * i + 32]))
    n = 4000
    items, want = [], []
    for k in range(n):
        p, c, z, y = base[rng.randrange(len(base))]
        kind = rng.randrange(8)
        if kind == 0:
            y = RV.fe(int.from_bytes(y, "big") + 1); want.append(False)
        elif kind == 1:
            q = base[rng.randrange(len(base))]
            want.append(q[0] == p and True); p = q[0]
        elif kind == 2:
            q = base[rng.randrange(len(base))]
            want.append(q[1] == c); c = q[1]
        else:
            want.append(True)
        items.append((p, c, z, y))
    cols = [b"".join(i[j] for i in items) for j in range(4)]
    ok, st = s.verify_proof_many(cols[0], cols[1], cols[2], cols[3], n)
    assert st == [0] * n
    assert ok == want, [k for k in range(n) if ok[k] != want[k]][:10]
    good = [k for k in range(n) if want[k]]
    gcols = [b"".join(items[k][j] for k in good) for j in range(4)]
    assert s.verify_proof_batch(gcols[0], gcols[1], gcols[2], gcols[3], len(good)) is True
    assert s.verify_proof_batch(cols[0], cols[1], cols[2], cols[3], n) is False


def test_sharded_point_batches_equal_single_call(engine_setup):
    """Setup::verify_proof_batch over shards (config 5's sweep is cut over the GPUs): records of contiguous shards with global
    index offsets, finished together, give the operands and the verdict of the single call; a shard's decoding error reaches the
    finisher in kateth's class order (proof, commitment, z, y)."""
    import ctypes as C
    import torch
    import kateth_b200 as K
    s = engine_setup
    n = 9
    blobs = [rand_blob(15000 + i) for i in range(3)]
    cat = b"".join(blobs[i % 3] for i in range(n))
    comm, _ = s.blob_to_commitment_many(cat, n)
    zs = b"".join(RV.fe(int.from_bytes(hashlib.sha256(b"sp%d" % i).digest(), "big")) for i in range(n))
    pr, ys, st = s.proof_many(cat, zs, n)
    assert st == [0] * n and s.verify_proof_batch(pr, comm, zs, ys, n) is True
    single = s.last_pairing_operands()
    RB = K.RECORD_BYTES
    recs = torch.zeros(3 * RB, dtype=torch.uint8, device="cuda")

    def run(pr_, comm_, zs_, ys_):
        rcs = []
        for g, (lo, hi) in enumerate([(0, 4), (4, 4), (4, 9)]):
            rcs.append(s.verify_proof_partial_host(pr_[48 * lo:48 * hi], comm_[48 * lo:48 * hi], zs_[32 * lo:32 * hi], ys_[32 * lo:32 * hi], hi - lo, lo, n,
                                                   recs.data_ptr() + g * RB))
        ok = C.c_int(-1)
        rc = K.lib().kb200_verify_batch_finish_device(s._h, C.c_void_p(recs.data_ptr()), 3, C.byref(ok))
        return rcs, rc, ok.value

    assert run(pr, comm, zs, ys) == ([0, 0, 0], 0, 1) and s.last_pairing_operands() == single
    bad_y = bytearray(ys); bad_y[32 * 7:32 * 8] = RV.fe(int.from_bytes(ys[32 * 7:32 * 8], "big") + 1)
    assert run(pr, comm, zs, bytes(bad_y)) == ([0, 0, 0], 0, 0)
    # a non-canonical y in shard 0 and a proof off the curve in shard 2: the proof class comes first
    ny = bytearray(ys); ny[0:32] = RV.be32(R)
    bp = bytearray(pr); bp[48 * 8:48 * 9] = RV.INVALID_G1[3]
    assert run(bytes(bp), comm, zs, bytes(ny)) == ([7, 0, 4], 4, 0)
    assert run(pr, comm, zs, bytes(ny)) == ([7, 0, 0], 7, 0)


def test_bucket_sums_equal_ladder_sums(engine_setup):
    """The two forms of a shard's sums (one 128-step ladder per term, k_rlc; bucket method over the batch, k_rlc_bucket) give the same
    336 record bytes: blob batches over 1 and 2 slices with a partial slice, items repeated hundreds of times (a bucket adds a point to
    itself: the doubling path of the mixed addition), points at infinity (the zero blob), an invalid item (left out of the sums, error in
    the record), a wrong proof (verdict false), a point batch (weights from z), and the single-proof form (weights 1)."""
    import torch
    import kateth_b200 as K
    s = engine_setup
    RB = K.RECORD_BYTES
    base = [rand_blob(17000 + i) for i in range(5)] + [bytes(131072)]
    try:
        for n in (700, 1500):
            cat = b"".join(base[(i * i) % 6] for i in range(n))
            comm, _ = s.blob_to_commitment_many(cat, n)
            pr, st = s.blob_proof_many(cat, comm, n)
            assert st == [0] * n
            d_b = torch.frombuffer(bytearray(cat), dtype=torch.uint8).cuda()
            rec = torch.zeros(2 * RB, dtype=torch.uint8, device="cuda"); d_st = torch.zeros(n, dtype=torch.int32, device="cuda")

            def records(comm_, pr_, off=3, total=n + 5):
                d_c = torch.frombuffer(bytearray(comm_), dtype=torch.uint8).cuda(); d_p = torch.frombuffer(bytearray(pr_), dtype=torch.uint8).cuda()
                out = []
                for g, mode in enumerate((1, 2)):
                    s.set_rlc_mode(mode)
                    out.append(s.verify_partial_device(d_b.data_ptr(), d_c.data_ptr(), d_p.data_ptr(), n, off, total, rec.data_ptr() + g * RB, d_st.data_ptr()))
                r = bytes(rec.cpu().numpy())
                return out, r[:RB], r[RB:]

            rcs, ladder, bucket = records(comm, pr)
            assert rcs == [0, 0] and ladder == bucket and any(ladder[:288])
            rcs, ladder, bucket = records(comm, pr, 0, n)                      # the whole batch: index 0 takes the pow(r, 0) == r weight
            assert rcs == [0, 0] and ladder == bucket
            assert s.verify_finish_device(rec.data_ptr() + RB, 1) is True
            swapped = bytearray(pr); swapped[0:48], swapped[48:96] = pr[48:96], pr[0:48]
            rcs, ladder, bucket = records(comm, bytes(swapped), 0, n)
            assert rcs == [0, 0] and ladder == bucket
            assert s.verify_finish_device(rec.data_ptr() + RB, 1) is False
            badc = bytearray(comm); badc[48 * 611:48 * 612] = RV.INVALID_G1[3]
            rcs, ladder, bucket = records(bytes(badc), pr)
            assert rcs[0] == rcs[1] != 0 and ladder == bucket
        # point batches (Setup::verify_proof_batch) and the single check
        n = 600
        cat = b"".join(base[i % 6] for i in range(n))
        comm, _ = s.blob_to_commitment_many(cat, n)
        zs = b"".join(RV.fe(int.from_bytes(hashlib.sha256(b"bk%d" % i).digest(), "big")) for i in range(n))
        pr, ys, st = s.proof_many(cat, zs, n)
        assert st == [0] * n
        ops = []
        for mode in (1, 2):
            s.set_rlc_mode(mode)
            assert s.verify_proof_batch(pr, comm, zs, ys, n) is True
            ops.append(s.last_pairing_operands())
            assert s.verify_proof(pr[:48], comm[:48], zs[:32], ys[:32]) is True
            ops.append(s.last_pairing_operands())
        assert ops[0] == ops[2] and ops[1] == ops[3]
        # the consensus-spec challenge (weights 1, r, r^2, ... from the transcript) through both forms
        pr, st = s.blob_proof_many(cat, comm, n)
        s.set_batch_challenge(1)
        spec_ops = []
        for mode in (1, 2):
            s.set_rlc_mode(mode)
            assert s.verify_blob_proof_batch_concat(cat, comm, pr, n) is True
            spec_ops.append(s.last_pairing_operands())
        assert spec_ops[0] == spec_ops[1] and spec_ops[0] != ops[0]
    finally:
        s.set_rlc_mode(0)
        s.set_batch_challenge(0)


# ---------------------------------------------------------------- loader extras
def test_g1_monomial_cross_check(engine_setup, tmp_path):
    """Files that also carry the monomial basis are cross-checked: X^k over the domain, in blob order, must commit to [tau^k]G1 for all
    4096 k (one batch commit).  The mainnet ceremony's own points pass; two swapped points, or a monomial array from another setup, do not."""
    import json
    import kateth_b200 as K
    mono = [bytes.fromhex(x[2:]) for x in json.load(open(os.path.join(ROOT, "tests/golden/g1_monomial_4096.json")))["g1_monomial"]]
    assert mono[0] == RV.G1
    engine_setup.check_g1_monomial(mono)
    engine_setup.check_g1_monomial(mono[:5])
    bad = list(mono); bad[7], bad[4000] = bad[4000], bad[7]
    with pytest.raises(K.KzgError) as e:
        engine_setup.check_g1_monomial(bad)
    assert e.value.status == -12 and "g1_monomial[7]" in str(e.value)


def test_loaders_cross_check_g1_monomial_and_default_window(tmp_path, gpu_free):
    import json
    import kateth_b200 as K
    d = json.load(open(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json")))
    mono = json.load(open(os.path.join(ROOT, "tests/golden/g1_monomial_4096.json")))["g1_monomial"]
    full = tmp_path / "presets.json"
    full.write_text(json.dumps({"g1_monomial": mono, "g1_lagrange": d["g1_lagrange"], "g2_monomial": d["g2_monomial"]}))
    s = K.Setup.load_json(str(full), window_bits=9)          # parses, builds, cross-checks
    assert s.blob_to_commitment(RV.valid_blobs()[2]).hex() == GOLD_COMM[2]
    s.close()
    wrong = tmp_path / "wrong.json"
    wrong.write_text(json.dumps({"g1_monomial": mono[1:] + mono[:1], "g1_lagrange": d["g1_lagrange"], "g2_monomial": d["g2_monomial"]}))
    with pytest.raises(K.KzgError) as e:
        K.Setup.load_json(str(wrong), window_bits=9)
    assert e.value.status == -12
    txt = tmp_path / "trusted_setup.txt"
    txt.write_text("\n".join(["4096", "65"] + [x[2:] for x in d["g1_lagrange"]] + [x[2:] for x in d["g2_monomial"]] + [x[2:] for x in mono]) + "\n")
    s = K.Setup.load_txt(str(txt), window_bits=9); s.close()
    # window_bits 0: on an idle B200 the default is the throughput table (GLV, 16-bit windows, 100 GiB) -- what bench.py measures
    s = K.Setup.load_json(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json"))
    try:
        info = s.info()
        assert (info["window_bits"], info["glv"]) == (16, 1) and info["table_bytes"] > 100 * 2 ** 30
        # a second handle beside it no longer has 100 GiB free and steps down instead of failing
        s2 = K.Setup.load_json(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json"))
        assert s2.info()["window_bits"] in (13, 9) and s2.blob_to_commitment(RV.valid_blobs()[2]).hex() == GOLD_COMM[2]
        s2.close()
    finally:
        s.close()


def test_table_cache_round_trip(tmp_path, gpu_free):
    """On-disk cache of the window table: first load builds and writes, second load reads (checksum verified on the GPU), a flipped byte
    in the file is noticed and the table is rebuilt and rewritten.  Same bytes out in every case."""
    import kateth_b200 as K
    path = os.path.join(ROOT, "tests/golden/trusted_setup_4096.json")
    K.set_table_cache(str(tmp_path))
    try:
        outs = []
        for expect_cached in (0, 1):
            s = K.Setup.load_json(path, window_bits=106)          # GLV, 6-bit windows: a 0.26 GiB table
            assert s.info()["table_from_cache"] == expect_cached
            outs.append(s.blob_to_commitment(RV.valid_blobs()[2]))
            s.close()
        files = [f for f in os.listdir(tmp_path) if f.startswith("kb200_table_") and f.endswith(".bin")]
        assert len(files) == 1 and outs[0] == outs[1] and outs[0].hex() == GOLD_COMM[2]
        fn = os.path.join(tmp_path, files[0])
        with open(fn, "r+b") as f:
            f.seek(os.path.getsize(fn) // 2); b = f.read(1); f.seek(-1, 1); f.write(bytes([b[0] ^ 0x40]))
        s = K.Setup.load_json(path, window_bits=106)
        assert s.info()["table_from_cache"] == 0 and s.blob_to_commitment(RV.valid_blobs()[2]) == outs[0]
        s.close()
        s = K.Setup.load_json(path, window_bits=106)
        assert s.info()["table_from_cache"] == 1
        s.close()
        # another window form is another file
        s = K.Setup.load_json(path, window_bits=6)
        assert s.info()["table_from_cache"] == 0
        s.close()
        assert len([f for f in os.listdir(tmp_path) if f.endswith(".bin")]) == 2
    finally:
        K.set_table_cache(None)
