"""CPU tests of the oracle: it is pinned against every golden the reference's own tests hold for this path
(SURVEY.md 8c): the 253 consensus-spec-tests vectors via their Git-LFS sha256 oids, Fr::ONE / Fr::MAX
(/root/reference/src/bls.rs:87-106, 600-613) and the self-checks of SURVEY.md 8c."""
import hashlib, json, os, sys
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import reconstruct_vectors as RV
from oracle import oracle as O


def test_all_253_vectors_match_lfs_oids():
    counts, bad, missing = RV.check_all("oracle")
    assert sum(counts.values()) == 253
    assert counts == {"blob_to_kzg_commitment": 11, "compute_kzg_proof": 52, "compute_blob_kzg_proof": 15, "verify_kzg_proof": 122,
                      "verify_blob_kzg_proof": 29, "verify_blob_kzg_proof_batch": 24}
    assert not bad and not missing


def test_fr_one_and_max_limbs():
    one, mx = O.fr_consts()
    # the literal limbs in /root/reference/src/bls.rs:87-106
    assert one == [8589934590, 6378425256633387010, 11064306276430008309, 1739710354780652911]
    assert mx == [18446744060824649731, 18102478225614246908, 11073656695919314959, 6613806504683796440]


def test_sha256_matches_hashlib():
    for n in (0, 1, 55, 56, 63, 64, 65, 119, 120, 131152):
        d = bytes((i * 7 + n) & 0xff for i in range(n))
        assert O.sha256(d) == hashlib.sha256(d).digest()


def test_sum_of_lagrange_is_generator(oracle_setup):
    # commit(all-ones) = sum L_i = [1]G1
    blob = RV.fe(1) * 4096
    assert oracle_setup.blob_to_commitment(blob) == RV.G1


def test_single_nonzero_blob_picks_brp_point(oracle_setup):
    d = json.load(open(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json")))
    brp = int(format(3211, "012b")[::-1], 2)
    blob = RV.valid_blobs()[6]
    assert oracle_setup.blob_to_commitment(blob).hex() == d["g1_lagrange"][brp][2:]


def test_pairing_bilinearity():
    # e([a]G1, [b]G2) == e([ab]G1, G2)
    a, b = 0x1234567, 0x89abcdef01
    g2 = O.g2_mul_gen(RV.fe(1)); bg2 = O.g2_mul_gen(RV.fe(b))
    ag1 = O.g1_lincomb([RV.G1], [RV.fe(a)]); abg1 = O.g1_lincomb([RV.G1], [RV.fe(a * b)])
    assert O.pairing_eq(ag1, bg2, abg1, g2)
    assert not O.pairing_eq(ag1, bg2, O.g1_lincomb([RV.G1], [RV.fe(a * b + 1)]), g2)


def test_decompress_error_classes():
    assert O.g1_decompress_status(RV.G1) == 0
    assert O.g1_decompress_status(RV.INF) == 0
    assert O.g1_decompress_status(RV.INVALID_G1[2]) == 5   # on curve, not in G1
    assert O.g1_decompress_status(RV.INVALID_G1[3]) == 4   # not on curve
    assert O.g1_decompress_status(bytes([0x17]) + RV.G1[1:]) == 3   # compression bit clear
    assert O.g1_decompress_status(b"\xe0" + bytes(47)) == 3         # infinity with sign bit
