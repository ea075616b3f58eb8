#!/usr/bin/env python3
"""Rebuild the 253 consensus-spec-tests KZG `data.yaml` files that kateth's conformance tests read
(/root/reference/src/kzg/setup.rs:305-463) and check each against the sha256 oid of its Git-LFS stub.

The stubs in /root/reference hold only `oid sha256:<hex>` + `size`; the generator's inputs are deterministic
(SURVEY.md Appendix A), so a file rebuilt from OUR outputs that hashes to the oid proves inputs, outputs and
verdicts byte-exact.  Outputs come from a pluggable backend:
    --backend oracle   (default)  oracle/kzg_oracle.c through oracle/oracle.py       [CPU, test infrastructure]
    --backend engine              kateth_b200 (the CUDA engine through the C ABI)    [needs a GPU]
The oids are read from tests/golden/lfs_oids.json (made by tools/make_lfs_manifest.py), so this runs without
/root/reference.  With --write DIR the rebuilt files are materialised under DIR/<handler>/<case>/data.yaml.

`cases()` is also the single source of truth for the parametrised vector tests in tests/.
"""
import argparse, hashlib, json, os, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

R = 0x73eda753299d7d483339d80809a1d80553bda402fffe5bfeffffffff00000001
N = 4096
OMEGA = pow(7, (R - 1) // N, R)
G1 = bytes.fromhex("97f1d3a73197d7942695638c4fa9ac0fc3688c4f9774b905a14e3a3f171bac586c55e83ff97a1aeffb3af00adb22c6bb")
INF = b"\xc0" + bytes(47)
H = lambda b: hashlib.sha256(b).digest()
hx = lambda b: "0x" + bytes(b).hex()


def fe(x):
    return (x % R).to_bytes(32, "big")


def be32(x):
    return x.to_bytes(32, "big")


def valid_blobs():
    return [
        bytes(131072),
        fe(2) * N,
        b"".join(fe(pow(2, n + 256, R)) for n in range(N)),
        b"".join(fe(pow(3, n + 256, R)) for n in range(N)),
        b"".join(fe(pow(5, n + 256, R)) for n in range(N)),
        fe(R - 1) * N,
        b"".join(fe(1) if n == 3211 else fe(0) for n in range(N)),
    ]


def invalid_blobs():
    vb = valid_blobs()
    z = bytearray(131072); z[2111 * 32:2112 * 32] = be32(R)
    return [b"\xff" * 131072, bytes(z), vb[2] + b"\x00", vb[2][:-1]]


FE_VALID = [fe(0), fe(1), fe(2), fe(pow(5, 1235, R)), fe(R - 1), fe(OMEGA)]
FE_INVALID = [be32(R), be32(R + 1), be32(2 ** 256 - 1), be32(2 ** 256 - 2 ** 128), fe(0) + b"\x00", fe(0)[:31]]
_NOTG1 = bytes.fromhex("8123456789abcdef0123456789abcdef0123456789abcdef0123456789abcdef0123456789abcdef0123456789abcdef")
INVALID_G1 = [G1[:47], G1 + b"\x00", _NOTG1, _NOTG1[:-1] + b"\xe0"]


def suffix(identifier):
    return H(identifier.encode("utf-8"))[:8].hex()


# ---------------------------------------------------------------- YAML emitter (Appendix A rule)
def _flow(prefix, items, indent, open_c, close_c):
    out = prefix + open_c
    col = len(out)
    for k, it in enumerate(items):
        if k:
            out += ","; col += 1
            if col > 80:
                out += "\n" + " " * indent; col = indent
            else:
                out += " "; col += 1
        out += it; col += len(it)
    return out + close_c + "\n"


def q(b):
    return "'" + hx(b) + "'"


def yaml_scalar(v):
    if v is None:
        return "null"
    if v is True:
        return "true"
    if v is False:
        return "false"
    return q(v)


def emit_simple(inp, output):
    """inp: ordered list of (key, bytes); output: None | bool | bytes | [bytes, bytes]."""
    s = _flow("input: ", ["%s: %s" % (k, q(v)) for k, v in inp], 2, "{", "}")
    if isinstance(output, (list, tuple)):
        s += _flow("output: ", [q(v) for v in output], 2, "[", "]")
    else:
        s += "output: " + yaml_scalar(output) + "\n"
    return s.encode()


def emit_batch(blobs, commitments, proofs, output):
    s = "input:\n"
    for key, lst in (("blobs", blobs), ("commitments", commitments), ("proofs", proofs)):
        s += _flow("  %s: " % key, [q(v) for v in lst], 4, "[", "]")
    s += "output: " + yaml_scalar(output) + "\n"
    return s.encode()


# ---------------------------------------------------------------- backends
class Err(Exception):
    pass


def get_backend(name, setup=None):
    if name == "engine" and setup is not None:       # a live engine handed in by the caller (tests/: one engine at a time fits HBM)
        import kateth_b200 as K
        from oracle import oracle as O
        return setup, K.KzgError, O.g1_add
    if name == "oracle":
        from oracle import oracle as O
        s = O.Setup.load_json(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json"))
        return s, O.OracleError, O.g1_add
    if name == "engine":
        import kateth_b200 as K
        from oracle import oracle as O  # only add_one() of the *generator recipe* (an input, not a checked output)
        s = K.Setup.load_json(os.path.join(ROOT, "tests/golden/trusted_setup_4096.json"), window_bits=int(os.environ.get("KB200_WINDOW_BITS", "0")))
        return s, K.KzgError, O.g1_add
    raise SystemExit("unknown backend")


def _try(errcls, fn, *a):
    """kateth harness semantics (src/kzg/setup.rs:330-342): Err => expected output is null."""
    try:
        return fn(*a)
    except errcls:
        return None


def cases(setup, errcls, g1_add):
    """Yield (handler, case_name, file_bytes) for all 253 cases, computing outputs with `setup`."""
    VB, IB = valid_blobs(), invalid_blobs()
    comm = [setup.blob_to_commitment(b) for b in VB]
    bproof = [setup.blob_proof(b, c) for b, c in zip(VB, comm)]
    add_one = lambda p: g1_add(p, G1)

    # blob_to_kzg_commitment
    h = "blob_to_kzg_commitment"
    for b, c in zip(VB, comm):
        yield h, "blob_to_kzg_commitment_case_valid_blob_" + suffix(hx(H(b))), emit_simple([("blob", b)], c)
    for b in IB:
        yield h, "blob_to_kzg_commitment_case_invalid_blob_" + suffix(hx(H(b))), emit_simple([("blob", b)], _try(errcls, setup.blob_to_commitment, b))

    # compute_kzg_proof
    h = "compute_kzg_proof"
    kproof = {}
    for bi, b in enumerate(VB):
        for zi, z in enumerate(FE_VALID):
            pr, y = setup.proof(b, z); kproof[(bi, zi)] = (pr, y)
            yield h, "compute_kzg_proof_case_valid_blob_" + suffix(hx(H(b)) + "_" + hx(z)), emit_simple([("blob", b), ("z", z)], [pr, y])
    for b in IB:
        r = _try(errcls, setup.proof, b, FE_VALID[0])
        yield h, "compute_kzg_proof_case_invalid_blob_" + suffix(hx(H(b))), emit_simple([("blob", b), ("z", FE_VALID[0])], list(r) if r else None)
    for z in FE_INVALID:
        r = _try(errcls, setup.proof, VB[4], z)
        yield h, "compute_kzg_proof_case_invalid_z_" + suffix(hx(H(z))), emit_simple([("blob", VB[4]), ("z", z)], list(r) if r else None)

    # compute_blob_kzg_proof
    h = "compute_blob_kzg_proof"
    for b, c, p in zip(VB, comm, bproof):
        yield h, "compute_blob_kzg_proof_case_valid_blob_" + suffix(hx(H(b))), emit_simple([("blob", b), ("commitment", c)], p)
    for b in IB:
        yield h, "compute_blob_kzg_proof_case_invalid_blob_" + suffix(hx(H(b))), emit_simple([("blob", b), ("commitment", G1)], _try(errcls, setup.blob_proof, b, G1))
    for c in INVALID_G1:
        yield h, "compute_blob_kzg_proof_case_invalid_commitment_" + suffix(hx(H(c))), emit_simple([("blob", VB[1]), ("commitment", c)], _try(errcls, setup.blob_proof, VB[1], c))

    # verify_kzg_proof   (yaml keys: commitment, z, y, proof;  kateth call order: proof, commitment, z, y)
    h = "verify_kzg_proof"
    def vk(c, z, y, p):
        return emit_simple([("commitment", c), ("z", z), ("y", y), ("proof", p)], _try(errcls, setup.verify_proof, p, c, z, y))
    for bi, b in enumerate(VB):
        for zi, z in enumerate(FE_VALID):
            pr, y = kproof[(bi, zi)]
            yield h, "verify_kzg_proof_case_correct_proof_" + suffix(hx(H(b)) + "_" + hx(z)), vk(comm[bi], z, y, pr)
    for bi, b in enumerate(VB):
        for zi, z in enumerate(FE_VALID):
            pr, y = kproof[(bi, zi)]
            yield h, "verify_kzg_proof_case_incorrect_proof_" + suffix(hx(H(b)) + "_" + hx(z)), vk(comm[bi], z, y, add_one(pr))
    for zi, z in enumerate(FE_VALID):
        _, y = kproof[(2, zi)]
        yield h, "verify_kzg_proof_case_incorrect_proof_point_at_infinity_" + suffix(hx(H(VB[2])) + "_" + hx(z)), vk(comm[2], z, y, INF)
    for zi, z in enumerate(FE_VALID):
        yield h, "verify_kzg_proof_case_correct_proof_point_at_infinity_for_zero_poly_" + suffix(hx(H(VB[0])) + "_" + hx(z)), vk(INF, z, fe(0), INF)
    for zi, z in enumerate(FE_VALID):
        yield h, "verify_kzg_proof_case_correct_proof_point_at_infinity_for_twos_poly_" + suffix(hx(H(VB[1])) + "_" + hx(z)), vk(comm[1], z, fe(2), INF)
    pr1, y1 = kproof[(2, 1)]
    for c in INVALID_G1:
        yield h, "verify_kzg_proof_case_invalid_commitment_" + suffix(hx(c)), vk(c, FE_VALID[1], y1, pr1)
    for p in INVALID_G1:
        yield h, "verify_kzg_proof_case_invalid_proof_" + suffix(hx(p)), vk(comm[2], FE_VALID[1], y1, p)
    pr4, y4 = kproof[(4, 1)]
    for z in FE_INVALID:
        yield h, "verify_kzg_proof_case_invalid_z_" + suffix(hx(z)), vk(comm[4], z, y4, pr4)
    for y in FE_INVALID:
        yield h, "verify_kzg_proof_case_invalid_y_" + suffix(hx(y)), vk(comm[4], FE_VALID[1], y, pr4)

    # verify_blob_kzg_proof
    h = "verify_blob_kzg_proof"
    def vb(b, c, p):
        return emit_simple([("blob", b), ("commitment", c), ("proof", p)], _try(errcls, setup.verify_blob_proof, b, c, p))
    for b, c, p in zip(VB, comm, bproof):
        yield h, "verify_blob_kzg_proof_case_correct_proof_" + suffix(hx(H(b))), vb(b, c, p)
    for b, c, p in zip(VB, comm, bproof):
        yield h, "verify_blob_kzg_proof_case_incorrect_proof_" + suffix(hx(H(b))), vb(b, c, add_one(p))
    yield h, "verify_blob_kzg_proof_case_incorrect_proof_point_at_infinity", vb(VB[2], comm[2], INF)
    yield h, "verify_blob_kzg_proof_case_correct_proof_point_at_infinity_for_zero_poly", vb(VB[0], INF, INF)
    yield h, "verify_blob_kzg_proof_case_correct_proof_point_at_infinity_for_twos_poly", vb(VB[1], comm[1], INF)
    for b in IB:
        yield h, "verify_blob_kzg_proof_case_invalid_blob_" + suffix(hx(H(b))), vb(b, G1, G1)
    for c in INVALID_G1:
        yield h, "verify_blob_kzg_proof_case_invalid_commitment_" + suffix(hx(H(c))), vb(VB[1], c, G1)
    for p in INVALID_G1:
        yield h, "verify_blob_kzg_proof_case_invalid_proof_" + suffix(hx(H(p))), vb(VB[1], G1, p)

    # verify_blob_kzg_proof_batch
    h = "verify_blob_kzg_proof_batch"
    def vbb(bs, cs, ps):
        if not (len(bs) == len(cs) == len(ps)):
            out = None   # harness maps mismatched lengths to a None input => expected null (src/kzg/spec.rs:208-215)
        else:
            out = _try(errcls, setup.verify_blob_proof_batch, bs, cs, ps)
        return emit_batch(bs, cs, ps, out)
    for i in range(len(VB)):
        yield h, "verify_blob_kzg_proof_batch_case_" + suffix(hx(H(b"".join(VB[:i])))), vbb(VB[:i], comm[:i], bproof[:i])
    yield h, "verify_blob_kzg_proof_batch_case_incorrect_proof_add_one", vbb(VB, comm, [add_one(bproof[0])] + bproof[1:])
    yield h, "verify_blob_kzg_proof_batch_case_incorrect_proof_point_at_infinity", vbb([VB[2]], [comm[2]], [INF])
    for b in IB:
        bs = list(VB); bs[4] = b
        yield h, "verify_blob_kzg_proof_batch_case_invalid_blob_" + suffix(hx(H(b))), vbb(bs, comm, bproof)
    for c in INVALID_G1:
        yield h, "verify_blob_kzg_proof_batch_case_invalid_commitment_" + suffix(hx(H(c))), vbb(VB, [c] + comm[1:], bproof)
    for p in INVALID_G1:
        yield h, "verify_blob_kzg_proof_batch_case_invalid_proof_" + suffix(hx(H(p))), vbb(VB, comm, [p] + bproof[1:])
    yield h, "verify_blob_kzg_proof_batch_case_blob_length_different", vbb(VB[:-1], comm, bproof)
    yield h, "verify_blob_kzg_proof_batch_case_commitment_length_different", vbb(VB, comm[:-1], bproof)
    yield h, "verify_blob_kzg_proof_batch_case_proof_length_different", vbb(VB, comm, bproof[:-1])


def check_all(backend="oracle", write=None, verbose=False, setup=None):
    oids = json.load(open(os.path.join(ROOT, "tests/golden/lfs_oids.json")))
    setup, errcls, g1_add = get_backend(backend, setup)
    seen, bad = {}, []
    for handler, case, data in cases(setup, errcls, g1_add):
        exp = oids.get(handler, {}).get(case)
        ok = exp is not None and hashlib.sha256(data).hexdigest() == exp["oid"] and len(data) == exp["size"]
        seen.setdefault(handler, []).append(case)
        if not ok:
            bad.append((handler, case, "unknown case name" if exp is None else "sha/size mismatch (%d vs %d)" % (len(data), exp["size"])))
        if verbose:
            print("ok " if ok else "BAD", handler, case)
        if write:
            d = os.path.join(write, handler, case); os.makedirs(d, exist_ok=True)
            open(os.path.join(d, "data.yaml"), "wb").write(data)
    missing = [(h, c) for h in oids for c in oids[h] if c not in seen.get(h, [])]
    return {h: len(v) for h, v in seen.items()}, bad, missing


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="oracle")
    ap.add_argument("--write")
    ap.add_argument("-v", action="store_true")
    a = ap.parse_args()
    counts, bad, missing = check_all(a.backend, a.write, a.v)
    total = sum(counts.values())
    print("rebuilt", counts, "total", total)
    for b in bad:
        print("MISMATCH", *b)
    for m in missing:
        print("MISSING", *m)
    print("%d/%d match their LFS sha256 oid" % (total - len(bad), total))
    sys.exit(1 if bad or missing else 0)
