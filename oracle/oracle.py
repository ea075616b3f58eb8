"""ctypes front end of the CPU oracle (oracle/kzg_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, tools/reconstruct_vectors.py, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The engine package kateth_b200 never imports it.

The method names mirror kateth's `Setup` (/root/reference/src/kzg/setup.rs:46,96,167,177,185,208,247); errors are
raised as OracleError(status) where status follows include/kateth_b200.h.
"""
import ctypes as C
import json
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libkzg_oracle.so")
BLOB_BYTES = 131072


class OracleError(Exception):
    def __init__(self, status):
        super().__init__("oracle status %d" % status)
        self.status = status


def build(force=False):
    if force or not os.path.exists(LIB) or any(
        os.path.getmtime(os.path.join(HERE, f)) > os.path.getmtime(LIB)
        for f in ("kzg_oracle.c", "kzo_field.h", "kzo_g1.h", "kzo_pairing.h", "constants.h")
    ):
        subprocess.check_call(["make", "-C", HERE, "-s"])
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB)
        _lib.kzo_setup_load.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(C.c_void_p)]
        for name in ("kzo_blob_to_commitment", "kzo_blob_proof", "kzo_proof", "kzo_verify_proof", "kzo_verify_blob_proof",
                     "kzo_verify_blob_proof_batch", "kzo_verify_blob_proof_batch_sums", "kzo_commit_prove_many", "kzo_verify_many",
                     "kzo_g1_lincomb", "kzo_g1_add", "kzo_blob_challenge_eval", "kzo_pairing_eq", "kzo_g1_decompress_check", "kzo_g2_mul_gen"):
            getattr(_lib, name).restype = C.c_int
    return _lib


def _hexlist(xs):
    return b"".join(bytes.fromhex(x[2:] if x.startswith("0x") else x) for x in xs)


class Setup:
    """Oracle-side stand-in for kateth's Setup<4096, 65>."""

    def __init__(self, handle):
        self._h = handle

    @classmethod
    def load_json(cls, path):
        d = json.load(open(path))
        g1, g2 = d["g1_lagrange"], d["g2_monomial"]
        if any(len(x) != 98 for x in g1) or any(len(x) != 194 for x in g2):
            raise OracleError(3)
        h = C.c_void_p()
        st = lib().kzo_setup_load(_hexlist(g1), len(g1), _hexlist(g2), len(g2), C.byref(h))
        if st:
            raise OracleError(st)
        return cls(h)

    def blob_to_commitment(self, blob):
        out = C.create_string_buffer(48)
        st = lib().kzo_blob_to_commitment(self._h, bytes(blob), C.c_size_t(len(blob)), out)
        if st:
            raise OracleError(st)
        return out.raw

    def blob_proof(self, blob, commitment):
        out = C.create_string_buffer(48)
        st = lib().kzo_blob_proof(self._h, bytes(blob), C.c_size_t(len(blob)), bytes(commitment), C.c_size_t(len(commitment)), out)
        if st:
            raise OracleError(st)
        return out.raw

    def proof(self, blob, z):
        out, y = C.create_string_buffer(48), C.create_string_buffer(32)
        st = lib().kzo_proof(self._h, bytes(blob), C.c_size_t(len(blob)), bytes(z), C.c_size_t(len(z)), out, y)
        if st:
            raise OracleError(st)
        return out.raw, y.raw

    def verify_proof(self, proof, commitment, z, y):
        ok = C.c_int(0)
        st = lib().kzo_verify_proof(self._h, bytes(proof), C.c_size_t(len(proof)), bytes(commitment), C.c_size_t(len(commitment)),
                                    bytes(z), C.c_size_t(len(z)), bytes(y), C.c_size_t(len(y)), C.byref(ok))
        if st:
            raise OracleError(st)
        return bool(ok.value)

    def verify_blob_proof(self, blob, commitment, proof):
        ok = C.c_int(0)
        st = lib().kzo_verify_blob_proof(self._h, bytes(blob), C.c_size_t(len(blob)), bytes(commitment), C.c_size_t(len(commitment)),
                                         bytes(proof), C.c_size_t(len(proof)), C.byref(ok))
        if st:
            raise OracleError(st)
        return bool(ok.value)

    def _batch_args(self, blobs, commitments, proofs):
        assert len(blobs) == len(commitments) == len(proofs)  # kateth: assert_eq! panics (src/kzg/setup.rs:256-257)
        for b in blobs:
            if len(b) != BLOB_BYTES:
                raise OracleError(1)
        for c in list(commitments) + list(proofs):
            if len(c) != 48:
                raise OracleError(3)
        return b"".join(bytes(b) for b in blobs), b"".join(bytes(c) for c in commitments), b"".join(bytes(p) for p in proofs)

    def verify_blob_proof_batch(self, blobs, commitments, proofs):
        b, c, p = self._batch_args(blobs, commitments, proofs)
        ok = C.c_int(0)
        st = lib().kzo_verify_blob_proof_batch(self._h, b, c, p, C.c_size_t(len(blobs)), C.byref(ok))
        if st:
            raise OracleError(st)
        return bool(ok.value)

    def verify_blob_proof_batch_sums(self, blobs, commitments, proofs):
        b, c, p = self._batch_args(blobs, commitments, proofs)
        a48, b48 = C.create_string_buffer(48), C.create_string_buffer(48)
        st = lib().kzo_verify_blob_proof_batch_sums(self._h, b, c, p, C.c_size_t(len(blobs)), a48, b48)
        if st:
            raise OracleError(st)
        return a48.raw, b48.raw

    def challenge_eval(self, blob, commitment):
        z, y = C.create_string_buffer(32), C.create_string_buffer(32)
        st = lib().kzo_blob_challenge_eval(self._h, bytes(blob), bytes(commitment), z, y)
        if st:
            raise OracleError(st)
        return z.raw, y.raw

    # ---- CPU-baseline legs (bench.py only) ----
    def commit_prove_many(self, blobs_concat, n, mode=2, nthreads=1, commitments=None):
        out_c, out_p = C.create_string_buffer(48 * n), C.create_string_buffer(48 * n)
        st = lib().kzo_commit_prove_many(self._h, blobs_concat, commitments, C.c_size_t(n), mode, nthreads, out_c, out_p)
        if st:
            raise OracleError(st)
        return out_c.raw, out_p.raw

    def verify_many(self, blobs_concat, comms, proofs, n, nthreads=1):
        ok = C.c_int(0)
        st = lib().kzo_verify_many(self._h, blobs_concat, comms, proofs, C.c_size_t(n), nthreads, C.byref(ok))
        if st:
            raise OracleError(st)
        return bool(ok.value)


# ---- free functions used by unit tests ----
def g1_lincomb(points48, scalars32):
    out = C.create_string_buffer(48)
    st = lib().kzo_g1_lincomb(b"".join(points48), b"".join(scalars32), C.c_size_t(len(points48)), out)
    if st:
        raise OracleError(st)
    return out.raw


def g1_add(a48, b48):
    out = C.create_string_buffer(48)
    st = lib().kzo_g1_add(bytes(a48), bytes(b48), out)
    if st:
        raise OracleError(st)
    return out.raw


def g1_decompress_status(p48):
    return lib().kzo_g1_decompress_check(bytes(p48))


def sha256(data):
    out = C.create_string_buffer(32)
    lib().kzo_sha256(bytes(data), C.c_size_t(len(data)), out)
    return out.raw


def fp_mul_raw(a_u64, b_u64):
    """a, b: numpy uint64 arrays of shape (n, 6) in Montgomery form -> (n, 6)."""
    import numpy as np
    a = np.ascontiguousarray(a_u64, dtype=np.uint64); b = np.ascontiguousarray(b_u64, dtype=np.uint64)
    out = np.empty_like(a)
    lib().kzo_fp_mul_raw(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), C.c_size_t(a.shape[0]))
    return out


def fr_mul_raw(a_u64, b_u64):
    import numpy as np
    a = np.ascontiguousarray(a_u64, dtype=np.uint64); b = np.ascontiguousarray(b_u64, dtype=np.uint64)
    out = np.empty_like(a)
    lib().kzo_fr_mul_raw(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), C.c_size_t(a.shape[0]))
    return out


def fr_consts():
    one, mx = (C.c_uint64 * 4)(), (C.c_uint64 * 4)()
    lib().kzo_fr_consts(one, mx)
    return list(one), list(mx)


def pairing_eq(a1, a2, b1, b2):
    ok = C.c_int(0)
    st = lib().kzo_pairing_eq(bytes(a1), bytes(a2), bytes(b1), bytes(b2), C.byref(ok))
    if st:
        raise OracleError(st)
    return bool(ok.value)


def g2_mul_gen(k32):
    out = C.create_string_buffer(96)
    st = lib().kzo_g2_mul_gen(bytes(k32), out)
    if st:
        raise OracleError(st)
    return out.raw
