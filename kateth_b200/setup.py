"""ctypes binding of include/kateth_b200.h and the `Setup` mirror class (see package docstring)."""
import ctypes as C
import os

BYTES_PER_BLOB = 131072
RECORD_BYTES = 336
NOALLOC_MAX_BLOBS = 64  # KB200_NOALLOC_MAX_BLOBS (include/kateth_b200.h)
_HERE = os.path.dirname(os.path.abspath(__file__))

_STATUS_NAMES = {
    1: "Blob(InvalidLen)", 2: "Blob(InvalidFieldElement)", 3: "Bls(ECGroup(InvalidEncoding))", 4: "Bls(ECGroup(NotOnCurve))",
    5: "Bls(ECGroup(NotInGroup))", 6: "Bls(FiniteField(InvalidEncoding))", 7: "Bls(FiniteField(NotInFiniteField))",
    -1: "CUDA", -2: "Io", -3: "Serde", -10: "InvalidLenG1Lagrange", -11: "InvalidLenG2Monomial", -12: "G1MonomialMismatch", -20: "BadArgument", -30: "Nccl",
}


class KzgError(Exception):
    """Mirror of kateth's kzg::Error / LoadSetupError (src/kzg/mod.rs:15-31, src/kzg/setup.rs:20-28)."""

    def __init__(self, status, detail=""):
        super().__init__("%s (status %d)%s" % (_STATUS_NAMES.get(status, "?"), status, (": " + detail) if detail else ""))
        self.status = status


def lib_path():
    # KB200_LIB: another build of the same library (A/B measurements of kernel variants on one box)
    return os.environ.get("KB200_LIB") or os.path.join(_HERE, "libkateth_b200.so")


_lib = None


class _Info(C.Structure):
    _fields_ = [("window_bits", C.c_int), ("windows", C.c_int), ("table_bytes", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("device", C.c_int), ("sm_count", C.c_int), ("setup_seconds", C.c_double),
                ("last_msm_batch_affine", C.c_int), ("ba_grid", C.c_int), ("glv", C.c_int), ("table_from_cache", C.c_int)]


# every symbol include/kateth_b200.h declares (tests/test_abi.py checks the header against this list and the .so)
SYMBOLS = [
    "kb200_setup_load_json", "kb200_setup_load_txt", "kb200_setup_load_raw", "kb200_setup_free", "kb200_setup_check_g1_monomial", "kb200_set_table_cache", "kb200_g1_validate", "kb200_set_stream", "kb200_last_error", "kb200_alloc_pinned", "kb200_free_pinned",
    "kb200_blob_to_commitment", "kb200_blob_to_commitment_device", "kb200_blob_to_commitment_device_noalloc", "kb200_check_blobs", "kb200_compute_blob_proof", "kb200_compute_blob_proof_device",
    "kb200_compute_proof", "kb200_compute_proof_device", "kb200_verify_proof", "kb200_verify_proof_batch", "kb200_verify_proof_many", "kb200_verify_blob_proof",
    "kb200_verify_blob_proof_batch", "kb200_verify_blob_proof_batch_device", "kb200_verify_blob_proof_batch_partial_device",
    "kb200_verify_blob_proof_batch_partial", "kb200_commit_prove_verify_partial", "kb200_verify_proof_batch_partial_device", "kb200_verify_proof_batch_partial", "kb200_verify_batch_finish_device", "kb200_set_profiling", "kb200_set_msm_mode", "kb200_set_rlc_mode", "kb200_set_batch_challenge", "kb200_get_msm_profile", "kb200_last_pairing_operands", "kb200_get_info", "kb200_test_fp_mul", "kb200_test_fr_mul",
    "kb200_test_g1_decompress", "kb200_test_mac_peak", "kb200_test_host_pairing",
    "kb200_group_create_json", "kb200_group_unique_id", "kb200_group_create_rank", "kb200_group_free", "kb200_group_info", "kb200_group_replica",
    "kb200_group_blob_to_commitment", "kb200_group_compute_blob_proof", "kb200_group_compute_proof", "kb200_group_verify_blob_proof_batch",
    "kb200_group_verify_blob_proof_batch_device", "kb200_group_verify_proof_batch", "kb200_group_verify_proof_many", "kb200_group_verify_proof_batch_device", "kb200_group_commit_prove_verify",
]


def lib():
    """Load libkateth_b200.so.  Fails loudly if it has not been built (python -c 'import __graft_entry__ as g; g.build()')."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise RuntimeError("libkateth_b200.so is not built (%s); run __graft_entry__.build(). There is no CPU fallback." % p)
        _lib = C.CDLL(p)
        _lib.kb200_last_error.restype = C.c_char_p
        for name in SYMBOLS:
            if name not in ("kb200_last_error", "kb200_setup_free", "kb200_group_free", "kb200_group_replica"):
                getattr(_lib, name).restype = C.c_int
        _lib.kb200_setup_free.restype = None
        _lib.kb200_setup_free.argtypes = [C.c_void_p]
        _lib.kb200_group_free.restype = None
        _lib.kb200_group_free.argtypes = [C.c_void_p]
        _lib.kb200_group_replica.restype = C.c_void_p
        _lib.kb200_group_replica.argtypes = [C.c_void_p, C.c_int]
    return _lib


def g1_validate(points48, device=0):
    """P1::decompress for a list of 48-byte encodings, no Setup needed: list of statuses (0, 3, 4 or 5)."""
    n = len(points48)
    st = (C.c_int32 * n)()
    _check(lib().kb200_g1_validate(b"".join(points48), C.c_size_t(n), int(device), st))
    return list(st)


def set_table_cache(directory):
    """Process-wide on-disk cache of the window table (None switches it off)."""
    _check(lib().kb200_set_table_cache(os.fsencode(directory) if directory else None))


def _vp(x):
    return C.c_void_p(int(x))


def _check(st):
    if st != 0:
        raise KzgError(st, lib().kb200_last_error().decode() if st < 0 else "")


class Setup:
    """kateth::kzg::Setup<4096, 65> (src/kzg/setup.rs:37-276) backed by the GPU engine."""

    def __init__(self, handle):
        self._h = handle

    # -- Setup::load_json (src/kzg/setup.rs:46-82)
    @classmethod
    def load_json(cls, path, device=0, window_bits=0):
        h = C.c_void_p()
        _check(lib().kb200_setup_load_json(os.fsencode(path), int(device), int(window_bits), C.byref(h)))
        return cls(h)

    @classmethod
    def load_txt(cls, path, device=0, window_bits=0):
        """c-kzg-4844 trusted_setup.txt (not a kateth format; same tables as load_json of the same points)."""
        h = C.c_void_p()
        _check(lib().kb200_setup_load_txt(os.fsencode(path), int(device), int(window_bits), C.byref(h)))
        return cls(h)

    def close(self):
        if self._h:
            lib().kb200_setup_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check_g1_monomial(self, points48):
        """Raises KzgError(-12) unless points48[k] == commit(X^k) for every k (kb200_setup_check_g1_monomial)."""
        _check(lib().kb200_setup_check_g1_monomial(self._h, b"".join(points48), C.c_size_t(len(points48))))

    def set_stream(self, cuda_stream_ptr):
        _check(lib().kb200_set_stream(self._h, _vp(cuda_stream_ptr)))

    def info(self):
        i = _Info()
        _check(lib().kb200_get_info(self._h, C.byref(i)))
        return {k: getattr(i, k) for k, _ in _Info._fields_}

    # -- Setup::blob_to_commitment (src/kzg/setup.rs:167-171), returns the compressed commitment
    def blob_to_commitment(self, blob):
        blob = bytes(blob)
        if len(blob) != BYTES_PER_BLOB:
            raise KzgError(1)
        out = C.create_string_buffer(48)
        _check(lib().kb200_blob_to_commitment(self._h, blob, C.c_size_t(1), out, None))
        return out.raw

    # -- Setup::blob_proof (src/kzg/setup.rs:177-183)
    def blob_proof(self, blob, commitment):
        blob, commitment = bytes(blob), bytes(commitment)
        if len(blob) != BYTES_PER_BLOB:
            raise KzgError(1)
        if len(commitment) != 48:      # &Bytes48 in kateth: unrepresentable; the vector harness maps it to an error
            raise KzgError(3)
        out = C.create_string_buffer(48)
        _check(lib().kb200_compute_blob_proof(self._h, blob, commitment, C.c_size_t(1), out, None))
        return out.raw

    # -- Setup::proof (src/kzg/setup.rs:185-194) -> (proof, y)
    def proof(self, blob, z):
        blob, z = bytes(blob), bytes(z)
        if len(blob) != BYTES_PER_BLOB:
            raise KzgError(1)
        if len(z) != 32:
            raise KzgError(6)
        out, y = C.create_string_buffer(48), C.create_string_buffer(32)
        _check(lib().kb200_compute_proof(self._h, blob, z, C.c_size_t(1), out, y, None))
        return out.raw, y.raw

    # -- Setup::verify_proof (src/kzg/setup.rs:96-113): order proof, commitment, point, eval
    def verify_proof(self, proof, commitment, z, y):
        proof, commitment, z, y = bytes(proof), bytes(commitment), bytes(z), bytes(y)
        if len(proof) != 48 or len(commitment) != 48:
            raise KzgError(3)
        if len(z) != 32 or len(y) != 32:
            raise KzgError(6)
        ok = C.c_int(0)
        _check(lib().kb200_verify_proof(self._h, proof, commitment, z, y, C.byref(ok)))
        return bool(ok.value)

    # -- Setup::verify_blob_proof (src/kzg/setup.rs:208-221)
    def verify_blob_proof(self, blob, commitment, proof):
        blob, commitment, proof = bytes(blob), bytes(commitment), bytes(proof)
        if len(blob) != BYTES_PER_BLOB:
            raise KzgError(1)
        if len(commitment) != 48 or len(proof) != 48:
            raise KzgError(3)
        ok = C.c_int(0)
        _check(lib().kb200_verify_blob_proof(self._h, blob, commitment, proof, C.byref(ok)))
        return bool(ok.value)

    # -- Setup::verify_blob_proof_batch (src/kzg/setup.rs:247-275)
    def verify_blob_proof_batch(self, blobs, commitments, proofs):
        assert len(blobs) == len(commitments) == len(proofs)   # kateth: assert_eq! (panics)
        blobs = [bytes(b) for b in blobs]
        # kateth decodes blob by blob (src/kzg/setup.rs:259-262, src/blob.rs:26-37): the first error in BLOB order wins, so an
        # element error of an earlier blob precedes the length error of a later one
        short = next((i for i, b in enumerate(blobs) if len(b) != BYTES_PER_BLOB), None)
        if short is not None:
            if short:
                st = (C.c_int32 * short)()
                rc = lib().kb200_check_blobs(self._h, b"".join(blobs[:short]), C.c_size_t(short), st)
                if rc:
                    _check(rc)
            raise KzgError(1)
        for c in list(commitments) + list(proofs):
            if len(c) != 48:
                raise KzgError(3)
        n = len(blobs)
        ok = C.c_int(0)
        _check(lib().kb200_verify_blob_proof_batch(self._h, b"".join(blobs), b"".join(bytes(c) for c in commitments),
                                                   b"".join(bytes(p) for p in proofs), C.c_size_t(n), C.byref(ok), None))
        return bool(ok.value)

    def last_pairing_operands(self):
        a, b = C.create_string_buffer(48), C.create_string_buffer(48)
        _check(lib().kb200_last_pairing_operands(self._h, a, b))
        return a.raw, b.raw

    # ---------------- batch / resident-data extensions (the throughput surface) ----------------
    def blob_to_commitment_many(self, blobs_concat, n):
        out, st = C.create_string_buffer(48 * n), (C.c_int32 * n)()
        rc = lib().kb200_blob_to_commitment(self._h, blobs_concat, C.c_size_t(n), out, st)
        if rc < 0:
            _check(rc)
        return out.raw, list(st)

    def blob_proof_many(self, blobs_concat, commitments_concat, n):
        out, st = C.create_string_buffer(48 * n), (C.c_int32 * n)()
        rc = lib().kb200_compute_blob_proof(self._h, blobs_concat, commitments_concat, C.c_size_t(n), out, st)
        if rc < 0:
            _check(rc)
        return out.raw, list(st)

    def proof_many(self, blobs_concat, z_concat, n):
        out, y, st = C.create_string_buffer(48 * n), C.create_string_buffer(32 * n), (C.c_int32 * n)()
        rc = lib().kb200_compute_proof(self._h, blobs_concat, z_concat, C.c_size_t(n), out, y, st)
        if rc < 0:
            _check(rc)
        return out.raw, y.raw, list(st)

    def verify_proof_batch(self, proofs_concat, commitments_concat, z_concat, y_concat, n):
        ok = C.c_int(0)
        _check(lib().kb200_verify_proof_batch(self._h, proofs_concat, commitments_concat, z_concat, y_concat, C.c_size_t(n), C.byref(ok), None))
        return bool(ok.value)

    def verify_proof_many(self, proofs_concat, commitments_concat, z_concat, y_concat, n):
        """n independent verify_proof calls, verdicts from the device pairing: returns (ok list of bool, status list)."""
        ok, st = (C.c_int32 * n)(), (C.c_int32 * n)()
        _check(lib().kb200_verify_proof_many(self._h, proofs_concat, commitments_concat, z_concat, y_concat, C.c_size_t(n), ok, st))
        return [bool(v) for v in ok], list(st)

    def verify_blob_proof_batch_concat(self, blobs_concat, commitments_concat, proofs_concat, n):
        ok = C.c_int(0)
        _check(lib().kb200_verify_blob_proof_batch(self._h, blobs_concat, commitments_concat, proofs_concat, C.c_size_t(n), C.byref(ok), None))
        return bool(ok.value)

    # device-pointer calls (integers = device addresses, e.g. torch.Tensor.data_ptr())
    def blob_to_commitment_device_noalloc(self, d_blobs, n, d_out48, d_status):
        """kb200_blob_to_commitment_device_noalloc: at most NOALLOC_MAX_BLOBS blobs, nothing allocated, enqueue-only (the caller
        synchronises the stream and reads d_status / d_out48)."""
        return lib().kb200_blob_to_commitment_device_noalloc(self._h, _vp(d_blobs), C.c_size_t(n), _vp(d_out48), _vp(d_status))

    def blob_to_commitment_device(self, d_blobs, n, d_out48, d_status):
        return lib().kb200_blob_to_commitment_device(self._h, _vp(d_blobs), C.c_size_t(n), _vp(d_out48), _vp(d_status))

    def blob_proof_device(self, d_blobs, d_comm48, n, d_out48, d_status):
        return lib().kb200_compute_blob_proof_device(self._h, _vp(d_blobs), _vp(d_comm48), C.c_size_t(n), _vp(d_out48), _vp(d_status))

    def proof_device(self, d_blobs, d_z32, n, d_out48, d_y32, d_status):
        return lib().kb200_compute_proof_device(self._h, _vp(d_blobs), _vp(d_z32), C.c_size_t(n), _vp(d_out48), _vp(d_y32), _vp(d_status))

    def verify_blob_proof_batch_device(self, d_blobs, d_comm48, d_proofs48, n, d_status):
        ok = C.c_int(0)
        _check(lib().kb200_verify_blob_proof_batch_device(self._h, _vp(d_blobs), _vp(d_comm48), _vp(d_proofs48), C.c_size_t(n), C.byref(ok), _vp(d_status)))
        return bool(ok.value)

    def verify_partial_device(self, d_blobs, d_comm48, d_proofs48, n_local, index_offset, n_total, d_record, d_status):
        return lib().kb200_verify_blob_proof_batch_partial_device(self._h, _vp(d_blobs), _vp(d_comm48), _vp(d_proofs48), C.c_size_t(n_local),
                                                                  C.c_size_t(index_offset), C.c_size_t(n_total), _vp(d_record), _vp(d_status))

    def verify_partial_host(self, blobs, comm48, proofs48, n_local, index_offset, n_total, d_record):
        """blobs / comm48 / proofs48: integer host addresses (e.g. pinned torch tensors' data_ptr()) or bytes."""
        conv = lambda x: x if isinstance(x, (bytes, bytearray)) else _vp(x)
        return lib().kb200_verify_blob_proof_batch_partial(self._h, conv(blobs), conv(comm48), conv(proofs48), C.c_size_t(n_local),
                                                           C.c_size_t(index_offset), C.c_size_t(n_total), _vp(d_record), None)

    def verify_proof_partial_host(self, proofs, comm48, z32, y32, n_local, index_offset, n_total, d_record):
        """One shard of a sharded batch of (proof, commitment, z, y) tuples (host bytes); the record goes to device memory."""
        return lib().kb200_verify_proof_batch_partial(self._h, proofs, comm48, z32, y32, C.c_size_t(n_local), C.c_size_t(index_offset), C.c_size_t(n_total),
                                                      _vp(d_record), None)

    def set_rlc_mode(self, mode):
        """0 = automatic, 1 = ladders (k_rlc), 2 = bucket method (k_rlc_bucket): same records."""
        _check(lib().kb200_set_rlc_mode(self._h, int(mode)))

    def set_msm_mode(self, mode):
        """0 = automatic, 1 = XYZZ accumulation (k_msm), 2 = batch-affine summation (k_msm_ba); same bytes either way."""
        _check(lib().kb200_set_msm_mode(self._h, int(mode)))

    def set_batch_challenge(self, mode):
        """0 = kateth's batch challenge (default, parity with the reference), 1 = consensus-spec transcript challenge."""
        _check(lib().kb200_set_batch_challenge(self._h, int(mode)))

    def set_profiling(self, on=True):
        _check(lib().kb200_set_profiling(self._h, 1 if on else 0))

    def msm_profile(self):
        ms, l, m = C.c_double(0), C.c_uint64(0), C.c_uint64(0)
        _check(lib().kb200_get_msm_profile(self._h, C.byref(ms), C.byref(l), C.byref(m)))
        return {"ms": ms.value, "launches": l.value, "msms": m.value}

    # host-buffer batch calls on raw addresses (pinned host memory): the end-to-end path of bench.py
    def blob_to_commitment_host(self, blobs_ptr, n, out48_ptr, status_ptr=0):
        return lib().kb200_blob_to_commitment(self._h, _vp(blobs_ptr), C.c_size_t(n), _vp(out48_ptr), _vp(status_ptr) if status_ptr else None)

    def blob_proof_host(self, blobs_ptr, comm48_ptr, n, out48_ptr, status_ptr=0):
        return lib().kb200_compute_blob_proof(self._h, _vp(blobs_ptr), _vp(comm48_ptr), C.c_size_t(n), _vp(out48_ptr), _vp(status_ptr) if status_ptr else None)

    def verify_blob_proof_batch_host(self, blobs_ptr, comm48_ptr, proofs48_ptr, n):
        ok = C.c_int(0)
        _check(lib().kb200_verify_blob_proof_batch(self._h, _vp(blobs_ptr), _vp(comm48_ptr), _vp(proofs48_ptr), C.c_size_t(n), C.byref(ok), None))
        return bool(ok.value)

    def proof_host(self, blobs_ptr, z32_ptr, n, out48_ptr, y32_ptr, status_ptr=0):
        return lib().kb200_compute_proof(self._h, _vp(blobs_ptr), _vp(z32_ptr), C.c_size_t(n), _vp(out48_ptr), _vp(y32_ptr), _vp(status_ptr) if status_ptr else None)

    def verify_finish_device(self, d_records, n_records):
        ok = C.c_int(0)
        _check(lib().kb200_verify_batch_finish_device(self._h, _vp(d_records), C.c_size_t(n_records), C.byref(ok)))
        return bool(ok.value)


class Group:
    """kateth's Setup over several B200s, driven through the C ABI's kb200_group_* entry points (one replica per GPU, contiguous
    shards, the all-gather of batch verification inside the library).  Two constructors: `load_json` (this process drives all the
    listed devices) and `join` (one process per GPU: rank 0 makes `unique_id()`, every rank joins with its own Setup)."""

    def __init__(self, handle, keep=None):
        self._g, self._keep = handle, keep

    @classmethod
    def load_json(cls, path, devices=None, window_bits=0):
        import torch
        devs = list(range(torch.cuda.device_count())) if devices is None else list(devices)
        arr = (C.c_int * len(devs))(*devs)
        g = C.c_void_p()
        _check(lib().kb200_group_create_json(os.fsencode(path), arr, len(devs), int(window_bits), C.byref(g)))
        return cls(g)

    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        _check(lib().kb200_group_unique_id(buf))
        return buf.raw

    @classmethod
    def join(cls, setup, rank, world, unique_id):
        g = C.c_void_p()
        _check(lib().kb200_group_create_rank(setup._h, int(rank), int(world), bytes(unique_id) if unique_id else None, C.byref(g)))
        return cls(g, keep=setup)

    def close(self):
        if self._g:
            lib().kb200_group_free(self._g)
            self._g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        a, b, c, d = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        _check(lib().kb200_group_info(self._g, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return {"local_replicas": a.value, "world": b.value, "first_rank": c.value, "nccl_version": d.value}

    def replica(self, i):
        """The i-th local handle as a (non-owning) Setup."""
        h = lib().kb200_group_replica(self._g, int(i))
        if not h:
            raise IndexError(i)
        s = Setup(C.c_void_p(h))
        s.close = lambda: None        # owned by the group (or by the caller who passed it to join)
        return s

    @staticmethod
    def _p(x):
        return x if isinstance(x, (bytes, bytearray)) or x is None else _vp(x)

    def blob_to_commitment_many(self, blobs, n):
        out, st = C.create_string_buffer(48 * n), (C.c_int32 * n)()
        rc = lib().kb200_group_blob_to_commitment(self._g, self._p(blobs), C.c_size_t(n), out, st)
        if rc < 0:
            _check(rc)
        return out.raw, list(st)

    def blob_proof_many(self, blobs, commitments, n):
        out, st = C.create_string_buffer(48 * n), (C.c_int32 * n)()
        rc = lib().kb200_group_compute_blob_proof(self._g, self._p(blobs), self._p(commitments), C.c_size_t(n), out, st)
        if rc < 0:
            _check(rc)
        return out.raw, list(st)

    def proof_many(self, blobs, zs, n):
        out, y, st = C.create_string_buffer(48 * n), C.create_string_buffer(32 * n), (C.c_int32 * n)()
        rc = lib().kb200_group_compute_proof(self._g, self._p(blobs), self._p(zs), C.c_size_t(n), out, y, st)
        if rc < 0:
            _check(rc)
        return out.raw, y.raw, list(st)

    def verify_blob_proof_batch(self, blobs, commitments, proofs, n_local, index_offset=0, n_total=None):
        """Raises KzgError with the batch's error (the same on every rank), else returns the verdict."""
        ok = C.c_int(0)
        _check(lib().kb200_group_verify_blob_proof_batch(self._g, self._p(blobs), self._p(commitments), self._p(proofs), C.c_size_t(n_local),
                                                         C.c_size_t(index_offset), C.c_size_t(n_local if n_total is None else n_total), C.byref(ok), None))
        return bool(ok.value)

    def verify_proof_batch(self, proofs, commitments, zs, ys, n_local, index_offset=0, n_total=None):
        ok = C.c_int(0)
        _check(lib().kb200_group_verify_proof_batch(self._g, self._p(proofs), self._p(commitments), self._p(zs), self._p(ys), C.c_size_t(n_local),
                                                    C.c_size_t(index_offset), C.c_size_t(n_local if n_total is None else n_total), C.byref(ok), None))
        return bool(ok.value)

    def verify_proof_many(self, proofs, commitments, zs, ys, n):
        ok, st = (C.c_int32 * n)(), (C.c_int32 * n)()
        _check(lib().kb200_group_verify_proof_many(self._g, self._p(proofs), self._p(commitments), self._p(zs), self._p(ys), C.c_size_t(n), ok, st))
        return [bool(v) for v in ok], list(st)

    # raw-address forms for the bench (pinned host buffers / device pointers); return (rc, ok)
    def blob_to_commitment_host(self, blobs_ptr, n, out48_ptr):
        return lib().kb200_group_blob_to_commitment(self._g, _vp(blobs_ptr), C.c_size_t(n), _vp(out48_ptr), None)

    def blob_proof_host(self, blobs_ptr, comm48_ptr, n, out48_ptr):
        return lib().kb200_group_compute_blob_proof(self._g, _vp(blobs_ptr), _vp(comm48_ptr), C.c_size_t(n), _vp(out48_ptr), None)

    def verify_host(self, blobs_ptr, comm48_ptr, proofs48_ptr, n_local, index_offset, n_total):
        ok = C.c_int(0)
        rc = lib().kb200_group_verify_blob_proof_batch(self._g, _vp(blobs_ptr), _vp(comm48_ptr), _vp(proofs48_ptr), C.c_size_t(n_local), C.c_size_t(index_offset),
                                                       C.c_size_t(n_total), C.byref(ok), None)
        return rc, bool(ok.value)

    def commit_prove_verify(self, blobs, n_local, index_offset=0, n_total=None):
        """Fresh batch: (commitments, proofs, verdict) over one upload of the blobs per GPU."""
        oc, op, ok = C.create_string_buffer(48 * n_local), C.create_string_buffer(48 * n_local), C.c_int(0)
        _check(lib().kb200_group_commit_prove_verify(self._g, self._p(blobs), C.c_size_t(n_local), C.c_size_t(index_offset),
                                                     C.c_size_t(n_local if n_total is None else n_total), oc, op, C.byref(ok), None))
        return oc.raw, op.raw, bool(ok.value)

    def commit_prove_verify_host(self, blobs_ptr, n_local, index_offset, n_total, out_comm48_ptr, out_proof48_ptr):
        ok = C.c_int(0)
        rc = lib().kb200_group_commit_prove_verify(self._g, _vp(blobs_ptr), C.c_size_t(n_local), C.c_size_t(index_offset), C.c_size_t(n_total), _vp(out_comm48_ptr),
                                                   _vp(out_proof48_ptr), C.byref(ok), None)
        return rc, bool(ok.value)

    def verify_points_host(self, proofs_ptr, comm48_ptr, z32_ptr, y32_ptr, n_local, index_offset, n_total):
        ok = C.c_int(0)
        rc = lib().kb200_group_verify_proof_batch(self._g, _vp(proofs_ptr), _vp(comm48_ptr), _vp(z32_ptr), _vp(y32_ptr), C.c_size_t(n_local), C.c_size_t(index_offset),
                                                  C.c_size_t(n_total), C.byref(ok), None)
        return rc, bool(ok.value)

    def verify_points_device(self, d_proofs48, d_comm48, d_z32, d_y32, n_per_replica, index_offset, n_total, d_status):
        k = len(n_per_replica)
        vp = lambda xs: (C.c_void_p * k)(*[int(x) for x in xs])
        ok = C.c_int(0)
        rc = lib().kb200_group_verify_proof_batch_device(self._g, vp(d_proofs48), vp(d_comm48), vp(d_z32), vp(d_y32), (C.c_size_t * k)(*n_per_replica),
                                                         C.c_size_t(index_offset), C.c_size_t(n_total), C.byref(ok), vp(d_status))
        return rc, bool(ok.value)

    def verify_device(self, d_blobs, d_comm48, d_proofs48, n_per_replica, index_offset, n_total, d_status):
        """Each argument: one entry per local replica (device addresses on that replica's GPU / item counts)."""
        k = len(n_per_replica)
        vp = lambda xs: (C.c_void_p * k)(*[int(x) for x in xs])
        ok = C.c_int(0)
        rc = lib().kb200_group_verify_blob_proof_batch_device(self._g, vp(d_blobs), vp(d_comm48), vp(d_proofs48), (C.c_size_t * k)(*n_per_replica),
                                                              C.c_size_t(index_offset), C.c_size_t(n_total), C.byref(ok), vp(d_status))
        return rc, bool(ok.value)


def shard_bounds(n_total, world_size, rank):
    """Contiguous shard [lo, hi) of rank `rank`."""
    return n_total * rank // world_size, n_total * (rank + 1) // world_size


def batch_error_of_records(gathered):
    """Host-side reading of what kb200_verify_batch_finish_device returns for a set of gathered records: each KB200_RECORD_BYTES
    record ends in four little-endian int32 {first blob error, first commitment error, first proof error, 0} of its shard; the
    batch's error is the first blob error in rank order, else the first commitment error, else the first proof error
    (kateth: src/kzg/setup.rs:259-271).  `gathered`: bytes-like, world x RECORD_BYTES.  Returns 0 when no shard failed."""
    import struct
    buf = bytes(gathered)
    errs = [struct.unpack_from("<4i", buf, g * RECORD_BYTES + 320) for g in range(len(buf) // RECORD_BYTES)]
    for k in range(3):
        for e in errs:
            if e[k]:
                return e[k]
    return 0


class BatchVerifier:
    """verify_blob_kzg_proof_batch sharded over the ranks of a torch.distributed process group: every rank computes the
    336-byte partial record (sums + first error of each class) of its contiguous shard on its GPU, ONE all_gather (NCCL over NVLink) exchanges the
    records, and every rank finishes (sum of records -> two-pairing check on the host)."""

    def __init__(self, setup, group=None):
        self.setup, self.group = setup, group

    def verify(self, d_blobs, d_comm48, d_proofs48, n_local, index_offset, n_total, d_status, record_tensor, gathered_tensor):
        """Returns (rc, ok): rc is the batch's error on EVERY rank (kateth's order: first blob error in rank order, else first
        commitment error, else first proof error -- the errors travel inside the records), ok the verdict when rc == 0."""
        import torch.distributed as dist
        rc = self.setup.verify_partial_device(d_blobs, d_comm48, d_proofs48, n_local, index_offset, n_total, record_tensor.data_ptr(), d_status)
        if rc < 0:
            _check(rc)
        dist.all_gather_into_tensor(gathered_tensor, record_tensor, group=self.group)
        world = gathered_tensor.numel() // RECORD_BYTES
        ok = C.c_int(0)
        rc = lib().kb200_verify_batch_finish_device(self.setup._h, _vp(gathered_tensor.data_ptr()), C.c_size_t(world), C.byref(ok))
        if rc < 0:
            _check(rc)
        return rc, bool(ok.value) if rc == 0 else None


class test_hooks:
    @staticmethod
    def fp_mul(a_u32, b_u32):
        import numpy as np
        a = np.ascontiguousarray(a_u32, dtype=np.uint32); b = np.ascontiguousarray(b_u32, dtype=np.uint32)
        out = np.empty_like(a)
        _check(lib().kb200_test_fp_mul(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), C.c_size_t(a.shape[0])))
        return out

    @staticmethod
    def fr_mul(a_u32, b_u32):
        import numpy as np
        a = np.ascontiguousarray(a_u32, dtype=np.uint32); b = np.ascontiguousarray(b_u32, dtype=np.uint32)
        out = np.empty_like(a)
        _check(lib().kb200_test_fr_mul(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), C.c_size_t(a.shape[0])))
        return out

    @staticmethod
    def g1_decompress(setup, points48, by_order=False):
        n = len(points48)
        st, out = (C.c_int32 * n)(), C.create_string_buffer(48 * n)
        _check(lib().kb200_test_g1_decompress(setup._h, b"".join(points48), C.c_size_t(n), st, out, 1 if by_order else 0))
        return list(st), [out.raw[48 * i:48 * i + 48] for i in range(n)]

    @staticmethod
    def mac_peak(iters=4096):
        a, b = C.c_double(0), C.c_double(0)
        _check(lib().kb200_test_mac_peak(int(iters), C.byref(a), C.byref(b)))
        return a.value, b.value
