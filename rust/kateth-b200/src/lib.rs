//! `kateth::kzg::Setup<4096, 65>` backed by the B200 engine (libkateth_b200.so, include/kateth_b200.h).
//!
//! Same method names, argument order and error enums as kateth (`/root/reference/src/kzg/setup.rs:37-276`,
//! `src/kzg/mod.rs:15-31`, `src/blob.rs:6-16`, `src/bls.rs:21-50`).  Points and scalars travel as the bytes kateth's own
//! tests compare (48-byte compressed G1, 32-byte big-endian scalars); inside kateth the returned `Bytes48` is turned into
//! `Commitment = Proof = P1` with `P1::decompress` (or `blst_p1_from_affine`, the engine's limb layout being blst's).
//!
//! NOT COMPILED HERE: the build image has no Rust toolchain.  The `ctypes` mirror `kateth_b200/setup.py` is the binding the
//! parity tests drive; this file is the same wrapper in the language `north_star` names.
#![allow(non_camel_case_types)]

use core::ffi::{c_char, c_int, c_void};
use std::ffi::CString;
use std::path::Path;

pub const FIELD_ELEMENTS_PER_BLOB: usize = 4096;
pub const BYTES_PER_BLOB: usize = 131072;
pub type Bytes32 = [u8; 32];
pub type Bytes48 = [u8; 48];

#[repr(C)]
pub struct kb200_setup {
    _private: [u8; 0],
}

extern "C" {
    fn kb200_setup_load_json(path: *const c_char, device: c_int, window_bits: c_int, out: *mut *mut kb200_setup) -> c_int;
    fn kb200_setup_free(s: *mut kb200_setup);
    fn kb200_last_error() -> *const c_char;
    fn kb200_blob_to_commitment(s: *mut kb200_setup, blobs: *const u8, n: usize, out48: *mut u8, status: *mut i32) -> c_int;
    fn kb200_compute_blob_proof(s: *mut kb200_setup, blobs: *const u8, commitments48: *const u8, n: usize, out48: *mut u8, status: *mut i32) -> c_int;
    fn kb200_compute_proof(s: *mut kb200_setup, blobs: *const u8, z32: *const u8, n: usize, out_proof48: *mut u8, out_y32: *mut u8, status: *mut i32) -> c_int;
    fn kb200_verify_proof(s: *mut kb200_setup, proof48: *const u8, commitment48: *const u8, z32: *const u8, y32: *const u8, ok: *mut c_int) -> c_int;
    fn kb200_verify_blob_proof(s: *mut kb200_setup, blob: *const u8, commitment48: *const u8, proof48: *const u8, ok: *mut c_int) -> c_int;
    fn kb200_verify_blob_proof_batch(s: *mut kb200_setup, blobs: *const u8, commitments48: *const u8, proofs48: *const u8, n: usize, ok: *mut c_int, status: *mut i32) -> c_int;
    // throughput forms: data already in HBM, sharded batch verification around one 320-byte all-gather
    pub fn kb200_blob_to_commitment_device(s: *mut kb200_setup, d_blobs: *const c_void, n: usize, d_out48: *mut c_void, d_status: *mut i32) -> c_int;
    pub fn kb200_compute_blob_proof_device(s: *mut kb200_setup, d_blobs: *const c_void, d_commitments48: *const c_void, n: usize, d_out48: *mut c_void, d_status: *mut i32) -> c_int;
    pub fn kb200_verify_blob_proof_batch_partial_device(s: *mut kb200_setup, d_blobs: *const c_void, d_commitments48: *const c_void, d_proofs48: *const c_void,
                                                        n_local: usize, index_offset: usize, n_total: usize, d_record: *mut c_void, d_status: *mut i32) -> c_int;
    pub fn kb200_verify_batch_finish_device(s: *mut kb200_setup, d_records: *const c_void, n_records: usize, ok: *mut c_int) -> c_int;
}

// ---- kateth's error enums (names and nesting as in the reference)
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum BlobError { InvalidFieldElement, InvalidLen }                       // src/blob.rs:6-16
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum ECGroupError { InvalidEncoding, NotInGroup, NotOnCurve }            // src/bls.rs:32-38
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum FiniteFieldError { InvalidEncoding, NotInFiniteField }              // src/bls.rs:21-30
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum BlsError { ECGroup(ECGroupError), FiniteField(FiniteFieldError) }   // src/bls.rs:40-50
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Error { Blob(BlobError), Bls(BlsError) }                            // src/kzg/mod.rs:15-31
#[derive(Debug)]
pub enum LoadSetupError { Bls(BlsError), Io(String), Serde(String), InvalidLenG1Lagrange, InvalidLenG2Monomial }   // src/kzg/setup.rs:20-28

fn last_error() -> String {
    unsafe { std::ffi::CStr::from_ptr(kb200_last_error()).to_string_lossy().into_owned() }
}

fn error_from_status(code: c_int) -> Error {
    match code {
        1 => Error::Blob(BlobError::InvalidLen),
        2 => Error::Blob(BlobError::InvalidFieldElement),
        3 => Error::Bls(BlsError::ECGroup(ECGroupError::InvalidEncoding)),
        4 => Error::Bls(BlsError::ECGroup(ECGroupError::NotOnCurve)),
        5 => Error::Bls(BlsError::ECGroup(ECGroupError::NotInGroup)),
        6 => Error::Bls(BlsError::FiniteField(FiniteFieldError::InvalidEncoding)),
        7 => Error::Bls(BlsError::FiniteField(FiniteFieldError::NotInFiniteField)),
        // CUDA / internal failures have no counterpart in kateth's enums: there is no CPU fallback to degrade to
        _ => panic!("kateth_b200: engine failure {code}: {}", last_error()),
    }
}

pub struct Setup<const G1: usize, const G2: usize> {
    handle: *mut kb200_setup,
}
// The handle owns immutable tables plus per-call scratch; calls on one handle must not overlap (include/kateth_b200.h).
unsafe impl<const G1: usize, const G2: usize> Send for Setup<G1, G2> {}

impl<const G1: usize, const G2: usize> Drop for Setup<G1, G2> {
    fn drop(&mut self) {
        unsafe { kb200_setup_free(self.handle) }
    }
}

impl Setup<4096, 65> {
    /// `Setup::load_json` (src/kzg/setup.rs:46-82).  Validates every point, builds the per-GPU tables (default window 13;
    /// `load_json_on(path, device, 15)` is the throughput configuration: 107 GiB of table).
    pub fn load_json(path: impl AsRef<Path>) -> Result<Self, LoadSetupError> {
        Self::load_json_on(path, 0, 0)
    }

    pub fn load_json_on(path: impl AsRef<Path>, device: i32, window_bits: i32) -> Result<Self, LoadSetupError> {
        let c = CString::new(path.as_ref().to_string_lossy().as_bytes()).map_err(|e| LoadSetupError::Io(e.to_string()))?;
        let mut handle: *mut kb200_setup = core::ptr::null_mut();
        let rc = unsafe { kb200_setup_load_json(c.as_ptr(), device, window_bits, &mut handle) };
        match rc {
            0 => Ok(Self { handle }),
            -2 => Err(LoadSetupError::Io(last_error())),
            -3 => Err(LoadSetupError::Serde(last_error())),
            -10 => Err(LoadSetupError::InvalidLenG1Lagrange),
            -11 => Err(LoadSetupError::InvalidLenG2Monomial),
            3 => Err(LoadSetupError::Bls(BlsError::ECGroup(ECGroupError::InvalidEncoding))),
            4 => Err(LoadSetupError::Bls(BlsError::ECGroup(ECGroupError::NotOnCurve))),
            5 => Err(LoadSetupError::Bls(BlsError::ECGroup(ECGroupError::NotInGroup))),
            _ => panic!("kateth_b200: engine failure {rc}: {}", last_error()),
        }
    }

    fn check_blob(blob: &[u8]) -> Result<(), Error> {
        // Blob::from_slice rejects the length before anything else (src/blob.rs:27-29); the canonical check is on the GPU
        if blob.len() != BYTES_PER_BLOB { Err(Error::Blob(BlobError::InvalidLen)) } else { Ok(()) }
    }

    /// src/kzg/setup.rs:167-171
    pub fn blob_to_commitment(&self, blob: impl AsRef<[u8]>) -> Result<Bytes48, Error> {
        let blob = blob.as_ref();
        Self::check_blob(blob)?;
        let mut out = [0u8; 48];
        let rc = unsafe { kb200_blob_to_commitment(self.handle, blob.as_ptr(), 1, out.as_mut_ptr(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(out)
    }

    /// src/kzg/setup.rs:177-183
    pub fn blob_proof(&self, blob: impl AsRef<[u8]>, commitment: &Bytes48) -> Result<Bytes48, Error> {
        let blob = blob.as_ref();
        Self::check_blob(blob)?;
        let mut out = [0u8; 48];
        let rc = unsafe { kb200_compute_blob_proof(self.handle, blob.as_ptr(), commitment.as_ptr(), 1, out.as_mut_ptr(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(out)
    }

    /// src/kzg/setup.rs:185-194: (proof, y)
    pub fn proof(&self, blob: impl AsRef<[u8]>, point: &Bytes32) -> Result<(Bytes48, Bytes32), Error> {
        let blob = blob.as_ref();
        Self::check_blob(blob)?;
        let (mut proof, mut y) = ([0u8; 48], [0u8; 32]);
        let rc = unsafe { kb200_compute_proof(self.handle, blob.as_ptr(), point.as_ptr(), 1, proof.as_mut_ptr(), y.as_mut_ptr(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok((proof, y))
    }

    /// src/kzg/setup.rs:96-113 -- argument order as in kateth: proof, commitment, point, eval
    pub fn verify_proof(&self, proof: &Bytes48, commitment: &Bytes48, point: &Bytes32, eval: &Bytes32) -> Result<bool, Error> {
        let mut ok: c_int = 0;
        let rc = unsafe { kb200_verify_proof(self.handle, proof.as_ptr(), commitment.as_ptr(), point.as_ptr(), eval.as_ptr(), &mut ok) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(ok != 0)
    }

    /// src/kzg/setup.rs:208-221
    pub fn verify_blob_proof(&self, blob: impl AsRef<[u8]>, commitment: &Bytes48, proof: &Bytes48) -> Result<bool, Error> {
        let blob = blob.as_ref();
        Self::check_blob(blob)?;
        let mut ok: c_int = 0;
        let rc = unsafe { kb200_verify_blob_proof(self.handle, blob.as_ptr(), commitment.as_ptr(), proof.as_ptr(), &mut ok) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(ok != 0)
    }

    /// src/kzg/setup.rs:247-275.  Length mismatch panics like the reference (`assert_eq!`, :256-257).
    pub fn verify_blob_proof_batch<B: AsRef<[u8]>>(&self, blobs: impl AsRef<[B]>, commitments: impl AsRef<[Bytes48]>,
                                                   proofs: impl AsRef<[Bytes48]>) -> Result<bool, Error> {
        let (blobs, commitments, proofs) = (blobs.as_ref(), commitments.as_ref(), proofs.as_ref());
        assert_eq!(blobs.len(), commitments.len());
        assert_eq!(commitments.len(), proofs.len());
        let mut flat = Vec::with_capacity(blobs.len() * BYTES_PER_BLOB);
        for b in blobs {
            Self::check_blob(b.as_ref())?;
            flat.extend_from_slice(b.as_ref());
        }
        let mut ok: c_int = 0;
        let rc = unsafe {
            kb200_verify_blob_proof_batch(self.handle, flat.as_ptr(), commitments.as_ptr().cast(), proofs.as_ptr().cast(), blobs.len(), &mut ok,
                                          core::ptr::null_mut())
        };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(ok != 0)
    }

    /// Batch forms of commit / prove over `n` contiguous blobs (the throughput extension of the two calls above).
    pub fn blob_to_commitment_many(&self, blobs: &[u8]) -> Result<Vec<Bytes48>, Error> {
        assert_eq!(blobs.len() % BYTES_PER_BLOB, 0);
        let n = blobs.len() / BYTES_PER_BLOB;
        let mut out = vec![[0u8; 48]; n];
        let rc = unsafe { kb200_blob_to_commitment(self.handle, blobs.as_ptr(), n, out.as_mut_ptr().cast(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(out)
    }

    pub fn blob_proof_many(&self, blobs: &[u8], commitments: &[Bytes48]) -> Result<Vec<Bytes48>, Error> {
        assert_eq!(blobs.len(), commitments.len() * BYTES_PER_BLOB);
        let n = commitments.len();
        let mut out = vec![[0u8; 48]; n];
        let rc = unsafe { kb200_compute_blob_proof(self.handle, blobs.as_ptr(), commitments.as_ptr().cast(), n, out.as_mut_ptr().cast(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(out)
    }

    /// Raw handle for the `_device` entry points (blobs resident in HBM, multi-GPU sharding).
    pub fn raw(&self) -> *mut kb200_setup {
        self.handle
    }
}
