//! `kateth::kzg::Setup<G1, G2>` (`/root/reference/src/kzg/setup.rs:37-276`) over the C ABI: same seven `pub fn`s, same argument
//! order, same return types (`Commitment` / `Proof` = `P1`, `Fr`), same errors; `Setup` is `Send + Sync` like kateth's (the handle is
//! safe for concurrent callers: include/kateth_b200.h).  `MultiSetup` is the same surface over several B200s (`kb200_group_*`).
use core::ffi::c_int;
use std::ffi::CString;
use std::path::Path;

use super::{Bytes32, Bytes48, Commitment, Error, Proof};
use crate::blob::Error as BlobError;
use crate::bls::{ECGroupError, Error as BlsError, FiniteFieldError, Fr, P1};
use crate::ffi;

pub const BYTES_PER_BLOB: usize = 131072;

/// src/kzg/setup.rs:20-28
#[derive(Debug)]
pub enum LoadSetupError {
    Bls(BlsError),
    Io(String),
    Serde(String),
    InvalidLenG1Lagrange,
    InvalidLenG2Monomial,
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
        // CUDA / NCCL / internal failures have no counterpart in kateth's enums, and there is no CPU path to degrade to
        _ => panic!("kateth_b200: engine failure {code}: {}", ffi::last_error()),
    }
}
fn load_error(rc: c_int) -> LoadSetupError {
    match rc {
        -2 => LoadSetupError::Io(ffi::last_error()),
        -3 => LoadSetupError::Serde(ffi::last_error()),
        -10 => LoadSetupError::InvalidLenG1Lagrange,
        -11 => LoadSetupError::InvalidLenG2Monomial,
        3 => LoadSetupError::Bls(BlsError::ECGroup(ECGroupError::InvalidEncoding)),
        4 => LoadSetupError::Bls(BlsError::ECGroup(ECGroupError::NotOnCurve)),
        5 => LoadSetupError::Bls(BlsError::ECGroup(ECGroupError::NotInGroup)),
        _ => panic!("kateth_b200: engine failure {rc}: {}", ffi::last_error()),
    }
}

pub struct Setup<const G1: usize, const G2: usize> {
    handle: *mut ffi::kb200_setup,
}
// kateth's Setup is immutable after load and every method takes &self (src/kzg/setup.rs:84-275); its tests share it through an Arc
// (:323).  The engine handle serialises concurrent callers internally, so both auto traits hold here too.
unsafe impl<const G1: usize, const G2: usize> Send for Setup<G1, G2> {}
unsafe impl<const G1: usize, const G2: usize> Sync for Setup<G1, G2> {}

impl<const G1: usize, const G2: usize> Drop for Setup<G1, G2> {
    fn drop(&mut self) {
        unsafe { ffi::kb200_setup_free(self.handle) }
    }
}

impl Setup<4096, 65> {
    /// src/kzg/setup.rs:46-82.  Validates every point, builds the tables on GPU 0; the window width is chosen from the free HBM
    /// (the 100 GiB throughput table on an idle B200).
    pub fn load_json(path: impl AsRef<Path>) -> Result<Self, LoadSetupError> {
        Self::load_json_on(path, 0, 0)
    }

    /// `window_bits`: 0 = automatic, 4..15 plain windows, 16 or 100 + c the GLV form (include/kateth_b200.h).
    pub fn load_json_on(path: impl AsRef<Path>, device: i32, window_bits: i32) -> Result<Self, LoadSetupError> {
        let c = CString::new(path.as_ref().to_string_lossy().as_bytes()).map_err(|e| LoadSetupError::Io(e.to_string()))?;
        let mut handle: *mut ffi::kb200_setup = core::ptr::null_mut();
        let rc = unsafe { ffi::kb200_setup_load_json(c.as_ptr(), device, window_bits, &mut handle) };
        if rc == 0 { Ok(Self { handle }) } else { Err(load_error(rc)) }
    }

    fn check_len(blob: &[u8]) -> Result<(), BlobError> {
        // Blob::from_slice rejects the length before anything else (src/blob.rs:27-29); the canonical check runs on the GPU
        if blob.len() != BYTES_PER_BLOB { Err(BlobError::InvalidLen) } else { Ok(()) }
    }

    /// src/kzg/setup.rs:96-113 -- argument order as in kateth: proof, commitment, point, eval
    pub fn verify_proof(&self, proof: &Bytes48, commitment: &Bytes48, point: &Bytes32, eval: &Bytes32) -> Result<bool, Error> {
        let mut ok: c_int = 0;
        let rc = unsafe { ffi::kb200_verify_proof(self.handle, proof.as_ptr(), commitment.as_ptr(), point.as_ptr(), eval.as_ptr(), &mut ok) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(ok != 0)
    }

    /// src/kzg/setup.rs:167-171
    pub fn blob_to_commitment(&self, blob: impl AsRef<[u8]>) -> Result<Commitment, BlobError> {
        let blob = blob.as_ref();
        Self::check_len(blob)?;
        let mut out = [0u8; 48];
        let rc = unsafe { ffi::kb200_blob_to_commitment(self.handle, blob.as_ptr(), 1, out.as_mut_ptr(), core::ptr::null_mut()) };
        match rc {
            0 => Ok(P1::from_engine(out)),
            2 => Err(BlobError::InvalidFieldElement),
            _ => panic!("kateth_b200: engine failure {rc}: {}", ffi::last_error()),
        }
    }

    /// src/kzg/setup.rs:177-183
    pub fn blob_proof(&self, blob: impl AsRef<[u8]>, commitment: &Bytes48) -> Result<Proof, Error> {
        let blob = blob.as_ref();
        Self::check_len(blob)?;
        let mut out = [0u8; 48];
        let rc = unsafe { ffi::kb200_compute_blob_proof(self.handle, blob.as_ptr(), commitment.as_ptr(), 1, out.as_mut_ptr(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(P1::from_engine(out))
    }

    /// src/kzg/setup.rs:185-194: (proof, y)
    pub fn proof(&self, blob: impl AsRef<[u8]>, point: &Bytes32) -> Result<(Proof, Fr), Error> {
        let blob = blob.as_ref();
        Self::check_len(blob)?;
        let (mut proof, mut y) = ([0u8; 48], [0u8; 32]);
        let rc = unsafe { ffi::kb200_compute_proof(self.handle, blob.as_ptr(), point.as_ptr(), 1, proof.as_mut_ptr(), y.as_mut_ptr(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok((P1::from_engine(proof), Fr { be: y }))
    }

    /// src/kzg/setup.rs:208-221
    pub fn verify_blob_proof(&self, blob: impl AsRef<[u8]>, commitment: &Bytes48, proof: &Bytes48) -> Result<bool, Error> {
        let blob = blob.as_ref();
        Self::check_len(blob)?;
        let mut ok: c_int = 0;
        let rc = unsafe { ffi::kb200_verify_blob_proof(self.handle, blob.as_ptr(), commitment.as_ptr(), proof.as_ptr(), &mut ok) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(ok != 0)
    }

    /// src/kzg/setup.rs:247-275.  Length mismatch panics like the reference (`assert_eq!`, :256-257); the first error follows kateth's
    /// order: blobs one by one (length, then elements), then commitments, then proofs.
    pub fn verify_blob_proof_batch<B: AsRef<[u8]>>(&self, blobs: impl AsRef<[B]>, commitments: impl AsRef<[Bytes48]>,
                                                   proofs: impl AsRef<[Bytes48]>) -> Result<bool, Error> {
        let (blobs, commitments, proofs) = (blobs.as_ref(), commitments.as_ref(), proofs.as_ref());
        assert_eq!(blobs.len(), commitments.len());
        assert_eq!(commitments.len(), proofs.len());
        let mut flat = Vec::with_capacity(blobs.len() * BYTES_PER_BLOB);
        for b in blobs {
            if b.as_ref().len() != BYTES_PER_BLOB {
                // an element error of an earlier blob precedes this length error
                if !flat.is_empty() {
                    let rc = unsafe { ffi::kb200_check_blobs(self.handle, flat.as_ptr(), flat.len() / BYTES_PER_BLOB, core::ptr::null_mut()) };
                    if rc != 0 { return Err(error_from_status(rc)); }
                }
                return Err(Error::Blob(BlobError::InvalidLen));
            }
            flat.extend_from_slice(b.as_ref());
        }
        let mut ok: c_int = 0;
        let rc = unsafe {
            ffi::kb200_verify_blob_proof_batch(self.handle, flat.as_ptr(), commitments.as_ptr().cast(), proofs.as_ptr().cast(), blobs.len(), &mut ok,
                                               core::ptr::null_mut())
        };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(ok != 0)
    }

    // ---- throughput extensions (not in kateth): n contiguous blobs per call, per-item verdicts from the device pairing ----
    pub fn blob_to_commitment_many(&self, blobs: &[u8]) -> Result<Vec<Commitment>, Error> {
        assert_eq!(blobs.len() % BYTES_PER_BLOB, 0);
        let n = blobs.len() / BYTES_PER_BLOB;
        let mut out = vec![[0u8; 48]; n];
        let rc = unsafe { ffi::kb200_blob_to_commitment(self.handle, blobs.as_ptr(), n, out.as_mut_ptr().cast(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(out.into_iter().map(P1::from_engine).collect())
    }

    pub fn blob_proof_many(&self, blobs: &[u8], commitments: &[Bytes48]) -> Result<Vec<Proof>, Error> {
        assert_eq!(blobs.len(), commitments.len() * BYTES_PER_BLOB);
        let n = commitments.len();
        let mut out = vec![[0u8; 48]; n];
        let rc = unsafe { ffi::kb200_compute_blob_proof(self.handle, blobs.as_ptr(), commitments.as_ptr().cast(), n, out.as_mut_ptr().cast(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(out.into_iter().map(P1::from_engine).collect())
    }

    /// `n` independent `verify_proof` calls in one launch: `Ok(verdict)` or that item's `Err`, per item.
    pub fn verify_proof_many(&self, proofs: &[Bytes48], commitments: &[Bytes48], points: &[Bytes32], evals: &[Bytes32]) -> Vec<Result<bool, Error>> {
        let n = proofs.len();
        assert!(commitments.len() == n && points.len() == n && evals.len() == n);
        let (mut ok, mut st) = (vec![0i32; n], vec![0i32; n]);
        let rc = unsafe {
            ffi::kb200_verify_proof_many(self.handle, proofs.as_ptr().cast(), commitments.as_ptr().cast(), points.as_ptr().cast(), evals.as_ptr().cast(), n,
                                         ok.as_mut_ptr(), st.as_mut_ptr())
        };
        if rc != 0 { panic!("kateth_b200: engine failure {rc}: {}", ffi::last_error()); }
        (0..n).map(|i| if st[i] != 0 { Err(error_from_status(st[i])) } else { Ok(ok[i] != 0) }).collect()
    }

    /// Raw handle for the `_device` entry points (blobs resident in HBM).
    pub fn raw(&self) -> *mut ffi::kb200_setup {
        self.handle
    }
}

/// The same surface over several B200s of one box: one replica of the setup per GPU, batches cut into contiguous shards, the only
/// traffic between GPUs an `ncclAllGather` of the 336-byte records of batch verification, done inside the library.
pub struct MultiSetup {
    group: *mut ffi::kb200_group,
}
unsafe impl Send for MultiSetup {}
unsafe impl Sync for MultiSetup {}
impl Drop for MultiSetup {
    fn drop(&mut self) {
        unsafe { ffi::kb200_group_free(self.group) }
    }
}
impl MultiSetup {
    /// One process drives `devices` (ncclCommInitAll); every replica runs `Setup::load_json`.
    pub fn load_json(path: impl AsRef<Path>, devices: &[i32], window_bits: i32) -> Result<Self, LoadSetupError> {
        let c = CString::new(path.as_ref().to_string_lossy().as_bytes()).map_err(|e| LoadSetupError::Io(e.to_string()))?;
        let mut group: *mut ffi::kb200_group = core::ptr::null_mut();
        let rc = unsafe { ffi::kb200_group_create_json(c.as_ptr(), devices.as_ptr(), devices.len() as c_int, window_bits, &mut group) };
        if rc == 0 { Ok(Self { group }) } else { Err(load_error(rc)) }
    }

    pub fn blob_to_commitment_many(&self, blobs: &[u8]) -> Result<Vec<Commitment>, Error> {
        assert_eq!(blobs.len() % BYTES_PER_BLOB, 0);
        let n = blobs.len() / BYTES_PER_BLOB;
        let mut out = vec![[0u8; 48]; n];
        let rc = unsafe { ffi::kb200_group_blob_to_commitment(self.group, blobs.as_ptr(), n, out.as_mut_ptr().cast(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(out.into_iter().map(P1::from_engine).collect())
    }

    pub fn blob_proof_many(&self, blobs: &[u8], commitments: &[Bytes48]) -> Result<Vec<Proof>, Error> {
        assert_eq!(blobs.len(), commitments.len() * BYTES_PER_BLOB);
        let n = commitments.len();
        let mut out = vec![[0u8; 48]; n];
        let rc = unsafe { ffi::kb200_group_compute_blob_proof(self.group, blobs.as_ptr(), commitments.as_ptr().cast(), n, out.as_mut_ptr().cast(), core::ptr::null_mut()) };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(out.into_iter().map(P1::from_engine).collect())
    }

    /// `Setup::verify_blob_proof_batch` over all the group's GPUs: one call, one verdict.
    pub fn verify_blob_proof_batch(&self, blobs: &[u8], commitments: &[Bytes48], proofs: &[Bytes48]) -> Result<bool, Error> {
        let n = commitments.len();
        assert_eq!(blobs.len(), n * BYTES_PER_BLOB);
        assert_eq!(n, proofs.len());
        let mut ok: c_int = 0;
        let rc = unsafe {
            ffi::kb200_group_verify_blob_proof_batch(self.group, blobs.as_ptr(), commitments.as_ptr().cast(), proofs.as_ptr().cast(), n, 0, n, &mut ok, core::ptr::null_mut())
        };
        if rc != 0 { return Err(error_from_status(rc)); }
        Ok(ok != 0)
    }
}
