//! `kateth::blob::Error` (`/root/reference/src/blob.rs:6-16`).  The blob codec itself lives on the GPU: the engine reads the
//! big-endian field elements straight out of the caller's bytes (`k_blob_check`, `k_msm_recode`).
use crate::bls::FiniteFieldError;

#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Error {
    InvalidFieldElement,
    InvalidLen,
}
impl From<FiniteFieldError> for Error {
    fn from(_err: FiniteFieldError) -> Self { Self::InvalidFieldElement }       // src/blob.rs:12-16
}
