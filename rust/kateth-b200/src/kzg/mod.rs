//! `kateth::kzg` (`/root/reference/src/kzg/mod.rs:1-35`): the type names and the error enum, unchanged.
use crate::{blob, bls};

mod setup;

pub type Proof = bls::P1;
pub type Commitment = bls::P1;

pub type Bytes32 = [u8; 32];
pub type Bytes48 = [u8; 48];

#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Error {
    Blob(blob::Error),
    Bls(bls::Error),
}
impl From<blob::Error> for Error { fn from(value: blob::Error) -> Self { Self::Blob(value) } }
impl From<bls::Error> for Error { fn from(value: bls::Error) -> Self { Self::Bls(value) } }

pub use setup::{LoadSetupError, MultiSetup, Setup};
