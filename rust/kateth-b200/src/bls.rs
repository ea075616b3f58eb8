//! The slice of kateth's private `bls` module that crosses its public API (`/root/reference/src/bls.rs`): the `Compress` /
//! `Decompress` traits (`:55-76`), the error enums (`:21-50`), `P1` (`Commitment = Proof = P1`, `src/kzg/mod.rs:9-10`) and `Fr`
//! as returned by `Setup::proof`.  kateth's `P1` wraps a Jacobian `blst_p1`; the engine returns the unique 48-byte compressed
//! encoding of the same group element, so this `P1` wraps those bytes -- `compress` is a copy, equality is byte equality, and
//! `decompress` validates through the engine (curve + subgroup check on the GPU, same error classes as blst).
use crate::ffi;

#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum FiniteFieldError { InvalidEncoding, NotInFiniteField }              // src/bls.rs:21-30
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum ECGroupError { InvalidEncoding, NotInGroup, NotOnCurve }            // src/bls.rs:32-38
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub enum Error { ECGroup(ECGroupError), FiniteField(FiniteFieldError) }      // src/bls.rs:40-50

impl From<ECGroupError> for Error { fn from(e: ECGroupError) -> Self { Self::ECGroup(e) } }
impl From<FiniteFieldError> for Error { fn from(e: FiniteFieldError) -> Self { Self::FiniteField(e) } }

/// src/bls.rs:55-65
pub trait Compress {
    const COMPRESSED: usize;
    fn compress(&self, buf: impl AsMut<[u8]>) -> Result<(), &'static str>;
}
/// src/bls.rs:70-76
pub trait Decompress: Sized {
    type Error;
    fn decompress(compressed: impl AsRef<[u8]>) -> Result<Self, Self::Error>;
}

/// A point of G1 in its canonical 48-byte ZCash encoding.
#[derive(Clone, Copy, Debug, PartialEq, Eq)]
pub struct P1 { pub(crate) bytes: [u8; 48] }

impl P1 {
    pub const BITS: usize = 384;
    pub const BYTES: usize = Self::BITS / 8;
    pub(crate) fn from_engine(bytes: [u8; 48]) -> Self { Self { bytes } }
}
impl Default for P1 {
    fn default() -> Self { let mut b = [0u8; 48]; b[0] = 0xc0; Self { bytes: b } }      // the point at infinity, like blst_p1::default()
}
impl Compress for P1 {
    const COMPRESSED: usize = 48;
    fn compress(&self, mut buf: impl AsMut<[u8]>) -> Result<(), &'static str> {
        if buf.as_mut().len() < Self::COMPRESSED { return Err("insufficient buffer length"); }       // src/bls.rs:494-497
        buf.as_mut()[..48].copy_from_slice(&self.bytes);
        Ok(())
    }
}
impl Decompress for P1 {
    type Error = ECGroupError;
    /// blst_p1_uncompress + blst_p1_affine_in_g1 (src/bls.rs:508-531) on the GPU.  (kateth reads 48 bytes behind the pointer without
    /// looking at the length, `:514`; a shorter slice is an InvalidEncoding here.)
    fn decompress(compressed: impl AsRef<[u8]>) -> Result<Self, Self::Error> {
        let c = compressed.as_ref();
        if c.len() < 48 { return Err(ECGroupError::InvalidEncoding); }
        let mut st: i32 = 0;
        let rc = unsafe { ffi::kb200_g1_validate(c.as_ptr(), 1, 0, &mut st) };
        if rc < 0 { panic!("kateth_b200: engine failure {rc}: {}", ffi::last_error()); }
        match st {
            0 => { let mut b = [0u8; 48]; b.copy_from_slice(&c[..48]); Ok(Self { bytes: b }) }
            3 => Err(ECGroupError::InvalidEncoding),
            4 => Err(ECGroupError::NotOnCurve),
            5 => Err(ECGroupError::NotInGroup),
            other => unreachable!("{other}"),
        }
    }
}

/// An element of the scalar field as its canonical 32-byte big-endian encoding (what kateth's tests read off `Setup::proof`'s `y`).
#[derive(Clone, Copy, Debug, Default, PartialEq, Eq)]
pub struct Fr { pub(crate) be: [u8; 32] }

impl Fr {
    pub const BITS: usize = 256;
    pub const BYTES: usize = Self::BITS / 8;
    const MODULUS_BE: [u8; 32] = [0x73, 0xed, 0xa7, 0x53, 0x29, 0x9d, 0x7d, 0x48, 0x33, 0x39, 0xd8, 0x08, 0x09, 0xa1, 0xd8, 0x05, 0x53, 0xbd, 0xa4, 0x02, 0xff, 0xfe,
                                  0x5b, 0xfe, 0xff, 0xff, 0xff, 0xff, 0x00, 0x00, 0x00, 0x01];
    /// src/bls.rs:130-139: the encoding must be 32 bytes and canonical (< r)
    pub fn from_be_slice(bytes: impl AsRef<[u8]>) -> Result<Self, FiniteFieldError> {
        let b = bytes.as_ref();
        if b.len() != Self::BYTES { return Err(FiniteFieldError::InvalidEncoding); }
        if b >= &Self::MODULUS_BE[..] { return Err(FiniteFieldError::NotInFiniteField); }
        let mut be = [0u8; 32]; be.copy_from_slice(b);
        Ok(Self { be })
    }
    /// src/bls.rs:141-149
    pub fn to_be_bytes(self) -> [u8; Self::BYTES] { self.be }
    pub fn to_le_bytes(self) -> [u8; Self::BYTES] { let mut o = self.be; o.reverse(); o }
}
