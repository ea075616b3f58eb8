//! kateth's public surface (`/root/reference/src/lib.rs:1-7`) backed by the B200 engine (libkateth_b200.so, include/kateth_b200.h):
//! `pub use bls::{Compress, Decompress}; pub mod blob; pub mod kzg;` -- the same three exports, the same type names
//! (`kzg::{Setup, Proof, Commitment, Bytes32, Bytes48, Error}`, `blob::Error`), the same method signatures, so kateth's own
//! conformance tests (`src/kzg/setup.rs:319-463`) and its criterion bench (`benches/kzg.rs`) compile against this crate unchanged:
//! `proof.compress(&mut bytes)`, `Proof::COMPRESSED`, `y.to_be_bytes()`, `Arc<Setup>` shared across threads.
//!
//! NOT COMPILED HERE: the build image has no Rust toolchain.  The `ctypes` mirror `kateth_b200/setup.py` is the
//! binding the parity tests drive; this crate is the same wrapper in the language `north_star` names, over the same C ABI.
mod bls;
mod ffi;

pub use bls::{Compress, Decompress};
pub mod blob;
pub mod kzg;
