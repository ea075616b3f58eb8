//! `extern "C"` declarations of include/kateth_b200.h (the part the wrapper uses).
#![allow(non_camel_case_types)]
use core::ffi::{c_char, c_int, c_void};

#[repr(C)]
pub struct kb200_setup {
    _private: [u8; 0],
}
#[repr(C)]
pub struct kb200_group {
    _private: [u8; 0],
}

pub const KB200_GROUP_ID_BYTES: usize = 128;

extern "C" {
    pub fn kb200_setup_load_json(path: *const c_char, device: c_int, window_bits: c_int, out: *mut *mut kb200_setup) -> c_int;
    pub fn kb200_setup_free(s: *mut kb200_setup);
    pub fn kb200_last_error() -> *const c_char;
    pub fn kb200_g1_validate(in48: *const u8, n: usize, device: c_int, status: *mut i32) -> c_int;
    pub fn kb200_blob_to_commitment(s: *mut kb200_setup, blobs: *const u8, n: usize, out48: *mut u8, status: *mut i32) -> c_int;
    pub fn kb200_check_blobs(s: *mut kb200_setup, blobs: *const u8, n: usize, status: *mut i32) -> c_int;
    pub fn kb200_compute_blob_proof(s: *mut kb200_setup, blobs: *const u8, commitments48: *const u8, n: usize, out48: *mut u8, status: *mut i32) -> c_int;
    pub fn kb200_compute_proof(s: *mut kb200_setup, blobs: *const u8, z32: *const u8, n: usize, out_proof48: *mut u8, out_y32: *mut u8, status: *mut i32) -> c_int;
    pub fn kb200_verify_proof(s: *mut kb200_setup, proof48: *const u8, commitment48: *const u8, z32: *const u8, y32: *const u8, ok: *mut c_int) -> c_int;
    pub fn kb200_verify_proof_many(s: *mut kb200_setup, proofs48: *const u8, commitments48: *const u8, z32: *const u8, y32: *const u8, n: usize, ok: *mut i32,
                                   status: *mut i32) -> c_int;
    pub fn kb200_verify_blob_proof(s: *mut kb200_setup, blob: *const u8, commitment48: *const u8, proof48: *const u8, ok: *mut c_int) -> c_int;
    pub fn kb200_verify_blob_proof_batch(s: *mut kb200_setup, blobs: *const u8, commitments48: *const u8, proofs48: *const u8, n: usize, ok: *mut c_int,
                                         status: *mut i32) -> c_int;
    // several B200s behind one call: replicas, contiguous shards, the NCCL all-gather inside the library
    pub fn kb200_group_create_json(path: *const c_char, devices: *const c_int, ndev: c_int, window_bits: c_int, out: *mut *mut kb200_group) -> c_int;
    pub fn kb200_group_unique_id(id128: *mut u8) -> c_int;
    pub fn kb200_group_create_rank(local: *mut kb200_setup, rank: c_int, world: c_int, id128: *const u8, out: *mut *mut kb200_group) -> c_int;
    pub fn kb200_group_free(g: *mut kb200_group);
    pub fn kb200_group_blob_to_commitment(g: *mut kb200_group, blobs: *const u8, n: usize, out48: *mut u8, status: *mut i32) -> c_int;
    pub fn kb200_group_compute_blob_proof(g: *mut kb200_group, blobs: *const u8, commitments48: *const u8, n: usize, out48: *mut u8, status: *mut i32) -> c_int;
    pub fn kb200_group_verify_blob_proof_batch(g: *mut kb200_group, blobs: *const u8, commitments48: *const u8, proofs48: *const u8, n_local: usize,
                                               index_offset: usize, n_total: usize, ok: *mut c_int, status: *mut i32) -> c_int;
    // data already in HBM (device pointers): see the header for the rest of the `_device` family
    pub fn kb200_blob_to_commitment_device(s: *mut kb200_setup, d_blobs: *const c_void, n: usize, d_out48: *mut c_void, d_status: *mut i32) -> c_int;
    pub fn kb200_compute_blob_proof_device(s: *mut kb200_setup, d_blobs: *const c_void, d_commitments48: *const c_void, n: usize, d_out48: *mut c_void,
                                           d_status: *mut i32) -> c_int;
}

pub fn last_error() -> String {
    unsafe { std::ffi::CStr::from_ptr(kb200_last_error()).to_string_lossy().into_owned() }
}
