/*
 * kateth_b200.h -- C ABI of the B200-native KZG engine (libkateth_b200.so).
 *
 * Drop-in boundary for kateth's EIP-4844 path.  kateth has no FFI of its own; its public surface is the seven
 * `pub fn`s of `Setup<4096, 65>` (/root/reference/src/kzg/setup.rs:46,96,167,177,185,208,247) plus
 * `Compress::compress` (/root/reference/src/bls.rs:55-65) that callers use to obtain bytes.  Each entry point
 * below names the reference function it replaces; INTEGRATION.md shows the Rust `extern "C"` block and the
 * `Setup` wrapper a kateth maintainer would add.
 *
 * Conventions
 *   - plain pointers and sizes only; the caller owns every host buffer, the library copies.
 *   - blobs: n contiguous 131072-byte blobs (4096 big-endian 32-byte field elements each).
 *   - G1 points travel as the 48-byte ZCash compressed encoding (the parity criterion of kateth's tests);
 *     scalars as 32-byte big-endian.
 *   - every function returns 0 on success, a positive KB200_E_* code mirroring kateth's error enums
 *     (kzg::Error / blob::Error / bls::Error, /root/reference/src/kzg/mod.rs:15-31, src/blob.rs:6-16,
 *     src/bls.rs:21-50), or a negative code for CUDA / internal failures.  `Ok(false)` is *ok = 0 with return 0.
 *   - batch calls also fill `status[i]` per item (0 or a KB200_E_* code); the function's return value is the first
 *     non-zero status in kateth's evaluation order (all blobs, then all commitments, then all proofs).
 *   - `*_device` variants take device pointers (data already resident in HBM, 16-byte aligned) and run on the
 *     stream set by kb200_set_stream (default: the legacy default stream); they return after enqueueing AND
 *     synchronising that stream unless noted.
 *   - no function aborts or throws across the ABI.
 *   - a handle may be shared by any number of host threads (kateth's Setup is Send + Sync, src/kzg/setup.rs:84-275): calls on one
 *     handle are serialised on an internal lock -- one batch call already fills the GPU; kb200_last_error is per thread.
 *   - there is NO CPU fallback: without a CUDA device kb200_setup_load_* returns KB200_E_CUDA.
 */
#ifndef KATETH_B200_H
#define KATETH_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KB200_FIELD_ELEMENTS_PER_BLOB 4096
#define KB200_BYTES_PER_BLOB 131072
#define KB200_BYTES_PER_G1 48
#define KB200_BYTES_PER_G2 96
#define KB200_BYTES_PER_SCALAR 32
#define KB200_RECORD_BYTES 336 /* per-GPU partial of batch verification: 3 affine G1 sums + 1 Fr + the shard's first blob / commitment / proof error */

/* status codes */
#define KB200_OK 0
#define KB200_E_INVALID_LEN 1              /* blob::Error::InvalidLen */
#define KB200_E_INVALID_FIELD_ELEMENT 2    /* blob::Error::InvalidFieldElement */
#define KB200_E_G1_INVALID_ENCODING 3      /* ECGroupError::InvalidEncoding */
#define KB200_E_G1_NOT_ON_CURVE 4          /* ECGroupError::NotOnCurve */
#define KB200_E_G1_NOT_IN_GROUP 5          /* ECGroupError::NotInGroup */
#define KB200_E_SCALAR_INVALID_ENCODING 6  /* FiniteFieldError::InvalidEncoding */
#define KB200_E_SCALAR_NOT_IN_FIELD 7      /* FiniteFieldError::NotInFiniteField */
#define KB200_E_CUDA (-1)                  /* CUDA runtime error (kb200_last_error has the text) */
#define KB200_E_IO (-2)                    /* LoadSetupError::Io */
#define KB200_E_PARSE (-3)                 /* LoadSetupError::Serde */
#define KB200_E_SETUP_G1_LEN (-10)         /* LoadSetupError::InvalidLenG1Lagrange */
#define KB200_E_SETUP_G2_LEN (-11)         /* LoadSetupError::InvalidLenG2Monomial */
#define KB200_E_SETUP_G1_MONOMIAL (-12)     /* the file's g1_monomial points do not match its g1_lagrange points (kb200_setup_check_g1_monomial) */
#define KB200_E_ARG (-20)                  /* null pointer / bad argument */
#define KB200_E_NCCL (-30)                 /* NCCL missing (dlopen libnccl.so.2) or a collective failed (kb200_last_error has the text) */

typedef struct kb200_setup kb200_setup;

/* ---- Setup::load_json (/root/reference/src/kzg/setup.rs:46-82) ----
 * Parses {"g1_lagrange": [4096 x "0x" + 96 hex], "g2_monomial": [65 x "0x" + 192 hex]}, validates every point
 * (curve + subgroup; G1 on the GPU, G2 on the host), bit-reverses the Lagrange points, builds the roots-of-unity
 * table and the fixed-base window table on `device`.  window_bits: 0 = default = the widest table the device's FREE memory takes with
 * room for the workspace of 4096-blob batches: 16 (the throughput configuration: GLV, 100 GiB) on an idle B200, else 13 (28.6 GiB), 9
 * or 6; kb200_get_info reports the choice.  4..15 = that many bits (15 -> 17 windows, 107 GiB).  The top window is unsigned; the
 * others hold signed-digit magnitudes.
 * If the file also has a "g1_monomial" array (consensus-spec presets) it is cross-checked: see kb200_setup_check_g1_monomial.
 * 16, or 100 + c for c in 4..16, selects the GLV form: every scalar is split along the G1 endomorphism into two 128-bit
 * halves that share one table of ceil(128 / c) windows (16 -> 8 windows, 100 GiB, 16 gathered points per scalar instead of 17). */
int kb200_setup_load_json(const char* path, int device, int window_bits, kb200_setup** out);
/* Same from c-kzg-4844's trusted_setup.txt ("4096", "65", 4096 G1 Lagrange lines, 65 G2 monomial lines, optionally 4096 G1 monomial
 * lines, which are cross-checked when all are there); kateth itself only reads the JSON form. */
int kb200_setup_load_txt(const char* path, int device, int window_bits, kb200_setup** out);
/* Same from decoded bytes: n_g1 x 48 bytes, n_g2 x 96 bytes. */
int kb200_setup_load_raw(const uint8_t* g1_lagrange48, size_t n_g1, const uint8_t* g2_monomial96, size_t n_g2, int device, int window_bits,
                         kb200_setup** out);
void kb200_setup_free(kb200_setup* s);
/* g1_lagrange against the monomial basis of the same ceremony: for every k < n the evaluations of X^k over the domain (in blob order) must
 * commit to g1_monomial[k] = [tau^k]G1.  One batch commit on the GPU; returns 0, or KB200_E_SETUP_G1_MONOMIAL with the first k in
 * kb200_last_error.  Called by the loaders when the file carries the array. */
int kb200_setup_check_g1_monomial(kb200_setup* s, const uint8_t* g1_monomial48, size_t n);
/* On-disk cache of the window table, process-wide, off by default (NULL / "" switches it off; the KB200_TABLE_CACHE environment variable
 * names a directory too).  File name = sha256(g1_lagrange | window form | layout version); a 64-bit checksum of the table, computed on the GPU
 * after the upload, is checked against the header, and a missing / torn / edited file is rebuilt and rewritten.  Whether it pays depends on
 * the storage: the GPU rebuilds the 100 GiB table in 4.8 s, which a file has to beat. */
int kb200_set_table_cache(const char* dir);
/* Stream for all later work on this handle (a cudaStream_t passed as void*; NULL = default stream). */
int kb200_set_stream(kb200_setup* s, void* cuda_stream);
const char* kb200_last_error(void);
/* Page-locked host buffers for the host-buffer entry points (cudaMemcpyAsync from pageable memory is staged and slower);
 * optional: any host pointer is accepted everywhere. */
int kb200_alloc_pinned(size_t bytes, void** out);
int kb200_free_pinned(void* p);

/* ---- P1::decompress (src/bls.rs:505-531: curve check, then subgroup check) for n 48-byte encodings, no setup needed: status[i] = 0 or
 *      KB200_E_G1_INVALID_ENCODING / _NOT_ON_CURVE / _NOT_IN_GROUP.  kateth's `Decompress` is an associated function of the point type
 *      (Commitment::decompress(bytes)); the Rust wrapper's P1 calls this. ---- */
int kb200_g1_validate(const uint8_t* in48, size_t n, int device, int32_t* status);

/* ---- Setup::blob_to_commitment (src/kzg/setup.rs:167-171) + Compress::compress; n = 1 is the reference call ---- */
int kb200_blob_to_commitment(kb200_setup* s, const uint8_t* blobs, size_t n, uint8_t* out48, int32_t* status);
int kb200_blob_to_commitment_device(kb200_setup* s, const void* d_blobs, size_t n, void* d_out48, int32_t* d_status);

/* NO counterpart in the reference: added for callers that cannot call cudaMalloc (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  Same bytes as kb200_blob_to_commitment_device for n <=
 * KB200_NOALLOC_MAX_BLOBS blobs (more: KB200_E_ARG and a message in kb200_last_error), but the call allocates nothing on host or device:
 * the MSM's partial sums live in a static 384 KiB device array that every GPU loading the library holds until exit.  Enqueue-only on the
 * handle's stream (kb200_set_stream): it returns 0 once the work is queued; d_status[i] (0, or the error class of blob i, whose
 * output is then 48 zero bytes) and d_out48 are valid when that stream has passed them.  All calls on one GPU share the array and it keeps
 * the last call's contents: calls that overlap on the device (two handles or two streams on the same GPU) race. */
#define KB200_NOALLOC_MAX_BLOBS 64
int kb200_blob_to_commitment_device_noalloc(kb200_setup* s, const void* d_blobs, size_t n, void* d_out48, int32_t* d_status);

/* Blob::from_slice alone (src/blob.rs:26-37) over n full-length blobs: status[i] = 0 or KB200_E_INVALID_FIELD_ELEMENT; returns the first
 * non-zero status.  For wrappers whose callers pass blobs of unequal byte lengths (kateth: `impl AsRef<[u8]>` per blob) and that must
 * report the first error in blob order -- an element error of blob 3 before the length error of blob 7. */
int kb200_check_blobs(kb200_setup* s, const uint8_t* blobs, size_t n, int32_t* status);

/* ---- Setup::blob_proof (src/kzg/setup.rs:177-183) + compress ---- */
int kb200_compute_blob_proof(kb200_setup* s, const uint8_t* blobs, const uint8_t* commitments48, size_t n, uint8_t* out48, int32_t* status);
int kb200_compute_blob_proof_device(kb200_setup* s, const void* d_blobs, const void* d_commitments48, size_t n, void* d_out48, int32_t* d_status);

/* ---- Setup::proof (src/kzg/setup.rs:185-194): (proof, y) at caller-supplied points z ---- */
int kb200_compute_proof(kb200_setup* s, const uint8_t* blobs, const uint8_t* z32, size_t n, uint8_t* out_proof48, uint8_t* out_y32, int32_t* status);
int kb200_compute_proof_device(kb200_setup* s, const void* d_blobs, const void* d_z32, size_t n, void* d_out_proof48, void* d_out_y32,
                               int32_t* d_status);

/* ---- Setup::verify_proof (src/kzg/setup.rs:96-113); argument order proof, commitment, point, eval as in kateth ---- */
int kb200_verify_proof(kb200_setup* s, const uint8_t proof48[48], const uint8_t commitment48[48], const uint8_t z32[32], const uint8_t y32[32],
                       int* ok);
/* ---- private Setup::verify_proof_batch (src/kzg/setup.rs:115-161): one RLC check over n (proof, commitment, z, y) ---- */
int kb200_verify_proof_batch(kb200_setup* s, const uint8_t* proofs48, const uint8_t* commitments48, const uint8_t* z32, const uint8_t* y32, size_t n,
                             int* ok, int32_t* status);
/* ---- n independent Setup::verify_proof calls in one launch: PER-ITEM verdicts from a two-pairing check run on the device (one thread
 *      per item: Miller loops against the line-precomputed [tau]_2 and G2, final exponentiation; bls::verify_pairings, src/bls.rs:572-598,
 *      rearranged to e(-proof, [tau]_2) e(C - [y]G1 + [z]proof, G2) == 1).  ok[i] = 1 / 0; status[i] = that item's decoding error (proof,
 *      commitment, point, eval order) with ok[i] = 0.  The return value is 0 unless the call itself failed. ---- */
int kb200_verify_proof_many(kb200_setup* s, const uint8_t* proofs48, const uint8_t* commitments48, const uint8_t* z32, const uint8_t* y32, size_t n,
                            int32_t* ok, int32_t* status);
/* ---- Setup::verify_blob_proof (src/kzg/setup.rs:208-221) ---- */
int kb200_verify_blob_proof(kb200_setup* s, const uint8_t* blob, const uint8_t commitment48[48], const uint8_t proof48[48], int* ok);
/* ---- Setup::verify_blob_proof_batch (src/kzg/setup.rs:247-275); n = 0 -> *ok = 1 ---- */
int kb200_verify_blob_proof_batch(kb200_setup* s, const uint8_t* blobs, const uint8_t* commitments48, const uint8_t* proofs48, size_t n, int* ok,
                                  int32_t* status);
int kb200_verify_blob_proof_batch_device(kb200_setup* s, const void* d_blobs, const void* d_commitments48, const void* d_proofs48, size_t n, int* ok,
                                         int32_t* d_status);

/* ---- multi-GPU batch verification: each rank computes the partial of its contiguous shard
 *      [index_offset, index_offset + n_local) of a batch of n_total (n_local may be 0), the KB200_RECORD_BYTES records are
 *      exchanged with one all-gather (NCCL over NVLink: kb200_group_* below does it inside the library; a caller with its own
 *      communicator may do it itself), and any rank finishes.  d_record: device memory.
 *      A shard with an invalid blob / commitment / proof returns that error AND writes it into its record, the failing items
 *      contribute nothing to the sums, and kb200_verify_batch_finish_device returns -- on every rank -- the error kateth would:
 *      the first blob error in rank order, else the first commitment error, else the first proof error. ---- */
int kb200_verify_blob_proof_batch_partial_device(kb200_setup* s, const void* d_blobs, const void* d_commitments48, const void* d_proofs48,
                                                 size_t n_local, size_t index_offset, size_t n_total, void* d_record, int32_t* d_status);
/* Same with the shard's inputs in host memory (any n_local; blobs are uploaded in chunks); the record is still written to device memory. */
int kb200_verify_blob_proof_batch_partial(kb200_setup* s, const uint8_t* blobs, const uint8_t* commitments48, const uint8_t* proofs48, size_t n_local,
                                          size_t index_offset, size_t n_total, void* d_record, int32_t* status);
/* The same exchange for batches of (proof, commitment, z, y) tuples (private Setup::verify_proof_batch, src/kzg/setup.rs:115-161): one record per
 * shard, error classes proof / commitment / z / y. */
int kb200_verify_proof_batch_partial_device(kb200_setup* s, const void* d_proofs48, const void* d_commitments48, const void* d_z32, const void* d_y32,
                                            size_t n_local, size_t index_offset, size_t n_total, void* d_record, int32_t* d_status);
int kb200_verify_proof_batch_partial(kb200_setup* s, const uint8_t* proofs48, const uint8_t* commitments48, const uint8_t* z32, const uint8_t* y32,
                                     size_t n_local, size_t index_offset, size_t n_total, void* d_record, int32_t* status);
/* Commit + blob proof + the batch-verification shard of a FRESH batch over one upload of the blobs: the three pipelines of
 * kb200_blob_to_commitment, kb200_compute_blob_proof and kb200_verify_blob_proof_batch_partial, each complete (the verification decodes the
 * 48-byte outputs again and recomputes every challenge and evaluation), run on the bytes of a single host-to-device copy per chunk.
 * Outputs: n commitments, n proofs (host), the shard's record (device). */
int kb200_commit_prove_verify_partial(kb200_setup* s, const uint8_t* blobs, size_t n_local, size_t index_offset, size_t n_total, uint8_t* out_commitments48,
                                      uint8_t* out_proofs48, void* d_record, int32_t* status);
/* d_records: n_records x KB200_RECORD_BYTES in device memory. */
int kb200_verify_batch_finish_device(kb200_setup* s, const void* d_records, size_t n_records, int* ok);
/* The two compressed G1 operands of the final pairing check of the last finish call (for diffing intermediate sums). */
int kb200_last_pairing_operands(kb200_setup* s, uint8_t lhs48[48], uint8_t rhs48[48]);

/* ---- the same surface over several B200s (SURVEY.md 8b "Setup: devices, ndev", 8e): one replica of the setup per GPU, batches cut
 *      into contiguous shards, one host thread + stream per GPU inside the library.  commit / prove need no communication; batch
 *      verification has ONE exchange step, an ncclAllGather of the KB200_RECORD_BYTES records over NVLink done here (NCCL is bound at
 *      run time with dlopen("libnccl.so.2"); a single-GPU caller never needs it), then ONE host pairing per process: one call,
 *      one verdict, as Setup::verify_blob_proof_batch (src/kzg/setup.rs:247-275).  A group is safe for concurrent callers (calls on
 *      one group are serialised: every rank must issue its collectives in the same order).
 *      Two ways to form a group:
 *        kb200_group_create_json   one process drives the `ndev` GPUs it names (devices == NULL: 0 .. ndev-1); the replicas are loaded
 *                                  concurrently (Setup::load_json per device) and owned by the group; ncclCommInitAll.
 *        kb200_group_create_rank   one process per GPU (torchrun / MPI / one service per device): rank 0 calls kb200_group_unique_id and
 *                                  ships the 128 bytes to the other ranks by any means; every rank passes its own loaded handle
 *                                  (not owned by the group) and joins with ncclCommInitRank. ---- */
typedef struct kb200_group kb200_group;
#define KB200_GROUP_ID_BYTES 128
int kb200_group_create_json(const char* path, const int* devices, int ndev, int window_bits, kb200_group** out);
int kb200_group_unique_id(uint8_t id128[KB200_GROUP_ID_BYTES]);
int kb200_group_create_rank(kb200_setup* local, int rank, int world, const uint8_t id128[KB200_GROUP_ID_BYTES], kb200_group** out);
void kb200_group_free(kb200_group* g);
/* local_replicas: GPUs driven by this process; world: ranks of the whole group; first_rank: rank of local replica 0; nccl_version: 0 when
 * the group has one rank (no communicator).  Any out pointer may be NULL. */
int kb200_group_info(kb200_group* g, int* local_replicas, int* world, int* first_rank, int* nccl_version);
kb200_setup* kb200_group_replica(kb200_group* g, int i); /* the i-th local handle (still owned by the group when it loaded it) */
/* Host-buffer calls: n items, cut contiguously over the replicas of THIS process; outputs and statuses in item order. */
int kb200_group_blob_to_commitment(kb200_group* g, const uint8_t* blobs, size_t n, uint8_t* out48, int32_t* status);
int kb200_group_compute_blob_proof(kb200_group* g, const uint8_t* blobs, const uint8_t* commitments48, size_t n, uint8_t* out48, int32_t* status);
int kb200_group_compute_proof(kb200_group* g, const uint8_t* blobs, const uint8_t* z32, size_t n, uint8_t* out_proof48, uint8_t* out_y32, int32_t* status);
/* One batch of n_total items over the whole group: this process holds items [index_offset, index_offset + n_local) (a one-process group:
 * index_offset = 0, n_local = n_total).  Every rank must make the call; every rank gets the verdict, or the batch's error in kateth's
 * order (the errors travel inside the records). */
int kb200_group_verify_blob_proof_batch(kb200_group* g, const uint8_t* blobs, const uint8_t* commitments48, const uint8_t* proofs48, size_t n_local,
                                        size_t index_offset, size_t n_total, int* ok, int32_t* status);
/* kb200_commit_prove_verify_partial on every replica + the exchange + one pairing: commitments, proofs and the verdict of a fresh batch. */
int kb200_group_commit_prove_verify(kb200_group* g, const uint8_t* blobs, size_t n_local, size_t index_offset, size_t n_total, uint8_t* out_commitments48,
                                    uint8_t* out_proofs48, int* ok, int32_t* status);
/* Setup::verify_proof_batch over the whole group (tuples instead of blobs), and n independent verify_proof calls with per-item verdicts from the
 * device pairing, cut over this process's replicas. */
int kb200_group_verify_proof_batch(kb200_group* g, const uint8_t* proofs48, const uint8_t* commitments48, const uint8_t* z32, const uint8_t* y32,
                                   size_t n_local, size_t index_offset, size_t n_total, int* ok, int32_t* status);
int kb200_group_verify_proof_many(kb200_group* g, const uint8_t* proofs48, const uint8_t* commitments48, const uint8_t* z32, const uint8_t* y32, size_t n,
                                  int32_t* ok, int32_t* status);
int kb200_group_verify_proof_batch_device(kb200_group* g, const void* const* d_proofs48, const void* const* d_commitments48, const void* const* d_z32,
                                          const void* const* d_y32, const size_t* n_per_replica, size_t index_offset, size_t n_total, int* ok,
                                          int32_t* const* d_status);
/* The same with the shards already in HBM: entry i of each array belongs to local replica i (device pointers on that replica's GPU),
 * n_per_replica[i] items each, laid out one after the other from index_offset on. */
int kb200_group_verify_blob_proof_batch_device(kb200_group* g, const void* const* d_blobs, const void* const* d_commitments48,
                                               const void* const* d_proofs48, const size_t* n_per_replica, size_t index_offset, size_t n_total, int* ok,
                                               int32_t* const* d_status);

/* ---- random challenge of the batch checks.  0 (default) = kateth's: r = SHA-256("RCKZGBATCH___V1_" | u128(4096) | u128(n))
 *      mod r and weights [r, r, r^2, ...] (src/kzg/setup.rs:127-136, src/bls.rs:169-187) -- bit-for-bit parity with the reference.
 *      1 = the consensus-spec verify_kzg_proof_batch transcript: r = hash_to_bls_field(domain | u64(4096) | u64(n) |
 *      (commitment | z | y | proof) per item), weights [1, r, r^2, ...]; sound against adaptive inputs.  Verdicts agree on every
 *      test vector.  In mode 1 a batch must be checked in one call of any size (the sharded *_partial entry points return
 *      KB200_E_ARG: the transcript needs every item in one place). ---- */
int kb200_set_batch_challenge(kb200_setup* s, int mode);

/* ---- introspection ---- */
typedef struct {
    int window_bits, windows;
    uint64_t table_bytes;      /* fixed-base table resident in HBM */
    uint64_t kernel_launches;  /* kernels launched by this handle since load */
    int device, sm_count;
    double setup_seconds;
    int last_msm_batch_affine; /* 1 when the most recent MSM ran as k_msm_ba (batch-affine), 0 for k_msm (XYZZ) */
    int ba_grid;               /* persistent grid of k_msm_ba (resident CTAs) */
    int glv;                   /* 1: GLV table (windows = windows per 128-bit half, 2 * windows points per scalar) */
    int table_from_cache;      /* 1: the table was read from the on-disk cache instead of being built */
} kb200_info;
int kb200_get_info(kb200_setup* s, kb200_info* info);
/* CUDA-event timing of the dominant kernel (the fixed-base MSM) for roofline reporting: switch on, run, read the summed
 * device time, launch count and number of 4096-point MSMs covered since the last read. */
int kb200_set_profiling(kb200_setup* s, int on);
/* Shape of the fixed-base MSM: 0 = automatic (batch-affine summation, msm_ba.cuh, once the batch fills the GPU with one
 * CTA per blob; XYZZ accumulation with several CTAs per blob below that), 1 = always XYZZ (k_msm), 2 = always batch-affine
 * (k_msm_ba).  Both produce the same bytes; the switch exists for the parity tests and for profiling. */
int kb200_set_msm_mode(kb200_setup* s, int mode);
/* How a shard of a batch check forms its three sums (verify.cuh): 0 = automatic (the bucket method: measured faster at every batch size,
 * a single verify_proof included), 1 = one 128-step ladder per term (k_rlc), 2 = buckets (k_rlc_bucket).  Same record bytes; for the
 * parity tests and profiling.
 * Environment: KB200_RLC_MODE. */
int kb200_set_rlc_mode(kb200_setup* s, int mode);
int kb200_get_msm_profile(kb200_setup* s, double* ms_total, uint64_t* launches, uint64_t* msms);

/* ---- unit-test hooks (device field arithmetic against the oracle; not part of the kateth surface) ---- */
int kb200_test_fp_mul(const uint32_t* a, const uint32_t* b, uint32_t* out, size_t n); /* n x 12 limbs, Montgomery */
int kb200_test_fr_mul(const uint32_t* a, const uint32_t* b, uint32_t* out, size_t n); /* n x 8 limbs */
/* by_order != 0: subgroup membership by the [r]P == infinity definition instead of the endomorphism test */
int kb200_test_g1_decompress(kb200_setup* s, const uint8_t* in48, size_t n, int32_t* status, uint8_t* recompressed48, int by_order);
int kb200_test_mac_peak(int iters, double* tmacs_wide, double* tmacs_lohi);            /* K0: IMAD microbenchmark */
/* host pairing alone: e(a1, [tau]_2) == e(b1, G2) for plain little-endian affine coordinates (12 x u32 each) */
int kb200_test_host_pairing(kb200_setup* s, const uint32_t* a_xy24, int a_inf, const uint32_t* b_xy24, int b_inf, int* ok);

#ifdef __cplusplus
}
#endif
#endif
