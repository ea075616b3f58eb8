// engine.cu -- C ABI of the B200-native KZG engine (include/kateth_b200.h): setup loading, workspace, kernel
// sequencing for commit / prove / verify, and the hand-off of the final two-pairing check to the host bls module.
// Mirrors the seven pub fns of kateth's Setup (/root/reference/src/kzg/setup.rs:46,96,167,177,185,208,247).
// There is no CPU fallback anywhere in this file: every path launches the sm_100a kernels or fails.
#include <cuda_runtime.h>
#include <stdio.h>
#include <ctype.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <chrono>
#include <mutex>

#include "../../include/kateth_b200.h"
#include "verify.cuh"
#include "pairing.cuh"
#include "msm_ba_api.hpp"
#include "host_bls.hpp"

using namespace kb200;

static thread_local char g_err[512] = "";
extern "C" const char* kb200_last_error(void) { return g_err; }
// group.cpp reports through the same per-thread buffer (not part of the public header)
extern "C" void kb200_internal_set_error(const char* msg) { snprintf(g_err, sizeof g_err, "%s", msg ? msg : ""); }
#define CU(call)                                                                                          \
    do {                                                                                                  \
        cudaError_t e_ = (call);                                                                          \
        if (e_ != cudaSuccess) {                                                                          \
            snprintf(g_err, sizeof g_err, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return KB200_E_CUDA;                                                                          \
        }                                                                                                 \
    } while (0)

// grow-only device buffer
struct DBuf {
    void* p = nullptr; size_t cap = 0;
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) { snprintf(g_err, sizeof g_err, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e)); return KB200_E_CUDA; }
        cap = bytes; return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() { return reinterpret_cast<T*>(p); }
};

struct kb200_setup {
    // kateth's Setup is Send + Sync (immutable tables, &self methods, src/kzg/setup.rs:84-275).  The tables here are immutable too, but
    // every call uses the handle's workspace and streams, so concurrent callers are serialised on this lock (recursive: the
    // host-buffer entry points call the device ones).
    std::recursive_mutex mu;
    int device = 0, sm_count = 148;
    cudaStream_t stream = 0;
    G1Affine* d_bases = nullptr;     // g1_lagrange_brp, affine Montgomery   (src/kzg/setup.rs:39)
    Fr* d_roots = nullptr;           // roots_of_unity_brp, Montgomery       (src/kzg/setup.rs:41)
    G1Affine* d_neg_g_pow = nullptr; // d * 2^(8w) * (-G1), w = 0..31, d = 0..255 (verify.cuh, k_rlc_combine)
    G1Affine* d_neg_g_bits = nullptr; // 2^k * (-G1), k = 0..254 (pairing.cuh, k_verify_proof_many)
    PairConsts* d_pair = nullptr;    // Miller-loop lines of [tau]_2 and G2 + Frobenius constants, device copy (pairing.cuh)
    uint4* d_table = nullptr;        // fixed-base window table (msm.cuh)
    MsmTable tab{};
    uint64_t table_bytes = 0;
    host::G2Prepared tau_prep, g2_prep;   // g2_monomial[1] and the G2 generator, Miller lines precomputed
    uint64_t launches = 0;
    double setup_seconds = 0;
    bool table_from_cache = false;
    // workspace (per chunk)
    DBuf blobs, c48, p48, zy32, out48, out32, st_blob, st_a, st_b, status, z, y, prod, indom, q, partials, aff_c, aff_p, rlc_part, rlc_s, rmont,
        records, combine, dig, scr_pts, scr_pre, st_tmp;
    // batch-affine MSM (msm_ba.cuh): 0 = automatic (batches that fill the GPU), 1 = always k_msm, 2 = always k_msm_ba
    int msm_mode = 0, ba_grid = 0, last_ba = 0;
    int rlc_mode = 0;                // sums of batch verification: 0 = automatic (buckets), 1 = ladders (k_rlc), 2 = buckets (k_rlc_bucket)
    DBuf rlc_dig, rlc_wsum, prodh;
    int challenge_mode = 0;          // batch-verification challenge: 0 = kateth's (parity with the reference), 1 = consensus-spec transcript
    DBuf transcript;
    unsigned int* d_counter = nullptr;
    G1* d_noalloc = nullptr;         // device address of g_noalloc_scratch on this handle's GPU (kb200_blob_to_commitment_device_noalloc)
    cudaStream_t aux_stream = nullptr;      // independent light kernels of one call run beside each other (verify_prepare)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    cudaStream_t copy_stream = nullptr;     // host-buffer entry points: blobs cross PCIe in two pieces (h2d_split)
    bool split_copy = true;                 // KB200_H2D_SPLIT=0 copies in one piece (measurement switch)
    cudaEvent_t ev_copy = nullptr, ev_head = nullptr, ev_tail = nullptr;
    cudaEvent_t ev_col[8] = {};             // column pieces of a hashed upload (h2d_blobs_hashed)
    DBuf sha_state;                         // chaining values between the pieces, [8][n] words
    cudaEvent_t ev_eval = nullptr;          // end of the evaluation that batch verification runs on the auxiliary stream beside k_rlc
    bool eval_pending = false;              // ... and whether the main stream still has to wait for it (run_rlc does)
    bool z_ready = false;                   // the challenges of the next proof / verify pass are already in s->z
    bool fork_recorded = false;             // ev_fork already marks the point of the main stream the point decompression may start from
    uint8_t last_lhs[48], last_rhs[48];
    bool have_last = false;
    // optional CUDA-event timing of the dominant kernel (k_msm), for bench.py's roofline
    bool prof = false;
    std::vector<cudaEvent_t> prof_ev;   // pairs (start, stop)
    size_t prof_used = 0;
    uint64_t prof_items = 0;            // MSMs (4096-point) covered by the recorded launches
};

#define LOCK(s) std::lock_guard<std::recursive_mutex> lock_((s)->mu)
// KB200_TRACE=1: CUDA-event marks on the main stream at the phase boundaries of a call, printed to stderr when the call ends (debugging aid)
struct Trace {
    bool on = getenv("KB200_TRACE") != nullptr;
    std::vector<std::pair<const char*, cudaEvent_t>> ev;
    void mark(cudaStream_t st, const char* what) { if (!on) return; cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, st); ev.emplace_back(what, e); }
    ~Trace() {
        if (!on || ev.empty()) return;
        cudaEventSynchronize(ev.back().second);
        for (size_t i = 1; i < ev.size(); i++) { float ms = 0; cudaEventElapsedTime(&ms, ev[i - 1].second, ev[i].second); fprintf(stderr, "[kb200 trace] %-28s %8.3f ms\n", ev[i].first, ms); }
        for (auto& p : ev) cudaEventDestroy(p.second);
    }
};
static thread_local Trace* g_trace = nullptr;
#define TRACE(s, what) do { if (g_trace) g_trace->mark((s)->stream, what); } while (0)
#define LAUNCH(s, kern, grid, block, ...)                   \
    do {                                                    \
        kern<<<grid, block, 0, (s)->stream>>>(__VA_ARGS__); \
        (s)->launches++;                                    \
    } while (0)

static const size_t RLC_BUCKET_MIN = 1;     // batch verification: shards from this size sum their terms with the bucket method (verify.cuh); measured faster
                                            // than one ladder per term at EVERY size (verify_proof 5.7 -> 4.8 ms, a 256-blob batch 8.4 -> 7.2 ms): the ladders stay for the parity tests
static const size_t CHUNK = 4096;   // blobs per internal pass (bounds the quotient scratch to 512 MiB)

static int init_consts() {
    DevConsts h;
    using namespace kb200::consts;
    memcpy(h.fp_mod, H_FP_P, 48); memcpy(h.fp_one, H_FP_ONE, 48); memcpy(h.fp_r2, H_FP_R2, 48); memcpy(h.fp_b4, H_FP_B4, 48);
    memcpy(h.fp_half, H_FP_HALF, 48); memcpy(h.fp_sqrt_exp, H_FP_SQRT_EXP, 48); memcpy(h.fp_pm2, H_FP_PM2, 48); memcpy(h.fp_beta, H_FP_BETA, 48);
    memcpy(h.fr_one, H_FR_ONE, 32); memcpy(h.fr_r2, H_FR_R2, 32); memcpy(h.fr_rm2, H_FR_RM2, 32); memcpy(h.fr_omega, H_FR_OMEGA, 32);
    memcpy(h.fr_ninv, H_FR_NINV, 32); memcpy(h.fr_mod, H_FR_R, 32);
    memcpy(h.g1_gen_x, H_G1_GEN_X, 48); memcpy(h.g1_gen_y, H_G1_GEN_Y, 48); memcpy(h.g1_neg_gen_y, H_G1_NEG_GEN_Y, 48);
    CU(cudaMemcpyToSymbol(c_k, &h, sizeof h));
    return 0;
}

// ------------------------------------------------------------------ small glue kernels
__global__ void k_brp_permute(const G1Affine* in, G1Affine* out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < KB_N) out[i] = in[brp12(i)];
}
// status[i] = first non-zero of a[i], b[i], c[i] (kateth's per-item evaluation order: blob, commitment, proof)
__global__ void k_merge_status(const int32_t* a, const int32_t* b, const int32_t* c, int n, int32_t* out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int32_t v = a ? a[i] : 0;
    if (v == 0 && b) v = b[i];
    if (v == 0 && c) v = c[i];
    out[i] = v;
}
__global__ void k_store_scalars_be(const Fr* in, int n, const int32_t* status, uint32_t* out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr v = in[i];
    if (status[i] != 0) fr_zero(v);
    fr_store_be(out + (size_t)i * 8, v);
}
// unit-test kernels (K1)
__global__ void k_test_fp_mul(const Fp* a, const Fp* b, Fp* o, int n) { int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) { Fp r; fp_mul(r, a[i], b[i]); o[i] = r; } }
__global__ void k_test_fr_mul(const Fr* a, const Fr* b, Fr* o, int n) { int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) { Fr r; fr_mul(r, a[i], b[i]); o[i] = r; } }
__global__ void k_test_compress(const G1Affine* in, int n, const int32_t* status, uint32_t* out48) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t w[12];
    if (status[i] != 0) { for (int k = 0; k < 12; k++) w[k] = 0; } else g1_compress(w, in[i]);
    for (int k = 0; k < 12; k++) out48[(size_t)i * 12 + k] = w[k];
}
// K0: integer multiplier peak.  (a) four independent carry chains of six IMAD.WIDE.U32.X per thread (the instruction the
// Montgomery products are made of; SASS checked: nothing but IMAD.WIDE in the loop); (b) a chain of Fp products.
__device__ __forceinline__ void k0_chain6(uint64_t (&t)[6], uint32_t a, uint32_t b) {
    asm volatile(
        "{\n\t.reg .u32 l0,h0,l1,h1,l2,h2,l3,h3,l4,h4,l5,h5;\n\t"
        "mov.b64 {l0,h0}, %0;\n\tmov.b64 {l1,h1}, %1;\n\tmov.b64 {l2,h2}, %2;\n\tmov.b64 {l3,h3}, %3;\n\tmov.b64 {l4,h4}, %4;\n\tmov.b64 {l5,h5}, %5;\n\t"
        "mad.lo.cc.u32 l0, %6, %7, l0;\n\tmadc.hi.cc.u32 h0, %6, %7, h0;\n\t"
        "madc.lo.cc.u32 l1, %6, %7, l1;\n\tmadc.hi.cc.u32 h1, %6, %7, h1;\n\t"
        "madc.lo.cc.u32 l2, %6, %7, l2;\n\tmadc.hi.cc.u32 h2, %6, %7, h2;\n\t"
        "madc.lo.cc.u32 l3, %6, %7, l3;\n\tmadc.hi.cc.u32 h3, %6, %7, h3;\n\t"
        "madc.lo.cc.u32 l4, %6, %7, l4;\n\tmadc.hi.cc.u32 h4, %6, %7, h4;\n\t"
        "madc.lo.cc.u32 l5, %6, %7, l5;\n\tmadc.hi.u32 h5, %6, %7, h5;\n\t"
        "mov.b64 %0, {l0,h0};\n\tmov.b64 %1, {l1,h1};\n\tmov.b64 %2, {l2,h2};\n\tmov.b64 %3, {l3,h3};\n\tmov.b64 %4, {l4,h4};\n\tmov.b64 %5, {l5,h5};\n\t}"
        : "+l"(t[0]), "+l"(t[1]), "+l"(t[2]), "+l"(t[3]), "+l"(t[4]), "+l"(t[5]) : "r"(a), "r"(b));
}
__global__ void __launch_bounds__(256) k_mac_wide(uint64_t* out, int iters, uint32_t seed) {
    uint64_t t[4][6];
    uint32_t a = seed + threadIdx.x, b = seed * 2654435761u + blockIdx.x;
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
        for (int k = 0; k < 6; k++) t[c][k] = threadIdx.x + k + c;
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int c = 0; c < 4; c++) k0_chain6(t[c], a + c, b);
    }
    uint64_t s = 0;
#pragma unroll
    for (int c = 0; c < 4; c++)
#pragma unroll
        for (int k = 0; k < 6; k++) s ^= t[c][k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) k_mac_fpmul(Fp* out, int iters, uint32_t seed) {
    Fp a, b;
    for (int k = 0; k < 12; k++) { a.l[k] = seed + threadIdx.x + k; b.l[k] = seed * 31 + blockIdx.x + k; }
    a.l[11] &= 0x0fffffff; b.l[11] &= 0x0fffffff;
#pragma unroll 1
    for (int it = 0; it < iters; it++) { fp_mul(a, a, b); fp_mul(b, b, a); }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a;
}

// ------------------------------------------------------------------ setup loading
static int hexval(int c) { if (c >= '0' && c <= '9') return c - '0'; if (c >= 'a' && c <= 'f') return c - 'a' + 10; if (c >= 'A' && c <= 'F') return c - 'A' + 10; return -1; }
// Pull the array of "0x.."-strings stored under `key`.  Each string is hex-decoded (src/bytes.rs:30-37).
static int json_hex_array(const std::string& doc, const char* key, std::vector<std::vector<uint8_t>>& out) {
    std::string k = std::string("\"") + key + "\"";
    size_t p = doc.find(k);
    if (p == std::string::npos) return KB200_E_PARSE;
    p = doc.find('[', p);
    if (p == std::string::npos) return KB200_E_PARSE;
    for (p++;;) {
        while (p < doc.size() && (doc[p] == ' ' || doc[p] == '\n' || doc[p] == '\r' || doc[p] == '\t' || doc[p] == ',')) p++;
        if (p >= doc.size()) return KB200_E_PARSE;
        if (doc[p] == ']') return 0;
        if (doc[p] != '"') return KB200_E_PARSE;
        size_t e = doc.find('"', p + 1);
        if (e == std::string::npos) return KB200_E_PARSE;
        size_t b = p + 1;
        if (e - b >= 2 && doc[b] == '0' && (doc[b + 1] == 'x' || doc[b + 1] == 'X')) b += 2;
        if ((e - b) & 1) return KB200_E_PARSE;
        std::vector<uint8_t> v((e - b) / 2);
        for (size_t i = 0; i < v.size(); i++) {
            int h = hexval(doc[b + 2 * i]), l = hexval(doc[b + 2 * i + 1]);
            if (h < 0 || l < 0) return KB200_E_PARSE;
            v[i] = (uint8_t)(h * 16 + l);
        }
        out.push_back(std::move(v));
        p = e + 1;
    }
}

static inline int blocks(size_t n, int t) { return (int)((n + t - 1) / t); }
static int first_nonzero(const std::vector<int32_t>& v) { for (int32_t x : v) if (x) return x; return 0; }

static void host_sha256(const uint8_t* data, size_t len, uint8_t out[32]);

// ---- optional on-disk cache of the fixed-base window table ----
// Off unless a directory is named (kb200_set_table_cache or KB200_TABLE_CACHE).  File: 64-byte header + the table bytes as they sit in HBM,
// named after sha256(g1_lagrange bytes | c | glv | layout version); a 64-bit checksum of the table is computed ON THE GPU after the upload
// and compared with the header, so a torn or edited file is rebuilt instead of trusted.
static std::mutex g_cache_mu;
static std::string g_cache_dir;
static bool g_cache_env_read = false;
static std::string cache_dir() {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    if (!g_cache_env_read) { g_cache_env_read = true; if (g_cache_dir.empty()) if (const char* e = getenv("KB200_TABLE_CACHE")) g_cache_dir = e; }
    return g_cache_dir;
}
extern "C" int kb200_set_table_cache(const char* dir) {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    g_cache_env_read = true;
    g_cache_dir = dir ? dir : "";
    return 0;
}
struct CacheHeader { char magic[8]; uint32_t version, c, nwin, dlast, glv, pad; uint64_t rows, bytes, checksum; uint8_t reserved[8]; };
static_assert(sizeof(CacheHeader) == 64, "cache header");
__global__ void k_table_checksum(const uint64_t* __restrict__ data, size_t nwords, unsigned long long* __restrict__ out) {
    unsigned long long acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (size_t)gridDim.x * blockDim.x) acc += data[i] * (2 * (unsigned long long)i + 1);
    for (int o = 16; o >= 1; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}
// scoped device allocation: freed on every exit path
struct DevTmp {
    void* p = nullptr;
    ~DevTmp() { if (p) cudaFree(p); }
    template <class T> T* as() { return reinterpret_cast<T*>(p); }
};

static int table_checksum(kb200_setup* s, const void* d_table, size_t bytes, uint64_t* out) {
    DevTmp acc; CU(cudaMalloc(&acc.p, 8));
    CU(cudaMemsetAsync(acc.p, 0, 8, s->stream));
    LAUNCH(s, k_table_checksum, s->sm_count * 8, 256, (const uint64_t*)d_table, bytes / 8, acc.as<unsigned long long>());
    CU(cudaMemcpyAsync(out, acc.p, 8, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return 0;
}
struct PinnedPair {       // two staging buffers for chunked disk <-> HBM streaming
    void* b[2] = {nullptr, nullptr}; cudaEvent_t ev[2] = {nullptr, nullptr};
    static const size_t CH = (size_t)64 << 20;
    int init() { for (int i = 0; i < 2; i++) { CU(cudaHostAlloc(&b[i], CH, cudaHostAllocDefault)); CU(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming)); } return 0; }
    ~PinnedPair() { for (int i = 0; i < 2; i++) { if (b[i]) cudaFreeHost(b[i]); if (ev[i]) cudaEventDestroy(ev[i]); } }
};
static std::string cache_path(const std::string& dir, const uint8_t* g1, int c, int glv) {
    std::vector<uint8_t> key(KB_N * 48 + 12);
    memcpy(key.data(), g1, KB_N * 48);
    const uint32_t tail[3] = {(uint32_t)c, (uint32_t)glv, 1u /* layout version */};
    memcpy(key.data() + KB_N * 48, tail, 12);
    uint8_t dg[32]; host_sha256(key.data(), key.size(), dg);
    char hex[65]; for (int i = 0; i < 32; i++) snprintf(hex + 2 * i, 3, "%02x", dg[i]);
    return dir + "/kb200_table_" + hex + ".bin";
}
// 0: table loaded from the cache; 1: no usable file (caller builds); < 0: CUDA error
static int cache_load(kb200_setup* s, const std::string& path, const CacheHeader& want, void* d_table) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return 1;
    CacheHeader h;
    bool ok = fread(&h, sizeof h, 1, f) == 1 && memcmp(h.magic, "KB200TBL", 8) == 0 && h.version == want.version && h.c == want.c && h.nwin == want.nwin &&
              h.dlast == want.dlast && h.glv == want.glv && h.rows == want.rows && h.bytes == want.bytes;
    PinnedPair pp;
    if (ok && pp.init() != 0) { fclose(f); return KB200_E_CUDA; }
    for (size_t off = 0, k = 0; ok && off < h.bytes; k++) {
        const size_t m = (h.bytes - off < PinnedPair::CH) ? h.bytes - off : PinnedPair::CH;
        const int i = (int)(k & 1);
        if (k >= 2) cudaEventSynchronize(pp.ev[i]);
        if (fread(pp.b[i], 1, m, f) != m) { ok = false; break; }
        if (cudaMemcpyAsync((char*)d_table + off, pp.b[i], m, cudaMemcpyHostToDevice, s->stream) != cudaSuccess) { fclose(f); return KB200_E_CUDA; }
        cudaEventRecord(pp.ev[i], s->stream);
        off += m;
    }
    fclose(f);
    cudaStreamSynchronize(s->stream);
    if (!ok) return 1;
    uint64_t sum = 0;
    if (int rc = table_checksum(s, d_table, h.bytes, &sum)) return rc;
    return sum == h.checksum ? 0 : 1;
}
static void cache_store(kb200_setup* s, const std::string& path, CacheHeader h, const void* d_table) {
    if (table_checksum(s, d_table, h.bytes, &h.checksum) != 0) return;
    const std::string tmp = path + ".tmp";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return;
    PinnedPair pp;
    bool ok = pp.init() == 0 && fwrite(&h, sizeof h, 1, f) == 1;
    for (size_t off = 0; ok && off < h.bytes;) {
        const size_t m = (h.bytes - off < PinnedPair::CH) ? h.bytes - off : PinnedPair::CH;
        ok = cudaMemcpy(pp.b[0], (const char*)d_table + off, m, cudaMemcpyDeviceToHost) == cudaSuccess && fwrite(pp.b[0], 1, m, f) == m;
        off += m;
    }
    ok = (fclose(f) == 0) && ok;
    if (ok) rename(tmp.c_str(), path.c_str()); else remove(tmp.c_str());
}

// ------------------------------------------------------------------ scratch of the allocation-free commit entry point
// kb200_blob_to_commitment_device_noalloc (below) keeps the per-(blob, slice) partial sums of k_msm in this file-scope array instead of
// the handle's growable workspace: KB200_NOALLOC_MAX_BLOBS blobs x 32 slices x one XYZZ point (192 B) = 384 KiB, held by every GPU that
// loads the library until the process exits.  In HOST code the name below is only a stand-in for the symbol: the address a kernel can
// use comes from cudaGetSymbolAddress, asked once per device with that device current.
#define NOALLOC_SLICES 32
static __device__ __align__(256) uint4 g_noalloc_scratch[(size_t)KB200_NOALLOC_MAX_BLOBS * NOALLOC_SLICES * (sizeof(G1) / sizeof(uint4))];
static_assert(sizeof(G1) % sizeof(uint4) == 0, "XYZZ point is a whole number of 16-byte words");
static std::mutex g_noalloc_mu;
static void* g_noalloc_ptr[64] = {};          // per device ordinal
static int noalloc_setup(kb200_setup* s) {    // called from build_setup with s->device current
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (s->device < 0 || s->device >= 64) return KB200_E_ARG;
    if (!g_noalloc_ptr[s->device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) { snprintf(g_err, sizeof g_err, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", s->device, cudaGetErrorString(e)); return KB200_E_CUDA; }
        g_noalloc_ptr[s->device] = p;
    }
    s->d_noalloc = (G1*)g_noalloc_ptr[s->device];
    return 0;
}

static int build_setup(kb200_setup* s, const uint8_t* g1, const uint8_t* g2, int window_bits) {
    auto t0 = std::chrono::steady_clock::now();
    CU(cudaSetDevice(s->device));
    cudaDeviceProp prop; CU(cudaGetDeviceProperties(&prop, s->device));
    s->sm_count = prop.multiProcessorCount;
    if (int rc = init_consts()) return rc;
    // G1: decompress + subgroup-check the 4096 Lagrange points on the GPU (src/kzg/setup.rs:59-65)
    DevTmp t_in, t_pts, t_st, t_scratch;
    CU(cudaMalloc(&t_in.p, KB_N * 48)); CU(cudaMalloc(&t_pts.p, KB_N * sizeof(G1Affine))); CU(cudaMalloc(&t_st.p, KB_N * 4));
    uint32_t* d_in = t_in.as<uint32_t>(); G1Affine* d_pts = t_pts.as<G1Affine>(); int32_t* d_st = t_st.as<int32_t>();
    CU(cudaMemcpy(d_in, g1, KB_N * 48, cudaMemcpyHostToDevice));
    CU(cudaMemset(d_st, 0, KB_N * 4));
    LAUNCH(s, k_g1_decompress, KB_N / 64, 64, d_in, (const uint32_t*)nullptr, KB_N, d_pts, (G1Affine*)nullptr, d_st, (int32_t*)nullptr, 0);
    std::vector<int32_t> st(KB_N);
    CU(cudaMemcpy(st.data(), d_st, KB_N * 4, cudaMemcpyDeviceToHost));
    if (int e = first_nonzero(st)) return e;
    CU(cudaMalloc(&s->d_bases, KB_N * sizeof(G1Affine)));
    LAUNCH(s, k_brp_permute, KB_N / 256, 256, d_pts, s->d_bases);
    // G2: all 65 points are validated on the host (src/kzg/setup.rs:68-72); only [1] is ever used (:91,:158)
    host::G2A tau, gen;
    for (int i = 0; i < 65; i++) {
        host::G2A q; int e = host::g2_decompress(q, g2 + 96 * i);
        if (e) return e;
        if (i == 1) tau = q;
    }
    host::g2_generator(gen);
    host::g2_prepare(s->tau_prep, tau); host::g2_prepare(s->g2_prep, gen);
    // roots of unity, bit-reversed (src/kzg/setup.rs:74-75)
    CU(cudaMalloc(&s->d_roots, KB_N * sizeof(Fr)));
    LAUNCH(s, k_roots, KB_N / 256, 256, s->d_roots);
    CU(cudaMalloc(&s->d_neg_g_pow, 32 * 256 * sizeof(G1Affine)));
    LAUNCH(s, k_neg_g_table, 32, 256, s->d_neg_g_pow);
    {   // device copy of what the per-item pairing check needs (pairing.cuh): the host module's prepared lines, bit for bit
        static_assert(sizeof(PairLine) == sizeof(host::Line) && sizeof(Fp2) == sizeof(host::Fp2) && sizeof(Fp) == sizeof(host::Fp), "host and device field layouts");
        if (s->tau_prep.inf || s->g2_prep.inf || s->tau_prep.n != s->g2_prep.n || s->tau_prep.n > 68) { snprintf(g_err, sizeof g_err, "setup: degenerate G2 points"); return KB200_E_ARG; }
        std::vector<uint8_t> hb(sizeof(PairConsts), 0);
        PairConsts* hp = reinterpret_cast<PairConsts*>(hb.data());
        memcpy(hp->tau, s->tau_prep.lines, sizeof(host::Line) * s->tau_prep.n);
        memcpy(hp->g2, s->g2_prep.lines, sizeof(host::Line) * s->g2_prep.n);
        host::Fp2 f1[6]; host::Fp f2[6]; host::frobenius_consts(f1, f2);
        memcpy(hp->frob1, f1, sizeof f1); memcpy(hp->frob2, f2, sizeof f2);
        hp->nlines = s->tau_prep.n;
        CU(cudaMalloc(&s->d_pair, sizeof(PairConsts)));
        CU(cudaMemcpy(s->d_pair, hp, sizeof(PairConsts), cudaMemcpyHostToDevice));
        CU(cudaMalloc(&s->d_neg_g_bits, 255 * sizeof(G1Affine)));
        LAUNCH(s, k_neg_g_bits, 4, 64, s->d_neg_g_bits);
    }
    // fixed-base window table: nwin = ceil(255 / c) windows; windows below the top hold the 2^(c-1) signed-digit magnitudes, the top
    // window is unsigned and holds (r >> c*(nwin-1)) + 1 magnitudes (scalars are canonical, so its digit never exceeds that)
    // GLV form (window_bits 16, or 100 + c for any c): the windows cover one 128-bit half of the scalar split k = k1 + k2*N
    // (glv.cuh), nwin = ceil(128 / c), and the top window holds (N >> c*(nwin-1)) + 1 magnitudes; both halves read the same table.
    if (window_bits == 0) {
        // default: the widest table the free HBM takes with room for the workspace of 4096-blob batches -- on an idle B200 the
        // throughput configuration (16 = GLV, 8 shared windows, 100 GiB); a second handle on the same GPU, or a smaller part, steps down
        size_t free_b = 0, total_b = 0;
        CU(cudaMemGetInfo(&free_b, &total_b));
        const double gib = (double)free_b / (double)((size_t)1 << 30);
        window_bits = gib >= 124.0 ? 16 : (gib >= 45.0 ? 13 : (gib >= 10.0 ? 9 : 6));
    }
    int c = window_bits;
    int glv = 0;
    if (c == 16) glv = 1;
    else if (c >= 100) { glv = 1; c -= 100; }
    if (c < 4 || c > (glv ? 16 : 15)) return KB200_E_ARG;
    int nwin = ((glv ? 128 : 255) + c - 1) / c;
    int dlast;
    {
        static const uint32_t GLV_N[4] = {0x00000000u, 0x00000001u, 0x0001a402u, 0xac45a401u};
        const uint32_t* top_of = glv ? GLV_N : kb200::consts::H_FR_R;
        const int limbs = glv ? 4 : 8;
        int sh = c * (nwin - 1), limb = sh >> 5, off = sh & 31;
        uint64_t top = top_of[limb];
        if (limb + 1 < limbs) top |= (uint64_t)top_of[limb + 1] << 32;
        dlast = (int)(top >> off) + 1;
    }
    size_t win_rows = (size_t)1 << (c - 1);
    size_t rows = (size_t)(nwin - 1) * win_rows + dlast;
    size_t bytes = rows * KB_N * sizeof(G1Affine);
    // build scratch (Z and prefix products) is bounded to ~16 GiB: windows are built in groups, the top window alone
    size_t row_bytes = (size_t)KB_N * sizeof(G1Affine);
    int group = (int)(((size_t)16 << 30) / (win_rows * row_bytes));
    if (group < 1) group = 1;
    if (group > nwin - 1) group = nwin - 1;
    size_t scratch_rows = (size_t)group * win_rows;
    if ((size_t)dlast > scratch_rows) scratch_rows = dlast;
    CU(cudaMalloc(&s->d_table, bytes));
    const std::string cdir = cache_dir();
    CacheHeader hdr; memset(&hdr, 0, sizeof hdr);
    memcpy(hdr.magic, "KB200TBL", 8); hdr.version = 1; hdr.c = (uint32_t)c; hdr.nwin = (uint32_t)nwin; hdr.dlast = (uint32_t)dlast; hdr.glv = (uint32_t)glv;
    hdr.rows = rows; hdr.bytes = bytes;
    const std::string cpath = cdir.empty() ? std::string() : cache_path(cdir, g1, c, glv);
    int cached = 1;
    if (!cpath.empty()) { cached = cache_load(s, cpath, hdr, s->d_table); if (cached < 0) return cached; }
    s->table_from_cache = cached == 0;
    if (cached != 0) {
        CU(cudaMalloc(&t_scratch.p, scratch_rows * row_bytes));
        uint4* scratch = t_scratch.as<uint4>();
        for (int j0 = 0; j0 < nwin;) {
            int cnt = (j0 == nwin - 1) ? 1 : ((nwin - 1 - j0 < group) ? nwin - 1 - j0 : group);
            LAUNCH(s, k_table_build, (KB_N * cnt + 31) / 32, 32, s->d_table, scratch, s->d_bases, c, nwin, dlast, j0, cnt);
            j0 += cnt;
        }
        CU(cudaStreamSynchronize(s->stream));
        CU(cudaGetLastError());
        if (!cpath.empty()) cache_store(s, cpath, hdr, s->d_table);
    }
    CU(cudaStreamSynchronize(s->stream));
    CU(cudaGetLastError());
    s->tab.data = s->d_table; s->tab.c = c; s->tab.nwin = nwin; s->tab.dlast = dlast; s->tab.glv = glv;
    {   // batch-affine MSM (its own translation unit, msm_ba.cu): persistent grid, work counter
        BaPlan plan;
        if (ba_init(&plan) != 0) { snprintf(g_err, sizeof g_err, "ba_init: %s", cudaGetErrorString(cudaGetLastError())); return KB200_E_CUDA; }
        s->ba_grid = plan.ctas_per_sm * s->sm_count;
        CU(cudaMalloc(&s->d_counter, sizeof(unsigned int)));
        if (int rc = noalloc_setup(s)) return rc;
        CU(cudaStreamCreateWithFlags(&s->aux_stream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s->ev_join, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s->ev_eval, cudaEventDisableTiming));
        CU(cudaStreamCreateWithFlags(&s->copy_stream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&s->ev_copy, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s->ev_head, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&s->ev_tail, cudaEventDisableTiming));
        for (int j = 0; j < 8; j++) CU(cudaEventCreateWithFlags(&s->ev_col[j], cudaEventDisableTiming));
        if (const char* e = getenv("KB200_MSM_MODE")) { int m = atoi(e); if (m >= 0 && m <= 2) s->msm_mode = m; }
        if (const char* e = getenv("KB200_RLC_MODE")) { int m = atoi(e); if (m >= 0 && m <= 2) s->rlc_mode = m; }
        if (const char* e = getenv("KB200_H2D_SPLIT")) s->split_copy = atoi(e) != 0;
    }
    s->table_bytes = bytes;
    s->setup_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return 0;
}

extern "C" int kb200_setup_load_raw(const uint8_t* g1, size_t n_g1, const uint8_t* g2, size_t n_g2, int device, int window_bits, kb200_setup** out) {
    if (!g1 || !g2 || !out) return KB200_E_ARG;
    if (n_g1 != KB_N) return KB200_E_SETUP_G1_LEN;
    if (n_g2 != 65) return KB200_E_SETUP_G2_LEN;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { snprintf(g_err, sizeof g_err, "no CUDA device: this engine has no CPU fallback"); return KB200_E_CUDA; }
    kb200_setup* s = new kb200_setup();
    s->device = device;
    int rc = build_setup(s, g1, g2, window_bits);
    if (rc) { kb200_setup_free(s); return rc; }
    *out = s; return 0;
}
// Cross-check of a loaded setup against the monomial-basis points the same ceremony publishes (consensus-spec presets, c-kzg's
// trusted_setup.txt): X^k evaluated over the domain, in blob order, must commit to [tau^k]G1 = g1_monomial[k].  All n <= 4096 powers are
// checked with the engine's own batch MSM (one 4096-blob commit, ~80 ms): a Lagrange array that was permuted, truncated or mixed with
// another ceremony's fails here instead of producing commitments nobody else can verify.
extern "C" int kb200_setup_check_g1_monomial(kb200_setup* s, const uint8_t* g1_monomial48, size_t n) {
    if (!s || !g1_monomial48 || n == 0 || n > KB_N) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (int rc = s->blobs.ensure(n * (size_t)KB_BLOB_BYTES)) return rc;
    if (int rc = s->out48.ensure(n * 48)) return rc;
    if (int rc = s->status.ensure(n * 4)) return rc;
    LAUNCH(s, k_vandermonde, blocks(n * KB_N, 256), 256, s->d_roots, 0, (int)n, s->blobs.as<uint32_t>());
    int rc = kb200_blob_to_commitment_device(s, s->blobs.p, n, s->out48.p, s->status.as<int32_t>());
    if (rc) return rc < 0 ? rc : KB200_E_SETUP_G1_MONOMIAL;
    std::vector<uint8_t> got(n * 48);
    CU(cudaMemcpy(got.data(), s->out48.p, n * 48, cudaMemcpyDeviceToHost));
    for (size_t k = 0; k < n; k++)
        if (memcmp(&got[48 * k], g1_monomial48 + 48 * k, 48) != 0) {
            snprintf(g_err, sizeof g_err, "g1_monomial[%zu] is not the commitment of X^%zu under g1_lagrange: the two bases belong to different setups", k, k);
            return KB200_E_SETUP_G1_MONOMIAL;
        }
    return 0;
}
static int check_monomial_if_present(kb200_setup** out, const std::vector<std::vector<uint8_t>>& g1m) {
    if (g1m.empty()) return 0;
    int rc = 0;
    if (g1m.size() != KB_N) { snprintf(g_err, sizeof g_err, "g1_monomial has %zu points, expected %d", g1m.size(), KB_N); rc = KB200_E_SETUP_G1_MONOMIAL; }
    std::vector<uint8_t> b(KB_N * 48);
    for (size_t i = 0; !rc && i < g1m.size(); i++) { if (g1m[i].size() != 48) rc = KB200_E_G1_INVALID_ENCODING; else memcpy(&b[48 * i], g1m[i].data(), 48); }
    if (!rc) rc = kb200_setup_check_g1_monomial(*out, b.data(), KB_N);
    if (rc) { kb200_setup_free(*out); *out = nullptr; }
    return rc;
}
extern "C" int kb200_setup_load_json(const char* path, int device, int window_bits, kb200_setup** out) {
    if (!path || !out) return KB200_E_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) { snprintf(g_err, sizeof g_err, "cannot open %s", path); return KB200_E_IO; }
    std::string doc; char buf[65536]; size_t k;
    while ((k = fread(buf, 1, sizeof buf, f)) > 0) doc.append(buf, k);
    fclose(f);
    std::vector<std::vector<uint8_t>> g1, g2;
    if (int rc = json_hex_array(doc, "g1_lagrange", g1)) return rc;
    if (int rc = json_hex_array(doc, "g2_monomial", g2)) return rc;
    if (g1.size() != KB_N) return KB200_E_SETUP_G1_LEN;
    if (g2.size() != 65) return KB200_E_SETUP_G2_LEN;
    std::vector<uint8_t> b1(KB_N * 48), b2(65 * 96);
    // kateth reads 48/96 bytes behind the pointer without a length check (src/bls.rs:514); the ABI checks.
    for (size_t i = 0; i < g1.size(); i++) { if (g1[i].size() != 48) return KB200_E_G1_INVALID_ENCODING; memcpy(&b1[48 * i], g1[i].data(), 48); }
    for (size_t i = 0; i < g2.size(); i++) { if (g2[i].size() != 96) return KB200_E_G1_INVALID_ENCODING; memcpy(&b2[96 * i], g2[i].data(), 96); }
    // files that also carry the monomial basis (consensus-spec presets) are cross-checked against it; kateth's own file has no such key
    std::vector<std::vector<uint8_t>> g1m;
    if (doc.find("\"g1_monomial\"") != std::string::npos) if (int rc = json_hex_array(doc, "g1_monomial", g1m)) return rc;
    if (int rc = kb200_setup_load_raw(b1.data(), KB_N, b2.data(), 65, device, window_bits, out)) return rc;
    return check_monomial_if_present(out, g1m);
}
// c-kzg-4844's trusted_setup.txt: "4096\n65\n", then 4096 lines of 96 hex digits (g1_lagrange), 65 lines of 192 hex digits
// (g2_monomial) and, in newer files, 4096 more G1 lines (g1_monomial, ignored here).  Whitespace-separated tokens, no "0x".
extern "C" int kb200_setup_load_txt(const char* path, int device, int window_bits, kb200_setup** out) {
    if (!path || !out) return KB200_E_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) { snprintf(g_err, sizeof g_err, "cannot open %s", path); return KB200_E_IO; }
    std::string doc; char buf[65536]; size_t k;
    while ((k = fread(buf, 1, sizeof buf, f)) > 0) doc.append(buf, k);
    fclose(f);
    std::vector<std::string> tok;
    for (size_t i = 0; i < doc.size();) {
        while (i < doc.size() && isspace((unsigned char)doc[i])) i++;
        size_t j = i;
        while (j < doc.size() && !isspace((unsigned char)doc[j])) j++;
        if (j > i) tok.emplace_back(doc, i, j - i);
        i = j;
    }
    if (tok.size() < 2) { snprintf(g_err, sizeof g_err, "trusted setup text: missing counts"); return KB200_E_PARSE; }
    char* end = nullptr;
    unsigned long n1 = strtoul(tok[0].c_str(), &end, 10); if (*end) return KB200_E_PARSE;
    unsigned long n2 = strtoul(tok[1].c_str(), &end, 10); if (*end) return KB200_E_PARSE;
    if (n1 != KB_N) return KB200_E_SETUP_G1_LEN;
    if (n2 != 65) return KB200_E_SETUP_G2_LEN;
    if (tok.size() < 2 + n1 + n2) { snprintf(g_err, sizeof g_err, "trusted setup text: %zu tokens, need %lu", tok.size(), 2 + n1 + n2); return KB200_E_PARSE; }
    std::vector<uint8_t> b1(KB_N * 48), b2(65 * 96);
    auto unhex = [&](const std::string& t, uint8_t* dst, size_t bytes) -> int {
        const char* p = t.c_str(); size_t len = t.size();
        if (len >= 2 && p[0] == '0' && (p[1] == 'x' || p[1] == 'X')) { p += 2; len -= 2; }
        if (len != 2 * bytes) return KB200_E_G1_INVALID_ENCODING;
        for (size_t i = 0; i < bytes; i++) {
            int hi = hexval(p[2 * i]), lo = hexval(p[2 * i + 1]);
            if (hi < 0 || lo < 0) return KB200_E_PARSE;
            dst[i] = (uint8_t)(hi << 4 | lo);
        }
        return 0;
    };
    for (size_t i = 0; i < n1; i++) if (int rc = unhex(tok[2 + i], &b1[48 * i], 48)) return rc;
    for (size_t i = 0; i < n2; i++) if (int rc = unhex(tok[2 + n1 + i], &b2[96 * i], 96)) return rc;
    // newer files append the 4096 monomial-basis G1 points: cross-check them (a partial tail is ignored, as before)
    std::vector<std::vector<uint8_t>> g1m;
    if (tok.size() >= 2 + n1 + n2 + KB_N)
        for (size_t i = 0; i < (size_t)KB_N; i++) { std::vector<uint8_t> v(48); if (int rc = unhex(tok[2 + n1 + n2 + i], v.data(), 48)) return rc; g1m.push_back(std::move(v)); }
    if (int rc = kb200_setup_load_raw(b1.data(), KB_N, b2.data(), 65, device, window_bits, out)) return rc;
    return check_monomial_if_present(out, g1m);
}
extern "C" void kb200_setup_free(kb200_setup* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->d_bases) cudaFree(s->d_bases);
    if (s->d_roots) cudaFree(s->d_roots);
    if (s->d_neg_g_pow) cudaFree(s->d_neg_g_pow);
    if (s->d_neg_g_bits) cudaFree(s->d_neg_g_bits);
    if (s->d_pair) cudaFree(s->d_pair);
    if (s->d_table) cudaFree(s->d_table);
    if (s->d_counter) cudaFree(s->d_counter);
    if (s->aux_stream) cudaStreamDestroy(s->aux_stream);
    if (s->ev_fork) cudaEventDestroy(s->ev_fork);
    if (s->ev_join) cudaEventDestroy(s->ev_join);
    if (s->ev_eval) cudaEventDestroy(s->ev_eval);
    if (s->copy_stream) cudaStreamDestroy(s->copy_stream);
    if (s->ev_copy) cudaEventDestroy(s->ev_copy);
    if (s->ev_head) cudaEventDestroy(s->ev_head);
    if (s->ev_tail) cudaEventDestroy(s->ev_tail);
    for (int j = 0; j < 8; j++) if (s->ev_col[j]) cudaEventDestroy(s->ev_col[j]);
    s->sha_state.release();
    DBuf* bufs[] = {&s->blobs, &s->c48, &s->p48, &s->zy32, &s->out48, &s->out32, &s->st_blob, &s->st_a, &s->st_b, &s->status, &s->z, &s->y, &s->prod,
                    &s->indom, &s->q, &s->partials, &s->aff_c, &s->aff_p, &s->rlc_part, &s->rlc_s, &s->rmont, &s->records, &s->combine, &s->dig, &s->scr_pts, &s->scr_pre, &s->transcript, &s->st_tmp, &s->rlc_dig, &s->rlc_wsum, &s->prodh};
    for (DBuf* b : bufs) b->release();
    for (cudaEvent_t e : s->prof_ev) cudaEventDestroy(e);
    delete s;
}
extern "C" int kb200_alloc_pinned(size_t bytes, void** out) {
    if (!out) return KB200_E_ARG;
    CU(cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault));
    return 0;
}
extern "C" int kb200_free_pinned(void* p) {
    if (p) CU(cudaFreeHost(p));
    return 0;
}
extern "C" int kb200_set_stream(kb200_setup* s, void* stream) { if (!s) return KB200_E_ARG; LOCK(s); s->stream = (cudaStream_t)stream; return 0; }
extern "C" int kb200_set_batch_challenge(kb200_setup* s, int mode) {
    if (!s || mode < 0 || mode > 1) return KB200_E_ARG;
    LOCK(s);
    s->challenge_mode = mode;
    return 0;
}
extern "C" int kb200_set_msm_mode(kb200_setup* s, int mode) {
    if (!s || mode < 0 || mode > 2) return KB200_E_ARG;
    LOCK(s);
    s->msm_mode = mode;
    return 0;
}
extern "C" int kb200_set_rlc_mode(kb200_setup* s, int mode) {
    if (!s || mode < 0 || mode > 2) return KB200_E_ARG;
    LOCK(s);
    s->rlc_mode = mode;
    return 0;
}
extern "C" int kb200_set_profiling(kb200_setup* s, int on) {
    if (!s) return KB200_E_ARG;
    LOCK(s);
    s->prof = on != 0; s->prof_used = 0; s->prof_items = 0;
    return 0;
}
// total device time (ms) of the k_msm launches since the last call / set_profiling, their count and the MSMs they covered
extern "C" int kb200_get_msm_profile(kb200_setup* s, double* ms_total, uint64_t* launches, uint64_t* msms) {
    if (!s || !ms_total || !launches || !msms) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    CU(cudaStreamSynchronize(s->stream));
    double tot = 0;
    for (size_t i = 0; i + 1 < s->prof_used; i += 2) { float ms = 0; CU(cudaEventElapsedTime(&ms, s->prof_ev[i], s->prof_ev[i + 1])); tot += ms; }
    *ms_total = tot; *launches = s->prof_used / 2; *msms = s->prof_items;
    s->prof_used = 0; s->prof_items = 0;
    return 0;
}
extern "C" int kb200_get_info(kb200_setup* s, kb200_info* info) {
    if (!s || !info) return KB200_E_ARG;
    LOCK(s);
    info->window_bits = s->tab.c; info->windows = s->tab.nwin; info->table_bytes = s->table_bytes; info->kernel_launches = s->launches;
    info->device = s->device; info->sm_count = s->sm_count; info->setup_seconds = s->setup_seconds;
    info->last_msm_batch_affine = s->last_ba; info->ba_grid = s->ba_grid; info->glv = s->tab.glv; info->table_from_cache = s->table_from_cache ? 1 : 0;
    return 0;
}

// ------------------------------------------------------------------ pipelines (all pointers are device pointers)
static inline int pick_slices(kb200_setup* s, size_t n) {
    // aim for >= 4 CTAs per SM; one CTA per blob once the batch alone fills the GPU
    size_t target = (size_t)s->sm_count * 4;
    int S = 1;
    while (S < 32 && n * S < target) S <<= 1;
    return S;
}

// MSM of n scalar vectors (FMT 0: blobs, FMT 1: LE limbs) -> n compressed points
template <int FMT>
static int run_msm(kb200_setup* s, const uint32_t* d_scalars, size_t n, const int32_t* d_status, uint32_t* d_out48) {
    int S = pick_slices(s, n);
    const bool ba = s->msm_mode == 2 || (s->msm_mode == 0 && S == 1);
    if (ba) S = 1;
    s->last_ba = ba ? 1 : 0;
    if (int rc = s->partials.ensure(n * S * sizeof(G1))) return rc;
    unsigned grid = 0;
    if (ba) {
        grid = (unsigned)((n < (size_t)s->ba_grid) ? n : (size_t)s->ba_grid);
        size_t dig_b, pts_b, pre_b;
        ba_workspace(s->tab, s->ba_grid, n, &dig_b, &pts_b, &pre_b);
        if (int rc = s->dig.ensure(dig_b)) return rc;
        if (int rc = s->scr_pts.ensure(pts_b)) return rc;
        if (int rc = s->scr_pre.ensure(pre_b)) return rc;
    }
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (s->prof) {
        if (s->prof_used + 2 > s->prof_ev.size()) { cudaEvent_t a, b; CU(cudaEventCreate(&a)); CU(cudaEventCreate(&b)); s->prof_ev.push_back(a); s->prof_ev.push_back(b); }
        e0 = s->prof_ev[s->prof_used]; e1 = s->prof_ev[s->prof_used + 1]; s->prof_used += 2; s->prof_items += n;
        CU(cudaEventRecord(e0, s->stream));
    }
    if (ba) {
        CU(cudaMemsetAsync(s->d_counter, 0, sizeof(unsigned int), s->stream));
        ba_launch(s->tab, FMT, d_scalars, d_status, s->dig.as<uint32_t>(), s->scr_pts.p, s->scr_pre.p, s->partials.as<G1>(), s->d_counter, (int)n, grid, s->stream);
        s->launches += 2;      // k_msm_recode + k_msm_ba
    } else {
        LAUNCH(s, (k_msm<FMT>), (unsigned)(n * S), MSM_THREADS, s->tab, d_scalars, d_status, s->partials.as<G1>(), S);
    }
    if (e1) CU(cudaEventRecord(e1, s->stream));
    LAUNCH(s, k_msm_finish, blocks(n, FIN_WARPS), 32 * FIN_WARPS, s->partials.as<G1>(), S, (int)n, d_status, d_out48, (G1Affine*)nullptr);
    return 0;
}
// evaluation (and optionally quotient) of n blobs at the canonical points d_z; y -> d_y (canonical)
template <bool WANT_Q>
static int run_eval(kb200_setup* s, const uint32_t* d_blobs, size_t n, const int32_t* d_status, const Fr* d_z, Fr* d_y) {
    if (int rc = s->prod.ensure(n * sizeof(Fr))) return rc;
    if (int rc = s->indom.ensure(n * 4)) return rc;
    if (WANT_Q) if (int rc = s->q.ensure(n * (size_t)KB_BLOB_BYTES)) return rc;
    // bit 0: evaluation, bit 1: evaluation + quotient, in two halves (k_eval_finish2).  Default 1: the quotient form measured no faster
    // than k_eval_finish<true> at 4096 blobs (prove stage 76.0 ms either way); it is kept, tested, behind the switch.
    static const int halves_env = getenv("KB200_EVAL2") ? atoi(getenv("KB200_EVAL2")) : 1;
    if ((halves_env >> (WANT_Q ? 1 : 0)) & 1) {
        if (int rc = s->prodh.ensure(2 * n * sizeof(Fr))) return rc;
        LAUNCH(s, k_eval_prod, (unsigned)n, EV_THREADS, d_z, s->d_roots, (int)n, s->prod.as<Fr>(), s->indom.as<int32_t>(), s->prodh.as<Fr>());
        LAUNCH(s, k_eval_inv, blocks(n, 32), 32, s->prod.as<Fr>(), (int)n);
        LAUNCH(s, (k_eval_finish2<WANT_Q>), (unsigned)n, EV2_THREADS, d_blobs, d_z, s->d_roots, s->prod.as<Fr>(), s->prodh.as<Fr>(), s->indom.as<int32_t>(), d_status, (int)n, d_y,
               s->q.as<uint32_t>());
        return 0;
    }
    LAUNCH(s, k_eval_prod, (unsigned)n, EV_THREADS, d_z, s->d_roots, (int)n, s->prod.as<Fr>(), s->indom.as<int32_t>(), (Fr*)nullptr);
    LAUNCH(s, k_eval_inv, blocks(n, 32), 32, s->prod.as<Fr>(), (int)n);
    LAUNCH(s, (k_eval_finish<WANT_Q>), (unsigned)n, EV_THREADS, d_blobs, d_z, s->d_roots, s->prod.as<Fr>(), s->indom.as<int32_t>(), d_status,
           (int)n, d_y, s->q.as<uint32_t>());
    return 0;
}
static int fetch_status(kb200_setup* s, const int32_t* d_status, size_t n, std::vector<int32_t>& h) {
    h.resize(n);
    if (n == 0) return 0;
    CU(cudaMemcpyAsync(h.data(), d_status, n * 4, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    CU(cudaGetLastError());
    return 0;
}

extern "C" int kb200_blob_to_commitment_device(kb200_setup* s, const void* d_blobs, size_t n, void* d_out48, int32_t* d_status) {
    if (!s || (n && (!d_blobs || !d_out48 || !d_status))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (n == 0) return 0;
    CU(cudaMemsetAsync(d_status, 0, n * 4, s->stream));
    LAUNCH(s, k_blob_check, blocks(n * KB_N, 256), 256, (const uint32_t*)d_blobs, (int)n, d_status);
    if (int rc = run_msm<0>(s, (const uint32_t*)d_blobs, n, d_status, (uint32_t*)d_out48)) return rc;
    std::vector<int32_t> h; if (int rc = fetch_status(s, d_status, n, h)) return rc;
    return first_nonzero(h);
}

// The same commitments with NO allocation of any kind (no counterpart in kateth; added for callers
// that cannot call cudaMalloc: graph capture, tight serving loops).  At most KB200_NOALLOC_MAX_BLOBS blobs per call; always the sliced XYZZ
// shape k_msm (the batch-affine shape needs gigabytes of workspace), its partial sums in g_noalloc_scratch.  Enqueue-only: blob check, MSM
// and compression go onto the handle's stream and the call returns; d_status[i] (0 or the blob's error class) and d_out48 are valid once
// that stream has passed them.  Every call on a GPU shares the one array: calls that overlap on the device (two handles or two streams
// on one GPU) race.  k_msm writes every partial slot it later reads, so nothing in the array has to start at zero; the slots of the call
// are cleared anyway so that a skipped launch could not pass on the previous call's sums.
extern "C" int kb200_blob_to_commitment_device_noalloc(kb200_setup* s, const void* d_blobs, size_t n, void* d_out48, int32_t* d_status) {
    if (!s || (n && (!d_blobs || !d_out48 || !d_status))) return KB200_E_ARG;
    if (n > (size_t)KB200_NOALLOC_MAX_BLOBS) {
        snprintf(g_err, sizeof g_err, "kb200_blob_to_commitment_device_noalloc: %zu blobs exceed the static scratch (KB200_NOALLOC_MAX_BLOBS = %d)", n, KB200_NOALLOC_MAX_BLOBS);
        return KB200_E_ARG;
    }
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (n == 0) return 0;
    if (!s->d_noalloc) { snprintf(g_err, sizeof g_err, "static scratch address missing (setup did not complete)"); return KB200_E_CUDA; }
    int S = pick_slices(s, n);
    if (S > NOALLOC_SLICES) S = NOALLOC_SLICES;
    CU(cudaMemsetAsync(d_status, 0, n * 4, s->stream));
    CU(cudaMemsetAsync(s->d_noalloc, 0, n * S * sizeof(G1), s->stream));
    LAUNCH(s, k_blob_check, blocks(n * KB_N, 256), 256, (const uint32_t*)d_blobs, (int)n, d_status);
    LAUNCH(s, (k_msm<0>), (unsigned)(n * S), MSM_THREADS, s->tab, (const uint32_t*)d_blobs, (const int32_t*)d_status, s->d_noalloc, S);
    LAUNCH(s, k_msm_finish, blocks(n, FIN_WARPS), 32 * FIN_WARPS, (const G1*)s->d_noalloc, S, (int)n, (const int32_t*)d_status, (uint32_t*)d_out48, (G1Affine*)nullptr);
    CU(cudaGetLastError());
    return 0;
}

// shared tail of blob_proof / proof: z in s->z (n items), status final -> quotient -> MSM -> compressed proofs; y is left in s->out32
static int prove_tail(kb200_setup* s, const uint32_t* d_blobs, size_t n, int32_t* d_status, uint32_t* d_out48) {
    if (int rc = s->out32.ensure(n * sizeof(Fr))) return rc;       // before the loop: growing it would drop the earlier chunks' y
    for (size_t off = 0; off < n; off += CHUNK) {
        size_t m = (n - off < CHUNK) ? n - off : CHUNK;
        if (int rc = run_eval<true>(s, d_blobs + off * (KB_N * 8), m, d_status + off, s->z.as<Fr>() + off, s->out32.as<Fr>() + off)) return rc;
        if (int rc = run_msm<1>(s, s->q.as<uint32_t>(), m, d_status + off, d_out48 + off * 12)) return rc;
    }
    return 0;
}

extern "C" int kb200_compute_blob_proof_device(kb200_setup* s, const void* d_blobs, const void* d_comm48, size_t n, void* d_out48, int32_t* d_status) {
    if (!s || (n && (!d_blobs || !d_comm48 || !d_out48 || !d_status))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (n == 0) return 0;
    if (int rc = s->st_blob.ensure(n * 4)) return rc;
    if (int rc = s->st_a.ensure(n * 4)) return rc;
    if (int rc = s->aff_c.ensure(n * sizeof(G1Affine))) return rc;
    if (int rc = s->z.ensure(n * sizeof(Fr))) return rc;
    CU(cudaMemsetAsync(s->st_blob.p, 0, n * 4, s->stream));
    // as in verify_prepare: the commitment check (a 380-step chain per point) runs on the auxiliary stream beside the blob check
    // and the serial SHA-256 of each blob; the challenge only needs the blob's own status
    if (!s->fork_recorded) CU(cudaEventRecord(s->ev_fork, s->stream));
    s->fork_recorded = false;
    CU(cudaStreamWaitEvent(s->aux_stream, s->ev_fork, 0));
    CU(cudaMemsetAsync(s->st_a.p, 0, n * 4, s->aux_stream));
    k_g1_decompress<<<blocks(n, 64), 64, 0, s->aux_stream>>>((const uint32_t*)d_comm48, (const uint32_t*)nullptr, (int)n, s->aff_c.as<G1Affine>(), (G1Affine*)nullptr,
                                                              s->st_a.as<int32_t>(), (int32_t*)nullptr, 0);
    s->launches++;
    CU(cudaEventRecord(s->ev_join, s->aux_stream));
    LAUNCH(s, k_blob_check, blocks(n * KB_N, 256), 256, (const uint32_t*)d_blobs, (int)n, s->st_blob.as<int32_t>());
    if (!s->z_ready) LAUNCH(s, k_challenge_w2, blocks(n, CH2_THREADS / 2), CH2_THREADS, (const uint32_t*)d_blobs, (const uint32_t*)d_comm48, (int)n, s->st_blob.as<int32_t>(), (uint32_t*)nullptr, s->z.as<Fr>(), 0, 2050);
    s->z_ready = false;
    CU(cudaStreamWaitEvent(s->stream, s->ev_join, 0));
    LAUNCH(s, k_merge_status, blocks(n, 256), 256, s->st_blob.as<int32_t>(), s->st_a.as<int32_t>(), (const int32_t*)nullptr, (int)n, d_status);
    if (int rc = prove_tail(s, (const uint32_t*)d_blobs, n, d_status, (uint32_t*)d_out48)) return rc;
    std::vector<int32_t> hb, ha;
    if (int rc = fetch_status(s, s->st_blob.as<int32_t>(), n, hb)) return rc;
    if (int rc = fetch_status(s, s->st_a.as<int32_t>(), n, ha)) return rc;
    if (int e = first_nonzero(hb)) return e;
    return first_nonzero(ha);
}

extern "C" int kb200_compute_proof_device(kb200_setup* s, const void* d_blobs, const void* d_z32, size_t n, void* d_out48, void* d_y32, int32_t* d_status) {
    if (!s || (n && (!d_blobs || !d_z32 || !d_out48 || !d_y32 || !d_status))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (n == 0) return 0;
    if (int rc = s->st_blob.ensure(n * 4)) return rc;
    if (int rc = s->st_a.ensure(n * 4)) return rc;
    if (int rc = s->z.ensure(n * sizeof(Fr))) return rc;
    CU(cudaMemsetAsync(s->st_blob.p, 0, n * 4, s->stream));
    CU(cudaMemsetAsync(s->st_a.p, 0, n * 4, s->stream));
    LAUNCH(s, k_blob_check, blocks(n * KB_N, 256), 256, (const uint32_t*)d_blobs, (int)n, s->st_blob.as<int32_t>());
    LAUNCH(s, k_scalar_check, blocks(n, 128), 128, (const uint32_t*)d_z32, (int)n, s->z.as<Fr>(), s->st_a.as<int32_t>());
    LAUNCH(s, k_merge_status, blocks(n, 256), 256, s->st_blob.as<int32_t>(), s->st_a.as<int32_t>(), (const int32_t*)nullptr, (int)n, d_status);
    if (int rc = prove_tail(s, (const uint32_t*)d_blobs, n, d_status, (uint32_t*)d_out48)) return rc;
    LAUNCH(s, k_store_scalars_be, blocks(n, 128), 128, s->out32.as<Fr>(), (int)n, d_status, (uint32_t*)d_y32);
    std::vector<int32_t> hb, ha;
    if (int rc = fetch_status(s, s->st_blob.as<int32_t>(), n, hb)) return rc;
    if (int rc = fetch_status(s, s->st_a.as<int32_t>(), n, ha)) return rc;
    if (int e = first_nonzero(hb)) return e;
    return first_nonzero(ha);
}

// ---- verification ----
// host SHA-256 (only for the optional consensus-spec batch challenge: one sequential hash over the whole transcript)
static void host_sha256(const uint8_t* data, size_t len, uint8_t out[32]) {
    static const uint32_t Kc[64] = {
        0x428a2f98, 0x71374491, 0xb5c0fbcf, 0xe9b5dba5, 0x3956c25b, 0x59f111f1, 0x923f82a4, 0xab1c5ed5, 0xd807aa98, 0x12835b01, 0x243185be, 0x550c7dc3,
        0x72be5d74, 0x80deb1fe, 0x9bdc06a7, 0xc19bf174, 0xe49b69c1, 0xefbe4786, 0x0fc19dc6, 0x240ca1cc, 0x2de92c6f, 0x4a7484aa, 0x5cb0a9dc, 0x76f988da,
        0x983e5152, 0xa831c66d, 0xb00327c8, 0xbf597fc7, 0xc6e00bf3, 0xd5a79147, 0x06ca6351, 0x14292967, 0x27b70a85, 0x2e1b2138, 0x4d2c6dfc, 0x53380d13,
        0x650a7354, 0x766a0abb, 0x81c2c92e, 0x92722c85, 0xa2bfe8a1, 0xa81a664b, 0xc24b8b70, 0xc76c51a3, 0xd192e819, 0xd6990624, 0xf40e3585, 0x106aa070,
        0x19a4c116, 0x1e376c08, 0x2748774c, 0x34b0bcb5, 0x391c0cb3, 0x4ed8aa4a, 0x5b9cca4f, 0x682e6ff3, 0x748f82ee, 0x78a5636f, 0x84c87814, 0x8cc70208,
        0x90befffa, 0xa4506ceb, 0xbef9a3f7, 0xc67178f2};
    uint32_t h[8] = {0x6a09e667, 0xbb67ae85, 0x3c6ef372, 0xa54ff53a, 0x510e527f, 0x9b05688c, 0x1f83d9ab, 0x5be0cd19};
    auto rotr = [](uint32_t x, int k) { return (x >> k) | (x << (32 - k)); };
    auto block = [&](const uint8_t* b) {
        uint32_t w[64];
        for (int i = 0; i < 16; i++) w[i] = (uint32_t)b[4 * i] << 24 | (uint32_t)b[4 * i + 1] << 16 | (uint32_t)b[4 * i + 2] << 8 | b[4 * i + 3];
        for (int i = 16; i < 64; i++) {
            uint32_t s0 = rotr(w[i - 15], 7) ^ rotr(w[i - 15], 18) ^ (w[i - 15] >> 3), s1 = rotr(w[i - 2], 17) ^ rotr(w[i - 2], 19) ^ (w[i - 2] >> 10);
            w[i] = w[i - 16] + s0 + w[i - 7] + s1;
        }
        uint32_t a = h[0], bb = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
        for (int i = 0; i < 64; i++) {
            uint32_t t1 = hh + (rotr(e, 6) ^ rotr(e, 11) ^ rotr(e, 25)) + ((e & f) ^ (~e & g)) + Kc[i] + w[i];
            uint32_t t2 = (rotr(a, 2) ^ rotr(a, 13) ^ rotr(a, 22)) + ((a & bb) ^ (a & c) ^ (bb & c));
            hh = g; g = f; f = e; e = d + t1; d = c; c = bb; bb = a; a = t1 + t2;
        }
        h[0] += a; h[1] += bb; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
    };
    size_t off = 0;
    for (; off + 64 <= len; off += 64) block(data + off);
    uint8_t tail[128] = {0};
    size_t rem = len - off;
    memcpy(tail, data + off, rem);
    tail[rem] = 0x80;
    size_t tl = (rem < 56) ? 64 : 128;
    uint64_t bits = (uint64_t)len * 8;
    for (int i = 0; i < 8; i++) tail[tl - 1 - i] = (uint8_t)(bits >> (8 * i));
    block(tail);
    if (tl == 128) block(tail + 64);
    for (int i = 0; i < 8; i++) { out[4 * i] = h[i] >> 24; out[4 * i + 1] = h[i] >> 16; out[4 * i + 2] = h[i] >> 8; out[4 * i + 3] = h[i]; }
}


// first error of each class over the items of a call, in kateth's order (src/kzg/setup.rs:259-271)
struct VErr {
    int blob = 0, comm = 0, proof = 0, fourth = 0;     // point batches (verify_proof_batch): proof, commitment, z, y in these four
    int first() const { return blob ? blob : (comm ? comm : (proof ? proof : fourth)); }
};
// workspace of a verification call over n items (must precede verify_prepare: growing a buffer drops its contents)
static int verify_reserve(kb200_setup* s, size_t n) {
    if (n == 0) n = 1;
    DBuf* st[] = {&s->st_blob, &s->st_a, &s->st_b};
    for (DBuf* b : st) if (int rc = b->ensure(n * 4)) return rc;
    if (int rc = s->aff_c.ensure(n * sizeof(G1Affine))) return rc;
    if (int rc = s->aff_p.ensure(n * sizeof(G1Affine))) return rc;
    if (int rc = s->z.ensure(n * sizeof(Fr))) return rc;
    return s->y.ensure(n * sizeof(Fr));
}

// RLC sums of a shard of n items (affine points in s->aff_p / s->aff_c, z and y in s->z / s->y, per-item status in d_status or null)
// -> one record.  d_c48 / d_p48 (the compressed inputs, device) are only read in consensus-spec challenge mode.
static int run_rlc(kb200_setup* s, size_t n, size_t index_offset, size_t n_total, int raw, RlcRecord* d_record, const int32_t* d_status,
                   const uint32_t* d_c48 = nullptr, const uint32_t* d_p48 = nullptr) {
    int nb = blocks(2 * n, RLC_THREADS);          // two threads per term (GLV halves); 0 for an empty shard
    const bool buckets = s->rlc_mode == 2 ? n > 0 : (s->rlc_mode == 0 && n >= RLC_BUCKET_MIN);
    const int nslices = (int)((n + RLCB_ITEMS - 1) / RLCB_ITEMS);
    if (int rc = s->rlc_part.ensure((size_t)3 * (nb > RLCB_WINDOWS ? nb : RLCB_WINDOWS) * sizeof(G1))) return rc;
    if (buckets) {
        if (int rc = s->rlc_dig.ensure((size_t)2 * RLCB_WINDOWS * n * sizeof(uint16_t))) return rc;
        if (int rc = s->rlc_wsum.ensure((size_t)3 * RLCB_WINDOWS * nslices * sizeof(G1))) return rc;
    }
    if (int rc = s->rlc_s.ensure((size_t)(nb ? nb : 1) * sizeof(Fr))) return rc;
    if (int rc = s->rmont.ensure(sizeof(Fr))) return rc;
    const int spec = (!raw && s->challenge_mode == 1) ? 1 : 0;
    if (spec) {
        if (s->eval_pending) { CU(cudaStreamWaitEvent(s->stream, s->ev_eval, 0)); s->eval_pending = false; }      // the transcript hashes every y
        // r = hash_to_bls_field(domain | u64(4096) | u64(n) | sum(commitment | z | y | proof)): needs the whole batch in one place
        if (n != n_total || index_offset != 0 || !d_c48 || !d_p48) {
            snprintf(g_err, sizeof g_err, "consensus-spec batch challenge: the batch must be verified in one call (no shards)");
            return KB200_E_ARG;
        }
        if (int rc = s->transcript.ensure(n * 160 + 32)) return rc;
        LAUNCH(s, k_transcript, blocks(n, 128), 128, d_c48, d_p48, s->z.as<Fr>(), s->y.as<Fr>(), (int)n, s->transcript.as<uint32_t>());
        std::vector<uint8_t> t(32 + n * 160);
        memcpy(t.data(), "RCKZGBATCH___V1_", 16);
        for (int i = 0; i < 8; i++) { t[16 + i] = (uint8_t)((uint64_t)KB_N >> (8 * (7 - i))); t[24 + i] = (uint8_t)((uint64_t)n >> (8 * (7 - i))); }
        CU(cudaMemcpyAsync(t.data() + 32, s->transcript.p, n * 160, cudaMemcpyDeviceToHost, s->stream));
        CU(cudaStreamSynchronize(s->stream));
        uint8_t dg[32]; host_sha256(t.data(), t.size(), dg);
        uint32_t plain[8];
        for (int k = 0; k < 8; k++) plain[7 - k] = (uint32_t)dg[4 * k] << 24 | (uint32_t)dg[4 * k + 1] << 16 | (uint32_t)dg[4 * k + 2] << 8 | dg[4 * k + 3];
        uint32_t* d_plain = s->transcript.as<uint32_t>();           // reuse the head of the transcript buffer
        CU(cudaMemcpyAsync(d_plain, plain, 32, cudaMemcpyHostToDevice, s->stream));
        LAUNCH(s, k_set_r, 1, 1, d_plain, s->rmont.as<Fr>());
    } else if (!raw) {
        LAUNCH(s, k_batch_r, 1, 1, (uint32_t)n_total, s->rmont.as<Fr>());
    }
    if (buckets) {
        LAUNCH(s, k_rlc_scalars, blocks(n, 128), 128, s->z.as<Fr>(), (int)n, (uint32_t)index_offset, s->rmont.as<Fr>(), raw, spec, d_status, s->rlc_dig.as<uint16_t>());
        LAUNCH(s, k_rlc_bucket, dim3(RLCB_WINDOWS * nslices, 3), RLCB_THREADS, s->aff_p.as<G1Affine>(), s->aff_c.as<G1Affine>(), s->rlc_dig.as<uint16_t>(), (int)n, nslices,
               s->rlc_wsum.as<G1>());
        LAUNCH(s, k_rlc_shift, dim3(RLCB_WINDOWS, 3), 32, s->rlc_wsum.as<G1>(), nslices, s->rlc_part.as<G1>());
    } else if (nb)
        LAUNCH(s, k_rlc, dim3(nb, 3), RLC_THREADS, s->aff_p.as<G1Affine>(), s->aff_c.as<G1Affine>(), s->z.as<Fr>(), s->y.as<Fr>(), (int)n, (uint32_t)index_offset,
               s->rmont.as<Fr>(), raw, spec, d_status, s->rlc_part.as<G1>(), (Fr*)nullptr);
    TRACE(s, "k_rlc");
    if (s->eval_pending) { CU(cudaStreamWaitEvent(s->stream, s->ev_eval, 0)); s->eval_pending = false; }      // y: see verify_prepare
    if (nb) LAUNCH(s, k_rlc_s, nb, RLC_THREADS / 2, s->y.as<Fr>(), (int)n, (uint32_t)index_offset, s->rmont.as<Fr>(), raw, spec, d_status, s->rlc_s.as<Fr>());
    TRACE(s, "join evaluation + k_rlc_s");
    LAUNCH(s, k_rlc_finish, 3, RLC_FIN_THREADS, s->rlc_part.as<G1>(), buckets ? RLCB_WINDOWS : nb, s->rlc_s.as<Fr>(), nb, d_record);     // nb == 0: the identity record
    TRACE(s, "k_rlc_finish");
    return 0;
}
// the shard's errors travel with its sums (k_rlc_finish wrote zeros)
static int set_record_err(kb200_setup* s, RlcRecord* d_record, const VErr& e) {
    if (!e.first()) return 0;
    const int32_t h[4] = {e.blob, e.comm, e.proof, e.fourth};
    CU(cudaMemcpyAsync(d_record->err, h, sizeof h, cudaMemcpyHostToDevice, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return 0;
}
static void u32_to_be48(uint8_t* out, const uint32_t* limbs12) { for (int i = 0; i < 12; i++) { uint32_t w = limbs12[11 - i]; out[4 * i] = w >> 24; out[4 * i + 1] = w >> 16; out[4 * i + 2] = w >> 8; out[4 * i + 3] = w; } }
static void compress_plain(uint8_t out[48], const uint32_t* xy24, bool inf) {
    if (inf) { memset(out, 0, 48); out[0] = 0xc0; return; }
    u32_to_be48(out, xy24);
    out[0] |= 0x80;
    // y > (p-1)/2 ?
    using namespace kb200::consts;
    bool gt = false;
    for (int i = 11; i >= 0; i--) { if (xy24[12 + i] > H_FP_HALF[i]) { gt = true; break; } if (xy24[12 + i] < H_FP_HALF[i]) break; }
    if (gt) out[0] |= 0x20;
}
// records of all shards -> error of the whole batch in kateth's order, or the verdict of the two-pairing check
static int finish_records(kb200_setup* s, const RlcRecord* d_records, size_t nrec, int* ok) {
    if (int rc = s->combine.ensure(64 * 4)) return rc;
    LAUNCH(s, k_rlc_combine, 1, 32, d_records, (int)nrec, s->d_neg_g_pow, s->combine.as<uint32_t>());
    uint32_t h[54];
    CU(cudaMemcpyAsync(h, s->combine.p, sizeof h, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    CU(cudaGetLastError());
    *ok = 0;
    for (int k = 50; k < 54; k++) if (h[k]) return (int)h[k];      // Err(..): no verdict (src/kzg/setup.rs:259-271)
    host::G1A a, b;
    a.inf = h[48] != 0; b.inf = h[49] != 0;
    host::fp_from_plain_u32(a.x, h); host::fp_from_plain_u32(a.y, h + 12);
    host::fp_from_plain_u32(b.x, h + 24); host::fp_from_plain_u32(b.y, h + 36);
    compress_plain(s->last_lhs, h, a.inf); compress_plain(s->last_rhs, h + 24, b.inf); s->have_last = true;
    // bls::verify_pairings((lhs, [tau]_2), (rhs, G2))  (src/kzg/setup.rs:157-160, src/bls.rs:572-598)
    *ok = host::verify_pairings(a, s->tau_prep, b, s->g2_prep) ? 1 : 0;
    return 0;
}
extern "C" int kb200_last_pairing_operands(kb200_setup* s, uint8_t lhs48[48], uint8_t rhs48[48]) {
    if (!s) return KB200_E_ARG;
    LOCK(s);
    if (!s->have_last) return KB200_E_ARG;
    memcpy(lhs48, s->last_lhs, 48); memcpy(rhs48, s->last_rhs, 48); return 0;
}

// decode + decompress + challenge + evaluate for items [off, off + m) of a call whose workspace verify_reserve sized: d_blobs,
// d_c48, d_p48, d_status point at item `off`; fills aff_c, aff_p, z, y at `off`.  Accumulates the first error of each class in e.
static int verify_prepare(kb200_setup* s, const uint32_t* d_blobs, const uint32_t* d_c48, const uint32_t* d_p48, size_t m, size_t off, int32_t* d_status,
                          VErr& e, bool last = true) {
    int32_t *st_blob = s->st_blob.as<int32_t>() + off, *st_a = s->st_a.as<int32_t>() + off, *st_b = s->st_b.as<int32_t>() + off;
    G1Affine *aff_c = s->aff_c.as<G1Affine>() + off, *aff_p = s->aff_p.as<G1Affine>() + off;
    Fr *z = s->z.as<Fr>() + off, *y = s->y.as<Fr>() + off;
    const int n = (int)m;
    CU(cudaMemsetAsync(st_blob, 0, m * 4, s->stream));
    TRACE(s, "start");
    if (!s->fork_recorded) CU(cudaEventRecord(s->ev_fork, s->stream));      // a host-buffer caller marked the fork point before the blob upload it queued; resident data: fork here
    s->fork_recorded = false;
    {
        // The point decompression (one ~1500-product dependent chain per point, 64 CTAs per 4096 blobs) runs on the auxiliary stream
        // beside the challenges (64 CTAs of SHA-256 chains) -- for a host-buffer call also under the upload of the blobs.  The two
        // kernels are shaped so that they land on DIFFERENT SMs: sharing an SM stretches the SHA chains from 2.4 to 5 ms (measured with
        // 128 + 128 smaller CTAs), on SMs of their own neither slows the other.
        CU(cudaStreamWaitEvent(s->aux_stream, s->ev_fork, 0));
        CU(cudaMemsetAsync(st_a, 0, m * 4, s->aux_stream));
        CU(cudaMemsetAsync(st_b, 0, m * 4, s->aux_stream));
        k_g1_decompress<<<blocks(2 * m, 128), 128, 0, s->aux_stream>>>(d_c48, d_p48, n, aff_c, aff_p, st_a, st_b, 0);
        s->launches++;
        CU(cudaEventRecord(s->ev_join, s->aux_stream));
    }
    LAUNCH(s, k_blob_check, blocks(m * KB_N, 256), 256, d_blobs, n, st_blob);
    if (!s->z_ready) LAUNCH(s, k_challenge_w2, blocks(m, CH2_THREADS / 2), CH2_THREADS, d_blobs, d_c48, n, st_blob, (uint32_t*)nullptr, z, 0, 2050);
    s->z_ready = false;
    TRACE(s, "blob check + challenge");
    CU(cudaStreamWaitEvent(s->stream, s->ev_join, 0));
    CU(cudaEventRecord(s->ev_fork, s->stream));            // z is known: from here the evaluation may run on the auxiliary stream
    auto launch_eval = [&]() -> int {
        CU(cudaStreamWaitEvent(s->aux_stream, s->ev_fork, 0));
        cudaStream_t keep = s->stream; s->stream = s->aux_stream;
        int rc = run_eval<false>(s, d_blobs, m, st_blob, z, y);
        s->stream = keep;
        if (rc) return rc;
        CU(cudaEventRecord(s->ev_eval, s->aux_stream));
        s->eval_pending = true;
        if (!last) { CU(cudaStreamWaitEvent(s->stream, s->ev_eval, 0)); s->eval_pending = false; }
        return 0;
    };
    // The evaluations y are only needed for the scalar s of the record; the scalar multiplications of k_rlc need z and the points.  So
    // the evaluation (multiplier-bound) goes to the auxiliary stream HERE, beside the decompression and k_rlc (128-step ladders on few warps),
    // and for the last (or only) chunk of a call run_rlc joins it after k_rlc; an earlier chunk joins at once, because the next chunk's
    // upload reuses the blob buffer the evaluation reads.  It skips blobs that failed their own check; a bad point does not matter to it.
    if (int rc = launch_eval()) return rc;      // measured (4096 blobs): 8.45 ms per shard this way, 9.1-9.4 forked after the decompression, 9.7 with no overlap
    TRACE(s, "decompress");
    LAUNCH(s, k_merge_status, blocks(m, 256), 256, st_blob, st_a, st_b, n, d_status);
    std::vector<int32_t> h0, h1, h2;
    TRACE(s, "merge status");
    if (int rc = fetch_status(s, st_blob, m, h0)) return rc;
    if (int rc = fetch_status(s, st_a, m, h1)) return rc;
    if (int rc = fetch_status(s, st_b, m, h2)) return rc;
    if (!e.blob) e.blob = first_nonzero(h0);
    if (!e.comm) e.comm = first_nonzero(h1);
    if (!e.proof) e.proof = first_nonzero(h2);
    TRACE(s, "fetch statuses (host sync)");
    return 0;
}
extern "C" int kb200_verify_blob_proof_batch_partial_device(kb200_setup* s, const void* d_blobs, const void* d_c48, const void* d_p48, size_t n,
                                                            size_t index_offset, size_t n_total, void* d_record, int32_t* d_status) {
    if (!s || !d_record || (n && (!d_blobs || !d_c48 || !d_p48 || !d_status))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    Trace trace; g_trace = trace.on ? &trace : nullptr;
    struct Untrace { ~Untrace() { g_trace = nullptr; } } untrace;
    VErr e;
    if (int rc = verify_reserve(s, n)) return rc;
    if (n) if (int rc = verify_prepare(s, (const uint32_t*)d_blobs, (const uint32_t*)d_c48, (const uint32_t*)d_p48, n, 0, d_status, e)) return rc;
    if (int rc = run_rlc(s, n, index_offset, n_total, 0, (RlcRecord*)d_record, n ? d_status : nullptr, (const uint32_t*)d_c48, (const uint32_t*)d_p48)) return rc;
    if (int rc = set_record_err(s, (RlcRecord*)d_record, e)) return rc;
    CU(cudaStreamSynchronize(s->stream));
    CU(cudaGetLastError());
    return e.first();
}
extern "C" int kb200_verify_batch_finish_device(kb200_setup* s, const void* d_records, size_t n_records, int* ok) {
    if (!s || !ok || (n_records && !d_records)) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    return finish_records(s, (const RlcRecord*)d_records, n_records, ok);
}
extern "C" int kb200_verify_blob_proof_batch_device(kb200_setup* s, const void* d_blobs, const void* d_c48, const void* d_p48, size_t n, int* ok,
                                                    int32_t* d_status) {
    if (!s || !ok) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (n == 0) { *ok = 1; return 0; }     // empty lincombs are the identity: e(O, .) e(O, .) == 1
    if (int rc = s->records.ensure(sizeof(RlcRecord))) return rc;
    int err = kb200_verify_blob_proof_batch_partial_device(s, d_blobs, d_c48, d_p48, n, 0, n, s->records.p, d_status);
    if (err < 0) return err;
    return finish_records(s, s->records.as<RlcRecord>(), 1, ok);     // returns the record's error, if any
}

// ------------------------------------------------------------------ host-buffer entry points (copy in, run, copy out)
static int h2d(kb200_setup* s, DBuf& b, const void* src, size_t bytes) {
    if (int rc = b.ensure(bytes ? bytes : 16)) return rc;
    if (bytes) CU(cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, s->stream));
    return 0;
}
static int d2h(kb200_setup* s, void* dst, const void* src, size_t bytes) {
    if (bytes) CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return 0;
}
// Large host-buffer commit batches cross PCIe in two pieces on the copy stream: a head of one MSM wave (ba_grid blobs: one blob per
// resident CTA of k_msm_ba) and the rest, so the rest's copy (~8 ms for 4096 blobs) runs under the head's MSM (~11 ms) instead
// of in front of everything.  The head is exactly one wave, so the second MSM launch starts as the first one drains.
static size_t head_items(kb200_setup* s, size_t m) { return (s->split_copy && s->msm_mode != 1 && m >= 2 * (size_t)s->ba_grid) ? (size_t)s->ba_grid : m; }
struct CopyGuard {      // the caller's host buffer must not be read after the entry point returns, whatever the exit path
    kb200_setup* s;
    ~CopyGuard() { if (s->copy_stream) cudaStreamSynchronize(s->copy_stream); s->z_ready = false; s->fork_recorded = false; }
};
// copies `bytes` into b in two pieces; the main stream waits for the first (head_bytes), ev_tail marks the second
static int h2d_split(kb200_setup* s, DBuf& b, const uint8_t* src, size_t bytes, size_t head_bytes) {
    if (int rc = b.ensure(bytes ? bytes : 16)) return rc;
    if (!bytes) return 0;
    CU(cudaEventRecord(s->ev_copy, s->stream));                  // b may still be read by work queued earlier on the main stream
    CU(cudaStreamWaitEvent(s->copy_stream, s->ev_copy, 0));
    CU(cudaMemcpyAsync(b.p, src, head_bytes, cudaMemcpyHostToDevice, s->copy_stream));
    CU(cudaEventRecord(s->ev_head, s->copy_stream));
    if (bytes > head_bytes) {
        CU(cudaMemcpyAsync((char*)b.p + head_bytes, src + head_bytes, bytes - head_bytes, cudaMemcpyHostToDevice, s->copy_stream));
        CU(cudaEventRecord(s->ev_tail, s->copy_stream));
    }
    CU(cudaStreamWaitEvent(s->stream, s->ev_head, 0));
    return 0;
}
// Upload for the calls that start with the Fiat-Shamir challenge (blob proof, blob verification).  The challenge is one serial
// SHA-256 stream per blob, ~4 ms per launch whatever the batch size, and it cannot start before the blob is there -- so the
// blobs are copied in NCOL = 8 column pieces (bytes [j*W, (j+1)*W) of every blob, a strided 2-D copy) and the hash advances over
// each piece while the next one is on the wire: the batch is hashed ~1 ms after its last byte arrives instead of ~4 ms.
// Commitments must already be queued to d_c48 on the main stream.  Leaves the challenges in d_z (n items) and sets z_ready.
#ifndef KB_NCOL
#define KB_NCOL 8
#endif
static const int NCOL = KB_NCOL;     // <= 8 (ev_col)
static const size_t HASHED_UPLOAD_MIN = 1024;     // below this the upload is shorter than one hash launch
static int h2d_blobs_hashed(kb200_setup* s, const uint8_t* blobs, size_t n, const uint32_t* d_c48, Fr* d_z) {
    if (int rc = s->blobs.ensure(n * KB_BLOB_BYTES)) return rc;
    if (!s->split_copy || n < HASHED_UPLOAD_MIN) {
        CU(cudaMemcpyAsync(s->blobs.p, blobs, n * KB_BLOB_BYTES, cudaMemcpyHostToDevice, s->stream));
        return 0;
    }
    if (int rc = s->sha_state.ensure(n * 32)) return rc;
    const size_t W = KB_BLOB_BYTES / NCOL;
    CU(cudaEventRecord(s->ev_fork, s->stream));      // commitments and proofs are queued: their decompression need not wait for the blobs
    s->fork_recorded = true;
    CU(cudaEventRecord(s->ev_copy, s->stream));
    CU(cudaStreamWaitEvent(s->copy_stream, s->ev_copy, 0));
    for (int j = 0; j < NCOL; j++) {
        CU(cudaMemcpy2DAsync((char*)s->blobs.p + j * W, KB_BLOB_BYTES, blobs + j * W, KB_BLOB_BYTES, W, n, cudaMemcpyHostToDevice, s->copy_stream));
        CU(cudaEventRecord(s->ev_col[j], s->copy_stream));
    }
    for (int j = 0; j < NCOL; j++) {
        CU(cudaStreamWaitEvent(s->stream, s->ev_col[j], 0));
        const int blk0 = (int)(j * W / 64), blk1 = (j == NCOL - 1) ? 2050 : (int)((j + 1) * W / 64);
        LAUNCH(s, k_challenge_w2, blocks(n, CH2_THREADS / 2), CH2_THREADS, s->blobs.as<uint32_t>(), d_c48, (int)n, (const int32_t*)nullptr, s->sha_state.as<uint32_t>(), d_z, blk0, blk1);
    }
    s->z_ready = true;
    return 0;
}
extern "C" int kb200_blob_to_commitment(kb200_setup* s, const uint8_t* blobs, size_t n, uint8_t* out48, int32_t* status) {
    if (!s || (n && (!blobs || !out48))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    CopyGuard guard{s};
    int first = 0;
    for (size_t off = 0; off < n; off += CHUNK) {
        const size_t m = (n - off < CHUNK) ? n - off : CHUNK, h = head_items(s, m);
        if (int rc = h2d_split(s, s->blobs, blobs + off * KB_BLOB_BYTES, m * KB_BLOB_BYTES, h * KB_BLOB_BYTES)) return rc;
        if (int rc = s->out48.ensure(m * 48)) return rc;
        if (int rc = s->status.ensure(m * 4)) return rc;
        for (size_t o = 0, k = h; o < m; o += k, k = m - o) {       // head, then the rest
            if (o) CU(cudaStreamWaitEvent(s->stream, s->ev_tail, 0));
            int e = kb200_blob_to_commitment_device(s, (char*)s->blobs.p + o * KB_BLOB_BYTES, k, (char*)s->out48.p + o * 48, s->status.as<int32_t>() + o);
            if (e < 0) return e;
            if (e && !first) first = e;
        }
        if (int rc = d2h(s, out48 + off * 48, s->out48.p, m * 48)) return rc;
        if (status) if (int rc = d2h(s, status + off, s->status.p, m * 4)) return rc;
    }
    return first;
}
// Blob::from_slice alone (src/blob.rs:26-37) over n blobs: status[i] = 0 or InvalidFieldElement.  For callers that hold blobs of
// unequal byte lengths and must report kateth's first error in blob order before the batch can be laid out contiguously.
extern "C" int kb200_check_blobs(kb200_setup* s, const uint8_t* blobs, size_t n, int32_t* status) {
    if (!s || (n && !blobs)) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    int first = 0;
    for (size_t off = 0; off < n; off += CHUNK) {
        const size_t m = (n - off < CHUNK) ? n - off : CHUNK;
        if (int rc = h2d(s, s->blobs, blobs + off * KB_BLOB_BYTES, m * KB_BLOB_BYTES)) return rc;
        if (int rc = s->status.ensure(m * 4)) return rc;
        CU(cudaMemsetAsync(s->status.p, 0, m * 4, s->stream));
        LAUNCH(s, k_blob_check, blocks(m * KB_N, 256), 256, s->blobs.as<uint32_t>(), (int)m, s->status.as<int32_t>());
        std::vector<int32_t> h; if (int rc = fetch_status(s, s->status.as<int32_t>(), m, h)) return rc;
        if (status) memcpy(status + off, h.data(), m * 4);
        if (!first) first = first_nonzero(h);
    }
    return first;
}
// (No head/rest copy here: the proof path runs latency-bound per-blob kernels in front of its MSM -- the serial SHA-256 of the
// challenge takes ~4 ms whether it hashes 592 or 4096 blobs -- so running them twice costs what the hidden copy saves;
// measured with tools/e2e_probe.py: 97.1 ms split against 95.9 ms in one piece for 4096 blobs.  Column pieces instead.)
extern "C" int kb200_compute_blob_proof(kb200_setup* s, const uint8_t* blobs, const uint8_t* comm48, size_t n, uint8_t* out48, int32_t* status) {
    if (!s || (n && (!blobs || !comm48 || !out48))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    CopyGuard guard{s};
    // kateth's order within one call is blob first, then commitment; across a batch we report the first item's error
    int first = 0;
    for (size_t off = 0; off < n; off += CHUNK) {
        size_t m = (n - off < CHUNK) ? n - off : CHUNK;
        if (int rc = h2d(s, s->c48, comm48 + off * 48, m * 48)) return rc;
        if (int rc = s->z.ensure(m * sizeof(Fr))) return rc;
        if (int rc = h2d_blobs_hashed(s, blobs + off * KB_BLOB_BYTES, m, s->c48.as<uint32_t>(), s->z.as<Fr>())) return rc;
        if (int rc = s->out48.ensure(m * 48)) return rc;
        if (int rc = s->status.ensure(m * 4)) return rc;
        int e = kb200_compute_blob_proof_device(s, s->blobs.p, s->c48.p, m, s->out48.p, s->status.as<int32_t>());
        if (e < 0) return e;
        if (e && !first) first = e;
        if (int rc = d2h(s, out48 + off * 48, s->out48.p, m * 48)) return rc;
        if (status) if (int rc = d2h(s, status + off, s->status.p, m * 4)) return rc;
    }
    return first;
}
extern "C" int kb200_compute_proof(kb200_setup* s, const uint8_t* blobs, const uint8_t* z32, size_t n, uint8_t* out48, uint8_t* y32, int32_t* status) {
    if (!s || (n && (!blobs || !z32 || !out48 || !y32))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    int first = 0;
    for (size_t off = 0; off < n; off += CHUNK) {
        size_t m = (n - off < CHUNK) ? n - off : CHUNK;
        if (int rc = h2d(s, s->blobs, blobs + off * KB_BLOB_BYTES, m * KB_BLOB_BYTES)) return rc;
        if (int rc = h2d(s, s->zy32, z32 + off * 32, m * 32)) return rc;
        if (int rc = s->out48.ensure(m * 48)) return rc;
        if (int rc = s->p48.ensure(m * 32)) return rc;     // reused as the y output staging
        if (int rc = s->status.ensure(m * 4)) return rc;
        int e = kb200_compute_proof_device(s, s->blobs.p, s->zy32.p, m, s->out48.p, s->p48.p, s->status.as<int32_t>());
        if (e < 0) return e;
        if (e && !first) first = e;
        if (int rc = d2h(s, out48 + off * 48, s->out48.p, m * 48)) return rc;
        if (int rc = d2h(s, y32 + off * 32, s->p48.p, m * 32)) return rc;
        if (status) if (int rc = d2h(s, status + off, s->status.p, m * 4)) return rc;
    }
    return first;
}
// A host-buffer shard of any size: commitments and proofs go up whole, the blobs in chunks of CHUNK (each chunk's challenge,
// evaluation and point checks run as it lands); then ONE pass of RLC sums over all n items, so a batch larger than a chunk
// still yields one record -- and, in consensus-spec challenge mode, one transcript.
static int verify_host_shard(kb200_setup* s, const uint8_t* blobs, const uint8_t* comm48, const uint8_t* proofs48, size_t n, size_t index_offset,
                             size_t n_total, RlcRecord* d_record, int32_t* status, VErr& e) {
    if (int rc = verify_reserve(s, n)) return rc;
    if (int rc = s->status.ensure(n ? n * 4 : 16)) return rc;
    if (int rc = h2d(s, s->c48, comm48, n * 48)) return rc;
    if (int rc = h2d(s, s->p48, proofs48, n * 48)) return rc;
    for (size_t off = 0; off < n; off += CHUNK) {
        const size_t m = (n - off < CHUNK) ? n - off : CHUNK;
        if (int rc = h2d_blobs_hashed(s, blobs + off * KB_BLOB_BYTES, m, s->c48.as<uint32_t>() + off * 12, s->z.as<Fr>() + off)) return rc;
        if (int rc = verify_prepare(s, s->blobs.as<uint32_t>(), s->c48.as<uint32_t>() + off * 12, s->p48.as<uint32_t>() + off * 12, m, off,
                                    s->status.as<int32_t>() + off, e, off + m >= n)) return rc;
    }
    if (status && n) if (int rc = d2h(s, status, s->status.p, n * 4)) return rc;
    if (int rc = run_rlc(s, n, index_offset, n_total, 0, d_record, n ? s->status.as<int32_t>() : nullptr, s->c48.as<uint32_t>(), s->p48.as<uint32_t>())) return rc;
    return set_record_err(s, d_record, e);
}
extern "C" int kb200_verify_blob_proof_batch(kb200_setup* s, const uint8_t* blobs, const uint8_t* comm48, const uint8_t* proofs48, size_t n, int* ok,
                                             int32_t* status) {
    if (!s || !ok || (n && (!blobs || !comm48 || !proofs48))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (n == 0) { *ok = 1; return 0; }
    CopyGuard guard{s};
    if (int rc = s->records.ensure(sizeof(RlcRecord))) return rc;
    VErr e;
    // kateth checks ALL blobs, then ALL commitments, then ALL proofs (src/kzg/setup.rs:259-271): e keeps the first error of each
    // class across the chunks, the record carries them, finish_records returns them in that order
    if (int rc = verify_host_shard(s, blobs, comm48, proofs48, n, 0, n, s->records.as<RlcRecord>(), status, e)) return rc;
    return finish_records(s, s->records.as<RlcRecord>(), 1, ok);
}
// host-buffer shard of a multi-GPU batch: inputs in (pinned) host memory, the record stays on the device
extern "C" int kb200_verify_blob_proof_batch_partial(kb200_setup* s, const uint8_t* blobs, const uint8_t* comm48, const uint8_t* proofs48, size_t n,
                                                     size_t index_offset, size_t n_total, void* d_record, int32_t* status) {
    if (!s || !d_record || (n && (!blobs || !comm48 || !proofs48))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    CopyGuard guard{s};
    VErr e;
    if (int rc = verify_host_shard(s, blobs, comm48, proofs48, n, index_offset, n_total, (RlcRecord*)d_record, status, e)) return rc;
    CU(cudaStreamSynchronize(s->stream));
    CU(cudaGetLastError());
    return e.first();
}
// ---- commit + blob proof + batch-verification shard over ONE upload of the blobs ----
// The three calls a producer makes for a fresh batch (Setup::blob_to_commitment, ::blob_proof, then ::verify_blob_proof_batch on what
// it is about to publish) each read the same 128 KiB per blob; done as three host-buffer calls the batch crosses PCIe three times.
// Here a chunk is uploaded once (head wave first, so the first MSM starts under the rest of the copy) and the three device-side
// pipelines run on the resident bytes -- each of them complete: the verification shard decodes the 48-byte commitments and proofs
// again, recomputes every challenge and evaluation, exactly as kb200_verify_blob_proof_batch_partial would.
extern "C" int kb200_commit_prove_verify_partial(kb200_setup* s, const uint8_t* blobs, size_t n, size_t index_offset, size_t n_total, uint8_t* out_comm48,
                                                 uint8_t* out_proof48, void* d_record, int32_t* status) {
    if (!s || !d_record || (n && (!blobs || !out_comm48 || !out_proof48))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    CopyGuard guard{s};
    if (int rc = verify_reserve(s, n)) return rc;
    if (int rc = s->status.ensure(n ? n * 4 : 16)) return rc;
    if (int rc = s->c48.ensure(n ? n * 48 : 16)) return rc;
    if (int rc = s->p48.ensure(n ? n * 48 : 16)) return rc;
    if (int rc = s->combine.ensure(64 * 4)) return rc;
    DBuf& st_tmp = s->st_tmp;                      // per-chunk statuses of commit / proof (the verification shard keeps the merged ones)
    VErr e;
    int first = 0;
    if (int rc = st_tmp.ensure((n < CHUNK ? n : CHUNK) * 4 + 16)) return rc;
    // Chunks are taken from the LAST to the first: the proof pipeline of a chunk works in items [0, m) of the shared workspace, the
    // verification data of a chunk is kept at its own offset, so a later (lower) chunk never overwrites what an earlier one left;
    // errors of lower chunks replace those of higher ones, which keeps kateth's first-in-item-order rule.
    const size_t nch = (n + CHUNK - 1) / CHUNK;
    for (size_t c = nch; c-- > 0;) {
        const size_t off = c * CHUNK, m = (n - off < CHUNK) ? n - off : CHUNK, h = head_items(s, m);
        if (int rc = h2d_split(s, s->blobs, blobs + off * KB_BLOB_BYTES, m * KB_BLOB_BYTES, h * KB_BLOB_BYTES)) return rc;
        uint32_t* d_c = s->c48.as<uint32_t>() + off * 12;
        uint32_t* d_p = s->p48.as<uint32_t>() + off * 12;
        int chunk_first = 0;
        for (size_t o = 0, k = h; o < m; o += k, k = m - o) {       // commit: head, then the rest
            if (o) CU(cudaStreamWaitEvent(s->stream, s->ev_tail, 0));
            int rc = kb200_blob_to_commitment_device(s, (char*)s->blobs.p + o * KB_BLOB_BYTES, k, d_c + o * 12, st_tmp.as<int32_t>() + o);
            if (rc < 0) return rc;
            if (rc && !chunk_first) chunk_first = rc;
        }
        if (int rc = d2h(s, out_comm48 + off * 48, d_c, m * 48)) return rc;
        int rc = kb200_compute_blob_proof_device(s, s->blobs.p, d_c, m, d_p, st_tmp.as<int32_t>());
        if (rc < 0) return rc;
        if (rc && !chunk_first) chunk_first = rc;
        if (int rc2 = d2h(s, out_proof48 + off * 48, d_p, m * 48)) return rc2;
        VErr ec;
        if (int rc2 = verify_prepare(s, s->blobs.as<uint32_t>(), d_c, d_p, m, off, s->status.as<int32_t>() + off, ec, c == 0)) return rc2;
        if (chunk_first) first = chunk_first;
        if (ec.blob) e.blob = ec.blob;
        if (ec.comm) e.comm = ec.comm;
        if (ec.proof) e.proof = ec.proof;
    }
    if (status && n) if (int rc = d2h(s, status, s->status.p, n * 4)) return rc;
    if (int rc = run_rlc(s, n, index_offset, n_total, 0, (RlcRecord*)d_record, n ? s->status.as<int32_t>() : nullptr, s->c48.as<uint32_t>(), s->p48.as<uint32_t>())) return rc;
    if (int rc = set_record_err(s, (RlcRecord*)d_record, e)) return rc;
    CU(cudaStreamSynchronize(s->stream));
    CU(cudaGetLastError());
    return first ? first : e.first();
}

// ---- (proof, commitment, z, y) tuples given directly: the private Setup::verify_proof_batch (src/kzg/setup.rs:115-161) ----
// decode the n tuples at the device pointers into s->aff_p / aff_c / z / y (workspace sized by verify_reserve); merged per-item
// status in kateth's order (proof, commitment, point, eval: src/kzg/setup.rs:103-110) -> d_status; first error of each class -> e
static int points_prepare(kb200_setup* s, const uint32_t* d_p48, const uint32_t* d_c48, const uint32_t* d_z32, const uint32_t* d_y32, size_t n, int32_t* d_status,
                          VErr& e, std::vector<int32_t>* host_status = nullptr) {
    if (int rc = s->prod.ensure(n * 4)) return rc;   // 4th status class (y)
    DBuf* cls[] = {&s->st_blob, &s->st_a, &s->st_b, &s->prod};
    for (DBuf* b : cls) CU(cudaMemsetAsync(b->p, 0, n * 4, s->stream));
    LAUNCH(s, k_g1_decompress, blocks(2 * n, 64), 64, d_p48, d_c48, (int)n, s->aff_p.as<G1Affine>(), s->aff_c.as<G1Affine>(), s->st_blob.as<int32_t>(), s->st_a.as<int32_t>(), 0);
    LAUNCH(s, k_scalar_check, blocks(n, 128), 128, d_z32, (int)n, s->z.as<Fr>(), s->st_b.as<int32_t>());
    LAUNCH(s, k_scalar_check, blocks(n, 128), 128, d_y32, (int)n, s->y.as<Fr>(), s->prod.as<int32_t>());
    LAUNCH(s, k_merge_status, blocks(n, 256), 256, s->st_blob.as<int32_t>(), s->st_a.as<int32_t>(), s->st_b.as<int32_t>(), (int)n, d_status);
    LAUNCH(s, k_merge_status, blocks(n, 256), 256, d_status, s->prod.as<int32_t>(), (const int32_t*)nullptr, (int)n, d_status);
    std::vector<int32_t> h[4];
    for (int k = 0; k < 4; k++) if (int rc = fetch_status(s, cls[k]->as<int32_t>(), n, h[k])) return rc;
    e.blob = first_nonzero(h[0]); e.comm = first_nonzero(h[1]); e.proof = first_nonzero(h[2]); e.fourth = first_nonzero(h[3]);
    if (host_status) { host_status->resize(n); for (size_t i = 0; i < n; i++) (*host_status)[i] = h[0][i] ? h[0][i] : h[1][i] ? h[1][i] : h[2][i] ? h[2][i] : h[3][i]; }
    return 0;
}
// one shard of a sharded point batch, inputs in HBM: record (with the shard's errors) -> d_record
extern "C" int kb200_verify_proof_batch_partial_device(kb200_setup* s, const void* d_proofs48, const void* d_c48, const void* d_z32, const void* d_y32, size_t n,
                                                       size_t index_offset, size_t n_total, void* d_record, int32_t* d_status) {
    if (!s || !d_record || (n && (!d_proofs48 || !d_c48 || !d_z32 || !d_y32 || !d_status))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    VErr e;
    if (int rc = verify_reserve(s, n)) return rc;
    if (n) if (int rc = points_prepare(s, (const uint32_t*)d_proofs48, (const uint32_t*)d_c48, (const uint32_t*)d_z32, (const uint32_t*)d_y32, n, d_status, e)) return rc;
    if (int rc = run_rlc(s, n, index_offset, n_total, 0, (RlcRecord*)d_record, n ? d_status : nullptr, (const uint32_t*)d_c48, (const uint32_t*)d_proofs48)) return rc;
    if (int rc = set_record_err(s, (RlcRecord*)d_record, e)) return rc;
    CU(cudaStreamSynchronize(s->stream));
    CU(cudaGetLastError());
    return e.first();
}
// the same with the shard's inputs in host memory
extern "C" int kb200_verify_proof_batch_partial(kb200_setup* s, const uint8_t* proofs48, const uint8_t* comm48, const uint8_t* z32, const uint8_t* y32, size_t n,
                                                size_t index_offset, size_t n_total, void* d_record, int32_t* status) {
    if (!s || !d_record || (n && (!proofs48 || !comm48 || !z32 || !y32))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (int rc = h2d(s, s->p48, proofs48, n * 48)) return rc;
    if (int rc = h2d(s, s->c48, comm48, n * 48)) return rc;
    if (int rc = h2d(s, s->zy32, z32, n * 32)) return rc;
    if (int rc = h2d(s, s->out32, y32, n * 32)) return rc;
    if (int rc = s->status.ensure(n ? n * 4 : 16)) return rc;
    int e = kb200_verify_proof_batch_partial_device(s, s->p48.p, s->c48.p, s->zy32.p, s->out32.p, n, index_offset, n_total, d_record, s->status.as<int32_t>());
    if (e < 0) return e;
    if (status && n) if (int rc = d2h(s, status, s->status.p, n * 4)) return rc;
    return e;
}
// single call: mirrors Setup::verify_proof_batch (raw = 0) and verify_proof_inner (raw = 1)
static int verify_points(kb200_setup* s, const uint8_t* proofs48, const uint8_t* comm48, const uint8_t* z32, const uint8_t* y32, size_t n, int raw, int* ok,
                         int32_t* status) {
    if (!s || !ok || (n && (!proofs48 || !comm48 || !z32 || !y32))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (n == 0) { *ok = 1; return 0; }
    if (int rc = h2d(s, s->p48, proofs48, n * 48)) return rc;
    if (int rc = h2d(s, s->c48, comm48, n * 48)) return rc;
    if (int rc = h2d(s, s->zy32, z32, n * 32)) return rc;
    if (int rc = h2d(s, s->out32, y32, n * 32)) return rc;
    if (int rc = verify_reserve(s, n)) return rc;
    if (int rc = s->status.ensure(n * 4)) return rc;
    VErr e;
    std::vector<int32_t> hs;
    if (int rc = points_prepare(s, s->p48.as<uint32_t>(), s->c48.as<uint32_t>(), s->zy32.as<uint32_t>(), s->out32.as<uint32_t>(), n, s->status.as<int32_t>(), e, &hs)) return rc;
    if (status) memcpy(status, hs.data(), n * 4);
    if (e.first()) return e.first();
    if (int rc = s->records.ensure(sizeof(RlcRecord))) return rc;
    if (int rc = run_rlc(s, n, 0, n, raw, s->records.as<RlcRecord>(), nullptr, s->c48.as<uint32_t>(), s->p48.as<uint32_t>())) return rc;
    return finish_records(s, s->records.as<RlcRecord>(), 1, ok);
}
extern "C" int kb200_verify_proof(kb200_setup* s, const uint8_t proof48[48], const uint8_t comm48[48], const uint8_t z32[32], const uint8_t y32[32], int* ok) {
    return verify_points(s, proof48, comm48, z32, y32, 1, 1, ok, nullptr);
}
extern "C" int kb200_verify_proof_batch(kb200_setup* s, const uint8_t* proofs48, const uint8_t* comm48, const uint8_t* z32, const uint8_t* y32, size_t n,
                                        int* ok, int32_t* status) {
    return verify_points(s, proofs48, comm48, z32, y32, n, 0, ok, status);
}
// n independent Setup::verify_proof calls (src/kzg/setup.rs:96-113) in one launch: per-item verdicts from the device pairing
// (pairing.cuh).  ok[i] = 1 / 0; status[i] = the item's decoding error in kateth's order (proof, commitment, point, eval) or 0, and
// ok[i] = 0 for such an item.  Returns 0 unless the call itself failed: an invalid item is that item's Err, not the batch's.
extern "C" int kb200_verify_proof_many(kb200_setup* s, const uint8_t* proofs48, const uint8_t* comm48, const uint8_t* z32, const uint8_t* y32, size_t n,
                                       int32_t* ok, int32_t* status) {
    if (!s || (n && (!proofs48 || !comm48 || !z32 || !y32 || !ok))) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (n == 0) return 0;
    if (int rc = h2d(s, s->p48, proofs48, n * 48)) return rc;
    if (int rc = h2d(s, s->c48, comm48, n * 48)) return rc;
    if (int rc = h2d(s, s->zy32, z32, n * 32)) return rc;
    if (int rc = h2d(s, s->out32, y32, n * 32)) return rc;
    if (int rc = verify_reserve(s, n)) return rc;
    if (int rc = s->status.ensure(n * 4)) return rc;
    if (int rc = s->indom.ensure(n * 4)) return rc;
    VErr e;
    if (int rc = points_prepare(s, s->p48.as<uint32_t>(), s->c48.as<uint32_t>(), s->zy32.as<uint32_t>(), s->out32.as<uint32_t>(), n, s->status.as<int32_t>(), e)) return rc;
    LAUNCH(s, k_verify_proof_many, blocks(n, PAIR_THREADS), PAIR_THREADS, s->aff_p.as<G1Affine>(), s->aff_c.as<G1Affine>(), s->z.as<Fr>(), s->y.as<Fr>(),
           s->status.as<int32_t>(), s->d_neg_g_bits, s->d_pair, (int)n, s->indom.as<int32_t>());
    if (int rc = d2h(s, ok, s->indom.p, n * 4)) return rc;
    if (status) if (int rc = d2h(s, status, s->status.p, n * 4)) return rc;
    CU(cudaGetLastError());
    return 0;
}
extern "C" int kb200_verify_blob_proof(kb200_setup* s, const uint8_t* blob, const uint8_t comm48[48], const uint8_t proof48[48], int* ok) {
    if (!s || !ok || !blob || !comm48 || !proof48) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (int rc = h2d(s, s->blobs, blob, KB_BLOB_BYTES)) return rc;
    if (int rc = h2d(s, s->c48, comm48, 48)) return rc;
    if (int rc = h2d(s, s->p48, proof48, 48)) return rc;
    if (int rc = s->status.ensure(4)) return rc;
    if (int rc = verify_reserve(s, 1)) return rc;
    VErr e;
    if (int rc = verify_prepare(s, s->blobs.as<uint32_t>(), s->c48.as<uint32_t>(), s->p48.as<uint32_t>(), 1, 0, s->status.as<int32_t>(), e)) return rc;
    if (e.first()) {
        if (s->eval_pending) { CU(cudaStreamWaitEvent(s->stream, s->ev_eval, 0)); s->eval_pending = false; }      // nothing of this call may outlive it
        return e.first();
    }
    if (int rc = s->records.ensure(sizeof(RlcRecord))) return rc;
    if (int rc = run_rlc(s, 1, 0, 1, 1, s->records.as<RlcRecord>(), nullptr)) return rc;
    return finish_records(s, s->records.as<RlcRecord>(), 1, ok);
}

// ------------------------------------------------------------------ unit-test hooks
extern "C" int kb200_test_fp_mul(const uint32_t* a, const uint32_t* b, uint32_t* out, size_t n) {
    if (int rc = init_consts()) return rc;
    DevTmp ta, tb, to; CU(cudaMalloc(&ta.p, n * 48)); CU(cudaMalloc(&tb.p, n * 48)); CU(cudaMalloc(&to.p, n * 48));
    Fp *da = ta.as<Fp>(), *db = tb.as<Fp>(), *dout = to.as<Fp>();
    CU(cudaMemcpy(da, a, n * 48, cudaMemcpyHostToDevice)); CU(cudaMemcpy(db, b, n * 48, cudaMemcpyHostToDevice));
    k_test_fp_mul<<<blocks(n, 128), 128>>>(da, db, dout, (int)n);
    CU(cudaMemcpy(out, dout, n * 48, cudaMemcpyDeviceToHost));
    return 0;
}
extern "C" int kb200_test_fr_mul(const uint32_t* a, const uint32_t* b, uint32_t* out, size_t n) {
    if (int rc = init_consts()) return rc;
    DevTmp ta, tb, to; CU(cudaMalloc(&ta.p, n * 32)); CU(cudaMalloc(&tb.p, n * 32)); CU(cudaMalloc(&to.p, n * 32));
    Fr *da = ta.as<Fr>(), *db = tb.as<Fr>(), *dout = to.as<Fr>();
    CU(cudaMemcpy(da, a, n * 32, cudaMemcpyHostToDevice)); CU(cudaMemcpy(db, b, n * 32, cudaMemcpyHostToDevice));
    k_test_fr_mul<<<blocks(n, 128), 128>>>(da, db, dout, (int)n);
    CU(cudaMemcpy(out, dout, n * 32, cudaMemcpyDeviceToHost));
    return 0;
}
extern "C" int kb200_test_g1_decompress(kb200_setup* s, const uint8_t* in48, size_t n, int32_t* status, uint8_t* re48, int by_order) {
    if (!s || !in48 || !status || !re48) return KB200_E_ARG;
    LOCK(s);
    CU(cudaSetDevice(s->device));
    if (int rc = h2d(s, s->c48, in48, n * 48)) return rc;
    if (int rc = s->aff_c.ensure(n * sizeof(G1Affine))) return rc;
    if (int rc = s->status.ensure(n * 4)) return rc;
    if (int rc = s->out48.ensure(n * 48)) return rc;
    CU(cudaMemsetAsync(s->status.p, 0, n * 4, s->stream));
    LAUNCH(s, k_g1_decompress, blocks(n, 64), 64, s->c48.as<uint32_t>(), (const uint32_t*)nullptr, (int)n, s->aff_c.as<G1Affine>(), (G1Affine*)nullptr, s->status.as<int32_t>(), (int32_t*)nullptr, by_order);
    LAUNCH(s, k_test_compress, blocks(n, 64), 64, s->aff_c.as<G1Affine>(), (int)n, s->status.as<int32_t>(), s->out48.as<uint32_t>());
    if (int rc = d2h(s, status, s->status.p, n * 4)) return rc;
    return d2h(s, re48, s->out48.p, n * 48);
}
// P1::decompress without a setup (src/bls.rs:505-531: blst_p1_uncompress + blst_p1_affine_in_g1): status[i] = 0 / 3 / 4 / 5 for n encodings.
// kateth's Decompress is an associated function of the point type; a wrapper that offers it has no handle to pass.
extern "C" int kb200_g1_validate(const uint8_t* in48, size_t n, int device, int32_t* status) {
    if ((n && !in48) || !status) return KB200_E_ARG;
    if (n == 0) return 0;
    CU(cudaSetDevice(device));
    if (int rc = init_consts()) return rc;
    DevTmp tin, tpt, tst;
    CU(cudaMalloc(&tin.p, n * 48)); CU(cudaMalloc(&tpt.p, n * sizeof(G1Affine))); CU(cudaMalloc(&tst.p, n * 4));
    CU(cudaMemcpy(tin.p, in48, n * 48, cudaMemcpyHostToDevice));
    CU(cudaMemset(tst.p, 0, n * 4));
    k_g1_decompress<<<blocks(n, 64), 64>>>(tin.as<uint32_t>(), (const uint32_t*)nullptr, (int)n, tpt.as<G1Affine>(), (G1Affine*)nullptr, tst.as<int32_t>(), (int32_t*)nullptr, 0);
    CU(cudaMemcpy(status, tst.p, n * 4, cudaMemcpyDeviceToHost));
    CU(cudaGetLastError());
    return 0;
}
extern "C" int kb200_test_mac_peak(int iters, double* tmacs_wide, double* tmacs_fpmul) {
    if (int rc = init_consts()) return rc;
    cudaDeviceProp prop; CU(cudaGetDeviceProperties(&prop, 0));
    int grid = prop.multiProcessorCount * 8, block = 256;
    DevTmp tbuf; CU(cudaMalloc(&tbuf.p, (size_t)grid * block * 48));
    void* buf = tbuf.p;
    struct Ev { cudaEvent_t e = nullptr; ~Ev() { if (e) cudaEventDestroy(e); } } ev0, ev1;
    CU(cudaEventCreate(&ev0.e)); CU(cudaEventCreate(&ev1.e));
    cudaEvent_t e0 = ev0.e, e1 = ev1.e;
    float ms;
    k_mac_wide<<<grid, block>>>((uint64_t*)buf, 64, 1);   // warm-up
    CU(cudaEventRecord(e0)); k_mac_wide<<<grid, block>>>((uint64_t*)buf, iters, 7); CU(cudaEventRecord(e1)); CU(cudaEventSynchronize(e1));
    CU(cudaEventElapsedTime(&ms, e0, e1));
    *tmacs_wide = (double)grid * block * iters * 24.0 / (ms * 1e-3) / 1e12;
    int it2 = iters / 16 > 0 ? iters / 16 : 1;
    k_mac_fpmul<<<grid, block>>>((Fp*)buf, 4, 1);
    CU(cudaEventRecord(e0)); k_mac_fpmul<<<grid, block>>>((Fp*)buf, it2, 7); CU(cudaEventRecord(e1)); CU(cudaEventSynchronize(e1));
    CU(cudaEventElapsedTime(&ms, e0, e1));
    *tmacs_fpmul = (double)grid * block * it2 * 2.0 * 300.0 / (ms * 1e-3) / 1e12;
    return 0;
}
extern "C" int kb200_test_host_pairing(kb200_setup* s, const uint32_t* a_xy24, int a_inf, const uint32_t* b_xy24, int b_inf, int* ok) {
    if (!s || !ok) return KB200_E_ARG;
    host::G1A a, b; a.inf = a_inf != 0; b.inf = b_inf != 0;
    if (!a.inf) { host::fp_from_plain_u32(a.x, a_xy24); host::fp_from_plain_u32(a.y, a_xy24 + 12); }
    if (!b.inf) { host::fp_from_plain_u32(b.x, b_xy24); host::fp_from_plain_u32(b.y, b_xy24 + 12); }
    *ok = host::verify_pairings(a, s->tau_prep, b, s->g2_prep) ? 1 : 0;
    return 0;
}
