// group.cpp -- the multi-GPU front end of the engine, INSIDE the C ABI (include/kateth_b200.h, kb200_group_*).
//
// kateth's Setup answers one call with one result (`&self`, src/kzg/setup.rs:167-275).  A group keeps that shape over several
// B200s: one replica of the setup per GPU (tables replicated), a batch is cut into contiguous shards, every replica
// works on its shard from its own host thread and stream, and batch verification has its one exchange step here -- an
// ncclAllGather of the 336-byte records over NVLink -- followed by ONE host pairing per process.  Two ways to form a group:
//   * kb200_group_create_json: one process drives all the GPUs it names (ncclCommInitAll);
//   * kb200_group_create_rank: one process per GPU (torchrun, MPI, a Rust service per device): rank 0 calls
//     kb200_group_unique_id, ships the 128 bytes to the others by any means, every rank joins with ncclCommInitRank.
// NCCL is bound at run time (dlopen of libnccl.so.2): the single-GPU library has no NCCL dependency, and inside a process that
// already carries an NCCL (PyTorch's) the same copy is used.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <stdio.h>
#include <string.h>
#include <stdarg.h>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/kateth_b200.h"

namespace {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;      // ncclSuccess == 0
enum { kNcclChar = 0 };        // ncclInt8 / ncclChar

struct Nccl {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};
Nccl g_nccl;
std::once_flag g_nccl_once;

// errors are reported through kb200_last_error() like everything else in the library (engine.cu owns the per-thread buffer)
extern "C" void kb200_internal_set_error(const char* msg);
void seterr(const char* fmt, ...) {
    char buf[512];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    kb200_internal_set_error(buf);
}

void load_nccl() {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (g_nccl.lib) break; }
    if (!g_nccl.lib) return;
#define SYM(field, name) *(void**)(&g_nccl.field) = dlsym(g_nccl.lib, name)
    SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommInitAll, "ncclCommInitAll");
    SYM(CommDestroy, "ncclCommDestroy"); SYM(AllGather, "ncclAllGather"); SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd");
    SYM(GetErrorString, "ncclGetErrorString"); SYM(GetVersion, "ncclGetVersion");
#undef SYM
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommInitAll || !g_nccl.CommDestroy || !g_nccl.AllGather || !g_nccl.GroupStart || !g_nccl.GroupEnd)
        g_nccl.lib = nullptr;
}
int need_nccl() {
    std::call_once(g_nccl_once, load_nccl);
    if (!g_nccl.lib) { seterr("NCCL not found (dlopen libnccl.so.2): %s", dlerror() ? dlerror() : "missing symbols"); return KB200_E_NCCL; }
    return 0;
}
#define NC(call)                                                                                                         \
    do {                                                                                                                 \
        ncclResult_t r_ = (call);                                                                                        \
        if (r_ != 0) {                                                                                                   \
            seterr("%s:%d %s: %s", __FILE__, __LINE__, #call, g_nccl.GetErrorString ? g_nccl.GetErrorString(r_) : "nccl error"); \
            return KB200_E_NCCL;                                                                                         \
        }                                                                                                                \
    } while (0)
#define CUG(call)                                                                                                \
    do {                                                                                                         \
        cudaError_t e_ = (call);                                                                                 \
        if (e_ != cudaSuccess) { seterr("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); return KB200_E_CUDA; } \
    } while (0)

}  // namespace

struct kb200_group {
    std::vector<kb200_setup*> reps;      // replicas held by this process, one per GPU
    std::vector<int> devs;
    std::vector<ncclComm_t> comms;       // one communicator rank per replica
    std::vector<cudaStream_t> streams;   // the collective's stream on each device
    std::vector<void*> d_rec, d_all;     // per replica: its own record, the gathered records of all ranks
    int world = 1, rank0 = 0;            // ranks of the local replicas: rank0 .. rank0 + reps.size() - 1
    bool owns = false;                   // the replicas were loaded by the group (create_json) and are freed with it
    std::mutex mu;                       // one group call at a time: every rank must issue its collectives in the same order
};

static int group_buffers(kb200_group* g) {
    const size_t nl = g->reps.size();
    g->streams.assign(nl, nullptr); g->d_rec.assign(nl, nullptr); g->d_all.assign(nl, nullptr);
    for (size_t i = 0; i < nl; i++) {
        CUG(cudaSetDevice(g->devs[i]));
        CUG(cudaStreamCreateWithFlags(&g->streams[i], cudaStreamNonBlocking));
        CUG(cudaMalloc(&g->d_rec[i], KB200_RECORD_BYTES));
        CUG(cudaMalloc(&g->d_all[i], (size_t)g->world * KB200_RECORD_BYTES));
    }
    return 0;
}

extern "C" void kb200_group_free(kb200_group* g) {
    if (!g) return;
    for (size_t i = 0; i < g->reps.size(); i++) {
        cudaSetDevice(g->devs[i]);
        if (i < g->comms.size() && g->comms[i] && g_nccl.CommDestroy) g_nccl.CommDestroy(g->comms[i]);
        if (i < g->streams.size() && g->streams[i]) cudaStreamDestroy(g->streams[i]);
        if (i < g->d_rec.size() && g->d_rec[i]) cudaFree(g->d_rec[i]);
        if (i < g->d_all.size() && g->d_all[i]) cudaFree(g->d_all[i]);
        if (g->owns && g->reps[i]) kb200_setup_free(g->reps[i]);
    }
    delete g;
}

// ---- one process, several GPUs: Setup::load_json once per device (tables are built concurrently), ncclCommInitAll ----
extern "C" int kb200_group_create_json(const char* path, const int* devices, int ndev, int window_bits, kb200_group** out) {
    if (!path || !out || ndev < 1 || ndev > 64) return KB200_E_ARG;
    int have = 0;
    if (cudaGetDeviceCount(&have) != cudaSuccess || have == 0) { seterr("no CUDA device: this engine has no CPU fallback"); return KB200_E_CUDA; }
    kb200_group* g = new kb200_group();
    g->owns = true; g->world = ndev; g->rank0 = 0;
    g->reps.assign(ndev, nullptr); g->devs.resize(ndev);
    for (int i = 0; i < ndev; i++) g->devs[i] = devices ? devices[i] : i;
    std::vector<int> rc(ndev, 0);
    std::vector<std::string> msg(ndev);
    {
        std::vector<std::thread> th;
        for (int i = 0; i < ndev; i++)
            th.emplace_back([&, i] { rc[i] = kb200_setup_load_json(path, g->devs[i], window_bits, &g->reps[i]); if (rc[i]) msg[i] = kb200_last_error(); });
        for (auto& t : th) t.join();
    }
    for (int i = 0; i < ndev; i++) if (rc[i]) { seterr("device %d: %s", g->devs[i], msg[i].c_str()); int r = rc[i]; kb200_group_free(g); return r; }
    if (int r = group_buffers(g)) { kb200_group_free(g); return r; }
    g->comms.assign(ndev, nullptr);
    if (ndev > 1) {
        if (int r = need_nccl()) { kb200_group_free(g); return r; }
        ncclResult_t nr = g_nccl.CommInitAll(g->comms.data(), ndev, g->devs.data());
        if (nr != 0) { seterr("ncclCommInitAll: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(nr) : "error"); kb200_group_free(g); return KB200_E_NCCL; }
    }
    *out = g; return 0;
}

// ---- one process per GPU ----
extern "C" int kb200_group_unique_id(uint8_t id128[KB200_GROUP_ID_BYTES]) {
    if (!id128) return KB200_E_ARG;
    if (int r = need_nccl()) return r;
    ncclUniqueId id;
    NC(g_nccl.GetUniqueId(&id));
    memcpy(id128, id.internal, 128);
    return 0;
}
extern "C" int kb200_group_create_rank(kb200_setup* local, int rank, int world, const uint8_t id128[KB200_GROUP_ID_BYTES], kb200_group** out) {
    if (!local || !out || world < 1 || rank < 0 || rank >= world || (world > 1 && !id128)) return KB200_E_ARG;
    kb200_info info;
    if (int r = kb200_get_info(local, &info)) return r;
    kb200_group* g = new kb200_group();
    g->owns = false; g->world = world; g->rank0 = rank;
    g->reps.assign(1, local); g->devs.assign(1, info.device); g->comms.assign(1, nullptr);
    if (int r = group_buffers(g)) { kb200_group_free(g); return r; }
    if (world > 1) {
        if (int r = need_nccl()) { kb200_group_free(g); return r; }
        ncclUniqueId id; memcpy(id.internal, id128, 128);
        cudaSetDevice(info.device);
        ncclResult_t nr = g_nccl.CommInitRank(&g->comms[0], world, id, rank);
        if (nr != 0) { seterr("ncclCommInitRank: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(nr) : "error"); kb200_group_free(g); return KB200_E_NCCL; }
    }
    *out = g; return 0;
}
extern "C" int kb200_group_info(kb200_group* g, int* local_replicas, int* world, int* first_rank, int* nccl_version) {
    if (!g) return KB200_E_ARG;
    if (local_replicas) *local_replicas = (int)g->reps.size();
    if (world) *world = g->world;
    if (first_rank) *first_rank = g->rank0;
    if (nccl_version) { *nccl_version = 0; if (g->world > 1 && g_nccl.GetVersion) g_nccl.GetVersion(nccl_version); }
    return 0;
}
extern "C" kb200_setup* kb200_group_replica(kb200_group* g, int i) { return (g && i >= 0 && (size_t)i < g->reps.size()) ? g->reps[i] : nullptr; }

// contiguous split of n items over the local replicas
static void split(size_t n, size_t parts, std::vector<size_t>& lo) {
    lo.resize(parts + 1);
    for (size_t i = 0; i <= parts; i++) lo[i] = n * i / parts;
}
// runs fn(i) for every local replica on its own host thread (in place when there is one); returns kateth's first error in item
// order: negative (infrastructure) codes win, then the lowest-index replica's positive status
template <class F>
static int for_replicas(kb200_group* g, F fn) {
    const size_t nl = g->reps.size();
    std::vector<int> rc(nl, 0);
    std::vector<std::string> msg(nl);
    if (nl == 1) { rc[0] = fn(0); if (rc[0] < 0) msg[0] = kb200_last_error(); }
    else {
        std::vector<std::thread> th;
        for (size_t i = 0; i < nl; i++) th.emplace_back([&, i] { rc[i] = fn(i); if (rc[i] < 0) msg[i] = kb200_last_error(); });
        for (auto& t : th) t.join();
    }
    for (size_t i = 0; i < nl; i++) if (rc[i] < 0) { seterr("device %d: %s", g->devs[i], msg[i].c_str()); return rc[i]; }
    for (size_t i = 0; i < nl; i++) if (rc[i]) return rc[i];
    return 0;
}

// ---- Setup::blob_to_commitment / blob_proof / proof over the group: shards are independent, no communication ----
extern "C" int kb200_group_blob_to_commitment(kb200_group* g, const uint8_t* blobs, size_t n, uint8_t* out48, int32_t* status) {
    if (!g || (n && (!blobs || !out48))) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    std::vector<size_t> lo; split(n, g->reps.size(), lo);
    return for_replicas(g, [&](size_t i) {
        const size_t a = lo[i], m = lo[i + 1] - a;
        return m ? kb200_blob_to_commitment(g->reps[i], blobs + a * KB200_BYTES_PER_BLOB, m, out48 + a * 48, status ? status + a : nullptr) : 0;
    });
}
extern "C" int kb200_group_compute_blob_proof(kb200_group* g, const uint8_t* blobs, const uint8_t* commitments48, size_t n, uint8_t* out48, int32_t* status) {
    if (!g || (n && (!blobs || !commitments48 || !out48))) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    std::vector<size_t> lo; split(n, g->reps.size(), lo);
    return for_replicas(g, [&](size_t i) {
        const size_t a = lo[i], m = lo[i + 1] - a;
        return m ? kb200_compute_blob_proof(g->reps[i], blobs + a * KB200_BYTES_PER_BLOB, commitments48 + a * 48, m, out48 + a * 48, status ? status + a : nullptr) : 0;
    });
}
extern "C" int kb200_group_compute_proof(kb200_group* g, const uint8_t* blobs, const uint8_t* z32, size_t n, uint8_t* out_proof48, uint8_t* out_y32, int32_t* status) {
    if (!g || (n && (!blobs || !z32 || !out_proof48 || !out_y32))) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    std::vector<size_t> lo; split(n, g->reps.size(), lo);
    return for_replicas(g, [&](size_t i) {
        const size_t a = lo[i], m = lo[i + 1] - a;
        return m ? kb200_compute_proof(g->reps[i], blobs + a * KB200_BYTES_PER_BLOB, z32 + a * 32, m, out_proof48 + a * 48, out_y32 + a * 32, status ? status + a : nullptr) : 0;
    });
}

// the exchange step: every replica's record -> every replica (ONE all-gather of KB200_RECORD_BYTES per rank), then one finish
static int exchange_and_finish(kb200_group* g, int* ok) {
    const size_t nl = g->reps.size();
    if (g->world == 1) return kb200_verify_batch_finish_device(g->reps[0], g->d_rec[0], 1, ok);
    NC(g_nccl.GroupStart());
    for (size_t i = 0; i < nl; i++) {
        ncclResult_t r = g_nccl.AllGather(g->d_rec[i], g->d_all[i], KB200_RECORD_BYTES, kNcclChar, g->comms[i], g->streams[i]);
        if (r != 0) { g_nccl.GroupEnd(); seterr("ncclAllGather: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error"); return KB200_E_NCCL; }
    }
    NC(g_nccl.GroupEnd());
    // only the replica that finishes has to wait for its copy; the other streams are drained so the buffers can be reused
    for (size_t i = 0; i < nl; i++) { CUG(cudaSetDevice(g->devs[i])); CUG(cudaStreamSynchronize(g->streams[i])); }
    return kb200_verify_batch_finish_device(g->reps[0], g->d_all[0], (size_t)g->world, ok);
}

// ---- Setup::verify_blob_proof_batch over the group (src/kzg/setup.rs:247-275): one call, one verdict ----
// This process contributes items [index_offset, index_offset + n_local) of a batch of n_total (one-process groups: the whole batch).
extern "C" int kb200_group_verify_blob_proof_batch(kb200_group* g, const uint8_t* blobs, const uint8_t* commitments48, const uint8_t* proofs48, size_t n_local,
                                                   size_t index_offset, size_t n_total, int* ok, int32_t* status) {
    if (!g || !ok || (n_local && (!blobs || !commitments48 || !proofs48)) || index_offset + n_local > n_total) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    if (n_total == 0) { *ok = 1; return 0; }
    std::vector<size_t> lo; split(n_local, g->reps.size(), lo);
    int rc = for_replicas(g, [&](size_t i) {
        const size_t a = lo[i], m = lo[i + 1] - a;
        int r = kb200_verify_blob_proof_batch_partial(g->reps[i], blobs + a * KB200_BYTES_PER_BLOB, commitments48 + a * 48, proofs48 + a * 48, m, index_offset + a, n_total,
                                                      g->d_rec[i], status ? status + a : nullptr);
        return r < 0 ? r : 0;        // item errors travel inside the record: every rank must still take part in the exchange
    });
    if (rc < 0) return rc;
    return exchange_and_finish(g, ok);
}
// Setup::verify_proof_batch over the group: (proof, commitment, z, y) tuples, this process holds [index_offset, index_offset + n_local)
extern "C" int kb200_group_verify_proof_batch(kb200_group* g, const uint8_t* proofs48, const uint8_t* commitments48, const uint8_t* z32, const uint8_t* y32,
                                              size_t n_local, size_t index_offset, size_t n_total, int* ok, int32_t* status) {
    if (!g || !ok || (n_local && (!proofs48 || !commitments48 || !z32 || !y32)) || index_offset + n_local > n_total) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    if (n_total == 0) { *ok = 1; return 0; }
    std::vector<size_t> lo; split(n_local, g->reps.size(), lo);
    int rc = for_replicas(g, [&](size_t i) {
        const size_t a = lo[i], m = lo[i + 1] - a;
        int r = kb200_verify_proof_batch_partial(g->reps[i], proofs48 + a * 48, commitments48 + a * 48, z32 + a * 32, y32 + a * 32, m, index_offset + a, n_total,
                                                 g->d_rec[i], status ? status + a : nullptr);
        return r < 0 ? r : 0;
    });
    if (rc < 0) return rc;
    return exchange_and_finish(g, ok);
}
// n independent verify_proof calls, per-item verdicts from the device pairing, cut over the local replicas (no communication)
extern "C" int kb200_group_verify_proof_many(kb200_group* g, const uint8_t* proofs48, const uint8_t* commitments48, const uint8_t* z32, const uint8_t* y32, size_t n,
                                             int32_t* ok, int32_t* status) {
    if (!g || (n && (!proofs48 || !commitments48 || !z32 || !y32 || !ok))) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    std::vector<size_t> lo; split(n, g->reps.size(), lo);
    return for_replicas(g, [&](size_t i) {
        const size_t a = lo[i], m = lo[i + 1] - a;
        return m ? kb200_verify_proof_many(g->reps[i], proofs48 + a * 48, commitments48 + a * 48, z32 + a * 32, y32 + a * 32, m, ok + a, status ? status + a : nullptr) : 0;
    });
}
// The same for shards that already sit in HBM: replica i's pointers at index i, its items follow those of replica i - 1.
extern "C" int kb200_group_verify_blob_proof_batch_device(kb200_group* g, const void* const* d_blobs, const void* const* d_commitments48, const void* const* d_proofs48,
                                                          const size_t* n_per_replica, size_t index_offset, size_t n_total, int* ok, int32_t* const* d_status) {
    if (!g || !ok || !d_blobs || !d_commitments48 || !d_proofs48 || !n_per_replica || !d_status) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    if (n_total == 0) { *ok = 1; return 0; }
    const size_t nl = g->reps.size();
    std::vector<size_t> off(nl + 1, index_offset);
    for (size_t i = 0; i < nl; i++) off[i + 1] = off[i] + n_per_replica[i];
    if (off[nl] > n_total) return KB200_E_ARG;
    int rc = for_replicas(g, [&](size_t i) {
        int r = kb200_verify_blob_proof_batch_partial_device(g->reps[i], d_blobs[i], d_commitments48[i], d_proofs48[i], n_per_replica[i], off[i], n_total, g->d_rec[i],
                                                             d_status[i]);
        return r < 0 ? r : 0;
    });
    if (rc < 0) return rc;
    return exchange_and_finish(g, ok);
}
// Setup::verify_proof_batch over the group with the tuples already in HBM (the outputs of compute_proof_device): entry i of each array
// belongs to local replica i.
extern "C" int kb200_group_verify_proof_batch_device(kb200_group* g, const void* const* d_proofs48, const void* const* d_commitments48, const void* const* d_z32,
                                                     const void* const* d_y32, const size_t* n_per_replica, size_t index_offset, size_t n_total, int* ok,
                                                     int32_t* const* d_status) {
    if (!g || !ok || !d_proofs48 || !d_commitments48 || !d_z32 || !d_y32 || !n_per_replica || !d_status) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    if (n_total == 0) { *ok = 1; return 0; }
    const size_t nl = g->reps.size();
    std::vector<size_t> off(nl + 1, index_offset);
    for (size_t i = 0; i < nl; i++) off[i + 1] = off[i] + n_per_replica[i];
    if (off[nl] > n_total) return KB200_E_ARG;
    int rc = for_replicas(g, [&](size_t i) {
        int r = kb200_verify_proof_batch_partial_device(g->reps[i], d_proofs48[i], d_commitments48[i], d_z32[i], d_y32[i], n_per_replica[i], off[i], n_total, g->d_rec[i],
                                                        d_status[i]);
        return r < 0 ? r : 0;
    });
    if (rc < 0) return rc;
    return exchange_and_finish(g, ok);
}
// commit + blob proof + batch verification of a fresh batch over ONE upload per shard (kb200_commit_prove_verify_partial), the verdict
// over the whole group: what a producer of blobs runs before publishing them.
extern "C" int kb200_group_commit_prove_verify(kb200_group* g, const uint8_t* blobs, size_t n_local, size_t index_offset, size_t n_total, uint8_t* out_comm48,
                                               uint8_t* out_proof48, int* ok, int32_t* status) {
    if (!g || !ok || (n_local && (!blobs || !out_comm48 || !out_proof48)) || index_offset + n_local > n_total) return KB200_E_ARG;
    std::lock_guard<std::mutex> lk(g->mu);
    if (n_total == 0) { *ok = 1; return 0; }
    std::vector<size_t> lo; split(n_local, g->reps.size(), lo);
    int rc = for_replicas(g, [&](size_t i) {
        const size_t a = lo[i], m = lo[i + 1] - a;
        int r = kb200_commit_prove_verify_partial(g->reps[i], blobs + a * KB200_BYTES_PER_BLOB, m, index_offset + a, n_total, out_comm48 + a * 48, out_proof48 + a * 48,
                                                  g->d_rec[i], status ? status + a : nullptr);
        return r < 0 ? r : 0;
    });
    if (rc < 0) return rc;
    return exchange_and_finish(g, ok);
}
