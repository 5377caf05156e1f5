// orbfe_sharded.cu — map-scale Hamming kNN (BASELINE configs[4]: 1M x 1M, k = 2) partitioned over the GPUs of one box
// (SURVEY.md §8e): query rows are sharded, the train set is broadcast with NCCL over NVLink, every rank matches its
// shard against the FULL train set (global train indices, so OpenCV's tie order is preserved and no index fix-up is
// needed) and the per-rank results are exchanged with NCCL.  One process per GPU; the communicator is created from a
// unique id the host application distributes (bench.py: torch.distributed).
//
// NCCL is bound at run time (dlopen of libnccl.so.2, or $ORBFE_NCCL_LIB): liborbfe.so has no link-time dependency on
// it, single-GPU users never load it, and inside a PyTorch process the copy PyTorch already loaded is re-used.
#include <dlfcn.h>

#include <cstdio>
#include <cstring>
#include <mutex>

#include "orbfe_internal.h"

namespace {

typedef struct ncclComm* ncclComm_t;
struct UniqueId { char internal[128]; };    // ncclUniqueId (NCCL_UNIQUE_ID_BYTES = 128)
enum { kNcclUint8 = 1, kNcclInt32 = 2 };    // ncclDataType_t values used here

struct NcclApi {
    void* lib = nullptr;
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, UniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    char err[256] = "";
};

NcclApi g_nccl;
std::mutex g_ncclMu;

bool nccl_load() {
    std::lock_guard<std::mutex> lk(g_ncclMu);
    if (g_nccl.lib) return true;
    const char* names[] = {getenv("ORBFE_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names) {
        if (!n || !*n) continue;
        h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) {
        snprintf(g_nccl.err, sizeof(g_nccl.err), "cannot load NCCL (libnccl.so.2; set ORBFE_NCCL_LIB): %s", dlerror());
        return false;
    }
    NcclApi a;
    a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId");
    a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
    a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy");
    a.Broadcast = (decltype(a.Broadcast))dlsym(h, "ncclBroadcast");
    a.GroupStart = (decltype(a.GroupStart))dlsym(h, "ncclGroupStart");
    a.GroupEnd = (decltype(a.GroupEnd))dlsym(h, "ncclGroupEnd");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.Broadcast || !a.GroupStart || !a.GroupEnd) {
        snprintf(g_nccl.err, sizeof(g_nccl.err), "NCCL library lacks a required symbol");
        dlclose(h);
        return false;
    }
    a.lib = h;
    g_nccl = a;
    return true;
}

}    // namespace

namespace orbfe {
// defined in orbfe_api.cu: one kNN launch on the matcher's stream with device buffers; records the kernel time
int matcher_knn_device(orbfe_matcher* m, const uint8_t* q, long long nq, const uint8_t* t, long long nt, int32_t* idx,
                       int32_t* dist, cudaStream_t* streamOut);
int matcher_fail(orbfe_matcher* m, int code, const char* msg);
}    // namespace orbfe

extern "C" {

int orbfe_shard_range(int64_t n, int world, int rank, int64_t* begin, int64_t* count) {
    if (n < 0 || world <= 0 || rank < 0 || rank >= world || !begin || !count) return ORBFE_ERR_ARG;
    const int64_t b = n * rank / world, e = n * (rank + 1) / world;    // [g Nq / G, (g + 1) Nq / G)
    *begin = b;
    *count = e - b;
    return ORBFE_OK;
}

const char* orbfe_nccl_last_error(void) { return g_nccl.err; }

int orbfe_nccl_unique_id(char out[128]) {
    if (!out) return ORBFE_ERR_ARG;
    if (!nccl_load()) return ORBFE_ERR_NCCL;
    UniqueId id;
    const int rc = g_nccl.GetUniqueId(&id);
    if (rc != 0) {
        snprintf(g_nccl.err, sizeof(g_nccl.err), "ncclGetUniqueId: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
        return ORBFE_ERR_NCCL;
    }
    memcpy(out, id.internal, 128);
    return ORBFE_OK;
}

int orbfe_nccl_comm_create(const char id[128], int rank, int world, int device, void** comm) {
    if (!id || !comm || world <= 0 || rank < 0 || rank >= world) return ORBFE_ERR_ARG;
    *comm = nullptr;
    if (!nccl_load()) return ORBFE_ERR_NCCL;
    if (cudaSetDevice(device) != cudaSuccess) return ORBFE_ERR_CUDA;
    UniqueId u;
    memcpy(u.internal, id, 128);
    ncclComm_t c = nullptr;
    const int rc = g_nccl.CommInitRank(&c, world, u, rank);
    if (rc != 0) {
        snprintf(g_nccl.err, sizeof(g_nccl.err), "ncclCommInitRank: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "?");
        return ORBFE_ERR_NCCL;
    }
    *comm = c;
    return ORBFE_OK;
}

int orbfe_nccl_comm_destroy(void* comm) {
    if (!comm) return ORBFE_ERR_ARG;
    if (!nccl_load()) return ORBFE_ERR_NCCL;
    return g_nccl.CommDestroy((ncclComm_t)comm) == 0 ? ORBFE_OK : ORBFE_ERR_NCCL;
}

int orbfe_knn2_sharded(orbfe_matcher* m, void* comm, int rank, int world, const uint8_t* q_shard, int64_t nq_total,
                       uint8_t* t, int64_t nt, int root, int32_t* idx_all, int32_t* dist_all, float* ms_out) {
    if (!m) return ORBFE_ERR_ARG;
    if (world <= 0 || rank < 0 || rank >= world || root < 0 || root >= world || nq_total < 0 || nt < 0 ||
        (nq_total > 0 && (!idx_all || !dist_all)) || (nt > 0 && !t))
        return orbfe::matcher_fail(m, ORBFE_ERR_ARG, "orbfe_knn2_sharded: bad arguments");
    if (world > 1 && (!comm || !nccl_load())) return orbfe::matcher_fail(m, ORBFE_ERR_NCCL, world > 1 && comm ? g_nccl.err : "no communicator");
    int64_t b0 = 0, cnt = 0;
    orbfe_shard_range(nq_total, world, rank, &b0, &cnt);
    if (cnt > 0 && !q_shard) return orbfe::matcher_fail(m, ORBFE_ERR_ARG, "orbfe_knn2_sharded: null query shard");
    cudaStream_t st = nullptr;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    // the kNN launcher hands back the matcher's stream; everything below is ordered on it
    int rc = orbfe::matcher_knn_device(m, nullptr, 0, nullptr, 0, nullptr, nullptr, &st);
    if (rc != ORBFE_OK) return rc;
    cudaEventRecord(e0, st);
    ncclComm_t c = (ncclComm_t)comm;
    char msg[256];
    if (world > 1 && nt > 0) {    // train set: root -> everyone (32 MB at C5; NVLink 5 / NVSwitch)
        const int nrc = g_nccl.Broadcast(t, t, (size_t)nt * 32, kNcclUint8, root, c, st);
        if (nrc != 0) {
            snprintf(msg, sizeof(msg), "ncclBroadcast(train): %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(nrc) : "?");
            return orbfe::matcher_fail(m, ORBFE_ERR_NCCL, msg);
        }
    }
    if (cnt > 0) {
        rc = orbfe::matcher_knn_device(m, q_shard, cnt, t, nt, idx_all + 2 * b0, dist_all + 2 * b0, nullptr);
        if (rc != ORBFE_OK) return rc;
    }
    if (world > 1 && nq_total > 0) {    // results: every shard to every rank (16 B per query row)
        g_nccl.GroupStart();
        int nrc = 0;
        for (int r = 0; r < world && nrc == 0; r++) {
            int64_t rb = 0, rn = 0;
            orbfe_shard_range(nq_total, world, r, &rb, &rn);
            if (rn == 0) continue;
            nrc = g_nccl.Broadcast(idx_all + 2 * rb, idx_all + 2 * rb, (size_t)rn * 2, kNcclInt32, r, c, st);
            if (nrc == 0) nrc = g_nccl.Broadcast(dist_all + 2 * rb, dist_all + 2 * rb, (size_t)rn * 2, kNcclInt32, r, c, st);
        }
        const int grc = g_nccl.GroupEnd();
        if (nrc != 0 || grc != 0) {
            snprintf(msg, sizeof(msg), "ncclBroadcast(results): %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(nrc ? nrc : grc) : "?");
            return orbfe::matcher_fail(m, ORBFE_ERR_NCCL, msg);
        }
    }
    cudaEventRecord(e1, st);
    if (cudaStreamSynchronize(st) != cudaSuccess) return orbfe::matcher_fail(m, ORBFE_ERR_CUDA, cudaGetErrorString(cudaGetLastError()));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms_out) *ms_out = ms;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return ORBFE_OK;
}

}    // extern "C"
