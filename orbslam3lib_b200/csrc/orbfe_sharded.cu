// orbfe_sharded.cu — map-scale Hamming kNN (BASELINE configs[4]: 1M x 1M, k = 2) partitioned over the GPUs of one box
// (SURVEY.md §8e): query rows are sharded, the train set is broadcast with NCCL over NVLink, every rank matches its
// shard against the FULL train set (global train indices, so OpenCV's tie order is preserved and no index fix-up is
// needed) and the per-rank results are exchanged with NCCL.  One process per GPU; the communicator is created from a
// unique id the host application distributes (bench.py: torch.distributed).
//
// NCCL is bound at run time (dlopen of libnccl.so.2, or $ORBFE_NCCL_LIB): liborbfe.so has no link-time dependency on
// it, single-GPU users never load it, and inside a PyTorch process the copy PyTorch already loaded is re-used.
#include <dlfcn.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
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
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
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
    a.AllGather = (decltype(a.AllGather))dlsym(h, "ncclAllGather");
    a.GroupStart = (decltype(a.GroupStart))dlsym(h, "ncclGroupStart");
    a.GroupEnd = (decltype(a.GroupEnd))dlsym(h, "ncclGroupEnd");
    a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.Broadcast || !a.AllGather || !a.GroupStart || !a.GroupEnd) {
        snprintf(g_nccl.err, sizeof(g_nccl.err), "NCCL library lacks a required symbol");
        dlclose(h);
        return false;
    }
    a.lib = h;
    g_nccl = a;
    return true;
}

// per-matcher state of the sharded matcher: a second stream for the train broadcast, events, the segment scratch of
// the multi-launch kNN and the int4 staging of the result exchange.  Lives in the matcher's aux slot.
constexpr int kMaxChunks = 16;
struct ShardCtx {
    cudaStream_t comm = nullptr;
    cudaEvent_t evStart = nullptr, evEnd = nullptr, evG0 = nullptr, evChunk[kMaxChunks] = {}, evK0[kMaxChunks] = {},
                evK1[kMaxChunks] = {};
    uint8_t* scratch = nullptr;
    size_t scratchBytes = 0;
    int4* res4 = nullptr;
    size_t res4Rows = 0;
    float times[4] = {0, 0, 0, 0};    // total, kernels, broadcast (on its stream), result exchange
    ~ShardCtx() {
        if (comm) cudaStreamDestroy(comm);
        for (cudaEvent_t e : {evStart, evEnd, evG0})
            if (e) cudaEventDestroy(e);
        for (int i = 0; i < kMaxChunks; i++)
            for (cudaEvent_t e : {evChunk[i], evK0[i], evK1[i]})
                if (e) cudaEventDestroy(e);
        if (scratch) cudaFree(scratch);
        if (res4) cudaFree(res4);
    }
};
void shard_ctx_free(void* p) { delete (ShardCtx*)p; }

// (idx0, idx1, d0, d1) records: one 16-byte element per query row for the all-gather
__global__ void shard_pack_kernel(const int2* __restrict__ idx, const int2* __restrict__ dist, int4* __restrict__ out, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int2 a = idx[i], b = dist[i];
    out[i] = make_int4(a.x, a.y, b.x, b.y);
}
// staging slot s = rank r's row j (r = s / maxCnt): global row = n r / world + j
__global__ void shard_unpack_kernel(const int4* __restrict__ in, long long maxCnt, int world, long long n, int skipRank,
                                    int2* __restrict__ idx, int2* __restrict__ dist) {
    const long long s = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= maxCnt * world) return;
    const int r = (int)(s / maxCnt);
    const long long j = s - (long long)r * maxCnt;
    const long long b = n * r / world, e = n * (r + 1) / world;
    if (r == skipRank || j >= e - b) return;
    const int4 v = in[s];
    idx[b + j] = make_int2(v.x, v.y);
    dist[b + j] = make_int2(v.z, v.w);
}

}    // namespace

namespace orbfe {
// defined in orbfe_api.cu: one kNN launch on the matcher's stream with device buffers; records the kernel time
int matcher_knn_device(orbfe_matcher* m, const uint8_t* q, long long nq, const uint8_t* t, long long nt, int32_t* idx,
                       int32_t* dist, cudaStream_t* streamOut);
int matcher_fail(orbfe_matcher* m, int code, const char* msg);
void** matcher_aux(orbfe_matcher* m, void (***freeFn)(void*));
void matcher_count_launches(orbfe_matcher* m, int launches);
// k_knn.cu / k_knn_tc.cu
bool knn_tc_enabled();
void knn2_tc_plan(long long nq, long long nt, int chunks, int* segsPerChunk, long long* segLen, int* zTotal,
                  size_t* scratchBytes);
int launch_knn2_tc_segments(const uint8_t* q, long long nq, long long planRows, const uint8_t* t, long long nt, int32_t* idx,
                            int32_t* dist, long long segLen, int zBase, int zCount, int zTotal, uint8_t* scratch,
                            cudaStream_t st);
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

#define SHARD_CUDA(call)                                                                              \
    do {                                                                                              \
        cudaError_t e__ = (call);                                                                     \
        if (e__ != cudaSuccess) return orbfe::matcher_fail(m, ORBFE_ERR_CUDA, cudaGetErrorString(e__)); \
    } while (0)
#define SHARD_NCCL(call, what)                                                                                       \
    do {                                                                                                             \
        const int n__ = (call);                                                                                      \
        if (n__ != 0) {                                                                                              \
            char msg__[256];                                                                                         \
            snprintf(msg__, sizeof(msg__), "%s: %s", what, g_nccl.GetErrorString ? g_nccl.GetErrorString(n__) : "?"); \
            return orbfe::matcher_fail(m, ORBFE_ERR_NCCL, msg__);                                                    \
        }                                                                                                            \
    } while (0)

// The train set travels on a second stream, optionally in ORBFE_SHARD_CHUNKS pieces; the kNN segments of piece c start
// as soon as it has landed (the root does not wait at all).  One piece is the measured optimum (see below).  Results:
// every rank packs its rows into (idx0, idx1, d0, d1) records and ONE ncclAllGather moves them.
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
    // the kNN launcher hands back the matcher's stream (and selects its device); everything below is ordered on it
    int rc = orbfe::matcher_knn_device(m, nullptr, 0, nullptr, 0, nullptr, nullptr, &st);
    if (rc != ORBFE_OK) return rc;
    void (**freeFn)(void*) = nullptr;
    void** slot = orbfe::matcher_aux(m, &freeFn);
    if (!*slot) {
        ShardCtx* c = new ShardCtx();
        *slot = c;
        *freeFn = shard_ctx_free;
        SHARD_CUDA(cudaStreamCreateWithFlags(&c->comm, cudaStreamNonBlocking));
        for (cudaEvent_t* e : {&c->evStart, &c->evEnd, &c->evG0}) SHARD_CUDA(cudaEventCreate(e));
        for (int i = 0; i < kMaxChunks; i++)
            for (cudaEvent_t* e : {&c->evChunk[i], &c->evK0[i], &c->evK1[i]}) SHARD_CUDA(cudaEventCreate(e));
    }
    ShardCtx* cx = (ShardCtx*)*slot;
    ncclComm_t c = (ncclComm_t)comm;
    const long long maxCnt = (nq_total + world - 1) / world;
    // Pieces of the train broadcast.  Default 1: measured on 2 B200s (1M x 1M), the whole 32 MB broadcast takes 0.08 ms
    // over NVLink and the result all-gather 0.06 ms of a 90.3 ms step, while broadcasting in pieces UNDER the kNN kernel
    // starves NCCL's copy kernel of SMs (23 - 45 ms per broadcast): 94.8 ms with 4 pieces, 102 ms with 2.
    static const int chunksEnv = getenv("ORBFE_SHARD_CHUNKS") ? atoi(getenv("ORBFE_SHARD_CHUNKS")) : 1;
    // segments need the tcgen05 kernel and a problem large enough for it (k_knn.cu: 296 CTAs); otherwise one launch
    // (ORBFE_SHARD_FORCE_SEG=1: the multi-launch path also at world == 1 and on small problems, for single-GPU tests)
    static const bool forceSeg = getenv("ORBFE_SHARD_FORCE_SEG") && atoi(getenv("ORBFE_SHARD_FORCE_SEG")) != 0;
    const bool segmented = orbfe::knn_tc_enabled() && nt >= 4 * 4096 &&
                           (forceSeg ? maxCnt >= 1 : (world > 1 && maxCnt >= 296 * 128));
    int chunks = segmented ? std::min(std::max(chunksEnv, 1), kMaxChunks) : 1, spc = 1, zTotal = 1;
    long long segLen = nt;
    size_t scratchBytes = 0;
    if (segmented) {
        orbfe::knn2_tc_plan(maxCnt, nt, chunks, &spc, &segLen, &zTotal, &scratchBytes);    // same plan on every rank
        chunks = (zTotal + spc - 1) / spc;
        if (cx->scratchBytes < scratchBytes) {
            if (cx->scratch) cudaFree(cx->scratch);
            cx->scratch = nullptr;
            cx->scratchBytes = 0;
            SHARD_CUDA(cudaMalloc(&cx->scratch, scratchBytes));
            cx->scratchBytes = scratchBytes;
        }
        // tickets of this plan: zero before the first launch (the kernel leaves them zero)
        const long long tiles = (maxCnt + 127) / 128;
        SHARD_CUDA(cudaMemsetAsync(cx->scratch + (size_t)zTotal * tiles * 128 * sizeof(int4), 0, (size_t)tiles * 4, st));
    }
    if (world > 1 && cx->res4Rows < (size_t)maxCnt * world) {
        if (cx->res4) cudaFree(cx->res4);
        cx->res4 = nullptr;
        cx->res4Rows = 0;
        SHARD_CUDA(cudaMalloc(&cx->res4, (size_t)maxCnt * world * sizeof(int4)));
        cx->res4Rows = (size_t)maxCnt * world;
    }
    SHARD_CUDA(cudaEventRecord(cx->evStart, st));
    int launches = 0;
    if (world > 1 && nt > 0) {    // train set: root -> everyone, piece by piece on the second stream (NVLink 5 / NVSwitch)
        SHARD_CUDA(cudaStreamWaitEvent(cx->comm, cx->evStart, 0));
        for (int k = 0; k < chunks; k++) {
            const long long r0 = std::min<long long>(nt, (long long)k * spc * segLen);
            const long long r1 = chunks == 1 ? nt : std::min<long long>(nt, (long long)(k + 1) * spc * segLen);
            if (r1 > r0) SHARD_NCCL(g_nccl.Broadcast(t + r0 * 32, t + r0 * 32, (size_t)(r1 - r0) * 32, kNcclUint8, root, c, cx->comm), "ncclBroadcast(train)");
            SHARD_CUDA(cudaEventRecord(cx->evChunk[k], cx->comm));
        }
    }
    for (int k = 0; k < chunks; k++) {
        if (world > 1 && nt > 0 && rank != root) SHARD_CUDA(cudaStreamWaitEvent(st, cx->evChunk[k], 0));
        SHARD_CUDA(cudaEventRecord(cx->evK0[k], st));
        if (cnt > 0) {
            if (segmented) {
                const int z0 = k * spc, zc = std::min(spc, zTotal - z0);
                // segments of a plan made for maxCnt rows; this rank may own one tile less: the scratch layout follows
                // the launch's own tile count, tickets are per tile, so that is consistent within the rank
                launches += orbfe::launch_knn2_tc_segments(q_shard, cnt, maxCnt, t, nt, idx_all + 2 * b0, dist_all + 2 * b0, segLen,
                                                           z0, zc, zTotal, cx->scratch, st);
            } else {
                rc = orbfe::matcher_knn_device(m, q_shard, cnt, t, nt, idx_all + 2 * b0, dist_all + 2 * b0, nullptr);
                if (rc != ORBFE_OK) return rc;
                launches += 1;
            }
        }
        SHARD_CUDA(cudaEventRecord(cx->evK1[k], st));
    }
    SHARD_CUDA(cudaGetLastError());
    SHARD_CUDA(cudaEventRecord(cx->evG0, st));
    if (world > 1 && nq_total > 0) {    // results: one 16-byte record per query row, one all-gather
        if (cnt > 0) {
            shard_pack_kernel<<<(unsigned)((cnt + 255) / 256), 256, 0, st>>>(reinterpret_cast<const int2*>(idx_all + 2 * b0),
                                                                             reinterpret_cast<const int2*>(dist_all + 2 * b0),
                                                                             cx->res4 + (size_t)rank * maxCnt, cnt);
            launches++;
        }
        SHARD_NCCL(g_nccl.AllGather(cx->res4 + (size_t)rank * maxCnt, cx->res4, (size_t)maxCnt * 4, kNcclInt32, c, st), "ncclAllGather(results)");
        const long long slots = maxCnt * world;
        shard_unpack_kernel<<<(unsigned)((slots + 255) / 256), 256, 0, st>>>(cx->res4, maxCnt, world, nq_total, rank,
                                                                             reinterpret_cast<int2*>(idx_all),
                                                                             reinterpret_cast<int2*>(dist_all));
        launches++;
    }
    SHARD_CUDA(cudaEventRecord(cx->evEnd, st));
    SHARD_CUDA(cudaStreamSynchronize(st));
    SHARD_CUDA(cudaStreamSynchronize(cx->comm));
    orbfe::matcher_count_launches(m, launches);
    float total = 0, kern = 0, bc = 0, ga = 0;
    cudaEventElapsedTime(&total, cx->evStart, cx->evEnd);
    for (int k = 0; k < chunks; k++) {
        float v = 0;
        cudaEventElapsedTime(&v, cx->evK0[k], cx->evK1[k]);
        kern += v;
    }
    if (world > 1 && nt > 0) cudaEventElapsedTime(&bc, cx->evStart, cx->evChunk[chunks - 1]);
    cudaEventElapsedTime(&ga, cx->evG0, cx->evEnd);
    cx->times[0] = total;
    cx->times[1] = kern;
    cx->times[2] = bc;
    cx->times[3] = ga;
    if (ms_out) *ms_out = total;
    return ORBFE_OK;
}

int orbfe_knn2_sharded_times(orbfe_matcher* m, float out[4]) {
    if (!m || !out) return ORBFE_ERR_ARG;
    void (**freeFn)(void*) = nullptr;
    void** slot = orbfe::matcher_aux(m, &freeFn);
    for (int i = 0; i < 4; i++) out[i] = *slot ? ((ShardCtx*)*slot)->times[i] : 0.0f;
    return ORBFE_OK;
}

}    // extern "C"
