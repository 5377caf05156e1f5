// k_knn.cu — cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, matches, k = 2)  (Frame.cc:45, 1227) on 256-bit
// ORB descriptors; stands in for dsp/src/orbslam_dsp_matcher.cpp:32-219 (knnMatchORB).
//
// One thread owns one query row (eight 32-bit words in registers); the CTA streams the train set through shared
// memory in tiles, every warp reading the same train row at a time (LDS.128 broadcast), so a train row loaded once
// serves 256 queries.  Per pair: 8 XOR + 8 POPC + adds, then a branch-free top-2 update on packed keys
// (distance << 23 | train index inside a 2^23-row super block): b2 = min(b2, max(b1, k)); b1 = min(b1, k).  Packed
// keys order by distance first and train index second, which is OpenCV's tie rule (equal distances keep the lower
// train index first).  Integer/POPC-pipe bound, not HBM-bound: the train set is read from L2.
//
// The small-problem variant (knn2_warp_kernel) gives a warp to every query and spreads the train rows over the
// lanes, finishing with a warp-shuffle top-2 merge: per-frame stereo matching (500..1200 rows) would otherwise
// occupy only a handful of CTAs.
#include <cstdlib>

#include "orbfe_internal.h"

namespace orbfe {

constexpr int KNN_THREADS = 256;
constexpr int KNN_TILE = 256;              // train rows per shared-memory tile
constexpr int KNN_SUPER_BITS = 23;
constexpr uint32_t KNN_EMPTY = 0xffffffffu;

__device__ __forceinline__ void top2_update(uint32_t& b1, uint32_t& b2, uint32_t k) {
    b2 = min(b2, max(b1, k));
    b1 = min(b1, k);
}

// merge a super block's packed result into the running (distance, index) pairs; later super blocks hold larger
// train indices, so strict "<" keeps the lower index on ties.
__device__ __forceinline__ void merge_super(int& d1, int& i1, int& d2, int& i2, uint32_t b, long long base) {
    if (b == KNN_EMPTY) return;
    const int d = (int)(b >> KNN_SUPER_BITS);
    const int i = (int)(base + (long long)(b & ((1u << KNN_SUPER_BITS) - 1u)));
    if (i1 < 0 || d < d1) { d2 = d1; i2 = i1; d1 = d; i1 = i; }
    else if (i2 < 0 || d < d2) { d2 = d; i2 = i; }
}

// grid (ceil(maxNq / KNN_THREADS), npairs)
__global__ void __launch_bounds__(KNN_THREADS) knn2_kernel(const uint8_t* __restrict__ q, const int* __restrict__ nqArr,
                                                           long long qCap, const uint8_t* __restrict__ t,
                                                           const int* __restrict__ ntArr, long long tCap,
                                                           long long nqSingle, long long ntSingle,
                                                           int32_t* __restrict__ idx, int32_t* __restrict__ dist) {
    __shared__ __align__(16) uint4 tile[2][KNN_TILE * 2];
    const int pair = blockIdx.y;
    const long long nq = nqArr ? nqArr[pair] : nqSingle;
    const long long nt = ntArr ? ntArr[pair] : ntSingle;
    if ((long long)blockIdx.x * KNN_THREADS >= nq) return;
    const long long row = (long long)blockIdx.x * KNN_THREADS + threadIdx.x;
    const bool active = row < nq;
    const uint4* qp = reinterpret_cast<const uint4*>(q + ((long long)pair * qCap + (active ? row : 0)) * 32);
    const uint4 qa = __ldg(qp), qb = __ldg(qp + 1);
    const uint4* tp = reinterpret_cast<const uint4*>(t + (long long)pair * tCap * 32);

    int d1 = -1, i1 = -1, d2 = -1, i2 = -1;
    for (long long sb = 0; sb < nt; sb += (1ll << KNN_SUPER_BITS)) {
        const int ns = (int)min(nt - sb, 1ll << KNN_SUPER_BITS);
        uint32_t b1 = KNN_EMPTY, b2 = KNN_EMPTY;
        const int ntiles = (ns + KNN_TILE - 1) / KNN_TILE;
        // prefetch tile 0
        {
            const int r = threadIdx.x;
            if (r < ns) {
                tile[0][2 * r] = __ldg(tp + (sb + r) * 2);
                tile[0][2 * r + 1] = __ldg(tp + (sb + r) * 2 + 1);
            }
        }
        __syncthreads();
        for (int ti = 0; ti < ntiles; ti++) {
            const int cur = ti & 1;
            const int j0 = ti * KNN_TILE;
            const int rows = min(KNN_TILE, ns - j0);
            // stage the next tile into the other buffer while this one is consumed
            uint4 na = make_uint4(0, 0, 0, 0), nb = na;
            const int nr = j0 + KNN_TILE + threadIdx.x;
            const bool haveNext = nr < ns;
            if (haveNext) {
                na = __ldg(tp + (sb + nr) * 2);
                nb = __ldg(tp + (sb + nr) * 2 + 1);
            }
            const uint4* tt = tile[cur];
            uint32_t key = (uint32_t)j0;
#pragma unroll 4
            for (int j = 0; j < rows; j++) {
                const uint4 ta = tt[2 * j], tb = tt[2 * j + 1];
                const uint32_t d = __popc(qa.x ^ ta.x) + __popc(qa.y ^ ta.y) + __popc(qa.z ^ ta.z) + __popc(qa.w ^ ta.w) +
                                   __popc(qb.x ^ tb.x) + __popc(qb.y ^ tb.y) + __popc(qb.z ^ tb.z) + __popc(qb.w ^ tb.w);
                top2_update(b1, b2, (d << KNN_SUPER_BITS) | key);
                key++;
            }
            if (haveNext) {
                tile[cur ^ 1][2 * threadIdx.x] = na;
                tile[cur ^ 1][2 * threadIdx.x + 1] = nb;
            }
            __syncthreads();
        }
        merge_super(d1, i1, d2, i2, b1, sb);
        merge_super(d1, i1, d2, i2, b2, sb);
    }
    if (active) {
        const long long o = ((long long)pair * qCap + row) * 2;
        *reinterpret_cast<int2*>(idx + o) = make_int2(i1, i2);
        *reinterpret_cast<int2*>(dist + o) = make_int2(d1, d2);
    }
}

// ---- tensor-core variant -------------------------------------------------------------------------------------------
// All-pairs Hamming is a binary GEMM: hamming(a, b) = |a| + sum_k b_k * (1 - 2 a_k).  knn2_kernel is bound by the POPC
// pipe (8 POPC per pair: ~0.55 Tpairs/s on a B200); here the sums run on the tensor cores as int8 IMMA (mma.sync
// m16n8k32 s8 with A' = 1 - 2a in {+1, -1} and B = b in {0, 1}: sm_100a has no native 1-bit MMA — ptxas emulates b1 with
// the same IMMA — and exact integer accumulation is wanted, so not fp8).  The accumulator is hamming - |a| directly.
// The k order of a dot product is free as long as both operands use the same one, so bits are expanded to bytes with a
// shift and a mask, never transposed: thread (g, tg) of a warp owns words 2tg, 2tg+1 of "its" row, and k-step s takes
// (word >> s) & 0x01010101.
//   * CTA = 8 warps x 32 query rows; the expanded A' fragments (2 m16 tiles x 8 k-steps x 4 registers) stay in
//     registers for the whole kernel;
//   * train rows are expanded ONCE per CTA while they are staged into shared memory (double buffered), already in
//     fragment order: [n8 tile][register quad][lane] x 16 B, so both the staging STS.128 and the four LDS.128 that
//     fetch the B fragments of an n8 tile are contiguous per warp; 16 IMMAs per n8 tile and warp;
//   * epilogue on the accumulator fragment: key = acc << S | j orders by (distance, train index) for a fixed row (|a|
//     is a per-row constant, added at the end), top-2 by min/max as in knn2_kernel, merged across the four lanes of a
//     quad by shuffles: 1 IMAD (fma pipe) + 2.5 min/max per pair instead of 8 XOR + 8 POPC + 7 ADD + 4.
constexpr int KM_THREADS = 256;
constexpr int KM_ROWS = 256;     // query rows per CTA
constexpr int KM_TILE = 128;     // train rows per shared-memory tile (32 KB expanded)
constexpr int KM_S = 22;         // index bits of a key; hamming - |a| is in [-256, 256]
constexpr int KM_EMPTY = 0x7fffffff;
constexpr int KM_SMEM = 2 * KM_TILE * 256;

__device__ __forceinline__ void top2_update_s(int& b1, int& b2, int k) {
    b2 = min(b2, max(b1, k));
    b1 = min(b1, k);
}

__device__ __forceinline__ void imma_16832(int (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void merge_super_s(int& d1, int& i1, int& d2, int& i2, int b, int na, long long base) {
    if (b == KM_EMPTY) return;
    const int d = (b >> KM_S) + na;
    const int i = (int)(base + (long long)(b & ((1 << KM_S) - 1)));
    if (i1 < 0 || d < d1) { d2 = d1; i2 = i1; d1 = d; i1 = i; }
    else if (i2 < 0 || d < d2) { d2 = d; i2 = i; }
}

// grid (ceil(maxNq / KM_ROWS), npairs); dynamic shared memory KM_SMEM
__global__ void __launch_bounds__(KM_THREADS, 2) knn2_mma_kernel(const uint8_t* __restrict__ q, const int* __restrict__ nqArr,
                                                                 long long qCap, const uint8_t* __restrict__ t,
                                                                 const int* __restrict__ ntArr, long long tCap,
                                                                 long long nqSingle, long long ntSingle,
                                                                 int32_t* __restrict__ idx, int32_t* __restrict__ dist) {
    extern __shared__ __align__(16) uint8_t kmSmem[];
    uint4* expB = reinterpret_cast<uint4*>(kmSmem);    // [2][KM_TILE / 8][4][32] uint4
    const int pair = blockIdx.y;
    const long long nq = nqArr ? nqArr[pair] : nqSingle;
    const long long nt = ntArr ? ntArr[pair] : ntSingle;
    if ((long long)blockIdx.x * KM_ROWS >= nq) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, tg = lane & 3;
    const long long rowBase = (long long)blockIdx.x * KM_ROWS + warp * 32;
    const uint2* tp = reinterpret_cast<const uint2*>(t + (long long)pair * tCap * 32);
    constexpr uint32_t M = 0x01010101u;

    // A' fragments: rows rowBase + mt * 16 + g + 8 * h, words 2tg and 2tg + 1; bytes +1 (bit clear) / -1 (bit set)
    uint32_t A[2][8][4];
    int na[4];
#pragma unroll
    for (int mt = 0; mt < 2; mt++)
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const long long row = rowBase + mt * 16 + g + 8 * h;
            uint2 w = make_uint2(0u, 0u);
            if (row < nq) w = __ldg(reinterpret_cast<const uint2*>(q + ((long long)pair * qCap + row) * 32) + tg);
            int pc = __popc(w.x) + __popc(w.y);
            pc += __shfl_xor_sync(0xffffffffu, pc, 1);
            pc += __shfl_xor_sync(0xffffffffu, pc, 2);
            na[mt * 2 + h] = pc;
#pragma unroll
            for (int s = 0; s < 8; s++) {
                A[mt][s][h] = (((w.x >> s) & M) * 0xffu) | M;
                A[mt][s][2 + h] = (((w.y >> s) & M) * 0xffu) | M;
            }
        }

    int d1[4], i1[4], d2[4], i2[4];
#pragma unroll
    for (int r = 0; r < 4; r++) d1[r] = i1[r] = d2[r] = i2[r] = -1;

    for (long long sb = 0; sb < nt; sb += (1ll << KM_S)) {
        const int ns = (int)min(nt - sb, 1ll << KM_S);
        int b1[4], b2[4];
#pragma unroll
        for (int r = 0; r < 4; r++) b1[r] = b2[r] = KM_EMPTY;
        const int ntiles = (ns + KM_TILE - 1) / KM_TILE;
        // a tile is 512 (row, tg) items of 8 bytes: thread t stages items t and t + 256; item e = row * 4 + tg is lane
        // (row & 7) * 4 + tg of n8 tile row >> 3, i.e. fragment slot e
        auto fetch = [&](int j0, uint2 (&v)[2]) {
#pragma unroll
            for (int u = 0; u < 2; u++) {
                const int e = tid + u * KM_THREADS;
                v[u] = make_uint2(0u, 0u);
                if (j0 + (e >> 2) < ns) v[u] = __ldg(tp + (sb + j0) * 4 + e);
            }
        };
        auto stage = [&](int buf, const uint2 (&v)[2]) {
#pragma unroll
            for (int u = 0; u < 2; u++) {
                const int e = tid + u * KM_THREADS;
                uint4* dst = expB + buf * (KM_TILE * 16) + (e >> 5) * 128 + (e & 31);
#pragma unroll
                for (int k = 0; k < 4; k++)
                    dst[k * 32] = make_uint4((v[u].x >> (2 * k)) & M, (v[u].y >> (2 * k)) & M, (v[u].x >> (2 * k + 1)) & M,
                                             (v[u].y >> (2 * k + 1)) & M);
            }
        };
        uint2 nx[2];
        fetch(0, nx);
        stage(0, nx);
        __syncthreads();
        for (int ti = 0; ti < ntiles; ti++) {
            const int cur = ti & 1;
            const int j0 = ti * KM_TILE;
            const bool haveNext = ti + 1 < ntiles;
            if (haveNext) fetch(j0 + KM_TILE, nx);
            const uint4* fb = expB + cur * (KM_TILE * 16) + lane;
            const int rows = min(KM_TILE, ns - j0);
            const int full = rows >> 3;    // n8 tiles without padded columns
            auto n8tile = [&](int n8, bool guard, bool filter) {
                const uint4 f0 = fb[n8 * 128], f1 = fb[n8 * 128 + 32], f2 = fb[n8 * 128 + 64], f3 = fb[n8 * 128 + 96];
                int acc[2][4] = {{0, 0, 0, 0}, {0, 0, 0, 0}};
                imma_16832(acc[0], A[0][0], f0.x, f0.y);
                imma_16832(acc[1], A[1][0], f0.x, f0.y);
                imma_16832(acc[0], A[0][1], f0.z, f0.w);
                imma_16832(acc[1], A[1][1], f0.z, f0.w);
                imma_16832(acc[0], A[0][2], f1.x, f1.y);
                imma_16832(acc[1], A[1][2], f1.x, f1.y);
                imma_16832(acc[0], A[0][3], f1.z, f1.w);
                imma_16832(acc[1], A[1][3], f1.z, f1.w);
                imma_16832(acc[0], A[0][4], f2.x, f2.y);
                imma_16832(acc[1], A[1][4], f2.x, f2.y);
                imma_16832(acc[0], A[0][5], f2.z, f2.w);
                imma_16832(acc[1], A[1][5], f2.z, f2.w);
                imma_16832(acc[0], A[0][6], f3.x, f3.y);
                imma_16832(acc[1], A[1][6], f3.x, f3.y);
                imma_16832(acc[0], A[0][7], f3.z, f3.w);
                imma_16832(acc[1], A[1][7], f3.z, f3.w);
                const int jc = j0 + n8 * 8 + 2 * tg;    // this thread's columns: jc, jc + 1
                const bool ok0 = !guard || jc < ns, ok1 = !guard || jc + 1 < ns;
                int k[2][4];
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    k[mt][0] = ok0 ? acc[mt][0] * (1 << KM_S) + jc : KM_EMPTY;
                    k[mt][1] = ok1 ? acc[mt][1] * (1 << KM_S) + (jc + 1) : KM_EMPTY;
                    k[mt][2] = ok0 ? acc[mt][2] * (1 << KM_S) + jc : KM_EMPTY;
                    k[mt][3] = ok1 ? acc[mt][3] * (1 << KM_S) + (jc + 1) : KM_EMPTY;
                }
                // long train sets: after the first few hundred columns a new candidate rarely enters a row's top 2, so
                // test before updating (the branch is almost never taken; keys are unique, so "<" loses nothing)
                if (filter && !(min(k[0][0], k[0][1]) < b2[0] || min(k[0][2], k[0][3]) < b2[1] ||
                                min(k[1][0], k[1][1]) < b2[2] || min(k[1][2], k[1][3]) < b2[3]))
                    return;
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    top2_update_s(b1[mt * 2], b2[mt * 2], k[mt][0]);
                    top2_update_s(b1[mt * 2], b2[mt * 2], k[mt][1]);
                    top2_update_s(b1[mt * 2 + 1], b2[mt * 2 + 1], k[mt][2]);
                    top2_update_s(b1[mt * 2 + 1], b2[mt * 2 + 1], k[mt][3]);
                }
            };
            if (j0 >= 4096) {
#pragma unroll 2
                for (int n8 = 0; n8 < full; n8++) n8tile(n8, false, true);
            } else {
#pragma unroll 2
                for (int n8 = 0; n8 < full; n8++) n8tile(n8, false, false);
            }
            if (rows & 7) n8tile(full, true, false);
            if (haveNext) stage(cur ^ 1, nx);
            __syncthreads();
        }
#pragma unroll
        for (int r = 0; r < 4; r++) {
#pragma unroll
            for (int o = 1; o <= 2; o <<= 1) {
                const int o1 = __shfl_xor_sync(0xffffffffu, b1[r], o), o2 = __shfl_xor_sync(0xffffffffu, b2[r], o);
                b2[r] = min(max(b1[r], o1), min(b2[r], o2));
                b1[r] = min(b1[r], o1);
            }
            merge_super_s(d1[r], i1[r], d2[r], i2[r], b1[r], na[r], sb);
            merge_super_s(d1[r], i1[r], d2[r], i2[r], b2[r], na[r], sb);
        }
    }
    if (tg == 0) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const long long row = rowBase + (r >> 1) * 16 + g + 8 * (r & 1);
            if (row < nq) {
                const long long o = ((long long)pair * qCap + row) * 2;
                *reinterpret_cast<int2*>(idx + o) = make_int2(i1[r], i2[r]);
                *reinterpret_cast<int2*>(dist + o) = make_int2(d1[r], d2[r]);
            }
        }
    }
}

// small problems: one warp per query, train rows strided over the lanes, warp-shuffle top-2 merge.
// grid (ceil(maxNq / 8), npairs); block 256.  nt < 2^23 per pair (the host dispatches).
__global__ void __launch_bounds__(KNN_THREADS) knn2_warp_kernel(const uint8_t* __restrict__ q,
                                                                const int* __restrict__ nqArr, long long qCap,
                                                                const uint8_t* __restrict__ t,
                                                                const int* __restrict__ ntArr, long long tCap,
                                                                long long nqSingle, long long ntSingle,
                                                                int32_t* __restrict__ idx, int32_t* __restrict__ dist) {
    const int pair = blockIdx.y, lane = threadIdx.x & 31;
    const int nq = nqArr ? nqArr[pair] : (int)nqSingle;
    const int nt = ntArr ? ntArr[pair] : (int)ntSingle;
    const int row = blockIdx.x * (KNN_THREADS / 32) + (threadIdx.x >> 5);
    if (row >= nq) return;
    const uint4* qp = reinterpret_cast<const uint4*>(q + ((long long)pair * qCap + row) * 32);
    const uint4 qa = __ldg(qp), qb = __ldg(qp + 1);
    const uint4* tp = reinterpret_cast<const uint4*>(t + (long long)pair * tCap * 32);
    uint32_t b1 = KNN_EMPTY, b2 = KNN_EMPTY;
    for (int j = lane; j < nt; j += 32) {
        const uint4 ta = __ldg(tp + 2 * j), tb = __ldg(tp + 2 * j + 1);
        const uint32_t d = __popc(qa.x ^ ta.x) + __popc(qa.y ^ ta.y) + __popc(qa.z ^ ta.z) + __popc(qa.w ^ ta.w) +
                           __popc(qb.x ^ tb.x) + __popc(qb.y ^ tb.y) + __popc(qb.z ^ tb.z) + __popc(qb.w ^ tb.w);
        top2_update(b1, b2, (d << KNN_SUPER_BITS) | (uint32_t)j);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t o1 = __shfl_xor_sync(0xffffffffu, b1, o), o2 = __shfl_xor_sync(0xffffffffu, b2, o);
        // top-2 of the union {b1, b2, o1, o2}; all keys are distinct (distinct train indices) or EMPTY
        const uint32_t lo = min(b1, o1), hi = max(b1, o1);
        b2 = min(hi, min(b2, o2));
        b1 = lo;
    }
    if (lane == 0) {
        int d1 = -1, i1 = -1, d2 = -1, i2 = -1;
        merge_super(d1, i1, d2, i2, b1, 0);
        merge_super(d1, i1, d2, i2, b2, 0);
        const long long o = ((long long)pair * qCap + row) * 2;
        *reinterpret_cast<int2*>(idx + o) = make_int2(i1, i2);
        *reinterpret_cast<int2*>(dist + o) = make_int2(d1, d2);
    }
}

int launch_knn2_tc(const uint8_t* q, const int* nq, long long qCap, const uint8_t* t, const int* nt, long long tCap,
                   int npairs, long long nqSingle, long long ntSingle, long long maxNq, int32_t* idx, int32_t* dist,
                   cudaStream_t st);

// the tcgen05 kernel (k_knn_tc.cu) is the default from 296 CTAs up; ORBFE_KNN_TC=0 falls back to the mma.sync kernel
static bool knn_use_tc() {
    static const int v = [] {
        const char* e = getenv("ORBFE_KNN_TC");
        return e ? atoi(e) : 1;
    }();
    return v != 0;
}

bool knn_tc_enabled() { return knn_use_tc(); }

// ORBFE_KNN_MMA=0 selects the POPC kernel (A/B measurements); default: tensor cores
static bool knn_use_mma() {
    static const int v = [] {
        const char* e = getenv("ORBFE_KNN_MMA");
        return e ? atoi(e) : 1;
    }();
    return v != 0;
}

int launch_knn2(const uint8_t* q, const int* nq, long long qCap, const uint8_t* t, const int* nt, long long tCap,
                int npairs, long long nqSingle, long long ntSingle, long long maxNq, long long maxNt, int32_t* idx,
                int32_t* dist, cudaStream_t st) {
    if (npairs <= 0 || maxNq <= 0) return 0;
    // thread-per-query needs >= ~2 waves of CTAs to fill 148 SMs; below that give every query a warp
    const long long ctasBig = ((maxNq + KNN_THREADS - 1) / KNN_THREADS) * npairs;
    if (ctasBig < 296 && maxNt < (1ll << KNN_SUPER_BITS)) {
        const dim3 grid((unsigned)((maxNq + 7) / 8), npairs);
        knn2_warp_kernel<<<grid, KNN_THREADS, 0, st>>>(q, nq, qCap, t, nt, tCap, nqSingle, ntSingle, idx, dist);
    } else if (knn_use_tc()) {
        return launch_knn2_tc(q, nq, qCap, t, nt, tCap, npairs, nqSingle, ntSingle, maxNq, idx, dist, st);
    } else if (knn_use_mma()) {
        const dim3 grid((unsigned)((maxNq + KM_ROWS - 1) / KM_ROWS), npairs);
        static SmemOptIn optIn;
        smem_opt_in(optIn, knn2_mma_kernel, (size_t)KM_SMEM);
        knn2_mma_kernel<<<grid, KM_THREADS, KM_SMEM, st>>>(q, nq, qCap, t, nt, tCap, nqSingle, ntSingle, idx, dist);
    } else {
        const dim3 grid((unsigned)((maxNq + KNN_THREADS - 1) / KNN_THREADS), npairs);
        knn2_kernel<<<grid, KNN_THREADS, 0, st>>>(q, nq, qCap, t, nt, tCap, nqSingle, ntSingle, idx, dist);
    }
    return 1;
}

}    // namespace orbfe
