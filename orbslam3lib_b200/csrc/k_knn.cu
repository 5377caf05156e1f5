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

int launch_knn2(const uint8_t* q, const int* nq, long long qCap, const uint8_t* t, const int* nt, long long tCap,
                int npairs, long long nqSingle, long long ntSingle, long long maxNq, long long maxNt, int32_t* idx,
                int32_t* dist, cudaStream_t st) {
    if (npairs <= 0 || maxNq <= 0) return 0;
    // thread-per-query needs >= ~2 waves of CTAs to fill 148 SMs; below that give every query a warp
    const long long ctasBig = ((maxNq + KNN_THREADS - 1) / KNN_THREADS) * npairs;
    if (ctasBig < 296 && maxNt < (1ll << KNN_SUPER_BITS)) {
        const dim3 grid((unsigned)((maxNq + 7) / 8), npairs);
        knn2_warp_kernel<<<grid, KNN_THREADS, 0, st>>>(q, nq, qCap, t, nt, tCap, nqSingle, ntSingle, idx, dist);
    } else {
        const dim3 grid((unsigned)((maxNq + KNN_THREADS - 1) / KNN_THREADS), npairs);
        knn2_kernel<<<grid, KNN_THREADS, 0, st>>>(q, nq, qCap, t, nt, tCap, nqSingle, ntSingle, idx, dist);
    }
    return 1;
}

}    // namespace orbfe
