// k_knn_tc.cu — Hamming kNN (k = 2) on the 5th-generation tensor cores: tcgen05.mma kind::i8 with the accumulators in
// tensor memory.  Same arithmetic as knn2_mma_kernel (k_knn.cu): hamming(a, b) - |a| = sum_k b_k * (1 - 2 a_k) with
// A' = 1 - 2a in {+1, -1} and B = b in {0, 1} as int8, exact in s32.
//
//   * CTA = 128 query rows (UMMA M = 128: accumulator row i = TMEM lane i) x tiles of KT_N train rows (UMMA N);
//     K = 256 = 8 instructions of K = 32.  Operands live in shared memory in the canonical K-major, no-swizzle layout:
//     core matrix = 8 rows x 16 bytes (128 contiguous bytes), SBO = 128 B between 8-row groups, LBO = rows / 8 * 128 B
//     between the two 16-byte K chunks of an instruction.  The k order of a dot product is free as long as A and B
//     agree, so a 16-byte chunk is simply four registers (word >> n) & 0x01010101 — bits are never transposed.
//   * the query tile is expanded once; every train tile is expanded by all threads from packed global rows (32 B per
//     row) straight into the canonical layout with conflict-free STS.128, made visible to the async proxy
//     (fence.proxy.async), then one thread issues the 8 MMAs and a tcgen05.commit onto an mbarrier;
//   * epilogue: warp w reads its 32 TMEM lanes with tcgen05.ld.32x32b.x32 (thread = query row, 32 train columns per
//     load), turns each accumulator into key = acc << S | column and keeps the two smallest per row in registers
//     (distance first, train index second = OpenCV's tie order); stored keys are re-based by -32 per chunk so the column
//     constants stay compile-time;
//   * no warp specialisation: phases of the 2-3 co-resident CTAs of an SM overlap each other (expansion and epilogue on
//     the ALU pipes of one CTA under the MMAs of another); the next tile's packed rows are prefetched into registers
//     before the wait;
//   * wave quantisation: one (query tile) CTA per 128 rows means a 125k-row shard (1M rows over 8 GPUs) is 977 CTAs =
//     2.2 waves of the 444 resident CTAs -> 27 % of the last wave idle.  For a single large problem the launcher can
//     cut the TRAIN set into `nseg` segments (grid.z): CTA (tile, z) keeps the top 2 of its segment, writes them to a
//     scratch array, and the last CTA of a tile to finish (atomic ticket) merges the nseg partial results in segment
//     order -- strict "<" keeps the lowest train index on ties, as OpenCV does.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <mutex>

#include "orbfe_internal.h"

namespace orbfe {

#ifndef ORBFE_KT_N
#define ORBFE_KT_N 128
#endif
constexpr int KT_M = 128;
constexpr int KT_N = ORBFE_KT_N;    // train rows per tile = TMEM columns
constexpr int KT_THREADS = 128;
constexpr int KT_S = 22;
constexpr int KT_EMPTY = 0x7fffffff;
constexpr int KT_RPT = KT_N / KT_THREADS;    // train rows expanded per thread and tile
static_assert(KT_N % 128 == 0 && KT_N <= 256, "tile = 1 or 2 rows per thread");
constexpr int KT_SMEM = KT_M * 256 + KT_N * 256 + 64;
constexpr uint32_t KT_IDESC = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(KT_N >> 3) << 17) | ((uint32_t)(KT_M >> 4) << 24);

// K-major, SWIZZLE_NONE shared-memory matrix descriptor (version 1 = sm_100)
__device__ __forceinline__ uint64_t kt_desc(uint32_t saddr, uint32_t lboBytes, uint32_t sboBytes) {
    return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lboBytes >> 4) & 0x3fffu) << 16) |
           ((uint64_t)((sboBytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}

__device__ __forceinline__ void kt_mma_i8(uint32_t dTmem, uint64_t aDesc, uint64_t bDesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(dTmem),
        "l"(aDesc), "l"(bDesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

__device__ __forceinline__ void kt_ld32(uint32_t taddr, int (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// expand one packed 256-bit row into its 16 chunks of the canonical layout; `rows` = rows of the operand tile
template <bool SIGNED_A>
__device__ __forceinline__ void kt_expand_row(uint8_t* base, int rows, int r, const uint4& lo, const uint4& hi) {
    constexpr uint32_t M = 0x01010101u;
    const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    uint8_t* rowp = base + (r >> 3) * 128 + (r & 7) * 16;
    const int chunkStride = (rows >> 3) * 128;
#pragma unroll
    for (int q = 0; q < 16; q++) {
        const int p = q & 3, j = q >> 2;
        uint32_t x0 = (w[2 * p] >> (2 * j)) & M, x1 = (w[2 * p] >> (2 * j + 1)) & M;
        uint32_t x2 = (w[2 * p + 1] >> (2 * j)) & M, x3 = (w[2 * p + 1] >> (2 * j + 1)) & M;
        if (SIGNED_A) {
            x0 = (x0 * 0xffu) | M;
            x1 = (x1 * 0xffu) | M;
            x2 = (x2 * 0xffu) | M;
            x3 = (x3 * 0xffu) | M;
        }
        *reinterpret_cast<uint4*>(rowp + q * chunkStride) = make_uint4(x0, x1, x2, x3);
    }
}

// grid (ceil(maxNq / KT_M), npairs); block KT_THREADS; dynamic shared memory KT_SMEM
__global__ void __launch_bounds__(KT_THREADS) knn2_tc_kernel(const uint8_t* __restrict__ q, const int* __restrict__ nqArr,
                                                            long long qCap, const uint8_t* __restrict__ t,
                                                            const int* __restrict__ ntArr, long long tCap,
                                                            long long nqSingle, long long ntSingle,
                                                            int32_t* __restrict__ idx, int32_t* __restrict__ dist,
                                                            long long segLen, int4* __restrict__ partial,
                                                            unsigned* __restrict__ ticket, int zBase, int zTotal) {
    extern __shared__ __align__(128) uint8_t ktSmem[];
    uint8_t* sA = ktSmem;
    uint8_t* sB = ktSmem + KT_M * 256;
    uint64_t* bar = reinterpret_cast<uint64_t*>(ktSmem + KT_M * 256 + KT_N * 256);
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(bar + 1);

    const int pair = blockIdx.y;
    const long long nq = nqArr ? nqArr[pair] : nqSingle;
    const long long nt = ntArr ? ntArr[pair] : ntSingle;
    if ((long long)blockIdx.x * KT_M >= nq) return;    // uniform: before any allocation
    const int tid = threadIdx.x, warp = tid >> 5;
    const long long row = (long long)blockIdx.x * KT_M + tid;
    const uint4* tp = reinterpret_cast<const uint4*>(t + (long long)pair * tCap * 32);

    if (tid == 0) mbar_init(bar, 1);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmemSlot)), "n"(KT_N)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // the query tile, +1 / -1
    int na = 0;
    {
        uint4 lo = make_uint4(0u, 0u, 0u, 0u), hi = lo;
        if (row < nq) {
            const uint4* qp = reinterpret_cast<const uint4*>(q + ((long long)pair * qCap + row) * 32);
            lo = __ldg(qp);
            hi = __ldg(qp + 1);
        }
        na = __popc(lo.x) + __popc(lo.y) + __popc(lo.z) + __popc(lo.w) + __popc(hi.x) + __popc(hi.y) + __popc(hi.z) + __popc(hi.w);
        kt_expand_row<true>(sA, KT_M, tid, lo, hi);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmemSlot;
    const uint32_t aBase = smem_u32(sA), bBase = smem_u32(sB);

    int d1 = -1, i1 = -1, d2 = -1, i2 = -1;
    uint32_t phase = 0;
    // train rows of this CTA: everything, or segment zg of zTotal (segLen is a multiple of KT_N).  The segments of one
    // problem may be spread over several launches (zBase = first segment of this launch): the sharded matcher starts
    // the segments of a train chunk as soon as that chunk has arrived (orbfe_sharded.cu)
    const int zg = zBase + (int)blockIdx.z;
    const long long tBeg = zTotal > 1 ? min(nt, (long long)zg * segLen) : 0;
    const long long tEnd = zTotal > 1 ? min(nt, tBeg + segLen) : nt;
    for (long long sb = tBeg; sb < tEnd; sb += (1ll << (KT_S - 1))) {
        const int ns = (int)min(tEnd - sb, 1ll << (KT_S - 1));
        int b1 = KT_EMPTY, b2 = KT_EMPTY;    // keys relative to the current chunk: (acc << S) + (column - chunk base)
        const int ntiles = (ns + KT_N - 1) / KT_N;
        uint4 nx[KT_RPT][2];
        auto fetch = [&](int j0) {
#pragma unroll
            for (int u = 0; u < KT_RPT; u++) {
                const int r = j0 + u * KT_THREADS + tid;
                nx[u][0] = nx[u][1] = make_uint4(0u, 0u, 0u, 0u);
                if (r < ns) {
                    nx[u][0] = __ldg(tp + (sb + r) * 2);
                    nx[u][1] = __ldg(tp + (sb + r) * 2 + 1);
                }
            }
        };
        fetch(0);
        for (int ti = 0; ti < ntiles; ti++) {
            const int j0 = ti * KT_N;
            // B tile -> shared memory (the previous tile's MMAs have completed: every thread waited on the barrier)
#pragma unroll
            for (int u = 0; u < KT_RPT; u++) kt_expand_row<false>(sB, KT_N, u * KT_THREADS + tid, nx[u][0], nx[u][1]);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");    // orders the previous tile's tcgen05.ld
            __syncthreads();
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int s = 0; s < 8; s++) {
                    const uint64_t ad = kt_desc(aBase + 2 * s * (KT_M >> 3) * 128, (KT_M >> 3) * 128, 128);
                    const uint64_t bd = kt_desc(bBase + 2 * s * (KT_N >> 3) * 128, (KT_N >> 3) * 128, 128);
                    kt_mma_i8(tmem, ad, bd, KT_IDESC, s > 0);
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                             : "memory");
            }
            if (ti + 1 < ntiles) fetch(j0 + KT_N);
            mbar_wait(bar, phase);
            phase ^= 1;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const bool guard = j0 + KT_N > ns;
#pragma unroll 1
            for (int c = 0; c < KT_N; c += 32) {
                int v[32];
                kt_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, v);
                if (guard) {
#pragma unroll
                    for (int i = 0; i < 32; i++) {
                        const int k = j0 + c + i < ns ? v[i] * (1 << KT_S) + i : KT_EMPTY;
                        b2 = min(b2, max(b1, k));
                        b1 = min(b1, k);
                    }
                } else {
                    // long train sets: after a few thousand columns a chunk rarely holds a candidate for a row's top 2,
                    // so look at the smallest accumulator first (16 three-input minima instead of ~80 operations)
                    bool need = true;
                    if (j0 >= 2048) {
                        int m = min(v[0], v[1]);
#pragma unroll
                        for (int i = 2; i < 32; i += 2) m = min(m, min(v[i], v[i + 1]));
                        need = m * (1 << KT_S) < b2;    // key_i >= acc_i << S
                    }
                    if (need) {
#pragma unroll
                        for (int i = 0; i < 32; i++) {
                            const int k = v[i] * (1 << KT_S) + i;
                            b2 = min(b2, max(b1, k));
                            b1 = min(b1, k);
                        }
                    }
                }
                b1 -= 32;    // re-base to the next chunk (EMPTY drifts by < 2^21 per super block: still above any real key)
                b2 -= 32;
            }
        }
        // decode: key = d * 2^S + (index - base), |index - base| < 2^(S-1); base = columns processed so far
        const long long base = sb + (long long)ntiles * KT_N;
        auto merge = [&](int b) {
            if (b > (400 << KT_S)) return;    // empty
            const int d = (b + (1 << (KT_S - 1))) >> KT_S;
            const int i = (int)(base + (b - d * (1 << KT_S)));
            const int dd = d + na;
            if (i1 < 0 || dd < d1) { d2 = d1; i2 = i1; d1 = dd; i1 = i; }
            else if (i2 < 0 || dd < d2) { d2 = dd; i2 = i; }
        };
        merge(b1);
        merge(b2);
    }
    if (zTotal > 1) {
        // partial top 2 of this segment -> scratch; the last CTA of the tile (over all launches) merges all segments in
        // train order
        const long long rowsPad = (long long)gridDim.x * KT_M;
        partial[(long long)zg * rowsPad + row] = make_int4(i1, i2, d1, d2);
        __threadfence();
        __syncthreads();
        if (tid == 0) *tmemSlot = atomicAdd(&ticket[blockIdx.x], 1u);    // tmemSlot is free: `tmem` holds the address
        __syncthreads();
        const bool last = *tmemSlot == (unsigned)(zTotal - 1);
        if (last) {
            __threadfence();
            i1 = i2 = d1 = d2 = -1;
            for (int z = 0; z < zTotal; z++) {
                const int4 pz = __ldcg(&partial[(long long)z * rowsPad + row]);
                if (pz.x >= 0 && (i1 < 0 || pz.z < d1)) { d2 = d1; i2 = i1; d1 = pz.z; i1 = pz.x; }
                else if (pz.x >= 0 && (i2 < 0 || pz.z < d2)) { d2 = pz.z; i2 = pz.x; }
                if (pz.y >= 0 && (i2 < 0 || pz.w < d2)) { d2 = pz.w; i2 = pz.y; }
            }
            if (tid == 0) ticket[blockIdx.x] = 0;    // ready for the next launch
        }
        if (last && row < nq) {
            const long long o = ((long long)pair * qCap + row) * 2;
            *reinterpret_cast<int2*>(idx + o) = make_int2(i1, i2);
            *reinterpret_cast<int2*>(dist + o) = make_int2(d1, d2);
        }
    } else if (row < nq) {
        const long long o = ((long long)pair * qCap + row) * 2;
        *reinterpret_cast<int2*>(idx + o) = make_int2(i1, i2);
        *reinterpret_cast<int2*>(dist + o) = make_int2(d1, d2);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(KT_N) : "memory");
}

// ---- knn2_tck_kernel: the per-pair (C2 / C3) shape with the top-2 KEY computed by the tensor core -----------------------------
// At 1200 x 1200 per pair a CTA runs 10 train tiles and the epilogue is 60 % of the kernel's instructions: per accumulator
// one LEA to build key = acc << S | column and ~2.2 VIMNMX for the top 2, plus a guarded copy of the loop for the last,
// partial tile.  Here the MMA delivers the key itself:
//   * A' = +-64 (instead of +-1) and B = 2 b (instead of b): every product is +-128, acc' = 128 (hamming - |a|);
//   * a ninth K = 32 instruction adds the column: A row = [1, 127, 127, 127, 127, 0 ...], B row of tile column j =
//     [j, inv, inv, inv, inv, 0 ...] with inv = 127 for columns past the end of the train set, so the accumulator is
//     128 (hamming - |a|) + j (+ 64516 for an invalid column: above every valid key);
//   * the epilogue is the bare top-2 insertion on the accumulators (no LEA, no guard), decoded once per tile:
//     distance = (key >> 7) + |a|, column = key & 127.  Equal distances keep the lower train index (lower j inside a
//     tile, earlier tile across tiles), as OpenCV does.
// 12.5 % more tensor work for ~40 % fewer instructions; used for batches of pairs and short train sets (the long-train-set
// kernel above keeps its prefilter and its train segments).
constexpr int KK_CHUNKS = 18;                               // 16 descriptor chunks + the key chunk pair
constexpr int KK_SMEM = (KT_M + KT_N) * 16 * KK_CHUNKS + 64;
constexpr int KK_EMPTY = 0x7fffffff;
static_assert(KT_N <= 128 || true, "the key holds a 7-bit tile column: the launcher uses this kernel for KT_N == 128 only");

template <bool SIGNED_A>
__device__ __forceinline__ void kk_expand_row(uint8_t* base, int rows, int r, const uint4& lo, const uint4& hi) {
    const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    uint8_t* rowp = base + (r >> 3) * 128 + (r & 7) * 16;
    const int chunkStride = (rows >> 3) * 128;
    auto ex = [](uint32_t v, int n) -> uint32_t {    // bit n of every byte of v -> A: 0x40 | bit << 7 (= +-64), B: bit << 1
        if (SIGNED_A) return ((v << (7 - n)) & 0x80808080u) | 0x40404040u;
        return (n >= 1 ? (v >> (n - 1)) : (v << 1)) & 0x02020202u;
    };
#pragma unroll
    for (int q = 0; q < 16; q++) {
        const int p = q & 3, j = q >> 2;
        *reinterpret_cast<uint4*>(rowp + q * chunkStride) =
            make_uint4(ex(w[2 * p], 2 * j), ex(w[2 * p], 2 * j + 1), ex(w[2 * p + 1], 2 * j), ex(w[2 * p + 1], 2 * j + 1));
    }
}

// grid (ceil(maxNq / KT_M), npairs); block KT_THREADS; dynamic shared memory KK_SMEM
__global__ void __launch_bounds__(KT_THREADS) knn2_tck_kernel(const uint8_t* __restrict__ q, const int* __restrict__ nqArr,
                                                             long long qCap, const uint8_t* __restrict__ t,
                                                             const int* __restrict__ ntArr, long long tCap,
                                                             long long nqSingle, long long ntSingle,
                                                             int32_t* __restrict__ idx, int32_t* __restrict__ dist) {
    extern __shared__ __align__(128) uint8_t ktSmem[];
    uint8_t* sA = ktSmem;
    uint8_t* sB = ktSmem + KT_M * 16 * KK_CHUNKS;
    uint64_t* bar = reinterpret_cast<uint64_t*>(ktSmem + (KT_M + KT_N) * 16 * KK_CHUNKS);
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(bar + 1);

    const int pair = blockIdx.y;
    const long long nq = nqArr ? nqArr[pair] : nqSingle;
    const long long nt = ntArr ? ntArr[pair] : ntSingle;
    if ((long long)blockIdx.x * KT_M >= nq) return;    // uniform: before any allocation
    const int tid = threadIdx.x, warp = tid >> 5;
    const long long row = (long long)blockIdx.x * KT_M + tid;
    const uint4* tp = reinterpret_cast<const uint4*>(t + (long long)pair * tCap * 32);

    if (tid == 0) mbar_init(bar, 1);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmemSlot)), "n"(KT_N)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    int na = 0;
    {
        uint4 lo = make_uint4(0u, 0u, 0u, 0u), hi = lo;
        if (row < nq) {
            const uint4* qp = reinterpret_cast<const uint4*>(q + ((long long)pair * qCap + row) * 32);
            lo = __ldg(qp);
            hi = __ldg(qp + 1);
        }
        na = __popc(lo.x) + __popc(lo.y) + __popc(lo.z) + __popc(lo.w) + __popc(hi.x) + __popc(hi.y) + __popc(hi.z) + __popc(hi.w);
        kk_expand_row<true>(sA, KT_M, tid, lo, hi);
        uint8_t* rowp = sA + (tid >> 3) * 128 + (tid & 7) * 16;    // key chunk pair of the A row: [1, 127 x 4, 0 ...]
        *reinterpret_cast<uint4*>(rowp + 16 * (KT_M >> 3) * 128) = make_uint4(0x7f7f7f01u, 0x0000007fu, 0u, 0u);
        *reinterpret_cast<uint4*>(rowp + 17 * (KT_M >> 3) * 128) = make_uint4(0u, 0u, 0u, 0u);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmemSlot;
    const uint32_t aBase = smem_u32(sA), bBase = smem_u32(sB);

    int d1 = -1, i1 = -1, d2 = -1, i2 = -1;
    uint32_t phase = 0;
    const long long ntiles = (nt + KT_N - 1) / KT_N;
    uint4 nx[KT_RPT][2];
    auto fetch = [&](long long j0) {
#pragma unroll
        for (int u = 0; u < KT_RPT; u++) {
            const long long r = j0 + u * KT_THREADS + tid;
            nx[u][0] = nx[u][1] = make_uint4(0u, 0u, 0u, 0u);
            if (r < nt) {
                nx[u][0] = __ldg(tp + r * 2);
                nx[u][1] = __ldg(tp + r * 2 + 1);
            }
        }
    };
    fetch(0);
    for (long long ti = 0; ti < ntiles; ti++) {
        const long long j0 = ti * KT_N;
        // B tile -> shared memory (the previous tile's MMAs have completed: every thread waited on the barrier)
#pragma unroll
        for (int u = 0; u < KT_RPT; u++) {
            const int r = u * KT_THREADS + tid;
            kk_expand_row<false>(sB, KT_N, r, nx[u][0], nx[u][1]);
            const uint32_t inv = j0 + r < nt ? 0u : 0x7fu;
            uint8_t* rowp = sB + (r >> 3) * 128 + (r & 7) * 16;
            *reinterpret_cast<uint4*>(rowp + 16 * (KT_N >> 3) * 128) = make_uint4((uint32_t)r | (inv * 0x01010100u), inv, 0u, 0u);
            *reinterpret_cast<uint4*>(rowp + 17 * (KT_N >> 3) * 128) = make_uint4(0u, 0u, 0u, 0u);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");    // orders the previous tile's tcgen05.ld
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int s = 0; s < KK_CHUNKS / 2; s++) {
                const uint64_t ad = kt_desc(aBase + 2 * s * (KT_M >> 3) * 128, (KT_M >> 3) * 128, 128);
                const uint64_t bd = kt_desc(bBase + 2 * s * (KT_N >> 3) * 128, (KT_N >> 3) * 128, 128);
                kt_mma_i8(tmem, ad, bd, KT_IDESC, s > 0);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                         : "memory");
        }
        if (ti + 1 < ntiles) fetch(j0 + KT_N);
        mbar_wait(bar, phase);
        phase ^= 1;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        int b1 = KK_EMPTY, b2 = KK_EMPTY;
#pragma unroll 1
        for (int c = 0; c < KT_N; c += 32) {
            int v[32];
            kt_ld32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c, v);
#pragma unroll
            for (int i = 0; i < 32; i++) {
                b2 = min(b2, max(b1, v[i]));
                b1 = min(b1, v[i]);
            }
        }
        // decode: key = 128 (hamming - |a|) + column of the tile; invalid columns sit above 64000
        auto merge = [&](int b) {
            if (b > 48000) return;
            const int dd = (b >> 7) + na;
            const int i = (int)(j0 + (b & 127));
            if (i1 < 0 || dd < d1) { d2 = d1; i2 = i1; d1 = dd; i1 = i; }
            else if (i2 < 0 || dd < d2) { d2 = dd; i2 = i; }
        };
        merge(b1);
        merge(b2);
    }
    if (row < nq) {
        const long long o = ((long long)pair * qCap + row) * 2;
        *reinterpret_cast<int2*>(idx + o) = make_int2(i1, i2);
        *reinterpret_cast<int2*>(dist + o) = make_int2(d1, d2);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(KT_N) : "memory");
}

// train segments for one large problem: the smallest count (<= 16) whose CTA total fills the resident-CTA waves to
// >= 98.5 %, else the best; segments stay >= 8k rows.  ORBFE_KNN_SEGS=n forces n (tests).
static int kt_pick_segments(long long tiles, long long nt, int resident) {
    static const int forced = [] {    // read once: the environment is not re-scanned per launch
        const char* e = getenv("ORBFE_KNN_SEGS");
        const int v = e ? atoi(e) : 0;
        return v >= 1 ? (v < 64 ? v : 64) : 0;
    }();
    if (forced) return forced;
    int best = 1;
    double bestEff = 0;
    for (int sgs = 1; sgs <= 16; sgs++) {    // a CTA's prologue (query tile) is ~0.1 % of a >= 8k-row segment: more segments are free
        if (sgs > 1 && nt / sgs < 8192) break;
        const long long ctas = tiles * sgs, waves = (ctas + resident - 1) / resident;
        const double eff = (double)ctas / (double)(waves * resident);
        if (eff > bestEff + 1e-9) { bestEff = eff; best = sgs; }
        if (eff >= 0.985) break;
    }
    return best;
}

int launch_knn2_tc(const uint8_t* q, const int* nq, long long qCap, const uint8_t* t, const int* nt, long long tCap,
                   int npairs, long long nqSingle, long long ntSingle, long long maxNq, int32_t* idx, int32_t* dist,
                   cudaStream_t st) {
    // per-device launch state (matcher handles on distinct devices / threads are independent: orbfe_c.h)
    static SmemOptIn optIn;
    static std::once_flag residentOnce[64];
    static int residentOf[64];
    smem_opt_in(optIn, knn2_tc_kernel, (size_t)KT_SMEM);
    int dev = 0;
    cudaGetDevice(&dev);
    std::call_once(residentOnce[dev & 63], [dev] {
        int sms = 0, perSm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, knn2_tc_kernel, KT_THREADS, KT_SMEM);
        residentOf[dev & 63] = sms * (perSm > 0 ? perSm : 1);
    });
    const int resident = residentOf[dev & 63];
    const long long tiles = (maxNq + KT_M - 1) / KT_M;
    // batches of pairs and short train sets: the keyed kernel (ORBFE_KNN_KEYED=0: the kernel above for everything)
    static const bool keyed = [] { const char* e = getenv("ORBFE_KNN_KEYED"); return !e || atoi(e) != 0; }();
    if (keyed && KT_N == 128 && (npairs > 1 || nq || nt || ntSingle <= 8192)) {
        static SmemOptIn optInK;
        smem_opt_in(optInK, knn2_tck_kernel, (size_t)KK_SMEM);
        knn2_tck_kernel<<<dim3((unsigned)tiles, npairs), KT_THREADS, KK_SMEM, st>>>(q, nq, qCap, t, nt, tCap, nqSingle, ntSingle, idx,
                                                                                dist);
        return 1;
    }
    int nseg = 1;
    if (npairs == 1 && !nq && !nt) nseg = kt_pick_segments(tiles, ntSingle, resident);
    // stream-ordered scratch for the segmented launch: partial results [nseg][tiles * 128] + one ticket per tile
    const long long segLen = (((ntSingle + (nseg > 1 ? nseg : 1) - 1) / (nseg > 1 ? nseg : 1)) + KT_N - 1) / KT_N * KT_N;
    const size_t partialBytes = (size_t)nseg * tiles * KT_M * sizeof(int4), ticketBytes = (size_t)tiles * sizeof(unsigned);
    uint8_t* scratch = nullptr;
    if (nseg > 1 && cudaMallocAsync(&scratch, partialBytes + ticketBytes, st) != cudaSuccess) {
        cudaGetLastError();    // no memory pool / no memory: the unsegmented launch needs no scratch
        nseg = 1;
    }
    if (nseg <= 1) {
        const dim3 grid((unsigned)tiles, npairs);
        knn2_tc_kernel<<<grid, KT_THREADS, KT_SMEM, st>>>(q, nq, qCap, t, nt, tCap, nqSingle, ntSingle, idx, dist, 0, nullptr,
                                                          nullptr, 0, 1);
        return 1;
    }
    cudaMemsetAsync(scratch + partialBytes, 0, ticketBytes, st);
    const dim3 grid((unsigned)tiles, 1, nseg);
    knn2_tc_kernel<<<grid, KT_THREADS, KT_SMEM, st>>>(q, nq, qCap, t, nt, tCap, nqSingle, ntSingle, idx, dist, segLen,
                                                      reinterpret_cast<int4*>(scratch),
                                                      reinterpret_cast<unsigned*>(scratch + partialBytes), 0, nseg);
    cudaFreeAsync(scratch, st);
    return 1;
}

// ---- multi-launch form (orbfe_sharded.cu) -----------------------------------------------------------------------------
// One nq x nt problem whose train set arrives in `chunks` pieces.  plan: segments per chunk (the smallest count <= 16
// whose CTA total fills the resident-CTA waves of ONE launch to >= 98.5 %, else the best), the segment length (a multiple
// of KT_N) and the scratch bytes (partial results of every segment + one ticket per query tile; tickets must be zero
// before the first launch and are zero again after the last).  Chunk c = segments [c spc, (c + 1) spc) = train rows
// [c spc segLen, (c + 1) spc segLen).
static int kt_resident_ctas() {
    static std::once_flag once[64];
    static int resident[64];
    static SmemOptIn optIn;
    smem_opt_in(optIn, knn2_tc_kernel, (size_t)KT_SMEM);
    int dev = 0;
    cudaGetDevice(&dev);
    std::call_once(once[dev & 63], [dev] {
        int sms = 0, perSm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, knn2_tc_kernel, KT_THREADS, KT_SMEM);
        resident[dev & 63] = sms * (perSm > 0 ? perSm : 1);
    });
    return resident[dev & 63];
}

void knn2_tc_plan(long long nq, long long nt, int chunks, int* segsPerChunk, long long* segLen, int* zTotal,
                  size_t* scratchBytes) {
    const long long tiles = (nq + KT_M - 1) / KT_M;
    const int resident = kt_resident_ctas();
    int best = 1;
    double bestEff = 0;
    for (int sgs = 1; sgs <= 16; sgs++) {
        if (sgs > 1 && nt / ((long long)chunks * sgs) < 4096) break;
        const long long ctas = tiles * sgs, waves = (ctas + resident - 1) / resident;
        const double eff = (double)ctas / (double)(waves * resident);
        if (eff > bestEff + 1e-9) { bestEff = eff; best = sgs; }
        if (eff >= 0.985) break;
    }
    const long long want = (long long)chunks * best;
    long long sl = ((nt + want - 1) / want + KT_N - 1) / KT_N * KT_N;
    if (sl < KT_N) sl = KT_N;
    *segsPerChunk = best;
    *segLen = sl;
    *zTotal = (int)std::max<long long>(1, (nt + sl - 1) / sl);
    *scratchBytes = (size_t)(*zTotal) * tiles * KT_M * sizeof(int4) + (size_t)tiles * sizeof(unsigned);
}

// launches the segments [zBase, zBase + zCount) of the plan; scratch = [zTotal x planRows/128 x 128 int4][tickets], laid
// out for the planRows the plan was made with (nq <= planRows: a rank may own fewer rows than the largest shard)
int launch_knn2_tc_segments(const uint8_t* q, long long nq, long long planRows, const uint8_t* t, long long nt, int32_t* idx,
                            int32_t* dist, long long segLen, int zBase, int zCount, int zTotal, uint8_t* scratch,
                            cudaStream_t st) {
    if (nq <= 0 || zCount <= 0) return 0;
    kt_resident_ctas();    // shared-memory opt-in of this device
    const long long tiles = (nq + KT_M - 1) / KT_M;
    const size_t partialBytes = (size_t)zTotal * ((planRows + KT_M - 1) / KT_M) * KT_M * sizeof(int4);
    const dim3 grid((unsigned)tiles, 1, (unsigned)zCount);
    // zTotal == 1 would take the kernel's direct-store path, which needs no scratch
    knn2_tc_kernel<<<grid, KT_THREADS, KT_SMEM, st>>>(q, nullptr, nq, t, nullptr, nt, nq, nt, idx, dist, segLen,
                                                      reinterpret_cast<int4*>(scratch),
                                                      reinterpret_cast<unsigned*>(scratch + partialBytes), zBase, zTotal);
    return 1;
}

// ---- orbfe_knn2_static_scratch: the segmented launch with NO allocation -------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  launch_knn2_tc above takes the scratch of
// its segmented form (partial top 2 of every train segment + one ticket per query tile) from cudaMallocAsync; this variant
// keeps it in a file-scope device array of a fixed, modest size instead.  Every GPU that loads the library holds the
// array until exit; every call on a device shares it, so calls that overlap on the GPU (different streams, different
// threads) race.  It keeps the last call's partial results (never read before they are rewritten); the tickets must start at
// zero and are cleared by every call.
constexpr long long KT_STATIC_MAX_NQ = ORBFE_KNN2_STATIC_MAX_NQ;     // 16384 query rows = 128 tiles
constexpr int KT_STATIC_MAX_SEGS = 8;
constexpr size_t KT_STATIC_PARTIAL_BYTES = (size_t)KT_STATIC_MAX_SEGS * KT_STATIC_MAX_NQ * sizeof(int4);    // 2 MiB
constexpr size_t KT_STATIC_TICKET_BYTES = (size_t)(KT_STATIC_MAX_NQ / KT_M) * sizeof(unsigned);            // 512 B
static_assert(KT_STATIC_MAX_NQ % KT_M == 0, "whole query tiles");
static __device__ __align__(256) uint8_t g_knnStaticScratch[KT_STATIC_PARTIAL_BYTES + KT_STATIC_TICKET_BYTES];

// In host code g_knnStaticScratch is a host-side stand-in, not a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current.
static int knn_static_scratch_ptr(uint8_t** out, char* err, size_t errLen) {
    static std::mutex mu;
    static uint8_t* ptrOf[64] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        if (err && errLen) snprintf(err, errLen, "orbfe_knn2_static_scratch: no usable current CUDA device (%s)", cudaGetErrorString(e));
        cudaGetLastError();
        return ORBFE_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrOf[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_knnStaticScratch);
        if (e != cudaSuccess || !p) {
            if (err && errLen) snprintf(err, errLen, "orbfe_knn2_static_scratch: cudaGetSymbolAddress failed on device %d (%s)", dev, cudaGetErrorString(e));
            cudaGetLastError();
            return ORBFE_ERR_CUDA;
        }
        ptrOf[dev] = static_cast<uint8_t*>(p);
    }
    *out = ptrOf[dev];
    return ORBFE_OK;
}

void library_count_launches(int launches);    // orbfe_api.cu

}    // namespace orbfe

extern "C" int orbfe_knn2_static_scratch(const uint8_t* q, int64_t nq, const uint8_t* t, int64_t nt, int32_t* idx, int32_t* dist,
                                         void* cuda_stream, char* err, size_t err_len) {
    using namespace orbfe;
    if (err && err_len) err[0] = 0;
    auto fail = [&](int code, const char* fmt, long long a, long long b) {
        if (err && err_len) snprintf(err, err_len, fmt, a, b);
        return code;
    };
    if (nq < 0 || nt < 0) return fail(ORBFE_ERR_ARG, "orbfe_knn2_static_scratch: negative size (nq %lld, nt %lld)", nq, nt);
    if (nq > KT_STATIC_MAX_NQ)
        return fail(ORBFE_ERR_CAPACITY, "orbfe_knn2_static_scratch: nq %lld exceeds the static scratch maximum of %lld query rows", nq,
                    KT_STATIC_MAX_NQ);
    if (nt > 0x7fffffffll) return fail(ORBFE_ERR_CAPACITY, "orbfe_knn2_static_scratch: nt %lld exceeds %lld train rows", nt, 0x7fffffffll);
    if (nq == 0) return ORBFE_OK;
    if (!q || !idx || !dist || (nt > 0 && !t)) return fail(ORBFE_ERR_ARG, "orbfe_knn2_static_scratch: null pointer (nq %lld, nt %lld)", nq, nt);
    uint8_t* scratch = nullptr;
    const int rc = knn_static_scratch_ptr(&scratch, err, err_len);
    if (rc != ORBFE_OK) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const int resident = kt_resident_ctas();    // shared-memory opt-in + occupancy of this device (no allocation)
    cudaGetLastError();    // an earlier, unrelated call of this thread may have left a non-sticky error behind: the check
                           // after the launch must report THIS launch
    const long long tiles = (nq + KT_M - 1) / KT_M;
    // segments: as many as fill the resident CTAs (<= KT_STATIC_MAX_SEGS, >= KT_N train rows each); two at least when the
    // train set has two tiles, so that the scratch path is the one that runs
    int nseg = (int)std::min<long long>(KT_STATIC_MAX_SEGS, std::max<long long>(1, nt / KT_N));
    while (nseg > 2 && tiles * (nseg - 1) >= resident) nseg--;
    long long segLen = ((nt + nseg - 1) / nseg + KT_N - 1) / KT_N * KT_N;
    if (segLen < KT_N) segLen = KT_N;
    const int zTotal = (int)std::max<long long>(1, (nt + segLen - 1) / segLen);
    if (zTotal <= 1) {
        knn2_tc_kernel<<<dim3((unsigned)tiles, 1), KT_THREADS, KT_SMEM, st>>>(q, nullptr, nq, t, nullptr, nt, nq, nt, idx, dist, 0,
                                                                               nullptr, nullptr, 0, 1);
    } else {
        const size_t partialBytes = (size_t)zTotal * tiles * KT_M * sizeof(int4);    // <= KT_STATIC_PARTIAL_BYTES by the limits above
        cudaMemsetAsync(scratch + partialBytes, 0, (size_t)tiles * sizeof(unsigned), st);
        knn2_tc_kernel<<<dim3((unsigned)tiles, 1, (unsigned)zTotal), KT_THREADS, KT_SMEM, st>>>(
            q, nullptr, nq, t, nullptr, nt, nq, nt, idx, dist, segLen, reinterpret_cast<int4*>(scratch),
            reinterpret_cast<unsigned*>(scratch + partialBytes), 0, zTotal);
    }
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        if (err && err_len) snprintf(err, err_len, "orbfe_knn2_static_scratch: launch failed (%s)", cudaGetErrorString(e));
        return ORBFE_ERR_CUDA;
    }
    library_count_launches(1);
    return ORBFE_OK;
}
