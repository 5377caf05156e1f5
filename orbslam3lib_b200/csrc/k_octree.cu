// k_octree.cu — ORBextractor::DistributeOctTree (ORBextractor_old.cc:557-781) with ExtractorNode::DivideNode
// (:482-538) and compareNodes (:540-555), one CTA per (image, level).
//
// The reference is a serial std::list / std::vector algorithm whose OUTPUT ORDER is observable (it fixes the row
// order of keypoints and descriptors), so the kernel reproduces the same sequence of list states, but computes each
// state transition in parallel:
//   * a node is (x0,y0,x1,y1) + a contiguous segment of the level's key array; splitting a node is a stable 4-way
//     partition of its segment (one warp per node, __ballot_sync ranks), so key order inside every node stays the
//     reference's vector order and "first key with the maximal response" (:765-775) is the first maximum of the
//     segment;
//   * a full pass of the first loop (:624-683) splits every node with more than one key.  Children are pushed to
//     the FRONT of the list in the order n1..n4, parents erased, so the new list is: all new children in reverse
//     creation order, then the untouched nodes in their old order.  Positions come from block-wide prefix sums;
//   * once size + 3*nToExpand > N (:691) the reference sorts the expandable nodes with std::sort(compareNodes) and
//     splits from the back until size >= N (:700-750).  compareNodes is not a total order (equal key count and
//     equal UL.x compare equivalent), so the result depends on the sort algorithm: lane 0 runs a literal port of
//     libstdc++'s introsort (median-of-3 + unguarded partition above 16 elements, heapsort at the depth limit,
//     final insertion sort), which is what the oracle's std::sort executes.  The early "break" at N is a prefix
//     sum over the sorted sequence (each split adds nonEmptyChildren-1 nodes).
// Keys are packed (x | y << 12 | response << 24), relative to (minBorderX, minBorderY).  Key segments live in shared
// memory when the level has <= keySmem candidates, otherwise in a global ping-pong scratch (same code, generic
// pointers).  Latency-bound, not bandwidth-bound: one CTA per tree, thousands of trees per batch.
#include "orbfe_internal.h"

namespace orbfe {

#ifndef ORBFE_OT_THREADS
#define ORBFE_OT_THREADS 256
#endif
constexpr int OT_THREADS = ORBFE_OT_THREADS;
constexpr int OT_WARPS = OT_THREADS / 32;

struct __align__(16) OtNode {
    short x0, y0, x1, y1;
    unsigned start;    // bit 31: which key buffer holds the segment
    int cnt;
};

struct OtSmem {
    OtNode* cur;
    OtNode* nxt;
    unsigned long long* ekey;    // (count << 32 | UL.x) of the expandable nodes, creation order
    int* epos;                   // their list positions
    int* P;                      // processing order of this pass (list positions)
    int* c4;                     // 4 quadrant counts per P entry
    int* tA;                     // scan temporaries
    int* tB;
    int* tC;
};

__host__ __device__ inline size_t octree_node_bytes(int nodeCap) {
    return ((sizeof(OtNode) * 2 + 8 + 16 + 4 + 4) * (size_t)nodeCap + 12 * (size_t)(nodeCap + 1) + 15) / 16 * 16;
}

// in-place exclusive scan of a[0..n) by the whole block; returns the total.  wsum: 2*OT_WARPS ints of shared memory.
__device__ int block_excl_scan(int* a, int n, int* wsum) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int per = (n + OT_THREADS - 1) / OT_THREADS;
    const int b = min(tid * per, n), e = min(b + per, n);
    int s = 0;
    for (int i = b; i < e; i++) s += a[i];
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int w = lane < OT_WARPS ? wsum[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < OT_WARPS; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += v;
        }
        if (lane < OT_WARPS) wsum[OT_WARPS + lane] = winc - w;
        if (lane == OT_WARPS - 1) wsum[2 * OT_WARPS] = winc;
    }
    __syncthreads();
    int run = wsum[OT_WARPS + warp] + inc - s;
    for (int i = b; i < e; i++) {
        const int v = a[i];
        a[i] = run;
        run += v;
    }
    const int total = wsum[2 * OT_WARPS];
    __syncthreads();
    return total;
}

// ---- literal port of libstdc++ std::sort on (key, pos) pairs; comp(a, b) = key[a] < key[b] --------------------
struct SortArr {
    unsigned long long* k;
    int* p;
    __device__ __forceinline__ bool less(int a, int b) const { return k[a] < k[b]; }
    __device__ __forceinline__ void swap(int a, int b) {
        const unsigned long long tk = k[a]; k[a] = k[b]; k[b] = tk;
        const int tp = p[a]; p[a] = p[b]; p[b] = tp;
    }
    __device__ __forceinline__ void move(int dst, int src) { k[dst] = k[src]; p[dst] = p[src]; }
};

__device__ void std_adjust_heap(SortArr& a, int first, int hole, int len, unsigned long long vk, int vp) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (a.less(first + child, first + child - 1)) child--;
        a.move(first + hole, first + child);
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        a.move(first + hole, first + child - 1);
        hole = child - 1;
    }
    int parent = (hole - 1) / 2;    // __push_heap
    while (hole > top && a.k[first + parent] < vk) {
        a.move(first + hole, first + parent);
        hole = parent;
        parent = (hole - 1) / 2;
    }
    a.k[first + hole] = vk;
    a.p[first + hole] = vp;
}

__device__ void std_heap_sort(SortArr& a, int first, int last) {    // __partial_sort(first, last, last)
    const int len = last - first;
    if (len >= 2) {
        int parent = (len - 2) / 2;
        while (true) {
            std_adjust_heap(a, first, parent, len, a.k[first + parent], a.p[first + parent]);
            if (parent == 0) break;
            parent--;
        }
    }
    while (last - first > 1) {
        --last;
        const unsigned long long vk = a.k[last];
        const int vp = a.p[last];
        a.move(last, first);
        std_adjust_heap(a, first, 0, last - first, vk, vp);
    }
}

__device__ void std_unguarded_linear_insert(SortArr& a, int last) {
    const unsigned long long vk = a.k[last];
    const int vp = a.p[last];
    int next = last - 1;
    while (vk < a.k[next]) {
        a.move(last, next);
        last = next;
        --next;
    }
    a.k[last] = vk;
    a.p[last] = vp;
}

__device__ void std_insertion_sort(SortArr& a, int first, int last) {
    if (first == last) return;
    for (int i = first + 1; i != last; ++i) {
        if (a.less(i, first)) {
            const unsigned long long vk = a.k[i];
            const int vp = a.p[i];
            for (int j = i; j > first; --j) a.move(j, j - 1);    // move_backward(first, i, i + 1)
            a.k[first] = vk;
            a.p[first] = vp;
        } else {
            std_unguarded_linear_insert(a, i);
        }
    }
}

__device__ void std_sort(SortArr& a, int n) {
    if (n <= 0) return;
    // __introsort_loop with an explicit stack for the recursion on the right part
    int stFirst[64], stLast[64], stDepth[64];
    int sp = 0;
    int lg = 0;
    for (int v = n; v > 1; v >>= 1) lg++;
    stFirst[sp] = 0; stLast[sp] = n; stDepth[sp] = 2 * lg; sp++;
    while (sp > 0) {
        --sp;
        int first = stFirst[sp], last = stLast[sp], depth = stDepth[sp];
        // libstdc++ recurses on [cut, last) FIRST and then loops on [first, cut).  The two ranges are disjoint, so
        // the order in which they are processed does not change the result; only the sequence of operations inside
        // a range does.
        while (last - first > 16) {
            if (depth == 0) {
                std_heap_sort(a, first, last);
                break;
            }
            --depth;
            const int mid = first + (last - first) / 2;
            {    // __move_median_to_first(first, first + 1, mid, last - 1)
                const int A = first + 1, B = mid, C = last - 1;
                if (a.less(A, B)) {
                    if (a.less(B, C)) a.swap(first, B);
                    else if (a.less(A, C)) a.swap(first, C);
                    else a.swap(first, A);
                } else if (a.less(A, C)) a.swap(first, A);
                else if (a.less(B, C)) a.swap(first, C);
                else a.swap(first, B);
            }
            int f = first + 1, l = last;    // __unguarded_partition(first + 1, last, first)
            while (true) {
                while (a.less(f, first)) ++f;
                --l;
                while (a.less(first, l)) --l;
                if (!(f < l)) break;
                a.swap(f, l);
                ++f;
            }
            const int cut = f;
            if (sp < 64) { stFirst[sp] = cut; stLast[sp] = last; stDepth[sp] = depth; sp++; }
            last = cut;
        }
    }
    // __final_insertion_sort
    if (n > 16) {
        std_insertion_sort(a, 0, 16);
        for (int i = 16; i < n; ++i) std_unguarded_linear_insert(a, i);
    } else {
        std_insertion_sort(a, 0, n);
    }
}

// ---- the same std::sort, computed by a warp + the block --------------------------------------------------------------------
// The serial port above costs ~40 % of the kernel (255 threads wait for lane 0).  Two observations make it parallel
// without changing a single result:
//  (1) __unguarded_partition is a Hoare partition whose swaps are known in advance: with A = the positions (ascending) whose
//      key is not less than the pivot and B = the positions (descending) whose key is not greater, the k-th swap exchanges
//      A_k and B_k as long as A_k < B_k (positions changed by earlier swaps lie behind the two scan pointers), and the
//      returned cut is min(A_{s+1}, B_s) after s swaps (the left scan stops at the next original stopper or at the last
//      swapped right position, which now holds a key >= pivot).  A warp builds both lists with ballots, swaps in parallel.
//      The median-of-3 step and the (rare) heapsort at the depth limit stay literal, on lane 0.
//  (2) __final_insertion_sort never moves an element past an equal one, so it is a STABLE sort of the array the introsort
//      loop leaves; that array consists of unsorted runs of <= 16 elements in sorted order (heapsorted runs are sorted), so
//      the final position of element i is i - #{j in [i-15, i): key_j > key_i} + #{j in (i, i+15]: key_j < key_i}: every
//      thread of the block computes one.
__device__ void warp_introsort_loop(SortArr& a, int n, int* listA, int* listB, int* stk) {
    const int lane = threadIdx.x & 31;
    const uint32_t ltMask = (1u << lane) - 1u;
    if (n <= 16) return;
    int lg = 0;
    for (int v = n; v > 1; v >>= 1) lg++;
    int sp = 0;
    if (lane == 0) { stk[0] = 0; stk[1] = n; stk[2] = 2 * lg; }
    sp = 1;
    __syncwarp();
    while (sp > 0) {
        --sp;
        int first = stk[3 * sp], last = stk[3 * sp + 1], depth = stk[3 * sp + 2];
        __syncwarp();
        while (last - first > 16) {
            if (depth == 0) {
                if (lane == 0) std_heap_sort(a, first, last);
                __syncwarp();
                break;
            }
            --depth;
            if (lane == 0) {    // __move_median_to_first(first, first + 1, mid, last - 1)
                const int A = first + 1, B = first + (last - first) / 2, C = last - 1;
                if (a.less(A, B)) {
                    if (a.less(B, C)) a.swap(first, B);
                    else if (a.less(A, C)) a.swap(first, C);
                    else a.swap(first, A);
                } else if (a.less(A, C)) a.swap(first, A);
                else if (a.less(B, C)) a.swap(first, C);
                else a.swap(first, B);
            }
            __syncwarp();
            const unsigned long long pivot = a.k[first];
            int cntA = 0, cntB = 0;
            for (int base = first + 1; base < last; base += 32) {    // A: ascending positions with !(key < pivot)
                const int p = base + lane;
                const bool geq = p < last && !(a.k[p] < pivot);
                const uint32_t m = __ballot_sync(0xffffffffu, geq);
                if (geq) listA[cntA + __popc(m & ltMask)] = p;
                cntA += __popc(m);
            }
            for (int base = last - 1; base > first; base -= 32) {    // B: descending positions with !(pivot < key)
                const int p = base - lane;
                const bool leq = p > first && !(pivot < a.k[p]);
                const uint32_t m = __ballot_sync(0xffffffffu, leq);
                if (leq) listB[cntB + __popc(m & ltMask)] = p;
                cntB += __popc(m);
            }
            __syncwarp();
            const int pairs = min(cntA, cntB);
            int sw = 0;
            for (int kb = 0; kb < pairs; kb += 32) {
                const int k = kb + lane;
                const bool ok = k < pairs && listA[k] < listB[k];
                const uint32_t m = __ballot_sync(0xffffffffu, ok);
                if (ok) a.swap(listA[k], listB[k]);
                sw += __popc(m);
                if (m != 0xffffffffu) break;
            }
            __syncwarp();
            const int nextA = sw < cntA ? listA[sw] : 0x7fffffff, prevB = sw > 0 ? listB[sw - 1] : 0x7fffffff;
            const int cut = min(nextA, prevB);
            __syncwarp();    // the lists are rewritten by the next partition
            if (lane == 0 && sp < 64) { stk[3 * sp] = cut; stk[3 * sp + 1] = last; stk[3 * sp + 2] = depth; }
            if (sp < 64) sp++;
            last = cut;
        }
        __syncwarp();
    }
}

__device__ __forceinline__ int key_quadrant(uint32_t key, int mx, int my) {
    const int x = key & 0xfff, y = (key >> 12) & 0xfff;
    return (x < mx ? 0 : 1) | (y < my ? 0 : 2);    // 0 = n1 (UL), 1 = n2 (UR), 2 = n3 (BL), 3 = n4 (BR), :516-526
}

// One (image, level) list with n > 0 candidates.  KEYS_SM: the two key buffers fit the shared-memory area behind the node
// arrays (n <= keySmem, the normal case) -- a template parameter so that every key access is an LDS / STS; with the buffer
// chosen at run time the pointers were generic and every access a generic LD / ST.  Otherwise the keys live in the
// per-image global scratch.
template <bool KEYS_SM>
__device__ __forceinline__ void octree_list(const Geom& g, const ExtractBuffers& b, int nodeCap, int keySmem, int n) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int wsum[2 * OT_WARPS + 2];
    __shared__ int sK, sTotal;
    __shared__ int sortStack[3 * 64];

    const int level = blockIdx.y, img = blockIdx.x;
    const LevelGeom& L = g.lv[level];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t ltMask = (1u << lane) - 1u;
    const int N = L.nfeat;

    OtSmem s;
    {
        uint8_t* p = smem;
        s.cur = reinterpret_cast<OtNode*>(p); p += sizeof(OtNode) * nodeCap;
        s.nxt = reinterpret_cast<OtNode*>(p); p += sizeof(OtNode) * nodeCap;
        s.ekey = reinterpret_cast<unsigned long long*>(p); p += 8 * nodeCap;
        s.c4 = reinterpret_cast<int*>(p); p += 16 * nodeCap;
        s.epos = reinterpret_cast<int*>(p); p += 4 * nodeCap;
        s.P = reinterpret_cast<int*>(p); p += 4 * nodeCap;
        s.tA = reinterpret_cast<int*>(p); p += 4 * (nodeCap + 1);
        s.tB = reinterpret_cast<int*>(p); p += 4 * (nodeCap + 1);
        s.tC = reinterpret_cast<int*>(p);
    }
    uint32_t* kS = reinterpret_cast<uint32_t*>(smem + octree_node_bytes(nodeCap));    // key buffers (when they fit)

    int* levelCount = b.levelCount + (size_t)img * g.nlevels + level;
    uint32_t* outKeys = b.levelKeys + (size_t)img * g.keyTotal + L.keyBase;
    const int nCellsL = L.nCols * L.nRows;
    const int* cellCnt = b.cellCount + (size_t)img * g.nCells + L.cellBase;
    const uint32_t* cellSlots = b.slots + ((size_t)img * g.nCells + L.cellBase) * g.slotCap;

    // two plain pointers and selects (not an indexed array): the address space of both survives to every access
    uint32_t *kb0, *kb1;
    if (KEYS_SM) {
        kb0 = kS;
        kb1 = kS + keySmem;
    } else {
        uint32_t* base = b.scratch + (size_t)img * b.scratchPerImage;
        unsigned long long off = 0;
        for (int l = 0; l < level; l++) off += 2ull * (unsigned long long)g.lv[l].maxCand;
        kb0 = base + off;
        kb1 = kb0 + L.maxCand;
    }

    // ---- gather the candidates in the reference's insertion order (cell row, cell column, raster) --------------
    {
        int running = 0;
        for (int c0 = 0; c0 < nCellsL; c0 += OT_THREADS) {
            const int c = c0 + tid;
            s.tA[tid] = c < nCellsL ? cellCnt[c] : 0;
            __syncthreads();
            const int tot = block_excl_scan(s.tA, OT_THREADS, wsum);
            const int m = min(OT_THREADS, nCellsL - c0);
            // four cells per step: their counts, then their first 32 candidates, are in flight together (a warp walks ~34
            // cells of a level-0 list; one dependent global round trip per cell was 8 % of the kernel's stall samples)
            for (int k4 = warp; k4 < m; k4 += 4 * OT_WARPS) {
                int cnt[4];
                uint32_t v[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int k = k4 + j * OT_WARPS;
                    cnt[j] = k < m ? cellCnt[c0 + k] : 0;
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int k = k4 + j * OT_WARPS;
                    v[j] = lane < cnt[j] ? cellSlots[(size_t)(c0 + k) * g.slotCap + lane] : 0u;
                }
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int k = k4 + j * OT_WARPS;
                    if (cnt[j] == 0) continue;    // warp-uniform
                    uint32_t* dst = kb0 + running + s.tA[k];
                    if (lane < cnt[j]) dst[lane] = v[j];
                    const uint32_t* src = cellSlots + (size_t)(c0 + k) * g.slotCap;
                    for (int i = lane + 32; i < cnt[j]; i += 32) dst[i] = src[i];
                }
            }
            running += tot;
            __syncthreads();
        }
    }

    // ---- root nodes (:561-603) ---------------------------------------------------------------------------------
    const int nIni = L.nIni;
    const float hX = L.hX;
    int size = 0;
    if (nIni == 1) {
        if (tid == 0) s.cur[0] = OtNode{0, 0, (short)(int)(hX * 1.0f), (short)(L.h - 2 * kMinBorder), 0u, n};
        size = 1;
    } else {
        // stable partition of all keys by root index (int)(x / hX) (:587), one root after the other
        if (warp == 0) {
            int pos = 0;
            for (int r = 0; r < nIni; r++) {
                const int start = pos;
                for (int base = 0; base < n; base += 32) {
                    const int i = base + lane;
                    uint32_t key = 0;
                    bool mine = false;
                    if (i < n) {
                        key = kb0[i];
                        int ri = (int)(__fdiv_rn((float)(key & 0xfff), hX));
                        ri = min(ri, nIni - 1);
                        mine = ri == r;
                    }
                    const uint32_t m = __ballot_sync(0xffffffffu, mine);
                    if (mine) kb1[pos + __popc(m & ltMask)] = key;
                    pos += __popc(m);
                }
                if (lane == 0) {
                    s.tA[r] = start;
                    s.tB[r] = pos - start;
                }
            }
        }
        __syncthreads();
        if (tid == 0) {
            int sz = 0;
            for (int r = 0; r < nIni; r++) {
                if (s.tB[r] == 0) continue;    // empty roots are erased (:597)
                s.cur[sz++] = OtNode{(short)(int)(hX * (float)r), 0, (short)(int)(hX * (float)(r + 1)),
                                     (short)(L.h - 2 * kMinBorder), (unsigned)s.tA[r] | 0x80000000u, s.tB[r]};
            }
            sTotal = sz;
        }
        __syncthreads();
        size = sTotal;
        __syncthreads();
    }
    __syncthreads();

    // ---- split passes ----------------------------------------------------------------------------------------
    int phase = 1, ne = 0;
    bool overflow = false;
    while (true) {
        const int prevSize = size;
        int np;
        if (phase == 1) {
            // P = nodes with more than one key, in list order (:631-636)
            for (int i = tid; i < size; i += OT_THREADS) s.tA[i] = s.cur[i].cnt > 1 ? 1 : 0;
            __syncthreads();
            for (int i = tid; i < size; i += OT_THREADS) s.tB[i] = s.tA[i];
            __syncthreads();
            np = block_excl_scan(s.tA, size, wsum);
            for (int i = tid; i < size; i += OT_THREADS)
                if (s.tB[i]) s.P[s.tA[i]] = i;
        } else {
            // std::sort(compareNodes) of the expandable nodes, then back to front (:700-705)
            np = ne;
#ifdef ORBFE_OT_SERIAL_SORT
            if (tid == 0) {
                SortArr a{s.ekey, s.epos};
                std_sort(a, ne);
            }
            __syncthreads();
            for (int k = tid; k < np; k += OT_THREADS) s.P[k] = s.epos[np - 1 - k];
#else
            if (warp == 0) {
                SortArr a{s.ekey, s.epos};
                warp_introsort_loop(a, ne, s.tA, s.tB, sortStack);
            }
            __syncthreads();
            for (int i = tid; i < ne; i += OT_THREADS) {    // the final insertion sort as a stable rank (see above)
                const unsigned long long ki = s.ekey[i];
                int rank = i;
                for (int j = max(0, i - 15); j < i; j++) rank -= s.ekey[j] > ki;
                for (int j = i + 1; j <= min(ne - 1, i + 15); j++) rank += s.ekey[j] < ki;
                s.P[np - 1 - rank] = s.epos[i];
            }
#endif
        }
        __syncthreads();

        // quadrant counts of every node to split (DivideNode :484-526)
        for (int k = warp; k < np; k += OT_WARPS) {
            const OtNode nd = s.cur[s.P[k]];
            const int mx = nd.x0 + ((nd.x1 - nd.x0 + 1) >> 1), my = nd.y0 + ((nd.y1 - nd.y0 + 1) >> 1);
            const uint32_t* src = ((nd.start >> 31) ? kb1 : kb0) + (nd.start & 0x7fffffffu);
            int c0 = 0, c1 = 0, c2 = 0, c3 = 0;
            for (int i = lane; i < nd.cnt; i += 32) {
                const int q = key_quadrant(src[i], mx, my);
                c0 += q == 0; c1 += q == 1; c2 += q == 2; c3 += q == 3;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                c0 += __shfl_xor_sync(0xffffffffu, c0, o);
                c1 += __shfl_xor_sync(0xffffffffu, c1, o);
                c2 += __shfl_xor_sync(0xffffffffu, c2, o);
                c3 += __shfl_xor_sync(0xffffffffu, c3, o);
            }
            if (lane == 0) {
                s.c4[4 * k] = c0; s.c4[4 * k + 1] = c1; s.c4[4 * k + 2] = c2; s.c4[4 * k + 3] = c3;
                s.tA[k] = (c0 > 0) + (c1 > 0) + (c2 > 0) + (c3 > 0);
            }
        }
        __syncthreads();
        for (int k = tid; k < np; k += OT_THREADS) s.tB[k] = s.tA[k];    // tB = non-empty children m_k
        if (tid == 0) sK = np - 1;
        __syncthreads();
        block_excl_scan(s.tA, np, wsum);    // tA = creation index of the first child of P[k]
        if (phase == 2) {
            // the loop of :703-750 stops after the first split that brings the list to N nodes
            for (int k = tid; k < np; k += OT_THREADS)
                if (prevSize + s.tA[k] + s.tB[k] - (k + 1) >= N) atomicMin(&sK, k);
        }
        __syncthreads();
        const int K = sK;
        const int M = np > 0 ? s.tA[K] + s.tB[K] : 0;
        const int newSize = prevSize + M - (np > 0 ? K + 1 : 0);
        if (newSize > nodeCap) { overflow = true; break; }

        // split: stable 4-way partition into the other key buffer, children to list positions M-1-creationIndex
        for (int i = tid; i < prevSize; i += OT_THREADS) s.tC[i] = 1;    // 1 = node survives this pass
        for (int i = tid; i < M; i += OT_THREADS) s.epos[i] = 0;         // reused: 1 = child is expandable
        __syncthreads();
        for (int k = warp; k <= K && k < np; k += OT_WARPS) {
            const int li = s.P[k];
            const OtNode nd = s.cur[li];
            const int hx = (nd.x1 - nd.x0 + 1) >> 1, hy = (nd.y1 - nd.y0 + 1) >> 1;    // ceil(float(d) / 2)
            const int mx = nd.x0 + hx, my = nd.y0 + hy;
            const unsigned sb = nd.start >> 31, st = nd.start & 0x7fffffffu;
            const uint32_t* src = (sb ? kb1 : kb0) + st;
            uint32_t* dst = (sb ? kb0 : kb1) + st;
            const int c0 = s.c4[4 * k], c1 = s.c4[4 * k + 1], c2 = s.c4[4 * k + 2], c3 = s.c4[4 * k + 3];
            int o0 = 0, o1 = c0, o2 = c0 + c1, o3 = c0 + c1 + c2;
            for (int base = 0; base < nd.cnt; base += 32) {
                const int i = base + lane;
                uint32_t key = 0;
                int q = -1;
                if (i < nd.cnt) {
                    key = src[i];
                    q = key_quadrant(key, mx, my);
                }
                const uint32_t m0 = __ballot_sync(0xffffffffu, q == 0), m1 = __ballot_sync(0xffffffffu, q == 1);
                const uint32_t m2 = __ballot_sync(0xffffffffu, q == 2), m3 = __ballot_sync(0xffffffffu, q == 3);
                if (q == 0) dst[o0 + __popc(m0 & ltMask)] = key;
                else if (q == 1) dst[o1 + __popc(m1 & ltMask)] = key;
                else if (q == 2) dst[o2 + __popc(m2 & ltMask)] = key;
                else if (q == 3) dst[o3 + __popc(m3 & ltMask)] = key;
                o0 += __popc(m0); o1 += __popc(m1); o2 += __popc(m2); o3 += __popc(m3);
            }
            if (lane == 0) {
                s.tC[li] = 0;
                int ci = s.tA[k];
                const unsigned nb = (sb ^ 1u) << 31;
                const int cc[4] = {c0, c1, c2, c3};
                const int cs[4] = {0, c0, c0 + c1, c0 + c1 + c2};
                // n1 = [x0,mx)x[y0,my), n2 = [mx,x1)x[y0,my), n3 = [x0,mx)x[my,y1), n4 = [mx,x1)x[my,y1)  (:487-510)
                const short rx0[4] = {nd.x0, (short)mx, nd.x0, (short)mx}, rx1[4] = {(short)mx, nd.x1, (short)mx, nd.x1};
                const short ry0[4] = {nd.y0, nd.y0, (short)my, (short)my}, ry1[4] = {(short)my, (short)my, nd.y1, nd.y1};
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    if (cc[q] == 0) continue;
                    s.nxt[M - 1 - ci] = OtNode{rx0[q], ry0[q], rx1[q], ry1[q], (st + cs[q]) | nb, cc[q]};
                    s.epos[ci] = cc[q] > 1 ? 1 : 0;
                    ci++;
                }
            }
        }
        __syncthreads();
        // survivors keep their relative order behind the new children
        for (int i = tid; i < prevSize; i += OT_THREADS) s.tB[i] = s.tC[i];
        __syncthreads();
        block_excl_scan(s.tC, prevSize, wsum);
        for (int i = tid; i < prevSize; i += OT_THREADS)
            if (s.tB[i]) s.nxt[M + s.tC[i]] = s.cur[i];
        // expandable children in creation order -> (count, UL.x) sort keys for a later phase-2 pass (:651-678)
        for (int i = tid; i < M; i += OT_THREADS) s.tA[i] = s.epos[i];
        __syncthreads();
        for (int i = tid; i < M; i += OT_THREADS) s.tC[i] = s.tA[i];
        __syncthreads();
        const int neNew = block_excl_scan(s.tC, M, wsum);
        for (int ci = tid; ci < M; ci += OT_THREADS)
            if (s.tA[ci]) {
                const int pos = M - 1 - ci;
                const OtNode c = s.nxt[pos];
                s.ekey[s.tC[ci]] = ((unsigned long long)(unsigned)c.cnt << 32) | (unsigned)(int)c.x0;
                s.epos[s.tC[ci]] = pos;
            }
        __syncthreads();
        // (epos served as the flag array above; its flags were copied to tA before entries [0, neNew) were
        // overwritten with list positions)
        {
            OtNode* t = s.cur; s.cur = s.nxt; s.nxt = t;
        }
        size = newSize;
        ne = neNew;
        if (size >= N || size == prevSize) break;                // :687, :753
        if (phase == 1 && size + 3 * ne > N) phase = 2;          // :691
        __syncthreads();
    }
    __syncthreads();

    if (overflow || size > L.keyCap) {
        if (tid == 0) {
            atomicExch(b.errFlag, 1);
            *levelCount = 0;
        }
        return;
    }

    // ---- keep the first key with the largest response of every node, in list order (:760-778) ----------------
    for (int i = warp; i < size; i += OT_WARPS) {
        const OtNode nd = s.cur[i];
        const uint32_t* src = ((nd.start >> 31) ? kb1 : kb0) + (nd.start & 0x7fffffffu);
        // maximise (response, -index): pack as response << 24 | (0xffffff - index) when the segment is short,
        // otherwise compare explicitly
        int bestR = -1, bestI = 0x7fffffff;
        for (int k = lane; k < nd.cnt; k += 32) {
            const int r = (int)(src[k] >> 24);
            if (r > bestR) { bestR = r; bestI = k; }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const int r2 = __shfl_xor_sync(0xffffffffu, bestR, o), i2 = __shfl_xor_sync(0xffffffffu, bestI, o);
            if (r2 > bestR || (r2 == bestR && i2 < bestI)) { bestR = r2; bestI = i2; }
        }
        if (lane == 0) outKeys[i] = src[bestI];
    }
    if (tid == 0) *levelCount = size;
}

// grid (n, nlevels); block OT_THREADS; dynamic shared memory laid out by the host (octree_smem_bytes).
__global__ void __launch_bounds__(OT_THREADS, 5) octree_kernel(const __grid_constant__ Geom g, const ExtractBuffers b,
                                                            int nodeCap, int keySmem) {
    __shared__ int wcount[OT_WARPS];
    const int level = blockIdx.y, img = blockIdx.x;
    const LevelGeom& L = g.lv[level];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // ---- total number of candidates -------------------------------------------------------------------------
    const int nCellsL = L.nCols * L.nRows;
    const int* cellCnt = b.cellCount + (size_t)img * g.nCells + L.cellBase;
    int part = 0;
    for (int c = tid; c < nCellsL; c += OT_THREADS) part += cellCnt[c];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) wcount[warp] = part;
    __syncthreads();
    int n = 0;
#pragma unroll
    for (int w = 0; w < OT_WARPS; w++) n += wcount[w];
    if (n == 0 || nCellsL == 0) {    // empty list stays empty (the loop of :624 ends at once with size == prevSize)
        if (tid == 0) b.levelCount[(size_t)img * g.nlevels + level] = 0;
        return;
    }
    if (n <= keySmem) octree_list<true>(g, b, nodeCap, keySmem, n);    // block-uniform
    else octree_list<false>(g, b, nodeCap, keySmem, n);
}

size_t octree_smem_bytes(int nodeCap, int keySmem) { return octree_node_bytes(nodeCap) + 8 * (size_t)keySmem; }

int launch_octree(const Geom& g, int n, const ExtractBuffers& b, int nodeCap, int keySmem, cudaStream_t st) {
    const size_t smem = octree_smem_bytes(nodeCap, keySmem);
    static SmemOptIn optIn;
    smem_opt_in(optIn, octree_kernel, smem);
    // level-major block order: the lists of level 0 (the most candidates, the longest CTAs) start first and the short
    // lists of the top levels fill the tail of the last wave
    octree_kernel<<<dim3(n, g.nlevels), OT_THREADS, smem, st>>>(g, b, nodeCap, keySmem);
    return 1;
}

}    // namespace orbfe
