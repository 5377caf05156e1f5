// k_search_bow.cu — ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches)
// (ORBmatcher.cc:224-425): the matcher of Tracking::TrackReferenceKeyFrame (Tracking.cc:2788) and of relocalisation
// (Tracking.cc:3709).
//
// Reference: the two DBoW2 feature vectors (std::map<NodeId, vector<feature index>>) are merged on equal node ids; inside
// a common node every key-frame feature that has a good MapPoint is compared with the frame's features of that node that
// are still unmatched; best and second-best distance, TH_LOW = 50 and the ratio test; stereo-camera frames keep a second
// best / second-best pair for the right-camera slots, accepted INSIDE the left acceptance branch and with `|| true` on its
// ratio (:355-357, reproduced); rotation histogram + ComputeThreeMaxima, losing bins are cleared (:400-420).
//
// A feature index sits in exactly one node, so nodes are independent; inside a node the key-frame features are
// sequential (a frame feature taken by an earlier one is skipped).  Mapping: one CTA per (key frame, frame) pair, one
// warp per common node (binary search of the node id in the frame's node list), the warp walks the node's key-frame
// features in order, lanes stride over the frame's features of the node: 256-bit XOR + POPC, best = first minimum in
// vector order = min (distance, position), second = second smallest distance with multiplicity, merged by shuffles.
// Feature vectors come as CSR arrays (node ids ascending, start[nnodes + 1], indices in vector order).
#include "orbfe_internal.h"

namespace orbfe {

constexpr int SB_THREADS = 256;
constexpr int SB_WARPS = SB_THREADS / 32;
constexpr int SB_TH_LOW = 50, SB_HISTO = 30;

struct BowParams {
    int cap, nodeCap;    // per-pair strides of the feature arrays and of the node arrays (start: nodeCap + 1)
    float nnRatio;
    int checkOrientation;
};

// (best key = dist << 16 | pos, second distance) of this lane merged over the warp
__device__ __forceinline__ void sb_merge(unsigned& k1, int& d2) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned ok = __shfl_xor_sync(0xffffffffu, k1, o);
        const int od = __shfl_xor_sync(0xffffffffu, d2, o);
        const unsigned loser = ok > k1 ? ok : k1;
        d2 = min(min(d2, od), (int)(loser >> 16));
        k1 = ok < k1 ? ok : k1;
    }
}

// grid: npairs; block SB_THREADS; dynamic shared memory: cap bytes (histogram bin per frame slot)
__global__ void __launch_bounds__(SB_THREADS) search_bow_kernel(
    const __grid_constant__ BowParams P, const uint8_t* __restrict__ descKFAll, const float* __restrict__ angleKFAll,
    const uint8_t* __restrict__ validKFAll, const int* __restrict__ kfIdsAll, const int* __restrict__ kfStartAll,
    const int* __restrict__ kfIdxAll, const int* __restrict__ kfNodes, const uint8_t* __restrict__ descFAll,
    const float* __restrict__ angleFAll, const int* __restrict__ nF, const int* __restrict__ nleftF,
    const int* __restrict__ fIdsAll, const int* __restrict__ fStartAll, const int* __restrict__ fIdxAll,
    const int* __restrict__ fNodes, int* __restrict__ matchesAll, int* __restrict__ nMatches) {
    extern __shared__ __align__(16) uint8_t binOf[];    // cap entries, 0xff = no histogram entry
    __shared__ int hist[32];
    __shared__ int sCount, sInd[3];
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint8_t* descKF = descKFAll + (size_t)f * P.cap * 32;
    const float* angleKF = angleKFAll + (size_t)f * P.cap;
    const uint8_t* validKF = validKFAll + (size_t)f * P.cap;
    const int* kfIds = kfIdsAll + (size_t)f * P.nodeCap;
    const int* kfStart = kfStartAll + (size_t)f * (P.nodeCap + 1);
    const int* kfIdx = kfIdxAll + (size_t)f * P.cap;
    const uint8_t* descF = descFAll + (size_t)f * P.cap * 32;
    const float* angleF = angleFAll + (size_t)f * P.cap;
    const int* fIds = fIdsAll + (size_t)f * P.nodeCap;
    const int* fStart = fStartAll + (size_t)f * (P.nodeCap + 1);
    const int* fIdx = fIdxAll + (size_t)f * P.cap;
    volatile int* matches = matchesAll + (size_t)f * P.cap;
    const int n = min(nF[f], P.cap);
    const int nk = min(kfNodes[f], P.nodeCap), nfn = min(fNodes[f], P.nodeCap);
    const int nl = nleftF ? nleftF[f] : -1;
    const int nleft = nl < 0 ? 0x7fffffff : nl;    // Nleft == -1: every slot is a "left" slot

    for (int i = tid; i < P.cap; i += SB_THREADS) binOf[i] = 0xffu;
    for (int i = tid; i < n; i += SB_THREADS) matches[i] = -1;    // vpMapPointMatches = vector<MapPoint*>(F.N, NULL)
    if (tid < 32) hist[tid] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();

    for (int jn = warp; jn < nk; jn += SB_WARPS) {
        const int id = kfIds[jn];
        int lo = 0, hi = nfn;    // lower_bound of id in the frame's node ids
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (fIds[mid] < id) lo = mid + 1; else hi = mid;
        }
        if (lo >= nfn || fIds[lo] != id) continue;
        const int fs = fStart[lo], fe = fStart[lo + 1];
        for (int iKF = kfStart[jn]; iKF < kfStart[jn + 1]; iKF++) {
            const int realIdxKF = kfIdx[iKF];
            if (!validKF[realIdxKF]) continue;    // !pMP || pMP->isBad()
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(descKF + (size_t)realIdxKF * 32));
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(descKF + (size_t)realIdxKF * 32) + 1);
            constexpr unsigned NONE = 256u << 16;
            unsigned k1 = NONE, k1R = NONE;
            int d2 = 256, d2R = 256;
            for (int t = fs + lane; t < fe; t += 32) {
                const int idxF = fIdx[t];
                if (matches[idxF] >= 0) continue;    // vpMapPointMatches[realIdxF]
                const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(descF + (size_t)idxF * 32));
                const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(descF + (size_t)idxF * 32) + 1);
                const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                              __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
                const unsigned key = ((unsigned)d << 16) | (unsigned)(t - fs);    // a node holds < 65536 features
                if (idxF < nleft) {
                    if (key < k1) { d2 = (int)(k1 >> 16); k1 = key; } else d2 = min(d2, d);
                } else {
                    if (key < k1R) { d2R = (int)(k1R >> 16); k1R = key; } else d2R = min(d2R, d);
                }
            }
            sb_merge(k1, d2);
            sb_merge(k1R, d2R);
            const int bestDist1 = (int)(k1 >> 16), bestDist1R = (int)(k1R >> 16);
            if (bestDist1 <= SB_TH_LOW) {    // warp-uniform
                if (lane == 0) {
                    const float angK = angleKF[realIdxKF];
                    auto push = [&](int slot) {
                        if (!P.checkOrientation) return;
                        float rot = angK - angleF[slot];
                        if (rot < 0.0f) rot += 360.0f;
                        int bin = (int)roundf(rot * (1.0f / SB_HISTO));
                        if (bin == SB_HISTO) bin = 0;
                        bin = min(max(bin, 0), SB_HISTO - 1);
                        binOf[slot] = (uint8_t)bin;
                        atomicAdd(&hist[bin], 1);
                    };
                    int got = 0;
                    if ((float)bestDist1 < P.nnRatio * (float)d2) {
                        const int slot = fIdx[fs + (int)(k1 & 0xffffu)];
                        matches[slot] = realIdxKF;
                        push(slot);
                        got++;
                    }
                    if (bestDist1R <= SB_TH_LOW) {    // ratio test `|| true` in the reference
                        const int slot = fIdx[fs + (int)(k1R & 0xffffu)];
                        matches[slot] = realIdxKF;
                        push(slot);
                        got++;
                    }
                    if (got) atomicAdd(&sCount, got);
                }
            }
            __syncwarp();    // the next key-frame feature of the node sees the slots just taken
        }
    }
    __syncthreads();
    if (P.checkOrientation) {
        if (tid == 0) {    // ComputeThreeMaxima (ORBmatcher.cc:2061-2102)
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int i = 0; i < SB_HISTO; i++) {
                const int sz = hist[i];
                if (sz > max1) {
                    max3 = max2; max2 = max1; max1 = sz;
                    ind3 = ind2; ind2 = ind1; ind1 = i;
                } else if (sz > max2) {
                    max3 = max2; max2 = sz;
                    ind3 = ind2; ind2 = i;
                } else if (sz > max3) {
                    max3 = sz;
                    ind3 = i;
                }
            }
            if ((float)max2 < 0.1f * (float)max1) {
                ind2 = -1;
                ind3 = -1;
            } else if ((float)max3 < 0.1f * (float)max1) {
                ind3 = -1;
            }
            sInd[0] = ind1;
            sInd[1] = ind2;
            sInd[2] = ind3;
        }
        __syncthreads();
        int removed = 0;
        for (int i = tid; i < n; i += SB_THREADS) {
            const int b = binOf[i];
            if (b != 0xff && b != sInd[0] && b != sInd[1] && b != sInd[2]) {
                matches[i] = -1;
                removed++;
            }
        }
        if (removed) atomicSub(&sCount, removed);
        __syncthreads();
    }
    if (tid == 0 && nMatches) nMatches[f] = sCount;
}

size_t search_bow_smem_bytes(int cap) { return ((size_t)cap + 15) & ~(size_t)15; }

int launch_search_bow(const uint8_t* descKF, const float* angleKF, const uint8_t* validKF, const int* kfIds, const int* kfStart,
                      const int* kfIdx, const int* kfNodes, const uint8_t* descF, const float* angleF, const int* nF,
                      const int* nleftF, const int* fIds, const int* fStart, const int* fIdx, const int* fNodes, int cap,
                      int nodeCap, int npairs, float nnRatio, int checkOrientation, int* matches, int* nMatches,
                      cudaStream_t st) {
    if (npairs <= 0) return 0;
    BowParams P;
    P.cap = cap;
    P.nodeCap = nodeCap;
    P.nnRatio = nnRatio;
    P.checkOrientation = checkOrientation;
    const size_t smem = search_bow_smem_bytes(cap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, search_bow_kernel, smem);
    search_bow_kernel<<<npairs, SB_THREADS, smem, st>>>(P, descKF, angleKF, validKF, kfIds, kfStart, kfIdx, kfNodes, descF, angleF,
                                                        nF, nleftF, fIds, fStart, fIdx, fNodes, matches, nMatches);
    return 1;
}

}    // namespace orbfe
