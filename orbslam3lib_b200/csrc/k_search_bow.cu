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
#include "kb8.cuh"

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

// ComputeThreeMaxima (ORBmatcher.cc:2061-2102) on the bin counts
__device__ void sb_three_maxima(const int* hist, int* ind) {
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
    ind[0] = ind1;
    ind[1] = ind2;
    ind[2] = ind3;
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
        if (tid == 0) sb_three_maxima(hist, sInd);
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

// ---- ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (ORBmatcher.cc:766-906) --------
// Loop detection / relocalisation between key frames.  Same node merge; inside a common node every KF1 feature with a good
// MapPoint looks at the KF2 features of the node that have a good MapPoint and are not in vbMatched2; best / second-best
// distance, accepted with bestDist1 < TH_LOW (strict) and the ratio test; the accepted KF2 feature enters vbMatched2.  The
// histogram stores idx1, the losing bins clear vpMatches12[idx1].  validN[i] = MapPoint i exists, is not bad, and (stereo-
// camera key frames) i < mvKeysUn.size() (:802-804, :821-823).
// grid: npairs; block SB_THREADS; dynamic shared memory: 2 * cap bytes (bin per KF1 feature, vbMatched2)
__global__ void __launch_bounds__(SB_THREADS) search_bow_kf_kernel(
    const __grid_constant__ BowParams P, const uint8_t* __restrict__ desc1All, const float* __restrict__ angle1All,
    const uint8_t* __restrict__ valid1All, const int* __restrict__ ids1All, const int* __restrict__ start1All,
    const int* __restrict__ idx1All, const int* __restrict__ nodes1, const uint8_t* __restrict__ desc2All,
    const float* __restrict__ angle2All, const uint8_t* __restrict__ valid2All, const int* __restrict__ ids2All,
    const int* __restrict__ start2All, const int* __restrict__ idx2All, const int* __restrict__ nodes2,
    int* __restrict__ matchesAll, int* __restrict__ nMatches) {
    extern __shared__ __align__(16) uint8_t binOf[];    // cap entries, 0xff = no histogram entry
    volatile uint8_t* matched2 = binOf + ((P.cap + 15) & ~15);
    __shared__ int hist[32];
    __shared__ int sCount, sInd[3];
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint8_t* desc1 = desc1All + (size_t)f * P.cap * 32;
    const float* angle1 = angle1All + (size_t)f * P.cap;
    const uint8_t* valid1 = valid1All + (size_t)f * P.cap;
    const int* ids1 = ids1All + (size_t)f * P.nodeCap;
    const int* start1 = start1All + (size_t)f * (P.nodeCap + 1);
    const int* idx1 = idx1All + (size_t)f * P.cap;
    const uint8_t* desc2 = desc2All + (size_t)f * P.cap * 32;
    const float* angle2 = angle2All + (size_t)f * P.cap;
    const uint8_t* valid2 = valid2All + (size_t)f * P.cap;
    const int* ids2 = ids2All + (size_t)f * P.nodeCap;
    const int* start2 = start2All + (size_t)f * (P.nodeCap + 1);
    const int* idx2 = idx2All + (size_t)f * P.cap;
    int* matches = matchesAll + (size_t)f * P.cap;
    const int nk = min(nodes1[f], P.nodeCap), nfn = min(nodes2[f], P.nodeCap);

    for (int i = tid; i < P.cap; i += SB_THREADS) {
        binOf[i] = 0xffu;
        matched2[i] = 0;
        matches[i] = -1;    // vpMatches12 = vector<MapPoint*>(vpMapPoints1.size(), NULL)
    }
    if (tid < 32) hist[tid] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();

    for (int jn = warp; jn < nk; jn += SB_WARPS) {
        const int id = ids1[jn];
        int lo = 0, hi = nfn;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (ids2[mid] < id) lo = mid + 1; else hi = mid;
        }
        if (lo >= nfn || ids2[lo] != id) continue;
        const int fs = start2[lo], fe = start2[lo + 1];
        for (int i1 = start1[jn]; i1 < start1[jn + 1]; i1++) {
            const int r1 = idx1[i1];
            if (r1 < 0 || r1 >= P.cap || !valid1[r1]) continue;
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(desc1 + (size_t)r1 * 32));
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(desc1 + (size_t)r1 * 32) + 1);
            constexpr unsigned NONE = 256u << 16;
            unsigned k1 = NONE;
            int d2 = 256;
            for (int t = fs + lane; t < fe; t += 32) {
                const int r2 = idx2[t];
                if (r2 < 0 || r2 >= P.cap || matched2[r2] || !valid2[r2]) continue;
                const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(desc2 + (size_t)r2 * 32));
                const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(desc2 + (size_t)r2 * 32) + 1);
                const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                              __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
                const unsigned key = ((unsigned)d << 16) | (unsigned)(t - fs);
                if (key < k1) { d2 = (int)(k1 >> 16); k1 = key; } else d2 = min(d2, d);
            }
            sb_merge(k1, d2);
            const int bestDist1 = (int)(k1 >> 16);
            if (lane == 0 && bestDist1 < SB_TH_LOW && (float)bestDist1 < P.nnRatio * (float)d2) {
                const int best2 = idx2[fs + (int)(k1 & 0xffffu)];
                matches[r1] = best2;
                matched2[best2] = 1;
                if (P.checkOrientation) {
                    float rot = angle1[r1] - angle2[best2];
                    if (rot < 0.0f) rot += 360.0f;
                    int bin = (int)roundf(rot * (1.0f / SB_HISTO));
                    if (bin == SB_HISTO) bin = 0;
                    bin = min(max(bin, 0), SB_HISTO - 1);
                    binOf[r1] = (uint8_t)bin;
                    atomicAdd(&hist[bin], 1);
                }
                atomicAdd(&sCount, 1);
            }
            __syncwarp();    // the next KF1 feature of the node sees vbMatched2
        }
    }
    __syncthreads();
    if (P.checkOrientation) {
        if (tid == 0) sb_three_maxima(hist, sInd);
        __syncthreads();
        int removed = 0;
        for (int i = tid; i < P.cap; i += SB_THREADS) {
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

// ---- ORBmatcher::SearchForTriangulation(pKF1, pKF2, vMatchedPairs, bOnlyStereo, bCoarse) (ORBmatcher.cc:908-1153) -----------
// Local mapping: matches between the features of two key frames that have NO MapPoint yet.  Node merge as above; inside a
// common node a KF1 feature looks at the KF2 features of the node, keeps dist <= TH_LOW && dist <= bestDist (the LAST of
// equal minima wins), skips features too close to the epipole (:1000-1008, only when neither side has a stereo measurement)
// and applies the epipolar constraint unless bCoarse (:1044; Pinhole::epipolarConstrain, Pinhole.cpp:102-124, with F12 from
// the caller).  vbMatched2 is never set in the reference, so the KF1 features are independent.  Rotation histogram on idx1.
// freeN[i] = feature i has no MapPoint (and, with bOnlyStereo, has a stereo measurement); stereoN[i] = !mpCamera2 && mvuRight[i] >= 0.
struct TriParams {
    int cap, nodeCap;
    int coarse, checkOrientation;
    float scale[ORBFE_MAX_LEVELS], sigma2[ORBFE_MAX_LEVELS];
    // two-camera (KannalaBrandt8) key frames (:1014-1043): mpCamera / mpCamera2 parameters; the features of a key frame are
    // its left keypoints followed by its right ones, the constraint is KannalaBrandt8::epipolarConstrain
    // (KannalaBrandt8.cpp:210-214: TriangulateMatches(...) > 0.0001) with the relative pose picked by the two sides
    int kb8;
    float camL[8], camR[8], precision;
};

__device__ __forceinline__ unsigned sb_min_u32(unsigned v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// grid: npairs; block SB_THREADS; dynamic shared memory: cap bytes (bin per KF1 feature)
__global__ void __launch_bounds__(SB_THREADS) search_triangulation_kernel(
    const __grid_constant__ TriParams P, const orbfe_keypoint* __restrict__ keys1All, const uint8_t* __restrict__ desc1All,
    const uint8_t* __restrict__ free1All, const uint8_t* __restrict__ stereo1All, const int* __restrict__ ids1All,
    const int* __restrict__ start1All, const int* __restrict__ idx1All, const int* __restrict__ nodes1,
    const orbfe_keypoint* __restrict__ keys2All, const uint8_t* __restrict__ desc2All, const uint8_t* __restrict__ free2All,
    const uint8_t* __restrict__ stereo2All, const int* __restrict__ ids2All, const int* __restrict__ start2All,
    const int* __restrict__ idx2All, const int* __restrict__ nodes2, const float* __restrict__ F12All,
    const float* __restrict__ epAll, const float* __restrict__ T12All, const int* __restrict__ nleft1Arr,
    const int* __restrict__ nleft2Arr, int* __restrict__ matchesAll, int* __restrict__ nMatches) {
    extern __shared__ __align__(16) uint8_t binOf[];
    __shared__ int hist[32];
    __shared__ int sCount, sInd[3];
    __shared__ float sF[9], sEp[2];
    __shared__ float sT[4][12];    // Tll, Tlr, Trl, Trr: R12 (row-major) and t12
    const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const orbfe_keypoint* keys1 = keys1All + (size_t)f * P.cap;
    const uint8_t* desc1 = desc1All + (size_t)f * P.cap * 32;
    const uint8_t* free1 = free1All + (size_t)f * P.cap;
    const uint8_t* stereo1 = stereo1All ? stereo1All + (size_t)f * P.cap : nullptr;
    const int* ids1 = ids1All + (size_t)f * P.nodeCap;
    const int* start1 = start1All + (size_t)f * (P.nodeCap + 1);
    const int* idx1 = idx1All + (size_t)f * P.cap;
    const orbfe_keypoint* keys2 = keys2All + (size_t)f * P.cap;
    const uint8_t* desc2 = desc2All + (size_t)f * P.cap * 32;
    const uint8_t* free2 = free2All + (size_t)f * P.cap;
    const uint8_t* stereo2 = stereo2All ? stereo2All + (size_t)f * P.cap : nullptr;
    const int* ids2 = ids2All + (size_t)f * P.nodeCap;
    const int* start2 = start2All + (size_t)f * (P.nodeCap + 1);
    const int* idx2 = idx2All + (size_t)f * P.cap;
    int* matches = matchesAll + (size_t)f * P.cap;
    const int nk = min(nodes1[f], P.nodeCap), nfn = min(nodes2[f], P.nodeCap);

    for (int i = tid; i < P.cap; i += SB_THREADS) {
        binOf[i] = 0xffu;
        matches[i] = -1;    // vMatches12(pKF1->N, -1)
    }
    if (tid < 32) hist[tid] = 0;
    if (tid < 9) sF[tid] = F12All ? F12All[(size_t)f * 9 + tid] : 0.f;
    if (tid < 2) sEp[tid] = epAll ? epAll[(size_t)f * 2 + tid] : 0.f;
    if (tid < 48) sT[tid / 12][tid % 12] = (P.kb8 && T12All) ? T12All[(size_t)f * 48 + tid] : 0.f;
    const int nleft1 = (P.kb8 && nleft1Arr) ? nleft1Arr[f] : 0x7fffffff, nleft2 = (P.kb8 && nleft2Arr) ? nleft2Arr[f] : 0x7fffffff;
    if (tid == 0) sCount = 0;
    __syncthreads();

    for (int jn = warp; jn < nk; jn += SB_WARPS) {
        const int id = ids1[jn];
        int lo = 0, hi = nfn;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (ids2[mid] < id) lo = mid + 1; else hi = mid;
        }
        if (lo >= nfn || ids2[lo] != id) continue;
        const int fs = start2[lo], fe = start2[lo + 1];
        for (int i1 = start1[jn]; i1 < start1[jn + 1]; i1++) {
            const int r1 = idx1[i1];
            if (r1 < 0 || r1 >= P.cap || !free1[r1]) continue;
            const bool bStereo1 = stereo1 && stereo1[r1];
            const orbfe_keypoint kp1 = keys1[r1];
            // epipolar line in the second image l = x1' F12 = [a b c] (Pinhole.cpp:110-112)
            const float a = kp1.x * sF[0] + kp1.y * sF[3] + sF[6];
            const float b = kp1.x * sF[1] + kp1.y * sF[4] + sF[7];
            const float c = kp1.x * sF[2] + kp1.y * sF[5] + sF[8];
            const float den = a * a + b * b;
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(desc1 + (size_t)r1 * 32));
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(desc1 + (size_t)r1 * 32) + 1);
            unsigned best = 0xffffffffu;    // (dist << 16) | (0xffff - position): smallest distance, LAST position
            for (int t = fs + lane; t < fe; t += 32) {
                const int r2 = idx2[t];
                if (r2 < 0 || r2 >= P.cap || !free2[r2]) continue;
                const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(desc2 + (size_t)r2 * 32));
                const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(desc2 + (size_t)r2 * 32) + 1);
                const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                              __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
                if (d > SB_TH_LOW) continue;
                const orbfe_keypoint kp2 = keys2[r2];
                const int o2 = min(max(kp2.octave, 0), ORBFE_MAX_LEVELS - 1);
                const bool bStereo2 = stereo2 && stereo2[r2];
                if (!P.kb8 && !bStereo1 && !bStereo2) {    // && !pKF1->mpCamera2
                    const float distex = sEp[0] - kp2.x, distey = sEp[1] - kp2.y;
                    if (distex * distex + distey * distey < 100.0f * P.scale[o2]) continue;
                }
                if (!P.coarse && P.kb8) {
                    const bool right1 = r1 >= nleft1, right2 = r2 >= nleft2;
                    const float* T = sT[(right1 ? 2 : 0) + (right2 ? 1 : 0)];
                    const int o1 = min(max(kp1.octave, 0), ORBFE_MAX_LEVELS - 1);
                    float p3D[3];
                    const float z = kb_triangulate(right1 ? P.camR : P.camL, right2 ? P.camR : P.camL, T, T + 9, P.precision, kp1.x,
                                                   kp1.y, kp2.x, kp2.y, P.sigma2[o1], P.sigma2[o2], p3D);
                    if (!(z > 0.0001f)) continue;
                } else if (!P.coarse) {
                    if (den == 0) continue;
                    const float num = a * kp2.x + b * kp2.y + c;
                    const float dsqr = num * num / den;
                    if (!((double)dsqr < 3.84 * (double)P.sigma2[o2])) continue;
                }
                best = min(best, ((unsigned)d << 16) | (0xffffu - (unsigned)(t - fs)));
            }
            best = sb_min_u32(best);
            if (lane == 0 && best != 0xffffffffu) {
                const int best2 = idx2[fs + (int)(0xffffu - (best & 0xffffu))];
                matches[r1] = best2;
                if (P.checkOrientation) {
                    float rot = kp1.angle - keys2[best2].angle;
                    if (rot < 0.0f) rot += 360.0f;
                    int bin = (int)roundf(rot * (1.0f / SB_HISTO));
                    if (bin == SB_HISTO) bin = 0;
                    bin = min(max(bin, 0), SB_HISTO - 1);
                    binOf[r1] = (uint8_t)bin;
                    atomicAdd(&hist[bin], 1);
                }
                atomicAdd(&sCount, 1);
            }
        }
    }
    __syncthreads();
    if (P.checkOrientation) {
        if (tid == 0) sb_three_maxima(hist, sInd);
        __syncthreads();
        int removed = 0;
        for (int i = tid; i < P.cap; i += SB_THREADS) {
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

int launch_search_bow_kf(const uint8_t* desc1, const float* angle1, const uint8_t* valid1, const int* ids1, const int* start1,
                         const int* idx1, const int* nodes1, const uint8_t* desc2, const float* angle2, const uint8_t* valid2,
                         const int* ids2, const int* start2, const int* idx2, const int* nodes2, int cap, int nodeCap, int npairs,
                         float nnRatio, int checkOrientation, int* matches, int* nMatches, cudaStream_t st) {
    if (npairs <= 0) return 0;
    BowParams P;
    P.cap = cap;
    P.nodeCap = nodeCap;
    P.nnRatio = nnRatio;
    P.checkOrientation = checkOrientation;
    const size_t smem = 2 * search_bow_smem_bytes(cap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, search_bow_kf_kernel, smem);
    search_bow_kf_kernel<<<npairs, SB_THREADS, smem, st>>>(P, desc1, angle1, valid1, ids1, start1, idx1, nodes1, desc2, angle2,
                                                          valid2, ids2, start2, idx2, nodes2, matches, nMatches);
    return 1;
}

int launch_search_triangulation(const orbfe_keypoint* keys1, const uint8_t* desc1, const uint8_t* free1, const uint8_t* stereo1,
                                const int* ids1, const int* start1, const int* idx1, const int* nodes1,
                                const orbfe_keypoint* keys2, const uint8_t* desc2, const uint8_t* free2, const uint8_t* stereo2,
                                const int* ids2, const int* start2, const int* idx2, const int* nodes2, const float* F12,
                                const float* ep, const float* scale, const float* sigma2, int nlevels, int cap, int nodeCap,
                                int npairs, int coarse, int checkOrientation, int* matches, int* nMatches, cudaStream_t st,
                                const float* kb8Cams, const float* T12, const int* nleft1, const int* nleft2) {
    if (npairs <= 0) return 0;
    TriParams P;
    P.kb8 = kb8Cams ? 1 : 0;    // kb8Cams: mpCamera parameters (8), mpCamera2 parameters (8), precision
    for (int i = 0; i < 8; i++) {
        P.camL[i] = kb8Cams ? kb8Cams[i] : 0.f;
        P.camR[i] = kb8Cams ? kb8Cams[8 + i] : 0.f;
    }
    P.precision = kb8Cams ? kb8Cams[16] : 0.f;
    P.cap = cap;
    P.nodeCap = nodeCap;
    P.coarse = coarse;
    P.checkOrientation = checkOrientation;
    for (int l = 0; l < ORBFE_MAX_LEVELS; l++) {
        P.scale[l] = l < nlevels ? scale[l] : 1.0f;
        P.sigma2[l] = l < nlevels ? sigma2[l] : 1.0f;
    }
    const size_t smem = search_bow_smem_bytes(cap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, search_triangulation_kernel, smem);
    search_triangulation_kernel<<<npairs, SB_THREADS, smem, st>>>(P, keys1, desc1, free1, stereo1, ids1, start1, idx1, nodes1, keys2,
                                                                 desc2, free2, stereo2, ids2, start2, idx2, nodes2, F12, ep,
                                                                 T12, nleft1, nleft2, matches, nMatches);
    return 1;
}

}    // namespace orbfe
