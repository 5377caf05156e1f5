// k_search_kf.cu — the matchers that project MapPoints into a KEY FRAME and search its grid:
//   ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming)                 ORBmatcher.cc:428-533
//   ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpPointsKFs, vpMatched, vpMatchedKF, ...)    :535-647   (loop closing / merging)
//   ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight)                                                  :1155-1345 (local mapping)
//   ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)                                        :1347-1462 (loop correction)
//   ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, S12, th)                                      :1464-1681 (two searches + agreement)
// with KeyFrame::GetFeaturesInArea (KeyFrame.cc:706-750) and KeyFrame::IsInImage (:752-755).
//
// All five share one inner loop: window of th * mvScaleFactors[nPredictedLevel] around the projection on the 64 x 48 grid
// (cells ix outer, iy inner, push_back order inside a cell), features of level nPredictedLevel - 1 or nPredictedLevel only,
// the FIRST minimum of DescriptorDistance in that order, accepted at <= a threshold (TH_LOW * ratioHamming, TH_LOW,
// TH_HIGH).  Differences: the Sim3 projection search skips features whose vpMatched slot is set and takes the slot it
// accepts, so later points depend on earlier ones (handled by the optimistic chunk + ordered prefix commit of
// k_search_stereo.cu); Fuse(pKF, vpMapPoints) drops candidates by the chi-square reprojection gate (:1274-1299).
// The pose algebra, PredictScale and the MapPoint gates that need the map (isBad, already found, distance range, viewing
// cosine) stay with the caller: a point arrives as orbfe_frame_point (u, v, invz, level, flags bit 0, descriptor).
// The map mutations of Fuse (Replace / AddObservation) are the caller's too: the device returns bestIdx per point.
#include "search_grid.cuh"

namespace orbfe {

struct KfPoint {    // == orbfe_frame_point
    float u, v, ur, vr, invz, angle;
    int level, flags;
    uint32_t desc[8];
};
static_assert(sizeof(KfPoint) == 64, "orbfe_frame_point layout");

struct KfSearchParams {
    StereoSearchParams g;    // bounds, scale, th, cap, ptsCap
    float invSigma2[ORBFE_MAX_LEVELS];
    float distTh;    // accepted when (float)bestDist <= distTh
    float bf;
    int blocks;      // an accepted point takes its slot (vpMatched[bestIdx] = pMP) and set slots are skipped
    int chi2;        // the reprojection gate of Fuse(pKF, vpMapPoints, th, bRight)
};

// best = min (distance, cell, index) over the features of window w that pass the gates: the first minimum in the
// reference's visiting order.  Warp-uniform result.
__device__ __forceinline__ unsigned long long kf_search(const SWindow& w, const int* cellStart, const unsigned short* items,
                                                        const orbfe_keypoint* keys, const uint8_t* desc,
                                                        const uint32_t* blocked, const KfPoint& mp, const float* uRight,
                                                        const KfSearchParams& Q) {
    constexpr unsigned long long NONE = 256ull << 32;
    unsigned long long k1 = NONE;
    if (w.cx0 <= w.cx1) {
        const int lane = threadIdx.x & 31;
        const int ncy = w.cy1 - w.cy0 + 1;
        const int ncell = (w.cx1 - w.cx0 + 1) * ncy;
        const float urProj = mp.u - Q.bf * mp.invz;    // :1221 (no FMA: --fmad=false)
        for (int c = lane; c < ncell; c += 32) {
            const int cell = (w.cx0 + c / ncy) * SS_ROWS + w.cy0 + c % ncy;
            for (int t = cellStart[cell]; t < cellStart[cell + 1]; t++) {
                const int idx = items[t];
                const orbfe_keypoint k = keys[idx];
                if (!ss_inside(w, k)) continue;
                if (blocked && ((blocked[idx >> 5] >> (idx & 31)) & 1u)) continue;
                if (Q.chi2) {
                    const float ex = mp.u - k.x, ey = mp.v - k.y;
                    const float kpr = uRight ? uRight[idx] : -1.0f;
                    const float inv = Q.invSigma2[min(max(k.octave, 0), ORBFE_MAX_LEVELS - 1)];
                    if (kpr >= 0) {
                        const float er = urProj - kpr;
                        const float e2 = ex * ex + ey * ey + er * er;
                        if ((double)(e2 * inv) > 7.8) continue;
                    } else {
                        const float e2 = ex * ex + ey * ey;
                        if ((double)(e2 * inv) > 5.99) continue;
                    }
                }
                const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(desc + (size_t)idx * 32));
                const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(desc + (size_t)idx * 32) + 1);
                const int d = __popc(mp.desc[0] ^ b0.x) + __popc(mp.desc[1] ^ b0.y) + __popc(mp.desc[2] ^ b0.z) +
                              __popc(mp.desc[3] ^ b0.w) + __popc(mp.desc[4] ^ b1.x) + __popc(mp.desc[5] ^ b1.y) +
                              __popc(mp.desc[6] ^ b1.z) + __popc(mp.desc[7] ^ b1.w);
                const unsigned long long key = ((unsigned long long)d << 32) | ((uint32_t)cell << 16) | (uint32_t)idx;
                k1 = key < k1 ? key : k1;
            }
        }
    }
    return ss_min_u64(k1);
}

__device__ __forceinline__ bool kf_eligible(const KfPoint& q, const StereoSearchParams& P) {
    // flags bit 0: the point reaches the projection; KeyFrame::IsInImage (a NaN projection fails every comparison)
    return (q.flags & 1) && q.level >= 0 && q.level < ORBFE_MAX_LEVELS && q.u >= P.minX && q.u < P.maxX && q.v >= P.minY &&
           q.v < P.maxY;
}

__device__ __forceinline__ void kf_window(SWindow& w, const KfPoint& q, const StereoSearchParams& P, float gwInv, float ghInv) {
    const float radius = P.th * P.scale[q.level];
    ss_window(w, P, gwInv, ghInv, q.u, q.v, radius, q.level);    // levels [level - 1, level]
}

// grid: nframes; block SS_THREADS; dynamic shared memory: search_kf_smem_bytes(cap)
__global__ void __launch_bounds__(SS_THREADS) search_kf_kernel(
    const __grid_constant__ KfSearchParams Q, const orbfe_keypoint* __restrict__ keysAll, const uint8_t* __restrict__ descAll,
    const float* __restrict__ uRightAll, const uint8_t* __restrict__ occupiedAll, const int* __restrict__ nFeat,
    const KfPoint* __restrict__ ptsAll, const int* __restrict__ nPts, int* __restrict__ matchAll, int* __restrict__ distAll,
    int* __restrict__ ownerAll, int* __restrict__ nMatches) {
    const StereoSearchParams& P = Q.g;
    extern __shared__ __align__(16) uint8_t smem[];
    int* cellStart = reinterpret_cast<int*>(smem);
    int* cellFill = cellStart + SS_CELLS + 1;
    uint32_t* blockedBits = reinterpret_cast<uint32_t*>(cellFill + SS_CELLS);
    unsigned short* items = reinterpret_cast<unsigned short*>(blockedBits + ((P.cap + 31) >> 5));
    __shared__ int sScan[SS_WARPS + 1];
    __shared__ int sTotal, sOpen, sCount;
    __shared__ int sRes[SS_WARPS], sDist[SS_WARPS];

    const int f = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(nFeat[f], P.cap);
    const int npts = min(nPts[f], P.ptsCap);
    const orbfe_keypoint* keys = keysAll + (size_t)f * P.cap;
    const uint8_t* desc = descAll + (size_t)f * P.cap * 32;
    const float* uRight = uRightAll ? uRightAll + (size_t)f * P.cap : nullptr;
    const uint8_t* occupied = occupiedAll ? occupiedAll + (size_t)f * P.cap : nullptr;
    const KfPoint* pts = ptsAll + (size_t)f * P.ptsCap;
    int* match = matchAll + (size_t)f * P.ptsCap;
    int* dist = distAll ? distAll + (size_t)f * P.ptsCap : nullptr;
    int* owner = ownerAll ? ownerAll + (size_t)f * P.cap : nullptr;
    const float gwInv = (float)SS_COLS / (P.maxX - P.minX);
    const float ghInv = (float)SS_ROWS / (P.maxY - P.minY);
    const uint32_t* blocked = Q.blocks ? blockedBits : nullptr;

    for (int i = tid; i < (P.cap + 31) >> 5; i += SS_THREADS) blockedBits[i] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();
    for (int i = tid; i < n; i += SS_THREADS) {
        if (Q.blocks && occupied && occupied[i]) atomicOr(&blockedBits[i >> 5], 1u << (i & 31));
        if (owner) owner[i] = -1;
    }
    ss_build_grid(keys, 0, n, P, gwInv, ghInv, cellStart, cellFill, items, sScan, &sTotal);

    if (!Q.blocks) {
        // independent points: one warp each, no commit protocol
        int accepted = 0;
        for (int iP = warp; iP < npts; iP += SS_WARPS) {
            const KfPoint mp = pts[iP];
            int res = -1, bd = -1;
            if (kf_eligible(mp, P)) {
                SWindow w;
                kf_window(w, mp, P, gwInv, ghInv);
                const unsigned long long k1 = kf_search(w, cellStart, items, keys, desc, nullptr, mp, uRight, Q);
                const int d = (int)(k1 >> 32);
                if (d < 256 && (float)d <= Q.distTh) {
                    res = (int)(k1 & 0xffffu);
                    bd = d;
                }
            }
            if (lane == 0) {
                match[iP] = res;
                if (dist) dist[iP] = bd;
                if (res >= 0) {
                    accepted++;
                    if (owner) atomicMax(&owner[res], iP);    // several points may choose one feature: the last one is reported
                }
            }
        }
        if (lane == 0 && accepted) atomicAdd(&sCount, accepted);
        __syncthreads();
        if (tid == 0 && nMatches) nMatches[f] = sCount;
        return;
    }

    for (int base = 0; base < npts; base += SS_WARPS) {
        const int iP = base + warp;
        const bool have = iP < npts;
        KfPoint mp;
        bool eligible = false;
        if (have) {
            mp = pts[iP];
            eligible = kf_eligible(mp, P);
        }
        if (tid == 0) sOpen = (int)(0xffffffffu >> (32 - min(npts - base, SS_WARPS)));
        __syncthreads();
        for (;;) {
            if (have && ((sOpen >> warp) & 1)) {
                int res = -1, bd = -1;
                if (eligible) {
                    SWindow w;
                    kf_window(w, mp, P, gwInv, ghInv);
                    const unsigned long long k1 = kf_search(w, cellStart, items, keys, desc, blocked, mp, uRight, Q);
                    const int d = (int)(k1 >> 32);
                    if (d < 256 && (float)d <= Q.distTh) {
                        res = (int)(k1 & 0xffffu);
                        bd = d;
                    }
                }
                if (lane == 0) {
                    sRes[warp] = res;
                    sDist[warp] = bd;
                }
            }
            __syncthreads();
            if (warp == 0) {
                const int open = sOpen;
                const bool mine = lane < SS_WARPS && ((open >> lane) & 1);
                SWindow wl;
                wl.cx0 = 1;
                wl.cx1 = 0;
                bool el = false;
                if (mine) {
                    const KfPoint q = pts[base + lane];
                    el = kf_eligible(q, P);
                    if (el) kf_window(wl, q, P, gwInv, ghInv);
                }
                const int take = mine ? sRes[lane] : -1;
                bool valid = true;
                for (int i = 0; i < SS_WARPS - 1; i++) {
                    const int tl = __shfl_sync(0xffffffffu, take, i);
                    if (i < lane && mine && el && tl >= 0 && ss_inside(wl, keys[tl])) valid = false;    // conservative
                }
                const unsigned bad = __ballot_sync(0xffffffffu, mine && !valid);
                const int firstBad = bad ? __ffs(bad) - 1 : 32;
                const bool commit = mine && lane < firstBad;
                const unsigned done = __ballot_sync(0xffffffffu, commit);
                if (commit) {    // committed points of a round chose different slots (a shared one invalidates the later point)
                    const int r = sRes[lane];
                    match[base + lane] = r;
                    if (dist) dist[base + lane] = sDist[lane];
                    if (r >= 0) {
                        if (owner) owner[r] = base + lane;
                        atomicOr(&blockedBits[r >> 5], 1u << (r & 31));
                        atomicAdd(&sCount, 1);
                    }
                }
                if (lane == 0) sOpen = open & ~(int)done;
            }
            __syncthreads();
            if (sOpen == 0) break;
            __syncthreads();
        }
        __syncthreads();
    }
    if (tid == 0 && nMatches) nMatches[f] = sCount;
}

// SearchBySim3's agreement (:1666-1679): vpMatches12[i1] = idx2 when vnMatch1[i1] == idx2 and vnMatch2[idx2] == i1
__global__ void __launch_bounds__(256) sim3_agree_kernel(const int* __restrict__ m1All, const int* __restrict__ m2All,
                                                         const int* __restrict__ n1Arr, const int* __restrict__ n2Arr, int cap,
                                                         int* __restrict__ outAll, int* __restrict__ nFound) {
    __shared__ int sCount;
    const int f = blockIdx.x, tid = threadIdx.x;
    const int n1 = min(n1Arr[f], cap), n2 = min(n2Arr[f], cap);
    const int* m1 = m1All + (size_t)f * cap;
    const int* m2 = m2All + (size_t)f * cap;
    int* out = outAll + (size_t)f * cap;
    if (tid == 0) sCount = 0;
    __syncthreads();
    int found = 0;
    for (int i1 = tid; i1 < cap; i1 += 256) {
        int r = -1;
        if (i1 < n1) {
            const int idx2 = m1[i1];
            if (idx2 >= 0 && idx2 < n2 && m2[idx2] == i1) {
                r = idx2;
                found++;
            }
        }
        out[i1] = r;
    }
    if (found) atomicAdd(&sCount, found);
    __syncthreads();
    if (tid == 0 && nFound) nFound[f] = sCount;
}

size_t search_kf_smem_bytes(int cap) {
    return (size_t)(2 * SS_CELLS + 1) * 4 + (size_t)((cap + 31) / 32) * 4 + (size_t)cap * 2 + 16;
}

int launch_search_kf(const float bounds[4], const float* scale, const float* invSigma2, int nlevels, float th, float distTh,
                     float bf, int blocks, int chi2, const orbfe_keypoint* keys, const uint8_t* desc, const float* uRight,
                     const uint8_t* occupied, const int* nFeat, int cap, const void* pts, const int* nPts, int ptsCap,
                     int nframes, int* match, int* dist, int* owner, int* nMatches, cudaStream_t st) {
    if (nframes <= 0) return 0;
    KfSearchParams Q;
    StereoSearchParams& P = Q.g;
    P.minX = bounds[0];
    P.minY = bounds[1];
    P.maxX = bounds[2];
    P.maxY = bounds[3];
    for (int l = 0; l < ORBFE_MAX_LEVELS; l++) {
        P.scale[l] = l < nlevels ? scale[l] : 1.0f;
        Q.invSigma2[l] = (invSigma2 && l < nlevels) ? invSigma2[l] : 1.0f;
    }
    P.th = th;
    P.thFar = 0.f;
    P.nnRatio = 0.f;
    P.farPoints = 0;
    P.cap = cap;
    P.ptsCap = ptsCap;
    Q.distTh = distTh;
    Q.bf = bf;
    Q.blocks = blocks;
    Q.chi2 = chi2;
    const size_t smem = search_kf_smem_bytes(cap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, search_kf_kernel, smem);
    search_kf_kernel<<<nframes, SS_THREADS, smem, st>>>(Q, keys, desc, uRight, occupied, nFeat, (const KfPoint*)pts, nPts, match,
                                                       dist, owner, nMatches);
    return 1;
}

int launch_sim3_agree(const int* m1, const int* m2, const int* n1, const int* n2, int cap, int nframes, int* out, int* nFound,
                      cudaStream_t st) {
    if (nframes <= 0) return 0;
    sim3_agree_kernel<<<nframes, 256, 0, st>>>(m1, m2, n1, n2, cap, out, nFound);
    return 1;
}

}    // namespace orbfe
