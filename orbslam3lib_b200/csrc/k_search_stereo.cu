// k_search_stereo.cu — ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFarPoints, thFarPoints) for
// stereo-camera frames (Nleft != -1, the Lynx fisheye rig): ORBmatcher.cc:44-214 with both branches.
//
// Features are the N = Nleft + Nright rows of mDescriptors / slots of F.mvpMapPoints, left first.  Per map point:
// the left search on mGrid (:59-143; no mvuRight gate in this mode), F.mvpMapPoints[bestIdx] = pMP and the same for the
// stereo partner mvLeftToRightMatch[bestIdx] + Nleft (:128-134); a ratio rejection `continue`s past the right search
// (:126-127); then the right search on mGridRight around (mTrackProjXR, mTrackProjYR) with mnTrackScaleLevelR and a radius
// that is NOT scaled by th (:148-151), F.mvpMapPoints[bestIdx + Nleft] = pMP and the partner mvRightToLeftMatch[bestIdx]
// (:196-205).  A slot holding a point with Observations() > 0 is skipped by later searches (:84-86, :171-173); an
// assignment of a point WITHOUT observations frees the slot again, which the partner assignments can do to an occupied
// slot, so every assignment both sets or clears the slot's bit.
//
// Same scheme as k_search.cu: one CTA per frame, two counting-sort grids in shared memory, one warp per map point, the
// order-free best / runner-up rule, and the optimistic chunk + ordered prefix commit for the dependence between
// points — here a point writes up to four slots, the right search of a point sees the writes of its own left search,
// and commits are applied in point order because two points of a round may write the same partner slot.
#include "search_grid.cuh"

namespace orbfe {

struct TrackPointStereo {    // == orbfe_track_point_stereo
    float projX, projY, projXR, projYR, viewCos, viewCosR, trackDepth;
    int level, levelR, flags;
    uint32_t desc[8];
};
static_assert(sizeof(TrackPointStereo) == 72, "orbfe_track_point_stereo layout");


// One search of the map-point loop by a warp: candidates of window w in the grid (cellStart, items) whose features are
// keys[first + idx]; slot first + idx is skipped when occupied (ovSlot overrides the shared bit with ovState).
// Returns the accepted feature index (relative to `first`) or -1; *rejected = the ratio test failed on equal levels.
__device__ __forceinline__ int ss_search(const SWindow& w, const int* cellStart, const unsigned short* items,
                                         const orbfe_keypoint* keys, const uint8_t* desc, int first,
                                         const uint32_t* blocked, int ovSlot, bool ovState, const uint32_t (&md)[8],
                                         float nnRatio, bool* rejected) {
    *rejected = false;
    if (w.cx0 > w.cx1) return -1;
    const int lane = threadIdx.x & 31;
    const int ncy = w.cy1 - w.cy0 + 1;
    const int ncell = (w.cx1 - w.cx0 + 1) * ncy;
    constexpr unsigned long long NONE = 256ull << 32;
    unsigned long long k1 = NONE, kP = NONE, kS = NONE;
    for (int pass = 0; pass < 2; pass++) {
        const uint32_t bestPos = (uint32_t)k1;
        unsigned long long kA = NONE, kB = NONE;
        for (int c = lane; c < ncell; c += 32) {
            const int ix = w.cx0 + c / ncy, iy = w.cy0 + c % ncy;
            const int cell = ix * SS_ROWS + iy;
            for (int t = cellStart[cell]; t < cellStart[cell + 1]; t++) {
                const int idx = items[t];
                const orbfe_keypoint k = keys[first + idx];
                if (!ss_inside(w, k)) continue;
                const int slot = first + idx;
                const bool occ = slot == ovSlot ? ovState : ((blocked[slot >> 5] >> (slot & 31)) & 1u);
                if (occ) continue;
                const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(desc + (size_t)slot * 32));
                const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(desc + (size_t)slot * 32) + 1);
                const int d = __popc(md[0] ^ b0.x) + __popc(md[1] ^ b0.y) + __popc(md[2] ^ b0.z) + __popc(md[3] ^ b0.w) +
                              __popc(md[4] ^ b1.x) + __popc(md[5] ^ b1.y) + __popc(md[6] ^ b1.z) + __popc(md[7] ^ b1.w);
                const uint32_t pos = ((uint32_t)cell << 16) | (uint32_t)idx;
                const unsigned long long key = ((unsigned long long)d << 32) | pos;
                if (pass == 0) {
                    kA = key < kA ? key : kA;
                } else if (pos < bestPos) {
                    kA = key < kA ? key : kA;
                } else if (pos > bestPos) {
                    kB = key < kB ? key : kB;
                }
            }
        }
        kA = ss_min_u64(kA);
        kB = ss_min_u64(kB);
        if (pass == 0) {
            k1 = kA;
            if ((int)(k1 >> 32) > SS_TH_HIGH) return -1;    // warp uniform (also covers "no candidate": 256)
        } else {
            kP = kA;
            kS = kB;
        }
    }
    const int bestDist = (int)(k1 >> 32);
    const int bestIdx = (int)(k1 & 0xffffu);
    const int bestLevel = keys[first + bestIdx].octave;
    const int dP = (int)(kP >> 32), dS = (int)(kS >> 32);
    int bestDist2 = 256, bestLevel2 = -1;
    if (dP < 256 && dP <= dS) {
        bestDist2 = dP;
        bestLevel2 = keys[first + (int)(kP & 0xffffu)].octave;
    } else if (dS < 256) {
        bestDist2 = dS;
        bestLevel2 = keys[first + (int)(kS & 0xffffu)].octave;
    }
    const float lim = nnRatio * (float)bestDist2;
    if (bestLevel == bestLevel2 && (float)bestDist > lim) {
        *rejected = true;
        return -1;
    }
    return bestIdx;    // the complement of the rejection: bestLevel != bestLevel2 || bestDist <= lim
}

// grid: nframes; block SS_THREADS; dynamic shared memory: see search_stereo_smem_bytes.
__global__ void __launch_bounds__(SS_THREADS) search_by_projection_stereo_kernel(
    const __grid_constant__ StereoSearchParams P, const orbfe_keypoint* __restrict__ keysAll,
    const uint8_t* __restrict__ descAll, const uint8_t* __restrict__ occupiedAll, const int* __restrict__ nFeat,
    const int* __restrict__ nLeftArr, const int* __restrict__ ltrAll, const int* __restrict__ rtlAll,
    const TrackPointStereo* __restrict__ ptsAll, const int* __restrict__ nPts, int* __restrict__ matchLAll,
    int* __restrict__ matchRAll, int* __restrict__ ownerAll, int* __restrict__ nMatches) {
    extern __shared__ __align__(16) uint8_t smem[];
    int* cellStartL = reinterpret_cast<int*>(smem);
    int* cellStartR = cellStartL + SS_CELLS + 1;
    int* cellFill = cellStartR + SS_CELLS + 1;
    uint32_t* blocked = reinterpret_cast<uint32_t*>(cellFill + SS_CELLS);
    unsigned short* itemsL = reinterpret_cast<unsigned short*>(blocked + ((P.cap + 31) >> 5));
    unsigned short* itemsR = itemsL + P.cap;
    __shared__ int sScan[SS_WARPS + 1];
    __shared__ int sTotal, sOpen, sCount;
    __shared__ int sResL[SS_WARPS], sResR[SS_WARPS], sCnt[SS_WARPS];
    __shared__ int sW[SS_WARPS][4];    // slots written by the chunk's points (-1 = none)

    const int f = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(nFeat[f], P.cap);
    const int nleft = min(max(nLeftArr[f], 0), n);
    const int nright = n - nleft;
    const int npts = min(nPts[f], P.ptsCap);
    const orbfe_keypoint* keys = keysAll + (size_t)f * P.cap;
    const uint8_t* desc = descAll + (size_t)f * P.cap * 32;
    const uint8_t* occupied = occupiedAll ? occupiedAll + (size_t)f * P.cap : nullptr;
    const int* ltr = ltrAll + (size_t)f * P.cap;    // nleft entries
    const int* rtl = rtlAll + (size_t)f * P.cap;    // nright entries
    const TrackPointStereo* pts = ptsAll + (size_t)f * P.ptsCap;
    int* matchL = matchLAll + (size_t)f * P.ptsCap;
    int* matchR = matchRAll + (size_t)f * P.ptsCap;
    int* owner = ownerAll ? ownerAll + (size_t)f * P.cap : nullptr;

    const float gwInv = (float)SS_COLS / (P.maxX - P.minX);
    const float ghInv = (float)SS_ROWS / (P.maxY - P.minY);

    for (int i = tid; i < (P.cap + 31) >> 5; i += SS_THREADS) blocked[i] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();
    for (int i = tid; i < n; i += SS_THREADS) {
        if (occupied && occupied[i]) atomicOr(&blocked[i >> 5], 1u << (i & 31));
        if (owner) owner[i] = -1;
    }
    ss_build_grid(keys, 0, nleft, P, gwInv, ghInv, cellStartL, cellFill, itemsL, sScan, &sTotal);
    ss_build_grid(keys, nleft, nright, P, gwInv, ghInv, cellStartR, cellFill, itemsR, sScan, &sTotal);

    const bool bFactor = P.th != 1.0f;
    auto level_ok = [](int l) { return l >= 0 && l < ORBFE_MAX_LEVELS; };
    for (int base = 0; base < npts; base += SS_WARPS) {
        const int iMP = base + warp;
        const bool have = iMP < npts;
        TrackPointStereo mp;
        bool eligible = false;
        if (have) {
            mp = pts[iMP];
            const bool inView = mp.flags & 1, inViewR = mp.flags & 8;
            eligible = (inView || inViewR) && !(P.farPoints && mp.trackDepth > P.thFar) && !(mp.flags & 2);
        }
        if (tid == 0) sOpen = (int)(0xffffffffu >> (32 - min(npts - base, SS_WARPS)));
        __syncthreads();
        for (;;) {
            if (have && ((sOpen >> warp) & 1)) {
                int resL = -1, resR = -1, cnt = 0;
                int w0 = -1, w1 = -1, w2 = -1, w3 = -1;
                if (eligible) {
                    const bool obs = mp.flags & 4;
                    bool skipRight = false;
                    if ((mp.flags & 1) && level_ok(mp.level)) {
                        float r = (double)mp.viewCos > 0.998 ? 2.5f : 4.0f;
                        if (bFactor) r *= P.th;
                        SWindow w;
                        ss_window(w, P, gwInv, ghInv, mp.projX, mp.projY, r * P.scale[mp.level], mp.level);
                        bool rej;
                        resL = ss_search(w, cellStartL, itemsL, keys, desc, 0, blocked, -1, false, mp.desc, P.nnRatio, &rej);
                        skipRight = rej;
                        if (resL >= 0) {
                            w0 = resL;
                            cnt = 1;
                            const int partner = ltr[resL];
                            if (partner != -1) {
                                w1 = partner + nleft;
                                cnt = 2;
                            }
                        }
                    }
                    if (!skipRight && (mp.flags & 8) && mp.levelR != -1 && level_ok(mp.levelR)) {
                        const float r = (double)mp.viewCosR > 0.998 ? 2.5f : 4.0f;    // no th factor here (:148-151)
                        SWindow w;
                        ss_window(w, P, gwInv, ghInv, mp.projXR, mp.projYR, r * P.scale[mp.levelR], mp.levelR);
                        bool rej;
                        resR = ss_search(w, cellStartR, itemsR, keys, desc, nleft, blocked, w1, obs, mp.desc, P.nnRatio, &rej);
                        if (resR >= 0) {
                            const int partner = rtl[resR];
                            if (partner != -1) {
                                w2 = partner;
                                cnt++;
                            }
                            w3 = resR + nleft;
                            cnt++;
                        }
                    }
                }
                if (lane == 0) {
                    sResL[warp] = resL;
                    sResR[warp] = resR;
                    sCnt[warp] = cnt;
                    sW[warp][0] = w0;
                    sW[warp][1] = w1;
                    sW[warp][2] = w2;
                    sW[warp][3] = w3;
                }
            }
            __syncthreads();
            if (warp == 0) {
                const int open = sOpen;
                const bool mine = (open >> lane) & 1;
                SWindow wl, wr;
                wl.cx0 = wr.cx0 = 1;
                wl.cx1 = wr.cx1 = 0;
                bool haveL = false, haveR = false, obs = false;
                if (mine) {
                    const TrackPointStereo q = pts[base + lane];
                    obs = q.flags & 4;
                    if (level_ok(q.level)) {
                        float r = (double)q.viewCos > 0.998 ? 2.5f : 4.0f;
                        if (bFactor) r *= P.th;
                        ss_window(wl, P, gwInv, ghInv, q.projX, q.projY, r * P.scale[q.level], q.level);
                        haveL = true;
                    }
                    if (level_ok(q.levelR)) {
                        const float r = (double)q.viewCosR > 0.998 ? 2.5f : 4.0f;
                        ss_window(wr, P, gwInv, ghInv, q.projXR, q.projYR, r * P.scale[q.levelR], q.levelR);
                        haveR = true;
                    }
                }
                bool valid = true;
                for (int i = 0; i < SS_WARPS; i++) {
                    if (!((open >> i) & 1)) continue;    // uniform
                    if (i < lane && mine) {
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            const int s = sW[i][q];
                            if (s < 0) continue;
                            // conservative: the slot lies inside this point's window of the same camera
                            if (s < nleft ? (haveL && ss_inside(wl, keys[s])) : (haveR && ss_inside(wr, keys[s]))) valid = false;
                        }
                    }
                }
                const unsigned bad = __ballot_sync(0xffffffffu, mine && !valid);
                const int firstBad = bad ? __ffs(bad) - 1 : 32;
                const bool commit = mine && lane < firstBad;
                unsigned todo = __ballot_sync(0xffffffffu, commit);
                const unsigned done = todo;
                while (todo) {    // in point order: two points of a round may write the same partner slot
                    const int i = __ffs(todo) - 1;
                    todo &= todo - 1;
                    if (lane == i) {
                        matchL[base + i] = sResL[i];
                        matchR[base + i] = sResR[i];
                        sCount += sCnt[i];
#pragma unroll
                        for (int q = 0; q < 4; q++) {
                            const int s = sW[i][q];
                            if (s < 0) continue;
                            if (owner) owner[s] = base + i;
                            if (obs) blocked[s >> 5] |= 1u << (s & 31);
                            else blocked[s >> 5] &= ~(1u << (s & 31));
                        }
                    }
                    __syncwarp();
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

size_t search_stereo_smem_bytes(int cap) {
    return (size_t)(3 * SS_CELLS + 2) * 4 + (size_t)((cap + 31) / 32) * 4 + (size_t)cap * 4 + 16;
}

int launch_search_by_projection_stereo(const float bounds[4], const float* scale, int nlevels, float th, int farPoints,
                                       float thFar, float nnRatio, const orbfe_keypoint* keys, const uint8_t* desc,
                                       const uint8_t* occupied, const int* nFeat, const int* nLeft, const int* ltr,
                                       const int* rtl, int cap, const void* pts, const int* nPts, int ptsCap, int nframes,
                                       int* matchL, int* matchR, int* owner, int* nMatches, cudaStream_t st) {
    if (nframes <= 0) return 0;
    StereoSearchParams P;
    P.minX = bounds[0];
    P.minY = bounds[1];
    P.maxX = bounds[2];
    P.maxY = bounds[3];
    for (int l = 0; l < ORBFE_MAX_LEVELS; l++) P.scale[l] = l < nlevels ? scale[l] : 1.0f;
    P.th = th;
    P.thFar = thFar;
    P.nnRatio = nnRatio;
    P.farPoints = farPoints;
    P.cap = cap;
    P.ptsCap = ptsCap;
    const size_t smem = search_stereo_smem_bytes(cap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, search_by_projection_stereo_kernel, smem);
    search_by_projection_stereo_kernel<<<nframes, SS_THREADS, smem, st>>>(
        P, keys, desc, occupied, nFeat, nLeft, ltr, rtl, (const TrackPointStereo*)pts, nPts, matchL, matchR, owner, nMatches);
    return 1;
}

// ---- ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (ORBmatcher.cc:1683-1921) and ----------------------
// ---- ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (:1923-2059) ----------------------------
// The per-frame tracking matcher: every MapPoint of the last frame (of a key frame) is projected into the current frame
// by the caller and searched in a window of th * scale[octave] on the frame grid; levels by motion direction (bForward:
// octave and above, bBackward: 0..octave, else octave +- 1); the first minimum in visiting order wins, accepted at
// <= TH_HIGH (<= ORBdist); stereo-camera frames repeat the search on mGridRight around the right-camera projection, but
// only when the left window held a feature at all (:1769-1770 `continue`); mono / rectified frames gate on mvuRight
// (:1785-1790).  A slot taken by a point with Observations() > 0 (KeyFrame form: by any point) is invisible to the later
// points: same optimistic chunk + ordered prefix commit as above.  Rotation histogram: one entry per accepted search
// (bins 0..12 only, because factor = 1 / HISTO_LENGTH multiplies degrees: reproduced), ComputeThreeMaxima, and every entry
// of a losing bin clears its slot and takes one off nmatches (:1697-1715) even when another point holds the slot too.
struct FramePoint {    // == orbfe_frame_point
    float u, v, ur, vr, invz, angle;
    int level, flags;
    uint32_t desc[8];
};
static_assert(sizeof(FramePoint) == 64, "orbfe_frame_point layout");

struct FrameSearchParams {
    StereoSearchParams g;    // bounds, scale, th, cap, ptsCap
    float mbf;
    int keyframe, distTh, checkOrientation;
};

// best = min (distance, cell, index) over the features of window w that are neither taken nor gated out; *any = the
// window holds a feature at all (vIndices2 is not empty).  Warp-uniform results.
__device__ __forceinline__ unsigned long long sf_search(const SWindow& w, const int* cellStart, const unsigned short* items,
                                                        const orbfe_keypoint* keys, const uint8_t* desc, int first,
                                                        const uint32_t* blocked, const uint32_t (&md)[8],
                                                        const float* uRight, float urProj, float radius, bool* any) {
    constexpr unsigned long long NONE = 256ull << 32;
    unsigned long long k1 = NONE;
    bool seen = false;
    if (w.cx0 <= w.cx1) {
        const int lane = threadIdx.x & 31;
        const int ncy = w.cy1 - w.cy0 + 1;
        const int ncell = (w.cx1 - w.cx0 + 1) * ncy;
        for (int c = lane; c < ncell; c += 32) {
            const int cell = (w.cx0 + c / ncy) * SS_ROWS + w.cy0 + c % ncy;
            for (int t = cellStart[cell]; t < cellStart[cell + 1]; t++) {
                const int idx = items[t];
                if (!ss_inside(w, keys[first + idx])) continue;
                seen = true;
                const int slot = first + idx;
                if ((blocked[slot >> 5] >> (slot & 31)) & 1u) continue;
                if (uRight) {
                    const float ur = uRight[slot];
                    if (ur > 0) {
                        const float er = fabsf(urProj - ur);
                        if (er > radius) continue;
                    }
                }
                const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(desc + (size_t)slot * 32));
                const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(desc + (size_t)slot * 32) + 1);
                const int d = __popc(md[0] ^ b0.x) + __popc(md[1] ^ b0.y) + __popc(md[2] ^ b0.z) + __popc(md[3] ^ b0.w) +
                              __popc(md[4] ^ b1.x) + __popc(md[5] ^ b1.y) + __popc(md[6] ^ b1.z) + __popc(md[7] ^ b1.w);
                const unsigned long long key = ((unsigned long long)d << 32) | ((uint32_t)cell << 16) | (uint32_t)idx;
                k1 = key < k1 ? key : k1;
            }
        }
    }
    *any = __any_sync(0xffffffffu, seen);
    return ss_min_u64(k1);
}

__device__ __forceinline__ void sf_levels(SWindow& w, int level, int dir, int keyframe) {
    if (!keyframe && dir == 1) {
        w.minLevel = level;
        w.maxLevel = -1;
    } else if (!keyframe && dir == 2) {
        w.minLevel = 0;
        w.maxLevel = level;
    } else {
        w.minLevel = level - 1;
        w.maxLevel = level + 1;
    }
}

__device__ __forceinline__ int sf_bin(float angleSrc, float angleCur, int keyframe) {
    float rot = angleSrc - angleCur;
    if (keyframe && (rot < 0.0f || rot > 360.0f)) rot = fmodf(rot, 360.0f);    // :2014-2018
    if (rot < 0.0f) rot += 360.0f;
    int bin = (int)roundf(rot * (1.0f / SI_HISTO_LEN));
    if (bin == SI_HISTO_LEN) bin = 0;
    return min(max(bin, 0), SI_HISTO_LEN - 1);
}

// grid: nframes; block SS_THREADS; dynamic shared memory: search_frame_smem_bytes(cap, ptsCap)
__global__ void __launch_bounds__(SS_THREADS) search_frame_kernel(
    const __grid_constant__ FrameSearchParams Q, const orbfe_keypoint* __restrict__ keysAll,
    const uint8_t* __restrict__ descAll, const float* __restrict__ uRightAll, const uint8_t* __restrict__ occupiedAll,
    const int* __restrict__ nFeat, const int* __restrict__ nLeftArr, const int* __restrict__ dirArr,
    const FramePoint* __restrict__ ptsAll, const int* __restrict__ nPts, int* __restrict__ matchLAll,
    int* __restrict__ matchRAll, int* __restrict__ ownerAll, int* __restrict__ nMatches) {
    const StereoSearchParams& P = Q.g;
    extern __shared__ __align__(16) uint8_t smem[];
    int* cellStartL = reinterpret_cast<int*>(smem);
    int* cellStartR = cellStartL + SS_CELLS + 1;
    int* cellFill = cellStartR + SS_CELLS + 1;
    int* hist = cellFill + SS_CELLS;    // 32
    uint32_t* blocked = reinterpret_cast<uint32_t*>(hist + 32);
    unsigned short* itemsL = reinterpret_cast<unsigned short*>(blocked + ((P.cap + 31) >> 5));
    unsigned short* itemsR = itemsL + P.cap;
    uint8_t* binL = reinterpret_cast<uint8_t*>(itemsR + P.cap);    // histogram bin of point i's left / right entry, 0xff none
    uint8_t* binR = binL + P.ptsCap;
    __shared__ int sScan[SS_WARPS + 1];
    __shared__ int sTotal, sOpen, sCount;
    __shared__ int sResL[SS_WARPS], sResR[SS_WARPS], sBin[SS_WARPS][2];
    __shared__ int sInd[3];

    const int f = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(nFeat[f], P.cap);
    const int nleft = nLeftArr ? min(max(nLeftArr[f], 0), n) : n;
    const bool stereoCam = nleft != n;    // Nleft != -1
    const int dir = dirArr ? dirArr[f] : 0;
    const int npts = min(nPts[f], P.ptsCap);
    const orbfe_keypoint* keys = keysAll + (size_t)f * P.cap;
    const uint8_t* desc = descAll + (size_t)f * P.cap * 32;
    const float* uRight = (uRightAll && !stereoCam && !Q.keyframe) ? uRightAll + (size_t)f * P.cap : nullptr;
    const uint8_t* occupied = occupiedAll ? occupiedAll + (size_t)f * P.cap : nullptr;
    const FramePoint* pts = ptsAll + (size_t)f * P.ptsCap;
    int* matchL = matchLAll + (size_t)f * P.ptsCap;
    int* matchR = matchRAll ? matchRAll + (size_t)f * P.ptsCap : nullptr;
    int* owner = ownerAll + (size_t)f * P.cap;
    const float gwInv = (float)SS_COLS / (P.maxX - P.minX);
    const float ghInv = (float)SS_ROWS / (P.maxY - P.minY);

    for (int i = tid; i < (P.cap + 31) >> 5; i += SS_THREADS) blocked[i] = 0;
    for (int i = tid; i < npts; i += SS_THREADS) {
        binL[i] = 0xffu;
        binR[i] = 0xffu;
    }
    if (tid < 32) hist[tid] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();
    for (int i = tid; i < n; i += SS_THREADS) {
        if (occupied && occupied[i]) atomicOr(&blocked[i >> 5], 1u << (i & 31));
        owner[i] = -1;
    }
    ss_build_grid(keys, 0, nleft, P, gwInv, ghInv, cellStartL, cellFill, itemsL, sScan, &sTotal);
    ss_build_grid(keys, nleft, n - nleft, P, gwInv, ghInv, cellStartR, cellFill, itemsR, sScan, &sTotal);

    auto eligible_of = [&](const FramePoint& q) {
        return (q.flags & 1) && q.level >= 0 && q.level < ORBFE_MAX_LEVELS && !(q.u < P.minX || q.u > P.maxX) &&
               !(q.v < P.minY || q.v > P.maxY);
    };
    for (int base = 0; base < npts; base += SS_WARPS) {
        const int iP = base + warp;
        const bool have = iP < npts;
        FramePoint mp;
        bool eligible = false;
        if (have) {
            mp = pts[iP];
            eligible = eligible_of(mp);
        }
        if (tid == 0) sOpen = (int)(0xffffffffu >> (32 - min(npts - base, SS_WARPS)));
        __syncthreads();
        for (;;) {
            if (have && ((sOpen >> warp) & 1)) {
                int resL = -1, resR = -1, bL = 0xff, bR = 0xff;
                if (eligible) {
                    const float radius = P.th * P.scale[mp.level];
                    SWindow w;
                    ss_window(w, P, gwInv, ghInv, mp.u, mp.v, radius, mp.level);
                    sf_levels(w, mp.level, dir, Q.keyframe);
                    const float urProj = mp.u - Q.mbf * mp.invz;    // :1786 (no FMA: --fmad=false)
                    bool any;
                    const unsigned long long k1 = sf_search(w, cellStartL, itemsL, keys, desc, 0, blocked, mp.desc, uRight, urProj, radius, &any);
                    if (any) {    // :1769-1770 an empty window skips the right camera too
                        if ((int)(k1 >> 32) <= Q.distTh) {
                            resL = (int)(k1 & 0xffffu);
                            if (Q.checkOrientation) bL = sf_bin(mp.angle, keys[resL].angle, Q.keyframe);
                        }
                        if (stereoCam && !Q.keyframe) {
                            SWindow wr;
                            ss_window(wr, P, gwInv, ghInv, mp.ur, mp.vr, radius, mp.level);
                            sf_levels(wr, mp.level, dir, 0);
                            bool anyR;
                            const unsigned long long k2 = sf_search(wr, cellStartR, itemsR, keys, desc, nleft, blocked, mp.desc, nullptr, 0.f, radius, &anyR);
                            if ((int)(k2 >> 32) <= Q.distTh) {
                                resR = (int)(k2 & 0xffffu);
                                if (Q.checkOrientation) bR = sf_bin(mp.angle, keys[nleft + resR].angle, 0);
                            }
                        }
                    }
                }
                if (lane == 0) {
                    sResL[warp] = resL;
                    sResR[warp] = resR;
                    sBin[warp][0] = bL;
                    sBin[warp][1] = bR;
                }
            }
            __syncthreads();
            if (warp == 0) {
                const int open = sOpen;
                const bool mine = lane < SS_WARPS && ((open >> lane) & 1);
                SWindow wl, wr;
                wl.cx0 = wr.cx0 = 1;
                wl.cx1 = wr.cx1 = 0;
                bool el = false, blocks = false;
                if (mine) {
                    const FramePoint q = pts[base + lane];
                    el = eligible_of(q);
                    blocks = Q.keyframe || (q.flags & 4);
                    if (el) {
                        const float radius = P.th * P.scale[q.level];
                        ss_window(wl, P, gwInv, ghInv, q.u, q.v, radius, q.level);
                        sf_levels(wl, q.level, dir, Q.keyframe);
                        ss_window(wr, P, gwInv, ghInv, q.ur, q.vr, radius, q.level);
                        sf_levels(wr, q.level, dir, 0);
                    }
                }
                // slots this point will make invisible to the later ones
                const int takeL = (mine && blocks) ? sResL[lane] : -1;
                const int takeR = (mine && blocks && sResR[lane] >= 0) ? sResR[lane] + nleft : -1;
                bool valid = true;
                for (int i = 0; i < SS_WARPS - 1; i++) {
                    const int tl = __shfl_sync(0xffffffffu, takeL, i), tr = __shfl_sync(0xffffffffu, takeR, i);
                    if (i < lane && mine && el) {
                        if (tl >= 0 && ss_inside(wl, keys[tl])) valid = false;    // conservative: ignores the other gates
                        if (tr >= 0 && stereoCam && ss_inside(wr, keys[tr])) valid = false;
                    }
                }
                const unsigned bad = __ballot_sync(0xffffffffu, mine && !valid);
                const int firstBad = bad ? __ffs(bad) - 1 : 32;
                const bool commit = mine && lane < firstBad;
                unsigned todo = __ballot_sync(0xffffffffu, commit);
                const unsigned done = todo;
                while (todo) {    // in point order: the last assignment of a slot wins (mvpMapPoints[slot] = pMP)
                    const int i = __ffs(todo) - 1;
                    todo &= todo - 1;
                    if (lane == i) {
                        const int rl = sResL[i], rr = sResR[i];
                        matchL[base + i] = rl;
                        if (matchR) matchR[base + i] = rr;
                        if (rl >= 0) {
                            owner[rl] = base + i;
                            if (blocks) blocked[rl >> 5] |= 1u << (rl & 31);
                            sCount++;
                            if (sBin[i][0] != 0xff) {
                                binL[base + i] = (uint8_t)sBin[i][0];
                                hist[sBin[i][0]]++;
                            }
                        }
                        if (rr >= 0) {
                            const int s = rr + nleft;
                            owner[s] = base + i;
                            if (blocks) blocked[s >> 5] |= 1u << (s & 31);
                            sCount++;
                            if (sBin[i][1] != 0xff) {
                                binR[base + i] = (uint8_t)sBin[i][1];
                                hist[sBin[i][1]]++;
                            }
                        }
                    }
                    __syncwarp();
                }
                if (lane == 0) sOpen = open & ~(int)done;
            }
            __syncthreads();
            if (sOpen == 0) break;
            __syncthreads();
        }
        __syncthreads();
    }
    if (Q.checkOrientation) {
        if (tid == 0) {    // ComputeThreeMaxima (:2061-2102)
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int i = 0; i < SI_HISTO_LEN; i++) {
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
        __syncthreads();    // every owner[] write of the loop is done: the removals below win
        int removed = 0;
        for (int i = tid; i < npts; i += SS_THREADS) {
            const int bl = binL[i], br = binR[i];
            if (bl != 0xff && bl != sInd[0] && bl != sInd[1] && bl != sInd[2]) {
                owner[matchL[i]] = -1;
                matchL[i] = -1;
                removed++;
            }
            if (br != 0xff && br != sInd[0] && br != sInd[1] && br != sInd[2]) {
                owner[matchR[i] + nleft] = -1;
                matchR[i] = -1;
                removed++;
            }
        }
        if (removed) atomicSub(&sCount, removed);
        __syncthreads();
    }
    if (tid == 0 && nMatches) nMatches[f] = sCount;
}

size_t search_frame_smem_bytes(int cap, int ptsCap) {
    return (size_t)(3 * SS_CELLS + 2 + 32) * 4 + (size_t)((cap + 31) / 32) * 4 + (size_t)cap * 4 + (size_t)ptsCap * 2 + 16;
}

// keyframe = 0: SearchByProjection(CurrentFrame, LastFrame, th, bMono); 1: SearchByProjection(CurrentFrame, pKF, ...)
int launch_search_frame(const float bounds[4], const float* scale, int nlevels, float th, float mbf, int keyframe, int distTh,
                        int checkOrientation, const orbfe_keypoint* keys, const uint8_t* desc, const float* uRight,
                        const uint8_t* occupied, const int* nFeat, const int* nLeft, const int* dir, int cap,
                        const void* pts, const int* nPts, int ptsCap, int nframes, int* matchL, int* matchR, int* owner,
                        int* nMatches, cudaStream_t st) {
    if (nframes <= 0) return 0;
    FrameSearchParams Q;
    StereoSearchParams& P = Q.g;
    P.minX = bounds[0];
    P.minY = bounds[1];
    P.maxX = bounds[2];
    P.maxY = bounds[3];
    for (int l = 0; l < ORBFE_MAX_LEVELS; l++) P.scale[l] = l < nlevels ? scale[l] : 1.0f;
    P.th = th;
    P.thFar = 0.f;
    P.nnRatio = 0.f;
    P.farPoints = 0;
    P.cap = cap;
    P.ptsCap = ptsCap;
    Q.mbf = mbf;
    Q.keyframe = keyframe;
    Q.distTh = distTh;
    Q.checkOrientation = checkOrientation;
    const size_t smem = search_frame_smem_bytes(cap, ptsCap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, search_frame_kernel, smem);
    search_frame_kernel<<<nframes, SS_THREADS, smem, st>>>(Q, keys, desc, uRight, occupied, nFeat, nLeft, dir,
                                                          (const FramePoint*)pts, nPts, matchL, matchR, owner, nMatches);
    return 1;
}

// ---- ORBmatcher::SearchForInitialization (ORBmatcher.cc:649-764) + ComputeThreeMaxima (:2061-2102) ----------------------
// Monocular initialisation: every level-0 keypoint of F1 searches F2's grid in a square of windowSize around
// vbPrevMatched[i1] (levels 0..0), keeps the best and the second-best DISTANCE, accepts with bestDist <= TH_LOW and
// bestDist < bestDist2 * mfNNratio, and a later keypoint steals a feature of F2 only with a strictly smaller distance
// (vMatchedDistance: candidates whose recorded distance is <= the current one are invisible, :690-691).  That state makes
// the loop order dependent; it is handled like the map-point loop above: one CTA per frame pair, one warp per keypoint,
// chunks searched against the current state, the longest prefix whose windows were not touched by an earlier commit of the
// same round is committed, the rest searches again.  best = min (dist, cell, index) is the first minimum in visiting
// order; bestDist2 is the second smallest distance with multiplicity, which is order free.
// The rotation histogram counts every accepted match, also the ones stolen later (rotHist keeps their i1, :724-733);
// bins are 0..12 only because factor = 1/HISTO_LENGTH (the reference's own arithmetic, reproduced).
constexpr int SI_TH_LOW = 50, SI_HISTO = 30;

__device__ __forceinline__ int ss_block_exclusive_scan(int s, int* sScan, int* sTotal) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    int inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
    }
    __syncthreads();    // sScan may still be read from an earlier use
    if (lane == 31) sScan[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const int w = lane < SS_WARPS ? sScan[lane] : 0;
        int winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, winc, o);
            if (lane >= o) winc += v;
        }
        if (lane < SS_WARPS) sScan[lane] = winc - w;
        if (lane == 31) *sTotal = winc;
    }
    __syncthreads();
    return sScan[warp] + inc - s;
}

// grid: nframes; block SS_THREADS; dynamic shared memory: search_init_smem_bytes(cap)
__global__ void __launch_bounds__(SS_THREADS) search_for_initialization_kernel(
    const __grid_constant__ StereoSearchParams P, float windowSize, int checkOrientation,
    const orbfe_keypoint* __restrict__ keys1All, const uint8_t* __restrict__ desc1All, const int* __restrict__ n1All,
    const orbfe_keypoint* __restrict__ keys2All, const uint8_t* __restrict__ desc2All, const int* __restrict__ n2All,
    float2* __restrict__ prevAll, int* __restrict__ match12All, int* __restrict__ nMatches) {
    extern __shared__ __align__(16) uint8_t smem[];
    int* cellStart = reinterpret_cast<int*>(smem);                                   // SS_CELLS + 1
    int* cellFill = cellStart + SS_CELLS + 1;                                        // SS_CELLS
    int* hist = cellFill + SS_CELLS;                                                 // 32
    unsigned short* items = reinterpret_cast<unsigned short*>(hist + 32);            // cap
    unsigned short* mdist = items + P.cap;                                           // vMatchedDistance (0xffff = INT_MAX)
    unsigned short* m21 = mdist + P.cap;                                             // vnMatches21 (0xffff = -1)
    unsigned short* list0 = m21 + P.cap;                                             // level-0 keypoints of F1, ascending
    uint8_t* binOf = reinterpret_cast<uint8_t*>(list0 + P.cap);                      // rotHist bin of i1 (0xff = none)
    __shared__ int sScan[SS_WARPS + 1];
    __shared__ int sRes[SS_WARPS], sDist[SS_WARPS];
    __shared__ int sOpen, sCount, sTotal;
    __shared__ int sInd[3];

    const int f = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n1 = min(n1All[f], P.cap), n2 = min(n2All[f], P.cap);
    const orbfe_keypoint* keys1 = keys1All + (size_t)f * P.cap;
    const orbfe_keypoint* keys2 = keys2All + (size_t)f * P.cap;
    const uint8_t* desc1 = desc1All + (size_t)f * P.cap * 32;
    const uint8_t* desc2 = desc2All + (size_t)f * P.cap * 32;
    float2* prev = prevAll + (size_t)f * P.cap;
    int* match12 = match12All + (size_t)f * P.cap;
    const float gwInv = (float)SS_COLS / (P.maxX - P.minX);
    const float ghInv = (float)SS_ROWS / (P.maxY - P.minY);

    ss_build_grid(keys2, 0, n2, P, gwInv, ghInv, cellStart, cellFill, items, sScan, &sTotal);
    for (int i = tid; i < P.cap; i += SS_THREADS) {
        mdist[i] = 0xffffu;
        m21[i] = 0xffffu;
        binOf[i] = 0xffu;
        match12[i] = -1;
    }
    if (tid < 32) hist[tid] = 0;
    if (tid == 0) sCount = 0;
    // level-0 keypoints of F1 in ascending index (:663-667)
    int n0;
    {
        const int per = (n1 + SS_THREADS - 1) / SS_THREADS;
        const int b = min(tid * per, n1), e = min(b + per, n1);
        int s = 0;
        for (int i = b; i < e; i++) s += keys1[i].octave <= 0;
        int at = ss_block_exclusive_scan(s, sScan, &sTotal);
        for (int i = b; i < e; i++)
            if (keys1[i].octave <= 0) list0[at++] = (unsigned short)i;
        n0 = sTotal;
    }
    __syncthreads();

    constexpr unsigned long long NONE = 0x7fffffffull << 32;
    for (int base = 0; base < n0; base += SS_WARPS) {
        const bool have = base + warp < n0;
        const int i1 = have ? list0[base + warp] : 0;
        SWindow w;
        w.cx0 = 1;
        w.cx1 = 0;
        uint32_t md[8];
        if (have) {
            const float2 c = prev[i1];
            ss_window(w, P, gwInv, ghInv, c.x, c.y, windowSize, 0);
            w.minLevel = 0;    // GetFeaturesInArea(x, y, windowSize, level1, level1), level1 == 0
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(desc1 + (size_t)i1 * 32));
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(desc1 + (size_t)i1 * 32) + 1);
            md[0] = a0.x; md[1] = a0.y; md[2] = a0.z; md[3] = a0.w;
            md[4] = a1.x; md[5] = a1.y; md[6] = a1.z; md[7] = a1.w;
        }
        if (tid == 0) sOpen = (int)(0xffffffffu >> (32 - min(n0 - base, SS_WARPS)));
        __syncthreads();
        for (;;) {
            if (have && ((sOpen >> warp) & 1)) {
                unsigned long long k1 = NONE;
                int d2 = 0x7fffffff;
                if (w.cx0 <= w.cx1) {
                    const int ncy = w.cy1 - w.cy0 + 1;
                    const int ncell = (w.cx1 - w.cx0 + 1) * ncy;
                    for (int c = lane; c < ncell; c += 32) {
                        const int cell = (w.cx0 + c / ncy) * SS_ROWS + w.cy0 + c % ncy;
                        for (int t = cellStart[cell]; t < cellStart[cell + 1]; t++) {
                            const int idx = items[t];
                            if (!ss_inside(w, keys2[idx])) continue;
                            const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(desc2 + (size_t)idx * 32));
                            const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(desc2 + (size_t)idx * 32) + 1);
                            const int d = __popc(md[0] ^ b0.x) + __popc(md[1] ^ b0.y) + __popc(md[2] ^ b0.z) +
                                          __popc(md[3] ^ b0.w) + __popc(md[4] ^ b1.x) + __popc(md[5] ^ b1.y) +
                                          __popc(md[6] ^ b1.z) + __popc(md[7] ^ b1.w);
                            if ((int)mdist[idx] <= d) continue;    // vMatchedDistance[i2] <= dist (:690-691)
                            const unsigned long long key = ((unsigned long long)d << 32) | ((uint32_t)cell << 16) | (uint32_t)idx;
                            if (key < k1) {
                                d2 = (int)(k1 >> 32);    // the old best is the new runner-up (it was <= d2)
                                k1 = key;
                            } else {
                                d2 = min(d2, d);
                            }
                        }
                    }
                }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const unsigned long long ok = __shfl_xor_sync(0xffffffffu, k1, o);
                    const int od = __shfl_xor_sync(0xffffffffu, d2, o);
                    const unsigned long long loser = ok > k1 ? ok : k1;
                    d2 = min(min(d2, od), (int)(loser >> 32));
                    k1 = ok < k1 ? ok : k1;
                }
                const int bestDist = (int)(k1 >> 32);
                int res = -1;
                if (bestDist <= SI_TH_LOW && (float)bestDist < (float)d2 * P.nnRatio) res = (int)(k1 & 0xffffu);
                if (lane == 0) {
                    sRes[warp] = res;
                    sDist[warp] = bestDist;
                }
            }
            __syncthreads();
            if (warp == 0) {
                const int open = sOpen;
                const bool mine = lane < SS_WARPS && ((open >> lane) & 1);
                SWindow lw;
                lw.x = lw.y = 0.f;
                lw.rr = windowSize;
                lw.minLevel = 0;
                lw.maxLevel = 0;
                int myRes = -1, my1 = 0;
                if (mine) {
                    my1 = list0[base + lane];
                    const float2 c = prev[my1];
                    lw.x = c.x;
                    lw.y = c.y;
                    myRes = sRes[lane];
                }
                bool valid = true;
                for (int i = 0; i < SS_WARPS - 1; i++) {
                    const int fi = __shfl_sync(0xffffffffu, myRes, i);
                    if (i < lane && fi >= 0 && mine && ss_inside(lw, keys2[fi])) valid = false;
                }
                const unsigned bad = __ballot_sync(0xffffffffu, mine && !valid);
                const int firstBad = bad ? __ffs(bad) - 1 : 32;
                const bool commit = mine && lane < firstBad;
                if (commit && myRes >= 0) {
                    const int old = m21[myRes];
                    if (old != 0xffff) match12[old] = -1;    // :705-709
                    match12[my1] = myRes;
                    m21[myRes] = (unsigned short)my1;
                    mdist[myRes] = (unsigned short)sDist[lane];
                    if (checkOrientation) {
                        float rot = keys1[my1].angle - keys2[myRes].angle;
                        if (rot < 0.0f) rot += 360.0f;
                        int bin = (int)roundf(rot * (1.0f / SI_HISTO));
                        if (bin == SI_HISTO) bin = 0;
                        bin = min(max(bin, 0), SI_HISTO - 1);
                        binOf[my1] = (uint8_t)bin;
                        atomicAdd(&hist[bin], 1);
                    }
                }
                const unsigned done = __ballot_sync(0xffffffffu, commit);
                if (lane == 0) sOpen = open & ~(int)done;
            }
            __syncthreads();
            if (sOpen == 0) break;
            __syncthreads();
        }
        __syncthreads();
    }

    if (checkOrientation) {
        if (tid == 0) {    // ComputeThreeMaxima
            int max1 = 0, max2 = 0, max3 = 0, ind1 = -1, ind2 = -1, ind3 = -1;
            for (int i = 0; i < SI_HISTO; i++) {
                const int s = hist[i];
                if (s > max1) {
                    max3 = max2; max2 = max1; max1 = s;
                    ind3 = ind2; ind2 = ind1; ind1 = i;
                } else if (s > max2) {
                    max3 = max2; max2 = s;
                    ind3 = ind2; ind2 = i;
                } else if (s > max3) {
                    max3 = s;
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
        for (int t = tid; t < n0; t += SS_THREADS) {
            const int i1 = list0[t];
            const int bin = binOf[i1];
            if (bin == 0xff || bin == sInd[0] || bin == sInd[1] || bin == sInd[2]) continue;
            if (match12[i1] >= 0) match12[i1] = -1;    // :740-747
        }
        __syncthreads();
    }
    int cnt = 0;
    for (int t = tid; t < n0; t += SS_THREADS) {    // :759-761
        const int i1 = list0[t];
        const int m = match12[i1];
        if (m >= 0) {
            prev[i1] = make_float2(keys2[m].x, keys2[m].y);
            cnt++;
        }
    }
    if (cnt) atomicAdd(&sCount, cnt);
    __syncthreads();
    if (tid == 0 && nMatches) nMatches[f] = sCount;
}

size_t search_init_smem_bytes(int cap) {
    return (size_t)(2 * SS_CELLS + 1 + 32) * 4 + (size_t)cap * 9 + 16;
}

int launch_search_for_initialization(const float bounds[4], int windowSize, float nnRatio, int checkOrientation,
                                     const orbfe_keypoint* keys1, const uint8_t* desc1, const int* n1,
                                     const orbfe_keypoint* keys2, const uint8_t* desc2, const int* n2, int cap, int nframes,
                                     float* prevMatched, int* match12, int* nMatches, cudaStream_t st) {
    if (nframes <= 0) return 0;
    StereoSearchParams P;
    P.minX = bounds[0];
    P.minY = bounds[1];
    P.maxX = bounds[2];
    P.maxY = bounds[3];
    for (int l = 0; l < ORBFE_MAX_LEVELS; l++) P.scale[l] = 1.0f;
    P.th = 1.0f;
    P.thFar = 0.f;
    P.nnRatio = nnRatio;
    P.farPoints = 0;
    P.cap = cap;
    P.ptsCap = cap;
    const size_t smem = search_init_smem_bytes(cap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, search_for_initialization_kernel, smem);
    search_for_initialization_kernel<<<nframes, SS_THREADS, smem, st>>>(P, (float)windowSize, checkOrientation, keys1, desc1,
                                                                        n1, keys2, desc2, n2,
                                                                        reinterpret_cast<float2*>(prevMatched), match12,
                                                                        nMatches);
    return 1;
}

}    // namespace orbfe
