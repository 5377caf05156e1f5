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
#include "orbfe_internal.h"

namespace orbfe {

constexpr int SS_THREADS = 256;
constexpr int SS_WARPS = SS_THREADS / 32;
constexpr int SS_ROWS = 48, SS_COLS = 64, SS_CELLS = SS_ROWS * SS_COLS;
constexpr int SS_TH_HIGH = 100;

struct StereoSearchParams {
    float minX, minY, maxX, maxY;
    float scale[ORBFE_MAX_LEVELS];
    float th, thFar, nnRatio;
    int farPoints;
    int cap, ptsCap;    // per-frame strides of the feature (N slots) and point arrays
};

struct TrackPointStereo {    // == orbfe_track_point_stereo
    float projX, projY, projXR, projYR, viewCos, viewCosR, trackDepth;
    int level, levelR, flags;
    uint32_t desc[8];
};
static_assert(sizeof(TrackPointStereo) == 72, "orbfe_track_point_stereo layout");

struct SWindow {
    float x, y, rr;
    int minLevel, maxLevel;
    int cx0, cx1, cy0, cy1;    // cx0 > cx1: empty
};

__device__ __forceinline__ unsigned long long ss_min_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, v, o);
        v = other < v ? other : v;
    }
    return v;
}

__device__ __forceinline__ bool ss_inside(const SWindow& w, const orbfe_keypoint& k) {
    if (k.octave < w.minLevel) return false;
    if (w.maxLevel >= 0 && k.octave > w.maxLevel) return false;
    const float distx = k.x - w.x, disty = k.y - w.y;
    return fabsf(distx) < w.rr && fabsf(disty) < w.rr;
}

// Frame::GetFeaturesInArea cell range (Frame.cc:681-703)
__device__ __forceinline__ void ss_window(SWindow& w, const StereoSearchParams& P, float gwInv, float ghInv, float x, float y,
                                          float rr, int level) {
    w.x = x;
    w.y = y;
    w.rr = rr;
    w.minLevel = level - 1;
    w.maxLevel = level;
    w.cx0 = 1;
    w.cx1 = 0;
    w.cy0 = 0;
    w.cy1 = 0;
    const int nMinCellX = max(0, (int)floorf((x - P.minX - rr) * gwInv));
    const int nMaxCellX = min(SS_COLS - 1, (int)ceilf((x - P.minX + rr) * gwInv));
    const int nMinCellY = max(0, (int)floorf((y - P.minY - rr) * ghInv));
    const int nMaxCellY = min(SS_ROWS - 1, (int)ceilf((y - P.minY + rr) * ghInv));
    if (!(nMinCellX >= SS_COLS || nMaxCellX < 0 || nMinCellY >= SS_ROWS || nMaxCellY < 0)) {
        w.cx0 = nMinCellX;
        w.cx1 = nMaxCellX;
        w.cy0 = nMinCellY;
        w.cy1 = nMaxCellY;
    }
}

// block-cooperative counting sort of features [first, first + count) into cells; items hold indices relative to `first`
__device__ void ss_build_grid(const orbfe_keypoint* keys, int first, int count, const StereoSearchParams& P, float gwInv,
                              float ghInv, int* cellStart, int* cellFill, unsigned short* items, int* sScan, int* sTotal) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i <= SS_CELLS; i += SS_THREADS) cellStart[i] = 0;
    for (int i = tid; i < SS_CELLS; i += SS_THREADS) cellFill[i] = 0;
    __syncthreads();
    auto cell_of = [&](int i) -> int {
        const orbfe_keypoint k = keys[first + i];
        const int posX = (int)roundf((k.x - P.minX) * gwInv);
        const int posY = (int)roundf((k.y - P.minY) * ghInv);
        if (posX < 0 || posX >= SS_COLS || posY < 0 || posY >= SS_ROWS) return -1;
        return posX * SS_ROWS + posY;
    };
    for (int i = tid; i < count; i += SS_THREADS) {
        const int c = cell_of(i);
        if (c >= 0) atomicAdd(&cellStart[c], 1);
    }
    __syncthreads();
    {
        constexpr int PER = (SS_CELLS + SS_THREADS - 1) / SS_THREADS;
        const int b = min(tid * PER, SS_CELLS), e = min(b + PER, SS_CELLS);
        int s = 0;
        for (int i = b; i < e; i++) s += cellStart[i];
        int inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
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
        int run = sScan[warp] + inc - s;
        for (int i = b; i < e; i++) {
            const int c = cellStart[i];
            cellStart[i] = run;
            run += c;
        }
        if (tid == SS_THREADS - 1) cellStart[SS_CELLS] = *sTotal;
    }
    __syncthreads();
    for (int i = tid; i < count; i += SS_THREADS) {
        const int c = cell_of(i);
        if (c >= 0) items[cellStart[c] + atomicAdd(&cellFill[c], 1)] = (unsigned short)i;
    }
    __syncthreads();
    for (int c = tid; c < SS_CELLS; c += SS_THREADS) {    // ascending index inside a cell (push_back order)
        const int b = cellStart[c], e = cellStart[c + 1];
        for (int i = b + 1; i < e; i++) {
            const unsigned short v = items[i];
            int j = i - 1;
            while (j >= b && items[j] > v) {
                items[j + 1] = items[j];
                j--;
            }
            items[j + 1] = v;
        }
    }
    __syncthreads();
}

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
    static size_t configured = 0;
    if (smem > 48 * 1024 && configured < smem) {
        cudaFuncSetAttribute(search_by_projection_stereo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        configured = smem;
    }
    search_by_projection_stereo_kernel<<<nframes, SS_THREADS, smem, st>>>(
        P, keys, desc, occupied, nFeat, nLeft, ltr, rtl, (const TrackPointStereo*)pts, nPts, matchL, matchR, owner, nMatches);
    return 1;
}

}    // namespace orbfe
