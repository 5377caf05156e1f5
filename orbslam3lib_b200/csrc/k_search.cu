// k_search.cu — the frame grid and ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFarPoints,
// thFarPoints) (ORBmatcher.cc:44-143) for frames with Nleft == -1, on the device.
//
// Reference pieces: Frame::AssignFeaturesToGrid + PosInGrid (Frame.cc:405-436, 741-751: 64 x 48 cells, cell =
// round((pt - min) * inv), per-cell vectors filled in ascending feature index), Frame::GetFeaturesInArea (:672-736:
// cell range from floor / ceil, then level and |dx| < r, |dy| < r tests, cells visited x-major), RadiusByViewingCos
// (ORBmatcher.cc:216-222), DescriptorDistance (:2107-2123 = 256-bit popcount), and the map-point loop: best / second
// best with their levels in visiting order, TH_HIGH and the same-level ratio test, F.mvpMapPoints[bestIdx] = pMP.
//
// Two things make the loop order dependent, and both are reproduced exactly:
//   * ties and the level of the runner-up depend on the visiting order.  With pos = (cell << 16 | feature index) the
//     sequential update rule is equivalent to: best = min (dist, pos); runner-up = min (dist, pos) over the candidates
//     visited BEFORE the best if that distance is <= the minimum over those visited AFTER it, else the latter
//     (checked exhaustively against the literal loop in tests/test_search_rule.py);
//   * a feature taken by an earlier map point with Observations() > 0 is skipped by the later ones (:84-86).  A CTA
//     owns a frame and walks the points in chunks of one point per warp: all warps of a chunk search against the
//     current blocked set, then one warp validates the chunk in order — a point is final if no earlier point of the
//     same round took a feature inside its window — commits the longest valid prefix and the rest of the chunk
//     searches again.  The first open point of a round is always final, so the loop terminates; conflicts are rare,
//     so it almost always takes one round.
#include "orbfe_internal.h"

namespace orbfe {

#ifndef ORBFE_SEARCH_THREADS
#define ORBFE_SEARCH_THREADS 256
#endif
constexpr int SR_THREADS = ORBFE_SEARCH_THREADS;
constexpr int SR_WARPS = SR_THREADS / 32;
constexpr int SR_ROWS = 48, SR_COLS = 64, SR_CELLS = SR_ROWS * SR_COLS;    // FRAME_GRID_ROWS / COLS, Frame.h:46-47
constexpr int SR_TH_HIGH = 100;                                           // ORBmatcher::TH_HIGH

struct SearchParams {
    float minX, minY, maxX, maxY;          // mnMinX ...
    float scale[ORBFE_MAX_LEVELS];         // mvScaleFactors
    float th, thFar, nnRatio;
    int farPoints;
    int cap, ptsCap;                       // per-frame strides of the feature / point arrays
};

struct TrackPoint {    // == orbfe_track_point
    float projX, projY, projXR, viewCos, trackDepth;
    int level, flags;
    uint32_t desc[8];
};
static_assert(sizeof(TrackPoint) == 60, "orbfe_track_point layout");

struct Window {
    float x, y, rr, rur;    // centre, r * scale (search radius), r * scale again for the uRight gate
    int minLevel, maxLevel;
    int cx0, cx1, cy0, cy1;    // cell range; cx0 > cx1 = empty
};

__device__ __forceinline__ unsigned long long shfl_min_u64(unsigned long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, v, o);
        v = other < v ? other : v;
    }
    return v;
}

__device__ __forceinline__ bool in_square_and_level(const Window& w, const orbfe_keypoint& k) {
    if (k.octave < w.minLevel) return false;    // bCheckLevels is always true here: maxLevel = level >= 0
    if (w.maxLevel >= 0 && k.octave > w.maxLevel) return false;
    const float distx = k.x - w.x, disty = k.y - w.y;
    return fabsf(distx) < w.rr && fabsf(disty) < w.rr;
}

// grid: nframes; block SR_THREADS; dynamic shared memory: see search_smem_bytes.
__global__ void __launch_bounds__(SR_THREADS) search_by_projection_kernel(
    const __grid_constant__ SearchParams P, const orbfe_keypoint* __restrict__ keysAll, const uint8_t* __restrict__ descAll,
    const float* __restrict__ uRightAll, const uint8_t* __restrict__ occupiedAll, const int* __restrict__ nFeat,
    const TrackPoint* __restrict__ ptsAll, const int* __restrict__ nPts, int* __restrict__ matchAll,
    int* __restrict__ ownerAll, int* __restrict__ nMatches) {
    extern __shared__ __align__(16) uint8_t smem[];
    int* cellStart = reinterpret_cast<int*>(smem);                        // SR_CELLS + 1
    int* cellFill = cellStart + SR_CELLS + 1;                             // SR_CELLS
    uint32_t* blocked = reinterpret_cast<uint32_t*>(cellFill + SR_CELLS);    // cap / 32 words
    unsigned short* items = reinterpret_cast<unsigned short*>(blocked + ((P.cap + 31) >> 5));    // cap
    __shared__ int sScan[SR_WARPS + 1];
    __shared__ int sRes[SR_WARPS];     // tentative feature of the chunk's points (-1 none)
    __shared__ int sOpen;              // bit w: point of warp w still open
    __shared__ int sCount, sTotal;

    const int f = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = min(nFeat[f], P.cap);
    const int npts = min(nPts[f], P.ptsCap);
    const orbfe_keypoint* keys = keysAll + (size_t)f * P.cap;
    const uint8_t* desc = descAll + (size_t)f * P.cap * 32;
    const float* uRight = uRightAll ? uRightAll + (size_t)f * P.cap : nullptr;
    const uint8_t* occupied = occupiedAll ? occupiedAll + (size_t)f * P.cap : nullptr;
    const TrackPoint* pts = ptsAll + (size_t)f * P.ptsCap;
    int* match = matchAll + (size_t)f * P.ptsCap;
    int* owner = ownerAll ? ownerAll + (size_t)f * P.cap : nullptr;

    const float gwInv = (float)SR_COLS / (P.maxX - P.minX);    // mfGridElementWidthInv, Frame.cc:179
    const float ghInv = (float)SR_ROWS / (P.maxY - P.minY);

    // ---- AssignFeaturesToGrid: counting sort by cell, ascending feature index inside a cell ----------------------------
    for (int i = tid; i <= SR_CELLS; i += SR_THREADS) cellStart[i] = 0;
    for (int i = tid; i < SR_CELLS; i += SR_THREADS) cellFill[i] = 0;
    for (int i = tid; i < (P.cap + 31) >> 5; i += SR_THREADS) blocked[i] = 0;
    if (tid == 0) sCount = 0;
    __syncthreads();
    auto cell_of = [&](int i) -> int {
        const orbfe_keypoint k = keys[i];
        const int posX = (int)roundf((k.x - P.minX) * gwInv);
        const int posY = (int)roundf((k.y - P.minY) * ghInv);
        if (posX < 0 || posX >= SR_COLS || posY < 0 || posY >= SR_ROWS) return -1;
        return posX * SR_ROWS + posY;
    };
    for (int i = tid; i < n; i += SR_THREADS) {
        const int c = cell_of(i);
        if (c >= 0) atomicAdd(&cellStart[c], 1);
        if (occupied && occupied[i]) atomicOr(&blocked[i >> 5], 1u << (i & 31));
        if (owner) owner[i] = -1;
    }
    __syncthreads();
    {    // exclusive scan of SR_CELLS counts: 3 cells per thread
        constexpr int PER = (SR_CELLS + SR_THREADS - 1) / SR_THREADS;
        const int b = min(tid * PER, SR_CELLS), e = min(b + PER, SR_CELLS);
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
            const int w = lane < SR_WARPS ? sScan[lane] : 0;
            int winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, winc, o);
                if (lane >= o) winc += v;
            }
            if (lane < SR_WARPS) sScan[lane] = winc - w;
            if (lane == 31) sTotal = winc;
        }
        __syncthreads();
        int run = sScan[warp] + inc - s;
        for (int i = b; i < e; i++) {
            const int c = cellStart[i];
            cellStart[i] = run;
            run += c;
        }
        if (tid == SR_THREADS - 1) cellStart[SR_CELLS] = sTotal;
    }
    __syncthreads();
    for (int i = tid; i < n; i += SR_THREADS) {
        const int c = cell_of(i);
        if (c >= 0) items[cellStart[c] + atomicAdd(&cellFill[c], 1)] = (unsigned short)i;
    }
    __syncthreads();
    for (int c = tid; c < SR_CELLS; c += SR_THREADS) {    // vectors are filled in ascending i (:420-434)
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

    // ---- the map-point loop ----------------------------------------------------------------------------------------
    const bool bFactor = P.th != 1.0f;
    for (int base = 0; base < npts; base += SR_WARPS) {
        const int iMP = base + warp;
        const bool have = iMP < npts;
        TrackPoint mp;
        Window w;
        w.cx0 = 1;
        w.cx1 = 0;
        bool searchable = false;
        float r = 0.f;
        if (have) {
            mp = pts[iMP];
            searchable = (mp.flags & 1) && !(P.farPoints && mp.trackDepth > P.thFar) && !(mp.flags & 2) && mp.level >= 0 &&
                         mp.level < ORBFE_MAX_LEVELS;
            if (searchable) {
                r = (double)mp.viewCos > 0.998 ? 2.5f : 4.0f;    // RadiusByViewingCos
                if (bFactor) r *= P.th;
                w.x = mp.projX;
                w.y = mp.projY;
                w.rr = r * P.scale[mp.level];
                w.rur = w.rr;
                w.minLevel = mp.level - 1;
                w.maxLevel = mp.level;
                const int nMinCellX = max(0, (int)floorf((w.x - P.minX - w.rr) * gwInv));
                const int nMaxCellX = min(SR_COLS - 1, (int)ceilf((w.x - P.minX + w.rr) * gwInv));
                const int nMinCellY = max(0, (int)floorf((w.y - P.minY - w.rr) * ghInv));
                const int nMaxCellY = min(SR_ROWS - 1, (int)ceilf((w.y - P.minY + w.rr) * ghInv));
                if (!(nMinCellX >= SR_COLS || nMaxCellX < 0 || nMinCellY >= SR_ROWS || nMaxCellY < 0)) {
                    w.cx0 = nMinCellX;
                    w.cx1 = nMaxCellX;
                    w.cy0 = nMinCellY;
                    w.cy1 = nMaxCellY;
                }
            }
        }
        if (tid == 0) sOpen = (int)(0xffffffffu >> (32 - min(npts - base, SR_WARPS)));
        __syncthreads();
        for (;;) {
            // -- search: every open point against the current blocked set --
            if (have && ((sOpen >> warp) & 1)) {
                int res = -1;
                if (searchable && w.cx0 <= w.cx1) {
                    const int ncy = w.cy1 - w.cy0 + 1;
                    const int ncell = (w.cx1 - w.cx0 + 1) * ncy;
                    constexpr unsigned long long NONE = 256ull << 32;
                    auto scan = [&](unsigned long long bestKey, unsigned long long& kA, unsigned long long& kB) {
                        // pass 1 (bestKey == NONE): kA = min key.  pass 2: kA = min key before bestKey's pos, kB = after.
                        kA = NONE;
                        kB = NONE;
                        const uint32_t bestPos = (uint32_t)bestKey;
                        for (int c = lane; c < ncell; c += 32) {
                            const int ix = w.cx0 + c / ncy, iy = w.cy0 + c % ncy;
                            const int cell = ix * SR_ROWS + iy;
                            for (int t = cellStart[cell]; t < cellStart[cell + 1]; t++) {
                                const int idx = items[t];
                                const orbfe_keypoint k = keys[idx];
                                if (!in_square_and_level(w, k)) continue;
                                if ((blocked[idx >> 5] >> (idx & 31)) & 1u) continue;    // :84-86
                                if (uRight) {
                                    const float ur = uRight[idx];
                                    if (ur > 0) {
                                        const float er = fabsf(mp.projXR - ur);    // :88-93
                                        if (er > w.rur) continue;
                                    }
                                }
                                const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(desc + (size_t)idx * 32));
                                const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(desc + (size_t)idx * 32) + 1);
                                const int d = __popc(mp.desc[0] ^ b0.x) + __popc(mp.desc[1] ^ b0.y) + __popc(mp.desc[2] ^ b0.z) +
                                              __popc(mp.desc[3] ^ b0.w) + __popc(mp.desc[4] ^ b1.x) + __popc(mp.desc[5] ^ b1.y) +
                                              __popc(mp.desc[6] ^ b1.z) + __popc(mp.desc[7] ^ b1.w);
                                const uint32_t pos = ((uint32_t)cell << 16) | (uint32_t)idx;
                                const unsigned long long key = ((unsigned long long)d << 32) | pos;
                                if (bestKey == NONE) {
                                    kA = key < kA ? key : kA;
                                } else if (pos < bestPos) {
                                    kA = key < kA ? key : kA;
                                } else if (pos > bestPos) {
                                    kB = key < kB ? key : kB;
                                }
                            }
                        }
                        kA = shfl_min_u64(kA);
                        kB = shfl_min_u64(kB);
                    };
                    unsigned long long k1, kP, kS, dummy;
                    scan(NONE, k1, dummy);
                    const int bestDist = (int)(k1 >> 32);
                    if (bestDist <= SR_TH_HIGH) {    // warp uniform; bestDist < 256
                        scan(k1, kP, kS);
                        const int bestIdx = (int)(k1 & 0xffffu);
                        const int bestLevel = keys[bestIdx].octave;
                        const int dP = (int)(kP >> 32), dS = (int)(kS >> 32);
                        int bestDist2 = 256, bestLevel2 = -1;
                        if (dP < 256 && dP <= dS) {
                            bestDist2 = dP;
                            bestLevel2 = keys[(int)(kP & 0xffffu)].octave;
                        } else if (dS < 256) {
                            bestDist2 = dS;
                            bestLevel2 = keys[(int)(kS & 0xffffu)].octave;
                        }
                        const float lim = P.nnRatio * (float)bestDist2;    // mfNNratio * bestDist2
                        const bool reject = bestLevel == bestLevel2 && (float)bestDist > lim;
                        if (!reject && (bestLevel != bestLevel2 || (float)bestDist <= lim)) res = bestIdx;
                    }
                }
                if (lane == 0) sRes[warp] = res;
            }
            __syncthreads();
            // -- validation + commit of the longest valid prefix (warp 0; lane l = point base + l) --
            if (warp == 0) {
                const int open = sOpen;
                const bool mine = (open >> lane) & 1;
                // every lane needs ITS point's window: reload (cheap, 32 points)
                Window lw;
                lw.cx0 = 1;
                lw.cx1 = 0;
                int myRes = -1, myFlags = 0;
                if (mine) {
                    const TrackPoint q = pts[base + lane];
                    myRes = sRes[lane];
                    myFlags = q.flags;
                    float rq = (double)q.viewCos > 0.998 ? 2.5f : 4.0f;
                    if (bFactor) rq *= P.th;
                    const int lv = min(max(q.level, 0), ORBFE_MAX_LEVELS - 1);
                    lw.x = q.projX;
                    lw.y = q.projY;
                    lw.rr = rq * P.scale[lv];
                    lw.minLevel = q.level - 1;
                    lw.maxLevel = q.level;
                }
                const int takes = (mine && myRes >= 0 && (myFlags & 4)) ? myRes : -1;    // feature this point will block
                bool valid = true;
                for (int i = 0; i < 31; i++) {
                    const int fi = __shfl_sync(0xffffffffu, takes, i);
                    if (i < lane && fi >= 0 && mine) {
                        if (in_square_and_level(lw, keys[fi])) valid = false;    // conservative: ignores the other gates
                    }
                }
                const unsigned bad = __ballot_sync(0xffffffffu, mine && !valid);
                const int firstBad = bad ? __ffs(bad) - 1 : 32;
                const bool commit = mine && lane < firstBad;
                if (commit) {
                    match[base + lane] = myRes;
                    if (myRes >= 0) {
                        if (owner) atomicMax(&owner[myRes], base + lane);    // the last assignment wins (:128)
                        if (myFlags & 4) atomicOr(&blocked[myRes >> 5], 1u << (myRes & 31));
                        atomicAdd(&sCount, 1);
                    }
                }
                const unsigned done = __ballot_sync(0xffffffffu, commit);
                if (lane == 0) sOpen = open & ~(int)done;
            }
            __syncthreads();
            if (sOpen == 0) break;
            __syncthreads();    // sOpen is rewritten by warp 0 in the next round only after everyone has read it
        }
        __syncthreads();
    }
    if (tid == 0 && nMatches) nMatches[f] = sCount;
}

size_t search_smem_bytes(int cap) {
    return (size_t)(2 * SR_CELLS + 1) * 4 + (size_t)((cap + 31) / 32) * 4 + (size_t)cap * 2 + 16;
}

// All pointers are device memory.  Returns the kernels launched.
int launch_search_by_projection(const float bounds[4], const float* scale, int nlevels, float th, int farPoints,
                                float thFar, float nnRatio, const orbfe_keypoint* keys, const uint8_t* desc,
                                const float* uRight, const uint8_t* occupied, const int* nFeat, int cap, const void* pts,
                                const int* nPts, int ptsCap, int nframes, int* match, int* owner, int* nMatches,
                                cudaStream_t st) {
    if (nframes <= 0) return 0;
    SearchParams P;
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
    const size_t smem = search_smem_bytes(cap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, search_by_projection_kernel, smem);
    search_by_projection_kernel<<<nframes, SR_THREADS, smem, st>>>(P, keys, desc, uRight, occupied, nFeat,
                                                                  (const TrackPoint*)pts, nPts, match, owner, nMatches);
    return 1;
}

// ---- Frame::UndistortKeyPoints (Frame.cc:763-796) --------------------------------------------------------------------
// cv::undistortPoints(pts, pts, K, mDistCoef, Mat(), mK) = OpenCV 4.x cvUndistortPointsInternal with the default
// TermCriteria(MAX_ITER, 5, 0.01): normalise with K, five fixed-point iterations of the radial / tangential model in
// double, re-project with P = mK, round to float.  One thread per keypoint; fp64, no FMA (--fmad=false), operation
// order as in OpenCV (the zero terms of P are kept: they are exact).
struct UndistortParams {
    double fx, fy, cx, cy;
    double k[14];
    int identity;    // mDistCoef[0] == 0: mvKeysUn = mvKeys (:765-769)
};

__global__ void __launch_bounds__(256) undistort_keypoints_kernel(const __grid_constant__ UndistortParams P,
                                                                  const orbfe_keypoint* __restrict__ keys, int n,
                                                                  orbfe_keypoint* __restrict__ keysUn) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= n) return;
    orbfe_keypoint kp = keys[i];
    if (!P.identity) {
        const double* k = P.k;
        const double ifx = 1. / P.fx, ify = 1. / P.fy;
        double x = kp.x, y = kp.y;
        const double u = x, v = y;
        x = (x - P.cx) * ifx;
        y = (y - P.cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) {
                x = (u - P.cx) * ifx;
                y = (v - P.cy) * ify;
                break;
            }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = P.fx * x + 0.0 * y + P.cx;
        const double yy = 0.0 * x + P.fy * y + P.cy;
        const double ww = 1. / (0.0 * x + 0.0 * y + 1.0);
        kp.x = (float)(xx * ww);
        kp.y = (float)(yy * ww);
    }
    keysUn[i] = kp;
}

int launch_undistort(const orbfe_keypoint* keys, int n, float fx, float fy, float cx, float cy, const float* dist,
                     int ndist, orbfe_keypoint* keysUn, cudaStream_t st) {
    if (n <= 0) return 0;
    UndistortParams P;
    P.fx = fx;
    P.fy = fy;
    P.cx = cx;
    P.cy = cy;
    for (int i = 0; i < 14; i++) P.k[i] = i < ndist ? (double)dist[i] : 0.0;
    P.identity = ndist <= 0 || dist[0] == 0.0;
    undistort_keypoints_kernel<<<(n + 255) / 256, 256, 0, st>>>(P, keys, n, keysUn);
    return 1;
}

}    // namespace orbfe
