// search_grid.cuh — shared pieces of the guided matchers that work on the 64 x 48 frame grid: launch shape, the frame
// parameters, the search window (Frame::GetFeaturesInArea / KeyFrame::GetFeaturesInArea cell range), the block-wide
// counting sort that rebuilds Frame::AssignFeaturesToGrid in shared memory, and the 64-bit warp minimum.
// Included by k_search_stereo.cu and k_search_kf.cu (device functions cannot be called across translation units
// without relocatable device code).
#pragma once
#include "orbfe_internal.h"

namespace orbfe {

constexpr int SS_THREADS = 256;
constexpr int SS_WARPS = SS_THREADS / 32;
constexpr int SS_ROWS = 48, SS_COLS = 64, SS_CELLS = SS_ROWS * SS_COLS;
constexpr int SS_TH_HIGH = 100;
constexpr int SI_HISTO_LEN = 30;    // HISTO_LENGTH, ORBmatcher.cc:38

struct StereoSearchParams {
    float minX, minY, maxX, maxY;
    float scale[ORBFE_MAX_LEVELS];
    float th, thFar, nnRatio;
    int farPoints;
    int cap, ptsCap;    // per-frame strides of the feature (N slots) and point arrays
};

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
static __device__ void ss_build_grid(const orbfe_keypoint* keys, int first, int count, const StereoSearchParams& P, float gwInv,
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

}    // namespace orbfe
