// k_stereo.cu — Frame::ComputeStereoMatches (Frame.cc:827-997) on the device-resident results of a stereo chunk.
//
// The reference builds a row table of the right keypoints (:836-854: keypoint iR is listed in rows
// floor(y - r) .. ceil(y + r), r = 2 * mvScaleFactors[octave]), then for every left keypoint walks the list of its row
// (int)vL in ascending iR order and keeps the first strictly smaller Hamming distance among the candidates with
// octave within +-1 and uR in [uL - maxD, uL] (:868-910); a best distance below (TH_HIGH + TH_LOW) / 2 = 75 is
// refined by an 11x11 SAD over +-5 px on the keypoint's pyramid level of both images (:914-949), a parabola fit
// (:954-962) and turned into (uRight, depth) (:964-979); finally matches whose SAD is >= 1.5 * 1.4 * median are
// dropped (:983-996).
//
// Mapping: the row table is never materialised.  "iR is in the list of row v" is the predicate minr <= v <= maxr, and
// "first strictly smaller in ascending iR" is the minimum of (distance << 16 | iR), so one warp per left keypoint scans
// the right keypoints of its pair (band / octave / x tables staged once per CTA in shared memory) 32 at a time and
// reduces with shuffles.  The SAD gives one window row to each of 11 lanes (11 + 21 byte loads from the level planes
// that the extractor left in HBM — mvImagePyramid never leaves the device), the 11 column sums are reduced across
// lanes, and lane 0 does the float arithmetic exactly as written in the reference (IEEE fp32, no FMA: the library is
// built with --fmad=false).  A second kernel per pair selects the median SAD with a two-level radix histogram and
// applies the cut.  Images of a chunk are interleaved: pair p = images 2p (left), 2p + 1 (right).
#include "orbfe_internal.h"

namespace orbfe {

constexpr int ST_THREADS = 256;
constexpr int ST_WARPS = ST_THREADS / 32;
constexpr int ST_LEFT_PER_WARP = 8;
constexpr int ST_LEFT_PER_CTA = ST_WARPS * ST_LEFT_PER_WARP;
constexpr int TH_HIGH = 100, TH_LOW = 50;    // ORBmatcher::TH_HIGH / TH_LOW, ORBmatcher.cc:35-36

struct StereoTables {
    float scale[ORBFE_MAX_LEVELS];       // mvScaleFactors
    float invScale[ORBFE_MAX_LEVELS];    // mvInvScaleFactors
};

__device__ __forceinline__ const uint8_t* level_px(const Geom& g, const uint8_t* pyr, int img, int level, int x, int y) {
    const LevelGeom& L = g.lv[level];
    return pyr + (size_t)img * g.slabBytes + L.planeOff + (size_t)(y + kEdge) * L.pitch + kXOff + x;
}

// grid (ceil(cap / ST_LEFT_PER_CTA), npairs); dynamic shared memory: 12 bytes per right keypoint slot (cap).
__global__ void __launch_bounds__(ST_THREADS) stereo_match_kernel(const __grid_constant__ Geom g,
                                                                  const __grid_constant__ StereoTables tb,
                                                                  const uint8_t* __restrict__ pyr,
                                                                  const orbfe_keypoint* __restrict__ kps,
                                                                  const uint8_t* __restrict__ desc,
                                                                  const ImageResult* __restrict__ result, int cap,
                                                                  float mb, float mbf, float* __restrict__ uRight,
                                                                  float* __restrict__ depth, int* __restrict__ sadOut) {
    extern __shared__ __align__(16) uint8_t smem[];
    float* sX = reinterpret_cast<float*>(smem);                    // kpR.pt.x
    int* sBand = reinterpret_cast<int*>(smem + (size_t)cap * 4);    // minr | maxr << 16 (both clamped to 16 bits)
    int* sOct = reinterpret_cast<int*>(smem + (size_t)cap * 8);
    const int pair = blockIdx.y;
    const int imgL = 2 * pair, imgR = 2 * pair + 1;
    const int nL = min(result[imgL].count, cap), nR = min(result[imgR].count, cap);
    const int l0 = blockIdx.x * ST_LEFT_PER_CTA;
    if (l0 >= nL) return;
    const orbfe_keypoint* kL = kps + (size_t)imgL * cap;
    const orbfe_keypoint* kR = kps + (size_t)imgR * cap;
    const uint8_t* dL = desc + (size_t)imgL * cap * 32;
    const uint8_t* dR = desc + (size_t)imgR * cap * 32;

    for (int i = threadIdx.x; i < nR; i += ST_THREADS) {    // :845-854
        const orbfe_keypoint k = kR[i];
        const float r = 2.0f * tb.scale[k.octave];
        const int maxr = (int)ceilf(k.y + r);
        const int minr = (int)floorf(k.y - r);
        sX[i] = k.x;
        sBand[i] = maxr < 0 ? 1 : (min(max(minr, 0), 0x7fff) | (min(maxr, 0x7fff) << 16));    // maxr < 0: empty band
        sOct[i] = k.octave;
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float maxD = mbf / mb;    // :857-859 (minZ = mb, minD = 0)
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    float* uOut = uRight + (size_t)pair * cap;
    float* zOut = depth + (size_t)pair * cap;
    int* sOutp = sadOut + (size_t)pair * cap;

    for (int k = 0; k < ST_LEFT_PER_WARP; k++) {
        const int iL = l0 + warp + k * ST_WARPS;
        if (iL >= nL) break;    // warp uniform
        const orbfe_keypoint kp = kL[iL];
        const int levelL = kp.octave;
        const float vL = kp.y, uL = kp.x;
        float outU = -1.0f, outZ = -1.0f;
        int outSad = -1;
        const int row = (int)vL;
        const float minU = uL - maxD, maxU = uL;
        uint32_t key = ((uint32_t)TH_HIGH << 16);
        if (!(maxU < 0)) {
            const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(dL + (size_t)iL * 32));
            const uint4 a1 = __ldg(reinterpret_cast<const uint4*>(dL + (size_t)iL * 32) + 1);
            for (int iR = lane; iR < nR; iR += 32) {
                const int band = sBand[iR];
                if (row < (band & 0xffff) || row > (band >> 16)) continue;
                const int o = sOct[iR];
                if (o < levelL - 1 || o > levelL + 1) continue;
                const float uR = sX[iR];
                if (!(uR >= minU && uR <= maxU)) continue;
                const uint4 b0 = __ldg(reinterpret_cast<const uint4*>(dR + (size_t)iR * 32));
                const uint4 b1 = __ldg(reinterpret_cast<const uint4*>(dR + (size_t)iR * 32) + 1);
                const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                              __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
                key = min(key, ((uint32_t)d << 16) | (uint32_t)iR);
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, o));
        }
        const int bestDist = (int)(key >> 16);
        if (bestDist < thOrbDist) {    // warp uniform from here on
            const int bestIdxR = (int)(key & 0xffffu);
            const float uR0 = sX[bestIdxR];
            const float sf = tb.invScale[levelL];
            const float scaleduL = roundf(uL * sf);
            const float scaledvL = roundf(vL * sf);
            const float scaleduR0 = roundf(uR0 * sf);
            const int w = 5, Lr = 5;
            const LevelGeom& LG = g.lv[levelL];
            const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w);
            // windows outside mvImagePyramid[level] are a cv::Exception in the reference: no match here
            const bool inside = y0 >= 0 && y0 + 2 * w + 1 <= LG.h && xl0 >= 0 && xl0 + 2 * w + 1 <= LG.w &&
                                (int)(scaleduR0 - Lr - w) >= 0;
            const float iniu = scaleduR0 + Lr - w;
            const float endu = scaleduR0 + Lr + w + 1;
            if (inside && !(iniu < 0 || endu >= (float)LG.w)) {
                const int xr0 = (int)(scaleduR0) - Lr - w;    // first column of the 21-px right row span
                int part[2 * 5 + 1];
#pragma unroll
                for (int i = 0; i < 11; i++) part[i] = 0;
                if (lane < 2 * w + 1) {
                    const uint8_t* pl = level_px(g, pyr, imgL, levelL, xl0, y0 + lane);
                    const uint8_t* pr = level_px(g, pyr, imgR, levelL, xr0, y0 + lane);
                    int lv[11], rv[21];
#pragma unroll
                    for (int i = 0; i < 11; i++) lv[i] = pl[i];
#pragma unroll
                    for (int i = 0; i < 21; i++) rv[i] = pr[i];
#pragma unroll
                    for (int inc = 0; inc < 11; inc++) {
                        int s = 0;
#pragma unroll
                        for (int i = 0; i < 11; i++) s += abs(lv[i] - rv[inc + i]);
                        part[inc] = s;
                    }
                }
#pragma unroll
                for (int inc = 0; inc < 11; inc++) {
#pragma unroll
                    for (int o = 8; o > 0; o >>= 1) part[inc] += __shfl_xor_sync(0xffffffffu, part[inc], o);
                }
                // lanes 0..15 now hold the 11 window SADs (cv::norm(IL, IR, NORM_L1): exact integers)
                int bestSad = 0x7fffffff, bestinc = 0;
#pragma unroll
                for (int inc = 0; inc < 11; inc++) {
                    if (part[inc] < bestSad) {
                        bestSad = part[inc];
                        bestinc = inc - Lr;
                    }
                }
                if (!(bestinc == -Lr || bestinc == Lr)) {
                    float dist1 = 0, dist2 = 0, dist3 = 0;
#pragma unroll
                    for (int inc = 1; inc < 10; inc++) {
                        if (inc - Lr == bestinc) {
                            dist1 = (float)part[inc - 1];
                            dist2 = (float)part[inc];
                            dist3 = (float)part[inc + 1];
                        }
                    }
                    const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));    // :960
                    if (!(deltaR < -1 || deltaR > 1)) {
                        float bestuR = tb.scale[levelL] * (scaleduR0 + (float)bestinc + deltaR);    // :966
                        float disparity = uL - bestuR;
                        if (disparity >= 0 && disparity < maxD) {
                            if (disparity <= 0) {
                                disparity = 0.01;
                                bestuR = (float)((double)uL - 0.01);
                            }
                            outZ = mbf / disparity;
                            outU = bestuR;
                            outSad = bestSad;
                        }
                    }
                }
            }
        }
        if (lane == 0) {
            uOut[iL] = outU;
            zOut[iL] = outZ;
            sOutp[iL] = outSad;
        }
    }
}

// :983-996 — median of the accepted SADs (element size/2 of the sorted list), cut at 1.5 * 1.4 * median.
// grid npairs; SAD <= 121 * 255 = 30855 < 2^15: two radix levels of 8 + 7 bits.
__global__ void __launch_bounds__(ST_THREADS) stereo_median_kernel(const ImageResult* __restrict__ result, int cap,
                                                                   float* __restrict__ uRight, float* __restrict__ depth,
                                                                   const int* __restrict__ sad) {
    __shared__ int hist[256];
    __shared__ int sSel[3];    // [0] total, [1] selected high bin, [2] rank inside the bin / selected low bits
    const int pair = blockIdx.x;
    const int nL = min(result[2 * pair].count, cap);
    const int* s = sad + (size_t)pair * cap;
    for (int i = nL + threadIdx.x; i < cap; i += ST_THREADS) {    // slots past the left keypoints: defined output
        uRight[(size_t)pair * cap + i] = -1.0f;
        depth[(size_t)pair * cap + i] = -1.0f;
    }
    hist[threadIdx.x] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < nL; i += ST_THREADS) {
        const int d = s[i];
        if (d >= 0) atomicAdd(&hist[d >> 7], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int total = 0;
        for (int b = 0; b < 256; b++) total += hist[b];
        int k = total / 2, b = 0;
        if (total > 0) {
            while (k >= hist[b]) k -= hist[b++];
        }
        sSel[0] = total;
        sSel[1] = b;
        sSel[2] = k;
    }
    __syncthreads();
    const int total = sSel[0], hiBin = sSel[1], k = sSel[2];
    if (total == 0) return;    // the reference indexes an empty vector here; nothing to cut
    __syncthreads();
    hist[threadIdx.x] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < nL; i += ST_THREADS) {
        const int d = s[i];
        if (d >= 0 && (d >> 7) == hiBin) atomicAdd(&hist[d & 127], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int kk = k, b = 0;
        while (kk >= hist[b]) kk -= hist[b++];
        sSel[2] = b;
    }
    __syncthreads();
    const float median = (float)((hiBin << 7) | sSel[2]);
    const float thDist = 1.5f * 1.4f * median;
    float* u = uRight + (size_t)pair * cap;
    float* z = depth + (size_t)pair * cap;
    for (int i = threadIdx.x; i < nL; i += ST_THREADS) {
        const int d = s[i];
        if (d >= 0 && !((float)d < thDist)) {
            u[i] = -1;
            z[i] = -1;
        }
    }
}

size_t stereo_smem_bytes(int cap) { return (size_t)cap * 12; }

// uRight / depth: npairs x cap floats; sad: npairs x cap ints of scratch.  Returns the kernels launched (0 on a
// configuration error, reported by the caller through stereo_smem_bytes).
int launch_stereo(const Geom& g, const float* scale, const float* invScale, const uint8_t* pyr, const orbfe_keypoint* kps,
                  const uint8_t* desc, const ImageResult* result, int npairs, int cap, float mb, float mbf, float* uRight,
                  float* depth, int* sad, cudaStream_t st) {
    if (npairs <= 0 || cap <= 0) return 0;
    StereoTables tb;
    for (int l = 0; l < ORBFE_MAX_LEVELS; l++) {
        tb.scale[l] = l < g.nlevels ? scale[l] : 1.0f;
        tb.invScale[l] = l < g.nlevels ? invScale[l] : 1.0f;
    }
    const size_t smem = stereo_smem_bytes(cap);
    static SmemOptIn optIn;
    smem_opt_in(optIn, stereo_match_kernel, smem);
    const dim3 grid((cap + ST_LEFT_PER_CTA - 1) / ST_LEFT_PER_CTA, npairs);
    stereo_match_kernel<<<grid, ST_THREADS, smem, st>>>(g, tb, pyr, kps, desc, result, cap, mb, mbf, uRight, depth, sad);
    stereo_median_kernel<<<npairs, ST_THREADS, 0, st>>>(result, cap, uRight, depth, sad);
    return 2;
}

}    // namespace orbfe
