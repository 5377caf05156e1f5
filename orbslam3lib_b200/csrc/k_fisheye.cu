// k_fisheye.cu — the half of Frame::ComputeStereoFishEyeMatches that follows BFMatchORB (Frame.cc:1164-1204) with
// KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:300-363), unproject (:107-136), project(Vector3f) (:60-79) and
// Triangulate (:380-392), on the device: SURVEY.md §8(f) rank 2.
//
// One thread per stereo-area left keypoint of a frame (grid.y = frame).  Per keypoint: the distance gates (0 = the
// reference's "bugged point", < 70), both rays by Newton iteration, the parallax gate, the 4x4 DLT system, its null
// vector, the two depth signs and the two chi-square reprojection gates; then mvLeftToRightMatch, mvRightToLeftMatch
// (several left keypoints may share a right one: the reference loop leaves the LAST, i.e. the largest left index ->
// atomicMax), mvDepth, mvStereo3Dpoints and the match count.
//
// Arithmetic: float exactly where the reference is float (no FMA: the library is built with --fmad=false).  The
// reference solves the DLT with Eigen::JacobiSVD<Matrix4f>; here the null vector comes from a one-sided Jacobi SVD in
// double on the same float matrix — a few hundred fp64 operations per match, nothing next to the extraction — so the
// result is the correctly rounded one and the float SVD noise of the reference (~4e-4 relative on far points) is the
// parity tolerance, not ours (tests/test_gpu_fisheye.py states it).  A few KB per frame: launch-latency bound.
#include "kb8.cuh"

namespace orbfe {

struct FisheyeRig {    // == orbfe_fisheye_rig
    float cam1[8], cam2[8];
    float Rlr[9], tlr[3];
    float precision;
    int nlevels;
    float levelSigma2[ORBFE_MAX_LEVELS];
};
static_assert(sizeof(FisheyeRig) == 184, "orbfe_fisheye_rig layout");

// KannalaBrandt8::TriangulateMatches of the rig's left camera against its right one: z1 or the negative reason
__device__ __forceinline__ float kb_triangulate_matches(const FisheyeRig& g, float x1, float y1, float x2, float y2, float sigmaLevel,
                                                        float unc, float p3D[3]) {
    return kb_triangulate(g.cam1, g.cam2, g.Rlr, g.tlr, g.precision, x1, y1, x2, y2, sigmaLevel, unc, p3D);
}

// clear the outputs of every frame: grid (ceil(cap / 256), nframes)
__global__ void __launch_bounds__(256) fisheye_clear_kernel(int cap, int* __restrict__ ltr, int* __restrict__ rtl,
                                                            float* __restrict__ depth, float* __restrict__ p3d,
                                                            int* __restrict__ status, int* __restrict__ nMatches) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    const size_t o = (size_t)blockIdx.y * cap + i;
    if (i == 0) nMatches[blockIdx.y] = 0;
    if (i >= cap) return;
    ltr[o] = -1;
    rtl[o] = -1;
    depth[o] = -1.0f;
    p3d[3 * o] = p3d[3 * o + 1] = p3d[3 * o + 2] = 0.f;
    if (status) status[o] = 0;
}

// grid (ceil(cap / 128), nframes); block 128
__global__ void __launch_bounds__(128) fisheye_matches_kernel(
    const __grid_constant__ FisheyeRig rig, int cap, int knnCap, const orbfe_keypoint* __restrict__ keysLAll,
    const orbfe_keypoint* __restrict__ keysRAll, const int* __restrict__ nLAll, const int* __restrict__ nRAll,
    const int* __restrict__ monoLAll, const int* __restrict__ monoRAll, const int* __restrict__ knnIdx,
    const int* __restrict__ knnDist, int* __restrict__ ltr, int* __restrict__ rtl, float* __restrict__ depth,
    float* __restrict__ p3d, int* __restrict__ status, int* __restrict__ nMatches) {
    const int f = blockIdx.y;
    const int i = blockIdx.x * 128 + threadIdx.x;    // stereo-area row
    const int nL = min(nLAll[f], cap), nR = min(nRAll[f], cap);
    const int monoL = monoLAll ? monoLAll[f] : 0, monoR = monoRAll ? monoRAll[f] : 0;
    bool matched = false;
    if (monoL >= 0 && monoR >= 0 && i < nL - monoL && i < knnCap) {
        const size_t kr = ((size_t)f * knnCap + i) * 2;
        const int d1 = knnDist[kr];
        const int rightPos = knnIdx[kr];
        if (d1 > 0 && d1 < 70 && rightPos >= 0 && rightPos + monoR < nR) {
            const size_t lo = (size_t)f * cap + i + monoL, ro = (size_t)f * cap + rightPos + monoR;
            const orbfe_keypoint kl = keysLAll[lo], kq = keysRAll[ro];
            const int ol = min(max(kl.octave, 0), ORBFE_MAX_LEVELS - 1), oq = min(max(kq.octave, 0), ORBFE_MAX_LEVELS - 1);
            float p[3] = {0.f, 0.f, 0.f};
            const float dep = kb_triangulate_matches(rig, kl.x, kl.y, kq.x, kq.y, rig.levelSigma2[ol], rig.levelSigma2[oq], p);
            matched = dep > 0.0001f;
            if (status) status[lo] = matched ? 1 : (dep < 0 ? (int)dep : -6);
            if (matched) {
                ltr[lo] = rightPos + monoR;
                atomicMax(&rtl[ro], i + monoL);
                depth[lo] = dep;
                p3d[3 * lo] = p[0];
                p3d[3 * lo + 1] = p[1];
                p3d[3 * lo + 2] = p[2];
            }
        }
    }
    const int cnt = __syncthreads_count(matched);
    if (threadIdx.x == 0 && cnt) atomicAdd(&nMatches[f], cnt);
}

// All pointers device memory except rig.  Returns the kernels launched.
int launch_fisheye_matches(const void* rig, int cap, int knnCap, int nframes, const orbfe_keypoint* keysL,
                           const orbfe_keypoint* keysR, const int* nL, const int* nR, const int* monoL, const int* monoR,
                           const int* knnIdx, const int* knnDist, int* ltr, int* rtl, float* depth, float* p3d, int* status,
                           int* nMatches, cudaStream_t st) {
    if (nframes <= 0 || cap <= 0) return 0;
    FisheyeRig g;
    memcpy(&g, rig, sizeof(g));
    fisheye_clear_kernel<<<dim3((cap + 255) / 256, nframes), 256, 0, st>>>(cap, ltr, rtl, depth, p3d, status, nMatches);
    fisheye_matches_kernel<<<dim3((cap + 127) / 128, nframes), 128, 0, st>>>(g, cap, knnCap, keysL, keysR, nL, nR, monoL,
                                                                            monoR, knnIdx, knnDist, ltr, rtl, depth, p3d,
                                                                            status, nMatches);
    return 2;
}

}    // namespace orbfe
