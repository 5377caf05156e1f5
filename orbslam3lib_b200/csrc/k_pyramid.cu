// k_pyramid.cu — ORBextractor::ComputePyramid (ORBextractor_old.cc:1331-1356, upstream body) on the device.
//
// Level 0 = copyMakeBorder(image, 19, BORDER_REFLECT_101); level l = resize(level l-1, INTER_LINEAR) followed by
// copyMakeBorder(19, BORDER_REFLECT_101 | BORDER_ISOLATED).  Both steps are one pass here: every byte of the padded
// destination plane (frame included) is produced directly from the source ROI by reflecting the destination
// coordinate first.  The resize arithmetic is OpenCV's 8-bit fixed-point INTER_LINEAR; the
// per-column / per-row taps are precomputed on the host exactly as cv::resize does (orbfe_api.cu: build_taps).
//
// HBM-bound ((P - P_last) + (P - P_0) bytes per image).  Each thread produces 4 adjacent bytes and
// issues one 32-bit store; source reads go through L1 (a 4-byte output group touches <= 6 source bytes per row).
#include "orbfe_internal.h"

namespace orbfe {

__device__ __forceinline__ int reflect101(int p, int len) {    // cv::borderInterpolate(p, len, BORDER_REFLECT_101)
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p; else p = 2 * len - p - 2;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

// grid: (ceil(words / 64), ceil(rows / 4), n); block (64, 4)
__global__ void __launch_bounds__(256) pad_level0_kernel(const uint8_t* __restrict__ images, int stride,
                                                         size_t imageStride, uint8_t* __restrict__ pyr,
                                                         unsigned long long slabBytes, int w, int h, int pitch) {
    const int wx = blockIdx.x * 64 + threadIdx.x;
    const int py = blockIdx.y * 4 + threadIdx.y;
    if (py >= h + 2 * kEdge) return;
    const int c0 = 4 * wx;
    if (c0 + 3 < kXOff - kEdge || c0 >= kXOff + w + kEdge) return;
    const uint8_t* src = images + (size_t)blockIdx.z * imageStride + (size_t)reflect101(py - kEdge, h) * stride;
    uint32_t out = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const int x = c0 + b - kXOff;
        uint32_t v = 0;
        if (x >= -kEdge && x < w + kEdge) v = __ldg(src + reflect101(x, w));
        out |= v << (8 * b);
    }
    *reinterpret_cast<uint32_t*>(pyr + (size_t)blockIdx.z * slabBytes + (size_t)py * pitch + c0) = out;
}

// grid: (ceil(words / 64), ceil(rows / 4), n); block (64, 4).  src/dst point at the planes of image 0.
__global__ void __launch_bounds__(256) resize_level_kernel(const uint8_t* __restrict__ srcPlane, int spitch,
                                                           uint8_t* __restrict__ dstPlane, int dpitch,
                                                           unsigned long long slabBytes, int dw, int dh,
                                                           const ResizeTap* __restrict__ xtab,
                                                           const ResizeTap* __restrict__ ytab) {
    const int wx = blockIdx.x * 64 + threadIdx.x;
    const int py = blockIdx.y * 4 + threadIdx.y;
    if (py >= dh + 2 * kEdge) return;
    const int c0 = 4 * wx;
    if (c0 + 3 < kXOff - kEdge || c0 >= kXOff + dw + kEdge) return;
    const ResizeTap yt = ytab[reflect101(py - kEdge, dh)];
    const uint8_t* s0 = srcPlane + (size_t)blockIdx.z * slabBytes + (size_t)(yt.s + kEdge) * spitch + kXOff;
    const uint8_t* s1 = srcPlane + (size_t)blockIdx.z * slabBytes + (size_t)(yt.s1 + kEdge) * spitch + kXOff;
    uint32_t out = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const int x = c0 + b - kXOff;
        uint32_t v = 0;
        if (x >= -kEdge && x < dw + kEdge) {
            const ResizeTap xt = xtab[reflect101(x, dw)];
            const int r0 = (int)__ldg(s0 + xt.s) * xt.a0 + (int)__ldg(s0 + xt.s1) * xt.a1;
            const int r1 = (int)__ldg(s1 + xt.s) * xt.a0 + (int)__ldg(s1 + xt.s1) * xt.a1;
            v = (uint32_t)((((yt.a0 * (r0 >> 4)) >> 16) + ((yt.a1 * (r1 >> 4)) >> 16) + 2) >> 2) & 0xffu;
        }
        out |= v << (8 * b);
    }
    *reinterpret_cast<uint32_t*>(dstPlane + (size_t)blockIdx.z * slabBytes + (size_t)py * dpitch + c0) = out;
}

int launch_pyramid(const Geom& g, const uint8_t* images, int n, int stride, size_t imageStride, uint8_t* pyr,
                   const ResizeTap* const* xtab, const ResizeTap* const* ytab, cudaStream_t st) {
    int launches = 0;
    const dim3 block(64, 4);
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& L = g.lv[l];
        const int words = (kXOff + L.w + kEdge + 3) / 4;
        const dim3 grid((words + 63) / 64, (L.h + 2 * kEdge + 3) / 4, n);
        if (l == 0) {
            pad_level0_kernel<<<grid, block, 0, st>>>(images, stride, imageStride, pyr + L.planeOff, g.slabBytes, L.w,
                                                      L.h, L.pitch);
        } else {
            const LevelGeom& P = g.lv[l - 1];
            resize_level_kernel<<<grid, block, 0, st>>>(pyr + P.planeOff, P.pitch, pyr + L.planeOff, L.pitch,
                                                        g.slabBytes, L.w, L.h, xtab[l], ytab[l]);
        }
        launches++;
    }
    return launches;
}

}    // namespace orbfe
