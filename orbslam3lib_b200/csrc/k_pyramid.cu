// k_pyramid.cu — ORBextractor::ComputePyramid (ORBextractor_old.cc:1331-1356, upstream body) on the device.
//
// Level 0 = copyMakeBorder(image, 19, BORDER_REFLECT_101); level l = resize(level l-1, INTER_LINEAR) followed by
// copyMakeBorder(19, BORDER_REFLECT_101 | BORDER_ISOLATED).  Both steps are one pass here: every byte of the padded
// destination plane (frame included) is produced directly from the source ROI by reflecting the destination
// coordinate first.  The resize arithmetic is OpenCV's 8-bit fixed-point INTER_LINEAR; the
// per-column / per-row taps are precomputed on the host exactly as cv::resize does (orbfe_api.cu: build_taps).
//
// HBM-bound by bytes ((P - P_last) + (P - P_0) bytes per image); the work is organised to spend as few
// instructions per pixel as possible:
//   * pad_level0_kernel: 16 bytes per thread, aligned 128-bit loads/stores in the interior, byte path on the frame;
//   * resize_tile_kernel: one CTA = a 256 x (4*RPT) tile of the padded destination plane.  The source window is staged
//     in shared memory with 128-bit loads, twice: as is and shifted by one byte, so that the horizontal pair
//     (S[s], S[s+1]) of any tap is ONE aligned 16-bit shared load.  A thread owns 4 adjacent destination columns
//     (taps in registers) and walks down its rows: horizontal pass = one IDP.2A per pixel and source row
//     (a0*S[s] + a1*S[s+1]); vertical pass in registers; one 32-bit store per 4 pixels.
#include <algorithm>
#include <cstdlib>

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

// range [lo, hi] of reflect101(x, len) for x in [xa, xb] (xa <= xb, |overhang| < len)
__device__ __forceinline__ void reflect_range(int xa, int xb, int len, int& lo, int& hi) {
    lo = len;
    hi = -1;
    const int ia = max(xa, 0), ib = min(xb, len - 1);
    if (ia <= ib) { lo = ia; hi = ib; }
    if (xa < 0) {
        const int a = reflect101(min(xb, -1), len), b = reflect101(xa, len);
        lo = min(lo, min(a, b));
        hi = max(hi, max(a, b));
    }
    if (xb >= len) {
        const int a = reflect101(xb, len), b = reflect101(max(xa, len), len);
        lo = min(lo, min(a, b));
        hi = max(hi, max(a, b));
    }
}

// grid: (ceil(chunks / 64), ceil(rows / 4), n); block (64, 4); one 16-byte chunk of the padded plane per thread.
// The chunk index is rotated by 3 so that the chunks touching the right frame and those touching the left frame sit
// in the same warp: only that warp takes the byte path, all the others copy 128 bits per thread.
__global__ void __launch_bounds__(256) pad_level0_kernel(const uint8_t* __restrict__ images, int stride,
                                                         size_t imageStride, uint8_t* __restrict__ pyr,
                                                         unsigned long long slabBytes, int w, int h, int pitch,
                                                         int nchunks) {
    const int cr = blockIdx.x * 64 + threadIdx.x;
    const int py = blockIdx.y * 4 + threadIdx.y;
    if (py >= h + 2 * kEdge || cr >= nchunks) return;
    int cx = cr + nchunks - 3;
    if (cx >= nchunks) cx -= nchunks;
    if (cx < 0) cx = cr;    // fewer than 3 chunks
    const int c0 = 16 * cx;
    const uint8_t* src = images + (size_t)blockIdx.z * imageStride + (size_t)reflect101(py - kEdge, h) * stride;
    uint4* dst = reinterpret_cast<uint4*>(pyr + (size_t)blockIdx.z * slabBytes + (size_t)py * pitch + c0);
    const int x0 = c0 - kXOff;
    if (x0 >= 0 && x0 + 16 <= w && ((reinterpret_cast<uintptr_t>(src + x0) & 15) == 0)) {
        *dst = __ldg(reinterpret_cast<const uint4*>(src + x0));
        return;
    }
    uint32_t o[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        uint32_t out = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int x = x0 + 4 * q + b;
            uint32_t v = 0;
            if (x >= -kEdge && x < w + kEdge) {
                int xr = x < 0 ? -x : (x >= w ? 2 * w - 2 - x : x);
                if ((unsigned)xr >= (unsigned)w) xr = reflect101(x, w);    // images narrower than the frame
                v = __ldg(src + xr);
            }
            out |= v << (8 * b);
        }
        o[q] = out;
    }
    *dst = make_uint4(o[0], o[1], o[2], o[3]);
}

// Level 0 for 16-byte aligned sources (the normal case), split so that no warp mixes the two paths:
//   pad_interior_kernel  one aligned 128-bit copy per thread for the chunks that lie fully inside the image columns
//                        (rows of the top / bottom frame are reflected rows: a scalar per thread row);
//   pad_frame_kernel     the few chunks per row that touch the left / right frame, byte path, one thread per chunk.
// grid interior: (ceil(nInterior / 64), ceil(rows / 4), n), block (64, 4); frame: (ceil(rows * nFrame / 256), n).
__global__ void __launch_bounds__(256) pad_interior_kernel(const uint8_t* __restrict__ images, int stride,
                                                           size_t imageStride, uint8_t* __restrict__ pyr,
                                                           unsigned long long slabBytes, int h, int pitch, int cFirst,
                                                           int nInterior) {
    const int ci = blockIdx.x * 64 + threadIdx.x;
    const int py = blockIdx.y * 4 + threadIdx.y;
    if (py >= h + 2 * kEdge || ci >= nInterior) return;
    const int c0 = 16 * (cFirst + ci);
    const uint8_t* src = images + (size_t)blockIdx.z * imageStride + (size_t)reflect101(py - kEdge, h) * stride + (c0 - kXOff);
    *reinterpret_cast<uint4*>(pyr + (size_t)blockIdx.z * slabBytes + (size_t)py * pitch + c0) =
        __ldg(reinterpret_cast<const uint4*>(src));
}

__global__ void __launch_bounds__(256) pad_frame_kernel(const uint8_t* __restrict__ images, int stride, size_t imageStride,
                                                        uint8_t* __restrict__ pyr, unsigned long long slabBytes, int w,
                                                        int h, int pitch, int cFirst, int nInterior, int nchunks) {
    const int nFrame = nchunks - nInterior;
    const int t = blockIdx.x * 256 + threadIdx.x;
    const int py = t / nFrame;
    if (py >= h + 2 * kEdge) return;
    int cx = t - py * nFrame;                    // frame chunks: [0, cFirst) then [cFirst + nInterior, nchunks)
    if (cx >= cFirst) cx += nInterior;
    const int c0 = 16 * cx, x0 = c0 - kXOff;
    const uint8_t* src = images + (size_t)blockIdx.y * imageStride + (size_t)reflect101(py - kEdge, h) * stride;
    uint32_t o[4];
#pragma unroll
    for (int q = 0; q < 4; q++) {
        uint32_t out = 0;
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const int x = x0 + 4 * q + b;
            uint32_t v = 0;
            if (x >= -kEdge && x < w + kEdge) v = __ldg(src + reflect101(x, w));
            out |= v << (8 * b);
        }
        o[q] = out;
    }
    *reinterpret_cast<uint4*>(pyr + (size_t)blockIdx.y * slabBytes + (size_t)py * pitch + c0) = make_uint4(o[0], o[1], o[2], o[3]);
}

constexpr int RS_TW = 256;    // destination columns per tile
constexpr int RS_C0 = 12;     // first plane column covered by tile 0 (kXOff - kEdge = 13, rounded down to a word)

// grid: (x tiles, y tiles, n); block 256 = 64 column groups x 4 row groups; dynamic shared memory 2 * SP * SH bytes.
// srcPlane / dstPlane point at the planes of image 0.
__global__ void __launch_bounds__(256) resize_tile_kernel(const uint8_t* __restrict__ srcPlane, int spitch,
                                                          uint8_t* __restrict__ dstPlane, int dpitch,
                                                          unsigned long long slabBytes, int dw, int dh,
                                                          const ResizeTap* __restrict__ xtab,
                                                          const ResizeTap* __restrict__ ytab, int SP, int SH, int rpt) {
    extern __shared__ __align__(16) uint8_t sm[];
    const int tid = threadIdx.x, tx = tid & 63, ty = tid >> 6;
    const int TH = 4 * rpt;
    const int cb = RS_C0 + blockIdx.x * RS_TW;    // plane column of the tile's first byte
    const int rb = blockIdx.y * TH;               // padded row of the tile's first row
    // destination ROI coordinates covered by the tile, reflected into the image: source window
    const int xa = max(cb - kXOff, -kEdge), xb = min(cb + RS_TW - 1 - kXOff, dw + kEdge - 1);
    const int ya = rb - kEdge, yb = min(rb + TH - 1, dh + 2 * kEdge - 1) - kEdge;
    if (xa > xb || ya > yb) return;
    int xlo, xhi, ylo, yhi;
    reflect_range(xa, xb, dw, xlo, xhi);
    reflect_range(ya, yb, dh, ylo, yhi);
    const int clo = xtab[xlo].s, chi = xtab[xhi].s + 1;    // source columns [clo, chi] (chi may be the first frame column)
    const int rlo = ytab[ylo].s, rhi = ytab[yhi].s1;       // source rows [rlo, rhi]
    const int colbase = (kXOff + clo) & ~15;
    const int nvec = ((kXOff + chi + 1) - colbase + 15) >> 4;
    const int nrows = rhi - rlo + 1;
    uint8_t* sA = sm;
    uint8_t* sB = sm + SP * SH;    // sB[j] = sA[j + 1]
    {
        const uint8_t* sp = srcPlane + (size_t)blockIdx.z * slabBytes + (size_t)(rlo + kEdge) * spitch + colbase;
        for (int i = tid; i < nvec * nrows; i += 256) {
            const int r = i / nvec, c = i - r * nvec;
            const uint8_t* p = sp + (size_t)r * spitch + 16 * c;
            const uint4 v = *reinterpret_cast<const uint4*>(p);
            const uint32_t nb = (colbase + 16 * c + 16 < spitch) ? p[16] : 0u;
            *reinterpret_cast<uint4*>(sA + r * SP + 16 * c) = v;
            *reinterpret_cast<uint4*>(sB + r * SP + 16 * c) =
                make_uint4(__funnelshift_r(v.x, v.y, 8), __funnelshift_r(v.y, v.z, 8), __funnelshift_r(v.z, v.w, 8),
                           __funnelshift_r(v.w, nb, 8));
        }
    }
    __syncthreads();
    const int c = cb + 4 * tx;
    if (c >= kXOff + dw + kEdge) return;
    uint32_t wt[4];
    int ad[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int x = c + k - kXOff;
        const bool valid = x >= -kEdge && x < dw + kEdge;
        const int2 raw = __ldg(reinterpret_cast<const int2*>(xtab + reflect101(valid ? x : 0, dw)));
        const int s = (short)(raw.x & 0xffff);
        const int off = kXOff + s - colbase;
        ad[k] = !valid ? 0 : ((off & 1) ? (SP * SH + off - 1) : off);
        wt[k] = valid ? (uint32_t)raw.y : 0u;    // a0 | a1 << 16
    }
    const int py0 = rb + ty * rpt;
    uint8_t* dp = dstPlane + (size_t)blockIdx.z * slabBytes + c;
    const int pyEnd = min(py0 + rpt, dh + 2 * kEdge);
    for (int py = py0; py < pyEnd; py++) {
        const int2 yraw = __ldg(reinterpret_cast<const int2*>(ytab + reflect101(py - kEdge, dh)));
        const int s0 = (short)(yraw.x & 0xffff) - rlo, s1 = (short)(yraw.x >> 16) - rlo;
        const uint32_t b0 = (uint32_t)yraw.y & 0xffffu, b1 = (uint32_t)yraw.y >> 16;
        // both horizontal passes every row: re-using the previous row's pass (1.2 instead of 2 per output row) costs
        // more register moves and branches than the 3 instructions per pixel it saves
        const uint8_t* row0 = sm + s0 * SP;
        const uint8_t* row1 = sm + s1 * SP;
        uint32_t A[4], B[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            A[k] = __dp2a_lo(wt[k], (uint32_t)*reinterpret_cast<const uint16_t*>(row0 + ad[k]), 0u) >> 4;
            B[k] = __dp2a_lo(wt[k], (uint32_t)*reinterpret_cast<const uint16_t*>(row1 + ad[k]), 0u) >> 4;
        }
        uint32_t out = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t v = (((b0 * A[k]) >> 16) + ((b1 * B[k]) >> 16) + 2u) >> 2;
            out |= v << (8 * k);    // v <= 255: the weights of each pass sum to 2048
        }
        *reinterpret_cast<uint32_t*>(dp + (size_t)py * dpitch) = out;
    }
}

// ---- resize_rows_kernel: the default resize -----------------------------------------------------------------------------
// Same tile as resize_tile_kernel (256 threads = 64 column groups x 4 row groups, a 256-column x 4 rpt-row tile of the
// padded destination plane, source window staged as is + shifted by one byte), with the row loop stripped down: all
// lanes of a warp are on the same destination row, so everything that depends on the row only -- the byte offsets of
// its two source rows in the window and the two vertical weights -- is ONE 16-byte record in shared memory, built once
// per CTA (no reflect / tap decode / row-offset multiplies in the loop), and the final (t + 2) >> 2 runs on two packed
// 16-bit pairs.  Everything that depends on the geometry only comes from host tables in padded plane coordinates
// (ResizeDims: the tile's staged window, the reflected taps per plane column / row, a reciprocal for the staging index).  Per 4 pixels and row: 1 LDS.128, 8 IADD3 + 8 LDS.U16 + 8 IDP.2A + 8 shifts (horizontal, both source
// rows), 8 IMAD + ~8 ALU (vertical), 5 ALU (rounding, PRMT), one 32-bit store: 64 SASS instructions against 92.
// Bit-exact with cv::resize(INTER_LINEAR) on CV_8UC1: same fixed-point steps in the same order.
__global__ void __launch_bounds__(256) resize_rows_kernel(const uint8_t* __restrict__ srcPlane, int spitch,
                                                          uint8_t* __restrict__ dstPlane, int dpitch,
                                                          unsigned long long slabBytes, int dw, int dh,
                                                          const ResizeTap* __restrict__ xpad,
                                                          const ResizeTap* __restrict__ ypad,
                                                          const ResizeTap* __restrict__ tileX,
                                                          const ResizeTap* __restrict__ tileY, int SP, int SH, int rpt) {
    extern __shared__ __align__(16) uint8_t sm[];
    const int tid = threadIdx.x, tx = tid & 63, ty = tid >> 6;
    const int TH = 4 * rpt;
    int4* sRow = reinterpret_cast<int4*>(sm + 2 * SP * SH);    // TH records
    const int cb = RS_C0 + blockIdx.x * RS_TW;    // plane column of the tile's first byte
    const int rb = blockIdx.y * TH;               // padded row of the tile's first row
    // the tile's staged source window, from the host tables (resize_pad_tables): no reflection or range search here
    const int2 tX = __ldg(reinterpret_cast<const int2*>(tileX + blockIdx.x));
    const int tY = __ldg(reinterpret_cast<const int*>(tileY + blockIdx.y));
    const int colbase = tX.x & 0xffff, nvec = tX.x >> 16;
    const unsigned vecMagic = (unsigned)tX.y;    // 2^31 / nvec + 1: i / nvec == umulhi(2 i, vecMagic) for i < 2^31 / nvec
    const int rlo = tY & 0xffff, nrows = tY >> 16;
    const int shift = SP * SH;    // sB[j] = sA[j + 1]
    for (int t = tid; t < TH; t += 256) {
        const int py = rb + t;
        int4 rec = make_int4(0, 0, 0, 0);
        if (py < dh + 2 * kEdge) {
            const int2 yraw = __ldg(reinterpret_cast<const int2*>(ypad + py));
            rec.x = ((yraw.x & 0xffff) - rlo) * SP;
            rec.y = ((int)((unsigned)yraw.x >> 16) - rlo) * SP;
            rec.z = yraw.y & 0xffff;
            rec.w = (int)((unsigned)yraw.y >> 16);
        }
        sRow[t] = rec;
    }
    {
        const uint8_t* sp = srcPlane + (size_t)blockIdx.z * slabBytes + (size_t)(rlo + kEdge) * spitch + colbase;
        const int lastVec = (spitch - colbase - 17) >> 4;    // vectors c <= lastVec have a byte behind them in the row
        for (int i = tid; i < nvec * nrows; i += 256) {
            const int r = (int)__umulhi(2u * (unsigned)i, vecMagic), c = i - r * nvec;
            const uint8_t* p = sp + r * spitch + 16 * c;
            const uint4 v = *reinterpret_cast<const uint4*>(p);
            const uint32_t nb = c <= lastVec ? p[16] : 0u;
            uint8_t* d = sm + r * SP + 16 * c;
            *reinterpret_cast<uint4*>(d) = v;
            *reinterpret_cast<uint4*>(d + shift) =
                make_uint4(__funnelshift_r(v.x, v.y, 8), __funnelshift_r(v.y, v.z, 8), __funnelshift_r(v.z, v.w, 8),
                           __funnelshift_r(v.w, nb, 8));
        }
    }
    __syncthreads();
    const int c = cb + 4 * tx;
    if (c >= kXOff + dw + kEdge) return;
    uint32_t wt[4];
    const uint8_t* ad[4];
    {
        const int4 q0 = __ldg(reinterpret_cast<const int4*>(xpad + c)), q1 = __ldg(reinterpret_cast<const int4*>(xpad + c + 2));
        const int offs[4] = {q0.x & 0xffff, q0.z & 0xffff, q1.x & 0xffff, q1.z & 0xffff};
        wt[0] = (uint32_t)q0.y; wt[1] = (uint32_t)q0.w; wt[2] = (uint32_t)q1.y; wt[3] = (uint32_t)q1.w;    // a0 | a1 << 16
#pragma unroll
        for (int k = 0; k < 4; k++) ad[k] = sm + ((offs[k] & 1) ? (shift + offs[k] - 1) : offs[k]);
    }
    const int r0 = ty * rpt, r1 = min(r0 + rpt, dh + 2 * kEdge - rb);
    uint8_t* dp = dstPlane + (size_t)blockIdx.z * slabBytes + (size_t)(rb + r0) * dpitch + c;
    // The horizontal pass of a source row is kept for the next destination row: when shrinking, the first source row of
    // a destination row is almost always the second one of the row above (1.2 horizontal passes per row instead of 2).
    // All lanes of a warp are on the same row, so the test is warp-uniform.
    int prevOff = -1;
    uint32_t Bp[4] = {0u, 0u, 0u, 0u};
    for (int r = r0; r < r1; r++, dp += dpitch) {
        const int4 rec = sRow[r];
        const uint32_t b0 = (uint32_t)rec.z, b1 = (uint32_t)rec.w;
        uint32_t A[4], t[4];
        if (rec.x == prevOff) {
#pragma unroll
            for (int k = 0; k < 4; k++) A[k] = Bp[k];
        } else {
#pragma unroll
            for (int k = 0; k < 4; k++)
                A[k] = __dp2a_lo(wt[k], (uint32_t)*reinterpret_cast<const uint16_t*>(ad[k] + rec.x), 0u) >> 4;
        }
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t B = __dp2a_lo(wt[k], (uint32_t)*reinterpret_cast<const uint16_t*>(ad[k] + rec.y), 0u) >> 4;
            t[k] = ((b0 * A[k]) >> 16) + ((b1 * B) >> 16);    // <= 1020: the weights of each pass sum to 2048
            Bp[k] = B;
        }
        prevOff = rec.y;
        // (t + 2) >> 2 on two packed pairs; bits that cross from the upper field land in bits 14-15, which the byte
        // selection below drops
        const uint32_t p01 = ((t[1] * 65536u + t[0]) + 0x00020002u) >> 2;
        const uint32_t p23 = ((t[3] * 65536u + t[2]) + 0x00020002u) >> 2;
        *reinterpret_cast<uint32_t*>(dp) = __byte_perm(p01, p23, 0x6420);
    }
}

// ---- halve_kernel: a level that is an exact 2x decimation of the one below -----------------------------------------------
// cv::resize(INTER_LINEAR) switches to INTER_AREA when both axes shrink by exactly 2 (imgproc/resize.cpp; scaleFactor 2.0 on
// even sizes): the 2x2 box average (s00 + s01 + s10 + s11 + 2) >> 2 of ResizeAreaFastVec.  The REFLECT_101 frame is
// written in the same pass by reflecting the destination coordinate first, like the bilinear kernels.  Not a hot path
// (no shipped configuration uses scale 2): one 32-bit word of the padded destination plane per thread, byte by byte.
// grid (ceil(words / 64), ceil(rows / 4), n); block (64, 4)
__global__ void __launch_bounds__(256) halve_kernel(const uint8_t* __restrict__ srcPlane, int spitch,
                                                    uint8_t* __restrict__ dstPlane, int dpitch, unsigned long long slabBytes,
                                                    int dw, int dh) {
    const int c0 = 4 * (blockIdx.x * 64 + threadIdx.x);    // plane column of the word
    const int py = blockIdx.y * 4 + threadIdx.y;           // padded row
    if (py >= dh + 2 * kEdge || c0 >= kXOff + dw + kEdge) return;
    const int y = reflect101(py - kEdge, dh);
    const uint8_t* r0 = srcPlane + (size_t)blockIdx.z * slabBytes + (size_t)(2 * y + kEdge) * spitch + kXOff;
    const uint8_t* r1 = r0 + spitch;
    uint32_t out = 0;
#pragma unroll
    for (int b = 0; b < 4; b++) {
        const int x = c0 + b - kXOff;
        if (x >= -kEdge && x < dw + kEdge) {
            const int xr = 2 * reflect101(x, dw);
            out |= (((uint32_t)r0[xr] + r0[xr + 1] + r1[xr] + r1[xr + 1] + 2u) >> 2) << (8 * b);
        }
    }
    *reinterpret_cast<uint32_t*>(dstPlane + (size_t)blockIdx.z * slabBytes + (size_t)py * dpitch + c0) = out;
}

// host: shared-memory window (pitch SP, rows SH) needed by the tiles of one level for `rpt` rows per thread.
// Mirrors the range computation of the kernel on the host copies of the tap tables.
static int h_reflect101(int p, int len) {
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p; else p = 2 * len - p - 2;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

void resize_tile_dims(const ResizeTap* xtab, const ResizeTap* ytab, int dw, int dh, int rpt, int* SP, int* SH) {
    int sp = 16, shh = 1;
    const int TH = 4 * rpt;
    for (int cb = RS_C0; cb < kXOff + dw + kEdge; cb += RS_TW) {
        const int xa = std::max(cb - kXOff, -kEdge), xb = std::min(cb + RS_TW - 1 - kXOff, dw + kEdge - 1);
        int lo = 1 << 30, hi = -1;
        for (int x = xa; x <= xb; x++) {
            const int r = h_reflect101(x, dw);
            lo = std::min(lo, r);
            hi = std::max(hi, r);
        }
        const int clo = xtab[lo].s, chi = xtab[hi].s + 1;
        const int colbase = (kXOff + clo) & ~15;
        sp = std::max(sp, (((kXOff + chi + 1) - colbase + 15) >> 4) * 16);
    }
    for (int rb = 0; rb < dh + 2 * kEdge; rb += TH) {
        const int ya = rb - kEdge, yb = std::min(rb + TH - 1, dh + 2 * kEdge - 1) - kEdge;
        int lo = 1 << 30, hi = -1;
        for (int y = ya; y <= yb; y++) {
            const int r = h_reflect101(y, dh);
            lo = std::min(lo, r);
            hi = std::max(hi, r);
        }
        shh = std::max(shh, ytab[hi].s1 - ytab[lo].s + 1);
    }
    *SP = sp;
    *SH = shh;
}

// host: the tables resize_rows_kernel reads (ResizeDims::xpad / ypad / tileX / tileY), for `rpt` rows per thread.
void resize_pad_counts(int dw, int dh, int rpt, int* nxpad, int* nypad, int* ntx, int* nty) {
    *ntx = (kXOff + dw + kEdge - RS_C0 + RS_TW - 1) / RS_TW;
    *nty = (dh + 2 * kEdge + 4 * rpt - 1) / (4 * rpt);
    *nxpad = RS_C0 + *ntx * RS_TW;    // indexed by plane column; a multiple of 4 (four taps are loaded as two 16-byte words)
    *nypad = dh + 2 * kEdge;
}

void resize_pad_tables(const ResizeTap* xtab, const ResizeTap* ytab, int dw, int dh, int rpt, ResizeTap* xpad, ResizeTap* ypad,
                       ResizeTap* tileX, ResizeTap* tileY) {
    int nxpad, nypad, ntx, nty;
    resize_pad_counts(dw, dh, rpt, &nxpad, &nypad, &ntx, &nty);
    const int TH = 4 * rpt;
    for (int i = 0; i < nxpad; i++) xpad[i] = ResizeTap{0, 0, 0, 0};    // columns outside the frame: weight 0, offset 0
    for (int bx = 0; bx < ntx; bx++) {
        const int cb = RS_C0 + bx * RS_TW;
        const int xa = std::max(cb - kXOff, -kEdge), xb = std::min(cb + RS_TW - 1 - kXOff, dw + kEdge - 1);
        int lo = 1 << 30, hi = -1;
        for (int x = xa; x <= xb; x++) {
            const int r = h_reflect101(x, dw);
            lo = std::min(lo, r);
            hi = std::max(hi, r);
        }
        const int clo = xtab[lo].s, chi = xtab[hi].s + 1;    // source columns [clo, chi] (chi may be the first frame column)
        const int colbase = (kXOff + clo) & ~15;
        const int nvec = ((kXOff + chi + 1) - colbase + 15) >> 4;
        const unsigned magic = 0x80000000u / (unsigned)nvec + 1u;    // nvec = 1 included (2^32 / nvec + 1 would not fit)
        tileX[bx] = ResizeTap{(short)colbase, (short)nvec, (short)(magic & 0xffffu), (short)(magic >> 16)};
        for (int x = xa; x <= xb; x++) {
            const ResizeTap& t = xtab[h_reflect101(x, dw)];
            xpad[x + kXOff] = ResizeTap{(short)(kXOff + t.s - colbase), 0, t.a0, t.a1};
        }
    }
    for (int by = 0; by < nty; by++) {
        const int rb = by * TH;
        const int ya = rb - kEdge, yb = std::min(rb + TH - 1, dh + 2 * kEdge - 1) - kEdge;
        int lo = 1 << 30, hi = -1;
        for (int y = ya; y <= yb; y++) {
            const int r = h_reflect101(y, dh);
            lo = std::min(lo, r);
            hi = std::max(hi, r);
        }
        const int rlo = ytab[lo].s, rhi = ytab[hi].s1;
        tileY[by] = ResizeTap{(short)rlo, (short)(rhi - rlo + 1), 0, 0};
    }
    for (int py = 0; py < nypad; py++) ypad[py] = ytab[h_reflect101(py - kEdge, dh)];
}

int launch_pyramid(const Geom& g, const uint8_t* images, int n, int stride, size_t imageStride, uint8_t* pyr,
                   const ResizeTap* const* xtab, const ResizeTap* const* ytab, const ResizeDims* dims,
                   const ResizeDims* rows16, cudaStream_t st) {
    int launches = 0;
    static SmemOptIn optIn;
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& L = g.lv[l];
        if (l == 0) {
            const int chunks = (kXOff + L.w + kEdge + 15) / 16;
            const int rows = L.h + 2 * kEdge;
            const int cFirst = kXOff / 16, nInterior = L.w / 16;    // chunks [cFirst, cFirst + nInterior) are pure image
            const bool aligned = stride % 16 == 0 && imageStride % 16 == 0 && (reinterpret_cast<uintptr_t>(images) & 15) == 0;
            if (aligned && nInterior > 0) {
                pad_interior_kernel<<<dim3((nInterior + 63) / 64, (rows + 3) / 4, n), dim3(64, 4), 0, st>>>(
                    images, stride, imageStride, pyr + L.planeOff, g.slabBytes, L.h, L.pitch, cFirst, nInterior);
                const int nFrame = chunks - nInterior;
                pad_frame_kernel<<<dim3((rows * nFrame + 255) / 256, n), 256, 0, st>>>(
                    images, stride, imageStride, pyr + L.planeOff, g.slabBytes, L.w, L.h, L.pitch, cFirst, nInterior, chunks);
                launches++;
            } else {
                const dim3 grid((chunks + 63) / 64, (rows + 3) / 4, n);
                pad_level0_kernel<<<grid, dim3(64, 4), 0, st>>>(images, stride, imageStride, pyr + L.planeOff, g.slabBytes,
                                                                L.w, L.h, L.pitch, chunks);
            }
        } else {
            const LevelGeom& P = g.lv[l - 1];
            if (P.w == 2 * L.w && P.h == 2 * L.h) {    // INTER_LINEAR is INTER_AREA here (see halve_kernel)
                const int words = (kXOff + L.w + kEdge + 3) / 4, rows = L.h + 2 * kEdge;
                halve_kernel<<<dim3((words + 63) / 64, (rows + 3) / 4, n), dim3(64, 4), 0, st>>>(pyr + P.planeOff, P.pitch, pyr + L.planeOff,
                                                                                           L.pitch, g.slabBytes, L.w, L.h);
                launches++;
                continue;
            }
            static const bool useTile = [] { const char* e = getenv("ORBFE_RESIZE_TILE"); return e && atoi(e) != 0; }();
            if (!useTile) {    // default: the stripped-down row loop on the same tiles
                const ResizeDims& d = dims[l];
                const int smem = 2 * d.SP * d.SH + 4 * d.rpt * 16;
                static SmemOptIn optRows;
                smem_opt_in(optRows, resize_rows_kernel, (size_t)smem);
                const dim3 grid((kXOff + L.w + kEdge - RS_C0 + RS_TW - 1) / RS_TW, (L.h + 2 * kEdge + 4 * d.rpt - 1) / (4 * d.rpt), n);
                resize_rows_kernel<<<grid, 256, smem, st>>>(pyr + P.planeOff, P.pitch, pyr + L.planeOff, L.pitch, g.slabBytes,
                                                            L.w, L.h, d.xpad, d.ypad, d.tileX, d.tileY, d.SP, d.SH, d.rpt);
                launches++;
                continue;
            }
            const ResizeDims& d = dims[l];
            const int smem = 2 * d.SP * d.SH;
            smem_opt_in(optIn, resize_tile_kernel, (size_t)smem);
            const dim3 grid((kXOff + L.w + kEdge - RS_C0 + RS_TW - 1) / RS_TW, (L.h + 2 * kEdge + 4 * d.rpt - 1) / (4 * d.rpt), n);
            resize_tile_kernel<<<grid, 256, smem, st>>>(pyr + P.planeOff, P.pitch, pyr + L.planeOff, L.pitch, g.slabBytes,
                                                        L.w, L.h, xtab[l], ytab[l], d.SP, d.SH, d.rpt);
        }
        launches++;
    }
    return launches;
}

}    // namespace orbfe
