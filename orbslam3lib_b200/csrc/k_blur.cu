// k_blur.cu — cv::GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) of ORBextractor_old.cc:1146-1147.
//
// OpenCV's 8-bit path is fixed point and bit exact: Q8 kernel k = [18,34,48,56,48,34,18]
// (sum 256), horizontal pass H = sum k_i * src (<= 255*256, fits u16), vertical pass V = sum k_j * H,
// dst = (V + 32768) >> 16.  The reference blurs a clone() of the level, i.e. the image WITHOUT its 19-px frame,
// reflecting about the true image edge; the padded pyramid plane holds exactly that reflection already (the frame
// is the REFLECT_101 extension), so the kernel reads its 3-px halo from the plane and never branches on borders.
//
// HBM-bound: read P + write P bytes per image.  One CTA = one 128 x 32 output tile of one level of
// one image: 16-byte coalesced loads of the (128+32) x (32+6) source window into shared memory, horizontal pass to a
// u16 tile in shared memory, vertical pass, one 32-bit store per 4 output pixels.
#include "orbfe_internal.h"

namespace orbfe {

constexpr int BT_W = 128, BT_H = 32;
constexpr int BT_SW = BT_W + 32;    // source window width (16-byte aligned halo on both sides)
constexpr int BT_SH = BT_H + 6;

struct BlurTile {
    short level, tx, ty, pad;
};

__global__ void __launch_bounds__(256) gauss7x7_kernel(const __grid_constant__ Geom g, const uint8_t* __restrict__ pyr,
                                                       uint8_t* __restrict__ blur, const BlurTile* __restrict__ tiles,
                                                       const int* __restrict__ levelCount) {
    __shared__ __align__(16) uint8_t sSrc[BT_SH * BT_SW];
    __shared__ __align__(16) uint16_t sH[BT_SH * BT_W];
    const BlurTile t = tiles[blockIdx.x];
    const int img = blockIdx.y;
    // ORBextractor_old.cc:1140-1141: levels without keypoints are not blurred
    if (levelCount && levelCount[img * g.nlevels + t.level] == 0) return;
    const LevelGeom& L = g.lv[t.level];
    const int x0 = t.tx * BT_W, y0 = t.ty * BT_H;
    const uint8_t* plane = pyr + (size_t)img * g.slabBytes + L.planeOff;
    const int planeRows = L.h + 2 * kEdge;

    // load rows y0-3 .. y0+BT_H+2, plane columns (kXOff + x0 - 16) .. +BT_SW
    const int pc0 = kXOff + x0 - 16;
    for (int i = threadIdx.x; i < BT_SH * (BT_SW / 16); i += 256) {
        const int r = i / (BT_SW / 16), c = i - r * (BT_SW / 16);
        const int pr = y0 - 3 + r + kEdge, pc = pc0 + 16 * c;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (pr < planeRows && pc + 16 <= L.pitch) v = __ldg(reinterpret_cast<const uint4*>(plane + (size_t)pr * L.pitch + pc));
        *reinterpret_cast<uint4*>(&sSrc[r * BT_SW + 16 * c]) = v;
    }
    __syncthreads();

    // horizontal pass: 4 adjacent outputs per thread
    for (int i = threadIdx.x; i < BT_SH * (BT_W / 4); i += 256) {
        const int r = i / (BT_W / 4), xg = (i - r * (BT_W / 4)) * 4;
        const uint8_t* s = &sSrc[r * BT_SW + 16 + xg - 3];    // s[0] = src(x - 3)
        uint32_t p[10];
#pragma unroll
        for (int k = 0; k < 10; k++) p[k] = s[k];
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            o[k] = 18u * (p[k] + p[k + 6]) + 34u * (p[k + 1] + p[k + 5]) + 48u * (p[k + 2] + p[k + 4]) + 56u * p[k + 3];
        *reinterpret_cast<uint2*>(&sH[r * BT_W + xg]) = make_uint2(o[0] | (o[1] << 16), o[2] | (o[3] << 16));
    }
    __syncthreads();

    // vertical pass
    uint8_t* out = blur + (size_t)img * g.slabBytes + L.planeOff;
    for (int i = threadIdx.x; i < BT_H * (BT_W / 4); i += 256) {
        const int r = i / (BT_W / 4), xg = (i - r * (BT_W / 4)) * 4;
        const int x = x0 + xg, y = y0 + r;
        if (x >= L.w || y >= L.h) continue;
        uint32_t acc[4] = {0, 0, 0, 0};
        const uint32_t kk[7] = {18, 34, 48, 56, 48, 34, 18};
#pragma unroll
        for (int j = 0; j < 7; j++) {
            const uint2 hv = *reinterpret_cast<const uint2*>(&sH[(r + j) * BT_W + xg]);
            acc[0] += kk[j] * (hv.x & 0xffffu);
            acc[1] += kk[j] * (hv.x >> 16);
            acc[2] += kk[j] * (hv.y & 0xffffu);
            acc[3] += kk[j] * (hv.y >> 16);
        }
        const uint32_t o = ((acc[0] + 32768u) >> 16) | (((acc[1] + 32768u) >> 16) << 8) |
                           (((acc[2] + 32768u) >> 16) << 16) | (((acc[3] + 32768u) >> 16) << 24);
        *reinterpret_cast<uint32_t*>(out + (size_t)(y + kEdge) * L.pitch + kXOff + x) = o;
    }
}

// host: tile table of one geometry (cached by the caller)
int blur_tile_table(const Geom& g, BlurTile* out, int cap) {
    int n = 0;
    for (int l = 0; l < g.nlevels; l++) {
        const int tx = (g.lv[l].w + BT_W - 1) / BT_W, ty = (g.lv[l].h + BT_H - 1) / BT_H;
        for (int y = 0; y < ty; y++)
            for (int x = 0; x < tx; x++) {
                if (out && n < cap) out[n] = BlurTile{(short)l, (short)x, (short)y, 0};
                n++;
            }
    }
    return n;
}

int launch_blur_tiles(const Geom& g, const uint8_t* pyr, uint8_t* blur, int n, const void* tiles, int nTiles,
                      const int* levelCount, cudaStream_t st) {
    gauss7x7_kernel<<<dim3(nTiles, n), 256, 0, st>>>(g, pyr, blur, (const BlurTile*)tiles, levelCount);
    return 1;
}

}    // namespace orbfe
