// k_blur.cu — cv::GaussianBlur(level, 7x7, sigma 2, BORDER_REFLECT_101) of ORBextractor_old.cc:1146-1147.
//
// OpenCV's 8-bit path is fixed point and bit exact: Q8 kernel k = [18,34,48,56,48,34,18]
// (sum 256), horizontal pass H = sum k_i * src (<= 255*256, exact in 16 bits), vertical pass V = sum k_j * H,
// dst = (V + 32768) >> 16.  Nothing is rounded before the final shift, so the result is the exact integer 2-D
// convolution with k (x) k and the two passes may run in either order.  The reference blurs a clone() of the level,
// i.e. the image WITHOUT its 19-px frame, reflecting about the true image edge; the padded pyramid plane holds exactly
// that reflection already, so the kernel reads its 3-px halo from the plane and never branches on borders.
//
// HBM-bound by bytes (read P + write P per image); organised for few instructions per pixel:
//   * one WARP = one strip of <= 30 output words (120 px) x BL_ROWS rows; a lane owns one 32-bit word column (4 px)
//     and walks down the rows with coalesced 128-byte row reads straight from global memory (no shared memory);
//   * vertical pass first, SWAR: the word is split into even / odd bytes (two 16-bit lanes per register), a 7-row
//     window lives in registers (fully unrolled, statically indexed), 3 adds + 4 multiply-adds per register give the
//     Q8.8 column sums of 2 pixels at once (<= 65280: no carry between the lanes);
//   * horizontal pass: the neighbour lanes' sums come by warp shuffle, adjacent 16-bit pairs are formed with PRMT and
//     every output is four IDP.2A (16-bit x 8-bit dot products) accumulated on the rounding constant 32768;
//   * lanes 0 and 31 of a strip are halo lanes (strips overlap by two words); one 32-bit store per 4 output pixels.
#include "orbfe_internal.h"

namespace orbfe {

constexpr int BL_ROWS = 32;       // output rows per warp tile
constexpr int BL_WORDS = 30;      // output words per strip (lanes 1..30)
constexpr int BL_WARPS = 8;
constexpr int BL_PF = 8;          // source rows requested ahead of use

struct BlurTile {
    short level, tx, ty, pad;
};

__host__ __device__ inline void blur_strips(int w, int* nStrips, int* stripWords) {
    const int words = (w + 3) >> 2;
    const int ns = (words + BL_WORDS - 1) / BL_WORDS;
    *nStrips = ns;
    *stripWords = (words + ns - 1) / ns;
}

__device__ __forceinline__ void store_word_if(uint8_t* p, uint32_t v, int on) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.s32 q, %2, 0;\n\t@q st.global.u32 [%0], %1;\n\t}" :: "l"(p), "r"(v), "r"(on) : "memory");
}

// The row loop of one warp tile.  FULL: all BL_ROWS output rows exist (every tile but the last tile row of a level), so the
// loop carries no row test and no clamp.  Otherwise requests past the last source row the outputs read (row rowsLeft + 5 of
// the tile, still inside the padded plane: plane row h + kEdge + 2) are clamped onto it instead of predicated -- one min
// per row, no compare / branch -- and the clamped rows only feed output rows >= rowsLeft, which are not computed.
template <bool FULL>
__device__ __forceinline__ void blur_rows(const uint8_t* src, uint8_t* dst, ptrdiff_t pitch, int rowsLeft, int outLane) {
    const int lastRow = rowsLeft + 5;
    constexpr uint32_t K1 = 18u | (34u << 8) | (48u << 16) | (56u << 24);    // lo: (18,34)  hi: (48,56)
    constexpr uint32_t K2 = 48u | (34u << 8) | (0u << 16) | (18u << 24);     // lo: (48,34)  hi: (0,18)
    constexpr uint32_t K3 = 18u;                                             // lo: (18,0)

    uint32_t E[7], O[7];    // even / odd bytes of the last 7 source rows (16-bit lanes)
    uint32_t raw[BL_PF];    // loads in flight: row i + BL_PF is requested while row i is consumed
#pragma unroll
    for (int j = 0; j < BL_PF; j++) raw[j] = __ldg(reinterpret_cast<const uint32_t*>(src + (FULL ? j : min(j, lastRow)) * pitch));
#pragma unroll
    for (int i = 0; i < BL_ROWS + 6; i++) {
        // source row i of the tile = plane row y0 - 3 + kEdge + i
        const uint32_t wv = raw[i % BL_PF];
        if (i + BL_PF < BL_ROWS + 6)
            raw[i % BL_PF] = __ldg(reinterpret_cast<const uint32_t*>(src + (FULL ? i + BL_PF : min(i + BL_PF, lastRow)) * pitch));
        E[i % 7] = wv & 0x00ff00ffu;
        O[i % 7] = __byte_perm(wv, 0u, 0x4341);
        if (i >= 6) {
            const int r = i - 6;    // output row of the tile; window rows r .. r + 6 live in slots (r + k) % 7
            if (FULL || r < rowsLeft) {     // warp-uniform
                const uint32_t VE = (E[r % 7] + E[(r + 6) % 7]) * 18u + (E[(r + 1) % 7] + E[(r + 5) % 7]) * 34u +
                                    (E[(r + 2) % 7] + E[(r + 4) % 7]) * 48u + E[(r + 3) % 7] * 56u;    // (v0, v2)
                const uint32_t VO = (O[r % 7] + O[(r + 6) % 7]) * 18u + (O[(r + 1) % 7] + O[(r + 5) % 7]) * 34u +
                                    (O[(r + 2) % 7] + O[(r + 4) % 7]) * 48u + O[(r + 3) % 7] * 56u;    // (v1, v3)
                const uint32_t LE = __shfl_up_sync(0xffffffffu, VE, 1), LO = __shfl_up_sync(0xffffffffu, VO, 1);      // (v-4, v-2), (v-3, v-1)
                const uint32_t RE = __shfl_down_sync(0xffffffffu, VE, 1), RO = __shfl_down_sync(0xffffffffu, VO, 1);  // (v4, v6), (v5, v7)
                // adjacent pairs (lo lane first)
                const uint32_t Pm3 = __byte_perm(LO, LE, 0x7610);    // (v-3, v-2)
                const uint32_t Pm2 = __byte_perm(LE, LO, 0x7632);    // (v-2, v-1)
                const uint32_t Pm1 = __byte_perm(LO, VE, 0x5432);    // (v-1, v0)
                const uint32_t P0 = __byte_perm(VE, VO, 0x5410);     // (v0, v1)
                const uint32_t P1 = __byte_perm(VO, VE, 0x7610);     // (v1, v2)
                const uint32_t P2 = __byte_perm(VE, VO, 0x7632);     // (v2, v3)
                const uint32_t P3 = __byte_perm(VO, RE, 0x5432);     // (v3, v4)
                const uint32_t P4 = __byte_perm(RE, RO, 0x5410);     // (v4, v5)
                uint32_t o0 = __dp2a_lo(Pm3, K1, 32768u);            // v-3*18 + v-2*34
                o0 = __dp2a_hi(Pm1, K1, o0);                         // v-1*48 + v0*56
                o0 = __dp2a_lo(P1, K2, o0);                          // v1*48 + v2*34
                o0 = __dp2a_hi(P2, K2, o0);                          // v3*18
                uint32_t o1 = __dp2a_lo(Pm2, K1, 32768u);            // v-2*18 + v-1*34
                o1 = __dp2a_hi(P0, K1, o1);                          // v0*48 + v1*56
                o1 = __dp2a_lo(P2, K2, o1);                          // v2*48 + v3*34
                o1 = __dp2a_lo(P4, K3, o1);                          // v4*18
                uint32_t o2 = __dp2a_lo(Pm1, K1, 32768u);            // v-1*18 + v0*34
                o2 = __dp2a_hi(P1, K1, o2);                          // v1*48 + v2*56
                o2 = __dp2a_lo(P3, K2, o2);                          // v3*48 + v4*34
                o2 = __dp2a_hi(P4, K2, o2);                          // v5*18
                uint32_t o3 = __dp2a_lo(P0, K1, 32768u);             // v0*18 + v1*34
                o3 = __dp2a_hi(P2, K1, o3);                          // v2*48 + v3*56
                o3 = __dp2a_lo(P4, K2, o3);                          // v4*48 + v5*34
                o3 = __dp2a_hi(RE, K2, o3);                          // v6*18
                // byte 2 of every accumulator = (V + 32768) >> 16
                const uint32_t out = __byte_perm(__byte_perm(o0, o1, 0x0062), __byte_perm(o2, o3, 0x0062), 0x5410);
                store_word_if(dst + r * pitch, out, outLane);
            }
        }
    }
}

// grid (ceil(nTiles / BL_WARPS), n); block 32 * BL_WARPS; one warp per tile
__global__ void __launch_bounds__(32 * BL_WARPS) gauss7x7_kernel(const __grid_constant__ Geom g,
                                                                 const uint8_t* __restrict__ pyr,
                                                                 uint8_t* __restrict__ blur,
                                                                 const BlurTile* __restrict__ tiles, int nTiles,
                                                                 const int* __restrict__ levelCount) {
    const int lane = threadIdx.x & 31;
    const int ti = blockIdx.x * BL_WARPS + (threadIdx.x >> 5);
    if (ti >= nTiles) return;
    const BlurTile t = tiles[ti];
    const int img = blockIdx.y;
    // ORBextractor_old.cc:1140-1141: levels without keypoints are not blurred
    if (levelCount && levelCount[img * g.nlevels + t.level] == 0) return;
    const LevelGeom& L = g.lv[t.level];
    int nStrips, sw;
    blur_strips(L.w, &nStrips, &sw);
    const int words = (L.w + 3) >> 2;
    const int wfirst = t.tx * sw;                          // first output word of the strip
    const int wlast = min(wfirst + sw, words);             // one past the last output word
    const int wi = wfirst - 1 + lane;                      // this lane's word (level x = 4 * wi)
    const int y0 = t.ty * BL_ROWS;
    const int pc = kXOff + 4 * wi;                         // plane column of the word
    const bool inPlane = pc >= 0 && pc + 4 <= L.pitch;    // false only for lanes right of every word an output reads
    const bool outLane = lane >= 1 && wi < wlast;
    const size_t planeBase = (size_t)img * g.slabBytes + L.planeOff;
    const ptrdiff_t pitch = L.pitch;
    // source row y - 3 of output row y is plane row y - 3 + kEdge
    const uint8_t* src = pyr + planeBase + (y0 - 3 + kEdge) * pitch + (inPlane ? pc : 0);
    uint8_t* dst = blur + planeBase + (y0 + kEdge) * pitch + pc;
    asm volatile("" : "+l"(src), "+l"(dst));               // opaque: two base registers, every access = base + k * pitch
    const int rowsLeft = L.h - y0;                         // output rows available in this tile (> 0)
    if (rowsLeft >= BL_ROWS) blur_rows<true>(src, dst, pitch, rowsLeft, outLane);
    else blur_rows<false>(src, dst, pitch, rowsLeft, outLane);
}

// host: tile table of one geometry (cached by the caller)
int blur_tile_table(const Geom& g, BlurTile* out, int cap) {
    int n = 0;
    for (int l = 0; l < g.nlevels; l++) {
        int tx, sw;
        blur_strips(g.lv[l].w, &tx, &sw);
        const int ty = (g.lv[l].h + BL_ROWS - 1) / BL_ROWS;
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
    if (nTiles <= 0) return 0;
    gauss7x7_kernel<<<dim3((nTiles + BL_WARPS - 1) / BL_WARPS, n), 32 * BL_WARPS, 0, st>>>(
        g, pyr, blur, (const BlurTile*)tiles, nTiles, levelCount);
    return 1;
}

}    // namespace orbfe
