// k_fast.cu — the per-cell FAST stage of ORBextractor::ComputeKeyPointsOctTree (ORBextractor_old.cc:783-874).
//
// For every W=35 cell of every level the reference runs cv::FAST(cell + 6 px, iniThFAST, nonmax=true) and, only if
// that returns nothing, again with minThFAST (:828-848).  cv::FAST semantics: FAST-9/16 on the
// Bresenham circle of radius 3, tested on the cell sub-image minus a 3-px border, score = the largest threshold at
// which the pixel is still a corner, minus one; 3x3 non-max suppression with strict ">" where everything outside
// the tested region scores 0; output in raster order.  Cells overlap by 6 px and the same pixel may be reported by
// two cells — reproduced, not de-duplicated.
//
// Mapping: one CTA stages a run of up to 8 cells of one cell row (a strip of the level plane) into shared memory with
// 16-byte coalesced loads.  A pixel's corner score depends only on its own neighbourhood, not on the cell asking, so
// (1) the whole CTA scores the strip ONCE at iniThFAST: every lane tests 4 pixels packed in a 32-bit word with SWAR
//     byte compares (opposite-pair rejection on ring pixels 0/8 and 4/12), survivors are compacted per warp with
//     __ballot_sync, and 32 survivors at a time get the full test (16 ring pixels classified into a "brighter" and a
//     "darker" 16-bit mask, 9-run detected with shifts/ANDs) and the score (min/max over the sixteen 9-arcs with a
//     log-step sliding window in registers), written to a strip-wide score map in shared memory;
// (2) one warp per cell runs the 3x3 NMS on the cell's tested window of that map (neighbours outside the window
//     count as 0, exactly what cv::FAST sees on the sub-image) and emits in raster order with __ballot_sync /
//     __popc ranks, so the output order is deterministic and equals OpenCV's;
// (3) a cell that comes back empty is searched again at minThFAST on a private score tile (rare path).  Candidates go to the cell's fixed slot array as packed
// (x | y << 12 | response << 24), coordinates relative to (minBorderX, minBorderY) as in :865-870.
//
// HBM-bound by design (reads P bytes per image once; the 6-px overlaps hit L1/L2); integer only.
#include "orbfe_internal.h"

namespace orbfe {

constexpr int FAST_WARPS = 8;
constexpr int FAST_ITEM_COLS = 8;

__device__ __forceinline__ bool has_run9(uint32_t m16) {    // 9 contiguous set bits in a circular 16-bit mask
    uint32_t m = m16 | (m16 << 16);
    uint32_t x = m & (m >> 1);
    x &= x >> 2;
    x &= x >> 4;     // bit k: bits k..k+7 set
    x &= m >> 8;     // bit k: bits k..k+8 set
    return (x & 0xffffu) != 0;
}

// c points at the centre pixel inside the shared strip tile (pitch tp).
__device__ __forceinline__ bool fast9_is_corner(const uint8_t* c, int tp, int t) {
    const int v = c[0];
    const int lo = v - t, hi = v + t;
    // opposite-pair rejection (a 9-arc contains one pixel of every opposite pair)
    const int p0 = c[3 * tp], p8 = c[-3 * tp];
    if (!((p0 < lo) | (p0 > hi) | (p8 < lo) | (p8 > hi))) return false;
    const int p4 = c[3], p12 = c[-3];
    if (!((p4 < lo) | (p4 > hi) | (p12 < lo) | (p12 > hi))) return false;
    uint32_t mb = 0, md = 0;
#define ORBFE_RING(k, dx, dy)                      \
    {                                              \
        const int x = c[(dy) * tp + (dx)];         \
        mb |= (uint32_t)(x > hi) << (k);           \
        md |= (uint32_t)(x < lo) << (k);           \
    }
    ORBFE_RING(0, 0, 3) ORBFE_RING(1, 1, 3) ORBFE_RING(2, 2, 2) ORBFE_RING(3, 3, 1)
    ORBFE_RING(4, 3, 0) ORBFE_RING(5, 3, -1) ORBFE_RING(6, 2, -2) ORBFE_RING(7, 1, -3)
    ORBFE_RING(8, 0, -3) ORBFE_RING(9, -1, -3) ORBFE_RING(10, -2, -2) ORBFE_RING(11, -3, -1)
    ORBFE_RING(12, -3, 0) ORBFE_RING(13, -3, 1) ORBFE_RING(14, -2, 2) ORBFE_RING(15, -1, 3)
#undef ORBFE_RING
    return has_run9(mb) || has_run9(md);
}

// cornerScore<16>: max over the sixteen 9-arcs of max(min d, -max d), d = v - ring; returns that value - 1.
__device__ __forceinline__ int fast9_score(const uint8_t* c, int tp) {
    const int v = c[0];
    int d[16];
    d[0] = v - c[3 * tp];       d[1] = v - c[3 * tp + 1];    d[2] = v - c[2 * tp + 2];    d[3] = v - c[tp + 3];
    d[4] = v - c[3];            d[5] = v - c[-tp + 3];       d[6] = v - c[-2 * tp + 2];   d[7] = v - c[-3 * tp + 1];
    d[8] = v - c[-3 * tp];      d[9] = v - c[-3 * tp - 1];   d[10] = v - c[-2 * tp - 2];  d[11] = v - c[-tp - 3];
    d[12] = v - c[-3];          d[13] = v - c[tp - 3];       d[14] = v - c[2 * tp - 2];   d[15] = v - c[3 * tp - 1];
    int mn[16], mx[16], a[16], b[16];
#pragma unroll
    for (int k = 0; k < 16; k++) { mn[k] = min(d[k], d[(k + 1) & 15]); mx[k] = max(d[k], d[(k + 1) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; k++) { a[k] = min(mn[k], mn[(k + 2) & 15]); b[k] = max(mx[k], mx[(k + 2) & 15]); }
#pragma unroll
    for (int k = 0; k < 16; k++) { mn[k] = min(a[k], a[(k + 4) & 15]); mx[k] = max(b[k], b[(k + 4) & 15]); }
    // max over arcs of min(d) and min over arcs of max(d) are reduced separately and combined once: the per-arc form
    // max(best, max(lo9, -hi9)) is miscompiled by nvcc 12.9 for sm_100a (unsigned VIMNMX on the negated operand;
    // tools/dbg_fast_unit.cu reproduces it), and this form is also cheaper.
    int dark = -1024, bright = 1024;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        dark = max(dark, min(mn[k], d[(k + 8) & 15]));       // min of d[k..k+8]
        bright = min(bright, max(mx[k], d[(k + 8) & 15]));   // max of d[k..k+8]
    }
    const int best = max(dark, -bright);
    return best - 1;
}

// bytes of x that are > t (t <= 127), as 0x80 flags: exact SWAR compare (the low 7 bits cannot carry)
__device__ __forceinline__ uint32_t bytes_gt(uint32_t x, uint32_t k127mt) {
    return (((x & 0x7f7f7f7fu) + k127mt) | x) & 0x80808080u;
}

// score `cnt` (<= 32) listed positions (r << 10 | col, tile coordinates) into the score map
__device__ __forceinline__ void score_listed(const uint8_t* tile, uint8_t* smap, int tp, const unsigned short* list, int cnt,
                                             int lane, int t) {
    if (lane < cnt) {
        const int e = list[lane];
        const int off = (e >> 10) * tp + (e & 1023);
        // corner at threshold t  <=>  the largest arc threshold (score + 1) exceeds t, so the score alone decides
        const int sc = fast9_score(tile + off, tp);
        if (sc >= t) smap[off] = (uint8_t)sc;
    }
}

// NMS + emission over the tested region (tw x th, origin `sc0`, pitch sp) of a score map, raster order.  Scores outside
// the tested region count as 0 (cv::FAST never scores the 3-px border of the sub-image it is given).
__device__ __forceinline__ int nms_emit(const uint8_t* sc0, int sp, int tw, int th, uint32_t* out, int slotCap, int xAdd,
                                        int yAdd, int lane, uint32_t ltMask) {
    int total = 0;
    int cy = lane / tw, cx = lane - cy * tw;
    const int npix = tw * th;
    for (int base = 0; base < npix; base += 32) {
        int s = 0;
        if (base + lane < npix) s = sc0[cy * sp + cx];
        if (__any_sync(0xffffffffu, s != 0)) {
            bool keep = false;
            if (s != 0) {
                const uint8_t* q = sc0 + cy * sp + cx;
                const bool l = cx > 0, r = cx < tw - 1, u = cy > 0, d = cy < th - 1;
                keep = (!l || s > q[-1]) && (!r || s > q[1]) && (!u || (s > q[-sp] && (!l || s > q[-sp - 1]) && (!r || s > q[-sp + 1]))) &&
                       (!d || (s > q[sp] && (!l || s > q[sp - 1]) && (!r || s > q[sp + 1])));
            }
            const uint32_t m = __ballot_sync(0xffffffffu, keep);
            if (keep) {
                const int k = total + __popc(m & ltMask);
                if (k < slotCap) out[k] = (uint32_t)(cx + xAdd) | ((uint32_t)(cy + yAdd) << 12) | ((uint32_t)s << 24);
            }
            total += __popc(m);
        }
        cx += 32;
        while (cx >= tw) { cx -= tw; cy++; }
    }
    return total;
}

// Same as nms_emit for the strip score map, but 4 pixels (one aligned 32-bit word) per lane: almost all words are 0.
// X0 = tile column of the first tested pixel of the cell; the tested rows start at tile row 3.
__device__ __forceinline__ int nms_emit_words(const uint8_t* smap, int tp, int X0, int tw, int th, uint32_t* out, int slotCap,
                                              int xAdd, int yAdd, int lane, uint32_t ltMask) {
    const int wpr = tp >> 2, wA = X0 >> 2, nwc = ((X0 + tw + 3) >> 2) - wA;
    const int nitems = nwc * th;
    const uint32_t firstMask = ~((1u << (8 * (X0 & 3))) - 1u);
    const int lastBytes = X0 + tw - 4 * (wA + nwc - 1);    // valid bytes in the last word of a row: 1..4
    const uint32_t lastMask = lastBytes >= 4 ? 0xffffffffu : ((1u << (8 * lastBytes)) - 1u);
    const int rStep = 32 / nwc, wStep = 32 - rStep * nwc;
    int row = lane / nwc, w = lane - row * nwc;
    int total = 0;
    for (int base = 0; base < nitems; base += 32) {
        uint32_t word = 0;
        if (base + lane < nitems) {
            word = reinterpret_cast<const uint32_t*>(smap)[(3 + row) * wpr + wA + w];
            if (w == 0) word &= firstMask;
            if (w == nwc - 1) word &= lastMask;
        }
        if (__any_sync(0xffffffffu, word != 0)) {
            uint32_t keep4 = 0;
            if (word != 0) {
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int s = (word >> (8 * j)) & 0xff;
                    if (s == 0) continue;
                    const int cx = 4 * (wA + w) + j - X0, cy = row;
                    const uint8_t* q = smap + (3 + row) * tp + X0 + cx;
                    const bool l = cx > 0, r = cx < tw - 1, u = cy > 0, d = cy < th - 1;
                    const bool keep = (!l || s > q[-1]) && (!r || s > q[1]) &&
                                      (!u || (s > q[-tp] && (!l || s > q[-tp - 1]) && (!r || s > q[-tp + 1]))) &&
                                      (!d || (s > q[tp] && (!l || s > q[tp - 1]) && (!r || s > q[tp + 1])));
                    if (keep) keep4 |= 1u << j;
                }
            }
            if (__any_sync(0xffffffffu, keep4 != 0)) {
                // raster order: lane-major, byte-minor.  rank = kept bytes in lower lanes + kept lower bytes of this lane
                int lower = 0, all = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const uint32_t bm = __ballot_sync(0xffffffffu, (keep4 >> j) & 1u);
                    lower += __popc(bm & ltMask);
                    all += __popc(bm);
                }
                int before = 0;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if ((keep4 >> j) & 1u) {
                        const int k = total + lower + before;
                        const int cx = 4 * (wA + w) + j - X0;
                        const int s = (word >> (8 * j)) & 0xff;
                        if (k < slotCap) out[k] = (uint32_t)(cx + xAdd) | ((uint32_t)(row + yAdd) << 12) | ((uint32_t)s << 24);
                        before++;
                    }
                }
                total += all;
            }
        }
        row += rStep;
        w += wStep;
        if (w >= nwc) { w -= nwc; row++; }
    }
    return total;
}

// grid (nItems, n); block 32 * FAST_WARPS.  dynamic shared memory: strip tile | strip score map | per warp: private
// score tile (minThFAST fallback) + 64-entry survivor list.
__global__ void __launch_bounds__(32 * FAST_WARPS, 4) fast_cells_kernel(const __grid_constant__ Geom g,
                                                                      const uint8_t* __restrict__ pyr,
                                                                      const FastItem* __restrict__ items,
                                                                      uint32_t* __restrict__ slots,
                                                                      int* __restrict__ cellCount, int tileBytes,
                                                                      int scoreBytes, int listBytes) {
    extern __shared__ __align__(16) uint8_t smem[];
    const FastItem it = items[blockIdx.x];
    const int img = blockIdx.y;
    const LevelGeom& L = g.lv[it.level];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ltMask = (1u << lane) - 1u;

    const int maxBX = L.w - kMinBorder, maxBY = L.h - kMinBorder;
    const int iniY = kMinBorder + it.row * L.hCell;
    const int maxY = min(iniY + L.hCell + 6, maxBY);
    const int ch = maxY - iniY;
    const int cellRowBase = L.cellBase + it.row * L.nCols;
    int* counts = cellCount + (size_t)img * g.nCells + cellRowBase;

    const bool rowSkipped = (iniY >= maxBY - 3) || ch < 7;    // :812 (and cv::FAST finds nothing in < 7 rows)
    if (rowSkipped) {
        for (int jc = threadIdx.x; jc < it.ncols; jc += 32 * FAST_WARPS) counts[it.col0 + jc] = 0;
        return;
    }
    // strip columns [xs, xe) of the level, as plane columns [ax0, ax1)
    const int xs = kMinBorder + it.col0 * L.wCell;
    const int xe = min(kMinBorder + (it.col0 + it.ncols - 1) * L.wCell + L.wCell + 6, maxBX);
    const int ax0 = (kXOff + xs) & ~15;
    const int ax1 = (kXOff + max(xe, xs) + 15) & ~15;
    const int tp = ax1 - ax0;
    uint8_t* tile = smem;
    uint8_t* smap = smem + tileBytes;
    uint8_t* pscore = smem + 2 * tileBytes + warp * (scoreBytes + listBytes);
    unsigned short* list = reinterpret_cast<unsigned short*>(pscore + scoreBytes);

    {
        const uint8_t* plane = pyr + (size_t)img * g.slabBytes + L.planeOff + (size_t)(iniY + kEdge) * L.pitch + ax0;
        const int vecPerRow = tp >> 4;
        for (int i = threadIdx.x; i < vecPerRow * ch; i += 32 * FAST_WARPS) {
            const int r = i / vecPerRow, c = i - r * vecPerRow;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (ax0 + 16 * c + 16 <= L.pitch) v = __ldg(reinterpret_cast<const uint4*>(plane + (size_t)r * L.pitch + 16 * c));
            *reinterpret_cast<uint4*>(tile + r * tp + 16 * c) = v;
            *reinterpret_cast<uint4*>(smap + r * tp + 16 * c) = make_uint4(0, 0, 0, 0);
        }
    }
    __syncthreads();

    // ---- phase 1 (whole CTA, whole strip, iniThFAST): 4 pixels per lane in one 32-bit word --------------------
    // opposite-pair rejection: a 9-arc contains one pixel of each opposite ring pair, so a corner needs
    // (|p0-v| > t or |p8-v| > t) and (|p4-v| > t or |p12-v| > t).  Survivors are listed per warp (64 entries), the
    // full FAST-9 test + score runs on 32 listed pixels at a time.
    const int tIni = min(max(g.iniTh, 0), 255);
    {
        const int wpr = tp >> 2;                          // words per tile row
        const int colLo = kXOff + xs - ax0 + 3;           // tile columns of tested pixels: [colLo, colHi)
        const int colHi = kXOff + xe - ax0 - 3;
        const int wLo = colLo >> 2, wHi = (colHi + 3) >> 2;
        const int nw = wHi - wLo, nitems = nw * (ch - 6);
        const uint32_t k127mt = (uint32_t)(127 - min(tIni, 127)) * 0x01010101u;
        int cnt = 0;
        int r = threadIdx.x / nw, wi = threadIdx.x - r * nw;
        const int rStep = (32 * FAST_WARPS) / nw, wStep = (32 * FAST_WARPS) - rStep * nw;
        for (int base = warp * 32; base < nitems; base += 32 * FAST_WARPS) {
            uint32_t m = 0;
            const int w = wLo + wi;
            if (base + lane < nitems) {
                const uint32_t* row = reinterpret_cast<const uint32_t*>(tile + (r + 3) * tp);
                const uint32_t c = row[w], cp = row[max(w - 1, 0)], cn = row[min(w + 1, wpr - 1)];
                const uint32_t up = row[w + 3 * wpr], dn = row[w - 3 * wpr];
                const uint32_t p4 = __byte_perm(c, cn, 0x6543), p12 = __byte_perm(cp, c, 0x4321);
                if (tIni <= 127)
                    m = (bytes_gt(__vabsdiffu4(up, c), k127mt) | bytes_gt(__vabsdiffu4(dn, c), k127mt)) &
                        (bytes_gt(__vabsdiffu4(p4, c), k127mt) | bytes_gt(__vabsdiffu4(p12, c), k127mt));
                else
                    m = 0x80808080u;    // no cheap rejection for thresholds above 127
                // bytes outside [colLo, colHi)
                const int c0 = 4 * w;
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if (c0 + j < colLo || c0 + j >= colHi) m &= ~(0x80u << (8 * j));
            }
            if (__any_sync(0xffffffffu, m != 0)) {
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const bool sv = (m >> (8 * j + 7)) & 1u;
                    const uint32_t bm = __ballot_sync(0xffffffffu, sv);
                    if (sv) list[cnt + __popc(bm & ltMask)] = (unsigned short)(((r + 3) << 10) | (4 * w + j));
                    cnt += __popc(bm);
                    __syncwarp();
                    if (cnt >= 32) {
                        score_listed(tile, smap, tp, list, 32, lane, tIni);
                        const int rest = cnt - 32;
                        const unsigned short mv = lane < rest ? list[32 + lane] : 0;
                        __syncwarp();
                        if (lane < rest) list[lane] = mv;
                        cnt = rest;
                        __syncwarp();
                    }
                }
            }
            r += rStep;
            wi += wStep;
            if (wi >= nw) { wi -= nw; r++; }
        }
        score_listed(tile, smap, tp, list, cnt, lane, tIni);
    }
    __syncthreads();

    // ---- phase 2: one warp per cell: NMS + emission on the strip score map; minThFAST fallback for empty cells ----
    for (int jc = warp; jc < it.ncols; jc += FAST_WARPS) {
        const int j = it.col0 + jc;
        const int iniX = kMinBorder + j * L.wCell;
        int total = 0;
        if (iniX < maxBX - 6) {    // :821
            const int maxX = min(iniX + L.wCell + 6, maxBX);
            const int cw = maxX - iniX;
            const int tw = cw - 6, th = ch - 6;
            const int off0 = 3 * tp + (kXOff + iniX - ax0) + 3;    // first tested pixel of the cell in tile coordinates
            uint32_t* out = slots + ((size_t)img * g.nCells + cellRowBase + j) * g.slotCap;
            const int xAdd = j * L.wCell + 3, yAdd = it.row * L.hCell + 3;
            total = nms_emit_words(smap, tp, kXOff + iniX - ax0 + 3, tw, th, out, g.slotCap, xAdd, yAdd, lane, ltMask);
            if (total == 0) {
                // :845-848: nothing at iniThFAST -> the cell is searched again at minThFAST, on a private score tile
                const int t = min(max(g.minTh, 0), 255);
                const uint8_t* cell = tile + off0;
                for (int i = lane; i < (tw * th + 3) / 4; i += 32) reinterpret_cast<uint32_t*>(pscore)[i] = 0;
                __syncwarp();
                const int npix = tw * th;
                int cnt = 0;
                int cy = lane / tw, cx = lane - cy * tw;
                for (int base = 0; base < npix; base += 32) {
                    bool sv = false;
                    if (base + lane < npix) {
                        const uint8_t* c = cell + cy * tp + cx;
                        const int v = c[0], lo = v - t, hi = v + t;
                        const int p0 = c[3 * tp], p8 = c[-3 * tp];
                        if ((p0 < lo) | (p0 > hi) | (p8 < lo) | (p8 > hi)) {
                            const int p4 = c[3], p12 = c[-3];
                            sv = (p4 < lo) | (p4 > hi) | (p12 < lo) | (p12 > hi);
                        }
                    }
                    const uint32_t bm = __ballot_sync(0xffffffffu, sv);
                    if (sv) list[cnt + __popc(bm & ltMask)] = (unsigned short)((cy << 8) | cx);
                    cnt += __popc(bm);
                    __syncwarp();
                    if (cnt >= 32) {
                        const int e = list[lane];
                        const uint8_t* c = cell + (e >> 8) * tp + (e & 0xff);
                        if (fast9_is_corner(c, tp, t)) pscore[(e >> 8) * tw + (e & 0xff)] = (uint8_t)fast9_score(c, tp);
                        const int rest = cnt - 32;
                        const unsigned short mv = lane < rest ? list[32 + lane] : 0;
                        __syncwarp();
                        if (lane < rest) list[lane] = mv;
                        cnt = rest;
                        __syncwarp();
                    }
                    cx += 32;
                    while (cx >= tw) { cx -= tw; cy++; }
                }
                if (lane < cnt) {
                    const int e = list[lane];
                    const uint8_t* c = cell + (e >> 8) * tp + (e & 0xff);
                    if (fast9_is_corner(c, tp, t)) pscore[(e >> 8) * tw + (e & 0xff)] = (uint8_t)fast9_score(c, tp);
                }
                __syncwarp();
                total = nms_emit(pscore, tw, tw, th, out, g.slotCap, xAdd, yAdd, lane, ltMask);
            }
        }
        if (lane == 0) counts[j] = total;
    }
}

// host: work items of one geometry.  Every cell row of every level is cut into runs of <= FAST_ITEM_COLS cells.
int fast_item_table(const Geom& g, FastItem* out, int cap, int* tileBytes, int* scoreBytes, int* listBytes) {
    int n = 0, maxTile = 16, maxScore = 4;
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& L = g.lv[l];
        for (int r = 0; r < L.nRows; r++)
            for (int c0 = 0; c0 < L.nCols; c0 += FAST_ITEM_COLS) {
                const int nc = L.nCols - c0 < FAST_ITEM_COLS ? L.nCols - c0 : FAST_ITEM_COLS;
                if (out && n < cap) out[n] = FastItem{(short)l, (short)r, (short)c0, (short)nc};
                n++;
                const int tp = ((nc * L.wCell + 6 + 15 + 15) & ~15) + 16;
                const int ch = L.hCell + 6;
                if (tp * ch > maxTile) maxTile = tp * ch;
            }
        const int sb = ((L.wCell + 6) * (L.hCell + 6) + 3) & ~3;
        if (L.nCols > 0 && L.nRows > 0 && sb > maxScore) maxScore = sb;
    }
    *tileBytes = (maxTile + 15) & ~15;
    *scoreBytes = (maxScore + 15) & ~15;
    *listBytes = 128;    // 64 u16 survivor entries per warp
    return n;
}

int fast_smem_bytes(int tileBytes, int scoreBytes, int listBytes) {
    return 2 * tileBytes + FAST_WARPS * (scoreBytes + listBytes);
}

int launch_fast(const Geom& g, const uint8_t* pyr, int n, const FastItem* items, int nItems, int tileBytes,
                int scoreBytes, int listBytes, uint32_t* slots, int* cellCount, cudaStream_t st) {
    if (nItems <= 0) return 0;
    const int smem = fast_smem_bytes(tileBytes, scoreBytes, listBytes);
    static int configured = 0;
    if (smem > 48 * 1024 && configured < smem) {
        cudaFuncSetAttribute(fast_cells_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        configured = smem;
    }
    fast_cells_kernel<<<dim3(nItems, n), 32 * FAST_WARPS, smem, st>>>(g, pyr, items, slots, cellCount, tileBytes,
                                                                      scoreBytes, listBytes);
    return 1;
}

}    // namespace orbfe
