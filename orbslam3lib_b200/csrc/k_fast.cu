// k_fast.cu — the per-cell FAST stage of ORBextractor::ComputeKeyPointsOctTree (ORBextractor_old.cc:783-874).
//
// For every W=35 cell of every level the reference runs cv::FAST(cell + 6 px, iniThFAST, nonmax=true) and, only if
// that returns nothing, again with minThFAST (:828-848).  cv::FAST semantics: FAST-9/16 on the
// Bresenham circle of radius 3, tested on the cell sub-image minus a 3-px border, score = the largest threshold at
// which the pixel is still a corner, minus one; 3x3 non-max suppression with strict ">" where everything outside
// the tested region scores 0; output in raster order.  Cells overlap by 6 px and the same pixel may be reported by
// two cells — reproduced, not de-duplicated.
//
// Mapping: one CTA stages a run of up to 16 cells of one cell row (a strip of the level plane) into shared memory
// with 16-byte coalesced loads; each warp then owns whole cells and works in three compaction stages, each 32
// positions at a time with __ballot_sync / __popc ranks so that raster order is preserved:
// (A) every tested pixel: opposite-pair rejection on ring pixels 0/8 and 4/12 (5 shared loads), survivors compacted;
// (B) survivors: the 16 ring pixels are classified into a "brighter" and a "darker" 16-bit mask, a 9-run is detected
//     with shifts/ANDs, corners are scored (min/max over the sixteen 9-arcs with a log-step sliding window in
//     registers) into a per-warp score tile and compacted in place;
// (C) corners: 3x3 NMS on the score tile + emission, so the output order is deterministic and equals OpenCV's.  Candidates go to the cell's fixed slot array as packed
// (x | y << 12 | response << 24), coordinates relative to (minBorderX, minBorderY) as in :865-870.
//
// HBM-bound by design (reads P bytes per image once; the 6-px overlaps hit L1/L2); integer only.
#include "orbfe_internal.h"

namespace orbfe {

constexpr int FAST_WARPS = 8;
constexpr int FAST_ITEM_COLS = 16;

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

// grid (nItems, n); block 32 * FAST_WARPS.  dynamic shared memory: strip tile + per-warp score tile + corner list.
__global__ void __launch_bounds__(32 * FAST_WARPS) fast_cells_kernel(const __grid_constant__ Geom g,
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
    // strip columns [xs, xe) of the level, as plane columns [ax0, ax1)
    const int xs = kMinBorder + it.col0 * L.wCell;
    const int xe = min(kMinBorder + (it.col0 + it.ncols - 1) * L.wCell + L.wCell + 6, maxBX);
    const int ax0 = (kXOff + xs) & ~15;
    const int ax1 = (kXOff + max(xe, xs) + 15) & ~15;
    const int tp = ax1 - ax0;
    uint8_t* tile = smem;

    if (!rowSkipped) {
        const uint8_t* plane = pyr + (size_t)img * g.slabBytes + L.planeOff + (size_t)(iniY + kEdge) * L.pitch + ax0;
        const int vecPerRow = tp >> 4;
        for (int i = threadIdx.x; i < vecPerRow * ch; i += 32 * FAST_WARPS) {
            const int r = i / vecPerRow, c = i - r * vecPerRow;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (ax0 + 16 * c + 16 <= L.pitch) v = __ldg(reinterpret_cast<const uint4*>(plane + (size_t)r * L.pitch + 16 * c));
            *reinterpret_cast<uint4*>(tile + r * tp + 16 * c) = v;
        }
    }
    __syncthreads();

    uint8_t* score = smem + tileBytes + warp * (scoreBytes + listBytes);
    unsigned short* list = reinterpret_cast<unsigned short*>(score + scoreBytes);    // (cy << 8 | cx) entries

    for (int jc = warp; jc < it.ncols; jc += FAST_WARPS) {
        const int j = it.col0 + jc;
        const int iniX = kMinBorder + j * L.wCell;
        int total = 0;
        if (!rowSkipped && iniX < maxBX - 6) {    // :821
            const int maxX = min(iniX + L.wCell + 6, maxBX);
            const int cw = maxX - iniX;
            const int tw = cw - 6, th = ch - 6;
            const int npix = tw * th;
            const uint8_t* cell = tile + (kXOff + iniX - ax0) + 3 * tp + 3;    // first tested pixel; rows tp apart
            uint8_t* sc0 = score + 3 * cw + 3;                                   // its score
            uint32_t* out = slots + ((size_t)img * g.nCells + cellRowBase + j) * g.slotCap;
            const int xAdd = j * L.wCell + 3, yAdd = it.row * L.hCell + 3;
            const int cy0 = lane / tw, cx0 = lane - cy0 * tw;
            for (int pass = 0; pass < 2; pass++) {
                const int t = min(max(pass == 0 ? g.iniTh : g.minTh, 0), 255);
                for (int i = lane; i < (cw * ch + 3) / 4; i += 32) reinterpret_cast<uint32_t*>(score)[i] = 0;
                // stage A: opposite-pair rejection on every tested pixel (a 9-arc contains one pixel of each
                // opposite pair), survivors compacted in raster order
                int nA = 0;
                {
                    int cx = cx0, cy = cy0;
                    for (int base = 0; base < npix; base += 32) {
                        bool pass2 = false;
                        if (base + lane < npix) {
                            const uint8_t* c = cell + cy * tp + cx;
                            const int v = c[0], lo = v - t, hi = v + t;
                            const int p0 = c[3 * tp], p8 = c[-3 * tp];
                            if ((p0 < lo) | (p0 > hi) | (p8 < lo) | (p8 > hi)) {
                                const int p4 = c[3], p12 = c[-3];
                                pass2 = (p4 < lo) | (p4 > hi) | (p12 < lo) | (p12 > hi);
                            }
                        }
                        const uint32_t m = __ballot_sync(0xffffffffu, pass2);
                        if (pass2) list[nA + __popc(m & ltMask)] = (unsigned short)((cy << 8) | cx);
                        nA += __popc(m);
                        cx += 32;
                        while (cx >= tw) { cx -= tw; cy++; }
                    }
                }
                __syncwarp();
                // stage B: full FAST-9 test + score on the survivors; corners compacted in place (writes trail reads)
                int nC = 0;
                for (int base = 0; base < nA; base += 32) {
                    const int i = base + lane;
                    bool corner = false;
                    unsigned short e = 0;
                    if (i < nA) {
                        e = list[i];
                        const int cy = e >> 8, cx = e & 0xff;
                        const uint8_t* c = cell + cy * tp + cx;
                        corner = fast9_is_corner(c, tp, t);
                        if (corner) sc0[cy * cw + cx] = (uint8_t)fast9_score(c, tp);
                    }
                    const uint32_t m = __ballot_sync(0xffffffffu, corner);
                    __syncwarp();
                    if (corner) list[nC + __popc(m & ltMask)] = e;
                    nC += __popc(m);
                }
                __syncwarp();
                // stage C: 3x3 non-max suppression on the corners + emission in raster order
                total = 0;
                for (int base = 0; base < nC; base += 32) {
                    const int i = base + lane;
                    bool keep = false;
                    int cx = 0, cy = 0, s = 0;
                    if (i < nC) {
                        const unsigned short e = list[i];
                        cy = e >> 8;
                        cx = e & 0xff;
                        const uint8_t* sp = sc0 + cy * cw + cx;
                        s = sp[0];
                        keep = s > 0 && s > sp[-1] && s > sp[1] && s > sp[-cw - 1] && s > sp[-cw] && s > sp[-cw + 1] &&
                               s > sp[cw - 1] && s > sp[cw] && s > sp[cw + 1];
                    }
                    const uint32_t m = __ballot_sync(0xffffffffu, keep);
                    if (keep) {
                        const int k = total + __popc(m & ltMask);
                        if (k < g.slotCap) out[k] = (uint32_t)(cx + xAdd) | ((uint32_t)(cy + yAdd) << 12) | ((uint32_t)s << 24);
                    }
                    total += __popc(m);
                }
                __syncwarp();
                if (total > 0) break;    // :845: the lower threshold is tried only when the cell came back empty
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
    *listBytes = (2 * maxScore + 15) & ~15;    // one u16 per tested pixel (< cell area)
    return n;
}

int fast_smem_bytes(int tileBytes, int scoreBytes, int listBytes) { return tileBytes + FAST_WARPS * (scoreBytes + listBytes); }

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
