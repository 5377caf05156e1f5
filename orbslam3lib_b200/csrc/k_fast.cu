// k_fast.cu — the per-cell FAST stage of ORBextractor::ComputeKeyPointsOctTree (ORBextractor_old.cc:783-874).
//
// For every W=35 cell of every level the reference runs cv::FAST(cell + 6 px, iniThFAST, nonmax=true) and, only if
// that returns nothing, again with minThFAST (:828-848).  cv::FAST semantics: FAST-9/16 on the
// Bresenham circle of radius 3, tested on the cell sub-image minus a 3-px border, score = the largest threshold at
// which the pixel is still a corner, minus one; 3x3 non-max suppression with strict ">" where everything outside
// the tested window of the cell scores 0; output in raster order.  The tested windows of adjacent cells are
// disjoint and contiguous (a cell is wCell + 6 wide and loses 3 px on each side), so every pixel is scored once.
//
// Mapping: one CTA stages a run of up to FAST_ITEM_COLS cells of one cell row (a strip of the level plane) into
// shared memory with ONE tiled TMA load (cp.async.bulk.tensor.3d on the level's CUtensorMap; fallback: one bulk copy
// per tile row) and runs up to two passes (iniThFAST for all cells, then
// minThFAST for the cells that came back empty) of a four-phase pipeline, every phase using all threads of the CTA.
// Everything that depends only on the geometry (tile pitch, tested columns, offsets, rows per unit) comes precomputed
// in the 48-byte work item (fast_item_table).
//   A  test      a thread owns one 32-bit word column (4 px) x R rows (R = 4..8, picked per item so that the units
//                fill the CTA's threads): SWAR opposite-pair rejection (a 9-arc contains
//                one pixel of each opposite ring pair, so a corner needs |p0-v|>t or |p8-v|>t, and |p4-v|>t or
//                |p12-v|>t) on 4 pixels at a time (VABSDIFF4 + carry-free byte compares); the survivor flags of the
//                thread are compacted into a shared survivor list (warp scan of the popcounts, one shared atomic
//                per warp);
//   B  score     one thread per listed pixel: the 16 ring differences are packed two per register (k, k+8) and the
//                sixteen 9-arc minima / maxima come from two levels of 3-input packed min/max (VIMNMX3.S16x2);
//                score >= t  <=>  corner at t, so the score alone decides; scores go to a strip-wide score map;
//   C  nms       the list is walked again: every scored pixel compares with its 8 neighbours inside its own cell's
//                tested window and sets its bit in a keep bitmap;
//   D  emit      one warp per cell: the window's column masks (<= 4 bitmap words) are built once, a lane owns window
//                rows (lane, lane + 32), both counts go through one packed warp scan for the raster-order ranks,
//                candidates go to the cell's slot array as packed (x | y << 12 | response << 24) relative to
//                (minBorderX, minBorderY) as in :865-870.
// Integer only; reads P bytes per image once (the 6-px overlaps hit L1/L2).
#include <cuda.h>

#include <cstdlib>
#include <cstring>

#include "orbfe_internal.h"

namespace orbfe {

#ifndef ORBFE_FAST_THREADS
#define ORBFE_FAST_THREADS 256
#endif
#ifndef ORBFE_FAST_COLS
#define ORBFE_FAST_COLS 5
#endif
constexpr int FAST_THREADS = ORBFE_FAST_THREADS;
constexpr int FAST_WARPS = FAST_THREADS / 32;
#ifndef ORBFE_FAST_MINB
#define ORBFE_FAST_MINB 5    // CTAs per SM the register allocation must allow (<= 51 registers at 256 threads)
#endif
constexpr int FAST_ITEM_COLS = ORBFE_FAST_COLS;    // <= FAST_WARPS (phase D gives a warp to every cell)
static_assert(FAST_ITEM_COLS <= FAST_WARPS, "phase D gives a warp to every cell");

__device__ __forceinline__ uint32_t swap16(uint32_t a) { return __byte_perm(a, a, 0x1032); }

// cornerScore<16> of cv::FAST: max over the sixteen 9-arcs of max(min d, -max d), d = v - ring, minus 1.
// c points at the centre pixel inside the shared strip tile (pitch tp).  Works on e = ring - v + 256 (biased, always
// positive, so two values pack into the 16-bit lanes of one register without borrows): lane lo of R[k] is ring
// position k, lane hi is position k + 8.
__device__ __forceinline__ int fast9_score(const uint8_t* c, int tp) {
    const uint32_t bias = (uint32_t)(256 - (int)c[0]) * 0x10001u;
    uint32_t R[10];
#define ORBFE_PAIR(k, dx0, dy0, dx1, dy1) \
    R[k] = (uint32_t)c[(dy0) * tp + (dx0)] + ((uint32_t)c[(dy1) * tp + (dx1)] << 16) + bias;
    ORBFE_PAIR(0, 0, 3, 0, -3)
    ORBFE_PAIR(1, 1, 3, -1, -3)
    ORBFE_PAIR(2, 2, 2, -2, -2)
    ORBFE_PAIR(3, 3, 1, -3, -1)
    ORBFE_PAIR(4, 3, 0, -3, 0)
    ORBFE_PAIR(5, 3, -1, -3, 1)
    ORBFE_PAIR(6, 2, -2, -2, 2)
    ORBFE_PAIR(7, 1, -3, -1, 3)
#undef ORBFE_PAIR
    R[8] = swap16(R[0]);
    R[9] = swap16(R[1]);
    uint32_t mn[14], mx[14];
#pragma unroll
    for (int k = 0; k < 8; k++) {    // 3 consecutive ring positions
        mn[k] = __vimin3_s16x2(R[k], R[k + 1], R[k + 2]);
        mx[k] = __vimax3_s16x2(R[k], R[k + 1], R[k + 2]);
    }
#pragma unroll
    for (int k = 0; k < 6; k++) {
        mn[8 + k] = swap16(mn[k]);
        mx[8 + k] = swap16(mx[k]);
    }
    uint32_t a9[8], b9[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {    // 9 consecutive ring positions: arc k (lo) and arc k + 8 (hi)
        a9[k] = __vimin3_s16x2(mn[k], mn[k + 3], mn[k + 6]);
        b9[k] = __vimax3_s16x2(mx[k], mx[k + 3], mx[k + 6]);
    }
    // max over arcs of min(e) ; min over arcs of max(e)
    uint32_t ma = __vimax3_s16x2(__vimax3_s16x2(a9[0], a9[1], a9[2]), __vimax3_s16x2(a9[3], a9[4], a9[5]),
                                 __vimax3_s16x2(a9[6], a9[7], a9[7]));
    uint32_t mb = __vimin3_s16x2(__vimin3_s16x2(b9[0], b9[1], b9[2]), __vimin3_s16x2(b9[3], b9[4], b9[5]),
                                 __vimin3_s16x2(b9[6], b9[7], b9[7]));
    const int maxMinE = max((int)(ma & 0xffffu), (int)(ma >> 16)) - 256;    // = -bright  (d = -e)
    const int minMaxE = min((int)(mb & 0xffffu), (int)(mb >> 16)) - 256;    // = -dark
    return max(maxMinE, -minMaxE) - 1;
}

// bytes of x that are > t (t <= 127), as 0x80 flags (other bits undefined): the low 7 bits cannot carry
__device__ __forceinline__ uint32_t bytes_gt_raw(uint32_t x, uint32_t k127mt) { return ((x & 0x7f7f7f7fu) + k127mt) | x; }

struct FastStrip {
    uint8_t* tile;           // strip pixels, pitch tp
    uint8_t* smap;           // strip score map, pitch tp
    uint32_t* kbits;         // keep bitmap, kw words per tile row
    unsigned short* list;    // survivor list: byte offsets into the tile (tile <= 64 KiB)
    int* cnt;                // [0] list length, [1] mask of cells that came back empty
    const uint8_t* cedge;    // per tile column: bit 0 = first column of its cell window, bit 1 = last column
    unsigned tpMagic;        // floor(2^32 / tp) + 1: off / tp == umulhi(off, tpMagic) for off < 65536
    int tp, kw, ch;          // tile pitch (multiple of 32), bitmap words per row (tp / 32), tile rows
    int colLo, colHi;        // tested tile columns [colLo, colHi)
    int wCell, ncols;        // cell pitch and number of cells in the strip
    int rpu;                 // rows per phase-A unit (4..8)
};

// cell (0 .. ncols-1) of a tested tile column
__device__ __forceinline__ int strip_cell(const FastStrip& s, int col) {
    const int rel = col - s.colLo;
    int j = 0;
#pragma unroll
    for (int q = 1; q < FAST_ITEM_COLS; q++) j += (q < s.ncols && rel >= q * s.wCell) ? 1 : 0;
    return j;
}

// phases A..C for threshold t over the cells of cellMask.  Ends with the keep bitmap complete (block synchronised).
__device__ __forceinline__ void fast_pass(const FastStrip& s, int t, uint32_t cellMask, bool allCells) {
    const int tid = threadIdx.x, lane = tid & 31;
    const int wpr = s.tp >> 2;
    const int wLo = s.colLo >> 2, nwc = ((s.colHi + 3) >> 2) - wLo;
    const int th = s.ch - 6;
    const int R = s.rpu;    // rows per unit: the host picks 4..8 so that the units fill the CTA's threads
    const int ngroups = (th + R - 1) / R;
    const uint32_t k127mt = (uint32_t)(127 - min(t, 127)) * 0x01010101u;
    const uint32_t* tw32 = reinterpret_cast<const uint32_t*>(s.tile);

    // ---- A: test + compaction --------------------------------------------------------------------------------
    const int nunits = nwc * ngroups;
    for (int u0 = 0; u0 < nunits; u0 += FAST_THREADS) {    // uniform trip count: the warp scan below needs all lanes
        const int u = u0 + tid;
        uint32_t acc = 0;
        int w = 0, r0 = 0;
        if (u < nunits) {
            const int g = u / nwc;
            w = wLo + (u - g * nwc);
            r0 = 3 + R * g;
            // bytes of this word column that are tested (and belong to a cell of cellMask)
            uint32_t bm = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int col = 4 * w + j;
                bool ok = col >= s.colLo && col < s.colHi;
                if (ok && !allCells) ok = (cellMask >> strip_cell(s, col)) & 1u;
                if (ok) bm |= 0x80u << (8 * j);
            }
            if (bm != 0) {
                const uint32_t* p = tw32 + r0 * wpr + w;
                if (t <= 127) {
#pragma unroll
                    for (int i = 0; i < 8; i++) {
                        if (i < 4 || i < R) {    // CTA-uniform
                            const uint32_t c = p[i * wpr], cp = p[i * wpr - 1], cn = p[i * wpr + 1];
                            const uint32_t up = p[(i + 3) * wpr], dn = p[(i - 3) * wpr];
                            const uint32_t p4 = __byte_perm(c, cn, 0x6543), p12 = __byte_perm(cp, c, 0x4321);
                            const uint32_t v = bytes_gt_raw(__vabsdiffu4(up, c), k127mt) | bytes_gt_raw(__vabsdiffu4(dn, c), k127mt);
                            const uint32_t h = bytes_gt_raw(__vabsdiffu4(p4, c), k127mt) | bytes_gt_raw(__vabsdiffu4(p12, c), k127mt);
                            acc = (acc >> 1) | (v & h & bm);
                        }
                    }
                    acc >>= 8 - R;    // row i of the unit -> bit i of its byte (flags never cross a byte: R <= 8 shifts)
                } else {
                    acc = bm | (bm >> 1) | (bm >> 2) | (bm >> 3) | (bm >> 4) | (bm >> 5) | (bm >> 6) | (bm >> 7);
                }
                // rows beyond the tested window (the last group may be partial)
                const int nvalid = min(R, th - R * g);
                acc &= ((1u << nvalid) - 1u) * 0x01010101u;
            }
        }
        // bit 8 j + i of acc: pixel (row r0 + i, column 4 w + j) survived
        const int n = __popc(acc);
        int inc = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        int base = 0;
        if (lane == 31 && inc > 0) base = atomicAdd(&s.cnt[0], inc);
        base = __shfl_sync(0xffffffffu, base, 31) + inc - n;
        const int o0 = r0 * s.tp + 4 * w;
        unsigned short* lp = s.list + base;
        if (__any_sync(0xffffffffu, n > 14)) {
            // survivors cluster along edges: when some lane holds many, 32 predicated steps (a predicated store and a
            // predicated pointer bump each) beat a loop whose trip count is the largest popcount of the warp
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int oi = o0 + i * s.tp;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if ((acc >> (8 * j + i)) & 1u) *lp++ = (unsigned short)(oi + j);
                }
            }
        } else {
            while (acc) {
                const int b = __ffs(acc) - 1;
                acc &= acc - 1;
                *lp++ = (unsigned short)(o0 + (b & 7) * s.tp + (b >> 3));
            }
        }
    }
    __syncthreads();

    // ---- B: score ----------------------------------------------------------------------------------------------
    const int nlist = s.cnt[0];
    for (int e = tid; e < nlist; e += FAST_THREADS) {
        const int off = s.list[e];
        const int sc = fast9_score(s.tile + off, s.tp);
        // corner at threshold t  <=>  the largest arc threshold (score + 1) exceeds t
        s.smap[off] = sc >= t ? (uint8_t)sc : (uint8_t)0;
    }
    __syncthreads();

    // ---- C: non-max suppression inside the pixel's own cell window ------------------------------------------------
    // Scores outside the tested rows / columns of the strip are 0 in the map, so only the borders BETWEEN the cell
    // windows of the strip need masking (cedge: bit 0 = first column of its window, bit 1 = last column).
    for (int e = tid; e < nlist; e += FAST_THREADS) {
        const int off = s.list[e];
        const uint8_t* q = s.smap + off;
        const int sc = q[0];
        if (sc == 0) continue;
        const int tp = s.tp;
        const int r = (int)__umulhi((unsigned)off, s.tpMagic);
        const int fl = s.cedge[off - r * tp];
        const int mc = max((int)q[-tp], (int)q[tp]);
        int mL = max(max((int)q[-tp - 1], (int)q[-1]), (int)q[tp - 1]);
        int mR = max(max((int)q[-tp + 1], (int)q[1]), (int)q[tp + 1]);
        if (fl & 1) mL = 0;
        if (fl & 2) mR = 0;
        if (sc > max(mc, max(mL, mR))) atomicOr(&s.kbits[off >> 5], 1u << (off & 31));
    }
    __syncthreads();
}

// phase D for the cells of cellMask: raster-order emission of the kept pixels of every cell window, one warp per cell.
// A window is at most wCell <= 97 columns wide, i.e. it spans at most four words of a keep-bitmap row: the four column
// masks are built once per warp.  A lane owns rows (lane, lane + 32) of a block of 64 window rows; both counts travel
// through ONE warp scan packed in 16-bit fields (<= 32 x 97 per field).  Totals are written to counts[]; empty cells
// are recorded in s.cnt[1] when markEmpty is set.
__device__ __forceinline__ void fast_emit(const FastStrip& s, uint32_t cellMask, bool markEmpty, uint32_t* slotsRow,
                                          int* counts, int slotCap, int col0, int xAddBase, int yAdd) {
    const int lane = threadIdx.x & 31, j = threadIdx.x >> 5;
    if (j >= s.ncols || !((cellMask >> j) & 1u)) return;
    const int x0 = s.colLo + j * s.wCell;
    const int x1 = min(x0 + s.wCell, s.colHi);
    int total = 0;
    if (x1 > x0) {
        const int th = s.ch - 6;
        const int wA = x0 >> 5, nw = ((x1 - 1) >> 5) - wA + 1;    // nw <= 4
        uint32_t m[4];
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const int lo = max(x0 - 32 * (wA + q), 0), hi = min(x1 - 32 * (wA + q), 32);
            m[q] = (q < nw && hi > lo) ? ((hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u)) : 0u;
        }
        uint32_t* out = slotsRow + (size_t)(col0 + j) * slotCap;
        const int xAdd = xAddBase + j * s.wCell - x0;
        for (int rbase = 0; rbase < th; rbase += 64) {
            const int rowA = rbase + lane, rowB = rowA + 32;
            const uint32_t* kA = s.kbits + (3 + rowA) * s.kw + wA;
            const uint32_t* kB = kA + 32 * s.kw;
            uint32_t n2 = 0;    // kept pixels of rowA | rowB << 16
#pragma unroll
            for (int q = 0; q < 4; q++) {
                if (q < nw) {
                    if (rowA < th) n2 += (uint32_t)__popc(kA[q] & m[q]);
                    if (rowB < th) n2 += (uint32_t)__popc(kB[q] & m[q]) << 16;
                }
            }
            uint32_t inc = n2;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += v;
            }
            const uint32_t tot2 = __shfl_sync(0xffffffffu, inc, 31);
            const int totA = (int)(tot2 & 0xffffu), totB = (int)(tot2 >> 16);
            if (n2 != 0) {
#pragma unroll
                for (int half = 0; half < 2; half++) {
                    const int n = half ? (int)(n2 >> 16) : (int)(n2 & 0xffffu);
                    if (n == 0) continue;
                    const int row = half ? rowB : rowA;
                    int k = total + (half ? totA + (int)(inc >> 16) : (int)(inc & 0xffffu)) - n;
                    const uint32_t* kr = half ? kB : kA;
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        uint32_t bits = q < nw ? kr[q] & m[q] : 0u;
                        while (bits) {
                            const int col = 32 * (wA + q) + __ffs(bits) - 1;
                            bits &= bits - 1;
                            const uint32_t sc = s.smap[(3 + row) * s.tp + col];
                            if (k < slotCap) out[k] = (uint32_t)(col + xAdd) | ((uint32_t)(row + yAdd) << 12) | (sc << 24);
                            k++;
                        }
                    }
                }
            }
            total += totA + totB;
        }
        if (markEmpty && total == 0 && lane == 0) atomicOr(&s.cnt[1], 1 << j);
    }
    if (lane == 0) counts[col0 + j] = total;
}

// grid (nItems, n); block FAST_THREADS.  dynamic shared memory: tile | score map | keep bitmap | survivor list.
// (A persistent variant of this kernel -- resident CTAs walking the (image, item) units with the next unit's tile
// prefetched by TMA into a second buffer -- was measured 35 % SLOWER (3.85 vs 2.80 ms per 1024 images): the block
// scheduler's staggered CTA starts interleave the phases of the CTAs sharing an SM better than lock-stepped loops.)
__global__ void __launch_bounds__(FAST_THREADS, ORBFE_FAST_MINB) fast_cells_kernel(const __grid_constant__ Geom g,
                                                                  const uint8_t* __restrict__ pyr,
                                                                  const FastItem* __restrict__ items,
                                                                  uint32_t* __restrict__ slots,
                                                                  int* __restrict__ cellCount, int tileBytes,
                                                                  int kbitsBytes, int listBytes,
                                                                  const __grid_constant__ FastMaps maps, int useMaps) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ int sCnt[2];
    __shared__ __align__(8) uint64_t sBar;
    __shared__ __align__(4) uint8_t sEdge[1024];
    const FastItem it = items[blockIdx.x];    // geometry-only quantities come precomputed from the host
    const int img = blockIdx.y;
    const LevelGeom& L = g.lv[it.level];
    const int ch = it.ch;
    const int cellRowBase = it.cellRowBase;
    int* counts = cellCount + (size_t)img * g.nCells + cellRowBase;

    if (ch == 0) {    // :812 (and cv::FAST finds nothing in < 7 rows)
        for (int jc = threadIdx.x; jc < it.ncols; jc += FAST_THREADS) counts[it.col0 + jc] = 0;
        return;
    }

    FastStrip s;
    s.tp = useMaps ? it.boxW : it.tp;    // tensor-map path: the tile has the level's box pitch
    s.kw = s.tp >> 5;
    s.tpMagic = useMaps ? it.boxMagic : it.tpMagic;
    s.cedge = sEdge;
    s.ch = ch;
    s.tile = smem;
    s.smap = smem + tileBytes;
    s.kbits = reinterpret_cast<uint32_t*>(smem + 2 * tileBytes);
    s.list = reinterpret_cast<unsigned short*>(smem + 2 * tileBytes + kbitsBytes);
    s.cnt = sCnt;
    s.colLo = it.colLo;
    s.colHi = it.colHi;
    s.wCell = it.wCell;
    s.ncols = it.ncols;
    s.rpu = it.rpu;

    {
        // staging: one tiled TMA load of the strip's box (or one bulk copy per tile row, issued by the lanes of warp 0),
        // completing on an mbarrier; meanwhile every thread clears the score map + keep bitmap (contiguous: one
        // 16-byte loop over both allocations) and the edge table.
        // Tile columns past the plane pitch are never copied: they lie beyond every tested pixel's ring.
        const uint8_t* plane = pyr + (size_t)img * g.slabBytes + L.planeOff + it.tileOff;
        const uint32_t rowBytes = it.rowBytes;
        if (threadIdx.x == 0) mbar_init(&sBar, 1);
        __syncthreads();
        if (useMaps) {
            // ONE tiled TMA load per strip (box = boxW x boxH x 1 image of the level's tensor map) instead of one bulk
            // copy per tile row: the TMA unit serves ~1 operation instead of ~44 per tile
            if (threadIdx.x == 0) {
                const int y0 = (int)(it.tileOff / (unsigned)L.pitch), x0 = (int)(it.tileOff - (unsigned)y0 * (unsigned)L.pitch);
                mbar_arrive_expect_tx(&sBar, (uint32_t)it.boxW * (uint32_t)it.boxH);
                tma_load_3d(s.tile, &maps.m[it.level], x0, y0, img, &sBar);
            }
        } else if (threadIdx.x < 32) {
            if (threadIdx.x == 0) mbar_arrive_expect_tx(&sBar, rowBytes * (uint32_t)ch);
            __syncwarp();
            for (int r = threadIdx.x; r < ch; r += 32) bulk_g2s(s.tile + r * s.tp, plane + (size_t)r * L.pitch, rowBytes, &sBar);
        }
        uint4* z = reinterpret_cast<uint4*>(s.smap);
        for (int i = threadIdx.x; i < (tileBytes + kbitsBytes) >> 4; i += FAST_THREADS) z[i] = make_uint4(0, 0, 0, 0);
        for (int i = threadIdx.x; i < (s.tp >> 2); i += FAST_THREADS) reinterpret_cast<uint32_t*>(sEdge)[i] = 0;
        if (threadIdx.x < 2) sCnt[threadIdx.x] = 0;
    }
    __syncthreads();
    if (threadIdx.x < it.ncols) {    // first / last column of every cell window (windows are >= 1 column wide)
        const int x0 = s.colLo + threadIdx.x * s.wCell, x1 = min(x0 + s.wCell, s.colHi);
        if (x1 > x0) {
            // a one-column window gets both flags: two byte stores from the same thread, no race between threads
            // because windows are disjoint
            sEdge[x0] = (uint8_t)(x1 - 1 == x0 ? 3 : 1);
            if (x1 - 1 != x0) sEdge[x1 - 1] = 2;
        }
    }
    mbar_wait(&sBar, 0);    // the strip tile has landed
    __syncthreads();

    uint32_t* slotsRow = slots + ((size_t)img * g.nCells + cellRowBase) * g.slotCap;
    const int xAddBase = it.xAddBase, yAdd = it.yAdd;
    const uint32_t all = (1u << it.ncols) - 1u;

    fast_pass(s, min(max(g.iniTh, 0), 255), all, true);
    fast_emit(s, all, true, slotsRow, counts, g.slotCap, it.col0, xAddBase, yAdd);
    __syncthreads();
    const uint32_t empty = (uint32_t)sCnt[1];
    if (empty == 0) return;
    // :845-848: nothing at iniThFAST -> those cells are searched again at minThFAST
    __syncthreads();
    if (threadIdx.x == 0) sCnt[0] = 0;
    __syncthreads();
    fast_pass(s, min(max(g.minTh, 0), 255), empty, false);
    fast_emit(s, empty, false, slotsRow, counts, g.slotCap, it.col0, xAddBase, yAdd);
}

// strip width (cells) of a level: at most FAST_ITEM_COLS; a TMA box is at most 256 elements wide; and the tile of a
// level with tall cells (small levels have few, tall cell rows) must not dictate the shared-memory allocation of every
// level -- its strips are narrowed until the tile fits ORBFE_FAST_TILE_BUDGET bytes
#ifndef ORBFE_FAST_TILE_BUDGET
#define ORBFE_FAST_TILE_BUDGET 12288
#endif
static int fast_level_cols(const LevelGeom& L) {
    int cols = FAST_ITEM_COLS;
    auto tpMax = [&](int c) { return ((c * L.wCell + 6 + 31 + 31) & ~31) + 32; };
    while (cols > 1 && (tpMax(cols) > 256 || tpMax(cols) * (L.hCell + 6) > ORBFE_FAST_TILE_BUDGET)) cols--;
    return cols;
}

// host: work items of one geometry.  Every cell row of every level is cut into runs of <= FAST_ITEM_COLS cells, and
// everything the kernel needs that depends only on the geometry is worked out here.
// tileBytes: strip tile (and score map) bytes; kbitsBytes: keep bitmap; listBytes: survivor list (worst case: every
// tested pixel survives).
int fast_item_table(const Geom& g, FastItem* out, int cap, int* tileBytes, int* kbitsBytes, int* listBytes) {
    int n = 0, maxTile = 16, maxK = 4, maxList = 16;
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& L = g.lv[l];
        const int maxBX = L.w - kMinBorder, maxBY = L.h - kMinBorder;
        if (L.wCell > 97) maxTile = 1 << 28;    // phase D: a cell window spans at most four bitmap words
        // strip width of the level: a TMA box is at most 256 elements wide
        const int cols = fast_level_cols(L);
        // the level's box: widest strip x tallest cell row (multiple of 32 columns)
        int boxW = 32;
        const int boxH = L.hCell + 6;
        for (int c0 = 0; c0 < L.nCols; c0 += cols) {
            const int nc = L.nCols - c0 < cols ? L.nCols - c0 : cols;
            const int xs = kMinBorder + c0 * L.wCell;
            int xe = kMinBorder + (c0 + nc - 1) * L.wCell + L.wCell + 6;
            if (xe > maxBX) xe = maxBX;
            const int tp = ((kXOff + (xe > xs ? xe : xs) + 31) & ~31) - ((kXOff + xs) & ~31);
            if (tp > boxW) boxW = tp;
        }
        if (boxW > 256 || boxH > 256) maxTile = 1 << 28;
        if (boxW * boxH > maxTile) maxTile = boxW * boxH;
        if ((boxW / 32) * 4 * boxH > maxK) maxK = (boxW / 32) * 4 * boxH;
        for (int r = 0; r < L.nRows; r++)
            for (int c0 = 0; c0 < L.nCols; c0 += cols) {
                const int nc = L.nCols - c0 < cols ? L.nCols - c0 : cols;
                const int tpMax = ((nc * L.wCell + 6 + 31 + 31) & ~31) + 32;
                if (out && n < cap) {
                    FastItem it{};
                    it.level = (short)l; it.row = (short)r; it.col0 = (short)c0; it.ncols = (short)nc;
                    const int iniY = kMinBorder + r * L.hCell;
                    const int maxY = iniY + L.hCell + 6 < maxBY ? iniY + L.hCell + 6 : maxBY;
                    int ch = maxY - iniY;
                    if (iniY >= maxBY - 3 || ch < 7) ch = 0;    // ORBextractor_old.cc:812
                    // strip columns [xs, xe) of the level, as plane columns [ax0, ax1)
                    const int xs = kMinBorder + c0 * L.wCell;
                    int xe = kMinBorder + (c0 + nc - 1) * L.wCell + L.wCell + 6;
                    if (xe > maxBX) xe = maxBX;
                    const int ax0 = (kXOff + xs) & ~31;
                    const int ax1 = (kXOff + (xe > xs ? xe : xs) + 31) & ~31;
                    const int tp = ax1 - ax0 > 32 ? ax1 - ax0 : 32;
                    it.ch = (short)ch;
                    it.tp = (short)tp;
                    it.colLo = (short)(kXOff + xs - ax0 + 3);
                    const int hi = kXOff + xe - ax0 - 3;
                    it.colHi = (short)(hi > it.colLo ? hi : it.colLo);
                    it.wCell = (short)L.wCell;
                    it.xAddBase = (short)(c0 * L.wCell + 3);
                    it.yAdd = (short)(r * L.hCell + 3);
                    it.tpMagic = 0xffffffffu / (unsigned)tp + 1u;
                    it.tileOff = (unsigned)((iniY + kEdge) * L.pitch + ax0);
                    it.rowBytes = (unsigned)(tp < L.pitch - ax0 ? tp : L.pitch - ax0);
                    it.cellRowBase = L.cellBase + r * L.nCols;
                    it.boxW = (short)boxW;
                    it.boxH = (short)boxH;
                    it.boxMagic = 0xffffffffu / (unsigned)boxW + 1u;
                    // rows per phase-A unit: a unit is one 4-px word column x R rows and costs ~R; the units are dealt
                    // to FAST_THREADS threads in rounds, so pick the R with the fewest rounds x R (ties: larger R)
                    const int th = ch - 6, nwc = ((it.colHi + 3) >> 2) - (it.colLo >> 2);
                    int bestR = 8, bestCost = 1 << 30;
                    for (int R = 8; R >= 4 && th > 0; R--) {
                        const int units = nwc * ((th + R - 1) / R);
                        const int cost = ((units + FAST_THREADS - 1) / FAST_THREADS) * (R + 2);    // +2: per-unit overhead
                        if (cost < bestCost) { bestCost = cost; bestR = R; }
                    }
                    it.rpu = (short)bestR;
                    if (tp > tpMax || tp > 1024) maxTile = 1 << 28;
                    out[n] = it;
                }
                n++;
                if (tpMax > 1024) maxTile = 1 << 28;    // sEdge[] holds 1024 columns
                const int chMax = L.hCell + 6;
                if (tpMax * chMax > maxTile) maxTile = tpMax * chMax;
                const int kb = (tpMax / 32) * 4 * chMax;
                if (kb > maxK) maxK = kb;
                const int lb = (nc * L.wCell + 8) * L.hCell * 2;
                if (lb > maxList) maxList = lb;
            }
    }
    *tileBytes = (maxTile + 15) & ~15;
    *kbitsBytes = (maxK + 15) & ~15;
    *listBytes = (maxList + 15) & ~15;
    return n;
}

int fast_smem_bytes(int tileBytes, int kbitsBytes, int listBytes) {
    // phase A reads up to 10 tile rows past the strip (masked afterwards): they must stay inside the allocation, and
    // list offsets are 16-bit
    if (tileBytes > 65536) return 1 << 30;
    return 2 * tileBytes + kbitsBytes + listBytes;
}

// ---- tensor maps ------------------------------------------------------------------------------------------------------
// One CUtensorMap per level over the pyramid slabs of a launch: u8, dims {pitch, plane rows, n images}, strides
// {pitch, slabBytes}, box {boxW, boxH, 1} (the level's box of fast_item_table), no swizzle, zero fill.  The driver entry
// point comes from cudaGetDriverEntryPoint, so liborbfe.so does not link libcuda.  ORBFE_FAST_TMAP=0 keeps the
// per-row bulk copies (A/B measurements); so does any failure to encode.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn fast_encode_fn() {
    static EncodeTiledFn fn = [] {
        const char* e = getenv("ORBFE_FAST_TMAP");
        if (e && atoi(e) == 0) return (EncodeTiledFn) nullptr;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess) {
            cudaGetLastError();
            return (EncodeTiledFn) nullptr;
        }
        return (EncodeTiledFn)p;
    }();
    return fn;
}

static bool fast_encode_maps(const Geom& g, const uint8_t* pyr, int n, FastMaps* out) {
    EncodeTiledFn enc = fast_encode_fn();
    if (!enc) return false;
    static_assert(sizeof(CUtensorMap) == 128, "FastMaps holds 128-byte tensor maps");
    for (int l = 0; l < g.nlevels; l++) {
        const LevelGeom& L = g.lv[l];
        const int maxBX = L.w - kMinBorder;
        const int cols = fast_level_cols(L);
        int boxW = 32;
        for (int c0 = 0; c0 < L.nCols; c0 += cols) {    // same box as fast_item_table
            const int nc = L.nCols - c0 < cols ? L.nCols - c0 : cols;
            const int xs = kMinBorder + c0 * L.wCell;
            int xe = kMinBorder + (c0 + nc - 1) * L.wCell + L.wCell + 6;
            if (xe > maxBX) xe = maxBX;
            const int tp = ((kXOff + (xe > xs ? xe : xs) + 31) & ~31) - ((kXOff + xs) & ~31);
            if (tp > boxW) boxW = tp;
        }
        const int boxH = L.hCell + 6;
        if (boxW > 256 || boxH > 256 || (L.pitch & 15) || (g.slabBytes & 15) || (L.planeOff & 15)) return false;
        const cuuint64_t dims[3] = {(cuuint64_t)L.pitch, (cuuint64_t)(L.h + 2 * kEdge), (cuuint64_t)n};
        const cuuint64_t strides[2] = {(cuuint64_t)L.pitch, (cuuint64_t)g.slabBytes};
        const cuuint32_t box[3] = {(cuuint32_t)boxW, (cuuint32_t)boxH, 1u};
        const cuuint32_t estr[3] = {1u, 1u, 1u};
        CUtensorMap tm;
        const CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<uint8_t*>(pyr) + L.planeOff, dims, strides, box,
                               estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return false;
        memcpy(out->m[l], &tm, 128);
    }
    return true;
}

int launch_fast(const Geom& g, const uint8_t* pyr, int n, const FastItem* items, int nItems, int tileBytes,
                int kbitsBytes, int listBytes, uint32_t* slots, int* cellCount, cudaStream_t st) {
    if (nItems <= 0 || n <= 0) return 0;
    const int smem = fast_smem_bytes(tileBytes, kbitsBytes, listBytes);
    static SmemOptIn optIn;    // static usage here is <= 16 KiB
    smem_opt_in(optIn, fast_cells_kernel, (size_t)smem);
    // tensor maps of this launch, cached per thread on (slab pointer, image count, level planes); four entries, oldest
    // replaced: the pipelined host path alternates between two workspace halves and a few stage sizes
    struct MapCache {
        const uint8_t* pyr = nullptr;
        int n = 0, ok = 0;
        unsigned long long key = 0;
        FastMaps maps;
    };
    static thread_local MapCache cache[4];
    static thread_local int cacheNext = 0;
    const unsigned long long key = geom_plane_key(g);
    MapCache* hit = nullptr;
    for (MapCache& c : cache)
        if (c.pyr == pyr && c.n == n && c.key == key) hit = &c;
    if (!hit) {
        hit = &cache[cacheNext];
        cacheNext = (cacheNext + 1) & 3;
        hit->pyr = pyr;
        hit->n = n;
        hit->key = key;
        memset(&hit->maps, 0, sizeof(hit->maps));
        hit->ok = fast_encode_maps(g, pyr, n, &hit->maps) ? 1 : 0;
    }
    MapCache& mc = *hit;
    fast_cells_kernel<<<dim3(nItems, n), FAST_THREADS, smem, st>>>(g, pyr, items, slots, cellCount, tileBytes, kbitsBytes,
                                                                   listBytes, mc.maps, mc.ok);
    return 1;
}

}    // namespace orbfe
