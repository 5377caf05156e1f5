// orbfe_internal.h — shared declarations of the liborbfe.so translation units (host + device).
//
// Data layout in HBM (DESIGN.md "Data layout"):
//   pyramid slab   one per image; level l is a padded u8 plane of (h_l + 38) rows, `pitch_l` bytes per row
//                  (multiple of 64).  The level's pixel (0,0) sits at row 19, column 32, so ROI rows start 16-byte
//                  aligned; the REFLECT_101 frame of 19 px (ORBextractor_old.cc:1346-1352) occupies columns 13..31
//                  and w+32..w+50.  Planes of one image are packed back to back (256-byte aligned).
//   blurred slab   same geometry; only the w x h ROI is written (GaussianBlur output, ORBextractor_old.cc:1147).
//   cell slots     per image, per FAST cell: `slotCap` packed candidates (x | y<<12 | response<<24, relative to
//                  (minBorderX, minBorderY)) + one count per cell; cells are numbered level-major, row-major,
//                  which is exactly the reference's insertion order (ORBextractor_old.cc:807-871).
//   level keys     per image, per level: the post-octree keypoints (same packing) in std::list order + count.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "../../include/orbfe_c.h"

namespace orbfe {

constexpr int kEdge = 19;         // EDGE_THRESHOLD, ORBextractor_old.cc:75
constexpr int kHalfPatch = 15;    // HALF_PATCH_SIZE, :74
constexpr int kPatch = 31;        // PATCH_SIZE, :73
constexpr int kXOff = 32;         // column of level pixel x = 0 inside a padded plane row
constexpr int kMinBorder = 16;    // EDGE_THRESHOLD - 3, :791
constexpr int kCellW = 35;        // W of ComputeKeyPointsOctTree, :787

struct LevelGeom {
    int w, h;            // level size (cvRound(cols * invScale), :1335)
    int pitch;           // bytes per plane row
    int nfeat;           // mnFeaturesPerLevel[l], :436-447
    int nCols, nRows;    // FAST cell grid, :802-803
    int wCell, hCell;    // :804-805
    int cellBase;        // index of cell (0,0) of this level inside the image's cell array
    int nIni;            // root nodes of DistributeOctTree, :561
    int keyCap;          // capacity of the level's post-octree keypoint list
    int keyBase;         // offset of that list inside the image's level-key array
    int maxCand;         // upper bound of FAST candidates of the level (sum of cell capacities)
    int scaledPatch;     // (int)(PATCH_SIZE * mvScaleFactor[l]), :891
    float scale;         // mvScaleFactor[l]
    float hX;            // (maxX - minX) / nIni, :563
    unsigned long long planeOff;    // byte offset of the plane inside a slab
};

struct Geom {
    int nlevels;
    int width, height;
    int iniTh, minTh;
    int nCells;          // cells per image (all levels)
    int slotCap;         // candidate capacity per cell
    int keyTotal;        // level-key entries per image (sum of keyCap)
    int outCap;          // == keyTotal: max keypoints per image
    int maxCellW, maxCellH;    // largest cell sub-image (wCell + 6, hCell + 6) over the levels
    unsigned long long slabBytes;    // bytes of one pyramid (or blurred) slab
    LevelGeom lv[ORBFE_MAX_LEVELS];
    int umax[kHalfPatch + 1];
};

// fingerprint of everything a tensor map of a level plane depends on besides the slab pointer and the image count
inline unsigned long long geom_plane_key(const Geom& g) {
    unsigned long long k = 1469598103934665603ull;
    for (int l = 0; l < g.nlevels; l++) {
        const unsigned long long v[4] = {(unsigned long long)g.lv[l].planeOff, (unsigned long long)g.lv[l].pitch,
                                         (unsigned long long)g.lv[l].w, (unsigned long long)g.lv[l].h};
        for (unsigned long long x : v) k = (k ^ x) * 1099511628211ull;
    }
    return (k ^ g.slabBytes) * 1099511628211ull;
}

// one FAST work item = a run of cells of one cell row (k_fast.cu).  Everything that depends only on the geometry is
// worked out once on the host (fast_item_table), so a CTA starts with three 16-byte loads instead of ~130 scalar
// instructions per thread.
struct alignas(16) FastItem {
    short level, row, col0, ncols;
    short ch, tp;             // tile rows (0: the cell row is skipped, ORBextractor_old.cc:812) and tile pitch
    short colLo, colHi;       // tested tile columns [colLo, colHi)
    short wCell, rpu;         // cell pitch; rows per phase-A unit (4..8), chosen to fill the CTA's threads
    short xAddBase, yAdd;     // candidate coordinates relative to (minBorderX, minBorderY): col0 * wCell + 3, row * hCell + 3
    unsigned tpMagic;         // floor(2^32 / tp) + 1
    unsigned tileOff;         // byte offset of the tile's first byte inside the level plane
    unsigned rowBytes;        // bytes per bulk-copied tile row
    int cellRowBase;          // index of the cell row's first cell inside the image's cell array
    unsigned boxMagic;        // floor(2^32 / boxW) + 1
    short boxW, boxH;         // tensor-map path: the level's TMA box (tile pitch and rows), same for all its items
};

// one CUtensorMap (128 bytes, 64-byte aligned) per pyramid level: u8 tensor {pitch, plane rows, images} over the
// pyramid slabs of one launch (k_fast.cu: fast_encode_maps); passed to the kernel as a __grid_constant__ parameter
struct alignas(64) FastMaps {
    unsigned char m[ORBFE_MAX_LEVELS][128];
};
static_assert(sizeof(FastItem) == 48, "FastItem is loaded as three 16-byte words");

// resize tables: per destination column / row (OpenCV linear resize coefficients)
struct ResizeTap {
    short s;         // source index (clamped)
    short s1;        // second tap index (clamped to src - 1)
    short a0, a1;    // weights * 2048
};

// shared-memory window of the resize tiles of one level (k_pyramid.cu)
struct ResizeDims {
    int SP, SH;    // pitch (multiple of 16) and rows of the staged source window
    int rpt;       // destination rows per thread (tile height = 4 * rpt)
    // resize_rows_kernel: tables in padded plane coordinates, built once per extractor (resize_pad_tables), so that the
    // kernel does no border reflection and no range search
    const ResizeTap* xpad;     // per plane column: s = byte offset of the first tap in the tile's staged window, a0, a1
    const ResizeTap* ypad;     // per padded plane row: the row's vertical tap (s, s1 absolute source rows)
    const ResizeTap* tileX;    // per x tile: s = first staged plane column, s1 = 16-byte vectors per row, (a0, a1) = 2^31 / s1 + 1
    const ResizeTap* tileY;    // per y tile: s = first staged source row, s1 = staged rows
};

// per-image results of the pack step
struct ImageResult {
    int count;     // nkeypoints
    int mono;      // monoIndex (ORBextractor_old.cc:1190)
};

// work record for the angle + descriptor kernel
struct KpWork {
    unsigned short x, y;    // level coordinates
    unsigned char level, response;
    unsigned short pad;
    int dst;                // final row in the caller's keypoint / descriptor arrays
};

struct ExtractBuffers {
    uint8_t* pyr;            // n slabs
    uint8_t* blur;           // n slabs
    uint32_t* slots;         // n * nCells * slotCap
    int* cellCount;          // n * nCells
    uint32_t* levelKeys;     // n * keyTotal
    int* levelCount;         // n * nlevels
    uint32_t* scratch;       // octree key ping-pong: n * scratchPerImage
    unsigned long long scratchPerImage;
    KpWork* work;            // n * outCap
    ImageResult* result;     // n
    int* errFlag;            // 1 int, device
};

// ---- kernel launchers (each returns the number of kernels it launched).  The ones whose argument lists carry
// per-stage tables (FAST items, blur tiles, octree capacities, kNN) are declared next to their only caller in
// orbfe_api.cu. ----
int launch_pyramid(const Geom& g, const uint8_t* images, int n, int stride, size_t imageStride, uint8_t* pyr,
                   const ResizeTap* const* xtab, const ResizeTap* const* ytab, const ResizeDims* dims,
                   const ResizeDims* rows16, cudaStream_t st);
void resize_tile_dims(const ResizeTap* xtab, const ResizeTap* ytab, int dw, int dh, int rpt, int* SP, int* SH);
void resize_pad_counts(int dw, int dh, int rpt, int* nxpad, int* nypad, int* ntx, int* nty);
void resize_pad_tables(const ResizeTap* xtab, const ResizeTap* ytab, int dw, int dh, int rpt, ResizeTap* xpad, ResizeTap* ypad,
                       ResizeTap* tileX, ResizeTap* tileY);
int launch_pack(const Geom& g, int n, const ExtractBuffers& b, int lap0, int lap1, cudaStream_t st);
int launch_describe(const Geom& g, int n, const ExtractBuffers& b, orbfe_keypoint* kps, uint8_t* desc, int cap,
                    cudaStream_t st);

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of the (kernel, device) pair: one of these per kernel
// remembers, per device, the largest dynamic shared-memory size this process has opted the kernel into.  Handles on
// different devices (and different threads) share it, hence the lock.
struct SmemOptIn {
    std::mutex mu;
    int have[64] = {0};
};

}    // namespace orbfe

#ifdef __CUDACC__
// ---- TMA bulk copies (cp.async.bulk, SASS UBLKCP) completing on a shared-memory mbarrier ---------------------------
// Users today: k_fast.cu (strip tiles: one tiled load each, or one bulk copy per row with ORBFE_FAST_TMAP=0).
// Bulk-copy addresses and sizes must be multiples of 16 bytes.
namespace orbfe {
template <typename Kernel>
inline void smem_opt_in(SmemOptIn& s, Kernel kernel, size_t bytes) {
    if (bytes <= 32 * 1024) return;    // the 48 KiB default limit counts static + dynamic: opt in early
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 63;
    std::lock_guard<std::mutex> lk(s.mu);
    if (s.have[dev] >= (int)bytes) return;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) == cudaSuccess)
        s.have[dev] = (int)bytes;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// one arrival that also announces `bytes` of bulk-copy traffic
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "ORBFE_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra ORBFE_DONE_%=;\n"
        "bra ORBFE_WAIT_%=;\n"
        "ORBFE_DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// one tiled TMA load (cp.async.bulk.tensor, SASS UTMALDG) of a 3-D box at (x, y, z); out-of-bounds elements are
// zero-filled and still counted: the transaction size is always the full box
__device__ __forceinline__ void tma_load_3d(void* dstSmem, const void* tmap, int x, int y, int z, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
            smem_u32(dstSmem)),
        "l"(tmap), "r"(x), "r"(y), "r"(z), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dstSmem, const void* srcGlobal, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dstSmem)),
                 "l"(srcGlobal), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
}    // namespace orbfe
#endif
