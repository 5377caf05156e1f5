// orb_oracle.cpp — CPU ORACLE (test infrastructure, NOT product code; see orb_oracle.h).
//
// Restates, with no OpenCV dependency, the CPU path of the reference:
//   ORBextractor_old.cc (ctor :411-471, DivideNode :482-538, compareNodes :540-555, DistributeOctTree :557-781,
//   ComputeKeyPointsOctTree :783-898, IC_Angle :78-105, computeOrbDescriptor :109-148, operator() :1088-1191,
//   ComputePyramid :1331-1356) and Frame.cc:45,1227-1231 (BFMatcher NORM_HAMMING knnMatch k=2).
// The OpenCV primitives (resize INTER_LINEAR 8U, copyMakeBorder REFLECT_101, FAST-9/16 + NMS, GaussianBlur 7x7
// sigma 2 fixed point, fastAtan2, cvRound) are restated from OpenCV 4.x's published algorithms; OpenCV is a
// third-party dependency of the reference (4.2.0 Android SDK, CMakeLists.txt:38-58) that is not vendored.
// Parity: pinned against cv2 4.13.0 in this container (tests/test_oracle_vs_cv2.py, tests/golden/).
//
// Floating point: built with -ffp-contract=off; every expression keeps the reference's types and order.
#include "orb_oracle.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cfloat>
#include <climits>
#include <cstring>
#include <list>
#include <thread>
#include <utility>
#include <vector>

#include "../../orbslam3lib_b200/csrc/orb_pattern.h"

namespace {

// ---- OpenCV scalar helpers -------------------------------------------------------------------------------------
inline int cvRoundD(double v) { return (int)std::lrint(v); }    // round-half-to-even (default FP environment)
inline int cvRoundF(float v) { return (int)std::lrintf(v); }
inline int cvFloorD(double v) { int i = (int)v; return i - (i > v); }
inline int cvCeilD(double v) { int i = (int)v; return i + (i < v); }
inline int reflect101(int p, int len) {
    // cv::borderInterpolate(p, len, BORDER_REFLECT_101)
    if ((unsigned)p < (unsigned)len) return p;
    if (len == 1) return 0;
    do {
        if (p < 0) p = -p; else p = 2 * len - p - 2;
    } while ((unsigned)p >= (unsigned)len);
    return p;
}

const int kPatch = 31, kHalfPatch = 15, kEdge = 19;

struct Timer {
    std::chrono::high_resolution_clock::time_point t0 = std::chrono::high_resolution_clock::now();
    double lap() {
        auto t1 = std::chrono::high_resolution_clock::now();
        double s = std::chrono::duration<double>(t1 - t0).count();
        t0 = t1;
        return s;
    }
};

// ---- cv::resize INTER_LINEAR, CV_8UC1 --------------------------------------------------------------------------
// imgproc/resize.cpp: coefficient setup of cv::resize + HResizeLinear<uchar,int,short,2048> +
// VResizeLinear<uchar,int,short,FixedPtCast<int,uchar,22>>.
void resizeLinearU8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
    // imgproc/resize.cpp, cv::resize: "if( interpolation == INTER_LINEAR && is_area_fast && iscale_x == 2 && iscale_y == 2 )
    // interpolation = INTER_AREA" -- an exact 2x decimation on both axes is the 2x2 box average of ResizeAreaFastVec for
    // uchar: (s00 + s01 + s10 + s11 + 2) >> 2
    if (sw == 2 * dw && sh == 2 * dh) {
        for (int y = 0; y < dh; y++) {
            const uint8_t* r0 = src + (size_t)(2 * y) * sstride;
            const uint8_t* r1 = r0 + sstride;
            for (int x = 0; x < dw; x++)
                dst[(size_t)y * dstride + x] = (uint8_t)((r0[2 * x] + r0[2 * x + 1] + r1[2 * x] + r1[2 * x + 1] + 2) >> 2);
        }
        return;
    }
    const double inv_x = (double)dw / sw, inv_y = (double)dh / sh;
    const double scale_x = 1. / inv_x, scale_y = 1. / inv_y;
    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> alpha(2 * dw), beta(2 * dh);
    for (int dx = 0; dx < dw; dx++) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cvFloorD(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        alpha[2 * dx] = (short)cvRoundF((1.f - fx) * 2048);
        alpha[2 * dx + 1] = (short)cvRoundF(fx * 2048);
    }
    for (int dy = 0; dy < dh; dy++) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cvFloorD(fy);
        fy -= sy;
        yofs[dy] = sy;
        beta[2 * dy] = (short)cvRoundF((1.f - fy) * 2048);
        beta[2 * dy + 1] = (short)cvRoundF(fy * 2048);
    }
    auto clipRow = [sh](int y) { return y < 0 ? 0 : (y < sh ? y : sh - 1); };
    std::vector<int> r0(dw), r1(dw);
    for (int dy = 0; dy < dh; dy++) {
        const uint8_t* s0 = src + (size_t)clipRow(yofs[dy]) * sstride;
        const uint8_t* s1 = src + (size_t)clipRow(yofs[dy] + 1) * sstride;
        for (int dx = 0; dx < dw; dx++) {
            int sx = xofs[dx], sx1 = sx + 1 < sw ? sx + 1 : sx;
            int a0 = alpha[2 * dx], a1 = alpha[2 * dx + 1];
            r0[dx] = s0[sx] * a0 + s0[sx1] * a1;
            r1[dx] = s1[sx] * a0 + s1[sx1] * a1;
        }
        int b0 = beta[2 * dy], b1 = beta[2 * dy + 1];
        uint8_t* d = dst + (size_t)dy * dstride;
        for (int dx = 0; dx < dw; dx++)
            d[dx] = (uint8_t)((((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2);
    }
}

void borderReflect101(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int b) {
    for (int y = -b; y < h + b; y++) {
        const uint8_t* s = src + (size_t)reflect101(y, h) * sstride;
        uint8_t* d = dst + (size_t)(y + b) * dstride;
        for (int x = -b; x < w + b; x++) d[x + b] = s[reflect101(x, w)];
    }
}

// ---- cv::GaussianBlur 7x7 sigma=2, CV_8U: fixed-point (Q8 kernel, ufixedpoint16/32) path of smooth.simd.hpp -----
void gaussian7x7U8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
    static const uint32_t k[7] = {18, 34, 48, 56, 48, 34, 18};    // getGaussianKernelBitExact(7, 2) * 256
    std::vector<uint16_t> hbuf((size_t)w * h);
    std::vector<int> xi(w + 6);
    for (int x = -3; x < w + 3; x++) xi[x + 3] = reflect101(x, w);
    for (int y = 0; y < h; y++) {
        const uint8_t* s = src + (size_t)y * sstride;
        uint16_t* hb = &hbuf[(size_t)y * w];
        for (int x = 0; x < w; x++) {
            uint32_t acc = 0;
            for (int i = 0; i < 7; i++) acc += k[i] * s[xi[x + i]];
            hb[x] = (uint16_t)acc;    // <= 255*256, no saturation
        }
    }
    for (int y = 0; y < h; y++) {
        const uint16_t* rows[7];
        for (int i = 0; i < 7; i++) rows[i] = &hbuf[(size_t)reflect101(y + i - 3, h) * w];
        uint8_t* d = dst + (size_t)y * dstride;
        for (int x = 0; x < w; x++) {
            uint32_t acc = 0;
            for (int i = 0; i < 7; i++) acc += k[i] * rows[i][x];
            d[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
}

// ---- cv::FAST TYPE_9_16 with non-max suppression (features2d/fast.cpp, fast_score.cpp) -------------------------
// Bresenham circle of radius 3, clockwise from (0,3).
const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// cornerScore<16> (fast_score.cpp): the largest threshold at which the pixel is still a FAST-9 corner, floored at
// `threshold`.  With d[k] = v - ring[k]:  max(threshold, max_arcs min(d), -min_arcs max(d)) - 1 over the sixteen
// 9-long circular arcs.  (OpenCV prunes arcs early; the value is the same.)
inline int fastCornerScore(const uint8_t* p, const int* off, int threshold) {
    int d[24];
    const int v = p[0];
    for (int k = 0; k < 24; k++) d[k] = v - p[off[k & 15]];
    int best = threshold;
    for (int k = 0; k < 16; k++) {
        int mn = d[k], mx = d[k];
        for (int i = 1; i < 9; i++) { mn = std::min(mn, d[k + i]); mx = std::max(mx, d[k + i]); }
        best = std::max(best, std::max(mn, -mx));
    }
    return best - 1;
}

struct XYR { int x, y, r; };

void fast9Nms(const uint8_t* img, int w, int h, int stride, int threshold, std::vector<XYR>& out) {
    out.clear();
    if (w < 7 || h < 7) return;
    threshold = std::min(std::max(threshold, 0), 255);
    int off[16];
    for (int k = 0; k < 16; k++) off[k] = kRingDy[k] * stride + kRingDx[k];
    // three rolling score rows (zero = not a corner) + corner position lists, as in FAST_t
    std::vector<uint8_t> sbuf((size_t)3 * w, 0);
    std::vector<int> cbuf((size_t)3 * (w + 1));
    uint8_t* buf[3] = {&sbuf[0], &sbuf[w], &sbuf[2 * (size_t)w]};
    int* cpbuf[3] = {&cbuf[0], &cbuf[w + 1], &cbuf[2 * (size_t)(w + 1)]};
    for (int i = 3; i < h - 2; i++) {
        const uint8_t* ptr = img + (size_t)i * stride + 3;
        uint8_t* curr = buf[(i - 3) % 3];
        int* cornerpos = cpbuf[(i - 3) % 3] + 1;
        std::memset(curr, 0, w);
        int ncorners = 0;
        if (i < h - 3) {
            for (int j = 3; j < w - 3; j++, ptr++) {
                int v = ptr[0];
                int lo = v - threshold, hi = v + threshold;
                // opposite-pair rejection cascade of FAST_t: a 9-arc contains one pixel of every opposite pair,
                // so the class (1 = darker, 2 = brighter) must survive the AND over all eight pairs.
                auto cls = [&](int k) { int x = ptr[off[k]]; return x < lo ? 1 : (x > hi ? 2 : 0); };
                int d = cls(0) | cls(8);
                if (d == 0) continue;
                d &= cls(2) | cls(10);
                d &= cls(4) | cls(12);
                d &= cls(6) | cls(14);
                if (d == 0) continue;
                d &= cls(1) | cls(9);
                d &= cls(3) | cls(11);
                d &= cls(5) | cls(13);
                d &= cls(7) | cls(15);
                bool corner = false;
                if (d & 1) {
                    int count = 0;
                    for (int k = 0; k < 25; k++) {
                        if (ptr[off[k & 15]] < lo) { if (++count > 8) { corner = true; break; } } else count = 0;
                    }
                }
                if (!corner && (d & 2)) {
                    int count = 0;
                    for (int k = 0; k < 25; k++) {
                        if (ptr[off[k & 15]] > hi) { if (++count > 8) { corner = true; break; } } else count = 0;
                    }
                }
                if (corner) {
                    cornerpos[ncorners++] = j;
                    curr[j] = (uint8_t)fastCornerScore(ptr, off, threshold);
                }
            }
        }
        cornerpos[-1] = ncorners;
        if (i == 3) continue;
        const uint8_t* prev = buf[(i - 4 + 3) % 3];
        const uint8_t* pprev = buf[(i - 5 + 3) % 3];
        cornerpos = cpbuf[(i - 4 + 3) % 3] + 1;
        ncorners = cornerpos[-1];
        for (int k = 0; k < ncorners; k++) {
            int j = cornerpos[k];
            int score = prev[j];
            if (score > prev[j + 1] && score > prev[j - 1] && score > pprev[j - 1] && score > pprev[j] &&
                score > pprev[j + 1] && score > curr[j - 1] && score > curr[j] && score > curr[j + 1])
                out.push_back({j, i - 1, score});
        }
    }
}

// ---- cv::fastAtan2 (core/mathfuncs_core: atan_f32) -------------------------------------------------------------
float fastAtan2f(float y, float x) {
    static const float p1 = 0.9997878412794807f * (float)(180 / 3.14159265358979323846);
    static const float p3 = -0.3258083974640975f * (float)(180 / 3.14159265358979323846);
    static const float p5 = 0.1555786518463281f * (float)(180 / 3.14159265358979323846);
    static const float p7 = -0.04432655554792128f * (float)(180 / 3.14159265358979323846);
    float ax = std::fabs(x), ay = std::fabs(y);
    float a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + (float)DBL_EPSILON);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + (float)DBL_EPSILON);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

// ---- Hamming kNN (k = 2), OpenCV tie rule: equal distances keep the lower train index first ---------------------
inline int hamming256(const uint64_t* a, const uint64_t* b) {
    return __builtin_popcountll(a[0] ^ b[0]) + __builtin_popcountll(a[1] ^ b[1]) +
           __builtin_popcountll(a[2] ^ b[2]) + __builtin_popcountll(a[3] ^ b[3]);
}
void knn2Range(const uint8_t* q, int64_t q0, int64_t q1, const uint8_t* t, int64_t nt, int32_t* idx, int32_t* dist) {
    for (int64_t i = q0; i < q1; i++) {
        uint64_t qa[4];
        std::memcpy(qa, q + i * 32, 32);
        int b1 = 1 << 30, b2 = 1 << 30;
        int64_t i1 = -1, i2 = -1;
        for (int64_t j = 0; j < nt; j++) {
            uint64_t tb[4];
            std::memcpy(tb, t + j * 32, 32);
            int d = hamming256(qa, tb);
            if (d < b1) { b2 = b1; i2 = i1; b1 = d; i1 = j; }
            else if (d < b2) { b2 = d; i2 = j; }
        }
        idx[2 * i] = (int32_t)i1; dist[2 * i] = i1 < 0 ? -1 : b1;
        idx[2 * i + 1] = (int32_t)i2; dist[2 * i + 1] = i2 < 0 ? -1 : b2;
    }
}

// ---- the extractor ---------------------------------------------------------------------------------------------
struct KeyPt {    // cv::KeyPoint
    float x = 0, y = 0, size = 0, angle = -1, response = 0;
    int octave = 0, class_id = -1;
};
struct Pt2i { int x = 0, y = 0; };

struct Node {    // ExtractorNode, ORBextractor_old.h:32-44
    std::vector<KeyPt> keys;
    Pt2i UL, UR, BL, BR;
    std::list<Node>::iterator lit;
    bool noMore = false;
    void divide(Node& n1, Node& n2, Node& n3, Node& n4) const;
};

void Node::divide(Node& n1, Node& n2, Node& n3, Node& n4) const {    // :482-538
    const int halfX = (int)std::ceil(static_cast<float>(UR.x - UL.x) / 2);
    const int halfY = (int)std::ceil(static_cast<float>(BR.y - UL.y) / 2);
    n1.UL = UL;
    n1.UR = {UL.x + halfX, UL.y};
    n1.BL = {UL.x, UL.y + halfY};
    n1.BR = {UL.x + halfX, UL.y + halfY};
    n2.UL = n1.UR;
    n2.UR = UR;
    n2.BL = n1.BR;
    n2.BR = {UR.x, UL.y + halfY};
    n3.UL = n1.BL;
    n3.UR = n1.BR;
    n3.BL = BL;
    n3.BR = {n1.BR.x, BL.y};
    n4.UL = n3.UR;
    n4.UR = n2.BR;
    n4.BL = n3.BR;
    n4.BR = BR;
    for (const KeyPt& kp : keys) {
        if (kp.x < n1.UR.x) {
            if (kp.y < n1.BR.y) n1.keys.push_back(kp); else n3.keys.push_back(kp);
        } else if (kp.y < n1.BR.y) n2.keys.push_back(kp);
        else n4.keys.push_back(kp);
    }
    n1.noMore = n1.keys.size() == 1;
    n2.noMore = n2.keys.size() == 1;
    n3.noMore = n3.keys.size() == 1;
    n4.noMore = n4.keys.size() == 1;
}

typedef std::pair<int, Node*> SizeNode;
thread_local int64_t g_sortTies = 0;
bool compareNodes(SizeNode& e1, SizeNode& e2) {    // :540-555
    if (e1.first < e2.first) return true;
    if (e1.first > e2.first) return false;
    if (e1.second != e2.second && e1.second->UL.x == e2.second->UL.x) g_sortTies++;
    return e1.second->UL.x < e2.second->UL.x;
}

}  // namespace

struct orc_extractor {
    int nfeatures, nlevels, iniTh, minTh;
    double scaleFactor;    // the reference stores the float ctor argument in a double member (ORBextractor_old.h:103)
    std::vector<float> scale, invScale, sigma2, invSigma2;
    std::vector<int> featuresPerLevel, umax;
    Pt2i pattern[512];
    // state of the last call
    struct Level {
        int w = 0, h = 0, pw = 0;
        std::vector<uint8_t> padded;     // (w+38) x (h+38), ROI at (19,19) — mvImagePyramid[level]
        std::vector<uint8_t> blurred;    // w x h, empty when the level produced no keypoint
        std::vector<XYR> candidates;     // vToDistributeKeys, relative to (minBorderX, minBorderY)
        std::vector<KeyPt> keypoints;    // after DistributeOctTree + border offset + angle (level coordinates)
        const uint8_t* roi() const { return &padded[(size_t)kEdge * pw + kEdge]; }
    };
    std::vector<Level> levels;
    double stage[7] = {0, 0, 0, 0, 0, 0, 0};
    int64_t sortTies = 0;

    orc_extractor(int nf, float sf, int nl, int ini, int mn);
    void computePyramid(const uint8_t* img, int w, int h, int stride);
    std::vector<KeyPt> distributeOctTree(const std::vector<KeyPt>& in, int minX, int maxX, int minY, int maxY, int N);
    void computeKeyPointsOctTree();
    int run(const uint8_t* img, int w, int h, int stride, int lap0, int lap1, orc_keypoint* kps, uint8_t* desc,
            int cap, int* count);
};

orc_extractor::orc_extractor(int nf, float sf, int nl, int ini, int mn)
    : nfeatures(nf), nlevels(nl), iniTh(ini), minTh(mn), scaleFactor(sf) {    // :411-471
    scale.resize(nlevels);
    sigma2.resize(nlevels);
    scale[0] = 1.0f;
    sigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        scale[i] = scale[i - 1] * scaleFactor;    // float * double -> double -> float, as in the reference
        sigma2[i] = scale[i] * scale[i];
    }
    invScale.resize(nlevels);
    invSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        invScale[i] = 1.0f / scale[i];
        invSigma2[i] = 1.0f / sigma2[i];
    }
    featuresPerLevel.resize(nlevels);
    float factor = 1.0f / scaleFactor;
    float nDesired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        featuresPerLevel[level] = cvRoundF(nDesired);
        sum += featuresPerLevel[level];
        nDesired *= factor;
    }
    featuresPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);
    for (int i = 0; i < 512; i++) {
        pattern[i].x = ORBFE_BIT_PATTERN_31[2 * i];
        pattern[i].y = ORBFE_BIT_PATTERN_31[2 * i + 1];
    }
    umax.resize(kHalfPatch + 1);
    int v, v0, vmax = cvFloorD(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    int vmin = cvCeilD(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (v = 0; v <= vmax; ++v) umax[v] = cvRoundD(std::sqrt(hp2 - v * v));
    for (v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
    levels.resize(nlevels);
}

void orc_extractor::computePyramid(const uint8_t* img, int w, int h, int stride) {    // :1331-1356
    for (int level = 0; level < nlevels; ++level) {
        float s = invScale[level];
        Level& L = levels[level];
        L.w = cvRoundF((float)w * s);
        L.h = cvRoundF((float)h * s);
        L.pw = L.w + 2 * kEdge;
        L.padded.assign((size_t)L.pw * (L.h + 2 * kEdge), 0);
        L.blurred.clear();
        L.candidates.clear();
        L.keypoints.clear();
        if (level != 0) {
            const Level& P = levels[level - 1];
            std::vector<uint8_t> tmp((size_t)L.w * L.h);
            resizeLinearU8(P.roi(), P.w, P.h, P.pw, tmp.data(), L.w, L.h, L.w);
            borderReflect101(tmp.data(), L.w, L.h, L.w, L.padded.data(), L.pw, kEdge);
        } else {
            borderReflect101(img, w, h, stride, L.padded.data(), L.pw, kEdge);
        }
    }
}

std::vector<KeyPt> orc_extractor::distributeOctTree(const std::vector<KeyPt>& in, int minX, int maxX, int minY,
                                                    int maxY, int N) {    // :557-781
    // the reference divides by nIni == 0 for images narrower than half their height (undefined behaviour); clamp.
    const int nIni = std::max(1, (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY)));
    const float hX = static_cast<float>(maxX - minX) / nIni;
    std::list<Node> nodes;
    std::vector<Node*> ini(nIni);
    for (int i = 0; i < nIni; i++) {
        Node ni;
        ni.UL = {(int)(hX * static_cast<float>(i)), 0};
        ni.UR = {(int)(hX * static_cast<float>(i + 1)), 0};
        ni.BL = {ni.UL.x, maxY - minY};
        ni.BR = {ni.UR.x, maxY - minY};
        nodes.push_back(ni);
        ini[i] = &nodes.back();
    }
    for (const KeyPt& kp : in) ini[(int)(kp.x / hX)]->keys.push_back(kp);

    auto lit = nodes.begin();
    while (lit != nodes.end()) {
        if (lit->keys.size() == 1) { lit->noMore = true; ++lit; }
        else if (lit->keys.empty()) lit = nodes.erase(lit);
        else ++lit;
    }

    bool finish = false;
    std::vector<SizeNode> sizeAndNode;
    sizeAndNode.reserve(nodes.size() * 4);

    // children go to the FRONT of the list in the order n1..n4; those with more than one key are recorded
    auto addChild = [&](Node& c, int* nToExpand) {
        if (c.keys.empty()) return;
        nodes.push_front(c);
        if (c.keys.size() > 1) {
            if (nToExpand) ++*nToExpand;
            sizeAndNode.push_back(std::make_pair((int)c.keys.size(), &nodes.front()));
            nodes.front().lit = nodes.begin();
        }
    };

    while (!finish) {
        int prevSize = (int)nodes.size();
        lit = nodes.begin();
        int nToExpand = 0;
        sizeAndNode.clear();
        while (lit != nodes.end()) {
            if (lit->noMore) { ++lit; continue; }
            Node n1, n2, n3, n4;
            lit->divide(n1, n2, n3, n4);
            addChild(n1, &nToExpand);
            addChild(n2, &nToExpand);
            addChild(n3, &nToExpand);
            addChild(n4, &nToExpand);
            lit = nodes.erase(lit);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) {
            finish = true;
        } else if (((int)nodes.size() + nToExpand * 3) > N) {
            while (!finish) {
                prevSize = (int)nodes.size();
                std::vector<SizeNode> prev = sizeAndNode;
                sizeAndNode.clear();
                std::sort(prev.begin(), prev.end(), compareNodes);    // libstdc++ introsort; ties are unordered
                for (int j = (int)prev.size() - 1; j >= 0; j--) {
                    Node n1, n2, n3, n4;
                    prev[j].second->divide(n1, n2, n3, n4);
                    addChild(n1, nullptr);
                    addChild(n2, nullptr);
                    addChild(n3, nullptr);
                    addChild(n4, nullptr);
                    nodes.erase(prev[j].second->lit);
                    if ((int)nodes.size() >= N) break;
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prevSize) finish = true;
            }
        }
    }

    std::vector<KeyPt> result;
    result.reserve(nfeatures);
    for (Node& n : nodes) {
        const KeyPt* best = &n.keys[0];
        float maxResponse = best->response;
        for (size_t k = 1; k < n.keys.size(); k++)
            if (n.keys[k].response > maxResponse) { best = &n.keys[k]; maxResponse = n.keys[k].response; }
        result.push_back(*best);
    }
    return result;
}

void orc_extractor::computeKeyPointsOctTree() {    // :783-898
    const float W = 35;
    Timer tm;
    std::vector<XYR> cell;
    for (int level = 0; level < nlevels; ++level) {
        Level& L = levels[level];
        const int minBorderX = kEdge - 3, minBorderY = minBorderX;
        const int maxBorderX = L.w - kEdge + 3, maxBorderY = L.h - kEdge + 3;
        std::vector<KeyPt> toDistribute;
        toDistribute.reserve(nfeatures * 10);
        const float width = (maxBorderX - minBorderX), height = (maxBorderY - minBorderY);
        const int nCols = width / W, nRows = height / W;
        const int wCell = nCols > 0 ? (int)std::ceil(width / nCols) : 0;
        const int hCell = nRows > 0 ? (int)std::ceil(height / nRows) : 0;
        for (int i = 0; i < nRows; i++) {
            const float iniY = minBorderY + i * hCell;
            float maxY = iniY + hCell + 6;
            if (iniY >= maxBorderY - 3) continue;
            if (maxY > maxBorderY) maxY = maxBorderY;
            for (int j = 0; j < nCols; j++) {
                const float iniX = minBorderX + j * wCell;
                float maxX = iniX + wCell + 6;
                if (iniX >= maxBorderX - 6) continue;
                if (maxX > maxBorderX) maxX = maxBorderX;
                const int y0 = (int)iniY, y1 = (int)maxY, x0 = (int)iniX, x1 = (int)maxX;
                const uint8_t* sub = L.roi() + (size_t)y0 * L.pw + x0;
                fast9Nms(sub, x1 - x0, y1 - y0, L.pw, iniTh, cell);
                if (cell.empty()) fast9Nms(sub, x1 - x0, y1 - y0, L.pw, minTh, cell);
                for (const XYR& c : cell) {
                    KeyPt kp;
                    kp.x = (float)c.x; kp.y = (float)c.y; kp.size = 7.f; kp.angle = -1; kp.response = (float)c.r;
                    kp.x += j * wCell;
                    kp.y += i * hCell;
                    toDistribute.push_back(kp);
                    L.candidates.push_back({(int)kp.x, (int)kp.y, c.r});
                }
            }
        }
        stage[1] += tm.lap();
        g_sortTies = 0;
        L.keypoints = (nCols > 0 && nRows > 0 && maxBorderX > minBorderX && maxBorderY > minBorderY)
                          ? distributeOctTree(toDistribute, minBorderX, maxBorderX, minBorderY, maxBorderY,
                                              featuresPerLevel[level])
                          : std::vector<KeyPt>();
        sortTies += g_sortTies;
        const int scaledPatchSize = kPatch * scale[level];
        for (KeyPt& kp : L.keypoints) {
            kp.x += minBorderX;
            kp.y += minBorderY;
            kp.octave = level;
            kp.size = scaledPatchSize;
        }
        stage[2] += tm.lap();
    }
    // computeOrientation -> IC_Angle (:78-105), on the unblurred level
    for (int level = 0; level < nlevels; ++level) {
        Level& L = levels[level];
        const int step = L.pw;
        for (KeyPt& kp : L.keypoints) {
            int m_01 = 0, m_10 = 0;
            const uint8_t* center = L.roi() + (size_t)cvRoundF(kp.y) * step + cvRoundF(kp.x);
            for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m_10 += u * center[u];
            for (int v = 1; v <= kHalfPatch; ++v) {
                int v_sum = 0, d = umax[v];
                for (int u = -d; u <= d; ++u) {
                    int val_plus = center[u + v * step], val_minus = center[u - v * step];
                    v_sum += (val_plus - val_minus);
                    m_10 += u * (val_plus + val_minus);
                }
                m_01 += v * v_sum;
            }
            kp.angle = fastAtan2f((float)m_01, (float)m_10);
        }
    }
    stage[3] += tm.lap();
}

int orc_extractor::run(const uint8_t* img, int w, int h, int stride, int lap0, int lap1, orc_keypoint* kps,
                       uint8_t* desc, int cap, int* count) {    // :1088-1191
    if (!img || w <= 0 || h <= 0) { if (count) *count = 0; return -1; }
    sortTies = 0;
    Timer tm;
    computePyramid(img, w, h, stride);
    stage[0] += tm.lap();
    computeKeyPointsOctTree();
    tm.lap();

    int nkeypoints = 0;
    for (int level = 0; level < nlevels; ++level) nkeypoints += (int)levels[level].keypoints.size();
    if (count) *count = nkeypoints;
    std::vector<KeyPt> outK(nkeypoints);
    std::vector<uint8_t> outD((size_t)nkeypoints * 32);
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    int monoIndex = 0, stereoIndex = nkeypoints - 1;
    for (int level = 0; level < nlevels; ++level) {
        Level& L = levels[level];
        if (L.keypoints.empty()) continue;
        // GaussianBlur(mvImagePyramid[level].clone(), 7x7, 2, 2, BORDER_REFLECT_101): the clone drops the frame
        L.blurred.resize((size_t)L.w * L.h);
        gaussian7x7U8(L.roi(), L.w, L.h, L.pw, L.blurred.data(), L.w);
        stage[4] += tm.lap();
        std::vector<uint8_t> d((size_t)L.keypoints.size() * 32);
        const int step = L.w;
        for (size_t i = 0; i < L.keypoints.size(); i++) {    // computeOrbDescriptor :109-148
            const KeyPt& kp = L.keypoints[i];
            float angle = (float)kp.angle * factorPI;
            // the reference calls cos/sin on a float (cosf/sinf); evaluated here in double and rounded to float
            // so that host libm and the CUDA math library agree (DESIGN.md "Floating point").
            float a = (float)std::cos((double)angle), b = (float)std::sin((double)angle);
            const uint8_t* center = &L.blurred[(size_t)cvRoundF(kp.y) * step + cvRoundF(kp.x)];
            const Pt2i* p = pattern;
            uint8_t* o = &d[i * 32];
            for (int k = 0; k < 32; ++k, p += 16) {
                int val = 0;
                for (int bit = 0; bit < 8; bit++) {
                    const Pt2i& p0 = p[2 * bit];
                    const Pt2i& p1 = p[2 * bit + 1];
                    int t0 = center[cvRoundF(p0.x * b + p0.y * a) * step + cvRoundF(p0.x * a - p0.y * b)];
                    int t1 = center[cvRoundF(p1.x * b + p1.y * a) * step + cvRoundF(p1.x * a - p1.y * b)];
                    val |= (t0 < t1) << bit;
                }
                o[k] = (uint8_t)val;
            }
        }
        stage[5] += tm.lap();
        float s = scale[level];
        int i = 0;
        for (KeyPt kp : L.keypoints) {
            if (level != 0) { kp.x *= s; kp.y *= s; }
            int dst;
            if (kp.x >= lap0 && kp.x <= lap1) dst = stereoIndex--; else dst = monoIndex++;
            outK[dst] = kp;
            std::memcpy(&outD[(size_t)dst * 32], &d[(size_t)i * 32], 32);
            i++;
        }
        stage[6] += tm.lap();
    }
    int ncopy = std::min(nkeypoints, cap);
    for (int i = 0; i < ncopy; i++) {
        const KeyPt& k = outK[i];
        if (kps) kps[i] = {k.x, k.y, k.size, k.angle, k.response, k.octave, k.class_id};
    }
    if (desc && ncopy > 0) std::memcpy(desc, outD.data(), (size_t)ncopy * 32);
    return monoIndex;
}

// ---- C interface -----------------------------------------------------------------------------------------------
extern "C" {

void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
    resizeLinearU8(src, sw, sh, sstride, dst, dw, dh, dstride);
}
void orc_border_reflect101_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int border) {
    borderReflect101(src, w, h, sstride, dst, dstride, border);
}
void orc_gaussian7x7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
    gaussian7x7U8(src, w, h, sstride, dst, dstride);
}
int orc_fast9_nms(const uint8_t* img, int w, int h, int stride, int threshold, int* out_xyr, int cap) {
    std::vector<XYR> v;
    fast9Nms(img, w, h, stride, threshold, v);
    for (int i = 0; i < (int)v.size() && i < cap; i++) {
        out_xyr[3 * i] = v[i].x; out_xyr[3 * i + 1] = v[i].y; out_xyr[3 * i + 2] = v[i].r;
    }
    return (int)v.size();
}
float orc_fast_atan2(float y, float x) { return fastAtan2f(y, x); }

void orc_knn2_hamming(const uint8_t* q, int64_t nq, const uint8_t* t, int64_t nt, int32_t* idx, int32_t* dist,
                      int threads) {
    if (threads <= 1 || nq < 2 * threads) { knn2Range(q, 0, nq, t, nt, idx, dist); return; }
    std::vector<std::thread> pool;
    std::atomic<int64_t> next(0);
    const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(256, nq / (threads * 4)));
    for (int k = 0; k < threads; k++)
        pool.emplace_back([&] {
            for (;;) {
                int64_t a = next.fetch_add(chunk);
                if (a >= nq) break;
                knn2Range(q, a, std::min(nq, a + chunk), t, nt, idx, dist);
            }
        });
    for (auto& th : pool) th.join();
}

orc_extractor* orc_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th) {
    if (nlevels < 1 || nlevels > 64 || nfeatures < 0) return nullptr;
    return new orc_extractor(nfeatures, scale_factor, nlevels, ini_th, min_th);
}
void orc_extractor_destroy(orc_extractor* e) { delete e; }
void orc_extractor_tables(const orc_extractor* e, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                          int* features_per_level, int* umax) {
    for (int i = 0; i < e->nlevels; i++) {
        if (scale) scale[i] = e->scale[i];
        if (inv_scale) inv_scale[i] = e->invScale[i];
        if (sigma2) sigma2[i] = e->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = e->invSigma2[i];
        if (features_per_level) features_per_level[i] = e->featuresPerLevel[i];
    }
    if (umax) for (int i = 0; i < 16; i++) umax[i] = e->umax[i];
}
int orc_extract(orc_extractor* e, const uint8_t* img, int w, int h, int stride, int lap0, int lap1,
                orc_keypoint* kps, uint8_t* desc, int cap, int* count) {
    return e->run(img, w, h, stride, lap0, lap1, kps, desc, cap, count);
}
void orc_level_size(const orc_extractor* e, int level, int* w, int* h) {
    *w = e->levels[level].w; *h = e->levels[level].h;
}
void orc_level_pixels(const orc_extractor* e, int level, uint8_t* dst, int dstride) {
    const auto& L = e->levels[level];
    for (int y = 0; y < L.h; y++) std::memcpy(dst + (size_t)y * dstride, L.roi() + (size_t)y * L.pw, L.w);
}
void orc_level_padded(const orc_extractor* e, int level, uint8_t* dst, int dstride) {
    const auto& L = e->levels[level];
    for (int y = 0; y < L.h + 2 * kEdge; y++) std::memcpy(dst + (size_t)y * dstride, &L.padded[(size_t)y * L.pw], L.pw);
}
void orc_level_blurred(const orc_extractor* e, int level, uint8_t* dst, int dstride) {
    const auto& L = e->levels[level];
    for (int y = 0; y < L.h; y++) {
        if (L.blurred.empty()) std::memset(dst + (size_t)y * dstride, 0, L.w);
        else std::memcpy(dst + (size_t)y * dstride, &L.blurred[(size_t)y * L.w], L.w);
    }
}
int orc_level_candidates(const orc_extractor* e, int level, int* out_xyr, int cap) {
    const auto& c = e->levels[level].candidates;
    for (int i = 0; i < (int)c.size() && i < cap; i++) {
        out_xyr[3 * i] = c[i].x; out_xyr[3 * i + 1] = c[i].y; out_xyr[3 * i + 2] = c[i].r;
    }
    return (int)c.size();
}
int orc_level_keypoints(const orc_extractor* e, int level, orc_keypoint* out, int cap) {
    const auto& k = e->levels[level].keypoints;
    for (int i = 0; i < (int)k.size() && i < cap; i++)
        out[i] = {k[i].x, k[i].y, k[i].size, k[i].angle, k[i].response, k[i].octave, k[i].class_id};
    return (int)k.size();
}
void orc_extractor_stage_seconds(const orc_extractor* e, double out[7]) {
    for (int i = 0; i < 7; i++) out[i] = e->stage[i];
}
int64_t orc_extractor_sort_ties(const orc_extractor* e) { return e->sortTies; }

int orc_distribute_octtree(orc_extractor* e, const orc_keypoint* keys, int n, int minX, int maxX, int minY, int maxY,
                           int N, orc_keypoint* out, int cap, int64_t* ties) {
    std::vector<KeyPt> in((size_t)n);
    for (int i = 0; i < n; ++i) {
        in[i].x = keys[i].x; in[i].y = keys[i].y; in[i].size = keys[i].size; in[i].angle = keys[i].angle;
        in[i].response = keys[i].response; in[i].octave = keys[i].octave; in[i].class_id = keys[i].class_id;
    }
    g_sortTies = 0;
    std::vector<KeyPt> r = e->distributeOctTree(in, minX, maxX, minY, maxY, N);
    if (ties) *ties = g_sortTies;
    const int m = (int)r.size();
    for (int i = 0; i < m && i < cap; ++i) {
        out[i].x = r[i].x; out[i].y = r[i].y; out[i].size = r[i].size; out[i].angle = r[i].angle;
        out[i].response = r[i].response; out[i].octave = r[i].octave; out[i].class_id = r[i].class_id;
    }
    return m;
}

int orc_extract_batch(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th, const uint8_t* imgs,
                      int n, int w, int h, int stride, int lap0, int lap1, orc_keypoint* kps, uint8_t* desc, int cap,
                      int* counts, int* monos, int threads) {
    if (threads < 1) threads = 1;
    std::atomic<int> next(0);
    auto work = [&] {
        orc_extractor e(nfeatures, scale_factor, nlevels, ini_th, min_th);
        for (;;) {
            int i = next.fetch_add(1);
            if (i >= n) break;
            int cnt = 0;
            int mono = e.run(imgs + (size_t)i * h * stride, w, h, stride, lap0, lap1,
                             kps ? kps + (size_t)i * cap : nullptr, desc ? desc + (size_t)i * cap * 32 : nullptr, cap,
                             &cnt);
            if (counts) counts[i] = cnt;
            if (monos) monos[i] = mono;
        }
    };
    if (threads == 1) { work(); return 0; }
    std::vector<std::thread> pool;
    for (int k = 0; k < threads; k++) pool.emplace_back(work);
    for (auto& th : pool) th.join();
    return 0;
}

// ---- Frame::ComputeStereoMatches (Frame.cc:827-997) --------------------------------------------------------------
// Restated statement by statement: row table (:836-854), Hamming best-of-band with octave +-1 and the disparity
// window (:868-910), 11x11 SAD over +-5 px on the keypoint's pyramid level (:914-949, cv::norm(IL, IR, NORM_L1) on
// CV_8U is an exact integer sum), parabola fit (:954-962), depth (:964-979), median cut (:983-996).
// ORBmatcher::TH_HIGH = 100, TH_LOW = 50 (ORBmatcher.cc:35-36).  Where the reference would index outside its row table
// or take a window outside mvImagePyramid[level] (undefined behaviour / cv::Exception) the keypoint gets no match and
// the event is counted in the return value.
int orc_stereo_matches(const orc_extractor* eL, const orc_extractor* eR, const orc_keypoint* kl, const uint8_t* dl,
                       int nl, const orc_keypoint* kr, const uint8_t* dr, int nr, float mb, float mbf, float* uRight,
                       float* depth, int32_t* sad) {
    const int TH_HIGH = 100, TH_LOW = 50;
    int faults = 0;
    for (int i = 0; i < nl; i++) {
        uRight[i] = -1.0f;
        depth[i] = -1.0f;
        if (sad) sad[i] = -1;
    }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = eL->levels[0].h;
    std::vector<std::vector<size_t>> vRowIndices(nRows);
    for (int iR = 0; iR < nr; iR++) {
        const float kpY = kr[iR].y;
        const float r = 2.0f * eL->scale[kr[iR].octave];
        const int maxr = (int)std::ceil(kpY + r);
        const int minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++) {
            if (yi < 0 || yi >= nRows) { faults++; continue; }
            vRowIndices[yi].push_back(iR);
        }
    }
    const float minZ = mb;
    const float minD = 0;
    const float maxD = mbf / minZ;
    std::vector<std::pair<int, int>> vDistIdx;
    vDistIdx.reserve(nl);
    for (int iL = 0; iL < nl; iL++) {
        const orc_keypoint& kpL = kl[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y;
        const float uL = kpL.x;
        const size_t row = (size_t)vL;
        if (row >= (size_t)nRows) { faults++; continue; }
        const std::vector<size_t>& vCandidates = vRowIndices[row];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD;
        const float maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        uint64_t a[4];
        std::memcpy(a, dl + (size_t)iL * 32, 32);
        for (size_t iC = 0; iC < vCandidates.size(); iC++) {
            const size_t iR = vCandidates[iC];
            const orc_keypoint& kpR = kr[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                uint64_t b[4];
                std::memcpy(b, dr + iR * 32, 32);
                const int dist = hamming256(a, b);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdxR = iR;
                }
            }
        }
        if (bestDist < thOrbDist) {
            const float uR0 = kr[bestIdxR].x;
            const float scaleFactor = eL->invScale[kpL.octave];
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);
            const int w = 5;
            const auto& LL = eL->levels[kpL.octave];
            const auto& LR = eR->levels[kpL.octave];
            // IL = level.rowRange(scaledvL-w, scaledvL+w+1).colRange(scaleduL-w, scaleduL+w+1)
            const int y0 = (int)(scaledvL - w), y1 = (int)(scaledvL + w + 1);
            const int xl0 = (int)(scaleduL - w), xl1 = (int)(scaleduL + w + 1);
            if (y0 < 0 || y1 > LL.h || xl0 < 0 || xl1 > LL.w) { faults++; continue; }
            int bestSad = INT_MAX;
            int bestincR = 0;
            const int L = 5;
            float vDists[2 * 5 + 1];
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= LR.w) continue;
            if ((int)(scaleduR0 - L - w) < 0 || y1 > LR.h) { faults++; continue; }
            for (int incR = -L; incR <= +L; incR++) {
                const int xr0 = (int)(scaleduR0 + incR - w);
                int64_t s = 0;
                for (int y = y0; y < y1; y++) {
                    const uint8_t* pl = LL.roi() + (size_t)y * LL.pw + xl0;
                    const uint8_t* pr = LR.roi() + (size_t)y * LR.pw + xr0;
                    for (int x = 0; x < 2 * w + 1; x++) s += std::abs((int)pl[x] - (int)pr[x]);
                }
                const float dist = (float)(double)s;
                if (dist < bestSad) {
                    bestSad = (int)dist;
                    bestincR = incR;
                }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;
            const float dist1 = vDists[L + bestincR - 1];
            const float dist2 = vDists[L + bestincR];
            const float dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = eL->scale[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) {
                    disparity = 0.01;
                    bestuR = uL - 0.01;
                }
                depth[iL] = mbf / disparity;
                uRight[iL] = bestuR;
                if (sad) sad[iL] = bestSad;
                vDistIdx.push_back(std::pair<int, int>(bestSad, iL));
            }
        }
    }
    if (vDistIdx.empty()) return faults;    // the reference indexes an empty vector here
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        uRight[vDistIdx[i].second] = -1;
        depth[vDistIdx[i].second] = -1;
    }
    return faults;
}

// ---- Frame grid + ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th, bFarPoints, thFarPoints) ------------
// Frame::AssignFeaturesToGrid (Frame.cc:405-436) with Frame::PosInGrid (:741-751), Frame::GetFeaturesInArea
// (:672-736), ORBmatcher::RadiusByViewingCos (ORBmatcher.cc:216-222), ORBmatcher::DescriptorDistance (:2107-2123) and
// the map-point loop of ORBmatcher.cc:44-143 for frames with Nleft == -1 (monocular / rectified stereo: the keypoints
// are mvKeysUn and the right-image branch :145-211 is not taken).  MapPoint state enters through orc_track_point:
// flags bit 0 = mbTrackInView, bit 1 = isBad(), bit 2 = Observations() > 0 (an assigned point then blocks its feature
// for the later points of the same call, :84-86).
namespace {
const int GRID_ROWS = 48, GRID_COLS = 64;    // FRAME_GRID_ROWS / FRAME_GRID_COLS, Frame.h:46-47
}

int orc_search_by_projection(const orc_keypoint* keysUn, const uint8_t* desc, const float* uRight,
                             const uint8_t* occupied, int n, float mnMinX, float mnMinY, float mnMaxX, float mnMaxY,
                             const float* mvScaleFactors, const orc_track_point* pts, int npts, float th, int bFarPoints,
                             float thFarPoints, float mfNNratio, int32_t* match, int32_t* owner) {
    const int TH_HIGH = 100;
    const float mfGridElementWidthInv = static_cast<float>(GRID_COLS) / (mnMaxX - mnMinX);      // Frame.cc:179-180
    const float mfGridElementHeightInv = static_cast<float>(GRID_ROWS) / (mnMaxY - mnMinY);
    static thread_local std::vector<size_t> mGrid[GRID_COLS][GRID_ROWS];
    for (int i = 0; i < GRID_COLS; i++)
        for (int j = 0; j < GRID_ROWS; j++) mGrid[i][j].clear();
    for (int i = 0; i < n; i++) {    // AssignFeaturesToGrid
        const orc_keypoint& kp = keysUn[i];
        const int posX = (int)std::round((kp.x - mnMinX) * mfGridElementWidthInv);
        const int posY = (int)std::round((kp.y - mnMinY) * mfGridElementHeightInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        mGrid[posX][posY].push_back(i);
    }
    // F.mvpMapPoints[idx] != NULL && Observations() > 0, and which point of this call owns the feature
    std::vector<uint8_t> blocked(n, 0);
    for (int i = 0; i < n; i++) {
        blocked[i] = occupied ? occupied[i] : 0;
        if (owner) owner[i] = -1;
    }
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<size_t> vIndices;
    for (int iMP = 0; iMP < npts; iMP++) {
        const orc_track_point& mp = pts[iMP];
        match[iMP] = -1;
        if (!(mp.flags & 1)) continue;                                  // !mbTrackInView (&& !mbTrackInViewR)
        if (bFarPoints && mp.track_depth > thFarPoints) continue;
        if (mp.flags & 2) continue;                                     // isBad()
        const int nPredictedLevel = mp.level;
        float r = mp.view_cos > 0.998 ? 2.5 : 4.0;                      // RadiusByViewingCos
        if (bFactor) r *= th;
        // GetFeaturesInArea(x, y, r * scale, level - 1, level)
        const float x = mp.proj_x, y = mp.proj_y;
        const float rr = r * mvScaleFactors[nPredictedLevel];
        const int minLevel = nPredictedLevel - 1, maxLevel = nPredictedLevel;
        vIndices.clear();
        {
            const float factorX = rr, factorY = rr;
            const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * mfGridElementWidthInv));
            const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
            const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * mfGridElementHeightInv));
            const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
            if (!(nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0)) {
                const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
                for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                    for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                        const std::vector<size_t>& vCell = mGrid[ix][iy];
                        for (size_t j = 0; j < vCell.size(); j++) {
                            const orc_keypoint& kpUn = keysUn[vCell[j]];
                            if (bCheckLevels) {
                                if (kpUn.octave < minLevel) continue;
                                if (maxLevel >= 0)
                                    if (kpUn.octave > maxLevel) continue;
                            }
                            const float distx = kpUn.x - x;
                            const float disty = kpUn.y - y;
                            if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) vIndices.push_back(vCell[j]);
                        }
                    }
            }
        }
        if (vIndices.empty()) continue;
        uint64_t a[4];
        std::memcpy(a, mp.desc, 32);
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (size_t k = 0; k < vIndices.size(); k++) {
            const size_t idx = vIndices[k];
            if (blocked[idx]) continue;
            if (uRight && uRight[idx] > 0) {
                const float er = std::fabs(mp.proj_xr - uRight[idx]);
                if (er > r * mvScaleFactors[nPredictedLevel]) continue;
            }
            uint64_t b[4];
            std::memcpy(b, desc + idx * 32, 32);
            const int dist = hamming256(a, b);
            if (dist < bestDist) {
                bestDist2 = bestDist;
                bestDist = dist;
                bestLevel2 = bestLevel;
                bestLevel = keysUn[idx].octave;
                bestIdx = (int)idx;
            } else if (dist < bestDist2) {
                bestLevel2 = keysUn[idx].octave;
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
            if (bestLevel != bestLevel2 || bestDist <= mfNNratio * bestDist2) {
                match[iMP] = bestIdx;                       // F.mvpMapPoints[bestIdx] = pMP
                if (owner) owner[bestIdx] = iMP;
                blocked[bestIdx] = (mp.flags & 4) ? 1 : 0;    // later points skip it iff pMP->Observations() > 0
                nmatches++;
            }
        }
    }
    return nmatches;
}

// ---- Frame::UndistortKeyPoints (Frame.cc:763-796) ------------------------------------------------------------------
// cv::undistortPoints(mat, mat, K, mDistCoef, cv::Mat(), mK): OpenCV 4.x cvUndistortPointsInternal with the default
// criteria TermCriteria(MAX_ITER, 5, 0.01): normalise, 5 fixed-point iterations of the radial / tangential model in
// double, re-project with P = mK, store as float.  k = (k1, k2, p1, p2[, k3]); mDistCoef[0] == 0 copies the keypoints
// (:765-769).  Pinned against cv2.undistortPoints in tests/test_oracle_vs_cv2.py.
void orc_undistort_keypoints(const orc_keypoint* keys, int n, float fx, float fy, float cx, float cy, const float* dist,
                             int ndist, orc_keypoint* keysUn) {
    if (ndist <= 0 || dist[0] == 0.0) {
        for (int i = 0; i < n; i++) keysUn[i] = keys[i];
        return;
    }
    double k[14] = {0};
    for (int i = 0; i < ndist && i < 14; i++) k[i] = dist[i];
    const double dfx = fx, dfy = fy, dcx = cx, dcy = cy;
    const double ifx = 1. / dfx, ify = 1. / dfy;
    for (int i = 0; i < n; i++) {
        double x = keys[i].x, y = keys[i].y;
        const double u = x, v = y;
        x = (x - dcx) * ifx;
        y = (y - dcy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) {
                x = (u - dcx) * ifx;
                y = (v - dcy) * ify;
                break;
            }
            const double deltaX = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double deltaY = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - deltaX) * icdist;
            y = (y0 - deltaY) * icdist;
        }
        const double xx = dfx * x + 0.0 * y + dcx;
        const double yy = 0.0 * x + dfy * y + dcy;
        const double ww = 1. / (0.0 * x + 0.0 * y + 1.0);
        keysUn[i] = keys[i];
        keysUn[i].x = (float)(xx * ww);
        keysUn[i].y = (float)(yy * ww);
    }
}

// ---- the same loop for stereo-camera frames (Nleft != -1: the Lynx fisheye rig), ORBmatcher.cc:44-214 -------------------
// keys / desc / occupied hold the N = Nleft + Nright features, left first (mvKeys, then mvKeysRight; mDescriptors rows;
// F.mvpMapPoints slots).  Left search on mGrid (:59-143, no mvuRight gate since Nleft != -1), its partner assignment
// through mvLeftToRightMatch (:130-134); a ratio rejection there skips the right search (:126-127 `continue`); right
// search on mGridRight with mTrackProjXR/YR, mnTrackScaleLevelR, mTrackViewCosR — the radius is NOT scaled by th there
// (:148-151) — and the partner assignment through mvRightToLeftMatch (:198-202).
int orc_search_by_projection_stereo(const orc_keypoint* keys, const uint8_t* desc, const uint8_t* occupied, int n,
                                    int nleft, const int32_t* leftToRight, const int32_t* rightToLeft, float mnMinX,
                                    float mnMinY, float mnMaxX, float mnMaxY, const float* mvScaleFactors,
                                    const orc_track_point_stereo* pts, int npts, float th, int bFarPoints,
                                    float thFarPoints, float mfNNratio, int32_t* matchL, int32_t* matchR, int32_t* owner) {
    const int TH_HIGH = 100;
    const float gwInv = static_cast<float>(GRID_COLS) / (mnMaxX - mnMinX);
    const float ghInv = static_cast<float>(GRID_ROWS) / (mnMaxY - mnMinY);
    static thread_local std::vector<size_t> mGrid[GRID_COLS][GRID_ROWS], mGridRight[GRID_COLS][GRID_ROWS];
    for (int i = 0; i < GRID_COLS; i++)
        for (int j = 0; j < GRID_ROWS; j++) {
            mGrid[i][j].clear();
            mGridRight[i][j].clear();
        }
    for (int i = 0; i < n; i++) {    // AssignFeaturesToGrid, Nleft != -1 (Frame.cc:420-434)
        const orc_keypoint& kp = keys[i];
        const int posX = (int)std::round((kp.x - mnMinX) * gwInv);
        const int posY = (int)std::round((kp.y - mnMinY) * ghInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        if (i < nleft) mGrid[posX][posY].push_back(i);
        else mGridRight[posX][posY].push_back(i - nleft);
    }
    std::vector<uint8_t> blocked(n, 0);
    for (int i = 0; i < n; i++) {
        blocked[i] = occupied ? occupied[i] : 0;
        if (owner) owner[i] = -1;
    }
    auto assign = [&](int slot, int iMP, bool obs) {    // F.mvpMapPoints[slot] = pMP
        if (owner) owner[slot] = iMP;
        blocked[slot] = obs ? 1 : 0;
    };
    auto area = [&](float x, float y, float r, int minLevel, int maxLevel, bool bRight, std::vector<size_t>& out) {
        out.clear();
        const float factorX = r, factorY = r;
        const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * gwInv));
        if (nMinCellX >= GRID_COLS) return;
        const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * gwInv));
        if (nMaxCellX < 0) return;
        const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * ghInv));
        if (nMinCellY >= GRID_ROWS) return;
        const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * ghInv));
        if (nMaxCellY < 0) return;
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<size_t>& vCell = (!bRight) ? mGrid[ix][iy] : mGridRight[ix][iy];
                for (size_t j = 0; j < vCell.size(); j++) {
                    const orc_keypoint& kpUn = (!bRight) ? keys[vCell[j]] : keys[nleft + vCell[j]];
                    if (bCheckLevels) {
                        if (kpUn.octave < minLevel) continue;
                        if (maxLevel >= 0)
                            if (kpUn.octave > maxLevel) continue;
                    }
                    const float distx = kpUn.x - x;
                    const float disty = kpUn.y - y;
                    if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) out.push_back(vCell[j]);
                }
            }
    };
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<size_t> vIndices;
    for (int iMP = 0; iMP < npts; iMP++) {
        const orc_track_point_stereo& mp = pts[iMP];
        matchL[iMP] = -1;
        matchR[iMP] = -1;
        const bool inView = mp.flags & 1, inViewR = mp.flags & 8, obs = mp.flags & 4;
        if (!inView && !inViewR) continue;
        if (bFarPoints && mp.track_depth > thFarPoints) continue;
        if (mp.flags & 2) continue;
        uint64_t a[4];
        std::memcpy(a, mp.desc, 32);
        bool skipRight = false;
        if (inView) {
            const int nPredictedLevel = mp.level;
            float r = mp.view_cos > 0.998 ? 2.5 : 4.0;
            if (bFactor) r *= th;
            area(mp.proj_x, mp.proj_y, r * mvScaleFactors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel, false, vIndices);
            if (!vIndices.empty()) {
                int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
                for (size_t k = 0; k < vIndices.size(); k++) {
                    const size_t idx = vIndices[k];
                    if (blocked[idx]) continue;
                    uint64_t b[4];
                    std::memcpy(b, desc + idx * 32, 32);
                    const int dist = hamming256(a, b);
                    if (dist < bestDist) {
                        bestDist2 = bestDist;
                        bestDist = dist;
                        bestLevel2 = bestLevel;
                        bestLevel = keys[idx].octave;
                        bestIdx = (int)idx;
                    } else if (dist < bestDist2) {
                        bestLevel2 = keys[idx].octave;
                        bestDist2 = dist;
                    }
                }
                if (bestDist <= TH_HIGH) {
                    if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) skipRight = true;    // `continue`
                    else if (bestLevel != bestLevel2 || bestDist <= mfNNratio * bestDist2) {
                        matchL[iMP] = bestIdx;
                        assign(bestIdx, iMP, obs);
                        if (leftToRight[bestIdx] != -1) {
                            assign(leftToRight[bestIdx] + nleft, iMP, obs);
                            nmatches++;
                        }
                        nmatches++;
                    }
                }
            }
        }
        if (skipRight) continue;
        if (inViewR) {
            const int nPredictedLevel = mp.level_r;
            if (nPredictedLevel != -1) {
                float r = mp.view_cos_r > 0.998 ? 2.5 : 4.0;
                area(mp.proj_xr, mp.proj_yr, r * mvScaleFactors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel, true, vIndices);
                if (vIndices.empty()) continue;
                int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
                for (size_t k = 0; k < vIndices.size(); k++) {
                    const size_t idx = vIndices[k];
                    if (blocked[idx + nleft]) continue;
                    uint64_t b[4];
                    std::memcpy(b, desc + (idx + nleft) * 32, 32);
                    const int dist = hamming256(a, b);
                    if (dist < bestDist) {
                        bestDist2 = bestDist;
                        bestDist = dist;
                        bestLevel2 = bestLevel;
                        bestLevel = keys[nleft + idx].octave;
                        bestIdx = (int)idx;
                    } else if (dist < bestDist2) {
                        bestLevel2 = keys[nleft + idx].octave;
                        bestDist2 = dist;
                    }
                }
                if (bestDist <= TH_HIGH) {
                    if (bestLevel == bestLevel2 && bestDist > mfNNratio * bestDist2) continue;
                    if (rightToLeft[bestIdx] != -1) {
                        assign(rightToLeft[bestIdx], iMP, obs);
                        nmatches++;
                    }
                    matchR[iMP] = bestIdx;
                    assign(bestIdx + nleft, iMP, obs);
                    nmatches++;
                }
            }
        }
    }
    return nmatches;
}

// ---- ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono) (ORBmatcher.cc:1683-1921) ----
// and ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, sAlreadyFound, th, ORBdist) (:1923-2059), with
// ComputeThreeMaxima (:2061-2102).  The pose algebra (Tcw * x3Dw, mpCamera->project, PredictScale, the distance gates of
// the KeyFrame form) stays with the caller: a point arrives as its projection(s), 1 / z, level, source-keypoint angle,
// descriptor and flags (bit 0: the point reaches the search, bit 2: pMP->Observations() > 0).
namespace {
struct FrameGrids {
    std::vector<size_t> mGrid[GRID_COLS][GRID_ROWS], mGridRight[GRID_COLS][GRID_ROWS];
};
void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {    // :2061-2102
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; i++) {
        const int s = (int)histo[i].size();
        if (s > max1) {
            max3 = max2; max2 = max1; max1 = s;
            ind3 = ind2; ind2 = ind1; ind1 = i;
        } else if (s > max2) {
            max3 = max2; max2 = s;
            ind3 = ind2; ind2 = i;
        } else if (s > max3) {
            max3 = s;
            ind3 = i;
        }
    }
    if (max2 < 0.1f * (float)max1) {
        ind2 = -1;
        ind3 = -1;
    } else if (max3 < 0.1f * (float)max1) {
        ind3 = -1;
    }
}
}    // namespace

int orc_search_by_projection_frame(const orc_keypoint* keys, const uint8_t* desc, const float* uRight,
                                   const uint8_t* occupied, int n, int nleft, float mnMinX, float mnMinY, float mnMaxX,
                                   float mnMaxY, const float* mvScaleFactors, float mbf, const orc_frame_point* pts,
                                   int npts, float th, int dir, int checkOrientation, int32_t* matchL, int32_t* matchR,
                                   int32_t* owner) {
    const int TH_HIGH = 100, HISTO_LENGTH = 30;
    const bool stereoCam = nleft != n;    // Nleft != -1: left and right features share the slot array, left first
    const float gwInv = static_cast<float>(GRID_COLS) / (mnMaxX - mnMinX);
    const float ghInv = static_cast<float>(GRID_ROWS) / (mnMaxY - mnMinY);
    static thread_local FrameGrids G;
    for (int i = 0; i < GRID_COLS; i++)
        for (int j = 0; j < GRID_ROWS; j++) {
            G.mGrid[i][j].clear();
            G.mGridRight[i][j].clear();
        }
    for (int i = 0; i < n; i++) {    // AssignFeaturesToGrid (Frame.cc:405-436)
        const orc_keypoint& kp = keys[i];
        const int posX = (int)std::round((kp.x - mnMinX) * gwInv);
        const int posY = (int)std::round((kp.y - mnMinY) * ghInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        if (i < nleft) G.mGrid[posX][posY].push_back(i);
        else G.mGridRight[posX][posY].push_back(i - nleft);
    }
    std::vector<uint8_t> blocked(n, 0);    // CurrentFrame.mvpMapPoints[i2] && ->Observations() > 0
    for (int i = 0; i < n; i++) {
        blocked[i] = occupied ? occupied[i] : 0;
        owner[i] = -1;
    }
    auto area = [&](float x, float y, float r, int minLevel, int maxLevel, bool bRight, std::vector<size_t>& out) {
        out.clear();    // Frame::GetFeaturesInArea (Frame.cc:672-736)
        const float factorX = r, factorY = r;
        const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * gwInv));
        if (nMinCellX >= GRID_COLS) return;
        const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * gwInv));
        if (nMaxCellX < 0) return;
        const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * ghInv));
        if (nMinCellY >= GRID_ROWS) return;
        const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * ghInv));
        if (nMaxCellY < 0) return;
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<size_t>& vCell = (!bRight) ? G.mGrid[ix][iy] : G.mGridRight[ix][iy];
                for (size_t j = 0; j < vCell.size(); j++) {
                    const orc_keypoint& kpUn = (!bRight) ? keys[vCell[j]] : keys[nleft + vCell[j]];
                    if (bCheckLevels) {
                        if (kpUn.octave < minLevel) continue;
                        if (maxLevel >= 0)
                            if (kpUn.octave > maxLevel) continue;
                    }
                    const float distx = kpUn.x - x;
                    const float disty = kpUn.y - y;
                    if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) out.push_back(vCell[j]);
                }
            }
    };
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    const bool bForward = dir == 1, bBackward = dir == 2;
    int nmatches = 0;
    std::vector<size_t> vIndices2;
    std::vector<std::pair<int, int>> entryPoint[HISTO_LENGTH];    // (point, 0 left / 1 right) of every histogram entry
    for (int i = 0; i < npts; i++) {
        matchL[i] = -1;
        matchR[i] = -1;
        const orc_frame_point& p = pts[i];
        if (!(p.flags & 1)) continue;    // pMP, !mvbOutlier[i], invzc >= 0
        const float invzc = p.invz;
        const float u = p.u, v = p.v;
        if (u < mnMinX || u > mnMaxX) continue;
        if (v < mnMinY || v > mnMaxY) continue;
        const int nLastOctave = p.level;
        float radius = th * mvScaleFactors[nLastOctave];
        if (bForward) area(u, v, radius, nLastOctave, -1, false, vIndices2);
        else if (bBackward) area(u, v, radius, 0, nLastOctave, false, vIndices2);
        else area(u, v, radius, nLastOctave - 1, nLastOctave + 1, false, vIndices2);
        if (vIndices2.empty()) continue;
        uint64_t a[4];
        std::memcpy(a, p.desc, 32);
        const bool obs = p.flags & 4;
        {
            int bestDist = 256, bestIdx2 = -1;
            for (size_t k = 0; k < vIndices2.size(); k++) {
                const size_t i2 = vIndices2[k];
                if (blocked[i2]) continue;
                if (!stereoCam && uRight && uRight[i2] > 0) {
                    const float ur = u - mbf * invzc;
                    const float er = std::fabs(ur - uRight[i2]);
                    if (er > radius) continue;
                }
                uint64_t b[4];
                std::memcpy(b, desc + i2 * 32, 32);
                const int dist = hamming256(a, b);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdx2 = (int)i2;
                }
            }
            if (bestDist <= TH_HIGH) {
                owner[bestIdx2] = i;    // CurrentFrame.mvpMapPoints[bestIdx2] = pMP
                blocked[bestIdx2] = obs ? 1 : 0;
                matchL[i] = bestIdx2;
                nmatches++;
                if (checkOrientation) {
                    float rot = p.angle - keys[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(bestIdx2);
                    entryPoint[bin].push_back({i, 0});
                }
            }
        }
        if (stereoCam) {
            const float ur = p.ur, vr = p.vr;
            radius = th * mvScaleFactors[nLastOctave];
            if (bForward) area(ur, vr, radius, nLastOctave, -1, true, vIndices2);
            else if (bBackward) area(ur, vr, radius, 0, nLastOctave, true, vIndices2);
            else area(ur, vr, radius, nLastOctave - 1, nLastOctave + 1, true, vIndices2);
            int bestDist = 256, bestIdx2 = -1;
            for (size_t k = 0; k < vIndices2.size(); k++) {
                const size_t i2 = vIndices2[k];
                if (blocked[i2 + nleft]) continue;
                uint64_t b[4];
                std::memcpy(b, desc + (i2 + nleft) * 32, 32);
                const int dist = hamming256(a, b);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdx2 = (int)i2;
                }
            }
            if (bestDist <= TH_HIGH) {
                owner[bestIdx2 + nleft] = i;
                blocked[bestIdx2 + nleft] = obs ? 1 : 0;
                matchR[i] = bestIdx2;
                nmatches++;
                if (checkOrientation) {
                    float rot = p.angle - keys[nleft + bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(bestIdx2 + nleft);
                    entryPoint[bin].push_back({i, 1});
                }
            }
        }
    }
    if (checkOrientation) {    // :1697-1715
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i != ind1 && i != ind2 && i != ind3) {
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    owner[rotHist[i][j]] = -1;    // CurrentFrame.mvpMapPoints[...] = NULL
                    nmatches--;
                    (entryPoint[i][j].second ? matchR : matchL)[entryPoint[i][j].first] = -1;
                }
            }
        }
    }
    return nmatches;
}

int orc_search_by_projection_keyframe(const orc_keypoint* keysUn, const uint8_t* desc, const uint8_t* occupied, int n,
                                      float mnMinX, float mnMinY, float mnMaxX, float mnMaxY, const float* mvScaleFactors,
                                      const orc_frame_point* pts, int npts, float th, int ORBdist, int checkOrientation,
                                      int32_t* match, int32_t* owner) {
    const int HISTO_LENGTH = 30;
    const float gwInv = static_cast<float>(GRID_COLS) / (mnMaxX - mnMinX);
    const float ghInv = static_cast<float>(GRID_ROWS) / (mnMaxY - mnMinY);
    static thread_local FrameGrids G;
    for (int i = 0; i < GRID_COLS; i++)
        for (int j = 0; j < GRID_ROWS; j++) G.mGrid[i][j].clear();
    for (int i = 0; i < n; i++) {
        const orc_keypoint& kp = keysUn[i];
        const int posX = (int)std::round((kp.x - mnMinX) * gwInv);
        const int posY = (int)std::round((kp.y - mnMinY) * ghInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        G.mGrid[posX][posY].push_back(i);
    }
    std::vector<uint8_t> taken(n, 0);    // CurrentFrame.mvpMapPoints[i2] != NULL
    for (int i = 0; i < n; i++) {
        taken[i] = occupied ? occupied[i] : 0;
        owner[i] = -1;
    }
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> entryPoint[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nmatches = 0;
    std::vector<size_t> vIndices2;
    for (int i = 0; i < npts; i++) {
        match[i] = -1;
        const orc_frame_point& p = pts[i];
        if (!(p.flags & 1)) continue;    // pMP && !isBad() && !sAlreadyFound.count(pMP) && dist3D inside [min, max]
        const float u = p.u, v = p.v;
        if (u < mnMinX || u > mnMaxX) continue;
        if (v < mnMinY || v > mnMaxY) continue;
        const int nPredictedLevel = p.level;
        const float radius = th * mvScaleFactors[nPredictedLevel];
        // GetFeaturesInArea(uv(0), uv(1), radius, nPredictedLevel - 1, nPredictedLevel + 1)
        vIndices2.clear();
        {
            const float factorX = radius, factorY = radius;
            const int minLevel = nPredictedLevel - 1, maxLevel = nPredictedLevel + 1;
            const int nMinCellX = std::max(0, (int)std::floor((u - mnMinX - factorX) * gwInv));
            const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((u - mnMinX + factorX) * gwInv));
            const int nMinCellY = std::max(0, (int)std::floor((v - mnMinY - factorY) * ghInv));
            const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((v - mnMinY + factorY) * ghInv));
            if (!(nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0)) {
                const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
                for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                    for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                        const std::vector<size_t>& vCell = G.mGrid[ix][iy];
                        for (size_t j = 0; j < vCell.size(); j++) {
                            const orc_keypoint& kpUn = keysUn[vCell[j]];
                            if (bCheckLevels) {
                                if (kpUn.octave < minLevel) continue;
                                if (maxLevel >= 0)
                                    if (kpUn.octave > maxLevel) continue;
                            }
                            const float distx = kpUn.x - u;
                            const float disty = kpUn.y - v;
                            if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) vIndices2.push_back(vCell[j]);
                        }
                    }
            }
        }
        if (vIndices2.empty()) continue;
        uint64_t a[4];
        std::memcpy(a, p.desc, 32);
        int bestDist = 256, bestIdx2 = -1;
        for (size_t k = 0; k < vIndices2.size(); k++) {
            const size_t i2 = vIndices2[k];
            if (taken[i2]) continue;
            uint64_t b[4];
            std::memcpy(b, desc + i2 * 32, 32);
            const int dist = hamming256(a, b);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx2 = (int)i2;
            }
        }
        if (bestDist <= ORBdist) {
            owner[bestIdx2] = i;
            taken[bestIdx2] = 1;
            match[i] = bestIdx2;
            nmatches++;
            if (checkOrientation) {
                float rot = p.angle - keysUn[bestIdx2].angle;
                if (rot < 0.0 || rot > 360.0) rot = std::fmod(rot, 360.0f);    // "REWRITTEN BUGGY PART" :2014-2018
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)std::round(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
                entryPoint[bin].push_back(i);
            }
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i != ind1 && i != ind2 && i != ind3) {
                for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                    owner[rotHist[i][j]] = -1;
                    nmatches--;
                    match[entryPoint[i][j]] = -1;
                }
            }
        }
    }
    return nmatches;
}

// ---- MapPoints projected into a key frame ---------------------------------------------------------------------------------
// KeyFrame keeps the Frame's grid (KeyFrame.cc ctor: mGrid copied cell by cell from F.mGrid), so it is rebuilt here with
// Frame::AssignFeaturesToGrid's rule.  KeyFrame::GetFeaturesInArea (KeyFrame.cc:706-750): no level filter, cells ix outer /
// iy inner, |dx| < r && |dy| < r.  KeyFrame::IsInImage (:752-755).
namespace {
struct KFGrid {
    std::vector<size_t> mGrid[GRID_COLS][GRID_ROWS];
    float mnMinX, mnMinY, mnMaxX, mnMaxY, mfGridElementWidthInv, mfGridElementHeightInv;
    const orc_keypoint* keys;
    void build(const orc_keypoint* k, int n, float minX, float minY, float maxX, float maxY) {
        keys = k;
        mnMinX = minX; mnMinY = minY; mnMaxX = maxX; mnMaxY = maxY;
        mfGridElementWidthInv = static_cast<float>(GRID_COLS) / (mnMaxX - mnMinX);
        mfGridElementHeightInv = static_cast<float>(GRID_ROWS) / (mnMaxY - mnMinY);
        for (int i = 0; i < GRID_COLS; i++)
            for (int j = 0; j < GRID_ROWS; j++) mGrid[i][j].clear();
        for (int i = 0; i < n; i++) {
            const int posX = (int)std::round((k[i].x - mnMinX) * mfGridElementWidthInv);
            const int posY = (int)std::round((k[i].y - mnMinY) * mfGridElementHeightInv);
            if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
            mGrid[posX][posY].push_back(i);
        }
    }
    bool IsInImage(const float& x, const float& y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }
    void GetFeaturesInArea(const float& x, const float& y, const float& r, std::vector<size_t>& vIndices) const {
        vIndices.clear();
        float factorX = r;
        float factorY = r;
        const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * mfGridElementWidthInv));
        if (nMinCellX >= GRID_COLS) return;
        const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * mfGridElementWidthInv));
        if (nMaxCellX < 0) return;
        const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * mfGridElementHeightInv));
        if (nMinCellY >= GRID_ROWS) return;
        const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * mfGridElementHeightInv));
        if (nMaxCellY < 0) return;
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const std::vector<size_t>& vCell = mGrid[ix][iy];
                for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
                    const orc_keypoint& kpUn = keys[vCell[j]];
                    const float distx = kpUn.x - x;
                    const float disty = kpUn.y - y;
                    if (std::fabs(distx) < r && std::fabs(disty) < r) vIndices.push_back(vCell[j]);
                }
            }
    }
};
int desc_distance(const uint8_t* a, const uint8_t* b) {
    uint64_t x[4], y[4];
    std::memcpy(x, a, 32);
    std::memcpy(y, b, 32);
    return hamming256(x, y);
}
}    // namespace

int orc_search_by_projection_sim3(const orc_keypoint* keysUn, const uint8_t* desc, const uint8_t* occupied, int n, float mnMinX,
                                  float mnMinY, float mnMaxX, float mnMaxY, const float* mvScaleFactors,
                                  const orc_frame_point* pts, int npts, float th, float ratioHamming, int32_t* match,
                                  int32_t* owner) {
    const int TH_LOW = 50;
    static thread_local KFGrid G;
    G.build(keysUn, n, mnMinX, mnMinY, mnMaxX, mnMaxY);
    std::vector<uint8_t> vpMatched(n, 0);
    for (int i = 0; i < n; i++) {
        vpMatched[i] = occupied ? (occupied[i] != 0) : 0;
        if (owner) owner[i] = -1;
    }
    int nmatches = 0;
    std::vector<size_t> vIndices;
    for (int iMP = 0; iMP < npts; iMP++) {
        match[iMP] = -1;
        const orc_frame_point& p = pts[iMP];
        if (!(p.flags & 1)) continue;    // isBad, spAlreadyFound, depth, distance range and viewing angle: the caller's
        if (!G.IsInImage(p.u, p.v)) continue;
        const int nPredictedLevel = p.level;
        const float radius = th * mvScaleFactors[nPredictedLevel];
        G.GetFeaturesInArea(p.u, p.v, radius, vIndices);
        if (vIndices.empty()) continue;
        int bestDist = 256;
        int bestIdx = -1;
        for (std::vector<size_t>::const_iterator vit = vIndices.begin(), vend = vIndices.end(); vit != vend; vit++) {
            const size_t idx = *vit;
            if (vpMatched[idx]) continue;
            const int& kpLevel = keysUn[idx].octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            const int dist = desc_distance(p.desc, desc + idx * 32);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx = (int)idx;
            }
        }
        if (bestDist <= TH_LOW * ratioHamming) {
            vpMatched[bestIdx] = 1;
            if (owner) owner[bestIdx] = iMP;
            match[iMP] = bestIdx;
            nmatches++;
        }
    }
    return nmatches;
}

int orc_fuse(const orc_keypoint* keys, const uint8_t* desc, const float* uRight, int n, float mnMinX, float mnMinY, float mnMaxX,
             float mnMaxY, const float* mvScaleFactors, const float* mvInvLevelSigma2, float bf, const orc_frame_point* pts,
             int npts, float th, int checkReprojection, int32_t* bestIdxOut, int32_t* bestDistOut) {
    const int TH_LOW = 50;
    static thread_local KFGrid G;
    G.build(keys, n, mnMinX, mnMinY, mnMaxX, mnMaxY);
    int nFused = 0;
    std::vector<size_t> vIndices;
    for (int i = 0; i < npts; i++) {
        bestIdxOut[i] = -1;
        if (bestDistOut) bestDistOut[i] = -1;
        const orc_frame_point& p = pts[i];
        if (!(p.flags & 1)) continue;
        const float invz = p.invz;
        const float uv0 = p.u, uv1 = p.v;
        if (!G.IsInImage(uv0, uv1)) continue;
        const float ur = uv0 - bf * invz;
        const int nPredictedLevel = p.level;
        const float radius = th * mvScaleFactors[nPredictedLevel];
        G.GetFeaturesInArea(uv0, uv1, radius, vIndices);
        if (vIndices.empty()) continue;
        int bestDist = 256;
        int bestIdx = -1;
        for (std::vector<size_t>::const_iterator vit = vIndices.begin(), vend = vIndices.end(); vit != vend; vit++) {
            size_t idx = *vit;
            const orc_keypoint& kp = keys[idx];
            const int& kpLevel = kp.octave;
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (checkReprojection) {
                const float mvuRight = uRight ? uRight[idx] : -1.0f;
                if (mvuRight >= 0) {
                    const float& kpx = kp.x;
                    const float& kpy = kp.y;
                    const float& kpr = mvuRight;
                    const float ex = uv0 - kpx;
                    const float ey = uv1 - kpy;
                    const float er = ur - kpr;
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * mvInvLevelSigma2[kpLevel] > 7.8) continue;
                } else {
                    const float& kpx = kp.x;
                    const float& kpy = kp.y;
                    const float ex = uv0 - kpx;
                    const float ey = uv1 - kpy;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * mvInvLevelSigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = desc_distance(p.desc, desc + idx * 32);
            if (dist < bestDist) {
                bestDist = dist;
                bestIdx = (int)idx;
            }
        }
        if (bestDist <= TH_LOW) {
            bestIdxOut[i] = bestIdx;
            if (bestDistOut) bestDistOut[i] = bestDist;
            nFused++;
        }
    }
    return nFused;
}

int orc_search_by_sim3(const orc_keypoint* keysUn1, const uint8_t* desc1, int n1, const orc_keypoint* keysUn2,
                       const uint8_t* desc2, int n2, const float* bounds1, const float* bounds2, const float* mvScaleFactors,
                       const orc_frame_point* pts1, const orc_frame_point* pts2, float th, int32_t* matches12,
                       int32_t* vnMatch1Out, int32_t* vnMatch2Out) {
    const int TH_HIGH = 100;
    static thread_local KFGrid G1, G2;
    G1.build(keysUn1, n1, bounds1[0], bounds1[1], bounds1[2], bounds1[3]);
    G2.build(keysUn2, n2, bounds2[0], bounds2[1], bounds2[2], bounds2[3]);
    std::vector<int> vnMatch1(n1, -1);
    std::vector<int> vnMatch2(n2, -1);
    std::vector<size_t> vIndices;
    // one direction of :1504-1583 / :1585-1664
    auto search = [&](const KFGrid& G, const orc_keypoint* keys, const uint8_t* desc, const orc_frame_point* pts, int npts,
                      std::vector<int>& vnMatch) {
        for (int i = 0; i < npts; i++) {
            const orc_frame_point& p = pts[i];
            if (!(p.flags & 1)) continue;
            const float u = p.u, v = p.v;
            if (!G.IsInImage(u, v)) continue;
            const int nPredictedLevel = p.level;
            const float radius = th * mvScaleFactors[nPredictedLevel];
            G.GetFeaturesInArea(u, v, radius, vIndices);
            if (vIndices.empty()) continue;
            int bestDist = INT_MAX;
            int bestIdx = -1;
            for (std::vector<size_t>::const_iterator vit = vIndices.begin(), vend = vIndices.end(); vit != vend; vit++) {
                const size_t idx = *vit;
                const orc_keypoint& kp = keys[idx];
                if (kp.octave < nPredictedLevel - 1 || kp.octave > nPredictedLevel) continue;
                const int dist = desc_distance(p.desc, desc + idx * 32);
                if (dist < bestDist) {
                    bestDist = dist;
                    bestIdx = (int)idx;
                }
            }
            if (bestDist <= TH_HIGH) vnMatch[i] = bestIdx;
        }
    };
    search(G2, keysUn2, desc2, pts1, n1, vnMatch1);
    search(G1, keysUn1, desc1, pts2, n2, vnMatch2);
    int nFound = 0;
    for (int i1 = 0; i1 < n1; i1++) {
        matches12[i1] = -1;
        int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            int idx1 = vnMatch2[idx2];
            if (idx1 == i1) {
                matches12[i1] = idx2;
                nFound++;
            }
        }
    }
    if (vnMatch1Out) std::copy(vnMatch1.begin(), vnMatch1.end(), vnMatch1Out);
    if (vnMatch2Out) std::copy(vnMatch2.begin(), vnMatch2.end(), vnMatch2Out);
    return nFound;
}

// ---- ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches) (ORBmatcher.cc:224-425) ------
// DBoW2::FeatureVector = std::map<NodeId, vector<unsigned>>: passed as CSR (node ids ascending, start[nnodes + 1], feature
// indices in each node's vector order).  validKF[i] = pMP && !pMP->isBad() of vpMapPointsKF[i]; angleKF / angleF: the
// angle of the keypoint the reference picks for that feature (:330-345).  matches[i] = KF feature whose MapPoint ends in
// vpMapPointMatches[i], -1 = NULL.  The right-camera branch sits INSIDE `if (bestDist1 <= TH_LOW)` and its ratio test is
// `|| true` (:355-357): both reproduced.
int orc_search_by_bow(const uint8_t* descKF, const float* angleKF, const uint8_t* validKF, int nKF, const int32_t* kfNodeIds,
                      const int32_t* kfNodeStart, const int32_t* kfIdx, int kfNodes, const uint8_t* descF,
                      const float* angleF, int nF, int nleftF, const int32_t* fNodeIds, const int32_t* fNodeStart,
                      const int32_t* fIdx, int fNodes, float mfNNratio, int checkOrientation, int32_t* matches) {
    const int TH_LOW = 50, HISTO_LENGTH = 30;
    (void)nKF;
    for (int i = 0; i < nF; i++) matches[i] = -1;    // vpMapPointMatches = vector<MapPoint*>(F.N, NULL)
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int KFit = 0, Fit = 0;
    while (KFit != kfNodes && Fit != fNodes) {
        if (kfNodeIds[KFit] == fNodeIds[Fit]) {
            for (int iKF = kfNodeStart[KFit]; iKF < kfNodeStart[KFit + 1]; iKF++) {
                const unsigned int realIdxKF = (unsigned)kfIdx[iKF];
                if (!validKF[realIdxKF]) continue;    // !pMP || pMP->isBad()
                uint64_t a[4];
                std::memcpy(a, descKF + (size_t)realIdxKF * 32, 32);
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                int bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
                for (int iF = fNodeStart[Fit]; iF < fNodeStart[Fit + 1]; iF++) {
                    const unsigned int realIdxF = (unsigned)fIdx[iF];
                    if (matches[realIdxF] >= 0) continue;    // vpMapPointMatches[realIdxF]
                    uint64_t b[4];
                    std::memcpy(b, descF + (size_t)realIdxF * 32, 32);
                    const int dist = hamming256(a, b);
                    if (nleftF == -1) {
                        if (dist < bestDist1) {
                            bestDist2 = bestDist1;
                            bestDist1 = dist;
                            bestIdxF = (int)realIdxF;
                        } else if (dist < bestDist2) {
                            bestDist2 = dist;
                        }
                    } else {
                        if (int(realIdxF) < nleftF && dist < bestDist1) {
                            bestDist2 = bestDist1;
                            bestDist1 = dist;
                            bestIdxF = (int)realIdxF;
                        } else if (int(realIdxF) < nleftF && dist < bestDist2) {
                            bestDist2 = dist;
                        }
                        if (int(realIdxF) >= nleftF && dist < bestDist1R) {
                            bestDist2R = bestDist1R;
                            bestDist1R = dist;
                            bestIdxFR = (int)realIdxF;
                        } else if (int(realIdxF) >= nleftF && dist < bestDist2R) {
                            bestDist2R = dist;
                        }
                    }
                }
                auto push = [&](int idxF) {
                    if (!checkOrientation) return;
                    float rot = angleKF[realIdxKF] - angleF[idxF];
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(idxF);
                };
                if (bestDist1 <= TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        matches[bestIdxF] = (int)realIdxKF;
                        push(bestIdxF);
                        nmatches++;
                    }
                    if (bestDist1R <= TH_LOW) {
                        if (static_cast<float>(bestDist1R) < mfNNratio * static_cast<float>(bestDist2R) || true) {
                            matches[bestIdxFR] = (int)realIdxKF;
                            push(bestIdxFR);
                            nmatches++;
                        }
                    }
                }
            }
            KFit++;
            Fit++;
        } else if (kfNodeIds[KFit] < fNodeIds[Fit]) {
            KFit = (int)(std::lower_bound(kfNodeIds, kfNodeIds + kfNodes, fNodeIds[Fit]) - kfNodeIds);
        } else {
            Fit = (int)(std::lower_bound(fNodeIds, fNodeIds + fNodes, kfNodeIds[KFit]) - fNodeIds);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                matches[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// ---- ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (ORBmatcher.cc:766-906) --------
int orc_search_by_bow_kf(const uint8_t* desc1, const float* angle1, const uint8_t* valid1, int n1, const int32_t* nodeIds1,
                         const int32_t* nodeStart1, const int32_t* idx1v, int nodes1, const uint8_t* desc2, const float* angle2,
                         const uint8_t* valid2, const int32_t* nodeIds2, const int32_t* nodeStart2, const int32_t* idx2v,
                         int nodes2, float mfNNratio, int checkOrientation, int32_t* matches12) {
    const int TH_LOW = 50, HISTO_LENGTH = 30;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    int maxIdx2 = 0;
    for (int t = 0; t < nodeStart2[nodes2]; t++) maxIdx2 = std::max(maxIdx2, idx2v[t] + 1);
    std::vector<bool> vbMatched2(maxIdx2, false);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int nmatches = 0;
    int f1it = 0, f2it = 0;
    while (f1it != nodes1 && f2it != nodes2) {
        if (nodeIds1[f1it] == nodeIds2[f2it]) {
            for (int i1 = nodeStart1[f1it]; i1 < nodeStart1[f1it + 1]; i1++) {
                const size_t idx1 = (size_t)idx1v[i1];
                if (!valid1[idx1]) continue;    // NLeft / mvKeysUn.size(), !pMP1, pMP1->isBad()
                int bestDist1 = 256;
                int bestIdx2 = -1;
                int bestDist2 = 256;
                for (int i2 = nodeStart2[f2it]; i2 < nodeStart2[f2it + 1]; i2++) {
                    const size_t idx2 = (size_t)idx2v[i2];
                    if (vbMatched2[idx2] || !valid2[idx2]) continue;
                    int dist = desc_distance(desc1 + idx1 * 32, desc2 + idx2 * 32);
                    if (dist < bestDist1) {
                        bestDist2 = bestDist1;
                        bestDist1 = dist;
                        bestIdx2 = (int)idx2;
                    } else if (dist < bestDist2) {
                        bestDist2 = dist;
                    }
                }
                if (bestDist1 < TH_LOW) {
                    if (static_cast<float>(bestDist1) < mfNNratio * static_cast<float>(bestDist2)) {
                        matches12[idx1] = bestIdx2;
                        vbMatched2[bestIdx2] = true;
                        if (checkOrientation) {
                            float rot = angle1[idx1] - angle2[bestIdx2];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)std::round(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back((int)idx1);
                        }
                        nmatches++;
                    }
                }
            }
            f1it++;
            f2it++;
        } else if (nodeIds1[f1it] < nodeIds2[f2it]) {
            f1it = (int)(std::lower_bound(nodeIds1, nodeIds1 + nodes1, nodeIds2[f2it]) - nodeIds1);
        } else {
            f2it = (int)(std::lower_bound(nodeIds2, nodeIds2 + nodes2, nodeIds1[f1it]) - nodeIds2);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                matches12[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// ---- ORBmatcher::SearchForTriangulation (ORBmatcher.cc:908-1153), one camera per key frame; Pinhole.cpp:102-124 ---------------
namespace {
bool pinhole_epipolar_constrain(const float* F12, const orc_keypoint& kp1, const orc_keypoint& kp2, const float unc) {
    // F12 row-major: F12(r, c) = F12[3 * r + c].  Epipolar line in second image l = x1'F12 = [a b c]
    const float a = kp1.x * F12[0] + kp1.y * F12[3] + F12[6];
    const float b = kp1.x * F12[1] + kp1.y * F12[4] + F12[7];
    const float c = kp1.x * F12[2] + kp1.y * F12[5] + F12[8];
    const float num = a * kp2.x + b * kp2.y + c;
    const float den = a * a + b * b;
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * unc;
}
}    // namespace

int orc_search_for_triangulation(const orc_keypoint* keys1, const uint8_t* desc1, const uint8_t* free1, const uint8_t* stereo1,
                                 int n1, const int32_t* nodeIds1, const int32_t* nodeStart1, const int32_t* idx1v, int nodes1,
                                 const orc_keypoint* keys2, const uint8_t* desc2, const uint8_t* free2, const uint8_t* stereo2,
                                 const int32_t* nodeIds2, const int32_t* nodeStart2, const int32_t* idx2v, int nodes2,
                                 const float* F12, const float* ep, const float* mvScaleFactors, const float* mvLevelSigma2,
                                 int bCoarse, int checkOrientation, int32_t* vMatches12) {
    const int TH_LOW = 50, HISTO_LENGTH = 30;
    int nmatches = 0;
    for (int i = 0; i < n1; i++) vMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int f1it = 0, f2it = 0;
    while (f1it != nodes1 && f2it != nodes2) {
        if (nodeIds1[f1it] == nodeIds2[f2it]) {
            for (int i1 = nodeStart1[f1it]; i1 < nodeStart1[f1it + 1]; i1++) {
                const size_t idx1 = (size_t)idx1v[i1];
                if (!free1[idx1]) continue;    // pMP1, or bOnlyStereo && !bStereo1
                const bool bStereo1 = stereo1 && stereo1[idx1];
                const orc_keypoint& kp1 = keys1[idx1];
                int bestDist = TH_LOW;
                int bestIdx2 = -1;
                for (int i2 = nodeStart2[f2it]; i2 < nodeStart2[f2it + 1]; i2++) {
                    size_t idx2 = (size_t)idx2v[i2];
                    if (!free2[idx2]) continue;    // vbMatched2 (never set) || pMP2, or bOnlyStereo && !bStereo2
                    const bool bStereo2 = stereo2 && stereo2[idx2];
                    const int dist = desc_distance(desc1 + idx1 * 32, desc2 + idx2 * 32);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const orc_keypoint& kp2 = keys2[idx2];
                    if (!bStereo1 && !bStereo2) {    // && !pKF1->mpCamera2
                        const float distex = ep[0] - kp2.x;
                        const float distey = ep[1] - kp2.y;
                        if (distex * distex + distey * distey < 100 * mvScaleFactors[kp2.octave]) continue;
                    }
                    if (bCoarse || pinhole_epipolar_constrain(F12, kp1, kp2, mvLevelSigma2[kp2.octave])) {
                        bestIdx2 = (int)idx2;
                        bestDist = dist;
                    }
                }
                if (bestIdx2 >= 0) {
                    const orc_keypoint& kp2 = keys2[bestIdx2];
                    vMatches12[idx1] = bestIdx2;
                    nmatches++;
                    if (checkOrientation) {
                        float rot = kp1.angle - kp2.angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back((int)idx1);
                    }
                }
            }
            f1it++;
            f2it++;
        } else if (nodeIds1[f1it] < nodeIds2[f2it]) {
            f1it = (int)(std::lower_bound(nodeIds1, nodeIds1 + nodes1, nodeIds2[f2it]) - nodeIds1);
        } else {
            f2it = (int)(std::lower_bound(nodeIds2, nodeIds2 + nodes2, nodeIds1[f1it]) - nodeIds2);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                vMatches12[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// ---- MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401), after the gathering of vDescriptors ------------------------
int orc_distinctive_descriptor(const uint8_t* vDescriptors, int Nin, int32_t* bestMedianOut) {
    const size_t N = (size_t)Nin;
    std::vector<float> Distances(N * N);    // float Distances[N][N]
    for (size_t i = 0; i < N; i++) {
        Distances[i * N + i] = 0;
        for (size_t j = i + 1; j < N; j++) {
            int distij = desc_distance(vDescriptors + i * 32, vDescriptors + j * 32);
            Distances[i * N + j] = distij;
            Distances[j * N + i] = distij;
        }
    }
    // Take the descriptor with least median distance to the rest
    int BestMedian = INT_MAX;
    int BestIdx = 0;
    for (size_t i = 0; i < N; i++) {
        std::vector<int> vDists(Distances.begin() + i * N, Distances.begin() + i * N + N);
        std::sort(vDists.begin(), vDists.end());
        int median = vDists[0.5 * (N - 1)];
        if (median < BestMedian) {
            BestMedian = median;
            BestIdx = (int)i;
        }
    }
    if (bestMedianOut) *bestMedianOut = BestMedian;
    return BestIdx;
}

// ---- ORBmatcher::SearchForInitialization (ORBmatcher.cc:649-764) + ComputeThreeMaxima (:2061-2102) ----------------------
// F1 / F2 are monocular frames (Nleft == -1); the grid is F2's.  vMatchedDistance / vnMatches21 make the loop order
// dependent (a later keypoint steals a match only with a strictly smaller distance, and matched candidates with a
// distance <= the current one are invisible to the best / second-best bookkeeping).  HISTO_LENGTH = 30, TH_LOW = 50.
int orc_search_for_initialization(const orc_keypoint* keys1, const uint8_t* desc1, int n1, const orc_keypoint* keys2,
                                  const uint8_t* desc2, int n2, float mnMinX, float mnMinY, float mnMaxX, float mnMaxY,
                                  float* prevMatchedXY, int windowSize, float mfNNratio, int checkOrientation,
                                  int32_t* vnMatches12) {
    const int TH_LOW = 50, HISTO_LENGTH = 30;
    const float gwInv = static_cast<float>(GRID_COLS) / (mnMaxX - mnMinX);
    const float ghInv = static_cast<float>(GRID_ROWS) / (mnMaxY - mnMinY);
    static thread_local std::vector<size_t> mGrid[GRID_COLS][GRID_ROWS];
    for (int i = 0; i < GRID_COLS; i++)
        for (int j = 0; j < GRID_ROWS; j++) mGrid[i][j].clear();
    for (int i = 0; i < n2; i++) {
        const int posX = (int)std::round((keys2[i].x - mnMinX) * gwInv);
        const int posY = (int)std::round((keys2[i].y - mnMinY) * ghInv);
        if (posX < 0 || posX >= GRID_COLS || posY < 0 || posY >= GRID_ROWS) continue;
        mGrid[posX][posY].push_back(i);
    }
    int nmatches = 0;
    for (int i = 0; i < n1; i++) vnMatches12[i] = -1;
    std::vector<int> rotHist[30];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> vMatchedDistance(n2, INT_MAX);
    std::vector<int> vnMatches21(n2, -1);
    std::vector<size_t> vIndices2;
    for (int i1 = 0; i1 < n1; i1++) {
        const orc_keypoint kp1 = keys1[i1];
        const int level1 = kp1.octave;
        if (level1 > 0) continue;
        // F2.GetFeaturesInArea(vbPrevMatched[i1].x, vbPrevMatched[i1].y, windowSize, level1, level1)
        vIndices2.clear();
        {
            const float x = prevMatchedXY[2 * i1], y = prevMatchedXY[2 * i1 + 1];
            const float factorX = windowSize, factorY = windowSize;
            const int minLevel = level1, maxLevel = level1;
            const int nMinCellX = std::max(0, (int)std::floor((x - mnMinX - factorX) * gwInv));
            const int nMaxCellX = std::min((int)GRID_COLS - 1, (int)std::ceil((x - mnMinX + factorX) * gwInv));
            const int nMinCellY = std::max(0, (int)std::floor((y - mnMinY - factorY) * ghInv));
            const int nMaxCellY = std::min((int)GRID_ROWS - 1, (int)std::ceil((y - mnMinY + factorY) * ghInv));
            if (!(nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0)) {
                const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
                for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
                    for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                        const std::vector<size_t>& vCell = mGrid[ix][iy];
                        for (size_t j = 0; j < vCell.size(); j++) {
                            const orc_keypoint& kpUn = keys2[vCell[j]];
                            if (bCheckLevels) {
                                if (kpUn.octave < minLevel) continue;
                                if (maxLevel >= 0)
                                    if (kpUn.octave > maxLevel) continue;
                            }
                            const float distx = kpUn.x - x;
                            const float disty = kpUn.y - y;
                            if (std::fabs(distx) < factorX && std::fabs(disty) < factorY) vIndices2.push_back(vCell[j]);
                        }
                    }
            }
        }
        if (vIndices2.empty()) continue;
        uint64_t a[4];
        std::memcpy(a, desc1 + (size_t)i1 * 32, 32);
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (size_t k = 0; k < vIndices2.size(); k++) {
            const size_t i2 = vIndices2[k];
            uint64_t b[4];
            std::memcpy(b, desc2 + i2 * 32, 32);
            const int dist = hamming256(a, b);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) {
                bestDist2 = bestDist;
                bestDist = dist;
                bestIdx2 = (int)i2;
            } else if (dist < bestDist2) {
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * mfNNratio) {
                if (vnMatches21[bestIdx2] >= 0) {
                    vnMatches12[vnMatches21[bestIdx2]] = -1;
                    nmatches--;
                }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOrientation) {
                    float rot = keys1[i1].angle - keys2[bestIdx2].angle;
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)std::round(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        {    // ComputeThreeMaxima
            int max1 = 0, max2 = 0, max3 = 0;
            for (int i = 0; i < HISTO_LENGTH; i++) {
                const int sz = (int)rotHist[i].size();
                if (sz > max1) {
                    max3 = max2;
                    max2 = max1;
                    max1 = sz;
                    ind3 = ind2;
                    ind2 = ind1;
                    ind1 = i;
                } else if (sz > max2) {
                    max3 = max2;
                    max2 = sz;
                    ind3 = ind2;
                    ind2 = i;
                } else if (sz > max3) {
                    max3 = sz;
                    ind3 = i;
                }
            }
            if (max2 < 0.1f * (float)max1) {
                ind2 = -1;
                ind3 = -1;
            } else if (max3 < 0.1f * (float)max1) {
                ind3 = -1;
            }
        }
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                const int idx1 = rotHist[i][j];
                if (vnMatches12[idx1] >= 0) {
                    vnMatches12[idx1] = -1;
                    nmatches--;
                }
            }
        }
    }
    for (int i1 = 0; i1 < n1; i1++)
        if (vnMatches12[i1] >= 0) {
            prevMatchedXY[2 * i1] = keys2[vnMatches12[i1]].x;
            prevMatchedXY[2 * i1 + 1] = keys2[vnMatches12[i1]].y;
        }
    return nmatches;
}

// ---- Frame::ComputeStereoFishEyeMatches, the part after BFMatchORB (Frame.cc:1164-1204) with
//      KannalaBrandt8::unproject / project(Vector3f) / Triangulate / TriangulateMatches (KannalaBrandt8.cpp:60-79,
//      107-136, 300-363, 380-392).  Everything is float as in the reference, except the 4x4 SVD: the reference calls
//      Eigen::JacobiSVD<Matrix4f> (Eigen 3.4.0, vendored under Thirdparty/); here the right singular vector of the
//      smallest singular value comes from a one-sided Jacobi SVD in double on the same float matrix, which is at least
//      as accurate — parity with the Eigen build is by tolerance (tests/golden/fisheye_*.npz, made with the real Eigen).
static void kb_unproject(const float* prm, float precision, float px, float py, float out[3]) {
    const float pwx = (px - prm[2]) / prm[0], pwy = (py - prm[3]) / prm[1];
    float scale = 1.f;
    float theta_d = sqrtf(pwx * pwx + pwy * pwy);
    theta_d = fminf(fmaxf(-(float)(3.1415926535897932384626433832795 / 2.f), theta_d), (float)(3.1415926535897932384626433832795 / 2.f));
    if (theta_d > 1e-8) {
        float theta = theta_d;
        for (int j = 0; j < 10; j++) {
            const float theta2 = theta * theta, theta4 = theta2 * theta2, theta6 = theta4 * theta2, theta8 = theta4 * theta4;
            const float k0_theta2 = prm[4] * theta2, k1_theta4 = prm[5] * theta4;
            const float k2_theta6 = prm[6] * theta6, k3_theta8 = prm[7] * theta8;
            const float theta_fix = (theta * (1 + k0_theta2 + k1_theta4 + k2_theta6 + k3_theta8) - theta_d) /
                                    (1 + 3 * k0_theta2 + 5 * k1_theta4 + 7 * k2_theta6 + 9 * k3_theta8);
            theta = theta - theta_fix;
            if (fabsf(theta_fix) < precision) break;
        }
        scale = std::tan(theta) / theta_d;
    }
    out[0] = pwx * scale;
    out[1] = pwy * scale;
    out[2] = 1.f;
}

static void kb_project(const float* prm, const float v[3], float uv[2]) {
    const float x2_plus_y2 = v[0] * v[0] + v[1] * v[1];
    const float theta = atan2f(sqrtf(x2_plus_y2), v[2]);
    const float psi = atan2f(v[1], v[0]);
    const float theta2 = theta * theta;
    const float theta3 = theta * theta2;
    const float theta5 = theta3 * theta2;
    const float theta7 = theta5 * theta2;
    const float theta9 = theta7 * theta2;
    const float r = theta + prm[4] * theta3 + prm[5] * theta5 + prm[6] * theta7 + prm[7] * theta9;
    uv[0] = prm[0] * r * std::cos(psi) + prm[2];
    uv[1] = prm[1] * r * std::sin(psi) + prm[3];
}

// right singular vector of the smallest singular value of a 4x4 matrix: one-sided (Hestenes) Jacobi, double
static orc_null_vector4_fn g_nullVectorHook = nullptr;
void orc_set_null_vector4_hook(orc_null_vector4_fn fn) { g_nullVectorHook = fn; }

static void null_vector4(const float Af[4][4], double v[4]) {
    if (g_nullVectorHook) {    // oracle/ref_eigen: the real Eigen::JacobiSVD<Matrix4f>, used to make the golden fixtures
        float vf[4];
        g_nullVectorHook(&Af[0][0], vf);
        for (int i = 0; i < 4; i++) v[i] = vf[i];
        return;
    }
    double U[4][4], V[4][4];
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            U[i][j] = Af[i][j];
            V[i][j] = i == j ? 1.0 : 0.0;
        }
    for (int sweep = 0; sweep < 30; sweep++) {
        bool rotated = false;
        for (int p = 0; p < 3; p++)
            for (int q = p + 1; q < 4; q++) {
                double alpha = 0, beta = 0, gamma = 0;
                for (int i = 0; i < 4; i++) {
                    alpha += U[i][p] * U[i][p];
                    beta += U[i][q] * U[i][q];
                    gamma += U[i][p] * U[i][q];
                }
                if (gamma == 0.0 || std::fabs(gamma) <= 1e-15 * std::sqrt(alpha * beta)) continue;
                rotated = true;
                const double zeta = (beta - alpha) / (2.0 * gamma);
                const double t = (zeta >= 0 ? 1.0 : -1.0) / (std::fabs(zeta) + std::sqrt(1.0 + zeta * zeta));
                const double c = 1.0 / std::sqrt(1.0 + t * t), sn = c * t;
                for (int i = 0; i < 4; i++) {
                    const double up = U[i][p], uq = U[i][q];
                    U[i][p] = c * up - sn * uq;
                    U[i][q] = sn * up + c * uq;
                    const double vp = V[i][p], vq = V[i][q];
                    V[i][p] = c * vp - sn * vq;
                    V[i][q] = sn * vp + c * vq;
                }
            }
        if (!rotated) break;
    }
    int best = 0;
    double bestN = 1e300;
    for (int j = 0; j < 4; j++) {
        double nrm = 0;
        for (int i = 0; i < 4; i++) nrm += U[i][j] * U[i][j];
        if (nrm < bestN) {
            bestN = nrm;
            best = j;
        }
    }
    for (int i = 0; i < 4; i++) v[i] = V[i][best];
}

// KannalaBrandt8::TriangulateMatches (:300-363); R12 row-major.  Returns z1 or the negative reason.
static float kb_triangulate_matches(const orc_fisheye_rig* g, float x1, float y1, float x2, float y2, float sigmaLevel,
                                    float unc, float p3D[3], float* slack) {
    *slack = 1e30f;    // smallest distance of a tested quantity from its threshold (relative for the chi-square gates)
    auto note = [&](double d) { if (std::fabs(d) < *slack) *slack = (float)std::fabs(d); };
    float r1[3], r2[3], r21[3];
    kb_unproject(g->cam1, g->precision, x1, y1, r1);
    kb_unproject(g->cam2, g->precision, x2, y2, r2);
    const float* R = g->Rlr;
    for (int i = 0; i < 3; i++) r21[i] = R[3 * i] * r2[0] + R[3 * i + 1] * r2[1] + R[3 * i + 2] * r2[2];
    const float dot = r1[0] * r21[0] + r1[1] * r21[1] + r1[2] * r21[2];
    const float n1 = std::sqrt(r1[0] * r1[0] + r1[1] * r1[1] + r1[2] * r1[2]);
    const float n2 = std::sqrt(r21[0] * r21[0] + r21[1] * r21[1] + r21[2] * r21[2]);
    const float cosParallaxRays = dot / (n1 * n2);
    note(cosParallaxRays - 0.99998);
    if (cosParallaxRays > 0.99998) return -1;
    // Tcw1 = [I | 0], Tcw2 = [R21 | -R21 * t12], R21 = R12^T
    float T2[3][4];
    for (int i = 0; i < 3; i++) {
        for (int j = 0; j < 3; j++) T2[i][j] = R[3 * j + i];
        T2[i][3] = -(R[i] * g->tlr[0] + R[3 + i] * g->tlr[1] + R[6 + i] * g->tlr[2]);
    }
    float A[4][4];
    const float T1[3][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}};
    for (int j = 0; j < 4; j++) {
        A[0][j] = r1[0] * T1[2][j] - T1[0][j];
        A[1][j] = r1[1] * T1[2][j] - T1[1][j];
        A[2][j] = r2[0] * T2[2][j] - T2[0][j];
        A[3][j] = r2[1] * T2[2][j] - T2[1][j];
    }
    double v[4];
    null_vector4(A, v);
    float x3D[3];
    if (g_nullVectorHook) {    // float division as x3Dh.head(3) / x3Dh(3)
        for (int i = 0; i < 3; i++) x3D[i] = (float)v[i] / (float)v[3];
    } else {
        for (int i = 0; i < 3; i++) x3D[i] = (float)(v[i] / v[3]);
    }
    const float z1 = x3D[2];
    note(z1);
    if (!(z1 > 0)) return -2;    // z1 <= 0, and a NaN from a zero w counts as no point
    const float z2 = T2[2][0] * x3D[0] + T2[2][1] * x3D[1] + T2[2][2] * x3D[2] + T2[2][3];
    note(z2);
    if (z2 <= 0) return -3;
    float uv1[2];
    kb_project(g->cam1, x3D, uv1);
    const float errX1 = uv1[0] - x1, errY1 = uv1[1] - y1;
    note((errX1 * errX1 + errY1 * errY1) / (5.991 * 8 * sigmaLevel) - 1.0);
    if ((errX1 * errX1 + errY1 * errY1) > 5.991 * 8 * sigmaLevel) return -4;
    float x3D2[3];
    for (int i = 0; i < 3; i++) x3D2[i] = T2[i][0] * x3D[0] + T2[i][1] * x3D[1] + T2[i][2] * x3D[2] + T2[i][3];
    float uv2[2];
    kb_project(g->cam2, x3D2, uv2);
    const float errX2 = uv2[0] - x2, errY2 = uv2[1] - y2;
    note((errX2 * errX2 + errY2 * errY2) / (5.991 * 8 * unc) - 1.0);
    if ((errX2 * errX2 + errY2 * errY2) > 5.991 * 8 * unc) return -5;
    p3D[0] = x3D[0];
    p3D[1] = x3D[1];
    p3D[2] = x3D[2];
    return z1;
}

// ---- ORBmatcher::SearchForTriangulation for two-camera key frames (ORBmatcher.cc:908-1153 with :1014-1043) ------------------
// Features = left keypoints then right keypoints; KannalaBrandt8::epipolarConstrain (KannalaBrandt8.cpp:210-214).  rig->cam1 /
// cam2 = mpCamera / mpCamera2 parameters; T12[c] = R12 (row-major) | t12 for c = 0 Tll, 1 Tlr, 2 Trl, 3 Trr.  slack[i1]
// (optional) = the smallest distance of any tested quantity from its gate over the candidates evaluated for feature i1.
int orc_search_for_triangulation_kb8(const orc_keypoint* keys1, const uint8_t* desc1, const uint8_t* free1, int n1, int nleft1,
                                     const int32_t* nodeIds1, const int32_t* nodeStart1, const int32_t* idx1v, int nodes1,
                                     const orc_keypoint* keys2, const uint8_t* desc2, const uint8_t* free2, int nleft2,
                                     const int32_t* nodeIds2, const int32_t* nodeStart2, const int32_t* idx2v, int nodes2,
                                     const orc_fisheye_rig* rig, const float* T12, const float* mvLevelSigma2, int bCoarse,
                                     int checkOrientation, int32_t* vMatches12, float* slack) {
    const int TH_LOW = 50, HISTO_LENGTH = 30;
    int nmatches = 0;
    for (int i = 0; i < n1; i++) {
        vMatches12[i] = -1;
        if (slack) slack[i] = 1e30f;
    }
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int f1it = 0, f2it = 0;
    while (f1it != nodes1 && f2it != nodes2) {
        if (nodeIds1[f1it] == nodeIds2[f2it]) {
            for (int i1 = nodeStart1[f1it]; i1 < nodeStart1[f1it + 1]; i1++) {
                const size_t idx1 = (size_t)idx1v[i1];
                if (!free1[idx1]) continue;    // pMP1 (bStereo1 is false with mpCamera2: bOnlyStereo would drop everything)
                const orc_keypoint& kp1 = keys1[idx1];    // mvKeys[idx1] or mvKeysRight[idx1 - NLeft]: concatenated here
                const bool bRight1 = (int)idx1 >= nleft1;
                int bestDist = TH_LOW;
                int bestIdx2 = -1;
                for (int i2 = nodeStart2[f2it]; i2 < nodeStart2[f2it + 1]; i2++) {
                    size_t idx2 = (size_t)idx2v[i2];
                    if (!free2[idx2]) continue;
                    const int dist = desc_distance(desc1 + idx1 * 32, desc2 + idx2 * 32);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    const orc_keypoint& kp2 = keys2[idx2];
                    const bool bRight2 = (int)idx2 >= nleft2;
                    // :1015-1042: R12 / t12 / pCamera1 / pCamera2 by the sides of the two features
                    const int c = (bRight1 ? 2 : 0) + (bRight2 ? 1 : 0);
                    bool ok = bCoarse != 0;
                    if (!ok) {
                        orc_fisheye_rig g = *rig;
                        std::memcpy(g.cam1, bRight1 ? rig->cam2 : rig->cam1, sizeof(g.cam1));
                        std::memcpy(g.cam2, bRight2 ? rig->cam2 : rig->cam1, sizeof(g.cam2));
                        std::memcpy(g.Rlr, T12 + 12 * c, 9 * sizeof(float));
                        std::memcpy(g.tlr, T12 + 12 * c + 9, 3 * sizeof(float));
                        float p3D[3], sl;
                        const float z = kb_triangulate_matches(&g, kp1.x, kp1.y, kp2.x, kp2.y, mvLevelSigma2[kp1.octave],
                                                               mvLevelSigma2[kp2.octave], p3D, &sl);
                        if (z > 0 && std::fabs(z - 0.0001f) < sl) sl = std::fabs(z - 0.0001f);
                        if (slack && sl < slack[idx1]) slack[idx1] = sl;
                        ok = z > 0.0001f;    // KannalaBrandt8::epipolarConstrain
                    }
                    if (ok) {
                        bestIdx2 = (int)idx2;
                        bestDist = dist;
                    }
                }
                if (bestIdx2 >= 0) {
                    const orc_keypoint& kp2 = keys2[bestIdx2];
                    vMatches12[idx1] = bestIdx2;
                    nmatches++;
                    if (checkOrientation) {
                        float rot = kp1.angle - kp2.angle;
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)std::round(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back((int)idx1);
                    }
                }
            }
            f1it++;
            f2it++;
        } else if (nodeIds1[f1it] < nodeIds2[f2it]) {
            f1it = (int)(std::lower_bound(nodeIds1, nodeIds1 + nodes1, nodeIds2[f2it]) - nodeIds1);
        } else {
            f2it = (int)(std::lower_bound(nodeIds2, nodeIds2 + nodes2, nodeIds1[f1it]) - nodeIds2);
        }
    }
    if (checkOrientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0, jend = rotHist[i].size(); j < jend; j++) {
                vMatches12[rotHist[i][j]] = -1;
                nmatches--;
            }
        }
    }
    return nmatches;
}

// test hook: KannalaBrandt8::TriangulateMatches for one pair (rig->cam1 / cam2 / Rlr / tlr as given)
float orc_kb_triangulate_matches(const orc_fisheye_rig* rig, float x1, float y1, float x2, float y2, float sigmaLevel, float unc) {
    float p3D[3], sl;
    return kb_triangulate_matches(rig, x1, y1, x2, y2, sigmaLevel, unc, p3D, &sl);
}

int orc_stereo_fisheye_matches(const orc_keypoint* keysL, int nL, int monoL, const orc_keypoint* keysR, int nR, int monoR,
                               const int32_t* knnIdx, const int32_t* knnDist, const orc_fisheye_rig* rig,
                               int32_t* leftToRight, int32_t* rightToLeft, float* depth, float* p3d, int32_t* status, float* slack) {
    for (int i = 0; i < nL; i++) {
        if (slack) slack[i] = 1e30f;
        leftToRight[i] = -1;
        depth[i] = -1.0f;
        p3d[3 * i] = p3d[3 * i + 1] = p3d[3 * i + 2] = 0.f;
        if (status) status[i] = 0;
    }
    for (int i = 0; i < nR; i++) rightToLeft[i] = -1;
    int nMatches = 0;
    const int max_distance = 70;
    for (int i = 0; i < nL - monoL; i++) {
        const int d1 = knnDist[2 * i];
        if (d1 == 0) continue;    // "Bugged point" (:1171-1175); d1 < 0 = no neighbour at all
        if (d1 > 0 && d1 < max_distance) {
            const int leftPos = i, rightPos = knnIdx[2 * i];
            if (rightPos + monoR >= nR || rightPos < 0) continue;
            const orc_keypoint& kl = keysL[leftPos + monoL];
            const orc_keypoint& kr = keysR[rightPos + monoR];
            const float sigma1 = rig->levelSigma2[kl.octave], sigma2 = rig->levelSigma2[kr.octave];
            float p[3] = {0, 0, 0};
            float sl;
            const float dep = kb_triangulate_matches(rig, kl.x, kl.y, kr.x, kr.y, sigma1, sigma2, p, &sl);
            if (slack) slack[leftPos + monoL] = sl;
            if (status) status[leftPos + monoL] = dep > 0.0001f ? 1 : (dep < 0 ? (int)dep : -6);
            if (dep > 0.0001f) {
                leftToRight[leftPos + monoL] = rightPos + monoR;
                rightToLeft[rightPos + monoR] = leftPos + monoL;
                p3d[3 * (leftPos + monoL)] = p[0];
                p3d[3 * (leftPos + monoL) + 1] = p[1];
                p3d[3 * (leftPos + monoL) + 2] = p[2];
                depth[leftPos + monoL] = dep;
                nMatches++;
            }
        }
    }
    return nMatches;
}

int orc_hardware_threads(void) { return (int)std::thread::hardware_concurrency(); }

}  // extern "C"
