// ref_ext_api.cpp — C entry points around the REFERENCE'S OWN ORB_SLAM3::ORBextractor (ORBextractor_old.cc compiled
// from /root/reference by make_ref_ext.py).  Test infrastructure: used by tests/test_oracle_ref.py and
// oracle/gen_golden.py to pin the restated oracle (oracle/cpp) and the golden fixtures against the reference's code.
// No algorithm lives here: every function marshals arguments and calls a member of the reference class.
#include "ORBextractor.h"    // generated from app/src/main/cpp/include/ORBextractor_old.h

#include <cstring>

namespace {
// protected members of the reference class are reachable from a derived class; nothing is overridden
struct Probe : public ORB_SLAM3::ORBextractor {
    using ORB_SLAM3::ORBextractor::ORBextractor;
    using ORB_SLAM3::ORBextractor::ComputeKeyPointsOctTree;
    using ORB_SLAM3::ORBextractor::DistributeOctTree;
    using ORB_SLAM3::ORBextractor::mnFeaturesPerLevel;
    using ORB_SLAM3::ORBextractor::umax;
};

void put(orc_keypoint* o, const cv::KeyPoint& k) {
    o->x = k.pt.x;
    o->y = k.pt.y;
    o->size = k.size;
    o->angle = k.angle;
    o->response = k.response;
    o->octave = k.octave;
    o->class_id = k.class_id;
}
}    // namespace

extern "C" {

void* refext_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th) {
    return new Probe(nfeatures, scale_factor, nlevels, ini_th, min_th);
}
void refext_destroy(void* h) { delete (Probe*)h; }

void refext_tables(void* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2, int* features_per_level,
                   int* umax) {
    Probe* e = (Probe*)h;
    const int n = e->GetLevels();
    std::vector<float> a = e->GetScaleFactors(), b = e->GetInverseScaleFactors(), c = e->GetScaleSigmaSquares(),
                       d = e->GetInverseScaleSigmaSquares();
    for (int i = 0; i < n; ++i) {
        scale[i] = a[i];
        inv_scale[i] = b[i];
        sigma2[i] = c[i];
        inv_sigma2[i] = d[i];
        features_per_level[i] = e->mnFeaturesPerLevel[i];
    }
    for (int i = 0; i < 16; ++i) umax[i] = e->umax[i];
}

/* ORBextractor::operator()(image, cv::Mat(), keypoints, descriptors, vLappingArea) -> monoIndex */
int refext_extract(void* h, const uint8_t* img, int w, int hgt, int stride, int lap0, int lap1, orc_keypoint* kps,
                   uint8_t* desc, int cap, int* count) {
    Probe* e = (Probe*)h;
    cv::Mat image(hgt, w, CV_8UC1, (void*)img, (size_t)stride), descriptors;
    std::vector<cv::KeyPoint> keypoints;
    std::vector<int> lap = {lap0, lap1};
    const int mono = (*e)(image, cv::Mat(), keypoints, descriptors, lap);
    const int n = (int)keypoints.size();
    *count = n;
    for (int i = 0; i < n && i < cap; ++i) {
        put(kps + i, keypoints[i]);
        std::memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
    }
    return mono;
}

void refext_level_size(void* h, int level, int* w, int* hgt) {
    const cv::Mat& m = ((Probe*)h)->mvImagePyramid[level];
    *w = m.cols;
    *hgt = m.rows;
}
void refext_level_pixels(void* h, int level, uint8_t* dst, int dstride) {
    const cv::Mat& m = ((Probe*)h)->mvImagePyramid[level];
    for (int r = 0; r < m.rows; ++r) std::memcpy(dst + (size_t)r * dstride, m.ptr(r), (size_t)m.cols);
}
/* the (w+38) x (h+38) `temp` image that mvImagePyramid[level] is the interior of (ComputePyramid, EDGE_THRESHOLD 19) */
void refext_level_padded(void* h, int level, uint8_t* dst, int dstride) {
    const cv::Mat& m = ((Probe*)h)->mvImagePyramid[level];
    const size_t step = m.step1();
    const uint8_t* base = m.data - 19 * step - 19;
    for (int r = 0; r < m.rows + 38; ++r) std::memcpy(dst + (size_t)r * dstride, base + (size_t)r * step, (size_t)m.cols + 38);
}
/* ComputeKeyPointsOctTree on the pyramid of the last refext_extract: the level's keypoints after the quadtree and
 * computeOrientation, level coordinates */
int refext_level_keypoints(void* h, int level, orc_keypoint* out, int cap) {
    std::vector<std::vector<cv::KeyPoint> > all;
    ((Probe*)h)->ComputeKeyPointsOctTree(all);
    const int n = (int)all[level].size();
    for (int i = 0; i < n && i < cap; ++i) put(out + i, all[level][i]);
    return n;
}
/* DistributeOctTree on caller-supplied keys (x, y, response used) */
int refext_distribute(void* h, const orc_keypoint* keys, int n, int minX, int maxX, int minY, int maxY, int N,
                      int level, orc_keypoint* out, int cap) {
    std::vector<cv::KeyPoint> v((size_t)n);
    for (int i = 0; i < n; ++i) {
        v[i].pt.x = keys[i].x;
        v[i].pt.y = keys[i].y;
        v[i].size = keys[i].size;
        v[i].angle = keys[i].angle;
        v[i].response = keys[i].response;
        v[i].octave = keys[i].octave;
        v[i].class_id = keys[i].class_id;
    }
    std::vector<cv::KeyPoint> r = ((Probe*)h)->DistributeOctTree(v, minX, maxX, minY, maxY, N, level);
    const int m = (int)r.size();
    for (int i = 0; i < m && i < cap; ++i) put(out + i, r[i]);
    return m;
}
}
