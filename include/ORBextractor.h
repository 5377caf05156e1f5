/* ORBextractor.h — header-compatible ORB_SLAM3::ORBextractor on top of liborbfe.so (include/orbfe_c.h).
 *
 * Drop-in for the class Tracking / Frame link against in the reference:
 *   app/src/main/cpp/include/ORBextractor.h:33-108      (class, ctor, getters, mvImagePyramid, Lynx stereo operator())
 *   app/src/main/cpp/include/ORBextractor_old.h:59-61   (classic operator()(image, mask, keypoints, descriptors, lapping))
 *   app/src/main/cpp/src/ORBextractor_old.cc:1088-1191  (semantics reproduced bit-exactly by the CUDA path)
 * Same namespace, class name, constructor arguments, getters and call operators.  The AHardwareBuffer* overload of
 * the Lynx build becomes a side-by-side 8-bit image pointer (left eye = columns [0, width), right eye = columns
 * [width, 2 width)); an AHardwareBuffer-typed overload would lock the buffer and forward to it.
 *
 * With OpenCV headers available define ORBFE_WITH_OPENCV and the real cv::KeyPoint / cv::Mat are used (cv::KeyPoint
 * is layout-identical to orbfe_keypoint: 28 bytes).  Without them (this repository's build image has no OpenCV C++
 * SDK) minimal stand-ins with the same member names are declared below so that the class can be compiled and tested.
 * Header-only: everything forwards to the C ABI; there is no CPU fallback (no CUDA device => the constructor throws).
 */
#ifndef ORBFE_ORBEXTRACTOR_H
#define ORBFE_ORBEXTRACTOR_H

#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "orbfe_c.h"

#ifdef ORBFE_WITH_OPENCV
#include <opencv2/core.hpp>
#else
#ifndef CV_8U
#define CV_8U 0
#define CV_8UC1 0
#endif
namespace cv {
struct Point2f {
    float x = 0, y = 0;
};
struct KeyPoint {    // modules/core/include/opencv2/core/types.hpp: pt, size, angle, response, octave, class_id
    Point2f pt;
    float size = 0, angle = -1, response = 0;
    int octave = 0, class_id = -1;
};
class Mat {    // CV_8UC1 only: rows x cols bytes, `step` bytes per row
public:
    int rows = 0, cols = 0;
    size_t step = 0;
    unsigned char* data = nullptr;
    Mat() {}
    Mat(int r, int c, int /*type*/) { create(r, c, CV_8U); }
    Mat(int r, int c, int /*type*/, void* ext, size_t stp = 0) : rows(r), cols(c), step(stp ? stp : (size_t)c), data((unsigned char*)ext) {}
    void create(int r, int c, int /*type*/) {
        own_ = std::make_shared<std::vector<unsigned char>>((size_t)r * c);
        rows = r;
        cols = c;
        step = (size_t)c;
        data = own_->data();
    }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    unsigned char* ptr(int r) { return data + (size_t)r * step; }
    const unsigned char* ptr(int r) const { return data + (size_t)r * step; }
private:
    std::shared_ptr<std::vector<unsigned char>> own_;
};
typedef const Mat& InputArray;
typedef Mat& OutputArray;
}    // namespace cv
#endif

static_assert(sizeof(cv::KeyPoint) == sizeof(orbfe_keypoint), "cv::KeyPoint must be the 28-byte layout of orbfe_keypoint");

namespace ORB_SLAM3 {

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

    // ORBextractor_old.cc:411-471; `device` selects the GPU (one extractor = one stream + workspace)
    ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST, int device = 0)
        : nfeatures(nfeatures), scaleFactor(scaleFactor), nlevels(nlevels), iniThFAST(iniThFAST), minThFAST(minThFAST) {
        orbfe_params p{nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST};
        const int rc = orbfe_extractor_create(device, &p, &mH);
        if (rc != ORBFE_OK) throw std::runtime_error("ORBextractor: orbfe_extractor_create failed (" + std::to_string(rc) + "): no CUDA device or bad parameters");
        mvScaleFactor.resize(nlevels);
        mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels);
        mvInvLevelSigma2.resize(nlevels);
        mnFeaturesPerLevel.resize(nlevels);
        umax.resize(16);
        orbfe_extractor_tables(mH, mvScaleFactor.data(), mvInvScaleFactor.data(), mvLevelSigma2.data(),
                               mvInvLevelSigma2.data(), mnFeaturesPerLevel.data(), umax.data());
        mCap = orbfe_extractor_max_keypoints(mH);
        mvImagePyramid.resize(nlevels);
    }
    ~ORBextractor() {
        if (mH) orbfe_extractor_destroy(mH);
    }
    ORBextractor(const ORBextractor&) = delete;
    ORBextractor& operator=(const ORBextractor&) = delete;

    // Classic call (ORBextractor_old.h:59-61): returns monoIndex; the mask is ignored like in the reference (:56-58).
    int operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint>& _keypoints,
                   cv::OutputArray _descriptors, std::vector<int>& vLappingArea) {
        const cv::Mat& image = _image;
        if (image.empty()) return -1;    // ORBextractor_old.cc:1091-1092
        const int lap0 = vLappingArea.size() > 0 ? vLappingArea[0] : 0, lap1 = vLappingArea.size() > 1 ? vLappingArea[1] : 0;
        int32_t n = 0, mono = 0;
        _keypoints.resize(mCap);
        mDescScratch.resize((size_t)mCap * 32);
        check(orbfe_extract(mH, image.data, image.cols, image.rows, (int)image.step, lap0, lap1,
                            reinterpret_cast<orbfe_keypoint*>(_keypoints.data()), mDescScratch.data(), mCap, &n, &mono));
        _keypoints.resize(n);
        _descriptors.create(n, 32, CV_8U);
        if (n > 0) std::memcpy(_descriptors.data, mDescScratch.data(), (size_t)n * 32);
        if (mbKeepPyramid) UpdateImagePyramid(0);
        return mono;
    }

    // Lynx stereo call (ORBextractor.h:52-57) on a side-by-side frame; both eyes go through one batched launch
    // sequence when their lapping areas agree.  Returns a frame id like the reference.
    int operator()(const uint8_t* sideBySide, int width, int height, int stride, std::vector<cv::KeyPoint>& leftKeypoints,
                   cv::OutputArray leftDescriptors, std::vector<int>& vLappingAreaLeft,
                   std::vector<cv::KeyPoint>& rightKeypoints, cv::OutputArray rightDescriptors,
                   std::vector<int>& vLappingAreaRight, int& monoLeft, int& monoRight) {
        if (!sideBySide || width <= 0 || height <= 0 || stride < 2 * width) throw std::invalid_argument("ORBextractor: bad stereo frame");
        mEyes.resize((size_t)2 * width * height);
        for (int y = 0; y < height; y++) {
            std::memcpy(&mEyes[(size_t)y * width], sideBySide + (size_t)y * stride, width);
            std::memcpy(&mEyes[(size_t)(height + y) * width], sideBySide + (size_t)y * stride + width, width);
        }
        std::vector<cv::KeyPoint>* K[2] = {&leftKeypoints, &rightKeypoints};
        cv::Mat* D[2] = {&leftDescriptors, &rightDescriptors};
        std::vector<int>* lap[2] = {&vLappingAreaLeft, &vLappingAreaRight};
        int* monoOut[2] = {&monoLeft, &monoRight};
        std::vector<orbfe_keypoint> kps((size_t)2 * mCap);
        mDescScratch.resize((size_t)2 * mCap * 32);
        int32_t counts[2] = {0, 0}, monos[2] = {0, 0};
        const size_t eye = (size_t)width * height;
        auto lapv = [&](int e, int i) { return (int)lap[e]->size() > i ? (*lap[e])[i] : 0; };
        if (lapv(0, 0) == lapv(1, 0) && lapv(0, 1) == lapv(1, 1)) {
            check(orbfe_extract_batch(mH, mEyes.data(), 2, width, height, width, eye, lapv(0, 0), lapv(0, 1), kps.data(),
                                      mDescScratch.data(), mCap, counts, monos, 0));
        } else {
            for (int e = 0; e < 2; e++)
                check(orbfe_extract_batch(mH, mEyes.data() + e * eye, 1, width, height, width, eye, lapv(e, 0), lapv(e, 1),
                                          kps.data() + (size_t)e * mCap, mDescScratch.data() + (size_t)e * mCap * 32, mCap,
                                          counts + e, monos + e, 0));
        }
        for (int e = 0; e < 2; e++) {
            K[e]->resize(counts[e]);
            if (counts[e] > 0) std::memcpy((void*)K[e]->data(), kps.data() + (size_t)e * mCap, (size_t)counts[e] * sizeof(orbfe_keypoint));
            D[e]->create(counts[e], 32, CV_8U);
            if (counts[e] > 0) std::memcpy(D[e]->data, mDescScratch.data() + (size_t)e * mCap * 32, (size_t)counts[e] * 32);
            *monoOut[e] = monos[e];
        }
        return ++mFrameId;
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    // mvImagePyramid[level] = the level's pixels (ROI of the padded plane) of image `index` of the last call; read
    // back from the device on demand (Frame::ComputeStereoMatches reads it, Frame.cc:834,924,934-939).
    std::vector<cv::Mat> mvImagePyramid;
    void UpdateImagePyramid(int index = 0) {
        for (int l = 0; l < nlevels; l++) {
            int w = 0, h = 0;
            check(orbfe_level_size(mH, l, &w, &h));
            if (mvImagePyramid[l].rows != h || mvImagePyramid[l].cols != w) mvImagePyramid[l].create(h, w, CV_8U);
            check(orbfe_level_read(mH, index, l, 0, mvImagePyramid[l].data, (int)mvImagePyramid[l].step));
        }
    }
    bool mbKeepPyramid = true;    // refresh mvImagePyramid after every classic call (one device->host copy per level)

    orbfe_extractor* handle() { return mH; }

protected:
    void check(int rc) {
        if (rc != ORBFE_OK) throw std::runtime_error(std::string("ORBextractor: ") + orbfe_last_error(mH));
    }

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<int> umax;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

    orbfe_extractor* mH = nullptr;
    int mCap = 0;
    int mFrameId = 0;
    std::vector<uint8_t> mDescScratch, mEyes;
};

}    // namespace ORB_SLAM3

#endif /* ORBFE_ORBEXTRACTOR_H */
