/* LynxHardwareAccelerator.h — header-compatible ORB_SLAM3::LynxHardwareAccelerator on top of liborbfe.so.
 *
 * Drop-in for the class Frame.cc / Tracking.cc call in the reference:
 *   app/src/main/cpp/include/LynxHardwareAcceleration/LynxHardwareAccelerator.h:31-93   (class, MATCHING_CACHE_SIZE,
 *       GetInstance, StoreInputBuffer, ExtractORB, BFMatchORB)
 *   app/src/main/cpp/src/LynxHardwareAcceleration/LynxHardwareAccelerator.cpp:40-118    (ctor / dtor: open the FastRPC
 *       session, rpcmem buffers of MAX_POINTS entries, three matching-cache slots)
 *   ...cpp:121-123 StoreInputBuffer, :133-204 ExtractORB, :206-213 BFMatchORB
 * Same namespace, class name, method names and argument lists.  The FastRPC stub calls (orbslam3_open / close /
 * extractFeatures, orbslam3.idl:14-22) go to the wire entry points of include/orbfe_c.h, rpcmem buffers become plain
 * host vectors.  Header-only; no CPU fallback (no CUDA device => the constructor throws where the reference exit(-1)s,
 * LHA.cpp:47,64).
 *
 * Deliberate differences (SURVEY.md §8a "known defects"):
 *   - frame size and extractor parameters are constructor arguments (the reference hard-codes 640 x 400, LHA.h:20-21,
 *     and the DSP hard-codes nfeatures / levels); defaults are the reference's;
 *   - totalPoints are true counts: the reference subtracts 1 (LHA.cpp:155-159) because the DSP reports count + 1;
 *   - BFMatchORB copies rows entries; the reference copies rows * sizeof(uint16_t) entries (LHA.cpp:210-212), of which
 *     Frame.cc:1166-1182 reads the first rows.
 * cv::KeyPoint / cv::Mat are OpenCV's with ORBFE_WITH_OPENCV, else the stand-ins of include/ORBextractor.h.
 */
#ifndef ORBFE_LYNXHARDWAREACCELERATOR_H
#define ORBFE_LYNXHARDWAREACCELERATOR_H

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "ORBextractor.h"    // cv:: types (or stand-ins) + orbfe_c.h

#ifndef DEFAULT_WIDTH
#define DEFAULT_WIDTH (640)
#define DEFAULT_HEIGHT (400)
#endif
#ifndef MAX_POINTS
#define MAX_POINTS 20000
#endif

namespace ORB_SLAM3 {

constexpr int MATCHING_CACHE_SIZE = 3;    // LHA.h:31: the next frame's extraction must not overwrite what the CPU thread still reads

class LynxHardwareAccelerator {
public:
    static LynxHardwareAccelerator& GetInstance() {    // LHA.h:38-41
        static LynxHardwareAccelerator instance;
        return instance;
    }

    explicit LynxHardwareAccelerator(int width = DEFAULT_WIDTH, int height = DEFAULT_HEIGHT, int nfeatures = 1000,
                                     float scaleFactor = 1.2f, int nlevels = 8, int iniThFAST = 20, int minThFAST = 7)
        : mWidth(width), mHeight(height), mThreshold(iniThFAST) {
        int rc = orbslam3_open("orbslam3", &mDSPHandle);    // LHA.cpp:62
        if (rc != ORBFE_OK) throw std::runtime_error("LynxHardwareAccelerator: orbslam3_open failed (" + std::to_string(rc) + "): no CUDA device");
        orbfe_params p{nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST};
        rc = orbslam3_configure(mDSPHandle, &p);
        if (rc != ORBFE_OK) {
            orbslam3_close(mDSPHandle);
            throw std::invalid_argument("LynxHardwareAccelerator: bad extractor parameters");
        }
        inputImageRPCMem.resize((size_t)mWidth * mHeight * 2);    // LHA.cpp:70
        for (std::vector<int32_t>* v : {&leftKeypointsX, &leftKeypointsY, &leftKeypointsAngle, &leftKeypointsLevels,
                                        &rightKeypointsX, &rightKeypointsY, &rightKeypointsAngle, &rightKeypointsLevels})
            v->resize(MAX_POINTS);    // LHA.cpp:72-83
        leftKeypointsORB.resize((size_t)MAX_POINTS * 32);
        rightKeypointsORB.resize((size_t)MAX_POINTS * 32);
        for (int i = 0; i < MATCHING_CACHE_SIZE; i++) {    // LHA.cpp:85-89
            matchingCacheIndices[i].resize(MAX_POINTS);
            matchingCacheDistances[i].resize(MAX_POINTS);
            matchingCacheDistances2[i].resize(MAX_POINTS);
        }
    }
    ~LynxHardwareAccelerator() {
        if (mDSPHandle) orbslam3_close(mDSPHandle);    // LHA.cpp:116
    }
    LynxHardwareAccelerator(const LynxHardwareAccelerator&) = delete;
    LynxHardwareAccelerator& operator=(const LynxHardwareAccelerator&) = delete;

    // LHA.cpp:121-123: the side-by-side frame, left eye in columns [0, width), right eye in [width, 2 width)
    void StoreInputBuffer(const uint8_t* frameData) { std::memcpy(inputImageRPCMem.data(), frameData, inputImageRPCMem.size()); }

    // LHA.cpp:133-204.  Returns the frame id that BFMatchORB takes.
    int ExtractORB(int& leftKeypointsCount, int& rightKeypointsCount, std::vector<cv::KeyPoint>& leftKeyPoints,
                   std::vector<cv::KeyPoint>& rightKeyPoints, cv::OutputArray leftDescriptors,
                   cv::OutputArray rightDescriptors, int vLappingLeft0, int vLappingLeft1, int vLappingRight0,
                   int vLappingRight1, int& monoLeft, int& monoRight) {
        mFrameCounter++;
        const int bufferId = mFrameCounter % MATCHING_CACHE_SIZE;    // LHA.cpp:136-137
        int32_t totalPointsLeft = 0, totalPointsRight = 0, monoLeftDSP = 0, monoRightDSP = 0;
        const int rc = orbslam3_extractFeatures(
            mDSPHandle, inputImageRPCMem.data(), (int)inputImageRPCMem.size(), mWidth, mHeight, mWidth * 2, mThreshold,
            vLappingLeft0, vLappingLeft1, vLappingRight0, vLappingRight1, &totalPointsLeft, leftKeypointsX.data(),
            MAX_POINTS, leftKeypointsY.data(), MAX_POINTS, leftKeypointsAngle.data(), MAX_POINTS,
            leftKeypointsLevels.data(), MAX_POINTS, leftKeypointsORB.data(), MAX_POINTS * 32, &totalPointsRight,
            rightKeypointsX.data(), MAX_POINTS, rightKeypointsY.data(), MAX_POINTS, rightKeypointsAngle.data(),
            MAX_POINTS, rightKeypointsLevels.data(), MAX_POINTS, rightKeypointsORB.data(), MAX_POINTS * 32, &monoLeftDSP,
            &monoRightDSP, matchingCacheIndices[bufferId].data(), MAX_POINTS, matchingCacheDistances[bufferId].data(),
            MAX_POINTS, matchingCacheDistances2[bufferId].data(), MAX_POINTS);    // LHA.cpp:146-150
        if (rc != ORBFE_OK) throw std::runtime_error("LynxHardwareAccelerator: orbslam3_extractFeatures failed (" + std::to_string(rc) + ")");
        monoLeft = monoLeftDSP;
        monoRight = monoRightDSP;
        leftKeypointsCount = totalPointsLeft;    // true counts (the reference: totalPoints - 1, LHA.cpp:155-156)
        rightKeypointsCount = totalPointsRight;
        const int finalPointsLeft = std::min(10000, totalPointsLeft), finalPointsRight = std::min(10000, totalPointsRight);
        convert(leftKeyPoints, leftKeypointsX, leftKeypointsY, leftKeypointsAngle, leftKeypointsLevels, finalPointsLeft);
        convert(rightKeyPoints, rightKeypointsX, rightKeypointsY, rightKeypointsAngle, rightKeypointsLevels, finalPointsRight);
        leftDescriptors.create(finalPointsLeft, 32, CV_8U);    // LHA.cpp:194-199: copies, the wire buffers are reused
        rightDescriptors.create(finalPointsRight, 32, CV_8U);
        if (finalPointsLeft) std::memcpy(leftDescriptors.data, leftKeypointsORB.data(), (size_t)finalPointsLeft * 32);
        if (finalPointsRight) std::memcpy(rightDescriptors.data, rightKeypointsORB.data(), (size_t)finalPointsRight * 32);
        return mFrameCounter;
    }

    // LHA.cpp:206-213, called as BFMatchORB(mnIdMatchingData, stereoDescRight, stereoDescLeft, ...) at Frame.cc:1164:
    // the cached left(query) -> right(train) stereo matches of frame nIdMatchingData; one entry per row of the SECOND matrix.
    void BFMatchORB(int nIdMatchingData, const cv::Mat& /*leftDescriptors*/, const cv::Mat& rightDescriptors,
                    std::vector<uint16_t>& indices, std::vector<uint16_t>& dist1, std::vector<uint16_t>& dist2) const {
        const int bufferId = (nIdMatchingData + 3) % MATCHING_CACHE_SIZE;
        const int n = std::min(rightDescriptors.rows, (int)MAX_POINTS);
        indices.assign(matchingCacheIndices[bufferId].begin(), matchingCacheIndices[bufferId].begin() + n);
        dist1.assign(matchingCacheDistances[bufferId].begin(), matchingCacheDistances[bufferId].begin() + n);
        dist2.assign(matchingCacheDistances2[bufferId].begin(), matchingCacheDistances2[bufferId].begin() + n);
    }

    orbfe_remote_handle64 handle() const { return mDSPHandle; }

private:
    static void convert(std::vector<cv::KeyPoint>& out, const std::vector<int32_t>& X, const std::vector<int32_t>& Y,
                        const std::vector<int32_t>& A, const std::vector<int32_t>& Lv, int n) {
        out = std::vector<cv::KeyPoint>();
        out.reserve(n);
        for (int i = 0; i < n; i++) {    // LHA.cpp:166-181
            cv::KeyPoint kp;
            kp.pt.x = (float)X[i];
            kp.pt.y = (float)Y[i];
            kp.octave = Lv[i];
            const int8_t c = (int8_t)(A[i] & 0xFF), s = (int8_t)((A[i] >> 8) & 0xFF);
            kp.angle = (float)(std::atan2(((float)s) / 64.0f, ((float)c) / 64.0f) * 180.0f / M_PI);
            out.push_back(kp);
        }
    }

    int mWidth, mHeight, mThreshold;
    orbfe_remote_handle64 mDSPHandle = 0;
    std::vector<uint8_t> inputImageRPCMem;
    std::vector<int32_t> leftKeypointsX, leftKeypointsY, leftKeypointsAngle, leftKeypointsLevels;
    std::vector<int32_t> rightKeypointsX, rightKeypointsY, rightKeypointsAngle, rightKeypointsLevels;
    std::vector<uint8_t> leftKeypointsORB, rightKeypointsORB;
    int mFrameCounter = 0;
    std::vector<int16_t> matchingCacheIndices[MATCHING_CACHE_SIZE], matchingCacheDistances[MATCHING_CACHE_SIZE],
        matchingCacheDistances2[MATCHING_CACHE_SIZE];
};

}    // namespace ORB_SLAM3

#endif /* ORBFE_LYNXHARDWAREACCELERATOR_H */
