/* orb_oracle.h — C interface of the CPU ORACLE (test infrastructure, NOT product code).
 *
 * Dependency-free C++ restatement of the reference's CPU ORB front-end
 * (app/src/main/cpp/src/ORBextractor_old.cc + the OpenCV 4.x primitives it calls) and of the
 * Frame.cc brute-force Hamming matcher.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this library.  The product (liborbfe.so) never does.
 *
 * Parity status: PINNED.  The reference ships no golden vectors for this path, so the
 * oracle is pinned against the reference itself run here: tests/test_oracle_ref.py compares it bit
 * for bit with oracle/_ref/libext_old.so, the reference's own ORBextractor_old.cc compiled from
 * /root/reference (oracle/ref_ext/make_ref_ext.py).  The restated OpenCV primitives are pinned
 * against the genuine OpenCV build in this container (cv2 4.13.0) by tests/test_oracle_vs_cv2.py
 * and tests/golden/primitives.npz.  Matcher-side call sites (ORBmatcher.cc / Frame.cc) cannot be
 * compiled here (g2o / DBoW2 / Sophus member types); they are pinned by literal Python
 * restatements (tests/test_search_rule.py) and by cv2.BFMatcher / cv2.norm / cv2.undistortPoints.
 */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* cv::KeyPoint memory layout (7 x 4 bytes). */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orc_keypoint;

/* ---- primitives (each restates one OpenCV call of the reference path) ---- */
/* cv::resize(..., INTER_LINEAR) on CV_8UC1 (ORBextractor_old.cc:1344). */
void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride,
                          uint8_t* dst, int dw, int dh, int dstride);
/* cv::copyMakeBorder(..., BORDER_REFLECT_101) (ORBextractor_old.cc:1346-1352). dst is (w+2b) x (h+2b). */
void orc_border_reflect101_u8(const uint8_t* src, int w, int h, int sstride,
                              uint8_t* dst, int dstride, int border);
/* cv::GaussianBlur(src, dst, Size(7,7), 2, 2, BORDER_REFLECT_101) on CV_8UC1 (ORBextractor_old.cc:1147). */
void orc_gaussian7x7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
/* cv::FAST(img, kps, threshold, true) TYPE_9_16 (ORBextractor_old.cc:828,845). out_xyr = triples
 * (x, y, response) in OpenCV output order; returns the number found (may exceed cap; only cap are written). */
int orc_fast9_nms(const uint8_t* img, int w, int h, int stride, int threshold, int* out_xyr, int cap);
/* cv::fastAtan2(y, x), degrees in [0,360). */
float orc_fast_atan2(float y, float x);
/* cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, k=2) (Frame.cc:45,1227).  idx/dist hold 2 entries per
 * query (best, second); missing neighbours (nt < 2) are idx = -1, dist = -1.  threads <= 1 runs single-threaded. */
void orc_knn2_hamming(const uint8_t* q, int64_t nq, const uint8_t* t, int64_t nt,
                      int32_t* idx, int32_t* dist, int threads);

/* ---- the extractor (ORBextractor_old.cc:411-471 ctor, :1088-1191 operator()) ---- */
typedef struct orc_extractor orc_extractor;
orc_extractor* orc_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th);
void orc_extractor_destroy(orc_extractor*);
/* ctor tables; out arrays hold nlevels entries (umax: 16). */
void orc_extractor_tables(const orc_extractor*, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                          int* features_per_level, int* umax);
/* operator()(image, mask, keypoints, descriptors, vLappingArea): returns monoIndex, or -1 for an empty image.
 * kps/desc must hold cap entries; *count receives the true keypoint count. */
int orc_extract(orc_extractor*, const uint8_t* img, int w, int h, int stride, int lap0, int lap1,
                orc_keypoint* kps, uint8_t* desc, int cap, int* count);
/* intermediates of the LAST orc_extract call on this object (for stage-by-stage parity). */
void orc_level_size(const orc_extractor*, int level, int* w, int* h);
void orc_level_pixels(const orc_extractor*, int level, uint8_t* dst, int dstride);          /* w x h ROI      */
void orc_level_padded(const orc_extractor*, int level, uint8_t* dst, int dstride);          /* (w+38)x(h+38)  */
void orc_level_blurred(const orc_extractor*, int level, uint8_t* dst, int dstride);         /* 0 if level had no kp */
int  orc_level_candidates(const orc_extractor*, int level, int* out_xyr, int cap);          /* pre-octree, rel. to (16,16) */
int  orc_level_keypoints(const orc_extractor*, int level, orc_keypoint* out, int cap);      /* post-octree + angle, level coords */
/* per-stage seconds accumulated since create: pyramid, fast, octree, angle, blur, desc, pack */
void orc_extractor_stage_seconds(const orc_extractor*, double out[7]);
/* number of compareNodes ties (equal count and UL.x, different node) seen by std::sort in the last call */
int64_t orc_extractor_sort_ties(const orc_extractor*);
/* ORBextractor::DistributeOctTree (ORBextractor_old.cc:557-781) alone, on caller-made keys (tests: stress inputs against
 * the reference build oracle/_ref/libext_old.so).  *ties (nullable) = compareNodes ties seen by std::sort. */
int orc_distribute_octtree(orc_extractor*, const orc_keypoint* keys, int n, int minX, int maxX, int minY, int maxY,
                           int N, orc_keypoint* out, int cap, int64_t* ties);

/* n images (contiguous, each h*stride bytes) over `threads` std::threads; outputs strided by cap.
 * returns 0.  counts[i], monos[i] per image. */
int orc_extract_batch(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th,
                      const uint8_t* imgs, int n, int w, int h, int stride, int lap0, int lap1,
                      orc_keypoint* kps, uint8_t* desc, int cap, int* counts, int* monos, int threads);

/* Frame::ComputeStereoMatches (Frame.cc:827-997) on the keypoints / descriptors of a rectified pair; eL / eR hold the
 * pyramids of the LAST orc_extract of the left / right image (mvImagePyramid) and the scale tables.  uRight/depth: nl
 * floats (-1 = no match, mvuRight / mvDepth); sad (optional): the 11x11 SAD of accepted matches before the median cut
 * (-1 = none).  Returns the number of places where the reference would have left its arrays (0 on extractor output). */
int orc_stereo_matches(const orc_extractor* eL, const orc_extractor* eR, const orc_keypoint* kl, const uint8_t* dl,
                       int nl, const orc_keypoint* kr, const uint8_t* dr, int nr, float mb, float mbf, float* uRight,
                       float* depth, int32_t* sad);

/* A map point as ORBmatcher::SearchByProjection sees it (MapPoint members mTrackProjX/Y, mTrackProjXR, mTrackViewCos,
 * mTrackDepth, mnTrackScaleLevel, GetDescriptor()); flags: 1 = mbTrackInView, 2 = isBad(), 4 = Observations() > 0. */
typedef struct {
    float proj_x, proj_y, proj_xr, view_cos, track_depth;
    int32_t level, flags;
    uint8_t desc[32];
} orc_track_point;
/* Frame::AssignFeaturesToGrid + GetFeaturesInArea (Frame.cc:405-436, 672-751) and ORBmatcher::SearchByProjection(F,
 * vpMapPoints, th, bFarPoints, thFarPoints) (ORBmatcher.cc:44-143) for Nleft == -1.  keysUn/desc/uRight/occupied: the
 * frame (uRight, occupied nullable); match[npts]: feature assigned to each point or -1; owner[n] (nullable): last
 * point assigned to each feature or -1.  Returns nmatches. */
int orc_search_by_projection(const orc_keypoint* keysUn, const uint8_t* desc, const float* uRight,
                             const uint8_t* occupied, int n, float mnMinX, float mnMinY, float mnMaxX, float mnMaxY,
                             const float* mvScaleFactors, const orc_track_point* pts, int npts, float th, int bFarPoints,
                             float thFarPoints, float mfNNratio, int32_t* match, int32_t* owner);

/* Frame::UndistortKeyPoints (Frame.cc:763-796): cv::undistortPoints(pts, pts, K, mDistCoef, Mat(), mK) on the keypoint
 * positions (5 iterations in double), everything else copied.  dist: k1, k2, p1, p2[, k3]; dist[0] == 0 copies. */
void orc_undistort_keypoints(const orc_keypoint* keys, int n, float fx, float fy, float cx, float cy, const float* dist,
                             int ndist, orc_keypoint* keysUn);

/* Stereo-camera frames (Nleft != -1): the map point also carries its right-camera projection (mTrackProjXR/YR,
 * mTrackViewCosR, mnTrackScaleLevelR); flags bit 3 = mbTrackInViewR. */
typedef struct {
    float proj_x, proj_y, proj_xr, proj_yr, view_cos, view_cos_r, track_depth;
    int32_t level, level_r, flags;
    uint8_t desc[32];
} orc_track_point_stereo;
/* ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) for Nleft != -1 (ORBmatcher.cc:44-214).
 * keys/desc/occupied: N features, the nleft left ones first; leftToRight[nleft] / rightToLeft[n - nleft]:
 * mvLeftToRightMatch / mvRightToLeftMatch (-1 = none).  matchL / matchR[npts]: feature found by the left / right
 * search (right: index into the right features) or -1; owner[n]: last point assigned to every slot of mvpMapPoints. */
int orc_search_by_projection_stereo(const orc_keypoint* keys, const uint8_t* desc, const uint8_t* occupied, int n,
                                    int nleft, const int32_t* leftToRight, const int32_t* rightToLeft, float mnMinX,
                                    float mnMinY, float mnMaxX, float mnMaxY, const float* mvScaleFactors,
                                    const orc_track_point_stereo* pts, int npts, float th, int bFarPoints,
                                    float thFarPoints, float mfNNratio, int32_t* matchL, int32_t* matchR, int32_t* owner);

/* A MapPoint of LastFrame / of a KeyFrame as the frame-to-frame searches see it: its projection into the current left
 * camera (u, v) and, for stereo-camera frames, into the right one (ur, vr); 1 / z in the current camera; the octave of
 * its keypoint in the last frame (or the predicted level for the KeyFrame form); the angle of that keypoint; flags: bit 0
 * the point reaches the search (all of the reference's `continue`s that need the pose are taken by the caller), bit 2
 * pMP->Observations() > 0. */
typedef struct {
    float u, v, ur, vr, invz, angle;
    int32_t level, flags;
    uint8_t desc[32];
} orc_frame_point;
/* ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) incl. SearchByProjectionThread and the rotation
 * histogram (ORBmatcher.cc:1683-1921).  keys / desc / occupied: the n slots of CurrentFrame, left first; nleft == n is
 * Nleft == -1 (keys = mvKeysUn, uRight = mvuRight applies), nleft < n a stereo-camera frame (keys = mvKeys then
 * mvKeysRight, uRight ignored).  dir: 0, 1 = bForward, 2 = bBackward.  matchL / matchR[npts]: slot chosen by the left /
 * right search (right: relative to nleft) or -1, after the rotation check; owner[n] = CurrentFrame.mvpMapPoints as point
 * indices (-1 = NULL).  Returns nmatches. */
int orc_search_by_projection_frame(const orc_keypoint* keys, const uint8_t* desc, const float* uRight,
                                   const uint8_t* occupied, int n, int nleft, float mnMinX, float mnMinY, float mnMaxX,
                                   float mnMaxY, const float* mvScaleFactors, float mbf, const orc_frame_point* pts,
                                   int npts, float th, int dir, int checkOrientation, int32_t* matchL, int32_t* matchR,
                                   int32_t* owner);
/* ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (ORBmatcher.cc:1923-2059), Nleft == -1.
 * occupied[i] = CurrentFrame.mvpMapPoints[i] != NULL. */
int orc_search_by_projection_keyframe(const orc_keypoint* keysUn, const uint8_t* desc, const uint8_t* occupied, int n,
                                      float mnMinX, float mnMinY, float mnMaxX, float mnMaxY, const float* mvScaleFactors,
                                      const orc_frame_point* pts, int npts, float th, int ORBdist, int checkOrientation,
                                      int32_t* match, int32_t* owner);

/* ---- MapPoints projected into a key frame.  KeyFrame::GetFeaturesInArea (KeyFrame.cc:706-750) + KeyFrame::IsInImage (:752-755).
 * A point = orc_frame_point (u, v, invz, level = nPredictedLevel, flags bit 0 = it reaches the projection, desc).
 * ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming) (ORBmatcher.cc:428-533; :535-647 is the same
 * loop).  occupied[i] = vpMatched[i] != NULL before the call; match[npts]; owner[n] = point placed in vpMatched[i] by this call
 * or -1.  Returns nmatches. */
int orc_search_by_projection_sim3(const orc_keypoint* keysUn, const uint8_t* desc, const uint8_t* occupied, int n, float mnMinX,
                                  float mnMinY, float mnMaxX, float mnMaxY, const float* mvScaleFactors,
                                  const orc_frame_point* pts, int npts, float th, float ratioHamming, int32_t* match,
                                  int32_t* owner);
/* The search of ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight) (ORBmatcher.cc:1155-1345, checkReprojection = 1) and of
 * ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1347-1462, checkReprojection = 0) without the map mutations:
 * bestIdx[i] = the feature the loop body reaches `if(bestDist<=TH_LOW)` with, else -1; bestDist likewise.  uRight = mvuRight
 * (NULL = all -1).  Returns the number of points with a match. */
int orc_fuse(const orc_keypoint* keys, const uint8_t* desc, const float* uRight, int n, float mnMinX, float mnMinY, float mnMaxX,
             float mnMaxY, const float* mvScaleFactors, const float* mvInvLevelSigma2, float bf, const orc_frame_point* pts,
             int npts, float th, int checkReprojection, int32_t* bestIdx, int32_t* bestDist);
/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, S12, th) (ORBmatcher.cc:1464-1681): pts1[n1] = KF1's MapPoints in KF2,
 * pts2[n2] = KF2's in KF1; bounds1 / bounds2 = mnMinX, mnMinY, mnMaxX, mnMaxY of KF1 / KF2.  matches12[n1] = idx2 or -1;
 * vnMatch1[n1] / vnMatch2[n2] nullable.  Returns nFound. */
int orc_search_by_sim3(const orc_keypoint* keysUn1, const uint8_t* desc1, int n1, const orc_keypoint* keysUn2,
                       const uint8_t* desc2, int n2, const float* bounds1, const float* bounds2, const float* mvScaleFactors,
                       const orc_frame_point* pts1, const orc_frame_point* pts2, float th, int32_t* matches12,
                       int32_t* vnMatch1, int32_t* vnMatch2);

/* ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vpMapPointMatches) (ORBmatcher.cc:224-425).  The DBoW2 feature vectors
 * arrive as CSR (node ids ascending, start[nnodes + 1], feature indices in vector order); validKF[i] = the key frame's map
 * point i exists and is not bad; angle*: angle of the keypoint the reference reads for that feature; nleftF = F.Nleft
 * (-1: monocular).  matches[nF]: key-frame feature whose MapPoint lands in vpMapPointMatches[i] or -1.  Returns nmatches. */
int orc_search_by_bow(const uint8_t* descKF, const float* angleKF, const uint8_t* validKF, int nKF, const int32_t* kfNodeIds,
                      const int32_t* kfNodeStart, const int32_t* kfIdx, int kfNodes, const uint8_t* descF,
                      const float* angleF, int nF, int nleftF, const int32_t* fNodeIds, const int32_t* fNodeStart,
                      const int32_t* fIdx, int fNodes, float mfNNratio, int checkOrientation, int32_t* matches);

/* ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (ORBmatcher.cc:766-906).  validN[i] = MapPoint i exists, is not bad and
 * (stereo-camera key frames) i < mvKeysUn.size().  matches12[n1] = idx2 or -1.  Returns nmatches. */
int orc_search_by_bow_kf(const uint8_t* desc1, const float* angle1, const uint8_t* valid1, int n1, const int32_t* nodeIds1,
                         const int32_t* nodeStart1, const int32_t* idx1, int nodes1, const uint8_t* desc2, const float* angle2,
                         const uint8_t* valid2, const int32_t* nodeIds2, const int32_t* nodeStart2, const int32_t* idx2,
                         int nodes2, float mfNNratio, int checkOrientation, int32_t* matches12);
/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, vMatchedPairs, bOnlyStereo, bCoarse) (ORBmatcher.cc:908-1153) for key frames
 * without a second camera, Pinhole::epipolarConstrain (Pinhole.cpp:102-124) with F12 given row-major.  freeN[i] = no MapPoint
 * (and a stereo measurement when bOnlyStereo); stereoN[i] = mvuRight[i] >= 0 (nullable).  matches12[n1] = vMatches12. */
int orc_search_for_triangulation(const orc_keypoint* keys1, const uint8_t* desc1, const uint8_t* free1, const uint8_t* stereo1,
                                 int n1, const int32_t* nodeIds1, const int32_t* nodeStart1, const int32_t* idx1, int nodes1,
                                 const orc_keypoint* keys2, const uint8_t* desc2, const uint8_t* free2, const uint8_t* stereo2,
                                 const int32_t* nodeIds2, const int32_t* nodeStart2, const int32_t* idx2, int nodes2,
                                 const float* F12, const float* ep, const float* mvScaleFactors, const float* mvLevelSigma2,
                                 int bCoarse, int checkOrientation, int32_t* matches12);

/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401) on one MapPoint's N gathered descriptors (N >= 1).
 * Returns BestIdx; *bestMedian (nullable) = BestMedian. */
int orc_distinctive_descriptor(const uint8_t* vDescriptors, int N, int32_t* bestMedian);

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (ORBmatcher.cc:649-764) for
 * monocular frames.  prevMatchedXY: n1 (x, y) pairs, updated in place (:759-761); vnMatches12: n1 ints. Returns nmatches. */
int orc_search_for_initialization(const orc_keypoint* keys1, const uint8_t* desc1, int n1, const orc_keypoint* keys2,
                                  const uint8_t* desc2, int n2, float mnMinX, float mnMinY, float mnMaxX, float mnMaxY,
                                  float* prevMatchedXY, int windowSize, float mfNNratio, int checkOrientation,
                                  int32_t* vnMatches12);

/* Frame::ComputeStereoFishEyeMatches after BFMatchORB (Frame.cc:1164-1204) + KannalaBrandt8::TriangulateMatches
 * (KannalaBrandt8.cpp:300-363).  knnIdx / knnDist: k = 2 rows for the nL - monoL stereo-area left keypoints (best right
 * index relative to monoR, distances).  status (optional, per left keypoint): 1 matched, 0 not tried, -1 .. -5 the
 * negative return of TriangulateMatches, -6 depth <= 0.0001.  slack (optional): how close the closest gate was to
 * flipping (absolute for cos / depths, relative for the chi-square gates) — lets a tolerance test tell a real
 * disagreement from a decision on the edge. */
typedef struct orc_fisheye_rig {
    float cam1[8], cam2[8]; /* KannalaBrandt8::mvParameters fx fy cx cy k0..k3 of mpCamera / mpCamera2 */
    float Rlr[9], tlr[3];   /* mRlr (row-major), mtlr */
    float precision;        /* KannalaBrandt8::precision (1e-6) */
    int32_t nlevels;
    float levelSigma2[16];  /* mvLevelSigma2 */
} orc_fisheye_rig;
/* test hook: replace the restated 4x4 null-vector solve (row-major A -> v) by an external one (oracle/ref_eigen) */
typedef void (*orc_null_vector4_fn)(const float* A, float* v);
void orc_set_null_vector4_hook(orc_null_vector4_fn fn);
int orc_stereo_fisheye_matches(const orc_keypoint* keysL, int nL, int monoL, const orc_keypoint* keysR, int nR, int monoR,
                               const int32_t* knnIdx, const int32_t* knnDist, const orc_fisheye_rig* rig,
                               int32_t* leftToRight, int32_t* rightToLeft, float* depth, float* p3d, int32_t* status,
                               float* slack);

/* ORBmatcher::SearchForTriangulation for two-camera (KannalaBrandt8) key frames: features = left then right keypoints,
 * rig->cam1 / cam2 = mpCamera / mpCamera2, T12[12 c ...] = R12 | t12 of Tll, Tlr, Trl, Trr; slack (optional, n1 floats). */
int orc_search_for_triangulation_kb8(const orc_keypoint* keys1, const uint8_t* desc1, const uint8_t* free1, int n1, int nleft1,
                                     const int32_t* nodeIds1, const int32_t* nodeStart1, const int32_t* idx1, int nodes1,
                                     const orc_keypoint* keys2, const uint8_t* desc2, const uint8_t* free2, int nleft2,
                                     const int32_t* nodeIds2, const int32_t* nodeStart2, const int32_t* idx2, int nodes2,
                                     const orc_fisheye_rig* rig, const float* T12, const float* mvLevelSigma2, int bCoarse,
                                     int checkOrientation, int32_t* matches12, float* slack);
float orc_kb_triangulate_matches(const orc_fisheye_rig* rig, float x1, float y1, float x2, float y2, float sigmaLevel, float unc);

int orc_hardware_threads(void);
#ifdef __cplusplus
}
#endif
#endif
