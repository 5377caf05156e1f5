/* orbfe_c.h — C ABI of liborbfe.so, the B200 (sm_100a) ORB front-end.
 *
 * This is the drop-in boundary.  It stands where the reference has its FastRPC stub
 * (app/src/main/cpp/inc/orbslam3.idl:14-22, expanded by QAIC to the skel signatures at
 * dsp/src/orbslam_dsp.cpp:866,931,567,1017) and the rpcmem / AHardwareBuffer plumbing of
 * LynxHardwareAccelerator.cpp:26-213.  Plain pointers and sizes only; no C++/torch types.
 *
 * Conventions:
 *   - every function returns an int status: 0 = ORBFE_OK, negative = error (never exit()s, never throws);
 *     the text of the last error of a handle is available through orbfe_last_error().
 *   - the caller owns every buffer; capacities are passed explicitly and the true counts are reported.
 *   - calls on one handle are serialised by an internal mutex; distinct handles are independent
 *     (one CUDA stream + one device workspace each).
 *   - buffers are host memory unless the corresponding ORBFE_*_DEVICE flag says otherwise.  Pinned host
 *     memory makes the copies asynchronous; pageable memory works but is slower.
 *   - there is no CPU fallback: without a CUDA device every entry point fails with ORBFE_ERR_CUDA.
 */
#ifndef ORBFE_C_H
#define ORBFE_C_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORBFE_OK 0
#define ORBFE_ERR_ARG (-1)      /* bad argument (null pointer, non-positive size, too many levels, ...) */
#define ORBFE_ERR_CAPACITY (-2) /* an output capacity was too small; counts hold the true sizes */
#define ORBFE_ERR_CUDA (-3)     /* CUDA runtime error; see orbfe_last_error() */
#define ORBFE_ERR_NCCL (-4)     /* NCCL error in the sharded matcher */
#define ORBFE_ERR_INTERNAL (-5) /* internal invariant violated (reported, never ignored) */

#define ORBFE_MAX_LEVELS 16

/* flags */
#define ORBFE_IN_DEVICE 1u  /* input pointers are device memory of the handle's GPU */
#define ORBFE_OUT_DEVICE 2u /* output pointers are device memory of the handle's GPU */

/* cv::KeyPoint memory layout (pt.x, pt.y, size, angle, response, octave, class_id) — 28 bytes. */
typedef struct orbfe_keypoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbfe_keypoint;

/* Arguments of ORB_SLAM3::ORBextractor::ORBextractor (include/ORBextractor.h:38-39, ORBextractor_old.cc:411-415). */
typedef struct orbfe_params {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;
    int32_t ini_th_fast;
    int32_t min_th_fast;
} orbfe_params;

typedef struct orbfe_extractor orbfe_extractor; /* opaque: params + stream + device workspace */
typedef struct orbfe_matcher orbfe_matcher;     /* opaque: stream + device workspace */

/* ---- library ---- */
const char* orbfe_version(void);
/* number of visible CUDA devices (0 when there is none; the library then cannot compute anything). */
int orbfe_device_count(void);
/* kernels launched by this library in this process so far (bench.py's gpu_launches). */
uint64_t orbfe_kernel_launches(void);

/* ---- extractor: replaces ORBextractor's ctor + LynxHardwareAccelerator ctor (LHA:26-93) ---- */
int orbfe_extractor_create(int device, const orbfe_params* params, orbfe_extractor** out);
int orbfe_extractor_destroy(orbfe_extractor* ex);
const char* orbfe_last_error(const orbfe_extractor* ex);
/* upper bound on device workspace (bytes) the extractor may hold; batches are processed in chunks that fit.
 * Default 24 GiB. */
int orbfe_extractor_set_workspace_limit(orbfe_extractor* ex, size_t bytes);
/* run this extractor's copies and kernels on the caller's CUDA stream (a cudaStream_t passed as void*; e.g. the
 * current PyTorch stream) instead of the private stream created by orbfe_extractor_create.  The caller keeps
 * ownership of the stream.  NULL restores the private stream. */
int orbfe_extractor_set_stream(orbfe_extractor* ex, void* cuda_stream);
/* ctor-derived tables (ORBextractor_old.cc:416-470; getters of include/ORBextractor.h:61-82).  Each array holds
 * nlevels entries (umax: 16); null pointers are skipped. */
int orbfe_extractor_tables(const orbfe_extractor* ex, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2,
                           int32_t* features_per_level, int32_t* umax);
/* max keypoints one image can produce with these params: the size to give `cap` below. */
int orbfe_extractor_max_keypoints(const orbfe_extractor* ex);

/* ORBextractor::operator()(image, mask, keypoints, descriptors, vLappingArea) (include/ORBextractor_old.h:59-61,
 * ORBextractor_old.cc:1088-1191) over a batch of n equally sized CV_8UC1 images.
 *   images       image i starts at images + i*image_stride bytes; rows are `stride` bytes apart
 *   lap0,lap1    vLappingArea; keypoints with lap0 <= x <= lap1 are packed from the back ("stereo"), the others
 *                from the front ("mono"), ORBextractor_old.cc:1171-1180
 *   kps, desc    image i writes kps[i*cap ...] and desc[(i*cap ...)*32]; cap = per-image capacity
 *   counts[i]    keypoints found in image i (true count, even when > cap -> ORBFE_ERR_CAPACITY)
 *   monos[i]     the operator's return value (monoIndex); -1 for an empty image
 */
int orbfe_extract_batch(orbfe_extractor* ex, const uint8_t* images, int n, int width, int height, int stride,
                        size_t image_stride, int lap0, int lap1, orbfe_keypoint* kps, uint8_t* desc, int cap,
                        int32_t* counts, int32_t* monos, unsigned flags);
/* single image convenience == ORBextractor::operator(); returns monoIndex through *mono. */
int orbfe_extract(orbfe_extractor* ex, const uint8_t* image, int width, int height, int stride, int lap0, int lap1,
                  orbfe_keypoint* kps, uint8_t* desc, int cap, int32_t* count, int32_t* mono);

/* Rectified stereo front-end: the two ORBextractor calls of Frame::Frame (Frame.cc:142-146, both with vLappingArea
 * (0,0)) followed by Frame::ComputeStereoMatches (Frame.cc:827-997) for npairs pairs, with mvImagePyramid staying on
 * the device between the two.
 *   images        2*npairs images, interleaved: image 2p = left, 2p+1 = right of pair p (strides as orbfe_extract_batch)
 *   kps/desc/counts   as orbfe_extract_batch, for the 2*npairs images (mvKeys / mvKeysRight, mDescriptors / ...Right)
 *   mb, mbf       stereo baseline in metres and baseline * fx (Frame members mb, mbf): minZ = mb, maxD = mbf / mb
 *   u_right/depth npairs x cap floats: mvuRight / mvDepth of the left keypoints of pair p at [p*cap ...), -1 = no match
 * ORBmatcher::TH_HIGH / TH_LOW are the reference's 100 / 50 (ORBmatcher.cc:35-36). */
int orbfe_extract_stereo_batch(orbfe_extractor* ex, const uint8_t* images, int npairs, int width, int height, int stride,
                               size_t image_stride, orbfe_keypoint* kps, uint8_t* desc, int cap, int32_t* counts,
                               float mb, float mbf, float* u_right, float* depth, unsigned flags);
/* The batched form of the Lynx wire call (orbslam3_extractFeatures: extraction of both eyes AND the left -> right matcher
 * in one round trip, dsp/src/orbslam_dsp.cpp:1017-1021): npairs stereo pairs, images interleaved L0, R0, L1, R1, ...;
 * ORBextractor::operator() on all 2 npairs images (vLappingArea = lap0, lap1) and, per pair, cv::BFMatcher(NORM_HAMMING)
 * .knnMatch(left descriptors, right descriptors, k = 2) (Frame.cc:45,1227) on the descriptors while they are still in
 * HBM.  kps / desc / counts / monos as orbfe_extract_batch (2 npairs images); knn_idx / knn_dist: [npairs][cap][2]
 * (best, second; -1 past the left count or when the right image has fewer than two keypoints).  With host buffers the copies
 * of stage k - 1 / k + 1 run under the kernels of stage k, the matches included. */
int orbfe_extract_match_batch(orbfe_extractor* ex, const uint8_t* images, int npairs, int width, int height, int stride,
                              size_t image_stride, int lap0, int lap1, orbfe_keypoint* kps, uint8_t* desc, int cap,
                              int32_t* counts, int32_t* monos, int32_t* knn_idx, int32_t* knn_dist, unsigned flags);
/* device time (ms) of the ComputeStereoMatches kernels of the last orbfe_extract_stereo_batch call (0 when the host
 * path was pipelined over sub-chunks and the stages were not timed individually) */
int orbfe_extractor_stereo_ms(const orbfe_extractor* ex, float* ms);

/* Stage outputs of image `index` of the LAST orbfe_extract* call on this extractor (parity tests; mvImagePyramid
 * readers such as Frame.cc:834,924).  All destinations are host memory. */
int orbfe_level_size(const orbfe_extractor* ex, int level, int* w, int* h);
/* which: 0 = level pixels w x h (mvImagePyramid[level] ROI), 1 = padded (w+38) x (h+38), 2 = blurred w x h */
int orbfe_level_read(orbfe_extractor* ex, int index, int level, int which, uint8_t* dst, int dstride);
/* pre-octree FAST candidates of a level as (x, y, response) int32 triples relative to (minBorderX, minBorderY),
 * in the reference's insertion order (ORBextractor_old.cc:865-870).  Returns the count (negative = error). */
int orbfe_level_candidates(orbfe_extractor* ex, int index, int level, int32_t* out_xyr, int cap);
/* post-octree keypoints of a level, level coordinates, with angle (ORBextractor_old.cc:879-897). */
int orbfe_level_keypoints(orbfe_extractor* ex, int index, int level, orbfe_keypoint* out, int cap);
/* Device-resident copies of the outputs of the LAST chunk of the last orbfe_extract* call (all of them when the batch
 * fitted one chunk, which is the normal case): keypoints and descriptors strided by *cap, int32 counts per image.
 * Valid until the next call on this extractor.  Lets the matcher consume descriptors without a host round trip:
 * orbfe_knn2_batch(m, desc, counts, cap, desc + P*cap*32, counts + P, cap, P, idx, dist, ORBFE_IN_DEVICE). */
int orbfe_extractor_device_outputs(orbfe_extractor* ex, const orbfe_keypoint** kps, const uint8_t** desc,
                                   const int32_t** counts, int* n, int* cap);
/* device time (ms) of each stage of the last call, measured with CUDA events:
 * [0] h2d, [1] pyramid, [2] fast, [3] octree, [4] blur, [5] angle+descriptors+pack, [6] d2h */
int orbfe_extractor_stage_ms(const orbfe_extractor* ex, float out[7]);
/* kernels launched by the last call on this extractor */
int orbfe_extractor_last_launches(const orbfe_extractor* ex);

/* ---- matcher: cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, matches, 2)  (Frame.cc:45,1227) ---- */
int orbfe_matcher_create(int device, orbfe_matcher** out);
int orbfe_matcher_destroy(orbfe_matcher* m);
const char* orbfe_matcher_last_error(const orbfe_matcher* m);
int orbfe_matcher_set_stream(orbfe_matcher* m, void* cuda_stream); /* as orbfe_extractor_set_stream */
/* q: nq x 32 bytes, t: nt x 32 bytes, row-major.  idx/dist: 2 entries per query row (best, second best);
 * equal distances keep the lower train index first (OpenCV's order); missing neighbours (nt < 2) are -1. */
int orbfe_knn2(orbfe_matcher* m, const uint8_t* q, int64_t nq, const uint8_t* t, int64_t nt, int32_t* idx,
               int32_t* dist, unsigned flags);
/* ---- kNN without any allocation.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The same tcgen05 kernel as orbfe_knn2 in its train-segmented form, whose scratch memory (partial top 2 per segment + one
 * ticket per query tile) lives in a file-scope `static __device__` array of 2 MiB + 512 B instead of coming from
 * cudaMallocAsync; every GPU that loads the library holds that array until exit.  The call allocates nothing (no
 * cudaMalloc / cudaFree / allocating library routine) and needs no handle: it runs on the CURRENT device, on the caller's
 * stream (a cudaStream_t as void*, NULL = the legacy default stream), and returns after enqueueing (it does not synchronise).
 *   q, t, idx, dist   DEVICE memory of the current device: nq x 32 and nt x 32 descriptor bytes in, 2 x int32 per query row
 *                     out (best, second best; -1 = missing), results identical to orbfe_knn2
 *   limits            nq <= ORBFE_KNN2_STATIC_MAX_NQ (beyond: ORBFE_ERR_CAPACITY and a message), nt < 2^31
 *   err / err_len     optional buffer for the text of an error (there is no handle to hold it)
 * All calls on a device share the one scratch array: calls that overlap on the GPU (two streams, two threads) RACE; serialise
 * them on one stream.  The first call on a device resolves the array's address (cudaGetSymbolAddress) and the kernel's
 * shared-memory opt-in: make it outside stream capture. */
#define ORBFE_KNN2_STATIC_MAX_NQ 16384
int orbfe_knn2_static_scratch(const uint8_t* q, int64_t nq, const uint8_t* t, int64_t nt, int32_t* idx, int32_t* dist,
                              void* cuda_stream, char* err, size_t err_len);
/* npairs independent problems: problem p uses rows [p*q_cap, p*q_cap + nq[p]) of q and [p*t_cap, p*t_cap + nt[p])
 * of t, and writes idx/dist rows [p*q_cap ...).  nq/nt are host arrays unless ORBFE_IN_DEVICE (then device int32). */
int orbfe_knn2_batch(orbfe_matcher* m, const uint8_t* q, const int32_t* nq, int q_cap, const uint8_t* t,
                     const int32_t* nt, int t_cap, int npairs, int32_t* idx, int32_t* dist, unsigned flags);
/* Lowe ratio test of Frame.cc:1230 on host arrays returned by orbfe_knn2: out[i] = idx[2i] if both neighbours exist
 * and (float)dist[2i] < (float)dist[2i+1] * ratio (evaluated in double, like the reference's `distance * 0.7`),
 * else -1.  Pure host code, no arithmetic of the hot path. */
int orbfe_ratio_test(const int32_t* idx, const int32_t* dist, int64_t nq, double ratio, int32_t* out);
/* ---- guided matcher: the frame grid + ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints)
 * (ORBmatcher.cc:44-143; Frame::AssignFeaturesToGrid / PosInGrid / GetFeaturesInArea, Frame.cc:405-436, 672-751;
 * ORBmatcher::DescriptorDistance, ORBmatcher.cc:2107-2123) for frames with Nleft == -1 (monocular, rectified stereo).
 * A map point as that loop reads it: mTrackProjX/Y, mTrackProjXR, mTrackViewCos, mTrackDepth, mnTrackScaleLevel,
 * GetDescriptor(); flags bit 0 = mbTrackInView, bit 1 = isBad(), bit 2 = Observations() > 0 (an assigned point then
 * blocks its feature for the later points of the call, ORBmatcher.cc:84-86 — the order dependence is reproduced). */
typedef struct orbfe_track_point {
    float proj_x, proj_y, proj_xr, view_cos, track_depth;
    int32_t level, flags;
    uint8_t desc[32];
} orbfe_track_point;
/*   keys_un/desc   mvKeysUn / mDescriptors of the frame (n features)
 *   u_right        mvuRight (n floats) or NULL for a monocular frame (no mTrackProjXR gate)
 *   occupied       n bytes, non-zero = F.mvpMapPoints[i] is set and has Observations() > 0 before the call; or NULL
 *   bounds         mnMinX, mnMinY, mnMaxX, mnMaxY;  scale_factors: mvScaleFactors (nlevels)
 *   nn_ratio       ORBmatcher::mfNNratio;  th / far_points / th_far: the call's arguments
 *   match[npts]    feature index assigned to each point (F.mvpMapPoints[match[i]] = point i) or -1
 *   owner[n]       (optional) index of the LAST point assigned to each feature or -1;  *nmatches: the return value */
int orbfe_search_by_projection(orbfe_matcher* m, const orbfe_keypoint* keys_un, const uint8_t* desc, const float* u_right,
                               const uint8_t* occupied, int n, const float bounds[4], const float* scale_factors,
                               int nlevels, const orbfe_track_point* pts, int npts, float th, int far_points, float th_far,
                               float nn_ratio, int32_t* match, int32_t* owner, int32_t* nmatches);
/* nframes independent frames, one CTA each: frame f uses features [f*cap, f*cap + n[f]) and points
 * [f*pts_cap, f*pts_cap + npts[f]); bounds / scale_factors / thresholds are shared (host memory always).  The other
 * arrays are host memory unless ORBFE_IN_DEVICE / ORBFE_OUT_DEVICE. */
int orbfe_search_by_projection_batch(orbfe_matcher* m, const orbfe_keypoint* keys_un, const uint8_t* desc,
                                     const float* u_right, const uint8_t* occupied, const int32_t* n, int cap,
                                     const float bounds[4], const float* scale_factors, int nlevels,
                                     const orbfe_track_point* pts, const int32_t* npts, int pts_cap, int nframes, float th,
                                     int far_points, float th_far, float nn_ratio, int32_t* match, int32_t* owner,
                                     int32_t* nmatches, unsigned flags);
/* The same call for stereo-camera frames (Nleft != -1, the Lynx fisheye rig): both branches of ORBmatcher.cc:44-214.
 * A map point also carries its right-camera projection (mTrackProjXR/YR, mTrackViewCosR, mnTrackScaleLevelR, -1 = none);
 * flags bit 3 = mbTrackInViewR. */
typedef struct orbfe_track_point_stereo {
    float proj_x, proj_y, proj_xr, proj_yr, view_cos, view_cos_r, track_depth;
    int32_t level, level_r, flags;
    uint8_t desc[32];
} orbfe_track_point_stereo;
/* Frame f: the n[f] features at [f*cap ...) are mvKeys (the first nleft[f]) followed by mvKeysRight, desc / occupied
 * likewise (= rows of mDescriptors / slots of F.mvpMapPoints); left_to_right[f*cap + i] = mvLeftToRightMatch[i]
 * (i < nleft[f]), right_to_left[f*cap + j] = mvRightToLeftMatch[j] (j < n[f] - nleft[f]), -1 = no partner.
 * match_left / match_right[f*pts_cap + i]: feature found for point i by the left / right search (right: index into the
 * right features) or -1; owner[f*cap + slot]: last point assigned to that slot of mvpMapPoints or -1; nmatches[f]: the
 * return value.  bounds / scale_factors are host memory; the rest follows ORBFE_IN_DEVICE / ORBFE_OUT_DEVICE. */
int orbfe_search_by_projection_stereo_batch(orbfe_matcher* m, const orbfe_keypoint* keys, const uint8_t* desc,
                                            const uint8_t* occupied, const int32_t* n, const int32_t* nleft,
                                            const int32_t* left_to_right, const int32_t* right_to_left, int cap,
                                            const float bounds[4], const float* scale_factors, int nlevels,
                                            const orbfe_track_point_stereo* pts, const int32_t* npts, int pts_cap,
                                            int nframes, float th, int far_points, float th_far, float nn_ratio,
                                            int32_t* match_left, int32_t* match_right, int32_t* owner, int32_t* nmatches,
                                            unsigned flags);
/* ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame &F, vector<MapPoint*> &vpMapPointMatches) (ORBmatcher.cc:224-425), the matcher
 * of Tracking::TrackReferenceKeyFrame (Tracking.cc:2788) and of relocalisation (Tracking.cc:3709); npairs (key frame, frame)
 * pairs per call.  The DBoW2::FeatureVector of each side (std::map<NodeId, vector<unsigned>>, computed by the vocabulary on
 * the CPU: out of scope) is passed as CSR: *_node_ids[pair*node_cap + j] ascending, *_node_start[pair*(node_cap+1) + j] the
 * offset of node j's features inside *_idx[pair*cap ...] (feature indices in the node's vector order), *_nnodes[pair].
 * Key frame: desc_kf rows = pKF->mDescriptors, valid_kf[i] = vpMapPointsKF[i] && !isBad(), angle_kf[i] = angle of the keypoint
 * :330-333 picks (mvKeysUn, or mvKeys / mvKeysRight with two cameras).  Frame: desc_f rows = F.mDescriptors (n_f[pair] of them),
 * angle_f[slot] = angle of F.mvKeys[slot] or F.mvKeysRight[slot - Nleft] (:337-340, :367-370), nleft_f[pair] = F.Nleft (NULL
 * or -1: monocular).  matches[pair*cap + i] = index of the key-frame feature whose MapPoint is vpMapPointMatches[i], -1 =
 * NULL, after the rotation check; nmatches[pair] = the return value.  TH_LOW = 50 (ORBmatcher.cc:36). */
int orbfe_search_by_bow_batch(orbfe_matcher* m, const uint8_t* desc_kf, const float* angle_kf, const uint8_t* valid_kf,
                              const int32_t* kf_node_ids, const int32_t* kf_node_start, const int32_t* kf_idx,
                              const int32_t* kf_nnodes, const uint8_t* desc_f, const float* angle_f, const int32_t* n_f,
                              const int32_t* nleft_f, const int32_t* f_node_ids, const int32_t* f_node_start,
                              const int32_t* f_idx, const int32_t* f_nnodes, int cap, int node_cap, int npairs,
                              float nn_ratio, int check_orientation, int32_t* matches, int32_t* nmatches, unsigned flags);
/* ORBmatcher::SearchByBoW(KeyFrame *pKF1, KeyFrame *pKF2, vector<MapPoint*> &vpMatches12) (ORBmatcher.cc:766-906): loop
 * detection and relocalisation between key frames.  Feature vectors as CSR like above, for both key frames.  valid1[i] /
 * valid2[i] = the MapPoint of feature i exists, is not bad, and (stereo-camera key frames) i < mvKeysUn.size() (:802-804,
 * :821-823); angle1 / angle2 = mvKeysUn[i].angle.  Accepted with bestDist1 < TH_LOW (strict) and the mfNNratio test; an
 * accepted KF2 feature enters vbMatched2.  matches12[pair*cap + i1] = idx2 (vpMatches12[i1] = vpMapPoints2[idx2]) or -1 after
 * the rotation check; every one of the cap entries is written; nmatches[pair] = the return value. */
int orbfe_search_by_bow_kf_batch(orbfe_matcher* m, const uint8_t* desc1, const float* angle1, const uint8_t* valid1,
                                 const int32_t* node_ids1, const int32_t* node_start1, const int32_t* idx1,
                                 const int32_t* nnodes1, const uint8_t* desc2, const float* angle2, const uint8_t* valid2,
                                 const int32_t* node_ids2, const int32_t* node_start2, const int32_t* idx2,
                                 const int32_t* nnodes2, int cap, int node_cap, int npairs, float nn_ratio, int check_orientation,
                                 int32_t* matches12, int32_t* nmatches, unsigned flags);
/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, vMatchedPairs, bOnlyStereo, bCoarse) (ORBmatcher.cc:908-1153) for key frames
 * with one camera (mpCamera2 == NULL: Pinhole::epipolarConstrain, Pinhole.cpp:102-124) or with coarse != 0 (any camera
 * model: the constraint is skipped, :1044).  keysN = the keypoint :973-975 / :1010-1012 picks for feature i (mvKeysUn);
 * freeN[i] = GetMapPoint(i) == NULL && (!bOnlyStereo || mvuRight[i] >= 0); stereoN[i] (nullable = 0) = mvuRight[i] >= 0;
 * f12[pair*9 + 3*r + c] = F12(r, c) = K1^-T [t12]x R12 K2^-1 (Pinhole.cpp:104-107; unread when coarse); epipole[pair*2 ...] =
 * pKF2->mpCamera->project(T2w * Cw1) (:930-932); scale_factors / level_sigma2 = pKF2's mvScaleFactors / mvLevelSigma2.
 * matches12[pair*cap + i1] = vMatches12[i1] after the rotation check (vMatchedPairs = the pairs (i1, matches12[i1]) with
 * matches12[i1] >= 0, in i1 order); nmatches[pair] = the return value.  Two-camera (KannalaBrandt8) key frames:
 * orbfe_search_for_triangulation_kb8_batch below. */
int orbfe_search_for_triangulation_batch(orbfe_matcher* m, const orbfe_keypoint* keys1, const uint8_t* desc1, const uint8_t* free1,
                                         const uint8_t* stereo1, const int32_t* node_ids1, const int32_t* node_start1,
                                         const int32_t* idx1, const int32_t* nnodes1, const orbfe_keypoint* keys2,
                                         const uint8_t* desc2, const uint8_t* free2, const uint8_t* stereo2,
                                         const int32_t* node_ids2, const int32_t* node_start2, const int32_t* idx2,
                                         const int32_t* nnodes2, int cap, int node_cap, int npairs, const float* f12,
                                         const float* epipole, const float* scale_factors, const float* level_sigma2,
                                         int nlevels, int coarse, int check_orientation, int32_t* matches12, int32_t* nmatches,
                                         unsigned flags);
/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:331-401) for npoints MapPoints per call.  desc = the rows that the
 * reference gathers into vDescriptors (observations in mObservations order, left index before right index, bad key frames
 * skipped), concatenated: point p owns rows [offsets[p], offsets[p + 1]), offsets[0] = 0, offsets[npoints] = total_rows.
 * best_idx[p] = BestIdx (row inside the point's block: the first row with the smallest median distance, median =
 * sorted[int(0.5 * (N - 1))] over the N distances of the row including its own 0), best_median[p] (nullable) = that median,
 * out_desc[p*32 ...] (nullable) = the chosen row = the new mDescriptor.  A point without rows, or with more than max_n rows,
 * gets best_idx = -1 and its out_desc row is left as it was (the reference returns before touching mDescriptor). */
int orbfe_distinctive_descriptors(orbfe_matcher* m, const uint8_t* desc, const int32_t* offsets, int npoints, int64_t total_rows,
                                  int max_n, int32_t* best_idx, int32_t* best_median, uint8_t* out_desc, unsigned flags);
/* ---- the per-frame tracking matcher: ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th,
 * bMono) with SearchByProjectionThread and the rotation histogram (ORBmatcher.cc:1683-1921), and
 * ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, sAlreadyFound, th, ORBdist) (:1923-2059), both with
 * ComputeThreeMaxima (:2061-2102).  The pose algebra stays with the caller (Sophus / the camera model): a MapPoint of the
 * last frame (of the key frame) arrives as
 *   u, v      CurrentFrame.mpCamera->project(Tcw * x3Dw)                       (:1735, :1949)
 *   ur, vr    mpCamera->project(GetRelativePoseTrl() * x3Dc), stereo-camera frames only (:1833-1834)
 *   invz      1 / x3Dc(2) (:1730); the mvuRight gate uses u - mbf * invz (:1786)
 *   angle     angle of the source keypoint: kpLF.angle (:1813-1819) / pKF->mvKeysUn[i].angle (:2011)
 *   level     nLastOctave (:1744-1746) / nPredictedLevel = pMP->PredictScale(dist3D, &CurrentFrame) (:1972)
 *   flags     bit 0: the point reaches the search (frame form: pMP && !mvbOutlier[i] && invzc >= 0; key-frame form: pMP &&
 *             !isBad() && !sAlreadyFound.count(pMP) && minDistance <= dist3D <= maxDistance); bit 2: Observations() > 0
 *   desc      pMP->GetDescriptor()
 * The image-bounds tests (:1737-1740), the window th * mvScaleFactors[level], the level range by direction, the grid
 * search, TH_HIGH / ORBdist, the rotation histogram and the removal of the losing bins run on the device. */
typedef struct orbfe_frame_point {
    float u, v, ur, vr, invz, angle;
    int32_t level, flags;
    uint8_t desc[32];
} orbfe_frame_point;
/* Frame f of the batch: n[f] slots at [f*cap ...), left features first; nleft == NULL or nleft[f] == n[f] is Nleft == -1
 * (keys = mvKeysUn, u_right = mvuRight (nullable) gates the candidates), nleft[f] < n[f] a stereo-camera frame (keys =
 * mvKeys then mvKeysRight; the right search runs on mGridRight).  occupied[slot] = CurrentFrame.mvpMapPoints[slot] &&
 * ->Observations() > 0 (nullable).  dir[f] (nullable = 0): 1 = bForward, 2 = bBackward (:1698-1699, from tlc(2) and mb).
 * match_left / match_right[f*pts_cap + i]: slot chosen for point i by the left / right search (right: relative to nleft;
 * match_right may be NULL when nleft is) or -1, AFTER the rotation check; owner[f*cap + slot] = the point that holds
 * CurrentFrame.mvpMapPoints[slot] at return (-1 = NULL); nmatches[f] = the return value. */
int orbfe_search_by_projection_frame_batch(orbfe_matcher* m, const orbfe_keypoint* keys, const uint8_t* desc,
                                           const float* u_right, const uint8_t* occupied, const int32_t* n,
                                           const int32_t* nleft, const int32_t* dir, int cap, const float bounds[4],
                                           const float* scale_factors, int nlevels, float mbf, const orbfe_frame_point* pts,
                                           const int32_t* npts, int pts_cap, int nframes, float th, int check_orientation,
                                           int32_t* match_left, int32_t* match_right, int32_t* owner, int32_t* nmatches,
                                           unsigned flags);
/* The key-frame form (relocalisation, Tracking.cc): Nleft == -1 frames; occupied[slot] = CurrentFrame.mvpMapPoints[slot]
 * != NULL; every accepted match takes its slot; levels nPredictedLevel +- 1; accepted at <= orb_dist. */
int orbfe_search_by_projection_keyframe_batch(orbfe_matcher* m, const orbfe_keypoint* keys_un, const uint8_t* desc,
                                              const uint8_t* occupied, const int32_t* n, int cap, const float bounds[4],
                                              const float* scale_factors, int nlevels, const orbfe_frame_point* pts,
                                              const int32_t* npts, int pts_cap, int nframes, float th, int orb_dist,
                                              int check_orientation, int32_t* match, int32_t* owner, int32_t* nmatches,
                                              unsigned flags);
/* ---- MapPoints projected into a KEY FRAME (loop closing, map merging, local mapping).  One inner loop, five callers:
 * KeyFrame::GetFeaturesInArea (KeyFrame.cc:706-750: the 64 x 48 grid copied from the Frame, cells ix outer / iy inner, no
 * level filter), features of level nPredictedLevel - 1 or nPredictedLevel, the first minimum of DescriptorDistance.  A point
 * is an orbfe_frame_point: u, v = pKF->mpCamera->project(Tcw * p3Dw) (or fx * x + cx by hand, :1527-1528), invz = 1 / z,
 * level = pMP->PredictScale(dist, pKF), desc = pMP->GetDescriptor(), flags bit 0 = the point passed every gate that needs
 * the map or the pose (pMP, !isBad(), not already found / not in the key frame, z >= 0, minDistance <= dist <= maxDistance,
 * PO.dot(Pn) >= 0.5 * dist where the caller has one); KeyFrame::IsInImage (x >= mnMinX && x < mnMaxX ..., :752-755) runs on
 * the device.  ur, vr, angle are not read.
 *
 * ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th, ratioHamming) (ORBmatcher.cc:428-533) and its twin that
 * also fills vpMatchedKF (:535-647): occupied[slot] = vpMatched[slot] != NULL before the call (nullable); an accepted point
 * (bestDist <= TH_LOW * ratioHamming, required < 256) takes its slot and later points skip it (order dependence
 * reproduced).  match[f*pts_cap + i] = bestIdx of point i or -1; owner[f*cap + slot] (nullable) = the point now in
 * vpMatched[slot] or -1 (a slot that was occupied before the call stays -1); nmatches[f] = the return value. */
int orbfe_search_by_projection_sim3_batch(orbfe_matcher* m, const orbfe_keypoint* keys_un, const uint8_t* desc,
                                          const uint8_t* occupied, const int32_t* n, int cap, const float bounds[4],
                                          const float* scale_factors, int nlevels, const orbfe_frame_point* pts,
                                          const int32_t* npts, int pts_cap, int nframes, float th, float ratio_hamming,
                                          int32_t* match, int32_t* owner, int32_t* nmatches, unsigned flags);
/* ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight) (ORBmatcher.cc:1155-1345; check_reprojection = 1: the chi-square gate of
 * :1274-1299 with u_right = mvuRight (nullable = all -1), inv_level_sigma2 = mvInvLevelSigma2, bf = mbf, and the
 * point's ur = u - bf * invz computed on the device) and ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)
 * (:1347-1462; check_reprojection = 0).  keys = mvKeysUn (Nleft == -1), or mvKeys / mvKeysRight with their descriptor rows
 * for the left / bRight call of a stereo-camera key frame (add NLeft to the returned index for bRight, :1302).  The search
 * does not depend on the map mutations of the loop, so every point is independent: best_idx[f*pts_cap + i] = bestIdx when
 * bestDist <= TH_LOW else -1, best_dist likewise (nullable); the caller applies Replace / AddObservation / AddMapPoint in
 * point order (:1316-1337, :1443-1457).  ncandidates[f] = number of points with a match = nFused when no point turns bad
 * inside the loop. */
int orbfe_fuse_batch(orbfe_matcher* m, const orbfe_keypoint* keys, const uint8_t* desc, const float* u_right, const int32_t* n,
                     int cap, const float bounds[4], const float* scale_factors, const float* inv_level_sigma2, int nlevels,
                     float bf, const orbfe_frame_point* pts, const int32_t* npts, int pts_cap, int nframes, float th,
                     int check_reprojection, int32_t* best_idx, int32_t* best_dist, int32_t* ncandidates, unsigned flags);
/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, S12, th) (ORBmatcher.cc:1464-1681).  pts1[f*cap + i1] = the MapPoint of
 * feature i1 of KF1 projected into KF2 with S21 (flags bit 0 = pMP && !vbAlreadyMatched1[i1] && !isBad() && z >= 0 &&
 * distance range), pts2[f*cap + i2] the reverse with S12; bounds1 / bounds2 = the image bounds of KF1 / KF2.  Both searches
 * accept at <= TH_HIGH = 100; matches12[f*cap + i1] = idx2 when vnMatch1[i1] == idx2 and vnMatch2[idx2] == i1 (:1666-1679:
 * vpMatches12[i1] = vpMapPoints2[idx2]), else -1; nfound[f] = the return value.  match1 / match2 (nullable) receive
 * vnMatch1 / vnMatch2. */
int orbfe_search_by_sim3_batch(orbfe_matcher* m, const orbfe_keypoint* keys_un1, const uint8_t* desc1, const int32_t* n1,
                               const orbfe_keypoint* keys_un2, const uint8_t* desc2, const int32_t* n2, int cap,
                               const float bounds1[4], const float bounds2[4], const float* scale_factors, int nlevels,
                               const orbfe_frame_point* pts1, const orbfe_frame_point* pts2, int nframes, float th,
                               int32_t* matches12, int32_t* match1, int32_t* match2, int32_t* nfound, unsigned flags);
/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize) (ORBmatcher.cc:649-764, with
 * ComputeThreeMaxima :2061-2102) for monocular frames, nframes independent (F1, F2) pairs, one CTA each.  Pair f:
 * keys1 / desc1 [f*cap, f*cap + n1[f]) = F1.mvKeysUn / mDescriptors, keys2 / desc2 / n2 likewise for F2; bounds =
 * F2's mnMinX, mnMinY, mnMaxX, mnMaxY (host memory).  prev_matched[f*cap + i] = vbPrevMatched[i] as (x, y), read as the
 * window centres and updated in place for the matched keypoints (:759-761); it follows ORBFE_IN_DEVICE.
 * matches12[f*cap + i] = vnMatches12[i] (index into F2 or -1; all cap slots are written); nmatches[f] = the return
 * value.  nn_ratio = mfNNratio, check_orientation = mbCheckOrientation (the constructor's arguments). */
int orbfe_search_for_initialization_batch(orbfe_matcher* m, const orbfe_keypoint* keys1, const uint8_t* desc1,
                                          const int32_t* n1, const orbfe_keypoint* keys2, const uint8_t* desc2,
                                          const int32_t* n2, int cap, int nframes, const float bounds[4], int window_size,
                                          float nn_ratio, int check_orientation, float* prev_matched, int32_t* matches12,
                                          int32_t* nmatches, unsigned flags);
/* The half of Frame::ComputeStereoFishEyeMatches that follows BFMatchORB (Frame.cc:1164-1204) with
 * KannalaBrandt8::TriangulateMatches (KannalaBrandt8.cpp:300-363: unproject both keypoints, parallax gate, 4x4 DLT, depth
 * signs, chi-square reprojection gates), nframes stereo-camera frames per call.
 *   rig            the two KannalaBrandt8 cameras and Frame::mRlr / mtlr / mvLevelSigma2 (host memory)
 *   keys_left / keys_right [f*cap ...)   mvKeys / mvKeysRight; n_left / n_right [f] = Nleft / Nright;
 *   mono_left / mono_right [f]           monoLeft / monoRight (NULL = 0): the stereo-area keypoints come after them
 *   knn_idx / knn_dist [f*knn_cap*2 ...) the k = 2 result of orbfe_knn2_batch (== BFMatchORB's indices / dist1 / dist2)
 *                  with the stereo-area LEFT descriptors as queries and the stereo-area RIGHT ones as train rows
 *   left_to_right / right_to_left / depth [f*cap ...)   mvLeftToRightMatch / mvRightToLeftMatch / mvDepth (-1 = none)
 *   points3d [f*cap*3 ...)               mvStereo3Dpoints (x, y, z; 0 where there is no match)
 *   status (optional) [f*cap ...)        1 matched, 0 not tried, -1 .. -5 = TriangulateMatches' return, -6 depth <= 1e-4
 *   nmatches [f]                         nMatches
 * Float arithmetic as the reference; the 4x4 null vector is solved in double (Jacobi) instead of Eigen's float
 * JacobiSVD: results agree with the reference build to its own float noise (<= 1e-3 relative on mvDepth). */
typedef struct orbfe_fisheye_rig {
    float cam1[8], cam2[8]; /* KannalaBrandt8::mvParameters (fx, fy, cx, cy, k0..k3) of mpCamera / mpCamera2 */
    float Rlr[9], tlr[3];   /* mRlr row-major, mtlr (Frame.cc:1115-1116) */
    float precision;        /* KannalaBrandt8::precision, 1e-6 */
    int32_t nlevels;
    float level_sigma2[ORBFE_MAX_LEVELS]; /* mvLevelSigma2 */
} orbfe_fisheye_rig;
int orbfe_stereo_fisheye_matches_batch(orbfe_matcher* m, const orbfe_keypoint* keys_left, const orbfe_keypoint* keys_right,
                                       const int32_t* n_left, const int32_t* n_right, const int32_t* mono_left,
                                       const int32_t* mono_right, int cap, const int32_t* knn_idx, const int32_t* knn_dist,
                                       int knn_cap, int nframes, const orbfe_fisheye_rig* rig, int32_t* left_to_right,
                                       int32_t* right_to_left, float* depth, float* points3d, int32_t* status,
                                       int32_t* nmatches, unsigned flags);
/* ORBmatcher::SearchForTriangulation for key frames of a TWO-CAMERA rig (mpCamera2 != NULL, the Lynx fisheye pair;
 * ORBmatcher.cc:908-1153 with the branch :1014-1043 and KannalaBrandt8::epipolarConstrain, KannalaBrandt8.cpp:210-214 =
 * TriangulateMatches(...) > 0.0001).  Features of a key frame = its left keypoints (mvKeys, nleftN[pair] of them) followed by
 * its right ones (mvKeysRight), descriptor rows likewise; feature vectors as CSR like orbfe_search_for_triangulation_batch;
 * freeN[i] = GetMapPoint(i) == NULL.  rig: cam1 / cam2 = the parameters of mpCamera / mpCamera2 and precision (Rlr, tlr and
 * the level table of the struct are not read).  t12[pair*48 + 12*c ...] = R12 (row-major, 9) and t12 (3) of the relative pose
 * c = 0: Tll = T1w * Tw2, 1: Tlr = T1w * Twr2, 2: Trl = Tr1w * Tw2, 3: Trr = Tr1w * Twr2 (:934-940); the pair of a candidate
 * picks c = 2 * (idx1 >= nleft1) + (idx2 >= nleft2) and the matching cameras (:1015-1042).  level_sigma2 = mvLevelSigma2 (both
 * key frames).  No epipole gate and no mvuRight in this mode (bStereo is false when mpCamera2 is set).  coarse != 0 skips the
 * constraint.  Results as orbfe_search_for_triangulation_batch.  The triangulation is the device code of
 * orbfe_stereo_fisheye_matches_batch: float as the reference, null vector in double (decisions equal the reference's away
 * from the gates). */
int orbfe_search_for_triangulation_kb8_batch(orbfe_matcher* m, const orbfe_keypoint* keys1, const uint8_t* desc1,
                                             const uint8_t* free1, const int32_t* nleft1, const int32_t* node_ids1,
                                             const int32_t* node_start1, const int32_t* idx1, const int32_t* nnodes1,
                                             const orbfe_keypoint* keys2, const uint8_t* desc2, const uint8_t* free2,
                                             const int32_t* nleft2, const int32_t* node_ids2, const int32_t* node_start2,
                                             const int32_t* idx2, const int32_t* nnodes2, int cap, int node_cap, int npairs,
                                             const orbfe_fisheye_rig* rig, const float* t12, const float* level_sigma2, int nlevels,
                                             int coarse, int check_orientation, int32_t* matches12, int32_t* nmatches,
                                             unsigned flags);
/* Frame::UndistortKeyPoints (Frame.cc:763-796): mvKeysUn = mvKeys with the positions passed through
 * cv::undistortPoints(pts, pts, K, mDistCoef, Mat(), mK) (OpenCV 4.x: 5 fixed-point iterations in double); K = (fx, fy,
 * cx, cy) of mK, dist_coef = k1, k2, p1, p2[, k3] (ndist 4 or 5, up to OpenCV's 14); dist_coef[0] == 0 copies the
 * keypoints (:765-769).  n keypoints of any number of frames of the same camera.  dist_coef is host memory. */
int orbfe_undistort_keypoints(orbfe_matcher* m, const orbfe_keypoint* keys, int64_t n, float fx, float fy, float cx,
                              float cy, const float* dist_coef, int ndist, orbfe_keypoint* keys_un, unsigned flags);
/* device time (ms) of the kernel(s) of the last call and their launch count */
int orbfe_matcher_last_ms(const orbfe_matcher* m, float* kernel_ms, int* launches);

/* ---- map-scale matcher over the GPUs of one box (BASELINE configs[4]) ------------------------------
 * One process per GPU.  Query rows are sharded: rank g owns rows [g*nq/G, (g+1)*nq/G) (orbfe_shard_range); the train
 * set is broadcast from `root` with NCCL, every rank matches its shard against the whole train set (global train
 * indices: the tie order of cv::BFMatcher is preserved), and the results are exchanged with NCCL so that every rank
 * ends with all nq_total rows.  NCCL is loaded at run time (libnccl.so.2 or $ORBFE_NCCL_LIB).  This replaces nothing
 * in the reference (its matcher is per frame, Frame.cc:1227); it is the partitioned form of the same knnMatch call. */
int orbfe_shard_range(int64_t n, int world, int rank, int64_t* begin, int64_t* count);
/* communicator plumbing: rank 0 calls orbfe_nccl_unique_id and ships the 128 bytes to the other ranks by any means
 * (bench.py: torch.distributed); every rank then creates its communicator (a ncclComm_t returned as void*). */
int orbfe_nccl_unique_id(char out[128]);
int orbfe_nccl_comm_create(const char id[128], int rank, int world, int device, void** comm);
int orbfe_nccl_comm_destroy(void* comm);
const char* orbfe_nccl_last_error(void);
/* All pointers are DEVICE memory of the matcher's GPU.  q_shard: this rank's query rows (count from
 * orbfe_shard_range); t: nt x 32 bytes, filled on `root`, overwritten by the broadcast elsewhere; idx_all / dist_all:
 * nq_total x 2 int32 on every rank.  world == 1 needs no communicator.  *ms_out: device time of broadcast + match +
 * exchange on this rank (CUDA events on the matcher's stream). */
int orbfe_knn2_sharded(orbfe_matcher* m, void* nccl_comm, int rank, int world, const uint8_t* q_shard,
                       int64_t nq_total, uint8_t* t, int64_t nt, int root, int32_t* idx_all, int32_t* dist_all,
                       float* ms_out);
/* device times (ms) of the last orbfe_knn2_sharded on this matcher: [0] whole call, [1] the kNN launches, [2] the train
 * broadcast as seen on its own stream (it overlaps the launches: pieces of the train set are matched as they land),
 * [3] the result exchange (pack + ncclAllGather + unpack).  [0] - [1] is the exposed communication. */
int orbfe_knn2_sharded_times(orbfe_matcher* m, float out[4]);

/* ---- measurement aid (bench.py "pipe_roofline") --------------------------------------------------------------
 * Measured issue rate, in warp instructions per second over the whole GPU, of one integer pipe of `device`:
 * ORBFE_PIPE_ALU (VABSDIFF4 chains; LOP3 / IADD3 / SHF / PRMT / VIMNMX issue on the same pipe) or ORBFE_PIPE_FMAHEAVY
 * (IDP.2A chains; IMAD / IDP.4A issue on the same pipe).  Replaces nothing in the reference. */
#define ORBFE_PIPE_ALU 0
#define ORBFE_PIPE_FMAHEAVY 1
int orbfe_ubench_pipe(int device, int pipe, double* warp_inst_per_s);

/* ---- wire-compatible entry points: the four skel functions of the FastRPC interface ----------------------
 * Same names, argument order and meaning as dsp/src/orbslam_dsp.cpp:866 (open), :931 (close), :1017-1021
 * (extractFeatures) and :567 (bfMatchStereo), so LynxHardwareAccelerator.cpp:62,116,146-150 links against
 * liborbfe.so instead of libcdsprpc + the generated stub.  Differences, all deliberate (SURVEY.md §8a "known
 * defects"): width/height/stride/threshold are honoured; the image is a side-by-side stereo frame (left eye =
 * columns [0,width), right eye = columns [width, 2*width), `stride` bytes per row); extraction follows the CPU
 * semantics of ORBextractor_old.cc with the parameters set by orbslam3_configure (default 1000,1.2,8,20,7; the
 * `threshold` argument, when > 0, replaces iniThFAST); angles are packed as (int8 sin*64)<<8 | (int8 cos*64)&0xff
 * (LHA:175-177); counts are the true counts (not count+1); matching is left(query) -> right(train) over the
 * stereo parts, indices/distances as int16 like the IDL. */
typedef uint64_t orbfe_remote_handle64;
int orbslam3_open(const char* uri, orbfe_remote_handle64* h);
int orbslam3_close(orbfe_remote_handle64 h);
int orbslam3_configure(orbfe_remote_handle64 h, const orbfe_params* params);
int orbslam3_extractFeatures(
    orbfe_remote_handle64 h, const unsigned char* image, int imageLen, int width, int height, int stride, int threshold,
    int vLappingAreaLeft0, int vLappingAreaLeft1, int vLappingAreaRight0, int vLappingAreaRight1, int* totalPointsLeft,
    int* leftKeypointsX, int leftKeypointsXLen, int* leftKeypointsY, int leftKeypointsYLen, int* leftKeypointsAngle,
    int leftKeypointsAngleLen, int* leftKeypointsLevel, int leftKeypointsLevelLen, uint8_t* leftKeypointsORB,
    int leftKeypointsORBLen, int* totalPointsRight, int* rightKeypointsX, int rightKeypointsXLen, int* rightKeypointsY,
    int rightKeypointsYLen, int* rightKeypointsAngle, int rightKeypointsAngleLen, int* rightKeypointsLevel,
    int rightKeypointsLevelLen, uint8_t* rightKeypointsORB, int rightKeypointsORBLen, int* monoLeft, int* monoRight,
    short* indices, int indicesLen, short* distances1, int distances1Len, short* distances2, int distances2Len);
int orbslam3_bfMatchStereo(orbfe_remote_handle64 h, const unsigned char* leftDescriptors, int leftDescriptorsLen,
                           const unsigned char* rightDescriptors, int rightDescriptorsLen, short* indices,
                           int indicesLen, short* distances1, int distances1Len, short* distances2, int distances2Len);

#ifdef __cplusplus
}
#endif
#endif /* ORBFE_C_H */
