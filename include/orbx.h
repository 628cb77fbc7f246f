/* orbx -- C-ABI of the B200-native ORB feature frontend (extract + Hamming searches).
 *
 * This is the drop-in boundary under ORB-SLAM3's feature frontend: plain C, plain pointers and sizes, no C++/torch/cv types.
 * Each entry point names the reference interface it replaces (file:line relative to wangyunpeng139/3_ORB_SLAM3_SelfNote).
 * The C++ adapters with the reference's own signatures (ORB_SLAM3::ORBextractor / ORBmatcher) live in
 * 3_orb_slam3_selfnote_b200/host/ and call only the functions declared here; see INTEGRATION.md.
 *
 * Conventions
 *   - every function returns an int status: ORBX_OK (0) or a negative error code; ORBX_EMPTY_IMAGE (-1) mirrors the
 *     reference's `return -1` on an empty input (src/ORBextractor.cc:1075-1076).  No exceptions cross the boundary.
 *   - outputs are caller-allocated; "d_" parameters are device pointers on the handle's device, all others host pointers.
 *   - a handle is not re-entrant (same rule as one ORBextractor object, src/Frame.cc:120-123); different handles may be
 *     used concurrently from different threads.  Each handle owns its stream(s) and workspaces.
 *   - there is NO CPU fallback: without a usable CUDA device every compute call fails with ORBX_ERR_CUDA.
 */
#ifndef ORBX_H_
#define ORBX_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_OK 0
#define ORBX_EMPTY_IMAGE (-1)  /* src/ORBextractor.cc:1075 */
#define ORBX_ERR_ARG (-2)
#define ORBX_ERR_CUDA (-3)
#define ORBX_ERR_CAPACITY (-4) /* an output or internal list would overflow; nothing is silently truncated */
#define ORBX_ERR_UNSUPPORTED (-5)

#define ORBX_TAPS_OCV346 0 /* 7-tap sigma-2 Gaussian of OpenCV 3.4.1-3.4.8: 18 34 49 55 49 34 18 (/256) -- the reference's pin */
#define ORBX_TAPS_OCV4 1   /* OpenCV >= 3.4.9 / 4.2 bit-exact kernel:        18 34 48 56 48 34 18 (/256) */

#define ORBX_MAX_LEVELS 16

/* Layout-compatible with cv::KeyPoint (pt.x, pt.y, size, angle, response, octave, class_id): 28 bytes. */
typedef struct orbx_keypoint {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} orbx_keypoint;

/* Constructor arguments of ORB_SLAM3::ORBextractor (include/ORBextractor.h:49-50) plus sizing of the device workspaces. */
typedef struct orbx_params {
  int nfeatures;
  float scale_factor;
  int nlevels;
  int ini_th_fast;
  int min_th_fast;
  int blur_taps;  /* ORBX_TAPS_* */
  int max_batch;  /* frames processed per device batch (>= 1); host batches of any size are chunked */
  int device;     /* CUDA device ordinal */
} orbx_params;

typedef struct orbx_extractor orbx_extractor;

/* ---- extractor ------------------------------------------------------------------------------------------------- */

/* ORBextractor::ORBextractor, src/ORBextractor.cc:408-468 (scale tables, per-level quotas, pattern, umax). */
int orbx_create(const orbx_params* params, orbx_extractor** out);
void orbx_destroy(orbx_extractor* h);

/* GetLevels / GetScaleFactor / GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares /
 * GetInverseScaleSigmaSquares, include/ORBextractor.h:61-82.  Arrays hold nlevels floats; any pointer may be NULL. */
int orbx_get_levels(const orbx_extractor* h);
float orbx_get_scale_factor(const orbx_extractor* h);
int orbx_get_scale_tables(const orbx_extractor* h, float* scale, float* inv_scale, float* sigma2, float* inv_sigma2);
/* mnFeaturesPerLevel (protected in the reference; exposed for tests). */
int orbx_get_features_per_level(const orbx_extractor* h, int* quota);
/* Upper bound of keypoints one frame can return (sum over levels of max(quota+3, 4*nIni)); size outputs with it. */
int orbx_max_keypoints(const orbx_extractor* h, int width, int height);

/* ORBextractor::operator()(image, mask, keypoints, descriptors, vLappingArea), src/ORBextractor.cc:1071-1184.
 * image: 8-bit single channel, `stride` bytes per row.  lap0/lap1 = vLappingArea[0..1].  The mask is ignored by the
 * reference (include/ORBextractor.h:56) and has no parameter here.  Writes *n keypoints (28 B each) and *n x 32 descriptor
 * bytes, *mono_index = the reference's return value.  Returns ORBX_EMPTY_IMAGE when image == NULL or width/height <= 0. */
int orbx_extract(orbx_extractor* h, const uint8_t* image, int width, int height, int stride, int lap0, int lap1,
                 orbx_keypoint* keypoints, uint8_t* descriptors, int capacity, int* n, int* mono_index);

/* Throughput form of the same call: `nframes` independent frames of one size, host buffers in and out (H2D and D2H copies
 * included; chunked by max_batch and double-buffered over two streams).  Frame f starts at image + f*frame_stride.
 * Per frame f: counts[f] keypoints at keypoints + f*capacity_per_frame (and descriptors + f*capacity_per_frame*32),
 * mono_index[f] = the reference's return value. */
int orbx_extract_batch_host(orbx_extractor* h, const uint8_t* images, int nframes, int width, int height, int stride,
                            size_t frame_stride, int lap0, int lap1, orbx_keypoint* keypoints, uint8_t* descriptors,
                            int capacity_per_frame, int* counts, int* mono_index);

/* Device-resident form: frames already in HBM (d_images, row pitch `pitch`, frame f at d_images + f*frame_stride), results
 * stay in handle-owned device buffers (see orbx_batch_results_device).  nframes <= max_batch.  Work is enqueued on
 * `stream` (a cudaStream_t; NULL = the handle's own stream) and NOT synchronised: the caller owns ordering. */
int orbx_extract_batch_device(orbx_extractor* h, const uint8_t* d_images, int nframes, int width, int height, int pitch,
                              size_t frame_stride, int lap0, int lap1, void* stream);
/* The same with the workspace slot named (0 or 1; orbx_extract_batch_device = slot 0).  A handle owns two complete sets of pyramid /
 * candidate / result buffers -- the host pipelines alternate them -- and consecutive device batches issued into alternating slots share
 * nothing, so whatever still reads batch i (orbx_stereo_match_batch on another stream, a download of its results) overlaps the extraction
 * of batch i+1.  The caller orders the reuse of a slot: batch i+2 after the readers of batch i (an event wait).  "Last batch" in the calls
 * below means the batch of the slot used last. */
int orbx_extract_batch_device_slot(orbx_extractor* h, int slot, const uint8_t* d_images, int nframes, int width, int height, int pitch,
                                   size_t frame_stride, int lap0, int lap1, void* stream);
/* Device views of the last batch: keypoints [nframes][kp_stride] (28 B), descriptors [nframes][kp_stride][32],
 * counts [nframes], mono_index [nframes], status word (0 = ok, else ORBX_ERR_CAPACITY raised by a kernel). */
int orbx_batch_results_device(orbx_extractor* h, const orbx_keypoint** d_keypoints, const uint8_t** d_descriptors,
                              const int** d_counts, const int** d_mono_index, const int** d_status, int* kp_stride);
/* Host copy of the last device batch (synchronises the device first): same layout as orbx_extract_batch_host's outputs.
 * Returns the status word raised by the kernels (ORBX_OK or ORBX_ERR_CAPACITY). */
int orbx_batch_results_host(orbx_extractor* h, orbx_keypoint* keypoints, uint8_t* descriptors, int capacity_per_frame,
                            int* counts, int* mono_index);
/* Profiling mode: when enabled, every batch records CUDA events between its stages on the launching stream;
 * orbx_stage_times then returns the per-stage device time (ms) of the last batch and the stage names (7 entries:
 * memset, pyramid, FAST/NMS, quadtree, blur, assemble, orientation+descriptor).  Returns the number of stages. */
int orbx_set_profiling(orbx_extractor* h, int enable);
int orbx_stage_times(orbx_extractor* h, float* ms, const char** names);
/* Number of kernel launches one orbx_extract_batch_device call enqueues for the current geometry (for bench accounting). */
int orbx_launches_per_batch(const orbx_extractor* h);

/* The public member mvImagePyramid (include/ORBextractor.h:84), read by Frame::ComputeStereoMatches (src/Frame.cc:908,998):
 * level `level` of frame `frame` of the last batch.  The host form copies the un-blurred level into dst (dst_stride bytes
 * per row; dst may be NULL to query the size only); the device form returns a pointer into the handle's pyramid slab.
 * The reference's 19-px replicated frame around each level is not materialised (nothing on the path reads it). */
int orbx_pyramid_level(orbx_extractor* h, int frame, int level, uint8_t* dst, int dst_stride, int* width, int* height);
int orbx_pyramid_level_device(orbx_extractor* h, int frame, int level, const uint8_t** d_ptr, int* width, int* height,
                              int* pitch);

/* Stage taps for parity tests (host copies of intermediate results of frame `frame` of the last batch).
 * candidates: the list ComputeKeyPointsOctTree hands to DistributeOctTree for `level` (src/ORBextractor.cc:845-859) as
 * (x, y, response) int triples relative to (minBorderX, minBorderY), in arbitrary order.  level_keypoints: the per-level
 * result after DistributeOctTree + orientation, in the reference's order, level coordinates. */
int orbx_debug_candidates(orbx_extractor* h, int frame, int level, int* xyr, int capacity, int* n);
int orbx_debug_level_keypoints(orbx_extractor* h, int frame, int level, orbx_keypoint* kps, int capacity, int* n);
int orbx_debug_blurred_level(orbx_extractor* h, int frame, int level, uint8_t* dst, int dst_stride);

/* ---- matcher --------------------------------------------------------------------------------------------------- */

/* ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:2463-2483 (host, 2 x 32 bytes). */
int orbx_hamming(const uint8_t* a, const uint8_t* b);

/* A matcher handle owns a stream and growable device workspaces.  Like ORBmatcher objects (stack objects at every call
 * site, used concurrently from Tracking / LocalMapping / LoopClosing) handles are independent; one handle is not re-entrant. */
typedef struct orbx_matcher orbx_matcher;
int orbx_matcher_create(int device, orbx_matcher** out);
void orbx_matcher_destroy(orbx_matcher* m);

/* The members of Frame the searches read (src/Frame.cc, include/Frame.h): N, mvKeysUn, mDescriptors, mvuRight (NULL for
 * monocular), mvScaleFactors, mnMinX/mnMinY/mnMaxX/mnMaxY.  The 64x48 grid (Frame::AssignFeaturesToGrid, Frame.cc:434-465)
 * is rebuilt on the device from these on every call. */
typedef struct orbx_frame_view {
  int n;
  const orbx_keypoint* keypoints;
  const uint8_t* descriptors;
  const float* u_right;
  const float* scale_factors;
  int nlevels;
  float min_x, min_y, max_x, max_y;
} orbx_frame_view;

/* n x DescriptorDistance(a[i], b[i]) on the device (host arrays in/out). */
int orbx_hamming_batch(orbx_matcher* m, const uint8_t* a, const uint8_t* b, int n, int* dist);

/* Frame::GetFeaturesInArea(x, y, r, minLevel, maxLevel), src/Frame.cc:744-813: indices in the reference's order
 * (grid column, grid row, insertion order).  *n = number found (may exceed capacity; only capacity are written). */
int orbx_features_in_area(orbx_matcher* m, const orbx_frame_view* F, float x, float y, float r, int min_level, int max_level,
                          int* indices, int capacity, int* n);

/* ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th, ...), src/ORBmatcher.cc:44-142 (the mbTrackInView /
 * Nleft == -1 branch).  Per map point i: in_view[i] = mbTrackInView && !isBad() && !(bFarPoints && far); proj_x/proj_y =
 * mTrackProjX/Y; level = mnTrackScaleLevel; view_cos = mTrackViewCos; mp_desc = GetDescriptor() (32 B); observed[i] =
 * Observations() > 0 (NULL: all observed); proj_xr = mTrackProjXR (read only when F->u_right != NULL).  kp_mp is
 * F.mvpMapPoints as indices, in/out: -1 none, -2 pre-existing map point with observations (blocks), -3 pre-existing
 * without observations; matched keypoints receive the map-point index.  *nmatches = the reference's return value. */
int orbx_search_by_projection_mappoints(orbx_matcher* m, const orbx_frame_view* F, int nmp, const uint8_t* in_view,
                                        const float* proj_x, const float* proj_y, const int* level, const float* view_cos,
                                        const uint8_t* mp_desc, const uint8_t* observed, const float* proj_xr, float th,
                                        float nnratio, int* kp_mp, int* nmatches);
/* The same search for `nframes` frames whose keypoints/descriptors are the device-resident results of the last batch of
 * extractor `ex` (mvKeysUn == mvKeys, bounds = the image): per-frame map-point arrays are [nframes][nmp] host arrays,
 * kp_mp is [nframes][capacity_per_frame] (in/out), nmatches [nframes]. */
int orbx_search_by_projection_mappoints_batch(orbx_matcher* m, orbx_extractor* ex, int nframes, int nmp, const uint8_t* in_view,
                                              const float* proj_x, const float* proj_y, const int* level, const float* view_cos,
                                              const uint8_t* mp_desc, const uint8_t* observed, float th, float nnratio,
                                              int* kp_mp, int capacity_per_frame, int* nmatches);

/* The same entry point on a fisheye-STEREO frame (F.Nleft != -1; Examples/Stereo/TUM_512.yaml-shaped rigs), src/ORBmatcher.cc:44-211:
 * per map point the left-camera pass (:63-141) over the left image's keypoints / grid, then the right-camera pass (:144-211)
 * over mvKeysRight / mGridRight; both claim F.mvpMapPoints, and a match also hands the map point to the keypoint's stereo
 * partner in the other image (mvLeftToRightMatch / mvRightToLeftMatch, :131-135, :196-200).  `left` = mvKeys + descriptor rows
 * [0, Nleft), `right` = mvKeysRight + rows [Nleft, N) (same bounds; u_right = NULL).  track_left = mbTrackInView, mTrackProjX/Y,
 * mnTrackScaleLevel, mTrackViewCos; track_right = mbTrackInViewR, mTrackProjXR/YR, mnTrackScaleLevelR (-1: no right pass),
 * mTrackViewCosR, as Frame::isInFrustumChecks wrote them (src/Frame.cc:1270-1343); in_view already carries the gates of :52-62
 * for both.  th scales the left window only (:70-71 vs :147-150).  kp_mp: [left->n + right->n], in/out, coded as above;
 * *nmatches counts partner assignments too, as the reference does. */
typedef struct orbx_mappoint_track {
  const uint8_t* in_view; const float* proj_x; const float* proj_y; const int* level; const float* view_cos;
} orbx_mappoint_track;
int orbx_search_by_projection_mappoints_stereo(orbx_matcher* m, const orbx_frame_view* left, const orbx_frame_view* right, int nmp,
                                               const orbx_mappoint_track* track_left, const orbx_mappoint_track* track_right,
                                               const uint8_t* mp_desc, const uint8_t* observed, const int* left_to_right,
                                               const int* right_to_left, float th, float nnratio, int* kp_mp, int* nmatches);

/* Window best-match search shared by ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)
 * (src/ORBmatcher.cc:2027-2289; mode 0 = levels [l-1,l+1], 1 = forward [l,inf), 2 = backward [0,l]; dist_th = TH_HIGH),
 * both SearchByProjection(KeyFrame*, Scw, vpPoints, [vpPointsKFs,] vpMatched, ...) overloads (:489-602, :604-720; mode 3 =
 * levels [l-1,l], any_mp_blocks = 1, dist_th = floor(TH_LOW*ratioHamming), check_orientation = 0, kp_mp = vpMatched) and
 * SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:2291-2413; mode 0, any_mp_blocks = 1,
 * dist_th = ORBdist).  The caller projects the points (pose algebra and camera model stay on the host, as in the
 * reference's loop prologue): valid[i] = point exists, is not an outlier / not already found, invzc >= 0 and projects
 * inside the image bounds; u,v = projection; level = nLastOctave / predicted level; ur = uv.x - mbf*invzc (read only when
 * F->u_right != NULL, may be NULL); angle = source keypoint angle.  Rotation-consistency filtering (30-bin histogram, three
 * maxima, src/ORBmatcher.cc:2416-2458) is applied when check_orientation != 0.
 * kp_mp (in/out): in -1 none / -2 blocking map point / -3 non-blocking map point; out >= 0 = index of the matched source
 * point, -1 = NULL (also for a -3 entry that was re-matched and then dropped by the rotation histogram: the reference writes
 * NULL there, :2270-2281), -2 / -3 = untouched. */
int orbx_search_by_projection_window(orbx_matcher* m, const orbx_frame_view* F, int n, const uint8_t* valid, const float* u,
                                     const float* v, const int* level, const float* ur, const float* angle,
                                     const uint8_t* src_desc, const uint8_t* observed, float th, int mode, int dist_th,
                                     int any_mp_blocks, int check_orientation, int* kp_mp, int* nmatches);

/* ---- the remaining ORBmatcher entry points ------------------------------------------------ */

/* The per-point search loop shared by ORBmatcher::Fuse(KeyFrame*, const vector<MapPoint*>&, th, bRight)
 * (src/ORBmatcher.cc:1524-1590), Fuse(KeyFrame*, Scw, ...) (:1728-1756) and both passes of SearchBySim3 (:1866-1895,
 * :1946-1975): KeyFrame::GetFeaturesInArea(u, v, radius) (src/KeyFrame.cc:796-838), octave in [level-1, level], optional
 * chi-square gate (gate != 0, Fuse only: e2 * inv_level_sigma2[octave] > 7.8 with a right coordinate, > 5.99 without; needs
 * F->u_right for stereo keyframes and ur = uv.x - bf*invz), strict argmin Hamming in the reference's traversal order.
 * Queries are independent.  best_idx[i] = -1 and best_dist[i] = 256 when nothing qualifies; the caller applies
 * bestDist <= TH_LOW / TH_HIGH and the MapPoint bookkeeping (Replace / AddObservation / the Sim3 agreement check). */
int orbx_search_window_best(orbx_matcher* m, const orbx_frame_view* F, int n, const uint8_t* valid, const float* u, const float* v,
                            const float* radius, const int* level, const float* ur, const uint8_t* desc, int gate,
                            const float* inv_level_sigma2, int* best_idx, int* best_dist);

/* DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned int>>, Thirdparty/DBoW2/DBoW2/FeatureVector.h) flattened:
 * node ids strictly ascending, node_start[nnodes + 1] offsets into indices. */
typedef struct orbx_featvec {
  int nnodes;
  const uint32_t* node_ids;
  const int* node_start;
  const uint32_t* indices;
} orbx_featvec;

/* Both ORBmatcher::SearchByBoW overloads: (KeyFrame*, Frame&, vpMapPointMatches) src/ORBmatcher.cc:273-475 (Nleft == -1
 * branch; max_dist = TH_LOW) and (KeyFrame*, KeyFrame*, vpMatches12) :839-979 (max_dist = TH_LOW - 1, the reference tests
 * "< TH_LOW").  Side 1 is walked node by node in the order of its lists, side 2 is claimed.  elig1[i] = feature i of side 1
 * has a good map point; elig2 (may be NULL) = static eligibility of side-2 features (KF-KF: has a good map point).
 * matches12[i] = claimed side-2 feature or -1, after the rotation filter; *nmatches = the reference's return value. */
int orbx_search_by_bow(orbx_matcher* m, int n1, const orbx_keypoint* kps1, const uint8_t* desc1, const orbx_featvec* fv1,
                       const uint8_t* elig1, int n2, const orbx_keypoint* kps2, const uint8_t* desc2, const orbx_featvec* fv2,
                       const uint8_t* elig2, int max_dist, float nnratio, int check_orientation, int* matches12, int* nmatches);

/* SearchByBoW(KeyFrame*, Frame&, ...) when the frame is a fisheye-stereo one (F.Nleft != -1), src/ORBmatcher.cc:338-431: side 2
 * holds N = Nleft + Nright features (keypoints mvKeys then mvKeysRight, descriptors as stacked in mDescriptors); within a node the
 * left-image candidates go through the best / second-best ratio test, the right-image ones keep their best only (the reference
 * voids that ratio test with "|| true") and only when the left best is within max_dist (:400 sits inside :373).  Both claim.
 * matches12_left[i] / matches12_right[i] = side-2 feature index (right ones >= n_left2) or -1, after the rotation filter. */
int orbx_search_by_bow_stereo(orbx_matcher* m, int n1, const orbx_keypoint* kps1, const uint8_t* desc1, const orbx_featvec* fv1,
                              const uint8_t* elig1, int n2, int n_left2, const orbx_keypoint* kps2, const uint8_t* desc2,
                              const orbx_featvec* fv2, int max_dist, float nnratio, int check_orientation, int* matches12_left,
                              int* matches12_right, int* nmatches);

/* ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo, bCoarse), src/ORBmatcher.cc:981-1222, for
 * pinhole keyframes without a second camera.  elig1[i] = !pKF1->GetMapPoint(i) && (!bOnlyStereo || stereo1[i]), elig2
 * likewise; stereo = mvuRight >= 0 (NULL = monocular).  F12 (row-major 3x3) is the matrix Pinhole::epipolarConstrain forms
 * (src/CameraModels/Pinhole.cpp:145-148: K1^-T [t12]x R12 K2^-1); ep = pKF2->mpCamera->project(R2w*Cw + t2w);
 * scale_factors2 / level_sigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2.  matches12[i] = index in pKF2 or -1. */
int orbx_search_for_triangulation(orbx_matcher* m, int n1, const orbx_keypoint* kps1, const uint8_t* desc1, const orbx_featvec* fv1,
                                  const uint8_t* elig1, const uint8_t* stereo1, int n2, const orbx_keypoint* kps2,
                                  const uint8_t* desc2, const orbx_featvec* fv2, const uint8_t* elig2, const uint8_t* stereo2,
                                  const float* F12, float epx, float epy, const float* scale_factors2,
                                  const float* level_sigma2_2, int nlevels, int coarse, int check_orientation, int* matches12,
                                  int* nmatches);

/* The same entry point for KannalaBrandt8 keyframes, where the gate of src/ORBmatcher.cc:1146 is
 * KannalaBrandt8::epipolarConstrain (src/CameraModels/KannalaBrandt8.cpp:244-247): the candidate pair is triangulated
 * (TriangulateMatches :343-418: unproject both keypoints, parallax test, the 4x4 homogeneous system solved by cv::SVD -- OpenCV's
 * one-sided Jacobi, restated on the device --, both depths positive, both reprojection errors within 5.991 sigma^2).
 * One record per (camera of pKF1, camera of pKF2) combination: R12 / t12 row-major as src/ORBmatcher.cc:1000-1019 forms them,
 * cam = fx fy cx cy k1 k2 k3 k4. */
typedef struct orbx_kb8_pair { float R12[9]; float t12[3]; float cam1[8]; float cam2[8]; } orbx_kb8_pair;
/* n_left1 = n_left2 = -1: one camera per keyframe (TUM-VI monocular), npairs = 1, kps = mvKeysUn, the epipole gate of
 * :1093-1101 applies.  Fisheye-stereo keyframes (mpCamera2 set, pKF->NLeft != -1): kps / desc = the left features followed by
 * the right ones (mvKeys then mvKeysRight), npairs = 4 in the order ll, lr, rl, rr (index 2 * bRight1 + bRight2, :1103-1140), no
 * feature counts as stereo and the epipole gate is off.  level_sigma2_1 / _2 = mvLevelSigma2 of pKF1 / pKF2. */
int orbx_search_for_triangulation_kb8(orbx_matcher* m, int n1, const orbx_keypoint* kps1, const uint8_t* desc1, const orbx_featvec* fv1,
                                      const uint8_t* elig1, const uint8_t* stereo1, int n_left1, int n2, const orbx_keypoint* kps2,
                                      const uint8_t* desc2, const orbx_featvec* fv2, const uint8_t* elig2, const uint8_t* stereo2,
                                      int n_left2, const orbx_kb8_pair* pairs, int npairs, float epx, float epy,
                                      const float* scale_factors2, const float* level_sigma2_1, const float* level_sigma2_2,
                                      int nlevels, int coarse, int check_orientation, int* matches12, int* nmatches);

/* KannalaBrandt8::TriangulateMatches (src/CameraModels/KannalaBrandt8.cpp:343-418) for n independent keypoint pairs: what
 * Frame::ComputeStereoFishEyeMatches does with every left-right match that survived orbx_knn_ratio_match (src/Frame.cc:1255-1262, pair =
 * {mRlr, mtlr, left camera, right camera}, sigma1 / sigma2 = mvLevelSigma2 of the two keypoints' octaves).  xy1 / xy2: n x 2 pixel
 * coordinates.  z1[i] = depth in camera 1 or -1 (the reference accepts z1 > 0.0001f); p3D[3i..] = the point when accepted, else 0. */
int orbx_kb8_triangulate_matches(orbx_matcher* m, const orbx_kb8_pair* pair, int n, const float* xy1, const float* xy2,
                                 const float* sigma1, const float* sigma2, float* z1, float* p3D);

/* ---- pure Hamming workloads on descriptors next to the path ---------------------------------- */

/* MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:350-436, for a batch of map points (LocalMapping calls it per
 * point in loops, src/LocalMapping.cc:539, :1011): the observations' descriptors of point p are rows start[p] .. start[p+1]-1
 * of desc (the vDescriptors the reference gathers at MapPoint.cc:366-382).  best_idx[p] = row (relative to start[p]) whose
 * median Hamming distance to the others is least, first row on ties; best_median[p] that median; -1 / INT_MAX when empty. */
int orbx_distinctive_descriptors(orbx_matcher* m, const uint8_t* desc, const int* start, int npoints, int* best_idx,
                                 int* best_median);

/* The DBoW2 vocabulary tree (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h, m_nodes) resident in HBM.  Node 0 is the root;
 * the children of node i are child_ids[child_start[i] .. child_start[i+1]) in stored order; node_desc is nnodes x 32 bytes
 * (FORB descriptors), node_weight the idf weights, node_word_id the word id of leaves (-1 otherwise).  ORBvoc (k = 10, L = 6,
 * 1 082 073 nodes) takes 52 MB. */
typedef struct orbx_vocabulary orbx_vocabulary;
int orbx_vocabulary_create(int device, int nnodes, int L, const int* child_start, const int* child_ids, const uint8_t* node_desc,
                           const double* node_weight, const int* node_word_id, orbx_vocabulary** out);
void orbx_vocabulary_destroy(orbx_vocabulary* v);
/* TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup), TemplatedVocabulary.h:1218-1259, for n features: the
 * tree descent of Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:828-835, levelsup = 4).  The caller folds the three
 * arrays into BowVector / FeatureVector in feature order (addWeight / addFeature / normalize, TemplatedVocabulary.h:1145-1195):
 * that part is a std::map insertion in a fixed order and stays on the host. */
int orbx_vocabulary_transform(orbx_matcher* m, const orbx_vocabulary* v, const uint8_t* desc, int n, int levelsup, int* word_id,
                              int* node_id, double* weight);

/* ---- searches that start from 3-D map points and a pose: the projection prologue runs on the device ---------------------------
 * ORBmatcher's Frame-Frame, Frame-KeyFrame, KeyFrame-Sim3, Fuse and SearchBySim3 entry points first transform every map point
 * into the target camera, project it, test bounds / distance range / viewing angle and pick a search radius and octave window
 * (src/ORBmatcher.cc:2068-2118, :2305-2342, :518-560, :1486-1525, :1844-1866), then run the window search.  orbx_projection
 * selects the reference's variant of each step; the 3x3 pose algebra in front of the loops (Rcw, tcw, Ow, the Sim3 pieces) is
 * done by the caller exactly as the reference's cv::Mat lines do (host/ORBmatcher.h has them).  Evaluation order and rounding
 * follow the reference's float / double code; KannalaBrandt8's atan2f / PredictScale's logf are (float) of the double function. */
typedef struct orbx_projection {
  float R[9], t[3];      /* x_c = R * x_w + t (row-major; cv::gemm's small-matrix float path) */
  float R2[9], t2[3];    /* chain != 0: x_c <- R2 * x_c + t2 (SearchBySim3: sR21 / t21 after R1w / t1w, :1845-1846) */
  int chain;
  float Ow[3];           /* camera centre (dist_mode 0) */
  float fx, fy, cx, cy;
  int camera_model;      /* ORBX_CAMERA_PINHOLE / ORBX_CAMERA_KB8 */
  float kb[4];
  int proj_mode;         /* 0: mpCamera->project(x_c); 1: u = fx * (x * invz) + cx (SearchBySim3 :1853-1858) */
  float min_x, max_x, min_y, max_y;
  int bounds_mode;       /* 0: Frame: u < min_x || u > max_x rejects (:2084-2087); 1: KeyFrame::IsInImage: min <= u < max;
                            2: no test (right-camera pass of the Frame-Frame search, :2189-2192) */
  int depth_mode;        /* 0: reject invz < 0 (:2077); 1: reject z < 0 (:524, :1488, :1848); 2: no test (:2310-2312) */
  int invz_mode;         /* 0: invz = (float)(1.0 / z) (:2075, :1852); 1: invz = 1 / z in float (Fuse :1493) */
  float bf;
  int want_xr;           /* xr = u - bf * invz (:2158, :1503): the mvuRight check of the window search / the Fuse gate */
  int range_mode;        /* 1: dist outside [0.8f * mfMinDistance, 1.2f * mfMaxDistance] rejects */
  int dist_mode;         /* 0: dist = |x_w - Ow|; 1: dist = |x_c| (SearchBySim3 :1862) */
  int normal_check;      /* 1: (x_w - Ow) . normal < 0.5 * dist rejects (:553, :1538) */
  int level_mode;        /* 0: level = octave[i] (:2089); 1: MapPoint::PredictScale(dist) */
  int window_mode;       /* octave window: 0 [l-1, l+1]; 1 [l, any) (bForward); 2 [0, l] (bBackward); 3 [l-1, l] (KeyFrame loops) */
  float th;              /* radius = th * scale_factors[level] */
  float log_scale_factor;
  int n_scale_levels;
} orbx_projection;
typedef struct orbx_projection_points {
  int n;
  const float* world_pos;      /* n x 3: MapPoint::GetWorldPos() */
  const float* normal;         /* n x 3 (normal_check) or NULL */
  const float* max_distance;   /* RAW mfMaxDistance / mfMinDistance (range_mode, level_mode 1) or NULL */
  const float* min_distance;
  const int* octave;           /* level_mode 0: the source keypoint's octave */
  const uint8_t* valid;        /* the point takes part at all (exists, !isBad, not an outlier, not already found ...) or NULL = all */
  const uint8_t* desc;         /* n x 32: MapPoint::GetDescriptor() */
} orbx_projection_points;
typedef struct orbx_projection_debug {   /* optional host outputs of the prologue, n entries each (any may be NULL) */
  uint8_t* valid; float* u; float* v; float* radius; int* level; float* xr;
} orbx_projection_debug;
/* The window search with sequential claiming and the rotation histogram (orbx_search_by_projection_window) behind the prologue:
 * SearchByProjection(Frame&, const Frame&, th, bMono), (Frame&, KeyFrame*, set, th, ORBdist) and both (KeyFrame*, Scw, ...). */
int orbx_search_projected_window(orbx_matcher* m, const orbx_frame_view* F, const orbx_projection* proj, const orbx_projection_points* pts,
                                 const float* angle, const uint8_t* observed, int dist_th, int any_mp_blocks, int check_orientation,
                                 int* kp_mp, int* nmatches, const orbx_projection_debug* dbg);
/* SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) when Cur is a fisheye-stereo frame (Nleft != -1), :2027-2289: per point
 * of the last frame the left-image search (proj_left, claims among left keypoints) and, only if that one got past its `continue`s
 * (projection valid, window not empty: :2077-2124), the right-image search (:2188-2255) of x3Dr = Rrl * x3Dc + trl -- proj_right =
 * proj_left with chain = 1, R2 / t2 = mTrl, depth_mode 2, bounds_mode 2 and, as the reference has it, the LEFT camera's model --
 * over mvKeysRight / mGridRight with its own claims.  Both feed ONE rotation histogram.  kp_mp: [left->n + right->n], coded as for
 * orbx_search_projected_window; any_mp_blocks = 0. */
int orbx_search_projected_window_stereo(orbx_matcher* m, const orbx_frame_view* left, const orbx_frame_view* right,
                                        const orbx_projection* proj_left, const orbx_projection* proj_right,
                                        const orbx_projection_points* pts, const float* angle, const uint8_t* observed, int dist_th,
                                        int check_orientation, int* kp_mp, int* nmatches);
/* The independent best-keypoint search (orbx_search_window_best) behind the prologue: Fuse x2 (gate != 0 for the first
 * overload's chi-square test) and both passes of SearchBySim3. */
int orbx_search_projected_best(orbx_matcher* m, const orbx_frame_view* F, const orbx_projection* proj, const orbx_projection_points* pts,
                               int gate, const float* inv_level_sigma2, int* best_idx, int* best_dist, const orbx_projection_debug* dbg);

/* ORBmatcher::SearchForInitialization(F1, F2, vbPrevMatched, vnMatches12, windowSize), src/ORBmatcher.cc:722-837.
 * prev_matched: F1->n x 2 floats (in/out), matches12: F1->n ints (out). */
int orbx_search_for_initialization(orbx_matcher* m, const orbx_frame_view* F1, const orbx_frame_view* F2, float* prev_matched,
                                   int window_size, float nnratio, int check_orientation, int* matches12, int* nmatches);

/* cv::BFMatcher(NORM_HAMMING).knnMatch(query, train, k = 2) as used by Frame::ComputeStereoFishEyeMatches
 * (src/Frame.cc:1246): idx/dist are nq x 2 (-1 where fewer than two train rows exist), ascending distance, lower train
 * index first on ties.  good (optional, nq bytes) = Lowe test of src/Frame.cc:1253 (two neighbours and d0 < 0.7*d1). */
int orbx_knn2(orbx_matcher* m, const uint8_t* query, int nq, const uint8_t* train, int nt, int* idx, int* dist, uint8_t* good);
/* Device building block of the all-pairs keyframe sweep: folds one train chunk (global index of its first row =
 * train_base) into the running best / second-best keys (dist << 32 | global index; init != 0 resets them).  d_exclude
 * (optional, nq x int2) = global train range to skip per query (its own keyframe).  Asynchronous on `stream`.
 * Two kernels fold, with identical results: the POPC kernel (thread = query) and, from 2^26 pairs per call on, the
 * tensor-core kernel (tcgen05 kind::i8 on the descriptors expanded to +-1 bytes; csrc/knn2_tc.cuh), which keeps the
 * expanded image of the current train chunk in a matcher-owned workspace: calls of ONE matcher on different streams must
 * be ordered by the caller.  orbx_matcher_set_knn2_kernel pins the choice (parity tests, A/B timing). */
#define ORBX_KNN2_AUTO 0
#define ORBX_KNN2_POPC 1
#define ORBX_KNN2_TENSOR 2
int orbx_matcher_set_knn2_kernel(orbx_matcher* m, int kind);
int orbx_knn2_chunk_device(orbx_matcher* m, const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, long long train_base,
                           const int* d_exclude, unsigned long long* d_key0, unsigned long long* d_key1, int init, void* stream);

/* All-pairs keyframe-descriptor 2-NN sweep (BASELINE.json config 5; SURVEY.md 8-e, the one step of the path with an exchange).
 * Every rank holds `rows` descriptors in HBM (rows / desc_per_kf keyframes; the same `rows` on every rank; global row =
 * rank * rows + local row) and gets, for each of them, the two nearest descriptors among all descriptors of all OTHER
 * keyframes of all ranks: idx[2*i+k] = global row, dist[2*i+k] = Hamming distance (k = 0 best, 1 second; -1 where fewer
 * exist), ties to the lower global row exactly like cv::BFMatcher.  `nccl_comm` is an ncclComm_t spanning the ranks (NULL =
 * single rank); the shards travel by ncclAllGather in `nchunks` pieces on a side stream while the previous piece is folded
 * by the 2-NN tile kernel.  idx / dist are HOST arrays of 2 * rows; the call blocks until they are written.
 * The communicator may come from the caller's own NCCL (same libnccl.so.2 in the process) or from the three helpers below,
 * which bind NCCL at run time (liborbx.so has no link-time NCCL dependency): rank 0 calls orbx_nccl_unique_id, ships the 128
 * bytes to the other ranks by any channel, every rank calls orbx_nccl_comm_init. */
typedef struct orbx_allpairs_stats {
  int world, rank, nchunks;
  double pairs;         /* descriptor pairs this rank evaluated */
  double total_ms;      /* device time of the sweep on this rank (first fold start to last fold end) */
  double fold_ms;       /* sum over chunks of the 2-NN folds */
  double gather_ms;     /* sum over chunks of the all-gathers (includes waiting for the slowest rank) */
  double gather_bytes;  /* bytes this rank received */
  double overlap_frac;  /* share of the all-gather time hidden behind folds */
} orbx_allpairs_stats;
int orbx_nccl_version(void);   /* e.g. 22809; 0 when libnccl.so.2 cannot be loaded */
int orbx_nccl_unique_id(uint8_t id[128]);
int orbx_nccl_comm_init(const uint8_t id[128], int nranks, int rank, int device, void** comm);
int orbx_nccl_comm_destroy(void* comm);
int orbx_allpairs_knn2(orbx_matcher* m, void* nccl_comm, const uint8_t* d_desc, long long rows, int desc_per_kf, int nchunks,
                       long long* idx, int* dist, orbx_allpairs_stats* stats);

/* Frame::ComputeStereoMatches, src/Frame.cc:901-1079, for every frame of the last batches of two extractor handles
 * (left / right image of each pair; same geometry and frame count; keypoints, descriptors and un-blurred pyramids are read
 * where they lie in HBM, exactly the members the reference reads at Frame.cc:908,998,1012).  mb / mbf as in Frame.
 * Asynchronous on `stream` (NULL = the matcher's stream); the call makes that stream wait on the device for the two extraction
 * batches wherever they were enqueued.  Results: per left keypoint mvuRight / mvDepth (-1 = no match). */
int orbx_stereo_match_batch(orbx_matcher* m, orbx_extractor* left, orbx_extractor* right, float mb, float mbf, void* stream);
int orbx_stereo_results_device(orbx_matcher* m, const float** d_u_right, const float** d_depth, int* stride);
/* Whole stereo frames from HOST buffers (the data-parallel part of Frame::Frame(stereo), src/Frame.cc:89-187): for each of
 * `nframes` rectified pairs, both extractions (vLappingArea {0,0}, Frame.cc:120-121) and ComputeStereoMatches; uploads,
 * kernels and downloads are chunked by max_batch and double-buffered.  Outputs per frame f at index f*capacity_per_frame.
 * keypoints_right and descriptors_right may both be NULL: the right image's records then stay on the device (counts_right is
 * still written).  After ComputeStereoMatches a RECTIFIED stereo Frame only reads mvuRight / mvDepth of the right image
 * (src/Frame.cc:141-187); mvKeysRight / mDescriptorsRight are consumed later only by fisheye rigs (Nleft != -1). */
int orbx_stereo_frames_host(orbx_matcher* m, orbx_extractor* left, orbx_extractor* right, const uint8_t* images_left,
                            const uint8_t* images_right, int nframes, int width, int height, int stride, size_t frame_stride,
                            float mb, float mbf, orbx_keypoint* keypoints_left, uint8_t* descriptors_left, int* counts_left,
                            orbx_keypoint* keypoints_right, uint8_t* descriptors_right, int* counts_right, float* u_right,
                            float* depth, int capacity_per_frame);
/* Host copy (synchronises): [nframes][capacity_per_frame]; best_dist / best_idx expose the Hamming stage (may be NULL). */
int orbx_stereo_results_host(orbx_matcher* m, float* u_right, float* depth, int* best_dist, int* best_idx,
                             int capacity_per_frame);

/* ---- pre-processing in front of the extractor ---------------------------------------------------
 * The 8-bit image operations the reference applies to a frame before ORBextractor::operator() sees it, bit-exact with the
 * OpenCV calls they replace.  Batched: nframes images, frame f at base + f * frame_stride, rows `stride` bytes apart.
 * mem = ORBX_MEM_HOST: host pointers, the call stages through the handle's device buffers and blocks until dst is written.
 * mem = ORBX_MEM_DEVICE: device pointers, the kernels are enqueued on `stream` (NULL = the handle's own stream) and the
 * call returns at once -- the form that feeds orbx_extract_batch_device without leaving HBM.
 * A handle is not re-entrant; handles are independent. */
#define ORBX_MEM_HOST 0
#define ORBX_MEM_DEVICE 1
#define ORBX_GRAY_OCV346 0 /* 14-bit weights R 4899 G 9617 B 1868 (OpenCV 3.4's table path; unverifiable here) */
#define ORBX_GRAY_OCV4 1   /* 15-bit weights R 9798 G 19235 B 3735 (verified bit-exact against cv2 4.13) */
typedef struct orbx_preproc orbx_preproc;
int orbx_preproc_create(int device, orbx_preproc** out);
void orbx_preproc_destroy(orbx_preproc* p);

/* cvtColor(im, im, CV_{RGB,BGR,RGBA,BGRA}2GRAY): src/Tracking.cc:989-1014 (GrabImageStereo), :1059-1073 (GrabImageRGBD), :1121-1135
 * (GrabImageMonocular).  channels 3 or 4; rgb_order != 0 when channel 0 is red (mbRGB).  src and dst must not overlap. */
int orbx_cvt_gray(orbx_preproc* p, const uint8_t* src, int nframes, int width, int height, int src_stride, long long src_frame_stride,
                  int channels, int rgb_order, int coeffs, uint8_t* dst, int dst_stride, long long dst_frame_stride, int mem, void* stream);

/* cv::remap(im, imRect, M1, M2, cv::INTER_LINEAR): Examples/Stereo/stereo_euroc.cc:166-167, maps (CV_32F, host) from
 * cv::initUndistortRectifyMap at :117-118.  The table is built once per camera: OpenCV's 1/32-px fixed-point form of the two
 * maps, 4 bytes per destination pixel.  BORDER_CONSTANT 0 like the reference's call.  Source images up to 2046 px. */
typedef struct orbx_remap_table orbx_remap_table;
int orbx_remap_create(orbx_preproc* p, const float* mapx, const float* mapy, int dst_width, int dst_height, int map_stride_floats,
                      int src_width, int src_height, orbx_remap_table** out);
void orbx_remap_destroy(orbx_remap_table* t);
int orbx_remap(orbx_preproc* p, const orbx_remap_table* t, const uint8_t* src, int nframes, int src_stride, long long src_frame_stride,
               uint8_t* dst, int dst_stride, long long dst_frame_stride, int mem, void* stream);

/* cv::createCLAHE(clip_limit, Size(tiles_x, tiles_y))->apply(im, im): Examples/Monocular/mono_tum_vi.cc:101,109 (3.0, 8x8).
 * dst may be src (same strides). */
int orbx_clahe(orbx_preproc* p, const uint8_t* src, int nframes, int width, int height, int src_stride, long long src_frame_stride,
               double clip_limit, int tiles_x, int tiles_y, uint8_t* dst, int dst_stride, long long dst_frame_stride, int mem, void* stream);

/* ---- Frame glue on either side of the path -----------------------------------------------------------
 * Same handle, same mem / stream convention as the pre-processing calls above. */

/* Frame::UndistortKeyPoints, src/Frame.cc:837-870 (and the four corners of ComputeImageBounds, :872-899):
 * cv::undistortPoints(mat, mat, K, mDistCoef, cv::Mat(), mK) -- 5 fixed-point iterations in double, then re-projection with
 * K.  dist5 = k1 k2 p1 p2 k3 (k3 = 0 for the 4-coefficient YAMLs).  Every other keypoint field is copied; when dist5[0] == 0
 * the keypoints are copied unchanged as the reference does (:839-843).  kps_un may be kps. */
int orbx_undistort_keypoints(orbx_preproc* p, const orbx_keypoint* kps, int n, float fx, float fy, float cx, float cy, const float* dist5,
                             orbx_keypoint* kps_un, int mem, void* stream);

/* Frame::isInFrustum (monocular / rectified-stereo branch, src/Frame.cc:572-642) with MapPoint::PredictScale
 * (src/MapPoint.cc:587-602) and Pinhole::project (src/CameraModels/Pinhole.cpp:46-49) over n map points.  The outputs are the
 * MapPoint fields the reference fills -- mbTrackInView, mTrackProjX, mTrackProjY, mTrackProjXR, mTrackDepth,
 * mnTrackScaleLevel, mTrackViewCos -- and are exactly the inputs of orbx_search_by_projection_mappoints
 * (Tracking::SearchLocalPoints, src/Tracking.cc:3490-3520).  Points that fail get in_view 0, level -1 and proj_x / proj_y -1
 * unless they passed the image-bounds test (:577-578, :604-605).
 * max_distance / min_distance are MapPoint::mfMaxDistance / mfMinDistance AS STORED: the range gate applies the invariance
 * factors itself (dist > 1.2f * mfMaxDistance or dist < 0.8f * mfMinDistance rejects, MapPoint.cc:555-565, Frame.cc:609-615),
 * PredictScale divides the raw mfMaxDistance (MapPoint.cc:587-602).  Do NOT pass Get{Max,Min}DistanceInvariance(). */
typedef struct orbx_frustum {
  float Rcw[9], tcw[3], Ow[3];          /* mRcw (row-major), mtcw, mOw */
  float fx, fy, cx, cy;                 /* Pinhole parameters */
  float min_x, max_x, min_y, max_y;     /* mnMinX .. mnMaxY */
  float bf, log_scale_factor, viewing_cos_limit; /* mbf, mfLogScaleFactor, 0.5 at the call site */
  int n_scale_levels;                   /* mnScaleLevels */
  int camera_model;                     /* ORBX_CAMERA_PINHOLE (Pinhole::project) or ORBX_CAMERA_KB8 (KannalaBrandt8::project,
                                           src/CameraModels/KannalaBrandt8.cpp:29-45: TUM-VI fisheye, Frame.cc:597 through mpCamera) */
  float kb[4];                          /* KannalaBrandt8 k1..k4 (mvParameters[4..7]); ignored for the pinhole model */
} orbx_frustum;
#define ORBX_CAMERA_PINHOLE 0
#define ORBX_CAMERA_KB8 1
int orbx_is_in_frustum(orbx_preproc* p, const orbx_frustum* cam, int n, const float* world_pos, const float* normal, const float* max_distance,
                       const float* min_distance, uint8_t* in_view, float* proj_x, float* proj_y, float* proj_xr, float* depth, int* level,
                       float* view_cos, int mem, void* stream);

/* Tracking::SearchLocalPoints (src/Tracking.cc:3441-3525) over a batch of frames without host staging: the local map is
 * uploaded once (orbx_local_map_create: MapPoint::GetWorldPos / GetNormal / mfMaxDistance / mfMinDistance (raw, as for
 * orbx_is_in_frustum) / GetDescriptor, optional Observations() > 0 flags), then for the nframes frames whose keypoints
 * `ex` holds on the device from its last orbx_extract_batch_device, frame f with pose / intrinsics frames[f]:
 * isInFrustum(pMP, viewing_cos_limit) for every map point (Frame.cc:572-642) and SearchByProjection(F, vpMapPoints, th)
 * (ORBmatcher.cc:44-142) with mvpMapPoints initially empty.  kp_mp[f * cap_per_frame + k] = index of the map point matched
 * to keypoint k of frame f, or -1.  All kernels run on `stream` (pass the extractor's stream; NULL = the matcher's own
 * stream after a device synchronisation); the call returns after the results have reached the host arrays. */
typedef struct orbx_local_map orbx_local_map;
int orbx_local_map_create(orbx_matcher* m, int nmp, const float* world_pos, const float* normal, const float* max_distance,
                          const float* min_distance, const uint8_t* desc, const uint8_t* observed, orbx_local_map** out);
void orbx_local_map_destroy(orbx_local_map* map);
int orbx_search_local_points_batch(orbx_matcher* m, orbx_extractor* ex, int nframes, const orbx_frustum* frames, const orbx_local_map* map,
                                   float th, float nnratio, int* kp_mp, int cap_per_frame, int* nmatches, void* stream);

/* Page-locked host staging buffers for the *_host entry points (cudaHostAlloc).  Frames handed over in such a buffer upload at
 * the full link rate; memory pinned after the fact (cudaHostRegister of pageable pages, which is what e.g. torch's
 * pin_memory() produced on the test box) was measured at half of it (27 vs 55 GB/s host-to-device on PCIe 5 x16).  The
 * reference has no counterpart: its cv::Mat frames live in pageable memory (Examples/Stereo/stereo_euroc.cc:166-167). */
int orbx_host_alloc(size_t bytes, void** ptr);
int orbx_host_free(void* ptr);

/* orbx_knn2 for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
 * Same answer as orbx_knn2 (cv::BFMatcher(NORM_HAMMING).knnMatch(k = 2) + the 0.7 ratio flag of src/Frame.cc:1246-1253) with
 * every array in DEVICE memory of the CURRENT device: d_query nq x 32 bytes, d_train nt x 32 bytes, d_idx / d_dist nq x 2
 * ints (-1 where fewer than two train rows exist), d_good nq bytes or NULL.  Asynchronous on `stream` (NULL = the default
 * stream); needs no matcher handle.  The running (best, second) keys live in a file-scope device array of
 * ORBX_KNN2_STATIC_MAX_QUERIES x 16 bytes = 256 KB that every GPU loading liborbx.so holds until exit; nq beyond that limit
 * is refused with ORBX_ERR_CAPACITY (orbx_last_error says so), nt is unlimited.  The call runs no cudaMalloc / cudaFree or
 * allocating library routine; the first call on each device looks the array's address up (cudaGetSymbolAddress), so make
 * one call per device at set-up, before capturing.  All calls share the array, which keeps the last call's contents (each
 * call overwrites the keys it reads first): calls that overlap on the GPU -- different streams without an ordering event,
 * or two host threads -- race and must be ordered by the caller. */
#define ORBX_KNN2_STATIC_MAX_QUERIES 16384
int orbx_knn2_device_static(const uint8_t* d_query, int nq, const uint8_t* d_train, int nt, int* d_idx, int* d_dist, uint8_t* d_good,
                            void* stream);

const char* orbx_last_error(void);
const char* orbx_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
