/* libcuorb — C ABI of the B200-native ORB front-end (hand-written sm_100a CUDA).
 *
 * This is the ONLY boundary between the untouched CPU SLAM code of the reference
 * (elkuno213/orbslam3_cpp_fork: Tracking / Frame / ORBmatcher callers) and the GPU.
 * Plain C: opaque handles, POD structs, raw pointers and sizes.  No C++/torch types.
 *
 * Every entry point cites the reference interface it replaces (paths relative to the
 * reference checkout).  Error convention (reference: src/ORBextractor.cc:1060-1062 returns -1
 * on an empty image; no exceptions on this path): functions return >= 0 on success and a
 * negative cuorb_status on failure; cuorb_last_error() gives a thread-local message.
 * There is NO CPU fallback: if no CUDA device / kernel image is usable the call fails.
 */
#ifndef CUORB_H
#define CUORB_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CUORB_VERSION 100

typedef enum cuorb_status {
  CUORB_OK            = 0,
  CUORB_E_EMPTY       = -1, /* empty image: same value operator() returns (ORBextractor.cc:1060) */
  CUORB_E_BADARG      = -2, /* null pointer, bad size, image larger than the handle was built for */
  CUORB_E_CAPACITY    = -3, /* caller buffers (or an internal candidate list) too small */
  CUORB_E_CUDA        = -4, /* CUDA runtime error (sticky) — see cuorb_last_error() */
  CUORB_E_UNSUPPORTED = -5, /* geometry the reference itself has undefined behaviour for */
  CUORB_E_STATE       = -6  /* call sequence error (e.g. fetch before extract) */
} cuorb_status;

/* Layout-identical to cv::KeyPoint (7 x 4 bytes, no padding): pt.x, pt.y, size, angle,
 * response, octave, class_id.  The host shim can memcpy into std::vector<cv::KeyPoint>. */
typedef struct cuorb_keypoint {
  float   x, y;
  float   size;
  float   angle;    /* degrees in [0,360) (cv::fastAtan2) */
  float   response; /* FAST score 1..254 */
  int32_t octave;
  int32_t class_id; /* always -1 */
} cuorb_keypoint;

/* ORBextractor::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
 * (include/ORBextractor.h:48, src/ORBextractor.cc:403-465) + capacity of the device arena. */
typedef struct cuorb_config {
  int32_t nfeatures;
  float   scale_factor;
  int32_t nlevels;      /* 1..CUORB_MAX_LEVELS */
  int32_t ini_th_fast;
  int32_t min_th_fast;
  int32_t max_width;    /* largest image the handle will be given */
  int32_t max_height;
  int32_t max_batch;    /* images per batched call (1 for the drop-in operator()) */
  int32_t device;       /* CUDA device ordinal */
} cuorb_config;

#define CUORB_MAX_LEVELS 16
#define CUORB_DESC_BYTES 32

typedef struct cuorb_extractor cuorb_extractor;

/* where a caller-supplied image / descriptor buffer lives */
typedef enum cuorb_mem { CUORB_MEM_HOST = 0, CUORB_MEM_DEVICE = 1 } cuorb_mem;

const char* cuorb_last_error(void);
int         cuorb_version(void);
int         cuorb_device_count(void);

/* pinned host memory helpers (so H2D/D2H copies of the batched calls are truly asynchronous) */
void* cuorb_host_alloc(size_t bytes);
void  cuorb_host_free(void* p);

/* ---- extractor: replaces class ORB_SLAM3::ORBextractor (include/ORBextractor.h:41-119) ---- */

/* ctor (src/ORBextractor.cc:403-465); allocates every device/pinned buffer up front. */
cuorb_extractor* cuorb_extractor_create(const cuorb_config* cfg);
void             cuorb_extractor_destroy(cuorb_extractor* ext);

/* GetScaleFactors / GetInverseScaleFactors / GetScaleSigmaSquares / GetInverseScaleSigmaSquares
 * (include/ORBextractor.h:63-85) and mnFeaturesPerLevel (src/ORBextractor.cc:427-440).
 * Each out pointer may be NULL; arrays hold nlevels entries. Returns nlevels. */
int cuorb_extractor_get_tables(const cuorb_extractor* ext, float* scale, float* inv_scale,
                               float* sigma2, float* inv_sigma2, int32_t* features_per_level);

/* Largest keypoint count one image can produce with this handle's current geometry
 * (sum over levels of quota + quadtree overshoot slack); device-side result rows are this long. */
int cuorb_extractor_capacity(const cuorb_extractor* ext);

/* int ORBextractor::operator()(image, mask, keypoints, descriptors, vLappingArea)
 * (include/ORBextractor.h:55-61, src/ORBextractor.cc:1052-1138), called by Frame::ExtractORB
 * (src/Frame.cc:555-562).  img: 8-bit single channel, host memory, row pitch `stride`.
 * kps/desc: host buffers of `capacity` entries / capacity*32 bytes; *n_out = keypoint count.
 * Returns monoIndex (>= 0) or a negative cuorb_status (CUORB_E_EMPTY == the reference's -1). */
int cuorb_extract(cuorb_extractor* ext, const uint8_t* img, int width, int height, size_t stride,
                  int lap0, int lap1, cuorb_keypoint* kps, uint8_t* desc, int capacity, int* n_out);

/* Batched / asynchronous form of the same operator: n_images images of identical size laid out
 * `image_stride` bytes apart in host (pinned recommended) or device memory.  Work is enqueued on
 * the handle's stream and returns immediately; results stay on the device until fetched. */
int cuorb_extract_batch_async(cuorb_extractor* ext, int n_images, const uint8_t* imgs, int width,
                              int height, size_t stride, size_t image_stride, cuorb_mem where,
                              int lap0, int lap1);

/* Copy the results of the last batch to host and wait.  kps: n_images*capacity entries, desc:
 * n_images*capacity*32 bytes (either may be NULL to skip), n_out/mono_index_out: n_images ints.
 * Returns the number of images, or a negative status (CUORB_E_CAPACITY if any image has more
 * keypoints than `capacity`; counts are still written). */
int cuorb_extract_fetch(cuorb_extractor* ext, cuorb_keypoint* kps, uint8_t* desc, int capacity,
                        int32_t* n_out, int32_t* mono_index_out);
/* Asynchronous variant: enqueues D2H into pinned caller buffers; pair with cuorb_sync(). */
int cuorb_extract_fetch_async(cuorb_extractor* ext, cuorb_keypoint* kps, uint8_t* desc,
                              int capacity, int32_t* n_out, int32_t* mono_index_out);
int cuorb_sync(cuorb_extractor* ext);
/* Status of the last batch fetched with cuorb_extract_fetch_async, to be called after cuorb_sync():
 * CUORB_OK, or CUORB_E_CAPACITY when an internal list overflowed (FAST candidates / quadtree nodes:
 * the batch is truncated, not silently wrong) or when an image produced more keypoints than the
 * `capacity` given to the fetch.  The synchronous calls perform the same check themselves.  (The
 * reference has no such failure mode: its std::vectors grow; src/ORBextractor.cc:767 only reserves.) */
int cuorb_extractor_status(cuorb_extractor* ext);
/* The handle's cudaStream_t (as void*): lets a caller order its own work / record CUDA events
 * (e.g. torch.cuda.ExternalStream) against the enqueued pipeline.  The library also runs part of an
 * extraction (the blur) on a second, internal stream; that work is forked from and joined back into
 * this stream inside every extract call, so anything ordered after the handle's stream is ordered
 * after the whole extraction. */
void* cuorb_extractor_stream(cuorb_extractor* ext);

/* std::vector<cv::Mat> mvImagePyramid (include/ORBextractor.h:87; read by
 * Frame::ComputeStereoMatches, src/Frame.cc:950,1039,1056).  Device pointer to pixel (0,0) of
 * `level` of image `image_index` of the last batch (the 19-px REFLECT_101 border lies at negative
 * offsets, as in the reference's ROI-of-temp layout); valid until the next extract call. */
int cuorb_extractor_pyramid(const cuorb_extractor* ext, int image_index, int level,
                            const uint8_t** dev_ptr, int* width, int* height, size_t* stride);

/* The device arrays behind mvKeys / mDescriptors (and, once cuorb_stereo_match has run on this handle, mvuRight /
 * mvDepth) of image `image_index` of the last batch, where the extractor left them: what
 * cuorb_localmap_search_resident reads in place instead of a host round trip.  Work that reads them must be ordered
 * after cuorb_extractor_stream(ext); they are valid until the next extract call on the handle.  The counts are the
 * ones cuorb_extract_fetch returns.  Returns the per-image capacity of the arrays, or a negative cuorb_status. */
int cuorb_extractor_device_features(const cuorb_extractor* ext, int image_index, const cuorb_keypoint** d_keys,
                                    const uint8_t** d_descriptors, const float** d_u_right, const float** d_depth);

/* Parity/debug downloads (synchronous).  plane: 0 = pyramid level incl. 19-px border
 * ((w+38)*(h+38) bytes, tightly packed), 1 = pyramid level without border (w*h),
 * 2 = 7x7 sigma-2 blurred level (w*h) (src/ORBextractor.cc:1103-1104).  Returns bytes written. */
int64_t cuorb_extractor_download_plane(cuorb_extractor* ext, int image_index, int level, int plane,
                                       uint8_t* out, size_t out_bytes, int* width, int* height);
/* Per-level FAST candidates handed to DistributeOctTree (vToDistributeKeys,
 * src/ORBextractor.cc:766-850): coordinates relative to (minBorderX,minBorderY), FAST response,
 * sorted into the reference's cell-row-major / row-major order.  Returns the count. */
int cuorb_extractor_download_candidates(cuorb_extractor* ext, int image_index, int level,
                                        int32_t* xs, int32_t* ys, int32_t* responses, int capacity);

/* ---- stereo association: replaces void Frame::ComputeStereoMatches()
 * (include/Frame.h:155, src/Frame.cc:944-1113); BASELINE.json calls it "SearchStereo". ---- */

/* Device-resident form: matches the last batch extracted on `left` against the last batch on
 * `right` (same n_images, kps/descriptors/pyramids still on the device).  Enqueued on left's
 * stream after right's work; whatever is enqueued on `right` afterwards (its next extraction
 * overwrites the buffers the match reads) is ordered behind the match, so no host synchronisation
 * is needed between steps.  mb/mbf: Frame::mb, Frame::mbf (src/Frame.cc:257). */
int cuorb_stereo_match_batch_async(cuorb_extractor* left, cuorb_extractor* right, float mb,
                                   float mbf);
/* mvuRight / mvDepth (src/Frame.cc:945-946) of the last stereo batch: n_images*capacity floats
 * each, -1 where unmatched.  best_idx_r / best_dist / sad are optional debug outputs (may be
 * NULL): coarse Hamming winner, its distance (src/Frame.cc:1001-1026), and the SAD pushed to
 * vDistIdx (src/Frame.cc:1096) or -1. */
int cuorb_stereo_fetch(cuorb_extractor* left, float* u_right, float* depth, int capacity,
                       int32_t* best_idx_r, int32_t* best_dist, int32_t* sad);
int cuorb_stereo_fetch_async(cuorb_extractor* left, float* u_right, float* depth, int capacity);

/* Host-pointer form with the exact member set the reference method reads: left/right keypoints
 * and descriptors (mvKeys, mvKeysRight, mDescriptors, mDescriptorsRight), mb, mbf; scale tables
 * and pyramids come from the two extractor handles (image 0 of their last extract). */
int cuorb_stereo_match(cuorb_extractor* left, cuorb_extractor* right, int n_left,
                       const cuorb_keypoint* kps_left, const uint8_t* desc_left, int n_right,
                       const cuorb_keypoint* kps_right, const uint8_t* desc_right, float mb,
                       float mbf, float* u_right, float* depth);

/* Fused stereo frame: the stereo Frame constructor's three steps - ExtractORB(0, imLeft, 0, 0) and
 * ExtractORB(1, imRight, 0, 0) on two threads (src/Frame.cc:202-205, 555-562) and ComputeStereoMatches()
 * (src/Frame.cc:231, 944-1113) - as ONE call: both extractions are enqueued on the two handles' own streams, the
 * association runs on the device on the resident keypoints / descriptors / pyramids (nothing is uploaded again) and
 * each handle's results return in a single copy.  Both handles must have been created with max_batch == 1.  Host
 * images (same size, `stride` bytes per row); lapping area {0, 0} as the reference's stereo constructor passes.
 * kps_* / desc_*: `capacity` entries each; u_right / depth: `capacity` floats (mvuRight / mvDepth of the left
 * keypoints, -1 where there is no match).  Results are bit-identical to cuorb_extract x 2 + cuorb_stereo_match.
 * Returns the left monoIndex (>= 0) or a negative status. */
int cuorb_extract_stereo(cuorb_extractor* left, cuorb_extractor* right, const uint8_t* img_left,
                         const uint8_t* img_right, int width, int height, size_t stride, float mb, float mbf,
                         cuorb_keypoint* kps_left, uint8_t* desc_left, int* n_left, cuorb_keypoint* kps_right,
                         uint8_t* desc_right, int* n_right, int capacity, float* u_right, float* depth);

/* ---- matcher: replaces the tracking-time subset of class ORB_SLAM3::ORBmatcher ---- */

/* static int ORBmatcher::DescriptorDistance(const cv::Mat&, const cv::Mat&)
 * (include/ORBmatcher.h:40, src/ORBmatcher.cc:2075-2089).  Host-side popcount: one pair is far
 * too small for a device round trip. */
int cuorb_hamming256(const void* a, const void* b);
/* Batched device form: out[i] = Hamming(a[ia[i]], b[ib[i]]) for n_pairs pairs; a, b are host
 * arrays of 32-byte descriptors, ia/ib may be NULL (then pair i = (i, i)). */
int cuorb_hamming_batch(int device, const uint8_t* a, int n_a, const uint8_t* b, int n_b,
                        const int32_t* ia, const int32_t* ib, int n_pairs, int32_t* out);

/* The Frame members the two SearchByProjection overloads read (flattened by the host shim). */
typedef struct cuorb_frame_view {
  int32_t               n;             /* Frame::N (Nleft == -1: pinhole, non-fisheye) */
  const cuorb_keypoint* keys_un;       /* mvKeysUn */
  const uint8_t*        descriptors;   /* mDescriptors, n x 32 */
  const float*          u_right;       /* mvuRight (may be NULL = all -1) */
  const uint8_t*        bound_with_obs;/* per feature: mvpMapPoints[i] && Observations()>0 */
  float                 min_x, min_y, max_x, max_y; /* mnMinX.. (src/Frame.cc:932-940) */
  float                 grid_w_inv, grid_h_inv;     /* mfGridElementWidthInv/HeightInv */
  float                 mb, mbf;
  int32_t               nlevels;
  const float*          scale_factors; /* mvScaleFactors */
} cuorb_frame_view;

/* int ORBmatcher::SearchByProjection(Frame&, const std::vector<MapPoint*>&, th, bFarPoints,
 * thFarPoints) (include/ORBmatcher.h:44-50, src/ORBmatcher.cc:38-150; pinhole branch).
 * Per MapPoint SoA: proj_x/proj_y/proj_xr = mTrackProjX/Y/XR, level = mnTrackScaleLevel,
 * view_cos = mTrackViewCos, depth = mTrackDepth, flags bit0 = mbTrackInView, bit1 = isBad(),
 * bit2 = Observations() > 0.  best_idx_out[i] = frame feature bound to MapPoint i, or -1; the
 * caller replays F.mvpMapPoints[best] = pMP in MapPoint order.  Returns nmatches. */
#define CUORB_MP_IN_VIEW 1u
#define CUORB_MP_BAD     2u
#define CUORB_MP_HAS_OBS 4u
int cuorb_search_by_projection(int device, const cuorb_frame_view* frame, int n_mappoints,
                               const uint8_t* mp_desc, const float* proj_x, const float* proj_y,
                               const float* proj_xr, const int32_t* level, const float* view_cos,
                               const float* depth, const uint8_t* flags, float th, int far_points,
                               float th_far_points, float nn_ratio, int32_t* best_idx_out);

/* int ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono)
 * (include/ORBmatcher.h:54-56, src/ORBmatcher.cc:1684-1811, 1889-1908; pinhole branch).
 * Per last-frame feature SoA (only features with a non-outlier MapPoint are passed as valid):
 * u/v = projected pixel, inv_z = 1/zc, last_octave, last_angle = kpLF.angle, flags as above with
 * bit0 = "has MapPoint and not outlier".  forward/backward = bForward/bBackward (1702-1703).
 * best_idx_out[i] = current-frame feature or -1 (BEFORE the rotation check); rot_bin_out[i] =
 * histogram bin or -1; keep_bins_out[3] = ComputeThreeMaxima result (src/ORBmatcher.cc:2038).
 * Returns nmatches after the rotation-consistency removal, as the reference does. */
int cuorb_search_by_projection_frame(int device, const cuorb_frame_view* cur, int n_last,
                                     const uint8_t* last_desc, const float* u, const float* v,
                                     const float* inv_z, const int32_t* last_octave,
                                     const float* last_angle, const uint8_t* flags, float th,
                                     int forward, int backward, int check_orientation,
                                     int32_t* best_idx_out, int32_t* rot_bin_out,
                                     int32_t* keep_bins_out);

/* ---- the same two searches on a two-camera rig (Frame::Nleft != -1, the "fisheye" branches) ----
 * A rig Frame keeps the left camera's features at indices [0, Nleft) (mvKeys, mGrid) and the right camera's at
 * [Nleft, N) (mvKeysRight, mGridRight, descriptor rows idx + Nleft).  It is passed as two cuorb_frame_views:
 * `left`  = {Nleft,  mvKeys,      mDescriptors rows [0, Nleft), bound_with_obs of mvpMapPoints[idx]},
 * `right` = {Nright, mvKeysRight, mDescriptors rows [Nleft, N), bound_with_obs of mvpMapPoints[idx + Nleft]};
 * both carry the Frame's image bounds, grid cell sizes and scale factors; u_right is ignored (the mvuRight gate is
 * pinhole-only, src/ORBmatcher.cc:103, :1770).  Results are indices local to each camera. */

/* int ORBmatcher::SearchByProjection(Frame&, const std::vector<MapPoint*>&, th, bFarPoints, thFarPoints) with
 * F.Nleft != -1 (src/ORBmatcher.cc:38-227: left part :63-150, right part :152-224).  left_to_right / right_to_left =
 * mvLeftToRightMatch [Nleft] / mvRightToLeftMatch [Nright] (-1: no stereo partner).  Per MapPoint: proj_x / proj_y /
 * level / view_cos = mTrackProjX / Y, mnTrackScaleLevel, mTrackViewCos; proj_xr / proj_yr / level_r / view_cos_r =
 * mTrackProjXR / YR, mnTrackScaleLevelR (-1 skips the right part, :154), mTrackViewCosR; flags as above plus
 * CUORB_MP_IN_VIEW_R = mbTrackInViewR.  Reference behaviour kept: a match on one side also writes the MapPoint into
 * the stereo partner's slot on the other side, unconditionally (:138-142, :212-216), so a slot's state is what its
 * LAST writer left (a MapPoint without observations un-blocks it); a ratio-test rejection on the left skips the
 * right part (:130-132); the right window has no `th` factor (:156-163).
 * best_l_out[i] / best_r_out[i] = left / right feature MapPoint i matched directly, or -1; the caller replays, in
 * MapPoint order, F.mvpMapPoints[bl] = pMP, F.mvpMapPoints[Nleft + l2r[bl]] = pMP, then F.mvpMapPoints[r2l[br]] = pMP,
 * F.mvpMapPoints[Nleft + br] = pMP.  Returns nmatches (one per slot written, as the reference counts). */
#define CUORB_MP_IN_VIEW_R 8u
int cuorb_search_by_projection_rig(int device, const cuorb_frame_view* left, const cuorb_frame_view* right,
                                   const int32_t* left_to_right, const int32_t* right_to_left, int n_mappoints,
                                   const uint8_t* mp_desc, const float* proj_x, const float* proj_y,
                                   const int32_t* level, const float* view_cos, const float* proj_xr,
                                   const float* proj_yr, const int32_t* level_r, const float* view_cos_r,
                                   const float* depth, const uint8_t* flags, float th, int far_points,
                                   float th_far_points, float nn_ratio, int32_t* best_l_out, int32_t* best_r_out);

/* int ORBmatcher::SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) with CurrentFrame.Nleft != -1
 * (src/ORBmatcher.cc:1684-1908, right part :1812-1886).  u / v / inv_z as for cuorb_search_by_projection_frame;
 * ur / vr = project(CurrentFrame.GetRelativePoseTrl() * x3Dc) (:1813-1814).  Reference behaviour kept: the right part
 * only runs for points that passed the left part's depth and bounds tests AND whose left window was not empty
 * (the `continue` at :1750-1752), has no bounds test of its own, makes no cross-writes, and shares the rotation
 * histogram with the left part.  best_l_out / best_r_out before the rotation check, rot_bin_l_out / rot_bin_r_out
 * the bins, keep_bins_out[3] the kept ones; the caller replays the writes in last-frame order and then NULLs every
 * (point, feature) entry whose bin was not kept (:1889-1905).  Returns nmatches after that removal. */
int cuorb_search_by_projection_frame_rig(int device, const cuorb_frame_view* left, const cuorb_frame_view* right,
                                         int n_last, const uint8_t* last_desc, const float* u, const float* v,
                                         const float* inv_z, const float* ur, const float* vr,
                                         const int32_t* last_octave, const float* last_angle, const uint8_t* flags,
                                         float th, int forward, int backward, int check_orientation,
                                         int32_t* best_l_out, int32_t* best_r_out, int32_t* rot_bin_l_out,
                                         int32_t* rot_bin_r_out, int32_t* keep_bins_out);

/* ---- benchmark input side ("cfg5") ----
 * Synthetic stereo pairs generated ON THE DEVICE: frames first_frame_id .. first_frame_id + n_pairs - 1 of
 * the counter-based generator defined by orbslam3_cpp_fork_b200/synth_cb.py (integer arithmetic only; the
 * CUDA and numpy statements agree byte for byte, tests/test_gpu_synth.py).  d_left / d_right: device, tightly
 * packed n_pairs x height x width u8; d_scratch: device, same size (the left eye before pixel noise).
 * Enqueued on `stream` (a cudaStream_t, may be NULL).  No counterpart in the reference (it reads datasets). */
int cuorb_synth_stereo(int device, uint64_t first_frame_id, int n_pairs, int width, int height,
                       uint8_t* d_left, uint8_t* d_right, uint8_t* d_scratch, void* stream);

/* ---- pre-extraction image operations ----
 * What System::TrackStereo / TrackRGBD / TrackMonocular (src/System.cc:280-300, 384-386, 476) and
 * Tracking::GrabImageStereo / RGBD / Monocular (src/Tracking.cc:1339-1357, 1449-1462, 1520-1533)
 * do to a camera frame before ORBextractor::operator() sees it, fused into one device pass:
 *   mode CUORB_PRE_REMAP : cv::remap(im, out, M1, M2, cv::INTER_LINEAR), M1/M2 = the CV_32FC1 maps of
 *                          cv::initUndistortRectifyMap (src/Settings.cc:492-510), BORDER_CONSTANT 0
 *   mode CUORB_PRE_RESIZE: cv::resize(im, out, newImSize)  (INTER_LINEAR; exact 2x2 -> INTER_AREA)
 *   then cv::cvtColor(... -> GRAY) when the frame has 3 or 4 channels.
 * Results are bit-identical to OpenCV 4.13's 8-bit fixed-point paths. */
typedef enum cuorb_pixfmt {
  CUORB_PIX_GRAY8 = 0, CUORB_PIX_RGB8 = 1, CUORB_PIX_BGR8 = 2, CUORB_PIX_RGBA8 = 3, CUORB_PIX_BGRA8 = 4
} cuorb_pixfmt;
typedef enum cuorb_pre_mode { CUORB_PRE_NONE = 0, CUORB_PRE_REMAP = 1, CUORB_PRE_RESIZE = 2 } cuorb_pre_mode;
typedef struct cuorb_preproc cuorb_preproc;

/* src_w x src_h: camera frame; dst_w x dst_h: what the extractor gets (== src for CUORB_PRE_NONE;
 * the map size for REMAP; newImSize for RESIZE).  map_x / map_y: host float arrays dst_h x dst_w
 * (REMAP only, else NULL).  max_batch frames per call.  NULL on error. */
cuorb_preproc* cuorb_preproc_create(int device, cuorb_pre_mode mode, int src_w, int src_h, int dst_w,
                                    int dst_h, const float* map_x, const float* map_y, int max_batch);
void cuorb_preproc_destroy(cuorb_preproc* pre);

/* Stand-alone (synchronous): n frames (row pitch `stride` bytes, `image_stride` bytes apart, host or
 * device) -> n gray planes dst_h x dst_w, tightly packed, in host or device memory. */
int cuorb_preprocess(cuorb_preproc* pre, int n_images, const uint8_t* src, size_t stride,
                     size_t image_stride, cuorb_mem src_where, cuorb_pixfmt fmt, uint8_t* dst_gray,
                     cuorb_mem dst_where);

/* Fused with extraction: the gray planes are written straight into the extractor's level-0 input on
 * the extractor's stream, then the cuorb_extract_batch_async pipeline runs on them (asynchronous;
 * fetch as usual).  The extractor must have been created for at least dst_w x dst_h. */
int cuorb_extract_batch_async_pre(cuorb_extractor* ext, cuorb_preproc* pre, int n_images,
                                  const uint8_t* src, size_t stride, size_t image_stride,
                                  cuorb_mem where, cuorb_pixfmt fmt, int lap0, int lap1);

/* ---- per-frame glue after extraction ---- */

/* void Frame::UndistortKeyPoints() (src/Frame.cc:886-913): cv::undistortPoints(mvKeys, K, D, cv::Mat(), K)
 * with OpenCV's double arithmetic (5 iterations), or a plain copy when dist5[0] == 0 as the reference
 * does.  K4 = fx, fy, cx, cy (the float entries of mK); dist5 = k1, k2, p1, p2, k3 (k3 = 0 if absent).
 * Host arrays; kps_un may alias kps.  Returns n. */
int cuorb_undistort_keypoints(int device, int n, const cuorb_keypoint* kps, const float* K4,
                              const float* dist5, cuorb_keypoint* kps_un);

/* bool Frame::isInFrustum(MapPoint* pMP, float viewingCosLimit) (src/Frame.cc:644-707, the Nleft == -1
 * branch; MapPoint::PredictScale src/MapPoint.cc:574-589; Pinhole::project src/CameraModels/Pinhole.cpp:70-76)
 * for n MapPoints at once: the producer of mbTrackInView / mTrackProjX / mTrackProjY / mTrackProjXR /
 * mnTrackScaleLevel / mTrackViewCos / mTrackDepth that cuorb_search_by_projection consumes
 * (Tracking::SearchLocalPoints, src/Tracking.cc).  camera: the Frame members the function reads.
 * world_pos / normal: n x 3 floats (GetWorldPos / GetNormal); min_distance / max_distance: mfMinDistance /
 * mfMaxDistance (the 0.8 / 1.2 invariance factors are applied here, as GetMin/MaxDistanceInvariance do).
 * Outputs (host arrays of n): in_view = the return value; proj_x / proj_y = -1 unless the projection falls
 * inside the image bounds (the reference writes them before the distance / angle tests); proj_xr,
 * scale_level, view_cos, depth are only meaningful where in_view (else -1, -1, 0, 0).
 * Float policy: every operation is a single round-to-nearest float operation in the reference's expression
 * order with Eigen's size-3 reduction order x0 + (x1 + x2) and no FMA contraction (the reference's own build,
 * -O3 -march=native, may contract); log() of PredictScale is evaluated in double and rounded.
 * Returns the number of MapPoints in view. */
typedef struct cuorb_frustum_camera {
  float rcw[9];   /* mRcw, row-major */
  float tcw[3];   /* mtcw */
  float ow[3];    /* mOw */
  float fx, fy, cx, cy, mbf;
  float min_x, max_x, min_y, max_y;   /* mnMinX .. mnMaxY */
  float log_scale_factor;             /* mfLogScaleFactor */
  int32_t n_levels;                   /* mnScaleLevels */
} cuorb_frustum_camera;
int cuorb_is_in_frustum(int device, const cuorb_frustum_camera* camera, int n, const float* world_pos,
                        const float* normal, const float* min_distance, const float* max_distance,
                        float viewing_cos_limit, uint8_t* in_view, float* proj_x, float* proj_y,
                        float* proj_xr, int32_t* scale_level, float* view_cos, float* depth);

/* Allocation-free variant of cuorb_is_in_frustum.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid allocation).  Same arguments, same kernel, same results, but the device
 * staging area is a file-scope `static __device__` array instead of a cudaMalloc'ed workspace:
 *   - at most CUORB_FRUSTUM_NOALLOC_MAX MapPoints per call; a larger n returns CUORB_E_CAPACITY with an error text;
 *   - every GPU that loads libcuorb holds the array (about 3.6 MiB) until the process exits;
 *   - the call performs no cudaMalloc / cudaFree and calls no allocating library routine;
 *   - all calls on one device share the array, which keeps the last call's contents (the call clears the one word
 *     that must start at zero): calls that overlap on the same GPU - other host threads, other streams - RACE on it
 *     and must be serialised by the caller.  The call itself is synchronous (returns after its stream drained). */
#define CUORB_FRUSTUM_NOALLOC_MAX 65536
int cuorb_is_in_frustum_noalloc(int device, const cuorb_frustum_camera* camera, int n, const float* world_pos,
                                const float* normal, const float* min_distance, const float* max_distance,
                                float viewing_cos_limit, uint8_t* in_view, float* proj_x, float* proj_y,
                                float* proj_xr, int32_t* scale_level, float* view_cos, float* depth);

/* ---- resident local map: Tracking::SearchLocalPoints' two hot loops in one call ----
 * What Tracking::SearchLocalPoints (src/Tracking.cc) does per frame with the local map: for every MapPoint that is
 * not already matched in this frame (mnLastFrameSeen == mCurrentFrame.mnId) and not bad, Frame::isInFrustum(pMP, 0.5)
 * (src/Frame.cc:644-721) fills the tracking fields; then ORBmatcher::SearchByProjection(F, mvpLocalMapPoints, th,
 * bFarPoints, thFarPoints) (src/ORBmatcher.cc:38-150) consumes them.  The local map changes by a few per cent per
 * frame while its world positions, normals, distance ranges and descriptors are what both loops read, so the handle
 * keeps them in HBM: per frame only the camera pose and the frame's own features go up and the matches come back.
 * Results are those of cuorb_is_in_frustum followed by cuorb_search_by_projection on the same data, bit for bit.
 * One handle serves one tracking thread; calls on one handle must not overlap. */
typedef struct cuorb_localmap cuorb_localmap;
cuorb_localmap* cuorb_localmap_create(int device, int capacity);   /* NULL + cuorb_last_error() on failure */
void cuorb_localmap_destroy(cuorb_localmap* map);
/* Write points [first, first + n) and set the number of points in use to new_count (<= capacity).  world_pos /
 * normal: n x 3 floats (GetWorldPos / GetNormal), min_distance / max_distance: mfMinDistance / mfMaxDistance (the
 * 0.8 / 1.2 invariance factors are applied inside), descriptors: n x 32 (GetDescriptor), flags: CUORB_MP_BAD |
 * CUORB_MP_HAS_OBS per point.  Slots never written count as bad points.  Host arrays; synchronous. */
int cuorb_localmap_update(cuorb_localmap* map, int first, int n, int new_count, const float* world_pos,
                          const float* normal, const float* min_distance, const float* max_distance,
                          const uint8_t* descriptors, const uint8_t* flags);
int cuorb_localmap_count(const cuorb_localmap* map);
/* One frame.  cam: the Frame's pose / intrinsics as for cuorb_is_in_frustum; frame: as for
 * cuorb_search_by_projection.  order[n_order] (may be NULL = slots 0 .. count-1 in storage order): the storage slots
 * of mvpLocalMapPoints in the order Tracking holds them this frame - the reference's loops run in that order and the
 * order decides who gets a contested feature, so storage can stay put while the vector is rebuilt every frame
 * (n_order <= capacity; a slot outside [0, count) counts as a bad point).  Everything below is indexed by POSITION in
 * that sequence.  skip_pos[n_skip]: positions to leave out (mnLastFrameSeen == this frame).  Outputs (host):
 * best_idx_out[n] (optional) = feature bound to each position or -1; in_view_out[n] (optional) = mbTrackInView as
 * isInFrustum leaves it (the caller's IncreaseVisible loop); match_pos_out / match_feature_out[match_capacity] = the
 * matches as (position, feature) pairs in sequence order, the order in which the caller replays
 * F.mvpMapPoints[feature] = pMP (CUORB_E_CAPACITY if there are more than match_capacity); n_in_view_out = number of
 * points isInFrustum accepted (the reference's nToMatch).  Returns nmatches. */
int cuorb_localmap_search(cuorb_localmap* map, const cuorb_frustum_camera* cam, const cuorb_frame_view* frame, int n_order,
                          const int32_t* order, int n_skip, const int32_t* skip_pos, float viewing_cos_limit, float th,
                          int far_points, float th_far_points, float nn_ratio, int32_t* best_idx_out,
                          uint8_t* in_view_out, int match_capacity, int32_t* match_pos_out, int32_t* match_feature_out,
                          int32_t* n_in_view_out);
/* The same search on a Frame whose arrays never left the device: frame->keys_un, frame->descriptors and (optional)
 * frame->u_right are DEVICE pointers on the map's device (cuorb_extractor_device_features), the remaining members
 * of the view are host data as before.  Replaces, on the device, the three steps the reference runs between
 * ORBextractor::operator() and SearchLocalPoints: Frame::UndistortKeyPoints (src/Frame.cc:886-913; K4 / dist5 as for
 * cuorb_undistort_keypoints, both NULL when keys_un already holds undistorted keypoints - a rectified stereo rig),
 * Frame::AssignFeaturesToGrid (src/Frame.cc:524-553) and the flattening of the Frame for the search.
 * producer_stream: the cudaStream_t that wrote the device arrays (cuorb_extractor_stream(ext)); the search is ordered
 * after it.  NULL = the arrays are already complete.  keys_un_out[frame->n] (host, optional): mvKeysUn as the search
 * used it.  Everything else as cuorb_localmap_search; results are identical to those of that call on the same Frame. */
int cuorb_localmap_search_resident(cuorb_localmap* map, const cuorb_frustum_camera* cam, const cuorb_frame_view* frame,
                                   void* producer_stream, const float* K4, const float* dist5,
                                   cuorb_keypoint* keys_un_out, int n_order, const int32_t* order, int n_skip,
                                   const int32_t* skip_pos, float viewing_cos_limit, float th, int far_points,
                                   float th_far_points, float nn_ratio, int32_t* best_idx_out, uint8_t* in_view_out,
                                   int match_capacity, int32_t* match_pos_out, int32_t* match_feature_out,
                                   int32_t* n_in_view_out);


/* void Frame::ComputeStereoFromRGBD(const cv::Mat& imDepth) (src/Frame.cc:1115-1133): depth = CV_32F
 * image (row pitch stride_bytes) in host or device memory; d = imDepth(v, u) at the truncated
 * keypoint position; if d > 0: depth_out = d, u_right = kps_un.x - mbf / d, else both -1. */
int cuorb_stereo_from_rgbd(int device, int n, const cuorb_keypoint* kps, const cuorb_keypoint* kps_un,
                           const float* depth, int width, int height, size_t stride_bytes,
                           cuorb_mem depth_where, float mbf, float* u_right, float* depth_out);

/* ---- vocabulary-node searches ---- */

/* int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches)
 * (src/ORBmatcher.cc:237-420, mode 0) and
 * int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12)
 * (src/ORBmatcher.cc:784-916, mode 1), mono branches, on flattened inputs.  Side 1 = pKF / pKF1,
 * side 2 = F / pKF2.  valid1[i] = side-1 feature i has a non-bad MapPoint; eligible2[j] = side-2 feature
 * j may be matched (all ones for a Frame; "has a non-bad MapPoint" for a KeyFrame); angle1/angle2 =
 * keypoint angles (may be NULL when check_orientation == 0).  The merge-join of the two
 * DBoW2::FeatureVectors is passed as its n_nodes common nodes: node k lists side-1 features
 * node_idx1[node_off1[k] .. node_off1[k+1]) and side-2 features node_idx2[node_off2[k] ..), both in the
 * FeatureVector's order.  match12_out[i] = side-2 feature matched to side-1 feature i, or -1, after
 * the rotation-consistency cut.  Returns nmatches. */
int cuorb_search_by_bow(int device, int n1, const uint8_t* desc1, const uint8_t* valid1,
                        const float* angle1, int n2, const uint8_t* desc2, const uint8_t* eligible2,
                        const float* angle2, int n_nodes, const int32_t* node_off1,
                        const int32_t* node_idx1, const int32_t* node_off2, const int32_t* node_idx2,
                        int mode, float nnratio, int check_orientation, int32_t* match12_out);

/* The same (KeyFrame, Frame) search when F is a two-camera rig Frame (F.Nleft != -1; src/ORBmatcher.cc:306-331,
 * :371-396).  Side-2 arrays cover F's features [0, nleft2) (left camera) and [nleft2, n2) (right camera); the node lists
 * index that combined range.  A KeyFrame feature may bind one feature per camera: the left one under the ratio test,
 * the right one inside the left one's `bestDist1 <= TH_LOW` block with its ratio test disabled (`|| true`, :373), as the
 * reference does.  match_left_out[i1] / match_right_out[i1] = F feature or -1, after the rotation cut; the caller sets
 * vpMapPointMatches[match] = pKF's MapPoint i1.  angle1 = kp.angle of the KeyFrame feature (:340-343), angle2 = the F
 * keypoint's (mvKeys for the left range, mvKeysRight for the right one).  Returns nmatches (both cameras). */
int cuorb_search_by_bow_rig(int device, int n1, const uint8_t* desc1, const uint8_t* valid1, const float* angle1, int n2,
                            int nleft2, const uint8_t* desc2, const float* angle2, int n_nodes, const int32_t* node_off1,
                            const int32_t* node_idx1, const int32_t* node_off2, const int32_t* node_idx2, float nnratio,
                            int check_orientation, int32_t* match_left_out, int32_t* match_right_out);

/* ---- the remaining ORBmatcher searches, pinhole / NLeft == -1 branches ----
 * The caller keeps what needs Sophus / Eigen / MapPoint state (pose transform, camera projection,
 * distance-invariance range, viewing-angle test, MapPoint::PredictScale) and passes one SoA entry per
 * query: active (all of those gates passed), the projected pixel (u, v), the predicted pyramid level
 * and the MapPoint descriptor.  The window search, the per-candidate gates, the reference's visiting
 * order and first/last-wins tie rules, the order-dependent binding and the rotation histogram run in
 * the library.  Views: keys_un = mvKeysUn, u_right = mvuRight; the meaning of bound_with_obs is stated
 * per function.  Host arrays throughout. */

/* int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>&
 * sAlreadyFound, const float th, const int ORBdist)  (include/ORBmatcher.h:60-66,
 * src/ORBmatcher.cc:1910-2036).  One entry per pKF->GetMapPointMatches()[i]: active = pMP && !isBad()
 * && !sAlreadyFound.count(pMP) && minDistance <= dist3D <= maxDistance; kf_angle = pKF->mvKeysUn[i].angle.
 * cur->bound_with_obs[i2] = CurrentFrame.mvpMapPoints[i2] != NULL on entry.  The image-bounds test
 * (:1937-1942) is applied here.  Outputs as cuorb_search_by_projection_frame.  Returns nmatches. */
int cuorb_search_by_projection_reloc(int device, const cuorb_frame_view* cur, int n_points,
                                     const uint8_t* mp_desc, const float* u, const float* v,
                                     const int32_t* level, const float* kf_angle, const uint8_t* active,
                                     float th, int orb_dist, int check_orientation, int32_t* best_idx_out,
                                     int32_t* rot_bin_out, int32_t* keep_bins_out);

/* int ORBmatcher::SearchByProjection(KeyFrame* pKF, Sophus::Sim3f& Scw, const std::vector<MapPoint*>&
 * vpPoints, std::vector<MapPoint*>& vpMatched, int th, float ratioHamming)  (include/ORBmatcher.h:70-77,
 * src/ORBmatcher.cc:423-538) and its overload that also fills vpMatchedKF (include/ORBmatcher.h:81-90,
 * src/ORBmatcher.cc:540-664; the search is the same, the caller copies vpPointsKFs[i] beside the match).
 * kf->bound_with_obs[idx] = vpMatched[idx] != NULL on entry.  best_idx_out[i] = KeyFrame feature now
 * matched to point i (the caller replays vpMatched[best] = vpPoints[i]) or -1.  Returns nmatches. */
int cuorb_search_by_projection_sim3(int device, const cuorb_frame_view* kf, int n_points,
                                    const uint8_t* mp_desc, const float* u, const float* v,
                                    const int32_t* level, const uint8_t* active, int th, float ratio_hamming,
                                    int32_t* best_idx_out);

/* int ORBmatcher::Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th, const
 * bool bRight = false)  (include/ORBmatcher.h:129-134, src/ORBmatcher.cc:1149-1328; check_reprojection
 * = 1: the 5.99 / 7.8 chi-square gates with inv_level_sigma2 = pKF->mvInvLevelSigma2 and ur = uv(0) -
 * bf * invz, :1208) and int ORBmatcher::Fuse(KeyFrame* pKF, Sophus::Sim3f& Scw, const
 * std::vector<MapPoint*>& vpPoints, float th, std::vector<MapPoint*>& vpReplacePoint)
 * (include/ORBmatcher.h:137-143, src/ORBmatcher.cc:1330-1452; check_reprojection = 0, ur unused).
 * best_idx_out[i] = the KeyFrame feature MapPoint i fuses into (bestDist <= TH_LOW) or -1; the caller
 * performs the Replace / AddObservation / AddMapPoint bookkeeping (:1301-1318).  Returns nFused. */
int cuorb_fuse(int device, const cuorb_frame_view* kf, const float* inv_level_sigma2, int n_points,
               const uint8_t* mp_desc, const float* u, const float* v, const float* ur, const int32_t* level,
               const uint8_t* active, float th, int check_reprojection, int32_t* best_idx_out);

/* int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const
 * Sophus::Sim3f& S12, const float th)  (include/ORBmatcher.h:120-126, src/ORBmatcher.cc:1454-1682).
 * Direction 1->2 has kf1->n entries (MapPoints of pKF1 projected into pKF2: mp_desc1, u12, v12,
 * level12, active1 = pMP && !vbAlreadyMatched1 && !isBad() && the geometric gates), direction 2->1 has
 * kf2->n entries.  match12_out[i1] = feature of pKF2 when both directions agree (:1667-1679), else -1.
 * Returns nFound. */
int cuorb_search_by_sim3(int device, const cuorb_frame_view* kf1, const cuorb_frame_view* kf2,
                         const uint8_t* mp_desc1, const float* u12, const float* v12, const int32_t* level12,
                         const uint8_t* active1, const uint8_t* mp_desc2, const float* u21, const float* v21,
                         const int32_t* level21, const uint8_t* active2, float th, int32_t* match12_out);

/* int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched,
 * std::vector<int>& vnMatches12, int windowSize = 10)  (include/ORBmatcher.h:99-105,
 * src/ORBmatcher.cc:666-782).  prev_matched = vbPrevMatched as f1->n (x, y) pairs, updated in place
 * (:771-776); matches12_out = vnMatches12 (f1->n).  Returns nmatches. */
int cuorb_search_for_initialization(int device, const cuorb_frame_view* f1, const cuorb_frame_view* f2,
                                    float* prev_matched, int window_size, float nnratio,
                                    int check_orientation, int32_t* matches12_out);

/* int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<std::pair<size_t,
 * size_t>>& vMatchedPairs, const bool bOnlyStereo, const bool bCoarse = false)
 * (include/ORBmatcher.h:108-114, src/ORBmatcher.cc:916-1147; mpCamera2 == NULL).  The common nodes of the
 * two FeatureVectors are passed as for cuorb_search_by_bow.  kfX->bound_with_obs[idx] =
 * pKFX->GetMapPoint(idx) != NULL; F12 = K1^-T [t12]x R12 K2^-1 row-major and epipole = project(T2w * Cw1)
 * come from the caller (:927-939, src/CameraModels/Pinhole.cpp:153-156); level_sigma2_2 =
 * pKF2->mvLevelSigma2.  match12_out[idx1] = idx2 or -1 after the rotation cut (the caller builds
 * vMatchedPairs from it, :1133-1144).  Returns nmatches. */
int cuorb_search_for_triangulation(int device, const cuorb_frame_view* kf1, const cuorb_frame_view* kf2,
                                   int n_nodes, const int32_t* node_off1, const int32_t* node_idx1,
                                   const int32_t* node_off2, const int32_t* node_idx2, const float* F12,
                                   const float* epipole, const float* level_sigma2_2, int only_stereo,
                                   int coarse, int check_orientation, int32_t* match12_out);

/* ---- dense all-pairs Hamming ---- */

/* cv::BFMatcher(cv::NORM_HAMMING).knnMatch(query, train, matches, 2) as Frame::ComputeStereoFishEyeMatches
 * uses it (src/Frame.cc:1304-1355), batched: n_sets independent (query, train) pairs of descriptor sets laid
 * out back to back; query_off / train_off hold n_sets + 1 ascending row offsets.  Per query row:
 * idx_out[2i..2i+1] / dist_out[2i..2i+1] = the two nearest train rows (index relative to the set's first
 * train row) in (distance, trainIdx) order, -1 where the train set has fewer than two rows.
 * ratio_match_out (optional, one int per query): idx0 if dist0 < dist1 * 0.7 (the Lowe test of :1329)
 * else -1.  Host arrays.  Returns the number of query rows. */
int cuorb_knn_match2(int device, int n_sets, const int32_t* query_off, const uint8_t* query_desc,
                     const int32_t* train_off, const uint8_t* train_desc, int32_t* idx_out,
                     int32_t* dist_out, int32_t* ratio_match_out);

/* void MapPoint::ComputeDistinctiveDescriptors() (src/MapPoint.cc:376-450) for n_points MapPoints at once:
 * desc = the observed descriptors of all points back to back (the reference's vDescriptors order),
 * off = n_points + 1 ascending row offsets.  best_out[p] = BestIdx (row inside point p's block; -1 for a
 * point without descriptors).  At most 1024 observations per point.  Host arrays. */
int cuorb_distinctive_descriptors(int device, int n_points, const int32_t* off, const uint8_t* desc,
                                  int32_t* best_out);

/* ---- instrumentation ---- */

/* number of libcuorb kernel launches issued by this process so far (bench.py's gpu_launches) */
uint64_t cuorb_kernel_launch_count(void);
/* CUDA-event time, in ms, of the stages of the last batch on this handle (requires
 * cuorb_set_profiling(ext, 1) before the extract call).  stage ids: CUORB_STAGE_*. */
enum {
  CUORB_STAGE_H2D = 0, CUORB_STAGE_PYRAMID, CUORB_STAGE_FAST, CUORB_STAGE_OCTREE,
  CUORB_STAGE_BLUR, CUORB_STAGE_DESCRIBE, CUORB_STAGE_STEREO, CUORB_STAGE_COUNT
};
int cuorb_set_profiling(cuorb_extractor* ext, int enabled);
int cuorb_stage_ms(cuorb_extractor* ext, float* ms_out /* CUORB_STAGE_COUNT floats */);

#ifdef __cplusplus
}
#endif
#endif /* CUORB_H */
