/*
 * orbgpu.h -- C ABI of liborbgpu, the B200 (sm_100a) implementation of ORB-SLAM2's
 * data-parallel front end.  This is the drop-in boundary for the hot path of
 * shishiteng/AR_ORBSLAM2: every entry point replaces one reference interface, cited as
 * "ref:" (paths relative to the reference checkout, ORB_SLAM2/...).
 *
 * Conventions
 *   - plain C types only: pointers, sizes, POD structs.  No C++ or torch types.
 *   - every function returns ORB_OK (0) or a negative orb_status; nothing throws.
 *     orb_last_error() returns a thread-local message for the last failure.
 *   - a handle may be used by one thread at a time; distinct handles are independent
 *     (each owns its CUDA stream and device buffers), matching the reference where the
 *     stereo Frame constructor runs two extractor objects concurrently (ref: src/Frame.cc:83-86).
 *   - there is NO CPU fallback: without a CUDA device every compute call fails with
 *     ORB_ERR_CUDA.
 *   - "loc" flags say where caller buffers live (ORB_HOST / ORB_DEVICE).
 */
#ifndef ORBGPU_H_
#define ORBGPU_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBGPU_ABI_VERSION 1

/* Streams.  Every handle (extractor, frame, database, vocabulary) owns a non-blocking CUDA stream.  The entry points that take
 * no handle (orb_descriptor_distance, orb_hamming_top2, orb_top2_merge, orb_search_matches_ratio, orb_is_in_frustum,
 * orb_undistort_*, orb_distinctive_descriptors, orb_rot_hist_filter, and orb_cvt_color_gray / orb_remap_linear called with
 * stream == NULL) run on the calling thread's CALL STREAM: by default a non-blocking stream the library creates per thread --
 * never CUDA's legacy default stream, so threads do not serialise against each other (the reference calls these from
 * Tracking, LocalMapping and LoopClosing concurrently).  Host-mode (ORB_HOST) calls block until their results are in the
 * caller's arrays.  Device-mode (ORB_DEVICE) calls are ASYNCHRONOUS on the call stream and are NOT ordered with any handle's
 * stream: to feed a helper's device output to a handle (e.g. orb_cvt_color_gray -> orb_extract_batch, orb_is_in_frustum ->
 * a matcher) either make the handle's stream the call stream first -- orb_set_call_stream(orb_extractor_stream(ex), 1) --
 * or call orb_call_stream_sync() in between.                                                                           */
int orb_set_call_stream(void *cuda_stream, int use_it);   /* use_it = 0: back to the thread's own stream                 */
void *orb_get_call_stream(void);                          /* cudaStream_t the next stream-less call of this thread uses  */
int orb_call_stream_sync(void);

typedef enum orb_status {
    ORB_OK = 0,
    ORB_ERR_INVALID = -1,      /* bad argument (null pointer, size out of range ...)            */
    ORB_ERR_CUDA = -2,         /* CUDA runtime error or no device                               */
    ORB_ERR_TOO_SMALL = -3,    /* image so small that the reference divides by zero
                                  (ref: src/ORBextractor.cc:530-532,753-756)                    */
    ORB_ERR_CAPACITY = -4,     /* caller buffer or internal pool too small; nothing truncated
                                  silently                                                      */
    ORB_ERR_EMPTY = -5,        /* stereo: no accepted match (ref: src/Frame.cc:630 indexes an
                                  empty vector there)                                           */
    ORB_ERR_NCCL = -6
} orb_status;

enum { ORB_HOST = 0, ORB_DEVICE = 1 };

/* ref: cv::KeyPoint as filled by src/ORBextractor.cc:786-813,1035-1043 (28 bytes). */
typedef struct orb_keypoint {
    float x, y;        /* pt, level-0 pixel coordinates (pt *= scale, ref :1040)             */
    float size;        /* int(31 * scale[octave]) stored as float (ref :804,812)             */
    float angle;       /* degrees in [0,360), cv::fastAtan2 (ref :97)                        */
    float response;    /* FAST score V-1                                                     */
    int32_t octave;    /* pyramid level                                                      */
    int32_t class_id;  /* -1                                                                 */
} orb_keypoint;

/* ref: ORBextractor::ORBextractor(int,float,int,int,int), include/ORBextractor.h:52-53,
 * configured from the ORBextractor.* yaml keys (src/Tracking.cc:449-464). */
typedef struct orb_params {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;       /* 1..ORB_MAX_LEVELS                                              */
    int32_t ini_th_fast;
    int32_t min_th_fast;
    int32_t max_width;     /* largest image this handle will see (buffers are sized once)    */
    int32_t max_height;
    int32_t max_batch;     /* frames processed per orb_extract_batch call (>=1)              */
    int32_t device;        /* CUDA ordinal, -1 = current device                              */
    int32_t blur_taps;     /* 0: OpenCV>=3.4/4.x 8.8 fixed-point taps {18,34,48,56,48,34,18} (pinned against cv2 4.13.0);
                              1: OpenCV 2.4 rounded-float taps {18,34,49,55,49,34,18} -- UNVERIFIED: no OpenCV 2.4
                                 exists on the build box, the taps are derived from its source, not from its output */
} orb_params;

#define ORB_MAX_LEVELS 16
#define ORB_EDGE_THRESHOLD 19   /* ref: src/ORBextractor.cc:71 */

typedef struct orb_extractor orb_extractor;

/* ---- extractor ------------------------------------------------------------------------- */

/* ref: constructor src/ORBextractor.cc:404-460 (scale tables, per-level quotas, umax). */
int orb_extractor_create(const orb_params *params, orb_extractor **out);
void orb_extractor_destroy(orb_extractor *ex);

/* ref: getters include/ORBextractor.h:64-86.  Each array has nlevels entries. */
int orb_extractor_tables(const orb_extractor *ex, float *scale, float *inv_scale,
                         float *sigma2, float *inv_sigma2, int32_t *features_per_level);

/* ref: ORBextractor::operator()(InputArray image, InputArray mask, vector<KeyPoint>&,
 * OutputArray descriptors), src/ORBextractor.cc:985-1045.  8-bit single channel image,
 * `pitch` bytes per row.  The mask is ignored by the reference (include/ORBextractor.h:60)
 * and has no parameter here.  Writes n keypoints (level-major order, exactly the
 * reference's) and n x 32 descriptor bytes; n can exceed nfeatures, so
 * `capacity` should be >= nfeatures + 4*nlevels; ORB_ERR_CAPACITY if n > capacity.
 * An empty image (width or height 0) returns ORB_OK with *n_out = 0 (ref :987-988).  */
int orb_extract(orb_extractor *ex, const uint8_t *image, int width, int height, int pitch,
                int image_loc, orb_keypoint *keypoints, uint8_t *descriptors, int capacity,
                int *n_out);

/* Batched form of the same call for independent frames (frame-sharded throughput, the
 * reference runs one operator() per Frame: src/Frame.cc:252-258).  `images` is one buffer
 * of `batch` frames, frame b at images + b*frame_stride.  Outputs: frame b's keypoints at
 * keypoints + b*capacity, descriptors at descriptors + b*capacity*32, count in n_out[b].
 * image_loc / out_loc choose host or device residency of the caller's buffers.  With
 * device outputs the call is asynchronous on the handle's stream (see orb_extractor_sync). */
int orb_extract_batch(orb_extractor *ex, const uint8_t *images, int batch, int width,
                      int height, int pitch, size_t frame_stride, int image_loc,
                      orb_keypoint *keypoints, uint8_t *descriptors, int capacity,
                      int32_t *n_out, int out_loc);
int orb_extractor_sync(orb_extractor *ex);
/* Device-output batches are asynchronous: overflow conditions they hit are latched on the device.
 * Synchronises, returns ORB_ERR_CAPACITY if any was raised since the last check, and clears it. */
int orb_extractor_check(orb_extractor *ex);
/* Per-stage device time, measured with CUDA events on the handle's stream between the stage
 * launches of orb_extract_batch.  Stages: 0 host->device copy of the frames, 1 pyramid,
 * 2 FAST cells, 3 quadtree, 4 orientation+blur+descriptor, 5 device->host copy.
 * set_profiling(1) resets the accumulators; stage_times returns accumulated milliseconds over
 * `calls` batches.                                                                          */
#define ORB_N_STAGES 6
int orb_extractor_set_profiling(orb_extractor *ex, int on);
int orb_extractor_stage_times(orb_extractor *ex, double *ms, int64_t *calls);
/* cudaStream_t of the handle as an opaque pointer (for event timing on the launching stream). */
void *orb_extractor_stream(orb_extractor *ex);
/* Number of kernels this handle has launched so far (bench.py's gpu_launches). */
int64_t orb_extractor_launch_count(const orb_extractor *ex);
/* Number of extractions that ran as one CUDA-graph launch: a call that repeats the previous call's arguments (same
 * device buffers, geometry, batch) is captured once and replayed from then on; kernels inside a replay are counted
 * by orb_extractor_launch_count as well.  ORBGPU_NO_GRAPH=1 in the environment keeps the plain launches. */
int64_t orb_extractor_graph_replays(const orb_extractor *ex);

/* ref: public member std::vector<cv::Mat> mvImagePyramid (include/ORBextractor.h:88), read
 * by Frame::ComputeStereoMatches (src/Frame.cc:476,566,578,583).  Geometry of level `level`
 * of batch slot `slot` after the last extract call: interior size, and the device pointer /
 * pitch of the interior's first pixel.  The plane carries the reference's 19-pixel
 * BORDER_REFLECT_101 frame (ref :1052-1068) around the interior.                          */
int orb_pyramid_level(const orb_extractor *ex, int slot, int level, const uint8_t **dev_interior,
                      int *width, int *height, int *pitch);
/* Host copy of that level.  with_border=1 copies the (w+38)x(h+38) padded plane.          */
int orb_pyramid_download(orb_extractor *ex, int slot, int level, int with_border,
                         uint8_t *dst, int dst_pitch);

/* Stage outputs for parity tests (not used by the shims).
 * Candidates = what ComputeKeyPointsOctTree feeds DistributeOctTree (ref :786-793): window-
 * relative x,y and FAST response, in the reference's order (cells row-major, row-major
 * inside a cell).                                                                          */
typedef struct orb_candidate { int16_t x, y; int32_t response; } orb_candidate;
int orb_debug_candidates(orb_extractor *ex, int slot, int level, orb_candidate *out,
                         int capacity, int *n_out);
/* 7x7 sigma=2 Gaussian of a level exactly as ref :1024-1026 (clone + GaussianBlur,
 * BORDER_REFLECT_101), produced by the same device routine the descriptor kernel uses.    */
int orb_debug_blur_level(orb_extractor *ex, int slot, int level, uint8_t *dst, int dst_pitch);

/* ---- Hamming matching ------------------------------------------------------------------ */

/* ref: ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1650-1666.  n pairs of 32-byte
 * descriptors -> n distances.  (Device kernel; the host shim's static DescriptorDistance
 * for a single pair is plain popcount.)                                                    */
int orb_descriptor_distance(const uint8_t *a, const uint8_t *b, int n, int32_t *dist, int loc);

/* Brute-force top-2 (best and second best with strict '<', first candidate wins ties; the
 * inner loop shared by ref src/ORBmatcher.cc:205-226,573-599, src/Marker.cc:314-349).
 * For each of nq queries scans db rows [0,ndb) in order.  Outputs are 4 arrays of nq:
 * best distance/index and second distance/index (256 / -1 when absent).  idx_base is added
 * to reported indices (keyframe-range shards).  loc applies to all buffers.               */
int orb_hamming_top2(const uint8_t *queries, int nq, const uint8_t *db, int64_t ndb,
                     int64_t idx_base, int32_t *best_dist, int64_t *best_idx,
                     int32_t *second_dist, int64_t *second_idx, int loc);

/* The scan of orb_hamming_top2 for callers that cannot allocate device memory.  (No counterpart in the reference: its
 * matchers run on the CPU and never own GPU memory.)
 *
 * Use it from code that may not call cudaMalloc while it runs: a query made inside a real-time tracking loop whose memory
 * was fixed at start-up, a host that gave the whole GPU heap to its own pool, a process near its memory limit where an
 * allocation in mid-frame can fail or stall the other streams.  orb_hamming_top2 grows a per-thread scratch block on
 * demand; this variant allocates nothing and calls no allocating routine.  Its chunk partials live in a 6 MiB array that
 * is part of the library image: every GPU that loads liborbgpu holds those 6 MiB until the process exits, used or not.
 *
 * All five buffers are DEVICE pointers on the current device; the result is bit-identical to orb_hamming_top2
 * (256 / -1 where absent, idx_base added).  The two kernels are enqueued on `cuda_stream` (a cudaStream_t; NULL = the
 * default stream) and the call returns without synchronising: wait on that stream before reading the outputs.
 * Limits: nq <= ORB_TOP2_STATIC_MAX_QUERIES, else ORB_ERR_CAPACITY and nothing is launched; ndb is unlimited.
 * The first call on each device looks up the array's address on that device (the device must be current, as for every
 * call); make that first call during set-up, before any stream capture is begun.
 * The array is shared by every call on a device and keeps the previous call's partials; each call overwrites all the
 * entries it reads, so nothing needs clearing -- but two calls whose kernels can overlap on one GPU (different streams or
 * threads without an ordering between them) race on it.  Serialise them on one stream or with events.               */
#define ORB_TOP2_STATIC_MAX_QUERIES 8192
#define ORB_TOP2_STATIC_MAX_PARTIALS (ORB_TOP2_STATIC_MAX_QUERIES * 32)   /* x 24 bytes = 6 MiB of scratch */
int orb_hamming_top2_static(const uint8_t *dev_queries, int nq, const uint8_t *dev_db, int64_t ndb,
                            int64_t idx_base, int32_t *dev_best_dist, int64_t *dev_best_idx,
                            int32_t *dev_second_dist, int64_t *dev_second_idx, void *cuda_stream);

/* Merge of per-shard top-2 lists in shard order, identical to one scan over the
 * concatenated database.  parts laid out [shard][nq].                         */
int orb_top2_merge(int nshards, int nq, const int32_t *best_dist, const int64_t *best_idx,
                   const int32_t *second_dist, const int64_t *second_idx,
                   int32_t *out_best_dist, int64_t *out_best_idx, int32_t *out_second_dist,
                   int64_t *out_second_idx, int loc);

/* Measured popc32 issue rate of the device (Gpopc32/s), the matching roofline denominator
 * (SURVEY 8d).                                                                             */
int orb_measure_popc_rate(double *gpopc_per_s);

/* ---- frame feature index + tracking matchers ------------------------------------------- */

/* ref: the parts of Frame that the matchers read: mvKeysUn, mDescriptors, mvuRight, the
 * 64x48 grid (include/Frame.h:39-40,161-163; src/Frame.cc:235-250,332-397) and the
 * static image bounds / grid scales.                                                       */
typedef struct orb_frame_desc {
    int32_t n;                     /* number of keypoints                                  */
    const orb_keypoint *keys_un;   /* mvKeysUn (host)                                      */
    const uint8_t *descriptors;    /* n x 32 (host)                                        */
    const float *u_right;          /* mvuRight or NULL (= all -1)                          */
    float min_x, max_x, min_y, max_y;      /* mnMinX ...                                   */
    int32_t nlevels;
    const float *scale_factors;    /* mvScaleFactors, nlevels entries                      */
    int32_t keyframe_view;         /* 1 = window queries see the bounds as KeyFrame does: its
                                      mnMinX/mnMinY are `int` copies of the Frame's floats
                                      (KeyFrame.h:190-193, KeyFrame.cc:41-46) while mGrid and the
                                      cell scales are the Frame's own (KeyFrame.cc:55-62)     */
} orb_frame_desc;

typedef struct orb_frame orb_frame;
int orb_frame_create(const orb_frame_desc *desc, orb_frame **out);
void orb_frame_destroy(orb_frame *f);

/* The same index built ON THE DEVICE from the results the extractor's last call left in HBM, with no host round trip
 * (ref: the Frame constructors run ExtractORB, UndistortKeyPoints and AssignFeaturesToGrid back to back,
 * src/Frame.cc:83-95,196-232).  slot = batch slot of the last orb_extract / orb_extract_batch / orb_stereo_frame call of
 * `ex` (0 = left, 1 = right for orb_stereo_frame).  The bounds are Frame::mnMinX.. (ComputeImageBounds, :441-468); key points
 * are undistorted on the device when ndist > 0 and dist_coef[0] != 0 (`mvKeysUn = mvKeys` otherwise, :411-415).  u_right =
 * mvuRight (n floats, host or device per u_right_loc; NULL = monocular, all -1; orb_extractor_stereo_result gives the device
 * array orb_stereo_frame left).  Asynchronous on the new frame's stream, ordered after the extraction; the extractor may
 * start its next frame only after the first matcher call on this frame returned, or after orb_frame_sync.
 * Lifetime: the call reads the key points / descriptors / counts where the extraction LEFT them.  For an extraction with
 * ORB_DEVICE outputs those are the caller's own device buffers -- keep them allocated and unmodified (no free, no reuse as
 * the output of another call) over the same span; for ORB_HOST outputs they are the handle's internal buffers, which the
 * handle's next extraction overwrites.  Only the LAST extraction of `ex` can be indexed: a second orb_extract* on `ex`
 * before this call makes the slot refer to the newer result.  Frames borrow
 * their stream and buffers from a pool: creating one per camera frame allocates nothing once the pool is warm.        */
typedef struct orb_frame_source {
    float min_x, max_x, min_y, max_y;
    float fx, fy, cx, cy;
    float dist_coef[5];
    int32_t ndist;
    const float *u_right;
    int32_t u_right_loc;
    int32_t keyframe_view;
} orb_frame_source;
int orb_frame_from_extractor(orb_extractor *ex, int slot, const orb_frame_source *src, orb_frame **out);
int orb_frame_sync(orb_frame *f);
int orb_frame_size(const orb_frame *f);                   /* number of key points                                       */
void *orb_frame_stream(orb_frame *f);
int64_t orb_frame_launch_count(const orb_frame *f);
/* Of the last matcher call on this frame: candidate (query, feature) pairs whose distance was computed, and the number of
 * claim-resolution sweeps (see csrc/orb_track.cu).                                                                   */
int orb_frame_last_stats(const orb_frame *f, int64_t *candidates, int32_t *sweeps);
/* Device arrays orb_stereo_frame left in the handle: mvuRight / mvDepth of the left frame (capacity floats each).       */
int orb_extractor_stereo_result(orb_extractor *ex, const float **dev_u_right, const float **dev_depth);

/* ref: Frame::GetFeaturesInArea, src/Frame.cc:332-385 -- indices in the reference's
 * traversal order (ix outer, iy inner, insertion order).                                   */
int orb_frame_features_in_area(orb_frame *f, float x, float y, float r, int min_level,
                               int max_level, int32_t *indices, int capacity, int *n_out);

/* Generic windowed matcher: the loop every projection matcher of the reference shares
 * (ORBmatcher.cc:51-126, 1354-1448, 1491-1578, 312-400 and the Fuse / SearchBySim3 loops):
 * for each query in order -> GetFeaturesInArea window -> skip claimed features -> optional
 * stereo gate -> best (and second) Hamming distance in traversal order -> threshold -> assign.
 * Assignments made by earlier queries are visible to later ones exactly as in the reference:
 * distances are computed in parallel, claims are resolved in query order on the device.      */
typedef struct orb_window_queries {
    int32_t n;
    const uint8_t *valid;                  /* 0 = the reference `continue`s before the search    */
    const float *x, *y, *radius;           /* GetFeaturesInArea(x, y, r, ...)                    */
    const int32_t *min_level, *max_level;  /* its level arguments                                */
    int32_t strict_levels;                 /* 0: Frame::GetFeaturesInArea rule (filter only if
                                              minLevel>0 || maxLevel>=0, Frame.cc:353-373);
                                              1: always skip octave<min || octave>max (the in-loop
                                              filter of ORBmatcher.cc:378-381, 912-915)          */
    const float *stereo_ref, *stereo_tol;  /* both NULL = no stereo gate; else a candidate with
                                              uRight>0 is skipped if |stereo_ref-uRight|>tol     */
    const uint8_t *descriptors;            /* n x 32                                             */
    const float *angle;                    /* query keypoint angle (rotation histogram)          */
    const uint8_t *claims;                 /* NULL = every accepted query blocks its feature for
                                              later queries; else only where claims[q] != 0
                                              (MapPoint::Observations()>0, ORBmatcher.cc:87-89)  */
} orb_window_queries;
enum { ORB_SELECT_BEST = 0,         /* accept if best <= th_dist (ORBmatcher.cc:1429, 1558, 394)  */
       ORB_SELECT_RATIO_LEVEL = 1   /* best <= th_dist and not (same octave && best > ratio*second),
                                       ORBmatcher.cc:118-125                                     */ };
/* claimed[i] != 0: feature i already holds a blocking map point on entry.  assigned[i] (out) =
 * index of the query whose map point ends up in mvpMapPoints[i]; -1 = the call never wrote
 * mvpMapPoints[i]; -2 = the call wrote it and the rotation filter then reset it to NULL.  With
 * check_orientation the 3-maxima rotation filter (ORBmatcher.cc:1451-1470) is applied; nmatches
 * follows the reference's counting (an overwritten assignment is still counted, F7).          */
int orb_search_windowed(orb_frame *f, const orb_window_queries *q, const uint8_t *claimed, int select,
                        int th_dist, float nnratio, int check_orientation, int32_t *assigned,
                        int *nmatches);

/* ref: ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:405-523.
 * prev_matched: n1 x 2 floats, updated in place (ref :518-520).  matches12: n1 ints.       */
int orb_search_for_initialization(orb_frame *f1, orb_frame *f2, float *prev_matched,
                                  int32_t *matches12, int window_size, float nnratio,
                                  int check_orientation, int *nmatches);

/* ref: ORBmatcher::SearchByProjection(Frame&, const vector<MapPoint*>&, th),
 * src/ORBmatcher.cc:45-129.  Map points are passed flat, as Frame::isInFrustum leaves them.  */
typedef struct orb_track_points {
    int32_t n;
    const uint8_t *in_view;        /* mbTrackInView && !isBad()                            */
    const float *proj_x, *proj_y, *proj_xr;   /* mTrackProjX/Y/XR                          */
    const int32_t *level;          /* mnTrackScaleLevel                                    */
    const float *view_cos;         /* mTrackViewCos                                        */
    const uint8_t *descriptors;    /* GetDescriptor(), n x 32                              */
    const uint8_t *obs_positive;   /* Observations()>0 per point; NULL = all               */
} orb_track_points;
/* claimed[i] != 0 : F.mvpMapPoints[i] && Observations()>0 on entry (ref :87-89).
 * assigned[i]: out, index of the map point written to F.mvpMapPoints[i] or -1.             */
int orb_search_by_projection_points(orb_frame *f, const orb_track_points *pts,
                                    const uint8_t *claimed, float th, float nnratio,
                                    int32_t *assigned, int *nmatches);

/* ref: ORBmatcher::SearchByProjection(Frame &Cur, const Frame &Last, th, bMono),
 * src/ORBmatcher.cc:1331-1473.  The pose algebra (ref :1341-1379) stays with the caller: each
 * Last feature i carries its projected (u,v), 1/z, octave, angle and descriptor.
 * forward/backward = bForward/bBackward of ref :1351-1352.                                  */
typedef struct orb_last_points {
    int32_t n;
    const uint8_t *valid;          /* pMP && !mvbOutlier && invzc>=0 && inside image bounds */
    const float *u, *v, *invz;
    const int32_t *octave;         /* LastFrame.mvKeys[i].octave                            */
    const float *angle;            /* LastFrame.mvKeysUn[i].angle                           */
    const uint8_t *descriptors;
    const uint8_t *obs_positive;   /* Observations()>0 per point; NULL = all               */
} orb_last_points;
int orb_search_by_projection_last(orb_frame *cur, const orb_last_points *pts,
                                  const uint8_t *claimed, float th, int forward, int backward,
                                  float mbf, int check_orientation, int32_t *assigned,
                                  int *nmatches);

/* ref: ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...), src/ORBmatcher.cc:159-288, and the
 * KeyFrame-KeyFrame form :525-658.  The two DBoW2::FeatureVectors are passed as CSR over the
 * common node ids in ascending order (the merge-join of ref :180-264 done by the caller):
 * group g holds KF feature indices kf_idx[kf_off[g]..kf_off[g+1]) and F indices
 * f_idx[f_off[g]..f_off[g+1]).  kf_valid[i] = map point exists and is not bad; f_valid (NULL =
 * all) = second-side validity of the KF-KF form (:577-583).  strict_low=1 selects the KF-KF
 * threshold `best < TH_LOW` (:601) instead of `<=` (:228).  matches[j] (size f->n) = KF feature
 * index matched to F feature j, or -1.                                                      */
int orb_search_by_bow(orb_frame *kf, orb_frame *f, int ngroups, const int32_t *kf_off,
                      const int32_t *kf_idx, const int32_t *f_off, const int32_t *f_idx,
                      const uint8_t *kf_valid, const uint8_t *f_valid, int strict_low,
                      float nnratio, int check_orientation, int32_t *matches, int *nmatches);

/* ref: ORBmatcher::ComputeThreeMaxima + the rotation binning, src/ORBmatcher.cc:238-245,
 * 1604-1645 (factor = 1/HISTO_LENGTH quirk kept).  keep[i]=1 if pair i survives.           */
int orb_rot_hist_filter(const float *angle_a, const float *angle_b, int n, uint8_t *keep);

/* ---- mapping / loop-closing matchers (LocalMapping and LoopClosing threads) ------------- */

/* ref: the candidate loops of ORBmatcher::Fuse(KeyFrame*, vector<MapPoint*>&, th)
 * (src/ORBmatcher.cc:895-951), Fuse(KeyFrame*, Scw, ...) (:1052-1072) and SearchBySim3
 * (:1191-1213, :1271-1293).  Every query is independent (these loops never look at what an
 * earlier query took): window -> in-loop level filter -> optional reprojection-error gate ->
 * closest descriptor, first one in traversal order on ties (`dist<bestDist`).  The caller keeps
 * the projection algebra, the final threshold (TH_LOW / TH_HIGH) and the map bookkeeping that
 * follows (:954-973, :1075-1088, :1301-1326).
 * best_idx[q] = -1, best_dist[q] = INT32_MAX when the window holds no admissible feature.      */
typedef struct orb_reproj_gate {   /* ref :917-945 */
    const float *ur;               /* per query: u - bf*invz                                   */
    const float *inv_level_sigma2; /* pKF->mvInvLevelSigma2, one per pyramid level             */
    double chi2_mono, chi2_stereo; /* 5.99 and 7.8 in the reference (compared in double)       */
} orb_reproj_gate;
int orb_search_window_best(orb_frame *f, const orb_window_queries *q, const orb_reproj_gate *gate,
                           int32_t *best_idx, int32_t *best_dist);

/* ref: ORBmatcher::SearchForTriangulation, src/ORBmatcher.cc:660-826, with
 * CheckDistEpipolarLine :140-157.  The FeatureVector merge-join (:695-786) is passed as CSR over
 * the common vocabulary nodes like orb_search_by_bow.  usable1[i] / usable2[i] = the key frame
 * holds no map point at that feature (:704-708, :726-730).  F12: 9 floats, row major.  (ex, ey):
 * projection of KF1's centre into KF2 (:667-673).  level_sigma2: pKF2->mvLevelSigma2;
 * pKF2->mvScaleFactors comes from kf2's orb_frame.  matches12[i1] (size kf1->n) = matched feature
 * of KF2 or -1 (the reference's vMatches12; vMatchedPairs is its non-negative entries in order).
 * This fork never sets vbMatched2 (:778-781), so a KF2 feature may serve several KF1 features;
 * that is reproduced.                                                                           */
int orb_search_for_triangulation(orb_frame *kf1, orb_frame *kf2, int ngroups, const int32_t *off1,
                                 const int32_t *idx1, const int32_t *off2, const int32_t *idx2,
                                 const uint8_t *usable1, const uint8_t *usable2, const float *F12,
                                 float ex, float ey, const float *level_sigma2, int only_stereo,
                                 int check_orientation, int32_t *matches12, int *nmatches);

/* ---- stereo ---------------------------------------------------------------------------- */

/* ref: Frame::ComputeStereoMatches, src/Frame.cc:471-643.  left/right are the two extractor
 * handles whose pyramids (slot `slot`) hold the current pair; keypoints/descriptors are the
 * outputs of the two orb_extract calls (host).  mb is Frame::mb (= mbf/fx) as the reference
 * intends it.  Outputs mvuRight / mvDepth (n_left floats, -1 where unmatched).             */
int orb_stereo_match(orb_extractor *left, orb_extractor *right, int slot,
                     const orb_keypoint *keys_left, const uint8_t *desc_left, int n_left,
                     const orb_keypoint *keys_right, const uint8_t *desc_right, int n_right,
                     float mbf, float mb, float *u_right, float *depth, int *n_matched);

/* ref: the stereo Frame constructor, src/Frame.cc:62-111: ExtractORB on the left and the right image (two threads, :83-86)
 * followed by ComputeStereoMatches (:95).  One call, one handle (created with max_batch >= 2), one host synchronisation:
 * both host images are uploaded into pyramid slots 0 (left) and 1 (right), extracted as one batch of two, associated on the
 * device-resident keypoints, and all results come back together.  Outputs are those of orb_extract (per camera, `capacity`
 * entries each) and of orb_stereo_match (u_right / depth: n_left floats).  Same error behaviour: ORB_ERR_CAPACITY with
 * n_left / n_right holding the required sizes, ORB_ERR_EMPTY when no match survives (ref :630).                      */
int orb_stereo_frame(orb_extractor *ex, const uint8_t *image_left, const uint8_t *image_right, int width, int height,
                     int pitch_left, int pitch_right, orb_keypoint *keys_left, uint8_t *desc_left, int *n_left,
                     orb_keypoint *keys_right, uint8_t *desc_right, int *n_right, int capacity, float mbf, float mb,
                     float *u_right, float *depth, int *n_matched);

/* Batched / device-resident form: pair b uses pyramid slot left_slot+b of `left` and
 * right_slot+b of `right` (the two may be the same handle holding 2*batch frames), keypoints at
 * keys_*[b*capacity], counts n_*[b].  loc = where all caller arrays live; with ORB_DEVICE the
 * call is asynchronous on left's stream.  n_matched[b] = matches surviving the median test.   */
int orb_stereo_match_batch(orb_extractor *left, int left_slot, orb_extractor *right, int right_slot,
                           int batch, const orb_keypoint *keys_left, const uint8_t *desc_left,
                           const int32_t *n_left, const orb_keypoint *keys_right,
                           const uint8_t *desc_right, const int32_t *n_right, int capacity,
                           float mbf, float mb, float *u_right, float *depth, int32_t *n_matched,
                           int loc);

/* ---- sharded descriptor database (relocalisation brute force) --------------- */

typedef struct orb_db orb_db;
/* Shard [row_begin,row_end) of a global database of 32-byte descriptors, resident in HBM. */
int orb_db_create(int64_t row_begin, int64_t row_end, int device, orb_db **out);
void orb_db_destroy(orb_db *db);
int orb_db_upload(orb_db *db, int64_t first_row, const uint8_t *rows, int64_t nrows, int loc);
/* Deterministic synthetic fill (splitmix64 of the global row index) directly on the device. */
int orb_db_fill_synthetic(orb_db *db, uint64_t seed);
int orb_db_row(orb_db *db, int64_t global_row, uint8_t *out32);
uint8_t *orb_db_device_ptr(orb_db *db);
void *orb_db_stream(orb_db *db);                 /* cudaStream_t the shard's kernels run on */
int64_t orb_db_launch_count(const orb_db *db);
/* Local top-2 of nq queries against the shard; outputs are device or host per loc.        */
int orb_db_query(orb_db *db, const uint8_t *queries, int nq, int32_t *best_dist,
                 int64_t *best_idx, int32_t *second_dist, int64_t *second_idx, int loc);

/* The same scan over a shard made of keyframes of `rows_per_keyframe` consecutive rows (BASELINE config 5: 50 000 keyframes x
 * 2000 descriptors; the shard must begin and end on keyframe boundaries).  Besides the shard's top-2 it returns, per
 * keyframe of the shard, how many queries that keyframe matches by the reference's acceptance rule for a (query, keyframe)
 * pair -- best distance inside the keyframe <= th_low and (float)best < ratio * (float)second, second = 256 with a single row
 * (ref: src/ORBmatcher.cc:216-231, TH_LOW = 50, mfNNratio): the relocalisation candidate score.  kf_counts holds
 * (row_end - row_begin) / rows_per_keyframe int32; counts are local to a shard, ranks concatenate them.   */
int orb_db_query_keyframes(orb_db *db, const uint8_t *queries, int nq, int rows_per_keyframe, int th_low, float ratio,
                           int32_t *best_dist, int64_t *best_idx, int32_t *second_dist, int64_t *second_idx,
                           int32_t *kf_counts, int loc);

/* Device-resident merge of [nshards][nq] partial lists (e.g. the result of an all-gather across ranks), asynchronous
 * on the shard's stream -- the exchange step of the sharded search.                                  */
int orb_db_merge_top2(orb_db *db, int nshards, int nq, const int32_t *best_dist, const int64_t *best_idx,
                      const int32_t *second_dist, const int64_t *second_idx, int32_t *out_best_dist,
                      int64_t *out_best_idx, int32_t *out_second_dist, int64_t *out_second_idx);

/* ---- RGB-D depth association ------------------------------------------------------------------ */

/* ref: Frame::ComputeStereoFromRGBD (src/Frame.cc:645-667), the RGB-D Frame constructor's counterpart of ComputeStereoMatches
 * (:151), with `imDepth.convertTo(imDepth, CV_32F, mDepthMapFactor)` (src/Tracking.cc:553-554) folded in: for key point i,
 * d = depth(v, u) at (int)mvKeys[i].pt; when d > 0: depth[i] = d and u_right[i] = mvKeysUn[i].pt.x - mbf / d, else both -1.
 * depth_type 0: CV_32F map (multiplied by depth_factor unless it is exactly 1), 1: CV_16U map (d = (float)raw * depth_factor,
 * OpenCV's convertTo arithmetic).  loc applies to keys, keys_un, the depth map and the outputs; ORB_DEVICE calls are
 * asynchronous on the thread's call stream.                                                                          */
int orb_depth_from_rgbd(const orb_keypoint *keys, const orb_keypoint *keys_un, int n, const void *depth_map, int depth_type,
                        int width, int height, int pitch_bytes, float depth_factor, float mbf, int loc, float *u_right,
                        float *depth);

/* ---- marker / AR brute-force matching ------------------------------------------ */

/* ref: Marker::searchMatches (src/Marker.cc:314-349) and naive_nn_search2 (AR-1.3/src/ORBMatcher.cpp:70-102): for every
 * query descriptor (keys2/descp2) the best and second-best train descriptor (keys1/descp1) with strict `<` updates in
 * train order, accepted when `min_dist <= (unsigned int)(sec_dist * ratio) && min_dist <= max_dist` (0.8 and 50 in the
 * reference; sec_dist stays INT_MAX with fewer than two train rows).  train_idx[i] = matched train row (DMatch.trainIdx)
 * or -1, dist[i] = min_dist (DMatch.distance; INT_MAX with no train row).  The caller pushes DMatch(i, train_idx[i], 0,
 * dist[i]) for the accepted i in ascending order.                                                                 */
int orb_search_matches_ratio(const uint8_t *train, int n_train, const uint8_t *query, int n_query, double ratio,
                             int max_dist, int loc, int32_t *train_idx, int32_t *dist);

/* ref: the `minD` / `maxD` bookkeeping inside naive_nn_search2 (AR-1.3/src/ORBMatcher.cpp:92-93), which ORBMatcher::Match uses
 * for its `distance <= 0.5 * maxD` filter (:147-152): min_max[0] / [1] = the smallest / largest Hamming distance over ALL
 * (query, train) pairs (INT_MAX / -1 when there is no pair).  The reference keeps them in globals across calls (initial 100 / 0,
 * `if (dist <= minD) minD = dist; if (dist > maxD) maxD = dist;`): the caller merges the same way.  min_max is a host array. */
int orb_hamming_minmax(const uint8_t *train, int n_train, const uint8_t *query, int n_query, int loc, int32_t *min_max);

/* ---- OpenCV-style cv::ORB (Harris score) for the marker pipeline --------------- */

/* ref: `ORB orb; orb(mImage1, Mat(), mvKeys1, mDescriptors1);` in Marker::setTargetImage (src/Marker.cc:83-84) and the same
 * call on the camera image in Marker::Match (:107-108).  This is OpenCV's own ORB (features2d/src/orb.cpp), NOT
 * ORBextractor: whole-layer FAST-9/16 with 3x3 non-maximum suppression at `fast_threshold`, key points closer than
 * `edge_threshold` to the layer edge dropped, the 2*quota best by FAST score kept (KeyPointsFilter::retainBest: everything
 * tied with the boundary response stays), Harris response (7x7 block, k = 0.04) of those, the quota best by Harris kept,
 * intensity-centroid angle, 7x7 sigma-2 blur, 256 rotated tests (double cos/sin of the angle).  Layers: level l has
 * cvRound(cols / (float)pow(scale, l)) columns, built by a cascade of resize(INTER_LINEAR_EXACT).  Arithmetic and key-point
 * ORDER (what std::nth_element + std::partition leave behind) are pinned against cv2 4.13.0's ORB_create(); OpenCV 2.4's
 * ORB (the version the reference was built with) resizes with INTER_LINEAR and is UNVERIFIED here (no OpenCV 2.4 on the
 * build box).  Defaults of `ORB orb;` : 500, 1.2f, 8, 31, score_type 0 (HARRIS_SCORE; 1 = FAST_SCORE), fast_threshold 20;
 * firstLevel 0, WTA_K 2 and patchSize 31 are fixed.                                                                   */
typedef struct orb_cvorb_params {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;         /* 1..ORB_MAX_LEVELS */
    int32_t edge_threshold;  /* >= 31 (the descriptor and moment windows must stay inside the layer) */
    int32_t score_type;      /* 0 HARRIS_SCORE, 1 FAST_SCORE */
    int32_t fast_threshold;
    int32_t max_width;       /* largest image this handle will see */
    int32_t max_height;
    int32_t device;          /* CUDA device ordinal, -1 = current */
} orb_cvorb_params;
typedef struct orb_cvorb orb_cvorb;
int orb_cvorb_create(const orb_cvorb_params *params, orb_cvorb **out);
void orb_cvorb_destroy(orb_cvorb *h);
/* capacity that always suffices for an image of the handle's maximum size (FAST ties can keep more than nfeatures) */
int orb_cvorb_capacity(const orb_cvorb *h);
/* `cv::ORB::operator()(image, noArray(), keypoints, descriptors)`: 8-bit single-channel image (image_loc = ORB_HOST or
 * ORB_DEVICE), results to host arrays; *n_out key points in OpenCV's order, descriptors n_out x 32.  An empty image or
 * one without key points gives *n_out = 0.  ORB_ERR_CAPACITY (with *n_out = the needed size) if capacity is too small. */
int orb_cvorb_detect_and_compute(orb_cvorb *h, const uint8_t *image, int width, int height, int pitch, int image_loc,
                                 orb_keypoint *keypoints, uint8_t *descriptors, int capacity, int *n_out);
/* debugging / tests: layer `level` of the last call as the descriptors sampled it (blurred), and its size; level | 256 selects the
 * unblurred layer, level | 512 the FAST score map (score = arc strength - 1 inside the region the detector evaluates, else stale) */
int orb_cvorb_level(orb_cvorb *h, int level, uint8_t *dst, int dst_pitch, int *width, int *height);
/* tests: KeyPointsFilter::retainBest alone, on `n` host responses: order_out receives the surviving original indices in the
 * order std::nth_element + std::partition leave them (depth_limit < 0: libstdc++'s own 2*floor(log2 n); >= 0 forces the
 * heap-select fallback of __introselect after that many partitions).                                                */
int orb_debug_retain_best(const float *responses, int n, int n_points, int depth_limit, int32_t *order_out, int *n_out);
void *orb_cvorb_stream(orb_cvorb *h);
int64_t orb_cvorb_launch_count(const orb_cvorb *h);

/* ---- colour -> grey in front of the extractor ---------------------------------- */

/* ref: cvtColor(mImGray, mImGray, CV_RGB2GRAY | CV_BGR2GRAY | CV_RGBA2GRAY | CV_BGRA2GRAY) in
 * Tracking::GrabImageStereo / GrabImageRGBD / GrabImageMonocular (src/Tracking.cc:510-526,545-556,567-577).  8-bit
 * interleaved input with `channels` = 3 or 4; rgb_order = 1 for RGB(A) (mbRGB), 0 for BGR(A).  variant 0 = OpenCV
 * 3.x/4.x arithmetic `(R*9798 + G*19235 + B*3735 + 16384) >> 15` (pinned against cv2 4.13.0), variant 1 = OpenCV
 * 2.4's `(R*4899 + G*9617 + B*1868 + 8192) >> 14` (the version the reference was built with; UNVERIFIED: no OpenCV 2.4
 * on the build box).  `batch` images, image b at src + b*src_stride / dst + b*dst_stride.  loc = where src and dst live;
 * `stream` is a cudaStream_t (NULL = the thread's call stream, see "Streams" at the top); ORB_DEVICE calls are asynchronous
 * on it.  To hand the grey image to orb_extract_batch(image_loc = ORB_DEVICE) without touching the host pass
 * stream = orb_extractor_stream(ex): the conversion is then ordered in front of the extraction.                      */
int orb_cvt_color_gray(const uint8_t *src, int width, int height, int src_pitch, size_t src_stride, int channels,
                       int rgb_order, int variant, int batch, uint8_t *dst, int dst_pitch, size_t dst_stride, int loc,
                       void *stream);

/* ref: cv::remap(image, rectified, M1, M2, cv::INTER_LINEAR), Examples/ROS/ORB_SLAM2/src/ros_stereo.cc:163-164, with the
 * CV_32F maps of cv::initUndistortRectifyMap (:107-112): the stereo rectification in front of TrackStereo.  8-bit single
 * channel; map_x / map_y hold one source coordinate per destination pixel (map_pitch floats per row).  Arithmetic = OpenCV's
 * fixed-point bilinear remap (1/32-pixel coordinates, 15-bit weights, zero outside the source), pinned against cv2 4.13.0.
 * `batch` images share the maps (image b at src + b*src_stride / dst + b*dst_stride); loc = where src, maps and dst live;
 * `stream` is a cudaStream_t (NULL = the thread's call stream), ORB_DEVICE calls are asynchronous on it; pass
 * orb_extractor_stream(ex) to order the rectification in front of that handle's extraction.                            */
int orb_remap_linear(const uint8_t *src, int src_width, int src_height, int src_pitch, size_t src_stride,
                     const float *map_x, const float *map_y, int map_pitch, uint8_t *dst, int dst_width, int dst_height,
                     int dst_pitch, size_t dst_stride, int batch, int loc, void *stream);

/* ref: Frame::UndistortKeyPoints, src/Frame.cc:409-439: cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) over
 * the key points' pt (everything else of the KeyPoint is copied); `mvKeysUn = mvKeys` when dist_coef[0] == 0.
 * fx, fy, cx, cy are mK's float entries, dist_coef = mDistCoef (k1, k2, p1, p2 [, k3]; src/Tracking.cc:427-438).
 * Arithmetic = OpenCV >= 3.x (double precision, 5 iterations, gives up on a point whose radial factor turns negative);
 * pinned against cv2 4.13.0.  orb_undistort_points is the same on bare (x,y) pairs: the four image corners of
 * Frame::ComputeImageBounds (src/Frame.cc:441-468).  loc = where the arrays live (ORB_DEVICE: asynchronous on the thread's
 * call stream, see "Streams").                                                                                        */
int orb_undistort_keypoints(const orb_keypoint *keys, int n, float fx, float fy, float cx, float cy,
                            const float *dist_coef, int ndist, orb_keypoint *keys_un, int loc);
int orb_undistort_points(const float *xy, int n, float fx, float fy, float cx, float cy,
                         const float *dist_coef, int ndist, float *xy_un, int loc);

/* ref: Frame::isInFrustum(MapPoint*, viewingCosLimit), src/Frame.cc:274-330, for all local map points of a frame in one
 * call (the loop of Tracking::SearchLocalPoints, src/Tracking.cc:1527-1539; the caller keeps its mnLastFrameSeen / isBad
 * pre-checks and passes consider[i] = 0 for the points it skips).  The camera block is the Frame's mRcw (row major), mtcw,
 * mOw, the static intrinsics and image bounds, mbf and mfLogScaleFactor; a point is its GetWorldPos(), GetNormal(),
 * mfMaxDistance and mfMinDistance (the 1.2 / 0.8 factors of Get*DistanceInvariance are applied here).  Outputs are
 * the MapPoint tracking fields the reference fills (:321-327): in_view = mbTrackInView, proj_x/proj_y/proj_xr =
 * mTrackProjX/Y/XR, level = mnTrackScaleLevel (MapPoint::PredictScale, src/MapPoint.cc:357-366: this fork does not clamp
 * it), view_cos = mTrackViewCos -- exactly the arrays orb_track_points takes.  Arithmetic: cv::Mat products and norms
 * accumulate in double and round to float like cv::gemm / cv::norm, float expressions are evaluated without FMA
 * contraction, logf is glibc's algorithm (bit-identical to the host libm on every finite float).                      */
typedef struct orb_frustum_camera {
    float Rcw[9], tcw[3], Ow[3];
    float fx, fy, cx, cy;
    float min_x, max_x, min_y, max_y;
    float mbf, log_scale_factor;
} orb_frustum_camera;
typedef struct orb_frustum_points {
    int32_t n;
    const float *world_pos;      /* [n][3] */
    const float *normal;         /* [n][3] */
    const float *max_distance;   /* [n] mfMaxDistance */
    const float *min_distance;   /* [n] mfMinDistance */
    const uint8_t *consider;     /* [n] or NULL = all */
} orb_frustum_points;
int orb_is_in_frustum(const orb_frustum_camera *cam, const orb_frustum_points *pts, float viewing_cos_limit, int loc,
                      uint8_t *in_view, float *proj_x, float *proj_y, float *proj_xr, int32_t *level, float *view_cos);

/* ref: the projection pre-loops of the mapping / loop-closing matchers -- SearchByProjection(Frame&, KeyFrame*, ...)
 * src/ORBmatcher.cc:1491-1531, SearchByProjection(KeyFrame*, Scw, ...) :311-360, Fuse :846-893 and :1004-1048, SearchBySim3
 * :1150-1226 / :1229-1305: per map point the transform into the camera, the pinhole projection, the image test, the distance
 * range and (where the reference has it) the viewing-angle test, with cv::Mat's rounding (3x3 products and norms accumulate in
 * double).  The sites differ in small ways that change the last bit; each is a rule:
 *   invz_rule          0: (float)(1.0 / z) (`1.0/z`, :1503, :1021, :1166)      1: 1.0f / z (`1/z`, :323, :861)
 *   uv_rule            0: fx*xc*invz + cx (:1505)                               1: x = xc*invz; fx*x + cx (:325-329 and the others)
 *   bounds_rule        0: reject u < min || u > max (Frame bounds, :1508-1511)  1: KeyFrame::IsInImage, min <= u < max (:332)
 *   reject_negative_z  the `z < 0` test (all sites but :1491-1531)
 *   dist_rule          0: |X - Ow|                                              1: |x| in the second camera (SearchBySim3 :1177)
 *   check_view_angle   PO.dot(Pn) < 0.5 * dist rejected (:350-353, :885-888, :1040-1043); needs dist_rule 0 and the normals
 *   two_stage          x = R2 * (R1 * X + t1) + t2 (SearchBySim3: camera a, then the similarity into camera b)
 * max_distance / min_distance are what MapPoint::GetMaxDistanceInvariance / GetMinDistanceInvariance return.  Outputs per point:
 * ok, u, v, u_right = u - bf * invz (:865), dist.  MapPoint::PredictScale(dist, logScaleFactor) stays with the caller (a public
 * method of the reference's class; the raw mfMaxDistance it reads is not public).  loc as in orb_is_in_frustum.             */
typedef struct orb_projection {
    float R1[9], t1[3], R2[9], t2[3], Ow[3];
    float fx, fy, cx, cy, bf;
    float min_x, max_x, min_y, max_y;
    int32_t two_stage, reject_negative_z, invz_rule, uv_rule, bounds_rule, dist_rule, check_view_angle;
} orb_projection;
typedef struct orb_projection_points {
    int32_t n;
    const float *world_pos;      /* [n][3] */
    const float *normal;         /* [n][3] or NULL */
    const float *max_distance;   /* [n] GetMaxDistanceInvariance() */
    const float *min_distance;   /* [n] GetMinDistanceInvariance() */
    const uint8_t *consider;     /* [n] or NULL = all */
} orb_projection_points;
int orb_project_points(const orb_projection *proj, const orb_projection_points *pts, int loc, uint8_t *ok, float *u, float *v,
                       float *u_right, float *dist);

/* ---- map-point descriptor selection ------------------------------------------ */

/* ref: MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:224-285: among the N descriptors observing a map
 * point (the caller gathers `pKF->mDescriptors.row(idx)` of the non-bad key frames in mObservations order, :243-249)
 * all N*N DescriptorDistances, per row the median `vDists[0.5*(N-1)]` of the sorted row (diagonal 0 included), and
 * the row with the smallest median, first row on ties (`median < BestMedian`, :272-275).  Any number of map points
 * per call: point p owns rows [offsets[p], offsets[p+1]).  best_row[p] = index of the chosen row inside its group
 * (-1 for an empty group: the reference returns early and keeps the old descriptor), best_median[p] = its median.
 * loc = where all four arrays live; ORB_DEVICE calls are asynchronous on the thread's call stream (see "Streams").  */
int orb_distinctive_descriptors(const uint8_t *descriptors, const int32_t *offsets, int npoints, int loc,
                                int32_t *best_row, int32_t *best_median);

/* ---- bag-of-words vocabulary (DBoW2 transform) --------------------------- */

/* ref: ORBVocabulary = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> (include/ORBVocabulary.h:29),
 * Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h.  The tree is resident in HBM; Frame::ComputeBoW
 * (src/Frame.cc:400-407) and KeyFrame::ComputeBoW (src/KeyFrame.cc:62-70) become one
 * orb_vocab_transform call whose FeatureVector output is already the CSR that orb_search_by_bow /
 * orb_search_for_triangulation consume.  A vocabulary handle may be shared by threads (the reference
 * calls the const transform() from Tracking, LocalMapping and LoopClosing): calls serialise on it.  */
typedef struct orb_vocab orb_vocab;

/* Flat tree, node 0 = root.  A node's children are the nodes naming it as parent, in ascending node id
 * -- the order loadFromTextFile (:1338-1425) and create() (:596-618) push them, which decides ties
 * in the descent.  A node without children is a leaf (Node::isLeaf, :328).                        */
typedef struct orb_vocab_desc {
    int32_t k, L;                /* branching factor and depth levels (m_k, m_L)                  */
    int32_t scoring;             /* DBoW2::ScoringType   0 L1_NORM 1 L2_NORM 2 CHI_SQUARE 3 KL 4 BHATTACHARYYA 5 DOT_PRODUCT */
    int32_t weighting;           /* DBoW2::WeightingType 0 TF_IDF 1 TF 2 IDF 3 BINARY              */
    int32_t n_nodes;             /* root included                                                  */
    const int32_t *parent;       /* [n_nodes], parent[0] ignored; parent[i] < i is NOT required    */
    const uint8_t *descriptors;  /* [n_nodes][32], row 0 ignored                                   */
    const double *weight;        /* [n_nodes] Node::weight                                         */
    const int32_t *word_id;      /* [n_nodes] Node::word_id (the reference leaves 0 on non-words)  */
} orb_vocab_desc;

int orb_vocab_create(const orb_vocab_desc *desc, int device, orb_vocab **out);
/* ref: TemplatedVocabulary::loadFromTextFile, :1338-1425 (the loader src/System.cc:67-75 uses).  One deliberate
 * difference: the reference's `while(!f.eof())` loop runs once more on the empty read after the final newline and
 * then uses `int pid` / `int nIsLeaf` that no extraction has written (undefined behaviour: the reference's own code,
 * compiled here, either attaches a phantom node to whatever the stack holds or crashes).  Here that empty final read
 * is not a node, so a file loads the same with or without a trailing newline; any other malformed line is
 * ORB_ERR_INVALID.                                                                                               */
int orb_vocab_load_text(const char *path, int device, orb_vocab **out);
void orb_vocab_destroy(orb_vocab *v);
/* info[0..5] = k, L, scoring, weighting, nodes (root included), words */
int orb_vocab_info(const orb_vocab *v, int32_t *info);
/* Read the tree back in the orb_vocab_desc layout (any pointer may be NULL). */
int orb_vocab_nodes(const orb_vocab *v, int32_t *parent, uint8_t *descriptors, double *weight, int32_t *word_id);

/* ref: transform(const TDescriptor&, WordId&, WordValue&, NodeId*, int levelsup), :1217-1255: descend from the
 * root taking at every level the child at the smallest FORB::distance (FORB.cpp:81-101), first child on ties.
 * node[i] = node reached at level L-levelsup (0 = root when that level is <= 0; when the leaf is shallower than
 * that level the reference leaves *nid unset -- here it is the leaf).  descriptors: n rows of 32 bytes.          */
int orb_vocab_transform_features(orb_vocab *v, const uint8_t *descriptors, int n, int loc, int levelsup,
                                 int32_t *word, int32_t *node, double *weight);

/* ref: transform(const vector<TDescriptor>&, BowVector&, FeatureVector&, int levelsup), :1126-1193, with
 * BowVector::addWeight / addIfNotExist / normalize (BowVector.cpp:33-88) and FeatureVector::addFeature
 * (FeatureVector.cpp:30-44).  Frame b's descriptors are rows [b*stride_rows, b*stride_rows + n[b]).  Outputs per
 * frame b at offset b*capacity (fv_off: b*(capacity+1)); capacity >= max n[b].
 *   bow_word/bow_value[0..n_bow)  the std::map<WordId, WordValue> in iteration order (ascending word id); values
 *                                 are accumulated and normalised in double in the reference's order, bit-exact
 *   fv_node[0..n_fv)              ascending node ids; features of group g = fv_feat[fv_off[g]..fv_off[g+1]),
 *                                 ascending feature index (push_back order)
 * Features whose word weight is 0 ("stopped", :1159,1186) appear in neither.  With out_loc == ORB_DEVICE the call
 * is asynchronous on the vocabulary's stream (orb_vocab_sync).                                                   */
typedef struct orb_bow_out {
    int32_t capacity;
    int32_t *bow_word;   /* [batch][capacity]   */
    double *bow_value;   /* [batch][capacity]   */
    int32_t *n_bow;      /* [batch]             */
    int32_t *fv_node;    /* [batch][capacity]   */
    int32_t *fv_off;     /* [batch][capacity+1] */
    int32_t *fv_feat;    /* [batch][capacity]   */
    int32_t *n_fv;       /* [batch]             */
} orb_bow_out;
int orb_vocab_transform(orb_vocab *v, const uint8_t *descriptors, const int32_t *n, int batch,
                        int stride_rows, int loc, int levelsup, const orb_bow_out *out, int out_loc);
int orb_vocab_sync(orb_vocab *v);
void *orb_vocab_stream(orb_vocab *v);
int64_t orb_vocab_launch_count(const orb_vocab *v);

const char *orb_last_error(void);
int orb_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* ORBGPU_H_ */
