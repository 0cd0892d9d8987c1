/* ORACLE -- TEST INFRASTRUCTURE ONLY.  C ABI of liborbref.so, the CPU restatement of the
 * reference's front end.  Same POD types as include/orbgpu.h so one harness drives both. */
#ifndef ORBREF_H_
#define ORBREF_H_
#include "../include/orbgpu.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct orbref_extractor orbref_extractor;
int orbref_extractor_create(const orb_params *p, orbref_extractor **out);
void orbref_extractor_destroy(orbref_extractor *ex);
/* 0 (default): restated glibc sincosf; 1: call the host libm's cosf/sinf */
int orbref_extractor_set_trig_mode(orbref_extractor *ex, int mode);
int orbref_extractor_tables(const orbref_extractor *ex, float *scale, float *inv_scale, float *sigma2,
                            float *inv_sigma2, int32_t *per_level, int32_t *umax16);
int orbref_extract(orbref_extractor *ex, const uint8_t *image, int width, int height, int pitch,
                   orb_keypoint *kps, uint8_t *desc, int capacity, int *n_out);
int orbref_level_size(const orbref_extractor *ex, int level, int *w, int *h);
int orbref_pyramid_download(orbref_extractor *ex, int level, int with_border, uint8_t *dst, int dst_pitch);
int orbref_debug_candidates(orbref_extractor *ex, int level, orb_candidate *out, int capacity, int *n_out);
int orbref_debug_blur_level(orbref_extractor *ex, int level, uint8_t *dst, int dst_pitch);
/* Frame::ComputeStereoMatches; sad_out (optional) = best SAD per left keypoint or -1 */
int orbref_stereo_match(orbref_extractor *left, orbref_extractor *right, const orb_keypoint *keysL,
                        const uint8_t *descL, int N, const orb_keypoint *keysR, const uint8_t *descR, int Nr,
                        float mbf, float mb, float *uRight, float *depth, int32_t *sad_out, int *n_matched);
int orbref_descriptor_distance(const uint8_t *a, const uint8_t *b, int n, int32_t *dist);
int orbref_hamming_top2(const uint8_t *queries, int nq, const uint8_t *db, int64_t ndb, int64_t idx_base,
                        int32_t *best_dist, int64_t *best_idx, int32_t *second_dist, int64_t *second_idx, int fast,
                        int threads);
void orbref_three_maxima(const int *sizes, int L, int *ind1, int *ind2, int *ind3);
int orbref_rot_bin(float angle_a, float angle_b);
int orbref_rot_hist_filter(const float *angle_a, const float *angle_b, int n, uint8_t *keep);
typedef struct orbref_frame orbref_frame;
int orbref_frame_create(const orb_frame_desc *d, orbref_frame **out);
void orbref_frame_destroy(orbref_frame *f);
int orbref_frame_features_in_area(orbref_frame *f, float x, float y, float r, int min_level, int max_level, int32_t *indices,
                                  int capacity, int *n_out);
int orbref_search_for_initialization(orbref_frame *F1, orbref_frame *F2, float *vbPrevMatched, int32_t *vnMatches12,
                                     int windowSize, float mfNNratio, int mbCheckOrientation, int *nmatches_out);
int orbref_search_windowed(orbref_frame *F, const orb_window_queries *Q, const uint8_t *claimed_in, int select, int th_dist,
                           float mfNNratio, int mbCheckOrientation, int32_t *assigned, int *nmatches_out);
int orbref_search_by_projection_points(orbref_frame *F, const orb_track_points *P, const uint8_t *claimed, float th, float nnratio,
                                       int32_t *assigned, int *nmatches);
int orbref_search_by_projection_last(orbref_frame *Cur, const orb_last_points *P, const uint8_t *claimed, float th, int forward,
                                     int backward, float mbf, int check_orientation, int32_t *assigned, int *nmatches);
int orbref_search_by_bow(orbref_frame *KF, orbref_frame *F, int ngroups, const int32_t *kf_off, const int32_t *kf_idx,
                         const int32_t *f_off, const int32_t *f_idx, const uint8_t *kf_valid, const uint8_t *f_valid,
                         int strict_low, float mfNNratio, int mbCheckOrientation, int32_t *matches, int *nmatches_out);
int orbref_search_window_best(orbref_frame *F, const orb_window_queries *Q, const orb_reproj_gate *gate, int32_t *best_idx,
                              int32_t *best_dist);
int orbref_search_for_triangulation(orbref_frame *KF1, orbref_frame *KF2, int ngroups, const int32_t *off1, const int32_t *idx1,
                                    const int32_t *off2, const int32_t *idx2, const uint8_t *usable1, const uint8_t *usable2,
                                    const float *F12, float ex, float ey, const float *level_sigma2, int only_stereo,
                                    int check_orientation, int32_t *matches12, int *nmatches);
/* DBoW2 vocabulary: text loader, tree descent, BowVector / FeatureVector assembly (orbref_bow.cpp) */
typedef struct orbref_vocab orbref_vocab;
int orbref_vocab_create(const orb_vocab_desc *d, orbref_vocab **out);
int orbref_vocab_load_text(const char *path, orbref_vocab **out);
void orbref_vocab_destroy(orbref_vocab *v);
int orbref_vocab_info(const orbref_vocab *v, int32_t *info);
int orbref_vocab_nodes(const orbref_vocab *v, int32_t *parent, uint8_t *desc, double *weight, int32_t *word_id);
int orbref_vocab_transform_features(const orbref_vocab *v, const uint8_t *desc, int n, int levelsup, int32_t *word,
                                    int32_t *node, double *weight);
int orbref_vocab_transform(const orbref_vocab *v, const uint8_t *desc, int n, int levelsup, int capacity, int32_t *bow_word,
                           double *bow_value, int32_t *n_bow, int32_t *fv_node, int32_t *fv_off, int32_t *fv_feat, int32_t *n_fv);
/* MapPoint::ComputeDistinctiveDescriptors (orbref_mappoint.cpp) */
int orbref_distinctive_descriptors(const uint8_t *desc, const int32_t *off, int npoints, int32_t *best_row, int32_t *best_median);
int orbref_search_matches_ratio(const uint8_t *train, int n_train, const uint8_t *query, int n_query, double ratio, int max_dist,
                                int32_t *train_idx, int32_t *dist);
/* Frame::isInFrustum + MapPoint::PredictScale (orbref_frustum.cpp) */
int orbref_is_in_frustum(const orb_frustum_camera *cam, const orb_frustum_points *pts, float limit, uint8_t *in_view, float *proj_x,
                         float *proj_y, float *proj_xr, int32_t *level, float *view_cos);
/* projection pre-loops of the mapping / loop matchers (orbref_frustum.cpp) */
int orbref_project_points(const orb_projection *P, const orb_projection_points *pts, uint8_t *ok, float *pu, float *pv, float *pur,
                          float *pdist);
/* cv::remap INTER_LINEAR, 8-bit, CV_32F maps (orbref_remap.cpp) */
int orbref_remap_linear(const uint8_t *src, int sw, int sh, int spitch, const float *mapx, const float *mapy, int mpitch, uint8_t *dst,
                        int dw, int dh, int dpitch);
/* cv::cvtColor X2GRAY, 8-bit (orbref_color.cpp) */
int orbref_cvt_color_gray(const uint8_t *src, int width, int height, int src_pitch, int channels, int rgb_order, int variant,
                          uint8_t *dst, int dst_pitch);
/* cv::undistortPoints(src, dst, K, D, noArray(), K) (orbref_undistort.cpp) */
int orbref_undistort_points(const float *xy, int n, float fx, float fy, float cx, float cy, const float *dist, int ndist, float *out);
/* OpenCV's cv::ORB with Harris score (orbref_cvorb.cpp), pinned to cv2 4.13 */
int orbref_cvorb_detect_and_compute(const orb_cvorb_params *p, const uint8_t *image, int width, int height, int pitch,
                                    orb_keypoint *kps, uint8_t *desc, int capacity, int *n_out);
int orbref_cvorb_level(const orb_cvorb_params *p, const uint8_t *image, int width, int height, int pitch, int level, int blurred,
                       uint8_t *dst, int dst_pitch, int *w, int *h);
int orbref_retain_best(const float *responses, int n, int n_points, int depth_limit, int32_t *order_out, int *n_out);
int orbref_gaussian_blur7_sep32f(const uint8_t *src, int w, int h, int spitch, uint8_t *dst, int dpitch);
int orbref_resize_linear_exact(const uint8_t *src, int sw, int sh, int spitch, uint8_t *dst, int dw, int dh, int dpitch);
/* primitives (golden-vector tests against cv2) */
int orbref_resize_linear(const uint8_t *src, int sw, int sh, int spitch, uint8_t *dst, int dw, int dh, int dpitch);
int orbref_gaussian_blur7(const uint8_t *src, int w, int h, int spitch, uint8_t *dst, int dpitch, int taps);
int orbref_fast(const uint8_t *img, int w, int h, int pitch, int threshold, orb_candidate *out, int capacity, int *n_out);
int orbref_fast_atan2(const float *y, const float *x, int n, float *out);
int orbref_sincos(const float *angle_rad, int n, float *s, float *c, int mode);
int orbref_make_border(uint8_t *plane, int w, int h, int pitch, int border);
#ifdef __cplusplus
}
#endif
#endif
