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
