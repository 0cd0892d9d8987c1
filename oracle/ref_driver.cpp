// ORACLE -- TEST INFRASTRUCTURE ONLY.  C entry points around the reference's own ORBextractor class, compiled from
// /root/reference/ORB_SLAM2/src/ORBextractor.cc against oracle/cvstub (see oracle/Makefile.ref).
//
// The reference breaks ties in DistributeOctTree's sort by node heap address (ORBextractor.cc:573,654-655).  This
// driver installs a monotonic (bump) operator new for the duration of an extraction, so addresses grow with
// allocation order and the reference itself realises the "creation order" tie rule the oracle fixes.
#include "ORBextractor.h"
#include "../include/orbgpu.h"

#include <cstdio>
#include <cstdlib>
#include <new>

namespace {
char *g_arena = nullptr;
size_t g_arena_size = 0, g_arena_used = 0;
bool g_bump = false;
}  // namespace

void *operator new(size_t n) {
    if (g_bump) {
        n = (n + 15) & ~(size_t)15;
        if (g_arena_used + n > g_arena_size) { fprintf(stderr, "ref_driver: arena exhausted\n"); abort(); }
        void *p = g_arena + g_arena_used;
        g_arena_used += n;
        return p;
    }
    void *p = malloc(n ? n : 1);
    if (!p) throw std::bad_alloc();
    return p;
}
void operator delete(void *p) noexcept {
    if (p >= (void *)g_arena && p < (void *)(g_arena + g_arena_size)) return;   // bump memory is never reused
    free(p);
}
void operator delete(void *p, size_t) noexcept { operator delete(p); }

extern "C" {

struct ref_extractor { ORB_SLAM2::ORBextractor *ex; };

int ref_extractor_create(int nfeatures, float scale, int nlevels, int ini, int min, ref_extractor **out) {
    ref_extractor *r = (ref_extractor *)malloc(sizeof(ref_extractor));
    r->ex = new ORB_SLAM2::ORBextractor(nfeatures, scale, nlevels, ini, min);
    *out = r;
    return 0;
}
void ref_extractor_destroy(ref_extractor *r) { delete r->ex; free(r); }

int ref_extract(ref_extractor *r, const uint8_t *image, int width, int height, int pitch, orb_keypoint *kps, uint8_t *desc,
                int capacity, int *n_out) {
    if (!g_arena) { g_arena_size = (size_t)3 << 30; g_arena = (char *)malloc(g_arena_size); }
    g_arena_used = 0;
    int n = 0;
    {
        g_bump = true;
        cv::Mat img(height, width, CV_8UC1);
        for (int y = 0; y < height; y++) memcpy(img.ptr(y), image + (size_t)y * pitch, width);
        std::vector<cv::KeyPoint> keys;
        cv::Mat descriptors;
        (*r->ex)(img, cv::Mat(), keys, descriptors);
        n = (int)keys.size();
        if (n <= capacity) {
            for (int i = 0; i < n; i++) {
                kps[i].x = keys[i].pt.x; kps[i].y = keys[i].pt.y; kps[i].size = keys[i].size; kps[i].angle = keys[i].angle;
                kps[i].response = keys[i].response; kps[i].octave = keys[i].octave; kps[i].class_id = keys[i].class_id;
                memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
            }
        }
        // every object allocated under the bump allocator dies here, before it is switched off
    }
    g_bump = false;
    *n_out = n;
    return n <= capacity ? 0 : -4;
}

// level l of mvImagePyramid (interior), after the last ref_extract
int ref_pyramid_level(ref_extractor *r, int level, uint8_t *dst, int dst_pitch, int *w, int *h) {
    cv::Mat &m = r->ex->mvImagePyramid[level];
    *w = m.cols; *h = m.rows;
    if (dst) for (int y = 0; y < m.rows; y++) memcpy(dst + (size_t)y * dst_pitch, m.ptr(y), m.cols);
    return 0;
}

}  // extern "C"
