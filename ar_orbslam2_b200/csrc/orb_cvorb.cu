// liborbgpu -- OpenCV's cv::ORB (Harris score) on the device, for the reference's marker pipeline.
//
// Replaces `ORB orb; orb(image, Mat(), keys, descriptors);` in Marker::setTargetImage and Marker::Match
// (ref: ORB_SLAM2/src/Marker.cc:83-84, 107-108).  The detector is OpenCV's (an un-vendored dependency of the reference), so the
// arithmetic follows OpenCV's published algorithm (features2d/src/orb.cpp, keypoint.cpp, imgproc/src/resize.cpp) and is
// pinned to cv2 4.13.0 by the golden vectors of tools/make_golden_cvorb.py: layers, key points (all fields, cv2's order)
// and descriptors are bit-identical.
//
// Stages (all device-resident; one handle = one image at a time, its own stream):
//   k_cv_resize      layer l from layer l-1, INTER_LINEAR_EXACT (8.8 fixed-point coefficients, 16.16 accumulation)
//   k_cv_blur_rows / k_cv_blur_cols   the fp32 separable Gaussian a submatrix gets (sequential-FMA rows, symmetric-FMA columns)
//   k_cv_fast_score  FAST-9/16 arc strength of every pixel that can matter (layer interior minus edgeThreshold - 1)
//   k_cv_nms_rows    3x3 strict non-maximum suppression, one warp per layer row: a bit mask + a count per row
//   k_cv_compact     row-major key-point lists (cv::FAST's order) from the masks
//   k_cv_retain      KeyPointsFilter::retainBest: std::nth_element + std::partition REPLAYED in parallel -- Hoare's
//                    partition swaps its k-th left stopper with its k-th right stopper whatever happened before, so the
//                    stoppers are ranked by ballot/prefix sums and all swaps of a pass are done at once; the result is the
//                    permutation libstdc++ leaves (cv::ORB's key-point order depends on it)
//   k_cv_harris      Harris response (7x7 block, Sobel-like 3x3 gradients) of the survivors, one warp each
//   k_cv_retain      again, on the Harris responses, down to the layer's quota
//   k_cv_describe    intensity-centroid angle (fastAtan2), 256 rotated tests on the blurred layer, one warp per key point
#include "orb_common.cuh"
#include "orb_devmath.cuh"
#include "orb_pattern.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace orbgpu {

constexpr int CV_MAXL = ORB_MAX_LEVELS;
constexpr int CV_RETAIN_T = 1024;

struct CvTap { int ofs, c1; };   // source index of the first sample, weight of the second (8.8); first weight = 256 - c1
struct CvElem { float resp; uint32_t xy; };   // key-point candidate: response, x | y << 16 (layer coordinates)

struct CvLevel {
    int w, h, pitch;
    long long off;        // byte offset of the layer in the u8 plane pools (image, blurred, score)
    long long foff;       // float offset of the layer in the row-filter scratch
    float scale;
    int quota;
    int tapx, tapy;       // offsets into the tap pool (level >= 1)
    int row0, nrows;      // flattened NMS rows [row0, row0 + nrows): layer rows edge .. h - edge - 1
    int mask_words;       // mask words per NMS row
    long long mask_off;   // first mask word of the layer
    int cand_off, cand_cap;
};
struct CvGeom {
    int nlevels, width, height, edge, fast_th, harris, nfeatures;
    int total_rows, out_cap;
    CvLevel lv[CV_MAXL];
};

__constant__ int8_t c_cv_pattern[ORB_PATTERN_INTS] = {ORB_PATTERN_VALUES};
__constant__ int c_cv_umax[16] = {15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3};
__constant__ float c_cv_gauss[7] = {0x1.1f5f62p-4f, 0x1.0c70fcp-3f, 0x1.869472p-3f, 0x1.ba95c0p-3f,
                                    0x1.869472p-3f, 0x1.0c70fcp-3f, 0x1.1f5f62p-4f};   // getGaussianKernel(7, 2, CV_32F)

// ---- pyramid: resize(INTER_LINEAR_EXACT) ------------------------------------------------------------------------------
// resize.cpp resize_bitExact<uint8_t, ufixedpoint16>: horizontal pass c0*p[ofs] + c1*p[ofs+1] in 8.8, vertical pass the same on
// the two row results in 16.16, + 0.5, >> 16.  Edge outputs (one source sample) carry c1 = 0, which gives the same value.
__global__ void __launch_bounds__(256)
k_cv_resize(const uint8_t *__restrict__ src, int sw, int sh, int spitch, uint8_t *__restrict__ dst, int dw, int dh, int dpitch,
            const CvTap *__restrict__ tx, const CvTap *__restrict__ ty) {
    const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y;
    if (x4 >= dw) return;
    const CvTap t = ty[y];
    const uint8_t *r0 = src + (size_t)t.ofs * spitch, *r1 = src + (size_t)min(t.ofs + 1, sh - 1) * spitch;
    const uint32_t m1 = (uint32_t)t.c1, m0 = 256u - m1;
    uint32_t out = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int x = x4 + i;
        if (x < dw) {
            const CvTap a = tx[x];
            const int x1 = min(a.ofs + 1, sw - 1);
            const uint32_t c1 = (uint32_t)a.c1, c0 = 256u - c1;
            const uint32_t h0 = c0 * r0[a.ofs] + c1 * r0[x1], h1 = c0 * r1[a.ofs] + c1 * r1[x1];
            const uint32_t v = (h0 * m0 + h1 * m1 + 32768u) >> 16;
            out |= min(v, 255u) << (8 * i);
        }
    }
    *(uint32_t *)(dst + (size_t)y * dpitch + x4) = out;   // pitch is a multiple of 4: the tail bytes of a row are padding
}

// ---- blur --------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_cv_blur_rows(const uint8_t *__restrict__ img, float *__restrict__ rows, const CvGeom *__restrict__ Gp) {
    __shared__ CvGeom G;
    for (int i = threadIdx.x; i < (int)(sizeof(CvGeom) / 4); i += blockDim.x) ((int *)&G)[i] = ((const int *)Gp)[i];
    __syncthreads();
    int l, base = 0;
    int row = blockIdx.y;
    for (l = 0; l < G.nlevels; l++) { if (row < base + G.lv[l].h) break; base += G.lv[l].h; }
    if (l >= G.nlevels) return;
    const CvLevel &L = G.lv[l];
    const int y = row - base, x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= L.w) return;
    const uint8_t *s = img + L.off + (size_t)y * L.pitch;
    // RowVec_32f: s = k[0]*p[x-3]; s = fma(p[x-3+j], k[j], s)
    float acc = __fmul_rn(c_cv_gauss[0], (float)s[reflect101(x - 3, L.w)]);
#pragma unroll
    for (int j = 1; j < 7; j++) acc = __fmaf_rn((float)s[reflect101(x - 3 + j, L.w)], c_cv_gauss[j], acc);
    rows[L.foff + (size_t)y * L.w + x] = acc;
}

__global__ void __launch_bounds__(256)
k_cv_blur_cols(const float *__restrict__ rows, uint8_t *__restrict__ blur, const CvGeom *__restrict__ Gp) {
    __shared__ CvGeom G;
    for (int i = threadIdx.x; i < (int)(sizeof(CvGeom) / 4); i += blockDim.x) ((int *)&G)[i] = ((const int *)Gp)[i];
    __syncthreads();
    int l, base = 0;
    int row = blockIdx.y;
    for (l = 0; l < G.nlevels; l++) { if (row < base + G.lv[l].h) break; base += G.lv[l].h; }
    if (l >= G.nlevels) return;
    const CvLevel &L = G.lv[l];
    const int y = row - base, x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= L.w) return;
    const float *r = rows + L.foff + x;
    // SymmColumnVec_32f: s = k[3]*r[y]; s = fma(r[y+j] + r[y-j], k[3+j], s)
    float acc = __fmul_rn(c_cv_gauss[3], r[(size_t)y * L.w]);
#pragma unroll
    for (int j = 1; j <= 3; j++) {
        const float pair = __fadd_rn(r[(size_t)reflect101(y + j, L.h) * L.w], r[(size_t)reflect101(y - j, L.h) * L.w]);
        acc = __fmaf_rn(pair, c_cv_gauss[3 + j], acc);
    }
    const int v = __float2int_rn(acc);
    blur[L.off + (size_t)y * L.pitch + x] = (uint8_t)min(max(v, 0), 255);
}

// ---- FAST-9/16 score ---------------------------------------------------------------------------------------------------
// cv::FAST(layer, threshold, nonmaxSuppression = true): a pixel's score is cornerScore = V - 1, V = the largest t for which it
// fails to be... precisely: V = max over the 16 arcs of 9 contiguous circle pixels of min(c - q) and of min(q - c); corner iff
// V > threshold.  Only pixels whose score can reach the output are computed: rows/columns edge-1 .. size-edge (the kept
// region plus the ring its suppression reads).
__device__ __forceinline__ int cv_arc_strength(const uint8_t *p, int pitch) {
    int d[16];
    const int c = p[0];
    d[0] = c - p[3 * pitch];       d[1] = c - p[3 * pitch + 1];   d[2] = c - p[2 * pitch + 2];   d[3] = c - p[pitch + 3];
    d[4] = c - p[3];               d[5] = c - p[-pitch + 3];      d[6] = c - p[-2 * pitch + 2];  d[7] = c - p[-3 * pitch + 1];
    d[8] = c - p[-3 * pitch];      d[9] = c - p[-3 * pitch - 1];  d[10] = c - p[-2 * pitch - 2]; d[11] = c - p[-pitch - 3];
    d[12] = c - p[-3];             d[13] = c - p[pitch - 3];      d[14] = c - p[2 * pitch - 2];  d[15] = c - p[3 * pitch - 1];
    // min over every window of 9 consecutive ring pixels by doubling (2, 4, 8, then the ninth), for c - q and for q - c
    int best = -256;
#pragma unroll
    for (int pol = 0; pol < 2; pol++) {
        int m2[16], m4[16], m8[16];
#pragma unroll
        for (int k = 0; k < 16; k++) m2[k] = min(d[k], d[(k + 1) & 15]);
#pragma unroll
        for (int k = 0; k < 16; k++) m4[k] = min(m2[k], m2[(k + 2) & 15]);
#pragma unroll
        for (int k = 0; k < 16; k++) m8[k] = min(m4[k], m4[(k + 4) & 15]);
#pragma unroll
        for (int k = 0; k < 16; k++) best = max(best, min(m8[k], d[(k + 8) & 15]));
#pragma unroll
        for (int k = 0; k < 16; k++) d[k] = -d[k];
    }
    return best;
}

__global__ void __launch_bounds__(256)
k_cv_fast_score(const uint8_t *__restrict__ img, uint8_t *__restrict__ score, const CvGeom *__restrict__ Gp) {
    __shared__ CvGeom G;
    for (int i = threadIdx.x; i < (int)(sizeof(CvGeom) / 4); i += blockDim.x) ((int *)&G)[i] = ((const int *)Gp)[i];
    __syncthreads();
    // flattened score rows: nrows + 2 per layer that has NMS rows
    int l, base = 0;
    const int row = blockIdx.y;
    for (l = 0; l < G.nlevels; l++) {
        const int n = G.lv[l].nrows > 0 ? G.lv[l].nrows + 2 : 0;
        if (row < base + n) break;
        base += n;
    }
    if (l >= G.nlevels) return;
    const CvLevel &L = G.lv[l];
    const int y = G.edge - 1 + (row - base), x = G.edge - 1 + blockIdx.x * blockDim.x + threadIdx.x;
    if (x > L.w - G.edge) return;
    const uint8_t *p = img + L.off + (size_t)y * L.pitch + x;
    const int t = G.fast_th, c = p[0];
    const int d0 = c - p[3 * L.pitch], d4 = c - p[3], d8 = c - p[-3 * L.pitch], d12 = c - p[-3];
    const int hi = (d0 > t) + (d4 > t) + (d8 > t) + (d12 > t), lo = (d0 < -t) + (d4 < -t) + (d8 < -t) + (d12 < -t);
    int s = 0;
    if (hi >= 2 || lo >= 2) {   // every 9-arc holds two adjacent compass points: exact early-out
        const int V = cv_arc_strength(p, L.pitch);
        s = V > t ? V - 1 : 0;
    }
    score[L.off + (size_t)y * L.pitch + x] = (uint8_t)s;
}

// one warp per NMS row: mask bit = the pixel's score beats its 8 neighbours strictly (cv::FAST's suppression)
__global__ void __launch_bounds__(256)
k_cv_nms_rows(const uint8_t *__restrict__ score, uint32_t *__restrict__ masks, int *__restrict__ rowcount, const CvGeom *__restrict__ Gp) {
    __shared__ CvGeom G;
    for (int i = threadIdx.x; i < (int)(sizeof(CvGeom) / 4); i += blockDim.x) ((int *)&G)[i] = ((const int *)Gp)[i];
    __syncthreads();
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= G.total_rows) return;
    int l, base = 0;
    for (l = 0; l < G.nlevels; l++) { if (row < base + G.lv[l].nrows) break; base += G.lv[l].nrows; }
    const CvLevel &L = G.lv[l];
    const int y = G.edge + (row - base), x_end = L.w - G.edge;
    const uint8_t *r = score + L.off + (size_t)y * L.pitch;
    uint32_t *m = masks + L.mask_off + (size_t)(row - base) * L.mask_words;
    int cnt = 0;
    for (int wdx = 0; wdx < L.mask_words; wdx++) {
        const int x = G.edge + wdx * 32 + lane;
        bool keep = false;
        if (x < x_end) {
            const int s = r[x];
            if (s > 0) {
                const uint8_t *u = r - L.pitch, *d = r + L.pitch;
                keep = s > r[x - 1] && s > r[x + 1] && s > u[x - 1] && s > u[x] && s > u[x + 1] && s > d[x - 1] && s > d[x] && s > d[x + 1];
            }
        }
        const uint32_t b = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) m[wdx] = b;
        cnt += __popc(b);
    }
    if (lane == 0) rowcount[row] = cnt;
}

// row-major lists: a warp per NMS row adds up the counts of the rows above it (same layer) and expands its mask
__global__ void __launch_bounds__(256)
k_cv_compact(const uint8_t *__restrict__ score, const uint32_t *__restrict__ masks, const int *__restrict__ rowcount,
             CvElem *__restrict__ cand, int *__restrict__ cand_n, int *__restrict__ err, const CvGeom *__restrict__ Gp) {
    __shared__ CvGeom G;
    for (int i = threadIdx.x; i < (int)(sizeof(CvGeom) / 4); i += blockDim.x) ((int *)&G)[i] = ((const int *)Gp)[i];
    __syncthreads();
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= G.total_rows) return;
    int l, base = 0;
    for (l = 0; l < G.nlevels; l++) { if (row < base + G.lv[l].nrows) break; base += G.lv[l].nrows; }
    const CvLevel &L = G.lv[l];
    const int ry = row - base;
    int ofs = 0;
    for (int i = lane; i < ry; i += 32) ofs += rowcount[base + i];
#pragma unroll
    for (int s = 16; s; s >>= 1) ofs += __shfl_xor_sync(0xffffffffu, ofs, s);
    const int mine = rowcount[row];
    if (ry == L.nrows - 1 && lane == 0) {
        int n = ofs + mine;
        if (n > L.cand_cap) { atomicOr(err, 1); n = L.cand_cap; }
        cand_n[l] = n;
    }
    if (!mine) return;
    const int y = G.edge + ry;
    const uint8_t *r = score + L.off + (size_t)y * L.pitch;
    const uint32_t *m = masks + L.mask_off + (size_t)ry * L.mask_words;
    for (int wdx = 0; wdx < L.mask_words; wdx++) {
        const uint32_t b = m[wdx];
        if (b >> lane & 1) {
            const int pos = ofs + __popc(b & ((1u << lane) - 1)), x = G.edge + wdx * 32 + lane;
            if (pos < L.cand_cap) cand[L.cand_off + pos] = CvElem{(float)r[x], (uint32_t)x | (uint32_t)y << 16};
        }
        ofs += __popc(b);
    }
}

// ---- KeyPointsFilter::retainBest, replayed ------------------------------------------------------------------------------
struct RetainShared {
    int wl[32], wr[32];   // per-warp stopper counts, then exclusive prefixes
    int nl, nr, k, cut;
};

// Hoare-style pass over E[lo, hi): left stoppers (scanning up) = elements with LEFT(e), right stoppers (scanning down) =
// elements with RIGHT(e); the k-th left stopper is swapped with the k-th right stopper while it lies left of it.  Ranks
// come from ballots inside a warp's contiguous segment and a scan over the warps; Lp / Rp receive the stopper positions.
// Returns K (number of swaps); S.nl / S.nr hold the stopper counts.  sentinel >= 0: a last-resort right stopper at that
// position (libstdc++'s unguarded partition stops its downward scan on the pivot, which sits just below the range).
template <class LeftF, class RightF>
__device__ int hoare_pass(CvElem *E, int lo, int hi, int *Lp, int *Rp, RetainShared &S, int sentinel, LeftF is_left, RightF is_right) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    const int m = hi - lo;
    const int seg = ((m + nw * 32 - 1) / (nw * 32)) * 32;   // elements per warp, multiple of 32
    const int b = lo + warp * seg, e = min(b + seg, hi);
    int cl = 0, cr = 0;
    for (int p = b + lane; p - lane < e; p += 32) {
        const bool in = p < e;
        CvElem v = in ? E[p] : CvElem{0.f, 0u};
        cl += __popc(__ballot_sync(0xffffffffu, in && is_left(v)));
        cr += __popc(__ballot_sync(0xffffffffu, in && is_right(v)));
    }
    if (lane == 0) { S.wl[warp] = cl; S.wr[warp] = cr; }
    __syncthreads();
    if (warp == 0) {
        int a = lane < nw ? S.wl[lane] : 0, c = lane < nw ? S.wr[lane] : 0;
        int ia = a, ic = c;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const int ta = __shfl_up_sync(0xffffffffu, ia, s), tc = __shfl_up_sync(0xffffffffu, ic, s);
            if (lane >= s) { ia += ta; ic += tc; }
        }
        if (lane < nw) { S.wl[lane] = ia - a; S.wr[lane] = ic - c; }
        if (lane == 31) { S.nl = ia; S.nr = ic; }
    }
    __syncthreads();
    const int NL = S.nl, NR = S.nr;
    int bl = S.wl[warp], br = S.wr[warp];
    for (int p = b + lane; p - lane < e; p += 32) {
        const bool in = p < e;
        CvElem v = in ? E[p] : CvElem{0.f, 0u};
        const bool fl = in && is_left(v), fr = in && is_right(v);
        const uint32_t ml = __ballot_sync(0xffffffffu, fl), mr = __ballot_sync(0xffffffffu, fr);
        const uint32_t lt = (1u << lane) - 1;
        if (fl) Lp[bl + __popc(ml & lt)] = p;
        if (fr) Rp[NR - 1 - (br + __popc(mr & lt))] = p;
        bl += __popc(ml);
        br += __popc(mr);
    }
    if (tid == 0) { S.k = 0; if (sentinel >= 0) Rp[NR] = sentinel; }
    __syncthreads();
    // K = number of k with Lp[k] < Rp[k] (a prefix: Lp ascends, Rp descends)
    const int pairs = min(NL, NR);
    int mine = 0;
    for (int k = tid; k < pairs; k += blockDim.x) mine += Lp[k] < Rp[k];
#pragma unroll
    for (int s = 16; s; s >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, s);
    if (lane == 0 && mine) atomicAdd(&S.k, mine);
    __syncthreads();
    const int K = S.k;
    for (int k = tid; k < K; k += blockDim.x) {
        const int a = Lp[k], c = Rp[k];
        const CvElem t = E[a];
        E[a] = E[c];
        E[c] = t;
    }
    __syncthreads();
    return K;
}

__device__ __forceinline__ bool cv_greater(const CvElem &a, const CvElem &b) { return a.resp > b.resp; }   // KeypointResponseGreater
__device__ __forceinline__ void cv_swap(CvElem &a, CvElem &b) { const CvElem t = a; a = b; b = t; }

// libstdc++ heap primitives (bits/stl_heap.h) for __introselect's depth-limit fallback; one thread.
__device__ void cv_push_heap(CvElem *first, int hole, int top, CvElem value) {
    int parent = (hole - 1) / 2;
    while (hole > top && cv_greater(first[parent], value)) {
        first[hole] = first[parent];
        hole = parent;
        parent = (hole - 1) / 2;
    }
    first[hole] = value;
}
__device__ void cv_adjust_heap(CvElem *first, int hole, int len, CvElem value) {
    const int top = hole;
    int child = hole;
    while (child < (len - 1) / 2) {
        child = 2 * (child + 1);
        if (cv_greater(first[child], first[child - 1])) child--;
        first[hole] = first[child];
        hole = child;
    }
    if ((len & 1) == 0 && child == (len - 2) / 2) {
        child = 2 * (child + 1);
        first[hole] = first[child - 1];
        hole = child - 1;
    }
    cv_push_heap(first, hole, top, value);
}
__device__ void cv_heap_select(CvElem *E, int first, int middle, int last) {
    CvElem *f = E + first;
    const int len = middle - first;
    if (len >= 2) {   // __make_heap
        int parent = (len - 2) / 2;
        while (true) {
            const CvElem v = f[parent];
            cv_adjust_heap(f, parent, len, v);
            if (parent == 0) break;
            parent--;
        }
    }
    for (int i = middle; i < last; ++i)
        if (cv_greater(E[i], f[0])) {   // __pop_heap(first, middle, i)
            const CvElem v = E[i];
            E[i] = f[0];
            cv_adjust_heap(f, 0, len, v);
        }
}

// std::nth_element(E, E + nth, E + n, KeypointResponseGreater()) then std::partition(E + nth + 1, E + n, resp >= E[nth].resp):
// libstdc++'s __introselect (median of first+1 / middle / last-1 to the front, unguarded Hoare partition around it, recurse
// into the side that holds nth, insertion sort of the last <= 3) with the partition passes done by the whole CTA.
// depth_limit < 0: the library's 2*floor(log2(n)).  Returns the new length.
__device__ int cv_retain_best(CvElem *E, int n, int n_points, int *Lp, int *Rp, RetainShared &S, int depth_limit) {
    if (n_points < 0 || n <= n_points) return n;
    if (n_points == 0) return 0;
    const int tid = threadIdx.x;
    const int nth = n_points - 1;
    int first = 0, last = n;
    int depth = depth_limit >= 0 ? depth_limit : 2 * (31 - __clz(n));
    bool done = false;
    while (last - first > 3) {
        if (depth == 0) {
            if (tid == 0) {
                cv_heap_select(E, first, nth + 1, last);
                cv_swap(E[first], E[nth]);
            }
            __syncthreads();
            done = true;
            break;
        }
        --depth;
        if (tid == 0) {   // __move_median_to_first(first, first + 1, mid, last - 1)
            const int a = first + 1, b = first + (last - first) / 2, c = last - 1;
            const CvElem ea = E[a], eb = E[b], ec = E[c];
            int pick;
            if (cv_greater(ea, eb)) pick = cv_greater(eb, ec) ? b : cv_greater(ea, ec) ? c : a;
            else pick = cv_greater(ea, ec) ? a : cv_greater(eb, ec) ? c : b;
            cv_swap(E[first], E[pick]);
        }
        __syncthreads();
        const float pivot = E[first].resp;
        // __unguarded_partition(first + 1, last, first): up-scan stops on !(e > pivot), down-scan on !(pivot > e)
        const int K = hoare_pass(E, first + 1, last, Lp, Rp, S, first,
                                 [pivot](const CvElem &v) { return !(v.resp > pivot); },
                                 [pivot](const CvElem &v) { return !(pivot > v.resp); });
        // the up-scan's K-th stop: the next original stopper, or the element the last swap put on the right
        int cut = last;
        if (K < S.nl) cut = Lp[K];
        if (K >= 1) cut = min(cut, Rp[K - 1]);
        __syncthreads();   // Lp / Rp are rewritten by the next pass
        if (cut <= nth) first = cut;
        else last = cut;
    }
    if (!done) {
        if (tid == 0) {   // __insertion_sort(first, last)
            for (int i = first + 1; i < last; ++i) {
                const CvElem val = E[i];
                if (cv_greater(val, E[first])) {
                    for (int j = i; j > first; --j) E[j] = E[j - 1];
                    E[first] = val;
                } else {
                    int j = i;
                    while (cv_greater(val, E[j - 1])) { E[j] = E[j - 1]; --j; }
                    E[j] = val;
                }
            }
        }
        __syncthreads();
    }
    const float amb = E[n_points - 1].resp;
    // std::partition (bidirectional): up-scan stops on !pred, down-scan on pred, swap while they have not met
    hoare_pass(E, n_points, n, Lp, Rp, S, -1,
               [amb](const CvElem &v) { return !(v.resp >= amb); },
               [amb](const CvElem &v) { return v.resp >= amb; });
    return n_points + S.nr;
}

// stage 0: the FAST-score cut (2 x quota with Harris, quota without); stage 1: the Harris cut (quota)
__global__ void __launch_bounds__(CV_RETAIN_T)
k_cv_retain(CvElem *__restrict__ cand, int *__restrict__ cand_n, int *__restrict__ scratch, const CvGeom *__restrict__ Gp, int stage) {
    __shared__ RetainShared S;
    const int l = blockIdx.x;
    const CvLevel &L = Gp->lv[l];
    if (L.nrows <= 0) { if (threadIdx.x == 0) cand_n[l] = 0; return; }
    const int n = cand_n[l];
    const int npts = stage == 0 ? (Gp->harris ? 2 * L.quota : L.quota) : L.quota;
    __syncthreads();
    const int out = cv_retain_best(cand + L.cand_off, n, npts, scratch + 2 * (size_t)L.cand_off + 2 * l,
                                   scratch + 2 * (size_t)L.cand_off + 2 * l + L.cand_cap + 1, S, -1);
    if (threadIdx.x == 0) cand_n[l] = out;
}

// tests: retainBest on caller-supplied responses (one CTA), with libstdc++'s depth limit or a forced one
__global__ void __launch_bounds__(CV_RETAIN_T)
k_cv_retain_debug(CvElem *E, int n, int n_points, int *scratch, int depth_limit, int *n_out) {
    __shared__ RetainShared S;
    const int out = cv_retain_best(E, n, n_points, scratch, scratch + n + 1, S, depth_limit);
    if (threadIdx.x == 0) *n_out = out;
}

// ---- Harris responses (orb.cpp HarrisResponses, blockSize 7, k = 0.04) ---------------------------------------------------
__global__ void __launch_bounds__(256)
k_cv_harris(const uint8_t *__restrict__ img, CvElem *__restrict__ cand, const int *__restrict__ cand_n, const CvGeom *__restrict__ Gp) {
    const int l = blockIdx.y, lane = threadIdx.x & 31;
    const CvLevel &L = Gp->lv[l];
    const int n = cand_n[l];
    for (int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n; i += gridDim.x * (blockDim.x >> 5)) {
    CvElem &e = cand[L.cand_off + i];
    const int x0 = e.xy & 0xffff, y0 = e.xy >> 16, step = L.pitch;
    int a = 0, b = 0, c = 0;
    for (int k = lane; k < 49; k += 32) {
        const uint8_t *p = img + L.off + (size_t)(y0 - 3 + k / 7) * step + (x0 - 3 + k % 7);
        const int Ix = (p[1] - p[-1]) * 2 + (p[-step + 1] - p[-step - 1]) + (p[step + 1] - p[step - 1]);
        const int Iy = (p[step] - p[-step]) * 2 + (p[step - 1] - p[-step - 1]) + (p[step + 1] - p[-step + 1]);
        a += Ix * Ix;
        b += Iy * Iy;
        c += Ix * Iy;
    }
#pragma unroll
    for (int s = 16; s; s >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, s);
        b += __shfl_xor_sync(0xffffffffu, b, s);
        c += __shfl_xor_sync(0xffffffffu, c, s);
    }
    if (lane == 0) {
        const float scale = __fdiv_rn(1.f, __fmul_rn(28.f, 255.f));   // 1.f / ((1 << 2) * blockSize * 255.f)
        const float ssss = __fmul_rn(__fmul_rn(__fmul_rn(scale, scale), scale), scale);
        const float fa = (float)a, fb = (float)b, fc = (float)c;
        const float sum = __fadd_rn(fa, fb);
        const float r = __fsub_rn(__fsub_rn(__fmul_rn(fa, fb), __fmul_rn(fc, fc)), __fmul_rn(__fmul_rn(0.04f, sum), sum));
        e.resp = __fmul_rn(r, ssss);
    }
    }
}

// ---- angles + descriptors ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_cv_describe(const uint8_t *__restrict__ img, const uint8_t *__restrict__ blur, const CvElem *__restrict__ cand,
              const int *__restrict__ cand_n, const CvGeom *__restrict__ Gp, orb_keypoint *__restrict__ kps,
              uint8_t *__restrict__ desc, int *__restrict__ n_out, int *__restrict__ err) {
    __shared__ int8_t pat[ORB_PATTERN_INTS];
    for (int i = threadIdx.x; i < ORB_PATTERN_INTS; i += blockDim.x) pat[i] = c_cv_pattern[i];
    __syncthreads();
    const int l = blockIdx.y, lane = threadIdx.x & 31;
    const CvLevel &L = Gp->lv[l];
    int before = 0, total = 0;
    for (int j = 0; j < Gp->nlevels; j++) {
        const int c = cand_n[j];
        if (j < l) before += c;
        total += c;
    }
    if (blockIdx.x == 0 && l == 0 && threadIdx.x == 0) {
        *n_out = total;
        if (total > Gp->out_cap) atomicOr(err, 2);
    }
    const int n = cand_n[l];
    for (int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n; i += gridDim.x * (blockDim.x >> 5)) {
    const int o = before + i;
    if (o >= Gp->out_cap) return;
    const CvElem e = cand[L.cand_off + i];
    const int x0 = e.xy & 0xffff, y0 = e.xy >> 16, step = L.pitch;
    // ICAngles: moments over the radius-15 disc of the UNBLURRED layer; lane = row v + 15
    int m10 = 0, m01 = 0;
    if (lane < 31) {
        const int v = lane - 15, d = c_cv_umax[abs(v)];
        const uint8_t *row = img + L.off + (size_t)(y0 + v) * step + x0;
        int s = 0;
        for (int u = -d; u <= d; ++u) {
            const int val = row[u];
            s += val;
            m10 += u * val;
        }
        m01 = v * s;
    }
#pragma unroll
    for (int s = 16; s; s >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, s);
        m01 += __shfl_xor_sync(0xffffffffu, m01, s);
    }
    const float angle = dev_fast_atan2((float)m01, (float)m10);
    // key point: pt *= layerScale; size = patchSize * layerScale
    const float px = __fmul_rn((float)x0, L.scale), py = __fmul_rn((float)y0, L.scale);
    if (lane == 0) {
        orb_keypoint k;
        k.x = px; k.y = py; k.size = __fmul_rn(31.f, L.scale); k.angle = angle; k.response = e.resp; k.octave = l; k.class_id = -1;
        kps[o] = k;
    }
    // computeOrbDescriptors: centre = cvRound(pt * (1.f / layerScale)); a = cosf, b = sinf of the angle in radians
    const float inv = __fdiv_rn(1.f, L.scale);
    const int cx = __float2int_rn(__fmul_rn(px, inv)), cy = __float2int_rn(__fmul_rn(py, inv));
    float a, b;
    dev_sincosf(__fmul_rn(angle, (float)(3.1415926535897932384626433832795 / 180.f)), &b, &a);
    const uint8_t *center = blur + L.off + (size_t)cy * step + cx;
    int val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int8_t *pp = pat + (lane * 16 + k * 2) * 2;
        int t[2];
#pragma unroll
        for (int s = 0; s < 2; s++) {
            const float qx = (float)pp[s * 2], qy = (float)pp[s * 2 + 1];
            const float x = __fsub_rn(__fmul_rn(qx, a), __fmul_rn(qy, b)), y = __fadd_rn(__fmul_rn(qx, b), __fmul_rn(qy, a));
            t[s] = center[__float2int_rn(y) * step + __float2int_rn(x)];
        }
        val |= (t[0] < t[1]) << k;
    }
    desc[(size_t)o * 32 + lane] = (uint8_t)val;
    }
}

}  // namespace orbgpu

using namespace orbgpu;

struct orb_cvorb {
    orb_cvorb_params p;
    int device = 0;
    cudaStream_t stream = nullptr, s_blur = nullptr;
    cudaEvent_t ev_pyr = nullptr, ev_blur = nullptr;
    CvGeom G;                       // geometry of the current image size
    int cur_w = 0, cur_h = 0;
    // device pools, sized for max_width x max_height
    uint8_t *d_img = nullptr, *d_blur = nullptr, *d_score = nullptr;
    float *d_rows = nullptr;
    CvTap *d_taps = nullptr;
    uint32_t *d_masks = nullptr;
    int *d_rowcount = nullptr, *d_cand_n = nullptr, *d_scratch = nullptr, *d_hdr = nullptr;   // hdr: [0] n_out, [1] error flags
    CvElem *d_cand = nullptr;
    CvGeom *d_geom = nullptr;
    uint8_t *h_pin = nullptr;       // pinned block the describe kernel writes straight into: header, key points [out_cap], descriptors
    uint8_t *h_in = nullptr;        // pinned staging of a host image (one memcpy, then a true DMA instead of the driver's pageable path)
    size_t plane_bytes = 0, rows_floats = 0, mask_words = 0, tap_count = 0, cand_total = 0, rows_total = 0;
    int out_cap = 0;
    int64_t launches = 0;
};

namespace {

int cv_round_host(double v) { return (int)std::nearbyint(v); }

// interpolation_linear<ufixedpoint16>::getCoeffs (resize.cpp); volatile keeps every step a rounded double (softdouble)
void linear_exact_taps(int srcsize, int dstsize, CvTap *t) {
    const double inv_scale = (double)dstsize / srcsize;
    const double scale = 1.0 / inv_scale;
    for (int v = 0; v < dstsize; v++) {
        volatile double m = scale * ((double)v + 0.5);
        const double fval = m - 0.5;
        const int ival = (int)std::floor(fval);
        if (ival >= 0 && srcsize > 1) {
            if (ival < srcsize - 1) {
                volatile double frac = fval - (double)ival;
                t[v].ofs = ival;
                t[v].c1 = cv_round_host(frac * 256.0);
            } else {
                t[v].ofs = srcsize - 1;
                t[v].c1 = 0;
            }
        } else {
            t[v].ofs = 0;
            t[v].c1 = 0;
        }
    }
}

// layer sizes, quotas and pool offsets for a width x height image (orb.cpp: detectAndCompute's layerInfo, computeKeyPoints'
// nfeaturesPerLevel).  Returns false when a layer would be empty.
bool cv_geometry(const orb_cvorb_params &p, int width, int height, CvGeom &G, std::vector<CvTap> *taps) {
    memset(&G, 0, sizeof(G));
    G.nlevels = p.nlevels; G.width = width; G.height = height; G.edge = p.edge_threshold; G.fast_th = p.fast_threshold;
    G.harris = p.score_type == 0; G.nfeatures = p.nfeatures;
    const double sf = (double)p.scale_factor;
    long long off = 0, foff = 0, moff = 0;
    int row0 = 0, cand_off = 0, tap_off = 0;
    for (int l = 0; l < p.nlevels; l++) {
        CvLevel &L = G.lv[l];
        L.scale = (float)std::pow(sf, (double)l);
        const float inv = 1.0f / L.scale;
        L.w = cv_round_host((double)((float)width * inv));
        L.h = cv_round_host((double)((float)height * inv));
        if (L.w < 4 || L.h < 4) return false;   // reflect101 of the 7-tap blur needs 4 pixels
        L.pitch = (L.w + 127) & ~127;
        L.off = off; off += (long long)L.pitch * L.h;
        L.foff = foff; foff += (long long)L.w * L.h;
        const bool room = !(L.h <= 2 * G.edge || L.w <= 2 * G.edge);
        L.nrows = room ? L.h - 2 * G.edge : 0;
        L.row0 = row0; row0 += L.nrows;
        L.mask_words = room ? (L.w - 2 * G.edge + 31) / 32 : 0;
        L.mask_off = moff; moff += (long long)L.mask_words * L.nrows;
        // cv::FAST's 3x3 suppression keeps at most one pixel per 2x2 block
        L.cand_cap = room ? ((L.w - 2 * G.edge + 1) / 2) * ((L.h - 2 * G.edge + 1) / 2) + 1 : 1;
        L.cand_off = cand_off; cand_off += L.cand_cap;
        L.tapx = tap_off; tap_off += L.w;
        L.tapy = tap_off; tap_off += L.h;
    }
    G.total_rows = row0;
    {   // computeKeyPoints
        const float factor = (float)(1.0 / sf);
        float nd = (float)p.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)p.nlevels));
        int sum = 0;
        for (int l = 0; l < p.nlevels - 1; l++) {
            G.lv[l].quota = cv_round_host((double)nd);
            sum += G.lv[l].quota;
            nd *= factor;
        }
        G.lv[p.nlevels - 1].quota = std::max(p.nfeatures - sum, 0);
    }
    if (taps) {
        taps->assign(tap_off, CvTap{0, 0});
        for (int l = 1; l < p.nlevels; l++) {
            linear_exact_taps(G.lv[l - 1].w, G.lv[l].w, taps->data() + G.lv[l].tapx);
            linear_exact_taps(G.lv[l - 1].h, G.lv[l].h, taps->data() + G.lv[l].tapy);
        }
    }
    return true;
}

void cv_free(orb_cvorb *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    if (h->s_blur) cudaStreamSynchronize(h->s_blur);
    cudaFree(h->d_img); cudaFree(h->d_blur); cudaFree(h->d_score); cudaFree(h->d_rows); cudaFree(h->d_taps); cudaFree(h->d_masks);
    cudaFree(h->d_rowcount); cudaFree(h->d_cand_n); cudaFree(h->d_scratch); cudaFree(h->d_hdr); cudaFree(h->d_cand); cudaFree(h->d_geom);
    if (h->h_pin) cudaFreeHost(h->h_pin);
    if (h->h_in) cudaFreeHost(h->h_in);
    if (h->ev_pyr) cudaEventDestroy(h->ev_pyr);
    if (h->ev_blur) cudaEventDestroy(h->ev_blur);
    if (h->s_blur) cudaStreamDestroy(h->s_blur);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int cv_set_geometry(orb_cvorb *h, int width, int height) {
    if (h->cur_w == width && h->cur_h == height) return ORB_OK;
    std::vector<CvTap> taps;
    ORB_REQUIRE(cv_geometry(h->p, width, height, h->G, &taps), ORB_ERR_TOO_SMALL, "image %dx%d leaves an empty pyramid layer", width, height);
    h->G.out_cap = h->out_cap;
    ORB_CUDA_TRY(cudaStreamSynchronize(h->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(h->d_geom, &h->G, sizeof(CvGeom), cudaMemcpyHostToDevice, h->stream));
    if (!taps.empty()) ORB_CUDA_TRY(cudaMemcpyAsync(h->d_taps, taps.data(), taps.size() * sizeof(CvTap), cudaMemcpyHostToDevice, h->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(h->stream));   // the sources are locals
    h->cur_w = width; h->cur_h = height;
    return ORB_OK;
}

}  // namespace

extern "C" {

int orb_cvorb_create(const orb_cvorb_params *params, orb_cvorb **out) {
    ORB_REQUIRE(params && out, ORB_ERR_INVALID, "null argument");
    const orb_cvorb_params &p = *params;
    ORB_REQUIRE(p.nfeatures >= 0 && p.nfeatures <= 100000, ORB_ERR_INVALID, "nfeatures %d out of range", p.nfeatures);
    ORB_REQUIRE(p.scale_factor > 1.f && p.scale_factor <= 4.f, ORB_ERR_INVALID, "scale_factor must be in (1, 4]");
    ORB_REQUIRE(p.nlevels >= 1 && p.nlevels <= CV_MAXL, ORB_ERR_INVALID, "nlevels must be 1..%d", CV_MAXL);
    ORB_REQUIRE(p.edge_threshold >= 31 && p.edge_threshold <= 1024, ORB_ERR_INVALID, "edge_threshold must be >= 31 (patch 31)");
    ORB_REQUIRE(p.score_type == 0 || p.score_type == 1, ORB_ERR_INVALID, "score_type must be 0 (Harris) or 1 (FAST)");
    ORB_REQUIRE(p.fast_threshold >= 1 && p.fast_threshold <= 254, ORB_ERR_INVALID, "fast_threshold must be 1..254");
    ORB_REQUIRE(p.max_width >= 1 && p.max_height >= 1 && p.max_width <= 32768 && p.max_height <= 32768, ORB_ERR_INVALID, "bad maximum image size");
    int dev = p.device;
    if (dev < 0) ORB_CUDA_TRY(cudaGetDevice(&dev));
    ORB_CUDA_TRY(cudaSetDevice(dev));
    orb_cvorb *h = new orb_cvorb();
    h->p = p; h->device = dev;
    CvGeom G;
    if (!cv_geometry(p, p.max_width, p.max_height, G, nullptr)) {
        delete h;
        set_error("maximum image %dx%d leaves an empty pyramid layer", p.max_width, p.max_height);
        return ORB_ERR_TOO_SMALL;
    }
    const CvLevel &T = G.lv[p.nlevels - 1];
    // pools are sized for the maximum image; a smaller image of another aspect ratio needs no more of anything
    // except padding of the pitches, covered by the slack below
    h->plane_bytes = (size_t)(T.off + (long long)T.pitch * T.h) + (size_t)p.nlevels * 128 * (p.max_height + 8);
    h->rows_floats = (size_t)(T.foff + (long long)T.w * T.h) + 64;
    h->mask_words = (size_t)(T.mask_off + (long long)T.mask_words * T.nrows) + (size_t)G.total_rows + 64;
    h->tap_count = (size_t)T.tapy + T.h + 64;
    h->cand_total = (size_t)T.cand_off + T.cand_cap + 64;
    h->rows_total = (size_t)G.total_rows + 64;
    h->out_cap = (int)std::min<size_t>(h->cand_total, (size_t)std::max(4 * p.nfeatures, p.nfeatures + 4096));
    auto fail = [&](cudaError_t e, const char *what) {
        set_error("orb_cvorb_create: %s -> %s", what, cudaGetErrorString(e));
        cv_free(h);
        return ORB_ERR_CUDA;
    };
    cudaError_t e;
#define CV_TRY(x) if ((e = (x)) != cudaSuccess) return fail(e, #x)
    CV_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CV_TRY(cudaStreamCreateWithFlags(&h->s_blur, cudaStreamNonBlocking));
    CV_TRY(cudaEventCreateWithFlags(&h->ev_pyr, cudaEventDisableTiming));
    CV_TRY(cudaEventCreateWithFlags(&h->ev_blur, cudaEventDisableTiming));
    CV_TRY(cudaMalloc(&h->d_img, h->plane_bytes));
    CV_TRY(cudaMalloc(&h->d_blur, h->plane_bytes));
    CV_TRY(cudaMalloc(&h->d_score, h->plane_bytes));
    CV_TRY(cudaMalloc(&h->d_rows, h->rows_floats * sizeof(float)));
    CV_TRY(cudaMalloc(&h->d_taps, h->tap_count * sizeof(CvTap)));
    CV_TRY(cudaMalloc(&h->d_masks, h->mask_words * sizeof(uint32_t)));
    CV_TRY(cudaMalloc(&h->d_rowcount, h->rows_total * sizeof(int)));
    CV_TRY(cudaMalloc(&h->d_cand_n, CV_MAXL * sizeof(int)));
    CV_TRY(cudaMalloc(&h->d_scratch, (2 * h->cand_total + 4 * CV_MAXL) * sizeof(int)));
    CV_TRY(cudaMalloc(&h->d_hdr, 4 * sizeof(int)));
    CV_TRY(cudaMalloc(&h->d_cand, h->cand_total * sizeof(CvElem)));
    CV_TRY(cudaMalloc(&h->d_geom, sizeof(CvGeom)));
    CV_TRY(cudaMallocHost(&h->h_pin, 16 + (size_t)h->out_cap * (sizeof(orb_keypoint) + 32)));
    CV_TRY(cudaMallocHost(&h->h_in, (size_t)p.max_width * p.max_height));
    CV_TRY(cudaMemsetAsync(h->d_score, 0, h->plane_bytes, h->stream));
    CV_TRY(cudaMemsetAsync(h->d_hdr, 0, 4 * sizeof(int), h->stream));
    CV_TRY(cudaStreamSynchronize(h->stream));
#undef CV_TRY
    *out = h;
    return ORB_OK;
}

void orb_cvorb_destroy(orb_cvorb *h) { cv_free(h); }
int orb_cvorb_capacity(const orb_cvorb *h) { return h ? h->out_cap : 0; }
void *orb_cvorb_stream(orb_cvorb *h) { return h ? (void *)h->stream : nullptr; }
int64_t orb_cvorb_launch_count(const orb_cvorb *h) { return h ? h->launches : 0; }

int orb_cvorb_detect_and_compute(orb_cvorb *h, const uint8_t *image, int width, int height, int pitch, int image_loc,
                                 orb_keypoint *keypoints, uint8_t *descriptors, int capacity, int *n_out) {
    ORB_REQUIRE(h && n_out, ORB_ERR_INVALID, "null argument");
    *n_out = 0;
    if (!image || width <= 0 || height <= 0) return ORB_OK;   // `_image.empty()` -> return (orb.cpp)
    ORB_REQUIRE(width <= h->p.max_width && height <= h->p.max_height, ORB_ERR_INVALID, "image %dx%d exceeds the handle's %dx%d", width, height,
                h->p.max_width, h->p.max_height);
    ORB_REQUIRE(pitch >= width, ORB_ERR_INVALID, "pitch < width");
    ORB_REQUIRE(image_loc == ORB_HOST || image_loc == ORB_DEVICE, ORB_ERR_INVALID, "bad image_loc");
    ORB_CUDA_TRY(cudaSetDevice(h->device));
    int rc = cv_set_geometry(h, width, height);
    if (rc != ORB_OK) return rc;
    const CvGeom &G = h->G;
    cudaStream_t st = h->stream;
    if (image_loc == ORB_HOST && (size_t)width * height <= (1u << 20)) {   // small frames: one memcpy + a true DMA beats the driver's pageable path (larger ones: measured slower)
        for (int y = 0; y < height; y++) memcpy(h->h_in + (size_t)y * width, image + (size_t)y * pitch, width);
        ORB_CUDA_TRY(cudaMemcpy2DAsync(h->d_img + G.lv[0].off, G.lv[0].pitch, h->h_in, width, width, height, cudaMemcpyHostToDevice, st));
    } else {
        ORB_CUDA_TRY(cudaMemcpy2DAsync(h->d_img + G.lv[0].off, G.lv[0].pitch, image, pitch, width, height,
                                       image_loc == ORB_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice, st));
    }
    for (int l = 1; l < G.nlevels; l++) {
        const CvLevel &S = G.lv[l - 1], &D = G.lv[l];
        dim3 grid((D.w + 4 * 256 - 1) / (4 * 256), D.h);
        k_cv_resize<<<grid, 256, 0, st>>>(h->d_img + S.off, S.w, S.h, S.pitch, h->d_img + D.off, D.w, D.h, D.pitch, h->d_taps + D.tapx,
                                          h->d_taps + D.tapy);
        h->launches++;
    }
    // blur of every layer beside the detector
    ORB_CUDA_TRY(cudaEventRecord(h->ev_pyr, st));
    ORB_CUDA_TRY(cudaStreamWaitEvent(h->s_blur, h->ev_pyr, 0));
    {
        int rows = 0;
        for (int l = 0; l < G.nlevels; l++) rows += G.lv[l].h;
        dim3 grid((G.lv[0].w + 255) / 256, rows);
        k_cv_blur_rows<<<grid, 256, 0, h->s_blur>>>(h->d_img, h->d_rows, h->d_geom);
        k_cv_blur_cols<<<grid, 256, 0, h->s_blur>>>(h->d_rows, h->d_blur, h->d_geom);
        h->launches += 2;
        ORB_CUDA_TRY(cudaEventRecord(h->ev_blur, h->s_blur));
    }
    ORB_CUDA_TRY(cudaMemsetAsync(h->d_cand_n, 0, CV_MAXL * sizeof(int), st));
    if (G.total_rows > 0) {
        int srows = 0, maxq = 0;
        for (int l = 0; l < G.nlevels; l++) srows += G.lv[l].nrows > 0 ? G.lv[l].nrows + 2 : 0;
        dim3 gs((G.lv[0].w - 2 * G.edge + 2 + 255) / 256, srows);
        k_cv_fast_score<<<gs, 256, 0, st>>>(h->d_img, h->d_score, h->d_geom);
        const int rb = (G.total_rows + 7) / 8;
        k_cv_nms_rows<<<rb, 256, 0, st>>>(h->d_score, h->d_masks, h->d_rowcount, h->d_geom);
        k_cv_compact<<<rb, 256, 0, st>>>(h->d_score, h->d_masks, h->d_rowcount, h->d_cand, h->d_cand_n, h->d_hdr + 1, h->d_geom);
        k_cv_retain<<<G.nlevels, CV_RETAIN_T, 0, st>>>(h->d_cand, h->d_cand_n, h->d_scratch, h->d_geom, 0);
        h->launches += 4;
        // after the first cut a layer holds at most its candidates; size the per-key-point grids by the quota plus ties
        // (grid-stride loops: ties can keep more than 2 x quota)
        for (int l = 0; l < G.nlevels; l++) maxq = std::max(maxq, std::min(G.lv[l].cand_cap, 2 * G.lv[l].quota + 64));
        if (G.harris) {
            dim3 gh((maxq + 7) / 8, G.nlevels);
            k_cv_harris<<<gh, 256, 0, st>>>(h->d_img, h->d_cand, h->d_cand_n, h->d_geom);
            k_cv_retain<<<G.nlevels, CV_RETAIN_T, 0, st>>>(h->d_cand, h->d_cand_n, h->d_scratch, h->d_geom, 1);
            h->launches += 2;
        }
        ORB_CUDA_TRY(cudaStreamWaitEvent(st, h->ev_blur, 0));
        dim3 gd((maxq + 7) / 8, G.nlevels);
        k_cv_describe<<<gd, 256, 0, st>>>(h->d_img, h->d_blur, h->d_cand, h->d_cand_n, h->d_geom, (orb_keypoint *)(h->h_pin + 16),
                                          h->h_pin + 16 + (size_t)h->out_cap * sizeof(orb_keypoint), h->d_hdr, h->d_hdr + 1);   // results go straight to pinned host memory
        h->launches++;
    } else {
        ORB_CUDA_TRY(cudaStreamWaitEvent(st, h->ev_blur, 0));
        ORB_CUDA_TRY(cudaMemsetAsync(h->d_hdr, 0, sizeof(int), st));
    }
    ORB_CUDA_TRY(cudaGetLastError());
    int *hdr = (int *)h->h_pin;
    ORB_CUDA_TRY(cudaMemcpyAsync(hdr, h->d_hdr, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    const int n = hdr[0], flags = hdr[1];
    if (flags) {
        ORB_CUDA_TRY(cudaMemsetAsync(h->d_hdr + 1, 0, sizeof(int), st));
        *n_out = n;
        ORB_REQUIRE(false, ORB_ERR_CAPACITY, "cv::ORB: %s", (flags & 1) ? "candidate list overflow" : "more key points than the handle's output capacity");
    }
    *n_out = n;
    if (n == 0) return ORB_OK;
    ORB_REQUIRE(n <= capacity && keypoints && descriptors, ORB_ERR_CAPACITY, "capacity %d < %d key points", capacity, n);
    const uint8_t *hk = h->h_pin + 16, *hd = hk + (size_t)h->out_cap * sizeof(orb_keypoint);   // written by k_cv_describe, complete before the header copy
    memcpy(keypoints, hk, (size_t)n * sizeof(orb_keypoint));
    memcpy(descriptors, hd, (size_t)n * 32);
    return ORB_OK;
}

int orb_cvorb_level(orb_cvorb *h, int level, uint8_t *dst, int dst_pitch, int *width, int *height) {
    ORB_REQUIRE(h && h->cur_w > 0 && level >= 0 && (level & 0xff) < h->G.nlevels, ORB_ERR_INVALID, "no such layer");
    const CvLevel &L = h->G.lv[level & 0xff];
    if (width) *width = L.w;
    if (height) *height = L.h;
    if (dst) {
        ORB_REQUIRE(dst_pitch >= L.w, ORB_ERR_INVALID, "dst_pitch < width");
        ORB_CUDA_TRY(cudaSetDevice(h->device));
        const uint8_t *pool = (level >> 8) == 1 ? h->d_img : (level >> 8) == 2 ? h->d_score : h->d_blur;   // debugging: level | 256 = unblurred, | 512 = FAST scores
        ORB_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_pitch, pool + L.off, L.pitch, L.w, L.h, cudaMemcpyDeviceToHost, h->stream));
        ORB_CUDA_TRY(cudaStreamSynchronize(h->stream));
    }
    return ORB_OK;
}

/* tests: KeyPointsFilter::retainBest on `n` responses (host arrays); order_out receives the surviving original indices in
 * the order libstdc++'s nth_element + partition leave them.  depth_limit < 0 = the library's own. */
int orb_debug_retain_best(const float *responses, int n, int n_points, int depth_limit, int32_t *order_out, int *n_out) {
    ORB_REQUIRE(responses && order_out && n_out && n >= 0, ORB_ERR_INVALID, "bad argument");
    ThreadCtx *ctx = nullptr;
    const size_t bytes = (size_t)(n + 8) * (sizeof(CvElem) + 2 * sizeof(int)) + 64;
    int rc = orbgpu_thread_ctx(&ctx, bytes, 0);
    if (rc != ORB_OK) return rc;
    std::vector<CvElem> e(n);
    for (int i = 0; i < n; i++) e[i] = CvElem{responses[i], (uint32_t)i};
    CvElem *dE = (CvElem *)ctx->d_buf;
    int *dS = (int *)(dE + n + 4), *dN = dS + 2 * n + 8;
    if (n) ORB_CUDA_TRY(cudaMemcpyAsync(dE, e.data(), (size_t)n * sizeof(CvElem), cudaMemcpyHostToDevice, ctx->stream));
    k_cv_retain_debug<<<1, CV_RETAIN_T, 0, ctx->stream>>>(dE, n, n_points, dS, depth_limit, dN);
    ORB_CUDA_TRY(cudaGetLastError());
    int m = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(&m, dN, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    if (m) ORB_CUDA_TRY(cudaMemcpyAsync(e.data(), dE, (size_t)m * sizeof(CvElem), cudaMemcpyDeviceToHost, ctx->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < m; i++) order_out[i] = (int32_t)e[i].xy;
    *n_out = m;
    return ORB_OK;
}

}  // extern "C"
