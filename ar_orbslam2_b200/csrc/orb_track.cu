// liborbgpu -- frame feature index and the tracking-time Hamming matchers for sm_100a.
//
// Replaces the hot loops of ORB_SLAM2::ORBmatcher (ref: ORB_SLAM2/src/ORBmatcher.cc) and the parts of
// Frame they read (ref: ORB_SLAM2/src/Frame.cc:235-250, 332-397):
//   orb_frame_*                      mvKeysUn / mDescriptors / mvuRight + the 64x48 grid on the device
//   orb_search_windowed              the projection-matcher loop (SearchByProjection x4, Fuse, SearchBySim3)
//   orb_search_for_initialization    ORBmatcher.cc:405-523
//   orb_search_by_bow                ORBmatcher.cc:159-288, 525-658
//   orb_rot_hist_filter              ORBmatcher.cc:1604-1645 + the 30-bin rotation histogram
//
// Structure: distances for every (query, candidate) pair are computed in parallel, one warp per query, with
// candidates emitted in the reference's grid traversal order; a single warp then replays the queries in order
// and resolves the "already claimed" dependencies with warp-level lexicographic (distance, rank) top-2
// reductions, which is exactly the reference's sequential strict-'<' best/second scan.
#include "orb_common.cuh"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

namespace orbgpu {

constexpr int GRID_COLS = 64, GRID_ROWS = 48;   // ref Frame.h:39-40
constexpr int TH_HIGH_M = 100, TH_LOW_M = 50, HISTO = 30;
constexpr uint32_t SENT = 0xffffffffu;

struct FrameDev {
    int n;
    const float *x, *y, *angle, *uRight;
    const int *octave;
    const uint8_t *desc;
    const int *cell_off;   // [GRID_COLS*GRID_ROWS + 1], cell = ix*GRID_ROWS + iy
    const int *cell_idx;   // keypoint indices, ascending inside a cell (= insertion order)
    float minX, minY, invW, invH;   // minX/minY as GetFeaturesInArea sees them (truncated to int for a KeyFrame view)
};

struct WinQ {   // device view of orb_window_queries
    int n;
    const uint8_t *valid;
    const float *x, *y, *radius;
    const int *min_level, *max_level;
    int strict_levels;
    const float *stereo_ref, *stereo_tol;
    const uint8_t *desc;
    const float *angle;
    const uint8_t *claims;
    const float *gate_ur;          // reprojection-error gate of Fuse (ref ORBmatcher.cc:917-945); gate_inv_sigma2 == NULL = off
    const float *gate_inv_sigma2;
    double gate_mono, gate_stereo;
};

__device__ __forceinline__ int warp_excl_scan(int v, int lane, int *total) {
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    *total = __shfl_sync(0xffffffffu, incl, 31);
    return incl - v;
}

__device__ __forceinline__ void top2_push(uint32_t &k1, uint32_t &k2, uint32_t key) {
    if (key < k1) { k2 = k1; k1 = key; }
    else if (key < k2) k2 = key;
}
__device__ __forceinline__ void warp_top2(uint32_t &k1, uint32_t &k2) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const uint32_t a1 = __shfl_xor_sync(0xffffffffu, k1, o), a2 = __shfl_xor_sync(0xffffffffu, k2, o);
        const uint32_t n1 = min(k1, a1), n2 = min(max(k1, a1), min(k2, a2));
        k1 = n1; k2 = n2;
    }
}

__device__ __forceinline__ int hamming_rows(const uint8_t *a, const uint8_t *b) {
    const uint4 *pa = reinterpret_cast<const uint4 *>(a), *pb = reinterpret_cast<const uint4 *>(b);
    const uint4 a0 = __ldg(pa), a1 = __ldg(pa + 1), b0 = __ldg(pb), b1 = __ldg(pb + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// Frame::GetFeaturesInArea (ref Frame.cc:332-385) for one query by one warp.  Lanes own grid cells in traversal
// order (ix outer, iy inner); a warp prefix sum per 32-cell round keeps the output in the reference's order.
// FILL=false: returns the count.  FILL=true: writes (index, distance) pairs at out_idx/out_dist.
struct WinTest {   // one query's admission test for a feature of the window's cells
    float x, y, r, sref, stol, gur;
    int minL, maxL;
    bool check, strict, stereo, gate;
    const float *inv_sigma2;
    double chi_mono, chi_stereo;
    __device__ __forceinline__ bool ok(const FrameDev &f, int idx) const {
        const int oct = f.octave[idx];
        if (strict) { if (oct < minL || oct > maxL) return false; }
        else if (check) { if (oct < minL || (maxL >= 0 && oct > maxL)) return false; }
        const float kx = f.x[idx], ky = f.y[idx];
        if (!(fabsf(__fsub_rn(kx, x)) < r && fabsf(__fsub_rn(ky, y)) < r)) return false;
        if (stereo) {
            const float ur = f.uRight[idx];
            if (ur > 0 && fabsf(__fsub_rn(sref, ur)) > stol) return false;
        }
        if (gate) {   // ref :917-945: squared reprojection error, scaled by the level's inverse sigma^2, against chi^2 (double compare)
            const float ex = __fsub_rn(x, kx), ey = __fsub_rn(y, ky), kr = f.uRight[idx];
            float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
            if (kr >= 0) {
                const float er = __fsub_rn(gur, kr);
                e2 = __fadd_rn(e2, __fmul_rn(er, er));
                if ((double)__fmul_rn(e2, inv_sigma2[oct]) > chi_stereo) return false;
            } else if ((double)__fmul_rn(e2, inv_sigma2[oct]) > chi_mono) return false;
        }
        return true;
    }
};

template <bool FILL>
__device__ int window_candidates(const FrameDev &f, const WinQ &Q, int q, int lane, int *out_idx, int *out_dist) {
    const float x = Q.x[q], y = Q.y[q], r = Q.radius[q];
    const int minL = Q.min_level[q], maxL = Q.max_level[q];
    const int cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, f.minX), r), f.invW)));
    if (cx0 >= GRID_COLS) return 0;
    const int cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, f.minX), r), f.invW)));
    if (cx1 < 0) return 0;
    const int cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, f.minY), r), f.invH)));
    if (cy0 >= GRID_ROWS) return 0;
    const int cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, f.minY), r), f.invH)));
    if (cy1 < 0) return 0;
    if (cx1 < cx0 || cy1 < cy0) return 0;
    WinTest T;
    T.x = x; T.y = y; T.r = r; T.minL = minL; T.maxL = maxL;
    T.check = (minL > 0) || (maxL >= 0);
    T.strict = Q.strict_levels != 0;
    T.stereo = Q.stereo_ref != nullptr;
    T.sref = T.stereo ? Q.stereo_ref[q] : 0.f; T.stol = T.stereo ? Q.stereo_tol[q] : 0.f;
    T.gate = Q.gate_inv_sigma2 != nullptr;
    T.gur = T.gate && Q.gate_ur ? Q.gate_ur[q] : 0.f;
    T.inv_sigma2 = Q.gate_inv_sigma2; T.chi_mono = Q.gate_mono; T.chi_stereo = Q.gate_stereo;
    const int ny = cy1 - cy0 + 1, ncells = (cx1 - cx0 + 1) * ny;
    const uint8_t *qd = Q.desc + (size_t)q * 32;
    int base = 0;
    for (int c0 = 0; c0 < ncells; c0 += 32) {
        const int c = c0 + lane;
        int beg = 0, end = 0;
        if (c < ncells) {
            const int cell = (cx0 + c / ny) * GRID_ROWS + cy0 + c % ny;
            beg = f.cell_off[cell]; end = f.cell_off[cell + 1];
        }
        int cnt = 0;
        unsigned pass = 0;   // cells hold few keypoints; longer cells fall back to re-testing
        for (int e = beg; e < end; e++) {
            if (T.ok(f, f.cell_idx[e])) { if (e - beg < 32) pass |= 1u << (e - beg); cnt++; }
        }
        int total;
        const int off = warp_excl_scan(cnt, lane, &total);
        if (FILL && cnt) {
            int w = base + off;
            for (int e = beg; e < end; e++) {
                bool ok;
                const int idx = f.cell_idx[e];
                if (e - beg < 32) ok = (pass >> (e - beg)) & 1;
                else ok = T.ok(f, idx);
                if (ok) {
                    out_idx[w] = idx;
                    out_dist[w] = hamming_rows(qd, f.desc + (size_t)idx * 32);
                    w++;
                }
            }
        }
        base += total;
    }
    return base;
}

__global__ void k_window_count(FrameDev f, WinQ Q, int *counts) {
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= Q.n) return;
    int n = 0;
    if (!Q.valid || Q.valid[q]) n = window_candidates<false>(f, Q, q, lane, nullptr, nullptr);
    if (lane == 0) counts[q] = n;
}
__global__ void k_window_fill(FrameDev f, WinQ Q, const int *off, int *cand_idx, int *cand_dist) {
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= Q.n) return;
    if (Q.valid && !Q.valid[q]) return;
    window_candidates<true>(f, Q, q, lane, cand_idx + off[q], cand_dist + off[q]);
}

// exclusive scan of counts[0..n) into off[0..n], one CTA
__global__ void __launch_bounds__(1024) k_scan_counts(const int *counts, int n, int *off) {
    __shared__ int warp_tot[33];
    __shared__ int carry;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const int v = i < n ? counts[i] : 0;
        int tot;
        int ex = warp_excl_scan(v, lane, &tot);
        if (lane == 31) warp_tot[tid >> 5] = ex + v;
        __syncthreads();
        if (tid < 32) {
            int t;
            const int w = warp_tot[tid];
            const int e = warp_excl_scan(w, tid, &t);
            warp_tot[tid] = e;
            if (tid == 31) warp_tot[32] = t;
        }
        __syncthreads();
        if (i < n) off[i] = carry + warp_tot[tid >> 5] + ex;
        __syncthreads();
        if (tid == 0) carry += warp_tot[32];
        __syncthreads();
    }
    if (tid == 0) off[n] = carry;
}

__device__ __forceinline__ int rot_bin(float a, float b) {   // ref ORBmatcher.cc:238-243, factor = 1/HISTO_LENGTH
    float rot = __fsub_rn(a, b);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 1.0f / HISTO));
    if (bin == HISTO) bin = 0;
    return bin;
}

__device__ __forceinline__ void three_maxima(const int *hist, int &ind1, int &ind2, int &ind3) {   // ref :1604-1645
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < HISTO; i++) {
        const int s = hist[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) { ind3 = -1; }
}

// One warp replays the queries in order (ORBmatcher.cc:51-126 / 1354-1448).
__global__ void __launch_bounds__(32)
k_resolve_windowed(FrameDev f, WinQ Q, const int *off, const int *cand_idx, const int *cand_dist, int *holder, uint8_t *blocking,
                   int select, int th_dist, float nnratio, int check_ori, int *acc_feat, int *acc_bin, int *assigned, int *nmatches) {
    __shared__ int hist[HISTO];
    __shared__ int s_nm;
    const int lane = threadIdx.x;
    if (lane < HISTO) hist[lane] = 0;
    if (lane == 0) s_nm = 0;
    __syncwarp();
    int nm = 0, na = 0;
    for (int q = 0; q < Q.n; q++) {
        if (Q.valid && !Q.valid[q]) continue;
        const int o = off[q], n = off[q + 1] - o;
        if (n == 0) continue;
        uint32_t k1 = SENT, k2 = SENT;
        for (int j = lane; j < n; j += 32) {
            const int idx = cand_idx[o + j];
            if (holder[idx] != -1 && blocking[idx]) continue;     // ref :87-89
            top2_push(k1, k2, ((uint32_t)cand_dist[o + j] << 16) | (uint32_t)j);
        }
        warp_top2(k1, k2);
        if (k1 == SENT) continue;
        const int bestDist = (int)(k1 >> 16);
        if (bestDist > th_dist) continue;
        const int bestIdx = cand_idx[o + (k1 & 0xffff)];
        if (select == ORB_SELECT_RATIO_LEVEL) {
            const int bestLevel = f.octave[bestIdx];
            const int bestLevel2 = k2 == SENT ? -1 : f.octave[cand_idx[o + (k2 & 0xffff)]];
            const int bestDist2 = k2 == SENT ? 256 : (int)(k2 >> 16);
            if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) continue;   // ref :120
        }
        if (lane == 0) {
            holder[bestIdx] = q;
            blocking[bestIdx] = Q.claims ? (Q.claims[q] != 0) : 1;
            if (check_ori) {
                const int bin = rot_bin(Q.angle[q], f.angle[bestIdx]);
                acc_feat[na] = bestIdx; acc_bin[na] = bin;
                hist[bin]++;
            }
        }
        nm++; na++;
        __syncwarp();
    }
    if (check_ori) {
        int i1, i2, i3;
        three_maxima(hist, i1, i2, i3);
        int dec = 0;
        for (int a = lane; a < na; a += 32) {
            const int b = acc_bin[a];
            if (b != i1 && b != i2 && b != i3) { holder[acc_feat[a]] = -3; dec++; }   // ref :1463-1467 (set to NULL)
        }
        if (dec) atomicAdd(&s_nm, dec);
        __syncwarp();
        nm -= s_nm;
    }
    __syncwarp();
    for (int i = lane; i < f.n; i += 32) assigned[i] = holder[i] >= 0 ? holder[i] : (holder[i] == -3 ? -2 : -1);
    if (lane == 0) *nmatches = nm;
}

// ORBmatcher::SearchForInitialization resolve pass (ref :418-520)
__global__ void __launch_bounds__(32)
k_resolve_init(FrameDev f1, FrameDev f2, const int *off, const int *cand_idx, const int *cand_dist, int *matched_dist, int *m21,
               int *m12, int *qbin, float nnratio, int check_ori, float *prev_matched, int *nmatches) {
    __shared__ int hist[HISTO];
    __shared__ int s_dec;
    const int lane = threadIdx.x;
    if (lane < HISTO) hist[lane] = 0;
    if (lane == 0) s_dec = 0;
    __syncwarp();
    int nm = 0;
    for (int i1 = 0; i1 < f1.n; i1++) {
        const int o = off[i1], n = off[i1 + 1] - o;
        if (n == 0) continue;
        uint32_t k1 = SENT, k2 = SENT;
        for (int j = lane; j < n; j += 32) {
            const int d = cand_dist[o + j];
            if (matched_dist[cand_idx[o + j]] <= d) continue;   // ref :444-445
            top2_push(k1, k2, ((uint32_t)d << 16) | (uint32_t)j);
        }
        warp_top2(k1, k2);
        if (k1 == SENT) continue;
        const int bestDist = (int)(k1 >> 16);
        if (bestDist > TH_LOW_M) continue;
        const float second = k2 == SENT ? (float)0x7fffffff : (float)(int)(k2 >> 16);
        if (!((float)bestDist < __fmul_rn(second, nnratio))) continue;   // ref :461
        const int b = cand_idx[o + (k1 & 0xffff)];
        const int prev = m21[b];
        __syncwarp();   // every lane has read m21[b] before lane 0 overwrites it
        if (lane == 0) {
            if (prev >= 0) m12[prev] = -1;
            m12[i1] = b; m21[b] = i1; matched_dist[b] = bestDist;
            if (check_ori) {
                const int bin = rot_bin(f1.angle[i1], f2.angle[b]);
                qbin[i1] = bin;
                hist[bin]++;
            }
        }
        if (prev >= 0) nm--;
        nm++;
        __syncwarp();
    }
    if (check_ori) {
        int a1, a2, a3;
        three_maxima(hist, a1, a2, a3);
        int dec = 0;
        for (int i = lane; i < f1.n; i += 32) {
            const int b = qbin[i];
            if (b >= 0 && b != a1 && b != a2 && b != a3 && m12[i] >= 0) { m12[i] = -1; dec++; }   // ref :500-513
        }
        if (dec) atomicAdd(&s_dec, dec);
        __syncwarp();
        nm -= s_dec;
    }
    __syncwarp();
    for (int i = lane; i < f1.n; i += 32)
        if (m12[i] >= 0) { prev_matched[2 * i] = f2.x[m12[i]]; prev_matched[2 * i + 1] = f2.y[m12[i]]; }   // ref :518-520
    if (lane == 0) *nmatches = nm;
}

// SearchByBoW: distances of every (KF entry, F entry of the same vocabulary node) pair, one warp per KF entry
__global__ void k_bow_dist(FrameDev kf, FrameDev f, int nentries, const int *kf_idx, const int *ent_fo, const int *ent_fn,
                           const int *ent_po, const int *f_idx, const uint8_t *kf_valid, int *dist) {
    const int a = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (a >= nentries) return;
    const int ki = kf_idx[a];
    if (!kf_valid[ki]) return;
    const uint8_t *d = kf.desc + (size_t)ki * 32;
    const int fo = ent_fo[a], fn = ent_fn[a], po = ent_po[a];
    for (int j = lane; j < fn; j += 32) dist[po + j] = hamming_rows(d, f.desc + (size_t)f_idx[fo + j] * 32);
}

__global__ void __launch_bounds__(32)
k_resolve_bow(FrameDev kf, FrameDev f, int nentries, const int *kf_idx, const int *ent_fo, const int *ent_fn, const int *ent_po,
              const int *f_idx, const uint8_t *kf_valid, const uint8_t *f_valid, const int *dist, int strict_low, float nnratio,
              int check_ori, int *acc_feat, int *acc_bin, int *matches, int *nmatches) {
    __shared__ int hist[HISTO];
    __shared__ int s_dec;
    const int lane = threadIdx.x;
    if (lane < HISTO) hist[lane] = 0;
    if (lane == 0) s_dec = 0;
    __syncwarp();
    int nm = 0, na = 0;
    for (int a = 0; a < nentries; a++) {
        const int ki = kf_idx[a];
        if (!kf_valid[ki]) continue;
        const int fo = ent_fo[a], fn = ent_fn[a], po = ent_po[a];
        uint32_t k1 = SENT, k2 = SENT;
        for (int j = lane; j < fn; j += 32) {
            const int fi = f_idx[fo + j];
            if (matches[fi] >= 0) continue;                       // ref :209-210
            if (f_valid && !f_valid[fi]) continue;                // ref :577-583
            top2_push(k1, k2, ((uint32_t)dist[po + j] << 16) | (uint32_t)j);
        }
        warp_top2(k1, k2);
        if (k1 == SENT) continue;
        const int best1 = (int)(k1 >> 16), best2 = k2 == SENT ? 256 : (int)(k2 >> 16);
        if (strict_low ? !(best1 < TH_LOW_M) : !(best1 <= TH_LOW_M)) continue;
        if (!((float)best1 < __fmul_rn(nnratio, (float)best2))) continue;   // ref :230
        const int fi = f_idx[fo + (k1 & 0xffff)];
        if (lane == 0) {
            matches[fi] = ki;
            if (check_ori) {
                const int bin = rot_bin(kf.angle[ki], f.angle[fi]);
                acc_feat[na] = fi; acc_bin[na] = bin;
                hist[bin]++;
            }
        }
        nm++; na++;
        __syncwarp();
    }
    if (check_ori) {
        int i1, i2, i3;
        three_maxima(hist, i1, i2, i3);
        int dec = 0;
        for (int a = lane; a < na; a += 32) {
            const int b = acc_bin[a];
            if (b != i1 && b != i2 && b != i3) { matches[acc_feat[a]] = -1; dec++; }
        }
        if (dec) atomicAdd(&s_dec, dec);
        __syncwarp();
        nm -= s_dec;
    }
    if (lane == 0) *nmatches = nm;
}

// Fuse / SearchBySim3 candidate loop (ref ORBmatcher.cc:895-951, 1052-1072, 1191-1213, 1271-1293): independent queries,
// closest descriptor of the window, the first one in traversal order on ties.  One warp per query over its CSR list.
__global__ void k_window_best(int nq, const int *off, const int *cand_idx, const int *cand_dist, int *best_idx, int *best_dist) {
    const int q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (q >= nq) return;
    const int o = off[q], n = off[q + 1] - o;
    unsigned long long key = ~0ull;
    for (int j = lane; j < n; j += 32) key = min(key, ((unsigned long long)(uint32_t)cand_dist[o + j] << 32) | (uint32_t)j);
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, s));
    if (lane == 0) {
        best_idx[q] = key == ~0ull ? -1 : cand_idx[o + (int)(uint32_t)key];
        best_dist[q] = key == ~0ull ? 0x7fffffff : (int)(key >> 32);
    }
}

// SearchForTriangulation (ref ORBmatcher.cc:660-826): one warp per KF1 feature of a shared vocabulary node, lanes over
// the node's KF2 features.  The running `dist>bestDist` test only ever rejects candidates that lose anyway, so the result
// is the admissible candidate of least distance, the LAST one in node order on ties (`dist>bestDist` lets equals through).
struct TriArgs {
    float F[9];
    float ex, ey;
    const float *sigma2, *scale2;   // pKF2->mvLevelSigma2, pKF2->mvScaleFactors
    int only_stereo;
};
__global__ void k_triangulate(FrameDev k1, FrameDev k2, int nent, const int *idx1, const int *ent_o2, const int *ent_n2,
                              const int *idx2, const uint8_t *usable1, const uint8_t *usable2, TriArgs A, int *m12) {
    const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (e >= nent) return;
    const int i1 = idx1[e];
    if (!usable1[i1]) return;                                            // ref :704-708
    const bool st1 = k1.uRight[i1] >= 0;
    if (A.only_stereo && !st1) return;                                   // ref :712-714
    const float x1 = k1.x[i1], y1 = k1.y[i1];
    // epipolar line l = x1' F12 (ref :143-145), plain fp32 multiplies and adds in source order
    const float a = __fadd_rn(__fadd_rn(__fmul_rn(x1, A.F[0]), __fmul_rn(y1, A.F[3])), A.F[6]);
    const float b = __fadd_rn(__fadd_rn(__fmul_rn(x1, A.F[1]), __fmul_rn(y1, A.F[4])), A.F[7]);
    const float c = __fadd_rn(__fadd_rn(__fmul_rn(x1, A.F[2]), __fmul_rn(y1, A.F[5])), A.F[8]);
    const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
    const uint8_t *d1 = k1.desc + (size_t)i1 * 32;
    const int o2 = ent_o2[e], n2 = ent_n2[e];
    uint32_t key = SENT;
    for (int j = lane; j < n2; j += 32) {
        const int i2 = idx2[o2 + j];
        if (!usable2[i2]) continue;                                      // ref :726-730
        const bool st2 = k2.uRight[i2] >= 0;
        if (A.only_stereo && !st2) continue;                             // ref :734-736
        const int dist = hamming_rows(d1, k2.desc + (size_t)i2 * 32);
        if (dist > TH_LOW_M) continue;                                   // ref :742
        const float x2 = k2.x[i2], y2 = k2.y[i2];
        const int oct = k2.octave[i2];
        if (!st1 && !st2) {                                              // ref :747-753: too close to the epipole
            const float dx = __fsub_rn(A.ex, x2), dy = __fsub_rn(A.ey, y2);
            if (__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)) < __fmul_rn(100.0f, A.scale2[oct])) continue;
        }
        if (den == 0) continue;                                          // ref :151-152
        const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, x2), __fmul_rn(b, y2)), c);
        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
        if (!((double)dsqr < __dmul_rn(3.84, (double)A.sigma2[oct]))) continue;   // ref :156
        key = min(key, ((uint32_t)dist << 16) | (uint32_t)(0xffff - j));
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, s));
    if (lane == 0 && key != SENT) m12[i1] = idx2[o2 + (0xffff - (int)(key & 0xffff))];
}

// rotation-consistency filter of SearchForTriangulation (ref :764-775, 796-814) and the final count
__global__ void __launch_bounds__(1024) k_tri_filter(FrameDev k1, FrameDev k2, int *m12, int check_ori, int *nmatches) {
    __shared__ int hist[HISTO];
    __shared__ int ind[3];
    __shared__ int s_n;
    const int tid = threadIdx.x;
    if (tid < HISTO) hist[tid] = 0;
    if (tid == 0) s_n = 0;
    __syncthreads();
    if (check_ori) {
        for (int i = tid; i < k1.n; i += 1024)
            if (m12[i] >= 0) atomicAdd(&hist[rot_bin(k1.angle[i], k2.angle[m12[i]])], 1);
        __syncthreads();
        if (tid == 0) three_maxima(hist, ind[0], ind[1], ind[2]);
        __syncthreads();
    }
    int cnt = 0;
    for (int i = tid; i < k1.n; i += 1024) {
        if (m12[i] < 0) continue;
        if (check_ori) {
            const int bin = rot_bin(k1.angle[i], k2.angle[m12[i]]);
            if (bin != ind[0] && bin != ind[1] && bin != ind[2]) { m12[i] = -1; continue; }
        }
        cnt++;
    }
    if (cnt) atomicAdd(&s_n, cnt);
    __syncthreads();
    if (tid == 0) *nmatches = s_n;
}

__global__ void __launch_bounds__(1024) k_rot_hist_filter(const float *a, const float *b, int n, uint8_t *keep) {
    __shared__ int hist[HISTO];
    __shared__ int ind[3];
    const int tid = threadIdx.x;
    if (tid < HISTO) hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += 1024) atomicAdd(&hist[rot_bin(a[i], b[i])], 1);
    __syncthreads();
    if (tid == 0) three_maxima(hist, ind[0], ind[1], ind[2]);
    __syncthreads();
    for (int i = tid; i < n; i += 1024) {
        const int bin = rot_bin(a[i], b[i]);
        keep[i] = (bin == ind[0] || bin == ind[1] || bin == ind[2]) ? 1 : 0;
    }
}

}  // namespace orbgpu

using namespace orbgpu;

struct orb_frame {
    int device = 0;
    cudaStream_t stream = nullptr;
    int n = 0, nlevels = 0;
    std::vector<orb_keypoint> keys;
    std::vector<float> scale;
    float minX, maxX, minY, maxY;
    void *d_static = nullptr;   // x,y,angle,uRight,octave,desc,cell_off,cell_idx
    FrameDev dev;
    void *d_scratch = nullptr;
    size_t scratch_bytes = 0;
    int64_t launches = 0;
};

namespace {

struct Arena {   // carve 256-byte aligned pieces out of the frame's scratch buffer (two-phase: size, then place)
    uint8_t *base = nullptr;
    size_t used = 0;
    template <class T> T *take(size_t count) {
        T *p = base ? reinterpret_cast<T *>(base + used) : nullptr;
        used += (count * sizeof(T) + 255) / 256 * 256;
        return p;
    }
};

int ensure_scratch(orb_frame *f, size_t bytes) {
    if (f->scratch_bytes >= bytes) return ORB_OK;
    if (f->d_scratch) cudaFree(f->d_scratch);
    f->d_scratch = nullptr; f->scratch_bytes = 0;
    bytes = bytes * 3 / 2 + 4096;
    ORB_CUDA_TRY(cudaMalloc(&f->d_scratch, bytes));
    f->scratch_bytes = bytes;
    return ORB_OK;
}

template <class T>
int upload(T *dst, const T *src, size_t count, cudaStream_t st) {
    if (count) ORB_CUDA_TRY(cudaMemcpyAsync(dst, src, count * sizeof(T), cudaMemcpyHostToDevice, st));
    return ORB_OK;
}
#define ORB_TRY(expr) do { int _rc = (expr); if (_rc != ORB_OK) return _rc; } while (0)

// Runs count -> scan -> fill for a query set living in `ar` (device pointers already placed).  Returns the
// total number of candidates; cand arrays are allocated after the count is known.
struct WindowLists { int *off = nullptr, *idx = nullptr, *dist = nullptr; int total = 0; };

}  // namespace

extern "C" {

int orb_frame_create(const orb_frame_desc *d, orb_frame **out) {
    ORB_REQUIRE(d && out && d->n >= 0 && (d->n == 0 || (d->keys_un && d->descriptors)) && d->scale_factors && d->nlevels >= 1,
                ORB_ERR_INVALID, "bad frame description");
    ORB_REQUIRE(d->max_x > d->min_x && d->max_y > d->min_y, ORB_ERR_INVALID, "empty image bounds");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    orb_frame *f = new orb_frame();
    cudaGetDevice(&f->device);
    f->n = d->n; f->nlevels = d->nlevels;
    f->keys.assign(d->keys_un, d->keys_un + d->n);
    f->scale.assign(d->scale_factors, d->scale_factors + d->nlevels);
    f->minX = d->min_x; f->maxX = d->max_x; f->minY = d->min_y; f->maxY = d->max_y;
    const int n = d->n, ncell = GRID_COLS * GRID_ROWS;
    // ref Frame.cc:217-218 and AssignFeaturesToGrid / PosInGrid :235-250, 387-397 (round(), not floor)
    const float invW = static_cast<float>(GRID_COLS) / static_cast<float>(f->maxX - f->minX);
    const float invH = static_cast<float>(GRID_ROWS) / static_cast<float>(f->maxY - f->minY);
    std::vector<float> hx(n), hy(n), ha(n), hu(n);
    std::vector<int> ho(n), cell(n), off(ncell + 1, 0), idx;
    for (int i = 0; i < n; i++) {
        const orb_keypoint &k = d->keys_un[i];
        hx[i] = k.x; hy[i] = k.y; ha[i] = k.angle; ho[i] = k.octave;
        hu[i] = d->u_right ? d->u_right[i] : -1.0f;
        const int px = (int)roundf((k.x - f->minX) * invW), py = (int)roundf((k.y - f->minY) * invH);
        cell[i] = (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) ? -1 : px * GRID_ROWS + py;
        if (cell[i] >= 0) off[cell[i] + 1]++;
    }
    for (int c = 0; c < ncell; c++) off[c + 1] += off[c];
    idx.resize(off[ncell] ? off[ncell] : 1);
    {
        std::vector<int> cur(off.begin(), off.end() - 1);
        for (int i = 0; i < n; i++) if (cell[i] >= 0) idx[cur[cell[i]]++] = i;
    }
    Arena ar;
    for (int phase = 0; phase < 2; phase++) {
        ar.used = 0;
        float *dx = ar.take<float>(n), *dy = ar.take<float>(n), *da = ar.take<float>(n), *du = ar.take<float>(n);
        int *doct = ar.take<int>(n);
        uint8_t *dd = ar.take<uint8_t>((size_t)n * 32);
        int *doff = ar.take<int>(ncell + 1), *didx = ar.take<int>(idx.size());
        if (phase == 0) {
            e = cudaMalloc(&f->d_static, ar.used + 256);
            if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&f->stream, cudaStreamNonBlocking);
            if (e != cudaSuccess) { set_error("orb_frame_create: %s", cudaGetErrorString(e)); if (f->d_static) cudaFree(f->d_static); delete f; return ORB_ERR_CUDA; }
            ar.base = (uint8_t *)f->d_static;
            continue;
        }
        cudaStream_t st = f->stream;
        int rc = upload(dx, hx.data(), n, st);
        if (rc == ORB_OK) rc = upload(dy, hy.data(), n, st);
        if (rc == ORB_OK) rc = upload(da, ha.data(), n, st);
        if (rc == ORB_OK) rc = upload(du, hu.data(), n, st);
        if (rc == ORB_OK) rc = upload(doct, ho.data(), n, st);
        if (rc == ORB_OK) rc = upload(dd, d->descriptors, (size_t)n * 32, st);
        if (rc == ORB_OK) rc = upload(doff, off.data(), ncell + 1, st);
        if (rc == ORB_OK) rc = upload(didx, idx.data(), idx.size(), st);
        if (rc == ORB_OK && cudaStreamSynchronize(st) != cudaSuccess) rc = ORB_ERR_CUDA;
        if (rc != ORB_OK) { cudaFree(f->d_static); cudaStreamDestroy(f->stream); delete f; return rc; }
        const bool kfv = d->keyframe_view != 0;   // the grid above was filled with the Frame's float bounds either way
        f->dev = FrameDev{n, dx, dy, da, du, doct, dd, doff, didx, kfv ? (float)(int)f->minX : f->minX, kfv ? (float)(int)f->minY : f->minY, invW, invH};
    }
    *out = f;
    return ORB_OK;
}

void orb_frame_destroy(orb_frame *f) {
    if (!f) return;
    cudaSetDevice(f->device);
    cudaStreamSynchronize(f->stream);
    cudaFree(f->d_static);
    if (f->d_scratch) cudaFree(f->d_scratch);
    cudaStreamDestroy(f->stream);
    delete f;
}

int64_t orb_frame_launch_count(const orb_frame *f) { return f ? f->launches : 0; }

}  // extern "C"

namespace {

// Host description of a query set (all pointers host); uploaded into the target frame's scratch.
struct HostQueries {
    int n = 0;
    const uint8_t *valid = nullptr;
    const float *x = nullptr, *y = nullptr, *radius = nullptr;
    const int32_t *min_level = nullptr, *max_level = nullptr;
    int strict_levels = 0;
    const float *stereo_ref = nullptr, *stereo_tol = nullptr;
    const uint8_t *desc = nullptr;
    const float *angle = nullptr;
    const uint8_t *claims = nullptr;
    const float *gate_ur = nullptr, *gate_inv_sigma2 = nullptr;   // optional reprojection gate (n and nlevels entries)
    int gate_levels = 0;
    double gate_mono = 0, gate_stereo = 0;
};

// Phase 1 of every windowed search: upload queries, count candidates, scan, fill (idx, dist) lists.
// Leaves everything in f->d_scratch; `extra_bytes` are reserved after the lists for the caller's state.
int build_window_lists(orb_frame *f, const HostQueries &H, size_t extra_bytes, WinQ *Qd, WindowLists *WL, uint8_t **extra) {
    const int n = H.n;
    cudaStream_t st = f->stream;
    // pass A: queries + counts + offsets
    Arena a;
    auto place = [&](Arena &ar) {
        ar.used = 0;
        WinQ Q;
        Q.n = n;
        Q.valid = H.valid ? ar.take<uint8_t>(n) : nullptr;
        Q.x = ar.take<float>(n); Q.y = ar.take<float>(n); Q.radius = ar.take<float>(n);
        Q.min_level = ar.take<int>(n); Q.max_level = ar.take<int>(n);
        Q.strict_levels = H.strict_levels;
        Q.stereo_ref = H.stereo_ref ? ar.take<float>(n) : nullptr;
        Q.stereo_tol = H.stereo_ref ? ar.take<float>(n) : nullptr;
        Q.desc = ar.take<uint8_t>((size_t)n * 32);
        Q.angle = H.angle ? ar.take<float>(n) : nullptr;
        Q.claims = H.claims ? ar.take<uint8_t>(n) : nullptr;
        Q.gate_ur = H.gate_inv_sigma2 ? ar.take<float>(n) : nullptr;
        Q.gate_inv_sigma2 = H.gate_inv_sigma2 ? ar.take<float>(H.gate_levels) : nullptr;
        Q.gate_mono = H.gate_mono; Q.gate_stereo = H.gate_stereo;
        WL->off = ar.take<int>(n + 1);
        return Q;
    };
    place(a);
    const size_t fixed = a.used + (size_t)(n + 1) * 4 + 256;
    // counts live temporarily where the candidate lists will go
    ORB_TRY(ensure_scratch(f, fixed + (size_t)n * 4 + extra_bytes + 1024));
    a.base = (uint8_t *)f->d_scratch;
    WinQ Q = place(a);
    const size_t lists_pos = a.used;
    int *counts = a.take<int>(n);
    if (H.valid) ORB_TRY(upload((uint8_t *)Q.valid, H.valid, n, st));
    ORB_TRY(upload((float *)Q.x, H.x, n, st)); ORB_TRY(upload((float *)Q.y, H.y, n, st)); ORB_TRY(upload((float *)Q.radius, H.radius, n, st));
    ORB_TRY(upload((int *)Q.min_level, H.min_level, n, st)); ORB_TRY(upload((int *)Q.max_level, H.max_level, n, st));
    if (H.stereo_ref) { ORB_TRY(upload((float *)Q.stereo_ref, H.stereo_ref, n, st)); ORB_TRY(upload((float *)Q.stereo_tol, H.stereo_tol, n, st)); }
    ORB_TRY(upload((uint8_t *)Q.desc, H.desc, (size_t)n * 32, st));
    if (H.angle) ORB_TRY(upload((float *)Q.angle, H.angle, n, st));
    if (H.claims) ORB_TRY(upload((uint8_t *)Q.claims, H.claims, n, st));
    if (H.gate_inv_sigma2) {
        ORB_TRY(upload((float *)Q.gate_ur, H.gate_ur, n, st));
        ORB_TRY(upload((float *)Q.gate_inv_sigma2, H.gate_inv_sigma2, H.gate_levels, st));
    }
    const int blocks = (n * 32 + 255) / 256;
    if (n) k_window_count<<<blocks, 256, 0, st>>>(f->dev, Q, counts);
    k_scan_counts<<<1, 1024, 0, st>>>(counts, n, WL->off);
    f->launches += 2;
    int total = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(&total, WL->off + n, 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    // pass B: lists.  Growing the scratch would move the queries, so re-run placement if it grows.
    const size_t need = fixed + (size_t)total * 8 + 512 + extra_bytes + 1024;
    if (f->scratch_bytes < need) {
        // preserve nothing: simply rebuild (rare: only when a frame sees its largest query set)
        ORB_TRY(ensure_scratch(f, need));
        return build_window_lists(f, H, extra_bytes, Qd, WL, extra);
    }
    Arena b;
    b.base = (uint8_t *)f->d_scratch;
    b.used = lists_pos;   // the candidate lists reuse the space of `counts`
    WL->idx = b.take<int>(total);
    WL->dist = b.take<int>(total);
    WL->total = total;
    *extra = b.base + b.used;
    if (n && total) { k_window_fill<<<blocks, 256, 0, st>>>(f->dev, Q, WL->off, WL->idx, WL->dist); f->launches++; }
    ORB_CUDA_TRY(cudaGetLastError());
    *Qd = Q;
    return ORB_OK;
}

}  // namespace

extern "C" {

int orb_frame_features_in_area(orb_frame *f, float x, float y, float r, int min_level, int max_level, int32_t *indices,
                               int capacity, int *n_out) {
    ORB_REQUIRE(f && indices && n_out, ORB_ERR_INVALID, "null argument");
    ORB_CUDA_TRY(cudaSetDevice(f->device));
    HostQueries H;
    uint8_t zero_desc[32] = {0};
    H.n = 1; H.x = &x; H.y = &y; H.radius = &r; H.min_level = &min_level; H.max_level = &max_level; H.desc = zero_desc;
    WinQ Q; WindowLists WL; uint8_t *extra;
    ORB_TRY(build_window_lists(f, H, 0, &Q, &WL, &extra));
    *n_out = WL.total;
    ORB_REQUIRE(WL.total <= capacity, ORB_ERR_CAPACITY, "index buffer too small");
    if (WL.total) {
        ORB_CUDA_TRY(cudaMemcpyAsync(indices, WL.idx, (size_t)WL.total * 4, cudaMemcpyDeviceToHost, f->stream));
        ORB_CUDA_TRY(cudaStreamSynchronize(f->stream));
    }
    return ORB_OK;
}

static int search_windowed_host(orb_frame *f, const HostQueries &H, const uint8_t *claimed, int select, int th_dist, float nnratio,
                                int check_ori, int32_t *assigned, int *nmatches) {
    ORB_CUDA_TRY(cudaSetDevice(f->device));
    const int n = H.n, nf = f->n;
    if (nmatches) *nmatches = 0;
    for (int i = 0; i < nf; i++) assigned[i] = -1;
    if (n == 0 || nf == 0) return ORB_OK;
    const size_t extra_bytes = (size_t)nf * 4 + nf + (size_t)n * 8 + (size_t)nf * 4 + 4 + 6 * 256;
    WinQ Q; WindowLists WL; uint8_t *extra;
    ORB_TRY(build_window_lists(f, H, extra_bytes, &Q, &WL, &extra));
    Arena ar; ar.base = extra;
    int *holder = ar.take<int>(nf);
    uint8_t *blocking = ar.take<uint8_t>(nf);
    int *acc_feat = ar.take<int>(n), *acc_bin = ar.take<int>(n), *d_assigned = ar.take<int>(nf), *d_nm = ar.take<int>(1);
    cudaStream_t st = f->stream;
    std::vector<int> h_holder(nf, -1);
    std::vector<uint8_t> h_block(nf, 0);
    if (claimed) for (int i = 0; i < nf; i++) if (claimed[i]) { h_holder[i] = -2; h_block[i] = 1; }
    ORB_TRY(upload(holder, h_holder.data(), nf, st));
    ORB_TRY(upload(blocking, h_block.data(), nf, st));
    k_resolve_windowed<<<1, 32, 0, st>>>(f->dev, Q, WL.off, WL.idx, WL.dist, holder, blocking, select, th_dist, nnratio, check_ori,
                                         acc_feat, acc_bin, d_assigned, d_nm);
    f->launches++;
    ORB_CUDA_TRY(cudaGetLastError());
    int nm = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(assigned, d_assigned, (size_t)nf * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(&nm, d_nm, 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    if (nmatches) *nmatches = nm;
    return ORB_OK;
}

int orb_search_windowed(orb_frame *f, const orb_window_queries *q, const uint8_t *claimed, int select, int th_dist, float nnratio,
                        int check_orientation, int32_t *assigned, int *nmatches) {
    ORB_REQUIRE(f && q && assigned, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(q->n >= 0 && (q->n == 0 || (q->x && q->y && q->radius && q->min_level && q->max_level && q->descriptors)),
                ORB_ERR_INVALID, "incomplete query set");
    ORB_REQUIRE((q->stereo_ref == nullptr) == (q->stereo_tol == nullptr), ORB_ERR_INVALID, "stereo_ref and stereo_tol go together");
    ORB_REQUIRE(!check_orientation || q->angle, ORB_ERR_INVALID, "check_orientation needs query angles");
    ORB_REQUIRE(select == ORB_SELECT_BEST || select == ORB_SELECT_RATIO_LEVEL, ORB_ERR_INVALID, "bad select mode");
    HostQueries H;
    H.n = q->n; H.valid = q->valid; H.x = q->x; H.y = q->y; H.radius = q->radius; H.min_level = q->min_level; H.max_level = q->max_level;
    H.strict_levels = q->strict_levels; H.stereo_ref = q->stereo_ref; H.stereo_tol = q->stereo_tol; H.desc = q->descriptors;
    H.angle = q->angle; H.claims = q->claims;
    return search_windowed_host(f, H, claimed, select, th_dist, nnratio, check_orientation, assigned, nmatches);
}

int orb_search_by_projection_points(orb_frame *f, const orb_track_points *P, const uint8_t *claimed, float th, float nnratio,
                                    int32_t *assigned, int *nmatches) {
    ORB_REQUIRE(f && P && assigned, ORB_ERR_INVALID, "null argument");
    const int n = P->n;
    ORB_REQUIRE(n >= 0 && (n == 0 || (P->in_view && P->proj_x && P->proj_y && P->proj_xr && P->level && P->view_cos && P->descriptors)),
                ORB_ERR_INVALID, "incomplete map point set");
    std::vector<float> radius(n);
    std::vector<int32_t> minl(n), maxl(n);
    const bool bFactor = th != 1.0;                              // ref :49
    for (int i = 0; i < n; i++) {
        const int lvl = P->level[i];
        float r = P->view_cos[i] > 0.998 ? 2.5f : 4.0f;          // RadiusByViewingCos, ref :131-137
        if (bFactor) r *= th;                                    // ref :65-66
        const int lv = std::min(std::max(lvl, 0), f->nlevels - 1);
        radius[i] = r * f->scale[lv];                            // ref :69, :94
        minl[i] = lvl - 1; maxl[i] = lvl;
    }
    HostQueries H;
    H.n = n; H.valid = P->in_view; H.x = P->proj_x; H.y = P->proj_y; H.radius = radius.data(); H.min_level = minl.data();
    H.max_level = maxl.data(); H.stereo_ref = P->proj_xr; H.stereo_tol = radius.data(); H.desc = P->descriptors; H.claims = P->obs_positive;
    return search_windowed_host(f, H, claimed, ORB_SELECT_RATIO_LEVEL, TH_HIGH_M, nnratio, 0, assigned, nmatches);
}

int orb_search_by_projection_last(orb_frame *cur, const orb_last_points *P, const uint8_t *claimed, float th, int forward,
                                  int backward, float mbf, int check_orientation, int32_t *assigned, int *nmatches) {
    ORB_REQUIRE(cur && P && assigned, ORB_ERR_INVALID, "null argument");
    const int n = P->n;
    ORB_REQUIRE(n >= 0 && (n == 0 || (P->valid && P->u && P->v && P->invz && P->octave && P->descriptors)), ORB_ERR_INVALID,
                "incomplete point set");
    ORB_REQUIRE(!check_orientation || P->angle, ORB_ERR_INVALID, "check_orientation needs angles");
    std::vector<float> radius(n), sref(n);
    std::vector<int32_t> minl(n), maxl(n);
    for (int i = 0; i < n; i++) {
        const int oct = std::min(std::max(P->octave[i], 0), cur->nlevels - 1);
        radius[i] = th * cur->scale[oct];                                   // ref :1384
        if (forward) { minl[i] = P->octave[i]; maxl[i] = -1; }              // ref :1388-1393
        else if (backward) { minl[i] = 0; maxl[i] = P->octave[i]; }
        else { minl[i] = P->octave[i] - 1; maxl[i] = P->octave[i] + 1; }
        sref[i] = P->u[i] - mbf * P->invz[i];                               // ref :1412
    }
    HostQueries H;
    H.n = n; H.valid = P->valid; H.x = P->u; H.y = P->v; H.radius = radius.data(); H.min_level = minl.data(); H.max_level = maxl.data();
    H.stereo_ref = sref.data(); H.stereo_tol = radius.data(); H.desc = P->descriptors; H.angle = P->angle; H.claims = P->obs_positive;
    return search_windowed_host(cur, H, claimed, ORB_SELECT_BEST, TH_HIGH_M, 0.f, check_orientation, assigned, nmatches);
}

int orb_search_for_initialization(orb_frame *f1, orb_frame *f2, float *prev_matched, int32_t *matches12, int window_size,
                                  float nnratio, int check_orientation, int *nmatches) {
    ORB_REQUIRE(f1 && f2 && prev_matched && matches12, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(f1->device == f2->device, ORB_ERR_INVALID, "frames on different devices");
    ORB_CUDA_TRY(cudaSetDevice(f2->device));
    const int n1 = f1->n, n2 = f2->n;
    if (nmatches) *nmatches = 0;
    for (int i = 0; i < n1; i++) matches12[i] = -1;
    if (n1 == 0 || n2 == 0) return ORB_OK;
    std::vector<uint8_t> valid(n1);
    std::vector<float> qx(n1), qy(n1), radius(n1, (float)window_size);
    std::vector<int32_t> lv(n1);
    std::vector<uint8_t> desc((size_t)n1 * 32);
    for (int i = 0; i < n1; i++) {
        valid[i] = f1->keys[i].octave > 0 ? 0 : 1;      // ref :421-423
        qx[i] = prev_matched[2 * i]; qy[i] = prev_matched[2 * i + 1];
        lv[i] = f1->keys[i].octave;                      // GetFeaturesInArea(..., level1, level1), ref :425
    }
    // query descriptors are F1's own rows: copy device-to-device after placement
    HostQueries H;
    H.n = n1; H.valid = valid.data(); H.x = qx.data(); H.y = qy.data(); H.radius = radius.data(); H.min_level = lv.data();
    H.max_level = lv.data(); H.desc = desc.data();
    cudaStream_t st = f2->stream;
    ORB_CUDA_TRY(cudaMemcpyAsync(desc.data(), f1->dev.desc, (size_t)n1 * 32, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    const size_t extra_bytes = (size_t)n2 * 8 + (size_t)n1 * 8 + (size_t)n1 * 8 + 4 + 6 * 256;
    WinQ Q; WindowLists WL; uint8_t *extra;
    ORB_TRY(build_window_lists(f2, H, extra_bytes, &Q, &WL, &extra));
    Arena ar; ar.base = extra;
    int *matched_dist = ar.take<int>(n2), *m21 = ar.take<int>(n2), *m12 = ar.take<int>(n1), *qbin = ar.take<int>(n1);
    float *d_prev = ar.take<float>((size_t)n1 * 2);
    int *d_nm = ar.take<int>(1);
    std::vector<int> init2(n2, 0x7fffffff);
    ORB_TRY(upload(matched_dist, init2.data(), n2, st));
    ORB_CUDA_TRY(cudaMemsetAsync(m21, 0xff, (size_t)n2 * 4, st));
    ORB_CUDA_TRY(cudaMemsetAsync(m12, 0xff, (size_t)n1 * 4, st));
    ORB_CUDA_TRY(cudaMemsetAsync(qbin, 0xff, (size_t)n1 * 4, st));
    ORB_TRY(upload(d_prev, prev_matched, (size_t)n1 * 2, st));
    k_resolve_init<<<1, 32, 0, st>>>(f1->dev, f2->dev, WL.off, WL.idx, WL.dist, matched_dist, m21, m12, qbin, nnratio,
                                     check_orientation, d_prev, d_nm);
    f2->launches++;
    ORB_CUDA_TRY(cudaGetLastError());
    int nm = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(matches12, m12, (size_t)n1 * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(prev_matched, d_prev, (size_t)n1 * 8, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(&nm, d_nm, 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    if (nmatches) *nmatches = nm;
    return ORB_OK;
}

int orb_search_by_bow(orb_frame *kf, orb_frame *f, int ngroups, const int32_t *kf_off, const int32_t *kf_idx, const int32_t *f_off,
                      const int32_t *f_idx, const uint8_t *kf_valid, const uint8_t *f_valid, int strict_low, float nnratio,
                      int check_orientation, int32_t *matches, int *nmatches) {
    ORB_REQUIRE(kf && f && matches && ngroups >= 0 && (ngroups == 0 || (kf_off && kf_idx && f_off && f_idx && kf_valid)), ORB_ERR_INVALID,
                "bad argument");
    ORB_REQUIRE(kf->device == f->device, ORB_ERR_INVALID, "frames on different devices");
    ORB_CUDA_TRY(cudaSetDevice(f->device));
    if (nmatches) *nmatches = 0;
    for (int i = 0; i < f->n; i++) matches[i] = -1;
    if (ngroups == 0 || f->n == 0 || kf->n == 0) return ORB_OK;
    const int nent = kf_off[ngroups], nfent = f_off[ngroups];
    std::vector<int> ent_fo(nent), ent_fn(nent), ent_po(nent);
    size_t npairs = 0;
    for (int g = 0; g < ngroups; g++)
        for (int a = kf_off[g]; a < kf_off[g + 1]; a++) {
            ORB_REQUIRE(kf_idx[a] >= 0 && kf_idx[a] < kf->n, ORB_ERR_INVALID, "KF feature index out of range");
            ent_fo[a] = f_off[g]; ent_fn[a] = f_off[g + 1] - f_off[g]; ent_po[a] = (int)npairs;
            npairs += ent_fn[a];
            ORB_REQUIRE(ent_fn[a] < 65536, ORB_ERR_INVALID, "vocabulary node with more than 65535 features");
        }
    for (int b = 0; b < nfent; b++) ORB_REQUIRE(f_idx[b] >= 0 && f_idx[b] < f->n, ORB_ERR_INVALID, "F feature index out of range");
    Arena ar;
    int *d_kfidx = nullptr, *d_fo = nullptr, *d_fn = nullptr, *d_po = nullptr, *d_fidx = nullptr, *d_dist = nullptr, *d_accf = nullptr,
        *d_accb = nullptr, *d_matches = nullptr, *d_nm = nullptr;
    uint8_t *d_kv = nullptr, *d_fv = nullptr;
    for (int phase = 0; phase < 2; phase++) {
        ar.used = 0;
        d_kfidx = ar.take<int>(nent); d_fo = ar.take<int>(nent); d_fn = ar.take<int>(nent); d_po = ar.take<int>(nent);
        d_fidx = ar.take<int>(nfent); d_dist = ar.take<int>(npairs); d_accf = ar.take<int>(nent); d_accb = ar.take<int>(nent);
        d_matches = ar.take<int>(f->n); d_nm = ar.take<int>(1);
        d_kv = ar.take<uint8_t>(kf->n); d_fv = ar.take<uint8_t>(f->n);
        if (phase == 0) { ORB_TRY(ensure_scratch(f, ar.used + 1024)); ar.base = (uint8_t *)f->d_scratch; }
    }
    cudaStream_t st = f->stream;
    ORB_TRY(upload(d_kfidx, kf_idx, nent, st)); ORB_TRY(upload(d_fo, ent_fo.data(), nent, st)); ORB_TRY(upload(d_fn, ent_fn.data(), nent, st));
    ORB_TRY(upload(d_po, ent_po.data(), nent, st)); ORB_TRY(upload(d_fidx, f_idx, nfent, st)); ORB_TRY(upload(d_kv, kf_valid, kf->n, st));
    if (f_valid) ORB_TRY(upload(d_fv, f_valid, f->n, st));
    ORB_CUDA_TRY(cudaMemsetAsync(d_matches, 0xff, (size_t)f->n * 4, st));
    if (nent) {
        k_bow_dist<<<(nent * 32 + 255) / 256, 256, 0, st>>>(kf->dev, f->dev, nent, d_kfidx, d_fo, d_fn, d_po, d_fidx, d_kv, d_dist);
        k_resolve_bow<<<1, 32, 0, st>>>(kf->dev, f->dev, nent, d_kfidx, d_fo, d_fn, d_po, d_fidx, d_kv, f_valid ? d_fv : nullptr, d_dist,
                                        strict_low, nnratio, check_orientation, d_accf, d_accb, d_matches, d_nm);
        f->launches += 2;
    }
    ORB_CUDA_TRY(cudaGetLastError());
    int nm = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(matches, d_matches, (size_t)f->n * 4, cudaMemcpyDeviceToHost, st));
    if (nent) ORB_CUDA_TRY(cudaMemcpyAsync(&nm, d_nm, 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    if (nmatches) *nmatches = nm;
    return ORB_OK;
}

int orb_search_window_best(orb_frame *f, const orb_window_queries *q, const orb_reproj_gate *gate, int32_t *best_idx,
                           int32_t *best_dist) {
    ORB_REQUIRE(f && q && best_idx && best_dist, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(q->n >= 0 && (q->n == 0 || (q->x && q->y && q->radius && q->min_level && q->max_level && q->descriptors)),
                ORB_ERR_INVALID, "incomplete query set");
    ORB_REQUIRE((q->stereo_ref == nullptr) == (q->stereo_tol == nullptr), ORB_ERR_INVALID, "stereo_ref and stereo_tol go together");
    ORB_REQUIRE(!gate || (gate->ur && gate->inv_level_sigma2), ORB_ERR_INVALID, "incomplete reprojection gate");
    ORB_CUDA_TRY(cudaSetDevice(f->device));
    const int n = q->n;
    for (int i = 0; i < n; i++) { best_idx[i] = -1; best_dist[i] = 0x7fffffff; }
    if (n == 0 || f->n == 0) return ORB_OK;
    HostQueries H;
    H.n = n; H.valid = q->valid; H.x = q->x; H.y = q->y; H.radius = q->radius; H.min_level = q->min_level; H.max_level = q->max_level;
    H.strict_levels = q->strict_levels; H.stereo_ref = q->stereo_ref; H.stereo_tol = q->stereo_tol; H.desc = q->descriptors;
    if (gate) {
        H.gate_ur = gate->ur; H.gate_inv_sigma2 = gate->inv_level_sigma2; H.gate_levels = f->nlevels;
        H.gate_mono = gate->chi2_mono; H.gate_stereo = gate->chi2_stereo;
    }
    WinQ Q; WindowLists WL; uint8_t *extra;
    ORB_TRY(build_window_lists(f, H, (size_t)n * 8 + 512, &Q, &WL, &extra));
    Arena ar; ar.base = extra;
    int *d_idx = ar.take<int>(n), *d_dist = ar.take<int>(n);
    cudaStream_t st = f->stream;
    k_window_best<<<(n * 32 + 255) / 256, 256, 0, st>>>(n, WL.off, WL.idx, WL.dist, d_idx, d_dist);
    f->launches++;
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(best_idx, d_idx, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(best_dist, d_dist, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

int orb_search_for_triangulation(orb_frame *kf1, orb_frame *kf2, int ngroups, const int32_t *off1, const int32_t *idx1,
                                 const int32_t *off2, const int32_t *idx2, const uint8_t *usable1, const uint8_t *usable2,
                                 const float *F12, float ex, float ey, const float *level_sigma2, int only_stereo,
                                 int check_orientation, int32_t *matches12, int *nmatches) {
    ORB_REQUIRE(kf1 && kf2 && matches12 && F12 && level_sigma2 && ngroups >= 0 &&
                    (ngroups == 0 || (off1 && idx1 && off2 && idx2 && usable1 && usable2)),
                ORB_ERR_INVALID, "bad argument");
    ORB_REQUIRE(kf1->device == kf2->device, ORB_ERR_INVALID, "frames on different devices");
    ORB_CUDA_TRY(cudaSetDevice(kf1->device));
    if (nmatches) *nmatches = 0;
    for (int i = 0; i < kf1->n; i++) matches12[i] = -1;
    if (ngroups == 0 || kf1->n == 0 || kf2->n == 0) return ORB_OK;
    const int nent = off1[ngroups], nent2 = off2[ngroups];
    if (nent == 0 || nent2 == 0) return ORB_OK;
    std::vector<int> ent_o2(nent), ent_n2(nent);
    for (int g = 0; g < ngroups; g++) {
        ORB_REQUIRE(off2[g + 1] - off2[g] < 65536, ORB_ERR_INVALID, "vocabulary node with more than 65535 features");
        for (int a = off1[g]; a < off1[g + 1]; a++) {
            ORB_REQUIRE(idx1[a] >= 0 && idx1[a] < kf1->n, ORB_ERR_INVALID, "KF1 feature index out of range");
            ent_o2[a] = off2[g]; ent_n2[a] = off2[g + 1] - off2[g];
        }
    }
    for (int b = 0; b < nent2; b++) ORB_REQUIRE(idx2[b] >= 0 && idx2[b] < kf2->n, ORB_ERR_INVALID, "KF2 feature index out of range");
    orb_frame *f = kf1;   // scratch and stream of the first key frame
    Arena ar;
    int *d_idx1 = nullptr, *d_o2 = nullptr, *d_n2 = nullptr, *d_idx2 = nullptr, *d_m12 = nullptr, *d_nm = nullptr;
    float *d_sigma = nullptr, *d_scale = nullptr;
    uint8_t *d_u1 = nullptr, *d_u2 = nullptr;
    for (int phase = 0; phase < 2; phase++) {
        ar.used = 0;
        d_idx1 = ar.take<int>(nent); d_o2 = ar.take<int>(nent); d_n2 = ar.take<int>(nent); d_idx2 = ar.take<int>(nent2);
        d_m12 = ar.take<int>(kf1->n); d_nm = ar.take<int>(1);
        d_sigma = ar.take<float>(kf2->nlevels); d_scale = ar.take<float>(kf2->nlevels);
        d_u1 = ar.take<uint8_t>(kf1->n); d_u2 = ar.take<uint8_t>(kf2->n);
        if (phase == 0) { ORB_TRY(ensure_scratch(f, ar.used + 1024)); ar.base = (uint8_t *)f->d_scratch; }
    }
    cudaStream_t st = f->stream;
    ORB_TRY(upload(d_idx1, idx1, nent, st)); ORB_TRY(upload(d_o2, ent_o2.data(), nent, st)); ORB_TRY(upload(d_n2, ent_n2.data(), nent, st));
    ORB_TRY(upload(d_idx2, idx2, nent2, st)); ORB_TRY(upload(d_u1, usable1, kf1->n, st)); ORB_TRY(upload(d_u2, usable2, kf2->n, st));
    ORB_TRY(upload(d_sigma, level_sigma2, kf2->nlevels, st)); ORB_TRY(upload(d_scale, kf2->scale.data(), kf2->nlevels, st));
    ORB_CUDA_TRY(cudaMemsetAsync(d_m12, 0xff, (size_t)kf1->n * 4, st));
    TriArgs A;
    for (int i = 0; i < 9; i++) A.F[i] = F12[i];
    A.ex = ex; A.ey = ey; A.sigma2 = d_sigma; A.scale2 = d_scale; A.only_stereo = only_stereo;
    k_triangulate<<<(nent * 32 + 255) / 256, 256, 0, st>>>(kf1->dev, kf2->dev, nent, d_idx1, d_o2, d_n2, d_idx2, d_u1, d_u2, A, d_m12);
    k_tri_filter<<<1, 1024, 0, st>>>(kf1->dev, kf2->dev, d_m12, check_orientation, d_nm);
    f->launches += 2;
    ORB_CUDA_TRY(cudaGetLastError());
    int nm = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(matches12, d_m12, (size_t)kf1->n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(&nm, d_nm, 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    if (nmatches) *nmatches = nm;
    return ORB_OK;
}

int orb_rot_hist_filter(const float *angle_a, const float *angle_b, int n, uint8_t *keep) {
    ORB_REQUIRE(n >= 0 && (n == 0 || (angle_a && angle_b && keep)), ORB_ERR_INVALID, "bad argument");
    if (n == 0) return ORB_OK;
    float *da = nullptr;
    ORB_CUDA_TRY(cudaMalloc(&da, (size_t)n * 9 + 64));
    float *db = da + n;
    uint8_t *dk = (uint8_t *)(db + n);
    cudaError_t e = cudaMemcpy(da, angle_a, (size_t)n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(db, angle_b, (size_t)n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        k_rot_hist_filter<<<1, 1024>>>(da, db, n, dk);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpy(keep, dk, n, cudaMemcpyDeviceToHost);
    cudaFree(da);
    ORB_CUDA_TRY(e);
    return ORB_OK;
}

}  // extern "C"
