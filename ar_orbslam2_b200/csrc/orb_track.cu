// liborbgpu -- frame feature index and the tracking-time Hamming matchers for sm_100a.
//
// Replaces the hot loops of ORB_SLAM2::ORBmatcher (ref: ORB_SLAM2/src/ORBmatcher.cc) and the parts of
// Frame they read (ref: ORB_SLAM2/src/Frame.cc:235-250, 332-397):
//   orb_frame_*                      mvKeysUn / mDescriptors / mvuRight + the 64x48 grid on the device
//   orb_frame_from_extractor         the same index built ON the device from an extractor's resident results
//   orb_search_windowed              the projection-matcher loop (SearchByProjection x4, Fuse, SearchBySim3)
//   orb_search_for_initialization    ORBmatcher.cc:405-523
//   orb_search_by_bow                ORBmatcher.cc:159-288, 525-658
//   orb_rot_hist_filter              ORBmatcher.cc:1604-1645 + the 30-bin rotation histogram
//
// Structure of one call (latency is what the Tracking thread waits for, so a call is ONE host->device copy, two or
// three kernels, ONE device->host copy and one synchronisation; no allocation, no stream creation):
//   1. the caller's flat arrays are packed into the frame's pinned block and go up as one copy (the same block zeroes
//      the candidate counter);
//   2. k_window_lists: one warp per query walks the query's grid cells in the reference's traversal order, reserves
//      a range of the candidate pool with one atomic and writes (feature, distance) pairs in that order;
//   3. k_resolve_*: the reference's loop is sequential -- a query skips features an EARLIER query took.  That is a
//      triangular dependency, so it has exactly one fixed point, and a Jacobi iteration "every query picks as if the
//      picks of the previous sweep were final" reaches it: after sweep k the first k queries are final.  In practice
//      two or three sweeps (conflicts are rare); every sweep is one thread per query over its short list;
//   4. rotation histogram / three maxima / bookkeeping in the same kernel, results into one block, one copy down.
#include "orb_common.cuh"
#include "orb_internal.cuh"

#include <cooperative_groups.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

namespace cg = cooperative_groups;

namespace orbgpu {

constexpr int GRID_COLS = 64, GRID_ROWS = 48;   // ref Frame.h:39-40
constexpr int NCELL = GRID_COLS * GRID_ROWS;
constexpr int TH_HIGH_M = 100, TH_LOW_M = 50, HISTO = 30;
constexpr int RES_THREADS = 1024;
constexpr int FREE_Q = 0x7fffffff;

struct FrameDev {
    int n;
    const float *x, *y, *angle, *uRight;
    const int *octave;
    const uint8_t *desc;
    const int *cell_off;   // [GRID_COLS*GRID_ROWS + 1], cell = ix*GRID_ROWS + iy
    const int *cell_idx;   // keypoint indices, ascending inside a cell (= insertion order)
    const int4 *rec;       // one record per grid entry, in cell_idx order: {x, y, uRight (float bits), index | octave << 24}
    float minX, minY, invW, invH;   // minX/minY as GetFeaturesInArea sees them (truncated to int for a KeyFrame view)
};

struct WinQ {   // device view of a query set
    int n;
    const uint8_t *valid;          // NULL = all
    const int *valid_if_zero;      // NULL, or query q is valid only where this is 0 (SearchForInitialization: octave 0 only)
    const float *x, *y;
    int xy_stride;                 // 1: separate arrays; 2: x[2q], y[2q] interleaved (vbPrevMatched)
    const float *radius;           // NULL = radius_const
    float radius_const;
    const int *min_level, *max_level;
    int strict_levels;
    const float *stereo_ref, *stereo_tol;
    const uint8_t *desc;           // NULL = no distances (GetFeaturesInArea only)
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

__device__ __forceinline__ int hamming_rows(const uint8_t *a, const uint8_t *b) {
    const uint4 *pa = reinterpret_cast<const uint4 *>(a), *pb = reinterpret_cast<const uint4 *>(b);
    const uint4 a0 = __ldg(pa), a1 = __ldg(pa + 1), b0 = __ldg(pb), b1 = __ldg(pb + 1);
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
           __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// One query's admission test for a feature of the window's cells (Frame::GetFeaturesInArea, ref Frame.cc:332-385, plus the
// in-loop gates of the matchers).
struct WinTest {
    float x, y, r, sref, stol, gur;
    int minL, maxL;
    bool check, strict, stereo, gate;
    const float *inv_sigma2;
    double chi_mono, chi_stereo;
    __device__ __forceinline__ bool ok(float kx, float ky, float ur, int oct) const {
        if (strict) { if (oct < minL || oct > maxL) return false; }
        else if (check) { if (oct < minL || (maxL >= 0 && oct > maxL)) return false; }
        if (!(fabsf(__fsub_rn(kx, x)) < r && fabsf(__fsub_rn(ky, y)) < r)) return false;
        if (stereo) {
            if (ur > 0 && fabsf(__fsub_rn(sref, ur)) > stol) return false;
        }
        if (gate) {   // ref :917-945: squared reprojection error, scaled by the level's inverse sigma^2, against chi^2 (double compare)
            const float ex = __fsub_rn(x, kx), ey = __fsub_rn(y, ky);
            float e2 = __fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey));
            if (ur >= 0) {
                const float er = __fsub_rn(gur, ur);
                e2 = __fadd_rn(e2, __fmul_rn(er, er));
                if ((double)__fmul_rn(e2, inv_sigma2[oct]) > chi_stereo) return false;
            } else if ((double)__fmul_rn(e2, inv_sigma2[oct]) > chi_mono) return false;
        }
        return true;
    }
};

// hdr[0] = candidate slots reserved so far (upper bounds), hdr[1] = overflow flag, hdr[2] = candidates written,
// hdr[3] = queries with a non-empty list, hdr[8..10] = the resolve kernels' "something changed" flags (all zero on entry)
//
// A group of G threads (a warp, or a whole CTA for wide windows) per query.  The grid stores the cells of one grid COLUMN
// back to back, so the window is <= 64 contiguous ranges of grid entries, one per column, and the reference's traversal
// order (ix outer, iy inner, insertion order) is simply "ranges in column order".  The group scans the range lengths, reserves
// their sum with one atomic, and then tests G entries per step -- one 16-byte record per entry, then the descriptor -- with a
// ballot / prefix compaction that keeps the order.  prune >= 0 drops candidates farther than `prune` bits (a matcher that only
// ever takes the best candidate under a threshold never looks at them).
template <int G>
__global__ void __launch_bounds__(256)
k_window_lists(FrameDev f, WinQ Q, int *__restrict__ q_off, int *__restrict__ q_cnt, int2 *__restrict__ lists, int cap, int *hdr, int prune) {
    constexpr int GPB = 256 / G, WPG = G / 32;
    __shared__ int s_pref[GPB][GRID_COLS + 1], s_beg[GPB][GRID_COLS];
    __shared__ int s_wcnt[GPB][WPG + 1];
    __shared__ int s_off[GPB];
    const int g = threadIdx.x / G, t = threadIdx.x % G, lane = threadIdx.x & 31, wg = t >> 5;
    const int q = blockIdx.x * GPB + g;
    auto gsync = [&]() { if (G == 32) __syncwarp(); else __syncthreads(); };
    if (G == 32 && q >= Q.n) return;                 // (a CTA-wide group never diverges on q: blockIdx.x < Q.n)
    bool valid = q < Q.n && (!Q.valid || Q.valid[q]);
    if (valid && Q.valid_if_zero) valid = Q.valid_if_zero[q] == 0;
    int cx0 = 0, cx1 = -1, cy0 = 0, cy1 = -1;
    WinTest T;
    if (valid) {
        const float x = Q.x[(size_t)q * Q.xy_stride], y = Q.y[(size_t)q * Q.xy_stride], r = Q.radius ? Q.radius[q] : Q.radius_const;
        T.x = x; T.y = y; T.r = r;
        cx0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, f.minX), r), f.invW)));
        cx1 = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, f.minX), r), f.invW)));
        cy0 = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, f.minY), r), f.invH)));
        cy1 = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, f.minY), r), f.invH)));
        if (cx0 >= GRID_COLS || cx1 < 0 || cy0 >= GRID_ROWS || cy1 < 0 || cx1 < cx0 || cy1 < cy0) valid = false;
    }
    if (!valid) {                                    // uniform over the group
        if (t == 0 && q < Q.n) { q_off[q] = 0; q_cnt[q] = 0; }
        return;
    }
    T.minL = Q.min_level[q]; T.maxL = Q.max_level[q];
    T.check = (T.minL > 0) || (T.maxL >= 0);
    T.strict = Q.strict_levels != 0;
    T.stereo = Q.stereo_ref != nullptr;
    T.sref = T.stereo ? Q.stereo_ref[q] : 0.f; T.stol = T.stereo ? Q.stereo_tol[q] : 0.f;
    T.gate = Q.gate_inv_sigma2 != nullptr;
    T.gur = T.gate && Q.gate_ur ? Q.gate_ur[q] : 0.f;
    T.inv_sigma2 = Q.gate_inv_sigma2; T.chi_mono = Q.gate_mono; T.chi_stereo = Q.gate_stereo;
    const int ny = cy1 - cy0 + 1, ncol = cx1 - cx0 + 1;
    // column ranges and their exclusive prefix sums: the group's first warp, two columns per lane
    if (t < 32) {
        const int c0 = 2 * t, c1 = c0 + 1;
        int b0 = 0, l0 = 0, b1 = 0, l1 = 0;
        if (c0 < ncol) { const int cell = (cx0 + c0) * GRID_ROWS + cy0; b0 = f.cell_off[cell]; l0 = f.cell_off[cell + ny] - b0; }
        if (c1 < ncol) { const int cell = (cx0 + c1) * GRID_ROWS + cy0; b1 = f.cell_off[cell]; l1 = f.cell_off[cell + ny] - b1; }
        int tot;
        const int ex = warp_excl_scan(l0 + l1, t, &tot);
        s_pref[g][c0] = ex; s_beg[g][c0] = b0;
        s_pref[g][c1] = ex + l0; s_beg[g][c1] = b1;
        if (t == 0) {
            s_pref[g][GRID_COLS] = tot;
            int off = 0;
            if (tot) off = atomicAdd(&hdr[0], tot);
            s_off[g] = off;
        }
    }
    gsync();
    const int ub = s_pref[g][GRID_COLS], off = s_off[g];
    if (ub == 0 || off + ub > cap) {
        if (t == 0) { if (ub) hdr[1] = 1; q_off[q] = 0; q_cnt[q] = 0; }
        return;
    }
    uint4 qa = make_uint4(0, 0, 0, 0), qb = qa;
    if (Q.desc) {
        const uint4 *qd = reinterpret_cast<const uint4 *>(Q.desc + (size_t)q * 32);
        qa = __ldg(qd); qb = __ldg(qd + 1);
    }
    int base = off;
    for (int e0 = 0; e0 < ub; e0 += G) {
        const int e = e0 + t;
        bool ok = false;
        int idx = 0, d = 0;
        if (e < ub) {
            int lo = 0, hi = ncol - 1;               // largest column c with pref[c] <= e
#pragma unroll
            for (int it = 0; it < 6; it++) {
                const int mid = (lo + hi + 1) >> 1;
                if (s_pref[g][mid] <= e) lo = mid; else hi = mid - 1;
            }
            const int4 rc = __ldg(f.rec + s_beg[g][lo] + (e - s_pref[g][lo]));
            idx = rc.w & 0xffffff;
            ok = T.ok(__int_as_float(rc.x), __int_as_float(rc.y), __int_as_float(rc.z), (int)((unsigned)rc.w >> 24));
            if (ok && Q.desc) {
                const uint4 *fd = reinterpret_cast<const uint4 *>(f.desc + (size_t)idx * 32);
                const uint4 fa = __ldg(fd), fb = __ldg(fd + 1);
                d = __popc(qa.x ^ fa.x) + __popc(qa.y ^ fa.y) + __popc(qa.z ^ fa.z) + __popc(qa.w ^ fa.w) + __popc(qb.x ^ fb.x) +
                    __popc(qb.y ^ fb.y) + __popc(qb.z ^ fb.z) + __popc(qb.w ^ fb.w);
                if (prune >= 0 && d > prune) ok = false;
            }
        }
        const unsigned bal = __ballot_sync(0xffffffffu, ok);
        int wbase = 0, total = __popc(bal);
        if (G > 32) {
            if (lane == 0) s_wcnt[g][wg] = total;
            __syncthreads();
            total = 0;
#pragma unroll
            for (int w = 0; w < WPG; w++) { if (w == wg) wbase = total; total += s_wcnt[g][w]; }
            __syncthreads();
        }
        if (ok) lists[base + wbase + __popc(bal & ((1u << lane) - 1u))] = make_int2(idx, d);
        base += total;
    }
    if (t == 0) {
        q_off[q] = off; q_cnt[q] = base - off;
        if (base > off) { atomicAdd(&hdr[2], base - off); atomicAdd(&hdr[3], 1); }
    }
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

__device__ __forceinline__ int block_sum(int v, int *s_acc) {   // every thread of the CTA calls it; result valid after the barrier
    __syncthreads();
    if (threadIdx.x == 0) *s_acc = 0;
    __syncthreads();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(s_acc, v);
    __syncthreads();
    return *s_acc;
}

// ---- claim resolution (see the file header for the sweep scheme) ---------------------------------------------------------
// Both kernels are COOPERATIVE launches: a sweep hands every query to a group of L lanes (L = 1..32, chosen from the
// average list length so that long lists -- SearchForInitialization's 100-px windows -- are scanned by several lanes and short
// ones by one), sweeps are separated by grid-wide barriers, and block 0 does the bookkeeping at the end.
// "Who holds feature i" is a 64-bit word (sweep tag << 32 | ~query): picks of sweep s are posted with tag s + 1 by atomicMax,
// so the newest tag wins, within it the smallest query, and nothing has to be cleared between sweeps.
constexpr int RES_BLOCK = 1024;

__device__ __forceinline__ int lanes_per_query(const int *hdr) {
    const int nonempty = max(hdr[3], 1), avg = hdr[2] / nonempty;
    int L = 1;
    while (L < 32 && 4 * L < avg) L <<= 1;   // ~4 candidates per lane
    return L;
}
// top-2 of (distance << 32 | rank) keys over the L lanes of a group (L a power of two, groups aligned inside a warp)
__device__ __forceinline__ void group_top2(unsigned long long &k1, unsigned long long &k2, int L) {
    for (int o = L >> 1; o > 0; o >>= 1) {
        const unsigned long long a1 = __shfl_xor_sync(0xffffffffu, k1, o), a2 = __shfl_xor_sync(0xffffffffu, k2, o);
        const unsigned long long n1 = min(k1, a1), n2 = min(max(k1, a1), min(k2, a2));
        k1 = n1; k2 = n2;
    }
}
constexpr unsigned long long NOKEY = ~0ull;
__device__ __forceinline__ unsigned long long holder_key(int tag, int q) { return ((unsigned long long)tag << 32) | (0xffffffffu - (unsigned)q); }
// query that holds the feature in sweep `tag`, or FREE_Q
__device__ __forceinline__ int holder_of(unsigned long long key, int tag) {   // tags start at 1: sweep 0 sees nobody
    return (tag != 0 && (int)(key >> 32) == tag) ? (int)(0xffffffffu - (unsigned)key) : FREE_Q;
}

// The query loop of ORBmatcher.cc:51-126 / 1354-1448 / 1491-1578 / 312-400.  Query q skips feature i iff it was blocked on
// entry (`claimed`) or a query before q holds it with a blocking map point -- exactly "an earlier query put a map point with
// Observations()>0 there" (ref :87-89, 1406-1408, 1544-1545).
// out[0] = nmatches, out[1] = sweeps, out[2..4] = hdr[0..2]
__global__ void __launch_bounds__(RES_BLOCK)
k_resolve_windowed(FrameDev f, WinQ Q, const int *__restrict__ q_off, const int *__restrict__ q_cnt, const int2 *__restrict__ lists,
                   const uint8_t *__restrict__ claimed, unsigned long long *held, int *res, int *bins, int *holder, uint8_t *bad, int select,
                   int th_dist, float nnratio, int check_ori, int *assigned, int *out, int *hdr) {
    cg::grid_group grid = cg::this_grid();
    const bool one_cta = gridDim.x == 1;   // small problems run as ONE CTA (plain launch): barriers are __syncthreads then
    auto gsync = [&]() { if (one_cta) __syncthreads(); else grid.sync(); };
    __shared__ int hist[HISTO];
    __shared__ int ind[3];
    __shared__ int s_acc;
    const int tid = threadIdx.x, n = Q.n, nf = f.n;
    const int gtid = blockIdx.x * RES_BLOCK + tid, nthreads = gridDim.x * RES_BLOCK;
    for (int i = gtid; i < nf; i += nthreads) held[i] = 0;
    for (int q = gtid; q < n; q += nthreads) res[q] = -1;
    const int L = lanes_per_query(hdr), ngroups = nthreads / L, gid = gtid / L, gl = gtid % L;
    int *changed = hdr + 8;
    gsync();
    int sweep = 0;
    for (;;) {
        if (gtid == 0) changed[(sweep + 1) % 3] = 0;
        for (int q0 = 0; q0 < n; q0 += ngroups) {          // uniform trip count: the shuffles below need whole warps
            const int q = q0 + gid;
            const int cnt = q < n ? q_cnt[q] : 0;
            unsigned long long k1 = NOKEY, k2 = NOKEY;
            const int2 *Lq = lists + (cnt ? q_off[q] : 0);
            for (int j = gl; j < cnt; j += L) {
                const int2 e = Lq[j];
                if (claimed && claimed[e.x]) continue;
                if (holder_of(held[e.x], sweep) < q) continue;
                const unsigned long long key = ((unsigned long long)(unsigned)e.y << 32) | (unsigned)j;
                if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
            }
            group_top2(k1, k2, L);
            if (gl == 0 && q < n) {
                int pick = -1;
                if (k1 != NOKEY) {
                    const int bestDist = (int)(k1 >> 32);
                    if (bestDist <= th_dist && bestDist < 256) {
                        pick = Lq[(unsigned)k1].x;
                        if (select == ORB_SELECT_RATIO_LEVEL) {   // ref :118-121
                            const int bestLevel = f.octave[pick], bestLevel2 = k2 != NOKEY ? f.octave[Lq[(unsigned)k2].x] : -1;
                            const int bestDist2 = k2 != NOKEY ? (int)(k2 >> 32) : 256;
                            if (bestLevel == bestLevel2 && (float)bestDist > __fmul_rn(nnratio, (float)bestDist2)) pick = -1;
                        }
                    }
                }
                if (pick != res[q]) { res[q] = pick; changed[sweep % 3] = 1; }
            }
        }
        gsync();
        const bool again = changed[sweep % 3] != 0;
        sweep++;
        if (!again) break;
        for (int q = gtid; q < n; q += nthreads) {
            const int p = res[q];
            if (p >= 0 && (!Q.claims || Q.claims[q])) atomicMax(&held[p], holder_key(sweep, q));
        }
        gsync();
    }
    if (blockIdx.x != 0) return;
    // ---- bookkeeping (block 0): who ends up in mvpMapPoints[i] (the last taker), rotation histogram over EVERY accepted query
    // (an overwritten assignment still counts, F7), three maxima, reset of the takers outside them (ref :1451-1470)
    if (tid < HISTO) hist[tid] = 0;
    for (int i = tid; i < nf; i += RES_BLOCK) { holder[i] = -1; bad[i] = 0; }
    __syncthreads();
    int acc = 0;
    for (int q = tid; q < n; q += RES_BLOCK) {
        const int p = res[q];
        if (p < 0) continue;
        acc++;
        atomicMax(&holder[p], q);
        if (check_ori) {
            const int b = rot_bin(Q.angle[q], f.angle[p]);
            bins[q] = b;
            atomicAdd(&hist[b], 1);
        }
    }
    int nm = block_sum(acc, &s_acc);
    if (check_ori) {
        if (tid == 0) three_maxima(hist, ind[0], ind[1], ind[2]);
        __syncthreads();
        int dec = 0;
        for (int q = tid; q < n; q += RES_BLOCK) {
            const int p = res[q];
            if (p < 0) continue;
            const int b = bins[q];
            if (b != ind[0] && b != ind[1] && b != ind[2]) { bad[p] = 1; dec++; }
        }
        nm -= block_sum(dec, &s_acc);
    }
    __syncthreads();
    for (int i = tid; i < nf; i += RES_BLOCK) assigned[i] = bad[i] ? -2 : holder[i];
    if (tid == 0) { out[0] = nm; out[1] = sweep; out[2] = hdr[0]; out[3] = hdr[1]; out[4] = hdr[2]; }
}

// ORBmatcher::SearchForInitialization (ref :418-520).  Query i1 ignores a candidate i2 at distance d when an earlier query
// holds i2 at a distance <= d (vMatchedDistance, ref :444-445); a later taker displaces the earlier one (:463-471).  After
// every sweep the takers of a feature are chained (head / next, with a snapshot of their distances) so that the gate of
// (q, i2) = min distance over the takers before q; a sweep only reads the chains and only writes its own res / res_dist.
__global__ void __launch_bounds__(RES_BLOCK)
k_resolve_init(FrameDev f1, FrameDev f2, const int *__restrict__ q_off, const int *__restrict__ q_cnt, const int2 *__restrict__ lists,
               unsigned long long *head, int *next, int *tdist, int *res, int *res_dist, int *last, float nnratio, int check_ori, int *m12,
               const float *prev_in, float *prev_out, int *out, int *hdr) {
    cg::grid_group grid = cg::this_grid();
    const bool one_cta = gridDim.x == 1;   // small problems run as ONE CTA (plain launch): barriers are __syncthreads then
    auto gsync = [&]() { if (one_cta) __syncthreads(); else grid.sync(); };
    __shared__ int hist[HISTO];
    __shared__ int ind[3];
    __shared__ int s_acc;
    const int tid = threadIdx.x, n1 = f1.n, n2 = f2.n;
    const int gtid = blockIdx.x * RES_BLOCK + tid, nthreads = gridDim.x * RES_BLOCK;
    for (int i = gtid; i < n2; i += nthreads) head[i] = 0;
    for (int q = gtid; q < n1; q += nthreads) { res[q] = -1; res_dist[q] = 0; }
    const int L = lanes_per_query(hdr), ngroups = nthreads / L, gid = gtid / L, gl = gtid % L;
    int *changed = hdr + 8;
    gsync();
    int sweep = 0;
    for (;;) {
        if (gtid == 0) changed[(sweep + 1) % 3] = 0;
        for (int q0 = 0; q0 < n1; q0 += ngroups) {
            const int q = q0 + gid;
            const int cnt = q < n1 ? q_cnt[q] : 0;
            unsigned long long k1 = NOKEY, k2 = NOKEY;
            const int2 *Lq = lists + (cnt ? q_off[q] : 0);
            for (int j = gl; j < cnt; j += L) {
                const int2 e = Lq[j];
                int gate = 0x7fffffff;
                for (int t = holder_of(head[e.x], sweep); t != FREE_Q && t >= 0; t = next[t])
                    if (t < q) gate = min(gate, tdist[t]);
                if (gate <= e.y) continue;                          // ref :444-445
                const unsigned long long key = ((unsigned long long)(unsigned)e.y << 32) | (unsigned)j;
                if (key < k1) { k2 = k1; k1 = key; } else if (key < k2) k2 = key;
            }
            group_top2(k1, k2, L);
            if (gl == 0 && q < n1) {
                int pick = -1, pick_d = 0;
                if (k1 != NOKEY) {
                    const int bestDist = (int)(k1 >> 32);
                    const float second = k2 != NOKEY ? (float)(int)(k2 >> 32) : (float)0x7fffffff;
                    if (bestDist <= TH_LOW_M && (float)bestDist < __fmul_rn(second, nnratio)) { pick = Lq[(unsigned)k1].x; pick_d = bestDist; }
                }
                if (pick != res[q] || pick_d != res_dist[q]) { res[q] = pick; res_dist[q] = pick_d; changed[sweep % 3] = 1; }
            }
        }
        gsync();
        const bool again = changed[sweep % 3] != 0;
        sweep++;
        if (!again) break;
        for (int q = gtid; q < n1; q += nthreads) {
            const int p = res[q];
            if (p >= 0) {
                tdist[q] = res_dist[q];
                const unsigned long long old = atomicExch(&head[p], holder_key(sweep, q));
                const int t = holder_of(old, sweep);
                next[q] = t == FREE_Q ? -1 : t;
            }
        }
        gsync();
    }
    if (blockIdx.x != 0) return;
    // ---- vnMatches21 = the last taker; an earlier taker of the same feature was displaced (m12 = -1) but stays in the
    // rotation histogram (ref :463-486); filter (:492-515); vbPrevMatched update (:518-520)
    if (tid < HISTO) hist[tid] = 0;
    for (int i = tid; i < n2; i += RES_BLOCK) last[i] = -1;
    __syncthreads();
    for (int q = tid; q < n1; q += RES_BLOCK) {
        const int p = res[q];
        if (p < 0) continue;
        atomicMax(&last[p], q);
        if (check_ori) atomicAdd(&hist[rot_bin(f1.angle[q], f2.angle[p])], 1);
    }
    __syncthreads();
    if (check_ori && tid == 0) three_maxima(hist, ind[0], ind[1], ind[2]);
    __syncthreads();
    int acc = 0;
    for (int q = tid; q < n1; q += RES_BLOCK) {
        const int p = res[q];
        int m = (p >= 0 && last[p] == q) ? p : -1;
        if (m >= 0 && check_ori) {
            const int b = rot_bin(f1.angle[q], f2.angle[p]);
            if (b != ind[0] && b != ind[1] && b != ind[2]) m = -1;
        }
        m12[q] = m;
        float px = prev_in[2 * q], py = prev_in[2 * q + 1];
        if (m >= 0) { px = f2.x[m]; py = f2.y[m]; acc++; }
        prev_out[2 * q] = px; prev_out[2 * q + 1] = py;
    }
    const int nm = block_sum(acc, &s_acc);
    if (tid == 0) { out[0] = nm; out[1] = sweep; out[2] = hdr[0]; out[3] = hdr[1]; out[4] = hdr[2]; }
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

// ref ORBmatcher.cc:196-262 / 565-636: entries (KF features in FeatureVector order) take F features of the same node;
// a taken F feature is skipped by later entries (:209-210).  Same sweep scheme: taken_by[fi] = smallest entry holding fi.
__global__ void __launch_bounds__(RES_THREADS)
k_resolve_bow(FrameDev kf, FrameDev f, int nentries, const int *__restrict__ kf_idx, const int *__restrict__ ent_fo,
              const int *__restrict__ ent_fn, const int *__restrict__ ent_po, const int *__restrict__ f_idx,
              const uint8_t *__restrict__ kf_valid, const uint8_t *__restrict__ f_valid, const int *__restrict__ dist, int strict_low,
              float nnratio, int check_ori, int *taken_by, int *res, int *matches, int *out) {
    __shared__ int hist[HISTO];
    __shared__ int ind[3];
    __shared__ int s_changed, s_acc;
    const int tid = threadIdx.x, nf = f.n;
    for (int i = tid; i < nf; i += RES_THREADS) taken_by[i] = FREE_Q;
    for (int a = tid; a < nentries; a += RES_THREADS) res[a] = -1;
    if (tid < HISTO) hist[tid] = 0;
    __syncthreads();
    int sweeps = 0;
    for (;;) {
        if (tid == 0) s_changed = 0;
        __syncthreads();
        bool changed = false;
        for (int a = tid; a < nentries; a += RES_THREADS) {
            const int ki = kf_idx[a];
            int pick = -1;
            if (kf_valid[ki]) {
                const int fo = ent_fo[a], fn = ent_fn[a], po = ent_po[a];
                int best1 = 256, best2 = 256, bestIdx = -1;
                for (int j = 0; j < fn; j++) {
                    const int fi = f_idx[fo + j];
                    if (taken_by[fi] < a) continue;                       // ref :209-210
                    if (f_valid && !f_valid[fi]) continue;                // ref :577-583
                    const int d = dist[po + j];
                    if (d < best1) { best2 = best1; best1 = d; bestIdx = fi; }
                    else if (d < best2) best2 = d;
                }
                if (bestIdx >= 0 && (strict_low ? best1 < TH_LOW_M : best1 <= TH_LOW_M) &&
                    (float)best1 < __fmul_rn(nnratio, (float)best2)) pick = bestIdx;   // ref :228-230
            }
            if (pick != res[a]) { res[a] = pick; changed = true; }
        }
        if (changed) s_changed = 1;
        sweeps++;
        __syncthreads();
        if (!s_changed) break;
        for (int i = tid; i < nf; i += RES_THREADS) taken_by[i] = FREE_Q;
        __syncthreads();
        for (int a = tid; a < nentries; a += RES_THREADS)
            if (res[a] >= 0) atomicMin(&taken_by[res[a]], a);
        __syncthreads();
    }
    // at the fixed point every taken feature has exactly one taker
    for (int i = tid; i < nf; i += RES_THREADS) matches[i] = -1;
    __syncthreads();
    int acc = 0;
    for (int a = tid; a < nentries; a += RES_THREADS) {
        const int p = res[a];
        if (p < 0) continue;
        acc++;
        if (check_ori) atomicAdd(&hist[rot_bin(kf.angle[kf_idx[a]], f.angle[p])], 1);
    }
    int nm = block_sum(acc, &s_acc);
    if (check_ori && tid == 0) three_maxima(hist, ind[0], ind[1], ind[2]);
    __syncthreads();
    int dec = 0;
    for (int a = tid; a < nentries; a += RES_THREADS) {
        const int p = res[a];
        if (p < 0) continue;
        const int ki = kf_idx[a];
        if (check_ori) {
            const int b = rot_bin(kf.angle[ki], f.angle[p]);
            if (b != ind[0] && b != ind[1] && b != ind[2]) { dec++; continue; }   // ref :270-281
        }
        matches[p] = ki;
    }
    nm -= block_sum(dec, &s_acc);
    if (tid == 0) { out[0] = nm; out[1] = sweeps; }
}

// Fuse / SearchBySim3 candidate loop (ref ORBmatcher.cc:895-951, 1052-1072, 1191-1213, 1271-1293): independent queries,
// closest descriptor of the window, the first one in traversal order on ties.
__global__ void k_window_best(int nq, const int *__restrict__ q_off, const int *__restrict__ q_cnt, const int2 *__restrict__ lists,
                              int *best_idx, int *best_dist) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const int2 *L = lists + q_off[q];
    const int n = q_cnt[q];
    int bd = 0x7fffffff, bi = -1;
    for (int j = 0; j < n; j++) {
        const int2 e = L[j];
        if (e.y < bd) { bd = e.y; bi = e.x; }
    }
    best_idx[q] = bi;
    best_dist[q] = bd;
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
    uint32_t key = 0xffffffffu;
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
    if (lane == 0 && key != 0xffffffffu) m12[i1] = idx2[o2 + (0xffff - (int)(key & 0xffff))];
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

// Frame::AssignFeaturesToGrid / PosInGrid (ref Frame.cc:235-250, 387-397) on the device, one CTA: key points (AoS, as the
// extractor leaves them) -> SoA + the 64x48 CSR grid.  Cell = round((x - minX) * invW) (round, not floor); indices inside a
// cell ascend (= the reference's insertion order): slots are handed out by atomics, then every cell's short list is sorted.
struct FrameBuild {
    float *x, *y, *angle, *uRight;
    int4 *rec;
    int *octave, *cell_off, *cell_idx, *cell_of;
    float minX, minY, invW, invH;
};
__global__ void __launch_bounds__(1024)
k_frame_build(const orb_keypoint *__restrict__ kp, int n, const float *__restrict__ u_right, FrameBuild B) {
    __shared__ int cnt[NCELL];
    __shared__ int warp_tot[33];
    const int tid = threadIdx.x, lane = tid & 31;
    for (int c = tid; c < NCELL; c += 1024) cnt[c] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += 1024) {
        const orb_keypoint k = kp[i];
        B.x[i] = k.x; B.y[i] = k.y; B.angle[i] = k.angle; B.octave[i] = k.octave;
        B.uRight[i] = u_right ? u_right[i] : -1.0f;
        const int px = (int)roundf(__fmul_rn(__fsub_rn(k.x, B.minX), B.invW)), py = (int)roundf(__fmul_rn(__fsub_rn(k.y, B.minY), B.invH));
        const int cell = (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) ? -1 : px * GRID_ROWS + py;
        B.cell_of[i] = cell;
        if (cell >= 0) atomicAdd(&cnt[cell], 1);
    }
    __syncthreads();
    // exclusive scan of the 3072 counts: 3 consecutive cells per thread
    const int c0 = tid * 3;
    const int a0 = cnt[c0], a1 = cnt[c0 + 1], a2 = cnt[c0 + 2];
    int tot;
    const int ex = warp_excl_scan(a0 + a1 + a2, lane, &tot);
    if (lane == 31) warp_tot[tid >> 5] = tot;
    __syncthreads();
    if (tid < 32) {
        int t;
        const int e = warp_excl_scan(warp_tot[tid], tid, &t);
        warp_tot[tid] = e;
        if (tid == 31) warp_tot[32] = t;
    }
    __syncthreads();
    const int base = warp_tot[tid >> 5] + ex;
    B.cell_off[c0] = base; B.cell_off[c0 + 1] = base + a0; B.cell_off[c0 + 2] = base + a0 + a1;
    if (tid == 1023) B.cell_off[NCELL] = warp_tot[32];
    __syncthreads();
    cnt[c0] = base; cnt[c0 + 1] = base + a0; cnt[c0 + 2] = base + a0 + a1;   // write cursors
    __syncthreads();
    for (int i = tid; i < n; i += 1024) {
        const int cell = B.cell_of[i];
        if (cell >= 0) B.cell_idx[atomicAdd(&cnt[cell], 1)] = i;
    }
    __syncthreads();
    for (int c = tid; c < NCELL; c += 1024) {
        const int beg = B.cell_off[c], end = cnt[c];
        for (int i = beg + 1; i < end; i++) {   // insertion sort, cells hold a handful of key points
            const int v = B.cell_idx[i];
            int j = i - 1;
            while (j >= beg && B.cell_idx[j] > v) { B.cell_idx[j + 1] = B.cell_idx[j]; j--; }
            B.cell_idx[j + 1] = v;
        }
        for (int i = beg; i < end; i++) {
            const int v = B.cell_idx[i];
            B.rec[i] = make_int4(__float_as_int(B.x[v]), __float_as_int(B.y[v]), __float_as_int(B.uRight[v]), v | (B.octave[v] << 24));
        }
    }
}

}  // namespace orbgpu

using namespace orbgpu;

// ------------------------------------------------------------------------------------------
// Host side.  Every orb_frame borrows a FrameRes (stream, event, device blocks, pinned blocks) from a per-process pool, so
// creating and destroying frames at camera rate allocates nothing once the pool is warm.
// ------------------------------------------------------------------------------------------
namespace {

struct FrameRes {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ready = nullptr;       // recorded after the frame's static data is complete
    uint8_t *d_static = nullptr; size_t static_bytes = 0;   // x, y, angle, uRight, octave, desc, cell_off, cell_idx
    uint8_t *d_work = nullptr;   size_t work_bytes = 0;     // per call: upload image | scratch | candidate lists | results
    uint8_t *h_up = nullptr;     size_t up_bytes = 0;       // pinned: what goes up
    uint8_t *h_dn = nullptr;     size_t dn_bytes = 0;       // pinned: what comes down
};

std::mutex g_pool_mu;
std::vector<FrameRes *> g_pool;
constexpr size_t POOL_MAX = 64;

int res_acquire(int device, FrameRes **out) {
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        for (size_t i = 0; i < g_pool.size(); i++)
            if (g_pool[i]->device == device) {
                *out = g_pool[i];
                g_pool.erase(g_pool.begin() + i);
                return ORB_OK;
            }
    }
    FrameRes *r = new FrameRes();
    r->device = device;
    cudaError_t e = cudaStreamCreateWithFlags(&r->stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&r->ready, cudaEventDisableTiming);
    if (e != cudaSuccess) {
        set_error("orb_frame: %s", cudaGetErrorString(e));
        if (r->stream) cudaStreamDestroy(r->stream);
        delete r;
        return ORB_ERR_CUDA;
    }
    *out = r;
    return ORB_OK;
}

void res_free(FrameRes *r) {
    cudaFree(r->d_static); cudaFree(r->d_work);
    if (r->h_up) cudaFreeHost(r->h_up);
    if (r->h_dn) cudaFreeHost(r->h_dn);
    cudaEventDestroy(r->ready);
    cudaStreamDestroy(r->stream);
    delete r;
}

void res_release(FrameRes *r) {
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        if (g_pool.size() < POOL_MAX) { g_pool.push_back(r); return; }
    }
    res_free(r);
}

int grow_dev(uint8_t **p, size_t *have, size_t need) {
    if (*have >= need) return ORB_OK;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    need = need * 3 / 2 + 65536;
    ORB_CUDA_TRY(cudaMalloc(p, need));
    *have = need;
    return ORB_OK;
}
int grow_host(uint8_t **p, size_t *have, size_t need) {
    if (*have >= need) return ORB_OK;
    if (*p) cudaFreeHost(*p);
    *p = nullptr; *have = 0;
    need = need * 3 / 2 + 65536;
    ORB_CUDA_TRY(cudaMallocHost(p, need));
    *have = need;
    return ORB_OK;
}

constexpr size_t ALIGN = 256;
inline size_t up256(size_t v) { return (v + ALIGN - 1) / ALIGN * ALIGN; }

struct Arena {   // carve 256-byte aligned pieces out of a device block
    uint8_t *base = nullptr;
    size_t used = 0;
    template <class T> T *take(size_t count) {
        T *p = reinterpret_cast<T *>(base + used);
        used += up256(count * sizeof(T));
        return p;
    }
};

// What goes up in one copy: pieces are written into the pinned block at the offset they will have in the device block.
struct Upload {
    uint8_t *h, *d;
    size_t used = 0;
    Upload(uint8_t *host, uint8_t *dev) : h(host), d(dev) {}
    template <class T> T *put(const T *src, size_t count) {   // returns the DEVICE address
        if (count) memcpy(h + used, src, count * sizeof(T));
        T *p = reinterpret_cast<T *>(d + used);
        used += up256(count * sizeof(T));
        return p;
    }
    template <class T> T *slot(size_t count, T **host) {      // caller fills *host
        *host = reinterpret_cast<T *>(h + used);
        T *p = reinterpret_cast<T *>(d + used);
        used += up256(count * sizeof(T));
        return p;
    }
};

#define ORB_TRY(expr) do { int _rc = (expr); if (_rc != ORB_OK) return _rc; } while (0)

}  // namespace

struct orb_frame {
    int device = 0;
    FrameRes *res = nullptr;
    cudaStream_t stream = nullptr;
    int n = 0, nlevels = 0;
    std::vector<float> scale;
    float minX = 0, maxX = 0, minY = 0, maxY = 0;
    FrameDev dev;
    size_t list_hint = 0;          // candidate pairs the largest call so far reserved
    bool up_inflight = false;      // an asynchronous build still reads h_up
    int64_t launches = 0;
    int64_t last_candidates = 0;
    int last_sweeps = 0;
};

namespace {

// carve the static block for n key points; returns the bytes needed (base may be null for sizing)
struct StaticLayout {
    float *x, *y, *angle, *uRight;
    int4 *rec;
    int *octave, *cell_off, *cell_idx, *cell_of;
    uint8_t *desc;
    size_t bytes;
};
StaticLayout static_layout(uint8_t *base, int n) {
    Arena ar;
    ar.base = base;
    StaticLayout L;
    const size_t m = (size_t)std::max(n, 1);
    L.x = ar.take<float>(m); L.y = ar.take<float>(m); L.angle = ar.take<float>(m); L.uRight = ar.take<float>(m);
    L.octave = ar.take<int>(m);
    L.desc = ar.take<uint8_t>(m * 32);
    L.cell_off = ar.take<int>(NCELL + 1); L.cell_idx = ar.take<int>(m); L.cell_of = ar.take<int>(m);
    L.rec = ar.take<int4>(m);
    L.bytes = ar.used;
    return L;
}

int frame_begin(int n, int nlevels, const float *scale_factors, float min_x, float max_x, float min_y, float max_y, orb_frame **out) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    orb_frame *f = new orb_frame();
    cudaGetDevice(&f->device);
    int rc = res_acquire(f->device, &f->res);
    if (rc != ORB_OK) { delete f; return rc; }
    f->stream = f->res->stream;
    f->n = n; f->nlevels = nlevels;
    f->scale.assign(scale_factors, scale_factors + nlevels);
    f->minX = min_x; f->maxX = max_x; f->minY = min_y; f->maxY = max_y;
    *out = f;
    return ORB_OK;
}

void frame_abort(orb_frame *f) {
    cudaStreamSynchronize(f->stream);
    res_release(f->res);
    delete f;
}

void frame_set_dev(orb_frame *f, const StaticLayout &L, bool keyframe_view) {
    // ref Frame.cc:217-218; a KeyFrame sees the bounds as `int` copies (KeyFrame.h:190-193) while the grid keeps the Frame's scales
    const float invW = static_cast<float>(GRID_COLS) / static_cast<float>(f->maxX - f->minX);
    const float invH = static_cast<float>(GRID_ROWS) / static_cast<float>(f->maxY - f->minY);
    f->dev = FrameDev{f->n, L.x, L.y, L.angle, L.uRight, L.octave, L.desc, L.cell_off, L.cell_idx, L.rec,
                      keyframe_view ? (float)(int)f->minX : f->minX, keyframe_view ? (float)(int)f->minY : f->minY, invW, invH};
}

// before the host touches h_up again: an asynchronous build may still be reading it
int settle(orb_frame *f) {
    if (f->up_inflight) {
        ORB_CUDA_TRY(cudaStreamSynchronize(f->stream));
        f->up_inflight = false;
    }
    return ORB_OK;
}

// another frame's static data is produced on that frame's stream
int wait_for(orb_frame *f, const orb_frame *other) {
    if (other != f) ORB_CUDA_TRY(cudaStreamWaitEvent(f->stream, other->res->ready, 0));
    return ORB_OK;
}

}  // namespace

extern "C" {

int orb_frame_create(const orb_frame_desc *d, orb_frame **out) {
    ORB_REQUIRE(d && out && d->n >= 0 && (d->n == 0 || (d->keys_un && d->descriptors)) && d->scale_factors && d->nlevels >= 1,
                ORB_ERR_INVALID, "bad frame description");
    ORB_REQUIRE(d->max_x > d->min_x && d->max_y > d->min_y, ORB_ERR_INVALID, "empty image bounds");
    ORB_REQUIRE(d->n < (1 << 24), ORB_ERR_INVALID, "more than 16M key points in one frame");
    for (int i = 0; i < d->n; i++) ORB_REQUIRE(d->keys_un[i].octave >= 0 && d->keys_un[i].octave < 128, ORB_ERR_INVALID, "key point octave out of range");
    orb_frame *f = nullptr;
    ORB_TRY(frame_begin(d->n, d->nlevels, d->scale_factors, d->min_x, d->max_x, d->min_y, d->max_y, &f));
    FrameRes *R = f->res;
    const int n = d->n;
    const size_t bytes = static_layout(nullptr, n).bytes;
    int rc = grow_dev(&R->d_static, &R->static_bytes, bytes);
    if (rc == ORB_OK) rc = grow_host(&R->h_up, &R->up_bytes, bytes);
    if (rc != ORB_OK) { frame_abort(f); return rc; }
    // the static block is assembled in pinned memory (SoA + the grid as CSR) and goes up in one copy.
    // ref Frame.cc:217-218 and AssignFeaturesToGrid / PosInGrid :235-250, 387-397 (round(), not floor)
    const StaticLayout H = static_layout(R->h_up, n), L = static_layout(R->d_static, n);
    const float invW = static_cast<float>(GRID_COLS) / static_cast<float>(f->maxX - f->minX);
    const float invH = static_cast<float>(GRID_ROWS) / static_cast<float>(f->maxY - f->minY);
    int *off = H.cell_off, *cell = H.cell_of;
    memset(off, 0, (NCELL + 1) * sizeof(int));
    for (int i = 0; i < n; i++) {
        const orb_keypoint &k = d->keys_un[i];
        H.x[i] = k.x; H.y[i] = k.y; H.angle[i] = k.angle; H.octave[i] = k.octave;
        H.uRight[i] = d->u_right ? d->u_right[i] : -1.0f;
        const int px = (int)roundf((k.x - f->minX) * invW), py = (int)roundf((k.y - f->minY) * invH);
        cell[i] = (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) ? -1 : px * GRID_ROWS + py;
        if (cell[i] >= 0) off[cell[i] + 1]++;
    }
    for (int c = 0; c < NCELL; c++) off[c + 1] += off[c];
    {
        std::vector<int> cur(off, off + NCELL);
        for (int i = 0; i < n; i++) if (cell[i] >= 0) H.cell_idx[cur[cell[i]]++] = i;
    }
    for (int e = 0; e < off[NCELL]; e++) {   // one 16-byte record per grid entry: what the window test reads
        const int i = H.cell_idx[e];
        int4 r;
        memcpy(&r.x, &H.x[i], 4); memcpy(&r.y, &H.y[i], 4); memcpy(&r.z, &H.uRight[i], 4);
        r.w = i | (H.octave[i] << 24);
        H.rec[e] = r;
    }
    if (n) memcpy(H.desc, d->descriptors, (size_t)n * 32);
    cudaError_t e = cudaMemcpyAsync(R->d_static, R->h_up, bytes, cudaMemcpyHostToDevice, f->stream);
    if (e == cudaSuccess) e = cudaEventRecord(R->ready, f->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(f->stream);
    if (e != cudaSuccess) { set_error("orb_frame_create: %s", cudaGetErrorString(e)); frame_abort(f); return ORB_ERR_CUDA; }
    frame_set_dev(f, L, d->keyframe_view != 0);
    *out = f;
    return ORB_OK;
}

int orb_frame_from_extractor(orb_extractor *ex, int slot, const orb_frame_source *src, orb_frame **out) {
    ORB_REQUIRE(ex && src && out, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(ex->last_kp && slot >= 0 && slot < ex->last_batch, ORB_ERR_INVALID, "no extraction result in slot %d", slot);
    ORB_REQUIRE(src->max_x > src->min_x && src->max_y > src->min_y, ORB_ERR_INVALID, "empty image bounds");
    ORB_REQUIRE(src->u_right == nullptr || src->u_right_loc == ORB_HOST || src->u_right_loc == ORB_DEVICE, ORB_ERR_INVALID, "bad u_right_loc");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    int n;
    if (ex->last_counts_valid) n = ex->last_counts[slot];
    else {
        ORB_CUDA_TRY(cudaMemcpyAsync(ex->h_pinned, ex->last_n + slot, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
        ORB_CUDA_TRY(cudaStreamSynchronize(ex->stream));
        n = ex->h_pinned[0];
    }
    n = std::min(std::max(n, 0), ex->last_capacity);
    orb_frame *f = nullptr;
    ORB_TRY(frame_begin(n, ex->prm.nlevels, ex->scale.data(), src->min_x, src->max_x, src->min_y, src->max_y, &f));
    FrameRes *R = f->res;
    const size_t bytes = static_layout(nullptr, n).bytes + up256((size_t)std::max(n, 1) * sizeof(orb_keypoint));
    int rc = grow_dev(&R->d_static, &R->static_bytes, bytes);
    if (rc != ORB_OK) { frame_abort(f); return rc; }
    const StaticLayout L = static_layout(R->d_static, n);
    orb_keypoint *d_un = reinterpret_cast<orb_keypoint *>(R->d_static + L.bytes);   // mvKeysUn when the camera has distortion
    cudaStream_t st = f->stream;
    auto fail = [&](cudaError_t e) { set_error("orb_frame_from_extractor: %s", cudaGetErrorString(e)); frame_abort(f); return ORB_ERR_CUDA; };
    // the extractor's results are produced on ITS stream
    cudaError_t e = cudaEventRecord(R->ready, ex->stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(st, R->ready, 0);
    if (e != cudaSuccess) return fail(e);
    const orb_keypoint *kp = ex->last_kp + (size_t)slot * ex->last_capacity;
    const uint8_t *desc = ex->last_desc + (size_t)slot * ex->last_capacity * 32;
    const float *d_ur = nullptr;
    if (src->u_right && n) {
        if (src->u_right_loc == ORB_DEVICE) d_ur = src->u_right;
        else {
            rc = grow_host(&R->h_up, &R->up_bytes, (size_t)n * 4);
            if (rc == ORB_OK) rc = grow_dev(&R->d_work, &R->work_bytes, (size_t)n * 4);
            if (rc != ORB_OK) { frame_abort(f); return rc; }
            memcpy(R->h_up, src->u_right, (size_t)n * 4);
            e = cudaMemcpyAsync(R->d_work, R->h_up, (size_t)n * 4, cudaMemcpyHostToDevice, st);
            if (e != cudaSuccess) return fail(e);
            d_ur = reinterpret_cast<const float *>(R->d_work);
            f->up_inflight = true;
        }
    }
    if (n && src->ndist > 0 && src->dist_coef[0] != 0.0f) {   // ref Frame.cc:409-439 (mvKeysUn = mvKeys otherwise, :411-415)
        rc = orbgpu_undistort_keypoints_async(kp, n, src->fx, src->fy, src->cx, src->cy, src->dist_coef, src->ndist, d_un, st);
        if (rc != ORB_OK) { frame_abort(f); return rc; }
        kp = d_un;
        f->launches++;
    }
    FrameBuild B;
    B.x = L.x; B.y = L.y; B.angle = L.angle; B.uRight = L.uRight; B.octave = L.octave;
    B.cell_off = L.cell_off; B.cell_idx = L.cell_idx; B.cell_of = L.cell_of; B.rec = L.rec;
    B.minX = f->minX; B.minY = f->minY;
    B.invW = static_cast<float>(GRID_COLS) / static_cast<float>(f->maxX - f->minX);
    B.invH = static_cast<float>(GRID_ROWS) / static_cast<float>(f->maxY - f->minY);
    k_frame_build<<<1, 1024, 0, st>>>(kp, n, d_ur, B);
    f->launches++;
    e = cudaGetLastError();
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(L.desc, desc, (size_t)n * 32, cudaMemcpyDeviceToDevice, st);
    if (e == cudaSuccess) e = cudaEventRecord(R->ready, st);
    if (e != cudaSuccess) return fail(e);
    frame_set_dev(f, L, src->keyframe_view != 0);
    *out = f;
    return ORB_OK;
}

void orb_frame_destroy(orb_frame *f) {
    if (!f) return;
    cudaSetDevice(f->device);
    cudaStreamSynchronize(f->stream);
    res_release(f->res);
    delete f;
}

int orb_frame_sync(orb_frame *f) {
    ORB_REQUIRE(f, ORB_ERR_INVALID, "null handle");
    ORB_CUDA_TRY(cudaSetDevice(f->device));
    ORB_CUDA_TRY(cudaStreamSynchronize(f->stream));
    f->up_inflight = false;
    return ORB_OK;
}
int orb_frame_size(const orb_frame *f) { return f ? f->n : 0; }
int64_t orb_frame_launch_count(const orb_frame *f) { return f ? f->launches : 0; }
void *orb_frame_stream(orb_frame *f) { return f ? (void *)f->stream : nullptr; }
int orb_frame_last_stats(const orb_frame *f, int64_t *candidates, int32_t *sweeps) {
    ORB_REQUIRE(f, ORB_ERR_INVALID, "null handle");
    if (candidates) *candidates = f->last_candidates;
    if (sweeps) *sweeps = f->last_sweeps;
    return ORB_OK;
}

}  // extern "C"

namespace {

// Host description of a query set (all pointers host unless noted); packed into the frame's upload block.
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

size_t queries_upload_bytes(const HostQueries &H) {
    const size_t n = (size_t)std::max(H.n, 1);
    return 16 * ALIGN + n * (1 + 4 * 3 + 4 * 2 + 4 * 2 + 32 + 4 + 1 + 4) + (size_t)H.gate_levels * 4 + ALIGN;
}

WinQ pack_queries(Upload &up, const HostQueries &H) {
    const int n = H.n;
    WinQ Q;
    memset(&Q, 0, sizeof(Q));
    Q.n = n;
    Q.valid = H.valid ? up.put(H.valid, n) : nullptr;
    Q.x = up.put(H.x, n); Q.y = up.put(H.y, n); Q.xy_stride = 1;
    Q.radius = up.put(H.radius, n);
    Q.min_level = up.put(H.min_level, n); Q.max_level = up.put(H.max_level, n);
    Q.strict_levels = H.strict_levels;
    Q.stereo_ref = H.stereo_ref ? up.put(H.stereo_ref, n) : nullptr;
    Q.stereo_tol = H.stereo_ref ? up.put(H.stereo_tol, n) : nullptr;
    Q.desc = H.desc ? up.put(H.desc, (size_t)n * 32) : nullptr;
    Q.angle = H.angle ? up.put(H.angle, n) : nullptr;
    Q.claims = H.claims ? up.put(H.claims, n) : nullptr;
    Q.gate_ur = H.gate_inv_sigma2 ? up.put(H.gate_ur, n) : nullptr;
    Q.gate_inv_sigma2 = H.gate_inv_sigma2 ? up.put(H.gate_inv_sigma2, H.gate_levels) : nullptr;
    Q.gate_mono = H.gate_mono; Q.gate_stereo = H.gate_stereo;
    return Q;
}

size_t list_capacity(const orb_frame *f, int n) { return std::max<size_t>(f->list_hint, (size_t)32 * n + 4096); }

// wide = the windows hold hundreds of grid entries (SearchForInitialization's 100-px windows): a whole CTA per query
void launch_lists(orb_frame *f, const WinQ &Q, int *q_off, int *q_cnt, int2 *lists, size_t cap, int *hdr, int prune, bool wide = false) {
    if (Q.n == 0) return;
    const int icap = (int)std::min<size_t>(cap, 0x7fffffff);
    if (wide) k_window_lists<256><<<Q.n, 256, 0, f->stream>>>(f->dev, Q, q_off, q_cnt, lists, icap, hdr, prune);
    else k_window_lists<32><<<(Q.n + 7) / 8, 256, 0, f->stream>>>(f->dev, Q, q_off, q_cnt, lists, icap, hdr, prune);
    f->launches++;
}

// CTAs of a resolve kernel: ~2 queries per thread with short lists (one CTA up to ~2000 queries: block barriers instead of
// grid barriers), 16 lanes per query with the long lists of wide windows
int resolve_grid(const orb_frame *f, int nqueries, bool wide) {
    static int sms[64] = {};
    int &n = sms[f->device & 63];
    if (!n) { if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, f->device) != cudaSuccess || n <= 0) n = 1; }
    const long long lanes = (long long)nqueries * (wide ? 16 : 1);
    return (int)std::max<long long>(1, std::min<long long>(n, (lanes + 2 * RES_BLOCK - 1) / (2 * RES_BLOCK)));
}
int launch_resolve(const void *kernel, int grid, void **args, cudaStream_t st) {
    // one CTA: an ordinary launch (the kernel then never touches the grid barrier); more: cooperative, all CTAs co-resident
    if (grid == 1) ORB_CUDA_TRY(cudaLaunchKernel(kernel, dim3(1), dim3(RES_BLOCK), args, 0, st));
    else ORB_CUDA_TRY(cudaLaunchCooperativeKernel(kernel, dim3(grid), dim3(RES_BLOCK), args, 0, st));
    return ORB_OK;
}

// The windowed matcher: upload, lists, resolve, download.  Retries once with a larger candidate pool when the pool overflowed.
int search_windowed_host(orb_frame *f, const HostQueries &H, const uint8_t *claimed, int select, int th_dist, float nnratio,
                         int check_ori, int32_t *assigned, int *nmatches) {
    ORB_CUDA_TRY(cudaSetDevice(f->device));
    const int n = H.n, nf = f->n;
    if (nmatches) *nmatches = 0;
    for (int i = 0; i < nf; i++) assigned[i] = -1;
    if (n == 0 || nf == 0) return ORB_OK;
    ORB_TRY(settle(f));
    FrameRes *R = f->res;
    cudaStream_t st = f->stream;
    for (int attempt = 0; attempt < 3; attempt++) {
        const size_t cap = list_capacity(f, n);
        const size_t up_need = queries_upload_bytes(H) + up256(nf);
        const size_t scratch = 4 * up256((size_t)n * 4) + 3 * up256((size_t)nf * 4) + up256(nf);
        const size_t dn_need = up256((size_t)nf * 4) + ALIGN;
        ORB_TRY(grow_host(&R->h_up, &R->up_bytes, up_need));
        ORB_TRY(grow_host(&R->h_dn, &R->dn_bytes, dn_need));
        ORB_TRY(grow_dev(&R->d_work, &R->work_bytes, up_need + scratch + dn_need + cap * 8 + ALIGN));
        Upload up(R->h_up, R->d_work);
        int *h_hdr;
        int *hdr = up.slot<int>(64, &h_hdr);
        memset(h_hdr, 0, 64 * sizeof(int));
        const WinQ Q = pack_queries(up, H);
        const uint8_t *d_claimed = claimed ? up.put(claimed, nf) : nullptr;
        Arena ar;
        ar.base = R->d_work; ar.used = up.used;
        int *q_off = ar.take<int>(n), *q_cnt = ar.take<int>(n), *res = ar.take<int>(n), *bins = ar.take<int>(n);
        unsigned long long *held = ar.take<unsigned long long>(nf);
        int *holder = ar.take<int>(nf);
        uint8_t *bad = ar.take<uint8_t>(nf);
        // results are written by the kernel straight into the pinned block (mapped into the device's address space): no copy down
        int *d_assigned = reinterpret_cast<int *>(R->h_dn), *d_out = reinterpret_cast<int *>(R->h_dn + up256((size_t)nf * 4));
        int2 *lists = ar.take<int2>(0);
        ORB_CUDA_TRY(cudaMemcpyAsync(R->d_work, R->h_up, up.used, cudaMemcpyHostToDevice, st));
        launch_lists(f, Q, q_off, q_cnt, lists, cap, hdr, select == ORB_SELECT_BEST ? th_dist : -1);
        {
            FrameDev fd = f->dev;
            WinQ Qa = Q;
            const int *a_off = q_off, *a_cnt = q_cnt;
            const int2 *a_lists = lists;
            float ratio = nnratio;
            int sel = select, th = th_dist, ori = check_ori;
            void *args[] = {&fd, &Qa, &a_off, &a_cnt, &a_lists, &d_claimed, &held, &res, &bins, &holder, &bad, &sel, &th, &ratio, &ori, &d_assigned,
                            &d_out, &hdr};
            ORB_TRY(launch_resolve((const void *)k_resolve_windowed, resolve_grid(f, n, false), args, st));
        }
        f->launches++;
        ORB_CUDA_TRY(cudaGetLastError());
        ORB_CUDA_TRY(cudaStreamSynchronize(st));
        const int *h_out = d_out;
        if (h_out[3]) {   // candidate pool too small: h_out[2] slots were asked for
            f->list_hint = (size_t)h_out[2] * 5 / 4 + 4096;
            continue;
        }
        f->list_hint = std::max(f->list_hint, (size_t)h_out[2] * 5 / 4 + 1024);
        memcpy(assigned, R->h_dn, (size_t)nf * 4);
        if (nmatches) *nmatches = h_out[0];
        f->last_sweeps = h_out[1];
        f->last_candidates = h_out[4];
        return ORB_OK;
    }
    set_error("candidate pool kept overflowing");
    return ORB_ERR_CAPACITY;
}

}  // namespace

extern "C" {

int orb_frame_features_in_area(orb_frame *f, float x, float y, float r, int min_level, int max_level, int32_t *indices,
                               int capacity, int *n_out) {
    ORB_REQUIRE(f && indices && n_out, ORB_ERR_INVALID, "null argument");
    ORB_CUDA_TRY(cudaSetDevice(f->device));
    *n_out = 0;
    if (f->n == 0) return ORB_OK;
    ORB_TRY(settle(f));
    FrameRes *R = f->res;
    cudaStream_t st = f->stream;
    HostQueries H;
    H.n = 1; H.x = &x; H.y = &y; H.radius = &r; H.min_level = &min_level; H.max_level = &max_level;
    const size_t cap = (size_t)f->n + 64;
    const size_t up_need = queries_upload_bytes(H), dn_need = up256(cap * 8) + 2 * ALIGN;
    ORB_TRY(grow_host(&R->h_up, &R->up_bytes, up_need));
    ORB_TRY(grow_host(&R->h_dn, &R->dn_bytes, dn_need));
    ORB_TRY(grow_dev(&R->d_work, &R->work_bytes, up_need + dn_need + 4 * ALIGN));
    Upload up(R->h_up, R->d_work);
    int *h_hdr;
    int *hdr = up.slot<int>(64, &h_hdr);
    memset(h_hdr, 0, 64 * sizeof(int));
    const WinQ Q = pack_queries(up, H);
    Arena ar;
    ar.base = R->d_work; ar.used = up.used;
    int *q_off = ar.take<int>(1), *q_cnt = ar.take<int>(1);   // one 256-byte piece each: [q_off | q_cnt | lists] go down together
    int2 *lists = ar.take<int2>(cap);
    ORB_CUDA_TRY(cudaMemcpyAsync(R->d_work, R->h_up, up.used, cudaMemcpyHostToDevice, st));
    launch_lists(f, Q, q_off, q_cnt, lists, cap, hdr, -1);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(R->h_dn, q_off, 2 * ALIGN + cap * 8, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    const int off = *reinterpret_cast<const int *>(R->h_dn), cnt = *reinterpret_cast<const int *>(R->h_dn + ALIGN);
    const int2 *hl = reinterpret_cast<const int2 *>(R->h_dn + 2 * ALIGN) + off;
    *n_out = cnt;
    ORB_REQUIRE(cnt <= capacity, ORB_ERR_CAPACITY, "index buffer too small");
    for (int i = 0; i < cnt; i++) indices[i] = hl[i].x;
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
    orb_frame *f = f2;   // runs on the second frame's stream and scratch
    ORB_TRY(settle(f));
    ORB_TRY(wait_for(f, f1));
    FrameRes *R = f->res;
    cudaStream_t st = f->stream;
    for (int attempt = 0; attempt < 3; attempt++) {
        const size_t cap = list_capacity(f, n1);
        const size_t up_need = 2 * ALIGN + up256((size_t)n1 * 8);
        const size_t scratch = 6 * up256((size_t)n1 * 4) + 3 * up256((size_t)n2 * 4);
        const size_t dn_need = up256((size_t)n1 * 4) + up256((size_t)n1 * 8) + ALIGN;
        ORB_TRY(grow_host(&R->h_up, &R->up_bytes, up_need));
        ORB_TRY(grow_host(&R->h_dn, &R->dn_bytes, dn_need));
        ORB_TRY(grow_dev(&R->d_work, &R->work_bytes, up_need + scratch + dn_need + cap * 8 + ALIGN));
        Upload up(R->h_up, R->d_work);
        int *h_hdr;
        int *hdr = up.slot<int>(64, &h_hdr);
        memset(h_hdr, 0, 64 * sizeof(int));
        const float *d_prev = up.put(prev_matched, (size_t)n1 * 2);
        // the queries are F1's own features: octave-0 ones only (ref :421-423), window around vbPrevMatched at that level
        // (:425), descriptors and levels straight from F1's device arrays
        WinQ Q;
        memset(&Q, 0, sizeof(Q));
        Q.n = n1; Q.valid_if_zero = f1->dev.octave;
        Q.x = d_prev; Q.y = d_prev + 1; Q.xy_stride = 2;
        Q.radius_const = (float)window_size;
        Q.min_level = f1->dev.octave; Q.max_level = f1->dev.octave;
        Q.desc = f1->dev.desc;
        Arena ar;
        ar.base = R->d_work; ar.used = up.used;
        int *q_off = ar.take<int>(n1), *q_cnt = ar.take<int>(n1), *next = ar.take<int>(n1), *tdist = ar.take<int>(n1);
        int *res = ar.take<int>(n1), *res_dist = ar.take<int>(n1), *last = ar.take<int>(n2);
        unsigned long long *head = ar.take<unsigned long long>(n2);
        // [m12 | prev | out]: written by the kernel straight into the pinned block (mapped into the device's address space)
        const size_t o_prev = up256((size_t)n1 * 4), o_out = o_prev + up256((size_t)n1 * 8);
        int *m12 = reinterpret_cast<int *>(R->h_dn);
        float *d_prev_out = reinterpret_cast<float *>(R->h_dn + o_prev);
        int *d_out = reinterpret_cast<int *>(R->h_dn + o_out);
        int2 *lists = ar.take<int2>(0);
        ORB_CUDA_TRY(cudaMemcpyAsync(R->d_work, R->h_up, up.used, cudaMemcpyHostToDevice, st));
        // a window of r px holds about n2 * (2r)^2 / (image area) grid entries
        const double pop = (double)n2 * 4.0 * window_size * window_size / ((double)(f2->maxX - f2->minX) * (f2->maxY - f2->minY));
        launch_lists(f, Q, q_off, q_cnt, lists, cap, hdr, -1, pop > 96.0);
        {
            FrameDev d1 = f1->dev, d2 = f2->dev;
            const int *a_off = q_off, *a_cnt = q_cnt;
            const int2 *a_lists = lists;
            float ratio = nnratio;
            int ori = check_orientation;
            void *args[] = {&d1, &d2, &a_off, &a_cnt, &a_lists, &head, &next, &tdist, &res, &res_dist, &last, &ratio, &ori, &m12, &d_prev,
                            &d_prev_out, &d_out, &hdr};
            ORB_TRY(launch_resolve((const void *)k_resolve_init, resolve_grid(f, n1, pop > 96.0), args, st));
        }
        f->launches++;
        ORB_CUDA_TRY(cudaGetLastError());
        ORB_CUDA_TRY(cudaStreamSynchronize(st));
        const int *h_out = d_out;
        if (h_out[3]) {
            f->list_hint = (size_t)h_out[2] * 5 / 4 + 4096;
            continue;
        }
        f->list_hint = std::max(f->list_hint, (size_t)h_out[2] * 5 / 4 + 1024);
        memcpy(matches12, R->h_dn, (size_t)n1 * 4);
        memcpy(prev_matched, R->h_dn + o_prev, (size_t)n1 * 8);
        if (nmatches) *nmatches = h_out[0];
        f->last_sweeps = h_out[1];
        f->last_candidates = h_out[4];
        return ORB_OK;
    }
    set_error("candidate pool kept overflowing");
    return ORB_ERR_CAPACITY;
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
    if (nent == 0) return ORB_OK;
    ORB_TRY(settle(f));
    ORB_TRY(wait_for(f, kf));
    FrameRes *R = f->res;
    cudaStream_t st = f->stream;
    size_t npairs = 0;
    for (int g = 0; g < ngroups; g++) npairs += (size_t)(kf_off[g + 1] - kf_off[g]) * (size_t)(f_off[g + 1] - f_off[g]);
    for (int a = 0; a < nent; a++) ORB_REQUIRE(kf_idx[a] >= 0 && kf_idx[a] < kf->n, ORB_ERR_INVALID, "KF feature index out of range");
    for (int b = 0; b < nfent; b++) ORB_REQUIRE(f_idx[b] >= 0 && f_idx[b] < f->n, ORB_ERR_INVALID, "F feature index out of range");
    const size_t up_need = 4 * up256((size_t)nent * 4) + up256((size_t)std::max(nfent, 1) * 4) + up256(kf->n) + up256(f->n) + ALIGN;
    const size_t scratch = up256(npairs * 4) + up256((size_t)nent * 4) + up256((size_t)f->n * 4);
    const size_t dn_need = up256((size_t)f->n * 4) + ALIGN;
    ORB_TRY(grow_host(&R->h_up, &R->up_bytes, up_need));
    ORB_TRY(grow_host(&R->h_dn, &R->dn_bytes, dn_need));
    ORB_TRY(grow_dev(&R->d_work, &R->work_bytes, up_need + scratch + dn_need + ALIGN));
    uint8_t *D = R->d_work;
    Upload up(R->h_up, D);
    const int *d_kfidx = up.put(kf_idx, nent);
    int *h_fo, *h_fn, *h_po;
    const int *d_fo = up.slot<int>(nent, &h_fo), *d_fn = up.slot<int>(nent, &h_fn), *d_po = up.slot<int>(nent, &h_po);
    {
        size_t pos = 0;
        for (int g = 0; g < ngroups; g++)
            for (int a = kf_off[g]; a < kf_off[g + 1]; a++) {
                h_fo[a] = f_off[g]; h_fn[a] = f_off[g + 1] - f_off[g]; h_po[a] = (int)pos;
                pos += h_fn[a];
            }
    }
    const int *d_fidx = up.put(f_idx, nfent);
    const uint8_t *d_kv = up.put(kf_valid, kf->n);
    const uint8_t *d_fv = f_valid ? up.put(f_valid, f->n) : nullptr;
    Arena ar;
    ar.base = D; ar.used = up.used;
    int *d_dist = ar.take<int>(npairs), *res = ar.take<int>(nent), *taken_by = ar.take<int>(f->n);
    int *d_matches = ar.take<int>(f->n), *d_out = ar.take<int>(8);
    ORB_CUDA_TRY(cudaMemcpyAsync(D, R->h_up, up.used, cudaMemcpyHostToDevice, st));
    k_bow_dist<<<(nent * 32 + 255) / 256, 256, 0, st>>>(kf->dev, f->dev, nent, d_kfidx, d_fo, d_fn, d_po, d_fidx, d_kv, d_dist);
    k_resolve_bow<<<1, RES_THREADS, 0, st>>>(kf->dev, f->dev, nent, d_kfidx, d_fo, d_fn, d_po, d_fidx, d_kv, d_fv, d_dist, strict_low, nnratio,
                                             check_orientation, taken_by, res, d_matches, d_out);
    f->launches += 2;
    ORB_CUDA_TRY(cudaGetLastError());
    const size_t o_out = up256((size_t)f->n * 4);
    ORB_CUDA_TRY(cudaMemcpyAsync(R->h_dn, d_matches, o_out + 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    const int *h_out = reinterpret_cast<const int *>(R->h_dn + o_out);
    memcpy(matches, R->h_dn, (size_t)f->n * 4);
    if (nmatches) *nmatches = h_out[0];
    f->last_sweeps = h_out[1];
    f->last_candidates = (int64_t)npairs;
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
    ORB_TRY(settle(f));
    FrameRes *R = f->res;
    cudaStream_t st = f->stream;
    for (int attempt = 0; attempt < 3; attempt++) {
        const size_t cap = list_capacity(f, n);
        const size_t up_need = queries_upload_bytes(H);
        const size_t scratch = 2 * up256((size_t)n * 4);
        const size_t dn_need = 2 * up256((size_t)n * 4) + ALIGN;
        ORB_TRY(grow_host(&R->h_up, &R->up_bytes, up_need));
        ORB_TRY(grow_host(&R->h_dn, &R->dn_bytes, dn_need));
        ORB_TRY(grow_dev(&R->d_work, &R->work_bytes, up_need + scratch + dn_need + cap * 8 + ALIGN));
        Upload up(R->h_up, R->d_work);
        int *h_hdr;
        int *hdr = up.slot<int>(64, &h_hdr);
        memset(h_hdr, 0, 64 * sizeof(int));
        const WinQ Q = pack_queries(up, H);
        Arena ar;
        ar.base = R->d_work; ar.used = up.used;
        int *q_off = ar.take<int>(n), *q_cnt = ar.take<int>(n);
        int *d_idx = ar.take<int>(n), *d_dist = ar.take<int>(n);
        int *d_hdr_copy = ar.take<int>(8);
        int2 *lists = ar.take<int2>(0);
        ORB_CUDA_TRY(cudaMemcpyAsync(R->d_work, R->h_up, up.used, cudaMemcpyHostToDevice, st));
        launch_lists(f, Q, q_off, q_cnt, lists, cap, hdr, -1);
        k_window_best<<<(n + 255) / 256, 256, 0, st>>>(n, q_off, q_cnt, lists, d_idx, d_dist);
        f->launches++;
        ORB_CUDA_TRY(cudaGetLastError());
        ORB_CUDA_TRY(cudaMemcpyAsync(d_hdr_copy, hdr, 8 * sizeof(int), cudaMemcpyDeviceToDevice, st));
        const size_t o_dist = up256((size_t)n * 4), o_hdr = 2 * o_dist;
        ORB_CUDA_TRY(cudaMemcpyAsync(R->h_dn, d_idx, o_hdr + 8 * sizeof(int), cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaStreamSynchronize(st));
        const int *h_hdr2 = reinterpret_cast<const int *>(R->h_dn + o_hdr);
        if (h_hdr2[1]) {
            f->list_hint = (size_t)h_hdr2[0] * 5 / 4 + 4096;
            continue;
        }
        f->list_hint = std::max(f->list_hint, (size_t)h_hdr2[0] * 5 / 4 + 1024);
        memcpy(best_idx, R->h_dn, (size_t)n * 4);
        memcpy(best_dist, R->h_dn + o_dist, (size_t)n * 4);
        f->last_candidates = h_hdr2[2];
        f->last_sweeps = 0;
        return ORB_OK;
    }
    set_error("candidate pool kept overflowing");
    return ORB_ERR_CAPACITY;
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
    orb_frame *f = kf1;   // scratch and stream of the first key frame
    ORB_TRY(settle(f));
    ORB_TRY(wait_for(f, kf2));
    FrameRes *R = f->res;
    cudaStream_t st = f->stream;
    const int nl = kf2->nlevels;
    const size_t up_need = 3 * up256((size_t)nent * 4) + up256((size_t)nent2 * 4) + up256(kf1->n) + up256(kf2->n) + 2 * up256((size_t)nl * 4) + ALIGN;
    const size_t dn_need = up256((size_t)kf1->n * 4) + ALIGN;
    ORB_TRY(grow_host(&R->h_up, &R->up_bytes, up_need));
    ORB_TRY(grow_host(&R->h_dn, &R->dn_bytes, dn_need));
    ORB_TRY(grow_dev(&R->d_work, &R->work_bytes, up_need + dn_need + ALIGN));
    Upload up(R->h_up, R->d_work);
    const int *d_idx1 = up.put(idx1, nent);
    int *h_o2, *h_n2;
    const int *d_o2 = up.slot<int>(nent, &h_o2), *d_n2 = up.slot<int>(nent, &h_n2);
    for (int g = 0; g < ngroups; g++) {
        ORB_REQUIRE(off2[g + 1] - off2[g] < 65536, ORB_ERR_INVALID, "vocabulary node with more than 65535 features");
        for (int a = off1[g]; a < off1[g + 1]; a++) {
            ORB_REQUIRE(idx1[a] >= 0 && idx1[a] < kf1->n, ORB_ERR_INVALID, "KF1 feature index out of range");
            h_o2[a] = off2[g]; h_n2[a] = off2[g + 1] - off2[g];
        }
    }
    for (int b = 0; b < nent2; b++) ORB_REQUIRE(idx2[b] >= 0 && idx2[b] < kf2->n, ORB_ERR_INVALID, "KF2 feature index out of range");
    const int *d_idx2 = up.put(idx2, nent2);
    const uint8_t *d_u1 = up.put(usable1, kf1->n), *d_u2 = up.put(usable2, kf2->n);
    const float *d_sigma = up.put(level_sigma2, nl), *d_scale = up.put(kf2->scale.data(), nl);
    Arena ar;
    ar.base = R->d_work; ar.used = up.used;
    int *d_m12 = ar.take<int>(kf1->n), *d_nm = ar.take<int>(8);
    ORB_CUDA_TRY(cudaMemcpyAsync(R->d_work, R->h_up, up.used, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemsetAsync(d_m12, 0xff, (size_t)kf1->n * 4, st));
    TriArgs A;
    for (int i = 0; i < 9; i++) A.F[i] = F12[i];
    A.ex = ex; A.ey = ey; A.sigma2 = d_sigma; A.scale2 = d_scale; A.only_stereo = only_stereo;
    k_triangulate<<<(nent * 32 + 255) / 256, 256, 0, st>>>(kf1->dev, kf2->dev, nent, d_idx1, d_o2, d_n2, d_idx2, d_u1, d_u2, A, d_m12);
    k_tri_filter<<<1, 1024, 0, st>>>(kf1->dev, kf2->dev, d_m12, check_orientation, d_nm);
    f->launches += 2;
    ORB_CUDA_TRY(cudaGetLastError());
    const size_t o_nm = up256((size_t)kf1->n * 4);
    ORB_CUDA_TRY(cudaMemcpyAsync(R->h_dn, d_m12, o_nm + sizeof(int), cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    memcpy(matches12, R->h_dn, (size_t)kf1->n * 4);
    if (nmatches) *nmatches = *reinterpret_cast<const int *>(R->h_dn + o_nm);
    return ORB_OK;
}

int orb_rot_hist_filter(const float *angle_a, const float *angle_b, int n, uint8_t *keep) {
    ORB_REQUIRE(n >= 0 && (n == 0 || (angle_a && angle_b && keep)), ORB_ERR_INVALID, "bad argument");
    if (n == 0) return ORB_OK;
    ThreadCtx *T = nullptr;
    ORB_TRY(orbgpu_thread_ctx(&T, (size_t)n * 9 + 4 * ALIGN, (size_t)n * 9 + 4 * ALIGN));
    float *ha = reinterpret_cast<float *>(T->h_buf), *hb = ha + n;
    memcpy(ha, angle_a, (size_t)n * 4);
    memcpy(hb, angle_b, (size_t)n * 4);
    float *da = reinterpret_cast<float *>(T->d_buf), *db = da + n;
    uint8_t *dk = reinterpret_cast<uint8_t *>(db + n), *hk = reinterpret_cast<uint8_t *>(hb + n);
    ORB_CUDA_TRY(cudaMemcpyAsync(da, ha, (size_t)n * 8, cudaMemcpyHostToDevice, T->stream));
    k_rot_hist_filter<<<1, 1024, 0, T->stream>>>(da, db, n, dk);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(hk, dk, n, cudaMemcpyDeviceToHost, T->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(T->stream));
    memcpy(keep, hk, n);
    return ORB_OK;
}

}  // extern "C"
