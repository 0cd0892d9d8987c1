// liborbgpu -- ORB extractor for sm_100a.
//
// Replaces ORB_SLAM2::ORBextractor (ref: ORB_SLAM2/src/ORBextractor.cc, include/ORBextractor.h)
// behind the C ABI of include/orbgpu.h.  A batch of B independent frames runs through
//   k_pyr_level0 / k_pyr_resize   8-level padded pyramid        (ref ComputePyramid        :1047-1072)
//   k_fast_cells                  FAST-9 per 30-px cell, 20/7    (ref ComputeKeyPointsOctTree:735-796)
//   k_octree                      quadtree distribution          (ref DistributeOctTree      :525-733)
//   k_describe                    IC_Angle + 7x7 blur + rBRIEF   (ref :73-144, 1016-1044)
// with every intermediate resident in HBM/L2; the only host traffic is the source frames in and
// 60 B per keypoint out.  Integer arithmetic reproduces OpenCV's fixed point exactly; the few fp32
// steps use explicit round-to-nearest intrinsics so nothing is contracted into an FMA.
#include "orb_common.cuh"
#include "orb_internal.cuh"
#include "orb_devmath.cuh"
#include <cuda.h>
#include <ctime>                // CUtensorMap (the encoder is fetched with cudaGetDriverEntryPoint: no link against libcuda)
#include <nvtx3/nvToolsExt.h>   // header-only; the ranges cost a function-pointer test when no tool is attached
#include "orb_pattern.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace orbgpu {

static thread_local char g_err[512] = "";
void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// the 256 test pairs; lane l of a descriptor warp reads its own 32 bytes, so this lives in global memory
// (per-lane addresses would serialise in the constant cache)
__device__ __align__(16) int8_t d_pattern[ORB_PATTERN_INTS];
__constant__ int c_umax[16];
// IC_Angle as word loads: the 31x31 disc of a key point spans 9 aligned words on each of its 31 rows = 279 (row, word)
// tasks, 9 per lane.  Entry [a][t] (a = alignment of the disc's left edge, t = task) = {u-weights, v-weights (signed bytes,
// 0 outside the disc)}; the word's address is arithmetic (row t/9 - 15, word t%9) so pixel loads do not wait for the table;
// tasks 279..287 are no-ops (zero weights, clamped to the last real word).
constexpr int MOM_TASKS = 288;
__device__ uint2 d_moment_tab[4 * MOM_TASKS];
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {   // unsigned pixels x signed weights
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// ------------------------------------------------------------------------------------------
// Pyramid.  Each thread produces one aligned 32-bit word (4 pixels) of a padded plane row, so
// stores are always aligned although the interior starts 19 pixels into the plane.
// ------------------------------------------------------------------------------------------
// Level 0: the caller's image (any pitch, rows not aligned) into the padded plane.  Columns of 16 output bytes that lie
// strictly inside the image are copied by the first grid.x - 1 CTAs of a row group: five aligned source words shifted
// into place, one 16-byte store per thread.  The columns that touch the reflected frame or the zero margin up to the pitch
// (three on the left, a handful on the right) go to the last CTA, word by word and byte by byte, so that no warp of the
// copy carries the slow path.
constexpr int L0_ROWS = 4;
__global__ void __launch_bounds__(128)
k_pyr_level0(const uint8_t *__restrict__ src, size_t frame_stride, int spitch, uint8_t *__restrict__ pyr, const Geom *__restrict__ G) {
    const LevelGeom &g = G->lv[0];
    const int lane = threadIdx.x & 31, b = blockIdx.z;
    const int py = (blockIdx.y * (blockDim.x >> 5) + (threadIdx.x >> 5)) * L0_ROWS;
    const int rows = g.h + 2 * EDGE, nq = g.pitch / 16;
    const int qL = (XOFF + 16) / 16;                               // first column with x0 >= 16
    const int qR = max(qL, (g.w - 19 + XOFF) / 16 + 1);            // first column with x0 + 19 > w
    const uint8_t *sframe = src + (size_t)b * frame_stride;
    uint8_t *plane = pyr + (size_t)b * G->pyr_frame + g.plane_off;
    if (blockIdx.x + 1 < gridDim.x) {
        const int q = qL + blockIdx.x * 32 + lane;
        if (q >= qR) return;
        const int x0 = q * 16 - XOFF;
#pragma unroll
        for (int r = 0; r < L0_ROWS; r++) {
            if (py + r >= rows) break;
            const uint8_t *p = sframe + (size_t)reflect101(py + r - EDGE, g.h) * spitch + x0;
            const unsigned sh = ((unsigned)(uintptr_t)p & 3u) * 8u;
            const uint32_t *w = reinterpret_cast<const uint32_t *>(p - ((uintptr_t)p & 3u));
            const uint32_t w0 = __ldg(w), w1 = __ldg(w + 1), w2 = __ldg(w + 2), w3 = __ldg(w + 3), w4 = sh ? __ldg(w + 4) : 0u;
            uint4 v;
            v.x = __funnelshift_r(w0, w1, sh); v.y = __funnelshift_r(w1, w2, sh);
            v.z = __funnelshift_r(w2, w3, sh); v.w = __funnelshift_r(w3, w4, sh);
            *reinterpret_cast<uint4 *>(plane + (size_t)(py + r) * g.pitch + (size_t)q * 16) = v;
        }
    } else {
        const int nwords = (qL + nq - qR) * 4;                     // words of the columns [0, qL) and [qR, nq)
        for (int i = lane; i < nwords; i += 32) {
            const int word = i < qL * 4 ? i : qR * 4 + (i - qL * 4);
#pragma unroll
            for (int r = 0; r < L0_ROWS; r++) {
                if (py + r >= rows) break;
                const uint8_t *srow = sframe + (size_t)reflect101(py + r - EDGE, g.h) * spitch;
                uint32_t v = 0;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const int x = word * 4 + k - XOFF;
                    if (x >= -EDGE && x < g.w + EDGE) v |= (uint32_t)srow[reflect101(x, g.w)] << (8 * k);
                }
                *reinterpret_cast<uint32_t *>(plane + (size_t)(py + r) * g.pitch + (size_t)word * 4) = v;
            }
        }
    }
}

__global__ void __launch_bounds__(128)
k_pyr_resize_generic(uint8_t *__restrict__ pyr, const Geom *__restrict__ G, const ResizeTap *__restrict__ taps, int level) {
    // cv::resize INTER_LINEAR straight into the padded plane (reflect-101 frame included); any scale factor.
    // A thread makes 4 horizontally adjacent pixels.  Their source pixels span < 12 bytes, so each of the two
    // source rows is fetched as three aligned words and the taps pick bytes out of registers (PRMT) instead of
    // issuing 16 byte loads.
    const LevelGeom &g = G->lv[level];
    const LevelGeom &p = G->lv[level - 1];
    const int word = blockIdx.x * blockDim.x + threadIdx.x;
    const int py = blockIdx.y, b = blockIdx.z;
    if (word * 4 >= g.pitch) return;
    uint8_t *frame = pyr + (size_t)b * G->pyr_frame;
    const ResizeTap ty = taps[g.tap_y_off + reflect101(py - EDGE, g.h)];
    const uint8_t *s0 = frame + p.interior_off + (size_t)ty.s0 * p.pitch;
    const uint8_t *s1 = frame + p.interior_off + (size_t)ty.s1 * p.pitch;
    ResizeTap tx[4];
    int lo = 1 << 30, hi_s0 = 0;
    bool any = false;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int x = word * 4 + k - XOFF;
        const bool in = x >= -EDGE && x < g.w + EDGE;
        tx[k] = taps[g.tap_x_off + reflect101(in ? x : 0, g.w)];
        if (!in) { tx[k].a0 = 0; tx[k].a1 = 0; }
        else { lo = min(lo, (int)tx[k].s0); hi_s0 = max(hi_s0, (int)tx[k].s0); any = true; }
    }
    uint32_t v = 0;
    const int base = lo & ~3;   // interior rows start 32-byte aligned, so base is a word boundary
    if (any && hi_s0 - base > 7) {
        // scale factors well above 1.2: the four pixels do not fit the 12-byte window, plain byte loads
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const int r0v = s0[tx[k].s0] * tx[k].a0 + s0[tx[k].s1] * tx[k].a1;
            const int r1v = s1[tx[k].s0] * tx[k].a0 + s1[tx[k].s1] * tx[k].a1;
            const int o8 = (((ty.a0 * (r0v >> 4)) >> 16) + ((ty.a1 * (r1v >> 4)) >> 16) + 2) >> 2;
            v |= (uint32_t)o8 << (8 * k);
        }
    } else if (any) {
        const uint32_t *r0 = reinterpret_cast<const uint32_t *>(s0 + base), *r1 = reinterpret_cast<const uint32_t *>(s1 + base);
        const uint32_t a0 = __ldg(r0), a1 = __ldg(r0 + 1), a2 = __ldg(r0 + 2);
        const uint32_t c0 = __ldg(r1), c1 = __ldg(r1 + 1), c2 = __ldg(r1 + 2);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            // byte offsets of the two taps inside the 12-byte window (s1 == s0 only where its weight is 0)
            const int o = (tx[k].a0 | tx[k].a1) ? tx[k].s0 - base : 0;
            const bool hi = o >= 4;
            const uint32_t ua = hi ? a1 : a0, ub = hi ? a2 : a1, uc = hi ? c1 : c0, ud = hi ? c2 : c1;
            const uint32_t sel = (uint32_t)(o & 3) | ((uint32_t)((o & 3) + 1) << 4) | 0x4400u;   // bytes o, o+1
            const uint32_t pa = __byte_perm(ua, ub, sel), pc = __byte_perm(uc, ud, sel);
            const int r0v = (int)(pa & 0xff) * tx[k].a0 + (int)((pa >> 8) & 0xff) * tx[k].a1;
            const int r1v = (int)(pc & 0xff) * tx[k].a0 + (int)((pc >> 8) & 0xff) * tx[k].a1;
            const int o8 = (((ty.a0 * (r0v >> 4)) >> 16) + ((ty.a1 * (r1v >> 4)) >> 16) + 2) >> 2;
            v |= (uint32_t)o8 << (8 * k);
        }
    }
    *reinterpret_cast<uint32_t *>(frame + g.plane_off + (size_t)py * g.pitch + word * 4) = v;
}

// Fast path of the same resize (every level of the reference's 1.2 pyramid; any scale factor up to ~2).  The host
// precomputes, per aligned output word (4 pixels of the padded row, reflection and plane limits already resolved), two
// 128-bit records: the four weight pairs as 16x2 words {a0 | a1 << 16}, and four PRMT selectors that pull the four
// pixels' two source bytes out of a 12-byte source window plus the window's word index.  A thread owns one output word
// column and walks PYR_ROWS plane rows.  cv::resize interpolates horizontally first, and a source row serves the output
// row above and the one below, so the horizontally interpolated source rows are kept in registers across output rows:
// ~1.2 source rows (3 aligned loads, 4 PRMT, 4 IDP2A) per output row instead of 2.
constexpr int PYR_ROWS = 8;
// Row records (one uint4 per plane row, after the word records in the same pool): x = byte offset of source row s0 | bit 0: s0 is the
// previous plane row's s1 | bit 1: s1 == s0;  y = byte offset of source row s1;  z, w = the vertical weights << 16
// (umulhi(a << 16, h) == (a * h) >> 16).  Reflection, row pitch and the reuse decision are resolved on the host: the row loop
// holds no address arithmetic beyond two additions.  The two horizontally interpolated source rows live in X and Y with
// ALTERNATING roles -- row r reads (n0, n1) = (X, Y), row r + 1 reads (Y, X) -- so "n0 is the previous row's n1" (4 rows in 5 at
// scale 1.2) costs nothing, and no register is ever moved (the previous version spent 40 of its 124 instructions per row on the
// moves and compare chains of a two-entry cache: profiles/r2c_resize_sass_counts.txt).
template <int ROWS>
__device__ __forceinline__ void pyr_resize_column(uint8_t *frame, const Geom *__restrict__ G, const ResizeTap *__restrict__ taps,
                                                  const uint4 *__restrict__ wtab, int level, int word, int py0) {
    const LevelGeom &g = G->lv[level];
    const LevelGeom &p = G->lv[level - 1];
    const uint4 Wt = __ldg(wtab + 2 * (g.wtab_off + word)), St = __ldg(wtab + 2 * (g.wtab_off + word) + 1);
    const uint8_t *srcw = frame + p.interior_off + (size_t)St.z * 4;       // the window's first word in source row 0
    const int nr = min(ROWS, g.h + 2 * EDGE - py0);
    const uint4 *rt = wtab + g.rtab_off + py0;
    uint8_t *dst = frame + g.plane_off + (size_t)word * 4 + (size_t)py0 * g.pitch;
    const int dpitch = g.pitch;
    const uint32_t selA = St.x, selB = St.x >> 16, selC = St.y, selD = St.y >> 16;
    uint32_t X[4], Y[4];
    auto hrow = [&](uint32_t off, uint32_t *h) {   // (S[x]*a0 + S[x+1]*a1) >> 4 of the source row at byte offset `off` for the four pixels
        const uint32_t *r = reinterpret_cast<const uint32_t *>(srcw + off);
        const uint32_t a0 = __ldg(r), a1 = __ldg(r + 1), a2 = __ldg(r + 2);
        const uint32_t x01 = __byte_perm(__byte_perm(a0, a1, selA), a2, selB);
        const uint32_t x23 = __byte_perm(__byte_perm(a0, a1, selC), a2, selD);
        h[0] = __dp2a_lo(Wt.x, x01, 0u) >> 4; h[1] = __dp2a_hi(Wt.y, x01, 0u) >> 4;
        h[2] = __dp2a_lo(Wt.z, x23, 0u) >> 4; h[3] = __dp2a_hi(Wt.w, x23, 0u) >> 4;
    };
    auto emit = [&](const uint32_t *n0, const uint32_t *n1, uint32_t A0, uint32_t A1) {
        uint32_t v = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) v |= ((__umulhi(A0, n0[k]) + __umulhi(A1, n1[k]) + 2u) >> 2) << (8 * k);
        *reinterpret_cast<uint32_t *>(dst) = v;
        dst += dpitch;
    };
#pragma unroll 1
    for (int rr = 0; rr < ROWS; rr += 2) {
        if (rr >= nr) break;
        {   // even row of the strip: (n0, n1) = (X, Y)
            const uint4 t = __ldg(rt + rr);
            if (rr == 0 || !(t.x & 1u)) hrow(t.x & ~3u, X);
            if (t.x & 2u) { Y[0] = X[0]; Y[1] = X[1]; Y[2] = X[2]; Y[3] = X[3]; }
            else hrow(t.y, Y);
            emit(X, Y, t.z, t.w);
        }
        if (ROWS == 1 || rr + 1 >= nr) break;
        {   // odd row: (n0, n1) = (Y, X)
            const uint4 t = __ldg(rt + rr + 1);
            if (!(t.x & 1u)) hrow(t.x & ~3u, Y);
            if (t.x & 2u) { X[0] = Y[0]; X[1] = Y[1]; X[2] = Y[2]; X[3] = Y[3]; }
            else hrow(t.y, X);
            emit(Y, X, t.z, t.w);
        }
    }
}

template <int ROWS>   // plane rows per thread: PYR_ROWS for batches (source rows reused in registers); fewer for single frames, where
                      // the seven dependent launches are a latency chain and a thread's rows are its serial part
__global__ void __launch_bounds__(128)
k_pyr_resize(uint8_t *__restrict__ pyr, const Geom *__restrict__ G, const ResizeTap *__restrict__ taps,
             const uint4 *__restrict__ wtab, int level) {
    const int word = blockIdx.x * blockDim.x + threadIdx.x;
    if (word * 4 >= G->lv[level].pitch) return;
    pyr_resize_column<ROWS>(pyr + (size_t)blockIdx.z * G->pyr_frame, G, taps, wtab, level, word, blockIdx.y * ROWS);
}

// ------------------------------------------------------------------------------------------
// FAST-9/16 per cell.  One CTA per (cell, frame), one launch per level.  V = arc strength: corner at t
// <=> V > t, response V-1, 3x3 non-max suppression inside the cell's computed region only.
//
// Stages inside the CTA (all in dynamic shared memory sized for the level's cell):
//   1. the cell's ROI is staged as aligned 16-byte chunks (cp.async);
//   2. every pixel takes the 4-compass test at the current threshold: a 9-arc always holds two ADJACENT compass
//      points, so a corner needs one of {up, down} and one of {left, right} brighter (type H) or darker (type L)
//      than the centre by more than t.  A thread tests one aligned word = 4 pixels with 16x2 SIMD min/max
//      (VIMNMX.S16x2) and leaves a flag word; a block scan then compacts the H and the L survivors into two lists;
//   3. survivors get the exact arc strength on dense warps, two same-type survivors per thread in the two 16-bit
//      lanes (VIMNMX3.S16x2); a pixel on both lists takes the larger of the two;
//   4. threshold + NMS; what passes is appended to the level's list on the spot (warp-aggregated atomic).
// The cell runs at iniThFAST first and repeats 2-4 at minThFAST only when it came up empty (ref :775-784): a pixel
// that fails the compass test at t has V <= t, so it can neither be a corner at t nor suppress one.
// ------------------------------------------------------------------------------------------
constexpr int FAST_T = FAST_THREADS;
constexpr int FT_PW = 20;                 // tile row pitch in words: 80 bytes >= 65 (cell < 60, + 6) + 15 bytes of alignment
constexpr int FT_PB = FT_PW * 4;          // slack; compile-time so that the 16 circle offsets are immediates

struct FastLevel {   // what k_fast_cells needs of one level, passed by value (constant bank)
    int level, nlevels, ini_th, min_th;
    int pitch, maxBX, maxBY, wCell, hCell, cand_off, cand_cap, cand_frame;
    int off_vmap, off_flags, off_survH, off_survL;   // byte offsets into dynamic shared memory (16-aligned)
    long long interior_off, pyr_frame;
    int w, h, ncx, ncy;   // level size and cell grid (k_fast_rows)
};

// magic[d] = floor((2^32 - 1) / d) + 1: n / d == umulhi(n, magic[d]) for the small n used here (d = 0, 1 unused)
struct FastMagic { uint32_t m[80]; };
static constexpr FastMagic make_fast_magic() {
    FastMagic t{};
    for (int d = 2; d < 80; d++) t.m[d] = 0xFFFFFFFFu / (uint32_t)d + 1u;
    return t;
}
__constant__ FastMagic c_fast_magic = make_fast_magic();

struct FastDiv {   // exact n / d for n < 2^32 / d via one multiply-high
    uint32_t M;
    int d;
    __device__ __forceinline__ explicit FastDiv(int dd) : M(c_fast_magic.m[dd]), d(dd) {}
    __device__ __forceinline__ int div(int n) const { return d > 1 ? (int)__umulhi((uint32_t)n, M) : n; }
};

static size_t fast_smem_layout(FastLevel &fl) {
    const int tw = fl.wCell + 6, th = fl.hCell + 6;
    (void)tw;
    const int nwc = (fl.wCell + 3) / 4 + 1;             // words per computed row, worst alignment
    size_t off = (size_t)th * FT_PB;
    auto carve = [&](size_t bytes) { off = (off + 15) & ~(size_t)15; const size_t o = off; off += bytes; return (int)o; };
    fl.off_vmap = carve((size_t)(fl.hCell + 3) * FT_PB);   // score map: rows -1 .. ch of the computed region, tile column geometry
    fl.off_flags = carve((size_t)nwc * fl.hCell * 4);   // flagged words of the compass pass, expanded into the lists afterwards
    (void)nwc;
    fl.off_survH = carve((size_t)fl.wCell * fl.hCell * 2);   // one-type survivors: brighter from the bottom, darker from the top
    fl.off_survL = carve((size_t)fl.wCell * fl.hCell * 2);   // survivors of both types
    return (off + 15) & ~(size_t)15;
}

// Two same-type survivors in the 16-bit lanes.  Lane value 256 + s*(I(p) - I(q_k)), s = +1 (brighter centre) or -1:
// always in [1, 511], so packed adds never carry between lanes.  Returns 256 + V per lane.  The sign is a run-time
// multiplier (sgn = 1 or -1 as a 32-bit word: q * -1 + base is base - q in packed arithmetic as well), so brighter-
// and darker-type pairs share one instruction stream and the multiplies go to the FMA pipe.
template <int PB = FT_PB>
__device__ __forceinline__ uint32_t fast_arc_pair(const uint8_t *p0, const uint8_t *p1, const uint32_t sgn) {
    const uint32_t cp = __byte_perm(p0[0], p1[0], 0x5410);
    const uint32_t base = 0x01000100u + cp * sgn, nsgn = 0u - sgn;
    const int off[16] = {3 * PB, 3 * PB + 1, 2 * PB + 2, PB + 3, 3, -PB + 3, -2 * PB + 2, -3 * PB + 1,
                         -3 * PB, -3 * PB - 1, -2 * PB - 2, -PB - 3, -3, PB - 3, 2 * PB - 2, 3 * PB - 1};
    uint32_t e[16];
#pragma unroll
    for (int k = 0; k < 16; k++) {
        const uint32_t q = __byte_perm(p0[off[k]], p1[off[k]], 0x5410);
        e[k] = q * nsgn + base;
    }
    uint32_t m3[16];
#pragma unroll
    for (int k = 0; k < 16; k++) m3[k] = __vimin3_s16x2(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
    uint32_t a = 0;
#pragma unroll
    for (int k = 0; k < 16; k += 2)
        a = __vimax3_s16x2(a, __vimin3_s16x2(m3[k], m3[(k + 3) & 15], m3[(k + 6) & 15]),
                           __vimin3_s16x2(m3[k + 1], m3[(k + 4) & 15], m3[(k + 7) & 15]));
    return a;
}

// out-of-line copy for the rare both-type pixels (a second inlined copy doubles the kernel's register count)
__device__ __noinline__ uint32_t fast_arc_pair_call(const uint8_t *p, const uint32_t sgn) { return fast_arc_pair(p, p, sgn); }

// One cell, start to finish, by the whole CTA; `thr0` is the threshold of the first pass (iniThFAST, or minThFAST when the
// caller already knows the cell is empty at iniThFAST).
template <int NT>
__device__ __forceinline__ void fast_cell(const uint8_t *__restrict__ pyr, const FastLevel &g, uint32_t *__restrict__ cand,
                                          uint32_t *__restrict__ cand_count, int *__restrict__ err, const int b, const int ci,
                                          const int cj, const int thr0) {
    extern __shared__ __align__(16) unsigned char fsm[];
    __shared__ int s_nkp, s_cnt, s_nB, s_nW;

    uint32_t *tileW = reinterpret_cast<uint32_t *>(fsm);
    uint32_t *vmapW = reinterpret_cast<uint32_t *>(fsm + g.off_vmap);   // scores; row r <-> computed row r - 1, tile columns
    uint16_t *surv = reinterpret_cast<uint16_t *>(fsm + g.off_survH);   // code = y << 7 | tile byte column; brighter-type
    uint16_t *survB = reinterpret_cast<uint16_t *>(fsm + g.off_survL);  // from 0 up, darker-type from the top down; both-type
    const int listTop = g.wCell * g.hCell - 1;
    uint32_t *wlist = reinterpret_cast<uint32_t *>(fsm + g.off_flags);   // code0 | brighter nibble << 16 | darker nibble << 20

    const int level = g.level;
    const int minB = EDGE - 3;
    // ROI handed to cv::FAST by the reference (ref :759-773)
    const int iniX = minB + cj * g.wCell, iniY = minB + ci * g.hCell;
    const int maxX = min(iniX + g.wCell + 6, g.maxBX), maxY = min(iniY + g.hCell + 6, g.maxBY);
    const int tw = maxX - iniX, th = maxY - iniY;      // ROI size
    const int cw = tw - 6, ch = th - 6;                // computed region (ROI minus 3 px rim)
    if (cw <= 0 || ch <= 0) return;
    const int tid = threadIdx.x, lane = tid & 31;
    constexpr int TPW = FT_PW, TPB = FT_PB;

    // ---- 1. stage the ROI; byte (y, x) of the ROI sits at tile byte y * TPB + x + sft
    const int sft = (iniX + XOFF) & 15;   // plane rows are 128-byte aligned and interior column 0 sits at byte XOFF
    {
        const int nQ = (tw + sft + 15) >> 4;   // <= 5 chunks of 16 bytes per row
        // thread -> (chunk q, row): 4 chunk slots per row (8 for the rare 5-chunk rows), rows NT / slots apart
        const int lg = nQ > 4 ? 3 : 2;
        const int q = tid & ((1 << lg) - 1), ystep = NT >> lg;
        if (q < nQ) {
            int y = tid >> lg;
            const uint8_t *src = pyr + (size_t)b * g.pyr_frame + g.interior_off + (ptrdiff_t)(iniY + y) * g.pitch + (iniX - sft) + q * 16;
            uint32_t dst = (uint32_t)__cvta_generic_to_shared(tileW) + (uint32_t)(y * TPB + q * 16);
            const size_t sstep = (size_t)ystep * g.pitch;
            for (; y < th; y += ystep, src += sstep, dst += ystep * TPB)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
        }
        asm volatile("cp.async.commit_group;");
        // score rows -1 and ch (the NMS apron); the rows between are cleared by the compass pass below
        if (tid < 2 * TPW) vmapW[tid < TPW ? tid : (ch + 1) * TPW + tid - TPW] = 0;
        if (tid == 0) { s_nkp = 0; s_cnt = 0; s_nB = 0; s_nW = 0; }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
    const uint8_t *tile = reinterpret_cast<const uint8_t *>(tileW);
    uint8_t *vmap = reinterpret_cast<uint8_t *>(vmapW);

    // words of a tile row that hold computed pixels: bytes [3 + sft, 3 + sft + cw)
    const int w0 = (3 + sft) >> 2, nWc = ((2 + sft + cw) >> 2) - w0 + 1;
    // thread -> (word column w, row group rg): nRg = NT / nWc row groups, the few threads beyond nRg * nWc idle in stage 2
    const FastDiv dU(nWc);
    const int nRg = dU.div(NT);
    const int rg0 = dU.div(tid), w = tid - rg0 * nWc + w0;
    const int rg = rg0 < nRg ? rg0 : ch;   // idle threads start past the last row
    uint32_t vm;                           // bytes of this column's words that are computed pixels: [lo, hi)
    {
        const int lo = 3 + sft - 4 * w, hi = lo + cw;
        vm = (0x80808080u << (8 * max(lo, 0))) & (0x80808080u >> (8 * max(4 - hi, 0)));
    }
    // the first / last word column also clears the score word beside it (the apron column may sit there)
    const bool edgeL = w == w0 && w0 > 0, edgeR = w == w0 + nWc - 1;

    int thr = thr0, nH = 0, nL = 0;
    for (;;) {
        // ---- 2. compass test at thr, 4 pixels per thread and step.  Lanes: A = (px0, px2), B = (px1, px3).
        //   H <=> c - max(min(U, D), min(L, R)) > t        L <=> min(max(U, D), max(L, R)) - c > t
        //   A word with survivors takes its ranges of the lists with one shared-memory atomic (both counters in one word,
        //   H low / L high) and writes the codes on the spot; the order of the lists is irrelevant.  The same pass zeroes
        //   the score map word by word.
        {
            const uint32_t tb = (uint32_t)thr * 0x00010001u + 0x80008000u;
            // thread = (word column, row group): the column's validity mask and the row pointer step are loop invariants
            for (int yy = rg; yy < ch; yy += nRg) {
                const uint32_t *r = tileW + (yy + 3) * TPW + w;
                const uint32_t C = r[0], P = r[-1], N = r[1], U = r[-3 * TPW], D = r[3 * TPW];
                uint32_t *vz = vmapW + (yy + 1) * TPW + w;
                vz[0] = 0;
                if (edgeL) vz[-1] = 0;
                if (edgeR) vz[1] = 0;
                const uint32_t cA = __byte_perm(C, 0, 0x4240), cB = __byte_perm(C, 0, 0x4341);
                const uint32_t uA = __byte_perm(U, 0, 0x4240), uB = __byte_perm(U, 0, 0x4341);
                const uint32_t dA = __byte_perm(D, 0, 0x4240), dB = __byte_perm(D, 0, 0x4341);
                const uint32_t lA = __byte_perm(P, 0, 0x4341), lB = __byte_perm(P, C, 0x0402) & 0x00ff00ffu;
                const uint32_t rA = __byte_perm(C, N, 0x0503) & 0x00ff00ffu, rB = __byte_perm(N, 0, 0x4240);
                const uint32_t mA = __vmaxs2(__vmins2(uA, dA), __vmins2(lA, rA)), mB = __vmaxs2(__vmins2(uB, dB), __vmins2(lB, rB));
                const uint32_t nA = __vmins2(__vmaxs2(uA, dA), __vmaxs2(lA, rA)), nB = __vmins2(__vmaxs2(uB, dB), __vmaxs2(lB, rB));
                // lane bit 15 set <=> NOT a survivor (no carries between lanes: every lane stays in [0x7f01, 0x81fe])
                const uint32_t xA = mA + tb - cA, xB = mB + tb - cB, yA = cA + tb - nA, yB = cB + tb - nB;
                const uint32_t Hb = __byte_perm(xA, xB, 0x7351), Lb = __byte_perm(yA, yB, 0x7351);   // byte j <- pixel j
                const uint32_t fH = ~Hb & vm, fL = ~Lb & vm;
                if (fH | fL)   // bit 7 of byte j -> nibble bit j (the four partial products land on distinct bits: no carries)
                    wlist[atomicAdd(&s_nW, 1)] = (uint32_t)((yy << 7) + 4 * w) | ((fH * 0x00204081u) >> 28) << 16 | ((fL * 0x00204081u) >> 28) << 20;
            }
        }
        __syncthreads();
        // ---- 2b. flagged words -> survivor lists, one word per thread (dense lanes; at compass time ~4 lanes of a warp hold one)
        for (int t = tid, nW = s_nW; t < nW; t += NT) {
            const uint32_t e = wlist[t];
            const int code0 = (int)(e & 0xffff);
            const uint32_t nH4 = (e >> 16) & 15u, nL4 = (e >> 20) & 15u, fB = nH4 & nL4, f1 = (nH4 | nL4) ^ fB, oL = nL4 & f1;
            if (f1) {   // one list array: brighter-type codes grow from the bottom, darker-type from the top
                const int pos = atomicAdd(&s_cnt, __popc(f1 ^ oL) + (__popc(oL) << 16));
                int pH = pos & 0xffff, pL = listTop - (pos >> 16);
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if ((f1 >> j) & 1) {
                        const int isL = (oL >> j) & 1;
                        surv[isL ? pL : pH] = (uint16_t)(code0 + j);
                        pH += 1 - isL; pL -= isL;
                    }
            }
            if (fB) {   // rare: brighter than one compass pair and darker than the other
                int pB = atomicAdd(&s_nB, __popc(fB));
#pragma unroll
                for (int j = 0; j < 4; j++)
                    if ((fB >> j) & 1) survB[pB++] = (uint16_t)(code0 + j);
            }
        }
        __syncthreads();
        nH = s_cnt & 0xffff; nL = s_cnt >> 16;
        const int nB = s_nB;

        // ---- 3. exact arc strength of the survivors in ONE pass: pair slots of the brighter-type list, then pair slots of
        // the darker-type list (two same-type survivors per thread, the type is a run-time sign), then the both-type pixels
        // one per thread.  Every pixel is on exactly one list, so every score byte has one writer.
        {
            const int nPH = (nH + 1) >> 1, nPP = nPH + ((nL + 1) >> 1);
            for (int s = tid; s < nPP + nB; s += NT) {
                int c0, c1;
                uint32_t sgn = 1u;
                if (s < nPH) { const int i0 = 2 * s; c0 = surv[i0]; c1 = i0 + 1 < nH ? surv[i0 + 1] : c0; }
                else if (s < nPP) { const int i0 = 2 * (s - nPH); sgn = 0xffffffffu; c0 = surv[listTop - i0]; c1 = i0 + 1 < nL ? surv[listTop - i0 - 1] : c0; }
                else { c0 = c1 = survB[s - nPP]; }   // both types: the same pixel in both lanes, once per sign
                const int o0 = (c0 >> 7) * TPB + (c0 & 127), o1 = (c1 >> 7) * TPB + (c1 & 127);
                uint32_t a = fast_arc_pair(tile + 3 * TPB + o0, tile + 3 * TPB + o1, sgn);
                if (s >= nPP) a = __vmaxs2(a, fast_arc_pair_call(tile + 3 * TPB + o0, 0xffffffffu));
                vmap[TPB + o0] = (uint8_t)max((int)(a & 0xffff) - 256, 0);
                vmap[TPB + o1] = (uint8_t)max((int)(a >> 16) - 256, 0);
            }
        }
        __syncthreads();

        // ---- 4. threshold + NMS over the lists.  Whatever passes is final -- a cell is only repeated when NOTHING passed --
        // so it goes straight to the level's candidate list: one global atomic per warp and step for the range,
        // window-relative coordinates (ref :789-790: ROI x + j*wCell).
        bool found = false;
        const int nS = nH + nL + nB;
        for (int s0 = 0; s0 < nS; s0 += NT) {
            const int s = s0 + tid;
            bool pass = false;
            int code = 0, V = 0;
            if (s < nS) {
                code = s < nH ? surv[s] : s < nH + nL ? surv[listTop - (s - nH)] : survB[s - nH - nL];
                const uint8_t *v = &vmap[((code >> 7) + 1) * TPB + (code & 127)];
                V = v[0];
                pass = V > thr && V > v[-1] && V > v[1] && V > v[-TPB - 1] && V > v[-TPB] && V > v[-TPB + 1] &&
                       V > v[TPB - 1] && V > v[TPB] && V > v[TPB + 1];
            }
            const unsigned bal = __ballot_sync(0xffffffffu, pass);
            if (bal) {
                found = true;
                int base = 0;
                if (lane == 0) base = (int)atomicAdd(cand_count + (size_t)b * g.nlevels + level, (uint32_t)__popc(bal));
                base = __shfl_sync(0xffffffffu, base, 0);
                if (pass) {
                    const int slot = base + __popc(bal & ((1u << lane) - 1u));
                    if (slot < g.cand_cap) {
                        const uint32_t wx = (code & 127) - sft + cj * g.wCell, wy = (code >> 7) + 3 + ci * g.hCell;
                        cand[(size_t)b * g.cand_frame + g.cand_off + slot] = wx | (wy << 12) | ((uint32_t)(V - 1) << 24);
                    } else {
                        atomicOr(err, ERRF_CAND_OVERFLOW);
                    }
                }
            }
        }
        if (found && lane == 0) s_nkp = 1;
        if (thr == g.min_th) break;
        __syncthreads();
        if (s_nkp) break;
        thr = g.min_th;   // empty cell: once more at the lower threshold (ref :780-784)
        if (tid == 0) { s_cnt = 0; s_nB = 0; s_nW = 0; }
        __syncthreads();
    }
}


__global__ void __launch_bounds__(FAST_T)
k_fast_cells(const uint8_t *__restrict__ pyr, const FastLevel g, uint32_t *__restrict__ cand, uint32_t *__restrict__ cand_count,
             int *__restrict__ err) {
    fast_cell<FAST_T>(pyr, g, cand, cand_count, err, blockIdx.z, blockIdx.y, blockIdx.x, g.ini_th);
}

// ------------------------------------------------------------------------------------------
// FAST for large batches: one CTA per (segment of a cell row, frame).  A segment is up to FR_COLS aligned word columns
// (<= 512 pixels) of whole cells; its borders are cell borders, so the 3x3 suppression never looks outside the CTA's tile
// (a neighbour outside the cell counts as 0, ref :759-790) and the cells only show up as column masks.  Same stages as
// fast_cell, with the per-cell fixed costs paid once per segment and a 3-px halo around ~14 cells instead of around each:
//   1. the tile (segment + halo) is staged as 16-byte chunks (cp.async);
//   2. compass test at iniThFAST: a thread owns one word column of one half of the rows and walks down with the rows
//      y - 3 .. y + 3 unpacked in registers (a row is unpacked once and serves as "down", "centre" and "up"; all shared
//      memory offsets are immediates).  The words of a warp's row that hold survivors go to a dense list with one
//      shared-memory atomic per warp and row (ballot + prefix);
//   3. flagged words -> the two survivor lists, one word per thread on dense lanes;
//   4. exact arc strength, one thread per pair of same-type survivors; 5. threshold + suppression per survivor, candidates
//      appended with one global atomic per warp; 6. the columns of cells that stayed empty repeat 2-5 at minThFAST on the
//      tile that is already staged.
// A segment with more survivors than the lists hold (dense noise) runs its cells through fast_cell instead.
// ------------------------------------------------------------------------------------------
constexpr int FR_LOGC = 5;
constexpr int FR_COLS = 1 << FR_LOGC;   // word columns of a segment
#ifndef ORBGPU_FR_PARTS
#define ORBGPU_FR_PARTS 2
#endif
constexpr int FR_PARTS = ORBGPU_FR_PARTS;   // row parts per word column (3 parts = 27 instead of 22 resident warps per SM but 16 % more compass
                                            // warm-up rows: measured 0.880 vs 0.870 ms per 288 C2 frames, so 2 stays)
constexpr int FR_T = FR_PARTS * FR_COLS;   // threads: FR_COLS word columns x FR_PARTS row parts
constexpr int FR_TPW = FR_COLS + 8;     // tile pitch in words: the columns + one word either side + 16-byte alignment slack
constexpr int FR_TPB = FR_TPW * 4;
constexpr int FR_ROWS = (60 + FR_PARTS - 1) / FR_PARTS;   // rows per thread (cells are < 60 px high)

struct FastRowsLayout { int off_score, off_slots, off_list, list_cap, seg_cells, use_tma, frame0; };   // byte offsets into dynamic shared memory

__device__ __noinline__ uint32_t fast_arc_pair_rows_call(const uint8_t *p0, const uint8_t *p1, const uint32_t sgn) {
    return fast_arc_pair<FR_TPB>(p0, p1, sgn);
}

// (out of line: the per-cell routine inlined twice drove k_fast_rows to 101 registers)
__device__ __noinline__ void fast_cell_rows_call(const uint8_t *pyr, const FastLevel &g, uint32_t *cand, uint32_t *cand_count, int *err,
                                                 int b, int ci, int cj, int thr) {
    fast_cell<FR_T>(pyr, g, cand, cand_count, err, b, ci, cj, thr);
}

__global__ void __launch_bounds__(FR_T, FR_PARTS == 2 ? 12 : 9)
k_fast_rows(const uint8_t *__restrict__ pyr, const FastLevel g, const FastRowsLayout lay, const __grid_constant__ CUtensorMap tmap,
            uint32_t *__restrict__ cand, uint32_t *__restrict__ cand_count, int *__restrict__ err) {
    extern __shared__ __align__(128) unsigned char fsm[];
    __shared__ int s_found[32], s_wcnt[FR_T / 32], s_cnt, s_over;
    __shared__ __align__(8) unsigned long long s_mbar;
    uint32_t *tileW = reinterpret_cast<uint32_t *>(fsm);
    uint32_t *scoreW = reinterpret_cast<uint32_t *>(fsm + lay.off_score);    // row r + 1 <-> computed row r, tile columns
    // flagged words: type bits of the four pixels (brighter 0 / 8 / 16 / 24, darker 1 / 9 / 17 / 25) | row << 2 | column (low six
    // bits << 10, bit 6 << 18); as many entries as the segment has words
    uint32_t *wlist = reinterpret_cast<uint32_t *>(fsm + lay.off_slots);
    uint16_t *list = reinterpret_cast<uint16_t *>(fsm + lay.off_list);       // x | row << 9 | both-type << 15; brighter up, darker down
    const int b = blockIdx.z, ci = blockIdx.y, cj0 = blockIdx.x * lay.seg_cells, tid = threadIdx.x, lane = tid & 31;
    const int ncell = min(lay.seg_cells, g.ncx - cj0);
    const int y0 = EDGE + ci * g.hCell, hc = min(g.hCell, g.h - EDGE - y0);                    // computed rows [y0, y0 + hc)
    const int xs = EDGE + cj0 * g.wCell, xe = min(xs + ncell * g.wCell, g.w - EDGE), rw = xe - xs;   // computed columns [xs, xe)
    if (hc <= 0 || rw <= 0) return;
    const int w0 = (XOFF + xs) >> 2, wl = (XOFF + xe - 1) >> 2, nW = wl - w0 + 1;              // plane words with computed pixels
    const int cbase = (w0 - 1) & ~3;                                                            // tile word 0 <-> plane word cbase
    const int lead = XOFF + xs - 4 * w0, xoff = 4 * (w0 - cbase);                               // list x = tile byte column - xoff
    // ---- 1. stage tile rows y0 - 3 .. y0 + hc + 2, words cbase .. cbase + FR_TPW - 1.  With a tensor map of the level (use_tma):
    // ONE 3-D tiled TMA load (cp.async.bulk.tensor, box = FR_TPB bytes x hCell + 6 rows x 1 frame; what lies outside the plane
    // arrives as zeros) issued by one thread, completion counted in bytes on an mbarrier -- no thread moves data.  Otherwise
    // 16-byte cp.async chunks, thread = (chunk of a row, row group).
    if (lay.use_tma) {
        const uint32_t mbar = (uint32_t)__cvta_generic_to_shared(&s_mbar);
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"((uint32_t)(FR_TPB * (g.hCell + 6))) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(
                             (uint32_t)__cvta_generic_to_shared(tileW)),
                         "l"(&tmap), "r"(4 * cbase), "r"(EDGE + y0 - 3), "r"(b + lay.frame0), "r"(mbar)
                         : "memory");
        }
        for (int i = tid; i < 2 * FR_TPW; i += FR_T) scoreW[i < FR_TPW ? i : (hc + 1) * FR_TPW + i - FR_TPW] = 0;   // apron rows -1 and hc
        if (tid < 32) s_found[tid] = 0;
        if (tid == 0) { s_over = 0; s_cnt = 0; }
        __syncthreads();   // the barrier's initialisation is visible to the waiting threads
        uint32_t done;
        do {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(mbar) : "memory");
        } while (!done);
    } else {
        const int nQ = ((wl + 1) >> 2) - (cbase >> 2) + 1;     // <= FR_COLS / 4 + 2 chunks per row
        const int lg = nQ > 32 ? 6 : nQ > 16 ? 5 : 4, q = tid & ((1 << lg) - 1), rstep = FR_T >> lg;
        if (q < nQ) {
            int r = tid >> lg;
            const uint8_t *src = pyr + (size_t)b * g.pyr_frame + g.interior_off - XOFF + (ptrdiff_t)(y0 - 3 + r) * g.pitch +
                                 (size_t)cbase * 4 + q * 16;
            uint32_t dst = (uint32_t)__cvta_generic_to_shared(tileW) + (uint32_t)(r * FR_TPB + q * 16);
            const size_t sstep = (size_t)rstep * g.pitch;
            for (; r < hc + 6; r += rstep, src += sstep, dst += rstep * FR_TPB)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
        }
        asm volatile("cp.async.commit_group;");
        for (int i = tid; i < 2 * FR_TPW; i += FR_T) scoreW[i < FR_TPW ? i : (hc + 1) * FR_TPW + i - FR_TPW] = 0;   // apron rows -1 and hc
        if (tid < 32) s_found[tid] = 0;
        if (tid == 0) { s_over = 0; s_cnt = 0; }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
    }
    const uint8_t *tile = reinterpret_cast<const uint8_t *>(tileW);
    uint8_t *score = reinterpret_cast<uint8_t *>(scoreW);
    const FastDiv dW(g.wCell);

    // thread = (word column, half of the rows); the rows of a warp are uniform (columns past the segment run masked)
    const int col = tid & (FR_COLS - 1), half = (hc + FR_PARTS - 1) / FR_PARTS, r0 = (tid >> FR_LOGC) * half;   // `half`: rows of one part
    const int nrows = min(half, hc - r0);
    const int tw = w0 + col - cbase;
    uint32_t vm;                          // bytes of this thread's words that are computed pixels (of cells still to do)
    {
        const int lo = lead - 4 * col, hi = lo + rw;
        vm = (0x80808080u << (8 * max(lo, 0))) & (0x80808080u >> (8 * max(4 - hi, 0)));
        if (col >= nW) vm = 0;
    }
    const uint32_t slot_tag = ((uint32_t)(col & 63) << 10) | ((uint32_t)(col >> 6) << 18) | ((uint32_t)r0 << 2);
    const uint32_t lt_mask = (1u << lane) - 1u;
    const int wcap = 32 * half;                       // flagged words a warp can produce: its 32 columns x its rows
    const int top = lay.list_cap - 1;

    for (int thr = g.ini_th;;) {
        // ---- 2. compass test at thr: the column walks down with the rows y - 3 .. y + 3 unpacked in registers; the flagged
        // words of a warp's row go to the dense list with one shared-memory atomic (ballot + prefix)
        int wcount = 0;
        uint32_t *wreg = wlist + (tid >> 5) * wcap;
        if (nrows > 0 && __any_sync(0xffffffffu, vm != 0)) {
            const uint32_t tb = (uint32_t)thr * 0x00010001u + 0x80008000u;
            const uint32_t *tp = tileW + r0 * FR_TPW + tw;      // tile row t <-> image row y0 - 3 + t: computed row r looks at rows r .. r + 6
            uint32_t *sp = scoreW + (r0 + 1) * FR_TPW + tw;
            uint32_t A[7], B[7], Pq[4], Nq[4];
#pragma unroll
            for (int i = 0; i < 6; i++) {
                const uint32_t C = tp[i * FR_TPW];
                A[i] = __byte_perm(C, 0, 0x4240); B[i] = __byte_perm(C, 0, 0x4341);
                if (i >= 3) { Pq[i & 3] = tp[i * FR_TPW - 1]; Nq[i & 3] = tp[i * FR_TPW + 1]; }
            }
#pragma unroll
            for (int k = 0; k < FR_ROWS; k++) {
                if (k >= nrows) break;
                const uint32_t Cn = tp[(k + 6) * FR_TPW];
                Pq[(k + 6) & 3] = tp[(k + 6) * FR_TPW - 1];
                Nq[(k + 6) & 3] = tp[(k + 6) * FR_TPW + 1];
                A[(k + 6) % 7] = __byte_perm(Cn, 0, 0x4240); B[(k + 6) % 7] = __byte_perm(Cn, 0, 0x4341);
                const uint32_t P = Pq[(k + 3) & 3], N = Nq[(k + 3) & 3];
                const uint32_t cA = A[(k + 3) % 7], cB = B[(k + 3) % 7], uA = A[k % 7], uB = B[k % 7], dA = A[(k + 6) % 7], dB = B[(k + 6) % 7];
                // left = pixels x - 3, right = x + 3: (P1, P3) (P2, C0) and (C3, N1) (N0, N2); byte 1 of an unpacked word is zero
                const uint32_t lA = __byte_perm(P, 0, 0x4341), lB = __byte_perm(P, cA, 0x5452);
                const uint32_t rA = __byte_perm(cB, N, 0x1512), rB = __byte_perm(N, 0, 0x4240);
                const uint32_t mA = __vmaxs2(__vmins2(uA, dA), __vmins2(lA, rA)), mB = __vmaxs2(__vmins2(uB, dB), __vmins2(lB, rB));
                const uint32_t nA = __vmins2(__vmaxs2(uA, dA), __vmaxs2(lA, rA)), nB = __vmins2(__vmaxs2(uB, dB), __vmaxs2(lB, rB));
                // lane bit 15 set <=> NOT a survivor (no carries between lanes: every lane stays in [0x7f01, 0x81fe])
                const uint32_t xA = mA + tb - cA, xB = mB + tb - cB, yA = cA + tb - nA, yB = cB + tb - nB;
                const uint32_t Hb = __byte_perm(xA, xB, 0x7351), Lb = __byte_perm(yA, yB, 0x7351);   // byte j <- pixel j
                const uint32_t fH = ~Hb & vm, fL = ~Lb & vm;
                if (vm) sp[k * FR_TPW] = 0;
                const bool f = (fH | fL) != 0;
                const unsigned bal = __ballot_sync(0xffffffffu, f);
                if (f) wreg[wcount + __popc(bal & lt_mask)] = ((fH >> 7) | (fL >> 6)) + (slot_tag + 4u * k);
                wcount += __popc(bal);   // warp-uniform: the warp appends to its own region of the list, no atomic
            }
        }
        if (lane == 0) s_wcnt[tid >> 5] = wcount;
        __syncthreads();
        int nFw = 0;
#pragma unroll
        for (int q = 0; q < FR_T / 32; q++) nFw += s_wcnt[q];
        // ---- 3b. flagged words -> survivor lists on dense lanes.  A pixel of both types goes on the brighter-type list, bit 15 set.
        {
            for (int t = tid; t < nFw; t += FR_T) {
                int u = t, q = 0;                      // entry t of the warps' regions taken one after the other
                while (u >= s_wcnt[q]) { u -= s_wcnt[q]; q++; }
                const uint32_t e = wlist[q * wcap + u], hm = e & 0x01010101u, lm = (e >> 1) & 0x01010101u, lo = lm & ~hm;
                const int p = atomicAdd(&s_cnt, __popc(hm) + (__popc(lo) << 16));
                int ph = p & 0xffff, pl = p >> 16;
                if (ph + __popc(hm) + pl + __popc(lo) > lay.list_cap) { s_over = 1; continue; }   // (a bound: both counts only grow)
                pl = top - pl;
                const uint32_t code0 = (((e >> 10) & 63u) | ((e >> 12) & 64u)) * 4u | ((e >> 2) & 63u) << 9;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if ((hm >> (8 * j)) & 1u) list[ph++] = (uint16_t)(code0 + j + (((lm >> (8 * j)) & 1u) << 15));
                    else if ((lm >> (8 * j)) & 1u) list[pl--] = (uint16_t)(code0 + j);
                }
            }
        }
        __syncthreads();
        const int nH = s_cnt & 0xffff, nL = s_cnt >> 16;
        if (s_over) {
            // dense frame: more survivors than the lists hold -> the cells concerned one by one through the per-cell routine
            for (int c = 0; c < ncell; c++) {
                if (thr != g.ini_th && s_found[c]) continue;
                __syncthreads();
                fast_cell_rows_call(pyr, g, cand, cand_count, err, b, ci, cj0 + c, thr);
            }
            return;
        }
        // ---- 4. exact arc strength, a thread per pair of same-type survivors
        {
            const int nPH = (nH + 1) >> 1, nPP = nPH + ((nL + 1) >> 1);
            for (int s = tid; s < nPP; s += FR_T) {
                uint32_t c0, c1, sgn = 1u;
                if (s < nPH) { const int i0 = 2 * s; c0 = list[i0]; c1 = i0 + 1 < nH ? list[i0 + 1] : c0; }
                else { const int i0 = 2 * (s - nPH); sgn = 0xffffffffu; c0 = list[top - i0]; c1 = i0 + 1 < nL ? list[top - i0 - 1] : c0; }
                const int o0 = (int)((c0 >> 9) & 63u) * FR_TPB + (int)(c0 & 511u) + xoff, o1 = (int)((c1 >> 9) & 63u) * FR_TPB + (int)(c1 & 511u) + xoff;
                uint32_t a = fast_arc_pair<FR_TPB>(tile + 3 * FR_TPB + o0, tile + 3 * FR_TPB + o1, sgn);
                if ((c0 | c1) >> 15) {   // rare: brighter than one compass pair and darker than the other -> the larger strength
                    const uint32_t a2 = fast_arc_pair_rows_call(tile + 3 * FR_TPB + o0, tile + 3 * FR_TPB + o1, 0xffffffffu);
                    a = __vmaxs2(a, a2 & (((c0 >> 15) ? 0xffffu : 0u) | ((c1 >> 15) ? 0xffff0000u : 0u)));
                }
                score[FR_TPB + o0] = (uint8_t)max((int)(a & 0xffff) - 256, 0);
                score[FR_TPB + o1] = (uint8_t)max((int)(a >> 16) - 256, 0);
            }
        }
        __syncthreads();
        // ---- 5. threshold + 3x3 suppression inside the survivor's cell; candidates straight to the level's list
        {
            const int n = nH + nL;
            for (int s0 = 0; s0 < n; s0 += FR_T) {
                const int s = s0 + tid;
                bool pass = false;
                int px = 0, row = 0, V = 0;
                if (s < n) {
                    const uint32_t c = s < nH ? list[s] : list[top - (s - nH)];
                    row = (int)((c >> 9) & 63u);
                    const int xl = (int)(c & 511u);
                    px = xl - lead;                                            // pixel offset inside the segment
                    const uint8_t *v = score + (row + 1) * FR_TPB + xl + xoff;
                    V = v[0];
                    if (V > thr) {
                        const int cjl = dW.div(px), lx = px - cjl * g.wCell, cww = min(g.wCell, rw - cjl * g.wCell);
                        // the rows above / below the cell are the zero apron rows; columns of the neighbouring cells count as 0
                        const int m = max(v[-FR_TPB], v[FR_TPB]);
                        const int mL = lx > 0 ? max(max(v[-1], v[-FR_TPB - 1]), v[FR_TPB - 1]) : 0;
                        const int mR = lx < cww - 1 ? max(max(v[1], v[-FR_TPB + 1]), v[FR_TPB + 1]) : 0;
                        pass = V > max(m, max(mL, mR));
                        if (pass) s_found[cjl] = 1;
                    }
                }
                const unsigned bal = __ballot_sync(0xffffffffu, pass);
                if (bal) {
                    int base = 0;
                    if (lane == 0) base = (int)atomicAdd(cand_count + (size_t)b * g.nlevels + g.level, (uint32_t)__popc(bal));
                    base = __shfl_sync(0xffffffffu, base, 0);
                    if (pass) {
                        const int slot = base + __popc(bal & ((1u << lane) - 1u));
                        if (slot < g.cand_cap)
                            cand[(size_t)b * g.cand_frame + g.cand_off + slot] =
                                (uint32_t)(xs + px - (EDGE - 3)) | ((uint32_t)(y0 + row - (EDGE - 3)) << 12) | ((uint32_t)(V - 1) << 24);
                        else
                            atomicOr(err, ERRF_CAND_OVERFLOW);
                    }
                }
            }
        }
        // ---- 6. cells that stayed empty once more at the lower threshold (ref :780-784): the same stages on their columns only
        if (thr == g.min_th) break;
        __syncthreads();
        {
            uint32_t keep = 0;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const int px = 4 * col + j - lead;
                if (((vm >> (8 * j + 7)) & 1u) && !s_found[dW.div(max(px, 0))]) keep |= 0x80u << (8 * j);
            }
            vm = keep;
        }
        if (!__syncthreads_or(vm != 0)) break;
        if (tid == 0) s_cnt = 0;
        thr = g.min_th;
        __syncthreads();
    }
}

static size_t fast_rows_layout(const FastLevel &fl, FastRowsLayout &lay) {
    lay.seg_cells = std::max(1, (4 * FR_COLS - 3) / fl.wCell);           // whole cells in <= 128 word columns at any alignment
    auto up = [](size_t v) { return (v + 15) & ~(size_t)15; };
    size_t off = up((size_t)(fl.hCell + 6) * FR_TPB);
    lay.off_score = (int)off; off = up(off + (size_t)(fl.hCell + 2) * FR_TPB);
    lay.off_slots = (int)off; off = up(off + (size_t)(fl.hCell + FR_PARTS - 1) * FR_COLS * 4);   // flagged words: a region of 32 columns x its rows per warp
    lay.list_cap = 2 * FR_COLS * fl.hCell;                               // half of the segment's pixels
    lay.off_list = (int)off; off = up(off + (size_t)lay.list_cap * 2);
    return off;
}

// ------------------------------------------------------------------------------------------
// Quadtree distribution, one CTA per (level, frame).  The reference's std::list is kept as an
// array in list order (front -> back).  One "pass" divides a chosen set of nodes in a chosen
// processing order; children are push_front'ed, so with K children created in the pass the
// t-th created child lands at position K-1-t and surviving old nodes follow in their old
// order.  Phase 1 processes every node with >1 keypoints in list order (ref :587-638);
// phase 2 processes them by (size desc, newest first) and stops as soon as the list holds
// N nodes (ref :646-709).
// ------------------------------------------------------------------------------------------
struct OctSmem {
    // sized [max_nodes] each; laid out manually in dynamic shared memory
    uint16_t *x0[2], *y0[2], *x1[2], *y1[2];
    uint32_t *cnt[2];
    uint32_t *ccount;     // [4*max_nodes]
    uint16_t *cpos;       // [4*max_nodes]
    uint16_t *keptpos;    // [max_nodes]
    uint8_t *divided;     // [max_nodes]
    int *scanA;           // [max_nodes] children-per-divided-node / creation base
    int *scanB;           // [max_nodes] undivided flags / rank
    unsigned long long *keys;   // [sort_n]
};

__device__ __forceinline__ int block_exclusive_scan(int *data, int n, int *warp_tot /*[32]*/) {
    // in-place exclusive scan of data[0..n) by the whole CTA; returns the total.
    const int tid = threadIdx.x, nt = blockDim.x;
    const int per = (n + nt - 1) / nt;
    const int beg = min(tid * per, n), end = min(beg + per, n);
    int sum = 0;
    for (int i = beg; i < end; i++) sum += data[i];
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, incl, o);
        if ((tid & 31) >= o) incl += v;
    }
    if ((tid & 31) == 31) warp_tot[tid >> 5] = incl;
    __syncthreads();
    if (tid < 32) {
        int w = tid < (nt >> 5) ? warp_tot[tid] : 0;
        int wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, wi, o);
            if (tid >= o) wi += v;
        }
        warp_tot[tid] = wi - w;   // exclusive warp offsets
        if (tid == 31) warp_tot[32] = wi;
    }
    __syncthreads();
    int run = warp_tot[tid >> 5] + incl - sum;
    for (int i = beg; i < end; i++) {
        int v = data[i];
        data[i] = run;
        run += v;
    }
    const int total = warp_tot[32];
    __syncthreads();
    return total;
}

__device__ __forceinline__ int quadrant_of(int x, int y, int nx0, int ny0, int nx1, int ny1) {
    // ref DivideNode :472-512.  halfX = ceil((UR.x-UL.x)/2) (exact for these small ints)
    const int hx = nx0 + ((nx1 - nx0 + 1) >> 1), hy = ny0 + ((ny1 - ny0 + 1) >> 1);
    return (x < hx ? 0 : 1) + (y < hy ? 0 : 2);
}

template <int NT>   // CTA size: 256 for batches (throughput), more for single frames (latency of the candidate loops)
__global__ void __launch_bounds__(NT)
k_octree(const Geom *__restrict__ G, const uint32_t *__restrict__ cand, const uint32_t *__restrict__ cand_count,
         uint16_t *__restrict__ node_of_pool, Kept *__restrict__ kept, uint32_t *__restrict__ kept_count,
         int *__restrict__ err, int level0) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int warp_tot[33];
    __shared__ int s_nexpand, s_jstar, s_phase, s_done;

    const int level = level0 + blockIdx.x, b = blockIdx.y;
    const LevelGeom &g = G->lv[level];
    const int MAXN = G->max_nodes;
    const int tid = threadIdx.x;
    const int N = g.quota;

    OctSmem S;
    {
        unsigned char *p = smem_raw;
        int sort_n = 1;
        while (sort_n < MAXN) sort_n <<= 1;
        S.keys = (unsigned long long *)p; p += sizeof(unsigned long long) * sort_n;
        S.ccount = (uint32_t *)p; p += 4 * 4 * MAXN;
        for (int k = 0; k < 2; k++) { S.cnt[k] = (uint32_t *)p; p += 4 * MAXN; }
        S.scanA = (int *)p; p += 4 * MAXN;
        S.scanB = (int *)p; p += 4 * MAXN;
        S.cpos = (uint16_t *)p; p += 2 * 4 * MAXN;
        for (int k = 0; k < 2; k++) {
            S.x0[k] = (uint16_t *)p; p += 2 * MAXN;
            S.y0[k] = (uint16_t *)p; p += 2 * MAXN;
            S.x1[k] = (uint16_t *)p; p += 2 * MAXN;
            S.y1[k] = (uint16_t *)p; p += 2 * MAXN;
        }
        S.keptpos = (uint16_t *)p; p += 2 * MAXN;
        S.divided = (uint8_t *)p; p += MAXN;
    }

    const uint32_t *C = cand + (size_t)b * G->cand_frame + g.cand_off;
    uint16_t *node_of = node_of_pool + (size_t)b * G->cand_frame + g.cand_off;
    const int n = min((int)cand_count[(size_t)b * G->nlevels + level], g.cand_cap);
    Kept *out = kept + (size_t)b * G->kept_frame + g.kept_off;
    uint32_t *out_n = kept_count + (size_t)b * G->nlevels + level;

    int cur = 0;
    // ---- initial nodes (ref :530-567)
    for (int i = tid; i < g.nIni; i += NT) {
        S.x0[0][i] = (uint16_t)g.ini_x[i]; S.x1[0][i] = (uint16_t)g.ini_x[i + 1];
        S.y0[0][i] = 0; S.y1[0][i] = (uint16_t)g.win_h;
        S.cnt[0][i] = 0;
    }
    __syncthreads();
    for (int c = tid; c < n; c += NT) {
        const int x = C[c] & 0xfff;
        int node = (int)__fdiv_rn((float)x, g.hX);   // vpIniNodes[kp.pt.x / hX]
        node = min(node, g.nIni - 1);
        node_of[c] = (uint16_t)node;
        atomicAdd(&S.cnt[0][node], 1u);
    }
    __syncthreads();
    // drop empty initial nodes, keep order
    for (int i = tid; i < g.nIni; i += NT) S.scanA[i] = S.cnt[0][i] > 0 ? 1 : 0;
    __syncthreads();
    int m = block_exclusive_scan(S.scanA, g.nIni, warp_tot);
    for (int i = tid; i < g.nIni; i += NT) {
        if (S.cnt[0][i] > 0) {
            const int p = S.scanA[i];
            S.x0[1][p] = S.x0[0][i]; S.x1[1][p] = S.x1[0][i]; S.y0[1][p] = S.y0[0][i]; S.y1[1][p] = S.y1[0][i];
            S.cnt[1][p] = S.cnt[0][i];
        }
    }
    __syncthreads();
    for (int c = tid; c < n; c += NT) node_of[c] = (uint16_t)S.scanA[node_of[c]];
    cur = 1;
    if (tid == 0) { s_phase = 1; s_done = 0; }
    __syncthreads();

    // ---- passes
    for (int iter = 0; iter < 64; iter++) {
        const int nxt = cur ^ 1;
        uint32_t *cnt = S.cnt[cur];
        const int phase = s_phase;
        for (int i = tid; i < 4 * m; i += NT) S.ccount[i] = 0;
        if (tid == 0) { s_nexpand = 0; s_jstar = 0x7fffffff; }
        __syncthreads();
        // child occupancy of every divisible node
        for (int c = tid; c < n; c += NT) {
            const int a = node_of[c];
            if (cnt[a] > 1) {
                const uint32_t pc = C[c];
                const int q = quadrant_of(pc & 0xfff, (pc >> 12) & 0xfff, S.x0[cur][a], S.y0[cur][a], S.x1[cur][a], S.y1[cur][a]);
                atomicAdd(&S.ccount[4 * a + q], 1u);
            }
        }
        __syncthreads();

        int K;   // children created in this pass
        if (phase == 1) {
            for (int a = tid; a < m; a += NT) {
                const bool dv = cnt[a] > 1;
                S.divided[a] = dv;
                const uint32_t *cc = &S.ccount[4 * a];
                S.scanA[a] = dv ? (cc[0] > 0) + (cc[1] > 0) + (cc[2] > 0) + (cc[3] > 0) : 0;
            }
            __syncthreads();
            K = block_exclusive_scan(S.scanA, m, warp_tot);
        } else {
            // expandable nodes sorted by (size desc, position asc) == reference's sorted vector
            // walked from the back with (size, creation order) keys.
            int sort_n = 1;
            while (sort_n < m) sort_n <<= 1;
            for (int a = tid; a < sort_n; a += NT) {
                unsigned long long key = 0;
                if (a < m) {
                    S.divided[a] = 0;
                    if (cnt[a] > 1) key = ((unsigned long long)cnt[a] << 16) | (unsigned long long)(0xffff - a);
                }
                S.keys[a] = key;
            }
            __syncthreads();
            for (int k = 2; k <= sort_n; k <<= 1)
                for (int j = k >> 1; j > 0; j >>= 1) {
                    for (int i = tid; i < sort_n; i += NT) {
                        const int ixj = i ^ j;
                        if (ixj > i) {
                            const unsigned long long A = S.keys[i], B = S.keys[ixj];
                            const bool desc = (i & k) == 0;
                            if (desc ? (A < B) : (A > B)) { S.keys[i] = B; S.keys[ixj] = A; }
                        }
                    }
                    __syncthreads();
                }
            // sorted index j -> node a_j; running list size after dividing j: m + sum_{i<=j}(k_i - 1)
            for (int j = tid; j < m; j += NT) {
                const unsigned long long key = S.keys[j];
                int kk = 0;
                if (key) {
                    const int a = 0xffff - (int)(key & 0xffff);
                    const uint32_t *cc = &S.ccount[4 * a];
                    kk = (cc[0] > 0) + (cc[1] > 0) + (cc[2] > 0) + (cc[3] > 0);
                    S.scanB[j] = kk - 1;
                } else
                    S.scanB[j] = 0;
                S.scanA[j] = kk;
            }
            __syncthreads();
            block_exclusive_scan(S.scanB, m, warp_tot);   // exclusive: growth before j
            for (int j = tid; j < m; j += NT) {
                const unsigned long long key = S.keys[j];
                if (key) {
                    const int after = m + S.scanB[j] + S.scanA[j] - 1;
                    if (after >= N) atomicMin(&s_jstar, j);
                }
            }
            __syncthreads();
            const int jstar = s_jstar;
            for (int j = tid; j < m; j += NT) {
                const unsigned long long key = S.keys[j];
                if (!key || j > jstar) S.scanA[j] = 0;
            }
            __syncthreads();
            K = block_exclusive_scan(S.scanA, m, warp_tot);   // creation base per sorted index
            // scatter the creation bases from sorted order back to list positions
            for (int j = tid; j < m; j += NT) {
                const unsigned long long key = S.keys[j];
                if (key && j <= jstar) {
                    const int a = 0xffff - (int)(key & 0xffff);
                    S.divided[a] = 1;
                    S.scanB[a] = S.scanA[j];
                }
            }
            __syncthreads();
            for (int a = tid; a < m; a += NT) S.scanA[a] = S.divided[a] ? S.scanB[a] : 0;
            __syncthreads();
        }
        // rank of undivided nodes
        for (int a = tid; a < m; a += NT) S.scanB[a] = S.divided[a] ? 0 : 1;
        __syncthreads();
        const int U = block_exclusive_scan(S.scanB, m, warp_tot);
        const int m_new = K + U;
        if (m_new > MAXN) {
            if (tid == 0) { atomicOr(err, ERRF_NODE_OVERFLOW); *out_n = 0; }
            return;
        }
        // build the new list
        for (int a = tid; a < m; a += NT) {
            const int ax0 = S.x0[cur][a], ay0 = S.y0[cur][a], ax1 = S.x1[cur][a], ay1 = S.y1[cur][a];
            if (S.divided[a]) {
                const int hx = ax0 + ((ax1 - ax0 + 1) >> 1), hy = ay0 + ((ay1 - ay0 + 1) >> 1);
                int t = S.scanA[a];
                int nexp = 0;
#pragma unroll
                for (int q = 0; q < 4; q++) {
                    const uint32_t cc = S.ccount[4 * a + q];
                    if (cc > 0) {
                        const int pos = K - 1 - t;
                        t++;
                        S.cpos[4 * a + q] = (uint16_t)pos;
                        S.x0[nxt][pos] = (uint16_t)((q & 1) ? hx : ax0);
                        S.x1[nxt][pos] = (uint16_t)((q & 1) ? ax1 : hx);
                        S.y0[nxt][pos] = (uint16_t)((q & 2) ? hy : ay0);
                        S.y1[nxt][pos] = (uint16_t)((q & 2) ? ay1 : hy);
                        S.cnt[nxt][pos] = cc;
                        nexp += cc > 1;
                    }
                }
                if (nexp) atomicAdd(&s_nexpand, nexp);
            } else {
                const int pos = K + S.scanB[a];
                S.keptpos[a] = (uint16_t)pos;
                S.x0[nxt][pos] = (uint16_t)ax0; S.x1[nxt][pos] = (uint16_t)ax1;
                S.y0[nxt][pos] = (uint16_t)ay0; S.y1[nxt][pos] = (uint16_t)ay1;
                S.cnt[nxt][pos] = cnt[a];
            }
        }
        __syncthreads();
        for (int c = tid; c < n; c += NT) {
            const int a = node_of[c];
            if (S.divided[a]) {
                const uint32_t pc = C[c];
                const int q = quadrant_of(pc & 0xfff, (pc >> 12) & 0xfff, S.x0[cur][a], S.y0[cur][a], S.x1[cur][a], S.y1[cur][a]);
                node_of[c] = S.cpos[4 * a + q];
            } else
                node_of[c] = S.keptpos[a];
        }
        __syncthreads();
        if (tid == 0) {
            // ref :640-644 and :706-707
            if (m_new >= N || m_new == m) s_done = 1;
            else if (phase == 1 && m_new + 3 * s_nexpand > N) s_phase = 2;
        }
        __syncthreads();
        m = m_new;
        cur = nxt;
        if (s_done) break;
    }

    // ---- best keypoint per node (ref :713-730): max response, first in candidate order wins.
    unsigned long long *best = S.keys;
    for (int a = tid; a < m; a += NT) best[a] = 0;
    __syncthreads();
    for (int c = tid; c < n; c += NT) {
        const uint32_t pc = C[c];
        const uint32_t x = pc & 0xfff, y = (pc >> 12) & 0xfff, r = pc >> 24;
        // candidate order = cells row-major, then row-major inside the cell (ref :758-793)
        const uint32_t ci = (y - 3) / g.hCell, cj = (x - 3) / g.wCell;
        const unsigned long long ord = ((unsigned long long)ci << 32) | ((unsigned long long)cj << 24) | ((unsigned long long)y << 12) | x;
        const unsigned long long key = ((unsigned long long)(r + 1) << 40) | (~ord & 0xffffffffffull);
        atomicMax(&best[node_of[c]], key);
    }
    __syncthreads();
    for (int a = tid; a < m; a += NT) {
        const unsigned long long key = best[a];
        const unsigned long long ord = ~key & 0xffffffffffull;
        if (a < g.kept_cap) {
            Kept k;
            k.x = (int16_t)((ord & 0xfff) + (EDGE - 3));          // += minBorderX (ref :809-810)
            k.y = (int16_t)(((ord >> 12) & 0xfff) + (EDGE - 3));
            k.response = (int32_t)(key >> 40) - 1;
            out[a] = k;
        }
    }
    if (tid == 0) {
        if (m > g.kept_cap) { atomicOr(err, ERRF_NODE_OVERFLOW); m = g.kept_cap; }
        *out_n = (uint32_t)m;
    }
}

// fp32 helpers of the descriptor stage (dev_fast_atan2, dev_sincosf): orb_devmath.cuh

// ------------------------------------------------------------------------------------------
// 7x7 Gaussian (sigma 2) of whole levels (ref :1024-1026: GaussianBlur on a clone of the level, BORDER_REFLECT_101 --
// the padded plane already holds that reflection).  SURVEY A.3: H = sum k*src (16 bit), out = (sum k*H + 32768) >> 16.
// A thread owns 4 horizontally adjacent pixels (one aligned word) and walks down BLUR_ROWS rows keeping the last
// seven horizontal sums in registers; every source row is fetched as three aligned words.
// ------------------------------------------------------------------------------------------
constexpr int BLUR_ROWS = 16;

// (16 resident CTAs of 128 threads = every warp slot, 32 registers without spills: 0.923 -> 0.915 ms for blur + describe against 12 CTAs
// at 40 registers; an explicit minimum of 1 made ptxas take 48 registers: 0.928)
#ifndef ORBGPU_BLUR_MINB
#define ORBGPU_BLUR_MINB 16
#endif
template <int TAPS>   // 0: OpenCV 3.x / 4.x fixed-point taps, 1: OpenCV 2.4's (compile-time: the tap vectors are immediates and the
                      // 2.4-only clamp disappears from the default build -- they were ~10 of 84 instructions per word-row)
__global__ void __launch_bounds__(128, ORBGPU_BLUR_MINB)
k_blur_planes(const uint8_t *__restrict__ pyr, uint8_t *__restrict__ blur, const Geom *__restrict__ G, int level) {
    const LevelGeom &g = G->lv[level];
    const int wx = blockIdx.x * blockDim.x + threadIdx.x;     // word column of the interior
    const int x0 = wx * 4;
    if (x0 >= g.w) return;
    const int y0 = blockIdx.y * BLUR_ROWS, b = blockIdx.z;
    constexpr uint32_t k0 = 18, k1 = 34, k2 = TAPS == 1 ? 49 : 48, k3 = TAPS == 1 ? 55 : 56;
    // taps as byte vectors: horizontal pass = two 4-byte dot products (IDP4A) over bytes j+1..j+4 and j+5..j+8 of the window;
    // vertical pass = 16-bit x 8-bit dot products (IDP2A) over row PAIRS (H[2m], H[2m+1]) packed into one register
    constexpr uint32_t KLO = k0 | (k1 << 8) | (k2 << 16) | (k3 << 24), KHI = k2 | (k1 << 8) | (k0 << 16);
    constexpr uint32_t T01 = k0 | (k1 << 8), T23 = k2 | (k3 << 8), T21 = k2 | (k1 << 8), T10 = k1 | (k0 << 8), T32 = k3 | (k2 << 8),
                       T12 = k1 | (k2 << 8), T_0 = k0 << 8;
    const uint8_t *src = pyr + (size_t)b * G->pyr_frame + g.interior_off + x0 - 4;
    uint8_t *dst = blur + (size_t)b * G->pyr_frame + g.interior_off + x0;
    uint32_t P[4][4];        // ring of row pairs: P[m & 3][j] = H[2m][j] | H[2m+1][j] << 16   (H <= 255 * 256 fits 16 bits)
    uint32_t hprev[4];       // H of the last even row, waiting for its partner
    // source rows -3.. and ..H+2 live in the 19-px frame.  Row indices are clamped (not predicated) and the next row's
    // words are requested before the current row is consumed; rows past H+2 only feed outputs that are never stored.
    // (source rows y0 - 3 .. y0 + BLUR_ROWS + 2 <= H + 17 all lie inside the padded plane, whose frame reaches row H + 18: no clamp)
    uint32_t n0, n1, n2;
    const int pitch = g.pitch;
    const uint8_t *sp = src + (ptrdiff_t)(y0 - 3) * pitch;
    uint8_t *dp = dst + (ptrdiff_t)y0 * pitch;
    {
        const uint32_t *p = reinterpret_cast<const uint32_t *>(sp);
        n0 = __ldg(p); n1 = __ldg(p + 1); n2 = __ldg(p + 2);
    }
#pragma unroll
    for (int r = 0; r < BLUR_ROWS + 6; r++) {
        const uint32_t w0 = n0, w1 = n1, w2 = n2;
        if (r + 1 < BLUR_ROWS + 6) {
            sp += pitch;
            const uint32_t *p = reinterpret_cast<const uint32_t *>(sp);
            n0 = __ldg(p); n1 = __ldg(p + 1); n2 = __ldg(p + 2);
        }
        uint32_t hc[4];      // horizontal pass of source row r for pixels x0 .. x0+3: pixel j reads window bytes j+1 .. j+7
        hc[0] = __dp4a(__funnelshift_r(w1, w2, 8), KHI, __dp4a(__funnelshift_r(w0, w1, 8), KLO, 0u));
        hc[1] = __dp4a(__funnelshift_r(w1, w2, 16), KHI, __dp4a(__funnelshift_r(w0, w1, 16), KLO, 0u));
        hc[2] = __dp4a(__funnelshift_r(w1, w2, 24), KHI, __dp4a(__funnelshift_r(w0, w1, 24), KLO, 0u));
        hc[3] = __dp4a(w2, KHI, __dp4a(w1, KLO, 0u));
        if (r & 1) {
#pragma unroll
            for (int j = 0; j < 4; j++) P[(r >> 1) & 3][j] = hprev[j] | (hc[j] << 16);
        } else {
#pragma unroll
            for (int j = 0; j < 4; j++) hprev[j] = hc[j];
        }
        if (r >= 6) {
            const int t = r - 6;                               // output row of the strip: source rows t .. t+6 = r
            const int yo = y0 + t;
            if (yo < g.h) {
                uint32_t v[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if ((t & 1) == 0) {                        // rows (t,t+1) (t+2,t+3) (t+4,t+5) are pairs, row t+6 is hc
                        const int m = t >> 1;
                        v[j] = __dp2a_lo(P[m & 3][j], T01, 32768u + k0 * hc[j]);
                        v[j] = __dp2a_lo(P[(m + 1) & 3][j], T23, v[j]);
                        v[j] = __dp2a_lo(P[(m + 2) & 3][j], T21, v[j]);
                    } else {                                   // row t is the upper half of pair (t-1,t); then three full pairs
                        const int m = (t - 1) >> 1;
                        v[j] = __dp2a_lo(P[m & 3][j], T_0, 32768u);
                        v[j] = __dp2a_lo(P[(m + 1) & 3][j], T12, v[j]);
                        v[j] = __dp2a_lo(P[(m + 2) & 3][j], T32, v[j]);
                        v[j] = __dp2a_lo(P[(m + 3) & 3][j], T10, v[j]);
                    }
                }
                // saturate_cast<uchar>((sum + 32768) >> 16) is byte 2 of each accumulator; the OpenCV 2.4 taps sum to 257, so a
                // white patch can reach 257 there and is clamped (the 3.x/4.x taps sum to 256 and cannot exceed 255)
                if (k2 == 49) {
#pragma unroll
                    for (int j = 0; j < 4; j++) v[j] = min(v[j], 0x00ffffffu);
                }
                const uint32_t o = __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
                *reinterpret_cast<uint32_t *>(dp) = o;
            }
            dp += pitch;
        }
    }
}

// ------------------------------------------------------------------------------------------
// Orientation + descriptor, one warp per kept keypoint, no shared memory: the intensity-centroid moments read the
// 31x31 disc of the level straight from L1/L2 (lane = column), the 256 rotated test pairs sample the blurred plane.
// ------------------------------------------------------------------------------------------
struct DescLevel {   // what k_describe needs of one level
    long long interior_off;
    int cap_end, kept_off, pitch;   // cap_end = sum of kept_cap over levels <= this one (flat slot -> level)
    float scale, size;
};
struct DescParams {   // by value: lives in the constant bank, no dependent global loads before the warp knows its key point
    int nlevels, kept_frame;
    long long pyr_frame;
    DescLevel lv[ORB_MAX_LEVELS];
};

// (8 resident CTAs = every warp slot of the SM: the kernel waits for its gathers -- 0.940 -> 0.903 ms for blur + describe per 288 frames
// against 6 CTAs at 38 registers; 32 registers, no spills)
__global__ void __launch_bounds__(DESC_WARPS * 32, 8)
k_describe(const uint8_t *__restrict__ pyr, const uint8_t *__restrict__ blur, const DescParams P,
           const Kept *__restrict__ kept, const uint32_t *__restrict__ kept_count, orb_keypoint *__restrict__ out_kp,
           uint8_t *__restrict__ out_desc, int32_t *__restrict__ out_n, int capacity, int *__restrict__ err) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.y;
    // flat slot -> (level, index) over the per-level kept capacities
    int slot = blockIdx.x * DESC_WARPS + warp;
    if (slot >= P.lv[P.nlevels - 1].cap_end) return;
    int level = 0;   // number of l in [0, 15) with slot >= cap_end[l] (cap_end ascends; beyond nlevels - 2 it is the total > slot): 4 probes
#pragma unroll
    for (int step = ORB_MAX_LEVELS / 2; step >= 1; step >>= 1)
        if (slot >= P.lv[level + step - 1].cap_end) level += step;
    if (level > 0) slot -= P.lv[level - 1].cap_end;
    // kept counts of all levels: lane l holds level l's, an inclusive warp scan gives this level's output base and the total
    const uint32_t *kc = kept_count + (size_t)b * P.nlevels;
    const int mine = lane < P.nlevels ? (int)kc[lane] : 0;
    int incl = mine;
#pragma unroll
    for (int o = 1; o < ORB_MAX_LEVELS; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (slot >= __shfl_sync(0xffffffffu, mine, level)) return;
    const int base = __shfl_sync(0xffffffffu, incl - mine, level), total = __shfl_sync(0xffffffffu, incl, P.nlevels - 1);
    if (base == 0 && slot == 0 && lane == 0) out_n[b] = total;   // first non-empty level's first warp
    const int oidx = base + slot;
    if (oidx >= capacity) {
        if (lane == 0) atomicOr(err, ERRF_OUT_OVERFLOW);
        return;
    }
    const DescLevel &g = P.lv[level];
    const Kept kp = kept[(size_t)b * P.kept_frame + g.kept_off + slot];
    const size_t plane = (size_t)b * P.pyr_frame + g.interior_off + (ptrdiff_t)kp.y * g.pitch + kp.x;

    // ---- IC_Angle (ref :73-98): lanes own columns u = lane-15 .. of the 31-wide disc
    int m10 = 0, m01 = 0;
    {
        const int a = (kp.x - 15) & 3;
        const uint8_t *c = pyr + plane - 15 - a;                      // 4-byte aligned: plane rows are, and so is interior column 0
        const uint2 *tab = d_moment_tab + a * MOM_TASKS + lane;
        uint32_t px[MOM_TASKS / 32];
#pragma unroll
        for (int i = 0; i < MOM_TASKS / 32; i++) {
            const int t = min(lane + 32 * i, 31 * 9 - 1), r = (t * 7282) >> 16, wi = t - 9 * r;   // r = t / 9 for t < 279
            px[i] = *reinterpret_cast<const uint32_t *>(c + (ptrdiff_t)(r - 15) * g.pitch + 4 * wi);
        }
#pragma unroll
        for (int i = 0; i < MOM_TASKS / 32; i++) {
            const uint2 e = __ldg(tab + 32 * i);
            m10 = dp4a_us(px[i], (int)e.x, m10);
            m01 = dp4a_us(px[i], (int)e.y, m01);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m10 += __shfl_xor_sync(0xffffffffu, m10, o);
        m01 += __shfl_xor_sync(0xffffffffu, m01, o);
    }
    const float angle = dev_fast_atan2((float)m01, (float)m10);

    // ---- steered BRIEF (ref :101-144): lane computes descriptor byte `lane`
    float sn, cs;
    const float factorPI = (float)(3.1415926535897932384626433832795 / 180.0);   // ref :100
    dev_sincosf(__fmul_rn(angle, factorPI), &sn, &cs);
    const uint8_t *ctr = blur + plane;
    const int BP = g.pitch;
    uint32_t pw[8];   // this lane's 8 test pairs: 32 signed bytes (x0,y0,x1,y1) x 8
    {
        const uint4 *pp = reinterpret_cast<const uint4 *>(d_pattern + lane * 32);
        const uint4 pa = __ldg(pp), pb = __ldg(pp + 1);
        pw[0] = pa.x; pw[1] = pa.y; pw[2] = pa.z; pw[3] = pa.w; pw[4] = pb.x; pw[5] = pb.y; pw[6] = pb.z; pw[7] = pb.w;
    }
    uint32_t val = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const float x0 = (float)(int8_t)(pw[k] & 0xff), y0 = (float)(int8_t)((pw[k] >> 8) & 0xff);
        const float x1 = (float)(int8_t)((pw[k] >> 16) & 0xff), y1 = (float)(int8_t)(pw[k] >> 24);
        const int r0 = __float2int_rn(__fadd_rn(__fmul_rn(x0, sn), __fmul_rn(y0, cs)));
        const int c0 = __float2int_rn(__fsub_rn(__fmul_rn(x0, cs), __fmul_rn(y0, sn)));
        const int r1 = __float2int_rn(__fadd_rn(__fmul_rn(x1, sn), __fmul_rn(y1, cs)));
        const int c1 = __float2int_rn(__fsub_rn(__fmul_rn(x1, cs), __fmul_rn(y1, sn)));
        const int t0 = ctr[r0 * BP + c0], t1 = ctr[r1 * BP + c1];
        val |= (uint32_t)(t0 < t1) << k;
    }
    out_desc[((size_t)b * capacity + oidx) * 32 + lane] = (uint8_t)val;
    if (lane == 0) {
        orb_keypoint o;
        o.x = (float)kp.x; o.y = (float)kp.y;
        if (level != 0) { o.x = __fmul_rn(o.x, g.scale); o.y = __fmul_rn(o.y, g.scale); }   // ref :1035-1041
        o.size = g.size;
        o.angle = angle;
        o.response = (float)kp.response;
        o.octave = level;
        o.class_id = -1;
        out_kp[(size_t)b * capacity + oidx] = o;
    }
}

}  // namespace orbgpu

using namespace orbgpu;

// ------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------
// Results of a small call to the host without DMA copies: every D2H cudaMemcpyAsync costs 6-8 us of latency on the stream and a
// blocking single-frame call made four of them (flags, counts, key points, descriptors).  One kernel writes the flags, the counts
// and the n valid entries of every frame into pinned host memory (coalesced 4-byte stores across PCIe) instead.
__global__ void __launch_bounds__(256)
k_export_results(const uint32_t *__restrict__ kp, const uint32_t *__restrict__ desc, const int32_t *__restrict__ n, const int *__restrict__ err,
                 int capacity, int batch, const uint32_t *__restrict__ ur, const uint32_t *__restrict__ dp, const int32_t *__restrict__ nm,
                 int *__restrict__ h_head, uint32_t *__restrict__ h_land) {
    const int b = blockIdx.y, t = blockIdx.x * blockDim.x + threadIdx.x, nt = gridDim.x * blockDim.x;
    const int nb = min(max(n[b], 0), capacity);
    const size_t kp_words = (size_t)batch * capacity * 7, desc_words = (size_t)batch * capacity * 8;
    const uint32_t *sk = kp + (size_t)b * capacity * 7, *sd = desc + (size_t)b * capacity * 8;
    uint32_t *dk = h_land + (size_t)b * capacity * 7, *dd = h_land + kp_words + (size_t)b * capacity * 8;
    for (int i = t; i < nb * 7; i += nt) dk[i] = sk[i];
    for (int i = t; i < nb * 8; i += nt) dd[i] = sd[i];
    if (b == 0) {
        if (ur) {   // mvuRight / mvDepth of the left frame and the accepted-match count
            uint32_t *du = h_land + kp_words + desc_words;
            for (int i = t; i < nb; i += nt) { du[i] = ur[i]; du[capacity + i] = dp[i]; }
            if (t == 0) du[2 * capacity] = (uint32_t)*nm;
        }
        if (t == 0) h_head[0] = *err;
        if (t >= 1 && t <= batch) h_head[t] = n[t - 1];
    }
}

static void prof_mark(orb_extractor *ex, int i) {
    if (ex->profiling) cudaEventRecord(ex->ev[i], ex->stream);
}
static void prof_collect(orb_extractor *ex) {
    if (!ex->profiling || !ex->ev_pending) return;
    cudaEventSynchronize(ex->ev[ORB_N_STAGES]);
    for (int i = 0; i < ORB_N_STAGES; i++) {
        float ms = 0;
        cudaEventElapsedTime(&ms, ex->ev[i], ex->ev[i + 1]);
        ex->stage_ms[i] += ms;
    }
    ex->stage_calls++;
    ex->ev_pending = false;
}

static int cv_round_host(double v) { return (int)std::nearbyint(v); }

static void build_tables(orb_extractor *ex) {
    // ref ORBextractor::ORBextractor :404-460 (same statements as the reference, float/double mix kept)
    const orb_params &p = ex->prm;
    const int nl = p.nlevels;
    const double scaleFactor = p.scale_factor;
    ex->scale.assign(nl, 1.f); ex->sigma2.assign(nl, 1.f); ex->inv_scale.assign(nl, 1.f); ex->inv_sigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; i++) {
        ex->scale[i] = (float)(ex->scale[i - 1] * scaleFactor);
        ex->sigma2[i] = ex->scale[i] * ex->scale[i];
    }
    for (int i = 0; i < nl; i++) {
        ex->inv_scale[i] = 1.0f / ex->scale[i];
        ex->inv_sigma2[i] = 1.0f / ex->sigma2[i];
    }
    ex->per_level.assign(nl, 0);
    const float factor = (float)(1.0f / scaleFactor);
    float nDesired = p.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; l++) {
        ex->per_level[l] = cv_round_host(nDesired);
        sum += ex->per_level[l];
        nDesired *= factor;
    }
    ex->per_level[nl - 1] = std::max(p.nfeatures - sum, 0);
    int v, v0;
    const int vmax = (int)std::floor(15 * std::sqrt(2.f) / 2 + 1), vmin = (int)std::ceil(15 * std::sqrt(2.f) / 2);
    for (v = 0; v < 16; v++) ex->umax[v] = 0;
    for (v = 0; v <= vmax; ++v) ex->umax[v] = cv_round_host(std::sqrt(225.0 - v * v));
    for (v = 15, v0 = 0; v >= vmin; --v) {
        while (ex->umax[v0] == ex->umax[v0 + 1]) ++v0;
        ex->umax[v] = v0;
        ++v0;
    }
}

// Level geometry for a (width,height) image; returns ORB_ERR_TOO_SMALL where the reference would
// divide by zero.  taps (optional) receives the resize tables.
static int compute_geometry(const orb_extractor *ex, int width, int height, Geom *G, std::vector<ResizeTap> *taps) {
    const orb_params &p = ex->prm;
    memset(G, 0, sizeof(*G));
    G->nlevels = p.nlevels; G->width = width; G->height = height;
    G->ini_th = p.ini_th_fast; G->min_th = p.min_th_fast; G->blur_taps = p.blur_taps;
    long long off = 0;
    int cells = 0, cand = 0, kept = 0, maxn = 0, tapoff = 0;
    if (taps) taps->clear();
    for (int l = 0; l < p.nlevels; l++) {
        LevelGeom &g = G->lv[l];
        g.w = cv_round_host((float)width * ex->inv_scale[l]);      // ref :1049-1051
        g.h = cv_round_host((float)height * ex->inv_scale[l]);
        g.pitch = (XOFF + g.w + EDGE + 127) / 128 * 128;
        g.plane_off = off;
        g.interior_off = off + (long long)EDGE * g.pitch + XOFF;
        off += (long long)g.pitch * (g.h + 2 * EDGE);
        off = (off + 255) / 256 * 256;
        const int minB = EDGE - 3;
        g.maxBX = g.w - EDGE + 3; g.maxBY = g.h - EDGE + 3;
        const float wf = (float)(g.maxBX - minB), hf = (float)(g.maxBY - minB);
        if (wf < 30.f || hf < 30.f) return ORB_ERR_TOO_SMALL;
        const int nCols = (int)(wf / 30.f), nRows = (int)(hf / 30.f);                 // ref :753-754
        g.wCell = (int)std::ceil(wf / nCols); g.hCell = (int)std::ceil(hf / nRows);   // ref :755-756
        if (g.wCell + 6 + 15 > FT_PB || g.hCell + 6 > FAST_TILE_MAX) return ORB_ERR_INVALID;
        g.ncx = 0; g.ncy = 0;
        for (int j = 0; j < nCols; j++) if (!(minB + j * g.wCell >= g.maxBX - 6)) g.ncx = j + 1;   // ref :770
        for (int i = 0; i < nRows; i++) if (!(minB + i * g.hCell >= g.maxBY - 3)) g.ncy = i + 1;   // ref :762
        g.cell_begin = cells;
        cells += g.ncx * g.ncy;
        g.win_w = g.maxBX - minB; g.win_h = g.maxBY - minB;
        g.quota = ex->per_level[l];
        g.nIni = (int)std::round(static_cast<float>(g.win_w) / g.win_h);              // ref :530
        if (g.nIni < 1) return ORB_ERR_TOO_SMALL;
        if (g.nIni > MAX_INI) return ORB_ERR_INVALID;
        g.hX = static_cast<float>(g.win_w) / g.nIni;                                  // ref :532
        for (int i = 0; i <= g.nIni; i++) g.ini_x[i] = (int)(g.hX * static_cast<float>(i));   // ref :541-542
        if (g.win_w > 4095 || g.win_h > 4095) return ORB_ERR_INVALID;
        g.cand_off = cand;
        g.cand_cap = ((g.win_w - 6) * (g.win_h - 6)) / 4 + 64;
        cand += (g.cand_cap + 63) / 64 * 64;
        g.kept_off = kept;
        g.kept_cap = std::max(g.quota + 3, 4 * g.nIni) + 1;
        kept += g.kept_cap;
        maxn = std::max(maxn, g.kept_cap);
        g.scale = ex->scale[l];
        g.inv_scale = ex->inv_scale[l];
        g.size = (float)(int)(31 * ex->scale[l]);                                     // ref :804
        if (l > 0) {
            const LevelGeom &s = G->lv[l - 1];
            g.tap_x_off = tapoff; g.tap_y_off = tapoff + g.w;
            tapoff += g.w + g.h;
            if (taps) {
                // cv::resize INTER_LINEAR coefficient tables
                for (int axis = 0; axis < 2; axis++) {
                    const int dn = axis ? g.h : g.w, sn = axis ? s.h : s.w;
                    const double sc = 1.0 / ((double)dn / sn);
                    for (int d = 0; d < dn; d++) {
                        float f = (float)((d + 0.5) * sc - 0.5);
                        int si = (int)std::floor(f);
                        f -= si;
                        if (si < 0) { f = 0; si = 0; }
                        if (si >= sn - 1) { f = 0; si = sn - 1; }
                        ResizeTap t;
                        t.s0 = (int16_t)si; t.s1 = (int16_t)std::min(si + 1, sn - 1);
                        t.a0 = (int16_t)cv_round_host((1.f - f) * 2048.f);
                        t.a1 = (int16_t)cv_round_host(f * 2048.f);
                        taps->push_back(t);
                    }
                }
            }
        }
    }
    G->total_cells = cells; G->cand_frame = cand; G->kept_frame = kept; G->pyr_frame = off; G->max_nodes = maxn;
    return ORB_OK;
}

struct NvtxRange {   // host-side range around the enqueue of one stage (ncu --nvtx / any NVTX consumer)
    explicit NvtxRange(const char *name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

static size_t octree_smem_bytes(int maxn) {
    int sort_n = 1;
    while (sort_n < maxn) sort_n <<= 1;
    return 8 * (size_t)sort_n + (size_t)maxn * (16 + 8 + 4 + 4 + 8 + 16 + 2 + 1) + 64;
}

static void free_all(orb_extractor *ex) {
    cudaFree(ex->d_geom); cudaFree(ex->d_taps); cudaFree(ex->d_wtab); cudaFree(ex->d_cells); cudaFree(ex->d_src); cudaFree(ex->d_pyr); cudaFree(ex->d_blur); cudaFree(ex->d_cand);
    cudaFree(ex->d_node_of); cudaFree(ex->d_counts); cudaFree(ex->d_kept); cudaFree(ex->d_out_kp);
    cudaFree(ex->d_out_desc); cudaFree(ex->d_out_n); cudaFree(ex->d_err); cudaFree(ex->stereo_scratch); cudaFree(ex->d_stereo_out);
    if (ex->h_pinned) cudaFreeHost(ex->h_pinned);
    if (ex->h_bounce) cudaFreeHost(ex->h_bounce);
    if (ex->h_src) cudaFreeHost(ex->h_src);
    if (ex->gexec) cudaGraphExecDestroy(ex->gexec);
    for (auto &e : ex->ev) if (e) cudaEventDestroy(e);
    for (auto &e : ex->ev_lvl) if (e) cudaEventDestroy(e);
    if (ex->ev_pair) cudaEventDestroy(ex->ev_pair);
    if (ex->ev_pair_done) cudaEventDestroy(ex->ev_pair_done);
    if (ex->ev_fast) cudaEventDestroy(ex->ev_fast);
    if (ex->ev_blur) cudaEventDestroy(ex->ev_blur);
    if (ex->s_fast) cudaStreamDestroy(ex->s_fast);
    if (ex->s_blur) cudaStreamDestroy(ex->s_blur);
    for (auto &e : ex->ev_fl) if (e) cudaEventDestroy(e);
    for (auto &e : ex->ev_ol) if (e) cudaEventDestroy(e);
    for (auto &e : ex->ev_chunk) if (e) cudaEventDestroy(e);
    for (auto &q : ex->s_oct) if (q) cudaStreamDestroy(q);
    for (auto &q : ex->s_bl) if (q) cudaStreamDestroy(q);
    if (ex->stream) cudaStreamDestroy(ex->stream);
}

static int upload_constants_once() {
    static std::mutex mu;
    static std::vector<int> done_devices;
    std::lock_guard<std::mutex> lk(mu);
    int dev = 0;
    ORB_CUDA_TRY(cudaGetDevice(&dev));
    if (std::find(done_devices.begin(), done_devices.end(), dev) != done_devices.end()) return ORB_OK;
    static const int pat[ORB_PATTERN_INTS] = {ORB_PATTERN_VALUES};
    int8_t p8[ORB_PATTERN_INTS];
    for (int i = 0; i < ORB_PATTERN_INTS; i++) p8[i] = (int8_t)pat[i];
    ORB_CUDA_TRY(cudaMemcpyToSymbol(d_pattern, p8, sizeof(p8)));
    done_devices.push_back(dev);
    return ORB_OK;
}

// One 3-D tensor map per level for k_fast_rows' tile loads: {plane row bytes, plane rows, frame slots} of u8, box = FR_TPB bytes x
// (hCell + 6) rows x 1 frame.  Levels where the driver refuses (or ORBGPU_NO_TMA=1) keep the cp.async staging.
static void build_tensor_maps(orb_extractor *ex, const Geom &G) {
    typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                 CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    static bool looked = false;
    if (!looked) {
        looked = true;
        void *fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (!getenv("ORBGPU_NO_TMA") && cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            encode = (EncodeFn)fn;
        (void)cudaGetLastError();
    }
    for (int l = 0; l < ORB_MAX_LEVELS; l++) ex->tmap_ok[l] = false;
    if (!encode) return;
    for (int l = 0; l < G.nlevels; l++) {
        const LevelGeom &g = G.lv[l];
        const cuuint64_t dims[3] = {(cuuint64_t)g.pitch, (cuuint64_t)(g.h + 2 * EDGE), (cuuint64_t)ex->prm.max_batch};
        const cuuint64_t strides[2] = {(cuuint64_t)g.pitch, (cuuint64_t)G.pyr_frame};
        const cuuint32_t box[3] = {(cuuint32_t)FR_TPB, (cuuint32_t)(g.hCell + 6), 1u}, estr[3] = {1u, 1u, 1u};
        if (g.hCell + 6 > 256 || (G.pyr_frame % 16) != 0) continue;
        ex->tmap_ok[l] = encode(&ex->tmap[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, ex->d_pyr + g.plane_off, dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
    }
}

static int set_geometry(orb_extractor *ex, int width, int height) {
    if (ex->geom.width == width && ex->geom.height == height && ex->geom.nlevels == ex->prm.nlevels) return ORB_OK;
    if (ex->gexec) { cudaGraphExecDestroy(ex->gexec); ex->gexec = nullptr; }   // the captured launches hold the old geometry
    ex->glast = {};
    Geom G;
    std::vector<ResizeTap> taps;
    int rc = compute_geometry(ex, width, height, &G, &taps);
    ORB_REQUIRE(rc == ORB_OK, rc, "image %dx%d: %s", width, height,
                rc == ORB_ERR_TOO_SMALL ? "a pyramid level is too small for the 30-px cell grid (reference divides by zero)"
                                        : "unsupported geometry");
    ORB_REQUIRE(taps.size() <= ex->taps_cap, ORB_ERR_INVALID, "internal: tap pool too small");
    build_tensor_maps(ex, G);
    // per-output-word resize records (fast path), two uint4 per word: weights and byte selectors (see k_pyr_resize).  A
    // level where some pixel's two source bytes do not fit the selectors' reach keeps the generic kernel.
    std::vector<uint4> wtab;
    for (int l = 1; l < G.nlevels; l++) {
        LevelGeom &g = G.lv[l];
        g.wtab_off = (int)wtab.size() / 2;
        g.fast_resize = 1;
        const int nwords = g.pitch / 4;
        for (int w = 0; w < nwords; w++) {
            int lo = 1 << 30;
            ResizeTap tx[4];
            bool in[4];
            for (int k = 0; k < 4; k++) {
                const int x = w * 4 + k - XOFF;
                in[k] = x >= -EDGE && x < g.w + EDGE;
                tx[k] = taps[g.tap_x_off + reflect101(in[k] ? x : 0, g.w)];
                if (in[k]) lo = std::min(lo, (int)tx[k].s0);
            }
            uint32_t wt[4] = {0, 0, 0, 0}, o[4] = {0, 0, 0, 0};
            int base = 0;
            if (lo != (1 << 30)) {
                base = lo & ~3;   // interior rows start 32-byte aligned, so this is a word boundary of the source row
                for (int k = 0; k < 4; k++)
                    if (in[k]) {
                        o[k] = (uint32_t)(tx[k].s0 - base);   // the taps are bytes o, o+1 of the window (s1 == s0 only where a1 == 0)
                        wt[k] = (uint32_t)(uint16_t)tx[k].a0 | ((uint32_t)(uint16_t)tx[k].a1 << 16);
                        if (o[k] + 1 > 11u) g.fast_resize = 0;
                    }
            }
            // x01 = PRMT(PRMT(w0, w1, selA), w2, selB) = window bytes {o0, o0+1, o1, o1+1}: a byte below 8 comes from the first
            // PRMT (kept in place by the second), one from 8 up straight from w2.  x23 likewise for pixels 2 and 3.  (Inside
            // the reflected frame the offsets are not monotone in k, so every pixel gets the full 12-byte reach.)
            auto selectors = [](uint32_t oa, uint32_t ob) {
                uint32_t selA = 0, selB = 0;
                const uint32_t want[4] = {oa, oa + 1, ob, ob + 1};
                for (int i = 0; i < 4; i++) {
                    if (want[i] < 8) { selA |= want[i] << (4 * i); selB |= (uint32_t)i << (4 * i); }
                    else selB |= (4u + ((want[i] - 8) & 3)) << (4 * i);
                }
                return selA | (selB << 16);
            };
            wtab.push_back(make_uint4(wt[0], wt[1], wt[2], wt[3]));
            wtab.push_back(make_uint4(selectors(o[0], o[1]), selectors(o[2], o[3]), (uint32_t)(base / 4), 0u));
        }
    }
    // row records of the fast path (see pyr_resize_column)
    for (int l = 1; l < G.nlevels; l++) {
        LevelGeom &g = G.lv[l];
        const LevelGeom &sl = G.lv[l - 1];
        g.rtab_off = (int)wtab.size();
        int prev_s1 = -1;
        for (int py = 0; py < g.h + 2 * EDGE; py++) {
            const ResizeTap ty = taps[g.tap_y_off + reflect101(py - EDGE, g.h)];
            uint32_t x = (uint32_t)ty.s0 * (uint32_t)sl.pitch, y = (uint32_t)ty.s1 * (uint32_t)sl.pitch;
            if (py > 0 && ty.s0 == prev_s1) x |= 1u;
            if (ty.s1 == ty.s0) x |= 2u;
            prev_s1 = ty.s1;
            wtab.push_back(make_uint4(x, y, (uint32_t)(uint16_t)ty.a0 << 16, (uint32_t)(uint16_t)ty.a1 << 16));
        }
    }
    ORB_REQUIRE(wtab.size() <= ex->wtab_cap, ORB_ERR_INVALID, "internal: resize word table too small");
    if (!wtab.empty())
        ORB_CUDA_TRY(cudaMemcpyAsync(ex->d_wtab, wtab.data(), wtab.size() * sizeof(uint4), cudaMemcpyHostToDevice, ex->stream));
    std::vector<uint32_t> cells;
    cells.reserve(G.total_cells);
    for (int l = 0; l < G.nlevels; l++)
        for (int ci = 0; ci < G.lv[l].ncy; ci++)
            for (int cj = 0; cj < G.lv[l].ncx; cj++) cells.push_back((uint32_t)l | ((uint32_t)ci << 4) | ((uint32_t)cj << 14));
    ORB_REQUIRE(cells.size() <= ex->cells_cap && G.lv[0].ncx < 1024 && G.lv[0].ncy < 1024, ORB_ERR_INVALID, "internal: cell table too small");
    if (!cells.empty())
        ORB_CUDA_TRY(cudaMemcpyAsync(ex->d_cells, cells.data(), cells.size() * 4, cudaMemcpyHostToDevice, ex->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(ex->d_geom, &G, sizeof(G), cudaMemcpyHostToDevice, ex->stream));
    if (!taps.empty())
        ORB_CUDA_TRY(cudaMemcpyAsync(ex->d_taps, taps.data(), taps.size() * sizeof(ResizeTap), cudaMemcpyHostToDevice, ex->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(ex->stream));   // `taps` and G are stack/heap temporaries
    ex->geom = G;
    return ORB_OK;
}

int orbgpu_ensure_out_staging(orb_extractor *ex, int batch, int capacity) {
    if ((size_t)batch * capacity <= ex->out_cap_elems) return ORB_OK;
    // grow the staging buffers to the caller's capacity
    cudaFree(ex->d_out_kp); cudaFree(ex->d_out_desc);
    ex->d_out_kp = nullptr; ex->d_out_desc = nullptr;
    if (ex->gexec) { cudaGraphExecDestroy(ex->gexec); ex->gexec = nullptr; }
    ex->glast = {};
    ex->out_cap_elems = (size_t)ex->prm.max_batch * capacity;
    ORB_CUDA_TRY(cudaMalloc(&ex->d_out_kp, ex->out_cap_elems * sizeof(orb_keypoint)));
    ORB_CUDA_TRY(cudaMalloc(&ex->d_out_desc, ex->out_cap_elems * 32));
    return ORB_OK;
}

void orbgpu_export_results(orb_extractor *ex, const orb_keypoint *kp, const uint8_t *desc, const int32_t *n, int batch, int capacity,
                           const float *u_right, const float *depth, const int32_t *n_matched, cudaStream_t st) {
    k_export_results<<<dim3(8, batch), 256, 0, st>>>((const uint32_t *)kp, (const uint32_t *)desc, n, ex->d_err, capacity, batch,
                                                     (const uint32_t *)u_right, (const uint32_t *)depth, n_matched, ex->h_pinned,
                                                     (uint32_t *)ex->h_bounce);
    ex->launches++;
}

extern "C" {

const char *orb_last_error(void) { return g_err; }
int orb_abi_version(void) { return ORBGPU_ABI_VERSION; }

int orb_extractor_create(const orb_params *params, orb_extractor **out) {
    ORB_REQUIRE(params && out, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(params->nlevels >= 1 && params->nlevels <= ORB_MAX_LEVELS, ORB_ERR_INVALID, "nlevels out of range");
    ORB_REQUIRE(params->nfeatures >= 1 && params->nfeatures <= 60000, ORB_ERR_INVALID, "nfeatures out of range");
    ORB_REQUIRE(params->scale_factor > 1.0f, ORB_ERR_INVALID, "scale_factor must be > 1");
    ORB_REQUIRE(params->max_width >= 1 && params->max_height >= 1 && params->max_width <= 4000 && params->max_height <= 4000,
                ORB_ERR_INVALID, "max_width/max_height out of range");
    ORB_REQUIRE(params->max_batch >= 1 && params->max_batch <= 4096, ORB_ERR_INVALID, "max_batch out of range");
    ORB_REQUIRE(params->ini_th_fast >= 1 && params->min_th_fast >= 1 && params->ini_th_fast < 255 && params->min_th_fast < 255,
                ORB_ERR_INVALID, "FAST thresholds out of range");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback",
                cudaGetErrorString(e));
    orb_extractor *ex = new orb_extractor();
    ex->prm = *params;
    if (params->device >= 0) {
        ORB_CUDA_TRY(cudaSetDevice(params->device));
        ex->device = params->device;
    } else {
        ORB_CUDA_TRY(cudaGetDevice(&ex->device));
    }
    build_tables(ex);
    memset(&ex->geom, 0, sizeof(ex->geom));
    Geom G;
    int rc = compute_geometry(ex, params->max_width, params->max_height, &G, nullptr);
    if (rc != ORB_OK) {
        delete ex;
        set_error("max image %dx%d: a pyramid level is too small or the geometry is unsupported", params->max_width, params->max_height);
        return rc;
    }
    const size_t B = params->max_batch;
    rc = upload_constants_once();
    if (rc != ORB_OK) { delete ex; return rc; }
#define TRY_ALLOC(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { set_error("%s -> %s", #expr, cudaGetErrorString(_e)); free_all(ex); delete ex; return ORB_ERR_CUDA; } } while (0)
    TRY_ALLOC(cudaStreamCreateWithFlags(&ex->stream, cudaStreamNonBlocking));
    TRY_ALLOC(cudaStreamCreateWithFlags(&ex->s_fast, cudaStreamNonBlocking));
    TRY_ALLOC(cudaStreamCreateWithFlags(&ex->s_blur, cudaStreamNonBlocking));
    for (int l = 0; l < params->nlevels; l++) TRY_ALLOC(cudaEventCreateWithFlags(&ex->ev_lvl[l], cudaEventDisableTiming));
    TRY_ALLOC(cudaEventCreateWithFlags(&ex->ev_fast, cudaEventDisableTiming));
    TRY_ALLOC(cudaEventCreateWithFlags(&ex->ev_blur, cudaEventDisableTiming));
    for (int l = 0; l < ORB_MAX_LEVELS; l++) {
        TRY_ALLOC(cudaStreamCreateWithFlags(&ex->s_oct[l], cudaStreamNonBlocking));
        TRY_ALLOC(cudaStreamCreateWithFlags(&ex->s_bl[l], cudaStreamNonBlocking));
        TRY_ALLOC(cudaEventCreateWithFlags(&ex->ev_fl[l], cudaEventDisableTiming));
        TRY_ALLOC(cudaEventCreateWithFlags(&ex->ev_ol[l], cudaEventDisableTiming));
    }
    ex->overlap = getenv("ORBGPU_SERIAL") == nullptr;
    ex->use_graph = getenv("ORBGPU_NO_GRAPH") == nullptr;
    TRY_ALLOC(cudaMemcpyToSymbol(c_umax, ex->umax, sizeof(ex->umax)));
    {
        std::vector<uint2> tab(4 * MOM_TASKS);
        for (int a = 0; a < 4; a++) {
            for (int t = 31 * 9; t < MOM_TASKS; t++) tab[a * MOM_TASKS + t] = make_uint2(0, 0);   // no-op: zero weights
            for (int t = 0; t < 31 * 9; t++) {
                const int v = t / 9 - 15, wi = t % 9, xoff = 4 * wi - 15 - a;
                uint32_t wu = 0, wv = 0;
                for (int j = 0; j < 4; j++) {
                    const int u = xoff + j;
                    if (abs(u) <= ex->umax[abs(v)]) { wu |= (uint32_t)(uint8_t)(int8_t)u << (8 * j); wv |= (uint32_t)(uint8_t)(int8_t)v << (8 * j); }
                }
                tab[a * MOM_TASKS + t] = make_uint2(wu, wv);
            }
        }
        TRY_ALLOC(cudaMemcpyToSymbol(d_moment_tab, tab.data(), tab.size() * sizeof(uint2)));
    }
    TRY_ALLOC(cudaMalloc(&ex->d_geom, sizeof(Geom)));
    // sizes below are monotone in the image size except the kept pool (aspect dependent): bound it
    size_t kept_frame = 0;
    int maxn = 0;
    for (int l = 0; l < params->nlevels; l++) {
        int cap = std::max(ex->per_level[l] + 3, 4 * MAX_INI) + 1;
        kept_frame += cap;
        maxn = std::max(maxn, cap);
    }
    ex->taps_cap = (size_t)(params->max_width + params->max_height + 64) * params->nlevels;
    TRY_ALLOC(cudaMalloc(&ex->d_taps, ex->taps_cap * sizeof(ResizeTap)));
    ex->wtab_cap = (2 * (size_t)((params->max_width + XOFF + EDGE + 127) / 128 * 32 + 32) + (size_t)params->max_height + 2 * EDGE) * params->nlevels;
    TRY_ALLOC(cudaMalloc(&ex->d_wtab, ex->wtab_cap * sizeof(uint4)));
    ex->cells_cap = (size_t)(params->max_width / 30 + 2) * (params->max_height / 30 + 2) * 4 + 64;   // sum over levels < 3.1x level 0
    TRY_ALLOC(cudaMalloc(&ex->d_cells, ex->cells_cap * 4));
    ex->src_bytes = B * (size_t)params->max_height * ((params->max_width + 127) / 128 * 128);
    TRY_ALLOC(cudaMalloc(&ex->d_src, ex->src_bytes));
    ex->pyr_bytes = B * (size_t)(G.pyr_frame + 4096 * params->nlevels);
    TRY_ALLOC(cudaMalloc(&ex->d_pyr, ex->pyr_bytes));
    TRY_ALLOC(cudaMemsetAsync(ex->d_pyr, 0, ex->pyr_bytes, ex->stream));
    TRY_ALLOC(cudaMalloc(&ex->d_blur, ex->pyr_bytes));   // blurred levels, same plane geometry
    TRY_ALLOC(cudaMemsetAsync(ex->d_blur, 0, ex->pyr_bytes, ex->stream));
    ex->cand_elems = B * (size_t)(G.cand_frame + 64 * params->nlevels);
    TRY_ALLOC(cudaMalloc(&ex->d_cand, ex->cand_elems * 4));
    TRY_ALLOC(cudaMalloc(&ex->d_node_of, ex->cand_elems * 2));
    TRY_ALLOC(cudaMalloc(&ex->d_counts, B * params->nlevels * 2 * sizeof(uint32_t)));
    ex->fast_rows = getenv("ORBGPU_FAST_CELLS") == nullptr;   // large batches: one CTA per cell-row segment on the big levels
    if (const char *e = getenv("ORBGPU_FAST_ROWS_MIN_PX")) ex->fast_rows_min_px = atoll(e);
    ex->host_trace = getenv("ORBGPU_HOST_TRACE") != nullptr;
    if (const char *e = getenv("ORBGPU_PYR_SMALL_ROWS")) ex->pyr_small_rows = atoi(e);
    if (const char *e = getenv("ORBGPU_OCT_SMALL_THREADS")) ex->oct_small_threads = atoi(e);
    if (const char *e = getenv("ORBGPU_CHUNK_MB")) ex->chunk_bytes = (size_t)std::max(0ll, atoll(e)) << 20;
    for (auto &e : ex->ev_chunk) TRY_ALLOC(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    {
        static std::mutex mu_rows;
        static bool rows_attr[64] = {};
        std::lock_guard<std::mutex> lk(mu_rows);
        if (!rows_attr[ex->device & 63]) {   // cells are < 60 px high: (65 + 61) * 544 + 30 * 512 + 59 * 512 bytes at most
            TRY_ALLOC(cudaFuncSetAttribute(k_fast_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024));
            TRY_ALLOC(cudaFuncSetAttribute(k_fast_rows, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
            rows_attr[ex->device & 63] = true;
        }
    }
    ex->kept_elems = B * kept_frame;
    TRY_ALLOC(cudaMalloc(&ex->d_kept, ex->kept_elems * sizeof(Kept)));
    ex->out_cap_elems = B * kept_frame;
    TRY_ALLOC(cudaMalloc(&ex->d_out_kp, ex->out_cap_elems * sizeof(orb_keypoint)));
    TRY_ALLOC(cudaMalloc(&ex->d_out_desc, ex->out_cap_elems * 32));
    TRY_ALLOC(cudaMalloc(&ex->d_out_n, B * sizeof(int32_t)));
    TRY_ALLOC(cudaMalloc(&ex->d_err, sizeof(int)));
    TRY_ALLOC(cudaMemsetAsync(ex->d_err, 0, sizeof(int), ex->stream));
    TRY_ALLOC(cudaMallocHost(&ex->h_pinned, (1 + B) * sizeof(int)));
    TRY_ALLOC(cudaMallocHost(&ex->h_bounce, ORB_BOUNCE_BYTES));
    if (getenv("ORBGPU_NO_MAPPED_INPUT") == nullptr) TRY_ALLOC(cudaMallocHost(&ex->h_src, ORB_HSRC_BYTES));
    ex->oct_smem = octree_smem_bytes(maxn);
    {
        // cudaFuncAttributeMaxDynamicSharedMemorySize is per function and per device, not per handle: keep the largest request
        // of all live and past handles (a later, smaller handle must not lower it under an earlier one's launches)
        int optin = 0;
        TRY_ALLOC(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, ex->device));
        if (ex->oct_smem + 256 > (size_t)optin) {
            set_error("nfeatures %d: the quadtree needs %zu bytes of shared memory per CTA, the device offers %d (about 15000 features "
                      "at 8 levels / scale 1.2)", params->nfeatures, ex->oct_smem, optin);
            free_all(ex); delete ex;
            return ORB_ERR_INVALID;
        }
        static std::mutex mu;
        static size_t dev_max[64] = {};
        std::lock_guard<std::mutex> lk(mu);
        size_t &cur = dev_max[ex->device & 63];
        if (ex->oct_smem > cur) {
            TRY_ALLOC(cudaFuncSetAttribute(k_octree<OCT_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ex->oct_smem));
            TRY_ALLOC(cudaFuncSetAttribute(k_octree<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ex->oct_smem));
            TRY_ALLOC(cudaFuncSetAttribute(k_octree<1024>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ex->oct_smem));
            cur = ex->oct_smem;
        }
    }
    TRY_ALLOC(cudaStreamSynchronize(ex->stream));
#undef TRY_ALLOC
    *out = ex;
    return ORB_OK;
}

void orb_extractor_destroy(orb_extractor *ex) {
    if (ex && ex->host_trace && ex->trace_calls)
        fprintf(stderr, "orbgpu host trace (us per host-output call over %lld calls): input %.1f | enqueue %.1f | wait %.1f | unpack %.1f\n",
                (long long)ex->trace_calls, ex->trace_us[0] / ex->trace_calls, ex->trace_us[1] / ex->trace_calls,
                ex->trace_us[2] / ex->trace_calls, ex->trace_us[3] / ex->trace_calls);
    if (!ex) return;
    cudaSetDevice(ex->device);
    cudaStreamSynchronize(ex->stream);
    free_all(ex);
    delete ex;
}

int orb_extractor_tables(const orb_extractor *ex, float *scale, float *inv_scale, float *sigma2, float *inv_sigma2,
                         int32_t *features_per_level) {
    ORB_REQUIRE(ex, ORB_ERR_INVALID, "null handle");
    for (int i = 0; i < ex->prm.nlevels; i++) {
        if (scale) scale[i] = ex->scale[i];
        if (inv_scale) inv_scale[i] = ex->inv_scale[i];
        if (sigma2) sigma2[i] = ex->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = ex->inv_sigma2[i];
        if (features_per_level) features_per_level[i] = ex->per_level[i];
    }
    return ORB_OK;
}

void *orb_extractor_stream(orb_extractor *ex) { return ex ? (void *)ex->stream : nullptr; }
int64_t orb_extractor_launch_count(const orb_extractor *ex) { return ex ? ex->launches : 0; }
int64_t orb_extractor_graph_replays(const orb_extractor *ex) { return ex ? ex->graph_replays : 0; }

int orb_extractor_sync(orb_extractor *ex) {
    ORB_REQUIRE(ex, ORB_ERR_INVALID, "null handle");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    ORB_CUDA_TRY(cudaStreamSynchronize(ex->stream));
    return ORB_OK;
}

static int check_err_flags(int flags) {
    if (flags & ERRF_CAND_OVERFLOW) { set_error("FAST candidate pool overflow"); return ORB_ERR_CAPACITY; }
    if (flags & ERRF_NODE_OVERFLOW) { set_error("quadtree node pool overflow"); return ORB_ERR_CAPACITY; }
    if (flags & ERRF_OUT_OVERFLOW) { set_error("more keypoints than `capacity`; n_out holds the required size"); return ORB_ERR_CAPACITY; }
    return ORB_OK;
}

int orb_extract_batch(orb_extractor *ex, const uint8_t *images, int batch, int width, int height, int pitch,
                      size_t frame_stride, int image_loc, orb_keypoint *keypoints, uint8_t *descriptors, int capacity,
                      int32_t *n_out, int out_loc) {
    ORB_REQUIRE(ex && n_out, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(batch >= 1 && batch <= ex->prm.max_batch, ORB_ERR_INVALID, "batch %d exceeds max_batch %d", batch, ex->prm.max_batch);
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    if (width == 0 || height == 0 || !images) {   // ref :987-988: empty image, silent return
        if (out_loc == ORB_HOST) for (int b = 0; b < batch; b++) n_out[b] = 0;
        else ORB_CUDA_TRY(cudaMemsetAsync(n_out, 0, batch * sizeof(int32_t), ex->stream));
        ex->last_batch = 0;
        ex->last_kp = nullptr;
        return ORB_OK;
    }
    ORB_REQUIRE(width > 0 && height > 0 && pitch >= width, ORB_ERR_INVALID, "bad image geometry");
    ORB_REQUIRE(width <= ex->prm.max_width && height <= ex->prm.max_height, ORB_ERR_INVALID,
                "image %dx%d exceeds the handle's max %dx%d", width, height, ex->prm.max_width, ex->prm.max_height);
    ORB_REQUIRE(capacity >= 1 && keypoints && descriptors, ORB_ERR_INVALID, "output buffers missing");
    int rc = set_geometry(ex, width, height);
    if (rc != ORB_OK) return rc;
    const Geom &G = ex->geom;
    cudaStream_t st = ex->stream;

    prof_collect(ex);
    prof_mark(ex, 0);
    const bool trace = ex->host_trace;
    auto now_us = []() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3; };
    const double tr0 = trace ? now_us() : 0;
    const uint8_t *d_images = images;
    size_t d_stride = frame_stride;
    int d_pitch = pitch;
    const bool small_call = (size_t)batch * width * height <= ((size_t)4 << 20);   // launches, not kernels, are the wait
    if (image_loc == ORB_HOST) {
        NvtxRange nv("orbgpu:h2d");
        const bool linear = (size_t)pitch * height == frame_stride || batch == 1;
        const size_t lin_bytes = (size_t)pitch * height * (batch - 1) + (size_t)pitch * (height - 1) + width;
        if (linear && small_call && out_loc == ORB_HOST && ex->h_src && lin_bytes <= ORB_HSRC_BYTES) {   // (a blocking call)
            // latency mode: the frame goes into the handle's pinned, device-mapped buffer with a plain memcpy and k_pyr_level0
            // reads it across PCIe -- no DMA copy in front of the first kernel (a pageable cudaMemcpyAsync of a VGA frame was
            // 29 us of the call: the driver's staging copy plus the copy's own latency on the stream)
            memcpy(ex->h_src, images, lin_bytes);
            d_pitch = pitch;
            d_stride = (size_t)pitch * height;
            d_images = ex->h_src;
        } else if (linear && (size_t)pitch * height * batch <= ex->src_bytes) {
            // the frames as they lie in host memory, row pitch included, in ONE linear copy (k_pyr_level0 takes any pitch): a
            // pitched copy from pageable memory goes row by row through the driver (measured 182 us for one 1241x376 frame)
            d_pitch = pitch;
            d_stride = (size_t)pitch * height;
            // (the last row of a view into a larger buffer holds `width` bytes, not `pitch`)
            ORB_CUDA_TRY(cudaMemcpyAsync(ex->d_src, images, lin_bytes, cudaMemcpyHostToDevice, st));
            d_images = ex->d_src;
        } else {
            d_pitch = (width + 127) / 128 * 128;
            d_stride = (size_t)d_pitch * height;
            for (int b = 0; b < batch; b++)
                ORB_CUDA_TRY(cudaMemcpy2DAsync(ex->d_src + b * d_stride, d_pitch, images + b * frame_stride, pitch, width, height,
                                               cudaMemcpyHostToDevice, st));
            d_images = ex->d_src;
        }
    }
    const double tr1 = trace ? now_us() : 0;
    orb_keypoint *o_kp = keypoints;
    uint8_t *o_desc = descriptors;
    int32_t *o_n = n_out;
    if (out_loc == ORB_HOST) {
        rc = orbgpu_ensure_out_staging(ex, batch, capacity);
        if (rc != ORB_OK) return rc;
        o_kp = ex->d_out_kp; o_desc = ex->d_out_desc; o_n = ex->d_out_n;
    }
    // small host-output calls: the results reach pinned host memory by one kernel at the end of the launch sequence
    const bool export_host = out_loc == ORB_HOST && ex->h_bounce && (size_t)batch * capacity * (sizeof(orb_keypoint) + 32) <= ORB_BOUNCE_BYTES;

    uint32_t *d_cand_count = ex->d_counts;
    uint32_t *d_kept_count = ex->d_counts + (size_t)ex->prm.max_batch * G.nlevels;

    // every per-frame array is addressed as base + frame * stride inside the kernels, so a sub-batch [b0, b0 + nb) of the call
    // is the same launch with the bases moved (only the TMA tile of k_fast_rows takes the first frame as a coordinate)
    auto launch_pyr = [&](int l, int b0, int nb) {
        NvtxRange nv("orbgpu:pyramid");
        const LevelGeom &g = G.lv[l];
        uint8_t *pyr = ex->d_pyr + (size_t)b0 * G.pyr_frame;
        if (l == 0) {
            const int qL = (XOFF + 16) / 16, qR = std::max(qL, (g.w - 19 + XOFF) / 16 + 1);   // as in the kernel
            dim3 grid((qR - qL + 31) / 32 + 1, (g.h + 2 * EDGE + 4 * L0_ROWS - 1) / (4 * L0_ROWS), nb);
            k_pyr_level0<<<grid, 128, 0, st>>>(d_images + (size_t)b0 * d_stride, d_stride, d_pitch, pyr, ex->d_geom);
        } else if (g.fast_resize) {
            const int nw = g.pitch / 4;
            int bx = 128;   // CTA width that wastes the fewest threads on this level's row
            for (int c = 96; c >= 32; c -= 32)
                if ((nw + c - 1) / c * c < (nw + bx - 1) / bx * bx) bx = c;
            if (small_call && ex->pyr_small_rows == 1) {
                dim3 grid((nw + bx - 1) / bx, g.h + 2 * EDGE, nb);
                k_pyr_resize<1><<<grid, bx, 0, st>>>(pyr, ex->d_geom, ex->d_taps, ex->d_wtab, l);
            } else if (small_call && ex->pyr_small_rows == 2) {
                dim3 grid((nw + bx - 1) / bx, (g.h + 2 * EDGE + 1) / 2, nb);
                k_pyr_resize<2><<<grid, bx, 0, st>>>(pyr, ex->d_geom, ex->d_taps, ex->d_wtab, l);
            } else if (small_call && ex->pyr_small_rows == 4) {
                dim3 grid((nw + bx - 1) / bx, (g.h + 2 * EDGE + 3) / 4, nb);
                k_pyr_resize<4><<<grid, bx, 0, st>>>(pyr, ex->d_geom, ex->d_taps, ex->d_wtab, l);
            } else {
                dim3 grid((nw + bx - 1) / bx, (g.h + 2 * EDGE + PYR_ROWS - 1) / PYR_ROWS, nb);
                k_pyr_resize<PYR_ROWS><<<grid, bx, 0, st>>>(pyr, ex->d_geom, ex->d_taps, ex->d_wtab, l);
            }
        } else {
            dim3 grid((g.pitch / 4 + 127) / 128, g.h + 2 * EDGE, nb);
            k_pyr_resize_generic<<<grid, 128, 0, st>>>(pyr, ex->d_geom, ex->d_taps, l);
        }
        ex->launches++;
    };
    // Large batches: levels with many pixels go through k_fast_rows (one CTA per segment of a cell row: 26 % fewer warp
    // instructions at level 0 of a 1241x376 frame, 236 us against 274 us per 288 frames); the small, corner-dense levels stay on
    // the per-cell kernel, whose many 2-warp CTAs keep more warps in flight (profiles/r2_fast_cells_vs_rows.txt).
    const bool rows_path = !small_call && ex->fast_rows;
    auto launch_fast = [&](int l, cudaStream_t s, int b0, int nb) {
        NvtxRange nv("orbgpu:fast_cells");
        const LevelGeom &g = G.lv[l];
        FastLevel fl;
        fl.level = l; fl.nlevels = G.nlevels; fl.ini_th = G.ini_th; fl.min_th = G.min_th;
        fl.pitch = g.pitch; fl.maxBX = g.maxBX; fl.maxBY = g.maxBY; fl.wCell = g.wCell; fl.hCell = g.hCell;
        fl.cand_off = g.cand_off; fl.cand_cap = g.cand_cap; fl.cand_frame = G.cand_frame;
        fl.interior_off = g.interior_off; fl.pyr_frame = G.pyr_frame;
        fl.w = g.w; fl.h = g.h; fl.ncx = g.ncx; fl.ncy = g.ncy;
        const size_t smem = fast_smem_layout(fl);
        const uint8_t *pyr = ex->d_pyr + (size_t)b0 * G.pyr_frame;
        uint32_t *cand = ex->d_cand + (size_t)b0 * G.cand_frame, *cnt = d_cand_count + (size_t)b0 * G.nlevels;
        if (rows_path && (long long)(g.w - 2 * EDGE) * (g.h - 2 * EDGE) >= ex->fast_rows_min_px) {
            FastRowsLayout lay;
            const size_t smem_rows = std::max(smem, fast_rows_layout(fl, lay));
            lay.use_tma = ex->tmap_ok[l] ? 1 : 0;
            lay.frame0 = b0;
            k_fast_rows<<<dim3((g.ncx + lay.seg_cells - 1) / lay.seg_cells, g.ncy, nb), FR_T, smem_rows, s>>>(
                pyr, fl, lay, ex->tmap[l], cand, cnt, ex->d_err);
        } else {
            k_fast_cells<<<dim3(g.ncx, g.ncy, nb), FAST_T, smem, s>>>(pyr, fl, cand, cnt, ex->d_err);
        }
        ex->launches++;
    };
    auto launch_blur = [&](int l, cudaStream_t s, int b0, int nb) {
        NvtxRange nv("orbgpu:blur");
        const LevelGeom &g = G.lv[l];
        dim3 grid(((g.w + 3) / 4 + 127) / 128, (g.h + BLUR_ROWS - 1) / BLUR_ROWS, nb);
        if (G.blur_taps == 1) k_blur_planes<1><<<grid, 128, 0, s>>>(ex->d_pyr + (size_t)b0 * G.pyr_frame, ex->d_blur + (size_t)b0 * G.pyr_frame, ex->d_geom, l);
        else k_blur_planes<0><<<grid, 128, 0, s>>>(ex->d_pyr + (size_t)b0 * G.pyr_frame, ex->d_blur + (size_t)b0 * G.pyr_frame, ex->d_geom, l);
        ex->launches++;
    };
    auto launch_octree = [&](int l0, int nl, cudaStream_t s, int b0, int nb) {
        NvtxRange nv("orbgpu:octree");
        auto kern = small_call && ex->oct_small_threads == 1024 ? k_octree<1024> : small_call && ex->oct_small_threads == 512 ? k_octree<512> : k_octree<OCT_THREADS>;
        const int nt = small_call && (ex->oct_small_threads == 1024 || ex->oct_small_threads == 512) ? ex->oct_small_threads : OCT_THREADS;
        kern<<<dim3(nl, nb), nt, octree_smem_bytes(G.max_nodes), s>>>(
            ex->d_geom, ex->d_cand + (size_t)b0 * G.cand_frame, d_cand_count + (size_t)b0 * G.nlevels,
            ex->d_node_of + (size_t)b0 * G.cand_frame, ex->d_kept + (size_t)b0 * G.kept_frame, d_kept_count + (size_t)b0 * G.nlevels,
            ex->d_err, l0);
        ex->launches++;
    };
    DescParams dp;
    dp.nlevels = G.nlevels; dp.kept_frame = G.kept_frame; dp.pyr_frame = G.pyr_frame;
    for (int l = 0, acc = 0; l < ORB_MAX_LEVELS; l++) {
        const LevelGeom &lg = G.lv[l < G.nlevels ? l : G.nlevels - 1];
        if (l < G.nlevels) acc += lg.kept_cap;
        dp.lv[l].interior_off = lg.interior_off; dp.lv[l].cap_end = acc; dp.lv[l].kept_off = lg.kept_off; dp.lv[l].pitch = lg.pitch;
        dp.lv[l].scale = lg.scale; dp.lv[l].size = lg.size;
    }
    auto launch_describe = [&](cudaStream_t s, int b0, int nb) {
        NvtxRange nv("orbgpu:describe");
        k_describe<<<dim3((G.kept_frame + DESC_WARPS - 1) / DESC_WARPS, nb), DESC_WARPS * 32, 0, s>>>(
            ex->d_pyr + (size_t)b0 * G.pyr_frame, ex->d_blur + (size_t)b0 * G.pyr_frame, dp, ex->d_kept + (size_t)b0 * G.kept_frame,
            d_kept_count + (size_t)b0 * G.nlevels, o_kp + (size_t)b0 * capacity, o_desc + (size_t)b0 * capacity * 32, o_n + b0, capacity,
            ex->d_err);
        ex->launches++;
    };
    // Large batches run as sub-batches whose planes (pyramid + blurred pyramid) stay in L2 from the kernel that writes them to
    // the kernels that read them: the resize cascade of sub-batch c+1 runs on the handle's stream while FAST, the quadtree and
    // blur + describe of sub-batch c run on three side streams; at most two sub-batches are in flight.
    int chunk = 0;
    if (!small_call && ex->overlap && !ex->profiling && ex->chunk_bytes > 0) {
        chunk = (int)std::max<size_t>(8, ex->chunk_bytes / (2 * (size_t)G.pyr_frame));
        if (chunk >= batch) chunk = 0;
        else chunk = (batch + (batch + chunk - 1) / chunk - 1) / ((batch + chunk - 1) / chunk);   // equal sub-batches
    }
    auto enqueue = [&]() -> int {   // everything of one extraction between the input copy and the output copy
        ORB_CUDA_TRY(cudaMemsetAsync(ex->d_counts, 0, (size_t)ex->prm.max_batch * G.nlevels * 2 * sizeof(uint32_t), st));
        ORB_CUDA_TRY(cudaMemsetAsync(o_n, 0, batch * sizeof(int32_t), st));
        if (chunk > 0) {
            cudaStream_t s_oct = ex->s_oct[0];
            for (int b0 = 0, c = 0; b0 < batch; b0 += chunk, c++) {
                const int nb = std::min(chunk, batch - b0);
                if (c >= 2) ORB_CUDA_TRY(cudaStreamWaitEvent(st, ex->ev_chunk[c & 1], 0));   // describe of sub-batch c-2 is done
                for (int l = 0; l < G.nlevels; l++) {
                    launch_pyr(l, b0, nb);
                    ORB_CUDA_TRY(cudaEventRecord(ex->ev_lvl[l], st));
                    ORB_CUDA_TRY(cudaStreamWaitEvent(ex->s_fast, ex->ev_lvl[l], 0));
                    launch_fast(l, ex->s_fast, b0, nb);
                    ORB_CUDA_TRY(cudaStreamWaitEvent(ex->s_blur, ex->ev_lvl[l], 0));
                    launch_blur(l, ex->s_blur, b0, nb);
                }
                ORB_CUDA_TRY(cudaEventRecord(ex->ev_fast, ex->s_fast));
                ORB_CUDA_TRY(cudaStreamWaitEvent(s_oct, ex->ev_fast, 0));
                launch_octree(0, G.nlevels, s_oct, b0, nb);
                ORB_CUDA_TRY(cudaEventRecord(ex->ev_ol[0], s_oct));
                ORB_CUDA_TRY(cudaStreamWaitEvent(ex->s_blur, ex->ev_ol[0], 0));
                launch_describe(ex->s_blur, b0, nb);
                ORB_CUDA_TRY(cudaEventRecord(ex->ev_chunk[c & 1], ex->s_blur));
            }
            ORB_CUDA_TRY(cudaEventRecord(ex->ev_blur, ex->s_blur));   // the last describe: everything before it is ordered in front
            ORB_CUDA_TRY(cudaStreamWaitEvent(st, ex->ev_blur, 0));
            // the host-output tail below copies nothing when export_host is set (it expects the pinned block to be filled here)
            if (export_host) orbgpu_export_results(ex, o_kp, o_desc, o_n, batch, capacity, nullptr, nullptr, nullptr, st);
            return ORB_OK;
        }
        if (ex->overlap && !ex->profiling) {
            // level l's FAST cells and blur depend only on level l: they go to two side streams behind an event per level and
            // overlap the rest of the resize cascade (and each other); the handle's stream joins them where their results are read
            if (small_call) {
                // latency mode: the launches of one level are a chain of their own -- FAST cells, then that level's quadtree (one CTA
                // per frame, 34 us for a VGA level 0) on stream l, the blur on stream l of a second set -- so a level waits for
                // nothing but its own plane; on one side stream the eight FAST launches alone were a 60-us serial chain.
                auto level_chains = [&](int l, cudaEvent_t ready) -> int {
                    ORB_CUDA_TRY(cudaStreamWaitEvent(ex->s_oct[l], ready, 0));
                    launch_fast(l, ex->s_oct[l], 0, batch);
                    launch_octree(l, 1, ex->s_oct[l], 0, batch);
                    ORB_CUDA_TRY(cudaEventRecord(ex->ev_ol[l], ex->s_oct[l]));
                    ORB_CUDA_TRY(cudaStreamWaitEvent(ex->s_bl[l], ready, 0));
                    launch_blur(l, ex->s_bl[l], 0, batch);
                    ORB_CUDA_TRY(cudaEventRecord(ex->ev_fl[l], ex->s_bl[l]));
                    return ORB_OK;
                };
                for (int l = 0; l < G.nlevels; l++) {
                    launch_pyr(l, 0, batch);
                    ORB_CUDA_TRY(cudaEventRecord(ex->ev_lvl[l], st));
                    if (int r = level_chains(l, ex->ev_lvl[l])) return r;
                }
                for (int l = 0; l < G.nlevels; l++) {
                    ORB_CUDA_TRY(cudaStreamWaitEvent(st, ex->ev_ol[l], 0));
                    ORB_CUDA_TRY(cudaStreamWaitEvent(st, ex->ev_fl[l], 0));
                }
            } else {
                for (int l = 0; l < G.nlevels; l++) {
                    launch_pyr(l, 0, batch);
                    ORB_CUDA_TRY(cudaEventRecord(ex->ev_lvl[l], st));
                    ORB_CUDA_TRY(cudaStreamWaitEvent(ex->s_fast, ex->ev_lvl[l], 0));
                    launch_fast(l, ex->s_fast, 0, batch);
                    ORB_CUDA_TRY(cudaStreamWaitEvent(ex->s_blur, ex->ev_lvl[l], 0));
                    launch_blur(l, ex->s_blur, 0, batch);
                }
                ORB_CUDA_TRY(cudaEventRecord(ex->ev_blur, ex->s_blur));
                ORB_CUDA_TRY(cudaEventRecord(ex->ev_fast, ex->s_fast));
                ORB_CUDA_TRY(cudaStreamWaitEvent(st, ex->ev_fast, 0));
                launch_octree(0, G.nlevels, st, 0, batch);   // (one launch per level on a side stream was measured at 288 frames: no gain)
                ORB_CUDA_TRY(cudaStreamWaitEvent(st, ex->ev_blur, 0));
            }
        } else {
            prof_mark(ex, 1);
            for (int l = 0; l < G.nlevels; l++) launch_pyr(l, 0, batch);
            prof_mark(ex, 2);
            for (int l = 0; l < G.nlevels; l++) launch_fast(l, st, 0, batch);
            prof_mark(ex, 3);
            launch_octree(0, G.nlevels, st, 0, batch);
            prof_mark(ex, 4);
            for (int l = 0; l < G.nlevels; l++) launch_blur(l, st, 0, batch);
        }
        launch_describe(st, 0, batch);
        if (export_host) orbgpu_export_results(ex, o_kp, o_desc, o_n, batch, capacity, nullptr, nullptr, nullptr, st);
        return ORB_OK;
    };
    const orb_extractor::GraphKey key = {d_images, d_stride, d_pitch, batch, width, height, capacity, o_kp, o_desc, o_n, export_host ? 1 : 0};
    // a graph pays where the launches themselves are the wait (single frames, stereo pairs); large batches measured 1 %
    // slower as a graph than as stream launches (92.5 -> 91.7 k frames/s at 288 frames), so they keep the streams
    const bool graphable = ex->use_graph && ex->overlap && !ex->profiling && small_call;
    if (graphable && ex->gexec && key == ex->gkey) {
        ORB_CUDA_TRY(cudaGraphLaunch(ex->gexec, st));
        ex->launches += ex->g_nodes;
        ex->graph_replays++;
    } else if (graphable && key == ex->glast) {
        // second call with these arguments: capture (nothing runs), instantiate, launch
        if (ex->gexec) { cudaGraphExecDestroy(ex->gexec); ex->gexec = nullptr; }
        const int64_t l0 = ex->launches;
        cudaGraph_t graph = nullptr;
        ORB_CUDA_TRY(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        const int erc = enqueue();
        const cudaError_t ce = cudaStreamEndCapture(st, &graph);
        if (erc != ORB_OK || ce != cudaSuccess || !graph) {
            if (graph) cudaGraphDestroy(graph);
            (void)cudaGetLastError();
            ex->use_graph = false;   // capture refused: stay on plain launches for the rest of the handle's life
            ex->launches = l0;
            rc = enqueue();
            if (rc != ORB_OK) return rc;
        } else {
            ex->g_nodes = (int)(ex->launches - l0);
            const cudaError_t ie = cudaGraphInstantiate(&ex->gexec, graph, 0);
            cudaGraphDestroy(graph);
            if (ie != cudaSuccess) {
                (void)cudaGetLastError();
                ex->gexec = nullptr; ex->use_graph = false; ex->launches = l0;
                rc = enqueue();
                if (rc != ORB_OK) return rc;
            } else {
                ex->gkey = key;
                ORB_CUDA_TRY(cudaGraphLaunch(ex->gexec, st));
            }
        }
    } else {
        rc = enqueue();
        if (rc != ORB_OK) return rc;
    }
    ex->glast = key;
    prof_mark(ex, 5);
    ORB_CUDA_TRY(cudaGetLastError());
    ex->last_batch = batch;
    ex->last_kp = o_kp; ex->last_desc = o_desc; ex->last_n = o_n; ex->last_capacity = capacity;
    ex->last_counts_valid = false;

    if (out_loc == ORB_HOST) {
        NvtxRange nv("orbgpu:d2h");
        // results: small ones (single frames, stereo pairs) are already on their way into the handle's pinned buffers (the export
        // kernel above) and only the n valid entries of each frame are handed on; copies straight into pageable caller memory
        // are synchronous staging round trips in the driver (measured 54 us of a 238 us VGA call).  Large batches copy straight
        // to the caller.
        const size_t kp_bytes = (size_t)batch * capacity * sizeof(orb_keypoint), desc_bytes = (size_t)batch * capacity * 32;
        const bool bounce = export_host;
        if (!bounce) {
            ORB_CUDA_TRY(cudaMemcpyAsync(ex->h_pinned, ex->d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
            ORB_CUDA_TRY(cudaMemcpyAsync(ex->h_pinned + 1, o_n, batch * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
            ORB_CUDA_TRY(cudaMemcpyAsync((void *)keypoints, o_kp, kp_bytes, cudaMemcpyDeviceToHost, st));
            ORB_CUDA_TRY(cudaMemcpyAsync((void *)descriptors, o_desc, desc_bytes, cudaMemcpyDeviceToHost, st));
        }
        prof_mark(ex, 6);
        ex->ev_pending = ex->profiling;
        const double tr2 = trace ? now_us() : 0;
        ORB_CUDA_TRY(cudaStreamSynchronize(st));
        const double tr3 = trace ? now_us() : 0;
        if (bounce)
            for (int b = 0; b < batch; b++) {
                const size_t n = (size_t)std::min(std::max(ex->h_pinned[1 + b], 0), capacity);
                memcpy(keypoints + (size_t)b * capacity, ex->h_bounce + (size_t)b * capacity * sizeof(orb_keypoint), n * sizeof(orb_keypoint));
                memcpy(descriptors + (size_t)b * capacity * 32, ex->h_bounce + kp_bytes + (size_t)b * capacity * 32, n * 32);
            }
        if (trace) {
            const double tr4 = now_us();
            ex->trace_us[0] += tr1 - tr0; ex->trace_us[1] += tr2 - tr1; ex->trace_us[2] += tr3 - tr2; ex->trace_us[3] += tr4 - tr3;
            ex->trace_calls++;
        }
        for (int b = 0; b < batch; b++) n_out[b] = ex->h_pinned[1 + b];
        ex->last_counts.assign(ex->h_pinned + 1, ex->h_pinned + 1 + batch);
        ex->last_counts_valid = true;
        const int flags = ex->h_pinned[0];
        if (flags) {
            cudaMemsetAsync(ex->d_err, 0, sizeof(int), st);
            return check_err_flags(flags);
        }
    }
    prof_mark(ex, 6);
    ex->ev_pending = ex->profiling;
    return ORB_OK;
}

int orb_extractor_set_profiling(orb_extractor *ex, int on) {
    ORB_REQUIRE(ex, ORB_ERR_INVALID, "null handle");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    if (on && !ex->ev[0])
        for (auto &e : ex->ev) ORB_CUDA_TRY(cudaEventCreate(&e));
    ex->profiling = on != 0;
    ex->ev_pending = false;
    for (auto &m : ex->stage_ms) m = 0;
    ex->stage_calls = 0;
    return ORB_OK;
}

int orb_extractor_stage_times(orb_extractor *ex, double *ms, int64_t *calls) {
    ORB_REQUIRE(ex && ms, ORB_ERR_INVALID, "null argument");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    prof_collect(ex);
    for (int i = 0; i < ORB_N_STAGES; i++) ms[i] = ex->stage_ms[i];
    if (calls) *calls = ex->stage_calls;
    return ORB_OK;
}

int orb_extract(orb_extractor *ex, const uint8_t *image, int width, int height, int pitch, int image_loc,
                orb_keypoint *keypoints, uint8_t *descriptors, int capacity, int *n_out) {
    int32_t n = 0;
    int rc = orb_extract_batch(ex, image, 1, width, height, pitch, (size_t)pitch * height, image_loc, keypoints, descriptors,
                               capacity, &n, ORB_HOST);
    if (n_out) *n_out = n;
    return rc;
}

/* Error flags raised by device-output (asynchronous) batches; clears them. */
int orb_extractor_check(orb_extractor *ex) {
    ORB_REQUIRE(ex, ORB_ERR_INVALID, "null handle");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    ORB_CUDA_TRY(cudaMemcpyAsync(ex->h_pinned, ex->d_err, sizeof(int), cudaMemcpyDeviceToHost, ex->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(ex->stream));
    const int flags = ex->h_pinned[0];
    if (flags) {
        cudaMemsetAsync(ex->d_err, 0, sizeof(int), ex->stream);
        return check_err_flags(flags);
    }
    return ORB_OK;
}

int orb_pyramid_level(const orb_extractor *ex, int slot, int level, const uint8_t **dev_interior, int *width, int *height,
                      int *pitch) {
    ORB_REQUIRE(ex, ORB_ERR_INVALID, "null handle");
    ORB_REQUIRE(ex->geom.width > 0, ORB_ERR_INVALID, "no frame extracted yet");
    ORB_REQUIRE(level >= 0 && level < ex->geom.nlevels && slot >= 0 && slot < ex->prm.max_batch, ORB_ERR_INVALID, "bad level/slot");
    const LevelGeom &g = ex->geom.lv[level];
    if (dev_interior) *dev_interior = ex->d_pyr + (size_t)slot * ex->geom.pyr_frame + g.interior_off;
    if (width) *width = g.w;
    if (height) *height = g.h;
    if (pitch) *pitch = g.pitch;
    return ORB_OK;
}

int orb_pyramid_download(orb_extractor *ex, int slot, int level, int with_border, uint8_t *dst, int dst_pitch) {
    const uint8_t *p;
    int w, h, pitch;
    int rc = orb_pyramid_level(ex, slot, level, &p, &w, &h, &pitch);
    if (rc != ORB_OK) return rc;
    ORB_REQUIRE(dst, ORB_ERR_INVALID, "null dst");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    if (with_border) {
        p -= (size_t)EDGE * pitch + EDGE;
        w += 2 * EDGE; h += 2 * EDGE;
    }
    ORB_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_pitch, p, pitch, w, h, cudaMemcpyDeviceToHost, ex->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(ex->stream));
    return ORB_OK;
}

int orb_debug_candidates(orb_extractor *ex, int slot, int level, orb_candidate *out, int capacity, int *n_out) {
    ORB_REQUIRE(ex && out && n_out, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(ex->geom.width > 0 && level >= 0 && level < ex->geom.nlevels && slot >= 0 && slot < ex->prm.max_batch,
                ORB_ERR_INVALID, "bad level/slot");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    const LevelGeom &g = ex->geom.lv[level];
    uint32_t n = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(&n, ex->d_counts + (size_t)slot * ex->geom.nlevels + level, 4, cudaMemcpyDeviceToHost, ex->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(ex->stream));
    n = std::min<uint32_t>(n, g.cand_cap);
    std::vector<uint32_t> raw(n);
    if (n) {
        ORB_CUDA_TRY(cudaMemcpyAsync(raw.data(), ex->d_cand + (size_t)slot * ex->geom.cand_frame + g.cand_off, n * 4ull,
                                     cudaMemcpyDeviceToHost, ex->stream));
        ORB_CUDA_TRY(cudaStreamSynchronize(ex->stream));
    }
    *n_out = (int)n;
    ORB_REQUIRE((int)n <= capacity, ORB_ERR_CAPACITY, "candidate buffer too small");
    // the device list is unordered; restore the reference's order (cells row-major, then row-major)
    std::vector<std::pair<uint64_t, uint32_t>> keyed(n);
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t x = raw[i] & 0xfff, y = (raw[i] >> 12) & 0xfff;
        const uint64_t ci = (y - 3) / g.hCell, cj = (x - 3) / g.wCell;
        keyed[i] = {(ci << 32) | (cj << 24) | ((uint64_t)y << 12) | x, raw[i]};
    }
    std::sort(keyed.begin(), keyed.end());
    for (uint32_t i = 0; i < n; i++) {
        out[i].x = (int16_t)(keyed[i].second & 0xfff);
        out[i].y = (int16_t)((keyed[i].second >> 12) & 0xfff);
        out[i].response = (int32_t)(keyed[i].second >> 24);
    }
    return ORB_OK;
}

int orb_debug_blur_level(orb_extractor *ex, int slot, int level, uint8_t *dst, int dst_pitch) {
    ORB_REQUIRE(ex && dst, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(ex->geom.width > 0 && level >= 0 && level < ex->geom.nlevels && slot >= 0 && slot < ex->prm.max_batch,
                ORB_ERR_INVALID, "bad level/slot");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    const LevelGeom &g = ex->geom.lv[level];
    // the blurred planes of the last extraction are resident; hand back the interior of the requested one
    ORB_CUDA_TRY(cudaMemcpy2DAsync(dst, dst_pitch, ex->d_blur + (size_t)slot * ex->geom.pyr_frame + g.interior_off, g.pitch, g.w, g.h,
                                   cudaMemcpyDeviceToHost, ex->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(ex->stream));
    return ORB_OK;
}

}  // extern "C"
