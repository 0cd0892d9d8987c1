// liborbgpu -- stereo association for sm_100a.
//
// Replaces Frame::ComputeStereoMatches (ref: ORB_SLAM2/src/Frame.cc:471-643).  One warp per left
// keypoint: row-band Hamming search over the right keypoints (lexicographic (distance, index)
// warp minimum == the reference's first-best scan), then the 11x11 SAD slide over +-5 px on the
// keypoint's pyramid level of both extractors, the parabola refinement and the depth.  A second
// one-CTA kernel per pair applies the median-SAD rejection (ref :629-642).
#include "orb_common.cuh"
#include "orb_internal.cuh"

#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

namespace orbgpu {

constexpr int ST_WARPS = 8;
constexpr int TH_HIGH = 100;   // ref ORBmatcher.cc:37

struct StereoArgs {
    const uint8_t *pyrL, *pyrR;
    const Geom *GL, *GR;
    int slotL0, slotR0;
    const orb_keypoint *keysL, *keysR;
    const uint8_t *descL, *descR;
    const int32_t *nL, *nR;
    int capacity;
    float mbf, mb;
    float *uRight, *depth;
    int4 *packR;           // [batch][capacity] right keypoints packed by k_stereo_prep
    int *rowoff;           // [batch][nrows + 1] CSR of the row table (ref vRowIndices, Frame.cc:479-496)
    int *rowlist;          // [batch][list_cap] right keypoint indices per image row
    int nrows, list_cap;
    int32_t *sad;          // [batch][capacity] best SAD or -1
    int32_t *n_matched;    // [batch]
};

// The reference's row table (ref :479-496): right keypoint iR is a candidate for every image row in
// [floor(y - r), ceil(y + r)], r = 2 * scale[octave].  One CTA per pair builds it as CSR in three passes over the right
// keypoints (count with shared atomics, block scan, fill) and packs {minr, maxr, octave, x} per keypoint.  The order
// inside a row is arbitrary: the search below takes the minimum of (distance, index), which is the reference's
// first-best-in-index-order result.
constexpr int ROWTAB_THREADS = 1024;
__global__ void __launch_bounds__(ROWTAB_THREADS) k_stereo_rowtable(StereoArgs A) {
    extern __shared__ int s_row[];          // [nrows + 1] counts, then offsets, then cursors
    __shared__ int s_warp[33];
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31;
    const int Nr = min(A.nR[b], A.capacity), nrows = A.nrows;
    for (int i = tid; i <= nrows; i += ROWTAB_THREADS) s_row[i] = 0;
    __syncthreads();
    for (int i = tid; i < Nr; i += ROWTAB_THREADS) {
        const orb_keypoint kr = A.keysR[(size_t)b * A.capacity + i];
        const float r = __fmul_rn(2.0f, A.GL->lv[kr.octave].scale);
        const int maxr = (int)ceilf(__fadd_rn(kr.y, r)), minr = (int)floorf(__fsub_rn(kr.y, r));
        A.packR[(size_t)b * A.capacity + i] = make_int4(minr, maxr, kr.octave, __float_as_int(kr.x));
        for (int y = max(minr, 0); y <= min(maxr, nrows - 1); y++) atomicAdd(&s_row[y], 1);
    }
    __syncthreads();
    // exclusive scan of s_row[0..nrows) in place
    int carry = 0;
    for (int base = 0; base < nrows; base += ROWTAB_THREADS) {
        const int i = base + tid;
        const int v = i < nrows ? s_row[i] : 0;
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        if (lane == 31) s_warp[tid >> 5] = incl;
        __syncthreads();
        if (tid < 32) {
            const int w = s_warp[tid];
            int wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, wi, o);
                if (tid >= o) wi += t;
            }
            s_warp[tid] = wi - w;
            if (tid == 31) s_warp[32] = wi;
        }
        __syncthreads();
        if (i < nrows) s_row[i] = carry + s_warp[tid >> 5] + incl - v;
        carry += s_warp[32];
        __syncthreads();
    }
    if (tid == 0) s_row[nrows] = carry;
    __syncthreads();
    int *off = A.rowoff + (size_t)b * (nrows + 1);
    for (int i = tid; i <= nrows; i += ROWTAB_THREADS) off[i] = s_row[i];
    __syncthreads();
    int *list = A.rowlist + (size_t)b * A.list_cap;
    for (int i = tid; i < Nr; i += ROWTAB_THREADS) {
        const int4 k = A.packR[(size_t)b * A.capacity + i];
        for (int y = max(k.x, 0); y <= min(k.y, nrows - 1); y++) {
            const int pos = atomicAdd(&s_row[y], 1);     // s_row doubles as the fill cursor
            if (pos < A.list_cap) list[pos] = i;
        }
    }
}

// latency-bound (long-scoreboard stalls on the candidate and patch loads): 6 resident CTAs of 8 warps (42 registers, 12 bytes
// spilled) instead of 5 (48 registers) measured 0.376 -> 0.314 ms per 144 pairs; 8 CTAs (32 registers, 72 bytes spilled) 0.316
#ifndef ORBGPU_ST_MINB
#define ORBGPU_ST_MINB 6
#endif
__global__ void __launch_bounds__(ST_WARPS * 32, ORBGPU_ST_MINB) k_stereo_match(StereoArgs A) {
    __shared__ uint8_t s_IR[ST_WARPS][11 * 24];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.y;
    const int iL = blockIdx.x * ST_WARPS + warp;
    const int N = min(A.nL[b], A.capacity);
    if (iL >= A.capacity) return;
    const size_t o = (size_t)b * A.capacity + iL;
    if (lane == 0) { A.uRight[o] = -1.0f; A.depth[o] = -1.0f; A.sad[o] = -1; }
    if (iL >= N) return;

    const orb_keypoint kpL = A.keysL[o];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const float minZ = A.mb, minD = -3.f;
    const float maxD = __fdiv_rn(A.mbf, minZ);                 // ref :499-501
    const float minU = __fsub_rn(uL, maxD), maxU = __fsub_rn(uL, minD);
    if (maxU < 0) return;                                      // ref :522-523
    const int row = (int)vL;                                   // vRowIndices[vL]

    uint32_t dl[8];
    {
        const uint32_t *p = reinterpret_cast<const uint32_t *>(A.descL + o * 32);
#pragma unroll
        for (int i = 0; i < 8; i++) dl[i] = __ldg(p + i);
    }
    // ---- Hamming search in the row band (ref :479-496, 525-552)
    uint32_t best = 0xffffffffu;   // (dist << 16) | iR
    const orb_keypoint *KR = A.keysR + (size_t)b * A.capacity;
    const uint8_t *DR = A.descR + (size_t)b * A.capacity * 32;
    const int4 *PR = A.packR + (size_t)b * A.capacity;
    if (row < 0 || row >= A.nrows) return;
    const int *off = A.rowoff + (size_t)b * (A.nrows + 1);
    const int lo = off[row], hi = min(off[row + 1], A.list_cap);
    const int *list = A.rowlist + (size_t)b * A.list_cap;
    for (int j = lo + lane; j < hi; j += 32) {                    // vCandidates = vRowIndices[vL]
        const int iR = list[j];
        const int4 kr = __ldg(PR + iR);                           // {minr, maxr, octave, x}
        if (kr.z < levelL - 1 || kr.z > levelL + 1) continue;
        const float uR = __int_as_float(kr.w);
        if (!(uR >= minU && uR <= maxU)) continue;
        const uint4 *p = reinterpret_cast<const uint4 *>(DR + (size_t)iR * 32);
        const uint4 q0 = __ldg(p), q1 = __ldg(p + 1);
        const int d = __popc(dl[0] ^ q0.x) + __popc(dl[1] ^ q0.y) + __popc(dl[2] ^ q0.z) + __popc(dl[3] ^ q0.w) +
                      __popc(dl[4] ^ q1.x) + __popc(dl[5] ^ q1.y) + __popc(dl[6] ^ q1.z) + __popc(dl[7] ^ q1.w);
        best = min(best, ((uint32_t)d << 16) | (uint32_t)iR);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, s));
    const int bestDist = best == 0xffffffffu ? TH_HIGH : (int)(best >> 16);
    if (bestDist >= TH_HIGH) return;
    const int bestIdxR = best & 0xffff;

    // ---- SAD slide on the pyramid level of the left keypoint (ref :557-595)
    const float uR0 = KR[bestIdxR].x;
    const LevelGeom &gl = A.GL->lv[levelL];
    const LevelGeom &gr = A.GR->lv[levelL];
    const float sf = gl.inv_scale;
    const float scaleduL = roundf(__fmul_rn(kpL.x, sf));
    const float scaledvL = roundf(__fmul_rn(kpL.y, sf));
    const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
    const int w = 5, L = 5;
    const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
    if (iniu < 0 || endu >= (float)gr.w) return;               // ref :576-579

    const uint8_t *imgL = A.pyrL + (size_t)(A.slotL0 + b) * A.GL->pyr_frame + gl.interior_off;
    const uint8_t *imgR = A.pyrR + (size_t)(A.slotR0 + b) * A.GR->pyr_frame + gr.interior_off;
    const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w), xr0 = (int)(scaleduR0 - L - w);
    // The reference's L1 norm of (IL - IL_centre) - (IR - IR_centre) over the 11x11 window (ref :566-586) is a sum of small
    // integers, exact in its fp32 arithmetic; here it is an integer SAD.  A lane owns window pixels lane, lane+32, ...; the 21
    // columns of the right strip go through shared memory because every lane slides over 11 of them.
    uint8_t *IR = s_IR[warp];
    int aL[4], irb[4];
#pragma unroll
    for (int m = 0; m < 4; m++) {
        const int i = lane + 32 * m, y = i / 11, x = i - y * 11;
        irb[m] = y * 24 + x;
        aL[m] = i < 121 ? (int)imgL[(ptrdiff_t)(y0 + y) * gl.pitch + xl0 + x] : 0;
    }
    if (lane < 21) {   // lane = column of the 21-wide right strip: 11 independent row loads, no index arithmetic
        const uint8_t *pr = imgR + (ptrdiff_t)y0 * gr.pitch + xr0 + lane;
        uint8_t v[11];
#pragma unroll
        for (int y = 0; y < 11; y++) v[y] = pr[(ptrdiff_t)y * gr.pitch];
#pragma unroll
        for (int y = 0; y < 11; y++) IR[y * 24 + lane] = v[y];
    }
    const int cL = __shfl_sync(0xffffffffu, aL[1], 60 - 32);   // window pixel 60 = (5, 5)
    __syncwarp();
    uint32_t acc[11];
#pragma unroll
    for (int k = 0; k < 11; k++) {   // incR = k - 5
        const int cRL = (int)IR[5 * 24 + k + 5] - cL;           // |(IL - cL) - (IR - cR)| = |(IL + cR - cL) - IR|
        uint32_t sum = 0;
#pragma unroll
        for (int m = 0; m < 4; m++)
            if (m < 3 || lane < 121 - 96) sum = __sad(aL[m] + cRL, (int)IR[irb[m] + k], sum);
        acc[k] = sum;
    }
    // a window sums to at most 121 * 510 < 2^16: reduce two offsets per register
    uint32_t pk[6];
#pragma unroll
    for (int k = 0; k < 5; k++) pk[k] = acc[2 * k] | (acc[2 * k + 1] << 16);
    pk[5] = acc[10];
#pragma unroll
    for (int sft = 16; sft > 0; sft >>= 1) {
#pragma unroll
        for (int k = 0; k < 6; k++) pk[k] += __shfl_xor_sync(0xffffffffu, pk[k], sft);
    }
    // every lane holds all 11 sums; lane k < 11 takes the k-th and the warp picks the first minimum (ref :588-592: strict '<'
    // over ascending incR) as the minimum of (sum << 4 | k); the parabola's three samples come back by shuffle
    uint32_t mine = pk[0];
#pragma unroll
    for (int k = 1; k < 6; k++) mine = (lane >> 1) == k ? pk[k] : mine;
    mine = (mine >> (16 * (lane & 1))) & 0xffffu;
    uint32_t key = lane < 11 ? (mine << 4) | (uint32_t)lane : 0xffffffffu;
#pragma unroll
    for (int sft = 8; sft > 0; sft >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, sft));
    key = __shfl_sync(0xffffffffu, key, 0);   // lanes 16..31 reduced among themselves
    const int bestS = (int)(key >> 4), bestk = (int)(key & 15u), bestinc = bestk - 5;
    if (bestinc == -L || bestinc == L) return;                 // ref :597-598
    const float fm = (float)mine;
    const float d1 = __shfl_sync(0xffffffffu, fm, bestk - 1), d2 = __shfl_sync(0xffffffffu, fm, bestk),
                d3 = __shfl_sync(0xffffffffu, fm, bestk + 1);
    // ref :605: (dist1-dist3)/(2.0f*(dist1+dist3-2.0f*dist2)); NaN (0/0) passes the +-1 test below
    const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
    if (deltaR < -1 || deltaR > 1) return;
    float bestuR = __fmul_rn(gl.scale, __fadd_rn(__fadd_rn(scaleduR0, (float)bestinc), deltaR));   // ref :611
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= 0 && disparity < maxD) {
        if (disparity <= 0) {
            disparity = 0.01f;
            bestuR = __double2float_rn(__dsub_rn((double)uL, 0.01));   // ref :620: float - double literal
        }
        if (lane == 0) {
            A.depth[o] = __fdiv_rn(A.mbf, disparity);
            A.uRight[o] = bestuR;
            A.sad[o] = bestS;
        }
    }
}

// Median-SAD rejection (ref :629-642): thDist = 1.5f*1.4f*median, where median is element size/2 of
// the ascending (SAD, iL) list.
// The reference sorts the (SAD, index) pairs only to read the median SAD and to drop everything at or above 2.1 x median (ref
// :629-642): an order statistic, found here by a two-level radix select over the 16-bit SAD range (121 pixels x 510 < 2^16) -- a
// histogram of the high bytes, a scan to the bucket that holds rank count / 2, a histogram of that bucket's low bytes -- instead of
// a bitonic sort of `capacity` keys (48 -> ~6 us per batch of 144 pairs; the kernel runs one CTA per pair, i.e. on a mostly
// empty GPU).
__global__ void __launch_bounds__(1024) k_stereo_filter(StereoArgs A) {
    __shared__ int s_hist[256], s_count, s_bucket, s_rank, s_median;
    const int b = blockIdx.x, tid = threadIdx.x;
    const int N = min(A.nL[b], A.capacity);
    const int32_t *sad = A.sad + (size_t)b * A.capacity;
    auto select = [&](int rank) {   // thread-uniform: bucket of s_hist that holds the element of 0-based rank `rank`; sets s_bucket, s_rank
        if (tid < 32) {
            int c[8], tot = 0;
#pragma unroll
            for (int i = 0; i < 8; i++) { c[i] = s_hist[tid * 8 + i]; tot += c[i]; }
            int incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (tid >= o) incl += t;
            }
            int before = incl - tot;
            if (rank >= before && rank < incl) {
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    if (rank >= before && rank < before + c[i]) { s_bucket = tid * 8 + i; s_rank = rank - before; }
                    before += c[i];
                }
            }
        }
    };
    if (tid < 256) s_hist[tid] = 0;
    if (tid == 0) s_count = 0;
    __syncthreads();
    int mine = 0;
    for (int i = tid; i < N; i += 1024) {
        const int v = sad[i];
        if (v >= 0) { atomicAdd(&s_hist[(v >> 8) & 255], 1); mine++; }
    }
    if (mine) atomicAdd(&s_count, mine);
    __syncthreads();
    const int count = s_count;
    if (count == 0) {
        if (tid == 0) A.n_matched[b] = 0;
        return;
    }
    select(count / 2);                       // median = vDistIdx[vDistIdx.size() / 2].first
    __syncthreads();
    const int hi = s_bucket, rank2 = s_rank;
    __syncthreads();
    if (tid < 256) s_hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < N; i += 1024) {
        const int v = sad[i];
        if (v >= 0 && ((v >> 8) & 255) == hi) atomicAdd(&s_hist[v & 255], 1);
    }
    __syncthreads();
    select(rank2);
    __syncthreads();
    if (tid == 0) { s_median = (hi << 8) | s_bucket; s_count = 0; }
    __syncthreads();
    const float median = (float)s_median;
    const float thDist = __fmul_rn(1.5f * 1.4f, median);
    int kept = 0;
    for (int i = tid; i < N; i += 1024) {
        if (sad[i] >= 0) {
            if ((float)sad[i] < thDist) kept++;
            else { A.uRight[(size_t)b * A.capacity + i] = -1.f; A.depth[(size_t)b * A.capacity + i] = -1.f; }
        }
    }
    if (kept) atomicAdd(&s_count, kept);
    __syncthreads();
    if (tid == 0) A.n_matched[b] = s_count;
}

}  // namespace orbgpu

using namespace orbgpu;

extern "C" {

int orb_stereo_match_batch(orb_extractor *left, int left_slot, orb_extractor *right, int right_slot, int batch,
                           const orb_keypoint *keys_left, const uint8_t *desc_left, const int32_t *n_left,
                           const orb_keypoint *keys_right, const uint8_t *desc_right, const int32_t *n_right, int capacity,
                           float mbf, float mb, float *u_right, float *depth, int32_t *n_matched, int loc) {
    ORB_REQUIRE(left && right && keys_left && desc_left && n_left && keys_right && desc_right && n_right && u_right && depth,
                ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(batch >= 1 && capacity >= 1, ORB_ERR_INVALID, "bad batch/capacity");
    ORB_REQUIRE(left->device == right->device, ORB_ERR_INVALID, "left and right extractors must live on one device");
    ORB_REQUIRE(left->geom.width > 0 && right->geom.width > 0, ORB_ERR_INVALID, "extract the pair before matching");
    ORB_REQUIRE(left->geom.nlevels == right->geom.nlevels && left->geom.width == right->geom.width &&
                    left->geom.height == right->geom.height, ORB_ERR_INVALID, "left/right pyramids differ in geometry");
    ORB_REQUIRE(left_slot >= 0 && left_slot + batch <= left->prm.max_batch && right_slot >= 0 &&
                    right_slot + batch <= right->prm.max_batch, ORB_ERR_INVALID, "slot range exceeds max_batch");
    ORB_CUDA_TRY(cudaSetDevice(left->device));
    cudaStream_t st = left->stream;
    if (right != left) {   // the right pyramid is produced on the other handle's stream
        if (!left->ev_pair) ORB_CUDA_TRY(cudaEventCreateWithFlags(&left->ev_pair, cudaEventDisableTiming));
        ORB_CUDA_TRY(cudaEventRecord(left->ev_pair, right->stream));
        ORB_CUDA_TRY(cudaStreamWaitEvent(st, left->ev_pair, 0));
    }
    const size_t per = (size_t)batch * capacity;
    // scratch layout: sad[per] | n_matched[batch] | (host mode) keysL,keysR,descL,descR,nL,nR,uRight,depth
    const int nrows = left->geom.lv[0].h;
    // a right keypoint is listed in rows floor(y-r)..ceil(y+r), r = 2*scale: at most 2*ceil(2*scale_max)+3 rows
    const int list_cap = capacity * (2 * (int)std::ceil(2.0f * left->scale.back()) + 3);
    size_t need = per * 4 + per * 16 + batch * 4 + (size_t)batch * (nrows + 1) * 4 + (size_t)batch * list_cap * 4 + 16 * 256;   // every carve() is rounded up to 256 B
    if (loc == ORB_HOST) need += 2 * per * sizeof(orb_keypoint) + 2 * per * 32 + 2 * batch * 4 + 2 * per * 4 + 1024;
    if (left->stereo_scratch_bytes < need) {
        if (left->stereo_scratch) cudaFree(left->stereo_scratch);
        left->stereo_scratch = nullptr;
        left->stereo_scratch_bytes = 0;
        ORB_CUDA_TRY(cudaMalloc(&left->stereo_scratch, need));
        left->stereo_scratch_bytes = need;
    }
    uint8_t *p = (uint8_t *)left->stereo_scratch;
    auto carve = [&](size_t bytes) { uint8_t *r = p; p += (bytes + 255) / 256 * 256; return r; };
    StereoArgs A;
    A.sad = (int32_t *)carve(per * 4);
    A.packR = (int4 *)carve(per * 16);
    A.rowoff = (int *)carve((size_t)batch * (nrows + 1) * 4);
    A.rowlist = (int *)carve((size_t)batch * list_cap * 4);
    A.nrows = nrows; A.list_cap = list_cap;
    A.n_matched = (int32_t *)carve(batch * 4);
    A.pyrL = left->d_pyr; A.pyrR = right->d_pyr;
    A.GL = left->d_geom; A.GR = right->d_geom;
    A.slotL0 = left_slot; A.slotR0 = right_slot;
    A.capacity = capacity; A.mbf = mbf; A.mb = mb;
    if (loc == ORB_HOST) {
        orb_keypoint *dkl = (orb_keypoint *)carve(per * sizeof(orb_keypoint));
        orb_keypoint *dkr = (orb_keypoint *)carve(per * sizeof(orb_keypoint));
        uint8_t *ddl = carve(per * 32), *ddr = carve(per * 32);
        int32_t *dnl = (int32_t *)carve(batch * 4), *dnr = (int32_t *)carve(batch * 4);
        float *dur = (float *)carve(per * 4), *ddp = (float *)carve(per * 4);
        ORB_CUDA_TRY(cudaMemcpyAsync(dkl, keys_left, per * sizeof(orb_keypoint), cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(dkr, keys_right, per * sizeof(orb_keypoint), cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(ddl, desc_left, per * 32, cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(ddr, desc_right, per * 32, cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(dnl, n_left, batch * 4, cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(dnr, n_right, batch * 4, cudaMemcpyHostToDevice, st));
        A.keysL = dkl; A.keysR = dkr; A.descL = ddl; A.descR = ddr; A.nL = dnl; A.nR = dnr; A.uRight = dur; A.depth = ddp;
    } else {
        A.keysL = keys_left; A.keysR = keys_right; A.descL = desc_left; A.descR = desc_right;
        A.nL = n_left; A.nR = n_right; A.uRight = u_right; A.depth = depth;
    }
    k_stereo_rowtable<<<batch, ROWTAB_THREADS, (nrows + 1) * sizeof(int), st>>>(A);
    k_stereo_match<<<dim3((capacity + ST_WARPS - 1) / ST_WARPS, batch), ST_WARPS * 32, 0, st>>>(A);
    k_stereo_filter<<<batch, 1024, 0, st>>>(A);
    left->launches += 3;
    ORB_CUDA_TRY(cudaGetLastError());
    if (right != left) {   // the right handle's next extraction must not overwrite its pyramid while these kernels read it
        if (!left->ev_pair_done) ORB_CUDA_TRY(cudaEventCreateWithFlags(&left->ev_pair_done, cudaEventDisableTiming));
        ORB_CUDA_TRY(cudaEventRecord(left->ev_pair_done, st));
        ORB_CUDA_TRY(cudaStreamWaitEvent(right->stream, left->ev_pair_done, 0));
    }
    if (loc == ORB_HOST) {
        ORB_CUDA_TRY(cudaMemcpyAsync(u_right, A.uRight, per * 4, cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(depth, A.depth, per * 4, cudaMemcpyDeviceToHost, st));
        if (n_matched) ORB_CUDA_TRY(cudaMemcpyAsync(n_matched, A.n_matched, batch * 4, cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaStreamSynchronize(st));
    } else if (n_matched) {
        ORB_CUDA_TRY(cudaMemcpyAsync(n_matched, A.n_matched, batch * 4, cudaMemcpyDeviceToDevice, st));
    }
    return ORB_OK;
}

int orb_stereo_match(orb_extractor *left, orb_extractor *right, int slot, const orb_keypoint *keys_left,
                     const uint8_t *desc_left, int n_left, const orb_keypoint *keys_right, const uint8_t *desc_right,
                     int n_right, float mbf, float mb, float *u_right, float *depth, int *n_matched) {
    ORB_REQUIRE(n_left >= 0 && n_right >= 0, ORB_ERR_INVALID, "negative count");
    if (n_matched) *n_matched = 0;
    if (n_left == 0) { set_error("no left keypoints"); return ORB_ERR_EMPTY; }
    // single pair: the arrays are exactly n long, so use separate strides via two capacities = max(n)
    const int cap = n_left > n_right ? n_left : n_right;
    ORB_REQUIRE(cap <= 65535, ORB_ERR_INVALID, "too many keypoints");
    std::vector<orb_keypoint> kl(cap), kr(cap);
    std::vector<uint8_t> dl((size_t)cap * 32), dr((size_t)cap * 32);
    std::vector<float> ur(cap), dp(cap);
    memcpy(kl.data(), keys_left, (size_t)n_left * sizeof(orb_keypoint));
    memcpy(dl.data(), desc_left, (size_t)n_left * 32);
    if (n_right) {
        memcpy(kr.data(), keys_right, (size_t)n_right * sizeof(orb_keypoint));
        memcpy(dr.data(), desc_right, (size_t)n_right * 32);
    }
    int32_t nl = n_left, nr = n_right, nm = 0;
    int rc = orb_stereo_match_batch(left, slot, right, slot, 1, kl.data(), dl.data(), &nl, kr.data(), dr.data(), &nr, cap, mbf, mb,
                                    ur.data(), dp.data(), &nm, ORB_HOST);
    if (rc != ORB_OK) return rc;
    memcpy(u_right, ur.data(), (size_t)n_left * 4);
    memcpy(depth, dp.data(), (size_t)n_left * 4);
    if (n_matched) *n_matched = nm;
    if (nm == 0) { set_error("no stereo match accepted (the reference indexes an empty vector here, Frame.cc:630)"); return ORB_ERR_EMPTY; }
    return ORB_OK;
}

/* One stereo frame through one handle with a single host synchronisation: both images up, both extractions as one batch of
 * two (a CUDA graph from the third frame on), the association on the device-resident keypoints, all results back through
 * the pinned landing buffer.  Same results as orb_extract x2 + orb_stereo_match. */
int orb_stereo_frame(orb_extractor *ex, const uint8_t *image_left, const uint8_t *image_right, int width, int height,
                     int pitch_left, int pitch_right, orb_keypoint *keys_left, uint8_t *desc_left, int *n_left,
                     orb_keypoint *keys_right, uint8_t *desc_right, int *n_right, int capacity, float mbf, float mb,
                     float *u_right, float *depth, int *n_matched) {
    ORB_REQUIRE(ex && image_left && image_right && keys_left && desc_left && n_left && keys_right && desc_right && n_right &&
                    u_right && depth, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(ex->prm.max_batch >= 2, ORB_ERR_INVALID, "orb_stereo_frame needs a handle with max_batch >= 2");
    ORB_REQUIRE(width > 0 && height > 0 && pitch_left >= width && pitch_right >= width, ORB_ERR_INVALID, "bad image geometry");
    ORB_REQUIRE(width <= ex->prm.max_width && height <= ex->prm.max_height, ORB_ERR_INVALID,
                "image %dx%d exceeds the handle's max %dx%d", width, height, ex->prm.max_width, ex->prm.max_height);
    ORB_REQUIRE(capacity >= 1 && capacity <= 65535, ORB_ERR_INVALID, "bad capacity");
    ORB_CUDA_TRY(cudaSetDevice(ex->device));
    cudaStream_t st = ex->stream;
    int rc = orbgpu_ensure_out_staging(ex, 2, capacity);
    if (rc != ORB_OK) return rc;
    if (ex->stereo_out_cap < (size_t)capacity) {
        cudaFree(ex->d_stereo_out);
        ex->d_stereo_out = nullptr; ex->stereo_out_cap = 0;
        ORB_CUDA_TRY(cudaMalloc(&ex->d_stereo_out, ((size_t)2 * capacity + 1) * 4));
        ex->stereo_out_cap = capacity;
    }
    // both frames into the handle's source slots 0 and 1 (linear copies when the two pitches agree and fit)
    int d_pitch = pitch_left;
    size_t d_stride = (size_t)pitch_left * height;
    const uint8_t *src = ex->d_src;
    const size_t used = (size_t)pitch_left * (height - 1) + width;   // the last row of a view holds `width` bytes, not `pitch`
    if (pitch_left == pitch_right && ex->h_src && d_stride + used <= ORB_HSRC_BYTES && 2 * (size_t)width * height <= ((size_t)4 << 20)) {
        // both frames into the pinned, device-mapped input buffer: k_pyr_level0 reads them across PCIe (see orb_extract_batch)
        memcpy(ex->h_src, image_left, used);
        memcpy(ex->h_src + d_stride, image_right, used);
        src = ex->h_src;
    } else if (pitch_left == pitch_right && 2 * d_stride <= ex->src_bytes) {
        ORB_CUDA_TRY(cudaMemcpyAsync(ex->d_src, image_left, used, cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(ex->d_src + d_stride, image_right, used, cudaMemcpyHostToDevice, st));
    } else {
        d_pitch = (width + 127) / 128 * 128;
        d_stride = (size_t)d_pitch * height;
        ORB_CUDA_TRY(cudaMemcpy2DAsync(ex->d_src, d_pitch, image_left, pitch_left, width, height, cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpy2DAsync(ex->d_src + d_stride, d_pitch, image_right, pitch_right, width, height, cudaMemcpyHostToDevice, st));
    }
    rc = orb_extract_batch(ex, src, 2, width, height, d_pitch, d_stride, ORB_DEVICE, ex->d_out_kp, ex->d_out_desc, capacity,
                           ex->d_out_n, ORB_DEVICE);
    if (rc != ORB_OK) return rc;
    float *d_ur = ex->d_stereo_out, *d_dp = d_ur + capacity;
    ex->stereo_last_capacity = capacity;
    int32_t *d_nm = (int32_t *)(d_dp + capacity);
    rc = orb_stereo_match_batch(ex, 0, ex, 1, 1, ex->d_out_kp, ex->d_out_desc, ex->d_out_n, ex->d_out_kp + capacity,
                                ex->d_out_desc + (size_t)capacity * 32, ex->d_out_n + 1, capacity, mbf, mb, d_ur, d_dp, d_nm, ORB_DEVICE);
    if (rc != ORB_OK) return rc;
    // everything back into pinned memory, one synchronisation
    const size_t kp_bytes = (size_t)2 * capacity * sizeof(orb_keypoint), desc_bytes = (size_t)2 * capacity * 32, f_bytes = (size_t)capacity * 4;
    const bool bounce = ex->h_bounce && kp_bytes + desc_bytes + 2 * f_bytes + 4 <= ORB_BOUNCE_BYTES;
    std::vector<uint8_t> big;
    uint8_t *land = ex->h_bounce;
    if (!bounce) { big.resize(kp_bytes + desc_bytes + 2 * f_bytes + 4); land = big.data(); }
    if (bounce) {   // one kernel writes flags, counts and the valid entries of all five arrays into the pinned buffers
        orbgpu_export_results(ex, ex->d_out_kp, ex->d_out_desc, ex->d_out_n, 2, capacity, d_ur, d_dp, d_nm, st);
        ORB_CUDA_TRY(cudaGetLastError());
    } else {
        ORB_CUDA_TRY(cudaMemcpyAsync(ex->h_pinned, ex->d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(ex->h_pinned + 1, ex->d_out_n, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(land, ex->d_out_kp, kp_bytes, cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(land + kp_bytes, ex->d_out_desc, desc_bytes, cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(land + kp_bytes + desc_bytes, d_ur, 2 * f_bytes + 4, cudaMemcpyDeviceToHost, st));
    }
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    const int flags = ex->h_pinned[0];
    *n_left = ex->h_pinned[1]; *n_right = ex->h_pinned[2];
    ex->last_counts.assign(ex->h_pinned + 1, ex->h_pinned + 3);   // orb_frame_from_extractor: counts known on the host
    ex->last_counts_valid = true;
    if (flags) {   // capacity / candidate overflow raised by the extraction kernels
        cudaMemsetAsync(ex->d_err, 0, sizeof(int), st);
        set_error("extraction raised error flags 0x%x (capacity %d, needed %d / %d)", flags, capacity, *n_left, *n_right);
        return ORB_ERR_CAPACITY;
    }
    const size_t nl = (size_t)std::min(*n_left, capacity), nr = (size_t)std::min(*n_right, capacity);
    memcpy(keys_left, land, nl * sizeof(orb_keypoint));
    memcpy(keys_right, land + (size_t)capacity * sizeof(orb_keypoint), nr * sizeof(orb_keypoint));
    memcpy(desc_left, land + kp_bytes, nl * 32);
    memcpy(desc_right, land + kp_bytes + (size_t)capacity * 32, nr * 32);
    memcpy(u_right, land + kp_bytes + desc_bytes, nl * 4);
    memcpy(depth, land + kp_bytes + desc_bytes + f_bytes, nl * 4);
    int32_t nm = 0;
    memcpy(&nm, land + kp_bytes + desc_bytes + 2 * f_bytes, 4);
    if (n_matched) *n_matched = nm;
    if (nm == 0) { set_error("no stereo match accepted (the reference indexes an empty vector here, Frame.cc:630)"); return ORB_ERR_EMPTY; }
    return ORB_OK;
}

int orb_extractor_stereo_result(orb_extractor *ex, const float **dev_u_right, const float **dev_depth) {
    ORB_REQUIRE(ex && ex->d_stereo_out, ORB_ERR_INVALID, "no orb_stereo_frame result in this handle");
    if (dev_u_right) *dev_u_right = ex->d_stereo_out;
    if (dev_depth) *dev_depth = ex->d_stereo_out + ex->stereo_last_capacity;
    return ORB_OK;
}

}  // extern "C"
