// liborbgpu: MapPoint::ComputeDistinctiveDescriptors on the device.
//
// ref: ORB_SLAM2/src/MapPoint.cc:224-285.  One warp per map point; lane = row of the N x N distance matrix.  The
// median of a row is found without storing or sorting it: distances are integers in [0,256], so a 9-step bisection on
// the value with a counting pass per step (`how many d(i,j) <= mid`) yields exactly sorted_row[(N-1)/2].  All lanes
// walk j together, so every descriptor load is one broadcast 2 x LDG.128; the inner loop is 8 XOR + 8 POPC.  The row
// with the smallest (median, row index) wins = the reference's strict `median < BestMedian`.
#include "orb_common.cuh"

using namespace orbgpu;

namespace {

constexpr int DD_WARPS = 4;

__global__ void __launch_bounds__(DD_WARPS * 32)
k_distinctive(const uint4 *__restrict__ desc, const int *__restrict__ off, int npoints, int *__restrict__ best_row,
              int *__restrict__ best_median) {
    const int p = blockIdx.x * DD_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (p >= npoints) return;
    const int beg = off[p], n = off[p + 1] - beg;
    if (n <= 0) { if (lane == 0) { best_row[p] = -1; best_median[p] = 0; } return; }
    const uint4 *rows = desc + 2 * (long long)beg;
    const int rank = (n - 1) >> 1;            // vDists[0.5*(N-1)] (MapPoint.cc:270): the double product truncates
    unsigned best = 0xffffffffu;              // (median << 20 | row)
    for (int i0 = 0; i0 < n; i0 += 32) {      // uniform trip count: the inner loads are warp-wide broadcasts
        const int i = i0 + lane;
        const bool live = i < n;
        uint32_t q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (live) {
            const uint4 a = __ldg(rows + 2 * i), c = __ldg(rows + 2 * i + 1);
            q[0] = a.x; q[1] = a.y; q[2] = a.z; q[3] = a.w; q[4] = c.x; q[5] = c.y; q[6] = c.z; q[7] = c.w;
        }
        int lo = 0, hi = 256;
#pragma unroll 1
        for (int step = 0; step < 9; step++) {   // ceil(log2(257)) = 9
            const int mid = (lo + hi) >> 1;
            int cnt = 0;
            for (int j = 0; j < n; j++) {
                const uint4 a = __ldg(rows + 2 * j), c = __ldg(rows + 2 * j + 1);
                const int d = __popc(q[0] ^ a.x) + __popc(q[1] ^ a.y) + __popc(q[2] ^ a.z) + __popc(q[3] ^ a.w) +
                              __popc(q[4] ^ c.x) + __popc(q[5] ^ c.y) + __popc(q[6] ^ c.z) + __popc(q[7] ^ c.w);
                cnt += d <= mid;
            }
            if (lo < hi) { if (cnt > rank) hi = mid; else lo = mid + 1; }
        }
        if (live) best = min(best, ((unsigned)lo << 20) | (unsigned)i);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) { best_row[p] = (int)(best & 0xfffffu); best_median[p] = (int)(best >> 20); }
}

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
};

}  // namespace

extern "C" int orb_distinctive_descriptors(const uint8_t *descriptors, const int32_t *offsets, int npoints, int loc,
                                           int32_t *best_row, int32_t *best_median) {
    ORB_REQUIRE(npoints >= 0 && (npoints == 0 || (descriptors && offsets && best_row && best_median)), ORB_ERR_INVALID, "bad argument");
    if (npoints == 0) return ORB_OK;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    const int blocks = (npoints + DD_WARPS - 1) / DD_WARPS;
    ThreadCtx *T = nullptr;
    { int rc = orbgpu_thread_ctx(&T, 0, 0); if (rc != ORB_OK) return rc; }
    cudaStream_t st = T->stream;   // the thread's call stream (orb_set_call_stream), never the legacy default stream
    if (loc == ORB_DEVICE) {
        k_distinctive<<<blocks, DD_WARPS * 32, 0, st>>>(reinterpret_cast<const uint4 *>(descriptors), offsets, npoints, best_row, best_median);
        ORB_CUDA_TRY(cudaGetLastError());
        return ORB_OK;
    }
    const int total = offsets[npoints];
    ORB_REQUIRE(offsets[0] == 0 && total >= 0, ORB_ERR_INVALID, "offsets must start at 0 and be non-decreasing");
    for (int p = 0; p < npoints; p++) {
        ORB_REQUIRE(offsets[p + 1] >= offsets[p], ORB_ERR_INVALID, "offsets must be non-decreasing");
        ORB_REQUIRE(offsets[p + 1] - offsets[p] < (1 << 20), ORB_ERR_INVALID, "map point %d has too many observations", p);
    }
    DevBuf d_desc, d_off, d_row, d_med;
    ORB_CUDA_TRY(d_desc.alloc((size_t)total * 32));
    ORB_CUDA_TRY(d_off.alloc((size_t)(npoints + 1) * 4));
    ORB_CUDA_TRY(d_row.alloc((size_t)npoints * 4));
    ORB_CUDA_TRY(d_med.alloc((size_t)npoints * 4));
    ORB_CUDA_TRY(cudaMemcpyAsync(d_desc.p, descriptors, (size_t)total * 32, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(d_off.p, offsets, (size_t)(npoints + 1) * 4, cudaMemcpyHostToDevice, st));
    k_distinctive<<<blocks, DD_WARPS * 32, 0, st>>>((const uint4 *)d_desc.p, (const int *)d_off.p, npoints, (int *)d_row.p, (int *)d_med.p);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(best_row, d_row.p, (size_t)npoints * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(best_median, d_med.p, (size_t)npoints * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}
