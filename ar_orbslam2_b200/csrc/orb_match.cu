// liborbgpu -- Hamming matching primitives for sm_100a.
//
//   orb_descriptor_distance   ORBmatcher::DescriptorDistance (ref: ORB_SLAM2/src/ORBmatcher.cc:1650-1666)
//   orb_hamming_top2          brute-force best / second-best scan (the inner loop of ref
//                             ORBmatcher.cc:205-226, 573-599 and Marker.cc:314-349) -- the C5 shard kernel
//   orb_top2_merge            merge of per-shard top-2 lists == one scan of the concatenated database
//   orb_db_*                  keyframe-range shard of a descriptor database resident in HBM
//   orb_measure_popc_rate     measured POPC issue rate (roofline denominator for matching)
//
// One compare = 8 x (XOR + POPC) on 32-bit words.  Queries live in registers (one query per thread),
// database rows stream through shared memory with cp.async and are read as broadcast 16-byte loads.
// Tensor cores are deliberately unused: this is bit counting, not a float contraction.
#include "orb_common.cuh"

#include <algorithm>
#include <cstdint>
#include <mutex>
#include <vector>

namespace orbgpu {

constexpr int T2_THREADS = 256;     // queries per CTA
constexpr int T2_TILE = 256;        // database rows per shared-memory stage (8 KB)
constexpr int T2_STAGES = 3;

struct Top2 {
    int32_t d1, d2;
    int64_t i1, i2;
};

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// grid = (query blocks, database chunks).  Each CTA scans rows [chunk*rows_per_chunk, ...) in order for
// its 256 queries and writes one partial top-2 per (chunk, query).
//
// KF = true: the rows are keyframes of `kf.rows` consecutive rows each (chunks start and end on keyframe boundaries).  The scan
// keeps the top-2 of the CURRENT keyframe; at a keyframe's last row it counts the query for that keyframe if its best is
// <= th_low and (float)best < ratio * (float)second (ref ORBmatcher.cc:228-231, second = 256 when the keyframe has one
// row), one atomic per warp, and folds the keyframe's top-2 into the chunk's (same lexicographic order as one scan).
struct KfParams { int rows, th_low; float ratio; int32_t *counts; };

template <bool KF>
__global__ void __launch_bounds__(T2_THREADS)
k_hamming_top2(const uint8_t *__restrict__ queries, int nq, const uint8_t *__restrict__ db, long long ndb,
               long long rows_per_chunk, int32_t *__restrict__ pd1, int32_t *__restrict__ pd2,
               int64_t *__restrict__ pi1, int64_t *__restrict__ pi2, long long idx_base, const KfParams kf) {
    __shared__ __align__(16) uint4 tile[T2_STAGES][T2_TILE * 2];
    const int tid = threadIdx.x;
    const int q = blockIdx.x * T2_THREADS + tid;
    const long long r0 = (long long)blockIdx.y * rows_per_chunk;
    const long long r1 = min(r0 + rows_per_chunk, ndb);
    uint32_t qa[8];
    {
        const uint4 *qp = reinterpret_cast<const uint4 *>(queries + (size_t)min(q, nq - 1) * 32);
        const uint4 a = __ldg(qp), b = __ldg(qp + 1);
        qa[0] = a.x; qa[1] = a.y; qa[2] = a.z; qa[3] = a.w; qa[4] = b.x; qa[5] = b.y; qa[6] = b.z; qa[7] = b.w;
    }
    int d1 = 256, d2 = 256;
    long long i1 = -1, i2 = -1;
    int k1 = 256, k2 = 256;                  // KF: top-2 inside the current keyframe
    long long j1 = -1, j2 = -1, kf_end = KF ? r0 + kf.rows : 0;
    const long long nrows = r1 - r0;
    const int ntiles = nrows > 0 ? (int)((nrows + T2_TILE - 1) / T2_TILE) : 0;
    const uint4 *src = reinterpret_cast<const uint4 *>(db) + r0 * 2;
    const long long n16 = nrows * 2;   // 16-byte units in this chunk

    auto issue = [&](int t) {
        if (t < ntiles) {
            uint4 *dst = tile[t % T2_STAGES];
            const long long base = (long long)t * T2_TILE * 2;
#pragma unroll
            for (int k = 0; k < 2; k++) {
                const long long u = base + tid + k * T2_THREADS;
                if (u < n16) cp_async16(dst + tid + k * T2_THREADS, src + u);
            }
        }
        cp_async_commit();
    };
    for (int t = 0; t < T2_STAGES - 1; t++) issue(t);
    for (int t = 0; t < ntiles; t++) {
        issue(t + T2_STAGES - 1);
        cp_async_wait<T2_STAGES - 1>();
        __syncthreads();
        const uint4 *rows = tile[t % T2_STAGES];
        const int nr = (int)min((long long)T2_TILE, nrows - (long long)t * T2_TILE);
        const long long rbase = r0 + (long long)t * T2_TILE;
        if (!KF) {
#pragma unroll 4
            for (int r = 0; r < nr; r++) {
                const uint4 a = rows[2 * r], b = rows[2 * r + 1];   // broadcast: every lane reads the same row
                const int d = __popc(qa[0] ^ a.x) + __popc(qa[1] ^ a.y) + __popc(qa[2] ^ a.z) + __popc(qa[3] ^ a.w) +
                              __popc(qa[4] ^ b.x) + __popc(qa[5] ^ b.y) + __popc(qa[6] ^ b.z) + __popc(qa[7] ^ b.w);
                if (d < d2) {   // strict '<': the first row seen wins ties, as the reference's sequential scan
                    if (d < d1) { d2 = d1; i2 = i1; d1 = d; i1 = rbase + r; }
                    else { d2 = d; i2 = rbase + r; }
                }
            }
        } else {
            int r = 0;
            while (r < nr) {
                const int seg = (int)min((long long)nr, kf_end - rbase);   // rows of this tile that belong to the current keyframe
#pragma unroll 4
                for (; r < seg; r++) {
                    const uint4 a = rows[2 * r], b = rows[2 * r + 1];
                    const int d = __popc(qa[0] ^ a.x) + __popc(qa[1] ^ a.y) + __popc(qa[2] ^ a.z) + __popc(qa[3] ^ a.w) +
                                  __popc(qa[4] ^ b.x) + __popc(qa[5] ^ b.y) + __popc(qa[6] ^ b.z) + __popc(qa[7] ^ b.w);
                    if (d < k2) {
                        if (d < k1) { k2 = k1; j2 = j1; k1 = d; j1 = rbase + r; }
                        else { k2 = d; j2 = rbase + r; }
                    }
                }
                if (rbase + r == kf_end) {   // the keyframe is complete (uniform over the CTA)
                    const bool ok = q < nq && j1 >= 0 && k1 <= kf.th_low && (float)k1 < kf.ratio * (float)k2;
                    const unsigned bal = __ballot_sync(0xffffffffu, ok);
                    if (bal && (tid & 31) == 0) atomicAdd(kf.counts + (kf_end / kf.rows - 1), __popc(bal));
                    if (k1 < d2) {
                        if (k1 < d1) { d2 = d1; i2 = i1; d1 = k1; i1 = j1; }
                        else { d2 = k1; i2 = j1; }
                    }
                    if (j2 >= 0 && k2 < d2) { d2 = k2; i2 = j2; }   // k2 >= k1 >= d1 here: it can only be the second
                    k1 = k2 = 256; j1 = j2 = -1;
                    kf_end += kf.rows;
                }
            }
        }
        __syncthreads();
    }
    if (q < nq) {
        const size_t o = (size_t)blockIdx.y * nq + q;
        pd1[o] = d1; pd2[o] = d2;
        pi1[o] = i1 < 0 ? -1 : i1 + idx_base;
        pi2[o] = i2 < 0 ? -1 : i2 + idx_base;
    }
}

// Lexicographic merge of `nparts` partial top-2 lists per query, parts in database order.  Ties on the
// distance keep the entry from the earlier part (= lower row index), so the result equals one scan.
__global__ void k_top2_merge(int nparts, int nq, const int32_t *__restrict__ pd1, const int64_t *__restrict__ pi1,
                             const int32_t *__restrict__ pd2, const int64_t *__restrict__ pi2, int32_t *__restrict__ od1,
                             int64_t *__restrict__ oi1, int32_t *__restrict__ od2, int64_t *__restrict__ oi2) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    int d1 = 256, d2 = 256;
    long long i1 = -1, i2 = -1;
    for (int p = 0; p < nparts; p++) {
        const size_t o = (size_t)p * nq + q;
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const int d = k ? pd2[o] : pd1[o];
            const long long i = k ? pi2[o] : pi1[o];
            if (i < 0) continue;
            if (d < d2) {
                if (d < d1) { d2 = d1; i2 = i1; d1 = d; i1 = i; }
                else { d2 = d; i2 = i; }
            }
        }
    }
    od1[q] = d1; oi1[q] = i1; od2[q] = d2; oi2[q] = i2;
}

__global__ void k_descriptor_distance(const uint8_t *__restrict__ a, const uint8_t *__restrict__ b, int n,
                                      int32_t *__restrict__ dist) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4 *pa = reinterpret_cast<const uint4 *>(a + (size_t)i * 32), *pb = reinterpret_cast<const uint4 *>(b + (size_t)i * 32);
    const uint4 a0 = __ldg(pa), a1 = __ldg(pa + 1), b0 = __ldg(pb), b1 = __ldg(pb + 1);
    dist[i] = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) + __popc(a1.x ^ b1.x) +
              __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// splitmix64 of the global 64-bit word index: the device twin of ar_orbslam2_b200/synth.py:synth_descriptors
__global__ void k_db_fill(unsigned long long *__restrict__ words, long long nwords, long long first_word,
                          unsigned long long seed) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += stride) {
        unsigned long long z = seed + (unsigned long long)(first_word + i + 1) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        words[i] = z ^ (z >> 31);
    }
}

// POPC micro-benchmark: 8 independent popc chains per thread.
__global__ void k_popc_rate(unsigned *out, int iters) {
    unsigned x0 = threadIdx.x, x1 = x0 * 3 + 1, x2 = x0 * 5 + 2, x3 = x0 * 7 + 3, x4 = x0 * 11 + 4, x5 = x0 * 13 + 5,
             x6 = x0 * 17 + 6, x7 = x0 * 19 + 7;
    for (int i = 0; i < iters; i++) {
        x0 = __popc(x0) + 0x10001u; x1 = __popc(x1) + 0x20003u; x2 = __popc(x2) + 0x40005u; x3 = __popc(x3) + 0x80007u;
        x4 = __popc(x4) + 0x10009u; x5 = __popc(x5) + 0x2000bu; x6 = __popc(x6) + 0x4000du; x7 = __popc(x7) + 0x8000fu;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = x0 ^ x1 ^ x2 ^ x3 ^ x4 ^ x5 ^ x6 ^ x7;
}

struct Top2Scratch {
    void *buf = nullptr;
    size_t bytes = 0;
    int device = -1;   // the device `buf` lives on: a thread that moves to another device gets a new block there
};

// scans db[0..ndb) for nq queries on `st`; all pointers are device pointers
static int top2_device(const uint8_t *d_q, int nq, const uint8_t *d_db, long long ndb, long long idx_base, int32_t *d_d1,
                       int64_t *d_i1, int32_t *d_d2, int64_t *d_i2, cudaStream_t st, Top2Scratch *scratch,
                       const KfParams *kf = nullptr) {
    int dev = 0, sms = 148;
    ORB_CUDA_TRY(cudaGetDevice(&dev));
    ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int qblocks = (nq + T2_THREADS - 1) / T2_THREADS;
    // enough chunks to fill the machine several times over, but never less than one tile per chunk
    long long chunks = std::max<long long>(1, (long long)sms * 8 / qblocks);
    chunks = std::min<long long>(chunks, std::max<long long>(1, (ndb + T2_TILE - 1) / T2_TILE));
    chunks = std::min<long long>(chunks, 65535);
    long long rows_per_chunk = (std::max<long long>(ndb, 1) + chunks - 1) / chunks;
    if (kf) rows_per_chunk = (rows_per_chunk + kf->rows - 1) / kf->rows * kf->rows;   // chunks hold whole keyframes
    chunks = std::max<long long>(1, (ndb + rows_per_chunk - 1) / rows_per_chunk);
    const size_t per = (size_t)chunks * nq;
    const size_t need = per * (4 + 4 + 8 + 8) + 1024;
    if (scratch->bytes < need || scratch->device != dev) {
        if (scratch->buf) cudaFree(scratch->buf);   // (unified addressing: frees the block on whichever device owns it)
        scratch->buf = nullptr; scratch->bytes = 0; scratch->device = -1;
        ORB_CUDA_TRY(cudaMalloc(&scratch->buf, need));
        scratch->bytes = need; scratch->device = dev;
    }
    int64_t *pi1 = (int64_t *)scratch->buf, *pi2 = pi1 + per;
    int32_t *pd1 = (int32_t *)(pi2 + per), *pd2 = pd1 + per;
    if (kf) {
        ORB_CUDA_TRY(cudaMemsetAsync(kf->counts, 0, (size_t)(ndb / kf->rows) * sizeof(int32_t), st));
        k_hamming_top2<true><<<dim3(qblocks, (unsigned)chunks), T2_THREADS, 0, st>>>(d_q, nq, d_db, ndb, rows_per_chunk, pd1, pd2, pi1,
                                                                                    pi2, idx_base, *kf);
    } else {
        k_hamming_top2<false><<<dim3(qblocks, (unsigned)chunks), T2_THREADS, 0, st>>>(d_q, nq, d_db, ndb, rows_per_chunk, pd1, pd2, pi1,
                                                                                     pi2, idx_base, KfParams{});
    }
    k_top2_merge<<<(nq + 127) / 128, 128, 0, st>>>((int)chunks, nq, pd1, pi1, pd2, pi2, d_d1, d_i1, d_d2, d_i2);
    ORB_CUDA_TRY(cudaGetLastError());
    return ORB_OK;
}

}  // namespace orbgpu

using namespace orbgpu;

struct orb_db {
    int device = 0;
    long long row_begin = 0, row_end = 0;
    uint8_t *d_rows = nullptr;
    cudaStream_t stream = nullptr;
    Top2Scratch scratch;
    void *d_io = nullptr;      // staging for host-mode queries/results
    size_t io_bytes = 0;
    int64_t launches = 0;
};

namespace {
struct DevBuf {   // RAII device allocation (the popc micro-benchmark only)
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t n) { return cudaMalloc(&p, n ? n : 1); }
};

// The one-shot entry points below run on the calling thread's call stream (orb_common.cuh: ThreadCtx) with that thread's
// reusable device block; host-mode arguments are copied on the same stream.  Nothing touches the legacy default stream.
thread_local Top2Scratch t_scratch;   // chunk partials of top2_device, per thread (kept across calls)

inline size_t a256(size_t v) { return (v + 255) / 256 * 256; }
}  // namespace

// smallest and largest distance over ALL (query, train) pairs: the `minD` / `maxD` bookkeeping of naive_nn_search2
// (ref AR-1.3/src/ORBMatcher.cpp:92-93).  One warp per query, lanes stride over the train rows.
__global__ void __launch_bounds__(256)
k_dist_minmax(const uint8_t *__restrict__ query, int nq, const uint8_t *__restrict__ train, int nt, int *__restrict__ out) {
    const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (q >= nq) return;
    const uint4 *pq = reinterpret_cast<const uint4 *>(query + (size_t)q * 32);
    const uint4 a0 = __ldg(pq), a1 = __ldg(pq + 1);
    int mn = 0x7fffffff, mx = -1;
    for (int j = lane; j < nt; j += 32) {
        const uint4 *pt = reinterpret_cast<const uint4 *>(train + (size_t)j * 32);
        const uint4 b0 = __ldg(pt), b1 = __ldg(pt + 1);
        const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                      __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
        mn = min(mn, d);
        mx = max(mx, d);
    }
#pragma unroll
    for (int s = 16; s; s >>= 1) {
        mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, s));
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, s));
    }
    if (lane == 0 && nt > 0) { atomicMin(out, mn); atomicMax(out + 1, mx); }
}

extern "C" {

int orb_descriptor_distance(const uint8_t *a, const uint8_t *b, int n, int32_t *dist, int loc) {
    ORB_REQUIRE(a && b && dist && n >= 0, ORB_ERR_INVALID, "bad argument");
    if (n == 0) return ORB_OK;
    ThreadCtx *T = nullptr;
    int rc = orbgpu_thread_ctx(&T, loc == ORB_HOST ? 2 * a256((size_t)n * 32) + a256((size_t)n * 4) : 0, 0);
    if (rc != ORB_OK) return rc;
    cudaStream_t st = T->stream;
    if (loc == ORB_DEVICE) {
        k_descriptor_distance<<<(n + 127) / 128, 128, 0, st>>>(a, b, n, dist);
        ORB_CUDA_TRY(cudaGetLastError());
        return ORB_OK;
    }
    uint8_t *da = T->d_buf, *db = da + a256((size_t)n * 32);
    int32_t *dd = (int32_t *)(db + a256((size_t)n * 32));
    ORB_CUDA_TRY(cudaMemcpyAsync(da, a, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(db, b, (size_t)n * 32, cudaMemcpyHostToDevice, st));
    k_descriptor_distance<<<(n + 127) / 128, 128, 0, st>>>(da, db, n, dd);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(dist, dd, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

int orb_hamming_top2(const uint8_t *queries, int nq, const uint8_t *db, int64_t ndb, int64_t idx_base, int32_t *best_dist,
                     int64_t *best_idx, int32_t *second_dist, int64_t *second_idx, int loc) {
    ORB_REQUIRE(queries && best_dist && best_idx && second_dist && second_idx, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(nq >= 1 && ndb >= 0 && (db || ndb == 0), ORB_ERR_INVALID, "bad sizes");
    ThreadCtx *T = nullptr;
    const size_t o_db = a256((size_t)nq * 32), o_out = o_db + a256((size_t)ndb * 32);
    int rc = orbgpu_thread_ctx(&T, loc == ORB_HOST ? o_out + a256((size_t)nq * 24) : 0, 0);
    if (rc != ORB_OK) return rc;
    cudaStream_t st = T->stream;
    if (loc == ORB_DEVICE) {
        rc = top2_device(queries, nq, db, ndb, idx_base, best_dist, best_idx, second_dist, second_idx, st, &t_scratch);
        if (rc == ORB_OK) ORB_CUDA_TRY(cudaStreamSynchronize(st));
        return rc;
    }
    uint8_t *dq = T->d_buf, *ddb = T->d_buf + o_db;
    int64_t *o_i1 = (int64_t *)(T->d_buf + o_out), *o_i2 = o_i1 + nq;
    int32_t *o_d1 = (int32_t *)(o_i2 + nq), *o_d2 = o_d1 + nq;
    ORB_CUDA_TRY(cudaMemcpyAsync(dq, queries, (size_t)nq * 32, cudaMemcpyHostToDevice, st));
    if (ndb) ORB_CUDA_TRY(cudaMemcpyAsync(ddb, db, (size_t)ndb * 32, cudaMemcpyHostToDevice, st));
    rc = top2_device(dq, nq, ddb, ndb, idx_base, o_d1, o_i1, o_d2, o_i2, st, &t_scratch);
    if (rc != ORB_OK) return rc;
    ORB_CUDA_TRY(cudaMemcpyAsync(best_idx, o_i1, (size_t)nq * 8, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(second_idx, o_i2, (size_t)nq * 8, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(best_dist, o_d1, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(second_dist, o_d2, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

}  // extern "C"

// ---- orb_hamming_top2_static: the same scan for callers that must not allocate ---------------------------------------
//
// Who this is for: code that runs where device allocation is off the table -- a relocalisation or loop-closing query issued from
// a real-time tracking loop whose memory was sized at start-up, a host that has handed the whole GPU heap to another allocator
// (a torch caching pool, a fixed arena) and does not want a second one growing beside it, or a process close to its memory limit
// where a cudaMalloc in the middle of a frame may fail or stall every stream on the device while the driver maps memory.
// orb_hamming_top2 grows a per-thread scratch block with cudaMalloc whenever a query needs more chunk partials than the last one;
// this variant keeps those partials in a fixed array that is part of the library image instead.
//
// The array below is 6 MiB and exists once per GPU that loads liborbgpu, from module load to process exit, whether or not this
// entry point is ever called.  Every call on a device uses the same array, so two calls whose kernels can run at the same time on
// that device (different streams, different host threads) overwrite each other's partials: order them on one stream or with events.
// The array keeps whatever the previous call left in it.  Nothing has to be cleared for the next call: the scan kernel writes
// every (chunk, query) partial that the merge kernel of the same call reads, and nothing else is read.
namespace {
constexpr size_t TOP2_STATIC_BYTES = (size_t)ORB_TOP2_STATIC_MAX_PARTIALS * (8 + 8 + 4 + 4);
static __device__ __align__(256) unsigned char g_top2_static[TOP2_STATIC_BYTES];

// In host code `g_top2_static` is only the host-side shadow the runtime registers for the symbol -- passing it to a kernel would
// hand the GPU a host address.  The device address differs per GPU and is looked up once per device, with that device current.
constexpr int TOP2_STATIC_MAX_DEVICES = 64;
std::mutex g_top2_static_mu;
void *g_top2_static_addr[TOP2_STATIC_MAX_DEVICES] = {};

int top2_static_address(int dev, void **out) {
    ORB_REQUIRE(dev >= 0 && dev < TOP2_STATIC_MAX_DEVICES, ORB_ERR_INVALID, "orb_hamming_top2_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_top2_static_mu);
    if (!g_top2_static_addr[dev]) {
        void *p = nullptr;
        ORB_CUDA_TRY(cudaGetSymbolAddress(&p, g_top2_static));   // device `dev` is current: the caller made it so
        ORB_REQUIRE(p && ((uintptr_t)p & 255) == 0, ORB_ERR_CUDA, "orb_hamming_top2_static: bad scratch address");
        g_top2_static_addr[dev] = p;
    }
    *out = g_top2_static_addr[dev];
    return ORB_OK;
}
}  // namespace

extern "C" {

int orb_hamming_top2_static(const uint8_t *dev_queries, int nq, const uint8_t *dev_db, int64_t ndb, int64_t idx_base,
                            int32_t *dev_best_dist, int64_t *dev_best_idx, int32_t *dev_second_dist, int64_t *dev_second_idx,
                            void *cuda_stream) {
    ORB_REQUIRE(dev_queries && dev_best_dist && dev_best_idx && dev_second_dist && dev_second_idx, ORB_ERR_INVALID, "null argument");
    ORB_REQUIRE(nq >= 1 && ndb >= 0 && (dev_db || ndb == 0), ORB_ERR_INVALID, "bad sizes");
    ORB_REQUIRE(nq <= ORB_TOP2_STATIC_MAX_QUERIES, ORB_ERR_CAPACITY,
                "orb_hamming_top2_static: %d queries exceed the fixed scratch (at most %d per call; split the query set, or use "
                "orb_hamming_top2, which allocates)", nq, ORB_TOP2_STATIC_MAX_QUERIES);
    int dev = 0, sms = 148;
    ORB_CUDA_TRY(cudaGetDevice(&dev));
    ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    void *scratch = nullptr;
    const int rc = top2_static_address(dev, &scratch);
    if (rc != ORB_OK) return rc;
    // same chunking rule as top2_device, with the chunk count additionally held to what the fixed array can take; the merged
    // result does not depend on how the rows are cut into chunks
    const int qblocks = (nq + T2_THREADS - 1) / T2_THREADS;
    long long chunks = std::max<long long>(1, (long long)sms * 8 / qblocks);
    chunks = std::min<long long>(chunks, std::max<long long>(1, (ndb + T2_TILE - 1) / T2_TILE));
    chunks = std::min<long long>(chunks, ORB_TOP2_STATIC_MAX_PARTIALS / nq);
    const long long rows_per_chunk = (std::max<long long>(ndb, 1) + chunks - 1) / chunks;
    chunks = std::max<long long>(1, (ndb + rows_per_chunk - 1) / rows_per_chunk);
    const size_t per = (size_t)chunks * nq;
    ORB_REQUIRE(per <= (size_t)ORB_TOP2_STATIC_MAX_PARTIALS, ORB_ERR_CAPACITY, "orb_hamming_top2_static: internal chunking overflow");
    int64_t *pi1 = (int64_t *)scratch, *pi2 = pi1 + per;
    int32_t *pd1 = (int32_t *)(pi2 + per), *pd2 = pd1 + per;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    k_hamming_top2<false><<<dim3(qblocks, (unsigned)chunks), T2_THREADS, 0, st>>>(dev_queries, nq, dev_db, ndb, rows_per_chunk, pd1, pd2,
                                                                                 pi1, pi2, idx_base, KfParams{});
    k_top2_merge<<<(nq + 127) / 128, 128, 0, st>>>((int)chunks, nq, pd1, pi1, pd2, pi2, dev_best_dist, dev_best_idx, dev_second_dist,
                                                   dev_second_idx);
    ORB_CUDA_TRY(cudaGetLastError());
    return ORB_OK;
}

// ref: Marker::searchMatches (ORB_SLAM2/src/Marker.cc:314-349) and naive_nn_search2 (AR-1.3/src/ORBMatcher.cpp:70-102):
// brute-force best / second-best of every query (keys2) over the train set (keys1), accepted when
// `min_dist <= (unsigned int)(sec_dist * 0.8) && min_dist <= 50`; sec_dist stays INT_MAX with fewer than two train rows.
__global__ void k_ratio_accept(int nq, const int32_t *__restrict__ d1, const int64_t *__restrict__ i1, const int32_t *__restrict__ d2,
                               const int64_t *__restrict__ i2, double ratio, int max_dist, int32_t *__restrict__ train_idx,
                               int32_t *__restrict__ dist) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    const unsigned int min_dist = i1[q] < 0 ? 0x7fffffffu : (unsigned int)d1[q];
    const unsigned int sec_dist = i2[q] < 0 ? 0x7fffffffu : (unsigned int)d2[q];
    const bool ok = i1[q] >= 0 && min_dist <= (unsigned int)(sec_dist * ratio) && min_dist <= (unsigned int)max_dist;
    train_idx[q] = ok ? (int32_t)i1[q] : -1;
    dist[q] = i1[q] < 0 ? 0x7fffffff : d1[q];
}

int orb_hamming_minmax(const uint8_t *train, int n_train, const uint8_t *query, int n_query, int loc, int32_t *min_max) {
    ORB_REQUIRE(n_query >= 0 && n_train >= 0 && min_max && (n_query == 0 || query) && (n_train == 0 || train), ORB_ERR_INVALID, "bad argument");
    min_max[0] = 0x7fffffff; min_max[1] = -1;
    if (n_query == 0 || n_train == 0) return ORB_OK;
    const size_t o_q = a256(16), o_t = o_q + a256((size_t)n_query * 32);
    ThreadCtx *T = nullptr;
    int rc = orbgpu_thread_ctx(&T, loc == ORB_HOST ? o_t + a256((size_t)n_train * 32) : o_q, 0);
    if (rc != ORB_OK) return rc;
    cudaStream_t st = T->stream;
    const uint8_t *d_q = query, *d_t = train;
    if (loc == ORB_HOST) {
        ORB_CUDA_TRY(cudaMemcpyAsync(T->d_buf + o_q, query, (size_t)n_query * 32, cudaMemcpyHostToDevice, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(T->d_buf + o_t, train, (size_t)n_train * 32, cudaMemcpyHostToDevice, st));
        d_q = T->d_buf + o_q; d_t = T->d_buf + o_t;
    }
    int *d_out = (int *)T->d_buf;
    ORB_CUDA_TRY(cudaMemcpyAsync(d_out, min_max, 2 * sizeof(int), cudaMemcpyHostToDevice, st));
    k_dist_minmax<<<(n_query + 7) / 8, 256, 0, st>>>(d_q, n_query, d_t, n_train, d_out);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(min_max, d_out, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

int orb_search_matches_ratio(const uint8_t *train, int n_train, const uint8_t *query, int n_query, double ratio, int max_dist,
                             int loc, int32_t *train_idx, int32_t *dist) {
    ORB_REQUIRE(n_query >= 0 && n_train >= 0 && (n_query == 0 || (query && train_idx && dist)) && (train || n_train == 0), ORB_ERR_INVALID,
                "bad argument");
    if (n_query == 0) return ORB_OK;
    const size_t o_tmp = 0, o_q = a256((size_t)n_query * 24), o_t = o_q + a256((size_t)n_query * 32),
                 o_out = o_t + a256((size_t)n_train * 32);
    ThreadCtx *T = nullptr;
    int rc = orbgpu_thread_ctx(&T, loc == ORB_HOST ? o_out + a256((size_t)n_query * 8) : o_q, 0);
    if (rc != ORB_OK) return rc;
    cudaStream_t st = T->stream;
    const uint8_t *d_q = query, *d_t = train;
    int32_t *o_idx = train_idx, *o_dist = dist;
    if (loc == ORB_HOST) {
        uint8_t *dq = T->d_buf + o_q, *dt = T->d_buf + o_t;
        ORB_CUDA_TRY(cudaMemcpyAsync(dq, query, (size_t)n_query * 32, cudaMemcpyHostToDevice, st));
        if (n_train) ORB_CUDA_TRY(cudaMemcpyAsync(dt, train, (size_t)n_train * 32, cudaMemcpyHostToDevice, st));
        d_q = dq; d_t = dt;
        o_idx = (int32_t *)(T->d_buf + o_out); o_dist = o_idx + n_query;
    }
    int64_t *t_i1 = (int64_t *)(T->d_buf + o_tmp), *t_i2 = t_i1 + n_query;
    int32_t *t_d1 = (int32_t *)(t_i2 + n_query), *t_d2 = t_d1 + n_query;
    rc = top2_device(d_q, n_query, d_t, n_train, 0, t_d1, t_i1, t_d2, t_i2, st, &t_scratch);
    if (rc != ORB_OK) return rc;
    k_ratio_accept<<<(n_query + 127) / 128, 128, 0, st>>>(n_query, t_d1, t_i1, t_d2, t_i2, ratio, max_dist, o_idx, o_dist);
    ORB_CUDA_TRY(cudaGetLastError());
    if (loc == ORB_HOST) {
        ORB_CUDA_TRY(cudaMemcpyAsync(train_idx, o_idx, (size_t)n_query * 4, cudaMemcpyDeviceToHost, st));
        ORB_CUDA_TRY(cudaMemcpyAsync(dist, o_dist, (size_t)n_query * 4, cudaMemcpyDeviceToHost, st));
    }
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

int orb_top2_merge(int nshards, int nq, const int32_t *best_dist, const int64_t *best_idx, const int32_t *second_dist,
                   const int64_t *second_idx, int32_t *out_best_dist, int64_t *out_best_idx, int32_t *out_second_dist,
                   int64_t *out_second_idx, int loc) {
    ORB_REQUIRE(nshards >= 1 && nq >= 1 && best_dist && best_idx && second_dist && second_idx && out_best_dist && out_best_idx &&
                    out_second_dist && out_second_idx, ORB_ERR_INVALID, "bad argument");
    const size_t per = (size_t)nshards * nq;
    ThreadCtx *T = nullptr;
    int rc = orbgpu_thread_ctx(&T, loc == ORB_HOST ? a256(per * 24) + a256((size_t)nq * 24) : 0, 0);
    if (rc != ORB_OK) return rc;
    cudaStream_t st = T->stream;
    if (loc == ORB_DEVICE) {
        k_top2_merge<<<(nq + 127) / 128, 128, 0, st>>>(nshards, nq, best_dist, best_idx, second_dist, second_idx, out_best_dist,
                                                       out_best_idx, out_second_dist, out_second_idx);
        ORB_CUDA_TRY(cudaGetLastError());
        return ORB_OK;
    }
    int64_t *i1 = (int64_t *)T->d_buf, *i2 = i1 + per;
    int32_t *d1 = (int32_t *)(i2 + per), *d2 = d1 + per;
    ORB_CUDA_TRY(cudaMemcpyAsync(i1, best_idx, per * 8, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(i2, second_idx, per * 8, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(d1, best_dist, per * 4, cudaMemcpyHostToDevice, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(d2, second_dist, per * 4, cudaMemcpyHostToDevice, st));
    int64_t *o_i1 = (int64_t *)(T->d_buf + a256(per * 24)), *o_i2 = o_i1 + nq;
    int32_t *o_d1 = (int32_t *)(o_i2 + nq), *o_d2 = o_d1 + nq;
    k_top2_merge<<<(nq + 127) / 128, 128, 0, st>>>(nshards, nq, d1, i1, d2, i2, o_d1, o_i1, o_d2, o_i2);
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaMemcpyAsync(out_best_idx, o_i1, (size_t)nq * 8, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(out_second_idx, o_i2, (size_t)nq * 8, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(out_best_dist, o_d1, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaMemcpyAsync(out_second_dist, o_d2, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA_TRY(cudaStreamSynchronize(st));
    return ORB_OK;
}

int orb_measure_popc_rate(double *gpopc_per_s) {
    ORB_REQUIRE(gpopc_per_s, ORB_ERR_INVALID, "null argument");
    int dev = 0, sms = 148;
    ORB_CUDA_TRY(cudaGetDevice(&dev));
    ORB_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int blocks = sms * 8, threads = 256, iters = 20000;
    DevBuf out;
    ORB_CUDA_TRY(out.alloc((size_t)blocks * threads * 4));
    cudaEvent_t e0, e1;
    ORB_CUDA_TRY(cudaEventCreate(&e0)); ORB_CUDA_TRY(cudaEventCreate(&e1));
    k_popc_rate<<<blocks, threads>>>((unsigned *)out.p, 2000);   // warm-up
    double best = 0;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        k_popc_rate<<<blocks, threads>>>((unsigned *)out.p, iters);
        cudaEventRecord(e1);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) { cudaEventDestroy(e0); cudaEventDestroy(e1); set_error("popc bench: %s", cudaGetErrorString(e)); return ORB_ERR_CUDA; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        best = std::max(best, (double)blocks * threads * iters * 8.0 / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    *gpopc_per_s = best;
    return ORB_OK;
}

// ---- sharded database ---------------------------------------------------------------------

int orb_db_create(int64_t row_begin, int64_t row_end, int device, orb_db **out) {
    ORB_REQUIRE(out && row_begin >= 0 && row_end >= row_begin, ORB_ERR_INVALID, "bad row range");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    orb_db *db = new orb_db();
    if (device >= 0) { if (cudaSetDevice(device) != cudaSuccess) { delete db; set_error("cudaSetDevice(%d) failed", device); return ORB_ERR_CUDA; } db->device = device; }
    else cudaGetDevice(&db->device);
    db->row_begin = row_begin; db->row_end = row_end;
    const size_t bytes = (size_t)(row_end - row_begin) * 32;
    e = cudaMalloc(&db->d_rows, bytes ? bytes : 32);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&db->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) { set_error("orb_db_create: %s", cudaGetErrorString(e)); if (db->d_rows) cudaFree(db->d_rows); delete db; return ORB_ERR_CUDA; }
    *out = db;
    return ORB_OK;
}

void orb_db_destroy(orb_db *db) {
    if (!db) return;
    cudaSetDevice(db->device);
    cudaStreamSynchronize(db->stream);
    cudaFree(db->d_rows);
    if (db->scratch.buf) cudaFree(db->scratch.buf);
    if (db->d_io) cudaFree(db->d_io);
    cudaStreamDestroy(db->stream);
    delete db;
}

int orb_db_upload(orb_db *db, int64_t first_row, const uint8_t *rows, int64_t nrows, int loc) {
    ORB_REQUIRE(db && rows && first_row >= db->row_begin && first_row + nrows <= db->row_end && nrows >= 0, ORB_ERR_INVALID,
                "rows outside the shard");
    ORB_CUDA_TRY(cudaSetDevice(db->device));
    ORB_CUDA_TRY(cudaMemcpyAsync(db->d_rows + (size_t)(first_row - db->row_begin) * 32, rows, (size_t)nrows * 32,
                                 loc == ORB_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, db->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(db->stream));
    return ORB_OK;
}

int orb_db_fill_synthetic(orb_db *db, uint64_t seed) {
    ORB_REQUIRE(db, ORB_ERR_INVALID, "null handle");
    ORB_CUDA_TRY(cudaSetDevice(db->device));
    const long long nwords = (db->row_end - db->row_begin) * 4;
    if (nwords > 0) {
        k_db_fill<<<2048, 256, 0, db->stream>>>((unsigned long long *)db->d_rows, nwords, db->row_begin * 4, seed);
        db->launches++;
    }
    ORB_CUDA_TRY(cudaGetLastError());
    ORB_CUDA_TRY(cudaStreamSynchronize(db->stream));
    return ORB_OK;
}

int orb_db_row(orb_db *db, int64_t global_row, uint8_t *out32) {
    ORB_REQUIRE(db && out32 && global_row >= db->row_begin && global_row < db->row_end, ORB_ERR_INVALID, "row outside the shard");
    ORB_CUDA_TRY(cudaSetDevice(db->device));
    ORB_CUDA_TRY(cudaMemcpy(out32, db->d_rows + (size_t)(global_row - db->row_begin) * 32, 32, cudaMemcpyDeviceToHost));
    return ORB_OK;
}

uint8_t *orb_db_device_ptr(orb_db *db) { return db ? db->d_rows : nullptr; }
void *orb_db_stream(orb_db *db) { return db ? (void *)db->stream : nullptr; }
int64_t orb_db_launch_count(const orb_db *db) { return db ? db->launches : 0; }

int orb_db_merge_top2(orb_db *db, int nshards, int nq, const int32_t *best_dist, const int64_t *best_idx, const int32_t *second_dist,
                      const int64_t *second_idx, int32_t *out_best_dist, int64_t *out_best_idx, int32_t *out_second_dist,
                      int64_t *out_second_idx) {
    ORB_REQUIRE(db && nshards >= 1 && nq >= 1 && best_dist && best_idx && second_dist && second_idx && out_best_dist && out_best_idx &&
                    out_second_dist && out_second_idx, ORB_ERR_INVALID, "bad argument");
    ORB_CUDA_TRY(cudaSetDevice(db->device));
    k_top2_merge<<<(nq + 127) / 128, 128, 0, db->stream>>>(nshards, nq, best_dist, best_idx, second_dist, second_idx, out_best_dist,
                                                           out_best_idx, out_second_dist, out_second_idx);
    db->launches++;
    ORB_CUDA_TRY(cudaGetLastError());
    return ORB_OK;
}

int orb_db_query(orb_db *db, const uint8_t *queries, int nq, int32_t *best_dist, int64_t *best_idx, int32_t *second_dist,
                 int64_t *second_idx, int loc) {
    ORB_REQUIRE(db && queries && nq >= 1 && best_dist && best_idx && second_dist && second_idx, ORB_ERR_INVALID, "bad argument");
    ORB_CUDA_TRY(cudaSetDevice(db->device));
    const long long n = db->row_end - db->row_begin;
    if (loc == ORB_DEVICE) {
        int rc = top2_device(queries, nq, db->d_rows, n, db->row_begin, best_dist, best_idx, second_dist, second_idx, db->stream,
                             &db->scratch);
        db->launches += 2;
        return rc;
    }
    const size_t need = (size_t)nq * (32 + 24) + 256;
    if (db->io_bytes < need) {
        if (db->d_io) cudaFree(db->d_io);
        db->d_io = nullptr; db->io_bytes = 0;
        ORB_CUDA_TRY(cudaMalloc(&db->d_io, need));
        db->io_bytes = need;
    }
    int64_t *o_i1 = (int64_t *)db->d_io, *o_i2 = o_i1 + nq;
    int32_t *o_d1 = (int32_t *)(o_i2 + nq), *o_d2 = o_d1 + nq;
    uint8_t *dq = (uint8_t *)(o_d2 + nq);
    dq = (uint8_t *)(((uintptr_t)dq + 15) & ~(uintptr_t)15);
    ORB_CUDA_TRY(cudaMemcpyAsync(dq, queries, (size_t)nq * 32, cudaMemcpyHostToDevice, db->stream));
    int rc = top2_device(dq, nq, db->d_rows, n, db->row_begin, o_d1, o_i1, o_d2, o_i2, db->stream, &db->scratch);
    db->launches += 2;
    if (rc != ORB_OK) return rc;
    ORB_CUDA_TRY(cudaMemcpyAsync(best_idx, o_i1, (size_t)nq * 8, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(second_idx, o_i2, (size_t)nq * 8, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(best_dist, o_d1, (size_t)nq * 4, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(second_dist, o_d2, (size_t)nq * 4, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(db->stream));
    return ORB_OK;
}

int orb_db_query_keyframes(orb_db *db, const uint8_t *queries, int nq, int rows_per_keyframe, int th_low, float ratio,
                           int32_t *best_dist, int64_t *best_idx, int32_t *second_dist, int64_t *second_idx, int32_t *kf_counts,
                           int loc) {
    ORB_REQUIRE(db && queries && nq >= 1 && best_dist && best_idx && second_dist && second_idx && kf_counts, ORB_ERR_INVALID, "bad argument");
    const long long n = db->row_end - db->row_begin;
    ORB_REQUIRE(rows_per_keyframe >= 1 && n % rows_per_keyframe == 0 && db->row_begin % rows_per_keyframe == 0, ORB_ERR_INVALID,
                "the shard [%lld, %lld) must hold whole keyframes of %d rows", db->row_begin, db->row_end, rows_per_keyframe);
    ORB_CUDA_TRY(cudaSetDevice(db->device));
    const long long nkf = n / rows_per_keyframe;
    KfParams kf{rows_per_keyframe, th_low, ratio, kf_counts};
    if (loc == ORB_DEVICE) {
        int rc = top2_device(queries, nq, db->d_rows, n, db->row_begin, best_dist, best_idx, second_dist, second_idx, db->stream,
                             &db->scratch, &kf);
        db->launches += 2;
        return rc;
    }
    const size_t need = (size_t)nq * (32 + 24) + (size_t)nkf * 4 + 512;
    if (db->io_bytes < need) {
        if (db->d_io) cudaFree(db->d_io);
        db->d_io = nullptr; db->io_bytes = 0;
        ORB_CUDA_TRY(cudaMalloc(&db->d_io, need));
        db->io_bytes = need;
    }
    int64_t *o_i1 = (int64_t *)db->d_io, *o_i2 = o_i1 + nq;
    int32_t *o_d1 = (int32_t *)(o_i2 + nq), *o_d2 = o_d1 + nq;
    uint8_t *dq = (uint8_t *)(o_d2 + nq);
    dq = (uint8_t *)(((uintptr_t)dq + 15) & ~(uintptr_t)15);
    kf.counts = (int32_t *)(((uintptr_t)(dq + (size_t)nq * 32) + 15) & ~(uintptr_t)15);
    ORB_CUDA_TRY(cudaMemcpyAsync(dq, queries, (size_t)nq * 32, cudaMemcpyHostToDevice, db->stream));
    int rc = top2_device(dq, nq, db->d_rows, n, db->row_begin, o_d1, o_i1, o_d2, o_i2, db->stream, &db->scratch, &kf);
    db->launches += 2;
    if (rc != ORB_OK) return rc;
    ORB_CUDA_TRY(cudaMemcpyAsync(best_idx, o_i1, (size_t)nq * 8, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(second_idx, o_i2, (size_t)nq * 8, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(best_dist, o_d1, (size_t)nq * 4, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(second_dist, o_d2, (size_t)nq * 4, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(kf_counts, kf.counts, (size_t)nkf * 4, cudaMemcpyDeviceToHost, db->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(db->stream));
    return ORB_OK;
}

}  // extern "C"
