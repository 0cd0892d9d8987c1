// liborbgpu: the bag-of-words vocabulary on the device -- DBoW2's transform().
//
// ref: ORB_SLAM2/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h (tree descent :1217-1255, BowVector/FeatureVector
// assembly :1126-1193, text loader :1338-1425), FORB.cpp:81-101 (distance), BowVector.cpp:33-88, FeatureVector.cpp:30-44;
// callers Frame::ComputeBoW (src/Frame.cc:400-407) and KeyFrame::ComputeBoW (src/KeyFrame.cc:62-70).
//
// Layout in HBM: the tree is renumbered breadth first so that the children of a node occupy CONSECUTIVE slots in
// the reference's child order; one slot = 32 descriptor bytes (two 16-byte loads) + a 32-byte record
// {first child slot, child count, reference node id, word id, weight}.  ORBvoc (k=10, L=6: 1.1 M nodes) is 71 MB and
// stays L2 resident.  Two kernels per call, any number of frames:
//   k_vocab_descend   16 lanes per feature, lane = child: 2x LDG.128 + 8 POPC per child, (distance, child) minimum by
//                     shuffles = the reference's strict `d < best_d`, first child wins
//   k_bow_assemble    one CTA per frame: bitonic sort of (word, feature) and (node, feature) keys in shared memory,
//                     segment heads by block scan; the double sums and the L1/L2 norm are accumulated sequentially in
//                     the reference's order (ascending word id, features in index order) so the values are bit-exact
#include "orb_common.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

using namespace orbgpu;

namespace {

struct __align__(16) SlotRec {
    int first, nchild, node, word;
    double weight, pad;
};
static_assert(sizeof(SlotRec) == 32, "SlotRec must be 32 bytes");

constexpr int GROUP = 16;             // lanes per feature in the descent
constexpr int DESCEND_THREADS = 128;
constexpr int ASM_THREADS = 1024;
constexpr int SMEM_SORT_MAX = 4096;   // keys sorted in shared memory up to this many (32 KB); beyond that in global scratch
constexpr unsigned long long KEY_NONE = ~0ull;

enum { VERR_COUNT = 1 };

__global__ void __launch_bounds__(DESCEND_THREADS)
k_vocab_descend(const uint4 *__restrict__ tdesc, const SlotRec *__restrict__ trec, const uint4 *__restrict__ feats,
                const int *__restrict__ n_per_frame, int batch, int stride_rows, int nid_level,
                int *__restrict__ word, int *__restrict__ node, double *__restrict__ weight, int *__restrict__ err) {
    const unsigned FULL = 0xffffffffu;
    const int gl = threadIdx.x & (GROUP - 1);
    const long long f = ((long long)blockIdx.x * DESCEND_THREADS + threadIdx.x) / GROUP;
    const int b = (int)(f / stride_rows), i = (int)(f % stride_rows);
    bool active = false;
    if (b < batch) {
        int n = n_per_frame[b];
        if (n > stride_rows || n < 0) { if (gl == 0 && i == 0) atomicOr(err, VERR_COUNT); n = max(0, min(n, stride_rows)); }
        active = i < n;
    }
    uint32_t q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (active) {
        const uint4 a = __ldg(feats + 2 * f), c = __ldg(feats + 2 * f + 1);
        q[0] = a.x; q[1] = a.y; q[2] = a.z; q[3] = a.w; q[4] = c.x; q[5] = c.y; q[6] = c.z; q[7] = c.w;
    }
    // root record
    int first = 0, nchild = 0, cur = 0, nid = 0, level = 0;
    bool nid_set = nid_level <= 0;   // ref :1227: root when the requested level is above the tree
    if (active) { const int4 r = __ldg(reinterpret_cast<const int4 *>(trec)); first = r.x; nchild = r.y; }
    bool done = !active || nchild == 0;
    while (__any_sync(FULL, !done)) {
        unsigned best = 0xffffffffu;
        for (int c0 = 0; __any_sync(FULL, !done && c0 < nchild); c0 += GROUP) {
            const int c = c0 + gl;
            if (!done && c < nchild) {
                const uint4 a = __ldg(tdesc + 2 * (long long)(first + c)), d = __ldg(tdesc + 2 * (long long)(first + c) + 1);
                const int dist = __popc(q[0] ^ a.x) + __popc(q[1] ^ a.y) + __popc(q[2] ^ a.z) + __popc(q[3] ^ a.w) +
                                 __popc(q[4] ^ d.x) + __popc(q[5] ^ d.y) + __popc(q[6] ^ d.z) + __popc(q[7] ^ d.w);
                best = min(best, ((unsigned)dist << 20) | (unsigned)c);   // (distance, child order): strict '<', first child wins
            }
        }
#pragma unroll
        for (int o = GROUP / 2; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(FULL, best, o));
        if (!done) {
            cur = first + (int)(best & 0xfffffu);
            const int4 r = __ldg(reinterpret_cast<const int4 *>(trec + cur));
            first = r.x; nchild = r.y;
            ++level;
            if (level == nid_level) { nid = r.z; nid_set = true; }
            if (nchild == 0) {
                done = true;
                if (gl == 0) {
                    word[f] = r.w;
                    node[f] = nid_set ? nid : r.z;
                    weight[f] = trec[cur].weight;
                }
            }
        }
    }
}

// ---- block helpers for the assembly kernel -------------------------------------------------------------------------

__device__ __forceinline__ void bitonic_sort(unsigned long long *keys, int P2) {
    for (int k = 2; k <= P2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < P2; t += ASM_THREADS) {
                const int p = t ^ j;
                if (p > t) {
                    const unsigned long long a = keys[t], b = keys[p];
                    const bool up = (t & k) == 0;
                    if ((a > b) == up) { keys[t] = b; keys[p] = a; }
                }
            }
            __syncthreads();
        }
    }
}

// exclusive scan of one int per thread over the CTA; *total gets the sum
__device__ __forceinline__ int block_exscan(int v, int *s_warp, int *total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_warp[w] = x;
    __syncthreads();
    if (w == 0) {
        int s = s_warp[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
        s_warp[lane] = s;   // inclusive over warps
    }
    __syncthreads();
    const int base = w ? s_warp[w - 1] : 0;
    *total = s_warp[ASM_THREADS / 32 - 1];
    __syncthreads();
    return base + x - v;
}


struct AsmOut {
    int capacity;
    int *bow_word; double *bow_value; int *n_bow;
    int *fv_node; int *fv_off; int *fv_feat; int *n_fv;
};

// One sorted key array -> its segments.  Returns (through *total) the number of distinct high words among the valid
// keys; calls emit(segment index, position of the segment head) for every head owned by this thread.
template <typename Emit>
__device__ __forceinline__ void for_each_segment(const unsigned long long *keys, int P2, int *s_warp, int *total, Emit emit) {
    const int items = max(1, P2 / ASM_THREADS);
    const int lo = min(P2, (int)threadIdx.x * items), hi = min(P2, lo + items);
    int cnt = 0;
    for (int i = lo; i < hi; i++) {
        const unsigned long long k = keys[i];
        cnt += (k != KEY_NONE) && (i == 0 || (unsigned)(keys[i - 1] >> 32) != (unsigned)(k >> 32));
    }
    int seg = block_exscan(cnt, s_warp, total);
    for (int i = lo; i < hi; i++) {
        const unsigned long long k = keys[i];
        if ((k != KEY_NONE) && (i == 0 || (unsigned)(keys[i - 1] >> 32) != (unsigned)(k >> 32))) emit(seg++, i);
    }
}

__global__ void __launch_bounds__(ASM_THREADS)
k_bow_assemble(const int *__restrict__ word, const int *__restrict__ node, const double *__restrict__ weight,
               const int *__restrict__ n_per_frame, int stride_rows, int P2, unsigned long long *gkeys, int tf, int must,
               int l2, AsmOut out) {
    extern __shared__ unsigned long long s_keys[];
    __shared__ int s_warp[32];
    __shared__ double s_norm;
    __shared__ int s_valid;
    const int b = blockIdx.x;
    unsigned long long *keys = gkeys ? gkeys + (size_t)b * P2 : s_keys;
    const int n = max(0, min(n_per_frame[b], stride_rows));
    const long long base = (long long)b * stride_rows;
    const double *wt = weight + base;
    int *bow_word = out.bow_word + (size_t)b * out.capacity;
    double *bow_value = out.bow_value + (size_t)b * out.capacity;
    int *fv_node = out.fv_node + (size_t)b * out.capacity;
    int *fv_feat = out.fv_feat + (size_t)b * out.capacity;
    int *fv_off = out.fv_off + (size_t)b * (out.capacity + 1);

    // ---- BowVector: std::map<WordId, WordValue> in iteration order (ref :1154-1161 / :1181-1188)
    for (int t = threadIdx.x; t < P2; t += ASM_THREADS)
        keys[t] = (t < n && wt[t] > 0) ? ((unsigned long long)(unsigned)word[base + t] << 32) | (unsigned)t : KEY_NONE;
    __syncthreads();
    bitonic_sort(keys, P2);
    int nbow = 0;
    for_each_segment(keys, P2, s_warp, &nbow, [&](int seg, int i) {
        const unsigned long long k = keys[i];
        const unsigned w = (unsigned)(k >> 32);
        double sum = wt[(unsigned)k];                       // addWeight / addIfNotExist insert the first value
        if (tf)                                             // addWeight: `vit->second += v` in feature order (BowVector.cpp:33-45)
            for (int j = i + 1; j < P2 && (unsigned)(keys[j] >> 32) == w; j++) sum += wt[(unsigned)keys[j]];
        bow_word[seg] = (int)w;
        bow_value[seg] = sum;
    });
    __syncthreads();
    if (tf && !must && nbow > 0) {                          // ref :1163-1169
        const double nd = (double)nbow;
        for (int t = threadIdx.x; t < nbow; t += ASM_THREADS) bow_value[t] /= nd;
    }
    if (must) {                                             // BowVector::normalize (BowVector.cpp:62-88): one running sum in map order
        if (threadIdx.x == 0) {
            double norm = 0.0;
            if (!l2) {
#pragma unroll 8
                for (int t = 0; t < nbow; t++) norm += fabs(bow_value[t]);
            } else {
#pragma unroll 8
                for (int t = 0; t < nbow; t++) norm += bow_value[t] * bow_value[t];
                norm = sqrt(norm);
            }
            s_norm = norm;
        }
        __syncthreads();
        const double norm = s_norm;
        if (norm > 0.0)
            for (int t = threadIdx.x; t < nbow; t += ASM_THREADS) bow_value[t] /= norm;
    }
    if (threadIdx.x == 0) { out.n_bow[b] = nbow; s_valid = 0; }
    __syncthreads();

    // ---- FeatureVector: std::map<NodeId, vector<feature>> (FeatureVector.cpp:30-44), features ascending = push_back order
    for (int t = threadIdx.x; t < P2; t += ASM_THREADS)
        keys[t] = (t < n && wt[t] > 0) ? ((unsigned long long)(unsigned)node[base + t] << 32) | (unsigned)t : KEY_NONE;
    __syncthreads();
    bitonic_sort(keys, P2);
    for (int t = threadIdx.x; t < P2; t += ASM_THREADS) {
        const unsigned long long k = keys[t];
        if (k != KEY_NONE) {
            fv_feat[t] = (int)(unsigned)k;
            if (t == P2 - 1 || keys[t + 1] == KEY_NONE) s_valid = t + 1;
        }
    }
    int nfv = 0;
    for_each_segment(keys, P2, s_warp, &nfv, [&](int seg, int i) {
        fv_node[seg] = (int)(unsigned)(keys[i] >> 32);
        fv_off[seg] = i;
    });
    __syncthreads();
    if (threadIdx.x == 0) { fv_off[nfv] = s_valid; out.n_fv[b] = nfv; }
}

}  // namespace

// ---- host side -------------------------------------------------------------------------------------------------------

struct orb_vocab {
    int device = 0;
    int k = 0, L = 0, scoring = 0, weighting = 0, n_nodes = 0, n_words = 0;
    // the tree as the caller gave it (orb_vocab_nodes) ...
    std::vector<int32_t> parent, word_id;
    std::vector<double> weight;
    std::vector<uint8_t> desc;
    // ... and as the kernels walk it
    uint4 *d_tdesc = nullptr;
    SlotRec *d_trec = nullptr;
    bool root_has_children = false;
    cudaStream_t stream = nullptr;
    std::mutex mu;                      // transform() is const in the reference and called from three threads
    uint8_t *d_scratch = nullptr; size_t scratch_bytes = 0;
    int *d_err = nullptr;
    int64_t launches = 0;
};

namespace {

int ensure_scratch(orb_vocab *v, size_t bytes) {
    if (bytes <= v->scratch_bytes) return ORB_OK;
    ORB_CUDA_TRY(cudaStreamSynchronize(v->stream));
    if (v->d_scratch) cudaFree(v->d_scratch);
    v->d_scratch = nullptr; v->scratch_bytes = 0;
    ORB_CUDA_TRY(cudaMalloc(&v->d_scratch, bytes));
    v->scratch_bytes = bytes;
    return ORB_OK;
}

size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

// Build the breadth-first slot layout and upload it.
int finish_create(orb_vocab *v) {
    const int n = v->n_nodes;
    std::vector<int> nchild(n, 0), start(n + 1, 0);
    for (int i = 1; i < n; i++) {
        ORB_REQUIRE(v->parent[i] >= 0 && v->parent[i] < n && v->parent[i] != i, ORB_ERR_INVALID, "node %d: bad parent %d", i, v->parent[i]);
        nchild[v->parent[i]]++;
    }
    for (int i = 0; i < n; i++) start[i + 1] = start[i] + nchild[i];
    std::vector<int> kids(std::max(n - 1, 1)), fill(start.begin(), start.end() - 1);
    for (int i = 1; i < n; i++) kids[fill[v->parent[i]]++] = i;   // ascending id within a parent = the reference's push_back order
    std::vector<int> slot_of(n, -1), order;
    order.reserve(n);
    order.push_back(0); slot_of[0] = 0;
    std::vector<SlotRec> rec(n);
    std::vector<uint8_t> sdesc((size_t)n * 32, 0);
    for (size_t h = 0; h < order.size(); h++) {
        const int id = order[h];
        SlotRec &r = rec[h];
        r.first = (int)order.size(); r.nchild = nchild[id]; r.node = id; r.word = v->word_id[id]; r.weight = v->weight[id]; r.pad = 0;
        ORB_REQUIRE(nchild[id] < (1 << 20), ORB_ERR_INVALID, "node %d has too many children", id);
        for (int c = 0; c < nchild[id]; c++) {
            const int ch = kids[start[id] + c];
            ORB_REQUIRE(slot_of[ch] < 0, ORB_ERR_INVALID, "node %d reached twice (cycle)", ch);
            slot_of[ch] = (int)order.size();
            order.push_back(ch);
        }
        std::memcpy(&sdesc[h * 32], &v->desc[(size_t)id * 32], 32);
    }
    ORB_REQUIRE((int)order.size() == n, ORB_ERR_INVALID, "%d of %d nodes are not reachable from the root", n - (int)order.size(), n);
    v->root_has_children = nchild[0] > 0;
    ORB_CUDA_TRY(cudaMalloc(&v->d_tdesc, (size_t)n * 32));
    ORB_CUDA_TRY(cudaMalloc(&v->d_trec, (size_t)n * sizeof(SlotRec)));
    ORB_CUDA_TRY(cudaMalloc(&v->d_err, sizeof(int)));
    ORB_CUDA_TRY(cudaMemset(v->d_err, 0, sizeof(int)));
    ORB_CUDA_TRY(cudaMemcpy(v->d_tdesc, sdesc.data(), (size_t)n * 32, cudaMemcpyHostToDevice));
    ORB_CUDA_TRY(cudaMemcpy(v->d_trec, rec.data(), (size_t)n * sizeof(SlotRec), cudaMemcpyHostToDevice));
    ORB_CUDA_TRY(cudaStreamCreateWithFlags(&v->stream, cudaStreamNonBlocking));
    return ORB_OK;
}

int new_vocab(int device, orb_vocab **pv) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    ORB_REQUIRE(e == cudaSuccess && ndev > 0, ORB_ERR_CUDA, "no CUDA device (%s); liborbgpu has no CPU fallback", cudaGetErrorString(e));
    orb_vocab *v = new orb_vocab();
    if (device >= 0) {
        if (cudaSetDevice(device) != cudaSuccess) { delete v; set_error("cudaSetDevice(%d) failed", device); return ORB_ERR_CUDA; }
        v->device = device;
    } else cudaGetDevice(&v->device);
    *pv = v;
    return ORB_OK;
}

}  // namespace

extern "C" {

void orb_vocab_destroy(orb_vocab *v) {
    if (!v) return;
    cudaSetDevice(v->device);
    if (v->stream) cudaStreamSynchronize(v->stream);
    if (v->d_tdesc) cudaFree(v->d_tdesc);
    if (v->d_trec) cudaFree(v->d_trec);
    if (v->d_err) cudaFree(v->d_err);
    if (v->d_scratch) cudaFree(v->d_scratch);
    if (v->stream) cudaStreamDestroy(v->stream);
    delete v;
}

int orb_vocab_create(const orb_vocab_desc *d, int device, orb_vocab **out) {
    ORB_REQUIRE(d && out && d->n_nodes >= 1 && d->parent && d->descriptors && d->weight && d->word_id, ORB_ERR_INVALID, "bad vocabulary description");
    ORB_REQUIRE(d->scoring >= 0 && d->scoring <= 5 && d->weighting >= 0 && d->weighting <= 3, ORB_ERR_INVALID, "bad scoring/weighting type");
    orb_vocab *v = nullptr;
    int rc = new_vocab(device, &v);
    if (rc != ORB_OK) return rc;
    v->k = d->k; v->L = d->L; v->scoring = d->scoring; v->weighting = d->weighting; v->n_nodes = d->n_nodes;
    v->parent.assign(d->parent, d->parent + d->n_nodes);
    v->word_id.assign(d->word_id, d->word_id + d->n_nodes);
    v->weight.assign(d->weight, d->weight + d->n_nodes);
    v->desc.assign(d->descriptors, d->descriptors + (size_t)32 * d->n_nodes);
    v->parent[0] = 0;
    std::vector<char> has_child(d->n_nodes, 0);
    for (int i = 1; i < d->n_nodes; i++) if (v->parent[i] >= 0 && v->parent[i] < d->n_nodes) has_child[v->parent[i]] = 1;
    for (int i = 1; i < d->n_nodes; i++) v->n_words += !has_child[i];
    rc = finish_create(v);
    if (rc != ORB_OK) { orb_vocab_destroy(v); return rc; }
    *out = v;
    return ORB_OK;
}

// ref: TemplatedVocabulary::loadFromTextFile, TemplatedVocabulary.h:1338-1425 (include/orbgpu.h: the empty read after the
// final newline, where the reference uses an unwritten `pid`, is not a node).
// Plain host parsing: strtol/strtod read what the reference's stringstream extractions read.
int orb_vocab_load_text(const char *path, int device, orb_vocab **out) {
    ORB_REQUIRE(path && out, ORB_ERR_INVALID, "bad argument");
    FILE *fp = fopen(path, "rb");   // C stdio: no iostream state shared with whatever libstdc++ the host process loaded
    ORB_REQUIRE(fp, ORB_ERR_INVALID, "cannot open %s", path);
    std::string text;
    {
        std::vector<char> buf(1 << 20);
        size_t got;
        while ((got = fread(buf.data(), 1, buf.size(), fp)) > 0) text.append(buf.data(), got);
        fclose(fp);
    }
    const char *p = text.c_str(), *end = p + text.size();
    auto next_line = [&](const char *&b, const char *&e) {   // getline: up to '\n' or the end of the file; sets eof like the stream would
        b = p;
        const char *nl = (const char *)memchr(p, '\n', end - p);
        if (nl) { e = nl; p = nl + 1; return false; }
        e = end; p = end;
        return true;   // hit eof while reading this line
    };
    const char *lb, *le;
    bool eof = next_line(lb, le);
    std::string line(lb, le);
    char *q = nullptr;
    const char *c = line.c_str();
    long hdr[4] = {0, 0, 0, 0};
    for (int i = 0; i < 4; i++) { hdr[i] = strtol(c, &q, 10); if (q == c) { hdr[i] = 0; break; } c = q; }
    ORB_REQUIRE(!(hdr[0] < 0 || hdr[0] > 20 || hdr[1] < 1 || hdr[1] > 10 || hdr[2] < 0 || hdr[2] > 5 || hdr[3] < 0 || hdr[3] > 3),
                ORB_ERR_INVALID, "%s: not a vocabulary text file (ref :1360-1364)", path);
    orb_vocab *v = nullptr;
    int rc = new_vocab(device, &v);
    if (rc != ORB_OK) return rc;
    v->k = (int)hdr[0]; v->L = (int)hdr[1]; v->scoring = (int)hdr[2]; v->weighting = (int)hdr[3];
    v->parent.push_back(0); v->word_id.push_back(0); v->weight.push_back(0.0); v->desc.resize(32, 0);
    while (!eof) {
        eof = next_line(lb, le);
        line.assign(lb, le);
        c = line.c_str();
        bool ok = true;   // once an extraction fails the reference's stream stays failed: everything after reads as 0 / is skipped
        auto read_long = [&]() -> long { if (!ok) return 0; long x = strtol(c, &q, 10); if (q == c) { ok = false; return 0; } c = q; return x; };
        const long pid = read_long();
        if (!ok) {
            if (eof && line.find_first_not_of(" \t\r") == std::string::npos) break;   // the empty read after the last newline
            set_error("%s: node line %d has no parent id", path, (int)v->parent.size());
            orb_vocab_destroy(v);
            return ORB_ERR_INVALID;
        }
        const long leaf = read_long();
        uint8_t d[32] = {0};
        for (int i = 0; i < 32; i++) { const char *c0 = c; const bool was = ok; const long b = read_long(); if (was && ok) d[i] = (uint8_t)b; (void)c0; }
        double w = 0.0;
        if (ok) { w = strtod(c, &q); if (q == c) w = 0.0; }
        const int nid = (int)v->parent.size();
        if (pid < 0 || pid > nid) { orb_vocab_destroy(v); set_error("%s: node %d names parent %ld", path, nid, pid); return ORB_ERR_INVALID; }
        v->parent.push_back((int32_t)pid);
        v->desc.insert(v->desc.end(), d, d + 32);
        v->weight.push_back(w);
        v->word_id.push_back(leaf > 0 ? v->n_words++ : 0);   // ref :1409-1416
    }
    v->n_nodes = (int)v->parent.size();
    rc = finish_create(v);
    if (rc != ORB_OK) { orb_vocab_destroy(v); return rc; }
    *out = v;
    return ORB_OK;
}

int orb_vocab_info(const orb_vocab *v, int32_t *info) {
    ORB_REQUIRE(v && info, ORB_ERR_INVALID, "bad argument");
    info[0] = v->k; info[1] = v->L; info[2] = v->scoring; info[3] = v->weighting; info[4] = v->n_nodes; info[5] = v->n_words;
    return ORB_OK;
}

int orb_vocab_nodes(const orb_vocab *v, int32_t *parent, uint8_t *descriptors, double *weight, int32_t *word_id) {
    ORB_REQUIRE(v, ORB_ERR_INVALID, "bad argument");
    const size_t n = v->n_nodes;
    if (parent) std::memcpy(parent, v->parent.data(), n * 4);
    if (descriptors) std::memcpy(descriptors, v->desc.data(), n * 32);
    if (weight) std::memcpy(weight, v->weight.data(), n * 8);
    if (word_id) std::memcpy(word_id, v->word_id.data(), n * 4);
    return ORB_OK;
}

void *orb_vocab_stream(orb_vocab *v) { return v ? (void *)v->stream : nullptr; }
int64_t orb_vocab_launch_count(const orb_vocab *v) { return v ? v->launches : 0; }

int orb_vocab_sync(orb_vocab *v) {
    ORB_REQUIRE(v, ORB_ERR_INVALID, "bad argument");
    std::lock_guard<std::mutex> lock(v->mu);
    ORB_CUDA_TRY(cudaSetDevice(v->device));
    int err = 0;
    ORB_CUDA_TRY(cudaMemcpyAsync(&err, v->d_err, sizeof(int), cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(v->stream));
    if (err) {
        cudaMemsetAsync(v->d_err, 0, sizeof(int), v->stream);
        set_error("orb_vocab_transform: a frame's feature count exceeds stride_rows");
        return ORB_ERR_CAPACITY;
    }
    return ORB_OK;
}

}  // extern "C"

namespace {

int next_pow2(int x) { int p = 32; while (p < x) p <<= 1; return p; }

// launches the descent over batch*stride_rows feature slots (device pointers only)
int launch_descend(orb_vocab *v, const uint8_t *d_desc, const int *d_n, int batch, int stride_rows, int levelsup, int *d_word,
                   int *d_node, double *d_weight) {
    const long long feats = (long long)batch * stride_rows;
    const long long blocks = (feats * GROUP + DESCEND_THREADS - 1) / DESCEND_THREADS;
    ORB_REQUIRE(blocks < 0x7fffffffLL, ORB_ERR_INVALID, "too many features in one call");
    k_vocab_descend<<<(unsigned)blocks, DESCEND_THREADS, 0, v->stream>>>(v->d_tdesc, v->d_trec, reinterpret_cast<const uint4 *>(d_desc), d_n,
                                                                         batch, stride_rows, v->L - levelsup, d_word, d_node, d_weight, v->d_err);
    ORB_CUDA_TRY(cudaGetLastError());
    v->launches++;
    return ORB_OK;
}

struct Carve {   // bump allocator over the handle's scratch buffer
    uint8_t *base; size_t off = 0;
    template <typename T> T *take(size_t count) { T *p = reinterpret_cast<T *>(base + off); off += align256(count * sizeof(T)); return p; }
};

}  // namespace

extern "C" {

int orb_vocab_transform_features(orb_vocab *v, const uint8_t *descriptors, int n, int loc, int levelsup, int32_t *word,
                                 int32_t *node, double *weight) {
    ORB_REQUIRE(v && n >= 0 && (n == 0 || (descriptors && word && node && weight)), ORB_ERR_INVALID, "bad argument");
    ORB_REQUIRE(v->n_words > 0 && v->root_has_children, ORB_ERR_INVALID, "empty vocabulary");
    if (n == 0) return ORB_OK;
    std::lock_guard<std::mutex> lock(v->mu);
    ORB_CUDA_TRY(cudaSetDevice(v->device));
    const bool host = loc == ORB_HOST;
    const size_t need = align256(4) + (host ? align256((size_t)n * 32) + 2 * align256((size_t)n * 4) + align256((size_t)n * 8) : 0);
    int rc = ensure_scratch(v, need);
    if (rc != ORB_OK) return rc;
    Carve c{v->d_scratch};
    int *d_n = c.take<int>(1);
    ORB_CUDA_TRY(cudaMemcpyAsync(d_n, &n, sizeof(int), cudaMemcpyHostToDevice, v->stream));
    if (!host) {
        rc = launch_descend(v, descriptors, d_n, 1, n, levelsup, word, node, weight);
        if (rc != ORB_OK) return rc;
        ORB_CUDA_TRY(cudaStreamSynchronize(v->stream));   // &n is a stack variable
        return ORB_OK;
    }
    uint8_t *d_desc = c.take<uint8_t>((size_t)n * 32);
    int *d_word = c.take<int>(n), *d_node = c.take<int>(n);
    double *d_w = c.take<double>(n);
    ORB_CUDA_TRY(cudaMemcpyAsync(d_desc, descriptors, (size_t)n * 32, cudaMemcpyHostToDevice, v->stream));
    rc = launch_descend(v, d_desc, d_n, 1, n, levelsup, d_word, d_node, d_w);
    if (rc != ORB_OK) return rc;
    ORB_CUDA_TRY(cudaMemcpyAsync(word, d_word, (size_t)n * 4, cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(node, d_node, (size_t)n * 4, cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA_TRY(cudaMemcpyAsync(weight, d_w, (size_t)n * 8, cudaMemcpyDeviceToHost, v->stream));
    ORB_CUDA_TRY(cudaStreamSynchronize(v->stream));
    return ORB_OK;
}

int orb_vocab_transform(orb_vocab *v, const uint8_t *descriptors, const int32_t *n, int batch, int stride_rows, int loc,
                        int levelsup, const orb_bow_out *out, int out_loc) {
    ORB_REQUIRE(v && n && out && batch >= 1 && stride_rows >= 1, ORB_ERR_INVALID, "bad argument");
    ORB_REQUIRE(out->bow_word && out->bow_value && out->n_bow && out->fv_node && out->fv_off && out->fv_feat && out->n_fv, ORB_ERR_INVALID,
                "orb_bow_out has a null array");
    ORB_REQUIRE(descriptors, ORB_ERR_INVALID, "descriptors is null");
    const bool in_host = loc == ORB_HOST, out_host = out_loc == ORB_HOST;
    int max_n = stride_rows;
    if (in_host) {
        max_n = 0;
        for (int b = 0; b < batch; b++) {
            ORB_REQUIRE(n[b] >= 0 && n[b] <= stride_rows, ORB_ERR_INVALID, "frame %d: %d features do not fit stride_rows=%d", b, n[b], stride_rows);
            max_n = std::max(max_n, n[b]);
        }
    }
    ORB_REQUIRE(out->capacity >= max_n && out->capacity >= 1, ORB_ERR_CAPACITY, "orb_bow_out.capacity %d < %d features", out->capacity, max_n);
    std::lock_guard<std::mutex> lock(v->mu);
    ORB_CUDA_TRY(cudaSetDevice(v->device));
    const size_t feats = (size_t)batch * stride_rows, cap = (size_t)out->capacity;
    const int P2 = next_pow2(stride_rows);
    const bool gsort = P2 > SMEM_SORT_MAX;
    size_t need = 2 * align256(feats * 4) + align256(feats * 8) + align256((size_t)batch * 4);
    if (in_host) need += align256(feats * 32);
    if (gsort) need += align256((size_t)batch * P2 * 8);
    if (out_host) need += 3 * align256(batch * cap * 4) + align256(batch * cap * 8) + align256(batch * (cap + 1) * 4) + 2 * align256((size_t)batch * 4);
    int rc = ensure_scratch(v, need);
    if (rc != ORB_OK) return rc;
    Carve c{v->d_scratch};
    int *d_word = c.take<int>(feats), *d_node = c.take<int>(feats);
    double *d_w = c.take<double>(feats);
    int *d_n = c.take<int>(batch);
    const uint8_t *d_desc = descriptors;
    const int *dn = n;
    if (in_host) {
        uint8_t *stage = c.take<uint8_t>(feats * 32);
        if (batch == 1) { if (n[0]) ORB_CUDA_TRY(cudaMemcpyAsync(stage, descriptors, (size_t)n[0] * 32, cudaMemcpyHostToDevice, v->stream)); }
        else ORB_CUDA_TRY(cudaMemcpyAsync(stage, descriptors, feats * 32, cudaMemcpyHostToDevice, v->stream));
        ORB_CUDA_TRY(cudaMemcpyAsync(d_n, n, (size_t)batch * 4, cudaMemcpyHostToDevice, v->stream));
        d_desc = stage; dn = d_n;
    }
    unsigned long long *d_keys = gsort ? c.take<unsigned long long>((size_t)batch * P2) : nullptr;
    AsmOut o;
    o.capacity = out->capacity;
    if (out_host) {
        o.bow_word = c.take<int>(batch * cap); o.bow_value = c.take<double>(batch * cap); o.n_bow = c.take<int>(batch);
        o.fv_node = c.take<int>(batch * cap); o.fv_off = c.take<int>(batch * (cap + 1)); o.fv_feat = c.take<int>(batch * cap);
        o.n_fv = c.take<int>(batch);
    } else {
        o.bow_word = out->bow_word; o.bow_value = out->bow_value; o.n_bow = out->n_bow;
        o.fv_node = out->fv_node; o.fv_off = out->fv_off; o.fv_feat = out->fv_feat; o.n_fv = out->n_fv;
    }
    if (v->n_words == 0 || !v->root_has_children) {   // empty(): both vectors cleared (ref :1133-1136)
        ORB_CUDA_TRY(cudaMemsetAsync(o.n_bow, 0, (size_t)batch * 4, v->stream));
        ORB_CUDA_TRY(cudaMemsetAsync(o.n_fv, 0, (size_t)batch * 4, v->stream));
        ORB_CUDA_TRY(cudaMemsetAsync(o.fv_off, 0, batch * (cap + 1) * 4, v->stream));
    } else {
        rc = launch_descend(v, d_desc, dn, batch, stride_rows, levelsup, d_word, d_node, d_w);
        if (rc != ORB_OK) return rc;
        const int tf = v->weighting == 0 || v->weighting == 1;      // TF_IDF, TF: addWeight; IDF, BINARY: addIfNotExist
        const int must = v->scoring != 5;                           // ScoringObject.h:74-89
        const int l2 = v->scoring == 1;
        const size_t smem = gsort ? 0 : (size_t)P2 * 8;
        k_bow_assemble<<<batch, ASM_THREADS, smem, v->stream>>>(d_word, d_node, d_w, dn, stride_rows, P2, d_keys, tf, must, l2, o);
        ORB_CUDA_TRY(cudaGetLastError());
        v->launches++;
    }
    if (out_host) {
        ORB_CUDA_TRY(cudaMemcpyAsync(out->bow_word, o.bow_word, batch * cap * 4, cudaMemcpyDeviceToHost, v->stream));
        ORB_CUDA_TRY(cudaMemcpyAsync(out->bow_value, o.bow_value, batch * cap * 8, cudaMemcpyDeviceToHost, v->stream));
        ORB_CUDA_TRY(cudaMemcpyAsync(out->n_bow, o.n_bow, (size_t)batch * 4, cudaMemcpyDeviceToHost, v->stream));
        ORB_CUDA_TRY(cudaMemcpyAsync(out->fv_node, o.fv_node, batch * cap * 4, cudaMemcpyDeviceToHost, v->stream));
        ORB_CUDA_TRY(cudaMemcpyAsync(out->fv_off, o.fv_off, batch * (cap + 1) * 4, cudaMemcpyDeviceToHost, v->stream));
        ORB_CUDA_TRY(cudaMemcpyAsync(out->fv_feat, o.fv_feat, batch * cap * 4, cudaMemcpyDeviceToHost, v->stream));
        ORB_CUDA_TRY(cudaMemcpyAsync(out->n_fv, o.n_fv, (size_t)batch * 4, cudaMemcpyDeviceToHost, v->stream));
    }
    if (in_host || out_host) ORB_CUDA_TRY(cudaStreamSynchronize(v->stream));
    return ORB_OK;
}

}  // extern "C"
