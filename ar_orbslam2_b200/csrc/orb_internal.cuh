// liborbgpu internals shared by the .cu translation units: level geometry, the extractor handle.
#pragma once
#include <cuda.h>
#include "orb_common.cuh"

#include <vector>

namespace orbgpu {

constexpr int EDGE = ORB_EDGE_THRESHOLD;   // 19
constexpr int XOFF = 32;                   // interior column 0 sits at byte 32 of a plane row
constexpr int MAX_INI = 64;                // max initial quadtree nodes (image aspect ratio)
constexpr int FAST_TILE_MAX = 72;          // cell (<60) + 6 halo, rounded up
constexpr int FAST_THREADS = 64;           // measured 32..256 on B200: 32-128 within 5%, 256 is 1.8x slower (per-warp fixed costs)
constexpr int OCT_THREADS = 256;           // measured 128 / 256 / 512 / 1024: 0.33 / 0.23 / 0.25 / 0.50 ms per 288 frames
constexpr int DESC_WARPS = 8;

enum { ERRF_CAND_OVERFLOW = 1, ERRF_NODE_OVERFLOW = 2, ERRF_OUT_OVERFLOW = 4 };

struct ResizeTap {   // one destination column/row of cv::resize INTER_LINEAR
    int16_t s0, s1, a0, a1;
};

struct LevelGeom {
    int w, h, pitch;             // interior size, bytes per padded row
    long long plane_off;         // byte offset of the padded plane inside a frame's pyramid block
    long long interior_off;      // byte offset of interior pixel (0,0)
    int tap_x_off, tap_y_off;    // offsets into the ResizeTap pool (level >= 1)
    int wtab_off, fast_resize;   // per-output-word resize records (fast path) and whether the level may use them
    int rtab_off;                // per-plane-row records of the fast path (uint4 index into the same pool)
    // FAST cell grid (ref :742-773)
    int maxBX, maxBY, wCell, hCell, ncx, ncy, cell_begin;
    // quadtree (ref :530-549)
    int quota, nIni;
    float hX;
    int ini_x[MAX_INI + 1];
    int win_w, win_h;            // maxBorder - minBorder
    int cand_off, cand_cap;      // into the per-frame candidate pool (elements)
    int kept_off, kept_cap;      // into the per-frame kept pool
    float scale;                 // mvScaleFactor[level]
    float inv_scale;             // mvInvScaleFactor[level]
    float size;                  // (float)(int)(31*scale)
};

struct Geom {
    int nlevels, width, height;
    int total_cells;
    int cand_frame, kept_frame;          // pool elements per frame
    long long pyr_frame;                 // pyramid bytes per frame
    int max_nodes;                       // smem sizing for k_octree
    int ini_th, min_th, blur_taps;
    LevelGeom lv[ORB_MAX_LEVELS];
};

struct Kept {   // a keypoint retained by the quadtree, level coordinates
    int16_t x, y;
    int32_t response;
};

}  // namespace orbgpu

constexpr size_t ORB_BOUNCE_BYTES = (size_t)1 << 20;
constexpr size_t ORB_HSRC_BYTES = ((size_t)4 << 20) + ((size_t)1 << 19);   // frames of a small call (<= 4 MB of pixels) incl. row pitch

struct orb_extractor {
    orb_params prm;
    int device = 0;
    cudaStream_t stream = nullptr;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> per_level;
    int umax[16];
    orbgpu::Geom geom;                 // host copy
    orbgpu::Geom *d_geom = nullptr;
    orbgpu::ResizeTap *d_taps = nullptr;
    size_t taps_cap = 0;
    uint4 *d_wtab = nullptr;           // per-output-word resize records
    size_t wtab_cap = 0;
    uint32_t *d_cells = nullptr;       // per FAST cell: level | ci << 4 | cj << 14
    size_t cells_cap = 0;
    uint8_t *d_src = nullptr;  size_t src_bytes = 0;
    uint8_t *d_pyr = nullptr;  size_t pyr_bytes = 0;
    uint8_t *d_blur = nullptr;         // 7x7 Gaussian of every level (same plane layout as d_pyr)
    uint32_t *d_cand = nullptr; uint16_t *d_node_of = nullptr; size_t cand_elems = 0;
    uint32_t *d_counts = nullptr;      // [B][nlevels] candidates, then [B][nlevels] kept
    bool fast_rows = true;             // large batches: k_fast_rows on levels of at least fast_rows_min_px computed pixels
    long long fast_rows_min_px = 100000;
    CUtensorMap tmap[ORB_MAX_LEVELS];  // per level: the pyramid planes of all frame slots as a 3-D tensor (k_fast_rows' tile loads)
    bool tmap_ok[ORB_MAX_LEVELS] = {};
    orbgpu::Kept *d_kept = nullptr; size_t kept_elems = 0;
    orb_keypoint *d_out_kp = nullptr; uint8_t *d_out_desc = nullptr; int32_t *d_out_n = nullptr;
    size_t out_cap_elems = 0;
    int *d_err = nullptr;
    int *h_pinned = nullptr;           // [0]=err flags, [1..] counts
    uint8_t *h_src = nullptr;          // pinned, device-mapped input buffer of small host-input calls (k_pyr_level0 reads it across PCIe)
    uint8_t *h_bounce = nullptr;       // pinned landing buffer for the results of small host-output calls
    float *d_stereo_out = nullptr;     // orb_stereo_frame: uRight[cap] | depth[cap] | n_matched
    size_t stereo_out_cap = 0;
    int stereo_last_capacity = 0;      // capacity of the last orb_stereo_frame call (depth starts that many floats in)
    int last_batch = 0;
    // where the last extraction left its results (device), for orb_frame_from_extractor: the caller's device buffers or the
    // handle's own staging; counts on the host when the call was a host-output one
    const orb_keypoint *last_kp = nullptr; const uint8_t *last_desc = nullptr; const int32_t *last_n = nullptr;
    int last_capacity = 0;
    std::vector<int> last_counts; bool last_counts_valid = false;
    int64_t launches = 0;
    size_t oct_smem = 0;
    cudaEvent_t ev_pair = nullptr, ev_pair_done = nullptr;   // orb_stereo_match_batch with two handles: ordering between their streams
    void *stereo_scratch = nullptr;    // device scratch of orb_stereo_match_batch
    size_t stereo_scratch_bytes = 0;
    // optional per-stage timing (CUDA events on the handle's stream)
    bool profiling = false;
    cudaEvent_t ev[ORB_N_STAGES + 1] = {};
    double stage_ms[ORB_N_STAGES] = {};
    int64_t stage_calls = 0;
    bool ev_pending = false;
    // side streams: FAST and the blur of a level start as soon as that level of the pyramid exists and run beside the
    // remaining resize launches; both are joined back into `stream` before the quadtree / the descriptors
    cudaStream_t s_fast = nullptr, s_blur = nullptr;
    cudaEvent_t ev_lvl[ORB_MAX_LEVELS] = {}, ev_fast = nullptr, ev_blur = nullptr;
    bool overlap = true;               // off while per-stage profiling is on (stage times need a serial schedule)
    // small calls (latency mode): the quadtree of level l runs on its own stream as soon as that level's FAST cells are done
    cudaStream_t s_oct[ORB_MAX_LEVELS] = {}, s_bl[ORB_MAX_LEVELS] = {};   // small calls: FAST + quadtree / blur of level l
    bool host_trace = false; double trace_us[4] = {}; int64_t trace_calls = 0;   // ORBGPU_HOST_TRACE=1: where a blocking call's host time goes
    int pyr_small_rows = 2;            // plane rows per thread of the resize in small calls (ORBGPU_PYR_SMALL_ROWS: 2 / 4 / 8)
    int oct_small_threads = 512;       // CTA size of the quadtree in small calls (ORBGPU_OCT_SMALL_THREADS: 256 / 512 / 1024)
    cudaEvent_t ev_chunk[2] = {};      // describe of a sub-batch is done (large batches run as L2-sized sub-batches)
    size_t chunk_bytes = 0;             // pyramid + blurred planes of one sub-batch (ORBGPU_CHUNK_MB); 0 = whole batch at once
    cudaEvent_t ev_fl[ORB_MAX_LEVELS] = {}, ev_ol[ORB_MAX_LEVELS] = {};
    // CUDA graph of one extraction (two memsets + the 26 launches with their fork/join on the side streams): captured the
    // second time a call repeats the previous call's arguments, replayed from then on; any other call runs the plain
    // launches.  One graph launch instead of 28 enqueues + 20 event operations is what a single-frame call waits for.
    struct GraphKey {
        const void *images; size_t stride; int pitch, batch, width, height, capacity; void *kp, *desc, *n; int export_host;
        bool operator==(const GraphKey &o) const {
            return images == o.images && stride == o.stride && pitch == o.pitch && batch == o.batch && width == o.width &&
                   height == o.height && capacity == o.capacity && kp == o.kp && desc == o.desc && n == o.n && export_host == o.export_host;
        }
    };
    GraphKey gkey = {}, glast = {};
    cudaGraphExec_t gexec = nullptr;
    int g_nodes = 0;                   // kernel launches inside the captured graph
    bool use_graph = true;             // ORBGPU_NO_GRAPH=1 switches it off
    int64_t graph_replays = 0;
};

// Frame::UndistortKeyPoints on device arrays, asynchronous on `stream` (orb_undistort.cu)
int orbgpu_undistort_keypoints_async(const orb_keypoint *keys, int n, float fx, float fy, float cx, float cy, const float *dist_coef,
                                     int ndist, orb_keypoint *keys_un, cudaStream_t stream);

// grows the handle's device staging for `batch` x `capacity` keypoints + descriptors (drops a captured graph that holds the old pointers)
int orbgpu_ensure_out_staging(orb_extractor *ex, int batch, int capacity);
/* Results of a small host-output call straight into the handle's pinned landing buffers by ONE kernel (the valid entries of every
 * frame only): h_pinned = {err flags, n[batch]}, h_bounce = key points | descriptors [| uRight | depth | n_matched], the layout the
 * host side unpacks.  u_right == nullptr: extraction only. */
void orbgpu_export_results(orb_extractor *ex, const orb_keypoint *kp, const uint8_t *desc, const int32_t *n, int batch, int capacity,
                           const float *u_right, const float *depth, const int32_t *n_matched, cudaStream_t st);
