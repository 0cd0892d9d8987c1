// ORACLE -- TEST INFRASTRUCTURE ONLY.  Never linked into or called by the product path.
//
// CPU restatement of MapPoint::ComputeDistinctiveDescriptors (ORB_SLAM2/src/MapPoint.cc:224-285) on flat arrays.  Pinned
// against the reference's own MapPoint.cc compiled unmodified (oracle/_ref/libmappoint_ref.so, tests/test_mappoint_cpu.py).
#include "orbref.h"

#include <algorithm>
#include <climits>
#include <cstring>
#include <vector>

static int dist256(const uint8_t *a, const uint8_t *b) {   // ORBmatcher::DescriptorDistance, src/ORBmatcher.cc:1650-1666
    int32_t pa[8], pb[8];
    std::memcpy(pa, a, 32); std::memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        unsigned int v = pa[i] ^ pb[i];
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}

extern "C" int orbref_distinctive_descriptors(const uint8_t *desc, const int32_t *off, int npoints, int32_t *best_row,
                                              int32_t *best_median) {
    for (int p = 0; p < npoints; p++) {
        const size_t N = (size_t)(off[p + 1] - off[p]);
        if (N == 0) { best_row[p] = -1; best_median[p] = 0; continue; }   // :251-252 returns early
        const uint8_t *rows = desc + (size_t)32 * off[p];
        std::vector<float> D(N * N);                                       // float Distances[N][N], :257
        for (size_t i = 0; i < N; i++) {
            D[i * N + i] = 0;
            for (size_t j = i + 1; j < N; j++) {
                const int dij = dist256(rows + 32 * i, rows + 32 * j);
                D[i * N + j] = (float)dij;
                D[j * N + i] = (float)dij;
            }
        }
        int BestMedian = INT_MAX, BestIdx = 0;
        for (size_t i = 0; i < N; i++) {
            std::vector<int> v(D.begin() + i * N, D.begin() + (i + 1) * N);
            std::sort(v.begin(), v.end());
            const int median = v[(size_t)(0.5 * (N - 1))];                 // :270
            if (median < BestMedian) { BestMedian = median; BestIdx = (int)i; }
        }
        best_row[p] = BestIdx;
        best_median[p] = BestMedian;
    }
    return ORB_OK;
}

// ref: Marker::searchMatches, ORB_SLAM2/src/Marker.cc:314-349 (= naive_nn_search2, AR-1.3/src/ORBMatcher.cpp:70-102), on flat
// arrays.  hamdist2 (Marker.cc:351-359) is cv::HammingLUT over 32 bytes = the bit count of the XOR.  PARITY UNPINNED for this
// function: Marker.cc needs OpenCV's legacy / features2d / calib3d modules and cannot be compiled here; this is a line-by-line
// restatement.
extern "C" int orbref_search_matches_ratio(const uint8_t *train, int n_train, const uint8_t *query, int n_query, double ratio,
                                           int max_dist, int32_t *train_idx, int32_t *dist) {
    for (int i = 0; i < n_query; i++) {
        unsigned int min_dist = INT_MAX, sec_dist = INT_MAX;
        int min_idx = -1, sec_idx = -1;
        for (int j = 0; j < n_train; j++) {
            const unsigned int d = (unsigned int)dist256(query + 32 * (size_t)i, train + 32 * (size_t)j);
            if (d < min_dist) { sec_dist = min_dist; sec_idx = min_idx; min_dist = d; min_idx = j; }
            else if (d < sec_dist) { sec_dist = d; sec_idx = j; }
        }
        (void)sec_idx;
        const bool ok = min_idx >= 0 && min_dist <= (unsigned int)(sec_dist * ratio) && min_dist <= (unsigned int)max_dist;
        train_idx[i] = ok ? min_idx : -1;
        dist[i] = (int32_t)min_dist;
    }
    return ORB_OK;
}
