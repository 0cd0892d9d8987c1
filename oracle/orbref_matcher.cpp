// ORACLE -- TEST INFRASTRUCTURE ONLY (see orbref_cv.h).  CPU restatement of the reference's
// Hamming matchers (ORB_SLAM2/src/ORBmatcher.cc) on flat arrays.
#include "orbref.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

// ref ORBmatcher::DescriptorDistance, ORBmatcher.cc:1650-1666
static inline int descriptor_distance(const uint8_t *a, const uint8_t *b) {
    const int32_t *pa = (const int32_t *)a, *pb = (const int32_t *)b;
    int dist = 0;
    for (int i = 0; i < 8; i++, pa++, pb++) {
        unsigned int v = *pa ^ *pb;
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}
// same distance with the hardware popcount (used only by the timed CPU baseline)
static inline int descriptor_distance_popcnt(const uint8_t *a, const uint8_t *b) {
    const uint64_t *pa = (const uint64_t *)a, *pb = (const uint64_t *)b;
    return __builtin_popcountll(pa[0] ^ pb[0]) + __builtin_popcountll(pa[1] ^ pb[1]) + __builtin_popcountll(pa[2] ^ pb[2]) +
           __builtin_popcountll(pa[3] ^ pb[3]);
}

extern "C" {

// Brute-force best / second-best scan with the reference's update rule (ORBmatcher.cc:216-225):
// strict '<' on both, so the first row seen wins ties.  fast=1 uses the hardware popcount and
// `threads` workers over query ranges (CPU baseline); fast=0 is the literal SWAR loop.
int orbref_hamming_top2(const uint8_t *queries, int nq, const uint8_t *db, int64_t ndb, int64_t idx_base,
                        int32_t *best_dist, int64_t *best_idx, int32_t *second_dist, int64_t *second_idx, int fast,
                        int threads) {
    auto work = [&](int q0, int q1) {
        for (int q = q0; q < q1; q++) {
            const uint8_t *dq = queries + (size_t)q * 32;
            int d1 = 256, d2 = 256;
            int64_t i1 = -1, i2 = -1;
            for (int64_t r = 0; r < ndb; r++) {
                const int d = fast ? descriptor_distance_popcnt(dq, db + (size_t)r * 32) : descriptor_distance(dq, db + (size_t)r * 32);
                if (d < d1) { d2 = d1; i2 = i1; d1 = d; i1 = r; }
                else if (d < d2) { d2 = d; i2 = r; }
            }
            best_dist[q] = d1; second_dist[q] = d2;
            best_idx[q] = i1 < 0 ? -1 : i1 + idx_base;
            second_idx[q] = i2 < 0 ? -1 : i2 + idx_base;
        }
    };
    if (threads <= 1) { work(0, nq); return ORB_OK; }
    std::vector<std::thread> th;
    for (int t = 0; t < threads; t++) th.emplace_back(work, (int)((int64_t)nq * t / threads), (int)((int64_t)nq * (t + 1) / threads));
    for (auto &t : th) t.join();
    return ORB_OK;
}

// ref ORBmatcher::ComputeThreeMaxima, ORBmatcher.cc:1604-1645, on bin counts
void orbref_three_maxima(const int *sizes, int L, int *ind1, int *ind2, int *ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    *ind1 = *ind2 = *ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = sizes[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; *ind3 = *ind2; *ind2 = *ind1; *ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; *ind3 = *ind2; *ind2 = i; }
        else if (s > max3) { max3 = s; *ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { *ind2 = -1; *ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { *ind3 = -1; }
}

// rotation bin of ORBmatcher.cc:238-243 (factor = 1/HISTO_LENGTH, the reference's quirk)
int orbref_rot_bin(float angle_a, float angle_b) {
    const int HISTO_LENGTH = 30;
    const float factor = 1.0f / HISTO_LENGTH;
    float rot = angle_a - angle_b;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * factor);
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

// keep[i] = pair i falls in one of the three dominant rotation bins
int orbref_rot_hist_filter(const float *angle_a, const float *angle_b, int n, uint8_t *keep) {
    int sizes[30] = {0};
    std::vector<int> bin(n);
    for (int i = 0; i < n; i++) { bin[i] = orbref_rot_bin(angle_a[i], angle_b[i]); sizes[bin[i]]++; }
    int i1, i2, i3;
    orbref_three_maxima(sizes, 30, &i1, &i2, &i3);
    for (int i = 0; i < n; i++) keep[i] = (bin[i] == i1 || bin[i] == i2 || bin[i] == i3) ? 1 : 0;
    return ORB_OK;
}

}  // extern "C"
