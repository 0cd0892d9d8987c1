// TEST INFRASTRUCTURE -- not part of the product path.
// C ABI around the reference's OWN Marker::searchMatches + Marker::hamdist2 (ORB_SLAM2/src/Marker.cc:314-358).  Marker.cc as a
// file needs OpenCV's ORB / findHomography / cvFitPlane, which are not on this image, so oracle/Makefile.ref lifts exactly
// those two member functions out of the reference's file at build time (sed line range -> _ref/marker_search.inc, deleted
// after the build; no reference text is kept in this repository) and compiles them here against the reference's
// include/Marker.h and oracle/cvstub.  Used to pin oracle/orbref_mappoint.cpp's orbref_search_matches_ratio and, through
// it, liborbgpu's orb_search_matches_ratio.
#include <climits>
#include <cstdio>
#include <cstring>
#include <vector>

#include "opencv2/opencv.hpp"
#include "Marker.h"   // the reference's header: class Marker with searchMatches / hamdist2 declared at :26-28

// Marker.cc:21-22 (file-scope statistics the function updates; outside the lifted range)
unsigned int maxD = 0;
unsigned int minD = 100;

// the reference's constructor lives in the part of Marker.cc that cannot be built here; searchMatches touches no member
Marker::Marker() : mThreshold(2.0f), mSize(0) {}

#include "marker_search.inc"

extern "C" int refmk_search_matches(const unsigned char *train, int n_train, const unsigned char *query, int n_query,
                                    int *query_idx, int *train_idx, float *distance, int capacity, int *n_matches) {
    std::vector<cv::KeyPoint> k1((size_t)n_train), k2((size_t)n_query);
    cv::Mat d1(n_train, 32, CV_8UC1, (void *)train), d2(n_query, 32, CV_8UC1, (void *)query);
    std::vector<cv::DMatch> m;
    Marker mk;
    FILE *keep = stdout;   // the function prints its running min / max distance (Marker.cc:345-346)
    FILE *nul = fopen("/dev/null", "w");
    if (nul) stdout = nul;
    mk.searchMatches(k1, d1, k2, d2, m);
    stdout = keep;
    if (nul) fclose(nul);
    *n_matches = (int)m.size();
    if ((int)m.size() > capacity) return -4;
    for (size_t i = 0; i < m.size(); i++) { query_idx[i] = m[i].queryIdx; train_idx[i] = m[i].trainIdx; distance[i] = m[i].distance; }
    return 0;
}
