// ORACLE -- TEST INFRASTRUCTURE ONLY.  Internal types shared by the oracle's translation units.
#pragma once
#include "orbref.h"

#include <cstdint>
#include <list>
#include <vector>

namespace orbref {
const int EDGE_THRESHOLD = 19;    // ref ORBextractor.cc:71

struct Cand {       // a cv::KeyPoint as produced by cv::FAST and shifted by ref :786-793
    float x, y;     // window-relative
    float response;
    float angle;
};

struct Level {
    int w = 0, h = 0, pitch = 0;          // interior size; pitch of the padded plane
    std::vector<uint8_t> plane;           // (w+38) x (h+38)
    std::vector<Cand> cands;              // vToDistributeKeys
    std::vector<Cand> kept;               // after DistributeOctTree, + minBorder, + angle
    std::vector<uint8_t> blurred;         // w x h, only if kept non-empty
    uint8_t *interior() { return plane.data() + (size_t)EDGE_THRESHOLD * pitch + EDGE_THRESHOLD; }
};

struct Node {   // ref ExtractorNode, include/ORBextractor.h:30-43
    std::vector<int> keys;        // indices into the candidate array, in vKeys order
    int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
    std::list<Node>::iterator lit;
    bool no_more = false;
    long seq = 0;                 // creation sequence number (oracle tie key)
};


}  // namespace orbref

struct orbref_extractor {
    orb_params prm;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> per_level;
    std::vector<int> umax;
    std::vector<orbref::Level> lv;
    int trig_mode = 0;            // 0: restated glibc sincosf; 1: host libm cosf/sinf
    long node_seq = 0;

    void divide(const orbref::Node &n, const std::vector<orbref::Cand> &c, orbref::Node out[4]);
    std::vector<orbref::Cand> distribute(const std::vector<orbref::Cand> &c, int minX, int maxX, int minY, int maxY, int N);
};
