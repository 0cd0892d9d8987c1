// Bodies of MarkerORB.h over the C ABI (include/orbgpu.h).  No arithmetic here.
#include "MarkerORB.h"

#include <algorithm>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <tuple>

#include "../../include/orbgpu.h"

namespace orbgpu_shim {

namespace {
// `ORB orb;` is a local object in Marker::setTargetImage / Match, built once per image: the device handle (14 buffers, two
// streams) outlives it in a per-thread cache keyed on the parameters and grown when a larger image arrives.
struct Cached {
    orb_cvorb *h = nullptr;
    int max_w = 0, max_h = 0;
};
typedef std::tuple<int, float, int, int, int> Key;
struct Cache {
    std::map<Key, Cached> m;
};
// never deleted: no CUDA call may run from a thread_local destructor at thread / process exit (see ORBmatcher_gpu.cc);
// ReleaseThreadHandles() frees the calling thread's handles explicitly
Cache &thread_cache() {
    thread_local Cache *c = new Cache;
    return *c;
}

orb_cvorb *handle_for(const Key &k, int w, int h) {
    Cached &c = thread_cache().m[k];
    if (c.h && w <= c.max_w && h <= c.max_h) return c.h;
    if (c.h) orb_cvorb_destroy(c.h);
    c.h = nullptr;
    orb_cvorb_params p;
    p.nfeatures = std::get<0>(k); p.scale_factor = std::get<1>(k); p.nlevels = std::get<2>(k); p.edge_threshold = std::get<3>(k);
    p.score_type = std::get<4>(k); p.fast_threshold = 20;   // OpenCV 2.4's ORB has no fastThreshold argument: cv::FAST at 20
    // sized for the larger of VGA / 1080p / what arrives, so that a marker image and camera frames share one handle
    c.max_w = p.max_width = std::max(std::max(w, c.max_w), w > 640 || h > 480 ? 1920 : 640);
    c.max_h = p.max_height = std::max(std::max(h, c.max_h), w > 640 || h > 480 ? 1080 : 480);
    p.device = -1;
    if (orb_cvorb_create(&p, &c.h) != ORB_OK) throw std::runtime_error(std::string("orbgpu_shim::ORB: ") + orb_last_error());
    return c.h;
}
}  // namespace

void ReleaseThreadHandles() {
    Cache &C = thread_cache();
    for (auto &kv : C.m) orb_cvorb_destroy(kv.second.h);
    C.m.clear();
}

ORB::ORB(int nf, float sf, int nl, int et, int firstLevel, int WTA_K, int st, int patchSize)
    : nfeatures(nf), nlevels(nl), edgeThreshold(et), scoreType(st), scaleFactor(sf) {
    if (firstLevel != 0 || WTA_K != 2 || patchSize != 31)
        throw std::invalid_argument("orbgpu_shim::ORB: firstLevel / WTA_K / patchSize are fixed at 0 / 2 / 31 (what the marker path uses)");
}

void ORB::operator()(cv::InputArray _image, cv::InputArray _mask, std::vector<cv::KeyPoint> &keypoints, cv::OutputArray _descriptors,
                     bool useProvidedKeypoints) const {
    if (useProvidedKeypoints || !_mask.empty()) throw std::invalid_argument("orbgpu_shim::ORB: masks / provided key points are not supported");
    keypoints.clear();
    if (_image.empty()) { _descriptors.release(); return; }
    cv::Mat image = _image.getMat();
    orb_cvorb *h = handle_for(Key(nfeatures, scaleFactor, nlevels, edgeThreshold, scoreType), image.cols, image.rows);
    int cap = orb_cvorb_capacity(h), n = 0;
    static_assert(sizeof(cv::KeyPoint) == sizeof(orb_keypoint), "cv::KeyPoint layout");
    std::vector<cv::KeyPoint> kps(cap);
    std::vector<unsigned char> desc((size_t)cap * 32);
    const int rc = orb_cvorb_detect_and_compute(h, image.data, image.cols, image.rows, (int)image.step, ORB_HOST, (orb_keypoint *)kps.data(),
                                                desc.data(), cap, &n);
    if (rc != ORB_OK) throw std::runtime_error(std::string("orbgpu_shim::ORB: ") + orb_last_error());
    if (n == 0) { _descriptors.release(); return; }
    kps.resize(n);
    keypoints.swap(kps);
    _descriptors.create(n, 32, CV_8U);
    cv::Mat d = _descriptors.getMat();
    for (int i = 0; i < n; i++) std::memcpy(d.ptr(i), &desc[(size_t)i * 32], 32);
}

void BruteForceMatchHamming(const cv::Mat &q, const cv::Mat &t, std::vector<cv::DMatch> &matches) {
    matches.clear();
    if (q.rows == 0 || t.rows == 0) return;
    cv::Mat qc = q.isContinuous() ? q : q.clone(), tc = t.isContinuous() ? t : t.clone();
    std::vector<int32_t> bd(q.rows), sd(q.rows);
    std::vector<int64_t> bi(q.rows), si(q.rows);
    if (orb_hamming_top2(qc.data, q.rows, tc.data, t.rows, 0, bd.data(), bi.data(), sd.data(), si.data(), ORB_HOST) != ORB_OK)
        throw std::runtime_error(std::string("orbgpu_shim::BruteForceMatchHamming: ") + orb_last_error());
    matches.reserve(q.rows);
    for (int i = 0; i < q.rows; i++) matches.push_back(cv::DMatch(i, (int)bi[i], 0, (float)bd[i]));
}

bool SearchMatches(const cv::Mat &descp1, const cv::Mat &descp2, std::vector<cv::DMatch> &matches) {
    cv::Mat tc = descp1.isContinuous() ? descp1 : descp1.clone(), qc = descp2.isContinuous() ? descp2 : descp2.clone();
    std::vector<int32_t> idx(descp2.rows), dist(descp2.rows);
    if (orb_search_matches_ratio(tc.data, descp1.rows, qc.data, descp2.rows, 0.8, 50, ORB_HOST, idx.data(), dist.data()) != ORB_OK)
        throw std::runtime_error(std::string("orbgpu_shim::SearchMatches: ") + orb_last_error());
    for (int i = 0; i < descp2.rows; i++)
        if (idx[i] >= 0) matches.push_back(cv::DMatch(i, idx[i], 0, (float)dist[i]));
    return true;
}

void NaiveNNSearch(const cv::Mat &descp1, const cv::Mat &descp2, std::vector<cv::DMatch> &matches) {
    std::vector<cv::DMatch> all;
    BruteForceMatchHamming(descp2, descp1, all);   // query = frame (keys2), train = marker (keys1)
    for (size_t i = 0; i < all.size(); i++)
        if (all[i].trainIdx >= 0 && all[i].distance <= 50.f) matches.push_back(all[i]);
}

void NaiveNNSearch2(const cv::Mat &descp1, const cv::Mat &descp2, std::vector<cv::DMatch> &matches, unsigned int &minD, unsigned int &maxD) {
    SearchMatches(descp1, descp2, matches);
    cv::Mat tc = descp1.isContinuous() ? descp1 : descp1.clone(), qc = descp2.isContinuous() ? descp2 : descp2.clone();
    int32_t mm[2];
    if (orb_hamming_minmax(tc.data, descp1.rows, qc.data, descp2.rows, ORB_HOST, mm) != ORB_OK)
        throw std::runtime_error(std::string("orbgpu_shim::NaiveNNSearch2: ") + orb_last_error());
    if (mm[1] >= 0) {   // `if (dist <= minD) minD = dist; if (dist > maxD) maxD = dist;` over every pair
        if ((unsigned int)mm[0] <= minD) minD = (unsigned int)mm[0];
        if ((unsigned int)mm[1] > maxD) maxD = (unsigned int)mm[1];
    }
}

}  // namespace orbgpu_shim
