// c3_bench -- per-call latency of the tracking matchers through the C ABI, from host arrays, as a C++ caller (the reference's
// Tracking thread) sees it.  BASELINE.json config 3: 752x480, frame B = frame A shifted by (+7,+4).
//
//   c3_bench <dir> <W> <H> <reps>
// <dir> holds the flat inputs bench.py wrote (little-endian raw arrays): a.raw / b.raw (the two frames), ka.bin / da.bin /
// kb.bin / db.bin (key points and descriptors of A and B), scale.bin, and for the two projection matchers last_*.bin
// (orb_last_points fields + claimed) and pts_*.bin (orb_track_points fields + claimed).  It writes the results of each call
// back into <dir> (init_m12.bin, init_prev.bin, last_assigned.bin, pts_assigned.bin) for bench.py to compare with the
// reference's own ORBmatcher.cc, and prints one JSON object with the timings.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/orbgpu.h"

static std::string g_dir;

template <class T> static std::vector<T> load(const char *name) {
    const std::string p = g_dir + "/" + name;
    FILE *f = fopen(p.c_str(), "rb");
    if (!f) { fprintf(stderr, "c3_bench: cannot open %s\n", p.c_str()); exit(2); }
    fseek(f, 0, SEEK_END);
    const long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<T> v((size_t)n / sizeof(T));
    if (n && fread(v.data(), 1, (size_t)n, f) != (size_t)n) { fprintf(stderr, "c3_bench: short read %s\n", p.c_str()); exit(2); }
    fclose(f);
    return v;
}
template <class T> static void save(const char *name, const std::vector<T> &v) {
    const std::string p = g_dir + "/" + name;
    FILE *f = fopen(p.c_str(), "wb");
    fwrite(v.data(), sizeof(T), v.size(), f);
    fclose(f);
}
static double now_us() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static void chk(int rc, const char *what) {
    if (rc != ORB_OK) { fprintf(stderr, "c3_bench: %s failed (%d): %s\n", what, rc, orb_last_error()); exit(3); }
}
struct Stat { double med, mn; };
static Stat stat(std::vector<double> v) {
    std::sort(v.begin(), v.end());
    return Stat{v[v.size() / 2], v[0]};
}

int main(int argc, char **argv) {
    if (argc < 5) { fprintf(stderr, "usage: c3_bench <dir> <W> <H> <reps>\n"); return 2; }
    g_dir = argv[1];
    const int W = atoi(argv[2]), H = atoi(argv[3]), reps = atoi(argv[4]), warm = 5;
    auto imgA = load<uint8_t>("a.raw"), imgB = load<uint8_t>("b.raw");
    auto ka = load<orb_keypoint>("ka.bin"), kb = load<orb_keypoint>("kb.bin");
    auto da = load<uint8_t>("da.bin"), db = load<uint8_t>("db.bin");
    auto scale = load<float>("scale.bin");
    const int na = (int)ka.size(), nb = (int)kb.size();

    // ---- frame index: from host arrays (what the C++ shim caches per Frame) and straight from the extractor's device results
    orb_frame_desc dA = {na, ka.data(), da.data(), nullptr, 0.f, (float)W, 0.f, (float)H, (int)scale.size(), scale.data(), 0};
    orb_frame_desc dB = {nb, kb.data(), db.data(), nullptr, 0.f, (float)W, 0.f, (float)H, (int)scale.size(), scale.data(), 0};
    orb_frame *fa = nullptr, *fb = nullptr;
    std::vector<double> t_create;
    for (int i = 0; i < warm + reps; i++) {
        orb_frame *t = nullptr;
        const double t0 = now_us();
        chk(orb_frame_create(&dB, &t), "orb_frame_create");
        const double t1 = now_us();
        orb_frame_destroy(t);
        if (i >= warm) t_create.push_back(t1 - t0);
    }
    chk(orb_frame_create(&dA, &fa), "orb_frame_create");
    chk(orb_frame_create(&dB, &fb), "orb_frame_create");

    orb_params prm;
    memset(&prm, 0, sizeof(prm));
    prm.nfeatures = 2000; prm.scale_factor = 1.2f; prm.nlevels = 8; prm.ini_th_fast = 20; prm.min_th_fast = 7;
    prm.max_width = W; prm.max_height = H; prm.max_batch = 1; prm.device = -1;
    orb_extractor *ex = nullptr;
    chk(orb_extractor_create(&prm, &ex), "orb_extractor_create");
    const int cap = 2000 + 4 * 8 + 256;
    std::vector<orb_keypoint> kk(cap);
    std::vector<uint8_t> dd((size_t)cap * 32);
    orb_frame_source src;
    memset(&src, 0, sizeof(src));
    src.min_x = 0; src.max_x = (float)W; src.min_y = 0; src.max_y = (float)H;
    std::vector<double> t_from, t_extract;
    int n_ex = 0;
    orb_frame *fb_dev = nullptr;
    for (int i = 0; i < warm + reps; i++) {
        const double t0 = now_us();
        chk(orb_extract(ex, imgB.data(), W, H, W, ORB_HOST, kk.data(), dd.data(), cap, &n_ex), "orb_extract");
        const double t1 = now_us();
        orb_frame *t = nullptr;
        chk(orb_frame_from_extractor(ex, 0, &src, &t), "orb_frame_from_extractor");
        chk(orb_frame_sync(t), "orb_frame_sync");
        const double t2 = now_us();
        if (i >= warm) { t_extract.push_back(t1 - t0); t_from.push_back(t2 - t1); }
        if (i + 1 == warm + reps) fb_dev = t; else orb_frame_destroy(t);
    }
    const bool same_extract = n_ex == nb && memcmp(kk.data(), kb.data(), (size_t)nb * sizeof(orb_keypoint)) == 0 &&
                              memcmp(dd.data(), db.data(), (size_t)nb * 32) == 0;

    // ---- SearchForInitialization(F1 = A, F2 = B, vbPrevMatched = A's key points, windowSize 100), ref ORBmatcher.cc:405-523
    std::vector<float> prev0((size_t)na * 2), prev;
    for (int i = 0; i < na; i++) { prev0[2 * i] = ka[i].x; prev0[2 * i + 1] = ka[i].y; }
    std::vector<int32_t> m12(na);
    std::vector<double> t_init, t_init_dev;
    int nm_init = 0, nm_init_dev = 0, sw_init = 0;
    int64_t cand_init = 0;
    for (int i = 0; i < warm + reps; i++) {
        prev = prev0;
        const double t0 = now_us();
        chk(orb_search_for_initialization(fa, fb, prev.data(), m12.data(), 100, 0.9f, 1, &nm_init), "orb_search_for_initialization");
        const double t1 = now_us();
        if (i >= warm) t_init.push_back(t1 - t0);
    }
    chk(orb_frame_last_stats(fb, &cand_init, &sw_init), "orb_frame_last_stats");
    save("init_m12.bin", m12);
    save("init_prev.bin", prev);
    {   // the same call against the index built on the device from the extractor's results
        std::vector<int32_t> m12d(na);
        std::vector<float> prevd;
        for (int i = 0; i < warm + reps; i++) {
            prevd = prev0;
            const double t0 = now_us();
            chk(orb_search_for_initialization(fa, fb_dev, prevd.data(), m12d.data(), 100, 0.9f, 1, &nm_init_dev), "orb_search_for_initialization");
            const double t1 = now_us();
            if (i >= warm) t_init_dev.push_back(t1 - t0);
        }
        if (m12d != m12 || prevd != prev) nm_init_dev = -1;
    }

    // ---- SearchByProjection(CurrentFrame, LastFrame, th = 15, bMono), ref :1331-1473.  Cur = A's features seen from the moved pose.
    auto l_valid = load<uint8_t>("last_valid.bin"), l_obs = load<uint8_t>("last_obs.bin"), l_claimed = load<uint8_t>("last_claimed.bin");
    auto l_u = load<float>("last_u.bin"), l_v = load<float>("last_v.bin"), l_invz = load<float>("last_invz.bin"), l_angle = load<float>("last_angle.bin");
    auto l_oct = load<int32_t>("last_octave.bin");
    auto l_ur = load<float>("last_uright.bin");
    orb_frame_desc dC = dA;
    dC.u_right = l_ur.data();
    orb_frame *fc = nullptr;
    chk(orb_frame_create(&dC, &fc), "orb_frame_create");
    orb_last_points LP = {na, l_valid.data(), l_u.data(), l_v.data(), l_invz.data(), l_oct.data(), l_angle.data(), da.data(), l_obs.data()};
    std::vector<int32_t> asg_last(na);
    std::vector<double> t_last;
    int nm_last = 0, sw_last = 0;
    int64_t cand_last = 0;
    for (int i = 0; i < warm + reps; i++) {
        const double t0 = now_us();
        chk(orb_search_by_projection_last(fc, &LP, l_claimed.data(), 15.0f, 0, 0, 40.0f, 1, asg_last.data(), &nm_last), "orb_search_by_projection_last");
        const double t1 = now_us();
        if (i >= warm) t_last.push_back(t1 - t0);
    }
    chk(orb_frame_last_stats(fc, &cand_last, &sw_last), "orb_frame_last_stats");
    save("last_assigned.bin", asg_last);

    // ---- SearchByProjection(Frame = B, vpMapPoints, th = 3), ref :45-129
    auto p_in = load<uint8_t>("pts_in_view.bin"), p_obs = load<uint8_t>("pts_obs.bin"), p_claimed = load<uint8_t>("pts_claimed.bin");
    auto p_x = load<float>("pts_x.bin"), p_y = load<float>("pts_y.bin"), p_xr = load<float>("pts_xr.bin"), p_vc = load<float>("pts_view_cos.bin");
    auto p_lvl = load<int32_t>("pts_level.bin");
    orb_track_points TP = {na, p_in.data(), p_x.data(), p_y.data(), p_xr.data(), p_lvl.data(), p_vc.data(), da.data(), p_obs.data()};
    std::vector<int32_t> asg_pts(nb);
    std::vector<double> t_pts;
    int nm_pts = 0, sw_pts = 0;
    int64_t cand_pts = 0;
    for (int i = 0; i < warm + reps; i++) {
        const double t0 = now_us();
        chk(orb_search_by_projection_points(fb, &TP, p_claimed.data(), 3.0f, 0.8f, asg_pts.data(), &nm_pts), "orb_search_by_projection_points");
        const double t1 = now_us();
        if (i >= warm) t_pts.push_back(t1 - t0);
    }
    chk(orb_frame_last_stats(fb, &cand_pts, &sw_pts), "orb_frame_last_stats");
    save("pts_assigned.bin", asg_pts);

    const Stat sc = stat(t_create), sf = stat(t_from), se = stat(t_extract), s1 = stat(t_init), s1d = stat(t_init_dev), s2 = stat(t_last), s3 = stat(t_pts);
    printf("{\"reps\": %d, \"n_a\": %d, \"n_b\": %d, \"extract_equals_input\": %s, "
           "\"frame_create_host_us\": {\"median\": %.2f, \"min\": %.2f}, \"frame_from_extractor_us\": {\"median\": %.2f, \"min\": %.2f}, "
           "\"extract_host_call_us\": {\"median\": %.2f, \"min\": %.2f}, "
           "\"search_for_initialization\": {\"us_median\": %.2f, \"us_min\": %.2f, \"nmatches\": %d, \"candidates\": %lld, \"sweeps\": %d, "
           "\"us_median_device_built_index\": %.2f, \"device_built_index_same_result\": %s}, "
           "\"search_by_projection_last\": {\"us_median\": %.2f, \"us_min\": %.2f, \"nmatches\": %d, \"candidates\": %lld, \"sweeps\": %d}, "
           "\"search_by_projection_points\": {\"us_median\": %.2f, \"us_min\": %.2f, \"nmatches\": %d, \"candidates\": %lld, \"sweeps\": %d}, "
           "\"launches\": {\"frame_b\": %lld, \"frame_cur\": %lld}}\n",
           reps, na, nb, same_extract ? "true" : "false", sc.med, sc.mn, sf.med, sf.mn, se.med, se.mn, s1.med, s1.mn, nm_init,
           (long long)cand_init, sw_init, s1d.med, nm_init_dev == nm_init ? "true" : "false", s2.med, s2.mn, nm_last, (long long)cand_last, sw_last,
           s3.med, s3.mn, nm_pts, (long long)cand_pts, sw_pts, (long long)orb_frame_launch_count(fb), (long long)orb_frame_launch_count(fc));
    orb_frame_destroy(fa); orb_frame_destroy(fb); orb_frame_destroy(fc); orb_frame_destroy(fb_dev);
    orb_extractor_destroy(ex);
    return 0;
}
