// Exercises the C++ drop-in shim the way Frame's stereo constructor does (ref Frame.cc:66-122): two extractors
// from two threads, then the stereo association.  Reads frames from raw files, writes results for the Python test
// to compare with the oracle.  usage: shim_selftest W H left.raw right.raw nfeatures out_prefix
#include <cstdio>
#include <cstdlib>
#include <string>
#include <algorithm>
#include <chrono>
#include <thread>
#include <vector>

#include "ORBextractor.h"
#include "MarkerORB.h"

static std::vector<unsigned char> slurp(const char *p, size_t n) {
    std::vector<unsigned char> b(n);
    FILE *f = fopen(p, "rb");
    if (!f || fread(b.data(), 1, n, f) != n) { fprintf(stderr, "cannot read %s\n", p); exit(2); }
    fclose(f);
    return b;
}
static void dump(const std::string &p, const void *d, size_t n) {
    FILE *f = fopen(p.c_str(), "wb");
    fwrite(d, 1, n, f);
    fclose(f);
}

int main(int argc, char **argv) {
    if (argc < 7) return 2;
    const int W = atoi(argv[1]), H = atoi(argv[2]), nf = atoi(argv[5]);
    std::vector<unsigned char> L = slurp(argv[3], (size_t)W * H), R = slurp(argv[4], (size_t)W * H);
    cv::Mat imLeft(H, W, CV_8UC1, L.data(), W), imRight(H, W, CV_8UC1, R.data(), W);
    ORB_SLAM2::ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);
    std::vector<cv::KeyPoint> kL, kR;
    cv::Mat dL, dR;
    std::thread tl([&] { exL(imLeft, cv::Mat(), kL, dL); });     // ref Frame.cc:83-86
    std::thread tr([&] { exR(imRight, cv::Mat(), kR, dR); });
    tl.join(); tr.join();
    std::vector<float> uR, depth;
    const float mbf = 386.1448f, mb = mbf / 718.856f;
    const bool ok = ORB_SLAM2::ORBGPU_ComputeStereoMatches(&exL, &exR, kL, dL, kR, dR, mbf, mb, uR, depth);
    const std::string pre = argv[6];
    dump(pre + "_kl.bin", kL.data(), kL.size() * sizeof(cv::KeyPoint));
    dump(pre + "_kr.bin", kR.data(), kR.size() * sizeof(cv::KeyPoint));
    dump(pre + "_dl.bin", dL.data, (size_t)dL.rows * 32);
    dump(pre + "_dr.bin", dR.data, (size_t)dR.rows * 32);
    dump(pre + "_ur.bin", uR.data(), uR.size() * 4);
    dump(pre + "_dp.bin", depth.data(), depth.size() * 4);
    // the host mirror of level 3 incl. the byte 7 px into the left border, as Frame.cc:583 may touch
    const cv::Mat &l3 = exL.mvImagePyramid[3];
    dump(pre + "_l3.bin", l3.data - 19 * l3.step - 19, (size_t)(l3.rows + 38) * l3.step);
    // the same frame through the single-call front end (orb_stereo_frame), twice more so that the CUDA-graph replay runs too
    std::vector<cv::KeyPoint> fkL, fkR;
    cv::Mat fdL, fdR;
    std::vector<float> fuR, fdepth;
    bool fok = true;
    for (int rep = 0; rep < 3; rep++)
        fok = ORB_SLAM2::ORBGPU_ExtractStereo(&exL, imLeft, imRight, fkL, fdL, fkR, fdR, mbf, mb, fuR, fdepth) && fok;
    dump(pre + "_fkl.bin", fkL.data(), fkL.size() * sizeof(cv::KeyPoint));
    dump(pre + "_fkr.bin", fkR.data(), fkR.size() * sizeof(cv::KeyPoint));
    dump(pre + "_fdl.bin", fdL.data, (size_t)fdL.rows * 32);
    dump(pre + "_fdr.bin", fdR.data, (size_t)fdR.rows * 32);
    dump(pre + "_fur.bin", fuR.data(), fuR.size() * 4);
    dump(pre + "_fdp.bin", fdepth.data(), fdepth.size() * 4);
    printf("fused ok=%d\n", (int)fok);
    {   // the marker pipeline's front end as Marker::setTargetImage / Match write it (ref Marker.cc:83-84,107-121,314-349)
        std::vector<cv::KeyPoint> mk1, mk2;
        cv::Mat md1, md2;
        { orbgpu_shim::ORB orb; orb(imLeft, cv::Mat(), mk1, md1); }    // `ORB orb; orb(mImage1, Mat(), mvKeys1, mDescriptors1);`
        { orbgpu_shim::ORB orb; orb(imRight, cv::Mat(), mk2, md2); }
        std::vector<cv::DMatch> bf, rm;
        orbgpu_shim::BruteForceMatchHamming(md1, md2, bf);
        orbgpu_shim::SearchMatches(md1, md2, rm);
        dump(pre + "_mk1.bin", mk1.data(), mk1.size() * sizeof(cv::KeyPoint));
        dump(pre + "_mk2.bin", mk2.data(), mk2.size() * sizeof(cv::KeyPoint));
        dump(pre + "_md1.bin", md1.data, (size_t)md1.rows * 32);
        dump(pre + "_md2.bin", md2.data, (size_t)md2.rows * 32);
        dump(pre + "_mbf.bin", bf.data(), bf.size() * sizeof(cv::DMatch));
        dump(pre + "_mrm.bin", rm.data(), rm.size() * sizeof(cv::DMatch));
        // AR-1.3's naive_nn_search / naive_nn_search2 with its file-scope minD / maxD (ORBMatcher.cpp:17-18)
        std::vector<cv::DMatch> n1, n2;
        unsigned int minD = 100, maxD = 0;
        orbgpu_shim::NaiveNNSearch(md1, md2, n1);
        orbgpu_shim::NaiveNNSearch2(md1, md2, n2, minD, maxD);
        dump(pre + "_mn1.bin", n1.data(), n1.size() * sizeof(cv::DMatch));
        dump(pre + "_mn2.bin", n2.data(), n2.size() * sizeof(cv::DMatch));
        printf("marker ok n1=%zu n2=%zu bf=%zu ratio=%zu nn=%zu nn2=%zu minD=%u maxD=%u\n", mk1.size(), mk2.size(), bf.size(), rm.size(), n1.size(),
               n2.size(), minD, maxD);
    }
    if (argc > 7) {   // latency of the calls a Tracking thread makes, without the host mirror of the pyramid (median of 50)
        auto med = [](std::vector<double> v) { std::sort(v.begin(), v.end()); return v[v.size() / 2]; };
        auto now = [] { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
        std::vector<double> t1, t2, t3;
        for (int i = 0; i < 55; i++) {
            double t0 = now();
            exL(imLeft, cv::Mat(), kL, dL);
            double ta = now();
            exR(imRight, cv::Mat(), kR, dR);
            ORB_SLAM2::ORBGPU_ComputeStereoMatches(&exL, &exR, kL, dL, kR, dR, mbf, mb, uR, depth);
            double tb = now();
            ORB_SLAM2::ORBGPU_ExtractStereo(&exL, imLeft, imRight, fkL, fdL, fkR, fdR, mbf, mb, fuR, fdepth);
            double tc = now();
            if (i >= 5) { t1.push_back(ta - t0); t2.push_back(tb - t0); t3.push_back(tc - tb); }
        }
        printf("latency_us extract=%.1f stereo_three_calls=%.1f stereo_single_call=%.1f\n", med(t1), med(t2), med(t3));
    }
    printf("shim ok=%d nL=%zu nR=%zu scale1=%.9g levels=%d l3=%dx%d step=%zu\n", (int)ok, kL.size(), kR.size(), exL.GetScaleFactors()[1],
           exL.GetLevels(), l3.cols, l3.rows, l3.step);
    return 0;
}
