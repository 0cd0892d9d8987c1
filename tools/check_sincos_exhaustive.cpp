// Exhaustive check of the oracle's restated cosf/sinf (oracle/orbref_cv.h: sincosf_libm) against the host
// libm on EVERY fp32 angle the descriptor can see: a * (float)(pi/180) for all floats a in [0, 360).
// Build+run:  g++ -O2 -ffp-contract=off -pthread tools/check_sincos_exhaustive.cpp -o /tmp/sc && /tmp/sc
// Result on the build box (glibc 2.39, x86-64): total 1135869952, sin mismatches 0, cos mismatches 0.
#include "../oracle/orbref_cv.h"

#include <atomic>
#include <cstdio>
#include <thread>
#include <vector>

int main() {
    const float factorPI = (float)(M_PI / 180.f);
    uint32_t hi;
    const float f360 = 360.f;
    memcpy(&hi, &f360, 4);
    const int NT = std::thread::hardware_concurrency() ? std::thread::hardware_concurrency() : 8;
    std::vector<std::thread> th;
    std::atomic<long> ms{0}, mc{0}, tot{0};
    for (int t = 0; t < NT; t++)
        th.emplace_back([&, t] {
            long a = 0, b = 0, c = 0;
            for (uint64_t u = t; u < hi; u += NT) {
                uint32_t uu = (uint32_t)u;
                float ang;
                memcpy(&ang, &uu, 4);
                const float x = ang * factorPI;
                volatile float vx = x;
                const float s = sinf(vx), co = cosf(vx);
                float ms_, mc_;
                orbref::sincosf_libm(x, &ms_, &mc_);
                if (s != ms_) a++;
                if (co != mc_) b++;
                c++;
            }
            ms += a; mc += b; tot += c;
        });
    for (auto &t : th) t.join();
    printf("total %ld sin mismatches %ld cos mismatches %ld\n", (long)tot, (long)ms, (long)mc);
    return (ms || mc) ? 1 : 0;
}
