// ORACLE -- TEST INFRASTRUCTURE ONLY.
//
// Flat C entry point around the REFERENCE'S OWN MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:224-285,
// compiled unmodified): a map point observed by N key frames, key frame j holding descriptor row j of the group at
// index 0.  Key frames are allocated in one array so that std::map<KeyFrame*, size_t> iterates them in group order
// (the reference's order is pointer order).  `protected` is opened so the driver can read mDescriptor back.
#include "MapPoint.h"
#include "ORBmatcher.h"

#include <cstdint>
#include <cstring>
#include <vector>

using namespace ORB_SLAM2;

extern "C" {

// desc: all groups' rows, group p = rows [off[p], off[p+1]); bad[row] = KeyFrame::isBad() of that observation.
// out_desc[p*32..] = the map point's mDescriptor afterwards (zeros when the reference returns early), best_row[p] =
// index inside the group of the first row equal to it among the non-bad ones in order -- only for reporting; parity is
// judged on the descriptor bytes.
int ref_distinctive_descriptors(const uint8_t *desc, const int32_t *off, int npoints, const uint8_t *bad, uint8_t *out_desc) {
    Map map;
    for (int p = 0; p < npoints; p++) {
        const int n = off[p + 1] - off[p];
        std::vector<KeyFrame> kfs(n > 0 ? n : 1);
        cv::Mat pos(3, 1, CV_32F);
        pos.at<float>(0) = pos.at<float>(1) = pos.at<float>(2) = 0.f;
        KeyFrame ref;
        MapPoint mp(pos, &ref, &map);
        for (int j = 0; j < n; j++) {
            kfs[j].mDescriptors = cv::Mat(1, 32, CV_8U);
            std::memcpy(kfs[j].mDescriptors.ptr(), desc + 32 * (size_t)(off[p] + j), 32);
            kfs[j].mvuRight.assign(1, -1.f);
            kfs[j].mbBad = bad && bad[off[p] + j];
            mp.AddObservation(&kfs[j], 0);
        }
        mp.ComputeDistinctiveDescriptors();
        cv::Mat d = mp.GetDescriptor();
        if (d.empty()) std::memset(out_desc + 32 * (size_t)p, 0, 32);
        else std::memcpy(out_desc + 32 * (size_t)p, d.ptr(), 32);
    }
    return 0;
}

}  // extern "C"
