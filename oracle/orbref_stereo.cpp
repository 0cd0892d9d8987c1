// ORACLE -- TEST INFRASTRUCTURE ONLY (see orbref_cv.h).  CPU restatement of
// Frame::ComputeStereoMatches (ORB_SLAM2/src/Frame.cc:471-643) and of
// ORBmatcher::DescriptorDistance (ORB_SLAM2/src/ORBmatcher.cc:1650-1666).
#include "orbref.h"
#include "orbref_internal.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <utility>
#include <vector>

using namespace orbref;

// ref ORBmatcher::DescriptorDistance, ORBmatcher.cc:1650-1666 (the SWAR bit-count, verbatim arithmetic)
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

extern "C" {

int orbref_descriptor_distance(const uint8_t *a, const uint8_t *b, int n, int32_t *dist) {
    for (int i = 0; i < n; i++) dist[i] = descriptor_distance(a + 32 * (size_t)i, b + 32 * (size_t)i);
    return ORB_OK;
}

// ref Frame::ComputeStereoMatches, Frame.cc:471-643.  left/right hold the pyramids of the pair
// (mpORBextractorLeft/Right->mvImagePyramid); mb is Frame::mb as the reference intends it (mbf/fx).
int orbref_stereo_match(orbref_extractor *left, orbref_extractor *right, const orb_keypoint *keysL,
                        const uint8_t *descL, int N, const orb_keypoint *keysR, const uint8_t *descR, int Nr,
                        float mbf, float mb, float *uRight, float *depth, int32_t *sad_out, int *n_matched) {
    const int TH_HIGH = 100;
    for (int i = 0; i < N; i++) { uRight[i] = -1.0f; depth[i] = -1.0f; if (sad_out) sad_out[i] = -1; }
    const int nRows = left->lv[0].h;

    std::vector<std::vector<size_t>> vRowIndices(nRows);
    for (int iR = 0; iR < Nr; iR++) {
        const orb_keypoint &kp = keysR[iR];
        const float kpY = kp.y;
        const float r = 2.0f * left->scale[kp.octave];   // Frame::mvScaleFactors come from the left extractor (Frame.cc:77)
        const int maxr = (int)std::ceil(kpY + r);
        const int minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);   // the reference does not guard; never out of range for extractor output
    }

    const float minZ = mb;
    const float minD = -3;
    const float maxD = mbf / minZ;

    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < N; iL++) {
        const orb_keypoint &kpL = keysL[iL];
        const int levelL = kpL.octave;
        const float vL = kpL.y, uL = kpL.x;
        const std::vector<size_t> &vCandidates = vRowIndices[(size_t)vL];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;

        int bestDist = TH_HIGH;
        size_t bestIdxR = 0;
        const uint8_t *dL = descL + 32 * (size_t)iL;
        for (size_t iC = 0; iC < vCandidates.size(); iC++) {
            const size_t iR = vCandidates[iC];
            const orb_keypoint &kpR = keysR[iR];
            if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
            const float uR = kpR.x;
            if (uR >= minU && uR <= maxU) {
                const int dist = descriptor_distance(dL, descR + 32 * iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }

        if (bestDist < TH_HIGH) {
            const float uR0 = keysR[bestIdxR].x;
            const float scaleFactor = left->inv_scale[kpL.octave];
            const float scaleduL = std::round(kpL.x * scaleFactor);
            const float scaledvL = std::round(kpL.y * scaleFactor);
            const float scaleduR0 = std::round(uR0 * scaleFactor);

            const int w = 5;
            Level &LL = left->lv[kpL.octave];
            Level &LR = right->lv[kpL.octave];
            // IL = patch - centre value (float), ref :566-568
            float IL[11][11];
            {
                const int y0 = (int)(scaledvL - w), x0 = (int)(scaleduL - w);
                for (int y = 0; y < 11; y++)
                    for (int x = 0; x < 11; x++) IL[y][x] = (float)LL.interior()[(ptrdiff_t)(y0 + y) * LL.pitch + x0 + x];
                const float c = IL[w][w];
                for (int y = 0; y < 11; y++)
                    for (int x = 0; x < 11; x++) IL[y][x] = IL[y][x] - c * 1.0f;
            }
            int bestDistS = INT_MAX;
            int bestincR = 0;
            const int L = 5;
            std::vector<float> vDists(2 * L + 1);
            const float iniu = scaleduR0 + L - w;
            const float endu = scaleduR0 + L + w + 1;
            if (iniu < 0 || endu >= LR.w) continue;

            for (int incR = -L; incR <= +L; incR++) {
                const int y0 = (int)(scaledvL - w), x0 = (int)(scaleduR0 + incR - w);
                float IR[11][11];
                for (int y = 0; y < 11; y++)
                    for (int x = 0; x < 11; x++) IR[y][x] = (float)LR.interior()[(ptrdiff_t)(y0 + y) * LR.pitch + x0 + x];
                const float c = IR[w][w];
                double s = 0;   // cv::norm(..., NORM_L1) accumulates in double; values are exact integers
                for (int y = 0; y < 11; y++)
                    for (int x = 0; x < 11; x++) s += std::fabs((double)(IL[y][x] - (IR[y][x] - c * 1.0f)));
                float dist = (float)s;
                if (dist < bestDistS) { bestDistS = (int)dist; bestincR = incR; }
                vDists[L + incR] = dist;
            }
            if (bestincR == -L || bestincR == L) continue;

            const float dist1 = vDists[L + bestincR - 1];
            const float dist2 = vDists[L + bestincR];
            const float dist3 = vDists[L + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;

            float bestuR = left->scale[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= 0 && disparity < maxD) {
                if (disparity <= 0) {
                    disparity = 0.01;
                    bestuR = uL - 0.01;
                }
                depth[iL] = mbf / disparity;
                uRight[iL] = bestuR;
                if (sad_out) sad_out[iL] = bestDistS;
                vDistIdx.push_back(std::pair<int, int>(bestDistS, iL));
            }
        }
    }
    if (vDistIdx.empty()) {   // the reference indexes vDistIdx[0] here (UB); defined as an explicit status
        if (n_matched) *n_matched = 0;
        return ORB_ERR_EMPTY;
    }
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    int kept = (int)vDistIdx.size();
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        uRight[vDistIdx[i].second] = -1;
        depth[vDistIdx[i].second] = -1;
        kept--;
    }
    if (n_matched) *n_matched = kept;
    return ORB_OK;
}

}  // extern "C"
