// ORACLE -- TEST INFRASTRUCTURE ONLY.
//
// Minimal stand-ins for ORB_SLAM2::Frame / KeyFrame / MapPoint so that the REFERENCE'S OWN
// ORB_SLAM2/src/ORBmatcher.cc (+ include/ORBmatcher.h) compiles unmodified: the build pre-defines the include
// guards FRAME_H / KEYFRAME_H / MAPPOINT_H (so the reference's headers, which drag in DBoW2 / g2o / Pangolin
// types, contribute nothing) and force-includes this file (oracle/Makefile.ref).  Only the members ORBmatcher.cc
// touches exist.  The few methods with behaviour are restated from the reference and cited; everything else is a
// plain data holder filled by oracle/ref_matcher_driver.cpp.
#pragma once
#include "../cvstub/cvstub.h"

#include <cmath>
#include <map>
#include <mutex>
#include <set>
#include <vector>

#include "Thirdparty/DBoW2/DBoW2/BowVector.h"       // the reference's own (pure STL) headers
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

using namespace std;   // the reference headers leak this into ORBmatcher.h (it declares `vector<MapPoint *>` unqualified)

#define FRAME_GRID_ROWS 48   // ref Frame.h:39
#define FRAME_GRID_COLS 64   // ref Frame.h:40

namespace ORB_SLAM2 {

class KeyFrame;
class Frame;

class MapPoint {
public:
    // variables used by the tracking (ref MapPoint.h:87-97)
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0;
    bool mbTrackInView = false;
    int mnTrackScaleLevel = 0;
    float mTrackViewCos = 0;

    // data holder
    int id = -1;
    bool bad = false;
    int nObs = 0;
    cv::Mat descriptor, worldPos, normal;
    float mfMaxDistance = 0, mfMinDistance = 0;
    std::map<KeyFrame *, size_t> observations;
    MapPoint *replaced = nullptr;

    bool isBad() { return bad; }
    cv::Mat GetDescriptor() { return descriptor.clone(); }
    int Observations() { return nObs; }
    cv::Mat GetWorldPos() { return worldPos.clone(); }
    cv::Mat GetNormal() { return normal.clone(); }
    float GetMaxDistanceInvariance() { return 1.2f * mfMaxDistance; }   // ref MapPoint.cc:352-355
    float GetMinDistanceInvariance() { return 0.8f * mfMinDistance; }   // ref MapPoint.cc:346-349
    int PredictScale(const float &currentDist, const float &logScaleFactor) {   // ref MapPoint.cc:357-366 (this fork does not clamp)
        float ratio = mfMaxDistance / currentDist;
        return ceil(log(ratio) / logScaleFactor);
    }
    bool IsInKeyFrame(KeyFrame *pKF) { return observations.count(pKF) != 0; }
    int GetIndexInKeyFrame(KeyFrame *pKF) { return observations.count(pKF) ? (int)observations[pKF] : -1; }
    void AddObservation(KeyFrame *pKF, size_t idx) { if (!observations.count(pKF)) { observations[pKF] = idx; nObs++; } }
    void Replace(MapPoint *pMP) { replaced = pMP; bad = true; }
};

// With ORBREF_REAL_FRAME the reference's own include/Frame.h + src/Frame.cc are compiled (its dependencies
// ORBVocabulary / Converter are the data-less stand-ins below); otherwise a stand-in Frame is used.
class ORBVocabulary {   // ref ORBVocabulary.h:28 (a DBoW2 vocabulary tree; the vocabulary file is not in the repository)
public:
    void transform(const std::vector<cv::Mat> &, DBoW2::BowVector &, DBoW2::FeatureVector &, int) {}
};
class Converter {       // ref Converter.h:32 (the real header needs Eigen and g2o)
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat &D) {
        std::vector<cv::Mat> v;
        for (int j = 0; j < D.rows; j++) v.push_back(D.row(j));
        return v;
    }
};

#ifndef ORBREF_REAL_FRAME
// the members of Frame that ORBmatcher.cc reads (ref Frame.h)
class Frame {
public:
    int N = 0;
    std::vector<cv::KeyPoint> mvKeys, mvKeysUn;
    std::vector<float> mvuRight;
    cv::Mat mDescriptors;
    std::vector<MapPoint *> mvpMapPoints;
    std::vector<bool> mvbOutlier;
    std::vector<float> mvScaleFactors;
    float mfLogScaleFactor = 0;
    DBoW2::FeatureVector mFeatVec;
    cv::Mat mTcw;
    float mbf = 0, mb = 0;
    static float fx, fy, cx, cy;
    static float mnMinX, mnMaxX, mnMinY, mnMaxY;
    static float mfGridElementWidthInv, mfGridElementHeightInv;
    std::vector<std::size_t> mGrid[FRAME_GRID_COLS][FRAME_GRID_ROWS];

    // restated from ref Frame.cc:387-397
    bool PosInGrid(const cv::KeyPoint &kp, int &posX, int &posY) {
        posX = round((kp.pt.x - mnMinX) * mfGridElementWidthInv);
        posY = round((kp.pt.y - mnMinY) * mfGridElementHeightInv);
        if (posX < 0 || posX >= FRAME_GRID_COLS || posY < 0 || posY >= FRAME_GRID_ROWS) return false;
        return true;
    }
    // restated from ref Frame.cc:235-250
    void AssignFeaturesToGrid() {
        for (int i = 0; i < N; i++) {
            const cv::KeyPoint &kp = mvKeysUn[i];
            int nGridPosX, nGridPosY;
            if (PosInGrid(kp, nGridPosX, nGridPosY)) mGrid[nGridPosX][nGridPosY].push_back(i);
        }
    }
    // restated from ref Frame.cc:332-385
    vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r, const int minLevel = -1, const int maxLevel = -1) const {
        vector<size_t> vIndices;
        vIndices.reserve(N);
        const int nMinCellX = max(0, (int)floor((x - mnMinX - r) * mfGridElementWidthInv));
        if (nMinCellX >= FRAME_GRID_COLS) return vIndices;
        const int nMaxCellX = min((int)FRAME_GRID_COLS - 1, (int)ceil((x - mnMinX + r) * mfGridElementWidthInv));
        if (nMaxCellX < 0) return vIndices;
        const int nMinCellY = max(0, (int)floor((y - mnMinY - r) * mfGridElementHeightInv));
        if (nMinCellY >= FRAME_GRID_ROWS) return vIndices;
        const int nMaxCellY = min((int)FRAME_GRID_ROWS - 1, (int)ceil((y - mnMinY + r) * mfGridElementHeightInv));
        if (nMaxCellY < 0) return vIndices;
        const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const vector<size_t> vCell = mGrid[ix][iy];
                if (vCell.empty()) continue;
                for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
                    const cv::KeyPoint &kpUn = mvKeysUn[vCell[j]];
                    if (bCheckLevels) {
                        if (kpUn.octave < minLevel) continue;
                        if (maxLevel >= 0)
                            if (kpUn.octave > maxLevel) continue;
                    }
                    const float distx = kpUn.pt.x - x;
                    const float disty = kpUn.pt.y - y;
                    if (fabs(distx) < r && fabs(disty) < r) vIndices.push_back(vCell[j]);
                }
            }
        }
        return vIndices;
    }
};

#endif  // ORBREF_REAL_FRAME

// the members of KeyFrame that ORBmatcher.cc reads (ref KeyFrame.h)
class KeyFrame {
public:
    int N = 0;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight;
    cv::Mat mDescriptors;
    DBoW2::FeatureVector mFeatVec;
    std::vector<MapPoint *> mvpMapPoints;
    std::vector<float> mvScaleFactors, mvLevelSigma2, mvInvLevelSigma2;
    float mfLogScaleFactor = 0;
    float fx = 0, fy = 0, cx = 0, cy = 0, mbf = 0;
    int mnMinX = 0, mnMinY = 0, mnMaxX = 0, mnMaxY = 0;
    int mnGridCols = FRAME_GRID_COLS, mnGridRows = FRAME_GRID_ROWS;
    float mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::vector<std::vector<std::vector<size_t>>> mGrid;
    cv::Mat Rcw, tcw, Ow;

    vector<MapPoint *> GetMapPointMatches() { return mvpMapPoints; }
    MapPoint *GetMapPoint(const size_t &idx) { return mvpMapPoints[idx]; }
    set<MapPoint *> GetMapPoints() {
        set<MapPoint *> s;
        for (MapPoint *p : mvpMapPoints) if (p && !p->isBad()) s.insert(p);
        return s;
    }
    void AddMapPoint(MapPoint *pMP, const size_t &idx) { mvpMapPoints[idx] = pMP; }
    cv::Mat GetRotation() { return Rcw.clone(); }
    cv::Mat GetTranslation() { return tcw.clone(); }
    cv::Mat GetCameraCenter() { return Ow.clone(); }
    bool IsInImage(const float &x, const float &y) const { return (x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY); }   // ref KeyFrame.cc:560-563
    // restated from ref KeyFrame.cc:518-558
    vector<size_t> GetFeaturesInArea(const float &x, const float &y, const float &r) const {
        vector<size_t> vIndices;
        vIndices.reserve(N);
        const int nMinCellX = max(0, (int)floor((x - mnMinX - r) * mfGridElementWidthInv));
        if (nMinCellX >= mnGridCols) return vIndices;
        const int nMaxCellX = min((int)mnGridCols - 1, (int)ceil((x - mnMinX + r) * mfGridElementWidthInv));
        if (nMaxCellX < 0) return vIndices;
        const int nMinCellY = max(0, (int)floor((y - mnMinY - r) * mfGridElementHeightInv));
        if (nMinCellY >= mnGridRows) return vIndices;
        const int nMaxCellY = min((int)mnGridRows - 1, (int)ceil((y - mnMinY + r) * mfGridElementHeightInv));
        if (nMaxCellY < 0) return vIndices;
        for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
            for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
                const vector<size_t> vCell = mGrid[ix][iy];
                for (size_t j = 0, jend = vCell.size(); j < jend; j++) {
                    const cv::KeyPoint &kpUn = mvKeysUn[vCell[j]];
                    const float distx = kpUn.pt.x - x;
                    const float disty = kpUn.pt.y - y;
                    if (fabs(distx) < r && fabs(disty) < r) vIndices.push_back(vCell[j]);
                }
            }
        }
        return vIndices;
    }
};

}  // namespace ORB_SLAM2
