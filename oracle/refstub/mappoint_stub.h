// ORACLE -- TEST INFRASTRUCTURE ONLY.
//
// Stand-ins that let the reference's OWN src/MapPoint.cc + include/MapPoint.h compile unmodified (oracle/Makefile.ref ->
// oracle/_ref/libmappoint_ref.so): KeyFrame.h, Frame.h and Map.h (which pull DBoW2 templates, g2o and the whole SLAM
// graph) are switched off by pre-defining their include guards; the classes below carry exactly the members
// MapPoint.cc touches.  ORBmatcher::DescriptorDistance is linked from liborbmatcher_ref.so (the reference's compiled
// ORBmatcher.cc).
#pragma once
#include "../cvstub/cvstub.h"

#include <map>
#include <mutex>
#include <set>
#include <vector>

#define KEYFRAME_H
#define FRAME_H
#define MAP_H
#define ORBVOCABULARY_H
#define KEYFRAMEDATABASE_H
#define ORBEXTRACTOR_H

using namespace std;   // the reference's headers rely on it leaking out of ORBextractor.h / the DBoW2 headers

namespace ORB_SLAM2 {

class MapPoint;
class KeyFrame;

class Frame {   // members read by MapPoint::MapPoint(Pos, Map*, Frame*, idxF), src/MapPoint.cc:46-72
public:
    long unsigned int mnId = 0;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvScaleFactors;
    int mnScaleLevels = 0;
    cv::Mat mDescriptors;
    cv::Mat mOw;
    cv::Mat GetCameraCenter() { return mOw.clone(); }
};

class KeyFrame {   // members read by src/MapPoint.cc (observations, descriptors, scale tables)
public:
    long unsigned int mnId = 0;
    long unsigned int mnFrameId = 0;
    std::vector<cv::KeyPoint> mvKeysUn;
    std::vector<float> mvuRight;
    std::vector<float> mvScaleFactors;
    int mnScaleLevels = 0;
    cv::Mat mDescriptors;
    cv::Mat mOw;
    bool mbBad = false;
    std::vector<MapPoint *> mvpMapPoints;
    cv::Mat GetCameraCenter() { return mOw.clone(); }
    bool isBad() { return mbBad; }
    void EraseMapPointMatch(const size_t &idx) { if (idx < mvpMapPoints.size()) mvpMapPoints[idx] = nullptr; }
    void EraseMapPointMatch(MapPoint *) {}
    void ReplaceMapPointMatch(const size_t &idx, MapPoint *pMP) { if (idx < mvpMapPoints.size()) mvpMapPoints[idx] = pMP; }
};

class Map {
public:
    std::mutex mMutexPointCreation;
    std::set<MapPoint *> erased;
    void EraseMapPoint(MapPoint *p) { erased.insert(p); }
};

}  // namespace ORB_SLAM2
