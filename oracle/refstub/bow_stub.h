// ORACLE -- TEST INFRASTRUCTURE ONLY.
//
// Stand-ins that let the reference's DBoW2 (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h, FORB.cpp, BowVector.cpp,
// FeatureVector.cpp, ScoringObject.cpp) compile unmodified against oracle/cvstub: cv::FileStorage / cv::FileNode are
// only reached from the YAML save()/load() members, which the tests never call (the reference itself loads its
// vocabulary with loadFromTextFile, src/System.cc:67-75), so they abort if they are ever executed.
#pragma once
#include "../cvstub/cvstub.h"

#include <cstdio>
#include <fstream>
#include <iostream>
#include <sstream>   // the reference gets these through the OpenCV headers
#include <string>

namespace cv {

class FileNode {
public:
    FileNode operator[](const char *) const { dead(); return FileNode(); }
    FileNode operator[](const std::string &) const { dead(); return FileNode(); }
    FileNode operator[](int) const { dead(); return FileNode(); }
    size_t size() const { dead(); return 0; }
    operator int() const { dead(); return 0; }
    operator double() const { dead(); return 0; }
    operator std::string() const { dead(); return std::string(); }
private:
    static void dead() { std::fprintf(stderr, "bow_stub: cv::FileNode is not available in the oracle build\n"); std::abort(); }
};

class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage(const char *, int) {}
    bool isOpened() const { return false; }
    FileNode operator[](const std::string &) const { return FileNode()[0]; }
    template <typename T> FileStorage &operator<<(const T &) {
        std::fprintf(stderr, "bow_stub: cv::FileStorage is not available in the oracle build\n");
        std::abort();
        return *this;
    }
};

}  // namespace cv
