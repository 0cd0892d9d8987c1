"""ar_orbslam2_b200 -- B200-native ORB-SLAM2 front end (pyramid, FAST + quadtree, orientation,
rBRIEF, Hamming matching, stereo SAD) behind the reference's ORBextractor / ORBmatcher /
Frame::ComputeStereoMatches interfaces.  The compute lives in liborbgpu.so (CUDA sm_100a, C ABI in
include/orbgpu.h); this package is the thin host-side mirror used by tests and benchmarks.
There is no CPU fallback: importing works anywhere, computing needs the built library and a GPU.
"""
__version__ = "0.1.0"
