// void ORB_SLAM2::Frame::ComputeStereoMatches() (ref: ORB_SLAM2/include/Frame.h:95, body ORB_SLAM2/src/Frame.cc:471-643)
// over liborbgpu.  Build it with the reference's own Frame.cc and drop that file's definition of the same member: either
// delete / #if 0 lines 471-643 of Frame.cc, or -- without touching the file -- weaken the symbol in its object,
//     objcopy --weaken-symbol=_ZN9ORB_SLAM25Frame20ComputeStereoMatchesEv Frame.o
// so that this strong definition wins at link time (oracle/Makefile.ref does exactly that for the parity test).
// The stereo constructor (Frame.cc:66-122) then runs unchanged: ExtractORB on both extractors in two threads (:83-86), this
// association (:95), ComputeStereoFromRGBD untouched.
#include "Frame.h"
#include "ORBextractor.h"

namespace ORB_SLAM2 {

void Frame::ComputeStereoMatches() {
    // mvuRight / mvDepth of N entries, -1 where unmatched (:473-474); ORBGPU_ComputeStereoMatches fills them from the two
    // extractors' device-resident pyramids (the reference reads mvImagePyramid[octave], :566,583) and returns false where
    // the reference would index an empty vector (:630): both stay all -1 then.
    ORBGPU_ComputeStereoMatches(mpORBextractorLeft, mpORBextractorRight, mvKeys, mDescriptors, mvKeysRight, mDescriptorsRight, mbf,
                                mb, mvuRight, mvDepth);
}

}  // namespace ORB_SLAM2
