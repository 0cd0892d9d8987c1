// ORACLE -- TEST INFRASTRUCTURE ONLY.  Stand-in for the OpenCV header of the same name; see cvstub.h.
#pragma once
#include "../cvstub.h"
