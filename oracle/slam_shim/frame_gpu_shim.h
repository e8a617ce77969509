// TEST INFRASTRUCTURE.  Force-included when the reference's src/Frame.cc is compiled VERBATIM against the PRODUCT's
// drop-in c_orb_slam_b200/csrc/adapter/ORBextractor.h (tests/cpp/Makefile): same stand-ins as frame_shim.h for
// MapPoint / KeyFrame / ORBVocabulary / Converter, but ORB_SLAM2::ORBextractor is the class a maintainer of the
// reference would drop in -- it defines the reference's own include guard, so `#include "ORBextractor.h"` in the
// reference's Frame.h is a no-op.  Frame's constructors then run the B200 extractor through the C ABI unchanged.
#pragma once
#define SLAM_SHIM_GPU_EXTRACTOR
#include "opencv2/core/core.hpp"
#include "adapter/ORBextractor.h"
#include "frame_shim.h"
