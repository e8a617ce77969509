// TEST INFRASTRUCTURE: part of the minimal OpenCV stand-in, see opencv2/core/core.hpp
#pragma once
#include "opencv2/core/core.hpp"
#include <cstdlib>
namespace cv {
// Only referenced by ComputeKeyPointsOld (ORBextractor.cc:855-1032), which is dead code in the
// reference (its call is commented out, :1057).  Declared so the file compiles; never runs.
struct KeyPointsFilter {
  static void retainBest(std::vector<KeyPoint>&, int) { std::abort(); }
};
}  // namespace cv
