// TEST INFRASTRUCTURE: part of the minimal OpenCV stand-in, see opencv2/core/core.hpp
#pragma once
#include "opencv2/core/core.hpp"
