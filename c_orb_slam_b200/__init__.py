"""c_orb_slam_b200 -- B200-native ORB front-end (extractor + Hamming matcher) behind ORB-SLAM2's
ORBextractor / ORBmatcher interface.  The compute path is liborb_gpu.so (include/orb_gpu.h)."""
from .api import (BUCKET_DTYPE, KP_DTYPE, PROJ_DTYPE, WINDOW_DTYPE, GpuFrame, ORBextractor, ORBmatcher, ORBVocabulary, OrbGpuError, imma_peak, umma_peak, launch_count, lib,  # noqa: F401
                  popc_peak)
