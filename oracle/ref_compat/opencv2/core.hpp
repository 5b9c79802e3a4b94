// Stand-in for <opencv2/core.hpp> (used to compile the VIORB_WITH_OPENCV branches of include/viorb_*.hpp in tests).
#pragma once
#include "../cv_compat.hpp"
