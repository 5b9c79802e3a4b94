// Stand-in for <opencv2/opencv.hpp> (reference include/vi_slam/common_include.h:27).
#pragma once
#include "../cv_compat.hpp"
