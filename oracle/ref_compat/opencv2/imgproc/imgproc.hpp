// Stand-in for the OpenCV header of the same name: the reference's
// include/vi_slam/common_include.h includes it.  Everything lives in cv_compat.hpp.
#pragma once
#include "../../cv_compat.hpp"
