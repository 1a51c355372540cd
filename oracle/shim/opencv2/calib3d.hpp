// Minimal header-only stand-in for <opencv2/calib3d.hpp>; see ../cvshim.hpp.
#pragma once
#include "../cvshim.hpp"
