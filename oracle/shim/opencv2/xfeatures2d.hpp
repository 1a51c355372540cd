// Minimal header-only stand-in for <opencv2/xfeatures2d.hpp>; see ../cvshim.hpp.
#pragma once
#include "../cvshim.hpp"
