// Minimal header-only stand-in for <opencv2/opencv.hpp>; see ../cvshim.hpp.
#pragma once
#include "../cvshim.hpp"
