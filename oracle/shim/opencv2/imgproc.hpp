// Minimal header-only stand-in for <opencv2/imgproc.hpp>; see ../cvshim.hpp.
#pragma once
#include "../cvshim.hpp"
