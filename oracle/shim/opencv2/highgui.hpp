// Minimal header-only stand-in for <opencv2/highgui.hpp>; see ../cvshim.hpp.
#pragma once
#include "../cvshim.hpp"
