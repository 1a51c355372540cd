// Minimal header-only stand-in for <opencv2/imgcodecs.hpp>; see ../cvshim.hpp.
#pragma once
#include "../cvshim.hpp"
