// Minimal header-only stand-in for <opencv2/features2d.hpp>; see ../cvshim.hpp.
#pragma once
#include "../cvshim.hpp"
