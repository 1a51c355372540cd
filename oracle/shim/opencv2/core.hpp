// Minimal header-only stand-in for <opencv2/core.hpp>; see ../cvshim.hpp.
#pragma once
#include "../cvshim.hpp"
