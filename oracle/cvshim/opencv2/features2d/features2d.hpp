// TEST INFRASTRUCTURE (oracle shim): cv::FAST lives in core.hpp of this shim.
#pragma once
#include "../core/core.hpp"
