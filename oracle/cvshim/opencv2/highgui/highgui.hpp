// TEST INFRASTRUCTURE (oracle shim): os1's ORBextractor.cc includes this header but uses nothing from it.
#pragma once
#include "../core/core.hpp"
