// TEST INFRASTRUCTURE (oracle shim): resize / GaussianBlur / copyMakeBorder live in core.hpp of this shim.
#pragma once
#include "../core/core.hpp"
