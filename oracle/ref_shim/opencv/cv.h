// CPU ORACLE support: stand-in for the OpenCV-3 legacy umbrella header named by include/ORBextractor.h:24.
#pragma once
#include "../opencv2/core/core.hpp"
