// CPU ORACLE support (test infrastructure, not product code).
// Pinhole::project / KannalaBrandt8::project: the reference's own bodies (oracle/_ref/camera_slice.inc, cut from
// src/CameraModels/Pinhole.cpp:46-56 and KannalaBrandt8.cpp:29-51 by oracle/Makefile), compiled in a translation unit WITHOUT a
// using-directive for std and with <cmath> only -- the state of KannalaBrandt8.cpp itself.  It matters for one line:
// `mvParameters[0] * r * cos(psi) + mvParameters[2]` (KannalaBrandt8.cpp:42-43) then binds to ::cos(double), i.e. the product and
// sum are double arithmetic rounded once to float.  (If a third-party header of the real build pulled in <math.h>, libstdc++
// would add the float overload and the line would be float arithmetic: the two readings differ by <= 1e-4 px, tests report it.)
#define SLAM_WORLD_NO_USING_STD
#include "slam_world.h"

namespace ORB_SLAM3 {
#include "camera_slice.inc"
}  // namespace ORB_SLAM3
