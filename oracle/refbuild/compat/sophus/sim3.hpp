// TEST INFRASTRUCTURE ONLY -- Sim3d is only a typedef target in the reference (util/NumType.h:49).
#ifndef DSO_COMPAT_SOPHUS_SIM3
#define DSO_COMPAT_SOPHUS_SIM3
#include "se3.hpp"
namespace Sophus { struct Sim3d { SE3d se3; double s; Sim3d() : s(1) {} }; typedef Sim3d Sim3; }
#endif
