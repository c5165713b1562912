// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for schaapcommon/math/ellipse.h (ours).
#ifndef SCHAAPCOMMON_MATH_ELLIPSE_H_
#define SCHAAPCOMMON_MATH_ELLIPSE_H_
namespace schaapcommon::math {
struct Ellipse {
  double major = 0.0, minor = 0.0, position_angle = 0.0;
};
}  // namespace schaapcommon::math
#endif
