// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/throwruntimeerror.h (ours).
#ifndef AOCOMMON_THROWRUNTIMEERROR_H_
#define AOCOMMON_THROWRUNTIMEERROR_H_
#include <sstream>
#include <stdexcept>
namespace aocommon {
template <typename... Args>
[[noreturn]] inline void ThrowRuntimeError(const Args&... args) {
  std::ostringstream s;
  (s << ... << args);
  throw std::runtime_error(s.str());
}
}  // namespace aocommon
#endif
