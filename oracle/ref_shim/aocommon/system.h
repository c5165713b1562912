// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/system.h (ours).
#ifndef AOCOMMON_SYSTEM_H_
#define AOCOMMON_SYSTEM_H_
#include <cstddef>
#include <thread>
namespace aocommon::system {
inline size_t ProcessorCount() {
  const unsigned n = std::thread::hardware_concurrency();
  return n ? n : 1;
}
}  // namespace aocommon::system
#endif
