// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/dynamicfor.h (ours).
#ifndef AOCOMMON_DYNAMICFOR_H_
#define AOCOMMON_DYNAMICFOR_H_
#include <cstddef>
#include <functional>

#include "../../oracle.h"
#include "recursivefor.h"

namespace aocommon {
template <typename Iter>
class DynamicFor {
 public:
  DynamicFor() = default;
  explicit DynamicFor(size_t) {}
  template <typename F>
  void Run(Iter begin, Iter end, F&& f) {
    if constexpr (std::is_invocable_v<F, Iter, size_t>) {
      oracle::DynamicFor(begin, end, [&](size_t i) { f(i, size_t(0)); });
    } else {
      oracle::DynamicFor(begin, end, [&](size_t i) { f(i); });
    }
  }
};
}  // namespace aocommon
#endif
