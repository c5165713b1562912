// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/staticfor.h (ours): contiguous
// chunk per thread, delegated to the oracle's own thread pool (oracle/core.cc).
#ifndef AOCOMMON_STATICFOR_H_
#define AOCOMMON_STATICFOR_H_
#include <cstddef>
#include <functional>

#include "../../oracle.h"

namespace aocommon {
template <typename Iter>
class StaticFor {
 public:
  StaticFor() = default;
  explicit StaticFor(size_t) {}
  template <typename F>
  void Run(Iter begin, Iter end, F&& f) {
    if constexpr (std::is_invocable_v<F, Iter, Iter, size_t>) {
      oracle::StaticFor(begin, end, [&](size_t lo, size_t hi) { f(lo, hi, size_t(0)); });
    } else {
      oracle::StaticFor(begin, end, [&](size_t lo, size_t hi) { f(lo, hi); });
    }
  }
  size_t NThreads() const { return oracle::ThreadCount(); }
};
}  // namespace aocommon
#endif
