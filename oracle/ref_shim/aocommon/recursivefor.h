// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/recursivefor.h (ours). Facet
// tasks run one after another on the calling thread; the loops nested inside each task use
// the oracle's thread pool. No result of the reference depends on the partition.
#ifndef AOCOMMON_RECURSIVEFOR_H_
#define AOCOMMON_RECURSIVEFOR_H_
#include <cstddef>

namespace aocommon {
class RecursiveFor {
 public:
  RecursiveFor() { instance_ = this; }
  ~RecursiveFor() {
    if (instance_ == this) instance_ = nullptr;
  }
  RecursiveFor(const RecursiveFor&) = delete;
  static RecursiveFor* GetInstance() { return instance_; }
  template <typename F>
  static void NestedRun(size_t begin, size_t end, F&& f) {
    for (size_t i = begin; i != end; ++i) f(i);
  }
  template <typename F>
  void Run(size_t begin, size_t end, F&& f) {
    for (size_t i = begin; i != end; ++i) f(i);
  }

 private:
  static inline RecursiveFor* instance_ = nullptr;
};
}  // namespace aocommon
#endif
