// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/threadpool.h (ours).
#ifndef AOCOMMON_THREADPOOL_H_
#define AOCOMMON_THREADPOOL_H_
#include <cstddef>

#include "../../oracle.h"
#include "recursivefor.h"

namespace aocommon {
class ThreadPool {
 public:
  static ThreadPool& GetInstance() {
    static ThreadPool pool;
    return pool;
  }
  void SetNThreads(size_t n) { oracle::SetThreadCount(n); }
  size_t NThreads() const { return oracle::ThreadCount(); }
};
}  // namespace aocommon
#endif
