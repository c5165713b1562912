// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/imageaccessor.h (ours): the
// abstract load/store interface WorkTableEntry holds (work_table_entry.h:70-83).
#ifndef AOCOMMON_IMAGEACCESSOR_H_
#define AOCOMMON_IMAGEACCESSOR_H_
#include <cstddef>
namespace aocommon {
class ImageAccessor {
 public:
  virtual ~ImageAccessor() = default;
  virtual size_t Width() const { return 0; }
  virtual size_t Height() const { return 0; }
  virtual void Load(float* data) const = 0;
  virtual void Store(const float* data) = 0;
};
}  // namespace aocommon
#endif
