// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/cloned_ptr.h (ours): a unique_ptr
// whose copy deep-copies the pointee.
#ifndef AOCOMMON_CLONED_PTR_H_
#define AOCOMMON_CLONED_PTR_H_
#include <memory>
namespace aocommon {
template <typename T>
class cloned_ptr {
 public:
  cloned_ptr() = default;
  cloned_ptr(std::nullptr_t) {}
  explicit cloned_ptr(T* p) : p_(p) {}
  cloned_ptr(std::unique_ptr<T>&& p) : p_(std::move(p)) {}
  cloned_ptr(const cloned_ptr& o) : p_(o.p_ ? new T(*o.p_) : nullptr) {}
  cloned_ptr(cloned_ptr&&) noexcept = default;
  cloned_ptr& operator=(const cloned_ptr& o) {
    if (this != &o) p_.reset(o.p_ ? new T(*o.p_) : nullptr);
    return *this;
  }
  cloned_ptr& operator=(cloned_ptr&&) noexcept = default;
  cloned_ptr& operator=(std::nullptr_t) {
    p_.reset();
    return *this;
  }
  void reset(T* p = nullptr) { p_.reset(p); }
  T* get() const { return p_.get(); }
  T& operator*() const { return *p_; }
  T* operator->() const { return p_.get(); }
  explicit operator bool() const { return bool(p_); }
  bool operator==(std::nullptr_t) const { return p_ == nullptr; }
  bool operator!=(std::nullptr_t) const { return p_ != nullptr; }

 private:
  std::unique_ptr<T> p_;
};
}  // namespace aocommon
#endif
