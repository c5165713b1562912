// ORACLE — TEST INFRASTRUCTURE ONLY.
// Minimal stand-in for aocommon/optionalnumber.h (aocommon @7798e81 is an empty
// submodule in /root/reference). Written from the call sites in
// cpp/math/peak_finder.{h,cc}: default-constructed = unset, explicit value ctor,
// HasValue / operator bool / operator* / ValueOr.
#ifndef RDL_REF_SHIM_OPTIONALNUMBER_H_
#define RDL_REF_SHIM_OPTIONALNUMBER_H_
#include <optional>
namespace aocommon {
template <typename T>
class OptionalNumber {
 public:
  constexpr OptionalNumber() = default;
  constexpr explicit OptionalNumber(T v) : v_(v) {}
  constexpr bool HasValue() const { return v_.has_value(); }
  constexpr explicit operator bool() const { return v_.has_value(); }
  constexpr T operator*() const { return *v_; }
  constexpr T ValueOr(T other) const { return v_.value_or(other); }
 private:
  std::optional<T> v_;
};
}  // namespace aocommon
#endif
