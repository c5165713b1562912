// ORACLE — TEST INFRASTRUCTURE ONLY.
// Stand-in for aocommon/optionalnumber.h (aocommon @7798e81 is an empty submodule in
// /root/reference). Written from the call sites in the reference: default-constructed =
// unset, value ctor / assignment, HasValue / operator bool / operator* / ValueOr / Reset,
// and comparisons against a plain number that are false while unset
// (e.g. `maxValue < 0.0f`, generic_clean.cc:171).
#ifndef RDL_REF_SHIM_OPTIONALNUMBER_H_
#define RDL_REF_SHIM_OPTIONALNUMBER_H_
#include <optional>
namespace aocommon {
template <typename T>
class OptionalNumber {
 public:
  constexpr OptionalNumber() = default;
  constexpr explicit OptionalNumber(T v) : v_(v) {}
  constexpr OptionalNumber(std::nullopt_t) {}
  constexpr OptionalNumber& operator=(T v) {
    v_ = v;
    return *this;
  }
  constexpr bool HasValue() const { return v_.has_value(); }
  constexpr explicit operator bool() const { return v_.has_value(); }
  constexpr T operator*() const { return *v_; }
  constexpr T& operator*() { return *v_; }
  constexpr T ValueOr(T other) const { return v_.value_or(other); }
  constexpr void Reset() { v_.reset(); }
  friend constexpr bool operator<(const OptionalNumber& a, T b) { return a.v_ && *a.v_ < b; }
  friend constexpr bool operator>(const OptionalNumber& a, T b) { return a.v_ && *a.v_ > b; }
  friend constexpr bool operator<=(const OptionalNumber& a, T b) { return a.v_ && *a.v_ <= b; }
  friend constexpr bool operator>=(const OptionalNumber& a, T b) { return a.v_ && *a.v_ >= b; }
  friend constexpr bool operator==(const OptionalNumber& a, T b) { return a.v_ && *a.v_ == b; }
  friend constexpr bool operator!=(const OptionalNumber& a, T b) { return !(a == b); }
  friend constexpr bool operator==(const OptionalNumber& a, const OptionalNumber& b) {
    return a.v_ == b.v_;
  }
 private:
  std::optional<T> v_;
};
}  // namespace aocommon
#endif
