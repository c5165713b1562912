// ORACLE — TEST INFRASTRUCTURE ONLY.
// Stand-in for aocommon/uvector.h (ours): a contiguous dynamic array with the std::vector
// surface the reference uses. Unlike std::vector<bool> it stores real bools, so data() works
// (the reference passes UVector<bool>::data() around as `const bool*` masks).
#ifndef RDL_REF_SHIM_UVECTOR_H_
#define RDL_REF_SHIM_UVECTOR_H_
#include <algorithm>
#include <cstddef>
#include <initializer_list>
#include <iterator>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <type_traits>

namespace aocommon {
template <typename T>
class UVector {
  static_assert(std::is_trivially_copyable_v<T>, "UVector holds trivially copyable types");

 public:
  using value_type = T;
  using iterator = T*;
  using const_iterator = const T*;
  using size_type = size_t;
  using reference = T&;
  using const_reference = const T&;

  UVector() = default;
  explicit UVector(size_t n) { resize(n); }
  UVector(size_t n, const T& v) { assign(n, v); }
  template <typename It, typename = typename std::iterator_traits<It>::iterator_category>
  UVector(It b, It e) { assign(b, e); }
  UVector(std::initializer_list<T> l) { assign(l.begin(), l.end()); }
  UVector(const UVector& o) { assign(o.begin(), o.end()); }
  UVector(UVector&& o) noexcept : d_(std::move(o.d_)), n_(o.n_), cap_(o.cap_) { o.n_ = o.cap_ = 0; }
  UVector& operator=(const UVector& o) {
    if (this != &o) assign(o.begin(), o.end());
    return *this;
  }
  UVector& operator=(UVector&& o) noexcept {
    d_ = std::move(o.d_);
    n_ = o.n_;
    cap_ = o.cap_;
    o.n_ = o.cap_ = 0;
    return *this;
  }
  UVector& operator=(std::initializer_list<T> l) {
    assign(l.begin(), l.end());
    return *this;
  }

  T* data() { return d_.get(); }
  const T* data() const { return d_.get(); }
  size_t size() const { return n_; }
  size_t capacity() const { return cap_; }
  bool empty() const { return n_ == 0; }
  iterator begin() { return d_.get(); }
  iterator end() { return d_.get() + n_; }
  const_iterator begin() const { return d_.get(); }
  const_iterator end() const { return d_.get() + n_; }
  const_iterator cbegin() const { return begin(); }
  const_iterator cend() const { return end(); }
  T& operator[](size_t i) { return d_[i]; }
  const T& operator[](size_t i) const { return d_[i]; }
  T& front() { return d_[0]; }
  const T& front() const { return d_[0]; }
  T& back() { return d_[n_ - 1]; }
  const T& back() const { return d_[n_ - 1]; }

  void reserve(size_t n) {
    if (n <= cap_) return;
    Storage nd(static_cast<T*>(std::calloc(n, sizeof(T))));
    if (n_) std::memcpy(static_cast<void*>(nd.get()), d_.get(), n_ * sizeof(T));
    d_ = std::move(nd);
    cap_ = n;
  }
  // new elements are zero bytes (the real uvector leaves them uninitialised)
  void resize(size_t n) {
    if (n > cap_) reserve(std::max(n, cap_ * 2));
    if (n > n_) std::memset(static_cast<void*>(d_.get() + n_), 0, (n - n_) * sizeof(T));
    n_ = n;
  }
  void resize(size_t n, const T& v) {
    const size_t old = n_;
    resize(n);
    if (n > old) std::fill(d_.get() + old, d_.get() + n, v);
  }
  void assign(size_t n, const T& v) {
    n_ = 0;
    resize(n);
    std::fill(begin(), end(), v);
  }
  template <typename It, typename = typename std::iterator_traits<It>::iterator_category>
  void assign(It b, It e) {
    n_ = 0;
    resize(size_t(std::distance(b, e)));
    std::copy(b, e, begin());
  }
  void push_back(const T& v) {
    const T copy = v;
    resize(n_ + 1);
    std::memcpy(static_cast<void*>(d_.get() + n_ - 1), &copy, sizeof(T));
  }
  // range append (an extension of the real uvector, used by component_list.h:71)
  template <typename It, typename = typename std::iterator_traits<It>::iterator_category>
  void push_back(It b, It e) {
    insert(end(), b, e);
  }
  template <typename... Args>
  T& emplace_back(Args&&... args) {
    push_back(T(std::forward<Args>(args)...));
    return back();
  }
  void pop_back() { --n_; }
  void clear() { n_ = 0; }
  void shrink_to_fit() {}
  void swap(UVector& o) noexcept {
    std::swap(d_, o.d_);
    std::swap(n_, o.n_);
    std::swap(cap_, o.cap_);
  }
  iterator insert(const_iterator pos, const T& v) {
    const size_t at = size_t(pos - begin());
    push_back(v);
    std::rotate(begin() + at, end() - 1, end());
    return begin() + at;
  }
  template <typename It, typename = typename std::iterator_traits<It>::iterator_category>
  iterator insert(const_iterator pos, It b, It e) {
    const size_t at = size_t(pos - begin()), cnt = size_t(std::distance(b, e)), old = n_;
    resize(n_ + cnt);
    std::copy(b, e, begin() + old);
    std::rotate(begin() + at, begin() + old, end());
    return begin() + at;
  }
  iterator erase(const_iterator pos) {
    const size_t at = size_t(pos - begin());
    std::copy(begin() + at + 1, end(), begin() + at);
    --n_;
    return begin() + at;
  }
  bool operator==(const UVector& o) const { return n_ == o.n_ && std::equal(begin(), end(), o.begin()); }
  bool operator!=(const UVector& o) const { return !(*this == o); }

 private:
  struct Free {
    void operator()(T* p) const { std::free(p); }
  };
  using Storage = std::unique_ptr<T[], Free>;
  Storage d_;
  size_t n_ = 0, cap_ = 0;
};
}  // namespace aocommon
#endif
