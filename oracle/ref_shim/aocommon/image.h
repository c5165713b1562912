// ORACLE — TEST INFRASTRUCTURE ONLY.
// Stand-in for aocommon/image.h (aocommon @ 7798e817, not vendored in the reference tree):
// just the ImageBase<T> surface the reference's hot-path sources use, so that those sources
// can be compiled where they lie under /root/reference (oracle/Makefile target `ref`).
// Semantics restated from the call sites and the reference's tests;
// this file is ours, nothing here is copied from aocommon.
#ifndef AOCOMMON_IMAGE_H_
#define AOCOMMON_IMAGE_H_

#include <math.h>

#include <algorithm>
#include <cassert>
#include <cmath>
#include <complex>
#include <cstddef>
#include <cstring>
#include <limits>
#include <utility>
#include <vector>

#include "uvector.h"

namespace aocommon {

template <typename NumT>
class ImageBase {
 public:
  using value_type = NumT;
  using num_t = NumT;
  using iterator = NumT*;
  using const_iterator = const NumT*;

  ImageBase() = default;
  ImageBase(size_t width, size_t height)
      : data_(new NumT[width * height]), width_(width), height_(height), owner_(true) {}
  ImageBase(size_t width, size_t height, NumT initial) : ImageBase(width, height) {
    std::fill(begin(), end(), initial);
  }
  // non-owning view on caller memory
  ImageBase(NumT* data, size_t width, size_t height)
      : data_(data), width_(width), height_(height), owner_(false) {}
  ImageBase(const ImageBase& o) : ImageBase(o.width_, o.height_) {
    std::copy(o.begin(), o.end(), begin());
  }
  ImageBase(ImageBase&& o) noexcept
      : data_(o.data_), width_(o.width_), height_(o.height_), owner_(o.owner_) {
    o.data_ = nullptr;
    o.width_ = o.height_ = 0;
    o.owner_ = true;
  }
  ~ImageBase() { Release(); }

  ImageBase& operator=(const ImageBase& o) {
    if (this == &o) return *this;
    if (Size() != o.Size() || !owner_) {
      Release();
      data_ = new NumT[o.Size()];
      owner_ = true;
    }
    width_ = o.width_;
    height_ = o.height_;
    std::copy(o.begin(), o.end(), begin());
    return *this;
  }
  ImageBase& operator=(ImageBase&& o) noexcept {
    if (this == &o) return *this;
    Release();
    data_ = o.data_;
    width_ = o.width_;
    height_ = o.height_;
    owner_ = o.owner_;
    o.data_ = nullptr;
    o.width_ = o.height_ = 0;
    o.owner_ = true;
    return *this;
  }
  ImageBase& operator=(NumT v) {
    std::fill(begin(), end(), v);
    return *this;
  }
  ImageBase& Assign(const NumT* b, const NumT* e) {
    std::copy(b, e, begin());
    return *this;
  }

  NumT* Data() { return data_; }
  const NumT* Data() const { return data_; }
  size_t Width() const { return width_; }
  size_t Height() const { return height_; }
  size_t Size() const { return width_ * height_; }
  bool Empty() const { return Size() == 0; }
  iterator begin() { return data_; }
  const_iterator begin() const { return data_; }
  iterator end() { return data_ + Size(); }
  const_iterator end() const { return data_ + Size(); }
  NumT& operator[](size_t i) { return data_[i]; }
  const NumT& operator[](size_t i) const { return data_[i]; }
  NumT& Value(size_t x, size_t y) { return data_[y * width_ + x]; }
  const NumT& Value(size_t x, size_t y) const { return data_[y * width_ + x]; }

  void Reset() {
    Release();
    data_ = nullptr;
    width_ = height_ = 0;
    owner_ = true;
  }

  ImageBase& operator+=(const ImageBase& o) {
    for (size_t i = 0; i != Size(); ++i) data_[i] += o.data_[i];
    return *this;
  }
  ImageBase& operator-=(const ImageBase& o) {
    for (size_t i = 0; i != Size(); ++i) data_[i] -= o.data_[i];
    return *this;
  }
  ImageBase& operator*=(const ImageBase& o) {
    for (size_t i = 0; i != Size(); ++i) data_[i] *= o.data_[i];
    return *this;
  }
  ImageBase& operator*=(NumT f) {
    for (size_t i = 0; i != Size(); ++i) data_[i] *= f;
    return *this;
  }
  ImageBase& operator/=(NumT f) { return (*this) *= NumT(1.0) / f; }
  ImageBase& operator+=(NumT f) {
    for (size_t i = 0; i != Size(); ++i) data_[i] += f;
    return *this;
  }

  ImageBase& Sqrt() {
    for (size_t i = 0; i != Size(); ++i) data_[i] = std::sqrt(data_[i]);
    return *this;
  }
  ImageBase& Square() {
    for (size_t i = 0; i != Size(); ++i) data_[i] *= data_[i];
    return *this;
  }
  ImageBase& SquareWithFactor(NumT f) {
    for (size_t i = 0; i != Size(); ++i) data_[i] *= data_[i] * f;
    return *this;
  }
  ImageBase& SqrtWithFactor(NumT f) {
    for (size_t i = 0; i != Size(); ++i) data_[i] = std::sqrt(data_[i]) * f;
    return *this;
  }
  // d += rhs * f. The reference build (-O3, GCC default -ffp-contract=fast) fuses this into
  // an FMA; the restated oracle and the CUDA path do so explicitly, so it is explicit here.
  ImageBase& AddWithFactor(const ImageBase& rhs, NumT f) {
    for (size_t i = 0; i != Size(); ++i) data_[i] = std::fma(rhs.data_[i], f, data_[i]);
    return *this;
  }
  ImageBase& AddSquared(const ImageBase& rhs) {
    for (size_t i = 0; i != Size(); ++i) data_[i] = std::fma(rhs.data_[i], rhs.data_[i], data_[i]);
    return *this;
  }
  ImageBase& AddSquared(const ImageBase& rhs, NumT f) {
    for (size_t i = 0; i != Size(); ++i)
      data_[i] = std::fma(rhs.data_[i] * rhs.data_[i], f, data_[i]);
    return *this;
  }
  ImageBase& Negate() {
    for (size_t i = 0; i != Size(); ++i) data_[i] = -data_[i];
    return *this;
  }

  // ---- centred crop / zero-pad: start = (outer - inner) / 2 -----------------------------
  template <typename T>
  static void TrimBox(T* output, size_t x1, size_t y1, size_t box_w, size_t box_h, const T* input,
                      size_t in_w, size_t /*in_h*/) {
    for (size_t y = 0; y != box_h; ++y)
      std::copy_n(input + (y + y1) * in_w + x1, box_w, output + y * box_w);
  }
  template <typename T>
  static void Trim(T* output, size_t out_w, size_t out_h, const T* input, size_t in_w,
                   size_t in_h) {
    TrimBox(output, (in_w - out_w) / 2, (in_h - out_h) / 2, out_w, out_h, input, in_w, in_h);
  }
  template <typename T>
  static void Untrim(T* output, size_t out_w, size_t out_h, const T* input, size_t in_w,
                     size_t in_h) {
    const size_t sx = (out_w - in_w) / 2, sy = (out_h - in_h) / 2;
    std::fill_n(output, out_w * out_h, T(0));
    for (size_t y = 0; y != in_h; ++y)
      std::copy_n(input + y * in_w, in_w, output + (y + sy) * out_w + sx);
  }
  ImageBase TrimBox(size_t x1, size_t y1, size_t box_w, size_t box_h) const {
    ImageBase out(box_w, box_h);
    TrimBox(out.Data(), x1, y1, box_w, box_h, Data(), Width(), Height());
    return out;
  }
  ImageBase Trim(size_t out_w, size_t out_h) const {
    ImageBase out(out_w, out_h);
    Trim(out.Data(), out_w, out_h, Data(), Width(), Height());
    return out;
  }
  ImageBase Untrim(size_t out_w, size_t out_h) const {
    ImageBase out(out_w, out_h);
    Untrim(out.Data(), out_w, out_h, Data(), Width(), Height());
    return out;
  }
  // per dimension: centred crop when smaller, centred zero-pad when larger
  ImageBase Resize(size_t new_w, size_t new_h) const {
    const size_t mid_w = std::min(new_w, Width()), mid_h = std::min(new_h, Height());
    ImageBase mid = (mid_w == Width() && mid_h == Height()) ? *this : Trim(mid_w, mid_h);
    if (mid_w == new_w && mid_h == new_h) return mid;
    return mid.Untrim(new_w, new_h);
  }

  static void CopyMasked(NumT* to, size_t to_x, size_t to_y, size_t to_w, const NumT* from,
                         size_t from_w, size_t from_h, const bool* from_mask) {
    for (size_t y = 0; y != from_h; ++y)
      for (size_t x = 0; x != from_w; ++x)
        if (from_mask[y * from_w + x]) to[(to_y + y) * to_w + to_x + x] = from[y * from_w + x];
  }
  static void AddSubImage(NumT* to, size_t to_x, size_t to_y, size_t to_w, const NumT* from,
                          size_t from_w, size_t from_h) {
    for (size_t y = 0; y != from_h; ++y)
      for (size_t x = 0; x != from_w; ++x)
        to[(to_y + y) * to_w + to_x + x] += from[y * from_w + x];
  }

  // ---- reductions (sequential, in the image's own type) ---------------------------------
  NumT Sum() const {
    NumT s = 0;
    for (size_t i = 0; i != Size(); ++i) s += data_[i];
    return s;
  }
  NumT Average() const { return Sum() / NumT(Size()); }
  NumT Min() const { return *std::min_element(begin(), end()); }
  NumT Max() const { return *std::max_element(begin(), end()); }
  NumT RMS() const {
    NumT s = 0;
    for (size_t i = 0; i != Size(); ++i) s += data_[i] * data_[i];
    return std::sqrt(s / NumT(Size()));
  }
  static NumT Median(const NumT* data, size_t size) {
    std::vector<NumT> copy;
    copy.reserve(size);
    for (size_t i = 0; i != size; ++i)
      if (std::isfinite(data[i])) copy.push_back(data[i]);
    return MedianOf(copy);
  }
  // median over finite values, MAD = median(|v - median|), stddev = 1.4826 * MAD
  void MedianAndStdDevFromMAD(NumT& median, NumT& stddev) const {
    std::vector<NumT> copy;
    copy.reserve(Size());
    for (size_t i = 0; i != Size(); ++i)
      if (std::isfinite(data_[i])) copy.push_back(data_[i]);
    median = MedianOf(copy);
    for (NumT& v : copy) v = std::fabs(v - median);
    stddev = NumT(1.4826) * MedianOf(copy);
  }
  std::pair<NumT, NumT> MedianAndStdDevFromMAD() const {
    NumT m, s;
    MedianAndStdDevFromMAD(m, s);
    return {m, s};
  }
  NumT StdDevFromMAD() const { return MedianAndStdDevFromMAD().second; }

 private:
  static NumT MedianOf(std::vector<NumT>& v) {
    if (v.empty()) return NumT(0);
    const size_t mid = v.size() / 2;
    std::nth_element(v.begin(), v.begin() + mid, v.end());
    if (v.size() % 2 == 1) return v[mid];
    const NumT hi = v[mid];
    const NumT lo = *std::max_element(v.begin(), v.begin() + mid);
    return (lo + hi) * NumT(0.5);
  }
  void Release() {
    if (owner_) delete[] data_;
  }
  NumT* data_ = nullptr;
  size_t width_ = 0, height_ = 0;
  bool owner_ = true;
};

using Image = ImageBase<float>;
using DImage = ImageBase<double>;
using ImageF = ImageBase<float>;

}  // namespace aocommon
#endif
