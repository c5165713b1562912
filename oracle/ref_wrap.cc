// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" wrappers around the reference's OWN leaf sources, compiled where they lie
// under /root/reference (see Makefile target `ref`): cpp/math/peak_finder.cc,
// cpp/algorithms/simple_clean.cc, cpp/math/dijkstra_splitter.cc and the header-only
// cpp/utils/fft_size_calculations.h. The rest of the reference cannot be built here
// (aocommon / schaapcommon submodules are empty, FFTW/Boost/GSL/casacore are absent).
#include <cstddef>
#include <cstdint>
#include <vector>

#include "algorithms/iuwt/image_analysis.h"
#include "algorithms/iuwt/iuwt_decomposition.h"
#include "algorithms/iuwt/iuwt_mask.h"
#include "math/rms_image.h"
// Mad and GetMaxAbs are private members of the reference's class; the wrapper below needs them
#define private public
#include "algorithms/iuwt_deconvolution_algorithm.h"
#undef private
#include "algorithms/multiscale/multiscale_transforms.h"
#include "algorithms/simple_clean.h"
#include "math/dijkstra_splitter.h"
#include "math/peak_finder.h"
#include "utils/fft_size_calculations.h"

using namespace radler;

extern "C" {

// variant: 0 = Simple, 1 = Avx (what Find dispatches to on an AVX build), 2 = FindWithMask
int ref_find_peak(int variant, const float* image, uint64_t w, uint64_t h, int allow_negative,
                  uint64_t start_y, uint64_t end_y, uint64_t hb, uint64_t vb,
                  const uint8_t* mask, uint64_t* x, uint64_t* y, float* value) {
  size_t px = 0, py = 0;
  aocommon::OptionalNumber<float> r;
  if (variant == 2)
    r = math::peak_finder::FindWithMask(image, w, h, px, py, allow_negative != 0, start_y, end_y,
                                        reinterpret_cast<const bool*>(mask), size_t(hb),
                                        size_t(vb));
  else if (variant == 1)
    r = math::peak_finder::Avx(image, w, h, px, py, allow_negative != 0, start_y, end_y, hb, vb);
  else
    r = math::peak_finder::Simple(image, w, h, px, py, allow_negative != 0, start_y, end_y, hb,
                                  vb);
  *x = px;
  *y = py;
  *value = r.ValueOr(0.0f);
  return r.HasValue() ? 1 : 0;
}

void ref_subtract_image(float* image, const float* psf, uint64_t w, uint64_t h, uint64_t x,
                        uint64_t y, float factor) {
  algorithms::simple_clean::SubtractImage(image, psf, w, h, x, y, factor);
}

void ref_partial_subtract_image(float* image, uint64_t iw, uint64_t ih, const float* psf,
                                uint64_t pw, uint64_t ph, uint64_t x, uint64_t y, float factor,
                                uint64_t start_y, uint64_t end_y) {
  if (iw == pw && ih == ph)
    algorithms::simple_clean::PartialSubtractImage(image, psf, iw, ih, x, y, factor, start_y,
                                                   end_y);
  else
    algorithms::simple_clean::PartialSubtractImage(image, iw, ih, psf, pw, ph, x, y, factor,
                                                   start_y, end_y);
}

uint64_t ref_good_fft_size(uint64_t n) { return utils::CalculateGoodFFTSize(n); }
uint64_t ref_convolution_size(double scale, uint64_t n, double padding) {
  return utils::GetConvolutionSize(scale, n, padding);
}

void ref_divide_vertically(const float* image, float* output, uint64_t w, uint64_t h,
                           uint64_t x1, uint64_t x2) {
  math::DijkstraSplitter(w, h).DivideVertically(image, output, x1, x2);
}
void ref_divide_horizontally(const float* image, float* output, uint64_t w, uint64_t h,
                             uint64_t y1, uint64_t y2) {
  math::DijkstraSplitter(w, h).DivideHorizontally(image, output, y1, y2);
}
void ref_add_vertical_divider(const float* image, float* scratch, float* output, uint64_t w,
                              uint64_t h, uint64_t x1, uint64_t x2) {
  math::DijkstraSplitter(w, h).AddVerticalDivider(image, scratch, output, x1, x2);
}
void ref_add_horizontal_divider(const float* image, float* scratch, float* output, uint64_t w,
                                uint64_t h, uint64_t y1, uint64_t y2) {
  math::DijkstraSplitter(w, h).AddHorizontalDivider(image, scratch, output, y1, y2);
}
void ref_flood_vertical_area(const float* subdivision, uint64_t w, uint64_t h, uint64_t sub_x,
                             uint8_t* mask, uint64_t* x, uint64_t* subwidth) {
  size_t ox = 0, ow = 0;
  math::DijkstraSplitter(w, h).FloodVerticalArea(subdivision, sub_x,
                                                 reinterpret_cast<bool*>(mask), ox, ow);
  *x = ox;
  *subwidth = ow;
}
void ref_flood_horizontal_area(const float* subdivision, uint64_t w, uint64_t h, uint64_t sub_y,
                               uint8_t* mask, uint64_t* y, uint64_t* subheight) {
  size_t oy = 0, oh = 0;
  math::DijkstraSplitter(w, h).FloodHorizontalArea(subdivision, sub_y,
                                                   reinterpret_cast<bool*>(mask), oy, oh);
  *y = oy;
  *subheight = oh;
}
void ref_get_bounding_mask(const uint8_t* vertical_mask, uint64_t w, uint64_t h, uint64_t vx,
                           uint64_t vw, const uint8_t* horizontal_mask, uint8_t* mask,
                           uint64_t* sub_x, uint64_t* sub_y, uint64_t* sub_w, uint64_t* sub_h) {
  size_t x = 0, y = 0, sw = 0, sh = 0;
  math::DijkstraSplitter(w, h).GetBoundingMask(reinterpret_cast<const bool*>(vertical_mask), vx,
                                               vw, reinterpret_cast<const bool*>(horizontal_mask),
                                               reinterpret_cast<bool*>(mask), x, y, sw, sh);
  *sub_x = x;
  *sub_y = y;
  *sub_w = sw;
  *sub_h = sh;
}

// multiscale_transforms.h:91-195 / multiscale_transforms.cc:11-23 of the reference itself
uint64_t ref_make_shape_function(float scale, uint64_t max_n, int shape, float* out) {
  size_t n = 0;
  const aocommon::Image k = algorithms::multiscale::MultiScaleTransforms::MakeShapeFunction(
      scale, n, max_n, shape == 1 ? MultiscaleShape::kGaussianShape
                                  : MultiscaleShape::kTaperedQuadraticShape);
  if (out) std::copy_n(k.Data(), n * n, out);
  return n;
}
void ref_ms_transform(float* images, uint64_t n_images, uint64_t w, uint64_t h, float scale,
                      int shape) {
  algorithms::multiscale::MultiScaleTransforms t(
      w, h, shape == 1 ? MultiscaleShape::kGaussianShape : MultiscaleShape::kTaperedQuadraticShape);
  std::vector<aocommon::Image> list;
  for (uint64_t i = 0; i != n_images; ++i) {
    list.emplace_back(w, h);
    std::copy_n(images + i * w * h, w * h, list.back().Data());
  }
  aocommon::Image scratch(w, h);
  t.Transform(list, scratch, scale);
  for (uint64_t i = 0; i != n_images; ++i) std::copy_n(list[i].Data(), w * h, images + i * w * h);
}
void ref_add_shape_component(float* image, uint64_t w, uint64_t h, float scale, uint64_t x,
                             uint64_t y, float gain, int shape) {
  algorithms::multiscale::MultiScaleTransforms::AddShapeComponent(
      image, w, h, scale, x, y, gain,
      shape == 1 ? MultiscaleShape::kGaussianShape : MultiscaleShape::kTaperedQuadraticShape);
}

// ---- IUWT decomposition (iuwt_decomposition.{h,cc}) of the reference itself ----------------
// coefficients: (n_scales + 1) planes; the last plane (the smooth remainder) is only filled
// when include_largest != 0.
void ref_iuwt_decompose(const float* input, uint64_t w, uint64_t h, int n_scales,
                        int include_largest, float* coefficients) {
  algorithms::iuwt::IuwtDecomposition d(n_scales, w, h);
  aocommon::Image scratch(w, h);
  d.Decompose(input, scratch.Data(), include_largest != 0);
  for (int s = 0; s != n_scales + (include_largest ? 1 : 0); ++s)
    std::copy_n(d[s].Coefficients().Data(), w * h, coefficients + size_t(s) * w * h);
}
// Decompose(x, x, false): the call IuwtDeconvolutionAlgorithm makes with the input aliased to the
// scratch image (iuwt_deconvolution_algorithm.cc:344,385,685,780). n_scales planes.
void ref_iuwt_decompose_aliased(float* x, uint64_t w, uint64_t h, int n_scales,
                                float* coefficients) {
  algorithms::iuwt::IuwtDecomposition d(n_scales, w, h);
  d.Decompose(x, x, false);
  for (int s = 0; s != n_scales; ++s)
    std::copy_n(d[s].Coefficients().Data(), w * h, coefficients + size_t(s) * w * h);
}
// masks (optional): n_scales planes of w*h bytes, applied with ApplyMask before Recompose.
void ref_iuwt_recompose(const float* coefficients, uint64_t w, uint64_t h, int n_scales,
                        int include_largest, const uint8_t* masks, float* out) {
  algorithms::iuwt::IuwtDecomposition d(n_scales, w, h);
  for (int s = 0; s != n_scales + 1; ++s) {
    d[s].Coefficients() = aocommon::Image(w, h, 0.0f);
    if (s < n_scales || include_largest)
      std::copy_n(coefficients + size_t(s) * w * h, w * h, d[s].Coefficients().Data());
  }
  if (masks) {
    algorithms::iuwt::IuwtMask m(n_scales, w, h);
    for (int s = 0; s != n_scales; ++s)
      for (size_t i = 0; i != w * h; ++i) m[s][i] = masks[size_t(s) * w * h + i] != 0;
    d.ApplyMask(m);
  }
  aocommon::Image o(w, h);
  d.Recompose(o, include_largest != 0 && !masks);
  std::copy_n(o.Data(), w * h, out);
}
// IuwtDecomposition::Convolve(image, to_scale): repeated B3 smoothing, "simple" tap order
void ref_iuwt_smooth(float* image, uint64_t w, uint64_t h, int to_scale) {
  algorithms::iuwt::IuwtDecomposition d(to_scale, w, h);
  aocommon::Image img(w, h);
  std::copy_n(image, w * h, img.Data());
  d.Convolve(img, to_scale);
  std::copy_n(img.Data(), w * h, image);
}

// ---- IUWT structure selection (iuwt/image_analysis.cc) and plane measurements
// (iuwt_deconvolution_algorithm.cc:104-167) of the reference itself --------------------------
namespace {
struct IuwtFixture {
  algorithms::iuwt::IuwtDecomposition d;
  algorithms::iuwt::IuwtMask m;
  aocommon::UVector<float> thresholds;
  IuwtFixture(const float* coefficients, const uint8_t* masks, uint64_t w, uint64_t h,
              int n_scales, const float* thr)
      : d(n_scales, w, h), m(n_scales, w, h), thresholds(thr, thr + n_scales) {
    for (int s = 0; s != n_scales; ++s) {
      d[s].Coefficients() = aocommon::Image(w, h);
      std::copy_n(coefficients + size_t(s) * w * h, w * h, d[s].Coefficients().Data());
      for (size_t i = 0; i != w * h; ++i) m[s][i] = masks[size_t(s) * w * h + i] != 0;
    }
  }
  void Store(uint8_t* masks, uint64_t w, uint64_t h, int n_scales) {
    for (int s = 0; s != n_scales; ++s)
      for (size_t i = 0; i != w * h; ++i) masks[size_t(s) * w * h + i] = m[s][i] ? 1 : 0;
  }
};
}  // namespace

uint64_t ref_iuwt_select_structures(const float* coefficients, uint8_t* masks, uint64_t w,
                                    uint64_t h, int n_scales, const float* thresholds,
                                    int min_scale, int end_scale, float clean_border,
                                    const uint8_t* prior_mask) {
  IuwtFixture f(coefficients, masks, w, h, n_scales, thresholds);
  size_t area = 0;
  algorithms::iuwt::image_analysis::SelectStructures(
      f.d, f.m, f.thresholds, min_scale, end_scale, clean_border,
      reinterpret_cast<const bool*>(prior_mask), area);
  f.Store(masks, w, h, n_scales);
  return area;
}
uint64_t ref_iuwt_floodfill(const float* coefficients, uint8_t* masks, uint64_t w, uint64_t h,
                            int n_scales, const float* thresholds, int min_scale, int end_scale,
                            uint64_t x, uint64_t y, int scale, float clean_border,
                            const uint8_t* prior_mask) {
  IuwtFixture f(coefficients, masks, w, h, n_scales, thresholds);
  size_t area = 0;
  const algorithms::iuwt::image_analysis::Component seed(x, y, scale);
  if (prior_mask)
    algorithms::iuwt::image_analysis::MaskedFloodfill(
        f.d, f.m, f.thresholds, min_scale, end_scale, seed, clean_border,
        reinterpret_cast<const bool*>(prior_mask), area);
  else
    algorithms::iuwt::image_analysis::Floodfill(f.d, f.m, f.thresholds, min_scale, end_scale, seed,
                                                clean_border, area);
  f.Store(masks, w, h, n_scales);
  return area;
}
uint64_t ref_iuwt_floodfill_2d(const float* image, uint8_t* mask, uint64_t w, uint64_t h,
                               float threshold, int use_abs, uint64_t x, uint64_t y) {
  const algorithms::iuwt::image_analysis::Component2D seed(x, y);
  bool* m = reinterpret_cast<bool*>(mask);
  if (use_abs) {
    std::vector<algorithms::iuwt::image_analysis::Component2D> area;
    algorithms::iuwt::image_analysis::FloodFill2D(image, m, threshold, seed, w, h, area);
    return area.size();
  }
  size_t area = 0;
  algorithms::iuwt::image_analysis::FloodFill2D(image, m, threshold, seed, w, h, area);
  return area;
}
float ref_iuwt_mad(const float* image, uint64_t w, uint64_t h) {
  algorithms::IuwtDeconvolutionAlgorithm a(w, h, 0.1f, 0.8f, 0.0f, true, nullptr, 0.0f);
  return a.Mad(image);
}
float ref_iuwt_max_abs(const float* image, uint64_t w, uint64_t h, float clean_border,
                       int allow_negative, const uint8_t* mask, uint64_t* x, uint64_t* y) {
  algorithms::IuwtDeconvolutionAlgorithm a(w, h, 0.1f, 0.8f, clean_border, allow_negative != 0,
                                           reinterpret_cast<const bool*>(mask), 0.0f);
  aocommon::Image img(w, h);
  std::copy_n(image, w * h, img.Data());
  size_t px = 0, py = 0;
  const float v = a.GetMaxAbs(img, px, py, w);
  *x = px;
  *y = py;
  return v;
}
// cpp/math/rms_image.cc: the sliding minimum / maximum and the RMS -> factor conversion
void ref_rms_sliding_minimum(const float* in, float* out, uint64_t w, uint64_t h, uint64_t window,
                             int maximum) {
  aocommon::Image input(w, h), output;
  std::copy_n(in, w * h, input.Data());
  if (maximum)
    math::rms_image::SlidingMaximum(output, input, window);
  else
    math::rms_image::SlidingMinimum(output, input, window);
  std::copy_n(output.Data(), w * h, out);
}
double ref_rms_make_factor_image(float* image, uint64_t w, uint64_t h, double strength) {
  aocommon::Image img(w, h);
  std::copy_n(image, w * h, img.Data());
  const double stddev = math::rms_image::MakeRmsFactorImage(img, strength);
  std::copy_n(img.Data(), w * h, image);
  return stddev;
}
}
