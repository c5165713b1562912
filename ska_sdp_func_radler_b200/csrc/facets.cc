// Facet geometry of ParallelDeconvolution on the host (see facets.h for the reference lines).
// Both orientations of the reference's splitter share one routine here, written in
// (lateral, primary) coordinates: a vertical divider advances along y (primary) and wanders in
// x (lateral); a horizontal one advances along x and wanders in y. The order in which
// neighbours enter the priority queue is the reference's, because std::priority_queue breaks
// equal-distance ties by insertion history and the chosen path must match pixel for pixel.
#include "facets.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <exception>
#include <limits>
#include <mutex>
#include <stdexcept>
#include <thread>

namespace rdl {
namespace {

struct Axes {
  size_t width;
  bool vertical;
  size_t operator()(size_t lat, size_t prim) const {
    return vertical ? lat + prim * width : prim + lat * width;
  }
};

void FillStrip(float* output, const Axes& at, size_t lo, size_t hi, size_t n_prim, float v) {
  for (size_t prim = 0; prim != n_prim; ++prim)
    for (size_t lat = lo; lat != hi; ++lat) output[at(lat, prim)] = v;
}

// The same path without a priority queue, when it is unique. The queue pops steps in order of
// distance, so a cell is settled by its cheapest incoming step: dist(c) = |image(c)| + min over
// the cells that can step to c (the three cells of the previous row and the two lateral
// neighbours). Costs are non-negative and float addition is monotone, so the lateral part of a
// path runs in one direction and the minimum over a row is two running sweeps whose sums are
// accumulated in path order, i.e. to the very same floats the queue arrives at. The queue's
// insertion history only matters where two incoming steps tie; the walk back from the exit
// checks that at every cell and gives up (returns false) on the first tie or non-finite value,
// and the caller then runs the queue.
bool UniqueDividingPath(const float* image, float* output, const Axes& at, size_t lo, size_t hi,
                        size_t n_prim) {
  const size_t span = hi - lo;
  const float inf = std::numeric_limits<float>::infinity();
  std::vector<float> dist(span * n_prim), entry(span), cost(span), back(span);
  for (size_t prim = 0; prim != n_prim; ++prim) {
    float* d = &dist[prim * span];
    const float* prev = d - span;
    for (size_t i = 0; i != span; ++i) {
      cost[i] = std::fabs(image[at(lo + i, prim)]);
      if (!std::isfinite(cost[i])) return false;
      if (prim == 0) {
        entry[i] = 0.0f;
      } else {
        float e = prev[i];
        if (i > 0) e = std::min(e, prev[i - 1]);
        if (i + 1 < span) e = std::min(e, prev[i + 1]);
        entry[i] = e;
      }
    }
    d[0] = entry[0] + cost[0];
    for (size_t i = 1; i != span; ++i) d[i] = std::min(entry[i], d[i - 1]) + cost[i];
    back[span - 1] = entry[span - 1] + cost[span - 1];
    for (size_t i = span - 1; i-- != 0;) back[i] = std::min(entry[i], back[i + 1]) + cost[i];
    for (size_t i = 0; i != span; ++i) d[i] = std::min(d[i], back[i]);
  }
  // the exit: the cheapest cell of the last row
  const float* last = &dist[(n_prim - 1) * span];
  size_t lat = 0, ties = 0;
  float best = inf;
  for (size_t i = 0; i != span; ++i) {
    if (last[i] < best) {
      best = last[i];
      lat = i;
      ties = 1;
    } else if (last[i] == best) {
      ++ties;
    }
  }
  if (ties != 1 || !std::isfinite(best)) return false;
  std::vector<uint32_t> path_lat(n_prim + span * 2);
  std::vector<uint32_t> path_prim(path_lat.size());
  size_t n_path = 0, prim = n_prim - 1;
  while (prim > 0) {
    if (n_path == path_lat.size()) {
      path_lat.resize(n_path * 2);
      path_prim.resize(n_path * 2);
    }
    path_lat[n_path] = uint32_t(lat);
    path_prim[n_path++] = uint32_t(prim);
    const float* d = &dist[prim * span];
    const float* prev = d - span;
    // the five cells that can step here; the cheapest of them did, if it is alone
    const size_t cand_lat[5] = {lat, lat - 1, lat + 1, lat - 1, lat + 1};
    const bool cand_prev[5] = {true, true, true, false, false};
    float low = inf;
    size_t low_k = 0, low_n = 0;
    for (size_t k = 0; k != 5; ++k) {
      const size_t cl = cand_lat[k];
      if (cl >= span) continue;  // also catches lat - 1 below zero
      const float v = cand_prev[k] ? prev[cl] : d[cl];
      if (v < low) {
        low = v;
        low_k = k;
        low_n = 1;
      } else if (v == low) {
        ++low_n;
      }
    }
    if (low_n != 1) return false;
    lat = cand_lat[low_k];
    if (cand_prev[low_k]) --prim;
  }
  FillStrip(output, at, lo, hi, n_prim, 0.0f);
  for (size_t k = 0; k != n_path; ++k) output[at(lo + path_lat[k], path_prim[k])] = 1.0f;
  output[at(lo + lat, 0)] = 1.0f;
  return true;
}

}  // namespace

namespace {

// The reference's walk (dijkstra_splitter.cc:34-121 / 123-205) with the same queue discipline:
// std::priority_queue is push_heap / pop_heap on a vector, and the comparator only looks at the
// distance, so a heap of compact steps goes through exactly the same sequence of moves and
// breaks equal-distance ties exactly like the reference's queue of (distance, x, y, prev_x,
// prev_y) records. A step is 8 bytes here: the distance, and the target cell of the strip
// (prim * span + lat) with the direction it was entered from in the low 3 bits. Costs and
// distances live in dense strip-local arrays, so the horizontal search does not stride through
// the frame.
struct CompactStep {
  float distance;
  uint32_t code;
  bool operator<(const CompactStep& rhs) const { return distance > rhs.distance; }  // min-heap
};

// where a step came from, relative to its target (lat, prim)
enum From : uint32_t {
  kFromHigherLatPrevPrim = 0,  // pushed as (s.lat - 1, s.prim + 1)
  kFromHigherLatSamePrim = 1,  // pushed as (s.lat - 1, s.prim)
  kFromSameLatPrevPrim = 2,    // pushed as (s.lat, s.prim + 1)
  kFromLowerLatPrevPrim = 3,   // pushed as (s.lat + 1, s.prim + 1)
  kFromLowerLatSamePrim = 4,   // pushed as (s.lat + 1, s.prim)
  kFromItself = 5              // the seeds of the first row
};

void StepBack(uint32_t from, size_t& lat, size_t& prim) {
  switch (from) {
    case kFromHigherLatPrevPrim: ++lat; --prim; break;
    case kFromHigherLatSamePrim: ++lat; break;
    case kFromSameLatPrevPrim: --prim; break;
    case kFromLowerLatPrevPrim: --lat; --prim; break;
    case kFromLowerLatSamePrim: --lat; break;
    default: break;
  }
}

void QueueDividingPath(const float* image, float* output, const Axes& at, size_t lo, size_t hi,
                       size_t n_prim) {
  const size_t span = hi - lo;
  if (span * (n_prim + 1) >= (size_t(1) << 29))
    throw std::invalid_argument("divider strip too large");
  std::vector<float> cost(span * n_prim), dist(span * n_prim, FLT_MAX);
  std::vector<uint8_t> back(span * n_prim, uint8_t(kFromItself));
  for (size_t prim = 0; prim != n_prim; ++prim)
    for (size_t i = 0; i != span; ++i) cost[prim * span + i] = std::fabs(image[at(lo + i, prim)]);
  std::vector<CompactStep> heap;
  heap.reserve(span * 64);
  auto push = [&heap](float distance, size_t cell, uint32_t from) {
    heap.push_back(CompactStep{distance, uint32_t(cell << 3) | from});
    std::push_heap(heap.begin(), heap.end());
  };
  for (size_t i = 0; i != span; ++i) push(0.0f, i, kFromItself);
  size_t end_lat = 0, end_prim = 0;
  while (!heap.empty()) {
    const CompactStep s = heap.front();
    std::pop_heap(heap.begin(), heap.end());
    heap.pop_back();
    const size_t cell = s.code >> 3, prim = cell / span, lat = cell - prim * span;
    if (prim == n_prim) {  // first path to leave the far edge is the cheapest
      end_lat = lat;
      end_prim = prim;
      StepBack(s.code & 7u, end_lat, end_prim);
      break;
    }
    const float reached = s.distance + cost[cell];
    if (reached < dist[cell]) {
      dist[cell] = reached;
      back[cell] = uint8_t(s.code & 7u);
      if (lat > 0) {
        push(reached, cell + span - 1, kFromHigherLatPrevPrim);
        push(reached, cell - 1, kFromHigherLatSamePrim);
      }
      push(reached, cell + span, kFromSameLatPrevPrim);
      if (lat + 1 < span) {
        push(reached, cell + span + 1, kFromLowerLatPrevPrim);
        push(reached, cell + 1, kFromLowerLatSamePrim);
      }
    }
  }
  FillStrip(output, at, lo, hi, n_prim, 0.0f);
  size_t lat = end_lat, prim = end_prim;
  while (prim > 0) {
    output[at(lo + lat, prim)] = 1.0f;
    StepBack(back[prim * span + lat], lat, prim);
  }
  output[at(lo + lat, 0)] = 1.0f;
}

}  // namespace

void DividingPath(const float* image, float* output, size_t width, size_t height, size_t lo,
                  size_t hi, bool vertical) {
  if (hi <= lo) throw std::invalid_argument("empty divider strip");
  const Axes at{width, vertical};
  const size_t n_prim = vertical ? height : width;
  if (UniqueDividingPath(image, output, at, lo, hi, n_prim)) return;
  QueueDividingPath(image, output, at, lo, hi, n_prim);
}

void FloodArea(const float* dividers, size_t width, size_t height, size_t inside, bool vertical,
               uint8_t* mask, size_t& start, size_t& extent) {
  const Axes at{width, vertical};
  const long n_lat = long(vertical ? width : height);
  const size_t n_prim = vertical ? height : width;
  std::fill(mask, mask + width * height, uint8_t(0));
  start = size_t(n_lat);
  size_t end = 0;
  for (size_t prim = 0; prim != n_prim; ++prim) {
    long lat = long(inside);
    // towards lower coordinates: the free area, then through the divider that bounds it
    while (lat >= 0 && dividers[at(size_t(lat), prim)] == 0.0f) mask[at(size_t(lat--), prim)] = 1;
    while (lat >= 0 && dividers[at(size_t(lat), prim)] != 0.0f) mask[at(size_t(lat--), prim)] = 1;
    start = std::min(start, size_t(lat + 1));
    // towards higher coordinates: up to (not including) the next divider
    lat = long(inside) + 1;
    while (lat < n_lat && dividers[at(size_t(lat), prim)] == 0.0f) mask[at(size_t(lat++), prim)] = 1;
    end = std::max(end, size_t(lat));
  }
  extent = end < start ? 0 : end - start;
}

void BoundingMask(const uint8_t* vertical_mask, size_t vertical_mask_x, size_t vertical_mask_width,
                  const uint8_t* horizontal_mask, uint8_t* mask, size_t width, size_t height,
                  size_t& sub_x, size_t& sub_y, size_t& sub_width, size_t& sub_height) {
  size_t x_lo = vertical_mask_x + vertical_mask_width, y_lo = height, x_hi = 0, y_hi = 0;
  // like the reference, pixels of `mask` outside the vertical area's columns are left as they are
  for (size_t y = 0; y != height; ++y) {
    for (size_t x = 0; x != vertical_mask_width; ++x) {
      const size_t fx = x + vertical_mask_x;
      const bool in = vertical_mask[y * vertical_mask_width + x] && horizontal_mask[y * width + fx];
      mask[y * width + fx] = in ? 1 : 0;
      if (in) {
        x_lo = std::min(x_lo, fx);
        y_lo = std::min(y_lo, y);
        x_hi = std::max(x_hi, fx);
        y_hi = y;
      }
    }
  }
  sub_x = x_lo;
  sub_y = y_lo;
  if (x_hi < x_lo) {
    sub_width = sub_height = 0;
  } else {
    sub_width = x_hi + 1 - x_lo;
    sub_height = y_hi + 1 - y_lo;
  }
  // even frames keep even facets: grow by one masked-out column / row
  if (width % 2 == 0 && sub_width % 2 != 0) {
    ++sub_width;
    size_t clear_x = sub_x + sub_width - 1;
    if (sub_width + sub_x >= width) clear_x = --sub_x;
    for (size_t y = sub_y; y != sub_y + sub_height; ++y) mask[clear_x + y * width] = 0;
  }
  if (height % 2 == 0 && sub_height % 2 != 0) {
    ++sub_height;
    size_t clear_y = sub_y + sub_height - 1;
    if (sub_height + sub_y >= height) clear_y = --sub_y;
    std::fill_n(mask + clear_y * width + sub_x, sub_width, uint8_t(0));
  }
}

namespace {
// aocommon::Image::TrimBox for byte masks
void CropBox(uint8_t* dst, size_t x, size_t y, size_t w, size_t h, const uint8_t* src,
             size_t src_width) {
  for (size_t row = 0; row != h; ++row)
    std::copy_n(src + (y + row) * src_width + x, w, dst + row * w);
}

// f(i) for i in [begin, end) on up to hardware_concurrency threads
template <typename F>
void RunOnThreads(size_t begin, size_t end, F&& f) {
  if (end <= begin) return;
  const size_t n_workers =
      std::min<size_t>(end - begin, std::max(1u, std::thread::hardware_concurrency()));
  std::atomic<size_t> next{begin};
  std::exception_ptr failure;
  std::mutex failure_lock;
  std::vector<std::thread> pool;
  for (size_t w = 0; w != n_workers; ++w)
    pool.emplace_back([&] {
      try {
        for (size_t i = next.fetch_add(1); i < end; i = next.fetch_add(1)) f(i);
      } catch (...) {
        std::lock_guard<std::mutex> guard(failure_lock);
        if (!failure) failure = std::current_exception();
        next.store(end);
      }
    });
  for (std::thread& t : pool) t.join();
  if (failure) std::rethrow_exception(failure);
}
}  // namespace

namespace {

// FloodArea without the frame-sized mask: along every line of the primary axis the area between
// two dividers is one run of pixels [lo, hi), so two numbers per line describe it.
struct AreaRuns {
  std::vector<uint32_t> lo, hi;  // per primary coordinate
  size_t start = 0, extent = 0;  // lateral range covered by any run
};

AreaRuns FloodRuns(const float* dividers, size_t width, size_t height, size_t inside,
                   bool vertical) {
  const Axes at{width, vertical};
  const long n_lat = long(vertical ? width : height);
  const size_t n_prim = vertical ? height : width;
  AreaRuns runs;
  runs.lo.resize(n_prim);
  runs.hi.resize(n_prim);
  size_t start = size_t(n_lat), end = 0;
  for (size_t prim = 0; prim != n_prim; ++prim) {
    long lat = long(inside);
    while (lat >= 0 && dividers[at(size_t(lat), prim)] == 0.0f) --lat;
    while (lat >= 0 && dividers[at(size_t(lat), prim)] != 0.0f) --lat;
    runs.lo[prim] = uint32_t(lat + 1);
    lat = long(inside) + 1;
    while (lat < n_lat && dividers[at(size_t(lat), prim)] == 0.0f) ++lat;
    runs.hi[prim] = uint32_t(lat);
    start = std::min(start, size_t(runs.lo[prim]));
    end = std::max(end, size_t(runs.hi[prim]));
  }
  runs.start = start;
  runs.extent = end < start ? 0 : end - start;
  return runs;
}

// GetBoundingMask for the facet where a column area and a row area cross, straight into the
// facet's own mask: the same box (grown to even sizes by one cleared column / row) and the same
// pixels as BoundingMask + TrimBox on frame-sized masks.
void FacetFromRuns(const AreaRuns& column, const AreaRuns& row, size_t width, size_t height,
                   Facet& f) {
  auto in = [&](size_t x, size_t y) {
    return x >= column.lo[y] && x < column.hi[y] && y >= row.lo[x] && y < row.hi[x];
  };
  const size_t cx0 = column.start, cx1 = column.start + column.extent;
  const size_t ry0 = row.start, ry1 = std::min(height, row.start + row.extent);
  size_t x_lo = cx1, y_lo = height, x_hi = 0, y_hi = 0;
  for (size_t y = ry0; y < ry1; ++y) {
    const size_t xa = std::max<size_t>(cx0, column.lo[y]), xb = std::min<size_t>(cx1, column.hi[y]);
    for (size_t x = xa; x < xb; ++x)
      if (y >= row.lo[x] && y < row.hi[x]) {
        x_lo = std::min(x_lo, x);
        x_hi = std::max(x_hi, x);
        y_lo = std::min(y_lo, y);
        y_hi = y;
      }
  }
  size_t sub_x = x_lo, sub_y = y_lo, sub_w = 0, sub_h = 0;
  if (x_hi >= x_lo) {
    sub_w = x_hi + 1 - x_lo;
    sub_h = y_hi + 1 - y_lo;
  }
  // even frames keep even facets: one more column / row, cleared (see BoundingMask)
  long clear_x = -1, clear_y = -1;
  size_t clear_x_rows_y = 0, clear_x_rows_h = 0;
  if (width % 2 == 0 && sub_w % 2 != 0) {
    ++sub_w;
    clear_x = long(sub_x + sub_w - 1);
    if (sub_w + sub_x >= width) clear_x = long(--sub_x);
    clear_x_rows_y = sub_y;
    clear_x_rows_h = sub_h;
  }
  if (height % 2 == 0 && sub_h % 2 != 0) {
    ++sub_h;
    clear_y = long(sub_y + sub_h - 1);
    if (sub_h + sub_y >= height) clear_y = long(--sub_y);
  }
  f.x = sub_x;
  f.y = sub_y;
  f.width = sub_w;
  f.height = sub_h;
  f.mask.assign(sub_w * sub_h, 0);
  for (size_t y = sub_y; y != sub_y + sub_h; ++y) {
    if (long(y) == clear_y) continue;
    uint8_t* out = &f.mask[(y - sub_y) * sub_w];
    for (size_t x = sub_x; x != sub_x + sub_w; ++x) {
      if (long(x) == clear_x && y >= clear_x_rows_y && y < clear_x_rows_y + clear_x_rows_h) continue;
      // BoundingMask only writes the columns of the column area; the rest of the box is one of
      // the cleared lines above
      if (x >= cx0 && x < cx1) out[x - sub_x] = in(x, y) ? 1 : 0;
    }
  }
}

}  // namespace

std::vector<Facet> MakeFacets(const float* image, size_t width, size_t height,
                              const uint8_t* user_mask, size_t grid_width, size_t grid_height,
                              const DividerSharing* sharing) {
  if (grid_width == 0 || grid_height == 0 || grid_width > width || grid_height > height)
    throw std::invalid_argument("bad facet grid");
  const size_t avg_w = width / grid_width, avg_h = height / grid_height;
  // The reference computes the vertical dividers, floods the columns, zeroes the divider image
  // and then computes the horizontal ones. The searches only read `image` and write disjoint
  // strips, so all of them run at once here, into one image per orientation.
  std::vector<float> dividers(width * height, 0.0f), dividers_h(width * height, 0.0f);
  const size_t n_vertical = grid_width - 1, n_dividers = n_vertical + grid_height - 1;
  const bool shared = sharing && sharing->world > 1;
  RunOnThreads(0, n_dividers, [&](size_t k) {
    if (shared && int(k % size_t(sharing->world)) != sharing->rank) return;
    if (k < n_vertical) {
      // each divider is searched inside the middle half of the strip between two facets
      const size_t mid = width * (k + 1) / grid_width;
      DividingPath(image, dividers.data(), width, height, mid - avg_w / 4, mid + avg_w / 4, true);
    } else {
      const size_t mid = height * (k - n_vertical + 1) / grid_height;
      DividingPath(image, dividers_h.data(), width, height, mid - avg_h / 4, mid + avg_h / 4,
                   false);
    }
  });
  if (shared) sharing->merge(dividers.data(), dividers_h.data(), width * height);
  // FloodVerticalArea / FloodHorizontalArea for every column / row of facets, as runs per line
  std::vector<AreaRuns> columns(grid_width), rows(grid_height);
  RunOnThreads(0, grid_width + grid_height, [&](size_t k) {
    if (k < grid_width)
      columns[k] = FloodRuns(dividers.data(), width, height, k * width / grid_width + avg_w / 2,
                             true);
    else
      rows[k - grid_width] = FloodRuns(dividers_h.data(), width, height,
                                       (k - grid_width) * height / grid_height + avg_h / 2, false);
  });
  std::vector<float>().swap(dividers);
  std::vector<float>().swap(dividers_h);

  std::vector<Facet> facets(grid_width * grid_height);
  RunOnThreads(0, facets.size(), [&](size_t index) {
    Facet& f = facets[index];
    f.index = index;
    FacetFromRuns(columns[index % grid_width], rows[index / grid_width], width, height, f);
    f.boundary_mask = f.mask;
    if (user_mask) {
      for (size_t row = 0; row != f.height; ++row)
        for (size_t col = 0; col != f.width; ++col)
          f.mask[row * f.width + col] &= user_mask[(f.y + row) * width + f.x + col] ? 1 : 0;
    }
  });
  return facets;
}

}  // namespace rdl
