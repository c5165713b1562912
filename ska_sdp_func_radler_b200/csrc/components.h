// The clean-component list of an algorithm (cpp/component_list.h): per scale, the positions and
// per-image values of every component, kept on the host between major iterations.
#ifndef RDL_COMPONENTS_H_
#define RDL_COMPONENTS_H_

#include <cstddef>
#include <cstdint>
#include <vector>

namespace rdl {

struct ComponentScaleList {
  std::vector<uint32_t> x, y;
  std::vector<float> values;  // n_values per component
};

struct ComponentList {
  size_t width = 0, height = 0, n_values = 0;  // n_values = images of the cube
  std::vector<ComponentScaleList> scales;
  bool merged = true;

  bool Empty() const { return scales.empty(); }
  // ComponentList::Add (component_list.h:70-77)
  void Add(size_t x, size_t y, size_t scale_index, const float* values);
  // ComponentList::MergeDuplicates (component_list.h:113-131, 222-262): the values of components
  // at one position are summed in the order they were added; positions come out grouped by the
  // first image whose sum is non-zero, row-major inside a group; all-zero positions are dropped.
  void MergeDuplicates();
};

}  // namespace rdl
#endif
