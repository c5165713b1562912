// ORACLE — TEST INFRASTRUCTURE ONLY.
// Stand-in for aocommon/uvector.h: an std::vector-compatible container (the real one
// only skips value-initialisation). Used by cpp/math/dijkstra_splitter.cc:46,101.
#ifndef RDL_REF_SHIM_UVECTOR_H_
#define RDL_REF_SHIM_UVECTOR_H_
#include <vector>
namespace aocommon {
template <typename T>
using UVector = std::vector<T>;
}
#endif
