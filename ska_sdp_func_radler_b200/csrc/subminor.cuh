// Sub-minor (Clark-like) loop on the device (algorithms/subminor_loop.{h,cc}).
#ifndef RDL_SUBMINOR_CUH_
#define RDL_SUBMINOR_CUH_

#include "handle.h"

namespace rdl {
rdl_result RunGenericCleanSubMinor(rdl_handle& h, DeviceCube& cube, cudaStream_t stream);
}  // namespace rdl
#endif
