// psim_cells.cuh — cell-binned force path.  Placeholder
// until the binned kernels land: selecting it reports PSIM_ERR_UNSUPPORTED.
#pragma once

#include "psim_common.cuh"

namespace psim {

template <class R>
struct CellPath {
  int64_t launches = 0;
  void invalidate() {}
  void destroy() {}
  template <class S>
  psim_status force(S&, int) { return PSIM_ERR_UNSUPPORTED; }
};

}  // namespace psim
