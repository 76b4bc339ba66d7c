// pyi_windowed.cu -- K4 translation unit: windowed 2-D interpolation (generic
// kernel pyi_windowed.cuh, 6 x 6 bicubic fast path pyi_fastwin.cuh).
#include "pyi_common.cuh"

namespace pyi {
// Debug knob for the parity tests: PYINTERP_B200_GENERIC=1 routes windowed calls
// through the generic kernel so the fast path can be compared with it.
static bool force_generic() {
  const char* e = getenv("PYINTERP_B200_GENERIC");
  return e != nullptr && e[0] == '1';
}
}  // namespace pyi
#include "pyi_windowed.cuh"
#include "pyi_fastwin.cuh"
