// pyi_windowed_nd.cu -- K6 for one window capacity (compiled three times with
// -DPYI_ND_MAXW=6 / 12 / 32; see pyi_windowed_nd.cuh).
#include "pyi_common.cuh"
#define PYI_WINDOWED_DEVICE_ONLY
#include "pyi_windowed.cuh"
#include "pyi_univariate.cuh"
#include "pyi_windowed_nd.cuh"

namespace pyi {
template cudaError_t launch_windowed_nd_w<PYI_ND_MAXW>(const WinNdParams&, cudaStream_t);
}  // namespace pyi
