// SoA-kernel instantiations of one (environment kind, robot) pair; see dm_soa_kernel.cuh.
#include <type_traits>
#define DM_SOA_LAUNCH_IMPL
#include "dm_soa_kernel.cuh"

namespace dm {
template <>
cudaError_t launch_soa<KIND_SINGLE, Robot_unitree_g1>(const SoaLaunch& c, const StepParams& q, const SoaExtra& ex) {
  return launch_soa_impl<KIND_SINGLE, Robot_unitree_g1>(c, q, ex);
}
}  // namespace dm
