// Team-kernel instantiations of one (environment kind, robot) pair; see dm_team_launch.cuh.
#include <type_traits>
#define DM_TEAM_LAUNCH_IMPL
#include "dm_team_launch.cuh"

namespace dm {
template <>
cudaError_t launch_team<KIND_SINGLE, Robot_unitree_g1>(const TeamLaunch& c, const StepParams& q, const TeamExtra& ex) {
  return launch_team_impl<KIND_SINGLE, Robot_unitree_g1>(c, q, ex);
}
}  // namespace dm
