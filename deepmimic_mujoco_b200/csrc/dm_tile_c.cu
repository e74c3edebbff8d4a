// Tile-kernel instantiations of one (environment kind, robot) pair; see dm_tile_launch.cuh.
#include <type_traits>
#define DM_TILE_LAUNCH_IMPL
#include "dm_tile_launch.cuh"

namespace dm {
template <>
cudaError_t launch_tile<KIND_COMBINED, Robot_unitree_g1>(const TileLaunch& c, const StepParams& q, const TileExtra& ex) {
  return launch_tile_impl<KIND_COMBINED, Robot_unitree_g1>(c, q, ex);
}
}  // namespace dm
