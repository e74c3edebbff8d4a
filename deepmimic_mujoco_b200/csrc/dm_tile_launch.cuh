// Host-side launcher of the tile kernel (K1b), one explicit specialisation per (environment kind, robot), each in its own
// translation unit (dm_tile_h.cu / dm_tile_g.cu / dm_tile_c.cu) so that the ~14 kernel instantiations of a robot compile
// side by side with the other robots' instead of one after another in dmhot.cu.  dmhot.cu only sees the declaration.
#pragma once
#include "dm_tile_kernel.cuh"

namespace dm {

struct TileLaunch {
  int grid, groups;    // CTAs, groups of TE threads per CTA
  size_t smem;         // dynamic shared memory
  cudaStream_t stream;
  bool pdl;            // programmatic dependent launch
  bool terms, stats;   // kernel template flags
  int table;           // 0 = rows in global memory, 1 = whole table in shared memory, 2 = per-environment rows staged
  int shape;           // 0 = KernelGroups::G groups (register cap), 1 = GWIDE, 2 = GMID
};

template <int KIND, class Rb>
cudaError_t launch_tile(const TileLaunch& c, const StepParams& q, const TileExtra& ex);

#ifdef DM_TILE_LAUNCH_IMPL
template <int KIND, class Rb>
static cudaError_t launch_tile_impl(const TileLaunch& c, const StepParams& q, const TileExtra& ex) {
  using TC = TileCfg<KIND, Rb>;
  auto launch = [&](auto kern) -> cudaError_t {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem);
    if (e != cudaSuccess) return e;
    // programmatic dependent launch: this kernel's launch latency and prologue (barriers, clip table, motor constants)
    // overlap the tail of whatever runs before it in the stream; its griddepcontrol.wait orders everything else
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(c.grid);
    cfg.blockDim = dim3(c.groups * TC::TE);
    cfg.dynamicSmemBytes = c.smem;
    cfg.stream = c.stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = c.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, q, ex);
  };
  constexpr int GMAX = KernelGroups<Rb>::G;
  auto pick = [&](auto terms, auto stats) -> cudaError_t {
    constexpr bool T = decltype(terms)::value, S = decltype(stats)::value;
    if constexpr (!T) {  // the uncapped shapes exist without reward terms only (terms are an N = 1 / debugging output)
      constexpr int GW = KernelGroups<Rb>::GWIDE;
      if (c.shape == 1 && c.table == 2) return launch(step_tile_kernel<KIND, Rb, T, S, 2, GW>);
      if (c.shape == 1) return c.table == 1 ? launch(step_tile_kernel<KIND, Rb, T, S, 1, GW>) : launch(step_tile_kernel<KIND, Rb, T, S, 0, GW>);
      if constexpr (KernelGroups<Rb>::GMID > 0) {
        constexpr int GM = KernelGroups<Rb>::GMID;
        if (c.shape == 2) return c.table == 1 ? launch(step_tile_kernel<KIND, Rb, T, S, 1, GM>) : launch(step_tile_kernel<KIND, Rb, T, S, 0, GM>);
      }
    }
    if (c.shape != 0 || c.table == 2) return cudaErrorInvalidValue;
    return c.table == 1 ? launch(step_tile_kernel<KIND, Rb, T, S, 1, GMAX>) : launch(step_tile_kernel<KIND, Rb, T, S, 0, GMAX>);
  };
  using std::true_type;
  using std::false_type;
  return c.terms ? (c.stats ? pick(true_type{}, true_type{}) : pick(true_type{}, false_type{}))
                 : (c.stats ? pick(false_type{}, true_type{}) : pick(false_type{}, false_type{}));
}
#endif

}  // namespace dm
