// Host-side launcher of the team kernel (K1c), one explicit specialisation per (environment kind, robot), each in its own
// translation unit (dm_team_h.cu / dm_team_g.cu / dm_team_c.cu) so they compile side by side.  dmhot.cu only sees the
// declaration.
#pragma once
#include "dm_team_kernel.cuh"

namespace dm {

struct TeamLaunch {
  int grid, teams;     // CTAs, teams of 128 threads per CTA
  int maxt;            // which instantiation: compiled for at most `maxt` teams per CTA (register budget 65536 / (128 maxt))
  size_t smem;         // dynamic shared memory
  cudaStream_t stream;
  bool pdl;            // programmatic dependent launch
  bool terms, stats;   // kernel template flags
  bool table_smem;     // the packed clip table is staged in shared memory
};

constexpr int kTeamMaxtWide = 4, kTeamMaxtMid = 6, kTeamMaxtFull = 8;

template <int KIND, class Rb>
cudaError_t launch_team(const TeamLaunch& c, const StepParams& q, const TeamExtra& ex);

#ifdef DM_TEAM_LAUNCH_IMPL
template <int KIND, class Rb>
static cudaError_t launch_team_impl(const TeamLaunch& c, const StepParams& q, const TeamExtra& ex) {
  auto launch = [&](auto kern) -> cudaError_t {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(c.grid);
    cfg.blockDim = dim3(c.teams * 128);
    cfg.dynamicSmemBytes = c.smem;
    cfg.stream = c.stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = c.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, q, ex);
  };
  auto pick = [&](auto terms, auto stats, auto table) -> cudaError_t {
    constexpr bool T = decltype(terms)::value, S = decltype(stats)::value, B = decltype(table)::value;
    if constexpr (!T) {  // the reward terms are an N = 1 / debugging output: one shape is enough
      if (c.maxt == kTeamMaxtWide) return launch(step_team_kernel<KIND, Rb, T, S, B, kTeamMaxtWide>);
      if (c.maxt == kTeamMaxtMid) return launch(step_team_kernel<KIND, Rb, T, S, B, kTeamMaxtMid>);
    }
    return launch(step_team_kernel<KIND, Rb, T, S, B, kTeamMaxtFull>);
  };
  using std::false_type;
  using std::true_type;
  auto pick2 = [&](auto terms, auto stats) -> cudaError_t {
    return c.table_smem ? pick(terms, stats, true_type{}) : pick(terms, stats, false_type{});
  };
  return c.terms ? (c.stats ? pick2(true_type{}, true_type{}) : pick2(true_type{}, false_type{}))
                 : (c.stats ? pick2(false_type{}, true_type{}) : pick2(false_type{}, false_type{}));
}
#endif

}  // namespace dm
