// K1s: the step kernel for STRUCTURE-OF-ARRAYS batches -- qpos [nq, ld], qvel [nv, ld], action [na, ld], obs [D, ld],
// torque [nu, ld]: one coalesced column per quantity, the layout batched simulators keep (north_star).
//
// Same shape as K1b (persistent CTAs, G groups of 128 threads, tiles of 128 consecutive environments, TMA in / TMA out,
// thread-per-environment env_step out of shared memory); what changes is the tile's geometry.  A tile of a SoA array is
// 128 consecutive floats of EVERY row: a [rows x 128] box of a 2-D tensor.  It arrives as TWO tensor-map bulk copies
// (cp.async.bulk.tensor.2d: the qpos box and the qvel box; the action rows are read directly by the torque pass) and
// leaves as three (the two joint blocks and the computed rows of the observation).  The first version moved every row
// with its own 512-byte 1-D bulk copy -- 69 loads + 67 stores per tile -- and was bound by TMA issue, not by bandwidth:
// 277 us for 1 M humanoid3d environments, DRAM 34 % busy (profiles/r02_soa_kernel.md).  The tile lives in shared
// memory as [row][128]: thread e reads element j of its environment at s[j * 128 + e] -- consecutive threads, consecutive
// words, no bank conflicts.  That geometry makes the observation tile nearly free:
//   obs rows [0, NJ)      ARE the staged qpos rows 7.. -- stored from where the load put them, no copy;
//   obs rows [NJ, 2 NJ)   are the staged qvel rows 6.. scaled by 0.1 in place by the thread that owns the column;
//   the computed rows     (torso, contacts, phase, player action) go to a small [NCOMP][128] block as they are computed
//                         (no registers held, unlike K1b);
// and every row leaves as one 512-byte bulk store.  Per-environment torques are written straight to global memory: for a
// fixed actuator the 128 threads of a group write 512 consecutive bytes.  Statistics: a warp per row, 128-bit shared-memory
// loads, shuffle reduction, fp64 running sums in shared memory.
// Batches whose size is not a multiple of 128, or whose leading dimension / base pointers are not 16-byte aligned, hand
// the remainder (or everything) to the strided kernel K1a.
#pragma once
#include <cuda.h>  // CUtensorMap (types only; the encoder is fetched through cudaGetDriverEntryPoint)

#include "dm_tile_kernel.cuh"

namespace dm {

template <int KIND, class Rb>
struct SoaCfg {
  using OL = ObsLayout<KIND, Rb>;
  static constexpr int TE = 128, NW = TE / 32;
  static constexpr int NCOMP = OL::DIM - 2 * Rb::NJ;         // computed observation rows
  static constexpr int IN_ROWS = Rb::NQ + Rb::NV;            // staged rows of one input tile (actions are read directly)
  static constexpr int SLOT_ROWS = IN_ROWS + NCOMP;          // + computed rows (the tile's rewards reuse qpos row 0, dead by then)
  static constexpr int SLOT_FLOATS = SLOT_ROWS * TE;
  static constexpr size_t SLOT_BYTES = sizeof(float) * SLOT_FLOATS;
  static constexpr int NCOLS = OL::DIM + 1;                  // statistics columns: observation entries, then the reward
  static constexpr size_t ACC_BYTES = sizeof(double) * (2 * NCOLS + 4);  // per CTA (shared-memory fp64 atomics)
  static constexpr int MAXG = 4;
  static constexpr size_t FIXED_BYTES = 128;                 // mbarriers: [0, MAXG) slots, [MAXG] table
  static constexpr size_t SMEM_BUDGET = 227 * 1024;
  static_assert(SLOT_BYTES % 128 == 0, "alignment");
};

struct SoaExtra {
  double* acc;       // statistics accumulators or NULL
  int n_tiles;       // full tiles in this launch
  int table_floats;  // packed clip table size (floats); staged in shared memory when TABLE_SMEM
  int table_pad;     // ... rounded up to a multiple of 32 floats: tensor-map copies want 128-byte aligned shared memory
  int prefetch;      // L2 prefetch of the group's next tile (DM_PREFETCH_NEXT=0 switches it off)
};

// tensor maps of one launch (encoded on the host, dmhot.cu): qpos [nq, ld] box {128, nq}; qvel box {128, nv};
// obs [D, ld] with a box of NJ rows (the two joint blocks) and one of NCOMP rows (the computed entries)
struct alignas(64) SoaMaps {
  CUtensorMap q, v, oj, oc;
};

__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap* tm, int c0, int c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global [%0, {%1, %2}];" ::"l"(tm), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* tm, int c0, int c1, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                   smem_u32(dst)),
               "l"(tm), "r"(c0), "r"(c1), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, int c0, int c1, const void* src) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(tm), "r"(c0), "r"(c1),
               "r"(smem_u32(src))
               : "memory");
}

// per-thread accessor over the [row][128] tile: inputs from the staged rows, computed observation entries to their block
template <int KIND, class Rb>
struct SoaIO {
  static constexpr bool kZeroOnGuard = false;  // the kernel zeroes the column when the guard fires
  static constexpr int TE = SoaCfg<KIND, Rb>::TE;
  const float* __restrict__ qp;  // &tile[qpos row 0][e]
  const float* __restrict__ qv;
  float* __restrict__ comp;      // &tile[computed row 0][e]
  __device__ __forceinline__ float qpos(int j) const { return qp[j * TE]; }
  __device__ __forceinline__ float qvel(int j) const { return qv[j * TE]; }
  __device__ __forceinline__ float act(int) const { return 0.0f; }
  __device__ __forceinline__ void obs(int j, float v) {  // j is a compile-time constant after unrolling
    if (j >= 2 * Rb::NJ) comp[(j - 2 * Rb::NJ) * TE] = v;
  }
  __device__ __forceinline__ void tau(int, float) {}
};

// groups per CTA the kernel is compiled for (register budget = 65536 / (128 G) per thread); SoA needs fewer registers
// than K1b: no joint blocks or computed entries are held across the tile barrier
template <int KIND, class Rb> struct SoaGroups { static constexpr int G = 4; };
// measured at 1 M / 262 144 envs: unitree_g1 DPEnv 272 us with four groups (128 registers, no spills) against 287 with three;
// the combined env 300 / 94.5 us with four against 298 / 83.5 with three (its 57 KB slots leave the fourth group no slack)
template <> struct SoaGroups<KIND_COMBINED, Robot_unitree_g1> { static constexpr int G = 3; };
// (humanoid3d with three uncapped groups -- 149 registers, no spills -- was measured after the packed-arithmetic changes:
// better at 262144 envs, 46.3 vs 49.1 us, equal at 1 M, worse at 65536 / 131072 / 524288: 19.2 / 29.4 / 85.2 vs 15.5 / 27.5 / 82.9 us)

template <int KIND, class Rb, bool TERMS, bool STATS, bool TABLE_SMEM, int G>
__global__ void __launch_bounds__(128 * G, 1) step_soa_kernel(const __grid_constant__ StepParams p, const SoaExtra x,
                                                              const __grid_constant__ SoaMaps tm) {
  using OL = ObsLayout<KIND, Rb>;
  using SC = SoaCfg<KIND, Rb>;
  using RL = RowLayout<Rb>;
  constexpr int TE = SC::TE, NJ = Rb::NJ, DIM = OL::DIM, NCOMP = SC::NCOMP, NQ = Rb::NQ, NV = Rb::NV, NA = Rb::NA;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);
  float* s_table = reinterpret_cast<float*>(smem_raw + SC::FIXED_BYTES);
  const int tid = threadIdx.x, group = tid / TE, gt = tid % TE, lane = tid & 31, warp = gt >> 5;
  const int n_groups = blockDim.x / TE;
  float* slot = s_table + (TABLE_SMEM ? x.table_pad : 0) + (size_t)group * SC::SLOT_FLOATS;
  float* s_q = slot;                       // [NQ][TE]
  float* s_v = s_q + NQ * TE;              // [NV][TE]
  float* s_c = s_v + NV * TE;              // [NCOMP][TE] computed observation rows
  float* s_rw = s_q;                       // [TE] rewards of the tile (statistics): qpos row 0 (root x), dead after the step
  double* g_acc = reinterpret_cast<double*>(s_table + (TABLE_SMEM ? x.table_pad : 0) + (size_t)n_groups * SC::SLOT_FLOATS);  // CTA-wide
  const long long ld = p.es;
  const bool want_tau = p.torque != nullptr;
  // this group's tiles: (j * grid + cta) * n_groups + group
  const int t_base = blockIdx.x * n_groups + group, t_stride = gridDim.x * n_groups;
  const int nt = t_base < x.n_tiles ? (x.n_tiles - 1 - t_base) / t_stride + 1 : 0;
  uint64_t* bar = bars + group;
  auto prefetch = [&](int j) {  // one thread: tile j's boxes towards L2 (a hint; safe before griddepcontrol.wait)
    const int e0 = (t_base + j * t_stride) * TE;
    tma_prefetch_2d(&tm.q, e0, 0);
    tma_prefetch_2d(&tm.v, e0, 0);
  };
  auto issue_load = [&](int j) {  // one thread: the qpos box and the qvel box of tile j
    const int e0 = (t_base + j * t_stride) * TE;
    mbar_expect_tx(bar, sizeof(float) * TE * (NQ + NV));
    tma_load_2d(s_q, &tm.q, e0, 0, bar);
    tma_load_2d(s_v, &tm.v, e0, 0, bar);
  };

  if (tid == 0) {
    for (int g = 0; g <= SC::MAXG; ++g) mbar_init(bars + g, 1);
    mbar_fence_init();
  }
  if (STATS) {
    for (int k = tid; k < 2 * SC::NCOLS + 4; k += blockDim.x) g_acc[k] = 0.0;
  }
  __syncthreads();
  if (TABLE_SMEM && tid == 0) {
    mbar_expect_tx(bars + SC::MAXG, sizeof(float) * x.table_floats);
    tma_load_1d(s_table, p.cs.rows, sizeof(float) * x.table_floats, bars + SC::MAXG);
  }
  const float* rows_tab = TABLE_SMEM ? s_table : p.cs.rows;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (gt == 0 && nt > 0 && x.prefetch) prefetch(0);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (gt == 0 && nt > 0) issue_load(0);
  bool table_ready = !TABLE_SMEM;

  for (int j = 0; j < nt; ++j) {
    const long long e0 = (long long)(t_base + j * t_stride) * TE;
    const int env = int(e0) + gt;
    EnvState st;
    st.idx = ld_stream(p.idx + env);
    st.ep_len = ld_stream(p.ep_len + env);
    st.clip = p.clip ? ld_stream(p.clip + env) : 0;
    const uint32_t contact = p.contact ? ld_stream(p.contact + env) : 0u;
    const float ep_ret0 = p.ep_return ? ld_stream(p.ep_return + env) : 0.0f;
    // action -> torque while the tile's bulk copies are in flight: for a fixed actuator the group's threads read and
    // write 512 consecutive bytes of the action / torque rows (the actions are not staged: 14 KB per slot that the clip
    // table needs -- with them in the slot the table fell out of shared memory and every step gathered its row from L2)
    // (ncu has 14 % of the G1 kernel's stall samples on the first use of these loads; loading them one tile ahead into
    // registers moved the stall elsewhere -- G1 combined, 1 M envs 272 -> 276 us -- and spilled in the four-group kernels)
    if (want_tau) {
      float av[NA];
#pragma unroll
      for (int k = 0; k < NA; ++k) av[k] = ld_stream(p.action + k * ld + env);
#pragma unroll
      for (int k = 0; k < Rb::NU; ++k) {
        float t = 0.0f;
        if (k < NA) {
          float c = av[k] * Rb::kActScale;
          c = dm_max_nan(dm_min_nan(c, Rb::ctrl_hi(k)), Rb::ctrl_lo(k));  // mju_clip; NaN passes through
          t = Rb::gear(k) * c;
        }
        p.torque[k * ld + env] = t;
      }
    }
    if (!table_ready) {
      mbar_wait(bars + SC::MAXG, 0);
      table_ready = true;
    }
    mbar_wait(bar, j & 1);
    // the next tile crosses HBM -> L2 while this one is computed (262144 humanoid3d envs 59.5 -> 52.6 us; nothing at 1 M envs).
    // Prefetching the NA action rows of the next tile the same way (512 bytes each) was measured far slower: G1 combined,
    // 1 M envs 272 -> 334 us.
    if (gt == 0 && j + 1 < nt && x.prefetch) prefetch(j + 1);
    SoaIO<KIND, Rb> io{s_q + gt, s_v + gt, s_c + gt};
    StepOut so;
    env_step<KIND, Rb, SoaIO<KIND, Rb>, TERMS>(io, p.cs, rows_tab, st, contact, so);
    finish_env<KIND, Rb, TERMS, false>(p, env, st, so);
    const float ep_ret1 = ep_ret0 + so.raw_reward;
    if (p.ep_return) p.ep_return[env] = ep_ret1;
    // the observation tile in place: this thread's column of the qvel rows scaled, everything zeroed on the guard
    const bool guard = so.reason == R_OBS_BOUNDS;
#pragma unroll
    for (int k = 0; k < NJ; ++k) s_v[(6 + k) * TE + gt] *= 0.1f;
    if (guard) {  // the row leaves as `obs * 0.` (zeros; NaN where an entry was infinite, deepmimic_env.py:476); rare
#pragma unroll 1
      for (int k = 0; k < NJ; ++k) { s_q[(7 + k) * TE + gt] *= 0.0f; s_v[(6 + k) * TE + gt] *= 0.0f; }
#pragma unroll 1
      for (int k = 0; k < NCOMP; ++k) s_c[k * TE + gt] *= 0.0f;
    }
    if (STATS) {
      s_rw[gt] = so.reward;
      if (so.done) {  // finished episodes are rare: shared-memory fp64 atomics
        atomicAdd(g_acc + 2 * SC::NCOLS + 0, 1.0);
        atomicAdd(g_acc + 2 * SC::NCOLS + 1, double(ep_ret1));
        atomicAdd(g_acc + 2 * SC::NCOLS + 2, double(st.ep_len));
      }
    }
    fence_async_smem();
    group_sync(group);
    // the observation tile leaves as three tensor-map stores: qpos rows 7.. -> obs rows [0, NJ), the scaled qvel rows
    // 6.. -> [NJ, 2 NJ), the computed rows -> [2 NJ, DIM)
    auto obs_row = [&](int r) -> const float* {
      return r < NJ ? s_q + (7 + r) * TE : (r < 2 * NJ ? s_v + (6 + r - NJ) * TE : s_c + (r - 2 * NJ) * TE);
    };
    if (gt == 0) {
      tma_store_2d(&tm.oj, int(e0), 0, s_q + 7 * TE);
      tma_store_2d(&tm.oj, int(e0), NJ, s_v + 6 * TE);
      tma_store_2d(&tm.oc, int(e0), 2 * NJ, s_c);
      bulk_commit();
    }
    if (STATS) {
      // column sums.  Warp w owns rows w, w + 4, ...: every lane sums its four environments of each of those rows
      // (one 128-bit shared-memory load per row), then a transposing butterfly reduces all the rows at once -- 31
      // shuffles per quantity instead of 5 per row -- and leaves lane L with the total of the warp's L-th row, so the
      // 32 lanes add to the fp64 running sums in parallel.
      constexpr int NI = (SC::NCOLS + SC::NW - 1) / SC::NW;
      static_assert(NI <= 32, "a warp reduces at most 32 rows at once");
      float va[32], vb[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        va[i] = vb[i] = 0.0f;
        if (i < NI) {
          const int r = warp + SC::NW * i;
          if (r <= DIM) {
            const float4 v = *reinterpret_cast<const float4*>((r < DIM ? obs_row(r) : s_rw) + 4 * lane);
            va[i] = (v.x + v.y) + (v.z + v.w);
            vb[i] = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, v.w * v.w)));
          }
        }
      }
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) {
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < o; ++i) {
          const float sa = up ? va[i] : va[i + o], ka = up ? va[i + o] : va[i];
          const float sb = up ? vb[i] : vb[i + o], kb = up ? vb[i + o] : vb[i];
          va[i] = ka + __shfl_xor_sync(0xffffffffu, sa, o);
          vb[i] = kb + __shfl_xor_sync(0xffffffffu, sb, o);
        }
      }
      const int r = warp + SC::NW * lane;  // lane L holds row index L of this warp's list
      if (lane < NI && r <= DIM) {  // the groups of a CTA share one set of running sums
        atomicAdd(g_acc + r, double(va[0]));
        atomicAdd(g_acc + SC::NCOLS + r, double(vb[0]));
      }
    }
    if (gt == 0) bulk_wait_read0();  // the stores have read the slot
    group_sync(group);
    if (gt == 0 && j + 1 < nt) issue_load(j + 1);
  }

  if (gt == 0) bulk_wait_read0();
  if (STATS && x.acc) {  // one set of fp64 global atomics per CTA
    if (gt == 0) atomicAdd(g_acc + 2 * SC::NCOLS + 3, double(nt));
    __syncthreads();
    for (int k = tid; k < 2 * DIM + 6; k += blockDim.x) {
      double v;
      if (k == 0) v = g_acc[2 * SC::NCOLS + 3] * TE;          // count
      else if (k == 1) v = g_acc[DIM];                         // sum r
      else if (k == 2) v = g_acc[SC::NCOLS + DIM];             // sum r^2
      else if (k < 6) v = g_acc[2 * SC::NCOLS + k - 3];        // n_done, sum ep_return(done), sum ep_len(done)
      else if (k < 6 + DIM) v = g_acc[k - 6];
      else v = g_acc[SC::NCOLS + k - 6 - DIM];
      if (v != 0.0) atomicAdd(x.acc + k, v);
    }
  }
}

struct SoaLaunch {
  int grid, groups;
  size_t smem;
  cudaStream_t stream;
  bool pdl, terms, stats, table_smem;
  const SoaMaps* maps;
};

template <int KIND, class Rb>
cudaError_t launch_soa(const SoaLaunch& c, const StepParams& q, const SoaExtra& ex);

#ifdef DM_SOA_LAUNCH_IMPL
template <int KIND, class Rb>
static cudaError_t launch_soa_impl(const SoaLaunch& c, const StepParams& q, const SoaExtra& ex) {
  auto launch = [&](auto kern) -> cudaError_t {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)c.smem);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(c.grid);
    cfg.blockDim = dim3(c.groups * 128);
    cfg.dynamicSmemBytes = c.smem;
    cfg.stream = c.stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = c.pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, q, ex, *c.maps);
  };
  constexpr int GK = SoaGroups<KIND, Rb>::G;
  auto pick = [&](auto terms, auto stats) -> cudaError_t {
    constexpr bool T = decltype(terms)::value, S = decltype(stats)::value;
    return c.table_smem ? launch(step_soa_kernel<KIND, Rb, T, S, true, GK>) : launch(step_soa_kernel<KIND, Rb, T, S, false, GK>);
  };
  using std::false_type;
  using std::true_type;
  return c.terms ? (c.stats ? pick(true_type{}, true_type{}) : pick(true_type{}, false_type{}))
                 : (c.stats ? pick(false_type{}, true_type{}) : pick(false_type{}, false_type{}));
}
#endif

}  // namespace dm
