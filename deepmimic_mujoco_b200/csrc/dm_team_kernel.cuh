// K1c: the team kernel -- one environment split over the four warps of a TEAM, 32 environments per team and tile.
//
// Why: thread-per-environment (K1b) runs ~3200 (humanoid3d) / ~5300 (unitree_g1) dependent-ish instructions on ONE
// thread at 128-240 registers, so an SM holds 8-16 warps and a small batch is bound by one thread's chain (8.7 us for
// 4096 environments), a large unitree_g1 batch by issue stalls (0.47-0.56 of the HBM roofline).  Here warp p of a team
// runs PART p of the kinematic tree (generated/parts_<robot>.cuh: trunk, arms, right leg, left leg / the two legs and
// the two arm + hand chains) for the team's 32 environments, lane = environment.  No intra-warp divergence (a warp runs
// one part), the critical path is one limb, a part needs a fraction of the registers, and the four warps of a team sit on
// the four SM sub-partitions, so every sub-partition's instruction stream is ONE part's code.
//
// Per team and tile (32 consecutive environments; row-major arrays, so every array's slice is one contiguous range):
//   in     three 1-D TMA bulk loads (qpos, qvel, action rows) into the team's input slot, completion on its mbarrier;
//   parts  every warp: forward kinematics (rotation matrices along the chain) of its bodies, the joint error sums,
//          limit counts, end effectors, floor contacts, its share of the two COM sums; its joints' observation entries
//          go straight into the team's output tile; its share of the flat action -> torque pass;
//   A      part sums -> shared memory, team barrier (named barrier, 128 threads);
//   comb.  warp 0: COM, torso cvel, torso / contact / phase / player-action observation entries, reward, termination,
//          motion state machine, counters (team_combine, dm_team_core.cuh), per-env scalars -> global memory;
//   B      team barrier; the observation tile leaves with ONE bulk store shared -> global; the statistics pass reads
//          it back column-wise (one column per thread, fp64 running sums in registers).
// Teams are independent: a CTA is T teams (T x 128 threads) that never synchronise with each other after the prologue,
// so one team's TMA wait or combine step overlaps the others' arithmetic.
#pragma once
#include "dm_tile_kernel.cuh"
#include "dm_team_core.cuh"
#include "generated/parts_humanoid3d.cuh"
#include "generated/parts_unitree_g1.cuh"

namespace dm {

__device__ __forceinline__ void team_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void team_sync(int team) {  // named barrier 1 + team, the 128 threads of one team
  asm volatile("bar.sync %0, 128;" ::"r"(team + 1) : "memory");
}

template <int KIND, class Rb>
struct TeamCfg {
  using OL = ObsLayout<KIND, Rb>;
  using PT = typename PartsOf<Rb>::type;
  static constexpr int TE = 32, NP = PT::NP, NT = 32 * NP;
  static_assert(NP == 4, "a team is four warps");
  static constexpr int MAXT = 8;                                          // teams per CTA (1024 threads)
  static constexpr int QF = TE * Rb::NQ, VF = TE * Rb::NV, AF = TE * Rb::NA;
  static constexpr int IN_FLOATS = QF + VF + AF;                          // one input slot: qpos | qvel | action rows
  static constexpr int OUT_FLOATS = TE * (OL::DIM + 1);                   // observation tile + the reward column
  static constexpr int PART_FLOATS = (NP - 1) * kPartVals * 32;           // sums of parts 1 .. NP-1, [part][value][lane]
  static constexpr int X_FLOATS = 8 * 32;                                 // [value][lane]: ep_len, contact, ep_return, root rpy
  static constexpr int STAT_FLOATS = 2 * (2 * NT + 4);                    // fp64 running sums: [2][thread] + finished-episode sums
  // a team's block: [output tile | part sums | scalars | statistics | nin input slots]
  static constexpr int OUT0 = 0, PART0 = OUT_FLOATS, X0 = PART0 + PART_FLOATS, STAT0 = X0 + X_FLOATS, IN0 = STAT0 + STAT_FLOATS;
  __host__ __device__ static constexpr int team_floats(int nin) { return IN0 + nin * IN_FLOATS; }
  static constexpr int NUP = (Rb::NU + 3) / 4 * 4;
  static constexpr size_t BAR_BYTES = 512;                                // 4 mbarriers per team (in0, in1, out-free, -) + table
  static constexpr size_t FIXED_BYTES = BAR_BYTES + sizeof(float) * 3 * NUP;
  static_assert(FIXED_BYTES % 16 == 0 && IN0 % 4 == 0, "alignment");
  static constexpr size_t SMEM_BUDGET = 227 * 1024;
};

struct TeamExtra {
  double* acc;       // statistics accumulators or NULL
  int n_tiles;       // full 32-environment tiles in this launch
  int table_floats;  // packed clip table size (floats); staged in shared memory when TABLE_SMEM
  int nin;           // input slots per team (1 or 2)
  int skip_compute;  // experiment: memory skeleton only
  unsigned long long* trace;  // debug (DM_TEAM_TRACE): %globaltimer marks of CTA 0, team 0, first tile: [0..15] warp 0, [16..31] warp 1
};

__device__ __forceinline__ unsigned long long team_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define DM_TR(k)                                                                                        \
  do {                                                                                                  \
    if (x.trace && blockIdx.x == 0 && team == 0 && lane == 0 && wt < 2) x.trace[wt * 16 + (k)] = team_now(); \
  } while (0)

template <class Rb>
struct TeamIn {  // a thread's view of its environment's staged rows
  const float* __restrict__ qp;
  const float* __restrict__ qv;
  __device__ __forceinline__ float qpos(int j) const { return qp[j]; }
  __device__ __forceinline__ float qvel(int j) const { return qv[j]; }
};
struct TeamOut {
  float* __restrict__ orow;
  __device__ __forceinline__ void obs(int j, float v) { orow[j] = v; }
};

template <int KIND, class Rb, bool TERMS, bool STATS, bool TABLE_SMEM, int MAXT>
__global__ void __launch_bounds__(128 * MAXT, 1) step_team_kernel(const __grid_constant__ StepParams p, const TeamExtra x) {
  using OL = ObsLayout<KIND, Rb>;
  using TC = TeamCfg<KIND, Rb>;
  using RL = RowLayout<Rb>;
  using PT = typename PartsOf<Rb>::type;
  constexpr int TE = TC::TE, DIM = OL::DIM;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int tid = threadIdx.x, team = tid >> 7, tt = tid & 127, wt = tt >> 5, lane = tid & 31;
  const int n_teams = blockDim.x >> 7, nin = x.nin;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);  // [team][4]: input slot 0, 1, out-free; [4 MAXT]: table
  uint64_t* tb = bars + team * 4;
  uint64_t* table_bar = bars + 4 * TC::MAXT;
  float* s_motor = reinterpret_cast<float*>(smem_raw + TC::BAR_BYTES);  // [3][NUP]: gear | ctrl lo | ctrl hi
  float* s_table = s_motor + 3 * TC::NUP;
  float* s_team = s_table + (TABLE_SMEM ? x.table_floats : 0) + (size_t)team * TC::team_floats(nin);
  float* s_out = s_team + TC::OUT0;
  float* s_rw = s_out + TE * DIM;
  float* s_part = s_team + TC::PART0;
  float* s_x = s_team + TC::X0;
  double* s_stat = reinterpret_cast<double*>(s_team + TC::STAT0);  // [0][tt] column sum, [1][tt] sum of squares, then 4 scalars

  // this team's tiles: (j * grid + cta) * n_teams + team, j = 0 .. nt-1 (the teams of a CTA work on adjacent tiles)
  const int t_base = blockIdx.x * n_teams + team, t_stride = gridDim.x * n_teams;
  const int nt = t_base < x.n_tiles ? (x.n_tiles - 1 - t_base) / t_stride + 1 : 0;
  const bool want_tau = p.torque != nullptr;
  auto issue_load = [&](int j) {  // one thread
    const int slot = j % nin;
    const long long f0 = (long long)(t_base + j * t_stride) * TE;
    float* dst = s_team + TC::IN0 + slot * TC::IN_FLOATS;
    mbar_expect_tx(tb + slot, sizeof(float) * (TC::QF + TC::VF + (want_tau ? TC::AF : 0)));
    tma_load_1d(dst, p.qpos + f0 * Rb::NQ, sizeof(float) * TC::QF, tb + slot);
    tma_load_1d(dst + TC::QF, p.qvel + f0 * Rb::NV, sizeof(float) * TC::VF, tb + slot);
    if (want_tau) tma_load_1d(dst + TC::QF + TC::VF, p.action + f0 * Rb::NA, sizeof(float) * TC::AF, tb + slot);
  };

  DM_TR(0);
  if (tid == 0) {
    for (int g = 0; g < n_teams; ++g) {
      mbar_init(bars + 4 * g + 0, 1);
      mbar_init(bars + 4 * g + 1, 1);
      mbar_init(bars + 4 * g + 2, TC::NP);  // out-free: one arrival per warp of the team
    }
    mbar_init(table_bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  // Nothing up to griddepcontrol.wait touches what an earlier launch writes (see dm_tile_kernel.cuh): under programmatic
  // dependent launch the launch latency, this prologue and the clip-table load overlap the previous kernel's tail.
  if (TABLE_SMEM && tid == 0) {
    mbar_expect_tx(table_bar, sizeof(float) * x.table_floats);
    tma_load_1d(s_table, p.cs.rows, sizeof(float) * x.table_floats, table_bar);
  }
  const float* rows = TABLE_SMEM ? s_table : p.cs.rows;
  for (int k = tid; k < TC::NUP; k += blockDim.x) {
    float g = 0.0f, lo = 0.0f, hi = 0.0f;
#pragma unroll
    for (int kk = 0; kk < Rb::NU; ++kk)
      if (kk == k) { g = Rb::gear(kk); lo = Rb::ctrl_lo(kk); hi = Rb::ctrl_hi(kk); }
    s_motor[k] = g;
    s_motor[TC::NUP + k] = lo;
    s_motor[2 * TC::NUP + k] = hi;
  }
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  DM_TR(1);
  if (tt == 0)
    for (int j = 0; j < nin && j < nt; ++j) issue_load(j);
  __syncthreads();  // motor table visible

  bool table_ready = !TABLE_SMEM;
  const bool cfk = p.cs.contact_from_fk != 0;
  if (STATS) {  // statistics: thread tt of a team owns column tt of the tile (observation columns, then the reward)
    s_stat[tt] = 0.0;
    s_stat[TC::NT + tt] = 0.0;
    if (tt < 4) s_stat[2 * TC::NT + tt] = 0.0;
  }

  for (int j = 0; j < nt; ++j) {
    const int slot = j % nin;
    const uint32_t parity = (j / nin) & 1;
    const int tile = t_base + j * t_stride;
    const int env = tile * TE + lane;
    const float* s_in = s_team + TC::IN0 + slot * TC::IN_FLOATS;
    // per-environment scalars (coalesced 128-byte loads; every warp needs the frame of its lane's environment).  What
    // only the combine step needs is parked in shared memory so that it does not occupy registers during the parts.
    EnvState st;
    st.idx = ld_stream(p.idx + env);
    st.clip = p.clip ? ld_stream(p.clip + env) : 0;
    if (wt == 0) {
      s_x[0 * 32 + lane] = __int_as_float(ld_stream(p.ep_len + env));
      s_x[1 * 32 + lane] = __uint_as_float(p.contact ? ld_stream(p.contact + env) : 0u);
      s_x[2 * 32 + lane] = p.ep_return ? ld_stream(p.ep_return + env) : 0.0f;
    }
    const ClipDesc cd = p.cs.clip[st.clip];
    const int frame = (KIND == KIND_SINGLE) ? st.idx : (st.idx % cd.len);
    const int row_i = cd.row0 + (cd.fixed_row >= 0 ? cd.fixed_row : frame);
    const float* __restrict__ row = rows + (size_t)row_i * RL::STRIDE;
    TeamIn<Rb> in{s_in + lane * Rb::NQ, s_in + TC::QF + lane * Rb::NV};
    TeamOut out{s_out + lane * DIM};
    typename PT::Keep keep;
    RootIn rin;
    PartSums mine;
    // the previous tile's observation store has finished reading the output tile: warp 0 reports it, then every warp
    // waits until all four have let go of the tile (statistics reads, the store)
    if (j > 0) {
      if (tt == 0) bulk_wait_read0();
      if (wt == 0) {
        __syncwarp();
        if (lane == 0) team_mbar_arrive(tb + 2);
      }
    }
    auto run_part = [&](auto P) {
      constexpr int PP = decltype(P)::value;
      PartAcc<KIND, Rb, PP, TeamOut> acc(out, row, cfk);
      if (!TABLE_SMEM) acc.load_row();  // L2 gathers in flight while the tile lands
      if (!table_ready) {
        mbar_wait(table_bar, 0);
        table_ready = true;
      }
      mbar_wait(tb + slot, parity);
      if (j > 0) mbar_wait(tb + 2, (j - 1) & 1);
      if (TABLE_SMEM) acc.load_row();
      if (j == 0) DM_TR(2);
      if (!x.skip_compute) {
        PT::template part<PP>(in, acc, keep);
      } else {
        acc.put(PT::j0(PP), in.qpos(7 + PT::j0(PP)));
      }
      if (PP == PT::RPY_PART) {  // roll / pitch / yaw of the free-joint quaternion for the combine step
        float r0, r1, r2;
        dm_rpy_raw(in.qpos(3), in.qpos(4), in.qpos(5), in.qpos(6), r0, r1, r2);
        s_x[3 * 32 + lane] = r0; s_x[4 * 32 + lane] = r1; s_x[5 * 32 + lane] = r2;
      }
      mine = acc.sums();
    };
    if (wt == 0) {
      run_part(std::integral_constant<int, 0>{});
#pragma unroll
      for (int k = 0; k < 3; ++k) rin.v[k] = in.qvel(k);
    } else if (wt == 1) {
      run_part(std::integral_constant<int, 1>{});
    } else if (wt == 2) {
      run_part(std::integral_constant<int, 2>{});
    } else {
      run_part(std::integral_constant<int, 3>{});
    }
    // action -> torque: a flat pass over the tile's [32, NU] outputs, a quarter per warp, coalesced stores
    if (want_tau) {
      const float* s_act = s_in + TC::QF + TC::VF;
      float* trow = p.torque + (long long)tile * TE * Rb::NU;
      constexpr int TOT = TE * Rb::NU, PER = (TOT + TC::NT - 1) / TC::NT;
#pragma unroll
      for (int m = 0; m < PER; ++m) {
        const int o = tt + m * TC::NT;
        if (o < TOT) {
          const int r = o / Rb::NU, k = o - r * Rb::NU;
          float c = (k < Rb::NA ? s_act[r * Rb::NA + k] : 0.0f) * Rb::kActScale;
          c = dm_max_nan(dm_min_nan(c, s_motor[2 * TC::NUP + k]), s_motor[TC::NUP + k]);  // mju_clip, NaN passes through
          trow[o] = k < Rb::NA ? s_motor[k] * c : 0.0f;
        }
      }
    }
    if (wt > 0) {
      float* sp = s_part + (wt - 1) * kPartVals * 32 + lane;
      sp[0 * 32] = mine.s1x; sp[1 * 32] = mine.s1y; sp[2 * 32] = mine.s1z;
      sp[3 * 32] = mine.s2x; sp[4 * 32] = mine.s2y; sp[5 * 32] = mine.s2z;
      sp[6 * 32] = mine.e_q; sp[7 * 32] = mine.e_v; sp[8 * 32] = mine.max_dq; sp[9 * 32] = mine.e_ee;
      sp[10 * 32] = mine.omax; sp[11 * 32] = mine.omin; sp[12 * 32] = __uint_as_float(mine.word);
    }
    if (j == 0) DM_TR(3);
    fence_async_smem();  // this thread's writes to the output tile become visible to the bulk store (async proxy)
    team_sync(team);     // A: every part's sums and observation entries are in shared memory, the input slot is dead
    if (tt == 0 && j + nin < nt) issue_load(j + nin);
    if (j == 0) DM_TR(4);
    if (wt == 0) {
#pragma unroll
      for (int q = 0; q < TC::NP - 1; ++q) {
        const float* sp = s_part + q * kPartVals * 32 + lane;
        PartSums o{sp[0], sp[32], sp[64], sp[96], sp[128], sp[160], sp[192], sp[224], sp[256], sp[288], sp[320], sp[352],
                   __float_as_uint(sp[384])};
        part_merge(mine, o);
      }
      st.ep_len = __float_as_int(s_x[0 * 32 + lane]);
      const uint32_t contact = __float_as_uint(s_x[1 * 32 + lane]);
      const float ep_ret0 = s_x[2 * 32 + lane];
      const float root_rpy[3] = {s_x[3 * 32 + lane], s_x[4 * 32 + lane], s_x[5 * 32 + lane]};
      Vec4 tail[5];
#pragma unroll
      for (int c = 0; c < 5; ++c) tail[c] = ld4(row + RL::ROLL + 4 * c);
      StepOut so;
      bool guard = false;
      if (!x.skip_compute) {
        guard = team_combine<KIND, Rb, TERMS>(out, keep, rin, root_rpy, mine, &tail[0].v[0], cd, frame, cfk, contact, st, so);
      } else {
        so.reward = so.raw_reward = mine.s1x; so.done = 0; so.reason = 0;
      }
      if (guard) {  // the row leaves as `obs * 0.` (zeros; NaN where an entry was infinite, deepmimic_env.py:476); rare
#pragma unroll 1
        for (int k = 0; k < DIM; ++k) out.orow[k] = out.orow[k] * 0.0f;
      }
      finish_env<KIND, Rb, TERMS, false>(p, env, st, so);
      const float ep_ret1 = ep_ret0 + so.raw_reward;
      if (p.ep_return) p.ep_return[env] = ep_ret1;
      if (STATS) {
        s_rw[lane] = so.reward;
        if (so.done) {  // finished episodes are rare: shared-memory fp64 atomics
          atomicAdd(s_stat + 2 * TC::NT + 0, 1.0);
          atomicAdd(s_stat + 2 * TC::NT + 1, double(ep_ret1));
          atomicAdd(s_stat + 2 * TC::NT + 2, double(st.ep_len));
        }
      }
      fence_async_smem();  // ... and so do the combine step's
    }
    if (j == 0) DM_TR(5);
    if (STATS) {
      team_sync(team);   // B: the tile is final
    } else if (wt == 0) {
      __syncwarp();
    }
    if (tt == 0) {  // (without statistics warps 1-3 passed A after their fence, warp 0 has just synchronised)
      tma_store_1d(p.obs + (long long)tile * TE * DIM, s_out, sizeof(float) * TE * DIM);
      bulk_commit();
    }
    if (STATS && tt <= DIM) {  // column tt of the finished tile; column DIM is the reward
      float cs = 0.0f, css = 0.0f;
      const float* src = tt < DIM ? s_out + tt : s_rw;
      const int stride = tt < DIM ? DIM : 1;
#pragma unroll 8
      for (int r = 0; r < TE; ++r) {
        const float v = src[r * stride];
        cs += v;
        css = fmaf(v, v, css);
      }
      s_stat[tt] += cs;
      s_stat[TC::NT + tt] += css;
    }
    if (j == 0) DM_TR(6);
    // this warp is done with the output tile (warp 0 arrives at the top of the next tile, after its store has read it)
    if (wt > 0 && j + 1 < nt) {
      __syncwarp();
      if (lane == 0) team_mbar_arrive(tb + 2);
    }
  }
  DM_TR(7);
  if (tt == 0) bulk_wait_read0();  // shared memory stays valid until the last store has read it
  DM_TR(8);

  if (STATS && x.acc) {
    // CTA-level reduction of the teams' running sums (team 0 adds them up), then ONE set of fp64 atomics per CTA
    if (tt == 0) s_stat[2 * TC::NT + 3] = double(nt);
    __syncthreads();
    const float* blk0 = s_table + (TABLE_SMEM ? x.table_floats : 0);
    auto stat_of = [&](int g) { return reinterpret_cast<const double*>(blk0 + (size_t)g * TC::team_floats(nin) + TC::STAT0); };
    if (team == 0) {
      if (tt <= DIM) {
        double s = 0.0, q = 0.0;
        for (int g = 0; g < n_teams; ++g) { s += stat_of(g)[tt]; q += stat_of(g)[TC::NT + tt]; }
        const int js = tt < DIM ? 6 + tt : 1, jq = tt < DIM ? 6 + DIM + tt : 2;
        if (s != 0.0) atomicAdd(x.acc + js, s);
        if (q != 0.0) atomicAdd(x.acc + jq, q);
      }
      if (tt >= 124) {  // 124: count, 125: n_done, 126: sum ep_return(done), 127: sum ep_len(done)
        const int k = tt - 124;
        double v = 0.0;
        for (int g = 0; g < n_teams; ++g) v += k == 0 ? stat_of(g)[2 * TC::NT + 3] * TE : stat_of(g)[2 * TC::NT + k - 1];
        if (v != 0.0) atomicAdd(x.acc + (k == 0 ? 0 : 2 + k), v);
      }
    }
  }
  DM_TR(9);
}

}  // namespace dm
