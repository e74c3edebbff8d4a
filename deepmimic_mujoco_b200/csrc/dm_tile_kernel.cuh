// K1b: row-major (AoS) step kernel -- persistent CTAs, shared-memory tiles, TMA bulk loads.
//
// Launch shape: one CTA per SM, G groups of 128 threads per CTA (G chosen from the shared-memory budget).
// Each group walks tiles of TE = 128 consecutive environments (tile = cta * G + group, stride grid * G):
//   * the tile's qpos / qvel rows are contiguous in the row-major arrays, so each is ONE 1-D TMA bulk copy
//     (cp.async.bulk global -> shared, completion on the group's mbarrier; SASS: UBLKCP);
//   * the packed clip table is staged ONCE per CTA in shared memory by the same mechanism (when it fits), so
//     the per-environment gather of a random mocap frame is a 128-bit shared-memory load, not an L1 gather;
//   * every thread steps its own environment out of shared memory (row stride nq / nv words);
//   * nothing is written to global memory with a row stride: the observation is emitted by a cooperative,
//     fully coalesced pass (warps walk rows, lanes walk columns) that takes the joint blocks straight from
//     the staged qpos / qvel tiles and the computed entries from a small per-env tile; action -> torque is a
//     second cooperative pass, global -> global, one lane per actuator, issued while the TMA is in flight;
//     reward / done / reason / counters are one coalesced store per array.
// Groups synchronise among themselves only (named barriers), so one group's TMA wait overlaps the others'
// arithmetic.  Shared memory per environment is ~320 B (humanoid3d), i.e. 4 tiles + the table per SM.
// The observation-normaliser moments are accumulated by the cooperative pass (per-lane column sums kept in
// registers across ALL tiles of the group) and flushed with one set of fp64 atomics per CTA.
#pragma once
#include "dm_kernels.cuh"

namespace dm {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// 1-D TMA bulk copy global -> shared (16-byte aligned, size multiple of 16), completes on `bar`
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void l2_prefetch(const void* src, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void group_sync(int group) {  // named barrier 1 + group, 128 threads
  asm volatile("bar.sync %0, 128;" ::"r"(group + 1) : "memory");
}

template <int KIND, class Rb>
struct TileCfg {
  using OL = ObsLayout<KIND, Rb>;
  static constexpr int NCOMP = OL::DIM - 2 * Rb::NJ;  // computed obs entries (torso, contacts, phase, player action)
  static constexpr int TE = 128;
  static constexpr int MAXG = 4;
  static constexpr int QP = TE * Rb::NQ, QV = TE * Rb::NV, CO = TE * NCOMP;
  static constexpr int GROUP_FLOATS = QP + QV + CO + TE / 4;  // + guard flags (1 byte per env)
  static constexpr size_t GROUP_BYTES = sizeof(float) * GROUP_FLOATS;
  static constexpr int NUP = (Rb::NU + 3) / 4 * 4;
  static constexpr size_t FIXED_BYTES = 128 + sizeof(float) * 3 * NUP;  // mbarriers + motor table (gear | lo | hi)
  static_assert(FIXED_BYTES % 16 == 0, "alignment");
  static constexpr size_t SMEM_BUDGET = 227 * 1024;
};

// per-thread accessor into the staged tiles
template <int KIND, class Rb>
struct TileIO {
  const float* __restrict__ qp;  // this env's qpos row in shared memory
  const float* __restrict__ qv;
  float* __restrict__ co;        // this env's computed-obs row in shared memory
  __device__ __forceinline__ float qpos(int j) const { return qp[j]; }
  __device__ __forceinline__ float qvel(int j) const { return qv[j]; }
  __device__ __forceinline__ float act(int) const { return 0.0f; }
  __device__ __forceinline__ void obs(int j, float v) {
    if (j >= 2 * Rb::NJ) co[j - 2 * Rb::NJ] = v;  // the joint blocks are copied by the cooperative pass
  }
  __device__ __forceinline__ void tau(int, float) {}
};

struct TileExtra {
  double* acc;       // statistics accumulators or NULL
  int n_tiles;       // full tiles in this launch
  int table_floats;  // packed clip table size (floats); staged in shared memory when TABLE_SMEM
};

template <int KIND, class Rb, bool TERMS, bool STATS, bool TABLE_SMEM>
__global__ void __launch_bounds__(512, 1) step_tile_kernel(const __grid_constant__ StepParams p, const TileExtra x) {
  using OL = ObsLayout<KIND, Rb>;
  using TC = TileCfg<KIND, Rb>;
  constexpr int TE = TC::TE, NJ = Rb::NJ, DIM = OL::DIM, NCOMP = TC::NCOMP;
  constexpr int NCOLIT = (DIM + 31) / 32;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_raw);  // [0..MAXG) group barriers, [MAXG] table barrier
  float* s_motor = reinterpret_cast<float*>(smem_raw + 128);  // [3][NUP]
  float* s_table = reinterpret_cast<float*>(smem_raw + TC::FIXED_BYTES);
  const int n_groups = blockDim.x / TE;
  const int tid = threadIdx.x, group = tid / TE, gt = tid % TE, lane = tid & 31, warp = gt >> 5;
  float* s_qp = s_table + (TABLE_SMEM ? x.table_floats : 0) + (size_t)group * TC::GROUP_FLOATS;
  float* s_qv = s_qp + TC::QP;
  float* s_co = s_qv + TC::QV;
  unsigned char* s_guard = reinterpret_cast<unsigned char*>(s_co + TC::CO);
  uint64_t* bar = bars + group;

  if (tid == 0) {
    for (int g = 0; g <= TC::MAXG; ++g) mbar_init(bars + g, 1);
    mbar_fence_init();
  }
  __syncthreads();
  if (TABLE_SMEM && tid == 0) {
    mbar_expect_tx(bars + TC::MAXG, sizeof(float) * x.table_floats);
    tma_load_1d(s_table, p.cs.rows, sizeof(float) * x.table_floats, bars + TC::MAXG);
  }
  const float* rows = TABLE_SMEM ? s_table : p.cs.rows;

  // motor constants (gear, ctrlrange) of every actuator, looked up by the flat torque pass
  for (int k = tid; k < TC::NUP; k += blockDim.x) {
    float g = 0.0f, lo = 0.0f, hi = 0.0f;
#pragma unroll
    for (int kk = 0; kk < Rb::NU; ++kk)
      if (kk == k) { g = Rb::gear(kk); lo = Rb::ctrl_lo(kk); hi = Rb::ctrl_hi(kk); }
    s_motor[k] = g;
    s_motor[TC::NUP + k] = lo;
    s_motor[2 * TC::NUP + k] = hi;
  }
  __syncthreads();
  // statistics carried across tiles
  double a_cs[NCOLIT], a_css[NCOLIT], a_sc[6];
#pragma unroll
  for (int it = 0; it < NCOLIT; ++it) a_cs[it] = a_css[it] = 0.0;
#pragma unroll
  for (int k = 0; k < 6; ++k) a_sc[k] = 0.0;

  uint32_t parity = 0;
  bool table_ready = !TABLE_SMEM;
  for (int tile = blockIdx.x * n_groups + group; tile < x.n_tiles; tile += gridDim.x * n_groups) {
    const long long e0 = (long long)tile * TE;
    if (gt == 0) {
      mbar_expect_tx(bar, sizeof(float) * (TC::QP + TC::QV));
      tma_load_1d(s_qp, p.qpos + e0 * Rb::NQ, sizeof(float) * TC::QP, bar);
      tma_load_1d(s_qv, p.qvel + e0 * Rb::NV, sizeof(float) * TC::QV, bar);
      const int nxt = tile + gridDim.x * n_groups;  // pull the group's next tile into L2 behind this one
      if (nxt < x.n_tiles) {
        const long long n0 = (long long)nxt * TE;
        l2_prefetch(p.qpos + n0 * Rb::NQ, sizeof(float) * TC::QP);
        l2_prefetch(p.qvel + n0 * Rb::NV, sizeof(float) * TC::QV);
        l2_prefetch(p.action + n0 * Rb::NA, sizeof(float) * TE * Rb::NA);
      }
    }
    // ---- per-env scalars: issue the loads now, they are consumed after the TMA wait -------------------------
    const int env = int(e0) + gt;
    EnvState st;
    st.idx = p.idx[env];
    st.ep_len = p.ep_len[env];
    st.clip = p.clip ? p.clip[env] : 0;
    const uint32_t contact = p.contact ? p.contact[env] : 0u;
    const float ep_ret0 = p.ep_return ? p.ep_return[env] : 0.0f;
    // ---- action -> torque while the tile is in flight: a flat, fully coalesced pass over the [TE, NA] tile with
    //      every load issued before the first use (the tile is contiguous in the row-major action array) --------
    if (p.torque) {
      if (Rb::NA == Rb::NU && Rb::NA % 4 == 0) {  // 128-bit path: chunk c covers actuators 4 (c mod NA/4) .. + 3
        constexpr int CH = TE * Rb::NA / 4, PER = (CH + TE - 1) / TE, K4 = Rb::NA / 4;
        const float4* a4 = reinterpret_cast<const float4*>(p.action + e0 * Rb::NA);
        float4* t4 = reinterpret_cast<float4*>(p.torque + e0 * Rb::NU);
        float4 v[PER];
#pragma unroll
        for (int m = 0; m < PER; ++m)
          if (gt + m * TE < CH) v[m] = __ldg(a4 + gt + m * TE);
#pragma unroll
        for (int m = 0; m < PER; ++m) {
          const int c = gt + m * TE;
          if (c < CH) {
            const int k0 = (c % K4) * 4;
            const float4 g = *reinterpret_cast<const float4*>(s_motor + k0);
            const float4 lo = *reinterpret_cast<const float4*>(s_motor + TC::NUP + k0);
            const float4 hi = *reinterpret_cast<const float4*>(s_motor + 2 * TC::NUP + k0);
            float4 o;
            float c0 = v[m].x * Rb::kActScale, c1 = v[m].y * Rb::kActScale, c2 = v[m].z * Rb::kActScale, c3 = v[m].w * Rb::kActScale;
            c0 = c0 > hi.x ? hi.x : c0; c0 = c0 < lo.x ? lo.x : c0; o.x = g.x * c0;
            c1 = c1 > hi.y ? hi.y : c1; c1 = c1 < lo.y ? lo.y : c1; o.y = g.y * c1;
            c2 = c2 > hi.z ? hi.z : c2; c2 = c2 < lo.z ? lo.z : c2; o.z = g.z * c2;
            c3 = c3 > hi.w ? hi.w : c3; c3 = c3 < lo.w ? lo.w : c3; o.w = g.w * c3;
            t4[c] = o;
          }
        }
      } else {  // scalar flat path: output element o = (row, actuator); hand actuators get zero torque
        constexpr int TOT = TE * Rb::NU, PER = (TOT + TE - 1) / TE, HALF = (PER + 1) / 2;
        const float* arow = p.action + e0 * Rb::NA;
        float* trow = p.torque + e0 * Rb::NU;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          float v[HALF];
#pragma unroll
          for (int m = 0; m < HALF; ++m) {
            const int o = gt + (h * HALF + m) * TE, r = o / Rb::NU, k = o - r * Rb::NU;
            v[m] = (h * HALF + m < PER && o < TOT && k < Rb::NA) ? __ldg(arow + r * Rb::NA + k) : 0.0f;
          }
#pragma unroll
          for (int m = 0; m < HALF; ++m) {
            const int o = gt + (h * HALF + m) * TE, r = o / Rb::NU, k = o - r * Rb::NU;
            if (h * HALF + m < PER && o < TOT) {
              float c = v[m] * Rb::kActScale;
              const float hi = s_motor[2 * TC::NUP + k], lo = s_motor[TC::NUP + k];
              c = c > hi ? hi : c;
              c = c < lo ? lo : c;
              trow[o] = k < Rb::NA ? s_motor[k] * c : 0.0f;
            }
          }
        }
      }
    }
    // ---- per-environment step out of shared memory --------------------------------------------------
    if (!table_ready) {
      mbar_wait(bars + TC::MAXG, 0);
      table_ready = true;
    }
    mbar_wait(bar, parity);
    parity ^= 1u;
    TileIO<KIND, Rb> io{s_qp + gt * Rb::NQ, s_qv + gt * Rb::NV, s_co + gt * NCOMP};
    StepOut so;
    env_step<KIND, Rb, TileIO<KIND, Rb>, TERMS>(io, p.cs, rows, st, contact, so);
    s_guard[gt] = (so.reason == R_OBS_BOUNDS) ? 1 : 0;
    finish_env<KIND, Rb, TERMS, false>(p, env, st, so);
    const float ep_ret1 = ep_ret0 + so.raw_reward;
    if (p.ep_return) p.ep_return[env] = ep_ret1;
    if (STATS) {
      a_sc[0] += 1.0;
      a_sc[1] += so.reward;
      a_sc[2] += double(so.reward) * so.reward;
      if (so.done) {
        a_sc[3] += 1.0;
        a_sc[4] += ep_ret1;
        a_sc[5] += st.ep_len;
      }
    }
    group_sync(group);
    // ---- observation: coalesced rows, joint blocks straight from the staged tiles -----------------------
    float cs32[NCOLIT], css32[NCOLIT];
#pragma unroll
    for (int it = 0; it < NCOLIT; ++it) cs32[it] = css32[it] = 0.0f;
#pragma unroll 2
    for (int r = warp; r < TE; r += 4) {
      const bool guard = s_guard[r] != 0;
      float* orow = p.obs + (e0 + r) * DIM;
#pragma unroll
      for (int it = 0; it < NCOLIT; ++it) {
        const int j = it * 32 + lane;
        if (j < DIM) {
          float v;
          if (j < NJ) v = s_qp[r * Rb::NQ + 7 + j];
          else if (j < 2 * NJ) v = s_qv[r * Rb::NV + 6 + (j - NJ)] * 0.1f;
          else v = s_co[r * NCOMP + (j - 2 * NJ)];
          if (guard) v = 0.0f;
          orow[j] = v;
          if (STATS) {
            cs32[it] += v;
            css32[it] = fmaf(v, v, css32[it]);
          }
        }
      }
    }
    if (STATS) {
#pragma unroll
      for (int it = 0; it < NCOLIT; ++it) {
        a_cs[it] += cs32[it];
        a_css[it] += css32[it];
      }
    }
    group_sync(group);  // everyone is done reading the tile before the next TMA overwrites it
  }

  if (STATS && x.acc) {
    // CTA-level reduction in shared memory (the tiles are dead by now), then ONE set of fp64 atomics per CTA
    __syncthreads();
    double* sc = reinterpret_cast<double*>(s_table + (TABLE_SMEM ? x.table_floats : 0));  // [n_warps][2 * NCOLIT * 32 + 6]
    constexpr int W = 2 * NCOLIT * 32 + 6;
    const int wid = tid >> 5, n_warps = blockDim.x >> 5;
#pragma unroll
    for (int it = 0; it < NCOLIT; ++it) {
      sc[wid * W + it * 32 + lane] = a_cs[it];
      sc[wid * W + NCOLIT * 32 + it * 32 + lane] = a_css[it];
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const double v = warp_sum(a_sc[k]);
      if (lane == 0) sc[wid * W + 2 * NCOLIT * 32 + k] = v;
    }
    __syncthreads();
    for (int j = tid; j < 2 * DIM + 6; j += blockDim.x) {
      // j < 6: scalars; then sum obs[DIM]; then sum obs^2[DIM]
      const int col = j < 6 ? 2 * NCOLIT * 32 + j : (j < 6 + DIM ? j - 6 : NCOLIT * 32 + (j - 6 - DIM));
      double v = 0.0;
      for (int w = 0; w < n_warps; ++w) v += sc[w * W + col];
      if (v != 0.0) atomicAdd(x.acc + j, v);
    }
  }
}

}  // namespace dm
