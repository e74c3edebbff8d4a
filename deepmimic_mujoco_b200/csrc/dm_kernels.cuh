// CUDA kernels of the hot path (sm_100a).  Thread-per-environment; see DESIGN.md for the layout
// and roofline discussion.  Kernel list:
//   step_strided_kernel  K1a  any layout through (env stride, element stride); coalesced for SoA
//   step_tile_kernel     K1b  row-major AoS: tiles of environments staged in shared memory with TMA
//                             bulk copies (cp.async.bulk + mbarrier), results written back the same way
//   reset_kernel         K2   masked reset + reference-state-init + reset observation
//   fk_table_kernel      K3   float64 FK of clip key frames (load time)
//   stats_kernel         K4   reward / observation moments -> fp64 accumulators
#pragma once
#include <cuda_runtime.h>

#include "dm_step.cuh"
#include "dm_fk_table.cuh"
#include "generated/robot_humanoid3d.cuh"
#include "generated/robot_unitree_g1.cuh"

namespace dm {

struct StepParams {
  int n;
  long long es;      // element stride: 1 (AoS) or ld (SoA)
  int aos;           // 1: env stride = dim, 0: env stride = 1
  const float* qpos;
  const float* qvel;
  const float* action;
  const uint32_t* contact;
  int* idx;
  int* ep_len;
  int* clip;
  float* obs;
  float* reward;
  uint8_t* done;
  uint8_t* reason;
  float* torque;
  float* terms;
  float* ep_return;
  ClipSet cs;
};

// ---- accessor over global memory with strides ------------------------------------------------
struct StridedIO {
  const float* __restrict__ qp;
  const float* __restrict__ qv;
  const float* __restrict__ ac;
  float* __restrict__ ob;
  float* __restrict__ ta;
  long long es;
  __device__ __forceinline__ float qpos(int j) const { return __ldg(qp + j * es); }
  __device__ __forceinline__ float qvel(int j) const { return __ldg(qv + j * es); }
  __device__ __forceinline__ float act(int j) const { return __ldg(ac + j * es); }
  __device__ __forceinline__ void obs(int j, float v) { ob[j * es] = v; }
  __device__ __forceinline__ void obs_times_zero(int j) { ob[j * es] = ob[j * es] * 0.0f; }
  __device__ __forceinline__ void tau(int j, float v) {
    if (ta) ta[j * es] = v;
  }
};

template <int KIND, class Rb, bool TERMS, bool EP_RETURN = true>
__device__ __forceinline__ void finish_env(const StepParams& p, int env, const EnvState& st, const StepOut& so) {
  p.idx[env] = st.idx;
  p.ep_len[env] = st.ep_len;
  if (KIND == KIND_COMBINED) p.clip[env] = st.clip;
  p.reward[env] = so.reward;
  p.done[env] = so.done;
  if (p.reason) p.reason[env] = so.reason;
  if (EP_RETURN && p.ep_return) p.ep_return[env] += so.raw_reward;
  if (TERMS) {
    const long long base = p.aos ? (long long)env * 8 : env;
#pragma unroll
    for (int k = 0; k < 8; ++k) p.terms[base + k * p.es] = so.terms[k];
  }
}

template <int KIND, class Rb, bool TERMS>
__global__ void __launch_bounds__(128) step_strided_kernel(const __grid_constant__ StepParams p) {
  using OL = ObsLayout<KIND, Rb>;
  for (int env = blockIdx.x * blockDim.x + threadIdx.x; env < p.n; env += gridDim.x * blockDim.x) {
    const long long e = env;
    StridedIO io;
    io.es = p.es;
    io.qp = p.qpos + (p.aos ? e * Rb::NQ : e);
    io.qv = p.qvel + (p.aos ? e * Rb::NV : e);
    io.ac = p.action + (p.aos ? e * Rb::NA : e);
    io.ob = p.obs + (p.aos ? e * OL::DIM : e);
    io.ta = p.torque ? p.torque + (p.aos ? e * Rb::NU : e) : nullptr;
    EnvState st;
    st.idx = p.idx[env];
    st.ep_len = p.ep_len[env];
    st.clip = p.clip ? p.clip[env] : 0;
    const uint32_t contact = p.contact ? p.contact[env] : 0u;
    StepOut so;
    env_step<KIND, Rb, StridedIO, TERMS>(io, p.cs, p.cs.rows, st, contact, so);
    finish_env<KIND, Rb, TERMS>(p, env, st, so);
  }
}

// ---- K2: reset -----------------------------------------------------------------------------------
// Philox-4x32-10 counter-based generator (Salmon et al. 2011): stateless, reproducible per (seed, counter, env).
__device__ __forceinline__ uint4 philox4x32(uint2 key, uint4 ctr) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, ctr.x), lo0 = 0xD2511F53u * ctr.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, ctr.z), lo1 = 0xCD9E8D57u * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += 0x9E3779B9u;
    key.y += 0xBB67AE85u;
  }
  return ctr;
}
__device__ __forceinline__ int randint_below(uint32_t r, int n) { return int(((unsigned long long)r * (unsigned)n) >> 32); }

struct ResetParams {
  int n;
  long long es;
  int aos;
  const uint8_t* mask;
  const int* frame;
  const int* coin;
  const uint32_t* contact;
  unsigned long long seed, counter;
  float* qpos;
  float* qvel;
  int* idx;
  int* ep_len;
  int* clip;
  float* obs;
  float* ep_return;
  float* terminal_obs;
  const float* state_rows;  // [total_rows, NQ + NV] fp32 data_config | data_vel
  ClipSet cs;
};

// ---- K8: get_abspos_obs (deepmimic_env.py:133-137, ENV_CFG.ADD_ABSPOS_OBS -- off in every configuration the reference
// ships, so this is an opt-in POST-PASS and the step kernels stay as they are).  The extended row is
//   [ obs[0 : PHASE] | geom_xpos.flatten() (3 NGEOM) | obs[PHASE : DIM] ]
// (the reference concatenates position, velocity, torso, contacts, joint_force, abs_pos, phase, player action).  In step
// mode (reward != NULL) the observation-bounds guard (deepmimic_env.py:465-476) is re-evaluated over the new block: a geom
// further than 100 m from the origin ends the episode with (obs * 0, 0, True, {}) exactly like an out-of-range standard
// entry; a row the step kernel has already zeroed gets its new block zeroed too.
struct AbsPosParams {
  int n;
  const float* qpos;     // [n, nq] row-major
  const float* obs_in;   // [n, DIM] the step's / reset's observation
  float* obs_out;        // [n, DIM + 3 NGEOM]
  float* reward;         // step mode: [n], else NULL (reset mode: rows are only extended)
  uint8_t* done;
  uint8_t* reason;
  const uint8_t* mask;   // NULL or [n]: rows to (re)compute
};
struct AbsPosIO {
  const float* __restrict__ qp;
  __device__ __forceinline__ float qpos(int j) const { return __ldg(qp + j); }
  __device__ __forceinline__ float qvel(int) const { return 0.0f; }
};
struct AbsPosSink {
  static constexpr bool kWantsGeoms = true;
  __device__ static constexpr bool contact_on() { return false; }
  float* __restrict__ out;  // this row's geom block
  float m;                  // max |coordinate|, NaN-propagating like numpy's max / min
  template <typename T> __device__ __forceinline__ void body(int, T, T, T) {}
  __device__ __forceinline__ void geom(int g, float x, float y, float z) {
    out[3 * g] = x; out[3 * g + 1] = y; out[3 * g + 2] = z;
    m = dm_max_nan(m, dm_max_nan(fabsf(x), dm_max_nan(fabsf(y), fabsf(z))));
  }
};
template <int KIND, class Rb>
__global__ void __launch_bounds__(128) obs_abspos_kernel(const AbsPosParams p) {
  using OL = ObsLayout<KIND, Rb>;
  constexpr int NG3 = 3 * Rb::NGEOM, EXT = OL::DIM + NG3;
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= p.n || (p.mask && !p.mask[env])) return;
  const float* __restrict__ src = p.obs_in + (size_t)env * OL::DIM;
  float* __restrict__ dst = p.obs_out + (size_t)env * EXT;
  for (int j = 0; j < OL::PHASE; ++j) dst[j] = src[j];
  for (int j = OL::PHASE; j < OL::DIM; ++j) dst[NG3 + j] = src[j];
  AbsPosSink sink{dst + OL::PHASE, 0.0f};
  for (int j = 0; j < NG3; ++j) sink.out[j] = 0.0f;  // the world body's geoms (the floor plane at the origin)
  AbsPosIO in{p.qpos + (size_t)env * Rb::NQ};
  FkOut<float> o;
  Rb::template fk<float>(in, o, sink);
  if (!p.reward) return;
  const bool already = p.reason && p.reason[env] == R_OBS_BOUNDS;
  if (already || sink.m > 100.0f) {
    for (int j = 0; j < EXT; ++j) dst[j] = dst[j] * 0.0f;  // `_get_obs() * 0.`: an infinite entry becomes NaN
    p.reward[env] = 0.0f;
    p.done[env] = 1;
    if (p.reason) p.reason[env] = R_OBS_BOUNDS;
  }
}

struct ResetIO {
  const float* __restrict__ row;  // state row: qpos | qvel
  float* __restrict__ ob;
  long long es;
  int nq;
  __device__ __forceinline__ float qpos(int j) const { return __ldg(row + j); }
  __device__ __forceinline__ float qvel(int j) const { return __ldg(row + nq + j); }
  __device__ __forceinline__ float act(int) const { return 0.0f; }
  __device__ __forceinline__ void obs(int j, float v) { ob[j * es] = v; }
  __device__ __forceinline__ void tau(int, float) {}
};

struct ResetSink {  // FK sink of the reset observation: only the floor-contact switch
  static constexpr bool kWantsGeoms = false;
  bool want_contact;
  __device__ __forceinline__ bool contact_on() const { return want_contact; }
  template <typename T> __device__ __forceinline__ void body(int, T, T, T) {}
  template <typename T> __device__ __forceinline__ void geom(int, T, T, T) {}
};

template <int KIND, class Rb>
__global__ void __launch_bounds__(128) reset_kernel(const __grid_constant__ ResetParams p) {
  using OL = ObsLayout<KIND, Rb>;
  const int env = blockIdx.x * blockDim.x + threadIdx.x;
  if (env >= p.n) return;
  if (p.mask && !p.mask[env]) return;
  const uint4 rnd = philox4x32(make_uint2((uint32_t)p.seed, (uint32_t)(p.seed >> 32)),
                               make_uint4((uint32_t)env, (uint32_t)p.counter, (uint32_t)(p.counter >> 32), 0x444d484fu));
  int clip, n_steps;
  if (KIND == KIND_SINGLE) {  // reference_state_init: idx_init = randint(0, len - 1)
    clip = p.clip ? p.clip[env] : 0;
    const int len = p.cs.clip[clip].len;
    n_steps = p.frame ? p.frame[env] : randint_below(rnd.x, len);
  } else {  // DPCombinedEnv.reset
    const int coin = p.coin ? p.coin[env] : int(rnd.y & 1u);
    if (coin == 0) {
      clip = C_WALK;
      const int len = p.cs.clip[C_WALK].len;
      n_steps = kAmnestySteps + 10 + (p.frame ? p.frame[env] : randint_below(rnd.x, len));
    } else if (coin == 1) {
      clip = C_GETUP;
      const int len = p.cs.clip[C_GETUP].len;
      n_steps = p.frame ? p.frame[env] : randint_below(rnd.x, len);
    } else {  // rsi = False
      clip = C_GETUP;
      n_steps = 0;
    }
    p.clip[env] = clip;
  }
  const ClipDesc cd = p.cs.clip[clip];
  const int frame = n_steps % cd.len;
  const int row_i = cd.row0 + (cd.fixed_row >= 0 ? cd.fixed_row : frame);
  const float* row = p.state_rows + (size_t)row_i * (Rb::NQ + Rb::NV);
  p.idx[env] = n_steps;
  p.ep_len[env] = 0;
  if (p.ep_return) p.ep_return[env] = 0.0f;
  const long long e = env;
  if (p.qpos) {
    float* q = p.qpos + (p.aos ? e * Rb::NQ : e);
    for (int j = 0; j < Rb::NQ; ++j) q[j * p.es] = row[j];
  }
  if (p.qvel) {
    float* v = p.qvel + (p.aos ? e * Rb::NV : e);
    for (int j = 0; j < Rb::NV; ++j) v[j * p.es] = row[Rb::NQ + j];
  }
  ResetIO io{row, p.obs + (p.aos ? e * OL::DIM : e), p.es, Rb::NQ};
  if (p.terminal_obs) {
    float* t = p.terminal_obs + (p.aos ? e * OL::DIM : e);
    for (int j = 0; j < OL::DIM; ++j) t[j * p.es] = io.ob[j * p.es];
  }
  FkOut<float> fk;
  ResetSink sink{p.cs.contact_from_fk != 0};
  Rb::template fk<float>(io, fk, sink);
  float root_rpy[3];
  dm_rpy_raw(io.qpos(3), io.qpos(4), io.qpos(5), io.qpos(6), root_rpy[0], root_rpy[1], root_rpy[2]);
  float omax, omin;
  const uint32_t contact = p.cs.contact_from_fk ? fk.contact : (p.contact ? p.contact[env] : 0u);
  write_obs<KIND, Rb>(io, fk, root_rpy, contact, frame, cd.len, clip, omax, omin);
}

// ---- K4: statistics ----------------------------------------------------------------------------
// acc = [count, sum r, sum r^2, n_done, sum ep_return(done), sum ep_len(done), sum obs[D], sum obs^2[D]]
struct StatsParams {
  int n;
  int obs_dim;
  long long es;
  int aos;
  const float* obs;
  const float* reward;
  const uint8_t* done;
  const float* ep_return;
  const int* ep_len;
  double* acc;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// blockDim = (32, 8): threadIdx.x walks environments, each block reduces a slab of environments
static __global__ void __launch_bounds__(256) stats_kernel(const StatsParams p) {
  const int lane = threadIdx.x, wy = threadIdx.y;
  const int per_block = (p.n + gridDim.x - 1) / gridDim.x;
  const int e0 = blockIdx.x * per_block, e1 = min(p.n, e0 + per_block);
  if (wy == 0) {  // scalars
    double c = 0, sr = 0, srr = 0, nd = 0, ser = 0, sel = 0;
    for (int e = e0 + lane; e < e1; e += 32) {
      const double r = p.reward[e];
      c += 1.0; sr += r; srr += r * r;
      if (p.done[e]) {
        nd += 1.0;
        if (p.ep_return) ser += p.ep_return[e];
        sel += p.ep_len[e];
      }
    }
    c = warp_sum(c); sr = warp_sum(sr); srr = warp_sum(srr); nd = warp_sum(nd); ser = warp_sum(ser); sel = warp_sum(sel);
    if (lane == 0) {
      atomicAdd(p.acc + 0, c); atomicAdd(p.acc + 1, sr); atomicAdd(p.acc + 2, srr);
      atomicAdd(p.acc + 3, nd); atomicAdd(p.acc + 4, ser); atomicAdd(p.acc + 5, sel);
    }
  }
  if (p.aos) {  // lanes walk columns of one row (coalesced), warps walk rows
    for (int j0 = 0; j0 < p.obs_dim; j0 += 32) {
      const int j = j0 + lane;
      double s = 0, ss = 0;
      if (j < p.obs_dim)
        for (int e = e0 + wy; e < e1; e += 8) {
          const double v = p.obs[(long long)e * p.obs_dim + j];
          s += v; ss += v * v;
        }
      if (j < p.obs_dim) {
        atomicAdd(p.acc + 6 + j, s);
        atomicAdd(p.acc + 6 + p.obs_dim + j, ss);
      }
    }
  } else {  // SoA: lanes walk environments of one column
    for (int j = wy; j < p.obs_dim; j += 8) {
      double s = 0, ss = 0;
      for (int e = e0 + lane; e < e1; e += 32) {
        const double v = p.obs[j * p.es + e];
        s += v; ss += v * v;
      }
      s = warp_sum(s); ss = warp_sum(ss);
      if (lane == 0) {
        atomicAdd(p.acc + 6 + j, s);
        atomicAdd(p.acc + 6 + p.obs_dim + j, ss);
      }
    }
  }
}

}  // namespace dm
