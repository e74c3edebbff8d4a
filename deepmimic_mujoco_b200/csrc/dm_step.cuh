// Per-environment step logic: observation, imitation reward, termination / motion state
// machine, action -> motor torque.  One call handles ONE environment; the kernels in
// dmhot.cu map environments onto threads and provide the `IO` accessor (global memory with
// strides, or shared-memory tiles).  Reference: src/deepmimic_env.py:33-173 (get_obs),
// :193-256 (calc_imitation_reward), :335-484 (DPEnv.step); src/combined_env.py:246-505
// (DPCombinedEnv.step / _get_obs); restated for float64 in oracle/envs.py.
#pragma once
#include <type_traits>

#include "dm_math.cuh"

namespace dm {

// ---- clip tables -------------------------------------------------------------------------
// One packed fp32 row per mocap frame holding only what a step reads (all target-side
// quantities are precomputed once per frame by the table builder):
//   [0, NSEL)            q*   selected joint angles          (deepmimic_env.py:200,208)   (block padded to NSELP)
//   [NSELP, NSELP+NSEL)  v*   selected joint velocities      (:202,210)                   (block padded to NSELP)
//   2 NSELP + 0, + 1     roll*, pitch* of the root quaternion (:218-219)
//   2 NSELP + 2 .. + 13  4 end-effector geom_xpos*           (:232)
//   2 NSELP + 14 .. + 16 COM* = sum(m * body_xpos*) / sum(m) (:235-237)
//   2 NSELP + 17, + 18   root linear velocity v*_xy          (combined_env.py:344)
// Rows are 16-byte aligned and a multiple of 16 bytes long.
template <class Rb>
struct RowLayout {
  // the two joint blocks are padded to a multiple of 4 floats so that every read of a row is a 128-bit load
  static constexpr int NSELP = (Rb::NSEL + 3) / 4 * 4;
  static constexpr int Q = 0, V = NSELP, ROLL = 2 * NSELP, PITCH = ROLL + 1, EE = ROLL + 2, COM = ROLL + 14,
                       VXY = ROLL + 17, USED = ROLL + 19, STRIDE = ROLL + 20;
};

struct alignas(16) Vec4 {
  float v[4];
};
DM_HD Vec4 ld4(const float* p) { return *reinterpret_cast<const Vec4*>(p); }

enum : int { CLIP_FLOOR = 1, CLIP_ACYCLIC = 2, CLIP_G1_RUN = 4 };
enum : int { KIND_SINGLE = 0, KIND_COMBINED = 1 };
enum : int { C_WALK = 0, C_RUN = 1, C_GETUP = 2, C_TO_GETUP = 3 };
enum : int { R_NONE = 0, R_LOW_Z, R_HIGH_Z, R_RUN_ROLL, R_RUN_PITCH, R_MAX_EP, R_ACYC_END, R_FALLEN, R_OBS_BOUNDS };

struct ClipDesc {
  int row0;       // first row of this clip in the packed table
  int len;        // get_length()
  int flags;      // CLIP_*
  int fixed_row;  // >= 0: constant-pose clip (MotionTransition: always frame 1), else -1
};
constexpr int kMaxClips = 8;
struct ClipSet {
  const float* rows;
  int n_clips;
  int contact_from_fk;  // != 0: floor-contact bits come from FK (generated plane-sphere / plane-box tests), not the caller
  ClipDesc clip[kMaxClips];
};

template <int KIND, class Rb>
struct ObsLayout;
template <class Rb>
struct ObsLayout<KIND_SINGLE, Rb> {  // DPEnvConfig: torso, foot contact, phase
  static constexpr int TORSO = 2 * Rb::NJ, CONTACT = TORSO + 8, NCONTACT = 2, PHASE = CONTACT + 2, PA = PHASE + 1,
                       DIM = PHASE + 1;
  static constexpr int MAX_EP = 1000;
};
template <class Rb>
struct ObsLayout<KIND_COMBINED, Rb> {  // DPCombinedEnvConfig: torso, extra contact, phase, player action
  static constexpr int TORSO = 2 * Rb::NJ, CONTACT = TORSO + 8, NCONTACT = 8, PHASE = CONTACT + 8, PA = PHASE + 1,
                       DIM = PA + 7;
  static constexpr int MAX_EP = 2000;
};

constexpr double kDeg60 = 1.0471975511965976;  // np.deg2rad(60.)
constexpr double kDeg15 = 0.2617993877991494;  // np.deg2rad(15)
constexpr int kAmnestySteps = 150;             // combined_env.py:35

struct EnvState {
  int idx;     // single: idx_curr;  combined: current_motion_n_steps
  int ep_len;  // episode_length
  int clip;    // clip id (single: fixed per env; combined: motion state)
};

struct StepOut {
  float reward;
  float raw_reward;  // reward before the obs-bounds guard zeroed it (the reference adds this to episode_reward)
  uint8_t done;
  uint8_t reason;
  float terms[8];  // reward_config, reward_qvel, reward_end_eff, reward_com, reward_joint_limit,
                   // imitation_reward, task_reward, z_com
};

// ---- observation (get_obs) -----------------------------------------------------------------
// The entries behind the two joint blocks: torso (deepmimic_env.py:47-76), contacts (:78-121), phase (:139-143),
// player action (:145-173).  torso roll / pitch / yaw, cvel = [angular; linear] of the torso body.  mx / mn carry the
// running max / min of the observation (numpy's NaN-propagating semantics) for the bounds guard.
template <int KIND, class Rb, class IO>
DM_HD void write_obs_tail(IO& io, float roll, float pitch, float yaw, const float* cvel, uint32_t contact, int frame, int len,
                          int clip, float& mx, float& mn) {
  using L = ObsLayout<KIND, Rb>;
  auto put = [&](int j, float v) {  // `max(obs) > 100 or min(obs) < -100` is `max |obs| > 100`: one NaN-propagating max per entry
    io.obs(j, v);
    mx = dm_max_nan(mx, fabsf(v));
  };
  float sy, cy;
  dm_sincos_f(-yaw, sy, cy);
  const float vx = cy * cvel[3] - sy * cvel[4];
  const float vy = sy * cvel[3] + cy * cvel[4];
  put(L::TORSO + 0, roll * 0.1f);
  put(L::TORSO + 1, pitch * 0.1f);
  put(L::TORSO + 2, vx * 0.1f);
  put(L::TORSO + 3, vy * 0.1f);
  put(L::TORSO + 4, cvel[5] * 0.1f);
  put(L::TORSO + 5, cvel[0] * 0.1f);
  put(L::TORSO + 6, cvel[1] * 0.1f);
  put(L::TORSO + 7, cvel[2] * 0.1f);
  if (KIND == KIND_SINGLE) {  // [rfoot & floor, lfoot & floor] (:78-105)
    put(L::CONTACT + 0, (contact & 1u) ? 1.0f : 0.0f);
    put(L::CONTACT + 1, (contact & 2u) ? 1.0f : 0.0f);
  } else {  // eight extra-contact geoms & floor (:107-121)
#pragma unroll
    for (int k = 0; k < 8; ++k) put(L::CONTACT + k, ((contact >> (2 + k)) & 1u) ? 1.0f : 0.0f);
  }
  float ph = float(frame) / float(len);  // :139-143
  ph = ph < 0.0f ? 0.0f : (ph > 1.0f ? 1.0f : ph);
  put(L::PHASE, ph);
  if (KIND == KIND_COMBINED) {  // :145-173 with PAWalk heading (1, 0, 0): hx = cos(-yaw), hy = sin(-yaw)
    put(L::PA + 0, cy);
    put(L::PA + 1, sy);
    put(L::PA + 2, 1.0f);  // onehot('walk'); PA resampling is disabled (combined_env.py:303)
    put(L::PA + 3, 0.0f);
    put(L::PA + 4, 0.0f);
    put(L::PA + 5, clip == C_TO_GETUP ? 1.0f : 0.0f);  // pa_getup_state (combined_env.py:499-504)
    put(L::PA + 6, clip == C_GETUP ? 1.0f : 0.0f);
  }
}

// Writes the DIM observation values through io.obs(j, v) and returns max / min with numpy's
// NaN-propagating semantics for the bounds guard.
template <int KIND, class Rb, class IO>
DM_HD void write_obs(IO& io, const FkOut<float>& fk, const float* root_rpy, uint32_t contact, int frame, int len,
                     int clip, float& omax, float& omin) {
  float mx = -INFINITY, mn = INFINITY;
  auto put = [&](int j, float v) {  // `max(obs) > 100 or min(obs) < -100` is `max |obs| > 100`: one NaN-propagating max per entry
    io.obs(j, v);
    mx = dm_max_nan(mx, fabsf(v));
  };
#pragma unroll
  for (int j = 0; j < Rb::NJ; ++j) put(j, io.qpos(7 + j));
#pragma unroll
  for (int j = 0; j < Rb::NJ; ++j) put(Rb::NJ + j, io.qvel(6 + j) * 0.1f);
  float roll, pitch, yaw;
  if (Rb::kTorsoIsRoot) {  // torso body carries the free joint: its xquat is the normalised qpos[3:7]
    roll = root_rpy[0]; pitch = root_rpy[1]; yaw = root_rpy[2];
  } else {
    dm_rpy(fk.tq[0], fk.tq[1], fk.tq[2], fk.tq[3], roll, pitch, yaw);
  }
  write_obs_tail<KIND, Rb>(io, roll, pitch, yaw, fk.cvel, contact, frame, len, clip, mx, mn);
  omax = mx;
  omin = mn;
}

// ---- action -> motor torque (deepmimic_env.py:347-352 + MuJoCo <motor>) ----------------------
template <class Rb, class IO>
DM_HD void write_torque(IO& io) {
#pragma unroll
  for (int k = 0; k < Rb::NU; ++k) {
    float t = 0.0f;
    if (k < Rb::NA) {
      float c = io.act(k) * Rb::kActScale;
      // mju_clip(ctrl, lo, hi) = max(lo, min(hi, ctrl)); NaN passes through
      c = dm_max_nan(dm_min_nan(c, Rb::ctrl_hi(k)), Rb::ctrl_lo(k));
      t = Rb::gear(k) * c;
    }
    io.tau(k, t);
  }
}

// ---- imitation reward (calc_imitation_reward) ---------------------------------------------
struct RewardMid {
  float sum_dq, max_dq, d_roll, d_pitch;
  float vx_ref, vy_ref;  // v*_xy of the frame (task reward of the combined env)
};

// FK sink of the per-step path: the squared end-effector error (deepmimic_env.py:229-233) is accumulated the
// moment each end-effector position exists, instead of keeping 12 coordinates live until the reward (at 128
// registers per thread those were exactly the values ptxas spilled).
template <class Rb>
struct EeSink {
  static constexpr bool kWantsGeoms = false;
  const float* __restrict__ tgt;  // the frame's 4 x 3 target positions (RowLayout::EE)
  float e2;
  bool want_contact;  // run the floor-contact tests of the FK (DM_CONTACT_FROM_FK)
  DM_HD bool contact_on() const { return want_contact; }
  template <typename T> DM_HD void body(int, T, T, T) {}
  DM_HD void geom(int g, float x, float y, float z) {  // g is a literal in the generated code
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (g == Rb::ee_geom(k)) {
        const float dx = x - tgt[3 * k], dy = y - tgt[3 * k + 1], dz = z - tgt[3 * k + 2];
        e2 = fmaf(dx, dx, e2);
        e2 = fmaf(dy, dy, e2);
        e2 = fmaf(dz, dz, e2);
      }
  }
};

// The reward once the per-joint sums exist (deepmimic_env.py:213-256): `tail` = the frame's row from RowLayout::ROLL on
// (roll*, pitch*, ee*[12], com*[3], vxy*[2]).  Shared by the thread-per-environment step and the team kernel's combine.
template <class Rb, bool WANT_COM>
DM_HD float reward_terms(float e_q, float e_v, float max_dq, int nlim, float e_ee, float roll, float pitch,
                         const float* __restrict__ tail, const float* com_xpos, StepOut& so, RewardMid& mid) {
  mid.vx_ref = tail[17];
  mid.vy_ref = tail[18];
  mid.sum_dq = e_q;
  mid.max_dq = max_dq;
  mid.d_roll = fabsf(roll - tail[0]);
  mid.d_pitch = fabsf(pitch - tail[1]);
  const float r_pose = expf(-(e_q + mid.d_pitch));
  const float r_vel = expf(-0.1f * e_v);
  const float r_ee = expf(-40.0f * e_ee);
  float r_com = 0.0f;
  if (WANT_COM) {
    const float dx = tail[14] - com_xpos[0], dy = tail[15] - com_xpos[1], dz = tail[16] - com_xpos[2];
    r_com = expf(-10.0f * (dx * dx + dy * dy + dz * dz));
  }
  const float qlim = float(nlim) / float(Rb::NSEL);
  so.terms[0] = r_pose;
  so.terms[1] = r_vel;
  so.terms[2] = r_ee;
  so.terms[3] = r_com;
  so.terms[4] = qlim;
  // wp 0.75, wv 0.1, we 0.15, wc 0.0, wj -0.1 (deepmimic_env.py:400-404); the wc term is exactly 0
  return 0.75f * r_pose + 0.1f * r_vel + 0.15f * r_ee + (-0.1f) * qlim;
}

template <class Rb, class IO, bool WANT_COM>
DM_HD float imitation_reward(const IO& io, const FkOut<float>& fk, const float* root_rpy, float e_ee,
                             const float* __restrict__ row, StepOut& so, RewardMid& mid) {
  using RL = RowLayout<Rb>;
  float e_q = 0.0f, e_v = 0.0f, mx = 0.0f;
  int nlim = 0;
#pragma unroll
  for (int c = 0; c < RL::NSELP / 4; ++c) {
    const Vec4 tq = ld4(row + RL::Q + 4 * c), tv = ld4(row + RL::V + 4 * c);
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int k = 4 * c + u;
      if (k < Rb::NSEL) {
        const float q = io.qpos(Rb::sel_q(k));
        const float d = fabsf(q - tq.v[u]);
        e_q += d;
        mx = dm_max_nan(mx, d);
        e_v += fabsf(tv.v[u] - io.qvel(Rb::sel_q(k) - 1));
        nlim += (q <= Rb::lim_lo(k) || q >= Rb::lim_hi(k)) ? 1 : 0;  // lo < hi: at most one of the two holds
      }
    }
  }
  Vec4 tail[5];  // roll*, pitch*, ee*[12], com*[3], vxy*[2], pad
#pragma unroll
  for (int c = 0; c < 5; ++c) tail[c] = ld4(row + RL::ROLL + 4 * c);
  return reward_terms<Rb, WANT_COM>(e_q, e_v, mx, nlim, e_ee, root_rpy[0], root_rpy[1], &tail[0].v[0], fk.com_xpos, so, mid);
}

// accessors that emit the observation themselves (the warp kernel assembles it in shared memory) declare
// `static constexpr bool kZeroOnGuard = false` and zero the row on R_OBS_BOUNDS when they write it
template <class IO, class = void>
struct io_zeroes_on_guard : std::true_type {};
template <class IO>
struct io_zeroes_on_guard<IO, std::enable_if_t<!IO::kZeroOnGuard>> : std::false_type {};

// ---- everything after the reward sums: task reward, termination, motion state machine, counters ------------------
// DPEnv.step (deepmimic_env.py:418-455) / DPCombinedEnv.step (combined_env.py:326-462); `z` = z of the mass-weighted
// xipos COM, qv0 / qv1 = root linear velocity (qvel[0:2]).  Shared by the thread-per-environment step and the team kernel.
template <int KIND, class Rb>
DM_HD void step_logic(const ClipDesc& cd, int frame, int clip0, float imit, const RewardMid& mid, float z, float qv0, float qv1,
                      EnvState& st, StepOut& so) {
  using OL = ObsLayout<KIND, Rb>;
  so.terms[7] = z;
  bool done = false;
  int reason = R_NONE;
  float reward;

  if (KIND == KIND_SINGLE) {
    reward = imit;
    so.terms[5] = imit;
    so.terms[6] = 0.0f;
    if (!(cd.flags & CLIP_FLOOR)) {  // deepmimic_env.py:420-424 (reason is written even when not done)
      const bool low = double(z) < double(Rb::kLowZd());
      done = low || (z > 2.0f);
      reason = low ? R_LOW_Z : R_HIGH_Z;
    }
    if (cd.flags & CLIP_G1_RUN) {  // :426-433
      if (double(mid.d_roll) > kDeg60) { done = true; reason = R_RUN_ROLL; }
      if (double(mid.d_pitch) > kDeg60) { done = true; reason = R_RUN_PITCH; }
    }
    if (st.ep_len >= OL::MAX_EP) { done = true; reason = R_MAX_EP; }                                    // :435-438
    if (frame + 1 == cd.len && (cd.flags & CLIP_ACYCLIC)) { done = true; reason = R_ACYC_END; }         // :440-442
    st.idx = (frame + 1) % cd.len;                                                                      // :452
    st.ep_len += 1;
  } else {
    float task = 0.0f, im = imit;
    if (clip0 == C_WALK || clip0 == C_RUN) {  // combined_env.py:340-347
      const float dx = mid.vx_ref - qv0, dy = mid.vy_ref - qv1;
      task = expf(-10.0f * sqrtf(dx * dx + dy * dy));
    }
    if (clip0 == C_TO_GETUP) {  // :348-352
      im = 0.0f;
      task = expf(-(mid.sum_dq + mid.d_pitch + mid.d_roll) / 5.0f) / 3.0f;
    }
    reward = im * 0.7f + task * 0.3f;
    so.terms[5] = im;
    so.terms[6] = task;
    int clip = clip0, n = st.idx;
    if (n >= cd.len - 1) {                       // :399-404
      if (clip == C_GETUP) { clip = C_RUN; n = 0; }  // Q3: identity comparison -> always the run clip
      if (clip == C_TO_GETUP) { clip = C_GETUP; n = 0; }
    }
    const bool ok = double(mid.d_pitch) < kDeg15 && double(mid.d_roll) < kDeg15 && double(mid.max_dq) < kDeg15;
    if (ok && clip == C_TO_GETUP) { clip = C_GETUP; n = 0; }  // :418-420
    if (clip == C_WALK || clip == C_RUN) {                    // :422-440
      bool fallen = double(z) < double(Rb::kLowZd()) || z > 2.0f;
      if (double(mid.d_roll) > kDeg60 || double(mid.d_pitch) > kDeg60) fallen = true;
      if (fallen) {
        if (!(n > kAmnestySteps)) { done = true; reason = R_FALLEN; }
        clip = C_TO_GETUP;
        n = 0;
      }
    }
    if (st.ep_len >= OL::MAX_EP) { done = true; reason = R_MAX_EP; }  // :442-445
    st.clip = clip;
    st.idx = n + 1;                                                    // :458-459
    st.ep_len += 1;
  }
  so.reward = reward;
  so.raw_reward = reward;
  so.done = done ? 1 : 0;
  so.reason = uint8_t(reason);
}

// ---- one environment step --------------------------------------------------------------------
// ROW_GIVEN: `rows` already points at THIS environment's packed row (the tile kernel staged it in shared memory)
template <int KIND, class Rb, class IO, bool WANT_TERMS, bool ROW_GIVEN = false>
DM_HD void env_step(IO& io, const ClipSet& cs, const float* __restrict__ rows, EnvState& st, uint32_t contact, StepOut& so) {
  using RL = RowLayout<Rb>;
  using OL = ObsLayout<KIND, Rb>;
  const int clip0 = st.clip;
  const ClipDesc cd = cs.clip[clip0];
  const int frame = (KIND == KIND_SINGLE) ? st.idx : (st.idx % cd.len);
  const int row_i = cd.row0 + (cd.fixed_row >= 0 ? cd.fixed_row : frame);
  const float* __restrict__ row = ROW_GIVEN ? rows : rows + size_t(row_i) * RL::STRIDE;

  write_torque<Rb>(io);

  FkOut<float> fk;
  EeSink<Rb> sink{row + RL::EE, 0.0f, cs.contact_from_fk != 0};
  Rb::template fk<float>(io, fk, sink);

  if (cs.contact_from_fk) contact = fk.contact;
  float root_rpy[3];
  dm_rpy_raw(io.qpos(3), io.qpos(4), io.qpos(5), io.qpos(6), root_rpy[0], root_rpy[1], root_rpy[2]);
  float omax, omin;
  write_obs<KIND, Rb>(io, fk, root_rpy, contact, frame, cd.len, clip0, omax, omin);

  RewardMid mid;
  const float imit = imitation_reward<Rb, IO, WANT_TERMS>(io, fk, root_rpy, sink.e2, row, so, mid);
  step_logic<KIND, Rb>(cd, frame, clip0, imit, mid, fk.com[2], io.qvel(0), io.qvel(1), st, so);
  // observation bounds guard (deepmimic_env.py:465-476, combined_env.py:474-485): zero observation,
  // zero reward, done, empty info -- evaluated after the counters advanced
  if (omax > 100.0f || omin < -100.0f) {
    if constexpr (io_zeroes_on_guard<IO>::value) {
#pragma unroll 1
      for (int j = 0; j < OL::DIM; ++j) io.obs_times_zero(j);  // `_get_obs() * 0.`: an infinite entry becomes NaN
    }
    so.reward = 0.0f;
    so.done = 1;
    so.reason = R_OBS_BOUNDS;
  }
}

}  // namespace dm
