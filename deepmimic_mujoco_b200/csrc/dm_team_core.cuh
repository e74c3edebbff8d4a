// One environment split over the NP = 4 warps of a TEAM (K1c): per-part accumulator, the partial sums the parts hand
// to the combine step, and the combine step itself.  Plain C++ apart from DM_HD, so tests/host_harness can run a team's
// arithmetic part by part on the CPU against the oracle.
//
// Reference arithmetic restated here (and nowhere changed): calc_imitation_reward (src/deepmimic_env.py:193-256) with its
// sums split by limb, get_obs (:33-173), DPEnv.step / DPCombinedEnv.step control flow (:418-476,
// src/combined_env.py:326-485) through step_logic / reward_terms / write_obs_tail of dm_step.cuh.
#pragma once
#include "dm_step.cuh"

namespace dm {

template <class Rb>
struct PartsOf;  // generated/parts_<robot>.cuh

// What a part hands over: its share of sum(m xpos) and sum(R m ipos), of the pose / velocity / end-effector error sums,
// the largest joint error, max / min of the observation entries it wrote, and one word = contact bits | limit count << 16.
constexpr int kPartVals = 13;
struct PartSums {
  float s1x, s1y, s1z, s2x, s2y, s2z, e_q, e_v, max_dq, e_ee, omax, omin;
  uint32_t word;
};

// Per-part accumulator handed to the generated part code.  `Out` writes observation entries (out.obs(j, v)).
template <int KIND, class Rb, int P, class Out>
struct PartAcc {
  using RL = RowLayout<Rb>;
  using PT = typename PartsOf<Rb>::type;
  static constexpr int K0 = PT::k0(P), K1 = PT::k1(P), C0 = K0 / 4, NC = (K1 + 3) / 4 - C0;
  float s1x = 0.0f, s1y = 0.0f, s1z = 0.0f, s2x = 0.0f, s2y = 0.0f, s2z = 0.0f;
  float e_q = 0.0f, e_v = 0.0f, max_dq = 0.0f, e_ee = 0.0f, omax = -INFINITY, omin = INFINITY;
  int nlim = 0;
  uint32_t contact = 0u;
  bool contact_fk = false;
  Out& out;
  const float* __restrict__ row;  // this environment's packed mocap row (RowLayout)
  Vec4 tq[NC], tv[NC];            // the row's q* / v* chunks that cover this part's joints
  DM_HD PartAcc(Out& o, const float* r, bool cfk) : contact_fk(cfk), out(o), row(r) {}
  DM_HD void load_row() {
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      tq[c] = ld4(row + RL::Q + 4 * (C0 + c));
      tv[c] = ld4(row + RL::V + 4 * (C0 + c));
    }
  }
  DM_HD bool want_contact() const { return contact_fk; }
  DM_HD void put(int j, float v) {
    out.obs(j, v);
    omax = dm_max_nan(omax, v);
    omin = dm_min_nan(omin, v);
  }
  // hinge J of the joint block (qpos[7 + J], qvel[6 + J]); K = its index in the reward's joint selection or -1
  template <int J, int K>
  DM_HD void joint(float q, float v) {
    put(J, q);
    put(Rb::NJ + J, v * 0.1f);
    if constexpr (K >= 0) {
      static_assert(K >= K0 && K < K1, "joint outside the part's selection range");
      const float d = fabsf(q - tq[K / 4 - C0].v[K % 4]);
      e_q += d;
      max_dq = dm_max_nan(max_dq, d);
      e_v += fabsf(tv[K / 4 - C0].v[K % 4] - v);
      nlim += (q <= Rb::lim_lo(K)) + (q >= Rb::lim_hi(K));
    }
  }
  template <int KE>
  DM_HD void ee(float x, float y, float z) {  // deepmimic_env.py:229-233
    const float* t = row + RL::EE + 3 * KE;
    const float dx = x - t[0], dy = y - t[1], dz = z - t[2];
    e_ee = fmaf(dx, dx, e_ee);
    e_ee = fmaf(dy, dy, e_ee);
    e_ee = fmaf(dz, dz, e_ee);
  }
  DM_HD PartSums sums() const {
    return PartSums{s1x, s1y, s1z, s2x, s2y, s2z, e_q, e_v, max_dq, e_ee, omax, omin, contact | (uint32_t(nlim) << 16)};
  }
};

DM_HD void part_merge(PartSums& a, const PartSums& b) {
  a.s1x += b.s1x; a.s1y += b.s1y; a.s1z += b.s1z;
  a.s2x += b.s2x; a.s2y += b.s2y; a.s2z += b.s2z;
  a.e_q += b.e_q; a.e_v += b.e_v; a.e_ee += b.e_ee;
  a.max_dq = dm_max_nan(a.max_dq, b.max_dq);
  a.omax = dm_max_nan(a.omax, b.omax);
  a.omin = dm_min_nan(a.omin, b.omin);
  a.word += b.word;  // disjoint contact bits, limit counts add (<= 2 NSEL < 2^16)
}

// What the combine step reads of the raw state (copied to registers before the team releases the input slot)
struct RootIn {
  float v[3];  // qvel[0:3], the free joint's linear velocity
  DM_HD float qvel(int j) const { return v[j]; }
};

// The combine step: one thread, after every part of its environment has delivered.  `all` = merged PartSums,
// root_rpy = roll / pitch / yaw of the free-joint quaternion (dm_rpy_raw; computed by part PT::RPY_PART's warp),
// `tail` = the frame's row from RowLayout::ROLL on.  Writes the computed observation entries through out.obs and
// returns whether the observation-bounds guard fired (the caller zeroes the row).
template <int KIND, class Rb, bool WANT_TERMS, class Out>
DM_HD bool team_combine(Out& out, const typename PartsOf<Rb>::type::Keep& keep, const RootIn& rin, const float* root_rpy,
                        const PartSums& all, const float* __restrict__ tail, const ClipDesc& cd, int frame, bool contact_fk, uint32_t contact,
                        EnvState& st, StepOut& so) {
  using PT = typename PartsOf<Rb>::type;
  const int clip0 = st.clip;
  const float inv_m = PT::inv_mass();
  const float com_xpos[3] = {all.s1x * inv_m, all.s1y * inv_m, all.s1z * inv_m};
  const float com[3] = {(all.s1x + all.s2x) * inv_m, (all.s1y + all.s2y) * inv_m, (all.s1z + all.s2z) * inv_m};
  float cvel[6];
  PT::torso_cvel(rin, keep, com, cvel);
  float roll, pitch, yaw;
  if (Rb::kTorsoIsRoot) {
    roll = root_rpy[0]; pitch = root_rpy[1]; yaw = root_rpy[2];
  } else {  // to_rpy of the torso frame read off its rotation matrix (py3dtf formulae in matrix entries)
    roll = dm_atan2_f(keep.t21, keep.t22);
    float sp = -keep.t20;
    sp = sp > 1.0f ? 1.0f : (sp < -1.0f ? -1.0f : sp);
    pitch = dm_atan2_f(sp, sqrtf(fmaf(-sp, sp, 1.0f)));
    yaw = dm_atan2_f(keep.t10, keep.t00);
  }
  if (contact_fk) contact = all.word & 0xffffu;
  float mx = all.omax, mn = all.omin;
  write_obs_tail<KIND, Rb>(out, roll, pitch, yaw, cvel, contact, frame, cd.len, clip0, mx, mn);
  RewardMid mid;
  const float imit = reward_terms<Rb, WANT_TERMS>(all.e_q, all.e_v, all.max_dq, int(all.word >> 16), all.e_ee, root_rpy[0],
                                                  root_rpy[1], tail, com_xpos, so, mid);
  step_logic<KIND, Rb>(cd, frame, clip0, imit, mid, com[2], rin.v[0], rin.v[1], st, so);
  // observation bounds guard (deepmimic_env.py:465-476, combined_env.py:474-485), after the counters advanced
  if (mx > 100.0f || mn < -100.0f) {
    so.reward = 0.0f;
    so.done = 1;
    so.reason = R_OBS_BOUNDS;
    return true;
  }
  return false;
}

// ---- a whole team on one thread (tests/host_harness; also the specification of what the kernel's warps do) ---------
template <int KIND, class Rb, bool WANT_TERMS, class In, class Out>
DM_HD void team_step_serial(const In& in, Out& out, const ClipSet& cs, EnvState& st, uint32_t contact, StepOut& so) {
  using RL = RowLayout<Rb>;
  using OL = ObsLayout<KIND, Rb>;
  using PT = typename PartsOf<Rb>::type;
  const ClipDesc cd = cs.clip[st.clip];
  const int frame = (KIND == KIND_SINGLE) ? st.idx : (st.idx % cd.len);
  const float* row = cs.rows + size_t(cd.row0 + (cd.fixed_row >= 0 ? cd.fixed_row : frame)) * RL::STRIDE;
  typename PT::Keep keep{};
  const bool cfk = cs.contact_from_fk != 0;
  PartSums all;
  {
    PartAcc<KIND, Rb, 0, Out> a(out, row, cfk);
    a.load_row();
    PT::template part<0>(in, a, keep);
    all = a.sums();
  }
  {
    PartAcc<KIND, Rb, 1, Out> a(out, row, cfk);
    a.load_row();
    PT::template part<1>(in, a, keep);
    part_merge(all, a.sums());
  }
  {
    PartAcc<KIND, Rb, 2, Out> a(out, row, cfk);
    a.load_row();
    PT::template part<2>(in, a, keep);
    part_merge(all, a.sums());
  }
  {
    PartAcc<KIND, Rb, 3, Out> a(out, row, cfk);
    a.load_row();
    PT::template part<3>(in, a, keep);
    part_merge(all, a.sums());
  }
  RootIn rin;
  for (int k = 0; k < 3; ++k) rin.v[k] = in.qvel(k);
  float root_rpy[3];  // part PT::RPY_PART's extra job in the kernel
  dm_rpy_raw(in.qpos(3), in.qpos(4), in.qpos(5), in.qpos(6), root_rpy[0], root_rpy[1], root_rpy[2]);
  float tail[20];
  for (int k = 0; k < 20; ++k) tail[k] = row[RL::ROLL + k];
  if (team_combine<KIND, Rb, WANT_TERMS>(out, keep, rin, root_rpy, all, tail, cd, frame, cfk, contact, st, so))
    for (int j = 0; j < OL::DIM; ++j) out.obs_times_zero(j);
}

}  // namespace dm
