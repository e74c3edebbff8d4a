// TEST-ONLY: compiles the kernel's per-environment arithmetic (csrc/dm_step.cuh + the generated FK)
// as plain C++ so that `pytest -m "not gpu"` can check it against the oracle on machines without a
// GPU.  Not part of libdmhot.so, not reachable from the product API -- the product has no CPU path.
#include <cstring>
#include <vector>

#include "../../deepmimic_mujoco_b200/csrc/dm_pack.h"
#include "../../deepmimic_mujoco_b200/csrc/generated/robot_humanoid3d.cuh"
#include "../../deepmimic_mujoco_b200/csrc/generated/robot_unitree_g1.cuh"
#include "../../deepmimic_mujoco_b200/csrc/dm_team_core.cuh"
#include "../../deepmimic_mujoco_b200/csrc/generated/parts_humanoid3d.cuh"
#include "../../deepmimic_mujoco_b200/csrc/generated/parts_unitree_g1.cuh"

using namespace dm;

struct ArrayIO {
  const float *qp, *qv, *ac;
  float *ob, *ta;
  float qpos(int j) const { return qp[j]; }
  float qvel(int j) const { return qv[j]; }
  float act(int j) const { return ac[j]; }
  void obs(int j, float v) { ob[j] = v; }
  void obs_times_zero(int j) { ob[j] = ob[j] * 0.0f; }
  void tau(int j, float v) { if (ta) ta[j] = v; }
};
struct DIO {
  const double *qp, *qv;
  double qpos(int j) const { return qp[j]; }
  double qvel(int j) const { return qv[j]; }
};
struct DSink {
  static constexpr bool kWantsGeoms = true;
  static constexpr bool contact_on() { return false; }
  double *bx, *gx;
  void body(int i, double x, double y, double z) { bx[3 * i] = x; bx[3 * i + 1] = y; bx[3 * i + 2] = z; }
  void geom(int g, double x, double y, double z) { gx[3 * g] = x; gx[3 * g + 1] = y; gx[3 * g + 2] = z; }
};

static int g_contact_from_fk = 0;
extern "C" void hh_set_contact_mode(int from_fk) { g_contact_from_fk = from_fk; }
static int g_team = 0;  // 1: the team kernel's arithmetic (dm_team_core.cuh + generated parts), its four parts run in turn
extern "C" void hh_set_team(int on) { g_team = on; }

template <int KIND, class Rb>
static int run(int n_clips, const dm_clip_t* clips, int n, const float* qpos, const float* qvel, const float* action,
               const uint32_t* contact, int* idx, int* ep_len, int* clip, float* obs, float* reward, uint8_t* done,
               uint8_t* reason, float* torque, float* terms) {
  ClipSet cs;
  std::vector<float> rows, srows;
  int total;
  if (pack_clip_rows<Rb>(clips, n_clips, cs, rows, srows, total)) return -1;
  cs.rows = rows.data();
  cs.contact_from_fk = g_contact_from_fk;
  using OL = ObsLayout<KIND, Rb>;
  for (int e = 0; e < n; ++e) {
    ArrayIO io{qpos + (size_t)e * Rb::NQ, qvel + (size_t)e * Rb::NV, action + (size_t)e * Rb::NA, obs + (size_t)e * OL::DIM,
               torque ? torque + (size_t)e * Rb::NU : nullptr};
    EnvState st{idx[e], ep_len[e], clip ? clip[e] : 0};
    StepOut so;
    if (g_team) {
      write_torque<Rb>(io);
      team_step_serial<KIND, Rb, true>(io, io, cs, st, contact ? contact[e] : 0u, so);
    } else {
      env_step<KIND, Rb, ArrayIO, true>(io, cs, cs.rows, st, contact ? contact[e] : 0u, so);
    }
    idx[e] = st.idx; ep_len[e] = st.ep_len;
    if (clip && KIND == KIND_COMBINED) clip[e] = st.clip;
    reward[e] = so.reward; done[e] = so.done; reason[e] = so.reason;
    if (terms) memcpy(terms + 8 * (size_t)e, so.terms, sizeof so.terms);
  }
  return 0;
}

extern "C" int hh_obs_dim(int robot, int kind) {
  if (robot == 0) return ObsLayout<KIND_SINGLE, Robot_humanoid3d>::DIM;
  return kind == 0 ? ObsLayout<KIND_SINGLE, Robot_unitree_g1>::DIM : ObsLayout<KIND_COMBINED, Robot_unitree_g1>::DIM;
}

extern "C" int hh_step(int robot, int kind, int n_clips, const dm_clip_t* clips, int n, const float* qpos, const float* qvel,
                       const float* action, const uint32_t* contact, int* idx, int* ep_len, int* clip, float* obs,
                       float* reward, uint8_t* done, uint8_t* reason, float* torque, float* terms) {
  if (robot == 0 && kind == 0)
    return run<KIND_SINGLE, Robot_humanoid3d>(n_clips, clips, n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason, torque, terms);
  if (robot == 1 && kind == 0)
    return run<KIND_SINGLE, Robot_unitree_g1>(n_clips, clips, n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason, torque, terms);
  if (robot == 1 && kind == 1)
    return run<KIND_COMBINED, Robot_unitree_g1>(n_clips, clips, n, qpos, qvel, action, contact, idx, ep_len, clip, obs, reward, done, reason, torque, terms);
  return -2;
}

template <class Rb>
static void fkt(int n, const double* qpos, const double* qvel, double* bx, double* gx) {
  for (int f = 0; f < n; ++f) {
    DIO in{qpos + (size_t)f * Rb::NQ, qvel + (size_t)f * Rb::NV};
    DSink sink{bx + (size_t)f * Rb::NBODY * 3, gx + (size_t)f * Rb::NGEOM * 3};
    sink.body(0, 0, 0, 0);
    for (int g = 0; g < Rb::NGEOM; ++g) sink.geom(g, 0, 0, 0);
    FkOut<double> o;
    Rb::template fk<double>(in, o, sink);
  }
}
extern "C" int hh_fk_tables(int robot, int n, const double* qpos, const double* qvel, double* bx, double* gx) {
  if (robot == 0) fkt<Robot_humanoid3d>(n, qpos, qvel, bx, gx); else fkt<Robot_unitree_g1>(n, qpos, qvel, bx, gx);
  return 0;
}

// ---- CopyPool self-test (csrc/dm_copy_pool.h): parallel copies of random sizes == memcpy ----
#include "../../deepmimic_mujoco_b200/csrc/dm_copy_pool.h"
extern "C" int hh_copy_pool_selftest(int helpers, int iters, unsigned seed) {
  dm::CopyPool pool(helpers);
  std::vector<unsigned char> src(5u << 20), dst(5u << 20), ref(5u << 20);
  unsigned x = seed * 2654435761u + 1u;
  auto rnd = [&] { x ^= x << 13; x ^= x >> 17; x ^= x << 5; return x; };
  for (auto& b : src) b = (unsigned char)rnd();
  for (int it = 0; it < iters; ++it) {
    const size_t bytes = 1 + rnd() % (it % 3 == 0 ? 4000000u : 300000u), off = rnd() % 4096;
    for (size_t i = 0; i < bytes + 64; ++i) dst[off + i] = ref[off + i] = (unsigned char)(it + i);
    pool.copy(dst.data() + off, src.data() + (it % 7), bytes);
    memcpy(ref.data() + off, src.data() + (it % 7), bytes);
    if (memcmp(dst.data() + off, ref.data() + off, bytes + 64) != 0) return it + 1;  // payload and the bytes behind it
  }
  // several segments as one job (the arrays of one chunk of a host step), a zero-fill among them
  for (int it = 0; it < iters / 4; ++it) {
    dm::CopyPool::Seg segs[6];
    size_t o = rnd() % 64;
    for (size_t i = 0; i < dst.size(); ++i) dst[i] = ref[i] = (unsigned char)(it + i);
    for (int k = 0; k < 6; ++k) {
      const size_t bytes = k == 4 ? 0 : 1 + rnd() % (k == 1 ? 900000u : 90000u);
      const unsigned char* sp = k == 3 ? nullptr : src.data() + (rnd() % 1000);
      segs[k] = dm::CopyPool::Seg{dst.data() + o, sp, bytes};
      if (sp) memcpy(ref.data() + o, sp, bytes); else memset(ref.data() + o, 0, bytes);
      o += bytes + rnd() % 300;
    }
    pool.copy_many(segs, 6);
    if (memcmp(dst.data(), ref.data(), dst.size()) != 0) return 100000 + it;
  }
  return 0;
}
