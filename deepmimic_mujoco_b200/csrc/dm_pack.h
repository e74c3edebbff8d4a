// Host-side packing of the mocap tables (float64, as MocapDM builds them) into the fp32 rows the
// step kernel reads (dm_step.cuh: RowLayout).  Plain C++; shared by libdmhot.so and the CPU unit-test
// harness under tests/host_harness.
#pragma once
#include <vector>

#include "../../include/dmhot.h"
#include "dm_step.cuh"

namespace dm {

// returns NULL on success, else a static message
template <class Rb>
const char* pack_clip_rows(const dm_clip_t* clips, int n_clips, ClipSet& cs, std::vector<float>& rows,
                           std::vector<float>& srows, int& total) {
  using RL = RowLayout<Rb>;
  cs = ClipSet{};
  cs.n_clips = n_clips;
  total = 0;
  for (int c = 0; c < n_clips; ++c) {
    const dm_clip_t& k = clips[c];
    if (k.len <= 0 || k.n_rows <= 0 || !k.qpos || !k.qvel || !k.body_xpos || !k.geom_xpos) return "a clip is empty or has NULL tables";
    if (k.fixed_row >= k.n_rows || (k.fixed_row < 0 && k.n_rows != k.len)) return "n_rows inconsistent with len / fixed_row";
    cs.clip[c] = ClipDesc{total, k.len, k.flags, k.fixed_row};
    total += k.n_rows;
  }
  rows.assign((size_t)total * RL::STRIDE, 0.0f);
  srows.assign((size_t)total * (Rb::NQ + Rb::NV), 0.0f);
  double mtot = 0;
  for (int i = 0; i < Rb::NBODY; ++i) mtot += Rb::body_mass(i);
  for (int c = 0; c < n_clips; ++c) {
    const dm_clip_t& k = clips[c];
    for (int f = 0; f < k.n_rows; ++f) {
      const double* q = k.qpos + (size_t)f * Rb::NQ;
      const double* v = k.qvel + (size_t)f * Rb::NV;
      float* r = rows.data() + (size_t)(cs.clip[c].row0 + f) * RL::STRIDE;
      for (int s = 0; s < Rb::NSEL; ++s) {
        r[RL::Q + s] = (float)q[Rb::sel_q(s)];
        r[RL::V + s] = (float)v[Rb::sel_q(s) - 1];
      }
      // to_rpy of the (possibly un-normalised, lerped) table quaternion, deepmimic_env.py:218-219
      double w = q[3], x = q[4], y = q[5], z = q[6], roll, pitch;
      dm_normalize4(w, x, y, z);
      dm_roll_pitch(w, x, y, z, roll, pitch);
      r[RL::ROLL] = (float)roll;
      r[RL::PITCH] = (float)pitch;
      const double* gx = k.geom_xpos + (size_t)f * Rb::NGEOM * 3;
      for (int e = 0; e < 4; ++e)
        for (int a = 0; a < 3; ++a) r[RL::EE + 3 * e + a] = (float)gx[3 * Rb::ee_geom(e) + a];
      const double* bx = k.body_xpos + (size_t)f * Rb::NBODY * 3;
      for (int a = 0; a < 3; ++a) {  // target COM, deepmimic_env.py:235-237
        double s = 0;
        for (int i = 0; i < Rb::NBODY; ++i) s += Rb::body_mass(i) * bx[3 * i + a];
        r[RL::COM + a] = (float)(s / mtot);
      }
      r[RL::VXY + 0] = (float)v[0];  // combined_env.py:344
      r[RL::VXY + 1] = (float)v[1];
      float* sr = srows.data() + (size_t)(cs.clip[c].row0 + f) * (Rb::NQ + Rb::NV);
      for (int j = 0; j < Rb::NQ; ++j) sr[j] = (float)q[j];
      for (int j = 0; j < Rb::NV; ++j) sr[Rb::NQ + j] = (float)v[j];
    }
  }
  return nullptr;
}

}  // namespace dm
