#!/usr/bin/env python3
"""Model compiler back end for the team kernel (K1c): one environment split over four warps.

Input : deepmimic_mujoco_b200/assets/models.json (compiled by mjcf.py from the reference's MJCF).
Output: deepmimic_mujoco_b200/csrc/generated/parts_<name>.cuh

The kinematic tree is cut into NP = 4 PARTS (limbs).  Warp p of a team runs part p for the team's 32 environments
(lane = environment): the forward kinematics of the part's bodies, the pose / velocity error sums, joint-limit counts and
observation entries of the part's joints, its end effectors and floor contacts, its share of the two centre-of-mass sums.
A part that hangs off the trunk recomputes the trunk bodies above it (root, chest / torso: a few dozen instructions)
instead of waiting for another warp.  What the reference gets from MuJoCo per body (mj_kinematics / mj_comPos / mj_comVel;
call sites src/deepmimic_env.py:29,51-54,232,236,357,421; algorithm restated in oracle/mjkin.py) is computed here with
ROTATION MATRICES along the chain,

    R_body = R_parent * C_body * Rot_axis(theta)        p_body = p_parent + R_parent * pos_body

(C = the constant body_quat as a matrix, zeros skipped at generation time), 12 multiply-adds per hinge where the
quaternion chain of robot_<name>.cuh (compose, normalise, quaternion -> matrix) needs ~40.  Same function in exact
arithmetic; fp32 round-off differs at the 1e-7 level (the parity tests hold both to the same 1e-5 bar).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, REPO)
sys.path.insert(0, HERE)
from deepmimic_mujoco_b200.mjcf import load_models  # noqa: E402
from gen_model_code import ROBOT_CFG, Emit  # noqa: E402

# part -> bodies it OWNS (COM sums, joint terms, observation entries, end effectors, contacts).  Part 0 is the team's
# combine warp: it owns the root and the chain to the torso body (whose cvel the observation needs).
PARTS = {
    "humanoid3d": [["root", "chest", "neck"],
                   ["right_shoulder", "right_elbow", "left_shoulder", "left_elbow"],
                   ["right_hip", "right_knee", "right_ankle"],
                   ["left_hip", "left_knee", "left_ankle"]],
    "unitree_g1": [["pelvis", "left_hip_pitch_link", "left_hip_roll_link", "left_hip_yaw_link", "left_knee_link",
                    "left_ankle_pitch_link", "left_ankle_roll_link"],
                   ["right_hip_pitch_link", "right_hip_roll_link", "right_hip_yaw_link", "right_knee_link",
                    "right_ankle_pitch_link", "right_ankle_roll_link", "torso_link"],
                   ["left_shoulder_pitch_link", "left_shoulder_roll_link", "left_shoulder_yaw_link", "left_elbow_pitch_link",
                    "left_elbow_roll_link", "left_zero_link", "left_one_link", "left_two_link", "left_three_link",
                    "left_four_link", "left_five_link", "left_six_link"],
                   ["right_shoulder_pitch_link", "right_shoulder_roll_link", "right_shoulder_yaw_link", "right_elbow_pitch_link",
                    "right_elbow_roll_link", "right_zero_link", "right_one_link", "right_two_link", "right_three_link",
                    "right_four_link", "right_five_link", "right_six_link"]],
}


RPY_PART = {"humanoid3d": 3, "unitree_g1": 1}  # the lightest part that is not the combine warp's


def flit(x):
    s = repr(float(np.float32(x)))
    return s + "f"


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([[w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def lincomb(terms, start=None):
    """fmaf chain for start + sum(name * const); terms = [(name, const)], zero constants already dropped."""
    ex = start
    for name, c in terms:
        if ex is None:
            ex = "%s * %s" % (name, flit(c))
        else:
            ex = "fmaf(%s, %s, %s)" % (name, flit(c), ex)
    return ex if ex is not None else "0.0f"


class PartEmitter:
    def __init__(self, name, m, cfg, e):
        self.name, self.m, self.cfg, self.e = name, m, cfg, e
        self.torso = m.body_names.index(cfg["torso"])
        self.ee_geoms = [m.geom_names.index(n) for n in cfg["ee"]]
        chain, b = [], self.torso
        while b != 0:
            chain.append(b)
            b = int(m.body_parentid[b])
        self.chain = chain[::-1]
        assert self.chain[0] == 1 and m.jnt_type[m.body_jntadr[1]] == 0, "body 1 must carry the free joint"
        self.sel_q = list(range(7, m.nq)) if name == "humanoid3d" else list(range(7, 25)) + list(range(32, 37))
        X = m.extra
        fl = m.geom_names.index(cfg["floor"])
        self.contact_slot = {}
        for bit, gname in enumerate([cfg["rfoot"], cfg["lfoot"]] + list(cfg["extra"])):
            g = m.geom_names.index(gname)
            collide = (m.geom_contype[fl] & m.geom_conaffinity[g]) or (m.geom_contype[g] & m.geom_conaffinity[fl])
            if collide:
                assert not X["geom_has_quat"][g] and X["geom_type"][g] in ("sphere", "box"), gname
                self.contact_slot[g] = (bit, max(X["geom_margin"][fl], X["geom_margin"][g]))

    def emit_part(self, p, own):
        m, e = self.m, self.e
        own = set(own)
        need = set()
        for b in own:
            while b != 0:
                need.add(b)
                b = int(m.body_parentid[b])
        bodies = sorted(need)  # id order = topological order
        e("  // part %d: owns %s%s" % (p, ", ".join("%d '%s'" % (b, m.body_names[b]) for b in sorted(own)),
                                       "" if need == own else "; recomputes " + ", ".join("'%s'" % m.body_names[b] for b in sorted(need - own))))
        e("  template <class In, class Acc>")
        e("  DM_HD static void part%d(const In& in, Acc& acc, Keep& keep) {" % p)
        R = {}  # body -> 3x3 variable names
        P = {}  # body -> 3 variable names
        for i in bodies:
            par = int(m.body_parentid[i])
            mine = i in own
            if i == 1:
                e("    // body 1 '%s' (free joint): mj_normalizeQuat, xmat" % m.body_names[1])
                e("    const float px_1 = in.qpos(0), py_1 = in.qpos(1), pz_1 = in.qpos(2);")
                e("    float qw_1 = in.qpos(3), qx_1 = in.qpos(4), qy_1 = in.qpos(5), qz_1 = in.qpos(6);")
                e("    dm_normalize4(qw_1, qx_1, qy_1, qz_1);")
                e("    float r00_1, r01_1, r02_1, r10_1, r11_1, r12_1, r20_1, r21_1, r22_1;")
                e("    dm_quat2mat(qw_1, qx_1, qy_1, qz_1, r00_1, r01_1, r02_1, r10_1, r11_1, r12_1, r20_1, r21_1, r22_1);")
                R[1] = [["r%d%d_1" % (a, k) for k in range(3)] for a in range(3)]
                P[1] = ["px_1", "py_1", "pz_1"]
            else:
                e("    // body %d '%s' (parent %d)%s" % (i, m.body_names[i], par, "" if mine else " -- recomputed, owned by another part"))
                Rp, Pp = R[par], P[par]
                for a, ax in enumerate("xyz"):
                    terms = [(Rp[a][k], m.body_pos[i][k]) for k in range(3) if m.body_pos[i][k] != 0.0]
                    if terms:
                        e("    const float p%s_%d = %s;" % (ax, i, lincomb(terms, Pp[a])))
                        Pp_new = "p%s_%d" % (ax, i)
                    else:
                        Pp_new = Pp[a]
                    P.setdefault(i, []).append(Pp_new)
                bq = m.body_quat[i]
                cur = [row[:] for row in Rp]
                if not np.array_equal(bq, [1, 0, 0, 0]):
                    C = quat_to_mat(bq / np.linalg.norm(bq))
                    C[np.abs(C) < 1e-15] = 0.0
                    new = [[None] * 3 for _ in range(3)]
                    for k in range(3):  # column k of R_parent * C
                        if C[k, k] == 1.0 and not np.any(np.delete(C[:, k], k)):
                            for a in range(3):
                                new[a][k] = cur[a][k]
                            continue
                        for a in range(3):
                            terms = [(cur[a][l], C[l, k]) for l in range(3) if C[l, k] != 0.0]
                            nm = "a%d%d_%d" % (a, k, i)
                            e("    const float %s = %s;" % (nm, lincomb(terms)))
                            new[a][k] = nm
                    cur = new
                on_chain = i in self.chain and p == 0
                if on_chain:
                    e("    float wx_%d = 0.0f, wy_%d = 0.0f, wz_%d = 0.0f;" % (i, i, i))
                for hn, j in enumerate(range(int(m.body_jntadr[i]), int(m.body_jntadr[i] + m.body_jntnum[i]))):
                    assert m.jnt_type[j] == 3
                    if np.abs(m.jnt_pos[j]).max() != 0:
                        raise NotImplementedError("off-centre hinge (jnt_pos != 0) is not emitted")
                    ax = m.jnt_axis[j]
                    k = int(np.argmax(np.abs(ax)))
                    if abs(abs(ax[k]) - 1.0) > 1e-12:
                        raise NotImplementedError("hinge axis must be a coordinate axis")
                    sign = 1.0 if ax[k] > 0 else -1.0
                    qa, da = int(m.jnt_qposadr[j]), int(m.jnt_dofadr[j])
                    J = qa - 7
                    K = self.sel_q.index(qa) if qa in self.sel_q else -1
                    e("    // hinge %d '%s' axis %s%s" % (j, m.jnt_names[j], "+" if sign > 0 else "-", "xyz"[k]))
                    e("    const float th_%d = in.qpos(%d);" % (j, qa))
                    if mine:
                        e("    acc.template joint<%d, %d>(th_%d, in.qvel(%d));" % (J, K, j, da))
                    if on_chain:  # world axis = column k of the frame the hinge acts in (mj_comPos: cdof = [xaxis; xaxis x offset])
                        dv = "dv_%d" % j
                        e("    const float %s = %sin.qvel(%d);" % (dv, "" if sign > 0 else "-", da))
                        e("    wx_%d = fmaf(%s, %s, wx_%d); wy_%d = fmaf(%s, %s, wy_%d); wz_%d = fmaf(%s, %s, wz_%d);"
                          % (i, cur[0][k], dv, i, i, cur[1][k], dv, i, i, cur[2][k], dv, i))
                    e("    float s_%d, c_%d; dm_sincos_f(th_%d, s_%d, c_%d);" % (j, j, j, j, j))
                    s = ("s_%d" % j) if sign > 0 else ("(-s_%d)" % j)
                    c = "c_%d" % j
                    u, v = [(1, 2), (2, 0), (0, 1)][k]  # the two columns the rotation mixes: u' = c u + s v, v' = c v - s u
                    new = [row[:] for row in cur]
                    for a in range(3):
                        nu, nv = "h%d%d_%d" % (a, u, j), "h%d%d_%d" % (a, v, j)
                        e("    const float %s = fmaf(%s, %s, %s * %s), %s = fmaf(%s, %s, -(%s * %s));"
                          % (nu, c, cur[a][u], s, cur[a][v], nv, c, cur[a][v], s, cur[a][u]))
                        new[a][u], new[a][v] = nu, nv
                    cur = new
                R[i] = cur
            Ri, Pi = R[i], P[i]
            if i in self.chain and p == 0:
                ci = self.chain.index(i)
                e("    keep.p[%d][0] = %s; keep.p[%d][1] = %s; keep.p[%d][2] = %s;" % (ci, Pi[0], ci, Pi[1], ci, Pi[2]))
                if i != 1:
                    e("    keep.w[%d][0] = wx_%d; keep.w[%d][1] = wy_%d; keep.w[%d][2] = wz_%d;" % (ci, i, ci, i, ci, i))
                if i == 1:  # free-joint rotational dofs: world angular velocity = xmat * qvel[3:6] (mj_comPos: axis = column k of xmat)
                    for a in range(3):
                        e("    keep.w[0][%d] = fmaf(%s, in.qvel(5), fmaf(%s, in.qvel(4), %s * in.qvel(3)));" % (a, Ri[a][2], Ri[a][1], Ri[a][0]))
                if i == self.torso and i != 1:  # roll / pitch / yaw of the torso body are read off these entries
                    e("    keep.t00 = %s; keep.t10 = %s; keep.t20 = %s; keep.t21 = %s; keep.t22 = %s;"
                      % (Ri[0][0], Ri[1][0], Ri[2][0], Ri[2][1], Ri[2][2]))
            if not mine:
                continue
            mass = float(m.body_mass[i])
            e("    acc.s1x = fmaf(%s, %s, acc.s1x); acc.s1y = fmaf(%s, %s, acc.s1y); acc.s1z = fmaf(%s, %s, acc.s1z);"
              % (Pi[0], flit(mass), Pi[1], flit(mass), Pi[2], flit(mass)))
            mi = m.body_ipos[i] * mass
            for a, ax in enumerate("xyz"):
                terms = [(Ri[a][k], mi[k]) for k in range(3) if mi[k] != 0.0]
                if terms:
                    e("    acc.s2%s = %s;" % (ax, lincomb(terms, "acc.s2%s" % ax)))
            X = m.extra
            for g in [g for g in range(m.ngeom) if m.geom_bodyid[g] == i]:
                if g not in self.contact_slot and g not in self.ee_geoms:
                    continue
                gp = m.geom_pos[g]
                ex = []
                for a in range(3):
                    terms = [(Ri[a][k], gp[k]) for k in range(3) if gp[k] != 0.0]
                    ex.append(lincomb(terms, Pi[a]) if terms else Pi[a])
                if g in self.contact_slot:
                    bit, margin = self.contact_slot[g]
                    size = X["geom_size"][g]
                    if X["geom_type"][g] == "sphere":
                        dist = "(%s) - %s" % (ex[2], flit(size[0]))
                    else:  # box: the lowest of its 8 corners
                        dist = "(%s) - (fabsf(%s) * %s + fabsf(%s) * %s + fabsf(%s) * %s)" % (
                            ex[2], Ri[2][0], flit(size[0]), Ri[2][1], flit(size[1]), Ri[2][2], flit(size[2]))
                    e("    if (acc.want_contact() && %s <= %s) acc.contact |= %du;  // floor & geom %d '%s'"
                      % (dist, flit(margin), 1 << bit, g, m.geom_names[g]))
                if g in self.ee_geoms:
                    e("    acc.template ee<%d>(%s, %s, %s);  // geom %d '%s'" % (self.ee_geoms.index(g), ex[0], ex[1], ex[2], g, m.geom_names[g]))
        e("  }")

    def emit_cvel(self):
        m, e = self.m, self.e
        e("  // mj_comVel of the torso body %d '%s' = [angular; linear velocity of its point at the robot COM], from what part 0 kept"
          % (self.torso, m.body_names[self.torso]))
        e("  template <class In>")
        e("  // `in.qvel(0..2)`: the free joint's linear velocity")
        e("  DM_HD static void torso_cvel(const In& in, const Keep& k, const float* com, float* cvel) {")
        e("    const float fwx = k.w[0][0], fwy = k.w[0][1], fwz = k.w[0][2];  // free joint: xmat * qvel[3:6]")
        e("    float ox = com[0] - k.p[0][0], oy = com[1] - k.p[0][1], oz = com[2] - k.p[0][2];")
        e("    float avx = fwx, avy = fwy, avz = fwz;")
        e("    float lvx = in.qvel(0) + (fwy * oz - fwz * oy);")
        e("    float lvy = in.qvel(1) + (fwz * ox - fwx * oz);")
        e("    float lvz = in.qvel(2) + (fwx * oy - fwy * ox);")
        for ci in range(1, len(self.chain)):
            e("    ox = com[0] - k.p[%d][0]; oy = com[1] - k.p[%d][1]; oz = com[2] - k.p[%d][2];" % (ci, ci, ci))
            e("    avx += k.w[%d][0]; avy += k.w[%d][1]; avz += k.w[%d][2];" % (ci, ci, ci))
            e("    lvx += k.w[%d][1] * oz - k.w[%d][2] * oy; lvy += k.w[%d][2] * ox - k.w[%d][0] * oz; lvz += k.w[%d][0] * oy - k.w[%d][1] * ox;"
              % (ci, ci, ci, ci, ci, ci))
        e("    cvel[0] = avx; cvel[1] = avy; cvel[2] = avz; cvel[3] = lvx; cvel[4] = lvy; cvel[5] = lvz;")
        e("  }")


def emit_parts(name, m, cfg):
    e = Emit()
    pe = PartEmitter(name, m, cfg, e)
    parts = [[m.body_names.index(b) for b in part] for part in PARTS[name]]
    owned = sorted(b for part in parts for b in part)
    assert owned == list(range(1, m.nbody)), "every body is owned by exactly one part"
    assert set(pe.chain) <= set(parts[0]), "part 0 owns the chain to the torso body"
    # hinge-block joints J = qposadr - 7 and reward-selection entries K owned by each part: contiguous ranges
    jr, kr = [], []
    for part in parts:
        js = sorted(int(m.jnt_qposadr[j]) - 7 for b in part for j in range(int(m.body_jntadr[b]), int(m.body_jntadr[b] + m.body_jntnum[b]))
                    if m.jnt_type[j] == 3)
        assert js == list(range(js[0], js[-1] + 1)), "a part's joints must be contiguous in qpos"
        ks = [pe.sel_q.index(j + 7) for j in js if j + 7 in pe.sel_q]
        assert ks == list(range(ks[0], ks[-1] + 1)), "a part's reward joints must be contiguous in the selection"
        jr.append((js[0], js[-1] + 1))
        kr.append((ks[0], ks[-1] + 1))
    e("// GENERATED by tools/gen_parts_code.py from assets/models.json -- do not edit.")
    e("// Robot '%s' cut into %d parts for the team kernel (one warp per part, lane = environment)." % (name, len(parts)))
    e("#pragma once")
    e('#include "robot_%s.cuh"' % name)
    e("namespace dm {")
    e("template <class Rb> struct PartsOf;")
    e("struct Parts_%s {" % name)
    e("  using Rb = Robot_%s;" % name)
    e("  static constexpr int NP = %d, NCHAIN = %d;  // parts; bodies on the chain root .. torso" % (len(parts), len(pe.chain)))

    def tab(fname, vals):
        e("  DM_HD static constexpr int %s(int p) { return %s; }" % (
            fname, " : ".join(["p == %d ? %d" % (i, v) for i, v in enumerate(vals[:-1])] + [str(vals[-1])])))

    tab("j0", [a for a, _ in jr])
    tab("j1", [b for _, b in jr])
    tab("k0", [a for a, _ in kr])
    tab("k1", [b for _, b in kr])
    e("  static constexpr int RPY_PART = %d;  // the part that also turns the free-joint quaternion into roll / pitch / yaw" % RPY_PART[name])
    e("  DM_HD static constexpr float inv_mass() { return %s; }" % flit(1.0 / float(m.body_mass.sum())))
    e("  struct Keep {  // what the combine step needs from part 0's forward kinematics")
    e("    float p[NCHAIN][3], w[NCHAIN][3];  // chain body positions; angular velocity each adds (w[0]: the free joint's)")
    e("    float t00, t10, t20, t21, t22;  // torso xmat entries (torso != root only)")
    e("  };")
    for p, own in enumerate(parts):
        pe.emit_part(p, own)
    pe.emit_cvel()
    e("  template <int P, class In, class Acc>")
    e("  DM_HD static void part(const In& in, Acc& acc, Keep& keep) {")
    for p in range(len(parts)):
        e("    if constexpr (P == %d) part%d(in, acc, keep);" % (p, p))
    e("  }")
    e("};")
    e("template <> struct PartsOf<Robot_%s> { using type = Parts_%s; };" % (name, name))
    e("}  // namespace dm")
    return "\n".join(e.lines) + "\n"


def main():
    models = load_models(os.path.join(REPO, "deepmimic_mujoco_b200", "assets", "models.json"))
    out = os.path.join(REPO, "deepmimic_mujoco_b200", "csrc", "generated")
    for name, m in models.items():
        src = emit_parts(name, m, ROBOT_CFG[name])
        with open(os.path.join(out, "parts_%s.cuh" % name), "w") as f:
            f.write(src)
        print(name, len(src.splitlines()), "lines")


if __name__ == "__main__":
    main()
