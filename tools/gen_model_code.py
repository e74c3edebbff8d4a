#!/usr/bin/env python3
"""Model compiler back end: emit straight-line CUDA for each robot's kinematic tree.

Input : deepmimic_mujoco_b200/assets/models.json (compiled by mjcf.py from the reference's MJCF).
Output: deepmimic_mujoco_b200/csrc/generated/robot_<name>.cuh

For every robot the header holds a traits struct (sizes, observation layout, reward joint
selection, exact fp32 joint-limit thresholds, motor gear / ctrlrange) and ONE fully unrolled
forward-kinematics routine `fk<T>(...)` restating MuJoCo's mj_kinematics / mj_comPos /
mj_comVel for exactly this tree (reference call sites src/deepmimic_env.py:29,51-54,232,236,
357,421; algorithm: see oracle/mjkin.py).  Unrolling at code-generation time is what lets the
kernel skip every zero of body_pos / body_ipos / joint axes and every identity body_quat --
a table-driven loop could not.  `T` is float in the step kernel and double in the clip-table
builder, so both share a single source of truth.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, REPO)
from deepmimic_mujoco_b200.mjcf import load_models  # noqa: E402

ROBOT_CFG = {  # reference src/config.py:3-22
    "humanoid3d": dict(torso="chest", ee=["left_ankle", "right_ankle", "left_wrist", "right_wrist"], low_z=0.7,
                       act_scale=1.0, n_hand=0, floor="floor", rfoot="right_ankle", lfoot="left_ankle", extra=[]),
    "unitree_g1": dict(torso="pelvis", ee=["left_foot", "right_foot", "left_hand", "right_hand"], low_z=0.4,
                       act_scale=20.0, n_hand=14, floor="floor", rfoot="right_foot", lfoot="left_foot",
                       extra=["left_foot_lheel", "left_foot_rheel", "left_foot_ltoe", "left_foot_rtoe",
                              "right_foot_lheel", "right_foot_rheel", "right_foot_ltoe", "right_foot_rtoe"]),
}


def lit(x):
    """C literal wrapped for the scalar type T."""
    return "T(%s)" % repr(float(x))


def f32_floor(x):
    f = np.float32(x)
    if float(f) > x:
        f = np.nextafter(f, np.float32(-np.inf))
    return f


def f32_ceil(x):
    f = np.float32(x)
    if float(f) < x:
        f = np.nextafter(f, np.float32(np.inf))
    return f


def c_float(f):
    s = repr(float(np.float32(f)))
    if "inf" in s or "nan" in s:
        raise ValueError(s)
    return s + "f"


class Emit:
    def __init__(self):
        self.lines = []

    def __call__(self, s=""):
        self.lines.append(s)


def mat_rows_times_const(R, vec, eps=0.0):
    """expressions for R @ vec, skipping zero constants; R is a 3x3 of variable names."""
    out = []
    for a in range(3):
        terms = ["%s * %s" % (R[a][k], lit(vec[k])) for k in range(3) if abs(vec[k]) > eps]
        out.append(terms)
    return out


def emit_fk(m, cfg, e):
    nb = m.nbody
    torso = m.body_names.index(cfg["torso"])
    ee_geoms = [m.geom_names.index(n) for n in cfg["ee"]]
    M = float(m.body_mass.sum())
    # ancestors of the torso (for cvel)
    chain = []
    b = torso
    while b != 0:
        chain.append(b)
        b = int(m.body_parentid[b])
    chain = chain[::-1]
    assert chain[0] == 1 and m.jnt_type[m.body_jntadr[1]] == 0, "body 1 must carry the free joint"
    geoms_of = {i: [g for g in range(m.ngeom) if m.geom_bodyid[g] == i] for i in range(nb)}
    # Floor contacts of the observation (deepmimic_env.py:78-121) from FK: MuJoCo's plane-sphere / plane-box narrow
    # phase against the floor plane (z = 0, normal +z), contact iff dist <= margin with margin = max of the two geom
    # margins (MuJoCo 2.1 mj_collideGeoms), and only for pairs that pass the contype / conaffinity filter.
    # bit 0 rfoot, bit 1 lfoot, bits 2.. the extra-contact geoms.
    X = m.extra
    fl = m.geom_names.index(cfg["floor"])
    assert X["geom_type"][fl] == "plane" and m.geom_bodyid[fl] == 0 and not np.any(m.geom_pos[fl]) and not X["geom_has_quat"][fl]
    contact_slot = {}
    for bit, name in enumerate([cfg["rfoot"], cfg["lfoot"]] + list(cfg["extra"])):
        g = m.geom_names.index(name)
        collide = (m.geom_contype[fl] & m.geom_conaffinity[g]) or (m.geom_contype[g] & m.geom_conaffinity[fl])
        if collide:
            assert not X["geom_has_quat"][g] and X["geom_type"][g] in ("sphere", "box"), name
            contact_slot[g] = (bit, max(X["geom_margin"][fl], X["geom_margin"][g]))

    e("  // Forward kinematics + COM + torso cvel, unrolled for this tree (see file header).")
    e("  template <typename T, typename In, typename Sink>")
    e("  DM_HD static void fk(const In& in, FkOut<T>& o, Sink& sink) {")
    # root body (free joint): mj_normalizeQuat + copy
    e("    // body 1 '%s' (free joint)" % m.body_names[1])
    e("    T px_1 = in.qpos(0), py_1 = in.qpos(1), pz_1 = in.qpos(2);")
    e("    T qw_1 = in.qpos(3), qx_1 = in.qpos(4), qy_1 = in.qpos(5), qz_1 = in.qpos(6);")
    e("    dm_normalize4(qw_1, qx_1, qy_1, qz_1);")
    e("    o.rq[0] = qw_1; o.rq[1] = qx_1; o.rq[2] = qy_1; o.rq[3] = qz_1;")

    def emit_mat(i):
        e("    T r00_%d, r01_%d, r02_%d, r10_%d, r11_%d, r12_%d, r20_%d, r21_%d, r22_%d;" % ((i,) * 9))
        e("    dm_quat2mat(qw_%d, qx_%d, qy_%d, qz_%d, r00_%d, r01_%d, r02_%d, r10_%d, r11_%d, r12_%d, r20_%d, r21_%d, r22_%d);"
          % ((i,) * 13))

    def Rn(i):
        return [["r%d%d_%d" % (a, k, i) for k in range(3)] for a in range(3)]

    def emit_body_tail(i):
        """xmat, COM accumulation, geoms, sink."""
        emit_mat(i)
        mass = float(m.body_mass[i])
        e("    s1x = dm_fma(px_%d, %s, s1x); s1y = dm_fma(py_%d, %s, s1y); s1z = dm_fma(pz_%d, %s, s1z);"
          % (i, lit(mass), i, lit(mass), i, lit(mass)))
        mi = m.body_ipos[i] * mass
        rows = mat_rows_times_const(Rn(i), mi)
        for a, ax in enumerate("xyz"):
            for t in rows[a]:
                e("    s2%s += %s;" % (ax, t))
        e("    sink.body(%d, px_%d, py_%d, pz_%d);" % (i, i, i, i))
        for g in geoms_of[i]:
            gp = m.geom_pos[g]
            rows = mat_rows_times_const(Rn(i), gp)
            ex = []
            for a, ax in enumerate("xyz"):
                ex.append(" + ".join(["p%s_%d" % (ax, i)] + rows[a]))
            if g in contact_slot:
                bit, margin = contact_slot[g]
                size = X["geom_size"][g]
                R = Rn(i)
                if X["geom_type"][g] == "sphere":
                    dist = "(%s) - %s" % (ex[2], lit(size[0]))
                else:  # box: the lowest of the 8 corners
                    dist = "(%s) - (dm_abs(%s) * %s + dm_abs(%s) * %s + dm_abs(%s) * %s)" % (
                        ex[2], R[2][0], lit(size[0]), R[2][1], lit(size[1]), R[2][2], lit(size[2]))
                e("    if (sink.contact_on() && %s <= %s) o.contact |= %du;  // floor & geom %d '%s'" % (dist, lit(margin), 1 << bit, g, m.geom_names[g]))
            if g in ee_geoms:
                k = ee_geoms.index(g)
                e("    o.ee[%d] = %s; o.ee[%d] = %s; o.ee[%d] = %s;  // geom %d '%s'"
                  % (3 * k, ex[0], 3 * k + 1, ex[1], 3 * k + 2, ex[2], g, m.geom_names[g]))
                e("    sink.geom(%d, o.ee[%d], o.ee[%d], o.ee[%d]);" % (g, 3 * k, 3 * k + 1, 3 * k + 2))
            else:
                e("    if (Sink::kWantsGeoms) sink.geom(%d, %s, %s, %s);" % (g, ex[0], ex[1], ex[2]))
        if i == torso:
            e("    o.tq[0] = qw_%d; o.tq[1] = qx_%d; o.tq[2] = qy_%d; o.tq[3] = qz_%d;" % ((i,) * 4))

    e("    o.contact = 0u;")
    e("    T s1x = T(0), s1y = T(0), s1z = T(0);  // sum m * xpos")
    e("    T s2x = T(0), s2y = T(0), s2z = T(0);  // sum R * (m * ipos)")
    emit_body_tail(1)
    chain_w = {}  # body -> names of (wx, wy, wz) angular contribution of its hinges
    hinge_order = [j for i in range(2, nb) for j in range(int(m.body_jntadr[i]), int(m.body_jntadr[i] + m.body_jntnum[i]))]
    sincos_done = set()
    for i in range(2, nb):
        p = int(m.body_parentid[i])
        e("    // body %d '%s' (parent %d)" % (i, m.body_names[i], p))
        rows = mat_rows_times_const(Rn(p), m.body_pos[i])
        for a, ax in enumerate("xyz"):
            e("    T p%s_%d = %s;" % (ax, i, " + ".join(["p%s_%d" % (ax, p)] + rows[a])))
        bq = m.body_quat[i]
        if np.allclose(bq, [1, 0, 0, 0], atol=0, rtol=0):
            e("    T qw_%d = qw_%d, qx_%d = qx_%d, qy_%d = qy_%d, qz_%d = qz_%d;" % (i, p, i, p, i, p, i, p))
        else:
            # xquat = xquat_parent * body_quat with the zero components of body_quat skipped
            a = ["qw_%d" % p, "qx_%d" % p, "qy_%d" % p, "qz_%d" % p]
            prod = {  # Hamilton product rows: (a index, b index, sign)
                "w": [(0, 0, 1), (1, 1, -1), (2, 2, -1), (3, 3, -1)],
                "x": [(0, 1, 1), (1, 0, 1), (2, 3, 1), (3, 2, -1)],
                "y": [(0, 2, 1), (1, 3, -1), (2, 0, 1), (3, 1, 1)],
                "z": [(0, 3, 1), (1, 2, 1), (2, 1, -1), (3, 0, 1)],
            }
            for comp, terms in prod.items():
                ex = ""
                for ai, bi, sg in terms:
                    if bq[bi] == 0.0:
                        continue
                    ex += " %s %s * %s" % ("+" if sg * bq[bi] > 0 else "-", a[ai], lit(abs(bq[bi])))
                ex = ex.strip()
                if ex.startswith("+ "):
                    ex = ex[2:]
                e("    T q%s_%d = %s;" % (comp, i, ex or "T(0)"))
        on_chain = i in chain
        if on_chain:
            e("    T wx_%d = T(0), wy_%d = T(0), wz_%d = T(0);" % (i, i, i))
            chain_w[i] = True
        for j in range(int(m.body_jntadr[i]), int(m.body_jntadr[i] + m.body_jntnum[i])):
            assert m.jnt_type[j] == 3
            if np.abs(m.jnt_pos[j]).max() != 0:
                raise NotImplementedError("off-centre hinge (jnt_pos != 0) is not emitted")
            ax = m.jnt_axis[j]
            k = int(np.argmax(np.abs(ax)))
            if abs(abs(ax[k]) - 1.0) > 1e-12:
                raise NotImplementedError("hinge axis must be a coordinate axis")
            sign = 1.0 if ax[k] > 0 else -1.0
            qa = int(m.jnt_qposadr[j])
            # half-angle sine / cosine: consecutive hinges (emission order) in pairs, one packed evaluation for both
            # (dm_sincos_half2, dm_math.cuh); the second hinge of a pair finds its values already there
            if j not in sincos_done:
                nxt = hinge_order[hinge_order.index(j) + 1] if hinge_order.index(j) + 1 < len(hinge_order) else None
                if nxt is not None and nxt not in sincos_done:
                    e("    T sh_%d, ch_%d, sh_%d, ch_%d; dm_sincos_half2(in.qpos(%d), in.qpos(%d), sh_%d, ch_%d, sh_%d, ch_%d);"
                      % (j, j, nxt, nxt, qa, int(m.jnt_qposadr[nxt]), j, j, nxt, nxt))
                    sincos_done.update((j, nxt))
                else:
                    e("    T sh_%d, ch_%d; dm_sincos_half(in.qpos(%d), sh_%d, ch_%d);" % (j, j, qa, j, j))
                    sincos_done.add(j)
            e("    {  // hinge %d '%s' axis %s%s" % (j, m.jnt_names[j], "+" if sign > 0 else "-", "xyz"[k]))
            if on_chain:
                da = int(m.jnt_dofadr[j])
                e("      T ax_, ay_, az_; dm_quat_col%d(qw_%d, qx_%d, qy_%d, qz_%d, ax_, ay_, az_);" % (k, i, i, i, i))
                e("      T dv_ = %sin.qvel(%d);" % ("" if sign > 0 else "-", da))
                e("      wx_%d = dm_fma(ax_, dv_, wx_%d); wy_%d = dm_fma(ay_, dv_, wy_%d); wz_%d = dm_fma(az_, dv_, wz_%d);"
                  % (i, i, i, i, i, i))
            e("      T s_ = %ssh_%d, c_ = ch_%d;" % ("-" if sign < 0 else "", j, j))
            e("      dm_quat_hinge_%s(qw_%d, qx_%d, qy_%d, qz_%d, s_, c_);" % ("xyz"[k], i, i, i, i))
            e("    }")
        emit_body_tail(i)
    # COM
    e("    const T inv_m = T(%s);" % repr(1.0 / M))
    e("    o.com_xpos[0] = s1x * inv_m; o.com_xpos[1] = s1y * inv_m; o.com_xpos[2] = s1z * inv_m;")
    e("    o.com[0] = (s1x + s2x) * inv_m; o.com[1] = (s1y + s2y) * inv_m; o.com[2] = (s1z + s2z) * inv_m;")
    # cvel of the torso: free joint + chain hinges
    e("    // mj_comVel for body %d '%s': [angular; linear at the robot COM]" % (torso, m.body_names[torso]))
    e("    T fwx = r00_1 * in.qvel(3) + r01_1 * in.qvel(4) + r02_1 * in.qvel(5);")
    e("    T fwy = r10_1 * in.qvel(3) + r11_1 * in.qvel(4) + r12_1 * in.qvel(5);")
    e("    T fwz = r20_1 * in.qvel(3) + r21_1 * in.qvel(4) + r22_1 * in.qvel(5);")
    e("    T ox = o.com[0] - px_1, oy = o.com[1] - py_1, oz = o.com[2] - pz_1;")
    e("    T avx = fwx, avy = fwy, avz = fwz;")
    e("    T lvx = in.qvel(0) + (fwy * oz - fwz * oy);")
    e("    T lvy = in.qvel(1) + (fwz * ox - fwx * oz);")
    e("    T lvz = in.qvel(2) + (fwx * oy - fwy * ox);")
    for i in chain[1:]:
        e("    ox = o.com[0] - px_%d; oy = o.com[1] - py_%d; oz = o.com[2] - pz_%d;" % (i, i, i))
        e("    avx += wx_%d; avy += wy_%d; avz += wz_%d;" % (i, i, i))
        e("    lvx += wy_%d * oz - wz_%d * oy; lvy += wz_%d * ox - wx_%d * oz; lvz += wx_%d * oy - wy_%d * ox;"
          % (i, i, i, i, i, i))
    e("    o.cvel[0] = avx; o.cvel[1] = avy; o.cvel[2] = avz; o.cvel[3] = lvx; o.cvel[4] = lvy; o.cvel[5] = lvz;")
    e("  }")


def emit_robot(name, m, cfg):
    e = Emit()
    nsel_q = list(range(7, m.nq)) if name == "humanoid3d" else list(range(7, 25)) + list(range(32, 37))
    nsel_v = [q - 1 for q in nsel_q]
    na = m.nu - cfg["n_hand"]
    tol = m.jnt_range[1:] * 0.99  # deepmimic_env.py:242
    sel_j = [q - 7 for q in nsel_q]
    e("// GENERATED by tools/gen_model_code.py from assets/models.json -- do not edit.")
    e("// Robot '%s': nq=%d nv=%d nu=%d nbody=%d ngeom=%d, total mass %.9g" %
      (name, m.nq, m.nv, m.nu, m.nbody, m.ngeom, m.body_mass.sum()))
    e("#pragma once")
    e('#include "../dm_math.cuh"')
    e("namespace dm {")
    e("struct Robot_%s {" % name)
    e("  static constexpr int kRobotId = %d;" % (0 if name == "humanoid3d" else 1))
    e("  static constexpr int NQ = %d, NV = %d, NU = %d, NA = %d, NBODY = %d, NGEOM = %d;" %
      (m.nq, m.nv, m.nu, na, m.nbody, m.ngeom))
    e("  static constexpr int NJ = NQ - 7;           // hinge joints = obs joint block")
    e("  static constexpr int NSEL = %d;              // joints in the reward (deepmimic_env.py:206-207)" % len(nsel_q))
    e("  static constexpr bool kIsG1 = %s;" % ("true" if name == "unitree_g1" else "false"))
    e("  static constexpr bool kTorsoIsRoot = %s;   // torso body == free-joint body" % ("true" if m.body_names.index(cfg["torso"]) == 1 else "false"))
    e("  static constexpr float kActScale = %s;      // deepmimic_env.py:347-349" % c_float(cfg["act_scale"]))
    e("  DM_HD static constexpr double kLowZd() { return %r; }  // config.py:13,22 (compared in double)" % cfg["low_z"])
    # selection as ranges
    runs = []
    for q in nsel_q:
        if runs and runs[-1][1] == q:
            runs[-1][1] = q + 1
        else:
            runs.append([q, q + 1])
    expr, off = "", 0
    for a, b in runs[:-1]:
        expr += "k < %d ? %d + (k - %d) : " % (off + b - a, a, off)
        off += b - a
    expr += "%d + (k - %d)" % (runs[-1][0], off)
    e("  DM_HD static constexpr int sel_q(int k) { return %s; }  // qpos index; qvel index = sel_q - 1" % expr)

    def arr(tname, cname, vals, fmt):
        e("  DM_HD static %s %s(int k) {" % (tname, cname))
        e("    switch (k) {")
        for i, v in enumerate(vals):
            e("      case %d: return %s;" % (i, fmt(v)))
        e("      default: return %s;" % fmt(0))
        e("    }")
        e("  }")

    # exact thresholds: (double)q <= tol_lo  <=>  q <= floor_f32(tol_lo);  q >= tol_hi <=> q >= ceil_f32(tol_hi)
    arr("float", "lim_lo", [f32_floor(tol[j, 0]) for j in sel_j], c_float)
    arr("float", "lim_hi", [f32_ceil(tol[j, 1]) for j in sel_j], c_float)
    arr("double", "body_mass", m.body_mass, lambda v: repr(float(v)))
    arr("int", "ee_geom", [m.geom_names.index(n) for n in cfg["ee"]], lambda v: str(int(v)))
    arr("float", "gear", m.actuator_gear, c_float)
    arr("float", "ctrl_lo", m.actuator_ctrlrange[:, 0], c_float)
    arr("float", "ctrl_hi", m.actuator_ctrlrange[:, 1], c_float)
    emit_fk(m, cfg, e)
    e("};")
    e("}  // namespace dm")
    return "\n".join(e.lines) + "\n"


def main():
    models = load_models(os.path.join(REPO, "deepmimic_mujoco_b200", "assets", "models.json"))
    out = os.path.join(REPO, "deepmimic_mujoco_b200", "csrc", "generated")
    os.makedirs(out, exist_ok=True)
    for name, m in models.items():
        src = emit_robot(name, m, ROBOT_CFG[name])
        with open(os.path.join(out, "robot_%s.cuh" % name), "w") as f:
            f.write(src)
        print(name, len(src.splitlines()), "lines")


if __name__ == "__main__":
    main()
