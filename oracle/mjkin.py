"""MuJoCo 2.1 kinematics restated in float64 numpy (oracle; test infrastructure).

The reference obtains every pose quantity from MuJoCo through mujoco-py:
`set_state()` -> `sim.forward()` (reference src/deepmimic_env.py:357,508) or
`sim.step()` (:362), then reads `data.qpos`, `data.body_xpos`, `data.body_xquat`,
`data.geom_xpos`, `data.xipos`, `data.cvel` (:29,51-54,162,216,232,236,421) and at
load time `body_xpos` / `geom_xpos` per key-frame (src/mujoco/mocap_v2.py:304-307).
MuJoCo is not installed here (PARITY UNPINNED at the value level); the functions
below follow the published algorithm of engine_core_smooth.c:

  mj_kinematics: normalise the quaternions stored in qpos IN PLACE; for each body in
    id order: free joint -> xpos = qpos[0:3], xquat = qpos[3:7]; otherwise
    xpos = xpos_parent + xmat_parent @ body_pos, xquat = xquat_parent * body_quat and
    for each hinge j: xaxis_j = R(xquat) axis_j, xanchor_j = R(xquat) jnt_pos_j + xpos,
    xquat = xquat * axisangle(axis_j, qpos_j - qpos0_j),
    xpos = xanchor_j - R(xquat) jnt_pos_j; then normalise xquat, xmat = mat(xquat);
    xipos = xpos + xmat @ body_ipos;  geom_xpos = xpos[body] + xmat[body] @ geom_pos.
  mj_comPos: subtree_com[root] = sum(m * xipos) / sum(m); cdof per dof with
    offset = subtree_com[root] - xanchor:  free translation k -> [0; e_k],
    free rotation k -> [xmat[:,k]; xmat[:,k] x offset], hinge -> [xaxis; xaxis x offset].
  mj_comVel: cvel[b] = cvel[parent] + sum_dofs cdof * qvel  ([angular; linear]).
  motor:     actuator_force = gear * clip(ctrl, ctrlrange)   (mj_fwdActuation).
"""
import numpy as np

from .rotations import axis_angle_to_quat, quat_mul, quat_normalize, quat_to_mat

JNT_FREE = 0


class MjDataLike:
    """The mujoco-py `sim.data` fields the reference reads."""

    __slots__ = ("qpos", "qvel", "xpos", "xquat", "xmat", "xipos", "geom_xpos", "xanchor", "xaxis",
                 "subtree_com", "cvel", "contact", "ctrl")

    # mujoco-py aliases (deepmimic_env.py:51-52,236)
    @property
    def body_xpos(self):
        return self.xpos

    @property
    def body_xquat(self):
        return self.xquat


def forward(model, qpos, qvel, contact_bits=0):
    """`set_state(qpos, qvel)` + `sim.forward()`, position/velocity stages only."""
    d = MjDataLike()
    qpos = np.array(qpos, dtype=np.float64)
    qvel = np.array(qvel, dtype=np.float64)
    nb = model.nbody
    xpos = np.zeros((nb, 3))
    xquat = np.zeros((nb, 4))
    xquat[0, 0] = 1.0
    xmat = np.zeros((nb, 3, 3))
    xmat[0] = np.eye(3)
    xanchor = np.zeros((model.njnt, 3))
    xaxis = np.zeros((model.njnt, 3))
    # mj_normalizeQuat(m, d->qpos)
    for j in range(model.njnt):
        if model.jnt_type[j] == JNT_FREE:
            a = model.jnt_qposadr[j] + 3
            qpos[a:a + 4] = quat_normalize(qpos[a:a + 4])
    for i in range(1, nb):
        jadr, jnum = model.body_jntadr[i], model.body_jntnum[i]
        if jnum == 1 and model.jnt_type[jadr] == JNT_FREE:
            a = model.jnt_qposadr[jadr]
            pos = qpos[a:a + 3].copy()
            quat = quat_normalize(qpos[a + 3:a + 7])
            xanchor[jadr] = pos
            xaxis[jadr] = model.jnt_axis[jadr]
        else:
            p = model.body_parentid[i]
            pos = xpos[p] + xmat[p] @ model.body_pos[i]
            quat = quat_mul(xquat[p], model.body_quat[i])
            for j in range(jadr, jadr + jnum):
                R = quat_to_mat(quat)
                xaxis[j] = R @ model.jnt_axis[j]
                xanchor[j] = R @ model.jnt_pos[j] + pos
                quat = quat_mul(quat, axis_angle_to_quat(model.jnt_axis[j], qpos[model.jnt_qposadr[j]]))
                pos = xanchor[j] - quat_to_mat(quat) @ model.jnt_pos[j]
        quat = quat_normalize(quat)
        xpos[i], xquat[i], xmat[i] = pos, quat, quat_to_mat(quat)
    xipos = xpos + np.einsum("bij,bj->bi", xmat, model.body_ipos)
    gb = model.geom_bodyid
    geom_xpos = xpos[gb] + np.einsum("gij,gj->gi", xmat[gb], model.geom_pos)
    # mj_comPos (single kinematic tree rooted at body 1)
    m = model.body_mass
    com = (m[:, None] * xipos).sum(0) / m.sum()
    cdof = np.zeros((model.nv, 6))
    for j in range(model.njnt):
        da = model.jnt_dofadr[j]
        off = com - xanchor[j]
        if model.jnt_type[j] == JNT_FREE:
            R = xmat[model.jnt_bodyid[j]]
            for k in range(3):
                cdof[da + k, 3 + k] = 1.0
                ax = R[:, k]
                cdof[da + 3 + k, :3] = ax
                cdof[da + 3 + k, 3:] = np.cross(ax, off)
        else:
            cdof[da, :3] = xaxis[j]
            cdof[da, 3:] = np.cross(xaxis[j], off)
    # mj_comVel
    cvel = np.zeros((nb, 6))
    dof_body = np.zeros(model.nv, dtype=np.int64)
    for j in range(model.njnt):
        n = 6 if model.jnt_type[j] == JNT_FREE else 1
        dof_body[model.jnt_dofadr[j]:model.jnt_dofadr[j] + n] = model.jnt_bodyid[j]
    for i in range(1, nb):
        v = cvel[model.body_parentid[i]].copy()
        for k in np.nonzero(dof_body == i)[0]:
            v += cdof[k] * qvel[k]
        cvel[i] = v
    d.qpos, d.qvel = qpos, qvel
    d.xpos, d.xquat, d.xmat, d.xipos, d.geom_xpos = xpos, xquat, xmat, xipos, geom_xpos
    d.xanchor, d.xaxis, d.subtree_com, d.cvel = xanchor, xaxis, com, cvel
    d.contact = contact_bits
    d.ctrl = np.zeros(model.nu)
    return d


def motor_forces(model, ctrl):
    """actuator_force of <motor>s: gear * clip(ctrl, ctrlrange) (clip only when ctrllimited)."""
    ctrl = np.asarray(ctrl, dtype=np.float64)
    lo, hi = model.actuator_ctrlrange[:, 0], model.actuator_ctrlrange[:, 1]
    clipped = np.where(model.actuator_ctrllimited, np.clip(ctrl, lo, hi), ctrl)
    return model.actuator_gear * clipped


def floor_contact_bits(model, d, rc):
    """Floor contacts of the observation from kinematics alone: the contact word (bit 0 rfoot & floor, bit 1 lfoot &
    floor, bits 2.. extra-contact geoms & floor) that get_foot_contact_obs / get_extra_contact_obs
    (src/deepmimic_env.py:78-121) would read off mjdata.contact.  Restates MuJoCo 2.1's narrow phase for the only
    pairs involved -- mjc_PlaneSphere: dist = (c - p0) . n - r;  mjc_PlaneBox: the box corners, i.e. the lowest one
    decides -- with a contact reported iff dist <= margin, margin = max(geom_margin) of the pair (mj_collideGeoms),
    after the contype / conaffinity filter.  PARITY UNPINNED: MuJoCo is not installable here."""
    X = model.extra
    fl = model.geom_name2id(rc["floor"])
    assert X["geom_type"][fl] == "plane"
    names = [rc["rfoot"], rc["lfoot"]] + list(rc["extra"] or [])
    word = 0
    for bit, name in enumerate(names):
        g = model.geom_name2id(name)
        if not ((model.geom_contype[fl] & model.geom_conaffinity[g]) or (model.geom_contype[g] & model.geom_conaffinity[fl])):
            continue
        b = model.geom_bodyid[g]
        size = X["geom_size"][g]
        if X["geom_type"][g] == "sphere":
            dist = d.geom_xpos[g][2] - size[0]
        elif X["geom_type"][g] == "box":
            dist = d.geom_xpos[g][2] - float(np.abs(d.xmat[b][2]) @ np.asarray(size))
        else:
            raise NotImplementedError(X["geom_type"][g])
        if dist <= max(X["geom_margin"][fl], X["geom_margin"][g]):
            word |= 1 << bit
    return word


def floor_contact_dists(model, d, rc):
    """The signed distances behind floor_contact_bits (tests use them to skip states sitting on the threshold)."""
    X = model.extra
    out = []
    for name in [rc["rfoot"], rc["lfoot"]] + list(rc["extra"] or []):
        g = model.geom_name2id(name)
        b = model.geom_bodyid[g]
        size = X["geom_size"][g]
        z = d.geom_xpos[g][2]
        out.append(z - size[0] if X["geom_type"][g] == "sphere" else z - float(np.abs(d.xmat[b][2]) @ np.asarray(size)))
    return np.asarray(out)
