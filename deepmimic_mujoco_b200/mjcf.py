"""MJCF mini-compiler: the subset of MuJoCo's model compiler the hot path needs.

The reference never parses MJCF itself -- it hands the XML to MuJoCo
(`mujoco_env.MujocoEnv.__init__(self, xml_file_path, 6)`, reference
src/deepmimic_env.py:301, src/combined_env.py:191) and then reads
`model.body_mass`, `model.jnt_range`, `model.geom_name2id`, ... from the
compiled model (src/deepmimic_env.py:28,50,196,231,242).  This module produces
those compiled arrays (kinematic tree, inertial frames, joints, named geoms,
motors) from the two MJCF files the reference ships
(src/mujoco/humanoid_deepmimic/envs/asset/deepmimic_humanoid3d.xml,
.../deepmimic_unitree_g1.xml) without meshes, dynamics or collision data.

Conventions follow MuJoCo's compiler: body / joint / geom ids are assigned in
depth-first document order with the world as body 0, geoms are grouped by body,
quaternions are (w, x, y, z), `inertiafromgeom` masses are the sum of the geom
masses and `body_ipos` the mass-weighted mean of the geom centres.
"""
from __future__ import annotations

import json
import math
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field

import numpy as np

JNT_FREE = 0
JNT_HINGE = 3  # mjJNT_HINGE


@dataclass
class RobotModel:
    """Compiled kinematic model (float64 numpy, MuJoCo field names)."""

    name: str
    nq: int
    nv: int
    nu: int
    nbody: int
    njnt: int
    ngeom: int
    body_names: list
    body_parentid: np.ndarray  # [nbody] int
    body_pos: np.ndarray  # [nbody,3]
    body_quat: np.ndarray  # [nbody,4] wxyz, normalised
    body_ipos: np.ndarray  # [nbody,3]
    body_mass: np.ndarray  # [nbody]
    body_jntadr: np.ndarray  # [nbody] first joint id or -1
    body_jntnum: np.ndarray  # [nbody]
    jnt_names: list
    jnt_type: np.ndarray  # [njnt]
    jnt_bodyid: np.ndarray
    jnt_qposadr: np.ndarray
    jnt_dofadr: np.ndarray
    jnt_axis: np.ndarray  # [njnt,3]
    jnt_pos: np.ndarray  # [njnt,3]
    jnt_range: np.ndarray  # [njnt,2]
    jnt_limited: np.ndarray  # [njnt] bool
    geom_names: list  # '' for unnamed
    geom_bodyid: np.ndarray
    geom_pos: np.ndarray  # [ngeom,3] in the body frame
    geom_contype: np.ndarray
    geom_conaffinity: np.ndarray
    actuator_names: list
    actuator_jntid: np.ndarray  # [nu]
    actuator_dofadr: np.ndarray  # [nu]
    actuator_gear: np.ndarray  # [nu]
    actuator_ctrlrange: np.ndarray  # [nu,2]
    actuator_ctrllimited: np.ndarray  # [nu] bool
    timestep: float = 0.0
    extra: dict = field(default_factory=dict)

    # MuJoCo-style name lookups used by the reference (deepmimic_env.py:50,231)
    def body_name2id(self, name):
        return self.body_names.index(name)

    def geom_name2id(self, name):
        return self.geom_names.index(name)

    def geom_id2name(self, i):
        n = self.geom_names[i]
        return n if n else None

    @property
    def joint_names(self):
        return tuple(self.jnt_names)

    def to_dict(self):
        out = {}
        for k, v in self.__dict__.items():
            out[k] = v.tolist() if isinstance(v, np.ndarray) else v
        return out

    @staticmethod
    def from_dict(d):
        kw = {}
        int_fields = {"body_parentid", "body_jntadr", "body_jntnum", "jnt_type", "jnt_bodyid",
                      "jnt_qposadr", "jnt_dofadr", "geom_bodyid", "geom_contype",
                      "geom_conaffinity", "actuator_jntid", "actuator_dofadr"}
        bool_fields = {"jnt_limited", "actuator_ctrllimited"}
        for k, v in d.items():
            if isinstance(v, list) and k not in ("body_names", "jnt_names", "geom_names", "actuator_names"):
                if k in int_fields:
                    v = np.asarray(v, dtype=np.int64)
                elif k in bool_fields:
                    v = np.asarray(v, dtype=bool)
                else:
                    v = np.asarray(v, dtype=np.float64)
            kw[k] = v
        return RobotModel(**kw)


def _floats(s, n=None):
    v = [float(x) for x in s.split()]
    if n is not None and len(v) != n:
        raise ValueError("expected %d numbers, got %r" % (n, s))
    return v


def _bool(s):
    return str(s).strip().lower() in ("true", "1")


class _Defaults:
    """MJCF default classes: nested <default class=...> blocks, inherited top-down."""

    def __init__(self, root):
        self.classes = {}  # name -> {tag: attrs}
        top = {}
        for d in root.findall("default"):
            self._walk(d, top, "main")

    def _walk(self, node, inherited, name):
        cur = {t: dict(a) for t, a in inherited.items()}
        for child in node:
            if child.tag == "default":
                continue
            cur.setdefault(child.tag, {}).update(child.attrib)
        cname = node.get("class", name)
        self.classes[cname] = cur
        for child in node.findall("default"):
            self._walk(child, cur, child.get("class"))

    def resolve(self, elem, tag, childclass):
        cls = elem.get("class", childclass or "main")
        base = dict(self.classes.get(cls, self.classes.get("main", {})).get(tag, {}))
        base.update(elem.attrib)
        return base


_GEOM_DENSITY = 1000.0


def _geom_volume(gtype, size):
    if gtype == "sphere":
        return 4.0 / 3.0 * math.pi * size[0] ** 3
    if gtype == "capsule":
        r, h = size[0], size[1]
        return math.pi * r * r * (2 * h) + 4.0 / 3.0 * math.pi * r ** 3
    if gtype == "cylinder":
        return math.pi * size[0] ** 2 * 2 * size[1]
    if gtype == "box":
        return 8.0 * size[0] * size[1] * size[2]
    raise ValueError("volume of geom type %r not supported (needs a mesh)" % gtype)


def compile_mjcf(path, name=None) -> RobotModel:
    """Compile an MJCF file into a RobotModel (kinematics/inertial subset)."""
    root = ET.parse(path).getroot()
    comp = root.find("compiler")
    if comp is not None and comp.get("angle", "degree") != "radian":
        raise ValueError("only angle=radian models are supported")
    inertiafromgeom = (comp.get("inertiafromgeom", "auto") if comp is not None else "auto").lower()
    defaults = _Defaults(root)
    opt = root.find("option")
    timestep = float(opt.get("timestep", "0.002")) if opt is not None else 0.002

    body_names, parent, bpos, bquat, bipos, bmass = ["world"], [0], [[0, 0, 0]], [[1, 0, 0, 0]], [[0, 0, 0]], [0.0]
    bjntadr, bjntnum = [-1], [0]
    jn, jt, jb, jq, jd, jax, jpos, jrng, jlim = [], [], [], [], [], [], [], [], []
    per_body_geoms = {0: []}
    state = {"nq": 0, "nv": 0}

    def add_geoms(elem, bid, childclass):
        for g in elem.findall("geom"):
            a = defaults.resolve(g, "geom", childclass)
            gtype = a.get("type", "sphere")
            if "fromto" in a:
                ft = _floats(a["fromto"], 6)
                pos = [(ft[i] + ft[i + 3]) / 2 for i in range(3)]
                half = 0.5 * math.sqrt(sum((ft[i + 3] - ft[i]) ** 2 for i in range(3)))
                size = _floats(a.get("size", "0"))
                size = [size[0], half]
            else:
                pos = _floats(a.get("pos", "0 0 0"), 3)
                size = _floats(a.get("size", "0"))
            per_body_geoms[bid].append({
                "name": a.get("name", ""), "pos": pos, "type": gtype, "size": size,
                "mass": float(a["mass"]) if "mass" in a else None,
                "density": float(a.get("density", _GEOM_DENSITY)),
                "contype": int(a.get("contype", 1)), "conaffinity": int(a.get("conaffinity", 1)),
                "margin": float(a.get("margin", 0.0)), "has_quat": any(k in a for k in ("quat", "euler", "axisangle", "xyaxes", "zaxis")),
            })

    def walk(elem, pid, childclass):
        for b in elem.findall("body"):
            bid = len(body_names)
            cc = b.get("childclass", childclass)
            body_names.append(b.get("name", "body%d" % bid))
            parent.append(pid)
            bpos.append(_floats(b.get("pos", "0 0 0"), 3))
            q = np.array(_floats(b.get("quat", "1 0 0 0"), 4))
            bquat.append((q / np.linalg.norm(q)).tolist())
            per_body_geoms[bid] = []
            # joints (document order); <freejoint/> is a free joint with no defaults
            jadr, jnum = len(jn), 0
            for j in b:
                if j.tag == "freejoint":
                    a = {"type": "free", "name": j.get("name", "")}
                elif j.tag == "joint":
                    a = defaults.resolve(j, "joint", cc)
                else:
                    continue
                jtype = a.get("type", "hinge")
                jn.append(a.get("name", ""))
                jb.append(bid)
                jq.append(state["nq"])
                jd.append(state["nv"])
                jpos.append(_floats(a.get("pos", "0 0 0"), 3))
                if jtype == "free":
                    jt.append(JNT_FREE)
                    jax.append([0.0, 0.0, 1.0])
                    jrng.append([0.0, 0.0])
                    jlim.append(False)
                    state["nq"] += 7
                    state["nv"] += 6
                elif jtype == "hinge":
                    jt.append(JNT_HINGE)
                    ax = np.array(_floats(a.get("axis", "0 0 1"), 3))
                    jax.append((ax / np.linalg.norm(ax)).tolist())
                    jrng.append(_floats(a.get("range", "0 0"), 2))
                    lim = a.get("limited", "auto")
                    jlim.append(_bool(lim) if lim != "auto" else ("range" in a))
                    state["nq"] += 1
                    state["nv"] += 1
                else:
                    raise ValueError("joint type %r not supported" % jtype)
                jnum += 1
            bjntadr.append(jadr if jnum else -1)
            bjntnum.append(jnum)
            add_geoms(b, bid, cc)
            # inertial frame
            inertial = b.find("inertial")
            use_geoms = inertiafromgeom == "true" or (inertiafromgeom == "auto" and inertial is None)
            if use_geoms:
                m_tot, c = 0.0, np.zeros(3)
                for g in per_body_geoms[bid]:
                    m = g["mass"] if g["mass"] is not None else g["density"] * _geom_volume(g["type"], g["size"])
                    m_tot += m
                    c += m * np.array(g["pos"])
                bmass.append(m_tot)
                bipos.append((c / m_tot).tolist() if m_tot > 0 else [0.0, 0.0, 0.0])
            else:
                bmass.append(float(inertial.get("mass")))
                bipos.append(_floats(inertial.get("pos", "0 0 0"), 3))
            walk(b, bid, cc)

    wb = root.find("worldbody")
    add_geoms(wb, 0, None)
    walk(wb, 0, None)

    gnames, gbody, gpos, gct, gca = [], [], [], [], []
    for bid in range(len(body_names)):
        for g in per_body_geoms[bid]:
            gnames.append(g["name"])
            gbody.append(bid)
            gpos.append(g["pos"])
            gct.append(g["contype"])
            gca.append(g["conaffinity"])

    an, ajid, adof, agear, arng, alim = [], [], [], [], [], []
    act = root.find("actuator")
    if act is not None:
        for m in act:
            if m.tag != "motor":
                raise ValueError("actuator type %r not supported" % m.tag)
            a = defaults.resolve(m, "motor", None)
            jid = jn.index(a["joint"])
            an.append(a.get("name", ""))
            ajid.append(jid)
            adof.append(jd[jid])
            agear.append(_floats(a.get("gear", "1"))[0])
            arng.append(_floats(a.get("ctrlrange", "0 0"), 2))
            lim = a.get("ctrllimited", "auto")
            alim.append(_bool(lim) if lim != "auto" else ("ctrlrange" in a))

    f64 = lambda x, shape=None: np.asarray(x, dtype=np.float64).reshape(shape) if shape else np.asarray(x, dtype=np.float64)
    i64 = lambda x: np.asarray(x, dtype=np.int64)
    return RobotModel(
        name=name or root.get("model", "model"),
        nq=state["nq"], nv=state["nv"], nu=len(an), nbody=len(body_names), njnt=len(jn), ngeom=len(gnames),
        body_names=body_names, body_parentid=i64(parent), body_pos=f64(bpos), body_quat=f64(bquat),
        body_ipos=f64(bipos), body_mass=f64(bmass), body_jntadr=i64(bjntadr), body_jntnum=i64(bjntnum),
        jnt_names=jn, jnt_type=i64(jt), jnt_bodyid=i64(jb), jnt_qposadr=i64(jq), jnt_dofadr=i64(jd),
        jnt_axis=f64(jax), jnt_pos=f64(jpos), jnt_range=f64(jrng), jnt_limited=np.asarray(jlim, dtype=bool),
        geom_names=gnames, geom_bodyid=i64(gbody), geom_pos=f64(gpos, (-1, 3)), geom_contype=i64(gct),
        geom_conaffinity=i64(gca),
        actuator_names=an, actuator_jntid=i64(ajid), actuator_dofadr=i64(adof), actuator_gear=f64(agear),
        actuator_ctrlrange=f64(arng, (-1, 2)), actuator_ctrllimited=np.asarray(alim, dtype=bool),
        timestep=timestep,
        # collision geometry of the floor-contact observation (deepmimic_env.py:78-121): type, half sizes, margin
        extra={"geom_type": [g["type"] for b in range(len(body_names)) for g in per_body_geoms[b]],
               "geom_size": [(list(g["size"]) + [0.0, 0.0, 0.0])[:3] for b in range(len(body_names)) for g in per_body_geoms[b]],
               "geom_margin": [g["margin"] for b in range(len(body_names)) for g in per_body_geoms[b]],
               "geom_has_quat": [bool(g["has_quat"]) for b in range(len(body_names)) for g in per_body_geoms[b]]},
    )


def dump_models(models, path):
    with open(path, "w") as f:
        json.dump({k: m.to_dict() for k, m in models.items()}, f)


def load_models(path):
    with open(path) as f:
        d = json.load(f)
    return {k: RobotModel.from_dict(v) for k, v in d.items()}
