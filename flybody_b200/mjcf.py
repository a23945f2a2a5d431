"""Minimal MJCF reader for the fly model (host side, numpy + xml.etree only).

Resolves `<default>` classes / `childclass` inheritance and returns a mutable
spec tree (`Spec`) that the walker/task surgery in `compile_fly.py` edits
before it is flattened into SoA arrays.  Covers the MJCF subset used by
`vnl_ray/fruitfly/assets/fruitfly.xml`: bodies, hinge/free/ball joints,
primitive + mesh geoms (`fromto`, `euler`, `quat`), sites, fixed tendons,
`general`/`adhesion` actuators, contact excludes, site sensors.
"""
from __future__ import annotations

import copy
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
import numpy as np


def _floats(s):
    return np.array([float(x) for x in s.split()], dtype=np.float64)


# ----------------------------------------------------------------------------
# quaternion helpers (w, x, y, z)
def quat_mul(a, b):
    aw, ax, ay, az = a
    bw, bx, by, bz = b
    return np.array([
        aw * bw - ax * bx - ay * by - az * bz,
        aw * bx + ax * bw + ay * bz - az * by,
        aw * by - ax * bz + ay * bw + az * bx,
        aw * bz + ax * by - ay * bx + az * bw,
    ])


def quat_conj(q):
    return np.array([q[0], -q[1], -q[2], -q[3]])


def quat_rot(q, v):
    m = quat_to_mat(q)
    return m @ np.asarray(v, dtype=np.float64)


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z],
    ])


def mat_to_quat(m):
    """Rotation matrix -> unit quaternion (w>=0 branch not enforced)."""
    t = np.trace(m)
    if t > 0:
        s = np.sqrt(t + 1.0) * 2
        q = np.array([0.25 * s, (m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s])
    elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
        s = np.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
        q = np.array([(m[2, 1] - m[1, 2]) / s, 0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s])
    elif m[1, 1] > m[2, 2]:
        s = np.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
        q = np.array([(m[0, 2] - m[2, 0]) / s, (m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s])
    else:
        s = np.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
        q = np.array([(m[1, 0] - m[0, 1]) / s, (m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s])
    return q / np.linalg.norm(q)


def quat_z2vec(vec):
    """Quaternion rotating the z axis onto `vec` (MJCF `fromto` semantics)."""
    v = np.asarray(vec, dtype=np.float64)
    v = v / np.linalg.norm(v)
    axis = np.cross([0.0, 0.0, 1.0], v)
    s = np.linalg.norm(axis)
    if s < 1e-10:
        axis = np.array([1.0, 0.0, 0.0])
    else:
        axis = axis / s
    ang = np.arctan2(s, v[2])
    return np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * axis])


def euler_to_quat(e):
    """Intrinsic xyz euler angles (MJCF default eulerseq) -> quaternion."""
    q = np.array([1.0, 0, 0, 0])
    for i in range(3):
        ax = np.zeros(3)
        ax[i] = 1.0
        q = quat_mul(q, np.concatenate([[np.cos(e[i] / 2)], np.sin(e[i] / 2) * ax]))
    return q


# ----------------------------------------------------------------------------
@dataclass
class Joint:
    name: str
    type: str = "hinge"
    pos: np.ndarray = field(default_factory=lambda: np.zeros(3))
    axis: np.ndarray = field(default_factory=lambda: np.array([0.0, 0.0, 1.0]))
    range: np.ndarray | None = None
    limited: bool = False
    stiffness: float = 0.0
    springref: float = 0.0
    damping: float = 0.0
    armature: float = 0.0
    springdamper: np.ndarray | None = None
    solreflimit: np.ndarray = field(default_factory=lambda: np.array([0.02, 1.0]))
    solimplimit: np.ndarray = field(default_factory=lambda: np.array([0.9, 0.95, 0.001, 0.5, 2.0]))
    dclass: str = "main"


@dataclass
class Geom:
    name: str
    type: str = "sphere"
    size: np.ndarray = field(default_factory=lambda: np.zeros(3))
    pos: np.ndarray = field(default_factory=lambda: np.zeros(3))
    quat: np.ndarray = field(default_factory=lambda: np.array([1.0, 0, 0, 0]))
    mass: float | None = None
    density: float = 1000.0
    contype: int = 1
    conaffinity: int = 1
    condim: int = 3
    friction: np.ndarray = field(default_factory=lambda: np.array([1.0, 0.005, 0.0001]))
    solref: np.ndarray = field(default_factory=lambda: np.array([0.02, 1.0]))
    solimp: np.ndarray = field(default_factory=lambda: np.array([0.9, 0.95, 0.001, 0.5, 2.0]))
    solmix: float = 1.0
    margin: float = 0.0
    gap: float = 0.0
    mesh: str | None = None
    fluidshape: str = "none"
    fluidcoef: np.ndarray = field(default_factory=lambda: np.array([0.5, 0.25, 1.5, 1.0, 1.0]))
    dclass: str = "main"


@dataclass
class Site:
    name: str
    type: str = "sphere"
    size: np.ndarray = field(default_factory=lambda: np.array([0.005, 0.005, 0.005]))
    pos: np.ndarray = field(default_factory=lambda: np.zeros(3))
    quat: np.ndarray = field(default_factory=lambda: np.array([1.0, 0, 0, 0]))


@dataclass
class Body:
    name: str
    pos: np.ndarray = field(default_factory=lambda: np.zeros(3))
    quat: np.ndarray = field(default_factory=lambda: np.array([1.0, 0, 0, 0]))
    joints: list = field(default_factory=list)
    geoms: list = field(default_factory=list)
    sites: list = field(default_factory=list)
    children: list = field(default_factory=list)
    parent: "Body | None" = None

    def walk(self):
        yield self
        for c in self.children:
            yield from c.walk()


@dataclass
class Tendon:
    name: str
    joints: list  # [(joint name, coef)]


@dataclass
class Actuator:
    name: str
    tag: str  # "general" | "adhesion"
    trntype: str  # "joint" | "tendon" | "body"
    target: str
    gainprm: float = 1.0
    biastype: str = "none"
    biasprm: np.ndarray = field(default_factory=lambda: np.zeros(3))
    dyntype: str = "none"
    dynprm: float = 1.0
    ctrllimited: bool = False
    ctrlrange: np.ndarray = field(default_factory=lambda: np.zeros(2))
    forcelimited: bool = False
    forcerange: np.ndarray = field(default_factory=lambda: np.zeros(2))
    gear: float = 1.0
    dclass: str = "main"


@dataclass
class Sensor:
    name: str
    type: str
    site: str


@dataclass
class Spec:
    world: Body
    tendons: list
    actuators: list
    sensors: list
    excludes: list  # [(body1, body2)]
    meshes: dict  # name -> (file, scale)
    option: dict
    defaults: dict  # class name -> {tag: {attr: str}}
    default_parent: dict  # class name -> parent class name

    def bodies(self):
        return list(self.world.walk())

    def class_is(self, cls, ancestor):
        """True if default class `cls` is `ancestor` or nested inside it."""
        while cls is not None:
            if cls == ancestor:
                return True
            cls = self.default_parent.get(cls)
        return False

    def find_body(self, name):
        for b in self.world.walk():
            if b.name == name:
                return b
        return None

    def all_joints(self):
        return [j for b in self.world.walk() for j in b.joints]

    def all_geoms(self):
        return [g for b in self.world.walk() for g in b.geoms]

    def find_site(self, name):
        for b in self.world.walk():
            for s in b.sites:
                if s.name == name:
                    return s, b
        return None, None


_ACT_TAGS = ("general", "adhesion", "motor", "position")


def _collect_defaults(node, parent_name, table, parents, inherited):
    name = node.get("class", "main")
    cur = copy.deepcopy(inherited)
    for child in node:
        if child.tag == "default":
            continue
        tag = "actuator" if child.tag in _ACT_TAGS else child.tag
        d = cur.setdefault(tag, {})
        attrs = dict(child.attrib)
        if child.tag == "adhesion":  # shortcut: gain -> gainprm[0]
            if "gain" in attrs:
                attrs["gainprm"] = attrs.pop("gain")
        d.update(attrs)
    table[name] = cur
    parents[name] = parent_name
    for child in node:
        if child.tag == "default":
            _collect_defaults(child, name, table, parents, cur)


def _pose(attrs):
    pos = _floats(attrs["pos"]) if "pos" in attrs else np.zeros(3)
    if "quat" in attrs:
        q = _floats(attrs["quat"])
        quat = q / np.linalg.norm(q)
    elif "euler" in attrs:
        quat = euler_to_quat(_floats(attrs["euler"]))
    elif "xyaxes" in attrs:
        a = _floats(attrs["xyaxes"])
        x = a[:3] / np.linalg.norm(a[:3])
        y = a[3:] - x * np.dot(x, a[3:])
        y /= np.linalg.norm(y)
        quat = mat_to_quat(np.stack([x, y, np.cross(x, y)], axis=1))
    else:
        quat = np.array([1.0, 0, 0, 0])
    return pos, quat


def _fromto(attrs):
    ft = _floats(attrs["fromto"])
    vec = ft[:3] - ft[3:]
    return 0.5 * (ft[:3] + ft[3:]), quat_z2vec(vec), 0.5 * np.linalg.norm(vec)


def parse(path: str) -> Spec:
    root = ET.parse(path).getroot()
    table, parents = {}, {}
    dnode = root.find("default")
    if dnode is not None:
        _collect_defaults(dnode, None, table, parents, {})
    else:
        table["main"] = {}
        parents["main"] = None
    option = dict(root.find("option").attrib) if root.find("option") is not None else {}
    compiler = dict(root.find("compiler").attrib) if root.find("compiler") is not None else {}
    autolimits = compiler.get("autolimits", "true") == "true"

    def resolved(tag, el, childclass):
        cls = el.get("class", childclass or "main")
        base = dict(table.get(cls, {}).get(tag, {}))
        own = dict(el.attrib)
        if el.tag == "adhesion" and "gain" in own:
            own["gainprm"] = own.pop("gain")
        base.update(own)
        return base, cls

    meshes = {}
    asset = root.find("asset")
    if asset is not None:
        for m in asset.findall("mesh"):
            a, _ = resolved("mesh", m, None)
            scale = _floats(a["scale"]) if "scale" in a else np.ones(3)
            nm = a.get("name") or a["file"].rsplit(".", 1)[0]
            meshes[nm] = (a["file"], scale)

    def make_body(el, parent, childclass):
        cc = el.get("childclass", childclass)
        pos, quat = _pose(el.attrib)
        body = Body(name=el.get("name", ""), pos=pos, quat=quat, parent=parent)
        for ch in el:
            if ch.tag in ("joint", "freejoint"):
                if ch.tag == "freejoint":
                    body.joints.append(Joint(name=ch.get("name", ""), type="free"))
                    continue
                a, jc = resolved("joint", ch, cc)
                j = Joint(name=a.get("name", ""), type=a.get("type", "hinge"), dclass=jc)
                if "pos" in a:
                    j.pos = _floats(a["pos"])
                if "axis" in a:
                    ax = _floats(a["axis"])
                    j.axis = ax / np.linalg.norm(ax)
                if "range" in a:
                    j.range = _floats(a["range"])
                if "limited" in a and a["limited"] != "auto":
                    j.limited = a["limited"] == "true"
                else:
                    j.limited = autolimits and j.range is not None
                j.stiffness = float(a.get("stiffness", 0))
                j.springref = float(a.get("springref", 0))
                j.damping = float(a.get("damping", 0))
                j.armature = float(a.get("armature", 0))
                if "springdamper" in a:
                    j.springdamper = _floats(a["springdamper"])
                if "solreflimit" in a:
                    j.solreflimit = _floats(a["solreflimit"])
                if "solimplimit" in a:
                    v = _floats(a["solimplimit"])
                    j.solimplimit[: len(v)] = v
                body.joints.append(j)
            elif ch.tag == "geom":
                a, gc = resolved("geom", ch, cc)
                g = Geom(name=a.get("name", ""), type=a.get("type", "sphere"), dclass=gc)
                size = _floats(a["size"]) if "size" in a else np.zeros(0)
                g.size = np.zeros(3)
                g.size[: len(size)] = size
                if "fromto" in a:
                    g.pos, g.quat, half = _fromto(a)
                    g.size[1] = half
                else:
                    g.pos, g.quat = _pose(a)
                if "mass" in a:
                    g.mass = float(a["mass"])
                g.density = float(a.get("density", 1000.0))
                g.contype = int(a.get("contype", 1))
                g.conaffinity = int(a.get("conaffinity", 1))
                g.condim = int(a.get("condim", 3))
                if "friction" in a:
                    v = _floats(a["friction"])
                    g.friction[: len(v)] = v
                if "solref" in a:
                    g.solref = _floats(a["solref"])
                if "solimp" in a:
                    v = _floats(a["solimp"])
                    g.solimp[: len(v)] = v
                g.solmix = float(a.get("solmix", 1.0))
                g.margin = float(a.get("margin", 0.0))
                g.gap = float(a.get("gap", 0.0))
                g.mesh = a.get("mesh")
                g.fluidshape = a.get("fluidshape", "none")
                if "fluidcoef" in a:
                    g.fluidcoef = _floats(a["fluidcoef"])
                body.geoms.append(g)
            elif ch.tag == "site":
                a, _ = resolved("site", ch, cc)
                s = Site(name=a.get("name", ""), type=a.get("type", "sphere"))
                if "size" in a:
                    v = _floats(a["size"])
                    s.size = np.zeros(3)
                    s.size[: len(v)] = v
                if "fromto" in a:
                    s.pos, s.quat, half = _fromto(a)
                    s.size[1] = half
                else:
                    s.pos, s.quat = _pose(a)
                body.sites.append(s)
            elif ch.tag == "body":
                body.children.append(make_body(ch, body, cc))
        return body

    world = Body(name="world")
    wb = root.find("worldbody")
    tmp = make_body(wb, None, None)
    world.geoms, world.sites, world.children = tmp.geoms, tmp.sites, tmp.children
    for c in world.children:
        c.parent = world

    tendons = []
    tn = root.find("tendon")
    if tn is not None:
        for t in tn.findall("fixed"):
            tendons.append(Tendon(t.get("name"), [(j.get("joint"), float(j.get("coef", 1))) for j in t.findall("joint")]))

    actuators = []
    an = root.find("actuator")
    if an is not None:
        for el in an:
            a, cls = resolved("actuator", el, None)
            if "joint" in a:
                trn, tgt = "joint", a["joint"]
            elif "tendon" in a:
                trn, tgt = "tendon", a["tendon"]
            else:
                trn, tgt = "body", a["body"]
            act = Actuator(name=a["name"], tag=el.tag, trntype=trn, target=tgt, dclass=cls)
            if "gainprm" in a:
                act.gainprm = _floats(a["gainprm"])[0]
            act.biastype = a.get("biastype", "none")
            if "biasprm" in a:
                v = _floats(a["biasprm"])
                act.biasprm[: len(v)] = v
            act.dyntype = a.get("dyntype", "none")
            if "dynprm" in a:
                act.dynprm = _floats(a["dynprm"])[0]
            if "ctrlrange" in a:
                act.ctrlrange = _floats(a["ctrlrange"])
            cl = a.get("ctrllimited", "auto")
            act.ctrllimited = (cl == "true") or (cl == "auto" and autolimits and "ctrlrange" in a)
            if el.tag == "adhesion":
                act.ctrllimited = True
                act.biastype = "none"
            if "forcerange" in a:
                act.forcerange = _floats(a["forcerange"])
            fl = a.get("forcelimited", "auto")
            act.forcelimited = (fl == "true") or (fl == "auto" and autolimits and "forcerange" in a)
            if "gear" in a:
                act.gear = _floats(a["gear"])[0]
            actuators.append(act)

    sensors = []
    sn = root.find("sensor")
    if sn is not None:
        for el in sn:
            sensors.append(Sensor(name=el.get("name"), type=el.tag, site=el.get("site")))

    excludes = []
    cn = root.find("contact")
    if cn is not None:
        for el in cn.findall("exclude"):
            excludes.append((el.get("body1"), el.get("body2")))

    return Spec(world=world, tendons=tendons, actuators=actuators, sensors=sensors, excludes=excludes,
                meshes=meshes, option=option, defaults=table, default_parent=parents)
