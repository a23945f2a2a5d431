"""Host model compiler: fly MJCF + walker/task surgery -> flat SoA model arrays.

Runs once (here, where `/root/reference` is mounted) and writes
`flybody_b200/assets/*.npz`; the GPU box only loads those files.

What is restated, with the reference lines it follows:
  * walker surgery        `vnl_ray/fruitfly/fruitfly.py:115-326` (`FruitFly._build`)
  * ghost construction    `vnl_ray/tasks/task_utils.py:79-114`   (`make_ghost_fly`)
  * task assembly         `vnl_ray/tasks/base.py:29-168,264-364`, `walk_imitation.py:27-84`,
                          `walk_on_ball.py:19-47`, `flight_imitation.py:24-85`, `fly_envs.py:29-157`
  * ball arena            `vnl_ray/tasks/arenas/ball.py:64-72`
MuJoCo's own model compilation (defaults, fromto, mesh inertia, inertial-frame
merge, invweight0, meaninertia, springdamper) is third-party and restated from
its documented behaviour.
"""
from __future__ import annotations

import copy
import os
import numpy as np

from . import mjcf
from .mjcf import Body, Geom, Joint, Site, quat_mul, quat_conj, quat_rot, quat_to_mat, mat_to_quat
from .meshes import load_mesh, mesh_mass_properties

_LEG = ("T1", "T2", "T3")
_NAME_SUBSTR = {  # fruitfly.py:174-183
    "adhesion": [],
    "head": ["head"],
    "mouth": ["rostrum", "haustellum", "labrum"],
    "antennae": ["antenna"],
    "wings": ["wing"],
    "abdomen": ["abdomen"],
    "legs": list(_LEG),
    "user": [],
}
_ACTION_CLASSES = ("adhesion", "head", "mouth", "antennae", "wings", "abdomen", "legs", "user")  # fruitfly.py:25
SPAWN_POS = np.array([0.0, 0.0, 0.1278])  # fruitfly.py:23

MJ_MINVAL = 1e-15
GEOM_TYPES = {"plane": 0, "sphere": 2, "capsule": 3, "ellipsoid": 4, "cylinder": 5, "box": 6, "mesh": 7}
JNT_FREE, JNT_BALL, JNT_HINGE = 0, 1, 3


def _any_in(subs, s):
    return any(x in s for x in subs)


def _neg_quat(q):
    r = np.array(q, dtype=np.float64)
    r[0] *= -1
    return r


def _body_quat_from_springrefs(body: Body):
    """fruitfly.py:60-80."""
    if not body.joints:
        return None
    quats = [np.concatenate([[np.cos(j.springref / 2)], np.sin(j.springref / 2) * j.axis]) for j in body.joints]
    quat = np.array([1.0, 0, 0, 0])
    for i in range(len(quats)):
        quat = quat_mul(quats[-1 - i], quat)
    return quat_mul(body.quat, quat)


def _change_body_frame(body: Body, frame_pos, frame_quat):
    """fruitfly.py:83-106: re-frame a body, keeping its children where they were."""
    dpos = body.pos - frame_pos
    body_quat = body.quat.copy()
    dquat = quat_mul(_neg_quat(frame_quat), body_quat)
    body.pos = np.array(frame_pos, dtype=np.float64)
    body.quat = np.array(frame_quat, dtype=np.float64)
    for child in list(body.joints) + list(body.geoms) + list(body.sites) + list(body.children):
        if hasattr(child, "quat"):
            child.quat = quat_mul(dquat, child.quat)
        pos_in_parent = quat_rot(body_quat, child.pos) + dpos
        child.pos = quat_rot(_neg_quat(frame_quat), pos_in_parent)
        if isinstance(child, Joint):  # joints carry an axis instead of a quat; all_children() in dm_control
            pass                      # leaves joint.axis untouched (no `quat` attribute)


class Walker:
    """Result of the `FruitFly._build` surgery on a parsed spec."""

    def __init__(self, xml_path, name="walker", use_legs=True, use_wings=False, use_mouth=False,
                 use_antennae=False, joint_filter=0.01, adhesion_filter=0.007, body_pitch_angle=47.5,
                 stroke_plane_angle=0.0, num_user_actions=0):
        spec = mjcf.parse(xml_path)
        self.spec = spec
        self.name = name
        self.use_wings = use_wings
        thorax = spec.find_body("thorax")
        thorax.joints = [j for j in thorax.joints if j.type != "free"]  # fruitfly.py:164
        observable = [j.name for j in spec.all_joints()]  # legacy_base: all joints, before removals

        def remove_actuator(nm):
            spec.actuators = [a for a in spec.actuators if a.name != nm]

        if not use_legs:  # fruitfly.py:188-219
            for b in spec.bodies():
                if _any_in(_LEG, b.name):
                    q = _body_quat_from_springrefs(b)
                    if q is not None:
                        b.quat = q
            for t in list(spec.tendons):
                if _any_in(_LEG, t.name):
                    remove_actuator(t.name)
                    spec.tendons.remove(t)
            for b in spec.bodies():
                for j in list(b.joints):
                    if _any_in(_LEG, j.name):
                        remove_actuator(j.name)
                        observable.remove(j.name)
                        b.joints.remove(j)
            spec.actuators = [a for a in spec.actuators if not ("adhere" in a.name and _any_in(_LEG, a.name))]
            spec.sensors = [s for s in spec.sensors if not _any_in(_LEG, s.name)]
        for flag, key in ((use_wings, "wings"), (use_mouth, "mouth"), (use_antennae, "antennae")):
            if flag:
                continue
            for j in spec.all_joints():  # fruitfly.py:221-247
                if _any_in(_NAME_SUBSTR[key], j.name):
                    remove_actuator(j.name)
                    observable.remove(j.name)
            if key == "mouth":
                spec.actuators = [a for a in spec.actuators
                                  if not ("adhere" in a.name and _any_in(_NAME_SUBSTR["mouth"], a.name))]
        self.observable_joints = observable

        if use_wings:  # fruitfly.py:249-269
            site, _ = spec.find_site("hover_up_dir")
            up_dir = site.quat
            delta = np.deg2rad(body_pitch_angle) - 2 * np.arccos(up_dir[0])
            dquat = np.array([np.cos(delta / 2), 0, np.sin(delta / 2), 0])
            up_dir = quat_mul(dquat, up_dir)
            site.quat = up_dir
            spa = np.deg2rad(stroke_plane_angle)
            sp_quat = np.array([np.cos(spa / 2), 0, np.sin(spa / 2), 0])
            for quat, wing in ((np.array([0.0, 0, 0, 1]), "wing_left"), (np.array([0.0, -1, 0, 0]), "wing_right")):
                dq = quat_mul(_neg_quat(sp_quat), quat)
                new_q = quat_mul(dq, _neg_quat(up_dir))
                body = spec.find_body(wing)
                _change_body_frame(body, body.pos.copy(), new_q)

        for a in spec.actuators:  # fruitfly.py:271-281
            if a.tag != "adhesion" and joint_filter > 0:
                a.dyntype, a.dynprm = "filter", joint_filter
            if a.tag == "adhesion" and adhesion_filter > 0:
                a.dyntype, a.dynprm = "filter", adhesion_filter

        names = [a.name for a in spec.actuators]  # fruitfly.py:283-318
        self.ctrl_indices = {}
        for k in _ACTION_CLASSES:
            idx = [i for i, n in enumerate(names) if _any_in(_NAME_SUBSTR[k], n) and "adhere" not in n]
            self.ctrl_indices[k] = idx or None
        idx = [i for i, n in enumerate(names) if "adhere" in n]
        self.ctrl_indices["adhesion"] = idx or None
        self.num_actions = {k: (len(self.ctrl_indices[k]) if self.ctrl_indices[k] else 0) for k in _ACTION_CLASSES}
        self.num_actions["user"] = num_user_actions
        self.action_indices, c = {}, 0
        for k in _ACTION_CLASSES:
            self.action_indices[k] = list(range(c, c + self.num_actions[k]))
            c += self.num_actions[k]
        self.action_dim = c


def make_ghost(xml_path):
    """`walker(name="ghost", use_wings=False, use_legs=False)` + `make_ghost_fly`
    (tasks/base.py:141-149, task_utils.py:79-114): mesh geoms only, no joints."""
    w = Walker(xml_path, name="ghost", use_legs=False, use_wings=False)
    spec = w.spec
    spec.tendons, spec.actuators, spec.excludes = [], [], []
    spec.sensors = [s for s in spec.sensors if s.type not in ("touch", "force")]

    def strip(b: Body):
        b.joints = []
        b.children = [c for c in b.children if not c.name.startswith("wing")]
        b.geoms = [g for g in b.geoms if g.mesh is not None]
        for g in b.geoms:
            g.contype = g.conaffinity = 0
        for c in b.children:
            strip(c)

    strip(spec.world)
    return w


# ----------------------------------------------------------------------------
# Task assembly (arena + walker + ghost), mirroring the reference constructors.
def _prefix(w: Walker, prefix: str):
    spec = w.spec
    for b in spec.bodies():
        if b is spec.world:
            continue
        b.name = prefix + b.name
        for e in b.joints + b.geoms + b.sites:
            e.name = prefix + e.name
    for t in spec.tendons:
        t.name = prefix + t.name
        t.joints = [(prefix + j, c) for j, c in t.joints]
    for a in spec.actuators:
        a.name = prefix + a.name
        a.target = prefix + a.target
    for s in spec.sensors:
        s.name = prefix + s.name
        s.site = prefix + s.site
    spec.excludes = [(prefix + a, prefix + b) for a, b in spec.excludes]


class TaskSpec:
    """Composite scene: world body tree + merged element lists + task metadata."""

    def __init__(self, task: str, xml_path: str):
        assert task in ("walk_imitation", "walk_on_ball", "flight_imitation", "template")
        self.task = task
        self.xml_path = xml_path
        flight = task == "flight_imitation"
        if flight:  # fly_envs.py:55-65, tasks/base.py:284-293
            self.walker = Walker(xml_path, use_legs=False, use_wings=True, joint_filter=0.0,
                                 adhesion_filter=0.007, num_user_actions=1)
            self.physics_dt, self.control_dt = 5e-5, 2e-4  # tasks/constants.py:22-23
            self.time_limit, self.future_steps = 0.6, 5
        else:  # fly_envs.py:104-115,144-150, tasks/base.py:343-352
            self.walker = Walker(xml_path, use_legs=True, use_wings=False, joint_filter=0.01, adhesion_filter=0.007)
            self.physics_dt, self.control_dt = 2e-4, 2e-3  # tasks/constants.py:16-17
            self.time_limit = 2.0 if task == "walk_on_ball" else 10.0
            self.future_steps = 64 if task == "walk_imitation" else 0
        self.n_substeps = int(round(self.control_dt / self.physics_dt))
        w = self.walker
        spec = w.spec
        _prefix(w, "walker/")

        world = Body(name="world")
        self.world = world
        self.excludes = list(spec.excludes)
        if task == "walk_on_ball":  # fly_envs.py:136-141, arenas/ball.py:64-72
            ball = Body(name="ball", pos=np.array([-0.05, 0.0, -0.419]), parent=world)
            ball.geoms.append(Geom(name="ball_geom", type="sphere", size=np.array([0.454, 0, 0]), density=0.0025))
            ball.joints.append(Joint(name="ball", type="ball"))
            world.children.append(ball)
            ground = ball.geoms[0]
        else:  # dm_control floors.Floor: an infinite plane at z=0 with default contact parameters [3P]
            ground = Geom(name="groundplane", type="plane", size=np.array([8.0, 8.0, 0.25]))
            world.geoms.append(ground)
        # attachment frame at the spawn site (tasks/base.py:127-133)
        frame = Body(name="walker/", pos=SPAWN_POS.copy(), parent=world)
        if task != "walk_on_ball":
            frame.joints.append(Joint(name="walker/", type="free"))  # create_root_joints
        frame.children = spec.world.children
        for c in frame.children:
            c.parent = frame
        world.children.append(frame)

        if not flight:  # tasks/base.py:354-358
            ground.friction = np.array([0.5, 0.005, 0.0001])
            ground.solref = np.array([0.001, 1.0])
            ground.solimp = np.array([0.95, 0.99, 0.01, 0.5, 2.0])
            for g in spec.all_geoms():  # walk_imitation.py:73-75, walk_on_ball.py:42-44 (claw_friction=1.0)
                if spec.class_is(g.dclass, "adhesion-collision"):
                    g.friction[0] = 1.0
        else:  # tasks/base.py:296-325
            ground.contype = ground.conaffinity = 0
            gains = {"yaw": 18.0, "roll": 18.0, "pitch": 18.0}  # tasks/constants.py:31
            for a in spec.actuators:
                if a.dclass in gains:
                    a.gainprm = gains[a.dclass]
            for g in spec.all_geoms():
                if "fluid" in g.name:
                    g.fluidshape = "ellipsoid"
                    g.fluidcoef = np.array([1.0, 0.5, 1.5, 1.7, 1.0])
            for j in spec.all_joints():
                if spec.class_is(j.dclass, "wing"):
                    j.stiffness, j.damping = 0.01, 0.007769230
        if task == "walk_on_ball":  # walk_on_ball.py:33-40
            thorax = spec.find_body("walker/thorax")
            for c in thorax.children:
                self.excludes.append(("walker/thorax", c.name))

        self.ghost = None
        if task in ("walk_imitation", "flight_imitation"):  # tasks/base.py:141-149
            g = make_ghost(xml_path)
            _prefix(g, "ghost/")
            gframe = Body(name="ghost/", pos=SPAWN_POS.copy(), parent=world)
            gframe.joints.append(Joint(name="ghost/", type="free", armature=1.0))
            gframe.children = g.spec.world.children
            for c in gframe.children:
                c.parent = gframe
            world.children.append(gframe)
            self.ghost = g
        self.tendons = spec.tendons
        self.actuators = spec.actuators
        self.sensors = spec.sensors
        self.meshes = spec.meshes
        self.option = spec.option


# ----------------------------------------------------------------------------
# Flattening: spec tree -> SoA arrays.  Bodies without joints are fused into
# their parent (rigidly welded => dynamically one body); their individual
# inertia boxes are kept as "fluid elements" so the per-body fluid drag of the
# unfused model is reproduced exactly.
_FALLBACK_MESH_DIR = "../build_fruitfly/assets"


class _MeshCache:
    def __init__(self, asset_dir, meshes):
        self.asset_dir, self.meshes, self.cache = asset_dir, meshes, {}
        self.missing = []

    def props(self, name):
        if name not in self.cache:
            fname, scale = self.meshes[name]
            m = load_mesh(self.asset_dir, fname, os.path.join(self.asset_dir, _FALLBACK_MESH_DIR))
            if m is None:
                self.missing.append(name)
                self.cache[name] = None
            else:
                self.cache[name] = mesh_mass_properties(m[0], m[1], scale)
        return self.cache[name]


def _geom_inertial(g: Geom, mc: _MeshCache):
    """(mass, com[3], inertia[3,3] about com) of one geom in its body's frame."""
    R = quat_to_mat(g.quat)
    if g.type == "mesh":
        p = mc.props(g.mesh)
        if p is None:
            return None
        vol, c, I = p
        com = g.pos + R @ c
    else:
        s = g.size
        com = g.pos.copy()
        if g.type == "sphere":
            vol = 4.0 / 3.0 * np.pi * s[0] ** 3
            I = np.eye(3) * (0.4 * s[0] ** 2) * vol
        elif g.type == "box":
            vol = 8 * s[0] * s[1] * s[2]
            I = np.diag([s[1] ** 2 + s[2] ** 2, s[0] ** 2 + s[2] ** 2, s[0] ** 2 + s[1] ** 2]) * vol / 3.0
        elif g.type == "ellipsoid":
            vol = 4.0 / 3.0 * np.pi * s[0] * s[1] * s[2]
            I = np.diag([s[1] ** 2 + s[2] ** 2, s[0] ** 2 + s[2] ** 2, s[0] ** 2 + s[1] ** 2]) * vol / 5.0
        elif g.type == "cylinder":
            r, h = s[0], 2 * s[1]
            vol = np.pi * r * r * h
            I = np.diag([(3 * r * r + h * h) / 12.0] * 2 + [r * r / 2.0]) * vol
        elif g.type == "capsule":
            r, h = s[0], 2 * s[1]
            vc, vs = np.pi * r * r * h, 4.0 / 3.0 * np.pi * r ** 3
            vol = vc + vs
            ixx = vc * (3 * r * r + h * h) / 12.0 + vs * (0.4 * r * r + 0.25 * h * h + 0.375 * r * h)
            I = np.diag([ixx, ixx, vc * r * r / 2.0 + vs * 0.4 * r * r])
        else:
            return None
    if g.mass is not None:
        if g.mass == 0.0:
            return None
        mass = g.mass
        I = I * (mass / vol)
    else:
        if g.density == 0.0:
            return None
        mass = g.density * vol
        I = I * g.density
    return mass, com, R @ I @ R.T


def _merge(parts):
    """Parallel-axis merge of [(mass, com, I_about_com)] -> (mass, com, I)."""
    mass = sum(p[0] for p in parts)
    com = sum(p[0] * p[1] for p in parts) / mass
    I = np.zeros((3, 3))
    for m, c, Ic in parts:
        d = c - com
        I += Ic + m * (np.dot(d, d) * np.eye(3) - np.outer(d, d))
    return mass, com, I


def _principal(I):
    """Eigen-decomposition -> (iquat, moments) with a right-handed frame, moments descending."""
    w, V = np.linalg.eigh(I)
    order = np.argsort(-w)
    w, V = w[order], V[:, order]
    if np.linalg.det(V) < 0:
        V[:, 2] *= -1
    return mat_to_quat(V), w


def _compose(pa, qa, pb, qb):
    """Pose b (given in frame a) expressed in a's parent frame."""
    return pa + quat_rot(qa, pb), quat_mul(qa, qb)


def flatten(ts: TaskSpec) -> dict:
    """TaskSpec -> dict of numpy arrays (the packed model; see include/flyb_model.h)."""
    mc = _MeshCache(os.path.dirname(ts.xml_path), ts.meshes)
    dt = ts.physics_dt
    opt = ts.option
    gravity = np.array([float(x) for x in opt.get("gravity", "0 0 -9.81").split()])
    density, viscosity = float(opt.get("density", 0)), float(opt.get("viscosity", 0))

    # --- pass 1: fused bodies -------------------------------------------------
    fb = []  # fused bodies: dict(name, parent, pos, quat, joints, parts, spec_names)
    owner_of = {}  # spec body name -> (fused id, rel pos, rel quat)
    geoms, sites, fluid_boxes, fluid_geoms = [], [], [], []
    head_red_fix = []

    def visit(b: Body, owner, rp, rq):
        # (rp, rq): pose of b's frame in its owner's frame
        if b.parent is None:
            owner = 0
            fb.append(dict(name="world", parent=-1, pos=np.zeros(3), quat=np.array([1.0, 0, 0, 0]), joints=[], parts=[]))
            rp, rq = np.zeros(3), np.array([1.0, 0, 0, 0])
        elif b.joints:
            pp, pq = _compose(rp, rq, b.pos, b.quat)  # pose in parent-owner frame
            fb.append(dict(name=b.name, parent=owner, pos=pp, quat=pq, joints=b.joints, parts=[]))
            owner = len(fb) - 1
            rp, rq = np.zeros(3), np.array([1.0, 0, 0, 0])
        else:
            rp, rq = _compose(rp, rq, b.pos, b.quat)
        owner_of[b.name] = (owner, rp, rq)
        R = quat_to_mat(rq)
        parts = []
        for g in b.geoms:
            gi = _geom_inertial(g, mc)
            if gi is None and g.type == "mesh" and g.mesh in mc.missing and g.mass != 0.0:
                head_red_fix.append((g, parts))
            if gi is not None:
                parts.append(gi)
            gp, gq = _compose(rp, rq, g.pos, g.quat)
            if g.type != "mesh" and (g.contype or g.conaffinity):
                geoms.append(dict(g=g, body=owner, pos=gp, quat=gq, spec_body=b.name))
            if g.fluidshape == "ellipsoid":
                fluid_geoms.append(dict(g=g, body=owner, pos=gp, quat=gq, spec_body=b.name))
        for g, plist in head_red_fix:
            if plist is parts:
                # `head_red.obj/.msh` is not in the checkout (.MISSING_LARGE_BLOBS).  Stand-in: a copy of the
                # head_body mass distribution scaled so that the head subtree (head, eyes, ocelli, mouth parts,
                # antennae) weighs 0.15 mg, the empirical head mass the model was built to
                # (build_fruitfly/make_fruitfly.py:24,1245-1246).
                have = 0.0
                for sb in b.walk():
                    for sg in sb.geoms:
                        si = _geom_inertial(sg, mc)
                        have += si[0] if si is not None else 0.0
                ref = parts[0]
                extra = max(0.00015 - have, 0.0)
                parts.append((extra, ref[1], ref[2] * (extra / ref[0])))
        if parts:
            m, c, I = _merge(parts)
            if owner != 0:
                fb[owner]["parts"].append((m, rp + R @ c, R @ I @ R.T))
            if m >= MJ_MINVAL and owner != 0:
                iq, mom = _principal(I)
                has_ell = any(g.fluidshape == "ellipsoid" for g in b.geoms)
                if not has_ell:
                    box = np.sqrt(np.maximum(MJ_MINVAL, [mom[1] + mom[2] - mom[0], mom[0] + mom[2] - mom[1],
                                                         mom[0] + mom[1] - mom[2]]) / m * 6.0)
                    p, q = _compose(rp, rq, c, iq)
                    fluid_boxes.append(dict(body=owner, pos=p, quat=q, box=box, mass=m, name=b.name))
        for s in b.sites:
            sp, sq = _compose(rp, rq, s.pos, s.quat)
            sites.append(dict(s=s, body=owner, pos=sp, quat=sq))
        for c in b.children:
            visit(c, owner, rp, rq)

    visit(ts.world, 0, None, None)
    nbody = len(fb)

    # --- joints / dofs ----------------------------------------------------------
    M = {}
    jn, dn = [], []  # joint records, dof records
    nq = nv = 0
    for bi, b in enumerate(fb):
        b["jntadr"], b["dofadr"] = len(jn), nv
        for j in b["joints"]:
            t = {"free": JNT_FREE, "ball": JNT_BALL, "hinge": JNT_HINGE}[j.type]
            ndof = {JNT_FREE: 6, JNT_BALL: 3, JNT_HINGE: 1}[t]
            nqj = {JNT_FREE: 7, JNT_BALL: 4, JNT_HINGE: 1}[t]
            jn.append(dict(j=j, type=t, body=bi, qposadr=nq, dofadr=nv))
            for k in range(ndof):
                if k == 0:
                    if len(dn) > b["dofadr"]:
                        par = len(dn) - 1
                    else:  # last dof of nearest ancestor with dofs
                        p, par = b["parent"], -1
                        while p > 0:
                            if fb[p]["dofnum"] > 0:
                                par = fb[p]["dofadr"] + fb[p]["dofnum"] - 1
                                break
                            p = fb[p]["parent"]
                else:
                    par = len(dn) - 1
                dn.append(dict(body=bi, jnt=len(jn) - 1, parent=par, armature=j.armature, damping=j.damping))
            nq += nqj
            nv += ndof
        b["jntnum"], b["dofnum"] = len(jn) - b["jntadr"], nv - b["dofadr"]
    njnt = len(jn)
    jid = {r["j"].name: i for i, r in enumerate(jn)}

    # --- body inertials ---------------------------------------------------------
    body_mass, body_ipos, body_iquat, body_inertia = (np.zeros(nbody), np.zeros((nbody, 3)),
                                                      np.tile([1.0, 0, 0, 0], (nbody, 1)), np.zeros((nbody, 3)))
    for bi, b in enumerate(fb):
        if b["parts"]:
            m, c, I = _merge(b["parts"])
            body_mass[bi], body_ipos[bi] = m, c
            body_iquat[bi], body_inertia[bi] = _principal(I)
    body_parent = np.array([b["parent"] for b in fb], dtype=np.int32)
    body_rootid = np.zeros(nbody, dtype=np.int32)
    for bi in range(1, nbody):
        body_rootid[bi] = bi if body_parent[bi] == 0 else body_rootid[body_parent[bi]]
    subtreemass = body_mass.copy()
    for bi in range(nbody - 1, 0, -1):
        subtreemass[body_parent[bi]] += subtreemass[bi]

    qpos0, qpos_spring = np.zeros(nq), np.zeros(nq)
    for r in jn:
        a, b = r["qposadr"], fb[r["body"]]
        if r["type"] == JNT_FREE:
            qpos0[a:a + 3], qpos0[a + 3:a + 7] = b["pos"], b["quat"]
            qpos_spring[a:a + 7] = qpos0[a:a + 7]
        elif r["type"] == JNT_BALL:
            qpos0[a:a + 4] = qpos_spring[a:a + 4] = [1, 0, 0, 0]
        else:
            qpos_spring[a] = r["j"].springref

    M["nq"], M["nv"], M["nbody"], M["njnt"] = nq, nv, nbody, njnt
    M["body_parent"], M["body_rootid"] = body_parent, body_rootid
    M["body_jntadr"] = np.array([b["jntadr"] for b in fb], np.int32)
    M["body_jntnum"] = np.array([b["jntnum"] for b in fb], np.int32)
    M["body_dofadr"] = np.array([b["dofadr"] for b in fb], np.int32)
    M["body_dofnum"] = np.array([b["dofnum"] for b in fb], np.int32)
    M["body_pos"] = np.array([b["pos"] for b in fb])
    M["body_quat"] = np.array([b["quat"] for b in fb])
    M["body_ipos"], M["body_iquat"], M["body_mass"], M["body_inertia"] = body_ipos, body_iquat, body_mass, body_inertia
    M["body_subtreemass"] = subtreemass
    M["jnt_type"] = np.array([r["type"] for r in jn], np.int32)
    M["jnt_qposadr"] = np.array([r["qposadr"] for r in jn], np.int32)
    M["jnt_dofadr"] = np.array([r["dofadr"] for r in jn], np.int32)
    M["jnt_bodyid"] = np.array([r["body"] for r in jn], np.int32)
    M["jnt_pos"] = np.array([r["j"].pos for r in jn])
    M["jnt_axis"] = np.array([r["j"].axis for r in jn])
    M["jnt_stiffness"] = np.array([r["j"].stiffness for r in jn])
    M["jnt_limited"] = np.array([int(r["j"].limited and r["type"] == JNT_HINGE) for r in jn], np.int32)
    M["jnt_range"] = np.array([r["j"].range if r["j"].range is not None else [0.0, 0.0] for r in jn])
    M["jnt_solref"] = np.array([r["j"].solreflimit for r in jn])
    M["jnt_solimp"] = np.array([r["j"].solimplimit for r in jn])
    M["qpos0"], M["qpos_spring"] = qpos0, qpos_spring
    M["dof_bodyid"] = np.array([d["body"] for d in dn], np.int32)
    M["dof_jntid"] = np.array([d["jnt"] for d in dn], np.int32)
    M["dof_parentid"] = np.array([d["parent"] for d in dn], np.int32)
    M["dof_armature"] = np.array([d["armature"] for d in dn])
    M["dof_damping"] = np.array([d["damping"] for d in dn])
    M["_names_body"] = [b["name"] for b in fb]
    M["_names_jnt"] = [r["j"].name for r in jn]
    M["_owner_of"] = owner_of
    M["_geoms"], M["_sites"], M["_fluid_boxes"], M["_fluid_geoms"], M["_jid"] = geoms, sites, fluid_boxes, fluid_geoms, jid
    M["_missing_meshes"] = list(mc.missing)
    M["opt_timestep"], M["opt_gravity"], M["opt_density"], M["opt_viscosity"] = dt, gravity, density, viscosity
    return M
