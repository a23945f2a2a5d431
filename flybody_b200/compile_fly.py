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
GEOM_TYPES = {"plane": 0, "hfield": 1, "sphere": 2, "capsule": 3, "ellipsoid": 4, "cylinder": 5, "box": 6, "mesh": 7}
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
        assert task in ("walk_imitation", "walk_on_ball", "flight_imitation", "template", "vision_flight")
        self.task = task
        self.xml_path = xml_path
        flight = task in ("flight_imitation", "vision_flight")
        self.flight, self.vision = flight, task == "vision_flight"
        self.hfield = None
        if flight:  # fly_envs.py:55-65, 191-200, tasks/base.py:284-293, tasks/vision_flight.py:49-55
            self.walker = Walker(xml_path, use_legs=False, use_wings=True, joint_filter=0.0,
                                 adhesion_filter=0.007, num_user_actions=1)
            self.physics_dt, self.control_dt = 5e-5, 2e-4  # tasks/constants.py:22-23
            self.time_limit, self.future_steps = (0.4, 0) if self.vision else (0.6, 5)
        else:  # fly_envs.py:104-115,144-150, tasks/base.py:343-352
            self.walker = Walker(xml_path, use_legs=True, use_wings=False, joint_filter=0.01, adhesion_filter=0.007)
            self.physics_dt, self.control_dt = 2e-4, 2e-3  # tasks/constants.py:16-17
            self.time_limit = {"walk_on_ball": 2.0, "template": 1.0}.get(task, 10.0)  # fly_envs.py:103,143,213
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
        elif self.vision:  # tasks/arenas/hills.py:160-221: the `terrain` height field at z = -0.01, then the ground plane
            dim, density = 20.0, 10
            n = int((2 * density * dim) // 2) * 2 + 1
            self.hfield = dict(nrow=n, ncol=n, size=np.array([dim, dim, 1.0, 0.05]))
            world.geoms.append(Geom(name="terrain", type="hfield", size=np.array([dim, dim, 1.0]), pos=np.array([0.0, 0.0, -0.01])))
            ground = Geom(name="groundplane", type="plane", size=np.array([dim, dim, 0.5]))
            world.geoms.append(ground)
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
            if not self.vision:   # floor_contacts=False; the vision task keeps them (fly_envs.py:198)
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

        if self.vision:
            # the two eye cameras of the head (fruitfly.xml:335-336) as sites: their world frames are the camera frames
            # (x right, y up, looking along -z) the ray caster starts from
            head = spec.find_body("walker/head")
            head.sites.append(mjcf.Site(name="walker/eye_right", pos=np.array([0.0219, 0.0131, 0.0]), quat=np.array([0.474, 0.688, -0.344, -0.429])))
            head.sites.append(mjcf.Site(name="walker/eye_left", pos=np.array([-0.0219, 0.0131, 0.0]), quat=np.array([0.474, 0.688, 0.344, 0.429])))
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


# ----------------------------------------------------------------------------
def _mix_pair(g1: Geom, g2: Geom):
    """MuJoCo contact-parameter mixing for equal priority [3P]: condim/friction/margin/gap = max,
    solref/solimp = solmix-weighted mean."""
    mix = g1.solmix / (g1.solmix + g2.solmix)
    fr = np.maximum(g1.friction, g2.friction)
    if g1.solref[0] > 0 and g2.solref[0] > 0:
        solref = mix * g1.solref + (1 - mix) * g2.solref
    else:
        solref = np.minimum(g1.solref, g2.solref)
    solimp = mix * g1.solimp + (1 - mix) * g2.solimp
    return (max(g1.condim, g2.condim), fr, solref, solimp, max(g1.margin, g2.margin), max(g1.gap, g2.gap))


def _rbound(g: Geom):
    s = g.size
    return {"plane": -1.0, "sphere": s[0], "capsule": s[0] + s[1], "ellipsoid": max(s),
            "cylinder": np.hypot(s[0], s[1]), "box": np.linalg.norm(s),
            "hfield": float(np.sqrt(s[0] ** 2 + s[1] ** 2 + 25.0))}[g.type]   # covers heights up to 5 (elevation_z_range)


def finalize(M: dict, ts: TaskSpec) -> dict:
    """Second half of the flattening: geoms, candidate pairs, sites, tendons,
    actuators, sensors, fluid elements, qpos0-derived constants, task tables."""
    fb_parent = M["body_parent"]
    owner_of = M["_owner_of"]
    order = {b.name: i for i, b in enumerate(ts.world.walk())}  # unfused (MuJoCo) body numbering
    geoms = M.pop("_geoms")
    ng = len(geoms)
    for i, r in enumerate(geoms):
        r["mjbody"] = order[r["spec_body"]]
    M["ngeom"] = ng
    M["geom_type"] = np.array([GEOM_TYPES[r["g"].type] for r in geoms], np.int32)
    M["geom_bodyid"] = np.array([r["body"] for r in geoms], np.int32)
    M["geom_pos"] = np.array([r["pos"] for r in geoms])
    M["geom_quat"] = np.array([r["quat"] for r in geoms])
    M["geom_size"] = np.array([r["g"].size for r in geoms])
    M["geom_rbound"] = np.array([_rbound(r["g"]) for r in geoms])
    M["_names_geom"] = [r["g"].name for r in geoms]

    excl = set()
    for a, b in ts.excludes:
        excl.add((a, b))
        excl.add((b, a))
    pairs = []
    for i in range(ng):
        for j in range(i + 1, ng):
            a, b = geoms[i], geoms[j]
            ga, gb = a["g"], b["g"]
            if not ((ga.contype & gb.conaffinity) or (gb.contype & ga.conaffinity)):
                continue
            w1, w2 = a["body"], b["body"]
            if w1 == w2:
                continue
            wp1 = fb_parent[w1] if w1 > 0 else -1
            wp2 = fb_parent[w2] if w2 > 0 else -1
            if w1 != 0 and w2 != 0 and (w1 == wp2 or w2 == wp1):
                continue
            if (a["spec_body"], b["spec_body"]) in excl:
                continue
            key = (min(a["mjbody"], b["mjbody"]), max(a["mjbody"], b["mjbody"]))
            first, second = (i, j) if a["mjbody"] <= b["mjbody"] else (j, i)
            pairs.append((key, first, second))
    pairs.sort(key=lambda p: (p[0], p[1], p[2]))
    pg1, pg2, pcd, pfr, psr, psi, pmg, pgap = [], [], [], [], [], [], [], []
    for _, i, j in pairs:
        if M["geom_type"][i] > M["geom_type"][j]:
            i, j = j, i
        cd, fr, sr, si, mg, gp = _mix_pair(geoms[i]["g"], geoms[j]["g"])
        pg1.append(i); pg2.append(j); pcd.append(cd); pfr.append(fr); psr.append(sr); psi.append(si)
        pmg.append(mg); pgap.append(gp)
    M["npair"] = len(pairs)
    M["pair_geom1"], M["pair_geom2"] = np.array(pg1, np.int32), np.array(pg2, np.int32)
    M["pair_condim"] = np.array(pcd, np.int32)
    M["pair_friction"], M["pair_solref"], M["pair_solimp"] = np.array(pfr), np.array(psr), np.array(psi)
    M["pair_margin"], M["pair_gap"] = np.array(pmg), np.array(pgap)

    sites = M.pop("_sites")
    M["nsite"] = len(sites)
    M["site_bodyid"] = np.array([r["body"] for r in sites], np.int32)
    M["site_pos"] = np.array([r["pos"] for r in sites])
    M["site_quat"] = np.array([r["quat"] for r in sites])
    M["site_type"] = np.array([GEOM_TYPES[r["s"].type] for r in sites], np.int32)
    M["site_size"] = np.array([r["s"].size for r in sites])
    M["_names_site"] = [r["s"].name for r in sites]
    sid = {n: i for i, n in enumerate(M["_names_site"])}
    jid = M.pop("_jid")

    # fixed tendons
    tadr, tnum, wdof, wqpos, wcoef, tid = [], [], [], [], [], {}
    for t in ts.tendons:
        tid[t.name] = len(tadr)
        tadr.append(len(wdof)); tnum.append(len(t.joints))
        for jn_, c in t.joints:
            j = jid[jn_]
            wdof.append(M["jnt_dofadr"][j]); wqpos.append(M["jnt_qposadr"][j]); wcoef.append(c)
    M["ntendon"], M["nwrap"] = len(tadr), len(wdof)
    M["tendon_adr"], M["tendon_num"] = np.array(tadr, np.int32), np.array(tnum, np.int32)
    M["wrap_dofadr"], M["wrap_qposadr"], M["wrap_coef"] = np.array(wdof, np.int32), np.array(wqpos, np.int32), np.array(wcoef)

    # actuators
    acts = ts.actuators
    nu = len(acts)
    trn = {"joint": 0, "tendon": 1, "body": 2}
    M["nu"] = nu
    M["act_trntype"] = np.array([trn[a.trntype] for a in acts], np.int32)
    tr = []
    for a in acts:
        if a.trntype == "joint":
            tr.append(jid[a.target])
        elif a.trntype == "tendon":
            tr.append(tid[a.target])
        else:
            tr.append(owner_of[a.target][0])
    M["act_trnid"] = np.array(tr, np.int32)
    M["act_gainprm"] = np.array([a.gainprm for a in acts], np.float64)
    M["act_biasprm"] = np.array([a.biasprm if a.biastype == "affine" else np.zeros(3) for a in acts]).reshape(nu, 3)
    M["act_dyntype"] = np.array([1 if a.dyntype == "filter" else 0 for a in acts], np.int32)
    M["act_dynprm"] = np.array([a.dynprm for a in acts], np.float64)
    M["act_ctrllimited"] = np.array([int(a.ctrllimited) for a in acts], np.int32)
    M["act_ctrlrange"] = np.array([a.ctrlrange for a in acts]).reshape(nu, 2)
    M["act_forcelimited"] = np.array([int(a.forcelimited) for a in acts], np.int32)
    M["act_forcerange"] = np.array([a.forcerange for a in acts]).reshape(nu, 2)
    M["act_gear"] = np.array([a.gear for a in acts], np.float64)
    adr, na = [], 0
    for a in acts:
        if a.dyntype == "filter":
            adr.append(na); na += 1
        else:
            adr.append(-1)
    M["na"] = na
    M["act_actadr"] = np.array(adr, np.int32)
    M["_names_act"] = [a.name for a in acts]

    # sensors (site sensors only); sensordata layout in MJCF order
    st = {"accelerometer": 0, "gyro": 1, "velocimeter": 2, "force": 3, "touch": 4}
    sens = [s for s in ts.sensors if s.site in sid]
    M["nsensor"] = len(sens)
    M["sensor_type"] = np.array([st[s.type] for s in sens], np.int32)
    M["sensor_site"] = np.array([sid[s.site] for s in sens], np.int32)
    sadr, n = [], 0
    for s in sens:
        sadr.append(n); n += 1 if s.type == "touch" else 3
    M["sensor_adr"], M["nsensordata"] = np.array(sadr, np.int32), n
    M["_names_sensor"] = [s.name for s in sens]

    # fluid elements
    fbx = M.pop("_fluid_boxes")
    use_fluid = M["opt_density"] > 0 or M["opt_viscosity"] > 0
    if not use_fluid:
        fbx = []
    M["nfluidbox"] = len(fbx)
    M["fbox_bodyid"] = np.array([r["body"] for r in fbx], np.int32)
    M["fbox_pos"] = np.array([r["pos"] for r in fbx]).reshape(-1, 3)
    M["fbox_quat"] = np.array([r["quat"] for r in fbx]).reshape(-1, 4)
    M["fbox_size"] = np.array([r["box"] for r in fbx]).reshape(-1, 3)
    fg = M.pop("_fluid_geoms")
    M["nfluidgeom"] = len(fg)
    M["fgeom_bodyid"] = np.array([r["body"] for r in fg], np.int32)
    M["fgeom_pos"] = np.array([r["pos"] for r in fg]).reshape(-1, 3)
    M["fgeom_quat"] = np.array([r["quat"] for r in fg]).reshape(-1, 4)
    M["fgeom_size"] = np.array([r["g"].size for r in fg]).reshape(-1, 3)
    M["fgeom_coef"] = np.array([_ellipsoid_fluid_coef(r["g"]) for r in fg]).reshape(-1, 12)
    return M


def _ellipsoid_fluid_coef(g: Geom):
    """geom_fluid[12] = [interaction, blunt, slender, angular, kutta, magnus, virtual_mass[3],
    virtual_inertia[3]] — added-mass coefficients of an ellipsoid in potential flow (Lamb) [3P]."""
    from scipy.integrate import quad
    a, b, c = g.size
    def kappa(x, y, z):
        f = lambda l: 1.0 / ((x * x + l) * np.sqrt((x * x + l) * (y * y + l) * (z * z + l)))
        return x * y * z * quad(f, 0, np.inf, limit=400)[0]
    k = np.array([kappa(a, b, c), kappa(b, c, a), kappa(c, a, b)])
    vol = 4.0 / 3.0 * np.pi * a * b * c
    vmass = vol * k / np.maximum(1e-12, 2 - k)
    d = np.array([a, b, c])
    vin = np.zeros(3)
    for i in range(3):
        j, l = (i + 1) % 3, (i + 2) % 3
        num = (d[j] ** 2 - d[l] ** 2) ** 2 * abs(k[l] - k[j])
        den = max(1e-12, abs(2 * (d[j] ** 2 - d[l] ** 2) + (d[j] ** 2 + d[l] ** 2) * (k[j] - k[l])))
        vin[i] = vol * num / den / 5.0
    return np.concatenate([[1.0], g.fluidcoef, vmass, vin])


# ----------------------------------------------------------------------------
def _kin0(M, qpos):
    """Plain numpy forward kinematics (host-side, for compile-time constants only)."""
    nb = M["nbody"]
    xpos, xquat = np.zeros((nb, 3)), np.tile([1.0, 0, 0, 0], (nb, 1))
    xanchor, xaxis = np.zeros((M["njnt"], 3)), np.zeros((M["njnt"], 3))
    for b in range(1, nb):
        p = M["body_parent"][b]
        pos, quat = _compose(xpos[p], xquat[p], M["body_pos"][b], M["body_quat"][b])
        for j in range(M["body_jntadr"][b], M["body_jntadr"][b] + M["body_jntnum"][b]):
            a, t = M["jnt_qposadr"][j], M["jnt_type"][j]
            if t == JNT_FREE:
                pos, quat = qpos[a:a + 3].copy(), qpos[a + 3:a + 7] / np.linalg.norm(qpos[a + 3:a + 7])
                xanchor[j], xaxis[j] = pos, quat_rot(quat, M["jnt_axis"][j])
                continue
            xanchor[j] = pos + quat_rot(quat, M["jnt_pos"][j])
            xaxis[j] = quat_rot(quat, M["jnt_axis"][j])
            if t == JNT_BALL:
                ql = qpos[a:a + 4] / np.linalg.norm(qpos[a:a + 4])
            else:
                ang = qpos[a] - M["qpos0"][a]
                ql = np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * M["jnt_axis"][j]])
            quat = quat_mul(quat, ql)
            pos = xanchor[j] - quat_rot(quat, M["jnt_pos"][j])
        xpos[b], xquat[b] = pos, quat
    return xpos, xquat, xanchor, xaxis


def _mass_matrix0(M, qpos):
    """Dense M(q) = sum_b J_b^T diag(I_b, m_b) J_b + armature, and the per-body 6xnv Jacobians at the
    body CoM ([rot; lin] rows)."""
    nb, nv = M["nbody"], M["nv"]
    xpos, xquat, xanchor, xaxis = _kin0(M, qpos)
    xipos = np.array([xpos[b] + quat_rot(xquat[b], M["body_ipos"][b]) for b in range(nb)])
    J = np.zeros((nb, 6, nv))
    for b in range(1, nb):
        c = b
        while c > 0:
            R = quat_to_mat(xquat[c])
            for j in range(M["body_jntadr"][c], M["body_jntadr"][c] + M["body_jntnum"][c]):
                d, t = M["jnt_dofadr"][j], M["jnt_type"][j]
                if t == JNT_HINGE:
                    J[b, :3, d] = xaxis[j]
                    J[b, 3:, d] = np.cross(xaxis[j], xipos[b] - xanchor[j])
                else:
                    if t == JNT_FREE:
                        J[b, 3:, d:d + 3] = np.eye(3)
                        d += 3
                    for k in range(3):
                        J[b, :3, d + k] = R[:, k]
                        J[b, 3:, d + k] = np.cross(R[:, k], xipos[b] - xanchor[j])
            c = M["body_parent"][c]
    H = np.diag(M["dof_armature"]).astype(np.float64)
    for b in range(1, nb):
        Ri = quat_to_mat(quat_mul(xquat[b], M["body_iquat"][b]))
        Iw = Ri @ np.diag(M["body_inertia"][b]) @ Ri.T
        H += J[b, :3].T @ Iw @ J[b, :3] + M["body_mass"][b] * J[b, 3:].T @ J[b, 3:]
    return H, J


def constants(M: dict, ts: TaskSpec) -> dict:
    """qpos0-derived constants [3P mj_setConst]: dof_invweight0, body_invweight0, meaninertia,
    `springdamper` -> stiffness/damping."""
    nv, nb = M["nv"], M["nbody"]
    H, J = _mass_matrix0(M, M["qpos0"])
    Hinv = np.linalg.inv(H)
    M["stat_meaninertia"] = float(np.trace(H) / nv)
    dinv = np.diag(Hinv).copy()
    dof_iw = np.zeros(nv)
    for j in range(M["njnt"]):
        d, t = M["jnt_dofadr"][j], M["jnt_type"][j]
        if t == JNT_HINGE:
            dof_iw[d] = dinv[d]
        elif t == JNT_BALL:
            dof_iw[d:d + 3] = dinv[d:d + 3].mean()
        else:
            dof_iw[d:d + 3] = dinv[d:d + 3].mean()
            dof_iw[d + 3:d + 6] = dinv[d + 3:d + 6].mean()
    M["dof_invweight0"] = dof_iw
    biw = np.zeros((nb, 2))
    for b in range(1, nb):
        if M["body_mass"][b] <= 0 and M["body_dofnum"][b] == 0:
            continue
        A = J[b] @ Hinv @ J[b].T
        biw[b, 0] = np.trace(A[3:, 3:]) / 3.0  # translational
        biw[b, 1] = np.trace(A[:3, :3]) / 3.0  # rotational
    M["body_invweight0"] = biw
    # springdamper (haltere class, fruitfly.xml:90): stiffness/damping from the joint's effective inertia
    stiff, damp = M["jnt_stiffness"].copy(), M["dof_damping"].copy()
    names = M["_names_jnt"]
    spec_j = {j.name: j for j in [jj for b in ts.world.walk() for jj in b.joints]}
    for j, nm in enumerate(names):
        sd = spec_j[nm].springdamper
        if sd is not None and sd[0] > 0 and sd[1] > 0:
            d = M["jnt_dofadr"][j]
            inertia = 1.0 / max(MJ_MINVAL, dof_iw[d])
            stiff[j] = inertia / max(MJ_MINVAL, sd[0] ** 2 * sd[1] ** 2)
            damp[d] = 2.0 * inertia / max(MJ_MINVAL, sd[0])
    M["jnt_stiffness"], M["dof_damping"] = stiff, damp
    return M


# ----------------------------------------------------------------------------
_TASK_TYPE = {"walk_on_ball": 0, "walk_imitation": 1, "flight_imitation": 2, "template": 3, "vision_flight": 4}
_APPENDAGES = ("claw_T1_left", "claw_T1_right", "claw_T2_left", "claw_T2_right", "claw_T3_left",
               "claw_T3_right", "head")  # fruitfly.py:422-446
_LEG_HINGES_PER_LEG = 11


def task_tables(M: dict, ts: TaskSpec) -> dict:
    """Index maps and constants of the task layer (action scatter, observation layout, reward/termination
    thresholds).  Cites: fruitfly.py:283-318,480-526; tasks/base.py:154-168,237-261; walk_imitation.py:60-84;
    walk_on_ball.py:47; flight_imitation.py:44-85; tasks/constants.py:16-27."""
    from .model_fields import OBS_KEYS
    w = ts.walker
    jn, jq, jd = M["_names_jnt"], M["jnt_qposadr"], M["jnt_dofadr"]
    jid = {n: i for i, n in enumerate(jn)}
    sid = {n: i for i, n in enumerate(M["_names_site"])}
    M["task_type"] = _TASK_TYPE[ts.task]
    M["n_substeps"], M["future_steps"] = ts.n_substeps, ts.future_steps
    M["control_dt"], M["time_limit"] = ts.control_dt, ts.time_limit
    M["max_episode_steps"] = int(round(ts.time_limit / ts.control_dt)) + (0 if ts.flight else 1)  # walk_imitation.py:60 ; flight_imitation.py:107
    # composer.Environment ends the episode when physics.time() >= time_limit; time is the float64 running sum of
    # the physics timestep, so the first control step at which that holds is found by the same accumulation.
    t, k = 0.0, 0
    while t < ts.time_limit:
        for _ in range(ts.n_substeps):
            t += ts.physics_dt
        k += 1
    M["time_limit_steps"] = k
    M["action_dim"] = w.action_dim
    ca = -np.ones(M["nu"], np.int32)
    for k, idx in w.action_indices.items():
        if w.ctrl_indices.get(k) and idx:
            for c, a in zip(w.ctrl_indices[k], idx):
                ca[c] = a
    M["ctrl_action"] = ca
    amin, amax = -np.ones(w.action_dim), np.ones(w.action_dim)
    for c in range(M["nu"]):
        if ca[c] >= 0:
            amin[ca[c]], amax[ca[c]] = M["act_ctrlrange"][c]
    M["action_min"], M["action_max"] = amin, amax
    M["user_action"] = w.action_indices["user"][0] if w.action_indices["user"] else -1
    root = M["_owner_of"]["walker/thorax"][0]
    M["root_body"] = root
    if "walker/" in jid:
        M["root_qposadr"], M["root_dofadr"] = int(jq[jid["walker/"]]), int(jd[jid["walker/"]])
    else:
        M["root_qposadr"] = M["root_dofadr"] = -1
    if "ghost/" in jid:
        M["ghost_qposadr"], M["ghost_dofadr"] = int(jq[jid["ghost/"]]), int(jd[jid["ghost/"]])
    else:
        M["ghost_qposadr"] = M["ghost_dofadr"] = -1
    M["ball_dofadr"] = int(jd[jid["ball"]]) if "ball" in jid else -1
    M["thorax_site"] = sid["walker/thorax"]
    oj = ["walker/" + n for n in w.observable_joints]
    M["n_obs_joints"] = len(oj)
    M["obs_joint_qposadr"] = np.array([jq[jid[n]] for n in oj], np.int32)
    M["obs_joint_dofadr"] = np.array([jd[jid[n]] for n in oj], np.int32)
    walking = not ts.flight
    app = [sid["walker/" + n] for n in _APPENDAGES] if walking else []
    M["n_appendages"], M["appendage_site"] = len(app), np.array(app, np.int32)
    # mocap joints/sites: the walking dataset's id2name lists are not shipped
    # (config/train_config_fly.yaml:12); default = the 66 leg hinges in model order, 6 claw sites.
    if ts.task == "walk_imitation":
        mj = [i for i, n in enumerate(jn) if n.startswith("walker/") and _any_in(_LEG, n)]
        ms = [sid["walker/" + n] for n in _APPENDAGES[:6]]
    else:
        mj, ms = [], []
    M["n_mocap_joints"], M["mocap_joint"] = len(mj), np.array(mj, np.int32)
    M["n_mocap_sites"], M["mocap_site"] = len(ms), np.array(ms, np.int32)
    wj = [f"walker/wing_{ax}_{side}" for side in ("left", "right") for ax in ("yaw", "roll", "pitch")]
    M["wing_joint"] = np.array([jid[n] for n in wj], np.int32)  # flight_imitation.py / base.py:315-320 order
    M["wing_action"] = np.array(w.action_indices["wings"] if w.action_indices["wings"] else [-1] * 6, np.int32)
    # reset pose: qpos0, wings at springref (fruitfly.py:335-340) == retract_wings values (task_utils.py:117-122)
    qr = M["qpos0"].copy()
    if walking:
        for n in wj:
            qr[jq[jid[n]]] = M["qpos_spring"][jq[jid[n]]]
    M["qpos_reset"] = qr
    # observation layout (alphabetical keys)
    F = ts.future_steps + 1
    dims = {"accelerometer": 3, "gyro": 3, "velocimeter": 3, "world_zaxis": 3, "joints_pos": len(oj),
            "joints_vel": len(oj), "actuator_activation": M["na"]}
    if walking:
        dims.update(appendages_pos=21, force=18, touch=6)
    if ts.task == "walk_on_ball":
        dims["ball_qvel"] = 3
    if ts.task in ("walk_imitation", "flight_imitation"):
        dims.update(ref_displacement=3 * F, ref_root_quat=4 * F)
    if ts.vision:   # tasks/vision_flight.py:243-252; the eye images are a separate uint8 output (flyb_render)
        dims["task_input"] = 2
    off, n = -np.ones(16, np.int32), 0
    for i, k in enumerate(OBS_KEYS):
        if k in dims:
            off[i] = n
            n += dims[k]
    M["obs_offset"], M["obs_dim"] = off, n
    M["_obs_dims"] = dims
    M["terminal_linvel"], M["terminal_angvel"] = 50.0, 200.0  # tasks/constants.py:18-19
    M["terminal_qacc"], M["terminal_height"] = 1e14, 0.2      # constants.py:25,27
    M["terminal_com_dist"] = {"walk_imitation": 0.3, "flight_imitation": 2.0}.get(ts.task, 0.0)  # fly_envs.py:33,78
    M["ghost_offset"] = np.zeros(3)
    M["pad_first_obs"], M["inference_mode"] = 0, 0
    # vision-guided flight (tasks/vision_flight.py:18-63, fly_envs.py:160-206): terrain geom, eye sites, episode draws
    gn = M["_names_geom"]
    M["hfield_geom"] = gn.index("terrain") if ts.vision else -1
    M["hfield_nrow"] = ts.hfield["nrow"] if ts.vision else 0
    M["hfield_ncol"] = ts.hfield["ncol"] if ts.vision else 0
    M["hfield_size"] = ts.hfield["size"] if ts.vision else np.zeros(4)
    M["eye_site"] = np.array([sid["walker/eye_right"], sid["walker/eye_left"]] if ts.vision else [-1, -1], np.int32)
    up = None
    if ts.flight:
        site, _ = w.spec.find_site("walker/hover_up_dir")
        up = site.quat
    M["vision_init_quat"] = _neg_quat(up) if ts.vision else np.array([1.0, 0, 0, 0])   # vision_flight.py:131: neg_quat(up_dir)
    th = np.deg2rad(47.5)                                                              # vision_flight.py:111-112, body pitch
    M["vision_target_zaxis"] = np.array([np.sin(th), 0.0, np.cos(th)])
    # target_height_range, target_speed_range, init_pos_x_range, init_pos_y_range (vision_flight.py:28-31)
    M["vision_ranges"] = np.array([0.5, 0.8, 20.0, 40.0, -5.0, -5.0, 0.0, 0.0])
    # solver options: MuJoCo defaults [3P] + fruitfly.xml:4
    M["opt_iterations"], M["opt_ls_iterations"] = 100, 50
    M["opt_noslip_iterations"] = int(ts.option.get("noslip_iterations", 0))
    M["opt_tolerance"], M["opt_ls_tolerance"], M["opt_noslip_tolerance"], M["opt_impratio"] = 1e-8, 0.01, 1e-6, 1.0
    return M


def compile_task(task: str, xml_path: str) -> dict:
    ts = TaskSpec(task, xml_path)
    M = flatten(ts)
    M = finalize(M, ts)
    M = constants(M, ts)
    M = task_tables(M, ts)
    M.pop("_owner_of")
    return M


def save_model(M: dict, path: str):
    import json
    arrs = {k: np.asarray(v) for k, v in M.items() if not k.startswith("_")}
    meta = {k: v for k, v in M.items() if k.startswith("_")}
    np.savez_compressed(path, _meta=np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8), **arrs)


def load_model(path: str) -> dict:
    import json
    z = np.load(path)
    M = {}
    for k in z.files:
        if k == "_meta":
            M.update(json.loads(bytes(z[k]).decode()))
        else:
            a = z[k]
            M[k] = a.item() if a.ndim == 0 else a
    return M
