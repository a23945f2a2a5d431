"""ctypes wrapper of the float64 CPU oracle (oracle/flyb_oracle.c).  TEST INFRASTRUCTURE ONLY:
imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs, never by flybody_b200/."""
from __future__ import annotations

import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libflyb_oracle.so")


def build(force: bool = False):
    src = os.path.join(_HERE, "flyb_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    return _LIB


class Contact(ctypes.Structure):
    _fields_ = [("pair", ctypes.c_int), ("geom1", ctypes.c_int), ("geom2", ctypes.c_int), ("dim", ctypes.c_int),
                ("exclude", ctypes.c_int), ("efc_address", ctypes.c_int), ("dist", ctypes.c_double),
                ("includemargin", ctypes.c_double), ("mu", ctypes.c_double), ("pos", ctypes.c_double * 3),
                ("frame", ctypes.c_double * 9), ("friction", ctypes.c_double * 3), ("solref", ctypes.c_double * 2),
                ("solimp", ctypes.c_double * 5)]


class Clip(ctypes.Structure):
    _fields_ = [("T", ctypes.c_int), ("qpos", ctypes.POINTER(ctypes.c_double)), ("qvel", ctypes.POINTER(ctypes.c_double)),
                ("root2site", ctypes.POINTER(ctypes.c_double)), ("joint_quat", ctypes.POINTER(ctypes.c_double))]


class Wbpg(ctypes.Structure):
    _fields_ = [("nfreq", ctypes.c_int), ("beat_freqs", ctypes.POINTER(ctypes.c_double)), ("offsets", ctypes.POINTER(ctypes.c_int)),
                ("traj", ctypes.POINTER(ctypes.c_double)), ("phase", ctypes.POINTER(ctypes.c_double)),
                ("base_freq", ctypes.c_double), ("rel_range", ctypes.c_double), ("rate", ctypes.c_double),
                ("dt_ctrl", ctypes.c_double)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        from flybody_b200.model_pack import FlybModel
        L = ctypes.CDLL(build())
        L.flyo_create.restype = ctypes.c_void_p
        L.flyo_create.argtypes = [ctypes.POINTER(FlybModel)]
        for fn in ("flyo_destroy", "flyo_reset_data", "flyo_forward", "flyo_step"):
            getattr(L, fn).argtypes = [ctypes.c_void_p]
            getattr(L, fn).restype = None
        L.flyo_set_option.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_double]
        L.flyo_dptr.restype = ctypes.POINTER(ctypes.c_double)
        L.flyo_dptr.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int)]
        L.flyo_iptr.restype = ctypes.POINTER(ctypes.c_int)
        L.flyo_iptr.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_int)]
        L.flyo_contacts.restype = ctypes.POINTER(Contact)
        L.flyo_contacts.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)]
        L.flyo_set_clip.argtypes = [ctypes.c_void_p, ctypes.POINTER(Clip)]
        dp = ctypes.POINTER(ctypes.c_double)
        L.flyo_set_wbpg.argtypes = [ctypes.c_void_p, ctypes.POINTER(Wbpg)]
        L.flyo_set_initial_phase.argtypes = [ctypes.c_void_p, ctypes.c_double]
        L.flyo_wbpg_state.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_int), dp]
        L.flyo_env_reset.argtypes = [ctypes.c_void_p, dp]
        L.flyo_env_reset.restype = ctypes.c_int
        L.flyo_env_step.argtypes = [ctypes.c_void_p, dp, dp, dp, dp]
        L.flyo_env_step.restype = ctypes.c_int
        L.flyo_narrow_pair.argtypes = [ctypes.c_int, dp, dp, dp, ctypes.c_int, dp, dp, dp, ctypes.c_double, ctypes.c_int, dp]
        L.flyo_narrow_pair.restype = ctypes.c_int
        L.flyo_narrow_hfield.argtypes = [dp, dp, dp, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_float), ctypes.c_int, dp, dp, dp,
                                         ctypes.c_double, ctypes.c_int, dp]
        L.flyo_narrow_hfield.restype = ctypes.c_int
        L.flyo_set_terrain.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_float), ctypes.c_int, dp, dp]
        L.flyo_set_vision_targets.argtypes = [ctypes.c_void_p, ctypes.c_double, ctypes.c_double]
        L.flyo_quat_op.argtypes = [ctypes.c_int, dp, dp, dp]
        L.flyo_quat_op.restype = ctypes.c_int
        _lib = L
    return _lib


class OracleEnv:
    """One float64 environment.  `d("qpos")` returns a live numpy view into the C arrays."""

    def __init__(self, model: dict):
        from flybody_b200.model_pack import pack_model
        self.M = model
        self._struct, self._keep = pack_model(model)
        self._h = lib().flyo_create(ctypes.byref(self._struct))
        self._clip = None

    def __del__(self):
        if getattr(self, "_h", None):
            lib().flyo_destroy(self._h)
            self._h = None

    def d(self, name: str) -> np.ndarray:
        n = ctypes.c_int(0)
        p = lib().flyo_dptr(self._h, name.encode(), ctypes.byref(n))
        if not p:
            raise KeyError(name)
        return np.ctypeslib.as_array(p, shape=(n.value,))

    def i(self, name: str) -> np.ndarray:
        n = ctypes.c_int(0)
        p = lib().flyo_iptr(self._h, name.encode(), ctypes.byref(n))
        if not p:
            raise KeyError(name)
        return np.ctypeslib.as_array(p, shape=(max(n.value, 0),))

    def option(self, name: str, value: float):
        lib().flyo_set_option(self._h, name.encode(), float(value))

    def reset_data(self):
        lib().flyo_reset_data(self._h)

    def forward(self):
        lib().flyo_forward(self._h)

    def step(self, n: int = 1):
        for _ in range(n):
            lib().flyo_step(self._h)

    def contacts(self):
        n = ctypes.c_int(0)
        p = lib().flyo_contacts(self._h, ctypes.byref(n))
        out = []
        for k in range(n.value):
            c = p[k]
            out.append(dict(pair=c.pair, geom1=c.geom1, geom2=c.geom2, dim=c.dim, exclude=c.exclude,
                            efc_address=c.efc_address, dist=c.dist, pos=np.array(c.pos), frame=np.array(c.frame),
                            mu=c.mu, friction=np.array(c.friction)))
        return out

    # ---- task level
    def set_clip(self, qpos, qvel, root2site=None, joint_quat=None):
        arrs = [np.ascontiguousarray(a, dtype=np.float64) if a is not None else None
                for a in (qpos, qvel, root2site, joint_quat)]
        dp = ctypes.POINTER(ctypes.c_double)
        c = Clip(T=arrs[0].shape[0], qpos=arrs[0].ctypes.data_as(dp), qvel=arrs[1].ctypes.data_as(dp),
                 root2site=arrs[2].ctypes.data_as(dp) if arrs[2] is not None else None,
                 joint_quat=arrs[3].ctypes.data_as(dp) if arrs[3] is not None else None)
        self._clip = (c, arrs)
        lib().flyo_set_clip(self._h, ctypes.byref(c))

    def set_wbpg(self, tables):
        """tables: dict from flybody_b200.wbpg.build_tables"""
        dp, ip = ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int)
        arrs = [np.ascontiguousarray(tables["beat_freqs"], np.float64), np.ascontiguousarray(tables["offsets"], np.int32),
                np.ascontiguousarray(tables["traj"], np.float64), np.ascontiguousarray(tables["phase"], np.float64)]
        wb = Wbpg(nfreq=len(arrs[0]), beat_freqs=arrs[0].ctypes.data_as(dp), offsets=arrs[1].ctypes.data_as(ip),
                  traj=arrs[2].ctypes.data_as(dp), phase=arrs[3].ctypes.data_as(dp), base_freq=tables["base_freq"],
                  rel_range=tables["rel_freq_range"], rate=tables["rate"], dt_ctrl=tables["dt_ctrl"])
        self._wbpg = (wb, arrs)
        lib().flyo_set_wbpg(self._h, ctypes.byref(wb))

    def set_initial_phase(self, phase):
        lib().flyo_set_initial_phase(self._h, float(phase))

    def set_terrain(self, heights, trench_x=None, trench_y=None):
        """vision_flight: the episode's height field [nrow, ncol] and, for a trench terrain, its centre line."""
        dp = ctypes.POINTER(ctypes.c_double)
        h = np.ascontiguousarray(heights, np.float32)
        tx = np.ascontiguousarray(trench_x if trench_x is not None else [], np.float64)
        ty = np.ascontiguousarray(trench_y if trench_y is not None else [], np.float64)
        self._terrain = (h, tx, ty)
        lib().flyo_set_terrain(self._h, h.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), len(tx), tx.ctypes.data_as(dp), ty.ctypes.data_as(dp))

    def set_vision_targets(self, height, speed):
        lib().flyo_set_vision_targets(self._h, float(height), float(speed))

    def wbpg_state(self):
        a, b, c = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_double(0)
        lib().flyo_wbpg_state(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
        return a.value, b.value, c.value

    def env_reset(self):
        obs = np.zeros(self.M["obs_dim"])
        st = lib().flyo_env_reset(self._h, obs.ctypes.data_as(ctypes.POINTER(ctypes.c_double)))
        return st, obs

    def env_step(self, action):
        dp = ctypes.POINTER(ctypes.c_double)
        a = np.ascontiguousarray(action, dtype=np.float64)
        obs = np.zeros(self.M["obs_dim"])
        r, g = ctypes.c_double(0), ctypes.c_double(0)
        st = lib().flyo_env_step(self._h, a.ctypes.data_as(dp), obs.ctypes.data_as(dp), ctypes.byref(r), ctypes.byref(g))
        return st, r.value, g.value, obs


GEOM_PLANE, GEOM_SPHERE, GEOM_CAPSULE, GEOM_ELLIPSOID, GEOM_CYLINDER = 0, 2, 3, 4, 5


def narrow_pair(t1, pos1, mat1, size1, t2, pos2, mat2, size2, margin=0.0, force_convex=False):
    """Narrow phase of one primitive pair (geom type, world position, 3x3 world orientation, size) in the oracle.
    Returns a list of (dist, pos[3], normal[3]); force_convex sends the pair through the MPR routine whatever its types."""
    a = [np.ascontiguousarray(x, np.float64).reshape(-1) for x in (pos1, mat1, size1, pos2, mat2, size2)]
    out = np.zeros(28)
    dp = ctypes.POINTER(ctypes.c_double)
    P = [x.ctypes.data_as(dp) for x in a]
    n = lib().flyo_narrow_pair(int(t1), P[0], P[1], P[2], int(t2), P[3], P[4], P[5], float(margin), int(force_convex),
                               out.ctypes.data_as(dp))
    return [(out[7 * i], out[7 * i + 1:7 * i + 4].copy(), out[7 * i + 4:7 * i + 7].copy()) for i in range(n)]


def quat_op(op: int, a, b=None):
    """The oracle's task-layer quaternion helpers on raw inputs (flyo_quat_op: 0 reciprocal, 1 rotate(v, q), 2 z2vec, 3 short-arc
    distance, 4 joint_orientation_quat(axis, angle), 5 get_dquat_local, 6 get_egocentric_vec([root, site], quat), 7 mult)."""
    dp = ctypes.POINTER(ctypes.c_double)
    a = np.ascontiguousarray(a, np.float64).reshape(-1)
    b = np.ascontiguousarray(b if b is not None else [0.0], np.float64).reshape(-1)
    out = np.zeros(4)
    n = lib().flyo_quat_op(int(op), a.ctypes.data_as(dp), b.ctypes.data_as(dp), out.ctypes.data_as(dp))
    return out[:n].copy()


def narrow_hfield(hpos, hmat, hsize, data, t2, pos2, mat2, size2, margin=0.0, maxcon=16):
    """Height field (geom pose hpos / hmat, hsize = [x half extent, y half extent, elevation scale, base], data [nrow, ncol]
    float32) against one convex geom in the oracle: list of (dist, pos[3], normal[3])."""
    dp = ctypes.POINTER(ctypes.c_double)
    a = [np.ascontiguousarray(x, np.float64).reshape(-1) for x in (hpos, hmat, hsize, pos2, mat2, size2)]
    d = np.ascontiguousarray(data, np.float32)
    out = np.zeros(7 * maxcon)
    n = lib().flyo_narrow_hfield(a[0].ctypes.data_as(dp), a[1].ctypes.data_as(dp), a[2].ctypes.data_as(dp), d.shape[0], d.shape[1],
                                 d.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), int(t2), a[3].ctypes.data_as(dp),
                                 a[4].ctypes.data_as(dp), a[5].ctypes.data_as(dp), float(margin), int(maxcon), out.ctypes.data_as(dp))
    return [(out[7 * i], out[7 * i + 1:7 * i + 4].copy(), out[7 * i + 4:7 * i + 7].copy()) for i in range(n)]
