"""ctypes wrapper of the CPU emulation of the engine source (tests/emu/flyb_emu.cpp).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libflyb_emu.so")
_CSRC = os.path.join(_HERE, "..", "..", "flybody_b200", "csrc")


def build(force: bool = False):
    srcs = [os.path.join(_HERE, "flyb_emu.cpp")] + [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".h"))]
    newest = max(os.path.getmtime(s) for s in srcs)
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < newest:
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-fPIC", "-shared", "-std=c++17", "-o", _LIB, srcs[0]])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        from flybody_b200.model_pack import FlybModel
        L = ctypes.CDLL(build())
        L.emu_create.restype = ctypes.c_void_p
        L.emu_create.argtypes = [ctypes.POINTER(FlybModel), ctypes.c_int, ctypes.c_int, ctypes.c_char_p, ctypes.c_int]
        L.emu_destroy.argtypes = [ctypes.c_void_p]
        fp, ip, dp = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_double)
        L.emu_set_clips.argtypes = [ctypes.c_void_p, fp, ip, ip, ctypes.c_int]
        L.emu_set_wbpg.argtypes = [ctypes.c_void_p, dp, ip, dp, dp, ctypes.c_int, dp]
        L.emu_set_phase.argtypes = [ctypes.c_void_p, ctypes.c_double]
        L.emu_wbpg_state.argtypes = [ctypes.c_void_p, ip, ip, dp]
        L.emu_set_terrain.argtypes = [ctypes.c_void_p, fp, ctypes.c_int, dp, fp, ctypes.c_int]
        L.emu_set_seed.argtypes = [ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int]
        L.emu_vision_draws.argtypes = [ctypes.c_void_p, ctypes.c_int, ip, dp, dp]
        L.emu_eyes.argtypes = [ctypes.c_void_p, fp, ip]
        L.emu_env_reset.argtypes = [ctypes.c_void_p, fp]
        L.emu_env_step.argtypes = [ctypes.c_void_p, fp, fp, dp, dp]
        for fn in ("emu_forward", "emu_physics_step"):
            getattr(L, fn).argtypes = [ctypes.c_void_p]
        L.emu_call.argtypes = [ctypes.c_void_p, ctypes.c_char_p]
        L.emu_get.argtypes = [ctypes.c_void_p, ctypes.c_char_p, dp, ctypes.c_int]
        L.emu_set.argtypes = [ctypes.c_void_p, ctypes.c_char_p, dp, ctypes.c_int]
        L.emu_geti.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ip, ctypes.c_int]
        L.emu_quat_op.argtypes = [ctypes.c_int, ctypes.c_int, dp, dp, dp]
        L.emu_hfield_contacts.argtypes = [fp, ctypes.c_int, ctypes.c_int, dp, dp, ctypes.c_int, ip, dp, ctypes.c_int, dp]
        _lib = L
    return _lib


class EmuEnv:
    def __init__(self, model: dict, flags: int = 0, double: bool = True):
        from flybody_b200.model_pack import pack_model
        self.M = model
        self._struct, self._keep = pack_model(model)
        err = ctypes.create_string_buffer(256)
        self._h = lib().emu_create(ctypes.byref(self._struct), flags, int(double), err, 256)
        if not self._h:
            raise RuntimeError(err.value.decode())

    def __del__(self):
        if getattr(self, "_h", None):
            lib().emu_destroy(self._h)
            self._h = None

    def get(self, name, cap=200000):
        out = np.zeros(cap)
        n = lib().emu_get(self._h, name.encode(), out.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), cap)
        if n < 0:
            raise KeyError(name)
        return out[:n].copy()

    def set(self, name, v):
        v = np.ascontiguousarray(v, dtype=np.float64)
        n = lib().emu_set(self._h, name.encode(), v.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), v.size)
        if n < 0:
            raise KeyError(name)

    def geti(self, name, cap=4096):
        out = np.zeros(cap, np.int32)
        n = lib().emu_geti(self._h, name.encode(), out.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), cap)
        if n < 0:
            raise KeyError(name)
        return out[:n].copy()

    def call(self, stage):
        lib().emu_call(self._h, stage.encode())

    def forward(self):
        lib().emu_forward(self._h)

    def physics_step(self, n=1):
        for _ in range(n):
            lib().emu_physics_step(self._h)

    def dense_m(self, name="qM"):
        nv = int(self.M["nv"])
        v, r, c = self.get(name), self.geti("mrow"), self.geti("mcol")
        A = np.zeros((nv, nv))
        A[r, c] = v[:len(r)]
        return A

    def set_clips(self, table, offsets, lengths):
        t = np.ascontiguousarray(table, np.float32)
        o = np.ascontiguousarray(offsets, np.int32)
        ln = np.ascontiguousarray(lengths, np.int32)
        self._clips = (t, o, ln)
        fp, ip = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int)
        lib().emu_set_clips(self._h, t.ctypes.data_as(fp), o.ctypes.data_as(ip), ln.ctypes.data_as(ip), len(o))

    def set_wbpg(self, tables):
        dp, ip = ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_int)
        a = [np.ascontiguousarray(tables["beat_freqs"], np.float64), np.ascontiguousarray(tables["offsets"], np.int32),
             np.ascontiguousarray(tables["traj"], np.float64), np.ascontiguousarray(tables["phase"], np.float64),
             np.array([tables["base_freq"], tables["rel_freq_range"], tables["rate"], tables["dt_ctrl"]], np.float64)]
        lib().emu_set_wbpg(self._h, a[0].ctypes.data_as(dp), a[1].ctypes.data_as(ip), a[2].ctypes.data_as(dp),
                           a[3].ctypes.data_as(dp), len(a[0]), a[4].ctypes.data_as(dp))

    def set_initial_phase(self, phase):
        lib().emu_set_phase(self._h, float(phase))

    def set_terrain(self, packed):
        """packed: flybody_b200.terrain.pack_pool(...)"""
        fp, dp = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_double)
        self._terrain = packed
        h, meta, ty = packed["height"], packed["trench_meta"], packed["trench_y"]
        lib().emu_set_terrain(self._h, h.ctypes.data_as(fp), h.shape[0], meta.ctypes.data_as(dp) if meta is not None else None,
                              ty.ctypes.data_as(fp) if ty is not None else None, packed["trench_stride"])

    def set_seed(self, seed, env_offset=0):
        lib().emu_set_seed(self._h, int(seed), int(env_offset))

    def vision_draws(self, episode):
        t, h, v = ctypes.c_int(0), ctypes.c_double(0), ctypes.c_double(0)
        lib().emu_vision_draws(self._h, int(episode), ctypes.byref(t), ctypes.byref(h), ctypes.byref(v))
        return t.value, h.value, v.value

    def eyes(self):
        out, t = np.zeros(24, np.float32), ctypes.c_int(0)
        lib().emu_eyes(self._h, out.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), ctypes.byref(t))
        return out.reshape(2, 12), t.value

    def wbpg_state(self):
        a, b, c = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_double(0)
        lib().emu_wbpg_state(self._h, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
        return a.value, b.value, c.value

    def env_reset(self):
        obs = np.zeros(int(self.M["obs_dim"]), np.float32)
        st = lib().emu_env_reset(self._h, obs.ctypes.data_as(ctypes.POINTER(ctypes.c_float)))
        return st, obs

    def env_step(self, action):
        a = np.ascontiguousarray(action, np.float32)
        obs = np.zeros(int(self.M["obs_dim"]), np.float32)
        r, g = ctypes.c_double(0), ctypes.c_double(0)
        fp = ctypes.POINTER(ctypes.c_float)
        st = lib().emu_env_step(self._h, a.ctypes.data_as(fp), obs.ctypes.data_as(fp), ctypes.byref(r), ctypes.byref(g))
        return st, r.value, g.value, obs


def quat_op(op: int, a, b=None, double: bool = False):
    """The engine's task-layer quaternion helpers (csrc/fly_env.cuh) on raw inputs, in float32 or float64 (ops: oracle.quat_op)."""
    dp = ctypes.POINTER(ctypes.c_double)
    aa = np.zeros(6); aa[:np.size(a)] = np.asarray(a, np.float64).reshape(-1)
    bb = np.zeros(4)
    if b is not None:
        bb[:np.size(b)] = np.asarray(b, np.float64).reshape(-1)
    out = np.zeros(4)
    n = lib().emu_quat_op(int(double), int(op), aa.ctypes.data_as(dp), bb.ctypes.data_as(dp), out.ctypes.data_as(dp))
    return out[:n].copy()


def hfield_contacts(data, pose12, size4, types, geoms16, maxcon: int = 16):
    """The engine's height-field narrow phase (csrc/fly_hfield.cuh) compiled for the CPU; arguments and result as
    flybody_b200.engine.hfield_contacts."""
    dp = ctypes.POINTER(ctypes.c_double)
    d = np.ascontiguousarray(data, np.float32)
    pose, size = np.ascontiguousarray(pose12, np.float64).reshape(12), np.ascontiguousarray(size4, np.float64).reshape(4)
    t = np.ascontiguousarray(types, np.int32)
    g = np.ascontiguousarray(geoms16, np.float64).reshape(len(t), 16)
    out = np.zeros((len(t), 1 + 7 * maxcon))
    lib().emu_hfield_contacts(d.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), d.shape[0], d.shape[1], pose.ctypes.data_as(dp),
                              size.ctypes.data_as(dp), len(t), t.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), g.ctypes.data_as(dp),
                              int(maxcon), out.ctypes.data_as(dp))
    return [[(o[1 + 7 * i], o[2 + 7 * i:5 + 7 * i].copy(), o[5 + 7 * i:8 + 7 * i].copy()) for i in range(int(o[0]))] for o in out]
