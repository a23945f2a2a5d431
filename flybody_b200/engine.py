"""ctypes binding of libflyb.so (include/flyb.h): the only way the Python host layer reaches the GPU.

There is no CPU fallback.  If the library has not been built (`python -c "import __graft_entry__ as g; g.build()"`)
or no CUDA device is present, constructing an engine raises."""
from __future__ import annotations

import ctypes
import os
import subprocess
import numpy as np

from .model_pack import FlybModel, pack_model

_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FLYB_LIB_PATH", os.path.join(_DIR, "libflyb.so"))  # the override is for kernel experiments (tools/)
CSRC = os.path.join(_DIR, "csrc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--expt-extended-lambda",
              "-Xcompiler", "-fPIC", "-shared"]

FLAG_NOCONTACT, FLAG_NOLIMIT, FLAG_NOFLUID, FLAG_NONOSLIP, FLAG_NOACTUATION, FLAG_NOWARMSTART = 1, 2, 4, 8, 16, 32
FLAG_TRACE = 1 << 16
FLAG_LANES16 = 1 << 17
(INFO_NUM_ENVS, INFO_OBS_DIM, INFO_ACTION_DIM, INFO_NQ, INFO_NV, INFO_NA, INFO_TRACE_INTS, INFO_TRACE_REALS,
 INFO_LAUNCHES, INFO_SMEM_BYTES, INFO_BULK_BYTES, INFO_GRID, INFO_MAXCON, INFO_MAXEFC, INFO_STATE_INTS,
 INFO_WARPS_PER_BLOCK, INFO_SMEM_LEVEL, INFO_SPEC_KERNEL, INFO_LANES) = range(19)

EXPORTS = ["flyb_create", "flyb_destroy", "flyb_last_error", "flyb_info", "flyb_configure", "flyb_load_clips",
           "flyb_set_clip_ids", "flyb_reset", "flyb_step", "flyb_reset_host", "flyb_step_host", "flyb_get_state",
           "flyb_set_state", "flyb_get_trace", "flyb_episode_stats", "flyb_load_wbpg", "flyb_set_wbpg_phase",
           "flyb_use_generic_kernel", "flyb_step_noalloc", "flyb_fma_peak", "flyb_hfield_contacts", "flyb_render_eyes",
           "flyb_load_terrain", "flyb_render",
           "flyb_policy_create", "flyb_policy_destroy", "flyb_policy_last_error", "flyb_policy_launches",
           "flyb_policy_set_weights", "flyb_policy_forward", "flyb_adder_step"]


N_SPEC_MODELS = 5   # csrc/fly_dims_gen.h: FLY_SPEC_COUNT
# Step-kernel instantiations built into the library: (lanes per env, warps per block, shared-memory level, model index or -1 = any
# model).  Each is a translation unit of its own (csrc/flyb_kern.cu) that registers its kernel when the library is loaded.
KERNEL_UNITS = ([(32, 28, 0, i) for i in range(N_SPEC_MODELS)]             # the default: one per shipped task model
                + [(32, 28, 0, -1), (32, 16, 3, -1), (32, 8, 4, -1)]          # any model; other launch configurations
                + [(16, 16, 0, -1)])                                          # two envs per warp (FLAG_LANES16)


def kernel_units():
    """KERNEL_UNITS, or the subset / extra units named in FLYB_BUILD_UNITS ("lanes:warps:level:spec,...") for kernel experiments."""
    e = os.environ.get("FLYB_BUILD_UNITS")
    if not e:
        return list(KERNEL_UNITS)
    return [tuple(int(x) for x in u.split(":")) for u in e.split(",") if u]


def compile_library(out_path: str, extra_flags=(), verbose: bool = False, units=None) -> str:
    """nvcc (sm_100a, -lineinfo): flyb_cuda.cu plus one translation unit per step-kernel instantiation (flyb_kern.cu), the
    compiles running in parallel, linked into `out_path`."""
    import tempfile
    from concurrent.futures import ThreadPoolExecutor
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + list(extra_flags) + (["-Xptxas", "-v"] if verbose else [])
    # flyb_cuda.cu also holds the height-field and eye-camera kernels.  The portal refinement against a prism takes support points
    # among the vertices of a face whenever its direction is that face's normal: ties decided by the last bit.  Without
    # multiply-add contraction the float64 device code rounds like the CPU code and picks the same vertex.
    tus = [("flyb_cuda", "flyb_cuda.cu", ["-fmad=false"]), ("flyb_policy", "flyb_policy.cu", []), ("flyb_actor", "flyb_actor.cu", [])]
    for nl, w, l, sp in (units if units is not None else kernel_units()):
        tus.append((f"k_{nl}_{w}_{l}_{'g' if sp < 0 else 's%d' % sp}", "flyb_kern.cu",
                    [f"-DFLY_NL={nl}", f"-DFLY_TU_WARPS={w}", f"-DFLY_TU_LEVEL={l}", f"-DFLY_TU_SPEC={sp}"]))
    with tempfile.TemporaryDirectory() as tmp:
        def one(u):
            # a stub source per unit so that every embedded cubin has a name of its own (cuobjdump -xelf, ncu tooling)
            src, obj = os.path.join(tmp, u[0] + ".cu"), os.path.join(tmp, u[0] + ".o")
            with open(src, "w") as f:
                f.write('#include "%s"\n' % os.path.join(CSRC, u[1]))
            subprocess.check_call(["nvcc"] + flags + u[2] + ["-I", CSRC, "-c", "-o", obj, src])
            return obj
        with ThreadPoolExecutor(max_workers=max(1, min(len(tus), os.cpu_count() or 1))) as ex:
            objs = list(ex.map(one, tus))
        subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out_path] + objs)
    return out_path


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Builds flybody_b200/libflyb.so (in-tree) when a source is newer than it."""
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(_DIR, "..", "include", f) for f in ("flyb.h", "flyb_model.h")]
    srcs.append(os.path.abspath(__file__))   # the build list and the compiler flags live here
    newest = max(os.path.getmtime(s) for s in srcs)
    if not force and os.path.exists(LIB_PATH) and os.path.getmtime(LIB_PATH) >= newest:
        return LIB_PATH
    return compile_library(LIB_PATH, verbose=verbose)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with __graft_entry__.build() (nvcc, sm_100a). "
                               "There is no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        vp, c_i, fp = ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_float)
        ip, u8p, dp = ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_uint8), ctypes.POINTER(ctypes.c_double)
        L.flyb_create.argtypes = [ctypes.POINTER(FlybModel), c_i, c_i, ctypes.c_uint64, c_i, c_i, ctypes.POINTER(vp)]
        L.flyb_destroy.argtypes = [vp]
        L.flyb_destroy.restype = None
        L.flyb_last_error.argtypes = [vp]
        L.flyb_last_error.restype = ctypes.c_char_p
        L.flyb_info.argtypes = [vp, c_i]
        L.flyb_info.restype = ctypes.c_int64
        L.flyb_configure.argtypes = [vp, c_i, c_i, c_i]
        L.flyb_use_generic_kernel.argtypes = [vp, c_i]
        L.flyb_load_clips.argtypes = [vp, fp, ip, ip, c_i]
        L.flyb_set_clip_ids.argtypes = [vp, ip]
        # device-pointer entry points take raw addresses (torch tensor.data_ptr()) and a raw cudaStream_t
        L.flyb_reset.argtypes = [vp, vp, vp, vp, vp]
        L.flyb_step.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.flyb_step_noalloc.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        L.flyb_fma_peak.argtypes = [c_i, dp]
        L.flyb_hfield_contacts.argtypes = [c_i, fp, c_i, c_i, dp, dp, c_i, ip, dp, c_i, dp]
        L.flyb_render_eyes.argtypes = [c_i, fp, c_i, c_i, c_i, ctypes.c_float, ctypes.c_float, c_i, ip, fp, c_i, c_i, ctypes.c_float, fp, u8p]
        L.flyb_reset_host.argtypes = [vp, u8p, fp, ip]
        L.flyb_step_host.argtypes = [vp, fp, fp, fp, fp, ip]
        L.flyb_get_state.argtypes = [vp, fp, fp, fp, fp, ip]
        L.flyb_set_state.argtypes = [vp, fp, fp, fp, fp, ip]
        L.flyb_get_trace.argtypes = [vp, ip, fp]
        L.flyb_episode_stats.argtypes = [vp, dp]
        L.flyb_load_wbpg.argtypes = [vp, dp, ip, dp, dp, c_i, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_double]
        L.flyb_set_wbpg_phase.argtypes = [vp, dp]
        L.flyb_load_terrain.argtypes = [vp, fp, c_i, dp, fp, c_i]
        L.flyb_render.argtypes = [vp, c_i, c_i, ctypes.c_float, fp, vp, vp]
        L.flyb_policy_create.argtypes = [c_i, c_i, c_i, c_i, ctypes.POINTER(vp)]
        L.flyb_policy_destroy.argtypes = [vp]
        L.flyb_policy_destroy.restype = None
        L.flyb_policy_last_error.argtypes = [vp]
        L.flyb_policy_last_error.restype = ctypes.c_char_p
        L.flyb_policy_launches.argtypes = [vp]
        L.flyb_policy_launches.restype = ctypes.c_longlong
        L.flyb_policy_set_weights.argtypes = [vp] + [vp] * 14 + [ctypes.c_float, ctypes.c_float, ctypes.c_float, c_i, vp]
        L.flyb_policy_forward.argtypes = [vp, vp, c_i, vp, vp, vp, vp, vp]
        L.flyb_adder_step.argtypes = [c_i] * 5 + [ctypes.c_float] + [vp] * 21
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float)) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32)) if a is not None else None


def fma_peak_tflops(device: int = 0) -> float:
    """Measured FP32 FMA peak of the device (flyb_fma_peak)."""
    out = ctypes.c_double(0)
    if lib().flyb_fma_peak(int(device), ctypes.byref(out)):
        raise RuntimeError("flyb_fma_peak failed")
    return out.value


def hfield_contacts(data, pose12, size4, types, geoms16, maxcon: int = 16, device: int = 0):
    """flyb_hfield_contacts: height field `data` [nrow, ncol] against the query geoms on the GPU; returns, per query, a list of
    (dist, pos[3], normal[3])."""
    dp = ctypes.POINTER(ctypes.c_double)
    d = np.ascontiguousarray(data, np.float32)
    pose, size = np.ascontiguousarray(pose12, np.float64).reshape(12), np.ascontiguousarray(size4, np.float64).reshape(4)
    t = np.ascontiguousarray(types, np.int32)
    g = np.ascontiguousarray(geoms16, np.float64).reshape(len(t), 16)
    out = np.zeros((len(t), 1 + 7 * maxcon))
    if lib().flyb_hfield_contacts(int(device), _fp(d), d.shape[0], d.shape[1], pose.ctypes.data_as(dp), size.ctypes.data_as(dp), len(t),
                                  _ip(t), g.ctypes.data_as(dp), int(maxcon), out.ctypes.data_as(dp)):
        raise RuntimeError(lib().flyb_last_error(None).decode())
    return [[(o[1 + 7 * i], o[2 + 7 * i:5 + 7 * i].copy(), o[5 + 7 * i:8 + 7 * i].copy()) for i in range(int(o[0]))] for o in out]


# eye cameras of the fly in the frame of its head body (vnl_ray/fruitfly/assets/fruitfly.xml:335-336)
EYE_POS = {"right": (0.0219, 0.0131, 0.0), "left": (-0.0219, 0.0131, 0.0)}
EYE_QUAT = {"right": (0.474, 0.688, -0.344, -0.429), "left": (0.474, 0.688, 0.344, 0.429)}
TERRAIN_RGB, HEADLIGHT = (0.2, 0.3, 0.4), (0.4, 0.8)   # hills.py:207,223-225 (aesthetic "default": flat colour)


def eye_cameras(head_pos, head_mat):
    """World poses [2, 12] (right eye, left eye: position, row-major orientation) from the head body's pose."""
    def q2m(q):
        w, x, y, z = np.asarray(q, np.float64) / np.linalg.norm(q)
        return np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)],
                         [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])
    hp, hm = np.asarray(head_pos, np.float64), np.asarray(head_mat, np.float64).reshape(3, 3)
    return np.stack([np.r_[hp + hm @ np.asarray(EYE_POS[k]), (hm @ q2m(EYE_QUAT[k])).ravel()] for k in ("right", "left")])


def render_eyes(heights, terrain_id, cams12, half_extent=20.0, z_offset=-0.01, size=32, fovy=150.0, rgb=TERRAIN_RGB,
                sky=(0.0, 0.0, 0.0), headlight=HEADLIGHT, device: int = 0):
    """flyb_render_eyes: uint8 images [ncam, size, size, 3] of the terrains `heights` [nterrain, n, n] seen from `cams12`."""
    h = np.ascontiguousarray(heights, np.float32)
    if h.ndim == 2:
        h = h[None]
    cams = np.ascontiguousarray(cams12, np.float32).reshape(-1, 12)
    tid = np.ascontiguousarray(terrain_id, np.int32).reshape(-1)
    colors = np.ascontiguousarray(list(rgb) + list(sky) + list(headlight), np.float32)
    out = np.zeros((len(cams), size, size, 3), np.uint8)
    if lib().flyb_render_eyes(int(device), _fp(h), h.shape[0], h.shape[1], h.shape[2], float(half_extent), float(z_offset), len(cams),
                              _ip(tid), _fp(cams), size, size, float(fovy), _fp(colors),
                              out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))):
        raise RuntimeError(lib().flyb_last_error(None).decode())
    return out


class Engine:
    """One libflyb handle: `num_envs` environments of one compiled task model on one GPU."""

    def __init__(self, model: dict, num_envs: int, device: int = 0, seed: int = 0, env_offset: int = 0, flags: int = 0):
        self.M = model
        self._struct, self._keep = pack_model(model)
        h = ctypes.c_void_p()
        L = lib()
        if L.flyb_create(ctypes.byref(self._struct), int(num_envs), int(device), int(seed) & (2**64 - 1), int(env_offset),
                         int(flags), ctypes.byref(h)):
            raise RuntimeError("flyb_create: " + L.flyb_last_error(None).decode())
        self._h = h
        self.num_envs, self.device, self.flags = int(num_envs), int(device), int(flags)
        self.obs_dim, self.action_dim = int(self.info(INFO_OBS_DIM)), int(self.info(INFO_ACTION_DIM))
        self.nq, self.nv, self.na = (int(self.info(k)) for k in (INFO_NQ, INFO_NV, INFO_NA))

    def close(self):
        if getattr(self, "_h", None):
            lib().flyb_destroy(self._h)
            self._h = None

    __del__ = close

    def _ck(self, rc):
        if rc:
            raise RuntimeError(lib().flyb_last_error(self._h).decode())

    def info(self, what: int) -> int:
        return int(lib().flyb_info(self._h, what))

    def configure(self, smem_level: int, warps_per_block: int, blocks_per_sm: int = 0):
        self._ck(lib().flyb_configure(self._h, smem_level, warps_per_block, blocks_per_sm))

    def use_generic_kernel(self, generic: bool = True):
        self._ck(lib().flyb_use_generic_kernel(self._h, int(generic)))

    def load_clips(self, rows, offsets, lengths):
        rows = np.ascontiguousarray(rows, np.float32)
        offsets = np.ascontiguousarray(offsets, np.int32)
        lengths = np.ascontiguousarray(lengths, np.int32)
        self._ck(lib().flyb_load_clips(self._h, _fp(rows), _ip(offsets), _ip(lengths), len(offsets)))
        self.n_clips = len(offsets)

    def load_wbpg(self, tables):
        """tables: dict from flybody_b200.wbpg.build_tables (flight_imitation only)."""
        dp = ctypes.POINTER(ctypes.c_double)
        f = np.ascontiguousarray(tables["beat_freqs"], np.float64); o = np.ascontiguousarray(tables["offsets"], np.int32)
        t = np.ascontiguousarray(tables["traj"], np.float64); p = np.ascontiguousarray(tables["phase"], np.float64)
        self._ck(lib().flyb_load_wbpg(self._h, f.ctypes.data_as(dp), _ip(o), t.ctypes.data_as(dp), p.ctypes.data_as(dp), len(f),
                                      float(tables["base_freq"]), float(tables["rel_freq_range"]), float(tables["rate"]),
                                      float(tables["dt_ctrl"])))

    def set_wbpg_phase(self, phase):
        """Per-env initial wing-beat phase for the next resets (None = hashed draw)."""
        if phase is None:
            self._ck(lib().flyb_set_wbpg_phase(self._h, None))
        else:
            p = np.ascontiguousarray(phase, np.float64)
            assert p.shape == (self.num_envs,)
            self._ck(lib().flyb_set_wbpg_phase(self._h, p.ctypes.data_as(ctypes.POINTER(ctypes.c_double))))

    def load_terrain(self, packed):
        """packed: flybody_b200.terrain.pack_pool(terrain_table(...)) (vision_flight only)."""
        h, meta, ty = packed["height"], packed["trench_meta"], packed["trench_y"]
        if h.shape[1:] != (int(self.M["hfield_nrow"]), int(self.M["hfield_ncol"])):
            raise ValueError(f"terrain grid {h.shape[1:]} does not match the model's height field")
        self._ck(lib().flyb_load_terrain(self._h, _fp(h), h.shape[0], meta.ctypes.data_as(ctypes.POINTER(ctypes.c_double)) if meta is not None else None,
                                         _fp(ty) if ty is not None else None, int(packed["trench_stride"])))
        self.n_terrains = h.shape[0]

    def render_ptr(self, out_ptr, size=32, fovy=150.0, rgb=TERRAIN_RGB, sky=(0.0, 0.0, 0.0), headlight=HEADLIGHT, stream=0):
        """flyb_render: the eye images of every env after the last reset / step into a device buffer [N, 2, size, size, 3] uint8."""
        colors = np.ascontiguousarray(list(rgb) + list(sky) + list(headlight), np.float32)
        self._ck(lib().flyb_render(self._h, int(size), int(size), float(fovy), _fp(colors), out_ptr, stream))

    def set_clip_ids(self, ids):
        ids = None if ids is None else np.ascontiguousarray(ids, np.int32)
        self._ck(lib().flyb_set_clip_ids(self._h, _ip(ids)))

    # ---- device-pointer calls (raw addresses)
    def reset_ptr(self, mask_ptr, obs_ptr, step_type_ptr, stream=0):
        self._ck(lib().flyb_reset(self._h, mask_ptr, obs_ptr, step_type_ptr, stream))

    def step_ptr(self, action_ptr, obs_ptr, reward_ptr, discount_ptr, step_type_ptr, stream=0):
        self._ck(lib().flyb_step(self._h, action_ptr, obs_ptr, reward_ptr, discount_ptr, step_type_ptr, stream))

    def step_noalloc_ptr(self, action_ptr, obs_ptr, reward_ptr, discount_ptr, step_type_ptr, stream=0):
        """flyb_step with the kernel's scratch in the library's static device array: the call allocates nothing (flyb.h)."""
        self._ck(lib().flyb_step_noalloc(self._h, action_ptr, obs_ptr, reward_ptr, discount_ptr, step_type_ptr, stream))

    # ---- host (numpy) calls
    def reset_host(self, mask=None):
        N = self.num_envs
        obs, st = np.zeros((N, self.obs_dim), np.float32), np.zeros(N, np.int32)
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        self._ck(lib().flyb_reset_host(self._h, m.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8)) if m is not None else None,
                                       _fp(obs), _ip(st)))
        return st, obs

    def step_host(self, action, out=None):
        N = self.num_envs
        a = np.ascontiguousarray(action, np.float32).reshape(N, self.action_dim)
        if out is None:
            out = (np.zeros((N, self.obs_dim), np.float32), np.zeros(N, np.float32), np.zeros(N, np.float32), np.zeros(N, np.int32))
        obs, r, g, st = out
        self._ck(lib().flyb_step_host(self._h, _fp(a), _fp(obs), _fp(r), _fp(g), _ip(st)))
        return st, r, g, obs

    def get_state(self):
        N = self.num_envs
        q, v = np.zeros((N, self.nq), np.float32), np.zeros((N, self.nv), np.float32)
        a, w = np.zeros((N, max(self.na, 1)), np.float32), np.zeros((N, self.nv), np.float32)
        si = np.zeros((N, int(self.info(INFO_STATE_INTS))), np.int32)
        self._ck(lib().flyb_get_state(self._h, _fp(q), _fp(v), _fp(a), _fp(w), _ip(si)))
        return dict(qpos=q, qvel=v, act=a[:, :self.na], warm=w, ints=si)

    def set_state(self, qpos=None, qvel=None, act=None, warm=None, ints=None):
        c = lambda x, dt: None if x is None else np.ascontiguousarray(x, dt)
        q, v, a, w, si = c(qpos, np.float32), c(qvel, np.float32), c(act, np.float32), c(warm, np.float32), c(ints, np.int32)
        self._ck(lib().flyb_set_state(self._h, _fp(q), _fp(v), _fp(a), _fp(w), _ip(si)))

    def get_trace(self):
        N = self.num_envs
        ti = np.zeros((N, int(self.info(INFO_TRACE_INTS))), np.int32)
        tr = np.zeros((N, int(self.info(INFO_TRACE_REALS))), np.float32)
        self._ck(lib().flyb_get_trace(self._h, _ip(ti), _fp(tr)))
        mc, nv, ns = int(self.info(INFO_MAXCON)), self.nv, int(self.M["nsensordata"])
        return dict(ncon=ti[:, 0], nefc=ti[:, 1], nlimit=ti[:, 2], niter=ti[:, 3], overflow=ti[:, 4], ncand=ti[:, 5],
                    ncon0=ti[:, 6], pairs=ti[:, 8:8 + mc], pairs0=ti[:, 8 + mc:8 + 2 * mc], dist=tr[:, :mc], qacc=tr[:, mc:mc + nv], sensordata=tr[:, mc + nv:mc + nv + ns],
                    efc_force=tr[:, mc + nv + ns:])

    def episode_stats(self):
        out = np.zeros(8, np.float64)
        self._ck(lib().flyb_episode_stats(self._h, out.ctypes.data_as(ctypes.POINTER(ctypes.c_double))))
        return dict(return_sum=out[0], length_sum=out[1], fatal=out[2], episodes=out[3], overflow_steps=out[4],
                    resets=out[5], steps=out[6])
