"""Batched, GPU-resident counterparts of the task constructors in `vnl_ray/fly_envs.py:29-157`.

Same names and task kwargs as the reference (`walk_on_ball`, `walk_imitation`, `flight_imitation`), plus
`num_envs`, `device`, `seed`.  The returned object keeps the dm_env contract of `composer.Environment`
(`reset() -> TimeStep(FIRST, None, None, obs)`, `step(action) -> TimeStep(MID|LAST, reward, discount, obs)`,
`action_spec()`, `observation_spec()`, `reward_spec()`, `discount_spec()`), with a leading batch dimension on
every field and float32 torch tensors that live on the GPU (what `SinglePrecisionWrapper`,
`vnl_ray/single_precision.py:72-86`, would hand the agent).  An env whose last step was LAST is reset by
its next `step` call, per env - dm_env's auto-reset, `reset_next_step` in composer.

All arithmetic happens in libflyb.so (CUDA, sm_100a); this module only owns buffers and specs.
"""
from __future__ import annotations

import collections
import os
from dataclasses import dataclass

import numpy as np

from . import clips as _clips
from .compile_fly import load_model
from .engine import Engine, FLAG_TRACE
from .model_fields import OBS_KEYS
from .model_pack import asset_path

TimeStep = collections.namedtuple("TimeStep", ["step_type", "reward", "discount", "observation"])
FIRST, MID, LAST = 0, 1, 2


@dataclass(frozen=True)
class Array:
    shape: tuple
    dtype: type
    name: str = ""


@dataclass(frozen=True)
class BoundedArray(Array):
    minimum: np.ndarray = None
    maximum: np.ndarray = None


def observation_layout(M: dict) -> "collections.OrderedDict[str, tuple[int, tuple]]":
    """`walker/<name>` -> (offset in the flat observation row, per-env shape); keys sorted as dm_control sorts them."""
    F = int(M["future_steps"]) + 1
    shapes = {
        "accelerometer": (3,), "actuator_activation": (int(M["na"]),), "appendages_pos": (3 * int(M["n_appendages"]),),
        "ball_qvel": (3,), "force": (3 * int(np.sum(np.asarray(M["sensor_type"]) == 3)),), "gyro": (3,),
        "joints_pos": (int(M["n_obs_joints"]),), "joints_vel": (int(M["n_obs_joints"]),),
        "ref_displacement": (F, 3), "ref_root_quat": (F, 4),
        "touch": (int(np.sum(np.asarray(M["sensor_type"]) == 4)),), "velocimeter": (3,), "world_zaxis": (3,),
    }
    out = collections.OrderedDict()
    for i, k in enumerate(OBS_KEYS):
        off = int(M["obs_offset"][i])
        if off >= 0 and int(np.prod(shapes[k])) > 0:
            out["walker/" + k] = (off, shapes[k])
    return out


class BatchedFlyEnv:
    """`num_envs` fruit-fly environments stepped by one CUDA launch per control step."""

    def __init__(self, task: str, model: dict, num_envs: int, device=0, seed: int = 0, env_offset: int = 0,
                 clips=None, trace: bool = False, canonical_actions: bool = False, wbpg=None):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("flybody_b200 needs a CUDA device (there is no CPU fallback)")
        self._torch = torch
        self.task_name, self.M, self.num_envs = task, model, int(num_envs)
        dev = torch.device(device if isinstance(device, (str, torch.device)) else f"cuda:{int(device)}")
        self.device = dev
        self.engine = Engine(model, num_envs, dev.index or 0, seed, env_offset, FLAG_TRACE if trace else 0)
        if clips is not None:
            self.engine.load_clips(*clips)
        if wbpg is not None:
            self.engine.load_wbpg(wbpg)
        N, O, A = self.num_envs, self.engine.obs_dim, self.engine.action_dim
        self._obs = torch.zeros((N, O), dtype=torch.float32, device=dev)
        self._reward = torch.zeros(N, dtype=torch.float32, device=dev)
        self._discount = torch.ones(N, dtype=torch.float32, device=dev)
        self._step_type = torch.zeros(N, dtype=torch.int32, device=dev)
        self._layout = observation_layout(model)
        self._canonical = bool(canonical_actions)
        self._amin = torch.as_tensor(np.asarray(model["action_min"], np.float32), device=dev)
        self._amax = torch.as_tensor(np.asarray(model["action_max"], np.float32), device=dev)
        self._A = A

    # ---- dm_env API
    def _obs_dict(self):
        N = self.num_envs
        return collections.OrderedDict(
            (k, self._obs[:, off:off + int(np.prod(shp))].view((N,) + shp)) for k, (off, shp) in self._layout.items())

    def _stream(self):
        return self._torch.cuda.current_stream(self.device).cuda_stream

    def reset(self, mask=None) -> TimeStep:
        mp = 0
        if mask is not None:
            mask = mask.to(device=self.device, dtype=self._torch.uint8).contiguous()
            mp = mask.data_ptr()
        self.engine.reset_ptr(mp, self._obs.data_ptr(), self._step_type.data_ptr(), self._stream())
        return TimeStep(self._step_type, None, None, self._obs_dict())

    def step(self, action) -> TimeStep:
        a = action.to(device=self.device, dtype=self._torch.float32)
        if self._canonical:  # CanonicalSpecWrapper / task_utils.canonical2real (vnl_ray/tasks/task_utils.py:53-76)
            a = self._amin + (a.clamp(-1.0, 1.0) + 1.0) * 0.5 * (self._amax - self._amin)
        a = a.contiguous()
        assert a.shape == (self.num_envs, self._A), f"action must be [{self.num_envs}, {self._A}]"
        self.engine.step_ptr(a.data_ptr(), self._obs.data_ptr(), self._reward.data_ptr(), self._discount.data_ptr(),
                             self._step_type.data_ptr(), self._stream())
        self._last_action = a  # keep alive until the launch has consumed it
        return TimeStep(self._step_type, self._reward, self._discount, self._obs_dict())

    def action_spec(self) -> BoundedArray:
        lo = np.asarray(self.M["action_min"], np.float32)
        hi = np.asarray(self.M["action_max"], np.float32)
        if self._canonical:
            lo, hi = -np.ones_like(lo), np.ones_like(hi)
        return BoundedArray((self._A,), np.float32, "action", lo, hi)

    def observation_spec(self):
        return collections.OrderedDict((k, Array(shp, np.float32, k)) for k, (off, shp) in self._layout.items())

    def reward_spec(self):
        return Array((), np.float32, "reward")

    def discount_spec(self):
        return BoundedArray((), np.float32, "discount", np.float32(0), np.float32(1))

    # ---- extras
    @property
    def flat_observation(self):
        """[num_envs, obs_dim] tensor the observation dict is a view of."""
        return self._obs

    def set_next_trajectory_index(self, idx):
        """`WalkImitation.set_next_trajectory_index` (walk_imitation.py:86-89), per env or one index for all."""
        ids = np.full(self.num_envs, int(idx), np.int32) if np.isscalar(idx) else np.asarray(idx, np.int32)
        self.engine.set_clip_ids(ids)

    def episode_stats(self):
        return self.engine.episode_stats()

    def close(self):
        self.engine.close()


def _load_clips_arg(M, ref_path):
    """ref_path: `.npz` written by `clips.save_clips`, a list of clip dicts, or 'synthetic:<n_clips>:<T>:<seed>'."""
    if ref_path is None:
        return None
    if isinstance(ref_path, (list, tuple)):
        return _clips.pack_clips(M, list(ref_path))
    if isinstance(ref_path, str) and ref_path.startswith("synthetic"):
        parts = ref_path.split(":")
        n, T, seed = (int(parts[1]) if len(parts) > 1 else 16), (int(parts[2]) if len(parts) > 2 else 600), \
            (int(parts[3]) if len(parts) > 3 else 0)
        return _clips.synthetic_table(M, n, T, seed)
    if isinstance(ref_path, str) and ref_path.endswith(".npz") and os.path.exists(ref_path):
        d = np.load(ref_path)
        return d["rows"], d["offsets"], d["lengths"]
    raise ValueError("ref_path must be a clip .npz, a list of clips or 'synthetic:<n>:<T>:<seed>' "
                     "(the HDF5 dataset of the reference cannot be read here: h5py is not installed)")


def walk_on_ball(random_state=None, num_envs: int = 1, device=0, seed: int | None = None, **kw) -> BatchedFlyEnv:
    """Tethered fly walking on a floating ball (`vnl_ray/fly_envs.py:125-157`)."""
    M = load_model(asset_path("walk_on_ball"))
    return BatchedFlyEnv("walk_on_ball", M, num_envs, device, _seed(random_state, seed), **kw)


def walk_imitation(ref_path=None, random_state=None, terminal_com_dist: float = 0.3, num_envs: int = 1, device=0,
                   seed: int | None = None, **kw) -> BatchedFlyEnv:
    """Fly tracking a reference walking fly (`vnl_ray/fly_envs.py:75-122`); `ref_path=None` = inference mode."""
    M = load_model(asset_path("walk_imitation"))
    M["terminal_com_dist"] = float(terminal_com_dist)
    M["inference_mode"] = int(ref_path is None)
    return BatchedFlyEnv("walk_imitation", M, num_envs, device, _seed(random_state, seed),
                         clips=_load_clips_arg(M, ref_path), **kw)


def flight_imitation(wpg_pattern_path=None, ref_path=None, random_state=None, terminal_com_dist: float = 2.0,
                     num_envs: int = 1, device=0, seed: int | None = None, **kw) -> BatchedFlyEnv:
    """Fly tracking a flying reference with the wing-beat pattern generator (`vnl_ray/fly_envs.py:29-72`).

    wpg_pattern_path: `.npy` with one wing-beat cycle `(timesteps, 3)` (yaw, roll, pitch) as the reference takes it, an
    array of that shape, or 'synthetic' (SURVEY.md section 8d, config 3).  ref_path: `.npz` of packed flight clips
    (`clips.save_clips`), a list of clip dicts from `clips.flight_clip_from_com`, or 'synthetic:<n_clips>:<T>:<seed>'."""
    from . import wbpg as _wbpg
    M = load_model(asset_path("flight_imitation"))
    M["terminal_com_dist"] = float(terminal_com_dist)
    if wpg_pattern_path is None or (isinstance(wpg_pattern_path, str) and wpg_pattern_path == "synthetic"):
        base = _wbpg.synthetic_base_pattern()
    elif isinstance(wpg_pattern_path, str):
        base = np.load(wpg_pattern_path)
    else:
        base = np.asarray(wpg_pattern_path)
    if isinstance(ref_path, str) and ref_path.startswith("synthetic"):
        parts = ref_path.split(":")
        n, T, cs = (int(parts[1]) if len(parts) > 1 else 16), (int(parts[2]) if len(parts) > 2 else 3006), \
            (int(parts[3]) if len(parts) > 3 else 0)
        clips = _clips.synthetic_flight_table(M, n, T, cs)
    else:
        clips = _load_clips_arg(M, ref_path)
    if clips is None:
        raise ValueError("flight_imitation needs reference clips (ref_path), as the reference does")
    return BatchedFlyEnv("flight_imitation", M, num_envs, device, _seed(random_state, seed), clips=clips,
                         wbpg=_wbpg.build_tables(base), **kw)


def template_task(random_state=None, joint_filter: float = 0.01, adhesion_filter: float = 0.007, time_limit: float = 1.0,
                  mjcb_control=None, observables_options=None, action_corruptor=None, num_envs: int = 1, device=0,
                  seed: int | None = None, **kw) -> BatchedFlyEnv:
    """The empty no-op walking task (`vnl_ray/fly_envs.py:209-259`, `tasks/template_task.py`): fly on the floor, reward 1,
    termination on |qacc| only.  The per-env Python hooks of the reference (`mjcb_control`, `action_corruptor`,
    `observables_options`) have no batched equivalent and must be None; the actuator filters are compiled into the model."""
    if mjcb_control is not None or action_corruptor is not None or observables_options is not None:
        raise NotImplementedError("template_task: Python callbacks / observable options are not supported by the batched engine")
    if (joint_filter, adhesion_filter) != (0.01, 0.007):
        raise NotImplementedError("template_task: the actuator filter time constants are fixed at model compile time (0.01, 0.007)")
    M = load_model(asset_path("template"))
    if float(time_limit) != 1.0:  # same accumulation as compile_fly.task_tables: first control step with time >= time_limit
        dt, t, k = float(M["opt_timestep"]), 0.0, 0
        while t < time_limit:
            for _ in range(int(M["n_substeps"])):
                t += dt
            k += 1
        M["time_limit_steps"] = k
        M["max_episode_steps"] = int(round(time_limit / float(M["control_dt"]))) + 1
    return BatchedFlyEnv("template", M, num_envs, device, _seed(random_state, seed), **kw)


def _seed(random_state, seed):
    if seed is not None:
        return int(seed)
    if random_state is None:
        return 0
    if isinstance(random_state, (int, np.integer)):
        return int(random_state)
    return int(random_state.randint(0, 2**31 - 1))
