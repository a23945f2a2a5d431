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
        "task_input": (2,),
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
                 clips=None, trace: bool = False, canonical_actions: bool = False, wbpg=None, action_corruptor=None,
                 observables_options=None, terrain=None, eye_camera_size: int = 32, eye_camera_fovy: float = 150.0):
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
        # vision_guided_flight: the terrain pool and the two eye images per env (uint8, rendered after every reset / step)
        self._eyes = None
        self._seed, self._env_offset = int(seed), int(env_offset)
        if terrain is not None:
            self.engine.load_terrain(terrain)
            self._eye_size, self._eye_fovy = int(eye_camera_size), float(eye_camera_fovy)
            self._eyes = torch.zeros((N, 2, self._eye_size, self._eye_size, 3), dtype=torch.uint8, device=dev)
        self._obs = torch.zeros((N, O), dtype=torch.float32, device=dev)
        self._reward = torch.zeros(N, dtype=torch.float32, device=dev)
        self._discount = torch.ones(N, dtype=torch.float32, device=dev)
        self._step_type = torch.zeros(N, dtype=torch.int32, device=dev)
        self._layout = observation_layout(model)
        # observables_options (tasks/base.py:127: walker.observables.set_options): of dm_control's per-observable options the
        # batched engine honours `enabled` (a disabled observable is left out of the observation dict and the spec); the kernel
        # still computes it.  Either {"enabled": ...} for all, or {"walker/<name>" | "<name>": {"enabled": ...}}.
        if observables_options:
            per = observables_options if all(isinstance(v, dict) for v in observables_options.values()) else \
                {k: observables_options for k in self._layout}
            for key, opts in per.items():
                unknown = set(opts) - {"enabled"}
                if unknown:
                    raise NotImplementedError(f"observables_options: only 'enabled' is supported by the batched engine, got {sorted(unknown)}")
                full = key if key in self._layout else "walker/" + key
                if full not in self._layout:
                    raise KeyError(f"observables_options: no observable {key!r}")
                if not opts.get("enabled", True):
                    del self._layout[full]
        # action_corruptor (tasks/template_task.py:68-70): called on the [num_envs, action_dim] CUDA tensor of real-valued
        # actions before every control step, as corruptor(action, generator) -> action (generator: a torch.Generator on the
        # env's device seeded from `seed`, standing in for the reference's random_state)
        self._corruptor = action_corruptor
        self._gen = torch.Generator(device=dev)
        self._gen.manual_seed(int(seed) + 7919 * int(env_offset))
        self._clip_tables = clips
        self._canonical = bool(canonical_actions)
        self._amin = torch.as_tensor(np.asarray(model["action_min"], np.float32), device=dev)
        self._amax = torch.as_tensor(np.asarray(model["action_max"], np.float32), device=dev)
        self._A = A

    # ---- dm_env API
    def _obs_dict(self):
        N = self.num_envs
        d = collections.OrderedDict(
            (k, self._obs[:, off:off + int(np.prod(shp))].view((N,) + shp)) for k, (off, shp) in self._layout.items())
        if self._eyes is not None:   # fruitfly.py:676-708: right_eye / left_eye, (size, size, 3) uint8 each
            self.engine.render_ptr(self._eyes.data_ptr(), self._eye_size, self._eye_fovy, stream=self._stream())
            d["walker/left_eye"], d["walker/right_eye"] = self._eyes[:, 1], self._eyes[:, 0]
            d = collections.OrderedDict(sorted(d.items()))
        return d

    def _stream(self):
        return self._torch.cuda.current_stream(self.device).cuda_stream

    def reset(self, mask=None) -> TimeStep:
        mp = 0
        if mask is not None:
            mask = mask.to(device=self.device, dtype=self._torch.uint8).contiguous()
            mp = mask.data_ptr()
        self._torch.cuda.nvtx.range_push("flyb.reset")
        self.engine.reset_ptr(mp, self._obs.data_ptr(), self._step_type.data_ptr(), self._stream())
        self._torch.cuda.nvtx.range_pop()
        return TimeStep(self._step_type, None, None, self._obs_dict())

    def step(self, action) -> TimeStep:
        a = action.to(device=self.device, dtype=self._torch.float32)
        if self._canonical:  # CanonicalSpecWrapper / task_utils.canonical2real (vnl_ray/tasks/task_utils.py:53-76)
            a = self._amin + (a.clamp(-1.0, 1.0) + 1.0) * 0.5 * (self._amax - self._amin)
        if self._corruptor is not None:
            a = self._corruptor(a, self._gen).to(device=self.device, dtype=self._torch.float32)
        a = a.contiguous()
        assert a.shape == (self.num_envs, self._A), f"action must be [{self.num_envs}, {self._A}]"
        self._torch.cuda.nvtx.range_push("flyb.step")     # NVTX range per control step (visible in nsys / ncu timelines)
        self.engine.step_ptr(a.data_ptr(), self._obs.data_ptr(), self._reward.data_ptr(), self._discount.data_ptr(),
                             self._step_type.data_ptr(), self._stream())
        self._torch.cuda.nvtx.range_pop()
        self._last_action = a  # keep alive until the launch has consumed it
        return TimeStep(self._step_type, self._reward, self._discount, self._obs_dict())

    def action_spec(self) -> BoundedArray:
        lo = np.asarray(self.M["action_min"], np.float32)
        hi = np.asarray(self.M["action_max"], np.float32)
        if self._canonical:
            lo, hi = -np.ones_like(lo), np.ones_like(hi)
        return BoundedArray((self._A,), np.float32, "action", lo, hi)

    def observation_spec(self):
        d = collections.OrderedDict((k, Array(shp, np.float32, k)) for k, (off, shp) in self._layout.items())
        if self._eyes is not None:
            for k in ("walker/left_eye", "walker/right_eye"):
                d[k] = Array((self._eye_size, self._eye_size, 3), np.uint8, k)
            d = collections.OrderedDict(sorted(d.items()))
        return d

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

    def set_next_trajectory(self, qpos, qvel, env=None):
        """`InferenceWalkingTrajectoryLoader.set_next_trajectory` (tasks/trajectory_loaders.py:241-248): a custom centre-of-mass
        trajectory, qpos (T, 7) and qvel (T, 6), becomes the reference of the next episode - of every env, or of the envs listed
        in `env`.  Joint and site features of the row are zero (in inference mode the reward is the constant 1,
        walk_imitation.py:153-154; the ghost, ref_displacement / ref_root_quat and the com-distance termination use the root
        pose only).  The trajectory is appended to the device clip table."""
        if self.task_name != "walk_imitation":
            raise ValueError("set_next_trajectory belongs to walk_imitation")
        qpos, qvel = np.asarray(qpos, np.float64), np.asarray(qvel, np.float64)
        if qpos.ndim != 2 or qpos.shape[1] != 7 or qvel.shape != (qpos.shape[0], 6):
            raise ValueError("qpos must be (T, 7) and qvel (T, 6)")
        M = self.M
        nj, ns, T = int(M["n_mocap_joints"]), int(M["n_mocap_sites"]), qpos.shape[0]
        if T < int(M["future_steps"]) + 2:
            raise ValueError(f"the trajectory needs at least future_steps + 2 = {int(M['future_steps']) + 2} rows")
        row = np.zeros((T, (7 + nj) + (6 + nj) + 3 * ns + 4 * nj), np.float32)
        row[:, :7] = qpos
        row[:, 7 + nj:13 + nj] = qvel
        row[:, 13 + 2 * nj + 3 * ns::4] = 1.0            # identity joint quaternions
        if self._clip_tables is None:
            rows, off, ln = row, np.zeros(1, np.int32), np.array([T], np.int32)
        else:
            r0, o0, l0 = self._clip_tables
            rows = np.concatenate([r0, row]); off = np.append(o0, len(r0)).astype(np.int32); ln = np.append(l0, T).astype(np.int32)
        self._clip_tables = (rows, off, ln)
        self.engine.load_clips(rows, off, ln)
        ids = np.full(self.num_envs, -1, np.int32)
        if env is None:
            ids[:] = len(off) - 1
        else:
            ids[np.asarray(env, np.int64)] = len(off) - 1
        self.engine.set_clip_ids(ids)

    def vision_targets(self):
        """(terrain index, target_height, target_speed) of every env's current episode: `task.target_height` / `target_speed`
        of the reference (tasks/vision_flight.py:235-241), recomputed on the host from the same counter-based draws as the device."""
        from . import terrain as _terrain
        if self._eyes is None:
            raise ValueError("vision_targets belongs to vision_guided_flight")
        ep = self.engine.get_state()["ints"][:, 4] - 1
        r = self.M["vision_ranges"]
        return [_terrain.vision_draws(self._seed, self._env_offset + e, int(max(ep[e], 0)), self.engine.n_terrains, (r[0], r[1]), (r[2], r[3]))
                for e in range(self.num_envs)]

    def episode_stats(self):
        return self.engine.episode_stats()

    def close(self):
        self.engine.close()


def _load_clips_arg(M, ref_path):
    """ref_path: the reference's HDF5 dataset (`*.hdf5` / `*.h5`, vnl_ray/tasks/trajectory_loaders.py: read with h5py where it
    exists, else with `hdf5_lite`), a `.npz` written by `clips.save_clips`, a list of clip dicts, or
    'synthetic:<n_clips>:<T>:<seed>'."""
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
    if isinstance(ref_path, str) and ref_path.lower().endswith((".hdf5", ".h5", ".hdf")):
        if not os.path.exists(ref_path):
            raise FileNotFoundError(ref_path)
        return _clips.load_hdf5_clips(M, ref_path)
    raise ValueError("ref_path must be the reference's HDF5 dataset, a clip .npz, a list of clips or 'synthetic:<n>:<T>:<seed>'")


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


def vision_guided_flight(wpg_pattern_path=None, bumps_or_trench: str = "bumps", random_state=None, num_envs: int = 1, device=0,
                         seed: int | None = None, n_terrains: int = 16, terrain_seed: int = 0, target_height_range=(0.5, 0.8),
                         target_speed_range=(20.0, 40.0), init_pos_x_range=(-5.0, -5.0), init_pos_y_range=(0.0, 0.0),
                         **kwargs_arena) -> BatchedFlyEnv:
    """Vision-guided flight over `SineBumps` / `SineTrench` terrain (`vnl_ray/fly_envs.py:160-206`, `tasks/vision_flight.py`).

    The reference regenerates its height field on the host at every episode start (`arenas/hills.py:232-258, 328-375, 427-458`);
    here `n_terrains` terrains are generated once with the reference's draws (`terrain.terrain_table`, `kwargs_arena` = the
    arena's `dim`, `grid_density` and range arguments) and every (env, episode) flies over one of them.  Observations: the
    flight observables, `walker/task_input` (target height, target speed) and the two eye cameras `walker/right_eye`,
    `walker/left_eye` as uint8 [num_envs, 32, 32, 3] tensors, ray-cast on the device (csrc/fly_eyes.cuh says what is drawn).
    Reward, termination (an active contact with the terrain or the ground plane is fatal) and the wing-beat pattern generator
    are the reference's; `initialize_qvel` is off as in the reference's constructor.  target_height_range, target_speed_range,
    init_pos_x_range, init_pos_y_range: the task's constructor arguments (tasks/vision_flight.py:28-31); the start position must
    be a single point (lo == hi), as it is in the reference's configuration."""
    from . import terrain as _terrain, wbpg as _wbpg
    if bumps_or_trench not in ("bumps", "trench"):
        raise ValueError("Only 'bumps' and 'trench' terrains are supported.")
    M = load_model(asset_path("vision_flight"))
    if init_pos_x_range[0] != init_pos_x_range[1] or init_pos_y_range[0] != init_pos_y_range[1]:
        raise NotImplementedError("vision_guided_flight: init_pos_x_range / init_pos_y_range must be single points (lo == hi)")
    M["vision_ranges"] = np.array([*target_height_range, *target_speed_range, *init_pos_x_range, *init_pos_y_range], np.float64)
    dim, density = float(kwargs_arena.get("dim", 20.0)), int(kwargs_arena.get("grid_density", 10))
    if _terrain.grid_shape(dim, density) != int(M["hfield_nrow"]) or dim != float(M["hfield_size"][0]):
        raise NotImplementedError("vision_guided_flight: dim / grid_density other than the compiled model's (20, 10) need a recompiled asset")
    if wpg_pattern_path is None or (isinstance(wpg_pattern_path, str) and wpg_pattern_path == "synthetic"):
        base = _wbpg.synthetic_base_pattern()
    elif isinstance(wpg_pattern_path, str):
        base = np.load(wpg_pattern_path)
    else:
        base = np.asarray(wpg_pattern_path)
    table = _terrain.terrain_table(bumps_or_trench, int(n_terrains), seed=int(terrain_seed), **kwargs_arena)
    return BatchedFlyEnv("vision_flight", M, num_envs, device, _seed(random_state, seed), wbpg=_wbpg.build_tables(base),
                         terrain=_terrain.pack_pool(table))


class GaussianActionNoise:
    """An `action_corruptor` (vnl_ray/fly_envs.py:231-233: "add random noise to the action"): zero-mean Gaussian noise with a
    standard deviation of `std` (scalar or per action dimension), on the device, clipped to the action bounds if given."""

    def __init__(self, std, minimum=None, maximum=None):
        self.std, self.lo, self.hi = std, minimum, maximum

    def __call__(self, action, generator):
        import torch
        std = torch.as_tensor(self.std, dtype=action.dtype, device=action.device)
        out = action + std * torch.randn(action.shape, dtype=action.dtype, device=action.device, generator=generator)
        if self.lo is not None:
            out = torch.maximum(out, torch.as_tensor(self.lo, dtype=action.dtype, device=action.device))
        if self.hi is not None:
            out = torch.minimum(out, torch.as_tensor(self.hi, dtype=action.dtype, device=action.device))
        return out


def template_task(random_state=None, joint_filter: float = 0.01, adhesion_filter: float = 0.007, time_limit: float = 1.0,
                  mjcb_control=None, observables_options=None, action_corruptor=None, num_envs: int = 1, device=0,
                  seed: int | None = None, **kw) -> BatchedFlyEnv:
    """The empty no-op walking task (`vnl_ray/fly_envs.py:209-259`, `tasks/template_task.py`): fly on the floor, reward 1,
    termination on |qacc| only.

    joint_filter / adhesion_filter: time constants of the actuator activation filters (fruitfly.py:271-281), run-time model
    fields here (0 = "filter disabled" would remove the activation states and change the observation layout: not supported).
    action_corruptor: batched form of the reference's hook, `corruptor(action [N, A] CUDA tensor, torch.Generator) -> action`
    (e.g. `GaussianActionNoise`).  observables_options: `enabled` only.  mjcb_control: MuJoCo calls it inside every physics
    step; the control step is one fused kernel launch here, so a per-physics-step Python callback cannot be honoured - use
    action_corruptor (once per control step, on the device) instead."""
    if mjcb_control is not None:
        raise NotImplementedError("template_task: mjcb_control runs inside every physics step; the batched engine fuses the control "
                                  "step into one launch - pass a per-control-step device hook as action_corruptor instead")
    M = load_model(asset_path("template"))
    if joint_filter <= 0 or adhesion_filter <= 0:
        raise NotImplementedError("template_task: a filter time constant of 0 (filter disabled) changes the model's state layout")
    if (joint_filter, adhesion_filter) != (0.01, 0.007):
        dyn = np.array(M["act_dynprm"], np.float64)
        for a, name in enumerate(M["_names_act"]):
            if int(M["act_actadr"][a]) >= 0:
                dyn[a] = adhesion_filter if "adhere" in name else joint_filter
        M["act_dynprm"] = dyn
    if float(time_limit) != 1.0:  # same accumulation as compile_fly.task_tables: first control step with time >= time_limit
        dt, t, k = float(M["opt_timestep"]), 0.0, 0
        while t < time_limit:
            for _ in range(int(M["n_substeps"])):
                t += dt
            k += 1
        M["time_limit_steps"] = k
        M["max_episode_steps"] = int(round(time_limit / float(M["control_dt"]))) + 1
    return BatchedFlyEnv("template", M, num_envs, device, _seed(random_state, seed), action_corruptor=action_corruptor,
                         observables_options=observables_options, **kw)


def _seed(random_state, seed):
    if seed is not None:
        return int(seed)
    if random_state is None:
        return 0
    if isinstance(random_state, (int, np.integer)):
        return int(random_state)
    return int(random_state.randint(0, 2**31 - 1))
