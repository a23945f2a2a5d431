"""Batched actor side of the DMPO loop: policy inference, n-step transition assembly and the environment loop, all on
the GPU the environments live on (BASELINE.json configs[3]).

The reference runs `num_actors` (90, `config/train_config_fly.yaml:8`) Ray processes, each stepping ONE dm_control env
through `acme.EnvironmentLoop` with a `DelayedFeedForwardActor` (`vnl_ray/agents/actors.py:21-101`) whose adder is acme's
`NStepTransitionAdder(n_step, discount)` (`agents/ray_distributed_dmpo.py:514-521`; n_step 50, `train_dmpo_ray.py:236`).
Here one process steps N environments per call (`fly_envs.BatchedFlyEnv`) and these classes are the same three roles with a
leading batch dimension:

  LayerNormMLPPolicy        the policy network of `agents/network_factory.py:73-101` (batch_concat -> LayerNormMLP ->
                            MultivariateNormalDiagHead), torch, random init (no checkpoints here): owner of the parameters
                            and fp32 reference
  FusedPolicy               the same network as one hand-written sm_100a kernel (tcgen05 / TMEM / TMA, `flyb_policy_*`)
  BatchedFeedForwardActor   select_action / observe_first / observe / update with the reference actor's action_delay and
                            observation_callback arguments
  NStepTransitionAdder      acme's n-step transitions (o_s, a_s, R, D, o_{s+h}) per env - growing horizons at the start of an
                            episode, full horizon n_step in between, shrinking horizons at its end - assembled from ring
                            buffers on the device and shipped to a `sink` callable in batches (the stand-in for the reverb
                            client, which is not installed here)
  BatchedEnvironmentLoop    acme.EnvironmentLoop over a batched, auto-resetting environment

acme / sonnet / reverb are third-party and absent: their semantics are restated (acme adders/reverb/transition.py, acme
tf/networks), not imported, and the restatement of the adder is tested against a per-env list implementation
(tests/test_actors.py).  The learner, replay and variable server stay out of scope."""
from __future__ import annotations

import math
from typing import Callable, Optional

import os
import sys

import torch
from torch import nn

FIRST, MID, LAST = 0, 1, 2


def batch_concat(observation) -> torch.Tensor:
    """acme tf2_utils.batch_concat: flatten every leaf to [N, -1] and concatenate, leaves in sorted-key order."""
    if isinstance(observation, torch.Tensor):
        return observation.reshape(observation.shape[0], -1)
    return torch.cat([observation[k].reshape(observation[k].shape[0], -1) for k in sorted(observation)], dim=1)


class LayerNormMLPPolicy(nn.Module):
    """Linear -> LayerNorm -> tanh -> (Linear -> ELU)* (acme LayerNormMLP, activate_final=True), then a diagonal Gaussian
    head: mean = Linear, scale = softplus(Linear) * init_scale / softplus(0) + min_scale (MultivariateNormalDiagHead)."""

    def __init__(self, obs_dim: int, action_dim: int, layer_sizes=(512, 512, 512, 512), min_scale: float = 1e-6,
                 tanh_mean: bool = False, init_scale: float = 0.7, fixed_scale: bool = False):
        super().__init__()
        sizes = list(layer_sizes)
        self.first = nn.Linear(obs_dim, sizes[0])
        self.norm = nn.LayerNorm(sizes[0])
        self.rest = nn.ModuleList([nn.Linear(a, b) for a, b in zip(sizes[:-1], sizes[1:])])
        self.mean = nn.Linear(sizes[-1], action_dim)
        self.scale = nn.Linear(sizes[-1], action_dim)
        self.min_scale, self.tanh_mean, self.init_scale, self.fixed_scale = min_scale, tanh_mean, init_scale, fixed_scale
        for lin in [self.first, *self.rest]:
            nn.init.trunc_normal_(lin.weight, std=1.0 / math.sqrt(lin.in_features))
            nn.init.zeros_(lin.bias)
        for lin in (self.mean, self.scale):   # VarianceScaling(1e-4): a near-zero head, as acme initialises it
            nn.init.normal_(lin.weight, std=math.sqrt(1e-4 / lin.in_features))
            nn.init.zeros_(lin.bias)

    def forward(self, observation):
        x = torch.tanh(self.norm(self.first(batch_concat(observation))))
        for lin in self.rest:
            x = nn.functional.elu(lin(x))
        mean = self.mean(x)
        if self.tanh_mean:
            mean = torch.tanh(mean)
        if self.fixed_scale:
            scale = torch.full_like(mean, self.init_scale)
        else:
            scale = nn.functional.softplus(self.scale(x)) * (self.init_scale / math.log(2.0)) + self.min_scale
        return mean, scale

    @torch.no_grad()
    def sample(self, observation, generator: Optional[torch.Generator] = None) -> torch.Tensor:
        mean, scale = self(observation)
        eps = torch.randn(mean.shape, device=mean.device, dtype=mean.dtype, generator=generator)
        return mean + scale * eps


class FusedPolicy:
    """A `LayerNormMLPPolicy` evaluated by the hand-written sm_100a kernel behind `flyb_policy_*` (include/flyb.h,
    csrc/flyb_policy.cu): bf16 operands on the tcgen05 tensor cores with fp32 accumulation in tensor memory, TMA-fed weight
    tiles, LayerNorm / tanh / ELU / softplus epilogues and the Gaussian sample fused into one launch per call (plus the bf16
    cast of the observations).  The torch module stays the owner of the parameters (a learner / variable client updates it);
    `refresh()` re-uploads them.  Same interface as the module: `policy(obs) -> (mean, scale)`, `policy.sample(obs, generator)`.
    No fallback: construction fails without the CUDA library or on a device without tcgen05."""

    def __init__(self, policy: LayerNormMLPPolicy, max_rows: int, device=None):
        import ctypes
        from . import engine
        sizes = [policy.first.out_features] + [lin.out_features for lin in policy.rest]
        if sizes != [512, 512, 512, 512]:
            raise ValueError(f"the fused policy kernel is built for layer_sizes (512, 512, 512, 512), got {tuple(sizes)}")
        self.module = policy
        dev = torch.device(device) if device is not None else policy.first.weight.device
        if dev.type != "cuda":
            raise RuntimeError("FusedPolicy needs the policy on a CUDA device (there is no CPU path)")
        self.device = dev
        self.obs_dim, self.act_dim = policy.first.in_features, policy.mean.out_features
        self.max_rows = int(max_rows)
        self._L = engine.lib()
        self._h = ctypes.c_void_p()
        index = dev.index if dev.index is not None else torch.cuda.current_device()
        if self._L.flyb_policy_create(index, self.obs_dim, self.act_dim, self.max_rows, ctypes.byref(self._h)):
            msg = self._L.flyb_policy_last_error(self._h).decode() if self._h else "allocation failed"
            self._L.flyb_policy_destroy(self._h); self._h = None
            raise RuntimeError("flyb_policy_create: " + msg)
        self.refresh()

    def _check(self, rc, what):
        if rc:
            raise RuntimeError(f"{what}: {self._L.flyb_policy_last_error(self._h).decode()}")

    def refresh(self):
        """Upload the module's current parameters (fp32 on the device) into the kernel's bf16 weight buffers."""
        m = self.module
        ts = [m.first.weight, m.first.bias, m.norm.weight, m.norm.bias]
        for lin in m.rest:
            ts += [lin.weight, lin.bias]
        ts += [m.mean.weight, m.mean.bias, m.scale.weight, m.scale.bias]
        ts = [t.detach().to(self.device, torch.float32).contiguous() for t in ts]
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream().cuda_stream
            self._check(self._L.flyb_policy_set_weights(self._h, *[t.data_ptr() for t in ts], float(m.norm.eps), float(m.init_scale),
                                                        float(m.min_scale), int(bool(m.tanh_mean)), st), "flyb_policy_set_weights")
        self._keep = ts   # alive until the copies enqueued above have run

    def _forward(self, observation, eps, want_stats: bool, want_action: bool):
        obs = batch_concat(observation).to(torch.float32).contiguous()
        n = obs.shape[0]
        mean = torch.empty((n, self.act_dim), device=self.device) if want_stats else None
        scale = torch.empty((n, self.act_dim), device=self.device) if want_stats else None
        action = torch.empty((n, self.act_dim), device=self.device) if want_action else None
        ptr = lambda t: t.data_ptr() if t is not None else None   # noqa: E731
        with torch.cuda.device(self.device):
            st = torch.cuda.current_stream().cuda_stream
            self._check(self._L.flyb_policy_forward(self._h, obs.data_ptr(), n, ptr(eps), ptr(mean), ptr(scale), ptr(action), st),
                        "flyb_policy_forward")
        return mean, scale, action

    def __call__(self, observation):
        mean, scale, _ = self._forward(observation, None, True, False)
        if self.module.fixed_scale:
            scale = torch.full_like(mean, self.module.init_scale)
        return mean, scale

    @torch.no_grad()
    def sample(self, observation, generator: Optional[torch.Generator] = None) -> torch.Tensor:
        n = next(iter(observation.values())).shape[0] if isinstance(observation, dict) else observation.shape[0]
        eps = torch.randn((n, self.act_dim), device=self.device, dtype=torch.float32, generator=generator)
        if self.module.fixed_scale:
            mean, _, _ = self._forward(observation, None, True, False)
            return mean + self.module.init_scale * eps
        return self._forward(observation, eps, False, True)[2]

    @property
    def launches(self) -> int:
        return int(self._L.flyb_policy_launches(self._h))

    def close(self):
        if self._h:
            self._L.flyb_policy_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class NStepTransitionAdder:
    """acme's NStepTransitionAdder (dm-acme adders/reverb/transition.py; the reference leaves dm-acme unpinned, this restates
    the current release) for N environments at once.

    For the step sequence (o_0, a_0, r_0, d_0, o_1, ...) of one episode acme writes one transition after EVERY step t (t = 1, 2,
    ... counts the steps taken): from s = max(0, t - n_step) to t, i.e. horizon h = min(t, n_step) - the growing horizons
    (o_0 -> o_1), (o_0 -> o_2), ... at the start of an episode, then the sliding full horizon (`add`: `_first_idx` only advances
    once `episode_steps >= n_step`, `_write` runs every step) - and, when the episode ends at t = L, the shrinking horizons from
    s = max(0, L - n_step) + 1 .. L - 1 to L (`_write_last`).  Each transition is
        (o_s, a_s, R = sum_{k<h} (prod_{j<k} discount * d_{s+j}) r_{s+k}, D = discount^(h-1) prod_{k<h} d_{s+k}, o_{s+h})
    (`_compute_cumulative_quantities`: the last factor of `discount` is left to the learner).
    Ring buffers live on `device`.  Every call ships two batches to `sink(dict)`: the per-step transitions as FIXED-SHAPE
    [N, ...] tensors with a boolean `valid` mask (no host synchronisation), and - only when an episode ended in this call, which
    costs the call's single device-to-host read - the shrinking-horizon transitions of the ended envs."""

    _force_torch = os.environ.get("FLYB_ADDER_TORCH") == "1"   # (tools/adder_probe.py: time the torch operations on CUDA tensors as well)

    def __init__(self, num_envs: int, obs_dim: int, action_dim: int, n_step: int, discount: float, device="cuda",
                 sink: Optional[Callable[[dict], None]] = None, count_on_host: bool = True):
        self.N, self.n, self.gamma = int(num_envs), int(n_step), float(discount)
        dev = torch.device(device)
        self.obs = torch.zeros((self.N, self.n + 1, obs_dim), device=dev)
        self.act = torch.zeros((self.N, self.n, action_dim), device=dev)
        self.rew = torch.zeros((self.N, self.n), device=dev)
        self.disc = torch.ones((self.N, self.n), device=dev)
        self.count = torch.zeros(self.N, dtype=torch.long, device=dev)   # transitions added to the running episode
        self.started = torch.zeros(self.N, dtype=torch.bool, device=dev)
        self._k = torch.arange(self.n, device=dev)
        self._env = torch.arange(self.N, device=dev)
        self.sink = sink
        self.written = torch.zeros((), dtype=torch.long, device=dev)     # transitions shipped so far (device-side counter)
        self._count_on_host = count_on_host
        if dev.type == "cuda":
            # The flush runs only when an episode ends; its first use of torch's indexing / scan kernels costs ~0.1 s of lazy
            # loading, which would land in some later control step.  Take it here, on an episode of env 0 that is thrown away.
            ended = torch.zeros(self.N, dtype=torch.bool, device=dev)
            ended[0] = True
            self._flush(ended, torch.full((self.N,), self.n + 1, dtype=torch.long, device=dev), ship=False)
            torch.cuda.synchronize(dev)

    @property
    def num_written(self) -> int:
        return int(self.written)

    def reset(self):
        self.count.zero_(); self.started.zero_()

    def add_first(self, timestep, mask: Optional[torch.Tensor] = None):
        """Start of an episode for the envs in `mask` (default: those whose step_type is FIRST).  Fixed shapes, no sync."""
        obs = batch_concat(timestep.observation)
        m = (timestep.step_type == FIRST) if mask is None else mask
        self.count = torch.where(m, torch.zeros_like(self.count), self.count)
        self.started = self.started | m
        self.obs[:, 0] = torch.where(m[:, None], obs, self.obs[:, 0])

    def _returns(self, env, start, horizon):
        """n-step return and discount of the transitions (env[i], start[i]) over horizon[i] <= n steps."""
        idx = (start[:, None] + self._k[None, :]) % self.n
        live = self._k[None, :] < horizon[:, None]
        r = torch.gather(self.rew[env], 1, idx)
        d = torch.where(live, torch.gather(self.disc[env], 1, idx), torch.ones((), device=r.device))
        g = torch.where(live, torch.full_like(d, self.gamma), torch.ones((), device=r.device))
        run = torch.cumprod(g * d, dim=1)
        before = torch.cat([torch.ones_like(run[:, :1]), run[:, :-1]], dim=1)     # prod_{j<k} discount * d_j
        R = (torch.where(live, r, torch.zeros((), device=r.device)) * before).sum(1)
        D = run[:, -1] / self.gamma                                                # discount^(h-1) * prod d
        return R, D

    def _batch(self, env, start, hor, next_obs=None):
        R, D = self._returns(env, start, hor)
        return {"observation": self.obs[env, start % (self.n + 1)], "action": self.act[env, start % self.n], "reward": R,
                "discount": D, "next_observation": self.obs[env, (start + hor) % (self.n + 1)] if next_obs is None else next_obs,
                "env": env, "start": start, "horizon": hor}

    def add(self, action: torch.Tensor, next_timestep):
        """Append (a_t, r_t, d_t, o_{t+1}) for every env that is inside an episode; envs whose `next_timestep` is FIRST
        (auto-reset inside env.step) start a new episode instead and their action is ignored, as the env ignored it.
        On CUDA tensors the whole step is one kernel (`flyb_adder_step`, csrc/flyb_actor.cu); the torch operations below are the
        same arithmetic for CPU tensors (and what tests/test_actors.py holds the kernel to)."""
        st = next_timestep.step_type
        obs = batch_concat(next_timestep.observation)
        e = self._env
        if self.obs.is_cuda and not self._force_torch:
            live, c = self._add_cuda(action, next_timestep, obs)
            if not int(self._any_ended.item()):              # the call's one host read
                return
        else:
            first = st == FIRST
            self.add_first(next_timestep, first)
            live = self.started & ~first
            c = self.count
            slot, oslot = c % self.n, (c + 1) % (self.n + 1)
            # unconditional writes: an env that is not live has count 0 (it was just (re)started, or never started), so these land
            # in slots 0 / 1, which its first live step overwrites before anything reads them - no gather / select needed
            self.act[e, slot] = action
            self.rew[e, slot] = next_timestep.reward
            self.disc[e, slot] = next_timestep.discount
            self.obs[e, oslot] = obs
            c = torch.where(live, c + 1, c)
            self.count = c
            # the transition acme writes after every step: from max(0, t - n) over min(t, n) steps (fixed shape, masked)
            start = torch.clamp(c - self.n, min=0)
            hor = torch.clamp(c, max=self.n).clamp(min=1)
            batch = self._batch(e, start, hor, next_obs=obs)     # the per-step transition ends at the observation just received
            batch["valid"] = live
            self.written += live.sum()
            if self.sink is not None:
                self.sink(batch)
            if not bool((live & (st == LAST)).any()):        # the call's one host read
                return
        self._flush(live & (st == LAST), c)

    def _flush(self, ended, c, ship: bool = True):
        """Shrinking-horizon transitions of the envs whose episode just ended: starts max(0, L - n) + 1 .. L - 1, all ending at L."""
        e = self._env
        last = e[ended]
        j = torch.arange(1, self.n, device=last.device)
        s = c[last][:, None] - self.n + j[None, :]
        ok = s >= 1
        env2 = last[:, None].expand(-1, self.n - 1)[ok]
        start2 = s[ok]
        hor2 = (self.n - j)[None, :].expand(last.numel(), -1)[ok]
        if env2.numel():
            tail = self._batch(env2, start2, hor2)
            tail["valid"] = torch.ones_like(env2, dtype=torch.bool)
            if ship:
                self.written += env2.numel()
                if self.sink is not None:
                    self.sink(tail)

    def _add_cuda(self, action, next_timestep, obs):
        """One launch: episode starts, ring appends, the per-step transition of every env and the `episode ended somewhere` flag
        (self._any_ended).  Returns (live, count) for the flush."""
        from . import engine as _engine
        import ctypes
        L = _engine.lib()
        N, O, A = self.N, self.obs.shape[2], self.act.shape[2]
        dev = self.obs.device
        f32 = lambda t: t.to(dtype=torch.float32).contiguous()
        action, obs = f32(action), f32(obs)
        reward, disc = f32(next_timestep.reward), f32(next_timestep.discount)
        st = next_timestep.step_type.to(dtype=torch.int32).contiguous()
        out_obs = torch.empty((N, O), device=dev); out_act = torch.empty((N, A), device=dev)
        out_R = torch.empty(N, device=dev); out_D = torch.empty(N, device=dev)
        valid = torch.empty(N, dtype=torch.bool, device=dev)
        start = torch.empty(N, dtype=torch.long, device=dev); hor = torch.empty(N, dtype=torch.long, device=dev)
        if not hasattr(self, "_any_ended"):
            self._any_ended = torch.zeros(1, dtype=torch.int32, device=dev)
        vp = ctypes.c_void_p
        stream = torch.cuda.current_stream(dev).cuda_stream
        rc = L.flyb_adder_step(dev.index if dev.index is not None else torch.cuda.current_device(), N, O, A, self.n, ctypes.c_float(self.gamma),
                               vp(self.obs.data_ptr()), vp(self.act.data_ptr()), vp(self.rew.data_ptr()), vp(self.disc.data_ptr()),
                               vp(self.count.data_ptr()), vp(self.started.data_ptr()), vp(action.data_ptr()), vp(obs.data_ptr()),
                               vp(reward.data_ptr()), vp(disc.data_ptr()), vp(st.data_ptr()), vp(out_obs.data_ptr()),
                               vp(out_act.data_ptr()), vp(out_R.data_ptr()), vp(out_D.data_ptr()), vp(valid.data_ptr()),
                               vp(start.data_ptr()), vp(hor.data_ptr()), vp(self.written.data_ptr()), vp(self._any_ended.data_ptr()),
                               vp(stream))
        if rc:
            raise RuntimeError("flyb_adder_step failed")
        if self.sink is not None:
            self.sink({"observation": out_obs, "action": out_act, "reward": out_R, "discount": out_D, "next_observation": obs,
                       "env": self._env, "start": start, "horizon": hor, "valid": valid})
        return valid, self.count


class BatchedFeedForwardActor:
    """`DelayedFeedForwardActor` (`vnl_ray/agents/actors.py:21-101`) over a batch of environments."""

    def __init__(self, policy_network: LayerNormMLPPolicy, adder: Optional[NStepTransitionAdder] = None, variable_client=None,
                 action_delay: Optional[int] = None, observation_callback: Optional[Callable] = None,
                 generator: Optional[torch.Generator] = None):
        self._policy_network, self._adder, self._variable_client = policy_network, adder, variable_client
        self._action_delay, self._observation_callback, self._generator = action_delay, observation_callback, generator
        self._action_queue = [] if action_delay is not None else None

    def select_action(self, observation) -> torch.Tensor:
        if self._observation_callback is not None:
            observation = self._observation_callback(observation)
        action = self._policy_network.sample(observation, self._generator)
        if self._action_delay is not None:                     # actors.py:81-88: zeros while the queue fills
            self._action_queue.append(action)
            action = 0 * action if len(self._action_queue) <= self._action_delay else self._action_queue.pop(0)
        return action

    def observe_first(self, timestep):
        if self._adder:
            self._adder.add_first(timestep)

    def observe(self, action, next_timestep):
        if self._adder:
            self._adder.add(action, next_timestep)

    def update(self, wait: bool = False):
        if self._variable_client:
            self._variable_client.update(wait)


class BatchedEnvironmentLoop:
    """acme.EnvironmentLoop for a batched environment that resets finished envs inside `step` (fly_envs.BatchedFlyEnv)."""

    def __init__(self, environment, actor: BatchedFeedForwardActor):
        self._environment, self._actor = environment, actor
        self._timestep = None
        self.steps = 0

    def run(self, num_steps: int):
        """`num_steps` batched control steps; returns (env-steps taken, sum of rewards) as python numbers."""
        env, actor = self._environment, self._actor
        if self._timestep is None:
            self._timestep = env.reset()
            actor.observe_first(self._timestep)
        reward_sum = torch.zeros((), device=self._timestep.step_type.device, dtype=torch.float64)
        for _ in range(num_steps):
            action = actor.select_action(self._timestep.observation)
            self._timestep = env.step(action)
            actor.observe(action, self._timestep)
            actor.update()
            reward_sum += self._timestep.reward.sum()
            self.steps += 1
        return num_steps * int(self._timestep.step_type.shape[0]), float(reward_sum)
