"""Wing-beat pattern generator tables for flight_imitation.

Host-side restatement of the table construction of the reference's `WingBeatPatternGenerator.__init__`
(vnl_ray/tasks/pattern_generators.py:18-119): one cycle of wing angles is resampled, for each of `num_freqs` beat
frequencies, into a multi-cycle sequence sampled at the control timestep, with the number of repeats chosen so that the
end of the sequence joins its start as smoothly as possible.  The per-step state machine (`reset :121-157`,
`step :159-191`) runs inside the CUDA kernel (csrc/fly_env.cuh) on the packed tables built here; `WbpgState` below is
the same machine in numpy, used by the tests to compare with the reference class.

Packed layout (float64, as the reference computes in float64 and the table indices must match exactly):
    beat_freqs [F]          frequencies, Hz
    offsets    [F + 1]      first row of frequency f in traj / phase
    traj       [sum T_f, 6] wing angles (yaw, roll, pitch) x (left, right)
    phase      [sum T_f]    phase (in cycles) of every row
"""
from __future__ import annotations

import numpy as np

BASE_FREQ = 218.0          # tasks/constants.py:29-37  _WING_PARAMS
REL_FREQ_RANGE = 0.05
NUM_FREQS = 201
FLY_CONTROL_TIMESTEP = 2e-4  # tasks/constants.py:21


def build_tables(base_pattern, base_beat_freq=BASE_FREQ, rel_freq_range=REL_FREQ_RANGE, num_freqs=NUM_FREQS,
                 min_repeats=10, max_repeats=20, dt_ctrl=FLY_CONTROL_TIMESTEP, ctrl_filter=0.5 / BASE_FREQ):
    """base_pattern: (n, 3) one wing-beat cycle (yaw, roll, pitch).  Returns the packed tables as a dict."""
    base = np.asarray(base_pattern, np.float64)
    if base.ndim != 2 or base.shape[1] != 3:
        raise ValueError("base pattern must have shape (timesteps, 3)")
    both = np.concatenate([base, base], axis=1)               # same pattern for the two wings (:60)
    n = both.shape[0]
    freqs = np.linspace((1 - rel_freq_range) * base_beat_freq, (1 + rel_freq_range) * base_beat_freq, num_freqs)
    reps = np.arange(min_repeats, max_repeats + 1)
    trajs, phases, offsets = [], [], [0]
    for f in freqs:
        period = 1 / f
        err = ((reps * period) % dt_ctrl) / dt_ctrl            # mismatch of the join, in control steps (:83)
        over, under = int(np.argmin(err)), int(np.argmin(np.abs(1 - err)))
        if err[over] < abs(1 - err[under]):
            pick, shift = over, dt_ctrl
        else:
            pick, shift = under, 0.0
        n_reps = pick + 1                                       # sic: index + 1, not reps[index] (:94)
        tiled = np.tile(both, (n_reps, 1))
        cyc = np.linspace(0, n_reps, n_reps * n, endpoint=False)
        duration = tiled.shape[0] * (period / n)
        t_data = np.linspace(0, duration, tiled.shape[0])
        t_ctrl = np.arange(0, duration - shift, dt_ctrl)
        trajs.append(np.stack([np.interp(t_ctrl, t_data, tiled[:, k]) for k in range(6)], axis=1))
        phases.append(np.interp(t_ctrl, t_data, cyc))
        offsets.append(offsets[-1] + t_ctrl.shape[0])
    return {
        "beat_freqs": freqs, "offsets": np.asarray(offsets, np.int32),
        "traj": np.concatenate(trajs, axis=0), "phase": np.concatenate(phases, axis=0),
        "base_freq": float(base_beat_freq), "rel_freq_range": float(rel_freq_range), "dt_ctrl": float(dt_ctrl),
        "rate": float(np.exp(-dt_ctrl / ctrl_filter)) if ctrl_filter != 0.0 else 0.0,
    }


def synthetic_base_pattern(n=64):
    """One synthetic wing-beat cycle (n, 3) inside the wing joint ranges of fruitfly.xml:59-70 (SURVEY.md section 8d,
    config 3): stroke (yaw) sinusoid, small deviation (roll) about the horizontal stroke plane, pitch flip in quadrature.
    The deviation is centred on 0: with the 0.6 rad offset an earlier version used, the wing blades swept through the
    abdomen for a fifth of every cycle (up to 0.46 mm deep once the convex narrow phase saw ellipsoid-cylinder pairs)."""
    t = np.arange(n) / n * 2 * np.pi
    yaw = 0.2 + 1.1 * np.cos(t)
    roll = 0.15 * np.sin(2 * t)
    pitch = -0.3 + 0.9 * np.sin(t)
    return np.stack([yaw, roll, pitch], axis=1)


def requested_freq(user_action, base_freq=BASE_FREQ, rel_freq_range=REL_FREQ_RANGE):
    """flight_imitation.py:151-154.  The reference pins numpy==1.26.4 (setup.py), where `base * (1 + rel * act)` with a
    float32 `act` and python-float constants is evaluated in float64 (legacy value-based promotion; numpy >= 2 would
    stay in float32).  The engine's action input is float32, so kernel and oracle widen it to float64 first."""
    return np.float64(base_freq) * (np.float64(1) + np.float64(rel_freq_range) * np.float64(np.float32(user_action)))


class WbpgState:
    """The reset/step state machine on the packed tables (what the kernel does per env), in numpy float64."""

    def __init__(self, tables):
        self.t = tables

    def _rows(self, f):
        o = self.t["offsets"]
        return self.t["traj"][o[f]:o[f + 1]], self.t["phase"][o[f]:o[f + 1]]

    def reset(self, initial_phase=0.0):
        self.ctrl_freq = np.float64(self.t["base_freq"])
        self.freq_idx = int(np.argmin(np.abs(self.t["beat_freqs"] - self.ctrl_freq)))
        traj, phase = self._rows(self.freq_idx)
        self.step_idx = int(np.argmin(np.abs(initial_phase - phase)))
        return traj[self.step_idx], (traj[self.step_idx + 1] - traj[self.step_idx]) / self.t["dt_ctrl"]

    def step(self, ctrl_freq):
        traj, phase = self._rows(self.freq_idx)
        self.step_idx = (self.step_idx + 1) % traj.shape[0]
        r = np.float64(self.t["rate"])   # the reference's _rate is np.float64: the low-pass runs in float64
        self.ctrl_freq = np.float64(ctrl_freq) if r == 0.0 else self.ctrl_freq * r + ctrl_freq * (1 - r)
        new = int(np.argmin(np.abs(self.t["beat_freqs"] - self.ctrl_freq)))
        if new != self.freq_idx:
            cur = phase[self.step_idx]
            traj, phase_new = self._rows(new)
            self.step_idx = int(np.argmin(np.abs(cur % 1 - phase_new % 1)))
            self.freq_idx = new
        return traj[self.step_idx]
