"""vision_guided_flight (vnl_ray/fly_envs.py:160-206, tasks/vision_flight.py, tasks/arenas/hills.py) wired into the
step: terrain contacts through the height-field prism walk inside the collision stage, the task's reset / reward / termination, the
per-episode draws, the eye camera frames.  CPU tests: the engine source under emulation (tests/emu) against the float64 oracle;
-m gpu: the CUDA path through fly_envs against per-env oracles, and the eye images against the stand-alone ray caster."""
import numpy as np
import pytest

from flybody_b200 import terrain as T, wbpg as W


def _oracle_for(M, tab, wb, terrain, height, speed, phase):
    from oracle.oracle import OracleEnv
    o = OracleEnv(M)
    o.set_wbpg(wb)
    trench = tab["kind"] == "trench"
    o.set_terrain(tab["height"][terrain], tab["trench_x"][terrain] if trench else None, tab["trench_y"][terrain] if trench else None)
    o.set_vision_targets(height, speed)
    o.set_initial_phase(phase)
    return o


def _high_ground(height, half=20.0, window=8.0):
    """(x, y) of the highest grid point within `window` of the centre: a start position over raised terrain (the reference's
    init_pos_x_range / init_pos_y_range arguments, vision_flight.py:30-31), so that the crash is onto the height field and not
    onto the ground plane that shows through the flat middle of the bowl."""
    n = height.shape[0]
    ax = np.linspace(-half, half, n)
    sel = np.abs(ax) < window
    sub = height[np.ix_(sel, sel)]
    r, c = np.unravel_index(np.argmax(sub), sub.shape)
    return float(ax[sel][c]), float(ax[sel][r])


def _with_start(M, x, y):
    M = dict(M)
    vr = np.array(M["vision_ranges"], np.float64)
    vr[4:8] = [x, x, y, y]
    M["vision_ranges"] = vr
    return M


def _block_err(M, obs, ref, only=None, skip=()):
    """largest error of an observation block relative to the block's own scale (max(1, largest reference entry))"""
    from flybody_b200.fly_envs import observation_layout
    worst = 0.0
    for k, (off, shp) in observation_layout(M).items():
        if (only is not None and k.split("/")[1] not in only) or k.split("/")[1] in skip:
            continue
        n = int(np.prod(shp))
        worst = max(worst, float(np.abs(obs[off:off + n] - ref[off:off + n]).max() / max(1.0, np.abs(ref[off:off + n]).max())))
    return worst


def _actions(M, rng, k, drop_after):
    """gentle random actions, then wings driven against the pattern so that the fly drops onto the terrain"""
    a = 0.3 * rng.uniform(M["action_min"], M["action_max"])
    if k >= drop_after:
        a[:] = 0
        a[np.asarray(M["wing_action"])] = -1.0
    return a


def test_model_has_the_terrain_the_ground_plane_and_the_eyes(models):
    M = models["vision_flight"]
    names = M["_names_geom"]
    assert names[int(M["hfield_geom"])] == "terrain" and names[1] == "groundplane"        # hills.py:186-221: terrain first
    assert int(M["geom_type"][0]) == 1 and int(M["geom_type"][1]) == 0
    assert np.allclose(M["geom_pos"][0], [0, 0, -0.01]) and np.allclose(M["hfield_size"], [20, 20, 1, 0.05])
    assert int(M["hfield_nrow"]) == int(M["hfield_ncol"]) == 401                            # hills.py:163-164
    fly = int(M["ngeom"]) - 2
    assert int(M["npair"]) == 842 + 2 * fly                                                # every fly geom against both world geoms
    assert int(M["ghost_qposadr"]) < 0 and int(M["user_action"]) == int(M["action_dim"]) - 1
    assert int(M["n_substeps"]) == 4 and abs(float(M["time_limit"]) - 0.4) < 1e-12          # fly_envs.py:190, constants.py:22-23
    from flybody_b200.fly_envs import observation_layout
    keys = [k.split("/")[1] for k in observation_layout(M)]        # (actuator_activation is empty: no filtered actuator is left)
    assert keys == ["accelerometer", "gyro", "joints_pos", "joints_vel", "task_input", "velocimeter", "world_zaxis"]
    assert [M["_names_site"][int(s)] for s in M["eye_site"]] == ["walker/eye_right", "walker/eye_left"]


def test_host_copy_of_the_episode_draws_is_the_devices(models):
    from tests.emu.emu import EmuEnv
    M = models["vision_flight"]
    e = EmuEnv(M, double=False)
    e.set_terrain(T.pack_pool({"kind": "bumps", "height": np.zeros((7, 401, 401), np.float32)}))
    seen = set()
    for seed, env in ((0, 0), (11, 5), (2**40 + 3, 16383), (7, 1 << 20)):
        e.set_seed(seed, env)
        for ep in (0, 1, 2, 977):
            got, want = e.vision_draws(ep), T.vision_draws(seed, env, ep, 7)
            assert got == want, (got, want)
            assert 0.5 <= got[1] <= 0.8 and 20 <= got[2] <= 40 and 0 <= got[0] < 7
            seen.add(got)
    assert len(seen) == 16


@pytest.mark.parametrize("kind", ["trench", "bumps"])
def test_emulated_engine_matches_the_oracle_through_a_crash(models, kind):
    """One episode under emulation (double: to rounding; float: the GPU's arithmetic): hover under gentle actions, then drop onto
    the terrain - height-field contacts enter the solver and the first active floor contact ends the episode with discount 0."""
    from tests.emu.emu import EmuEnv
    tab = T.terrain_table(kind, 3, seed=0)
    wb = W.build_tables(W.synthetic_base_pattern())
    x0, y0 = _high_ground(tab["height"][T.vision_draws(11, 5, 0, 3)[0]])
    M = _with_start(models["vision_flight"], x0, y0)
    for dbl in (True, False):
        e = EmuEnv(M, double=dbl)
        e.set_wbpg(wb); e.set_terrain(T.pack_pool(tab)); e.set_seed(11, 5); e.set_initial_phase(0.3)
        t, h, v = e.vision_draws(0)
        o = _oracle_for(M, tab, wb, t, h, v, 0.3)
        st, obs = e.env_reset()
        ost, oobs = o.env_reset()
        assert st == ost == 0
        ti = int(M["obs_offset"][10])
        assert np.allclose(obs[ti:ti + 2], [h, v], atol=1e-6)
        z0 = T.nearest_height(tab["height"][t], (20, 20), x0, y0) + h              # vision_flight.py:128-131
        assert abs(o.d("qpos")[2] - z0) < 1e-6 and np.allclose(o.d("qpos")[:2], [x0, y0]) and z0 - h > 0.4
        eyes, et = e.eyes()
        assert et == t and abs(eyes[0][1] + eyes[1][1] - 2 * o.d("qpos")[1]) < 0.01   # the eyes sit symmetrically about the mid-plane
        rng = np.random.RandomState(0)
        ended, hf_contacts = False, 0
        for k in range(400):
            a = _actions(M, rng, k, 30)
            st, r, g, obs = e.env_step(a)
            ost, orr, og, oobs = o.env_step(a)
            hf_contacts += sum(c["geom1"] == int(M["hfield_geom"]) for c in o.contacts())
            assert st == ost and g == og, (k, st, ost, g, og)
            assert abs(r - orr) < (1e-6 if dbl else 2e-4), (k, r, orr)
            assert _block_err(M, obs, oobs) < (1e-5 if dbl else 2e-3), (k, _block_err(M, obs, oobs))
            if st == 2:
                ended = True
                assert g == 0.0 and hf_contacts > 0          # ended by the terrain, not by the clock
                break
        assert ended


def _tolerance(x, bounds, margin):
    """dm_control.utils.rewards.tolerance(x, bounds, margin, sigmoid="linear", value_at_margin=0) - third party, restated: 1 inside
    the bounds, falling linearly to 0 at `margin` outside them"""
    lo, hi = bounds
    if lo <= x <= hi:
        return 1.0
    if margin == 0:
        return 0.0
    d = (lo - x if x < lo else x - hi) / margin
    return 1.0 - d if abs(d) < 1 else 0.0


def _reference_reward(M, o, tab, terrain, height, speed):
    """get_reward_factors of tasks/vision_flight.py:157-212 written out in numpy on the oracle's state"""
    q, v = o.d("qpos")[int(M["root_qposadr"]):][:7], o.d("qvel")[int(M["root_dofadr"]):][:6]
    f = [_tolerance(q[2] - T.nearest_height(tab["height"][terrain], (20, 20), q[0], q[1]), (height, height), 0.15)]
    f.append(_tolerance(v[0], (speed, np.inf), 1.1 * speed))
    f.append(_tolerance(np.linalg.norm(v[:3]), (speed, speed), 1.1 * speed))
    vel = [o.d("sensordata")[int(M["sensor_adr"][s]):][:3] for s in range(int(M["nsensor"])) if int(M["sensor_type"][s]) == 2][0]
    f.append(_tolerance(vel[1], (0, 0), 10))
    zaxis = o.d("xmat").reshape(-1, 9)[int(M["root_body"])][6:9]
    th = np.deg2rad(47.5)
    f.append(_tolerance(np.arccos(np.dot([np.sin(th), 0, np.cos(th)], zaxis)), (0, 0), np.pi))
    c = 1.0
    if tab["kind"] == "trench":
        tx, ty = tab["trench_x"][terrain], tab["trench_y"][terrain]
        if tx[0] <= q[0] <= tx[-1]:
            c = _tolerance(q[1], (ty[np.abs(tx - q[0]).argmin()],) * 2, 0.15)
    f.append(c)
    return float(np.prod(f)), f


@pytest.mark.parametrize("kind", ["trench", "bumps"])
def test_oracle_reward_is_the_reference_formula(models, kind):
    """The oracle's C restatement of the six reward factors against the formula read off the reference source, on a flight that
    crosses the trench's x range (centre-of-trench factor active) at various heights and speeds."""
    tab = T.terrain_table(kind, 2, seed=5)
    wb = W.build_tables(W.synthetic_base_pattern())
    x0 = float(tab["trench_x"][0][3]) if kind == "trench" else -5.0
    y0 = float(tab["trench_y"][0][3]) + 0.05 if kind == "trench" else 0.0
    M = _with_start(models["vision_flight"], x0, y0)
    o = _oracle_for(M, tab, wb, 0, 0.55, 25.0, 0.1)
    o.env_reset()
    o.d("qvel")[int(M["root_dofadr"])] = 22.0          # forward, a little below the target speed
    rng = np.random.RandomState(2)
    seen_center, nonzero = 0, 0
    for k in range(120):
        st, r, g, obs = o.env_step(_actions(M, rng, k, 1000))
        want, f = _reference_reward(M, o, tab, 0, 0.55, 25.0)
        assert abs(r - want) < 1e-12, (k, r, want, f)
        seen_center += f[5] < 1.0
        nonzero += r > 0
        if st == 2:
            break
    assert nonzero > 10 and (kind != "trench" or seen_center > 5)


# ------------------------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_gpu_vision_flight_against_per_env_oracles(models):
    import torch
    from flybody_b200 import fly_envs
    M = models["vision_flight"]
    N, seed = 256, 3
    env = fly_envs.vision_guided_flight(bumps_or_trench="trench", num_envs=N, seed=seed, n_terrains=4, terrain_seed=1)
    tab = T.terrain_table("trench", 4, seed=1)
    wb = W.build_tables(W.synthetic_base_pattern())
    spec = env.observation_spec()
    assert list(spec) == ["walker/accelerometer", "walker/gyro", "walker/joints_pos", "walker/joints_vel", "walker/left_eye",
                          "walker/right_eye", "walker/task_input", "walker/velocimeter", "walker/world_zaxis"]
    assert spec["walker/right_eye"].shape == (32, 32, 3) and spec["walker/right_eye"].dtype == np.uint8
    ts = env.reset()
    draws = env.vision_targets()
    ti = ts.observation["walker/task_input"].cpu().numpy()
    assert np.allclose(ti, [[d[1], d[2]] for d in draws], atol=1e-6) and len({d[0] for d in draws}) == 4
    sample = [0, 1, 17, 100, 255]
    oracles = []
    for e in sample:
        t, h, v = draws[e]
        o = _oracle_for(M, tab, wb, t, h, v, T.wbpg_phase0(seed, e, 0))
        ost, oobs = o.env_reset()
        oracles.append(o)
        assert _block_err(M, env.flat_observation[e].cpu().numpy(), oobs) < 2e-3
    rng = np.random.RandomState(5)
    alive, ended_by_terrain = set(sample), 0
    worst_r = 0.0
    for k in range(260):
        a = np.stack([_actions(M, rng, k, 20 + 3 * (e % 7)) for e in range(N)]).astype(np.float32)
        ts = env.step(torch.as_tensor(a, device=env.device))
        st, r, g = ts.step_type.cpu().numpy(), ts.reward.cpu().numpy(), ts.discount.cpu().numpy()
        obs = env.flat_observation.cpu().numpy()
        for e, o in zip(sample, oracles):
            if e not in alive:
                continue
            ost, orr, og, oobs = o.env_step(a[e].astype(np.float64))
            assert st[e] == ost and g[e] == og, (k, e, st[e], ost, g[e], og)
            worst_r = max(worst_r, abs(r[e] - orr))
            assert abs(r[e] - orr) < 5e-4, (k, e, r[e], orr)
            assert _block_err(M, obs[e], oobs) < 5e-3, (k, e, _block_err(M, obs[e], oobs))
            if ost == 2:
                alive.discard(e)
                ended_by_terrain += int(og == 0.0)
        if not alive:
            break
    assert not alive and ended_by_terrain == len(sample)
    stats = env.episode_stats()
    assert stats["overflow_steps"] == 0 and stats["episodes"] >= len(sample)
    print(f"\nvision flight: {len(sample)} sampled envs of {N} equal their oracles to the crash; worst reward difference {worst_r:.2e}")
    env.close()


@pytest.mark.gpu
def test_gpu_vision_flight_at_the_bench_size(models):
    """16 384 envs over a pool of 16 trench terrains, 200 control steps with auto-resets (gentle actions, then the wings are driven
    against the pattern and the flies drop): sampled envs equal their oracles step for step, their episodes end on the floor and
    restart on other terrains with other targets; no capacity overflow, real images."""
    import torch
    from flybody_b200 import fly_envs
    M = models["vision_flight"]
    N, seed, NT = 16384, 17, 16
    env = fly_envs.vision_guided_flight(bumps_or_trench="trench", num_envs=N, seed=seed, n_terrains=NT, terrain_seed=3,
                                        target_height_range=(0.2, 0.3))
    M = env.M
    tab = T.terrain_table("trench", NT, seed=3)
    wb = W.build_tables(W.synthetic_base_pattern())
    ts = env.reset()
    sample = [0, 4143, 4144, 9999, 16383]
    episode = {e: 0 for e in sample}

    def oracle_of(e):
        t, h, v = T.vision_draws(seed, e, episode[e], NT, height_range=(0.2, 0.3))
        o = _oracle_for(M, tab, wb, t, h, v, T.wbpg_phase0(seed, e, episode[e]))
        return o

    oracles = {e: oracle_of(e) for e in sample}
    for e in sample:
        oracles[e].env_reset()
    gen = torch.Generator(device=env.device); gen.manual_seed(3)
    lo = torch.as_tensor(np.asarray(M["action_min"], np.float32), device=env.device)
    hi = torch.as_tensor(np.asarray(M["action_max"], np.float32), device=env.device)
    restarts = 0
    since = torch.zeros(N, dtype=torch.int64, device=env.device)       # control steps since the env's last reset
    wing = torch.as_tensor(np.asarray(M["wing_action"]), device=env.device)
    for k in range(200):
        a = 0.3 * (lo + torch.rand((N, lo.shape[0]), device=env.device, generator=gen) * (hi - lo))
        drop = since >= 5 + (torch.arange(N, device=env.device) % 11)
        a[drop] = 0.0
        a[drop[:, None] & torch.isin(torch.arange(lo.shape[0], device=env.device), wing)[None, :]] = -1.0
        ts = env.step(a)
        since = torch.where(ts.step_type == 0, torch.zeros_like(since), since + 1)
        st, r, g = ts.step_type.cpu().numpy(), ts.reward.cpu().numpy(), ts.discount.cpu().numpy()
        obs, ah = env.flat_observation.cpu().numpy(), a.cpu().numpy()
        for e in sample:
            o = oracles[e]
            ost, orr, og, oobs = o.env_step(ah[e].astype(np.float64))
            if ost == 0:          # the oracle restarted: new episode, new draws - rebuild it on the new terrain
                episode[e] += 1
                restarts += 1
                o = oracles[e] = oracle_of(e)
                ost, oobs = o.env_reset()
                orr, og = 0.0, 1.0
                assert st[e] == 0
            assert st[e] == ost, (k, e, st[e], ost)
            if ost != 0:
                assert g[e] == og and abs(r[e] - orr) < 5e-4, (k, e, g[e], og, r[e], orr)
            assert _block_err(M, obs[e], oobs, skip=("accelerometer",)) < 5e-3, (k, e, _block_err(M, obs[e], oobs, skip=("accelerometer",)))
    stats = env.episode_stats()
    assert stats["overflow_steps"] == 0 and restarts >= 3, (stats, restarts)
    eyes = ts.observation["walker/right_eye"]
    assert eyes.shape == (N, 32, 32, 3) and int(eyes.max()) > 50
    env.close()


@pytest.mark.gpu
def test_gpu_crash_onto_the_height_field(models):
    """Start over the highest ground of the one terrain in the pool (init_pos_x_range / init_pos_y_range, vision_flight.py:30-31):
    the contacts that end the episodes are the height field's (prism walk inside the step kernel), step for step with the oracle."""
    import torch
    from flybody_b200 import fly_envs
    tab = T.terrain_table("bumps", 1, seed=2)
    x0, y0 = _high_ground(tab["height"][0])
    N, seed = 64, 21
    env = fly_envs.vision_guided_flight(bumps_or_trench="bumps", num_envs=N, seed=seed, n_terrains=1, terrain_seed=2,
                                        init_pos_x_range=(x0, x0), init_pos_y_range=(y0, y0), target_height_range=(0.3, 0.5))
    M = env.M
    wb = W.build_tables(W.synthetic_base_pattern())
    env.reset()
    draws = env.vision_targets()
    sample = [0, 9, 63]
    oracles = [_oracle_for(M, tab, wb, *draws[e], T.wbpg_phase0(seed, e, 0)) for e in sample]
    for o in oracles:
        o.env_reset()
    rng = np.random.RandomState(1)
    alive, on_field = set(sample), 0
    for k in range(300):
        a = np.stack([_actions(M, rng, k, 5 + e % 4) for e in range(N)]).astype(np.float32)
        ts = env.step(torch.as_tensor(a, device=env.device))
        st, g = ts.step_type.cpu().numpy(), ts.discount.cpu().numpy()
        r, obs = ts.reward.cpu().numpy(), env.flat_observation.cpu().numpy()
        for e, o in zip(sample, oracles):
            if e not in alive:
                continue
            ost, orr, og, oobs = o.env_step(a[e].astype(np.float64))
            assert st[e] == ost and g[e] == og and abs(r[e] - orr) < 5e-4, (k, e, st[e], ost, g[e], og, r[e], orr)
            # the accelerometer averages the control step's four physics steps: once a geom touches the terrain in one of them
            # its reading is the stiff contact force (solref 0.0002 clamped to 2 dt) over a 1e-3 mg body, where a float32
            # penetration depth moves it by per cents (measured 2.9e-2 of the block's scale); everything else stays at 5e-3
            assert _block_err(M, obs[e], oobs, skip=("accelerometer",)) < 5e-3, (k, e)
            assert _block_err(M, obs[e], oobs, only=("accelerometer",)) < 1e-1, (k, e)
            if ost == 2:
                alive.discard(e)
                on_field += any(c["geom1"] == int(M["hfield_geom"]) and c["efc_address"] >= 0 for c in o.contacts())
        if not alive:
            break
    assert not alive and on_field == len(sample), (alive, on_field)
    assert env.episode_stats()["overflow_steps"] == 0
    env.close()


@pytest.mark.gpu
def test_gpu_eye_images_of_the_env_match_the_standalone_ray_caster(models):
    """The eye observables of the batched env = flyb_render_eyes (tests/test_eyes.py holds that one to the brute-force oracle) fed
    with the eye sites' world frames of a float64 oracle in the same state."""
    from flybody_b200 import fly_envs
    from flybody_b200.engine import render_eyes
    M = models["vision_flight"]
    N, seed = 8, 9
    env = fly_envs.vision_guided_flight(bumps_or_trench="bumps", num_envs=N, seed=seed, n_terrains=3)
    tab = T.terrain_table("bumps", 3, seed=0)
    wb = W.build_tables(W.synthetic_base_pattern())
    ts = env.reset()
    right, left = ts.observation["walker/right_eye"].cpu().numpy(), ts.observation["walker/left_eye"].cpu().numpy()
    assert right.shape == (N, 32, 32, 3) and right.dtype == np.uint8 and not np.array_equal(right, left)
    draws = env.vision_targets()
    for e in (0, 3, 7):
        t, h, v = draws[e]
        o = _oracle_for(M, tab, wb, t, h, v, T.wbpg_phase0(seed, e, 0))
        o.env_reset()
        sx, sm = o.d("site_xpos").reshape(-1, 3), o.d("site_xmat").reshape(-1, 9)
        cams = np.stack([np.r_[sx[int(s)], sm[int(s)]] for s in M["eye_site"]])
        ref = render_eyes(tab["height"], np.array([t, t], np.int32), cams)
        for img, want in ((right[e], ref[0]), (left[e], ref[1])):
            diff = np.abs(img.astype(int) - want.astype(int)).max(-1)
            assert (diff > 1).mean() < 0.02, (e, (diff > 1).mean())
        assert len(np.unique(right[e].reshape(-1, 3), axis=0)) > 10          # a real image
    env.close()
