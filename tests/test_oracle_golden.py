"""The float64 oracle against fixtures produced by the reference's own numpy code (tools/make_golden.py)."""
import os

import numpy as np

from tests.conftest import clip_arrays

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_walk_reward_and_reference_observations_match_reference_code(models):
    from flybody_b200 import clips as C
    from oracle.oracle import OracleEnv
    g = np.load(os.path.join(GOLD, "walk_reward.npz"))
    M = models["walk_imitation"]
    rows, off, ln = C.pack_clips(M, [C.synth_walk_clip(M, T=int(g["T"]), seed=int(g["seed"]))])
    o = OracleEnv(M)
    o.set_clip(*clip_arrays(M, rows))
    o.env_reset()
    offd, offq = int(M["obs_offset"][8]), int(M["obs_offset"][9])
    F = int(M["future_steps"]) + 1
    for k, a in enumerate(g["actions"]):
        st, r, disc, obs = o.env_step(a.astype(np.float64))
        assert abs(r - g["reward"][k]) <= 1e-10 * max(1.0, abs(g["reward"][k]))
        np.testing.assert_allclose(obs[offd:offd + 3 * F].reshape(F, 3), g["ref_displacement"][k], atol=1e-12)
        np.testing.assert_allclose(obs[offq:offq + 4 * F].reshape(F, 4), g["ref_root_quat"][k], atol=1e-12)
    assert st == 2 and disc == 0.0  # the seeded random-action rollout ends by drifting off the reference


def test_reference_quaternion_fixture_is_selfconsistent():
    """Sanity of the fixture itself (the restatements are exercised through the rollout test above)."""
    g = np.load(os.path.join(GOLD, "quaternions.npz"))
    q1, q2 = g["q1"], g["q2"]
    n2 = (q1 ** 2).sum(1, keepdims=True)
    np.testing.assert_allclose(g["recip"], q1 * [1, -1, -1, -1] / n2, atol=1e-14)
    assert g["z2vec"][0].tolist() == [1, 0, 0, 0] and g["z2vec"][1].tolist() == [0, 1, 0, 0]
    d = np.abs((q1 * q2).sum(1)) / np.linalg.norm(q1, axis=1) / np.linalg.norm(q2, axis=1)
    np.testing.assert_allclose(g["dist"], np.arccos(np.minimum(1.0, 2 * d * d - 1)), atol=1e-12)


def test_wbpg_tables_and_state_machine_match_reference_class():
    """tests/golden/wbpg.npz was produced by the reference's WingBeatPatternGenerator (tools/make_golden.py): the host
    table builder must reproduce its tables bit for bit, and the numpy state machine its (step, frequency index, angles)
    trace for a wandering frequency command with 81 frequency switches."""
    import os
    from flybody_b200 import wbpg as W
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "wbpg.npz"))
    t = W.build_tables(G["base"])
    assert np.array_equal(np.diff(t["offsets"]), G["lens"])
    assert np.array_equal(t["beat_freqs"], G["beat_freqs"])
    o = t["offsets"]
    assert np.array_equal(np.array([t["traj"][o[f]:o[f + 1]].sum() for f in range(len(G["lens"]))]), G["traj_sum"])
    assert np.array_equal(np.array([t["phase"][o[f]:o[f + 1]].sum() for f in range(len(G["lens"]))]), G["phase_sum"])
    assert np.array_equal(t["traj"][:4], G["row0_f0"]) and np.array_equal(t["traj"][o[200]:o[200] + 4], G["row0_f200"])
    s = W.WbpgState(t)
    q0, v0 = s.reset(float(G["phase0"]))
    assert np.array_equal(q0, G["q0"]) and np.array_equal(v0, G["v0"])
    for k, a in enumerate(G["act"]):
        ang = s.step(W.requested_freq(a))
        assert (s.step_idx, s.freq_idx) == (int(G["steps"][k]), int(G["freq_idx"][k])), k
        assert np.array_equal(ang, G["angles"][k]) and s.ctrl_freq == G["ctrl_freq"][k]


def test_oracle_flight_env_drives_the_generator_like_the_reference(models):
    """The C oracle's generator (inside flyo_env_step) fed with the golden user-action sequence must visit exactly the
    golden (step, frequency index, filtered frequency) states, and command wing torques from the golden angles."""
    import os
    from flybody_b200 import clips as C, wbpg as W
    from oracle.oracle import OracleEnv
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "wbpg.npz"))
    M = models["flight_imitation"]
    t = W.build_tables(G["base"])
    rows, off, ln = C.pack_clips(M, [C.synth_flight_clip(T=700, seed=3)])
    r = rows.astype(np.float64)
    o = OracleEnv(M)
    o.set_clip(r[:, :7], r[:, 7:13]); o.set_wbpg(t); o.set_initial_phase(float(G["phase0"]))
    o.env_reset()
    jq = [int(M["jnt_qposadr"][j]) for j in M["wing_joint"]]
    jd = [int(M["jnt_dofadr"][j]) for j in M["wing_joint"]]
    assert np.allclose(o.d("qpos")[jq], G["q0"], rtol=0, atol=0) and np.allclose(o.d("qvel")[jd], G["v0"], rtol=0, atol=0)
    for k in range(300):
        a = np.zeros(int(M["action_dim"]))
        a[int(M["user_action"])] = G["act"][k]
        wing_before = o.d("qpos")[jq].copy()
        st, rew, g, obs = o.env_step(a)
        step, fidx, cf = o.wbpg_state()
        assert (step, fidx) == (int(G["steps"][k]), int(G["freq_idx"][k])) and cf == G["ctrl_freq"][k]
        # ctrl of the wing actuators = clip(action + target - qpos) (flight_imitation.py:157-159)
        ctrl = o.d("ctrl")
        for i in range(6):
            c = [c for c in range(int(M["nu"])) if M["ctrl_action"][c] == M["wing_action"][i]][0]
            assert abs(ctrl[c] - (G["angles"][k][i] - wing_before[i])) < 1e-12
        if st == 2:
            break
    assert k >= 50
