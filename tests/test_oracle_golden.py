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
