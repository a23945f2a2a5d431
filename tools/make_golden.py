"""Generates tests/golden/*.npz by running the importable pure-numpy pieces of the REFERENCE
(/root/reference/vnl_ray/{quaternions.py,tasks/rewards.py}) on states produced by the float64 oracle.
Run in the build container only (needs /root/reference); the fixtures are committed and travel to the GPU box.

What the fixtures pin:
  walk_reward.npz  - per control step of a seeded walk_imitation rollout: the reward that
                     rewards.reward_factors_deep_mimic(get_walker_features-equivalent, get_reference_features, weights=(20,1,1,1))
                     returns, and the ref_displacement / ref_root_quat observations computed with the reference's
                     quaternions.get_dquat_local, from the oracle's qpos / qvel / site_xpos / xaxis at that step.
  quaternions.npz  - input/output pairs of the reference quaternion helpers the task layer restates.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")
import numpy as np

from vnl_ray import quaternions as Q                      # reference code
from vnl_ray.tasks import rewards as R                    # reference code
from flybody_b200 import clips as C
from flybody_b200.compile_fly import load_model
from flybody_b200.model_pack import asset_path
from oracle.oracle import OracleEnv

OUT = os.path.join(ROOT, "tests", "golden")


def walk_reward(n_steps=40, seed=0, T=120):
    M = load_model(asset_path("walk_imitation"))
    clip = C.synth_walk_clip(M, T=T, seed=seed)
    rows, off, ln = C.pack_clips(M, [clip])
    r = rows.astype(np.float64)   # the engine stores clips in float32; the oracle gets the same rounded values
    nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
    ref = dict(qpos=r[:, :7 + nj], qvel=r[:, 7 + nj:13 + 2 * nj], root2site=r[:, 13 + 2 * nj:13 + 2 * nj + 3 * ns].reshape(-1, ns, 3),
               joint_quat=r[:, 13 + 2 * nj + 3 * ns:].reshape(-1, nj, 4))
    o = OracleEnv(M)
    o.set_clip(ref["qpos"], ref["qvel"], ref["root2site"], ref["joint_quat"])
    o.env_reset()
    rng = np.random.RandomState(seed)
    F = int(M["future_steps"]) + 1
    ra, rd = int(M["root_qposadr"]), int(M["root_dofadr"])
    actions, rewards, ref_disp, ref_quat, oracle_rewards = [], [], [], [], []
    for k in range(n_steps):
        a = rng.uniform(M["action_min"], M["action_max"]).astype(np.float32)
        st, orew, og, oobs = o.env_step(a.astype(np.float64))
        step = k + 1
        qpos, qvel = o.d("qpos").copy(), o.d("qvel").copy()
        jq = [int(M["jnt_qposadr"][j]) for j in M["mocap_joint"]]
        jd = [int(M["jnt_dofadr"][j]) for j in M["mocap_joint"]]
        mq = np.concatenate([qpos[ra:ra + 7], qpos[jq]])          # physics.bind(mocap_joints).qpos
        mv = np.concatenate([qvel[rd:rd + 6], qvel[jd]])
        sites = o.d("site_xpos").reshape(-1, 3)[np.asarray(M["mocap_site"])]
        xaxis1 = o.d("xaxis").reshape(-1, 3)[np.asarray(M["mocap_joint"])]
        # rewards.get_walker_features, lines 39-60, with physics.bind(...) replaced by the arrays above
        root2site = Q.get_egocentric_vec(mq[:3], sites, mq[3:7])
        root_quat = mq[3:7]
        xaxis1 = Q.rotate_vec_with_quat(xaxis1, Q.reciprocal_quat(root_quat))
        joint_quat = np.vstack((root_quat, Q.joint_orientation_quat(xaxis1, mq[7:])))
        walker_ft = {"com": mq[:3], "qvel": mv, "root2site": root2site, "joint_quat": joint_quat}
        reference_ft = R.get_reference_features(ref, step)
        rew = float(np.prod(R.reward_factors_deep_mimic(walker_ft, reference_ft, weights=(20, 1, 1, 1))))
        # tasks/base.py:237-261
        xmat = o.d("xmat").reshape(-1, 3, 3)[int(M["root_body"])]
        ref_pos = ref["qpos"][step:step + F, :3]
        disp = (ref_pos - qpos[ra:ra + 3]) @ xmat
        dq = Q.get_dquat_local(qpos[ra + 3:ra + 7], ref["qpos"][step:step + F, 3:7])
        actions.append(a); rewards.append(rew); ref_disp.append(disp); ref_quat.append(dq); oracle_rewards.append(orew)
        if st == 2:
            break
    np.savez(os.path.join(OUT, "walk_reward.npz"), seed=seed, T=T, actions=np.array(actions), reward=np.array(rewards),
             ref_displacement=np.array(ref_disp), ref_root_quat=np.array(ref_quat))
    print("walk_reward: steps", len(rewards), "max |oracle - reference| reward", np.abs(np.array(rewards) - np.array(oracle_rewards)).max())


def quaternion_vectors(seed=1):
    rng = np.random.RandomState(seed)
    q1, q2 = rng.normal(size=(64, 4)), rng.normal(size=(64, 4))
    v = rng.normal(size=(64, 3))
    v[:4] = [[0, 0, 1.0], [0, 0, -1.0], [0, 0, 2.5], [1e-30, 0, 1.0]]   # quat_z2vec edge cases
    ang = rng.uniform(-3, 3, 64)
    np.savez(os.path.join(OUT, "quaternions.npz"), q1=q1, q2=q2, v=v, ang=ang,
             mult=Q.mult_quat(q1, q2), recip=Q.reciprocal_quat(q1), rot=Q.rotate_vec_with_quat(v, q1),
             dquat_local=Q.get_dquat_local(q1, q2), z2vec=Q.quat_z2vec(v), dist=Q.quat_dist_short_arc(q1, q2),
             joint_quat=Q.joint_orientation_quat(v, ang), egocentric=Q.get_egocentric_vec(v[0], v, q1[0]))
    print("quaternions: ok")


def terrain_vectors():
    """Terrains of the vision-flight arenas from the reference's OWN functions.  vnl_ray/tasks/arenas/hills.py imports
    dm_control at module level, so its four module-level functions (pure numpy / scipy) are lifted out of the source file
    with `ast` and executed as they stand; the arenas' `initialize_episode` bodies (hills.py:357-390, 439-452) are replayed
    here call by call with the same RandomState."""
    import ast
    from scipy import ndimage
    src = open("/root/reference/vnl_ray/tasks/arenas/hills.py").read()
    tree = ast.parse(src)
    want = ("pos_to_terrain_idx", "terrain_bowl", "add_sine_bumps", "add_sine_trench")
    ns = {"np": np, "ndimage": ndimage}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in want:
            exec(compile(ast.Module([node], []), "hills.py", "exec"), ns)

    class _Model:   # what terrain_bowl reads from physics.model (hills.py:160-166: dim 20, grid_density 10)
        hfield_size = np.array([[20.0, 20.0, 1.0, 0.05]])
        hfield_nrow = np.array([401])
        hfield_ncol = np.array([401])

    class _Physics:
        model = _Model()
    out = {}
    for seed in (0, 1):
        rs = np.random.RandomState(seed)      # SineBumps.initialize_episode
        bowl = ns["terrain_bowl"](_Physics(), elevation_z=rs.uniform(4.0, 5.0), random_state=rs)
        size = _Model.hfield_size[0, :2]
        t = ns["add_sine_bumps"](bowl, size, wavelength=rs.uniform(10.0, 15.0), phase=rs.uniform(0.0, 2 * np.pi), height=rs.uniform(0.5, 1.0))
        out[f"bumps{seed}_sub"] = t[::8, ::8].copy()
        out[f"bumps{seed}_stats"] = np.array([t.min(), t.max(), t.mean(), (t ** 2).sum()])
        rs = np.random.RandomState(seed)      # SineTrench.initialize_episode
        bowl = ns["terrain_bowl"](_Physics(), elevation_z=rs.uniform(4.0, 5.0), random_state=rs)
        start_x = rs.uniform(-5.0, -3.0)
        end_x = start_x + rs.uniform(4.0, 10.0)
        amplitude = rs.uniform(0.7 / 2, 1.2 / 2)
        width = 2 * amplitude + 0.604 * rs.uniform(0.5, 1)
        t, sine = ns["add_sine_trench"](bowl, size, start_x=start_x, end_x=end_x, phase=rs.uniform(0.0, 2 * np.pi),
                                        wavelength=rs.uniform(5.0, 8.0), amplitude=amplitude, width=width,
                                        height=rs.uniform(1.3, 1.3), sigma=rs.uniform(0.2, 0.2))
        out[f"trench{seed}_sub"] = t[::8, ::8].copy()
        out[f"trench{seed}_stats"] = np.array([t.min(), t.max(), t.mean(), (t ** 2).sum()])
        out[f"trench{seed}_x"] = np.linspace(start_x, end_x, sine.shape[0])
        out[f"trench{seed}_y"] = sine
        out[f"trench{seed}_row"] = t[200].copy()       # the row through the arena's centre: walls and trench floor
    out["idx"] = np.array([ns["pos_to_terrain_idx"](x, y, size, 401, 401) for x, y in ((-5.0, 0.0), (3.3, -1.2), (19.99, 7.0), (0.04, 0.06))])
    np.savez_compressed(os.path.join(OUT, "terrain.npz"), **out)
    print("terrain: ok")


def wbpg_vectors(seed=2, n_steps=1500):
    """wbpg.npz - the reference WingBeatPatternGenerator (tasks/pattern_generators.py) run on the synthetic base
    pattern: table sizes and checksums, and a reset + n_steps trace (requested frequency, table step, frequency index,
    wing angles) for a frequency command that wanders over the whole range."""
    import tempfile
    from vnl_ray.tasks.pattern_generators import WingBeatPatternGenerator   # reference code
    from flybody_b200 import wbpg as W
    base = W.synthetic_base_pattern()
    with tempfile.NamedTemporaryFile(suffix=".npy") as f:
        np.save(f.name, base)
        g = WingBeatPatternGenerator(base_pattern_path=f.name)
    lens = np.array([t["traj"].shape[0] for t in g.traj_ctrl], np.int32)
    traj_sum = np.array([t["traj"].sum() for t in g.traj_ctrl])
    phase_sum = np.array([t["phase"].sum() for t in g.traj_ctrl])
    rng = np.random.RandomState(seed)
    phase0 = float(rng.uniform())
    q0, v0 = g.reset(initial_phase=phase0, return_qvel=True)
    # user action in [-1, 1]: slow random walk with occasional jumps, as float32 like the engine's input
    act = np.clip(np.cumsum(rng.normal(scale=0.05, size=n_steps)) + (rng.uniform(size=n_steps) < 0.01) * rng.uniform(-1, 1, n_steps), -1, 1).astype(np.float32)
    steps, fidx, angles, cf = [], [], [], []
    for k in range(n_steps):
        # flight_imitation.py:153-154; float64(act): what the reference's pinned numpy 1.26.4 computes for a float32
        # action (legacy promotion), independent of the numpy running this script
        freq = g.base_beat_freq * (1 + g.rel_freq_range * np.float64(act[k]))
        a = g.step(ctrl_freq=freq)
        steps.append(g._step); fidx.append(g._freq_idx); angles.append(a.copy()); cf.append(g._ctrl_freq)
    np.savez_compressed(os.path.join(OUT, "wbpg.npz"), base=base, lens=lens, traj_sum=traj_sum, phase_sum=phase_sum,
                        beat_freqs=g.beat_freqs, phase0=phase0, q0=q0, v0=v0, act=act, steps=np.array(steps, np.int32),
                        freq_idx=np.array(fidx, np.int32), angles=np.array(angles), ctrl_freq=np.array(cf),
                        row0_f0=g.traj_ctrl[0]["traj"][:4], row0_f200=g.traj_ctrl[200]["traj"][:4])
    print("wbpg: table rows", int(lens.sum()), "frequency switches", int((np.diff(fidx) != 0).sum()))


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    walk_reward()
    quaternion_vectors()
    terrain_vectors()
    wbpg_vectors()
