"""The fly_envs wrappers' extras: template_task filters / action corruptor / observable options, and the
inference-mode trajectory hook of walk_imitation.  Run with -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_template_task_filter_time_constants_are_runtime_fields():
    """joint_filter / adhesion_filter (vnl_ray/fly_envs.py:211-212, fruitfly.py:271-281): act' = act + dt (ctrl - act) / tau per
    physics step, so after one control step (10 physics steps) from rest act = ctrl (1 - (1 - dt / tau)^10)."""
    import torch
    from flybody_b200 import fly_envs
    for jf, af in ((0.01, 0.007), (0.02, 0.003)):
        env = fly_envs.template_task(joint_filter=jf, adhesion_filter=af, num_envs=3, seed=1)
        M = env.M
        env.reset()
        spec = env.action_spec()
        a = torch.as_tensor(0.5 * np.asarray(spec.maximum, np.float32), device=env.device).repeat(3, 1)
        ts = env.step(a)
        act = ts.observation["walker/actuator_activation"][0].cpu().numpy()
        dt, n = float(M["opt_timestep"]), int(M["n_substeps"])
        ctrl = np.zeros(int(M["nu"]))
        for c in range(int(M["nu"])):
            if M["ctrl_action"][c] >= 0:
                ctrl[c] = a[0, int(M["ctrl_action"][c])].item()
        for c, name in enumerate(M["_names_act"]):
            aa = int(M["act_actadr"][c])
            if aa >= 0:
                tau = af if "adhere" in name else jf
                assert abs(act[aa] - ctrl[c] * (1 - (1 - dt / tau) ** n)) < 2e-6 * max(1.0, abs(ctrl[c])), (name, tau)
        env.close()
    with pytest.raises(NotImplementedError):
        fly_envs.template_task(mjcb_control=lambda m, d: None, num_envs=1)


def test_action_corruptor_and_observable_options():
    import torch
    from flybody_b200 import fly_envs
    spec_lo = None
    outs = []
    for noise in (None, 0.05, 0.05):
        env = fly_envs.template_task(num_envs=4, seed=3,
                                     action_corruptor=None if noise is None else fly_envs.GaussianActionNoise(noise),
                                     observables_options={"touch": {"enabled": False}, "walker/force": {"enabled": False}})
        assert "walker/touch" not in env.observation_spec() and "walker/force" not in env.observation_spec()
        assert "walker/accelerometer" in env.observation_spec()
        ts = env.reset()
        assert "walker/touch" not in ts.observation
        a = torch.zeros((4, env.action_spec().shape[0]), device=env.device)
        for _ in range(3):
            ts = env.step(a)
        outs.append(ts.observation["walker/joints_pos"].cpu().numpy().copy())
        env.close()
    assert (outs[0] == outs[0][0]).all()                          # no corruptor: identical envs stay identical
    assert np.abs(outs[1] - outs[0]).max() > 1e-5                   # the corruptor reaches the actuators ...
    assert not (outs[1] == outs[1][0]).all()                      # ... with noise of its own per env ...
    assert (outs[1] == outs[2]).all()                             # ... reproducibly for a given seed
    with pytest.raises(NotImplementedError):
        fly_envs.template_task(num_envs=1, observables_options={"touch": {"delay": 2}})


def test_inference_mode_takes_a_custom_trajectory(models):
    """env.task._traj_generator.set_next_trajectory(qpos, qvel) (tasks/trajectory_loaders.py:226-260): walk_imitation without
    ref_path, a straight-line CoM trajectory supplied at run time; the reference observations must follow it
    (tasks/base.py:237-261) and the reward is the constant 1 (walk_imitation.py:153-154)."""
    import torch
    from flybody_b200 import fly_envs
    M = models["walk_imitation"]
    env = fly_envs.walk_imitation(ref_path=None, num_envs=2, seed=0)
    T = 120
    q = np.zeros((T, 7)); q[:, 0] = 0.002 * np.arange(T); q[:, 2] = 0.1278; q[:, 3] = 1.0     # 1 cm/s along x at stance height
    v = np.zeros((T, 6)); v[:, 0] = 1.0
    env.set_next_trajectory(q, v)
    ts = env.reset()
    F = int(M["future_steps"]) + 1
    disp = ts.observation["walker/ref_displacement"][0].cpu().numpy().reshape(F, 3)
    root = env.engine.get_state()["qpos"][0][:7]
    assert np.abs(root - q[0]).max() < 1e-6                                                # walk_imitation.py:119: starts on the clip
    np.testing.assert_allclose(disp[:, 0], q[:F, 0] - q[0, 0], atol=1e-6)                    # identity orientation: plain differences
    a = torch.zeros((2, env.action_spec().shape[0]), device=env.device)
    for k in range(5):
        ts = env.step(a)
        assert (ts.reward == 1).all()
    d5 = ts.observation["walker/ref_displacement"][0].cpu().numpy().reshape(F, 3)
    s = env.engine.get_state()["qpos"][0]
    R = np.eye(3)                                                                              # world -> root frame of the current pose
    w, x, y, z = s[3:7] / np.linalg.norm(s[3:7])
    R = np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)],
                  [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])
    np.testing.assert_allclose(d5, (q[5:5 + F, :3] - s[:3]) @ R, atol=2e-6)
    with pytest.raises(ValueError):
        env.set_next_trajectory(q[:10], v[:10])
    env.close()


def test_walk_imitation_reads_the_reference_hdf5_layout(models, tmp_path):
    """`walk_imitation(ref_path="*.hdf5")` (vnl_ray/fly_envs.py:75-122 with trajectory_loaders.py:163-223): a dataset in the
    reference's layout, read without h5py, drives the same rollout as the same clips handed over as a list."""
    import torch
    from flybody_b200 import clips as C, fly_envs
    M = models["walk_imitation"]
    cl = [C.synth_walk_clip(M, T=100 + i, seed=i) for i in range(4)]
    p = str(tmp_path / "walk.hdf5")
    C.export_hdf5(p, M, cl, dataset_options={"qpos": dict(chunks=(50, 66), deflate=2)})
    outs = []
    for ref in (p, cl):
        env = fly_envs.walk_imitation(ref_path=ref, num_envs=8, seed=5)
        ts = env.reset()
        a = torch.zeros((8, env.action_spec().shape[0]), device=env.device)
        for _ in range(3):
            ts = env.step(a)
        outs.append((ts.reward.cpu().numpy().copy(), ts.observation["walker/ref_displacement"].cpu().numpy().copy()))
        env.close()
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    assert len(set(outs[0][0].tolist())) > 1          # the envs drew different clips
