"""Host-side logic that needs no GPU: observation layout, clip packing, sharding arithmetic, stats reduction (gloo)."""
import os

import numpy as np
import pytest


def test_observation_layout_matches_survey(models):
    from flybody_b200.fly_envs import observation_layout
    lay = observation_layout(models["walk_imitation"])
    dims = {k: int(np.prod(s)) for k, (o, s) in lay.items()}
    assert sum(dims.values()) == 741 == int(models["walk_imitation"]["obs_dim"])
    assert lay["walker/ref_displacement"][1] == (65, 3) and lay["walker/ref_root_quat"][1] == (65, 4)
    assert list(lay) == sorted(lay)  # dm_control orders observation keys
    assert sum(int(np.prod(s)) for _, s in observation_layout(models["walk_on_ball"]).values()) == 289
    assert "walker/ball_qvel" in observation_layout(models["walk_on_ball"])
    lay = observation_layout(models["template"])  # Walking base observables only (tasks/base.py:359-364)
    assert sum(int(np.prod(s)) for _, s in lay.values()) == 286 and "walker/ref_displacement" not in lay and "walker/touch" in lay
    offs = sorted((o, int(np.prod(s))) for o, s in lay.values())
    assert all(offs[i][0] + offs[i][1] == offs[i + 1][0] for i in range(len(offs) - 1))  # blocks tile the row


def test_action_spec_order(models):
    M = models["walk_imitation"]
    assert int(M["action_dim"]) == 59 and (M["action_min"][:6] == 0).all() and (M["action_max"][:6] == 1).all()
    assert sorted(int(i) for i in M["ctrl_action"]) == list(range(59))


def test_clip_packing_roundtrip(models):
    from flybody_b200 import clips as C
    M = models["walk_imitation"]
    cl = [C.synth_walk_clip(M, T=20 + 3 * i, seed=i, features=(i == 0)) for i in range(3)]
    for c in cl[1:]:
        c["root2site"] = np.zeros((c["qpos"].shape[0], 6, 3)); c["joint_quat"] = np.zeros((c["qpos"].shape[0], 66, 4))
    rows, off, ln = C.pack_clips(M, cl)
    assert rows.shape == (20 + 23 + 26, C.clip_row_width(M)) and off.tolist() == [0, 20, 43] and ln.tolist() == [20, 23, 26]
    np.testing.assert_allclose(rows[off[1]:off[1] + ln[1], :73], cl[1]["qpos"], rtol=1e-6, atol=1e-7)
    r2s, jq = C.batch_walker_features(M, _full_qpos(M, cl[0]["qpos"]))
    np.testing.assert_allclose(r2s, cl[0]["root2site"], atol=1e-12)
    np.testing.assert_allclose(jq, cl[0]["joint_quat"], atol=1e-12)


def _full_qpos(M, q):
    full = np.tile(np.asarray(M["qpos_reset"], np.float64), (q.shape[0], 1))
    full[:, :7] = q[:, :7]
    for i, j in enumerate(M["mocap_joint"]):
        full[:, int(M["jnt_qposadr"][j])] = q[:, 7 + i]
    return full


def _worker(rank, world, port, q):
    import torch.distributed as dist
    import torch
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from flybody_b200.sharding import shard_range, reduce_stats
    lo, hi = shard_range(10, rank, world)
    stats = dict(return_sum=float(rank + 1), length_sum=10.0 * (rank + 1), fatal=1.0, episodes=2.0)
    out = reduce_stats(stats, torch.device("cpu"))
    q.put((rank, lo, hi, out))
    dist.destroy_process_group()


def test_sharding_and_stats_allreduce_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, 29512, q)) for r in range(2)]
    [p.start() for p in ps]
    res = sorted(q.get(timeout=120) for _ in ps)
    [p.join() for p in ps]
    assert (res[0][1], res[0][2]) == (0, 5) and (res[1][1], res[1][2]) == (5, 10)
    for r in res:
        assert r[3] == dict(return_sum=3.0, length_sum=30.0, fatal=2.0, episodes=4.0)


def test_hdf5_converter_logic_on_a_dataset_shaped_mapping(models, tmp_path):
    """tools/convert_hdf5_clips.py reads the flybody dataset layout (trajectory_loaders.py:10-223); h5py is absent here, so
    the same code is driven with nested dicts of the same shape."""
    import importlib.util, os
    from flybody_b200 import clips as C
    spec = importlib.util.spec_from_file_location("conv", os.path.join(os.path.dirname(__file__), "..", "tools", "convert_hdf5_clips.py"))
    conv = importlib.util.module_from_spec(spec); spec.loader.exec_module(conv)
    M = models["walk_imitation"]
    nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
    rng = np.random.RandomState(0)
    def walk(T):
        return dict(root_qpos=rng.normal(size=(T, 7)) + [3, -2, 0.1, 0, 0, 0, 0], qpos=rng.normal(size=(T, nj)),
                    root_qvel=rng.normal(size=(T, 6)), qvel=rng.normal(size=(T, nj)),
                    root2site=rng.normal(size=(T, ns, 3)), joint_quat=rng.normal(size=(T, nj, 4)))
    f = {"trajectories": {"00001": walk(9), "00000": walk(7)}, "id2name": {"joints": [b"j"] * nj}}
    cl = conv.walking_clips(f, M)
    assert [c["qpos"].shape[0] for c in cl] == [7, 9]                       # numeric order of the zero-padded keys
    assert np.all(cl[0]["qpos"][0, :2] == 0) and cl[0]["qpos"].shape[1] == 7 + nj
    rows, off, ln = C.pack_clips(M, cl)
    assert rows.shape == (16, C.clip_row_width(M)) and off.tolist() == [0, 7] and ln.tolist() == [7, 9]
    com = np.concatenate([rng.normal(size=(5, 3)) + [1, 2, 3], np.tile([1.0, 0, 0, 0], (5, 1))], axis=1)
    fl = conv.flight_clips({"trajectories": {"0": dict(com_qpos=com, com_qvel=rng.normal(size=(5, 6)))}})
    # identity orientation: root = com - offset, after the xy origin moved to the first frame
    assert np.allclose(fl[0]["qpos"][0, :3], [0, 0, com[0, 2]] - C._COM_OFFSET)
    assert np.allclose(C.root2com(fl[0]["qpos"][2]), com[2, :3] - [com[0, 0], com[0, 1], 0])


def test_generated_model_sizes_match_the_shipped_assets(models):
    """csrc/fly_dims_gen.h (tools/gen_dims.py) drives the per-model step kernels: a stale table would silently send a shipped
    model to the generic kernel (25 % slower).  Every row must equal the sizes of its asset, derived ones included."""
    import os, re
    from tests.emu.emu import EmuEnv
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    text = open(os.path.join(root, "flybody_b200", "csrc", "fly_dims_gen.h")).read()
    fields = re.search(r"X\(index, ([^)]*)\)", text).group(1).replace(" ", "").split(",")
    rows = {m.group(2): [int(x) for x in m.group(1).split(",")[1:]]
            for m in re.finditer(r"X\(([\d, ]+)\) /\* (\w+) \*/", text)}
    assert set(rows) == {"walk_imitation", "walk_on_ball", "template", "flight_imitation", "vision_flight"}
    for task, vals in rows.items():
        M = models[task]
        e = EmuEnv(M, double=False)
        for k, v in zip(fields, vals):
            have = int(e.geti(k)[0]) if k in ("nM", "ntrunk", "gsteps", "nlevel") else (int(M["hfield_geom"] >= 0) if k == "nhfield" else int(M[k]))
            assert have == v, (task, k, have, v)


@pytest.mark.parametrize("task", ["walk_on_ball", "walk_imitation", "flight_imitation", "template", "vision_flight"])
def test_shipped_assets_are_what_the_reference_mjcf_compiles_to(task):
    """flybody_b200/assets/<task>.npz must be bit for bit what compile_fly.compile_task makes of the reference's
    vnl_ray/fruitfly/assets/fruitfly.xml (fruitfly.py:115-326 + tasks/base.py surgery restated in compile_fly.py).  Skipped
    where the reference is not mounted (the GPU box)."""
    import os
    import numpy as np
    xml = "/root/reference/vnl_ray/fruitfly/assets/fruitfly.xml"
    if not os.path.exists(xml):
        pytest.skip("reference not mounted")
    from flybody_b200.compile_fly import compile_task, load_model
    from flybody_b200.model_pack import asset_path
    fresh, shipped = compile_task(task, xml), load_model(asset_path(task))
    assert set(fresh) == set(shipped)
    for k in shipped:
        a, b = shipped[k], fresh[k]
        if isinstance(a, (list, tuple, str)):
            assert list(a) == list(b), k
        else:
            assert np.array_equal(np.asarray(a), np.asarray(b)), k
