"""Offline converter: flybody HDF5 reference datasets -> packed clip tables (.npz) for `fly_envs.walk_imitation(ref_path=...)`
/ `flight_imitation(ref_path=...)`.

Run it wherever `h5py` exists (it is not in the GPU image); the .npz it writes is what `flyb_load_clips` takes.  Layout read
(vnl_ray/tasks/trajectory_loaders.py:10-223): `trajectories/<zero-padded idx>/{root_qpos, qpos, root_qvel, qvel, root2site,
joint_quat}` for walking and `.../{com_qpos, com_qvel}` for flight, `trajectory_lengths`, `timestep_seconds`,
`id2name/{joints,sites}`.  Per clip: walking rows are `[root_qpos | qpos]`, `[root_qvel | qvel]`, root2site, joint_quat with the xy
origin moved to the first frame (:206); flight clips go through `clips.flight_clip_from_com` (xy re-origin :130, com2root).

    python tools/convert_hdf5_clips.py walking  wt_berlin_walk.hdf5  walk_clips.npz  [--max-clips N]
    python tools/convert_hdf5_clips.py flight   flight-dataset.hdf5  flight_clips.npz
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from flybody_b200 import clips as C
from flybody_b200.compile_fly import load_model
from flybody_b200.model_pack import asset_path


def _traj_keys(f):
    return sorted(f["trajectories"].keys(), key=lambda k: int(k))


def walking_clips(f, M, max_clips=None):
    """f: an open h5py.File (or any mapping with the same layout).  Checks the dataset's joint / site lists against the
    model's mocap joints / sites when `id2name` is present."""
    nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
    if "id2name" in f and "joints" in f["id2name"]:
        names = [n.decode() if isinstance(n, bytes) else str(n) for n in np.asarray(f["id2name"]["joints"])]
        if len(names) != nj:
            raise ValueError(f"dataset has {len(names)} joints, the compiled model tracks {nj}")
    out = []
    for k in _traj_keys(f)[:max_clips]:
        t = f["trajectories"][k]
        qpos = np.concatenate([np.asarray(t["root_qpos"], np.float64), np.asarray(t["qpos"], np.float64)], axis=1)
        qvel = np.concatenate([np.asarray(t["root_qvel"], np.float64), np.asarray(t["qvel"], np.float64)], axis=1)
        qpos[:, :2] -= qpos[0, :2]                                     # trajectory_loaders.py:206
        r2s, jq = np.asarray(t["root2site"], np.float64), np.asarray(t["joint_quat"], np.float64)
        if qpos.shape[1] != 7 + nj or r2s.shape[1:] != (ns, 3) or jq.shape[1:] != (nj, 4):
            raise ValueError(f"clip {k}: shapes {qpos.shape} {r2s.shape} {jq.shape} do not match the model ({nj} joints, {ns} sites)")
        out.append(dict(qpos=qpos, qvel=qvel, root2site=r2s, joint_quat=jq))
    return out


def flight_clips(f, max_clips=None):
    return [C.flight_clip_from_com(np.asarray(f["trajectories"][k]["com_qpos"], np.float64),
                                   np.asarray(f["trajectories"][k]["com_qvel"], np.float64)) for k in _traj_keys(f)[:max_clips]]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=["walking", "flight"])
    ap.add_argument("hdf5")
    ap.add_argument("out")
    ap.add_argument("--max-clips", type=int, default=None)
    a = ap.parse_args()
    try:
        import h5py
    except ImportError:
        sys.exit("h5py is not installed here: run this converter where it is, then copy the .npz")
    with h5py.File(a.hdf5, "r") as f:
        if a.kind == "walking":
            M = load_model(asset_path("walk_imitation"))
            cl = walking_clips(f, M, a.max_clips)
        else:
            M = load_model(asset_path("flight_imitation"))
            cl = flight_clips(f, a.max_clips)
    C.save_clips(a.out, M, cl)
    print(f"{len(cl)} clips, {sum(c['qpos'].shape[0] for c in cl)} rows -> {a.out}")


if __name__ == "__main__":
    main()
