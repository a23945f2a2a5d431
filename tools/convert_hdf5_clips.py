"""Offline converter: flybody HDF5 reference datasets -> packed clip tables (.npz) for `fly_envs.walk_imitation(ref_path=...)`
/ `flight_imitation(ref_path=...)`.

Not needed to use a dataset - `ref_path="*.hdf5"` is read directly (`flybody_b200/clips.py: load_hdf5_clips`, through h5py where
it is installed and `flybody_b200/hdf5_lite.py` where it is not) - but a 16 252-clip file parses faster once it is packed.
Layout read (vnl_ray/tasks/trajectory_loaders.py:10-223): `trajectories/<zero-padded idx>/{root_qpos, qpos, root_qvel, qvel,
root2site, joint_quat}` for walking and `.../{com_qpos, com_qvel}` for flight, `trajectory_lengths`, `timestep_seconds`,
`id2name/{joints,sites}`.

    python tools/convert_hdf5_clips.py walking  wt_berlin_walk.hdf5  walk_clips.npz  [--max-clips N]
    python tools/convert_hdf5_clips.py flight   flight-dataset.hdf5  flight_clips.npz
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from flybody_b200 import clips as C
from flybody_b200.compile_fly import load_model
from flybody_b200.model_pack import asset_path

walking_clips = C.hdf5_walking_clips      # (f, M, max_clips=None, traj_indices=None)
flight_clips = C.hdf5_flight_clips        # (f, max_clips=None, traj_indices=None)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("kind", choices=["walking", "flight"])
    ap.add_argument("hdf5")
    ap.add_argument("out")
    ap.add_argument("--max-clips", type=int, default=None)
    a = ap.parse_args()
    M = load_model(asset_path("walk_imitation" if a.kind == "walking" else "flight_imitation"))
    with C._open_hdf5(a.hdf5) as f:
        cl = walking_clips(f, M, a.max_clips) if a.kind == "walking" else flight_clips(f, a.max_clips)
    C.save_clips(a.out, M, cl)
    print(f"{len(cl)} clips, {sum(c['qpos'].shape[0] for c in cl)} rows -> {a.out}")


if __name__ == "__main__":
    main()
