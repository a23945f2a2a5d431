"""Compiles the three task models from the reference MJCF into flybody_b200/assets/*.npz.
Run here (where /root/reference is mounted); the GPU box only loads the .npz files."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from flybody_b200.compile_fly import compile_task, save_model  # noqa: E402

XML = "/root/reference/vnl_ray/fruitfly/assets/fruitfly.xml"
OUT = os.path.join(os.path.dirname(__file__), "..", "flybody_b200", "assets")
for task in ("walk_on_ball", "walk_imitation", "flight_imitation", "template", "vision_flight"):
    M = compile_task(task, XML)
    save_model(M, os.path.join(OUT, f"{task}.npz"))
    print(task, {k: M[k] for k in ("nq", "nv", "nu", "na", "nbody", "ngeom", "npair", "obs_dim", "action_dim")})
