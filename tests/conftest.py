import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def models():
    from flybody_b200.compile_fly import load_model
    from flybody_b200.model_pack import asset_path
    return {t: load_model(asset_path(t)) for t in ("walk_on_ball", "walk_imitation", "flight_imitation", "template", "vision_flight")}


def clip_arrays(M, rows):
    """float32 packed rows -> the float64 arrays the oracle takes (same rounded values the engine sees)."""
    import numpy as np
    r = rows.astype(np.float64)
    nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
    return r[:, :7 + nj], r[:, 7 + nj:13 + 2 * nj], r[:, 13 + 2 * nj:13 + 2 * nj + 3 * ns], r[:, 13 + 2 * nj + 3 * ns:]


@pytest.fixture(scope="session")
def walk_clip(models):
    from flybody_b200 import clips as C
    M = models["walk_imitation"]
    clip = C.synth_walk_clip(M, T=120, seed=0)
    return C.pack_clips(M, [clip])
