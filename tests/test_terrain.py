"""The vision-flight terrains (flybody_b200/terrain.py) against tests/golden/terrain.npz, which tools/make_golden.py produced by
executing the reference's own functions (vnl_ray/tasks/arenas/hills.py) with the same RandomState."""
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(__file__), "golden", "terrain.npz")


def test_sine_bumps_and_trench_reproduce_the_reference_terrains():
    from flybody_b200 import terrain as T
    g = np.load(GOLD)
    for seed in (0, 1):
        b = T.sine_bumps(np.random.RandomState(seed))
        h = b["height"]
        assert h.shape == (401, 401)
        np.testing.assert_allclose(h[::8, ::8], g[f"bumps{seed}_sub"], rtol=0, atol=1e-12)
        np.testing.assert_allclose([h.min(), h.max(), h.mean(), (h ** 2).sum()], g[f"bumps{seed}_stats"], rtol=1e-12)
        t = T.sine_trench(np.random.RandomState(seed))
        h = t["height"]
        np.testing.assert_allclose(h[::8, ::8], g[f"trench{seed}_sub"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(h[200], g[f"trench{seed}_row"], rtol=0, atol=1e-12)
        np.testing.assert_allclose([h.min(), h.max(), h.mean(), (h ** 2).sum()], g[f"trench{seed}_stats"], rtol=1e-12)
        np.testing.assert_allclose(t["trench_x"], g[f"trench{seed}_x"], atol=1e-14)
        np.testing.assert_allclose(t["trench_y"], g[f"trench{seed}_y"], atol=1e-14)
    got = [T.to_index(x, y, (20.0, 20.0), 401, 401) for x, y in ((-5.0, 0.0), (3.3, -1.2), (19.99, 7.0), (0.04, 0.06))]
    assert np.array_equal(np.array(got), g["idx"])


def test_trench_geometry_and_nearest_height():
    """The trench floor follows its centre line between walls of the drawn height (1.3 cm), and get_hfield_height's nearest-grid
    lookup (tasks/vision_flight.py:92-104) returns the grid value."""
    from flybody_b200 import terrain as T
    t = T.sine_trench(np.random.RandomState(3))
    h, half = t["height"], t["half"]
    k = len(t["trench_x"]) // 2
    x, y = t["trench_x"][k], t["trench_y"][k]
    assert T.nearest_height(h, half, x, y) < 0.05                       # on the centre line: trench floor
    assert T.nearest_height(h, half, x, y + 1.5) > 1.2                  # 1.5 cm to the side: on top of the (smoothed) wall
    axis = np.linspace(-20, 20, 401)
    assert T.nearest_height(h, half, axis[37] + 0.04, axis[250] - 0.04) == h[250, 37]
    tab = T.terrain_table("bumps", 3, seed=1)
    assert tab["height"].shape == (3, 401, 401) and tab["height"].dtype == np.float32
    assert not np.array_equal(tab["height"][0], tab["height"][1])
