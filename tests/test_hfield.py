"""Height-field collision of the vision-flight terrain (first slice): the oracle's restatement of MuJoCo's
mjc_ConvexHField against closed forms, and the engine's device code (compiled for the CPU; on the GPU with -m gpu) against it."""
import numpy as np
import pytest

I3 = np.eye(3)
SPHERE, CAPSULE, ELLIPSOID, CYLINDER = 2, 3, 4, 5


def _queries(field, half, n, seed):
    """Random fly-sized geoms hovering within +-0.03 cm of the terrain surface (some touching, some not)."""
    rng = np.random.RandomState(seed)
    axis = np.linspace(-half, half, field.shape[1])
    types, geoms = [], []
    for _ in range(n):
        t = int(rng.choice([SPHERE, CAPSULE, ELLIPSOID, CYLINDER]))
        x, y = rng.uniform(-6, 6, 2)
        h = field[int(np.argmin(np.abs(axis - y))), int(np.argmin(np.abs(axis - x)))]
        size = np.array([rng.uniform(0.01, 0.06), rng.uniform(0.02, 0.1), rng.uniform(0.01, 0.06)])
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        w, a, b, c = q
        R = np.array([[w*w+a*a-b*b-c*c, 2*(a*b-w*c), 2*(a*c+w*b)], [2*(a*b+w*c), w*w-a*a+b*b-c*c, 2*(b*c-w*a)],
                      [2*(a*c-w*b), 2*(b*c+w*a), w*w-a*a-b*b+c*c]])
        z = h + size.max() * rng.uniform(0.3, 1.6)
        types.append(t)
        geoms.append(np.r_[x, y, z, R.ravel(), size, rng.choice([0.0, 0.0005])])
    return np.array(types, np.int32), np.array(geoms)


def test_oracle_sphere_on_flat_and_tilted_height_fields():
    """Known answers: over a flat field the contact is the plane's (dist = gap, normal +z); over a field z = a + b x the
    prism the sphere's foot point falls in gives dist = (vertical gap) / sqrt(1 + b^2) - r and the plane's normal."""
    from oracle.oracle import narrow_hfield
    flat = np.full((21, 21), 0.5, np.float32)
    size = [2, 2, 1, 0.05]
    assert narrow_hfield([0, 0, 0], I3, size, flat, SPHERE, [0.03, 0.04, 0.56], I3, [0.05, 0, 0]) == []
    (d, p, n), = narrow_hfield([0, 0, 0], I3, size, flat, SPHERE, [0.03, 0.04, 0.54], I3, [0.05, 0, 0])
    assert abs(d + 0.01) < 1e-6 and np.abs(n - [0, 0, 1]).max() < 1e-5 and 0.49 <= p[2] <= 0.5
    # hfield geom offset and margin: the same contact 0.01 lower, found 0.0005 earlier
    (d, p, n), = narrow_hfield([0, 0, -0.01], I3, size, flat, SPHERE, [0.03, 0.04, 0.5403], I3, [0.05, 0, 0], margin=0.0005)
    assert abs(d - 0.0003) < 1e-6
    x = np.linspace(-2, 2, 21)
    tilt = (0.3 + 0.2 * x)[None, :].repeat(21, 0).astype(np.float32)
    hits = narrow_hfield([0, 0, 0], I3, size, tilt, SPHERE, [0.13, 0.07, 0.3 + 0.2 * 0.13 + 0.04], I3, [0.05, 0, 0])
    want_n = np.array([-0.2, 0, 1]) / np.hypot(0.2, 1)
    best = min(hits, key=lambda h: np.abs(h[2] - want_n).max())
    assert np.abs(best[2] - want_n).max() < 1e-5 and abs(best[0] - (0.04 / np.hypot(0.2, 1) - 0.05)) < 1e-6


def test_engine_height_field_code_matches_the_oracle_on_a_sine_bumps_terrain():
    from flybody_b200 import terrain as T
    from oracle.oracle import narrow_hfield
    from tests.emu.emu import hfield_contacts
    t = T.sine_bumps(np.random.RandomState(0))
    field = t["height"].astype(np.float32)
    pose, size = np.r_[0, 0, -0.01, I3.ravel()], [20, 20, 1, 0.05]
    types, geoms = _queries(field, 20.0, 120, 1)
    got = hfield_contacts(field, pose, size, types, geoms)
    touching = 0
    for q in range(len(types)):
        g = geoms[q]
        want = narrow_hfield(pose[:3], I3, size, field, types[q], g[:3], g[3:12].reshape(3, 3), g[12:15], margin=g[15])
        assert len(got[q]) == len(want), q
        for (d1, p1, n1), (d2, p2, n2) in zip(got[q], want):
            assert abs(d1 - d2) < 1e-9 and np.abs(p1 - p2).max() < 1e-9 and np.abs(n1 - n2).max() < 1e-9
        touching += len(want) > 0
    assert 20 < touching < 110


@pytest.mark.gpu
def test_gpu_height_field_contacts_match_the_oracle():
    from flybody_b200 import terrain as T
    from flybody_b200.engine import hfield_contacts
    from oracle.oracle import narrow_hfield
    t = T.sine_trench(np.random.RandomState(2))
    field = t["height"].astype(np.float32)
    pose, size = np.r_[0, 0, -0.01, I3.ravel()], [20, 20, 1, 0.05]
    types, geoms = _queries(field, 20.0, 2000, 3)
    got = hfield_contacts(field, pose, size, types, geoms)
    # Same float64 algorithm, and the translation unit is compiled without multiply-add contraction (engine.compile_library), so
    # that the support-point ties of the portal refinement against a prism fall the way they do on the CPU: the contact lists
    # must agree contact by contact.
    touching, ncon = 0, 0
    for q in range(len(types)):
        g = geoms[q]
        want = narrow_hfield(pose[:3], I3, size, field, types[q], g[:3], g[3:12].reshape(3, 3), g[12:15], margin=g[15])
        assert len(got[q]) == len(want), q
        for (d1, p1, n1), (d2, p2, n2) in zip(got[q], want):
            assert abs(d1 - d2) < 1e-9 and np.abs(p1 - p2).max() < 1e-9 and np.abs(n1 - n2).max() < 1e-9, q
        touching += len(want) > 0
        ncon += len(want)
    print(f"\nheight field on the GPU: {touching} of {len(types)} geoms touching, {ncon} contacts, all equal to the oracle's within 1e-9")
    assert touching > 400
