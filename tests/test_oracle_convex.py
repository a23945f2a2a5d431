"""Known-answer tests of the oracle's general convex narrow phase (MPR restatement of MuJoCo's mjc_Convex / libccd,
oracle/flyb_oracle.c: mpr_penetration).  No MuJoCo build exists here, so the MPR path is pinned against the oracle's
analytic routines on pairs both can do (sphere-like ellipsoids, capsules, cylinders, spheres) and against closed forms."""
import numpy as np
import pytest

from oracle import oracle as O


def _rot(rng):
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    w, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                     [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                     [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])


def _compare(t1, p1, m1, s1, t2, p2, m2, s2, margin, tol_d=3e-6, tol_n=2e-3):
    """MPR vs the analytic routine.  MPR measures the overlap where the ray from (centre1 - centre2) through the origin leaves
    the Minkowski difference, so it equals the true signed distance when the contact normal lies on the line of centres and
    overestimates the penetration otherwise (a property of the algorithm MuJoCo uses for these pairs, kept as is)."""
    a = O.narrow_pair(t1, p1, m1, s1, t2, p2, m2, s2, margin)
    c = O.narrow_pair(t1, p1, m1, s1, t2, p2, m2, s2, margin, force_convex=True)
    if not a or a[0][0] > margin - 1e-5:   # analytic: no contact, or right at the margin where either answer is fine
        return 0
    assert len(c) == 1, (a, c)
    assert c[0][0] <= a[0][0] + tol_d, (a, c)            # never less overlap than the true minimum
    centres = (np.asarray(p2) - np.asarray(p1)) / np.linalg.norm(np.asarray(p2) - np.asarray(p1))
    if a[0][2] @ centres < 1 - 1e-9:
        return 0
    assert abs(a[0][0] - c[0][0]) < tol_d, (a, c)
    assert np.abs(a[0][2] - c[0][2]).max() < tol_n, (a, c)
    # the MPR contact point is a barycentric estimate inside the overlap of the margin-inflated shapes
    assert np.abs(a[0][1] - c[0][1]).max() < 1e-5 + 0.6 * (margin - a[0][0]), (a, c)
    return 1


def test_convex_matches_sphere_sphere_and_margin_band():
    rng = np.random.RandomState(0)
    hits = 0
    for _ in range(300):
        r1, r2 = rng.uniform(0.02, 0.2, 2)
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        gap = rng.uniform(-0.3, 0.3) * min(r1, r2)      # shallow penetration ... separated inside the margin
        p2 = d * (r1 + r2 + gap)
        hits += _compare(O.GEOM_SPHERE, np.zeros(3), _rot(rng), [r1, 0, 0], O.GEOM_SPHERE, p2, _rot(rng), [r2, 0, 0], 0.05)
    assert hits > 100


def test_convex_matches_sphere_capsule_and_capsule_capsule():
    rng = np.random.RandomState(1)
    hits = 0
    for i in range(400):
        r1, r2, l2 = rng.uniform(0.02, 0.1), rng.uniform(0.02, 0.1), rng.uniform(0.05, 0.3)
        m2 = _rot(rng)
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        if i % 2:       # anywhere near the surface: bound only (unless it happens to be aligned)
            t = rng.uniform(-1.3, 1.3) * l2
            if abs(t) < l2:
                d -= m2[:, 2] * (d @ m2[:, 2]); d /= np.linalg.norm(d)
            c1 = m2[:, 2] * np.clip(t, -l2, l2) + d * (r1 + r2 - rng.uniform(-0.2, 0.3) * min(r1, r2))
        elif i % 4:     # beside the middle of the capsule: normal on the line of centres
            d -= m2[:, 2] * (d @ m2[:, 2]); d /= np.linalg.norm(d)
            c1 = d * (r1 + r2 - rng.uniform(-0.2, 0.3) * min(r1, r2))
        else:           # on the axis beyond a cap
            c1 = m2[:, 2] * rng.choice([-1.0, 1.0]) * (l2 + r1 + r2 - rng.uniform(-0.2, 0.3) * min(r1, r2))
        hits += _compare(O.GEOM_SPHERE, c1, _rot(rng), [r1, 0, 0], O.GEOM_CAPSULE, np.zeros(3), m2, [r2, l2, 0], 0.01)
        if i % 4 == 2:  # crossed capsules touching at their middles
            l1 = rng.uniform(0.05, 0.3)
            ax1 = np.cross(d, m2[:, 2]); ax1 /= np.linalg.norm(ax1)
            m1 = np.stack([d, np.cross(ax1, d), ax1], axis=1)
            hits += _compare(O.GEOM_CAPSULE, c1, m1, [r1, l1, 0], O.GEOM_CAPSULE, np.zeros(3), m2, [r2, l2, 0], 0.01)
    assert hits > 250


def test_convex_matches_sphere_cylinder():
    rng = np.random.RandomState(2)
    hits = 0
    for i in range(400):
        r1, R, H = rng.uniform(0.02, 0.1), rng.uniform(0.05, 0.2), rng.uniform(0.05, 0.2)
        m2 = _rot(rng)
        ang = rng.uniform(0, 2 * np.pi)
        radial = m2[:, 0] * np.cos(ang) + m2[:, 1] * np.sin(ang)
        aligned = i % 2 == 0
        if rng.randint(2):   # side (aligned: at mid height)
            c1 = m2[:, 2] * (0 if aligned else rng.uniform(-0.8, 0.8) * H) + radial * (R + r1 - rng.uniform(-0.1, 0.2) * r1)
        else:                # cap (aligned: on the axis)
            sg = rng.choice([-1.0, 1.0])
            c1 = radial * (0 if aligned else rng.uniform(0, 0.8) * R) + m2[:, 2] * sg * (H + r1 - rng.uniform(-0.1, 0.2) * r1)
        hits += _compare(O.GEOM_SPHERE, c1, _rot(rng), [r1, 0, 0], O.GEOM_CYLINDER, np.zeros(3), m2, [R, H, 0], 0.005)
    assert hits > 150


def test_ellipsoid_against_small_sphere():
    """A tiny sphere just inside/outside an ellipsoid along a surface normal: the overlap is never underestimated, it is the
    signed offset exactly on the principal axes (normal on the line of centres), and the reported normal is the ellipsoid's
    surface normal at the contact point (mjc_fixNormal)."""
    rng = np.random.RandomState(3)
    for i in range(240):
        s = rng.uniform(0.03, 0.3, 3)
        m = _rot(rng)
        if i % 2:
            u = rng.normal(size=3); u /= np.linalg.norm(u)
        else:
            u = np.zeros(3); u[rng.randint(3)] = rng.choice([-1.0, 1.0])
        surf = s * u                                         # point on the ellipsoid (local)
        nl = surf / (s * s); nl /= np.linalg.norm(nl)
        r, off = 1e-3, rng.uniform(-4e-4, 4e-4)
        c_local = surf + nl * (r + off)
        hit = O.narrow_pair(O.GEOM_SPHERE, m @ c_local, np.eye(3), [r, 0, 0], O.GEOM_ELLIPSOID, np.zeros(3), m, s, 1e-3)
        assert len(hit) == 1
        dist, pos, n = hit[0]
        assert dist <= off + 3e-6, (dist, off)
        if i % 2 == 0:
            assert abs(dist - off) < 3e-6, (dist, off)
            assert np.abs(n - (-(m @ nl))).max() < 2e-3      # normal points from the sphere (geom 1) to the ellipsoid (geom 2)
        pl = m.T @ pos
        g = pl / (s * s); g /= np.linalg.norm(g)
        sph = (pos - m @ c_local); sph /= np.linalg.norm(sph)
        expect = sph - m @ g; expect /= np.linalg.norm(expect)
        assert np.abs(n - expect).max() < 1e-9


def test_separated_pairs_make_no_contact_and_results_are_frame_invariant():
    rng = np.random.RandomState(4)
    for _ in range(100):
        s1, s2 = rng.uniform(0.03, 0.2, 3), rng.uniform(0.03, 0.2, 3)
        m1, m2 = _rot(rng), _rot(rng)
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        far = d * (s1.max() + s2.max() + 0.01)
        assert O.narrow_pair(O.GEOM_ELLIPSOID, np.zeros(3), m1, s1, O.GEOM_ELLIPSOID, far, m2, s2, 0.005) == []
        near = d * 0.6 * (s1.min() + s2.min())
        h = O.narrow_pair(O.GEOM_ELLIPSOID, np.zeros(3), m1, s1, O.GEOM_CYLINDER, near, m2, s2, 0.0)
        assert len(h) == 1 and h[0][0] < 0
        R, t = _rot(rng), rng.normal(size=3)
        g = O.narrow_pair(O.GEOM_ELLIPSOID, t, R @ m1, s1, O.GEOM_CYLINDER, R @ near + t, R @ m2, s2, 0.0)
        assert abs(g[0][0] - h[0][0]) < 5e-6
        assert np.abs(g[0][2] - R @ h[0][2]).max() < 5e-3
