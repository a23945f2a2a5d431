"""Known-answer tests of the float64 oracle's physics: nothing in the reference pins MuJoCo's
arithmetic, so the oracle is checked against closed forms and invariants."""
import numpy as np
import pytest


def _free_env(M):
    from oracle.oracle import OracleEnv
    o = OracleEnv(M)
    for k in ("disable_contact", "disable_limit", "disable_fluid", "disable_actuation"):
        o.option(k, 1)
    return o


def test_free_fall_com_acceleration_is_gravity(models):
    M = models["walk_imitation"]
    o = _free_env(M)
    o.d("qpos")[:] = M["qpos_reset"]; o.d("qvel")[:] = 0
    o.d("qpos")[2] += 1.0
    o.forward()
    # passive joint springs move the limbs (so the root's own acceleration is not -g), but the CoM of the
    # walker tree falls with g: check its displacement over 5 steps
    nb = int(M["nbody"]); mass = M["body_mass"]
    walker = np.asarray(M["body_rootid"]) == int(M["root_body"])
    def com():
        return (o.d("xipos").reshape(nb, 3)[walker] * mass[walker, None]).sum(0) / mass[walker].sum()
    c0 = com(); dt = float(M["opt_timestep"])
    o.step(5)
    c1 = com()
    expect = 981.0 * dt * dt * 15  # semi-implicit Euler: sum_{k=1..5} k dt^2 g
    assert abs((c1[2] - c0[2]) + expect) < 1e-5 * expect
    assert abs(c1[0] - c0[0]) < 1e-5 * expect and abs(c1[1] - c0[1]) < 1e-5 * expect


def test_mass_matrix_is_symmetric_positive_and_ldl_reconstructs(models):
    M = models["walk_imitation"]
    from oracle.oracle import OracleEnv
    o = OracleEnv(M)
    rng = np.random.RandomState(0)
    q = M["qpos_reset"].copy()
    q[7:109] += rng.uniform(-0.2, 0.2, 102)
    o.d("qpos")[:] = q
    o.forward()
    nv = int(M["nv"])
    A = np.tril(o.d("qM").reshape(nv, nv)); A = A + np.tril(A, -1).T
    assert np.linalg.eigvalsh(A).min() > 0
    LD = np.tril(o.d("qLD").reshape(nv, nv))
    L = np.tril(LD, -1) + np.eye(nv); D = np.diag(np.diag(LD))
    np.testing.assert_allclose(L.T @ D @ L, A, atol=1e-12 * np.abs(A).max())
    total = M["body_mass"][np.asarray(M["body_rootid"]) == int(M["root_body"])].sum()
    assert abs(A[0, 0] - total) < 1e-12  # translational block of the free root = total walker mass


def test_actuator_filter_step_response(models):
    M = models["walk_on_ball"]
    from oracle.oracle import OracleEnv
    o = OracleEnv(M)
    o.d("ctrl")[:] = M["act_ctrlrange"][:, 1]
    n = 50
    o.step(n)
    dt, tau = float(M["opt_timestep"]), M["act_dynprm"]
    np.testing.assert_allclose(o.d("act"), M["act_ctrlrange"][:, 1] * (1 - (1 - dt / tau) ** n), rtol=1e-10)


def test_resting_contact_force_balances_weight(models):
    """Fly standing on the floor, settled: sum of normal contact forces = weight (soft-constraint equilibrium)."""
    M = models["walk_imitation"]
    from oracle.oracle import OracleEnv
    o = OracleEnv(M)
    o.option("disable_fluid", 1)
    o.d("qpos")[:] = M["qpos_reset"]
    o.step(1500)
    walker = np.asarray(M["body_rootid"]) == int(M["root_body"])
    weight = 981.0 * M["body_mass"][walker].sum()
    f = 0.0
    for c in o.contacts():
        if c["efc_address"] >= 0 and c["geom1"] == 0:
            f += o.d("efc_force")[c["efc_address"]] * c["frame"][2]
    assert np.abs(o.d("qvel")[:6]).max() < 5.0
    assert abs(f - weight) / weight < 0.05


def test_friction_cone_is_respected(models):
    M = models["walk_imitation"]
    from oracle.oracle import OracleEnv
    o = OracleEnv(M)
    o.d("qpos")[:] = M["qpos_reset"]
    o.d("qvel")[0] = 3.0
    o.step(30)
    fr = o.d("efc_force")
    n = 0
    for c in o.contacts():
        if c["efc_address"] >= 0 and c["dim"] == 3:
            a = c["efc_address"]
            assert np.hypot(fr[a + 1], fr[a + 2]) <= c["friction"][0] * fr[a] * (1 + 1e-6) + 1e-9
            n += 1
    assert n > 0


def test_momentum_drift_of_internal_forces_vanishes_linearly_with_the_timestep(models):
    """Free flight with actuators, joint limits, springs and dampers on (all internal forces; contacts and fluid off):
    gravity changes the linear momentum by exactly m g t and nothing changes the angular momentum about the CoM.  The
    semi-implicit Euler step keeps that only to first order, so the drift over a fixed time must halve when the
    timestep halves (a spurious external force - an inconsistent mass matrix, bias force or actuator moment - would give
    a drift that does not vanish).  Checks CRB / RNE / actuation / limits / implicit damping against one another."""
    from oracle.oracle import OracleEnv
    M = models["flight_imitation"]
    nb, root = int(M["nbody"]), int(M["root_body"])
    mass, inertia = np.asarray(M["body_mass"]), np.asarray(M["body_inertia"]).reshape(nb, 3)
    iquat = np.asarray(M["body_iquat"]).reshape(nb, 4)
    idx = np.where(np.asarray(M["body_rootid"]) == root)[0]      # the fly without the ghost
    mt = mass[idx].sum()
    nvw = int(M["ghost_dofadr"])

    def q2m(q):
        w, x, y, z = q
        return np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)],
                         [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])

    def qmul(a, b):
        return np.array([a[0]*b[0]-a[1]*b[1]-a[2]*b[2]-a[3]*b[3], a[0]*b[1]+a[1]*b[0]+a[2]*b[3]-a[3]*b[2],
                         a[0]*b[2]-a[1]*b[3]+a[2]*b[0]+a[3]*b[1], a[0]*b[3]+a[1]*b[2]-a[2]*b[1]+a[3]*b[0]])

    def momentum(o):
        xipos, cvel, xq = o.d("xipos").reshape(nb, 3), o.d("cvel").reshape(nb, 6), o.d("xquat").reshape(nb, 4)
        sc = o.d("subtree_com").reshape(nb, 3)[root]
        c = (xipos[idx] * mass[idx, None]).sum(0) / mt
        v = {b: cvel[b, 3:] + np.cross(cvel[b, :3], xipos[b] - sc) for b in idx}   # cvel is taken at the root's subtree CoM
        P = sum(mass[b] * v[b] for b in idx)
        L = np.zeros(3)
        for b in idx:
            R = q2m(qmul(xq[b], iquat[b]))
            L += mass[b] * np.cross(xipos[b] - c, v[b] - P / mt) + R @ np.diag(inertia[b]) @ R.T @ cvel[b, :3]
        return P, L

    drift = []
    for sub in (1, 2, 4):
        o = OracleEnv(M)
        o.option("disable_contact", 1); o.option("disable_fluid", 1)
        o.option("timestep", 5e-5 / sub)
        rng = np.random.RandomState(0)
        o.d("qpos")[:] = M["qpos_reset"]; o.d("qpos")[2] += 1.0
        qv = o.d("qvel"); qv[:] = 0; qv[6:nvw] = rng.normal(scale=20, size=nvw - 6)
        o.d("ctrl")[:] = rng.uniform(-1, 1, int(M["nu"]))
        o.forward()
        P0, L0 = momentum(o)
        o.step(20 * sub)
        P1, L1 = momentum(o)
        dP = P1 - P0 + np.array([0, 0, 981.0 * mt * 20 * 5e-5])
        drift.append((np.linalg.norm(dP), np.linalg.norm(L1 - L0)))
    for k in range(2):
        assert 1.9 < drift[k][0] / drift[k + 1][0] < 2.1 and 1.9 < drift[k][1] / drift[k + 1][1] < 2.1, drift
