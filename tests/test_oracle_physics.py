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
