"""Known-answer tests of the float64 oracle, second batch (SURVEY.md 8c items 2, 3, 4, 5b, 6b, 8, 9, 12 + touch containment,
part masses, the reset pass, and the reference quaternion fixture applied to the oracle's and the engine's helpers).

MuJoCo is not installable here, so nothing can pin the oracle's physics against MuJoCo's own numbers.  What these tests do
instead: each stage is compared with a closed form (or a finite-difference identity) that is written here independently in
numpy from the published formulation - a wrong coefficient, sign, frame or clamp in the oracle makes the test fail."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _oracle(M, *off):
    from oracle.oracle import OracleEnv
    o = OracleEnv(M)
    for k in off:
        o.option("disable_" + k, 1)
    return o


def _hinge_dofs(M):
    return [int(M["jnt_dofadr"][j]) for j in range(int(M["njnt"])) if M["jnt_type"][j] == 3 and M["jnt_dofadr"][j] < M["ghost_dofadr"]]


def _posed(M, seed=0, scale=0.25):
    rng = np.random.RandomState(seed)
    q = np.array(M["qpos_reset"], np.float64)
    for j in range(int(M["njnt"])):
        if M["jnt_type"][j] == 3:
            lo, hi = M["jnt_range"][j]
            a = int(M["jnt_qposadr"][j])
            q[a] = np.clip(q[a] + rng.uniform(-scale, scale) * (hi - lo), lo, hi)
    ra = int(M["root_qposadr"])
    q[ra + 2] += 1.0
    qq = q[ra + 3:ra + 7] + rng.normal(0, 0.2, 4)
    q[ra + 3:ra + 7] = qq / np.linalg.norm(qq)
    return q


# --------------------------------------------------------------------------------------------------- reset pass
def test_first_observation_comes_from_an_unactuated_forward_pass(models, walk_clip):
    """dm_control's Physics.after_reset() runs mj_forward with actuation disabled: at a pose where the position servos
    (force = kp (ctrl - q), ctrl = 0) would pull - the first row of a walking clip - the sensors of the FIRST time step must be
    those of the unactuated model."""
    from tests.conftest import clip_arrays
    M = models["walk_imitation"]
    o = _oracle(M)
    o.set_clip(*clip_arrays(M, walk_clip[0]))
    o.env_reset()
    first = o.d("sensordata").copy()
    assert np.abs(o.d("qfrc_actuator")).max() == 0.0 and np.array_equal(o.d("qacc_warmstart"), o.d("qacc"))
    q = o.d("qpos").copy()
    un, ac = _oracle(M, "actuation"), _oracle(M)
    for e in (un, ac):
        e.d("qpos")[:] = q
        e.forward()
    np.testing.assert_allclose(first, un.d("sensordata"), rtol=0, atol=1e-12 * np.abs(first).max())
    assert np.abs(ac.d("qfrc_actuator")).max() > 1e-6                      # the servos do pull at this pose ...
    assert np.abs(ac.d("sensordata") - first).max() > 1e-3               # ... and the accelerometer / force sensors see it


# --------------------------------------------------------------------------------------------------- m4: mass matrix
def test_mass_matrix_equals_the_kinetic_energy_form_of_finite_difference_jacobians(models):
    """M_ij = sum_b m_b Jv_bi . Jv_bj + Jw_bi^T I_b Jw_bj with the body Jacobians taken by central differences of the
    kinematics (CoM positions and inertial-frame orientations): pins CRB against nothing but kinematics + body inertias."""
    M = models["walk_imitation"]
    o = _oracle(M, "contact", "limit", "fluid", "actuation")
    q = _posed(M, 1)
    nb, nv = int(M["nbody"]), int(M["nv"])
    hd = _hinge_dofs(M)
    qadr = {int(M["jnt_dofadr"][j]): int(M["jnt_qposadr"][j]) for j in range(int(M["njnt"]))}
    eps = 1e-6

    def kin(qq):
        o.d("qpos")[:] = qq
        o.forward()
        return o.d("xipos").reshape(nb, 3).copy(), o.d("ximat").reshape(nb, 3, 3).copy()
    Jv, Jw = np.zeros((nb, 3, len(hd))), np.zeros((nb, 3, len(hd)))
    for c, d in enumerate(hd):
        qp, qm = q.copy(), q.copy()
        qp[qadr[d]] += eps; qm[qadr[d]] -= eps
        (xp, Rp), (xm, Rm) = kin(qp), kin(qm)
        Jv[:, :, c] = (xp - xm) / (2 * eps)
        W = np.einsum("bij,bkj->bik", Rp, Rm)                       # R+ R-^T = I + 2 eps [w]x
        Jw[:, :, c] = np.stack([W[:, 2, 1] - W[:, 1, 2], W[:, 0, 2] - W[:, 2, 0], W[:, 1, 0] - W[:, 0, 1]], 1) / (4 * eps)
    x0, R0 = kin(q)
    A = np.tril(o.d("qM").reshape(nv, nv)); A = A + np.tril(A, -1).T
    mass, inertia = np.asarray(M["body_mass"]), np.asarray(M["body_inertia"]).reshape(nb, 3)
    Iw = np.einsum("bij,bj,bkj->bik", R0, inertia, R0)
    Mfd = np.einsum("b,bxi,bxj->ij", mass, Jv, Jv) + np.einsum("bxi,bxy,byj->ij", Jw, Iw, Jw)
    Mfd[np.diag_indices(len(hd))] += np.asarray(M["dof_armature"])[hd]
    Ah = A[np.ix_(hd, hd)]
    scale = np.sqrt(np.outer(np.diag(Ah), np.diag(Ah)))
    assert (np.abs(Mfd - Ah) / scale).max() < 1e-6


# --------------------------------------------------------------------------------------------------- m8: bias forces
def test_bias_forces_equal_the_lagrangian_terms_of_the_mass_matrix(models):
    """qfrc_bias_i = sum_jk (dM_ij/dq_k - 1/2 dM_jk/dq_i) v_j v_k + dV/dq_i for hinge coordinates (root at rest), with the
    derivatives of the oracle's own mass matrix (pinned above) and of the potential g sum m z by central differences."""
    M = models["walk_imitation"]
    o = _oracle(M, "contact", "limit", "fluid", "actuation")
    q = _posed(M, 2)
    nb, nv = int(M["nbody"]), int(M["nv"])
    hd = _hinge_dofs(M)
    qadr = {int(M["jnt_dofadr"][j]): int(M["jnt_qposadr"][j]) for j in range(int(M["njnt"]))}
    rng = np.random.RandomState(3)
    S = sorted(rng.choice(hd, 14, replace=False).tolist())            # moving coordinates = coordinates whose equation is checked
    v = np.zeros(nv); v[S] = rng.normal(0, 400.0, len(S))   # rad/s: fast enough for the velocity terms to match gravity's
    mass = np.asarray(M["body_mass"])
    walker = np.asarray(M["body_rootid"]) == int(M["root_body"])

    def mv(qq):
        o.d("qpos")[:] = qq; o.d("qvel")[:] = 0
        o.forward()
        A = np.tril(o.d("qM").reshape(nv, nv)); A = A + np.tril(A, -1).T
        V = 981.0 * (mass[walker] * o.d("xipos").reshape(nb, 3)[walker, 2]).sum()
        return A[np.ix_(S, S)], V
    eps = 1e-5
    dM, dV = [], []
    for d in S:
        qp, qm = q.copy(), q.copy()
        qp[qadr[d]] += eps; qm[qadr[d]] -= eps
        (Ap, Vp), (Am, Vm) = mv(qp), mv(qm)
        dM.append((Ap - Am) / (2 * eps)); dV.append((Vp - Vm) / (2 * eps))
    dM, dV, vs = np.array(dM), np.array(dV), v[S]                      # dM[k, i, j] = dM_ij / dq_k
    expect = np.einsum("kij,j,k->i", dM, vs, vs) - 0.5 * np.einsum("ijk,j,k->i", dM, vs, vs) + dV
    o.d("qpos")[:] = q; o.d("qvel")[:] = v
    o.forward()
    got = o.d("qfrc_bias")[S]
    assert np.abs(got - expect).max() < 1e-6 * np.abs(expect).max()
    # the limbs are so light that gravity dominates the sum: hold the error against the velocity terms alone as well
    assert np.abs(got - expect).max() < 1e-4 * np.abs(expect - dV).max()


# --------------------------------------------------------------------------------------------------- m13: Euler step
def test_euler_step_with_implicit_joint_damping_closed_form(models):
    """qvel' = qvel + dt (M + dt diag(damping))^-1 (qfrc_smooth + qfrc_constraint); hinge qpos' = qpos + dt qvel'; free-joint
    quaternion' = q * exp(dt w' / 2) with w' in the body frame; act' = act + dt act_dot; warm start' = qacc."""
    M = models["walk_imitation"]
    o = _oracle(M)
    rng = np.random.RandomState(4)
    nv, dt = int(M["nv"]), float(M["opt_timestep"])
    q = _posed(M, 4); q[int(M["root_qposadr"]) + 2] -= 1.0              # back on the floor: contacts and limits take part
    o.d("qpos")[:] = q; o.d("qvel")[:] = rng.normal(0, 2.0, nv)
    o.d("ctrl")[:] = rng.uniform(M["act_ctrlrange"][:, 0], M["act_ctrlrange"][:, 1])
    o.d("act")[:] = rng.uniform(-0.2, 0.2, int(M["na"]))
    o.forward()
    A = np.tril(o.d("qM").reshape(nv, nv)); A = A + np.tril(A, -1).T
    rhs = o.d("qfrc_smooth") + o.d("qfrc_constraint")
    assert np.abs(o.d("qfrc_constraint")).max() > 0
    v0, q0, act0, adot, qacc = o.d("qvel").copy(), o.d("qpos").copy(), o.d("act").copy(), o.d("act_dot").copy(), o.d("qacc").copy()
    v1 = v0 + dt * np.linalg.solve(A + dt * np.diag(M["dof_damping"]), rhs)
    o.step(1)
    np.testing.assert_allclose(o.d("qvel"), v1, rtol=0, atol=1e-10 * np.abs(v1).max())
    np.testing.assert_allclose(o.d("act"), act0 + dt * adot, rtol=0, atol=1e-14)
    np.testing.assert_allclose(o.d("qacc_warmstart"), qacc, rtol=0, atol=1e-9 * np.abs(qacc).max())
    for j in range(int(M["njnt"])):
        qa, da = int(M["jnt_qposadr"][j]), int(M["jnt_dofadr"][j])
        if M["jnt_type"][j] == 3:
            assert abs(o.d("qpos")[qa] - (q0[qa] + dt * v1[da])) < 1e-13
        elif M["jnt_type"][j] == 0:
            np.testing.assert_allclose(o.d("qpos")[qa:qa + 3], q0[qa:qa + 3] + dt * v1[da:da + 3], atol=1e-13)
            w = v1[da + 3:da + 6]; th = np.linalg.norm(w) * dt
            dq = np.r_[np.cos(th / 2), np.sin(th / 2) * w / np.linalg.norm(w)]
            a = q0[qa + 3:qa + 7]
            e = np.r_[a[0] * dq[0] - a[1:] @ dq[1:], a[0] * dq[1:] + dq[0] * a[1:] + np.cross(a[1:], dq[1:])]
            np.testing.assert_allclose(o.d("qpos")[qa + 3:qa + 7], e / np.linalg.norm(e), atol=1e-13)


# --------------------------------------------------------------------------------------------------- m6: constraint rows
def _impedance(solimp, x):
    d0, dw, width, mid, power = solimp
    x = min(1.0, abs(x) / width)
    y = x ** power / mid ** (power - 1) if x <= mid else 1 - (1 - x) ** power / (1 - mid) ** (power - 1)
    return d0 + y * (dw - d0)


def _check_contact_rows(M, o, want_refsafe):
    dt, bw = float(M["opt_timestep"]), np.asarray(M["body_invweight0"]).reshape(-1, 2)
    KBIP, R, aref, vel = o.d("efc_KBIP").reshape(-1, 4), o.d("efc_R"), o.d("efc_aref"), o.d("efc_vel")
    n = clamped = 0
    for c in o.contacts():
        r = c["efc_address"]
        if r < 0:
            assert c["dist"] >= M["pair_margin"][c["pair"]] - M["pair_gap"][c["pair"]] - 1e-15      # gap band: detected, no row
            continue
        p = c["pair"]
        tau, zeta = M["pair_solref"][p]
        clamped += tau < 2 * dt
        tau = max(tau, 2 * dt)                                                                       # refsafe
        solimp = M["pair_solimp"][p]
        incl = M["pair_margin"][p] - M["pair_gap"][p]
        imp = _impedance(solimp, c["dist"] - incl)
        k, b = 1.0 / (solimp[1] ** 2 * tau ** 2 * zeta ** 2), 2.0 / (solimp[1] * tau)
        np.testing.assert_allclose(KBIP[r, :3], [k, b, imp], rtol=1e-12)
        assert abs(aref[r] - (-b * vel[r] - k * imp * (c["dist"] - incl))) <= 1e-10 * max(1.0, abs(aref[r]))
        b1, b2 = int(M["geom_bodyid"][c["geom1"]]), int(M["geom_bodyid"][c["geom2"]])
        R0 = (1 - imp) / imp * (bw[b1, 0] + bw[b2, 0])
        assert abs(R[r] - R0) <= 1e-12 * R0
        for kk in range(1, c["dim"]):                                                                # elliptic cone, condim 3
            assert abs(R[r + kk] - R0 / float(M["opt_impratio"])) <= 1e-12 * R0
            assert abs(aref[r + kk] + b * vel[r + kk]) <= 1e-10 * max(1.0, abs(aref[r + kk]))          # friction rows: no position term
        assert abs(c["mu"] - c["friction"][0] * np.sqrt(R[r + 1] / R[r]) if c["dim"] > 1 else 0.0) < 1e-12
        n += 1
    assert n > 0 and (clamped > 0) == want_refsafe
    return n


def test_contact_row_parameters_closed_form(models):
    """Settled standing fly: six claw-floor contacts (solref 0.0006 = mix of claw 0.0002 and floor 0.001, above 2 dt) and the
    fly-fly contacts it rests in (haustellum-head, femur_T3-abdomen: solref 0.0002 < 2 dt = 0.0004, so the refsafe clamp
    bites; solimp 0.95 0.99 0.01 throughout): K, B, impedance, aref, the regulariser R and the cone's friction rows of every
    row against the soft-constraint formulas."""
    M = models["walk_imitation"]
    o = _oracle(M, "fluid")
    o.d("qpos")[:] = M["qpos_reset"]
    o.step(1500)
    o.forward()                                    # rows, velocities and reference accelerations of one and the same state
    assert _check_contact_rows(M, o, want_refsafe=True) >= 8
    assert sum(c["geom1"] == 0 and c["efc_address"] >= 0 for c in o.contacts()) == 6


def test_elliptic_cone_force_law_row_by_row(models):
    """Forces of the converged Newton solve (before the noslip pass) from the residual jar = J qacc - aref, zone by zone
    (MuJoCo's elliptic cone: satisfied N >= mu T -> 0; bottom mu N + T <= 0 -> -D jar per row; middle -> the cone surface
    with D_m = D_0 / (mu^2 (1 + mu^2))), written out here in numpy; limit and condim-1 rows: -D min(jar, 0).
    Together with test_contact_row_parameters_closed_form (aref and R from solref / solimp / refsafe / diagApprox) this is the
    soft-constraint equilibrium: at rest jar = -aref = -k imp depth, so a frictionless row carries k imp depth / R.  (The
    settled fly itself never is exactly at rest - its splayed legs sit on their friction cones and creep at ~0.03 rad/s, and
    b J qvel is then a third of k imp depth on the 1e-5 cm deep fly-fly contacts - so the identity is checked per row.)"""
    M = models["walk_imitation"]
    o = _oracle(M, "fluid", "noslip")
    rng = np.random.RandomState(8)
    o.d("qpos")[:] = M["qpos_reset"]
    o.step(400)
    o.d("qvel")[:int(M["ghost_dofadr"])] += rng.normal(0, 0.5, int(M["ghost_dofadr"]))     # load the friction rows
    o.forward()
    nv = int(M["nv"])
    J, aref, R, f = o.d("efc_J").reshape(-1, nv), o.d("efc_aref"), o.d("efc_R"), o.d("efc_force")
    jar = J @ o.d("qacc") - aref[:J.shape[0]]
    zones = set()
    scale = np.abs(f).max()
    for c in o.contacts():
        r = c["efc_address"]
        if r < 0:
            continue
        if c["dim"] == 1:
            assert abs(f[r] + min(jar[r], 0.0) / R[r]) < 1e-6 * scale
            continue
        mu, f0 = c["mu"], c["friction"][0]
        fr = np.array([mu, f0, f0])
        U = jar[r:r + 3] * fr
        N, T = U[0], np.hypot(U[1], U[2])
        if N >= mu * T:
            want = np.zeros(3); zones.add("satisfied")
        elif mu * N + T <= 0:
            want = -jar[r:r + 3] / R[r:r + 3]; zones.add("bottom")
        else:
            Dm = (1.0 / R[r]) / (mu * mu * (1 + mu * mu))
            NT = N - mu * T
            want = np.r_[-Dm * NT * mu, Dm * NT * mu / T * U[1:] * fr[1:]]; zones.add("middle")
        # the solver stops at tolerance 1e-8 (scaled): forces are those of its last iterate
        assert np.abs(f[r:r + 3] - want).max() < 1e-6 * scale, (r, f[r:r + 3], want)
    assert "middle" in zones and len(zones) >= 2, zones


# --------------------------------------------------------------------------------------------------- m11: noslip
def test_noslip_pass_stops_the_slip_the_soft_friction_rows_allow(models):
    """Fly held to the floor by its six adhesion actuators (friction reserve on every claw), nudged sideways: the regularised
    friction rows let the claws slip, the noslip pass (3 iterations, fruitfly.xml:4; friction rows re-solved without
    regularisation) holds them - the tangential contact velocity after the nudge is several times lower (measured 3.3x)."""
    M = models["walk_imitation"]
    adh = [i for i, n in enumerate(M["_names_act"]) if "adhere" in n]
    nv = int(M["nv"])
    slip = {}
    for mode in ("noslip", "soft"):
        o = _oracle(M, "fluid", *(["noslip"] if mode == "soft" else []))
        o.d("qpos")[:] = M["qpos_reset"]
        o.d("ctrl")[adh] = 1.0
        o.step(1500)
        o.d("qvel")[1] += 0.3
        o.step(5)
        J = o.d("efc_J").reshape(-1, nv)
        s = [np.hypot(J[c["efc_address"] + 1] @ o.d("qvel"), J[c["efc_address"] + 2] @ o.d("qvel"))
             for c in o.contacts() if c["efc_address"] >= 0 and c["dim"] == 3 and c["geom1"] == 0]
        assert len(s) == 6
        slip[mode] = float(np.max(s))
    assert slip["noslip"] < 0.5 * slip["soft"], slip


# --------------------------------------------------------------------------------------------------- m3: adhesion
def test_adhesion_pulls_with_gain_times_activation_also_through_a_gap_band_contact(models):
    """Adhesion actuator on a claw: moment = -(mean normal Jacobian of the body's contacts), force = gain * act.  For a floor
    contact the normal is +z and the normal Jacobian of the root's z translation is 1, so qfrc_actuator[root z] = -gain act
    per adhering claw - with the claw pressed into the floor, and with the claw hovering inside the margin-gap band
    (0 <= dist < margin = gap = 0.0005: detected, no constraint row, still seen by the transmission)."""
    M = models["walk_imitation"]
    names = list(M["_names_act"])
    adh = [i for i, n in enumerate(names) if "adhere" in n]
    assert len(adh) == 6 and all(M["act_trntype"][a] == 2 for a in adh)
    o = _oracle(M, "fluid")
    o.d("qpos")[:] = M["qpos_reset"]
    o.step(1500)
    rz = int(M["root_dofadr"]) + 2
    act = np.zeros(int(M["na"]))
    for a in adh:
        act[int(M["act_actadr"][a])] = 0.7
    zero_ctrl = o.d("ctrl").copy()
    o.d("act")[:] = 0; o.forward(); base = o.d("qfrc_actuator")[rz]
    o.d("act")[:] = act; o.forward()
    pull = o.d("qfrc_actuator")[rz] - base
    expect = -sum(M["act_gainprm"][a] * 0.7 for a in adh)
    assert abs(pull - expect) < 1e-9 * abs(expect)
    nv = int(M["nv"])
    mom = o.d("act_moment").reshape(-1, nv)
    J = o.d("efc_J").reshape(-1, nv)
    for a in adh:                               # moment = - normal row of the claw's one floor contact
        body = int(M["act_trnid"][a])
        rows = [c["efc_address"] for c in o.contacts() if int(M["geom_bodyid"][c["geom2"]]) == body and c["efc_address"] >= 0]
        assert len(rows) == 1
        np.testing.assert_allclose(mom[a], -J[rows[0]], atol=1e-12)
    # lift the whole fly until the deepest claw is 0.0002 above the floor: every claw within 0.0005 is in the gap band
    dists = [c["dist"] for c in o.contacts() if c["geom1"] == 0]
    q = o.d("qpos").copy(); q[int(M["root_qposadr"]) + 2] += -min(dists) + 0.0002
    o.d("qpos")[:] = q; o.d("qvel")[:] = 0; o.forward()
    band = [c for c in o.contacts() if c["geom1"] == 0]
    assert band and all(c["efc_address"] < 0 and 0 <= c["dist"] < 0.0005 for c in band)
    bodies = {int(M["geom_bodyid"][c["geom2"]]) for c in band}
    expect = -sum(M["act_gainprm"][a] * 0.7 for a in adh if int(M["act_trnid"][a]) in bodies)
    o.d("act")[:] = 0; o.forward(); base = o.d("qfrc_actuator")[rz]
    o.d("act")[:] = act; o.forward()
    assert abs((o.d("qfrc_actuator")[rz] - base) - expect) < 1e-9 * abs(expect) and expect < 0
    assert np.array_equal(zero_ctrl, o.d("ctrl"))


# --------------------------------------------------------------------------------------------------- m12: touch
def test_touch_sensors_read_the_normal_force_of_the_claw_contacts(models):
    """Touch site (capsule 0.00219) contains the claw geom (0.00199, fruitfly.xml:496): every claw-floor contact point lies
    inside it, so each touch sensor reads exactly the normal force of its claw's contact."""
    M = models["walk_imitation"]
    o = _oracle(M, "fluid")
    o.d("qpos")[:] = M["qpos_reset"]
    o.step(800)
    o.forward()
    names = list(M["_names_sensor"])
    f = o.d("efc_force")
    n = 0
    for s, name in enumerate(names):
        if "touch" not in name:
            continue
        body = int(M["site_bodyid"][int(M["sensor_site"][s])])
        want = sum(f[c["efc_address"]] for c in o.contacts()
                   if c["efc_address"] >= 0 and body in (int(M["geom_bodyid"][c["geom1"]]), int(M["geom_bodyid"][c["geom2"]])))
        got = o.d("sensordata")[int(M["sensor_adr"][s])]
        assert want > 0 and abs(got - want) < 1e-12 * want
        n += 1
    assert n == 6


# --------------------------------------------------------------------------------------------------- m7: ellipsoid fluid
def _wing_force(M, qvel_root):
    """World force and torque (about the root position) that the ellipsoid model puts on the wings for a rigid-body root
    twist: difference of qfrc_passive with the model's ellipsoid coefficients on and off."""
    out = []
    for scale in (1.0, 0.0):
        Md = dict(M)
        coef = np.array(M["fgeom_coef"], np.float64).reshape(-1, 12).copy()
        coef[:, 0] *= scale
        Md["fgeom_coef"] = coef.reshape(np.asarray(M["fgeom_coef"]).shape)
        o = _oracle(Md, "contact", "limit", "actuation")
        o.d("qpos")[:] = M["qpos_reset"]; o.d("qvel")[:] = 0
        o.d("qvel")[:6] = qvel_root
        o.forward()
        out.append((o.d("qfrc_passive")[:6].copy(), o.d("xmat").reshape(-1, 3, 3)[int(M["root_body"])].copy(), o))
    return out[0][0] - out[1][0], out[0][1], out[0][2]


def test_ellipsoid_fluid_forces_closed_form(models):
    """wing_*_fluid ellipsoids (fluidcoef blunt 1.0? no: [1.0, 0.5, 1.5, 1.7, 1.0] = blunt drag, slender drag, angular drag,
    Kutta lift, Magnus lift; tasks/base.py:309-313).  Whole fly translating rigidly with velocity v along a principal axis
    of a wing ellipsoid: no circulation (n x v = 0), no Magnus, no added-mass term (p x v = 0), what remains is
    F = -(3 pi beta eqD + rho |v| (A_proj C_blunt + C_slender (A_max - A_proj))) v.  Oblique in-plane velocity adds the
    Kutta force C_K rho cos(alpha) A_proj (n x v) x v.  Checked on the summed force of the two (mirror-image) wings."""
    M = models["flight_imitation"]
    assert int(M["nfluidgeom"]) == 2
    rho, beta = float(M["opt_density"]), float(M["opt_viscosity"])
    coef = np.asarray(M["fgeom_coef"]).reshape(2, 12)
    size = np.asarray(M["fgeom_size"]).reshape(2, 3)
    assert np.allclose(coef[:, 1:6], [1.0, 0.5, 1.5, 1.7, 1.0]) and np.allclose(coef[:, 0], 1.0)
    o0 = _oracle(M, "contact", "limit", "actuation")
    o0.d("qpos")[:] = M["qpos_reset"]; o0.forward()
    nb = int(M["nbody"])
    xmat, xpos = o0.d("xmat").reshape(nb, 3, 3), o0.d("xpos").reshape(nb, 3)
    gmat = np.asarray(M["fgeom_quat"]).reshape(2, 4)

    def q2m(q):
        w, x, y, z = q
        return np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)],
                         [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])
    Rg = [xmat[int(M["fgeom_bodyid"][g])] @ q2m(gmat[g]) for g in range(2)]     # world orientation of each ellipsoid

    def closed_form(g, v_world, kutta=True):
        s = size[g]
        v = Rg[g].T @ v_world
        blunt, slender, angc, ck, cm = coef[g, 1:6]
        speed = np.linalg.norm(v)
        dmax, dmin = s.max(), s.min(); dmid = s.sum() - dmax - dmin
        A_max = np.pi * dmax * dmid
        bc, ca, ab = s[1] * s[2], s[2] * s[0], s[0] * s[1]
        pd = bc ** 4 * v[0] ** 2 + ca ** 4 * v[1] ** 2 + ab ** 4 * v[2] ** 2
        pn = bc ** 2 * v[0] ** 2 + ca ** 2 * v[1] ** 2 + ab ** 2 * v[2] ** 2
        A_proj = np.pi * np.sqrt(pd / pn)
        eqD = 2.0 / 3.0 * s.sum()
        F = -(3 * np.pi * beta * eqD + rho * speed * (A_proj * blunt + slender * (A_max - A_proj))) * v
        if kutta:
            n = np.array([bc ** 2 * v[0], ca ** 2 * v[1], ab ** 2 * v[2]])
            cosa = pn / (speed * pd)
            F = F + np.cross(ck * rho * cosa * A_proj * np.cross(n, v), v)
        # added mass: p x w and (p x v + L x w) vanish without rotation except p x v (torque) - forces only here
        return Rg[g] @ F
    speed = 40.0
    for axis in range(3):                                   # along a principal axis of wing 0 (= mirrored axis of wing 1)
        v_world = Rg[0][:, axis] * speed
        got, _, _ = _wing_force(M, np.r_[v_world, 0, 0, 0])
        want = closed_form(0, v_world) + closed_form(1, v_world)
        assert np.abs(got[:3] - want).max() < 1e-9 * np.abs(want).max(), (axis, got[:3], want)
    v_world = (Rg[0][:, 0] * 0.8 + Rg[0][:, 2] * 0.6) * speed   # oblique: angle of attack, circulation, Kutta lift
    got, _, _ = _wing_force(M, np.r_[v_world, 0, 0, 0])
    want = closed_form(0, v_world) + closed_form(1, v_world)
    drag_only = closed_form(0, v_world, False) + closed_form(1, v_world, False)
    assert np.abs(got[:3] - want).max() < 1e-9 * np.abs(want).max()
    assert np.abs(want - drag_only).max() > 0.05 * np.abs(want).max()       # the lift term is a real part of it


# --------------------------------------------------------------------------------------------------- m15: meshes, masses
def _cube():
    v = np.array([[x, y, z] for x in (-0.5, 0.5) for y in (-0.5, 0.5) for z in (-0.5, 0.5)], float)
    f = [[0, 1, 3], [0, 3, 2], [4, 6, 7], [4, 7, 5], [0, 4, 5], [0, 5, 1], [2, 3, 7], [2, 7, 6], [0, 2, 6], [0, 6, 4], [1, 5, 7], [1, 7, 3]]
    return v, np.array(f)


def _icosphere(level=4):
    t = (1 + 5 ** 0.5) / 2
    v = [[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]]
    f = [[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
         [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]]
    v = [np.array(p, float) / np.linalg.norm(p) for p in v]
    for _ in range(level):
        cache, nf = {}, []
        def mid(a, b):
            k = (min(a, b), max(a, b))
            if k not in cache:
                p = v[a] + v[b]; v.append(p / np.linalg.norm(p)); cache[k] = len(v) - 1
            return cache[k]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        f = nf
    return np.array(v), np.array(f)


def test_mesh_mass_properties_of_a_cube_and_a_sphere():
    """Mesh volume, centre of mass and inertia (host compiler, m15) against the analytic solids, off-centre and scaled."""
    from flybody_b200.meshes import mesh_mass_properties
    v, f = _cube()
    for mode in ("exact", "legacy"):
        vol, com, I = mesh_mass_properties(v + [0.3, -0.2, 0.1], f, scale=(2.0, 1.0, 0.5), mode=mode)[:3]
        assert abs(vol - 1.0) < 1e-12
        np.testing.assert_allclose(com, [0.6, -0.2, 0.05], atol=1e-12)
        a, b, c = 2.0, 1.0, 0.5
        np.testing.assert_allclose(np.diag(I) if np.ndim(I) == 2 else I[:3], np.array([b*b + c*c, a*a + c*c, a*a + b*b]) / 12.0, rtol=1e-10)
    v, f = _icosphere(4)
    vol, com, I = mesh_mass_properties(0.7 * v, f, mode="exact")[:3]
    assert abs(vol - 4 / 3 * np.pi * 0.7 ** 3) < 4e-3 * vol and np.abs(com).max() < 1e-12     # 5120 flat faces inside the sphere
    Id = np.diag(I) if np.ndim(I) == 2 else I[:3]
    np.testing.assert_allclose(Id / vol, 0.4 * 0.7 ** 2, rtol=4e-3)


def test_compiled_part_masses_match_the_reference_empirical_values(models):
    """vnl_ray/fruitfly/build_fruitfly/make_fruitfly.py:24,1240-1254: _MASS (milligrams) head 0.15, thorax 0.34, abdomen 0.38,
    leg 0.0162, wing 0.008, as that script prints them: thorax = body mass, head / abdomen / leg / wing = subtree masses
    (model units are grams).  The reference model reproduces them to the script's printed precision; the head subtree carries
    the stand-in for the missing head_red mesh and still meets its 0.15."""
    M = models["walk_imitation"]            # legs articulated, wings folded but present; the root body "walker/" is the thorax
    names = [n.split("/")[-1] for n in M["_names_body"]]
    sub = np.asarray(M["body_subtreemass"]) * 1e3
    mass = np.asarray(M["body_mass"]) * 1e3
    walker = [i for i, n in enumerate(M["_names_body"]) if n.startswith("walker/")]
    def idx(name):
        return next(i for i in walker if names[i] == name)
    assert abs(mass[int(M["root_body"])] - 0.34) < 1e-3 * 0.34
    assert abs(sub[idx("head")] - 0.15) < 1e-3 * 0.15
    assert abs(sub[idx("abdomen")] - 0.38) < 5e-3 * 0.38
    wings = [sub[idx(n)] for n in ("wing_left", "wing_right")]
    assert abs(np.mean(wings) - 0.008) < 1e-3 * 0.008
    legs = [sub[idx(f"coxa_{t}_{s}")] for t in ("T1", "T2", "T3") for s in ("left", "right")]
    assert abs(np.mean(legs) - 0.0162) < 1e-2 * 0.0162          # 0.01629: the script prints 4 significant digits of 0.0163
    total = 0.15 + 0.34 + 0.38 + 6 * 0.0162 + 2 * 0.008           # make_fruitfly.py:1258
    assert abs(sub[int(M["root_body"])] - total) < 5e-3 * total


# --------------------------------------------------------------------------------------------------- quaternion fixture
@pytest.mark.parametrize("impl", ["oracle", "engine_f64", "engine_f32"])
def test_reference_quaternion_fixture_pins_the_oracle_and_engine_helpers(impl):
    """tests/golden/quaternions.npz holds inputs and the outputs of the reference's vnl_ray/quaternions.py (tools/make_golden.py),
    edge-case vectors of quat_z2vec (quaternions.py:220-247: +z, -z, scaled z, almost z) first.  The oracle's C helpers and the
    engine's (csrc/fly_env.cuh, compiled for the CPU in tests/emu) must reproduce them."""
    g = np.load(os.path.join(GOLD, "quaternions.npz"))
    if impl == "oracle":
        from oracle.oracle import quat_op as op
        tol = 1e-12
    else:
        from tests.emu.emu import quat_op as eop
        op = lambda o, a, b=None: eop(o, a, b, double=impl == "engine_f64")
        tol = 1e-12 if impl == "engine_f64" else 2e-6
    q1, q2, v, ang = g["q1"], g["q2"], g["v"], g["ang"]
    for i in range(len(q1)):
        np.testing.assert_allclose(op(0, q1[i]), g["recip"][i], atol=tol * np.abs(g["recip"][i]).max())
        np.testing.assert_allclose(op(7, q1[i], q2[i]), g["mult"][i], atol=tol * np.abs(g["mult"][i]).max())
        np.testing.assert_allclose(op(1, v[i], q1[i]), g["rot"][i], atol=tol * max(1.0, np.abs(g["rot"][i]).max()))
        np.testing.assert_allclose(op(5, q1[i], q2[i]), g["dquat_local"][i], atol=tol * np.abs(g["dquat_local"][i]).max())
        if impl == "engine_f32" and i == 3:
            continue    # v = (1e-30, 0, 1): float32 cannot hold the 1e-30 tilt; the reference's float64 result is its axis
        np.testing.assert_allclose(op(2, v[i]), g["z2vec"][i], atol=tol)
        # arccos near 1 loses half the digits: the distance is compared through its cosine
        np.testing.assert_allclose(np.cos(op(3, q1[i], q2[i])), np.cos(g["dist"][i]), atol=tol)
        np.testing.assert_allclose(op(4, v[i], [ang[i]]), g["joint_quat"][i], atol=tol)
        np.testing.assert_allclose(op(6, np.r_[v[0], v[i]], q1[0]), g["egocentric"][i], atol=tol * max(1.0, np.abs(g["egocentric"][i]).max()))
