"""CPU emulation of the engine source (tests/emu) against the float64 oracle.

The emulation compiles flybody_b200/csrc/fly_*.cuh - the code the CUDA kernels are built from - with a warp
mapped to one thread.  In double it must agree with the oracle to rounding (proves the warp-cooperative
decomposition computes the same mathematics); in float it must meet the north-star tolerances
(one-step qpos/qvel relative error <= 1e-5, reward <= 1e-5) - the same bar the -m gpu tests hold the GPU to."""
import numpy as np
import pytest

from tests.conftest import clip_arrays


def _rel(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return np.abs(a - b).max() / (np.abs(b).max() + 1e-30)


def _rand_state(M, rng, scale=0.3):
    q = M["qpos_reset"].copy()
    for j in range(int(M["njnt"])):
        if M["jnt_type"][j] == 3:
            lo, hi = M["jnt_range"][j]
            a = M["jnt_qposadr"][j]
            q[a] = np.clip(q[a] + rng.uniform(-scale, scale) * (hi - lo), lo - 0.02, hi + 0.02)
    if int(M["task_type"]) != 0:
        ra = int(M["root_qposadr"])
        q[ra + 2] -= 0.02 * rng.uniform(0, 1)
        qq = q[ra + 3:ra + 7] + rng.normal(0, 0.05, 4)
        q[ra + 3:ra + 7] = qq / np.linalg.norm(qq)
    return q, rng.normal(0, 1.0, int(M["nv"]))


def _pair(M, double, flags=0, opts=()):
    from oracle.oracle import OracleEnv
    from tests.emu.emu import EmuEnv
    o, e = OracleEnv(M), EmuEnv(M, flags=flags, double=double)
    for k in opts:
        o.option(k, 1)
    return o, e


def _load_state(M, o, e, rng, scale):
    q, v = _rand_state(M, rng, scale)
    ctrl = rng.uniform(M["act_ctrlrange"][:, 0], M["act_ctrlrange"][:, 1])
    act = rng.uniform(-0.3, 0.3, int(M["na"]))
    o.d("qpos")[:] = q; o.d("qvel")[:] = v; o.d("ctrl")[:] = ctrl; o.d("qacc_warmstart")[:] = 0
    if int(M["na"]):
        o.d("act")[:] = act
    e.set("qpos", q); e.set("qvel", v); e.set("ctrl", ctrl); e.set("act", act); e.set("warm", np.zeros(int(M["nv"])))


@pytest.mark.parametrize("task", ["walk_imitation", "walk_on_ball", "flight_imitation"])
def test_double_emulation_matches_oracle_stage_by_stage(models, task):
    M = models[task]
    o, e = _pair(M, True)
    rng = np.random.RandomState(1)
    nv = int(M["nv"])
    for trial, scale in enumerate((0.05, 0.3, 0.3)):
        _load_state(M, o, e, rng, scale)
        o.forward(); e.forward()
        for nm in ("xpos", "xquat", "xipos", "geom_xpos", "site_xpos", "cinert", "cdof", "cvel", "cdof_dot", "qfrc_bias",
                   "qfrc_passive", "qfrc_actuator", "qacc_smooth", "sensordata"):
            a, b = e.get(nm), o.d(nm)
            n = min(len(a), len(b))
            assert _rel(a[:n], b[:n]) < 1e-11, (task, trial, nm)
        assert _rel(e.dense_m("qLD"), np.tril(o.d("qLD").reshape(nv, nv))) < 1e-11
        e.call("crb")  # fly_crb leaves M in qLD, which is then factorised in place
        assert _rel(e.dense_m("qLD"), np.tril(o.d("qM").reshape(nv, nv))) < 1e-13
        e.forward()
        # contact selection: same pairs, same order (bit-exact integer work)
        oc = o.contacts()
        assert e.geti("counts")[0] == len(oc)
        assert e.geti("con_pair")[:len(oc)].tolist() == [c["pair"] for c in oc]
        assert e.geti("con_efc")[:len(oc)].tolist() == [c["efc_address"] for c in oc]
        nefc = int(o.i("nefc")[0])
        assert e.geti("counts")[1] == nefc and e.geti("counts")[6] == 0
        if nefc:
            assert _rel(e.get("efc_aref")[:nefc], o.d("efc_aref")[:nefc]) < 1e-11
            assert _rel(e.get("efc_force")[:nefc], o.d("efc_force")[:nefc]) < 1e-7
        assert _rel(e.get("qacc"), o.d("qacc")) < 1e-8
        o.step(3); e.physics_step(3)
        assert _rel(e.get("qpos"), o.d("qpos")) < 1e-10 and _rel(e.get("qvel"), o.d("qvel")) < 1e-8


@pytest.mark.parametrize("task", ["walk_imitation", "walk_on_ball", "flight_imitation"])
def test_float_emulation_meets_one_step_tolerance(models, task):
    M = models[task]
    o, e = _pair(M, False)
    rng = np.random.RandomState(2)
    for scale in (0.05, 0.2, 0.3):
        _load_state(M, o, e, rng, scale)
        o.forward(); e.forward()
        oc = o.contacts()
        # pair selection can only differ for a contact within float rounding of its margin
        margins = [abs(c["dist"] - float(M["pair_margin"][c["pair"]])) for c in oc]
        if not margins or min(margins) > 1e-6:
            assert e.geti("con_pair")[:len(oc)].tolist() == [c["pair"] for c in oc]
        o.step(1); e.physics_step(1)
        assert _rel(e.get("qpos"), o.d("qpos")) <= 1e-5
        assert _rel(e.get("qvel"), o.d("qvel")) <= 1e-5


@pytest.mark.parametrize("flags,opts", [(1, ("disable_contact",)), (2, ("disable_limit",)), (4, ("disable_fluid",)),
                                        (8, ("disable_noslip",)), (32, ("disable_warmstart",)), (1 | 2 | 4 | 16, ("disable_contact", "disable_limit", "disable_fluid", "disable_actuation"))])
def test_stage_switches_match(models, flags, opts):
    M = models["walk_imitation"]
    o, e = _pair(M, True, flags, opts)
    rng = np.random.RandomState(3)
    _load_state(M, o, e, rng, 0.2)
    o.forward(); e.forward()
    o.step(2); e.physics_step(2)
    assert _rel(e.get("qvel"), o.d("qvel")) < 1e-8


def test_env_level_rollout_matches_oracle(models, walk_clip):
    """reset + control steps through the task layer: obs, reward, discount, step_type, termination step."""
    from oracle.oracle import OracleEnv
    from tests.emu.emu import EmuEnv
    M = models["walk_imitation"]
    rows, off, ln = walk_clip
    for double, tol_obs, tol_r in ((True, 2e-7, 1e-6), (False, 2e-4, 1e-5)):
        o, e = OracleEnv(M), EmuEnv(M, double=double)
        o.set_clip(*clip_arrays(M, rows)); e.set_clips(rows, off, ln)
        st, oobs = o.env_reset(); st2, eobs = e.env_reset()
        assert st == st2 == 0 and _rel(eobs, oobs) < tol_obs
        rng = np.random.RandomState(0)
        for k in range(45):
            a = rng.uniform(M["action_min"], M["action_max"]).astype(np.float32)
            if k == 3:
                a[5] = np.nan  # walk_imitation.py:147
            st, r, g, oobs = o.env_step(a.astype(np.float64)); st2, r2, g2, eobs = e.env_step(a)
            assert st == st2 and g == g2
            assert abs(r - r2) <= tol_r * max(1.0, abs(r))
            assert _rel(eobs, oobs) < tol_obs
            if st == 2:
                break
        assert st == 2
        # the next step() is the auto-reset: FIRST, reward 0, discount 1
        st, r, g, oobs = o.env_step(a.astype(np.float64)); st2, r2, g2, eobs = e.env_step(a)
        assert st == st2 == 0 and r2 == 0 and g2 == 1 and _rel(eobs, oobs) < tol_obs


@pytest.mark.parametrize("task", ["walk_on_ball", "template"])
def test_clipless_task_rollout_matches_oracle(models, task):
    """walk_on_ball (tolerance reward on the ball velocity) and the no-op template task (reward 1, qacc termination only)."""
    from oracle.oracle import OracleEnv
    from tests.emu.emu import EmuEnv
    M = models[task]
    o, e = OracleEnv(M), EmuEnv(M, double=False)
    o.env_reset(); e.env_reset()
    rng = np.random.RandomState(0)
    for k in range(8):
        a = rng.uniform(M["action_min"], M["action_max"]).astype(np.float32)
        st, r, g, oobs = o.env_step(a.astype(np.float64)); st2, r2, g2, eobs = e.env_step(a)
        assert st == st2 and g == g2 and abs(r - r2) <= 1e-5
        assert _rel(eobs, oobs) < 5e-4


def test_flight_imitation_rollout_matches_oracle(models):
    """flight task layer: generator indices bit-exact, obs / reward in tolerance."""
    from flybody_b200 import clips as C, wbpg as W
    from oracle.oracle import OracleEnv
    from tests.emu.emu import EmuEnv
    M = models["flight_imitation"]
    rows, off, ln = C.pack_clips(M, [C.synth_flight_clip(T=120, seed=5)])
    r = rows.astype(np.float64)
    t = W.build_tables(W.synthetic_base_pattern())
    for double, tol_obs, tol_r in ((True, 2e-7, 1e-6), (False, 5e-4, 1e-5)):
        o, e = OracleEnv(M), EmuEnv(M, double=double)
        o.set_clip(r[:, :7], r[:, 7:13]); o.set_wbpg(t); o.set_initial_phase(0.61)
        e.set_clips(rows, off, ln); e.set_wbpg(t); e.set_initial_phase(0.61)
        st, oobs = o.env_reset(); st2, eobs = e.env_reset()
        assert st == st2 == 0 and _rel(eobs, oobs) < tol_obs and o.wbpg_state() == e.wbpg_state()
        rng = np.random.RandomState(1)
        user = int(M["user_action"])
        u = 0.0
        for k in range(130):
            a = (0.4 * rng.uniform(M["action_min"], M["action_max"])).astype(np.float32)
            u = float(np.clip(u + rng.normal(scale=0.3), -1, 1))
            a[user] = u
            st, rw, g, oobs = o.env_step(a.astype(np.float64)); st2, rw2, g2, eobs = e.env_step(a)
            assert st == st2 and g == g2
            assert o.wbpg_state() == e.wbpg_state(), k          # (table step, frequency index, filtered frequency)
            assert abs(rw - rw2) <= tol_r * max(1.0, abs(rw))
            assert _rel(eobs, oobs) < tol_obs
            if st == 2:
                break
        # the clip has 120 rows: the episode ends "well" at step 120 - 6 with discount 1 (flight_imitation.py:106-108,211-220)
        assert st == 2 and k + 1 == 120 - 6 and g == 1.0
