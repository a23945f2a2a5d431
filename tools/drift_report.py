"""North-star item "100-step rollout drift reported": the float32 kernel against the float64 oracle over 100 control steps
(1000 physics steps walking, 400 flying) from the same reset and the same actions, per task.  Prints one JSON line per
task with the relative state error after 1, 10, 30 and 100 control steps (max |x - x_ref| / max |x_ref| over qpos, qvel),
the reward error and whether contact counts still agree.  Test infrastructure use of oracle/ (a checker)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from flybody_b200 import clips as C, wbpg as W
from flybody_b200.compile_fly import load_model
from flybody_b200.engine import Engine, FLAG_TRACE
from flybody_b200.model_pack import asset_path
from oracle.oracle import OracleEnv

rel = lambda a, b: float(np.abs(np.asarray(a, np.float64) - b).max() / (np.abs(b).max() + 1e-12))
for task, scale in (("walk_imitation", 0.3), ("walk_on_ball", 0.3), ("flight_imitation", 0.3)):
    M = load_model(asset_path(task))
    eng = Engine(M, 2, 0, seed=0, flags=FLAG_TRACE)
    o = OracleEnv(M)
    if task == "walk_imitation":
        rows, off, ln = C.pack_clips(M, [C.synth_walk_clip(M, T=240, seed=0)])
        eng.load_clips(rows, off, ln); eng.set_clip_ids(np.zeros(2, np.int32))
        r = rows.astype(np.float64); nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
        o.set_clip(r[:, :7 + nj], r[:, 7 + nj:13 + 2 * nj], r[:, 13 + 2 * nj:13 + 2 * nj + 3 * ns], r[:, 13 + 2 * nj + 3 * ns:])
    elif task == "flight_imitation":
        rows, off, ln = C.pack_clips(M, [C.synth_flight_clip(T=240, seed=7)])
        t = W.build_tables(W.synthetic_base_pattern())
        eng.load_clips(rows, off, ln); eng.load_wbpg(t); eng.set_clip_ids(np.zeros(2, np.int32)); eng.set_wbpg_phase(np.array([0.2, 0.2]))
        r = rows.astype(np.float64)
        o.set_clip(r[:, :7], r[:, 7:13]); o.set_wbpg(t); o.set_initial_phase(0.2)
    eng.reset_host(); o.env_reset()
    rng = np.random.RandomState(1)
    lo, hi = np.asarray(M["action_min"]), np.asarray(M["action_max"])
    mid, half = 0.5 * (lo + hi), 0.5 * (hi - lo)
    out = {"task": task, "actions": f"mid-range + {scale} x half-range x U(-1,1), fresh every control step", "steps": {}}
    alive = True
    for k in range(1, 101):
        a = (mid + scale * half * rng.uniform(-1, 1, lo.shape)).astype(np.float32)
        st, rw, g, obs = eng.step_host(np.tile(a, (2, 1)))
        ost, orw, og, oobs = o.env_step(a.astype(np.float64))
        s, tr = eng.get_state(), eng.get_trace()
        if k in (1, 10, 30, 100) or ost == 2 or st[0] != ost:
            out["steps"][k] = {"qpos": rel(s["qpos"][0], o.d("qpos")), "qvel": rel(s["qvel"][0], o.d("qvel")),
                               "reward_abs": abs(float(rw[0]) - orw), "step_type": [int(st[0]), int(ost)],
                               "ncon": [int(tr["ncon"][0]), len(o.contacts())]}
        if ost == 2 or st[0] != ost:
            out["ended_at"] = k
            break
    print(json.dumps(out), flush=True)
    eng.close()
