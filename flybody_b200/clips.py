"""Reference clips for the imitation tasks (host side).

The 16 252-snippet walking dataset named by `vnl_ray/config/train_config_fly.yaml:12` is not shipped, and
h5py is not installed, so clips are generated synthetically in the *format* the reference loaders produce
(`vnl_ray/tasks/trajectory_loaders.py:163-215`): per clip
    qpos       [T, 7 + n_j]   root pos, root quat, mocap joint angles (xy re-origined, loaders :130,206)
    qvel       [T, 6 + n_j]   root linear (world) / angular (local) velocity, joint velocities
    root2site  [T, n_s, 3]    egocentric end-effector vectors   (semantics of rewards.py:36-60)
    joint_quat [T, n_j, 4]    egocentric joint orientation quats (quaternions.py:298-321)
`pack_clips` lays a set of clips out as one float32 table + offsets for the device.
"""
from __future__ import annotations

import numpy as np

from .mjcf import quat_mul, quat_conj, quat_to_mat, quat_rot
from .compile_fly import _kin0, JNT_HINGE


def _quat_z2vec(v):
    v = v / np.linalg.norm(v)
    if v[0] == 0.0 and v[1] == 0.0:
        return np.array([0.0, 1, 0, 0]) if v[2] < 0 else np.array([1.0, 0, 0, 0])
    ax = np.array([-v[1], v[0], 0.0])
    ax /= np.linalg.norm(ax)
    ang = np.arccos(v[2])
    return np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * ax])


def walker_features(M: dict, qpos: np.ndarray):
    """root2site [n_s,3] and joint_quat [n_j,4] of the walker at `qpos` (rewards.get_walker_features)."""
    xpos, xquat, xanchor, xaxis = _kin0(M, qpos)
    ra = M["root_qposadr"]
    p, q = qpos[ra:ra + 3], qpos[ra + 3:ra + 7] / np.linalg.norm(qpos[ra + 3:ra + 7])
    Rt = quat_to_mat(q).T
    r2s = []
    for s in M["mocap_site"]:
        b = M["site_bodyid"][s]
        sp = xpos[b] + quat_rot(xquat[b], M["site_pos"][s])
        r2s.append(Rt @ (sp - p))
    jq = []
    for j in M["mocap_joint"]:
        ax = Rt @ xaxis[j]
        ang = qpos[M["jnt_qposadr"][j]]
        q2 = np.concatenate([[np.cos(ang / 2)], np.sin(ang / 2) * ax / np.linalg.norm(ax)])
        jq.append(quat_mul(q2, _quat_z2vec(ax)))
    return np.array(r2s).reshape(-1, 3), np.array(jq).reshape(-1, 4)


def synth_walk_clip(M: dict, T: int = 600, seed: int = 0, speed=(1.0, 3.0), features: bool = True):
    """One synthetic walking clip: smooth random-walk root path at 1-3 cm/s, z at the spawn height
    (fruitfly.py:23), integrated yaw, tripod-gait leg sinusoids inside each joint's range."""
    rng = np.random.RandomState(seed)
    dt = float(M["control_dt"])
    nj = int(M["n_mocap_joints"])
    v = rng.uniform(*speed)
    yaw_rate = np.cumsum(rng.normal(0, 0.02, T)) * 0.5
    yaw_rate = np.clip(yaw_rate, -1.0, 1.0)
    yaw = rng.uniform(-np.pi, np.pi) + np.cumsum(yaw_rate) * dt
    pos = np.zeros((T, 3))
    pos[:, 0] = np.cumsum(v * np.cos(yaw)) * dt
    pos[:, 1] = np.cumsum(v * np.sin(yaw)) * dt
    pos[:, :2] -= pos[0, :2]
    pos[:, 2] = M["qpos0"][M["root_qposadr"] + 2]
    quat = np.stack([np.cos(yaw / 2), 0 * yaw, 0 * yaw, np.sin(yaw / 2)], axis=1)
    qpos = np.zeros((T, 7 + nj))
    qpos[:, :3], qpos[:, 3:7] = pos, quat
    names = M["_names_jnt"]
    t = np.arange(T) * dt
    freq = rng.uniform(8.0, 12.0)
    for i, j in enumerate(M["mocap_joint"]):
        lo, hi = M["jnt_range"][j]
        n = names[j]
        leg = ["T1_left", "T1_right", "T2_left", "T2_right", "T3_left", "T3_right"]
        li = next((k for k, s in enumerate(leg) if s in n), 0)
        phase = 0.0 if li in (0, 3, 4) else np.pi  # tripod groups
        amp = 0.0
        if "coxa_T" in n and "twist" not in n and "abduct" not in n:
            amp = 0.12
        elif "femur_T" in n and "twist" not in n:
            amp = 0.15
        elif "tibia" in n:
            amp = 0.15
        a = amp * np.sin(2 * np.pi * freq * t + phase + (0.5 if "tibia" in n else 0.0))
        qpos[:, 7 + i] = np.clip(M["qpos0"][M["jnt_qposadr"][j]] + a, lo + 1e-3, hi - 1e-3)
    qvel = np.zeros((T, 6 + nj))
    qvel[:-1, :3] = np.diff(pos, axis=0) / dt
    for k in range(T - 1):
        dq = quat_mul(quat_conj(quat[k]), quat[k + 1])
        s = np.linalg.norm(dq[1:])
        ang = 2 * np.arctan2(s, dq[0])
        qvel[k, 3:6] = (dq[1:] / s * ang / dt) if s > 1e-12 else 0.0
    qvel[:-1, 6:] = np.diff(qpos[:, 7:], axis=0) / dt
    qvel[-1] = qvel[-2]
    clip = dict(qpos=qpos, qvel=qvel)
    if features:
        r2s, jq = np.zeros((T, int(M["n_mocap_sites"]), 3)), np.zeros((T, nj, 4))
        full = M["qpos_reset"].copy()
        for k in range(T):
            full[M["root_qposadr"]:M["root_qposadr"] + 7] = qpos[k, :7]
            for i, j in enumerate(M["mocap_joint"]):
                full[M["jnt_qposadr"][j]] = qpos[k, 7 + i]
            r2s[k], jq[k] = walker_features(M, full)
        clip.update(root2site=r2s, joint_quat=jq)
    return clip


def clip_row_width(M: dict) -> int:
    """Floats per packed clip row: qpos (7+n_j) | qvel (6+n_j) | root2site (3 n_s) | joint_quat (4 n_j)."""
    nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
    return (7 + nj) + (6 + nj) + 3 * ns + 4 * nj


def pack_clips(M: dict, clips: list):
    """-> (table float32 [sum T, row_width], offsets int32 [n], lengths int32 [n])."""
    W = clip_row_width(M)
    nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
    rows, off, ln, n = [], [], [], 0
    for c in clips:
        T = c["qpos"].shape[0]
        r = np.zeros((T, W), np.float32)
        r[:, :7 + nj] = c["qpos"]
        r[:, 7 + nj:13 + 2 * nj] = c["qvel"]
        if ns:
            r[:, 13 + 2 * nj:13 + 2 * nj + 3 * ns] = c["root2site"].reshape(T, -1)
        if nj:
            r[:, 13 + 2 * nj + 3 * ns:] = c["joint_quat"].reshape(T, -1)
        rows.append(r); off.append(n); ln.append(T); n += T
    return np.concatenate(rows), np.array(off, np.int32), np.array(ln, np.int32)


# ---------------------------------------------------------------------------- batched (vectorised) generation
def _bq_mul(a, b):
    aw, ax, ay, az = a[..., 0], a[..., 1], a[..., 2], a[..., 3]
    bw, bx, by, bz = b[..., 0], b[..., 1], b[..., 2], b[..., 3]
    return np.stack([aw * bw - ax * bx - ay * by - az * bz, aw * bx + ax * bw + ay * bz - az * by,
                     aw * by - ax * bz + ay * bw + az * bx, aw * bz + ax * by - ay * bx + az * bw], axis=-1)


def _bq_rot(q, v):
    """rotate v [...,3] by unit quaternion q [...,4]"""
    w, u = q[..., :1], q[..., 1:]
    t = 2.0 * np.cross(u, v)
    return v + w * t + np.cross(u, t)


def batch_walker_features(M: dict, qpos: np.ndarray):
    """`walker_features` for a batch of full qpos vectors [B, nq] -> root2site [B, n_s, 3], joint_quat [B, n_j, 4].
    Same arithmetic as rewards.get_walker_features (vnl_ray/tasks/rewards.py:36-60), vectorised over frames."""
    B, nb = qpos.shape[0], int(M["nbody"])
    xpos = np.zeros((B, nb, 3))
    xquat = np.zeros((B, nb, 4)); xquat[..., 0] = 1.0
    xaxis = np.zeros((B, int(M["njnt"]), 3))
    for b in range(1, nb):
        p = int(M["body_parent"][b])
        pos = xpos[:, p] + _bq_rot(xquat[:, p], np.broadcast_to(M["body_pos"][b], (B, 3)))
        quat = _bq_mul(xquat[:, p], np.broadcast_to(M["body_quat"][b], (B, 4)))
        for j in range(int(M["body_jntadr"][b]), int(M["body_jntadr"][b] + M["body_jntnum"][b])):
            a, t = int(M["jnt_qposadr"][j]), int(M["jnt_type"][j])
            ax0 = np.broadcast_to(M["jnt_axis"][j], (B, 3))
            if t == 0:
                pos = qpos[:, a:a + 3].copy()
                quat = qpos[:, a + 3:a + 7] / np.linalg.norm(qpos[:, a + 3:a + 7], axis=1, keepdims=True)
                xaxis[:, j] = _bq_rot(quat, ax0)
                continue
            jp = np.broadcast_to(M["jnt_pos"][j], (B, 3))
            anchor = pos + _bq_rot(quat, jp)
            xaxis[:, j] = _bq_rot(quat, ax0)
            if t == 1:
                ql = qpos[:, a:a + 4] / np.linalg.norm(qpos[:, a:a + 4], axis=1, keepdims=True)
            else:
                ang = qpos[:, a] - M["qpos0"][a]
                ql = np.concatenate([np.cos(ang / 2)[:, None], np.sin(ang / 2)[:, None] * ax0], axis=1)
            quat = _bq_mul(quat, ql)
            pos = anchor - _bq_rot(quat, jp)
        xpos[:, b], xquat[:, b] = pos, quat
    ra = int(M["root_qposadr"])
    p = qpos[:, ra:ra + 3]
    q = qpos[:, ra + 3:ra + 7] / np.linalg.norm(qpos[:, ra + 3:ra + 7], axis=1, keepdims=True)
    qc = q * np.array([1.0, -1, -1, -1])
    ns, nj = int(M["n_mocap_sites"]), int(M["n_mocap_joints"])
    r2s = np.zeros((B, ns, 3))
    for i, s in enumerate(M["mocap_site"]):
        b = int(M["site_bodyid"][s])
        sp = xpos[:, b] + _bq_rot(xquat[:, b], np.broadcast_to(M["site_pos"][s], (B, 3)))
        r2s[:, i] = _bq_rot(qc, sp - p)
    jq = np.zeros((B, nj, 4))
    for i, j in enumerate(M["mocap_joint"]):
        ax = _bq_rot(qc, xaxis[:, j])
        ax = ax / np.linalg.norm(ax, axis=1, keepdims=True)
        ang = qpos[:, int(M["jnt_qposadr"][j])]
        q2 = np.concatenate([np.cos(ang / 2)[:, None], np.sin(ang / 2)[:, None] * ax], axis=1)
        # quat_z2vec (quaternions.py:205-252), vectorised incl. the axis-parallel-to-z edge case
        edge = (ax[:, 0] == 0.0) & (ax[:, 1] == 0.0)
        rot = np.stack([-ax[:, 1], ax[:, 0]], axis=1)
        rn = np.linalg.norm(rot, axis=1, keepdims=True)
        rot = rot / np.where(rn > 0, rn, 1.0)
        a2 = np.arccos(np.clip(ax[:, 2], -1.0, 1.0))
        q1 = np.stack([np.cos(a2 / 2), np.sin(a2 / 2) * rot[:, 0], np.sin(a2 / 2) * rot[:, 1], 0 * a2], axis=1)
        q1[edge & (ax[:, 2] < 0)] = [0.0, 1, 0, 0]
        q1[edge & (ax[:, 2] >= 0)] = [1.0, 0, 0, 0]
        jq[:, i] = _bq_mul(q2, q1)
    return r2s, jq


def synthetic_table(M: dict, n_clips: int = 1024, T: int = 600, seed: int = 0, cache: bool = True):
    """Packed synthetic walking clip table: -> (rows float32 [n T, W], offsets, lengths).
    Cached under flybody_b200/assets/_cache/ because generation takes a while for 1024 x 600 frames."""
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets", "_cache",
                        f"{['walk_on_ball', 'walk_imitation', 'flight_imitation'][int(M['task_type'])]}_{n_clips}x{T}_s{seed}.npz")
    if cache and os.path.exists(path):
        d = np.load(path)
        return d["rows"], d["offsets"], d["lengths"]
    cl = [synth_walk_clip(M, T=T, seed=seed * 100003 + c, features=False) for c in range(n_clips)]
    nj = int(M["n_mocap_joints"])
    full = np.tile(np.asarray(M["qpos_reset"], np.float64), (n_clips * T, 1))
    allq = np.concatenate([c["qpos"] for c in cl])
    ra = int(M["root_qposadr"])
    full[:, ra:ra + 7] = allq[:, :7]
    for i, j in enumerate(M["mocap_joint"]):
        full[:, int(M["jnt_qposadr"][j])] = allq[:, 7 + i]
    r2s = np.zeros((n_clips * T, int(M["n_mocap_sites"]), 3)); jq = np.zeros((n_clips * T, nj, 4))
    for s in range(0, n_clips * T, 32768):
        r2s[s:s + 32768], jq[s:s + 32768] = batch_walker_features(M, full[s:s + 32768])
    for c in range(n_clips):
        cl[c]["root2site"], cl[c]["joint_quat"] = r2s[c * T:(c + 1) * T], jq[c * T:(c + 1) * T]
    out = pack_clips(M, cl)
    if cache:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        np.savez(path, rows=out[0], offsets=out[1], lengths=out[2])
    return out


def save_clips(path: str, M: dict, clips: list):
    rows, off, ln = pack_clips(M, clips)
    np.savez(path, rows=rows, offsets=off, lengths=ln)


# ---------------------------------------------------------------------------- flight reference clips
_COM_OFFSET = np.array([-0.03697732, 0.00029205, -0.0142447])  # task_utils.py:187,209: CoM in thorax coordinates


def com2root(com, quat):
    """Root-joint position from CoM position and orientation, batched (task_utils.py:194-213)."""
    return np.asarray(com, np.float64) + _bq_rot(np.asarray(quat, np.float64), np.broadcast_to(-_COM_OFFSET, np.shape(com)))


def root2com(root_qpos):
    """Inverse of com2root for one pose (task_utils.py:174-191)."""
    root_qpos = np.asarray(root_qpos, np.float64)
    return root_qpos[:3] + _bq_rot(root_qpos[3:7][None], _COM_OFFSET[None])[0]


def flight_clip_from_com(com_qpos, com_qvel):
    """(com_qpos [T,7], com_qvel [T,6]) as stored in the flight dataset -> clip dict with the ROOT pose:
    xy re-origin (trajectory_loaders.py:130) then com2root (flight_imitation.py:101-103)."""
    q = np.array(com_qpos, np.float64)
    q[:, :2] -= q[0, :2]
    return dict(qpos=np.concatenate([com2root(q[:, :3], q[:, 3:7]), q[:, 3:7]], axis=1), qvel=np.array(com_qvel, np.float64))


def synth_flight_clip(T: int = 3006, seed: int = 0, dt: float = 2e-4):
    """Synthetic CoM clip: hover with slow drift (|v| <= 30 cm/s) and yaw saccades."""
    rng = np.random.RandomState(seed)
    t = np.arange(T) * dt
    vel = np.zeros((T, 3))
    for k in range(3):
        for _ in range(3):
            vel[:, k] += rng.uniform(-5, 5) * np.sin(2 * np.pi * rng.uniform(1, 8) * t + rng.uniform(0, 2 * np.pi))
    pos = np.cumsum(vel, axis=0) * dt
    pos -= pos[0]
    pos[:, 2] += 1.5 + rng.uniform(-0.2, 0.2)
    yaw_rate = sum(rng.uniform(-6, 6) * np.sin(2 * np.pi * rng.uniform(1, 6) * t + rng.uniform(0, 2 * np.pi)) for _ in range(2))
    yaw = np.cumsum(yaw_rate) * dt + rng.uniform(-np.pi, np.pi)
    quat = np.stack([np.cos(yaw / 2), np.zeros(T), np.zeros(T), np.sin(yaw / 2)], axis=1)
    ang = np.stack([np.zeros(T), np.zeros(T), yaw_rate], axis=1)
    return flight_clip_from_com(np.concatenate([pos, quat], axis=1), np.concatenate([vel, ang], axis=1))


def synthetic_flight_table(M: dict, n_clips: int = 64, T: int = 3006, seed: int = 0):
    """Packed synthetic flight clip table -> (rows float32 [n T, 13], offsets, lengths)."""
    return pack_clips(M, [synth_flight_clip(T=T, seed=seed * 100003 + c) for c in range(n_clips)])


# ---------------------------------------------------------------------------- the reference's HDF5 datasets
def _open_hdf5(path):
    """h5py where it exists; otherwise the pure-python reader of `hdf5_lite` (same mapping interface)."""
    try:
        import h5py
        return h5py.File(path, "r")
    except ImportError:
        from . import hdf5_lite
        return hdf5_lite.File(path)


def _traj_keys(f):
    return sorted(f["trajectories"].keys(), key=lambda k: int(k))


def _decode(names):
    return [n.decode("utf-8") if isinstance(n, (bytes, np.bytes_)) else str(n) for n in np.asarray(names).tolist()]


def hdf5_walking_clips(f, M: dict, max_clips=None, traj_indices=None):
    """Clip dicts from an open walking dataset (`h5py.File`, `hdf5_lite.File` or a mapping of the same shape), as
    `HDF5WalkingTrajectoryLoader.get_trajectory` returns them (trajectory_loaders.py:189-215): `qpos = [root_qpos | qpos]`,
    `qvel = [root_qvel | qvel]`, xy origin moved to the first frame (:206), `root2site`, `joint_quat`.

    The reference takes its mocap joints / sites from the dataset's `id2name` lists (walk_imitation.py:63-71); the compiled model
    fixes them (`M["mocap_joint"]`, `M["mocap_site"]`), so when `id2name` is present its names are matched to the model's and the
    dataset's columns are reordered into the model's order; a name the model tracks but the dataset lacks is an error."""
    nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
    jsel = ssel = None
    if "id2name" in f:
        strip = lambda n: n.split("/", 1)[1] if n.startswith("walker/") else n   # noqa: E731
        if "joints" in f["id2name"] and "_names_jnt" in M:
            names = _decode(f["id2name"]["joints"])
            have = {n: i for i, n in enumerate(names)}
            want = [strip(M["_names_jnt"][int(j)]) for j in M["mocap_joint"]]
            if all(n in have for n in want):
                jsel = np.array([have[n] for n in want], np.int64)
            elif len(names) != nj:       # unknown names but the right count: taken in the order stored
                raise ValueError(f"dataset lists {len(names)} joints that do not cover the {nj} joints the compiled model tracks "
                                 f"(missing e.g. {[n for n in want if n not in have][:3]})")
        if "sites" in f["id2name"] and "_names_site" in M:
            names = _decode(f["id2name"]["sites"])
            have = {n: i for i, n in enumerate(names)}
            want = [strip(M["_names_site"][int(s)]) for s in M["mocap_site"]]
            if all(n in have for n in want):
                ssel = np.array([have[n] for n in want], np.int64)
            elif len(names) != ns:
                raise ValueError(f"dataset lists {len(names)} sites that do not cover the {ns} sites the compiled model tracks")
    keys = _traj_keys(f)
    if traj_indices is not None:
        keys = [keys[int(i)] for i in traj_indices]
    out = []
    for k in keys[:max_clips]:
        t = f["trajectories"][k]
        jq_pos, jq_vel = np.asarray(t["qpos"], np.float64), np.asarray(t["qvel"], np.float64)
        r2s, jq = np.asarray(t["root2site"], np.float64), np.asarray(t["joint_quat"], np.float64)
        if jsel is not None:
            jq_pos, jq_vel, jq = jq_pos[:, jsel], jq_vel[:, jsel], jq[:, jsel]
        if ssel is not None:
            r2s = r2s[:, ssel]
        qpos = np.concatenate([np.asarray(t["root_qpos"], np.float64), jq_pos], axis=1)
        qvel = np.concatenate([np.asarray(t["root_qvel"], np.float64), jq_vel], axis=1)
        qpos[:, :2] -= qpos[0, :2]
        if qpos.shape[1] != 7 + nj or r2s.shape[1:] != (ns, 3) or jq.shape[1:] != (nj, 4):
            raise ValueError(f"clip {k}: shapes {qpos.shape} {r2s.shape} {jq.shape} do not match the model ({nj} joints, {ns} sites)")
        out.append(dict(qpos=qpos, qvel=qvel, root2site=r2s, joint_quat=jq))
    return out


def hdf5_flight_clips(f, max_clips=None, traj_indices=None):
    """Clip dicts from an open flight dataset (`com_qpos`, `com_qvel` per trajectory; trajectory_loaders.py:90-132)."""
    keys = _traj_keys(f)
    if traj_indices is not None:
        keys = [keys[int(i)] for i in traj_indices]
    return [flight_clip_from_com(np.asarray(f["trajectories"][k]["com_qpos"], np.float64),
                                 np.asarray(f["trajectories"][k]["com_qvel"], np.float64)) for k in keys[:max_clips]]


def load_hdf5_clips(M: dict, path: str, max_clips=None, traj_indices=None):
    """`ref_path="*.hdf5"` of `fly_envs.walk_imitation` / `flight_imitation`: the reference's dataset file -> packed clip table.
    Walking or flight is told from the datasets the first trajectory holds."""
    with _open_hdf5(path) as f:
        first = f["trajectories"][_traj_keys(f)[0]]
        if "com_qpos" in first:
            cl = hdf5_flight_clips(f, max_clips, traj_indices)
        else:
            cl = hdf5_walking_clips(f, M, max_clips, traj_indices)
        dt = float(np.asarray(f["timestep_seconds"][()])) if "timestep_seconds" in f else None
    if dt is not None and abs(dt - float(M["control_dt"])) > 1e-9 * max(1.0, dt):
        raise ValueError(f"dataset timestep {dt} s differs from the task's control timestep {float(M['control_dt'])} s")
    return pack_clips(M, cl)


def export_hdf5(path: str, M: dict, clips: list, kind: str = "walking", **write_options):
    """Write clip dicts in the reference's dataset layout (trajectory_loaders.py:33-35, 90-97, 155-157, 189-223) with
    `hdf5_lite.write_file`: `trajectories/<zero-padded idx>/…`, `trajectory_lengths`, `timestep_seconds`, `id2name/{joints,sites}`.
    Flight clips are stored as CoM poses (`root2com`), walking clips as they are (their xy origin is already at the first frame)."""
    from . import hdf5_lite
    nz = len(str(len(clips)))
    traj = {}
    for i, c in enumerate(clips):
        if kind == "flight":
            com = np.stack([root2com(r) for r in c["qpos"]])
            traj[str(i).zfill(nz)] = dict(com_qpos=np.concatenate([com, c["qpos"][:, 3:7]], axis=1), com_qvel=np.asarray(c["qvel"], np.float64))
        else:
            traj[str(i).zfill(nz)] = dict(root_qpos=c["qpos"][:, :7], qpos=c["qpos"][:, 7:], root_qvel=c["qvel"][:, :6], qvel=c["qvel"][:, 6:],
                                          root2site=c["root2site"], joint_quat=c["joint_quat"])
    tree = {"trajectories": traj, "trajectory_lengths": np.array([c["qpos"].shape[0] for c in clips], np.int64),
            "timestep_seconds": np.float64(M["control_dt"])}
    if kind != "flight":
        strip = lambda n: n.split("/", 1)[1] if n.startswith("walker/") else n   # noqa: E731
        tree["id2name"] = {"joints": np.array([strip(M["_names_jnt"][int(j)]).encode() for j in M["mocap_joint"]]),
                           "sites": np.array([strip(M["_names_site"][int(s)]).encode() for s in M["mocap_site"]])}
    hdf5_lite.write_file(path, tree, **write_options)
