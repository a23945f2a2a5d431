"""Single source of truth for the packed fly model (`FlybModel`).

`include/flyb_model.h` is generated from these lists (tools/gen_header.py) and
the ctypes mirror in `model_pack.py` is built from them at import, so the C
struct and the Python struct cannot drift apart (tests/test_abi.py checks the
header is current and that sizeof agrees on both sides).

Every array is host memory: int32 or float64, C-contiguous.  The CUDA library
converts to float32 when it uploads; the float64 oracle uses them as they are.
"""

# scalar int fields, in struct order
SCALAR_INT = [
    "nq", "nv", "nu", "na", "nbody", "njnt", "ngeom", "npair", "nsite", "ntendon", "nwrap",
    "nsensor", "nsensordata", "nfluidbox", "nfluidgeom",
    "opt_iterations", "opt_ls_iterations", "opt_noslip_iterations",
    # task block
    "task_type",            # 0 walk_on_ball, 1 walk_imitation, 2 flight_imitation, 3 template, 4 vision_flight
    "n_substeps", "future_steps", "action_dim", "obs_dim", "max_episode_steps", "time_limit_steps",
    "root_body", "root_qposadr", "root_dofadr", "ghost_qposadr", "ghost_dofadr", "ball_dofadr",
    "thorax_site", "n_obs_joints", "n_appendages", "n_mocap_joints", "n_mocap_sites",
    "user_action", "pad_first_obs", "inference_mode",
    "hfield_geom", "hfield_nrow", "hfield_ncol",   # vision_flight: the `terrain` geom (-1: none) and its grid
]

# scalar double fields
SCALAR_DBL = [
    "opt_timestep", "opt_density", "opt_viscosity", "opt_tolerance", "opt_ls_tolerance",
    "opt_noslip_tolerance", "opt_impratio", "stat_meaninertia",
    "control_dt", "time_limit", "terminal_linvel", "terminal_angvel", "terminal_com_dist",
    "terminal_qacc", "terminal_height",
]

# (name, dtype, shape) — shape entries are field names or ints
ARRAYS = [
    ("opt_gravity", "d", (3,)),
    # bodies (fused: every body except world carries at least one joint)
    ("body_parent", "i", ("nbody",)), ("body_rootid", "i", ("nbody",)),
    ("body_jntadr", "i", ("nbody",)), ("body_jntnum", "i", ("nbody",)),
    ("body_dofadr", "i", ("nbody",)), ("body_dofnum", "i", ("nbody",)),
    ("body_pos", "d", ("nbody", 3)), ("body_quat", "d", ("nbody", 4)),
    ("body_ipos", "d", ("nbody", 3)), ("body_iquat", "d", ("nbody", 4)),
    ("body_mass", "d", ("nbody",)), ("body_inertia", "d", ("nbody", 3)),
    ("body_invweight0", "d", ("nbody", 2)), ("body_subtreemass", "d", ("nbody",)),
    # joints / dofs
    ("jnt_type", "i", ("njnt",)), ("jnt_qposadr", "i", ("njnt",)), ("jnt_dofadr", "i", ("njnt",)),
    ("jnt_bodyid", "i", ("njnt",)), ("jnt_limited", "i", ("njnt",)),
    ("jnt_pos", "d", ("njnt", 3)), ("jnt_axis", "d", ("njnt", 3)), ("jnt_stiffness", "d", ("njnt",)),
    ("jnt_range", "d", ("njnt", 2)), ("jnt_solref", "d", ("njnt", 2)), ("jnt_solimp", "d", ("njnt", 5)),
    ("qpos0", "d", ("nq",)), ("qpos_spring", "d", ("nq",)), ("qpos_reset", "d", ("nq",)),
    ("dof_bodyid", "i", ("nv",)), ("dof_jntid", "i", ("nv",)), ("dof_parentid", "i", ("nv",)),
    ("dof_armature", "d", ("nv",)), ("dof_damping", "d", ("nv",)), ("dof_invweight0", "d", ("nv",)),
    # collision geoms and the static candidate pair list (filters applied on the host)
    ("geom_type", "i", ("ngeom",)), ("geom_bodyid", "i", ("ngeom",)),
    ("geom_pos", "d", ("ngeom", 3)), ("geom_quat", "d", ("ngeom", 4)), ("geom_size", "d", ("ngeom", 3)),
    ("geom_rbound", "d", ("ngeom",)),
    ("pair_geom1", "i", ("npair",)), ("pair_geom2", "i", ("npair",)), ("pair_condim", "i", ("npair",)),
    ("pair_friction", "d", ("npair", 3)), ("pair_solref", "d", ("npair", 2)), ("pair_solimp", "d", ("npair", 5)),
    ("pair_margin", "d", ("npair",)), ("pair_gap", "d", ("npair",)),
    # sites
    ("site_bodyid", "i", ("nsite",)), ("site_type", "i", ("nsite",)),
    ("site_pos", "d", ("nsite", 3)), ("site_quat", "d", ("nsite", 4)), ("site_size", "d", ("nsite", 3)),
    # fixed tendons
    ("tendon_adr", "i", ("ntendon",)), ("tendon_num", "i", ("ntendon",)),
    ("wrap_dofadr", "i", ("nwrap",)), ("wrap_qposadr", "i", ("nwrap",)), ("wrap_coef", "d", ("nwrap",)),
    # actuators
    ("act_trntype", "i", ("nu",)), ("act_trnid", "i", ("nu",)), ("act_dyntype", "i", ("nu",)),
    ("act_actadr", "i", ("nu",)), ("act_ctrllimited", "i", ("nu",)), ("act_forcelimited", "i", ("nu",)),
    ("act_gainprm", "d", ("nu",)), ("act_biasprm", "d", ("nu", 3)), ("act_dynprm", "d", ("nu",)),
    ("act_ctrlrange", "d", ("nu", 2)), ("act_forcerange", "d", ("nu", 2)), ("act_gear", "d", ("nu",)),
    # sensors
    ("sensor_type", "i", ("nsensor",)), ("sensor_site", "i", ("nsensor",)), ("sensor_adr", "i", ("nsensor",)),
    # fluid elements
    ("fbox_bodyid", "i", ("nfluidbox",)), ("fbox_pos", "d", ("nfluidbox", 3)),
    ("fbox_quat", "d", ("nfluidbox", 4)), ("fbox_size", "d", ("nfluidbox", 3)),
    ("fgeom_bodyid", "i", ("nfluidgeom",)), ("fgeom_pos", "d", ("nfluidgeom", 3)),
    ("fgeom_quat", "d", ("nfluidgeom", 4)), ("fgeom_size", "d", ("nfluidgeom", 3)),
    ("fgeom_coef", "d", ("nfluidgeom", 12)),
    # task tables
    ("ctrl_action", "i", ("nu",)),                 # ctrl[i] = action[ctrl_action[i]]  (fruitfly.py:480-492)
    ("action_min", "d", ("action_dim",)), ("action_max", "d", ("action_dim",)),
    ("obs_joint_qposadr", "i", ("n_obs_joints",)), ("obs_joint_dofadr", "i", ("n_obs_joints",)),
    ("appendage_site", "i", ("n_appendages",)),
    ("mocap_joint", "i", ("n_mocap_joints",)),     # joint ids (hinges) tracked by the imitation reward
    ("mocap_site", "i", ("n_mocap_sites",)),
    ("wing_joint", "i", (6,)), ("wing_action", "i", (6,)),
    ("obs_offset", "i", (16,)),                    # offsets of the observation blocks (OBS_* order), -1 = absent
    ("ghost_offset", "d", (3,)),
    # vision-guided flight (tasks/vision_flight.py)
    ("hfield_size", "d", (4,)),                    # x half extent, y half extent, elevation scale, base thickness
    ("eye_site", "i", (2,)),                       # sites carrying the right / left eye camera frames (-1: none)
    ("vision_init_quat", "d", (4,)), ("vision_target_zaxis", "d", (3,)),
    ("vision_ranges", "d", (8,)),                  # target height lo hi, target speed lo hi, init x lo hi, init y lo hi
]

# observation blocks, alphabetical by key (`walker/<name>`); index into obs_offset
OBS_KEYS = [
    "accelerometer", "actuator_activation", "appendages_pos", "ball_qvel", "force", "gyro", "joints_pos",
    "joints_vel", "ref_displacement", "ref_root_quat", "task_input", "touch", "velocimeter", "world_zaxis",
]
