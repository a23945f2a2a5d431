/* flyb_oracle.h - float64 single-environment CPU restatement of the fly env-step hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (flybody_b200/, include/flyb.h) may link,
 * import or call this library; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, as the checker and as the timed CPU baseline.
 *
 * PARITY UNPINNED: the arithmetic restated here lives in third-party packages that are neither vendored
 * in /root/reference nor installable in this image - `mujoco` (requirements.txt:3, unpinned) and
 * `dm_control` (requirements.txt:6 / setup.py:24, unpinned) - and the reference ships no tests or golden
 * vectors.  The physics follows MuJoCo's published pipeline (mj_step2 + mj_step1 in
 * dm_control's legacy step order); the task layer follows the reference
 * files cited per function and IS pinned against the importable reference modules
 * (vnl_ray/quaternions.py, vnl_ray/tasks/rewards.py, vnl_ray/tasks/pattern_generators.py) through
 * tests/golden/.
 */
#ifndef FLYB_ORACLE_H_
#define FLYB_ORACLE_H_
#include "../include/flyb_model.h"

#ifdef __cplusplus
extern "C" {
#endif

#define FLYO_MAXCON 160
#define FLYO_MAXHFCON 16   /* contacts kept per geom - height field pair */
#define FLYO_MAXEFC 480

typedef struct FlyoContact {
  int pair, geom1, geom2, dim, exclude, efc_address;
  double dist, includemargin, mu;
  double pos[3], frame[9], friction[3], solref[2], solimp[5];
} FlyoContact;

typedef struct FlyoClip {  /* one reference clip, row-major float64 */
  int T;                   /* rows */
  const double* qpos;      /* [T, 7 + n_mocap_joints]  (flight: [T,7] root pose after com2root) */
  const double* qvel;      /* [T, 6 + n_mocap_joints]  (flight: [T,6]) */
  const double* root2site; /* [T, n_mocap_sites, 3] or NULL */
  const double* joint_quat;/* [T, n_mocap_joints, 4] or NULL */
} FlyoClip;

typedef struct FlyoWbpg {   /* packed wing-beat pattern generator tables (flybody_b200/wbpg.py), float64 */
  int nfreq;
  const double* beat_freqs; /* [nfreq] */
  const int* offsets;       /* [nfreq + 1] first row of each frequency */
  const double* traj;       /* [rows, 6] */
  const double* phase;      /* [rows] */
  double base_freq, rel_range, rate, dt_ctrl;
} FlyoWbpg;

typedef struct FlyoEnv FlyoEnv;

FlyoEnv* flyo_create(const FlybModel* m);
void flyo_destroy(FlyoEnv* e);

/* physics level */
void flyo_reset_data(FlyoEnv* e);                    /* qpos<-qpos0, everything else 0, then forward */
void flyo_forward(FlyoEnv* e);                       /* mj_forward */
void flyo_step(FlyoEnv* e);                          /* legacy physics.step(): mj_step2 then mj_step1 */
void flyo_set_option(FlyoEnv* e, const char* name, double value);
/* named views into the env's arrays (tests): returns pointer and writes the element count */
double* flyo_dptr(FlyoEnv* e, const char* name, int* count);
int* flyo_iptr(FlyoEnv* e, const char* name, int* count);
const FlyoContact* flyo_contacts(FlyoEnv* e, int* ncon);
/* narrow phase of one primitive pair (unit tests): out[7*i] = dist, pos(3), normal(3) of hit i; returns the hit count */
int flyo_narrow_hfield(const double* hpos, const double* hmat, const double* hsize, int nrow, int ncol, const float* data,
                       int t2, const double* c2, const double* m2, const double* s2, double margin, int maxcon, double* out);
int flyo_quat_op(int op, const double* a, const double* b, double* out);   /* test hook: task-layer quaternion helpers */
int flyo_narrow_pair(int t1, const double* c1, const double* m1, const double* s1, int t2, const double* c2,
                     const double* m2, const double* s2, double margin, int force_convex, double* out);

/* task level (composer.Environment semantics) */
void flyo_set_clip(FlyoEnv* e, const FlyoClip* clip);
void flyo_set_wbpg(FlyoEnv* e, const FlyoWbpg* w);   /* flight_imitation only */
void flyo_set_initial_phase(FlyoEnv* e, double phase); /* WBPG phase drawn at the next reset (random_state.uniform()) */
void flyo_wbpg_state(const FlyoEnv* e, int* step, int* freq_idx, double* ctrl_freq);
/* vision_flight: the episode's terrain ([hfield_nrow, hfield_ncol] heights; trench centre line or n_trench = 0) and draws */
void flyo_set_terrain(FlyoEnv* e, const float* data, int n_trench, const double* trench_x, const double* trench_y);
void flyo_set_vision_targets(FlyoEnv* e, double height, double speed);
int flyo_env_reset(FlyoEnv* e, double* obs);         /* returns step_type (0 = FIRST) */
int flyo_env_step(FlyoEnv* e, const double* action, double* obs, double* reward, double* discount);

#ifdef __cplusplus
}
#endif
#endif
