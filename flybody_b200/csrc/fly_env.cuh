/* fly_env.cuh - one control step (or reset) of one environment: composer.Environment.step semantics
 *  with the task hooks of vnl_ray/tasks fused in. */
#ifndef FLY_ENV_CUH_
#define FLY_ENV_CUH_
#include "fly_task.cuh"
#include "fly_hfield.cuh"

/* per-env persistent integers */
enum { FLY_SI_STEP = 0, FLY_SI_EPISODE_STEPS, FLY_SI_RESET_NEXT, FLY_SI_CLIP, FLY_SI_EPISODE, FLY_SI_FORCED_CLIP,
       FLY_SI_EPLEN, FLY_SI_WB_STEP, FLY_SI_WB_FIDX, FLY_SI_WB_FREQ_LO, FLY_SI_WB_FREQ_HI, FLY_SI_COUNT = 12 };
/* per-env trace integers (optional debug output) */
enum { FLY_TR_NCON = 0, FLY_TR_NEFC, FLY_TR_NLIMIT, FLY_TR_NITER, FLY_TR_OVERFLOW, FLY_TR_NCAND, FLY_TR_NCON0, FLY_TR_PAIRS = 8,
       FLY_TR_PAIRS0 = 8 + FLY_MAXCON,   /* contact pair list of the state the control step STARTED from (its first collision pass) */
       FLY_TR_INTS = 8 + 2 * FLY_MAXCON };

#define FLY_NSTAT 8
template <class T>
struct FlyIO {
  T* sreal;            /* [N, sreal_stride]: qpos | qvel | act | qacc_warmstart | episode return */
  int* sint;           /* [N, FLY_SI_COUNT] */
  int sreal_stride;
  const float* action; /* [N, action_dim] */
  float* obs;          /* [N, obs_dim] */
  float* reward;       /* [N] */
  float* discount;     /* [N] */
  int* step_type;      /* [N] 0 FIRST, 1 MID, 2 LAST */
  int* trace_int;      /* nullable [N, FLY_TR_INTS] */
  T* trace_real;       /* nullable [N, trace_real_stride]: con_dist[MAXCON] | qacc[nv] | sensordata | efc_force[MAXEFC] */
  int trace_real_stride;
  float* stats;        /* nullable [N, FLY_NSTAT] per-env sums: return, length, fatal, count of finished episodes; control steps with a
                          capacity overflow (contacts / rows dropped); reset passes; control steps */
  unsigned long long seed;
  int env_offset;      /* global id of env 0 of this shard */
  int force_reset;     /* 1: reset every env in mask regardless of reset_next */
  const unsigned char* reset_mask; /* nullable */
  const double* wbpg_phase;        /* nullable [N]: initial wing-beat phase to use at reset instead of the hashed draw */
  int* work;                       /* nullable [N]: cycles/64 the env spent in its constraint solves this step (load-balancing key) */
  float* eyes;                     /* nullable [N, 24]: world frames of the right / left eye camera (vision_flight) */
  int* eye_terrain;                /* [N] with eyes: index of the env's terrain in the pool */
  /* nullable, flyb_step_host: device-visible addresses of page-locked HOST buffers.  A warp copies its env's finished row (and
     reward, discount, step type) there itself - posted writes over PCIe, under the other envs' steps - so that no device -> host
     copy is left after the launch */
  float* host_obs; float* host_reward; float* host_discount; int* host_step_type;
};

FLY_INL unsigned long long fly_hash(unsigned long long x) {  /* splitmix64 finaliser */
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

template <class T> FLY_INL void quat_recip(T* r, const T* q) {  /* quaternions.py:88-102 */
  const T n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  r[0] = q[0] / n2; r[1] = -q[1] / n2; r[2] = -q[2] / n2; r[3] = -q[3] / n2;
}
template <class T> FLY_INL void quat_rotate(T* r, const T* v, const T* q) {  /* quaternions.py:105-134 */
  T qi[4], va[4] = {0, v[0], v[1], v[2]}, t[4], o[4];
  quat_recip(qi, q);
  quat_mul(t, va, qi); quat_mul(o, q, t);
  r[0] = o[1]; r[1] = o[2]; r[2] = o[3];
}
template <class T> FLY_INL void quat_z2vec(T* q, const T* vin) {  /* quaternions.py:205-252 */
  T v[3] = {vin[0], vin[1], vin[2]};
  const bool edge = (v[0] == 0 && v[1] == 0);
  if (edge) v[0] = 1;
  const T n = norm3(v);
  v[0] /= n; v[1] /= n; v[2] /= n;
  if (edge) { if (v[2] < 0) { q[0] = 0; q[1] = 1; q[2] = 0; q[3] = 0; } else { q[0] = 1; q[1] = q[2] = q[3] = 0; } return; }
  T ax[2] = {-v[1], v[0]};
  const T an = sqrt(ax[0] * ax[0] + ax[1] * ax[1]);
  ax[0] /= an; ax[1] /= an;
  const T ang = acos(v[2]);
  q[0] = cos(ang / 2); q[1] = sin(ang / 2) * ax[0]; q[2] = sin(ang / 2) * ax[1]; q[3] = 0;
}
template <class T> FLY_INL T quat_dist_short_arc(const T* a, const T* b) {  /* quaternions.py:273-295 */
  const T na = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2] + a[3] * a[3]);
  const T nb = sqrt(b[0] * b[0] + b[1] * b[1] + b[2] * b[2] + b[3] * b[3]);
  const T d = (a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3]) / (na * nb);
  T x = 2 * d * d - 1;
  if (x > 1) x = 1;
  return acos(x);
}

struct FlyClipView { const float* rows; int T; };
template <class T>
FLY_INL const float* fly_clip_row(const DModel<T>& m, const FlyClipView& c, int row) {
  if (row >= c.T) row = c.T - 1;
  return c.rows + (size_t)row * m.clip_width;
}

/* observation assembly: fruitfly.py:532-669, tasks/base.py:237-261, walk_on_ball.py:81-88 */
template <class T>
FLY_FN void fly_write_obs(const DModel<T>& m, const Ws<T>& w, const FlyClipView& clip, int step_counter, float* obs,
                          T sensor_scale) {
  const int* off = m.obs_offset;
  const int rb = m.root_body;
  const T *rp = w.xpos + 3 * rb, *rm = w.xmat + 9 * rb;
  FLY_FOR(s, FLY_DIM(m, nsensor)) {
    const int key = m.sensor_obs[2 * s], idx = m.sensor_obs[2 * s + 1];
    const int dim = m.sensor_type[s] == FLYB_SENS_TOUCH ? 1 : 3;
    if (off[key] >= 0)
      for (int k = 0; k < dim; k++) obs[off[key] + dim * idx + k] = (float)(w.sens_sum[m.sensor_adr[s] + k] * sensor_scale);
  }
  if (off[FLYB_OBS_ACTUATOR_ACTIVATION] >= 0) FLY_FOR(i, FLY_DIM(m, na)) obs[off[FLYB_OBS_ACTUATOR_ACTIVATION] + i] = (float)w.act[i];
  if (off[FLYB_OBS_APPENDAGES_POS] >= 0)
    FLY_FOR(i, m.n_appendages) {
      const T* sp = w.site_xpos + 3 * m.appendage_site[i];
      T d[3] = {sp[0] - rp[0], sp[1] - rp[1], sp[2] - rp[2]}, o[3];
      matT_vec(o, rm, d);
      for (int k = 0; k < 3; k++) obs[off[FLYB_OBS_APPENDAGES_POS] + 3 * i + k] = (float)o[k];
    }
  if (off[FLYB_OBS_BALL_QVEL] >= 0) FLY_FOR(k, 3) obs[off[FLYB_OBS_BALL_QVEL] + k] = (float)w.qvel[m.ball_dofadr + k];
  FLY_FOR(i, m.n_obs_joints) {
    obs[off[FLYB_OBS_JOINTS_POS] + i] = (float)w.qpos[m.obs_joint_qposadr[i]];
    obs[off[FLYB_OBS_JOINTS_VEL] + i] = (float)w.qvel[m.obs_joint_dofadr[i]];
  }
  FLY_FOR(k, 3) obs[off[FLYB_OBS_WORLD_ZAXIS] + k] = (float)rm[6 + k];
  if (off[FLYB_OBS_REF_DISPLACEMENT] >= 0 && clip.rows) {
    const T* rq = w.qpos + m.root_qposadr + 3;
    T qi[4];
    quat_recip(qi, rq);
    FLY_FOR(i, m.future_steps + 1) {
      const float* r = fly_clip_row(m, clip, step_counter + i);
      T d[3], o[3], q[4], rqt[4] = {(T)r[3], (T)r[4], (T)r[5], (T)r[6]};
      for (int k = 0; k < 3; k++) d[k] = (T)r[k] - w.qpos[m.root_qposadr + k];
      matT_vec(o, rm, d);
      quat_mul(q, qi, rqt);  /* quaternions.py:13-17 */
      for (int k = 0; k < 3; k++) obs[off[FLYB_OBS_REF_DISPLACEMENT] + 3 * i + k] = (float)o[k];
      for (int k = 0; k < 4; k++) obs[off[FLYB_OBS_REF_ROOT_QUAT] + 4 * i + k] = (float)q[k];
      if (i == 0) { for (int k = 0; k < 3; k++) w.obs_tmp[k] = o[k]; for (int k = 0; k < 4; k++) w.obs_tmp[4 + k] = q[k]; }
    }
  }
  FLY_SYNC();
}

/* DeepMimic-style reward: walk_imitation.py:151-163, rewards.py:36-111 */
template <class T>
FLY_FN T fly_walk_imitation_reward(const DModel<T>& m, const Ws<T>& w, const FlyClipView& clip, int step) {
  if (m.inference_mode || !clip.rows) return T(1);
  const int nj = m.n_mocap_joints, ns = m.n_mocap_sites;
  const float* row = fly_clip_row(m, clip, step);
  const float *rq = row, *rv = row + 7 + nj, *r2s = rv + 6 + nj, *rjq = r2s + 3 * ns;
  const T* q = w.qpos + m.root_qposadr;
  const T* v = w.qvel + m.root_dofadr;
  T d_com = 0, d_qvel = 0, d_r2s = 0, d_jq = 0;
  for (int k = 0; k < 3; k++) d_com += (q[k] - (T)rq[k]) * (q[k] - (T)rq[k]);
  T part_qvel = 0, part_jq = 0, part_r2s = 0;
  const T qc[4] = {q[3], -q[4], -q[5], -q[6]};
  T qr[4];
  quat_recip(qr, q + 3);
  FLY_FOR(i, nj) {
    const int j = m.mocap_joint[i];
    const T qj = w.qpos[m.jnt_qposadr[j]], vj = w.qvel[m.jnt_dofadr[j]];
    part_qvel += (vj - (T)rv[6 + i]) * (vj - (T)rv[6 + i]);
    T ax[3], q1[4], q2[4], jq[4];
    quat_rotate(ax, w.xaxis + 3 * j, qr);   /* rewards.py:47-48 */
    quat_z2vec(q1, ax);                       /* quaternions.py:298-321 */
    const T n = norm3(ax);
    const T a2[3] = {ax[0] / n, ax[1] / n, ax[2] / n};
    axis_angle_quat(q2, a2, qj);
    quat_mul(jq, q2, q1);
    const T ref[4] = {(T)rjq[4 * i], (T)rjq[4 * i + 1], (T)rjq[4 * i + 2], (T)rjq[4 * i + 3]};
    const T a = quat_dist_short_arc(jq, ref);
    part_jq += a * a;
  }
  FLY_FOR(i, ns) {
    const T* sp = w.site_xpos + 3 * m.mocap_site[i];
    T d[3] = {sp[0] - q[0], sp[1] - q[1], sp[2] - q[2]}, r[3];
    quat_rotate(r, d, qc);                    /* quaternions.py:137-159 */
    for (int k = 0; k < 3; k++) part_r2s += (r[k] - (T)r2s[3 * i + k]) * (r[k] - (T)r2s[3 * i + k]);
  }
  for (int k = 0; k < 6; k++) d_qvel += (v[k] - (T)rv[k]) * (v[k] - (T)rv[k]);
  {
    const T rqt[4] = {(T)rq[3], (T)rq[4], (T)rq[5], (T)rq[6]};
    const T a = quat_dist_short_arc(q + 3, rqt);
    d_jq += a * a;
  }
  d_qvel += warp_sum(part_qvel); d_jq += warp_sum(part_jq); d_r2s += warp_sum(part_r2s);
  const T s_com = T(0.078487), s_qvel = T(53.7801), s_r2s = T(0.0735), s_jq = T(1.2247);  /* rewards.py:99-104 */
  return T(20) * exp(T(-0.5) / (s_com * s_com) * d_com) * exp(T(-0.5) / (s_qvel * s_qvel) * d_qvel)
         * exp(T(-0.5) / (s_r2s * s_r2s) * d_r2s) * exp(T(-0.5) / (s_jq * s_jq) * d_jq);
}

template <class T>
FLY_INL void fly_set_ghost(const DModel<T>& m, const Ws<T>& w, const FlyClipView& clip, int row, bool with_vel) {
  if (m.ghost_qposadr < 0 || !clip.rows) return;  /* walk_imitation.py:119-136,138-144 */
  const float* r = fly_clip_row(m, clip, row);
  const int nj = m.task_type == FLYB_TASK_WALK_IMITATION ? m.n_mocap_joints : 0;
  FLY_FOR(k, 7) w.qpos[m.ghost_qposadr + k] = (T)r[k] + (k < 3 ? m.ghost_offset[k] : T(0));
  if (with_vel) FLY_FOR(k, 6) w.qvel[m.ghost_dofadr + k] = (T)r[7 + nj + k];
  FLY_SYNC();
}

template <class T>
FLY_INL FlyClipView fly_clip_view(const DModel<T>& m, const FlyClips& clips, int id) {
  FlyClipView v;
  if (!clips.data || id < 0 || id >= clips.nclips) { v.rows = 0; v.T = 0; return v; }
  v.rows = clips.data + (size_t)clips.offset[id] * m.clip_width;
  v.T = clips.length[id];
  return v;
}

/* ---- wing-beat pattern generator (tasks/pattern_generators.py:121-191), per-env state in sint[FLY_SI_WB_*].
 * All of it in float64 with numpy's first-minimum argmin, so the table step and frequency index are bit-identical to
 * the reference's for the same float32 action. */
FLY_INL void fly_put_double(int* lo_hi, double v) {
  unsigned long long b;
  memcpy(&b, &v, 8);
  lo_hi[0] = (int)(unsigned)(b & 0xffffffffull); lo_hi[1] = (int)(unsigned)(b >> 32);
}
FLY_INL double fly_get_double(const int* lo_hi) {
  const unsigned long long b = (unsigned long long)(unsigned)lo_hi[0] | ((unsigned long long)(unsigned)lo_hi[1] << 32);
  double v;
  memcpy(&v, &b, 8);
  return v;
}
/* separately rounded multiply / add: nvcc would otherwise contract a*b + c*d into an FMA, which differs from numpy's
 * result in the last bit and, accumulated in the low-pass state, would eventually pick another table row */
FLY_INL double fly_mul_rn(double a, double b) {
#ifdef FLY_GPU
  return __dmul_rn(a, b);
#else
  return a * b;
#endif
}
FLY_INL double fly_add_rn(double a, double b) {
#ifdef FLY_GPU
  return __dadd_rn(a, b);
#else
  return a + b;
#endif
}
FLY_INL int fly_wbpg_nearest_freq(const FlyWbpg& wb, double f) {
  double bv = 1e300; int bi = 0x7fffffff;
  FLY_FOR(i, wb.nfreq) { const double d = fabs(wb.beat_freqs[i] - f); if (d < bv) { bv = d; bi = i; } }
  warp_argmin(bv, bi);
  return bi;
}
/* row of frequency table `fidx` whose phase (optionally modulo 1) is nearest to `target` */
FLY_INL int fly_wbpg_seek(const FlyWbpg& wb, int fidx, double target, bool mod1) {
  const int o = wb.offsets[fidx], n = wb.offsets[fidx + 1] - o;
  double bv = 1e300; int bi = 0x7fffffff;
  FLY_FOR(i, n) {
    const double ph = mod1 ? fmod(wb.phase[o + i], 1.0) : wb.phase[o + i];
    const double d = fabs(target - ph);
    if (d < bv) { bv = d; bi = i; }
  }
  warp_argmin(bv, bi);
  return bi;
}
/* uniform [0,1) initial phase of (env, episode): the counter-based stand-in for random_state.uniform() (flight_imitation.py:137) */
FLY_INL double fly_wbpg_phase0(unsigned long long seed, int global_env, int episode) {
  const unsigned long long h = fly_hash(seed ^ fly_hash((((unsigned long long)global_env << 24) ^ (unsigned long long)episode) + 0x51ED270B7ull));
  return (double)(h >> 11) * (1.0 / 9007199254740992.0);
}

/* ---- vision-guided flight (tasks/vision_flight.py).  The episode's draws - which terrain of the pool, target height, target
 * speed (vision_flight.py:106-109; arenas/hills.py regenerates the terrain every episode) - are counter-based hashes of (seed,
 * global env id, episode), recomputed by every control step instead of stored (flybody_b200/fly_envs.py: vision_draws is the
 * host-side copy). */
template <class T>
FLY_INL void fly_vision_draws(const DModel<T>& m, unsigned long long seed, int global_env, int episode, int* terrain, T* height, T* speed) {
  const unsigned long long h = fly_hash(seed ^ fly_hash((((unsigned long long)global_env << 24) ^ (unsigned long long)episode) + 0x7E44A1ull));
  *terrain = m.nterrain > 0 ? (int)(h % (unsigned long long)m.nterrain) : 0;
  const double u1 = (double)(fly_hash(h + 1) >> 11) * (1.0 / 9007199254740992.0);
  const double u2 = (double)(fly_hash(h + 2) >> 11) * (1.0 / 9007199254740992.0);
  *height = (T)((double)m.vision_ranges[0] + ((double)m.vision_ranges[1] - (double)m.vision_ranges[0]) * u1);
  *speed = (T)((double)m.vision_ranges[2] + ((double)m.vision_ranges[3] - (double)m.vision_ranges[2]) * u2);
}
/* index of the first minimum of |linspace(lo, hi, n) - x|, as numpy.argmin returns it (vision_flight.py:100-101, 181): only the
 * grid points next to x can hold it */
FLY_INL int fly_nearest_linspace(double lo, double hi, int n, double x) {
  const double step = (hi - lo) / (n - 1);
  int k = (int)floor((x - lo) / step);
  int best = 0;
  double bv = 1e300;
  for (int i = k - 1; i <= k + 2; i++) {
    const int j = i < 0 ? 0 : (i > n - 1 ? n - 1 : i);
    const double a = j == n - 1 ? hi : j * step + lo, d = fabs(a - x);
    if (d < bv || (d == bv && j < best)) { bv = d; best = j; }
  }
  return best;
}
template <class T>
FLY_INL T fly_hfield_height(const DModel<T>& m, int terrain, T x, T y) {  /* get_hfield_height, vision_flight.py:92-104 */
  if (!m.hf_pool) return T(0);
  const double half = m.hfield_size[0];
  const int xi = fly_nearest_linspace(-half, half, m.hfield_ncol, (double)x), yi = fly_nearest_linspace(-half, half, m.hfield_ncol, (double)y);
  return (T)(m.hfield_size[2] * (double)m.hf_pool[((size_t)terrain * m.hfield_nrow + yi) * m.hfield_ncol + xi]);
}
template <class T> FLY_INL T fly_tol_linear(T x, T lo, T hi, T margin) {  /* dm_control rewards.tolerance, "linear", value_at_margin 0 */
  if (lo <= x && x <= hi) return T(1);
  if (margin == 0) return T(0);
  const T d = (x < lo ? lo - x : x - hi) / margin;
  return d < 1 ? 1 - d : T(0);
}
template <class T>
FLY_FN T fly_vision_reward(const DModel<T>& m, const Ws<T>& w, int terrain, T t_height, T t_speed) {  /* vision_flight.py:157-212 */
  const T* xpos = w.qpos + m.root_qposadr;
  const T* vel = w.qvel + m.root_dofadr;
  const T height = fly_tol_linear(xpos[2] - fly_hfield_height(m, terrain, xpos[0], xpos[1]), t_height, t_height, T(0.15));
  T center = 1;
  if (m.trench_meta) {
    const double* tm = m.trench_meta + 3 * terrain;
    const int nt = (int)tm[0];
    if (nt > 1 && tm[1] <= (double)xpos[0] && (double)xpos[0] <= tm[2]) {
      const int idx = fly_nearest_linspace(tm[1], tm[2], nt, (double)xpos[0]);
      const T c = (T)m.trench_y[(size_t)terrain * m.trench_stride + idx];
      center = fly_tol_linear(xpos[1], c, c, T(0.15));
    }
  }
  const T x_speed = vel[0] >= t_speed ? T(1) : fly_tol_linear(vel[0], t_speed, t_speed, T(1.1) * t_speed);
  const T speed = fly_tol_linear(norm3(vel), t_speed, t_speed, T(1.1) * t_speed);
  T side = 0;
  for (int s = 0; s < FLY_DIM(m, nsensor); s++)
    if (m.sensor_type[s] == FLYB_SENS_VELOCIMETER) side = w.sensordata[m.sensor_adr[s] + 1];
  const T side_speed = fly_tol_linear(side, T(0), T(0), T(10));
  const T* zax = w.xmat + 9 * m.root_body + 6;
  T zd = m.vision_target_zaxis[0] * zax[0] + m.vision_target_zaxis[1] * zax[1] + m.vision_target_zaxis[2] * zax[2];
  if (zd > 1) zd = 1;
  if (zd < -1) zd = -1;
  const T world_zaxis = fly_tol_linear(acos(zd), T(0), T(0), T(FLY_PI));
  return height * x_speed * speed * side_speed * world_zaxis * center;
}

/* The whole job of one warp for one env and one call: load the state, reset the episode OR advance it by one control
 * step, store state and outputs.  Reset and step share one pass through the pipeline (each stage has a single call
 * site, so that the fully inlined kernel holds one copy of it):
 *   reset: step1, forward_acc                                 (mj_forward of the initial state)
 *   step : step1, n_substeps x [forward_acc, integrate, step1] (dm_control's legacy order: mj_step2 then mj_step1) */
template <class T>
FLY_FN void fly_env_run(const DModel<T>& m, const Ws<T>& w, const FlyClips& clips, const FlyWbpg& wb, const FlyIO<T>& io,
                        int env) {
  if (env < 0) { fly_block_pad(fly_nb_control_step(m)); return; }  /* idle warp of a block that still has work */
  T* sr = io.sreal + (size_t)env * io.sreal_stride;
  int* sint = io.sint + (size_t)env * FLY_SI_COUNT;
  float* obs = io.obs + (size_t)env * m.obs_dim;
  FLY_LANE0 for (int k = 0; k < 16; k++) w.counts[k] = 0;
  const bool do_reset = io.force_reset ? (!io.reset_mask || io.reset_mask[env]) : sint[FLY_SI_RESET_NEXT] != 0;
  if (io.force_reset && !do_reset) { fly_block_pad(fly_nb_control_step(m)); return; }
  T *s_qpos = sr, *s_qvel = sr + FLY_DIM(m, nq), *s_act = s_qvel + FLY_DIM(m, nv), *s_warm = s_act + FLY_DIM(m, na), *s_ret = s_warm + FLY_DIM(m, nv);
  const int episode = sint[FLY_SI_EPISODE];
  int terrain = 0;
  T v_height = 0, v_speed = 0;
  if (m.task_type == FLYB_TASK_VISION_FLIGHT) {   /* the draws of the episode being started / continued */
    fly_vision_draws(m, io.seed, io.env_offset + env, do_reset ? episode : episode - 1, &terrain, &v_height, &v_speed);
    FLY_LANE0 w.counts[FLY_C_TERRAIN] = terrain;
  }
  FLY_SYNC();
  int clip_id, step = 0, episode_steps = m.max_episode_steps;
  FlyClipView clip;
  if (do_reset) {
    clip_id = sint[FLY_SI_FORCED_CLIP];
    if (clip_id < 0 && clips.nclips > 0)  /* trajectory_loaders.py:122,184: random_state.choice over clips */
      clip_id = (int)(fly_hash(io.seed ^ fly_hash(((unsigned long long)(io.env_offset + env) << 24) ^ (unsigned long long)episode))
                      % (unsigned long long)clips.nclips);
    clip = fly_clip_view(m, clips, clip_id);
    FLY_FOR(i, FLY_DIM(m, nq)) w.qpos[i] = m.qpos_reset[i];
    FLY_FOR(i, FLY_DIM(m, nv)) { w.qvel[i] = 0; w.warm[i] = 0; w.qacc[i] = 0; }
    FLY_FOR(i, FLY_DIM(m, na)) w.act[i] = 0;
    FLY_FOR(i, FLY_DIM(m, nu)) w.ctrl[i] = 0;
    FLY_FOR(i, FLY_DIM(m, nsensordata)) { w.sens_sum[i] = 0; w.sensordata[i] = 0; }
    FLY_SYNC();
    if (m.task_type == FLYB_TASK_WALK_IMITATION && clip.rows) {
      const int snippet = clip.T - m.future_steps - 1;  /* walk_imitation.py:101-102 */
      if (snippet < episode_steps) episode_steps = snippet;
      const float* r = clip.rows;                        /* walk_imitation.py:119 */
      FLY_FOR(k, 7) w.qpos[m.root_qposadr + k] = (T)r[k];
      FLY_FOR(i, m.n_mocap_joints) w.qpos[m.jnt_qposadr[m.mocap_joint[i]]] = (T)r[7 + i];
      FLY_SYNC();
    }
    if (m.task_type == FLYB_TASK_FLIGHT_IMITATION && clip.rows) {
      /* flight_imitation.py:106-108: traj_timesteps = min(T, round(time_limit / control_dt)) - (future_steps + 1) */
      episode_steps = (clip.T < m.max_episode_steps ? clip.T : m.max_episode_steps) - (m.future_steps + 1);
      const float* r = clip.rows;
      FLY_FOR(k, 7) w.qpos[m.root_qposadr + k] = (T)r[k];        /* :138 walker.set_pose */
      FLY_FOR(k, 3) w.qvel[m.root_dofadr + k] = (T)r[7 + k];     /* :145-147 linear velocity only */
      if (wb.nfreq > 0) {                                        /* :135-143 wbpg.reset(initial_phase, return_qvel) */
        const double phase0 = io.wbpg_phase ? io.wbpg_phase[env] : fly_wbpg_phase0(io.seed, io.env_offset + env, episode);
        const int fidx = fly_wbpg_nearest_freq(wb, wb.base_freq);
        const int ws = fly_wbpg_seek(wb, fidx, phase0, false);
        const double* row = wb.traj + (size_t)(wb.offsets[fidx] + ws) * 6;
        FLY_FOR(k, 6) {
          const int j = m.wing_joint[k];
          w.qpos[m.jnt_qposadr[j]] = (T)row[k];
          w.qvel[m.jnt_dofadr[j]] = (T)((row[6 + k] - row[k]) / wb.dt_ctrl);
        }
        FLY_LANE0 { sint[FLY_SI_WB_STEP] = ws; sint[FLY_SI_WB_FIDX] = fidx; fly_put_double(sint + FLY_SI_WB_FREQ_LO, wb.base_freq); }
      }
      FLY_SYNC();
    }
    if (m.task_type == FLYB_TASK_VISION_FLIGHT) {   /* vision_flight.py:114-146 */
      /* init_pos_x / y ranges are single points in the reference's configuration; initialize_qvel is off: the fly starts at rest */
      const T x0 = m.vision_ranges[4], y0 = m.vision_ranges[6];
      const T z0 = fly_hfield_height(m, terrain, x0, y0) + v_height;
      FLY_LANE0 {
        T* q = w.qpos + m.root_qposadr;
        q[0] = x0; q[1] = y0; q[2] = z0;
        for (int k = 0; k < 4; k++) q[3 + k] = m.vision_init_quat[k];
      }
      if (wb.nfreq > 0) {                                        /* :125-126, 134: wbpg.reset(initial_phase) -> wing angles */
        const double phase0 = io.wbpg_phase ? io.wbpg_phase[env] : fly_wbpg_phase0(io.seed, io.env_offset + env, episode);
        const int fidx = fly_wbpg_nearest_freq(wb, wb.base_freq);
        const int ws = fly_wbpg_seek(wb, fidx, phase0, false);
        const double* row = wb.traj + (size_t)(wb.offsets[fidx] + ws) * 6;
        FLY_FOR(k, 6) w.qpos[m.jnt_qposadr[m.wing_joint[k]]] = (T)row[k];
        FLY_LANE0 { sint[FLY_SI_WB_STEP] = ws; sint[FLY_SI_WB_FIDX] = fidx; fly_put_double(sint + FLY_SI_WB_FREQ_LO, wb.base_freq); }
      }
      FLY_SYNC();
    }
  } else {
    FLY_FOR(i, FLY_DIM(m, nq)) w.qpos[i] = s_qpos[i];
    FLY_FOR(i, FLY_DIM(m, nv)) { w.qvel[i] = s_qvel[i]; w.warm[i] = s_warm[i]; }
    FLY_FOR(i, FLY_DIM(m, na)) w.act[i] = s_act[i];
    const float* action_in = io.action + (size_t)env * m.action_dim;
    FLY_FOR(a, FLY_DIM(m, nu)) {  /* fruitfly.py:480-492 ; NaN -> 0 walk_imitation.py:147 */
      const int ai = m.ctrl_action[a];
      T v = ai >= 0 ? (T)action_in[ai] : T(0);
      if (v != v) v = 0;
      w.ctrl[a] = v;
    }
    FLY_FOR(k, FLY_DIM(m, nsensordata)) w.sens_sum[k] = 0;
    if ((m.task_type == FLYB_TASK_FLIGHT_IMITATION || m.task_type == FLYB_TASK_VISION_FLIGHT) && wb.nfreq > 0 && m.user_action >= 0) {
      /* flight_imitation.py:149-159, vision_flight.py:148-155: requested beat frequency from the user action, generator step, and the wing
         position targets turned into force commands: action[wings] += target - wing qpos */
      float au = action_in[m.user_action];
      if (au != au) au = 0;
      /* float64 from the float32 action, as numpy 1.26.4 (the reference's pin) evaluates it (flybody_b200/wbpg.py) */
      const double req = fly_mul_rn(wb.base_freq, fly_add_rn(1.0, fly_mul_rn(wb.rel_range, (double)au)));
      int fidx = sint[FLY_SI_WB_FIDX], ws = sint[FLY_SI_WB_STEP];
      double cf = fly_get_double(sint + FLY_SI_WB_FREQ_LO);
      const int o = wb.offsets[fidx], n = wb.offsets[fidx + 1] - o;
      ws = (ws + 1) % n;
      cf = wb.rate == 0.0 ? req : fly_add_rn(fly_mul_rn(cf, wb.rate), fly_mul_rn(req, 1 - wb.rate));
      const int nf = fly_wbpg_nearest_freq(wb, cf);
      if (nf != fidx) {
        ws = fly_wbpg_seek(wb, nf, fmod(wb.phase[o + ws], 1.0), true);
        fidx = nf;
      }
      const double* row = wb.traj + (size_t)(wb.offsets[fidx] + ws) * 6;
      FLY_SYNC();
      FLY_FOR(k, 6) {
        const T delta = (T)(row[k] - (double)w.qpos[m.jnt_qposadr[m.wing_joint[k]]]);
        for (int a = 0; a < FLY_DIM(m, nu); a++) if (m.ctrl_action[a] == m.wing_action[k]) w.ctrl[a] += delta;
      }
      FLY_SYNC();
      FLY_LANE0 { sint[FLY_SI_WB_STEP] = ws; sint[FLY_SI_WB_FIDX] = fidx; fly_put_double(sint + FLY_SI_WB_FREQ_LO, cf); }
    }
    clip_id = sint[FLY_SI_CLIP];
    clip = fly_clip_view(m, clips, clip_id);
    step = sint[FLY_SI_STEP];
    episode_steps = sint[FLY_SI_EPISODE_STEPS];
    FLY_SYNC();
  }
  if (m.task_type != FLYB_TASK_WALK_ON_BALL) fly_set_ghost(m, w, clip, step, !do_reset);  /* no-op without ghost or clip */
  if (!do_reset) step += 1;  /* tasks/base.py:192 */
  const int nsub = do_reset ? 0 : m.n_substeps;
  for (int it = 0;; it++) {
    /* the position stage of the state a control step ends in is the one its successor starts with, so collision, factorisation
       and constraint rows are left to the successor - except where the task reads the final state's contacts: the vision task
       ends the episode on an active floor contact (vision_flight.py:214-233) */
    fly_step1(m, w, it == 0 || it < nsub || m.task_type == FLYB_TASK_VISION_FLIGHT);
    if (it == 0 && io.trace_int) {
      int* ti = io.trace_int + (size_t)env * FLY_TR_INTS;
      FLY_LANE0 ti[FLY_TR_NCON0] = w.counts[FLY_C_NCON];
      FLY_FOR(k, FLY_MAXCON) ti[FLY_TR_PAIRS0 + k] = k < w.counts[FLY_C_NCON] ? w.con_pair[k] : -1;
    }
    if (it > 0) fly_accumulate(m, w);
    if (it == nsub && !do_reset) break;
    /* the forward pass of a reset runs with actuation disabled (dm_control Physics.after_reset) and, like every
       mj_fwdConstraint, leaves its qacc as the next warm start */
    fly_forward_acc(m, w, !do_reset);
    if (do_reset) {
      FLY_FOR(d, FLY_DIM(m, nv)) w.warm[d] = w.qacc[d];
      fly_accumulate(m, w);
      break;
    }
    fly_integrate(m, w);
  }
  if (do_reset) fly_block_pad(fly_nb_control_step(m) - FLY_NB_STEP1 - FLY_NB_FWD);
  int st = 0, fatal = 0;
  T reward = 0, discount = 1;
  fly_write_obs(m, w, clip, step, obs, (do_reset && m.pad_first_obs) ? T(1) : T(1) / T(m.n_substeps));
  if (m.task_type == FLYB_TASK_VISION_FLIGHT) {
    if (m.obs_offset[FLYB_OBS_TASK_INPUT] >= 0) FLY_LANE0 {   /* vision_flight.py:243-252 */
      obs[m.obs_offset[FLYB_OBS_TASK_INPUT]] = (float)v_height; obs[m.obs_offset[FLYB_OBS_TASK_INPUT] + 1] = (float)v_speed;
    }
    if (io.eyes) {   /* camera frames of the two eyes for the ray caster (flyb_render): position, row-major orientation */
      float* eo = io.eyes + (size_t)env * 24;
      FLY_FOR(k, 24) {
        const int s = m.eye_site[k / 12], c = k % 12;
        eo[k] = (float)(c < 3 ? w.site_xpos[3 * s + c] : w.site_xmat[9 * s + c - 3]);
      }
      FLY_LANE0 io.eye_terrain[env] = terrain;
    }
  }
  if (do_reset) {
    FLY_LANE0 {
      sint[FLY_SI_STEP] = 0; sint[FLY_SI_EPISODE_STEPS] = episode_steps; sint[FLY_SI_RESET_NEXT] = 0;
      sint[FLY_SI_CLIP] = clip_id; sint[FLY_SI_EPISODE] = episode + 1; sint[FLY_SI_EPLEN] = 0;
      sint[FLY_SI_FORCED_CLIP] = -1;   /* walk_imitation.py:86-89: the forced index holds for the next episode only */
      s_ret[0] = 0;
      if (io.stats) {
        float* o = io.stats + FLY_NSTAT * (size_t)env;
        o[5] += 1.0f;
        if (w.counts[FLY_C_OVERFLOW]) o[4] += 1.0f;
      }
    }
  } else {
    /* termination: tasks/base.py:214-217, walk_imitation.py:165-191, walk_on_ball.py:75-79 */
    T qn = 0;
    FLY_FOR(d, FLY_DIM(m, nv)) qn += w.qacc[d] * w.qacc[d];
    qn = warp_sum(qn);
    int reached_end = 0;
    fatal = !(sqrt(qn) <= m.terminal_qacc);  /* NaN counts as divergence */
    if (m.task_type == FLYB_TASK_WALK_IMITATION || m.task_type == FLYB_TASK_WALK_ON_BALL) {  /* not flight, not template_task.py:78-83 */
      T lv[6];
      const int site = m.thorax_site;
      fly_local_velocity(m, w, lv, m.site_bodyid[site], w.site_xpos + 3 * site, w.site_xmat + 9 * site);
      if (norm3(lv + 3) > m.terminal_linvel || norm3(lv) > m.terminal_angvel) fatal = 1;
    }
    T r = 1;
    if (m.task_type == FLYB_TASK_WALK_IMITATION) {
      reached_end = step == episode_steps;
      if (clip.rows && norm3(w.obs_tmp) > m.terminal_com_dist) fatal = 1;
      r = fly_walk_imitation_reward(m, w, clip, step);
    } else if (m.task_type == FLYB_TASK_FLIGHT_IMITATION) {  /* flight_imitation.py:169-209 */
      reached_end = step == episode_steps;
      if (w.site_xpos[3 * m.thorax_site + 2] < m.terminal_height) fatal = 1;
      if (clip.rows) {
        if (norm3(w.obs_tmp) > m.terminal_com_dist) fatal = 1;
        if (m.ghost_body >= 0) {
          /* CoM term: ghost CoM = root2com(ghost pose read back from the physics) vs subtree CoM of the fly; orientation
             term: short-arc angle of ref_root_quat[0]; both linear tolerances (margin 0.4 cm / pi) */
          const T coff[3] = {T(-0.03697732), T(0.00029205), T(-0.0142447)};  /* task_utils.py:187 */
          T og[3], d[3];
          quat_rotate(og, coff, w.xquat + 4 * m.ghost_body);
          for (int k = 0; k < 3; k++) d[k] = w.xpos[3 * m.ghost_body + k] + og[k] - w.scom[3 * m.root_body + k];
          const T x = norm3(d) / T(0.4);
          /* quat_dist_short_arc([1,0,0,0], q) = arccos(2 (q0/|q|)^2 - 1) (quaternions.py:273-295) evaluated as
             2 atan2(|q_vec|, |q0|): the same angle, but well conditioned near 0 where fp32 arccos loses half its digits */
          const T* qd = w.obs_tmp + 4;
          const T th = 2 * atan2(norm3(qd + 1), fabs_(qd[0])) / T(3.14159265358979323846);
          r = (x < 1 ? 1 - x : T(0)) * (th < 1 ? 1 - th : T(0));
        }
      }
    } else if (m.task_type == FLYB_TASK_WALK_ON_BALL) {  /* walk_on_ball.py:61-73: tolerance(linear, margin 6) */
      const T tgt[3] = {0, -5, 0};
      for (int k = 0; k < 3; k++) {
        const T x = fabs_(w.qvel[m.ball_dofadr + k] - tgt[k]) / T(6);
        r *= x < 1 ? 1 - x : T(0);
      }
    } else if (m.task_type == FLYB_TASK_VISION_FLIGHT) {  /* vision_flight.py:214-233: an active contact with a world geom is fatal */
      int touch = 0;
      FLY_FOR(ci, w.counts[FLY_C_NCON]) {
        const int pg = m.pair_geom[w.con_pair[ci]];
        if (w.con_efc[ci] >= 0 && (m.geom_bodyid[pg & 255] == 0 || m.geom_bodyid[(pg >> 8) & 255] == 0)) touch = 1;
      }
      if (warp_sum_i(touch) > 0) fatal = 1;
      r = fly_vision_reward(m, w, terrain, v_height, v_speed);
    }
    reward = r;
    const int terminate = fatal || reached_end;
    discount = (m.task_type == FLYB_TASK_WALK_ON_BALL || m.task_type == FLYB_TASK_VISION_FLIGHT) ? (terminate ? T(0) : T(1))
                                                                                               : ((terminate && !reached_end) ? T(0) : T(1));
    st = (terminate || step >= m.time_limit_steps) ? 2 : 1;
    FLY_SYNC();
    FLY_LANE0 {
      sint[FLY_SI_STEP] = step;
      sint[FLY_SI_EPLEN] += 1;
      if (st == 2) sint[FLY_SI_RESET_NEXT] = 1;
      s_ret[0] += reward;
      if (io.stats) {
        float* o = io.stats + FLY_NSTAT * (size_t)env;
        if (st == 2) { o[0] += (float)s_ret[0]; o[1] += (float)sint[FLY_SI_EPLEN]; o[2] += (float)fatal; o[3] += 1.0f; }
        if (w.counts[FLY_C_OVERFLOW]) o[4] += 1.0f;   /* contacts or constraint rows were dropped somewhere in this control step */
        o[6] += 1.0f;
      }
    }
  }
  FLY_SYNC();
  FLY_FOR(i, FLY_DIM(m, nq)) s_qpos[i] = w.qpos[i];
  FLY_FOR(i, FLY_DIM(m, nv)) { s_qvel[i] = w.qvel[i]; s_warm[i] = w.warm[i]; }
  FLY_FOR(i, FLY_DIM(m, na)) s_act[i] = w.act[i];
  FLY_LANE0 {
    if (io.reward) io.reward[env] = (float)reward;
    if (io.discount) io.discount[env] = (float)discount;
    io.step_type[env] = st;
    if (io.host_reward) io.host_reward[env] = (float)reward;
    if (io.host_discount) io.host_discount[env] = (float)discount;
    if (io.host_step_type) io.host_step_type[env] = st;
  }
  if (io.host_obs) {   /* whole 128-byte segments per warp store: the row was assembled in device memory above */
    float* ho = io.host_obs + (size_t)env * m.obs_dim;
    FLY_FOR(i, m.obs_dim) ho[i] = obs[i];
  }
  if (io.work) FLY_LANE0 io.work[env] = w.counts[FLY_C_WORK];
  if (io.trace_int) {
    int* ti = io.trace_int + (size_t)env * FLY_TR_INTS;
    FLY_LANE0 {
      ti[FLY_TR_NCON] = w.counts[FLY_C_NCON]; ti[FLY_TR_NEFC] = w.counts[FLY_C_NEFC];
      ti[FLY_TR_NLIMIT] = w.counts[FLY_C_NLIMIT]; ti[FLY_TR_NITER] = w.counts[FLY_C_NITER];
      ti[FLY_TR_OVERFLOW] = w.counts[FLY_C_OVERFLOW]; ti[FLY_TR_NCAND] = w.counts[FLY_C_NCAND];
    }
    FLY_FOR(k, FLY_MAXCON) ti[FLY_TR_PAIRS + k] = k < w.counts[FLY_C_NCON] ? w.con_pair[k] : -1;
  }
  if (io.trace_real) {
    T* tr = io.trace_real + (size_t)env * io.trace_real_stride;
    FLY_FOR(k, FLY_MAXCON) tr[k] = k < w.counts[FLY_C_NCON] ? w.con_dist[k] : T(0);
    FLY_FOR(k, FLY_DIM(m, nv)) tr[FLY_MAXCON + k] = w.qacc[k];
    FLY_FOR(k, FLY_DIM(m, nsensordata)) tr[FLY_MAXCON + FLY_DIM(m, nv) + k] = w.sensordata[k];
    FLY_FOR(k, FLY_MAXEFC) tr[FLY_MAXCON + FLY_DIM(m, nv) + FLY_DIM(m, nsensordata) + k] = k < w.counts[FLY_C_NEFC] ? w.efc_force[k] : T(0);
  }
  FLY_SYNC();
}
#endif
