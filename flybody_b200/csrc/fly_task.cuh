/* fly_task.cuh - sensors (m12, m14), integration (m13), the physics substep in dm_control's legacy order, and
 * the task layer (observation assembly, imitation reward, termination, per-env reset) of one environment.
 *
 * Task functions restate vnl_ray/tasks/{base,walk_imitation,walk_on_ball}.py, vnl_ray/tasks/rewards.py and
 * vnl_ray/fruitfly/fruitfly.py; each cites the lines it follows. */
#ifndef FLY_TASK_CUH_
#define FLY_TASK_CUH_
#include "fly_solver.cuh"

/* reference clips in HBM: packed float rows [qpos 7+nj | qvel 6+nj | root2site 3 ns | joint_quat 4 nj] */
struct FlyClips {
  const float* data;
  const int* offset;   /* first row of clip c */
  const int* length;   /* rows of clip c */
  int nclips;
};

/* wing-beat pattern generator tables in HBM (flybody_b200/wbpg.py; float64 so that table indices match the reference) */
struct FlyWbpg {
  int nfreq;                /* 0: no generator loaded */
  const double* beat_freqs; /* [nfreq] */
  const int* offsets;       /* [nfreq + 1] */
  const double* traj;       /* [rows, 6] */
  const double* phase;      /* [rows] */
  double base_freq, rel_range, rate, dt_ctrl;
};

/* ---------------------------------------------------------------- m12 acceleration-stage sensors */
template <class T>
FLY_INL T fly_ray_capsule(const T* pos, const T* mat, const T* size, const T* pnt, const T* vec) {
  T d[3] = {pnt[0] - pos[0], pnt[1] - pos[1], pnt[2] - pos[2]}, lp[3], lv[3];
  matT_vec(lp, mat, d); matT_vec(lv, mat, vec);
  const T r = size[0], h = size[1];
  T best = -1;
  const T a = lv[0] * lv[0] + lv[1] * lv[1], b = lp[0] * lv[0] + lp[1] * lv[1], c = lp[0] * lp[0] + lp[1] * lp[1] - r * r;
  const T zc = fmin_(h, fmax_(-h, lp[2]));
  if (lp[0] * lp[0] + lp[1] * lp[1] + (lp[2] - zc) * (lp[2] - zc) <= r * r) return 0;
  if (a > minval<T>()) {
    const T det = b * b - a * c;
    if (det >= 0) {
      const T x = (-b - sqrt(det)) / a;
      if (x >= 0 && fabs_(lp[2] + x * lv[2]) <= h) best = x;
    }
  }
  for (int sgn = -1; sgn <= 1; sgn += 2) {
    const T q[3] = {lp[0], lp[1], lp[2] - sgn * h};
    const T aa = dot3(lv, lv), bb = dot3(q, lv), cc = dot3(q, q) - r * r, det = bb * bb - aa * cc;
    if (det < 0 || aa < minval<T>()) continue;
    const T x = (-bb - sqrt(det)) / aa;
    if (x >= 0 && sgn * (lp[2] + x * lv[2]) >= h && (best < 0 || x < best)) best = x;
  }
  return best;
}

template <class T>
FLY_FN void fly_sensors_acc(const DModel<T>& m, const Ws<T>& w) {
  const int ncon = w.counts[FLY_C_NCON];
  bool need_rne = false;
  for (int s = 0; s < FLY_DIM(m, nsensor); s++) {
    const int t = m.sensor_type[s];
    if (t == FLYB_SENS_ACCELEROMETER || t == FLYB_SENS_FORCE) need_rne = true;
  }
  if (need_rne) {
    FLY_FOR(it, 6 * FLY_DIM(m, nbody)) w.cfrc_ext[it] = 0;
    FLY_SYNC();
    for (int ci = 0; ci < ncon; ci++) {
      const int ea = w.con_efc[ci];
      if (ea < 0) continue;
      const int dim = w.con_dim[ci], pg = m.pair_geom[w.con_pair[ci]];
      T fw[3] = {0, 0, 0};
      for (int k = 0; k < dim; k++) for (int a = 0; a < 3; a++) fw[a] += w.con_frame[9 * ci + 3 * k + a] * w.efc_force[ea + k];
      FLY_FOR(it, 12) {
        const int side = it / 6, c = it % 6;
        const int b = side ? m.geom_bodyid[(pg >> 8) & 255] : m.geom_bodyid[pg & 255];
        if (b > 0) {
          const T sg = side ? T(1) : T(-1);
          const T* c0 = w.scom + 3 * m.body_rootid[b];
          const T off[3] = {w.con_pos[3 * ci] - c0[0], w.con_pos[3 * ci + 1] - c0[1], w.con_pos[3 * ci + 2] - c0[2]};
          T tq[3];
          cross3(tq, off, fw);
          w.cfrc_ext[6 * b + c] += sg * (c < 3 ? tq[c] : fw[c - 3]);
        }
      }
      FLY_SYNC();
    }
    fly_rne(m, w, (T*)0, 1, 1);
  }
  FLY_FOR(s, FLY_DIM(m, nsensor)) {
    const int site = m.sensor_site[s], b = m.site_bodyid[site], type = m.sensor_type[s];
    T* out = w.sensordata + m.sensor_adr[s];
    const T *sp = w.site_xpos + 3 * site, *sm = w.site_xmat + 9 * site;
    if (type == FLYB_SENS_ACCELEROMETER) {
      const T* c0 = w.scom + 3 * m.body_rootid[b];
      const T* ca = w.cacc + 6 * b;
      T off[3] = {sp[0] - c0[0], sp[1] - c0[1], sp[2] - c0[2]}, t[3], lin[3], lv[6], cor[3], la[3];
      cross3(t, ca, off);
      for (int k = 0; k < 3; k++) lin[k] = ca[3 + k] + t[k];
      matT_vec(la, sm, lin);
      fly_local_velocity(m, w, lv, b, sp, sm);
      cross3(cor, lv, lv + 3);
      for (int k = 0; k < 3; k++) out[k] = la[k] + cor[k];
    } else if (type == FLYB_SENS_FORCE) {
      matT_vec(out, sm, w.cfrc_int + 6 * b + 3);
    } else if (type == FLYB_SENS_TOUCH) {
      T sum = 0;
      for (int ci = 0; ci < ncon; ci++) {
        const int ea = w.con_efc[ci];
        if (ea < 0) continue;
        const int pg = m.pair_geom[w.con_pair[ci]];
        const int b1 = m.geom_bodyid[pg & 255], b2 = m.geom_bodyid[(pg >> 8) & 255];
        if (b1 != b && b2 != b) continue;
        const T fn = w.efc_force[ea];
        if (fn <= 0) continue;
        T ray[3] = {w.con_frame[9 * ci], w.con_frame[9 * ci + 1], w.con_frame[9 * ci + 2]};
        if (b2 == b) { ray[0] = -ray[0]; ray[1] = -ray[1]; ray[2] = -ray[2]; }
        if (fly_ray_capsule(sp, sm, m.site_size + 3 * site, w.con_pos + 3 * ci, ray) >= 0) sum += fn;
      }
      out[0] = sum;
    }
  }
  FLY_SYNC();
}

/* m14: gyro, velocimeter */
template <class T>
FLY_FN void fly_sensors_vel(const DModel<T>& m, const Ws<T>& w) {
  FLY_FOR(s, FLY_DIM(m, nsensor)) {
    const int site = m.sensor_site[s], type = m.sensor_type[s];
    if (type == FLYB_SENS_GYRO || type == FLYB_SENS_VELOCIMETER) {
      T lv[6];
      fly_local_velocity(m, w, lv, m.site_bodyid[site], w.site_xpos + 3 * site, w.site_xmat + 9 * site);
      const T* src = type == FLYB_SENS_GYRO ? lv : lv + 3;
      for (int k = 0; k < 3; k++) w.sensordata[m.sensor_adr[s] + k] = src[k];
    }
  }
  FLY_SYNC();
}

/* ---------------------------------------------------------------- pipeline pieces
 * Block rendezvous points (FLY_STAGE_SYNC, one bit of FLY_SYNC_MASK each): between all stages of fly_step1 and
 * fly_forward_acc and before fly_integrate; every warp passes each of them once per call whatever its env does.  A warp
 * that does less than a full control step (reset, masked or idle) makes the difference up with fly_block_pad.
 * Measured on B200 (16 warps per SM): without any rendezvous the kernel is 3.4x slower than with seven per substep. */
/* rendezvous bits: fly_step1 0,7,8,1,9,10,2,11,12,13 (in program order) ; fly_forward_acc 3,14,4,15,16,17,5 ; fly_integrate 6 */
#define FLY_SYNC_BITS_STEP1 0x3f87u
#define FLY_SYNC_BITS_FWD 0x3c038u
#define FLY_SYNC_BITS_INT 0x40u
FLY_INL constexpr int fly_popc(unsigned x) { return x ? (int)(x & 1u) + fly_popc(x >> 1) : 0; }
#define FLY_NB_STEP1 fly_popc(FLY_SYNC_MASK & FLY_SYNC_BITS_STEP1)
#define FLY_NB_FWD fly_popc(FLY_SYNC_MASK & FLY_SYNC_BITS_FWD)
#define FLY_NB_INT fly_popc(FLY_SYNC_MASK & FLY_SYNC_BITS_INT)
template <class T>
FLY_INL int fly_nb_control_step(const DModel<T>& m) {
  return (1 + m.n_substeps) * FLY_NB_STEP1 + m.n_substeps * (FLY_NB_FWD + FLY_NB_INT);
}
FLY_INL void fly_block_pad(int n) {
  for (int k = 0; k < n; k++) FLY_BLOCK_SYNC();
}

template <class T>
FLY_FN void fly_step1(const DModel<T>& m, const Ws<T>& w, bool full) {
  FLY_STAGE_SYNC(0);
  fly_kinematics(m, w);
  FLY_STAGE_SYNC(7);
  fly_com_pos(m, w);
  FLY_STAGE_SYNC(8);
  if (full) fly_collision(m, w);  /* before the factorisations: its convex search borrows their storage (qLD | qH) */
  FLY_STAGE_SYNC(1);
  if (full) fly_crb(m, w);
  FLY_STAGE_SYNC(9);
  if (full) {
    {  /* M and M + dt * damping (both written by fly_crb) are factorised together */
      T* LD[2] = {w.qLD, w.qH};
      T* iv[2] = {w.qDinv, w.qHinv};
      fly_factor<T, 2>(m, LD, iv);
    }
  }
  FLY_STAGE_SYNC(10);
  if (full) fly_make_constraint(m, w);
  FLY_STAGE_SYNC(2);
  fly_com_vel(m, w);
  FLY_STAGE_SYNC(11);
  if (full) fly_passive(m, w);
  FLY_STAGE_SYNC(12);
  if (full) fly_rne(m, w, w.qfrc_bias, 0, 0);
  FLY_STAGE_SYNC(13);
  fly_sensors_vel(m, w);
}

/* actuation, smooth acceleration, constraint solve, acceleration sensors */
template <class T>
FLY_FN void fly_forward_acc(const DModel<T>& m, const Ws<T>& w, bool actuate = true) {
  FLY_STAGE_SYNC(3);
  fly_actuation(m, w, actuate);
  FLY_FOR(d, FLY_DIM(m, nv)) {
    const T f = w.qfrc_passive[d] - w.qfrc_bias[d] + w.qfrc_actuator[d];
    w.qfrc_smooth[d] = f; w.qacc_smooth[d] = f;
  }
  FLY_SYNC();
  FLY_STAGE_SYNC(14);
  {
    const T* LD[1] = {w.qLD};
    const T* iv[1] = {w.qDinv};
    T* x[1] = {w.qacc_smooth};
    fly_solve<T, 1>(m, w, LD, iv, x);
  }
  FLY_STAGE_SYNC(4);
#ifdef FLY_GPU
  const long long t0 = clock64();
#endif
  fly_fwd_constraint(m, w);
#ifdef FLY_GPU
  /* cycles this env spent in its constraint solve, summed over the control step: the load-balancing key of the next
     step's launch order (flyb_cuda.cu: fly_bin_kernel) */
  FLY_LANE0 w.counts[FLY_C_WORK] += (int)((clock64() - t0) >> 6);
#endif
  FLY_STAGE_SYNC(5);
  fly_sensors_acc(m, w);
}

/* m13: Euler with implicit joint damping */
template <class T>
FLY_FN void fly_integrate(const DModel<T>& m, const Ws<T>& w) {
  const T dt = m.timestep;
  FLY_STAGE_SYNC(6);
  /* sMa = (M + dt B)^-1 (qfrc_smooth + qfrc_constraint), solved together with qacc in fly_fwd_constraint */
  FLY_FOR(i, FLY_DIM(m, na)) w.act[i] += dt * w.act_dot[i];
  FLY_FOR(d, FLY_DIM(m, nv)) { w.qvel[d] += dt * w.sMa[d]; w.warm[d] = w.qacc[d]; }
  FLY_SYNC();
  FLY_FOR(j, FLY_DIM(m, njnt)) {
    int qa = m.jnt_qposadr[j], da = m.jnt_dofadr[j];
    const int type = m.jnt_type[j];
    if (type == FLYB_JNT_HINGE) w.qpos[qa] += dt * w.qvel[da];
    else {
      if (type == FLYB_JNT_FREE) { for (int k = 0; k < 3; k++) w.qpos[qa + k] += dt * w.qvel[da + k]; qa += 3; da += 3; }
      quat_integrate(w.qpos + qa, w.qvel + da, dt);
    }
  }
  FLY_SYNC();
}

template <class T>
FLY_INL void fly_accumulate(const DModel<T>& m, const Ws<T>& w) {
  FLY_FOR(k, FLY_DIM(m, nsensordata)) w.sens_sum[k] += w.sensordata[k];
  FLY_SYNC();
}
#endif
