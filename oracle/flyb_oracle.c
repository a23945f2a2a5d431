/* flyb_oracle.c - see flyb_oracle.h.  TEST INFRASTRUCTURE ONLY; "parity unpinned" at the MuJoCo boundary.
 *
 * Restates, for one environment in float64, the path
 *   composer.Environment.step -> task hooks -> physics.step() x n_substeps -> observables -> reward
 * of vnl_ray/fly_envs.py:29-157.  Physics stages follow MuJoCo's documented pipeline (third-party,
 * not vendored); each task function cites the reference file:line.
 * Algorithms are the plain sequential textbook forms (dense-ish, tree-ordered loops); the CUDA kernels
 * in flybody_b200/csrc use a different, warp-cooperative decomposition of the same mathematics. */
#include "flyb_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#define MINVAL 1e-15
#define PI 3.14159265358979323846

struct FlyoEnv {
  const FlybModel* m;
  FlybModel mcopy;
  /* options that tests may toggle */
  int disable_contact, disable_limit, disable_fluid, disable_noslip, disable_actuation, disable_warmstart;
  /* state */
  double time;
  double *qpos, *qvel, *act, *ctrl, *qacc_warmstart;
  /* position stage */
  double *xpos, *xquat, *xmat, *xipos, *ximat, *xanchor, *xaxis, *geom_xpos, *geom_xmat, *site_xpos, *site_xmat;
  double *subtree_com, *cinert, *cdof, *crb, *qM, *qLD;
  double *ten_length, *act_length, *act_moment;
  int ncon;
  FlyoContact con[FLYO_MAXCON];
  int nefc, nlimit;
  double *efc_J, *efc_pos, *efc_margin, *efc_diagApprox, *efc_R, *efc_D, *efc_KBIP, *efc_vel, *efc_aref, *efc_force, *efc_b;
  int *efc_type, *efc_id, *efc_state;
  /* velocity stage */
  double *cvel, *cdof_dot, *qfrc_bias, *qfrc_passive, *act_velocity;
  /* acceleration stage */
  double *act_dot, *act_force, *qfrc_actuator, *qfrc_smooth, *qacc_smooth, *qacc, *qfrc_constraint;
  double *cacc, *cfrc_int, *cfrc_ext, *sensordata;
  int solver_niter;
  /* task layer */
  FlyoClip clip;
  int has_clip, step_counter, episode_steps, reset_next, nsub_done;
  double *sens_sum, *prev_action;
  FlyoWbpg wb;
  int has_wbpg, wb_step, wb_fidx;
  double wb_freq, init_phase;
  /* vision-guided flight: the episode's terrain and draws (tasks/vision_flight.py:106-112, arenas/hills.py) */
  const float* hf_data;
  int n_trench;
  const double *trench_x, *trench_y;
  double target_height, target_speed;
};

/* ---------------------------------------------------------------- small math */
static void cross3(double* r, const double* a, const double* b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static double norm3(const double* a) { return sqrt(dot3(a, a)); }
static double normalize3(double* a) {
  double n = norm3(a);
  if (n < MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; return 0; }
  a[0] /= n; a[1] /= n; a[2] /= n; return n;
}
static void quat_mul(double* r, const double* a, const double* b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
static void quat_normalize(double* q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n;
}
static void quat2mat(double* m, const double* q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
static void rot_vec_quat(double* r, const double* v, const double* q) {
  double m[9]; quat2mat(m, q);
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2],
         z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void mat_vec(double* r, const double* m, const double* v) {
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2],
         z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void matT_vec(double* r, const double* m, const double* v) {
  double x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2], y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2],
         z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void axis_angle_quat(double* q, const double* axis, double angle) {
  double s = sin(0.5 * angle);
  q[0] = cos(0.5 * angle); q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
/* q <- q * exp(0.5 * h * w), w in the local frame (quaternion integration of body-frame angular velocity) */
static void quat_integrate(double* q, const double* w, double h) {
  double ax[3] = {w[0], w[1], w[2]};
  double n = normalize3(ax);
  if (n * h == 0 || n < MINVAL) return;
  double dq[4], r[4];
  axis_angle_quat(dq, ax, n * h);
  quat_mul(r, q, dq);
  memcpy(q, r, sizeof(r));
  quat_normalize(q);
}

/* spatial algebra on [angular; linear] 6-vectors expressed about a common point */
static void cross_motion(double* r, const double* v, const double* s) { /* v x_m s */
  double a[3], b[3], c[3];
  cross3(a, v, s); cross3(b, v, s + 3); cross3(c, v + 3, s);
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
static void cross_force(double* r, const double* v, const double* f) { /* v x_f f */
  double a[3], b[3], c[3];
  cross3(a, v, f); cross3(b, v + 3, f + 3); cross3(c, v, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}
/* 10-number spatial inertia about a reference point: [Ixx Iyy Izz Ixy Ixz Iyz, m*c(3), m] */
static void inert_vec(double* r, const double* I, const double* v) {
  const double* h = I + 6; double m = I[9];
  double hv[3], wh[3];
  cross3(hv, h, v + 3); cross3(wh, v, h);
  r[0] = I[0] * v[0] + I[3] * v[1] + I[4] * v[2] + hv[0];
  r[1] = I[3] * v[0] + I[1] * v[1] + I[5] * v[2] + hv[1];
  r[2] = I[4] * v[0] + I[5] * v[1] + I[2] * v[2] + hv[2];
  r[3] = m * v[3] + wh[0]; r[4] = m * v[4] + wh[1]; r[5] = m * v[5] + wh[2];
}

/* ---------------------------------------------------------------- m1 kinematics */
static void kinematics(FlyoEnv* e) {
  const FlybModel* m = e->m;
  for (int b = 1; b < m->nbody; b++) {
    int p = m->body_parent[b];
    double pos[3], quat[4], t[3];
    mat_vec(t, e->xmat + 9 * p, m->body_pos + 3 * b);
    for (int k = 0; k < 3; k++) pos[k] = e->xpos[3 * p + k] + t[k];
    quat_mul(quat, e->xquat + 4 * p, m->body_quat + 4 * b);
    for (int j = m->body_jntadr[b]; j < m->body_jntadr[b] + m->body_jntnum[b]; j++) {
      int qa = m->jnt_qposadr[j], type = m->jnt_type[j];
      if (type == FLYB_JNT_FREE) {
        quat_normalize(e->qpos + qa + 3);
        memcpy(pos, e->qpos + qa, 3 * sizeof(double));
        memcpy(quat, e->qpos + qa + 3, 4 * sizeof(double));
        memcpy(e->xanchor + 3 * j, pos, 3 * sizeof(double));
        rot_vec_quat(e->xaxis + 3 * j, m->jnt_axis + 3 * j, quat);
        continue;
      }
      rot_vec_quat(t, m->jnt_pos + 3 * j, quat);
      for (int k = 0; k < 3; k++) e->xanchor[3 * j + k] = pos[k] + t[k];
      rot_vec_quat(e->xaxis + 3 * j, m->jnt_axis + 3 * j, quat);
      double ql[4], nq[4];
      if (type == FLYB_JNT_BALL) {
        quat_normalize(e->qpos + qa);
        memcpy(ql, e->qpos + qa, sizeof(ql));
      } else {
        axis_angle_quat(ql, m->jnt_axis + 3 * j, e->qpos[qa] - m->qpos0[qa]);
      }
      quat_mul(nq, quat, ql);
      memcpy(quat, nq, sizeof(nq));
      rot_vec_quat(t, m->jnt_pos + 3 * j, quat);  /* keep the anchor fixed */
      for (int k = 0; k < 3; k++) pos[k] = e->xanchor[3 * j + k] - t[k];
    }
    quat_normalize(quat);
    memcpy(e->xpos + 3 * b, pos, sizeof(pos));
    memcpy(e->xquat + 4 * b, quat, sizeof(quat));
    quat2mat(e->xmat + 9 * b, quat);
    mat_vec(t, e->xmat + 9 * b, m->body_ipos + 3 * b);
    for (int k = 0; k < 3; k++) e->xipos[3 * b + k] = pos[k] + t[k];
    double iq[4];
    quat_mul(iq, quat, m->body_iquat + 4 * b);
    quat2mat(e->ximat + 9 * b, iq);
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_bodyid[g];
    double t[3], q[4];
    mat_vec(t, e->xmat + 9 * b, m->geom_pos + 3 * g);
    for (int k = 0; k < 3; k++) e->geom_xpos[3 * g + k] = e->xpos[3 * b + k] + t[k];
    quat_mul(q, e->xquat + 4 * b, m->geom_quat + 4 * g);
    quat2mat(e->geom_xmat + 9 * g, q);
  }
  for (int s = 0; s < m->nsite; s++) {
    int b = m->site_bodyid[s];
    double t[3], q[4];
    mat_vec(t, e->xmat + 9 * b, m->site_pos + 3 * s);
    for (int k = 0; k < 3; k++) e->site_xpos[3 * s + k] = e->xpos[3 * b + k] + t[k];
    quat_mul(q, e->xquat + 4 * b, m->site_quat + 4 * s);
    quat2mat(e->site_xmat + 9 * s, q);
  }
}

/* ---------------------------------------------------------------- m2 comPos */
static void com_pos(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nb = m->nbody;
  for (int b = 0; b < nb; b++)
    for (int k = 0; k < 3; k++) e->subtree_com[3 * b + k] = m->body_mass[b] * e->xipos[3 * b + k];
  for (int b = nb - 1; b > 0; b--)
    for (int k = 0; k < 3; k++) e->subtree_com[3 * m->body_parent[b] + k] += e->subtree_com[3 * b + k];
  for (int b = 0; b < nb; b++) {
    if (m->body_subtreemass[b] < MINVAL) memcpy(e->subtree_com + 3 * b, e->xipos + 3 * b, 3 * sizeof(double));
    else for (int k = 0; k < 3; k++) e->subtree_com[3 * b + k] /= m->body_subtreemass[b];
  }
  for (int b = 1; b < nb; b++) {
    const double* R = e->ximat + 9 * b; const double* I = m->body_inertia + 3 * b;
    const double* c0 = e->subtree_com + 3 * m->body_rootid[b];
    double d[3], mass = m->body_mass[b], *ci = e->cinert + 10 * b;
    for (int k = 0; k < 3; k++) d[k] = e->xipos[3 * b + k] - c0[k];
    double dd = dot3(d, d);
    /* R diag(I) R^T + m (|d|^2 1 - d d^T) */
    double W[9];
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++)
      W[3 * r + c] = R[3 * r] * I[0] * R[3 * c] + R[3 * r + 1] * I[1] * R[3 * c + 1] + R[3 * r + 2] * I[2] * R[3 * c + 2]
                     + mass * ((r == c ? dd : 0) - d[r] * d[c]);
    ci[0] = W[0]; ci[1] = W[4]; ci[2] = W[8]; ci[3] = W[1]; ci[4] = W[2]; ci[5] = W[5];
    ci[6] = mass * d[0]; ci[7] = mass * d[1]; ci[8] = mass * d[2]; ci[9] = mass;
  }
  for (int j = 0; j < m->njnt; j++) {
    int b = m->jnt_bodyid[j], da = m->jnt_dofadr[j];
    const double* c0 = e->subtree_com + 3 * m->body_rootid[b];
    double off[3];
    for (int k = 0; k < 3; k++) off[k] = c0[k] - e->xanchor[3 * j + k];
    if (m->jnt_type[j] == FLYB_JNT_HINGE) {
      double* cd = e->cdof + 6 * da;
      memcpy(cd, e->xaxis + 3 * j, 3 * sizeof(double));
      cross3(cd + 3, e->xaxis + 3 * j, off);
    } else {
      if (m->jnt_type[j] == FLYB_JNT_FREE) {
        for (int k = 0; k < 3; k++) { double* cd = e->cdof + 6 * (da + k); memset(cd, 0, 6 * sizeof(double)); cd[3 + k] = 1; }
        da += 3;
      }
      for (int k = 0; k < 3; k++) {  /* rotational dofs: body-frame axes */
        double ax[3] = {e->xmat[9 * b + k], e->xmat[9 * b + 3 + k], e->xmat[9 * b + 6 + k]};
        double* cd = e->cdof + 6 * (da + k);
        memcpy(cd, ax, sizeof(ax));
        cross3(cd + 3, ax, off);
      }
    }
  }
}

/* ---------------------------------------------------------------- m4 CRB + factorisation */
static void crb(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nv = m->nv;
  memcpy(e->crb, e->cinert, 10 * m->nbody * sizeof(double));
  for (int b = m->nbody - 1; b > 0; b--)
    if (m->body_parent[b] > 0) for (int k = 0; k < 10; k++) e->crb[10 * m->body_parent[b] + k] += e->crb[10 * b + k];
  memset(e->qM, 0, (size_t)nv * nv * sizeof(double));
  for (int i = 0; i < nv; i++) {
    double f[6];
    inert_vec(f, e->crb + 10 * m->dof_bodyid[i], e->cdof + 6 * i);
    for (int j = i; j >= 0; j = m->dof_parentid[j]) {
      double s = 0;
      for (int k = 0; k < 6; k++) s += e->cdof[6 * j + k] * f[k];
      e->qM[i * nv + j] = s;
    }
    e->qM[i * nv + i] += m->dof_armature[i];
  }
}
/* in-place M = L^T D L over the tree sparsity; LD holds L below the diagonal, D on it */
static void factor_tree(const FlybModel* m, double* LD) {
  int nv = m->nv;
  for (int k = nv - 1; k >= 0; k--) {
    double dk = LD[k * nv + k];
    for (int i = m->dof_parentid[k]; i >= 0; i = m->dof_parentid[i]) {
      double t = LD[k * nv + i] / dk;
      for (int j = i; j >= 0; j = m->dof_parentid[j]) LD[i * nv + j] -= t * LD[k * nv + j];
      LD[k * nv + i] = t;
    }
  }
}
static void solve_tree(const FlybModel* m, const double* LD, double* x) {
  int nv = m->nv;
  for (int i = nv - 1; i >= 0; i--)
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) x[j] -= LD[i * nv + j] * x[i];
  for (int i = 0; i < nv; i++) x[i] /= LD[i * nv + i];
  for (int i = 0; i < nv; i++)
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) x[i] -= LD[i * nv + j] * x[j];
}
static void mul_M(const FlybModel* m, const double* M, double* r, const double* v) {
  int nv = m->nv;
  for (int i = 0; i < nv; i++) r[i] = 0;
  for (int i = 0; i < nv; i++) {
    r[i] += M[i * nv + i] * v[i];
    for (int j = m->dof_parentid[i]; j >= 0; j = m->dof_parentid[j]) {
      r[i] += M[i * nv + j] * v[j];
      r[j] += M[i * nv + j] * v[i];
    }
  }
}

/* ---------------------------------------------------------------- m5 collision (narrow phase primitives) */
typedef struct { double dist, pos[3], n[3], t[3]; int has_t; } Hit;

static void make_frame(double* fr, const double* n, const double* hint, int has_hint) {
  /* x = n; y = hint orthogonalised (or a canonical choice); z = x cross y */
  double x[3] = {n[0], n[1], n[2]}, y[3];
  normalize3(x);
  if (has_hint) { memcpy(y, hint, sizeof(y)); }
  if (!has_hint || norm3(y) < 0.5) {
    y[0] = 0; y[1] = 0; y[2] = 0;
    if (x[1] < 0.5 && x[1] > -0.5) y[1] = 1; else y[2] = 1;
  }
  double d = dot3(x, y);
  for (int k = 0; k < 3; k++) y[k] -= d * x[k];
  normalize3(y);
  memcpy(fr, x, sizeof(x)); memcpy(fr + 3, y, sizeof(y));
  cross3(fr + 6, x, y);
}

static int hit_sphere_sphere(Hit* h, const double* c1, double r1, const double* c2, double r2, double margin) {
  double d[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
  double len = norm3(d), dist = len - r1 - r2;
  if (dist > margin) return 0;
  if (len < MINVAL) { d[0] = 1; d[1] = 0; d[2] = 0; } else { d[0] /= len; d[1] /= len; d[2] /= len; }
  h->dist = dist; h->has_t = 0;
  for (int k = 0; k < 3; k++) { h->n[k] = d[k]; h->pos[k] = c1[k] + d[k] * (r1 + 0.5 * dist); }
  return 1;
}
static int hit_plane_sphere(Hit* h, const double* pp, const double* pn, const double* c, double r, double margin) {
  double d[3] = {c[0] - pp[0], c[1] - pp[1], c[2] - pp[2]};
  double dist = dot3(d, pn) - r;
  if (dist > margin) return 0;
  h->dist = dist; h->has_t = 0;
  for (int k = 0; k < 3; k++) { h->n[k] = pn[k]; h->pos[k] = c[k] - pn[k] * (r + 0.5 * dist); }
  return 1;
}
static int hit_plane_capsule(Hit* h, const double* pp, const double* pn, const double* c, const double* mat,
                             const double* size, double margin) {
  double ax[3] = {mat[2], mat[5], mat[8]}, e1[3], e2[3];
  for (int k = 0; k < 3; k++) { e1[k] = c[k] + ax[k] * size[1]; e2[k] = c[k] - ax[k] * size[1]; }
  int n = 0;
  n += hit_plane_sphere(h + n, pp, pn, e1, size[0], margin);
  n += hit_plane_sphere(h + n, pp, pn, e2, size[0], margin);
  for (int i = 0; i < n; i++) { memcpy(h[i].t, ax, sizeof(ax)); h[i].has_t = 1; }
  return n;
}
static int hit_plane_ellipsoid(Hit* h, const double* pp, const double* pn, const double* c, const double* mat,
                               const double* size, double margin) {
  /* deepest point = support point of the ellipsoid along -n */
  double nl[3], s[3], w[3];
  matT_vec(nl, mat, pn);
  for (int k = 0; k < 3; k++) s[k] = size[k] * size[k] * nl[k];
  double den = sqrt(size[0] * size[0] * nl[0] * nl[0] + size[1] * size[1] * nl[1] * nl[1] + size[2] * size[2] * nl[2] * nl[2]);
  if (den < MINVAL) return 0;
  for (int k = 0; k < 3; k++) s[k] = -s[k] / den;
  mat_vec(w, mat, s);
  double p[3] = {c[0] + w[0], c[1] + w[1], c[2] + w[2]};
  double d[3] = {p[0] - pp[0], p[1] - pp[1], p[2] - pp[2]};
  double dist = dot3(d, pn);
  if (dist > margin) return 0;
  h->dist = dist; h->has_t = 0;
  for (int k = 0; k < 3; k++) { h->n[k] = pn[k]; h->pos[k] = p[k] - pn[k] * 0.5 * dist; }
  return 1;
}
static int hit_plane_cylinder(Hit* h, const double* pp, const double* pn, const double* c, const double* mat,
                              const double* size, double margin) {
  double ax[3] = {mat[2], mat[5], mat[8]};
  double prjaxis = dot3(pn, ax);
  if (prjaxis > 0) { ax[0] = -ax[0]; ax[1] = -ax[1]; ax[2] = -ax[2]; prjaxis = -prjaxis; }
  double d0[3] = {c[0] - pp[0], c[1] - pp[1], c[2] - pp[2]};
  double dist0 = dot3(d0, pn);
  double vec[3];
  for (int k = 0; k < 3; k++) vec[k] = ax[k] * prjaxis - pn[k];
  double len = norm3(vec);
  if (len < 1e-12) { vec[0] = mat[0] * size[0]; vec[1] = mat[3] * size[0]; vec[2] = mat[6] * size[0]; }
  else for (int k = 0; k < 3; k++) vec[k] *= size[0] / len;
  double prjvec = dot3(vec, pn);
  double a[3] = {ax[0] * size[1], ax[1] * size[1], ax[2] * size[1]};
  prjaxis *= size[1];
  int n = 0;
  double dist = dist0 + prjaxis + prjvec;
  if (dist > margin) return 0;
  h[n].dist = dist; h[n].has_t = 0;
  for (int k = 0; k < 3; k++) { h[n].n[k] = pn[k]; h[n].pos[k] = c[k] + vec[k] + a[k] - pn[k] * 0.5 * dist; }
  n++;
  dist = dist0 - prjaxis + prjvec;
  if (dist <= margin) {
    h[n].dist = dist; h[n].has_t = 0;
    for (int k = 0; k < 3; k++) { h[n].n[k] = pn[k]; h[n].pos[k] = c[k] + vec[k] - a[k] - pn[k] * 0.5 * dist; }
    n++;
  }
  dist = dist0 + prjaxis - 0.5 * prjvec;
  if (dist <= margin) {
    double v1[3];
    cross3(v1, vec, ax);
    normalize3(v1);
    for (int k = 0; k < 3; k++) v1[k] *= size[0] * sqrt(3.0) * 0.5;
    for (int sgn = 1; sgn >= -1; sgn -= 2) {
      h[n].dist = dist; h[n].has_t = 0;
      for (int k = 0; k < 3; k++) { h[n].n[k] = pn[k]; h[n].pos[k] = c[k] + sgn * v1[k] + a[k] - 0.5 * vec[k] - pn[k] * 0.5 * dist; }
      n++;
    }
  }
  return n;
}
static int hit_sphere_capsule(Hit* h, const double* c1, double r1, const double* c2, const double* mat2,
                              const double* size2, double margin) {
  double ax[3] = {mat2[2], mat2[5], mat2[8]}, d[3] = {c1[0] - c2[0], c1[1] - c2[1], c1[2] - c2[2]};
  double x = dot3(ax, d);
  if (x > size2[1]) x = size2[1]; else if (x < -size2[1]) x = -size2[1];
  double p[3] = {c2[0] + ax[0] * x, c2[1] + ax[1] * x, c2[2] + ax[2] * x};
  return hit_sphere_sphere(h, c1, r1, p, size2[0], margin);
}
static int hit_capsule_capsule(Hit* h, const double* c1, const double* mat1, const double* size1, const double* c2,
                               const double* mat2, const double* size2, double margin) {
  double a1[3] = {mat1[2], mat1[5], mat1[8]}, a2[3] = {mat2[2], mat2[5], mat2[8]};
  double dif[3] = {c1[0] - c2[0], c1[1] - c2[1], c1[2] - c2[2]};
  double l1 = size1[1], l2 = size2[1];
  double ma = 1.0, mb = -dot3(a1, a2), mc = 1.0, u = -dot3(a1, dif), v = dot3(a2, dif);
  double det = ma * mc - mb * mb;
  double p1[3], p2[3];
  if (fabs(det) >= 1e-12) {
    double x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
    if (x1 > l1) { x1 = l1; x2 = (v - mb * l1) / mc; }
    else if (x1 < -l1) { x1 = -l1; x2 = (v + mb * l1) / mc; }
    if (x2 > l2) { x2 = l2; x1 = (u - mb * l2) / ma; if (x1 > l1) x1 = l1; else if (x1 < -l1) x1 = -l1; }
    else if (x2 < -l2) { x2 = -l2; x1 = (u + mb * l2) / ma; if (x1 > l1) x1 = l1; else if (x1 < -l1) x1 = -l1; }
    for (int k = 0; k < 3; k++) { p1[k] = c1[k] + a1[k] * x1; p2[k] = c2[k] + a2[k] * x2; }
    return hit_sphere_sphere(h, p1, size1[0], p2, size2[0], margin);
  }
  /* parallel axes: up to two contacts at the ends of the overlap */
  int n = 0;
  double x1s[2] = {l1, -l1};
  for (int i = 0; i < 2 && n < 2; i++) {
    double x1 = x1s[i], x2 = (v - mb * x1) / mc;
    if (x2 <= l2 && x2 >= -l2) {
      for (int k = 0; k < 3; k++) { p1[k] = c1[k] + a1[k] * x1; p2[k] = c2[k] + a2[k] * x2; }
      n += hit_sphere_sphere(h + n, p1, size1[0], p2, size2[0], margin);
    }
  }
  double x2s[2] = {l2, -l2};
  for (int i = 0; i < 2 && n < 2; i++) {
    double x2 = x2s[i], x1 = (u - mb * x2) / ma;
    if (x1 < l1 && x1 > -l1) {
      for (int k = 0; k < 3; k++) { p1[k] = c1[k] + a1[k] * x1; p2[k] = c2[k] + a2[k] * x2; }
      n += hit_sphere_sphere(h + n, p1, size1[0], p2, size2[0], margin);
    }
  }
  return n;
}
static int hit_sphere_cylinder(Hit* h, const double* c1, double r1, const double* c2, const double* mat2,
                               const double* size2, double margin) {
  double ax[3] = {mat2[2], mat2[5], mat2[8]}, d[3] = {c1[0] - c2[0], c1[1] - c2[1], c1[2] - c2[2]};
  double x = dot3(ax, d), pr[3];
  for (int k = 0; k < 3; k++) pr[k] = d[k] - ax[k] * x;
  double psq = dot3(pr, pr), R = size2[0], H = size2[1];
  if (fabs(x) <= H && psq >= R * R) {  /* side */
    double p[3] = {c2[0] + ax[0] * x, c2[1] + ax[1] * x, c2[2] + ax[2] * x};
    return hit_sphere_sphere(h, c1, r1, p, R, margin);
  }
  double sg = x >= 0 ? 1.0 : -1.0;
  if (psq <= R * R) {  /* cap: plane-sphere against the cap plane, normal flipped to point sphere -> cylinder */
    double pp[3], pn[3];
    for (int k = 0; k < 3; k++) { pn[k] = sg * ax[k]; pp[k] = c2[k] + pn[k] * H; }
    int n = hit_plane_sphere(h, pp, pn, c1, r1, margin);
    if (n) for (int k = 0; k < 3; k++) h->n[k] = -h->n[k];
    return n;
  }
  double pl = sqrt(psq), p[3];  /* rim */
  for (int k = 0; k < 3; k++) p[k] = c2[k] + sg * ax[k] * H + pr[k] * (R / pl);
  return hit_sphere_sphere(h, c1, r1, p, 0.0, margin);
}

/* ---- general convex pairs (MuJoCo mjc_Convex): ellipsoid and cylinder geoms off the plane.
 * MuJoCo (engine_collision_convex.c, third party, not vendored; the reference leaves the version unpinned, 3.1-3.2 by its
 * dates) hands these pairs to libccd's Minkowski Portal Refinement (ccdMPRPenetration, libccd src/mpr.c, vendored inside
 * MuJoCo) with mpr_tolerance 1e-6 and mpr_iterations 50, support functions per geom type, each object inflated by
 * margin/2, dist = margin - depth, then adjusts the normal from the geom shapes (mjc_fixNormal).  Restated here from the
 * published algorithm (G. Snethen, "XenoCollide", Game Programming Gems 7; libccd's structure: discoverPortal,
 * refinePortal, findPenetr, findPos).  "parity unpinned": no MuJoCo build is available to check against; the
 * known-answer tests compare it with the analytic sphere/capsule/cylinder routines above on degenerate ellipsoids. */
#define CCD_EPS 2.220446049250313e-16
#define MPR_TOLERANCE 1e-6
#define MPR_ITERATIONS 50
#define MPR_PORTAL_CAP 100  /* guard on libccd's open-ended portal loops (never reached in exact arithmetic) */
typedef struct { double v[3], v1[3], v2[3]; } Supp;
typedef struct { int type; const double *pos, *mat, *size; double margin; } CcdObj;
#define GEOM_PRISM 100   /* internal: one prism of a height field (size -> 6 vertices relative to pos) */

static int ccd_is_zero(double x) { return fabs(x) < CCD_EPS; }
static int ccd_eq(double a, double b) {
  double ab = fabs(a - b);
  if (ab < CCD_EPS) return 1;
  a = fabs(a); b = fabs(b);
  return b > a ? ab < CCD_EPS * b : ab < CCD_EPS * a;
}
static double sign0(double x) { return x > 0 ? 1.0 : (x < 0 ? -1.0 : 0.0); }
/* support point of one geom in world direction dir (mjccd_support) */
static void ccd_support_geom(const CcdObj* o, const double* dir, double* res) {
  double ld[3], r[3] = {0, 0, 0}, w[3];
  matT_vec(ld, o->mat, dir);
  switch (o->type) {
    case FLYB_GEOM_SPHERE: for (int k = 0; k < 3; k++) r[k] = ld[k] * o->size[0]; break;
    case FLYB_GEOM_CAPSULE:
      for (int k = 0; k < 3; k++) r[k] = ld[k] * o->size[0];
      r[2] += sign0(ld[2]) * o->size[1];
      break;
    case FLYB_GEOM_ELLIPSOID: {
      double t[3] = {ld[0] * o->size[0], ld[1] * o->size[1], ld[2] * o->size[2]};
      normalize3(t);
      for (int k = 0; k < 3; k++) r[k] = t[k] * o->size[k];
      break;
    }
    case FLYB_GEOM_CYLINDER: {
      double t = sqrt(ld[0] * ld[0] + ld[1] * ld[1]);
      if (t > MINVAL) { double sc = o->size[0] / t; r[0] = ld[0] * sc; r[1] = ld[1] * sc; }
      r[2] = sign0(ld[2]) * o->size[1];
      break;
    }
    case GEOM_PRISM: {   /* six vertices in size[0..18), given relative to pos: the farthest one */
      double best = -1e300;
      for (int v = 0; v < 6; v++) {
        const double* q = o->size + 3 * v;
        const double d = q[0] * ld[0] + q[1] * ld[1] + q[2] * ld[2];
        if (d > best) { best = d; r[0] = q[0]; r[1] = q[1]; r[2] = q[2]; }
      }
      break;
    }
    default: break;
  }
  mat_vec(w, o->mat, r);
  for (int k = 0; k < 3; k++) res[k] = w[k] + o->pos[k] + dir[k] * o->margin;
}
static void ccd_support(const CcdObj* o1, const CcdObj* o2, const double* dir, Supp* s) {
  double nd[3] = {-dir[0], -dir[1], -dir[2]};
  ccd_support_geom(o1, dir, s->v1);
  ccd_support_geom(o2, nd, s->v2);
  for (int k = 0; k < 3; k++) s->v[k] = s->v1[k] - s->v2[k];
}
static void ccd_normalize(double* v) { double k = 1.0 / norm3(v); v[0] *= k; v[1] *= k; v[2] *= k; }  /* ccdVec3Normalize: scale by 1/len, no guard */
static void portal_dir(const Supp* P, double* dir) {
  double a[3], b[3];
  for (int k = 0; k < 3; k++) { a[k] = P[2].v[k] - P[1].v[k]; b[k] = P[3].v[k] - P[1].v[k]; }
  cross3(dir, a, b);
  ccd_normalize(dir);
}
static int portal_reach_tolerance(const Supp* P, const Supp* v4, const double* dir) {
  double dv1 = dot3(P[1].v, dir), dv2 = dot3(P[2].v, dir), dv3 = dot3(P[3].v, dir), dv4 = dot3(v4->v, dir);
  double d = fmin(fmin(dv4 - dv1, dv4 - dv2), dv4 - dv3);
  return ccd_eq(d, MPR_TOLERANCE) || d < MPR_TOLERANCE;
}
static void expand_portal(Supp* P, const Supp* v4) {
  double v4v0[3];
  cross3(v4v0, v4->v, P[0].v);
  if (dot3(P[1].v, v4v0) > 0) {
    if (dot3(P[2].v, v4v0) > 0) P[1] = *v4; else P[3] = *v4;
  } else {
    if (dot3(P[3].v, v4v0) > 0) P[2] = *v4; else P[1] = *v4;
  }
}
static double point_seg_dist2(const double* x0, const double* b, double* wit) {  /* P = origin */
  double d[3] = {b[0] - x0[0], b[1] - x0[1], b[2] - x0[2]};
  double t = -dot3(x0, d) / dot3(d, d);
  if (t < 0 || ccd_is_zero(t)) { memcpy(wit, x0, 24); }
  else if (t > 1 || ccd_eq(t, 1)) { memcpy(wit, b, 24); }
  else for (int k = 0; k < 3; k++) wit[k] = x0[k] + t * d[k];
  return dot3(wit, wit);
}
static double origin_tri_dist2(const double* x0, const double* B, const double* C, double* wit) {  /* ccdVec3PointTriDist2 */
  double d1[3], d2[3];
  for (int k = 0; k < 3; k++) { d1[k] = B[k] - x0[k]; d2[k] = C[k] - x0[k]; }
  double v = dot3(d1, d1), w = dot3(d2, d2), p = dot3(x0, d1), q = dot3(x0, d2), r = dot3(d1, d2);
  double s = (q * r - w * p) / (w * v - r * r), t = (-s * r - q) / w;
  if ((ccd_is_zero(s) || s > 0) && (ccd_eq(s, 1) || s < 1) && (ccd_is_zero(t) || t > 0) && (ccd_eq(t, 1) || t < 1) &&
      (ccd_eq(t + s, 1) || t + s < 1)) {
    for (int k = 0; k < 3; k++) wit[k] = x0[k] + s * d1[k] + t * d2[k];
    return dot3(wit, wit);
  }
  double w2[3], dist = point_seg_dist2(x0, B, wit), d;
  d = point_seg_dist2(x0, C, w2); if (d < dist) { dist = d; memcpy(wit, w2, 24); }
  d = point_seg_dist2(B, C, w2);  if (d < dist) { dist = d; memcpy(wit, w2, 24); }
  return dist;
}
/* returns 0 and (depth, dir, pos) when the inflated objects intersect, -1 otherwise (ccdMPRPenetration) */
static int mpr_penetration(const CcdObj* o1, const CcdObj* o2, double* depth, double* pdir, double* pos) {
  Supp P[4], v4;
  double dir[3], va[3], vb[3], dot;
  /* discoverPortal */
  for (int k = 0; k < 3; k++) { P[0].v1[k] = o1->pos[k]; P[0].v2[k] = o2->pos[k]; P[0].v[k] = o1->pos[k] - o2->pos[k]; }
  if (ccd_eq(P[0].v[0], 0) && ccd_eq(P[0].v[1], 0) && ccd_eq(P[0].v[2], 0)) P[0].v[0] += CCD_EPS * 10;
  for (int k = 0; k < 3; k++) dir[k] = -P[0].v[k];
  ccd_normalize(dir);
  ccd_support(o1, o2, dir, &P[1]);
  dot = dot3(P[1].v, dir);
  if (ccd_is_zero(dot) || dot < 0) return -1;
  cross3(dir, P[0].v, P[1].v);
  if (ccd_is_zero(dot3(dir, dir))) {
    if (ccd_eq(P[1].v[0], 0) && ccd_eq(P[1].v[1], 0) && ccd_eq(P[1].v[2], 0)) {  /* findPenetrTouch: normal undefined */
      *depth = 0; pdir[0] = pdir[1] = pdir[2] = 0;
    } else {                                                                       /* findPenetrSegment */
      *depth = norm3(P[1].v);
      memcpy(pdir, P[1].v, 24); ccd_normalize(pdir);
    }
    for (int k = 0; k < 3; k++) pos[k] = 0.5 * (P[1].v1[k] + P[1].v2[k]);
    return 0;
  }
  ccd_normalize(dir);
  ccd_support(o1, o2, dir, &P[2]);
  dot = dot3(P[2].v, dir);
  if (ccd_is_zero(dot) || dot < 0) return -1;
  for (int k = 0; k < 3; k++) { va[k] = P[1].v[k] - P[0].v[k]; vb[k] = P[2].v[k] - P[0].v[k]; }
  cross3(dir, va, vb);
  ccd_normalize(dir);
  if (dot3(dir, P[0].v) > 0) { Supp t = P[1]; P[1] = P[2]; P[2] = t; dir[0] = -dir[0]; dir[1] = -dir[1]; dir[2] = -dir[2]; }
  for (int guard = 0;; guard++) {
    if (guard > MPR_PORTAL_CAP) return -1;
    ccd_support(o1, o2, dir, &P[3]);
    dot = dot3(P[3].v, dir);
    if (ccd_is_zero(dot) || dot < 0) return -1;
    int cont = 0;
    cross3(va, P[1].v, P[3].v);
    dot = dot3(va, P[0].v);
    if (dot < 0 && !ccd_is_zero(dot)) { P[2] = P[3]; cont = 1; }
    if (!cont) {
      cross3(va, P[3].v, P[2].v);
      dot = dot3(va, P[0].v);
      if (dot < 0 && !ccd_is_zero(dot)) { P[1] = P[3]; cont = 1; }
    }
    if (!cont) break;
    for (int k = 0; k < 3; k++) { va[k] = P[1].v[k] - P[0].v[k]; vb[k] = P[2].v[k] - P[0].v[k]; }
    cross3(dir, va, vb);
    ccd_normalize(dir);
  }
  /* refinePortal */
  for (int guard = 0;; guard++) {
    if (guard > MPR_PORTAL_CAP) return -1;
    portal_dir(P, dir);
    dot = dot3(dir, P[1].v);
    if (ccd_is_zero(dot) || dot > 0) break;  /* portal encapsulates the origin */
    ccd_support(o1, o2, dir, &v4);
    dot = dot3(v4.v, dir);
    if (!(ccd_is_zero(dot) || dot > 0) || portal_reach_tolerance(P, &v4, dir)) return -1;
    expand_portal(P, &v4);
  }
  /* findPenetr */
  for (unsigned long it = 0;; it++) {
    portal_dir(P, dir);
    ccd_support(o1, o2, dir, &v4);
    if (portal_reach_tolerance(P, &v4, dir) || it > MPR_ITERATIONS) {
      *depth = sqrt(origin_tri_dist2(P[1].v, P[2].v, P[3].v, pdir));
      if (ccd_is_zero(pdir[0]) && ccd_is_zero(pdir[1]) && ccd_is_zero(pdir[2])) memcpy(pdir, dir, 24);  /* origin on the portal */
      ccd_normalize(pdir);
      /* findPos: barycentric coordinates of the origin in the portal tetrahedron */
      double b[4], vec[3], sum;
      portal_dir(P, dir);
      cross3(vec, P[1].v, P[2].v); b[0] = dot3(vec, P[3].v);
      cross3(vec, P[3].v, P[2].v); b[1] = dot3(vec, P[0].v);
      cross3(vec, P[0].v, P[1].v); b[2] = dot3(vec, P[3].v);
      cross3(vec, P[2].v, P[1].v); b[3] = dot3(vec, P[0].v);
      sum = b[0] + b[1] + b[2] + b[3];
      if (ccd_is_zero(sum) || sum < 0) {
        b[0] = 0;
        cross3(vec, P[2].v, P[3].v); b[1] = dot3(vec, dir);
        cross3(vec, P[3].v, P[1].v); b[2] = dot3(vec, dir);
        cross3(vec, P[1].v, P[2].v); b[3] = dot3(vec, dir);
        sum = b[1] + b[2] + b[3];
      }
      for (int k = 0; k < 3; k++) {
        double p1 = 0, p2 = 0;
        for (int i = 0; i < 4; i++) { p1 += b[i] * P[i].v1[k]; p2 += b[i] * P[i].v2[k]; }
        pos[k] = 0.5 * (p1 + p2) / sum;
      }
      return 0;
    }
    expand_portal(P, &v4);
  }
}
/* mjc_fixNormal: normal of a smooth geom at the contact point, from its shape; n is the MPR direction on entry */
static void convex_geom_normal(const CcdObj* o, const double* cpos, double* n) {
  double dif[3] = {cpos[0] - o->pos[0], cpos[1] - o->pos[1], cpos[2] - o->pos[2]}, p[3], nl[3];
  matT_vec(p, o->mat, dif);
  matT_vec(nl, o->mat, n);
  switch (o->type) {
    case FLYB_GEOM_SPHERE: memcpy(nl, p, 24); break;
    case FLYB_GEOM_CAPSULE:
      nl[0] = p[0]; nl[1] = p[1];
      nl[2] = p[2] < -o->size[1] ? p[2] + o->size[1] : (p[2] > o->size[1] ? p[2] - o->size[1] : 0.0);
      break;
    case FLYB_GEOM_ELLIPSOID: for (int k = 0; k < 3; k++) nl[k] = p[k] / (o->size[k] * o->size[k]); break;
    case FLYB_GEOM_CYLINDER: {
      double dside = fabs(sqrt(p[0] * p[0] + p[1] * p[1]) - o->size[0]), dcap = fabs(fabs(p[2]) - o->size[1]);
      if (dside < 0.5 * dcap) { nl[0] = p[0]; nl[1] = p[1]; nl[2] = 0; }
      else if (dcap < 0.5 * dside) { nl[0] = 0; nl[1] = 0; nl[2] = p[2]; }
      break;
    }
    default: break;
  }
  normalize3(nl);
  mat_vec(n, o->mat, nl);
}
static int hit_convex(Hit* h, int t1, const double* c1, const double* m1, const double* s1, int t2, const double* c2,
                      const double* m2, const double* s2, double margin) {
  CcdObj o1 = {t1, c1, m1, s1, 0.5 * margin}, o2 = {t2, c2, m2, s2, 0.5 * margin};
  double depth, dir[3], pos[3];
  if (mpr_penetration(&o1, &o2, &depth, dir, pos) != 0) return 0;
  if (dir[0] == 0 && dir[1] == 0 && dir[2] == 0) return 0;  /* touching contact without a normal */
  double n1[3] = {dir[0], dir[1], dir[2]}, n2[3] = {-dir[0], -dir[1], -dir[2]};
  convex_geom_normal(&o1, pos, n1);
  convex_geom_normal(&o2, pos, n2);
  for (int k = 0; k < 3; k++) h->n[k] = n1[k] - n2[k];
  normalize3(h->n);
  h->dist = margin - depth; h->has_t = 0;
  memcpy(h->pos, pos, 24);
  return 1;
}

/* world-frame translational Jacobian row of a point fixed to body b, projected on direction dir:
   out[d] += sign * dir . (cdof_lin[d] + cdof_ang[d] x (point - subtree_com[root])) for every dof d above b */
static void add_jac_dir(FlyoEnv* e, double* out, int b, const double* point, const double* dir, double sign) {
  const FlybModel* m = e->m;
  if (b <= 0) return;
  const double* c0 = e->subtree_com + 3 * m->body_rootid[b];
  double off[3] = {point[0] - c0[0], point[1] - c0[1], point[2] - c0[2]};
  for (int d = m->body_dofadr[b] + m->body_dofnum[b] - 1; d >= 0; d = m->dof_parentid[d]) {
    const double* cd = e->cdof + 6 * d;
    double t[3];
    cross3(t, cd, off);
    out[d] += sign * (dir[0] * (cd[3] + t[0]) + dir[1] * (cd[4] + t[1]) + dir[2] * (cd[5] + t[2]));
  }
}

/* narrow phase of one geom pair (t1 <= t2 in MuJoCo's type order); the collision driver and the unit tests share it */
static int narrow_pair(Hit* h, int t1, const double* c1, const double* m1, const double* s1, int t2, const double* c2,
                       const double* m2, const double* s2, double margin) {
  if (t1 == FLYB_GEOM_PLANE) {
    double pn[3] = {m1[2], m1[5], m1[8]};
    if (t2 == FLYB_GEOM_SPHERE) return hit_plane_sphere(h, c1, pn, c2, s2[0], margin);
    if (t2 == FLYB_GEOM_CAPSULE) return hit_plane_capsule(h, c1, pn, c2, m2, s2, margin);
    if (t2 == FLYB_GEOM_ELLIPSOID) return hit_plane_ellipsoid(h, c1, pn, c2, m2, s2, margin);
    if (t2 == FLYB_GEOM_CYLINDER) return hit_plane_cylinder(h, c1, pn, c2, m2, s2, margin);
    return 0;
  }
  if (t1 == FLYB_GEOM_SPHERE) {
    if (t2 == FLYB_GEOM_SPHERE) return hit_sphere_sphere(h, c1, s1[0], c2, s2[0], margin);
    if (t2 == FLYB_GEOM_CAPSULE) return hit_sphere_capsule(h, c1, s1[0], c2, m2, s2, margin);
    if (t2 == FLYB_GEOM_CYLINDER) return hit_sphere_cylinder(h, c1, s1[0], c2, m2, s2, margin);
  }
  if (t1 == FLYB_GEOM_CAPSULE && t2 == FLYB_GEOM_CAPSULE) return hit_capsule_capsule(h, c1, m1, s1, c2, m2, s2, margin);
  return hit_convex(h, t1, c1, m1, s1, t2, c2, m2, s2, margin);  /* every pair type without an analytic routine */
}
/* test hook: out = [n, then per hit: dist, pos(3), normal(3)]; force_convex routes the pair through the MPR path */
int flyo_narrow_pair(int t1, const double* c1, const double* m1, const double* s1, int t2, const double* c2,
                     const double* m2, const double* s2, double margin, int force_convex, double* out) {
  Hit h[4];
  int n = force_convex ? hit_convex(h, t1, c1, m1, s1, t2, c2, m2, s2, margin)
                       : narrow_pair(h, t1, c1, m1, s1, t2, c2, m2, s2, margin);
  for (int i = 0; i < n; i++) {
    out[7 * i] = h[i].dist;
    for (int k = 0; k < 3; k++) { out[7 * i + 1 + k] = h[i].pos[k]; out[7 * i + 4 + k] = h[i].n[k]; }
  }
  return n;
}

/* ---- height field against a convex geom (MuJoCo mjc_ConvexHField, engine_collision_convex.c; third party, restated from the
 * published description and memory of its structure - "parity unpinned" like the rest of the convex code).
 * The geom is expressed in the height field's frame; the grid cells under its axis-aligned bounds (taken with its support
 * function) are walked row by row as a zig-zag strip of vertices (column c of row r, column c of row r + 1, column c + 1 of row
 * r, ...): every three consecutive strip vertices are the top triangle of a prism whose bottom lies at -base, its top raised by
 * the margin.  Prisms whose top stays below the geom's lowest point are skipped; the others go through the same portal
 * refinement as every convex pair (prism as object 1, its centroid as centre).  A penetration of depth d gives a contact with
 * dist = margin - d, normal = the refinement's direction (prism -> geom), position = its contact point lowered by margin / 2.
 * hsize: x half extent, y half extent, elevation scale, base thickness; data[r * ncol + c] x hsize[2] is the height at
 * x = -hsize[0] + c dx, y = -hsize[1] + r dy.  Returns the number of contacts written (at most maxcon): dist, pos(3), normal(3). */
int flyo_narrow_hfield(const double* hpos, const double* hmat, const double* hsize, int nrow, int ncol, const float* data,
                       int t2, const double* c2, const double* m2, const double* s2, double margin, int maxcon, double* out) {
  double pos[3], mat[9], d[3] = {c2[0] - hpos[0], c2[1] - hpos[1], c2[2] - hpos[2]};
  matT_vec(pos, hmat, d);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) mat[3 * i + j] = hmat[i] * m2[j] + hmat[3 + i] * m2[3 + j] + hmat[6 + i] * m2[6 + j];
  const CcdObj geom = {t2, pos, mat, s2, 0.0};
  /* bounds of the geom in the height field's frame */
  double lo[3], hi[3];
  for (int a = 0; a < 3; a++) {
    double dir[3] = {0, 0, 0}, p[3];
    dir[a] = 1; ccd_support_geom(&geom, dir, p); hi[a] = p[a];
    dir[a] = -1; ccd_support_geom(&geom, dir, p); lo[a] = p[a];
  }
  const double sx = hsize[0], sy = hsize[1];
  if (hi[0] < -sx || lo[0] > sx || hi[1] < -sy || lo[1] > sy || hi[2] < -hsize[3]) return 0;
  const double dx = 2 * sx / (ncol - 1), dy = 2 * sy / (nrow - 1);
  int cmin = (int)floor((lo[0] + sx) / dx), cmax = (int)ceil((hi[0] + sx) / dx);
  int rmin = (int)floor((lo[1] + sy) / dy), rmax = (int)ceil((hi[1] + sy) / dy);
  if (cmin < 0) cmin = 0; if (rmin < 0) rmin = 0;
  if (cmax > ncol - 1) cmax = ncol - 1; if (rmax > nrow - 1) rmax = nrow - 1;
  static const double ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  int ncon = 0;
  for (int r = rmin; r < rmax; r++) {
    double top[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};   /* sliding window over the strip */
    int nvert = 0;
    for (int c = cmin; c <= cmax; c++)
      for (int i = 0; i < 2; i++) {
        memcpy(top[0], top[1], 24); memcpy(top[1], top[2], 24);
        top[2][0] = dx * c - sx; top[2][1] = dy * (r + i) - sy;
        top[2][2] = (double)data[(size_t)(r + i) * ncol + c] * hsize[2] + margin;
        if (++nvert < 3) continue;
        if (top[0][2] < lo[2] && top[1][2] < lo[2] && top[2][2] < lo[2]) continue;
        double cen[3] = {0, 0, 0}, verts[18];
        for (int v = 0; v < 3; v++) { cen[0] += top[v][0] / 3; cen[1] += top[v][1] / 3; cen[2] += 0.5 * (top[v][2] - hsize[3]) / 3; }
        for (int v = 0; v < 3; v++) {
          for (int k = 0; k < 2; k++) { verts[3 * v + k] = top[v][k] - cen[k]; verts[9 + 3 * v + k] = top[v][k] - cen[k]; }
          verts[3 * v + 2] = top[v][2] - cen[2]; verts[9 + 3 * v + 2] = -hsize[3] - cen[2];
        }
        const CcdObj prism = {GEOM_PRISM, cen, ident, verts, 0.0};
        double depth, dir[3], cp[3];
        if (mpr_penetration(&prism, &geom, &depth, dir, cp) != 0) continue;
        if ((dir[0] == 0 && dir[1] == 0 && dir[2] == 0) || ncon >= maxcon) continue;
        double lp[3] = {cp[0] - 0.5 * margin * dir[0], cp[1] - 0.5 * margin * dir[1], cp[2] - 0.5 * margin * dir[2]}, wp[3], wn[3];
        mat_vec(wp, hmat, lp); mat_vec(wn, hmat, dir);
        out[7 * ncon] = margin - depth;
        for (int k = 0; k < 3; k++) { out[7 * ncon + 1 + k] = wp[k] + hpos[k]; out[7 * ncon + 4 + k] = wn[k]; }
        ncon++;
      }
  }
  return ncon;
}

static void collision(FlyoEnv* e) {
  const FlybModel* m = e->m;
  e->ncon = 0;
  if (e->disable_contact) return;
  for (int p = 0; p < m->npair; p++) {
    int g1 = m->pair_geom1[p], g2 = m->pair_geom2[p];
    int t1 = m->geom_type[g1], t2 = m->geom_type[g2];
    const double *c1 = e->geom_xpos + 3 * g1, *c2 = e->geom_xpos + 3 * g2;
    const double *m1 = e->geom_xmat + 9 * g1, *m2 = e->geom_xmat + 9 * g2;
    const double *s1 = m->geom_size + 3 * g1, *s2 = m->geom_size + 3 * g2;
    double margin = m->pair_margin[p];
    /* bounding-sphere rejection (cannot change the result: narrow phase only reports dist <= margin) */
    if (t1 == FLYB_GEOM_PLANE) {
      double n[3] = {m1[2], m1[5], m1[8]}, d[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
      if (dot3(d, n) > m->geom_rbound[g2] + margin) continue;
    } else if (t1 == FLYB_GEOM_HFIELD) {   /* the routine's own bounds test decides */
    } else {
      double d[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]}, bound = m->geom_rbound[g1] + m->geom_rbound[g2] + margin;
      if (dot3(d, d) > bound * bound) continue;
    }
    Hit h[FLYO_MAXHFCON];
    int n;
    if (t1 == FLYB_GEOM_HFIELD) {
      /* mjc_ConvexHField keeps up to mjMAXCONPAIR = 50 contacts of one pair; oracle and engine keep FLYO_MAXHFCON = 16
         (a fly geom of at most 0.1 cm over a 0.1 cm grid touches a handful of prisms; the vision task ends the episode
         on the first active floor contact, tasks/vision_flight.py:214-232) */
      double out[7 * FLYO_MAXHFCON];
      n = e->hf_data ? flyo_narrow_hfield(c1, m1, m->hfield_size, m->hfield_nrow, m->hfield_ncol, e->hf_data, t2, c2, m2, s2,
                                          margin, FLYO_MAXHFCON, out) : 0;
      for (int i = 0; i < n; i++) {
        h[i].dist = out[7 * i]; h[i].has_t = 0;
        memcpy(h[i].pos, out + 7 * i + 1, 24); memcpy(h[i].n, out + 7 * i + 4, 24);
      }
    } else n = narrow_pair(h, t1, c1, m1, s1, t2, c2, m2, s2, margin);
    for (int i = 0; i < n && e->ncon < FLYO_MAXCON; i++) {
      FlyoContact* c = e->con + e->ncon++;
      c->pair = p; c->geom1 = g1; c->geom2 = g2; c->dim = m->pair_condim[p];
      c->dist = h[i].dist; c->includemargin = margin - m->pair_gap[p];
      memcpy(c->pos, h[i].pos, sizeof(c->pos));
      make_frame(c->frame, h[i].n, h[i].t, h[i].has_t);
      memcpy(c->friction, m->pair_friction + 3 * p, sizeof(c->friction));
      memcpy(c->solref, m->pair_solref + 2 * p, sizeof(c->solref));
      memcpy(c->solimp, m->pair_solimp + 5 * p, sizeof(c->solimp));
      c->exclude = (c->dist >= c->includemargin) ? 1 : 0;
      c->efc_address = -1; c->mu = c->friction[0];
    }
  }
}

/* ---------------------------------------------------------------- m6 constraint rows */
enum { CT_LIMIT = 0, CT_CONTACT1 = 1, CT_ELLIPTIC = 2 };
enum { ST_SATISFIED = 0, ST_QUADRATIC = 1, ST_CONE = 2 };

static void impedance(const double* solimp, double pos, double margin, double* imp) {
  double d0 = solimp[0], dw = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  d0 = fmin(0.9999, fmax(0.0001, d0)); dw = fmin(0.9999, fmax(0.0001, dw));
  if (power < 1) power = 1;
  mid = fmin(0.9999, fmax(0.0001, mid));
  if (d0 == dw || width <= MINVAL) { *imp = 0.5 * (d0 + dw); return; }
  double x = fabs((pos - margin) / width), y;
  if (x >= 1) { *imp = dw; return; }
  if (x == 0) { *imp = d0; return; }
  if (power == 1) y = x;
  else if (x <= mid) y = pow(x, power) / pow(mid, power - 1);
  else y = 1 - pow(1 - x, power) / pow(1 - mid, power - 1);
  *imp = d0 + y * (dw - d0);
}

static void finish_row(FlyoEnv* e, int r, const double* solref_in, const double* solimp, int friction_row) {
  const FlybModel* m = e->m;
  double solref[2] = {solref_in[0], solref_in[1]};
  if (solref[0] > 0 && solref[0] < 2 * m->opt_timestep) solref[0] = 2 * m->opt_timestep;  /* refsafe */
  double imp;
  impedance(solimp, e->efc_pos[r], e->efc_margin[r], &imp);
  double dmax = fmin(0.9999, fmax(0.0001, solimp[1]));
  double K, B;
  if (solref[0] > 0) {
    K = 1.0 / fmax(MINVAL, dmax * dmax * solref[0] * solref[0] * solref[1] * solref[1]);
    B = 2.0 / fmax(MINVAL, dmax * solref[0]);
  } else { K = -solref[0] / fmax(MINVAL, dmax * dmax); B = -solref[1] / fmax(MINVAL, dmax); }
  if (friction_row) K = 0;
  e->efc_KBIP[4 * r] = K; e->efc_KBIP[4 * r + 1] = B; e->efc_KBIP[4 * r + 2] = imp; e->efc_KBIP[4 * r + 3] = 0;
  e->efc_R[r] = fmax(MINVAL, (1 - imp) * e->efc_diagApprox[r] / imp);
}

static void make_constraint(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nv = m->nv, r = 0;
  /* joint limits (hinges), lower side then upper side */
  if (!e->disable_limit)
    for (int j = 0; j < m->njnt; j++) {
      if (!m->jnt_limited[j]) continue;
      double q = e->qpos[m->jnt_qposadr[j]];
      for (int side = -1; side <= 1; side += 2) {
        double dist = side * (m->jnt_range[2 * j + (side + 1) / 2] - q);
        if (dist < 0 && r < FLYO_MAXEFC) {  /* jnt_margin = 0 */
          memset(e->efc_J + (size_t)r * nv, 0, nv * sizeof(double));
          e->efc_J[(size_t)r * nv + m->jnt_dofadr[j]] = -side;
          e->efc_pos[r] = dist; e->efc_margin[r] = 0;
          e->efc_type[r] = CT_LIMIT; e->efc_id[r] = j;
          e->efc_diagApprox[r] = m->dof_invweight0[m->jnt_dofadr[j]];
          finish_row(e, r, m->jnt_solref + 2 * j, m->jnt_solimp + 5 * j, 0);
          r++;
        }
      }
    }
  e->nlimit = r;
  for (int i = 0; i < e->ncon; i++) {
    FlyoContact* c = e->con + i;
    c->efc_address = -1;
    if (c->exclude) continue;
    int dim = c->dim;
    if (r + dim > FLYO_MAXEFC) break;
    int b1 = m->geom_bodyid[c->geom1], b2 = m->geom_bodyid[c->geom2];
    c->efc_address = r;
    double tran = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
    for (int k = 0; k < dim; k++) {
      double* J = e->efc_J + (size_t)(r + k) * nv;
      memset(J, 0, nv * sizeof(double));
      add_jac_dir(e, J, b2, c->pos, c->frame + 3 * k, 1.0);
      add_jac_dir(e, J, b1, c->pos, c->frame + 3 * k, -1.0);
      e->efc_pos[r + k] = k == 0 ? c->dist : 0;
      e->efc_margin[r + k] = k == 0 ? c->includemargin : 0;
      e->efc_type[r + k] = dim == 1 ? CT_CONTACT1 : CT_ELLIPTIC;
      e->efc_id[r + k] = i;
      e->efc_diagApprox[r + k] = tran;
    }
    finish_row(e, r, c->solref, c->solimp, 0);
    if (dim > 1) {
      /* elliptic cone: friction rows share B and impedance; R1 = R0/impratio, Rj = R1 mu1^2/muj^2;
         regularised cone coefficient mu = mu1 sqrt(R1/R0) */
      e->efc_R[r + 1] = e->efc_R[r] / fmax(MINVAL, m->opt_impratio);
      for (int k = 2; k < dim; k++) {
        double f1 = c->friction[0], fk = c->friction[0];  /* condim 3: both tangential dims use friction[0] */
        e->efc_R[r + k] = e->efc_R[r + 1] * f1 * f1 / (fk * fk);
      }
      c->mu = c->friction[0] * sqrt(e->efc_R[r + 1] / e->efc_R[r]);
      for (int k = 1; k < dim; k++) {
        memcpy(e->efc_KBIP + 4 * (r + k), e->efc_KBIP + 4 * r, 4 * sizeof(double));
        e->efc_KBIP[4 * (r + k)] = 0;
      }
    }
    r += dim;
  }
  e->nefc = r;
  for (int i = 0; i < r; i++) e->efc_D[i] = 1.0 / e->efc_R[i];
}

/* ---------------------------------------------------------------- m3 tendon + transmission */
static void transmission(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nv = m->nv;
  for (int t = 0; t < m->ntendon; t++) {
    double L = 0;
    for (int w = m->tendon_adr[t]; w < m->tendon_adr[t] + m->tendon_num[t]; w++) L += m->wrap_coef[w] * e->qpos[m->wrap_qposadr[w]];
    e->ten_length[t] = L;
  }
  memset(e->act_moment, 0, (size_t)m->nu * nv * sizeof(double));
  for (int a = 0; a < m->nu; a++) {
    double* mom = e->act_moment + (size_t)a * nv;
    int id = m->act_trnid[a];
    double gear = m->act_gear[a];
    if (m->act_trntype[a] == FLYB_TRN_JOINT) {
      e->act_length[a] = gear * e->qpos[m->jnt_qposadr[id]];
      mom[m->jnt_dofadr[id]] = gear;
    } else if (m->act_trntype[a] == FLYB_TRN_TENDON) {
      e->act_length[a] = gear * e->ten_length[id];
      for (int w = m->tendon_adr[id]; w < m->tendon_adr[id] + m->tendon_num[id]; w++) mom[m->wrap_dofadr[w]] += gear * m->wrap_coef[w];
    } else {
      /* adhesion: moment = -(mean of the contact-normal Jacobians over the body's contacts, gap band included) */
      e->act_length[a] = 0;
      int cnt = 0;
      for (int i = 0; i < e->ncon; i++) {
        FlyoContact* c = e->con + i;
        int b1 = m->geom_bodyid[c->geom1], b2 = m->geom_bodyid[c->geom2];
        if (b1 != id && b2 != id) continue;
        cnt++;
        add_jac_dir(e, mom, b2, c->pos, c->frame, 1.0);
        add_jac_dir(e, mom, b1, c->pos, c->frame, -1.0);
      }
      if (cnt) for (int d = 0; d < nv; d++) mom[d] *= -gear / cnt;
    }
  }
}

/* ---------------------------------------------------------------- m7/m8 velocity stage */
static void com_vel(FlyoEnv* e) {
  const FlybModel* m = e->m;
  memset(e->cvel, 0, 6 * sizeof(double));
  for (int b = 1; b < m->nbody; b++) {
    double v[6];
    memcpy(v, e->cvel + 6 * m->body_parent[b], sizeof(v));
    for (int j = m->body_jntadr[b]; j < m->body_jntadr[b] + m->body_jntnum[b]; j++) {
      int d = m->jnt_dofadr[j], type = m->jnt_type[j], n = 1;
      if (type == FLYB_JNT_FREE) {
        memset(e->cdof_dot + 6 * d, 0, 18 * sizeof(double));
        for (int k = 0; k < 3; k++) for (int c = 0; c < 6; c++) v[c] += e->cdof[6 * (d + k) + c] * e->qvel[d + k];
        d += 3; n = 3;
      } else if (type == FLYB_JNT_BALL) n = 3;
      for (int k = 0; k < n; k++) cross_motion(e->cdof_dot + 6 * (d + k), v, e->cdof + 6 * (d + k));
      for (int k = 0; k < n; k++) for (int c = 0; c < 6; c++) v[c] += e->cdof[6 * (d + k) + c] * e->qvel[d + k];
    }
    memcpy(e->cvel + 6 * b, v, sizeof(v));
  }
}

/* generalized force of a world-frame force/torque applied at `point` on body b */
static void apply_ft(FlyoEnv* e, double* qfrc, int b, const double* point, const double* force, const double* torque) {
  const FlybModel* m = e->m;
  const double* c0 = e->subtree_com + 3 * m->body_rootid[b];
  double off[3] = {point[0] - c0[0], point[1] - c0[1], point[2] - c0[2]};
  for (int d = m->body_dofadr[b] + m->body_dofnum[b] - 1; d >= 0; d = m->dof_parentid[d]) {
    const double* cd = e->cdof + 6 * d;
    double t[3];
    cross3(t, cd, off);
    qfrc[d] += force[0] * (cd[3] + t[0]) + force[1] * (cd[4] + t[1]) + force[2] * (cd[5] + t[2]) + dot3(torque, cd);
  }
}
/* 6D velocity [ang; lin] of a frame (pos, mat) rigidly attached to body b, expressed in that frame */
static void local_velocity(FlyoEnv* e, double* lv, int b, const double* pos, const double* mat) {
  const FlybModel* m = e->m;
  const double* c0 = e->subtree_com + 3 * m->body_rootid[b];
  const double* cv = e->cvel + 6 * b;
  double off[3] = {pos[0] - c0[0], pos[1] - c0[1], pos[2] - c0[2]}, t[3], lin[3];
  cross3(t, cv, off);
  for (int k = 0; k < 3; k++) lin[k] = cv[3 + k] + t[k];
  matT_vec(lv, mat, cv);
  matT_vec(lv + 3, mat, lin);
}

static double max3(const double* s) { return fmax(s[0], fmax(s[1], s[2])); }
static double min3(const double* s) { return fmin(s[0], fmin(s[1], s[2])); }

static void passive(FlyoEnv* e) {
  const FlybModel* m = e->m;
  memset(e->qfrc_passive, 0, m->nv * sizeof(double));
  for (int j = 0; j < m->njnt; j++) {
    if (m->jnt_type[j] != FLYB_JNT_HINGE) continue;
    int qa = m->jnt_qposadr[j], d = m->jnt_dofadr[j];
    e->qfrc_passive[d] -= m->jnt_stiffness[j] * (e->qpos[qa] - m->qpos_spring[qa]);
  }
  for (int d = 0; d < m->nv; d++) e->qfrc_passive[d] -= m->dof_damping[d] * e->qvel[d];
  if (e->disable_fluid) return;
  double rho = m->opt_density, beta = m->opt_viscosity;
  /* inertia-box model, one element per (unfused) massive body */
  for (int i = 0; i < m->nfluidbox; i++) {
    int b = m->fbox_bodyid[i];
    double pos[3], q[4], mat[9], t[3], lv[6], lf[6];
    mat_vec(t, e->xmat + 9 * b, m->fbox_pos + 3 * i);
    for (int k = 0; k < 3; k++) pos[k] = e->xpos[3 * b + k] + t[k];
    quat_mul(q, e->xquat + 4 * b, m->fbox_quat + 4 * i);
    quat2mat(mat, q);
    local_velocity(e, lv, b, pos, mat);
    const double* bx = m->fbox_size + 3 * i;
    memset(lf, 0, sizeof(lf));
    if (beta > 0) {
      double diam = (bx[0] + bx[1] + bx[2]) / 3.0;
      for (int k = 0; k < 3; k++) { lf[k] = -PI * diam * diam * diam * beta * lv[k]; lf[3 + k] = -3.0 * PI * diam * beta * lv[3 + k]; }
    }
    if (rho > 0) {
      lf[3] -= 0.5 * rho * bx[1] * bx[2] * fabs(lv[3]) * lv[3];
      lf[4] -= 0.5 * rho * bx[0] * bx[2] * fabs(lv[4]) * lv[4];
      lf[5] -= 0.5 * rho * bx[0] * bx[1] * fabs(lv[5]) * lv[5];
      lf[0] -= rho * bx[0] * (pow(bx[1], 4) + pow(bx[2], 4)) * fabs(lv[0]) * lv[0] / 64.0;
      lf[1] -= rho * bx[1] * (pow(bx[0], 4) + pow(bx[2], 4)) * fabs(lv[1]) * lv[1] / 64.0;
      lf[2] -= rho * bx[2] * (pow(bx[0], 4) + pow(bx[1], 4)) * fabs(lv[2]) * lv[2] / 64.0;
    }
    double f[3], tq[3];
    mat_vec(tq, mat, lf); mat_vec(f, mat, lf + 3);
    apply_ft(e, e->qfrc_passive, b, pos, f, tq);
  }
  /* ellipsoid model (wing_*_fluid geoms, tasks/base.py:309-313) */
  for (int i = 0; i < m->nfluidgeom; i++) {
    int b = m->fgeom_bodyid[i];
    double pos[3], q[4], mat[9], t[3], lv[6], lf[6];
    mat_vec(t, e->xmat + 9 * b, m->fgeom_pos + 3 * i);
    for (int k = 0; k < 3; k++) pos[k] = e->xpos[3 * b + k] + t[k];
    quat_mul(q, e->xquat + 4 * b, m->fgeom_quat + 4 * i);
    quat2mat(mat, q);
    local_velocity(e, lv, b, pos, mat);
    const double* s = m->fgeom_size + 3 * i; const double* c = m->fgeom_coef + 12 * i;
    double blunt = c[1], slender = c[2], angc = c[3], kutta = c[4], magnus = c[5];
    const double *vm = c + 6, *vi = c + 9, *w = lv, *v = lv + 3;
    memset(lf, 0, sizeof(lf));
    /* added mass (Kirchhoff): force = p x w, torque = p x v + L x w */
    double p[3] = {rho * vm[0] * v[0], rho * vm[1] * v[1], rho * vm[2] * v[2]};
    double L[3] = {rho * vi[0] * w[0], rho * vi[1] * w[1], rho * vi[2] * w[2]}, a1[3], a2[3], a3[3];
    cross3(a1, p, w); cross3(a2, p, v); cross3(a3, L, w);
    for (int k = 0; k < 3; k++) { lf[3 + k] += a1[k]; lf[k] += a2[k] + a3[k]; }
    /* viscous / lift terms */
    double vol = 4.0 / 3.0 * PI * s[0] * s[1] * s[2];
    double dmax = max3(s), dmin = min3(s), dmid = s[0] + s[1] + s[2] - dmax - dmin;
    double A_max = PI * dmax * dmid;
    double mg[3]; cross3(mg, w, v);
    for (int k = 0; k < 3; k++) lf[3 + k] += magnus * rho * vol * mg[k];
    double bc = s[1] * s[2], ca = s[2] * s[0], ab = s[0] * s[1];
    double pd = bc * bc * bc * bc * v[0] * v[0] + ca * ca * ca * ca * v[1] * v[1] + ab * ab * ab * ab * v[2] * v[2];
    double pn = bc * bc * v[0] * v[0] + ca * ca * v[1] * v[1] + ab * ab * v[2] * v[2];
    double A_proj = PI * sqrt(pd / fmax(MINVAL, pn));
    double speed = norm3(v);
    double nrm[3] = {bc * bc * v[0], ca * ca * v[1], ab * ab * v[2]};
    double cos_alpha = pn / fmax(MINVAL, speed * pd);
    double circ[3], kf[3];
    cross3(circ, nrm, v);
    for (int k = 0; k < 3; k++) circ[k] *= kutta * rho * cos_alpha * A_proj;
    cross3(kf, circ, v);
    for (int k = 0; k < 3; k++) lf[3 + k] += kf[k];
    double eqD = 2.0 / 3.0 * (s[0] + s[1] + s[2]);
    double I_max = 8.0 / 15.0 * PI * dmid * pow(dmax, 4);
    double II[3], mv[3];
    for (int k = 0; k < 3; k++) {
      double o = fmax(s[(k + 1) % 3], s[(k + 2) % 3]);
      II[k] = 8.0 / 15.0 * PI * s[k] * pow(o, 4);
      mv[k] = w[k] * (angc * II[k] + slender * (I_max - II[k]));
    }
    double drag_lin = beta * 3.0 * PI * eqD + rho * speed * (A_proj * blunt + slender * (A_max - A_proj));
    double drag_ang = beta * PI * eqD * eqD * eqD + rho * norm3(mv);
    for (int k = 0; k < 3; k++) { lf[3 + k] -= drag_lin * v[k]; lf[k] -= drag_ang * w[k]; }
    for (int k = 0; k < 6; k++) lf[k] *= c[0];
    double f[3], tq[3];
    mat_vec(tq, mat, lf); mat_vec(f, mat, lf + 3);
    apply_ft(e, e->qfrc_passive, b, pos, f, tq);
  }
}

/* recursive Newton-Euler: flg_acc=0 -> bias forces; also fills cacc/cfrc_int when post=1 (m12) */
static void rne(FlyoEnv* e, double* result, int flg_acc, int post) {
  const FlybModel* m = e->m;
  int nb = m->nbody;
  double* cacc = e->cacc; double* cf = e->cfrc_int;
  memset(cacc, 0, 6 * sizeof(double));
  for (int k = 0; k < 3; k++) cacc[3 + k] = -m->opt_gravity[k];
  memset(cf, 0, 6 * sizeof(double));
  for (int b = 1; b < nb; b++) {
    double* a = cacc + 6 * b;
    memcpy(a, cacc + 6 * m->body_parent[b], 6 * sizeof(double));
    for (int d = m->body_dofadr[b]; d < m->body_dofadr[b] + m->body_dofnum[b]; d++)
      for (int c = 0; c < 6; c++) {
        a[c] += e->cdof_dot[6 * d + c] * e->qvel[d];
        if (flg_acc) a[c] += e->cdof[6 * d + c] * e->qacc[d];
      }
    double Ia[6], Iv[6], x[6];
    inert_vec(Ia, e->cinert + 10 * b, a);
    inert_vec(Iv, e->cinert + 10 * b, e->cvel + 6 * b);
    cross_force(x, e->cvel + 6 * b, Iv);
    for (int c = 0; c < 6; c++) cf[6 * b + c] = Ia[c] + x[c] - (post ? e->cfrc_ext[6 * b + c] : 0);
  }
  for (int b = nb - 1; b > 0; b--)
    for (int c = 0; c < 6; c++) cf[6 * m->body_parent[b] + c] += cf[6 * b + c];
  if (result)
    for (int d = 0; d < m->nv; d++) {
      double s = 0;
      for (int c = 0; c < 6; c++) s += e->cdof[6 * d + c] * cf[6 * m->dof_bodyid[d] + c];
      result[d] = s;
    }
}

/* ---------------------------------------------------------------- m9/m10 actuation + smooth acceleration */
static void fwd_actuation(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nv = m->nv;
  memset(e->qfrc_actuator, 0, nv * sizeof(double));
  for (int a = 0; a < m->nu; a++) {
    double ctrl = e->ctrl[a];
    if (m->act_ctrllimited[a]) ctrl = fmin(m->act_ctrlrange[2 * a + 1], fmax(m->act_ctrlrange[2 * a], ctrl));
    double input = ctrl;
    int aa = m->act_actadr[a];
    if (aa >= 0) {
      e->act_dot[aa] = (ctrl - e->act[aa]) / fmax(MINVAL, m->act_dynprm[a]);
      input = e->act[aa];
    }
    double f = m->act_gainprm[a] * input + m->act_biasprm[3 * a] + m->act_biasprm[3 * a + 1] * e->act_length[a]
               + m->act_biasprm[3 * a + 2] * e->act_velocity[a];
    if (m->act_forcelimited[a]) f = fmin(m->act_forcerange[2 * a + 1], fmax(m->act_forcerange[2 * a], f));
    if (e->disable_actuation) f = 0;
    e->act_force[a] = f;
    const double* mom = e->act_moment + (size_t)a * nv;
    for (int d = 0; d < nv; d++) e->qfrc_actuator[d] += mom[d] * f;
  }
}
static void fwd_acceleration(FlyoEnv* e) {
  const FlybModel* m = e->m;
  for (int d = 0; d < m->nv; d++) e->qfrc_smooth[d] = e->qfrc_passive[d] - e->qfrc_bias[d] + e->qfrc_actuator[d];
  memcpy(e->qacc_smooth, e->qfrc_smooth, m->nv * sizeof(double));
  solve_tree(m, e->qLD, e->qacc_smooth);
}

/* ---------------------------------------------------------------- m11 constraint solver (Newton, primal) */
typedef struct {
  int nv, nefc;
  double *Ma, *jar, *grad, *Mgrad, *search, *Mv, *jv, *H, *hcone; /* hcone: 9 per contact */
  double cost, gauss;
} Solver;

static double update_constraint(FlyoEnv* e, const double* jar, double* force, int* state, double* hcone) {
  double cost = 0;
  for (int r = 0; r < e->nefc;) {
    if (e->efc_type[r] != CT_ELLIPTIC) {
      if (jar[r] < 0) { force[r] = -e->efc_D[r] * jar[r]; state[r] = ST_QUADRATIC; cost += 0.5 * e->efc_D[r] * jar[r] * jar[r]; }
      else { force[r] = 0; state[r] = ST_SATISFIED; }
      r++;
      continue;
    }
    const FlyoContact* c = e->con + e->efc_id[r];
    int dim = c->dim;
    double mu = c->mu, fr[3] = {mu, c->friction[0], c->friction[0]};
    double U[3], T2 = 0;
    for (int k = 0; k < dim; k++) U[k] = jar[r + k] * fr[k];
    for (int k = 1; k < dim; k++) T2 += U[k] * U[k];
    double N = U[0], T = sqrt(T2);
    if (N >= mu * T || (T <= 0 && N >= 0)) {
      for (int k = 0; k < dim; k++) { force[r + k] = 0; state[r + k] = ST_SATISFIED; }
    } else if (mu * N + T <= 0 || (T <= 0 && N < 0)) {
      for (int k = 0; k < dim; k++) {
        force[r + k] = -e->efc_D[r + k] * jar[r + k]; state[r + k] = ST_QUADRATIC;
        cost += 0.5 * e->efc_D[r + k] * jar[r + k] * jar[r + k];
      }
    } else {
      double Dm = e->efc_D[r] / fmax(MINVAL, mu * mu * (1 + mu * mu));
      double NT = N - mu * T;
      cost += 0.5 * Dm * NT * NT;
      force[r] = -Dm * NT * mu;
      for (int k = 1; k < dim; k++) force[r + k] = Dm * NT * mu / T * U[k] * fr[k];
      for (int k = 0; k < dim; k++) state[r + k] = ST_CONE;
      if (hcone) {
        double* h = hcone + 9 * e->efc_id[r];
        double Hu[9];
        Hu[0] = 1;
        for (int i = 1; i < dim; i++) Hu[i] = Hu[3 * i] = -mu * U[i] / T;
        for (int i = 1; i < dim; i++)
          for (int j = 1; j < dim; j++)
            Hu[3 * i + j] = mu * N / (T * T * T) * U[i] * U[j] + (i == j ? mu * mu - mu * N / T : 0);
        for (int i = 0; i < dim; i++) for (int j = 0; j < dim; j++) h[3 * i + j] = Dm * fr[i] * fr[j] * Hu[3 * i + j];
      }
    }
    r += dim;
  }
  return cost;
}

static int cholesky(double* A, int n) { /* lower, in place */
  for (int j = 0; j < n; j++) {
    double s = A[j * n + j];
    for (int k = 0; k < j; k++) s -= A[j * n + k] * A[j * n + k];
    if (s < MINVAL) s = MINVAL;
    double d = sqrt(s);
    A[j * n + j] = d;
    for (int i = j + 1; i < n; i++) {
      double t = A[i * n + j];
      for (int k = 0; k < j; k++) t -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = t / d;
    }
  }
  return 0;
}
static void chol_solve(const double* L, int n, double* x) {
  for (int i = 0; i < n; i++) { double s = x[i]; for (int k = 0; k < i; k++) s -= L[i * n + k] * x[k]; x[i] = s / L[i * n + i]; }
  for (int i = n - 1; i >= 0; i--) { double s = x[i]; for (int k = i + 1; k < n; k++) s -= L[k * n + i] * x[k]; x[i] = s / L[i * n + i]; }
}

/* phi(alpha) along the search direction: value and first two derivatives */
static void ls_eval(FlyoEnv* e, const Solver* S, double alpha, const double* quadGauss, double* out) {
  double c = quadGauss[0] + alpha * quadGauss[1] + alpha * alpha * quadGauss[2];
  double d1 = quadGauss[1] + 2 * alpha * quadGauss[2], d2 = 2 * quadGauss[2];
  for (int r = 0; r < e->nefc;) {
    if (e->efc_type[r] != CT_ELLIPTIC) {
      double x = S->jar[r] + alpha * S->jv[r];
      if (x < 0) { double D = e->efc_D[r]; c += 0.5 * D * x * x; d1 += D * x * S->jv[r]; d2 += D * S->jv[r] * S->jv[r]; }
      r++;
      continue;
    }
    const FlyoContact* cn = e->con + e->efc_id[r];
    int dim = cn->dim;
    double mu = cn->mu, fr[3] = {mu, cn->friction[0], cn->friction[0]};
    double x[3], v[3], T2 = 0, TV = 0, VV = 0;
    for (int k = 0; k < dim; k++) { x[k] = (S->jar[r + k] + alpha * S->jv[r + k]) * fr[k]; v[k] = S->jv[r + k] * fr[k]; }
    for (int k = 1; k < dim; k++) { T2 += x[k] * x[k]; TV += x[k] * v[k]; VV += v[k] * v[k]; }
    double N = x[0], T = sqrt(T2);
    if (N >= mu * T || (T <= 0 && N >= 0)) {
    } else if (mu * N + T <= 0 || (T <= 0 && N < 0)) {
      for (int k = 0; k < dim; k++) {
        double D = e->efc_D[r + k], xx = S->jar[r + k] + alpha * S->jv[r + k];
        c += 0.5 * D * xx * xx; d1 += D * xx * S->jv[r + k]; d2 += D * S->jv[r + k] * S->jv[r + k];
      }
    } else {
      double Dm = e->efc_D[r] / fmax(MINVAL, mu * mu * (1 + mu * mu));
      double NT = N - mu * T, N1 = v[0], T1 = TV / T, Tdd = VV / T - T1 * T1 / T;
      double g1 = N1 - mu * T1;
      c += 0.5 * Dm * NT * NT; d1 += Dm * NT * g1; d2 += Dm * (g1 * g1 - NT * mu * Tdd);
    }
    r += dim;
  }
  out[0] = c; out[1] = d1; out[2] = d2;
}

static double line_search(FlyoEnv* e, Solver* S) {
  const FlybModel* m = e->m;
  int nv = S->nv;
  double snorm = 0, q1 = 0, q2 = 0;
  for (int d = 0; d < nv; d++) {
    snorm += S->search[d] * S->search[d];
    q1 += S->search[d] * (S->Ma[d] - e->qfrc_smooth[d]);
    q2 += 0.5 * S->search[d] * S->Mv[d];
  }
  snorm = sqrt(snorm);
  if (snorm < MINVAL) return 0;
  double quadGauss[3] = {S->gauss, q1, q2};
  double gtol = m->opt_tolerance * m->opt_ls_tolerance * snorm * m->stat_meaninertia * (nv > 1 ? nv : 1);
  double p[3], lo = 0, hi = -1, alpha = 0;
  ls_eval(e, S, 0, quadGauss, p);
  if (p[1] >= 0) return 0;  /* not a descent direction */
  double dlo = p[1];
  for (int it = 0; it < m->opt_ls_iterations; it++) {
    double step = (p[2] > MINVAL) ? -p[1] / p[2] : 0;
    double next = alpha + step;
    if (hi < 0) { if (!(next > lo)) next = alpha > 0 ? 2 * alpha : 1.0; }
    else if (!(next > lo && next < hi)) next = 0.5 * (lo + hi);
    alpha = next;
    ls_eval(e, S, alpha, quadGauss, p);
    if (fabs(p[1]) < gtol) break;
    if (p[1] < 0) { lo = alpha; dlo = p[1]; } else hi = alpha;
  }
  (void)dlo;
  return alpha;
}

static void jt_mul(FlyoEnv* e, double* out, const double* f) {
  int nv = e->m->nv;
  memset(out, 0, nv * sizeof(double));
  for (int r = 0; r < e->nefc; r++) {
    if (f[r] == 0) continue;
    const double* J = e->efc_J + (size_t)r * nv;
    for (int d = 0; d < nv; d++) out[d] += J[d] * f[r];
  }
}
static void j_mul(FlyoEnv* e, double* out, const double* v) {
  int nv = e->m->nv;
  for (int r = 0; r < e->nefc; r++) {
    const double* J = e->efc_J + (size_t)r * nv;
    double s = 0;
    for (int d = 0; d < nv; d++) s += J[d] * v[d];
    out[r] = s;
  }
}

static void newton_solve(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nv = m->nv, nefc = e->nefc;
  e->solver_niter = 0;
  if (nefc == 0) {
    memcpy(e->qacc, e->qacc_smooth, nv * sizeof(double));
    memset(e->qfrc_constraint, 0, nv * sizeof(double));
    return;
  }
  Solver S; S.nv = nv; S.nefc = nefc;
  size_t n1 = nv, n2 = nefc;
  double* buf = (double*)calloc(6 * n1 + 2 * n2 + n1 * n1 + 9 * FLYO_MAXCON, sizeof(double));
  S.Ma = buf; S.grad = S.Ma + n1; S.Mgrad = S.grad + n1; S.search = S.Mgrad + n1; S.Mv = S.search + n1;
  double* tmp = S.Mv + n1; S.jar = tmp + n1; S.jv = S.jar + n2; S.H = S.jv + n2; S.hcone = S.H + n1 * n1;
  double scale = 1.0 / (m->stat_meaninertia * (nv > 1 ? nv : 1));
  /* warm start: the better of qacc_warmstart and qacc_smooth */
  {
    double *f = (double*)calloc(n2, sizeof(double)); int* st = (int*)calloc(n2, sizeof(int));
    memcpy(e->qacc, e->qacc_warmstart, nv * sizeof(double));
    j_mul(e, S.jar, e->qacc);
    for (int r = 0; r < nefc; r++) S.jar[r] -= e->efc_aref[r];
    mul_M(m, e->qM, S.Ma, e->qacc);
    double cw = update_constraint(e, S.jar, f, st, NULL);
    for (int d = 0; d < nv; d++) cw += 0.5 * (S.Ma[d] - e->qfrc_smooth[d]) * (e->qacc[d] - e->qacc_smooth[d]);
    j_mul(e, S.jar, e->qacc_smooth);
    for (int r = 0; r < nefc; r++) S.jar[r] -= e->efc_aref[r];
    double cs = update_constraint(e, S.jar, f, st, NULL);
    if (e->disable_warmstart || cs < cw) memcpy(e->qacc, e->qacc_smooth, nv * sizeof(double));
    free(f); free(st);
  }
  mul_M(m, e->qM, S.Ma, e->qacc);
  j_mul(e, S.jar, e->qacc);
  for (int r = 0; r < nefc; r++) S.jar[r] -= e->efc_aref[r];
  for (int it = 0;; it++) {
    /* cost, forces, gradient */
    S.cost = update_constraint(e, S.jar, e->efc_force, e->efc_state, S.hcone);
    S.gauss = 0;
    for (int d = 0; d < nv; d++) S.gauss += 0.5 * (S.Ma[d] - e->qfrc_smooth[d]) * (e->qacc[d] - e->qacc_smooth[d]);
    S.cost += S.gauss;
    jt_mul(e, e->qfrc_constraint, e->efc_force);
    double gn = 0;
    for (int d = 0; d < nv; d++) { S.grad[d] = S.Ma[d] - e->qfrc_smooth[d] - e->qfrc_constraint[d]; gn += S.grad[d] * S.grad[d]; }
    if (it > 0 && scale * sqrt(gn) < m->opt_tolerance) break;
    if (it >= m->opt_iterations) break;
    /* Hessian H = M + J^T D J (+ cone blocks), dense Cholesky */
    memset(S.H, 0, n1 * n1 * sizeof(double));
    for (int i = 0; i < nv; i++)
      for (int j = i; j >= 0; j = m->dof_parentid[j]) { S.H[i * n1 + j] = e->qM[i * n1 + j]; S.H[j * n1 + i] = e->qM[i * n1 + j]; }
    for (int r = 0; r < nefc;) {
      if (e->efc_state[r] == ST_QUADRATIC) {
        const double* J = e->efc_J + (size_t)r * nv; double D = e->efc_D[r];
        for (int i = 0; i < nv; i++) if (J[i] != 0) for (int j = 0; j <= i; j++) S.H[i * n1 + j] += D * J[i] * J[j];
        r++;
      } else if (e->efc_state[r] == ST_CONE) {
        int dim = e->con[e->efc_id[r]].dim; const double* h = S.hcone + 9 * e->efc_id[r];
        for (int a = 0; a < dim; a++) for (int b = 0; b < dim; b++) {
          const double *Ja = e->efc_J + (size_t)(r + a) * nv, *Jb = e->efc_J + (size_t)(r + b) * nv; double w = h[3 * a + b];
          for (int i = 0; i < nv; i++) if (Ja[i] != 0) for (int j = 0; j <= i; j++) S.H[i * n1 + j] += w * Ja[i] * Jb[j];
        }
        r += dim;
      } else r++;
    }
    for (int i = 0; i < nv; i++) for (int j = 0; j < i; j++) S.H[j * n1 + i] = S.H[i * n1 + j];
    cholesky(S.H, nv);
    memcpy(S.Mgrad, S.grad, nv * sizeof(double));
    chol_solve(S.H, nv, S.Mgrad);
    for (int d = 0; d < nv; d++) S.search[d] = -S.Mgrad[d];
    mul_M(m, e->qM, S.Mv, S.search);
    j_mul(e, S.jv, S.search);
    double alpha = line_search(e, &S);
    if (alpha == 0) break;
    for (int d = 0; d < nv; d++) { e->qacc[d] += alpha * S.search[d]; S.Ma[d] += alpha * S.Mv[d]; }
    for (int r = 0; r < nefc; r++) S.jar[r] += alpha * S.jv[r];
    double old = S.cost;
    double nc = update_constraint(e, S.jar, e->efc_force, e->efc_state, NULL);
    for (int d = 0; d < nv; d++) nc += 0.5 * (S.Ma[d] - e->qfrc_smooth[d]) * (e->qacc[d] - e->qacc_smooth[d]);
    e->solver_niter = it + 1;
    if (scale * (old - nc) < m->opt_tolerance) {
      S.cost = update_constraint(e, S.jar, e->efc_force, e->efc_state, NULL);
      jt_mul(e, e->qfrc_constraint, e->efc_force);
      break;
    }
  }
  free(buf);
}

/* ---------------------------------------------------------------- noslip (m11, second half) */
static int qcqp2(double* res, const double* A, const double* b, const double* d, double r) {
  double b1 = b[0] * d[0], b2 = b[1] * d[1];
  double A11 = A[0] * d[0] * d[0], A22 = A[3] * d[1] * d[1], A12 = A[1] * d[0] * d[1];
  double la = 0, v1 = 0, v2 = 0;
  for (int it = 0; it < 20; it++) {
    double det = (A11 + la) * (A22 + la) - A12 * A12;
    if (det < 1e-10) { res[0] = res[1] = 0; return 0; }
    double P11 = (A22 + la) / det, P22 = (A11 + la) / det, P12 = -A12 / det;
    v1 = -P11 * b1 - P12 * b2; v2 = -P12 * b1 - P22 * b2;
    double val = v1 * v1 + v2 * v2 - r * r;
    if (val < 1e-10) break;
    double deriv = -2 * (P11 * v1 * v1 + 2 * P12 * v1 * v2 + P22 * v2 * v2);
    double delta = -val / deriv;
    if (delta < 1e-10) break;
    la += delta;
  }
  res[0] = v1 * d[0]; res[1] = v2 * d[1];
  return la != 0;
}

static void noslip(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nv = m->nv, nefc = e->nefc, nfr = 0;
  if (e->disable_noslip || m->opt_noslip_iterations <= 0) return;
  for (int r = 0; r < nefc; r++) if (e->efc_type[r] == CT_ELLIPTIC && e->efc_id[r] >= 0 && r != e->con[e->efc_id[r]].efc_address) nfr++;
  if (!nfr) return;
  /* A = J M^-1 J^T rows for the friction rows only */
  double* X = (double*)calloc((size_t)nefc * nv, sizeof(double));   /* M^-1 J_r^T for friction rows */
  double* A = (double*)calloc((size_t)nefc * nefc, sizeof(double));
  for (int r = 0; r < nefc; r++) {
    if (!(e->efc_type[r] == CT_ELLIPTIC && r != e->con[e->efc_id[r]].efc_address)) continue;
    double* x = X + (size_t)r * nv;
    memcpy(x, e->efc_J + (size_t)r * nv, nv * sizeof(double));
    solve_tree(m, e->qLD, x);
    for (int c = 0; c < nefc; c++) {
      const double* J = e->efc_J + (size_t)c * nv; double s = 0;
      for (int d = 0; d < nv; d++) s += J[d] * x[d];
      A[(size_t)r * nefc + c] = s;
    }
  }
  j_mul(e, e->efc_b, e->qacc_smooth);
  for (int r = 0; r < nefc; r++) e->efc_b[r] -= e->efc_aref[r];
  double scale = 1.0 / (m->stat_meaninertia * (nv > 1 ? nv : 1));
  double* f = e->efc_force;
  for (int it = 0; it < m->opt_noslip_iterations; it++) {
    double improvement = 0;
    for (int i = 0; i < e->ncon; i++) {
      const FlyoContact* c = e->con + i;
      if (c->efc_address < 0 || c->dim < 3) continue;
      int r0 = c->efc_address;
      double res[2], old[2] = {f[r0 + 1], f[r0 + 2]}, Ac[4], bc[2], v[2];
      for (int j = 0; j < 2; j++) {
        double s = e->efc_b[r0 + 1 + j];
        for (int k = 0; k < nefc; k++) s += A[(size_t)(r0 + 1 + j) * nefc + k] * f[k];
        res[j] = s;
      }
      for (int a = 0; a < 2; a++) for (int b = 0; b < 2; b++) Ac[2 * a + b] = A[(size_t)(r0 + 1 + a) * nefc + r0 + 1 + b];
      if (f[r0] < MINVAL) { v[0] = v[1] = 0; }
      else {
        for (int a = 0; a < 2; a++) bc[a] = res[a] - Ac[2 * a] * old[0] - Ac[2 * a + 1] * old[1];
        double mu[2] = {c->friction[0], c->friction[0]};
        int active = qcqp2(v, Ac, bc, mu, f[r0]);
        if (active) {
          double s = v[0] * v[0] / (mu[0] * mu[0]) + v[1] * v[1] / (mu[1] * mu[1]);
          s = sqrt(f[r0] * f[r0] / fmax(MINVAL, s));
          v[0] *= s; v[1] *= s;
        }
      }
      f[r0 + 1] = v[0]; f[r0 + 2] = v[1];
      double dl[2] = {v[0] - old[0], v[1] - old[1]};
      double change = 0.5 * (dl[0] * (Ac[0] * dl[0] + Ac[1] * dl[1]) + dl[1] * (Ac[2] * dl[0] + Ac[3] * dl[1])) + dl[0] * res[0] + dl[1] * res[1];
      improvement -= change;
    }
    if (improvement * scale < m->opt_noslip_tolerance) break;
  }
  jt_mul(e, e->qfrc_constraint, f);
  free(X); free(A);
}

static void fwd_constraint(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nv = m->nv;
  j_mul(e, e->efc_vel, e->qvel);
  for (int r = 0; r < e->nefc; r++)
    e->efc_aref[r] = -e->efc_KBIP[4 * r + 1] * e->efc_vel[r] - e->efc_KBIP[4 * r] * e->efc_KBIP[4 * r + 2] * (e->efc_pos[r] - e->efc_margin[r]);
  newton_solve(e);
  if (e->nefc) {
    noslip(e);
    /* qacc = qacc_smooth + M^-1 J^T f (exact after noslip, consistent with the primal optimum before it) */
    if (!e->disable_noslip && m->opt_noslip_iterations > 0) {
      double* t = (double*)malloc(nv * sizeof(double));
      memcpy(t, e->qfrc_constraint, nv * sizeof(double));
      solve_tree(m, e->qLD, t);
      for (int d = 0; d < nv; d++) e->qacc[d] = e->qacc_smooth[d] + t[d];
      free(t);
    }
  }
}

/* ---------------------------------------------------------------- m12/m14 sensors */
static double ray_capsule(const double* pos, const double* mat, const double* size, const double* pnt, const double* vec) {
  /* nearest non-negative ray parameter hitting the capsule surface (or -1); origin inside counts as a hit */
  double d[3] = {pnt[0] - pos[0], pnt[1] - pos[1], pnt[2] - pos[2]}, lp[3], lv[3];
  matT_vec(lp, mat, d); matT_vec(lv, mat, vec);
  double r = size[0], h = size[1], best = -1;
  double a = lv[0] * lv[0] + lv[1] * lv[1], b = lp[0] * lv[0] + lp[1] * lv[1], c = lp[0] * lp[0] + lp[1] * lp[1] - r * r;
  /* inside test */
  double zc = fmin(h, fmax(-h, lp[2]));
  if (lp[0] * lp[0] + lp[1] * lp[1] + (lp[2] - zc) * (lp[2] - zc) <= r * r) return 0;
  if (a > MINVAL) {
    double det = b * b - a * c;
    if (det >= 0) {
      double x = (-b - sqrt(det)) / a;
      if (x >= 0 && fabs(lp[2] + x * lv[2]) <= h) best = x;
    }
  }
  for (int sgn = -1; sgn <= 1; sgn += 2) {
    double q[3] = {lp[0], lp[1], lp[2] - sgn * h};
    double aa = dot3(lv, lv), bb = dot3(q, lv), cc = dot3(q, q) - r * r, det = bb * bb - aa * cc;
    if (det < 0 || aa < MINVAL) continue;
    double x = (-bb - sqrt(det)) / aa;
    if (x >= 0 && sgn * (lp[2] + x * lv[2]) >= h && (best < 0 || x < best)) best = x;
  }
  return best;
}

static void sensors_acc(FlyoEnv* e) {
  const FlybModel* m = e->m;
  memset(e->cfrc_ext, 0, 6 * m->nbody * sizeof(double));
  for (int i = 0; i < e->ncon; i++) {
    const FlyoContact* c = e->con + i;
    if (c->efc_address < 0) continue;
    double fw[3] = {0, 0, 0};
    for (int k = 0; k < c->dim; k++) for (int a = 0; a < 3; a++) fw[a] += c->frame[3 * k + a] * e->efc_force[c->efc_address + k];
    int bs[2] = {m->geom_bodyid[c->geom1], m->geom_bodyid[c->geom2]};
    for (int s = 0; s < 2; s++) {
      int b = bs[s]; if (b <= 0) continue;
      double sg = s ? 1.0 : -1.0;
      const double* c0 = e->subtree_com + 3 * m->body_rootid[b];
      double off[3] = {c->pos[0] - c0[0], c->pos[1] - c0[1], c->pos[2] - c0[2]}, tq[3];
      cross3(tq, off, fw);
      for (int a = 0; a < 3; a++) { e->cfrc_ext[6 * b + a] += sg * tq[a]; e->cfrc_ext[6 * b + 3 + a] += sg * fw[a]; }
    }
  }
  rne(e, NULL, 1, 1);
  for (int s = 0; s < m->nsensor; s++) {
    int site = m->sensor_site[s], b = m->site_bodyid[site], type = m->sensor_type[s];
    double* out = e->sensordata + m->sensor_adr[s];
    const double *sp = e->site_xpos + 3 * site, *sm = e->site_xmat + 9 * site;
    if (type == FLYB_SENS_ACCELEROMETER) {
      const double* c0 = e->subtree_com + 3 * m->body_rootid[b];
      const double* ca = e->cacc + 6 * b;
      double off[3] = {sp[0] - c0[0], sp[1] - c0[1], sp[2] - c0[2]}, t[3], lin[3], lv[6], cor[3], la[3];
      cross3(t, ca, off);
      for (int k = 0; k < 3; k++) lin[k] = ca[3 + k] + t[k];
      matT_vec(la, sm, lin);
      local_velocity(e, lv, b, sp, sm);
      cross3(cor, lv, lv + 3);
      for (int k = 0; k < 3; k++) out[k] = la[k] + cor[k];
    } else if (type == FLYB_SENS_FORCE) {
      matT_vec(out, sm, e->cfrc_int + 6 * b + 3);
    } else if (type == FLYB_SENS_TOUCH) {
      double sum = 0;
      for (int i = 0; i < e->ncon; i++) {
        const FlyoContact* c = e->con + i;
        if (c->efc_address < 0) continue;
        int b1 = m->geom_bodyid[c->geom1], b2 = m->geom_bodyid[c->geom2];
        if (b1 != b && b2 != b) continue;
        double fn = e->efc_force[c->efc_address];
        if (fn <= 0) continue;
        double ray[3] = {c->frame[0], c->frame[1], c->frame[2]};
        if (b2 == b) { ray[0] = -ray[0]; ray[1] = -ray[1]; ray[2] = -ray[2]; }
        if (ray_capsule(sp, sm, m->site_size + 3 * site, c->pos, ray) >= 0) sum += fn;
      }
      out[0] = sum;
    }
  }
}
static void sensors_vel(FlyoEnv* e) {
  const FlybModel* m = e->m;
  for (int s = 0; s < m->nsensor; s++) {
    int site = m->sensor_site[s], type = m->sensor_type[s];
    if (type != FLYB_SENS_GYRO && type != FLYB_SENS_VELOCIMETER) continue;
    double lv[6];
    local_velocity(e, lv, m->site_bodyid[site], e->site_xpos + 3 * site, e->site_xmat + 9 * site);
    memcpy(e->sensordata + m->sensor_adr[s], type == FLYB_SENS_GYRO ? lv : lv + 3, 3 * sizeof(double));
  }
}

/* ---------------------------------------------------------------- pipeline */
static void fwd_position(FlyoEnv* e) {
  kinematics(e); com_pos(e); crb(e);
  memcpy(e->qLD, e->qM, (size_t)e->m->nv * e->m->nv * sizeof(double));
  factor_tree(e->m, e->qLD);
  collision(e); make_constraint(e); transmission(e);
}
static void fwd_velocity(FlyoEnv* e) {
  const FlybModel* m = e->m;
  for (int a = 0; a < m->nu; a++) {
    double s = 0; const double* mom = e->act_moment + (size_t)a * m->nv;
    for (int d = 0; d < m->nv; d++) s += mom[d] * e->qvel[d];
    e->act_velocity[a] = s;
  }
  com_vel(e); passive(e); rne(e, e->qfrc_bias, 0, 0);
}
static void step1(FlyoEnv* e) { fwd_position(e); fwd_velocity(e); sensors_vel(e); }
static void step2(FlyoEnv* e) {
  const FlybModel* m = e->m;
  int nv = m->nv;
  double dt = m->opt_timestep;
  fwd_actuation(e); fwd_acceleration(e); fwd_constraint(e); sensors_acc(e);
  /* Euler with implicit joint damping (m13) */
  double* H = (double*)malloc((size_t)nv * nv * sizeof(double));
  double* a = (double*)malloc(nv * sizeof(double));
  memcpy(H, e->qM, (size_t)nv * nv * sizeof(double));
  for (int d = 0; d < nv; d++) { H[d * nv + d] += dt * m->dof_damping[d]; a[d] = e->qfrc_smooth[d] + e->qfrc_constraint[d]; }
  factor_tree(m, H); solve_tree(m, H, a);
  for (int i = 0; i < m->na; i++) e->act[i] += dt * e->act_dot[i];
  for (int d = 0; d < nv; d++) e->qvel[d] += dt * a[d];
  for (int j = 0; j < m->njnt; j++) {
    int qa = m->jnt_qposadr[j], da = m->jnt_dofadr[j];
    if (m->jnt_type[j] == FLYB_JNT_HINGE) e->qpos[qa] += dt * e->qvel[da];
    else {
      if (m->jnt_type[j] == FLYB_JNT_FREE) { for (int k = 0; k < 3; k++) e->qpos[qa + k] += dt * e->qvel[da + k]; qa += 3; da += 3; }
      quat_integrate(e->qpos + qa, e->qvel + da, dt);
    }
  }
  e->time += dt;
  memcpy(e->qacc_warmstart, e->qacc, nv * sizeof(double));
  free(H); free(a);
}
void flyo_forward(FlyoEnv* e) { step1(e); fwd_actuation(e); fwd_acceleration(e); fwd_constraint(e); sensors_acc(e); }
void flyo_step(FlyoEnv* e) { step2(e); step1(e); }

/* ---------------------------------------------------------------- allocation / access */
typedef struct { const char* name; double** p; size_t n; } DField;
static int dfields(FlyoEnv* e, DField* f) {
  const FlybModel* m = e->m;
  size_t nq = m->nq, nv = m->nv, nb = m->nbody, nj = m->njnt, ng = m->ngeom, ns = m->nsite, nu = m->nu, ne = FLYO_MAXEFC;
  int k = 0;
#define F(nm, cnt) f[k].name = #nm; f[k].p = &e->nm; f[k].n = (cnt); k++;
  F(qpos, nq) F(qvel, nv) F(act, m->na) F(ctrl, nu) F(qacc_warmstart, nv)
  F(xpos, 3 * nb) F(xquat, 4 * nb) F(xmat, 9 * nb) F(xipos, 3 * nb) F(ximat, 9 * nb) F(xanchor, 3 * nj) F(xaxis, 3 * nj)
  F(geom_xpos, 3 * ng) F(geom_xmat, 9 * ng) F(site_xpos, 3 * ns) F(site_xmat, 9 * ns)
  F(subtree_com, 3 * nb) F(cinert, 10 * nb) F(cdof, 6 * nv) F(crb, 10 * nb) F(qM, nv * nv) F(qLD, nv * nv)
  F(ten_length, m->ntendon) F(act_length, nu) F(act_moment, nu * nv)
  F(efc_J, ne * nv) F(efc_pos, ne) F(efc_margin, ne) F(efc_diagApprox, ne) F(efc_R, ne) F(efc_D, ne) F(efc_KBIP, 4 * ne)
  F(efc_vel, ne) F(efc_aref, ne) F(efc_force, ne) F(efc_b, ne)
  F(cvel, 6 * nb) F(cdof_dot, 6 * nv) F(qfrc_bias, nv) F(qfrc_passive, nv) F(act_velocity, nu)
  F(act_dot, m->na) F(act_force, nu) F(qfrc_actuator, nv) F(qfrc_smooth, nv) F(qacc_smooth, nv) F(qacc, nv)
  F(qfrc_constraint, nv) F(cacc, 6 * nb) F(cfrc_int, 6 * nb) F(cfrc_ext, 6 * nb) F(sensordata, m->nsensordata)
  F(sens_sum, m->nsensordata) F(prev_action, m->action_dim)
#undef F
  return k;
}

FlyoEnv* flyo_create(const FlybModel* m) {
  FlyoEnv* e = (FlyoEnv*)calloc(1, sizeof(FlyoEnv));
  e->mcopy = *m;       /* scalar fields may be changed per env through flyo_set_option; arrays are shared */
  e->m = &e->mcopy;
  DField f[96];
  int n = dfields(e, f);
  for (int i = 0; i < n; i++) *f[i].p = (double*)calloc(f[i].n ? f[i].n : 1, sizeof(double));
  e->efc_type = (int*)calloc(FLYO_MAXEFC, sizeof(int));
  e->efc_id = (int*)calloc(FLYO_MAXEFC, sizeof(int));
  e->efc_state = (int*)calloc(FLYO_MAXEFC, sizeof(int));
  flyo_reset_data(e);
  return e;
}
void flyo_destroy(FlyoEnv* e) {
  if (!e) return;
  DField f[96];
  int n = dfields(e, f);
  for (int i = 0; i < n; i++) free(*f[i].p);
  free(e->efc_type); free(e->efc_id); free(e->efc_state);
  free(e);
}
static void init_world(FlyoEnv* e) {
  e->xquat[0] = 1; e->xmat[0] = e->xmat[4] = e->xmat[8] = 1; e->ximat[0] = e->ximat[4] = e->ximat[8] = 1;
}
void flyo_reset_data(FlyoEnv* e) {
  const FlybModel* m = e->m;
  DField f[96];
  int n = dfields(e, f);
  for (int i = 0; i < n; i++) memset(*f[i].p, 0, f[i].n * sizeof(double));
  memcpy(e->qpos, m->qpos0, m->nq * sizeof(double));
  e->time = 0; e->ncon = 0; e->nefc = 0;
  init_world(e);
  flyo_forward(e);
}
double* flyo_dptr(FlyoEnv* e, const char* name, int* count) {
  DField f[96];
  int n = dfields(e, f);
  for (int i = 0; i < n; i++)
    if (!strcmp(f[i].name, name)) { if (count) *count = (int)f[i].n; return *f[i].p; }
  if (!strcmp(name, "time")) { if (count) *count = 1; return &e->time; }
  return NULL;
}
int* flyo_iptr(FlyoEnv* e, const char* name, int* count) {
  if (!strcmp(name, "nefc")) { if (count) *count = 1; return &e->nefc; }
  if (!strcmp(name, "ncon")) { if (count) *count = 1; return &e->ncon; }
  if (!strcmp(name, "nlimit")) { if (count) *count = 1; return &e->nlimit; }
  if (!strcmp(name, "solver_niter")) { if (count) *count = 1; return &e->solver_niter; }
  if (!strcmp(name, "efc_type")) { if (count) *count = e->nefc; return e->efc_type; }
  if (!strcmp(name, "efc_id")) { if (count) *count = e->nefc; return e->efc_id; }
  if (!strcmp(name, "efc_state")) { if (count) *count = e->nefc; return e->efc_state; }
  if (!strcmp(name, "step_counter")) { if (count) *count = 1; return &e->step_counter; }
  return NULL;
}
const FlyoContact* flyo_contacts(FlyoEnv* e, int* ncon) { *ncon = e->ncon; return e->con; }
void flyo_set_option(FlyoEnv* e, const char* name, double v) {
  FlybModel* m = &e->mcopy;
  if (!strcmp(name, "disable_contact")) e->disable_contact = (int)v;
  else if (!strcmp(name, "disable_limit")) e->disable_limit = (int)v;
  else if (!strcmp(name, "disable_fluid")) e->disable_fluid = (int)v;
  else if (!strcmp(name, "disable_noslip")) e->disable_noslip = (int)v;
  else if (!strcmp(name, "disable_actuation")) e->disable_actuation = (int)v;
  else if (!strcmp(name, "disable_warmstart")) e->disable_warmstart = (int)v;
  else if (!strcmp(name, "timestep")) m->opt_timestep = v;
  else if (!strcmp(name, "tolerance")) m->opt_tolerance = v;
  else if (!strcmp(name, "iterations")) m->opt_iterations = (int)v;
  else if (!strcmp(name, "noslip_iterations")) m->opt_noslip_iterations = (int)v;
  else if (!strcmp(name, "pad_first_obs")) m->pad_first_obs = (int)v;
  else if (!strcmp(name, "terminal_com_dist")) m->terminal_com_dist = v;
  else fprintf(stderr, "flyo_set_option: unknown option %s\n", name);
}

/* ---------------------------------------------------------------- task layer (stubs filled below) */
void flyo_set_clip(FlyoEnv* e, const FlyoClip* clip) { e->clip = *clip; e->has_clip = 1; }
void flyo_set_wbpg(FlyoEnv* e, const FlyoWbpg* w) { e->wb = *w; e->has_wbpg = 1; }
void flyo_set_initial_phase(FlyoEnv* e, double phase) { e->init_phase = phase; }
void flyo_wbpg_state(const FlyoEnv* e, int* step, int* freq_idx, double* ctrl_freq) {
  *step = e->wb_step; *freq_idx = e->wb_fidx; *ctrl_freq = e->wb_freq;
}

/* ---- wing-beat pattern generator state machine: tasks/pattern_generators.py:121-191 (numpy argmin = first minimum) */
static int wbpg_nearest_freq(const FlyoWbpg* w, double f) {
  int best = 0;
  double bd = fabs(w->beat_freqs[0] - f);
  for (int i = 1; i < w->nfreq; i++) { double d = fabs(w->beat_freqs[i] - f); if (d < bd) { bd = d; best = i; } }
  return best;
}
static void wbpg_reset(FlyoEnv* e, double phase0, double* ang, double* vel) {  /* :121-157 */
  const FlyoWbpg* w = &e->wb;
  e->wb_freq = w->base_freq;
  e->wb_fidx = wbpg_nearest_freq(w, e->wb_freq);
  int o = w->offsets[e->wb_fidx], n = w->offsets[e->wb_fidx + 1] - o, best = 0;
  double bd = fabs(phase0 - w->phase[o]);
  for (int i = 1; i < n; i++) { double d = fabs(phase0 - w->phase[o + i]); if (d < bd) { bd = d; best = i; } }
  e->wb_step = best;
  for (int k = 0; k < 6; k++) {
    ang[k] = w->traj[(size_t)(o + best) * 6 + k];
    vel[k] = (w->traj[(size_t)(o + best + 1) * 6 + k] - ang[k]) / w->dt_ctrl;
  }
}
static void wbpg_step(FlyoEnv* e, double ctrl_freq, double* ang) {  /* :159-191 */
  const FlyoWbpg* w = &e->wb;
  int o = w->offsets[e->wb_fidx], n = w->offsets[e->wb_fidx + 1] - o;
  e->wb_step = (e->wb_step + 1) % n;
  e->wb_freq = w->rate == 0.0 ? ctrl_freq : e->wb_freq * w->rate + ctrl_freq * (1 - w->rate);
  int nf = wbpg_nearest_freq(w, e->wb_freq);
  if (nf != e->wb_fidx) {
    double cur = fmod(w->phase[o + e->wb_step], 1.0);
    int o2 = w->offsets[nf], n2 = w->offsets[nf + 1] - o2, best = 0;
    double bd = fabs(cur - fmod(w->phase[o2], 1.0));
    for (int i = 1; i < n2; i++) { double d = fabs(cur - fmod(w->phase[o2 + i], 1.0)); if (d < bd) { bd = d; best = i; } }
    e->wb_step = best; e->wb_fidx = nf; o = o2;
  }
  for (int k = 0; k < 6; k++) ang[k] = w->traj[(size_t)(o + e->wb_step) * 6 + k];
}

/* quaternion helpers for the task math (vnl_ray/quaternions.py) */
static void quat_recip(double* r, const double* q) {  /* quaternions.py:88-102 */
  double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  r[0] = q[0] / n2; r[1] = -q[1] / n2; r[2] = -q[2] / n2; r[3] = -q[3] / n2;
}
static void quat_rotate(double* r, const double* v, const double* q) {  /* quaternions.py:105-134: q v q^-1 */
  double qi[4], va[4] = {0, v[0], v[1], v[2]}, t[4], o[4];
  quat_recip(qi, q);
  quat_mul(t, va, qi); quat_mul(o, q, t);
  r[0] = o[1]; r[1] = o[2]; r[2] = o[3];
}
static void quat_z2vec(double* q, const double* vin) {  /* quaternions.py:205-252 */
  double v[3] = {vin[0], vin[1], vin[2]};
  int edge = (v[0] == 0.0 && v[1] == 0.0);
  if (edge) v[0] = 1.0;
  double n = norm3(v);
  v[0] /= n; v[1] /= n; v[2] /= n;
  if (edge) { if (v[2] < 0) { q[0] = 0; q[1] = 1; q[2] = 0; q[3] = 0; } else { q[0] = 1; q[1] = q[2] = q[3] = 0; } return; }
  double ax[3] = {-v[1], v[0], 0}, an = sqrt(ax[0] * ax[0] + ax[1] * ax[1]);
  ax[0] /= an; ax[1] /= an;
  double ang = acos(v[2]);
  q[0] = cos(ang / 2); q[1] = sin(ang / 2) * ax[0]; q[2] = sin(ang / 2) * ax[1]; q[3] = 0;
}
static double quat_dist_short_arc(const double* a, const double* b) {  /* quaternions.py:273-295 */
  double na = sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2] + a[3] * a[3]), nb = sqrt(b[0] * b[0] + b[1] * b[1] + b[2] * b[2] + b[3] * b[3]);
  double d = (a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3]) / (na * nb);
  double x = 2 * d * d - 1;
  if (x > 1.0) x = 1.0;
  return acos(x);
}

/* test hook: the quaternion helpers above on raw inputs, so that the reference-generated fixture (tests/golden/quaternions.npz,
 * vnl_ray/quaternions.py) checks them directly, edge cases included.  op: 0 reciprocal(a) 1 rotate(v = a, q = b) 2 z2vec(a)
 * 3 dist_short_arc(a, b) 4 joint_orientation_quat(axis = a, angle = b[0]) 5 get_dquat_local(a, b) 6 get_egocentric_vec(root =
 * a[0..3), site = a[3..6), root_quat = b) 7 mult(a, b).  Returns the number of doubles written. */
int flyo_quat_op(int op, const double* a, const double* b, double* out) {
  switch (op) {
    case 0: quat_recip(out, a); return 4;
    case 1: quat_rotate(out, a, b); return 3;
    case 2: quat_z2vec(out, a); return 4;
    case 3: out[0] = quat_dist_short_arc(a, b); return 1;
    case 4: {   /* rewards.py:47-51 as the reward code evaluates it */
      double q1[4], q2[4], n = norm3(a), a2[3] = {a[0] / n, a[1] / n, a[2] / n};
      quat_z2vec(q1, a); axis_angle_quat(q2, a2, b[0]); quat_mul(out, q2, q1); return 4;
    }
    case 5: { double qi[4]; quat_recip(qi, a); quat_mul(out, qi, b); return 4; }
    case 6: { double d[3] = {a[3] - a[0], a[4] - a[1], a[5] - a[2]}, qc[4] = {b[0], -b[1], -b[2], -b[3]}; quat_rotate(out, d, qc); return 3; }
    case 7: quat_mul(out, a, b); return 4;
  }
  return 0;
}

static const double* clip_row(const double* base, int row, int width) { return base + (size_t)row * width; }

/* observation assembly: fruitfly.py:532-669, tasks/base.py:237-261, walk_on_ball.py:81-88 */
static void write_obs(FlyoEnv* e, double* obs, double sensor_scale) {
  const FlybModel* m = e->m;
  const int* off = m->obs_offset;
  int rb = m->root_body;
  const double *rp = e->xpos + 3 * rb, *rm = e->xmat + 9 * rb;
  for (int s = 0; s < m->nsensor; s++) {
    int t = m->sensor_type[s], key, idx = 0, dim = t == FLYB_SENS_TOUCH ? 1 : 3;
    if (t == FLYB_SENS_ACCELEROMETER) key = FLYB_OBS_ACCELEROMETER;
    else if (t == FLYB_SENS_GYRO) key = FLYB_OBS_GYRO;
    else if (t == FLYB_SENS_VELOCIMETER) key = FLYB_OBS_VELOCIMETER;
    else if (t == FLYB_SENS_FORCE) key = FLYB_OBS_FORCE;
    else key = FLYB_OBS_TOUCH;
    for (int s2 = 0; s2 < s; s2++) if (m->sensor_type[s2] == t) idx++;
    if (off[key] < 0) continue;
    for (int k = 0; k < dim; k++) obs[off[key] + dim * idx + k] = e->sens_sum[m->sensor_adr[s] + k] * sensor_scale;
  }
  if (off[FLYB_OBS_ACTUATOR_ACTIVATION] >= 0) memcpy(obs + off[FLYB_OBS_ACTUATOR_ACTIVATION], e->act, m->na * sizeof(double));
  if (off[FLYB_OBS_APPENDAGES_POS] >= 0)
    for (int i = 0; i < m->n_appendages; i++) {
      const double* sp = e->site_xpos + 3 * m->appendage_site[i];
      double d[3] = {sp[0] - rp[0], sp[1] - rp[1], sp[2] - rp[2]};
      matT_vec(obs + off[FLYB_OBS_APPENDAGES_POS] + 3 * i, rm, d);
    }
  if (off[FLYB_OBS_BALL_QVEL] >= 0) memcpy(obs + off[FLYB_OBS_BALL_QVEL], e->qvel + m->ball_dofadr, 3 * sizeof(double));
  for (int i = 0; i < m->n_obs_joints; i++) {
    obs[off[FLYB_OBS_JOINTS_POS] + i] = e->qpos[m->obs_joint_qposadr[i]];
    obs[off[FLYB_OBS_JOINTS_VEL] + i] = e->qvel[m->obs_joint_dofadr[i]];
  }
  memcpy(obs + off[FLYB_OBS_WORLD_ZAXIS], rm + 6, 3 * sizeof(double));
  if (off[FLYB_OBS_TASK_INPUT] >= 0) {   /* tasks/vision_flight.py:243-252 */
    obs[off[FLYB_OBS_TASK_INPUT]] = e->target_height; obs[off[FLYB_OBS_TASK_INPUT] + 1] = e->target_speed;
  }
  if (off[FLYB_OBS_REF_DISPLACEMENT] >= 0 && e->has_clip) {
    int F = m->future_steps + 1, W = 7 + (m->task_type == FLYB_TASK_WALK_IMITATION ? m->n_mocap_joints : 0);
    const double* rq = e->qpos + m->root_qposadr + 3;
    double qi[4];
    quat_recip(qi, rq);
    for (int i = 0; i < F; i++) {
      int row = e->step_counter + i;
      if (row >= e->clip.T) row = e->clip.T - 1;
      const double* r = clip_row(e->clip.qpos, row, W);
      double d[3] = {r[0] - e->qpos[m->root_qposadr], r[1] - e->qpos[m->root_qposadr + 1], r[2] - e->qpos[m->root_qposadr + 2]};
      matT_vec(obs + off[FLYB_OBS_REF_DISPLACEMENT] + 3 * i, rm, d);
      quat_mul(obs + off[FLYB_OBS_REF_ROOT_QUAT] + 4 * i, qi, r + 3);  /* quaternions.py:13-17 */
    }
  }
}

/* DeepMimic-style reward: walk_imitation.py:151-163, rewards.py:36-111 */
static double walk_imitation_reward(FlyoEnv* e, int step) {
  const FlybModel* m = e->m;
  if (m->inference_mode || !e->has_clip) return 1.0;
  int nj = m->n_mocap_joints, ns = m->n_mocap_sites, Wq = 7 + nj, Wv = 6 + nj;
  if (step >= e->clip.T) step = e->clip.T - 1;
  const double* rq = clip_row(e->clip.qpos, step, Wq);
  const double* rv = clip_row(e->clip.qvel, step, Wv);
  const double* q = e->qpos + m->root_qposadr;
  const double* v = e->qvel + m->root_dofadr;
  double d_com = 0, d_qvel = 0, d_r2s = 0, d_jq = 0;
  for (int k = 0; k < 3; k++) d_com += (q[k] - rq[k]) * (q[k] - rq[k]);
  for (int k = 0; k < 6; k++) d_qvel += (v[k] - rv[k]) * (v[k] - rv[k]);
  double qc[4] = {q[3], -q[4], -q[5], -q[6]}, qr[4];
  quat_recip(qr, q + 3);
  { double a = quat_dist_short_arc(q + 3, rq + 3); d_jq += a * a; }
  for (int i = 0; i < nj; i++) {
    int j = m->mocap_joint[i];
    double qj = e->qpos[m->jnt_qposadr[j]], vj = e->qvel[m->jnt_dofadr[j]];
    d_qvel += (vj - rv[6 + i]) * (vj - rv[6 + i]);
    double ax[3], q1[4], q2[4], jq[4];
    quat_rotate(ax, e->xaxis + 3 * j, qr);            /* rewards.py:47-48 */
    quat_z2vec(q1, ax);                                 /* quaternions.py:298-321 */
    double n = norm3(ax), a2[3] = {ax[0] / n, ax[1] / n, ax[2] / n};
    axis_angle_quat(q2, a2, qj);
    quat_mul(jq, q2, q1);
    double a = quat_dist_short_arc(jq, e->clip.joint_quat + ((size_t)step * nj + i) * 4);
    d_jq += a * a;
  }
  for (int i = 0; i < ns; i++) {
    const double* sp = e->site_xpos + 3 * m->mocap_site[i];
    double d[3] = {sp[0] - q[0], sp[1] - q[1], sp[2] - q[2]}, r[3];
    quat_rotate(r, d, qc);                              /* quaternions.py:137-159 */
    const double* ref = e->clip.root2site + ((size_t)step * ns + i) * 3;
    for (int k = 0; k < 3; k++) d_r2s += (r[k] - ref[k]) * (r[k] - ref[k]);
  }
  const double s_com = 0.078487, s_qvel = 53.7801, s_r2s = 0.0735, s_jq = 1.2247;  /* rewards.py:99-104 */
  return 20.0 * exp(-0.5 / (s_com * s_com) * d_com) * exp(-0.5 / (s_qvel * s_qvel) * d_qvel)
         * exp(-0.5 / (s_r2s * s_r2s) * d_r2s) * exp(-0.5 / (s_jq * s_jq) * d_jq);
}

/* flight_imitation.py:169-196: CoM-distance and root-orientation factors, each a linear tolerance */
static double flight_imitation_reward(FlyoEnv* e, const double* obs) {
  const FlybModel* m = e->m;
  int gb = -1;
  for (int j = 0; j < m->njnt; j++) if (m->jnt_qposadr[j] == m->ghost_qposadr) gb = m->jnt_bodyid[j];
  if (gb < 0) return 1.0;
  const double off[3] = {-0.03697732, 0.00029205, -0.0142447};  /* task_utils.py:174-191 root2com */
  double og[3], d[3];
  quat_rotate(og, off, e->xquat + 4 * gb);
  for (int k = 0; k < 3; k++) d[k] = e->xpos[3 * gb + k] + og[k] - e->subtree_com[3 * m->root_body + k];
  double x = norm3(d) / 0.4, r = x < 1 ? 1 - x : 0.0;
  const double one[4] = {1, 0, 0, 0};
  double th = quat_dist_short_arc(one, obs + m->obs_offset[FLYB_OBS_REF_ROOT_QUAT]) / 3.14159265358979323846;
  return r * (th < 1 ? 1 - th : 0.0);
}

/* ---- vision-guided flight (tasks/vision_flight.py) */
/* index of the first minimum of |linspace(lo, hi, n) - x| (numpy.argmin of the reference: vision_flight.py:100-101, 181) */
static int nearest_linspace(double lo, double hi, int n, double x) {
  const double step = (hi - lo) / (n - 1);
  int best = 0;
  double bv = 1e300;
  for (int i = 0; i < n; i++) {
    const double a = i == n - 1 ? hi : i * step + lo, d = fabs(a - x);
    if (d < bv) { bv = d; best = i; }
  }
  return best;
}
static double hfield_height_at(const FlyoEnv* e, double x, double y) {  /* get_hfield_height, vision_flight.py:92-104 */
  const FlybModel* m = e->m;
  if (!e->hf_data) return 0.0;
  const double half = m->hfield_size[0];
  const int xi = nearest_linspace(-half, half, m->hfield_ncol, x), yi = nearest_linspace(-half, half, m->hfield_ncol, y);
  return m->hfield_size[2] * (double)e->hf_data[(size_t)yi * m->hfield_ncol + xi];
}
static double tol_linear(double x, double lo, double hi, double margin) {  /* dm_control rewards.tolerance, sigmoid "linear", value_at_margin 0 */
  if (lo <= x && x <= hi) return 1.0;
  if (margin == 0) return 0.0;
  const double d = (x < lo ? lo - x : x - hi) / margin;
  return fabs(d) < 1 ? 1 - d : 0.0;
}
static int vision_floor_contact(const FlyoEnv* e) {  /* check_floor_contact, vision_flight.py:214-227 */
  const FlybModel* m = e->m;
  for (int i = 0; i < e->ncon; i++) {
    const FlyoContact* c = e->con + i;
    if (c->efc_address < 0) continue;
    if (m->geom_bodyid[c->geom1] == 0 || m->geom_bodyid[c->geom2] == 0) return 1;
  }
  return 0;
}
static double vision_reward(FlyoEnv* e) {  /* get_reward_factors, vision_flight.py:157-212 */
  const FlybModel* m = e->m;
  const double* xpos = e->qpos + m->root_qposadr;
  const double* vel = e->qvel + m->root_dofadr;
  const double height = tol_linear(xpos[2] - hfield_height_at(e, xpos[0], xpos[1]), e->target_height, e->target_height, 0.15);
  double center = 1.0;
  if (e->n_trench > 0 && e->trench_x[0] <= xpos[0] && xpos[0] <= e->trench_x[e->n_trench - 1]) {
    int idx = 0;
    double bv = 1e300;
    for (int i = 0; i < e->n_trench; i++) { const double d = fabs(e->trench_x[i] - xpos[0]); if (d < bv) { bv = d; idx = i; } }
    center = tol_linear(xpos[1], e->trench_y[idx], e->trench_y[idx], 0.15);
  }
  const double ts = e->target_speed;
  const double x_speed = tol_linear(vel[0], ts, INFINITY, 1.1 * ts);
  const double speed = tol_linear(norm3(vel), ts, ts, 1.1 * ts);
  double side = 0, zdot = 0;
  for (int s = 0; s < m->nsensor; s++)
    if (m->sensor_type[s] == FLYB_SENS_VELOCIMETER) side = e->sensordata[m->sensor_adr[s] + 1];   /* the observable's current value */
  const double side_speed = tol_linear(side, 0, 0, 10);
  const double* zax = e->xmat + 9 * m->root_body + 6;   /* world_zaxis observable */
  for (int k = 0; k < 3; k++) zdot += m->vision_target_zaxis[k] * zax[k];
  const double world_zaxis = tol_linear(acos(zdot), 0, 0, PI);
  return height * x_speed * speed * side_speed * world_zaxis * center;
}

void flyo_set_terrain(FlyoEnv* e, const float* data, int n_trench, const double* trench_x, const double* trench_y) {
  e->hf_data = data; e->n_trench = n_trench; e->trench_x = trench_x; e->trench_y = trench_y;
}
void flyo_set_vision_targets(FlyoEnv* e, double height, double speed) { e->target_height = height; e->target_speed = speed; }

static void set_ghost(FlyoEnv* e, int row, int with_vel) {  /* walk_imitation.py:119-136,138-144 */
  const FlybModel* m = e->m;
  if (m->ghost_qposadr < 0 || !e->has_clip) return;
  int Wq = 7 + (m->task_type == FLYB_TASK_WALK_IMITATION ? m->n_mocap_joints : 0);
  int Wv = 6 + (m->task_type == FLYB_TASK_WALK_IMITATION ? m->n_mocap_joints : 0);
  if (row >= e->clip.T) row = e->clip.T - 1;
  const double* r = clip_row(e->clip.qpos, row, Wq);
  for (int k = 0; k < 7; k++) e->qpos[m->ghost_qposadr + k] = r[k] + (k < 3 ? m->ghost_offset[k] : 0);
  if (with_vel) memcpy(e->qvel + m->ghost_dofadr, clip_row(e->clip.qvel, row, Wv), 6 * sizeof(double));
}

static void accumulate_sensors(FlyoEnv* e) {
  for (int k = 0; k < e->m->nsensordata; k++) e->sens_sum[k] += e->sensordata[k];
}

int flyo_env_reset(FlyoEnv* e, double* obs) {
  const FlybModel* m = e->m;
  DField f[96];
  int n = dfields(e, f);
  for (int i = 0; i < n; i++) memset(*f[i].p, 0, f[i].n * sizeof(double));
  init_world(e);
  memcpy(e->qpos, m->qpos_reset, m->nq * sizeof(double));
  e->time = 0; e->step_counter = 0; e->reset_next = 0; e->ncon = e->nefc = 0;
  e->episode_steps = m->max_episode_steps;
  if (m->task_type == FLYB_TASK_WALK_IMITATION && e->has_clip) {
    int snippet = e->clip.T - m->future_steps - 1;          /* walk_imitation.py:101-102 */
    if (snippet < e->episode_steps) e->episode_steps = snippet;
    const double* r = e->clip.qpos;                          /* walk_imitation.py:119: full initial qpos */
    memcpy(e->qpos + m->root_qposadr, r, 7 * sizeof(double));
    for (int i = 0; i < m->n_mocap_joints; i++) e->qpos[m->jnt_qposadr[m->mocap_joint[i]]] = r[7 + i];
    set_ghost(e, 0, 0);
  }
  if (m->task_type == FLYB_TASK_FLIGHT_IMITATION && e->has_clip) {
    /* flight_imitation.py:106-108: traj_timesteps = min(T, round(time_limit / control_dt)) - (future_steps + 1) */
    e->episode_steps = (e->clip.T < m->max_episode_steps ? e->clip.T : m->max_episode_steps) - (m->future_steps + 1);
    set_ghost(e, 0, 0);                                                   /* :131-132 */
    memcpy(e->qpos + m->root_qposadr, e->clip.qpos, 7 * sizeof(double));  /* :138 walker.set_pose */
    if (e->has_wbpg) {                                                    /* :135-143 */
      double ang[6], vel[6];
      wbpg_reset(e, e->init_phase, ang, vel);
      for (int k = 0; k < 6; k++) {
        int j = m->wing_joint[k];
        e->qpos[m->jnt_qposadr[j]] = ang[k];
        e->qvel[m->jnt_dofadr[j]] = vel[k];
      }
    }
    memcpy(e->qvel + m->root_dofadr, e->clip.qvel, 3 * sizeof(double));   /* :145-147 linear velocity only */
  }
  if (m->task_type == FLYB_TASK_VISION_FLIGHT) {   /* vision_flight.py:114-146 */
    /* init_pos ranges are single points in the reference's configuration (init_pos_x_range (-5, -5), y (0, 0)) */
    const double x = m->vision_ranges[4], y = m->vision_ranges[6];
    double* q = e->qpos + m->root_qposadr;
    q[0] = x; q[1] = y; q[2] = hfield_height_at(e, x, y) + e->target_height;
    memcpy(q + 3, m->vision_init_quat, 4 * sizeof(double));
    if (e->has_wbpg) {                               /* :125-126,134: wbpg.reset(initial_phase) -> wing angles only */
      double ang[6], vel[6];
      wbpg_reset(e, e->init_phase, ang, vel);
      for (int k = 0; k < 6; k++) e->qpos[m->jnt_qposadr[m->wing_joint[k]]] = ang[k];
    }
    /* initialize_qvel is False (tasks/base.py:51; fly_envs.py:191-200 does not set it): the fly starts at rest */
  }
  /* dm_control's Physics.after_reset(), reached when composer's _reset_attempt leaves reset_context(), runs mj_forward with
     actuation disabled ("we don't yet have meaningful control inputs"): the affine position servos (biasprm 0 -kp) would
     otherwise pull at the initial pose, and the FIRST observation's accelerometer / force / touch would see it */
  {
    const int keep = e->disable_actuation;
    e->disable_actuation = 1;
    flyo_forward(e);
    e->disable_actuation = keep;
  }
  memcpy(e->qacc_warmstart, e->qacc, m->nv * sizeof(double));   /* mj_fwdConstraint leaves its result as the next warm start */
  accumulate_sensors(e);
  /* composer observation buffers: first sample then zero padding unless pad_with_initial_value */
  write_obs(e, obs, m->pad_first_obs ? 1.0 : 1.0 / m->n_substeps);
  return 0;
}

int flyo_env_step(FlyoEnv* e, const double* action_in, double* obs, double* reward, double* discount) {
  const FlybModel* m = e->m;
  if (e->reset_next) { *reward = 0; *discount = 1; return flyo_env_reset(e, obs); }
  double action[256];
  for (int i = 0; i < m->action_dim; i++) action[i] = isnan(action_in[i]) ? 0.0 : action_in[i];  /* walk_imitation.py:147 */
  int step = e->step_counter;                      /* == round(time / control_dt) before stepping */
  if (m->task_type == FLYB_TASK_WALK_IMITATION) set_ghost(e, step, 1);
  if (m->task_type == FLYB_TASK_FLIGHT_IMITATION || m->task_type == FLYB_TASK_VISION_FLIGHT) {  /* flight_imitation.py:149-167, vision_flight.py:148-155 */
    if (e->has_wbpg && m->user_action >= 0) {
      double ang[6];
      /* float64, as numpy 1.26.4 (the reference's pin) evaluates base * (1 + rel * act) for a float32 act */
      wbpg_step(e, e->wb.base_freq * (1 + e->wb.rel_range * (double)(float)action[m->user_action]), ang);
      for (int k = 0; k < 6; k++)   /* position target -> force command */
        action[m->wing_action[k]] += ang[k] - e->qpos[m->jnt_qposadr[m->wing_joint[k]]];
    }
    set_ghost(e, step, 1);
  }
  e->step_counter++;                               /* tasks/base.py:192 */
  memcpy(e->prev_action, action, m->action_dim * sizeof(double));
  for (int a = 0; a < m->nu; a++) e->ctrl[a] = m->ctrl_action[a] >= 0 ? action[m->ctrl_action[a]] : 0.0;  /* fruitfly.py:480-492 */
  memset(e->sens_sum, 0, m->nsensordata * sizeof(double));
  for (int s = 0; s < m->n_substeps; s++) { flyo_step(e); accumulate_sensors(e); }
  step = e->step_counter;
  /* termination: tasks/base.py:214-217, walk_imitation.py:165-191, walk_on_ball.py:75-79 */
  double qn = 0;
  for (int d = 0; d < m->nv; d++) qn += e->qacc[d] * e->qacc[d];
  int fatal = sqrt(qn) > m->terminal_qacc, reached_end = 0;
  if (m->task_type == FLYB_TASK_WALK_IMITATION || m->task_type == FLYB_TASK_WALK_ON_BALL) {  /* not flight, not template_task.py:78-83 */
    double lv[6];
    int site = m->thorax_site;
    local_velocity(e, lv, m->site_bodyid[site], e->site_xpos + 3 * site, e->site_xmat + 9 * site);
    if (norm3(lv + 3) > m->terminal_linvel || norm3(lv) > m->terminal_angvel) fatal = 1;
  }
  write_obs(e, obs, 1.0 / m->n_substeps);
  double r = 1.0;
  if (m->task_type == FLYB_TASK_WALK_IMITATION) {
    reached_end = step == e->episode_steps;
    if (e->has_clip) {
      const double* d = obs + m->obs_offset[FLYB_OBS_REF_DISPLACEMENT];
      if (norm3(d) > m->terminal_com_dist) fatal = 1;
    }
    r = walk_imitation_reward(e, step);
  } else if (m->task_type == FLYB_TASK_FLIGHT_IMITATION) {
    reached_end = step == e->episode_steps;          /* flight_imitation.py:198-209 */
    if (e->site_xpos[3 * m->thorax_site + 2] < m->terminal_height) fatal = 1;
    if (e->has_clip) {
      const double* d = obs + m->obs_offset[FLYB_OBS_REF_DISPLACEMENT];
      if (norm3(d) > m->terminal_com_dist) fatal = 1;
      r = flight_imitation_reward(e, obs);
    }
  } else if (m->task_type == FLYB_TASK_WALK_ON_BALL) {   /* walk_on_ball.py:61-73: tolerance(linear, margin 6) */
    const double tgt[3] = {0.0, -5.0, 0.0};
    for (int k = 0; k < 3; k++) {
      double x = fabs(e->qvel[m->ball_dofadr + k] - tgt[k]) / 6.0;
      r *= x < 1 ? 1 - x : 0.0;
    }
  } else if (m->task_type == FLYB_TASK_VISION_FLIGHT) {  /* vision_flight.py:157-233: floor contacts are fatal */
    if (vision_floor_contact(e)) fatal = 1;
    r = vision_reward(e);
  }
  *reward = r;
  int terminate = fatal || reached_end;
  *discount = (m->task_type == FLYB_TASK_WALK_ON_BALL || m->task_type == FLYB_TASK_VISION_FLIGHT) ? (terminate ? 0.0 : 1.0)
                                                                                                  : ((terminate && !reached_end) ? 0.0 : 1.0);
  if (terminate || step >= m->time_limit_steps) { e->reset_next = 1; return 2; }
  return 1;
}
