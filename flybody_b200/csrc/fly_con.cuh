/* fly_con.cuh - constraint rows (m6), transmissions (m3) and actuation (m9) for one environment. */
#ifndef FLY_CON_CUH_
#define FLY_CON_CUH_
#include "fly_col.cuh"

enum { FLY_CT_LIMIT = 0, FLY_CT_CONTACT1 = 1, FLY_CT_ELLIPTIC = 2 };
enum { FLY_ST_SATISFIED = 0, FLY_ST_QUADRATIC = 1, FLY_ST_CONE = 2 };
/* efc_xid[r]: -2 joint-limit row (single entry), -1 single-chain contact row, >= 0 index of the second chain */

template <class T> FLY_INL bool fly_is_anc(const DModel<T>& m, int d, int e) { return d <= e && e <= m.hsublast[d]; }

/* translational Jacobian of a point fixed to body b, projected on dir, written along b's dof chain */
template <class T>
FLY_INL void fly_chain_jac(const DModel<T>& m, const Ws<T>& w, int b, const T* point, const T* dir, T sign, T* out,
                           bool accumulate, int offset) {
  const int s = m.body_lastdof[b], a = m.madr[s], nk = m.body_kchain[b];
  const T* c0 = w.scom + 3 * m.body_rootid[b];
  const T off[3] = {point[0] - c0[0], point[1] - c0[1], point[2] - c0[2]};
  for (int k = 0; k < nk; k++) {
    const T* cd = w.cdof + 6 * m.mcol[a + k];
    T t[3];
    cross3(t, cd, off);
    const T v = sign * (dir[0] * (cd[3] + t[0]) + dir[1] * (cd[4] + t[1]) + dir[2] * (cd[5] + t[2]));
    if (accumulate) out[offset + k] += v; else out[offset + k] = v;
  }
}

template <class T>
FLY_INL T fly_impedance(const T* solimp, T pos, T margin) {
  T d0 = solimp[0], dw = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  d0 = fmin_(T(0.9999), fmax_(T(0.0001), d0)); dw = fmin_(T(0.9999), fmax_(T(0.0001), dw));
  if (power < 1) power = 1;
  mid = fmin_(T(0.9999), fmax_(T(0.0001), mid));
  if (d0 == dw || width <= minval<T>()) return T(0.5) * (d0 + dw);
  T x = fabs_((pos - margin) / width), y;
  if (x >= 1) return dw;
  if (x == 0) return d0;
  if (power == 1) y = x;
  else if (power == 2) y = x <= mid ? x * x / mid : 1 - (1 - x) * (1 - x) / (1 - mid);
  else if (x <= mid) y = pow(x, power) / pow(mid, power - 1);
  else y = 1 - pow(1 - x, power) / pow(1 - mid, power - 1);
  return d0 + y * (dw - d0);
}

/* K, B, impedance and regulariser of row r (refsafe applied) */
template <class T>
FLY_INL void fly_finish_row(const DModel<T>& m, const Ws<T>& w, int r, const T* solref_in, const T* solimp) {
  T sr0 = solref_in[0], sr1 = solref_in[1];
  if (sr0 > 0 && sr0 < 2 * m.timestep) sr0 = 2 * m.timestep;
  const T imp = fly_impedance(solimp, w.efc_pos[r], w.efc_margin[r]);
  const T dmax = fmin_(T(0.9999), fmax_(T(0.0001), solimp[1]));
  T K, B;
  if (sr0 > 0) {
    K = T(1) / fmax_(minval<T>(), dmax * dmax * sr0 * sr0 * sr1 * sr1);
    B = T(2) / fmax_(minval<T>(), dmax * sr0);
  } else { K = -sr0 / fmax_(minval<T>(), dmax * dmax); B = -sr1 / fmax_(minval<T>(), dmax); }
  w.efc_K0[r] = K; w.efc_B[r] = B; w.efc_imp[r] = imp;
  w.efc_R[r] = fmax_(minval<T>(), (1 - imp) * w.efc_diag[r] / imp);
}

template <class T>
FLY_FN void fly_make_constraint(const DModel<T>& m, const Ws<T>& w) {
  const int ncon = w.counts[FLY_C_NCON];
  /* joint limits: item 2j + side, lower side first; ordered compaction keeps the oracle's row order */
  int nlimit = 0;
  if (!(m.flags & FLY_OPT_NOLIMIT)) {
    nlimit = warp_compact(2 * FLY_DIM(m, njnt), [&](int it) -> bool {
      const int j = it >> 1, side = (it & 1) ? 1 : -1;
      if (!m.jnt_limited[j] || m.jnt_type[j] != FLYB_JNT_HINGE) return false;
      return side * (m.jnt_range[2 * j + (it & 1)] - w.qpos[m.jnt_qposadr[j]]) < 0;
    }, w.lim_item, FLY_MAXEFC / 2);
    if (nlimit > FLY_MAXEFC / 2) { nlimit = FLY_MAXEFC / 2; FLY_LANE0 w.counts[FLY_C_OVERFLOW] |= 4; }
  }
  FLY_SYNC();
  FLY_FOR(r, nlimit) {
    const int it = w.lim_item[r], j = it >> 1, side = (it & 1) ? 1 : -1, d = m.jnt_dofadr[j];
    T* J = w.efc_J + r * FLY_MAXCHAIN;
    const int len = m.hdepth[d] + 1;
    for (int k = 0; k < len; k++) J[k] = 0;
    J[0] = T(-side);
    w.efc_start[r] = d; w.efc_xid[r] = -2;
    w.efc_pos[r] = side * (m.jnt_range[2 * j + (it & 1)] - w.qpos[m.jnt_qposadr[j]]);
    w.efc_margin[r] = 0;
    w.efc_diag[r] = m.dof_invweight0[d];
    fly_finish_row(m, w, r, m.jnt_solref + 2 * j, m.jnt_solimp + 5 * j);
    w.grp_type[r] = FLY_CT_LIMIT; w.grp_row[r] = r; w.grp_con[r] = j; w.efc_g[r] = r;
  }
  /* contact rows: addresses are assigned in contact order by one lane */
  FLY_LANE0 {
    int r = nlimit, ngrp = nlimit, nx = 0;
    for (int ci = 0; ci < ncon; ci++) {
      w.con_efc[ci] = -1;
      if (w.con_excl[ci]) continue;
      const int dim = w.con_dim[ci];
      if (r + dim > FLY_MAXEFC || ngrp >= FLY_MAXGRP) { w.counts[FLY_C_OVERFLOW] |= 8; break; }
      const int pg = m.pair_geom[w.con_pair[ci]];
      const int b1 = m.geom_bodyid[pg & 255], b2 = m.geom_bodyid[(pg >> 8) & 255];
      const int s1 = b1 > 0 ? m.body_lastdof[b1] : -1, s2 = b2 > 0 ? m.body_lastdof[b2] : -1;
      const bool cross = s1 >= 0 && s2 >= 0 && !fly_is_anc(m, s1, s2) && !fly_is_anc(m, s2, s1);
      if (cross && nx + dim > FLY_MAXCROSS) { w.counts[FLY_C_OVERFLOW] |= 16; continue; }
      w.con_efc[ci] = r;
      w.grp_type[ngrp] = dim == 1 ? FLY_CT_CONTACT1 : FLY_CT_ELLIPTIC;
      w.grp_row[ngrp] = r; w.grp_con[ngrp] = ci;
      for (int k = 0; k < dim; k++) {
        w.efc_g[r + k] = ngrp;
        w.efc_xid[r + k] = cross ? nx : -1;
        if (cross) { w.x_row[nx] = r + k; nx++; }
      }
      ngrp++;
      r += dim;
    }
    w.counts[FLY_C_NEFC] = r; w.counts[FLY_C_NLIMIT] = nlimit; w.counts[FLY_C_NGRP] = ngrp;
    w.counts[FLY_C_NCROSSROW] = nx;
  }
  FLY_SYNC();
  const int ngrp = w.counts[FLY_C_NGRP];
  /* Jacobian chains: one (contact, row) per lane */
  FLY_FOR(it, (ngrp - nlimit) * 3) {
    const int g = nlimit + it / 3, k = it % 3, ci = w.grp_con[g], dim = w.con_dim[ci];
    if (k < dim) {
      const int r = w.grp_row[g] + k;
      const int pg = m.pair_geom[w.con_pair[ci]];
      const int b1 = m.geom_bodyid[pg & 255], b2 = m.geom_bodyid[(pg >> 8) & 255];
      const T *pos = w.con_pos + 3 * ci, *dir = w.con_frame + 9 * ci + 3 * k;
      T* J = w.efc_J + r * FLY_MAXCHAIN;
      const int s1 = b1 > 0 ? m.body_lastdof[b1] : -1, s2 = b2 > 0 ? m.body_lastdof[b2] : -1;
      const int xid = w.efc_xid[r];
      if (xid >= 0) {           /* two branches: chain of b2 (+) and chain of b1 (-) kept apart */
        T* JB = w.x_J + xid * FLY_MAXCHAIN;
        for (int q = 0; q <= m.hdepth[s2]; q++) J[q] = 0;
        for (int q = 0; q <= m.hdepth[s1]; q++) JB[q] = 0;
        fly_chain_jac(m, w, b2, pos, dir, T(1), J, false, 0);
        fly_chain_jac(m, w, b1, pos, dir, T(-1), JB, false, 0);
        w.efc_start[r] = s2; w.x_start[xid] = s1;
      } else {
        /* single chain: the deeper body's chain contains the other one as its root-side suffix */
        int sa, sb, ba, bb; T sga;
        if (s2 >= 0 && (s1 < 0 || fly_is_anc(m, s1, s2))) { sa = s2; ba = b2; sga = 1; sb = s1; bb = b1; }
        else { sa = s1; ba = b1; sga = -1; sb = s2; bb = b2; }
        for (int q = 0; q <= m.hdepth[sa]; q++) J[q] = 0;
        fly_chain_jac(m, w, ba, pos, dir, sga, J, false, 0);
        if (sb >= 0) fly_chain_jac(m, w, bb, pos, dir, -sga, J, true, m.hdepth[sa] - m.hdepth[sb]);
        w.efc_start[r] = sa;
      }
      w.efc_pos[r] = k == 0 ? w.con_dist[ci] : T(0);
      w.efc_margin[r] = k == 0 ? w.con_incl[ci] : T(0);
      w.efc_diag[r] = m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2];
    }
  }
  FLY_SYNC();
  FLY_FOR(gg, ngrp - nlimit) {
    const int g = nlimit + gg, ci = w.grp_con[g], dim = w.con_dim[ci], r = w.grp_row[g], p = w.con_pair[ci];
    fly_finish_row(m, w, r, m.pair_solref + 2 * p, m.pair_solimp + 5 * p);
    if (dim > 1) {
      /* elliptic cone: R1 = R0 / impratio, Rj = R1 mu1^2 / muj^2 (condim 3: equal), mu = mu1 sqrt(R1 / R0) */
      w.efc_R[r + 1] = w.efc_R[r] / fmax_(minval<T>(), m.impratio);
      for (int k = 2; k < dim; k++) w.efc_R[r + k] = w.efc_R[r + 1];
      w.con_mu[ci] = w.con_fric[ci] * sqrt(w.efc_R[r + 1] / w.efc_R[r]);
      for (int k = 1; k < dim; k++) { w.efc_K0[r + k] = 0; w.efc_B[r + k] = w.efc_B[r]; w.efc_imp[r + k] = w.efc_imp[r]; }
    }
  }
  FLY_SYNC();
  FLY_FOR(r, w.counts[FLY_C_NEFC]) w.efc_D[r] = T(1) / w.efc_R[r];
  FLY_SYNC();
}

/* J_r . v */
template <class T>
FLY_INL T fly_row_dot(const DModel<T>& m, const Ws<T>& w, int r, const T* v) {
  const int s = w.efc_start[r], xid = w.efc_xid[r];
  const T* J = w.efc_J + r * FLY_MAXCHAIN;
  if (xid == -2) return J[0] * v[s];
  const int a = m.madr[s], len = m.hdepth[s] + 1;
  T sum = 0;
  for (int k = 0; k < len; k++) sum += J[k] * v[m.mcol[a + k]];
  if (xid >= 0) {
    const int s1 = w.x_start[xid], a1 = m.madr[s1], len1 = m.hdepth[s1] + 1;
    const T* JB = w.x_J + xid * FLY_MAXCHAIN;
    for (int k = 0; k < len1; k++) sum += JB[k] * v[m.mcol[a1 + k]];
  }
  return sum;
}

/* out <- J v over all rows */
template <class T>
FLY_INL void fly_j_mul(const DModel<T>& m, const Ws<T>& w, T* out, const T* v) {
  FLY_FOR(r, w.counts[FLY_C_NEFC]) out[r] = fly_row_dot(m, w, r, v);
  FLY_SYNC();
}

/* out <- J^T f : limit rows lane-parallel (distinct dofs), contact rows one at a time with the chain over lanes */
template <class T>
FLY_FN void fly_jt_mul(const DModel<T>& m, const Ws<T>& w, T* out, const T* f) {
  const int nefc = w.counts[FLY_C_NEFC], nlimit = w.counts[FLY_C_NLIMIT];
  FLY_FOR(d, FLY_DIM(m, nv)) out[d] = 0;
  FLY_SYNC();
  FLY_FOR(r, nlimit) out[w.efc_start[r]] += w.efc_J[r * FLY_MAXCHAIN] * f[r];
  FLY_SYNC();
  for (int r = nlimit; r < nefc; r++) {
    const T fr = f[r];
    if (fr == 0) continue;
    const int s = w.efc_start[r], a = m.madr[s], len = m.hdepth[s] + 1;
    FLY_FOR(k, len) out[m.mcol[a + k]] += w.efc_J[r * FLY_MAXCHAIN + k] * fr;
    FLY_SYNC();
    const int xid = w.efc_xid[r];
    if (xid >= 0) {
      const int s1 = w.x_start[xid], a1 = m.madr[s1], len1 = m.hdepth[s1] + 1;
      FLY_FOR(k, len1) out[m.mcol[a1 + k]] += w.x_J[xid * FLY_MAXCHAIN + k] * fr;
      FLY_SYNC();
    }
  }
}

/* ---------------------------------------------------------------- m3 + m9 transmission and actuation */
template <class T>
FLY_FN void fly_actuation(const DModel<T>& m, const Ws<T>& w, bool actuate) {
  FLY_FOR(a, FLY_DIM(m, nu)) {
    const int id = m.act_trnid[a], trn = m.act_trntype[a];
    const T gear = m.act_gear[a];
    T len = 0, vel = 0;
    if (trn == FLYB_TRN_JOINT) { len = gear * w.qpos[m.jnt_qposadr[id]]; vel = gear * w.qvel[m.jnt_dofadr[id]]; }
    else if (trn == FLYB_TRN_TENDON) {
      for (int q = m.tendon_adr[id]; q < m.tendon_adr[id] + m.tendon_num[id]; q++) {
        len += m.wrap_coef[q] * w.qpos[m.wrap_qposadr[q]];
        vel += m.wrap_coef[q] * w.qvel[m.wrap_dofadr[q]];
      }
      len *= gear; vel *= gear;
    }
    w.act_length[a] = len;
    T ctrl = w.ctrl[a];
    if (m.act_ctrllimited[a]) ctrl = fmin_(m.act_ctrlrange[2 * a + 1], fmax_(m.act_ctrlrange[2 * a], ctrl));
    T input = ctrl;
    const int aa = m.act_actadr[a];
    if (aa >= 0) {
      w.act_dot[aa] = (ctrl - w.act[aa]) / fmax_(minval<T>(), m.act_dynprm[a]);
      input = w.act[aa];
    }
    T f = m.act_gainprm[a] * input + m.act_biasprm[3 * a] + m.act_biasprm[3 * a + 1] * len + m.act_biasprm[3 * a + 2] * vel;
    if (m.act_forcelimited[a]) f = fmin_(m.act_forcerange[2 * a + 1], fmax_(m.act_forcerange[2 * a], f));
    if ((m.flags & FLY_OPT_NOACTUATION) || !actuate) f = 0;   /* mjDSBL_ACTUATION: the forward pass of a reset */
    w.act_force[a] = f;
  }
  FLY_FOR(d, FLY_DIM(m, nv)) w.qfrc_actuator[d] = 0;
  FLY_SYNC();
  FLY_FOR(a, FLY_DIM(m, nu))
    if (m.act_trntype[a] == FLYB_TRN_JOINT) w.qfrc_actuator[m.jnt_dofadr[m.act_trnid[a]]] += m.act_gear[a] * w.act_force[a];
  FLY_SYNC();
  const int ncon = w.counts[FLY_C_NCON];
  for (int oa = 0; oa < m.nact_other; oa++) {  /* tendon and adhesion actuators, in actuator order */
    const int a = m.act_other[oa];
    const int trn = m.act_trntype[a], id = m.act_trnid[a];
    const T f = w.act_force[a], gear = m.act_gear[a];
    if (trn == FLYB_TRN_TENDON) {
      const int q0 = m.tendon_adr[id];
      FLY_FOR(q, m.tendon_num[id]) w.qfrc_actuator[m.wrap_dofadr[q0 + q]] += gear * m.wrap_coef[q0 + q] * f;
      FLY_SYNC();
      continue;
    }
    /* adhesion: moment = -(mean contact-normal Jacobian over the body's contacts, gap band included) */
    int cnt = 0;
    FLY_FOR(ci, ncon) {
      const int pg = m.pair_geom[w.con_pair[ci]];
      if (m.geom_bodyid[pg & 255] == id || m.geom_bodyid[(pg >> 8) & 255] == id) cnt++;
    }
    cnt = warp_sum_i(cnt);
    if (!cnt || f == 0) continue;
    const T coef = -gear * f / T(cnt);
    for (int ci = 0; ci < ncon; ci++) {
      const int pg = m.pair_geom[w.con_pair[ci]];
      const int b1 = m.geom_bodyid[pg & 255], b2 = m.geom_bodyid[(pg >> 8) & 255];
      if (b1 != id && b2 != id) continue;
      const T *pos = w.con_pos + 3 * ci, *dir = w.con_frame + 9 * ci;
      for (int side = 0; side < 2; side++) {
        const int b = side ? b1 : b2;
        if (b <= 0) continue;
        const T sg = side ? -coef : coef;
        const int s = m.body_lastdof[b], ad = m.madr[s];
        const T* c0 = w.scom + 3 * m.body_rootid[b];
        const T off[3] = {pos[0] - c0[0], pos[1] - c0[1], pos[2] - c0[2]};
        FLY_FOR(k, m.body_kchain[b]) {
          const int d = m.mcol[ad + k];
          const T* cd = w.cdof + 6 * d;
          T t[3];
          cross3(t, cd, off);
          w.qfrc_actuator[d] += sg * (dir[0] * (cd[3] + t[0]) + dir[1] * (cd[4] + t[1]) + dir[2] * (cd[5] + t[2]));
        }
        FLY_SYNC();
      }
    }
  }
  FLY_SYNC();
}
#endif
