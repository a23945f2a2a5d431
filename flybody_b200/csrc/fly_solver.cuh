/* fly_solver.cuh - constraint solve (m11) for one environment: primal Newton with elliptic cones and exact
 * line search, followed by the noslip Gauss-Seidel pass.
 *
 * Same iteration as the float64 oracle (and as MuJoCo's Newton solver), different linear algebra: the
 * Hessian H = M + J^T D J is kept in the TREE-SPARSE layout of M.  A contact row whose support is one dof
 * chain (everything that touches the floor or the ball, and self-contacts along one limb) only adds to
 * entries M already has, so H factorises with the same fill-free L^T D L as M.  The few rows that couple
 * two branches (fly-on-fly contacts) enter through a small Woodbury correction.  The Hessian only steers the
 * search direction; cost and gradient are exact, so the optimum is the oracle's. */
#ifndef FLY_SOLVER_CUH_
#define FLY_SOLVER_CUH_
#include "fly_con.cuh"

/* forces, states, cone Hessians for the residual `jar`; returns the constraint cost */
template <class T>
FLY_INL T fly_update_constraint(const DModel<T>& m, const Ws<T>& w, const T* jar, bool write, bool cone_hess) {
  T cost = 0;
  FLY_FOR(g, w.counts[FLY_C_NGRP]) {
    const int r = w.grp_row[g];
    if (w.grp_type[g] != FLY_CT_ELLIPTIC) {
      const T x = jar[r], D = w.efc_D[r];
      if (x < 0) { cost += T(0.5) * D * x * x; if (write) { w.efc_force[r] = -D * x; w.efc_state[r] = FLY_ST_QUADRATIC; } }
      else if (write) { w.efc_force[r] = 0; w.efc_state[r] = FLY_ST_SATISFIED; }
      continue;
    }
    const int ci = w.grp_con[g];
    const T mu = w.con_mu[ci], f0 = w.con_fric[ci];
    const T fr[3] = {mu, f0, f0};
    T U[3], T2 = 0;
    for (int k = 0; k < 3; k++) U[k] = jar[r + k] * fr[k];
    T2 = U[1] * U[1] + U[2] * U[2];
    const T N = U[0], Tn = sqrt(T2);
    if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
      if (write) for (int k = 0; k < 3; k++) { w.efc_force[r + k] = 0; w.efc_state[r + k] = FLY_ST_SATISFIED; }
    } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
      for (int k = 0; k < 3; k++) {
        const T D = w.efc_D[r + k], x = jar[r + k];
        cost += T(0.5) * D * x * x;
        if (write) { w.efc_force[r + k] = -D * x; w.efc_state[r + k] = FLY_ST_QUADRATIC; }
      }
    } else {
      const T Dm = w.efc_D[r] / fmax_(minval<T>(), mu * mu * (1 + mu * mu));
      const T NT = N - mu * Tn;
      cost += T(0.5) * Dm * NT * NT;
      if (write) {
        w.efc_force[r] = -Dm * NT * mu;
        for (int k = 1; k < 3; k++) w.efc_force[r + k] = Dm * NT * mu / Tn * U[k] * fr[k];
        for (int k = 0; k < 3; k++) w.efc_state[r + k] = FLY_ST_CONE;
      }
      if (cone_hess) {
        T* h = w.con_hc + 9 * ci;
        T Hu[9];
        Hu[0] = 1;
        for (int i = 1; i < 3; i++) Hu[i] = Hu[3 * i] = -mu * U[i] / Tn;
        for (int i = 1; i < 3; i++)
          for (int j = 1; j < 3; j++)
            Hu[3 * i + j] = mu * N / (Tn * Tn * Tn) * U[i] * U[j] + (i == j ? mu * mu - mu * N / Tn : T(0));
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) h[3 * i + j] = Dm * fr[i] * fr[j] * Hu[3 * i + j];
      }
    }
  }
  cost = warp_sum(cost);
  FLY_SYNC();
  return cost;
}

/* 1/2 (Ma - qfrc_smooth) . (qacc - qacc_smooth) */
template <class T>
FLY_INL T fly_gauss(const DModel<T>& m, const Ws<T>& w) {
  T s = 0;
  FLY_FOR(d, m.nv) s += T(0.5) * (w.sMa[d] - w.qfrc_smooth[d]) * (w.qacc[d] - w.qacc_smooth[d]);
  return warp_sum(s);
}

/* phi(alpha) along the search direction: value and first two derivatives */
template <class T>
FLY_INL void fly_ls_eval(const DModel<T>& m, const Ws<T>& w, T alpha, const T* quad, T* out) {
  T c = 0, d1 = 0, d2 = 0;
  FLY_FOR(g, w.counts[FLY_C_NGRP]) {
    const int r = w.grp_row[g];
    if (w.grp_type[g] != FLY_CT_ELLIPTIC) {
      const T jv = w.efc_jv[r], x = w.efc_jar[r] + alpha * jv;
      if (x < 0) { const T D = w.efc_D[r]; c += T(0.5) * D * x * x; d1 += D * x * jv; d2 += D * jv * jv; }
      continue;
    }
    const int ci = w.grp_con[g];
    const T mu = w.con_mu[ci], f0 = w.con_fric[ci];
    const T fr[3] = {mu, f0, f0};
    T x[3], v[3];
    for (int k = 0; k < 3; k++) { x[k] = (w.efc_jar[r + k] + alpha * w.efc_jv[r + k]) * fr[k]; v[k] = w.efc_jv[r + k] * fr[k]; }
    const T T2 = x[1] * x[1] + x[2] * x[2], TV = x[1] * v[1] + x[2] * v[2], VV = v[1] * v[1] + v[2] * v[2];
    const T N = x[0], Tn = sqrt(T2);
    if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
    } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
      for (int k = 0; k < 3; k++) {
        const T D = w.efc_D[r + k], jv = w.efc_jv[r + k], xx = w.efc_jar[r + k] + alpha * jv;
        c += T(0.5) * D * xx * xx; d1 += D * xx * jv; d2 += D * jv * jv;
      }
    } else {
      const T Dm = w.efc_D[r] / fmax_(minval<T>(), mu * mu * (1 + mu * mu));
      const T NT = N - mu * Tn, N1 = v[0], T1 = TV / Tn, Tdd = VV / Tn - T1 * T1 / Tn;
      const T g1 = N1 - mu * T1;
      c += T(0.5) * Dm * NT * NT; d1 += Dm * NT * g1; d2 += Dm * (g1 * g1 - NT * mu * Tdd);
    }
  }
  c = warp_sum(c); d1 = warp_sum(d1); d2 = warp_sum(d2);
  out[0] = c + quad[0] + alpha * quad[1] + alpha * alpha * quad[2];
  out[1] = d1 + quad[1] + 2 * alpha * quad[2];
  out[2] = d2 + 2 * quad[2];
}

template <class T>
FLY_FN T fly_line_search(const DModel<T>& m, const Ws<T>& w, T gauss) {
  T snorm = 0, q1 = 0, q2 = 0;
  FLY_FOR(d, m.nv) {
    const T s = w.ssearch[d];
    snorm += s * s;
    q1 += s * (w.sMa[d] - w.qfrc_smooth[d]);
    q2 += T(0.5) * s * w.sMv[d];
  }
  snorm = sqrt(warp_sum(snorm)); q1 = warp_sum(q1); q2 = warp_sum(q2);
  if (snorm < minval<T>()) return 0;
  const T quad[3] = {gauss, q1, q2};
  const T gtol = m.tolerance * m.ls_tolerance * snorm * m.meaninertia * T(m.nv > 1 ? m.nv : 1);
  T p[3], lo = 0, hi = -1, alpha = 0;
  fly_ls_eval(m, w, T(0), quad, p);
  if (p[1] >= 0) return 0;
  for (int it = 0; it < m.opt_ls_iterations; it++) {
    const T step = (p[2] > minval<T>()) ? -p[1] / p[2] : T(0);
    T next = alpha + step;
    if (hi < 0) { if (!(next > lo)) next = alpha > 0 ? 2 * alpha : T(1); }
    else if (!(next > lo && next < hi)) next = T(0.5) * (lo + hi);
    alpha = next;
    fly_ls_eval(m, w, alpha, quad, p);
    if (fabs_(p[1]) < gtol) break;
    if (p[1] < 0) lo = alpha; else hi = alpha;
  }
  return alpha;
}

/* H <- M + J^T D J (+ cone blocks) for single-chain rows, in the tree-sparse layout; returns the number of
 * active two-chain rows, listed in x_z-adjacent scratch (w.ns_fr reused as the active list) */
template <class T>
FLY_FN int fly_hessian(const DModel<T>& m, const Ws<T>& w) {
  const int ngrp = w.counts[FLY_C_NGRP], nlimit = w.counts[FLY_C_NLIMIT];
  FLY_FOR(e, m.nM) w.qH[e] = w.qM[e];
  FLY_SYNC();
  FLY_FOR(r, nlimit) if (w.efc_state[r] == FLY_ST_QUADRATIC) w.qH[m.madr[w.efc_start[r]]] += w.efc_D[r];
  FLY_SYNC();
  int nact = 0;
  for (int g = nlimit; g < ngrp; g++) {
    const int r = w.grp_row[g], st = w.efc_state[r];
    if (st == FLY_ST_SATISFIED) continue;
    const int dim = w.grp_type[g] == FLY_CT_ELLIPTIC ? 3 : 1;
    if (w.efc_xid[r] >= 0) {   /* two-chain rows go to the Woodbury list (diagonal weights) */
      for (int k = 0; k < dim; k++) {
        const T wgt = st == FLY_ST_QUADRATIC ? w.efc_D[r + k] : w.con_hc[9 * w.grp_con[g] + 4 * k];
        if (nact < FLY_MAXCROSS && wgt > 0) {
          FLY_LANE0 { w.ns_fr[nact] = r + k; w.x_z[nact] = wgt; }
          nact++;
        }
      }
      continue;
    }
    const int s = w.efc_start[r], a = m.madr[s], len = m.hdepth[s] + 1;
    const T* J = w.efc_J + r * FLY_MAXCHAIN;
    const T* hc = w.con_hc + 9 * w.grp_con[g];
    FLY_FOR(e, len * (len + 1) / 2) {
      const int pa = m.tri_a[e], pb = m.tri_b[e];   /* pa <= pb: row = deeper dof, column = its ancestor */
      T v;
      if (dim == 1) v = w.efc_D[r] * J[pa] * J[pb];
      else if (st == FLY_ST_QUADRATIC)
        v = w.efc_D[r] * J[pa] * J[pb] + w.efc_D[r + 1] * J[FLY_MAXCHAIN + pa] * J[FLY_MAXCHAIN + pb]
            + w.efc_D[r + 2] * J[2 * FLY_MAXCHAIN + pa] * J[2 * FLY_MAXCHAIN + pb];
      else {
        v = 0;
        for (int k = 0; k < 3; k++) for (int l = 0; l < 3; l++) v += hc[3 * k + l] * J[k * FLY_MAXCHAIN + pa] * J[l * FLY_MAXCHAIN + pb];
      }
      w.qH[m.madr[m.mcol[a + pa]] + (pb - pa)] += v;
    }
    FLY_SYNC();
  }
  FLY_SYNC();
  return nact;
}

/* x <- H^-1 x with H = Ht + sum_c wc Jc^T Jc ; Ht factorised in w.qH ; nact two-chain rows in w.ns_fr / w.x_z */
template <class T>
FLY_FN void fly_cross_prepare(const DModel<T>& m, const Ws<T>& w, int nact) {
  for (int c = 0; c < nact; c++) {
    const int r = w.ns_fr[c], xid = w.efc_xid[r];
    T* y = w.x_dense + c * m.nv;
    FLY_FOR(d, m.nv) y[d] = 0;
    FLY_SYNC();
    { const int s = w.efc_start[r], a = m.madr[s]; FLY_FOR(k, m.hdepth[s] + 1) y[m.mcol[a + k]] += w.efc_J[r * FLY_MAXCHAIN + k]; }
    FLY_SYNC();
    { const int s = w.x_start[xid], a = m.madr[s]; FLY_FOR(k, m.hdepth[s] + 1) y[m.mcol[a + k]] += w.x_J[xid * FLY_MAXCHAIN + k]; }
    FLY_SYNC();
    fly_solve(m, w.qH, y);
  }
  FLY_FOR(it, nact * nact) {
    const int c = it / nact, c2 = it % nact;
    w.x_S[c * FLY_MAXCROSS + c2] = fly_row_dot(m, w, w.ns_fr[c], w.x_dense + c2 * m.nv) + (c == c2 ? T(1) / w.x_z[c] : T(0));
  }
  FLY_SYNC();
}
template <class T>
FLY_FN void fly_cross_apply(const DModel<T>& m, const Ws<T>& w, int nact, T* x) {
  T* z = w.x_z + FLY_MAXCROSS / 2 * 0;  /* weights stay in x_z; rhs/solution go to obs_tmp-free scratch below */
  T* rhs = w.efc_b;                      /* efc_b is free during the Newton phase */
  FLY_FOR(c, nact) rhs[c] = fly_row_dot(m, w, w.ns_fr[c], x);
  FLY_SYNC();
  FLY_LANE0 {   /* small dense solve S u = rhs, Gaussian elimination with partial pivoting on a copy */
    T* S = w.x_Y;  /* FLY_MAXCROSS * FLY_MAXCHAIN >= nact^2 scratch */
    for (int i = 0; i < nact; i++) for (int j = 0; j < nact; j++) S[i * nact + j] = w.x_S[i * FLY_MAXCROSS + j];
    for (int k = 0; k < nact; k++) {
      int piv = k; T best = fabs_(S[k * nact + k]);
      for (int i = k + 1; i < nact; i++) if (fabs_(S[i * nact + k]) > best) { best = fabs_(S[i * nact + k]); piv = i; }
      if (piv != k) {
        for (int j = 0; j < nact; j++) { T t = S[k * nact + j]; S[k * nact + j] = S[piv * nact + j]; S[piv * nact + j] = t; }
        T t = rhs[k]; rhs[k] = rhs[piv]; rhs[piv] = t;
      }
      const T dk = S[k * nact + k];
      if (fabs_(dk) < minval<T>()) continue;
      for (int i = k + 1; i < nact; i++) {
        const T f = S[i * nact + k] / dk;
        for (int j = k; j < nact; j++) S[i * nact + j] -= f * S[k * nact + j];
        rhs[i] -= f * rhs[k];
      }
    }
    for (int k = nact - 1; k >= 0; k--) {
      T s = rhs[k];
      for (int j = k + 1; j < nact; j++) s -= S[k * nact + j] * rhs[j];
      const T dk = S[k * nact + k];
      rhs[k] = fabs_(dk) < minval<T>() ? T(0) : s / dk;
    }
  }
  FLY_SYNC();
  FLY_FOR(d, m.nv) {
    T s = 0;
    for (int c = 0; c < nact; c++) s += rhs[c] * w.x_dense[c * m.nv + d];
    x[d] -= s;
  }
  FLY_SYNC();
  (void)z;
}

template <class T>
FLY_FN void fly_newton(const DModel<T>& m, const Ws<T>& w) {
  const int nv = m.nv, nefc = w.counts[FLY_C_NEFC];
  int niter = 0;
  if (nefc == 0) {
    FLY_FOR(d, nv) { w.qacc[d] = w.qacc_smooth[d]; w.qfrc_constraint[d] = 0; }
    FLY_LANE0 w.counts[FLY_C_NITER] = 0;
    FLY_SYNC();
    return;
  }
  const T scale = T(1) / (m.meaninertia * T(nv > 1 ? nv : 1));
  /* warm start: the better of qacc_warmstart and qacc_smooth */
  FLY_FOR(d, nv) w.qacc[d] = w.warm[d];
  FLY_SYNC();
  FLY_FOR(r, nefc) w.efc_jar[r] = fly_row_dot(m, w, r, w.qacc) - w.efc_aref[r];
  FLY_SYNC();
  fly_mul_m(m, w.qM, w.sMa, w.qacc);
  T cw = fly_update_constraint(m, w, w.efc_jar, false, false) + fly_gauss(m, w);
  FLY_FOR(r, nefc) w.efc_jv[r] = fly_row_dot(m, w, r, w.qacc_smooth) - w.efc_aref[r];
  FLY_SYNC();
  const T cs = fly_update_constraint(m, w, w.efc_jv, false, false);
  if ((m.flags & FLY_OPT_NOWARMSTART) || cs < cw) {
    FLY_FOR(d, nv) w.qacc[d] = w.qacc_smooth[d];
    FLY_FOR(r, nefc) w.efc_jar[r] = w.efc_jv[r];
    FLY_SYNC();
    fly_mul_m(m, w.qM, w.sMa, w.qacc);
  }
  T cost = fly_update_constraint(m, w, w.efc_jar, true, true);
  T gauss = fly_gauss(m, w);
  cost += gauss;
  for (int it = 0;; it++) {
    fly_jt_mul(m, w, w.qfrc_constraint, w.efc_force);
    T gn = 0;
    FLY_FOR(d, nv) { const T g = w.sMa[d] - w.qfrc_smooth[d] - w.qfrc_constraint[d]; w.sgrad[d] = g; gn += g * g; }
    gn = warp_sum(gn);
    FLY_SYNC();
    if (it > 0 && scale * sqrt(gn) < m.tolerance) break;
    if (it >= m.opt_iterations) break;
    const int nact = fly_hessian(m, w);
    fly_factor(m, w.qH);
    FLY_FOR(d, nv) w.ssearch[d] = w.sgrad[d];
    FLY_SYNC();
    fly_solve(m, w.qH, w.ssearch);
    if (nact) { fly_cross_prepare(m, w, nact); fly_cross_apply(m, w, nact, w.ssearch); }
    FLY_FOR(d, nv) w.ssearch[d] = -w.ssearch[d];
    FLY_SYNC();
    fly_mul_m(m, w.qM, w.sMv, w.ssearch);
    fly_j_mul(m, w, w.efc_jv, w.ssearch);
    const T alpha = fly_line_search(m, w, gauss);
    if (alpha == 0) break;
    FLY_FOR(d, nv) { w.qacc[d] += alpha * w.ssearch[d]; w.sMa[d] += alpha * w.sMv[d]; }
    FLY_FOR(r, nefc) w.efc_jar[r] += alpha * w.efc_jv[r];
    FLY_SYNC();
    const T old = cost;
    cost = fly_update_constraint(m, w, w.efc_jar, true, true);
    gauss = fly_gauss(m, w);
    cost += gauss;
    niter = it + 1;
    if (scale * (old - cost) < m.tolerance) {
      fly_jt_mul(m, w, w.qfrc_constraint, w.efc_force);
      break;
    }
  }
  FLY_LANE0 w.counts[FLY_C_NITER] = niter;
  FLY_SYNC();
}

/* ---------------------------------------------------------------- noslip */
template <class T>
FLY_INL int fly_qcqp2(T* res, const T* A, const T* b, const T* d, T r) {
  const T b1 = b[0] * d[0], b2 = b[1] * d[1];
  const T A11 = A[0] * d[0] * d[0], A22 = A[3] * d[1] * d[1], A12 = A[1] * d[0] * d[1];
  T la = 0, v1 = 0, v2 = 0;
  for (int it = 0; it < 20; it++) {
    const T det = (A11 + la) * (A22 + la) - A12 * A12;
    if (det < T(1e-10)) { res[0] = res[1] = 0; return 0; }
    const T P11 = (A22 + la) / det, P22 = (A11 + la) / det, P12 = -A12 / det;
    v1 = -P11 * b1 - P12 * b2; v2 = -P12 * b1 - P22 * b2;
    const T val = v1 * v1 + v2 * v2 - r * r;
    if (val < T(1e-10)) break;
    const T deriv = -2 * (P11 * v1 * v1 + 2 * P12 * v1 * v2 + P22 * v2 * v2);
    const T delta = -val / deriv;
    if (delta < T(1e-10)) break;
    la += delta;
  }
  res[0] = v1 * d[0]; res[1] = v2 * d[1];
  return la != 0;
}

/* y <- rows of (L^-T J^T) D^-1/2 restricted to the chain that starts at dof s (in place, chain-local) */
template <class T>
FLY_INL void fly_half_solve(const DModel<T>& m, const T* LD, int s, T* x) {
  const int a0 = m.madr[s], len = m.hdepth[s] + 1;
  for (int p = 0; p < len; p++) {
    const int a = m.madr[m.mcol[a0 + p]];
    const T xp = x[p];
    if (xp != 0) for (int q = p + 1; q < len; q++) x[q] -= LD[a + (q - p)] * xp;
    x[p] = xp / sqrt(LD[a]);
  }
}
template <class T>
FLY_INL T fly_part_dot(const DModel<T>& m, int sa, const T* ya, int sb, const T* yb) {
  const int nc = m.ncommon[sa * m.nv + sb];
  const T *pa = ya + (m.hdepth[sa] + 1 - nc), *pb = yb + (m.hdepth[sb] + 1 - nc);
  T s = 0;
  for (int t = 0; t < nc; t++) s += pa[t] * pb[t];
  return s;
}

template <class T>
FLY_FN void fly_noslip(const DModel<T>& m, const Ws<T>& w) {
  const int nefc = w.counts[FLY_C_NEFC], ngrp = w.counts[FLY_C_NGRP], nlimit = w.counts[FLY_C_NLIMIT];
  const int nx = w.counts[FLY_C_NCROSSROW];
  /* elliptic contacts, in contact order */
  FLY_LANE0 {
    int n3 = 0;
    for (int g = nlimit; g < ngrp; g++) if (w.grp_type[g] == FLY_CT_ELLIPTIC) { w.ns_fr[n3] = w.grp_row[g]; w.ns_con[n3] = w.grp_con[g]; n3++; }
    w.counts[FLY_C_NFR] = n3;
  }
  FLY_SYNC();
  const int n3 = w.counts[FLY_C_NFR];
  if (!n3) return;
  /* A = J M^-1 J^T, friction rows only, through chain-local half solves */
  FLY_FOR(it, nefc + nx) {
    if (it < nefc) {
      const int s = w.efc_start[it], len = m.hdepth[s] + 1;
      T* y = w.efc_Y + it * FLY_MAXCHAIN;
      for (int k = 0; k < len; k++) y[k] = w.efc_J[it * FLY_MAXCHAIN + k];
      fly_half_solve(m, w.qLD, s, y);
    } else {
      const int xi = it - nefc, s = w.x_start[xi], len = m.hdepth[s] + 1;
      T* y = w.x_Y + xi * FLY_MAXCHAIN;
      for (int k = 0; k < len; k++) y[k] = w.x_J[xi * FLY_MAXCHAIN + k];
      fly_half_solve(m, w.qLD, s, y);
    }
  }
  FLY_SYNC();
  FLY_FOR(it, 2 * n3 * nefc) {
    const int fi = it / nefc, c = it % nefc, rf = w.ns_fr[fi >> 1] + 1 + (fi & 1);
    const int sa = w.efc_start[rf], sb = w.efc_start[c], xa = w.efc_xid[rf], xb = w.efc_xid[c];
    const T *ya = w.efc_Y + rf * FLY_MAXCHAIN, *yb = w.efc_Y + c * FLY_MAXCHAIN;
    T s = fly_part_dot(m, sa, ya, sb, yb);
    if (xa >= 0) s += fly_part_dot(m, w.x_start[xa], w.x_Y + xa * FLY_MAXCHAIN, sb, yb);
    if (xb >= 0) s += fly_part_dot(m, sa, ya, w.x_start[xb], w.x_Y + xb * FLY_MAXCHAIN);
    if (xa >= 0 && xb >= 0) s += fly_part_dot(m, w.x_start[xa], w.x_Y + xa * FLY_MAXCHAIN, w.x_start[xb], w.x_Y + xb * FLY_MAXCHAIN);
    w.ns_A[it] = s;
  }
  FLY_FOR(r, nefc) w.efc_b[r] = fly_row_dot(m, w, r, w.qacc_smooth) - w.efc_aref[r];
  FLY_SYNC();
  const T scale = T(1) / (m.meaninertia * T(m.nv > 1 ? m.nv : 1));
  T* f = w.efc_force;
  for (int it = 0; it < m.opt_noslip_iterations; it++) {
    T improvement = 0;
    for (int i = 0; i < n3; i++) {
      const int r0 = w.ns_fr[i];
      const T *A0 = w.ns_A + (2 * i) * nefc, *A1 = A0 + nefc;
      T res[2] = {0, 0};
      FLY_FOR(k, nefc) { const T fk = f[k]; res[0] += A0[k] * fk; res[1] += A1[k] * fk; }
      res[0] = warp_sum(res[0]) + w.efc_b[r0 + 1];
      res[1] = warp_sum(res[1]) + w.efc_b[r0 + 2];
      const T old[2] = {f[r0 + 1], f[r0 + 2]};
      const T Ac[4] = {A0[r0 + 1], A0[r0 + 2], A1[r0 + 1], A1[r0 + 2]};
      const T fn = f[r0];
      T v[2];
      if (fn < minval<T>()) { v[0] = v[1] = 0; }
      else {
        T bc[2];
        for (int a = 0; a < 2; a++) bc[a] = res[a] - Ac[2 * a] * old[0] - Ac[2 * a + 1] * old[1];
        const T fric = w.con_fric[w.ns_con[i]];
        const T mu[2] = {fric, fric};
        const int active = fly_qcqp2(v, Ac, bc, mu, fn);
        if (active) {
          T s = v[0] * v[0] / (mu[0] * mu[0]) + v[1] * v[1] / (mu[1] * mu[1]);
          s = sqrt(fn * fn / fmax_(minval<T>(), s));
          v[0] *= s; v[1] *= s;
        }
      }
      FLY_SYNC();   /* every lane has read f before it changes */
      FLY_LANE0 { f[r0 + 1] = v[0]; f[r0 + 2] = v[1]; }
      FLY_SYNC();
      const T dl[2] = {v[0] - old[0], v[1] - old[1]};
      const T change = T(0.5) * (dl[0] * (Ac[0] * dl[0] + Ac[1] * dl[1]) + dl[1] * (Ac[2] * dl[0] + Ac[3] * dl[1]))
                       + dl[0] * res[0] + dl[1] * res[1];
      improvement -= change;
    }
    if (improvement * scale < m.noslip_tolerance) break;
  }
  fly_jt_mul(m, w, w.qfrc_constraint, f);
}

/* ---------------------------------------------------------------- m11 driver */
template <class T>
FLY_FN void fly_fwd_constraint(const DModel<T>& m, const Ws<T>& w) {
  const int nefc = w.counts[FLY_C_NEFC];
  FLY_FOR(r, nefc) {
    const T vel = fly_row_dot(m, w, r, w.qvel);
    w.efc_aref[r] = -w.efc_B[r] * vel - w.efc_K[r] * w.efc_imp[r] * (w.efc_pos[r] - w.efc_margin[r]);
  }
  FLY_SYNC();
  fly_newton(m, w);
  if (nefc && !(m.flags & FLY_OPT_NONOSLIP) && m.opt_noslip_iterations > 0) {
    fly_noslip(m, w);
    /* qacc = qacc_smooth + M^-1 J^T f */
    FLY_FOR(d, m.nv) w.stmp[d] = w.qfrc_constraint[d];
    FLY_SYNC();
    fly_solve(m, w.qLD, w.stmp);
    FLY_FOR(d, m.nv) w.qacc[d] = w.qacc_smooth[d] + w.stmp[d];
    FLY_SYNC();
  }
}
#endif
