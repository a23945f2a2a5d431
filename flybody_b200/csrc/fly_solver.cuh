/* fly_solver.cuh - constraint solve (m11) for one environment: primal Newton with elliptic cones and exact
 * line search, followed by the noslip Gauss-Seidel pass.
 *
 * Same Newton iteration as the float64 oracle (and as MuJoCo's Newton solver), but carried out in CONSTRAINT
 * SPACE.  With M = L^T D L, u = D^1/2 L qacc and y_r = D^-1/2 L^-T J_r^T (a chain-local half solve, so y_r is
 * as sparse as J_r) the problem reads  min 1/2 |u - u_s|^2 + sum_r s_r(y_r . u - aref_r):  the Hessian is
 * I + Y^T W Y, identity plus rank nefc, and its inverse needs only the small dense matrix I + W A with
 * A = Y Y^T = J M^-1 J^T (nefc x nefc, typically ~10-20).  Every iterate is kept as coordinates (theta, c) of
 * qacc - qacc_smooth = theta (qacc_warmstart - qacc_smooth) + M^-1 J^T c, so one Newton iteration costs two
 * mat-vecs with A, one small LU solve and the line search - no nv-sized factorisation.  M is factorised once
 * per physics step.  The search direction equals -H^-1 grad of the primal formulation exactly (push-through
 * identity); the only deliberate deviation is the norm used to scale the line-search tolerance (see below). */
#ifndef FLY_SOLVER_CUH_
#define FLY_SOLVER_CUH_
#include "fly_con.cuh"

/* forces, states, cone Hessians for the residual `jar`; returns the constraint cost */
typedef double S;  /* accumulator type of the constraint-space iterate (see FLY_WS_DBL) */

template <class T>
FLY_INL S fly_update_constraint(const DModel<T>& m, const Ws<T>& w, const S* jar, bool write, bool cone_hess) {
  S cost = 0;
  FLY_FOR(g, w.counts[FLY_C_NGRP]) {
    const int r = w.grp_row[g];
    if (w.grp_type[g] != FLY_CT_ELLIPTIC) {
      const S x = jar[r], D = w.efc_D[r];
      if (x < 0) { cost += S(0.5) * D * x * x; if (write) { w.efc_force[r] = (T)(-D * x); w.efc_state[r] = FLY_ST_QUADRATIC; } }
      else if (write) { w.efc_force[r] = 0; w.efc_state[r] = FLY_ST_SATISFIED; }
      continue;
    }
    const int ci = w.grp_con[g];
    const S mu = w.con_mu[ci], f0 = w.con_fric[ci];
    const S fr[3] = {mu, f0, f0};
    S U[3], T2 = 0;
    for (int k = 0; k < 3; k++) U[k] = jar[r + k] * fr[k];
    T2 = U[1] * U[1] + U[2] * U[2];
    const S N = U[0], Tn = sqrt(T2);
    if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
      if (write) for (int k = 0; k < 3; k++) { w.efc_force[r + k] = 0; w.efc_state[r + k] = FLY_ST_SATISFIED; }
    } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
      for (int k = 0; k < 3; k++) {
        const S D = w.efc_D[r + k], x = jar[r + k];
        cost += S(0.5) * D * x * x;
        if (write) { w.efc_force[r + k] = (T)(-D * x); w.efc_state[r + k] = FLY_ST_QUADRATIC; }
      }
    } else {
      const S Dm = w.efc_D[r] / fmax_(minval<S>(), mu * mu * (1 + mu * mu));
      const S NT = N - mu * Tn;
      cost += S(0.5) * Dm * NT * NT;
      if (write) {
        w.efc_force[r] = (T)(-Dm * NT * mu);
        for (int k = 1; k < 3; k++) w.efc_force[r + k] = (T)(Dm * NT * mu / Tn * U[k] * fr[k]);
        for (int k = 0; k < 3; k++) w.efc_state[r + k] = FLY_ST_CONE;
      }
      if (cone_hess) {
        /* Hessian block of the cone cost 1/2 Dm (N - mu T)^2 in the rows' own units: Dm F (g g^T + c p p^T) F with F = diag(fr),
           g = (1, -mu U1/T, -mu U2/T), p = (0, -U2/T, U1/T), c = mu (mu T - N) / T >= 0 in this zone - positive semi-definite of
           rank 2.  Its factor S (3 x 2, H = S S^T) is what the Newton step needs: columns sqrt(Dm) F g and sqrt(Dm c) F p. */
        T* h = w.con_hc + 9 * ci;
        const S sd = sqrt(Dm), sc = sqrt(Dm * fmax_(S(0), mu * (mu * Tn - N) / Tn));
        h[0] = (T)(sd * fr[0]); h[1] = (T)(-sd * fr[1] * mu * U[1] / Tn); h[2] = (T)(-sd * fr[2] * mu * U[2] / Tn);
        h[3] = 0; h[4] = (T)(-sc * fr[1] * U[2] / Tn); h[5] = (T)(sc * fr[2] * U[1] / Tn);
      }
    }
  }
  cost = warp_sum(cost);
  FLY_SYNC();
  return cost;
}

/* phi(alpha) along the search direction: value and first two derivatives */
template <class T>
FLY_INL void fly_ls_eval(const DModel<T>& m, const Ws<T>& w, S alpha, const S* quad, S* out) {
  S c = 0, d1 = 0, d2 = 0;
  FLY_FOR(g, w.counts[FLY_C_NGRP]) {
    const int r = w.grp_row[g];
    if (w.grp_type[g] != FLY_CT_ELLIPTIC) {
      const S jv = w.efc_jv[r], x = w.efc_jar[r] + alpha * jv;
      if (x < 0) { const S D = w.efc_D[r]; c += S(0.5) * D * x * x; d1 += D * x * jv; d2 += D * jv * jv; }
      continue;
    }
    const int ci = w.grp_con[g];
    const S mu = w.con_mu[ci], f0 = w.con_fric[ci];
    const S fr[3] = {mu, f0, f0};
    S x[3], v[3];
    for (int k = 0; k < 3; k++) { x[k] = (w.efc_jar[r + k] + alpha * w.efc_jv[r + k]) * fr[k]; v[k] = w.efc_jv[r + k] * fr[k]; }
    const S T2 = x[1] * x[1] + x[2] * x[2], TV = x[1] * v[1] + x[2] * v[2], VV = v[1] * v[1] + v[2] * v[2];
    const S N = x[0], Tn = sqrt(T2);
    if (N >= mu * Tn || (Tn <= 0 && N >= 0)) {
    } else if (mu * N + Tn <= 0 || (Tn <= 0 && N < 0)) {
      for (int k = 0; k < 3; k++) {
        const S D = w.efc_D[r + k], jv = w.efc_jv[r + k], xx = w.efc_jar[r + k] + alpha * jv;
        c += S(0.5) * D * xx * xx; d1 += D * xx * jv; d2 += D * jv * jv;
      }
    } else {
      const S Dm = w.efc_D[r] / fmax_(minval<S>(), mu * mu * (1 + mu * mu));
      const S NT = N - mu * Tn, N1 = v[0], T1 = TV / Tn, Tdd = VV / Tn - T1 * T1 / Tn;
      const S g1 = N1 - mu * T1;
      c += S(0.5) * Dm * NT * NT; d1 += Dm * NT * g1; d2 += Dm * (g1 * g1 - NT * mu * Tdd);
    }
  }
  c = warp_sum(c); d1 = warp_sum(d1); d2 = warp_sum(d2);
  out[0] = c + quad[0] + alpha * quad[1] + alpha * alpha * quad[2];
  out[1] = d1 + quad[1] + 2 * alpha * quad[2];
  out[2] = d2 + 2 * quad[2];
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
  const int nc = m.ncommon[sa * FLY_DIM(m, nv) + sb];
  const T *pa = ya + (m.hdepth[sa] + 1 - nc), *pb = yb + (m.hdepth[sb] + 1 - nc);
  T s = 0;
  for (int t = 0; t < nc; t++) s += pa[t] * pb[t];
  return s;
}


/* exact 1-D Newton line search along (efc_jar, efc_jv); quad = {gauss, q1, q2} of the quadratic part */
template <class T>
FLY_FN S fly_line_search(const DModel<T>& m, const Ws<T>& w, const S* quad, S gtol) {
  S p[3], lo = 0, hi = -1, alpha = 0;
  fly_ls_eval(m, w, S(0), quad, p);
  if (p[1] >= 0) return 0;
  for (int it = 0; it < m.opt_ls_iterations; it++) {
    const S step = (p[2] > minval<S>()) ? -p[1] / p[2] : S(0);
    S next = alpha + step;
    if (hi < 0) { if (!(next > lo)) next = alpha > 0 ? 2 * alpha : S(1); }
    else if (!(next > lo && next < hi)) next = S(0.5) * (lo + hi);
    alpha = next;
    fly_ls_eval(m, w, alpha, quad, p);
    if (fabs_(p[1]) < gtol) break;
    if (p[1] < 0) lo = alpha; else hi = alpha;
  }
  return alpha;
}

/* (value, index) arg-max over lanes; ties go to the lowest index so that every lane agrees with a serial scan */
template <class T>
FLY_INL void warp_argmax(T& v, int& idx) {
#ifdef FLY_GPU
#pragma unroll
  for (int o = FLY_NL / 2; o > 0; o >>= 1) {
    const T ov = __shfl_xor_sync(FLY_MASK(), v, o);
    const int oi = __shfl_xor_sync(FLY_MASK(), idx, o);
    if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
  }
#endif
}

/* dense n x n solve K z = rhs for a symmetric positive definite K (lower triangle given, row-major, destroyed; rhs <- z):
 * Cholesky K = L L^T without pivoting, the forward substitution folded into the elimination, then the back substitution
 * column by column.  Two barriers per column going down, one coming back; no pivot search, no shuffles.
 * scratch: 2 n numbers (1 / L_kk and the intermediate vector y). */
template <class T>
FLY_FN void fly_chol_solve(T* K, int n, T* rhs, T* scratch) {
  T *dinv = scratch, *y = scratch + n;
  for (int k = 0; k < n; k++) {
    FLY_SYNC();                       /* the trailing update of column k - 1 (or the caller's fill) is visible */
    const T inv = T(1) / sqrt(fmax_(K[k * n + k], minval<T>()));
    const T yk = rhs[k] * inv;
    const int nr = n - k - 1;
    FLY_FOR(i, nr + 1) {
      if (i == nr) { dinv[k] = inv; y[k] = yk; }
      else {
        const int r = k + 1 + i;
        const T l = K[r * n + k] * inv;
        K[r * n + k] = l;
        rhs[r] -= l * yk;
      }
    }
    FLY_SYNC();
    /* trailing block, lower triangle: rows across lane / 8, columns across lane % 8 */
    FLY_FOR_ROW(i, nr) {
      const int r = k + 1 + i;
      const T l = K[r * n + k];
      FLY_FOR_COL(j, k + 1, r + 1) K[r * n + j] -= l * K[j * n + k];
    }
  }
  FLY_SYNC();
  for (int k = n - 1; k >= 0; k--) {   /* L^T z = y: z_k = y_k / L_kk, then y_i -= L_ki z_k for i < k (row k is contiguous) */
    const T zk = y[k] * dinv[k];
    FLY_FOR(i, k + 1) {
      if (i == k) rhs[k] = zk;
      else y[i] -= K[k * n + i] * zk;
    }
    FLY_SYNC();
  }
}

/* Y = D^-1/2 L^-T J^T row by row (chain-local), then A = Y Y^T = J M^-1 J^T.  Needs qLD of this step. */
template <class T>
FLY_FN void fly_constraint_prepare(const DModel<T>& m, const Ws<T>& w) {
  const int nefc = w.counts[FLY_C_NEFC], nx = w.counts[FLY_C_NCROSSROW];
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
  FLY_FOR(it, nefc * nefc) {
    const int ra = it / nefc, rb = it % nefc;
    const int sa = w.efc_start[ra], sb = w.efc_start[rb], xa = w.efc_xid[ra], xb = w.efc_xid[rb];
    const T *ya = w.efc_Y + ra * FLY_MAXCHAIN, *yb = w.efc_Y + rb * FLY_MAXCHAIN;
    T s = fly_part_dot(m, sa, ya, sb, yb);
    if (xa >= 0) s += fly_part_dot(m, w.x_start[xa], w.x_Y + xa * FLY_MAXCHAIN, sb, yb);
    if (xb >= 0) s += fly_part_dot(m, sa, ya, w.x_start[xb], w.x_Y + xb * FLY_MAXCHAIN);
    if (xa >= 0 && xb >= 0) s += fly_part_dot(m, w.x_start[xa], w.x_Y + xa * FLY_MAXCHAIN, w.x_start[xb], w.x_Y + xb * FLY_MAXCHAIN);
    w.efc_A[it] = s;
  }
  FLY_SYNC();
}

/* out <- A v (dense nefc x nefc) */
template <class T>
FLY_INL void fly_a_mul(const Ws<T>& w, int nefc, S* out, const S* v) {
  FLY_FOR(r, nefc) {
    const T* a = w.efc_A + r * nefc;
    S s = 0;
    for (int k = 0; k < nefc; k++) s += a[k] * v[k];
    out[r] = s;
  }
  FLY_SYNC();
}

template <class T>
FLY_FN void fly_newton(const DModel<T>& m, const Ws<T>& w) {
  const int nv = FLY_DIM(m, nv), nefc = w.counts[FLY_C_NEFC];
  int niter = 0;
  if (nefc == 0) {
    FLY_FOR(d, nv) { w.qfrc_constraint[d] = 0; w.stmp[d] = 0; w.sgrad[d] = 0; }
    FLY_LANE0 w.counts[FLY_C_NITER] = 0;
    FLY_SYNC();
    return;
  }
  const S scale = S(1) / ((S)m.meaninertia * S(nv > 1 ? nv : 1));
  S *bs = w.efc_b, *b0 = w.efc_b0, *c = w.efc_c, *cp = w.efc_cp, *Ac = w.efc_Ac, *Acp = w.efc_Acp, *h = w.efc_h;
  S *jar = w.efc_jar, *jv = w.efc_jv;
  T* f = w.efc_force;
  T* dvec = w.ssearch;  /* qacc_warmstart - qacc_smooth */
  T* m0 = w.sMa;        /* M (qacc_warmstart - qacc_smooth) */
  /* warm start: the better of qacc_warmstart and qacc_smooth */
  FLY_FOR(d, nv) dvec[d] = w.warm[d] - w.qacc_smooth[d];
  FLY_SYNC();
  fly_mul_m(m, w.qLD, m0, dvec, w.stmp);
  S dd = 0;
  FLY_FOR(d, nv) dd += (S)dvec[d] * (S)m0[d];
  dd = warp_sum(dd);
  FLY_FOR(r, nefc) {
    bs[r] = (S)fly_row_dot(m, w, r, w.qacc_smooth) - (S)w.efc_aref[r];
    b0[r] = fly_row_dot(m, w, r, dvec);
    jar[r] = bs[r] + b0[r];
    c[r] = 0; Ac[r] = 0;
  }
  FLY_SYNC();
  const S cw = fly_update_constraint(m, w, jar, false, false) + S(0.5) * dd;
  const S cs = fly_update_constraint(m, w, bs, false, false);
  S theta = ((m.flags & FLY_OPT_NOWARMSTART) || cs < cw) ? S(0) : S(1);
  if (theta == 0) { FLY_FOR(r, nefc) jar[r] = bs[r]; FLY_SYNC(); }
  S gauss = S(0.5) * theta * theta * dd;
  S cost = fly_update_constraint(m, w, jar, true, true) + gauss;
  for (int it = 0;; it++) {
    /* gradient in generalized coordinates: theta m0 + J^T (c - f) */
    FLY_FOR(r, nefc) h[r] = c[r] - f[r];
    FLY_SYNC();
    FLY_FOR(r, nefc) w.efc_lam[r] = (T)h[r];
    FLY_SYNC();
    fly_jt_mul(m, w, w.sgrad, w.efc_lam);
    T gn = 0;
    FLY_FOR(d, nv) { const T g = w.sgrad[d] + (T)theta * m0[d]; gn += g * g; }
    gn = warp_sum(gn);
    if (it > 0 && scale * sqrt((S)gn) < m.tolerance) break;
    if (it >= m.opt_iterations) break;
    /* h = Y g~ = theta b0 + A (c - f) */
    fly_a_mul(w, nefc, Acp, h);
    FLY_FOR(r, nefc) h[r] = theta * b0[r] + Acp[r];
    FLY_SYNC();
    /* Newton step (I + W A) x = W h with W the block-diagonal Hessian of the constraint cost (D_r on quadratic rows, the
       rank-2 cone block S_c S_c^T on contacts in the cone zone).  With W = S S^T and x = S z this is the symmetric positive
       definite system (I + S^T A S) z = S^T h: one column of S per quadratic row (a single entry sqrt(D_r)) and two per
       cone contact - smaller than the row system, Cholesky instead of pivoted LU.  lim_item lists the columns by their
       first row: a quadratic row, or row r0 / r0 + 1 of a cone contact for its first / second column. */
    const int nz = warp_compact(nefc, [&](int r) -> bool {
      const int stt = w.efc_state[r];
      return stt == FLY_ST_QUADRATIC || (stt == FLY_ST_CONE && r - w.grp_row[w.efc_g[r]] < 2);
    }, w.lim_item, FLY_MAXEFC);
    FLY_SYNC();
    T* K = w.efc_K;
    T* lam = w.efc_lam;
    /* column a: rows ra .. ra + na - 1 with entries sa[] */
#define FLY_SCOL(a, ra, na, sa)                                                                          \
    int ra, na; T sa[3];                                                                                 \
    {                                                                                                    \
      const int r_ = w.lim_item[a];                                                                      \
      if (w.efc_state[r_] == FLY_ST_CONE) {                                                              \
        const int g_ = w.efc_g[r_], r0_ = w.grp_row[g_];                                                 \
        const T* hc_ = w.con_hc + 9 * w.grp_con[g_] + 3 * (r_ - r0_);                                    \
        ra = r0_; na = 3; sa[0] = hc_[0]; sa[1] = hc_[1]; sa[2] = hc_[2];                                \
      } else { ra = r_; na = 1; sa[0] = sqrt(w.efc_D[r_]); sa[1] = sa[2] = 0; }                          \
    }
    FLY_FOR_ROW(a, nz) {
      FLY_SCOL(a, ra, na, sa)
      FLY_FOR_COL(b, 0, a + 1) {
        FLY_SCOL(b, rb, nb, sb)
        T acc = a == b ? T(1) : T(0);
        for (int i = 0; i < na; i++) {
          T t = 0;
          for (int j = 0; j < nb; j++) t += w.efc_A[(ra + i) * nefc + rb + j] * sb[j];
          acc += sa[i] * t;
        }
        K[a * nz + b] = acc;
      }
    }
    FLY_FOR(a, nz) {
      FLY_SCOL(a, ra, na, sa)
      S t = 0;
      for (int i = 0; i < na; i++) t += (S)sa[i] * h[ra + i];
      lam[a] = (T)t;
    }
    FLY_SYNC();
    fly_chol_solve(K, nz, lam, w.hit_tmp);   /* hit_tmp: scratch of the collision stage, idle here */
    /* search direction in (theta, c) coordinates: c_p = f + x - c with x = S z, theta_p = -theta */
    FLY_FOR(r, nefc) cp[r] = f[r] - c[r];
    FLY_SYNC();
    FLY_FOR(a, nz) {
      const int r_ = w.lim_item[a];
      if (w.efc_state[r_] != FLY_ST_CONE) cp[r_] += (S)(sqrt(w.efc_D[r_]) * lam[a]);
      else {
        const int g_ = w.efc_g[r_], r0_ = w.grp_row[g_];
        if (r_ == r0_) {   /* the lane of a contact's first column adds both columns: they share the three rows */
          const T* hc_ = w.con_hc + 9 * w.grp_con[g_];
          for (int k = 0; k < 3; k++) cp[r0_ + k] += (S)(hc_[k] * lam[a] + hc_[3 + k] * lam[a + 1]);
        }
      }
    }
#undef FLY_SCOL
    FLY_SYNC();
    const S thp = -theta;
    fly_a_mul(w, nefc, Acp, cp);
    S s1 = 0, s2 = 0, s3 = 0, s4 = 0;
    FLY_FOR(r, nefc) {
      jv[r] = thp * b0[r] + Acp[r];
      s1 += b0[r] * c[r]; s2 += b0[r] * cp[r]; s3 += cp[r] * Ac[r]; s4 += cp[r] * Acp[r];
    }
    s1 = warp_sum(s1); s2 = warp_sum(s2); s3 = warp_sum(s3); s4 = warp_sum(s4);
    FLY_SYNC();
    const S q1 = thp * theta * dd + thp * s1 + theta * s2 + s3;
    const S q2 = S(0.5) * (thp * thp * dd + 2 * thp * s2 + s4);
    if (!(q2 > 0)) break;
    /* MuJoCo scales the line-search tolerance with |search|_2; here |search|_M / sqrt(meaninertia) stands in
       for it (the Euclidean norm would cost one more tree solve per iteration) */
    const S snorm = sqrt(2 * q2 / (S)m.meaninertia);
    if (snorm < minval<S>()) break;
    const S gtol = (S)m.tolerance * (S)m.ls_tolerance * snorm * (S)m.meaninertia * S(nv > 1 ? nv : 1);
    const S quad[3] = {gauss, q1, q2};
    const S alpha = fly_line_search(m, w, quad, gtol);
    if (alpha == 0) break;
    S g1 = 0, g2 = 0;
    FLY_FOR(r, nefc) {
      const S cn = c[r] + alpha * cp[r], acn = Ac[r] + alpha * Acp[r];
      c[r] = cn; Ac[r] = acn; jar[r] += alpha * jv[r];
      g1 += b0[r] * cn; g2 += cn * acn;
    }
    g1 = warp_sum(g1); g2 = warp_sum(g2);
    FLY_SYNC();
    theta += alpha * thp;
    gauss = S(0.5) * (theta * theta * dd + 2 * theta * g1 + g2);
    const S old = cost;
    cost = fly_update_constraint(m, w, jar, true, true) + gauss;
    niter = it + 1;
    if (scale * (old - cost) < m.tolerance) break;
  }
  /* the iterate is qacc = qacc_smooth + theta d + M^-1 J^T c: leave J^T c in stmp and theta d in sgrad for the caller's
     solve (with noslip the caller recomputes qacc from the final forces instead) */
  FLY_FOR(r, nefc) w.efc_lam[r] = (T)c[r];
  FLY_FOR(d, nv) w.sgrad[d] = (T)theta * dvec[d];
  FLY_SYNC();
  fly_jt_mul(m, w, w.stmp, w.efc_lam);
  fly_jt_mul(m, w, w.qfrc_constraint, f);
  FLY_LANE0 w.counts[FLY_C_NITER] = niter;
  FLY_SYNC();
}

/* noslip: Gauss-Seidel over the friction dimensions with the unregularised A, normal forces held fixed */
template <class T>
FLY_FN void fly_noslip(const DModel<T>& m, const Ws<T>& w) {
  const int nefc = w.counts[FLY_C_NEFC], ngrp = w.counts[FLY_C_NGRP], nlimit = w.counts[FLY_C_NLIMIT];
  FLY_LANE0 {
    int n3 = 0;
    for (int g = nlimit; g < ngrp; g++) if (w.grp_type[g] == FLY_CT_ELLIPTIC) { w.ns_fr[n3] = w.grp_row[g]; w.ns_con[n3] = w.grp_con[g]; n3++; }
    w.counts[FLY_C_NFR] = n3;
  }
  FLY_SYNC();
  const int n3 = w.counts[FLY_C_NFR];
  if (!n3) return;
  const T scale = T(1) / (m.meaninertia * T(FLY_DIM(m, nv) > 1 ? FLY_DIM(m, nv) : 1));
  T* f = w.efc_force;
  for (int it = 0; it < m.opt_noslip_iterations; it++) {
    T improvement = 0;
    for (int i = 0; i < n3; i++) {
      const int r0 = w.ns_fr[i];
      const T *A0 = w.efc_A + (r0 + 1) * nefc, *A1 = A0 + nefc;
      T res[2] = {0, 0};
      FLY_FOR(k, nefc) { const T fk = f[k]; res[0] += A0[k] * fk; res[1] += A1[k] * fk; }
      res[0] = warp_sum(res[0]) + (T)w.efc_b[r0 + 1];
      res[1] = warp_sum(res[1]) + (T)w.efc_b[r0 + 2];
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
    w.efc_aref[r] = -w.efc_B[r] * vel - w.efc_K0[r] * w.efc_imp[r] * (w.efc_pos[r] - w.efc_margin[r]);
  }
  FLY_SYNC();
  if (nefc) fly_constraint_prepare(m, w);
  FLY_STAGE_SYNC(15);
  fly_newton(m, w);
  FLY_STAGE_SYNC(16);
  if (nefc && !(m.flags & FLY_OPT_NONOSLIP) && m.opt_noslip_iterations > 0) {
    fly_noslip(m, w);
    /* qacc = qacc_smooth + M^-1 J^T f */
    FLY_FOR(d, FLY_DIM(m, nv)) { w.stmp[d] = w.qfrc_constraint[d]; w.sgrad[d] = 0; }
    FLY_SYNC();
  }
  /* Two solves with one walk over the tree: M^-1 stmp for qacc, and the implicit-damping solve of the Euler step
     (m13), (M + dt B)^-1 (qfrc_smooth + qfrc_constraint), whose right-hand side is final here as well. */
  FLY_FOR(d, FLY_DIM(m, nv)) w.sMa[d] = w.qfrc_smooth[d] + w.qfrc_constraint[d];
  FLY_SYNC();
  FLY_STAGE_SYNC(17);
  {
    const T* LD[2] = {w.qLD, w.qH};
    const T* iv[2] = {w.qDinv, w.qHinv};
    T* x[2] = {w.stmp, w.sMa};
    fly_solve<T, 2>(m, w, LD, iv, x);
  }
  FLY_FOR(d, FLY_DIM(m, nv)) w.qacc[d] = w.qacc_smooth[d] + w.sgrad[d] + w.stmp[d];
  FLY_SYNC();
}
#endif
