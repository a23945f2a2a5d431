/* fly_dyn.cuh - smooth dynamics of one environment, warp-cooperative.
 *
 * Stages m1-m4, m7-m10 of SURVEY.md section 8(a2): kinematics, subtree CoM / cinert / cdof, composite
 * inertia -> tree-sparse mass matrix, L^T D L factorisation and solves on the tree sparsity, body
 * velocities, passive forces (springs, dampers, inertia-box and ellipsoid fluid models), RNE bias,
 * actuation.  The tree recursions run one depth level at a time with the bodies (or dofs, or matrix
 * entries) of the level spread over the lanes; sums toward the root are gathers over precomputed
 * child / descendant lists so that no atomics are needed and results are deterministic. */
#ifndef FLY_DYN_CUH_
#define FLY_DYN_CUH_
#include "fly_ws.cuh"

/* ---------------------------------------------------------------- m1 kinematics */
template <class T>
FLY_FN void fly_kinematics(const DModel<T>& m, const Ws<T>& w) {
  FLY_LANE0 {
    for (int k = 0; k < 3; k++) { w.xpos[k] = 0; w.xipos[k] = 0; }
    w.xquat[0] = 1; w.xquat[1] = w.xquat[2] = w.xquat[3] = 0;
    for (int k = 0; k < 9; k++) w.xmat[k] = (k % 4 == 0) ? T(1) : T(0);
  }
  FLY_SYNC();
  for (int L = 1; L <= FLY_DIM(m, nlevel); L++) {
    const int adr = m.lvl_adr[L], n = m.lvl_adr[L + 1] - adr;
    FLY_FOR(k, n) {
      const int b = m.lvl_body[adr + k], p = m.body_parent[b];
      T pos[3], quat[4], t[3];
      mat_vec(t, w.xmat + 9 * p, m.body_pos + 3 * b);
      for (int c = 0; c < 3; c++) pos[c] = w.xpos[3 * p + c] + t[c];
      quat_mul(quat, w.xquat + 4 * p, m.body_quat + 4 * b);
      const int j0 = m.body_jntadr[b], j1 = j0 + m.body_jntnum[b];
      for (int j = j0; j < j1; j++) {
        const int qa = m.jnt_qposadr[j], type = m.jnt_type[j];
        if (type == FLYB_JNT_FREE) {
          quat_normalize(w.qpos + qa + 3);
          for (int c = 0; c < 3; c++) { pos[c] = w.qpos[qa + c]; w.xanchor[3 * j + c] = pos[c]; }
          for (int c = 0; c < 4; c++) quat[c] = w.qpos[qa + 3 + c];
          rot_vec_quat(w.xaxis + 3 * j, m.jnt_axis + 3 * j, quat);
          continue;
        }
        T anchor[3];
        rot_vec_quat(t, m.jnt_pos + 3 * j, quat);
        for (int c = 0; c < 3; c++) { anchor[c] = pos[c] + t[c]; w.xanchor[3 * j + c] = anchor[c]; }
        rot_vec_quat(w.xaxis + 3 * j, m.jnt_axis + 3 * j, quat);
        T ql[4], nq[4];
        if (type == FLYB_JNT_BALL) {
          quat_normalize(w.qpos + qa);
          for (int c = 0; c < 4; c++) ql[c] = w.qpos[qa + c];
        } else {
          axis_angle_quat(ql, m.jnt_axis + 3 * j, w.qpos[qa] - m.qpos0[qa]);
        }
        quat_mul(nq, quat, ql);
        for (int c = 0; c < 4; c++) quat[c] = nq[c];
        rot_vec_quat(t, m.jnt_pos + 3 * j, quat);
        for (int c = 0; c < 3; c++) pos[c] = anchor[c] - t[c];
      }
      quat_normalize(quat);
      T mat[9];
      quat2mat(mat, quat);
      for (int c = 0; c < 3; c++) w.xpos[3 * b + c] = pos[c];
      for (int c = 0; c < 4; c++) w.xquat[4 * b + c] = quat[c];
      for (int c = 0; c < 9; c++) w.xmat[9 * b + c] = mat[c];
      mat_vec(t, mat, m.body_ipos + 3 * b);
      for (int c = 0; c < 3; c++) w.xipos[3 * b + c] = pos[c] + t[c];
    }
    FLY_SYNC();
  }
  FLY_FOR(g, FLY_DIM(m, ngeom)) {
    const int b = m.geom_bodyid[g];
    T t[3];
    mat_vec(t, w.xmat + 9 * b, m.geom_pos + 3 * g);
    for (int c = 0; c < 3; c++) w.geom_xpos[3 * g + c] = w.xpos[3 * b + c] + t[c];
  }
  FLY_FOR(s, FLY_DIM(m, nsite)) {
    const int b = m.site_bodyid[s];
    T t[3];
    mat_vec(t, w.xmat + 9 * b, m.site_pos + 3 * s);
    for (int c = 0; c < 3; c++) w.site_xpos[3 * s + c] = w.xpos[3 * b + c] + t[c];
    mat_mul(w.site_xmat + 9 * s, w.xmat + 9 * b, m.site_mat + 9 * s);
  }
  FLY_SYNC();
}

/* world-frame rotation of a collision geom */
template <class T>
FLY_INL void fly_geom_xmat(const DModel<T>& m, const Ws<T>& w, int g, T* mat) {
  mat_mul(mat, w.xmat + 9 * m.geom_bodyid[g], m.geom_mat + 9 * g);
}

/* acc[b] (K numbers per body) <- acc[b] + sum over children, for every non-world body, leaves first */
template <class T>
FLY_INL void fly_tree_upsweep(const DModel<T>& m, T* acc, int K) {
  for (int L = FLY_DIM(m, nlevel) - 1; L >= 1; L--) {
    const int adr = m.lvl_adr[L], n = m.lvl_adr[L + 1] - adr;
    FLY_FOR(it, n * K) {
      const int b = m.lvl_body[adr + it / K], c = it % K;
      const int c0 = m.child_adr[b], c1 = m.child_adr[b + 1];
      if (c1 > c0) {
        T s = acc[K * b + c];
        for (int q = c0; q < c1; q++) s += acc[K * m.child[q] + c];
        acc[K * b + c] = s;
      }
    }
    FLY_SYNC();
  }
}

/* ---------------------------------------------------------------- m2 comPos */
template <class T>
FLY_FN void fly_com_pos(const DModel<T>& m, const Ws<T>& w) {
  FLY_FOR(it, 3 * FLY_DIM(m, nbody)) w.scom[it] = m.body_mass[it / 3] * w.xipos[it];
  FLY_SYNC();
  fly_tree_upsweep(m, w.scom, 3);
  FLY_FOR(it, 3 * FLY_DIM(m, nbody)) {
    const int b = it / 3;
    const T sm = m.body_subtreemass[b];
    w.scom[it] = sm < minval<T>() ? w.xipos[it] : w.scom[it] / sm;
  }
  FLY_SYNC();
  FLY_FOR(bb, FLY_DIM(m, nbody) - 1) {
    const int b = bb + 1;
    T iq[4], R[9];  /* inertial frame orientation (ximat), rebuilt here instead of being stored by the kinematics */
    quat_mul(iq, w.xquat + 4 * b, m.body_iquat + 4 * b);
    quat2mat(R, iq);
    const T* I = m.body_inertia + 3 * b;
    const T* c0 = w.scom + 3 * m.body_rootid[b];
    T d[3], mass = m.body_mass[b];
    T* ci = w.cinert + 10 * b;
    for (int k = 0; k < 3; k++) d[k] = w.xipos[3 * b + k] - c0[k];
    const T dd = dot3(d, d);
    T W[6];
    const int rr[6] = {0, 1, 2, 0, 0, 1}, cc[6] = {0, 1, 2, 1, 2, 2};
    for (int e = 0; e < 6; e++) {
      const int r = rr[e], c = cc[e];
      W[e] = R[3 * r] * I[0] * R[3 * c] + R[3 * r + 1] * I[1] * R[3 * c + 1] + R[3 * r + 2] * I[2] * R[3 * c + 2]
             + mass * ((r == c ? dd : T(0)) - d[r] * d[c]);
    }
    for (int e = 0; e < 6; e++) ci[e] = W[e];
    ci[6] = mass * d[0]; ci[7] = mass * d[1]; ci[8] = mass * d[2]; ci[9] = mass;
  }
  FLY_FOR(j, FLY_DIM(m, njnt)) {
    const int b = m.jnt_bodyid[j];
    int da = m.jnt_dofadr[j];
    const T* c0 = w.scom + 3 * m.body_rootid[b];
    T off[3];
    for (int k = 0; k < 3; k++) off[k] = c0[k] - w.xanchor[3 * j + k];
    const int type = m.jnt_type[j];
    if (type == FLYB_JNT_HINGE) {
      T* cd = w.cdof + 6 * da;
      for (int k = 0; k < 3; k++) cd[k] = w.xaxis[3 * j + k];
      cross3(cd + 3, w.xaxis + 3 * j, off);
    } else {
      if (type == FLYB_JNT_FREE) {
        for (int k = 0; k < 3; k++) {
          T* cd = w.cdof + 6 * (da + k);
          for (int c = 0; c < 6; c++) cd[c] = 0;
          cd[3 + k] = 1;
        }
        da += 3;
      }
      for (int k = 0; k < 3; k++) {
        T ax[3] = {w.xmat[9 * b + k], w.xmat[9 * b + 3 + k], w.xmat[9 * b + 6 + k]};
        T* cd = w.cdof + 6 * (da + k);
        for (int c = 0; c < 3; c++) cd[c] = ax[c];
        cross3(cd + 3, ax, off);
      }
    }
  }
  FLY_SYNC();
}

/* ---------------------------------------------------------------- m4 CRB -> tree-sparse M */
template <class T>
FLY_FN void fly_crb(const DModel<T>& m, const Ws<T>& w) {
  FLY_FOR(it, 10 * FLY_DIM(m, nbody)) w.crb[it] = it < 10 ? T(0) : w.cinert[it];
  FLY_SYNC();
  fly_tree_upsweep(m, w.crb, 10);
  T* f = w.cdof_dot; /* scratch: cdof_dot is rebuilt later, in the velocity stage */
  FLY_FOR(i, FLY_DIM(m, nv)) inert_vec(f + 6 * i, w.crb + 10 * m.dof_bodyid[i], w.cdof + 6 * i);
  FLY_SYNC();
  const T dt = m.timestep;
  FLY_FOR(e, FLY_DIM(m, nM)) {
    const int i = m.mrow[e], j = m.mcol[e], k = e - m.madr[i];
    T s = 0;
    if (k <= m.kdepth[i]) {
      const T *a = w.cdof + 6 * j, *b = f + 6 * i;
      s = a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] + a[5] * b[5];
      if (k == 0) s += m.dof_armature[i];
    }
    /* M goes straight into the two arrays that are factorised in place: qLD <- M, qH <- M + dt * damping (the matrix of
       the implicit-damping Euler step, m13).  M itself is not kept: products with it use the factors (fly_mul_m). */
    w.qLD[e] = s;
    w.qH[e] = k == 0 ? s + dt * m.dof_damping[i] : s;
  }
  FLY_SYNC();
}

/* ---------------------------------------------------------------- tree-sparse L^T D L, branch per lane
 * NS matrices of the same sparsity are processed together (M and M + dt*damping share every index and branch, so
 * the second one rides along in the shadow of the first one's memory latency).
 *
 * LD[s][ai..ai+n) -= t[s] * LD[s][as..as+n) for two rows of one array.  The loads of a block are issued before its
 * stores: the compiler cannot prove the rows do not overlap, and a load-FMA-store chain per element would expose the
 * full shared-memory latency n times. */
template <class T, int NS>
FLY_INL void fly_row_axpy(T* const* LD, int ai, int as, const T* t, int n) {
  int q = 0;
  for (; q + 4 <= n; q += 4) {
    T sv[NS][4], dv[NS][4];
#pragma unroll
    for (int s = 0; s < NS; s++)
#pragma unroll
      for (int e = 0; e < 4; e++) { sv[s][e] = LD[s][as + q + e]; dv[s][e] = LD[s][ai + q + e]; }
#pragma unroll
    for (int s = 0; s < NS; s++)
#pragma unroll
      for (int e = 0; e < 4; e++) LD[s][ai + q + e] = dv[s][e] - t[s] * sv[s][e];
  }
  const int r = n - q;
  if (r > 0) {
    T sv[NS][3], dv[NS][3];
#pragma unroll
    for (int s = 0; s < NS; s++)
#pragma unroll
      for (int e = 0; e < 3; e++) { sv[s][e] = e < r ? LD[s][as + q + e] : T(0); dv[s][e] = e < r ? LD[s][ai + q + e] : T(0); }
#pragma unroll
    for (int s = 0; s < NS; s++)
#pragma unroll
      for (int e = 0; e < 3; e++) if (e < r) LD[s][ai + q + e] = dv[s][e] - t[s] * sv[s][e];
  }
}

/* In place: LD[s] holds a matrix on entry (layout of qM: row i = A(i,i), A(i,parent), A(i,grandparent) ...) and
 * D_i on the diagonal / L(i,ancestor) off it on return; invD[s][i] = 1 / D_i.
 * Phase 1: each lane eliminates the rows of its own branches, deepest dof first (the row updates stay inside the
 * branch: no barrier, no index stream).  Phase 2: one lane per entry of the small trunk block sums the outer products
 * of the branch rows' trunk entries into it, then those entries are normalised.  Phase 3: the trunk chain (at most
 * FLY_MAXTRUNK dofs) is eliminated with one lane per entry. */
template <class T, int NS>
FLY_FN void fly_factor(const DModel<T>& m, T* const* LD, T* const* invD) {
  for (int st = 0; st < FLY_DIM(m, gsteps); st++) {
    FLY_FOR(l, FLY_NL) {
      const int word = FLY_LDG(m.gstep_b + st * FLY_NL + l);
      if (word >= 0) {
        const int k = word & 255, j = (word >> 8) & 31, g = (word >> 13) & 31, ak = word >> 18;
        const int hk = FLY_LDG(m.hdepth + k), dk = FLY_LDG(m.dof_din + k);
        T inv[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) inv[s] = T(1) / LD[s][ak];
        if (j == 0) {
#pragma unroll
          for (int s = 0; s < NS; s++) invD[s][k] = inv[s];
        }
        /* row k (final: all its descendants are done) updates the rows of its in-branch ancestors; this lane takes
           the entries mm = j, j + g, ... of each of them */
        int ai_next = FLY_LDG(m.mdiag + ak + 1);
        for (int r = 1; r <= dk; r++) {
          const int ai = ai_next, n = hk - r + 1;
          ai_next = FLY_LDG(m.mdiag + ak + r + 1);
          T t[NS];
#pragma unroll
          for (int s = 0; s < NS; s++) t[s] = LD[s][ak + r] * inv[s];
          for (int m0 = j; m0 < n; m0 += 4 * g) {
            T sv[NS][4], dv[NS][4];
#pragma unroll
            for (int s = 0; s < NS; s++)
#pragma unroll
              for (int e = 0; e < 4; e++) {
                const int mm = m0 + e * g;
                sv[s][e] = mm < n ? LD[s][ak + r + mm] : T(0);
                dv[s][e] = mm < n ? LD[s][ai + mm] : T(0);
              }
#pragma unroll
            for (int s = 0; s < NS; s++)
#pragma unroll
              for (int e = 0; e < 4; e++) {
                const int mm = m0 + e * g;
                if (mm < n) LD[s][ai + mm] = dv[s][e] - t[s] * sv[s][e];
              }
          }
        }
      }
    }
    FLY_SYNC();
    FLY_FOR(l, FLY_NL) {  /* normalise row k: U(k, ancestor) -> L(k, ancestor), in-branch part */
      const int word = FLY_LDG(m.gstep_b + st * FLY_NL + l);
      if (word >= 0) {
        const int k = word & 255, j = (word >> 8) & 31, g = (word >> 13) & 31, ak = word >> 18, dk = FLY_LDG(m.dof_din + k);
        for (int r = 1 + j; r <= dk; r += g) {
#pragma unroll
          for (int s = 0; s < NS; s++) LD[s][ak + r] *= invD[s][k];
        }
      }
    }
    FLY_SYNC();
  }
  const int nt = FLY_DIM(m, ntrunk), nb = FLY_DIM(m, nv) - nt;
  if (nt == 0) return;
  FLY_FOR(p, nt * (nt + 1) / 2) {
    const int j = FLY_LDG(m.tri_a + p), i = FLY_LDG(m.tri_b + p);
    T acc[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) acc[s] = 0;
#pragma unroll 2
    for (int e = 0; e < nb; e++) {
      const int word = FLY_LDG(m.bt_list + e), ae = word & 2047, ntk = (word >> 11) & 15, k = word >> 15;
      if (i < ntk) {
#pragma unroll
        for (int s = 0; s < NS; s++) acc[s] += LD[s][ae - i] * LD[s][ae - j] * invD[s][k];
      }
    }
    const int ti = FLY_LDG(m.trunk_adr + i);
#pragma unroll
    for (int s = 0; s < NS; s++) LD[s][ti + i - j] -= acc[s];
  }
  FLY_SYNC();
  FLY_FOR(e, nb) {
    const int word = FLY_LDG(m.bt_list + e), ae = word & 2047, ntk = (word >> 11) & 15, k = word >> 15;
    for (int q = 0; q < ntk; q++) {
#pragma unroll
      for (int s = 0; s < NS; s++) LD[s][ae - q] *= invD[s][k];
    }
  }
  for (int k = nt - 1; k >= 1; k--) {
    const int ak = FLY_LDG(m.trunk_adr + k);
    FLY_SYNC();
    T dinv[NS];
#pragma unroll
    for (int s = 0; s < NS; s++) dinv[s] = T(1) / LD[s][ak];
    FLY_FOR(p, k * (k + 1) / 2) {
      const int j = FLY_LDG(m.tri_a + p), i = FLY_LDG(m.tri_b + p), ti = FLY_LDG(m.trunk_adr + i);
#pragma unroll
      for (int s = 0; s < NS; s++) LD[s][ti + i - j] -= LD[s][ak + k - i] * LD[s][ak + k - j] * dinv[s];
    }
    FLY_SYNC();
    FLY_FOR(i, k) {
#pragma unroll
      for (int s = 0; s < NS; s++) LD[s][ak + k - i] *= dinv[s];
    }
  }
  FLY_SYNC();
  FLY_FOR(i, nt) {
#pragma unroll
    for (int s = 0; s < NS; s++) invD[s][FLY_LDG(m.trunk_dof + i)] = T(1) / LD[s][FLY_LDG(m.trunk_adr + i)];
  }
  FLY_SYNC();
}

/* x[s] <- (L^T D L)[s]^-1 x[s] with the same ownership: branch rows by one lane (backward: scatter of the finished
 * x[c] to its in-branch ancestors; forward: gather over them), the trunk chain in registers in between. */
template <class T, int NS>
FLY_FN void fly_solve(const DModel<T>& m, const Ws<T>& w, const T* const* LD, const T* const* invD, T* const* x) {
  for (int st = 0; st < FLY_DIM(m, gsteps); st++) {
    FLY_FOR(l, FLY_NL) {
      const int word = FLY_LDG(m.gstep_b + st * FLY_NL + l);
      if (word >= 0) {
        const int c = word & 255, j = (word >> 8) & 31, g = (word >> 13) & 31, a = word >> 18, dk = FLY_LDG(m.dof_din + c);
        T xc[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) xc[s] = x[s][c];
        for (int k = 1 + j; k <= dk; k += g) {
          const int jx = FLY_LDG(m.mcol + a + k);
#pragma unroll
          for (int s = 0; s < NS; s++) x[s][jx] -= LD[s][a + k] * xc[s];
        }
      }
    }
    FLY_SYNC();
  }
  const int nt = FLY_DIM(m, ntrunk), nb = FLY_DIM(m, nv) - nt;
  if (nt > 0) {
    T acc[NS][FLY_MAXTRUNK], y[NS][FLY_MAXTRUNK];
#pragma unroll
    for (int s = 0; s < NS; s++)
#pragma unroll
      for (int q = 0; q < FLY_MAXTRUNK; q++) acc[s][q] = 0;
    FLY_FOR(e, nb) {
      const int word = FLY_LDG(m.bt_list + e), ae = word & 2047, ntk = (word >> 11) & 15, c = word >> 15;
#pragma unroll
      for (int s = 0; s < NS; s++) {
        const T xc = x[s][c];
#pragma unroll
        for (int q = 0; q < FLY_MAXTRUNK; q++) if (q < ntk) acc[s][q] += LD[s][ae - q] * xc;
      }
    }
    int ta[FLY_MAXTRUNK], td[FLY_MAXTRUNK];
#pragma unroll
    for (int i = 0; i < FLY_MAXTRUNK; i++) {
      td[i] = i < nt ? FLY_LDG(m.trunk_dof + i) : 0;
      ta[i] = i < nt ? FLY_LDG(m.trunk_adr + i) : 0;
    }
#pragma unroll
    for (int s = 0; s < NS; s++) {
      T xt[FLY_MAXTRUNK];
#pragma unroll
      for (int k = FLY_MAXTRUNK - 1; k >= 0; k--) {
        xt[k] = 0;
        if (k < nt) {
          T sum = warp_sum(acc[s][k]);
#pragma unroll
          for (int k2 = k + 1; k2 < FLY_MAXTRUNK; k2++) if (k2 < nt) sum += LD[s][ta[k2] + k2 - k] * xt[k2];
          xt[k] = x[s][td[k]] - sum;
        }
      }
#pragma unroll
      for (int k = 0; k < FLY_MAXTRUNK; k++) {
        y[s][k] = 0;
        if (k < nt) {
          T sum = 0;
#pragma unroll
          for (int k2 = 0; k2 < k; k2++) sum += LD[s][ta[k] + k - k2] * y[s][k2];
          y[s][k] = xt[k] * invD[s][td[k]] - sum;
        }
      }
    }
    FLY_SYNC();
    FLY_LANE0 {
#pragma unroll
      for (int s = 0; s < NS; s++)
#pragma unroll
        for (int k = 0; k < FLY_MAXTRUNK; k++) if (k < nt) x[s][td[k]] = y[s][k];
    }
    FLY_FOR(e, nb) {
      const int word = FLY_LDG(m.bt_list + e), ae = word & 2047, ntk = (word >> 11) & 15, c = word >> 15;
      const int a = FLY_LDG(m.madr + c);
#pragma unroll
      for (int s = 0; s < NS; s++) {
        T sum = 0;
#pragma unroll
        for (int q = 0; q < FLY_MAXTRUNK; q++) if (q < ntk) sum += LD[s][ae - q] * y[s][q];
        x[s][c] -= sum * LD[s][a];
      }
    }
    FLY_SYNC();
  }
  for (int st = 0; st < FLY_DIM(m, gsteps); st++) {
    FLY_FOR(l, FLY_NL) {
      const int word = FLY_LDG(m.gstep_f + st * FLY_NL + l);
      if (word >= 0) {
        const int i = word & 255, j = (word >> 8) & 31, g = (word >> 13) & 31, a = word >> 18, dk = FLY_LDG(m.dof_din + i);
        T sum[NS];
#pragma unroll
        for (int s = 0; s < NS; s++) sum[s] = 0;
        for (int k = 1 + j; k <= dk; k += g) {
          const int jx = FLY_LDG(m.mcol + a + k);
#pragma unroll
          for (int s = 0; s < NS; s++) sum[s] += LD[s][a + k] * x[s][jx];
        }
        if (g > 1) {
#pragma unroll
          for (int s = 0; s < NS; s++) w.gpart[s * FLY_NL + l] = sum[s];
        } else {
#pragma unroll
          for (int s = 0; s < NS; s++) x[s][i] = x[s][i] * invD[s][i] - sum[s];
        }
      }
    }
    FLY_SYNC();
    FLY_FOR(l, FLY_NL) {
      const int word = FLY_LDG(m.gstep_f + st * FLY_NL + l);
      if (word >= 0 && ((word >> 8) & 31) == 0 && ((word >> 13) & 31) > 1) {
        const int i = word & 255, g = (word >> 13) & 31;
#pragma unroll
        for (int s = 0; s < NS; s++) {
          T tot = 0;
          for (int q = 0; q < g; q++) tot += w.gpart[s * FLY_NL + l + q];
          x[s][i] = x[s][i] * invD[s][i] - tot;
        }
      }
    }
    FLY_SYNC();
  }
}

/* r <- M v from the factors, M = L^T D L:  z = D (L v) by a gather along each row, then r = L^T z by a gather over the
 * descendants (warp-cooperative for the few dofs with many descendants).  z is an nv scratch vector. */
template <class T>
FLY_FN void fly_mul_m(const DModel<T>& m, const T* LD, T* r, const T* v, T* z) {
  FLY_FOR(i, FLY_DIM(m, nv)) {
    const int a = m.madr[i], nrow = m.madr[i + 1] - a;
    T s = v[i];
    for (int q = 1; q < nrow; q++) s += LD[a + q] * v[m.mcol[a + q]];
    z[i] = LD[a] * s;
  }
  FLY_SYNC();
  FLY_FOR(i, FLY_DIM(m, nv)) {
    T s = z[i];
    if (m.hsublast[i] - i <= FLY_HEAVY)
      for (int q = m.desc_adr[i]; q < m.desc_adr[i + 1]; q++) {
        const int pk = m.desc[q];
        s += LD[pk >> 8] * z[pk & 255];
      }
    r[i] = s;
  }
  FLY_SYNC();
  for (int h = 0; h < m.nheavy; h++) {
    const int i = m.heavy_dof[h], q0 = m.desc_adr[i], nd = m.desc_adr[i + 1] - q0;
    T s = 0;
    FLY_FOR(q, nd) {
      const int pk = m.desc[q0 + q];
      s += LD[pk >> 8] * z[pk & 255];
    }
    s = warp_sum(s);
    FLY_SYNC();
    FLY_LANE0 r[i] += s;
  }
  FLY_SYNC();
}

/* ---------------------------------------------------------------- m7 comVel */
template <class T>
FLY_FN void fly_com_vel(const DModel<T>& m, const Ws<T>& w) {
  /* cvel of a body = sum of cdof * qvel over its kinematic dof chain, cdof_dot of a dof = (velocity accumulated before its
     joint) x cdof.  Each lane walks the chain of its own dof / body from the root down - the additions a recursion over
     the body levels would make, in its order, so the bits are the same - instead of nlevel dependent rounds of a few
     bodies each. */
  FLY_FOR(d, FLY_DIM(m, nv)) {
    const int j = m.dof_jntid[d], type = m.jnt_type[j], d0 = m.jnt_dofadr[j];
    T* out = w.cdof_dot + 6 * d;
    if (type == FLYB_JNT_FREE && d < d0 + 3) { for (int c = 0; c < 6; c++) out[c] = 0; continue; }
    /* chain entries (leaf first) that belong to this joint and do not count: a ball's three dofs and a free joint's
       three rotations all see the velocity before them */
    const int skip = type == FLYB_JNT_FREE ? d - d0 - 2 : d - d0 + 1;
    const int a = m.madr[d];
    T v[6] = {0, 0, 0, 0, 0, 0};
    for (int k = m.kdepth[d]; k >= skip; k--) {
      const int e = m.mcol[a + k];
      const T qv = w.qvel[e];
      for (int c = 0; c < 6; c++) v[c] += w.cdof[6 * e + c] * qv;
    }
    cross_motion(out, v, w.cdof + 6 * d);
  }
  FLY_FOR(b, FLY_DIM(m, nbody)) {
    T v[6] = {0, 0, 0, 0, 0, 0};
    if (b > 0) {
      const int d = m.body_lastdof[b], a = m.madr[d];
      for (int k = m.kdepth[d]; k >= 0; k--) {
        const int e = m.mcol[a + k];
        const T qv = w.qvel[e];
        for (int c = 0; c < 6; c++) v[c] += w.cdof[6 * e + c] * qv;
      }
    }
    for (int c = 0; c < 6; c++) w.cvel[6 * b + c] = v[c];
  }
  FLY_SYNC();
}

/* 6D velocity [ang; lin] of a frame (pos, mat) rigidly attached to body b, expressed in that frame */
template <class T>
FLY_INL void fly_local_velocity(const DModel<T>& m, const Ws<T>& w, T* lv, int b, const T* pos, const T* mat) {
  const T* c0 = w.scom + 3 * m.body_rootid[b];
  const T* cv = w.cvel + 6 * b;
  T off[3] = {pos[0] - c0[0], pos[1] - c0[1], pos[2] - c0[2]}, t[3], lin[3];
  cross3(t, cv, off);
  for (int k = 0; k < 3; k++) lin[k] = cv[3 + k] + t[k];
  matT_vec(lv, mat, cv);
  matT_vec(lv + 3, mat, lin);
}

/* ---------------------------------------------------------------- m7 passive forces */
template <class T> FLY_INL T pow4(T x) { T y = x * x; return y * y; }

template <class T>
FLY_FN void fly_passive(const DModel<T>& m, const Ws<T>& w) {
  FLY_FOR(d, FLY_DIM(m, nv)) {
    T f = -m.dof_damping[d] * w.qvel[d];
    const int j = m.dof_jntid[d];
    if (m.jnt_type[j] == FLYB_JNT_HINGE) {
      const int qa = m.jnt_qposadr[j];
      f -= m.jnt_stiffness[j] * (w.qpos[qa] - m.qpos_spring[qa]);
    }
    w.qfrc_passive[d] = f;
  }
  FLY_SYNC();
  if (m.flags & FLY_OPT_NOFLUID) return;
  const T rho = m.density, beta = m.viscosity, PI = T(FLY_PI);
  const int nbox = FLY_DIM(m, nfluidbox), nel = nbox + FLY_DIM(m, nfluidgeom);
  /* phase 1: each fluid element -> world-frame spatial force [torque; force] about its tree's CoM */
  FLY_FOR(i, nel) {
    const bool isbox = i < nbox;
    const int e = isbox ? i : i - nbox;
    const int b = isbox ? m.fbox_bodyid[e] : m.fgeom_bodyid[e];
    const T* lp = isbox ? m.fbox_pos + 3 * e : m.fgeom_pos + 3 * e;
    const T* lm = isbox ? m.fbox_mat + 9 * e : m.fgeom_mat + 9 * e;
    T pos[3], mat[9], t[3], lv[6], lf[6];
    mat_vec(t, w.xmat + 9 * b, lp);
    for (int k = 0; k < 3; k++) pos[k] = w.xpos[3 * b + k] + t[k];
    mat_mul(mat, w.xmat + 9 * b, lm);
    fly_local_velocity(m, w, lv, b, pos, mat);
    for (int k = 0; k < 6; k++) lf[k] = 0;
    if (isbox) {
      const T* bx = m.fbox_size + 3 * e;
      if (beta > 0) {
        const T diam = (bx[0] + bx[1] + bx[2]) / T(3);
        for (int k = 0; k < 3; k++) {
          lf[k] = -PI * diam * diam * diam * beta * lv[k];
          lf[3 + k] = -T(3) * PI * diam * beta * lv[3 + k];
        }
      }
      if (rho > 0) {
        lf[3] -= T(0.5) * rho * bx[1] * bx[2] * fabs_(lv[3]) * lv[3];
        lf[4] -= T(0.5) * rho * bx[0] * bx[2] * fabs_(lv[4]) * lv[4];
        lf[5] -= T(0.5) * rho * bx[0] * bx[1] * fabs_(lv[5]) * lv[5];
        lf[0] -= rho * bx[0] * (pow4(bx[1]) + pow4(bx[2])) * fabs_(lv[0]) * lv[0] / T(64);
        lf[1] -= rho * bx[1] * (pow4(bx[0]) + pow4(bx[2])) * fabs_(lv[1]) * lv[1] / T(64);
        lf[2] -= rho * bx[2] * (pow4(bx[0]) + pow4(bx[1])) * fabs_(lv[2]) * lv[2] / T(64);
      }
    } else {
      /* ellipsoid model (wing_*_fluid geoms, vnl_ray/tasks/base.py:309-313) */
      const T* s = m.fgeom_size + 3 * e;
      const T* c = m.fgeom_coef + 12 * e;
      const T blunt = c[1], slender = c[2], angc = c[3], kutta = c[4], magnus = c[5];
      const T *vm = c + 6, *vi = c + 9, *wv = lv, *v = lv + 3;
      T p[3] = {rho * vm[0] * v[0], rho * vm[1] * v[1], rho * vm[2] * v[2]};
      T Lm[3] = {rho * vi[0] * wv[0], rho * vi[1] * wv[1], rho * vi[2] * wv[2]}, a1[3], a2[3], a3[3];
      cross3(a1, p, wv); cross3(a2, p, v); cross3(a3, Lm, wv);
      for (int k = 0; k < 3; k++) { lf[3 + k] += a1[k]; lf[k] += a2[k] + a3[k]; }
      const T vol = T(4.0 / 3.0) * PI * s[0] * s[1] * s[2];
      const T dmax = fmax_(s[0], fmax_(s[1], s[2])), dmin = fmin_(s[0], fmin_(s[1], s[2]));
      const T dmid = s[0] + s[1] + s[2] - dmax - dmin;
      const T A_max = PI * dmax * dmid;
      T mg[3];
      cross3(mg, wv, v);
      for (int k = 0; k < 3; k++) lf[3 + k] += magnus * rho * vol * mg[k];
      const T bc = s[1] * s[2], ca = s[2] * s[0], ab = s[0] * s[1];
      const T pd = pow4(bc) * v[0] * v[0] + pow4(ca) * v[1] * v[1] + pow4(ab) * v[2] * v[2];
      const T pn = bc * bc * v[0] * v[0] + ca * ca * v[1] * v[1] + ab * ab * v[2] * v[2];
      const T A_proj = PI * sqrt(pd / fmax_(minval<T>(), pn));
      const T speed = norm3(v);
      T nrm[3] = {bc * bc * v[0], ca * ca * v[1], ab * ab * v[2]};
      const T cos_alpha = pn / fmax_(minval<T>(), speed * pd);
      T circ[3], kf[3];
      cross3(circ, nrm, v);
      for (int k = 0; k < 3; k++) circ[k] *= kutta * rho * cos_alpha * A_proj;
      cross3(kf, circ, v);
      for (int k = 0; k < 3; k++) lf[3 + k] += kf[k];
      const T eqD = T(2.0 / 3.0) * (s[0] + s[1] + s[2]);
      const T I_max = T(8.0 / 15.0) * PI * dmid * pow4(dmax);
      T mv[3];
      for (int k = 0; k < 3; k++) {
        const T o = fmax_(s[(k + 1) % 3], s[(k + 2) % 3]);
        const T II = T(8.0 / 15.0) * PI * s[k] * pow4(o);
        mv[k] = wv[k] * (angc * II + slender * (I_max - II));
      }
      const T drag_lin = beta * T(3) * PI * eqD + rho * speed * (A_proj * blunt + slender * (A_max - A_proj));
      const T drag_ang = beta * PI * eqD * eqD * eqD + rho * norm3(mv);
      for (int k = 0; k < 3; k++) { lf[3 + k] -= drag_lin * v[k]; lf[k] -= drag_ang * wv[k]; }
      for (int k = 0; k < 6; k++) lf[k] *= c[0];
    }
    T f[3], tq[3], off[3], x[3];
    mat_vec(tq, mat, lf); mat_vec(f, mat, lf + 3);
    const T* c0 = w.scom + 3 * m.body_rootid[b];
    for (int k = 0; k < 3; k++) off[k] = pos[k] - c0[k];
    cross3(x, off, f);
    for (int k = 0; k < 3; k++) { w.boxf[6 * i + k] = tq[k] + x[k]; w.boxf[6 * i + 3 + k] = f[k]; }
  }
  FLY_SYNC();
  /* phase 2: sum per body, then toward the root; qfrc[d] = cdof[d] . F(subtree of d's body) */
  FLY_FOR(it, 6 * FLY_DIM(m, nbody)) {
    const int b = it / 6, c = it % 6;
    T s = 0;
    for (int q = m.body_fbox_adr[b]; q < m.body_fbox_adr[b + 1]; q++) s += w.boxf[6 * m.body_fbox[q] + c];
    for (int g = 0; g < FLY_DIM(m, nfluidgeom); g++) if (m.fgeom_bodyid[g] == b) s += w.boxf[6 * (nbox + g) + c];
    w.bodyf[it] = s;
  }
  FLY_SYNC();
  fly_tree_upsweep(m, w.bodyf, 6);
  FLY_FOR(d, FLY_DIM(m, nv)) {
    const T *cd = w.cdof + 6 * d, *F = w.bodyf + 6 * m.dof_bodyid[d];
    w.qfrc_passive[d] += cd[0] * F[0] + cd[1] * F[1] + cd[2] * F[2] + cd[3] * F[3] + cd[4] * F[4] + cd[5] * F[5];
  }
  FLY_SYNC();
}

/* ---------------------------------------------------------------- m8 / m12 recursive Newton-Euler
 * flg_acc = 0: bias forces into `result`.  flg_acc = 1, post = 1: cacc / cfrc_int with contact forces (sensors). */
template <class T>
FLY_FN void fly_rne(const DModel<T>& m, const Ws<T>& w, T* result, int flg_acc, int post) {
  FLY_LANE0 {
    for (int c = 0; c < 3; c++) { w.cacc[c] = 0; w.cacc[3 + c] = -m.gravity[c]; }
    for (int c = 0; c < 6; c++) w.cfrc_int[c] = 0;
  }
  FLY_SYNC();
  for (int L = 1; L <= FLY_DIM(m, nlevel); L++) {
    const int adr = m.lvl_adr[L], n = m.lvl_adr[L + 1] - adr;
    FLY_FOR(k, n) {
      const int b = m.lvl_body[adr + k];
      T a[6];
      for (int c = 0; c < 6; c++) a[c] = w.cacc[6 * m.body_parent[b] + c];
      const int d0 = m.body_dofadr[b], d1 = d0 + m.body_dofnum[b];
      for (int d = d0; d < d1; d++) {
        const T qv = w.qvel[d], qa = flg_acc ? w.qacc[d] : T(0);
        for (int c = 0; c < 6; c++) a[c] += w.cdof_dot[6 * d + c] * qv + w.cdof[6 * d + c] * qa;
      }
      for (int c = 0; c < 6; c++) w.cacc[6 * b + c] = a[c];
      T Ia[6], Iv[6], x[6];
      inert_vec(Ia, w.cinert + 10 * b, a);
      inert_vec(Iv, w.cinert + 10 * b, w.cvel + 6 * b);
      cross_force(x, w.cvel + 6 * b, Iv);
      for (int c = 0; c < 6; c++) w.cfrc_int[6 * b + c] = Ia[c] + x[c] - (post ? w.cfrc_ext[6 * b + c] : T(0));
    }
    FLY_SYNC();
  }
  fly_tree_upsweep(m, w.cfrc_int, 6);
  if (result) {
    FLY_FOR(d, FLY_DIM(m, nv)) {
      const T *cd = w.cdof + 6 * d, *F = w.cfrc_int + 6 * m.dof_bodyid[d];
      result[d] = cd[0] * F[0] + cd[1] * F[1] + cd[2] * F[2] + cd[3] * F[3] + cd[4] * F[4] + cd[5] * F[5];
    }
    FLY_SYNC();
  }
}
#endif
