/* fly_col.cuh - collision detection (m5) and constraint-row construction (m6, m3) for one environment.
 *
 * The candidate pair list is static (all filters are applied by the host compiler), so pair selection is a
 * warp-wide ordered stream compaction of the bounding-sphere test followed by the analytic narrow phase, one
 * candidate per lane.  Ordered compaction keeps the contact list in pair order, which is what makes contact
 * selection and indices bit-reproducible against the float64 oracle.
 *
 * Constraint Jacobians are never stored dense.  A row is one "dof chain" (values along the path from a body
 * up to the root of the Hessian tree, FLY_MAXCHAIN numbers) or, for the rare contact between two different
 * branches of the fly, two chains. */
#ifndef FLY_COL_CUH_
#define FLY_COL_CUH_
#include "fly_dyn.cuh"

template <class T> struct Hit { T dist, pos[3], n[3], t[3]; int has_t; };

template <class T>
FLY_INL int hit_sphere_sphere(Hit<T>* h, const T* c1, T r1, const T* c2, T r2, T margin) {
  T d[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
  T len = norm3(d), dist = len - r1 - r2;
  if (dist > margin) return 0;
  if (len < minval<T>()) { d[0] = 1; d[1] = 0; d[2] = 0; } else { T s = T(1) / len; d[0] *= s; d[1] *= s; d[2] *= s; }
  h->dist = dist; h->has_t = 0;
  for (int k = 0; k < 3; k++) { h->n[k] = d[k]; h->pos[k] = c1[k] + d[k] * (r1 + T(0.5) * dist); }
  return 1;
}
template <class T>
FLY_INL int hit_plane_sphere(Hit<T>* h, const T* pp, const T* pn, const T* c, T r, T margin) {
  T d[3] = {c[0] - pp[0], c[1] - pp[1], c[2] - pp[2]};
  T dist = dot3(d, pn) - r;
  if (dist > margin) return 0;
  h->dist = dist; h->has_t = 0;
  for (int k = 0; k < 3; k++) { h->n[k] = pn[k]; h->pos[k] = c[k] - pn[k] * (r + T(0.5) * dist); }
  return 1;
}
template <class T>
FLY_INL int hit_plane_capsule(Hit<T>* h, const T* pp, const T* pn, const T* c, const T* ax, const T* size, T margin) {
  T e1[3], e2[3];
  for (int k = 0; k < 3; k++) { e1[k] = c[k] + ax[k] * size[1]; e2[k] = c[k] - ax[k] * size[1]; }
  int n = 0;
  n += hit_plane_sphere(h + n, pp, pn, e1, size[0], margin);
  n += hit_plane_sphere(h + n, pp, pn, e2, size[0], margin);
  for (int i = 0; i < n; i++) { h[i].t[0] = ax[0]; h[i].t[1] = ax[1]; h[i].t[2] = ax[2]; h[i].has_t = 1; }
  return n;
}
template <class T>
FLY_INL int hit_plane_ellipsoid(Hit<T>* h, const T* pp, const T* pn, const T* c, const T* mat, const T* size, T margin) {
  T nl[3], s[3], wv[3];
  matT_vec(nl, mat, pn);
  for (int k = 0; k < 3; k++) s[k] = size[k] * size[k] * nl[k];
  T den = sqrt(size[0] * size[0] * nl[0] * nl[0] + size[1] * size[1] * nl[1] * nl[1] + size[2] * size[2] * nl[2] * nl[2]);
  if (den < minval<T>()) return 0;
  for (int k = 0; k < 3; k++) s[k] = -s[k] / den;
  mat_vec(wv, mat, s);
  T p[3] = {c[0] + wv[0], c[1] + wv[1], c[2] + wv[2]};
  T d[3] = {p[0] - pp[0], p[1] - pp[1], p[2] - pp[2]};
  T dist = dot3(d, pn);
  if (dist > margin) return 0;
  h->dist = dist; h->has_t = 0;
  for (int k = 0; k < 3; k++) { h->n[k] = pn[k]; h->pos[k] = p[k] - pn[k] * T(0.5) * dist; }
  return 1;
}
template <class T>
FLY_INL int hit_plane_cylinder(Hit<T>* h, const T* pp, const T* pn, const T* c, const T* mat, const T* size, T margin) {
  T ax[3] = {mat[2], mat[5], mat[8]};
  T prjaxis = dot3(pn, ax);
  if (prjaxis > 0) { ax[0] = -ax[0]; ax[1] = -ax[1]; ax[2] = -ax[2]; prjaxis = -prjaxis; }
  T d0[3] = {c[0] - pp[0], c[1] - pp[1], c[2] - pp[2]};
  T dist0 = dot3(d0, pn), vec[3];
  for (int k = 0; k < 3; k++) vec[k] = ax[k] * prjaxis - pn[k];
  T len = norm3(vec);
  if (len < T(1e-12)) { vec[0] = mat[0] * size[0]; vec[1] = mat[3] * size[0]; vec[2] = mat[6] * size[0]; }
  else for (int k = 0; k < 3; k++) vec[k] *= size[0] / len;
  T prjvec = dot3(vec, pn);
  T a[3] = {ax[0] * size[1], ax[1] * size[1], ax[2] * size[1]};
  prjaxis *= size[1];
  int n = 0;
  T dist = dist0 + prjaxis + prjvec;
  if (dist > margin) return 0;
  h[n].dist = dist; h[n].has_t = 0;
  for (int k = 0; k < 3; k++) { h[n].n[k] = pn[k]; h[n].pos[k] = c[k] + vec[k] + a[k] - pn[k] * T(0.5) * dist; }
  n++;
  dist = dist0 - prjaxis + prjvec;
  if (dist <= margin) {
    h[n].dist = dist; h[n].has_t = 0;
    for (int k = 0; k < 3; k++) { h[n].n[k] = pn[k]; h[n].pos[k] = c[k] + vec[k] - a[k] - pn[k] * T(0.5) * dist; }
    n++;
  }
  dist = dist0 + prjaxis - T(0.5) * prjvec;
  if (dist <= margin) {
    T v1[3];
    cross3(v1, vec, ax);
    normalize3(v1);
    for (int k = 0; k < 3; k++) v1[k] *= size[0] * T(0.86602540378443864676);
    for (int sgn = 1; sgn >= -1; sgn -= 2) {
      h[n].dist = dist; h[n].has_t = 0;
      for (int k = 0; k < 3; k++) { h[n].n[k] = pn[k]; h[n].pos[k] = c[k] + T(sgn) * v1[k] + a[k] - T(0.5) * vec[k] - pn[k] * T(0.5) * dist; }
      n++;
    }
  }
  return n;
}
template <class T>
FLY_INL int hit_sphere_capsule(Hit<T>* h, const T* c1, T r1, const T* c2, const T* ax, const T* size2, T margin) {
  T d[3] = {c1[0] - c2[0], c1[1] - c2[1], c1[2] - c2[2]};
  T x = dot3(ax, d);
  if (x > size2[1]) x = size2[1]; else if (x < -size2[1]) x = -size2[1];
  T p[3] = {c2[0] + ax[0] * x, c2[1] + ax[1] * x, c2[2] + ax[2] * x};
  return hit_sphere_sphere(h, c1, r1, p, size2[0], margin);
}
template <class T>
FLY_INL int hit_capsule_capsule(Hit<T>* h, const T* c1, const T* a1, const T* size1, const T* c2, const T* a2,
                                const T* size2, T margin) {
  T dif[3] = {c1[0] - c2[0], c1[1] - c2[1], c1[2] - c2[2]};
  T l1 = size1[1], l2 = size2[1];
  T ma = 1, mb = -dot3(a1, a2), mc = 1, u = -dot3(a1, dif), v = dot3(a2, dif);
  T det = ma * mc - mb * mb;
  T p1[3], p2[3];
  if (fabs_(det) >= T(1e-12)) {
    T x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
    if (x1 > l1) { x1 = l1; x2 = (v - mb * l1) / mc; }
    else if (x1 < -l1) { x1 = -l1; x2 = (v + mb * l1) / mc; }
    if (x2 > l2) { x2 = l2; x1 = (u - mb * l2) / ma; if (x1 > l1) x1 = l1; else if (x1 < -l1) x1 = -l1; }
    else if (x2 < -l2) { x2 = -l2; x1 = (u + mb * l2) / ma; if (x1 > l1) x1 = l1; else if (x1 < -l1) x1 = -l1; }
    for (int k = 0; k < 3; k++) { p1[k] = c1[k] + a1[k] * x1; p2[k] = c2[k] + a2[k] * x2; }
    return hit_sphere_sphere(h, p1, size1[0], p2, size2[0], margin);
  }
  int n = 0;
  T x1s[2] = {l1, -l1};
  for (int i = 0; i < 2 && n < 2; i++) {
    T x1 = x1s[i], x2 = (v - mb * x1) / mc;
    if (x2 <= l2 && x2 >= -l2) {
      for (int k = 0; k < 3; k++) { p1[k] = c1[k] + a1[k] * x1; p2[k] = c2[k] + a2[k] * x2; }
      n += hit_sphere_sphere(h + n, p1, size1[0], p2, size2[0], margin);
    }
  }
  T x2s[2] = {l2, -l2};
  for (int i = 0; i < 2 && n < 2; i++) {
    T x2 = x2s[i], x1 = (u - mb * x2) / ma;
    if (x1 < l1 && x1 > -l1) {
      for (int k = 0; k < 3; k++) { p1[k] = c1[k] + a1[k] * x1; p2[k] = c2[k] + a2[k] * x2; }
      n += hit_sphere_sphere(h + n, p1, size1[0], p2, size2[0], margin);
    }
  }
  return n;
}
template <class T>
FLY_INL int hit_sphere_cylinder(Hit<T>* h, const T* c1, T r1, const T* c2, const T* ax, const T* size2, T margin) {
  T d[3] = {c1[0] - c2[0], c1[1] - c2[1], c1[2] - c2[2]};
  T x = dot3(ax, d), pr[3];
  for (int k = 0; k < 3; k++) pr[k] = d[k] - ax[k] * x;
  T psq = dot3(pr, pr), R = size2[0], H = size2[1];
  if (fabs_(x) <= H && psq >= R * R) {
    T p[3] = {c2[0] + ax[0] * x, c2[1] + ax[1] * x, c2[2] + ax[2] * x};
    return hit_sphere_sphere(h, c1, r1, p, R, margin);
  }
  T sg = x >= 0 ? T(1) : T(-1);
  if (psq <= R * R) {
    T pp[3], pn[3];
    for (int k = 0; k < 3; k++) { pn[k] = sg * ax[k]; pp[k] = c2[k] + pn[k] * H; }
    int n = hit_plane_sphere(h, pp, pn, c1, r1, margin);
    if (n) for (int k = 0; k < 3; k++) h->n[k] = -h->n[k];
    return n;
  }
  T pl = sqrt(psq), p[3];
  for (int k = 0; k < 3; k++) p[k] = c2[k] + sg * ax[k] * H + pr[k] * (R / pl);
  return hit_sphere_sphere(h, c1, r1, p, T(0), margin);
}

template <class T>
FLY_INL void make_frame(T* fr, const T* n, const T* hint, int has_hint) {
  T x[3] = {n[0], n[1], n[2]}, y[3] = {0, 0, 0};
  normalize3(x);
  if (has_hint) { y[0] = hint[0]; y[1] = hint[1]; y[2] = hint[2]; }
  if (!has_hint || norm3(y) < T(0.5)) {
    y[0] = 0; y[1] = 0; y[2] = 0;
    if (x[1] < T(0.5) && x[1] > T(-0.5)) y[1] = 1; else y[2] = 1;
  }
  T d = dot3(x, y);
  for (int k = 0; k < 3; k++) y[k] -= d * x[k];
  normalize3(y);
  for (int k = 0; k < 3; k++) { fr[k] = x[k]; fr[3 + k] = y[k]; }
  cross3(fr + 6, x, y);
}

/* ---------------------------------------------------------------- m5 collision driver */
template <class T>
FLY_FN void fly_collision(const DModel<T>& m, const Ws<T>& w) {
  int ncon = 0;
  if (m.flags & FLY_OPT_NOCONTACT) {
    FLY_LANE0 { w.counts[FLY_C_NCON] = 0; w.counts[FLY_C_NCAND] = 0; }
    FLY_SYNC();
    return;
  }
  /* broad phase: bounding spheres (cannot change the result: the narrow phase only reports dist <= margin) */
  int ncand = warp_compact(m.npair, [&](int p) -> bool {
    const int pg = FLY_LDG(m.pair_geom + p), g1 = pg & 255, g2 = pg >> 8;
    const T *c1 = w.geom_xpos + 3 * g1, *c2 = w.geom_xpos + 3 * g2;
    const T d[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
    const T bound = FLY_LDG(m.pair_bound + p);
    if (m.geom_type[g1] == FLYB_GEOM_PLANE) {
      T mat[9];
      fly_geom_xmat(m, w, g1, mat);
      return !(d[0] * mat[2] + d[1] * mat[5] + d[2] * mat[8] > bound);
    }
    return !(dot3(d, d) > bound * bound);
  }, w.cand, FLY_MAXCAND);
  if (ncand > FLY_MAXCAND) { ncand = FLY_MAXCAND; FLY_LANE0 w.counts[FLY_C_OVERFLOW] |= 1; }
  FLY_SYNC();
  /* narrow phase, 32 candidates per round; hits are appended in candidate order */
  for (int base = 0; base < ncand; base += 32) {
    const int nround = ncand - base < 32 ? ncand - base : 32;
    FLY_FOR(l, nround) {
      const int p = w.cand[base + l];
      const int pg = m.pair_geom[p], g1 = pg & 255, g2 = pg >> 8;
      const int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
      const T *c1 = w.geom_xpos + 3 * g1, *c2 = w.geom_xpos + 3 * g2;
      const T *s1 = m.geom_size + 3 * g1, *s2 = m.geom_size + 3 * g2;
      const T margin = m.pair_margin[p];
      T m1[9], m2[9];
      fly_geom_xmat(m, w, g1, m1);
      fly_geom_xmat(m, w, g2, m2);
      const T a1[3] = {m1[2], m1[5], m1[8]}, a2[3] = {m2[2], m2[5], m2[8]};
      Hit<T> h[FLY_NHIT];
      int n = 0;
      if (t1 == FLYB_GEOM_PLANE) {
        if (t2 == FLYB_GEOM_SPHERE) n = hit_plane_sphere(h, c1, a1, c2, s2[0], margin);
        else if (t2 == FLYB_GEOM_CAPSULE) n = hit_plane_capsule(h, c1, a1, c2, a2, s2, margin);
        else if (t2 == FLYB_GEOM_ELLIPSOID) n = hit_plane_ellipsoid(h, c1, a1, c2, m2, s2, margin);
        else if (t2 == FLYB_GEOM_CYLINDER) n = hit_plane_cylinder(h, c1, a1, c2, m2, s2, margin);
      } else if (t1 == FLYB_GEOM_SPHERE) {
        if (t2 == FLYB_GEOM_SPHERE) n = hit_sphere_sphere(h, c1, s1[0], c2, s2[0], margin);
        else if (t2 == FLYB_GEOM_CAPSULE) n = hit_sphere_capsule(h, c1, s1[0], c2, a2, s2, margin);
        else if (t2 == FLYB_GEOM_CYLINDER) n = hit_sphere_cylinder(h, c1, s1[0], c2, a2, s2, margin);
      } else if (t1 == FLYB_GEOM_CAPSULE && t2 == FLYB_GEOM_CAPSULE) {
        n = hit_capsule_capsule(h, c1, a1, s1, c2, a2, s2, margin);
      }
      /* ellipsoid / cylinder pairs off the plane need a general convex routine: not built */
      w.hit_cnt[l] = n;
      for (int i = 0; i < n; i++) {
        T* o = w.hit_tmp + (l * FLY_NHIT + i) * 11;
        o[0] = h[i].dist;
        for (int k = 0; k < 3; k++) { o[1 + k] = h[i].pos[k]; o[4 + k] = h[i].n[k]; o[7 + k] = h[i].has_t ? h[i].t[k] : T(0); }
        o[10] = (T)h[i].has_t;
      }
    }
    FLY_SYNC();
    FLY_LANE0 {
      int s = 0;
      for (int l = 0; l < nround; l++) { w.hit_off[l] = s; s += w.hit_cnt[l]; }
      w.hit_off[32] = s;
    }
    FLY_SYNC();
    const int total = w.hit_off[32];
    FLY_FOR(l, nround) {
      const int p = w.cand[base + l], n = w.hit_cnt[l];
      for (int i = 0; i < n; i++) {
        const int ci = ncon + w.hit_off[l] + i;
        if (ci >= FLY_MAXCON) break;
        const T* o = w.hit_tmp + (l * FLY_NHIT + i) * 11;
        w.con_pair[ci] = p;
        w.con_dim[ci] = m.pair_condim[p];
        w.con_dist[ci] = o[0];
        const T incl = m.pair_margin[p] - m.pair_gap[p];
        w.con_incl[ci] = incl;
        for (int k = 0; k < 3; k++) w.con_pos[3 * ci + k] = o[1 + k];
        make_frame(w.con_frame + 9 * ci, o + 4, o + 7, o[10] != 0);
        w.con_fric[ci] = m.pair_friction[3 * p];
        w.con_mu[ci] = m.pair_friction[3 * p];
        w.con_excl[ci] = o[0] >= incl ? 1 : 0;
        w.con_efc[ci] = -1;
      }
    }
    ncon += total;
    if (ncon > FLY_MAXCON) { ncon = FLY_MAXCON; FLY_LANE0 w.counts[FLY_C_OVERFLOW] |= 2; }
    FLY_SYNC();
  }
  FLY_LANE0 { w.counts[FLY_C_NCON] = ncon; w.counts[FLY_C_NCAND] = ncand; }
  FLY_SYNC();
}
#endif
