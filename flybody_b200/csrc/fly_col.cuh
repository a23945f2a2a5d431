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

/* ---- general convex pairs (ellipsoid / cylinder geoms off the plane): Minkowski Portal Refinement, the routine MuJoCo's
 * mjc_Convex runs through libccd (third party; restated from the published algorithm, see oracle/flyb_oracle.c:
 * mpr_penetration for the notes).  Each object is inflated by margin/2 through its support function, dist = margin - depth,
 * and the normal is rebuilt from the geom shapes at the contact point. */
/* one geom of a convex pair: arithmetic in T, geometry stored as G (the engine's working precision) with element stride gs
 * (1: plain arrays; 32: the lane-interleaved warp scratch) */
#define FLY_GEOM_PRISM 100   /* internal geom type: one prism of a height field */
template <class T, class G> struct CcdObj {
  int type; const G *pos, *mat, *size; int gs; T margin;
  FLY_MEM T p(int k) const { return (T)pos[k * gs]; }
  FLY_MEM T r(int k) const { return (T)mat[k * gs]; }
  FLY_MEM T s(int k) const { return (T)size[k]; }
  FLY_MEM void to_local(T* o, const T* v) const {  /* o = R^T v */
    for (int j = 0; j < 3; j++) o[j] = r(j) * v[0] + r(3 + j) * v[1] + r(6 + j) * v[2];
  }
  FLY_MEM void to_world(T* o, const T* v) const {  /* o = R v */
    for (int i = 0; i < 3; i++) o[i] = r(3 * i) * v[0] + r(3 * i + 1) * v[1] + r(3 * i + 2) * v[2];
  }
};
template <class T> struct Supp { T v[3], v1[3]; };  /* Minkowski-difference point v = v1 - v2 and its point on object 1 */
/* The portal (centre point + three vertices) of one lane's search lives in a scratch array shared by the warp, element e of
 * lane l at [e * 32 + l]: the vertices are replaced through a computed index, which in thread-local arrays means local
 * memory; in shared memory it is an LDS.  The geoms' poses follow it in the same scratch (FLY_GEOM_WORDS). */
#define FLY_PORTAL_WORDS (4 * 6 * FLY_NL)   /* float64 words of portal scratch per warp */
#define FLY_GEOM_WORDS (24 * FLY_NL)        /* working-precision words after it: pos1 mat1 pos2 mat2 per lane */
template <class T> struct Portal {
  T* s;
  FLY_MEM T& at(int i, int e) const { return s[(i * 6 + e) * FLY_NL]; }
  FLY_MEM void get_v(int i, T* o) const { o[0] = at(i, 0); o[1] = at(i, 1); o[2] = at(i, 2); }
  FLY_MEM void get_v1(int i, T* o) const { o[0] = at(i, 3); o[1] = at(i, 4); o[2] = at(i, 5); }
  FLY_MEM void set(int i, const Supp<T>& p) const { for (int k = 0; k < 3; k++) { at(i, k) = p.v[k]; at(i, 3 + k) = p.v1[k]; } }
  FLY_MEM Supp<T> get(int i) const { Supp<T> p; get_v(i, p.v); get_v1(i, p.v1); return p; }
};
template <class T> FLY_INL T ccd_eps() { return sizeof(T) == 4 ? T(1.1920929e-7) : T(2.220446049250313e-16); }  /* CCD_EPS */
#define FLY_MPR_TOLERANCE 1e-6
#ifndef FLY_MPR_ITERATIONS
#define FLY_MPR_ITERATIONS 50
#endif
#define FLY_MPR_PORTAL_CAP 100  /* guard on libccd's open-ended portal loops (never reached in exact arithmetic) */
template <class T> FLY_INL bool ccd_is_zero(T x) { return fabs_(x) < ccd_eps<T>(); }
template <class T> FLY_INL bool ccd_eq(T a, T b) {
  const T ab = fabs_(a - b);
  if (ab < ccd_eps<T>()) return true;
  a = fabs_(a); b = fabs_(b);
  return b > a ? ab < ccd_eps<T>() * b : ab < ccd_eps<T>() * a;
}
template <class T> FLY_INL T sign0(T x) { return x > 0 ? T(1) : (x < 0 ? T(-1) : T(0)); }
template <class T, class G> FLY_INL void ccd_support_geom(const CcdObj<T, G>& o, const T* dir, T* res) {
  T ld[3], r[3] = {0, 0, 0}, w[3];
  o.to_local(ld, dir);
  if (o.type == FLYB_GEOM_SPHERE || o.type == FLYB_GEOM_CAPSULE) {
    for (int k = 0; k < 3; k++) r[k] = ld[k] * o.s(0);
    if (o.type == FLYB_GEOM_CAPSULE) r[2] += sign0(ld[2]) * o.s(1);
  } else if (o.type == FLYB_GEOM_ELLIPSOID) {
    T t[3] = {ld[0] * o.s(0), ld[1] * o.s(1), ld[2] * o.s(2)};
    normalize3(t);
    for (int k = 0; k < 3; k++) r[k] = t[k] * o.s(k);
  } else if (o.type == FLYB_GEOM_CYLINDER) {
    const T t = sqrt(ld[0] * ld[0] + ld[1] * ld[1]);
    if (t > minval<T>()) { const T sc = o.s(0) / t; r[0] = ld[0] * sc; r[1] = ld[1] * sc; }
    r[2] = sign0(ld[2]) * o.s(1);
  } else if (o.type == FLY_GEOM_PRISM) {   /* height-field prism (fly_hfield.cuh): six vertices relative to pos in size[0..18) */
    T best = T(-1e300);
    for (int v = 0; v < 6; v++) {
      const T q0 = o.s(3 * v), q1 = o.s(3 * v + 1), q2 = o.s(3 * v + 2);
      const T d = q0 * ld[0] + q1 * ld[1] + q2 * ld[2];
      if (d > best) { best = d; r[0] = q0; r[1] = q1; r[2] = q2; }
    }
  }
  o.to_world(w, r);
  for (int k = 0; k < 3; k++) res[k] = w[k] + o.p(k) + dir[k] * o.margin;
}
template <class T, class G>
FLY_INL void ccd_support(const CcdObj<T, G>& o1, const CcdObj<T, G>& o2, const T* dir, Supp<T>& s) {
  const T nd[3] = {-dir[0], -dir[1], -dir[2]};
  T v2[3];
  ccd_support_geom(o1, dir, s.v1);
  ccd_support_geom(o2, nd, v2);
  for (int k = 0; k < 3; k++) s.v[k] = s.v1[k] - v2[k];
}
template <class T> FLY_INL void ccd_normalize(T* v) { const T k = T(1) / norm3(v); v[0] *= k; v[1] *= k; v[2] *= k; }  /* ccdVec3Normalize */
template <class T> FLY_INL void portal_dir(const Portal<T>& P, T* dir) {
  T p1[3], p2[3], p3[3], a[3], b[3];
  P.get_v(1, p1); P.get_v(2, p2); P.get_v(3, p3);
  for (int k = 0; k < 3; k++) { a[k] = p2[k] - p1[k]; b[k] = p3[k] - p1[k]; }
  cross3(dir, a, b);
  ccd_normalize(dir);
}
template <class T> FLY_INL bool portal_reach_tolerance(const Portal<T>& P, const Supp<T>& v4, const T* dir) {
  T p1[3], p2[3], p3[3];
  P.get_v(1, p1); P.get_v(2, p2); P.get_v(3, p3);
  const T dv1 = dot3(p1, dir), dv2 = dot3(p2, dir), dv3 = dot3(p3, dir), dv4 = dot3(v4.v, dir);
  const T d = fmin_(fmin_(dv4 - dv1, dv4 - dv2), dv4 - dv3);
  return ccd_eq(d, T(FLY_MPR_TOLERANCE)) || d < T(FLY_MPR_TOLERANCE);
}
template <class T> FLY_INL void expand_portal(const Portal<T>& P, const Supp<T>& v4) {
  T v4v0[3], p0[3], q[3];
  P.get_v(0, p0);
  cross3(v4v0, v4.v, p0);
  int slot;
  P.get_v(1, q);
  if (dot3(q, v4v0) > 0) { P.get_v(2, q); slot = dot3(q, v4v0) > 0 ? 1 : 3; }
  else { P.get_v(3, q); slot = dot3(q, v4v0) > 0 ? 2 : 1; }
  P.set(slot, v4);
}
template <class T> FLY_INL T point_seg_dist2(const T* x0, const T* b, T* wit) {  /* from the origin */
  const T d[3] = {b[0] - x0[0], b[1] - x0[1], b[2] - x0[2]};
  const T t = -dot3(x0, d) / dot3(d, d);
  if (t < 0 || ccd_is_zero(t)) { wit[0] = x0[0]; wit[1] = x0[1]; wit[2] = x0[2]; }
  else if (t > 1 || ccd_eq(t, T(1))) { wit[0] = b[0]; wit[1] = b[1]; wit[2] = b[2]; }
  else for (int k = 0; k < 3; k++) wit[k] = x0[k] + t * d[k];
  return dot3(wit, wit);
}
template <class T> FLY_INL T origin_tri_dist2(const T* x0, const T* B, const T* C, T* wit) {
  T d1[3], d2[3];
  for (int k = 0; k < 3; k++) { d1[k] = B[k] - x0[k]; d2[k] = C[k] - x0[k]; }
  const T v = dot3(d1, d1), w = dot3(d2, d2), p = dot3(x0, d1), q = dot3(x0, d2), r = dot3(d1, d2);
  const T s = (q * r - w * p) / (w * v - r * r), t = (-s * r - q) / w;
  if ((ccd_is_zero(s) || s > 0) && (ccd_eq(s, T(1)) || s < 1) && (ccd_is_zero(t) || t > 0) && (ccd_eq(t, T(1)) || t < 1) &&
      (ccd_eq(t + s, T(1)) || t + s < 1)) {
    for (int k = 0; k < 3; k++) wit[k] = x0[k] + s * d1[k] + t * d2[k];
    return dot3(wit, wit);
  }
  T w2[3], dist = point_seg_dist2(x0, B, wit), d;
  d = point_seg_dist2(x0, C, w2); if (d < dist) { dist = d; wit[0] = w2[0]; wit[1] = w2[1]; wit[2] = w2[2]; }
  d = point_seg_dist2(B, C, w2);  if (d < dist) { dist = d; wit[0] = w2[0]; wit[1] = w2[1]; wit[2] = w2[2]; }
  return dist;
}
/* true and (depth, dir, pos) when the inflated objects intersect; P: this lane's portal scratch.
 * libccd's discoverPortal / refinePortal / findPenetr are three loops around the same step (a support point in some
 * direction, then a decision).  Here they are ONE loop over a per-lane phase, so that the lanes of a warp - each on its own
 * candidate, in whatever phase - share the instructions of the support evaluation instead of running their loops one
 * lane after the other (on the GPU the search is a float64 dependency chain executed by two or three lanes per warp).
 * Decisions and arithmetic are those of the three loops, in their order. */
enum { FLY_PH_V1 = 0, FLY_PH_V2, FLY_PH_V3, FLY_PH_REFINE, FLY_PH_PENETR, FLY_PH_DONE };
template <class T, class G>
FLY_INL bool mpr_penetration(const CcdObj<T, G>& o1, const CcdObj<T, G>& o2, const Portal<T>& P, T* depth, T* pdir, T* pos) {
  Supp<T> v4;
  T dir[3], va[3], vb[3], p0[3], p1[3], p2[3], p3[3], dot;
  int phase = FLY_PH_V1, guard = 0, it = 0, res = 0;  /* res: 0 no intersection, 1 portal found, 2 touching / segment case */
  for (int k = 0; k < 3; k++) p0[k] = o1.p(k) - o2.p(k);
  if (ccd_eq(p0[0], T(0)) && ccd_eq(p0[1], T(0)) && ccd_eq(p0[2], T(0))) p0[0] += ccd_eps<T>() * 10;
  for (int k = 0; k < 3; k++) { P.at(0, k) = p0[k]; P.at(0, 3 + k) = o1.p(k); dir[k] = -p0[k]; }
  ccd_normalize(dir);
  while (phase != FLY_PH_DONE) {
    if (phase >= FLY_PH_REFINE) {
      portal_dir(P, dir);
      if (phase == FLY_PH_REFINE) {
        if (guard++ > FLY_MPR_PORTAL_CAP) { phase = FLY_PH_DONE; continue; }
        P.get_v(1, p1);
        dot = dot3(dir, p1);
        if (ccd_is_zero(dot) || dot > 0) phase = FLY_PH_PENETR;  /* the portal encapsulates the origin: same dir goes on */
      }
    } else if (phase == FLY_PH_V3 && guard++ > FLY_MPR_PORTAL_CAP) { phase = FLY_PH_DONE; continue; }
    ccd_support(o1, o2, dir, v4);
    dot = dot3(v4.v, dir);
    if (phase <= FLY_PH_V3) {
      if (ccd_is_zero(dot) || dot < 0) { phase = FLY_PH_DONE; continue; }  /* the origin is outside the support plane */
      P.get_v(0, p0);
      if (phase == FLY_PH_V1) {
        cross3(dir, p0, v4.v);
        if (ccd_is_zero(dot3(dir, dir))) {  /* origin on the centre ray: findPenetrTouch / findPenetrSegment */
          if (ccd_eq(v4.v[0], T(0)) && ccd_eq(v4.v[1], T(0)) && ccd_eq(v4.v[2], T(0))) {
            *depth = 0; pdir[0] = pdir[1] = pdir[2] = 0;
          } else {
            *depth = norm3(v4.v);
            pdir[0] = v4.v[0]; pdir[1] = v4.v[1]; pdir[2] = v4.v[2];
            ccd_normalize(pdir);
          }
          for (int k = 0; k < 3; k++) pos[k] = v4.v1[k] - T(0.5) * v4.v[k];  /* midpoint of v1 and v2 = v1 - v */
          res = 2; phase = FLY_PH_DONE; continue;
        }
        P.set(1, v4);
        ccd_normalize(dir);
        phase = FLY_PH_V2;
      } else if (phase == FLY_PH_V2) {
        P.get_v(1, p1);
        for (int k = 0; k < 3; k++) { va[k] = p1[k] - p0[k]; vb[k] = v4.v[k] - p0[k]; }
        cross3(dir, va, vb);
        ccd_normalize(dir);
        if (dot3(dir, p0) > 0) {  /* swap vertices 1 and 2 */
          const Supp<T> t = P.get(1);
          P.set(1, v4); P.set(2, t);
          dir[0] = -dir[0]; dir[1] = -dir[1]; dir[2] = -dir[2];
        } else P.set(2, v4);
        phase = FLY_PH_V3; guard = 0;
      } else {
        P.get_v(1, p1); P.get_v(2, p2);
        int cont = 0;
        cross3(va, p1, v4.v);
        T d2 = dot3(va, p0);
        if (d2 < 0 && !ccd_is_zero(d2)) { P.set(2, v4); for (int k = 0; k < 3; k++) p2[k] = v4.v[k]; cont = 1; }
        if (!cont) {
          cross3(va, v4.v, p2);
          d2 = dot3(va, p0);
          if (d2 < 0 && !ccd_is_zero(d2)) { P.set(1, v4); for (int k = 0; k < 3; k++) p1[k] = v4.v[k]; cont = 1; }
        }
        if (!cont) { P.set(3, v4); phase = FLY_PH_REFINE; guard = 0; }
        else {
          for (int k = 0; k < 3; k++) { va[k] = p1[k] - p0[k]; vb[k] = p2[k] - p0[k]; }
          cross3(dir, va, vb);
          ccd_normalize(dir);
        }
      }
    } else if (phase == FLY_PH_REFINE) {
      if (!(ccd_is_zero(dot) || dot > 0) || portal_reach_tolerance(P, v4, dir)) { phase = FLY_PH_DONE; continue; }
      expand_portal(P, v4);
    } else {  /* findPenetr */
      if (portal_reach_tolerance(P, v4, dir) || it > FLY_MPR_ITERATIONS) { res = 1; phase = FLY_PH_DONE; continue; }
      expand_portal(P, v4);
      it++;
    }
  }
  if (res != 1) return res == 2;
  P.get_v(0, p0); P.get_v(1, p1); P.get_v(2, p2); P.get_v(3, p3);
  *depth = sqrt(origin_tri_dist2(p1, p2, p3, pdir));
  if (ccd_is_zero(pdir[0]) && ccd_is_zero(pdir[1]) && ccd_is_zero(pdir[2])) { pdir[0] = dir[0]; pdir[1] = dir[1]; pdir[2] = dir[2]; }
  ccd_normalize(pdir);
  T b[4], vec[3], sum;  /* findPos */
  portal_dir(P, dir);
  cross3(vec, p1, p2); b[0] = dot3(vec, p3);
  cross3(vec, p3, p2); b[1] = dot3(vec, p0);
  cross3(vec, p0, p1); b[2] = dot3(vec, p3);
  cross3(vec, p2, p1); b[3] = dot3(vec, p0);
  sum = b[0] + b[1] + b[2] + b[3];
  if (ccd_is_zero(sum) || sum < 0) {
    b[0] = 0;
    cross3(vec, p2, p3); b[1] = dot3(vec, dir);
    cross3(vec, p3, p1); b[2] = dot3(vec, dir);
    cross3(vec, p1, p2); b[3] = dot3(vec, dir);
    sum = b[1] + b[2] + b[3];
  }
  for (int k = 0; k < 3; k++) {
    T q1 = 0, q2 = 0;
    for (int i = 0; i < 4; i++) { const T a1 = P.at(i, 3 + k); q1 += b[i] * a1; q2 += b[i] * (a1 - P.at(i, k)); }
    pos[k] = T(0.5) * (q1 + q2) / sum;
  }
  return true;
}
template <class T, class G> FLY_INL void convex_geom_normal(const CcdObj<T, G>& o, const T* cpos, T* n) {
  const T dif[3] = {cpos[0] - o.p(0), cpos[1] - o.p(1), cpos[2] - o.p(2)};
  T p[3], nl[3];
  o.to_local(p, dif);
  o.to_local(nl, n);
  if (o.type == FLYB_GEOM_SPHERE) { nl[0] = p[0]; nl[1] = p[1]; nl[2] = p[2]; }
  else if (o.type == FLYB_GEOM_CAPSULE) {
    nl[0] = p[0]; nl[1] = p[1];
    nl[2] = p[2] < -o.s(1) ? p[2] + o.s(1) : (p[2] > o.s(1) ? p[2] - o.s(1) : T(0));
  } else if (o.type == FLYB_GEOM_ELLIPSOID) {
    for (int k = 0; k < 3; k++) nl[k] = p[k] / (o.s(k) * o.s(k));
  } else if (o.type == FLYB_GEOM_CYLINDER) {
    const T dside = fabs_(sqrt(p[0] * p[0] + p[1] * p[1]) - o.s(0)), dcap = fabs_(fabs_(p[2]) - o.s(1));
    if (dside < T(0.5) * dcap) { nl[0] = p[0]; nl[1] = p[1]; nl[2] = 0; }
    else if (dcap < T(0.5) * dside) { nl[0] = 0; nl[1] = 0; nl[2] = p[2]; }
  }
  normalize3(nl);
  o.to_world(n, nl);
}
/* T: arithmetic of the search (float64 in the engine), G: storage of the geometry */
template <class T, class G>
FLY_INL int hit_convex(Hit<T>* h, const CcdObj<T, G>& o1, const CcdObj<T, G>& o2, T margin, const Portal<T>& P) {
  T depth, dir[3], pos[3];
  if (!mpr_penetration(o1, o2, P, &depth, dir, pos)) return 0;
  if (dir[0] == 0 && dir[1] == 0 && dir[2] == 0) return 0;
  T n1[3] = {dir[0], dir[1], dir[2]}, n2[3] = {-dir[0], -dir[1], -dir[2]};
  convex_geom_normal(o1, pos, n1);
  convex_geom_normal(o2, pos, n2);
  for (int k = 0; k < 3; k++) { h->n[k] = n1[k] - n2[k]; h->pos[k] = pos[k]; }
  normalize3(h->n);
  h->dist = margin - depth; h->has_t = 0;
  return 1;
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

/* pair types (t1 <= t2, MuJoCo's geom type order) that have no analytic routine and go to the convex search */
FLY_INL bool fly_pair_is_convex(int t1, int t2) {
  if (t1 == FLYB_GEOM_PLANE) return false;
  if (t1 == FLYB_GEOM_SPHERE) return t2 == FLYB_GEOM_ELLIPSOID;
  return !(t1 == FLYB_GEOM_CAPSULE && t2 == FLYB_GEOM_CAPSULE);
}

/* ---------------------------------------------------------------- m5 collision driver */
#define FLY_MAXHFCON 16   /* contacts kept per geom - height field pair (oracle: FLYO_MAXHFCON) */
#define FLY_HF_OFF (FLY_PORTAL_WORDS + 24 * FLY_NL)   /* float64 words of efc_A in front of the height-field contact slices */
template <class T> FLY_INL int fly_hfield_pair(const DModel<T>& m, const Ws<T>& w, const T* c2, const T* m2, const T* s2, int t2,
                                               double margin, double* out, int g2id);

template <class T>
FLY_FN void fly_collision(const DModel<T>& m, const Ws<T>& w) {
  int ncon = 0;
  if (m.flags & FLY_OPT_NOCONTACT) {
    FLY_LANE0 { w.counts[FLY_C_NCON] = 0; w.counts[FLY_C_NCAND] = 0; }
    FLY_SYNC();
    return;
  }
  /* broad phase: bounding spheres (cannot change the result: the narrow phase only reports dist <= margin).  The geom
     centres are staged in qLD (free until fly_crb; shared memory when the launch keeps it there) so that the 2 x npair position
     reads are close; bit 16 of the pair word says "geom 1 is a plane" (no geom_type lookup on the way) */
  T* gx = (T*)w.qLD;
  FLY_FOR(i, 3 * FLY_DIM(m, ngeom)) gx[i] = w.geom_xpos[i];
  FLY_SYNC();
  int ncand = warp_compact(FLY_DIM(m, npair), [&](int p) -> bool {
    const int pg = FLY_LDG(m.pair_geom + p), g1 = pg & 255, g2 = (pg >> 8) & 255;
    const T bound = FLY_LDG(m.pair_bound + p);
    const T *c1 = gx + 3 * g1, *c2 = gx + 3 * g2;
    const T d[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
    if (FLY_DIM(m, nhfield) && ((pg >> 17) & 1)) return true;   /* height field: its narrow phase tests the bounds */
    if ((pg >> 16) & 1) {
      T mat[9];
      fly_geom_xmat(m, w, g1, mat);
      return !(d[0] * mat[2] + d[1] * mat[5] + d[2] * mat[8] > bound);
    }
    return !(dot3(d, d) > bound * bound);
  }, w.cand, FLY_MAXCAND);
  if (ncand > FLY_MAXCAND) { ncand = FLY_MAXCAND; FLY_LANE0 w.counts[FLY_C_OVERFLOW] |= 1; }
  FLY_SYNC();
  /* mid phase for the convex pairs: the portal search starts with the support point along the line of centres and
     reports "no contact" when that already separates the (margin-inflated) shapes - most candidates end there.  The same
     test in working precision, with a slack far above its rounding error, drops the clearly separated ones before the
     float64 search (cannot change the result); the list is compacted in place, order kept. */
  ncand = warp_filter(ncand, [&](int i) -> bool {
    const int p = w.cand[i];
    const int pg = m.pair_geom[p], g1 = pg & 255, g2 = (pg >> 8) & 255;
    const int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
    if (t1 == FLYB_GEOM_HFIELD || !fly_pair_is_convex(t1, t2)) return true;
    T m1[9], m2[9], dir[3], nd[3], p1[3], p2[3];
    fly_geom_xmat(m, w, g1, m1);
    fly_geom_xmat(m, w, g2, m2);
    const CcdObj<T, T> o1 = {t1, w.geom_xpos + 3 * g1, m1, m.geom_size + 3 * g1, 1, T(0.5) * m.pair_margin[p]};
    const CcdObj<T, T> o2 = {t2, w.geom_xpos + 3 * g2, m2, m.geom_size + 3 * g2, 1, T(0.5) * m.pair_margin[p]};
    for (int k = 0; k < 3; k++) dir[k] = o2.pos[k] - o1.pos[k];
    const T len = norm3(dir);
    if (!(len > T(1e-6))) return true;
    for (int k = 0; k < 3; k++) { dir[k] /= len; nd[k] = -dir[k]; }
    ccd_support_geom(o1, dir, p1);
    ccd_support_geom(o2, nd, p2);
    const T reach = (p1[0] - p2[0]) * dir[0] + (p1[1] - p2[1]) * dir[1] + (p1[2] - p2[2]) * dir[2];
    return !(reach < T(-1e-4));
  }, w.cand);
  FLY_SYNC();
  /* narrow phase, FLY_NL candidates per round; hits are appended in candidate order */
  for (int base = 0; base < ncand; base += FLY_NL) {
    const int nround = ncand - base < FLY_NL ? ncand - base : FLY_NL;
    FLY_FOR(l, nround) {
      const int p = w.cand[base + l];
      const int pg = m.pair_geom[p], g1 = pg & 255, g2 = (pg >> 8) & 255;
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
      bool convex = false;
      if (FLY_DIM(m, nhfield) && t1 == FLYB_GEOM_HFIELD) {
        /* terrain of the vision-guided flight task (fly_hfield.cuh): up to FLY_MAXHFCON contacts, kept in this lane's slice of
           the scratch behind the portal words (float64, like every convex pair) */
        double* hb = (double*)w.efc_A + FLY_HF_OFF + (size_t)l * (7 * FLY_MAXHFCON);
        n = fly_hfield_pair(m, w, c2, m2, s2, t2, (double)margin, hb, g2);
      } else if (t1 == FLYB_GEOM_PLANE) {
        if (t2 == FLYB_GEOM_SPHERE) n = hit_plane_sphere(h, c1, a1, c2, s2[0], margin);
        else if (t2 == FLYB_GEOM_CAPSULE) n = hit_plane_capsule(h, c1, a1, c2, a2, s2, margin);
        else if (t2 == FLYB_GEOM_ELLIPSOID) n = hit_plane_ellipsoid(h, c1, a1, c2, m2, s2, margin);
        else if (t2 == FLYB_GEOM_CYLINDER) n = hit_plane_cylinder(h, c1, a1, c2, m2, s2, margin);
      } else if (t1 == FLYB_GEOM_SPHERE) {
        if (t2 == FLYB_GEOM_SPHERE) n = hit_sphere_sphere(h, c1, s1[0], c2, s2[0], margin);
        else if (t2 == FLYB_GEOM_CAPSULE) n = hit_sphere_capsule(h, c1, s1[0], c2, a2, s2, margin);
        else if (t2 == FLYB_GEOM_CYLINDER) n = hit_sphere_cylinder(h, c1, s1[0], c2, a2, s2, margin);
        else convex = true;
      } else if (t1 == FLYB_GEOM_CAPSULE && t2 == FLYB_GEOM_CAPSULE) {
        n = hit_capsule_capsule(h, c1, a1, s1, c2, a2, s2, margin);
      } else convex = true;
#ifdef FLY_NO_CONVEX   /* kernel experiment: the engine as it was before the convex narrow phase */
      convex = false;
#endif
      if (convex) {
        /* pair types without an analytic routine.  Always in float64: the portal search branches on signs of small
           differences and its contact point is a barycentric estimate on the last portal, so a float32 run takes other
           branches and lands up to a portal width away (measured: normals off by O(1) on the thin wing ellipsoids) */
        /* this lane's slice of the env's scratch: portal (float64) | pos1 mat1 pos2 mat2 (working precision) */
        const Portal<double> P = {(double*)w.efc_A + FLY_LANE()};  /* efc_A: dead between the constraint solve and fly_make_constraint */
        T* gs = (T*)((double*)w.efc_A + FLY_PORTAL_WORDS) + FLY_LANE();
        for (int k = 0; k < 3; k++) { gs[k * FLY_NL] = c1[k]; gs[(12 + k) * FLY_NL] = c2[k]; }
        for (int k = 0; k < 9; k++) { gs[(3 + k) * FLY_NL] = m1[k]; gs[(15 + k) * FLY_NL] = m2[k]; }
        const CcdObj<double, T> o1 = {t1, gs, gs + 3 * FLY_NL, s1, FLY_NL, 0.5 * (double)margin};
        const CcdObj<double, T> o2 = {t2, gs + 12 * FLY_NL, gs + 15 * FLY_NL, s2, FLY_NL, 0.5 * (double)margin};
        Hit<double> hd;
#ifdef FLY_MPR_SKIP  /* kernel experiment: everything but the search itself */
        n = 0;
#else
        n = hit_convex(&hd, o1, o2, (double)margin, P);
#endif
#ifdef FLY_MPR_DISCARD  /* kernel experiment: pay for the search, drop its contacts */
        n = 0;
#endif
        if (n) {
          h[0].dist = (T)hd.dist; h[0].has_t = 0;
          for (int k = 0; k < 3; k++) { h[0].pos[k] = (T)hd.pos[k]; h[0].n[k] = (T)hd.n[k]; }
        }
      }
      w.hit_cnt[l] = n;
      if (!(FLY_DIM(m, nhfield) && t1 == FLYB_GEOM_HFIELD))
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
      w.hit_off[FLY_NL] = s;
    }
    FLY_SYNC();
    const int total = w.hit_off[FLY_NL];
    FLY_FOR(l, nround) {
      const int p = w.cand[base + l], n = w.hit_cnt[l];
      const bool hf = FLY_DIM(m, nhfield) && ((m.pair_geom[p] >> 17) & 1);
      for (int i = 0; i < n; i++) {
        const int ci = ncon + w.hit_off[l] + i;
        if (ci >= FLY_MAXCON) break;
        T ob[11];
        if (hf) {
          const double* hb = (const double*)w.efc_A + FLY_HF_OFF + (size_t)l * (7 * FLY_MAXHFCON) + 7 * i;
          for (int k = 0; k < 7; k++) ob[k] = (T)hb[k];
          ob[7] = ob[8] = ob[9] = ob[10] = 0;
        }
        const T* o = hf ? ob : w.hit_tmp + (l * FLY_NHIT + i) * 11;
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
