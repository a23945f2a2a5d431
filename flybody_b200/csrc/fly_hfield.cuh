/* fly_hfield.cuh - height field against a convex geom (terrain of the vision-guided flight tasks,
 * vnl_ray/tasks/arenas/hills.py; MuJoCo mjc_ConvexHField restated, see oracle/flyb_oracle.c: flyo_narrow_hfield for the notes).
 *
 * One query = one geom, handled by ONE lane: the geom is expressed in the height field's frame, the cells under its bounds are
 * walked as zig-zag strips, every three consecutive strip vertices top a prism (bottom at -base, top raised by the margin) that
 * goes through the portal refinement of fly_col.cuh as object 1.  Arithmetic in float64, like every convex pair. */
#ifndef FLY_HFIELD_CUH_
#define FLY_HFIELD_CUH_
#include "fly_col.cuh"

struct FlyHfield {
  const float* data;   /* [nrow, ncol] heights before scaling */
  int nrow, ncol;
  double pos[3], mat[9];   /* pose of the height-field geom */
  double size[4];          /* x half extent, y half extent, elevation scale, base thickness */
};

/* contacts of geom (type, world pos / mat, size) with the field; out: dist, pos(3), normal(3) per contact; P: this lane's
 * portal scratch.  Returns the number of contacts written (at most maxcon). */
FLY_INL int fly_hfield_narrow(const FlyHfield& hf, int type, const double* c2, const double* m2, const double* s2, double margin,
                              int maxcon, double* out, const Portal<double>& P) {
  typedef double T;
  T pos[3], mat[9];
  {
    const T d[3] = {c2[0] - hf.pos[0], c2[1] - hf.pos[1], c2[2] - hf.pos[2]};
    matT_vec(pos, hf.mat, d);
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) mat[3 * i + j] = hf.mat[i] * m2[j] + hf.mat[3 + i] * m2[3 + j] + hf.mat[6 + i] * m2[6 + j];
  }
  const CcdObj<T, T> geom = {type, pos, mat, s2, 1, T(0)};
  T lo[3], hi[3];
  for (int a = 0; a < 3; a++) {
    T dir[3] = {0, 0, 0}, p[3];
    dir[a] = 1; ccd_support_geom(geom, dir, p); hi[a] = p[a];
    dir[a] = -1; ccd_support_geom(geom, dir, p); lo[a] = p[a];
  }
  const T sx = hf.size[0], sy = hf.size[1];
  if (hi[0] < -sx || lo[0] > sx || hi[1] < -sy || lo[1] > sy || hi[2] < -hf.size[3]) return 0;
  const T dx = 2 * sx / (hf.ncol - 1), dy = 2 * sy / (hf.nrow - 1);
  int cmin = (int)floor((lo[0] + sx) / dx), cmax = (int)ceil((hi[0] + sx) / dx);
  int rmin = (int)floor((lo[1] + sy) / dy), rmax = (int)ceil((hi[1] + sy) / dy);
  if (cmin < 0) cmin = 0;
  if (rmin < 0) rmin = 0;
  if (cmax > hf.ncol - 1) cmax = hf.ncol - 1;
  if (rmax > hf.nrow - 1) rmax = hf.nrow - 1;
  const T ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  int ncon = 0;
  for (int r = rmin; r < rmax; r++) {
    T top[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};   /* the last three strip vertices (shifted before the first is set: start defined) */
    int nvert = 0;
    for (int c = cmin; c <= cmax; c++)
      for (int i = 0; i < 2; i++) {
        for (int k = 0; k < 3; k++) { top[0][k] = top[1][k]; top[1][k] = top[2][k]; }
        top[2][0] = dx * c - sx; top[2][1] = dy * (r + i) - sy;
        top[2][2] = (T)hf.data[(size_t)(r + i) * hf.ncol + c] * hf.size[2] + margin;
        if (++nvert < 3) continue;
        if (top[0][2] < lo[2] && top[1][2] < lo[2] && top[2][2] < lo[2]) continue;
        T cen[3] = {0, 0, 0}, verts[18];
        for (int v = 0; v < 3; v++) { cen[0] += top[v][0] / 3; cen[1] += top[v][1] / 3; cen[2] += T(0.5) * (top[v][2] - hf.size[3]) / 3; }
        for (int v = 0; v < 3; v++) {
          for (int k = 0; k < 2; k++) { verts[3 * v + k] = top[v][k] - cen[k]; verts[9 + 3 * v + k] = top[v][k] - cen[k]; }
          verts[3 * v + 2] = top[v][2] - cen[2]; verts[9 + 3 * v + 2] = -hf.size[3] - cen[2];
        }
        const CcdObj<T, T> prism = {FLY_GEOM_PRISM, cen, ident, verts, 1, T(0)};
        T depth, dir[3], cp[3];
        if (!mpr_penetration(prism, geom, P, &depth, dir, cp)) continue;
        if ((dir[0] == 0 && dir[1] == 0 && dir[2] == 0) || ncon >= maxcon) continue;
        const T lp[3] = {cp[0] - T(0.5) * margin * dir[0], cp[1] - T(0.5) * margin * dir[1], cp[2] - T(0.5) * margin * dir[2]};
        T wp[3], wn[3];
        mat_vec(wp, hf.mat, lp); mat_vec(wn, hf.mat, dir);
        out[7 * ncon] = margin - depth;
        for (int k = 0; k < 3; k++) { out[7 * ncon + 1 + k] = wp[k] + hf.pos[k]; out[7 * ncon + 4 + k] = wn[k]; }
        ncon++;
      }
  }
  return ncon;
}

/* the height-field pair of the step kernel's narrow phase (fly_col.cuh): geom (type t2, pose c2 / m2, size s2) against the
 * terrain this env flies over - w.counts[FLY_C_TERRAIN] of the pool m.hf_pool; out: dist, pos(3), normal(3) per contact */
template <class T>
FLY_INL int fly_hfield_pair(const DModel<T>& m, const Ws<T>& w, const T* c2, const T* m2, const T* s2, int t2, double margin,
                            double* out, int g2id) {
  if (!m.hf_pool) return 0;
  const int g = m.hfield_geom;
  if (m.hf_bmax && m.hfield_size[2] == 1.0) {
    /* working-precision reject (cannot change the result): the prism tops reach at most the block maximum (+ margin), so a geom
       whose bounding sphere, margin included, ends above the maxima of the blocks under it has no contact.  The table holds
       max(height + z of the geom, 0) per block of cells of an axis-aligned field */
    T hm[9];
    fly_geom_xmat(m, w, g, hm);
    if (hm[0] == T(1) && hm[4] == T(1) && hm[8] == T(1) && fabs_(w.geom_xpos[3 * g + 2] + T(0.01)) < T(1e-6)) {   /* the table's frame */
      const T rb = m.geom_rbound[g2id] + (T)margin + T(1e-5);
      const T sx = (T)m.hfield_size[0], sy = (T)m.hfield_size[1];
      const T dx = 2 * sx / (m.hfield_ncol - 1), dy = 2 * sy / (m.hfield_nrow - 1);
      const T px = c2[0] - w.geom_xpos[3 * g], py = c2[1] - w.geom_xpos[3 * g + 1];
      int c0 = (int)floor((px - rb + sx) / dx) - 1, c1 = (int)ceil((px + rb + sx) / dx) + 1;
      int r0 = (int)floor((py - rb + sy) / dy) - 1, r1 = (int)ceil((py + rb + sy) / dy) + 1;
      if (c1 >= 0 && r1 >= 0 && c0 <= m.hfield_ncol - 1 && r0 <= m.hfield_nrow - 1) {
        c0 = c0 < 0 ? 0 : c0; r0 = r0 < 0 ? 0 : r0;
        c1 = c1 > m.hfield_ncol - 2 ? m.hfield_ncol - 2 : c1; r1 = r1 > m.hfield_nrow - 2 ? m.hfield_nrow - 2 : r1;
        /* the finer of the two tables: blocks of 4 x 4 cells (0.4 cm), behind the [hf_nbr, hf_nbc] table of the 16 x 16 blocks */
        const float* tb = m.hf_bmax + (size_t)w.counts[FLY_C_TERRAIN] * m.hf_bstride + (size_t)m.hf_nbr * m.hf_nbc;
        float top = 0.f;
        for (int br = r0 >> 2; br <= (r1 >> 2); br++)
          for (int bc = c0 >> 2; bc <= (c1 >> 2); bc++) top = fmax_(top, tb[br * m.hf_nbc2 + bc]);
        if (c2[2] - rb > (T)top) return 0;
      }
    }
  }
  FlyHfield hf;
  hf.data = m.hf_pool + (size_t)w.counts[FLY_C_TERRAIN] * m.hfield_nrow * m.hfield_ncol;
  hf.nrow = m.hfield_nrow; hf.ncol = m.hfield_ncol;
  for (int k = 0; k < 3; k++) hf.pos[k] = (double)w.geom_xpos[3 * g + k];
  T hm[9];
  fly_geom_xmat(m, w, g, hm);
  for (int k = 0; k < 9; k++) hf.mat[k] = (double)hm[k];
  for (int k = 0; k < 4; k++) hf.size[k] = m.hfield_size[k];
  double c[3], mm[9], sz[3];
  for (int k = 0; k < 3; k++) { c[k] = (double)c2[k]; sz[k] = (double)s2[k]; }
  for (int k = 0; k < 9; k++) mm[k] = (double)m2[k];
  const Portal<double> P = {(double*)w.efc_A + FLY_LANE()};
  return fly_hfield_narrow(hf, t2, c, mm, sz, margin, FLY_MAXHFCON, out, P);
}
#endif
