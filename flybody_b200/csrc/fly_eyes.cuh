/* fly_eyes.cuh - the fly's two eye cameras as a ray caster over the height-field terrain (first slice;
 * vnl_ray/fruitfly/fruitfly.py:676-708: MJCFCamera observables of `eye_right` / `eye_left`, 32 x 32 pixels, fovy 150 degrees,
 * tasks/vision_flight.py:25-26,83-84).  The reference renders them with MuJoCo's OpenGL renderer (unavailable on this box); this
 * is the camera and shading arithmetic as plain device code, one thread per pixel:
 *   camera frame x right, y up, looking along -z; pixel (row i from the top, column j): direction
 *   (tan(fovy/2) (W/H) (2 (j + 1/2) / W - 1), tan(fovy/2) (1 - 2 (i + 1/2) / H), -1), rotated into the world;
 *   scene: the terrain surface max(height - 0.01, 0) over [-half, half]^2 (the `terrain` hfield geom at z = -0.01 together
 *   with the `groundplane` at z = 0, hills.py:186-221), each grid cell split into the two triangles of the collision strip;
 *   everything else is background;
 *   shading: flat colour x (ambient + diffuse max(0, n . l)) with the headlight - the only light the task keeps
 *   (vision_flight.py:66-68; hills.py:223-225 ambient 0.4, diffuse 0.8) - shining along the camera axis, rounded to 8 bits.
 * Not rendered in this slice: the fly's own body (antennae and legs in view), textures and skybox of the "outdoor_natural"
 * aesthetic (dm_control assets, not available here), specular highlights.
 * The ray walks the grid cell by cell (2-D DDA), skips cells whose corner heights lie below the ray, and intersects the two
 * triangles of the others.  R is float on the device and double in the oracle's restatement (oracle/flyb_oracle.c:
 * flyo_render_eyes). */
#ifndef FLY_EYES_CUH_
#define FLY_EYES_CUH_
#include "fly_math.cuh"

struct FlyScene {
  const float* data;      /* [nterrain, nrow, ncol] heights */
  int nrow, ncol;
  float half;             /* half extent in x and y */
  float zoff;             /* z of the hfield geom (-0.01) */
  float rgb[3], sky[3];   /* terrain / ground colour, background */
  float ambient, diffuse; /* headlight */
  /* nullable: per terrain, the largest surface height over each block of FLY_EYE_BLOCK x FLY_EYE_BLOCK cells
     [nterrain, nbr, nbc] - cells of a block the ray stays above are stepped over without reading their heights */
  const float* bmax;      /* per terrain: [nbr, nbc] maxima of the 16 x 16 blocks, then [nbr2, nbc2] of the 4 x 4 blocks */
  int nbr, nbc, nbr2, nbc2;
};
#define FLY_EYE_SHIFT 4
#define FLY_EYE_SHIFT2 2
#define FLY_EYE_BLOCK (1 << FLY_EYE_SHIFT)

template <class R> FLY_INL R eye_height(const float* h, int ncol, int r, int c, R zoff) {
  const R z = (R)h[(size_t)r * ncol + c] + zoff;
  return z > 0 ? z : R(0);
}

/* ray / triangle (Moeller-Trumbore), front or back; returns t or -1 */
template <class R> FLY_INL R eye_tri(const R* o, const R* d, const R* a, const R* b, const R* c, R* nrm) {
  const R e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
  R p[3], q[3];
  cross3(p, d, e2);
  const R det = dot3(e1, p);
  if (det > R(-1e-12) && det < R(1e-12)) return R(-1);
  const R inv = R(1) / det;
  const R s[3] = {o[0] - a[0], o[1] - a[1], o[2] - a[2]};
  const R u = dot3(s, p) * inv;
  if (u < 0 || u > 1) return R(-1);
  cross3(q, s, e1);
  const R v = dot3(d, q) * inv;
  if (v < 0 || u + v > 1) return R(-1);
  const R t = dot3(e2, q) * inv;
  if (t <= 0) return R(-1);
  cross3(nrm, e1, e2);
  return t;
}

/* colour of the ray (o, d) in 0..1; fwd: the camera's viewing direction (headlight) */
template <class R> FLY_INL void eye_trace(const FlyScene& sc, const float* h, const R* o, const R* d, const R* fwd, R* rgb,
                                          const float* bmax = nullptr) {
  const R half = sc.half, cell = 2 * half / (sc.ncol - 1);
  for (int k = 0; k < 3; k++) rgb[k] = sc.sky[k];
  /* clip the ray to the slab |x| <= half, |y| <= half */
  R t0 = 0, t1 = R(1e30);
  for (int a = 0; a < 2; a++) {
    if (d[a] > R(-1e-12) && d[a] < R(1e-12)) { if (o[a] < -half || o[a] > half) return; continue; }
    R ta = (-half - o[a]) / d[a], tb = (half - o[a]) / d[a];
    if (ta > tb) { const R x = ta; ta = tb; tb = x; }
    if (ta > t0) t0 = ta;
    if (tb < t1) t1 = tb;
  }
  if (t0 >= t1) return;
  /* cell of the entry point */
  R px = o[0] + t0 * d[0], py = o[1] + t0 * d[1];
  int c = (int)floor((px + half) / cell), r = (int)floor((py + half) / cell);
  if (c < 0) c = 0; if (c > sc.ncol - 2) c = sc.ncol - 2;
  if (r < 0) r = 0; if (r > sc.nrow - 2) r = sc.nrow - 2;
  const int sc_ = d[0] > 0 ? 1 : -1, sr_ = d[1] > 0 ? 1 : -1;
  const R big = R(1e30);
  R tx = d[0] != 0 ? ((-half + (c + (sc_ > 0 ? 1 : 0)) * cell) - o[0]) / d[0] : big;   /* t of the next column / row boundary */
  R ty = d[1] != 0 ? ((-half + (r + (sr_ > 0 ? 1 : 0)) * cell) - o[1]) / d[1] : big;
  const R dtx = d[0] != 0 ? cell / (d[0] > 0 ? d[0] : -d[0]) : big, dty = d[1] != 0 ? cell / (d[1] > 0 ? d[1] : -d[1]) : big;
  R tin = t0;
  const R idx = d[0] != 0 ? 1 / d[0] : big, idy = d[1] != 0 ? 1 / d[1] : big;
  int blk[2][2] = {{-1, -1}, {-1, -1}};   /* per level: the block found NOT skippable last (its sub-blocks / cells are walked) */
  for (int step = 0; step < 2 * (sc.nrow + sc.ncol); step++) {
    if (bmax) {
      /* empty-space skipping, two levels (blocks of 16 x 16 cells, then of 4 x 4): if the ray stays above the highest point of a
         block until it leaves the block, no cell of the block can pass the height test below, and the walk resumes in the first
         cell behind the block's far edge */
      bool jumped = false;
#pragma unroll
      for (int lv = 0; lv < 2; lv++) {
        const int S = lv == 0 ? FLY_EYE_SHIFT : FLY_EYE_SHIFT2;
        const int br = r >> S, bc = c >> S;
        if (br == blk[lv][0] && bc == blk[lv][1]) continue;
        const float* tbl = lv == 0 ? bmax : bmax + sc.nbr * sc.nbc;
        const int nbc = lv == 0 ? sc.nbc : sc.nbc2;
        const int ce = sc_ > 0 ? ((bc + 1) << S) : (bc << S);   /* grid line of the far edge */
        const int re = sr_ > 0 ? ((br + 1) << S) : (br << S);
        const R txe = d[0] != 0 ? ((-half + (R)ce * cell) - o[0]) * idx : big;
        const R tye = d[1] != 0 ? ((-half + (R)re * cell) - o[1]) * idy : big;
        const R te = txe < tye ? txe : tye;
        const R tz = (te < t1 ? te : t1) * (1 + R(1e-5)) + R(1e-5);
        const R za = o[2] + tin * d[2], zb = o[2] + tz * d[2];
        if (!((za < zb ? za : zb) > (R)tbl[br * nbc + bc] + R(1e-4))) { blk[lv][0] = br; blk[lv][1] = bc; continue; }
        if (te >= t1) return;
        if (txe < tye) {   /* leaves through a column line: the row follows from y there, kept inside the block's rows */
          c = sc_ > 0 ? ce : ce - 1;
          if (c < 0 || c > sc.ncol - 2) return;
          int rr = (int)floor((o[1] + te * d[1] + half) / cell);
          const int rlo = br << S, rhi = ((br + 1) << S) - 1;
          rr = rr < rlo ? rlo : (rr > rhi ? rhi : rr);
          r = rr > sc.nrow - 2 ? sc.nrow - 2 : rr;
        } else {
          r = sr_ > 0 ? re : re - 1;
          if (r < 0 || r > sc.nrow - 2) return;
          int cc = (int)floor((o[0] + te * d[0] + half) / cell);
          const int clo = bc << S, chi = ((bc + 1) << S) - 1;
          cc = cc < clo ? clo : (cc > chi ? chi : cc);
          c = cc > sc.ncol - 2 ? sc.ncol - 2 : cc;
        }
        tin = te;
        tx = d[0] != 0 ? ((-half + (c + (sc_ > 0 ? 1 : 0)) * cell) - o[0]) / d[0] : big;
        ty = d[1] != 0 ? ((-half + (r + (sr_ > 0 ? 1 : 0)) * cell) - o[1]) / d[1] : big;
        jumped = true;
        break;
      }
      if (jumped) continue;
    }
    const R tout = tx < ty ? tx : ty;
    const R h00 = eye_height<R>(h, sc.ncol, r, c, sc.zoff), h01 = eye_height<R>(h, sc.ncol, r, c + 1, sc.zoff);
    const R h10 = eye_height<R>(h, sc.ncol, r + 1, c, sc.zoff), h11 = eye_height<R>(h, sc.ncol, r + 1, c + 1, sc.zoff);
    const R hmax = fmax_(fmax_(h00, h01), fmax_(h10, h11));
    const R za = o[2] + tin * d[2], zb = o[2] + (tout < t1 ? tout : t1) * d[2];
    if ((za < zb ? za : zb) <= hmax) {
      const R x0 = -half + c * cell, y0 = -half + r * cell;
      const R A[3] = {x0, y0, h00}, B[3] = {x0, y0 + cell, h10}, C[3] = {x0 + cell, y0, h01}, D[3] = {x0 + cell, y0 + cell, h11};
      R n1[3], n2[3];
      const R ta = eye_tri(o, d, A, B, C, n1), tb = eye_tri(o, d, B, C, D, n2);
      R t = -1; const R* n = n1;
      if (ta > 0 && (tb <= 0 || ta <= tb)) { t = ta; n = n1; }
      else if (tb > 0) { t = tb; n = n2; }
      if (t > 0) {
        R nn[3] = {n[0], n[1], n[2]};
        normalize3(nn);
        R lam = -(nn[0] * fwd[0] + nn[1] * fwd[1] + nn[2] * fwd[2]);      /* headlight shines along fwd; surfaces are two-sided */
        if (lam < 0) lam = -lam;
        for (int k = 0; k < 3; k++) { const R v = sc.rgb[k] * (sc.ambient + sc.diffuse * lam); rgb[k] = v > 1 ? R(1) : v; }
        return;
      }
    }
    if (tout >= t1) return;
    tin = tout;
    if (tx < ty) { c += sc_; tx += dtx; if (c < 0 || c > sc.ncol - 2) return; }
    else { r += sr_; ty += dty; if (r < 0 || r > sc.nrow - 2) return; }
  }
}

/* pixel (i, j) of a camera at pos with orientation mat (row-major, columns = camera axes in the world) */
template <class R> FLY_INL void eye_pixel(const FlyScene& sc, const float* h, const R* pos, const R* mat, int W, int H, R tanhalf, int i, int j,
                                          unsigned char* out, const float* bmax = nullptr) {
  const R dc[3] = {tanhalf * ((R)W / (R)H) * (2 * ((R)j + R(0.5)) / (R)W - 1), tanhalf * (1 - 2 * ((R)i + R(0.5)) / (R)H), R(-1)};
  R d[3], fwd[3] = {-mat[2], -mat[5], -mat[8]}, rgb[3];
  mat_vec(d, mat, dc);
  normalize3(d);
  eye_trace<R>(sc, h, pos, d, fwd, rgb, bmax);
  for (int k = 0; k < 3; k++) out[k] = (unsigned char)(int)(rgb[k] * R(255) + R(0.5));
}
#endif
