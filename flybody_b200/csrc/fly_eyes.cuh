/* fly_eyes.cuh - the fly's two eye cameras as a ray caster over the height-field terrain (SURVEY.md 8f rank 4;
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
 * Not rendered: the fly's own body (antennae and legs in view), textures and skybox of the "outdoor_natural"
 * aesthetic (dm_control assets, not available here), specular highlights.
 * The ray walks the grid cell by cell (2-D DDA), skips cells whose corner heights lie below the ray, and intersects the two
 * triangles of the others; with block maxima (flyb_render) it jumps over blocks of cells it stays above.  R is float on the
 * device; the oracle is the brute-force float64 renderer oracle/eyes_ref.py (every triangle against every ray). */
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
  const float* bmax;      /* per terrain: [nbr, nbc] maxima of the 16 x 16 blocks, [nbr2, nbc2] of the 4 x 4 blocks, the terrain's maximum */
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

/* One ray as a state machine: eye_ray_init clips it to the terrain's slab and finds its first cell, eye_ray_step takes one step of
 * the walk (a jump over a block of cells the ray stays above, or one cell with its height test and two triangles).  eye_trace
 * loops over the steps; the persistent kernel of flyb_render interleaves the steps of 32 rays per warp and hands a lane its next
 * ray as soon as its current one ends (rays of one image row take 5 to 200 steps: without that 6 of 32 lanes were busy). */
template <class R> struct EyeRay {
  R o[3], d[3], fwd[3];
  R t1, tin, tx, ty, dtx, dty, idx, idy;
  int c, r, sc_, sr_, steps;
  int blk[2][2];   /* per level: the block found NOT skippable last (its sub-blocks / cells are walked) */
  R hc[4];         /* corner heights of the candidate cell (eye_ray_advance -> eye_ray_test) */
  R gmax;          /* with block maxima: the terrain's highest point + 1e-4 (eye_ray_set_gmax), above which a rising ray is done */
};
template <class R> FLY_INL void eye_ray_set_gmax(const FlyScene& sc, const float* bmax, EyeRay<R>& y) {
  y.gmax = bmax ? (R)bmax[sc.nbr * sc.nbc + sc.nbr2 * sc.nbc2] + R(1e-4) : R(1e30);
}

/* false: the ray misses the slab |x|, |y| <= half (background) */
template <class R> FLY_INL bool eye_ray_init(const FlyScene& sc, EyeRay<R>& y) {
  const R half = sc.half, cell = 2 * half / (sc.ncol - 1), big = R(1e30);
  const R *o = y.o, *d = y.d;
  R t0 = 0, t1 = big;
  for (int a = 0; a < 2; a++) {
    if (d[a] > R(-1e-12) && d[a] < R(1e-12)) { if (o[a] < -half || o[a] > half) return false; continue; }
    R ta = (-half - o[a]) / d[a], tb = (half - o[a]) / d[a];
    if (ta > tb) { const R x = ta; ta = tb; tb = x; }
    if (ta > t0) t0 = ta;
    if (tb < t1) t1 = tb;
  }
  if (t0 >= t1) return false;
  const R px = o[0] + t0 * d[0], py = o[1] + t0 * d[1];
  int c = (int)floor((px + half) / cell), r = (int)floor((py + half) / cell);
  if (c < 0) c = 0; if (c > sc.ncol - 2) c = sc.ncol - 2;
  if (r < 0) r = 0; if (r > sc.nrow - 2) r = sc.nrow - 2;
  y.sc_ = d[0] > 0 ? 1 : -1; y.sr_ = d[1] > 0 ? 1 : -1;
  y.tx = d[0] != 0 ? ((-half + (c + (y.sc_ > 0 ? 1 : 0)) * cell) - o[0]) / d[0] : big;   /* t of the next column / row boundary */
  y.ty = d[1] != 0 ? ((-half + (r + (y.sr_ > 0 ? 1 : 0)) * cell) - o[1]) / d[1] : big;
  y.dtx = d[0] != 0 ? cell / (d[0] > 0 ? d[0] : -d[0]) : big; y.dty = d[1] != 0 ? cell / (d[1] > 0 ? d[1] : -d[1]) : big;
  y.idx = d[0] != 0 ? 1 / d[0] : big; y.idy = d[1] != 0 ? 1 / d[1] : big;
  y.c = c; y.r = r; y.t1 = t1; y.tin = t0; y.steps = 0;
  y.blk[0][0] = y.blk[0][1] = y.blk[1][0] = y.blk[1][1] = -1;
  return true;
}

/* the DDA step to the neighbouring cell; returns 0: go on, 1: left the terrain */
template <class R> FLY_INL int eye_ray_next_cell(const FlyScene& sc, EyeRay<R>& y) {
  const R tout = y.tx < y.ty ? y.tx : y.ty;
  if (tout >= y.t1) return 1;
  y.tin = tout;
  if (y.tx < y.ty) { y.c += y.sc_; y.tx += y.dtx; if (y.c < 0 || y.c > sc.ncol - 2) return 1; }
  else { y.r += y.sr_; y.ty += y.dty; if (y.r < 0 || y.r > sc.nrow - 2) return 1; }
  return 0;
}

template <class R> FLY_INL bool eye_ray_in_known_blocks(const EyeRay<R>& y) {
  return (y.r >> FLY_EYE_SHIFT) == y.blk[0][0] && (y.c >> FLY_EYE_SHIFT) == y.blk[0][1] &&
         (y.r >> FLY_EYE_SHIFT2) == y.blk[1][0] && (y.c >> FLY_EYE_SHIFT2) == y.blk[1][1];
}

/* the walk up to the next cell the ray may hit: jumps over blocks, cells with the height test.  Returns 0: go on, 1: left the
 * terrain (background), 3: the current cell (y.c, y.r, corner heights in y.hc) passes the height test - eye_ray_test decides.
 * Split from the triangle tests so that the lanes of a warp can run the expensive part together (flyb_render). */
template <class R> FLY_INL int eye_ray_advance(const FlyScene& sc, const float* h, const float* bmax, EyeRay<R>& y) {
  const R half = sc.half, cell = 2 * half / (sc.ncol - 1), icell = (sc.ncol - 1) / (2 * half), big = R(1e30);
  const R *o = y.o, *d = y.d;
  if (++y.steps > 2 * (sc.nrow + sc.ncol)) return 1;
  if (bmax) {
    /* empty-space skipping, two levels (blocks of 16 x 16 cells, then of 4 x 4): if the ray stays above the highest point of a
       block until it leaves the block, no cell of the block can pass the height test below, and the walk resumes in the first
       cell behind the block's far edge.  Up to four jumps per step, so that the lanes of a warp reach their next cell test
       together.  A rising ray above the terrain's highest point is done. */
    if (d[2] >= 0 && o[2] + y.tin * d[2] > y.gmax) return 1;
    for (int attempt = 0;; attempt++) {
      int lv = -1;
      R te = 0, txe = 0, tye = 0;
      int br = 0, bc = 0, ce = 0, re = 0, S = 0;
#pragma unroll
      for (int v = 0; v < 2; v++) {
        if (lv >= 0) continue;
        S = v == 0 ? FLY_EYE_SHIFT : FLY_EYE_SHIFT2;
        br = y.r >> S; bc = y.c >> S;
        if (br == y.blk[v][0] && bc == y.blk[v][1]) continue;
        const float* tbl = v == 0 ? bmax : bmax + sc.nbr * sc.nbc;
        const int nbc = v == 0 ? sc.nbc : sc.nbc2;
        ce = y.sc_ > 0 ? ((bc + 1) << S) : (bc << S);   /* grid line of the far edge */
        re = y.sr_ > 0 ? ((br + 1) << S) : (br << S);
        txe = d[0] != 0 ? ((-half + (R)ce * cell) - o[0]) * y.idx : big;
        tye = d[1] != 0 ? ((-half + (R)re * cell) - o[1]) * y.idy : big;
        te = txe < tye ? txe : tye;
        const R tz = (te < y.t1 ? te : y.t1) * (1 + R(1e-5)) + R(1e-5);
        const R za = o[2] + y.tin * d[2], zb = o[2] + tz * d[2];
        if (!((za < zb ? za : zb) > (R)tbl[br * nbc + bc] + R(1e-4))) { y.blk[v][0] = br; y.blk[v][1] = bc; continue; }
        lv = v;
      }
      if (lv < 0) break;           /* both levels need a closer look: test the cell */
      if (te >= y.t1) return 1;
      int c, r;
      if (txe < tye) {   /* leaves through a column line: the row follows from y there, kept inside the block's rows */
        c = y.sc_ > 0 ? ce : ce - 1;
        if (c < 0 || c > sc.ncol - 2) return 1;
        int rr = (int)floor((o[1] + te * d[1] + half) * icell);
        const int rlo = br << S, rhi = ((br + 1) << S) - 1;
        rr = rr < rlo ? rlo : (rr > rhi ? rhi : rr);
        r = rr > sc.nrow - 2 ? sc.nrow - 2 : rr;
      } else {
        r = y.sr_ > 0 ? re : re - 1;
        if (r < 0 || r > sc.nrow - 2) return 1;
        int cc = (int)floor((o[0] + te * d[0] + half) * icell);
        const int clo = bc << S, chi = ((bc + 1) << S) - 1;
        cc = cc < clo ? clo : (cc > chi ? chi : cc);
        c = cc > sc.ncol - 2 ? sc.ncol - 2 : cc;
      }
      y.c = c; y.r = r; y.tin = te;
      /* (reciprocals instead of divisions in the jump path: an IEEE division is ~10 instructions, and these two lines were 12 % of
         the renderer's; where the walk resumes differs by an ulp at most, which the cell tests that follow absorb) */
      y.tx = d[0] != 0 ? ((-half + (c + (y.sc_ > 0 ? 1 : 0)) * cell) - o[0]) * y.idx : big;
      y.ty = d[1] != 0 ? ((-half + (r + (y.sr_ > 0 ? 1 : 0)) * cell) - o[1]) * y.idy : big;
      if (d[2] >= 0 && o[2] + te * d[2] > y.gmax) return 1;
      if (attempt >= 3) return 0;
    }
  }
  const int c = y.c, r = y.r;
  const R tout = y.tx < y.ty ? y.tx : y.ty;
  const R h00 = eye_height<R>(h, sc.ncol, r, c, sc.zoff), h01 = eye_height<R>(h, sc.ncol, r, c + 1, sc.zoff);
  const R h10 = eye_height<R>(h, sc.ncol, r + 1, c, sc.zoff), h11 = eye_height<R>(h, sc.ncol, r + 1, c + 1, sc.zoff);
  const R hmax = fmax_(fmax_(h00, h01), fmax_(h10, h11));
  const R za = o[2] + y.tin * d[2], zb = o[2] + (tout < y.t1 ? tout : y.t1) * d[2];
  if ((za < zb ? za : zb) <= hmax) { y.hc[0] = h00; y.hc[1] = h01; y.hc[2] = h10; y.hc[3] = h11; return 3; }
  return eye_ray_next_cell(sc, y);
}

/* The walk as separate moves for a warp that schedules them (flyb_render): eye_ray_jump tries ONE jump over a block the ray stays above -
 * returns 0: jumped, 1: left the terrain, 2: no block to jump (the cell has to be looked at); eye_ray_cell tests the current cell -
 * returns 3: candidate (corner heights in y.hc), 1: left the terrain, 0: moved on to the next cell, which still lies in the blocks
 * found not skippable, 4: moved on and a jump may be possible again. */
template <class R> FLY_INL int eye_ray_jump(const FlyScene& sc, const float* bmax, EyeRay<R>& y) {
  const R half = sc.half, cell = 2 * half / (sc.ncol - 1), icell = (sc.ncol - 1) / (2 * half), big = R(1e30);
  const R *o = y.o, *d = y.d;
  if (++y.steps > 4 * (sc.nrow + sc.ncol)) return 1;
  if (d[2] >= 0 && o[2] + y.tin * d[2] > y.gmax) return 1;
  int lv = -1;
  R te = 0, txe = 0, tye = 0;
  int br = 0, bc = 0, ce = 0, re = 0, S = 0;
#pragma unroll
  for (int v = 0; v < 2; v++) {
    if (lv >= 0) continue;
    S = v == 0 ? FLY_EYE_SHIFT : FLY_EYE_SHIFT2;
    br = y.r >> S; bc = y.c >> S;
    if (br == y.blk[v][0] && bc == y.blk[v][1]) continue;
    const float* tbl = v == 0 ? bmax : bmax + sc.nbr * sc.nbc;
    const int nbc = v == 0 ? sc.nbc : sc.nbc2;
    ce = y.sc_ > 0 ? ((bc + 1) << S) : (bc << S);
    re = y.sr_ > 0 ? ((br + 1) << S) : (br << S);
    txe = d[0] != 0 ? ((-half + (R)ce * cell) - o[0]) * y.idx : big;
    tye = d[1] != 0 ? ((-half + (R)re * cell) - o[1]) * y.idy : big;
    te = txe < tye ? txe : tye;
    const R tz = (te < y.t1 ? te : y.t1) * (1 + R(1e-5)) + R(1e-5);
    const R za = o[2] + y.tin * d[2], zb = o[2] + tz * d[2];
    if (!((za < zb ? za : zb) > (R)tbl[br * nbc + bc] + R(1e-4))) { y.blk[v][0] = br; y.blk[v][1] = bc; continue; }
    lv = v;
  }
  if (lv < 0) return 2;
  if (te >= y.t1) return 1;
  int c, r;
  if (txe < tye) {
    c = y.sc_ > 0 ? ce : ce - 1;
    if (c < 0 || c > sc.ncol - 2) return 1;
    int rr = (int)floor((o[1] + te * d[1] + half) * icell);
    const int rlo = br << S, rhi = ((br + 1) << S) - 1;
    rr = rr < rlo ? rlo : (rr > rhi ? rhi : rr);
    r = rr > sc.nrow - 2 ? sc.nrow - 2 : rr;
  } else {
    r = y.sr_ > 0 ? re : re - 1;
    if (r < 0 || r > sc.nrow - 2) return 1;
    int cc = (int)floor((o[0] + te * d[0] + half) * icell);
    const int clo = bc << S, chi = ((bc + 1) << S) - 1;
    cc = cc < clo ? clo : (cc > chi ? chi : cc);
    c = cc > sc.ncol - 2 ? sc.ncol - 2 : cc;
  }
  y.c = c; y.r = r; y.tin = te;
  y.tx = d[0] != 0 ? ((-half + (c + (y.sc_ > 0 ? 1 : 0)) * cell) - o[0]) * y.idx : big;
  y.ty = d[1] != 0 ? ((-half + (r + (y.sr_ > 0 ? 1 : 0)) * cell) - o[1]) * y.idy : big;
  if (d[2] >= 0 && o[2] + te * d[2] > y.gmax) return 1;
  return 0;
}
template <class R> FLY_INL int eye_ray_cell(const FlyScene& sc, const float* h, EyeRay<R>& y) {
  const R *o = y.o, *d = y.d;
  if (++y.steps > 4 * (sc.nrow + sc.ncol)) return 1;
  const int c = y.c, r = y.r;
  const R tout = y.tx < y.ty ? y.tx : y.ty;
  const R h00 = eye_height<R>(h, sc.ncol, r, c, sc.zoff), h01 = eye_height<R>(h, sc.ncol, r, c + 1, sc.zoff);
  const R h10 = eye_height<R>(h, sc.ncol, r + 1, c, sc.zoff), h11 = eye_height<R>(h, sc.ncol, r + 1, c + 1, sc.zoff);
  const R hmax = fmax_(fmax_(h00, h01), fmax_(h10, h11));
  const R za = o[2] + y.tin * d[2], zb = o[2] + (tout < y.t1 ? tout : y.t1) * d[2];
  if ((za < zb ? za : zb) <= hmax) { y.hc[0] = h00; y.hc[1] = h01; y.hc[2] = h10; y.hc[3] = h11; return 3; }
  if (eye_ray_next_cell(sc, y)) return 1;
  return eye_ray_in_known_blocks(y) ? 0 : 4;
}

/* the two triangles of the candidate cell; returns 2: hit (rgb written), else what eye_ray_next_cell says */
template <class R> FLY_INL int eye_ray_test(const FlyScene& sc, EyeRay<R>& y, R* rgb) {
  const R half = sc.half, cell = 2 * half / (sc.ncol - 1);
  const R *o = y.o, *d = y.d;
  const R x0 = -half + y.c * cell, y0 = -half + y.r * cell;
  const R A[3] = {x0, y0, y.hc[0]}, B[3] = {x0, y0 + cell, y.hc[2]}, C[3] = {x0 + cell, y0, y.hc[1]}, D[3] = {x0 + cell, y0 + cell, y.hc[3]};
  R n1[3], n2[3];
  const R ta = eye_tri(o, d, A, B, C, n1), tb = eye_tri(o, d, B, C, D, n2);
  R t = -1; const R* n = n1;
  if (ta > 0 && (tb <= 0 || ta <= tb)) { t = ta; n = n1; }
  else if (tb > 0) { t = tb; n = n2; }
  if (t > 0) {
    R nn[3] = {n[0], n[1], n[2]};
    normalize3(nn);
    R lam = -(nn[0] * y.fwd[0] + nn[1] * y.fwd[1] + nn[2] * y.fwd[2]);      /* headlight shines along fwd; surfaces are two-sided */
    if (lam < 0) lam = -lam;
    for (int k = 0; k < 3; k++) { const R v = sc.rgb[k] * (sc.ambient + sc.diffuse * lam); rgb[k] = v > 1 ? R(1) : v; }
    return 2;
  }
  return eye_ray_next_cell(sc, y);
}

/* one step of the whole walk; returns 0: go on, 1: left the terrain (background), 2: hit (rgb written) */
template <class R> FLY_INL int eye_ray_step(const FlyScene& sc, const float* h, const float* bmax, EyeRay<R>& y, R* rgb) {
  const int s = eye_ray_advance(sc, h, bmax, y);
  return s == 3 ? eye_ray_test(sc, y, rgb) : s;
}

/* direction and headlight axis of pixel (i, j) of a camera at pos with orientation mat (row-major, columns = camera axes in the world) */
template <class R> FLY_INL void eye_ray_of_pixel(EyeRay<R>& y, const R* pos, const R* mat, int W, int H, R tanhalf, int i, int j) {
  const R dc[3] = {tanhalf * ((R)W / (R)H) * (2 * ((R)j + R(0.5)) / (R)W - 1), tanhalf * (1 - 2 * ((R)i + R(0.5)) / (R)H), R(-1)};
  mat_vec(y.d, mat, dc);
  normalize3(y.d);
  for (int k = 0; k < 3; k++) { y.o[k] = pos[k]; y.fwd[k] = -mat[3 * k + 2]; }
}
FLY_INL void eye_store(unsigned char* out, const float* rgb) {
  for (int k = 0; k < 3; k++) out[k] = (unsigned char)(int)(rgb[k] * 255.0f + 0.5f);
}

/* pixel (i, j) of a camera at pos with orientation mat (row-major, columns = camera axes in the world): the whole ray at once */
template <class R> FLY_INL void eye_pixel(const FlyScene& sc, const float* h, const R* pos, const R* mat, int W, int H, R tanhalf, int i, int j,
                                          unsigned char* out, const float* bmax = nullptr) {
  EyeRay<R> y;
  R rgb[3] = {(R)sc.sky[0], (R)sc.sky[1], (R)sc.sky[2]};
  eye_ray_of_pixel(y, pos, mat, W, H, tanhalf, i, j);
  if (eye_ray_init(sc, y)) {
    eye_ray_set_gmax(sc, bmax, y);
    while (eye_ray_step(sc, h, bmax, y, rgb) == 0) {}
  }
  for (int k = 0; k < 3; k++) out[k] = (unsigned char)(int)(rgb[k] * R(255) + R(0.5));
}
#endif
