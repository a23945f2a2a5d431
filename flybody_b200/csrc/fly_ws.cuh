/* fly_ws.cuh - per-environment workspace of one warp.
 *
 * Every intermediate of the step (positions, spatial quantities, tree-sparse matrices, contact and
 * constraint rows, solver vectors) lives in one of two per-warp arenas: `fast` (shared memory on the GPU)
 * and `bulk` (a global-memory slot owned by the resident warp, hot in L1/L2).  The table below is the single
 * place that decides which array goes where; ws_bind() walks it to carve both arenas. */
#ifndef FLY_WS_CUH_
#define FLY_WS_CUH_
#include "fly_model.h"
#include "fly_math.cuh"

/* tier of an array: 0 = always bulk; 1..6 = goes to the fast arena when the launch's `level` is >= the tier
 * (1: the factorisation of M; 2: the vectors its solves run on; 3: the factorisation of M + dt B; 4: state and solver
 * vectors; 5, 6: kinematics).  `nl` (lanes per env) is part of the signatures for arrays sized per lane. */
#if defined(__CUDACC__)
#define FLY_HD __host__ __device__ inline
#else
#define FLY_HD inline
#endif
#define FLY_NHIT 4   /* most contacts a primitive pair can produce */

/* X(name, element-count expression over `m`, tier) ; arrays of T */
#define FLY_WS_REAL(X)                                                                                       \
  X(qpos, m.nq, 4) X(qvel, m.nv, 4) X(act, m.na, 4) X(ctrl, m.nu, 4)             \
  X(warm, m.nv, 4) X(qacc, m.nv, 4)                                                            \
  X(xpos, 3 * m.nbody, 5) X(xquat, 4 * m.nbody, 5) X(xmat, 9 * m.nbody, 5)              \
  X(xipos, 3 * m.nbody, 6)                                            \
  X(xanchor, 3 * m.njnt, 0) X(xaxis, 3 * m.njnt, 0)                                            \
  X(geom_xpos, 3 * m.ngeom, 5) X(site_xpos, 3 * m.nsite, 0) X(site_xmat, 9 * m.nsite, 0)\
  X(scom, 3 * m.nbody, 5) X(cinert, 10 * m.nbody, 6) X(crb, 0, 0)            \
  X(cdof, 6 * m.nv, 5) X(cdof_dot, 6 * m.nv, 6) X(cvel, 6 * m.nbody, 6)                 \
  X(qLD, m.nM + 1, 1) X(qH, m.nM + 1, 3)                    \
  X(qDinv, m.nv, 1) X(qHinv, m.nv, 3) X(gpart, 64, 0)                                         \
  X(act_length, m.nu, 0) X(act_force, m.nu, 0) X(act_dot, m.na, 0)                      \
  X(qfrc_bias, m.nv, 0) X(qfrc_passive, m.nv, 0) X(qfrc_actuator, m.nv, 0)              \
  X(qfrc_smooth, m.nv, 4) X(qacc_smooth, m.nv, 2) X(qfrc_constraint, m.nv, 4)           \
  X(sMa, m.nv, 2) X(sgrad, m.nv, 4) X(ssearch, m.nv, 4)          \
  X(stmp, m.nv, 2)                                                                                    \
  X(bodyf, 0, 0) X(boxf, 6 * (m.nfluidbox + m.nfluidgeom), 0)                        \
  X(cacc, 6 * m.nbody, 0) X(cfrc_int, 6 * m.nbody, 0) X(cfrc_ext, 6 * m.nbody, 0)       \
  X(sensordata, m.nsensordata, 0) X(sens_sum, m.nsensordata, 0)                                \
  X(con_dist, FLY_MAXCON, 0) X(con_incl, FLY_MAXCON, 0) X(con_mu, FLY_MAXCON, 0)        \
  X(con_fric, FLY_MAXCON, 0) X(con_pos, 3 * FLY_MAXCON, 0) X(con_frame, 9 * FLY_MAXCON, 0)\
  X(con_hc, 9 * FLY_MAXCON, 0) X(hit_tmp, 32 * FLY_NHIT * 11, 0)                               \
  X(efc_J, FLY_MAXEFC * FLY_MAXCHAIN, 0) X(efc_Y, FLY_MAXEFC * FLY_MAXCHAIN, 0)                \
  X(x_J, FLY_MAXCROSS * FLY_MAXCHAIN, 0) X(x_Y, FLY_MAXCROSS * FLY_MAXCHAIN, 0)                \
  X(efc_pos, FLY_MAXEFC, 0) X(efc_margin, FLY_MAXEFC, 0) X(efc_diag, FLY_MAXEFC, 0)     \
  X(efc_R, FLY_MAXEFC, 0) X(efc_D, FLY_MAXEFC, 4) X(efc_K0, FLY_MAXEFC, 0)               \
  X(efc_B, FLY_MAXEFC, 0) X(efc_imp, FLY_MAXEFC, 0) X(efc_aref, FLY_MAXEFC, 0)          \
  X(efc_force, FLY_MAXEFC, 4)        \
  X(efc_A, FLY_MAXEFC * FLY_MAXEFC, 0) X(efc_K, FLY_MAXEFC * FLY_MAXEFC, 0)             \
          \
  X(efc_lam, FLY_MAXEFC, 0)                              \
  X(obs_tmp, 8, 0)

/* arrays of double: the Newton iterate in constraint space (c, jar = b + A c, ...) cancels to a few parts in 1e4
 * of its terms, so these nefc-sized vectors are kept in float64 even when T is float */
#define FLY_WS_DBL(X)                                                                                        \
  X(efc_jar, FLY_MAXEFC, 4) X(efc_jv, FLY_MAXEFC, 4) X(efc_b, FLY_MAXEFC, 0) X(efc_b0, FLY_MAXEFC, 0)        \
  X(efc_c, FLY_MAXEFC, 0) X(efc_cp, FLY_MAXEFC, 0) X(efc_Ac, FLY_MAXEFC, 0) X(efc_Acp, FLY_MAXEFC, 0)        \
  X(efc_h, FLY_MAXEFC, 0)

/* arrays of int */
#define FLY_WS_INT(X)                                                                                        \
  X(con_pair, FLY_MAXCON, 0) X(con_dim, FLY_MAXCON, 0) X(con_excl, FLY_MAXCON, 0)       \
  X(con_efc, FLY_MAXCON, 0) X(hit_cnt, 32, 0) X(hit_off, 33, 0)                         \
  X(efc_start, FLY_MAXEFC, 0) X(efc_g, FLY_MAXEFC, 0) X(efc_xid, FLY_MAXEFC, 0) X(efc_state, FLY_MAXEFC, 0)     \
  X(grp_type, FLY_MAXGRP, 0) X(grp_row, FLY_MAXGRP, 0) X(grp_con, FLY_MAXGRP, 0)        \
  X(x_start, FLY_MAXCROSS, 0) X(x_row, FLY_MAXCROSS, 0) X(ns_fr, FLY_MAXEFC, 0) X(ns_con, FLY_MAXCON, 0)         \
  X(counts, 16, 0)

/* arrays of unsigned short */
#define FLY_WS_U16(X) X(cand, FLY_MAXCAND, 0) X(lim_item, FLY_MAXEFC, 0)

/* indices into counts[] */
enum { FLY_C_NCON = 0, FLY_C_NEFC, FLY_C_NLIMIT, FLY_C_NGRP, FLY_C_NCAND, FLY_C_NITER, FLY_C_OVERFLOW, FLY_C_NCROSS,
       FLY_C_NFR, FLY_C_NCROSSROW, FLY_C_WORK, FLY_C_TERRAIN /* index of the env's terrain in the pool */ };

/* Arrays that only live inside one stage share storage with arrays of another stage (the bulk footprint of all
 * resident warps has to fit in L2):  crb (CRB pass of step1, 10 per body) lies over cacc + cfrc_int (6 + 6 per body),
 * bodyf (fly_passive) over cfrc_ext; those three are rebuilt by every RNE pass that uses them. */
#define FLY_WS_ALIASES(w) \
  (w).crb = (w).cacc;     \
  (w).bodyf = (w).cfrc_ext;

template <class T>
struct Ws {
#define X(n, cnt, tier) T* n;
  FLY_WS_REAL(X)
#undef X
#define X(n, cnt, tier) double* n;
  FLY_WS_DBL(X)
#undef X
#define X(n, cnt, tier) int* n;
  FLY_WS_INT(X)
#undef X
#define X(n, cnt, tier) unsigned short* n;
  FLY_WS_U16(X)
#undef X
};


/* Carves the two arenas.  Pass null bases to only measure; sizes (bytes, 16-aligned) are returned. */
template <class T>
FLY_HD void ws_bind(const DModel<T>& m, Ws<T>* w, char* fast, char* bulk, size_t* fast_bytes, size_t* bulk_bytes,
                    int level, int nl = 32) {
  size_t off[2] = {0, 0};
  char* base[2] = {bulk, fast};
#define PLACE(n, cnt, tier, type)                                                    \
  {                                                                                  \
    int t = ((tier) > 0 && (tier) <= level) ? 1 : 0;                                            \
    size_t bytes = (((size_t)(cnt) * sizeof(type)) + 15) & ~(size_t)15;              \
    if (w) w->n = base[t] ? (type*)(base[t] + off[t]) : (type*)0;                    \
    off[t] += bytes;                                                                 \
  }
#define X(n, cnt, tier) PLACE(n, cnt, tier, T)
  FLY_WS_REAL(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, double)
  FLY_WS_DBL(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, int)
  FLY_WS_INT(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, unsigned short)
  FLY_WS_U16(X)
#undef X
#undef PLACE
  if (w) { FLY_WS_ALIASES(*w) }
  if (fast_bytes) *fast_bytes = off[1];
  if (bulk_bytes) *bulk_bytes = off[0];
}

/* Byte offset of every array inside its arena.  The host fills it once per launch configuration (ws_layout) and
 * passes it as a kernel parameter; each thread then forms its pointers as arena base + constant-bank offset
 * (ws_attach), with the arena of every array fixed at compile time by LEVEL so that the compiler can tell shared
 * from global addresses. */
struct WsOff {
#define X(n, cnt, tier) unsigned n;
  FLY_WS_REAL(X) FLY_WS_DBL(X) FLY_WS_INT(X) FLY_WS_U16(X)
#undef X
};

template <class T>
inline void ws_layout(const DModel<T>& m, int level, WsOff* o, size_t* fast_bytes, size_t* bulk_bytes, int nl = 32) {
  size_t off[2] = {0, 0};
#define PLACE(n, cnt, tier, type)                                          \
  {                                                                        \
    int t = ((tier) > 0 && (tier) <= level) ? 1 : 0;                       \
    size_t bytes = (((size_t)(cnt) * sizeof(type)) + 15) & ~(size_t)15;    \
    o->n = (unsigned)off[t];                                               \
    off[t] += bytes;                                                       \
  }
#define X(n, cnt, tier) PLACE(n, cnt, tier, T)
  FLY_WS_REAL(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, double)
  FLY_WS_DBL(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, int)
  FLY_WS_INT(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, unsigned short)
  FLY_WS_U16(X)
#undef X
#undef PLACE
  if (fast_bytes) *fast_bytes = off[1];
  if (bulk_bytes) *bulk_bytes = off[0];
}

/* The sizes the layout depends on.  For the shipped task models they are known when the library is compiled
 * (fly_dims_gen.h), and the step kernel instantiated for such a model takes its offsets from ws_layout_c at compile time:
 * every array address is then arena base + immediate (measured on the generic kernel: 10.6 % of all executed instructions
 * only added a constant-bank offset to an arena base). */
struct WsDims {
  int nq, nv, na, nu, nbody, njnt, ngeom, nsite, nM, nsensordata, nfluidbox, nfluidgeom;  /* the layout depends on these */
  int npair, nsensor, ntrunk, gsteps, nlevel;  /* further loop bounds the per-model kernels read as constants (FLY_DIM) */
  int nhfield;                                 /* 1: the model has a height-field terrain (compiles the prism walk in) */
};
template <class M> inline WsDims ws_dims_of(const M& m) {
  return WsDims{m.nq, m.nv, m.na, m.nu, m.nbody, m.njnt, m.ngeom, m.nsite, m.nM, m.nsensordata, m.nfluidbox, m.nfluidgeom,
                m.npair, m.nsensor, m.ntrunk, m.gsteps, m.nlevel, m.nhfield};
}
inline bool ws_dims_equal(const WsDims& a, const WsDims& b) {
  return a.nq == b.nq && a.nv == b.nv && a.na == b.na && a.nu == b.nu && a.nbody == b.nbody && a.njnt == b.njnt &&
         a.ngeom == b.ngeom && a.nsite == b.nsite && a.nM == b.nM && a.nsensordata == b.nsensordata &&
         a.nfluidbox == b.nfluidbox && a.nfluidgeom == b.nfluidgeom && a.npair == b.npair && a.nsensor == b.nsensor &&
         a.ntrunk == b.ntrunk && a.gsteps == b.gsteps && a.nlevel == b.nlevel && a.nhfield == b.nhfield;
}
template <class T>
FLY_HD constexpr WsOff ws_layout_c(const WsDims m, int level, int nl = 32) {
  WsOff o{};
  size_t off[2] = {0, 0};
#define PLACE(n, cnt, tier, type)                                          \
  {                                                                        \
    int t = ((tier) > 0 && (tier) <= level) ? 1 : 0;                       \
    size_t bytes = (((size_t)(cnt) * sizeof(type)) + 15) & ~(size_t)15;    \
    o.n = (unsigned)off[t];                                                \
    off[t] += bytes;                                                       \
  }
#define X(n, cnt, tier) PLACE(n, cnt, tier, T)
  FLY_WS_REAL(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, double)
  FLY_WS_DBL(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, int)
  FLY_WS_INT(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, unsigned short)
  FLY_WS_U16(X)
#undef X
#undef PLACE
  return o;
}

#if defined(FLY_GPU)
template <class T, int LEVEL>
FLY_INL void ws_attach(Ws<T>& w, const WsOff& o, char* fast, char* bulk) {
#define X(n, cnt, tier) w.n = (T*)((((tier) > 0 && (tier) <= LEVEL) ? fast : bulk) + o.n);
  FLY_WS_REAL(X)
#undef X
#define X(n, cnt, tier) w.n = (double*)((((tier) > 0 && (tier) <= LEVEL) ? fast : bulk) + o.n);
  FLY_WS_DBL(X)
#undef X
#define X(n, cnt, tier) w.n = (int*)((((tier) > 0 && (tier) <= LEVEL) ? fast : bulk) + o.n);
  FLY_WS_INT(X)
#undef X
#define X(n, cnt, tier) w.n = (unsigned short*)((((tier) > 0 && (tier) <= LEVEL) ? fast : bulk) + o.n);
  FLY_WS_U16(X)
#undef X
  FLY_WS_ALIASES(w)
}
#endif
#endif
