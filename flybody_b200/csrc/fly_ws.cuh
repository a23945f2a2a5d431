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

#define FLY_FAST 1
#define FLY_BULK 0
#define FLY_NHIT 4   /* most contacts a primitive pair can produce */

/* X(name, element-count expression over `m`, tier) ; arrays of T */
#define FLY_WS_REAL(X)                                                                                       \
  X(qpos, m.nq, FLY_FAST) X(qvel, m.nv, FLY_FAST) X(act, m.na, FLY_FAST) X(ctrl, m.nu, FLY_FAST)             \
  X(warm, m.nv, FLY_FAST) X(qacc, m.nv, FLY_FAST)                                                            \
  X(xpos, 3 * m.nbody, FLY_FAST) X(xquat, 4 * m.nbody, FLY_FAST) X(xmat, 9 * m.nbody, FLY_FAST)              \
  X(xipos, 3 * m.nbody, FLY_BULK) X(ximat, 9 * m.nbody, FLY_BULK)                                            \
  X(xanchor, 3 * m.njnt, FLY_BULK) X(xaxis, 3 * m.njnt, FLY_BULK)                                            \
  X(geom_xpos, 3 * m.ngeom, FLY_FAST) X(site_xpos, 3 * m.nsite, FLY_BULK) X(site_xmat, 9 * m.nsite, FLY_BULK)\
  X(scom, 3 * m.nbody, FLY_FAST) X(cinert, 10 * m.nbody, FLY_BULK) X(crb, 10 * m.nbody, FLY_BULK)            \
  X(cdof, 6 * m.nv, FLY_FAST) X(cdof_dot, 6 * m.nv, FLY_BULK) X(cvel, 6 * m.nbody, FLY_FAST)                 \
  X(qM, m.nM, FLY_FAST) X(qLD, m.nM, FLY_FAST) X(qH, m.nM, FLY_FAST)                                         \
  X(act_length, m.nu, FLY_BULK) X(act_force, m.nu, FLY_BULK) X(act_dot, m.na, FLY_BULK)                      \
  X(qfrc_bias, m.nv, FLY_BULK) X(qfrc_passive, m.nv, FLY_BULK) X(qfrc_actuator, m.nv, FLY_BULK)              \
  X(qfrc_smooth, m.nv, FLY_FAST) X(qacc_smooth, m.nv, FLY_FAST) X(qfrc_constraint, m.nv, FLY_FAST)           \
  X(sMa, m.nv, FLY_FAST) X(sgrad, m.nv, FLY_FAST) X(ssearch, m.nv, FLY_FAST) X(sMv, m.nv, FLY_FAST)          \
  X(stmp, m.nv, FLY_FAST)                                                                                    \
  X(bodyf, 6 * m.nbody, FLY_BULK) X(boxf, 6 * (m.nfluidbox + m.nfluidgeom), FLY_BULK)                        \
  X(cacc, 6 * m.nbody, FLY_BULK) X(cfrc_int, 6 * m.nbody, FLY_BULK) X(cfrc_ext, 6 * m.nbody, FLY_BULK)       \
  X(sensordata, m.nsensordata, FLY_BULK) X(sens_sum, m.nsensordata, FLY_BULK)                                \
  X(con_dist, FLY_MAXCON, FLY_BULK) X(con_incl, FLY_MAXCON, FLY_BULK) X(con_mu, FLY_MAXCON, FLY_BULK)        \
  X(con_fric, FLY_MAXCON, FLY_BULK) X(con_pos, 3 * FLY_MAXCON, FLY_BULK) X(con_frame, 9 * FLY_MAXCON, FLY_BULK)\
  X(con_hc, 9 * FLY_MAXCON, FLY_BULK) X(hit_tmp, 32 * FLY_NHIT * 11, FLY_BULK)                               \
  X(efc_J, FLY_MAXEFC * FLY_MAXCHAIN, FLY_BULK) X(efc_Y, FLY_MAXEFC * FLY_MAXCHAIN, FLY_BULK)                \
  X(x_J, FLY_MAXCROSS * FLY_MAXCHAIN, FLY_BULK) X(x_Y, FLY_MAXCROSS * FLY_MAXCHAIN, FLY_BULK)                \
  X(x_dense, FLY_MAXCROSS * m.nv, FLY_BULK) X(x_S, FLY_MAXCROSS * FLY_MAXCROSS, FLY_BULK)                    \
  X(x_z, FLY_MAXCROSS, FLY_BULK)                                                                             \
  X(efc_pos, FLY_MAXEFC, FLY_BULK) X(efc_margin, FLY_MAXEFC, FLY_BULK) X(efc_diag, FLY_MAXEFC, FLY_BULK)     \
  X(efc_R, FLY_MAXEFC, FLY_BULK) X(efc_D, FLY_MAXEFC, FLY_FAST) X(efc_K, FLY_MAXEFC, FLY_BULK)               \
  X(efc_B, FLY_MAXEFC, FLY_BULK) X(efc_imp, FLY_MAXEFC, FLY_BULK) X(efc_aref, FLY_MAXEFC, FLY_BULK)          \
  X(efc_force, FLY_MAXEFC, FLY_FAST) X(efc_jar, FLY_MAXEFC, FLY_FAST) X(efc_jv, FLY_MAXEFC, FLY_FAST)        \
  X(efc_b, FLY_MAXEFC, FLY_BULK) X(ns_A, FLY_MAXEFC * FLY_MAXEFC * 2 / 3, FLY_BULK)                              \
  X(obs_tmp, 8, FLY_BULK)

/* arrays of int */
#define FLY_WS_INT(X)                                                                                        \
  X(con_pair, FLY_MAXCON, FLY_BULK) X(con_dim, FLY_MAXCON, FLY_BULK) X(con_excl, FLY_MAXCON, FLY_BULK)       \
  X(con_efc, FLY_MAXCON, FLY_BULK) X(hit_cnt, 32, FLY_BULK) X(hit_off, 33, FLY_BULK)                         \
  X(efc_start, FLY_MAXEFC, FLY_BULK) X(efc_xid, FLY_MAXEFC, FLY_BULK) X(efc_state, FLY_MAXEFC, FLY_BULK)     \
  X(grp_type, FLY_MAXGRP, FLY_BULK) X(grp_row, FLY_MAXGRP, FLY_BULK) X(grp_con, FLY_MAXGRP, FLY_BULK)        \
  X(x_start, FLY_MAXCROSS, FLY_BULK) X(x_row, FLY_MAXCROSS, FLY_BULK) X(ns_fr, FLY_MAXEFC, FLY_BULK) X(ns_con, FLY_MAXCON, FLY_BULK)         \
  X(counts, 16, FLY_BULK)

/* arrays of unsigned short */
#define FLY_WS_U16(X) X(cand, FLY_MAXCAND, FLY_BULK) X(lim_item, FLY_MAXEFC, FLY_BULK)

/* indices into counts[] */
enum { FLY_C_NCON = 0, FLY_C_NEFC, FLY_C_NLIMIT, FLY_C_NGRP, FLY_C_NCAND, FLY_C_NITER, FLY_C_OVERFLOW, FLY_C_NCROSS,
       FLY_C_NFR, FLY_C_NCROSSROW };

template <class T>
struct Ws {
#define X(n, cnt, tier) T* n;
  FLY_WS_REAL(X)
#undef X
#define X(n, cnt, tier) int* n;
  FLY_WS_INT(X)
#undef X
#define X(n, cnt, tier) unsigned short* n;
  FLY_WS_U16(X)
#undef X
};

#if defined(__CUDACC__)
#define FLY_HD __host__ __device__ inline
#else
#define FLY_HD inline
#endif

/* Carves the two arenas.  Pass null bases to only measure; sizes (bytes, 16-aligned) are returned. */
template <class T>
FLY_HD void ws_bind(const DModel<T>& m, Ws<T>* w, char* fast, char* bulk, size_t* fast_bytes, size_t* bulk_bytes,
                    bool all_bulk) {
  size_t off[2] = {0, 0};
  char* base[2] = {bulk, fast};
#define PLACE(n, cnt, tier, type)                                                    \
  {                                                                                  \
    int t = all_bulk ? FLY_BULK : (tier);                                            \
    size_t bytes = (((size_t)(cnt) * sizeof(type)) + 15) & ~(size_t)15;              \
    if (w) w->n = base[t] ? (type*)(base[t] + off[t]) : (type*)0;                    \
    off[t] += bytes;                                                                 \
  }
#define X(n, cnt, tier) PLACE(n, cnt, tier, T)
  FLY_WS_REAL(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, int)
  FLY_WS_INT(X)
#undef X
#define X(n, cnt, tier) PLACE(n, cnt, tier, unsigned short)
  FLY_WS_U16(X)
#undef X
#undef PLACE
  if (fast_bytes) *fast_bytes = off[FLY_FAST];
  if (bulk_bytes) *bulk_bytes = off[FLY_BULK];
}
#endif
