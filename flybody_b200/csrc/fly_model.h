/* fly_model.h - device-side model ("DModel") and the host code that derives it from a FlybModel.
 *
 * The FlybModel (include/flyb_model.h) is the flat float64 model the host compiler emits.  The CUDA
 * engine wants more than that: the kinematic tree cut into depth levels (one warp walks a level at a
 * time), the tree-sparse layout of the mass matrix, per-body dof chains for contact Jacobians, descendant
 * lists for gather-style sparse factorisation, and the static collision-pair table with its broad-phase
 * bounds.  build_dmodel() computes all of those once on the host and packs them, converted to the
 * arithmetic type T, into one arena that is copied to HBM in a single cudaMemcpy.
 *
 * Shared by the CUDA library (flyb_cuda.cu) and the host emulation used by the CPU unit tests
 * (tests/emu): both run the same fly_step.cuh source on the same DModel. */
#ifndef FLY_MODEL_H_
#define FLY_MODEL_H_
#include <stdint.h>
#include <string.h>
#include <string>
#include <vector>
#include <algorithm>
#include "../../include/flyb_model.h"

/* Array sizes of the model as the engine source reads them.  In a translation unit compiled for one shipped task model
 * (flyb_kern.cu with FLY_TU_SPEC >= 0) they are that model's compile-time constants: loop bounds the compiler can unroll
 * and schedule across; everywhere else they are the fields of the DModel. */
#ifndef FLY_DIM
#define FLY_DIM(m, f) ((m).f)
#endif

#define FLY_MAXCHAIN 24   /* longest dof chain root->leaf in the Hessian tree (walk: 6 + 14 abdomen = 20) */
#define FLY_MAXCON 64     /* detected contacts per env (gap-band contacts included) */
#define FLY_MAXEFC 160    /* constraint rows per env */
#define FLY_MAXGRP 144    /* constraint groups (limit row | condim-1 contact | 3-row elliptic contact) */
#define FLY_MAXCAND 256   /* narrow-phase candidates after the bounding-sphere test */
#define FLY_MAXCROSS 64   /* two-chain (cross-branch) contact rows */
#define FLY_HEAVY 24      /* dofs with more descendants than this are reduced warp-cooperatively */
#define FLY_MAXTRUNK 8    /* dofs of the trunk chain (held in registers by the factorisation and the solves) */

enum { FLY_OPT_NOCONTACT = 1, FLY_OPT_NOLIMIT = 2, FLY_OPT_NOFLUID = 4, FLY_OPT_NONOSLIP = 8,
       FLY_OPT_NOACTUATION = 16, FLY_OPT_NOWARMSTART = 32 };

/* int32 arrays */
#define FLY_INT_FIELDS(X)                                                                                  \
  X(body_parent) X(body_rootid) X(body_jntadr) X(body_jntnum) X(body_dofadr) X(body_dofnum)               \
  X(body_lastdof) X(body_kchain) X(lvl_adr) X(lvl_body) X(child_adr) X(child)                              \
  X(jnt_type) X(jnt_qposadr) X(jnt_dofadr) X(jnt_bodyid) X(jnt_limited)                                    \
  X(dof_bodyid) X(dof_jntid) X(dof_parentid) X(hpar) X(hdepth) X(kdepth) X(hsublast) X(madr) X(mrow) X(mcol) X(mdiag)\
  X(dlvl_adr) X(dlvl_dof) X(dlvl_nlight) X(heavy_dof) X(desc_adr) X(desc) X(tri_a) X(tri_b)                  \
  X(gstep_b) X(gstep_f) X(dof_din) X(trunk_dof) X(trunk_adr) X(bt_list)                             \
  X(geom_type) X(geom_bodyid) X(pair_geom) X(pair_condim)                                                  \
  X(site_bodyid) X(tendon_adr) X(tendon_num) X(wrap_dofadr) X(wrap_qposadr)                                \
  X(act_trntype) X(act_trnid) X(act_actadr) X(act_ctrllimited) X(act_forcelimited) X(act_other)            \
  X(sensor_type) X(sensor_site) X(sensor_adr) X(sensor_obs)                                                \
  X(fbox_bodyid) X(body_fbox_adr) X(body_fbox) X(fgeom_bodyid)                                             \
  X(ctrl_action) X(obs_joint_qposadr) X(obs_joint_dofadr) X(appendage_site) X(mocap_joint) X(mocap_site)   \
  X(wing_joint) X(wing_action) X(obs_offset)

/* arrays of T */
#define FLY_REAL_FIELDS(X)                                                                                 \
  X(body_pos) X(body_quat) X(body_ipos) X(body_iquat) X(body_mass) X(body_inertia) X(body_invweight0)      \
  X(body_subtreemass) X(jnt_pos) X(jnt_axis) X(jnt_stiffness) X(jnt_range) X(jnt_solref) X(jnt_solimp)     \
  X(qpos0) X(qpos_spring) X(qpos_reset) X(dof_armature) X(dof_damping) X(dof_invweight0)                   \
  X(geom_pos) X(geom_mat) X(geom_size) X(geom_rbound) X(pair_bound) X(pair_friction) X(pair_solref)        \
  X(pair_solimp) X(pair_margin) X(pair_gap)                                                                \
  X(site_pos) X(site_mat) X(site_size) X(wrap_coef)                                                        \
  X(act_gainprm) X(act_biasprm) X(act_dynprm) X(act_ctrlrange) X(act_forcerange) X(act_gear)               \
  X(fbox_pos) X(fbox_mat) X(fbox_size) X(fgeom_pos) X(fgeom_mat) X(fgeom_size) X(fgeom_coef)               \
  X(action_min) X(action_max) X(ghost_offset) X(gravity)

/* byte arrays */
#define FLY_BYTE_FIELDS(X) X(ncommon)

template <class T>
struct DModel {
  /* sizes */
  int nq, nv, nu, na, nbody, njnt, ngeom, npair, nsite, ntendon, nwrap, nsensor, nsensordata, nfluidbox, nfluidgeom;
  int nM, nlevel, ndlevel, ntri, nheavy;
  int nact_other;  /* actuators whose transmission is not a joint (tendon, adhesion body): act_other[] lists them */
  int ntrunk, gsteps;
  int opt_iterations, opt_ls_iterations, opt_noslip_iterations, flags;
  /* task */
  int task_type, n_substeps, future_steps, action_dim, obs_dim, max_episode_steps, time_limit_steps;
  int root_body, root_qposadr, root_dofadr, ghost_qposadr, ghost_dofadr, ghost_body, ball_dofadr, thorax_site;
  int n_obs_joints, n_appendages, n_mocap_joints, n_mocap_sites, user_action, pad_first_obs, inference_mode;
  int clip_width;
  /* vision-guided flight: the terrain geom (nhfield 0 / 1), its grid, the eye sites; the terrain pool in HBM is attached by the
     host after build_dmodel (plain device pointers, not part of the arena) */
  int nhfield, hfield_geom, hfield_nrow, hfield_ncol, eye_site[2], nterrain, trench_stride;
  double hfield_size[4];
  T vision_init_quat[4], vision_target_zaxis[3], vision_ranges[8];
  const float* hf_pool;        /* [nterrain, hfield_nrow, hfield_ncol] heights */
  const double* trench_meta;   /* [nterrain, 3]: points of the centre line (0: no trench), x of the first, x of the last */
  const float* trench_y;       /* [nterrain, trench_stride] centre line y */
  const float* hf_bmax;        /* nullable: per terrain, max(height + geom z, 0) over blocks of 16 x 16 cells [hf_nbr, hf_nbc] (the eye
                                  renderer's table): lets the narrow phase drop a geom that flies above everything under it */
  int hf_nbr, hf_nbc, hf_nbc2;   /* blocks of 16 x 16 cells; columns of the table of 4 x 4 blocks that follows it */
  long long hf_bstride;        /* floats per terrain in hf_bmax */
  T timestep, density, viscosity, tolerance, ls_tolerance, noslip_tolerance, impratio, meaninertia;
  T control_dt, terminal_linvel, terminal_angvel, terminal_com_dist, terminal_qacc, terminal_height;
#define X(n) const int* n;
  FLY_INT_FIELDS(X)
#undef X
#define X(n) const T* n;
  FLY_REAL_FIELDS(X)
#undef X
#define X(n) const unsigned char* n;
  FLY_BYTE_FIELDS(X)
#undef X
};

/* ------------------------------------------------------------------ host: FlybModel -> DModel arena */
struct FlyArena {
  std::vector<char> buf;
  size_t put(const void* src, size_t bytes) {
    size_t off = (buf.size() + 15) & ~(size_t)15;
    buf.resize(off + (bytes ? bytes : 16));
    if (bytes) memcpy(buf.data() + off, src, bytes);
    return off;
  }
};

static inline void fly_quat2mat(const double* q, double* m) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}

/* Builds the model with pointer fields holding OFFSETS into `arena`; call fly_relocate() with the base
 * address the arena finally lives at (host or device).  Returns "" or an error message. */
template <class T>
std::string build_dmodel(const FlybModel* m, int flags, DModel<T>* out, FlyArena* arena, int nlanes = 32) {
  DModel<T>& d = *out;
  memset(&d, 0, sizeof(d));
#define CP(n) d.n = m->n;
#define CP2(n) d.n = m->n;
  CP(nq) CP(nv) CP(nu) CP(na) CP(nbody) CP(njnt) CP(ngeom) CP(npair) CP(nsite) CP(ntendon) CP(nwrap) CP(nsensor)
  CP(nsensordata) CP(nfluidbox) CP(nfluidgeom) CP(opt_iterations) CP(opt_ls_iterations) CP(opt_noslip_iterations)
  CP(task_type) CP(n_substeps) CP(future_steps) CP(action_dim) CP(obs_dim) CP(max_episode_steps) CP(time_limit_steps)
  CP(root_body) CP(root_qposadr) CP(root_dofadr) CP(ghost_qposadr) CP(ghost_dofadr) CP(ball_dofadr) CP(thorax_site)
  CP(n_obs_joints) CP(n_appendages) CP(n_mocap_joints) CP(n_mocap_sites) CP(user_action) CP(pad_first_obs)
  CP(inference_mode)
#undef CP
  d.flags = flags;
  CP2(hfield_geom) CP2(hfield_nrow) CP2(hfield_ncol)
  d.nhfield = m->hfield_geom >= 0 ? 1 : 0;
  d.eye_site[0] = m->eye_site[0]; d.eye_site[1] = m->eye_site[1];
  for (int k = 0; k < 4; k++) { d.hfield_size[k] = m->hfield_size[k]; d.vision_init_quat[k] = (T)m->vision_init_quat[k]; }
  for (int k = 0; k < 3; k++) d.vision_target_zaxis[k] = (T)m->vision_target_zaxis[k];
  for (int k = 0; k < 8; k++) d.vision_ranges[k] = (T)m->vision_ranges[k];
  d.ghost_body = -1;
  for (int j = 0; j < m->njnt; j++) if (m->ghost_qposadr >= 0 && m->jnt_qposadr[j] == m->ghost_qposadr) d.ghost_body = m->jnt_bodyid[j];
  d.timestep = (T)m->opt_timestep; d.density = (T)m->opt_density; d.viscosity = (T)m->opt_viscosity;
  d.tolerance = (T)m->opt_tolerance; d.ls_tolerance = (T)m->opt_ls_tolerance;
  d.noslip_tolerance = (T)m->opt_noslip_tolerance; d.impratio = (T)m->opt_impratio;
  d.meaninertia = (T)m->stat_meaninertia; d.control_dt = (T)m->control_dt;
  d.terminal_linvel = (T)m->terminal_linvel; d.terminal_angvel = (T)m->terminal_angvel;
  d.terminal_com_dist = (T)m->terminal_com_dist; d.terminal_qacc = (T)m->terminal_qacc;
  d.terminal_height = (T)m->terminal_height;
  {
    int nj = m->n_mocap_joints, ns = m->n_mocap_sites;
    d.clip_width = m->task_type == FLYB_TASK_WALK_IMITATION ? (7 + nj) + (6 + nj) + 3 * ns + 4 * nj : 13;
  }
  const int nb = m->nbody, nv = m->nv;
  if (nb > 255 || nv > 255 || m->ngeom > 255) return "model too large for 8-bit ids";

  auto puti = [&](const std::vector<int>& v) { return (const int*)(uintptr_t)arena->put(v.data(), v.size() * sizeof(int)); };
  auto putr = [&](const double* src, size_t n) {
    std::vector<T> t(n);
    for (size_t i = 0; i < n; i++) t[i] = (T)src[i];
    return (const T*)(uintptr_t)arena->put(t.data(), n * sizeof(T));
  };
  auto vi = [](const int32_t* p, size_t n) { return std::vector<int>(p, p + n); };

  /* --- body tree */
  std::vector<int> depth(nb, 0), lastdof(nb, -1), kchain(nb, 0);
  int maxdepth = 0;
  for (int b = 1; b < nb; b++) {
    if (m->body_parent[b] >= b) return "bodies must be in depth-first order";
    if (m->body_dofnum[b] <= 0) return "every non-world body must carry a joint (fuse static bodies on the host)";
    depth[b] = depth[m->body_parent[b]] + 1;
    if (depth[b] > maxdepth) maxdepth = depth[b];
    lastdof[b] = m->body_dofadr[b] + m->body_dofnum[b] - 1;
  }
  d.nlevel = maxdepth;  /* levels 1..maxdepth hold bodies */
  std::vector<int> lvl_adr(maxdepth + 2, 0), lvl_body;
  for (int L = 1; L <= maxdepth; L++) {
    lvl_adr[L] = (int)lvl_body.size();
    for (int b = 1; b < nb; b++) if (depth[b] == L) lvl_body.push_back(b);
  }
  lvl_adr[maxdepth + 1] = (int)lvl_body.size();
  std::vector<int> child_adr(nb + 1, 0), child;
  for (int b = 0; b < nb; b++) {
    child_adr[b] = (int)child.size();
    for (int c = 1; c < nb; c++) if (m->body_parent[c] == b) child.push_back(c);
  }
  child_adr[nb] = (int)child.size();

  /* --- Hessian tree over dofs: the kinematic dof tree, except that for walk_on_ball the ball's dofs are
     made ancestors of every other tree so that leg-ball contact rows stay single-chain (no fill-in). */
  std::vector<int> hpar(m->dof_parentid, m->dof_parentid + nv), hdepth(nv, 0), hsublast(nv, 0), kdepth(nv, 0);
  for (int i = 0; i < nv; i++) kdepth[i] = hpar[i] < 0 ? 0 : kdepth[hpar[i]] + 1;
  if (m->task_type == FLYB_TASK_WALK_ON_BALL && m->ball_dofadr >= 0) {
    int top = m->ball_dofadr + 2;
    if (m->ball_dofadr != 0) return "ball dofs must come first";
    for (int i = top + 1; i < nv; i++) if (hpar[i] < 0) hpar[i] = top;
  }
  for (int i = 0; i < nv; i++) {
    if (hpar[i] >= i) return "dofs must be in depth-first order";
    hdepth[i] = hpar[i] < 0 ? 0 : hdepth[hpar[i]] + 1;
    if (hdepth[i] + 1 > FLY_MAXCHAIN) return "dof chain longer than FLY_MAXCHAIN";
  }
  for (int i = nv - 1; i >= 0; i--) {
    if (hsublast[i] < i) hsublast[i] = i;
    if (hpar[i] >= 0 && hsublast[hpar[i]] < hsublast[i]) hsublast[hpar[i]] = hsublast[i];
  }
  for (int i = 0; i < nv; i++)  /* subtree of i must be the contiguous range [i, hsublast[i]] */
    for (int k = i + 1; k <= hsublast[i]; k++) {
      int a = k; while (a > i) a = hpar[a];
      if (a != i) return "dof subtrees are not contiguous";
    }
  for (int b = 1; b < nb; b++) kchain[b] = kdepth[lastdof[b]] + 1;
  std::vector<int> madr(nv + 1, 0), mrow, mcol;
  for (int i = 0; i < nv; i++) {
    madr[i] = (int)mrow.size();
    for (int j = i; j >= 0; j = hpar[j]) { mrow.push_back(i); mcol.push_back(j); }
  }
  madr[nv] = (int)mrow.size();
  d.nM = (int)mrow.size();
  int maxhd = 0;
  for (int i = 0; i < nv; i++) if (hdepth[i] > maxhd) maxhd = hdepth[i];
  d.ndlevel = maxhd + 1;
  std::vector<int> desc_adr(nv + 1, 0), desc;
  for (int i = 0; i < nv; i++) {
    desc_adr[i] = (int)desc.size();
    /* packed: descendant dof c in bits 0-7, address of the tree-sparse entry (c, i) in bits 8+ */
    for (int k = i + 1; k <= hsublast[i]; k++) desc.push_back(k | ((madr[k] + hdepth[k] - hdepth[i]) << 8));
  }
  desc_adr[nv] = (int)desc.size();
  /* dof levels: within a level the "light" dofs (few descendants) come first, heavy ones after */
  std::vector<int> dlvl_adr(maxhd + 2, 0), dlvl_dof, dlvl_nlight(maxhd + 1, 0);
  for (int L = 0; L <= maxhd; L++) {
    dlvl_adr[L] = (int)dlvl_dof.size();
    for (int i = 0; i < nv; i++) if (hdepth[i] == L && hsublast[i] - i <= FLY_HEAVY) { dlvl_dof.push_back(i); dlvl_nlight[L]++; }
    for (int i = 0; i < nv; i++) if (hdepth[i] == L && hsublast[i] - i > FLY_HEAVY) dlvl_dof.push_back(i);
  }
  dlvl_adr[maxhd + 1] = (int)dlvl_dof.size();
  std::vector<int> heavy_dof;
  for (int i = 0; i < nv; i++) if (hsublast[i] - i > FLY_HEAVY) heavy_dof.push_back(i);
  d.nheavy = (int)heavy_dof.size();
  /* ncommon[a*nv+b]: number of shared entries at the root end of the dof chains starting at a and at b */
  std::vector<unsigned char> ncommon((size_t)nv * nv, 0);
  for (int a = 0; a < nv; a++)
    for (int b = 0; b < nv; b++) {
      int x = a, y = b;
      while (x != y && x >= 0 && y >= 0) { if (x > y) x = hpar[x]; else y = hpar[y]; }
      ncommon[(size_t)a * nv + b] = (x == y && x >= 0) ? (unsigned char)(hdepth[x] + 1) : 0;
    }
  /* ---- branch schedule of the tree-sparse factorisation and solves (fly_dyn.cuh: fly_factor / fly_solve).
     Trunk = the dofs with many descendants (free root, or the ball): a chain at the top of the Hessian tree, kept in
     registers by every lane.  Every other subtree hanging off the trunk is a "branch"; a branch is owned by ONE lane,
     which walks its rows with no barrier inside.  Row descriptor: dof | in-branch depth << 8 | madr << 13. */
  if (d.nM >= 2047) return "tree-sparse matrix too large for the packed row descriptors";
  std::vector<int> trunk;
  for (int i = 0; i < nv; i++) if (hsublast[i] - i > FLY_HEAVY) trunk.push_back(i);
  if ((int)trunk.size() > FLY_MAXTRUNK) return "more than FLY_MAXTRUNK trunk dofs";
  for (size_t t = 0; t < trunk.size(); t++)
    if (hpar[trunk[t]] != (t ? trunk[t - 1] : -1) || hdepth[trunk[t]] != (int)t) return "trunk dofs must form a chain from the root";
  d.ntrunk = (int)trunk.size();
  std::vector<char> is_trunk(nv, 0);
  for (size_t t = 0; t < trunk.size(); t++) is_trunk[trunk[t]] = 1;
  std::vector<int> dof_din(nv, -1);
  for (int i = 0; i < nv; i++) if (!is_trunk[i]) dof_din[i] = (hpar[i] < 0 || is_trunk[hpar[i]]) ? 0 : dof_din[hpar[i]] + 1;
  struct Branch { int root; long cost; };
  std::vector<Branch> branches;
  for (int i = 0; i < nv; i++) if (!is_trunk[i] && dof_din[i] == 0) {
    long cost = 0;  /* multiply-adds of the factorisation inside the branch */
    for (int c = i; c <= hsublast[i]; c++)
      for (int a = hpar[c]; a >= 0 && !is_trunk[a]; a = hpar[a]) cost += hdepth[a] + 1;
    Branch b; b.root = i; b.cost = cost + 1; branches.push_back(b);
  }
  /* Lane groups.  The branches are packed into "groups" (a big branch alone, small ones chained); a group owns a set of
     consecutive lanes and walks its dofs one per step, all groups in step.  Within a step the lanes of a group share the
     entries of the row being eliminated / substituted.  Step word: dof | lane in group << 8 | group size << 13 |
     madr << 18, or -1 for an idle lane. */
  std::vector<std::vector<int> > grp_dofs;   /* dofs of each group, ascending */
  std::vector<long> grp_cost;
  {
    std::vector<int> order(branches.size());
    for (size_t i = 0; i < order.size(); i++) order[i] = (int)i;
    for (size_t a = 0; a < order.size(); a++) for (size_t b = a + 1; b < order.size(); b++) if (branches[order[b]].cost > branches[order[a]].cost) std::swap(order[a], order[b]);
    size_t max_steps = 1;
    for (size_t a = 0; a < branches.size(); a++) { size_t n = (size_t)(hsublast[branches[a].root] - branches[a].root + 1); if (n > max_steps) max_steps = n; }
    for (size_t a = 0; a < order.size(); a++) {
      const Branch& br = branches[order[a]];
      const size_t n = (size_t)(hsublast[br.root] - br.root + 1);
      /* a branch longer than half the longest one gets its own group; smaller ones are chained onto the cheapest group
         that stays within the longest branch's step count (everything is chained once three quarters of the lanes own a group) */
      int into = -1;
      if (n * 2 <= max_steps || (int)grp_dofs.size() >= nlanes * 3 / 4)
        for (size_t g = 0; g < grp_dofs.size(); g++)
          if (grp_dofs[g].size() + n <= max_steps && (into < 0 || grp_cost[g] < grp_cost[into])) into = (int)g;
      if (into < 0) {
        if ((int)grp_dofs.size() < nlanes) { grp_dofs.push_back(std::vector<int>()); grp_cost.push_back(0); into = (int)grp_dofs.size() - 1; }
        else { into = 0; for (size_t g = 1; g < grp_dofs.size(); g++) if (grp_dofs[g].size() < grp_dofs[into].size()) into = (int)g; }
      }
      for (int c = br.root; c <= hsublast[br.root]; c++) grp_dofs[into].push_back(c);
      grp_cost[into] += br.cost;
    }
  }
  const int ngrp = (int)grp_dofs.size();
  std::vector<int> grp_lanes(ngrp, 1);
  for (int spare = nlanes - ngrp; spare > 0; spare--) {  /* remaining lanes go where the per-lane work is largest */
    int best = 0;
    for (int g = 1; g < ngrp; g++) if (grp_cost[g] * grp_lanes[best] > grp_cost[best] * grp_lanes[g]) best = g;
    if (grp_lanes[best] >= 8) {  /* more than 8 lanes on one row does not pay: look for another taker */
      int alt = -1;
      for (int g = 0; g < ngrp; g++) if (grp_lanes[g] < 8 && (alt < 0 || grp_cost[g] * grp_lanes[alt] > grp_cost[alt] * grp_lanes[g])) alt = g;
      if (alt < 0) break;
      best = alt;
    }
    grp_lanes[best]++;
  }
  size_t gsteps = 0;
  for (int g = 0; g < ngrp; g++) { std::sort(grp_dofs[g].begin(), grp_dofs[g].end()); if (grp_dofs[g].size() > gsteps) gsteps = grp_dofs[g].size(); }
  std::vector<int> gstep_f(gsteps * nlanes + 1, -1), gstep_b(gsteps * nlanes + 1, -1);
  {
    int lane0 = 0;
    for (int g = 0; g < ngrp; g++) {
      const size_t n = grp_dofs[g].size();
      for (size_t q = 0; q < n; q++)
        for (int j = 0; j < grp_lanes[g]; j++) {
          const int i = grp_dofs[g][q], word = i | (j << 8) | (grp_lanes[g] << 13) | (madr[i] << 18);
          gstep_f[q * nlanes + lane0 + j] = word; gstep_b[(n - 1 - q) * nlanes + lane0 + j] = word;
        }
      lane0 += grp_lanes[g];
    }
    if (lane0 > nlanes) return "internal: lane groups exceed the lanes of an env";
  }
  d.gsteps = (int)gsteps;
  /* non-trunk dofs: address of the LAST entry of the row (the one towards trunk dof 0) | trunk ancestors << 11 | dof << 15 */
  std::vector<int> bt_list, trunk_adr;
  for (int i = 0; i < nv; i++) if (!is_trunk[i]) bt_list.push_back((madr[i] + hdepth[i]) | ((hdepth[i] - dof_din[i]) << 11) | (i << 15));
  for (size_t t = 0; t < trunk.size(); t++) trunk_adr.push_back(madr[trunk[t]]);
  bt_list.push_back(0); trunk_adr.push_back(0);

  /* triangle enumeration (a <= b < FLY_MAXCHAIN) used to spread a chain row's outer product over lanes */
  std::vector<int> tri_a, tri_b;
  /* b-major so that the pairs with b < len are exactly the first len (len + 1) / 2 entries */
  for (int b = 0; b < FLY_MAXCHAIN; b++) for (int a = 0; a <= b; a++) { tri_a.push_back(a); tri_b.push_back(b); }
  d.ntri = (int)tri_a.size();

  /* --- collision pairs */
  std::vector<int> pair_geom(m->npair);
  std::vector<double> pair_bound(m->npair), geom_mat(9 * (size_t)m->ngeom), site_mat(9 * (size_t)m->nsite);
  for (int p = 0; p < m->npair; p++) {
    int g1 = m->pair_geom1[p], g2 = m->pair_geom2[p];
    /* bit 16: geom 1 is a plane; bit 17: geom 1 is the height field (its narrow phase does the bounds test) */
    pair_geom[p] = g1 | (g2 << 8) | (m->geom_type[g1] == FLYB_GEOM_PLANE ? 1 << 16 : 0) | (m->geom_type[g1] == FLYB_GEOM_HFIELD ? 1 << 17 : 0);
    if (m->geom_type[g2] == FLYB_GEOM_PLANE || m->geom_type[g2] == FLYB_GEOM_HFIELD) return "plane / height field must be geom1 of a pair";
    if (m->geom_type[g1] == FLYB_GEOM_HFIELD && g1 != m->hfield_geom) return "one height field per model";
    pair_bound[p] = (m->geom_type[g1] == FLYB_GEOM_PLANE ? 0.0 : m->geom_rbound[g1]) + m->geom_rbound[g2] + m->pair_margin[p];
  }
  for (int g = 0; g < m->ngeom; g++) fly_quat2mat(m->geom_quat + 4 * g, geom_mat.data() + 9 * g);
  for (int s = 0; s < m->nsite; s++) fly_quat2mat(m->site_quat + 4 * s, site_mat.data() + 9 * s);

  /* --- actuators: joint transmissions must not share a dof (they are scattered lane-parallel) */
  {
    std::vector<int> used(nv, 0);
    for (int a = 0; a < m->nu; a++)
      if (m->act_trntype[a] == FLYB_TRN_JOINT) {
        int dd = m->jnt_dofadr[m->act_trnid[a]];
        if (m->jnt_type[m->act_trnid[a]] != FLYB_JNT_HINGE) return "joint actuators must drive hinges";
        if (used[dd]++) return "two joint actuators share a dof";
      }
  }
  /* --- sensors: observation block and index inside the block */
  std::vector<int> sensor_obs(2 * (size_t)m->nsensor);
  for (int s = 0; s < m->nsensor; s++) {
    int t = m->sensor_type[s], key, idx = 0;
    key = t == FLYB_SENS_ACCELEROMETER ? FLYB_OBS_ACCELEROMETER : t == FLYB_SENS_GYRO ? FLYB_OBS_GYRO
          : t == FLYB_SENS_VELOCIMETER ? FLYB_OBS_VELOCIMETER : t == FLYB_SENS_FORCE ? FLYB_OBS_FORCE : FLYB_OBS_TOUCH;
    for (int s2 = 0; s2 < s; s2++) if (m->sensor_type[s2] == t) idx++;
    sensor_obs[2 * s] = key; sensor_obs[2 * s + 1] = idx;
  }
  /* --- fluid boxes grouped by body */
  std::vector<int> body_fbox_adr(nb + 1, 0), body_fbox;
  for (int b = 0; b < nb; b++) {
    body_fbox_adr[b] = (int)body_fbox.size();
    for (int i = 0; i < m->nfluidbox; i++) if (m->fbox_bodyid[i] == b) body_fbox.push_back(i);
  }
  body_fbox_adr[nb] = (int)body_fbox.size();
  std::vector<double> fbox_mat(9 * (size_t)m->nfluidbox), fgeom_mat(9 * (size_t)m->nfluidgeom);
  for (int i = 0; i < m->nfluidbox; i++) fly_quat2mat(m->fbox_quat + 4 * i, fbox_mat.data() + 9 * i);
  for (int i = 0; i < m->nfluidgeom; i++) fly_quat2mat(m->fgeom_quat + 4 * i, fgeom_mat.data() + 9 * i);

  /* --- pack */
  d.body_parent = puti(vi(m->body_parent, nb)); d.body_rootid = puti(vi(m->body_rootid, nb));
  d.body_jntadr = puti(vi(m->body_jntadr, nb)); d.body_jntnum = puti(vi(m->body_jntnum, nb));
  d.body_dofadr = puti(vi(m->body_dofadr, nb)); d.body_dofnum = puti(vi(m->body_dofnum, nb));
  d.body_lastdof = puti(lastdof); d.body_kchain = puti(kchain);
  d.lvl_adr = puti(lvl_adr); d.lvl_body = puti(lvl_body); d.child_adr = puti(child_adr); d.child = puti(child);
  d.jnt_type = puti(vi(m->jnt_type, m->njnt)); d.jnt_qposadr = puti(vi(m->jnt_qposadr, m->njnt));
  d.jnt_dofadr = puti(vi(m->jnt_dofadr, m->njnt)); d.jnt_bodyid = puti(vi(m->jnt_bodyid, m->njnt));
  d.jnt_limited = puti(vi(m->jnt_limited, m->njnt));
  d.dof_bodyid = puti(vi(m->dof_bodyid, nv)); d.dof_jntid = puti(vi(m->dof_jntid, nv));
  d.dof_parentid = puti(vi(m->dof_parentid, nv));
  d.hpar = puti(hpar); d.hdepth = puti(hdepth); d.kdepth = puti(kdepth); d.hsublast = puti(hsublast);
  d.heavy_dof = puti(heavy_dof);
  d.gstep_b = puti(gstep_b); d.gstep_f = puti(gstep_f); d.dof_din = puti(dof_din); d.trunk_dof = puti(trunk);
  d.trunk_adr = puti(trunk_adr); d.bt_list = puti(bt_list);
  d.ncommon = (const unsigned char*)(uintptr_t)arena->put(ncommon.data(), ncommon.size());
  {
    std::vector<int> mdiag(mcol.size() + 4, 0), mcolp(mcol);
    for (size_t e = 0; e < mcol.size(); e++) mdiag[e] = madr[mcol[e]];  /* address of the diagonal of entry e's column */
    mcolp.resize(mcol.size() + 4, 0);
    d.madr = puti(madr); d.mrow = puti(mrow); d.mcol = puti(mcolp); d.mdiag = puti(mdiag);
  }
  d.dlvl_adr = puti(dlvl_adr); d.dlvl_dof = puti(dlvl_dof); d.dlvl_nlight = puti(dlvl_nlight);
  d.desc_adr = puti(desc_adr); d.desc = puti(desc); d.tri_a = puti(tri_a); d.tri_b = puti(tri_b);
  d.geom_type = puti(vi(m->geom_type, m->ngeom)); d.geom_bodyid = puti(vi(m->geom_bodyid, m->ngeom));
  d.pair_geom = puti(pair_geom); d.pair_condim = puti(vi(m->pair_condim, m->npair));
  d.site_bodyid = puti(vi(m->site_bodyid, m->nsite));
  d.tendon_adr = puti(vi(m->tendon_adr, m->ntendon)); d.tendon_num = puti(vi(m->tendon_num, m->ntendon));
  d.wrap_dofadr = puti(vi(m->wrap_dofadr, m->nwrap)); d.wrap_qposadr = puti(vi(m->wrap_qposadr, m->nwrap));
  d.act_trntype = puti(vi(m->act_trntype, m->nu)); d.act_trnid = puti(vi(m->act_trnid, m->nu));
  {
    std::vector<int> other;
    for (int a = 0; a < m->nu; a++) if (m->act_trntype[a] != FLYB_TRN_JOINT) other.push_back(a);
    d.nact_other = (int)other.size();
    if (other.empty()) other.push_back(0);
    d.act_other = puti(other);
  }
  d.act_actadr = puti(vi(m->act_actadr, m->nu)); d.act_ctrllimited = puti(vi(m->act_ctrllimited, m->nu));
  d.act_forcelimited = puti(vi(m->act_forcelimited, m->nu));
  d.sensor_type = puti(vi(m->sensor_type, m->nsensor)); d.sensor_site = puti(vi(m->sensor_site, m->nsensor));
  d.sensor_adr = puti(vi(m->sensor_adr, m->nsensor)); d.sensor_obs = puti(sensor_obs);
  d.fbox_bodyid = puti(vi(m->fbox_bodyid, m->nfluidbox)); d.body_fbox_adr = puti(body_fbox_adr);
  d.body_fbox = puti(body_fbox); d.fgeom_bodyid = puti(vi(m->fgeom_bodyid, m->nfluidgeom));
  d.ctrl_action = puti(vi(m->ctrl_action, m->nu));
  d.obs_joint_qposadr = puti(vi(m->obs_joint_qposadr, m->n_obs_joints));
  d.obs_joint_dofadr = puti(vi(m->obs_joint_dofadr, m->n_obs_joints));
  d.appendage_site = puti(vi(m->appendage_site, m->n_appendages));
  d.mocap_joint = puti(vi(m->mocap_joint, m->n_mocap_joints)); d.mocap_site = puti(vi(m->mocap_site, m->n_mocap_sites));
  d.wing_joint = puti(vi(m->wing_joint, 6)); d.wing_action = puti(vi(m->wing_action, 6));
  d.obs_offset = puti(vi(m->obs_offset, 16));

  d.body_pos = putr(m->body_pos, 3 * nb); d.body_quat = putr(m->body_quat, 4 * nb);
  d.body_ipos = putr(m->body_ipos, 3 * nb); d.body_iquat = putr(m->body_iquat, 4 * nb);
  d.body_mass = putr(m->body_mass, nb); d.body_inertia = putr(m->body_inertia, 3 * nb);
  d.body_invweight0 = putr(m->body_invweight0, 2 * nb); d.body_subtreemass = putr(m->body_subtreemass, nb);
  d.jnt_pos = putr(m->jnt_pos, 3 * m->njnt); d.jnt_axis = putr(m->jnt_axis, 3 * m->njnt);
  d.jnt_stiffness = putr(m->jnt_stiffness, m->njnt); d.jnt_range = putr(m->jnt_range, 2 * m->njnt);
  d.jnt_solref = putr(m->jnt_solref, 2 * m->njnt); d.jnt_solimp = putr(m->jnt_solimp, 5 * m->njnt);
  d.qpos0 = putr(m->qpos0, m->nq); d.qpos_spring = putr(m->qpos_spring, m->nq); d.qpos_reset = putr(m->qpos_reset, m->nq);
  d.dof_armature = putr(m->dof_armature, nv); d.dof_damping = putr(m->dof_damping, nv);
  d.dof_invweight0 = putr(m->dof_invweight0, nv);
  d.geom_pos = putr(m->geom_pos, 3 * m->ngeom); d.geom_mat = putr(geom_mat.data(), geom_mat.size());
  d.geom_size = putr(m->geom_size, 3 * m->ngeom); d.geom_rbound = putr(m->geom_rbound, m->ngeom);
  d.pair_bound = putr(pair_bound.data(), pair_bound.size());
  d.pair_friction = putr(m->pair_friction, 3 * m->npair); d.pair_solref = putr(m->pair_solref, 2 * m->npair);
  d.pair_solimp = putr(m->pair_solimp, 5 * m->npair); d.pair_margin = putr(m->pair_margin, m->npair);
  d.pair_gap = putr(m->pair_gap, m->npair);
  d.site_pos = putr(m->site_pos, 3 * m->nsite); d.site_mat = putr(site_mat.data(), site_mat.size());
  d.site_size = putr(m->site_size, 3 * m->nsite); d.wrap_coef = putr(m->wrap_coef, m->nwrap);
  d.act_gainprm = putr(m->act_gainprm, m->nu); d.act_biasprm = putr(m->act_biasprm, 3 * m->nu);
  d.act_dynprm = putr(m->act_dynprm, m->nu); d.act_ctrlrange = putr(m->act_ctrlrange, 2 * m->nu);
  d.act_forcerange = putr(m->act_forcerange, 2 * m->nu); d.act_gear = putr(m->act_gear, m->nu);
  d.fbox_pos = putr(m->fbox_pos, 3 * m->nfluidbox); d.fbox_mat = putr(fbox_mat.data(), fbox_mat.size());
  d.fbox_size = putr(m->fbox_size, 3 * m->nfluidbox);
  d.fgeom_pos = putr(m->fgeom_pos, 3 * m->nfluidgeom); d.fgeom_mat = putr(fgeom_mat.data(), fgeom_mat.size());
  d.fgeom_size = putr(m->fgeom_size, 3 * m->nfluidgeom); d.fgeom_coef = putr(m->fgeom_coef, 12 * m->nfluidgeom);
  d.action_min = putr(m->action_min, m->action_dim); d.action_max = putr(m->action_max, m->action_dim);
  d.ghost_offset = putr(m->ghost_offset, 3); d.gravity = putr(m->opt_gravity, 3);
  return "";
}

template <class T>
void fly_relocate(DModel<T>* d, const void* base) {
  uintptr_t b = (uintptr_t)base;
#define X(n) d->n = (const int*)(b + (uintptr_t)d->n);
  FLY_INT_FIELDS(X)
#undef X
#define X(n) d->n = (const T*)(b + (uintptr_t)d->n);
  FLY_REAL_FIELDS(X)
#undef X
#define X(n) d->n = (const unsigned char*)(b + (uintptr_t)d->n);
  FLY_BYTE_FIELDS(X)
#undef X
}
#endif  /* FLY_MODEL_H_ */
