/* flyb_emu.cpp - TEST INFRASTRUCTURE: runs the engine source (flybody_b200/csrc/fly_*.cuh) on the CPU.
 *
 * With FLY_EMU defined a warp is a single thread, so the very code the CUDA kernels are built from can be
 * unit-tested here (no GPU in the build container) against the float64 oracle, in float and in double.
 * Never linked into, nor reachable from, the product library. */
#define FLY_EMU 1
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../flybody_b200/csrc/fly_env.cuh"
#include "../../flybody_b200/csrc/fly_hfield.cuh"

struct EmuBase {
  virtual ~EmuBase() {}
  virtual void set_clips(const float* data, const int* off, const int* len, int n) = 0;
  virtual void set_wbpg(const double* freqs, const int* off, const double* traj, const double* phase, int nf, const double* prm) = 0;
  virtual void set_phase(double p) = 0;
  virtual void set_terrain(const float* data, int n, const double* meta, const float* ty, int stride) = 0;
  virtual void set_seed(unsigned long long seed, int env_offset) = 0;
  virtual void vision_draws(int episode, int* terrain, double* height, double* speed) = 0;
  virtual void eyes(float* out24, int* terrain) = 0;
  virtual void wbpg_state(int* step, int* fidx, double* freq) = 0;
  virtual int env_reset(float* obs) = 0;
  virtual int env_step(const float* action, float* obs, double* reward, double* discount) = 0;
  virtual void forward() = 0;
  virtual void physics_step() = 0;
  virtual int get(const char* name, double* out, int cap) = 0;
  virtual int set(const char* name, const double* in, int n) = 0;
  virtual int geti(const char* name, int* out, int cap) = 0;
  virtual void call(const char* stage) = 0;
};

template <class T>
struct Emu : EmuBase {
  DModel<T> m;
  FlyArena arena;
  Ws<T> w;
  std::vector<char> mem;
  std::vector<T> sreal;
  std::vector<int> sint;
  std::vector<float> clip_data;
  std::vector<int> clip_off, clip_len;
  FlyClips clips;
  FlyWbpg wb;
  std::vector<double> wb_freqs, wb_traj, wb_phase;
  std::vector<int> wb_off;
  double phase0 = 0;
  std::vector<float> hf, trench_y, eye_out = std::vector<float>(24, 0.f);
  std::vector<double> trench_meta;
  unsigned long long seed = 0;
  int env_offset = 0, eye_terrain = 0;
  std::string err;

  Emu(const FlybModel* fm, int flags) {
    err = build_dmodel<T>(fm, flags, &m, &arena);
    if (!err.empty()) return;
    fly_relocate(&m, arena.buf.data());
    size_t fb = 0, bb = 0;
    ws_bind<T>(m, (Ws<T>*)0, 0, 0, &fb, &bb, 0);
    mem.assign(bb + 64, 0);
    ws_bind<T>(m, &w, 0, mem.data(), &fb, &bb, 0);
    sreal.assign(m.nq + 2 * m.nv + m.na + 4, 0);
    sint.assign(FLY_SI_COUNT, 0);
    sint[FLY_SI_FORCED_CLIP] = -1;
    clips.data = 0; clips.offset = 0; clips.length = 0; clips.nclips = 0;
    memset(&wb, 0, sizeof(wb));
    for (int i = 0; i < m.nq; i++) w.qpos[i] = m.qpos0[i];
  }
  void set_clips(const float* data, const int* off, const int* len, int n) override {
    int rows = 0;
    for (int i = 0; i < n; i++) rows = off[i] + len[i] > rows ? off[i] + len[i] : rows;
    clip_data.assign(data, data + (size_t)rows * m.clip_width);
    clip_off.assign(off, off + n); clip_len.assign(len, len + n);
    clips.data = clip_data.data(); clips.offset = clip_off.data(); clips.length = clip_len.data(); clips.nclips = n;
    sint[FLY_SI_FORCED_CLIP] = 0;
  }
  void set_wbpg(const double* freqs, const int* off, const double* traj, const double* phase, int nf, const double* prm) override {
    const int rows = off[nf];
    wb_freqs.assign(freqs, freqs + nf); wb_off.assign(off, off + nf + 1);
    wb_traj.assign(traj, traj + (size_t)rows * 6); wb_traj.resize((size_t)(rows + 1) * 6, 0.0);
    wb_phase.assign(phase, phase + rows);
    wb.nfreq = nf; wb.beat_freqs = wb_freqs.data(); wb.offsets = wb_off.data(); wb.traj = wb_traj.data(); wb.phase = wb_phase.data();
    wb.base_freq = prm[0]; wb.rel_range = prm[1]; wb.rate = prm[2]; wb.dt_ctrl = prm[3];
  }
  void set_phase(double p) override { phase0 = p; }
  void set_terrain(const float* data, int n, const double* meta, const float* ty, int stride) override {
    hf.assign(data, data + (size_t)n * m.hfield_nrow * m.hfield_ncol);
    m.hf_pool = hf.data(); m.nterrain = n; m.trench_meta = 0; m.trench_y = 0; m.trench_stride = stride;
    if (meta) {
      trench_meta.assign(meta, meta + 3 * n); trench_y.assign(ty, ty + (size_t)n * stride);
      m.trench_meta = trench_meta.data(); m.trench_y = trench_y.data();
    }
  }
  void set_seed(unsigned long long s, int off) override { seed = s; env_offset = off; }
  void vision_draws(int episode, int* terrain, double* height, double* speed) override {
    T h, v;
    fly_vision_draws(m, seed, env_offset, episode, terrain, &h, &v);
    *height = (double)h; *speed = (double)v;
  }
  void eyes(float* out24, int* terrain) override { memcpy(out24, eye_out.data(), 24 * sizeof(float)); *terrain = eye_terrain; }
  void wbpg_state(int* step, int* fidx, double* freq) override {
    *step = sint[FLY_SI_WB_STEP]; *fidx = sint[FLY_SI_WB_FIDX]; *freq = fly_get_double(sint.data() + FLY_SI_WB_FREQ_LO);
  }
  FlyIO<T> io(const float* action, float* obs, float* rdst) {
    FlyIO<T> o;
    memset(&o, 0, sizeof(o));
    o.sreal = sreal.data(); o.sint = sint.data(); o.sreal_stride = (int)sreal.size();
    o.action = action; o.obs = obs; o.reward = rdst; o.discount = rdst + 1; o.step_type = (int*)(rdst + 2);
    o.wbpg_phase = &phase0;
    o.seed = seed; o.env_offset = env_offset;
    o.eyes = eye_out.data(); o.eye_terrain = &eye_terrain;
    return o;
  }
  int env_reset(float* obs) override {
    float rd[4];
    FlyIO<T> o = io(0, obs, rd);
    o.force_reset = 1;
    fly_env_run(m, w, clips, wb, o, 0);
    return *(int*)(rd + 2);
  }
  int env_step(const float* action, float* obs, double* reward, double* discount) override {
    float rd[4];
    FlyIO<T> o = io(action, obs, rd);
    /* full-precision reward: rerun through the typed path */
    fly_env_run(m, w, clips, wb, o, 0);
    *reward = rd[0]; *discount = rd[1];
    return *(int*)(rd + 2);
  }
  void forward() override { fly_step1(m, w, true); fly_forward_acc(m, w); }
  void physics_step() override { fly_forward_acc(m, w); fly_integrate(m, w); fly_step1(m, w, true); }
  void call(const char* stage) override {
    std::string s(stage);
    if (s == "kinematics") fly_kinematics(m, w);
    else if (s == "com_pos") fly_com_pos(m, w);
    else if (s == "crb") fly_crb(m, w);
    else if (s == "factor") {
      T* LD[2] = {w.qLD, w.qH}; T* iv[2] = {w.qDinv, w.qHinv};
      fly_factor<T, 2>(m, LD, iv);
    }
    else if (s == "collision") fly_collision(m, w);
    else if (s == "make_constraint") fly_make_constraint(m, w);
    else if (s == "step1") fly_step1(m, w, true);
    else if (s == "forward_acc") fly_forward_acc(m, w);
    else if (s == "integrate") fly_integrate(m, w);
    else if (s == "solve_stmp") { const T* LD[1] = {w.qLD}; const T* iv[1] = {w.qDinv}; T* x[1] = {w.stmp}; fly_solve<T, 1>(m, w, LD, iv, x); }
    else fprintf(stderr, "emu: unknown stage %s\n", stage);
  }
  int get(const char* name, double* out, int cap) override {
    size_t fb, bb;
    (void)fb; (void)bb;
    const int nl = 32; (void)nl;   /* lanes per env the workspace was sized for (fly_ws.cuh) */
#define X(n, cnt, tier) if (!strcmp(name, #n)) { int c = (int)(cnt); if (c > cap) c = cap; for (int i = 0; i < c; i++) out[i] = (double)w.n[i]; return c; }
    FLY_WS_REAL(X)
    FLY_WS_DBL(X)
#undef X
    if (!strcmp(name, "sreal")) { int c = (int)sreal.size() < cap ? (int)sreal.size() : cap; for (int i = 0; i < c; i++) out[i] = (double)sreal[i]; return c; }
    return -1;
  }
  int set(const char* name, const double* in, int n) override {
    const int nl = 32; (void)nl;
#define X(nm, cnt, tier) if (!strcmp(name, #nm)) { int c = (int)(cnt); if (c > n) c = n; for (int i = 0; i < c; i++) w.nm[i] = (T)in[i]; return c; }
    FLY_WS_REAL(X)
#undef X
    return -1;
  }
  int geti(const char* name, int* out, int cap) override {
#define X(n, cnt, tier) if (!strcmp(name, #n)) { int c = (int)(cnt); if (c > cap) c = cap; for (int i = 0; i < c; i++) out[i] = (int)w.n[i]; return c; }
    FLY_WS_INT(X)
    FLY_WS_U16(X)
#undef X
    if (!strcmp(name, "sint")) { int c = FLY_SI_COUNT < cap ? FLY_SI_COUNT : cap; for (int i = 0; i < c; i++) out[i] = sint[i]; return c; }
    if (!strcmp(name, "mrow")) { int c = m.nM < cap ? m.nM : cap; for (int i = 0; i < c; i++) out[i] = m.mrow[i]; return c; }
    if (!strcmp(name, "mcol")) { int c = m.nM < cap ? m.nM : cap; for (int i = 0; i < c; i++) out[i] = m.mcol[i]; return c; }
    if (!strcmp(name, "nM")) { out[0] = m.nM; return 1; }
    if (!strcmp(name, "ntrunk")) { out[0] = m.ntrunk; return 1; }
    if (!strcmp(name, "gsteps")) { out[0] = m.gsteps; return 1; }
    if (!strcmp(name, "nlevel")) { out[0] = m.nlevel; return 1; }
    return -1;
  }
};

/* test hook: the engine's task-layer quaternion helpers (fly_env.cuh, fly_math.cuh) on raw inputs, in T; ops as flyo_quat_op */
template <class T> static int quat_op_t(int op, const double* ad, const double* bd, double* out) {
  T a[6], b[4], o[4] = {0, 0, 0, 0};
  for (int k = 0; k < 6; k++) a[k] = (T)ad[k];
  for (int k = 0; k < 4; k++) b[k] = (T)bd[k];
  int n = 0;
  switch (op) {
    case 0: quat_recip(o, a); n = 4; break;
    case 1: quat_rotate(o, a, b); n = 3; break;
    case 2: quat_z2vec(o, a); n = 4; break;
    case 3: o[0] = quat_dist_short_arc(a, b); n = 1; break;
    case 4: {   /* fly_walk_imitation_reward's joint quaternion */
      T q1[4], q2[4];
      quat_z2vec(q1, a);
      const T nn = norm3(a);
      const T a2[3] = {a[0] / nn, a[1] / nn, a[2] / nn};
      axis_angle_quat(q2, a2, b[0]);
      quat_mul(o, q2, q1); n = 4; break;
    }
    case 5: { T qi[4]; quat_recip(qi, a); quat_mul(o, qi, b); n = 4; break; }
    case 6: { T d[3] = {a[3] - a[0], a[4] - a[1], a[5] - a[2]}, qc[4] = {b[0], -b[1], -b[2], -b[3]}; quat_rotate(o, d, qc); n = 3; break; }
    case 7: quat_mul(o, a, b); n = 4; break;
  }
  for (int k = 0; k < n; k++) out[k] = (double)o[k];
  return n;
}
extern "C" {
void* emu_create(const FlybModel* m, int flags, int use_double, char* errbuf, int errcap) {
  EmuBase* e;
  std::string err;
  if (use_double) { Emu<double>* x = new Emu<double>(m, flags); err = x->err; e = x; }
  else { Emu<float>* x = new Emu<float>(m, flags); err = x->err; e = x; }
  if (!err.empty()) { snprintf(errbuf, errcap, "%s", err.c_str()); delete e; return 0; }
  return e;
}
void emu_destroy(void* h) { delete (EmuBase*)h; }
/* the engine's height-field narrow phase (csrc/fly_hfield.cuh) on the CPU: same arguments as flyb_hfield_contacts */
int emu_hfield_contacts(const float* data, int nrow, int ncol, const double* pose12, const double* size4, int nquery,
                        const int* types, const double* geoms16, int maxcon, double* out) {
  FlyHfield hf;
  hf.data = data; hf.nrow = nrow; hf.ncol = ncol;
  for (int k = 0; k < 3; k++) hf.pos[k] = pose12[k];
  for (int k = 0; k < 9; k++) hf.mat[k] = pose12[3 + k];
  for (int k = 0; k < 4; k++) hf.size[k] = size4[k];
  std::vector<double> scratch(4 * 6 * 32);
  const Portal<double> P = {scratch.data()};
  for (int q = 0; q < nquery; q++) {
    const double* g = geoms16 + 16 * (size_t)q;
    double* o = out + (size_t)q * (1 + 7 * maxcon);
    o[0] = (double)fly_hfield_narrow(hf, types[q], g, g + 3, g + 12, g[15], maxcon, o + 1, P);
  }
  return 0;
}
int emu_quat_op(int use_double, int op, const double* a, const double* b, double* out) {
  return use_double ? quat_op_t<double>(op, a, b, out) : quat_op_t<float>(op, a, b, out);
}
void emu_set_clips(void* h, const float* data, const int* off, const int* len, int n) { ((EmuBase*)h)->set_clips(data, off, len, n); }
void emu_set_wbpg(void* h, const double* freqs, const int* off, const double* traj, const double* phase, int nf, const double* prm) {
  ((EmuBase*)h)->set_wbpg(freqs, off, traj, phase, nf, prm);
}
void emu_set_phase(void* h, double p) { ((EmuBase*)h)->set_phase(p); }
void emu_set_terrain(void* h, const float* data, int n, const double* meta, const float* ty, int stride) { ((EmuBase*)h)->set_terrain(data, n, meta, ty, stride); }
void emu_set_seed(void* h, unsigned long long seed, int env_offset) { ((EmuBase*)h)->set_seed(seed, env_offset); }
void emu_vision_draws(void* h, int episode, int* terrain, double* height, double* speed) { ((EmuBase*)h)->vision_draws(episode, terrain, height, speed); }
void emu_eyes(void* h, float* out24, int* terrain) { ((EmuBase*)h)->eyes(out24, terrain); }
void emu_wbpg_state(void* h, int* step, int* fidx, double* freq) { ((EmuBase*)h)->wbpg_state(step, fidx, freq); }
int emu_env_reset(void* h, float* obs) { return ((EmuBase*)h)->env_reset(obs); }
int emu_env_step(void* h, const float* a, float* obs, double* r, double* g) { return ((EmuBase*)h)->env_step(a, obs, r, g); }
void emu_forward(void* h) { ((EmuBase*)h)->forward(); }
void emu_physics_step(void* h) { ((EmuBase*)h)->physics_step(); }
void emu_call(void* h, const char* stage) { ((EmuBase*)h)->call(stage); }
int emu_get(void* h, const char* name, double* out, int cap) { return ((EmuBase*)h)->get(name, out, cap); }
int emu_set(void* h, const char* name, const double* in, int n) { return ((EmuBase*)h)->set(name, in, n); }
int emu_geti(void* h, const char* name, int* out, int cap) { return ((EmuBase*)h)->geti(name, out, cap); }
}
