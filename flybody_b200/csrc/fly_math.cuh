/* fly_math.cuh - execution-model macros and small vector/quaternion helpers for the fly engine.
 *
 * One environment is stepped by ONE WARP.  All engine code is written against four primitives:
 *   FLY_FOR(i, n)   work items 0..n-1 spread over the 32 lanes (items must be independent)
 *   FLY_SYNC()      warp barrier + memory ordering between phases
 *   FLY_LANE0       section executed by a single lane
 *   warp_sum/max    butterfly all-reduce (identical bits on every lane, so control flow stays uniform)
 * Under nvcc these map to lane-strided loops, __syncwarp and shuffles.  With FLY_EMU defined (tests/emu,
 * built by g++ for the CPU unit tests) a "warp" is a single thread: FLY_FOR is a plain loop and the rest
 * are no-ops, which is exactly equivalent as long as FLY_FOR bodies are independent. */
#ifndef FLY_MATH_CUH_
#define FLY_MATH_CUH_
#include <math.h>
#include <stdint.h>

#ifndef FLY_SYNC_MASK
/* bit per block rendezvous point (fly_task.cuh).  All eighteen except the four between the constraint solve and the next
 * collision pass (5: before the acceleration sensors, 6: integration, 0: kinematics, 7: subtree CoM): warps that finish
 * their solve early run those short stages instead of waiting (measured 60.2 -> 59.7 ms per step; fewer points cost more:
 * profiles/r1_summary.md) */
#define FLY_SYNC_MASK 0x3ff1eu
#endif
#define FLY_STAGE_SYNC(bit) do { if (FLY_SYNC_MASK & (1u << (bit))) FLY_BLOCK_SYNC(); } while (0)

/* Lanes per environment.  FLY_NL = 32: one warp steps one env.  FLY_NL = 16 (8): a warp steps two (four) envs side by side,
 * each on its own group of FLY_NL consecutive lanes - the semantics of cooperative_groups::tiled_partition<FLY_NL>: barriers,
 * shuffles and ballots carry the group's lane mask, so the groups of a warp may run different control flow (different contact
 * counts, solver iterations) and still share every instruction they execute in common.  Most stages have fewer than 16
 * independent work items (profiles/r1_by_function_final.txt: 16.0 of 32 lanes active), and the register file, not the lane
 * count, is what limits the envs resident on an SM: narrower groups put more envs behind the same registers. */
#ifndef FLY_NL
#define FLY_NL 32
#endif
#define FLY_SUBS (32 / FLY_NL)   /* envs per warp */

#if defined(__CUDACC__) && !defined(FLY_EMU)
#define FLY_GPU 1
#define FLY_FN __device__ __forceinline__
#define FLY_INL __device__ __forceinline__
#define FLY_MEM __device__ __forceinline__   /* member functions */
#define FLY_NOINL __device__ __noinline__     /* the few routines that want their own register allocation */
#define FLY_LANE() ((int)(threadIdx.x & (FLY_NL - 1)))
#if FLY_NL == 32
#define FLY_SHIFT() 0
#define FLY_MASK() 0xffffffffu
#else
#define FLY_SHIFT() ((int)(threadIdx.x & 31 & ~(FLY_NL - 1)))          /* first lane of this env's group */
#define FLY_MASK() (((1u << FLY_NL) - 1u) << FLY_SHIFT())             /* the group's lanes */
#endif
#define FLY_SUB() ((int)((threadIdx.x & 31) / FLY_NL))                 /* which group of the warp */
#define FLY_FOR(i, n) for (int i = FLY_LANE(); i < (n); i += FLY_NL)
#define FLY_SYNC() __syncwarp(FLY_MASK())
/* ballot over the group, bit l = lane l of the group */
#define FLY_BALLOT(p) ((__ballot_sync(FLY_MASK(), (p)) >> FLY_SHIFT()) & (FLY_NL == 32 ? 0xffffffffu : ((1u << (FLY_NL & 31)) - 1u)))
/* 2-D work items over a (FLY_NL / 8) x 8 lane tile: rows i in [0,n) across lane / 8, columns j in [j0,n) across lane % 8 */
#define FLY_FOR_ROW(i, n) for (int i = FLY_LANE() >> 3; i < (n); i += FLY_NL / 8)
#define FLY_FOR_COL(j, j0, n) for (int j = (j0) + (FLY_LANE() & 7); j < (n); j += 8)
/* block-wide rendezvous of the warps of a block (each steps its own envs): keeps them in the same stage so that
 * instruction fetches and model-table lines are shared in the SM's caches.  Every env must pass the same number.
 * With several envs per warp the groups of a warp may arrive apart: the barrier is the per-thread (non-aligned) form. */
#ifndef FLY_SYNC_GROUP
#if FLY_NL == 32
#define FLY_BLOCK_SYNC() __syncthreads()
#else
#define FLY_BLOCK_SYNC() asm volatile("barrier.sync 0;" ::: "memory")
#endif
#else  /* lockstep groups of FLY_SYNC_GROUP warps (named barriers 1..): smaller groups wait less for their slowest env */
#define FLY_BLOCK_SYNC() \
  asm volatile("barrier.sync %0, %1;" ::"r"(1 + (int)(threadIdx.x >> 5) / FLY_SYNC_GROUP), "n"(FLY_SYNC_GROUP * 32) : "memory")
#endif
#define FLY_LANE0 if (FLY_LANE() == 0)
#define FLY_LDG(p) __ldg(p)
template <class T> FLY_INL T warp_sum(T v) {
#pragma unroll
  for (int o = FLY_NL / 2; o > 0; o >>= 1) v += __shfl_xor_sync(FLY_MASK(), v, o);
  return v;
}
template <class T> FLY_INL T warp_max(T v) {
#pragma unroll
  for (int o = FLY_NL / 2; o > 0; o >>= 1) { T u = __shfl_xor_sync(FLY_MASK(), v, o); v = u > v ? u : v; }
  return v;
}
FLY_INL int warp_sum_i(int v) {
#pragma unroll
  for (int o = FLY_NL / 2; o > 0; o >>= 1) v += __shfl_xor_sync(FLY_MASK(), v, o);
  return v;
}
/* all-reduce of (value, index) pairs to the smallest value, lowest index on ties (numpy.argmin) */
FLY_INL void warp_argmin(double& v, int& i) {
#pragma unroll
  for (int o = FLY_NL / 2; o > 0; o >>= 1) {
    const double ov = __shfl_xor_sync(FLY_MASK(), v, o);
    const int oi = __shfl_xor_sync(FLY_MASK(), i, o);
    if (ov < v || (ov == v && oi < i)) { v = ov; i = oi; }
  }
}
/* ordered stream compaction of the items i in [0,n) with pred(i): writes their indices to out[], in order.  Four batches
 * of FLY_NL items are evaluated before their ballots, so that the loads of four independent predicates are in flight together */
template <class P> FLY_INL int warp_compact(int n, P pred, unsigned short* out, int maxout) {
  int cnt = 0;
  const int lane = FLY_LANE();
  for (int base = 0; base < n; base += 4 * FLY_NL) {
    bool p[4];
#pragma unroll
    for (int u = 0; u < 4; u++) { const int i = base + FLY_NL * u + lane; p[u] = i < n && pred(i); }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const unsigned mask = FLY_BALLOT(p[u]);
      const int pos = cnt + __popc(mask & ((1u << lane) - 1u));
      if (p[u] && pos < maxout) out[pos] = (unsigned short)(base + FLY_NL * u + lane);
      cnt += __popc(mask);
    }
  }
  FLY_SYNC();
  return cnt;
}
/* ordered in-place filter of list[0..n): keeps the entries with pred(i), returns how many (a kept entry never moves up) */
template <class P> FLY_INL int warp_filter(int n, P pred, unsigned short* list) {
  int cnt = 0;
  const int lane = FLY_LANE();
  for (int base = 0; base < n; base += FLY_NL) {
    const int i = base + lane;
    const unsigned short v = i < n ? list[i] : (unsigned short)0;
    const bool p = i < n && pred(i);
    const unsigned mask = FLY_BALLOT(p);   /* every lane has read its entry before any lane writes */
    if (p) list[cnt + __popc(mask & ((1u << lane) - 1u))] = v;
    cnt += __popc(mask);
  }
  FLY_SYNC();
  return cnt;
}
#else
#define FLY_FN static
#define FLY_INL static inline
#define FLY_MEM inline
#define FLY_NOINL static
#define FLY_LANE() 0
#define FLY_FOR(i, n) for (int i = 0; i < (n); i++)
#define FLY_FOR_ROW(i, n) for (int i = 0; i < (n); i++)
#define FLY_FOR_COL(j, j0, n) for (int j = (j0); j < (n); j++)
#define FLY_SYNC()
#define FLY_BLOCK_SYNC()
#define FLY_LANE0
#define FLY_LDG(p) (*(p))
template <class T> FLY_INL T warp_sum(T v) { return v; }
template <class T> FLY_INL T warp_max(T v) { return v; }
FLY_INL int warp_sum_i(int v) { return v; }
FLY_INL void warp_argmin(double& v, int& i) { (void)v; (void)i; }
template <class P> FLY_INL int warp_compact(int n, P pred, unsigned short* out, int maxout) {
  int cnt = 0;
  for (int i = 0; i < n; i++)
    if (pred(i)) { if (cnt < maxout) out[cnt] = (unsigned short)i; cnt++; }
  return cnt;
}
template <class P> FLY_INL int warp_filter(int n, P pred, unsigned short* list) {
  int cnt = 0;
  for (int i = 0; i < n; i++)
    if (pred(i)) list[cnt++] = list[i];
  return cnt;
}
#endif

#define FLY_MINVAL 1e-15
#define FLY_PI 3.14159265358979323846

template <class T> FLY_INL T fmin_(T a, T b) { return a < b ? a : b; }
template <class T> FLY_INL T fmax_(T a, T b) { return a > b ? a : b; }
template <class T> FLY_INL T fabs_(T a) { return a < 0 ? -a : a; }
/* smallest positive value treated as zero; float cannot represent 1e-15 relative sizes so keep it tiny */
template <class T> FLY_INL T minval() { return (T)FLY_MINVAL; }

template <class T> FLY_INL void cross3(T* r, const T* a, const T* b) {
  T x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
template <class T> FLY_INL T dot3(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <class T> FLY_INL T norm3(const T* a) { return sqrt(dot3(a, a)); }
template <class T> FLY_INL T normalize3(T* a) {
  T n = norm3(a);
  if (n < minval<T>()) { a[0] = 1; a[1] = 0; a[2] = 0; return 0; }
  T s = T(1) / n;
  a[0] *= s; a[1] *= s; a[2] *= s;
  return n;
}
template <class T> FLY_INL void quat_mul(T* r, const T* a, const T* b) {
  T w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  T x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  T y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  T z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
template <class T> FLY_INL void quat_normalize(T* q) {
  T n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < minval<T>()) { q[0] = 1; q[1] = q[2] = q[3] = 0; return; }
  T s = T(1) / n;
  q[0] *= s; q[1] *= s; q[2] *= s; q[3] *= s;
}
template <class T> FLY_INL void quat2mat(T* m, const T* q) {
  T w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
template <class T> FLY_INL void mat_vec(T* r, const T* m, const T* v) {
  T x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2],
    z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
template <class T> FLY_INL void matT_vec(T* r, const T* m, const T* v) {
  T x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2], y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2],
    z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
template <class T> FLY_INL void mat_mul(T* r, const T* a, const T* b) { /* r = a b, 3x3 row-major */
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) r[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
}
template <class T> FLY_INL void rot_vec_quat(T* r, const T* v, const T* q) {
  T m[9];
  quat2mat(m, q);
  mat_vec(r, m, v);
}
template <class T> FLY_INL void axis_angle_quat(T* q, const T* axis, T angle) {
  T s = sin(T(0.5) * angle);
  q[0] = cos(T(0.5) * angle); q[1] = axis[0] * s; q[2] = axis[1] * s; q[3] = axis[2] * s;
}
/* q <- q * exp(h w / 2), w in the local frame */
template <class T> FLY_INL void quat_integrate(T* q, const T* w, T h) {
  T ax[3] = {w[0], w[1], w[2]};
  T n = normalize3(ax);
  if (n * h == 0 || n < minval<T>()) return;
  T dq[4], r[4];
  axis_angle_quat(dq, ax, n * h);
  quat_mul(r, q, dq);
  q[0] = r[0]; q[1] = r[1]; q[2] = r[2]; q[3] = r[3];
  quat_normalize(q);
}
/* spatial algebra on [angular; linear] 6-vectors about a common point */
template <class T> FLY_INL void cross_motion(T* r, const T* v, const T* s) {
  T a[3], b[3], c[3];
  cross3(a, v, s); cross3(b, v, s + 3); cross3(c, v + 3, s);
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
template <class T> FLY_INL void cross_force(T* r, const T* v, const T* f) {
  T a[3], b[3], c[3];
  cross3(a, v, f); cross3(b, v + 3, f + 3); cross3(c, v, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}
/* 10-number spatial inertia about a reference point: [Ixx Iyy Izz Ixy Ixz Iyz, m*c(3), m] */
template <class T> FLY_INL void inert_vec(T* r, const T* I, const T* v) {
  const T* h = I + 6;
  T m = I[9], hv[3], wh[3];
  cross3(hv, h, v + 3); cross3(wh, v, h);
  r[0] = I[0] * v[0] + I[3] * v[1] + I[4] * v[2] + hv[0];
  r[1] = I[3] * v[0] + I[1] * v[1] + I[5] * v[2] + hv[1];
  r[2] = I[4] * v[0] + I[5] * v[1] + I[2] * v[2] + hv[2];
  r[3] = m * v[3] + wh[0]; r[4] = m * v[4] + wh[1]; r[5] = m * v[5] + wh[2];
}
#endif
