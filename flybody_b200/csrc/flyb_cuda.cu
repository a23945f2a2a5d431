/* flyb_cuda.cu - the CUDA engine behind include/flyb.h (sm_100a).
 *
 * One persistent warp steps one environment at a time: warps pull env indices from a global counter until
 * the batch is done, so envs with many contacts do not stall their neighbours.  Each resident warp owns a
 * slice of shared memory (the hot tiers of the workspace, fly_ws.cuh) and a scratch slot in global memory
 * that stays hot in L1/L2 because it is indexed by the resident warp, not by the env.  Per-env persistent
 * state (qpos | qvel | act | qacc_warmstart) is one contiguous row per env in HBM, read and written once per
 * control step with coalesced lane-strided accesses. */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/flyb.h"
#include "fly_kernel.cuh"
#include "fly_hfield.cuh"
#include "fly_eyes.cuh"

struct FlybEngine {
  int device = 0, num_envs = 0, flags = 0, env_offset = 0;
  unsigned long long seed = 0;
  DModel<float> m;        /* pointers relocated to device addresses */
  void* d_arena = nullptr;
  float* d_sreal = nullptr; int* d_sint = nullptr; int sreal_stride = 0;
  float* d_stats = nullptr;
  int* d_trace_int = nullptr; float* d_trace_real = nullptr; int trace_real_stride = 0;
  char* d_bulk = nullptr; size_t bulk_stride = 0, fast_stride = 0; int bulk_slots = 0;
  int* d_counter = nullptr;
  int *d_work = nullptr, *d_perm = nullptr; int binning = 0;   /* FLYB_BINNING=1: cost-sorted env -> block assignment (measured: no gain, off) */
  float* d_clip_rows = nullptr; int* d_clip_off = nullptr; int* d_clip_len = nullptr; int n_clips = 0;
  FlyWbpg wb = {};        /* device pointers; nfreq 0 until flyb_load_wbpg */
  void* d_wbpg = nullptr; double* d_wbpg_phase = nullptr;
  /* vision_flight: terrain pool, trench centre lines, and the per-env eye camera frames the step kernel leaves for flyb_render */
  float* d_hf = nullptr; double* d_trench_meta = nullptr; float* d_trench_y = nullptr;
  float* d_eyes = nullptr; int* d_eye_terrain = nullptr; float* d_hf_bmax = nullptr; unsigned* d_eye_counter = nullptr;
  /* staging for the host-pointer API */
  float *h_action = nullptr, *h_obs = nullptr, *h_rd = nullptr; int* h_st = nullptr; unsigned char* h_mask = nullptr;
  float *d_action = nullptr, *d_obs = nullptr, *d_rd = nullptr; int* d_st = nullptr; unsigned char* d_mask = nullptr;
  cudaStream_t host_stream = nullptr;
  cudaEvent_t last_launch = nullptr;   /* recorded after every launch on the caller's stream: the *_host calls wait for it */
  int smem_level = 0, warps = 28, blocks_per_sm = 0, sm_count = 148, grid = 0;
  int lanes = 32;               /* lanes per env (FLY_NL of the kernels this handle launches): fixed at flyb_create */
  int spec = -1, no_spec = 0;   /* index of the model-specialised kernel in use (-1: generic); FLYB_NO_SPEC=1 forces generic */
  long long launches = 0;
  unsigned char* noalloc_scratch = nullptr;   /* device address of g_noalloc_scratch on h->device (flyb_step_noalloc) */
  std::string err;
};

static std::string g_create_error;

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      h->err = std::string(#call) + ": " + cudaGetErrorString(e_);                                 \
      return 1;                                                                                    \
    }                                                                                              \
  } while (0)

/* the kernels live in their own translation units (flyb_kern.cu); each registers itself at load time */
struct FlyKernelEntry { int lanes, warps, level, spec; const void* fn; };
static std::vector<FlyKernelEntry>& kernel_registry() { static std::vector<FlyKernelEntry> r; return r; }
void fly_register_kernel(int lanes, int warps, int level, int spec, const void* fn) {
  kernel_registry().push_back(FlyKernelEntry{lanes, warps, level, spec, fn});
}
static int spec_of(const DModel<float>& m) {
  const WsDims d = ws_dims_of(m);
#define X(i, ...) { const WsDims t = {__VA_ARGS__}; if (ws_dims_equal(d, t)) return i; }
  FLY_SPEC_MODELS(X)
#undef X
  return -1;
}
static const void* pick_kernel(int lanes, int warps, int level, int spec) {
  for (const FlyKernelEntry& e : kernel_registry())
    if (e.lanes == lanes && e.warps == warps && e.level == level && e.spec == spec) return e.fn;
  return nullptr;
}

/* Launch order of the envs for the next step: counting sort by the time their constraint solves took in the last step,
 * slowest first.  The warps of a block walk the pipeline in step (FLY_STAGE_SYNC), so a block is as slow as its busiest
 * env; the cost of an env persists from step to step (it follows its contact state), so with envs of similar cost side by
 * side the wait at the rendezvous after the constraint solve shrinks, and the light envs end up in the tail of the grid.
 * Keys are linear in cost up to 4x the mean (1024 bins), everything above shares the first bin.  Which warp steps an
 * env never changes that env's result. */
#define FLY_NBINS 1024
__global__ void __launch_bounds__(1024) fly_bin_kernel(const int* __restrict__ work, int* __restrict__ perm, int n) {
  __shared__ int cnt[FLY_NBINS], start[FLY_NBINS];
  __shared__ unsigned long long total;
  if (threadIdx.x == 0) total = 0;
  cnt[threadIdx.x] = 0;
  __syncthreads();
  unsigned long long mine = 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) mine += (unsigned long long)(work[i] > 0 ? work[i] : 0);
  atomicAdd(&total, mine);
  __syncthreads();
  const unsigned long long mean = total / (unsigned long long)n + 1;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const unsigned long long k = (unsigned long long)(work[i] > 0 ? work[i] : 0) * 256ull / mean;
    atomicAdd(&cnt[FLY_NBINS - 1 - (int)(k < FLY_NBINS ? k : FLY_NBINS - 1)], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) { int s = 0; for (int b = 0; b < FLY_NBINS; b++) { start[b] = s; s += cnt[b]; } }
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const unsigned long long k = (unsigned long long)(work[i] > 0 ? work[i] : 0) * 256ull / mean;
    perm[atomicAdd(&start[FLY_NBINS - 1 - (int)(k < FLY_NBINS ? k : FLY_NBINS - 1)], 1)] = i;
  }
}

/* ---- height field against convex geoms, one query per thread (flyb_hfield_contacts: first slice of the vision-guided flight
 * terrain).  The portal scratch of the convex search is lane-interleaved per warp in shared memory. */
__global__ void __launch_bounds__(128) fly_hfield_kernel(FlyHfield hf, int nq, const int* __restrict__ types,
                                                          const double* __restrict__ geoms, int maxcon, double* __restrict__ out) {
  __shared__ double scratch[4][4 * 6 * 32];
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  const Portal<double> P = {scratch[threadIdx.x >> 5] + (threadIdx.x & 31)};
  if (q >= nq) return;
  const double* g = geoms + 16 * (size_t)q;
  double* o = out + (size_t)q * (1 + 7 * maxcon);
  o[0] = (double)fly_hfield_narrow(hf, types[q], g, g + 3, g + 12, g[15], maxcon, o + 1, P);
}

/* ---- eye cameras: one thread per pixel (flyb_render_eyes, csrc/fly_eyes.cuh) */
__global__ void __launch_bounds__(256) fly_eyes_kernel(FlyScene sc, int ncam, const int* __restrict__ terrain_id,
                                                        const float* __restrict__ cams, int W, int H, float tanhalf,
                                                        unsigned char* __restrict__ out) {
  const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= (size_t)ncam * W * H) return;
  const int cam = (int)(p / ((size_t)W * H)), i = (int)((p / W) % H), j = (int)(p % W);
  const float* c = cams + 12 * (size_t)cam;
  const float* h = sc.data + (size_t)terrain_id[cam] * sc.nrow * sc.ncol;
  eye_pixel<float>(sc, h, c, c + 3, W, H, tanhalf, i, j, out + 3 * p);
}

/* largest surface height of every FLY_EYE_BLOCK x FLY_EYE_BLOCK block of cells (corner heights as the ray caster reads them) */
__global__ void fly_eyes_blockmax_kernel(const float* __restrict__ data, int nterrain, int nrow, int ncol, float zoff, int shift, int nbr,
                                         int nbc, size_t stride, float* __restrict__ bmax) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nterrain * nbr * nbc) return;
  const int t = i / (nbr * nbc), br = (i / nbc) % nbr, bc = i % nbc, B = 1 << shift;
  const float* h = data + (size_t)t * nrow * ncol;
  float mx = 0.f;
  for (int r = br * B; r <= (br + 1) * B && r < nrow; r++)
    for (int c = bc * B; c <= (bc + 1) * B && c < ncol; c++) mx = fmaxf(mx, eye_height<float>(h, ncol, r, c, zoff));
  bmax[(size_t)t * stride + (size_t)br * nbc + bc] = mx;
}

/* the terrain's maximum = the largest of its 16 x 16 block maxima, stored behind the two block tables */
__global__ void fly_eyes_gmax_kernel(float* __restrict__ bmax, int nterrain, int nb, size_t stride) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nterrain) return;
  float mx = 0.f;
  for (int i = 0; i < nb; i++) mx = fmaxf(mx, bmax[(size_t)t * stride + i]);
  bmax[(size_t)t * stride + stride - 1] = mx;
}

/* the same for the eyes of a handle's envs (flyb_render): camera 2 e + k is eye k of env e, flying over terrain terrain_id[e].
 * Persistent: every lane walks one ray a step at a time and takes its next pixel from a global counter as soon as its ray has
 * ended, so the 32 lanes of a warp stay busy although rays take between 5 and 200 steps. */
__global__ void __launch_bounds__(256) fly_eyes_env_kernel(FlyScene sc, int ncam, const int* __restrict__ terrain_id,
                                                            const float* __restrict__ cams, int W, int H, float tanhalf,
                                                            unsigned char* __restrict__ out, unsigned* __restrict__ counter) {
  const unsigned total = (unsigned)ncam * W * H, lane = threadIdx.x & 31;
  const size_t tstride = (size_t)sc.nbr * sc.nbc + (size_t)sc.nbr2 * sc.nbc2 + 1;
  EyeRay<float> y;
  const float *h = nullptr, *bm = nullptr;
  long long pix = -1;          /* -1: needs a ray, -2: no rays left */
  bool cand = false;           /* the ray stands in a cell that passed the height test */
  for (;;) {
    const unsigned need = __ballot_sync(0xffffffffu, pix == -1);
    if (need) {
      unsigned base = 0;
      const int leader = __ffs(need) - 1;
      if ((int)lane == leader) base = atomicAdd(counter, (unsigned)__popc(need));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (pix == -1) {
        const unsigned p = base + __popc(need & ((1u << lane) - 1u));
        if (p >= total) pix = -2;
        else {
          pix = p;
          const int cam = (int)(p / ((unsigned)W * H)), i = (int)((p / W) % H), j = (int)(p % W);
          const float* c = cams + 12 * (size_t)cam;
          const int t = terrain_id[cam >> 1];
          h = sc.data + (size_t)t * sc.nrow * sc.ncol;
          bm = sc.bmax ? sc.bmax + (size_t)t * tstride : nullptr;
          eye_ray_of_pixel(y, c, c + 3, W, H, tanhalf, i, j);
          if (!eye_ray_init(sc, y)) { eye_store(out + 3 * (size_t)p, sc.sky); pix = -1; }
          else eye_ray_set_gmax(sc, bm, y);
        }
      }
    }
    if (__all_sync(0xffffffffu, pix == -2)) break;
    /* the cheap part of the walk (jumps, height tests) a few times over, then the triangle tests of every lane that has found a
       candidate cell in one go: the expensive code runs with as many lanes as possible */
#pragma unroll 1
    for (int it = 0; it < 4; it++) {
      if (pix >= 0 && !cand) {
        const int st = eye_ray_advance(sc, h, bm, y);
        if (st == 1) { eye_store(out + 3 * (size_t)pix, sc.sky); pix = -1; }
        cand = st == 3;
      }
    }
    if (pix >= 0 && cand) {
      float rgb[3] = {sc.sky[0], sc.sky[1], sc.sky[2]};
      const int st = eye_ray_test(sc, y, rgb);
      cand = false;
      if (st) { eye_store(out + 3 * (size_t)pix, rgb); pix = -1; }
    }
  }
}

/* The same renderer with the warp choosing its next move: every lane is in one of the states "needs a ray", "try a jump", "test the
 * cell", "test the two triangles"; each round the warp runs the move most of its lanes wait for (ballots), the other lanes keep their
 * state.  The fixed order of the kernel above (four advances, one triangle test) ran its code with 8 of 32 lanes; this one is the default
 * (FLYB_EYES_SCHED=0 selects the other). */
__global__ void __launch_bounds__(256) fly_eyes_sched_kernel(FlyScene sc, int ncam, const int* __restrict__ terrain_id,
                                                              const float* __restrict__ cams, int W, int H, float tanhalf,
                                                              unsigned char* __restrict__ out, unsigned* __restrict__ counter) {
  /* moves per round and thresholds, swept on B200 (16 384 envs, trench terrains): (5, 4, 8, 6) 7.05 ms; (3, 2) 7.40; (8, 6) 7.38;
     one move per round 7.97; the fixed order of the kernel above 7.45 */
  constexpr int rep_j = 5, rep_c = 4, min_idle = 8, min_tri = 6;
  const unsigned total = (unsigned)ncam * W * H, lane = threadIdx.x & 31, full = 0xffffffffu;
  const size_t tstride = (size_t)sc.nbr * sc.nbc + (size_t)sc.nbr2 * sc.nbc2 + 1;
  EyeRay<float> y;
  const float *h = nullptr, *bm = nullptr;
  unsigned pix = 0;
  int state = 0;   /* 0: needs a ray, 1: jump, 2: cell, 3: triangles, 4: no rays left */
  for (;;) {
    const unsigned mI = __ballot_sync(full, state == 0), mJ = __ballot_sync(full, state == 1), mC = __ballot_sync(full, state == 2),
                   mT = __ballot_sync(full, state == 3);
    if (!(mI | mJ | mC | mT)) break;
    const int nI = __popc(mI), nJ = __popc(mJ), nC = __popc(mC), nT = __popc(mT);
    if (nI && (nI >= min_idle || !(mJ | mC | mT))) {
      unsigned base = 0;
      const int leader = __ffs(mI) - 1;
      if ((int)lane == leader) base = atomicAdd(counter, (unsigned)nI);
      base = __shfl_sync(full, base, leader);
      if (state == 0) {
        const unsigned p = base + __popc(mI & ((1u << lane) - 1u));
        if (p >= total) state = 4;
        else {
          pix = p;
          const int cam = (int)(p / ((unsigned)W * H)), i = (int)((p / W) % H), j = (int)(p % W);
          const float* c = cams + 12 * (size_t)cam;
          const int t = terrain_id[cam >> 1];
          h = sc.data + (size_t)t * sc.nrow * sc.ncol;
          bm = sc.bmax + (size_t)t * tstride;
          eye_ray_of_pixel(y, c, c + 3, W, H, tanhalf, i, j);
          if (!eye_ray_init(sc, y)) eye_store(out + 3 * (size_t)p, sc.sky);
          else { eye_ray_set_gmax(sc, bm, y); state = 1; }
        }
      }
      continue;
    }
    if (nT && ((nT >= nJ && nT >= nC) || nT >= min_tri)) {
      if (state == 3) {
        float rgb[3] = {sc.sky[0], sc.sky[1], sc.sky[2]};
        const int st = eye_ray_test(sc, y, rgb);
        if (st) { eye_store(out + 3 * (size_t)pix, rgb); state = 0; }
        else state = eye_ray_in_known_blocks(y) ? 2 : 1;
      }
    } else if (nC >= nJ) {
#pragma unroll 1
      for (int rep = 0; rep < rep_c; rep++)
        if (state == 2) {
          const int st = eye_ray_cell(sc, h, y);
          if (st == 1) { eye_store(out + 3 * (size_t)pix, sc.sky); state = 0; }
          else state = st == 3 ? 3 : (st == 4 ? 1 : 2);
        }
    } else {
#pragma unroll 1
      for (int rep = 0; rep < rep_j; rep++)
        if (state == 1) {
          const int st = eye_ray_jump(sc, bm, y);
          if (st == 1) { eye_store(out + 3 * (size_t)pix, sc.sky); state = 0; }
          else if (st == 2) state = 2;
        }
    }
  }
}

/* ---- FP32 FMA peak of the device, measured (flyb_fma_peak): every thread runs 16 independent FMA chains; 8 blocks of 256
 * threads per SM keep the FMA pipes full.  The roofline denominator bench.py quotes FP32-pipe fractions against. */
__global__ void __launch_bounds__(256) fly_fma_peak_kernel(float* out, int iters) {
  float a[16];
#pragma unroll
  for (int k = 0; k < 16; k++) a[k] = 1.0f + 0.001f * (float)(threadIdx.x + k);
  const float m = 0.9999f, c = 1e-4f;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < 16; k++) a[k] = fmaf(a[k], m, c);
  }
  float s = 0;
#pragma unroll
  for (int k = 0; k < 16; k++) s += a[k];
  if (s == 12345.678f) out[0] = s;   /* keeps the chains alive without a store on the timed path */
}

/* ---- scratch of flyb_step_noalloc (include/flyb.h): the step kernel's per-warp workspace slots and its work counter in
 * a file-scope array instead of the cudaMalloc'ed d_bulk / d_counter, for callers that must not allocate.  Every GPU that
 * loads the library holds these FLYB_NOALLOC_SCRATCH_BYTES until exit.  Layout: [0, 256) the work counter (cleared by
 * every call), then workspace slots of bulk_stride bytes each.  All calls on a device share the array: two flyb_step_noalloc
 * calls that overlap on the GPU (different streams, different handles) race on it.
 * In host code the name g_noalloc_scratch is only a handle for the runtime; the device address comes from
 * cudaGetSymbolAddress, resolved once per device by noalloc_scratch_of() with that device current. */
static __device__ __align__(256) unsigned char g_noalloc_scratch[FLYB_NOALLOC_SCRATCH_BYTES];
#define FLYB_NOALLOC_HEADER 256
#define FLYB_MAX_DEVICES 64
static unsigned char* g_noalloc_ptr[FLYB_MAX_DEVICES];   /* device address of the array on each device, 0 = not resolved yet */

/* call with `device` current; returns the device address of the scratch array or null (message in *err) */
static unsigned char* noalloc_scratch_of(int device, std::string* err) {
  if (device < 0 || device >= FLYB_MAX_DEVICES) { *err = "device index beyond FLYB_MAX_DEVICES"; return nullptr; }
  if (!g_noalloc_ptr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) { *err = std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e); return nullptr; }
    g_noalloc_ptr[device] = (unsigned char*)p;
  }
  return g_noalloc_ptr[device];
}

/* where a launch takes its scratch from: null = the handle's own cudaMalloc'ed buffers */
struct LaunchScratch { char* bulk; int slots; int* counter; };

static int launch(FlybEngine* h, const FlyIO<float>& io, cudaStream_t st, const LaunchScratch* sc = nullptr) {
  CK(cudaSetDevice(h->device));
  size_t fb = 0, bb = 0;
  WsOff off;
  ws_layout<float>(h->m, h->smem_level, &off, &fb, &bb, h->lanes);
  h->fast_stride = fb;
  const int W = h->warps, EPB = W * (32 / h->lanes);   /* envs per block */
  const size_t smem = fb * EPB;
  if (smem > 227 * 1024 - 2048) { h->err = "shared memory request too large for this smem_level / warps_per_block"; return 1; }
  if (bb > h->bulk_stride) { h->err = "internal: bulk slot too small"; return 1; }
  int spec = h->no_spec ? -1 : spec_of(h->m);
  const void* kern = pick_kernel(h->lanes, W, h->smem_level, spec);
  if (!kern && spec >= 0) { spec = -1; kern = pick_kernel(h->lanes, W, h->smem_level, -1); }
  h->spec = spec;
  if (h->spec >= 0) {  /* the compile-time layout must be the one the host sized the arenas with */
    WsOff c = {};
#define X(i, ...) if (h->spec == i) { const WsDims t = {__VA_ARGS__}; c = ws_layout_c<float>(t, h->smem_level, h->lanes); }
    FLY_SPEC_MODELS(X)
#undef X
    if (memcmp(&c, &off, sizeof(WsOff)) != 0) { h->err = "internal: compile-time workspace layout differs from the run-time one"; return 1; }
  }
  if (!kern) { h->err = "no step kernel compiled for this (lanes, warps_per_block, smem_level); see KERNEL_UNITS in flybody_b200/engine.py"; return 1; }
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W * 32, smem));
  if (per_sm < 1) { h->err = "kernel does not fit on an SM"; return 1; }
  if (h->blocks_per_sm > 0 && h->blocks_per_sm < per_sm) per_sm = h->blocks_per_sm;
  int grid = h->sm_count * per_sm;
  const int need = (h->num_envs + EPB - 1) / EPB;
  if (grid > need) grid = need;
  const int slots = sc ? sc->slots : h->bulk_slots;
  if (grid * EPB > slots) grid = slots / EPB;
  if (grid < 1) { h->err = "scratch too small for one block of this launch configuration"; return 1; }
  h->grid = grid;
  KArgs a;
  a.clips.data = h->d_clip_rows; a.clips.offset = h->d_clip_off; a.clips.length = h->d_clip_len; a.clips.nclips = h->n_clips;
  a.io = io;
  a.wb = h->wb;
  a.bulk = sc ? sc->bulk : h->d_bulk; a.bulk_stride = h->bulk_stride; a.fast_stride = fb;
  a.counter = sc ? sc->counter : h->d_counter;
  a.num_envs = h->num_envs; a.level = h->smem_level;
  a.perm = nullptr;
  a.io.work = h->d_work;
  if (h->binning && !sc && !io.force_reset && h->num_envs > 2 * EPB) {
    fly_bin_kernel<<<1, 1024, 0, st>>>(h->d_work, h->d_perm, h->num_envs);
    a.perm = h->d_perm;
    h->launches++;
  }
  CK(cudaMemsetAsync(a.counter, 0, sizeof(int), st));
  void* kargs[3] = {(void*)&h->m, (void*)&a, (void*)&off};
  CK(cudaLaunchKernel(kern, dim3(grid), dim3(W * 32), kargs, smem, st));
  CK(cudaGetLastError());
  if (!sc) CK(cudaEventRecord(h->last_launch, st));   /* (flyb_step_noalloc stays free of anything but the clear and the launch) */
  h->launches++;
  return 0;
}

static FlyIO<float> make_io(FlybEngine* h, const float* action, float* obs, float* reward, float* discount, int* st) {
  FlyIO<float> io;
  memset(&io, 0, sizeof(io));
  io.sreal = h->d_sreal; io.sint = h->d_sint; io.sreal_stride = h->sreal_stride;
  io.action = action; io.obs = obs; io.reward = reward; io.discount = discount; io.step_type = st;
  io.trace_int = h->d_trace_int; io.trace_real = h->d_trace_real; io.trace_real_stride = h->trace_real_stride;
  io.stats = h->d_stats; io.seed = h->seed; io.env_offset = h->env_offset;
  io.wbpg_phase = h->d_wbpg_phase;
  io.eyes = h->d_eyes; io.eye_terrain = h->d_eye_terrain;
  return io;
}

extern "C" {

const char* flyb_last_error(flyb_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int flyb_create(const FlybModel* model, int num_envs, int device, uint64_t seed, int env_offset, int flags,
                flyb_handle* out) {
  *out = nullptr;
  FlybEngine* h = new FlybEngine();
  auto fail = [&](const std::string& msg) { g_create_error = msg; delete h; return 1; };
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0)
    return fail("no CUDA device: this library has no CPU fallback");
  if (device < 0 || device >= ndev) return fail("bad device index");
  if (num_envs <= 0) return fail("num_envs must be positive");
  if (cudaSetDevice(device) != cudaSuccess) return fail("cudaSetDevice failed");
  h->device = device; h->num_envs = num_envs; h->flags = flags; h->seed = seed; h->env_offset = env_offset;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return fail("cudaGetDeviceProperties failed");
  h->sm_count = prop.multiProcessorCount;
  { const char* e = getenv("FLYB_SM_LIMIT"); if (e && atoi(e) > 0 && atoi(e) < h->sm_count) h->sm_count = atoi(e); }   /* kernel experiments: blocks on fewer SMs */
  {  /* set-up of flyb_step_noalloc: device address of the static scratch, once per device, with that device current */
    std::string e;
    h->noalloc_scratch = noalloc_scratch_of(device, &e);
    if (!h->noalloc_scratch) return fail(e);
  }
  FlyArena arena;
  if (flags & FLYB_FLAG_LANES16) h->lanes = 16;
  { const char* e = getenv("FLYB_LANES"); if (e) h->lanes = atoi(e); }   /* kernel experiments */
  if (h->lanes != 32 && h->lanes != 16 && h->lanes != 8) return fail("FLYB_LANES must be 32, 16 or 8");
  std::string e = build_dmodel<float>(model, flags & 0xffff, &h->m, &arena, h->lanes);
  if (!e.empty()) return fail(e);
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)
  CKC(cudaMalloc(&h->d_arena, arena.buf.size()));
  CKC(cudaMemcpy(h->d_arena, arena.buf.data(), arena.buf.size(), cudaMemcpyHostToDevice));
  fly_relocate(&h->m, h->d_arena);
  const DModel<float>& m = h->m;
  const size_t N = (size_t)num_envs;
  h->sreal_stride = ((m.nq + 2 * m.nv + m.na + 1) + 3) & ~3;
  CKC(cudaMalloc(&h->d_sreal, N * h->sreal_stride * sizeof(float)));
  CKC(cudaMemset(h->d_sreal, 0, N * h->sreal_stride * sizeof(float)));
  CKC(cudaMalloc(&h->d_sint, N * FLY_SI_COUNT * sizeof(int)));
  {
    std::vector<int> si(N * FLY_SI_COUNT, 0);
    for (size_t i = 0; i < N; i++) { si[i * FLY_SI_COUNT + FLY_SI_FORCED_CLIP] = -1; si[i * FLY_SI_COUNT + FLY_SI_RESET_NEXT] = 1; }
    CKC(cudaMemcpy(h->d_sint, si.data(), si.size() * sizeof(int), cudaMemcpyHostToDevice));
  }
  CKC(cudaMalloc(&h->d_stats, N * FLY_NSTAT * sizeof(float)));
  CKC(cudaMemset(h->d_stats, 0, N * FLY_NSTAT * sizeof(float)));
  CKC(cudaEventCreateWithFlags(&h->last_launch, cudaEventDisableTiming));
  if (flags & FLYB_FLAG_TRACE) {
    h->trace_real_stride = FLY_MAXCON + m.nv + m.nsensordata + FLY_MAXEFC;
    CKC(cudaMalloc(&h->d_trace_int, N * FLY_TR_INTS * sizeof(int)));
    CKC(cudaMalloc(&h->d_trace_real, N * h->trace_real_stride * sizeof(float)));
    CKC(cudaMemset(h->d_trace_int, 0, N * FLY_TR_INTS * sizeof(int)));
    CKC(cudaMemset(h->d_trace_real, 0, N * h->trace_real_stride * sizeof(float)));
  }
  /* scratch slots: one per resident warp, sized for the all-bulk layout so any smem_level fits */
  size_t fb = 0, bb = 0;
  ws_bind<float>(m, nullptr, nullptr, nullptr, &fb, &bb, 0, h->lanes);
  h->bulk_stride = (bb + 255) & ~(size_t)255;
  h->bulk_slots = h->sm_count * 32 * (32 / h->lanes);
  if ((size_t)h->bulk_slots > N + 128) h->bulk_slots = (int)((N + 127) & ~(size_t)127);  /* whole blocks of up to 128 envs */
  CKC(cudaMalloc(&h->d_bulk, (size_t)h->bulk_slots * h->bulk_stride));
  CKC(cudaMemset(h->d_bulk, 0, (size_t)h->bulk_slots * h->bulk_stride));
  CKC(cudaMalloc(&h->d_counter, sizeof(int)));
  CKC(cudaMalloc(&h->d_work, N * sizeof(int)));
  CKC(cudaMalloc(&h->d_perm, N * sizeof(int)));
  CKC(cudaMemset(h->d_work, 0, N * sizeof(int)));
  { const char* e = getenv("FLYB_BINNING"); if (e) h->binning = atoi(e); }
  /* staging */
  CKC(cudaStreamCreateWithFlags(&h->host_stream, cudaStreamNonBlocking));
  CKC(cudaMallocHost(&h->h_action, N * m.action_dim * sizeof(float)));
  CKC(cudaMallocHost(&h->h_obs, N * m.obs_dim * sizeof(float)));
  CKC(cudaMallocHost(&h->h_rd, N * 2 * sizeof(float)));
  CKC(cudaMallocHost(&h->h_st, N * sizeof(int)));
  CKC(cudaMallocHost(&h->h_mask, N));
  CKC(cudaMalloc(&h->d_action, N * m.action_dim * sizeof(float)));
  CKC(cudaMalloc(&h->d_obs, N * m.obs_dim * sizeof(float)));
  CKC(cudaMalloc(&h->d_rd, N * 2 * sizeof(float)));
  CKC(cudaMalloc(&h->d_st, N * sizeof(int)));
  CKC(cudaMalloc(&h->d_mask, N));
  CKC(cudaMemset(h->d_action, 0, N * m.action_dim * sizeof(float)));
#undef CKC
  const char* lv = getenv("FLYB_SMEM_LEVEL");
  if (lv) h->smem_level = atoi(lv);
  { const char* e = getenv("FLYB_NO_SPEC"); if (e) h->no_spec = atoi(e); }
  const char* wp = getenv("FLYB_WARPS");
  if (wp) h->warps = atoi(wp);
  *out = h;
  return 0;
}

void flyb_destroy(flyb_handle h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaDeviceSynchronize();
  cudaFree(h->d_arena); cudaFree(h->d_sreal); cudaFree(h->d_sint); cudaFree(h->d_stats); cudaFree(h->d_trace_int);
  cudaFree(h->d_trace_real); cudaFree(h->d_bulk); cudaFree(h->d_counter); cudaFree(h->d_work); cudaFree(h->d_perm); cudaFree(h->d_clip_rows);
  cudaFree(h->d_wbpg); cudaFree(h->d_wbpg_phase);
  cudaFree(h->d_hf); cudaFree(h->d_trench_meta); cudaFree(h->d_trench_y); cudaFree(h->d_eyes); cudaFree(h->d_eye_terrain); cudaFree(h->d_hf_bmax); cudaFree(h->d_eye_counter);
  cudaFree(h->d_clip_off); cudaFree(h->d_clip_len); cudaFree(h->d_action); cudaFree(h->d_obs); cudaFree(h->d_rd);
  cudaFree(h->d_st); cudaFree(h->d_mask);
  cudaFreeHost(h->h_action); cudaFreeHost(h->h_obs); cudaFreeHost(h->h_rd); cudaFreeHost(h->h_st); cudaFreeHost(h->h_mask);
  if (h->host_stream) cudaStreamDestroy(h->host_stream);
  if (h->last_launch) cudaEventDestroy(h->last_launch);
  delete h;
}

int64_t flyb_info(flyb_handle h, int what) {
  switch (what) {
    case FLYB_INFO_NUM_ENVS: return h->num_envs;
    case FLYB_INFO_OBS_DIM: return h->m.obs_dim;
    case FLYB_INFO_ACTION_DIM: return h->m.action_dim;
    case FLYB_INFO_NQ: return h->m.nq;
    case FLYB_INFO_NV: return h->m.nv;
    case FLYB_INFO_NA: return h->m.na;
    case FLYB_INFO_TRACE_INTS: return FLY_TR_INTS;
    case FLYB_INFO_TRACE_REALS: return h->trace_real_stride;
    case FLYB_INFO_LAUNCHES: return h->launches;
    case FLYB_INFO_SMEM_BYTES: return (int64_t)(h->fast_stride * h->warps * (32 / h->lanes));
    case FLYB_INFO_BULK_BYTES: return (int64_t)h->bulk_stride;
    case FLYB_INFO_GRID: return h->grid;
    case FLYB_INFO_MAXCON: return FLY_MAXCON;
    case FLYB_INFO_MAXEFC: return FLY_MAXEFC;
    case FLYB_INFO_STATE_INTS: return FLY_SI_COUNT;
    case FLYB_INFO_WARPS_PER_BLOCK: return h->warps;
    case FLYB_INFO_SMEM_LEVEL: return h->smem_level;
    case FLYB_INFO_SPEC_KERNEL: return h->spec;
    case FLYB_INFO_LANES: return h->lanes;
  }
  return -1;
}

int flyb_configure(flyb_handle h, int smem_level, int warps_per_block, int blocks_per_sm) {
  if (smem_level < 0 || smem_level > 6) { h->err = "smem_level must be 0..6"; return 1; }
  if (!pick_kernel(h->lanes, warps_per_block, smem_level, -1) && !pick_kernel(h->lanes, warps_per_block, smem_level, spec_of(h->m))) {
    h->err = "no step kernel compiled for this (lanes, warps_per_block, smem_level); see KERNEL_UNITS in flybody_b200/engine.py"; return 1;
  }
  h->smem_level = smem_level; h->warps = warps_per_block; h->blocks_per_sm = blocks_per_sm;
  return 0;
}

int flyb_use_generic_kernel(flyb_handle h, int generic) { h->no_spec = generic != 0; return 0; }

int flyb_load_clips(flyb_handle h, const float* rows, const int32_t* offsets, const int32_t* lengths, int n_clips) {
  CK(cudaSetDevice(h->device));
  if (n_clips <= 0) { h->err = "n_clips must be positive"; return 1; }
  size_t nrows = 0;
  for (int i = 0; i < n_clips; i++) {
    if (lengths[i] <= 0 || offsets[i] < 0) { h->err = "bad clip offsets/lengths"; return 1; }
    size_t e = (size_t)offsets[i] + lengths[i];
    if (e > nrows) nrows = e;
  }
  cudaFree(h->d_clip_rows); cudaFree(h->d_clip_off); cudaFree(h->d_clip_len);
  h->d_clip_rows = nullptr; h->d_clip_off = nullptr; h->d_clip_len = nullptr; h->n_clips = 0;
  CK(cudaMalloc(&h->d_clip_rows, nrows * h->m.clip_width * sizeof(float)));
  CK(cudaMalloc(&h->d_clip_off, n_clips * sizeof(int)));
  CK(cudaMalloc(&h->d_clip_len, n_clips * sizeof(int)));
  CK(cudaMemcpy(h->d_clip_rows, rows, nrows * h->m.clip_width * sizeof(float), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(h->d_clip_off, offsets, n_clips * sizeof(int), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(h->d_clip_len, lengths, n_clips * sizeof(int), cudaMemcpyHostToDevice));
  h->n_clips = n_clips;
  return 0;
}

int flyb_load_wbpg(flyb_handle h, const double* beat_freqs, const int32_t* offsets, const double* traj, const double* phase,
                   int n_freqs, double base_freq, double rel_freq_range, double rate, double dt_ctrl) {
  CK(cudaSetDevice(h->device));
  if (h->m.task_type != FLYB_TASK_FLIGHT_IMITATION && h->m.task_type != FLYB_TASK_VISION_FLIGHT) {
    h->err = "the wing-beat pattern generator belongs to the flight tasks"; return 1;
  }
  if (n_freqs <= 0 || offsets[0] != 0) { h->err = "bad wing-beat tables"; return 1; }
  for (int f = 0; f < n_freqs; f++)
    if (offsets[f + 1] - offsets[f] < 2) { h->err = "every wing-beat table needs at least two rows"; return 1; }
  const size_t rows = (size_t)offsets[n_freqs];
  /* one allocation: freqs | phase | traj (+1 padding row: reset reads row step+1) | offsets */
  const size_t b_freq = (size_t)n_freqs * 8, b_phase = rows * 8, b_traj = (rows + 1) * 6 * 8, b_off = (size_t)(n_freqs + 1) * 4;
  cudaFree(h->d_wbpg); h->d_wbpg = nullptr; h->wb.nfreq = 0;
  CK(cudaMalloc(&h->d_wbpg, b_freq + b_phase + b_traj + b_off));
  CK(cudaMemset(h->d_wbpg, 0, b_freq + b_phase + b_traj + b_off));
  char* base = (char*)h->d_wbpg;
  CK(cudaMemcpy(base, beat_freqs, b_freq, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(base + b_freq, phase, b_phase, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(base + b_freq + b_phase, traj, rows * 6 * 8, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(base + b_freq + b_phase + b_traj, offsets, b_off, cudaMemcpyHostToDevice));
  h->wb.beat_freqs = (const double*)base; h->wb.phase = (const double*)(base + b_freq);
  h->wb.traj = (const double*)(base + b_freq + b_phase); h->wb.offsets = (const int*)(base + b_freq + b_phase + b_traj);
  h->wb.base_freq = base_freq; h->wb.rel_range = rel_freq_range; h->wb.rate = rate; h->wb.dt_ctrl = dt_ctrl;
  h->wb.nfreq = n_freqs;
  return 0;
}

int flyb_set_wbpg_phase(flyb_handle h, const double* phase) {
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  if (!phase) { cudaFree(h->d_wbpg_phase); h->d_wbpg_phase = nullptr; return 0; }
  if (!h->d_wbpg_phase) CK(cudaMalloc(&h->d_wbpg_phase, (size_t)h->num_envs * sizeof(double)));
  CK(cudaMemcpy(h->d_wbpg_phase, phase, (size_t)h->num_envs * sizeof(double), cudaMemcpyHostToDevice));
  return 0;
}

int flyb_set_clip_ids(flyb_handle h, const int32_t* ids) {
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  if (ids && h->n_clips > 0)
    for (int i = 0; i < h->num_envs; i++)
      if (ids[i] >= h->n_clips) { h->err = "flyb_set_clip_ids: clip id " + std::to_string(ids[i]) + " of env " + std::to_string(i) + " is beyond the loaded table (" + std::to_string(h->n_clips) + " clips)"; return 1; }
  std::vector<int> si((size_t)h->num_envs * FLY_SI_COUNT);
  CK(cudaMemcpy(si.data(), h->d_sint, si.size() * sizeof(int), cudaMemcpyDeviceToHost));
  for (int i = 0; i < h->num_envs; i++) si[(size_t)i * FLY_SI_COUNT + FLY_SI_FORCED_CLIP] = ids ? ids[i] : -1;
  CK(cudaMemcpy(h->d_sint, si.data(), si.size() * sizeof(int), cudaMemcpyHostToDevice));
  return 0;
}

int flyb_reset(flyb_handle h, const uint8_t* mask, float* obs, int32_t* step_type, void* stream) {
  FlyIO<float> io = make_io(h, nullptr, obs, nullptr, nullptr, step_type);
  io.force_reset = 1; io.reset_mask = mask;
  return launch(h, io, (cudaStream_t)stream);
}

int flyb_step(flyb_handle h, const float* action, float* obs, float* reward, float* discount, int32_t* step_type,
              void* stream) {
  FlyIO<float> io = make_io(h, action, obs, reward, discount, step_type);
  return launch(h, io, (cudaStream_t)stream);
}

int flyb_hfield_contacts(int device, const float* data, int nrow, int ncol, const double* pose12, const double* size4, int nquery,
                         const int32_t* types, const double* geoms16, int maxcon, double* out) {
  auto fail = [&](const std::string& msg) { g_create_error = msg; return 1; };
  if (nrow < 2 || ncol < 2 || nquery <= 0 || maxcon <= 0) return fail("flyb_hfield_contacts: bad sizes");
  if (cudaSetDevice(device) != cudaSuccess) return fail("flyb_hfield_contacts: no CUDA device (there is no CPU fallback)");
  FlyHfield hf;
  float* d_data = nullptr; int* d_types = nullptr; double *d_geoms = nullptr, *d_out = nullptr;
  const size_t nd = (size_t)nrow * ncol, no = (size_t)nquery * (1 + 7 * maxcon);
  cudaError_t e = cudaSuccess;
#define CKH(call) if (e == cudaSuccess) e = (call)
  CKH(cudaMalloc(&d_data, nd * sizeof(float)));
  CKH(cudaMalloc(&d_types, nquery * sizeof(int)));
  CKH(cudaMalloc(&d_geoms, (size_t)nquery * 16 * sizeof(double)));
  CKH(cudaMalloc(&d_out, no * sizeof(double)));
  CKH(cudaMemcpy(d_data, data, nd * sizeof(float), cudaMemcpyHostToDevice));
  CKH(cudaMemcpy(d_types, types, nquery * sizeof(int), cudaMemcpyHostToDevice));
  CKH(cudaMemcpy(d_geoms, geoms16, (size_t)nquery * 16 * sizeof(double), cudaMemcpyHostToDevice));
  CKH(cudaMemset(d_out, 0, no * sizeof(double)));
  if (e == cudaSuccess) {
    hf.data = d_data; hf.nrow = nrow; hf.ncol = ncol;
    for (int k = 0; k < 3; k++) hf.pos[k] = pose12[k];
    for (int k = 0; k < 9; k++) hf.mat[k] = pose12[3 + k];
    for (int k = 0; k < 4; k++) hf.size[k] = size4[k];
    fly_hfield_kernel<<<(nquery + 127) / 128, 128>>>(hf, nquery, d_types, d_geoms, maxcon, d_out);
    e = cudaGetLastError();
  }
  CKH(cudaMemcpy(out, d_out, no * sizeof(double), cudaMemcpyDeviceToHost));
#undef CKH
  cudaFree(d_data); cudaFree(d_types); cudaFree(d_geoms); cudaFree(d_out);
  if (e != cudaSuccess) return fail(std::string("flyb_hfield_contacts: ") + cudaGetErrorString(e));
  return 0;
}

int flyb_load_terrain(flyb_handle h, const float* heights, int n_terrains, const double* trench_meta, const float* trench_y,
                      int trench_stride) {
  CK(cudaSetDevice(h->device));
  if (h->m.hfield_geom < 0) { h->err = "flyb_load_terrain: the model has no height-field geom (vision_flight only)"; return 1; }
  if (n_terrains <= 0 || !heights) { h->err = "flyb_load_terrain: no terrains"; return 1; }
  if ((trench_meta != nullptr) != (trench_y != nullptr) || (trench_meta && trench_stride < 1)) { h->err = "flyb_load_terrain: trench_meta and trench_y go together"; return 1; }
  if (trench_meta)
    for (int t = 0; t < n_terrains; t++)
      if (trench_meta[3 * t] < 0 || trench_meta[3 * t] > trench_stride) { h->err = "flyb_load_terrain: trench centre line longer than trench_stride"; return 1; }
  cudaFree(h->d_hf); cudaFree(h->d_trench_meta); cudaFree(h->d_trench_y); cudaFree(h->d_hf_bmax);
  h->d_hf = nullptr; h->d_trench_meta = nullptr; h->d_trench_y = nullptr; h->d_hf_bmax = nullptr;
  h->m.hf_pool = nullptr; h->m.trench_meta = nullptr; h->m.trench_y = nullptr; h->m.nterrain = 0; h->m.hf_bmax = nullptr;
  const size_t nd = (size_t)n_terrains * h->m.hfield_nrow * h->m.hfield_ncol;
  CK(cudaMalloc(&h->d_hf, nd * sizeof(float)));
  CK(cudaMemcpy(h->d_hf, heights, nd * sizeof(float), cudaMemcpyHostToDevice));
  if (trench_meta) {
    CK(cudaMalloc(&h->d_trench_meta, (size_t)n_terrains * 3 * sizeof(double)));
    CK(cudaMalloc(&h->d_trench_y, (size_t)n_terrains * trench_stride * sizeof(float)));
    CK(cudaMemcpy(h->d_trench_meta, trench_meta, (size_t)n_terrains * 3 * sizeof(double), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(h->d_trench_y, trench_y, (size_t)n_terrains * trench_stride * sizeof(float), cudaMemcpyHostToDevice));
  }
  {  /* block maxima for the eye renderer's empty-space skipping */
    const int nr = h->m.hfield_nrow - 1, nc = h->m.hfield_ncol - 1;
    const int nbr = (nr + FLY_EYE_BLOCK - 1) >> FLY_EYE_SHIFT, nbc = (nc + FLY_EYE_BLOCK - 1) >> FLY_EYE_SHIFT;
    const int nbr2 = (nr + (1 << FLY_EYE_SHIFT2) - 1) >> FLY_EYE_SHIFT2, nbc2 = (nc + (1 << FLY_EYE_SHIFT2) - 1) >> FLY_EYE_SHIFT2;
    const size_t stride = (size_t)nbr * nbc + (size_t)nbr2 * nbc2 + 1;
    CK(cudaMalloc(&h->d_hf_bmax, (size_t)n_terrains * stride * sizeof(float)));
    fly_eyes_blockmax_kernel<<<(n_terrains * nbr * nbc + 127) / 128, 128>>>(h->d_hf, n_terrains, h->m.hfield_nrow, h->m.hfield_ncol, -0.01f,
                                                                          FLY_EYE_SHIFT, nbr, nbc, stride, h->d_hf_bmax);
    fly_eyes_blockmax_kernel<<<(n_terrains * nbr2 * nbc2 + 127) / 128, 128>>>(h->d_hf, n_terrains, h->m.hfield_nrow, h->m.hfield_ncol, -0.01f,
                                                                            FLY_EYE_SHIFT2, nbr2, nbc2, stride, h->d_hf_bmax + (size_t)nbr * nbc);
    fly_eyes_gmax_kernel<<<(n_terrains + 127) / 128, 128>>>(h->d_hf_bmax, n_terrains, nbr * nbc, stride);
    CK(cudaGetLastError());
    CK(cudaDeviceSynchronize());
    h->m.hf_bmax = h->d_hf_bmax; h->m.hf_nbr = nbr; h->m.hf_nbc = nbc; h->m.hf_nbc2 = nbc2; h->m.hf_bstride = (long long)stride;
  }
  if (!h->d_eyes) {
    CK(cudaMalloc(&h->d_eyes, (size_t)h->num_envs * 24 * sizeof(float)));
    CK(cudaMalloc(&h->d_eye_terrain, (size_t)h->num_envs * sizeof(int)));
    CK(cudaMalloc(&h->d_eye_counter, sizeof(unsigned)));
    CK(cudaMemset(h->d_eyes, 0, (size_t)h->num_envs * 24 * sizeof(float)));
    CK(cudaMemset(h->d_eye_terrain, 0, (size_t)h->num_envs * sizeof(int)));
  }
  h->m.hf_pool = h->d_hf; h->m.trench_meta = h->d_trench_meta; h->m.trench_y = h->d_trench_y;
  h->m.nterrain = n_terrains; h->m.trench_stride = trench_stride;
  return 0;
}

int flyb_render(flyb_handle h, int width, int height, float fovy_deg, const float* colors8, uint8_t* out, void* stream) {
  CK(cudaSetDevice(h->device));
  if (!h->d_hf || !h->d_eyes) { h->err = "flyb_render: no terrain loaded (flyb_load_terrain)"; return 1; }
  if (width <= 0 || height <= 0 || !out || !colors8) { h->err = "flyb_render: bad arguments"; return 1; }
  FlyScene sc;
  sc.data = h->d_hf; sc.nrow = h->m.hfield_nrow; sc.ncol = h->m.hfield_ncol; sc.half = (float)h->m.hfield_size[0];
  sc.zoff = -0.01f;   /* z of the terrain geom (tasks/arenas/hills.py:190, 215) */
  sc.bmax = h->d_hf_bmax;
  sc.nbr = (h->m.hfield_nrow - 2 + FLY_EYE_BLOCK) >> FLY_EYE_SHIFT; sc.nbc = (h->m.hfield_ncol - 2 + FLY_EYE_BLOCK) >> FLY_EYE_SHIFT;
  sc.nbr2 = (h->m.hfield_nrow - 2 + (1 << FLY_EYE_SHIFT2)) >> FLY_EYE_SHIFT2; sc.nbc2 = (h->m.hfield_ncol - 2 + (1 << FLY_EYE_SHIFT2)) >> FLY_EYE_SHIFT2;
  for (int k = 0; k < 3; k++) { sc.rgb[k] = colors8[k]; sc.sky[k] = colors8[3 + k]; }
  sc.ambient = colors8[6]; sc.diffuse = colors8[7];
  const int ncam = 2 * h->num_envs;
  const size_t npix = (size_t)ncam * width * height;
  unsigned blocks = (unsigned)((npix + 255) / 256), full = (unsigned)h->sm_count * 8u;
  if (blocks > full) blocks = full;
  CK(cudaMemsetAsync(h->d_eye_counter, 0, sizeof(unsigned), (cudaStream_t)stream));
  static const bool sched = []() { const char* e = getenv("FLYB_EYES_SCHED"); return !(e && atoi(e) == 0); }();   /* 0: the fixed-order kernel */
  if (sched)
    fly_eyes_sched_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(sc, ncam, h->d_eye_terrain, h->d_eyes, width, height,
                                                                   tanf(0.5f * fovy_deg * 3.14159265358979f / 180.0f), out, h->d_eye_counter);
  else
  fly_eyes_env_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(sc, ncam, h->d_eye_terrain, h->d_eyes, width, height,
                                                               tanf(0.5f * fovy_deg * 3.14159265358979f / 180.0f), out, h->d_eye_counter);
  CK(cudaGetLastError());
  h->launches++;
  return 0;
}

int flyb_render_eyes(int device, const float* data, int nterrain, int nrow, int ncol, float half_extent, float z_offset, int ncam,
                     const int32_t* terrain_id, const float* cams12, int width, int height, float fovy_deg, const float* colors8,
                     uint8_t* out) {
  auto fail = [&](const std::string& msg) { g_create_error = msg; return 1; };
  if (nterrain <= 0 || nrow < 2 || ncol < 2 || ncam <= 0 || width <= 0 || height <= 0) return fail("flyb_render_eyes: bad sizes");
  for (int i = 0; i < ncam; i++) if (terrain_id[i] < 0 || terrain_id[i] >= nterrain) return fail("flyb_render_eyes: terrain id out of range");
  if (cudaSetDevice(device) != cudaSuccess) return fail("flyb_render_eyes: no CUDA device (there is no CPU fallback)");
  float *d_data = nullptr, *d_cams = nullptr; int* d_tid = nullptr; unsigned char* d_out = nullptr;
  const size_t nd = (size_t)nterrain * nrow * ncol, no = (size_t)ncam * width * height * 3;
  cudaError_t e = cudaSuccess;
#define CKH(call) if (e == cudaSuccess) e = (call)
  CKH(cudaMalloc(&d_data, nd * sizeof(float)));
  CKH(cudaMalloc(&d_cams, (size_t)ncam * 12 * sizeof(float)));
  CKH(cudaMalloc(&d_tid, ncam * sizeof(int)));
  CKH(cudaMalloc(&d_out, no));
  CKH(cudaMemcpy(d_data, data, nd * sizeof(float), cudaMemcpyHostToDevice));
  CKH(cudaMemcpy(d_cams, cams12, (size_t)ncam * 12 * sizeof(float), cudaMemcpyHostToDevice));
  CKH(cudaMemcpy(d_tid, terrain_id, ncam * sizeof(int), cudaMemcpyHostToDevice));
  if (e == cudaSuccess) {
    FlyScene sc;
    sc.data = d_data; sc.nrow = nrow; sc.ncol = ncol; sc.half = half_extent; sc.zoff = z_offset;
    for (int k = 0; k < 3; k++) { sc.rgb[k] = colors8[k]; sc.sky[k] = colors8[3 + k]; }
    sc.ambient = colors8[6]; sc.diffuse = colors8[7];
    sc.bmax = nullptr; sc.nbr = sc.nbc = sc.nbr2 = sc.nbc2 = 0;   /* the stand-alone entry point walks every cell (the reference path of the tests) */
    const size_t npix = (size_t)ncam * width * height;
    fly_eyes_kernel<<<(unsigned)((npix + 255) / 256), 256>>>(sc, ncam, d_tid, d_cams, width, height,
                                                            tanf(0.5f * fovy_deg * 3.14159265358979f / 180.0f), d_out);
    e = cudaGetLastError();
  }
  CKH(cudaMemcpy(out, d_out, no, cudaMemcpyDeviceToHost));
#undef CKH
  cudaFree(d_data); cudaFree(d_cams); cudaFree(d_tid); cudaFree(d_out);
  if (e != cudaSuccess) return fail(std::string("flyb_render_eyes: ") + cudaGetErrorString(e));
  return 0;
}

int flyb_fma_peak(int device, double* tflops) {
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 1;
  float* d = nullptr;
  if (cudaMalloc(&d, 256) != cudaSuccess) return 1;
  const int blocks = prop.multiProcessorCount * 8, iters = 1 << 16;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 8; rep++) {   /* first repetitions warm the clocks; the best of the rest is the peak */
    cudaEventRecord(e0);
    fly_fma_peak_kernel<<<blocks, 256>>>(d, iters);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return 1; }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep >= 2 && ms < best) best = ms;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(d);
  *tflops = 2.0 * 16.0 * (double)iters * 256.0 * (double)blocks / ((double)best * 1e-3) / 1e12;
  return 0;
}

int flyb_step_noalloc(flyb_handle h, const float* action, float* obs, float* reward, float* discount, int32_t* step_type,
                      void* stream) {
  /* no cudaMalloc / cudaFree / allocating library call on this path: the workspace slots and the work counter live in
     g_noalloc_scratch (address resolved at flyb_create); launch() only sets a function attribute, queries occupancy,
     clears the counter (cudaMemsetAsync) and launches */
  if (h->num_envs > FLYB_NOALLOC_MAX_ENVS) {
    h->err = "flyb_step_noalloc: num_envs " + std::to_string(h->num_envs) + " exceeds FLYB_NOALLOC_MAX_ENVS (" +
             std::to_string(FLYB_NOALLOC_MAX_ENVS) + "); use flyb_step";
    return 1;
  }
  if (!h->noalloc_scratch) { h->err = "flyb_step_noalloc: static scratch was not resolved at flyb_create"; return 1; }
  const size_t room = (size_t)FLYB_NOALLOC_SCRATCH_BYTES - FLYB_NOALLOC_HEADER;
  const int slots = (int)(room / h->bulk_stride);
  if (slots < h->warps * (32 / h->lanes)) {
    h->err = "flyb_step_noalloc: the workspace of one block (" + std::to_string((size_t)h->warps * (32 / h->lanes) * h->bulk_stride) +
             " bytes) exceeds FLYB_NOALLOC_SCRATCH_BYTES (" + std::to_string((size_t)FLYB_NOALLOC_SCRATCH_BYTES) + ")";
    return 1;
  }
  LaunchScratch sc;
  sc.counter = (int*)h->noalloc_scratch;
  sc.bulk = (char*)h->noalloc_scratch + FLYB_NOALLOC_HEADER;
  sc.slots = slots;
  FlyIO<float> io = make_io(h, action, obs, reward, discount, step_type);
  return launch(h, io, (cudaStream_t)stream, &sc);
}

int flyb_reset_host(flyb_handle h, const uint8_t* mask, float* obs, int32_t* step_type) {
  CK(cudaSetDevice(h->device));
  const size_t N = h->num_envs;
  cudaStream_t s = h->host_stream;
  CK(cudaStreamWaitEvent(s, h->last_launch, 0));   /* a device-pointer call still in flight on the caller's stream goes first */
  if (mask) { memcpy(h->h_mask, mask, N); CK(cudaMemcpyAsync(h->d_mask, h->h_mask, N, cudaMemcpyHostToDevice, s)); }
  if (flyb_reset(h, mask ? h->d_mask : nullptr, h->d_obs, h->d_st, s)) return 1;
  CK(cudaMemcpyAsync(h->h_obs, h->d_obs, N * h->m.obs_dim * sizeof(float), cudaMemcpyDeviceToHost, s));
  CK(cudaMemcpyAsync(h->h_st, h->d_st, N * sizeof(int), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  for (size_t i = 0; i < N; i++) {
    if (mask && !mask[i]) continue;
    if (obs) memcpy(obs + i * h->m.obs_dim, h->h_obs + i * h->m.obs_dim, h->m.obs_dim * sizeof(float));
    if (step_type) step_type[i] = h->h_st[i];
  }
  return 0;
}

/* page-locked host memory (cudaHostAlloc / cudaHostRegister, e.g. a pinned torch tensor) can be the DMA target itself */
static bool is_pinned(const void* p) {
  cudaPointerAttributes at;
  if (!p || cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost;
}

/* device-visible address of page-locked host memory (null: not mapped for this device) */
static void* host_dev_ptr(void* p) {
  void* d = nullptr;
  if (!p) return nullptr;
  if (cudaHostGetDevicePointer(&d, p, 0) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return d;
}

int flyb_step_host(flyb_handle h, const float* action, float* obs, float* reward, float* discount, int32_t* step_type) {
  CK(cudaSetDevice(h->device));
  const size_t N = h->num_envs, A = h->m.action_dim, O = h->m.obs_dim;
  if (!action) { h->err = "flyb_step_host: action must not be NULL"; return 1; }
  cudaStream_t s = h->host_stream;
  CK(cudaStreamWaitEvent(s, h->last_launch, 0));   /* a device-pointer call still in flight on the caller's stream goes first */
  /* pageable caller buffers go through the library's pinned staging buffers (one extra host copy each way); page-locked
     ones are read and written directly */
  const bool pa = is_pinned(action), po = is_pinned(obs), pr = is_pinned(reward), pd = is_pinned(discount), ps = is_pinned(step_type);
  if (!pa) memcpy(h->h_action, action, N * A * sizeof(float));
  CK(cudaMemcpyAsync(h->d_action, pa ? action : h->h_action, N * A * sizeof(float), cudaMemcpyHostToDevice, s));
  /* outputs: every warp writes its env's finished observation row, reward, discount and step type straight into the page-locked
     host buffers (FlyIO::host_*), so the 3 KB per env cross PCIe while the other envs are still stepping and nothing is left to
     copy when the kernel ends (measured: 1.2 ms of 48.7 per control step at 16 384 walking envs).  FLYB_HOST_DIRECT=0, or a
     buffer the device cannot address, falls back to device -> host copies after the launch. */
  static const bool direct_ok = []() { const char* e = getenv("FLYB_HOST_DIRECT"); return !(e && atoi(e) == 0); }();
  float* t_obs = obs ? (po ? obs : h->h_obs) : nullptr;
  float* t_rew = reward ? (pr ? reward : h->h_rd) : nullptr;
  float* t_dis = discount ? (pd ? discount : h->h_rd + N) : nullptr;
  int* t_st = step_type ? (ps ? (int*)step_type : h->h_st) : nullptr;
  FlyIO<float> io = make_io(h, h->d_action, h->d_obs, h->d_rd, h->d_rd + N, h->d_st);
  bool direct = direct_ok;
  if (direct) {
    io.host_obs = (float*)host_dev_ptr(t_obs); io.host_reward = (float*)host_dev_ptr(t_rew);
    io.host_discount = (float*)host_dev_ptr(t_dis); io.host_step_type = (int*)host_dev_ptr(t_st);
    if ((t_obs && !io.host_obs) || (t_rew && !io.host_reward) || (t_dis && !io.host_discount) || (t_st && !io.host_step_type)) {
      direct = false;
      io.host_obs = io.host_reward = io.host_discount = nullptr; io.host_step_type = nullptr;
    }
  }
  if (launch(h, io, s)) return 1;
  if (!direct) {
    if (t_obs) CK(cudaMemcpyAsync(t_obs, h->d_obs, N * O * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (t_rew) CK(cudaMemcpyAsync(t_rew, h->d_rd, N * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (t_dis) CK(cudaMemcpyAsync(t_dis, h->d_rd + N, N * sizeof(float), cudaMemcpyDeviceToHost, s));
    if (t_st) CK(cudaMemcpyAsync(t_st, h->d_st, N * sizeof(int), cudaMemcpyDeviceToHost, s));
  }
  CK(cudaStreamSynchronize(s));   /* (the kernel's writes to host memory are complete and visible once it has ended) */
  if (obs && !po) memcpy(obs, h->h_obs, N * O * sizeof(float));
  if (reward && !pr) memcpy(reward, h->h_rd, N * sizeof(float));
  if (discount && !pd) memcpy(discount, h->h_rd + N, N * sizeof(float));
  if (step_type && !ps) memcpy(step_type, h->h_st, N * sizeof(int));
  return 0;
}

int flyb_get_state(flyb_handle h, float* qpos, float* qvel, float* act, float* warm, int32_t* ints) {
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  const DModel<float>& m = h->m;
  const size_t N = h->num_envs, S = h->sreal_stride;
  std::vector<float> sr(N * S);
  CK(cudaMemcpy(sr.data(), h->d_sreal, sr.size() * sizeof(float), cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < N; i++) {
    const float* r = sr.data() + i * S;
    if (qpos) memcpy(qpos + i * m.nq, r, m.nq * sizeof(float));
    if (qvel) memcpy(qvel + i * m.nv, r + m.nq, m.nv * sizeof(float));
    if (act && m.na) memcpy(act + i * m.na, r + m.nq + m.nv, m.na * sizeof(float));
    if (warm) memcpy(warm + i * m.nv, r + m.nq + m.nv + m.na, m.nv * sizeof(float));
  }
  if (ints) CK(cudaMemcpy(ints, h->d_sint, N * FLY_SI_COUNT * sizeof(int), cudaMemcpyDeviceToHost));
  return 0;
}

int flyb_set_state(flyb_handle h, const float* qpos, const float* qvel, const float* act, const float* warm,
                   const int32_t* ints) {
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  const DModel<float>& m = h->m;
  const size_t N = h->num_envs, S = h->sreal_stride;
  std::vector<float> sr(N * S);
  CK(cudaMemcpy(sr.data(), h->d_sreal, sr.size() * sizeof(float), cudaMemcpyDeviceToHost));
  for (size_t i = 0; i < N; i++) {
    float* r = sr.data() + i * S;
    if (qpos) memcpy(r, qpos + i * m.nq, m.nq * sizeof(float));
    if (qvel) memcpy(r + m.nq, qvel + i * m.nv, m.nv * sizeof(float));
    if (act && m.na) memcpy(r + m.nq + m.nv, act + i * m.na, m.na * sizeof(float));
    if (warm) memcpy(r + m.nq + m.nv + m.na, warm + i * m.nv, m.nv * sizeof(float));
  }
  CK(cudaMemcpy(h->d_sreal, sr.data(), sr.size() * sizeof(float), cudaMemcpyHostToDevice));
  if (ints) CK(cudaMemcpy(h->d_sint, ints, N * FLY_SI_COUNT * sizeof(int), cudaMemcpyHostToDevice));
  return 0;
}

int flyb_get_trace(flyb_handle h, int32_t* ints, float* reals) {
  if (!h->d_trace_int) { h->err = "create the engine with FLYB_FLAG_TRACE to record traces"; return 1; }
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  const size_t N = h->num_envs;
  if (ints) CK(cudaMemcpy(ints, h->d_trace_int, N * FLY_TR_INTS * sizeof(int), cudaMemcpyDeviceToHost));
  if (reals) CK(cudaMemcpy(reals, h->d_trace_real, N * h->trace_real_stride * sizeof(float), cudaMemcpyDeviceToHost));
  return 0;
}

int flyb_episode_stats(flyb_handle h, double* out8) {
  CK(cudaSetDevice(h->device));
  CK(cudaDeviceSynchronize());
  const size_t N = h->num_envs;
  std::vector<float> s(N * FLY_NSTAT);
  CK(cudaMemcpy(s.data(), h->d_stats, s.size() * sizeof(float), cudaMemcpyDeviceToHost));
  CK(cudaMemset(h->d_stats, 0, s.size() * sizeof(float)));
  for (int k = 0; k < FLY_NSTAT; k++) out8[k] = 0;
  for (size_t i = 0; i < N; i++) for (int k = 0; k < FLY_NSTAT; k++) out8[k] += s[i * FLY_NSTAT + k];
  return 0;
}

}  /* extern "C" */
