/* fly_kernel.cuh - the step kernel template and its argument block.
 *
 * Every instantiation (warps per block, shared-memory level, model specialisation) is compiled in a translation unit of
 * its own (flyb_kern.cu, one nvcc process each, in parallel - flybody_b200/engine.py: build_library) and handed to the
 * host code in flyb_cuda.cu as a function pointer: the fully inlined kernel takes ~45 s to compile and there are nine. */
#ifndef FLY_KERNEL_CUH_
#define FLY_KERNEL_CUH_
#include <cuda_runtime.h>
#include "fly_env.cuh"
#include "fly_dims_gen.h"

/* sizes of the shipped task models: kernels with a compile-time workspace layout (fly_ws.cuh: ws_layout_c) */
template <int S> struct SpecDims;
#define X(i, ...) template <> struct SpecDims<i> { static constexpr WsDims value = {__VA_ARGS__}; };
FLY_SPEC_MODELS(X)
#undef X

struct KArgs {
  FlyClips clips;
  FlyWbpg wb;
  FlyIO<float> io;
  char* bulk;
  size_t bulk_stride, fast_stride;
  int* counter;
  const int* perm;   /* nullable: position in the launch order -> env id */
  int num_envs, level;
};

/* SPEC >= 0: the model is entry SPEC of fly_dims_gen.h and array offsets are compile-time constants; SPEC < 0: any model,
 * offsets from the kernel parameter `off` */
template <int WARPS, int LEVEL, int SPEC>
__global__ void __launch_bounds__(WARPS * 32, 1) fly_step_kernel(const __grid_constant__ DModel<float> m,
                                                                const __grid_constant__ KArgs a,
                                                                const __grid_constant__ WsOff off) {
  extern __shared__ __align__(16) char smem[];
#ifdef FLY_SYNC_GROUP
  constexpr int GW = FLY_SYNC_GROUP;   /* warps of one lockstep group: fetches its own envs, runs on its own barrier */
#else
  constexpr int GW = WARPS;
#endif
  static_assert(WARPS % GW == 0, "lockstep group must divide the block");
  __shared__ int s_base[WARPS / GW];
  const int warp = threadIdx.x >> 5;
  const int slot = blockIdx.x * WARPS + warp;
  const int grp = warp / GW, gwarp = warp % GW;
  /* every thread keeps its warp's array pointers as (arena base + constant offset): registers / immediates, with the
     address space of each array known to the compiler */
  Ws<float> w;
  if constexpr (SPEC >= 0) {
    constexpr WsOff OFF = ws_layout_c<float>(SpecDims<SPEC>::value, LEVEL);
    ws_attach<float, LEVEL>(w, OFF, smem + warp * a.fast_stride, a.bulk + (size_t)slot * a.bulk_stride);
  } else {
    ws_attach<float, LEVEL>(w, off, smem + warp * a.fast_stride, a.bulk + (size_t)slot * a.bulk_stride);
  }
  for (;;) {  /* the group takes GW consecutive envs at a time and its warps step them stage by stage together */
    FLY_BLOCK_SYNC();
    if (gwarp == 0 && (threadIdx.x & 31) == 0) s_base[grp] = atomicAdd(a.counter, GW);
    FLY_BLOCK_SYNC();
    const int base = s_base[grp];
    const int pos = base + gwarp;
    if (base >= a.num_envs) break;
    fly_env_run<float>(m, w, a.clips, a.wb, a.io, pos < a.num_envs ? (a.perm ? a.perm[pos] : pos) : -1);
  }
}


/* compiled (warps per block, shared-memory level) combinations of the generic kernel; the default one (16, 1) is also
 * compiled per shipped task model (FLY_SPEC_MODELS) */
#define FLY_KERNEL_CONFIGS(X) X(16, 1) X(14, 1) X(12, 1) X(8, 2) X(6, 3)
#define FLY_SPEC_WARPS 16
#define FLY_SPEC_LEVEL 1
#endif
